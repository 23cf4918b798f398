"""Decode an Annex-B H.264 byte string with libavcodec (OpenCV's bundled FFmpeg) -> list of (Y,U,V)."""
import os
import tempfile
import numpy as np


def decode(stream: bytes, width: int, height: int):
    import cv2
    fd, path = tempfile.mkstemp(suffix=".h264")
    try:
        with os.fdopen(fd, "wb") as f:
            f.write(stream)
        cap = cv2.VideoCapture(path, cv2.CAP_FFMPEG)
        cap.set(cv2.CAP_PROP_CONVERT_RGB, 0)
        frames = []
        while True:
            ok, fr = cap.read()
            if not ok:
                break
            fr = np.asarray(fr)
            if fr.ndim == 3:
                fr = fr[:, :, 0]
            flat = fr.reshape(-1)
            n = width * height
            if flat.size >= n * 3 // 2:
                y = flat[:n].reshape(height, width)
                u = flat[n:n + n // 4].reshape(height // 2, width // 2)
                v = flat[n + n // 4:n + n // 2].reshape(height // 2, width // 2)
            else:
                y = flat[:n].reshape(height, width); u = v = None
            frames.append((y.copy(), None if u is None else u.copy(), None if v is None else v.copy()))
        cap.release()
        return frames
    finally:
        os.unlink(path)
