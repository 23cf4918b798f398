"""Decode Annex-B H.264 access units with libavcodec's native `h264` decoder -> (Y,U,V) per frame.

This is the decoder `avdec_h264` wraps (north_star's "decode cleanly with avdec_h264" gate). We bind
the libavcodec that ships inside OpenCV's wheel through ctypes, so that all three planes come back
exactly (cv2.VideoCapture only hands out Y) and decode errors are visible as return codes.
Falls back to cv2.VideoCapture (Y only) if the bundled libraries cannot be bound.
"""
import ctypes as C
import glob
import os
import tempfile
import numpy as np

AV_CODEC_ID_H264 = 27
AVERROR_EAGAIN = -11
AV_EF_EXPLODE = 1 << 3


class _AVFrameHead(C.Structure):  # stable prefix of AVFrame
    _fields_ = [("data", C.c_void_p * 8), ("linesize", C.c_int * 8), ("extended_data", C.c_void_p),
                ("width", C.c_int), ("height", C.c_int), ("nb_samples", C.c_int), ("format", C.c_int)]


class _AVPacketHead(C.Structure):  # stable prefix of AVPacket
    _fields_ = [("buf", C.c_void_p), ("pts", C.c_int64), ("dts", C.c_int64), ("data", C.c_void_p), ("size", C.c_int)]


_libs = None


def _load():
    global _libs
    if _libs is not None:
        return _libs
    import cv2
    d = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)), "opencv_python_headless.libs")
    if not os.path.isdir(d):
        d = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)), "opencv_python.libs")
    mode = C.RTLD_GLOBAL
    for pat in ("libcrypto-*", "libssl-*", "libdrm-*", "libavutil-*", "libswresample-*", "libaom-*", "libvpx-*"):
        for p in glob.glob(os.path.join(d, pat)):
            try:
                C.CDLL(p, mode=mode)
            except OSError:
                pass
    avu = C.CDLL(glob.glob(os.path.join(d, "libavutil-*"))[0], mode=mode)
    avc = C.CDLL(glob.glob(os.path.join(d, "libavcodec-*"))[0], mode=mode)
    avc.avcodec_find_decoder.restype = C.c_void_p
    avc.avcodec_find_decoder.argtypes = [C.c_int]
    avc.avcodec_alloc_context3.restype = C.c_void_p
    avc.avcodec_alloc_context3.argtypes = [C.c_void_p]
    avc.avcodec_open2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    avc.avcodec_free_context.argtypes = [C.c_void_p]
    avc.av_packet_alloc.restype = C.c_void_p
    avc.av_packet_free.argtypes = [C.c_void_p]
    avc.avcodec_send_packet.argtypes = [C.c_void_p, C.c_void_p]
    avc.avcodec_receive_frame.argtypes = [C.c_void_p, C.c_void_p]
    avu.av_frame_alloc.restype = C.c_void_p
    avu.av_frame_free.argtypes = [C.c_void_p]
    avu.av_frame_unref.argtypes = [C.c_void_p]
    avu.av_opt_set_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_int]
    avu.av_log_set_level.argtypes = [C.c_int]
    _libs = (avc, avu)
    return _libs


def _grab(fr, out):
    h = C.cast(fr, C.POINTER(_AVFrameHead)).contents
    planes = []
    for i, (w, hh) in enumerate(((h.width, h.height), (h.width // 2, h.height // 2), (h.width // 2, h.height // 2))):
        ls = h.linesize[i]
        buf = np.ctypeslib.as_array(C.cast(h.data[i], C.POINTER(C.c_uint8)), shape=(hh, ls))
        planes.append(buf[:, :w].copy())
    out.append(tuple(planes))


def decode_aus(aus, strict=True):
    """aus: list of bytes, one access unit each. Returns list of (Y,U,V) uint8 arrays (cropped size).
    Raises RuntimeError on any decoder error when strict."""
    avc, avu = _load()
    avu.av_log_set_level(16 if strict else -8)
    codec = avc.avcodec_find_decoder(AV_CODEC_ID_H264)
    assert codec, "libavcodec has no h264 decoder"
    ctx = avc.avcodec_alloc_context3(codec)
    avu.av_opt_set_int(ctx, b"err_detect", AV_EF_EXPLODE, 0)
    avu.av_opt_set_int(ctx, b"threads", 1, 0)
    assert avc.avcodec_open2(ctx, codec, None) == 0
    pkt = avc.av_packet_alloc()
    fr = avu.av_frame_alloc()
    out = []
    try:
        for k, au in enumerate(list(aus) + [None]):
            if au is None:
                rc = avc.avcodec_send_packet(ctx, None)
            else:
                pad = au + b"\0" * 64
                buf = C.create_string_buffer(pad, len(pad))
                ph = C.cast(pkt, C.POINTER(_AVPacketHead)).contents
                ph.data = C.cast(buf, C.c_void_p).value
                ph.size = len(au)
                rc = avc.avcodec_send_packet(ctx, pkt)
            if rc < 0 and au is not None and strict:
                raise RuntimeError(f"avcodec_send_packet failed on AU {k}: {rc}")
            while True:
                rc = avc.avcodec_receive_frame(ctx, fr)
                if rc < 0:
                    break
                _grab(fr, out)
                avu.av_frame_unref(fr)
    finally:
        p = C.c_void_p(fr); avu.av_frame_free(C.byref(p))
        p = C.c_void_p(pkt); avc.av_packet_free(C.byref(p))
        p = C.c_void_p(ctx); avc.avcodec_free_context(C.byref(p))
    return out


def split_aus(stream: bytes):
    """Split an Annex-B stream of this repo's encoders into access units (an AU starts at an SPS, or at a
    slice NAL whose first_mb_in_slice is 0, i.e. whose first payload bit is 1)."""
    pos = []
    i = 0
    n = len(stream)
    while True:
        j = stream.find(b"\0\0\1", i)
        if j < 0 or j + 3 >= n:
            break
        pos.append(j - 1 if j > 0 and stream[j - 1] == 0 else j)
        i = j + 3
    aus, start, prev_sps = [], None, False
    for k, p in enumerate(pos):
        q = stream.find(b"\1", p) + 1
        t = stream[q] & 31
        new_au = t == 7 or (t in (1, 5) and (stream[q + 1] & 0x80) and not prev_sps)
        if t == 7:
            prev_sps = True
        elif t in (1, 5):
            prev_sps = False
        if new_au:
            if start is not None:
                aus.append(stream[start:p])
            start = p
    if start is not None:
        aus.append(stream[start:])
    return aus


def decode(stream, width=None, height=None):
    """stream: bytes (Annex-B, split here) or list of AU bytes."""
    aus = split_aus(stream) if isinstance(stream, (bytes, bytearray)) else list(stream)
    try:
        return decode_aus(aus)
    except (OSError, IndexError, AttributeError):
        return _decode_cv2(b"".join(aus), width, height)


def _decode_cv2(stream, width, height):
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
            flat = np.asarray(fr).reshape(-1)
            frames.append((flat[:width * height].reshape(height, width).copy(), None, None))
        cap.release()
        return frames
    finally:
        os.unlink(path)
