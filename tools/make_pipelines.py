#!/usr/bin/env python
"""Write pipeline/generic/cuda_h264_* — ceracoder pipeline files that use `cudah264enc` where the reference's
generic files use `x264enc` (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:1-11).

The named-element contract ceracoder relies on is kept: `appsink` (src/ceracoder.c new_buf_cb), `venc_kbps`
(src/gst/encoder_control.c:29-32), `overlay`, `v_delay`, `a_delay` (README.md:249-259).  The encoder line keeps
x264enc's property spelling (speed-preset, key-int-max) so the edit against a reference file is one token.
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "pipeline", "generic")
TAIL = [
    "h264parse config-interval=-1 ! queue max-size-time=10000000000 max-size-buffers=1000 max-size-bytes=41943040 ! mux. ",
]
AUDIO_ALSA = [
    "alsasrc device=hw:2 ! identity name=a_delay signal-handoffs=TRUE ! volume volume=1.0 ! ",
    "audioconvert ! avenc_aac bitrate=131072 ! aacparse ! queue max-size-time=10000000000 max-size-buffers=1000 ! mux. ",
]
AUDIO_TEST = [
    "audiotestsrc ! audio/x-raw,channels=2,rate=48000 ! identity name=a_delay signal-handoffs=TRUE ! ",
    "audioconvert ! avenc_aac bitrate=131072 ! aacparse ! queue max-size-time=10000000000 max-size-buffers=1000 ! mux. ",
]
MUX = ["mpegtsmux name=mux ! ", "appsink name=appsink"]
OVERLAY = "textoverlay text='' valignment=top halignment=right font-desc=\"Monospace, 5\" name=overlay ! queue ! "


def encoder(preset):
    return f"cudah264enc speed-preset={preset} key-int-max=60 name=venc_kbps ! "


def v4l_mjpeg(w, h, fps, preset):
    caps = f"image/jpeg,width={w},height={h}" + (f",framerate={fps}/1" if fps else "")
    return ["v4l2src ! ", "identity name=v_delay signal-handoffs=TRUE ! ", f"{caps} ! jpegparse ! jpegdec ! ", OVERLAY,
            "videoconvert ! video/x-raw,format=I420 ! ", encoder(preset)] + TAIL + AUDIO_ALSA + MUX


def camlink(preset, direct=False):
    """direct: the capture format (YUY2) goes straight into the encoder, which converts it inside its ingest kernel
     — no videoconvert pass on the CPU"""
    return ["v4l2src ! ", "identity name=v_delay signal-handoffs=TRUE ! ", "videorate ! video/x-raw,framerate=30000/1001 ! ", OVERLAY,
            "video/x-raw,format=YUY2 ! " if direct else "videoconvert ! video/x-raw,format=NV12 ! ", encoder(preset)] + TAIL + AUDIO_ALSA + MUX


def test_pattern(w, h, fps, preset):
    return [f"videotestsrc ! video/x-raw,width={w},height={h},framerate={fps}/1 ! identity name=v_delay signal-handoffs=TRUE ! queue ! ", OVERLAY,
            "videoconvert ! video/x-raw,format=I420 ! ", encoder(preset)] + TAIL + AUDIO_TEST + MUX


FILES = {
    "cuda_h264_v4l_mjpeg_1080p30": v4l_mjpeg(1920, 1080, 30, 3),
    "cuda_h264_v4l_mjpeg_1080p": v4l_mjpeg(1920, 1080, 0, 3),
    "cuda_h264_v4l_mjpeg_720p30": v4l_mjpeg(1280, 720, 30, 3),
    "cuda_h264_v4l_mjpeg_720p": v4l_mjpeg(1280, 720, 0, 3),
    "cuda_h264_camlink": camlink(3),
    "cuda_h264_camlink_yuy2": camlink(3, direct=True),
    "cuda_h264_test_1080p30": test_pattern(1920, 1080, 30, 3),
    "cuda_h264_test_1080p60": test_pattern(1920, 1080, 60, 3),
    "cuda_h264_test_720p60": test_pattern(1280, 720, 60, 3),
    "cuda_h264_test_2160p30": test_pattern(3840, 2160, 30, 3),
}

if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for name, lines in FILES.items():
        with open(os.path.join(OUT, name), "w") as f:
            f.write("\n".join(lines) + "\n")
    print(f"wrote {len(FILES)} files to {OUT}")
