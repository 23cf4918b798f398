"""`-m gpu`: the benched kind of stream — rate control on, QP moving, IDR restarts, a bitrate step, scene cuts — held to the oracle
byte for byte at the BASELINE sizes.  The oracle replays the encoder's own (frame type, QP) trace."""
import threading
import numpy as np
import pytest
from parity_util import run_rc_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from ceracoder_b200 import load_library
    return load_library()


def test_1080p60_rate_controlled_130_frames_match_oracle(lib):
    """config 2: 1080p60, 12 -> 2 -> 12 Mbps, GOP 60 (IDR at 0, 60, 120), scene cuts at 50 and 100"""
    from ceracoder_b200.synth import PanSprites
    from h264_decode import decode_aus
    w, h = 1920, 1080
    src = PanSprites(w, h, cut_every=50)
    aus, trace = run_rc_parity(lib, src, w, h, 60, 130, lambda i: 12000 if i < 45 or i >= 100 else 2000, gop=60, recon_every=13)
    idr = [i for i, t in enumerate(trace) if t[0]]
    qps = [t[1] for t in trace]
    assert idr == [0, 60, 120]
    assert max(qps) - min(qps) >= 6, "the bitrate step should have moved the QP"
    assert len(decode_aus(aus)) == 130


def test_2160p30_rate_controlled_12_frames_match_oracle(lib):
    """config 4: 2160p30 @ 20 Mbps with a step to 8 Mbps, GOP 6 so that the run crosses an IDR"""
    from ceracoder_b200.synth import PanSprites
    w, h = 3840, 2160
    src = PanSprites(w, h, cut_every=9)
    aus, trace = run_rc_parity(lib, src, w, h, 30, 12, lambda i: 20000 if i < 5 else 8000, gop=6, recon_every=6)
    assert [i for i, t in enumerate(trace) if t[0]] == [0, 6]


def test_720p60_low_rate_ball_and_pan_match_oracle(lib):
    """config 3: 720p60 @ 4 Mbps"""
    from ceracoder_b200.synth import make_source
    run_rc_parity(lib, make_source("pan", 1280, 720), 1280, 720, 60, 24, lambda i: 4000, gop=10, recon_every=8)
    run_rc_parity(lib, make_source("ball", 1280, 720), 1280, 720, 60, 12, lambda i: 4000, gop=60, recon_every=6)


def test_eight_concurrent_1080p_streams_equal_solo_runs(lib):
    """the bench's shape: 8 encoders, 8 host threads, 8 CUDA streams, full-size pictures, rate control on"""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h, n = 1920, 1080, 8
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(n)]

    def run(out, k):
        e = CudaH264Enc(w, h, fps=(60, 1), lib=lib, bitrate=3000 + 1500 * (k % 4), key_int_max=5)
        out[k] = [e.encode(*f) for f in frames]
        e.close()
    res, solo = {}, {}
    ths = [threading.Thread(target=run, args=(res, k)) for k in range(8)]
    [t.start() for t in ths]; [t.join() for t in ths]
    for k in range(4):
        run(solo, k)
    for k in range(8):
        assert res[k] == solo[k % 4], f"stream {k} differs from its solo run"


def test_pageable_frames_take_the_pinned_staging_copy(lib):
    """numpy-owned (pageable) frames, as a mapped GstBuffer would be, go through the handle's pinned staging frame and give the
    same access units as page-locked frames used in place"""
    import ctypes as C
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h = 1920, 1080
    src = make_source("pan", w, h)
    a = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    b = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    sz = w * h * 3 // 2
    pin = lib.cc_host_alloc(sz)
    buf = np.ctypeslib.as_array(C.cast(pin, C.POINTER(C.c_uint8)), shape=(sz,))
    for i in range(3):
        y, u, v = src.frame(i)
        buf[:w * h] = y.ravel(); buf[w * h:w * h * 5 // 4] = u.ravel(); buf[w * h * 5 // 4:] = v.ravel()
        x = a.encode(y, u, v)
        z = b.encode((pin, w), (pin + w * h, w // 2), (pin + w * h * 5 // 4, w // 2))
        assert x == z
    assert a.staged_frames() == 3 and b.staged_frames() == 0
    a.close(); b.close()
    lib.cc_host_free(pin)
