"""`-m gpu` parity proper: the real libceracuda.so (sm_100a kernels) through the C ABI against the CPU oracle on the
same seeded frames and QPs — access units byte-identical, reconstructions identical, libavcodec decodes to the same
pictures — plus size-independent properties at the full BASELINE sizes."""
import numpy as np
import pytest
from parity_util import run_parity, psnr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from ceracoder_b200 import load_library
    return load_library()


@pytest.mark.parametrize("name,w,h,rps,qps", [
    ("pan", 64, 48, 3, [28, 29, 30]),
    ("pan", 176, 144, 4, [24, 25, 26, 40, 18]),
    ("pan", 200, 120, 2, [34, 35, 36, 51, 12]),
    ("testsrc", 160, 96, 6, [28, 28, 30, 27]),
    ("ball", 128, 80, 1, [22, 23, 24, 25]),
    ("pan", 96, 64, 8, [30, 31]),
    ("pan", 640, 360, 4, [26, 27, 28, 33, 22, 30]),
    ("ball", 1280, 720, 4, [30, 28, 26]),
])
def test_bitstream_identical_to_oracle(lib, name, w, h, rps, qps):
    run_parity(lib, name, w, h, qps, rps=rps, gop=60)


def test_gop_restart_nv12_noise(lib):
    run_parity(lib, "pan", 80, 64, [30, 30, 30, 30, 30], rps=2, gop=2)
    run_parity(lib, "pan", 112, 80, [27, 28, 29], rps=3, fmt="NV12")


def test_partitions_off(lib):
    """cc_enc_params.partitions = 0: the search never chooses P_8x8 (16x16 partitions only)."""
    run_parity(lib, "pan", 176, 144, [26, 27, 28], rps=4, partitions=False)
    rng = np.random.default_rng(5)
    w, h = 64, 64
    frames = [(rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8),
               rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)) for _ in range(4)]
    run_parity(lib, None, w, h, [10, 11, 51, 30], rps=2, frames=frames)


def test_1080p_matches_oracle_two_frames(lib):
    run_parity(lib, "pan", 1920, 1080, [28, 30], rps=4)


@pytest.mark.parametrize("w,h,fps,kbps", [(1920, 1080, 60, 12000), (1280, 720, 60, 4000), (3840, 2160, 30, 20000)])
def test_full_size_stream_decodes_to_recon(lib, w, h, fps, kbps):
    """BASELINE sizes, rate control on: decoder output == encoder reconstruction for every frame, PSNR sane."""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    from h264_decode import decode_aus
    src = make_source("pan", w, h)
    enc = CudaH264Enc(w, h, fps=(fps, 1), bitrate=kbps, lib=lib)
    aus, recs, srcs = [], [], []
    n = 5 if w > 2000 else 8
    for i in range(n):
        y, u, v = src.frame(i)
        aus.append(enc.encode(y, u, v)); recs.append(enc.recon()); srcs.append(y)
    dec = decode_aus(aus)
    assert len(dec) == n
    for (y, u, v), (ry, ru, rv), sy in zip(dec, recs, srcs):
        assert (y == ry[:h, :w]).all() and (u == ru[:h // 2, :w // 2]).all() and (v == rv[:h // 2, :w // 2]).all()
        assert psnr(y, sy) > 28
    enc.close()


def test_concurrent_streams_are_independent(lib):
    """8 encoders on 8 CUDA streams driven from 8 threads give the same bytes as one encoder alone."""
    import threading
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h, n = 320, 192, 6
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(n)]

    def run(out, k):
        e = CudaH264Enc(w, h, lib=lib, qp_fixed=26 + k % 3)
        out[k] = [e.encode(*f) for f in frames]
        e.close()
    res = {}
    ths = [threading.Thread(target=run, args=(res, k)) for k in range(8)]
    [t.start() for t in ths]; [t.join() for t in ths]
    solo = {}
    for k in range(3):
        run(solo, k)
    for k in range(8):
        assert res[k] == solo[k % 3]


def test_submit_collect_and_device_frames(lib):
    import ctypes as C
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h = 256, 144
    y, u, v = make_source("pan", w, h).frame(0)
    a = CudaH264Enc(w, h, lib=lib, qp_fixed=30)
    ref = a.encode(y, u, v)
    b = CudaH264Enc(w, h, lib=lib, qp_fixed=30)
    L = lib
    ptrs = []
    for pl in (y, u, v):
        d = L.cc_dev_alloc(pl.size); assert d
        assert L.cc_dev_upload(d, pl.ctypes.data, pl.size) == 0
        ptrs.append((d, pl.shape[1]))
    b.submit(*ptrs, device_frame=True)
    assert b.collect() == ref
    with pytest.raises(Exception):
        b.collect()
    for d, _ in ptrs:
        L.cc_dev_free(d)
    a.close(); b.close()


@pytest.mark.gpu
def test_slice_writer_threads_give_the_same_access_units(lib):
    """cc_enc_params.threads: 1080p, 17 slices written by 6 host threads = the bytes of the single-threaded writer"""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h = 1920, 1080
    src = make_source("pan", w, h)
    a = CudaH264Enc(w, h, lib=lib, bitrate=12000)
    b = CudaH264Enc(w, h, lib=lib, bitrate=12000, threads=6)
    for i in range(12):
        y, u, v = src.frame(i)
        assert a.encode(y, u, v) == b.encode(y, u, v)
    a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,rps", [(16, 16, 4), (18, 16, 1), (32, 18, 2), (48, 16, 8), (8192 // 4, 16, 4)])
def test_minimum_and_odd_sizes(lib, w, h, rps):
    """smallest legal pictures, sizes that are not macroblock multiples, one very wide macroblock row: the TMA window tiles of the
    coarse search hang over every edge of the padded plane here"""
    run_parity(lib, "pan", w, h, [30, 31, 22], rps=rps)


def test_device_entropy_equals_host_entropy_at_full_size(lib):
    """SURVEY 8f N3 on the B200: 1080p60 with rate control on (QP moves, IDR + P frames, every macroblock type) — the device CAVLC
    kernels and the host C writer produce the same access units; 2160p likewise on a few frames"""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    for w, h, n, kbps in ((1920, 1080, 16, 12000), (3840, 2160, 4, 20000), (1280, 720, 8, 1000)):
        src = make_source("pan", w, h)
        a = CudaH264Enc(w, h, fps=(60, 1), lib=lib, bitrate=kbps, key_int_max=10)
        b = CudaH264Enc(w, h, fps=(60, 1), lib=lib, bitrate=kbps, key_int_max=10, host_entropy=True, threads=4)
        for i in range(n):
            y, u, v = src.frame(i)
            assert a.encode(y, u, v) == b.encode(y, u, v), f"{w}x{h} frame {i}"
            assert a.last.qp == b.last.qp
        a.close(); b.close()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,rps,qps", [
    (64, 320, 16, [28, 30, 26, 34]),
    (96, 400, 24, [30, 27, 33]),
    (80, 272, 100, [26, 29, 31, 24]),
    (640, 368, 16, [27, 30, 33, 24, 29]),
    (1920, 1080, 68, [28, 30, 32]),       # one slice per 1080p picture: nine bands chained across CTAs
    (1920, 1080, 32, [30, 26]),
])
def test_slices_taller_than_a_band(lib, w, h, rps, qps):
    """rows_per_slice > 8 on the B200: bands of a slice run as concurrent CTAs that wait on one another through global memory —
    the case the sequential CPU emulation cannot show.  Bytes, reconstruction (after the cross-band in-loop filter) and decoder
    output are the oracle's."""
    run_parity(lib, "pan", w, h, qps, rps=rps, gop=60)


@pytest.mark.gpu
def test_one_slice_per_picture_under_rate_control_many_frames(lib):
    """60 rate-controlled 1080p frames with one slice per picture, eight streams at once (so the bands' CTAs share the GPU with other
    streams' kernels): identical to a solo run"""
    import threading
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h, n = 1920, 1080, 24
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(8)]

    def run(out, k):
        e = CudaH264Enc(w, h, fps=(60, 1), lib=lib, bitrate=4000 + 2000 * (k % 2), key_int_max=7, rows_per_slice=68)
        out[k] = [e.encode(*frames[i % 8]) for i in range(n)]
        e.close()
    res, solo = {}, {}
    ths = [threading.Thread(target=run, args=(res, k)) for k in range(8)]
    [t.start() for t in ths]; [t.join() for t in ths]
    run(solo, 0); run(solo, 1)
    for k in range(8):
        assert res[k] == solo[k % 2], f"stream {k} differs from its solo run"


@pytest.mark.gpu
@pytest.mark.parametrize("where", ["pinned", "device", "pageable"])
def test_encode_sequence_with_uploads_running_ahead_equals_frame_by_frame(where):
    """cc_enc_encode_sequence uploads frame i + 1 into a second set of source planes while frame i is coded (page-locked host or device
    frames): same access units and reconstruction as one pageable frame at a time, and every frame but the first arrives that way."""
    import ctypes as C
    import zlib
    from ceracoder_b200 import CudaH264Enc, load_library
    from ceracoder_b200.synth import make_source
    L = load_library()
    L.cc_enc_prefetched_frames.restype = C.c_long
    L.cc_enc_prefetched_frames.argtypes = [C.c_void_p]
    L.cc_dev_alloc.restype = C.c_void_p
    L.cc_dev_alloc.argtypes = [C.c_size_t]
    L.cc_dev_free.argtypes = [C.c_void_p]
    L.cc_dev_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    w, h, n = 640, 368, 9
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(n)]
    rates = [3000] * 4 + [800] * 5
    fsz = w * h * 3 // 2
    keep = None
    if where == "pinned":
        base = L.cc_host_alloc(n * fsz)
    elif where == "pageable":  # what gst_video_frame_map hands the element: staged through the second pinned frame
        keep = np.zeros(n * fsz, np.uint8)
        base = keep.ctypes.data
    else:
        base = L.cc_dev_alloc(n * fsz)
    assert base
    planes = []
    for i, (y, u, v) in enumerate(frames):
        tight = np.concatenate([y.ravel(), u.ravel(), v.ravel()])
        p = base + i * fsz
        if where != "device":
            C.memmove(p, tight.ctypes.data, fsz)
        else:
            assert L.cc_dev_upload(p, tight.ctypes.data, fsz) == 0
        planes.append((p, p + w * h, p + w * h * 5 // 4))
    a = CudaH264Enc(w, h, fps=(60, 1), lib=L)
    from ceracoder_b200.encoder import Frame
    arr = (Frame * n)()
    for i, pl in enumerate(planes):
        for j in range(3):
            arr[i].plane[j] = pl[j]
            arr[i].stride[j] = w if j == 0 else w // 2
        arr[i].pts = i
    br = (C.c_int * n)(*[r * 1000 for r in rates])
    seq = (arr, br, n, None)
    st = a.encode_sequence(seq, device_frame=(where == "device"))
    last_crc = zlib.crc32(C.string_at(a.last.data, a.last.size))
    assert a.last.size > 0
    assert L.cc_enc_prefetched_frames(a._h) == n - 1
    b = CudaH264Enc(w, h, fps=(60, 1), lib=L)
    sizes = []
    for (y, u, v), r in zip(frames, rates):
        b.set_property("bitrate", r)
        au = b.encode(y, u, v)
        sizes.append(len(au))
    assert st.frames == n and st.au_bytes == sum(sizes) and zlib.crc32(au) == last_crc
    (ya, ua, va), (yb, ub, vb) = a.recon(), b.recon()
    assert (ya == yb).all() and (ua == ub).all() and (va == vb).all()
    a.close(); b.close()
    if where == "pinned":
        L.cc_host_free(C.c_void_p(base))
    elif where == "device":
        L.cc_dev_free(C.c_void_p(base))


@pytest.mark.gpu
@pytest.mark.parametrize("rps,pct", [(0, 70), (0, 0), (68, 30), (16, 100)])
def test_filter_running_next_to_the_encoder_equals_filter_after_it(lib, monkeypatch, rps, pct):
    """k_deblock launched behind k_mb_encode with programmatic stream serialisation (it follows the encoder's progress words while the
    encoder is still running: CC_OVERLAP=1, released pct % into the rows) against the same kernels one after the other (CC_OVERLAP=0):
    identical access units and reconstruction, 1080p under rate control, four streams at once so that the CTAs really share the GPU;
    default slices, one slice per picture and slices of two bands (the filter then also chains across its own CTAs)."""
    import threading
    import zlib
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h, n = 1920, 1080, 14
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(7)]

    def run(out, k):
        e = CudaH264Enc(w, h, fps=(60, 1), lib=lib, bitrate=3000 + 4000 * (k % 2), key_int_max=9, rows_per_slice=rps)
        aus = [e.encode(*frames[i % 7]) for i in range(n)]
        out[k] = (aus, [zlib.crc32(np.ascontiguousarray(p).tobytes()) for p in e.recon()])
        e.close()

    monkeypatch.setenv("CC_TRAIL_PCT", str(pct))
    monkeypatch.setenv("CC_OVERLAP", "1")  # read by cc_enc_open
    res = {}
    ths = [threading.Thread(target=run, args=(res, k)) for k in range(4)]
    [t.start() for t in ths]; [t.join() for t in ths]
    monkeypatch.setenv("CC_OVERLAP", "0")
    solo = {}
    run(solo, 0); run(solo, 1)
    for k in range(4):
        assert res[k][0] == solo[k % 2][0], f"stream {k}: access units differ"
        assert res[k][1] == solo[k % 2][1], f"stream {k}: reconstruction differs"
