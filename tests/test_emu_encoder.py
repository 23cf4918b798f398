"""CPU-only parity: the product's own sources (host C + kernel algorithms) built against the SIMT emulator must
produce access units byte-identical to the oracle's, and libavcodec must decode them to the same pictures.
(The `-m gpu` twin, tests/test_gpu_encoder.py, runs the same checks through the real libceracuda.so on a B200.)"""
import numpy as np
import pytest
from emu_lib import emu_library
from parity_util import run_parity


@pytest.fixture(scope="module")
def lib():
    return emu_library()


@pytest.mark.parametrize("name,w,h,rps,qps", [
    ("pan", 64, 48, 3, [28, 29, 30]),
    ("pan", 176, 144, 4, [24, 25, 26, 40, 18]),
    ("pan", 200, 120, 2, [34, 35, 36, 51, 12]),      # cropping; qp extremes
    ("testsrc", 160, 96, 6, [28, 28, 30, 27]),       # mostly P_Skip
    ("ball", 128, 80, 1, [22, 23, 24, 25]),          # one row per slice: no top neighbours at all
    ("pan", 96, 64, 8, [30, 31]),                    # slice taller than the picture
])
def test_bitstream_identical_to_oracle(lib, name, w, h, rps, qps):
    run_parity(lib, name, w, h, qps, rps=rps, gop=60)


def test_gop_restart_and_nv12(lib):
    run_parity(lib, "pan", 80, 64, [30, 30, 30, 30, 30], rps=2, gop=2)
    run_parity(lib, "pan", 112, 80, [27, 28, 29], rps=3, fmt="NV12")


def test_partitions_off(lib):
    """cc_enc_params.partitions = 0: the search never chooses P_8x8 (16x16 partitions only)."""
    run_parity(lib, "pan", 176, 144, [26, 27, 28], rps=4, partitions=False)


def test_random_noise_frames(lib):
    rng = np.random.default_rng(5)
    w, h = 64, 64
    frames = [(rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8),
               rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)) for _ in range(4)]
    run_parity(lib, None, w, h, [10, 11, 51, 30], rps=2, frames=frames)


def test_strided_input(lib):
    from ceracoder_b200 import CudaH264Enc
    from oracle_lib import OracleEncoder
    from ceracoder_b200.synth import make_source
    w, h = 80, 48
    y, u, v = make_source("pan", w, h).frame(0)
    big = np.zeros((h, w + 24), np.uint8); big[:, :w] = y
    bu = np.zeros((h // 2, w // 2 + 12), np.uint8); bu[:, :w // 2] = u
    bv = np.zeros((h // 2, w // 2 + 12), np.uint8); bv[:, :w // 2] = v
    enc = CudaH264Enc(w, h, lib=lib)
    a = enc.encode(big[:, :w], bu[:, :w // 2], bv[:, :w // 2], qp=30)
    orc = OracleEncoder(w, h, 30, 4)
    assert a == orc.encode(y, u, v, True, 30)
    enc.close()


def test_encode_sequence_equals_frame_by_frame(lib):
    """cc_enc_encode_sequence (the streaming-thread loop in C) = set_property + encode per frame, under rate control."""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h = 96, 64
    src = make_source("pan", w, h)
    frames = [src.frame(i) for i in range(5)]
    rates = [6000, 6000, 2000, 2000, 9000]
    a = CudaH264Enc(w, h, fps=(60, 1), lib=lib)
    st = a.encode_sequence(frames, rates)
    b = CudaH264Enc(w, h, fps=(60, 1), lib=lib)
    sizes = []
    for (y, u, v), r in zip(frames, rates):
        b.set_property("bitrate", r)
        sizes.append(len(b.encode(y, u, v)))
    assert st.frames == 5 and st.au_bytes == sum(sizes) and st.last_qp == b.last.qp
    assert a.last.size == sizes[-1] and a.get_property("bitrate") == 9000
    (ya, ua, va), (yb, ub, vb) = a.recon(), b.recon()
    assert (ya == yb).all() and (ua == ub).all() and (va == vb).all()
    a.close(); b.close()


@pytest.mark.parametrize("w,h,rps", [(16, 16, 4), (18, 16, 1), (32, 18, 2), (48, 16, 8)])
def test_minimum_and_odd_sizes(lib, w, h, rps):
    """smallest legal pictures (even, >= 16), sizes that are not macroblock multiples, a single macroblock"""
    run_parity(lib, "pan", w, h, [30, 31, 22], rps=rps)


def test_force_idr_flag_and_pts(lib):
    """CC_FORCE_IDR (upstream force-key-unit) restarts the GOP; pts passes through; SPS/PPS precede every IDR."""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    from h264_decode import decode_aus
    w, h = 64, 48
    src = make_source("pan", w, h)
    e = CudaH264Enc(w, h, lib=lib, qp_fixed=30)
    aus = []
    for i in range(5):
        y, u, v = src.frame(i)
        aus.append(e.encode(y, u, v, pts=1000 + i, force_idr=(i == 3)))
        assert e.last.pts == 1000 + i
        assert e.last.is_idr == (i in (0, 3))
        assert (aus[-1][:5] == b"\x00\x00\x00\x01\x67") == bool(e.last.is_idr)
    assert len(decode_aus(aus)) == 5
    e.close()


def test_bitrate_property_is_clamped_to_the_reference_range(lib):
    """/root/reference/src/core/bitrate_control.h:30-31: 300 kbps .. 30 Mbps"""
    from ceracoder_b200 import CudaH264Enc
    e = CudaH264Enc(64, 48, lib=lib)
    e.set_property("bitrate", 1)
    assert e.get_property("bitrate") == 300
    e.set_property("bitrate", 10 ** 6)
    assert e.get_property("bitrate") == 30000
    e.close()


def test_slice_writer_threads_give_the_same_access_units(lib):
    """cc_enc_params.threads (x264enc's `threads`): the slices of an access unit written by a pool of host threads are the
    bytes the calling thread alone writes."""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    w, h = 176, 144
    src = make_source("pan", w, h)
    a = CudaH264Enc(w, h, lib=lib, rows_per_slice=1, threads=0)
    b = CudaH264Enc(w, h, lib=lib, rows_per_slice=1, threads=4)
    c = CudaH264Enc(w, h, lib=lib, rows_per_slice=4, threads=16)  # more threads than slices: clamped
    d = CudaH264Enc(w, h, lib=lib, rows_per_slice=4)
    for i in range(8):
        y, u, v = src.frame(i)
        assert a.encode(y, u, v, qp=26 + i) == b.encode(y, u, v, qp=26 + i)
        assert c.encode(y, u, v, qp=26 + i) == d.encode(y, u, v, qp=26 + i)
    [e.close() for e in (a, b, c, d)]


def test_device_and_host_entropy_coders_write_the_same_bytes(lib):
    """the device CAVLC kernels (default) and the host C writer (cc_enc_params.host_entropy) produce identical access
    units — on texture, on noise at the QP extremes (escape-coded levels, every nC class) and on all-skip pictures"""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    rng = np.random.default_rng(11)
    w, h = 80, 64
    noise = [(rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8),
              rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)) for _ in range(3)]
    still = [noise[0]] * 3                                                      # P frames of nothing but P_Skip: one mb_skip_run per slice
    pan = [make_source("pan", w, h).frame(i) for i in range(4)]
    for frames, qps, rps in ((noise, [10, 10, 51], 2), (still, [30, 30, 30], 4), (pan, [24, 26, 28, 40], 3), (pan, [33, 30, 27, 24], 1)):
        a = CudaH264Enc(w, h, lib=lib, rows_per_slice=rps)
        b = CudaH264Enc(w, h, lib=lib, rows_per_slice=rps, host_entropy=True)
        for f, qp in zip(frames, qps):
            x, y = a.encode(*f, qp=qp), b.encode(*f, qp=qp)
            assert x == y, f"device and host entropy coders differ at qp {qp} ({len(x)} vs {len(y)} bytes)"
            assert a.last.sym_bytes < b.last.sym_bytes or qp == 10       # finished bits cross PCIe, not symbols
        a.close(); b.close()


def test_rate_controlled_stream_replayed_by_the_oracle(lib):
    """the method of tests/test_gpu_parity_long.py at a size the emulator finishes quickly: rate control on, a bitrate step, two
    IDR restarts, a scene cut; the oracle replays the (frame type, QP) trace and every access unit is byte-identical"""
    from ceracoder_b200.synth import PanSprites
    from parity_util import run_rc_parity
    w, h = 176, 144
    aus, trace = run_rc_parity(lib, PanSprites(w, h, cut_every=11), w, h, 30, 26, lambda i: 900 if i < 9 else 300, gop=12, rps=3, recon_every=5)
    assert [i for i, t in enumerate(trace) if t[0]] == [0, 12, 24]
    assert len({t[1] for t in trace}) > 3


@pytest.mark.parametrize("w,h,rps,qps", [
    (64, 320, 16, [28, 30, 26, 34]),      # 20 rows: a two-band slice of 16 rows + a slice of 4
    (96, 400, 24, [30, 27, 33]),          # 25 rows: three bands + one row
    (80, 272, 100, [26, 29, 31, 24]),     # one slice per picture (17 rows: 8 + 8 + 1)
    (48, 144, 9, [30, 28]),               # rps >= picture rows, 9 rows: 8 + 1
])
def test_slices_taller_than_a_band(lib, w, h, rps, qps):
    """rows_per_slice > 8: the slice's wavefront chains 8-row bands across CTAs (frame-tagged progress words, neighbour records and
    unfiltered lines through global memory) in the encode and the filter kernel; bytes and reconstruction stay the oracle's"""
    run_parity(lib, "pan", w, h, qps, rps=rps, gop=60)


@pytest.mark.parametrize("w,h,rps,qps", [(176, 144, 4, [30, 27, 33, 25]), (64, 320, 16, [28, 30, 26])])
def test_filter_after_the_encoder_instead_of_next_to_it(lib, monkeypatch, w, h, rps, qps):
    """CC_OVERLAP=0: k_deblock<false> (launched after k_mb_encode: no progress words of the encoder to follow, pixel loads through
    the L1) — what a handle runs while it is profiled or while more than 8 handles are open; the other tests run the default, the
    filter as it is launched next to the encoder."""
    monkeypatch.setenv("CC_OVERLAP", "0")  # read by cc_enc_open
    run_parity(lib, "pan", w, h, qps, rps=rps, gop=60)


def test_launch_order_follows_the_number_of_open_handles(lib, monkeypatch):
    """cc_enc_submit's default: the in-loop filter is launched next to the macroblock kernel while the device has at most 8 handles
    open in the process (CC_OVERLAP_MAX_HANDLES), after it beyond that; CC_OVERLAP forces either; stage timing turns it off.  The
    bytes do not depend on it."""
    from ceracoder_b200 import CudaH264Enc
    from ceracoder_b200.synth import make_source
    monkeypatch.delenv("CC_OVERLAP", raising=False)
    monkeypatch.setenv("CC_OVERLAP_MAX_HANDLES", "2")
    w, h = 64, 48
    f = [make_source("pan", w, h).frame(i) for i in range(3)]
    a = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    b = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    au = [a.encode(*f[0])]
    assert a.overlapped_frames() == 1          # two handles open: within the limit
    c = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    au.append(a.encode(*f[1]))
    assert a.overlapped_frames() == 1          # three: the filter goes after the encoder
    c.close()
    a.set_profile(True)
    au.append(a.encode(*f[2]))
    assert a.overlapped_frames() == 1          # per-stage events: never next to it
    a.close()
    monkeypatch.setenv("CC_OVERLAP", "0")
    d = CudaH264Enc(w, h, lib=lib, qp_fixed=28)
    ref = [d.encode(*x) for x in f]
    assert d.overlapped_frames() == 0 and ref == au
    d.close(); b.close()
