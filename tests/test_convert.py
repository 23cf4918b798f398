"""Packed input (YUY2 / BGRx) converted inside the ingest kernel — SURVEY.md §8f N2, the `videoconvert` in front of x264enc in
/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:3-5.

Parity chain: oracle/ref_convert.c states the arithmetic; an encoder fed the packed frames must emit exactly the access units of
an encoder fed the oracle-converted I420 planes (which tests/test_*_encoder.py already pin to the oracle encoder), and its
reconstruction must match.  CPU run: kernel sources through the SIMT emulator; `-m gpu`: the real library."""
import ctypes as C
import numpy as np
import pytest
from oracle_lib import lib as oracle_handle


def _convert(fmt, packed, w, h):
    L = oracle_handle()
    y = np.zeros((h, w), np.uint8); u = np.zeros((h // 2, w // 2), np.uint8); v = np.zeros_like(u)
    fn = {"YUY2": L.ref_convert_yuy2, "BGRx": L.ref_convert_bgrx}[fmt]
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    fn.restype = None
    fn(packed.ctypes.data, packed.strides[0], w, h, y.ctypes.data, w, u.ctypes.data, v.ctypes.data, w // 2)
    return y, u, v


def _packed_frames(fmt, w, h, n, seed):
    """moving gradient + noise so that P frames have motion, residuals and some intra blocks"""
    rng = np.random.default_rng(seed)
    bpp = 2 if fmt == "YUY2" else 4
    base = rng.integers(0, 256, (h + 16, (w + 16) * bpp), dtype=np.uint8)
    base = (base // 4 + np.linspace(0, 180, (w + 16) * bpp, dtype=np.float32)[None, :].astype(np.uint8)).astype(np.uint8)
    out = []
    for i in range(n):
        ox = 2 * i * bpp if fmt == "BGRx" else 4 * i  # whole pixels (BGRx) / whole pixel pairs (YUY2)
        f = np.ascontiguousarray(base[i:i + h, ox:ox + w * bpp])
        out.append(f)
    return out


def _run(lib, fmt, w, h, qps, rps=2):
    from ceracoder_b200 import CudaH264Enc
    frames = _packed_frames(fmt, w, h, len(qps), seed=w + h)
    a = CudaH264Enc(w, h, lib=lib, fmt=fmt, rows_per_slice=rps)
    b = CudaH264Enc(w, h, lib=lib, fmt="I420", rows_per_slice=rps)
    for f, qp in zip(frames, qps):
        au_a = a.encode(f, qp=qp)
        y, u, v = _convert(fmt, f, w, h)
        au_b = b.encode(y, u, v, qp=qp)
        assert au_a == au_b, f"{fmt}: access unit differs from the one of the oracle-converted I420 frame"
        ra, rb = a.recon(), b.recon()
        assert all(np.array_equal(p, q) for p, q in zip(ra, rb))
    a.close(); b.close()


def test_oracle_conversion_known_values():
    """limited-range anchors: black, white and the primaries land on the nominal BT.601 / BT.709 code points (within 1)"""
    for h, exp in ((16, {"white": (235, 128, 128), "black": (16, 128, 128), "red": (81, 90, 240), "blue": (41, 240, 110)}),
                   (720, {"white": (235, 128, 128), "black": (16, 128, 128), "red": (63, 102, 240), "blue": (32, 240, 118)})):
        for name, bgr in (("white", (255, 255, 255)), ("black", (0, 0, 0)), ("red", (0, 0, 255)), ("blue", (255, 0, 0))):
            px = np.zeros((h, 16, 4), np.uint8); px[..., 0], px[..., 1], px[..., 2] = bgr
            y, u, v = _convert("BGRx", px.reshape(h, 64), 16, h)
            got = (int(y[0, 0]), int(u[0, 0]), int(v[0, 0]))
            assert all(abs(g - e) <= 1 for g, e in zip(got, exp[name])), (h, name, got, exp[name])
    yuy2 = np.array([[10, 100, 20, 200] * 8, [30, 101, 40, 203] * 8] * 8, np.uint8)
    y, u, v = _convert("YUY2", yuy2, 16, 16)
    assert (y[0, :2].tolist(), y[1, :2].tolist(), int(u[0, 0]), int(v[0, 0])) == ([10, 20], [30, 40], 101, 202)


@pytest.mark.parametrize("fmt,w,h", [("YUY2", 64, 48), ("BGRx", 64, 48), ("BGRx", 40, 24), ("YUY2", 72, 40)])
def test_packed_input_equals_converted_i420_emulator(fmt, w, h):
    from emu_lib import emu_library
    _run(emu_library(), fmt, w, h, [28, 29, 30])


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,w,h", [("YUY2", 176, 144), ("BGRx", 200, 120), ("BGRx", 1280, 720), ("YUY2", 1920, 1080)])
def test_packed_input_equals_converted_i420_gpu(fmt, w, h):
    from ceracoder_b200 import load_library
    _run(load_library(), fmt, w, h, [27, 28, 29, 30], rps=4)


def test_packed_stride_is_validated():
    from emu_lib import emu_library
    from ceracoder_b200 import CudaH264Enc, CeraCudaError
    e = CudaH264Enc(32, 32, lib=emu_library(), fmt="BGRx")
    with pytest.raises(CeraCudaError):
        e.encode((np.zeros((32, 64), np.uint8).ctypes.data, 64))  # 32 pixels of BGRx need 128 bytes per row
    e.close()
