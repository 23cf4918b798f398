"""Bit-exact parity of the batch primitives (include/ceracuda.h cc_prim_*) with the oracle on identical seeded
inputs — the stand-ins for libx264's pixel_sad / pixel_satd / sub4x4_dct / quant_4x4 / dequant_4x4 / add4x4_idct
C functions named by north_star [EXT].  Runs on the CPU emulation build everywhere and on the real library on a GPU."""
import ctypes as C
import numpy as np
import pytest
from oracle_lib import lib as oracle

U16 = C.c_uint16 * 16
I16 = C.c_int16 * 16


def _libs():
    out = [pytest.param("emu", id="emu")]
    out.append(pytest.param("gpu", id="gpu", marks=pytest.mark.gpu))
    return out


@pytest.fixture(params=_libs())
def L(request):
    if request.param == "emu":
        from emu_lib import emu_library
        return emu_library()
    from ceracoder_b200 import load_library
    return load_library()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("w,h", [(16, 16), (8, 8), (16, 8)])
def test_sad(L, w, h):
    O = oracle()
    rng = np.random.default_rng(w * h)
    n = 300
    a = rng.integers(0, 256, (n, h, w), dtype=np.uint8); b = rng.integers(0, 256, (n, h, w), dtype=np.uint8)
    b[0] = a[0]; a[1] = 0; b[1] = 255
    out = np.zeros(n, np.int32)
    assert L.cc_prim_sad(_p(a), _p(b), w, h, n, _p(out)) == 0
    ref = [O.ref_sad(_p(a[i]), w, _p(b[i]), w, w, h) for i in range(n)]
    assert out.tolist() == ref and out[0] == 0 and out[1] == 255 * w * h


def test_satd(L):
    O = oracle()
    rng = np.random.default_rng(9)
    n = 500
    a = rng.integers(0, 256, (n, 4, 4), dtype=np.uint8); b = rng.integers(0, 256, (n, 4, 4), dtype=np.uint8)
    a[0] = 200; b[0] = 197
    out = np.zeros(n, np.int32)
    assert L.cc_prim_satd4x4(_p(a), _p(b), n, _p(out)) == 0
    assert out.tolist() == [O.ref_satd_4x4(_p(a[i]), 4, _p(b[i]), 4) for i in range(n)] and out[0] == 24
    n = 64
    a = rng.integers(0, 256, (n, 16, 16), dtype=np.uint8); b = rng.integers(0, 256, (n, 16, 16), dtype=np.uint8)
    out = np.zeros(n, np.int32)
    assert L.cc_prim_satd16x16(_p(a), _p(b), n, _p(out)) == 0
    assert out.tolist() == [O.ref_satd(_p(a[i]), 16, _p(b[i]), 16, 16, 16) for i in range(n)]


def test_dct(L):
    O = oracle()
    rng = np.random.default_rng(10)
    n = 400
    a = rng.integers(0, 256, (n, 4, 4), dtype=np.uint8); b = rng.integers(0, 256, (n, 4, 4), dtype=np.uint8)
    a[0] = 255; b[0] = 0
    out = np.zeros((n, 16), np.int16)
    assert L.cc_prim_dct4x4(_p(a), _p(b), n, _p(out)) == 0
    d = I16()
    for i in range(n):
        O.ref_sub4x4_dct(d, _p(a[i]), 4, _p(b[i]), 4)
        assert out[i].tolist() == list(d)
    assert out[0][0] == 255 * 16


@pytest.mark.parametrize("intra", [0, 1])
def test_transform_quant_recon_all_qp(L, intra):
    O = oracle()
    O.ref_quant_tables.argtypes = [C.c_int, C.c_int, U16, U16]
    rng = np.random.default_rng(11 + intra)
    n = 96
    for qp in range(0, 52):
        a = rng.integers(0, 256, (n, 4, 4), dtype=np.uint8)
        b = np.clip(a.astype(int) + rng.integers(-40, 41, (n, 4, 4)), 0, 255).astype(np.uint8) if qp % 2 else rng.integers(0, 256, (n, 4, 4), dtype=np.uint8)
        lv = np.zeros((n, 16), np.int16); rec = np.zeros((n, 4, 4), np.uint8)
        assert L.cc_prim_tq4x4(_p(a), _p(b), n, qp, intra, _p(lv), _p(rec)) == 0
        mf, bias = U16(), U16()
        O.ref_quant_tables(qp, intra, mf, bias)
        d = I16(); r = np.zeros((4, 4), np.uint8)
        for i in range(n):
            O.ref_sub4x4_dct(d, _p(a[i]), 4, _p(b[i]), 4)
            O.ref_quant_4x4(d, mf, bias)
            assert lv[i].tolist() == list(d), (qp, i)
            O.ref_dequant_4x4(d, qp)
            O.ref_add4x4_idct(_p(r), 4, _p(b[i]), 4, d)
            assert (rec[i] == r).all(), (qp, i)


def test_pred4x4(L):
    O = oracle()
    rng = np.random.default_rng(12)
    n = 200
    edge = rng.integers(0, 256, (n, 13), dtype=np.uint8)
    for mode in range(9):
        out = np.zeros((n, 16), np.uint8)
        assert L.cc_prim_pred4x4(_p(edge), mode, n, _p(out)) == 0
        o = (C.c_uint8 * 16)()
        for i in range(n):
            top = np.ascontiguousarray(edge[i, 4:13]); left = np.ascontiguousarray(edge[i, 3::-1])
            O.ref_pred4x4(o, mode, C.c_void_p(top.ctypes.data + 1), _p(left), 1, 1)
            assert out[i].tolist() == list(o), (mode, i)


def test_mc_luma_all_fractions(L):
    O = oracle()
    rng = np.random.default_rng(13)
    w, h = 64, 48
    plane = rng.integers(0, 256, (h, w), dtype=np.uint8)
    pad = 48
    ext = np.pad(plane, pad, mode="edge")
    xy, mv = [], []
    for f in range(16):
        for (x, y, ix, iy) in ((0, 0, -9, -7), (16, 16, 3, 2), (48, 32, 8, 6), (32, 0, -1, 0), (0, 32, 0, 5)):
            xy.append((x, y)); mv.append((4 * ix + (f & 3), 4 * iy + (f >> 2)))
    xy = np.array(xy, np.int16); mv = np.array(mv, np.int16)
    n = len(xy)
    out = np.zeros((n, 16, 16), np.uint8)
    assert L.cc_prim_mc_luma(_p(plane), w, h, _p(xy), _p(mv), n, _p(out)) == 0
    r = np.zeros((16, 16), np.uint8)
    s = ext.strides[0]
    for i in range(n):
        base = ext.ctypes.data + (pad + int(xy[i, 1])) * s + pad + int(xy[i, 0])
        O.ref_mc_luma(_p(r), 16, C.c_void_p(base), s, int(mv[i, 0]), int(mv[i, 1]), 16, 16)
        assert (out[i] == r).all(), (i, mv[i])


@pytest.mark.parametrize("qp", [10, 23, 30, 36, 44, 51])
def test_transform_chain_8x8(L, qp):
    """sub8x8_dct8 -> quant_8x8 -> dequant_8x8 -> add8x8_idct8 and the sa8d transform-size metric, per macroblock."""
    O = oracle()
    rng = np.random.default_rng(800 + qp)
    n = 64
    cur = rng.integers(0, 256, (n, 16, 16), dtype=np.uint8)
    pred = np.clip(cur.astype(np.int32) + rng.integers(-40, 41, (n, 16, 16)) * (rng.random((n, 1, 1)) < 0.8), 0, 255).astype(np.uint8)
    pred[0] = cur[0]; cur[1] = 255; pred[1] = 0            # zero residual; extreme residual
    lev = np.zeros((n, 4, 64), np.int16); rec = np.zeros((n, 16, 16), np.uint8); sa = np.zeros(n, np.int32)
    assert L.cc_prim_tq8x8(_p(cur), _p(pred), n, qp, _p(lev), _p(rec), _p(sa)) == 0
    for i in range(n):
        ref_rec = np.zeros((16, 16), np.uint8); s = 0
        for b in range(4):
            oy, ox = (b >> 1) * 8, (b & 1) * 8
            c = np.ascontiguousarray(cur[i, oy:oy + 8, ox:ox + 8]); p = np.ascontiguousarray(pred[i, oy:oy + 8, ox:ox + 8])
            d = np.zeros(64, np.int16)
            O.ref_sub8x8_dct8(_p(d), _p(c), 8, _p(p), 8)
            O.ref_quant_8x8(_p(d), qp, 0)
            assert (d == lev[i, b]).all(), f"levels differ: mb {i} block {b} qp {qp}"
            O.ref_dequant_8x8(_p(d), qp)
            o = np.zeros((8, 8), np.uint8)
            O.ref_add8x8_idct8(_p(o), 8, _p(p), 8, _p(d))
            ref_rec[oy:oy + 8, ox:ox + 8] = o
            s += O.ref_sa8d_8x8_raw(_p(c), 8, _p(p), 8)
        assert (ref_rec == rec[i]).all(), f"reconstruction differs: mb {i} qp {qp}"
        assert sa[i] == (s + 2) >> 2
    assert (lev[0] == 0).all() and (rec[0] == cur[0]).all() and sa[0] == 0   # zero residual reproduces the prediction


def test_oracle_8x8_known_answers():
    """Definition-level pins of the 8x8 oracle primitives (the decoder pins dequant + inverse transform separately)."""
    O = oracle()
    c = np.full((8, 8), 130, np.uint8); p = np.full((8, 8), 120, np.uint8)
    d = np.zeros(64, np.int16)
    O.ref_sub8x8_dct8(_p(d), _p(c), 8, _p(p), 8)
    assert d[0] == 64 * 10 and (d[1:] == 0).all()                      # constant residual -> DC only, gain 64
    assert O.ref_sa8d_8x8_raw(_p(c), 8, _p(p), 8) == 64 * 10           # Hadamard DC = 64 * c
    z = np.zeros(64, np.int16); z[0] = 1
    assert O.ref_decimate_score64(_p(z)) == 3
    z[0] = 0; z[20] = 1
    assert O.ref_decimate_score64(_p(z)) == 1                          # run of 20 zeros before it
    z[21] = -2
    assert O.ref_decimate_score64(_p(z)) == 9                          # any |level| > 1
    # round trip at a fine quantiser stays close to the source
    rng = np.random.default_rng(3)
    c = rng.integers(0, 256, (8, 8), dtype=np.uint8); p = rng.integers(0, 256, (8, 8), dtype=np.uint8)
    O.ref_sub8x8_dct8(_p(d), _p(c), 8, _p(p), 8); O.ref_quant_8x8(_p(d), 10, 0); O.ref_dequant_8x8(_p(d), 10)
    o = np.zeros((8, 8), np.uint8); O.ref_add8x8_idct8(_p(o), 8, _p(p), 8, _p(d))
    assert np.abs(o.astype(int) - c.astype(int)).max() <= 3
