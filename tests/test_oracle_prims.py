"""Definition-level known answers for the CPU oracle's primitives.

The reference holds no encoder vectors (parity unpinned at reference level), so the non-normative
primitives are pinned by hand-derivable identities and the normative ones by the decoder
(test_oracle_decode.py)."""
import ctypes as C
import numpy as np
import pytest
from oracle_lib import lib

I16 = C.c_int16 * 16
U16 = C.c_uint16 * 16


def _u8(a):
    a = np.ascontiguousarray(a, np.uint8)
    return a, a.ctypes.data_as(C.c_void_p)


def test_sad_identity_and_value():
    L = lib()
    rng = np.random.default_rng(1)
    a, pa = _u8(rng.integers(0, 256, (16, 16)))
    b, pb = _u8(rng.integers(0, 256, (16, 16)))
    assert L.ref_sad(pa, 16, pa, 16, 16, 16) == 0
    assert L.ref_sad(pa, 16, pb, 16, 16, 16) == int(np.abs(a.astype(int) - b.astype(int)).sum())
    assert L.ref_sad(pa, 16, pb, 16, 8, 8) == int(np.abs(a[:8, :8].astype(int) - b[:8, :8].astype(int)).sum())


@pytest.mark.parametrize("c", [1, -3, 17, 255])
def test_satd_constant_difference(c):
    L = lib()
    a, pa = _u8(np.full((4, 4), max(c, 0)))
    b, pb = _u8(np.full((4, 4), max(-c, 0)))
    assert L.ref_satd_4x4(pa, 4, pb, 4) == 8 * abs(c)


def test_satd_matches_matrix_definition():
    L = lib()
    rng = np.random.default_rng(2)
    Hd = np.array([[1, 1, 1, 1], [1, 1, -1, -1], [1, -1, -1, 1], [1, -1, 1, -1]])
    for _ in range(50):
        a, pa = _u8(rng.integers(0, 256, (4, 4)))
        b, pb = _u8(rng.integers(0, 256, (4, 4)))
        d = a.astype(int) - b.astype(int)
        assert L.ref_satd_4x4(pa, 4, pb, 4) == int(np.abs(Hd @ d @ Hd.T).sum()) >> 1


def test_dct_matrix_definition_and_dc():
    L = lib()
    rng = np.random.default_rng(3)
    Cf = np.array([[1, 1, 1, 1], [2, 1, -1, -2], [1, -1, -1, 1], [1, -2, 2, -1]])
    out = I16()
    a, pa = _u8(np.full((4, 4), 9)); z, pz = _u8(np.zeros((4, 4)))
    L.ref_sub4x4_dct(out, pa, 4, pz, 4)
    assert list(out) == [144] + [0] * 15
    for _ in range(50):
        a, pa = _u8(rng.integers(0, 256, (4, 4))); b, pb = _u8(rng.integers(0, 256, (4, 4)))
        L.ref_sub4x4_dct(out, pa, 4, pb, 4)
        d = a.astype(int) - b.astype(int)
        assert (np.array(out).reshape(4, 4) == Cf @ d @ Cf.T).all()


@pytest.mark.parametrize("qp", [10, 17, 24, 30, 36, 45, 51])
def test_quant_matches_textbook_form(qp):
    """x264's (bias+|c|)*mf>>16 equals level=(|Y|*MF+f)>>(15+qp/6) with f=dz/64 of the step, for qp>=6."""
    L = lib()
    L.ref_quant_tables.argtypes = [C.c_int, C.c_int, U16, U16]
    mf, bias = U16(), U16()
    L.ref_quant_tables(qp, 1, mf, bias)
    rng = np.random.default_rng(qp)
    c = rng.integers(-4000, 4000, 16).astype(np.int16)
    d = I16(*c)
    L.ref_quant_4x4(d, mf, bias)
    MF = np.array([[13107, 5243, 8066], [11916, 4660, 7490], [10082, 4194, 6554], [9362, 3647, 5825], [8192, 3355, 5243], [7282, 2893, 4559]])
    for i in range(16):
        x, y = i & 3, i >> 2
        cls = 1 if (x & 1 and y & 1) else (0 if not (x & 1 or y & 1) else 2)
        step_q = (1 << (15 + qp // 6)) / MF[qp % 6][cls]
        lv = abs(int(c[i])) / step_q
        got = abs(d[i])
        assert abs(got - (lv + 21 / 64)) <= 1.02 and np.sign(d[i]) in (0, np.sign(c[i]))


@pytest.mark.parametrize("qp", [12, 22, 28, 37, 51])
def test_zero_residual_reproduces_pred(qp):
    L = lib()
    rng = np.random.default_rng(5)
    p, pp = _u8(rng.integers(0, 256, (4, 4)))
    dst, pd = _u8(np.zeros((4, 4)))
    d = I16()
    L.ref_sub4x4_dct(d, pp, 4, pp, 4)
    L.ref_dequant_4x4(d, qp)
    L.ref_add4x4_idct(pd, 4, pp, 4, d)
    assert (dst == p).all()


def test_transform_roundtrip_small_qp_close():
    L = lib()
    L.ref_quant_tables.argtypes = [C.c_int, C.c_int, U16, U16]
    rng = np.random.default_rng(6)
    mf, bias = U16(), U16()
    L.ref_quant_tables(12, 1, mf, bias)
    for _ in range(20):
        a, pa = _u8(rng.integers(0, 256, (4, 4))); b, pb = _u8(rng.integers(0, 256, (4, 4)))
        dst, pd = _u8(np.zeros((4, 4)))
        d = I16()
        L.ref_sub4x4_dct(d, pa, 4, pb, 4); L.ref_quant_4x4(d, mf, bias); L.ref_dequant_4x4(d, 12); L.ref_add4x4_idct(pd, 4, pb, 4, d)
        assert np.abs(dst.astype(int) - a.astype(int)).max() <= 2


def test_hadamard_dc_pairs():
    L = lib()
    d = I16(*([7] * 16))
    L.ref_dct4x4dc(d)
    assert d[0] == (7 * 16 + 1) >> 1 and all(v == 0 for v in list(d)[1:])
    e = (C.c_int16 * 4)(5, 5, 5, 5)
    L.ref_dct2x2dc(e)
    assert list(e) == [20, 0, 0, 0]


def test_decimate_score():
    L = lib()
    z = I16(); assert L.ref_decimate_score(z, 16) == 0
    z = I16(1); assert L.ref_decimate_score(z, 16) == 3
    z = I16(2); assert L.ref_decimate_score(z, 16) == 9
    z = I16(0, 0, 0, 1); assert L.ref_decimate_score(z, 16) == 1
    z = I16(1, 0, 1); assert L.ref_decimate_score(z, 16) == 2 + 3


def test_mc_integer_and_halfpel_flat():
    L = lib()
    rng = np.random.default_rng(7)
    ref, pr = _u8(rng.integers(0, 256, (40, 40)))
    dst, pd = _u8(np.zeros((16, 16)))
    base = C.c_void_p(pr.value + 12 * 40 + 12)
    L.ref_mc_luma(pd, 16, base, 40, 8, -4, 16, 16)
    assert (dst == ref[11:27, 14:30]).all()
    flat, pf = _u8(np.full((40, 40), 77))
    for mv in range(16):
        L.ref_mc_luma(pd, 16, C.c_void_p(pf.value + 12 * 40 + 12), 40, mv & 3, mv >> 2, 16, 16)
        assert (dst == 77).all()
    L.ref_mc_chroma(pd, 16, C.c_void_p(pf.value + 12 * 40 + 12), 40, 5, 3, 8, 8)
    assert (dst[:8, :8] == 77).all()


def test_mv_bits_and_median():
    L = lib()
    assert [L.ref_mv_bits(v) for v in (0, 1, -1, 2, -2, 3, 4, -8)] == [1, 3, 3, 5, 5, 5, 7, 9]
    assert L.ref_median3(1, 5, 3) == 3 and L.ref_median3(-4, -4, 9) == -4


def test_cavlc_known_block():
    """H.264 worked example: zigzag 0,3,0,1,-1,-1,0,1,0... with nC=0 -> 0000100 0111 0010 1000 1 01 1 (JVT tutorial)."""
    L = lib()
    out = (C.c_uint8 * 64)()
    bp = C.c_int(0)
    zz = I16(0, 3, 0, 1, -1, -1, 0, 1)
    n = L.ref_cavlc_block(out, C.byref(bp), zz, 16, 0)
    bits = "".join(f"{b:08b}" for b in bytes(out)[:(n + 7) // 8])[:n]
    assert bits == "000010001110010111101101"
