"""cc_prim_satd16x16_noalloc (include/ceracuda.h): the allocation-free variant of cc_prim_satd16x16 — device staging in a fixed
device array inside the library instead of cudaMalloc'ed buffers.  No counterpart in the reference.
Checked through the C ABI against the existing entry point (bit-exact: integer SATD) and against the oracle, on
two calls in a row (the array keeps the previous call's contents).  Collected last on purpose (file name)."""
import ctypes as C
import numpy as np
import pytest
from oracle_lib import lib as oracle

MAX_BLOCKS = 1024  # CC_PRIM_NOALLOC_MAX_BLOCKS


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _lib(kind):
    if kind == "emu":
        from emu_lib import emu_library
        L = emu_library()
    else:
        from ceracoder_b200 import load_library
        L = load_library()
    L.cc_prim_satd16x16_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.cc_prim_satd16x16_noalloc.restype = C.c_int
    return L


def _check_two_calls(L):
    O = oracle()
    rng = np.random.default_rng(77)
    # first call fills the whole scratch, the second is shorter and on other data: stale contents must not show through
    for n in (MAX_BLOCKS, 37):
        a = rng.integers(0, 256, (n, 16, 16), dtype=np.uint8)
        b = rng.integers(0, 256, (n, 16, 16), dtype=np.uint8)
        b[0] = a[0]
        want = np.full(n, -1, np.int32)
        got = np.full(n, -2, np.int32)
        assert L.cc_prim_satd16x16(_p(a), _p(b), n, _p(want)) == 0
        assert L.cc_prim_satd16x16_noalloc(_p(a), _p(b), n, _p(got)) == 0
        assert (got == want).all() and got[0] == 0  # integer work: bit-exact, no tolerance
        for i in range(0, n, 97):
            assert int(got[i]) == O.ref_satd(_p(a[i]), 16, _p(b[i]), 16, 16, 16)


def _check_limit(L):
    n = MAX_BLOCKS + 1
    a = np.zeros((n, 16, 16), np.uint8)
    out = np.zeros(n, np.int32)
    assert L.cc_prim_satd16x16_noalloc(_p(a), _p(a), n, _p(out)) == -1  # CC_ERR_PARAM
    assert L.cc_prim_satd16x16_noalloc(_p(a), _p(a), 0, _p(out)) == -1
    assert L.cc_prim_satd16x16_noalloc(None, _p(a), 4, _p(out)) == -1


def test_noalloc_satd_emu():
    L = _lib("emu")
    _check_two_calls(L)
    _check_limit(L)


@pytest.mark.gpu
def test_noalloc_satd_gpu_matches_existing_entry_point_twice():
    L = _lib("gpu")
    _check_two_calls(L)
    _check_two_calls(L)


@pytest.mark.gpu
def test_noalloc_satd_gpu_rejects_sizes_beyond_the_scratch():
    _check_limit(_lib("gpu"))
