"""Host-side checks of generated tables against the oracle (no GPU)."""
import ctypes as C
import os
import sys
import numpy as np
from oracle_lib import lib, ROOT

sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_i4_tap_table_matches_oracle_predictors():
    import gen_i4_table
    tab = gen_i4_table.table()
    L = lib()
    rng = np.random.default_rng(3)
    for _ in range(40):
        top = rng.integers(0, 256, 9).astype(np.uint8)   # [0] = top-left, [1..8] = top[0..7]
        left = rng.integers(0, 256, 4).astype(np.uint8)
        for ht, hl in ((1, 1), (1, 0), (0, 1), (0, 0)):
            st, sl = int(top[1:5].sum()), int(left.sum())
            dc = (st + sl + 4) >> 3 if ht and hl else ((st + 2) >> 2 if ht else ((sl + 2) >> 2 if hl else 128))
            E = [int(left[3]), int(left[2]), int(left[1]), int(left[0]), int(top[0])] + [int(v) for v in top[1:]] + [dc]
            for mode in range(9):
                ok = mode == 2 or (mode in (0, 3, 7) and ht) or (mode in (1, 8) and hl) or (mode in (4, 5, 6) and ht and hl)
                if not ok:
                    continue
                out = (C.c_uint8 * 16)()
                L.ref_pred4x4(out, mode, C.c_void_p(top.ctypes.data + 1), C.c_void_p(left.ctypes.data), ht, hl)
                mine = [(E[t & 15] + 2 * E[(t >> 4) & 15] + E[(t >> 8) & 15] + 2) >> 2 for t in tab[mode]]
                assert mine == list(out), (mode, ht, hl)
                S = gen_i4_table.build_S(E)   # the one-lookup-per-pixel form used by the wavefront kernel
                st = gen_i4_table.sidx_table()[mode]
                assert [S[(st[i >> 2] >> (8 * (i & 3))) & 255] for i in range(16)] == list(out), (mode, ht, hl)


def test_generated_headers_are_current():
    import gen_i4_table
    txt = open(os.path.join(ROOT, "ceracoder_b200", "csrc", "cc_i4_table.h")).read()
    for r in gen_i4_table.table():
        assert ", ".join("0x%03x" % v for v in r) in txt
    for r in gen_i4_table.sidx_table():
        assert ", ".join("0x%08xu" % v for v in r) in txt
