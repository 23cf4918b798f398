#!/usr/bin/env python
"""What a change of a decision rule costs (CPU only): the oracle — the product bit for bit — with other constants in the P-frame
intra gate / Intra4x4 start rule, on S2 1080p at fixed QPs.  JSON lines for tools/bd_rate.py, plus the analysis counts.
usage: python tools/rule_eval.py [frames] > rules.jsonl"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from ceracoder_b200.synth import PanSprites  # noqa: E402
from oracle_lib import OracleEncoder, lib, MbInfo  # noqa: E402

W, H = 1920, 1080
QPS = (22, 26, 30, 34, 38)
RULES = {"gate 9/8, i4 5/4 (flat 25 %)": (9, 8, 5, 4, 0), "i4 by QP: 10/8 below 28, 9/8 below 34, 8/8 above (product)": (9, 8, 0, 0, 0), "i4 9/8": (9, 8, 9, 8, 0), "i4 1/1": (9, 8, 1, 1, 0),
         "gate 17/16": (17, 16, 5, 4, 0), "gate 1/1": (1, 1, 5, 4, 0), "product + no intra analysis when the inter coding has cbp 0": (9, 8, 0, 0, 1)}


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    src = PanSprites(W, H, cut_every=3 * n // 4)
    frames = [src.frame(i) for i in range(n)]
    L = lib()
    L.ref_set_threads(os.cpu_count())
    for name, r in RULES.items():
        L.ref_x_set_rules(*r[:4]); L.ref_x_set_nocbp(r[4])
        for qp in QPS:
            enc = OracleEncoder(W, H, 60, 8)
            bits, psnr, n_i16 = [], [], 0
            n_gate = 0
            for i, (y, u, v) in enumerate(frames):
                au = enc.encode(y, u, v, i == 0, qp)
                bits.append(8 * len(au))
                ry = enc.recon()[0][:H, :W]
                mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                psnr.append(10 * np.log10(255.0 ** 2 / max(mse, 1e-9)))
                if i:
                    raw = enc.mb_info()
                    o = MbInfo.dbg_cost_i16.offset
                    n_gate += int((raw[:, o:o + 4].copy().view(np.int32)[:, 0] < (1 << 30)).sum())
            enc.close()
            print(json.dumps({"config": name, "qp": qp, "frames": n, "kbps_at_60fps": float(np.mean(bits) * 60 / 1e3), "y_psnr_db": float(np.mean(psnr)),
                              "intra_analyses_per_p_frame": n_gate / (n - 1)}), flush=True)
    L.ref_x_set_rules(9, 8, 0, 0); L.ref_x_set_nocbp(0)


if __name__ == "__main__":
    main()
