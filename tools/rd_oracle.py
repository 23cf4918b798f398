#!/usr/bin/env python
"""How far is the product from the best its own toolset can do?  (CPU only, no GPU needed.)
x264 cannot be had on this image, so the Y-PSNR gap to x264 veryfast stays unmeasured; this is the proxy: the CPU oracle — which
is the product bit for bit in its normal mode (parity tests) — against its own upper-bound mode (ref_enc_set_effort: exhaustive
+-24 full-pel search, 8x8 partitions always refined, per macroblock the candidate with the smallest SSD + lambda * real CAVLC bits)
and against one slice per picture, on S2 1080p at fixed QPs.  Prints JSON lines for tools/bd_rate.py.
  product        rows_per_slice 8, fast decisions      (= libceracuda)
  one_slice      one slice per picture, fast decisions (what the slice structure costs)
  rd             rows_per_slice 8, upper-bound decisions
  rd_one_slice   one slice, upper-bound decisions      (the bound)
The upper-bound streams are checked with libavcodec (decoder output == reconstruction) before their numbers are used.
usage: python tools/rd_oracle.py [frames] > rd_oracle.jsonl"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from ceracoder_b200.synth import PanSprites  # noqa: E402
from oracle_lib import OracleEncoder, lib  # noqa: E402
from h264_decode import decode_aus  # noqa: E402

W, H = 1920, 1080
QPS = (22, 26, 30, 34, 38)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    src = PanSprites(W, H, cut_every=2 * n // 3)  # a scene cut two thirds in
    frames = [src.frame(i) for i in range(n)]
    L = lib()
    import ctypes as C
    L.ref_enc_set_effort.argtypes = [C.c_void_p, C.c_int]
    L.ref_set_threads(os.cpu_count())
    for name, rps, effort in (("product", 8, 0), ("one_slice", 68, 0), ("rd", 8, 1), ("rd_one_slice", 68, 1)):
        for qp in QPS:
            enc = OracleEncoder(W, H, 60, rps)
            L.ref_enc_set_effort(enc.h_, effort)
            bits, psnr, aus, last = [], [], [], None
            for i, (y, u, v) in enumerate(frames):
                au = enc.encode(y, u, v, i == 0, qp)
                bits.append(8 * len(au)); aus.append(au)
                ry = enc.recon()[0][:H, :W]
                last = ry
                mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                psnr.append(10 * np.log10(255.0 ** 2 / max(mse, 1e-9)))
            ok = None
            if effort and qp == QPS[2]:  # the upper-bound encoder is not covered by the parity tests: check its stream here
                dec = decode_aus(aus)
                ok = bool(len(dec) == n and (dec[-1][0] == last).all())
            enc.close()
            print(json.dumps({"config": name, "qp": qp, "frames": n, "kbps_at_60fps": float(np.mean(bits) * 60 / 1e3), "y_psnr_db": float(np.mean(psnr)),
                              "decoder_equals_recon": ok}), flush=True)


if __name__ == "__main__":
    main()
