#!/usr/bin/env python
"""Rate / distortion table of the encoder on S2 1080p content at fixed QPs (run under gpurun): what the slice structure and the
partition search cost or buy, in bits and Y-PSNR — the quality price of the latency design.
Frames: 0..89 and 270..329 of the S2 sequence (the second stretch crosses the scene cut at 300), GOP 60, one stream.
For every configuration and QP: mean bits per frame, Y-PSNR of the reconstruction against the source (the decoder reproduces the
reconstruction bit for bit - parity tests).  Prints JSON lines; tools/bd_rate.py turns them into BD-rate against the default.
usage: python tools/rd_report.py > rd.jsonl"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

QPS = (20, 24, 28, 32, 36, 40)
CONFIGS = [("rps4_part (default)", dict(rows_per_slice=4, partitions=True)),
           ("rps1_part", dict(rows_per_slice=1, partitions=True)),
           ("rps2_part", dict(rows_per_slice=2, partitions=True)),
           ("rps8_part", dict(rows_per_slice=8, partitions=True)),
           ("rps16_part", dict(rows_per_slice=16, partitions=True)),
           ("rps24_part", dict(rows_per_slice=24, partitions=True)),
           ("rps32_part", dict(rows_per_slice=32, partitions=True)),
           ("one_slice_part", dict(rows_per_slice=68, partitions=True)),
           ("rps4_16x16only", dict(rows_per_slice=4, partitions=False))]


def main():
    from ceracoder_b200 import CudaH264Enc, load_library
    store = bench.FrameStore(bench.POOL, 0)
    L = load_library()
    idx = list(range(0, 90)) + list(range(270, 330))
    W, H = bench.W, bench.H
    for name, kw in CONFIGS:
        for qp in QPS:
            enc = CudaH264Enc(W, H, fps=(60, 1), key_int_max=60, qp_fixed=qp, lib=L, **kw)
            bits, psnr, skip, intra = [], [], 0, 0
            for j, i in enumerate(idx):
                y, u, v = store.planes(0, i)
                au = enc.encode(y, u, v, force_idr=(j == 90))  # the second stretch starts a new GOP
                bits.append(8 * len(au)); skip += enc.last.n_skip; intra += enc.last.n_intra
                if j % 3 == 2:
                    ry = enc.recon()[0][:H, :W]
                    mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                    psnr.append(10 * np.log10(255.0 ** 2 / max(mse, 1e-9)))
            enc.close()
            print(json.dumps({"config": name, "qp": qp, "frames": len(idx), "kbps_at_60fps": float(np.mean(bits) * 60 / 1e3), "bits_per_frame": float(np.mean(bits)),
                              "y_psnr_db": float(np.mean(psnr)), "skip_per_frame": skip / len(idx), "intra_per_frame": intra / len(idx)}), flush=True)


if __name__ == "__main__":
    main()
