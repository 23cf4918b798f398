#!/usr/bin/env python
"""Rate / quality view of the single-stream configs BASELINE.json names, each on the synthetic source SURVEY.md 8(d) assigns to it:
  config 1  1080p30 @  6 Mbps  S1 testsrc (static bars + noise patch: almost every macroblock P_Skip)
  config 3   720p60 @  4 Mbps  S3 ball
  config 4  2160p30 @ 20 Mbps  S2 pan-sprites
One `cudah264enc` stream alone on the GPU at a constant `bitrate`, frame by frame through cc_enc_encode (what the element's
handle_frame does).  Prints one JSON object per config: achieved kbit/s, bitrate error over 1 s windows (mean / p95, first second
left out of the error: the rate control starts from its initial QP), Y-PSNR of the reconstruction against the source (the decoder
reproduces the reconstruction bit for bit - tests/test_gpu_encoder.py), host milliseconds per frame.  Not a bench.py line
(bench.py reports the 8-stream throughput workload and the 2 <-> 12 Mbps step response of config 2).
With a second argument, per-frame `frame,bits,qp,is_idr` lines of every config are written to that file (rate-control analysis).
usage (under gpurun): python tools/config_report.py [seconds [frames.csv]]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ceracoder_b200 import CudaH264Enc, load_library  # noqa: E402
from ceracoder_b200.synth import make_source  # noqa: E402

CASES = [("config1_1080p30_6mbps", "testsrc", 1920, 1080, 30, 6000),
         ("config3_720p60_4mbps", "ball", 1280, 720, 60, 4000),
         ("config4_2160p30_20mbps", "pan", 3840, 2160, 30, 20000)]


def main():
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 6.0
    L = load_library()
    dump = open(sys.argv[2], "w") if len(sys.argv) > 2 else None
    for name, source, w, h, fps, kbps in CASES:
        src = make_source(source, w, h)
        n = int(secs * fps)
        pool = [src.frame(i) for i in range(24)] if source == "pan" else None  # 4K pan frames cost 1.3 s each to draw: 24 played ping-pong
        enc = CudaH264Enc(w, h, fps=(fps, 1), bitrate=kbps, key_int_max=60, lib=L)
        bits, psnr, ms, qps = [], [], [], []
        for i in range(n):
            if pool:
                j = i % (2 * len(pool) - 2)
                y, u, v = pool[j if j < len(pool) else 2 * len(pool) - 2 - j]
            else:
                y, u, v = src.frame(i)
            t0 = time.perf_counter()
            au = enc.encode(y, u, v, pts=i)
            ms.append(1e3 * (time.perf_counter() - t0))
            bits.append(8 * len(au)); qps.append(int(enc.last.qp))
            if dump:
                dump.write(f"{name},{i},{8 * len(au)},{int(enc.last.qp)},{int(enc.last.is_idr)}\n")
            if i % 6 == 5:
                ry = enc.recon()[0][:h, :w]
                mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                psnr.append(10 * np.log10(255.0 ** 2 / max(mse, 1e-9)))
        enc.close()
        b = np.array(bits, float)
        tgt = kbps * 1000.0
        err = [abs(b[i:i + fps].sum() - tgt) / tgt for i in range(fps, n - fps + 1, max(fps // 4, 1))]
        print(json.dumps({"case": name, "source": source, "frames": n, "target_kbps": kbps,
                          "achieved_kbps": float(b.sum() / (n / fps) / 1e3),
                          "achieved_kbps_after_first_second": float(b[fps:].sum() / ((n - fps) / fps) / 1e3),
                          "bitrate_error_1s_windows": {"mean": float(np.mean(err)), "p95": float(np.percentile(err, 95))},
                          "y_psnr_db": float(np.mean(psnr)), "qp": {"min": min(qps), "median": float(np.median(qps)), "max": max(qps)},
                          "ms_per_frame": {"mean": float(np.mean(ms)), "p95": float(np.percentile(ms, 95))}}), flush=True)


if __name__ == "__main__":
    main()
