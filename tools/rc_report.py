#!/usr/bin/env python
"""Rate-control view of every BASELINE config (run under gpurun): one `cudah264enc` stream alone, frame by frame through
cc_enc_encode with the bitrate property written as ceracoder's 20 ms tick would, on the content SURVEY.md 8(d) assigns.  Prints one
JSON object per case (achieved kbit/s, 1 s-window bitrate error mean / p95 with the first second left out, QP range) and, with an
argument, writes per-frame `case,frame,target_bps,bits,qp,is_idr` lines — tools/rc_sim.py replays those complexity traces offline.
usage: python tools/rc_report.py [frames.csv]"""
import csv
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ceracoder_b200 import CudaH264Enc, load_library  # noqa: E402
from ceracoder_b200.synth import make_source  # noqa: E402


def balancer():
    rows = list(csv.reader(open(os.path.join(ROOT, "tests", "golden", "balancer_2_12mbps.csv"))))[1:]
    s = [(int(t) / 1e3, int(b)) for t, b in rows]

    def f(t):
        b = s[0][1]
        for tt, v in s:
            if tt <= t:
                b = v
        return b
    return f


CASES = [("config2_1080p60_balancer", "pan", 1920, 1080, 60, balancer(), 36.0, 40),
         ("config2_1080p60_square", "pan", 1920, 1080, 60, lambda t: 12e6 if int(t / 2) % 2 == 0 else 2e6, 12.0, 40),
         ("config1_1080p30_6mbps_testsrc", "testsrc", 1920, 1080, 30, lambda t: 6e6, 7.0, 0),
         ("config1_1080p30_6mbps_pan", "pan", 1920, 1080, 30, lambda t: 6e6, 7.0, 40),
         ("config3_720p60_4mbps_pan", "pan", 1280, 720, 60, lambda t: 4e6, 7.0, 40),
         ("config4_2160p30_20mbps_pan", "pan", 3840, 2160, 30, lambda t: 20e6, 7.0, 24)]


def main():
    L = load_library()
    dump = open(sys.argv[1], "w") if len(sys.argv) > 1 else None
    grid = [("", {})]
    if os.environ.get("RC_GRID"):  # e.g. RC_GRID="EMA=0.5,STEP=3;EMA=0.25,STEP=3;EMA=0.12,STEP=2": the cases once per setting
        grid = [(g, dict(kv.split("=") for kv in g.split(","))) for g in os.environ["RC_GRID"].split(";")]
    cache = {}
    for tag, env in grid:
      for k_, v_ in env.items():
        os.environ["CERACUDA_RC_" + k_] = v_
      for name, source, w, h, fps, sched, secs, pool in CASES:
        if tag:
            name = name + "[" + tag + "]"
        key = (source, w, h, pool)
        if key not in cache:
            src0 = make_source(source, w, h)
            cache[key] = (src0, [src0.frame(i) for i in range(pool)] if pool else None)
        src, frames = cache[key]
        enc = CudaH264Enc(w, h, fps=(fps, 1), bitrate=int(sched(0)) // 1000, key_int_max=60, lib=L)
        rows = []
        psnr = []
        for i in range(int(secs * fps)):
            bps = int(sched(i / fps))
            enc.set_property("bps", bps // 1000)
            if frames:
                p = i % (2 * pool - 2)
                y, u, v = frames[p if p < pool else 2 * pool - 2 - p]
            else:
                y, u, v = src.frame(i)
            au = enc.encode(y, u, v)
            rows.append((bps, 8 * len(au), int(enc.last.qp), int(enc.last.is_idr)))
            if i % 7 == 6:
                ry = enc.recon()[0][:h, :w]
                psnr.append(10 * np.log10(255.0 ** 2 / max(float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean()), 1e-9)))
            if dump:
                dump.write(f"{name},{i},{bps},{8 * len(au)},{int(enc.last.qp)},{int(enc.last.is_idr)}\n")
        enc.close()
        a = np.array(rows, float)
        tgt = np.convolve(a[:, 0] / fps, np.ones(fps), "valid")[fps:]
        got = np.convolve(a[:, 1], np.ones(fps), "valid")[fps:]
        err = np.abs(got - tgt) / tgt
        print(json.dumps({"case": name, "y_psnr_db": float(np.mean(psnr)), "frames": len(rows), "achieved_kbps": float(a[fps:, 1].sum() / ((len(rows) - fps) / fps) / 1e3),
                          "target_kbps": float((a[fps:, 0] / fps).sum() / ((len(rows) - fps) / fps) / 1e3),
                          "bitrate_error_1s_windows": {"mean": float(err.mean()), "p95": float(np.percentile(err, 95)), "max": float(err.max())},
                          "qp": {"min": int(a[:, 2].min()), "median": float(np.median(a[:, 2])), "max": int(a[:, 2].max())}}), flush=True)


if __name__ == "__main__":
    main()
