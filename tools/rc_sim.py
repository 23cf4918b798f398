#!/usr/bin/env python
"""Offline check of cc_rc.c against a synthetic encoder model (no GPU): bits = C(frame) / qscale(qp)^1.1 with noise, key frames 8x a
P frame, a scene cut (one P frame at 6x) every 300 frames.  Drives the real cc_rc_* functions through ctypes and prints the
1 s-window bitrate error (mean / p95) for the BASELINE configs.  A design aid: the numbers that count come from the encoder on
the B200 (tests/test_rate_control.py, tools/config_report.py).
usage: python tools/rc_sim.py [path-to-.so-with-cc_rc_*] [old]"""
import ctypes as C
import csv
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def qscale(qp):
    return 0.85 * 2.0 ** ((qp - 12.0) / 6.0)


def run(L, old, fps, gop, nmb, sched, seconds, c_p, seed=1):
    rc = C.create_string_buffer(512)
    if old:
        L.cc_rc_init(rc, C.c_double(fps), int(sched(0)), 10, 51)
    else:
        L.cc_rc_init(rc, C.c_double(fps), int(sched(0)), 10, 51, gop)
    rng = np.random.default_rng(seed)
    out = []
    for i in range(int(seconds * fps)):
        bps = int(sched(i / fps))
        idr = i % gop == 0
        qp = L.cc_rc_frame_qp(rc, bps, int(idr), nmb)
        c = c_p * (8.0 if idr else (6.0 if i % 300 == 0 else 1.0)) * float(np.exp(rng.normal(0, 0.08))) * (1 + 0.3 * np.sin(i / 90.0))
        bits = int(c / qscale(qp) ** 1.1)
        L.cc_rc_update(rc, int(idr), qp, bits)
        out.append((bps, bits, qp))
    return np.array(out, float)


def errors(tb, fps):
    w = int(round(fps))
    tgt = np.convolve(tb[:, 0] / fps, np.ones(w), "valid")[w:]
    got = np.convolve(tb[:, 1], np.ones(w), "valid")[w:]
    e = np.abs(got - tgt) / tgt
    return e.mean(), np.percentile(e, 95)


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ceracoder_b200", "libceracuda.so")
    old = len(sys.argv) > 2
    L = C.CDLL(path)
    L.cc_rc_frame_qp.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    L.cc_rc_update.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_long]
    rows = list(csv.reader(open(os.path.join(ROOT, "tests", "golden", "balancer_2_12mbps.csv"))))[1:]
    bal = [(int(t) / 1e3, int(b)) for t, b in rows]

    def balancer(t):
        b = bal[0][1]
        for tt, v in bal:
            if tt <= t:
                b = v
        return b
    cases = [("1080p60 balancer 12->2->12", 60, 60, 8160, balancer, 36, 200000 * qscale(24) ** 1.1),
             ("1080p60 square 12<->2", 60, 60, 8160, lambda t: 12e6 if int(t / 2) % 2 == 0 else 2e6, 12, 200000 * qscale(24) ** 1.1),
             ("1080p30 @6", 30, 60, 8160, lambda t: 6e6, 12, 200000 * qscale(28) ** 1.1),
             ("720p60 @4", 60, 60, 3600, lambda t: 4e6, 12, 66000 * qscale(26) ** 1.1),
             ("2160p30 @20", 30, 60, 32400, lambda t: 20e6, 12, 666000 * qscale(24) ** 1.1)]
    for name, fps, gop, nmb, sched, secs, c_p in cases:
        tb = run(L, old, fps, gop, nmb, sched, secs, c_p)
        m, p = errors(tb, fps)
        print(f"{name:30s} mean {100 * m:5.2f} %  p95 {100 * p:5.2f} %  achieved/target {tb[:, 1].sum() / (tb[:, 0] / fps).sum():.3f}  qp {tb[:, 2].min():.0f}..{tb[:, 2].max():.0f}")


if __name__ == "__main__":
    main()
