#!/usr/bin/env python
"""Bjontegaard delta rate between RD curves (piecewise-cubic fit of log-rate over PSNR, integrated over the common PSNR range).
Input: JSON lines with config / kbps_at_60fps / y_psnr_db (tools/rd_report.py, tools/rd_oracle.py).
usage: python tools/bd_rate.py rd.jsonl [anchor-config-name]"""
import json
import sys

import numpy as np


def bd_rate(anchor, test):
    """average bitrate difference (%) of `test` against `anchor` at equal PSNR; curves = [(kbps, psnr)]"""
    ra, pa = np.log(np.array([r for r, _ in anchor])), np.array([p for _, p in anchor])
    rt, pt = np.log(np.array([r for r, _ in test])), np.array([p for _, p in test])
    lo, hi = max(pa.min(), pt.min()), min(pa.max(), pt.max())
    if hi <= lo:
        return float("nan")
    fa, ft = np.polyfit(pa, ra, 3), np.polyfit(pt, rt, 3)
    ia, it = np.polyint(fa), np.polyint(ft)
    avg = ((np.polyval(it, hi) - np.polyval(it, lo)) - (np.polyval(ia, hi) - np.polyval(ia, lo))) / (hi - lo)
    return float((np.exp(avg) - 1) * 100)


def main():
    rows = [json.loads(l) for l in open(sys.argv[1]) if l.strip().startswith("{")]
    curves = {}
    for r in rows:
        curves.setdefault(r["config"], []).append((r["kbps_at_60fps"], r["y_psnr_db"]))
    anchor = sys.argv[2] if len(sys.argv) > 2 else next(iter(curves))
    print(f"BD-rate against '{anchor}' (negative = fewer bits at equal Y-PSNR)")
    for name, c in curves.items():
        c.sort(key=lambda t: t[1])
        print(f"  {name:28s} {bd_rate(sorted(curves[anchor], key=lambda t: t[1]), c):+7.2f} %   " + "  ".join(f"{k:.0f}k/{p:.2f}dB" for k, p in c))


if __name__ == "__main__":
    main()
