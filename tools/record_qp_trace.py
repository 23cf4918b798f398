#!/usr/bin/env python
"""Record tests/golden/qp_trace_1080p60.csv (run under gpurun): stream 0 of bench.py's workload — its frames, its bitrate
schedule, GOP 60, rate control on — coded by libceracuda alone on the GPU, one line per frame:
    frame, is_idr, qp, au_bytes, crc32(access unit)
`bench.py --impl reference` replays (is_idr, qp) through the CPU oracle on the same frames — equal bitrate by construction, and
byte-identical access units (checked against au_bytes / crc32) since GPU == oracle at equal QP (tests/test_gpu_parity_long.py).
The trace depends on the rate control (cc_rc.c), the schedule and the content: re-record after changing any of them;
tests/test_qp_trace.py (-m gpu) fails when the committed trace is stale.
usage: python tools/record_qp_trace.py [n_frames [out.csv]]"""
import os
import sys
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def record(n, store=None, lib=None):
    from ceracoder_b200 import CudaH264Enc, load_library
    store = store or bench.FrameStore(bench.POOL, 0)
    L = lib or load_library()
    enc = CudaH264Enc(bench.W, bench.H, fps=(bench.FPS, 1), bitrate=12000, key_int_max=60, lib=L)
    rows = []
    for i in range(n):
        enc.set_property("bitrate", bench.schedule_kbps(i))
        au = enc.encode(*store.planes(0, i))
        rows.append((i, int(enc.last.is_idr), int(enc.last.qp), len(au), zlib.crc32(au) & 0xffffffff))
    enc.close()
    return rows


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 900
    out = sys.argv[2] if len(sys.argv) > 2 else bench.TRACE
    rows = record(n)
    with open(out, "w") as f:
        f.write("# stream 0 of bench.py's workload as libceracuda codes it; written by tools/record_qp_trace.py\n")
        f.write("frame,is_idr,qp,au_bytes,crc32\n")
        for r in rows:
            f.write(f"{r[0]},{r[1]},{r[2]},{r[3]},{r[4]:08x}\n")
    kbps = sum(r[3] for r in rows) * 8 / (n / bench.FPS) / 1e3
    print(f"wrote {out}: {n} frames, {sum(r[1] for r in rows)} IDR, QP {min(r[2] for r in rows)}..{max(r[2] for r in rows)}, {kbps:.0f} kbit/s")


if __name__ == "__main__":
    main()
