#!/usr/bin/env python
"""Single-stream frame latency at the sizes BASELINE.json's configs name (720p60 @ 4 Mbps, 1080p60 @ 6 Mbps, 2160p30 @ 20 Mbps):
one `cudah264enc` stream alone on the GPU, pinned host frames in, access unit out (cc_enc_encode, what the element's
handle_frame does).  Prints one JSON object per size: mean / p50 / p95 / max milliseconds per frame (host wall clock around the
call), the device share (CUDA events) and the host entropy-coding share, with the access unit written by the calling thread
alone (threads 1) and by four slice-writer threads (cc_enc_params.threads = 4).  Not a bench.py line: bench.py reports the 8-stream
throughput workload; this is the latency view of the same path.
usage (under gpurun): python tools/latency_probe.py [frames]"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ceracoder_b200 import CudaH264Enc, load_library  # noqa: E402
from ceracoder_b200.synth import make_source  # noqa: E402

CASES = [("720p60", 1280, 720, 60, 4000), ("1080p60", 1920, 1080, 60, 6000), ("2160p30", 3840, 2160, 30, 20000)]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 240
    L = load_library()
    for name, w, h, fps, kbps, threads in [c + (t,) for c in CASES for t in (1, 4)]:
        src = make_source("pan", w, h)
        pool = 12
        fsz = w * h * 3 // 2
        base = L.cc_host_alloc(fsz * pool)
        arr = np.ctypeslib.as_array(C.cast(base, C.POINTER(C.c_uint8)), shape=(pool, fsz))
        for i in range(pool):
            y, u, v = src.frame(i)
            arr[i, :w * h] = y.ravel(); arr[i, w * h:w * h * 5 // 4] = u.ravel(); arr[i, w * h * 5 // 4:] = v.ravel()
        e = CudaH264Enc(w, h, fps=(fps, 1), bitrate=kbps, key_int_max=60, lib=L, threads=threads)
        ms, gpu, host, idr = [], [], [], []
        for i in range(n + 30):
            j = i % (2 * pool - 2); j = j if j < pool else 2 * pool - 2 - j
            p = base + j * fsz
            t0 = time.perf_counter()
            e.encode((p, w), (p + w * h, w // 2), (p + w * h * 5 // 4, w // 2))
            dt = 1e3 * (time.perf_counter() - t0)
            if i >= 30:  # warm-up: first IDR, rate control settling
                ms.append(dt); gpu.append(e.last.gpu_ms); host.append(e.last.host_ms); idr.append(e.last.is_idr)
        e.close()
        L.cc_host_free(base)
        a = np.array(ms); p_only = a[~np.array(idr, bool)]
        print(json.dumps({"case": name, "threads": threads, "target_kbps": kbps, "frames": n, "ms_per_frame": {"mean": float(a.mean()), "p50": float(np.median(a)),
                          "p95": float(np.percentile(a, 95)), "max": float(a.max()), "p_frames_mean": float(p_only.mean())},
                          "device_ms_mean": float(np.mean(gpu)), "host_cavlc_ms_mean": float(np.mean(host)),
                          "frame_period_ms": 1e3 / fps, "idr_frames": int(np.sum(idr))}))


if __name__ == "__main__":
    main()
