import sys, time, threading, numpy as np
sys.path.insert(0, '/root/repo')
from ceracoder_b200 import CudaH264Enc, load_library
from ceracoder_b200.synth import make_source
L = load_library()
W, H = 1920, 1080
src = make_source("pan", W, H)
pool = [src.frame(i) for i in range(12)]
def run(ns, kbps, n=90):
    encs = [CudaH264Enc(W, H, fps=(60, 1), bitrate=kbps, lib=L) for _ in range(ns)]
    res = [None] * ns
    def work(k):
        e = encs[k]; g = h = 0.0; sz = 0
        for i in range(n):
            y, u, v = pool[i % 12 if (i // 12) % 2 == 0 else 11 - i % 12]
            e.encode(y, u, v)
            if i >= 30: g += e.last.gpu_ms; h += e.last.host_ms; sz += e.last.size
        res[k] = (g / (n - 30), h / (n - 30), sz / (n - 30), e.last.qp)
    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(k,)) for k in range(ns)]
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.perf_counter() - t0
    print(f"streams {ns} kbps {kbps}: fps {ns * n / dt:.0f}  gpu_ms {np.mean([r[0] for r in res]):.3f} host_ms {np.mean([r[1] for r in res]):.3f} bytes {np.mean([r[2] for r in res]):.0f} qp {res[0][3]}", flush=True)
    [e.close() for e in encs]
for ns, kb in ((1, 12000), (1, 3000), (2, 12000), (4, 12000), (8, 12000), (8, 3000), (16, 6000)):
    run(ns, kb)
