"""All-IDR 1080p stream with CC_DEBUG_CYCLES=1: cycle counters of the full intra path of the wavefront kernel (run under gpurun)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CC_DEBUG_CYCLES"] = "1"
from ceracoder_b200 import CudaH264Enc
from ceracoder_b200.synth import make_source
w, h = 1920, 1080
src = make_source("pan", w, h)
e = CudaH264Enc(w, h, key_int_max=1, bitrate=12000)
for i in range(20):
    y, u, v = src.frame(i)
    e.encode(y, u, v)
print("qp", e.last.qp, "gpu_ms", e.last.gpu_ms)
e.close()
