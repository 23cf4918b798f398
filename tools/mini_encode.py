"""Three frames of a small picture through libceracuda: the shortest run that exercises every kernel (CC_DEBUG_SYNC=1 names a faulting stage)."""
import sys
sys.path.insert(0,'/root/repo')
from ceracoder_b200 import CudaH264Enc, load_library
from ceracoder_b200.synth import make_source
L=load_library()
w,h=176,144
src=make_source("pan",w,h)
e=CudaH264Enc(w,h,lib=L,qp_fixed=28)
for i in range(3):
    au=e.encode(*src.frame(i)); print(i,len(au))
e.close()
