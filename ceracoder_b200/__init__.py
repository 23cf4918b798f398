"""ceracoder_b200 — host-side mirror of the encoder slot of a ceracoder pipeline, backed by libceracuda.so
(hand-written sm_100a CUDA kernels + host C).  There is no CPU fallback: importing works anywhere, but
opening an encoder needs the built library and a CUDA device."""
from .encoder import CudaH264Enc, load_library, LIB_PATH, CeraCudaError  # noqa: F401
