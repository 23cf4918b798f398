"""Loads the CPU SIMT-emulation build of libceracuda (tests/emu/_build/libceracuda_emu.so): the product's own
sources compiled with g++ against tests/emu/cuda_emu.h.  TEST INFRASTRUCTURE: lets CPU-only tests exercise the host
logic (rate control, CAVLC writer, NAL packaging) and the kernels' algorithms.  The product never loads this."""
import os
import subprocess
from ceracoder_b200.encoder import load_library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "emu", "_build", "libceracuda_emu.so")


def emu_library():
    subprocess.check_call(["make", "-s", "-C", ROOT, "emu"], stdout=subprocess.DEVNULL)
    return load_library(SO)
