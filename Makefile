# Builds libceracuda.so (the product: nvcc, sm_100a), the oracle (test infrastructure, gcc) and the CPU
# SIMT-emulation build of the same sources (test infrastructure: lets `-m "not gpu"` tests run the kernels' algorithms).
NVCC ?= /usr/local/cuda/bin/nvcc
CC = /usr/bin/gcc
CXX = /usr/bin/g++
SRC = ceracoder_b200/csrc
CU = $(SRC)/cc_k_frame.cu $(SRC)/cc_k_motion.cu $(SRC)/cc_k_mb.cu $(SRC)/cc_k_cavlc.cu $(SRC)/cc_k_prims.cu
CS = $(SRC)/cc_api.c $(SRC)/cc_bitstream.c $(SRC)/cc_cavlc.c $(SRC)/cc_rc.c
HDR = $(wildcard $(SRC)/*.h $(SRC)/*.cuh include/*.h)
NVFLAGS = -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O2 -Xptxas -v
LIB = ceracoder_b200/libceracuda.so
EMU = tests/emu/_build/libceracuda_emu.so

all: $(LIB) oracle
lib: $(LIB)
$(LIB): $(CU) $(CS) $(HDR)
	mkdir -p build
	for f in $(CS); do $(CC) -O2 -fPIC -std=gnu11 -Wall -I/usr/local/cuda/include -c $$f -o build/$$(basename $$f .c).o || exit 1; done
	for f in $(CU); do $(NVCC) $(NVFLAGS) -c $$f -o build/$$(basename $$f .cu).o 2> build/$$(basename $$f .cu).ptxas.log || { cat build/$$(basename $$f .cu).ptxas.log; exit 1; }; done
	$(NVCC) -shared -o $@ build/cc_api.o build/cc_bitstream.o build/cc_cavlc.o build/cc_rc.o build/cc_k_frame.o build/cc_k_motion.o build/cc_k_mb.o build/cc_k_cavlc.o build/cc_k_prims.o -lm -lpthread
oracle:
	$(MAKE) -C oracle
emu: $(EMU)
$(EMU): $(CU) $(CS) $(HDR) tests/emu/cuda_emu.h tests/emu/cuda_emu_rt.h tests/emu/cuda_emu.cpp
	mkdir -p tests/emu/_build
	for f in $(CS); do $(CC) -O2 -g -fPIC -std=gnu11 -Wall -DCC_EMU_BUILD -Itests/emu -c $$f -o tests/emu/_build/$$(basename $$f .c).o || exit 1; done
	for f in $(CU); do $(CXX) -O2 -g -fPIC -std=c++17 -Wall -Wno-unused-function -Wno-unknown-pragmas -DCC_EMU_BUILD -Itests/emu -include tests/emu/cuda_emu.h -x c++ -c $$f -o tests/emu/_build/$$(basename $$f .cu).o || exit 1; done
	$(CXX) -O2 -g -fPIC -std=c++17 -Itests/emu -c tests/emu/cuda_emu.cpp -o tests/emu/_build/cuda_emu.o
	$(CXX) -shared -o $@ tests/emu/_build/*.o -lm -lpthread
clean:
	rm -rf build tests/emu/_build $(LIB)
	$(MAKE) -C oracle clean
.PHONY: all lib oracle emu clean

# GStreamer element (outer boundary). Built only where GStreamer development files exist.
gst:
	@pkg-config --exists gstreamer-video-1.0 || { echo "gstreamer-video-1.0 not found: cudah264enc is compile-gated (see INTEGRATION.md)"; exit 0; }; \
	$(CC) -O2 -fPIC -shared -std=gnu11 -Wall $$(pkg-config --cflags gstreamer-video-1.0) ceracoder_b200/gst/gstcudah264enc.c -o ceracoder_b200/gst/libgstcudah264enc.so \
	  -Lceracoder_b200 -lceracuda -Wl,-rpath,'$$ORIGIN/..' $$(pkg-config --libs gstreamer-video-1.0)
.PHONY: gst
