// Does a kernel on another stream start BEFORE the kernel it depends on through a programmatic event (trigger at block start) ends?
// And: programmatic stream serialisation in one stream with griddepcontrol.launch_dependents at the top of the primary.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o pdl_probe pdl_probe.cu ; prints times in microseconds relative to A's start
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
__device__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__global__ void A(unsigned long long *t, int trigger) {
    if (trigger) asm volatile("griddepcontrol.launch_dependents;");
    const unsigned long long t0 = gt();
    if (threadIdx.x == 0 && blockIdx.x == 0) t[0] = t0;
    while (gt() - t0 < 2000000ull) __nanosleep(1000);
    if (threadIdx.x == 0 && blockIdx.x == 0) t[1] = gt();
}
__global__ void B(unsigned long long *t) { if (threadIdx.x == 0 && blockIdx.x == 0) t[2] = gt(); }
int main() {
    unsigned long long *t, h[3];
    cudaMalloc(&t, 24);
    cudaStream_t s1, s2; cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    cudaEvent_t ev; cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    for (int mode = 0; mode < 4; mode++) {
        cudaMemset(t, 0, 24); cudaDeviceSynchronize();
        cudaLaunchAttribute at[1]; cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3(9); cfg.blockDim = dim3(256); cfg.attrs = at; cfg.numAttrs = 1;
        if (mode < 2) { // event across streams; mode 1: the dependent opts in with the serialisation attribute as well
            at[0].id = cudaLaunchAttributeProgrammaticEvent; at[0].val.programmaticEvent.event = ev; at[0].val.programmaticEvent.flags = 0; at[0].val.programmaticEvent.triggerAtBlockStart = 1;
            cfg.stream = s1; cudaLaunchKernelEx(&cfg, A, t, 0);
            cudaStreamWaitEvent(s2, ev, 0);
            if (mode == 0) B<<<9, 512, 0, s2>>>(t);
            else { at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1; cfg.stream = s2; cfg.blockDim = dim3(512); cudaLaunchKernelEx(&cfg, B, t); }
        } else { // same stream; mode 2: A triggers at its top, mode 3: it does not (implicit trigger at exit)
            A<<<9, 256, 0, s1>>>(t, mode == 2);
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1; cfg.stream = s1; cfg.blockDim = dim3(512);
            cudaLaunchKernelEx(&cfg, B, t);
        }
        cudaError_t er = cudaDeviceSynchronize();
        cudaMemcpy(h, t, 24, cudaMemcpyDeviceToHost);
        printf("mode %d (%s): A ran %.1f us, B started at %.1f us  [%s]\n", mode,
               mode == 0 ? "programmatic event, plain dependent" : mode == 1 ? "programmatic event, dependent with serialisation attr" : mode == 2 ? "stream serialisation, early trigger" : "stream serialisation, no trigger",
               (h[1] - h[0]) / 1e3, ((long long)h[2] - (long long)h[0]) / 1e3, cudaGetErrorString(er));
    }
    return 0;
}
