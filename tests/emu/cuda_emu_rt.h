/*
 * tests/emu/cuda_emu_rt.h — C-compatible stand-in for the slice of the CUDA runtime API that
 * libceracuda's host code uses, for the CPU SIMT-emulation build (tests only, never shipped).
 * Device memory is host memory; streams are synchronous; events are wall-clock stamps.
 */
#ifndef CUDA_EMU_RT_H
#define CUDA_EMU_RT_H
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <time.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef struct { double t; } *cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaHostAllocDefault = 0, cudaHostAllocMapped = 2, cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaEventDefault = 0 };
static inline const char *cudaGetErrorString(cudaError_t e) { return e ? "emu error" : "no error"; }
static inline cudaError_t cudaGetLastError(void) { return 0; }
static inline cudaError_t cudaSetDevice(int d) { (void)d; return 0; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return 0; }
static inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return 0; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = malloc(n ? n : 1); if (*p) memset(*p, 0xCD, n); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaHostAlloc(void **p, size_t n, unsigned f) { (void)f; *p = malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaMallocHost(void **p, size_t n) { return cudaHostAlloc(p, n, 0); }
static inline cudaError_t cudaFreeHost(void *p) { free(p); return 0; }
enum { cudaHostRegisterDefault = 0 };
static inline cudaError_t cudaHostRegister(void *p, size_t n, unsigned f) { (void)p; (void)n; (void)f; return 0; }
static inline cudaError_t cudaHostUnregister(void *p) { (void)p; return 0; }
static inline cudaError_t cudaHostGetDevicePointer(void **d, void *h, unsigned f) { (void)f; *d = h; return 0; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, enum cudaMemcpyKind k) { (void)k; memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, enum cudaMemcpyKind k, cudaStream_t st) { (void)k; (void)st; memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, enum cudaMemcpyKind k, cudaStream_t st) {
    (void)k; (void)st;
    for (size_t y = 0; y < h; y++) memmove((char *)d + y * dp, (const char *)s + y * sp, w);
    return 0;
}
static inline cudaError_t cudaMemcpy2D(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, enum cudaMemcpyKind k) { return cudaMemcpy2DAsync(d, dp, s, sp, w, h, k, 0); }
enum cudaMemoryType { cudaMemoryTypeUnregistered = 0, cudaMemoryTypeHost = 1, cudaMemoryTypeDevice = 2, cudaMemoryTypeManaged = 3 };
struct cudaPointerAttributes { enum cudaMemoryType type; int device; void *devicePointer, *hostPointer; };
/* the emulator cannot tell pinned from pageable memory: report pageable, so the staging copy of cc_enc_submit is exercised */
static inline cudaError_t cudaPointerGetAttributes(struct cudaPointerAttributes *a, const void *p) { (void)p; a->type = cudaMemoryTypeUnregistered; a->device = 0; a->devicePointer = a->hostPointer = 0; return 0; }
/* a "device" symbol of the emulation build is a host object */
#define cudaGetSymbolAddress(pp, sym) (*(pp) = (void *)&(sym), 0)
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t st) { (void)st; memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned f) { (void)f; *s = 0; return 0; }
static inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t *s, unsigned f, int p) { (void)f; (void)p; *s = 0; return 0; }
static inline cudaError_t cudaDeviceGetStreamPriorityRange(int *lo, int *hi) { *lo = 0; *hi = 0; return 0; }
static inline cudaError_t cudaStreamCreate(cudaStream_t *s) { *s = 0; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { (void)s; return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t s) { (void)s; return 0; }
static inline cudaError_t cudaDeviceSynchronize(void) { return 0; }
static inline double emu_now_ms(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = (cudaEvent_t)calloc(1, sizeof(**e)); return 0; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned f) { (void)f; return cudaEventCreate(e); }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { free(e); return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s) { (void)s; e->t = emu_now_ms(); return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t e) { (void)e; return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned f) { (void)s; (void)e; (void)f; return 0; } /* launches run to completion in order */
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return 0; }
#ifdef __cplusplus
}
#endif
#endif
