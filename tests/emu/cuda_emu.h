/*
 * tests/emu/cuda_emu.h — a tiny single-threaded SIMT emulator so that the sm_100a kernel sources of
 * libceracuda can be compiled with g++ and executed on a CPU-only machine (this container has no GPU).
 *
 * TEST INFRASTRUCTURE ONLY: it exists so the host logic (rate control, CAVLC writer, NAL packaging) and
 * the kernels' warp-level algorithms can be checked bit-exactly against the oracle in `-m "not gpu"`
 * tests before GPU time is spent.  The product (libceracuda.so) is built by nvcc from the same sources
 * and never includes this file.
 *
 * Model: every CUDA thread of one CTA is a fiber (own stack, hand-written x86-64 context switch); CTAs run one
 * after another (the kernels never wait across CTAs).  Warp collectives must be called with the full
 * mask by all 32 lanes — anything else aborts, which doubles as a divergence check.
 */
#ifndef CUDA_EMU_H
#define CUDA_EMU_H
#ifndef __x86_64__
#error "cuda_emu.h: x86-64 only"
#endif
#include "cuda_emu_rt.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <functional>
#include <vector>
#include <algorithm>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __shared__ static
#define __constant__ static const
#define __restrict__ __restrict
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
struct int2 { int x, y; };
struct short2 { short x, y; };
struct uchar4 { unsigned char x, y, z, w; };
static inline uint2 make_uint2(unsigned a, unsigned b) { return uint2{a, b}; }
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return uint4{a, b, c, d}; }
static inline int2 make_int2(int a, int b) { return int2{a, b}; }
static inline short2 make_short2(short a, short b) { return short2{a, b}; }

namespace emu {
struct Warp { int count = 0; unsigned gen = 0; uint64_t slot[32]; unsigned ballot = 0; };
struct Fiber {
    void *sp = nullptr; char *stack = nullptr; bool done = false;
    uint3 tid; int lane, warp;
};
struct State {
    std::vector<Fiber> fibers; std::vector<Warp> warps;
    Fiber *cur = nullptr; void *sched_sp = nullptr;
    uint3 bid; dim3 bdim, gdim;
    int cta_count = 0; unsigned cta_gen = 0;
    unsigned long progress = 0;
    const std::function<void()> *body = nullptr;
    std::vector<unsigned char> dyn_smem;
};
State &S();
void yield();
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
void warp_barrier();
[[noreturn]] void die(const char *msg);
static inline Warp &W() { return S().warps[S().cur->warp]; }
static inline unsigned char *dyn_smem() { return S().dyn_smem.data(); }
static const size_t STACK = 256 * 1024;
}

#define threadIdx (emu::S().cur->tid)
#define blockIdx (emu::S().bid)
#define blockDim (emu::S().bdim)
#define gridDim (emu::S().gdim)
#define warpSize 32

static inline void __syncwarp(unsigned m = 0xffffffffu) { if (m != 0xffffffffu) emu::die("__syncwarp: partial mask"); emu::warp_barrier(); }
void __syncthreads();
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __nanosleep(unsigned) { emu::yield(); }
static inline long long clock64() { return 0; }

template <typename T> static inline T emu_xchg(unsigned m, T v, int src_lane, bool valid) {
    static_assert(sizeof(T) <= 8, "shfl: type too wide");
    if (m != 0xffffffffu) emu::die("warp collective: partial mask");
    uint64_t raw = 0; memcpy(&raw, &v, sizeof(T));
    emu::Fiber *f = emu::S().cur;
    emu::W().slot[f->lane] = raw;
    emu::warp_barrier();
    uint64_t got = emu::W().slot[valid ? (src_lane & 31) : f->lane];
    emu::warp_barrier();
    T out; memcpy(&out, &got, sizeof(T)); return out;
}
template <typename T> static inline T __shfl_sync(unsigned m, T v, int src, int width = 32) {
    int lane = emu::S().cur->lane, base = lane & ~(width - 1);
    return emu_xchg(m, v, base + (src & (width - 1)), true);
}
template <typename T> static inline T __shfl_xor_sync(unsigned m, T v, int x, int width = 32) {
    int lane = emu::S().cur->lane, src = lane ^ x;
    return emu_xchg(m, v, src, (src & ~(width - 1)) == (lane & ~(width - 1)));
}
template <typename T> static inline T __shfl_down_sync(unsigned m, T v, unsigned d, int width = 32) {
    int lane = emu::S().cur->lane, src = lane + (int)d;
    return emu_xchg(m, v, src, (src & ~(width - 1)) == (lane & ~(width - 1)) && src < 32);
}
template <typename T> static inline T __shfl_up_sync(unsigned m, T v, unsigned d, int width = 32) {
    int lane = emu::S().cur->lane, src = lane - (int)d;
    return emu_xchg(m, v, src, src >= (lane & ~(width - 1)));
}
static inline unsigned __ballot_sync(unsigned m, int pred) {
    if (m != 0xffffffffu) emu::die("ballot: partial mask");
    emu::Fiber *f = emu::S().cur;
    emu::W().slot[f->lane] = (uint64_t)(pred != 0);
    emu::warp_barrier();
    unsigned r = 0;
    for (int i = 0; i < 32; i++) r |= (unsigned)(emu::W().slot[i] & 1) << i;
    emu::warp_barrier();
    return r;
}
static inline unsigned __reduce_min_sync(unsigned m, unsigned v) { for (int o = 16; o; o >>= 1) { unsigned t = __shfl_xor_sync(m, v, o); v = t < v ? t : v; } return v; }
static inline unsigned __reduce_add_sync(unsigned m, unsigned v) { for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(m, v, o); return v; }
static inline int __reduce_add_sync(unsigned m, int v) { for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(m, v, o); return v; }
static inline unsigned __reduce_or_sync(unsigned m, unsigned v) { for (int o = 16; o; o >>= 1) v |= __shfl_xor_sync(m, v, o); return v; }
static inline int __any_sync(unsigned m, int p) { return __ballot_sync(m, p) != 0; }
static inline int __all_sync(unsigned m, int p) { return __ballot_sync(m, p) == 0xffffffffu; }
static inline unsigned __activemask() { return 0xffffffffu; }

static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; i++) r |= ((x >> i) & 1u) << (31 - i); return r; }
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned s) {
    uint64_t v = ((uint64_t)b << 32) | a; unsigned r = 0;
    for (int i = 0; i < 4; i++) {
        unsigned sel = (s >> (4 * i)) & 0xf; unsigned byte = (unsigned)(v >> (8 * (sel & 7))) & 0xff;
        if (sel & 8) byte = (byte & 0x80) ? 0xff : 0;
        r |= byte << (8 * i);
    }
    return r;
}
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { uint64_t v = ((uint64_t)hi << 32) | lo; return (unsigned)(v >> (sh & 31)); }
static inline unsigned __funnelshift_rc(unsigned lo, unsigned hi, unsigned sh) { uint64_t v = ((uint64_t)hi << 32) | lo; return (unsigned)(v >> (sh > 32 ? 32 : sh)); }
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned sh) { uint64_t v = ((uint64_t)hi << 32) | lo; return (unsigned)((v << (sh & 31)) >> 32); }
static inline unsigned __vsadu4(unsigned a, unsigned b) { unsigned s = 0; for (int i = 0; i < 4; i++) { int x = (a >> (8 * i)) & 255, y = (b >> (8 * i)) & 255; s += (unsigned)(x > y ? x - y : y - x); } return s; }
static inline unsigned __vabsdiffu4(unsigned a, unsigned b) { unsigned r = 0; for (int i = 0; i < 4; i++) { int x = (a >> (8 * i)) & 255, y = (b >> (8 * i)) & 255; r |= (unsigned)(x > y ? x - y : y - x) << (8 * i); } return r; }
static inline unsigned __vavgu4(unsigned a, unsigned b) { unsigned r = 0; for (int i = 0; i < 4; i++) { unsigned x = (a >> (8 * i)) & 255, y = (b >> (8 * i)) & 255; r |= ((x + y + 1) >> 1) << (8 * i); } return r; }
static inline int __dp4a(unsigned a, unsigned b, int c) { for (int i = 0; i < 4; i++) c += (int)((a >> (8 * i)) & 255) * (int)((b >> (8 * i)) & 255); return c; }
static inline unsigned __sad(int a, int b, unsigned c) { return c + (unsigned)(a > b ? a - b : b - a); }
static inline unsigned __usad(unsigned a, unsigned b, unsigned c) { return c + (a > b ? a - b : b - a); }
template <typename T> static inline T __ldg(const T *p) { return *p; }
template <typename T> static inline T __ldcg(const T *p) { T v; memcpy(&v, (const void *)p, sizeof(T)); return v; } /* the emulator runs one fiber at a time: plain loads */
template <typename T> static inline T __ldcv(const T *p) { T v; memcpy(&v, (const void *)p, sizeof(T)); return v; }
template <typename T> static inline void __stcg(T *p, T v) { *p = v; }
static inline int atomicAdd(int *p, int v) { int o = *p; *p += v; return o; }
static inline unsigned atomicAdd(unsigned *p, unsigned v) { unsigned o = *p; *p += v; return o; }
static inline unsigned atomicOr(unsigned *p, unsigned v) { unsigned o = *p; *p |= v; return o; }
static inline unsigned atomicMin(unsigned *p, unsigned v) { unsigned o = *p; if (v < o) *p = v; return o; }
static inline unsigned atomicExch(unsigned *p, unsigned v) { unsigned o = *p; *p = v; return o; }
using std::min;
using std::max;
static inline int abs_i(int x) { return x < 0 ? -x : x; }

#define CC_EMU 1
#endif
