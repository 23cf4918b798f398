/*
 * cc_k_motion.cu — motion estimation: fully parallel over macroblocks, one warp
 * per macroblock, integer pipes only.
 *   k_me_l1  coarse level: exhaustive +-16 search at half resolution on the MB's 8x8 footprint.  Eight macroblocks
 *            share one TMA-staged window tile, expanded into four byte-phase planes; a thread scores the 33 candidates
 *            of one column of one macroblock's window with VABSDIFF4.ACC on aligned 64-bit loads.
 *            cost = 4*SAD + |dx| + |dy|, lowest candidate index wins ties.
 *   k_me_l0  fine level: 27 full-pel candidates (SAD + lambda*mvd bits) around twice the coarse vector, then
 *            9 half-pel and 8 quarter-pel candidates scored by 4x4-Hadamard SATD.  Quarter-pel predictors are
 *            averages (__vavgu4) of two samples from the {G,b,h,j} planes built by k_hpel.
 * The predictor used for the mvd cost is the median of the neighbours' COARSE vectors, which removes the
 * raster dependency of true MV prediction from the search (the exact mvd is charged later, in k_mb_encode).
 * Oracle counterparts: oracle/ref_enc.c me_l1 / me_l0.
 */
#include "cc_dev.cuh"

#define L1_MBS 8           /* k_me_l1: macroblocks per CTA (they share one window tile) */
#define L1_PITCHW 28       /* words per staged row: 8 macroblocks x 8 pixels + 32 of search range = 96 bytes, + the bytes the last phase-plane pair
                            * reads past them, rounded up to the 16 bytes a TMA box row is a multiple of */
#define L1_ROWS 40         /* 8 + 2 * 16 */
#ifndef CC_EMU_BUILD
#include <cuda.h>          /* CUtensorMap */
#endif
#ifndef ME0_WARPS
#define ME0_WARPS 4 /* k_me_l0: warps (macroblocks) per CTA */
#endif

/* CTA = eight horizontally adjacent macroblocks (grid = (mb_w / 8, mb_h)): their search windows overlap, so the CTA stages ONE
 * 112-byte x 40-row tile of the half-resolution reference with a single TMA load (cp.async.bulk.tensor.2d into shared memory,
 * completion on an mbarrier).
 *
 * The kernel is bound by the VABSDIFF4 pipe (half rate: the measured peak is one per two issue slots), so everything else has to
 * fit into the other half of the slots.  Two things make it fit:
 *   - the tile is expanded once per CTA into four byte-phase planes of (word, next word) pairs: a candidate whose window starts
 *     at byte X of a row reads its 8 pixels with ONE aligned 64-bit load from plane X & 3 — no funnel shifts in the search loop
 *     (the first version spent 54 of them per 128 abs-diffs);
 *   - work item = (macroblock, dx, all 33 dy): each of the 40 window rows it reads is used by up to 8 of its 33 candidates, whose
 *     SADs build up in 33 registers: 40 loads + 528 VABSDIFF4.ACC + 3 instructions of cost / minimum per candidate, all offsets
 *     and candidate constants immediate.
 * 8 macroblocks x 33 dx = 264 items = one per thread of a 9-warp CTA.  The winner per macroblock is the minimum of
 * (cost << 11 | candidate index) over its 33 threads (shared-memory atomicMin): the lowest index wins ties, as in the oracle. */
#define L1_NTHR 288        /* 9 warps: 264 work items */
#define L1_QROW 23         /* pairs per plane row: window start words 0 .. 22 (X = 8 * 7 + 32 = 88 at most) */
#define L1_QPLANE 932      /* pairs per phase plane: 40 * 23 = 920, padded so that the four phases of a half-warp's loads fall on different banks (932 = 4 mod 16) */
#ifndef CC_EMU_BUILD
__global__ void __launch_bounds__(L1_NTHR, 7) k_me_l1(cc_kparams k, const __grid_constant__ CUtensorMap tmap) {
#else
__global__ void __launch_bounds__(L1_NTHR, 7) k_me_l1(cc_kparams k) {
#endif
    /* the tile lives in the tail of the planes' own storage (it is dead once they are built): 30 KB of shared memory and 32 registers
     * let seven CTAs share an SM, which makes the 1020 CTAs of a 1080p frame one wave on 148 SMs */
    __shared__ __align__(128) uint2 Q[4 * L1_QPLANE];
    uint32_t *const win = (uint32_t *)(Q + 4 * L1_QPLANE) - L1_ROWS * L1_PITCHW;
    __shared__ __align__(16) uint32_t scur[L1_MBS][16]; /* the macroblocks' 8x8 half-resolution source blocks: 8 rows x 2 words */
    __shared__ unsigned smin[L1_MBS];
    const int tid = threadIdx.x;
    const int mbx0 = blockIdx.x * L1_MBS, mby = blockIdx.y;
#ifndef CC_EMU_BUILD
    __shared__ __align__(8) unsigned long long mbar;
    const uint32_t mbar_a = (uint32_t)__cvta_generic_to_shared(&mbar), win_a = (uint32_t)__cvta_generic_to_shared(win);
    if (tid == 0) { /* tile origin in allocation coordinates: the plane starts CC_PADC rows / columns into it */
        const int tx = CC_PADC + mbx0 * 8 - CC_L1R, ty = CC_PADC + mby * 8 - CC_L1R;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_a) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar_a), "r"(L1_ROWS * L1_PITCHW * 4) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(win_a), "l"((unsigned long long)&tmap), "r"(tx), "r"(ty), "r"(mbar_a) : "memory");
    }
#endif
    if (tid < L1_MBS * 16) { /* source blocks; a macroblock past the right edge repeats the last one (its result is not stored) */
        const int w = tid >> 4, i = tid & 15, mbx = min(mbx0 + w, k.mb_w - 1);
        scur[w][i] = *(const uint32_t *)(k.l1cur + (size_t)(mby * 8 + (i >> 1)) * k.sc + mbx * 8 + 4 * (i & 1));
    }
    if (tid < L1_MBS) smin[tid] = 0xffffffffu;
    __syncthreads();
#ifndef CC_EMU_BUILD
    if (tid < 4 * L1_ROWS) { /* the tile has landed when the mbarrier's phase 0 completes (only the threads that read it wait) */
        uint32_t ok;
        do {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(mbar_a) : "memory");
        } while (!ok);
    }
#else
    { /* bytes past the allocation's row are never used by a stored result: clamp the read instead of a TMA's zero fill */
        const uint8_t *ref = k.l1ref + (ptrdiff_t)(mby * 8 - CC_L1R) * k.sc + (mbx0 * 8 - CC_L1R); /* 4-byte aligned */
        const int avail = (k.sc - (CC_PADC + mbx0 * 8 - CC_L1R)) / 4; /* words left in the allocation's row */
        for (int i = tid; i < L1_ROWS * L1_PITCHW; i += L1_NTHR) {
            const int r = i / L1_PITCHW, c = i - r * L1_PITCHW;
            win[i] = c < avail ? *(const uint32_t *)(ref + (ptrdiff_t)r * k.sc + 4 * c) : 0u;
        }
        __syncthreads();
    }
#endif
    /* phase planes: Q[p][row][j] = bytes 4 j + p .. 4 j + p + 7 of the tile row; thread = (phase, row), sliding along the row */
    {
        const int p = tid & 3, row = tid >> 2;
        uint32_t r[L1_QROW + 2];
        if (tid < 4 * L1_ROWS) {
            const uint32_t *wp = win + row * L1_PITCHW;
#pragma unroll
            for (int j = 0; j < L1_QROW + 2; j++) r[j] = wp[j];
        }
        __syncthreads(); /* every row is in registers: the planes may now overwrite the tile */
        if (tid < 4 * L1_ROWS) {
            uint2 *qp = Q + p * L1_QPLANE + row * L1_QROW;
            uint32_t a = __funnelshift_r(r[0], r[1], 8 * p);
#pragma unroll
            for (int j = 0; j < L1_QROW; j++) {
                const uint32_t b = __funnelshift_r(r[j + 1], r[j + 2], 8 * p);
                qp[j] = make_uint2(a, b);
                a = b;
            }
        }
    }
    __syncthreads();
    if (tid < L1_MBS * 33) {
        const int w = tid / 33, dx = tid - w * 33; /* window coordinates: dx, dy in 0..32 */
        const int X = 8 * w + dx;
        const uint2 *q = Q + (X & 3) * L1_QPLANE + (X >> 2);
        uint32_t c0[8], c1[8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint4 v = *(const uint4 *)&scur[w][4 * i];
            c0[2 * i] = v.x; c1[2 * i] = v.y; c0[2 * i + 1] = v.z; c1[2 * i + 1] = v.w;
        }
        uint32_t acc[33];
#pragma unroll
        for (int d = 0; d < 33; d++) acc[d] = 0;
#pragma unroll
        for (int rr = 0; rr < L1_ROWS; rr++) {
            const uint2 ab = q[rr * L1_QROW];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int d = rr - i;
                if (d >= 0 && d < 33) acc[d] = cc_sad4_acc(ab.x, c0[i], cc_sad4_acc(ab.y, c1[i], acc[d]));
            }
        }
        /* key = (4 SAD + |dx - 16| + |dy - 16|) << 11 | (33 dy + dx) = SAD * 8192 + (the thread's part) + (a constant per dy) */
        const unsigned base = ((unsigned)cc_iabs(dx - 16) << 11) + (unsigned)dx;
        unsigned best = 0xffffffffu;
#pragma unroll
        for (int d = 0; d < 33; d++)
            best = min(best, acc[d] * 8192u + (base + (unsigned)(((d < 16 ? 16 - d : d - 16) << 11) + 33 * d)));
        atomicMin(&smin[w], best);
    }
    __syncthreads();
    if (tid < L1_MBS && mbx0 + tid < k.mb_w) {
        const int idx = (int)(smin[tid] & 2047u), mb = mby * k.mb_w + mbx0 + tid;
        k.me[mb].l1_mv[0] = (int16_t)(idx % 33 - 16);
        k.me[mb].l1_mv[1] = (int16_t)(idx / 33 - 16);
    }
}

/* predicted 4x4 block (4 rows of 4 pixels) at quarter-pel vector (mvx,mvy) relative to pixel (px,py) */
__device__ __forceinline__ void cc_qpel4x4(const cc_kparams &k, int px, int py, int mvx, int mvy, uint32_t out[4]) {
    int e = c_qpel[(mvy & 3) * 4 + (mvx & 3)];
    const int pa = e & 3, pb = (e >> 4) & 3; /* plane selects, not an indexed pointer array (that would live on the stack) */
    const int base = (py + (mvy >> 2)) * k.sy + px + (mvx >> 2);
    const uint8_t *A = (pa == 0 ? k.ref[0] : (pa == 1 ? k.hb : (pa == 2 ? k.hh : k.hj))) + (base + ((e >> 2) & 1) + ((e >> 3) & 1) * k.sy);
    const uint8_t *B = (pb == 0 ? k.ref[0] : (pb == 1 ? k.hb : (pb == 2 ? k.hh : k.hj))) + (base + ((e >> 6) & 1) + ((e >> 7) & 1) * k.sy);
#pragma unroll
    for (int r = 0; r < 4; r++) out[r] = __vavgu4(cc_ld32u(A + (ptrdiff_t)r * k.sy), cc_ld32u(B + (ptrdiff_t)r * k.sy));
}

/* The same from the warp's staged window: tiles[p] = plane p (G, b, h, j) over SP_W x SP_H bytes whose top-left sample is
 * (wx, wy) relative to the macroblock origin; (bx, by) = the block's offset inside the macroblock */
#define SP_W 48 /* 32 columns are needed; the window starts on a 16-byte boundary of the plane (TMA), up to 15 columns further left */
#define SP_H 28
#define SP_ORG 4 /* the window starts (at least) 4 samples left of / above the vector it is staged around (the coarse vector minus one: 5 left of / above it) */
#define SP_PLANE 1408 /* bytes from one plane's tile to the next: SP_W * SP_H rounded up to the 128 bytes a TMA destination must be aligned to */
__device__ __forceinline__ void cc_qpel4x4_s(const uint8_t *tiles, int wx, int wy, int bx, int by, int mvx, int mvy, uint32_t out[4]) {
    const int e = c_qpel[(mvy & 3) * 4 + (mvx & 3)];
    const int pa = e & 3, pb = (e >> 4) & 3;
    const int base = (by + (mvy >> 2) - wy) * SP_W + bx + (mvx >> 2) - wx;
    const uint8_t *A = tiles + pa * SP_PLANE + base + ((e >> 2) & 1) + ((e >> 3) & 1) * SP_W;
    const uint8_t *B = tiles + pb * SP_PLANE + base + ((e >> 6) & 1) + ((e >> 7) & 1) * SP_W;
#pragma unroll
    for (int r = 0; r < 4; r++) out[r] = __vavgu4(cc_ld32u(A + r * SP_W), cc_ld32u(B + r * SP_W));
}
/* the same for a vector whose components are both even (integer and half-pel positions): the sample IS one plane's sample */
__device__ __forceinline__ void cc_hpel4x4_s(const uint8_t *tiles, int wx, int wy, int bx, int by, int mvx, int mvy, uint32_t out[4]) {
    const int pa = ((mvx >> 1) & 1) | (mvy & 2); /* G, b, h, j */
    const uint8_t *A = tiles + pa * SP_PLANE + (by + (mvy >> 2) - wy) * SP_W + bx + (mvx >> 2) - wx;
    const uintptr_t a = (uintptr_t)A;
    const uint32_t *q = (const uint32_t *)(a & ~(uintptr_t)3);
    const unsigned sh = (unsigned)(a & 3) * 8u;
#pragma unroll
    for (int r = 0; r < 4; r++) out[r] = __funnelshift_r(q[r * (SP_W / 4)], q[r * (SP_W / 4) + 1], sh);
}

/* Full-pel SADs of five horizontally adjacent candidates at once.  A lane takes four source rows (4 g .. 4 g + 3) of one
 * candidate row: it loads the 24-byte strip of the reference row once (six aligned words, realigned with five funnel shifts) and
 * slides the 16-pixel window over it in registers — a quarter of the loads of "lane = candidate", and every load is used by five
 * candidates.  r points at the pixel of candidate dx = 0 in the first of the lane's rows; acc[dx][h] += SAD of the left (h = 0) /
 * right (h = 1) 8 columns. */
__device__ __forceinline__ void sad_strip5(const uint8_t *r, int sy, const uint32_t *sc4, unsigned acc[5][2]) {
    const uintptr_t a = (uintptr_t)r;
    const unsigned sh = (unsigned)(a & 3) * 8u;
    const uint32_t *q = (const uint32_t *)(a & ~(uintptr_t)3);
#pragma unroll
    for (int y = 0; y < 4; y++) {
        const uint32_t *w = (const uint32_t *)((const uint8_t *)q + (ptrdiff_t)y * sy);
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4], w5 = w[5];
        uint32_t v[6];
        v[0] = __funnelshift_r(w0, w1, sh); v[1] = __funnelshift_r(w1, w2, sh); v[2] = __funnelshift_r(w2, w3, sh);
        v[3] = __funnelshift_r(w3, w4, sh); v[4] = __funnelshift_r(w4, w5, sh); v[5] = 0;
        const uint32_t c0 = sc4[y * 4], c1 = sc4[y * 4 + 1], c2 = sc4[y * 4 + 2], c3 = sc4[y * 4 + 3];
#pragma unroll
        for (int dx = 0; dx < 5; dx++) {
            const int j = dx >> 2;
            const unsigned s8 = (unsigned)(dx & 3) * 8u;
            const uint32_t p0 = s8 ? __funnelshift_r(v[j], v[j + 1], s8) : v[j], p1 = s8 ? __funnelshift_r(v[j + 1], v[j + 2], s8) : v[j + 1];
            const uint32_t p2 = s8 ? __funnelshift_r(v[j + 2], v[j + 3], s8) : v[j + 2], p3 = s8 ? __funnelshift_r(v[j + 3], v[j + 4], s8) : v[j + 3];
            acc[dx][0] = cc_sad4_acc(p1, c1, cc_sad4_acc(p0, c0, acc[dx][0]));
            acc[dx][1] = cc_sad4_acc(p3, c3, cc_sad4_acc(p2, c2, acc[dx][1]));
        }
    }
}

/* 8x8 partitions (north_star: "motion estimation over 16x16/8x8 partitions"; oracle/ref_enc.c me_l0, second half).
 * Every quadrant takes its best of the 5 x 5 full-pel candidates around the coarse vector by SAD — the strips of the 16x16 search
 * hold the four quadrant SADs of every candidate already (key4: per quadrant the lane's best (cost << 5 | candidate)) — then refines
 * to half and quarter-pel by SATD exactly like the 16x16 search, all four quadrants at once: 8 lanes per quadrant = 2 candidates x
 * four 4x4 blocks (skipped unless the full-pel quadrant costs already beat the full-pel 16x16 cost).
 * (cx8, cy8) = the centre of the 5 x 5 window, full-pel.
 * Returns the P_8x8 cost (sum of the quadrant costs + 8 lambda of mb_type / sub_mb_type overhead) in all lanes. */
__device__ __forceinline__ int me_8x8(const cc_kparams &k, const uint32_t *sc, const uint8_t *tiles, int wx, int wy, int mb, const unsigned key4[4], int cx8, int cy8, unsigned full16, const int pm[2], int lambda, int lane) {
    const int p = lane >> 3, slot = (lane >> 2) & 1, b4 = lane & 3;
    int fi = 0;
    unsigned full8 = (unsigned)(8 * lambda);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const unsigned bq = __reduce_min_sync(FULL, key4[q]);
        full8 += bq >> 5;
        if (q == p) fi = (int)(bq & 31u);
    }
    int qx = 4 * (cx8 + fi % 5 - 2), qy = 4 * (cy8 + fi / 5 - 2);
    /* partitions pay only where parts of the macroblock move differently: unless the four full-pel quadrant costs plus the
     * partition overhead already beat the full-pel 16x16 cost, the sub-pel refinement is not run and the macroblock stays one
     * partition */
    if (full8 >= full16) {
        if (lane < 8) { ((uint32_t *)k.me8[mb].mv)[lane & 3] = 0u; ((uint32_t *)k.me8[mb].satd)[lane & 3] = 0u; }
        return 1 << 30;
    }
    const int ox = (p & 1) * 8 + (b4 & 1) * 4, oy = (p >> 1) * 8 + (b4 >> 1) * 4;
    uint32_t cw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = sc[(oy + r) * 4 + (ox >> 2)];
    unsigned bestkey = 0xffffffffu;
    int bestsatd = 0, bestx = qx, besty = qy;
    for (int step = 2; step >= 1; step--) {
        const int cx = qx, cy = qy;
        const int npass = step == 2 ? 5 : 4;
        for (int ps = 0; ps < npass; ps++) {
            const int o = 2 * ps + slot + (step == 2 ? 0 : 1);
            const bool valid = o <= 8;
            int dx = 0, dy = 0;
            if (o >= 1) { const int i = o <= 4 ? o - 1 : o, i3 = (i * 11) >> 5; dx = (i - 3 * i3 - 1) * step; dy = (i3 - 1) * step; }
            int satd = 0;
            if (valid) {
                uint32_t pw[4];
                if (step == 2) cc_hpel4x4_s(tiles, wx, wy, ox, oy, cx + dx, cy + dy, pw);
                else cc_qpel4x4_s(tiles, wx, wy, ox, oy, cx + dx, cy + dy, pw);
                satd = cc_satd4x4_words(cw, pw);
            }
            satd += __shfl_xor_sync(FULL, satd, 1);
            satd += __shfl_xor_sync(FULL, satd, 2);
            if (valid) {
                const int mvx = cx + dx, mvy = cy + dy;
                const unsigned cost = (unsigned)(satd + lambda * (cc_se_bits(mvx - 4 * pm[0]) + cc_se_bits(mvy - 4 * pm[1])));
                const unsigned kk = (cost << 5) | (unsigned)o;
                if (kk < bestkey) { bestkey = kk; bestsatd = satd; bestx = mvx; besty = mvy; }
            }
        }
        const unsigned other = __shfl_xor_sync(FULL, bestkey, 4);
        const int osatd = __shfl_xor_sync(FULL, bestsatd, 4), oxv = __shfl_xor_sync(FULL, bestx, 4), oyv = __shfl_xor_sync(FULL, besty, 4);
        if (other < bestkey) { bestkey = other; bestsatd = osatd; bestx = oxv; besty = oyv; }
        qx = bestx; qy = besty;
        bestkey &= ~31u;
    }
    if ((lane & 7) == 0) {
        k.me8[mb].mv[p][0] = (int16_t)qx; k.me8[mb].mv[p][1] = (int16_t)qy; k.me8[mb].satd[p] = bestsatd;
    }
    int cost8 = (int)(bestkey >> 5);
    cost8 += __shfl_xor_sync(FULL, cost8, 8);
    cost8 += __shfl_xor_sync(FULL, cost8, 16);
    return cost8 + 8 * lambda;
}

/* Sub-pel stages (north_star: "Hadamard SATD on TMA/shared-memory-staged reference tiles"): once the best full-pel vector is
 * known, the warp stages the 32 x 28 window around it of all four planes (G, b, h, j: 3.5 KB) with four TMA loads
 * (cp.async.bulk.tensor.2d, one mbarrier per warp), and all 17 + 4 x 17 quarter-pel predictors of the 16x16 and 8x8 refinements are
 * built from shared memory — instead of two unaligned global word pairs per predictor row from four planes. */
#ifndef CC_EMU_BUILD
__global__ void __launch_bounds__(ME0_WARPS * 32) k_me_l0(cc_kparams k, const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmB,
                                                          const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmJ) {
#else
__global__ void __launch_bounds__(ME0_WARPS * 32) k_me_l0(cc_kparams k) {
#endif
#if ME0_WARPS > 8
    CC_DYN_SMEM(me_dsm0);
    uint8_t *const me_dsm = (uint8_t *)(((uintptr_t)me_dsm0 + 127) & ~(uintptr_t)127);
    uint8_t (*swin)[4 * SP_PLANE] = (uint8_t (*)[4 * SP_PLANE])me_dsm;
    uint32_t (*scur)[64] = (uint32_t (*)[64])(me_dsm + ME0_WARPS * 4 * SP_PLANE);
#else
    __shared__ __align__(128) uint8_t swin[ME0_WARPS][4 * SP_PLANE];
    __shared__ uint32_t scur[ME0_WARPS][64]; /* current MB, 16 rows x 4 words */
#endif
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int mb = blockIdx.x * ME0_WARPS + wid, nmb = k.mb_w * k.mb_h;
    if (mb >= nmb) return;
    const int mw = k.mb_w, mbx = mb % mw, mby = mb / mw, lambda = c_lambda[k.qp];
    const int px = mbx * 16, py = mby * 16;
    {
        const uint8_t *cur = k.cur[0] + (size_t)py * k.sy + px;
        scur[wid][lane] = *(const uint32_t *)(cur + (size_t)(lane >> 2) * k.sy + 4 * (lane & 3));
        scur[wid][lane + 32] = *(const uint32_t *)(cur + (size_t)(8 + (lane >> 2)) * k.sy + 4 * (lane & 3));
    }
    /* approximate predictor: median of the neighbours' coarse vectors, in full-pel units */
    const cc_me *me = k.me + mb;
    const int L = mbx > 0, T = mby > (mby / k.rps) * k.rps, C = T && mbx + 1 < mw, TL = T && L;
    const int l1x = 2 * me->l1_mv[0], l1y = 2 * me->l1_mv[1];
    /* The window of the four planes (G, b, h, j) is requested as soon as the coarse vector is known, around it: columns from the
     * 16-byte boundary of the plane at or left of l1x - 5, rows from l1y - 5.  48 x 28 samples from there hold everything the search
     * reads when the best full-pel vector lies within +-2 of the coarse one — the 5 x 5 full-pel strips, the sub-pel predictors, the
     * +-2 strips and predictors of the 8x8 quadrants — so the full-pel search no longer goes to global memory either.  When the
     * predictor or the zero vector wins from further away, the window is staged again around the winner. */
    uint8_t *const tiles = swin[wid];
    int wx, wy;
#ifndef CC_EMU_BUILD
    __shared__ __align__(8) unsigned long long mbar[ME0_WARPS];
    const uint32_t mbar_a = (uint32_t)__cvta_generic_to_shared(&mbar[wid]), win_a = (uint32_t)__cvta_generic_to_shared(swin[wid]);
    uint32_t phase = 0;
    if (lane == 0 && k.dbg != 1) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_a) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#endif
    auto stage = [&](int cx, int cy) { /* window around the full-pel vector (cx, cy); origin relative to the macroblock in (wx, wy) */
        wy = cy - SP_ORG;
        wx = cx - SP_ORG - ((px + cx - SP_ORG) & 15); /* the plane's origin is 16-byte aligned and CC_PAD is a multiple of 16: the same boundary in allocation coordinates */
#ifndef CC_EMU_BUILD
        if (k.dbg != 1) {
            const int tx = CC_PAD + px + wx, ty = CC_PAD + py + wy; /* allocation coordinates: the plane starts CC_PAD in */
            __syncwarp(); /* nobody still reads the previous window */
            if (lane == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar_a), "r"(4 * SP_W * SP_H) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(win_a), "l"((unsigned long long)&tmG), "r"(tx), "r"(ty), "r"(mbar_a) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(win_a + SP_PLANE), "l"((unsigned long long)&tmB), "r"(tx), "r"(ty), "r"(mbar_a) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(win_a + 2 * SP_PLANE), "l"((unsigned long long)&tmH), "r"(tx), "r"(ty), "r"(mbar_a) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(win_a + 3 * SP_PLANE), "l"((unsigned long long)&tmJ), "r"(tx), "r"(ty), "r"(mbar_a) : "memory");
            }
            return;
        }
#endif
        { /* the same window by plain loads (the CPU emulation build; CC_DBG_ME=1 on the device) */
            const uint8_t *pl[4] = {k.ref[0], k.hb, k.hh, k.hj};
            __syncwarp();
            for (int p = 0; p < 4; p++)
                for (int i = lane; i < SP_W * SP_H / 4; i += 32) {
                    const int r = i / (SP_W / 4), cw4 = i - r * (SP_W / 4);
                    *(uint32_t *)(swin[wid] + p * SP_PLANE + r * SP_W + 4 * cw4) =
                        *(const uint32_t *)(pl[p] + (ptrdiff_t)(py + wy + r) * k.sy + px + wx + 4 * cw4);
                }
            __syncwarp();
        }
    };
    auto stage_wait = [&]() {
#ifndef CC_EMU_BUILD
        if (k.dbg != 1) {
            uint32_t ok;
            do {
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(mbar_a), "r"(phase) : "memory");
            } while (!ok);
            phase ^= 1u;
        }
#endif
    };
#ifndef CC_EMU_BUILD
    __syncwarp(); /* the mbarrier is initialised before anybody waits on it */
#endif
    stage(l1x - 1, l1y - 1); /* origin l1 - 5 */
    int pm[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
        int a = L ? 2 * me[-1].l1_mv[c] : 0, b = T ? 2 * me[-mw].l1_mv[c] : 0;
        int cc = C ? 2 * me[-mw + 1].l1_mv[c] : (TL ? 2 * me[-mw - 1].l1_mv[c] : 0);
        pm[c] = (!T && !(C || TL) && L) ? a : cc_median3(a, b, cc);
    }
    __syncwarp();
    const uint32_t *sc = scur[wid];
    stage_wait();

    /* full-pel, 27 candidates: the predictor (0), the zero vector (1), 5 x 5 around the coarse vector (2 + 5 dy + dx).
     * lanes 0..19 = (candidate row, four source rows) of the 5 x 5 window, five candidates each (sad_strip5), from the staged
     * window; lanes 20..23 / 24..27 = the predictor / the zero vector as the centre of a strip of their own, from global memory */
    unsigned key = 0xffffffffu;
    unsigned key4[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu}; /* 8x8 quadrants: best full-pel candidate of the window, per lane */
    {
        const int g = lane & 3, dyi = lane >> 2;
        int sx = l1x - 2, syy = l1y + dyi - 2;
        if (lane >= 24) { sx = -2; syy = 0; }
        else if (lane >= 20) { sx = pm[0] - 2; syy = pm[1]; }
        unsigned acc[5][2];
#pragma unroll
        for (int dx = 0; dx < 5; dx++) acc[dx][0] = acc[dx][1] = 0;
        const uint8_t *sp = lane < 20 ? tiles + (syy + 4 * g - wy) * SP_W + sx - wx : k.ref[0] + (ptrdiff_t)(py + syy + 4 * g) * k.sy + px + sx;
        if (lane < 28) sad_strip5(sp, lane < 20 ? SP_W : k.sy, sc + 16 * g, acc);
        const int bits_y = cc_se_bits(4 * (syy - pm[1]));
#pragma unroll
        unsigned kl = 0xffffffffu, kr = 0xffffffffu;
        for (int dx = 0; dx < 5; dx++) {
            /* rows 4 g .. 4 g + 3 of the left / right eight columns; with the lane next door (g ^ 1) that is a quadrant */
            const unsigned sl = acc[dx][0] + __shfl_xor_sync(FULL, acc[dx][0], 1), sr = acc[dx][1] + __shfl_xor_sync(FULL, acc[dx][1], 1);
            unsigned sad = sl + sr;
            sad += __shfl_xor_sync(FULL, sad, 2);
            const unsigned bits = (unsigned)(lambda * (cc_se_bits(4 * (sx + dx - pm[0])) + bits_y));
            const unsigned cost = sad + bits;
            const bool ok = lane < 20 || (lane < 28 && dx == 2);
            const unsigned cand = lane < 20 ? (unsigned)(2 + dyi * 5 + dx) : (lane < 24 ? 0u : 1u);
            if (ok) key = min(key, (cost << 5) | cand);
            kl = min(kl, ((sl + bits) << 5) | (unsigned)(dyi * 5 + dx));
            kr = min(kr, ((sr + bits) << 5) | (unsigned)(dyi * 5 + dx));
        }
        if (lane < 20) {
            if (g < 2) { key4[0] = kl; key4[1] = kr; }
            else { key4[2] = kl; key4[3] = kr; }
        }
    }
    const unsigned bk = __reduce_min_sync(FULL, key);
    const int bl = (int)(bk & 31u);
    /* best full-pel vector: centre of the 8x8 searches */
    const int fbx = bl == 0 ? pm[0] : (bl == 1 ? 0 : l1x + (bl - 2) % 5 - 2), fby = bl == 0 ? pm[1] : (bl == 1 ? 0 : l1y + (bl - 2) / 5 - 2);
    int bx = 4 * fbx, by = 4 * fby;
    if (cc_iabs(fbx - l1x) > 2 || cc_iabs(fby - l1y) > 2) { /* the predictor or the zero vector won from outside the window (warp-uniform) */
        stage(fbx - 1, fby - 1);
        stage_wait();
    }

    /* sub-pel: half-warp = candidate, lane&15 = 4x4 block (raster within the MB) */
    const int blk = lane & 15, half = lane >> 4, ox = (blk & 3) * 4, oy = (blk >> 2) * 4;
    uint32_t cw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = sc[(oy + r) * 4 + (blk & 3)];
    int tc[16]; /* the source block's share of every candidate's SATD */
    cc_satd4x4_src(cw, tc);
    unsigned bestkey = 0xffffffffu; /* cost<<5 | order; this half-warp's best so far */
    int bestsatd = 0, bestx = bx, besty = by;
    for (int step = 2; step >= 1; step--) {
        const int cx = bx, cy = by;
        const int npass = step == 2 ? 5 : 4;
        for (int p = 0; p < npass; p++) {
            /* order o: step 2: 0 = centre, 1..8 = ring; step 1: 1..8 = ring (incumbent holds order 0) */
            int o = 2 * p + half + (step == 2 ? 0 : 1);
            bool valid = o <= 8;
            int dx = 0, dy = 0;
            if (o >= 1) { const int i = o <= 4 ? o - 1 : o, i3 = (i * 11) >> 5; /* i / 3 for i < 9 */ dx = (i - 3 * i3 - 1) * step; dy = (i3 - 1) * step; }
            int satd = 0;
            if (valid) { /* uniform within each half-warp; no collectives inside */
                uint32_t pw[4];
                if (step == 2) cc_hpel4x4_s(tiles, wx, wy, ox, oy, cx + dx, cy + dy, pw); /* the half-pel ring around a full-pel centre */
                else cc_qpel4x4_s(tiles, wx, wy, ox, oy, cx + dx, cy + dy, pw);
                satd = cc_satd4x4_pre(tc, pw);
            }
#pragma unroll
            for (int s = 8; s; s >>= 1) satd += __shfl_xor_sync(FULL, satd, s); /* 16-lane sum (both halves at once) */
            if (valid) {
                int mvx = cx + dx, mvy = cy + dy;
                unsigned cost = (unsigned)(satd + lambda * (cc_se_bits(mvx - 4 * pm[0]) + cc_se_bits(mvy - 4 * pm[1])));
                unsigned kk = (cost << 5) | (unsigned)o;
                if (kk < bestkey) { bestkey = kk; bestsatd = satd; bestx = mvx; besty = mvy; }
            }
        }
        /* merge the two half-warps: lower key wins */
        unsigned other = __shfl_xor_sync(FULL, bestkey, 16);
        int osatd = __shfl_xor_sync(FULL, bestsatd, 16), oxv = __shfl_xor_sync(FULL, bestx, 16), oyv = __shfl_xor_sync(FULL, besty, 16);
        if (other < bestkey) { bestkey = other; bestsatd = osatd; bestx = oxv; besty = oyv; }
        bx = bestx; by = besty;
        bestkey &= ~31u; /* the winner becomes the incumbent (order 0) of the next step */
    }
    const int cost16 = (int)(bestkey >> 5);
    int part = 0;
    if (k.part8) {
        /* the quadrants search the same 5 x 5 window as the macroblock; when the predictor or the zero vector won from outside it,
         * the macroblock stays one partition (oracle/ref_enc.c me_l0) */
        if (cc_iabs(fbx - l1x) > 2 || cc_iabs(fby - l1y) > 2) {
            if (lane < 8) { ((uint32_t *)k.me8[mb].mv)[lane & 3] = 0u; ((uint32_t *)k.me8[mb].satd)[lane & 3] = 0u; }
        } else part = me_8x8(k, sc, tiles, wx, wy, mb, key4, l1x, l1y, bk >> 5, pm, lambda, lane) < cost16;
    }
    if (lane == 0) {
        k.me[mb].mv[0] = (int16_t)bx; k.me[mb].mv[1] = (int16_t)by; k.me[mb].satd = bestsatd; k.me[mb].part = part;
    }
}

extern "C" int cck_me_l1(const cc_kparams *k, cudaStream_t st) {
    const dim3 grid((unsigned)((k->mb_w + L1_MBS - 1) / L1_MBS), (unsigned)k->mb_h);
#ifndef CC_EMU_BUILD
    CUtensorMap tm;
    memcpy(&tm, k->tmap_l1, sizeof tm);
    CC_LAUNCH(k_me_l1, grid, dim3(L1_NTHR), st, *k, tm);
#else
    CC_LAUNCH(k_me_l1, grid, dim3(L1_NTHR), st, *k);
#endif
    return 1;
}
/* Tensor map over the whole allocation of the half-resolution reference plane (pitch sc, CC_PADC border rows / columns around
 * the picture): box = one CTA's window tile.  cuTensorMapEncodeTiled comes from the driver through the runtime's entry-point
 * query, so the library does not link libcuda. */
extern "C" int cck_me_l1_prepare(cc_kparams *k) {
#ifndef CC_EMU_BUILD
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = NULL;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return -2; /* CC_ERR_CUDA */
    static_assert(sizeof(CUtensorMap) <= sizeof(((cc_kparams *)0)->tmap_l1), "tensor map does not fit its slot");
    CUtensorMap tm;
    uint8_t *base = k->l1ref - ((size_t)CC_PADC * k->sc + CC_PADC);
    const cuuint64_t dims[2] = {(cuuint64_t)k->sc, (cuuint64_t)(k->H / 2 + 2 * CC_PADC)};
    const cuuint64_t strides[1] = {(cuuint64_t)k->sc};
    const cuuint32_t box[2] = {L1_PITCHW * 4, L1_ROWS}, es[2] = {1, 1};
    if (((encode_fn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -2; /* CC_ERR_CUDA */
    memcpy(k->tmap_l1, &tm, sizeof tm);
#endif
    k->tmap_l1_ok = 1;
    return 0;
}
extern "C" int cck_me_l0(const cc_kparams *k, cudaStream_t st) {
    int nmb = k->mb_w * k->mb_h;
#ifndef CC_EMU_BUILD
    const CUtensorMap *tm = (const CUtensorMap *)k->tmaps; /* [0], [1]: the two reconstruction planes; [2..4]: b, h, j */
    const CUtensorMap *g = k->ref[0] == k->tmap_plane[0] ? &tm[0] : &tm[1];
#if ME0_WARPS > 8
    const size_t dyn = (size_t)ME0_WARPS * (4 * SP_PLANE + 256) + 128;
    cudaFuncSetAttribute(k_me_l0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    CC_LAUNCH_SMEM(k_me_l0, dim3((unsigned)((nmb + ME0_WARPS - 1) / ME0_WARPS)), dim3(ME0_WARPS * 32), dyn, st, *k, *g, tm[2], tm[3], tm[4]);
#else
    CC_LAUNCH(k_me_l0, dim3((unsigned)((nmb + ME0_WARPS - 1) / ME0_WARPS)), dim3(ME0_WARPS * 32), st, *k, *g, tm[2], tm[3], tm[4]);
#endif
#else
    CC_LAUNCH(k_me_l0, dim3((unsigned)((nmb + ME0_WARPS - 1) / ME0_WARPS)), dim3(ME0_WARPS * 32), st, *k);
#endif
    return 1;
}
/* Tensor maps over the allocations of the five luma-sized planes the sub-pel search reads (pitch sy, CC_PAD border rows / columns):
 * box = one warp's SP_W x SP_H window.  planes[i] = origin (pixel 0,0) of plane i; maps = room for five CUtensorMap (128 B each). */
extern "C" int cck_me_l0_prepare(cc_kparams *k, uint8_t *const planes[5], void *maps) {
    k->tmaps = maps;
    k->tmap_plane[0] = planes[0]; k->tmap_plane[1] = planes[1];
#ifndef CC_EMU_BUILD
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = NULL;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return -2; /* CC_ERR_CUDA */
    const cuuint64_t dims[2] = {(cuuint64_t)k->sy, (cuuint64_t)(k->H + 2 * CC_PAD)};
    const cuuint64_t strides[1] = {(cuuint64_t)k->sy};
    const cuuint32_t box[2] = {SP_W, SP_H}, es[2] = {1, 1};
    for (int i = 0; i < 5; i++) {
        uint8_t *base = planes[i] - ((size_t)CC_PAD * k->sy + CC_PAD);
        if (((encode_fn)fn)((CUtensorMap *)maps + i, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -2;
    }
#endif
    return 0;
}
extern "C" int cck_motion(const cc_kparams *k, cudaStream_t st) { return cck_me_l1(k, st) + cck_me_l0(k, st); }
