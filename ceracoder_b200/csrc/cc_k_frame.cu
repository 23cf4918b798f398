/*
 * cc_k_frame.cu — frame-level, fully parallel kernels (SURVEY.md §8a rows K1 and the reference-frame
 * preparation that x264 calls frame_expand_border / hpel_filter [EXT]):
 *   k_ingest      pad the uploaded picture to a macroblock multiple, de-interleave NV12 chroma,
 *                 build the half-resolution luma plane of the current picture
 *   k_border      replicate the borders of a finished reconstruction (unrestricted motion vectors)
 *   k_hpel        half-pel planes b (horizontal), h (vertical), j (centre) of the reconstruction with the
 *                 6-tap filter of H.264 8.4.2.2.1; every quarter-pel predictor is then the average of
 *                 two samples of {G,b,h,j}
 *   k_l1border    border of the half-resolution reference (the plane itself is written with the reconstruction, cc_k_mb.cu)
 * All are HBM/L2 streaming kernels: one pass, 4-byte or wider accesses, grid >> 148 SMs.
 * Oracle counterparts: oracle/ref_enc.c load_plane / extend_border / downsample2, ref_prims.c hpel_*.
 */
#include "cc_dev.cuh"

/* ---- k_ingest --------------------------------------------------------------------------------------- */
/* one thread = one half-res pixel = 2x2 luma pixels + one chroma sample position */
__global__ void k_ingest(cc_kparams k, int nv12) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    int W2 = k.W >> 1, w2 = k.w >> 1, h2 = k.h >> 1;
    if (x >= W2) return;
    /* luma: clamp source coordinates to the displayed picture; write back only in the padding */
    int sx0 = min(2 * x, k.w - 1), sx1 = min(2 * x + 1, k.w - 1);
    int sy0 = min(2 * y, k.h - 1), sy1 = min(2 * y + 1, k.h - 1);
    uint8_t *Y = k.cur[0];
    int a = Y[(size_t)sy0 * k.sy + sx0], b = Y[(size_t)sy0 * k.sy + sx1];
    int c = Y[(size_t)sy1 * k.sy + sx0], d = Y[(size_t)sy1 * k.sy + sx1];
    if (2 * x + 1 >= k.w || 2 * y + 1 >= k.h) {
        Y[(size_t)(2 * y) * k.sy + 2 * x] = (uint8_t)a; Y[(size_t)(2 * y) * k.sy + 2 * x + 1] = (uint8_t)b;
        Y[(size_t)(2 * y + 1) * k.sy + 2 * x] = (uint8_t)c; Y[(size_t)(2 * y + 1) * k.sy + 2 * x + 1] = (uint8_t)d;
    }
    k.l1cur[(size_t)y * k.sc + x] = (uint8_t)((a + b + c + d + 2) >> 2);
    /* chroma */
    int cx = min(x, w2 - 1), cy = min(y, h2 - 1);
    if (nv12) {
        const uint8_t *uv = k.uvstage + (size_t)cy * k.sy + 2 * cx;
        k.cur[1][(size_t)y * k.sc + x] = uv[0];
        k.cur[2][(size_t)y * k.sc + x] = uv[1];
    } else if (x >= w2 || y >= h2) {
        k.cur[1][(size_t)y * k.sc + x] = k.cur[1][(size_t)cy * k.sc + cx];
        k.cur[2][(size_t)y * k.sc + x] = k.cur[2][(size_t)cy * k.sc + cx];
    }
}


/* ---- k_ingest_packed: colour conversion fused into the ingest --------------------------------- */
/* Packed input (YUY2 or BGRx, uploaded as is into k.pkstage) -> padded 4:2:0 planes + half-resolution luma in one pass: the
 * `videoconvert` in front of the encoder in the reference's pipeline text.  One thread = one 2x2 luma block; interior blocks
 * read one 32-bit (YUY2) or 64-bit (BGRx) word per row, blocks in the padding clamp their source coordinates.
 * Arithmetic: oracle/ref_convert.c (bit-exact, tests/test_convert.py). */
__constant__ int c_rgb2yuv[2][9] = {
    {66, 129, 25, -38, -74, 112, 112, -94, -18},   /* BT.601, pictures of up to 576 lines */
    {47, 157, 16, -26, -87, 112, 112, -102, -10},  /* BT.709 */
};
template <int FMT> /* 2 = YUY2, 3 = BGRx */
__global__ void k_ingest_packed(cc_kparams k) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= (k.W >> 1)) return;
    const int sx0 = min(2 * x, k.w - 2), sy0 = min(2 * y, k.h - 2); /* w, h are even: the block clamps as a whole ... */
    const bool edge = 2 * x >= k.w || 2 * y >= k.h;              /* ... and replicates the last column / row in the padding */
    const uint8_t *r0 = k.pkstage + (size_t)sy0 * k.pks, *r1 = r0 + k.pks;
    int Y[4], U, V;
    if (FMT == 2) {
        const uint32_t a = *(const uint32_t *)(r0 + 2 * sx0), b = *(const uint32_t *)(r1 + 2 * sx0);
        Y[0] = a & 255; Y[1] = (a >> 16) & 255; Y[2] = b & 255; Y[3] = (b >> 16) & 255;
        U = (int)(((a >> 8) & 255) + ((b >> 8) & 255) + 1) >> 1;
        V = (int)((a >> 24) + (b >> 24) + 1) >> 1;
    } else {
        const int *m = c_rgb2yuv[k.h > 576];
        const uint2 a = *(const uint2 *)(r0 + 4 * sx0), b = *(const uint2 *)(r1 + 4 * sx0);
        const uint32_t px[4] = {a.x, a.y, b.x, b.y};
        int sr = 0, sg = 0, sb = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int B = px[i] & 255, G = (px[i] >> 8) & 255, R = (px[i] >> 16) & 255;
            Y[i] = (m[0] * R + m[1] * G + m[2] * B + 4224) >> 8;
            sr += R; sg += G; sb += B;
        }
        const int R = (sr + 2) >> 2, G = (sg + 2) >> 2, B = (sb + 2) >> 2;
        U = (m[3] * R + m[4] * G + m[5] * B + 32896) >> 8;
        V = (m[6] * R + m[7] * G + m[8] * B + 32896) >> 8;
    }
    if (edge) { /* padding: replicate the picture's last column / row (what k_ingest does for planar input) */
        if (2 * x >= k.w) { Y[0] = Y[1]; Y[2] = Y[3]; }
        if (2 * y >= k.h) { Y[0] = Y[2]; Y[1] = Y[3]; }
        if (2 * x >= k.w && 2 * y >= k.h) Y[0] = Y[1] = Y[2] = Y[3];
    }
    uint8_t *dy = k.cur[0] + (size_t)(2 * y) * k.sy + 2 * x;
    *(uint16_t *)dy = (uint16_t)(Y[0] | (Y[1] << 8));
    *(uint16_t *)(dy + k.sy) = (uint16_t)(Y[2] | (Y[3] << 8));
    k.l1cur[(size_t)y * k.sc + x] = (uint8_t)((Y[0] + Y[1] + Y[2] + Y[3] + 2) >> 2);
    k.cur[1][(size_t)y * k.sc + x] = (uint8_t)U;
    k.cur[2][(size_t)y * k.sc + x] = (uint8_t)V;
}

/* ---- k_border ---------------------------------------------------------------------------------------- */
/* grid.y = plane (0 luma, 1 Cb, 2 Cr); one thread per border byte-quad */
/* blk of nblk blocks of plane pl (256 threads each, grid-stride) */
__device__ __forceinline__ void border_work(const cc_kparams &k, int pl, int blk, int nblk) {
    int W = pl ? k.W >> 1 : k.W, H = pl ? k.H >> 1 : k.H, pad = pl ? CC_PADC : CC_PAD, s = pl ? k.sc : k.sy;
    uint8_t *P = k.rec[pl];
    int fullw = W + 2 * pad;
    /* work items: rows -pad..H+pad-1, for each row the 4-byte groups that lie in the border */
    int groups_lr = (2 * pad) >> 2;   /* left+right groups in a picture row */
    int groups_full = fullw >> 2;     /* groups of a row entirely in the top/bottom border */
    int n_mid = H * groups_lr, n_tb = 2 * pad * groups_full;
    for (int i = blk * 256 + threadIdx.x; i < n_mid + n_tb; i += nblk * 256) {
        int y, x;
        if (i < n_mid) {
            y = i / groups_lr;
            int g = i % groups_lr;
            x = g < (pad >> 2) ? -pad + 4 * g : W + 4 * (g - (pad >> 2));
        } else {
            int j = i - n_mid, r = j / groups_full;
            y = r < pad ? r - pad : H + (r - pad);
            x = -pad + 4 * (j % groups_full);
        }
        int sy_ = min(max(y, 0), H - 1);
        const uint8_t *row = P + (size_t)sy_ * s;
        uint32_t v;
        if (x < 0) v = row[0] * 0x01010101u;
        else if (x >= W) v = row[W - 1] * 0x01010101u;
        else v = *(const uint32_t *)(row + x);
        *(uint32_t *)(P + (ptrdiff_t)y * s + x) = v;
    }
}
__global__ void __launch_bounds__(256) k_border(cc_kparams k) { border_work(k, blockIdx.y, blockIdx.x, gridDim.x); }

/* ---- k_hpel ------------------------------------------------------------------------------------------ */
#define HP_TW 64
#define HP_TH 16
__device__ __forceinline__ int cc_tap6(int a, int b, int c, int d, int e, int f) { return a - 5 * b + 20 * c + 20 * d - 5 * e + f; }

/* CTA = 256 threads, output tile 64x16; reads the border-extended reconstruction (run after k_border).
 * The 6-tap filter (1, -5, 20, 20, -5, 1) runs as dot products on packed samples: horizontally the six bytes of an output sit in
 * two (funnel-shifted) words -> two IDP.4A; vertically a 4x4 byte transpose (PRMT) puts a column's first four rows into one word ->
 * two IDP.4A per output; the centre plane filters the 16-bit horizontal sums vertically with three IDP.2A per output.
 * thread = four horizontally adjacent outputs of each plane. */
#define HP_A 0x1414fb01u  /* ( 1, -5, 20, 20) as signed bytes, first tap in the low byte */
#define HP_B0 0x01fb0000u /* ( 0,  0, -5,  1) */
#define HP_B1 0x000001fbu /* (-5,  1,  0,  0) */
#define HP_B3 0x0001fb00u /* ( 0, -5,  1,  0) */
__device__ __forceinline__ void hp_row4(const uint32_t *gw, int t[4]) { /* gw: the aligned word left of the four outputs, then theirs, then the next */
    const uint32_t w0 = gw[0], w1 = gw[1], w2 = gw[2];
    t[0] = cc_dp4a_us(w1, HP_B0, cc_dp4a_us(__funnelshift_r(w0, w1, 16), HP_A, 0));
    t[1] = cc_dp4a_us(__funnelshift_r(w1, w2, 24), HP_B1, cc_dp4a_us(__funnelshift_r(w0, w1, 24), HP_A, 0));
    t[2] = cc_dp4a_us(w2, HP_B1, cc_dp4a_us(w1, HP_A, 0));
    t[3] = cc_dp4a_us(w2, HP_B3, cc_dp4a_us(__funnelshift_r(w1, w2, 8), HP_A, 0));
}
__device__ __forceinline__ void hpel_tile(const cc_kparams &k, int bx, int by) {
    __shared__ uint32_t g[HP_TH + 5][18];        /* rows y-2..y+18, bytes x-4..x+67 */
    __shared__ uint32_t t2[HP_TH + 5][HP_TW / 2]; /* unrounded horizontal 6-tap of every staged row, two int16 per word */
    const int x0 = -CC_HPAD + bx * HP_TW, y0 = -CC_HPAD + by * HP_TH;
    const uint8_t *G = k.rec[0];
    /* samples outside the picture are the nearest picture sample (what k_border replicates into the plane's border): the reads are
     * clamped to the picture, so this kernel does not wait for k_border */
    for (int i = threadIdx.x; i < (HP_TH + 5) * 18; i += 256) { /* 18 words per staged row */
        int r = i / 18, c = i % 18;
        int y = min(max(y0 - 2 + r, 0), k.H - 1), x = x0 - 4 + 4 * c; /* staged cols start at x0-4 (aligned); col index = x-(x0-4) */
        uint32_t v = 0;
        if (x >= 0 && x + 3 <= k.W - 1) v = *(const uint32_t *)(G + (ptrdiff_t)y * k.sy + x);
        else for (int b = 0; b < 4; b++) v |= (uint32_t)G[(ptrdiff_t)y * k.sy + min(max(x + b, 0), k.W - 1)] << (8 * b);
        g[r][c] = v;
    }
    __syncthreads();
    /* g[r] byte cx holds sample (x0-4+cx, y0-2+r) */
    const int r = threadIdx.x >> 4, gi = threadIdx.x & 15;
    uint32_t vb;
    { /* horizontal sums of the thread's own output row (staged row r + 2): plane b comes straight from them */
        int t[4];
        hp_row4(&g[r + 2][gi], t);
        t2[r + 2][2 * gi] = ((uint32_t)t[0] & 0xffffu) | ((uint32_t)t[1] << 16);
        t2[r + 2][2 * gi + 1] = ((uint32_t)t[2] & 0xffffu) | ((uint32_t)t[3] << 16);
        vb = (uint32_t)cc_clip255((t[0] + 16) >> 5) | ((uint32_t)cc_clip255((t[1] + 16) >> 5) << 8) |
             ((uint32_t)cc_clip255((t[2] + 16) >> 5) << 16) | ((uint32_t)cc_clip255((t[3] + 16) >> 5) << 24);
    }
    if (threadIdx.x < 80) { /* the five staged rows above / below the tile's own sixteen */
        const int q = threadIdx.x >> 4, rr = q < 2 ? q : 16 + q;
        int t[4];
        hp_row4(&g[rr][gi], t);
        t2[rr][2 * gi] = ((uint32_t)t[0] & 0xffffu) | ((uint32_t)t[1] << 16);
        t2[rr][2 * gi + 1] = ((uint32_t)t[2] & 0xffffu) | ((uint32_t)t[3] << 16);
    }
    __syncthreads();
    const int y = y0 + r, x = x0 + 4 * gi;
    if (x >= k.W + CC_HPAD || y >= k.H + CC_HPAD) return;
    uint32_t vh, vj;
    { /* plane h: vertical filter of the samples; rows r .. r + 5 of the thread's own word */
        const uint32_t R0 = g[r][gi + 1], R1 = g[r + 1][gi + 1], R2 = g[r + 2][gi + 1], R3 = g[r + 3][gi + 1], R4 = g[r + 4][gi + 1], R5 = g[r + 5][gi + 1];
        const uint32_t a0 = __byte_perm(R0, R1, 0x5140), a1 = __byte_perm(R2, R3, 0x5140), a2 = __byte_perm(R0, R1, 0x7362), a3 = __byte_perm(R2, R3, 0x7362);
        const uint32_t u0 = __byte_perm(R4, R5, 0x5140), u1 = __byte_perm(R4, R5, 0x7362);
        const int h0 = cc_dp4a_us(u0, HP_B1, cc_dp4a_us(__byte_perm(a0, a1, 0x5410), HP_A, 16));
        const int h1 = cc_dp4a_us(u0, HP_B0, cc_dp4a_us(__byte_perm(a0, a1, 0x7632), HP_A, 16));
        const int h2 = cc_dp4a_us(u1, HP_B1, cc_dp4a_us(__byte_perm(a2, a3, 0x5410), HP_A, 16));
        const int h3 = cc_dp4a_us(u1, HP_B0, cc_dp4a_us(__byte_perm(a2, a3, 0x7632), HP_A, 16));
        vh = (uint32_t)cc_clip255(h0 >> 5) | ((uint32_t)cc_clip255(h1 >> 5) << 8) | ((uint32_t)cc_clip255(h2 >> 5) << 16) | ((uint32_t)cc_clip255(h3 >> 5) << 24);
    }
    { /* plane j: vertical filter of the horizontal sums */
        int j[4];
#pragma unroll
        for (int w = 0; w < 2; w++) {
            const uint32_t T0 = t2[r][2 * gi + w], T1 = t2[r + 1][2 * gi + w], T2 = t2[r + 2][2 * gi + w], T3 = t2[r + 3][2 * gi + w], T4 = t2[r + 4][2 * gi + w], T5 = t2[r + 5][2 * gi + w];
            j[2 * w] = cc_dp2a_lo(__byte_perm(T4, T5, 0x5410), HP_B1, cc_dp2a_hi(__byte_perm(T2, T3, 0x5410), HP_A, cc_dp2a_lo(__byte_perm(T0, T1, 0x5410), HP_A, 512)));
            j[2 * w + 1] = cc_dp2a_lo(__byte_perm(T4, T5, 0x7632), HP_B1, cc_dp2a_hi(__byte_perm(T2, T3, 0x7632), HP_A, cc_dp2a_lo(__byte_perm(T0, T1, 0x7632), HP_A, 512)));
        }
        vj = (uint32_t)cc_clip255(j[0] >> 10) | ((uint32_t)cc_clip255(j[1] >> 10) << 8) | ((uint32_t)cc_clip255(j[2] >> 10) << 16) | ((uint32_t)cc_clip255(j[3] >> 10) << 24);
    }
    const ptrdiff_t o = (ptrdiff_t)y * k.sy + x;
    *(uint32_t *)(k.hb + o) = vb;
    *(uint32_t *)(k.hh + o) = vh;
    *(uint32_t *)(k.hj + o) = vj;
}
__global__ void __launch_bounds__(256) k_hpel(cc_kparams k) { hpel_tile(k, (int)blockIdx.x, (int)blockIdx.y); }
/* What finishes a reconstruction into a reference in ONE launch: the half-pel tiles (which read the picture only) and, in the CTAs
 * behind them, the border replication of the three planes — nothing in either depends on the other, and the in-loop filter's tail
 * is on every stream's critical cycle (a launch less is 0.03 ms of it).  gx x gy tiles, then HB_BLOCKS border CTAs per plane. */
#define HB_BLOCKS 48
__global__ void __launch_bounds__(256) k_hpel_border(cc_kparams k, int gx, int ntiles) {
    const int id = (int)blockIdx.x;
    if (id < ntiles) hpel_tile(k, id % gx, id / gx);
    else border_work(k, (id - ntiles) / HB_BLOCKS, (id - ntiles) % HB_BLOCKS, HB_BLOCKS);
}

/* ---- k_l1border -------------------------------------------------------------------------------------- */
/* The half-resolution reference is written macroblock by macroblock where the reconstruction is published (cc_k_mb.cu
 * publish_pixels); this replicates its edge samples into the 24-pixel border the coarse search windows reach into.
 * One thread = one border sample. */
__global__ void k_l1border(cc_kparams k) {
    const int B = 24, W2 = k.W >> 1, H2 = k.H >> 1;
    const int nside = 2 * B * H2, ntb = 2 * B * (W2 + 2 * B);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int x, y;
    if (i < nside) { y = i / (2 * B); const int j = i - y * 2 * B; x = j < B ? j - B : W2 + j - B; }        /* left / right strips */
    else if (i < nside + ntb) { const int j = i - nside, w = W2 + 2 * B; const int r = j / w; x = j - r * w - B; y = r < B ? r - B : H2 + r - B; } /* top / bottom */
    else return;
    const int sx = x < 0 ? 0 : (x >= W2 ? W2 - 1 : x), sy = y < 0 ? 0 : (y >= H2 ? H2 - 1 : y);
    k.l1ref[(ptrdiff_t)y * k.sc + x] = k.l1ref[(ptrdiff_t)sy * k.sc + sx];
}

/* ---- launch wrappers --------------------------------------------------------------------------------- */
extern "C" int cck_ingest(const cc_kparams *k, int fmt, cudaStream_t st) { /* fmt: CC_FMT_* of include/ceracuda.h */
    dim3 grid((unsigned)((k->W / 2 + 127) / 128), (unsigned)(k->H / 2));
    if (fmt == 2) CC_LAUNCH(k_ingest_packed<2>, grid, dim3(128), st, *k);
    else if (fmt == 3) CC_LAUNCH(k_ingest_packed<3>, grid, dim3(128), st, *k);
    else CC_LAUNCH(k_ingest, grid, dim3(128), st, *k, fmt == 1);
    return 1;
}
extern "C" int cck_border(const cc_kparams *k, cudaStream_t st) {
    CC_LAUNCH(k_border, dim3(148, 3), dim3(256), st, *k);
    return 1;
}
extern "C" int cck_hpel(const cc_kparams *k, cudaStream_t st) {
    dim3 gh((unsigned)((k->W + 2 * CC_HPAD + HP_TW - 1) / HP_TW), (unsigned)((k->H + 2 * CC_HPAD + HP_TH - 1) / HP_TH));
    CC_LAUNCH(k_hpel, gh, dim3(256), st, *k);
    return 1;
}
extern "C" int cck_hpel_border(const cc_kparams *k, cudaStream_t st) {
    const int gx = (k->W + 2 * CC_HPAD + HP_TW - 1) / HP_TW, gy = (k->H + 2 * CC_HPAD + HP_TH - 1) / HP_TH;
    CC_LAUNCH(k_hpel_border, dim3((unsigned)(gx * gy + 3 * HB_BLOCKS)), dim3(256), st, *k, gx, gx * gy);
    return 1;
}
extern "C" int cck_l1border(const cc_kparams *k, cudaStream_t st) {
    const int n = 48 * (k->H / 2) + 48 * (k->W / 2 + 48);
    CC_LAUNCH(k_l1border, dim3((unsigned)((n + 255) / 256)), dim3(256), st, *k);
    return 1;
}
extern "C" int cck_finish_ref(const cc_kparams *k, cudaStream_t st) { return cck_border(k, st) + cck_hpel(k, st); }
