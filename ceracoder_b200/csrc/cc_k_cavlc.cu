/*
 * cc_k_cavlc.cu — device-side CAVLC: macroblock_layer() / residual_block_cavlc() of ITU-T H.264
 * 7.3.5 / 9.2 written by the GPU, so that what crosses PCIe is the finished slice RBSP and the host only adds start codes, NAL
 * headers and emulation-prevention bytes (cc_api.c).  The consumers it stays byte-compatible with are
 * `h264parse config-interval=-1 ! queue ! mux.` (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6-7) and the
 * 1316-byte chunker behind the appsink (/root/reference/src/ceracoder.c:297-339).  Same bits as the host writer
 * (cc_cavlc.c, cc_enc_params.host_entropy) and as the oracle (oracle/ref_mb.c ref_mb_write).
 *
 * Entropy coding is serial in the bit position only.  Everything that determines a macroblock's bits is local: its own record
 * and symbol slot, the total_coeff of the blocks left of / above it (nC, 9.2.1) — read from the neighbours' significance
 * masks — and the number of skipped macroblocks before it (mb_skip_run), found by looking back over the type bytes.  So:
 *
 *   k_cavlc_size   one warp per macroblock, lane = residual block (syntax order): bit length of every macroblock.
 *   k_cavlc_write  one CTA per macroblock row: bit offset of the row inside its slice (sum over the rows above it, prefix sum
 *                  inside the row), then the same per-macroblock coder again, this time OR-ing its code words into the slice's
 *                  bit buffer in HBM (RED.OR; lanes of a warp, warps of a CTA and the row CTAs of a slice all write disjoint
 *                  bits of shared words).  The first row of a slice also writes the slice header (7.3.3), the last one
 *                  rbsp_slice_trailing_bits; whichever row CTA of a slice finishes last copies the slice's bytes to mapped host
 *                  memory (16-byte stores over PCIe) and zeroes the bit buffer behind itself for the next frame.
 *
 * HBM-bound integer/bit work, no staging: per frame the kernels read the symbol slots once each (twice for the levels) and
 * write the bitstream once.
 */
#define CC_VLC_TABLE __constant__
#include "cc_dev.cuh"
#include "cc_vlc_tables.h"

struct VlcTab { /* the code tables in shared memory: lanes look up different entries (constant memory would serialise them) */
    uint16_t coeff_token[4][68], coeff_token_cdc[20], total_zeros[15][16], total_zeros_cdc[3][4], run_before[7][16];
    uint8_t cbp_code[2][48];
};
__device__ __forceinline__ void vlc_load(VlcTab *t) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < 4 * 68; i += nt) t->coeff_token[i / 68][i % 68] = cc_vlc_coeff_token[i / 68][i % 68];
    for (int i = tid; i < 20; i += nt) t->coeff_token_cdc[i] = cc_vlc_coeff_token_cdc[0][i];
    for (int i = tid; i < 15 * 16; i += nt) t->total_zeros[i >> 4][i & 15] = cc_vlc_total_zeros[i >> 4][i & 15];
    for (int i = tid; i < 12; i += nt) t->total_zeros_cdc[i >> 2][i & 3] = cc_vlc_total_zeros_cdc[i >> 2][i & 3];
    for (int i = tid; i < 7 * 16; i += nt) t->run_before[i >> 4][i & 15] = cc_vlc_run_before[i >> 4][i & 15];
    for (int i = tid; i < 96; i += nt) t->cbp_code[i / 48][i % 48] = cc_vlc_cbp_code[i / 48][i % 48];
}

/* ---- bit sinks: the coder below is written once and instantiated with each ----------------------------------------- */
struct BitCount {
    unsigned n;
    __device__ __forceinline__ void put(int len, uint32_t) { n += (unsigned)len; }
};
/* MSB-first into 32-bit words (bit 31 of word 0 = first bit of the slice); the buffer is zero where nobody wrote yet */
struct BitRed {
    uint32_t *w;
    unsigned pos;
    __device__ __forceinline__ void put(int len, uint32_t v) { /* 0 <= len <= 32, v < 2^len */
        if (len == 0) return;
        const unsigned long long x = (unsigned long long)v << (64 - (int)(pos & 31u) - len);
        const uint32_t hi = (uint32_t)(x >> 32), lo = (uint32_t)x;
        if (hi) atomicOr(w + (pos >> 5), hi);
        if (lo) atomicOr(w + (pos >> 5) + 1, lo);
        pos += (unsigned)len;
    }
};
template <class W> __device__ __forceinline__ void put_ue(W &w, unsigned v) {
    const unsigned k = v + 1;
    const int n = 31 - __clz((int)k);
    w.put(n, 0);
    w.put(n + 1, k);
}
template <class W> __device__ __forceinline__ void put_se(W &w, int v) { put_ue(w, v > 0 ? 2u * (unsigned)v - 1u : 2u * (unsigned)(-v)); }
template <class W> __device__ __forceinline__ void put_vlc(W &w, uint16_t e) { w.put(e >> 8, e & 0xffu); }

/* slice_header(), 7.3.3 — the same fields as cc_bitstream.c cc_write_slice_header */
template <class W> __device__ __forceinline__ void slice_header(W &w, int first_mb, int is_idr, int frame_num, int idr_pic_id, int qp) {
    put_ue(w, (unsigned)first_mb);
    put_ue(w, is_idr ? 7u : 5u);
    put_ue(w, 0);
    w.put(4, (uint32_t)frame_num & 15u);
    if (is_idr) { put_ue(w, (unsigned)idr_pic_id); w.put(2, 0); }
    else w.put(3, 0);
    put_se(w, qp - 26);
    put_ue(w, 2);
    put_se(w, 0); put_se(w, 0);
}

/* residual_block_cavlc(), 9.2: mask bit i = scan position i holds a non-zero level, lv = those levels in ascending scan order */
template <class W>
__device__ __forceinline__ void cavlc_block(W &w, const VlcTab *t, unsigned mask, const int16_t *lv, int max_coeff, int nC) {
    const int total = __popc(mask);
    const int cls = nC < 2 ? 0 : (nC < 4 ? 1 : (nC < 8 ? 2 : 3));
    if (!total) { put_vlc(w, nC < 0 ? t->coeff_token_cdc[0] : t->coeff_token[cls][0]); return; }
    int t1 = 0;
    unsigned signs = 0;
    while (t1 < 3 && t1 < total) {
        const int v = lv[total - 1 - t1];
        if (v != 1 && v != -1) break;
        signs = (signs << 1) | (unsigned)(v < 0);
        t1++;
    }
    put_vlc(w, nC < 0 ? t->coeff_token_cdc[4 * total + t1] : t->coeff_token[cls][4 * total + t1]);
    w.put(t1, signs);
    int sl = (total > 10 && t1 < 3) ? 1 : 0;
    for (int i = t1; i < total; i++) { /* levels, highest frequency first (9.2.2.1 read backwards) */
        const int v = lv[total - 1 - i], a = v < 0 ? -v : v;
        int code = 2 * a - 2 + (v < 0);
        if (i == t1 && t1 < 3) code -= 2;
        if (sl == 0) {
            if (code < 14) w.put(code + 1, 1);
            else if (code < 30) { w.put(15, 1); w.put(4, (uint32_t)(code - 14)); }
            else { w.put(16, 1); w.put(12, (uint32_t)(code - 30)); }
        } else {
            const int pre = code >> sl;
            if (pre < 15) { w.put(pre + 1, 1); w.put(sl, (uint32_t)(code & ((1 << sl) - 1))); }
            else { w.put(16, 1); w.put(12, (uint32_t)(code - (15 << sl))); }
        }
        if (sl == 0) sl = 1;
        if (a > (3 << (sl - 1)) && sl < 6) sl++;
    }
    const int last = 31 - __clz((int)mask);
    int zeros_left = last + 1 - total;
    if (total < max_coeff) put_vlc(w, max_coeff == 4 ? t->total_zeros_cdc[total - 1][zeros_left] : t->total_zeros[total - 1][zeros_left]);
    unsigned m = mask;
    for (int i = 0; i < total - 1 && zeros_left > 0; i++) {
        const int hi = 31 - __clz((int)m);
        m &= ~(1u << hi);
        const int run = hi - (31 - __clz((int)m)) - 1;
        put_vlc(w, t->run_before[(zeros_left > 7 ? 7 : zeros_left) - 1][run]);
        zeros_left -= run;
    }
}

/* ---- where a block's mask sits in its macroblock's slot -------------------------------------------------------------- */
/* fields of a cc_mbhost record's first word (type | modes << 8 | cbp << 16 | t8 << 24) */
__device__ __forceinline__ int slot_prefix(int type, int t8) { return type == CC_MB_I4 ? 4 : ((type == CC_MB_P16 && (t8 & 2)) ? 8 : 0); }
/* index of block b (0..15 luma in decoding order, 16..23 chroma AC) among the coded blocks of a macroblock, or -1 */
__device__ __forceinline__ int block_rank(int type, int cbp, int b) {
    const int cl = cbp & 15, cc = cbp >> 4, j0 = type == CC_MB_I16;
    if (b < 16) return ((cl >> (b >> 2)) & 1) ? j0 + 4 * __popc((unsigned)cl & ((1u << (b >> 2)) - 1u)) + (b & 3) : -1;
    return cc == 2 ? j0 + 4 * __popc((unsigned)cl) + 2 + (b - 16) : -1;
}
/* total_coeff of block b of macroblock n as its neighbours see it (9.2.1: 0 for P_Skip and for blocks that are not coded) */
__device__ __forceinline__ int nb_total(const cc_kparams &k, int n, uint32_t h, int b) {
    const int type = (int)(h & 255u), cbp = (int)((h >> 16) & 255u), t8 = (int)(h >> 24);
    if (type == CC_MB_PSKIP) return 0;
    const int r = block_rank(type, cbp, b);
    if (r < 0) return 0;
    const uint16_t *slot = (const uint16_t *)(k.rowbuf + (size_t)n * CC_ROW_PAYLOAD_PER_MB);
    return __popc((unsigned)slot[slot_prefix(type, t8) + r]);
}
__device__ __forceinline__ int blk_of_xy(int bx, int by) { return ((by >> 1) << 3) | ((bx >> 1) << 2) | ((by & 1) << 1) | (bx & 1); }

/* which residual block lane j of a macroblock's warp holds (syntax order): kind 0 luma (incl. the Intra16x16 DC block, which sits
 * at position 0,0), 1 chroma DC, 2 chroma AC; (bx, by) in 4x4 units inside the macroblock / the plane pl */
struct LaneBlock { int kind, bx, by, pl, maxc; };
__device__ __forceinline__ LaneBlock lane_block(int i16, int cl, int nl, int lane) {
    LaneBlock b = {0, 0, 0, 0, 16};
    const int j = lane - i16;
    if (i16 && lane == 0) return b;
    if (j < nl) {
        int q = 0, nth = j >> 2; /* the (j >> 2)-th coded 8x8 */
        for (int i = 0; i < 4; i++) { if ((cl >> i) & 1) { if (nth == 0) { q = i; break; } nth--; } }
        const int blk = q * 4 + (j & 3);
        b.maxc = i16 ? 15 : 16;
        b.bx = ((blk >> 2) & 1) * 2 + (blk & 1); b.by = ((blk >> 3) & 1) * 2 + ((blk >> 1) & 1);
    } else if (j < nl + 2) { b.kind = 1; b.maxc = 4; b.pl = j - nl; }
    else { const int i = j - nl - 2; b.kind = 2; b.maxc = 15; b.pl = i >> 2; b.bx = i & 1; b.by = (i >> 1) & 1; }
    return b;
}

/* What k_cavlc_size leaves per macroblock for k_cavlc_write (32 bytes): everything that needs other macroblocks, so that the
 * writer's loads are all independent of each other.  nC[j] = the nC of 9.2.1 for the j-th coded block (-1 = chroma DC) */
struct __align__(16) CvRec { int8_t nC[28]; uint16_t run, slot_words; };

/* macroblock_layer() of one macroblock by one warp, from its record h, its symbol slot (staged in shared memory), the
 * mb_skip_run before it and its lanes' nC.  Returns the bit length in all lanes; WRITE: also ORs the bits into bs at bitpos.
 * Every lane must call it (warp collectives inside). */
template <bool WRITE>
__device__ __forceinline__ unsigned mb_code(const VlcTab *t, const uint16_t *slot, uint2 h, int P, int run, int nC, int lane, uint32_t *bs, unsigned bitpos) {
    const int type = (int)(h.x & 255u), modes = (int)((h.x >> 8) & 255u), cbp = (int)((h.x >> 16) & 255u), t8 = (int)(h.x >> 24);
    const int cl = cbp & 15, cc = cbp >> 4, i16 = type == CC_MB_I16;
    if (type == CC_MB_PSKIP) { /* the slice ends in skipped macroblocks: the last one carries the final mb_skip_run */
        BitCount c = {0};
        put_ue(c, (unsigned)run + 1u);
        if (WRITE && lane == 0) { BitRed w = {bs, bitpos}; put_ue(w, (unsigned)run + 1u); }
        return c.n;
    }
    auto header = [&](auto &w) { /* up to residual(), 7.3.5 */
        if (P) put_ue(w, (unsigned)run);
        if (type == CC_MB_P16 && (t8 & 2)) { /* P_8x8: four sub_mb_type P_L0_8x8 (ue 0 each), no ref_idx (one reference), four mvd */
            put_ue(w, 3);
            w.put(4, 15);
            for (int i = 0; i < 8; i++) put_se(w, (int)(int16_t)slot[i]);
        } else if (type == CC_MB_P16) {
            put_ue(w, 0);
            put_se(w, (int)(int16_t)(h.y & 0xffffu)); put_se(w, (int)(int16_t)(h.y >> 16));
        } else if (type == CC_MB_I4) {
            put_ue(w, P ? 5u : 0u);
            w.put(1, 0); /* transform_size_8x8_flag: the PPS enables transform_8x8_mode, Intra8x8 is not used */
            for (int i = 0; i < 16; i++) { /* nibble: 8 = prev_intra4x4_pred_mode_flag, else rem_intra4x4_pred_mode */
                const unsigned nib = ((unsigned)slot[i >> 2] >> (4 * (i & 3))) & 15u;
                if (nib & 8u) w.put(1, 1); else w.put(4, nib);
            }
        } else put_ue(w, (unsigned)((P ? 5 : 0) + 1 + (modes & 3) + 4 * cc + (cl ? 12 : 0)));
        if (type >= CC_MB_I16) put_ue(w, (unsigned)(modes >> 2) & 3u);
        if (!i16) put_ue(w, t->cbp_code[type == CC_MB_I4 ? 0 : 1][cbp]);
        if (type == CC_MB_P16 && cl) w.put(1, (uint32_t)t8 & 1u);
        if (i16 || cbp) w.put(1, 1); /* mb_qp_delta = 0 */
    };
    BitCount hc = {0};
    header(hc);
    if (WRITE && lane == 0) { BitRed w = {bs, bitpos}; header(w); }
    if (!i16 && !cbp) return hc.n;
    /* residual(): lane j = the j-th coded block in syntax order */
    const int nl = i16 ? (cl ? 16 : 0) : 4 * __popc((unsigned)cl), nb = i16 + nl + (cc ? 2 : 0) + (cc == 2 ? 8 : 0);
    const uint16_t *masks = slot + slot_prefix(type, t8);
    const bool act = lane < nb;
    const unsigned mask = act ? masks[lane] : 0u;
    const int total = __popc(mask);
    int lo = total; /* exclusive prefix of total over the lanes = offset of this block's levels */
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, lo, d); if (lane >= d) lo += v; }
    lo -= total;
    const int maxc = lane_block(i16, cl, nl, lane).maxc;
    const int16_t *lv = (const int16_t *)(masks + nb) + lo;
    BitCount bc = {0};
    if (act) cavlc_block(bc, t, mask, lv, maxc, nC);
    unsigned incl = bc.n; /* inclusive prefix of the blocks' bit lengths */
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
    if (WRITE && act) {
        BitRed w = {bs, bitpos + hc.n + incl - bc.n};
        cavlc_block(w, t, mask, lv, maxc, nC);
    }
    return hc.n + __shfl_sync(FULL, incl, 31);
}

/* a macroblock's symbol slot into the warp's shared memory, coalesced (the coder reads the levels one by one, twice when writing) */
__device__ __forceinline__ void stage_slot(const cc_kparams &k, uint32_t *stage, int mb, unsigned slot_words, int lane) {
    const uint32_t *src = (const uint32_t *)(k.rowbuf + (size_t)mb * CC_ROW_PAYLOAD_PER_MB);
    __syncwarp();
    for (unsigned i = lane; i < slot_words; i += 32) stage[i] = src[i];
    __syncwarp();
}

/* ---- kernels ---------------------------------------------------------------------------------------------------------- */
#define CV_WARPS 8
#define CV_MB_PER_CTA 16
/* one warp per macroblock: what its bits depend on outside itself (mb_skip_run, nC), then its bit length */
__global__ void __launch_bounds__(CV_WARPS * 32) k_cavlc_size(cc_kparams k) {
    __shared__ VlcTab tab;
    __shared__ __align__(16) uint32_t stage_all[CV_WARPS][CC_ROW_PAYLOAD_PER_MB / 4];
    vlc_load(&tab);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nmb = k.mb_w * k.mb_h, mw = k.mb_w, P = !k.is_idr;
    uint32_t *stage = stage_all[wid];
    for (int it = wid; it < CV_MB_PER_CTA; it += CV_WARPS) {
        const int mb = blockIdx.x * CV_MB_PER_CTA + it;
        if (mb >= nmb) break;
        const int mbx = mb % mw, row = mb / mw, row0 = (row / k.rps) * k.rps;
        const int slice_first = row0 * mw, slice_last = min(row0 + k.rps, k.mb_h) * mw - 1;
        const int has_left = mbx > 0, has_top = row > row0;
        /* everything that does not depend on another load goes out first: the record, the neighbours' records, the slot length */
        const uint2 h = *(const uint2 *)(k.mbh + mb);
        const uint32_t h_left = has_left ? *(const uint32_t *)(k.mbh + mb - 1) : 0u, h_top = has_top ? *(const uint32_t *)(k.mbh + mb - mw) : 0u;
        const unsigned slot_words = (unsigned)k.symlen[mb] >> 1;
        const int type = (int)(h.x & 255u), cbp = (int)((h.x >> 16) & 255u);
        if (type == CC_MB_PSKIP && mb != slice_last) { if (lane == 0) k.mbbits[mb] = 0; continue; }
        /* mb_skip_run: the skipped macroblocks straight before this one inside the slice */
        int run = 0;
        if (P) {
            int p = mb - 1;
            bool more = true;
            while (more && p >= slice_first) {
                const int idx = p - lane;
                const int stop = idx < slice_first ? 1 : (*(const uint8_t *)(k.mbh + idx) != CC_MB_PSKIP);
                const unsigned bal = __ballot_sync(FULL, stop);
                if (bal) { run += __ffs((int)bal) - 1; more = false; }
                else { run += 32; p -= 32; }
            }
        }
        int nC = 0;
        if (type != CC_MB_PSKIP) {
            stage_slot(k, stage, mb, slot_words, lane);
            /* total_coeff of the blocks across the macroblock's left and top edges: lanes 0-7 the left neighbour's right column
             * {luma 5, 7, 13, 15, Cb 1, 3, Cr 1, 3}, lanes 8-15 the top neighbour's bottom row {luma 10, 11, 14, 15, Cb 2, 3, Cr 2, 3} */
            int ext = 0;
            if (lane < 16) {
                const int top = lane >> 3, i = lane & 7;
                if (top ? has_top : has_left) {
                    const int b = i < 4 ? (top ? (i < 2 ? 10 + i : 12 + i) : (i < 2 ? 5 + 2 * i : 9 + 2 * i)) : 16 + 4 * ((i - 4) >> 1) + (top ? 2 + (i & 1) : 1 + 2 * (i & 1));
                    ext = nb_total(k, mb - (top ? mw : 1), top ? h_top : h_left, b);
                }
            }
            const int t8 = (int)(h.x >> 24), cl = cbp & 15, cc = cbp >> 4, i16 = type == CC_MB_I16;
            const int nl = i16 ? (cl ? 16 : 0) : 4 * __popc((unsigned)cl), nb = i16 + nl + (cc ? 2 : 0) + (cc == 2 ? 8 : 0);
            const uint16_t *masks = (const uint16_t *)stage + slot_prefix(type, t8);
            const int total = lane < nb ? __popc((unsigned)masks[lane]) : 0;
            const LaneBlock lb = lane_block(i16, cl, nl, lane);
            int srcA, srcB, extA, extB; /* lane holding the in-macroblock neighbour (or -1), lane holding the outside one */
            if (lb.kind == 2) {
                srcA = lb.bx > 0 ? block_rank(type, cbp, 16 + 4 * lb.pl + lb.by * 2) : -1;
                srcB = lb.by > 0 ? block_rank(type, cbp, 16 + 4 * lb.pl + lb.bx) : -1;
                extA = 4 + 2 * lb.pl + lb.by; extB = 12 + 2 * lb.pl + lb.bx;
            } else {
                srcA = lb.bx > 0 ? block_rank(type, cbp, blk_of_xy(lb.bx - 1, lb.by)) : -1;
                srcB = lb.by > 0 ? block_rank(type, cbp, blk_of_xy(lb.bx, lb.by - 1)) : -1;
                extA = lb.by; extB = 8 + lb.bx;
            }
            const int inA = __shfl_sync(FULL, total, srcA & 31), inB = __shfl_sync(FULL, total, srcB & 31);
            const int outA = __shfl_sync(FULL, ext, extA), outB = __shfl_sync(FULL, ext, extB);
            const int availA = lb.bx > 0 || has_left, availB = lb.by > 0 || has_top;
            const int nA = lb.bx > 0 ? (srcA >= 0 ? inA : 0) : outA, nB = lb.by > 0 ? (srcB >= 0 ? inB : 0) : outB;
            nC = (availA && availB) ? (nA + nB + 1) >> 1 : (availA ? nA : (availB ? nB : 0)); /* 9.2.1 */
            if (lb.kind == 1) nC = -1;
        }
        const unsigned n = mb_code<false>(&tab, (const uint16_t *)stage, h, P, run, nC, lane, nullptr, 0);
        CvRec *rec = k.cvrec + mb;
        if (lane < 28) rec->nC[lane] = (int8_t)nC;
        else if (lane == 28) rec->run = (uint16_t)run;
        else if (lane == 29) rec->slot_words = (uint16_t)slot_words;
        else if (lane == 30) k.mbbits[mb] = n;
    }
}

/* layout of the mapped host buffer in device-entropy mode: [cc_rowhdr x mb_h][u32 slice bytes x ns][slice RBSPs, bs_slice_words
 * words apart] (cc_api.c collect reads the same) */
#define CW_WARPS 16
__global__ void __launch_bounds__(CW_WARPS * 32) k_cavlc_write(cc_kparams k) {
    __shared__ VlcTab tab;
    __shared__ __align__(16) uint32_t stage_all[CW_WARPS][CC_ROW_PAYLOAD_PER_MB / 4];
    __shared__ unsigned red[CW_WARPS], mboff[513], s_last, n_work;
    __shared__ uint16_t work[512]; /* the row's macroblocks that have bits, in order */
    vlc_load(&tab);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, mw = k.mb_w, P = !k.is_idr;
    const int row = blockIdx.x, s = row / k.rps, row0 = s * k.rps, rows = min(k.rps, k.mb_h - row0), ns = (k.mb_h + k.rps - 1) / k.rps;
    uint32_t *bs = k.bs + (size_t)s * k.bs_slice_words;
    uint32_t *stage = stage_all[wid];
    { /* bits of the slice's rows above this one */
        unsigned v = 0;
        const uint32_t *src = k.mbbits + (size_t)row0 * mw;
        for (int i = tid; i < (row - row0) * mw; i += CW_WARPS * 32) v += src[i];
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
        if (lane == 0) red[wid] = v;
    }
    if (wid == 0) { /* exclusive prefix of this row's macroblock lengths (mboff[mw] = the row's total) and the list of coded macroblocks */
        unsigned run = 0, nw = 0;
        for (int base = 0; base < mw; base += 32) {
            const int i = base + lane;
            const unsigned len = i < mw ? k.mbbits[(size_t)row * mw + i] : 0u;
            unsigned inc = len;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const unsigned v = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += v; }
            if (i < mw) mboff[i] = run + inc - len;
            run += __shfl_sync(FULL, inc, 31);
            const unsigned bal = __ballot_sync(FULL, len != 0u);
            if (len) work[nw + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
            nw += (unsigned)__popc(bal);
        }
        if (lane == 0) { mboff[mw] = run; n_work = nw; }
    }
    __syncthreads();
    BitCount hc = {0};
    slice_header(hc, row0 * mw, k.is_idr, k.frame_num, k.idr_pic_id, k.qp);
    unsigned base = hc.n;
#pragma unroll
    for (int i = 0; i < CW_WARPS; i++) base += red[i];
    for (unsigned j = wid; j < n_work; j += CW_WARPS) {
        const int x = work[j], mb = row * mw + x;
        /* three independent loads: the record, what k_cavlc_size left, the head of the slot (the rest only if the slot is longer) */
        const uint2 h = *(const uint2 *)(k.mbh + mb);
        const uint32_t rw = ((const uint32_t *)(k.cvrec + mb))[lane & 7];
        const uint32_t *src = (const uint32_t *)(k.rowbuf + (size_t)mb * CC_ROW_PAYLOAD_PER_MB);
        __syncwarp();
        stage[lane] = src[lane];
        const uint32_t tail = __shfl_sync(FULL, rw, 7);
        const unsigned slot_words = tail >> 16;
        for (unsigned i = 32 + lane; i < slot_words; i += 32) stage[i] = src[i];
        __syncwarp();
        const int nC = (int)(int8_t)(__shfl_sync(FULL, rw, lane >> 2) >> (8 * (lane & 3)));
        mb_code<true>(&tab, (const uint16_t *)stage, h, P, (int)(tail & 0xffffu), nC, lane, bs, base + mboff[x]);
    }
    if (tid == 0) {
        if (row == row0) { BitRed w = {bs, 0}; slice_header(w, row0 * mw, k.is_idr, k.frame_num, k.idr_pic_id, k.qp); }
        if (row == row0 + rows - 1) { /* rbsp_slice_trailing_bits: a one, then zeros (already there) up to the byte boundary */
            BitRed w = {bs, base + mboff[mw]};
            w.put(1, 1);
            k.bs_ctl[ns + s] = (w.pos + 7u) >> 3;
        }
    }
    if (tid < (int)(sizeof(cc_rowhdr) / 4)) ((uint32_t *)k.out_host_dev)[row * (sizeof(cc_rowhdr) / 4) + tid] = ((const uint32_t *)(k.rowhdr + row))[tid];
    /* whichever row CTA of the slice gets here last hands the finished slice to the host */
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(k.bs_ctl + s, 1u) == (unsigned)rows - 1u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const unsigned bytes = __ldcg(k.bs_ctl + ns + s), nq = (bytes + 15u) >> 4;
    const unsigned hdr_bytes = (unsigned)((sizeof(cc_rowhdr) * k.mb_h + 15) / 16 * 16), len_bytes = (unsigned)((4 * ns + 15) / 16 * 16);
    uint4 *dst = (uint4 *)(k.out_host_dev + hdr_bytes + len_bytes + (size_t)s * k.bs_slice_words * 4);
    uint4 *src = (uint4 *)bs;
    for (unsigned i = tid; i < nq; i += CW_WARPS * 32) {
        uint4 v = __ldcg(src + i);
        src[i] = make_uint4(0, 0, 0, 0);
        v.x = __byte_perm(v.x, 0, 0x0123); v.y = __byte_perm(v.y, 0, 0x0123); v.z = __byte_perm(v.z, 0, 0x0123); v.w = __byte_perm(v.w, 0, 0x0123);
        dst[i] = v;
    }
    if (tid == 0) {
        ((uint32_t *)(k.out_host_dev + hdr_bytes))[s] = bytes;
        k.bs_ctl[s] = 0;
    }
}

extern "C" int cck_cavlc(const cc_kparams *k, cudaStream_t st) {
    const int nmb = k->mb_w * k->mb_h;
    CC_LAUNCH(k_cavlc_size, dim3((unsigned)((nmb + CV_MB_PER_CTA - 1) / CV_MB_PER_CTA)), dim3(CV_WARPS * 32), st, *k);
    CC_LAUNCH(k_cavlc_write, dim3((unsigned)k->mb_h), dim3(CW_WARPS * 32), st, *k);
    return 2;
}
