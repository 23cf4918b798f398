/*
 * cc_bitstream.c — sequence/picture parameter sets, slice header and Annex-B NAL packaging
 * (ITU-T H.264 7.3.2.1, 7.3.2.2, 7.3.3, 7.4.1, Annex B).  Host C.
 * Stream shape: High profile (profile_idc 100: transform_8x8_mode on, entropy coding CAVLC), one reference frame, POC type 2,
 * slices of whole MB rows,
 * in-loop deblocking confined to the slice (disable_deblocking_filter_idc = 2) so that slices are
 * independent work units for the GPU.  Downstream in the reference pipeline is
 * `h264parse config-interval=-1 ! mpegtsmux` (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6-7),
 * so SPS/PPS are repeated in-band at every IDR.
 */
#include "cc_host.h"
#include <string.h>

/* Table A-1: {level_idc, MaxMBPS, MaxFS, MaxBR}; MaxBR in units of 1250 bit/s for High profile (cpbBrVclFactor, Table A-2).
 * The bitrate the level must cover is the largest one the `bitrate` property can be set to at run time: 30 Mbit/s
 * (/root/reference/src/core/bitrate_control.h:31), not the value the stream starts with. */
static const int cc_levels[][4] = {{30, 40500, 1620, 10000},   {31, 108000, 3600, 14000},  {32, 216000, 5120, 20000},
                                   {40, 245760, 8192, 20000},  {41, 245760, 8192, 50000},  {42, 522240, 8704, 50000},
                                   {50, 589824, 22080, 135000}, {51, 983040, 36864, 240000}, {52, 2073600, 36864, 240000}};
#define CC_LEVEL_MAX_BPS 30000000L
int cc_level_idc_for(long fs, long mbps) {
    for (int i = 0; i < 9; i++)
        if (fs <= cc_levels[i][2] && mbps <= cc_levels[i][1] && CC_LEVEL_MAX_BPS <= 1250L * cc_levels[i][3]) return cc_levels[i][0];
    return 0; /* beyond level 5.2 */
}
static int cc_level_idc(const cc_seq *s) {
    long fs = (long)s->mb_w * s->mb_h, mbps = (fs * s->fps_n + (s->fps_d ? s->fps_d : 1) - 1) / (s->fps_d ? s->fps_d : 1);
    int l = cc_level_idc_for(fs, mbps);
    return l ? l : 52; /* cc_enc_open rejects such streams; kept total for callers that bypass it */
}

size_t cc_write_sps(const cc_seq *s, uint8_t *rbsp, size_t cap) {
    cc_bw b;
    int crop_r = (s->mb_w * 16 - s->w) / 2, crop_b = (s->mb_h * 16 - s->h) / 2;
    cc_bw_init(&b, rbsp, cap);
    cc_bw_put(&b, 8, 100);  /* profile_idc: High (8x8 transform; entropy coding stays CAVLC) */
    cc_bw_put(&b, 8, 0);    /* constraint_set flags */
    cc_bw_put(&b, 8, (uint32_t)cc_level_idc(s));
    cc_bw_ue(&b, 0);        /* seq_parameter_set_id */
    cc_bw_ue(&b, 1);        /* chroma_format_idc: 4:2:0 */
    cc_bw_ue(&b, 0); cc_bw_ue(&b, 0); /* bit_depth_luma/chroma_minus8 */
    cc_bw_put(&b, 1, 0);    /* qpprime_y_zero_transform_bypass_flag */
    cc_bw_put(&b, 1, 0);    /* seq_scaling_matrix_present_flag */
    cc_bw_ue(&b, 0);        /* log2_max_frame_num_minus4 */
    cc_bw_ue(&b, 2);        /* pic_order_cnt_type: output order = decoding order */
    cc_bw_ue(&b, 1);        /* max_num_ref_frames */
    cc_bw_put(&b, 1, 0);    /* gaps_in_frame_num_value_allowed_flag */
    cc_bw_ue(&b, (unsigned)s->mb_w - 1);
    cc_bw_ue(&b, (unsigned)s->mb_h - 1);
    cc_bw_put(&b, 1, 1);    /* frame_mbs_only_flag */
    cc_bw_put(&b, 1, 1);    /* direct_8x8_inference_flag */
    if (crop_r || crop_b) {
        cc_bw_put(&b, 1, 1);
        cc_bw_ue(&b, 0); cc_bw_ue(&b, (unsigned)crop_r); cc_bw_ue(&b, 0); cc_bw_ue(&b, (unsigned)crop_b);
    } else cc_bw_put(&b, 1, 0);
    cc_bw_put(&b, 1, 1);    /* vui_parameters_present_flag */
    cc_bw_put(&b, 4, 0);    /* aspect_ratio, overscan, video_signal_type, chroma_loc: absent */
    cc_bw_put(&b, 1, 1);    /* timing_info_present_flag */
    cc_bw_put(&b, 32, (uint32_t)s->fps_d);
    cc_bw_put(&b, 32, (uint32_t)(2 * s->fps_n));
    cc_bw_put(&b, 1, 1);    /* fixed_frame_rate_flag */
    cc_bw_put(&b, 3, 0);    /* nal_hrd, vcl_hrd, pic_struct_present */
    cc_bw_put(&b, 1, 1);    /* bitstream_restriction_flag */
    cc_bw_put(&b, 1, 1);    /* motion_vectors_over_pic_boundaries_flag */
    cc_bw_ue(&b, 0); cc_bw_ue(&b, 0); /* max_bytes_per_pic_denom, max_bits_per_mb_denom */
    cc_bw_ue(&b, 9); cc_bw_ue(&b, 9); /* log2_max_mv_length_horizontal / vertical */
    cc_bw_ue(&b, 0);        /* max_num_reorder_frames */
    cc_bw_ue(&b, 1);        /* max_dec_frame_buffering */
    size_t n = cc_bw_finish(&b);
    return b.overflow ? 0 : n;
}

size_t cc_write_pps(uint8_t *rbsp, size_t cap) {
    cc_bw b;
    cc_bw_init(&b, rbsp, cap);
    cc_bw_ue(&b, 0); cc_bw_ue(&b, 0); /* pps id, sps id */
    cc_bw_put(&b, 1, 0);    /* entropy_coding_mode_flag: CAVLC */
    cc_bw_put(&b, 1, 0);    /* bottom_field_pic_order_in_frame_present_flag */
    cc_bw_ue(&b, 0);        /* num_slice_groups_minus1 */
    cc_bw_ue(&b, 0); cc_bw_ue(&b, 0); /* num_ref_idx_l0/l1_default_active_minus1 */
    cc_bw_put(&b, 1, 0);    /* weighted_pred_flag */
    cc_bw_put(&b, 2, 0);    /* weighted_bipred_idc */
    cc_bw_se(&b, 0); cc_bw_se(&b, 0); cc_bw_se(&b, 0); /* pic_init_qp-26, pic_init_qs-26, chroma_qp_index_offset */
    cc_bw_put(&b, 1, 1);    /* deblocking_filter_control_present_flag */
    cc_bw_put(&b, 1, 0);    /* constrained_intra_pred_flag */
    cc_bw_put(&b, 1, 0);    /* redundant_pic_cnt_present_flag */
    cc_bw_put(&b, 1, 1);    /* transform_8x8_mode_flag */
    cc_bw_put(&b, 1, 0);    /* pic_scaling_matrix_present_flag */
    cc_bw_se(&b, 0);        /* second_chroma_qp_index_offset */
    size_t n = cc_bw_finish(&b);
    return b.overflow ? 0 : n;
}

void cc_write_slice_header(cc_bw *b, int first_mb, int is_idr, int frame_num, int idr_pic_id, int qp) {
    cc_bw_ue(b, (unsigned)first_mb);
    cc_bw_ue(b, is_idr ? 7 : 5); /* slice_type: all slices of the picture are I / P */
    cc_bw_ue(b, 0);              /* pic_parameter_set_id */
    cc_bw_put(b, 4, (uint32_t)frame_num & 15);
    if (is_idr) {
        cc_bw_ue(b, (unsigned)idr_pic_id);
        cc_bw_put(b, 2, 0);      /* no_output_of_prior_pics_flag, long_term_reference_flag */
    } else {
        cc_bw_put(b, 1, 0);      /* num_ref_idx_active_override_flag */
        cc_bw_put(b, 1, 0);      /* ref_pic_list_modification_flag_l0 */
        cc_bw_put(b, 1, 0);      /* adaptive_ref_pic_marking_mode_flag */
    }
    cc_bw_se(b, qp - 26);        /* slice_qp_delta */
    cc_bw_ue(b, 2);              /* disable_deblocking_filter_idc: filter, but not across slice edges */
    cc_bw_se(b, 0); cc_bw_se(b, 0);
}

size_t cc_write_nal(uint8_t *out, size_t cap, int nal_hdr, const uint8_t *rbsp, size_t n) {
    if (cap < n + n / 2 + 8) return 0;
    size_t k = 0, i = 0;
    int zeros = 0;
    out[k++] = 0; out[k++] = 0; out[k++] = 0; out[k++] = 1; out[k++] = (uint8_t)nal_hdr;
    while (i < n) { /* emulation prevention (7.4.1): 00 00 0x -> 00 00 03 0x for x <= 3 */
        if (!zeros) { /* no zero pending: everything up to the next zero byte is copied as it is (memchr / memcpy run at SIMD speed) */
            const uint8_t *z = (const uint8_t *)memchr(rbsp + i, 0, n - i);
            const size_t j = z ? (size_t)(z - rbsp) : n;
            memcpy(out + k, rbsp + i, j - i);
            k += j - i; i = j;
            if (i >= n) break;
        }
        const uint8_t v = rbsp[i++];
        if (zeros >= 2 && v <= 3) { out[k++] = 3; zeros = 0; }
        out[k++] = v;
        zeros = v ? 0 : zeros + 1;
    }
    return k;
}

/* flat-matrix quantiser multipliers and dead-zone bias in the (bias + |c|) * mf >> 16 form [EXT: x264 cqm/quant];
 * dead zone 21/64 of a step for intra, 11/64 for inter */
void cc_quant_tables(int qp, int intra, uint16_t mf[3], uint16_t bias[3]) {
    static const uint16_t MF[6][3] = {{13107, 5243, 8066}, {11916, 4660, 7490}, {10082, 4194, 6554},
                                      {9362, 3647, 5825},  {8192, 3355, 5243},  {7282, 2893, 4559}};
    int dz = intra ? 21 : 11, sh = qp / 6 - 1;
    for (int c = 0; c < 3; c++) {
        int m = MF[qp % 6][c];
        int j = sh <= 0 ? m << -sh : (m + (1 << (sh - 1))) >> sh;
        int b = ((dz << 10) + (j >> 1)) / j, lim = (1 << 15) / j;
        mf[c] = (uint16_t)j;
        bias[c] = (uint16_t)(b < lim ? b : lim);
    }
}

/* 8x8 transform quantiser (inter): mf[6], bias[6] per position class of H.264 8.5.9 (normAdjust8x8 classes);
 * level = sign * ((|c| + bias) * mf >> 16) [EXT: x264 quant_8x8, flat matrix, inter dead zone 11/64] */
void cc_quant8_tables(int qp, uint16_t mf[6], uint16_t bias[6]) {
    static const uint16_t MF8[6][6] = {{13107, 11428, 20972, 12222, 16777, 15481}, {11916, 10826, 19174, 11058, 14980, 14290},
                                       {10082, 8943, 15978, 9675, 12710, 11985},   {9362, 8228, 14913, 8931, 11984, 11259},
                                       {8192, 7346, 13159, 7740, 10486, 9777},     {7282, 6428, 11570, 6830, 9118, 8640}};
    const int dz = 11, sh = qp / 6;
    for (int c = 0; c < 6; c++) {
        int m = MF8[qp % 6][c];
        int j = sh <= 0 ? m : (m + (1 << (sh - 1))) >> sh;
        int b = ((dz << 10) + (j >> 1)) / j, lim = (1 << 15) / j;
        mf[c] = (uint16_t)j;
        bias[c] = (uint16_t)(b < lim ? b : lim);
    }
}
