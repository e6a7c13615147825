// Few-channel first-layer ("stem") convolution (ResNet-50's 7x7 s2 on 3 channels, VGG16's 3x3 s1, LeNet's 5x5 s1: any
// stride_w <= 8 with k_w <= 7) on the TMA + tcgen05 path. The text below describes the stride-2 case; for stride s the
// "pixel" pitch is 4*s floats.
//
// With C_in <= 4 a K-chunk of 32 values spans 8 filter taps, which im2col-mode TMA (one tap per instruction, >= 32
// channels) cannot deliver. But in NHWC4 memory the taps of one filter ROW are contiguous: for output column ow and
// filter row kh, pixels 2ow-pL .. 2ow-pL+kW-1 of input row (oh*sH-pT+kh) are <= 8 pixels x 4 floats = one 128-byte
// run. So the input is described to TMA as a tensor whose "pixel" pitch is 8 floats (two real pixels) and whose
// "channel" extent is 32 floats -- consecutive pixels OVERLAP -- and the convolution becomes kH taps of 32 channels
// with unit stride along W: exactly what the generic TMA kernels consume. The caller stores the image row-packed:
// [B][H][row_pitch_px][4] with the image starting at pixel pL+1 (so every window start is pixel-pair aligned with one
// leading dummy tap) and zeros in the padding columns; padding in H is resolved by the TMA unit.
//
// Filter rows wider than 7 taps (AlexNet's 11x11/4) take two 32-float K-chunks per row: chunk c holds window pixels
// 8c .. 8c+7, fetched from the same tensor map with a W offset of 8c packed pixels = 8c / stride_w window starts -- which is
// why k_w in 8..15 needs a stride_w that divides 8. The second chunk of an 11-tap row is half padding (4 of its 8 taps).
#include "conv_common.cuh"
#include "conv_tma.cuh"

int conv_tc_wgrad_splits(dl4jb200_ctx *ctx, const ConvGeom &g, int c_out);

// chunks of 8 window pixels per filter row: the window is one dummy tap + k_w real ones
static inline int stem_chunks(const dl4jb200_conv_desc *d) { return (d->k_w + 1 + 7) / 8; }

// W[co][ci][kh][kw] -> Wp[n][(kh*nchunk + c)*32 + p8*4 + ci],  kw = 8*c + p8 - 1  (kw == -1 / >= kW / ci >= Cin: zero)
__global__ void __launch_bounds__(256) k_pack_stem_filter(const float *__restrict__ w, float *__restrict__ wp, int Cout, int Cin,
                                                          int kH, int kW, int Kp, int rows_p, int nchunk) {
    size_t total = (size_t)rows_p * Kp, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int n = (int)(i / Kp), k = (int)(i % Kp);
        int kc = k >> 5, r = k & 31, kh = kc / nchunk, c = kc - kh * nchunk, ci = r & 3, kw = 8 * c + (r >> 2) - 1;
        float v = 0.f;
        if (n < Cout && kh < kH && kw >= 0 && kw < kW && ci < Cin) v = w[(((size_t)n * Cin + ci) * kH + kh) * kW + kw];
        wp[i] = to_tf32_rna(v);
    }
}

// dW[co][ci][kh][kw] = sum_s partial[s][co][(kh*nchunk + (kw+1)/8)*32 + ((kw+1)%8)*4 + ci]
__global__ void __launch_bounds__(256) k_stem_wgrad_finalize(const float *__restrict__ partial, float *__restrict__ dw, int Cout,
                                                             int Cin, int kH, int kW, int Kp, int rows_p, int splits,
                                                             const float *__restrict__ db_partial, float *__restrict__ db, int nchunk) {
    size_t total = (size_t)Cout * Cin * kH * kW, stride = (size_t)gridDim.x * blockDim.x;
    pdl_trigger();
    pdl_wait();
    if (db_partial) {  // [splits][rows_p] sums from the GEMM's ones column
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)Cout; i += stride) {
            float a0 = 0.f, a1 = 0.f;
            int s = 0;
            for (; s + 1 < splits; s += 2) { a0 += db_partial[(size_t)s * rows_p + i]; a1 += db_partial[(size_t)(s + 1) * rows_p + i]; }
            if (s < splits) a0 += db_partial[(size_t)s * rows_p + i];
            db[i] = a0 + a1;
        }
    }
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int kw = (int)(i % kW);
        size_t t = i / kW;
        int kh = (int)(t % kH); t /= kH;
        int ci = (int)(t % Cin), co = (int)(t / Cin);
        int k = (kh * nchunk + ((kw + 1) >> 3)) * 32 + ((kw + 1) & 7) * 4 + ci;
        const float *p = partial + (size_t)co * Kp + k;
        const size_t ss = (size_t)rows_p * Kp;
        // four independent partial sums: the loads of a round are in flight together (the fold over ~148 splits is a latency
        // chain otherwise, and this kernel is the tail of the step); fixed combination order, deterministic
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int s = 0;
        for (; s + 3 < splits; s += 4) {
            const float v0 = p[(size_t)s * ss], v1 = p[(size_t)(s + 1) * ss], v2 = p[(size_t)(s + 2) * ss], v3 = p[(size_t)(s + 3) * ss];
            a0 += v0; a1 += v1; a2 += v2; a3 += v3;
        }
        for (; s < splits; s++) a0 += p[(size_t)s * ss];
        dw[i] = (a0 + a1) + (a2 + a3);
    }
}

// NCHW image -> row-packed NHWC4: dst[b][y][x_off + x][c], c < C <= 4 (padding columns / channel 3 untouched: keep them zero)
__global__ void __launch_bounds__(256) k_nchw_to_rowpack4(const float *__restrict__ src, float *__restrict__ dst, int B, int C, int H,
                                                          int W, int row_pitch_px, int x_off, int rnd) {
    size_t total = (size_t)B * H * W, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int x = (int)(i % W);
        size_t t = i / W;
        int y = (int)(t % H), b = (int)(t / H);
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        for (int c = 0; c < C; c++) v[c] = dl4j_rnd(src[(((size_t)b * C + c) * H + y) * W + x], rnd);
        *reinterpret_cast<float4 *>(dst + (((size_t)b * H + y) * row_pitch_px + x_off + x) * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// Four pixels of a row per thread when W % 4 == 0 and the planes are 16-byte aligned: one 128-bit load per plane, four 128-bit
// stores (the one-pixel form runs at 3.5 TB/s on the 155 MB ResNet-50 minibatch, and this pass is inside the timed step).
__global__ void __launch_bounds__(256) k_nchw_to_rowpack4_v4(const float *__restrict__ src, float *__restrict__ dst, int B, int C, int H,
                                                             int W4, int row_pitch_px, int x_off, int rnd, FastDiv d_w4, FastDiv d_h) {
    const size_t total = (size_t)B * H * W4, stride = (size_t)gridDim.x * blockDim.x, plane = (size_t)H * W4 * 4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        uint32_t t, x4, b, y;
        d_w4.divmod((uint32_t)i, t, x4);
        d_h.divmod(t, b, y);
        float4 v[4] = {make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0), make_float4(0, 0, 0, 0)};
        const float *s = src + (size_t)b * C * plane + ((size_t)y * W4 + x4) * 4;
#pragma unroll
        for (int c = 0; c < 4; c++)
            if (c < C) v[c] = dl4j_rnd4(*reinterpret_cast<const float4 *>(s + (size_t)c * plane), rnd);
        float4 *d = reinterpret_cast<float4 *>(dst + (((size_t)b * H + y) * row_pitch_px + x_off + (size_t)x4 * 4) * 4);
        d[0] = make_float4(v[0].x, v[1].x, v[2].x, v[3].x);
        d[1] = make_float4(v[0].y, v[1].y, v[2].y, v[3].y);
        d[2] = make_float4(v[0].z, v[1].z, v[2].z, v[3].z);
        d[3] = make_float4(v[0].w, v[1].w, v[2].w, v[3].w);
    }
}

static int stem_check(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int row_pitch_px, const char *who) {
    DL4J_REQUIRE(ctx && d, "%s: NULL argument", who);
    if (d->c_in > 4 || d->stride_w < 1 || d->stride_w > 8 || d->k_w > 15 || (d->k_w > 7 && 8 % d->stride_w != 0) || d->k_h > 16 ||
        d->dil_h != 1 || d->dil_w != 1 || !tma_driver_ready() || d->pad_top > 127 || d->stride_h > 8) {
        dl4j_set_error("%s: row-packed path needs c_in <= 4, stride_w <= 8, k_w <= 7 (<= 15 when stride_w divides 8), no dilation", who);
        return DL4JB200_NOT_SUPPORTED;
    }
    DL4J_REQUIRE(row_pitch_px % 2 == 0 && row_pitch_px >= d->stride_w * (d->out_w - 1) + 8 * stem_chunks(d) && row_pitch_px >= d->pad_left + 1 + d->in_w,
                 "%s: row_pitch_px must be even and >= max(stride_w*(out_w-1) + 8*chunks, pad_left + 1 + in_w)", who);
    return DL4JB200_OK;
}

static int stem_xmap(CUtensorMap *tm, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px, uint32_t pixels,
                     CUtensorMapSwizzle swz) {
    // tensor {C = 32 floats, W = out_w window starts (pitch stride_w pixels = 4*stride_w floats), H = in_h rows, N}; taps only
    // along H. Window ow covers packed pixels stride_w*ow .. +7, i.e. image columns stride_w*ow - pad_left - 1 .. : one
    // leading dummy tap, then the k_w <= 7 real ones (the filter is packed accordingly, independent of the stride).
    // (a second chunk of a wide filter row starts 8 / stride_w windows further right: the W extent includes those starts)
    const int extra = (stem_chunks(d) - 1) * (8 / d->stride_w);
    return tma_make_im2col_pitched(tm, xp, d->batch, d->in_h, d->out_w + extra, 32, 4 * d->stride_w, (long long)row_pitch_px * 4,
                                   (long long)d->in_h * row_pitch_px * 4, 0, -d->pad_top, d->out_w, d->out_h, 1, d->stride_h, pixels, swz);
}

extern "C" {

int dl4jb200_nchw_to_rowpack4(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int row_pitch_px,
                              int x_off) {
    DL4J_REQUIRE(ctx && src && dst, "nchw_to_rowpack4: NULL argument");
    DL4J_REQUIRE(C >= 1 && C <= 4 && B > 0 && H > 0 && W > 0 && x_off >= 0 && x_off + W <= row_pitch_px, "nchw_to_rowpack4: bad shape");
    size_t total = (size_t)B * H * W;
    if (W % 4 == 0 && ((uintptr_t)src & 15) == 0 && total / 4 < (1ull << 32)) {
        const size_t items = total / 4, blocks4 = (items + 255) / 256, cap4 = (size_t)ctx->sm_count * 16;
        k_nchw_to_rowpack4_v4<<<(int)(blocks4 > cap4 ? cap4 : blocks4), 256, 0, ctx->stream>>>(src, dst, B, C, H, W / 4, row_pitch_px, x_off,
                                                                                              ctx->round_act, FastDiv((uint32_t)(W / 4)),
                                                                                              FastDiv((uint32_t)H));
        DL4J_CHECK_LAUNCH();
        return DL4JB200_OK;
    }
    size_t blocks = (total + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    k_nchw_to_rowpack4<<<(int)(blocks > cap ? cap : blocks), 256, 0, ctx->stream>>>(src, dst, B, C, H, W, row_pitch_px, x_off, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

static int rowpack_fwd_impl(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px, const float *w,
                            const float *bias, float *z, int z_cs, int act, const float *pivot, float *stat_partial, int *rows_out) {
    int rc = stem_check(ctx, d, row_pitch_px, "conv_rowpack_fwd");
    if (rc) return rc;
    DL4J_REQUIRE(xp && w && z && z_cs >= d->c_out && z_cs % 4 == 0 && ((uintptr_t)z & 15) == 0 && ((uintptr_t)xp & 15) == 0,
                 "conv_rowpack_fwd: bad output / input buffer");
    if (act < 0 || act > DL4JB200_ACT_LEAKYRELU) { dl4j_set_error("conv_rowpack_fwd: activation %d not handled", act); return DL4JB200_NOT_SUPPORTED; }
    if (bias && ((uintptr_t)bias & 15)) { dl4j_set_error("conv_rowpack_fwd: bias must be 16-byte aligned"); return DL4JB200_NOT_SUPPORTED; }
    const int nchunk = stem_chunks(d), Kp = d->k_h * nchunk * 32, rows_p = dl4j_round_up(z_cs, 256);
    float *wp = (float *)dl4j_workspace(ctx, 0, (size_t)rows_p * Kp * sizeof(float));
    if (!wp) return DL4JB200_ERR_CUDA;
    k_pack_stem_filter<<<dl4j_ceil_div((long long)rows_p * Kp, 256), 256, 0, ctx->stream>>>(w, wp, d->c_out, d->c_in, d->k_h, d->k_w, Kp, rows_p,
                                                                                            nchunk);
    DL4J_CHECK_LAUNCH();
    alignas(64) CUtensorMap tmA, tmB;
    if ((rc = stem_xmap(&tmA, d, xp, row_pitch_px, 128, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    const int bn = z_cs > 64 ? 128 : (z_cs > 32 ? 64 : 32);
    if ((rc = tma_make_2d(&tmB, wp, (uint64_t)rows_p, (uint64_t)Kp, (uint64_t)Kp, (uint32_t)bn, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    tcgeom::TmaGemmGeom t;
    // K-chunk it = (filter row it / nchunk, chunk it % nchunk): the kernels read it as a "tap" (kh, kw) with W dilation 8 / stride_w
    t.M = d->batch * d->out_h * d->out_w; t.num_k = d->k_h * nchunk; t.kW = nchunk; t.dH = 1; t.dW = nchunk > 1 ? 8 / d->stride_w : 1;
    t.sH = d->stride_h; t.sW = 1;
    t.lowH = -d->pad_top; t.lowW = 0;
    t.div_Wm = FastDiv((uint32_t)d->out_w); t.div_HWm = FastDiv((uint32_t)(d->out_h * d->out_w));
    t.div_Cg = FastDiv(32u); t.div_kW = FastDiv((uint32_t)nchunk);
    t.remap = 0; t.rm_sy = t.rm_sx = 1; t.rm_H = d->out_h; t.rm_W = d->out_w;
    const int act_code = act | ((ctx->round_act && act != DL4JB200_ACT_IDENTITY) ? 0x100 : 0);
    if (stat_partial) {  // batch-norm statistics in the epilogue (see dl4jb200_conv_fwd_bnstats)
        const int n_tiles = dl4j_ceil_div(z_cs, bn), total = dl4j_ceil_div(t.M, 128) * n_tiles;
        const int grid = (total < ctx->sm_count ? total : ctx->sm_count) / n_tiles * n_tiles;
        if (act != DL4JB200_ACT_IDENTITY || d->c_out % bn != 0 || z_cs != d->c_out || grid <= 0 || !pivot || ((uintptr_t)pivot & 15) ||
            ((uintptr_t)stat_partial & 15)) {
            dl4j_set_error("conv_rowpack_fwd_bnstats: this convolution cannot carry the statistics");
            return DL4JB200_NOT_SUPPORTED;
        }
        t.stat_out = stat_partial;
        t.stat_pivot = pivot;
        *rows_out = grid / n_tiles * (bn >= 64 ? 1 : 4);  // persist_stat_rows
    }
    return conv_tc_gemm_tma_raw(ctx, tmA, tmB, t, bias, z, d->c_out, z_cs, act_code, bn);
}

int dl4jb200_conv_rowpack_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px, const float *w,
                              const float *bias, float *z, int z_cs, int act) {
    return rowpack_fwd_impl(ctx, d, xp, row_pitch_px, w, bias, z, z_cs, act, nullptr, nullptr, nullptr);
}

int dl4jb200_conv_rowpack_fwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px,
                                      const float *w, const float *bias, float *z, int z_cs, const float *pivot,
                                      float *stat_partial, int *rows_out) {
    DL4J_REQUIRE(pivot && stat_partial && rows_out, "conv_rowpack_fwd_bnstats: NULL argument");
    return rowpack_fwd_impl(ctx, d, xp, row_pitch_px, w, bias, z, z_cs, DL4JB200_ACT_IDENTITY, pivot, stat_partial, rows_out);
}

int dl4jb200_conv_rowpack_wgrad(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px,
                                const float *delta, int delta_cs, float *dw, float *db) {
    int rc = stem_check(ctx, d, row_pitch_px, "conv_rowpack_wgrad");
    if (rc) return rc;
    DL4J_REQUIRE(xp && delta && dw && delta_cs >= d->c_out && delta_cs % 4 == 0 && ((uintptr_t)delta & 15) == 0 && ((uintptr_t)xp & 15) == 0,
                 "conv_rowpack_wgrad: bad buffer");
    const int nchunk = stem_chunks(d), Kp = d->k_h * nchunk * 32, rows_w = dl4j_round_up(d->c_out, 128), M = d->batch * d->out_h * d->out_w;
    ConvGeom g;  // only the fields the split heuristic reads
    g.Kp = Kp; g.M = M;
    const int splits = conv_tc_wgrad_splits(ctx, g, d->c_out);
    const int mps = dl4j_round_up(dl4j_ceil_div(M, splits), 32);
    // db: with TF32-representable deltas (DL4JB200_OPT_TF32_ACTIVATIONS) it rides in the GEMM as a ones column, else it is
    // the plain column-sum pass
    const bool db_fused = db && ctx->round_act;
    float *partial = (float *)dl4j_workspace(ctx, 0, ((size_t)splits * rows_w * Kp + (db_fused ? (size_t)splits * rows_w : 0)) * sizeof(float));
    if (!partial) return DL4JB200_ERR_CUDA;
    float *dbp = db_fused ? partial + (size_t)splits * rows_w * Kp : nullptr;
    alignas(64) CUtensorMap tmD, tmX;
    if ((rc = tma_make_2d(&tmD, delta, (uint64_t)M, (uint64_t)delta_cs, (uint64_t)delta_cs, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))) return rc;
    if ((rc = stem_xmap(&tmX, d, xp, row_pitch_px, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))) return rc;
    tcgeom::TmaWgradGeom t;
    t.M = M; t.kW = nchunk; t.dH = 1; t.dW = nchunk > 1 ? 8 / d->stride_w : 1; t.sH = d->stride_h; t.sW = 1; t.lowH = -d->pad_top; t.lowW = 0;
    t.Kp = Kp;
    t.div_Wm = FastDiv((uint32_t)d->out_w); t.div_HWm = FastDiv((uint32_t)(d->out_h * d->out_w));
    t.div_Cg = FastDiv(32u); t.div_kW = FastDiv((uint32_t)nchunk);
    if ((rc = conv_tc_wgrad_tma_raw(ctx, tmD, tmX, t, partial, rows_w, splits, mps, dbp))) return rc;
    size_t total = (size_t)d->c_out * d->c_in * d->k_h * d->k_w;
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_stem_wgrad_finalize, dim3((unsigned)dl4j_ceil_div((long long)total, 256)), dim3(256), 0, ctx->stream,
                                  (const float *)partial, dw, d->c_out, d->c_in, d->k_h, d->k_w, Kp, rows_w, splits, (const float *)dbp, db, nchunk));
    DL4J_CHECK_LAUNCH();
    if (db && !db_fused) return dl4jb200_bias_grad(ctx, delta, M, delta_cs, d->c_out, db);
    return DL4JB200_OK;
}

}  // extern "C"
