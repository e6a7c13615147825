// Implicit-GEMM view of DL4J's convolution passes, shared by the FP32 (CUDA-core) and TF32 (tcgen05) kernels.
//
// All three passes gather 16-byte channel quads from an NHWC tensor ("src") into GEMM operand tiles:
//   forward      z[m, co]  = sum_k  A(m,k) * Wf[co,k]     m=(b,oh,ow)  k=(kh,kw,ci)   src = x
//   backprop-data dx[m, ci] = sum_k  A'(m,k) * Wb[ci,k]    m=(b,h,w)    k=(kh,kw,co)   src = delta (transposed gather)
//   backprop-filter dW[co,k] = sum_m delta[m,co] * A(m,k)  reduction over m, split across CTAs, src = x
// K is indexed tap-major with Cg = round_up(C,4) channels per tap, so a 16-byte quad never straddles taps
// and zero-filled channel padding keeps the arithmetic exact. (ConvolutionLayer.java:202-257, 385-417.)
#pragma once
#include "common.cuh"

struct ConvGeom {
    int B, Hs, Ws, cs_src;  // gathered tensor: spatial dims, pixel stride (floats)
    int Hm, Wm;             // pixel grid the GEMM rows enumerate; M = B*Hm*Wm
    int kH, kW, sH, sW, pT, pL, dH, dW;
    int Cg, taps, K, Kp, M;
    int transposed;         // 0: iy = y*sH - pT + kh*dH      1: iy = (y + pT - kh*dH)/sH when divisible
    int fast;               // taps are an affine offset from a per-row base pointer (see conv_row_setup)
    int use_tma;            // operands fetched by TMA (im2col mode); backprop-data then needs the filter packed tap-flipped
    int remap, rm_sy, rm_sx, rm_H, rm_W;  // output-row scatter (see TmaGemmGeom)
    int rm_keep = 0;        // the pixels between the scattered rows belong to other launches: neither zero nor clear them
    FastDiv div_Wm, div_HWm, div_Cg, div_kW;
    // fused epsilon fan-in + activation backprop of the persistent TMA kernel's epilogue (see TmaGemmGeom)
    const float *epi_add = nullptr, *epi_mask = nullptr;
    float *stat_out = nullptr;          // batch-norm statistics of the output (see TmaGemmGeom)
    const float *stat_pivot = nullptr;
    // backward batch-norm sums of the vertex that owns the output of a backprop-data GEMM (see TmaGemmGeom)
    const float *bst_x = nullptr, *bst_mean = nullptr, *bst_std = nullptr, *bst_gamma = nullptr, *bst_beta = nullptr;
    int bst_relu = 0;
};

// tap/channel of a K index (k % 4 == 0 in all callers)
__device__ __forceinline__ void conv_k_decode(const ConvGeom &g, int k, int &kh, int &kw, int &c, bool &valid) {
    uint32_t tap, cc, a, b;
    g.div_Cg.divmod((uint32_t)k, tap, cc);
    g.div_kW.divmod(tap, a, b);
    kh = (int)a; kw = (int)b; c = (int)cc;
    valid = (int)tap < g.taps;
}

// pixel of a GEMM row: returns base coordinates used by conv_src_offset
__device__ __forceinline__ void conv_m_decode(const ConvGeom &g, int m, int &b, int &y0, int &x0) {
    uint32_t bb, r, yy, xx;
    g.div_HWm.divmod((uint32_t)m, bb, r);
    g.div_Wm.divmod(r, yy, xx);
    b = (int)bb;
    if (g.transposed) { y0 = (int)yy + g.pT; x0 = (int)xx + g.pL; }
    else { y0 = (int)yy * g.sH - g.pT; x0 = (int)xx * g.sW - g.pL; }
}

// element offset into src of (row pixel, tap) or -1 when the tap falls into padding / between strides
__device__ __forceinline__ long long conv_src_offset(const ConvGeom &g, int b, int y0, int x0, int kh, int kw) {
    int iy, ix;
    if (g.transposed) {
        int ty = y0 - kh * g.dH, tx = x0 - kw * g.dW;
        if (ty < 0 || tx < 0) return -1;
        if (g.sH == 1) iy = ty; else { iy = ty / g.sH; if (iy * g.sH != ty) return -1; }
        if (g.sW == 1) ix = tx; else { ix = tx / g.sW; if (ix * g.sW != tx) return -1; }
    } else {
        iy = y0 + kh * g.dH; ix = x0 + kw * g.dW;
        if (iy < 0 || ix < 0) return -1;
    }
    if (iy >= g.Hs || ix >= g.Ws) return -1;
    return (((long long)b * g.Hs + iy) * g.Ws + ix) * g.cs_src;
}

// Fast gather set-up for one GEMM row (g.fast): element offset of the row's tap-(0,0) source pixel and a validity
// mask -- bit kh says "tap row kh is inside the image", bit 16+kw the same for tap column kw -- so that per K-chunk a
// producer needs only `rowoff + tapoff` and one mask test per 16-byte copy.
__device__ __forceinline__ void conv_row_setup(const ConvGeom &g, int m, long long &rowoff, uint32_t &mask) {
    int b, y0, x0;
    conv_m_decode(g, m, b, y0, x0);
    mask = 0;
    if (g.transposed && (g.sH != 1 || g.sW != 1)) {  // 1x1 strided backprop-data: a single tap
        int iy = y0 / g.sH, ix = x0 / g.sW;
        bool ok = y0 >= 0 && x0 >= 0 && iy * g.sH == y0 && ix * g.sW == x0 && iy < g.Hs && ix < g.Ws;
        rowoff = ok ? (((long long)b * g.Hs + iy) * g.Ws + ix) * g.cs_src : 0;
        mask = ok ? 0x00010001u : 0u;
        return;
    }
    const int sgn = g.transposed ? -1 : 1;
    for (int kh = 0; kh < g.kH; kh++) { int iy = y0 + sgn * kh * g.dH; if (iy >= 0 && iy < g.Hs) mask |= 1u << kh; }
    for (int kw = 0; kw < g.kW; kw++) { int ix = x0 + sgn * kw * g.dW; if (ix >= 0 && ix < g.Ws) mask |= 1u << (16 + kw); }
    rowoff = (((long long)b * g.Hs + y0) * g.Ws + x0) * g.cs_src;
}
// per-chunk part: offset of tap (kh,kw) channel c relative to the row base, and the mask bits that must be set
__device__ __forceinline__ void conv_tap_setup(const ConvGeom &g, int kh, int kw, int c, int &tapoff, uint32_t &need) {
    const int sgn = g.transposed ? -1 : 1;
    tapoff = sgn * (kh * g.dH * g.Ws + kw * g.dW) * g.cs_src + c;
    need = (1u << kh) | (1u << (16 + kw));
}

__device__ __forceinline__ float to_tf32_rna(float v) { return dl4j_tf32(v); }

// host helpers implemented in conv.cu
ConvGeom conv_geom_fwd(const dl4jb200_conv_desc *d, int x_cs);
ConvGeom conv_geom_bwd_data(const dl4jb200_conv_desc *d, int delta_cs);

// kernels implemented in conv_ffma.cu / conv_tc.cu (host launchers)
int conv_ffma_gemm(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, const float *bias,
                   float *out, int n_valid, int out_cs, int act);
int conv_ffma_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs, int c_out,
                    float *partial, int rows_p, int splits);
int conv_tc_gemm(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, int n_rows_packed,
                 const float *bias, float *out, int n_valid, int out_cs, int act);
int conv_tc_stat_rows(dl4jb200_ctx *ctx, const ConvGeom &g, int n_rows_packed, int n_valid, int out_cs);
int conv_tc_bst_rows(dl4jb200_ctx *ctx, const ConvGeom &g, int n_rows_packed, int n_valid, int out_cs);
int conv_tc_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs, int c_out,
                  float *partial, int rows_p, int splits, float *db_partial = nullptr);
int conv_tc_wgrad_splits(dl4jb200_ctx *ctx, const ConvGeom &g, int c_out);
bool conv_tc_gemm_can_tma(const ConvGeom &g);
