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
    FastDiv div_Wm, div_HWm, div_Cg, div_kW;
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

__device__ __forceinline__ float to_tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}

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
int conv_tc_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs, int c_out,
                  float *partial, int rows_p, int splits);
int conv_tc_wgrad_splits(dl4jb200_ctx *ctx, const ConvGeom &g, int c_out);
