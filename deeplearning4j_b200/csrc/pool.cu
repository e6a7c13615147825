// Max / average / p-norm pooling forward and backward (SubsamplingLayer.java:112-268, 285-363).
// Bandwidth kernels: one thread per output element (NCHW, coalesced along W) or per 4 channels of one
// output pixel (NHWC, 128-bit loads/stores along C). Backward is a GATHER over the windows that contain
// an input cell -- no atomics, fixed summation order (kh, kw ascending), so results are deterministic and
// the first-maximum tie-break of the reference's IsMax (SubsamplingLayerTest.java:85-101) is bit-exact.
#include <float.h>

#include "common.cuh"

struct PoolGeom {
    int B, C, H, W, Ho, Wo, kH, kW, sH, sW, pT, pL, dH, dW;
    int xcs, ycs;  // NHWC pixel strides
    FastDiv d_sH, d_sW;            // stride divisions of the backward gather
    FastDiv d_q, d_w, d_h, d_c;    // index decode of the tensor the kernel enumerates (set per launch)
    int rnd;                       // store outputs rounded to TF32 (ctx->round_act)
};

template <int V>
struct Vec {
    float v[V];
};
template <int V>
__device__ __forceinline__ Vec<V> vload(const float *p) {
    Vec<V> r;
    if (V == 4) {
        float4 t = *reinterpret_cast<const float4 *>(p);
        r.v[0] = t.x; r.v[1 % V] = t.y; r.v[2 % V] = t.z; r.v[3 % V] = t.w;
    } else {
        r.v[0] = *p;
    }
    return r;
}
template <int V>
__device__ __forceinline__ void vstore(float *p, const Vec<V> &r) {
    if (V == 4) *reinterpret_cast<float4 *>(p) = make_float4(r.v[0], r.v[1 % V], r.v[2 % V], r.v[3 % V]);
    else *p = r.v[0];
}
template <int V>
__device__ __forceinline__ void vstore_i(int32_t *p, const int (&r)[V]) {
    if (V == 4) *reinterpret_cast<int4 *>(p) = make_int4(r[0], r[1 % V], r[2 % V], r[3 % V]);
    else *p = r[0];
}

// NHWC: V = 4, a work item is (b, y, x, channel quad). NCHW: V = 1, a work item is (b, c, y, x).
template <bool NHWC>
__device__ __forceinline__ void decode(size_t idx, const PoolGeom &g, int &b, int &c, int &y, int &x) {
    uint32_t i = (uint32_t)idx, q, r;  // work-item counts are < 2^31 (checked on the host)
    if (NHWC) {
        g.d_q.divmod(i, q, r); c = (int)r * 4; i = q;
        g.d_w.divmod(i, q, r); x = (int)r; i = q;
        g.d_h.divmod(i, q, r); y = (int)r; b = (int)q;
    } else {
        g.d_w.divmod(i, q, r); x = (int)r; i = q;
        g.d_h.divmod(i, q, r); y = (int)r; i = q;
        g.d_c.divmod(i, q, r); c = (int)r; b = (int)q;
    }
}
template <bool NHWC>
__device__ __forceinline__ size_t offs(int b, int c, int y, int x, int C, int Hh, int Ww, int cs) {
    return NHWC ? (((size_t)b * Hh + y) * Ww + x) * cs + c : (((size_t)b * C + c) * Hh + y) * Ww + x;
}

// KS > 0: square KS x KS window known at compile time (tap loops unrolled: the generic loop is instruction-bound on the
// 3x3/2 stem pool); KS == 0: window from the geometry.
template <bool NHWC, int KS = 0>
__global__ void __launch_bounds__(256) k_pool_fwd(const float *__restrict__ x, float *__restrict__ y,
                                                  int32_t *__restrict__ argmax, PoolGeom g, int type, int pnorm,
                                                  size_t total) {
    constexpr int V = NHWC ? 4 : 1;
    const int kH = KS ? KS : g.kH, kW = KS ? KS : g.kW;
    pdl_trigger();
    pdl_wait();
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        int b, c, oh, ow;
        decode<NHWC>(idx, g, b, c, oh, ow);
        float acc[V];
        int best[V];
#pragma unroll
        for (int i = 0; i < V; i++) { acc[i] = type == DL4JB200_POOL_MAX ? -FLT_MAX : 0.f; best[i] = 0; }
        auto tap = [&](const Vec<V> &v, int t) {
#pragma unroll
            for (int i = 0; i < V; i++) {
                if (type == DL4JB200_POOL_MAX) {
                    if (v.v[i] > acc[i]) { acc[i] = v.v[i]; best[i] = t; }
                } else if (type == DL4JB200_POOL_PNORM) {
                    acc[i] += powf(fabsf(v.v[i]), (float)pnorm);
                } else {
                    acc[i] += v.v[i];
                }
            }
        };
        const int ih0 = oh * g.sH - g.pT, iw0 = ow * g.sW - g.pL;
        if (KS && ih0 >= 0 && iw0 >= 0 && ih0 + (KS - 1) * g.dH < g.H && iw0 + (KS - 1) * g.dW < g.W) {
            // window entirely inside the image: issue all KS*KS loads before the first use (the bounds-checked loop
            // serialises them behind their branches and runs latency-bound)
            Vec<V> v[KS ? KS * KS : 1];
#pragma unroll
            for (int kh = 0; kh < KS; kh++)
#pragma unroll
                for (int kw = 0; kw < KS; kw++)
                    v[kh * KS + kw] = vload<V>(x + offs<NHWC>(b, c, ih0 + kh * g.dH, iw0 + kw * g.dW, g.C, g.H, g.W, g.xcs));
#pragma unroll
            for (int t = 0; t < KS * KS; t++) tap(v[t], t);
        } else {
#pragma unroll KS ? KS : 1
            for (int kh = 0; kh < kH; kh++) {
                int ih = ih0 + kh * g.dH;
                if (ih < 0 || ih >= g.H) continue;  // padding: -FLT_MAX for max (never wins), 0 for sums
#pragma unroll KS ? KS : 1
                for (int kw = 0; kw < kW; kw++) {
                    int iw = iw0 + kw * g.dW;
                    if (iw < 0 || iw >= g.W) continue;
                    tap(vload<V>(x + offs<NHWC>(b, c, ih, iw, g.C, g.H, g.W, g.xcs)), kh * kW + kw);
                }
            }
        }
        Vec<V> out;
#pragma unroll
        for (int i = 0; i < V; i++) {
            if (type == DL4JB200_POOL_AVG) out.v[i] = dl4j_rnd(acc[i] / (float)(kH * kW), g.rnd);
            else if (type == DL4JB200_POOL_PNORM) out.v[i] = dl4j_rnd(powf(acc[i], 1.f / (float)pnorm), g.rnd);
            else out.v[i] = dl4j_rnd(acc[i], g.rnd);  // no-op when the producer already rounded (idempotent)
        }
        size_t o = offs<NHWC>(b, c, oh, ow, g.C, g.Ho, g.Wo, g.ycs);
        vstore<V>(y + o, out);
        if (argmax && type == DL4JB200_POOL_MAX) vstore_i<V>(argmax + o, best);
    }
}

template <bool NHWC>
__global__ void __launch_bounds__(256) k_pool_bwd(const float *__restrict__ x, const float *__restrict__ y,
                                                  const float *__restrict__ eps, const int32_t *__restrict__ argmax,
                                                  float *__restrict__ dx, PoolGeom g, int type, int pnorm,
                                                  float pnorm_eps, size_t total) {
    constexpr int V = NHWC ? 4 : 1;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        int b, c, h, w;
        decode<NHWC>(idx, g, b, c, h, w);
        Vec<V> xv;
        if (type != DL4JB200_POOL_AVG && !(type == DL4JB200_POOL_MAX && argmax))
            xv = vload<V>(x + offs<NHWC>(b, c, h, w, g.C, g.H, g.W, g.xcs));
        float acc[V];
#pragma unroll
        for (int i = 0; i < V; i++) acc[i] = 0.f;
        for (int kh = 0; kh < g.kH; kh++) {
            int t = h + g.pT - kh * g.dH;
            if (t < 0) continue;
            uint32_t qq, rr;
            g.d_sH.divmod((uint32_t)t, qq, rr);
            if (rr) continue;
            int oh = (int)qq;
            if (oh >= g.Ho) continue;
            for (int kw = 0; kw < g.kW; kw++) {
                int u = w + g.pL - kw * g.dW;
                if (u < 0) continue;
                g.d_sW.divmod((uint32_t)u, qq, rr);
                if (rr) continue;
                int ow = (int)qq;
                if (ow >= g.Wo) continue;
                size_t o = offs<NHWC>(b, c, oh, ow, g.C, g.Ho, g.Wo, g.ycs);
                Vec<V> e = vload<V>(eps + o);
                if (type == DL4JB200_POOL_AVG) {
#pragma unroll
                    for (int i = 0; i < V; i++) acc[i] += e.v[i];
                } else if (type == DL4JB200_POOL_MAX) {
                    if (argmax) {
                        int me = kh * g.kW + kw;
                        if (V == 4) {
                            int4 a = *reinterpret_cast<const int4 *>(argmax + o);
                            acc[0] += a.x == me ? e.v[0] : 0.f; acc[1 % V] += a.y == me ? e.v[1 % V] : 0.f;
                            acc[2 % V] += a.z == me ? e.v[2 % V] : 0.f; acc[3 % V] += a.w == me ? e.v[3 % V] : 0.f;
                        } else {
                            acc[0] += argmax[o] == me ? e.v[0] : 0.f;
                        }
                    } else {
                        // IsMax recomputed: this cell is the window's first maximum iff it is strictly greater
                        // than every earlier cell (row-major) and >= every later one; padding cells are -FLT_MAX.
                        bool win[V];
#pragma unroll
                        for (int i = 0; i < V; i++) win[i] = true;
                        for (int k2 = 0; k2 < g.kH; k2++) {
                            int ih = oh * g.sH - g.pT + k2 * g.dH;
                            for (int k3 = 0; k3 < g.kW; k3++) {
                                if (k2 == kh && k3 == kw) continue;
                                int iw = ow * g.sW - g.pL + k3 * g.dW;
                                bool inside = ih >= 0 && ih < g.H && iw >= 0 && iw < g.W;
                                Vec<V> o2;
                                if (inside) o2 = vload<V>(x + offs<NHWC>(b, c, ih, iw, g.C, g.H, g.W, g.xcs));
                                bool earlier = (k2 * g.kW + k3) < (kh * g.kW + kw);
#pragma unroll
                                for (int i = 0; i < V; i++) {
                                    float ov = inside ? o2.v[i] : -FLT_MAX;
                                    win[i] = win[i] && (earlier ? (xv.v[i] > ov) : (xv.v[i] >= ov));
                                }
                            }
                        }
#pragma unroll
                        for (int i = 0; i < V; i++) acc[i] += win[i] ? e.v[i] : 0.f;
                    }
                } else {  // PNORM: dx += eps * x|x|^(p-2) / max(y^(p-1), eps_p)   (SubsamplingLayer.java:225-250)
                    Vec<V> yv = vload<V>(y + o);
#pragma unroll
                    for (int i = 0; i < V; i++) {
                        float den = fmaxf(powf(yv.v[i], (float)(pnorm - 1)), pnorm_eps);
                        acc[i] += e.v[i] * xv.v[i] * powf(fabsf(xv.v[i]), (float)(pnorm - 2)) / den;
                    }
                }
            }
        }
        Vec<V> out;
#pragma unroll
        for (int i = 0; i < V; i++)
            out.v[i] = dl4j_rnd(type == DL4JB200_POOL_AVG ? acc[i] / (float)(g.kH * g.kW) : acc[i], g.rnd);
        vstore<V>(dx + offs<NHWC>(b, c, h, w, g.C, g.H, g.W, g.xcs), out);
    }
}

// Hot case of the backward pass (MAX, saved argmax, NHWC, no dilation): walk only the <= ceil(k/s)^2 windows that can
// contain the cell instead of testing all k*k taps for stride divisibility. Same summation order (kh, kw ascending).
__global__ void __launch_bounds__(256, 4) k_maxpool_bwd_nhwc_fast(const float *__restrict__ eps, const int32_t *__restrict__ argmax,
                                                                 float *__restrict__ dx, PoolGeom g, size_t total) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        int b, c, h, w;
        decode<true>(idx, g, b, c, h, w);
        const int th = h + g.pT, tw = w + g.pL;
        // SAME mode with stride > kernel yields NEGATIVE padding (ConvolutionUtils.java:243-253): th / tw may be < 0
        const int oh_hi = th < 0 ? -1 : min((int)g.d_sH.div((uint32_t)th), g.Ho - 1);
        const int ow_hi = tw < 0 ? -1 : min((int)g.d_sW.div((uint32_t)tw), g.Wo - 1);
        const int lh = th - g.kH + 1, lw = tw - g.kW + 1;
        const int oh_lo = lh <= 0 ? 0 : (int)g.d_sH.div((uint32_t)(lh + g.sH - 1));
        const int ow_lo = lw <= 0 ? 0 : (int)g.d_sW.div((uint32_t)(lw + g.sW - 1));
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int oh = oh_hi; oh >= oh_lo; oh--) {       // kh = th - oh*sH ascending
            const int kh = th - oh * g.sH;
            for (int ow = ow_hi; ow >= ow_lo; ow--) {   // kw ascending
                const int me = kh * g.kW + (tw - ow * g.sW);
                const size_t o = (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.ycs + c;
                const int4 a = *reinterpret_cast<const int4 *>(argmax + o);
                const float4 e = *reinterpret_cast<const float4 *>(eps + o);
                acc.x += a.x == me ? e.x : 0.f; acc.y += a.y == me ? e.y : 0.f;
                acc.z += a.z == me ? e.z : 0.f; acc.w += a.w == me ? e.w : 0.f;
            }
        }
        *reinterpret_cast<float4 *>(dx + (((size_t)b * g.H + h) * g.W + w) * g.xcs + c) = dl4j_rnd4(acc, g.rnd);
    }
}

// Stride-2 windows (ResNet's 3x3/2 stem pool, every 2x2/2 pool): one thread owns a 2x2 patch of dx (x 4 channels).
// The patch's four cells share their candidate windows, so eps / argmax are fetched once per window and patch instead
// of once per window and cell: 4 window reads per 4 cells instead of 9 for k = 3 (the kernel is bound by those L2
// reads). Per cell the contributions are added in the same order as k_maxpool_bwd_nhwc_fast (kh, kw ascending).
// SMALL (kH, kW <= 3): a patch meets at most 2 x 2 windows, all fetched up front (8 independent loads in flight instead of
// one dependent pair per loop iteration; same summation order per cell: window rows / columns descending = taps ascending).
template <bool SMALL>
__global__ void __launch_bounds__(256, 4) k_maxpool_bwd_nhwc_patch2(const float *__restrict__ eps, const int32_t *__restrict__ argmax,
                                                                   float *__restrict__ dx, PoolGeom g, int Hp, int Wp, size_t total) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    pdl_trigger();
    pdl_wait();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        uint32_t i = (uint32_t)idx, q, r;
        g.d_q.divmod(i, q, r); const int c = (int)r * 4; i = q;
        g.d_w.divmod(i, q, r); const int w0 = (int)r * 2; i = q;   // d_w / d_h hold the PATCH grid here
        g.d_h.divmod(i, q, r); const int h0 = (int)r * 2; const int b = (int)q;
        const int th0 = h0 + g.pT, tw0 = w0 + g.pL;  // >= 0: this kernel is only used with non-negative padding
        const int oh_hi = min((th0 + 1) >> 1, g.Ho - 1), ow_hi = min((tw0 + 1) >> 1, g.Wo - 1);
        const int lh = th0 - g.kH + 1, lw = tw0 - g.kW + 1;
        const int oh_lo = lh <= 0 ? 0 : (lh + 1) >> 1, ow_lo = lw <= 0 ? 0 : (lw + 1) >> 1;
        float4 acc[2][2];
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int d = 0; d < 2; d++) acc[a][d] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (SMALL) {
            int4 am[2][2];
            float4 e[2][2];
#pragma unroll
            for (int i = 0; i < 2; i++)
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const int oh = oh_hi - i, ow = ow_hi - j;
                    am[i][j] = make_int4(-1, -1, -1, -1);  // matches no tap
                    e[i][j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (oh >= oh_lo && ow >= ow_lo) {
                        const size_t o = (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.ycs + c;
                        am[i][j] = *reinterpret_cast<const int4 *>(argmax + o);
                        e[i][j] = *reinterpret_cast<const float4 *>(eps + o);
                    }
                }
#pragma unroll
            for (int i = 0; i < 2; i++)
#pragma unroll
                for (int j = 0; j < 2; j++)
#pragma unroll
                    for (int a = 0; a < 2; a++) {
                        const int kh = th0 - 2 * (oh_hi - i) + a;
#pragma unroll
                        for (int d = 0; d < 2; d++) {
                            const int kw = tw0 - 2 * (ow_hi - j) + d;
                            const int me = ((unsigned)kh < (unsigned)g.kH && (unsigned)kw < (unsigned)g.kW) ? kh * g.kW + kw : -2;
                            acc[a][d].x += am[i][j].x == me ? e[i][j].x : 0.f; acc[a][d].y += am[i][j].y == me ? e[i][j].y : 0.f;
                            acc[a][d].z += am[i][j].z == me ? e[i][j].z : 0.f; acc[a][d].w += am[i][j].w == me ? e[i][j].w : 0.f;
                        }
                    }
        } else {
            for (int oh = oh_hi; oh >= oh_lo; oh--) {
                // tap row of window oh that lands on patch row a: kh = th0 + a - 2*oh, valid when 0 <= kh < kH
                const int kh0 = th0 - 2 * oh;
                for (int ow = ow_hi; ow >= ow_lo; ow--) {
                    const int kw0 = tw0 - 2 * ow;
                    const size_t o = (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.ycs + c;
                    const int4 am = *reinterpret_cast<const int4 *>(argmax + o);
                    const float4 e = *reinterpret_cast<const float4 *>(eps + o);
#pragma unroll
                    for (int a = 0; a < 2; a++) {
                        const int kh = kh0 + a;
                        if (kh < 0 || kh >= g.kH) continue;
#pragma unroll
                        for (int d = 0; d < 2; d++) {
                            const int kw = kw0 + d;
                            if (kw < 0 || kw >= g.kW) continue;
                            const int me = kh * g.kW + kw;
                            acc[a][d].x += am.x == me ? e.x : 0.f; acc[a][d].y += am.y == me ? e.y : 0.f;
                            acc[a][d].z += am.z == me ? e.z : 0.f; acc[a][d].w += am.w == me ? e.w : 0.f;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int d = 0; d < 2; d++)
                if (h0 + a < g.H && w0 + d < g.W)
                    *reinterpret_cast<float4 *>(dx + (((size_t)b * g.H + h0 + a) * g.W + w0 + d) * g.xcs + c) = dl4j_rnd4(acc[a][d], g.rnd);
    }
}

static int pool_geom(const dl4jb200_conv_desc *d, int layout, int x_cs, int y_cs, PoolGeom &g, const char *who) {
    DL4J_REQUIRE(d, "%s: desc is NULL", who);
    DL4J_REQUIRE(d->batch > 0 && d->c_in > 0 && d->in_h > 0 && d->in_w > 0 && d->out_h > 0 && d->out_w > 0 &&
                     d->k_h > 0 && d->k_w > 0 && d->stride_h > 0 && d->stride_w > 0 && d->dil_h > 0 && d->dil_w > 0,
                 "%s: non-positive dimension in descriptor", who);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "%s: unknown layout %d", who, layout);
    if (layout == DL4JB200_NHWC)
        DL4J_REQUIRE(x_cs >= d->c_in && y_cs >= d->c_in && x_cs % 4 == 0 && y_cs % 4 == 0 && x_cs == y_cs,
                     "%s: NHWC pixel strides must be equal multiples of 4 and >= C", who);
    g = PoolGeom{d->batch, d->c_in, d->in_h, d->in_w, d->out_h, d->out_w, d->k_h, d->k_w, d->stride_h, d->stride_w,
                 d->pad_top, d->pad_left, d->dil_h, d->dil_w, x_cs, y_cs};
    g.d_sH = FastDiv((uint32_t)d->stride_h);
    g.d_sW = FastDiv((uint32_t)d->stride_w);
    DL4J_REQUIRE((long long)d->batch * d->c_in * d->in_h * d->in_w < (1LL << 31), "%s: tensor too large", who);
    return DL4JB200_OK;
}

static inline int pool_grid(dl4jb200_ctx *ctx, size_t total) {
    size_t blocks = (total + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    return (int)(blocks > cap ? cap : (blocks ? blocks : 1));
}

extern "C" {

int dl4jb200_pool_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, float *y, int32_t *argmax,
                      int layout, int x_cs, int y_cs, int pool_type, int pnorm) {
    DL4J_REQUIRE(ctx && x && y, "pool_fwd: NULL argument");
    PoolGeom g;
    int rc = pool_geom(d, layout, x_cs, y_cs, g, "pool_fwd");
    g.rnd = ctx->round_act;
    if (rc) return rc;
    DL4J_REQUIRE(pool_type >= DL4JB200_POOL_MAX && pool_type <= DL4JB200_POOL_PNORM, "pool_fwd: unknown pooling type %d", pool_type);
    DL4J_REQUIRE(pool_type != DL4JB200_POOL_PNORM || pnorm >= 1, "pool_fwd: pnorm must be >= 1");
    g.d_q = FastDiv((uint32_t)(layout == DL4JB200_NHWC ? g.ycs / 4 : 1));
    g.d_w = FastDiv((uint32_t)g.Wo); g.d_h = FastDiv((uint32_t)g.Ho); g.d_c = FastDiv((uint32_t)g.C);
    if (layout == DL4JB200_NHWC) {
        size_t total = (size_t)g.B * g.Ho * g.Wo * (g.ycs / 4);
        if (g.kH == 3 && g.kW == 3) k_pool_fwd<true, 3><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, argmax, g, pool_type, pnorm, total);
        else if (g.kH == 2 && g.kW == 2) k_pool_fwd<true, 2><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, argmax, g, pool_type, pnorm, total);
        else k_pool_fwd<true><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, argmax, g, pool_type, pnorm, total);
    } else {
        size_t total = (size_t)g.B * g.C * g.Ho * g.Wo;
        k_pool_fwd<false><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, argmax, g, pool_type, pnorm, total);
    }
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_pool_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, const float *y,
                      const float *eps, const int32_t *argmax, float *dx, int layout, int x_cs, int y_cs,
                      int pool_type, int pnorm, double pnorm_eps) {
    DL4J_REQUIRE(ctx && eps && dx, "pool_bwd: NULL argument");
    PoolGeom g;
    int rc = pool_geom(d, layout, x_cs, y_cs, g, "pool_bwd");
    g.rnd = ctx->round_act;
    if (rc) return rc;
    if (pool_type == DL4JB200_POOL_SUM) {  // SubsamplingLayer.java:251-253 throws; the helper declines
        dl4j_set_error("pool_bwd: SUM pooling has no backprop in the reference");
        return DL4JB200_NOT_SUPPORTED;
    }
    DL4J_REQUIRE(pool_type == DL4JB200_POOL_MAX || pool_type == DL4JB200_POOL_AVG || pool_type == DL4JB200_POOL_PNORM,
                 "pool_bwd: unknown pooling type %d", pool_type);
    DL4J_REQUIRE(pool_type == DL4JB200_POOL_AVG || argmax || x, "pool_bwd: x is required when argmax is NULL");
    DL4J_REQUIRE(pool_type != DL4JB200_POOL_PNORM || (x && y && pnorm >= 1), "pool_bwd: PNORM needs x, y and pnorm >= 1");
    g.d_q = FastDiv((uint32_t)(layout == DL4JB200_NHWC ? g.xcs / 4 : 1));
    g.d_w = FastDiv((uint32_t)g.W); g.d_h = FastDiv((uint32_t)g.H); g.d_c = FastDiv((uint32_t)g.C);
    if (layout == DL4JB200_NHWC && pool_type == DL4JB200_POOL_MAX && argmax && g.dH == 1 && g.dW == 1 && g.sH == 2 && g.sW == 2 &&
        g.pT >= 0 && g.pL >= 0) {
        const int Hp = (g.H + 1) / 2, Wp = (g.W + 1) / 2;
        g.d_w = FastDiv((uint32_t)Wp); g.d_h = FastDiv((uint32_t)Hp);
        size_t total = (size_t)g.B * Hp * Wp * (g.xcs / 4);
        DL4J_CUDA_TRY(dl4j_launch_pdl((g.kH <= 3 && g.kW <= 3) ? k_maxpool_bwd_nhwc_patch2<true> : k_maxpool_bwd_nhwc_patch2<false>,
                                      dim3(pool_grid(ctx, total)), dim3(256), 0, ctx->stream, eps, argmax, dx, g, Hp, Wp, total));
    } else if (layout == DL4JB200_NHWC && pool_type == DL4JB200_POOL_MAX && argmax && g.dH == 1 && g.dW == 1) {
        size_t total = (size_t)g.B * g.H * g.W * (g.xcs / 4);
        k_maxpool_bwd_nhwc_fast<<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(eps, argmax, dx, g, total);
    } else if (layout == DL4JB200_NHWC) {
        size_t total = (size_t)g.B * g.H * g.W * (g.xcs / 4);
        k_pool_bwd<true><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, eps, argmax, dx, g, pool_type, pnorm,
                                                                         (float)pnorm_eps, total);
    } else {
        size_t total = (size_t)g.B * g.C * g.H * g.W;
        k_pool_bwd<false><<<pool_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, eps, argmax, dx, g, pool_type, pnorm,
                                                                          (float)pnorm_eps, total);
    }
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

}  // extern "C"
