// Cross-channel local response normalisation (LocalResponseNormalization.java:117-190).
// One thread per element; the 2*halfN+1 window reads along C hit L1/L2 (the same lines are fetched by
// neighbouring threads), so DRAM traffic stays at the compulsory 8 B/elem (fwd) / 16 B/elem (bwd).
// NCHW threads run along W (coalesced, channel stride H*W); NHWC threads run along C.
#include "common.cuh"

struct LrnGeom {
    int B, C, H, W, cs, halfN;
    float k, alpha, beta;
    int rnd;  // ctx->round_act
};

template <bool NHWC>
__device__ __forceinline__ void lrn_decode(size_t idx, const LrnGeom &g, int &c, size_t &base, size_t &cstride) {
    if (NHWC) {
        c = (int)(idx % g.cs);
        base = idx - c;  // pixel start
        cstride = 1;
    } else {
        size_t HW = (size_t)g.H * g.W;
        size_t p = idx % HW, t = idx / HW;
        c = (int)(t % g.C);
        base = (t - c) * HW + p;  // (b, 0, y, x)
        cstride = HW;
    }
}

template <bool NHWC>
__global__ void __launch_bounds__(256) k_lrn_fwd(const float *__restrict__ x, float *__restrict__ y, LrnGeom g, size_t total) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        int c; size_t base, cst;
        lrn_decode<NHWC>(idx, g, c, base, cst);
        if (NHWC && c >= g.C) { y[idx] = 0.f; continue; }
        int j0 = max(0, c - g.halfN), j1 = min(g.C - 1, c + g.halfN);
        float s = 0.f;
        for (int j = j0; j <= j1; j++) { float v = x[base + (size_t)j * cst]; s += v * v; }
        float u = g.k + g.alpha * s;
        y[idx] = dl4j_rnd(x[idx] * powf(u, -g.beta), g.rnd);
    }
}

template <bool NHWC>
__global__ void __launch_bounds__(256) k_lrn_bwd(const float *__restrict__ x, const float *__restrict__ e,
                                                 float *__restrict__ dx, LrnGeom g, size_t total) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        int c; size_t base, cst;
        lrn_decode<NHWC>(idx, g, c, base, cst);
        if (NHWC && c >= g.C) { dx[idx] = 0.f; continue; }
        // unitScale of every channel in [c-2h, c+2h] is needed: R = sum_{j in win(c)} y_j*e_j, y_j = x_j*u_j^-beta
        int j0 = max(0, c - g.halfN), j1 = min(g.C - 1, c + g.halfN);
        float R = 0.f, uc = 0.f;
        for (int j = j0; j <= j1; j++) {
            int i0 = max(0, j - g.halfN), i1 = min(g.C - 1, j + g.halfN);
            float s = 0.f;
            for (int i = i0; i <= i1; i++) { float v = x[base + (size_t)i * cst]; s += v * v; }
            float u = g.k + g.alpha * s;
            if (j == c) uc = u;
            size_t o = base + (size_t)j * cst;
            R += x[o] * powf(u, -g.beta) * e[o];
        }
        dx[idx] = dl4j_rnd(e[idx] * powf(uc, -g.beta) - 2.f * g.alpha * g.beta * x[idx] * R / uc, g.rnd);
    }
}

static int lrn_geom(int B, int C, int H, int W, int layout, int cs, double k, double n, double alpha, double beta,
                    LrnGeom &g, const char *who) {
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "%s: bad shape", who);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "%s: unknown layout %d", who, layout);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || cs >= C, "%s: cs < C", who);
    DL4J_REQUIRE(n >= 1, "%s: n must be >= 1", who);
    g = LrnGeom{B, C, H, W, layout == DL4JB200_NHWC ? cs : C, (int)n / 2, (float)k, (float)alpha, (float)beta, 0};
    return DL4JB200_OK;
}

static inline int lrn_grid(dl4jb200_ctx *ctx, size_t total) {
    size_t blocks = (total + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    return (int)(blocks > cap ? cap : (blocks ? blocks : 1));
}

extern "C" {

int dl4jb200_lrn_fwd(dl4jb200_ctx *ctx, const float *x, float *y, int B, int C, int H, int W, int layout, int cs,
                     double k, double n, double alpha, double beta) {
    DL4J_REQUIRE(ctx && x && y, "lrn_fwd: NULL argument");
    LrnGeom g;
    int rc = lrn_geom(B, C, H, W, layout, cs, k, n, alpha, beta, g, "lrn_fwd");
    if (rc) return rc;
    g.rnd = ctx->round_act;
    size_t total = (size_t)B * H * W * g.cs;
    if (layout == DL4JB200_NHWC) k_lrn_fwd<true><<<lrn_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, g, total);
    else k_lrn_fwd<false><<<lrn_grid(ctx, total), 256, 0, ctx->stream>>>(x, y, g, total);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_lrn_bwd(dl4jb200_ctx *ctx, const float *x, const float *eps, float *dx, int B, int C, int H, int W,
                     int layout, int cs, double k, double n, double alpha, double beta) {
    DL4J_REQUIRE(ctx && x && eps && dx, "lrn_bwd: NULL argument");
    LrnGeom g;
    int rc = lrn_geom(B, C, H, W, layout, cs, k, n, alpha, beta, g, "lrn_bwd");
    if (rc) return rc;
    g.rnd = ctx->round_act;
    size_t total = (size_t)B * H * W * g.cs;
    if (layout == DL4JB200_NHWC) k_lrn_bwd<true><<<lrn_grid(ctx, total), 256, 0, ctx->stream>>>(x, eps, dx, g, total);
    else k_lrn_bwd<false><<<lrn_grid(ctx, total), 256, 0, ctx->stream>>>(x, eps, dx, g, total);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

}  // extern "C"
