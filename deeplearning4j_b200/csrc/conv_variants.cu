// The layers that inherit ConvolutionLayer's helper field but run different arithmetic:
//   Deconvolution2DLayer        -> the existing tcgen05 GEMMs with the roles of input and epsilon exchanged
//   SeparableConvolution2DLayer -> a depth-wise convolution (bandwidth kernels below) followed by the 1x1 tcgen05 GEMM
// The reference runs both through libnd4j custom ops ("deconv2d", "sconv2d" and their _bp forms,
// nn/layers/convolution/Deconvolution2DLayer.java:48-233, SeparableConvolution2DLayer.java:60-246).
#include "conv_common.cuh"

// ------------------------------------------------------------------ bias + activation in place
// NHWC: z[m][cs]; NCHW: z[b][c][hw]
__global__ void __launch_bounds__(256) k_bias_act(float *__restrict__ z, const float *__restrict__ bias, size_t n, int C, int cs,
                                                  int HW, int nhwc, int act, int rnd) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int c = nhwc ? (int)(i % cs) : (int)((i / HW) % C);
        if (c >= C) continue;  // NHWC padding channels stay zero
        z[i] = dl4j_rnd(dl4j_act(z[i] + (bias ? __ldg(bias + c) : 0.f), act), rnd);
    }
}

static inline int vgrid(dl4jb200_ctx *ctx, size_t items) {
    size_t blocks = (items + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    return (int)(blocks > cap ? cap : (blocks ? blocks : 1));
}

// ------------------------------------------------------------------ depth-wise convolution, NHWC
// u[b,oh,ow, ci*mult+m] = sum_{kh,kw} x[b, oh*sH-pT+kh*dH, ow*sW-pL+kw*dW, ci] * wd[m][ci][kh][kw]
// One thread per output element; neighbouring threads are neighbouring channels (coalesced), window re-reads hit L1/L2.
struct DwGeom {
    int B, C, H, W, mult, kH, kW, sH, sW, pT, pL, dH, dW, Ho, Wo, xcs, ucs, rnd;
};

__global__ void __launch_bounds__(256) k_depthwise_fwd(const float *__restrict__ x, const float *__restrict__ wd,
                                                       float *__restrict__ u, DwGeom g) {
    const size_t total = (size_t)g.B * g.Ho * g.Wo * g.ucs, stride = (size_t)gridDim.x * blockDim.x;
    const int Cm = g.C * g.mult;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int j = (int)(i % g.ucs);
        size_t p = i / g.ucs;
        if (j >= Cm) { u[i] = 0.f; continue; }
        const int ow = (int)(p % g.Wo); p /= g.Wo;
        const int oh = (int)(p % g.Ho);
        const int b = (int)(p / g.Ho);
        const int ci = j / g.mult, m = j - ci * g.mult;
        const float *wr = wd + ((size_t)m * g.C + ci) * g.kH * g.kW;
        float acc = 0.f;
        for (int kh = 0; kh < g.kH; kh++) {
            const int ih = oh * g.sH - g.pT + kh * g.dH;
            if ((unsigned)ih >= (unsigned)g.H) continue;
            for (int kw = 0; kw < g.kW; kw++) {
                const int iw = ow * g.sW - g.pL + kw * g.dW;
                if ((unsigned)iw >= (unsigned)g.W) continue;
                acc += x[(((size_t)b * g.H + ih) * g.W + iw) * g.xcs + ci] * __ldg(wr + kh * g.kW + kw);
            }
        }
        u[i] = dl4j_rnd(acc, g.rnd);
    }
}

// dx[b,ih,iw,ci] = sum_{m,kh,kw : (ih+pT-kh*dH) % sH == 0 ...} du[b,oh,ow, ci*mult+m] * wd[m][ci][kh][kw]   (gather, no atomics)
__global__ void __launch_bounds__(256) k_depthwise_bwd_data(const float *__restrict__ du, const float *__restrict__ wd,
                                                            float *__restrict__ dx, DwGeom g) {
    const size_t total = (size_t)g.B * g.H * g.W * g.xcs, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int ci = (int)(i % g.xcs);
        size_t p = i / g.xcs;
        if (ci >= g.C) { dx[i] = 0.f; continue; }
        const int iw = (int)(p % g.W); p /= g.W;
        const int ih = (int)(p % g.H);
        const int b = (int)(p / g.H);
        float acc = 0.f;
        for (int kh = 0; kh < g.kH; kh++) {
            const int th = ih + g.pT - kh * g.dH;
            if (th < 0 || th % g.sH) continue;
            const int oh = th / g.sH;
            if (oh >= g.Ho) continue;
            for (int kw = 0; kw < g.kW; kw++) {
                const int tw = iw + g.pL - kw * g.dW;
                if (tw < 0 || tw % g.sW) continue;
                const int ow = tw / g.sW;
                if (ow >= g.Wo) continue;
                const float *dp = du + (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.ucs + ci * g.mult;
                for (int m = 0; m < g.mult; m++) acc += dp[m] * __ldg(wd + (((size_t)m * g.C + ci) * g.kH + kh) * g.kW + kw);
            }
        }
        dx[i] = dl4j_rnd(acc, g.rnd);
    }
}

// dwd[m][ci][kh][kw] = sum over output pixels of x * du. Grid (channel groups of 32, taps, S pixel ranges); block = 32
// channels x 8 pixel lanes; per-block partial rows are folded in double by k_depthwise_wgrad_final (fixed order).
__global__ void __launch_bounds__(256) k_depthwise_wgrad_partial(const float *__restrict__ x, const float *__restrict__ du,
                                                                 float *__restrict__ partial, DwGeom g) {
    __shared__ float red[8][33];
    const int Cm = g.C * g.mult, taps = g.kH * g.kW;
    const int j = blockIdx.x * 32 + (threadIdx.x & 31), lane = threadIdx.x >> 5;
    const int tap = blockIdx.y, kh = tap / g.kW, kw = tap - kh * g.kW;
    const long long M = (long long)g.B * g.Ho * g.Wo, per = (M + gridDim.z - 1) / gridDim.z;
    const long long p0 = (long long)blockIdx.z * per, p1 = p0 + per < M ? p0 + per : M;
    float acc = 0.f;
    if (j < Cm) {
        const int ci = j / g.mult;
        for (long long p = p0 + lane; p < p1; p += 8) {
            const int ow = (int)(p % g.Wo);
            const long long q = p / g.Wo;
            const int oh = (int)(q % g.Ho), b = (int)(q / g.Ho);
            const int ih = oh * g.sH - g.pT + kh * g.dH, iw = ow * g.sW - g.pL + kw * g.dW;
            if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) continue;
            acc += x[(((size_t)b * g.H + ih) * g.W + iw) * g.xcs + ci] * du[(size_t)p * g.ucs + j];
        }
    }
    red[lane][threadIdx.x & 31] = acc;
    __syncthreads();
    if (lane == 0 && j < Cm) {
        float s = 0.f;
#pragma unroll
        for (int l = 0; l < 8; l++) s += red[l][threadIdx.x & 31];
        partial[((size_t)blockIdx.z * taps + tap) * Cm + j] = s;
    }
}

__global__ void __launch_bounds__(256) k_depthwise_wgrad_final(const float *__restrict__ partial, float *__restrict__ dwd, int C,
                                                               int mult, int taps, int S) {
    const int Cm = C * mult, total = Cm * taps;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int tap = i / Cm, j = i - tap * Cm, ci = j / mult, m = j - ci * mult;
        double a = 0.0;
        for (int s = 0; s < S; s++) a += (double)partial[((size_t)s * taps + tap) * Cm + j];
        dwd[((size_t)m * C + ci) * taps + tap] = (float)a;
    }
}

static int dw_geom(const dl4jb200_conv_desc *d, int mult, int xcs, int ucs, int rnd, DwGeom &g, const char *who) {
    DL4J_REQUIRE(d && d->batch > 0 && d->c_in > 0 && d->in_h > 0 && d->in_w > 0 && d->k_h > 0 && d->k_w > 0 && d->stride_h > 0 &&
                     d->stride_w > 0 && d->dil_h > 0 && d->dil_w > 0 && d->out_h > 0 && d->out_w > 0 && mult > 0,
                 "%s: non-positive dimension", who);
    g = DwGeom{d->batch, d->c_in, d->in_h, d->in_w, mult, d->k_h, d->k_w, d->stride_h, d->stride_w, d->pad_top, d->pad_left,
               d->dil_h, d->dil_w, d->out_h, d->out_w, xcs, ucs, rnd};
    return DL4JB200_OK;
}

extern "C" {

int dl4jb200_bias_act(dl4jb200_ctx *ctx, float *z, const float *bias, int B, int C, int H, int W, int layout, int cs, int act) {
    DL4J_REQUIRE(ctx && z, "bias_act: NULL argument");
    DL4J_REQUIRE(layout == DL4JB200_NCHW || (layout == DL4JB200_NHWC && cs >= C), "bias_act: bad layout");
    if (act < 0 || act > DL4JB200_ACT_LEAKYRELU) { dl4j_set_error("bias_act: activation %d not handled", act); return DL4JB200_NOT_SUPPORTED; }
    const size_t n = (size_t)B * H * W * (layout == DL4JB200_NHWC ? cs : C);
    k_bias_act<<<vgrid(ctx, n), 256, 0, ctx->stream>>>(z, bias, n, C, cs, H * W, layout == DL4JB200_NHWC, act,
                                                       ctx->round_act && act != DL4JB200_ACT_IDENTITY);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// The convolution whose backprop-data pass IS this transposed convolution: it maps the deconvolution's OUTPUT
// [B, nOut, out_h, out_w] to its INPUT [B, nIn, in_h, in_w] with the same kernel / stride / padding / dilation, and
// its filter layout [c_out = nIn][c_in = nOut][kH][kW] is exactly DeconvolutionParamInitializer's [nIn, nOut, kH, kW].
static dl4jb200_conv_desc deconv_as_conv(const dl4jb200_conv_desc *d) {
    dl4jb200_conv_desc c = *d;
    c.c_in = d->c_out; c.in_h = d->out_h; c.in_w = d->out_w;
    c.c_out = d->c_in; c.out_h = d->in_h; c.out_w = d->in_w;
    return c;
}

int dl4jb200_deconv2d_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs, const float *w,
                          const float *bias, float *z, int z_layout, int z_cs, int act, int algo) {
    DL4J_REQUIRE(ctx && d && x && w && z, "deconv2d_fwd: NULL argument");
    DL4J_REQUIRE(d->out_h == d->stride_h * (d->in_h - 1) + d->k_h + (d->k_h - 1) * (d->dil_h - 1) - 2 * d->pad_top &&
                     d->out_w == d->stride_w * (d->in_w - 1) + d->k_w + (d->k_w - 1) * (d->dil_w - 1) - 2 * d->pad_left,
                 "deconv2d_fwd: out_h/out_w must be stride*(in-1) + kEff - 2*pad (ConvolutionUtils.getDeconvolutionOutputSize)");
    const dl4jb200_conv_desc c = deconv_as_conv(d);
    // z = backprop-data of that convolution applied to x (x plays the epsilon; the convolution's own input is not read)
    int rc = dl4jb200_conv_bwd(ctx, &c, z, z_layout, z_cs, w, x, x_layout, x_cs, nullptr, nullptr, z, z_layout, z_cs, algo);
    if (rc) return rc;
    if (!bias && act == DL4JB200_ACT_IDENTITY) return DL4JB200_OK;
    return dl4jb200_bias_act(ctx, z, bias, d->batch, d->c_out, d->out_h, d->out_w, z_layout, z_cs, act);
}

int dl4jb200_deconv2d_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs, const float *w,
                          const float *delta, int delta_layout, int delta_cs, float *dw, float *db, float *dx, int dx_layout,
                          int dx_cs, int algo) {
    DL4J_REQUIRE(ctx && d && x && w && delta, "deconv2d_bwd: NULL argument");
    DL4J_REQUIRE(dw || db || dx, "deconv2d_bwd: nothing to compute");
    const dl4jb200_conv_desc c = deconv_as_conv(d);
    int rc;
    // dW[nIn][nOut][kh][kw]: the convolution's filter gradient with delta as its input and x as its epsilon
    if (dw && (rc = dl4jb200_conv_bwd(ctx, &c, delta, delta_layout, delta_cs, w, x, x_layout, x_cs, dw, nullptr, nullptr, DL4JB200_NHWC,
                                      0, algo)))
        return rc;
    if (db) {  // sum of delta over batch and pixels
        const float *dn = delta;
        int cs = delta_cs;
        if (delta_layout == DL4JB200_NCHW) {
            cs = dl4j_round_up(d->c_out, 4);
            float *t = (float *)dl4j_workspace(ctx, 3, (size_t)d->batch * d->out_h * d->out_w * cs * sizeof(float));
            if (!t) return DL4JB200_ERR_CUDA;
            if ((rc = dl4jb200_nchw_to_nhwc(ctx, delta, t, d->batch, d->c_out, d->out_h, d->out_w, cs))) return rc;
            dn = t;
        }
        if ((rc = dl4jb200_bias_grad(ctx, dn, d->batch * d->out_h * d->out_w, cs, d->c_out, db))) return rc;
    }
    // epsilon for the layer below: the forward pass of that convolution applied to delta
    if (dx) return dl4jb200_conv_fwd(ctx, &c, delta, delta_layout, delta_cs, w, nullptr, dx, dx_layout, dx_cs, DL4JB200_ACT_IDENTITY, algo);
    return DL4JB200_OK;
}

// ---- depth-wise stage of SeparableConvolution2D (NHWC, any pixel stride >= channels)
int dl4jb200_depthwise_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int depth_mult, const float *x, int x_cs, const float *wd,
                           float *u, int u_cs) {
    DL4J_REQUIRE(ctx && x && wd && u, "depthwise_fwd: NULL argument");
    DwGeom g;
    int rc = dw_geom(d, depth_mult, x_cs, u_cs, ctx->round_act, g, "depthwise_fwd");
    if (rc) return rc;
    DL4J_REQUIRE(x_cs >= d->c_in && u_cs >= d->c_in * depth_mult, "depthwise_fwd: pixel stride below the channel count");
    k_depthwise_fwd<<<vgrid(ctx, (size_t)g.B * g.Ho * g.Wo * g.ucs), 256, 0, ctx->stream>>>(x, wd, u, g);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_depthwise_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int depth_mult, const float *x, int x_cs, const float *wd,
                           const float *du, int u_cs, float *dwd, float *dx, int dx_cs) {
    DL4J_REQUIRE(ctx && x && wd && du && (dwd || dx), "depthwise_bwd: NULL argument");
    DwGeom g;
    int rc = dw_geom(d, depth_mult, x_cs, u_cs, ctx->round_act, g, "depthwise_bwd");
    if (rc) return rc;
    DL4J_REQUIRE(x_cs >= d->c_in && u_cs >= d->c_in * depth_mult && (!dx || dx_cs >= d->c_in), "depthwise_bwd: pixel stride below the channel count");
    if (dwd) {
        const int Cm = g.C * g.mult, taps = g.kH * g.kW;
        const long long M = (long long)g.B * g.Ho * g.Wo;
        int S = (int)(M / 256);
        const int cap = dl4j_ceil_div(4 * ctx->sm_count, dl4j_ceil_div(Cm, 32) * taps);
        if (S > cap) S = cap;
        if (S < 1) S = 1;
        float *partial = (float *)dl4j_workspace(ctx, 0, (size_t)S * taps * Cm * sizeof(float));
        if (!partial) return DL4JB200_ERR_CUDA;
        k_depthwise_wgrad_partial<<<dim3(dl4j_ceil_div(Cm, 32), taps, S), 256, 0, ctx->stream>>>(x, du, partial, g);
        DL4J_CHECK_LAUNCH();
        k_depthwise_wgrad_final<<<dl4j_ceil_div(Cm * taps, 256), 256, 0, ctx->stream>>>(partial, dwd, g.C, g.mult, taps, S);
        DL4J_CHECK_LAUNCH();
    }
    if (dx) {
        DwGeom gx = g;
        gx.xcs = dx_cs;
        k_depthwise_bwd_data<<<vgrid(ctx, (size_t)g.B * g.H * g.W * dx_cs), 256, 0, ctx->stream>>>(du, wd, dx, gx);
        DL4J_CHECK_LAUNCH();
    }
    return DL4JB200_OK;
}

}  // extern "C"
