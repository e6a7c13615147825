// FP32 CUDA-core implicit-GEMM convolution (DL4JB200_CONV_ALGO_FP32): the parity-mode arithmetic
// (exact fp32 products, fp32 accumulate), selected the way AlgoMode.NO_WORKSPACE is in the reference's
// helper-vs-builtin tests (deeplearning4j-cuda/src/test/.../TestConvolution.java:71-151).
// 64x64x16 shared-memory tiles, 256 threads, 4x4 register micro-tiles. Not the throughput path.
#include "conv_common.cuh"

#define FB_M 64
#define FB_N 64
#define FB_K 16

// out[m, n] = act(bias[n] + sum_k gather(src)[m,k] * wpack[n,k])   (forward and backprop-data)
__global__ void __launch_bounds__(256) k_conv_ffma(const float *__restrict__ src, const float *__restrict__ wpack,
                                                   const float *__restrict__ bias, float *__restrict__ out, ConvGeom g,
                                                   int n_valid, int out_cs, int act) {
    __shared__ __align__(16) float As[FB_K][FB_M + 4];
    __shared__ __align__(16) float Bs[FB_K][FB_N + 4];
    const int t = threadIdx.x, m0 = blockIdx.x * FB_M, n0 = blockIdx.y * FB_N;
    const int lrow = t >> 2, lk = (t & 3) * 4;
    const int ty = t >> 4, tx = t & 15;
    int b, y0, x0;
    const int m = m0 + lrow;
    const bool mvalid = m < g.M;
    if (mvalid) conv_m_decode(g, m, b, y0, x0); else { b = 0; y0 = 0; x0 = 0; }
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < g.Kp; k0 += FB_K) {
        float4 av = make_float4(0, 0, 0, 0);
        {
            int kh, kw, c; bool kvalid;
            conv_k_decode(g, k0 + lk, kh, kw, c, kvalid);
            if (mvalid && kvalid) {
                long long o = conv_src_offset(g, b, y0, x0, kh, kw);
                if (o >= 0) av = *reinterpret_cast<const float4 *>(src + o + c);
            }
        }
        float4 bv = *reinterpret_cast<const float4 *>(wpack + (size_t)(n0 + lrow) * g.Kp + k0 + lk);
        __syncthreads();
        As[lk + 0][lrow] = av.x; As[lk + 1][lrow] = av.y; As[lk + 2][lrow] = av.z; As[lk + 3][lrow] = av.w;
        Bs[lk + 0][lrow] = bv.x; Bs[lk + 1][lrow] = bv.y; Bs[lk + 2][lrow] = bv.z; Bs[lk + 3][lrow] = bv.w;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < FB_K; kk++) {
            float4 a = *reinterpret_cast<const float4 *>(&As[kk][ty * 4]);
            float4 bq = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
            float ar[4] = {a.x, a.y, a.z, a.w}, br[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
    }
    const int n = n0 + tx * 4;
    if (n >= out_cs) return;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int mm = m0 + ty * 4 + i;
        if (mm >= g.M) continue;
        float r[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int nn = n + j;
            float v = 0.f;
            if (nn < n_valid) v = dl4j_act(acc[i][j] + (bias ? __ldg(bias + nn) : 0.f), act);
            r[j] = v;
        }
        *reinterpret_cast<float4 *>(out + (size_t)mm * out_cs + n) = make_float4(r[0], r[1], r[2], r[3]);
    }
}

// partial[split][co, k] = sum_{m in split} delta[m, co] * gather(x)[m, k]
__global__ void __launch_bounds__(256) k_conv_ffma_wgrad(const float *__restrict__ x, const float *__restrict__ delta,
                                                         float *__restrict__ partial, ConvGeom g, int delta_cs,
                                                         int rows_p, int m_per_split) {
    __shared__ __align__(16) float Ds[FB_K][FB_M + 4];
    __shared__ __align__(16) float As[FB_K][FB_N + 4];
    const int t = threadIdx.x, k0 = blockIdx.x * FB_N, co0 = blockIdx.y * FB_M;
    const int lr = t >> 4, lq = (t & 15) * 4;
    const int ty = t >> 4, tx = t & 15;
    int kh, kw, c; bool kvalid;
    conv_k_decode(g, k0 + lq, kh, kw, c, kvalid);
    kvalid = kvalid && (k0 + lq) < g.Kp;
    const int ms = blockIdx.z * m_per_split, me = min(g.M, ms + m_per_split);
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    for (int mb = ms; mb < me; mb += FB_K) {
        const int m = mb + lr;
        float4 dv = make_float4(0, 0, 0, 0), av = dv;
        if (m < me) {
            if (co0 + lq < delta_cs) dv = *reinterpret_cast<const float4 *>(delta + (size_t)m * delta_cs + co0 + lq);
            if (kvalid) {
                int b, y0, x0;
                conv_m_decode(g, m, b, y0, x0);
                long long o = conv_src_offset(g, b, y0, x0, kh, kw);
                if (o >= 0) av = *reinterpret_cast<const float4 *>(x + o + c);
            }
        }
        __syncthreads();
        *reinterpret_cast<float4 *>(&Ds[lr][lq]) = dv;
        *reinterpret_cast<float4 *>(&As[lr][lq]) = av;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < FB_K; kk++) {
            float4 a = *reinterpret_cast<const float4 *>(&Ds[kk][ty * 4]);
            float4 bq = *reinterpret_cast<const float4 *>(&As[kk][tx * 4]);
            float ar[4] = {a.x, a.y, a.z, a.w}, br[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
    }
    float *p = partial + (size_t)blockIdx.z * rows_p * g.Kp;
    const int kc = k0 + tx * 4;
    if (kc >= g.Kp) return;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int co = co0 + ty * 4 + i;
        if (co >= rows_p) continue;
        *reinterpret_cast<float4 *>(p + (size_t)co * g.Kp + kc) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
}

int conv_ffma_gemm(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, const float *bias,
                   float *out, int n_valid, int out_cs, int act) {
    dim3 grid(dl4j_ceil_div(g.M, FB_M), dl4j_ceil_div(out_cs, FB_N));
    k_conv_ffma<<<grid, 256, 0, ctx->stream>>>(src, wpack, bias, out, g, n_valid, out_cs, act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int conv_ffma_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs, int c_out,
                    float *partial, int rows_p, int splits) {
    int m_per_split = dl4j_round_up(dl4j_ceil_div(g.M, splits), FB_K);
    dim3 grid(dl4j_ceil_div(g.Kp, FB_N), dl4j_ceil_div(rows_p, FB_M), splits);
    k_conv_ffma_wgrad<<<grid, 256, 0, ctx->stream>>>(x, delta, partial, g, delta_cs, rows_p, m_per_split);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}
