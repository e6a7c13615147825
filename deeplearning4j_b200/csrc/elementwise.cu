// HBM-bandwidth kernels without reductions: re-layout, activations, residual add, epsilon fan-in,
// softmax-NLL output delta, updater. All vectorised 128-bit where alignment allows; grids are sized
// in multiples of the SM count with a grid-stride loop.
#include "common.cuh"

static inline int ew_grid(dl4jb200_ctx *ctx, size_t work_items, int threads) {
    size_t blocks = (work_items + threads - 1) / threads;
    size_t cap = (size_t)ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ---------------------------------------------------------------- NCHW <-> NHWC
// One block transposes a 32(channels) x 32(pixels) tile of one image through shared memory so that both
// the global read (contiguous pixels) and the global write (contiguous channels) are coalesced.
__global__ void __launch_bounds__(256) k_nchw_to_nhwc(const float *__restrict__ src, float *__restrict__ dst, int C,
                                                      int HW, int cs, int rnd) {
    __shared__ float tile[32][33];
    int b = blockIdx.z;
    int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const float *s = src + (size_t)b * C * HW;
    float *d = dst + (size_t)b * HW * cs;
    int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int c = c0 + ty + i * 8, p = p0 + tx;
        tile[ty + i * 8][tx] = (c < C && p < HW) ? s[(size_t)c * HW + p] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int p = p0 + ty + i * 8, c = c0 + tx;
        if (p < HW && c < cs) d[(size_t)p * cs + c] = dl4j_rnd(tile[tx][ty + i * 8], rnd);
    }
}

// C <= 4 into 4-float pixels (the network input: 3 x 224 x 224 images): one thread per pixel reads its <= 4 planes (coalesced
// across the warp) and writes one 128-bit pixel. The 32 x 32 tile transpose above spends 29 of its 32 channel lanes on nothing
// here (AlexNet batch 128: 299 us -> a copy).
__global__ void __launch_bounds__(256) k_nchw_to_nhwc4(const float *__restrict__ src, float *__restrict__ dst, int C, int HW, size_t total,
                                                       int rnd) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t b = i / HW, p = i - b * HW;
        const float *s = src + b * C * HW + p;
        float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 4; c++)
            if (c < C) v[c] = dl4j_rnd(s[(size_t)c * HW], rnd);
        reinterpret_cast<float4 *>(dst)[i] = make_float4(v[0], v[1], v[2], v[3]);
    }
}

__global__ void __launch_bounds__(256) k_nhwc_to_nchw(const float *__restrict__ src, float *__restrict__ dst, int C,
                                                      int HW, int cs) {
    __shared__ float tile[32][33];
    int b = blockIdx.z;
    int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const float *s = src + (size_t)b * HW * cs;
    float *d = dst + (size_t)b * C * HW;
    int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int p = p0 + ty + i * 8, c = c0 + tx;
        tile[ty + i * 8][tx] = (p < HW && c < C) ? s[(size_t)p * cs + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int c = c0 + ty + i * 8, p = p0 + tx;
        if (c < C && p < HW) d[(size_t)c * HW + p] = tile[tx][ty + i * 8];
    }
}

// ---------------------------------------------------------------- activations
template <bool VEC>
__global__ void __launch_bounds__(256) k_act_fwd(const float *__restrict__ z, float *__restrict__ a, size_t n, int act, int rnd) {
    size_t stride = (size_t)gridDim.x * blockDim.x, i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        size_t n4 = n >> 2;
        for (; i < n4; i += stride) {
            float4 v = reinterpret_cast<const float4 *>(z)[i];
            v.x = dl4j_act(v.x, act); v.y = dl4j_act(v.y, act); v.z = dl4j_act(v.z, act); v.w = dl4j_act(v.w, act);
            reinterpret_cast<float4 *>(a)[i] = dl4j_rnd4(v, rnd);
        }
    } else {
        for (; i < n; i += stride) a[i] = dl4j_rnd(dl4j_act(z[i], act), rnd);
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256) k_act_bwd(const float *__restrict__ z, const float *__restrict__ e,
                                                 float *__restrict__ d, size_t n, int act, int rnd) {
    const bool from_out = (act & DL4JB200_ACT_FROM_OUTPUT) != 0;  // z holds act(z), not z
    act &= 0xff;
    auto grad = [&](float v) { return from_out ? dl4j_act_grad_out(v, act) : dl4j_act_grad(v, act); };
    size_t stride = (size_t)gridDim.x * blockDim.x, i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        size_t n4 = n >> 2;
        for (; i < n4; i += stride) {
            float4 v = reinterpret_cast<const float4 *>(z)[i], g = reinterpret_cast<const float4 *>(e)[i];
            g.x *= grad(v.x); g.y *= grad(v.y); g.z *= grad(v.z); g.w *= grad(v.w);
            reinterpret_cast<float4 *>(d)[i] = dl4j_rnd4(g, rnd);
        }
    } else {
        for (; i < n; i += stride) d[i] = dl4j_rnd(e[i] * grad(z[i]), rnd);
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256) k_add_act(const float *__restrict__ a, const float *__restrict__ b,
                                                 float *__restrict__ o, size_t n, int act, int rnd) {
    size_t stride = (size_t)gridDim.x * blockDim.x, i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        size_t n4 = n >> 2;
        for (; i < n4; i += stride) {
            float4 x = reinterpret_cast<const float4 *>(a)[i], y = reinterpret_cast<const float4 *>(b)[i];
            x.x = dl4j_act(x.x + y.x, act); x.y = dl4j_act(x.y + y.y, act);
            x.z = dl4j_act(x.z + y.z, act); x.w = dl4j_act(x.w + y.w, act);
            reinterpret_cast<float4 *>(o)[i] = dl4j_rnd4(x, rnd);
        }
    } else {
        for (; i < n; i += stride) o[i] = dl4j_rnd(dl4j_act(a[i] + b[i], act), rnd);
    }
}

// inverted dropout with a caller-supplied 0/1 mask: y = x * mask / p   (nd4j DropOutInverted as Dropout.applyDropout uses it)
template <bool VEC>
__global__ void __launch_bounds__(256) k_dropout(const float *__restrict__ x, const float *__restrict__ m, float *__restrict__ y,
                                                 size_t n, float inv_p, int rnd) {
    size_t stride = (size_t)gridDim.x * blockDim.x, i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        size_t n4 = n >> 2;
        for (; i < n4; i += stride) {
            float4 v = reinterpret_cast<const float4 *>(x)[i], k = reinterpret_cast<const float4 *>(m)[i];
            v.x = v.x * k.x * inv_p; v.y = v.y * k.y * inv_p; v.z = v.z * k.z * inv_p; v.w = v.w * k.w * inv_p;
            reinterpret_cast<float4 *>(y)[i] = dl4j_rnd4(v, rnd);
        }
    } else {
        for (; i < n; i += stride) y[i] = dl4j_rnd(x[i] * m[i] * inv_p, rnd);
    }
}

// epsilon fan-in + activation backprop in one pass: d = (a + b) * act'(y)
template <bool VEC>
__global__ void __launch_bounds__(256) k_sum_act_bwd(const float *__restrict__ a, const float *__restrict__ b,
                                                     const float *__restrict__ y, float *__restrict__ d, size_t n, int act,
                                                     int rnd) {
    size_t stride = (size_t)gridDim.x * blockDim.x, i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        size_t n4 = n >> 2;
        for (; i < n4; i += stride) {
            float4 p = reinterpret_cast<const float4 *>(a)[i], q = reinterpret_cast<const float4 *>(b)[i];
            float4 v = reinterpret_cast<const float4 *>(y)[i];
            p.x = (p.x + q.x) * dl4j_act_grad(v.x, act); p.y = (p.y + q.y) * dl4j_act_grad(v.y, act);
            p.z = (p.z + q.z) * dl4j_act_grad(v.z, act); p.w = (p.w + q.w) * dl4j_act_grad(v.w, act);
            reinterpret_cast<float4 *>(d)[i] = dl4j_rnd4(p, rnd);
        }
    } else {
        for (; i < n; i += stride) d[i] = dl4j_rnd((a[i] + b[i]) * dl4j_act_grad(y[i], act), rnd);
    }
}

static inline bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

// ---------------------------------------------------------------- softmax + NLL (one warp per row)
__global__ void __launch_bounds__(256) k_softmax_nll(const float *__restrict__ z, int z_ld, const float *__restrict__ labels,
                                                     float *__restrict__ delta, int d_ld, float *loss_rows, int B, int K,
                                                     int rnd) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= B) return;
    const float *zr = z + (size_t)row * z_ld;
    const float *lr = labels + (size_t)row * K;
    float m = -INFINITY;
    for (int k = lane; k < K; k += 32) m = fmaxf(m, zr[k]);
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int k = lane; k < K; k += 32) s += expf(zr[k] - m);
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float logs = logf(s), loss = 0.f, inv = 1.f / s;
    for (int k = lane; k < K; k += 32) {
        float zz = zr[k] - m, l = lr[k];
        delta[(size_t)row * d_ld + k] = dl4j_rnd(expf(zz) * inv - l, rnd);
        if (l != 0.f) loss -= l * (zz - logs);
    }
    for (int k = K + lane; k < d_ld; k += 32) delta[(size_t)row * d_ld + k] = 0.f;
    for (int o = 16; o; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
    if (lane == 0) loss_rows[row] = loss;
}

// deterministic single-block sum of the per-row losses
__global__ void __launch_bounds__(256) k_sum_rows(const float *__restrict__ v, float *out, int n) {
    __shared__ double sm[256];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += (double)v[i];
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = (float)sm[0];
}

// score regularisation terms: per-block partials, then a deterministic single-block fold into score_acc
__global__ void __launch_bounds__(256) k_l1l2_partial(const float *__restrict__ w, size_t n, double *partial) {
    __shared__ double s1[256], s2[256];
    double a = 0.0, b = 0.0;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float v = w[i];
        a += fabsf(v);
        b += (double)v * (double)v;
    }
    s1[threadIdx.x] = a; s2[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[2 * blockIdx.x] = s1[0]; partial[2 * blockIdx.x + 1] = s2[0]; }
}
__global__ void __launch_bounds__(128) k_l1l2_final(const double *partial, int nblk, double l1, double l2, float *score_acc) {
    __shared__ double s1[128], s2[128];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblk; i += 128) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    s1[threadIdx.x] = a; s2[threadIdx.x] = b;
    __syncthreads();
    for (int o = 64; o; o >>= 1) {
        if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) score_acc[0] += (float)(l1 * s1[0] + 0.5 * l2 * s2[0]);
}

__global__ void __launch_bounds__(256) k_copy2d(const float *__restrict__ src, int src_ld, float *__restrict__ dst, int dst_ld,
                                                int rows, int cols) {
    size_t total = (size_t)rows * dst_ld, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int r = (int)(i / dst_ld), c = (int)(i % dst_ld);
        dst[i] = c < cols ? src[(size_t)r * src_ld + c] : 0.f;
    }
}

// ---------------------------------------------------------------- updater (UpdaterBlock.java:142,170-193)
__global__ void __launch_bounds__(256) k_updater(float *__restrict__ w, const float *__restrict__ g,
                                                 float *__restrict__ state, size_t n, int kind, float lr, float h1,
                                                 float h2, float l1, float l2, float inv_batch) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float gi = g[i], wi = w[i], u;
        switch (kind) {
        case DL4JB200_UPD_NESTEROVS: {
            float v = state[i], vn = h1 * v - lr * gi;
            u = h1 * v - (1.f + h1) * vn;
            state[i] = vn;
        } break;
        case DL4JB200_UPD_RMSPROP: {
            float c = h1 * state[i] + (1.f - h1) * gi * gi;
            state[i] = c;
            u = lr * gi / sqrtf(c + h2);
        } break;
        case DL4JB200_UPD_ADADELTA: {
            float msg = h1 * state[i] + (1.f - h1) * gi * gi, msdx = state[n + i];
            u = gi * sqrtf(msdx + h2) / sqrtf(msg + h2);
            state[i] = msg;
            state[n + i] = h1 * msdx + (1.f - h1) * u * u;
        } break;
        case DL4JB200_UPD_NOOP: u = 0.f; break;
        default: u = lr * gi;
        }
        if (l2 > 0.f) u += l2 * wi;
        if (l1 > 0.f) u += l1 * (wi > 0.f ? 1.f : (wi < 0.f ? -1.f : 0.f));
        w[i] = wi - u * inv_batch;
    }
}

// ---------------------------------------------------------------- whole-vector updater (all UpdaterBlocks in one launch)
// The flat parameter vector is a sequence of segments (one per parameter array), each with its own updater
// configuration. One launch walks the whole vector; a thread finds its segment by binary search over the
// segment starts, applies updater + L1/L2 + minibatch division, and the L1/L2 score terms of the regularised
// segments are reduced per block (double) for a second, tiny launch to fold into the score.
struct UpdSegCfg {  // per-configuration constants (few distinct ones per network)
    int kind;
    float lr, h1, h2, l1, l2;
};
#define DL4J_MAX_UPD_CFGS 8
struct UpdCfgTable { UpdSegCfg c[DL4J_MAX_UPD_CFGS]; };

// one element of the segmented updater; returns the new parameter value (state updated in place)
__device__ __forceinline__ float upd_one(const UpdSegCfg &c, float wi, float gi, float *state, long long i, long long s0,
                                         long long len, float inv_batch, double &score) {
    float u;
    switch (c.kind) {
    case DL4JB200_UPD_NESTEROVS: {
        float v = state[i], vn = c.h1 * v - c.lr * gi;
        u = c.h1 * v - (1.f + c.h1) * vn;
        state[i] = vn;
    } break;
    case DL4JB200_UPD_RMSPROP: {
        float r = c.h1 * state[i] + (1.f - c.h1) * gi * gi;
        state[i] = r;
        u = c.lr * gi / sqrtf(r + c.h2);
    } break;
    case DL4JB200_UPD_ADADELTA: {
        // block-local msg | msdx layout: state[2*start + j] and state[2*start + len + j]
        const long long j = i - s0;
        float msg = c.h1 * state[2 * s0 + j] + (1.f - c.h1) * gi * gi, msdx = state[2 * s0 + len + j];
        u = gi * sqrtf(msdx + c.h2) / sqrtf(msg + c.h2);
        state[2 * s0 + j] = msg;
        state[2 * s0 + len + j] = c.h1 * msdx + (1.f - c.h1) * u * u;
    } break;
    case DL4JB200_UPD_NOOP: u = 0.f; break;
    default: u = c.lr * gi;
    }
    if (c.l2 > 0.f) { u += c.l2 * wi; score += 0.5 * (double)c.l2 * (double)wi * (double)wi; }
    if (c.l1 > 0.f) { u += c.l1 * (wi > 0.f ? 1.f : (wi < 0.f ? -1.f : 0.f)); score += (double)c.l1 * fabs((double)wi); }
    return wi - u * inv_batch;
}

// dst[off_i + j] = scale * src[off_i + j] over a table of spans (one block per span): moves the BatchNormalization running
// statistics -- scattered through the flat parameter vector -- into the (always zero) slots their gradients occupy in the
// flat gradient vector and back, so that ONE all-reduce of the gradient also averages them across replicas.
__global__ void __launch_bounds__(256) k_spans_scale_copy(float *__restrict__ dst, const float *__restrict__ src,
                                                          const long long *__restrict__ off, const long long *__restrict__ cnt,
                                                          float scale) {
    const long long o = off[blockIdx.x], n = cnt[blockIdx.x];
    for (long long i = threadIdx.x; i < n; i += blockDim.x) dst[o + i] = scale == 0.f ? 0.f : scale * src[o + i];
}

// A thread owns 4 consecutive elements: one segment search per quad, and when the quad lies inside one segment of a
// stateless / flat-state updater (everything but AdaDelta's block-local layout) the five streams move as 128-bit accesses.
__global__ void __launch_bounds__(256) k_updater_segments(float *__restrict__ w, const float *__restrict__ g,
                                                          float *__restrict__ state, long long n,
                                                          const long long *__restrict__ seg_start,
                                                          const int *__restrict__ seg_cfg, int nseg, UpdCfgTable tab,
                                                          float inv_batch, int vec_ok, double *score_partial, long long q_lo,
                                                          long long q_hi) {
    __shared__ double sred[256];
    double score = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;  // quads [q_lo, q_hi) of the flat vector
    for (long long qi = q_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; qi < q_hi; qi += stride) {
        const long long i0 = qi << 2;
        int lo = 0, hi = nseg - 1;  // last segment whose start <= i0
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (__ldg(seg_start + mid) <= i0) lo = mid; else hi = mid - 1;
        }
        const long long s0 = __ldg(seg_start + lo), s1 = __ldg(seg_start + lo + 1);
        const int ci = __ldg(seg_cfg + lo);
        if (i0 + 4 <= s1 && i0 + 4 <= n && vec_ok && (ci < 0 || tab.c[ci].kind != DL4JB200_UPD_ADADELTA)) {
            if (ci < 0) continue;  // NoOp segment (BatchNormalization running statistics)
            const UpdSegCfg c = tab.c[ci];
            float4 wv = *reinterpret_cast<const float4 *>(w + i0);
            const float4 gv = *reinterpret_cast<const float4 *>(g + i0);
            const bool has_state = c.kind == DL4JB200_UPD_NESTEROVS || c.kind == DL4JB200_UPD_RMSPROP;
            float st[4] = {0.f, 0.f, 0.f, 0.f};
            if (has_state) {
                const float4 sv = *reinterpret_cast<const float4 *>(state + i0);
                st[0] = sv.x; st[1] = sv.y; st[2] = sv.z; st[3] = sv.w;
            }
            // upd_one indexes state[i]: hand it the register copy with i = lane index
            wv.x = upd_one(c, wv.x, gv.x, st, 0, 0, 0, inv_batch, score);
            wv.y = upd_one(c, wv.y, gv.y, st, 1, 0, 0, inv_batch, score);
            wv.z = upd_one(c, wv.z, gv.z, st, 2, 0, 0, inv_batch, score);
            wv.w = upd_one(c, wv.w, gv.w, st, 3, 0, 0, inv_batch, score);
            if (has_state) *reinterpret_cast<float4 *>(state + i0) = make_float4(st[0], st[1], st[2], st[3]);
            *reinterpret_cast<float4 *>(w + i0) = wv;
        } else {
            int seg = lo;
            for (long long i = i0; i < i0 + 4 && i < n; i++) {
                while (seg + 1 < nseg && __ldg(seg_start + seg + 1) <= i) seg++;
                const int cj = __ldg(seg_cfg + seg);
                if (cj < 0) continue;
                const long long t0 = __ldg(seg_start + seg), len = __ldg(seg_start + seg + 1) - t0;
                w[i] = upd_one(tab.c[cj], w[i], g[i], state, i, t0, len, inv_batch, score);
            }
        }
    }
    if (score_partial) {
        sred[threadIdx.x] = score;
        __syncthreads();
        for (int o = 128; o; o >>= 1) {
            if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) score_partial[blockIdx.x] = sred[0];
    }
}
__global__ void __launch_bounds__(256) k_score_fold(const double *partial, int nblk, float *score_out, int accumulate) {
    __shared__ double s[256];
    double a = 0.0;
    for (int i = threadIdx.x; i < nblk; i += 256) a += partial[i];
    s[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) score_out[0] = (float)(s[0] + (accumulate ? (double)score_out[0] : 0.0));
}

// ---------------------------------------------------------------- gradient normalisation (BaseMultiLayerUpdater.preApply)
// (a gradient block starts anywhere in the flat vector: scalar head up to the first 16-byte boundary, 128-bit body with two
// loads in flight and four independent double accumulators per thread, scalar tail -- the one-float-per-iteration form ran
// at 1.3 TB/s on its dependent double-add chain)
__global__ void __launch_bounds__(256) k_sumsq_partial(const float *__restrict__ g, size_t n, double *partial) {
    __shared__ double s[256];
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    size_t head = ((16 - ((uintptr_t)g & 15)) & 15) >> 2;
    if (head > n) head = n;
    const size_t n4 = (n - head) >> 2, tail0 = head + (n4 << 2);
    const float4 *g4 = reinterpret_cast<const float4 *>(g + head);
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    size_t i = tid;
    for (; i + stride < n4; i += 2 * stride) {
        const float4 u = g4[i], w = g4[i + stride];
        a0 += (double)u.x * u.x + (double)w.x * w.x; a1 += (double)u.y * u.y + (double)w.y * w.y;
        a2 += (double)u.z * u.z + (double)w.z * w.z; a3 += (double)u.w * u.w + (double)w.w * w.w;
    }
    if (i < n4) {
        const float4 u = g4[i];
        a0 += (double)u.x * u.x; a1 += (double)u.y * u.y; a2 += (double)u.z * u.z; a3 += (double)u.w * u.w;
    }
    if (tid < head) { const double v = g[tid]; a0 += v * v; }
    if (tail0 + tid < n) { const double v = g[tail0 + tid]; a1 += v * v; }  // < 4 elements
    const double a = (a0 + a1) + (a2 + a3);
    s[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}
// mode 1 RenormalizeL2PerLayer: g /= ||g||2 ; mode 2 ClipL2PerLayer: g *= thr/||g||2 if ||g||2 > thr ; mode 3 clip elementwise
__global__ void __launch_bounds__(256) k_grad_scale(float *__restrict__ g, size_t n, const double *partial, int nblk, int mode,
                                                    float thr) {
    __shared__ float scale_s;
    if (mode != 3) {
        __shared__ double s[256];
        double a = 0.0;
        for (int i = threadIdx.x; i < nblk; i += 256) a += partial[i];
        s[threadIdx.x] = a;
        __syncthreads();
        for (int o = 128; o; o >>= 1) {
            if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            double l2 = sqrt(s[0]);
            scale_s = mode == 1 ? (float)(1.0 / l2) : (l2 > (double)thr ? (float)((double)thr / l2) : 1.f);
        }
        __syncthreads();
    }
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        g[i] = mode == 3 ? fminf(fmaxf(g[i], -thr), thr) : g[i] * scale_s;
}

extern "C" {

int dl4jb200_grad_normalize(dl4jb200_ctx *ctx, float *g, size_t n, int mode, double threshold) {
    DL4J_REQUIRE(ctx && g, "grad_normalize: NULL argument");
    DL4J_REQUIRE(mode >= 1 && mode <= 3, "grad_normalize: mode 1 (RenormalizeL2PerLayer), 2 (ClipL2PerLayer) or 3 (ClipElementWiseAbsoluteValue)");
    if (n == 0) return DL4JB200_OK;
    int nblk = ew_grid(ctx, n, 256);
    if (nblk > 256) nblk = 256;
    double *partial = (double *)dl4j_workspace(ctx, 0, (size_t)nblk * sizeof(double));
    if (!partial) return DL4JB200_ERR_CUDA;
    if (mode != 3) {
        k_sumsq_partial<<<nblk, 256, 0, ctx->stream>>>(g, n, partial);
        DL4J_CHECK_LAUNCH();
    }
    k_grad_scale<<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(g, n, partial, nblk, mode, (float)threshold);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_updater_apply_segments_range(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n,
                                          const int64_t *seg_start_dev, const int32_t *seg_cfg_dev, int nseg,
                                          const double *cfg_host, int ncfg, double batch_div, float *score_reg_out, size_t lo,
                                          size_t hi, int score_accumulate) {
    DL4J_REQUIRE(ctx && w && g && seg_start_dev && seg_cfg_dev && cfg_host, "updater_apply_segments: NULL argument");
    DL4J_REQUIRE(nseg > 0 && ncfg > 0 && ncfg <= DL4J_MAX_UPD_CFGS, "updater_apply_segments: 1..%d configurations supported", DL4J_MAX_UPD_CFGS);
    DL4J_REQUIRE(batch_div > 0, "updater_apply_segments: batch_div must be positive");
    DL4J_REQUIRE(lo <= hi && hi <= n && lo % 4 == 0 && (hi % 4 == 0 || hi == n), "updater_apply_segments: range [%zu, %zu) must lie in [0, n] on multiples of 4", lo, hi);
    UpdCfgTable tab;
    int any_state = 0;
    for (int i = 0; i < ncfg; i++) {
        const double *c = cfg_host + 6 * i;
        tab.c[i] = UpdSegCfg{(int)c[0], (float)c[1], (float)c[2], (float)c[3], (float)c[4], (float)c[5]};
        DL4J_REQUIRE(tab.c[i].kind >= 0 && tab.c[i].kind <= DL4JB200_UPD_NOOP, "updater_apply_segments: unknown updater %d", tab.c[i].kind);
        if (tab.c[i].kind != DL4JB200_UPD_SGD && tab.c[i].kind != DL4JB200_UPD_NOOP) any_state = 1;
    }
    DL4J_REQUIRE(state || !any_state, "updater_apply_segments: state is NULL");
    if (hi == lo) return DL4JB200_OK;
    const long long q_lo = (long long)(lo >> 2), q_hi = (long long)((hi + 3) >> 2);
    int nblk = ew_grid(ctx, (size_t)(q_hi - q_lo), 256);
    double *partial = nullptr;
    if (score_reg_out) {
        partial = (double *)dl4j_workspace(ctx, 0, (size_t)nblk * sizeof(double));
        if (!partial) return DL4JB200_ERR_CUDA;
    }
    const int vec_ok = aligned16(w) && aligned16(g) && (!state || aligned16(state));
    k_updater_segments<<<nblk, 256, 0, ctx->stream>>>(w, g, state, (long long)n, (const long long *)seg_start_dev, seg_cfg_dev, nseg,
                                                      tab, (float)(1.0 / batch_div), vec_ok, partial, q_lo, q_hi);
    DL4J_CHECK_LAUNCH();
    if (score_reg_out) {
        k_score_fold<<<1, 256, 0, ctx->stream>>>(partial, nblk, score_reg_out, score_accumulate);
        DL4J_CHECK_LAUNCH();
    }
    return DL4JB200_OK;
}

int dl4jb200_updater_apply_segments(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n,
                                    const int64_t *seg_start_dev, const int32_t *seg_cfg_dev, int nseg,
                                    const double *cfg_host, int ncfg, double batch_div, float *score_reg_out) {
    return dl4jb200_updater_apply_segments_range(ctx, w, g, state, n, seg_start_dev, seg_cfg_dev, nseg, cfg_host, ncfg, batch_div,
                                                 score_reg_out, 0, n, 0);
}

int dl4jb200_spans_scale_copy(dl4jb200_ctx *ctx, float *dst, const float *src, const long long *offsets_dev,
                              const long long *counts_dev, int nspans, double scale) {
    DL4J_REQUIRE(ctx && dst && src && (nspans == 0 || (offsets_dev && counts_dev)), "spans_scale_copy: NULL argument");
    if (nspans == 0) return DL4JB200_OK;
    k_spans_scale_copy<<<nspans, 256, 0, ctx->stream>>>(dst, src, offsets_dev, counts_dev, (float)scale);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_nchw_to_nhwc(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int cs) {
    DL4J_REQUIRE(ctx && src && dst, "nchw_to_nhwc: NULL argument");
    DL4J_REQUIRE(cs >= C && B > 0 && C > 0 && H > 0 && W > 0, "nchw_to_nhwc: bad shape");
    DL4J_REQUIRE(B <= 65535, "nchw_to_nhwc: batch > 65535");
    int HW = H * W;
    if (cs == 4 && aligned16(dst)) {
        const size_t total = (size_t)B * HW;
        k_nchw_to_nhwc4<<<ew_grid(ctx, total, 256), 256, 0, ctx->stream>>>(src, dst, C, HW, total, ctx->round_act);
        DL4J_CHECK_LAUNCH();
        return DL4JB200_OK;
    }
    dim3 grid(dl4j_ceil_div(HW, 32), dl4j_ceil_div(cs, 32), B);
    k_nchw_to_nhwc<<<grid, 256, 0, ctx->stream>>>(src, dst, C, HW, cs, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_nhwc_to_nchw(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int cs) {
    DL4J_REQUIRE(ctx && src && dst, "nhwc_to_nchw: NULL argument");
    DL4J_REQUIRE(cs >= C && B > 0 && C > 0 && H > 0 && W > 0, "nhwc_to_nchw: bad shape");
    DL4J_REQUIRE(B <= 65535, "nhwc_to_nchw: batch > 65535");
    int HW = H * W;
    dim3 grid(dl4j_ceil_div(HW, 32), dl4j_ceil_div(C, 32), B);
    k_nhwc_to_nchw<<<grid, 256, 0, ctx->stream>>>(src, dst, C, HW, cs);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_act_fwd(dl4jb200_ctx *ctx, const float *z, float *a, size_t n, int act) {
    DL4J_REQUIRE(ctx && z && a, "act_fwd: NULL argument");
    if (act < 0 || act > DL4JB200_ACT_LEAKYRELU) { dl4j_set_error("act_fwd: activation %d not handled by the helper", act); return DL4JB200_NOT_SUPPORTED; }
    if (n == 0) return DL4JB200_OK;
    if ((n & 3) == 0 && aligned16(z) && aligned16(a))
        k_act_fwd<true><<<ew_grid(ctx, n / 4, 256), 256, 0, ctx->stream>>>(z, a, n, act, ctx->round_act);
    else
        k_act_fwd<false><<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(z, a, n, act, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_act_bwd(dl4jb200_ctx *ctx, const float *z, const float *eps, float *delta, size_t n, int act) {
    DL4J_REQUIRE(ctx && z && eps && delta, "act_bwd: NULL argument");
    if (act < 0 || (act & 0xff) > DL4JB200_ACT_LEAKYRELU || (act & ~(0xff | DL4JB200_ACT_FROM_OUTPUT))) { dl4j_set_error("act_bwd: activation %d not handled by the helper", act); return DL4JB200_NOT_SUPPORTED; }
    if (n == 0) return DL4JB200_OK;
    if ((n & 3) == 0 && aligned16(z) && aligned16(eps) && aligned16(delta))
        k_act_bwd<true><<<ew_grid(ctx, n / 4, 256), 256, 0, ctx->stream>>>(z, eps, delta, n, act, ctx->round_act);
    else
        k_act_bwd<false><<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(z, eps, delta, n, act, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_add_act_fwd(dl4jb200_ctx *ctx, const float *a, const float *b, float *out, size_t n, int act) {
    DL4J_REQUIRE(ctx && a && b && out, "add_act_fwd: NULL argument");
    if (n == 0) return DL4JB200_OK;
    if ((n & 3) == 0 && aligned16(a) && aligned16(b) && aligned16(out))
        k_add_act<true><<<ew_grid(ctx, n / 4, 256), 256, 0, ctx->stream>>>(a, b, out, n, act, ctx->round_act);
    else
        k_add_act<false><<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(a, b, out, n, act, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_dropout_fwd(dl4jb200_ctx *ctx, const float *x, const float *mask, float *y, size_t n, double retain_prob) {
    DL4J_REQUIRE(ctx && x && mask && y, "dropout_fwd: NULL argument");
    DL4J_REQUIRE(retain_prob > 0.0 && retain_prob <= 1.0, "dropout_fwd: retain probability must be in (0, 1]");
    if (n == 0) return DL4JB200_OK;
    const float inv_p = (float)(1.0 / retain_prob);
    if ((n & 3) == 0 && aligned16(x) && aligned16(mask) && aligned16(y))
        k_dropout<true><<<ew_grid(ctx, n / 4, 256), 256, 0, ctx->stream>>>(x, mask, y, n, inv_p, ctx->round_act);
    else
        k_dropout<false><<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(x, mask, y, n, inv_p, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_eps_sum_act_bwd(dl4jb200_ctx *ctx, const float *a, const float *b, const float *act_out, float *delta, size_t n,
                             int act) {
    DL4J_REQUIRE(ctx && a && b && act_out && delta, "eps_sum_act_bwd: NULL argument");
    if (act < 0 || act > DL4JB200_ACT_LEAKYRELU) { dl4j_set_error("eps_sum_act_bwd: activation %d not handled by the helper", act); return DL4JB200_NOT_SUPPORTED; }
    if (n == 0) return DL4JB200_OK;
    if ((n & 3) == 0 && aligned16(a) && aligned16(b) && aligned16(act_out) && aligned16(delta))
        k_sum_act_bwd<true><<<ew_grid(ctx, n / 4, 256), 256, 0, ctx->stream>>>(a, b, act_out, delta, n, act, ctx->round_act);
    else
        k_sum_act_bwd<false><<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(a, b, act_out, delta, n, act, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_accumulate(dl4jb200_ctx *ctx, float *dst, const float *src, size_t n) {
    return dl4jb200_add_act_fwd(ctx, dst, src, dst, n, DL4JB200_ACT_IDENTITY);
}

int dl4jb200_softmax_nll(dl4jb200_ctx *ctx, const float *z, int z_ld, const float *labels, float *delta, int delta_ld,
                         float *loss_sum_out, int B, int K) {
    DL4J_REQUIRE(ctx && z && labels && delta && loss_sum_out, "softmax_nll: NULL argument");
    DL4J_REQUIRE(B > 0 && K > 0 && z_ld >= K && delta_ld >= K, "softmax_nll: bad shape");
    float *rows = (float *)dl4j_workspace(ctx, 0, (size_t)B * sizeof(float));
    if (!rows) return DL4JB200_ERR_CUDA;
    k_softmax_nll<<<dl4j_ceil_div(B, 8), 256, 0, ctx->stream>>>(z, z_ld, labels, delta, delta_ld, rows, B, K, ctx->round_act);
    DL4J_CHECK_LAUNCH();
    k_sum_rows<<<1, 256, 0, ctx->stream>>>(rows, loss_sum_out, B);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

__global__ void k_score_combine(const float *__restrict__ loss_sum, const float *__restrict__ reg, float *__restrict__ score, float b) {
    score[0] = (loss_sum[0] + (reg ? reg[0] : 0.f)) / b;
}

int dl4jb200_score_combine(dl4jb200_ctx *ctx, const float *loss_sum, const float *reg, float *score, int minibatch) {
    DL4J_REQUIRE(ctx && loss_sum && score && minibatch > 0, "score_combine: bad argument");
    k_score_combine<<<1, 1, 0, ctx->stream>>>(loss_sum, reg, score, (float)minibatch);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_l1l2_score(dl4jb200_ctx *ctx, const float *w, size_t n, double l1, double l2, float *score_acc) {
    DL4J_REQUIRE(ctx && w && score_acc, "l1l2_score: NULL argument");
    if (n == 0 || (l1 == 0.0 && l2 == 0.0)) return DL4JB200_OK;
    int nblk = ew_grid(ctx, n, 256);
    if (nblk > 1024) nblk = 1024;
    double *partial = (double *)dl4j_workspace(ctx, 0, (size_t)nblk * 2 * sizeof(double));
    if (!partial) return DL4JB200_ERR_CUDA;
    k_l1l2_partial<<<nblk, 256, 0, ctx->stream>>>(w, n, partial);
    DL4J_CHECK_LAUNCH();
    k_l1l2_final<<<1, 128, 0, ctx->stream>>>(partial, nblk, l1, l2, score_acc);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_copy2d(dl4jb200_ctx *ctx, const float *src, int src_ld, float *dst, int dst_ld, int rows, int cols) {
    DL4J_REQUIRE(ctx && src && dst, "copy2d: NULL argument");
    DL4J_REQUIRE(rows > 0 && cols > 0 && src_ld >= cols && dst_ld >= cols, "copy2d: bad shape");
    k_copy2d<<<ew_grid(ctx, (size_t)rows * dst_ld, 256), 256, 0, ctx->stream>>>(src, src_ld, dst, dst_ld, rows, cols);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_updater_apply(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n, int kind, double lr,
                           double h1, double h2, double l1, double l2, double batch_div) {
    DL4J_REQUIRE(ctx && w && g, "updater_apply: NULL argument");
    DL4J_REQUIRE(kind >= 0 && kind <= DL4JB200_UPD_NOOP, "updater_apply: unknown updater %d", kind);
    DL4J_REQUIRE(state || kind == DL4JB200_UPD_SGD || kind == DL4JB200_UPD_NOOP, "updater_apply: state is NULL");
    DL4J_REQUIRE(batch_div > 0, "updater_apply: batch_div must be positive");
    if (n == 0) return DL4JB200_OK;
    k_updater<<<ew_grid(ctx, n, 256), 256, 0, ctx->stream>>>(w, g, state, n, kind, (float)lr, (float)h1, (float)h2,
                                                           (float)l1, (float)l2, (float)(1.0 / batch_div));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

}  // extern "C"
