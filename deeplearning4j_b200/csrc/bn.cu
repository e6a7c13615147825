// Batch normalisation forward (train / inference) and backward (BatchNormalization.java:93-219, 239-395).
//
// HBM-bound. Each direction is: (1) one streaming pass producing per-block partial sums for every channel
// (128-bit loads, per-thread accumulation, shared-memory fold; partials go to context scratch), (2) a tiny
// finalize kernel that folds the partials in double, applies the reference's exact formulas (biased
// variance, eps stored into the running variance, ...) and emits per-channel coefficients, (3) one
// streaming pass applying them. Forward sums are taken about a per-channel pivot (the channel's first
// element) so that var = E[(x-K)^2] - E[x-K]^2 does not cancel catastrophically in fp32.
// Algorithmic traffic: forward 8 B/elem + one extra read (12 moved), backward 12 B/elem + extra read of x
// and eps (20 moved); the extra reads are what a conv-epilogue fusion removes later.
#include <float.h>

#include "common.cuh"

#define BN_THREADS 256

// ------------------------------------------------------------------ pass 1: partial sums
// MODE 0 (forward):  s1 = sum (x-K),  s2 = sum (x-K)^2          K = pivot[c]
// MODE 1 (backward): s1 = sum e,      s2 = sum e*(x-mean[c])    e = eps, zeroed where the fused ReLU was inactive; the
//                    mask is RECOMPUTED as (x-mean)*gamma/std + beta > 0 (same expression as the forward apply), which
//                    saves reading the activation tensor back from HBM
struct BnMask {  // fused-ReLU mask inputs for the backward kernels (relu == 0: no mask)
    const float *gamma, *beta, *std;
    int relu;
};
__device__ __forceinline__ void bn_mask_coef(const BnMask &mk, int c, int C, float &a, float &b) {
    if (c < C) {
        a = (float)((mk.gamma ? (double)mk.gamma[c] : 1.0) / (double)mk.std[c]);  // identical to k_bn_finalize_fwd's coef
        b = mk.beta ? mk.beta[c] : 0.f;
    } else { a = 0.f; b = 0.f; }
}

// NHWC: grid (G, S). Block (g, s) reduces rows [s*rows_per, (s+1)*rows_per) of the 32-channel group g (8 channel
// quads = one 128-byte line per pixel). 256 threads = 8 quads x 32 row lanes, 4 independent 128-bit loads in flight per
// thread. A channel therefore has only S <= ~600/G partial rows for the finalize kernel to fold (the fold is a
// latency-bound tail on the critical path), instead of one row per block of the whole launch.
static inline int bn_nhwc_groups(int cs) { return (cs / 4 + 7) / 8; }
static inline int bn_nhwc_splits(int sm_count, int cs, long long M) {
    long long G = bn_nhwc_groups(cs), cap = ((long long)sm_count * 4 + G - 1) / G, by_rows = M / 128;
    long long S = cap < by_rows ? cap : by_rows;
    return (int)(S < 1 ? 1 : S);
}

template <int MODE>
__global__ void __launch_bounds__(BN_THREADS) k_bn_partial_nhwc(const float *__restrict__ x, const float *__restrict__ e,
                                                                BnMask mk,
                                                                const float *__restrict__ center, int center_stride,
                                                                float *__restrict__ partial, int M, int cs) {
    __shared__ float4 red[2][BN_THREADS];
    pdl_trigger();
    pdl_wait();
    const int Q = cs >> 2;
    constexpr int R = BN_THREADS / 8;  // row lanes
    const int q = blockIdx.x * 8 + (threadIdx.x & 7), rl = threadIdx.x >> 3;
    const int rows_per_blk = (M + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows_per_blk, r1 = min(M, r0 + rows_per_blk);
    float4 a1 = make_float4(0, 0, 0, 0), a2 = a1;
    if (q < Q) {
        // forward pivot = first row of the tensor; backward centre = saved mean (stride 1 array)
        float4 k;
        if (MODE == 0) k = *reinterpret_cast<const float4 *>(center + (size_t)q * 4);
        else {
            int c = q * 4;
            k.x = c + 0 < center_stride ? center[c + 0] : 0.f; k.y = c + 1 < center_stride ? center[c + 1] : 0.f;
            k.z = c + 2 < center_stride ? center[c + 2] : 0.f; k.w = c + 3 < center_stride ? center[c + 3] : 0.f;
        }
        float4 ma = make_float4(0, 0, 0, 0), mb = ma;
        if (MODE == 1 && mk.relu) {
            int c = q * 4;
            bn_mask_coef(mk, c, center_stride, ma.x, mb.x); bn_mask_coef(mk, c + 1, center_stride, ma.y, mb.y);
            bn_mask_coef(mk, c + 2, center_stride, ma.z, mb.z); bn_mask_coef(mk, c + 3, center_stride, ma.w, mb.w);
        }
        constexpr int U = 4;  // independent loads in flight per thread (latency hiding)
        for (int rb = r0 + rl; rb < r1; rb += U * R) {
            float4 v[U], g[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                int r = rb + u * R;
                size_t o = (size_t)r * cs + (size_t)q * 4;
                if (r < r1) {
                    v[u] = *reinterpret_cast<const float4 *>(x + o);
                    if (MODE == 1) g[u] = *reinterpret_cast<const float4 *>(e + o);
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (rb + u * R >= r1) break;
                float4 w = v[u];
                w.x -= k.x; w.y -= k.y; w.z -= k.z; w.w -= k.w;
                if (MODE == 0) {
                    a1.x += w.x; a1.y += w.y; a1.z += w.z; a1.w += w.w;
                    a2.x += w.x * w.x; a2.y += w.y * w.y; a2.z += w.z * w.z; a2.w += w.w * w.w;
                } else {
                    float4 gg = g[u];
                    if (mk.relu) {
                        gg.x = w.x * ma.x + mb.x > 0.f ? gg.x : 0.f; gg.y = w.y * ma.y + mb.y > 0.f ? gg.y : 0.f;
                        gg.z = w.z * ma.z + mb.z > 0.f ? gg.z : 0.f; gg.w = w.w * ma.w + mb.w > 0.f ? gg.w : 0.f;
                    }
                    a1.x += gg.x; a1.y += gg.y; a1.z += gg.z; a1.w += gg.w;
                    a2.x += gg.x * w.x; a2.y += gg.y * w.y; a2.z += gg.z * w.z; a2.w += gg.w * w.w;
                }
            }
        }
    }
    red[0][threadIdx.x] = a1;
    red[1][threadIdx.x] = a2;
    __syncthreads();
    // tree over the 32 row lanes (offsets are multiples of 8, so a quad lane only ever meets itself); fixed shape
    for (int o = BN_THREADS / 2; o >= 8; o >>= 1) {
        if ((int)threadIdx.x < o) {
            float4 b1 = red[0][threadIdx.x + o], b2 = red[1][threadIdx.x + o];
            float4 c1 = red[0][threadIdx.x], c2 = red[1][threadIdx.x];
            c1.x += b1.x; c1.y += b1.y; c1.z += b1.z; c1.w += b1.w;
            c2.x += b2.x; c2.y += b2.y; c2.z += b2.z; c2.w += b2.w;
            red[0][threadIdx.x] = c1; red[1][threadIdx.x] = c2;
        }
        __syncthreads();
    }
    if (threadIdx.x < 8 && q < Q) {
        float *p = partial + (size_t)blockIdx.y * 2 * cs;
        *reinterpret_cast<float4 *>(p + (size_t)q * 4) = red[0][threadIdx.x];
        *reinterpret_cast<float4 *>(p + cs + (size_t)q * 4) = red[1][threadIdx.x];
    }
}

// NCHW: grid (C, S); block (c, s) reduces elements [s*chunk, (s+1)*chunk) of the B*HW values of channel c.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(BN_THREADS) k_bn_partial_nchw(const float *__restrict__ x, const float *__restrict__ e,
                                                                BnMask mk,
                                                                const float *__restrict__ center, float *__restrict__ partial,
                                                                int B, int C, int HW, long long chunk) {
    __shared__ float red[2][BN_THREADS];
    const int c = blockIdx.x;
    const long long total = (long long)B * HW;
    const long long j0 = (long long)blockIdx.y * chunk, j1 = min(total, j0 + chunk);
    const float k = MODE == 0 ? x[(size_t)c * HW] : center[c];
    float ma = 0.f, mb = 0.f;
    if (MODE == 1 && mk.relu) bn_mask_coef(mk, c, C, ma, mb);
    float a1 = 0.f, a2 = 0.f;
    constexpr int V = VEC ? 4 : 1;
    for (long long j = j0 + (long long)threadIdx.x * V; j < j1; j += (long long)BN_THREADS * V) {
        long long b = j / HW;
        size_t o = ((size_t)b * C + c) * HW + (size_t)(j - b * HW);
        float xv[V], ev[V];
        if (VEC) {
            float4 t = *reinterpret_cast<const float4 *>(x + o);
            xv[0] = t.x; xv[1 % V] = t.y; xv[2 % V] = t.z; xv[3 % V] = t.w;
            if (MODE == 1) {
                t = *reinterpret_cast<const float4 *>(e + o);
                ev[0] = t.x; ev[1 % V] = t.y; ev[2 % V] = t.z; ev[3 % V] = t.w;
            }
        } else {
            xv[0] = x[o];
            if (MODE == 1) ev[0] = e[o];
        }
#pragma unroll
        for (int i = 0; i < V; i++) {
            float d = xv[i] - k;
            if (MODE == 0) { a1 += d; a2 += d * d; }
            else {
                float g = (mk.relu && !(d * ma + mb > 0.f)) ? 0.f : ev[i];
                a1 += g; a2 += g * d;
            }
        }
    }
    red[0][threadIdx.x] = a1;
    red[1][threadIdx.x] = a2;
    __syncthreads();
    for (int o = BN_THREADS / 2; o; o >>= 1) {
        if (threadIdx.x < o) { red[0][threadIdx.x] += red[0][threadIdx.x + o]; red[1][threadIdx.x] += red[1][threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float *p = partial + (size_t)blockIdx.y * 2 * C;
        p[c] = red[0][0];
        p[C + c] = red[1][0];
    }
}

// ------------------------------------------------------------------ pass 2: finalize (one thread per channel)
// coef layout: [4][cpad] = centre, a, b, c   with  y = (x-centre)*a + c            (forward; b unused)
//                                              dx = a*e + b*(x-centre) + c          (backward)
// Folds partial[nblk][2][cpad] for channel c over the BN_FY threads (threadIdx.y) that share it; result valid for y == 0.
// A block is 32 channels x BN_FY = 256 threads, not 1024: the finalize kernels run next to the side stream's filter-gradient
// CTAs, and a 1024-thread block does not fit into the registers those leave free on an SM (it waited ~10 us for one to
// drain, on the critical path, 102 times per ResNet-50 step); 256 threads are resident at once.
#define BN_FY 8
__device__ __forceinline__ void bn_fold_partials(const float *__restrict__ partial, int nblk, int cpad, int c, bool active,
                                                 double &s1, double &s2) {
    __shared__ double sh1[BN_FY][33], sh2[BN_FY][33];
    double a = 0.0, b = 0.0;
    if (active) {
        // the fold is latency-bound (a few hundred values per channel): issue every load of a 16-row batch before the
        // first add. Row order per thread is fixed (y, y+BN_FY, y+2*BN_FY, ...), so the result is deterministic.
        constexpr int U = 8;
        for (int i = threadIdx.y; i < nblk; i += BN_FY * U) {
            float p[U], q[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int r = i + BN_FY * u;
                const bool ok = r < nblk;
                p[u] = ok ? partial[(size_t)r * 2 * cpad + c] : 0.f;
                q[u] = ok ? partial[(size_t)r * 2 * cpad + cpad + c] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < U; u++) { a += (double)p[u]; b += (double)q[u]; }
        }
    }
    sh1[threadIdx.y][threadIdx.x] = a;
    sh2[threadIdx.y][threadIdx.x] = b;
    __syncthreads();
    // tree fold over the BN_FY y-slots (fixed shape => deterministic)
    for (int h = BN_FY / 2; h >= 1; h >>= 1) {
        if ((int)threadIdx.y < h) {
            sh1[threadIdx.y][threadIdx.x] += sh1[threadIdx.y + h][threadIdx.x];
            sh2[threadIdx.y][threadIdx.x] += sh2[threadIdx.y + h][threadIdx.x];
        }
        __syncthreads();
    }
    s1 = sh1[0][threadIdx.x];
    s2 = sh2[0][threadIdx.x];
}

__global__ void k_bn_finalize_fwd(const float *__restrict__ partial, int nblk, const float *__restrict__ pivot,
                                  int pivot_stride, const float *__restrict__ gamma, const float *__restrict__ beta,
                                  float *gmean, float *gvar, float *save_mean, float *save_std, float *coef, int C,
                                  int cpad, double N, double decay, double eps, int training, int minibatch) {
    int c = blockIdx.x * 32 + threadIdx.x;
    double s1, s2;
    pdl_trigger();
    pdl_wait();
    bn_fold_partials(partial, training ? nblk : 0, cpad, c, c < C, s1, s2);
    if (threadIdx.y != 0 || c >= cpad) return;
    if (c >= C) {  // NHWC padding channel: stays zero
        coef[c] = 0.f; coef[cpad + c] = 0.f; coef[2 * cpad + c] = 0.f; coef[3 * cpad + c] = 0.f;
        return;
    }
    double mean, std;
    if (training) {
        double k = (double)pivot[(size_t)c * pivot_stride];
        double m1 = s1 / N;
        mean = k + m1;                              // BatchNormalization.java:264
        double var = s2 / N - m1 * m1;              // biased, :265
        if (var < 0.0) var = 0.0;
        double v = var + eps;                       // :273 (in place: eps becomes part of the stored variance)
        std = sqrt(v);                              // :279
        if (minibatch) {                            // :377-384
            gmean[c] = (float)(decay * (double)gmean[c] + (1.0 - decay) * mean);
            gvar[c] = (float)(decay * (double)gvar[c] + (1.0 - decay) * v);
        } else {                                    // :386-390
            gmean[c] = (float)mean;
            gvar[c] = (float)v;
        }
    } else {                                        // :274-279, no eps added in test mode
        mean = (double)gmean[c];
        std = sqrt((double)gvar[c]);
    }
    if (save_mean) save_mean[c] = (float)mean;
    if (save_std) save_std[c] = (float)std;
    double g = gamma ? (double)gamma[c] : 1.0, bt = beta ? (double)beta[c] : 0.0;
    coef[c] = (float)mean;
    coef[cpad + c] = (float)(g / std);
    coef[2 * cpad + c] = 0.f;
    coef[3 * cpad + c] = (float)bt;
}

__global__ void k_bn_finalize_bwd(const float *__restrict__ partial, int nblk, const float *__restrict__ gamma,
                                  const float *__restrict__ save_mean, const float *__restrict__ save_std,
                                  const float *__restrict__ beta, float *dgamma, float *dbeta, float *coef, int C,
                                  int cpad, double N) {
    int c = blockIdx.x * 32 + threadIdx.x;
    double s1, s2;
    pdl_trigger();
    pdl_wait();
    bn_fold_partials(partial, nblk, cpad, c, c < C, s1, s2);
    if (threadIdx.y != 0 || c >= cpad) return;
    coef[4 * cpad + c] = (c < C && beta) ? beta[c] : 0.f;  // fused-ReLU mask: (x-mean)*a + beta > 0
    if (c >= C) {
        coef[c] = 0.f; coef[cpad + c] = 0.f; coef[2 * cpad + c] = 0.f; coef[3 * cpad + c] = 0.f;
        return;
    }
    double std = (double)save_std[c], g = gamma ? (double)gamma[c] : 1.0;
    // BatchNormalization.java:172-196 with dxhat = e*gamma:
    //   dVar = sum(dxhat*xMu) * -0.5 * std^-3 ; dMu = -sum(dxhat)/std (+ dVar*(-2/N)*sum xMu, sum xMu == 0)
    //   dx = dxhat/std + xMu*dVar*2/N + dMu/N
    if (dbeta) dbeta[c] = (float)s1;
    if (dgamma) dgamma[c] = (float)(s2 / std);
    coef[c] = save_mean[c];
    coef[cpad + c] = (float)(g / std);
    coef[2 * cpad + c] = (float)(-g * s2 / (N * std * std * std));
    coef[3 * cpad + c] = (float)(-g * s1 / (N * std));
}

// ------------------------------------------------------------------ pass 3: apply
// MODE 0: y = act((x-centre)*a + c).  MODE 1: dx = a*e' + b*(x-centre) + c, e' = eps masked by act_out.
// FIXQ: the launch's stride (gridDim.x * 256) is a multiple of the channel-quad count, so a thread meets the same four
// channels in every iteration and keeps their coefficients in registers (no per-element modulo, no coefficient re-loads).
template <int MODE, bool FIXQ = false>
__global__ void __launch_bounds__(256) k_bn_apply_nhwc(const float *__restrict__ x, const float *__restrict__ e,
                                                       const float *__restrict__ mask_beta, float *__restrict__ out,
                                                       const float *__restrict__ coef, size_t total4, int cs, int act) {
    const int rnd = act >> 8;  // host packs ctx->round_act above the activation code
    act &= 0xff;
    const int Q = cs >> 2;
    pdl_trigger();
    pdl_wait();
    size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t first = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float4 k, a, b, c, m;
    auto load_coef = [&](int q) {
        k = __ldg(reinterpret_cast<const float4 *>(coef) + q);
        a = __ldg(reinterpret_cast<const float4 *>(coef + cs) + q);
        c = __ldg(reinterpret_cast<const float4 *>(coef + 3 * (size_t)cs) + q);
        if (MODE == 1) {
            b = __ldg(reinterpret_cast<const float4 *>(coef + 2 * (size_t)cs) + q);
            if (mask_beta) m = __ldg(reinterpret_cast<const float4 *>(mask_beta) + q);
        }
    };
    if (FIXQ && first < total4) load_coef((int)((total4 - 1 - first) % Q));
    // The tensor is walked from its END to its start. The producer of x (a convolution epilogue, or the partial-sum pass
    // that has just read x and eps) finished with the high addresses, so those lines are the ones still in the 126 MB L2;
    // and the consumer of `out` (the next convolution, whose tiles run from the first pixel up) in turn finds the lines
    // this kernel wrote last. For tensors up to the L2 size the second read never reaches DRAM.
    for (size_t ii = first; ii < total4; ii += stride) {
        const size_t i = total4 - 1 - ii;
        if (!FIXQ) load_coef((int)(i % Q));
        float4 v = reinterpret_cast<const float4 *>(x)[i];
        v.x -= k.x; v.y -= k.y; v.z -= k.z; v.w -= k.w;
        float4 r;
        if (MODE == 0) {
            float4 res = make_float4(0, 0, 0, 0);
            if (e) res = reinterpret_cast<const float4 *>(e)[i];  // fused ElementWiseVertex(Add): y = act(bn(x) + residual)
            r.x = dl4j_act(v.x * a.x + c.x + res.x, act); r.y = dl4j_act(v.y * a.y + c.y + res.y, act);
            r.z = dl4j_act(v.z * a.z + c.z + res.z, act); r.w = dl4j_act(v.w * a.w + c.w + res.w, act);
        } else {
            float4 g = reinterpret_cast<const float4 *>(e)[i];
            if (mask_beta) {  // fused ReLU of the forward: recompute its mask instead of re-reading the activation
                g.x = v.x * a.x + m.x > 0.f ? g.x : 0.f; g.y = v.y * a.y + m.y > 0.f ? g.y : 0.f;
                g.z = v.z * a.z + m.z > 0.f ? g.z : 0.f; g.w = v.w * a.w + m.w > 0.f ? g.w : 0.f;
            }
            r.x = a.x * g.x + b.x * v.x + c.x; r.y = a.y * g.y + b.y * v.y + c.y;
            r.z = a.z * g.z + b.z * v.z + c.z; r.w = a.w * g.w + b.w * v.w + c.w;
        }
        reinterpret_cast<float4 *>(out)[i] = dl4j_rnd4(r, rnd);
    }
}

// Forward apply + the fused activation + the max-pooling layer that is the batch norm's only consumer, in one pass:
// y_pool[b,oh,ow,c] = max over the window of rnd(act((x - mean) * a + beta)), argmax = first maximum in (kh, kw) row-major
// order exactly as SubsamplingLayer's IsMax picks it (SubsamplingLayer.java:205-218). The normalised tensor itself is
// never written or read back (ResNet-50's stem: 822 MB each way per step). Every tap value is rounded exactly as the
// stand-alone apply pass would have stored it, so values, ties and indices are those of the two-kernel path.
struct BnPoolGeom {
    int B, H, W, Ho, Wo, kH, kW, sH, sW, pT, pL, xcs, ycs;
    FastDiv d_q, d_w, d_h;
};

template <int KS>
__global__ void __launch_bounds__(256) k_bn_act_maxpool_nhwc(const float *__restrict__ x, float *__restrict__ y, int32_t *__restrict__ argmax,
                                                             const float *__restrict__ coef, BnPoolGeom g, size_t total, int act) {
    const int rnd = act >> 8;
    act &= 0xff;
    const int kH = KS ? KS : g.kH, kW = KS ? KS : g.kW;
    pdl_trigger();
    pdl_wait();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        uint32_t q, ow, oh, b, r;
        g.d_q.divmod((uint32_t)idx, r, q);  // total < 2^32 (checked on the host)
        g.d_w.divmod(r, r, ow);
        g.d_h.divmod(r, b, oh);
        const float4 k = __ldg(reinterpret_cast<const float4 *>(coef) + q);
        const float4 a = __ldg(reinterpret_cast<const float4 *>(coef + g.xcs) + q);
        const float4 c = __ldg(reinterpret_cast<const float4 *>(coef + 3 * (size_t)g.xcs) + q);
        float4 best = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
        int bi[4] = {0, 0, 0, 0};
        auto tap = [&](float4 v, int t) {
            v.x = dl4j_act((v.x - k.x) * a.x + c.x, act); v.y = dl4j_act((v.y - k.y) * a.y + c.y, act);
            v.z = dl4j_act((v.z - k.z) * a.z + c.z, act); v.w = dl4j_act((v.w - k.w) * a.w + c.w, act);
            v = dl4j_rnd4(v, rnd);
            if (v.x > best.x) { best.x = v.x; bi[0] = t; }
            if (v.y > best.y) { best.y = v.y; bi[1] = t; }
            if (v.z > best.z) { best.z = v.z; bi[2] = t; }
            if (v.w > best.w) { best.w = v.w; bi[3] = t; }
        };
        const int ih0 = (int)oh * g.sH - g.pT, iw0 = (int)ow * g.sW - g.pL;
        if (KS && ih0 >= 0 && iw0 >= 0 && ih0 + KS <= g.H && iw0 + KS <= g.W) {
            // window entirely inside the image (every window of an unpadded layer): all KS*KS loads are issued before the
            // first use -- the bounds-checked loop below serialises them behind their branches and runs latency-bound
            float4 v[KS ? KS * KS : 1];
            const float *xp = x + (((size_t)b * g.H + ih0) * g.W + iw0) * g.xcs + (size_t)q * 4;
#pragma unroll
            for (int kh = 0; kh < KS; kh++)
#pragma unroll
                for (int kw = 0; kw < KS; kw++) v[kh * KS + kw] = *reinterpret_cast<const float4 *>(xp + ((size_t)kh * g.W + kw) * g.xcs);
#pragma unroll
            for (int t = 0; t < KS * KS; t++) tap(v[t], t);
        } else {
#pragma unroll KS ? KS : 1
            for (int kh = 0; kh < kH; kh++) {
                const int ih = ih0 + kh;
                if (ih < 0 || ih >= g.H) continue;
#pragma unroll KS ? KS : 1
                for (int kw = 0; kw < kW; kw++) {
                    const int iw = iw0 + kw;
                    if (iw < 0 || iw >= g.W) continue;
                    tap(*reinterpret_cast<const float4 *>(x + (((size_t)b * g.H + ih) * g.W + iw) * g.xcs + (size_t)q * 4), kh * kW + kw);
                }
            }
        }
        const size_t o = (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.ycs + (size_t)q * 4;
        *reinterpret_cast<float4 *>(y + o) = best;
        if (argmax) *reinterpret_cast<int4 *>(argmax + o) = make_int4(bi[0], bi[1], bi[2], bi[3]);
    }
}

// ------------------------------------------------------------------ MAX-pooling backward that also takes the BN backward sums
// The epsilon of a BatchNormalization whose only consumer is a stride-2 MAX pooling layer (ResNet-50's stem) is produced by
// that layer's backprop. This is k_maxpool_bwd_nhwc_patch2 (pool.cu: one thread per 2x2 patch of dx x 4 channels, the <= 2x2
// candidate windows fetched up front, same per-cell summation order and rounding) which, while the patch is in registers,
// also reads the batch norm's input x, recomputes the fused-ReLU gate exactly as k_bn_partial_nhwc does and accumulates
// sum e', sum e' * (x - mean) for its channel quad -- the launch's stride is a multiple of the quad count, so a thread keeps
// its channels. One partial row per block in k_bn_finalize_bwd's layout: the separate partial-sum pass over (x, eps)
// (1.6 GB for the stem) disappears.
struct PoolBstGeom {
    int B, H, W, Ho, Wo, kH, kW, pT, pL, cs, Hp, Wp, rnd;
    FastDiv d_q, d_wp, d_hp;
};

__global__ void __launch_bounds__(256) k_maxpool_bwd_bnsums_nhwc(const float *__restrict__ eps, const int32_t *__restrict__ argmax,
                                                                 float *__restrict__ dx, const float *__restrict__ x, BnMask mk,
                                                                 const float *__restrict__ mean, float *__restrict__ partial, PoolBstGeom g,
                                                                 size_t total) {
    __shared__ float4 red[2][256];
    pdl_trigger();
    pdl_wait();
    const int Q = g.cs >> 2, q = (int)(threadIdx.x % (unsigned)Q), c = q * 4;  // (gridDim.x * 256) % Q == 0: fixed per thread
    const float4 k = *reinterpret_cast<const float4 *>(mean + c);
    float4 ma = make_float4(0, 0, 0, 0), mb = ma;
    if (mk.relu) {
        bn_mask_coef(mk, c, g.cs, ma.x, mb.x); bn_mask_coef(mk, c + 1, g.cs, ma.y, mb.y);
        bn_mask_coef(mk, c + 2, g.cs, ma.z, mb.z); bn_mask_coef(mk, c + 3, g.cs, ma.w, mb.w);
    }
    float4 a1 = make_float4(0, 0, 0, 0), a2 = a1;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        uint32_t t, qq, wp, hp, b;
        g.d_q.divmod((uint32_t)idx, t, qq);  // qq == q (see above)
        (void)qq;
        g.d_wp.divmod(t, t, wp);
        g.d_hp.divmod(t, b, hp);
        const int h0 = (int)hp * 2, w0 = (int)wp * 2;
        const int th0 = h0 + g.pT, tw0 = w0 + g.pL;
        const int oh_hi = min((th0 + 1) >> 1, g.Ho - 1), ow_hi = min((tw0 + 1) >> 1, g.Wo - 1);
        const int lh = th0 - g.kH + 1, lw = tw0 - g.kW + 1;
        const int oh_lo = lh <= 0 ? 0 : (lh + 1) >> 1, ow_lo = lw <= 0 ? 0 : (lw + 1) >> 1;
        float4 v[2][2];
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int d = 0; d < 2; d++) {
                v[a][d] = make_float4(0, 0, 0, 0);
                if (h0 + a < g.H && w0 + d < g.W) v[a][d] = *reinterpret_cast<const float4 *>(x + (((size_t)b * g.H + h0 + a) * g.W + w0 + d) * g.cs + c);
            }
        int4 am[2][2];
        float4 e[2][2];
#pragma unroll
        for (int i = 0; i < 2; i++)
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int oh = oh_hi - i, ow = ow_hi - j;
                am[i][j] = make_int4(-1, -1, -1, -1);
                e[i][j] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (oh >= oh_lo && ow >= ow_lo) {
                    const size_t o = (((size_t)b * g.Ho + oh) * g.Wo + ow) * g.cs + c;
                    am[i][j] = *reinterpret_cast<const int4 *>(argmax + o);
                    e[i][j] = *reinterpret_cast<const float4 *>(eps + o);
                }
            }
        float4 acc[2][2];
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int d = 0; d < 2; d++) acc[a][d] = make_float4(0.f, 0.f, 0.f, 0.f);
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
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int d = 0; d < 2; d++) {
                if (h0 + a >= g.H || w0 + d >= g.W) continue;
                const float4 r = dl4j_rnd4(acc[a][d], g.rnd);
                *reinterpret_cast<float4 *>(dx + (((size_t)b * g.H + h0 + a) * g.W + w0 + d) * g.cs + c) = r;
                float4 w = v[a][d], gg = r;
                w.x -= k.x; w.y -= k.y; w.z -= k.z; w.w -= k.w;
                if (mk.relu) {
                    gg.x = w.x * ma.x + mb.x > 0.f ? gg.x : 0.f; gg.y = w.y * ma.y + mb.y > 0.f ? gg.y : 0.f;
                    gg.z = w.z * ma.z + mb.z > 0.f ? gg.z : 0.f; gg.w = w.w * ma.w + mb.w > 0.f ? gg.w : 0.f;
                }
                a1.x += gg.x; a1.y += gg.y; a1.z += gg.z; a1.w += gg.w;
                a2.x += gg.x * w.x; a2.y += gg.y * w.y; a2.z += gg.z * w.z; a2.w += gg.w * w.w;
            }
    }
    red[0][threadIdx.x] = a1;
    red[1][threadIdx.x] = a2;
    __syncthreads();
    if ((int)threadIdx.x < Q) {  // fold the 256 / Q threads that share this quad, in thread order (deterministic)
        float4 s1 = red[0][threadIdx.x], s2 = red[1][threadIdx.x];
        for (int t = threadIdx.x + Q; t < 256; t += Q) {
            const float4 b1 = red[0][t], b2 = red[1][t];
            s1.x += b1.x; s1.y += b1.y; s1.z += b1.z; s1.w += b1.w;
            s2.x += b2.x; s2.y += b2.y; s2.z += b2.z; s2.w += b2.w;
        }
        float *p = partial + (size_t)blockIdx.x * 2 * g.cs;
        *reinterpret_cast<float4 *>(p + c) = s1;
        *reinterpret_cast<float4 *>(p + g.cs + c) = s2;
    }
}

// launches k_bn_apply_nhwc, with per-thread fixed channels whenever a grid close to the usual one makes the stride a multiple
// of the channel-quad count
template <int MODE>
static cudaError_t bn_launch_apply_nhwc(dl4jb200_ctx *ctx, int grid, const float *x, const float *e, const float *mask_beta, float *out,
                                        const float *coef, size_t total4, int cs, int act) {
    const int Q = cs >> 2;
    int g = grid;
    bool fix = false;
    for (int t = 0; t < 4 && !fix; t++) {  // grid, grid-1, ... : 256 * g divisible by Q (Q is a power of two in every zoo model)
        if (g - t >= 1 && ((long long)(g - t) * 256) % Q == 0) { g -= t; fix = true; }
    }
    if (fix) return dl4j_launch_pdl(k_bn_apply_nhwc<MODE, true>, dim3(g), dim3(256), 0, ctx->stream, x, e, mask_beta, out, coef, total4, cs, act);
    return dl4j_launch_pdl(k_bn_apply_nhwc<MODE, false>, dim3(grid), dim3(256), 0, ctx->stream, x, e, mask_beta, out, coef, total4, cs, act);
}

template <int MODE, bool VEC>
__global__ void __launch_bounds__(256) k_bn_apply_nchw(const float *__restrict__ x, const float *__restrict__ e,
                                                       const float *__restrict__ mask_beta, float *__restrict__ out,
                                                       const float *__restrict__ coef, size_t total, int C, int HW,
                                                       int act) {
    constexpr int V = VEC ? 4 : 1;
    const int rnd = act >> 8;
    act &= 0xff;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i * V < total; i += stride) {
        size_t o = i * V;
        int ch = (int)((o / HW) % C);
        float k = __ldg(coef + ch), a = __ldg(coef + C + ch), b = __ldg(coef + 2 * C + ch), c = __ldg(coef + 3 * C + ch);
        const float mb = (MODE == 1 && mask_beta) ? __ldg(mask_beta + ch) : 0.f;
        float xv[V], ev[V], r[V];
        if (VEC) {
            float4 t = *reinterpret_cast<const float4 *>(x + o);
            xv[0] = t.x; xv[1 % V] = t.y; xv[2 % V] = t.z; xv[3 % V] = t.w;
            if (MODE == 1) {
                t = *reinterpret_cast<const float4 *>(e + o);
                ev[0] = t.x; ev[1 % V] = t.y; ev[2 % V] = t.z; ev[3 % V] = t.w;
            }
        } else {
            xv[0] = x[o];
            if (MODE == 1) ev[0] = e[o];
        }
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < V; j++) ev[j] = e ? e[o + j] : 0.f;  // optional residual
        }
#pragma unroll
        for (int j = 0; j < V; j++) {
            float d = xv[j] - k;
            if (MODE == 0) r[j] = dl4j_act(d * a + c + ev[j], act);
            else {
                float g = (mask_beta && !(d * a + mb > 0.f)) ? 0.f : ev[j];
                r[j] = a * g + b * d + c;
            }
            r[j] = dl4j_rnd(r[j], rnd);
        }
        if (VEC) *reinterpret_cast<float4 *>(out + o) = make_float4(r[0], r[1 % V], r[2 % V], r[3 % V]);
        else out[o] = r[0];
    }
}

static inline bool al16(const void *p) { return ((uintptr_t)p & 15) == 0; }

struct BnPlan {
    int nblk;         // partial blocks
    long long chunk;  // NCHW: elements per split
    float *partial, *coef;
    int cpad;
};

static int bn_plan(dl4jb200_ctx *ctx, int B, int C, int H, int W, int layout, int cs, BnPlan &p) {
    long long HW = (long long)H * W, M = (long long)B * HW;
    if (layout == DL4JB200_NHWC) {
        p.cpad = cs;
        p.nblk = bn_nhwc_splits(ctx->sm_count, cs, M);
        p.chunk = 0;
    } else {
        p.cpad = C;
        long long per = M;  // elements per channel
        long long cap = ((long long)ctx->sm_count * 8 + C - 1) / C;  // aim at ~8 blocks per SM overall
        long long s = per / 4096;
        if (s > cap) s = cap;
        if (s < 1) s = 1;
        long long chunk = (per + s - 1) / s;
        chunk = (chunk + 3) / 4 * 4;
        p.nblk = (int)((per + chunk - 1) / chunk);
        p.chunk = chunk;
    }
    size_t bytes = ((size_t)p.nblk * 2 + 5) * p.cpad * sizeof(float);
    float *ws = (float *)dl4j_workspace(ctx, 0, bytes);
    if (!ws) return DL4JB200_ERR_CUDA;
    p.coef = ws;  // first: 16-byte aligned coefficient block (cpad % 4 == 0 for NHWC)
    p.partial = ws + 5 * (size_t)p.cpad;  // coef rows: centre, a, b, c, mask beta
    return DL4JB200_OK;
}

static inline int bn_grid(dl4jb200_ctx *ctx, size_t items) {
    size_t blocks = (items + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    return (int)(blocks > cap ? cap : (blocks ? blocks : 1));
}

extern "C" {

int dl4jb200_bn_fwd(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta,
                    float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C, int H, int W,
                    int layout, int cs, double decay, double eps, int training, int minibatch, int act,
                    const float *residual) {
    DL4J_REQUIRE(ctx && x && y && gmean && gvar, "bn_fwd: NULL argument");
    DL4J_REQUIRE(!residual || layout == DL4JB200_NCHW || al16(residual), "bn_fwd: residual must be 16-byte aligned");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "bn_fwd: bad shape");
    DL4J_REQUIRE((long long)B * H * W < (1LL << 31), "bn_fwd: B*H*W must stay below 2^31");
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "bn_fwd: unknown layout %d", layout);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || (cs >= C && cs % 4 == 0 && al16(x) && al16(y)), "bn_fwd: NHWC needs cs %% 4 == 0, cs >= C and 16-byte aligned tensors");
    DL4J_REQUIRE(C <= 65535 * 8, "bn_fwd: too many channels");
    BnPlan p;
    int rc = bn_plan(ctx, B, C, H, W, layout, cs, p);
    if (rc) return rc;
    const int HW = H * W;
    const long long M = (long long)B * HW;
    const size_t total = (size_t)M * (layout == DL4JB200_NHWC ? cs : C);
    if (training) {
        if (layout == DL4JB200_NHWC) {
            DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_partial_nhwc<0>, dim3(bn_nhwc_groups(cs), p.nblk), dim3(BN_THREADS), 0, ctx->stream, x, nullptr,
                                          BnMask{nullptr, nullptr, nullptr, 0}, x, cs, p.partial, (int)M, cs));
        } else {
            dim3 grid(C, p.nblk);
            bool vec = (HW % 4 == 0) && al16(x);
            if (vec) k_bn_partial_nchw<0, true><<<grid, BN_THREADS, 0, ctx->stream>>>(x, nullptr, BnMask{nullptr, nullptr, nullptr, 0}, nullptr, p.partial, B, C, HW, p.chunk);
            else k_bn_partial_nchw<0, false><<<grid, BN_THREADS, 0, ctx->stream>>>(x, nullptr, BnMask{nullptr, nullptr, nullptr, 0}, nullptr, p.partial, B, C, HW, p.chunk);
        }
        DL4J_CHECK_LAUNCH();
    }
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_finalize_fwd, dim3(dl4j_ceil_div(p.cpad, 32)), dim3(32, BN_FY), 0, ctx->stream, p.partial, p.nblk, x,
                                  layout == DL4JB200_NHWC ? 1 : HW, gamma, beta, gmean, gvar, save_mean, save_std, p.coef, C, p.cpad,
                                  (double)M, decay, eps, training, minibatch));
    DL4J_CHECK_LAUNCH();
    if (layout == DL4JB200_NHWC) {
        DL4J_CUDA_TRY(bn_launch_apply_nhwc<0>(ctx, bn_grid(ctx, total / 4), x, residual, nullptr, y,
                                      p.coef, total / 4, cs, act | (ctx->round_act << 8)));
    } else {
        bool vec = (HW % 4 == 0) && al16(x) && al16(y);
        if (vec) k_bn_apply_nchw<0, true><<<bn_grid(ctx, total / 4), 256, 0, ctx->stream>>>(x, residual, nullptr, y, p.coef, total, C, HW, act | (ctx->round_act << 8));
        else k_bn_apply_nchw<0, false><<<bn_grid(ctx, total), 256, 0, ctx->stream>>>(x, residual, nullptr, y, p.coef, total, C, HW, act | (ctx->round_act << 8));
    }
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// Training forward on statistics some producer already took (dl4jb200_conv_fwd_bnstats): finalize + apply only.
int dl4jb200_bn_fwd_prestats(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta, float *gmean,
                             float *gvar, float *save_mean, float *save_std, int B, int C, int H, int W, int cs, double decay,
                             double eps, int minibatch, int act, const float *residual, const float *pivot,
                             const float *stat_partial, int stat_rows) {
    DL4J_REQUIRE(ctx && x && y && gmean && gvar && pivot && stat_partial, "bn_fwd_prestats: NULL argument");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && stat_rows > 0, "bn_fwd_prestats: bad shape");
    DL4J_REQUIRE(cs == C && cs % 4 == 0 && al16(x) && al16(y) && (!residual || al16(residual)),
                 "bn_fwd_prestats: NHWC with cs == C, cs %% 4 == 0 and 16-byte aligned tensors");
    // coefficients live in the context scratch, like bn_fwd's
    float *coef = (float *)dl4j_workspace(ctx, 0, 5 * (size_t)cs * sizeof(float));
    if (!coef) return DL4JB200_ERR_CUDA;
    const long long M = (long long)B * H * W;
    const size_t total = (size_t)M * cs;
    // pivot may be gmean itself: k_bn_finalize_fwd reads the pivot of a channel before it updates that channel's gmean
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_finalize_fwd, dim3(dl4j_ceil_div(cs, 32)), dim3(32, BN_FY), 0, ctx->stream, stat_partial, stat_rows,
                                  pivot, 1, gamma, beta, gmean, gvar, save_mean, save_std, coef, C, cs, (double)M, decay, eps, 1,
                                  minibatch));
    DL4J_CHECK_LAUNCH();
    DL4J_CUDA_TRY(bn_launch_apply_nhwc<0>(ctx, bn_grid(ctx, total / 4), x, residual, nullptr, y,
                                  coef, total / 4, cs, act | (ctx->round_act << 8)));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// bn_fwd_prestats whose apply pass also performs the MAX pooling layer that consumes the normalised (and activated) tensor.
int dl4jb200_bn_fwd_prestats_maxpool(dl4jb200_ctx *ctx, const float *x, float *y_pool, int32_t *argmax, const float *gamma,
                                     const float *beta, float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C,
                                     int H, int W, int cs, double decay, double eps, int minibatch, int act, const float *pivot,
                                     const float *stat_partial, int stat_rows, const dl4jb200_conv_desc *pool) {
    DL4J_REQUIRE(ctx && x && y_pool && gmean && gvar && pivot && stat_partial && pool, "bn_fwd_prestats_maxpool: NULL argument");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && stat_rows > 0, "bn_fwd_prestats_maxpool: bad shape");
    DL4J_REQUIRE(cs == C && cs % 4 == 0 && al16(x) && al16(y_pool) && (!argmax || al16(argmax)),
                 "bn_fwd_prestats_maxpool: NHWC with cs == C, cs %% 4 == 0 and 16-byte aligned tensors");
    DL4J_REQUIRE(pool->batch == B && pool->c_in == C && pool->in_h == H && pool->in_w == W && pool->dil_h == 1 && pool->dil_w == 1 &&
                     pool->k_h > 0 && pool->k_w > 0 && pool->stride_h > 0 && pool->stride_w > 0 && pool->out_h > 0 && pool->out_w > 0,
                 "bn_fwd_prestats_maxpool: the pooling descriptor does not describe a window over this tensor (no dilation)");
    const size_t total = (size_t)B * pool->out_h * pool->out_w * (cs / 4);
    DL4J_REQUIRE(total < (1ull << 32), "bn_fwd_prestats_maxpool: more than 2^32 output quads");
    float *coef = (float *)dl4j_workspace(ctx, 0, 5 * (size_t)cs * sizeof(float));
    if (!coef) return DL4JB200_ERR_CUDA;
    const long long M = (long long)B * H * W;
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_finalize_fwd, dim3(dl4j_ceil_div(cs, 32)), dim3(32, BN_FY), 0, ctx->stream, stat_partial, stat_rows,
                                  pivot, 1, gamma, beta, gmean, gvar, save_mean, save_std, coef, C, cs, (double)M, decay, eps, 1,
                                  minibatch));
    DL4J_CHECK_LAUNCH();
    BnPoolGeom g{B, H, W, pool->out_h, pool->out_w, pool->k_h, pool->k_w, pool->stride_h, pool->stride_w, pool->pad_top, pool->pad_left, cs, cs,
                 FastDiv((uint32_t)(cs / 4)), FastDiv((uint32_t)pool->out_w), FastDiv((uint32_t)pool->out_h)};
    const int grid = bn_grid(ctx, total), code = act | (ctx->round_act << 8);
    if (g.kH == 3 && g.kW == 3)
        DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_act_maxpool_nhwc<3>, dim3(grid), dim3(256), 0, ctx->stream, x, y_pool, argmax, (const float *)coef, g, total, code));
    else if (g.kH == 2 && g.kW == 2)
        DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_act_maxpool_nhwc<2>, dim3(grid), dim3(256), 0, ctx->stream, x, y_pool, argmax, (const float *)coef, g, total, code));
    else
        DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_act_maxpool_nhwc<0>, dim3(grid), dim3(256), 0, ctx->stream, x, y_pool, argmax, (const float *)coef, g, total, code));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_bn_bwd(dl4jb200_ctx *ctx, const float *x, const float *eps, const float *gamma, const float *beta,
                    const float *save_mean, const float *save_std, int act, float *dx, float *dgamma, float *dbeta, int B,
                    int C, int H, int W, int layout, int cs) {
    DL4J_REQUIRE(ctx && x && eps && save_mean && save_std && dx, "bn_bwd: NULL argument");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "bn_bwd: bad shape");
    DL4J_REQUIRE((long long)B * H * W < (1LL << 31), "bn_bwd: B*H*W must stay below 2^31");
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "bn_bwd: unknown layout %d", layout);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || (cs >= C && cs % 4 == 0 && al16(x) && al16(eps) && al16(dx)),
                 "bn_bwd: NHWC needs cs %% 4 == 0, cs >= C and 16-byte aligned tensors");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act == DL4JB200_ACT_RELU, "bn_bwd: only IDENTITY / RELU can be fused (got %d)", act);
    BnPlan p;
    int rc = bn_plan(ctx, B, C, H, W, layout, cs, p);
    if (rc) return rc;
    const int HW = H * W;
    const long long M = (long long)B * HW;
    const size_t total = (size_t)M * (layout == DL4JB200_NHWC ? cs : C);
    const bool relu = act == DL4JB200_ACT_RELU;
    const BnMask mk{gamma, beta, save_std, relu ? 1 : 0};
    const float *mask_beta = relu ? p.coef + 4 * (size_t)p.cpad : nullptr;
    bool vec = (HW % 4 == 0) && al16(x) && al16(eps) && al16(dx);
    if (layout == DL4JB200_NHWC) {
        DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_partial_nhwc<1>, dim3(bn_nhwc_groups(cs), p.nblk), dim3(BN_THREADS), 0, ctx->stream, x, eps, mk, save_mean, C, p.partial,
                                      (int)M, cs));
    } else {
        dim3 grid(C, p.nblk);
        if (vec) k_bn_partial_nchw<1, true><<<grid, BN_THREADS, 0, ctx->stream>>>(x, eps, mk, save_mean, p.partial, B, C, HW, p.chunk);
        else k_bn_partial_nchw<1, false><<<grid, BN_THREADS, 0, ctx->stream>>>(x, eps, mk, save_mean, p.partial, B, C, HW, p.chunk);
    }
    DL4J_CHECK_LAUNCH();
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_finalize_bwd, dim3(dl4j_ceil_div(p.cpad, 32)), dim3(32, BN_FY), 0, ctx->stream, p.partial, p.nblk, gamma,
                                  save_mean, save_std, relu ? beta : nullptr, dgamma, dbeta, p.coef, C, p.cpad, (double)M));
    DL4J_CHECK_LAUNCH();
    if (layout == DL4JB200_NHWC) {
        DL4J_CUDA_TRY(bn_launch_apply_nhwc<1>(ctx, bn_grid(ctx, total / 4), x, eps, mask_beta, dx, p.coef,
                                      total / 4, cs, ctx->round_act << 8));
    } else {
        if (vec) k_bn_apply_nchw<1, true><<<bn_grid(ctx, total / 4), 256, 0, ctx->stream>>>(x, eps, mask_beta, dx, p.coef, total, C, HW, ctx->round_act << 8);
        else k_bn_apply_nchw<1, false><<<bn_grid(ctx, total), 256, 0, ctx->stream>>>(x, eps, mask_beta, dx, p.coef, total, C, HW, ctx->round_act << 8);
    }
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// Backward on sums a producer already took (dl4jb200_conv_bwd_data_bnstats): finalize + apply only.
int dl4jb200_bn_bwd_prestats(dl4jb200_ctx *ctx, const float *x, const float *eps, const float *gamma, const float *beta,
                             const float *save_mean, const float *save_std, int act, float *dx, float *dgamma, float *dbeta, int B,
                             int C, int H, int W, int cs, const float *stat_partial, int stat_rows) {
    DL4J_REQUIRE(ctx && x && eps && save_mean && save_std && dx && stat_partial, "bn_bwd_prestats: NULL argument");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && stat_rows > 0, "bn_bwd_prestats: bad shape");
    DL4J_REQUIRE((long long)B * H * W < (1LL << 31), "bn_bwd_prestats: B*H*W must stay below 2^31");
    DL4J_REQUIRE(cs == C && cs % 4 == 0 && al16(x) && al16(eps) && al16(dx), "bn_bwd_prestats: NHWC with cs == C, cs %% 4 == 0, 16-byte aligned tensors");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act == DL4JB200_ACT_RELU, "bn_bwd_prestats: only IDENTITY / RELU can be fused (got %d)", act);
    float *coef = (float *)dl4j_workspace(ctx, 0, 5 * (size_t)cs * sizeof(float));
    if (!coef) return DL4JB200_ERR_CUDA;
    const long long M = (long long)B * H * W;
    const size_t total = (size_t)M * cs;
    const bool relu = act == DL4JB200_ACT_RELU;
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_bn_finalize_bwd, dim3(dl4j_ceil_div(cs, 32)), dim3(32, BN_FY), 0, ctx->stream, stat_partial, stat_rows, gamma,
                                  save_mean, save_std, relu ? beta : nullptr, dgamma, dbeta, coef, C, cs, (double)M));
    DL4J_CHECK_LAUNCH();
    DL4J_CUDA_TRY(bn_launch_apply_nhwc<1>(ctx, bn_grid(ctx, total / 4), x, eps,
                                  relu ? (const float *)(coef + 4 * (size_t)cs) : (const float *)nullptr, dx, (const float *)coef, total / 4, cs,
                                  ctx->round_act << 8));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// MAX-pooling backward (stride-2 windows of at most 3x3, NHWC, saved taps) that also leaves the backward sums of the
// BatchNormalization whose output the pooling layer consumed: dx is what dl4jb200_pool_bwd writes; stat_partial receives
// *rows_out rows ([2][cs] each) for dl4jb200_bn_bwd_prestats. stat_partial must hold 4 * (number of SMs) rows.
int dl4jb200_pool_bwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *pool, const float *eps, const int32_t *argmax, float *dx,
                              int cs, const float *bn_x, const float *gamma, const float *beta, const float *save_mean,
                              const float *save_std, int bn_relu, float *stat_partial, int *rows_out) {
    DL4J_REQUIRE(ctx && pool && eps && argmax && dx && bn_x && save_mean && save_std && stat_partial && rows_out,
                 "pool_bwd_bnstats: NULL argument");
    const int C = pool->c_in, H = pool->in_h, W = pool->in_w, B = pool->batch;
    if (pool->stride_h != 2 || pool->stride_w != 2 || pool->pad_top < 0 || pool->pad_left < 0 || pool->dil_h != 1 || pool->dil_w != 1 ||
        pool->k_h > 3 || pool->k_w > 3 || cs != C || C % 4 != 0 || 256 % (C / 4) != 0) {
        dl4j_set_error("pool_bwd_bnstats: needs stride-2 windows of at most 3x3 without dilation, non-negative padding, cs == C and "
                       "C/4 dividing 256 (got k %dx%d stride %dx%d C %d cs %d)", pool->k_h, pool->k_w, pool->stride_h, pool->stride_w, C, cs);
        return DL4JB200_NOT_SUPPORTED;
    }
    DL4J_REQUIRE(B > 0 && H > 0 && W > 0 && pool->out_h > 0 && pool->out_w > 0 && pool->k_h > 0 && pool->k_w > 0, "pool_bwd_bnstats: bad shape");
    DL4J_REQUIRE(al16(eps) && al16(argmax) && al16(dx) && al16(bn_x) && al16(save_mean) && al16(stat_partial),
                 "pool_bwd_bnstats: tensors must be 16-byte aligned");
    const int Hp = (H + 1) / 2, Wp = (W + 1) / 2;
    const size_t total = (size_t)B * Hp * Wp * (C / 4);
    DL4J_REQUIRE(total < (1ull << 32), "pool_bwd_bnstats: tensor too large");
    size_t blocks = (total + 255) / 256;
    const size_t cap = (size_t)ctx->sm_count * 4;  // one partial row per block: keep the finalize fold short
    if (blocks > cap) blocks = cap;
    PoolBstGeom g{B, H, W, pool->out_h, pool->out_w, pool->k_h, pool->k_w, pool->pad_top, pool->pad_left, cs, Hp, Wp, ctx->round_act,
                  FastDiv((uint32_t)(C / 4)), FastDiv((uint32_t)Wp), FastDiv((uint32_t)Hp)};
    const BnMask mk{gamma, beta, save_std, bn_relu ? 1 : 0};
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_maxpool_bwd_bnsums_nhwc, dim3((unsigned)blocks), dim3(256), 0, ctx->stream, eps, argmax, dx, bn_x, mk, save_mean,
                                  stat_partial, g, total));
    DL4J_CHECK_LAUNCH();
    *rows_out = (int)blocks;
    return DL4JB200_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the batch-norm forward (no counterpart in the reference).
//
// dl4jb200_bn_fwd keeps its per-block partial sums and per-channel coefficients in the context's scratch, which is
// obtained from cudaMalloc the first time (and whenever a larger tensor comes along). Some callers must never
// trigger a device allocation: a serving or training loop that cannot afford the device-wide synchronisation a
// cudaMalloc implies, a process that has handed the whole card to another allocator (nd4j workspaces or a pool
// sized to the GPU) so that a driver allocation may simply fail, or a host layer whose rules forbid cudaMalloc /
// cudaFree outside start-up. For them the same three kernels run with their scratch in the fixed array below instead.
//
// DL4JB200_BN_NOALLOC_FLOATS floats (4 MiB) are reserved on every GPU that loads this library, until the process
// exits. That bounds the channel count (see DL4JB200_BN_NOALLOC_MAX_CHANNELS); B, H and W are bounded only by
// B*H*W < 2^31 (the same limit as dl4jb200_bn_fwd), because the number of partial blocks is reduced to fit. All calls share the one array: two calls that overlap on the GPU
// (different streams, different contexts on the same device) race on it -- keep such calls on one stream or
// serialise them. Nothing in the array has to start at zero: every slot a call reads is written earlier in that
// same call, so leftovers from the previous call are harmless.
#define DL4JB200_BN_NOALLOC_FLOATS (1u << 20)
#define DL4JB200_BN_NOALLOC_MAX_CHANNELS 4096

static __device__ __align__(256) float g_bn_noalloc_scratch[DL4JB200_BN_NOALLOC_FLOATS];

// In host code `g_bn_noalloc_scratch` is only a handle for the symbol, not a device address; the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current. The table is filled without a
// lock: two threads racing on a device's first call both store the same address (the lookup is idempotent), so the
// worst case is one redundant lookup.
static float *g_bn_noalloc_ptr[64] = {nullptr};

static int bn_noalloc_scratch(dl4jb200_ctx *ctx, float **out) {
    DL4J_REQUIRE(ctx->device >= 0 && ctx->device < 64, "bn_fwd_noalloc: device index %d out of range", ctx->device);
    if (!g_bn_noalloc_ptr[ctx->device]) {
        int cur = -1;
        DL4J_CUDA_TRY(cudaGetDevice(&cur));
        if (cur != ctx->device) DL4J_CUDA_TRY(cudaSetDevice(ctx->device));
        float *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_bn_noalloc_scratch);
        if (cur != ctx->device && cur >= 0) cudaSetDevice(cur);
        if (e != cudaSuccess || !p) {
            dl4j_set_error("bn_fwd_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return DL4JB200_ERR_CUDA;
        }
        g_bn_noalloc_ptr[ctx->device] = p;
    }
    *out = g_bn_noalloc_ptr[ctx->device];
    return DL4JB200_OK;
}

extern "C" int dl4jb200_bn_fwd_noalloc(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta,
                                       float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C, int H,
                                       int W, int layout, int cs, double decay, double eps, int training, int minibatch,
                                       int act) {
    DL4J_REQUIRE(ctx && x && y && gmean && gvar, "bn_fwd_noalloc: NULL argument");
    DL4J_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "bn_fwd_noalloc: bad shape");
    DL4J_REQUIRE((long long)B * H * W < (1LL << 31), "bn_fwd_noalloc: B*H*W must stay below 2^31");
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "bn_fwd_noalloc: unknown layout %d", layout);
    DL4J_REQUIRE(layout == DL4JB200_NCHW || (cs >= C && cs % 4 == 0 && al16(x) && al16(y)),
                 "bn_fwd_noalloc: NHWC needs cs %% 4 == 0, cs >= C and 16-byte aligned tensors");
    const int cpad = layout == DL4JB200_NHWC ? cs : C;
    if (cpad > DL4JB200_BN_NOALLOC_MAX_CHANNELS) {
        dl4j_set_error("bn_fwd_noalloc: %d channels exceed the fixed scratch (maximum %d channels; use dl4jb200_bn_fwd, "
                       "which sizes its scratch on demand)", cpad, DL4JB200_BN_NOALLOC_MAX_CHANNELS);
        return DL4JB200_ERR_INVALID_ARG;
    }
    float *scratch = nullptr;
    int rc = bn_noalloc_scratch(ctx, &scratch);
    if (rc) return rc;
    // same plan as bn_plan(), with the number of partial blocks capped so that 4*cpad + nblk*2*cpad floats fit
    const int HW = H * W;
    const long long M = (long long)B * HW;
    const int max_blk = (int)((DL4JB200_BN_NOALLOC_FLOATS - 4u * (unsigned)cpad) / (2u * (unsigned)cpad));
    int nblk;
    long long chunk = 0;
    if (layout == DL4JB200_NHWC) {
        nblk = bn_nhwc_splits(ctx->sm_count, cs, M);
        if (nblk > max_blk) nblk = max_blk;
    } else {
        long long cap = ((long long)ctx->sm_count * 8 + C - 1) / C, s = M / 4096;
        if (s > cap) s = cap;
        if (s > max_blk) s = max_blk;
        if (s < 1) s = 1;
        chunk = ((M + s - 1) / s + 3) / 4 * 4;
        nblk = (int)((M + chunk - 1) / chunk);
    }
    float *coef = scratch, *partial = scratch + 4 * (size_t)cpad;
    const size_t total = (size_t)M * cpad;
    if (training) {
        if (layout == DL4JB200_NHWC) {
            k_bn_partial_nhwc<0><<<dim3(bn_nhwc_groups(cs), nblk), BN_THREADS, 0, ctx->stream>>>(x, nullptr, BnMask{nullptr, nullptr, nullptr, 0}, x, cs, partial, (int)M, cs);
        } else {
            dim3 grid(C, nblk);
            if ((HW % 4 == 0) && al16(x)) k_bn_partial_nchw<0, true><<<grid, BN_THREADS, 0, ctx->stream>>>(x, nullptr, BnMask{nullptr, nullptr, nullptr, 0}, nullptr, partial, B, C, HW, chunk);
            else k_bn_partial_nchw<0, false><<<grid, BN_THREADS, 0, ctx->stream>>>(x, nullptr, BnMask{nullptr, nullptr, nullptr, 0}, nullptr, partial, B, C, HW, chunk);
        }
        DL4J_CHECK_LAUNCH();
    }
    k_bn_finalize_fwd<<<dl4j_ceil_div(cpad, 32), dim3(32, BN_FY), 0, ctx->stream>>>(
        partial, nblk, x, layout == DL4JB200_NHWC ? 1 : HW, gamma, beta, gmean, gvar, save_mean, save_std, coef, C, cpad,
        (double)M, decay, eps, training, minibatch);
    DL4J_CHECK_LAUNCH();
    if (layout == DL4JB200_NHWC) {
        k_bn_apply_nhwc<0><<<bn_grid(ctx, total / 4), 256, 0, ctx->stream>>>(x, nullptr, nullptr, y, coef, total / 4, cs, act | (ctx->round_act << 8));
    } else {
        if ((HW % 4 == 0) && al16(x) && al16(y)) k_bn_apply_nchw<0, true><<<bn_grid(ctx, total / 4), 256, 0, ctx->stream>>>(x, nullptr, nullptr, y, coef, total, C, HW, act | (ctx->round_act << 8));
        else k_bn_apply_nchw<0, false><<<bn_grid(ctx, total), 256, 0, ctx->stream>>>(x, nullptr, nullptr, y, coef, total, C, HW, act | (ctx->round_act << 8));
    }
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}
