// Convolution entry points: geometry, filter packing, algorithm dispatch, split-K reduction, bias gradient,
// and NCHW staging for callers that hold DL4J-layout ('c' order NCHW) tensors.
#include <algorithm>
#include <unordered_map>
#include <vector>

#include "conv_common.cuh"

// ------------------------------------------------------------------ geometry
static void geom_finish(ConvGeom &g) {
    g.taps = g.kH * g.kW;
    g.K = g.taps * g.Cg;
    g.Kp = dl4j_round_up(g.K, 32);
    g.M = g.B * g.Hm * g.Wm;
    g.div_Wm = FastDiv((uint32_t)g.Wm);
    g.div_HWm = FastDiv((uint32_t)(g.Hm * g.Wm));
    g.div_Cg = FastDiv((uint32_t)g.Cg);
    g.div_kW = FastDiv((uint32_t)g.kW);
    g.fast = g.kH <= 16 && g.kW <= 16 && (!g.transposed || (g.sH == 1 && g.sW == 1) || g.taps == 1);
    g.use_tma = 0;
    g.remap = 0; g.rm_sy = g.rm_sx = 1; g.rm_H = g.Hm; g.rm_W = g.Wm;
}

ConvGeom conv_geom_fwd(const dl4jb200_conv_desc *d, int x_cs) {
    ConvGeom g;
    g.B = d->batch; g.Hs = d->in_h; g.Ws = d->in_w; g.cs_src = x_cs;
    g.Hm = d->out_h; g.Wm = d->out_w;
    g.kH = d->k_h; g.kW = d->k_w; g.sH = d->stride_h; g.sW = d->stride_w;
    g.pT = d->pad_top; g.pL = d->pad_left; g.dH = d->dil_h; g.dW = d->dil_w;
    g.Cg = dl4j_round_up(d->c_in, 4);
    g.transposed = 0;
    geom_finish(g);
    return g;
}

ConvGeom conv_geom_bwd_data(const dl4jb200_conv_desc *d, int delta_cs) {
    ConvGeom g;
    g.B = d->batch; g.Hs = d->out_h; g.Ws = d->out_w; g.cs_src = delta_cs;
    g.Hm = d->in_h; g.Wm = d->in_w;
    g.kH = d->k_h; g.kW = d->k_w; g.sH = d->stride_h; g.sW = d->stride_w;
    g.pT = d->pad_top; g.pL = d->pad_left; g.dH = d->dil_h; g.dW = d->dil_w;
    g.Cg = dl4j_round_up(d->c_out, 4);
    g.transposed = 1;
    geom_finish(g);
    return g;
}

// ------------------------------------------------------------------ filter packing
// DL4J filter W[co][ci][kh][kw] ('c' order, ConvolutionParamInitializer.java:144-155) ->
//   forward operand  Wf[n = co][k = (kh*kW+kw)*Cg + ci]       (transposed = 0)
//   backprop-data    Wb[n = ci][k = (kh*kW+kw)*Cg + co]       (transposed = 1)
// rows padded with zeros to `rows_p`, K padded to Kp. round_tf32: round-to-nearest into TF32 so that the
// tensor core's truncation of operand mantissas does not bias the filter side.
__global__ void __launch_bounds__(256) k_pack_filter(const float *__restrict__ w, float *__restrict__ wp, int Cout, int Cin,
                                                     int kH, int kW, int Cg, int Kp, int rows_p, int transposed,
                                                     int round_tf32, int flip) {
    size_t total = (size_t)rows_p * Kp, stride = (size_t)gridDim.x * blockDim.x;
    const int taps = kH * kW;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int n = (int)(i / Kp), k = (int)(i % Kp);
        int tap = k / Cg, c = k % Cg;
        float v = 0.f;
        if (tap < taps) {
            int co = transposed ? c : n, ci = transposed ? n : c;
            if (co < Cout && ci < Cin) v = w[((size_t)co * Cin + ci) * taps + (flip ? taps - 1 - tap : tap)];
        }
        wp[i] = round_tf32 ? to_tf32_rna(v) : v;
    }
}

// Backprop-data operand of ONE stride phase of a strided convolution (see bwd_data_phases): the sub-filter made of the
// taps kh = a + sH*j, kw = b + sW*j', transposed and tap-flipped:  Wb[n = ci][k = (J*J' - 1 - (j*J' + j'))*Cg + co]
__global__ void __launch_bounds__(256) k_pack_filter_phase(const float *__restrict__ w, float *__restrict__ wp, int Cout, int Cin,
                                                           int kH, int kW, int Cg, int Kp, int rows_p, int a, int b, int sH, int sW,
                                                           int Ja, int Jb) {
    size_t total = (size_t)rows_p * Kp, stride = (size_t)gridDim.x * blockDim.x;
    const int taps = Ja * Jb;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int n = (int)(i / Kp), k = (int)(i % Kp);
        int tap = k / Cg, co = k % Cg;
        float v = 0.f;
        if (tap < taps && co < Cout && n < Cin) {
            const int ft = taps - 1 - tap, j = ft / Jb, jj = ft - j * Jb;
            v = w[(((size_t)co * Cin + n) * kH + (a + sH * j)) * kW + (b + sW * jj)];
        }
        wp[i] = to_tf32_rna(v);
    }
}

// ------------------------------------------------------------------ pre-packed filters
// conv_fwd / conv_bwd pack the filter on every call (one small kernel each: 106 launches per ResNet-50 step). A caller
// that knows when the parameters change registers each filter once and re-packs all of them with ONE launch after
// every parameter update (dl4jb200_filters_repack); convolution calls on a registered `w` then use the stored operand.
struct PackJob {  // device-visible
    const float *w;
    float *wp;
    unsigned long long start;  // unused by the kernel (kept so that a job is self-describing in a debugger)
    int Cout, Cin, kH, kW, Cg, Kp, rows_p, transposed, round_tf32, flip;
    // divisors of the pack kernel's index arithmetic (it is ALU-bound on integer division otherwise): taps*Cg, Cg, taps,
    // Cin/4, 32*taps
    FastDiv d_kk, d_Cg, d_taps, d_c4, d_span32;
};
static void pack_job_divs(PackJob &j) {
    const uint32_t taps = (uint32_t)(j.kH * j.kW);
    j.d_kk = FastDiv(taps * (uint32_t)j.Cg);
    j.d_Cg = FastDiv((uint32_t)j.Cg);
    j.d_taps = FastDiv(taps);
    j.d_c4 = FastDiv((uint32_t)(j.Cin >> 2 ? j.Cin >> 2 : 1));
    j.d_span32 = FastDiv(32u * taps);
}
struct PackUnit;
struct FilterEntry {
    PackJob fwd, bwd;  // bwd.wp == nullptr: no backprop-data operand registered
};
struct FilterRegistry {
    std::unordered_map<const float *, FilterEntry> entries;
    PackJob *jobs_dev = nullptr;
    PackUnit *units_dev = nullptr;
    int njobs = 0, nunits = 0;
    bool table_dirty = true, valid = false;
};

void conv_filters_free(dl4jb200_ctx *ctx) {
    FilterRegistry *r = ctx->filters;
    if (!r) return;
    for (auto &kv : r->entries) {
        if (kv.second.fwd.wp) cudaFree(kv.second.fwd.wp);
        if (kv.second.bwd.wp) cudaFree(kv.second.bwd.wp);
    }
    if (r->jobs_dev) cudaFree(r->jobs_dev);
    if (r->units_dev) cudaFree(r->units_dev);
    delete r;
    ctx->filters = nullptr;
}

// One block per work unit. Forward operand: a unit is a run of filter rows (co); the source W[co][ci][tap] is read as one
// contiguous span into shared memory and written out in operand order Wf[co][tap*Cg + ci] (both sides coalesced).
// Backprop-data operand: a unit is a 32(co) x 32(ci) tile with all its taps; each co row of the tile is a contiguous
// 32*taps span in W, and the operand rows Wb[ci][tap'*Cg + co] receive 32 consecutive co (128 bytes) per (ci, tap).
// Padding rows / columns of the operands are zeroed once at registration and never written again.
struct PackUnit {
    int job, a, b, pad;  // fwd: a = first row, b = row count; bwd: a = co0, b = ci0
};
constexpr int PACK_SMEM_FLOATS = 12288;  // 48 KB

__global__ void __launch_bounds__(256) k_pack_filters_batched(const PackJob *__restrict__ jobs, const PackUnit *__restrict__ units) {
    extern __shared__ float sm[];
    const PackUnit u = units[blockIdx.x];
    const PackJob j = jobs[u.job];
    const int taps = j.kH * j.kW, row_src = j.Cin * taps;
    if (!j.transposed) {
        const int n0 = u.a, rows = u.b;
        const float *src = j.w + (size_t)n0 * row_src;
        const int span = rows * row_src;
        if (taps == 1) {  // operand row = filter row, only the pitch differs
            if ((j.Cin & 3) == 0 && (j.Kp & 3) == 0 && ((uintptr_t)src & 15) == 0) {  // 128-bit both ways
                const int c4 = j.Cin >> 2;
#pragma unroll 4
                for (int e = threadIdx.x; e < rows * c4; e += 256) {
                    const int r = (int)j.d_c4.div((uint32_t)e), q = e - r * c4;
                    float4 v = reinterpret_cast<const float4 *>(src)[e];
                    if (j.round_tf32) { v.x = to_tf32_rna(v.x); v.y = to_tf32_rna(v.y); v.z = to_tf32_rna(v.z); v.w = to_tf32_rna(v.w); }
                    *reinterpret_cast<float4 *>(j.wp + (size_t)(n0 + r) * j.Kp + q * 4) = v;
                }
                return;
            }
#pragma unroll 4
            for (int e = threadIdx.x; e < span; e += 256) {
                const int r = e / j.Cin, ci = e - r * j.Cin;
                const float v = src[e];
                j.wp[(size_t)(n0 + r) * j.Kp + ci] = j.round_tf32 ? to_tf32_rna(v) : v;
            }
            return;
        }
        const bool staged = span <= PACK_SMEM_FLOATS;
        if (staged) {
#pragma unroll 8
            for (int e = threadIdx.x; e < span; e += 256) sm[e] = src[e];
            __syncthreads();
        }
        const int kk = taps * j.Cg;
#pragma unroll 4
        for (int e = threadIdx.x; e < rows * kk; e += 256) {
            const int r = (int)j.d_kk.div((uint32_t)e), k = e - r * kk, tap = (int)j.d_Cg.div((uint32_t)k), ci = k - tap * j.Cg;
            float v = 0.f;
            if (ci < j.Cin) {
                const int o = (r * j.Cin + ci) * taps + (j.flip ? taps - 1 - tap : tap);
                v = staged ? sm[o] : src[o];
            }
            j.wp[(size_t)(n0 + r) * j.Kp + k] = j.round_tf32 ? to_tf32_rna(v) : v;
        }
    } else {
        const int co0 = u.a, nco = min(32, j.Cout - co0);
        // a unit is a run of 32-wide ci tiles (u.pad of them): enough work per block for the small-tap filters
        for (int sub = 0; sub < u.pad; sub++) {
        const int ci0 = u.b + 32 * sub;
        if (ci0 >= j.Cin) break;
        const int nci = min(32, j.Cin - ci0);
        if (sub) __syncthreads();  // the staging buffer is reused
        const int span = nci * taps;  // contiguous floats per co row of the tile
        const bool staged = 32 * span <= PACK_SMEM_FLOATS;
        if (staged) {
#pragma unroll 8
            for (int e = threadIdx.x; e < nco * span; e += 256) {
                const int r = nci == 32 ? (int)j.d_span32.div((uint32_t)e) : e / span, t = e - r * span;
                sm[r * span + t] = j.w[((size_t)(co0 + r) * j.Cin + ci0) * taps + t];
            }
            __syncthreads();
        }
        // destination: row ci, column tap'*Cg + co; lanes run along co
#pragma unroll 4
        for (int e = threadIdx.x; e < nci * taps * 32; e += 256) {
            const int cl = e & 31, t2 = e >> 5, il = (int)j.d_taps.div((uint32_t)t2), tap = t2 - il * taps;
            if (cl >= nco) continue;
            const int st = j.flip ? taps - 1 - tap : tap;
            const float v = staged ? sm[cl * span + il * taps + st] : j.w[((size_t)(co0 + cl) * j.Cin + ci0 + il) * taps + st];
            j.wp[(size_t)(ci0 + il) * j.Kp + (size_t)tap * j.Cg + co0 + cl] = j.round_tf32 ? to_tf32_rna(v) : v;
        }
        }
    }
}

// the stored operand for (w, direction) if one is registered, current, and laid out exactly as this call needs it
static const float *filters_lookup(dl4jb200_ctx *ctx, const float *w, bool bwd, int Cg, int Kp, int rows_p, int round_tf32, int flip) {
    FilterRegistry *r = ctx->filters;
    if (!r || !r->valid) return nullptr;
    auto it = r->entries.find(w);
    if (it == r->entries.end()) return nullptr;
    const PackJob &j = bwd ? it->second.bwd : it->second.fwd;
    if (!j.wp || j.Cg != Cg || j.Kp != Kp || j.rows_p != rows_p || j.round_tf32 != round_tf32) return nullptr;
    if (j.kH * j.kW > 1 && j.flip != flip) return nullptr;
    return j.wp;
}

// dW[co][ci][kh][kw] = sum_s partial[s][co][(kh*kW+kw)*Cg + ci]   (fixed order: deterministic)
// db_partial (optional): [splits][rows_p] per-split bias-gradient sums from the filter-gradient GEMM's extra column.
__global__ void __launch_bounds__(256) k_wgrad_finalize(const float *__restrict__ partial, float *__restrict__ dw, int Cout,
                                                        int Cin, int taps, int Cg, int Kp, int rows_p, int splits,
                                                        const float *__restrict__ db_partial, float *__restrict__ db) {
    size_t total = (size_t)Cout * Cin * taps, stride = (size_t)gridDim.x * blockDim.x;
    pdl_trigger();
    pdl_wait();
    if (db_partial) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)Cout; i += stride) {
            float acc = 0.f;
            for (int s = 0; s < splits; s++) acc += db_partial[(size_t)s * rows_p + i];
            db[i] = acc;
        }
    }
    // (output order; enumerating in the partial layout's order instead -- coalesced reads, strided writes -- measured
    // 40 % slower under ncu: neighbouring taps share sectors here and L1 absorbs the re-reads)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int tap = (int)(i % taps);
        size_t t = i / taps;
        int ci = (int)(t % Cin), co = (int)(t / Cin);
        const float *p = partial + (size_t)co * Kp + (size_t)tap * Cg + ci;
        const size_t ss = (size_t)rows_p * Kp;
        // four independent partial sums: the loads of one round are all in flight together (the fold is latency-bound);
        // the combination order is fixed, so the result stays deterministic
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

// Many-split form (small filters on large images: 59-146 splits of a few thousand outputs). A block reduces 32 outputs with
// 8 threads each: thread (o, l) sums splits l, l+8, ... and the eight lane sums are combined in lane order through shared
// memory -- a fixed tree, so still deterministic -- instead of one thread walking 100+ strided partials on its own.
__global__ void __launch_bounds__(256) k_wgrad_finalize_wide(const float *__restrict__ partial, float *__restrict__ dw, int Cout,
                                                             int Cin, int taps, int Cg, int Kp, int rows_p, int splits,
                                                             const float *__restrict__ db_partial, float *__restrict__ db) {
    __shared__ float red[8][33];
    pdl_trigger();
    pdl_wait();
    const int o = threadIdx.x & 31, l = threadIdx.x >> 5;
    const size_t nw = (size_t)Cout * Cin * taps, total = nw + (db_partial ? (size_t)Cout : 0);
    for (size_t base = (size_t)blockIdx.x * 32; base < total; base += (size_t)gridDim.x * 32) {
        const size_t i = base + o;
        float acc = 0.f;
        float *dst = nullptr;
        if (i < total) {
            const float *p;
            size_t ss;
            if (i < nw) {
                const int tap = (int)(i % taps);
                const size_t t = i / taps;
                const int ci = (int)(t % Cin), co = (int)(t / Cin);
                p = partial + (size_t)co * Kp + (size_t)tap * Cg + ci;
                ss = (size_t)rows_p * Kp;
                dst = dw + i;
            } else {
                p = db_partial + (i - nw);
                ss = (size_t)rows_p;
                dst = db + (i - nw);
            }
            float a0 = 0.f, a1 = 0.f;
            int sp = l;
            for (; sp + 8 < splits; sp += 16) { a0 += p[(size_t)sp * ss]; a1 += p[(size_t)(sp + 8) * ss]; }
            if (sp < splits) a0 += p[(size_t)sp * ss];
            acc = a0 + a1;
        }
        red[l][o] = acc;
        __syncthreads();
        if (l == 0 && dst) {
            float r = red[0][o];
#pragma unroll
            for (int q = 1; q < 8; q++) r += red[q][o];
            *dst = r;
        }
        __syncthreads();
    }
}

// bias gradient: column sums of delta[M, cs] (ConvolutionLayer.java:257). Two deterministic stages.
__global__ void __launch_bounds__(256) k_colsum_partial(const float *__restrict__ d, float *__restrict__ partial, int M, int cs) {
    __shared__ float4 red[256];
    const int Q = cs >> 2;
    const int rows_per_blk = (M + gridDim.x - 1) / gridDim.x;
    const int r0 = blockIdx.x * rows_per_blk, r1 = min(M, r0 + rows_per_blk);
    for (int qb = 0; qb < Q; qb += 256) {
        const int qc = min(256, Q - qb), R = 256 / qc;
        const int q = qb + (threadIdx.x % qc), rl = threadIdx.x / qc;
        float4 a = make_float4(0, 0, 0, 0);
        if (rl < R)
            for (int rb = r0 + rl; rb < r1; rb += 4 * R) {  // 4 independent 128-bit loads in flight per thread
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    int r = rb + u * R;
                    v[u] = r < r1 ? *reinterpret_cast<const float4 *>(d + (size_t)r * cs + (size_t)q * 4) : make_float4(0, 0, 0, 0);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) { a.x += v[u].x; a.y += v[u].y; a.z += v[u].z; a.w += v[u].w; }
            }
        red[threadIdx.x] = a;
        __syncthreads();
        if (rl == 0) {
            for (int j = 1; j < R; j++) { float4 b = red[threadIdx.x + j * qc]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
            *reinterpret_cast<float4 *>(partial + (size_t)blockIdx.x * cs + (size_t)q * 4) = a;
        }
        __syncthreads();
    }
}
// blockDim = (32 channels, 32 partial-row groups)
__global__ void k_colsum_final(const float *__restrict__ partial, float *__restrict__ out, int nblk, int cs, int C) {
    __shared__ double sh[32][33];
    int c = blockIdx.x * 32 + threadIdx.x;
    double acc = 0.0;
    if (c < C)
        for (int b = threadIdx.y; b < nblk; b += 32) acc += (double)partial[(size_t)b * cs + c];
    sh[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        double t = 0.0;
        for (int j = 0; j < 32; j++) t += sh[j][threadIdx.x];
        out[c] = (float)t;
    }
}

static inline int grid_for(dl4jb200_ctx *ctx, size_t items) {
    size_t blocks = (items + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    return (int)(blocks > cap ? cap : (blocks ? blocks : 1));
}

static int check_desc(const dl4jb200_conv_desc *d, const char *who) {
    DL4J_REQUIRE(d, "%s: desc is NULL", who);
    DL4J_REQUIRE(d->batch > 0 && d->c_in > 0 && d->in_h > 0 && d->in_w > 0 && d->c_out > 0 && d->k_h > 0 && d->k_w > 0 &&
                     d->stride_h > 0 && d->stride_w > 0 && d->dil_h > 0 && d->dil_w > 0 && d->out_h > 0 && d->out_w > 0,
                 "%s: non-positive dimension in descriptor", who);
    DL4J_REQUIRE((long long)d->batch * d->in_h * d->in_w < (1LL << 31) / 4 && (long long)d->batch * d->out_h * d->out_w < (1LL << 31) / 4,
                 "%s: more than 2^29 pixels per tensor is not supported", who);
    return DL4JB200_OK;
}

static int check_layout(int layout, int cs, int C, const void *p, const char *who, const char *name) {
    DL4J_REQUIRE(layout == DL4JB200_NCHW || layout == DL4JB200_NHWC, "%s: unknown layout for %s", who, name);
    if (layout == DL4JB200_NHWC)
        DL4J_REQUIRE(cs >= C && cs % 4 == 0 && ((uintptr_t)p & 15) == 0, "%s: NHWC %s needs cs %% 4 == 0, cs >= C, 16-byte alignment", who, name);
    return DL4JB200_OK;
}

// Backprop-data of a STRIDED convolution with a kernel larger than 1x1 on the TMA / tcgen05 path. The transposed walk
// iy = (y + pT - kh) / sH is only defined where the division is exact, which im2col-mode TMA cannot express -- the gather
// kernel handled it at ~20 TFLOP/s (AlexNet's 5x5 stride-2 layer: 1.4 ms of a 3.7 ms step). Decomposed by stride phase it
// is sH*sW ordinary unit-stride problems: the input pixels with (y + pT) mod sH == a only ever meet the taps kh = a + sH*j,
// and for them oy = (y + pT - a)/sH - j walks delta with unit stride. Phase (a, b) is therefore the backprop-data of a
// unit-stride convolution with the J_a x J_b sub-filter, padding (c_a, c_b) = ((y0 + pT - a)/sH, ...), run on the compact
// pixel set y = y0 + sH*i and scattered into dx by the output-row remap; the phases tile dx exactly, nothing is zeroed.
// Returns 1 when the shape does not qualify (caller falls back to the gather kernel).
static int bwd_data_phases(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, const float *ds, int delta_cs, float *dxs,
                           int dx_cs, float *wb, size_t wb_floats) {
    const int sH = d->stride_h, sW = d->stride_w;
    if (d->dil_h != 1 || d->dil_w != 1 || d->k_h < sH || d->k_w < sW) return 1;
    const int rows_b = dl4j_round_up(dx_cs > d->c_in ? dx_cs : d->c_in, 256);
    struct Phase { dl4jb200_conv_desc pd; ConvGeom g; int a, b, y0, x0, Ja, Jb; };
    Phase ph[64];
    int np = 0;
    if (sH * sW > 64) return 1;
    for (int a = 0; a < sH; a++)
        for (int b = 0; b < sW; b++) {
            Phase &p = ph[np];
            p.a = a; p.b = b;
            p.y0 = ((a - d->pad_top) % sH + sH) % sH; p.x0 = ((b - d->pad_left) % sW + sW) % sW;
            const int na = p.y0 < d->in_h ? (d->in_h - 1 - p.y0) / sH + 1 : 0, nb = p.x0 < d->in_w ? (d->in_w - 1 - p.x0) / sW + 1 : 0;
            if (na == 0 || nb == 0) continue;
            p.Ja = (d->k_h - a + sH - 1) / sH; p.Jb = (d->k_w - b + sW - 1) / sW;
            p.pd = *d;
            p.pd.in_h = na; p.pd.in_w = nb; p.pd.k_h = p.Ja; p.pd.k_w = p.Jb; p.pd.stride_h = p.pd.stride_w = 1;
            p.pd.pad_top = (p.y0 + d->pad_top - a) / sH; p.pd.pad_left = (p.x0 + d->pad_left - b) / sW;
            p.g = conv_geom_bwd_data(&p.pd, delta_cs);
            if (delta_cs < p.g.Cg || !conv_tc_gemm_can_tma(p.g) || (size_t)rows_b * p.g.Kp > wb_floats) return 1;
            p.g.use_tma = 1;
            p.g.remap = 1; p.g.rm_sy = sH; p.g.rm_sx = sW; p.g.rm_H = d->in_h; p.g.rm_W = d->in_w; p.g.rm_keep = 1;
            np++;
        }
    const int dx_act = DL4JB200_ACT_IDENTITY | (ctx->round_act ? 0x100 : 0);
    for (int i = 0; i < np; i++) {
        Phase &p = ph[i];
        const size_t n = (size_t)rows_b * p.g.Kp;
        size_t blocks = (n + 255) / 256, cap = (size_t)ctx->sm_count * 8;
        k_pack_filter_phase<<<(int)(blocks > cap ? cap : blocks), 256, 0, ctx->stream>>>(w, wb, d->c_out, d->c_in, d->k_h, d->k_w, p.g.Cg, p.g.Kp,
                                                                                      rows_b, p.a, p.b, sH, sW, p.Ja, p.Jb);
        DL4J_CHECK_LAUNCH();
        float *base = dxs + ((size_t)p.y0 * d->in_w + p.x0) * dx_cs;  // pixel (y0, x0) of image 0: the remap strides from there
        int rc = conv_tc_gemm(ctx, p.g, ds, wb, rows_b, nullptr, base, d->c_in, dx_cs, dx_act);
        if (rc) return rc;
    }
    return DL4JB200_OK;
}

extern "C" {

int dl4jb200_bias_grad(dl4jb200_ctx *ctx, const float *delta, int M, int cs, int C, float *db) {
    DL4J_REQUIRE(ctx && delta && db, "bias_grad: NULL argument");
    DL4J_REQUIRE(M > 0 && C > 0 && cs >= C && cs % 4 == 0 && ((uintptr_t)delta & 15) == 0, "bias_grad: bad shape / alignment");
    int nblk = M / 8;
    if (nblk > ctx->sm_count * 4) nblk = ctx->sm_count * 4;
    if (nblk < 1) nblk = 1;
    float *colp = (float *)dl4j_workspace(ctx, 0, (size_t)nblk * cs * sizeof(float));
    if (!colp) return DL4JB200_ERR_CUDA;
    k_colsum_partial<<<nblk, 256, 0, ctx->stream>>>(delta, colp, M, cs);
    DL4J_CHECK_LAUNCH();
    k_colsum_final<<<dl4j_ceil_div(C, 32), dim3(32, 32), 0, ctx->stream>>>(colp, db, nblk, cs, C);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

int dl4jb200_conv_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs,
                      const float *w, const float *bias, float *z, int z_layout, int z_cs, int act, int algo) {
    DL4J_REQUIRE(ctx && x && w && z, "conv_fwd: NULL argument");
    int rc = check_desc(d, "conv_fwd");
    if (rc) return rc;
    if ((rc = check_layout(x_layout, x_cs, d->c_in, x, "conv_fwd", "x"))) return rc;
    if ((rc = check_layout(z_layout, z_cs, d->c_out, z, "conv_fwd", "z"))) return rc;
    DL4J_REQUIRE(algo == DL4JB200_CONV_ALGO_TC_TF32 || algo == DL4JB200_CONV_ALGO_FP32, "conv_fwd: unknown algo %d", algo);
    if (act < 0 || act > DL4JB200_ACT_LEAKYRELU) { dl4j_set_error("conv_fwd: activation %d not handled by the helper", act); return DL4JB200_NOT_SUPPORTED; }

    // stage DL4J-layout tensors into the resident NHWC layout
    const float *xs = x;
    if (x_layout == DL4JB200_NCHW) {
        x_cs = dl4j_round_up(d->c_in, 4);
        float *t = (float *)dl4j_workspace(ctx, 1, (size_t)d->batch * d->in_h * d->in_w * x_cs * sizeof(float));
        if (!t) return DL4JB200_ERR_CUDA;
        if ((rc = dl4jb200_nchw_to_nhwc(ctx, x, t, d->batch, d->c_in, d->in_h, d->in_w, x_cs))) return rc;
        xs = t;
    }
    float *zs = z;
    if (z_layout == DL4JB200_NCHW) {
        z_cs = dl4j_round_up(d->c_out, 4);
        zs = (float *)dl4j_workspace(ctx, 2, (size_t)d->batch * d->out_h * d->out_w * z_cs * sizeof(float));
        if (!zs) return DL4JB200_ERR_CUDA;
    }
    ConvGeom g = conv_geom_fwd(d, x_cs);
    DL4J_REQUIRE(x_cs >= g.Cg, "conv_fwd: x pixel stride %d < %d", x_cs, g.Cg);
    g.use_tma = algo == DL4JB200_CONV_ALGO_TC_TF32 && conv_tc_gemm_can_tma(g);
    const int rows_p = dl4j_round_up(z_cs > d->c_out ? z_cs : d->c_out, 256);
    const int tf = algo == DL4JB200_CONV_ALGO_TC_TF32;
    const float *wp = filters_lookup(ctx, w, false, g.Cg, g.Kp, rows_p, tf, 0);
    if (!wp) {
        float *t = (float *)dl4j_workspace(ctx, 0, (size_t)rows_p * g.Kp * sizeof(float));
        if (!t) return DL4JB200_ERR_CUDA;
        k_pack_filter<<<grid_for(ctx, (size_t)rows_p * g.Kp), 256, 0, ctx->stream>>>(w, t, d->c_out, d->c_in, d->k_h, d->k_w, g.Cg, g.Kp,
                                                                                  rows_p, 0, tf, 0);
        DL4J_CHECK_LAUNCH();
        wp = t;
    }
    // with DL4JB200_OPT_TF32_ACTIVATIONS an activated output (it feeds the next convolution directly) is stored rounded
    const int act_code = act | ((ctx->round_act && act != DL4JB200_ACT_IDENTITY) ? 0x100 : 0);
    if (algo == DL4JB200_CONV_ALGO_TC_TF32) rc = conv_tc_gemm(ctx, g, xs, wp, rows_p, bias, zs, d->c_out, z_cs, act_code);
    else rc = conv_ffma_gemm(ctx, g, xs, wp, bias, zs, d->c_out, z_cs, act);
    if (rc) return rc;
    if (z_layout == DL4JB200_NCHW) return dl4jb200_nhwc_to_nchw(ctx, zs, z, d->batch, d->c_out, d->out_h, d->out_w, z_cs);
    return DL4JB200_OK;
}

int dl4jb200_conv_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs,
                      const float *w, const float *delta, int delta_layout, int delta_cs, float *dw, float *db,
                      float *dx, int dx_layout, int dx_cs, int algo) {
    DL4J_REQUIRE(ctx && x && w && delta, "conv_bwd: NULL argument");
    DL4J_REQUIRE(dw || db || dx, "conv_bwd: nothing to compute (dw, db and dx are all NULL)");
    int rc = check_desc(d, "conv_bwd");
    if (rc) return rc;
    if ((rc = check_layout(x_layout, x_cs, d->c_in, x, "conv_bwd", "x"))) return rc;
    if ((rc = check_layout(delta_layout, delta_cs, d->c_out, delta, "conv_bwd", "delta"))) return rc;
    if (dx && (rc = check_layout(dx_layout, dx_cs, d->c_in, dx, "conv_bwd", "dx"))) return rc;
    DL4J_REQUIRE(algo == DL4JB200_CONV_ALGO_TC_TF32 || algo == DL4JB200_CONV_ALGO_FP32, "conv_bwd: unknown algo %d", algo);
    const bool tcore = algo == DL4JB200_CONV_ALGO_TC_TF32;

    const float *xs = x, *ds = delta;
    if (x_layout == DL4JB200_NCHW) {
        x_cs = dl4j_round_up(d->c_in, 4);
        if (dw) {  // only the filter gradient reads x
            float *t = (float *)dl4j_workspace(ctx, 1, (size_t)d->batch * d->in_h * d->in_w * x_cs * sizeof(float));
            if (!t) return DL4JB200_ERR_CUDA;
            if ((rc = dl4jb200_nchw_to_nhwc(ctx, x, t, d->batch, d->c_in, d->in_h, d->in_w, x_cs))) return rc;
            xs = t;
        }
    }
    if (delta_layout == DL4JB200_NCHW) {
        delta_cs = dl4j_round_up(d->c_out, 4);
        float *t = (float *)dl4j_workspace(ctx, 2, (size_t)d->batch * d->out_h * d->out_w * delta_cs * sizeof(float));
        if (!t) return DL4JB200_ERR_CUDA;
        if ((rc = dl4jb200_nchw_to_nhwc(ctx, delta, t, d->batch, d->c_out, d->out_h, d->out_w, delta_cs))) return rc;
        ds = t;
    }
    float *dxs = dx;
    if (dx && dx_layout == DL4JB200_NCHW) {
        dx_cs = dl4j_round_up(d->c_in, 4);
        dxs = (float *)dl4j_workspace(ctx, 3, (size_t)d->batch * d->in_h * d->in_w * dx_cs * sizeof(float));
        if (!dxs) return DL4JB200_ERR_CUDA;
    }

    // ---- filter gradient: split over pixel ranges, partials in scratch, fixed-order reduction into dw
    ConvGeom gf = conv_geom_fwd(d, x_cs);
    DL4J_REQUIRE(x_cs >= gf.Cg, "conv_bwd: x pixel stride %d < %d", x_cs, gf.Cg);
    gf.use_tma = tcore && conv_tc_gemm_can_tma(gf);
    const int rows_w = dl4j_round_up(d->c_out, 128);
    int splits;
    if (tcore) splits = conv_tc_wgrad_splits(ctx, gf, d->c_out);
    else {
        int tiles = dl4j_ceil_div(gf.Kp, 64) * dl4j_ceil_div(rows_w, 64);
        splits = dl4j_ceil_div(4 * ctx->sm_count, tiles);
        int maxs = dl4j_ceil_div(gf.M, 256);
        if (splits > maxs) splits = maxs;
        if (splits < 1) splits = 1;
    }
    // ---- data gradient operand sizes (computed first so that one scratch request covers everything)
    ConvGeom gb = conv_geom_bwd_data(d, delta_cs);
    gb.use_tma = tcore && dx && conv_tc_gemm_can_tma(gb);
    const int rows_b = dx ? dl4j_round_up(dx_cs > d->c_in ? dx_cs : d->c_in, 256) : 0;
    int nblk = gf.M / 8;  // bias-gradient partial blocks: >= 8 pixel rows each, up to 4 per SM
    if (nblk > ctx->sm_count * 4) nblk = ctx->sm_count * 4;
    if (nblk < 1) nblk = 1;
    // (a dx-only call needs neither the split-K partials nor the bias-gradient rows: do not inflate the scratch for them)
    size_t part_f = dw ? (size_t)splits * rows_w * gf.Kp : 0, pack_f = (size_t)rows_b * gb.Kp, col_f = db ? (size_t)nblk * delta_cs : 0;
    if (dw && db && col_f < (size_t)splits * rows_w) col_f = (size_t)splits * rows_w;  // or the per-split bias-gradient column
    float *ws = (float *)dl4j_workspace(ctx, 0, (part_f + pack_f + col_f) * sizeof(float));
    if (!ws) return DL4JB200_ERR_CUDA;
    float *partial = ws, *wb = ws + part_f, *colp = wb + pack_f;

    // With DL4JB200_OPT_TF32_ACTIVATIONS the caller's delta is TF32-representable, so the bias gradient can ride in the
    // filter-gradient GEMM as an extra accumulator column (delta^T x ones: exact products, fp32 accumulation) instead
    // of a separate column-sum pass over delta.
    const bool db_fused = dw && db && tcore && gf.use_tma && ctx->round_act;
    // One split of a 1x1 filter whose K needs no padding: the GEMM's [c_out][K] output IS dW (and its ones column IS db), so it
    // writes them in place and the reduce / un-pack pass -- a 2 x |W| copy for the dense layers, whose M is the minibatch --
    // is skipped (AlexNet's three dense layers: 0.2 ms of a 2.5 ms step).
    const bool dw_direct = dw && tcore && gf.use_tma && splits == 1 && gf.taps == 1 && gf.Kp == d->c_in && ((uintptr_t)dw & 15) == 0 &&
                           (!db || db_fused);
    if (dw_direct) {
        if ((rc = conv_tc_wgrad(ctx, gf, xs, ds, delta_cs, d->c_out, dw, d->c_out, 1, db_fused ? db : nullptr))) return rc;
    } else if (dw) {
        float *dbp = db_fused ? colp : nullptr;  // [splits][rows_w], carved from the same scratch request below
        if (tcore) rc = conv_tc_wgrad(ctx, gf, xs, ds, delta_cs, d->c_out, partial, rows_w, splits, dbp);
        else rc = conv_ffma_wgrad(ctx, gf, xs, ds, delta_cs, d->c_out, partial, rows_w, splits);
        if (rc) return rc;
        const size_t n_out = (size_t)d->c_out * d->c_in * gf.taps;
        if (splits >= 32) {
            size_t blocks = (n_out + d->c_out + 31) / 32, cap = (size_t)ctx->sm_count * 16;
            DL4J_CUDA_TRY(dl4j_launch_pdl(k_wgrad_finalize_wide, dim3((int)(blocks < cap ? blocks : cap)), dim3(256), 0, ctx->stream,
                                          (const float *)partial, dw, d->c_out, d->c_in, gf.taps, gf.Cg, gf.Kp, rows_w, splits, (const float *)dbp, db));
        } else {
            DL4J_CUDA_TRY(dl4j_launch_pdl(k_wgrad_finalize, dim3(grid_for(ctx, n_out)), dim3(256), 0, ctx->stream, (const float *)partial, dw,
                                          d->c_out, d->c_in, gf.taps, gf.Cg, gf.Kp, rows_w, splits, (const float *)dbp, db));
        }
        DL4J_CHECK_LAUNCH();
    }

    if (db && !db_fused) {
        k_colsum_partial<<<nblk, 256, 0, ctx->stream>>>(ds, colp, gf.M, delta_cs);
        DL4J_CHECK_LAUNCH();
        k_colsum_final<<<dl4j_ceil_div(d->c_out, 32), dim3(32, 32), 0, ctx->stream>>>(colp, db, nblk, delta_cs, d->c_out);
        DL4J_CHECK_LAUNCH();
    }

    // the epsilon handed to the previous layer is an operand of that layer's GEMMs: with DL4JB200_OPT_TF32_ACTIVATIONS it is
    // stored rounded to nearest TF32 like every other producer's output (the tensor core would otherwise truncate it)
    const int dx_act = DL4JB200_ACT_IDENTITY | (ctx->round_act ? 0x100 : 0);
    if (dx && tcore && d->k_h == 1 && d->k_w == 1 && (d->stride_h > 1 || d->stride_w > 1) && d->pad_top == 0 && d->pad_left == 0) {
        // strided 1x1: dx is non-zero only at the sampled pixels. Run the GEMM on the compact pixel set (a plain
        // unit-stride 1x1 walk over delta, TMA-able) and scatter its rows; everything else is zero.
        ConvGeom gc = gb;
        gc.transposed = 0; gc.Hm = d->out_h; gc.Wm = d->out_w; gc.sH = gc.sW = 1;
        geom_finish(gc);
        gc.remap = 1; gc.rm_sy = d->stride_h; gc.rm_sx = d->stride_w; gc.rm_H = d->in_h; gc.rm_W = d->in_w;
        gc.use_tma = conv_tc_gemm_can_tma(gc);
        if (gc.use_tma) {
            // (dx is zero away from the sampled pixels: the GEMM's epilogue writes those zeros itself, or clears dx first)
            const float *wq = filters_lookup(ctx, w, true, gc.Cg, gc.Kp, rows_b, 1, 0);
            if (!wq) {
                k_pack_filter<<<grid_for(ctx, pack_f), 256, 0, ctx->stream>>>(w, wb, d->c_out, d->c_in, 1, 1, gc.Cg, gc.Kp, rows_b, 1, 1, 0);
                DL4J_CHECK_LAUNCH();
                wq = wb;
            }
            if ((rc = conv_tc_gemm(ctx, gc, ds, wq, rows_b, nullptr, dxs, d->c_in, dx_cs, dx_act))) return rc;
            if (dx_layout == DL4JB200_NCHW) return dl4jb200_nhwc_to_nchw(ctx, dxs, dx, d->batch, d->c_in, d->in_h, d->in_w, dx_cs);
            return DL4JB200_OK;
        }
    }
    if (dx && tcore && (d->stride_h > 1 || d->stride_w > 1) && (d->k_h > 1 || d->k_w > 1)) {
        rc = bwd_data_phases(ctx, d, w, ds, delta_cs, dxs, dx_cs, wb, pack_f);
        if (rc < 0) return rc;
        if (rc == DL4JB200_OK) {
            if (dx_layout == DL4JB200_NCHW) return dl4jb200_nhwc_to_nchw(ctx, dxs, dx, d->batch, d->c_in, d->in_h, d->in_w, dx_cs);
            return DL4JB200_OK;
        }
    }
    if (dx) {
        DL4J_REQUIRE(delta_cs >= gb.Cg, "conv_bwd: delta pixel stride %d < %d", delta_cs, gb.Cg);
        const float *wq = filters_lookup(ctx, w, true, gb.Cg, gb.Kp, rows_b, tcore, gb.use_tma);
        if (!wq) {
            k_pack_filter<<<grid_for(ctx, pack_f), 256, 0, ctx->stream>>>(w, wb, d->c_out, d->c_in, d->k_h, d->k_w, gb.Cg, gb.Kp, rows_b, 1, tcore, gb.use_tma);
            DL4J_CHECK_LAUNCH();
            wq = wb;
        }
        if (tcore) rc = conv_tc_gemm(ctx, gb, ds, wq, rows_b, nullptr, dxs, d->c_in, dx_cs, dx_act);
        else rc = conv_ffma_gemm(ctx, gb, ds, wq, nullptr, dxs, d->c_in, dx_cs, DL4JB200_ACT_IDENTITY);
        if (rc) return rc;
        if (dx_layout == DL4JB200_NCHW) return dl4jb200_nhwc_to_nchw(ctx, dxs, dx, d->batch, d->c_in, d->in_h, d->in_w, dx_cs);
    }
    return DL4JB200_OK;
}

// Forward convolution (identity activation) that also leaves the batch-norm statistics of its output, taken in the GEMM
// epilogue while the tile is on chip: the following BatchNormalization then skips its own partial-sum pass over z.
static int conv_fwd_stats_geom(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int x_cs, int z_cs, int algo, ConvGeom &g, int &rows_p) {
    g = conv_geom_fwd(d, x_cs);
    g.use_tma = algo == DL4JB200_CONV_ALGO_TC_TF32 && x_cs >= g.Cg && conv_tc_gemm_can_tma(g);
    rows_p = dl4j_round_up(z_cs > d->c_out ? z_cs : d->c_out, 256);
    return conv_tc_stat_rows(ctx, g, rows_p, d->c_out, z_cs);
}

int dl4jb200_conv_fwd_bnstats_rows(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int x_cs, int z_cs, int algo, int *rows_out) {
    DL4J_REQUIRE(ctx && rows_out, "conv_fwd_bnstats_rows: NULL argument");
    int rc = check_desc(d, "conv_fwd_bnstats_rows");
    if (rc) return rc;
    ConvGeom g;
    int rows_p;
    *rows_out = (x_cs % 4 == 0 && z_cs % 4 == 0) ? conv_fwd_stats_geom(ctx, d, x_cs, z_cs, algo, g, rows_p) : 0;
    return DL4JB200_OK;
}

int dl4jb200_conv_fwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_cs, const float *w,
                              const float *bias, float *z, int z_cs, int algo, const float *pivot, float *stat_partial) {
    DL4J_REQUIRE(ctx && x && w && z && pivot && stat_partial, "conv_fwd_bnstats: NULL argument");
    int rc = check_desc(d, "conv_fwd_bnstats");
    if (rc) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, x_cs, d->c_in, x, "conv_fwd_bnstats", "x"))) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, z_cs, d->c_out, z, "conv_fwd_bnstats", "z"))) return rc;
    DL4J_REQUIRE(((uintptr_t)pivot & 15) == 0 && ((uintptr_t)stat_partial & 15) == 0 && (!bias || ((uintptr_t)bias & 15) == 0),
                 "conv_fwd_bnstats: pivot, stat_partial and bias must be 16-byte aligned");
    ConvGeom g;
    int rows_p;
    if (conv_fwd_stats_geom(ctx, d, x_cs, z_cs, algo, g, rows_p) <= 0) {
        dl4j_set_error("conv_fwd_bnstats: this convolution cannot carry the statistics (see conv_fwd_bnstats_rows)");
        return DL4JB200_NOT_SUPPORTED;
    }
    const float *wp = filters_lookup(ctx, w, false, g.Cg, g.Kp, rows_p, 1, 0);
    if (!wp) {
        float *t = (float *)dl4j_workspace(ctx, 0, (size_t)rows_p * g.Kp * sizeof(float));
        if (!t) return DL4JB200_ERR_CUDA;
        k_pack_filter<<<grid_for(ctx, (size_t)rows_p * g.Kp), 256, 0, ctx->stream>>>(w, t, d->c_out, d->c_in, d->k_h, d->k_w, g.Cg, g.Kp,
                                                                                  rows_p, 0, 1, 0);
        DL4J_CHECK_LAUNCH();
        wp = t;
    }
    g.stat_out = stat_partial;
    g.stat_pivot = pivot;
    return conv_tc_gemm(ctx, g, x, wp, rows_p, bias, z, d->c_out, z_cs, DL4JB200_ACT_IDENTITY);
}

int dl4jb200_filter_register(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, int x_cs, int z_cs, int algo,
                             int with_bwd_data) {
    DL4J_REQUIRE(ctx && w, "filter_register: NULL argument");
    int rc = check_desc(d, "filter_register");
    if (rc) return rc;
    DL4J_REQUIRE(algo == DL4JB200_CONV_ALGO_TC_TF32 || algo == DL4JB200_CONV_ALGO_FP32, "filter_register: unknown algo %d", algo);
    DL4J_REQUIRE(x_cs >= d->c_in && z_cs >= d->c_out && x_cs % 4 == 0 && z_cs % 4 == 0, "filter_register: bad pixel strides");
    if (!ctx->filters) ctx->filters = new FilterRegistry();
    FilterRegistry *r = ctx->filters;
    DL4J_REQUIRE(r->entries.find(w) == r->entries.end(), "filter_register: this filter is already registered");
    const int tf = algo == DL4JB200_CONV_ALGO_TC_TF32;
    FilterEntry e{};
    ConvGeom gf = conv_geom_fwd(d, x_cs);
    e.fwd = PackJob{w, nullptr, 0ULL, d->c_out, d->c_in, d->k_h, d->k_w, gf.Cg, gf.Kp, dl4j_round_up(z_cs > d->c_out ? z_cs : d->c_out, 256), 0, tf, 0};
    pack_job_divs(e.fwd);
    DL4J_CUDA_TRY(cudaMalloc(&e.fwd.wp, (size_t)e.fwd.rows_p * e.fwd.Kp * sizeof(float)));
    DL4J_CUDA_TRY(cudaMemsetAsync(e.fwd.wp, 0, (size_t)e.fwd.rows_p * e.fwd.Kp * sizeof(float), ctx->stream));  // padding stays zero
    if (with_bwd_data) {
        ConvGeom gb = conv_geom_bwd_data(d, z_cs);
        const int flip = tf && conv_tc_gemm_can_tma(gb);
        e.bwd = PackJob{w, nullptr, 0ULL, d->c_out, d->c_in, d->k_h, d->k_w, gb.Cg, gb.Kp, dl4j_round_up(x_cs > d->c_in ? x_cs : d->c_in, 256), 1, tf, flip};
        pack_job_divs(e.bwd);
        DL4J_CUDA_TRY(cudaMalloc(&e.bwd.wp, (size_t)e.bwd.rows_p * e.bwd.Kp * sizeof(float)));
        DL4J_CUDA_TRY(cudaMemsetAsync(e.bwd.wp, 0, (size_t)e.bwd.rows_p * e.bwd.Kp * sizeof(float), ctx->stream));
    }
    r->entries[w] = e;
    r->table_dirty = true;
    r->valid = false;  // nothing is packed until the next dl4jb200_filters_repack
    return DL4JB200_OK;
}

int dl4jb200_filters_repack(dl4jb200_ctx *ctx) {
    DL4J_REQUIRE(ctx, "filters_repack: ctx is NULL");
    FilterRegistry *r = ctx->filters;
    if (!r || r->entries.empty()) return DL4JB200_OK;
    if (r->table_dirty) {
        std::vector<PackJob> jobs;
        std::vector<PackUnit> units;
        for (auto &kv : r->entries)
            for (PackJob *j : {&kv.second.fwd, &kv.second.bwd}) {
                if (!j->wp) continue;
                const int ji = (int)jobs.size(), taps = j->kH * j->kW;
                j->start = (unsigned long long)units.size();
                jobs.push_back(*j);
                if (!j->transposed) {
                    long long row = (long long)j->Cin * taps;
                    int rows_per = (int)(8192 / row);
                    if (rows_per < 1) rows_per = 1;
                    for (int n0 = 0; n0 < j->Cout; n0 += rows_per)
                        units.push_back(PackUnit{ji, n0, j->Cout - n0 < rows_per ? j->Cout - n0 : rows_per, 0});
                } else {
                    int subs = 8192 / (32 * 32 * taps);  // ~8 K elements per block
                    if (subs < 1) subs = 1;
                    for (int co0 = 0; co0 < j->Cout; co0 += 32)
                        for (int ci0 = 0; ci0 < j->Cin; ci0 += 32 * subs) units.push_back(PackUnit{ji, co0, ci0, subs});
                }
            }
        DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        if (r->jobs_dev) { cudaFree(r->jobs_dev); r->jobs_dev = nullptr; }
        if (r->units_dev) { cudaFree(r->units_dev); r->units_dev = nullptr; }
        DL4J_CUDA_TRY(cudaMalloc(&r->jobs_dev, jobs.size() * sizeof(PackJob)));
        DL4J_CUDA_TRY(cudaMalloc(&r->units_dev, units.size() * sizeof(PackUnit)));
        DL4J_CUDA_TRY(cudaMemcpyAsync(r->jobs_dev, jobs.data(), jobs.size() * sizeof(PackJob), cudaMemcpyHostToDevice, ctx->stream));
        DL4J_CUDA_TRY(cudaMemcpyAsync(r->units_dev, units.data(), units.size() * sizeof(PackUnit), cudaMemcpyHostToDevice, ctx->stream));
        DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));  // the tables are stack vectors
        r->njobs = (int)jobs.size();
        r->nunits = (int)units.size();
        r->table_dirty = false;
        static bool configured[16] = {false};
        if (!configured[ctx->device & 15]) {
            DL4J_CUDA_TRY(cudaFuncSetAttribute(k_pack_filters_batched, cudaFuncAttributeMaxDynamicSharedMemorySize, PACK_SMEM_FLOATS * 4));
            configured[ctx->device & 15] = true;
        }
    }
    k_pack_filters_batched<<<r->nunits, 256, PACK_SMEM_FLOATS * 4, ctx->stream>>>(r->jobs_dev, r->units_dev);
    DL4J_CHECK_LAUNCH();
    r->valid = true;
    return DL4JB200_OK;
}

int dl4jb200_filters_invalidate(dl4jb200_ctx *ctx) {
    DL4J_REQUIRE(ctx, "filters_invalidate: ctx is NULL");
    if (ctx->filters) ctx->filters->valid = false;
    return DL4JB200_OK;
}

// Backprop-data with the consumer-side bookkeeping of the vertex that owns dx folded into the GEMM epilogue:
//   dx = relu'(act_out) * (conv_bwd_data(delta) + eps_sum)
// (ComputationGraph.java:2044-2063 epsilon fan-in, then ActivationLayer.java:68-92 / the Add vertex' ReLU backprop).
int dl4jb200_conv_bwd_data_fused(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, const float *delta, int delta_cs,
                                 float *dx, int dx_cs, const float *eps_sum, const float *act_out, int act, int algo) {
    DL4J_REQUIRE(ctx && w && delta && dx, "conv_bwd_data_fused: NULL argument");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act == DL4JB200_ACT_RELU, "conv_bwd_data_fused: only identity / relu masks");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act_out, "conv_bwd_data_fused: relu needs act_out");
    int rc = check_desc(d, "conv_bwd_data_fused");
    if (rc) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, delta_cs, d->c_out, delta, "conv_bwd_data_fused", "delta"))) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, dx_cs, d->c_in, dx, "conv_bwd_data_fused", "dx"))) return rc;
    DL4J_REQUIRE(algo == DL4JB200_CONV_ALGO_TC_TF32 || algo == DL4JB200_CONV_ALGO_FP32, "conv_bwd_data_fused: unknown algo %d", algo);
    const float *mask = act == DL4JB200_ACT_RELU ? act_out : nullptr;
    const size_t n = (size_t)d->batch * d->in_h * d->in_w * dx_cs;
    const bool strided_1x1 = d->k_h == 1 && d->k_w == 1 && (d->stride_h > 1 || d->stride_w > 1);
    ConvGeom gb = conv_geom_bwd_data(d, delta_cs);
    gb.use_tma = algo == DL4JB200_CONV_ALGO_TC_TF32 && conv_tc_gemm_can_tma(gb);
    if (!gb.use_tma || strided_1x1 || (!eps_sum && !mask)) {
        // no fused epilogue on this path: plain backprop-data into scratch, then one fused elementwise pass
        float *tmp = dx;
        if (eps_sum || mask) {
            tmp = (float *)dl4j_workspace(ctx, 3, n * sizeof(float));
            if (!tmp) return DL4JB200_ERR_CUDA;
        }
        // (x is not read when dw == db == NULL; pass delta to satisfy the argument check)
        if ((rc = dl4jb200_conv_bwd(ctx, d, dx, DL4JB200_NHWC, dx_cs, w, delta, DL4JB200_NHWC, delta_cs, nullptr, nullptr, tmp,
                                    DL4JB200_NHWC, dx_cs, algo)))
            return rc;
        if (tmp == dx) return DL4JB200_OK;
        if (eps_sum) return dl4jb200_eps_sum_act_bwd(ctx, tmp, eps_sum, mask ? mask : tmp, dx, n, mask ? DL4JB200_ACT_RELU : DL4JB200_ACT_IDENTITY);
        return dl4jb200_act_bwd(ctx, mask, tmp, dx, n, DL4JB200_ACT_RELU);
    }
    DL4J_REQUIRE(delta_cs >= gb.Cg, "conv_bwd_data_fused: delta pixel stride %d < %d", delta_cs, gb.Cg);
    const int rows_b = dl4j_round_up(dx_cs > d->c_in ? dx_cs : d->c_in, 256);
    const size_t pack_f = (size_t)rows_b * gb.Kp;
    const float *wq = filters_lookup(ctx, w, true, gb.Cg, gb.Kp, rows_b, 1, 1);
    if (!wq) {
        float *wb = (float *)dl4j_workspace(ctx, 0, pack_f * sizeof(float));
        if (!wb) return DL4JB200_ERR_CUDA;
        k_pack_filter<<<grid_for(ctx, pack_f), 256, 0, ctx->stream>>>(w, wb, d->c_out, d->c_in, d->k_h, d->k_w, gb.Cg, gb.Kp, rows_b, 1, 1, 1);
        DL4J_CHECK_LAUNCH();
        wq = wb;
    }
    gb.epi_add = eps_sum;
    gb.epi_mask = mask;
    return conv_tc_gemm(ctx, gb, delta, wq, rows_b, nullptr, dx, d->c_in, dx_cs, DL4JB200_ACT_IDENTITY | (ctx->round_act ? 0x100 : 0));
}

// Backprop-data (optionally with the fused fan-in sum / ReLU backprop of conv_bwd_data_fused) whose epilogue also takes the
// backward sums of the BatchNormalization that owns dx -- directly (dx is that layer's epsilon) or through the Add vertex
// the layer is folded into. The BatchNormalization's backward then needs no pass of its own over (x, eps) for the sums.
static int bwd_bnstats_geom(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int delta_cs, int dx_cs, int algo, ConvGeom &gb, int &rows_b) {
    gb = conv_geom_bwd_data(d, delta_cs);
    const bool strided_1x1 = d->k_h == 1 && d->k_w == 1 && (d->stride_h > 1 || d->stride_w > 1);
    gb.use_tma = algo == DL4JB200_CONV_ALGO_TC_TF32 && !strided_1x1 && delta_cs >= gb.Cg && conv_tc_gemm_can_tma(gb);
    rows_b = dl4j_round_up(dx_cs > d->c_in ? dx_cs : d->c_in, 256);
    return conv_tc_bst_rows(ctx, gb, rows_b, d->c_in, dx_cs);
}

int dl4jb200_conv_bwd_data_bnstats_rows(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int delta_cs, int dx_cs, int algo, int *rows_out) {
    DL4J_REQUIRE(ctx && rows_out, "conv_bwd_data_bnstats_rows: NULL argument");
    int rc = check_desc(d, "conv_bwd_data_bnstats_rows");
    if (rc) return rc;
    ConvGeom g;
    int rows_b;
    *rows_out = (delta_cs % 4 == 0 && dx_cs % 4 == 0) ? bwd_bnstats_geom(ctx, d, delta_cs, dx_cs, algo, g, rows_b) : 0;
    return DL4JB200_OK;
}

int dl4jb200_conv_bwd_data_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, const float *delta, int delta_cs,
                                   float *dx, int dx_cs, const float *eps_sum, const float *act_out, int act, int algo,
                                   const float *bn_x, const float *bn_gamma, const float *bn_beta, const float *bn_mean,
                                   const float *bn_std, int bn_relu, float *stat_partial) {
    DL4J_REQUIRE(ctx && w && delta && dx && bn_x && bn_mean && bn_std && stat_partial, "conv_bwd_data_bnstats: NULL argument");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act == DL4JB200_ACT_RELU, "conv_bwd_data_bnstats: only identity / relu masks");
    DL4J_REQUIRE(act == DL4JB200_ACT_IDENTITY || act_out, "conv_bwd_data_bnstats: relu needs act_out");
    int rc = check_desc(d, "conv_bwd_data_bnstats");
    if (rc) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, delta_cs, d->c_out, delta, "conv_bwd_data_bnstats", "delta"))) return rc;
    if ((rc = check_layout(DL4JB200_NHWC, dx_cs, d->c_in, dx, "conv_bwd_data_bnstats", "dx"))) return rc;
    auto al = [](const void *p) { return !p || ((uintptr_t)p & 15) == 0; };
    DL4J_REQUIRE(al(bn_x) && al(bn_gamma) && al(bn_beta) && al(bn_mean) && al(bn_std) && al(stat_partial) && al(eps_sum) && al(act_out),
                 "conv_bwd_data_bnstats: every tensor / per-channel array must be 16-byte aligned");
    ConvGeom gb;
    int rows_b;
    if (bwd_bnstats_geom(ctx, d, delta_cs, dx_cs, algo, gb, rows_b) <= 0) {
        dl4j_set_error("conv_bwd_data_bnstats: this backprop-data cannot carry the sums (see conv_bwd_data_bnstats_rows)");
        return DL4JB200_NOT_SUPPORTED;
    }
    const size_t pack_f = (size_t)rows_b * gb.Kp;
    const float *wq = filters_lookup(ctx, w, true, gb.Cg, gb.Kp, rows_b, 1, 1);
    if (!wq) {
        float *wb = (float *)dl4j_workspace(ctx, 0, pack_f * sizeof(float));
        if (!wb) return DL4JB200_ERR_CUDA;
        k_pack_filter<<<grid_for(ctx, pack_f), 256, 0, ctx->stream>>>(w, wb, d->c_out, d->c_in, d->k_h, d->k_w, gb.Cg, gb.Kp, rows_b, 1, 1, 1);
        DL4J_CHECK_LAUNCH();
        wq = wb;
    }
    gb.epi_add = eps_sum;
    gb.epi_mask = act == DL4JB200_ACT_RELU ? act_out : nullptr;
    gb.stat_out = stat_partial;
    gb.bst_x = bn_x; gb.bst_mean = bn_mean; gb.bst_std = bn_std; gb.bst_gamma = bn_gamma; gb.bst_beta = bn_beta; gb.bst_relu = bn_relu ? 1 : 0;
    return conv_tc_gemm(ctx, gb, delta, wq, rows_b, nullptr, dx, d->c_in, dx_cs, DL4JB200_ACT_IDENTITY | (ctx->round_act ? 0x100 : 0));
}

}  // extern "C"
