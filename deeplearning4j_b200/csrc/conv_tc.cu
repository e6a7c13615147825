// TF32 tensor-core implicit-GEMM convolution for sm_100a: tcgen05.mma with the accumulator in TMEM,
// operands staged in 128-byte-swizzled shared memory through a multi-stage mbarrier pipeline.
//
// Warp roles per CTA (160 threads): warps 0-3 gather operand tiles (16-byte cp.async with zero-fill for
// padding / stride holes / ragged edges -- the im2col matrix is never materialised) and later run the
// epilogue (tcgen05.ld -> bias + activation -> global); warp 4 allocates TMEM and its lane 0 issues the MMAs.
// Two CTAs fit per SM (<= 100 KB smem, <= 128 TMEM columns each) so one CTA's epilogue overlaps the
// other's main loop.
//
//   k_conv_tc_gemm : forward / backprop-data.  D[128 pixels, BN channels] += A[128, 32] * B[BN, 32]^T,
//                    both operands K-major (a row = 32 consecutive K values = 128 bytes).
//   k_conv_tc_wgrad: backprop-filter. D[128 co, BN k] += delta[32 px, 128 co]^T * im2col[32 px, BN k]:
//                    the reduction runs over pixels, so both operands are MN-major (a row = 32 consecutive
//                    M/N values of one pixel); split over pixel ranges, partials reduced by conv.cu.
#include "conv_common.cuh"
#include "conv_tma.cuh"

using tcgeom::TmaGemmGeom;
using tcgeom::TmaWgradGeom;

namespace tc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 32;  // fp32 elements: one 128-byte swizzle row
constexpr int PRODUCERS = 128;
constexpr int THREADS = 160;
constexpr int LOOKAHEAD = 2;  // k-chunks in flight per producer thread beyond the one being published

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must surface as a CUDA error, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    long long t0 = clock64();
    uint32_t spins = 0;
    while (!mbar_try(bar, parity)) {
        if ((++spins & 1023u) == 0 && clock64() - t0 > 6000000000LL) {
            printf("dl4jb200: mbarrier wait timed out (block %d,%d thread %d)\n", blockIdx.x, blockIdx.y, threadIdx.x);
            __trap();
        }
    }
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "n"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
// ---- CTA pairs (cluster of two CTAs on one TPC, tcgen05 cta_group::2): one UMMA of M = 256 spans both SMs -- each CTA holds
// its own 128 rows of A and its own 128 accumulator lanes, and only HALF of the B tile (the instruction reads the other
// half out of the partner's shared memory), so a K-step brings 16 KB + BN * 64 B into each SM instead of 16 KB + BN * 128 B.
// The even CTA (rank 0) issues the MMAs; both CTAs' TMA loads complete on ITS full barrier; its commits are multicast to
// both CTAs' barriers.
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_alloc2(uint32_t dst_smem) {  // the same warp of BOTH CTAs, same dst offset
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "n"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void umma_commit2(uint32_t bar) {  // arrives on `bar` of both CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void umma_tf32_2(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives row (lane base + i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

// UMMA shared-memory descriptor, 128-byte swizzle (cute/arch/mma_sm100_desc.hpp SmemDescriptor):
// [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout=2 (SWIZZLE_128B)
// layout: 2 = SWIZZLE_128B (16-byte chunks XOR row%8; K-major operands), 1 = SWIZZLE_128B_BASE32B (32-byte chunks
// XOR row%4, 4-row atoms) -- the only swizzle the hardware offers for MN-major 32-bit operands
// (cutlass/gemm/collective/builders/sm100_common.inl:89-93).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 2) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)layout << 61);
}
// byte offset of 16-byte piece `piece` (0..7) of pixel row r inside an MN-major tile whose MN extent is NATOMS
// 32-float atoms: atoms are 4 rows x 128 B = 512 B, ordered [row group r/4][atom]; within a row the 32-byte
// chunk index is XORed with r%4.
template <int NATOMS>
__device__ __forceinline__ uint32_t mn_tile_off(int r, int atom, int piece) {
    return (uint32_t)(((r >> 2) * NATOMS + atom) << 9) + ((uint32_t)(r & 3) << 7) +
           ((uint32_t)(((piece >> 1) ^ (r & 3))) << 5) + ((uint32_t)(piece & 1) << 4);
}
// Epilogue store of one warp's 32 rows x 32 fp32 columns (thread = row, v = its 32 columns) through a 4 KB
// shared-memory staging tile so that every global store instruction writes four complete 128-byte row segments
// (lane = 16-byte chunk c of row 4i + lane/8) instead of 32 scattered 16-byte pieces. 16-byte chunks are
// XOR-swizzled with the row so that both the row-wise writes and the transposed reads are bank-conflict free.
__device__ __forceinline__ void store_tile_32x32(uint32_t stage_smem, const float (&v)[32], int lane, float *out_base,
                                                 size_t row_stride, int rows_valid) {
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; q++) {
        const uint32_t a = stage_smem + (uint32_t)lane * 128u + (uint32_t)((q ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]),
                     "f"(v[q * 4 + 3]) : "memory");
    }
    __syncwarp();
    const int c = lane & 7;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int row = 4 * i + (lane >> 3);
        float4 t;
        const uint32_t a = stage_smem + (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4);
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(a));
        if (row < rows_valid) *reinterpret_cast<float4 *>(out_base + (size_t)row * row_stride + c * 4) = t;
    }
}

// Same as store_tile_32x32 with the fused backprop epilogue: t = (t + add) * (mask > 0), optionally rounded to TF32.
__device__ __forceinline__ void store_tile_32x32_fused(uint32_t stage_smem, const float (&v)[32], int lane, size_t off_base,
                                                       size_t row_stride, int rows_valid, float *out,
                                                       const float *__restrict__ add, const float *__restrict__ mask, int rnd) {
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; q++) {
        const uint32_t a = stage_smem + (uint32_t)lane * 128u + (uint32_t)((q ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]),
                     "f"(v[q * 4 + 3]) : "memory");
    }
    __syncwarp();
    const int c = lane & 7;
    float4 ad[8], mk[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {  // issue all global reads first: 16 independent 128-bit loads in flight per lane
        const int row = 4 * i + (lane >> 3);
        const size_t o = off_base + (size_t)row * row_stride + c * 4;
        ad[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        mk[i] = make_float4(1.f, 1.f, 1.f, 1.f);
        if (row < rows_valid) {
            if (add) ad[i] = *reinterpret_cast<const float4 *>(add + o);
            if (mask) mk[i] = __ldg(reinterpret_cast<const float4 *>(mask + o));
        }
    }
    // keep the consumers (mask compares) behind ALL sixteen loads: folded next to its own load, each compare would
    // stall the in-order issue for a full DRAM latency before the next load leaves the SM
#pragma unroll
    for (int i = 0; i < 8; i++) {
        asm volatile("" : "+f"(mk[i].x), "+f"(mk[i].y), "+f"(mk[i].z), "+f"(mk[i].w));
        asm volatile("" : "+f"(ad[i].x), "+f"(ad[i].y), "+f"(ad[i].z), "+f"(ad[i].w));
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int row = 4 * i + (lane >> 3);
        float4 t;
        const uint32_t a = stage_smem + (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4);
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(a));
        t.x = mk[i].x > 0.f ? t.x + ad[i].x : 0.f; t.y = mk[i].y > 0.f ? t.y + ad[i].y : 0.f;
        t.z = mk[i].z > 0.f ? t.z + ad[i].z : 0.f; t.w = mk[i].w > 0.f ? t.w + ad[i].w : 0.f;
        if (row < rows_valid) *reinterpret_cast<float4 *>(out + off_base + (size_t)row * row_stride + c * 4) = dl4j_rnd4(t, rnd);
    }
}

// store_tile_32x32 that also accumulates, per lane, sum(t - K) and sum((t - K)^2) of the four columns it stores, over the
// valid rows: the batch-norm statistics of the convolution's output taken while the tile is still on chip.
__device__ __forceinline__ void store_tile_32x32_stats(uint32_t stage_smem, const float (&v)[32], int lane, float *out_base,
                                                       size_t row_stride, int rows_valid, const float4 &K, float4 &s, float4 &q) {
    __syncwarp();
#pragma unroll
    for (int p = 0; p < 8; p++) {
        const uint32_t a = stage_smem + (uint32_t)lane * 128u + (uint32_t)((p ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v[p * 4]), "f"(v[p * 4 + 1]), "f"(v[p * 4 + 2]),
                     "f"(v[p * 4 + 3]) : "memory");
    }
    __syncwarp();
    const int c = lane & 7;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int row = 4 * i + (lane >> 3);
        float4 t;
        const uint32_t a = stage_smem + (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4);
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(a));
        if (row < rows_valid) {
            *reinterpret_cast<float4 *>(out_base + (size_t)row * row_stride + c * 4) = t;
            const float dx = t.x - K.x, dy = t.y - K.y, dz = t.z - K.z, dw = t.w - K.w;
            s.x += dx; s.y += dy; s.z += dz; s.w += dw;
            q.x += dx * dx; q.y += dy * dy; q.z += dz * dz; q.w += dw * dw;
        }
    }
}

// Backprop-data epilogue that also takes the BACKWARD batch-norm sums of the layer that owns `out` (TmaGemmGeom::bst_*):
// t = (mask > 0 ? t + add : 0) when the fused fan-in / ReLU-backprop operands are given, stored (rounded to TF32 if asked),
// then with x = bx[...]: xm = x - mean, e' = (relu && !(xm*a + beta > 0)) ? 0 : t, s1 += e', s2 += e' * xm for the lane's
// four columns. Rows are handled in two batches of four so that at most twelve 128-bit loads are in flight per lane.
__device__ __forceinline__ void store_tile_32x32_bst(uint32_t stage_smem, const float (&v)[32], int lane, size_t off_base,
                                                     size_t row_stride, int rows_valid, float *out,
                                                     const float *__restrict__ add, const float *__restrict__ mask, int rnd,
                                                     const float *__restrict__ bx, const float4 &mean, const float4 &a,
                                                     const float4 &beta, int relu, float4 &s1, float4 &s2) {
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; q++) {
        const uint32_t sa = stage_smem + (uint32_t)lane * 128u + (uint32_t)((q ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(sa), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]),
                     "f"(v[q * 4 + 3]) : "memory");
    }
    __syncwarp();
    const int c = lane & 7;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        float4 ad[4], mk[4], xs[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int row = 4 * (half * 4 + i) + (lane >> 3);
            const size_t o = off_base + (size_t)row * row_stride + c * 4;
            ad[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            mk[i] = make_float4(1.f, 1.f, 1.f, 1.f);
            xs[i] = ad[i];
            if (row < rows_valid) {
                xs[i] = __ldg(reinterpret_cast<const float4 *>(bx + o));
                if (add) ad[i] = *reinterpret_cast<const float4 *>(add + o);
                if (mask) mk[i] = __ldg(reinterpret_cast<const float4 *>(mask + o));
            }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {  // keep every consumer behind all the loads of the batch (see store_tile_32x32_fused)
            asm volatile("" : "+f"(mk[i].x), "+f"(mk[i].y), "+f"(mk[i].z), "+f"(mk[i].w));
            asm volatile("" : "+f"(ad[i].x), "+f"(ad[i].y), "+f"(ad[i].z), "+f"(ad[i].w));
            asm volatile("" : "+f"(xs[i].x), "+f"(xs[i].y), "+f"(xs[i].z), "+f"(xs[i].w));
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int row = 4 * (half * 4 + i) + (lane >> 3);
            float4 t;
            const uint32_t sa = stage_smem + (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4);
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(sa));
            if (add || mask) {
                t.x = mk[i].x > 0.f ? t.x + ad[i].x : 0.f; t.y = mk[i].y > 0.f ? t.y + ad[i].y : 0.f;
                t.z = mk[i].z > 0.f ? t.z + ad[i].z : 0.f; t.w = mk[i].w > 0.f ? t.w + ad[i].w : 0.f;
            }
            t = dl4j_rnd4(t, rnd);  // the sums are taken of what is stored: what a separate pass would have read back
            if (row < rows_valid) {
                *reinterpret_cast<float4 *>(out + off_base + (size_t)row * row_stride + c * 4) = t;
                float4 xm = xs[i];
                xm.x -= mean.x; xm.y -= mean.y; xm.z -= mean.z; xm.w -= mean.w;
                if (relu) {
                    t.x = xm.x * a.x + beta.x > 0.f ? t.x : 0.f; t.y = xm.y * a.y + beta.y > 0.f ? t.y : 0.f;
                    t.z = xm.z * a.z + beta.z > 0.f ? t.z : 0.f; t.w = xm.w * a.w + beta.w > 0.f ? t.w : 0.f;
                }
                s1.x += t.x; s1.y += t.y; s1.z += t.z; s1.w += t.w;
                s2.x += t.x * xm.x; s2.y += t.y * xm.y; s2.z += t.z * xm.z; s2.w += t.w * xm.w;
            }
        }
    }
}

// Same, but every row has its own destination (element offset from `out`, or < 0 to skip): output-row scatter.
// zmask (per row, 0 when the caller pre-zeroed the tensor): the scattered rows are the stride-2 sample points of a larger
// image whose other pixels are zero; bit 0 / 1 say that the pixel to the right / below exists, and the lane then also
// writes zeros to the up-to-three hole pixels of its 2x2 cell, so that no separate memset pass over the tensor is needed.
__device__ __forceinline__ void store_tile_32x32_rows(uint32_t stage_smem, const float (&v)[32], int lane, float *out,
                                                      const long long (&rowoff)[8], int col0, const uint32_t (&zmask)[8],
                                                      long long pix_stride, long long line_stride) {
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; q++) {
        const uint32_t a = stage_smem + (uint32_t)lane * 128u + (uint32_t)((q ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]),
                     "f"(v[q * 4 + 3]) : "memory");
    }
    __syncwarp();
    const int c = lane & 7;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int row = 4 * i + (lane >> 3);
        float4 t;
        const uint32_t a = stage_smem + (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4);
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(a));
        if (rowoff[i] >= 0) {
            float *p = out + rowoff[i] + col0 + c * 4;
            *reinterpret_cast<float4 *>(p) = t;
            if (zmask[i] & 1u) *reinterpret_cast<float4 *>(p + pix_stride) = zero;
            if (zmask[i] & 2u) *reinterpret_cast<float4 *>(p + line_stride) = zero;
            if ((zmask[i] & 3u) == 3u) *reinterpret_cast<float4 *>(p + line_stride + pix_stride) = zero;
        }
    }
}

// Instruction descriptor (InstrDescriptor): c_format F32 [4,6)=1, a/b_format TF32 [7,10)=[10,13)=2,
// a_major bit 15, b_major bit 16 (1 = MN-major), N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <int BN, int STAGES>
struct SmemLayout {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 4;  // 16 KB (gemm: 128 rows x 128 B; wgrad: 32 px x 512 B)
    static constexpr int B_BYTES = BN * BLOCK_K * 4;
    static constexpr int BAR_OFF = STAGES * (A_BYTES + B_BYTES);
    static constexpr int BIAS_OFF = BAR_OFF + 256;          // BN floats: the tile's bias slice
    static constexpr int BYTES = BIAS_OFF + BN * 4 + 1024;  // + slack for 1024-byte alignment
};

// ------------------------------------------------------------------------------------------------
template <int BN, int STAGES, bool FAST>
__global__ void __launch_bounds__(THREADS) k_conv_tc_gemm(const float *__restrict__ src, const float *__restrict__ wpack,
                                                          const float *__restrict__ bias, float *__restrict__ out,
                                                          ConvGeom g, int n_valid, int out_cs, int act) {
    using L = SmemLayout<BN, STAGES>;
    const int rnd = act >> 8;  // bit 8: store the activated output rounded to TF32 (ctx->round_act)
    act &= 0xff;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + STAGES * L::A_BYTES, bar_base = base + L::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t acc_bar = bar_base + 8u * (2 * STAGES), tmem_slot = bar_base + 8u * (2 * STAGES + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BLOCK_M, n0 = blockIdx.y * BN;
    const int num_k = g.Kp / BLOCK_K;

    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), PRODUCERS); mbar_init(empty_bar(s), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) tmem_alloc<BN>(tmem_slot);
    float *bias_s = reinterpret_cast<float *>(smem_raw + (base - smem_u32(smem_raw)) + L::BIAS_OFF);
    if (tid < BN) bias_s[tid] = (bias && n0 + tid < n_valid) ? __ldg(bias + n0 + tid) : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_d;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_d) : "r"(tmem_slot));

    if (warp < 4) {
        // ------------------------------ producers
        const int j = tid & 7, rb = tid >> 3;
        const uint32_t off0 = ((uint32_t)(rb >> 3) << 10) + ((uint32_t)(rb & 7) << 7) + ((uint32_t)(j ^ (rb & 7)) << 4);
        // per-row state, set up once per tile. FAST: base offset + tap validity mask (a copy then costs ~5
        // instructions); otherwise the generic (b, y, x) decode is replayed per copy.
        long long rowoff[8];
        uint32_t rowmask[8];
        int rowB[8], rowY[8], rowX[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            int m = m0 + rb + 16 * i;
            if (FAST) {
                if (m < g.M) conv_row_setup(g, m, rowoff[i], rowmask[i]);
                else { rowoff[i] = 0; rowmask[i] = 0; }
            } else {
                if (m < g.M) conv_m_decode(g, m, rowB[i], rowY[i], rowX[i]);
                else { rowB[i] = -1; rowY[i] = 0; rowX[i] = 0; }
            }
        }
        const float *wrow = wpack + (size_t)(n0 + rb) * g.Kp + j * 4;
        const size_t wstep = (size_t)16 * g.Kp;
        for (int it = 0; it < num_k + LOOKAHEAD; ++it) {
            if (it < num_k) {
                const int s = it % STAGES;
                mbar_wait(empty_bar(s), ((it / STAGES) & 1) ^ 1);
                int kh, kw, c; bool kvalid;
                conv_k_decode(g, it * BLOCK_K + j * 4, kh, kw, c, kvalid);
                const uint32_t a_dst = a_base + s * L::A_BYTES + off0;
                if (FAST) {
                    int tapoff = 0; uint32_t need = 0xFFFFFFFFu;
                    if (kvalid) conv_tap_setup(g, kh, kw, c, tapoff, need);
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        const bool ok = (rowmask[i] & need) == need;
                        cp_async16(a_dst + i * 2048u, ok ? src + (rowoff[i] + tapoff) : src, ok ? 16u : 0u);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        long long o = (kvalid && rowB[i] >= 0) ? conv_src_offset(g, rowB[i], rowY[i], rowX[i], kh, kw) : -1;
                        const float *p = o >= 0 ? src + o + c : src;
                        cp_async16(a_dst + i * 2048u, p, o >= 0 ? 16u : 0u);
                    }
                }
                const uint32_t b_dst = b_base + s * L::B_BYTES + off0;
                const float *wp = wrow + (size_t)it * BLOCK_K;
#pragma unroll
                for (int i = 0; i < BN / 16; i++, wp += wstep) cp_async16(b_dst + i * 2048u, wp, 16u);
            }
            cp_async_commit();
            if (it >= LOOKAHEAD) {
                cp_async_wait<LOOKAHEAD>();
                fence_proxy_async();
                mbar_arrive(full_bar((it - LOOKAHEAD) % STAGES));
            }
        }
        // ------------------------------ epilogue: thread = one output pixel row, 32 channels per TMEM load
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int m = m0 + warp * 32 + lane;
        float *orow = out + (size_t)m * out_cs;
        const bool full_n = n0 + BN <= n_valid;  // whole tile inside the valid channels: no per-element guards
        const uint32_t stage_smem = a_base + warp * 4096u;  // operand stages are idle once acc_bar has fired
        const int rows_valid = g.M - (m0 + warp * 32);
#pragma unroll 1
        for (int cb = 0; cb < BN / 32; cb++) {
            float v[32];
            tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cb * 32), v);
            const int nb = n0 + cb * 32;
            if (full_n) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const float4 bq = *reinterpret_cast<const float4 *>(bias_s + cb * 32 + q * 4);
                    v[q * 4 + 0] += bq.x; v[q * 4 + 1] += bq.y; v[q * 4 + 2] += bq.z; v[q * 4 + 3] += bq.w;
                }
                if (act == DL4JB200_ACT_RELU) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = fmaxf(v[e], 0.f);
                } else if (act != DL4JB200_ACT_IDENTITY) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = dl4j_act(v[e], act);  // static indices keep v[] in registers
                }
                if (rnd) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = dl4j_tf32(v[e]);
                }
                store_tile_32x32(stage_smem, v, lane, out + (size_t)(m0 + warp * 32) * out_cs + nb, (size_t)out_cs, rows_valid);
            } else if (m < g.M) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int n = nb + q * 4;
                    if (n >= out_cs) continue;
                    float r[4];
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        r[e] = n + e < n_valid ? dl4j_rnd(dl4j_act(v[q * 4 + e] + bias_s[cb * 32 + q * 4 + e], act), rnd) : 0.f;
                    *reinterpret_cast<float4 *>(orow + n) = make_float4(r[0], r[1], r[2], r[3]);
                }
            }
        }
    } else {
        // ------------------------------ MMA issuer
        constexpr uint32_t idesc = umma_idesc(BLOCK_M, BN, 0, 0);
        for (int it = 0; it < num_k; ++it) {
            const int s = it % STAGES;
            mbar_wait(full_bar(s), (it / STAGES) & 1);
            tc_fence_after();
            if (lane == 0) {
                const uint64_t ad = umma_desc(a_base + s * L::A_BYTES, 16, 1024);
                const uint64_t bd = umma_desc(b_base + s * L::B_BYTES, 16, 1024);
#pragma unroll
                for (int k = 0; k < BLOCK_K / 8; k++)  // UMMA_K = 8 tf32 = 32 bytes = +2 in the address field
                    umma_tf32(tmem_d, ad + 2u * k, bd + 2u * k, idesc, (it | k) ? 1u : 0u);
                umma_commit(empty_bar(s));
            }
            __syncwarp();
        }
        if (lane == 0) umma_commit(acc_bar);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc<BN>(tmem_d);
}

// ------------------------------------------------------------------------------------------------
template <int BN, int STAGES>
__global__ void __launch_bounds__(THREADS) k_conv_tc_wgrad(const float *__restrict__ x, const float *__restrict__ delta,
                                                           float *__restrict__ partial, ConvGeom g, int delta_cs,
                                                           int rows_p, int m_per_split, int k_tiles) {
    using L = SmemLayout<BN, STAGES>;
    constexpr int NB_ATOMS = BN / 32;  // 32-float MN atoms across the im2col tile
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + STAGES * L::A_BYTES, bar_base = base + L::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t acc_bar = bar_base + 8u * (2 * STAGES), tmem_slot = bar_base + 8u * (2 * STAGES + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kt = blockIdx.x % k_tiles, ct = blockIdx.x / k_tiles;
    const int kcol0 = kt * BN, co0 = ct * BLOCK_M;
    const int ms = blockIdx.y * m_per_split, me = min(g.M, ms + m_per_split);
    const int num_k = (me - ms + BLOCK_K - 1) / BLOCK_K;  // >= 1 by construction of the split count

    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), PRODUCERS); mbar_init(empty_bar(s), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) tmem_alloc<BN>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_d;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_d) : "r"(tmem_slot));

    if (warp < 4) {
        // A' = delta[px, co0 .. co0+127]: 32 pieces per pixel row, 4 rows per pass, 8 passes
        const int pa = tid & 31, ra = tid >> 5;
        const bool a_col_ok = co0 + pa * 4 < delta_cs;
        // B' = im2col[px, kcol0 .. kcol0+BN-1]: BN/4 pieces per row
        constexpr int PPR = BN / 4, RPP = PRODUCERS / PPR, PASSES = BLOCK_K / RPP;
        const int pb = tid % PPR, rbb = tid / PPR;
        int kh, kw, c; bool kvalid;
        conv_k_decode(g, kcol0 + pb * 4, kh, kw, c, kvalid);
        kvalid = kvalid && (kcol0 + pb * 4) < g.Kp;
        // this thread's im2col column is one fixed (tap, channel quad): fold the tap into the pixel arithmetic
        const int dy = kh * g.dH - g.pT, dx = kw * g.dW - g.pL;
        const float *xcol = x + c;
        const float *dcol = delta + co0 + pa * 4;
        for (int it = 0; it < num_k + LOOKAHEAD; ++it) {
            if (it < num_k) {
                const int s = it % STAGES;
                mbar_wait(empty_bar(s), ((it / STAGES) & 1) ^ 1);
                const int mrow0 = ms + it * BLOCK_K;
                const uint32_t a_stage = a_base + s * L::A_BYTES, b_stage = b_base + s * L::B_BYTES;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int r = ra + 4 * i, m = mrow0 + r;
                    const bool ok = a_col_ok && m < me;
                    const uint32_t dst = a_stage + mn_tile_off<4>(r, pa >> 3, pa & 7);
                    cp_async16(dst, ok ? dcol + (size_t)m * delta_cs : delta, ok ? 16u : 0u);
                }
#pragma unroll
                for (int i = 0; i < PASSES; i++) {
                    const int r = rbb + RPP * i, m = mrow0 + r;
                    bool ok = kvalid && m < me;
                    size_t o = 0;
                    if (ok) {
                        uint32_t bb, rr, yy, xx;
                        g.div_HWm.divmod((uint32_t)m, bb, rr);
                        g.div_Wm.divmod(rr, yy, xx);
                        const int iy = (int)yy * g.sH + dy, ix = (int)xx * g.sW + dx;
                        ok = (unsigned)iy < (unsigned)g.Hs && (unsigned)ix < (unsigned)g.Ws;
                        o = (((size_t)bb * g.Hs + iy) * g.Ws + ix) * g.cs_src;
                    }
                    const uint32_t dst = b_stage + mn_tile_off<NB_ATOMS>(r, pb >> 3, pb & 7);
                    cp_async16(dst, ok ? xcol + o : x, ok ? 16u : 0u);
                }
            }
            cp_async_commit();
            if (it >= LOOKAHEAD) {
                cp_async_wait<LOOKAHEAD>();
                fence_proxy_async();
                mbar_arrive(full_bar((it - LOOKAHEAD) % STAGES));
            }
        }
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int co = co0 + warp * 32 + lane;
        float *prow = partial + ((size_t)blockIdx.y * rows_p + co) * g.Kp;
        const uint32_t stage_smem = a_base + warp * 4096u;
#pragma unroll 1
        for (int cb = 0; cb < BN / 32; cb++) {
            float v[32];
            tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cb * 32), v);
            if (kcol0 + cb * 32 + 32 <= g.Kp) {
                store_tile_32x32(stage_smem, v, lane, partial + ((size_t)blockIdx.y * rows_p + co0 + warp * 32) * g.Kp + kcol0 + cb * 32,
                                 (size_t)g.Kp, rows_p - (co0 + warp * 32));
            } else if (co < rows_p) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int k = kcol0 + cb * 32 + q * 4;
                    if (k < g.Kp) *reinterpret_cast<float4 *>(prow + k) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
                }
            }
        }
    } else {
        constexpr uint32_t idesc = umma_idesc(BLOCK_M, BN, 1, 1);
        for (int it = 0; it < num_k; ++it) {
            const int s = it % STAGES;
            mbar_wait(full_bar(s), (it / STAGES) & 1);
            tc_fence_after();
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / 8; k++) {  // 8 pixels = two 4-row groups per MMA; LBO = atom stride along
                                                         // MN (512 B), SBO = stride between 4-row groups along K
                    const uint64_t ad = umma_desc(a_base + s * L::A_BYTES + k * 2 * 4 * 512, 512, 4 * 512, 1);
                    const uint64_t bd = umma_desc(b_base + s * L::B_BYTES + k * 2 * NB_ATOMS * 512, 512, NB_ATOMS * 512, 1);
                    umma_tf32(tmem_d, ad, bd, idesc, (it | k) ? 1u : 0u);
                }
                umma_commit(empty_bar(s));
            }
            __syncwarp();
        }
        if (lane == 0) umma_commit(acc_bar);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc<BN>(tmem_d);
}

// ------------------------------------------------------------------------------------------------
// TMA-fed variant of k_conv_tc_gemm (channels per tap a multiple of 32, unit stride for backprop-data):
// one elected thread issues, per K-chunk, ONE im2col-mode TMA for the 128-pixel x 32-channel A tile and ONE tiled
// TMA for the BN x 32 filter tile; both signal the stage's mbarrier with their byte count. No per-element address
// arithmetic is left on the SM; padding, stride and image borders are resolved by the TMA unit (zero fill).

// Split-K (gridDim.z > 1): CTA (x, y, z) accumulates K-chunks [z * k_per_split, (z+1) * k_per_split) of its tile and stores the raw
// accumulator at out + z * split_stride (the caller passes no bias, identity activation, and folds the splits with
// k_splitk_finalize) -- for GEMMs with a handful of tiles and a long K, i.e. the dense layers at minibatch-sized M.
template <int BN, int STAGES>
__global__ void __launch_bounds__(THREADS) k_conv_tc_gemm_tma(const __grid_constant__ CUtensorMap tmA,
                                                              const __grid_constant__ CUtensorMap tmB,
                                                              const float *__restrict__ bias, float *__restrict__ out,
                                                              TmaGemmGeom g, int n_valid, int out_cs, int act, int k_per_split,
                                                              size_t split_stride) {
    using L = SmemLayout<BN, STAGES>;
    const int rnd = act >> 8;
    act &= 0xff;
    const int k_lo = (int)blockIdx.z * k_per_split, k_hi = min(g.num_k, k_lo + k_per_split);
    out += (size_t)blockIdx.z * split_stride;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + STAGES * L::A_BYTES, bar_base = base + L::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t acc_bar = bar_base + 8u * (2 * STAGES), tmem_slot = bar_base + 8u * (2 * STAGES + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BLOCK_M, n0 = blockIdx.y * BN;

    if (tid == 0) {
        tma::prefetch_desc(&tmA);
        tma::prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) tmem_alloc<BN>(tmem_slot);
    __syncthreads();
    pdl_trigger();  // after the TMEM allocation (see k_conv_tc_gemm_tma_persist); no-ops under an ordinary launch
    pdl_wait();
    float *bias_s = reinterpret_cast<float *>(smem_raw + (base - smem_u32(smem_raw)) + L::BIAS_OFF);
    if (tid < BN) bias_s[tid] = (bias && n0 + tid < n_valid) ? __ldg(bias + n0 + tid) : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_d;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_d) : "r"(tmem_slot));

    if (warp < 4) {
        if (tid == 0) {
            // ------------------------------ TMA producer (one thread)
            uint32_t img, r, oy, ox;
            g.div_HWm.divmod((uint32_t)m0, img, r);
            g.div_Wm.divmod(r, oy, ox);
            const int w0 = (int)ox * g.sW + g.lowW, h0 = (int)oy * g.sH + g.lowH;
            for (int it = k_lo; it < k_hi; ++it) {
                const int j = it - k_lo, s = j % STAGES;
                mbar_wait(empty_bar(s), ((j / STAGES) & 1) ^ 1);
                uint32_t tap, c0, kh, kw;
                g.div_Cg.divmod((uint32_t)(it * BLOCK_K), tap, c0);
                g.div_kW.divmod(tap, kh, kw);
                tma::expect_tx(full_bar(s), (uint32_t)(L::A_BYTES + L::B_BYTES));
                tma::load_im2col_4d(a_base + s * L::A_BYTES, &tmA, full_bar(s), (int)c0, w0, h0, (int)img,
                                    (uint16_t)(kw * g.dW), (uint16_t)(kh * g.dH));
                tma::load_2d(b_base + s * L::B_BYTES, &tmB, full_bar(s), it * BLOCK_K, n0);
            }
        }
        __syncwarp();
        // ------------------------------ epilogue (same as k_conv_tc_gemm)
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int m = m0 + warp * 32 + lane;
        float *orow = out + (size_t)m * out_cs;
        const bool full_n = n0 + BN <= n_valid;
        const uint32_t stage_smem = a_base + warp * 4096u;
        const int rows_valid = g.M - (m0 + warp * 32);
#pragma unroll 1
        for (int cb = 0; cb < BN / 32; cb++) {
            float v[32];
            tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cb * 32), v);
            const int nb = n0 + cb * 32;
            if (full_n) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const float4 bq = *reinterpret_cast<const float4 *>(bias_s + cb * 32 + q * 4);
                    v[q * 4 + 0] += bq.x; v[q * 4 + 1] += bq.y; v[q * 4 + 2] += bq.z; v[q * 4 + 3] += bq.w;
                }
                if (act == DL4JB200_ACT_RELU) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = fmaxf(v[e], 0.f);
                } else if (act != DL4JB200_ACT_IDENTITY) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = dl4j_act(v[e], act);  // static indices keep v[] in registers
                }
                if (rnd) {
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = dl4j_tf32(v[e]);
                }
                store_tile_32x32(stage_smem, v, lane, out + (size_t)(m0 + warp * 32) * out_cs + nb, (size_t)out_cs, rows_valid);
            } else if (m < g.M) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int n = nb + q * 4;
                    if (n >= out_cs) continue;
                    float r4[4];
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        r4[e] = n + e < n_valid ? dl4j_rnd(dl4j_act(v[q * 4 + e] + bias_s[cb * 32 + q * 4 + e], act), rnd) : 0.f;
                    *reinterpret_cast<float4 *>(orow + n) = make_float4(r4[0], r4[1], r4[2], r4[3]);
                }
            }
        }
    } else {
        constexpr uint32_t idesc = umma_idesc(BLOCK_M, BN, 0, 0);
        for (int j = 0; j < k_hi - k_lo; ++j) {
            const int s = j % STAGES;
            mbar_wait(full_bar(s), (j / STAGES) & 1);
            tc_fence_after();
            if (lane == 0) {
                const uint64_t ad = umma_desc(a_base + s * L::A_BYTES, 16, 1024);
                const uint64_t bd = umma_desc(b_base + s * L::B_BYTES, 16, 1024);
#pragma unroll
                for (int k = 0; k < BLOCK_K / 8; k++) umma_tf32(tmem_d, ad + 2u * k, bd + 2u * k, idesc, (j | k) ? 1u : 0u);
                umma_commit(empty_bar(s));
            }
            __syncwarp();
        }
        if (lane == 0) umma_commit(acc_bar);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc<BN>(tmem_d);
}

// Folds the split-K partials of k_conv_tc_gemm_tma in split order (deterministic) and applies what the single-pass epilogue
// would have: bias, activation, the optional fused backprop epilogue out = (acc + add) masked by mask > 0, TF32 rounding.
__global__ void __launch_bounds__(256) k_splitk_finalize(const float *__restrict__ partial, float *__restrict__ out,
                                                         const float *__restrict__ bias, const float *__restrict__ epi_add,
                                                         const float *__restrict__ epi_mask, int M, int n_valid, int out_cs,
                                                         int splits, size_t split_stride, int act) {
    const int rnd = act >> 8;
    act &= 0xff;
    pdl_trigger();
    pdl_wait();
    const int Q = out_cs >> 2;
    const size_t total = (size_t)M * Q, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int q = (int)(i % Q);
        const size_t o = (i / Q) * out_cs + (size_t)q * 4;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        int sp = 0;
        for (; sp + 1 < splits; sp += 2) {  // two independent chains: both loads of a round in flight; fixed combination order
            const float4 u = *reinterpret_cast<const float4 *>(partial + (size_t)sp * split_stride + o);
            const float4 w = *reinterpret_cast<const float4 *>(partial + (size_t)(sp + 1) * split_stride + o);
            a0.x += u.x; a0.y += u.y; a0.z += u.z; a0.w += u.w;
            a1.x += w.x; a1.y += w.y; a1.z += w.z; a1.w += w.w;
        }
        if (sp < splits) {
            const float4 u = *reinterpret_cast<const float4 *>(partial + (size_t)sp * split_stride + o);
            a0.x += u.x; a0.y += u.y; a0.z += u.z; a0.w += u.w;
        }
        float r[4] = {a0.x + a1.x, a0.y + a1.y, a0.z + a1.z, a0.w + a1.w};
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int n = q * 4 + e;
            float v = 0.f;
            if (n < n_valid) {
                v = dl4j_act(r[e] + (bias ? __ldg(bias + n) : 0.f), act);
                if (epi_add) v += epi_add[o + e];
                if (epi_mask && !(epi_mask[o + e] > 0.f)) v = 0.f;
                v = dl4j_rnd(v, rnd);
            }
            r[e] = v;
        }
        *reinterpret_cast<float4 *>(out + o) = make_float4(r[0], r[1], r[2], r[3]);
    }
}

// ------------------------------------------------------------------------------------------------
// Persistent form of the TMA-fed GEMM: one CTA per SM loops over output tiles. Ten warps: 0-3 and 6-9 epilogue, 4 MMA
// issuer, 5 TMA producer. The accumulator is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i (TMEM ->
// registers -> staging smem -> coalesced global stores) overlaps the TMA loads and MMAs of tile i+1, and the operand
// pipeline never drains between tiles. This is what the small-K 1x1 convolutions need: their per-tile main loop is
// only 2-8 K-chunks long, so set-up, drain and epilogue dominated the one-tile-per-CTA kernel.
constexpr int PTHREADS = 320;

template <int BN, int STAGES, bool PAIR = false>
struct PersistSmem {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 4;
    static constexpr int B_BYTES = (PAIR ? BN / 2 : BN) * BLOCK_K * 4;  // a CTA of a pair holds half of the B tile
    static constexpr int STAGING_OFF = STAGES * (A_BYTES + B_BYTES);  // 8 warps x 4 KB epilogue staging
    static constexpr int BAR_OFF = STAGING_OFF + 8 * 4096;
    static constexpr int BYTES = BAR_OFF + 256 + 1024;
};

// STATS: 0 none, 1 forward batch-norm statistics of the output, 2 backward batch-norm sums (TmaGemmGeom::bst_*)
// PAIR: launched as clusters of two CTAs; a tile is 256 rows x BN, CTA `rank` owns rows [128 * rank, 128 * rank + 128) -- its
// own A stream, its own accumulator lanes and epilogue -- and loads B rows [rank * BN/2, +BN/2) only (see the cta_group::2
// helpers above). total_tiles counts those 256-row tiles; tmB's box is BN/2 rows.
template <int BN, int STAGES, bool REMAP, int STATS = 0, bool PAIR = false>
__global__ void __launch_bounds__(PTHREADS, 1) k_conv_tc_gemm_tma_persist(const __grid_constant__ CUtensorMap tmA,
                                                                          const __grid_constant__ CUtensorMap tmB,
                                                                          const float *__restrict__ bias,
                                                                          float *__restrict__ out, TmaGemmGeom g,
                                                                          int n_valid, int out_cs, int act, int n_tiles,
                                                                          int total_tiles) {
    using L = PersistSmem<BN, STAGES, PAIR>;
    const int rnd = act >> 8;
    act &= 0xff;
    const bool fused = !REMAP && STATS != 2 && (g.epi_add != nullptr || g.epi_mask != nullptr);
    // tile-loop identity: a lone CTA, or one CTA pair (cluster dims (2,1,1): rank == blockIdx.x & 1)
    const int rank = PAIR ? (int)(blockIdx.x & 1u) : 0;
    const int unit = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, units = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    constexpr int TILE_M = PAIR ? 2 * BLOCK_M : BLOCK_M;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + STAGES * L::A_BYTES, bar_base = base + L::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
    auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4), bres_bar = bar_base + 8u * (2 * STAGES + 5);
    // B-resident mode: every tile of this CTA uses the same filter columns (the grid is a multiple of the N-tile count, so
    // t % n_tiles is constant per CTA) and the whole K extent of the filter fits the ring's B slots: B is fetched ONCE
    // per CTA instead of once per tile and only A streams (small-K 1x1 layers, the 7-row stem).
    const bool bres = g.num_k <= STAGES && units % n_tiles == 0;  // grid multiple of n_tiles: a CTA keeps its N tile

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        tma::prefetch_desc(&tmA);
        tma::prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        // accumulator hand-back: every epilogue lane arrives; in a pair one lane per epilogue warp of BOTH CTAs arrives on
        // the leader's barrier
        for (int b = 0; b < 2; b++) { mbar_init(tfull_bar(b), 1); mbar_init(tempty_bar(b), PAIR ? 16 : 256); }
        mbar_init(bres_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        if (PAIR) tmem_alloc2<2 * BN>(tmem_slot);
        else tmem_alloc<2 * BN>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (PAIR) cluster_sync_all();  // the partner's barriers exist before anything is signalled across
    uint32_t tmem_d;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_d) : "r"(tmem_slot));
    // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) ran while the previous kernel
    // of the stream was still draining; global memory is touched only below. The trigger comes AFTER this CTA holds its
    // TMEM: a dependent CTA scheduled onto this SM may then block in its own tcgen05.alloc until we free ours, but can
    // never take columns this CTA is still waiting for.
    pdl_trigger();
    pdl_wait();

    if (warp == 5) {
        // ------------------------------ TMA producer
        if (lane == 0) {
            int kit = 0;
            const int nb_off = PAIR ? rank * (BN / 2) : 0;  // this CTA's share of the B tile
            if (bres && unit < total_tiles) {
                if (rank == 0) tma::expect_tx(bres_bar, (uint32_t)(g.num_k * L::B_BYTES * (PAIR ? 2 : 1)));
                const uint32_t lead = PAIR ? mapa_rank(bres_bar, 0) : bres_bar;
                for (int it = 0; it < g.num_k; ++it) {
                    if (PAIR) tma::load_2d_pair(b_base + it * L::B_BYTES, &tmB, lead, it * BLOCK_K, (unit % n_tiles) * BN + nb_off);
                    else tma::load_2d(b_base + it * L::B_BYTES, &tmB, bres_bar, it * BLOCK_K, (unit % n_tiles) * BN);
                }
            }
            for (int t = unit; t < total_tiles; t += units) {
                const int m0 = (t / n_tiles) * TILE_M + rank * BLOCK_M, n0 = (t % n_tiles) * BN + nb_off;
                uint32_t img, r, oy, ox;
                g.div_HWm.divmod((uint32_t)m0, img, r);
                g.div_Wm.divmod(r, oy, ox);
                const int w0 = (int)ox * g.sW + g.lowW, h0 = (int)oy * g.sH + g.lowH;
                for (int it = 0; it < g.num_k; ++it, ++kit) {
                    const int s = kit % STAGES;
                    mbar_wait(empty_bar(s), ((kit / STAGES) & 1) ^ 1);
                    uint32_t tap, c0, kh, kw;
                    g.div_Cg.divmod((uint32_t)(it * BLOCK_K), tap, c0);
                    g.div_kW.divmod(tap, kh, kw);
                    if (PAIR) {
                        // both CTAs' bytes complete on the leader's barrier (its expect_tx may come after the partner's
                        // first bytes: the phase cannot complete before the leader's own arrive)
                        if (rank == 0) tma::expect_tx(full_bar(s), 2u * (uint32_t)(bres ? L::A_BYTES : L::A_BYTES + L::B_BYTES));
                        const uint32_t lead = mapa_rank(full_bar(s), 0);
                        tma::load_im2col_4d_pair(a_base + s * L::A_BYTES, &tmA, lead, (int)c0, w0, h0, (int)img, (uint16_t)(kw * g.dW),
                                                 (uint16_t)(kh * g.dH));
                        if (!bres) tma::load_2d_pair(b_base + s * L::B_BYTES, &tmB, lead, it * BLOCK_K, n0);
                    } else {
                        tma::expect_tx(full_bar(s), (uint32_t)(bres ? L::A_BYTES : L::A_BYTES + L::B_BYTES));
                        tma::load_im2col_4d(a_base + s * L::A_BYTES, &tmA, full_bar(s), (int)c0, w0, h0, (int)img,
                                            (uint16_t)(kw * g.dW), (uint16_t)(kh * g.dH));
                        if (!bres) tma::load_2d(b_base + s * L::B_BYTES, &tmB, full_bar(s), it * BLOCK_K, n0);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 4) {
        // ------------------------------ MMA issuer
        constexpr uint32_t idesc = umma_idesc(TILE_M, BN, 0, 0);
        int kit = 0, i = 0;
        if (bres && unit < total_tiles && rank == 0) mbar_wait(bres_bar, 0);
        for (int t = unit; t < total_tiles && rank == 0; t += units, ++i) {  // in a pair only the leader issues
            const int buf = i & 1;
            mbar_wait(tempty_bar(buf), ((i >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
            for (int it = 0; it < g.num_k; ++it, ++kit) {
                const int s = kit % STAGES;
                mbar_wait(full_bar(s), (kit / STAGES) & 1);
                tc_fence_after();
                if (lane == 0) {
                    const uint64_t ad = umma_desc(a_base + s * L::A_BYTES, 16, 1024);
                    const uint64_t bd = umma_desc(b_base + (bres ? it : s) * L::B_BYTES, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / 8; k++) {
                        if (PAIR) umma_tf32_2(acc, ad + 2u * k, bd + 2u * k, idesc, (it | k) ? 1u : 0u);
                        else umma_tf32(acc, ad + 2u * k, bd + 2u * k, idesc, (it | k) ? 1u : 0u);
                    }
                    if (PAIR) umma_commit2(empty_bar(s));  // frees the slot in both CTAs
                    else umma_commit(empty_bar(s));
                }
                __syncwarp();
            }
            if (lane == 0) {
                if (PAIR) umma_commit2(tfull_bar(buf));
                else umma_commit(tfull_bar(buf));
            }
            __syncwarp();
        }
    } else {
        // ------------------------------ epilogue: warps 0-3 take the even 32-column chunks of a tile, warps 6-9 the odd
        // ones (a warp may only read the TMEM lane quadrant warp % 4, so each quadrant gets two warps). Twice the
        // loads / stores in flight per SM: the epilogue of the HBM-bound layers is latency-, not issue-limited.
        const int quad = warp & 3, grp = warp >= 6 ? 1 : 0;
        const uint32_t stage_smem = base + L::STAGING_OFF + (uint32_t)(grp * 4 + quad) * 4096u;
        // STATS: per-warp running sums of (z - K), (z - K)^2 for the columns of its chunks. The launch is sized so that
        // gridDim.x is a multiple of n_tiles: every tile of this CTA then has the same column range and the sums stay in
        // registers until the single flush after the tile loop.
        constexpr int NCH = STATS ? (BN / 32 + 1) / 2 : 1;
        constexpr int NCH2 = STATS == 2 ? NCH : 1;
        float4 st_s[NCH], st_q[NCH], st_k[NCH];
        float4 st_a[NCH2], st_b[NCH2];  // STATS == 2: gamma / std and beta of the gate; st_k holds the saved mean
        if (STATS) {
#pragma unroll
            for (int j = 0; j < NCH; j++) {
                st_s[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                st_q[j] = st_s[j];
                const int cb = grp + 2 * j;
                const int col = (unit % n_tiles) * BN + cb * 32;
                if constexpr (STATS == 1) {
                    st_k[j] = cb < BN / 32 ? __ldg(reinterpret_cast<const float4 *>(g.stat_pivot + col) + (lane & 7)) : st_s[j];
                } else if constexpr (STATS == 2) {
                    st_k[j] = st_s[j];
                    st_a[j] = st_s[j];
                    st_b[j] = st_s[j];
                    if (cb < BN / 32) {
                        const int c4 = col + (lane & 7) * 4;
                        st_k[j] = __ldg(reinterpret_cast<const float4 *>(g.bst_mean + c4));
                        if (g.bst_relu) {  // identical to bn_mask_coef (bn.cu): (float)(gamma / std) in double, beta as is
                            const float4 sd = __ldg(reinterpret_cast<const float4 *>(g.bst_std + c4));
                            float4 gm = make_float4(1.f, 1.f, 1.f, 1.f);
                            if (g.bst_gamma) gm = __ldg(reinterpret_cast<const float4 *>(g.bst_gamma + c4));
                            st_a[j] = make_float4((float)((double)gm.x / (double)sd.x), (float)((double)gm.y / (double)sd.y),
                                                  (float)((double)gm.z / (double)sd.z), (float)((double)gm.w / (double)sd.w));
                            if (g.bst_beta) st_b[j] = __ldg(reinterpret_cast<const float4 *>(g.bst_beta + c4));
                        }
                    }
                }
            }
        }
        // hand an accumulator buffer back to the MMA issuer (the caller has fenced its TMEM reads)
        const uint32_t lead_tempty[2] = {PAIR ? mapa_rank(tempty_bar(0), 0) : 0u, PAIR ? mapa_rank(tempty_bar(1), 0) : 0u};
        auto release_acc = [&](int b) {
            if (PAIR) {
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(lead_tempty[b]);
            } else {
                mbar_arrive(tempty_bar(b));
            }
        };
        int i = 0;
        for (int t = unit; t < total_tiles; t += units, ++i) {
            const int buf = i & 1;
            const int m0 = (t / n_tiles) * TILE_M + rank * BLOCK_M, n0 = (t % n_tiles) * BN;
            const bool full_n = n0 + BN <= n_valid;
            const int m = m0 + quad * 32 + lane;
            const int rows_valid = g.M - (m0 + quad * 32);
            long long rowoff[8];
            uint32_t zmask[8];
            if (REMAP) {
#pragma unroll
                for (int r8 = 0; r8 < 8; r8++) {
                    const int mm = m0 + quad * 32 + 4 * r8 + (lane >> 3);
                    rowoff[r8] = -1;
                    zmask[r8] = 0;
                    if (mm < g.M) {
                        uint32_t img, rr, oy, ox;
                        g.div_HWm.divmod((uint32_t)mm, img, rr);
                        g.div_Wm.divmod(rr, oy, ox);
                        rowoff[r8] = (((long long)img * g.rm_H + (long long)oy * g.rm_sy) * g.rm_W + (long long)ox * g.rm_sx) * out_cs;
                        if (g.rm_zero)  // the hole pixels of this sample's 2x2 cell that lie inside the image
                            zmask[r8] = ((g.rm_sx == 2 && (int)ox * 2 + 1 < g.rm_W) ? 1u : 0u) | ((g.rm_sy == 2 && (int)oy * 2 + 1 < g.rm_H) ? 2u : 0u);
                    }
                }
            }
            if (fused && grp == 0) {
                // pull the epilogue's extra operands of the NEXT tile of this CTA (and, on the first pass, of this one)
                // into L2 now: one bulk prefetch per row and tensor, a whole tile of lead time
                for (int tp = (i == 0 ? t : t + units); tp <= t + units && tp < total_tiles; tp += units) {
                    const int pm = (tp / n_tiles) * TILE_M + rank * BLOCK_M + quad * 32 + lane, pn = (tp % n_tiles) * BN;
                    if (pm < g.M) {
                        const size_t o = (size_t)pm * out_cs + pn;
                        const uint32_t bytes = (uint32_t)(min(BN, out_cs - pn) * 4);
                        if (g.epi_add) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g.epi_add + o), "r"(bytes) : "memory");
                        if (g.epi_mask) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g.epi_mask + o), "r"(bytes) : "memory");
                    }
                }
            }
            mbar_wait(tfull_bar(buf), (i >> 1) & 1);
            tc_fence_after();
            const uint32_t acc = tmem_d + (uint32_t)(buf * BN) + ((uint32_t)(quad * 32) << 16);
            constexpr int NCB = BN / 32;
            if (grp >= NCB) {  // BN == 32: one chunk only, the second warp group just keeps the barrier phase moving
                tc_fence_before();
                release_acc(buf);
                continue;
            }
            const int cb_last = NCB == 1 ? 0 : ((NCB - 1 - grp) & ~1) + grp;  // last chunk of this group's parity
            if constexpr (STATS == 2) {  // backprop-data + backward batch-norm sums; whole tiles only (checked on the host)
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    const int cb = grp + 2 * j;
                    if (cb < NCB) {
                        float v[32];
                        tmem_ld32(acc + (uint32_t)(cb * 32), v);
                        if (cb == cb_last) {
                            tc_fence_before();
                            release_acc(buf);
                        }
                        store_tile_32x32_bst(stage_smem, v, lane, (size_t)(m0 + quad * 32) * out_cs + n0 + cb * 32, (size_t)out_cs, rows_valid, out,
                                             g.epi_add, g.epi_mask, rnd, g.bst_x, st_k[j], st_a[j], st_b[j], g.bst_relu, st_s[j], st_q[j]);
                    }
                }
                continue;
            }
            if constexpr (STATS == 1) {  // identity activation, whole tiles only (checked on the host); chunk loop unrolled: static accumulators
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    const int cb = grp + 2 * j;
                    if (cb < NCB) {
                        float v[32];
                        tmem_ld32(acc + (uint32_t)(cb * 32), v);
                        if (cb == cb_last) {
                            tc_fence_before();
                            release_acc(buf);
                        }
                        const int nb = n0 + cb * 32;
                        if (bias) {
#pragma unroll
                            for (int q = 0; q < 8; q++) {
                                const float4 bq = __ldg(reinterpret_cast<const float4 *>(bias + nb) + q);
                                v[q * 4 + 0] += bq.x; v[q * 4 + 1] += bq.y; v[q * 4 + 2] += bq.z; v[q * 4 + 3] += bq.w;
                            }
                        }
                        store_tile_32x32_stats(stage_smem, v, lane, out + (size_t)(m0 + quad * 32) * out_cs + nb, (size_t)out_cs, rows_valid,
                                               st_k[j], st_s[j], st_q[j]);
                    }
                }
                continue;
            }
#pragma unroll 1
            for (int cb = grp; cb < NCB; cb += (NCB == 1 ? 1 : 2)) {
                float v[32];
                tmem_ld32(acc + (uint32_t)(cb * 32), v);
                if (cb == cb_last) {  // last TMEM read of this warp in this tile: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    release_acc(buf);
                }
                const int nb = n0 + cb * 32;
                if (full_n) {
                    if (bias) {
#pragma unroll
                        for (int q = 0; q < 8; q++) {
                            const float4 bq = __ldg(reinterpret_cast<const float4 *>(bias + nb) + q);
                            v[q * 4 + 0] += bq.x; v[q * 4 + 1] += bq.y; v[q * 4 + 2] += bq.z; v[q * 4 + 3] += bq.w;
                        }
                    }
                    if (act == DL4JB200_ACT_RELU) {
#pragma unroll
                        for (int e = 0; e < 32; e++) v[e] = fmaxf(v[e], 0.f);
                    } else if (act != DL4JB200_ACT_IDENTITY) {
#pragma unroll
                        for (int e = 0; e < 32; e++) v[e] = dl4j_act(v[e], act);  // static indices keep v[] in registers
                    }
                    if (rnd && !fused) {
#pragma unroll
                        for (int e = 0; e < 32; e++) v[e] = dl4j_tf32(v[e]);
                    }
                    if (REMAP) store_tile_32x32_rows(stage_smem, v, lane, out, rowoff, nb, zmask, (long long)out_cs, (long long)g.rm_W * out_cs);
                    else if (fused)
                        store_tile_32x32_fused(stage_smem, v, lane, (size_t)(m0 + quad * 32) * out_cs + nb, (size_t)out_cs, rows_valid,
                                               out, g.epi_add, g.epi_mask, rnd);
                    else store_tile_32x32(stage_smem, v, lane, out + (size_t)(m0 + quad * 32) * out_cs + nb, (size_t)out_cs, rows_valid);
                } else if (m < g.M) {
                    long long ro = (long long)m * out_cs;
                    if (REMAP) {
                        uint32_t img, rr, oy, ox;
                        g.div_HWm.divmod((uint32_t)m, img, rr);
                        g.div_Wm.divmod(rr, oy, ox);
                        ro = (((long long)img * g.rm_H + (long long)oy * g.rm_sy) * g.rm_W + (long long)ox * g.rm_sx) * out_cs;
                    }
                    float *orow = out + ro;
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        const int n = nb + q * 4;
                        if (n >= out_cs) continue;
                        float r4[4];
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const int nn = n + e;
                            float val = dl4j_act(v[q * 4 + e] + (bias ? __ldg(bias + nn) : 0.f), act);
                            if (fused && nn < n_valid) {
                                if (g.epi_add) val += g.epi_add[ro + nn];
                                if (g.epi_mask && !(g.epi_mask[ro + nn] > 0.f)) val = 0.f;
                            }
                            r4[e] = nn < n_valid ? dl4j_rnd(val, rnd) : 0.f;
                        }
                        *reinterpret_cast<float4 *>(orow + n) = make_float4(r4[0], r4[1], r4[2], r4[3]);
                    }
                }
            }
        }
        if (STATS && grp < BN / 32) {
            // flush: fold the four lanes that hold the same columns (rows r, r+1, r+2, r+3 mod 4 of every group); then, when
            // both warp groups are active (BN >= 64), the four lane quadrants of the CTA are combined through the staging
            // smem so that the CTA writes ONE partial row (the finalize fold is a latency-bound tail: fewer rows, shorter)
            constexpr bool COMBINE = BN >= 64;
            const int ncol0 = (unit % n_tiles) * BN;
            const int srow = PAIR ? (unit / n_tiles) * 2 + rank : unit / n_tiles;  // this CTA's partial row
            __syncwarp();
#pragma unroll
            for (int j = 0; j < NCH; j++) {
                const int cb = grp + 2 * j;
                if (cb < BN / 32) {
                    float4 a = st_s[j], b = st_q[j];
#pragma unroll
                    for (int o = 8; o <= 16; o <<= 1) {
                        a.x += __shfl_xor_sync(0xffffffffu, a.x, o); a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
                        a.z += __shfl_xor_sync(0xffffffffu, a.z, o); a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
                        b.x += __shfl_xor_sync(0xffffffffu, b.x, o); b.y += __shfl_xor_sync(0xffffffffu, b.y, o);
                        b.z += __shfl_xor_sync(0xffffffffu, b.z, o); b.w += __shfl_xor_sync(0xffffffffu, b.w, o);
                    }
                    if (lane < 8) {
                        if (COMBINE) {
                            const uint32_t sa = stage_smem + (uint32_t)j * 256u + (uint32_t)lane * 16u;
                            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(sa), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
                            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(sa + 128u), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
                        } else {
                            float *p = g.stat_out + (size_t)(srow * 4 + quad) * 2 * out_cs + ncol0 + cb * 32 + lane * 4;
                            *reinterpret_cast<float4 *>(p) = a;
                            *reinterpret_cast<float4 *>(p + out_cs) = b;
                        }
                    }
                }
            }
            if (COMBINE) {
                asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight epilogue warps
                if (quad == 0 && lane < 8) {
#pragma unroll
                    for (int j = 0; j < NCH; j++) {
                        const int cb = grp + 2 * j;
                        if (cb < BN / 32) {
                            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll
                            for (int qq = 0; qq < 4; qq++) {  // fixed order: deterministic
                                const uint32_t sa = base + L::STAGING_OFF + (uint32_t)(grp * 4 + qq) * 4096u + (uint32_t)j * 256u + (uint32_t)lane * 16u;
                                float4 u, w;
                                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(u.x), "=f"(u.y), "=f"(u.z), "=f"(u.w) : "r"(sa));
                                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(w.x), "=f"(w.y), "=f"(w.z), "=f"(w.w) : "r"(sa + 128u));
                                a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w;
                                b.x += w.x; b.y += w.y; b.z += w.z; b.w += w.w;
                            }
                            float *p = g.stat_out + (size_t)srow * 2 * out_cs + ncol0 + cb * 32 + lane * 4;
                            *reinterpret_cast<float4 *>(p) = a;
                            *reinterpret_cast<float4 *>(p + out_cs) = b;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();  // the partner has finished with the paired allocation (and with this CTA's barriers)
    if (warp == 4) {
        if (PAIR) tmem_dealloc2<2 * BN>(tmem_d);
        else tmem_dealloc<2 * BN>(tmem_d);
    }
}

// grid of a statistics-carrying launch: <= #SMs, <= total tiles, a multiple of n_tiles (0: impossible)
static inline int persist_stat_grid(int sm_count, int total, int n_tiles) {
    int g = total < sm_count ? total : sm_count;
    return g / n_tiles * n_tiles;
}

// partial rows such a launch writes: one per CTA column group when the CTA combines its quadrants (bn >= 64), else four
static inline int persist_stat_rows(int grid, int n_tiles, int bn) { return grid / n_tiles * (bn >= 64 ? 1 : 4); }

// one instantiation of the persistent kernel: shared-memory opt-in once per device, then the (clustered) PDL launch
template <int BN, int STAGES, bool REMAP, int STATS, bool PAIR>
static int launch_persist_variant(dl4jb200_ctx *ctx, int grid, const CUtensorMap &tmA, const CUtensorMap &tmB, const TmaGemmGeom &g,
                                  const float *bias, float *out, int n_valid, int out_cs, int act, int n_tiles, int total) {
    using L = PersistSmem<BN, STAGES, PAIR>;
    auto kern = k_conv_tc_gemm_tma_persist<BN, STAGES, REMAP, STATS, PAIR>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
        configured[ctx->device & 15] = true;
    }
    if (PAIR)
        DL4J_CUDA_TRY(dl4j_launch_pdl_cluster(kern, dim3(grid), dim3(PTHREADS), L::BYTES, ctx->stream, 2, tmA, tmB, bias, out, g, n_valid, out_cs,
                                              act, n_tiles, total));
    else
        DL4J_CUDA_TRY(dl4j_launch_pdl(kern, dim3(grid), dim3(PTHREADS), L::BYTES, ctx->stream, tmA, tmB, bias, out, g, n_valid, out_cs, act,
                                      n_tiles, total));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// pipeline depth of the pair kernels: a stage is 16 KB of A + BN * 64 B of B
template <int BN>
struct PairStages { static constexpr int N = BN == 256 ? 5 : (BN == 128 ? 7 : 8); };

// pair: tiles are 256 rows, tmB was built with a BN/2-row box, the grid is twice the number of pairs
template <int BN, int STAGES>
static int launch_gemm_tma_persist(dl4jb200_ctx *ctx, const CUtensorMap &tmA, const CUtensorMap &tmB, const TmaGemmGeom &g,
                                   const float *bias, float *out, int n_valid, int out_cs, int act, bool pair = false) {
    constexpr int PS = PairStages<BN>::N;
    const int n_tiles = dl4j_ceil_div(out_cs, BN), tile_m = pair ? 2 * BLOCK_M : BLOCK_M, total = dl4j_ceil_div(g.M, tile_m) * n_tiles;
    const int slots = pair ? ctx->sm_count / 2 : ctx->sm_count, stages = pair ? PS : STAGES, mult = pair ? 2 : 1;
    int units = total < slots ? total : slots;
    if (n_tiles > 1 && g.num_k <= stages && units >= 2 * n_tiles) units = units / n_tiles * n_tiles;  // enables the B-resident mode
    if constexpr (BN < 64) {
        if (pair) { dl4j_set_error("conv_tc_gemm_tma: CTA pairs need tiles of at least 64 columns"); return DL4JB200_ERR_INVALID_ARG; }
    }
    if (g.stat_out && g.bst_x) {  // backward batch-norm sums in a backprop-data epilogue
        if constexpr (BN == 64 || BN == 128) {
            const int gs = persist_stat_grid(slots, total, n_tiles);
            if (g.remap || gs <= 0 || n_valid % BN != 0 || out_cs != n_valid || (act & 0xff) != DL4JB200_ACT_IDENTITY || bias) {
                dl4j_set_error("conv_tc_gemm_tma: this launch cannot carry the backward batch-norm sums");
                return DL4JB200_ERR_INVALID_ARG;
            }
            if (pair) return launch_persist_variant<BN, PS, false, 2, true>(ctx, gs * mult, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
            return launch_persist_variant<BN, STAGES, false, 2, false>(ctx, gs, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
        } else {
            dl4j_set_error("conv_tc_gemm_tma: backward batch-norm sums need 64- or 128-wide tiles");
            return DL4JB200_ERR_INVALID_ARG;
        }
    }
    if (g.stat_out) {  // batch-norm statistics in the epilogue: grid a multiple of n_tiles (see the kernel)
        const int gs = persist_stat_grid(slots, total, n_tiles);
        if (g.remap || gs <= 0 || n_valid % BN != 0 || (act & 0xff) != DL4JB200_ACT_IDENTITY || g.epi_add || g.epi_mask) {
            dl4j_set_error("conv_tc_gemm_tma: this launch cannot carry batch-norm statistics");
            return DL4JB200_ERR_INVALID_ARG;
        }
        if constexpr (BN >= 64) {
            if (pair) return launch_persist_variant<BN, PS, false, 1, true>(ctx, gs * mult, tmA, tmB, g, bias, out, n_valid, out_cs, act & 0xff, n_tiles, total);
        }
        return launch_persist_variant<BN, STAGES, false, 1, false>(ctx, gs, tmA, tmB, g, bias, out, n_valid, out_cs, act & 0xff, n_tiles, total);
    }
    if constexpr (BN >= 64) {
        if (pair) {
            if (g.remap) return launch_persist_variant<BN, PS, true, 0, true>(ctx, units * 2, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
            return launch_persist_variant<BN, PS, false, 0, true>(ctx, units * 2, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
        }
    }
    if (g.remap) return launch_persist_variant<BN, STAGES, true, 0, false>(ctx, units, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
    return launch_persist_variant<BN, STAGES, false, 0, false>(ctx, units, tmA, tmB, g, bias, out, n_valid, out_cs, act, n_tiles, total);
}

template <int BN, int STAGES>
static int launch_gemm_tma(dl4jb200_ctx *ctx, const CUtensorMap &tmA, const CUtensorMap &tmB, const TmaGemmGeom &g,
                           const float *bias, float *out, int n_valid, int out_cs, int act) {
    using L = SmemLayout<BN, STAGES>;
    auto kern = k_conv_tc_gemm_tma<BN, STAGES>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
        configured[ctx->device & 15] = true;
    }
    dim3 grid(dl4j_ceil_div(g.M, BLOCK_M), dl4j_ceil_div(out_cs, BN));
    kern<<<grid, THREADS, L::BYTES, ctx->stream>>>(tmA, tmB, bias, out, g, n_valid, out_cs, act, g.num_k, (size_t)0);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// split-K form: `splits` CTAs per tile write raw partials (row pitch out_cs, rows padded to whole tiles) into scratch slot 4,
// k_splitk_finalize folds them into out
template <int BN, int STAGES>
static int launch_gemm_tma_splitk(dl4jb200_ctx *ctx, const CUtensorMap &tmA, const CUtensorMap &tmB, const TmaGemmGeom &g,
                                  const float *bias, float *out, int n_valid, int out_cs, int act, int splits) {
    using L = SmemLayout<BN, STAGES>;
    auto kern = k_conv_tc_gemm_tma<BN, STAGES>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
        configured[ctx->device & 15] = true;
    }
    const int m_tiles = dl4j_ceil_div(g.M, BLOCK_M), kps = dl4j_ceil_div(g.num_k, splits);
    splits = dl4j_ceil_div(g.num_k, kps);  // no empty split
    const size_t split_stride = (size_t)m_tiles * BLOCK_M * out_cs;
    float *partial = (float *)dl4j_workspace(ctx, 4, (size_t)splits * split_stride * sizeof(float));
    if (!partial) return DL4JB200_ERR_CUDA;
    dim3 grid(m_tiles, dl4j_ceil_div(out_cs, BN), splits);
    // n_valid = out_cs for the partial pass: every column of the tile is stored (the padding columns hold zeros)
    DL4J_CUDA_TRY(dl4j_launch_pdl(kern, grid, dim3(THREADS), L::BYTES, ctx->stream, tmA, tmB, (const float *)nullptr, partial, g, out_cs, out_cs,
                                  (int)DL4JB200_ACT_IDENTITY, kps, split_stride));
    DL4J_CHECK_LAUNCH();
    const size_t items = (size_t)g.M * (out_cs / 4);
    const size_t blocks = (items + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    DL4J_CUDA_TRY(dl4j_launch_pdl(k_splitk_finalize, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, ctx->stream, (const float *)partial,
                                  out, bias, g.epi_add, g.epi_mask, g.M, n_valid, out_cs, splits, split_stride, act));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// ------------------------------------------------------------------------------------------------
// TMA-fed backprop-filter. Per 32-pixel K-chunk: 4 tiled TMA loads bring delta[32 px, 128 co] (one 32-channel MN atom
// each) and BN/32 im2col-mode loads bring the 32-pixel x 32-channel slabs of the taps this CTA's k-columns cover.
// TMA lays a 32-row x 128-byte box out as 8 consecutive 4-row / 512-byte groups with SWIZZLE_128B_ATOM_32B, which is
// exactly the MN-major UMMA atom: LBO (next 32-float MN atom) = 4096 B, SBO (next 4-row group) = 512 B.

template <int BN, int STAGES>
__global__ void __launch_bounds__(THREADS) k_conv_tc_wgrad_tma(const __grid_constant__ CUtensorMap tmD,
                                                               const __grid_constant__ CUtensorMap tmX,
                                                               float *__restrict__ partial, TmaWgradGeom g, int rows_p,
                                                               int m_per_split, int k_tiles, float *__restrict__ db_partial) {
    using L = SmemLayout<BN, STAGES>;
    constexpr int NB_ATOMS = BN / 32;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + STAGES * L::A_BYTES, bar_base = base + L::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t acc_bar = bar_base + 8u * (2 * STAGES), tmem_slot = bar_base + 8u * (2 * STAGES + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kt = blockIdx.x % k_tiles, ct = blockIdx.x / k_tiles;
    const int kcol0 = kt * BN, co0 = ct * BLOCK_M;
    const int ms = blockIdx.y * m_per_split, me = min(g.M, ms + m_per_split);
    const int num_k = (me - ms + BLOCK_K - 1) / BLOCK_K;
    // Bias gradient db[co] = sum over pixels of delta (ConvolutionLayer.java:257) as one more accumulator block: the CTAs
    // of k-tile 0 multiply the same delta operand with an all-ones B tile (N = 16, the minimum for M = 128) into TMEM
    // columns [BN, BN+16). 1 KB of ones serves every K-step; being constant, its swizzle does not matter.
    const bool do_db = db_partial != nullptr && kt == 0;
    const uint32_t ones_base = base + L::BAR_OFF + 1024;
    if (do_db) {
        for (int i = tid; i < 256; i += THREADS) asm volatile("st.shared.f32 [%0], %1;" ::"r"(ones_base + 4u * i), "f"(1.0f) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }

    if (tid == 0) {
        tma::prefetch_desc(&tmD);
        tma::prefetch_desc(&tmX);
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) tmem_alloc<2 * BN>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_d;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_d) : "r"(tmem_slot));
    pdl_trigger();  // after the TMEM allocation, see k_conv_tc_gemm_tma_persist
    pdl_wait();

    if (warp < 4) {
        if (tid == 0) {
            // taps / channel offsets of this CTA's k-column chunks are fixed for the whole kernel
            int c0[NB_ATOMS];
            uint16_t offw[NB_ATOMS], offh[NB_ATOMS];
            bool live[NB_ATOMS];
#pragma unroll
            for (int a = 0; a < NB_ATOMS; a++) {
                const int k = kcol0 + a * 32;
                live[a] = k < g.Kp;
                uint32_t tap, cc, kh, kw;
                g.div_Cg.divmod((uint32_t)k, tap, cc);
                g.div_kW.divmod(tap, kh, kw);
                c0[a] = (int)cc; offw[a] = (uint16_t)(kw * g.dW); offh[a] = (uint16_t)(kh * g.dH);
            }
            for (int it = 0; it < num_k; ++it) {
                const int s = it % STAGES;
                mbar_wait(empty_bar(s), ((it / STAGES) & 1) ^ 1);
                const int mrow0 = ms + it * BLOCK_K;
                uint32_t img, r, oy, ox;
                g.div_HWm.divmod((uint32_t)mrow0, img, r);
                g.div_Wm.divmod(r, oy, ox);
                const int w0 = (int)ox * g.sW + g.lowW, h0 = (int)oy * g.sH + g.lowH;
                tma::expect_tx(full_bar(s), (uint32_t)(L::A_BYTES + L::B_BYTES));
                const uint32_t a_stage = a_base + s * L::A_BYTES, b_stage = b_base + s * L::B_BYTES;
#pragma unroll
                for (int a = 0; a < 4; a++) tma::load_2d(a_stage + a * 4096u, &tmD, full_bar(s), co0 + a * 32, mrow0);
#pragma unroll
                for (int a = 0; a < NB_ATOMS; a++) {
                    // a chunk past Kp still has to deliver its bytes: fetch channel block 0 far outside the image => zeros
                    if (live[a]) tma::load_im2col_4d(b_stage + a * 4096u, &tmX, full_bar(s), c0[a], w0, h0, (int)img, offw[a], offh[a]);
                    else tma::load_2d(b_stage + a * 4096u, &tmD, full_bar(s), 0, 0x3fffffff);
                }
            }
        }
        __syncwarp();
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int co = co0 + warp * 32 + lane;
        float *prow = partial + ((size_t)blockIdx.y * rows_p + co) * g.Kp;
        const uint32_t stage_smem = a_base + warp * 4096u;
#pragma unroll 1
        for (int cb = 0; cb < BN / 32; cb++) {
            float v[32];
            tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cb * 32), v);
            if (kcol0 + cb * 32 + 32 <= g.Kp) {
                store_tile_32x32(stage_smem, v, lane, partial + ((size_t)blockIdx.y * rows_p + co0 + warp * 32) * g.Kp + kcol0 + cb * 32,
                                 (size_t)g.Kp, rows_p - (co0 + warp * 32));
            } else if (co < rows_p) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int k = kcol0 + cb * 32 + q * 4;
                    if (k < g.Kp) *reinterpret_cast<float4 *>(prow + k) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
                }
            }
        }
        if (do_db) {
            float v[32];
            tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)BN, v);  // columns BN..BN+15 all hold the row's sum
            if (co < rows_p) db_partial[(size_t)blockIdx.y * rows_p + co] = v[0];
        }
    } else {
        constexpr uint32_t idesc = umma_idesc(BLOCK_M, BN, 1, 1);
        constexpr uint32_t idesc_db = umma_idesc(BLOCK_M, 16, 1, 1);
        const uint64_t od = umma_desc(ones_base, 4096, 512, 1);
        for (int it = 0; it < num_k; ++it) {
            const int s = it % STAGES;
            mbar_wait(full_bar(s), (it / STAGES) & 1);
            tc_fence_after();
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / 8; k++) {  // 8 pixels = two 4-row groups = 1024 B further down every atom
                    const uint64_t ad = umma_desc(a_base + s * L::A_BYTES + k * 1024, 4096, 512, 1);
                    const uint64_t bd = umma_desc(b_base + s * L::B_BYTES + k * 1024, 4096, 512, 1);
                    umma_tf32(tmem_d, ad, bd, idesc, (it | k) ? 1u : 0u);
                    if (do_db) umma_tf32(tmem_d + (uint32_t)BN, ad, od, idesc_db, (it | k) ? 1u : 0u);
                }
                umma_commit(empty_bar(s));
            }
            __syncwarp();
        }
        if (lane == 0) umma_commit(acc_bar);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc<2 * BN>(tmem_d);
}

template <int BN, int STAGES>
static int launch_wgrad_tma(dl4jb200_ctx *ctx, const CUtensorMap &tmD, const CUtensorMap &tmX, const TmaWgradGeom &g,
                            float *partial, int rows_p, int splits, int m_per_split, float *db_partial) {
    using L = SmemLayout<BN, STAGES>;
    constexpr int SMEM = L::BAR_OFF + 1024 + 1024 + 1024;  // stages | barriers | 1 KB ones tile | alignment slack
    auto kern = k_conv_tc_wgrad_tma<BN, STAGES>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        configured[ctx->device & 15] = true;
    }
    int k_tiles = dl4j_ceil_div(g.Kp, BN), c_tiles = dl4j_ceil_div(rows_p, BLOCK_M);
    dim3 grid(k_tiles * c_tiles, splits);
    DL4J_CUDA_TRY(dl4j_launch_pdl(kern, grid, dim3(THREADS), SMEM, ctx->stream, tmD, tmX, partial, g, rows_p, m_per_split, k_tiles, db_partial));
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

template <int BN, int STAGES, bool FAST>
static int launch_gemm_impl(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, const float *bias,
                            float *out, int n_valid, int out_cs, int act) {
    using L = SmemLayout<BN, STAGES>;
    auto kern = k_conv_tc_gemm<BN, STAGES, FAST>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
        configured[ctx->device & 15] = true;
    }
    dim3 grid(dl4j_ceil_div(g.M, BLOCK_M), dl4j_ceil_div(out_cs, BN));
    kern<<<grid, THREADS, L::BYTES, ctx->stream>>>(src, wpack, bias, out, g, n_valid, out_cs, act);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

template <int BN, int STAGES>
static int launch_gemm(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, const float *bias,
                       float *out, int n_valid, int out_cs, int act) {
    return g.fast ? launch_gemm_impl<BN, STAGES, true>(ctx, g, src, wpack, bias, out, n_valid, out_cs, act)
                  : launch_gemm_impl<BN, STAGES, false>(ctx, g, src, wpack, bias, out, n_valid, out_cs, act);
}

template <int BN, int STAGES>
static int launch_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs,
                        float *partial, int rows_p, int splits, int m_per_split) {
    using L = SmemLayout<BN, STAGES>;
    auto kern = k_conv_tc_wgrad<BN, STAGES>;
    static bool configured[16] = {false};
    if (!configured[ctx->device & 15]) {
        DL4J_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
        configured[ctx->device & 15] = true;
    }
    int k_tiles = dl4j_ceil_div(g.Kp, BN), c_tiles = dl4j_ceil_div(rows_p, BLOCK_M);
    dim3 grid(k_tiles * c_tiles, splits);
    kern<<<grid, THREADS, L::BYTES, ctx->stream>>>(x, delta, partial, g, delta_cs, rows_p, m_per_split, k_tiles);
    DL4J_CHECK_LAUNCH();
    return DL4JB200_OK;
}

// k-columns per filter-gradient tile (256-wide tiles were measured 5-15 % slower on every ResNet-50 shape: one CTA per SM
// instead of two loses more overlap than the smaller operand traffic per MMA step wins)
static inline int wgrad_bn(const ConvGeom &g) { return g.Kp >= 128 ? 128 : (g.Kp >= 64 ? 64 : 32); }

}  // namespace tc

// TMA path applies when every K-chunk is 32 channels of one tap and the gather is a plain (strided) window walk.
bool conv_tc_gemm_can_tma(const ConvGeom &g) {
    // (a 1x1 window has a single tap, so a 32-channel chunk can never straddle two taps: any channel count of at least 32 that
    // keeps pixels 16-byte aligned will do -- the last chunk's channels beyond C are zero-filled by the TMA unit and meet the
    // zero columns [K, Kp) of the packed filter; dense layers with 500 / 1000 inputs or outputs take this path)
    if (g.taps == 1 ? (g.Cg % 4 != 0 || g.Cg < 32 || g.cs_src % 4 != 0) : (g.Cg % 32 != 0 || g.Kp != g.K)) return false;
    if (g.transposed && (g.sH != 1 || g.sW != 1)) return false;
    if (g.kH > 16 || g.kW > 16 || g.dH * (g.kH - 1) > 255 || g.dW * (g.kW - 1) > 255) return false;
    const int lowH = g.transposed ? g.pT - (g.kH - 1) * g.dH : -g.pT, lowW = g.transposed ? g.pL - (g.kW - 1) * g.dW : -g.pL;
    const int sH = g.transposed ? 1 : g.sH, sW = g.transposed ? 1 : g.sW;
    const int upH = lowH + (g.Hm - 1) * sH + 1 - g.Hs, upW = lowW + (g.Wm - 1) * sW + 1 - g.Ws;
    if (lowH < -128 || lowH > 127 || lowW < -128 || lowW > 127 || upH < -128 || upH > 127 || upW < -128 || upW > 127) return false;
    if (sH > 8 || sW > 8) return false;
    return tma_driver_ready();
}

// N-tile width of the TMA GEMM: 256-wide tiles (one UMMA of N=256, the whole TMEM as a double-buffered accumulator) when the
// layer is wide and there are still at least ~2 tiles per SM
static int tma_gemm_bn(dl4jb200_ctx *ctx, const ConvGeom &g, int n_rows_packed, int out_cs, bool aligned_bias) {
    // (the backward batch-norm sums keep five per-column values per lane and chunk in registers: 128-wide tiles at most)
    const bool can256 = aligned_bias && out_cs >= 256 && n_rows_packed % 256 == 0 && g.Kp >= 256 && !g.bst_x;
    const long long m_tiles = dl4j_ceil_div(g.M, tc::BLOCK_M);
    const int narrow = out_cs > 64 ? 128 : (out_cs > 32 ? 64 : 32);
    if (can256 && g.Kp >= 512) {
        // long-K layers are bound by the per-SM TMA fill: a K-step costs ~(16 KB of A + bn * 128 B of B), and the launch
        // takes ceil(tiles / #SMs) rounds of tiles. Pick the width with the smaller product (e.g. 98 M tiles x N = 256:
        // one round of 256-wide tiles, 48 KB per step, beats two rounds of 128-wide ones at 32 KB).
        auto cost = [&](int bn) {
            const long long tiles = m_tiles * dl4j_ceil_div(out_cs, bn);
            return (double)((tiles + ctx->sm_count - 1) / ctx->sm_count) * (16.0 + bn / 8.0);
        };
        return cost(256) < cost(narrow) ? 256 : narrow;
    }
    const bool wide = can256 && m_tiles * dl4j_ceil_div(out_cs, 256) >= 2LL * ctx->sm_count;
    return wide ? 256 : narrow;
}

// 256-row tiles on CTA pairs (see the cta_group::2 helpers): wherever the persistent kernel runs with tiles of at least 64
// columns and the layer has two or more 128-row tiles
static bool tma_gemm_pair(dl4jb200_ctx *ctx, const ConvGeom &g, int bn, bool aligned_bias) {
    return ctx->pair_tiles && aligned_bias && bn >= 64 && g.M > tc::BLOCK_M && ctx->sm_count >= 2;
}

// partial rows a statistics-carrying launch writes: tiles / grid follow launch_gemm_tma_persist
static int tma_gemm_stat_rows(dl4jb200_ctx *ctx, const ConvGeom &g, int bn, int out_cs) {
    const bool pair = tma_gemm_pair(ctx, g, bn, true);
    const int n_tiles = out_cs / bn, total = dl4j_ceil_div(g.M, pair ? 2 * tc::BLOCK_M : tc::BLOCK_M) * n_tiles;
    const int grid = tc::persist_stat_grid(pair ? ctx->sm_count / 2 : ctx->sm_count, total, n_tiles) * (pair ? 2 : 1);
    return tc::persist_stat_rows(grid, n_tiles, bn);
}

// partial rows a statistics-carrying forward launch of this geometry writes (0: the launch cannot carry statistics)
int conv_tc_stat_rows(dl4jb200_ctx *ctx, const ConvGeom &g, int n_rows_packed, int n_valid, int out_cs) {
    if (!g.use_tma || g.remap) return 0;
    const int bn = tma_gemm_bn(ctx, g, n_rows_packed, out_cs, true);
    if (n_valid % bn != 0 || out_cs != n_valid) return 0;
    return tma_gemm_stat_rows(ctx, g, bn, out_cs);
}

// partial rows a backprop-data launch carrying the backward batch-norm sums writes (0: it cannot carry them)
int conv_tc_bst_rows(dl4jb200_ctx *ctx, const ConvGeom &g0, int n_rows_packed, int n_valid, int out_cs) {
    if (!g0.use_tma || g0.remap) return 0;
    ConvGeom g = g0;
    g.bst_x = reinterpret_cast<const float *>(16);  // any non-null value: selects the narrow tile widths
    const int bn = tma_gemm_bn(ctx, g, n_rows_packed, out_cs, true);
    if (bn < 64 || n_valid % bn != 0 || out_cs != n_valid) return 0;
    return tma_gemm_stat_rows(ctx, g, bn, out_cs);
}

static int conv_tc_gemm_tma(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, int n_rows_packed,
                            const float *bias, float *out, int n_valid, int out_cs, int act) {
    // backprop-data runs as a forward walk over delta with flipped taps (the filter is packed flipped, conv.cu)
    const int lowH = g.transposed ? g.pT - (g.kH - 1) * g.dH : -g.pT, lowW = g.transposed ? g.pL - (g.kW - 1) * g.dW : -g.pL;
    const int sH = g.transposed ? 1 : g.sH, sW = g.transposed ? 1 : g.sW;
    alignas(64) CUtensorMap tmA, tmB;
    const bool aligned_bias = !bias || ((uintptr_t)bias & 15) == 0;
    int rc = tma_make_im2col(&tmA, src, g.B, g.Hs, g.Ws, g.Cg, g.cs_src, lowW, lowH, g.Wm, g.Hm, sW, sH, tc::BLOCK_M,
                             CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    const int bn = tma_gemm_bn(ctx, g, n_rows_packed, out_cs, aligned_bias);
    // A handful of tiles and a long K (the dense layers: M = minibatch, i.e. one or two row tiles): split K over enough CTAs to
    // fill the machine and fold the partial accumulators afterwards. Otherwise 4-32 CTAs walk hundreds of K-chunks one TMA
    // round trip at a time (VGG16's 25088 -> 1000 output layer: 247 us for a 100 MB weight read).
    int splits = 0;
    if (!g.stat_out && !g.remap && out_cs % 4 == 0 && bn <= 128 && dl4j_ceil_div(g.M, tc::BLOCK_M) <= 2 && g.Kp / tc::BLOCK_K >= 32) {
        const long long tiles = (long long)dl4j_ceil_div(g.M, tc::BLOCK_M) * dl4j_ceil_div(out_cs, bn);
        splits = (int)(ctx->sm_count / tiles);
        if (splits > g.Kp / tc::BLOCK_K / 8) splits = g.Kp / tc::BLOCK_K / 8;  // at least eight K-chunks per split
        if (tiles * 4 > ctx->sm_count || splits < 2) splits = 0;
    }
    const bool pair = !splits && tma_gemm_pair(ctx, g, bn, aligned_bias);  // each CTA of a pair fetches half of the filter tile
    rc = tma_make_2d(&tmB, wpack, (uint64_t)n_rows_packed, (uint64_t)g.Kp, (uint64_t)g.Kp, (uint32_t)(pair ? bn / 2 : bn), CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    TmaGemmGeom t;
    t.M = g.M; t.num_k = g.Kp / tc::BLOCK_K; t.kW = g.kW; t.dH = g.dH; t.dW = g.dW; t.sH = sH; t.sW = sW; t.lowH = lowH; t.lowW = lowW;
    t.div_Wm = g.div_Wm; t.div_HWm = g.div_HWm; t.div_Cg = g.div_Cg; t.div_kW = g.div_kW;
    t.remap = g.remap; t.rm_sy = g.rm_sy; t.rm_sx = g.rm_sx; t.rm_H = g.rm_H; t.rm_W = g.rm_W;
    t.epi_add = g.epi_add; t.epi_mask = g.epi_mask;
    t.stat_out = g.stat_out; t.stat_pivot = g.stat_pivot;
    t.bst_x = g.bst_x; t.bst_mean = g.bst_mean; t.bst_std = g.bst_std; t.bst_gamma = g.bst_gamma; t.bst_beta = g.bst_beta;
    t.bst_relu = g.bst_relu;
    if (g.stat_out && !aligned_bias) { dl4j_set_error("conv_tc_gemm_tma: statistics need a 16-byte aligned bias"); return DL4JB200_ERR_INVALID_ARG; }
    if (g.bst_x && (!g.stat_out || g.remap || bias)) {
        dl4j_set_error("conv_tc_gemm_tma: backward batch-norm sums need a stat buffer, no bias and no row scatter");
        return DL4JB200_ERR_INVALID_ARG;
    }
    if ((g.epi_add || g.epi_mask) && (g.remap || !aligned_bias)) {
        dl4j_set_error("conv_tc_gemm_tma: fused epilogue needs the unscattered persistent kernel");
        return DL4JB200_ERR_INVALID_ARG;
    }
    if (g.remap && !aligned_bias) { dl4j_set_error("conv_tc_gemm_tma: row scatter needs the persistent kernel"); return DL4JB200_ERR_INVALID_ARG; }
    if (g.remap) {
        // Everything but the scattered rows is zero. When the stride is 1 or 2 per axis and every tile is a whole tile the
        // epilogue writes those zeros itself (each sample point owns its 2x2 cell); otherwise the tensor is cleared first.
        t.rm_zero = !g.rm_keep && g.rm_sy <= 2 && g.rm_sx <= 2 && n_valid % bn == 0 && out_cs == n_valid && g.rm_sx * (g.Wm - 1) + g.rm_sx >= g.rm_W &&
                    g.rm_sy * (g.Hm - 1) + g.rm_sy >= g.rm_H;
        if (!t.rm_zero && !g.rm_keep) DL4J_CUDA_TRY(cudaMemsetAsync(out, 0, (size_t)g.B * g.rm_H * g.rm_W * out_cs * sizeof(float), ctx->stream));
    }
    if (splits) {
        if (bn == 128) return tc::launch_gemm_tma_splitk<128, 3>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, splits);
        if (bn == 64) return tc::launch_gemm_tma_splitk<64, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, splits);
        return tc::launch_gemm_tma_splitk<32, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, splits);
    }
    if (bn == 256) return tc::launch_gemm_tma_persist<256, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, pair);
    if (!aligned_bias) {  // the persistent epilogue reads the bias with 128-bit loads
        if (bn == 128) return tc::launch_gemm_tma<128, 3>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
        if (bn == 64) return tc::launch_gemm_tma<64, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
        return tc::launch_gemm_tma<32, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
    }
    if (bn == 128) return tc::launch_gemm_tma_persist<128, 5>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, pair);
    if (bn == 64) return tc::launch_gemm_tma_persist<64, 6>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act, pair);
    return tc::launch_gemm_tma_persist<32, 6>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
}

int conv_tc_gemm(dl4jb200_ctx *ctx, const ConvGeom &g, const float *src, const float *wpack, int n_rows_packed,
                 const float *bias, float *out, int n_valid, int out_cs, int act) {
    if (g.use_tma) return conv_tc_gemm_tma(ctx, g, src, wpack, n_rows_packed, bias, out, n_valid, out_cs, act);
    if (out_cs > 64) return tc::launch_gemm<128, 3>(ctx, g, src, wpack, bias, out, n_valid, out_cs, act);
    if (out_cs > 32) return tc::launch_gemm<64, 4>(ctx, g, src, wpack, bias, out, n_valid, out_cs, act);
    return tc::launch_gemm<32, 4>(ctx, g, src, wpack, bias, out, n_valid, out_cs, act);
}

// number of pixel-range splits: enough CTAs for ~2 waves, every split non-empty
int conv_tc_wgrad_splits(dl4jb200_ctx *ctx, const ConvGeom &g, int c_out) {
    int bn = tc::wgrad_bn(g);
    int tiles = dl4j_ceil_div(g.Kp, bn) * dl4j_ceil_div(c_out, tc::BLOCK_M);
    int max_splits = dl4j_ceil_div(g.M, 8 * tc::BLOCK_K);  // at least 8 k-chunks per split
    if (max_splits < 1) max_splits = 1;
    // Two CTAs of this kernel fit an SM: 2 * #SMs slots per round. A launch of tiles * s CTAs takes ceil(tiles*s / slots)
    // rounds of ceil(chunks / s) K-chunks each; take the s with the smallest product (ties: fewer splits = less partial
    // traffic for the finalize). Rounding 2*#SMs / tiles UP, as before, put most layers a few CTAs past one full round.
    const long long slots = 2LL * ctx->sm_count, chunks = dl4j_ceil_div(g.M, tc::BLOCK_K);
    int want = 1;
    long long best = -1;
    const int s_hi = (int)(2 * slots / tiles) + 1;
    for (int sp = 1; sp <= max_splits && sp <= s_hi; sp++) {
        const long long rounds = ((long long)tiles * sp + slots - 1) / slots, cost = rounds * ((chunks + sp - 1) / sp);
        if (best < 0 || cost < best) { best = cost; want = sp; }
    }
    int mps = dl4j_round_up(dl4j_ceil_div(g.M, want), tc::BLOCK_K);
    return dl4j_ceil_div(g.M, mps);
}

int conv_tc_wgrad(dl4jb200_ctx *ctx, const ConvGeom &g, const float *x, const float *delta, int delta_cs, int c_out,
                  float *partial, int rows_p, int splits, float *db_partial) {
    (void)c_out;
    int mps = dl4j_round_up(dl4j_ceil_div(g.M, splits), tc::BLOCK_K);
    if (dl4j_ceil_div(g.M, mps) != splits) {
        dl4j_set_error("conv_tc_wgrad: split count %d leaves an empty split (M=%d)", splits, g.M);
        return DL4JB200_ERR_INVALID_ARG;
    }
    if (g.use_tma) {
        alignas(64) CUtensorMap tmD, tmX;
        int rc = tma_make_2d(&tmD, delta, (uint64_t)g.M, (uint64_t)delta_cs, (uint64_t)delta_cs, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (rc) return rc;
        rc = tma_make_im2col(&tmX, x, g.B, g.Hs, g.Ws, g.Cg, g.cs_src, -g.pL, -g.pT, g.Wm, g.Hm, g.sW, g.sH, tc::BLOCK_K,
                             CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (rc) return rc;
        TmaWgradGeom t;
        t.M = g.M; t.kW = g.kW; t.dH = g.dH; t.dW = g.dW; t.sH = g.sH; t.sW = g.sW; t.lowH = -g.pT; t.lowW = -g.pL; t.Kp = g.Kp;
        t.div_Wm = g.div_Wm; t.div_HWm = g.div_HWm; t.div_Cg = g.div_Cg; t.div_kW = g.div_kW;
        switch (tc::wgrad_bn(g)) {
        case 128: return tc::launch_wgrad_tma<128, 3>(ctx, tmD, tmX, t, partial, rows_p, splits, mps, db_partial);
        case 64: return tc::launch_wgrad_tma<64, 4>(ctx, tmD, tmX, t, partial, rows_p, splits, mps, db_partial);
        default: return tc::launch_wgrad_tma<32, 4>(ctx, tmD, tmX, t, partial, rows_p, splits, mps, db_partial);
        }
    }
    if (db_partial) { dl4j_set_error("conv_tc_wgrad: the bias-gradient column needs the TMA kernel"); return DL4JB200_ERR_INVALID_ARG; }
    switch (tc::wgrad_bn(g)) {
    case 128: return tc::launch_wgrad<128, 3>(ctx, g, x, delta, delta_cs, partial, rows_p, splits, mps);
    case 64: return tc::launch_wgrad<64, 4>(ctx, g, x, delta, delta_cs, partial, rows_p, splits, mps);
    default: return tc::launch_wgrad<32, 4>(ctx, g, x, delta, delta_cs, partial, rows_p, splits, mps);
    }
}

// ---- raw launchers for callers that build their own tensor maps (conv_stem.cu)
int conv_tc_gemm_tma_raw(dl4jb200_ctx *ctx, const CUtensorMap &tmA, const CUtensorMap &tmB, const TmaGemmGeom &t, const float *bias,
                         float *out, int n_valid, int out_cs, int act, int bn) {
    if (bias && ((uintptr_t)bias & 15)) {
        dl4j_set_error("conv_tc_gemm_tma_raw: bias must be 16-byte aligned");
        return DL4JB200_NOT_SUPPORTED;
    }
    if (bn == 256) return tc::launch_gemm_tma_persist<256, 4>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
    if (bn == 128) return tc::launch_gemm_tma_persist<128, 5>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
    if (bn == 64 && t.num_k == 7 && out_cs <= 64)  // 7-row stem filters: one more stage lets the filter stay resident
        return tc::launch_gemm_tma_persist<64, 7>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
    if (bn == 64) return tc::launch_gemm_tma_persist<64, 6>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
    return tc::launch_gemm_tma_persist<32, 6>(ctx, tmA, tmB, t, bias, out, n_valid, out_cs, act);
}

int conv_tc_wgrad_tma_raw(dl4jb200_ctx *ctx, const CUtensorMap &tmD, const CUtensorMap &tmX, const TmaWgradGeom &t, float *partial,
                          int rows_p, int splits, int m_per_split, float *db_partial) {
    const int bn = t.Kp >= 128 ? 128 : (t.Kp >= 64 ? 64 : 32);
    switch (bn) {
    case 128: return tc::launch_wgrad_tma<128, 3>(ctx, tmD, tmX, t, partial, rows_p, splits, m_per_split, db_partial);
    case 64: return tc::launch_wgrad_tma<64, 4>(ctx, tmD, tmX, t, partial, rows_p, splits, m_per_split, db_partial);
    default: return tc::launch_wgrad_tma<32, 4>(ctx, tmD, tmX, t, partial, rows_p, splits, m_per_split, db_partial);
    }
}
