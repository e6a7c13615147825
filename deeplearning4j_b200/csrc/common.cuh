// Shared host/device plumbing of libdl4jb200: context, error reporting, scratch workspace.
#pragma once
#include <utility>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <unordered_map>

#include "../../include/dl4jb200.h"

struct MirrorBuf {
    void *dev;
    size_t bytes;
    int state;  // DL4JB200_BUF_HOST_NEWER / SYNCED / DEVICE_NEWER: which side holds the current contents
};

#define DL4J_WS_SLOTS 5
struct FilterRegistry;
void conv_filters_free(struct dl4jb200_ctx *ctx);
struct dl4jb200_ctx {
    int device;
    cudaStream_t stream;
    bool own_stream;
    int sm_count;
    int round_act;  // DL4JB200_OPT_TF32_ACTIVATIONS: activation / epsilon producers store values rounded to TF32
    int pair_tiles = 0;  // DL4JB200_OPT_PAIR_TILES: 256-row tiles on CTA pairs (tcgen05 cta_group::2) where they apply
    // scratch, grown on demand (never shrunk); slot 0: kernel partials / packed filters, slots 1-3: NHWC staging,
    // slot 4: split-K partial accumulators of the dense-layer GEMMs
    void *ws[DL4J_WS_SLOTS];
    size_t ws_bytes[DL4J_WS_SLOTS];
    // A captured CUDA graph bakes scratch addresses (and tensor maps built on them) into its nodes: once frozen, a
    // request that would have to grow a region fails instead of freeing memory the graph still references.
    bool ws_frozen = false;
    unsigned long long ws_generation = 0;  // bumped whenever a region is (re)allocated
    std::unordered_map<const void *, MirrorBuf> mirrors;
    std::unordered_multimap<size_t, void *> mirror_pool;  // released device buffers by size, reused by the next acquire
    unsigned long long h2d_bytes = 0, d2h_bytes = 0;      // bytes this context moved across the host link
    struct FilterRegistry *filters = nullptr;  // pre-packed convolution filters (conv.cu); freed by conv_filters_free
};

void dl4j_set_error(const char *fmt, ...);
// Returns a scratch region of at least `bytes` in slot `slot`; nullptr (and error set) on failure.
void *dl4j_workspace(dl4jb200_ctx *ctx, int slot, size_t bytes);

#define DL4J_CUDA_TRY(expr)                                                                       \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            dl4j_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return DL4JB200_ERR_CUDA;                                                             \
        }                                                                                         \
    } while (0)

// every kernel launch in the library is followed by this macro: it also feeds dl4jb200_launch_count()
extern unsigned long long g_dl4j_launches;
#define DL4J_CHECK_LAUNCH()                  \
    do {                                     \
        ++g_dl4j_launches;                   \
        DL4J_CUDA_TRY(cudaGetLastError());   \
    } while (0)

// Programmatic dependent launch (sm_90+). A kernel that starts with pdl_trigger(); pdl_wait(); may be launched with
// dl4j_launch_pdl: its CTAs are then scheduled while the preceding kernel of the stream drains (launch latency and
// prologue overlap the tail), and pdl_wait() returns only when that kernel has completed and its writes are visible.
// Both instructions are no-ops under an ordinary launch. Rule: no global memory access before pdl_wait().
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t dl4j_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                   Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
// the same, as thread-block clusters of cluster_x CTAs along x (gridDim.x must be a multiple of it)
template <typename... KArgs, typename... Args>
inline cudaError_t dl4j_launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster_x,
                                           Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = (unsigned)cluster_x;
    at[1].val.clusterDim.y = 1;
    at[1].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
#endif

#define DL4J_REQUIRE(cond, ...)            \
    do {                                   \
        if (!(cond)) {                     \
            dl4j_set_error(__VA_ARGS__);   \
            return DL4JB200_ERR_INVALID_ARG; \
        }                                  \
    } while (0)

static inline int dl4j_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline int dl4j_round_up(int a, int b) { return (a + b - 1) / b * b; }

// Unsigned division by a runtime constant via multiply-high (divisor < 2^31, dividend < 2^31).
struct FastDiv {
    uint32_t d, mul, shr;
    FastDiv() : d(1), mul(0), shr(0) {}
    explicit FastDiv(uint32_t div) : d(div) {
        if (div == 1) { mul = 0; shr = 0; return; }
        uint32_t l = 0;
        while ((1u << l) < div) l++;  // ceil(log2(d))
        uint64_t m = ((uint64_t)1 << (32 + l)) / div + 1;  // fits: result in [2^32, 2^33)
        mul = (uint32_t)(m - ((uint64_t)1 << 32));
        shr = l;
    }
    __host__ __device__ __forceinline__ uint32_t div(uint32_t n) const {
#ifdef __CUDA_ARCH__
        if (d == 1) return n;
        uint32_t t = __umulhi(n, mul);
        return (t + ((n - t) >> 1)) >> (shr - 1);
#else
        return n / d;
#endif
    }
    __host__ __device__ __forceinline__ void divmod(uint32_t n, uint32_t &q, uint32_t &r) const {
        q = div(n);
        r = n - q * d;
    }
};

// round-to-nearest into TF32 (10-bit mantissa); used when the context stores activations TF32-representable so that the
// tensor core's operand truncation becomes exact
__device__ __forceinline__ float dl4j_tf32(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}
__device__ __forceinline__ float dl4j_rnd(float v, int rnd) { return rnd ? dl4j_tf32(v) : v; }
__device__ __forceinline__ float4 dl4j_rnd4(float4 v, int rnd) {
    if (rnd) { v.x = dl4j_tf32(v.x); v.y = dl4j_tf32(v.y); v.z = dl4j_tf32(v.z); v.w = dl4j_tf32(v.w); }
    return v;
}

// activation codes shared by every epilogue (DL4JB200_ACT_*)
__device__ __forceinline__ float dl4j_act(float z, int kind) {
    switch (kind) {
    case DL4JB200_ACT_RELU: return z > 0.f ? z : 0.f;
    case DL4JB200_ACT_SIGMOID: return 1.f / (1.f + expf(-z));
    case DL4JB200_ACT_TANH: return tanhf(z);
    case DL4JB200_ACT_LEAKYRELU: return z > 0.f ? z : 0.01f * z;
    default: return z;
    }
}
__device__ __forceinline__ float dl4j_act_grad(float z, int kind) {
    switch (kind) {
    case DL4JB200_ACT_RELU: return z > 0.f ? 1.f : 0.f;
    case DL4JB200_ACT_SIGMOID: { float s = 1.f / (1.f + expf(-z)); return s * (1.f - s); }
    case DL4JB200_ACT_TANH: { float t = tanhf(z); return 1.f - t * t; }
    case DL4JB200_ACT_LEAKYRELU: return z > 0.f ? 1.f : 0.01f;
    default: return 1.f;
    }
}
// the same derivative written in terms of the activation OUTPUT a = act(z), for callers that kept only a
// (DL4JB200_ACT_FROM_OUTPUT): relu'(z) = [a > 0], sigmoid' = a(1-a), tanh' = 1-a^2, leakyrelu keeps the sign of z
__device__ __forceinline__ float dl4j_act_grad_out(float a, int kind) {
    switch (kind) {
    case DL4JB200_ACT_RELU: return a > 0.f ? 1.f : 0.f;
    case DL4JB200_ACT_SIGMOID: return a * (1.f - a);
    case DL4JB200_ACT_TANH: return 1.f - a * a;
    case DL4JB200_ACT_LEAKYRELU: return a > 0.f ? 1.f : 0.01f;
    default: return 1.f;
    }
}
