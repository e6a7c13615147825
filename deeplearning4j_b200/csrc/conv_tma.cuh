// TMA plumbing for the convolution kernels: tensor-map construction on the host (driver entry points are fetched
// through the runtime, so the library does not link libcuda) and the cp.async.bulk.tensor PTX wrappers.
//
// Activations are NHWC with pixel stride cs, i.e. a rank-4 tensor {C, W, H, N} (innermost first). The A operand of
// the implicit GEMM is loaded in TMA *im2col* mode: one instruction fetches 128 consecutive output pixels
// (traversing w -> h -> n inside the padded bounding box, with the convolution stride as traversal stride) x 32
// channels of one filter tap (tap offset passed per instruction), zero-filling whatever falls outside the image.
#pragma once
#include <cuda.h>

#include "conv_common.cuh"

typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                        const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*PFN_tmapEncodeIm2col)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                         const cuuint64_t *, const int *, const int *, cuuint32_t, cuuint32_t,
                                         const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                         CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// returns false (and sets the error string) if the driver does not expose the encoders
bool tma_driver_ready();
// rows x cols fp32 row-major matrix (row stride ld floats); box = box_rows x 32 floats, given swizzle
int tma_make_2d(CUtensorMap *out, const float *base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                CUtensorMapSwizzle swizzle);
// NHWC activation tensor for im2col loads: `pixels` base pixels x 32 channels per instruction.
// lower_w/h: bounding-box lower corner (= -pad for a forward convolution); out_w/h: number of base pixels per row/col;
// stride_w/h: traversal stride.
int tma_make_im2col(CUtensorMap *out, const float *base, int N, int H, int W, int C, int cs, int lower_w, int lower_h,
                    int out_w, int out_h, int stride_w, int stride_h, uint32_t pixels, CUtensorMapSwizzle swizzle);

// same, with explicit pixel / row / image pitches (floats): lets a caller describe a tensor whose "pixels" overlap
// (pixel pitch < channel extent), which is how the 3-channel stem convolution is fed (conv_stem.cu)
int tma_make_im2col_pitched(CUtensorMap *out, const float *base, int N, int H, int W, int C, long long pix_pitch,
                            long long row_pitch, long long img_pitch, int lower_w, int lower_h, int out_w, int out_h,
                            int stride_w, int stride_h, uint32_t pixels, CUtensorMapSwizzle swizzle);

namespace tcgeom {
struct TmaGemmGeom {
    int M, num_k, kW, dH, dW, sH, sW, lowH, lowW;
    FastDiv div_Wm, div_HWm, div_Cg, div_kW;
    // optional output-row scatter: GEMM row (b, oy, ox) is stored at pixel (b, oy*rm_sy, ox*rm_sx) of an rm_H x rm_W image
    // (backprop-data of a strided 1x1 convolution computed on the compact pixel set; the rest of dx is pre-zeroed)
    int remap, rm_sy, rm_sx, rm_H, rm_W;
    int rm_zero = 0;  // the epilogue also writes the zeros of the skipped pixels (strides <= 2), no memset before the launch
    // optional fused epilogue of a backprop-data GEMM, both laid out exactly like `out`:
    //   out = (acc + epi_add) masked by epi_mask > 0
    // i.e. the ComputationGraph epsilon fan-in sum followed by the ReLU backprop of the vertex that owns `out`
    // (epi_add may alias out: every element is read and written by the same thread)
    const float *epi_add = nullptr, *epi_mask = nullptr;
    // optional batch-norm statistics of the output (forward, identity activation): per-warp partial rows
    //   stat_out[row][0][c] = sum (z - stat_pivot[c]),  stat_out[row][1][c] = sum (z - stat_pivot[c])^2
    // rows = gridDim.x / n_tiles (x 4 for 32-wide tiles), row pitch 2 * out_cs floats (the layout k_bn_finalize folds)
    float *stat_out = nullptr;
    const float *stat_pivot = nullptr;
    // optional BACKWARD batch-norm sums (backprop-data GEMM whose output e is the epsilon of a BatchNormalization, or of the
    // Add vertex a BatchNormalization is folded into): while the tile is on chip the epilogue also reads that layer's input
    // x (laid out like `out`) and leaves, in stat_out's row layout,
    //   stat_out[row][0][c] = sum e',   stat_out[row][1][c] = sum e' * (x - mean[c]),   e' = e gated by the layer's fused ReLU
    // (gate recomputed as (x-mean)*gamma/std + beta > 0, the expression k_bn_partial_nhwc / k_bn_apply_nhwc use) -- the
    // sums BatchNormalization.backpropGradient needs (dBeta, dGamma*std; BatchNormalization.java:172-196), so the
    // separate partial-sum pass over (x, e) disappears. bst_x != nullptr selects this mode.
    const float *bst_x = nullptr, *bst_mean = nullptr, *bst_std = nullptr, *bst_gamma = nullptr, *bst_beta = nullptr;
    int bst_relu = 0;
};
struct TmaWgradGeom {
    int M, kW, dH, dW, sH, sW, lowH, lowW, Kp;
    FastDiv div_Wm, div_HWm, div_Cg, div_kW;
};
}  // namespace tcgeom
int conv_tc_gemm_tma_raw(dl4jb200_ctx *ctx, const CUtensorMap &tmA, const CUtensorMap &tmB, const tcgeom::TmaGemmGeom &t,
                         const float *bias, float *out, int n_valid, int out_cs, int act, int bn);
int conv_tc_wgrad_tma_raw(dl4jb200_ctx *ctx, const CUtensorMap &tmD, const CUtensorMap &tmX, const tcgeom::TmaWgradGeom &t,
                          float *partial, int rows_p, int splits, int m_per_split, float *db_partial = nullptr);

#ifdef __CUDACC__
namespace tma {
__device__ __forceinline__ void prefetch_desc(const CUtensorMap *m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void load_2d(uint32_t dst, const CUtensorMap *m, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void load_im2col_4d(uint32_t dst, const CUtensorMap *m, uint32_t bar, int c, int w, int h, int n,
                                               uint16_t off_w, uint16_t off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
        : "memory");
}
// CTA-pair forms (cta_group::2): the data lands in THIS CTA's shared memory, the completion is signalled on an mbarrier
// given as a shared::cluster address (the pair leader's full barrier). Plain unicast loads of this GPU's own tensors.
__device__ __forceinline__ void load_2d_pair(uint32_t dst, const CUtensorMap *m, uint32_t cluster_bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(m)), "r"(cluster_bar), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void load_im2col_4d_pair(uint32_t dst, const CUtensorMap *m, uint32_t cluster_bar, int c, int w, int h, int n,
                                                    uint16_t off_w, uint16_t off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(cluster_bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
        : "memory");
}
}  // namespace tma
#endif
