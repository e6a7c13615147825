// Host side of the TMA path: tensor-map encoders fetched from the driver through the runtime.
#include "conv_tma.cuh"

static PFN_tmapEncodeTiled g_encode_tiled = nullptr;
static PFN_tmapEncodeIm2col g_encode_im2col = nullptr;
static int g_tma_state = 0;  // 0 unknown, 1 ready, -1 unavailable

bool tma_driver_ready() {
    if (g_tma_state == 0) {
        void *f1 = nullptr, *f2 = nullptr;
        cudaDriverEntryPointQueryResult q1, q2;
        cudaError_t e1 = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f1, cudaEnableDefault, &q1);
        cudaError_t e2 = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f2, cudaEnableDefault, &q2);
        if (e1 == cudaSuccess && e2 == cudaSuccess && f1 && f2 && q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess) {
            g_encode_tiled = (PFN_tmapEncodeTiled)f1;
            g_encode_im2col = (PFN_tmapEncodeIm2col)f2;
            g_tma_state = 1;
        } else {
            (void)cudaGetLastError();
            g_tma_state = -1;
        }
    }
    return g_tma_state == 1;
}

int tma_make_2d(CUtensorMap *out, const float *base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                CUtensorMapSwizzle swizzle) {
    DL4J_REQUIRE(tma_driver_ready(), "TMA: driver does not expose cuTensorMapEncodeTiled");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode_tiled(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)base, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        dl4j_set_error("cuTensorMapEncodeTiled failed (%d): rows=%llu cols=%llu ld=%llu box_rows=%u", (int)r,
                       (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows);
        return DL4JB200_ERR_CUDA;
    }
    return DL4JB200_OK;
}

int tma_make_im2col(CUtensorMap *out, const float *base, int N, int H, int W, int C, int cs, int lower_w, int lower_h,
                    int out_w, int out_h, int stride_w, int stride_h, uint32_t pixels, CUtensorMapSwizzle swizzle) {
    return tma_make_im2col_pitched(out, base, N, H, W, C, cs, (long long)W * cs, (long long)H * W * cs, lower_w, lower_h, out_w,
                                   out_h, stride_w, stride_h, pixels, swizzle);
}

int tma_make_im2col_pitched(CUtensorMap *out, const float *base, int N, int H, int W, int C, long long pix_pitch,
                            long long row_pitch, long long img_pitch, int lower_w, int lower_h, int out_w, int out_h,
                            int stride_w, int stride_h, uint32_t pixels, CUtensorMapSwizzle swizzle) {
    DL4J_REQUIRE(tma_driver_ready(), "TMA: driver does not expose cuTensorMapEncodeIm2col");
    const int cs = (int)pix_pitch;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)pix_pitch * 4, (cuuint64_t)row_pitch * 4, (cuuint64_t)img_pitch * 4};
    // number of base pixels along an axis = (dim + upper - lower - 1) / stride + 1  ==> upper chosen for `out` pixels
    int lower[2] = {lower_w, lower_h};
    int upper[2] = {lower_w + (out_w - 1) * stride_w + 1 - W, lower_h + (out_h - 1) * stride_h + 1 - H};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)stride_h, 1};
    CUresult r = g_encode_im2col(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void *)base, dims, strides, lower, upper, 32, pixels,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        dl4j_set_error("cuTensorMapEncodeIm2col failed (%d): N=%d H=%d W=%d C=%d cs=%d lower=(%d,%d) upper=(%d,%d) stride=(%d,%d)",
                       (int)r, N, H, W, C, cs, lower[0], lower[1], upper[0], upper[1], stride_w, stride_h);
        return DL4JB200_ERR_CUDA;
    }
    return DL4JB200_OK;
}
