/*
 * dl4jb200.h -- C ABI of the B200-native back-end for Deeplearning4j's CNN layer helpers.
 *
 * This is the drop-in boundary: the four Java helper classes that DL4J's layers look up by
 * reflection each marshal their INDArray arguments and make exactly one of the
 * calls below (INTEGRATION.md shows the JavaCPP/JNI stubs). Paths cited are relative to
 * /root/reference/deeplearning4j-nn/src/main/java/org/deeplearning4j/ ("NN/") unless noted.
 *
 * Conventions
 *   - every function returns an int status: 0 OK; 1 NOT_SUPPORTED (Java helper returns null, the
 *     layer runs its built-in path: NN/nn/layers/convolution/ConvolutionLayer.java:197-199,372-374);
 *     <0 error (Java helper throws; conv/pool layers then fall back when cudnnAllowFallback,
 *     ConvolutionLayer.java:189-196). dl4jb200_last_error() describes the last failure of the
 *     calling thread.
 *   - tensors are fp32. Every data pointer is a DEVICE pointer valid on the context's device
 *     (obtained from dl4jb200_malloc, from the host-mirror registry, or from any other CUDA
 *     allocator such as torch) unless the parameter name ends in _host.
 *   - activations are rank 4, described by (B, C, H, W), a layout tag and, for NHWC, the pixel
 *     stride in floats ("cs", >= C, multiple of 4; channels [C,cs) are padding and hold zeros).
 *     NCHW is DL4J's layout ('c' order INDArray); NHWC is the resident layout the tensor-core
 *     kernels consume. Functions taking a layout accept either and produce the same layout.
 *   - work is enqueued on the context's stream and is asynchronous with respect to the host, as
 *     the cuDNN helper is on nd4j's stream (deeplearning4j-cuda/.../CudnnConvolutionHelper.java:476-477).
 *   - there is no CPU fallback anywhere behind this header.
 */
#ifndef DL4JB200_H
#define DL4JB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DL4JB200_OK 0
#define DL4JB200_NOT_SUPPORTED 1
#define DL4JB200_ERR_INVALID_ARG (-1)
#define DL4JB200_ERR_CUDA (-2)
#define DL4JB200_ERR_INVALID_INPUT (-3)  /* DL4JInvalidInputException, NN/util/ConvolutionUtils.java:136-167 */
#define DL4JB200_ERR_INVALID_CONFIG (-4) /* DL4JInvalidConfigException (Strict mode), ConvolutionUtils.java:169-211 */

/* NN/nn/conf/ConvolutionMode.java */
#define DL4JB200_MODE_STRICT 0
#define DL4JB200_MODE_TRUNCATE 1
#define DL4JB200_MODE_SAME 2

#define DL4JB200_NCHW 0
#define DL4JB200_NHWC 1

/* NN/nn/conf/layers/PoolingType.java */
#define DL4JB200_POOL_MAX 0
#define DL4JB200_POOL_AVG 1
#define DL4JB200_POOL_SUM 2
#define DL4JB200_POOL_PNORM 3

/* IActivation kinds the helper accepts (the cuDNN helper handles sigmoid/relu/tanh/identity,
 * deeplearning4j-cuda/.../CudnnConvolutionHelper.java:527-558; anything else => NOT_SUPPORTED). */
#define DL4JB200_ACT_IDENTITY 0
#define DL4JB200_ACT_RELU 1
#define DL4JB200_ACT_SIGMOID 2
#define DL4JB200_ACT_TANH 3
#define DL4JB200_ACT_LEAKYRELU 4
/* OR-ed into `act` of dl4jb200_act_bwd: the `z` argument holds the activation OUTPUT a = act(z) (what a caller that
 * activated in place still has); the derivative is then taken as relu: [a > 0], sigmoid: a(1-a), tanh: 1-a^2. */
#define DL4JB200_ACT_FROM_OUTPUT 0x200

/* convolution arithmetic: TF32 tensor cores with FP32 accumulate (throughput), or FP32 CUDA cores
 * (parity mode, plays the role AlgoMode.NO_WORKSPACE plays in the reference's tests). */
#define DL4JB200_CONV_ALGO_TC_TF32 0
#define DL4JB200_CONV_ALGO_FP32 1

/* IUpdater kinds on the flat parameter vector (NN/nn/updater/UpdaterBlock.java:115-193) */
#define DL4JB200_UPD_SGD 0
#define DL4JB200_UPD_NESTEROVS 1
#define DL4JB200_UPD_RMSPROP 2
#define DL4JB200_UPD_ADADELTA 3
#define DL4JB200_UPD_NOOP 4

typedef struct dl4jb200_ctx dl4jb200_ctx;

/* Geometry of one conv / pooling call. out_h/out_w/pad_* are what dl4jb200_out_size returns for the
 * layer's ConvolutionMode (pad_top/pad_left: ConvolutionLayer.java:148-150; the extra bottom/right cell
 * of Same mode is implied by out_h/out_w). For pooling c_out is ignored. */
typedef struct dl4jb200_conv_desc {
    int32_t batch, c_in, in_h, in_w;
    int32_t c_out, k_h, k_w;
    int32_t stride_h, stride_w;
    int32_t pad_top, pad_left;
    int32_t dil_h, dil_w;
    int32_t out_h, out_w;
} dl4jb200_conv_desc;

/* ------------------------------------------------------------------ runtime ------------------ */
const char *dl4jb200_last_error(void);
int dl4jb200_version(void);
/* number of CUDA kernels this library has launched in this process (diagnostics / bench accounting) */
unsigned long long dl4jb200_launch_count(void);

/* One context per (worker thread, device): ParallelWrapper pins each worker thread to one device for
 * life (deeplearning4j-scaleout/.../parallelism/ParallelWrapper.java:125-141). stream == NULL makes the
 * context create its own non-blocking stream; otherwise it enqueues on the caller's cudaStream_t. */
int dl4jb200_ctx_create(int device, void *stream, dl4jb200_ctx **out);
int dl4jb200_ctx_destroy(dl4jb200_ctx *ctx);
int dl4jb200_ctx_set_stream(dl4jb200_ctx *ctx, void *stream);
/* Options. DL4JB200_OPT_TF32_ACTIVATIONS (default 0): every kernel that PRODUCES a tensor a convolution will consume
 * (re-layout of the minibatch, BN / activation / add / pooling outputs forward and backward, softmax delta, and the
 * outputs of convolutions with a fused activation) stores it rounded-to-nearest into TF32. The tensor cores truncate
 * fp32 operands to TF32; with representable operands that truncation is exact, so the only TF32 error left is one
 * unbiased rounding per stored value (2^-12 relative) instead of a biased truncation per product. Meant to be
 * switched on together with DL4JB200_CONV_ALGO_TC_TF32; leave it off for FP32 parity runs. */
#define DL4JB200_OPT_TF32_ACTIVATIONS 1
/* DL4JB200_OPT_PAIR_TILES (default 0, or 1 when the environment holds DL4JB200_PAIR_TILES=1 at ctx_create): the tensor-core
 * convolution GEMMs run 256-row tiles on pairs of CTAs (two SMs of one TPC, tcgen05 cta_group::2) wherever the layer has at
 * least two 128-row tiles and 64 output columns, so that each SM streams only half of the filter tile. Convolution outputs
 * are bit-identical either way (same products, same accumulation order per output element); epilogue batch-norm statistics
 * are grouped per CTA and therefore differ in fp32 summation order. Off by default: measured on B200 it moves 25 % fewer
 * bytes into the SMs but is 0-10 % slower on every ResNet-50 shape (the layers are bound by L2 sector reads of the im2col
 * operand and by tile quantisation, not by per-SM fill) and 0-3 % faster on VGG16's (profiles/r02_pair_tiles.md). */
#define DL4JB200_OPT_PAIR_TILES 2
int dl4jb200_ctx_set_option(dl4jb200_ctx *ctx, int option, int value);
void *dl4jb200_ctx_stream(dl4jb200_ctx *ctx);
int dl4jb200_ctx_sync(dl4jb200_ctx *ctx);
/* bytes of internal scratch (split-K partials, BN partial sums, packed filters) currently held */
size_t dl4jb200_ctx_workspace_bytes(dl4jb200_ctx *ctx);
int dl4jb200_ctx_sm_count(dl4jb200_ctx *ctx);
/* The scratch regions grow on demand (cudaFree + cudaMalloc after draining the stream). A CUDA graph captured from calls
 * on this context has their addresses baked in, so a caller that captures freezes the scratch afterwards: while frozen, a
 * call that would need a larger region fails with ERR_CUDA and a clear message instead of freeing memory the graph still
 * uses. workspace_generation counts (re)allocations; equal values before and after mean every baked address is still live.
 * (Plays the role of nd4j's workspace scope checks, deeplearning4j-core/src/test/.../BaseDL4JTest.java:8-25.) */
int dl4jb200_ctx_workspace_freeze(dl4jb200_ctx *ctx, int frozen);
unsigned long long dl4jb200_ctx_workspace_generation(dl4jb200_ctx *ctx);

int dl4jb200_malloc(dl4jb200_ctx *ctx, size_t bytes, void **out);
int dl4jb200_free(dl4jb200_ctx *ctx, void *ptr);
int dl4jb200_memset(dl4jb200_ctx *ctx, void *dst, int value, size_t bytes);
int dl4jb200_memcpy_h2d(dl4jb200_ctx *ctx, void *dst, const void *src_host, size_t bytes);
int dl4jb200_memcpy_d2h(dl4jb200_ctx *ctx, void *dst_host, const void *src, size_t bytes);
int dl4jb200_memcpy_d2d(dl4jb200_ctx *ctx, void *dst, const void *src, size_t bytes);

/* Host-mirror registry: maps an INDArray's host address (INDArray.data().addressPointer() under
 * nd4j-native) to a device buffer owned by the context, with a state that says which side holds the current contents,
 * so that chained helper calls stay on the device and bytes cross PCIe only when the host side really wrote or reads
 * (replaces what the cuDNN helper gets from nd4j-cuda's AtomicAllocator.getPointer, CudnnConvolutionHelper.java:442-447).
 *   buf_acquire(READ)   an operand a kernel will read : registered if new; uploaded only if the host copy is newer
 *   buf_acquire(WRITE)  an operand a kernel will write: registered if new; becomes "device newer", nothing moves
 *   buf_mark_host_dirty the host wrote the array (a non-helper layer, the updater): next READ acquire uploads
 *   buf_sync_to_host    the host is about to read it : downloaded only if the device copy is newer (synchronous)
 *   buf_release         the array died: its device buffer goes to a pool reused by later acquires of the same size
 * buf_register / buf_upload / buf_download are the unconditional forms. ctx_transfer_bytes reports the bytes the context
 * has moved across the host link (registry traffic and memcpy_h2d / memcpy_d2h). */
#define DL4JB200_BUF_HOST_NEWER 0
#define DL4JB200_BUF_SYNCED 1
#define DL4JB200_BUF_DEVICE_NEWER 2
#define DL4JB200_BUF_READ 1
#define DL4JB200_BUF_WRITE 2
int dl4jb200_buf_register(dl4jb200_ctx *ctx, const void *host_ptr, size_t bytes, void **device_out);
int dl4jb200_buf_acquire(dl4jb200_ctx *ctx, const void *host_ptr, size_t bytes, int access, void **device_out);
int dl4jb200_buf_mark_host_dirty(dl4jb200_ctx *ctx, const void *host_ptr);
int dl4jb200_buf_sync_to_host(dl4jb200_ctx *ctx, void *host_ptr);
int dl4jb200_buf_state(dl4jb200_ctx *ctx, const void *host_ptr, int *state_out); /* -1: not registered */
int dl4jb200_ctx_transfer_bytes(dl4jb200_ctx *ctx, unsigned long long *h2d_out, unsigned long long *d2h_out);
int dl4jb200_buf_lookup(dl4jb200_ctx *ctx, const void *host_ptr, void **device_out, size_t *bytes_out);
int dl4jb200_buf_upload(dl4jb200_ctx *ctx, const void *host_ptr);  /* host -> device mirror */
int dl4jb200_buf_download(dl4jb200_ctx *ctx, void *host_ptr);      /* device mirror -> host */
int dl4jb200_buf_release(dl4jb200_ctx *ctx, const void *host_ptr);

/* ------------------------------------------------------------------ shapes ------------------- */
/* NN/util/ConvolutionUtils.java:96-131 getOutputSize, :213-221 effectiveKernelSize, :243-266
 * getSameModeTopLeftPadding / getSameModeBottomRightPadding, :133-211 validateShapes. One spatial axis.
 * Returns DL4JB200_ERR_INVALID_INPUT / _CONFIG where the reference throws. */
int dl4jb200_out_size(int in, int kernel, int stride, int pad, int dilation, int mode, int *out,
                      int *pad_lo, int *pad_hi);

/* NCHW <-> NHWC (pixel stride cs) re-layout on the device; pad channels are written as zeros. */
int dl4jb200_nchw_to_nhwc(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int cs);
int dl4jb200_nhwc_to_nchw(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int cs);

/* ------------------------------------------------------------------ convolution -------------- */
/* ConvolutionHelper.preOutput (NN/nn/layers/convolution/ConvolutionHelper.java:42-43; CPU semantics
 * ConvolutionLayer.java:385-417): z = conv2d(x, W) + b, cross-correlation, zero padding.
 * w: DL4J layout [c_out, c_in, k_h, k_w] 'c' order (a view into the flat parameter vector);
 * bias [c_out] or NULL (hasBias=false, ConvolutionLayer.java:351-358). act: activation fused into the
 * epilogue (ConvolutionHelper.activate applied to z in place, ConvolutionLayer.java:456-461);
 * pass DL4JB200_ACT_IDENTITY for a plain preOutput. x_cs/z_cs: NHWC pixel strides (ignored for NCHW). */
int dl4jb200_conv_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs,
                      const float *w, const float *bias, float *z, int z_layout, int z_cs, int act, int algo);

/* ConvolutionHelper.backpropGradient (ConvolutionHelper.java:37-40; CPU semantics
 * ConvolutionLayer.java:202-257). delta = dL/dz (the layer applies afn.backprop before calling the helper,
 * ConvolutionLayer.java:170-171). dw / db are the weightGradView / biasGradView: written IN PLACE,
 * overwritten (not accumulated), summed over the batch, DL4J layout. Any of dw, db, dx may be NULL (that part is then
 * skipped), which lets a caller run the filter/bias gradients and the data gradient of one layer on two contexts /
 * streams concurrently: they are independent of each other. */
int dl4jb200_conv_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs,
                      const float *w, const float *delta, int delta_layout, int delta_cs, float *dw, float *db,
                      float *dx, int dx_layout, int dx_cs, int algo);

/* Bias gradient alone: db[c] = sum over the M = B*Ho*Wo pixels of delta[m, c] (delta NHWC, pixel stride cs)
 * (ConvolutionLayer.java:257). dl4jb200_conv_bwd computes it too; this is for callers that split the backward. */
int dl4jb200_bias_grad(dl4jb200_ctx *ctx, const float *delta, int M, int cs, int C, float *db);

/* Row-packed first-layer ("stem") convolution: c_in <= 4, stride_w <= 8, k_w <= 7 (k_w <= 15 when stride_w divides 8: the
 * filter row then takes two 8-pixel K-chunks), no dilation -- ResNet-50's 7x7/2 on 3 channels
 * (deeplearning4j-zoo/.../zoo/model/ResNet50.java:181-186), VGG16's 3x3/1, LeNet's 5x5/1, AlexNet's 11x11/4. Same arithmetic as
 * dl4jb200_conv_fwd / the filter-gradient half of dl4jb200_conv_bwd, but the input is held row-packed so that TMA can
 * feed the tensor cores: xp = [B][in_h][row_pitch_px][4] floats, image pixel x at column pad_left + 1 + x, every other
 * column (and channel 3) zero; row_pitch_px even and >= max(stride_w*(out_w-1) + 8*ceil((k_w+1)/8), pad_left + 1 + in_w). Padding rows are not
 * stored (the TMA unit zero-fills them). Returns NOT_SUPPORTED when the shape does not qualify.
 * dl4jb200_nchw_to_rowpack4 writes a DL4J-layout minibatch into such a buffer (pads must already be zero). */
int dl4jb200_nchw_to_rowpack4(dl4jb200_ctx *ctx, const float *src, float *dst, int B, int C, int H, int W, int row_pitch_px,
                              int x_off);
int dl4jb200_conv_rowpack_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px, const float *w,
                              const float *bias, float *z, int z_cs, int act);
/* conv_rowpack_fwd with the following BatchNormalization's statistics taken in the epilogue (see conv_fwd_bnstats):
 * stat_partial must hold 4 * (number of SMs) rows of 2 * z_cs floats; *rows_out receives the rows written. */
int dl4jb200_conv_rowpack_fwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px,
                                      const float *w, const float *bias, float *z, int z_cs, const float *pivot,
                                      float *stat_partial, int *rows_out);
/* db may be NULL (filter gradient only); otherwise the bias gradient is produced too (inside the GEMM when
 * DL4JB200_OPT_TF32_ACTIVATIONS is on, by a column-sum pass otherwise). */
int dl4jb200_conv_rowpack_wgrad(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *xp, int row_pitch_px,
                                const float *delta, int delta_cs, float *dw, float *db);

/* ConvolutionHelper.activate (ConvolutionHelper.java:45) and IActivation.backprop
 * (ConvolutionLayer.java:171): elementwise over n floats; in place allowed (a == z, delta == eps).
 * act_bwd: act may carry DL4JB200_ACT_FROM_OUTPUT when only the activated tensor was kept. */
int dl4jb200_act_fwd(dl4jb200_ctx *ctx, const float *z, float *a, size_t n, int act);
int dl4jb200_act_bwd(dl4jb200_ctx *ctx, const float *z, const float *eps, float *delta, size_t n, int act);

/* ------------------------------------------------------------------ pooling ------------------ */
/* SubsamplingHelper.activate (NN/nn/layers/convolution/subsampling/SubsamplingHelper.java:36-38; CPU
 * semantics SubsamplingLayer.java:339-360). MAX ignores padding; AVG divides by k_h*k_w including
 * padding; PNORM uses exponent pnorm. argmax (int32, same shape/layout as y, MAX only) may be NULL;
 * it receives kh*k_w+kw of the first maximum in row-major window order. SUM => NOT_SUPPORTED on bwd. */
int dl4jb200_pool_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, float *y, int32_t *argmax,
                      int layout, int x_cs, int y_cs, int pool_type, int pnorm);
/* SubsamplingHelper.backpropGradient (SubsamplingHelper.java:33-34; CPU semantics
 * SubsamplingLayer.java:205-266). argmax NULL => recomputed from x exactly as the reference's
 * im2col + IsMax does. y (forward output) is required for PNORM only. */
int dl4jb200_pool_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, const float *y,
                      const float *eps, const int32_t *argmax, float *dx, int layout, int x_cs, int y_cs,
                      int pool_type, int pnorm, double pnorm_eps);

/* ------------------------------------------------------------------ batch normalisation ------ */
/* BatchNormalizationHelper.preOutput (NN/nn/layers/normalization/BatchNormalizationHelper.java:34-35; CPU
 * semantics BatchNormalization.java:255-390). training: batch mean / biased variance over (B,H,W),
 * std = sqrt(var+eps); running stats gmean = decay*gmean + (1-decay)*mean and
 * gvar = decay*gvar + (1-decay)*(var+eps) updated IN PLACE (minibatch==0: plain assignment);
 * save_mean / save_std [C] receive what the layer keeps for backward. inference (training==0):
 * std = sqrt(gvar) with no eps added. gamma/beta NULL => 1 / 0 (lockGammaBeta). act: optional fused
 * activation of the following ActivationLayer (IDENTITY or RELU). residual: NULL, or a tensor of x's shape/layout
 * added before the activation, y = act(bn(x) + residual) -- the ElementWiseVertex(Add) + ActivationLayer that
 * close a ResNet block (NN/nn/graph/vertex/impl/ElementWiseVertex.java:81-86) folded into the BN apply pass. */
int dl4jb200_bn_fwd(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta,
                    float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C, int H, int W,
                    int layout, int cs, double decay, double eps, int training, int minibatch, int act,
                    const float *residual);
/* BatchNormalizationHelper.backpropGradient (BatchNormalizationHelper.java:31-32; CPU semantics
 * BatchNormalization.java:171-210). dgamma / dbeta are the gradient views, written in place. act: IDENTITY, or RELU
 * when the forward call fused the following ActivationLayer -- eps is then masked with relu'(gamma*xhat+beta), the
 * mask being recomputed from x (beta is read only in that case) instead of re-reading the activation. */
int dl4jb200_bn_bwd(dl4jb200_ctx *ctx, const float *x, const float *eps, const float *gamma, const float *beta,
                    const float *save_mean, const float *save_std, int act, float *dx, float *dgamma, float *dbeta,
                    int B, int C, int H, int W, int layout, int cs);

/* Allocation-free batch-norm forward. No counterpart in the reference: it exists for callers of this library that
 * must not cause a device allocation while they run -- a loop that cannot take the device-wide stall of a cudaMalloc,
 * a process whose GPU memory already belongs to another allocator (so a driver allocation may fail), or code
 * running where CUDA memory management is off limits. Same arguments, results and status codes as dl4jb200_bn_fwd,
 * but the per-block partial sums and per-channel coefficients live in a fixed 4 MiB array inside the library (held on
 * every GPU that loads it, until exit) instead of the context's on-demand scratch; it calls no cudaMalloc / cudaFree.
 * Limit: at most 4096 channels (cs for NHWC), otherwise DL4JB200_ERR_INVALID_ARG; B*H*W < 2^31 as for dl4jb200_bn_fwd.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress): make that first call
 * before starting any stream capture. All calls share the array, so calls that overlap on the GPU race: keep them on
 * one stream or serialise them. Nothing needs clearing between calls. */
int dl4jb200_bn_fwd_noalloc(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta,
                            float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C, int H, int W,
                            int layout, int cs, double decay, double eps, int training, int minibatch, int act);

/* ------------------------------------------------------------------ LRN ---------------------- */
/* LocalResponseNormalizationHelper.activate / backpropGradient
 * (NN/nn/layers/normalization/LocalResponseNormalizationHelper.java:31-35; CPU semantics
 * LocalResponseNormalization.java:150-190 / 117-147): halfN=(int)n/2, unitScale = k + alpha*sum x^2
 * (alpha NOT divided by n), y = x*unitScale^-beta. */
int dl4jb200_lrn_fwd(dl4jb200_ctx *ctx, const float *x, float *y, int B, int C, int H, int W, int layout, int cs,
                     double k, double n, double alpha, double beta);
int dl4jb200_lrn_bwd(dl4jb200_ctx *ctx, const float *x, const float *eps, float *dx, int B, int C, int H, int W,
                     int layout, int cs, double k, double n, double alpha, double beta);

/* ------------------------------------------------------------------ callers either side (§8f) - */
/* ElementWiseVertex Add (+ following ActivationLayer ReLU): out = act(a + b); backward mask from out.
 * NN/nn/graph/vertex/impl/ElementWiseVertex.java:81-86,126-131; NN/nn/layers/ActivationLayer.java:68-92 */
int dl4jb200_add_act_fwd(dl4jb200_ctx *ctx, const float *a, const float *b, float *out, size_t n, int act);
/* Batch-norm statistics taken in the producing convolution's epilogue (TF32 TMA path, NHWC, identity activation).
 * BatchNormalization.preOutput computes mean / biased variance of its input with separate reduction passes
 * (BatchNormalization.java:262-266); when that input is the output of a convolution helper call, the sums can be taken
 * while the GEMM tile is still on chip. conv_fwd_bnstats writes z AND per-warp partial sums of (z - pivot[c]) and
 * (z - pivot[c])^2 into stat_partial ([rows][2][z_cs] floats, rows from conv_fwd_bnstats_rows: 0 = this convolution
 * cannot carry them, use conv_fwd + bn_fwd); bn_fwd_prestats is bn_fwd(training) on those sums: finalize + apply only.
 * pivot: any per-channel value near the batch mean (it only conditions the fp32 sums; the running mean is a good
 * choice and may be passed directly -- it is read before it is updated); the same pointer goes to both calls. */
int dl4jb200_conv_fwd_bnstats_rows(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int x_cs, int z_cs, int algo, int *rows_out);
int dl4jb200_conv_fwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_cs, const float *w,
                              const float *bias, float *z, int z_cs, int algo, const float *pivot, float *stat_partial);
int dl4jb200_bn_fwd_prestats(dl4jb200_ctx *ctx, const float *x, float *y, const float *gamma, const float *beta, float *gmean,
                             float *gvar, float *save_mean, float *save_std, int B, int C, int H, int W, int cs, double decay,
                             double eps, int minibatch, int act, const float *residual, const float *pivot,
                             const float *stat_partial, int stat_rows);

/* Pre-packed filters. conv_fwd / conv_bwd re-pack the DL4J-layout filter W[c_out,c_in,kH,kW]
 * (ConvolutionParamInitializer.java:144-155) into GEMM operand layout on every call. A caller that knows when the
 * parameters change (once per iteration: StochasticGradientDescent.java:78, params.subi(update)) registers each filter
 * view once -- with the pixel strides its convolution calls use -- and calls dl4jb200_filters_repack after every
 * parameter update: one launch re-packs every registered filter, and conv_fwd / conv_bwd / conv_bwd_data_fused calls
 * whose `w` is registered use the stored operand instead of launching their own pack kernel. Registration allocates
 * (cudaMalloc); repack does not, except for the job table the first time. Until the first repack, after a new
 * registration, or after dl4jb200_filters_invalidate, calls pack on the fly as before. Not using these three entry
 * points changes nothing. */
int dl4jb200_filter_register(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, int x_cs, int z_cs, int algo,
                             int with_bwd_data);
int dl4jb200_filters_repack(dl4jb200_ctx *ctx);
int dl4jb200_filters_invalidate(dl4jb200_ctx *ctx);

/* Inverted dropout on a layer's input, AbstractLayer.applyDropOutIfNecessary (NN/nn/layers/AbstractLayer.java:360-374) ->
 * Dropout.applyDropout (NN/nn/conf/dropout/Dropout.java:75): y = x * mask / retain_prob with a caller-supplied 0/1 mask in
 * the tensor's own layout (nd4j's DropOutInverted RNG stream is not reproducible outside nd4j, so the mask is an input).
 * The reference has no dropout backprop step in this version: the layer's gradients are taken at the dropped input and
 * the epsilon it returns travels on unmasked. In place allowed. */
int dl4jb200_dropout_fwd(dl4jb200_ctx *ctx, const float *x, const float *mask, float *y, size_t n, double retain_prob);

/* Epsilon fan-in and activation backprop of one vertex in a single pass: delta = act'(act_out) * (a + b)
 * (NN/nn/graph/ComputationGraph.java:2044-2063 followed by NN/nn/layers/ActivationLayer.java:68-92). In place allowed. */
int dl4jb200_eps_sum_act_bwd(dl4jb200_ctx *ctx, const float *a, const float *b, const float *act_out, float *delta, size_t n,
                             int act);
/* The same bookkeeping folded into the epilogue of the backprop-data GEMM that produces the LAST contribution to a
 * vertex' epsilon: dx = act'(act_out) * (conv_bwd_data(delta) + eps_sum). NHWC only; eps_sum may be NULL or alias dx;
 * act is IDENTITY (act_out ignored) or RELU. Where the tensor-core TMA path cannot take the epilogue (FP32 algo, the
 * compact strided-1x1 walk) the same result is produced by plain backprop-data into scratch + eps_sum_act_bwd.
 * Replaces: ConvolutionLayer.backpropGradient's epsNext (ConvolutionLayer.java:205-216) + the two calls above. */
int dl4jb200_conv_bwd_data_fused(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, const float *delta, int delta_cs,
                                 float *dx, int dx_cs, const float *eps_sum, const float *act_out, int act, int algo);
/* ComputationGraph epsilon fan-in: dst += src (NN/nn/graph/ComputationGraph.java:2044-2063) */
int dl4jb200_accumulate(dl4jb200_ctx *ctx, float *dst, const float *src, size_t n);
/* softmax + NEGATIVELOGLIKELIHOOD / MCXENT output layer: p = softmax(z) row-wise, delta = p - labels,
 * loss_sum_out[0] = -sum labels*log p (device scalar). z / delta rows are z_ld / delta_ld floats apart
 * (>= K; delta's padding columns are zeroed); labels is dense [B,K].
 * NN/nn/layers/BaseOutputLayer.java:82-100,147-190 */
int dl4jb200_softmax_nll(dl4jb200_ctx *ctx, const float *z, int z_ld, const float *labels, float *delta, int delta_ld,
                         float *loss_sum_out, int B, int K);
/* Regularisation part of the score: score_acc[0] += l1*sum|w| + 0.5*l2*sum w^2 over n weights
 * (ConvolutionLayer.calcL1/calcL2, NN/nn/layers/convolution/ConvolutionLayer.java:98-123; BaseLayer likewise). */
int dl4jb200_l1l2_score(dl4jb200_ctx *ctx, const float *w, size_t n, double l1, double l2, float *score_acc);
/* The iteration's score as the network reports it: score[0] = (loss_sum[0] + reg[0]) / minibatch, loss_sum from
 * dl4jb200_softmax_nll, reg the L1/L2 terms (NULL: none) -- BaseOutputLayer.computeScore, NN/nn/layers/BaseOutputLayer.java:82-100.
 * One more node of the captured step instead of host-framework arithmetic on device scalars. */
int dl4jb200_score_combine(dl4jb200_ctx *ctx, const float *loss_sum, const float *reg, float *score, int minibatch);
/* Row-strided copy dst[r, 0:cols] = src[r, 0:cols], dst[r, cols:dst_ld] = 0 (CnnToFeedForwardPreProcessor
 * reshape when the row length is not a multiple of the 4-float channel granule). */
int dl4jb200_copy2d(dl4jb200_ctx *ctx, const float *src, int src_ld, float *dst, int dst_ld, int rows, int cols);
/* One updater block over n contiguous floats of the flat vectors: u = Updater(g); u += l2*w;
 * u += l1*sign(w); u /= batch_div; w -= u. state: n floats (2n for ADADELTA), may be NULL for SGD/NOOP.
 * NN/nn/updater/UpdaterBlock.java:142,170-193; BaseMultiLayerUpdater.java:296-307;
 * NN/optimize/solvers/StochasticGradientDescent.java:78 */
int dl4jb200_updater_apply(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n, int kind, double lr,
                           double h1, double h2, double l1, double l2, double batch_div);

/* Gradient normalisation of one layer's gradient view, in place, before the updater
 * (BaseMultiLayerUpdater.preApply, NN/nn/updater/BaseMultiLayerUpdater.java:322-380). mode 1 RenormalizeL2PerLayer:
 * g /= ||g||_2; mode 2 ClipL2PerLayer: g *= threshold/||g||_2 when the norm exceeds threshold; mode 3
 * ClipElementWiseAbsoluteValue: clamp to [-threshold, threshold]. */
int dl4jb200_grad_normalize(dl4jb200_ctx *ctx, float *g, size_t n, int mode, double threshold);

/* All updater blocks of a network in ONE launch (BaseMultiLayerUpdater.update walks its UpdaterBlocks,
 * NN/nn/updater/BaseMultiLayerUpdater.java:223-309): the flat vectors w / g / state of n floats are split into nseg
 * segments; seg_start_dev[0..nseg] (device, int64, seg_start[nseg] == n) gives their bounds and seg_cfg_dev[s]
 * (device, int32) the index of the segment's configuration in cfg_host, or -1 for a NoOp segment (BatchNormalization
 * running statistics). cfg_host: ncfg x 6 doubles {kind, lr, h1, h2, l1, l2} (host). Same arithmetic as
 * dl4jb200_updater_apply per element (AdaDelta state is block-local: 2*start .. 2*start+2*len). If score_reg_out is
 * not NULL it receives sum over segments of l1*|w|_1 + 0.5*l2*|w|_2^2 evaluated on the PRE-update weights (the
 * regularisation part of the score, BaseOutputLayer.java:82-100), replacing the per-layer dl4jb200_l1l2_score calls. */
int dl4jb200_updater_apply_segments(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n,
                                    const int64_t *seg_start_dev, const int32_t *seg_cfg_dev, int nseg,
                                    const double *cfg_host, int ncfg, double batch_div, float *score_reg_out);
/* The same on the elements [lo, hi) of the flat vectors only (lo, hi multiples of 4, or hi == n); score_accumulate != 0 adds
 * this range's regularisation term to *score_reg_out instead of overwriting it. Lets a caller update everything whose
 * gradient is final while the last filter gradient (the first layer's: backward reaches it last) is still being computed,
 * and finish that layer's slice afterwards. */
int dl4jb200_updater_apply_segments_range(dl4jb200_ctx *ctx, float *w, const float *g, float *state, size_t n,
                                          const int64_t *seg_start_dev, const int32_t *seg_cfg_dev, int nseg,
                                          const double *cfg_host, int ncfg, double batch_div, float *score_reg_out, size_t lo,
                                          size_t hi, int score_accumulate);

/* bn_fwd_prestats with the MAX pooling layer that is the normalised tensor's only consumer folded into the apply pass
 * (BatchNormalization -> ActivationLayer -> SubsamplingLayer(MAX), ResNet-50's stem,
 * deeplearning4j-zoo/.../zoo/model/ResNet50.java:181-189): y_pool / argmax are what dl4jb200_pool_fwd would have produced from
 * the BN's (activated, TF32-rounded if the context rounds) output -- same values, same first-maximum window indices -- but that
 * output itself is never written. `pool` is the pooling descriptor (dl4jb200_pool_fwd's; no dilation); NHWC, cs == C. The
 * backward pass needs nothing else: pool_bwd takes argmax, bn_bwd recomputes the activation gate from x. */
int dl4jb200_bn_fwd_prestats_maxpool(dl4jb200_ctx *ctx, const float *x, float *y_pool, int32_t *argmax, const float *gamma,
                                     const float *beta, float *gmean, float *gvar, float *save_mean, float *save_std, int B, int C,
                                     int H, int W, int cs, double decay, double eps, int minibatch, int act, const float *pivot,
                                     const float *stat_partial, int stat_rows, const dl4jb200_conv_desc *pool);

/* SubsamplingLayer.backpropGradient for MAX pooling (SubsamplingLayer.java:205-218, 263; saved taps, NHWC, stride-2 windows of
 * at most 3x3, non-negative padding) that also takes the backward sums of the BatchNormalization whose output the pooling
 * layer consumed (BatchNormalization.java:172-196): dx is exactly what dl4jb200_pool_bwd writes -- the epsilon of that
 * batch norm -- and while each 2x2 patch of it is in registers the kernel reads the batch norm's input bn_x (laid out like
 * dx), recomputes the fused-ReLU gate as dl4jb200_bn_bwd does and leaves sum e', sum e' * (x - mean) in stat_partial
 * (*rows_out rows of 2 * cs floats; the buffer must hold 4 * #SMs rows) for dl4jb200_bn_bwd_prestats. The partial-sum pass
 * over (x, eps) -- 1.6 GB for ResNet-50's stem -- disappears. cs == C, C/4 must divide 256; NOT_SUPPORTED otherwise. */
int dl4jb200_pool_bwd_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *pool, const float *eps, const int32_t *argmax, float *dx,
                              int cs, const float *bn_x, const float *gamma, const float *beta, const float *save_mean,
                              const float *save_std, int bn_relu, float *stat_partial, int *rows_out);

/* Backward counterpart of conv_fwd_bnstats. BatchNormalization.backpropGradient needs dBeta = sum eps and
 * sum eps * (x - mean) (BatchNormalization.java:172-196) before it can form dx: a separate reduction pass over (x, eps).
 * When eps is produced by a helper call -- the backprop-data of the convolution that consumed the layer's output, with or
 * without the fused fan-in sum / ReLU backprop of conv_bwd_data_fused -- the sums can be taken in that GEMM's epilogue
 * while the tile is on chip: it additionally reads bn_x (the BatchNormalization's INPUT, laid out like dx) and writes
 * [rows][2][dx_cs] partial sums (rows from conv_bwd_data_bnstats_rows; 0 = this call cannot carry them: TF32 TMA path,
 * dx_cs == c_in a multiple of the tile width, no strided 1x1). bn_relu: the layer's output passed through a fused ReLU,
 * eps is gated by (x - mean) * gamma / std + beta > 0 exactly as bn_bwd does. bn_bwd_prestats is then bn_bwd without its
 * partial-sum pass (finalize + apply). */
int dl4jb200_conv_bwd_data_bnstats_rows(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int delta_cs, int dx_cs, int algo, int *rows_out);
int dl4jb200_conv_bwd_data_bnstats(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *w, const float *delta, int delta_cs,
                                   float *dx, int dx_cs, const float *eps_sum, const float *act_out, int act, int algo,
                                   const float *bn_x, const float *bn_gamma, const float *bn_beta, const float *bn_mean,
                                   const float *bn_std, int bn_relu, float *stat_partial);
int dl4jb200_bn_bwd_prestats(dl4jb200_ctx *ctx, const float *x, const float *eps, const float *gamma, const float *beta,
                             const float *save_mean, const float *save_std, int act, float *dx, float *dgamma, float *dbeta, int B,
                             int C, int H, int W, int cs, const float *stat_partial, int stat_rows);

/* ------------------------------------------------------------------ layers that share the conv helper field (f4) - */
/* Deconvolution2DLayer (NN/nn/layers/convolution/Deconvolution2DLayer.java:145-233 forward, :48-138 backward; the reference
 * runs libnd4j's "deconv2d" / "deconv2d_bp" custom ops). d describes the DECONVOLUTION: c_in = nIn, in_h/in_w its input,
 * c_out = nOut, out_h/out_w = stride*(in-1) + kEff - 2*pad (NN/util/ConvolutionUtils.java:63-83; Same mode is refused by the
 * layer itself, :147-149). w is [nIn, nOut, kH, kW] 'c' (NN/nn/params/DeconvolutionParamInitializer.java:75-80). Runs on the
 * same tcgen05 GEMMs with input and epsilon exchanged: forward = backprop-data of the mirrored convolution (+ bias,
 * activation), filter gradient = its filter gradient, epsilon = its forward pass. */
int dl4jb200_deconv2d_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs, const float *w,
                          const float *bias, float *z, int z_layout, int z_cs, int act, int algo);
int dl4jb200_deconv2d_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, const float *x, int x_layout, int x_cs, const float *w,
                          const float *delta, int delta_layout, int delta_cs, float *dw, float *db, float *dx, int dx_layout,
                          int dx_cs, int algo);
/* z = act(z + bias[c]) in place (the bias / activation step of the layers above; NCHW or NHWC) */
int dl4jb200_bias_act(dl4jb200_ctx *ctx, float *z, const float *bias, int B, int C, int H, int W, int layout, int cs, int act);
/* Depth-wise stage of SeparableConvolution2DLayer (NN/nn/layers/convolution/SeparableConvolution2DLayer.java:160-246 / :60-152;
 * libnd4j "sconv2d" / "sconv2d_bp"): wd is [depthMultiplier, nIn, kH, kW] 'c', u channel = ci*depthMultiplier + m
 * (NN/nn/params/SeparableConvolutionParamInitializer.java). NHWC. The point-wise stage is dl4jb200_conv_fwd / conv_bwd with a
 * 1x1 kernel on u: preOutput = depthwise_fwd -> conv_fwd(1x1, bias); backprop = conv_bwd(1x1) -> depthwise_bwd. d: c_in = nIn,
 * geometry of the depth-wise kernel; c_out is ignored. */
int dl4jb200_depthwise_fwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int depth_mult, const float *x, int x_cs, const float *wd,
                           float *u, int u_cs);
int dl4jb200_depthwise_bwd(dl4jb200_ctx *ctx, const dl4jb200_conv_desc *d, int depth_mult, const float *x, int x_cs, const float *wd,
                           const float *du, int u_cs, float *dwd, float *dx, int dx_cs);

/* ------------------------------------------------------------------ data-parallel exchange --- */
/* The plug-in point for single-box data parallelism is GradientsAccumulator.storeUpdate / applyUpdate
 * (NN/optimize/solvers/accumulation/GradientsAccumulator.java:27-41), installed into every replica's optimizer by
 * ParallelWrapper's TrainingMode.CUSTOM (deeplearning4j-scaleout/.../parallelism/ParallelWrapper.java:838-843) and called
 * from StochasticGradientDescent.optimize (NN/optimize/solvers/StochasticGradientDescent.java:66-79). The reference's own
 * implementations move threshold-encoded updates through host queues (EncodedGradientsAccumulator.java:244-315, 440-520) or
 * average parameters with Nd4j.averageAndPropagate (ParallelWrapper.java:316-376); here the exchange is an NCCL all-reduce
 * of the flat fp32 vectors over NVLink, enqueued on a CUDA stream (stream == NULL: the context's). One communicator per
 * (replica, device): rank 0 creates the 128-byte id (comm_unique_id) and hands it to the other replicas -- worker threads of
 * one JVM share it in memory, separate processes send it over whatever channel started them -- then every replica calls
 * comm_init. NCCL is bound at run time (dlopen); without it these calls return NOT_SUPPORTED and nothing else is affected. */
typedef struct dl4jb200_comm dl4jb200_comm;
int dl4jb200_comm_unique_id(void *id_out, size_t bytes);
int dl4jb200_comm_init(dl4jb200_ctx *ctx, int nranks, int rank, const void *id, dl4jb200_comm **out);
int dl4jb200_comm_destroy(dl4jb200_comm *comm);
int dl4jb200_comm_nranks(dl4jb200_comm *comm);
int dl4jb200_comm_nccl_version(void);
/* in place over n floats: sum (the raw gradient: replicas then run the identical updater with batchSize = workers * B) or
 * average (parameters / updater state of the reference's AVERAGING mode) */
int dl4jb200_comm_allreduce_sum_f32(dl4jb200_comm *comm, float *buf, size_t n, void *stream);
int dl4jb200_comm_allreduce_avg_f32(dl4jb200_comm *comm, float *buf, size_t n, void *stream);
/* average several disjoint spans of one buffer as one NCCL group (host arrays of float offsets / counts) */
int dl4jb200_comm_allreduce_avg_spans_f32(dl4jb200_comm *comm, float *base, const long long *offsets, const long long *counts,
                                          int nspans, void *stream);
/* DefaultTrainer copies the master's parameters into every replica (parallelism/trainer/DefaultTrainer.java:254-311) */
int dl4jb200_comm_broadcast_f32(dl4jb200_comm *comm, float *buf, size_t n, int root, void *stream);
/* dst[off_i + j] = scale * src[off_i + j] for a DEVICE table of spans (scale 0 writes zeros): parks the BatchNormalization
 * running statistics (parameters with a NoOp updater whose gradient view is identically zero,
 * BatchNormalization.java:205-208) in their own gradient slots, scaled by 1 / workers, so that the one gradient all-reduce
 * averages them too, and moves them back afterwards. */
int dl4jb200_spans_scale_copy(dl4jb200_ctx *ctx, float *dst, const float *src, const long long *offsets_dev,
                              const long long *counts_dev, int nspans, double scale);

#ifdef __cplusplus
}
#endif
#endif /* DL4JB200_H */
