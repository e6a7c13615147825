package org.deeplearning4j.b200;

import org.bytedeco.javacpp.BytePointer;
import org.bytedeco.javacpp.FloatPointer;
import org.bytedeco.javacpp.IntPointer;
import org.bytedeco.javacpp.Loader;
import org.bytedeco.javacpp.Pointer;
import org.bytedeco.javacpp.PointerPointer;
import org.bytedeco.javacpp.annotation.Cast;
import org.bytedeco.javacpp.annotation.Const;
import org.bytedeco.javacpp.annotation.Platform;

/**
 * JavaCPP binding of include/dl4jb200.h (libdl4jb200.so). One native method per C entry point, same names and
 * argument order. NOT COMPILED in this repository's CI image (no JDK); see INTEGRATION.md.
 */
@Platform(include = "dl4jb200.h", link = "dl4jb200")
public class Dl4jB200 {
    static { Loader.load(); }

    public static final int OK = 0, NOT_SUPPORTED = 1, NCHW = 0, NHWC = 1;
    public static final int MODE_STRICT = 0, MODE_TRUNCATE = 1, MODE_SAME = 2;
    public static final int POOL_MAX = 0, POOL_AVG = 1, POOL_SUM = 2, POOL_PNORM = 3;
    public static final int ACT_IDENTITY = 0, ACT_RELU = 1, ACT_SIGMOID = 2, ACT_TANH = 3, ACT_LEAKYRELU = 4;
    public static final int ALGO_TC_TF32 = 0, ALGO_FP32 = 1;

    /** dl4jb200_conv_desc: 15 consecutive int32 fields. */
    public static class ConvDesc extends IntPointer {
        public ConvDesc() { super(15); }
        public ConvDesc set(int batch, int cIn, int inH, int inW, int cOut, int kH, int kW, int sH, int sW, int padTop,
                        int padLeft, int dH, int dW, int outH, int outW) {
            put(new int[] {batch, cIn, inH, inW, cOut, kH, kW, sH, sW, padTop, padLeft, dH, dW, outH, outW});
            return this;
        }
    }

    public static native @Cast("const char*") BytePointer dl4jb200_last_error();
    public static native int dl4jb200_ctx_create(int device, Pointer stream, @Cast("dl4jb200_ctx**") PointerPointer out);
    public static native int dl4jb200_ctx_destroy(@Cast("dl4jb200_ctx*") Pointer ctx);
    public static native int dl4jb200_ctx_sync(@Cast("dl4jb200_ctx*") Pointer ctx);
    public static final int OPT_TF32_ACTIVATIONS = 1;
    public static final int OPT_PAIR_TILES = 2;   // 256-row tiles on CTA pairs (off by default, see dl4jb200.h)
    public static native int dl4jb200_ctx_set_option(@Cast("dl4jb200_ctx*") Pointer ctx, int option, int value);
    // optional fast paths (INTEGRATION.md): pre-packed filters, fused epsilon fan-in + ReLU backprop, mask-fed dropout
    public static native int dl4jb200_filter_register(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer w, int xCs, int zCs, int algo, int withBwdData);
    public static native int dl4jb200_filters_repack(@Cast("dl4jb200_ctx*") Pointer ctx);
    public static native int dl4jb200_filters_invalidate(@Cast("dl4jb200_ctx*") Pointer ctx);
    public static native int dl4jb200_conv_bwd_data_fused(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer w, @Const FloatPointer delta, int deltaCs, FloatPointer dx, int dxCs,
                    @Const FloatPointer epsSum, @Const FloatPointer actOut, int act, int algo);
    public static native int dl4jb200_eps_sum_act_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer a, @Const FloatPointer b,
                    @Const FloatPointer actOut, FloatPointer delta, @Cast("size_t") long n, int act);
    public static native int dl4jb200_conv_fwd_bnstats_rows(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    int xCs, int zCs, int algo, IntPointer rowsOut);
    public static native int dl4jb200_conv_fwd_bnstats(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, int xCs, @Const FloatPointer w, @Const FloatPointer bias, FloatPointer z, int zCs, int algo,
                    @Const FloatPointer pivot, FloatPointer statPartial);
    public static native int dl4jb200_bn_fwd_prestats(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, FloatPointer y,
                    @Const FloatPointer gamma, @Const FloatPointer beta, FloatPointer gMean, FloatPointer gVar, FloatPointer saveMean,
                    FloatPointer saveStd, int b, int c, int h, int w, int cs, double decay, double eps, int minibatch, int act,
                    @Const FloatPointer residual, @Const FloatPointer pivot, @Const FloatPointer statPartial, int statRows);
    public static native int dl4jb200_bn_fwd_prestats_maxpool(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, FloatPointer yPool,
                    IntPointer argmax, @Const FloatPointer gamma, @Const FloatPointer beta, FloatPointer gMean, FloatPointer gVar,
                    FloatPointer saveMean, FloatPointer saveStd, int b, int c, int h, int w, int cs, double decay, double eps,
                    int minibatch, int act, @Const FloatPointer pivot, @Const FloatPointer statPartial, int statRows,
                    @Cast("const dl4jb200_conv_desc*") IntPointer pool);
    public static native int dl4jb200_pool_bwd_bnstats(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer pool,
                    @Const FloatPointer eps, @Const IntPointer argmax, FloatPointer dx, int cs, @Const FloatPointer bnX,
                    @Const FloatPointer gamma, @Const FloatPointer beta, @Const FloatPointer saveMean, @Const FloatPointer saveStd,
                    int bnRelu, FloatPointer statPartial, IntPointer rowsOut);
    public static native int dl4jb200_dropout_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, @Const FloatPointer mask,
                    FloatPointer y, @Cast("size_t") long n, double retainProb);
    public static native int dl4jb200_buf_register(@Cast("dl4jb200_ctx*") Pointer ctx, @Const Pointer host, long bytes,
                    @Cast("void**") PointerPointer deviceOut);
    public static final int BUF_HOST_NEWER = 0, BUF_SYNCED = 1, BUF_DEVICE_NEWER = 2, BUF_READ = 1, BUF_WRITE = 2;
    public static native int dl4jb200_buf_acquire(@Cast("dl4jb200_ctx*") Pointer ctx, @Const Pointer host, @Cast("size_t") long bytes,
                    int access, @Cast("void**") PointerPointer deviceOut);
    public static native int dl4jb200_buf_mark_host_dirty(@Cast("dl4jb200_ctx*") Pointer ctx, @Const Pointer host);
    public static native int dl4jb200_buf_sync_to_host(@Cast("dl4jb200_ctx*") Pointer ctx, Pointer host);
    public static native int dl4jb200_buf_release(@Cast("dl4jb200_ctx*") Pointer ctx, @Const Pointer host);
    public static native int dl4jb200_ctx_transfer_bytes(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("unsigned long long*") long[] h2d,
                    @Cast("unsigned long long*") long[] d2h);
    // data-parallel exchange (B200GradientsAccumulator)
    public static native int dl4jb200_comm_unique_id(Pointer idOut, @Cast("size_t") long bytes);
    public static native int dl4jb200_comm_init(@Cast("dl4jb200_ctx*") Pointer ctx, int nranks, int rank, @Const Pointer id,
                    @Cast("dl4jb200_comm**") PointerPointer out);
    public static native int dl4jb200_comm_destroy(@Cast("dl4jb200_comm*") Pointer comm);
    public static native int dl4jb200_comm_allreduce_sum_f32(@Cast("dl4jb200_comm*") Pointer comm, FloatPointer buf, @Cast("size_t") long n,
                    Pointer stream);
    public static native int dl4jb200_comm_allreduce_avg_f32(@Cast("dl4jb200_comm*") Pointer comm, FloatPointer buf, @Cast("size_t") long n,
                    Pointer stream);
    public static native int dl4jb200_comm_broadcast_f32(@Cast("dl4jb200_comm*") Pointer comm, FloatPointer buf, @Cast("size_t") long n,
                    int root, Pointer stream);
    public static native int dl4jb200_ctx_workspace_freeze(@Cast("dl4jb200_ctx*") Pointer ctx, int frozen);
    public static native int dl4jb200_buf_upload(@Cast("dl4jb200_ctx*") Pointer ctx, @Const Pointer host);
    public static native int dl4jb200_buf_download(@Cast("dl4jb200_ctx*") Pointer ctx, Pointer host);
    public static native int dl4jb200_out_size(int in, int kernel, int stride, int pad, int dilation, int mode, IntPointer out,
                    IntPointer padLo, IntPointer padHi);
    public static native int dl4jb200_conv_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, int xLayout, int xCs, @Const FloatPointer w, @Const FloatPointer bias, FloatPointer z,
                    int zLayout, int zCs, int act, int algo);
    public static native int dl4jb200_conv_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, int xLayout, int xCs, @Const FloatPointer w, @Const FloatPointer delta, int deltaLayout,
                    int deltaCs, FloatPointer dw, FloatPointer db, FloatPointer dx, int dxLayout, int dxCs, int algo);
    public static native int dl4jb200_act_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer z, FloatPointer a, long n, int act);
    public static native int dl4jb200_pool_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, FloatPointer y, IntPointer argmax, int layout, int xCs, int yCs, int poolType, int pnorm);
    public static native int dl4jb200_pool_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, @Const FloatPointer y, @Const FloatPointer eps, @Const IntPointer argmax, FloatPointer dx,
                    int layout, int xCs, int yCs, int poolType, int pnorm, double pnormEps);
    public static native int dl4jb200_bn_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, FloatPointer y,
                    @Const FloatPointer gamma, @Const FloatPointer beta, FloatPointer gmean, FloatPointer gvar, FloatPointer saveMean,
                    FloatPointer saveStd, int b, int c, int h, int w, int layout, int cs, double decay, double eps, int training,
                    int minibatch, int act, @Const FloatPointer residual);
    public static native int dl4jb200_bn_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, @Const FloatPointer eps,
                    @Const FloatPointer gamma, @Const FloatPointer beta, @Const FloatPointer saveMean, @Const FloatPointer saveStd, int act,
                    FloatPointer dx, FloatPointer dgamma, FloatPointer dbeta, int b, int c, int h, int w, int layout, int cs);
    public static native int dl4jb200_lrn_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, FloatPointer y, int b, int c,
                    int h, int w, int layout, int cs, double k, double n, double alpha, double beta);
    public static native int dl4jb200_lrn_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, @Const FloatPointer eps,
                    FloatPointer dx, int b, int c, int h, int w, int layout, int cs, double k, double n, double alpha, double beta);

    // ---- layers that share the convolution helper field (Deconvolution2D, SeparableConvolution2D), fused backward sums,
    // activation backprop, allocation-free batch norm. The entry points of include/dl4jb200.h that only the device-resident
    // step driver uses (updaters, re-layouts, softmax loss, raw memory) are not bound by hand here: JavaCPP's parser
    // generates them from the header when they are needed.
    public static final int ACT_FROM_OUTPUT = 0x200;
    public static native int dl4jb200_act_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer z, @Const FloatPointer eps,
                    FloatPointer delta, @Cast("size_t") long n, int act);
    public static native int dl4jb200_bias_grad(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer delta, int m, int cs, int c,
                    FloatPointer db);
    public static native int dl4jb200_bias_act(@Cast("dl4jb200_ctx*") Pointer ctx, FloatPointer z, @Const FloatPointer bias, int b, int c,
                    int h, int w, int layout, int cs, int act);
    public static native int dl4jb200_deconv2d_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, int xLayout, int xCs, @Const FloatPointer w, @Const FloatPointer bias, FloatPointer z,
                    int zLayout, int zCs, int act, int algo);
    public static native int dl4jb200_deconv2d_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer x, int xLayout, int xCs, @Const FloatPointer w, @Const FloatPointer delta, int deltaLayout,
                    int deltaCs, FloatPointer dw, FloatPointer db, FloatPointer dx, int dxLayout, int dxCs, int algo);
    public static native int dl4jb200_depthwise_fwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    int depthMult, @Const FloatPointer x, int xCs, @Const FloatPointer wd, FloatPointer u, int uCs);
    public static native int dl4jb200_depthwise_bwd(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    int depthMult, @Const FloatPointer x, int xCs, @Const FloatPointer wd, @Const FloatPointer du, int uCs,
                    FloatPointer dwd, FloatPointer dx, int dxCs);
    public static native int dl4jb200_conv_bwd_data_bnstats_rows(@Cast("dl4jb200_ctx*") Pointer ctx,
                    @Cast("const dl4jb200_conv_desc*") IntPointer d, int deltaCs, int dxCs, int algo, IntPointer rowsOut);
    public static native int dl4jb200_conv_bwd_data_bnstats(@Cast("dl4jb200_ctx*") Pointer ctx, @Cast("const dl4jb200_conv_desc*") IntPointer d,
                    @Const FloatPointer w, @Const FloatPointer delta, int deltaCs, FloatPointer dx, int dxCs, @Const FloatPointer epsSum,
                    @Const FloatPointer actOut, int act, int algo, @Const FloatPointer bnX, @Const FloatPointer bnGamma,
                    @Const FloatPointer bnBeta, @Const FloatPointer bnMean, @Const FloatPointer bnStd, int bnRelu, FloatPointer statPartial);
    public static native int dl4jb200_bn_bwd_prestats(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, @Const FloatPointer eps,
                    @Const FloatPointer gamma, @Const FloatPointer beta, @Const FloatPointer saveMean, @Const FloatPointer saveStd, int act,
                    FloatPointer dx, FloatPointer dgamma, FloatPointer dbeta, int b, int c, int h, int w, int cs,
                    @Const FloatPointer statPartial, int statRows);
    public static native int dl4jb200_bn_fwd_noalloc(@Cast("dl4jb200_ctx*") Pointer ctx, @Const FloatPointer x, FloatPointer y,
                    @Const FloatPointer gamma, @Const FloatPointer beta, FloatPointer gmean, FloatPointer gvar, FloatPointer saveMean,
                    FloatPointer saveStd, int b, int c, int h, int w, int layout, int cs, double decay, double eps, int training,
                    int minibatch, int act);
}
