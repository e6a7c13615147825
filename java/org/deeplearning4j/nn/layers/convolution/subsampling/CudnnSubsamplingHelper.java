package org.deeplearning4j.nn.layers.convolution.subsampling;

import org.deeplearning4j.b200.B200Context;
import org.deeplearning4j.b200.Dl4jB200;
import org.deeplearning4j.nn.conf.ConvolutionMode;
import org.deeplearning4j.nn.conf.layers.PoolingType;
import org.deeplearning4j.nn.gradient.DefaultGradient;
import org.deeplearning4j.nn.gradient.Gradient;
import org.nd4j.linalg.api.ndarray.INDArray;
import org.nd4j.linalg.primitives.Pair;

/** Drop-in found by SubsamplingLayer.java:74-94. Mirrored by helpers.CudnnSubsamplingHelper. */
public class CudnnSubsamplingHelper implements SubsamplingHelper {
    private final B200Context ctx = B200Context.forCurrentThread();
    private INDArray reduced; // forward output, kept for PNORM backprop like the cuDNN helper's `reduced`

    @Override
    public boolean checkSupported() { return ctx.isUsable(); }

    private static int type(PoolingType t) {
        switch (t) {
            case MAX: return Dl4jB200.POOL_MAX;
            case AVG: return Dl4jB200.POOL_AVG;
            case SUM: return Dl4jB200.POOL_SUM;
            default: return Dl4jB200.POOL_PNORM;
        }
    }

    @Override
    public INDArray activate(INDArray input, boolean training, int[] kernel, int[] strides, int[] pad, PoolingType poolingType,
                    ConvolutionMode convolutionMode, int[] dilation) {
        input = B200Context.dense(input);
        Dl4jB200.ConvDesc d = B200Context.poolDesc(input, kernel, strides, pad, dilation, convolutionMode);
        reduced = B200Context.createOutput(new int[] {d.get(0), d.get(1), d.get(13), d.get(14)});
        int rc = Dl4jB200.dl4jb200_pool_fwd(ctx.handle(), d, ctx.in(input), ctx.out(reduced), null, Dl4jB200.NCHW, 0, 0,
                        type(poolingType), B200Context.pnorm(this));
        if (rc == Dl4jB200.NOT_SUPPORTED) return null;
        B200Context.check(rc);
        return ctx.done(reduced);
    }

    @Override
    public Pair<Gradient, INDArray> backpropGradient(INDArray input, INDArray epsilon, int[] kernel, int[] strides, int[] pad,
                    PoolingType poolingType, ConvolutionMode convolutionMode, int[] dilation) {
        input = B200Context.dense(input);
        epsilon = B200Context.dense(epsilon);
        Dl4jB200.ConvDesc d = B200Context.poolDesc(input, kernel, strides, pad, dilation, convolutionMode);
        INDArray epsNext = B200Context.createOutput(input.shape());
        // argmax == NULL: IsMax is re-derived from x on the device, exactly as SubsamplingLayer.java:205-218 does
        int rc = Dl4jB200.dl4jb200_pool_bwd(ctx.handle(), d, ctx.in(input), reduced == null ? null : ctx.in(reduced), ctx.in(epsilon),
                        null, ctx.out(epsNext), Dl4jB200.NCHW, 0, 0, type(poolingType), B200Context.pnorm(this), 1e-8);
        if (rc == Dl4jB200.NOT_SUPPORTED) return null; // SUM: the layer's own code throws (SubsamplingLayer.java:251-253)
        B200Context.check(rc);
        return new Pair<Gradient, INDArray>(new DefaultGradient(), ctx.done(epsNext));
    }
}
