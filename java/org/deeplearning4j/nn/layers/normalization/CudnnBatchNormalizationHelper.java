package org.deeplearning4j.nn.layers.normalization;

import org.deeplearning4j.b200.B200Context;
import org.deeplearning4j.b200.Dl4jB200;
import org.deeplearning4j.nn.gradient.DefaultGradient;
import org.deeplearning4j.nn.gradient.Gradient;
import org.deeplearning4j.nn.params.BatchNormalizationParamInitializer;
import org.nd4j.linalg.api.ndarray.INDArray;
import org.nd4j.linalg.factory.Nd4j;
import org.nd4j.linalg.primitives.Pair;

/** Drop-in found by BatchNormalization.java:55-75. Mirrored by helpers.CudnnBatchNormalizationHelper. */
public class CudnnBatchNormalizationHelper implements BatchNormalizationHelper {
    private final B200Context ctx = B200Context.forCurrentThread();
    private INDArray saveMean, saveStd; // meanCache / varCache of the cuDNN helper

    @Override
    public boolean checkSupported(double eps) { return ctx.isUsable(); } // no CUDNN_BN_MIN_EPSILON restriction

    @Override
    public INDArray preOutput(INDArray x, boolean training, int[] shape, INDArray gamma, INDArray beta, INDArray mean, INDArray var,
                    double decay, double eps) {
        x = B200Context.dense(x);
        INDArray y = B200Context.createOutput(x.shape());
        if (saveMean == null || saveMean.length() != shape[1]) { saveMean = Nd4j.create(1, shape[1]); saveStd = Nd4j.create(1, shape[1]); }
        // mean / var are the GLOBAL_MEAN / GLOBAL_VAR parameter views: updated in place with the reference's CPU rule
        // gVar = decay*gVar + (1-decay)*(var+eps) (BatchNormalization.java:377-384), NOT cuDNN's exponentialAverageFactor
        B200Context.check(Dl4jB200.dl4jb200_bn_fwd(ctx.handle(), ctx.in(x), ctx.out(y), ctx.in(gamma), ctx.in(beta), ctx.inout(mean),
                        ctx.inout(var), ctx.out(saveMean), ctx.out(saveStd), shape[0], shape[1], shape[2], shape[3], Dl4jB200.NCHW, 0,
                        decay, eps, training ? 1 : 0, 1, Dl4jB200.ACT_IDENTITY, null));
        if (training) { ctx.done(mean); ctx.done(var); }
        return ctx.done(y);
    }

    @Override
    public Pair<Gradient, INDArray> backpropGradient(INDArray input, INDArray epsilon, int[] shape, INDArray gamma, INDArray dGammaView,
                    INDArray dBetaView, double eps) {
        input = B200Context.dense(input);
        epsilon = B200Context.dense(epsilon);
        INDArray epsNext = B200Context.createOutput(input.shape());
        B200Context.check(Dl4jB200.dl4jb200_bn_bwd(ctx.handle(), ctx.in(input), ctx.in(epsilon), ctx.in(gamma), null, ctx.in(saveMean),
                        ctx.in(saveStd), Dl4jB200.ACT_IDENTITY, ctx.out(epsNext), ctx.out(dGammaView), ctx.out(dBetaView), shape[0], shape[1], shape[2],
                        shape[3], Dl4jB200.NCHW, 0));
        Gradient g = new DefaultGradient();
        g.setGradientFor(BatchNormalizationParamInitializer.GAMMA, ctx.done(dGammaView));
        g.setGradientFor(BatchNormalizationParamInitializer.BETA, ctx.done(dBetaView));
        return new Pair<>(g, ctx.done(epsNext));
    }
}
