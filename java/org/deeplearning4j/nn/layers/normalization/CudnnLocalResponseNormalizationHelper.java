package org.deeplearning4j.nn.layers.normalization;

import org.deeplearning4j.b200.B200Context;
import org.deeplearning4j.b200.Dl4jB200;
import org.deeplearning4j.nn.gradient.DefaultGradient;
import org.deeplearning4j.nn.gradient.Gradient;
import org.nd4j.linalg.api.ndarray.INDArray;
import org.nd4j.linalg.primitives.Pair;

/** Drop-in found by LocalResponseNormalization.java:76-97. Mirrored by helpers.CudnnLocalResponseNormalizationHelper. */
public class CudnnLocalResponseNormalizationHelper implements LocalResponseNormalizationHelper {
    private final B200Context ctx = B200Context.forCurrentThread();

    @Override
    public boolean checkSupported(double k, double n, double alpha, double beta) { return ctx.isUsable() && n >= 1; }

    @Override
    public INDArray activate(INDArray x, boolean training, double k, double n, double alpha, double beta) {
        x = B200Context.dense(x);
        INDArray y = B200Context.createOutput(x.shape());
        // alpha is passed through unchanged: the CPU path does NOT divide it by n (LocalResponseNormalization.java:167-188)
        B200Context.check(Dl4jB200.dl4jb200_lrn_fwd(ctx.handle(), ctx.in(x), ctx.out(y), x.size(0), x.size(1), x.size(2), x.size(3),
                        Dl4jB200.NCHW, 0, k, n, alpha, beta));
        return ctx.done(y);
    }

    @Override
    public Pair<Gradient, INDArray> backpropGradient(INDArray input, INDArray epsilon, double k, double n, double alpha, double beta) {
        input = B200Context.dense(input);
        epsilon = B200Context.dense(epsilon);
        INDArray epsNext = B200Context.createOutput(input.shape());
        B200Context.check(Dl4jB200.dl4jb200_lrn_bwd(ctx.handle(), ctx.in(input), ctx.in(epsilon), ctx.out(epsNext), input.size(0),
                        input.size(1), input.size(2), input.size(3), Dl4jB200.NCHW, 0, k, n, alpha, beta));
        return new Pair<Gradient, INDArray>(new DefaultGradient(), ctx.done(epsNext));
    }
}
