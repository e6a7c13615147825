package org.deeplearning4j.nn.layers.convolution;

import org.bytedeco.javacpp.FloatPointer;
import org.bytedeco.javacpp.IntPointer;
import org.bytedeco.javacpp.Pointer;
import org.deeplearning4j.b200.B200Context;
import org.deeplearning4j.b200.Dl4jB200;
import org.deeplearning4j.nn.conf.ConvolutionMode;
import org.deeplearning4j.nn.conf.layers.ConvolutionLayer.AlgoMode;
import org.deeplearning4j.nn.conf.layers.ConvolutionLayer.BwdDataAlgo;
import org.deeplearning4j.nn.conf.layers.ConvolutionLayer.BwdFilterAlgo;
import org.deeplearning4j.nn.conf.layers.ConvolutionLayer.FwdAlgo;
import org.deeplearning4j.nn.gradient.DefaultGradient;
import org.deeplearning4j.nn.gradient.Gradient;
import org.deeplearning4j.nn.params.ConvolutionParamInitializer;
import org.nd4j.linalg.activations.IActivation;
import org.nd4j.linalg.api.buffer.DataBuffer;
import org.nd4j.linalg.api.ndarray.INDArray;
import org.nd4j.linalg.factory.Nd4j;
import org.nd4j.linalg.primitives.Pair;

/**
 * Drop-in for deeplearning4j-cuda's class of the same name: ConvolutionLayer finds it through
 * Class.forName("org.deeplearning4j.nn.layers.convolution.CudnnConvolutionHelper") (ConvolutionLayer.java:64-95).
 * Every method = marshal + one call into libdl4jb200.so. Mirrored 1:1 by deeplearning4j_b200/helpers.py, which is
 * what this repository's tests execute (no JDK in the build image).
 */
public class CudnnConvolutionHelper implements ConvolutionHelper {
    private final B200Context ctx = B200Context.forCurrentThread();

    @Override
    public boolean checkSupported() {
        // DOUBLE (the reference's -Ddtype=double test runs) and HALF fall back to the built-in path
        return ctx.isUsable() && Nd4j.dataType() == DataBuffer.Type.FLOAT;
    }

    private static Dl4jB200.ConvDesc desc(INDArray input, int nOut, int[] k, int[] s, int[] pad, int[] dil, ConvolutionMode cm) {
        int mode = cm == ConvolutionMode.Same ? Dl4jB200.MODE_SAME : cm == ConvolutionMode.Strict ? Dl4jB200.MODE_STRICT : Dl4jB200.MODE_TRUNCATE;
        IntPointer o = new IntPointer(1), lo = new IntPointer(1), hi = new IntPointer(1);
        B200Context.check(Dl4jB200.dl4jb200_out_size(input.size(2), k[0], s[0], pad[0], dil[0], mode, o, lo, hi));
        int outH = o.get(), padT = lo.get();
        B200Context.check(Dl4jB200.dl4jb200_out_size(input.size(3), k[1], s[1], pad[1], dil[1], mode, o, lo, hi));
        return new Dl4jB200.ConvDesc().set(input.size(0), input.size(1), input.size(2), input.size(3), nOut, k[0], k[1], s[0], s[1],
                        padT, lo.get(), dil[0], dil[1], outH, o.get());
    }

    @Override
    public INDArray preOutput(INDArray input, INDArray weights, INDArray bias, int[] kernel, int[] strides, int[] pad,
                    AlgoMode mode, FwdAlgo fwdAlgo, ConvolutionMode convolutionMode, int[] dilation) {
        input = B200Context.dense(input); // views / permuted arrays are dup'ed, as CudnnConvolutionHelper.java:574-576 does
        Dl4jB200.ConvDesc d = desc(input, weights.size(0), kernel, strides, pad, dilation, convolutionMode);
        INDArray z = B200Context.createOutput(new int[] {d.get(0), d.get(4), d.get(13), d.get(14)}); // WORKSPACE_EXTERNAL aware
        int algo = mode == AlgoMode.NO_WORKSPACE ? Dl4jB200.ALGO_FP32 : Dl4jB200.ALGO_TC_TF32;
        int rc = Dl4jB200.dl4jb200_conv_fwd(ctx.handle(), d, ctx.in(input), Dl4jB200.NCHW, 0, ctx.in(weights), ctx.in(bias),
                        ctx.out(z), Dl4jB200.NCHW, 0, Dl4jB200.ACT_IDENTITY, algo);
        if (rc == Dl4jB200.NOT_SUPPORTED) return null; // layer runs its built-in path (ConvolutionLayer.java:372-374)
        B200Context.check(rc);
        return ctx.done(z);
    }

    @Override
    public INDArray activate(INDArray z, IActivation afn) {
        int act = B200Context.activationCode(afn); // -1 for anything but identity/relu/sigmoid/tanh/leakyrelu
        if (act < 0) return null;
        FloatPointer p = ctx.inout(z);
        B200Context.check(Dl4jB200.dl4jb200_act_fwd(ctx.handle(), p, p, z.length(), act));
        return ctx.done(z);
    }

    @Override
    public Pair<Gradient, INDArray> backpropGradient(INDArray input, INDArray weights, INDArray delta, int[] kernel,
                    int[] strides, int[] pad, INDArray biasGradView, INDArray weightGradView, IActivation afn, AlgoMode mode,
                    BwdFilterAlgo bwdFilterAlgo, BwdDataAlgo bwdDataAlgo, ConvolutionMode convolutionMode, int[] dilation) {
        input = B200Context.dense(input);
        delta = B200Context.dense(delta); // delta already holds afn.backprop(z, eps) (ConvolutionLayer.java:170-171)
        Dl4jB200.ConvDesc d = desc(input, weights.size(0), kernel, strides, pad, dilation, convolutionMode);
        INDArray epsNext = B200Context.createOutput(input.shape());
        int algo = mode == AlgoMode.NO_WORKSPACE ? Dl4jB200.ALGO_FP32 : Dl4jB200.ALGO_TC_TF32;
        // weightGradView / biasGradView are views into the flat gradient vector: written in place
        int rc = Dl4jB200.dl4jb200_conv_bwd(ctx.handle(), d, ctx.in(input), Dl4jB200.NCHW, 0, ctx.in(weights), ctx.in(delta),
                        Dl4jB200.NCHW, 0, ctx.out(weightGradView), ctx.out(biasGradView), ctx.out(epsNext), Dl4jB200.NCHW, 0, algo);
        if (rc == Dl4jB200.NOT_SUPPORTED) return null;
        B200Context.check(rc);
        ctx.done(weightGradView);
        ctx.done(biasGradView);
        Gradient g = new DefaultGradient();
        g.setGradientFor(ConvolutionParamInitializer.BIAS_KEY, biasGradView);
        g.setGradientFor(ConvolutionParamInitializer.WEIGHT_KEY, weightGradView, 'c');
        return new Pair<>(g, ctx.done(epsNext));
    }
}
