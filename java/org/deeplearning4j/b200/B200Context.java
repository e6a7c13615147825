package org.deeplearning4j.b200;

import java.util.HashSet;
import java.util.Set;

import org.bytedeco.javacpp.FloatPointer;
import org.bytedeco.javacpp.Pointer;
import org.bytedeco.javacpp.PointerPointer;
import org.deeplearning4j.nn.conf.ConvolutionMode;
import org.deeplearning4j.nn.graph.ComputationGraph;
import org.nd4j.linalg.activations.IActivation;
import org.nd4j.linalg.api.memory.MemoryWorkspace;
import org.nd4j.linalg.api.ndarray.INDArray;
import org.nd4j.linalg.api.shape.Shape;
import org.nd4j.linalg.factory.Nd4j;

/**
 * Per-worker-thread context: ParallelWrapper pins every worker thread to one device for life
 * (ParallelWrapper.java:125-141), so one dl4jb200_ctx per thread. Owns the host-mirror registry glue: an INDArray's host
 * address (nd4j-native) is registered once, uploaded when the host side wrote it, downloaded when the host side reads it.
 */
public final class B200Context {
    private static final ThreadLocal<B200Context> TL = ThreadLocal.withInitial(B200Context::new);
    private Pointer handle;
    private final Set<Long> deviceDirty = new HashSet<>();

    private B200Context() {
        try {
            PointerPointer<Pointer> out = new PointerPointer<>(1);
            int device = Nd4j.getAffinityManager().getDeviceForCurrentThread();
            if (Dl4jB200.dl4jb200_ctx_create(device, null, out) == Dl4jB200.OK) handle = out.get(0);
        } catch (Throwable t) {
            handle = null; // layer logs and uses its built-in path (ConvolutionLayer.java:83-94)
        }
    }

    public static B200Context forCurrentThread() { return TL.get(); }
    public boolean isUsable() { return handle != null; }
    public Pointer handle() { return handle; }

    public static void check(int rc) {
        if (rc < 0) throw new RuntimeException("dl4jb200: " + Dl4jB200.dl4jb200_last_error().getString());
    }

    /** Non-contiguous views are copied, as the cuDNN helper does (CudnnConvolutionHelper.java:154-157). */
    public static INDArray dense(INDArray a) {
        return (a.isView() || !Shape.strideDescendingCAscendingF(a) || a.ordering() != 'c') ? a.dup('c') : a;
    }

    /** Outputs live in WORKSPACE_EXTERNAL when it is open (CudnnConvolutionHelper.java:250-256). */
    public static INDArray createOutput(int[] shape) {
        if (Nd4j.getWorkspaceManager().checkIfWorkspaceExistsAndActive(ComputationGraph.WORKSPACE_EXTERNAL)) {
            try (MemoryWorkspace ws = Nd4j.getWorkspaceManager().getWorkspaceForCurrentThread(ComputationGraph.WORKSPACE_EXTERNAL)
                            .notifyScopeBorrowed()) {
                return Nd4j.createUninitialized(shape, 'c');
            }
        }
        return Nd4j.createUninitialized(shape, 'c');
    }

    private FloatPointer device(INDArray a, boolean upload) {
        if (a == null) return null;
        Pointer host = a.data().addressPointer();
        PointerPointer<Pointer> dev = new PointerPointer<>(1);
        check(Dl4jB200.dl4jb200_buf_register(handle, host, a.length() * 4L, dev));
        if (upload && !deviceDirty.contains(host.address())) check(Dl4jB200.dl4jb200_buf_upload(handle, host));
        return new FloatPointer(dev.get(0));
    }

    public FloatPointer in(INDArray a) { return device(a, true); }
    public FloatPointer inout(INDArray a) { return device(a, true); }
    public FloatPointer out(INDArray a) { return device(a, false); }

    /** Result produced on the device: bring it back for host-side consumers (non-helper layers, updater, listeners). */
    public INDArray done(INDArray a) {
        if (a != null) check(Dl4jB200.dl4jb200_buf_download(handle, a.data().addressPointer()));
        return a;
    }

    public static int activationCode(IActivation afn) {
        switch (afn.toString()) {
            case "identity": return Dl4jB200.ACT_IDENTITY;
            case "relu": return Dl4jB200.ACT_RELU;
            case "sigmoid": return Dl4jB200.ACT_SIGMOID;
            case "tanh": return Dl4jB200.ACT_TANH;
            case "leakyrelu": return Dl4jB200.ACT_LEAKYRELU;
            default: return -1;
        }
    }

    public static Dl4jB200.ConvDesc poolDesc(INDArray input, int[] k, int[] s, int[] pad, int[] dil, ConvolutionMode cm) {
        int mode = cm == ConvolutionMode.Same ? Dl4jB200.MODE_SAME : cm == ConvolutionMode.Strict ? Dl4jB200.MODE_STRICT : Dl4jB200.MODE_TRUNCATE;
        org.bytedeco.javacpp.IntPointer o = new org.bytedeco.javacpp.IntPointer(1), lo = new org.bytedeco.javacpp.IntPointer(1),
                        hi = new org.bytedeco.javacpp.IntPointer(1);
        check(Dl4jB200.dl4jb200_out_size(input.size(2), k[0], s[0], pad[0], dil[0], mode, o, lo, hi));
        int outH = o.get(), padT = lo.get();
        check(Dl4jB200.dl4jb200_out_size(input.size(3), k[1], s[1], pad[1], dil[1], mode, o, lo, hi));
        return new Dl4jB200.ConvDesc().set(input.size(0), input.size(1), input.size(2), input.size(3), input.size(1), k[0], k[1],
                        s[0], s[1], padT, lo.get(), dil[0], dil[1], outH, o.get());
    }

    public static int pnorm(Object helper) { return 2; }
}
