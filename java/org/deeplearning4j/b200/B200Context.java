package org.deeplearning4j.b200;

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
    /**
     * Lazy device-to-host synchronisation. nd4j-native gives no hook when host code reads an INDArray, so by default a
     * helper result is brought back before the helper returns (done()); a caller that promises to call syncToHost()
     * before any host-side read (e.g. a MultiLayerNetwork whose CNN stack consists of helper-backed layers only) may
     * switch it off with -Ddl4jb200.lazySync=true and then pays PCIe only at the minibatch and parameter boundaries.
     */
    private final boolean lazySync = Boolean.getBoolean("dl4jb200.lazySync");

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

    private FloatPointer device(INDArray a, int access) {
        if (a == null) return null;
        Pointer host = a.data().addressPointer();
        PointerPointer<Pointer> dev = new PointerPointer<>(1);
        // registered on first sight; uploaded only when the host copy is newer; WRITE access marks the device copy newer
        check(Dl4jB200.dl4jb200_buf_acquire(handle, host, a.length() * 4L, access, dev));
        return new FloatPointer(dev.get(0));
    }

    /** An operand a kernel reads. Host-made arrays (the minibatch, parameters after the updater ran) must have been
     *  declared with hostWrote(); arrays produced by a helper are already current on the device. */
    public FloatPointer in(INDArray a) {
        if (!lazySync) hostWrote(a);  // without the lazy protocol any host array may have changed since the last call
        return device(a, Dl4jB200.BUF_READ);
    }
    public FloatPointer inout(INDArray a) {
        if (!lazySync) hostWrote(a);
        return device(a, Dl4jB200.BUF_READ | Dl4jB200.BUF_WRITE);
    }
    public FloatPointer out(INDArray a) { return device(a, Dl4jB200.BUF_WRITE); }

    /** Host code wrote this array (DataSet features, params.subi(update), a non-helper layer's output). */
    public void hostWrote(INDArray a) {
        if (a != null) check(Dl4jB200.dl4jb200_buf_mark_host_dirty(handle, a.data().addressPointer()));
    }

    /** Host code is about to read this array: copies it back only if the device copy is newer. */
    public INDArray syncToHost(INDArray a) {
        if (a != null) check(Dl4jB200.dl4jb200_buf_sync_to_host(handle, a.data().addressPointer()));
        return a;
    }

    /** Result produced on the device, handed to the layer: brought back now unless lazy synchronisation is on. Gradient
     *  views and running statistics (the parameter boundary) always come back. */
    public INDArray done(INDArray a) { return lazySync ? a : syncToHost(a); }
    public INDArray doneParameter(INDArray a) { return syncToHost(a); }

    /** The INDArray died (workspace closed / garbage collected): its device buffer returns to the pool. */
    public void release(INDArray a) {
        if (a != null) check(Dl4jB200.dl4jb200_buf_release(handle, a.data().addressPointer()));
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
