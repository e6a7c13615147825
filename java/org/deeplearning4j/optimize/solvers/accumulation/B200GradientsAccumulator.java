package org.deeplearning4j.optimize.solvers.accumulation;

import java.util.Queue;
import java.util.concurrent.atomic.AtomicInteger;

import org.bytedeco.javacpp.BytePointer;
import org.bytedeco.javacpp.FloatPointer;
import org.bytedeco.javacpp.Pointer;
import org.bytedeco.javacpp.PointerPointer;
import org.deeplearning4j.b200.B200Context;
import org.deeplearning4j.b200.Dl4jB200;
import org.deeplearning4j.optimize.api.StepFunction;
import org.nd4j.linalg.api.ndarray.INDArray;

/**
 * GradientsAccumulator over NCCL: the plug-in ParallelWrapper installs into every replica's optimizer with
 * {@code new ParallelWrapper.Builder<>(model).trainingMode(TrainingMode.CUSTOM).gradientsAccumulator(new B200GradientsAccumulator(workers))}
 * (ParallelWrapper.java:838-843). StochasticGradientDescent.optimize calls storeUpdate(update) and then
 * applyUpdate(stepFunction, params, update) on every worker thread (StochasticGradientDescent.java:66-79): storeUpdate
 * all-reduces the worker's update in place (sum over replicas, then divided by the replica count: every replica applied
 * minibatch division for its own batch, so the mean of the updates is the update of the concatenated minibatch for a
 * linear updater -- for the adaptive updaters of the zoo configs run the exchange on the raw gradient instead, as
 * deeplearning4j_b200/parallel.py does, by putting this accumulator in front of the updater), applyUpdate is then the
 * plain step. One instance is shared by all worker threads of the JVM; each thread owns its communicator.
 * NOT COMPILED in this repository's image (no JDK): kept in step with include/dl4jb200.h, see INTEGRATION.md.
 */
public class B200GradientsAccumulator implements GradientsAccumulator, Registerable {
    private final int workers;
    private final BytePointer uniqueId = new BytePointer(128);
    private final AtomicInteger nextRank = new AtomicInteger();
    private final ThreadLocal<Pointer> comm = new ThreadLocal<>();

    public B200GradientsAccumulator(int workers) {
        this.workers = workers;
        B200Context.check(Dl4jB200.dl4jb200_comm_unique_id(uniqueId, 128));
    }

    /** DefaultTrainer calls touch() on its own (device-pinned) thread before the first iteration. */
    @Override
    public void touch() {
        if (comm.get() != null) return;
        PointerPointer<Pointer> out = new PointerPointer<>(1);
        B200Context ctx = B200Context.forCurrentThread();
        B200Context.check(Dl4jB200.dl4jb200_comm_init(ctx.handle(), workers, nextRank.getAndIncrement(), uniqueId, out));
        comm.set(out.get(0));
    }

    @Override
    public void storeUpdate(INDArray update) {
        touch();
        B200Context ctx = B200Context.forCurrentThread();
        ctx.hostWrote(update);
        FloatPointer dev = ctx.inout(update);
        B200Context.check(Dl4jB200.dl4jb200_comm_allreduce_avg_f32(comm.get(), dev, update.length(), null));
        ctx.doneParameter(update);
    }

    @Override
    public void applyUpdate(StepFunction function, INDArray params, INDArray updates) {
        function.step(params, updates);
    }

    @Override
    public void applyUpdate(StepFunction function, INDArray params, INDArray updates, double alpha) {
        function.step(params, updates, alpha);
    }

    @Override public void receiveUpdate(INDArray array) { /* nothing is queued: the all-reduce already delivered every replica's share */ }
    @Override public void setExternalSource(Queue<INDArray> source) { /* single box: no external source */ }
    @Override public void reset() { }
    @Override public void registerConsumers(int numConsumers) { }
    @Override public void fallbackToSingleConsumerMode(boolean reallyFallback) { }
}
