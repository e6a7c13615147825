"""Data-parallel training over the GPUs of one box: the ParallelWrapper hot path.

The reference runs one worker thread + model replica per device and exchanges through the host: parameter
averaging (`Nd4j.averageAndPropagate`, deeplearning4j-scaleout/.../parallelism/ParallelWrapper.java:316-376) or
threshold-encoded update sharing (nn/optimize/solvers/accumulation/EncodedGradientsAccumulator.java:244-315,
440-520). Here a replica is one process per GPU (torchrun) and the exchange is one NCCL collective over
NVLink/NVSwitch on the flat [1,P] vectors, issued on the replica's own CUDA stream -- this class is what a
`GradientsAccumulator` plugged in through `TrainingMode.CUSTOM` (ParallelWrapper.java:838-843) does.

On CUDA the collective is issued through the library's own C ABI (dl4jb200_comm_*: NCCL bound at run time, enqueued on
a CUDA stream) -- the entry points a Java GradientsAccumulator would call -- and is overlapped with the backward pass:
the flat gradient is laid out in layer order (MultiLayerNetwork.java:691-727) and backward fills it from the END, so it is
cut into a few buckets that are all-reduced on a communication stream as soon as their last gradient block is final.
torch.distributed only bootstraps (it carries the 128-byte NCCL id to the other ranks); on CPU (gloo, the unit tests)
the same arithmetic runs through torch.distributed itself.

Modes
  "ALLREDUCE" (default): replicas all-reduce(sum) the raw flat gradient after backward and run the identical
      updater with batchSize = workers * B -- single-device training on the concatenated minibatch, except that
      BatchNormalization statistics stay per replica; the running mean/var (parameters with a NoOp updater,
      BatchNormalization.java:205-208) are all-reduce-averaged so replicas stay bit-identical.
  "AVERAGING": reference-faithful parameter averaging: every `averaging_frequency` local steps all-reduce(avg)
      the parameters and, if `average_updaters`, the updater state (ParallelWrapper.java:326,353,370).
"""
import ctypes as C

import numpy as np


def bn_stat_spans(layout):
    """(offsets, counts) of the BatchNormalization running mean|var blocks (adjacent in the flat vector) -> merged spans."""
    spans = []
    for name, key, off, shp, pk in layout:
        if pk == "bn_stat":
            n = int(np.prod(shp))
            if spans and spans[-1][0] + spans[-1][1] == off:
                spans[-1][1] += n
            else:
                spans.append([off, n])
    return np.array([s[0] for s in spans], np.int64), np.array([s[1] for s in spans], np.int64)


def plan_buckets(model, layout, P, nbuckets):
    """Cuts the flat gradient into <= nbuckets contiguous ranges along vertex boundaries, in the order backward completes
    them (last layer first). -> {vertex name: (lo, hi)}: the range that is final once backward has processed that vertex."""
    spans = {}
    for name, key, off, shp, pk in layout:
        lo, hi = spans.get(name, (off, off))
        spans[name] = (min(lo, off), max(hi, off + int(np.prod(shp))))
    order = [v["name"] for v in reversed(model["vertices"]) if v["name"] in spans]
    target = max(1, -(-P // max(1, nbuckets)))
    close, hi_open = {}, P
    for i, name in enumerate(order):
        lo = spans[name][0]
        if hi_open - lo >= target or i == len(order) - 1:
            close[name] = (lo if i < len(order) - 1 else 0, hi_open)
            hi_open = lo
    return close


def bn_stat_index(layout):
    """Flat-vector positions of every BatchNormalization running mean/var element."""
    idx = []
    for name, key, off, shp, pk in layout:
        if pk == "bn_stat":
            idx.append(np.arange(off, off + int(np.prod(shp)), dtype=np.int64))
    return np.concatenate(idx) if idx else np.zeros(0, np.int64)


def _avg(dist, t, world):
    """all-reduce average (NCCL has a native AVG; gloo -- used by the CPU tests -- does not)."""
    if dist.get_backend() == "nccl":
        dist.all_reduce(t, op=dist.ReduceOp.AVG)
    else:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t.div_(world)


def exchange_gradients(dist, grads, params, stat_idx, world):
    """ALLREDUCE mode exchange on flat tensors (device-agnostic so it is testable with gloo on CPU)."""
    dist.all_reduce(grads, op=dist.ReduceOp.SUM)
    if stat_idx is not None and stat_idx.numel():
        stats = params.index_select(0, stat_idx)
        _avg(dist, stats, world)
        params.index_copy_(0, stat_idx, stats)


def average_parameters(dist, params, state, world, average_updaters=True):
    """AVERAGING mode exchange (Nd4j.averageAndPropagate on params and updater state)."""
    _avg(dist, params, world)
    if average_updaters and state is not None:
        _avg(dist, state, world)


class ParallelWrapper:
    """ParallelWrapper.Builder(model).workers(world).averagingFrequency(f).averageUpdaters(b).build()"""

    def __init__(self, net, mode="ALLREDUCE", averaging_frequency=1, average_updaters=True, buckets=6, use_c_abi=None):
        import torch
        import torch.distributed as dist
        assert dist.is_initialized(), "torch.distributed must be initialised (one process per GPU)"
        assert mode in ("ALLREDUCE", "AVERAGING")
        self.net, self.mode, self.dist, self.torch = net, mode, dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.freq, self.avg_upd, self.iter = int(averaging_frequency), bool(average_updaters), 0
        self.stat_idx = torch.from_numpy(bn_stat_index(net.layout)).to(net.dev)
        self.use_c_abi = (dist.get_backend() == "nccl") if use_c_abi is None else bool(use_c_abi)
        self.comm = None
        if self.use_c_abi:
            self._init_comm(buckets)
        if mode == "ALLREDUCE":
            net.batch_div_scale = self.world  # the updater divides by the global minibatch size
        # replicas start from identical parameters (DefaultTrainer.java:254-311 copies params into each replica)
        if self.comm is not None:
            for t in (net.params, net.state):
                _lib().call("dl4jb200_comm_broadcast_f32", self.comm, C.c_void_p(t.data_ptr()), t.numel(), 0, None)
        else:
            with torch.cuda.stream(net.stream):
                dist.broadcast(net.params, src=0)
                dist.broadcast(net.state, src=0)
        net.params_changed()
        # ALLREDUCE through the C ABI is driven from inside backward (bucketed, overlapped); otherwise a post-backward hook
        self.hook = None if (self.comm is not None and mode == "ALLREDUCE") else (self._hook if mode == "ALLREDUCE" else None)
        if self.comm is not None and mode == "ALLREDUCE":
            net.dp = self

    # ------------------------------------------------------------------ C-ABI communicator
    def _init_comm(self, buckets):
        t, net, L = self.torch, self.net, _lib()
        ident = (C.c_uint8 * 128)()
        if self.rank == 0:
            L.call("dl4jb200_comm_unique_id", ident, 128)
        box = [bytes(ident)]
        self.dist.broadcast_object_list(box, src=0)  # bootstrap only: the id travels over the existing process group
        ident = (C.c_uint8 * 128).from_buffer_copy(box[0])
        comm = C.c_void_p()
        L.call("dl4jb200_comm_init", net.h, self.world, self.rank, ident, C.byref(comm))
        self.comm = comm
        self.comm_stream = t.cuda.Stream(net.dev)
        off, cnt = bn_stat_spans(net.layout)
        self.n_spans = int(off.size)
        self.span_off, self.span_cnt = t.from_numpy(off).to(net.dev), t.from_numpy(cnt).to(net.dev)
        self.close_at = plan_buckets(net.model, net.layout, net.P, buckets)
        t.cuda.synchronize(net.dev)

    def _allreduce(self, tensor, lo, hi, avg=False, stream=None):
        fn = "dl4jb200_comm_allreduce_avg_f32" if avg else "dl4jb200_comm_allreduce_sum_f32"
        _lib().call(fn, self.comm, C.c_void_p(tensor.data_ptr() + 4 * lo), hi - lo,
                    C.c_void_p(stream.cuda_stream) if stream is not None else None)

    # callbacks from Network.backward (ALLREDUCE mode)
    def backward_begin(self, net):
        # the running statistics were updated by the forward pass: park them, pre-divided by the replica count, in their
        # own (identically zero) gradient slots so that the bucket all-reduces below average them as well
        if self.n_spans:
            net._call("dl4jb200_spans_scale_copy", net._ptr(net.grads), net._ptr(net.params), net._ptr(self.span_off),
                      net._ptr(self.span_cnt), self.n_spans, 1.0 / self.world)

    def vertex_done(self, net, name):
        rng = self.close_at.get(name)
        if rng is None:
            return
        t = self.torch
        ev = t.cuda.Event()
        ev.record(net.stream)              # batch-norm parameter gradients, the parked statistics
        self.comm_stream.wait_event(ev)
        if net._side_used:
            ev2 = t.cuda.Event()
            ev2.record(net.side_stream)    # filter / bias gradients
            self.comm_stream.wait_event(ev2)
        self._allreduce(net.grads, rng[0], rng[1], stream=self.comm_stream)

    def backward_end(self, net):
        t = self.torch
        done = t.cuda.Event()
        done.record(self.comm_stream)
        net.stream.wait_event(done)
        if self.n_spans:  # averaged statistics back into the parameters; their gradient slots are zero again
            net._call("dl4jb200_spans_scale_copy", net._ptr(net.params), net._ptr(net.grads), net._ptr(self.span_off),
                      net._ptr(self.span_cnt), self.n_spans, 1.0)
            net._call("dl4jb200_spans_scale_copy", net._ptr(net.grads), net._ptr(net.grads), net._ptr(self.span_off),
                      net._ptr(self.span_cnt), self.n_spans, 0.0)

    # ------------------------------------------------------------------ torch.distributed path (gloo on CPU; opt-out on CUDA)
    def _hook(self, net):
        with self.torch.cuda.stream(net.stream):
            exchange_gradients(self.dist, net.grads, net.params, self.stat_idx, self.world)

    def fit(self, features=None, labels=None):
        net = self.net
        if self.mode == "ALLREDUCE":
            return net.fit(features, labels, hook=self.hook)
        score = net.fit(features, labels)
        self.iter += 1
        if self.iter % self.freq == 0:
            if self.comm is not None:
                self._allreduce(net.params, 0, net.P, avg=True)
                if self.avg_upd:
                    self._allreduce(net.state, 0, net.state.numel(), avg=True)
            else:
                with self.torch.cuda.stream(net.stream):
                    average_parameters(self.dist, net.params, net.state, self.world, self.avg_upd)
            net.params_changed()
        return score

    def close(self):
        if self.comm is not None:
            self.net.dp = None
            self.torch.cuda.synchronize(self.net.dev)
            _lib().call("dl4jb200_comm_destroy", self.comm)
            self.comm = None


def _lib():
    from . import _lib as L
    return L
