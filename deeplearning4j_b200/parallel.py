"""Data-parallel training over the GPUs of one box: the ParallelWrapper hot path.

The reference runs one worker thread + model replica per device and exchanges through the host: parameter
averaging (`Nd4j.averageAndPropagate`, deeplearning4j-scaleout/.../parallelism/ParallelWrapper.java:316-376) or
threshold-encoded update sharing (nn/optimize/solvers/accumulation/EncodedGradientsAccumulator.java:244-315,
440-520). Here a replica is one process per GPU (torchrun) and the exchange is one NCCL collective over
NVLink/NVSwitch on the flat [1,P] vectors, issued on the replica's own CUDA stream -- this class is what a
`GradientsAccumulator` plugged in through `TrainingMode.CUSTOM` (ParallelWrapper.java:838-843) does.

Modes
  "ALLREDUCE" (default): replicas all-reduce(sum) the raw flat gradient after backward and run the identical
      updater with batchSize = workers * B -- single-device training on the concatenated minibatch, except that
      BatchNormalization statistics stay per replica; the running mean/var (parameters with a NoOp updater,
      BatchNormalization.java:205-208) are all-reduce-averaged so replicas stay bit-identical.
  "AVERAGING": reference-faithful parameter averaging: every `averaging_frequency` local steps all-reduce(avg)
      the parameters and, if `average_updaters`, the updater state (ParallelWrapper.java:326,353,370).
"""
import numpy as np


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

    def __init__(self, net, mode="ALLREDUCE", averaging_frequency=1, average_updaters=True):
        import torch
        import torch.distributed as dist
        assert dist.is_initialized(), "torch.distributed must be initialised (one process per GPU)"
        assert mode in ("ALLREDUCE", "AVERAGING")
        self.net, self.mode, self.dist, self.torch = net, mode, dist, torch
        self.world = dist.get_world_size()
        self.freq, self.avg_upd, self.iter = int(averaging_frequency), bool(average_updaters), 0
        self.stat_idx = torch.from_numpy(bn_stat_index(net.layout)).to(net.dev)
        if mode == "ALLREDUCE":
            net.batch_div_scale = self.world  # the updater divides by the global minibatch size
        # replicas start from identical parameters (DefaultTrainer.java:254-311 copies params into each replica)
        with torch.cuda.stream(net.stream):
            dist.broadcast(net.params, src=0)
            dist.broadcast(net.state, src=0)
        net.params_changed()

    def _hook(self, net):
        with self.torch.cuda.stream(net.stream):
            exchange_gradients(self.dist, net.grads, net.params, self.stat_idx, self.world)

    def fit(self, features=None, labels=None):
        net = self.net
        if self.mode == "ALLREDUCE":
            return net.fit(features, labels, hook=self._hook)
        score = net.fit(features, labels)
        self.iter += 1
        if self.iter % self.freq == 0:
            with self.torch.cuda.stream(net.stream):
                average_parameters(self.dist, net.params, net.state, self.world, self.avg_upd)
            net.params_changed()
        return score
