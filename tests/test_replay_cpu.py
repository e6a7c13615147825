"""CPU check of the test infrastructure itself: tests/step_replay.py run against a stand-in "network" whose tensors come
from OracleNet. Oracle-vs-oracle must pass with (near) zero error, which pins the replay's graph walking (epsilon fan-in,
flatten aliases, zero-padding, gradient blocks, updater, score) without a GPU; the -m gpu tests then swap in the CUDA path."""
import numpy as np
import torch

from step_replay import replay_step


class _StandIn:
    """Mimics the attributes of deeplearning4j_b200.network.Network that replay_step reads."""

    def __init__(self, model, B, tf32):
        from deeplearning4j_b200 import zoo
        from oracle import oracle as o
        from oracle.net_oracle import OracleNet
        self.model, self.B, self.round_tf32 = model, B, tf32
        self.shapes = zoo.infer_shapes(model, o.out_size)
        self.P, self.layout = zoo.param_layout(model, o.out_size)
        self.pidx = {(n, k): (off, int(np.prod(shp)), shp, pk) for n, k, off, shp, pk in self.layout}
        self.ref = OracleNet(model, self.layout, self.P, zoo.init_params(model, out_size=o.out_size), fast=True, tf32=tf32)
        self.params, self.state, self.grads = (torch.from_numpy(a) for a in (self.ref.params, self.ref.state, self.ref.grads))
        self.drop, self.fused_bn, self.aux, self.trace, self.out = {}, {}, {}, None, {}
        for v in model["vertices"]:
            s = self.shapes[v["name"]]
            alias = v["op"] == "flatten" and self.shapes[v["in"][0]][1] * self.shapes[v["in"][0]][2] == 1
            self.out[v["name"]] = None if alias else 1

    def fit(self, x, lab):
        r = self.ref
        s = r.fit(x, lab)
        self.loss = torch.tensor([r.loss], dtype=torch.float64)
        for n, (mean, std, _) in r.bn_saved.items():
            self.aux[n] = (torch.from_numpy(mean), torch.from_numpy(std))
        if self.trace is not None:
            for n, d in r.deltas.items():
                if self.out.get(n, 0) is None or self.ref.vmap[n]["op"] == "zeropad":
                    continue
                self.trace[n] = torch.from_numpy(np.ascontiguousarray(d.transpose(0, 2, 3, 1)) if d.ndim == 4 else d)
        return torch.tensor([s], dtype=torch.float64)

    def sync(self):
        pass

    def output_host(self, n):
        return self.ref.act[n]


def _data(model, B, K):
    rng = np.random.default_rng(7)
    x = rng.random((B,) + tuple(model["input"]), dtype=np.float32)
    lab = np.zeros((B, K), np.float32)
    lab[:, 1] = 1.0
    return x, lab


def _run(model, B, K, tf32):
    net = _StandIn(model, B, tf32)
    x, lab = _data(model, B, K)
    rep = replay_step(net, model, x, lab)
    assert len(rep.rows) > 10
    return rep


def test_replay_walks_a_residual_graph():
    from test_gpu_network import _tiny_resnet
    for tf32 in (False, True):
        rep = _run(_tiny_resnet(), 4, 10, tf32)
        assert rep.worst("grad.conv_W")[2] < 1e-5 and rep.worst("update.params")[2] < 1e-6


def test_replay_walks_a_sequential_net_with_dense_layers():
    from deeplearning4j_b200 import zoo
    rep = _run(zoo.lenet(), 8, 10, True)
    assert rep.worst("grad.dense_W")[2] < 1e-5


def test_tap_routing_matches_the_pooling_oracle():
    """_gather_taps / _scatter_taps (used where BatchNormalization is folded into its max-pooling consumer) against the oracle's
    own SubsamplingLayer forward / backward, padded and unpadded windows."""
    from oracle import oracle as o
    from step_replay import _gather_taps, _scatter_taps
    rng = np.random.default_rng(3)
    for shape, k, s, pad, mode in (((2, 3, 11, 12), (3, 3), (2, 2), (0, 0), "Truncate"), ((2, 4, 9, 9), (3, 3), (2, 2), (1, 1), "Truncate"),
                                   ((1, 2, 7, 8), (2, 3), (1, 2), (0, 0), "Same")):
        x = rng.standard_normal(shape).astype(np.float32)
        v = {"kernel": k, "stride": s, "pad": pad, "mode": mode}
        m = {"Strict": 0, "Truncate": 1, "Same": 2}[mode]
        y, am = o.pool_fwd(x, k, s, pad, (1, 1), m, o.POOL_MAX, want_argmax=True)
        np.testing.assert_array_equal(_gather_taps(o, x, am, v), y)
        d = rng.standard_normal(y.shape).astype(np.float32)
        want = o.pool_bwd(x, d, k, s, pad, (1, 1), m, o.POOL_MAX)
        np.testing.assert_allclose(_scatter_taps(o, d, am, v, x.shape), want, rtol=0, atol=1e-6)
