"""World-size-2 gloo tests (CPU) of the data-parallel exchange logic in deeplearning4j_b200/parallel.py, plus the
DP parity rule of SURVEY.md §8e: all-reduce(sum) of per-replica raw gradients followed by one updater step with the
global batch size equals single-device training on the concatenated minibatch (for nets without batch statistics)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from deeplearning4j_b200 import parallel, zoo
        from oracle.net_oracle import OracleNet
        # --- exchange_gradients: grads summed, BN running stats averaged, everything else untouched
        m = zoo.resnet50()
        P, layout = zoo.param_layout(m)
        idx = torch.from_numpy(parallel.bn_stat_index(layout))
        assert idx.numel() == 2 * sum(int(np.prod(s)) for n, k, o, s, pk in layout if k == "mean")
        n = 5000  # a window of the flat vector that contains BN stats of the stem
        first = int(idx[0])
        sel = idx[(idx >= first) & (idx < first + n)] - first
        grads = torch.full((n,), float(rank + 1))
        params = torch.arange(n, dtype=torch.float32) * (rank + 1)
        before = params.clone()
        parallel.exchange_gradients(dist, grads, params, sel, world)
        assert torch.equal(grads, torch.full((n,), 3.0))
        exp = before.clone()
        exp[sel] = torch.arange(n, dtype=torch.float32)[sel] * 1.5
        assert torch.equal(params, exp)
        # --- average_parameters (AVERAGING mode)
        p, st = torch.full((10,), float(rank)), torch.full((10,), 2.0 * rank)
        parallel.average_parameters(dist, p, st, world, True)
        assert torch.allclose(p, torch.full((10,), 0.5)) and torch.allclose(st, torch.full((10,), 1.0))
        # --- DP parity: LeNet, each replica its own half of the minibatch
        lm = zoo.lenet()
        LP, ll = zoo.param_layout(lm)
        p0 = zoo.init_params(lm)
        rng = np.random.default_rng(12345)
        x = rng.random((8, 1, 28, 28), dtype=np.float32)
        lab = np.zeros((8, 10), np.float32)
        lab[np.arange(8), np.arange(8) % 10] = 1
        rep = OracleNet(lm, ll, LP, p0)
        xs, ls = x[rank * 4:(rank + 1) * 4], lab[rank * 4:(rank + 1) * 4]
        rep.forward(xs, ls, True)
        rep.backward()
        g = torch.from_numpy(rep.grads)
        parallel.exchange_gradients(dist, g, torch.from_numpy(rep.params), torch.zeros(0, dtype=torch.int64), world)
        rep.update(8)  # global batch size
        full = OracleNet(lm, ll, LP, p0)
        full.forward(x, lab, True)
        full.backward()
        np.testing.assert_allclose(rep.grads, full.grads, rtol=1e-4, atol=1e-5)
        full.update(8)
        np.testing.assert_allclose(rep.params, full.params, rtol=1e-5, atol=1e-6)
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_dp_exchange_world2_gloo():
    from oracle import oracle as o
    o.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_bucket_plan_covers_the_flat_gradient_in_backward_order():
    """plan_buckets: contiguous ranges along vertex boundaries, closing from the end of the flat vector to its start."""
    from deeplearning4j_b200 import parallel, zoo
    from oracle import oracle as o
    for model in (zoo.resnet50(), zoo.vgg16(), zoo.lenet()):
        P, layout = zoo.param_layout(model, o.out_size)
        for nb in (1, 4, 6, 1000):
            close = parallel.plan_buckets(model, layout, P, nb)
            order = [v["name"] for v in reversed(model["vertices"]) if v["name"] in close]
            hi = P
            for name in order:  # in the order backward reaches them
                lo, h = close[name]
                assert h == hi and lo < h
                hi = lo
            assert hi == 0
            assert len(close) <= max(1, min(nb, sum(1 for v in model["vertices"] if v["op"] in ("conv", "bn", "dense", "output"))))
        off, cnt = parallel.bn_stat_spans(layout)
        assert int(cnt.sum()) == parallel.bn_stat_index(layout).size
