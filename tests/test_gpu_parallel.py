"""-m gpu, needs >= 2 GPUs (skipped otherwise): the data-parallel exchange on hardware, through the library's own
dl4jb200_comm_* entry points (NCCL over NVLink) with the bucketed all-reduce overlapped with backward.

Checked, per tests/test_parallel_cpu.py's rule re-asserted on NCCL: (1) after two steps every replica holds bit-identical
parameters, updater state and running statistics; (2) for a net without batch statistics (LeNet) the two replicas, each on
its own half of the minibatch, reproduce single-device training on the concatenated minibatch (the CPU oracle, FP32
arithmetic on both sides) -- gradients and post-update parameters; (3) with BatchNormalization the running mean / var are
the average of the replicas' own statistics; (4) the captured CUDA graph of the step gives the same parameters as the eager
chain. The reference has no numeric multi-replica test (ParallelWrapperTest.java:33-121 is a smoke test)."""
import os
import socket

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    import faulthandler
    import torch
    import torch.distributed as dist
    logdir = os.path.join(ROOT, "gpurun_out")
    logf = open(os.path.join(logdir, f"dp_rank{rank}.log"), "w") if os.path.isdir(logdir) else open(os.devnull, "w")
    faulthandler.dump_traceback_later(150, file=logf, exit=False)  # a hang leaves its stack in the log

    def stage(msg):
        logf.write(msg + "\n")
        logf.flush()

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from deeplearning4j_b200 import _lib, network, parallel, zoo
        from oracle.net_oracle import OracleNet
        from test_gpu_network import _tiny_resnet

        def gathered(t):
            out = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(out, t)
            return out

        # ---- (2) LeNet, FP32 arithmetic: two replicas on half a minibatch each == one device on the whole minibatch
        lm = zoo.lenet()
        LP, ll = zoo.param_layout(lm)
        p0 = zoo.init_params(lm)
        rng = np.random.default_rng(12345)
        x = rng.random((16, 1, 28, 28), dtype=np.float32)
        lab = np.zeros((16, 10), np.float32)
        lab[np.arange(16), np.arange(16) % 10] = 1
        net = network.Network(lm, 8, device=rank, algo=_lib.ALGO_FP32, params=p0)
        stage("lenet: network built")
        pw = parallel.ParallelWrapper(net, buckets=3)
        stage("lenet: wrapper built")
        assert pw.comm is not None and net.dp is pw, "the exchange must go through dl4jb200_comm_*"
        full = OracleNet(lm, ll, LP, p0)
        for it in range(2):
            pw.fit(x[rank * 8:(rank + 1) * 8], lab[rank * 8:(rank + 1) * 8])
            net.sync()
            stage(f"lenet: step {it} done")
            full.forward(x, lab, True)
            full.backward()
            g = net.grads.cpu().numpy()
            scale = np.abs(full.grads).max()
            assert np.abs(g - full.grads).max() <= 1e-4 * scale, ("grads", it, float(np.abs(g - full.grads).max() / scale))
            full.update(16)
            p = net.params.cpu().numpy()
            assert np.abs(p - full.params).max() <= 1e-4 * np.abs(full.params).max(), ("params", it)
        for t in (net.params, net.state):
            a, b = gathered(t)
            assert torch.equal(a, b), "replicas diverged (LeNet)"
        stage("lenet: replicas identical")
        pw.close()
        net.close()
        stage("lenet: closed")

        # ---- (1), (3), (4) tiny ResNet with BatchNormalization, TF32 tensor cores, bucketed exchange, graph replay
        m = _tiny_resnet()
        P, layout = zoo.param_layout(m)
        p0 = zoo.init_params(m)
        rng = np.random.default_rng(7 + rank)  # each replica its own minibatch
        xs = rng.random((8, 3, 33, 33), dtype=np.float32)
        ls = np.zeros((8, 10), np.float32)
        ls[:, 1] = 1
        results = []
        for use_graph in (False, True):
            net = network.Network(m, 8, device=rank, params=p0)
            pw = parallel.ParallelWrapper(net, buckets=4)
            stage(f"resnet graph={use_graph}: wrapper built")
            pw.fit(xs, ls)
            net.sync()
            stage(f"resnet graph={use_graph}: first step done")
            # (3) running statistics = average over replicas of (decay * old + (1-decay) * own batch statistic)
            lone = network.Network(m, 8, device=rank, params=p0)
            lone.set_batch(xs, ls)
            lone.forward(True)
            lone.sync()
            idx = torch.from_numpy(parallel.bn_stat_index(layout)).to(net.dev)
            own = lone.params.index_select(0, idx)
            mean_own = sum(gathered(own)) / world
            got = net.params.index_select(0, idx)
            assert torch.allclose(got, mean_own, rtol=1e-5, atol=1e-6), "running statistics are not the replica average"
            assert not net.grads.index_select(0, idx).any(), "statistic slots of the gradient must be zero again"
            lone.close()
            stage(f"resnet graph={use_graph}: statistics checked")
            if use_graph:
                net.capture(hook=pw.hook, include_input=True)
                for _ in range(2):
                    net.fit_graph(xs, ls)
            else:
                for _ in range(2):
                    pw.fit(xs, ls)
            net.sync()
            stage(f"resnet graph={use_graph}: steps done")
            for t in (net.params, net.state):
                a, b = gathered(t)
                assert torch.equal(a, b), "replicas diverged (tiny ResNet, graph=%s)" % use_graph
            results.append(net.params.clone())
            assert torch.isfinite(net.params).all()
            stage(f"resnet graph={use_graph}: replicas identical")
            if not use_graph:  # (a communicator captured into a CUDA graph blocks its own teardown: left to process exit)
                pw.close()
                net.close()
        assert torch.equal(results[0], results[1]), "graph replay of the data-parallel step differs from the eager chain"
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        q.close()
        q.join_thread()  # the result must have left this process before the hard exit below
        try:
            torch.cuda.synchronize()
        except Exception:  # noqa: BLE001
            pass
        logf.flush()
        os._exit(0)  # a communicator that was captured into a CUDA graph blocks the collective teardown


def test_dp_two_gpus_nccl_through_c_abi(oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=200) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
