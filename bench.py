#!/usr/bin/env python
"""bench.py -- ResNet-50 (DL4J zoo variant) training images/sec on N B200s, BASELINE.json's metric.

  python bench.py --gpus N --steps K --warmup W            this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference ...                     the reference's CPU algorithm (im2col + BLAS sgemm + col2im,
                                                           multi-pass BN/pool; oracle/ -- the reference itself needs a JVM
                                                           and un-vendored nd4j, unavailable on this box) on the host cores

A step = one fit() iteration (forward, backward, score, updater, params -= update) on one synthetic minibatch
(BenchmarkDataSetIterator semantics: one fixed U[0,1) feature tensor, every label = class 1). `value` is measured
with the minibatch already resident in HBM; `e2e` goes through Network.fit(host features, host labels) with the
pinned host->device copy and the device->host read of the score inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# CPU legs (the reference arm, cpu_baseline): the oracle's C loops are OpenMP, its GEMMs are numpy's OpenBLAS (own
# thread pool). OpenMP workers that spin after a parallel region fight the BLAS threads for the same cores -- measured
# here: 4.7 images/s spinning vs 8.4 passive on 8 cores -- so waiting threads sleep; and torchrun's OMP_NUM_THREADS=1 is
# lifted for the rank that runs the CPU arm. Must be set before libgomp / OpenBLAS initialise (i.e. before numpy).
os.environ["OMP_WAIT_POLICY"] = "PASSIVE"
if "reference" in sys.argv or int(os.environ.get("WORLD_SIZE", "1")) == 1:
    os.environ["OMP_NUM_THREADS"] = str(_host_threads())
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(min(_host_threads(), 64)))

import numpy as np  # noqa: E402

# model -> (per-GPU batch, labels, reference-arm sample batch, workload text); batches are BASELINE.json's configs
CONFIGS = {
    "resnet50": (256, 1000, 32, "ResNet-50 (DL4J zoo ComputationGraph, BatchNormalization) 3x224x224 full training step "
                                "(forward, backward, score, RmsProp+L1/L2 updater)"),
    "vgg16": (64, 1000, 8, "VGG16 (DL4J zoo, 13 conv3x3+ReLU, 5 max-pool, softmax 25088->1000) 3x224x224 full training step "
                           "(forward, backward, score, Nesterovs updater)"),
    "alexnet": (128, 1000, 32, "AlexNet (DL4J zoo) 3x224x224 full training step (forward with inverted dropout 0.5 on every "
                               "layer input, backward, score, RenormalizeL2PerLayer + Nesterovs + L2 updater)"),
    "lenet": (64, 10, 64, "LeNet (DL4J zoo) 1x28x28 full training step (forward, backward, score, AdaDelta updater)"),
}


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}
    f = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(f):
        try:
            j = json.load(open(f))
            p.update({k: j[k] for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in j})
            p["src"] = "measured"
        except Exception:
            pass
    return p


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def synthetic(model, B, K):
    rng = np.random.default_rng(12345)
    x = rng.random((B,) + tuple(model["input"]), dtype=np.float32)
    lab = np.zeros((B, K), np.float32)
    lab[:, 1] = 1.0
    return x, lab


def conv_flops(net):
    """{vertex: (fwd_flops, bwd_flops)} algorithmic FLOPs 2*B*Ho*Wo*Cout*Cin*kH*kW per pass."""
    out = {}
    first = net.model["vertices"][0]["name"]
    for v in net.model["vertices"]:
        if v["op"] == "conv":
            src, d = net._conv_desc(v)
            f = 2.0 * d.batch * d.out_h * d.out_w * d.c_out * d.c_in * d.k_h * d.k_w
            out[v["name"]] = (f, f * (1 if src == first else 2))
        elif v["op"] in ("dense", "output"):
            d = net._dense_desc(v)
            f = 2.0 * d.batch * d.c_out * d.c_in
            out[v["name"]] = (f, 2 * f)
    return out


def conv_alg_bytes(net):
    """Algorithmic DRAM bytes of all convolution / dense passes of one step: forward reads x and writes z; backward reads
    x and delta for the filter gradient and delta again, writing dx, for the data gradient (filters are negligible)."""
    B, tot = net.B, 0.0
    for v in net.model["vertices"]:
        if v["op"] not in ("conv", "dense", "output"):
            continue
        so, si = net.shapes[v["name"]], net.shapes[v["in"][0]]
        nout, nin = float(B) * float(np.prod(so)), float(B) * float(np.prod(si))
        first = net.vmap[v["in"][0]]["op"] in ("input", "zeropad")
        tot += 4.0 * (nin + nout) * (2 if first else 3)
    return tot


def blas_threads():
    try:
        from threadpoolctl import threadpool_info
        return {f"{i.get('user_api')}:{i.get('internal_api')}": i.get("num_threads") for i in threadpool_info()}
    except Exception:  # noqa: BLE001
        return {}


def make_masks(model, shapes, B, seed=777):
    """Host-generated 0/1 masks for every vertex with a dropout retain probability (nd4j's DropOutInverted stream cannot be
    reproduced outside nd4j, so the masks are inputs of the step; the bench reuses one fixed set)."""
    rng = np.random.default_rng(seed)
    masks = {}
    for v in model["vertices"]:
        p = v.get("dropout")
        if p:
            s = shapes[v["in"][0]]
            masks[v["name"]] = (rng.random((B,) + tuple(s)) < p).astype(np.float32)
    return masks or None


def cpu_reference_step(model_name, Bs, steps, warm):
    """Times `steps` full fit() iterations of the reference's CPU algorithm (oracle/: im2col + BLAS sgemm + col2im, the
    multi-pass BN / pooling, updater) at batch Bs on the host cores. Imports nothing of the CUDA library."""
    from deeplearning4j_b200 import zoo  # model tables only: shape arithmetic is routed to the oracle below
    from oracle import oracle as o
    from oracle.net_oracle import OracleNet
    o.build()
    labels = CONFIGS[model_name][1]
    m = zoo.MODELS[model_name](dropout=0.5) if model_name == "alexnet" else zoo.MODELS[model_name]()
    P, layout = zoo.param_layout(m, o.out_size)
    x, lab = synthetic(m, Bs, labels)
    p0 = zoo.init_params(m, out_size=o.out_size)
    net = OracleNet(m, layout, P, p0, fast=True)
    net.masks = make_masks(m, zoo.infer_shapes(m, o.out_size), Bs)
    # Every step starts from the initial parameters: after one RmsProp(0.1) update of this untrained net most activations are
    # denormal, and x86 denormal arithmetic made later steps 9x slower here (4 s -> 38 s at batch 32) -- a value-dependent
    # CPU artefact the GPU does not have. Resetting keeps the CPU number reproducible and errs in the CPU's favour.
    total = 0.0
    for i in range(warm + steps):
        net.params[:] = p0
        net.state[:] = 0.0
        t0 = time.perf_counter()
        net.fit(x, lab)
        if i >= warm:
            total += time.perf_counter() - t0
    return total / steps


def run_reference(args, rank, world):
    """The reference arm: the reference's CPU algorithm for the same step, timed on the host cores (rank 0 only)."""
    if rank != 0:
        return
    B, labels, Bs, workload = CONFIGS[args.model]
    Bs = min(Bs, args.batch or B)
    steps, warm = max(1, min(args.steps, 3)), 1
    dt = cpu_reference_step(args.model, Bs, steps, warm)
    v = Bs / dt
    cores = _host_threads()
    print(json.dumps({
        "impl": "reference", "metric": f"{args.model}_train_images_per_sec", "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "global_batch": (args.batch or B) * world, "per_gpu_batch": args.batch or B, "sample_batch": Bs,
                   "sample": f"reference CPU algorithm on the host cores; each timed step is one full fit() of the same network at "
                             f"batch {Bs} (a bounded sample of the batch-{args.batch or B} step: images/s = {Bs} / step time; every step starts from the "
                             f"initial parameters so that x86 denormal slow-downs of later steps do not handicap the CPU)"},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": "port", "threads": blas_threads(),
                         "sample": f"{steps} full fit() iterations at batch {Bs} after {warm} warm-up (im2col + OpenBLAS sgemm + col2im; "
                                   f"OpenMP loops on {os.environ.get('OMP_NUM_THREADS')} threads, passive wait policy)"},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def tf32_peak(pk):
    """(TFLOP/s, note): the TF32 dense roof. Measured on this pool's B200s by scripts/measure_tf32_peak.py (torch.matmul with
    allow_tf32 = cuBLAS, 8192^3, burst and 4 s sustained -- the same method MEASURED_PEAKS.json uses for bf16) and committed
    under profiles/; until that file exists the provisional 0.5 x bf16 sustained figure is used. Both are reported."""
    half = pk["bf16_tflops_sustained"] / 2.0
    f = os.path.join(ROOT, "profiles", "r02_tf32_peak.json")
    if os.path.exists(f):
        try:
            j = json.load(open(f))
            return float(j["tf32_tflops_sustained"]), (f"measured cuBLAS TF32 8192^3 sustained (profiles/r02_tf32_peak.json; burst "
                                                       f"{j['tf32_tflops']:.0f}); 0.5 x bf16 sustained would be {half:.0f}")
        except Exception:  # noqa: BLE001
            pass
    return half, f"TF32 dense = 0.5 x bf16_tflops_sustained ({pk['src']}), provisional: no measured TF32 peak file"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--model", default="resnet50", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: BASELINE.json's for the model)")
    ap.add_argument("--algo", default="tf32", choices=["tf32", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time the eager kernel chain instead of the captured CUDA graph")
    ap.add_argument("--dump-profile", default=None, help="write the per-layer event timings to this JSON file")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(args.warmup, 3)
    # stdout carries exactly one JSON line: native libraries (the NCCL version banner) write to fd 1 directly, so fd 1 is
    # pointed at stderr for the run and the line goes to a private duplicate of the original stdout
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from deeplearning4j_b200 import _lib, network, parallel, zoo
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B0, LABELS, ref_batch, WORKLOAD = CONFIGS[args.model]
    m = zoo.MODELS[args.model](dropout=0.5) if args.model == "alexnet" else zoo.MODELS[args.model]()
    B = args.batch or B0
    algo = _lib.ALGO_TC_TF32 if args.algo == "tf32" else _lib.ALGO_FP32
    net = network.Network(m, B, device=local, algo=algo)
    masks = make_masks(m, net.shapes, B)
    if masks:
        net.set_dropout_masks(masks)
    pw = parallel.ParallelWrapper(net) if world > 1 else None
    step = (lambda f=None, l=None: pw.fit(f, l)) if pw else (lambda f=None, l=None: net.fit(f, l))
    x, lab = synthetic(m, B, LABELS)
    xh, labh = torch.from_numpy(x).pin_memory(), torch.from_numpy(lab).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also grows the library's scratch to its final size), then capture the step in a CUDA graph
    net.set_batch(xh, labh)
    for _ in range(args.warmup):
        step()
    barrier()
    use_graph = not args.no_graph
    if use_graph:
        try:
            net.capture(hook=pw.hook if pw else None, include_input=False)
            for _ in range(2):
                net._relayout_input()
                net.fit_graph()
            barrier()
        except Exception as e:  # noqa: BLE001 -- capture is an optimisation: report and time the eager chain instead
            if rank == 0:
                print(f"graph capture failed ({type(e).__name__}: {e}); timing the eager chain", file=sys.stderr)
            use_graph = False

    def run():
        # one step on the resident minibatch: the NCHW -> NHWC (row-packed) re-layout of the features is part of it
        if use_graph:
            net._relayout_input()
            net.fit_graph()
        else:
            net._relayout_input()
            step()

    # ---- timed region 1: device-resident minibatch (CUDA events on the launching stream)
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(net.stream)
    for _ in range(args.steps):
        run()
    e1.record(net.stream)
    barrier()
    ms = e0.elapsed_time(e1)

    # ---- timed region 2: end to end through the public API with host buffers
    hook = pw.hook if pw else None
    for _ in range(2):  # untimed: the double buffers of the input pipeline are allocated on first use
        net.prefetch(xh, labh)
        net.fit_prefetched(hook=hook)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    f0.record(net.stream)
    last = None
    net.prefetch(xh, labh)  # every step's pinned host->device copy happens inside the timed region ...
    for k in range(args.steps):
        if k + 1 < args.steps:
            net.prefetch(xh, labh)  # ... overlapped with the previous step on a copy stream (AsyncDataSetIterator-style)
        s = net.fit_prefetched(hook=hook)
        net.sync()
        last = float(s.cpu()[0])  # device -> host read of the step's score
    f1.record(net.stream)
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- data-parallel check: after the timed steps every replica must hold bit-identical parameters
    dp_check = None
    if world > 1:
        with torch.cuda.stream(net.stream):
            pv = net.params.view(torch.int32)
            chk = torch.stack([pv.sum(dtype=torch.int64), (pv.to(torch.int64) * 31 % 1000003).sum()])
            lo, hi = chk.clone(), chk.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        net.sync()
        dp_check = {"replicas_bit_identical": bool((lo == hi).all().item()), "ranks": world,
                    "how": "all-reduce MIN == MAX of two integer checksums of the flat parameter vector after the timed steps"}

    # what the box's host link gives a lone pinned copy of one minibatch (explains e2e on boxes where it is below
    # bytes / step time: the prefetch then cannot hide the copy)
    dev_x = torch.empty_like(xh, device=net.dev)
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(net.stream):
        dev_x.copy_(xh, non_blocking=True)
        c0.record(net.stream)
        for _ in range(3):
            dev_x.copy_(xh, non_blocking=True)
        c1.record(net.stream)
    net.sync()
    h2d_gbs = 3 * x.nbytes / (c0.elapsed_time(c1) * 1e-3) / 1e9
    del dev_x

    # ---- kernel-level pass: the same step issued eagerly with a CUDA-event pair around every conv / BN / pool call
    # (a replayed graph cannot carry per-kernel events); feeds the roofline and the launch count. The captured graph is
    # dropped first: on one stream the filter gradients use the main context's scratch, which may have to grow.
    if use_graph:
        net.release_graph()
    net.overlap_wgrad = False  # one stream: every call is timed alone (in the timed regions above the filter gradients overlap the backward chain)
    net.profile(["dl4jb200_conv_fwd", "dl4jb200_conv_fwd_bnstats", "dl4jb200_conv_bwd", "dl4jb200_conv_bwd_data_fused", "dl4jb200_conv_bwd_data_bnstats",
                 "dl4jb200_conv_rowpack_fwd", "dl4jb200_conv_rowpack_wgrad", "dl4jb200_conv_rowpack_fwd_bnstats", "dl4jb200_bn_fwd",
                 "dl4jb200_bn_fwd_prestats", "dl4jb200_bn_bwd", "dl4jb200_bn_bwd_prestats", "dl4jb200_pool_fwd", "dl4jb200_pool_bwd",
                 "dl4jb200_bn_fwd_prestats_maxpool", "dl4jb200_pool_bwd_bnstats"])
    psteps = min(args.steps, 10)
    step()  # untimed: lets the single-stream variant size its scratch
    barrier()
    net.prof = []
    l0 = _lib.cdll().dl4jb200_launch_count()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    p0.record(net.stream)
    for _ in range(psteps):
        step()
    p1.record(net.stream)
    barrier()
    ms_eager = p0.elapsed_time(p1) / psteps
    launches = (_lib.cdll().dl4jb200_launch_count() - l0) // psteps * args.steps
    prof = net.profile_results()
    net.prof = None

    t = torch.tensor([ms, ms_e2e], device=net.dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])

    def finish():
        # NCCL communicators referenced by a captured CUDA graph make destroy_process_group() block: drop the graph,
        # drain, flush and leave without the collective teardown
        sys.stdout.flush()
        sys.stderr.flush()
        if world > 1:
            net._graph = None
            torch.cuda.synchronize()
            os._exit(0)

    if rank != 0:
        return finish()

    pk = peaks()
    total_imgs = B * world * args.steps
    value, e2e = total_imgs / (ms / 1e3), total_imgs / (ms_e2e / 1e3)
    # ---- roofline of the dominant kernel family: the tcgen05 implicit-GEMM convolutions
    fl = conv_flops(net)
    conv_ms = conv_fl = 0.0
    n_conv_calls = 0
    per_layer = {}
    conv_fns = ("dl4jb200_conv_fwd", "dl4jb200_conv_fwd_bnstats", "dl4jb200_conv_bwd", "dl4jb200_conv_bwd_data_fused", "dl4jb200_conv_bwd_data_bnstats",
                "dl4jb200_conv_rowpack_fwd", "dl4jb200_conv_rowpack_fwd_bnstats", "dl4jb200_conv_rowpack_wgrad")
    passes = {}  # (layer, "fwd"|"bwd") -> [ms over psteps, calls]: a layer's backward is several calls (filter / data gradient)
    for (fn, tag), times in prof.items():
        if fn in conv_fns and tag in fl:
            e = passes.setdefault((tag, "fwd" if "_fwd" in fn else "bwd"), [0.0, 0])
            e[0] += sum(times)
            e[1] += len(times)
    for (tag, ps), (tms, calls) in passes.items():
        f = fl[tag][0 if ps == "fwd" else 1]
        conv_ms += tms
        conv_fl += f * psteps
        n_conv_calls += calls
        per_layer[f"{tag}:conv_{ps}"] = {"ms": tms / psteps, "tflops": f / (tms / psteps * 1e-3) / 1e12}
    # DRAM bytes of the same kernel family from the committed ncu capture (cold-cache, serialised): per call, like `achieved`
    traffic, traffic_note = None, "no ncu capture for this model"
    if args.model == "resnet50" and B == 256:
        for name in ("r02_conv_traffic.json", "r01_conv_traffic.json"):
            try:
                tr = json.load(open(os.path.join(ROOT, "profiles", name)))
                traffic = (tr["dram_read_bytes"] + tr["dram_write_bytes"]) / max(1, n_conv_calls // psteps)
                traffic_note = ("dram__bytes_read.sum + dram__bytes_write.sum of the %d convolution-family launches of one step "
                                "(profiles/%s, batch 256) / calls per step; algorithmic bytes per step: %.2f GB"
                                % (tr["launches"], name, conv_alg_bytes(net) / 1e9))
                break
            except Exception:  # noqa: BLE001
                pass
    roof, roof_note = tf32_peak(pk)
    ach = conv_fl / (conv_ms * 1e-3) / 1e12 if conv_ms else 0.0
    roofline = {"bound": "tensor", "kernel": "k_conv_tc_gemm_tma_persist / k_conv_tc_wgrad_tma (all convolution + dense calls of the step)",
                "achieved": ach, "peak": roof, "unit": "TFLOP/s", "frac": ach / roof, "traffic": traffic,
                "traffic_note": traffic_note, "peak_note": roof_note + "; kernel timed inside a long step",
                "frac_of_half_bf16_sustained": ach / (pk["bf16_tflops_sustained"] / 2.0),
                "share_of_step": conv_ms / (ms_eager * psteps), "calls_per_step": n_conv_calls // psteps,
                "timing": "CUDA events per call in an eager single-stream pass of the same step (%.2f ms/step eager serial vs %.2f ms/step graph with overlapped filter gradients)" % (ms_eager, ms / args.steps)}
    # ---- HBM kernels (BN / pool): achieved GB/s on algorithmic bytes
    bw = {}
    # (with the statistics taken in the convolution epilogue the forward is finalize + apply only: same 8 algorithmic B/elem)
    for key, fns, bpe in (("bn_fwd", ("dl4jb200_bn_fwd", "dl4jb200_bn_fwd_prestats"), 8), ("bn_bwd", ("dl4jb200_bn_bwd", "dl4jb200_bn_bwd_prestats"), 12)):
        tot_ms = tot_b = 0.0
        for (f2, tag), times in prof.items():
            if f2 in fns:
                sh = net.shapes[tag if tag in net.shapes and net.vmap[tag]["op"] == "bn" else net.fused_add.get(tag, (tag,))[0]]
                tot_ms += sum(times)
                tot_b += float(B) * sh[0] * sh[1] * sh[2] * bpe * len(times)
            elif key == "bn_fwd" and f2 == "dl4jb200_bn_fwd_prestats_maxpool":
                # BN + ReLU + max-pool in one pass: reads the BN's input, writes the pooled tensor and its taps
                sh, so = net.shapes[tag], net.shapes[net.pool_fuse[tag]]
                tot_ms += sum(times)
                tot_b += (4.0 * B * sh[0] * sh[1] * sh[2] + 8.0 * B * so[0] * so[1] * so[2]) * len(times)
        if tot_ms:
            bw[key] = {"achieved_gbs": tot_b / (tot_ms * 1e-3) / 1e9, "frac_of_hbm": tot_b / (tot_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                       "share_of_step": tot_ms / (ms_eager * psteps)}
    for fn in ("dl4jb200_pool_fwd", "dl4jb200_pool_bwd"):
        tot_ms = tot_b = 0.0
        for (f2, tag), times in prof.items():
            if f2 == fn or (fn == "dl4jb200_pool_bwd" and f2 == "dl4jb200_pool_bwd_bnstats"):
                so, si = net.shapes[tag], net.shapes[net.vmap[tag]["in"][0]]
                n_in, n_out = float(B) * si[0] * si[1] * si[2], float(B) * so[0] * so[1] * so[2]
                tot_ms += sum(times)
                # (the statistics-carrying backward also reads the batch norm's input once: 4 more bytes per input element)
                tot_b += (4 * (n_in + n_out) + 4 * n_out + (4 * n_in if f2 == "dl4jb200_pool_bwd_bnstats" else 0)) * len(times)
        if tot_ms:
            bw[fn[9:]] = {"achieved_gbs": tot_b / (tot_ms * 1e-3) / 1e9, "frac_of_hbm": tot_b / (tot_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                          "share_of_step": tot_ms / (ms_eager * psteps)}
    if args.dump_profile:
        os.makedirs(os.path.dirname(os.path.abspath(args.dump_profile)), exist_ok=True)
        json.dump({"per_layer": per_layer, "bandwidth": bw, "roofline": roofline, "ms_per_step": ms / args.steps},
                  open(args.dump_profile, "w"), indent=1)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        Bs = min(ref_batch, B)
        nref = 2
        dt = cpu_reference_step(args.model, Bs, nref, 1)
        cpu = {"value": Bs / dt, "unit": "images/s", "cores": _host_threads(), "kind": "port", "threads": blas_threads(),
               "sample": f"{nref} full fit() iterations of the same network at batch {Bs} after 1 warm-up "
                         "(reference CPU algorithm: im2col + OpenBLAS sgemm + col2im, OpenMP over all cores, passive wait policy)"}

    line = {
        "metric": f"{args.model}_train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "tf32" if args.algo == "tf32" else "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world,
                   "per_gpu_batch": B, "parallelism": f"dp{world}" if world > 1 else "single",
                   "l2_flush": "working set (>= 10 GB of activations per step) exceeds the 126 MB L2" if args.model != "lenet"
                               else "none: the whole LeNet step fits L2 (latency-bound config)",
                   "launch": ("input re-layout + CUDA graph replay of the whole step" if use_graph else "eager kernel chain"),
                   "dropout": "fixed host-generated masks (inputs of the step)" if masks else "none"},
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": int(x.nbytes + lab.nbytes), "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps, "last_score": last, "h2d_link_gbs": h2d_gbs},
        "gpu_launches": int(launches), "clocks": sampler.summary(), "roofline": roofline, "hbm_kernels": bw,
    }
    if dp_check:
        line["dp_check"] = dp_check
    if cpu:
        line["cpu_baseline"] = cpu
    json_out.write(json.dumps(line) + "\n")
    json_out.flush()
    finish()


if __name__ == "__main__":
    main()
