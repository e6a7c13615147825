"""Measures the TF32 dense roof on this box, the same way MEASURED_PEAKS.json measures bf16: torch.matmul (cuBLAS) with
allow_tf32 on 8192^3, best of 10 (burst) and back to back for 4 s (sustained) -- plus this library's own tcgen05 kernel on
a GEMM-shaped 1x1 convolution (M = 65536 pixels, K = N = 4096). Prints one JSON object (committed as profiles/r02_tf32_peak.json)."""
import ctypes as C
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def clocks():
    import subprocess
    try:
        out = subprocess.run(["nvidia-smi", "--id=0", "--query-gpu=clocks.sm,clocks.max.sm,power.draw", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        return [float(v) for v in out.split(",")]
    except Exception:  # noqa: BLE001
        return None


def matmul_peak(dtype, allow_tf32, n=8192):
    torch.backends.cuda.matmul.allow_tf32 = allow_tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype)
    b = torch.randn(n, n, device="cuda", dtype=dtype)
    c = torch.empty(n, n, device="cuda", dtype=dtype)
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    fl = 2.0 * n ** 3
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, fl / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    # sustained: back to back for ~4 s
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, t0 = 0, time.time()
    e0.record()
    mid = None
    while time.time() - t0 < 4.0:
        for _ in range(20):
            torch.matmul(a, b, out=c)
        reps += 20
        torch.cuda.synchronize()
        if mid is None and time.time() - t0 > 2.0:
            mid = clocks()
    e1.record()
    torch.cuda.synchronize()
    return best, reps * fl / (e0.elapsed_time(e1) * 1e-3) / 1e12, mid


def own_kernel():
    from deeplearning4j_b200 import _lib
    B, H, W, Cin, Cout = 64, 32, 32, 4096, 4096
    stream = torch.cuda.Stream()
    h = C.c_void_p()
    _lib.call("dl4jb200_ctx_create", 0, C.c_void_p(stream.cuda_stream), C.byref(h))
    _lib.call("dl4jb200_ctx_set_option", h, _lib.OPT_TF32_ACTIVATIONS, 1)
    x = torch.randn(B, H, W, Cin, device="cuda")
    w = torch.randn(Cout, Cin, device="cuda") * 0.02
    z = torch.empty(B, H, W, Cout, device="cuda")
    d = _lib.ConvDesc(B, Cin, H, W, Cout, 1, 1, 1, 1, 0, 0, 1, 1, H, W)
    torch.cuda.synchronize()
    _lib.call("dl4jb200_filter_register", h, C.byref(d), C.c_void_p(w.data_ptr()), Cin, Cout, _lib.ALGO_TC_TF32, 0)
    _lib.call("dl4jb200_filters_repack", h)
    args = (h, C.byref(d), C.c_void_p(x.data_ptr()), _lib.NHWC, Cin, C.c_void_p(w.data_ptr()), None, C.c_void_p(z.data_ptr()),
            _lib.NHWC, Cout, _lib.ACT_IDENTITY, _lib.ALGO_TC_TF32)
    for _ in range(3):
        _lib.call("dl4jb200_conv_fwd", *args)
    stream.synchronize()
    fl = 2.0 * B * H * W * Cin * Cout
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        _lib.call("dl4jb200_conv_fwd", *args)
        e1.record(stream)
        stream.synchronize()
        best = max(best, fl / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, t0 = 0, time.time()
    e0.record(stream)
    while time.time() - t0 < 4.0:
        for _ in range(20):
            _lib.call("dl4jb200_conv_fwd", *args)
        reps += 20
        stream.synchronize()
    e1.record(stream)
    stream.synchronize()
    sus = reps * fl / (e0.elapsed_time(e1) * 1e-3) / 1e12
    _lib.call("dl4jb200_ctx_destroy", h)
    return best, sus


if __name__ == "__main__":
    tf_b, tf_s, tf_clk = matmul_peak(torch.float32, True)
    bf_b, bf_s, bf_clk = matmul_peak(torch.bfloat16, False)
    fp_b, _, _ = matmul_peak(torch.float32, False, n=4096)
    ob, osus = own_kernel()
    print(json.dumps({
        "tf32_tflops": tf_b, "tf32_tflops_sustained": tf_s, "bf16_tflops": bf_b, "bf16_tflops_sustained": bf_s,
        "fp32_simt_tflops_4096": fp_b, "own_tcgen05_1x1conv_tflops": ob, "own_tcgen05_1x1conv_tflops_sustained": osus,
        "clocks_mid_tf32": tf_clk, "clocks_mid_bf16": bf_clk, "gpu": torch.cuda.get_device_name(0),
        "how": "torch.matmul (cuBLAS) fp32 inputs with allow_tf32=True, 8192^3 (2 N^3 flops): best of 10 single calls (burst) and "
               "back to back for 4 s (sustained); bf16 the same way for comparison with MEASURED_PEAKS.json; own kernel = "
               "dl4jb200_conv_fwd 1x1 NHWC, M=65536 K=4096 N=4096, pre-packed filter, CUDA events on the launching stream"}))
