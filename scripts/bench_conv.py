"""Micro-benchmark of one convolution layer through the C ABI (NHWC device tensors), CUDA-event timed.

usage: python scripts/bench_conv.py B Cin H W Cout k stride mode [iters] [what]   (what: fwd|bwd|both|dgrad)
what=dgrad times backprop-data alone, plain and with the fused epsilon-sum + ReLU-mask epilogue.
Prints ms and TFLOP/s per pass. Also the command profiled with `ncu --set full` for profiles/.
"""
import ctypes as C
import sys

sys.path.insert(0, ".")
import numpy as np
import torch

from deeplearning4j_b200 import _lib, helpers as H

B, Cin, Hh, W, Cout, k, s = (int(a) for a in sys.argv[1:8])
mode = sys.argv[8] if len(sys.argv) > 8 else "Truncate"
iters = int(sys.argv[9]) if len(sys.argv) > 9 else 20
what = sys.argv[10] if len(sys.argv) > 10 else "both"
stream = torch.cuda.Stream()
h = C.c_void_p()
_lib.call("dl4jb200_ctx_create", 0, C.c_void_p(stream.cuda_stream), C.byref(h))
d = H.conv_desc((B, Cin, Hh, W), Cout, (k, k), (s, s), (0, 0), (1, 1), mode)
xcs, zcs = (Cin + 3) // 4 * 4, (Cout + 3) // 4 * 4
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.rand(B, Hh, W, xcs, device="cuda", generator=g)
w = torch.randn(Cout, Cin, k, k, device="cuda", generator=g) * 0.1
b = torch.randn(Cout, device="cuda", generator=g)
z = torch.empty(B, d.out_h, d.out_w, zcs, device="cuda")
delta = torch.randn(B, d.out_h, d.out_w, zcs, device="cuda", generator=g)
dw, db, dx = torch.empty_like(w), torch.empty_like(b), torch.empty_like(x)
torch.cuda.synchronize()
p = lambda t: C.c_void_p(t.data_ptr())
flops = 2.0 * B * d.out_h * d.out_w * Cout * Cin * k * k


def fwd():
    _lib.call("dl4jb200_conv_fwd", h, C.byref(d), p(x), 1, xcs, p(w), p(b), p(z), 1, zcs, 0, 0)


def bwd():
    _lib.call("dl4jb200_conv_bwd", h, C.byref(d), p(x), 1, xcs, p(w), p(delta), 1, zcs, p(dw), p(db), p(dx), 1, xcs, 0)


def dgrad():
    _lib.call("dl4jb200_conv_bwd", h, C.byref(d), p(x), 1, xcs, p(w), p(delta), 1, zcs, None, None, p(dx), 1, xcs, 0)


prior, dx2 = torch.randn_like(x), torch.empty_like(x)


def dgrad_fused():
    _lib.call("dl4jb200_conv_bwd_data_fused", h, C.byref(d), p(w), p(delta), zcs, p(dx2), xcs, p(prior), p(x), 1, 0)


def unfused_chain():
    dgrad()
    _lib.call("dl4jb200_accumulate", h, p(dx), p(prior), dx.numel())
    _lib.call("dl4jb200_act_bwd", h, p(x), p(dx), p(dx), dx.numel(), 1)


cases = (("fwd", fwd, 1), ("bwd", bwd, 2))
if what == "dgrad":
    cases = (("dgrad", dgrad, 1), ("dgrad_fused", dgrad_fused, 1), ("dgrad+accumulate+act_bwd", unfused_chain, 1))
for name, fn, passes in cases:
    if what not in (name, "both", "dgrad"):
        continue
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream.synchronize()
    e0.record(stream)
    for _ in range(iters):
        fn()
    e1.record(stream)
    stream.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"{name}: {ms:.4f} ms  {passes * flops / ms / 1e9:.1f} TFLOP/s  (M={B * d.out_h * d.out_w} N={Cout} K={Cin * k * k})")
