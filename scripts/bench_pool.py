"""Micro-benchmark of max-pool forward/backward (NHWC) through the C ABI. usage: bench_pool.py B C H W k s [iters]"""
import ctypes as C, sys
sys.path.insert(0, ".")
import torch
from deeplearning4j_b200 import _lib, helpers as H
B, Cc, Hh, W, k, s = (int(a) for a in sys.argv[1:7])
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 20
stream = torch.cuda.Stream()
h = C.c_void_p(); _lib.call("dl4jb200_ctx_create", 0, C.c_void_p(stream.cuda_stream), C.byref(h))
d = H.conv_desc((B, Cc, Hh, W), Cc, (k, k), (s, s), (0, 0), (1, 1), "Truncate")
x = torch.randn(B, Hh, W, Cc, device="cuda"); y = torch.empty(B, d.out_h, d.out_w, Cc, device="cuda")
am = torch.empty(B, d.out_h, d.out_w, Cc, device="cuda", dtype=torch.int32); e = torch.randn_like(y); dx = torch.empty_like(x)
p = lambda t: C.c_void_p(t.data_ptr())
torch.cuda.synchronize()
def fwd(): _lib.call("dl4jb200_pool_fwd", h, C.byref(d), p(x), p(y), p(am), 1, Cc, Cc, 0, 2)
def bwd(): _lib.call("dl4jb200_pool_bwd", h, C.byref(d), p(x), p(y), p(e), p(am), p(dx), 1, Cc, Cc, 0, 2, 1e-8)
for name, fn, nbytes in (("fwd", fwd, 4*(x.numel()+2*y.numel())), ("bwd", bwd, 4*(x.numel()+2*y.numel()))):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream.synchronize(); e0.record(stream)
    for _ in range(iters): fn()
    e1.record(stream); stream.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"pool {name}: {ms:.4f} ms  {nbytes/ms/1e6:.0f} GB/s (algorithmic)")
