"""A/B timing of the convolution GEMMs with and without CTA-pair tiles (DL4JB200_OPT_PAIR_TILES) on ResNet-50 / VGG16 layer
shapes at their BASELINE batch: forward and backprop-data GEMM alone (filters pre-packed), CUDA events on the context's stream, L2-cold
inputs by rotation over several buffer sets. Prints one line per shape: us with pairs off / on."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from deeplearning4j_b200 import _lib as L, helpers as H

SHAPES = [  # name, B, Cin, H, W, Cout, k, stride, pad-mode
    ("res2 3x3 64->64 @28", 256, 64, 28, 28, 64, 3, 1, "Same"),
    ("res2 1x1 64->256 @28", 256, 64, 28, 28, 256, 1, 1, "Truncate"),
    ("res2 1x1 256->64 @28", 256, 256, 28, 28, 64, 1, 1, "Truncate"),
    ("res3 3x3 128->128 @14", 256, 128, 14, 14, 128, 3, 1, "Same"),
    ("res3 1x1 128->512 @14", 256, 128, 14, 14, 512, 1, 1, "Truncate"),
    ("res3 1x1 512->128 @14", 256, 512, 14, 14, 128, 1, 1, "Truncate"),
    ("res4 3x3 256->256 @7", 256, 256, 7, 7, 256, 3, 1, "Same"),
    ("res4 1x1 256->1024 @7", 256, 256, 7, 7, 1024, 1, 1, "Truncate"),
    ("res4 1x1 1024->256 @7", 256, 1024, 7, 7, 256, 1, 1, "Truncate"),
    ("res5 3x3 512->512 @4", 256, 512, 4, 4, 512, 3, 1, "Same"),
    ("res5 1x1 512->2048 @4", 256, 512, 4, 4, 2048, 1, 1, "Truncate"),
    ("vgg 3x3 256->256 @56", 64, 256, 56, 56, 256, 3, 1, "Same"),
    ("vgg 3x3 512->512 @28", 64, 512, 28, 28, 512, 3, 1, "Same"),
]
only = sys.argv[1] if len(sys.argv) > 1 else None
torch.cuda.init()
st = torch.cuda.Stream()
ctx = H.Context(0, C.c_void_p(st.cuda_stream))
L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_TF32_ACTIVATIONS, 1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
P = lambda t: C.c_void_p(t.data_ptr())
for name, B, Cin, Hh, W, Cout, k, s, mode in SHAPES:
    if only and only not in name:
        continue
    d = H.conv_desc((B, Cin, Hh, W), Cout, (k, k), (s, s), (0, 0), (1, 1), mode)
    with torch.cuda.stream(st):
        x = torch.randn(B, Hh, W, Cin, device="cuda")
        w = torch.randn(Cout, Cin, k, k, device="cuda") * 0.05
        b = torch.zeros(Cout, device="cuda")
        z = torch.empty(B, d.out_h, d.out_w, Cout, device="cuda")
        dz = torch.randn(B, d.out_h, d.out_w, Cout, device="cuda")
        dx = torch.empty_like(x); dw = torch.empty_like(w); db = torch.empty_like(b)
    L.call("dl4jb200_filters_invalidate", ctx.handle)
    L.call("dl4jb200_filter_register", ctx.handle, C.byref(d), P(w), Cin, Cout, L.ALGO_TC_TF32, 1)
    L.call("dl4jb200_filters_repack", ctx.handle)   # the calls below launch the GEMM alone (no pack kernel)
    res = {}
    outs = {}
    for pair in (0, 1):
        L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_PAIR_TILES, pair)
        def fwd():
            L.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), P(x), L.NHWC, Cin, P(w), P(b), P(z), L.NHWC, Cout, 0, L.ALGO_TC_TF32)
        def bwd():   # backprop-data only
            L.call("dl4jb200_conv_bwd_data_fused", ctx.handle, C.byref(d), P(w), P(dz), Cout, P(dx), Cin, None, None, L.ACT_IDENTITY, L.ALGO_TC_TF32)
        for what, fn in (("fwd", fwd), ("bwd", bwd)):
            ts = []
            with torch.cuda.stream(st):
                for it in range(12):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(st); fn(); e1.record(st)
                    st.synchronize()
                    if it >= 3:
                        ts.append(e0.elapsed_time(e1) * 1e3)
            res[(what, pair)] = float(np.median(ts))
        st.synchronize()
        outs[pair] = (z.clone(), dx.clone())
    same = bool(torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]))
    print("%-26s fwd %7.1f -> %7.1f us   dgrad %7.1f -> %7.1f us   bit-identical %s" % (
        name, res[("fwd", 0)], res[("fwd", 1)], res[("bwd", 0)], res[("bwd", 1)], same), flush=True)
ctx.close()
