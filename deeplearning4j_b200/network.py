"""Device-resident training-step driver for the BASELINE configs.

Replays, per minibatch, the call sequence that MultiLayerNetwork / ComputationGraph issue
(computeGradientAndScore -> feedForward -> calcBackpropGradients -> updater -> params.subi:
nn/multilayer/MultiLayerNetwork.java:2360-2420, nn/graph/ComputationGraph.java:1321-1360,1604-1739,1977-2097,
optimize/solvers/StochasticGradientDescent.java:58-91) as a chain of C-ABI calls on device-resident NHWC
buffers: activations, the flat parameter vector, the flat gradient vector (layout of zoo.param_layout ==
DL4J's flat views) and the updater state never leave HBM. torch is used only to own device memory and the
stream (and, in parallel.py, for NCCL); all arithmetic runs in libdl4jb200's hand-written kernels.
"""
import ctypes as C

import numpy as np

from . import _lib, zoo
from ._lib import ConvDesc

_ACT = {"identity": _lib.ACT_IDENTITY, "relu": _lib.ACT_RELU, "sigmoid": _lib.ACT_SIGMOID, "tanh": _lib.ACT_TANH}
_POOL = {"MAX": _lib.POOL_MAX, "AVG": _lib.POOL_AVG, "PNORM": _lib.POOL_PNORM}
_MODE = {"Strict": 0, "Truncate": 1, "Same": 2}


def _r4(n):
    return (n + 3) // 4 * 4


class Network:
    def __init__(self, model, batch, device=0, algo=_lib.ALGO_TC_TF32, params=None, with_score=True, round_tf32=None):
        import torch
        self.torch = torch
        self.model, self.B, self.algo, self.with_score = model, int(batch), algo, with_score
        self.dev = torch.device("cuda", device)
        torch.cuda.set_device(self.dev)
        # a dedicated stream shared by torch (memory ops, NCCL) and the C-ABI context: one ordered queue
        self.stream = torch.cuda.Stream(self.dev)
        h = C.c_void_p()
        _lib.call("dl4jb200_ctx_create", device, C.c_void_p(self.stream.cuda_stream), C.byref(h))
        self.h = h
        # Filter / bias gradients do not feed the backward chain: they run on a second stream + context (own scratch),
        # forked from the main stream once a layer's delta is final and joined before the gradient exchange / updater.
        self.side_stream = torch.cuda.Stream(self.dev)
        h2 = C.c_void_p()
        _lib.call("dl4jb200_ctx_create", device, C.c_void_p(self.side_stream.cuda_stream), C.byref(h2))
        self.h2 = h2
        # TF32 mode: producers store activations / epsilons TF32-representable, so the tensor cores' operand truncation
        # is exact and the whole step sees one unbiased rounding per stored value (include/dl4jb200.h, ctx_set_option)
        self.round_tf32 = (algo == _lib.ALGO_TC_TF32) if round_tf32 is None else bool(round_tf32)
        for ctx in (h, h2):
            _lib.call("dl4jb200_ctx_set_option", ctx, _lib.OPT_TF32_ACTIVATIONS, int(self.round_tf32))
        self.drop, self.mask = {}, {}
        self.overlap_wgrad = True
        self.fuse_dgrad = True  # epsilon fan-in + ReLU backprop in the backprop-data epilogue (_emit_dgrad)
        self.launches = 0
        self.prof, self.prof_fns, self._cur = None, set(), None
        # debugging / parity aid: when set to a dict, backward() leaves in it a device copy of every vertex' final delta
        # (the epsilon summed over its consumers, after the vertex' own activation backprop; raw epsilon for BN / pooling)
        self.trace = None
        # data-parallel exchange overlapped with backward (parallel.ParallelWrapper sets itself here): it is told when
        # backward starts, after every vertex (its gradient blocks are then enqueued on both streams) and when it ends
        self.dp = None
        self.batch_div_scale = 1  # data-parallel replicas: global minibatch = scale * B (parallel.py)
        self.shapes = zoo.infer_shapes(model)
        self.P, self.layout = zoo.param_layout(model)
        self.vmap = {v["name"]: v for v in model["vertices"]}
        host = zoo.init_params(model) if params is None else np.ascontiguousarray(params, np.float32)
        assert host.size == self.P
        self.params = torch.from_numpy(host.copy()).to(self.dev)
        self.grads = torch.zeros(self.P, device=self.dev)
        kind = model["updater"]["kind"]
        self.state = torch.zeros(self.P * (2 if kind == "adadelta" else 1), device=self.dev)
        self.pidx = {}
        for name, key, off, shp, pk in self.layout:
            self.pidx[(name, key)] = (off, int(np.prod(shp)), shp, pk)
        self._plan()
        torch.cuda.synchronize(self.dev)  # allocations / fills above ran on torch's default stream
        self.prepack = True
        self._register_filters()
        self.params_changed()
        self.fuse_bn_stats = True
        # backward sums of a BatchNormalization taken in the epilogue of the backprop-data GEMM that produces its epsilon
        # (dl4jb200_conv_bwd_data_bnstats): the layer's backward then runs finalize + apply only
        self.fuse_bn_bwd_stats = algo == _lib.ALGO_TC_TF32
        self._bst_plan, self._bst, self._pool_bst = {}, {}, {}
        self._plan_bn_stats()
        self._plan_forward_branches()

    # ------------------------------------------------------------------ pre-packed filters
    def _register_filters(self):
        """Every convolution / dense filter view of the flat parameter vector is registered once with the context; one
        dl4jb200_filters_repack launch after each parameter update then replaces the per-call pack kernels."""
        for v in self.model["vertices"]:
            n, op = v["name"], v["op"]
            if op not in ("conv", "dense", "output") or (self.stem and self.stem[0] == n):
                continue
            src, d = self._conv_desc(v) if op == "conv" else (v["in"][0], self._dense_desc(v))
            _, xcs = self._src(src)
            pw, _, _, _ = self._p(n, "W")
            self._call("dl4jb200_filter_register", C.byref(d), pw, xcs, self.cs[n], self.algo, int(self._needs_eps(src)))

    def _plan_bn_stats(self):
        """Convolutions whose only consumer is a BatchNormalization take that layer's batch statistics in their GEMM
        epilogue (dl4jb200_conv_fwd_bnstats) when the launch can carry them; the BN then runs finalize + apply only."""
        self.stat_conv, self.stat_bn = {}, {}
        if self.algo != _lib.ALGO_TC_TF32:
            return
        sm_count = self.torch.cuda.get_device_properties(self.dev).multi_processor_count
        for v in self.model["vertices"]:
            n = v["name"]
            if v["op"] != "conv" or v["act"] != "identity" or v.get("dropout"):
                continue
            cons = self.consumers[n]
            if len(cons) != 1 or self.vmap[cons[0]]["op"] != "bn":
                continue
            if self.stem and self.stem[0] == n:
                # row-packed stem: the call itself reports the rows it wrote (or NOT_SUPPORTED, probed in the first forward)
                if self.cs[n] == self.shapes[n][0] and self.cs[n] % 32 == 0:
                    self.stat_conv[n] = (cons[0], self._buf(4 * sm_count * 2 * self.cs[n]))
                continue
            src, d = self._conv_desc(v)
            _, xcs = self._src(src)
            rows = C.c_int(0)
            self._call("dl4jb200_conv_fwd_bnstats_rows", C.byref(d), xcs, self.cs[n], self.algo, C.byref(rows))
            if rows.value > 0:
                buf = self._buf(rows.value * 2 * self.cs[n])
                self.stat_conv[n] = (cons[0], buf)
                self.stat_bn[cons[0]] = (buf, rows.value)

    def _plan_forward_branches(self):
        """Projection shortcuts (conv -> BN feeding an Add whose other input's BN is folded into it) do not depend on the
        residual branch: in the forward pass they are issued on the side stream, forked when the block input is ready and
        joined before the Add, so they fill the SMs the small-tile convolutions of the main branch leave idle."""
        self.fwd_side, self.fwd_fork, self.fwd_join, self.fwd_mark = set(), {}, {}, set()
        for add, (bname, other) in self.fused_add.items():
            ov = self.vmap[other]
            if ov["op"] != "bn" or self.consumers[other] != [add] or other in self.fused_bn:
                continue
            c = ov["in"][0]
            cv = self.vmap[c]
            if cv["op"] != "conv" or self.consumers[c] != [other] or (self.stem and self.stem[0] == c) or c in self.fold_pad:
                continue
            u = cv["in"][0]
            if self.vmap[u]["op"] in ("input", "zeropad") or self.out.get(u, 0) is None or len(self.consumers[u]) < 2:
                continue
            self.fwd_side.update((c, other))
            self.fwd_fork[c] = u        # side waits until u has been produced
            self.fwd_mark.add(u)        # record an event on the main stream right after producing u
            self.fwd_join[add] = other  # main waits for the side chain before the Add

    def params_changed(self):
        """Call after anything other than update() wrote the flat parameter vector (broadcast, parameter averaging)."""
        if self.prepack:
            self._call("dl4jb200_filters_repack")

    # ------------------------------------------------------------------ planning
    def _buf(self, *shape, dtype=None):
        t = self.torch
        return t.zeros(*shape, device=self.dev, dtype=dtype or t.float32)

    def _plan(self):
        B = self.B
        self.consumers = {v["name"]: [] for v in self.model["vertices"]}
        for v in self.model["vertices"]:
            for u in v["in"]:
                self.consumers[u].append(v["name"])
        self.out, self.cs, self.eps, self.aux = {}, {}, {}, {}
        self.fold_pad = {}
        # BatchNormalization (identity activation) consumed only by an ElementWiseVertex(Add): the add (+ its ReLU) is
        # folded into the BN apply pass, y = act(bn(x) + other). fused_bn[bn] = add vertex; fused_add[add] = (bn, other)
        self.fused_bn, self.fused_add = {}, {}
        for v in self.model["vertices"]:
            if v["op"] == "add" and len(v["in"]) == 2:
                for i, u in enumerate(v["in"]):
                    uv = self.vmap[u]
                    if uv["op"] == "bn" and uv["act"] == "identity" and self.consumers[u] == [v["name"]] and v["in"][1 - i] != u:
                        self.fused_bn[u] = v["name"]
                        self.fused_add[v["name"]] = (u, v["in"][1 - i])
                        break
        for v in self.model["vertices"]:
            n, op, s = v["name"], v["op"], self.shapes[v["name"]]
            if len(s) == 3:
                cs = _r4(s[0])
                self.cs[n] = cs
                shape = (B, s[1], s[2], cs)
            else:
                cs = _r4(s[0])
                self.cs[n] = cs
                shape = (B, cs)
            if op == "zeropad":
                # ZeroPaddingLayer feeding exactly one Truncate-mode conv is the same arithmetic as that conv with
                # explicit zero padding (nn/layers/convolution/ZeroPaddingLayer.java:50-75): fold, no copy.
                cons = self.consumers[n]
                t, b, l, r = v["pad"]
                ok = len(cons) == 1 and self.vmap[cons[0]]["op"] == "conv" and self.vmap[cons[0]]["mode"] != "Same" and t == b and l == r
                if not ok:
                    raise NotImplementedError("stand-alone ZeroPaddingLayer")
                self.fold_pad[cons[0]] = (v["in"][0], t, l)
                continue
            if n in self.fused_bn:
                self.aux[n] = (self._buf(s[0]), self._buf(s[0]))  # saved mean / std; the output itself is never stored
                continue
            if op == "flatten":
                src = self.shapes[v["in"][0]]
                if src[1] * src[2] == 1 and self.cs[v["in"][0]] == src[0]:
                    self.out[n] = None  # alias of the input buffer ([B,1,1,C] NHWC == [B,C])
                    continue
                self.aux[n] = self._buf(B, src[0] * src[1] * src[2])
            self.out[n] = self._buf(*shape)
            if op == "pool" and v["type"] == "MAX":
                self.aux[n] = self._buf(*shape, dtype=self.torch.int32)
            if op == "bn":
                self.aux[n] = (self._buf(s[0]), self._buf(s[0]))  # saved mean / std
            if op == "output":
                self.aux[n] = self._buf(B, cs)  # delta = p - labels
        self._plan_stem()
        # BatchNormalization (+ folded ReLU) whose only consumer is a MAX pooling layer: when the batch statistics come from
        # the producing convolution's epilogue, the apply pass pools on the fly and the normalised tensor is never written
        # (dl4jb200_bn_fwd_prestats_maxpool; ResNet-50's stem). pool_fuse[bn] = pool vertex; decided per forward pass.
        self.pool_fuse, self.fused_pool_active, self.fuse_bn_pool = {}, set(), True
        for v in self.model["vertices"]:
            if v["op"] != "bn" or v["name"] in self.fused_bn or v["act"] not in ("identity", "relu"):
                continue
            cons = self.consumers[v["name"]]
            if len(cons) != 1:
                continue
            pv = self.vmap[cons[0]]
            s3 = self.shapes[v["name"]]
            if (pv["op"] == "pool" and pv["type"] == "MAX" and not pv.get("dropout") and len(s3) == 3 and s3[0] % 4 == 0
                    and self.cs[v["name"]] == s3[0] and self.cs[pv["name"]] == s3[0] and tuple(pv.get("dilation", (1, 1))) == (1, 1)):
                self.pool_fuse[v["name"]] = pv["name"]
        self.loss = self._buf(1)
        self.score_reg = self._buf(1)
        self.score = self._buf(1)
        self.tmp_eps = {}
        self.labels = None

    def _plan_stem(self):
        """First-layer convolution on <= 4 channels (ResNet-50's 7x7/2, VGG16's 3x3/1, LeNet's 5x5/1, AlexNet's 11x11/4): keep
        the minibatch row-packed so the TMA-fed kernels can run it (dl4jb200_conv_rowpack_*)."""
        self.stem = None
        if self.algo != _lib.ALGO_TC_TF32:
            return
        inp = self.model["vertices"][0]["name"]
        users = [c for c in self.consumers[inp]]
        if len(users) == 1 and self.vmap[users[0]]["op"] == "zeropad":
            users = self.consumers[users[0]]
        if len(users) != 1 or self.vmap[users[0]]["op"] != "conv":
            return
        v = self.vmap[users[0]]
        src, d = self._conv_desc(v)
        if src != inp or d.c_in > 4 or d.stride_w > 8 or d.k_w > 15 or (d.k_w > 7 and 8 % d.stride_w) or d.k_h > 16 \
                or d.dil_h != 1 or d.dil_w != 1 or v["act"] not in ("identity", "relu"):
            return
        chunks = (d.k_w + 1 + 7) // 8  # 8-pixel K-chunks per filter row (AlexNet's 11x11/4: two)
        pitch = max(d.stride_w * (d.out_w - 1) + 8 * chunks, d.pad_left + 1 + d.in_w)
        pitch += pitch % 2
        self.stem = (v["name"], pitch, d.pad_left + 1)
        self.out[inp] = self._buf(self.B, d.in_h, pitch, 4)  # padding columns stay zero

    def _ptr(self, t, off=0):
        return C.c_void_p(t.data_ptr() + 4 * off)

    def _p(self, name, key):
        off, n, shp, pk = self.pidx[(name, key)]
        return self._ptr(self.params, off), self._ptr(self.grads, off), off, n

    def _call(self, fn, *args):
        self.launches += 1
        if self.prof is not None and fn in self.prof_fns:
            ev = self.torch.cuda.Event
            a, b = ev(enable_timing=True), ev(enable_timing=True)
            a.record(self.stream)
            _lib.call(fn, self.h, *args)
            b.record(self.stream)
            self.prof.append((fn, self._cur, a, b))
        else:
            _lib.call(fn, self.h, *args)

    def _call_side(self, fn, *args):
        """Same as _call on the side (filter-gradient) context / stream."""
        self.launches += 1
        if self.prof is not None and fn in self.prof_fns:
            ev = self.torch.cuda.Event
            a, b = ev(enable_timing=True), ev(enable_timing=True)
            a.record(self.side_stream)
            _lib.call(fn, self.h2, *args)
            b.record(self.side_stream)
            self.prof.append((fn, self._cur, a, b))
        else:
            _lib.call(fn, self.h2, *args)

    def profile(self, fns):
        """Start recording CUDA-event pairs (on the network's stream) around the given C-ABI calls."""
        self.prof, self.prof_fns = [], set(fns)

    def profile_results(self):
        """-> {(fn, vertex): [ms, ...]} ; call after sync()."""
        out = {}
        for fn, tag, a, b in self.prof or []:
            out.setdefault((fn, tag), []).append(a.elapsed_time(b))
        return out

    def _src(self, u):
        """(tensor, cs) of vertex u's output, resolving flatten aliases."""
        while self.out.get(u, 0) is None:
            u = self.vmap[u]["in"][0]
        return self.out[u], self.cs[u]

    def _lin(self, n, src):
        """(tensor, cs) of the INPUT of layer n: the output of src, or its dropped copy when n has dropout and masks are
        set (AbstractLayer.applyDropOutIfNecessary replaces the layer's input; backprop then uses that dropped input)."""
        x, xcs = self._src(src)
        d = self.drop.get(n)
        return (d, xcs) if d is not None else (x, xcs)

    def set_dropout_masks(self, masks):
        """masks: {layer name: 0/1 float array shaped like that layer's input in DL4J layout ([B,C,H,W] or [B,K])}, for
        every vertex with a "dropout" retain probability; None switches dropout off again. The masks are inputs because
        nd4j's DropOutInverted random stream cannot be reproduced outside nd4j."""
        t = self.torch
        self.drop, self.mask = {}, {}
        if not masks:
            return
        for v in self.model["vertices"]:
            n, op = v["name"], v["op"]
            if not v.get("dropout"):
                continue
            if op not in ("conv", "dense", "output", "pool") or n in self.fold_pad:
                raise NotImplementedError(f"dropout on {op} vertex {n}")
            x, xcs = self._src(v["in"][0])
            m = np.ascontiguousarray(masks[n], np.float32)
            with t.cuda.stream(self.stream):
                dev = t.zeros_like(x)
                if self.stem and self.stem[0] == n:  # the first layer's input is held row-packed: so is its mask
                    Bb, Cc, Hh, Ww = m.shape
                    tmp = t.from_numpy(m).to(self.dev)
                    self._call("dl4jb200_nchw_to_rowpack4", self._ptr(tmp), self._ptr(dev), Bb, Cc, Hh, Ww, self.stem[1], self.stem[2])
                    self.stream.synchronize()
                elif m.ndim == 4:
                    Bb, Cc, Hh, Ww = m.shape
                    tmp = t.from_numpy(m).to(self.dev)
                    self._call("dl4jb200_nchw_to_nhwc", self._ptr(tmp), self._ptr(dev), Bb, Cc, Hh, Ww, xcs)
                    self.stream.synchronize()  # tmp is released when this scope ends
                else:
                    dev.view(self.B, -1)[:, :m.shape[1]].copy_(t.from_numpy(m).to(self.dev))
                self.mask[n] = dev
                self.drop[n] = t.zeros_like(x)

    def _apply_dropout(self, v, train):
        n = v["name"]
        if train and n in self.drop:
            x, _ = self._src(v["in"][0])
            self._call("dl4jb200_dropout_fwd", self._ptr(x), self._ptr(self.mask[n]), self._ptr(self.drop[n]), x.numel(),
                       float(v["dropout"]))

    def _conv_desc(self, v):
        n = v["name"]
        if n in self.fold_pad:
            src, pt, pl = self.fold_pad[n]
        else:
            src, pt, pl = v["in"][0], None, None
        s_in, s_out = self.shapes[src], self.shapes[n]
        k, st, d = v["kernel"], v["stride"], v["dilation"]
        if pt is None:
            m = _MODE[v["mode"]]
            pt = _lib.out_size(s_in[1], k[0], st[0], v["pad"][0], d[0], m)[1]
            pl = _lib.out_size(s_in[2], k[1], st[1], v["pad"][1], d[1], m)[1]
        return src, ConvDesc(self.B, s_in[0], s_in[1], s_in[2], v["nOut"], k[0], k[1], st[0], st[1], pt, pl, d[0], d[1],
                             s_out[1], s_out[2])

    def _dense_desc(self, v):
        nin = self.shapes[v["in"][0]][0]
        return ConvDesc(self.B, nin, 1, 1, v["nOut"], 1, 1, 1, 1, 0, 0, 1, 1, 1, 1)

    # ------------------------------------------------------------------ data
    def set_batch(self, features, labels):
        """features: float32 [B,C,H,W] (host numpy / pinned or device torch tensor, NCHW as DL4J feeds it);
        labels: [B,K]. Copies to the device (async on the context's stream) and re-lays out to NHWC."""
        t = self.torch
        inp = self.model["vertices"][0]["name"]
        Cc, Hh, W = self.shapes[inp]
        with t.cuda.stream(self.stream):
            f = t.as_tensor(features) if not isinstance(features, t.Tensor) else features
            if f.device != self.dev:
                f = f.to(self.dev, non_blocking=True)
            lab = t.as_tensor(labels) if not isinstance(labels, t.Tensor) else labels
            self.labels = lab.to(self.dev, non_blocking=True).contiguous()
            self._feat = f.contiguous()
        self._relayout_input()

    def _relayout_input(self, src=None):
        inp = self.model["vertices"][0]["name"]
        Cc, Hh, W = self.shapes[inp]
        src = self._feat if src is None else src
        if self.stem:
            self._call("dl4jb200_nchw_to_rowpack4", self._ptr(src), self._ptr(self.out[inp]), self.B, Cc, Hh, W,
                       self.stem[1], self.stem[2])
        else:
            self._call("dl4jb200_nchw_to_nhwc", self._ptr(src), self._ptr(self.out[inp]), self.B, Cc, Hh, W, self.cs[inp])

    # ------------------------------------------------------------------ forward (feedForward, train mode)
    def forward(self, train=True):
        try:
            self._forward(train)
        finally:
            self.__dict__.pop("_call", None)  # never leave the side-stream redirection of a shortcut branch in place

    def _forward(self, train):
        L, B = _lib, self.B
        self._pool_done, self.fused_pool_active = set(), set()
        branches = self.overlap_wgrad and bool(self.fwd_side)
        ready, joined = {}, {}
        for v in self.model["vertices"]:
            n, op = v["name"], v["op"]
            self._cur = n
            if op in ("input", "zeropad") or self.out.get(n, 0) is None or n in self.fused_bn:
                continue
            if branches:
                self.__dict__.pop("_call", None)
                if n in self.fwd_side:
                    if n in self.fwd_fork:
                        self.side_stream.wait_event(ready[self.fwd_fork[n]])
                    self._call = self._call_side  # this vertex' C-ABI calls go to the side context / stream
                elif n in self.fwd_join:
                    self.stream.wait_event(joined[self.fwd_join[n]])
            self._forward_vertex(v, train)
            if branches:
                if n in self.fwd_mark:
                    ready[n] = self.torch.cuda.Event()
                    ready[n].record(self.stream)
                if n in self.fwd_side and n in self.fwd_join.values():
                    joined[n] = self.torch.cuda.Event()
                    joined[n].record(self.side_stream)

    def _forward_vertex(self, v, train):
        L, B = _lib, self.B
        n, op = v["name"], v["op"]
        if op == "conv":
            src, d = self._conv_desc(v)
            self._apply_dropout(v, train)
            x, xcs = self._lin(n, src)
            pw, _, _, _ = self._p(n, "W")
            pb, _, _, _ = self._p(n, "b")
            if self.stem and self.stem[0] == n and train and self.fuse_bn_stats and n in self.stat_conv:
                bn, buf = self.stat_conv[n]
                pm, _, _, _ = self._p(bn, "mean")
                rows = C.c_int(0)
                self._call("dl4jb200_conv_rowpack_fwd_bnstats", C.byref(d), self._ptr(x), self.stem[1], pw, pb,
                           self._ptr(self.out[n]), self.cs[n], pm, self._ptr(buf), C.byref(rows))
                self.stat_bn[bn] = (buf, rows.value)
            elif self.stem and self.stem[0] == n:
                self.stat_bn.pop(self.stat_conv.get(n, (None,))[0], None)
                self._call("dl4jb200_conv_rowpack_fwd", C.byref(d), self._ptr(x), self.stem[1], pw, pb, self._ptr(self.out[n]),
                           self.cs[n], _ACT[v["act"]])
            elif train and self.fuse_bn_stats and n in self.stat_conv:
                bn, buf = self.stat_conv[n]
                pm, _, _, _ = self._p(bn, "mean")  # pivot of the fp32 sums: the running mean (read before it is updated)
                self._call("dl4jb200_conv_fwd_bnstats", C.byref(d), self._ptr(x), xcs, pw, pb, self._ptr(self.out[n]), self.cs[n],
                           self.algo, pm, self._ptr(buf))
            else:
                self._call("dl4jb200_conv_fwd", C.byref(d), self._ptr(x), L.NHWC, xcs, pw, pb, self._ptr(self.out[n]), L.NHWC,
                           self.cs[n], _ACT[v["act"]], self.algo)
        elif op == "bn" or n in self.fused_add:
            if op == "bn":
                bnv, res, act = v, None, _ACT[v["act"]]
            else:  # Add(+ReLU) folded into the BatchNormalization that feeds it
                bname, other = self.fused_add[n]
                bnv, res, act = self.vmap[bname], self._ptr(self._src(other)[0]), _ACT[v["act"]]
            bn = bnv["name"]
            x, xcs = self._src(bnv["in"][0])
            s = self.shapes[bn]
            pg, _, _, _ = self._p(bn, "gamma")
            pbeta, _, _, _ = self._p(bn, "beta")
            pm, _, _, _ = self._p(bn, "mean")
            pv, _, _, _ = self._p(bn, "var")
            sm, ss = self.aux[bn]
            if (train and self.fuse_bn_stats and bn in self.stat_bn and op == "bn" and self.fuse_bn_pool
                    and bn in self.pool_fuse and self.algo == L.ALGO_TC_TF32):
                buf, rows = self.stat_bn[bn]
                pn = self.pool_fuse[bn]
                pd = self._pool_desc(self.vmap[pn])
                self._call("dl4jb200_bn_fwd_prestats_maxpool", self._ptr(x), self._ptr(self.out[pn]), self._ptr(self.aux[pn]), pg, pbeta,
                           pm, pv, self._ptr(sm), self._ptr(ss), B, s[0], s[1], s[2], xcs, float(bnv["decay"]), float(bnv["eps"]),
                           int(self.model["minibatch"]), act, pm, self._ptr(buf), rows, C.byref(pd))
                self._pool_done.add(pn)
                self.fused_pool_active.add(bn)  # out[bn] is NOT written in this pass
            elif train and self.fuse_bn_stats and bn in self.stat_bn:
                buf, rows = self.stat_bn[bn]
                self._call("dl4jb200_bn_fwd_prestats", self._ptr(x), self._ptr(self.out[n]), pg, pbeta, pm, pv, self._ptr(sm),
                           self._ptr(ss), B, s[0], s[1], s[2], xcs, float(bnv["decay"]), float(bnv["eps"]),
                           int(self.model["minibatch"]), act, res, pm, self._ptr(buf), rows)
            else:
                self._call("dl4jb200_bn_fwd", self._ptr(x), self._ptr(self.out[n]), pg, pbeta, pm, pv, self._ptr(sm), self._ptr(ss),
                           B, s[0], s[1], s[2], L.NHWC, xcs, float(bnv["decay"]), float(bnv["eps"]), int(train),
                           int(self.model["minibatch"]), act, res)
        elif op == "pool":
            if n in self._pool_done:  # already produced by the fused batch-norm apply pass of its input
                return
            d = self._pool_desc(v)
            self._apply_dropout(v, train)
            x, xcs = self._lin(n, v["in"][0])
            am = self._ptr(self.aux[n]) if v["type"] == "MAX" else None
            self._call("dl4jb200_pool_fwd", C.byref(d), self._ptr(x), self._ptr(self.out[n]), am, L.NHWC, xcs, self.cs[n],
                       _POOL[v["type"]], int(v.get("pnorm", 2)))
        elif op == "add":
            a, _ = self._src(v["in"][0])
            b, _ = self._src(v["in"][1])
            self._call("dl4jb200_add_act_fwd", self._ptr(a), self._ptr(b), self._ptr(self.out[n]), self.out[n].numel(),
                       _ACT[v["act"]])
        elif op == "act":
            a, _ = self._src(v["in"][0])
            self._call("dl4jb200_act_fwd", self._ptr(a), self._ptr(self.out[n]), a.numel(), _ACT[v["act"]])
        elif op == "flatten":
            x, xcs = self._src(v["in"][0])
            s = self.shapes[v["in"][0]]
            chw = s[0] * s[1] * s[2]
            dst = self.out[n] if self.cs[n] == chw else self.aux[n]
            self._call("dl4jb200_nhwc_to_nchw", self._ptr(x), self._ptr(dst), B, s[0], s[1], s[2], xcs)
            if dst is not self.out[n]:
                self._call("dl4jb200_copy2d", self._ptr(dst), chw, self._ptr(self.out[n]), self.cs[n], B, chw)
        elif op in ("dense", "output"):
            self._apply_dropout(v, train)
            x, xcs = self._lin(n, v["in"][0])
            d = self._dense_desc(v)
            pw, _, _, _ = self._p(n, "W")
            pb, _, _, _ = self._p(n, "b")
            act = _ACT[v["act"]] if op == "dense" else L.ACT_IDENTITY
            self._call("dl4jb200_conv_fwd", C.byref(d), self._ptr(x), L.NHWC, xcs, pw, pb, self._ptr(self.out[n]), L.NHWC,
                       self.cs[n], act, self.algo)
            if op == "output" and self.labels is not None:
                self._call("dl4jb200_softmax_nll", self._ptr(self.out[n]), self.cs[n], self._ptr(self.labels),
                           self._ptr(self.aux[n]), self.cs[n], self._ptr(self.loss), B, v["nOut"])
        else:
            raise NotImplementedError(op)

    def _pool_desc(self, v):
        s_in, s_out = self.shapes[v["in"][0]], self.shapes[v["name"]]
        k, st = v["kernel"], v["stride"]
        m = _MODE[v["mode"]]
        pt = _lib.out_size(s_in[1], k[0], st[0], v["pad"][0], 1, m)[1]
        pl = _lib.out_size(s_in[2], k[1], st[1], v["pad"][1], 1, m)[1]
        return ConvDesc(self.B, s_in[0], s_in[1], s_in[2], s_in[0], k[0], k[1], st[0], st[1], pt, pl, 1, 1, s_out[1], s_out[2])

    # ------------------------------------------------------------------ backward (calcBackpropGradients)
    def _eps_buf(self, u):
        if u not in self.eps:
            with self.torch.cuda.stream(self.stream):  # the fill is ordered before the kernels that write the buffer
                self.eps[u] = self.torch.zeros_like(self.out[u])
        return self.eps[u]

    def _needs_eps(self, u):
        return self.vmap[u]["op"] not in ("input", "zeropad")

    def _emit(self, u, fn):
        """Run fn(ptr) producing dL/d(out[u]); the first producer writes eps[u] directly, later ones are summed
        (ComputationGraph.java:2044-2063 epsilon fan-in)."""
        if not self._needs_eps(u):
            return
        if self.out.get(u, 0) is None:  # flatten alias: pass straight through to its input
            return self._emit(self.vmap[u]["in"][0], fn)
        e = self._eps_buf(u)
        self._contrib[u] = self._contrib.get(u, 0) + 1
        if u not in self._written:
            fn(self._ptr(e))
            self._written.add(u)
        else:
            key = tuple(e.shape)
            if key not in self.tmp_eps:
                with self.torch.cuda.stream(self.stream):
                    self.tmp_eps[key] = self.torch.zeros_like(e)
            t = self.tmp_eps[key]
            fn(self._ptr(t))
            self._call("dl4jb200_accumulate", self._ptr(e), self._ptr(t), e.numel())

    def _lend(self, u, tensor):
        if not self._needs_eps(u):
            return
        if self.out.get(u, 0) is None:
            return self._lend(self.vmap[u]["in"][0], tensor)
        self._borrowed.setdefault(u, []).append(tensor)
        self._contrib[u] = self._contrib.get(u, 0) + 1

    def _bst_buffer(self, n, d, bn):
        """Partial-sum buffer for the backward batch-norm sums carried by convolution n's backprop-data (None: it cannot)."""
        if n not in self._bst_plan:
            rows = C.c_int(0)
            cs_in = self.cs[self.vmap[bn]["in"][0]] if self.out.get(self.vmap[bn]["in"][0], 0) is not None else 0
            Cb = self.shapes[bn][0]
            ok = (self.fuse_bn_bwd_stats and len(self.shapes[bn]) == 3 and cs_in == Cb and Cb % 4 == 0
                  and all(self.pidx[(bn, k)][0] % 4 == 0 for k in ("gamma", "beta")))
            if ok:
                self._call("dl4jb200_conv_bwd_data_bnstats_rows", C.byref(d), self.cs[n], Cb, self.algo, C.byref(rows))
            with self.torch.cuda.stream(self.stream):  # the fill is ordered before the kernels that write the buffer
                self._bst_plan[n] = (rows.value, self._buf(rows.value * 2 * Cb)) if rows.value > 0 else None
        return self._bst_plan[n]

    def _pool_bst_plan(self, v, d):
        """Partial-sum buffer when MAX pooling layer v can carry the backward sums of the BatchNormalization feeding it
        (dl4jb200_pool_bwd_bnstats), else None."""
        n, u = v["name"], v["in"][0]
        if n not in self._pool_bst:
            uv = self.vmap[u]
            Cb = self.shapes[u][0]
            ok = (self.fuse_bn_bwd_stats and uv["op"] == "bn" and u not in self.fused_bn and self.consumers[u] == [n]
                  and uv["act"] in ("identity", "relu") and v["type"] == "MAX" and not v.get("dropout")
                  and len(self.shapes[u]) == 3 and self.cs[u] == Cb and self.cs[n] == Cb and Cb % 4 == 0 and 256 % (Cb // 4) == 0
                  and self.out.get(uv["in"][0], 0) is not None and self.cs[uv["in"][0]] == Cb
                  and d.stride_h == 2 and d.stride_w == 2 and d.k_h <= 3 and d.k_w <= 3 and d.pad_top >= 0 and d.pad_left >= 0
                  and all(self.pidx[(u, k)][0] % 4 == 0 for k in ("gamma", "beta")))
            buf = None
            if ok:
                sms = self.torch.cuda.get_device_properties(self.dev).multi_processor_count
                with self.torch.cuda.stream(self.stream):
                    buf = self._buf(4 * sms * 2 * Cb)
            self._pool_bst[n] = buf
        return self._pool_bst[n]

    def _bst_args(self, bn):
        v = self.vmap[bn]
        bx, _ = self._src(v["in"][0])
        pg, _, _, _ = self._p(bn, "gamma")
        pbeta, _, _, _ = self._p(bn, "beta")
        sm, ss = self.aux[bn]
        return self._ptr(bx), pg, pbeta, self._ptr(sm), self._ptr(ss), int(v["act"] == "relu")

    def _emit_dgrad(self, u, d, pw, e, n, x, xcs):
        """Backprop-data of convolution n into the epsilon of its input u. When this is the LAST contribution u will
        receive and u is an Add / Activation vertex with ReLU, the fan-in sum and u's ReLU backprop ride in the GEMM
        epilogue (dl4jb200_conv_bwd_data_fused) instead of running as an accumulate + act_bwd pass each."""
        L = _lib
        if not self._needs_eps(u):
            return
        while self.out.get(u, 0) is None:  # flatten alias
            u = self.vmap[u]["in"][0]
            if not self._needs_eps(u):
                return
        uv = self.vmap[u]
        lent = self._borrowed.get(u, [])
        last = self._contrib.get(u, 0) + 1 == len(self.consumers[u])
        fusable = (self.fuse_dgrad and last and uv["op"] in ("add", "act") and uv.get("act") == "relu"
                   and not (u in self._written and lent) and len(lent) <= 1)
        if not fusable:
            # u is a BatchNormalization (with or without its folded ReLU) that only this convolution consumes: dx IS that
            # layer's complete epsilon, so its backward sums can ride in this GEMM's epilogue
            if (uv["op"] == "bn" and u not in self.fused_bn and self.consumers[u] == [n] and u not in self._written and not lent
                    and uv["act"] in ("identity", "relu")):
                plan = self._bst_buffer(n, d, u)
                if plan is not None:
                    dst = self._eps_buf(u)
                    self._call("dl4jb200_conv_bwd_data_bnstats", C.byref(d), pw, self._ptr(e), self.cs[n], self._ptr(dst), xcs, None,
                               None, L.ACT_IDENTITY, self.algo, *self._bst_args(u), self._ptr(plan[1]))
                    self._written.add(u)
                    self._contrib[u] = self._contrib.get(u, 0) + 1
                    self._bst[u] = (plan[1], plan[0])
                    return
            return self._emit(u, lambda dx: self._call(
                "dl4jb200_conv_bwd", C.byref(d), self._ptr(x), L.NHWC, xcs, pw, self._ptr(e), L.NHWC, self.cs[n],
                None, None, dx, L.NHWC, xcs, self.algo))
        dst = self._eps_buf(u)
        prior = self._ptr(dst) if u in self._written else (self._ptr(lent[0]) if lent else None)
        # the delta of an Add vertex is the epsilon of the BatchNormalization folded into it: same trick
        bn = self.fused_add[u][0] if u in self.fused_add else None
        plan = self._bst_buffer(n, d, bn) if bn is not None else None
        if plan is not None:
            self._call("dl4jb200_conv_bwd_data_bnstats", C.byref(d), pw, self._ptr(e), self.cs[n], self._ptr(dst), xcs, prior,
                       self._ptr(self.out[u]), L.ACT_RELU, self.algo, *self._bst_args(bn), self._ptr(plan[1]))
            self._bst[bn] = (plan[1], plan[0])
        else:
            self._call("dl4jb200_conv_bwd_data_fused", C.byref(d), pw, self._ptr(e), self.cs[n], self._ptr(dst), xcs, prior,
                       self._ptr(self.out[u]), L.ACT_RELU, self.algo)
        self._written.add(u)
        self._borrowed.pop(u, None)
        self._contrib[u] = self._contrib.get(u, 0) + 1
        self._delta_done.add(u)

    def _trace(self, n, e):
        if self.trace is not None:
            with self.torch.cuda.stream(self.stream):
                self.trace[n] = e.clone()

    def _resolve(self, n):
        lent = self._borrowed.get(n, [])
        if n in self._written:
            e = self.eps[n]
            for t in lent:
                self._call("dl4jb200_accumulate", self._ptr(e), self._ptr(t), e.numel())
            return e
        if len(lent) == 1:
            return lent[0]
        if not lent:
            raise RuntimeError(f"no epsilon reached vertex {n}")
        e = self._eps_buf(n)
        self._call("dl4jb200_memcpy_d2d", self._ptr(e), self._ptr(lent[0]), e.numel() * 4)
        for t in lent[1:]:
            self._call("dl4jb200_accumulate", self._ptr(e), self._ptr(t), e.numel())
        return e

    def backward(self, join_side=True):
        L, B = _lib, self.B
        self._written, self._borrowed = set(), {}
        self._contrib, self._delta_done = {}, set()
        self._bst = {}
        self._side_used = False
        self._side_before_last, self._last_wgrad = None, None
        if self.dp is not None:
            self.dp.backward_begin(self)
        for v in reversed(self.model["vertices"]):
            n, op = v["name"], v["op"]
            self._cur = n
            if op in ("input", "zeropad") or (self.out.get(n, 0) is None and n not in self.fused_bn):
                continue
            self._backward_vertex(v)
            if self.dp is not None:
                self.dp.vertex_done(self, n)
        self._join_pending = None
        if self._side_used:  # join: gradient exchange / updater must see every filter gradient
            join = self.torch.cuda.Event()
            join.record(self.side_stream)
            if join_side:
                self.stream.wait_event(join)
            else:
                self._join_pending = join  # update() joins, after it has done everything that does not need the last one
        if self.dp is not None:
            self.dp.backward_end(self)

    def _backward_vertex(self, v):
        L, B = _lib, self.B
        n, op = v["name"], v["op"]
        if op == "output":
            e = self.aux[n]
        else:
            e = self._resolve(n)
        if op in ("conv", "dense", "output"):
            if op != "output" and v["act"] != "identity":
                # delta = afn.backprop(z, eps) (ConvolutionLayer.java:171, BaseLayer.java:71-102); only the activated
                # tensor is resident, so the derivative is taken from it: relu' = [a > 0], sigmoid' = a(1-a), tanh' = 1-a^2
                self._call("dl4jb200_act_bwd", self._ptr(self.out[n]), self._ptr(e), self._ptr(e), e.numel(),
                           _ACT[v["act"]] | L.ACT_FROM_OUTPUT)
            self._trace(n, e)
            if op == "conv":
                src, d = self._conv_desc(v)
            else:
                src, d = v["in"][0], self._dense_desc(v)
            x, xcs = self._lin(n, src)
            pw, gw, _, _ = self._p(n, "W")
            _, gb, _, _ = self._p(n, "b")

            side = self.overlap_wgrad
            wcall = self._call_side if side else self._call
            if side:  # fork: the side stream may start once everything queued so far (delta is final) has run
                # (remember where the side stream stood before the LAST filter gradient of the pass -- the first layer's,
                # which depends on the very last epsilon: update() does not have to wait for that one kernel, see there)
                self._side_before_last = self.torch.cuda.Event()
                self._side_before_last.record(self.side_stream)
                self._last_wgrad = n
                fork = self.torch.cuda.Event()
                fork.record(self.stream)
                self.side_stream.wait_event(fork)
                self._side_used = True
            if self.stem and self.stem[0] == n:
                wcall("dl4jb200_conv_rowpack_wgrad", C.byref(d), self._ptr(x), self.stem[1], self._ptr(e), self.cs[n], gw, gb)
            else:
                wcall("dl4jb200_conv_bwd", C.byref(d), self._ptr(x), L.NHWC, xcs, pw, self._ptr(e), L.NHWC, self.cs[n],
                      gw, gb, None, L.NHWC, xcs, self.algo)
                self._emit_dgrad(src, d, pw, e, n, x, xcs)
        elif op == "bn":
            self._trace(n, e)
            u = v["in"][0]
            x, xcs = self._src(u)
            s = self.shapes[n]
            pg, gg, _, _ = self._p(n, "gamma")
            pbeta, gbeta, _, _ = self._p(n, "beta")
            sm, ss = self.aux[n]
            pre = self._bst.pop(n, None)
            if pre is not None:  # the sums were taken by the GEMM that produced e: finalize + apply only
                self._emit(u, lambda dx: self._call("dl4jb200_bn_bwd_prestats", self._ptr(x), self._ptr(e), pg, pbeta, self._ptr(sm),
                                                    self._ptr(ss), _ACT[v["act"]], dx, gg, gbeta, B, s[0], s[1], s[2], xcs,
                                                    self._ptr(pre[0]), pre[1]))
            else:
                self._emit(u, lambda dx: self._call("dl4jb200_bn_bwd", self._ptr(x), self._ptr(e), pg, pbeta, self._ptr(sm),
                                                    self._ptr(ss), _ACT[v["act"]], dx, gg, gbeta, B, s[0], s[1], s[2], L.NHWC, xcs))
        elif op == "pool":
            self._trace(n, e)
            u = v["in"][0]
            x, xcs = self._lin(n, u)
            d = self._pool_desc(v)
            am = self._ptr(self.aux[n]) if v["type"] == "MAX" else None
            plan = self._pool_bst_plan(v, d)
            if plan is not None:  # this layer's backprop IS the epsilon of the batch norm below it: take that layer's sums here
                rows = C.c_int(0)
                dst = self._eps_buf(u)
                self._call("dl4jb200_pool_bwd_bnstats", C.byref(d), self._ptr(e), am, self._ptr(dst), self.cs[n], *self._bst_args(u),
                           self._ptr(plan), C.byref(rows))
                self._written.add(u)
                self._contrib[u] = self._contrib.get(u, 0) + 1
                self._bst[u] = (plan, rows.value)
                return
            self._emit(u, lambda dx: self._call("dl4jb200_pool_bwd", C.byref(d), self._ptr(x), self._ptr(self.out[n]), self._ptr(e), am,
                                                dx, L.NHWC, xcs, self.cs[n], _POOL[v["type"]], int(v.get("pnorm", 2)), 1e-8))
        elif op in ("add", "act"):
            if v["act"] != "identity" and n not in self._delta_done:
                self._call("dl4jb200_act_bwd", self._ptr(self.out[n]), self._ptr(e), self._ptr(e), e.numel(),
                           _ACT[v["act"]] | L.ACT_FROM_OUTPUT)
            self._trace(n, e)
            for u in v["in"]:
                self._lend(u, e)
        elif op == "flatten":
            self._trace(n, e)
            u = v["in"][0]
            s = self.shapes[u]
            chw = s[0] * s[1] * s[2]
            srcbuf = e
            if self.cs[n] != chw:
                self._call("dl4jb200_copy2d", self._ptr(e), self.cs[n], self._ptr(self.aux[n]), chw, B, chw)
                srcbuf = self.aux[n]
            self._emit(u, lambda dx: self._call("dl4jb200_nchw_to_nhwc", self._ptr(srcbuf), dx, B, s[0], s[1], s[2], self.cs[u]))
        else:
            raise NotImplementedError(op)

    # ------------------------------------------------------------------ updater + step (BaseMultiLayerUpdater.update)
    def _upd_args(self, cfg):
        k = cfg["kind"]
        if k == "rmsprop":
            return _lib.UPD_RMSPROP, cfg["lr"], cfg["decay"], cfg["eps"]
        if k == "nesterovs":
            return _lib.UPD_NESTEROVS, cfg["lr"], cfg["momentum"], 0.0
        if k == "adadelta":
            return _lib.UPD_ADADELTA, 0.0, cfg["rho"], cfg["eps"]
        if k == "sgd":
            return _lib.UPD_SGD, cfg["lr"], 0.0, 0.0
        raise NotImplementedError(k)

    def _build_segments(self):
        """Segment / configuration tables for dl4jb200_updater_apply_segments (one UpdaterBlock per parameter array)."""
        m, t = self.model, self.torch
        main, bias = self._upd_args(m["updater"]), self._upd_args(m.get("bias_updater", m["updater"]))
        cfgs, starts, idx = [], [], []

        def cfg_id(c):
            if c not in cfgs:
                cfgs.append(c)
            return cfgs.index(c)

        for name, key, off, shp, pk in self.layout:
            starts.append(off)
            if pk == "bn_stat":
                idx.append(-1)
                continue
            kind, lr, h1, h2 = bias if pk == "b" else main
            l1, l2 = (m["l1"], m["l2"]) if pk == "w" else (0.0, 0.0)
            idx.append(cfg_id((float(kind), float(lr), float(h1), float(h2), float(l1), float(l2))))
        starts.append(self.P)
        self.seg_start = t.tensor(starts, dtype=t.int64, device=self.dev)
        self.seg_cfg = t.tensor(idx, dtype=t.int32, device=self.dev)
        self.seg_cfg_host = (C.c_double * (6 * len(cfgs)))(*[x for c in cfgs for x in c])
        self.n_cfg = len(cfgs)

    def update(self):
        """All updater blocks + the L1/L2 score terms in one launch (update_per_block is the per-array form).
        After backward(join_side=False) the filter gradients may still be running on the side stream: everything except the
        LAST layer backward reached (the first convolution -- its filter gradient depends on the very last epsilon and is
        the tail of the step) is updated and re-packed as soon as the other gradients are final, and that layer's slice
        of the flat vector follows when its gradient arrives."""
        m = self.model
        if not hasattr(self, "seg_start"):
            self._build_segments()
        div = float(self.B * self.batch_div_scale) if m["minibatch"] else 1.0
        score = self._ptr(self.score_reg) if self.with_score else None
        args = (self._ptr(self.params), self._ptr(self.grads), self._ptr(self.state), self.P, self._ptr(self.seg_start),
                self._ptr(self.seg_cfg), len(self.layout), self.seg_cfg_host, self.n_cfg, div, score)
        pending = getattr(self, "_join_pending", None)
        self._join_pending = None
        late = self._late_span() if pending is not None else None
        if late is None:
            if pending is not None:
                self.stream.wait_event(pending)
            self._grad_normalize()
            self._call("dl4jb200_updater_apply_segments", *args)
            self.params_changed()
            return
        lo, hi = late
        self.stream.wait_event(self._side_before_last)  # every filter gradient but the last one
        first = True
        for a, b in ((0, lo), (hi, self.P)):
            if b > a:
                self._call("dl4jb200_updater_apply_segments_range", *args, a, b, 0 if first else 1)
                first = False
        self.params_changed()                            # the registered filters do not include the late layer's
        self.stream.wait_event(pending)
        self._call("dl4jb200_updater_apply_segments_range", *args, lo, hi, 0 if first else 1)

    def _late_span(self):
        """Flat-vector range of the layer whose filter gradient is issued last, when the split update applies."""
        n = self._last_wgrad
        if (n is None or self._side_before_last is None or self.model.get("grad_norm") or self.dp is not None
                or not (self.stem and self.stem[0] == n)):  # (its filter must not be among the re-packed ones)
            return None
        offs = [(off, off + cnt) for (name, key), (off, cnt, shp, pk) in self.pidx.items() if name == n]
        lo, hi = min(o[0] for o in offs), max(o[1] for o in offs)
        if lo % 4 or (hi % 4 and hi != self.P):
            return None
        return lo, hi

    def _grad_normalize(self):
        """GradientNormalization over each layer's gradient view (BaseMultiLayerUpdater.java:322-380)."""
        gn = self.model.get("grad_norm")
        if not gn:
            return
        mode = {"RenormalizeL2PerLayer": 1, "ClipL2PerLayer": 2, "ClipElementWiseAbsoluteValue": 3}.get(gn)
        if mode is None:
            raise NotImplementedError("gradient normalisation " + str(gn))
        thr = float(self.model.get("grad_norm_threshold", 1.0))
        spans = {}
        for name, key, off, shp, pk in self.layout:  # a layer's parameters are contiguous in the flat vector
            lo, hi = spans.get(name, (off, off))
            spans[name] = (min(lo, off), max(hi, off + int(np.prod(shp))))
        for name, (lo, hi) in spans.items():
            self._call("dl4jb200_grad_normalize", self._ptr(self.grads, lo), hi - lo, mode, thr)

    def update_per_block(self):
        m = self.model
        self._grad_normalize()
        div = float(self.B * self.batch_div_scale) if m["minibatch"] else 1.0
        main = self._upd_args(m["updater"])
        bias = self._upd_args(m.get("bias_updater", m["updater"]))
        adadelta = m["updater"]["kind"] == "adadelta"
        for name, key, off, shp, pk in self.layout:
            if pk == "bn_stat":  # NoOp updater, gradient view is zero (BatchNormalization.java:205-208)
                continue
            n = int(np.prod(shp))
            kind, lr, h1, h2 = bias if pk == "b" else main
            l1, l2 = (m["l1"], m["l2"]) if pk == "w" else (0.0, 0.0)
            if adadelta:
                # AdaDelta keeps msg|msdx per updater block: block-local 2n layout inside the 2P state vector
                st = self._ptr(self.state, 2 * off)
            else:
                st = self._ptr(self.state, off)
            self._call("dl4jb200_updater_apply", self._ptr(self.params, off), self._ptr(self.grads, off), st, n, kind,
                       float(lr), float(h1), float(h2), float(l1), float(l2), div)
        self.params_changed()

    def score_terms(self):
        """Queues the L1/L2 part of the score (BaseOutputLayer.java:82-100): call before update()."""
        m = self.model
        self._call("dl4jb200_memset", self._ptr(self.score_reg), 0, 4)
        if m["l1"] or m["l2"]:
            for name, key, off, shp, pk in self.layout:
                if pk == "w":
                    self._call("dl4jb200_l1l2_score", self._ptr(self.params, off), int(np.prod(shp)), float(m["l1"]),
                               float(m["l2"]), self._ptr(self.score_reg))

    def fit(self, features=None, labels=None, hook=None):
        """One MultiLayerNetwork/ComputationGraph.fit(DataSet) iteration. Returns the score tensor (device)."""
        if features is not None:
            self.set_batch(features, labels)
        self.forward(True)
        self.backward(join_side=hook is not None)
        if hook is not None:
            hook(self)  # data-parallel gradient exchange (parallel.py)
        self.update()  # also leaves the L1/L2 part of the score (pre-update weights) in score_reg
        return self._score()

    def _score(self):
        """(loss sum + L1/L2 terms) / minibatch into self.score (a library kernel: the captured step holds no framework kernels)."""
        self._call("dl4jb200_score_combine", self._ptr(self.loss), self._ptr(self.score_reg), self._ptr(self.score), self.B)
        return self.score

    # ------------------------------------------------------------------ CUDA graph of one iteration
    def capture(self, hook=None, include_input=True):
        """Record one fit() iteration (input re-layout, forward, backward, [gradient exchange], updater, score) into a
        CUDA graph on the network's stream. Every buffer the step touches is static, so the graph can be replayed for
        every minibatch; this removes the ~800 per-kernel launch gaps of the eager chain. Call after at least one eager
        fit() (scratch regions and epsilon buffers must already exist: nothing may be allocated while capturing)."""
        t = self.torch
        assert self.labels is not None and self._feat is not None, "run one eager fit(features, labels) first"
        self.stream.synchronize()
        self._graph = t.cuda.CUDAGraph()
        self._graph_has_input = include_input
        with t.cuda.graph(self._graph, stream=self.stream):
            if include_input:
                self._relayout_input()
            self.forward(True)
            self.backward(join_side=hook is not None)
            if hook is not None:
                hook(self)
            self.update()
            self._graph_score = self._score()
        # the graph's nodes hold scratch addresses of both contexts: from here on a call that would have to grow a
        # region fails loudly instead of freeing memory the graph still references (release_graph() lifts it)
        for ctx in (self.h, self.h2):
            _lib.call("dl4jb200_ctx_workspace_freeze", ctx, 1)
        self._graph_gen = self._ws_generation()
        return self._graph

    def _ws_generation(self):
        lib = _lib.cdll()
        return (lib.dl4jb200_ctx_workspace_generation(self.h), lib.dl4jb200_ctx_workspace_generation(self.h2))

    def release_graph(self):
        """Drop the captured step and unfreeze the scratch (call before running other shapes eagerly on this network)."""
        self.stream.synchronize()
        self._graph = None
        for ctx in (self.h, self.h2):
            _lib.call("dl4jb200_ctx_workspace_freeze", ctx, 0)

    def _replay(self):
        assert self._ws_generation() == self._graph_gen, "scratch was re-allocated after capture: re-capture the step"
        self._graph.replay()

    def fit_graph(self, features=None, labels=None):
        """fit() through the captured graph. New minibatches are copied into the static input buffers first."""
        t = self.torch
        if features is not None:
            with t.cuda.stream(self.stream):
                f = t.as_tensor(features) if not isinstance(features, t.Tensor) else features
                lab = t.as_tensor(labels) if not isinstance(labels, t.Tensor) else labels
                self._feat.copy_(f, non_blocking=True)
                self.labels.copy_(lab, non_blocking=True)
            if not self._graph_has_input:
                self._relayout_input()
        with t.cuda.stream(self.stream):
            self._replay()
        return self._graph_score

    # ------------------------------------------------------------------ prefetching input pipeline
    def prefetch(self, features, labels):
        """Start the host->device copy of the NEXT minibatch on a separate copy stream (double-buffered), so it overlaps
        the step that is running -- what MultiLayerNetwork.fit(DataSetIterator) gets from wrapping its iterator in an
        AsyncDataSetIterator (nn/multilayer/MultiLayerNetwork.java:1272-1275). Pair with fit_prefetched()."""
        t = self.torch
        if not hasattr(self, "_pf"):
            inp = self.model["vertices"][0]["name"]
            Cc, Hh, W = self.shapes[inp]
            K = self.shapes[self.model["vertices"][-1]["name"]][0]
            self._pf = {"stream": t.cuda.Stream(self.dev), "head": 0, "tail": 0,
                        "x": [t.empty(self.B, Cc, Hh, W, device=self.dev) for _ in range(2)],
                        "y": [t.empty(self.B, K, device=self.dev) for _ in range(2)],
                        "ready": [t.cuda.Event() for _ in range(2)], "free": [None, None]}
        pf = self._pf
        slot = pf["head"] % 2
        pf["head"] += 1
        f = t.as_tensor(features) if not isinstance(features, t.Tensor) else features
        lab = t.as_tensor(labels) if not isinstance(labels, t.Tensor) else labels
        with t.cuda.stream(pf["stream"]):
            if pf["free"][slot] is not None:
                pf["stream"].wait_event(pf["free"][slot])  # the step that consumed this slot has finished reading it
            pf["x"][slot].copy_(f, non_blocking=True)
            pf["y"][slot].copy_(lab, non_blocking=True)
            pf["ready"][slot].record(pf["stream"])

    def fit_prefetched(self, hook=None):
        """One iteration on the oldest prefetched minibatch (graph replay when capture(include_input=False) was made)."""
        t = self.torch
        pf = self._pf
        assert pf["tail"] < pf["head"], "nothing prefetched"
        slot = pf["tail"] % 2
        pf["tail"] += 1
        self.stream.wait_event(pf["ready"][slot])
        self._relayout_input(pf["x"][slot])
        with t.cuda.stream(self.stream):
            if self.labels is None:
                self.labels = t.empty_like(pf["y"][slot])
            self.labels.copy_(pf["y"][slot], non_blocking=True)
            ev = t.cuda.Event()
            ev.record(self.stream)
        pf["free"][slot] = ev
        if getattr(self, "_graph", None) is not None and not self._graph_has_input:
            with t.cuda.stream(self.stream):
                self._replay()
            return self._graph_score
        return self.fit(hook=hook)

    def sync(self):
        """Host waits for everything queued on the network's stream (results are then safe to read from any stream)."""
        self.stream.synchronize()

    def output_host(self, name):
        """NCHW host copy of a vertex activation (tests)."""
        t, cs = self._src(name)
        s = self.shapes[name]
        self.stream.synchronize()
        a = t.cpu().numpy()
        if len(s) == 3:
            return np.ascontiguousarray(a[..., :s[0]].transpose(0, 3, 1, 2))
        return a[:, :s[0]].copy()

    # ------------------------------------------------------------------ checkpoint / resume (ModelSerializer layout)
    def save_checkpoint(self, path, configuration_json="{}", save_updater=True):
        """ModelSerializer.writeModel: the device-resident flat vectors are brought to the host (the parameter boundary) and
        written as coefficients.bin / updaterState.bin (util/ModelSerializer.java:120-163)."""
        from . import modelserializer as ms
        self.stream.synchronize()
        ms.write_model(path, configuration_json, self.params.cpu().numpy(), self.state.cpu().numpy() if save_updater else None)

    def load_checkpoint(self, path):
        """ModelSerializer.restore*: parameters (and updater state when present) go back to the device."""
        from . import modelserializer as ms
        t = self.torch
        conf, params, state = ms.restore_model(path)
        assert params.size == self.P, f"checkpoint holds {params.size} parameters, the network has {self.P}"
        with t.cuda.stream(self.stream):
            self.params.copy_(t.from_numpy(params), non_blocking=False)
            if state is not None:
                assert state.size == self.state.numel()
                self.state.copy_(t.from_numpy(state), non_blocking=False)
        self.params_changed()
        return conf

    def close(self):
        if self.h:
            self.torch.cuda.synchronize(self.dev)
            _lib.call("dl4jb200_ctx_destroy", self.h)
            _lib.call("dl4jb200_ctx_destroy", self.h2)
            self.h = None
