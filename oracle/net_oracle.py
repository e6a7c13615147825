"""CPU restatement of one DL4J training iteration on a zoo-style model graph (TEST INFRASTRUCTURE ONLY).

Follows computeGradientAndScore -> updater -> params.subi exactly as SURVEY.md Appendix A states it
(nn/multilayer/MultiLayerNetwork.java:2360-2420; nn/graph/ComputationGraph.java:1321-1360, 2044-2063 for the
epsilon fan-in; nn/updater/BaseMultiLayerUpdater.java:223-309; optimize/solvers/StochasticGradientDescent.java:78).
The model dict and the flat parameter layout are passed in by the caller (tests / bench); nothing here is
imported by the product package. fast=True runs the reference's literal im2col + BLAS sgemm + col2im pipeline in
fp32 (the form timed as the CPU baseline); fast=False runs the direct double-accumulate form (parity oracle).

tf32=True additionally restates the OPERAND QUANTISATION of the CUDA path's TF32 mode (include/dl4jb200.h,
DL4JB200_OPT_TF32_ACTIVATIONS): filters are rounded to nearest TF32 when packed, convolution operands are truncated to
TF32 by the tensor core, and every producer of a tensor that a convolution consumes stores it rounded to nearest TF32.
All arithmetic stays the reference's; only those quantisation points are added, so the TF32 path can be checked at
fp32-accumulation tolerance instead of at "TF32 is ~1e-3 per layer" tolerance. `unrounded` names the vertices whose
output the CUDA path never materialises (a BatchNormalization folded into the following Add).
"""
import numpy as np

from . import oracle as o

_ACT = {"identity": o.ACT_IDENTITY, "relu": o.ACT_RELU, "sigmoid": o.ACT_SIGMOID, "tanh": o.ACT_TANH}
_MODE = {"Strict": o.STRICT, "Truncate": o.TRUNCATE, "Same": o.SAME}
_POOL = {"MAX": o.POOL_MAX, "AVG": o.POOL_AVG, "PNORM": o.POOL_PNORM}


def tf32_round(a):
    """cvt.rna.tf32.f32: round the magnitude to 10 explicit mantissa bits, ties away from zero."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    return ((u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)


def tf32_trunc(a):
    """What tcgen05.mma kind::tf32 does to an fp32 operand: the low 13 mantissa bits are ignored."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    return (u & np.uint32(0xFFFFE000)).view(np.float32)


class OracleNet:
    def __init__(self, model, layout, P, params, fast=False, tf32=False, unrounded=()):
        self.model, self.layout, self.P, self.fast = model, layout, P, fast
        self.tf32, self.unrounded = bool(tf32), set(unrounded)
        self._q = tf32_round if tf32 else (lambda a: a)   # producer-side store
        self._t = tf32_trunc if tf32 else (lambda a: a)   # tensor-core operand fetch
        self.params = np.ascontiguousarray(params, np.float32).copy()
        self.grads = np.zeros(P, np.float32)
        k = model["updater"]["kind"]
        self.state = np.zeros(P * (2 if k == "adadelta" else 1), np.float32)
        self.pidx = {(n, key): (off, int(np.prod(shp)), shp, pk) for n, key, off, shp, pk in layout}
        self.vmap = {v["name"]: v for v in model["vertices"]}
        self.masks = None  # {layer name: 0/1 array shaped like the layer's input}: dropout is on when set

    def _layer_input(self, v, a, train):
        """AbstractLayer.applyDropOutIfNecessary (AbstractLayer.java:360-374): in training the layer's input becomes
        x * mask / p (nd4j DropOutInverted, Dropout.java:75); backprop uses that dropped input, and there is no dropout
        backprop step -- the epsilon the layer returns goes to the previous layer unmasked."""
        p = v.get("dropout")
        if train and p and self.masks is not None:
            a = self._q((a * self.masks[v["name"]].reshape(a.shape).astype(np.float32) * np.float32(1.0 / p)).astype(np.float32))
        self.lin[v["name"]] = a
        return a

    def _p(self, n, key, arr=None):
        off, cnt, shp, _ = self.pidx[(n, key)]
        a = self.params if arr is None else arr
        return a[off:off + cnt].reshape(shp)

    def forward(self, x, labels=None, train=True):
        m = self.model
        act = {}
        self.bn_saved = {}
        self.lin = {}
        self.z = {}  # pre-activation of every vertex with a non-identity activation: IActivation.backprop(z, eps) needs it
        for v in m["vertices"]:
            n, op = v["name"], v["op"]
            if op == "input":
                act[n] = self._q(np.ascontiguousarray(x, np.float32))
                continue
            a = act[v["in"][0]]
            if op == "zeropad":
                t, b, l, r = v["pad"]
                act[n] = np.pad(a, ((0, 0), (0, 0), (t, b), (l, r)))
            elif op == "conv":
                a = self._layer_input(v, a, train)
                f = o.conv_fwd_im2col if self.fast else o.conv_fwd
                z = f(self._t(a), self._q(self._p(n, "W")), self._p(n, "b"), stride=v["stride"], pad=v["pad"],
                      dilation=v["dilation"], mode=_MODE[v["mode"]])
                self.z[n] = z
                act[n] = self._q(o.act_fwd(z, _ACT[v["act"]])) if v["act"] != "identity" else z
            elif op == "bn":
                y, mean, std = o.bn_fwd(a, self._p(n, "gamma"), self._p(n, "beta"), self._p(n, "mean"), self._p(n, "var"),
                                        v["decay"], v["eps"], train, m["minibatch"])
                self.bn_saved[n] = (mean, std, y)
                y = o.act_fwd(y, _ACT[v["act"]]) if v["act"] != "identity" else y
                act[n] = y if n in self.unrounded else self._q(y)
            elif op == "pool":
                a = self._layer_input(v, a, train)
                act[n] = self._q(o.pool_fwd(a, v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]], _POOL[v["type"]]))
            elif op == "add":
                s = a + act[v["in"][1]]
                self.z[n] = s
                act[n] = self._q(o.act_fwd(s, _ACT[v["act"]]) if v["act"] != "identity" else s)
            elif op == "act":
                self.z[n] = a
                act[n] = self._q(o.act_fwd(a, _ACT[v["act"]]))
            elif op == "flatten":
                act[n] = a.reshape(a.shape[0], -1)
            elif op in ("dense", "output"):
                a = self._layer_input(v, a, train)
                if self.fast:
                    z = self._t(a) @ self._q(self._p(n, "W")).T + self._p(n, "b")[None, :]
                else:
                    z = o.dense_fwd(self._t(a), self._q(self._p(n, "W")), self._p(n, "b"))
                if op == "dense":
                    self.z[n] = z
                    act[n] = self._q(o.act_fwd(z, _ACT[v["act"]])) if v["act"] != "identity" else z
                else:
                    act[n] = z
                    if labels is not None:
                        self.loss, self.probs, self.delta = o.softmax_nll(z, labels)
                        self.delta = self._q(self.delta)
            else:
                raise NotImplementedError(op)
        self.act = act
        return act

    def backward(self):
        m, act = self.model, self.act
        eps = {}
        self.deltas = {}  # per vertex: the summed epsilon after the vertex' own activation backprop (raw for BN / pooling)

        def give(u, e):
            if self.vmap[u]["op"] in ("input",):
                return
            eps[u] = e if u not in eps else eps[u] + e

        for v in reversed(m["vertices"]):
            n, op = v["name"], v["op"]
            if op == "input":
                continue
            u = v["in"][0]
            e = self.delta if op == "output" else eps.get(n)
            if e is None:
                continue
            self.deltas[n] = e
            if op == "zeropad":
                t, b, l, r = v["pad"]
                give(u, np.ascontiguousarray(e[:, :, t:e.shape[2] - b, l:e.shape[3] - r]))
            elif op == "conv":
                d = self._q(o.act_bwd(self.z[n], e, _ACT[v["act"]])) if v["act"] != "identity" else e
                self.deltas[n] = d
                f = o.conv_bwd_im2col if self.fast else o.conv_bwd
                dw, db, dx = f(self._t(self.lin[n]), self._q(self._p(n, "W")), self._t(d), stride=v["stride"], pad=v["pad"],
                               dilation=v["dilation"], mode=_MODE[v["mode"]])
                if self.tf32:  # the bias gradient is a plain fp32 column sum of the stored delta, no tensor core
                    db = d.sum(axis=(0, 2, 3), dtype=np.float64).astype(np.float32)
                self._p(n, "W", self.grads)[...] = dw
                self._p(n, "b", self.grads)[...] = db
                give(u, self._q(dx))  # backprop-data stores its epsilon rounded like every other producer
            elif op in ("dense", "output"):
                d = self._q(o.act_bwd(self.z[n], e, _ACT[v["act"]])) if (op == "dense" and v["act"] != "identity") else e
                self.deltas[n] = d
                if self.fast:
                    dwf, db, dx = self._t(d).T @ self._t(self.lin[n]), d.sum(0), self._t(d) @ self._q(self._p(n, "W"))
                else:
                    dwf, db, dx = o.dense_bwd(self._t(self.lin[n]), self._q(self._p(n, "W")), self._t(d))
                    if self.tf32:
                        db = d.sum(axis=0, dtype=np.float64).astype(np.float32)
                self._p(n, "W", self.grads)[...] = dwf
                self._p(n, "b", self.grads)[...] = db
                give(u, self._q(dx))
            elif op == "bn":
                mean, std, y = self.bn_saved[n]
                d = o.act_bwd(y, e, _ACT[v["act"]]) if v["act"] != "identity" else e  # y: the pre-activation batch-norm output
                dx, dg, dbt = o.bn_bwd(act[u], d, self._p(n, "gamma"), mean, std)
                self._p(n, "gamma", self.grads)[...] = dg
                self._p(n, "beta", self.grads)[...] = dbt
                self._p(n, "mean", self.grads)[...] = 0  # BatchNormalization.java:205-208
                self._p(n, "var", self.grads)[...] = 0
                give(u, self._q(dx))
            elif op == "pool":
                give(u, self._q(o.pool_bwd(self.lin[n], e, v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]], _POOL[v["type"]])))
            elif op in ("add", "act"):
                d = self._q(o.act_bwd(self.z[n], e, _ACT[v["act"]])) if v["act"] != "identity" else e
                self.deltas[n] = d
                for uu in v["in"]:
                    give(uu, d)
            elif op == "flatten":
                give(u, self._q(e.reshape(act[u].shape)))  # the CUDA path re-lays the epsilon out (a producer store)
            else:
                raise NotImplementedError(op)

    def _upd(self, cfg):
        k = cfg["kind"]
        if k == "rmsprop":
            return o.UPD_RMSPROP, dict(lr=cfg["lr"], h1=cfg["decay"], h2=cfg["eps"])
        if k == "nesterovs":
            return o.UPD_NESTEROVS, dict(lr=cfg["lr"], h1=cfg["momentum"])
        if k == "adadelta":
            return o.UPD_ADADELTA, dict(h1=cfg["rho"], h2=cfg["eps"])
        return o.UPD_SGD, dict(lr=cfg["lr"])

    def score(self, B):
        m = self.model
        reg = 0.0
        for n, key, off, shp, pk in self.layout:
            if pk == "w":
                w = self.params[off:off + int(np.prod(shp))].astype(np.float64)
                reg += m["l1"] * np.abs(w).sum() + 0.5 * m["l2"] * (w * w).sum()
        return (self.loss + reg) / B

    def grad_normalize(self):
        # BaseMultiLayerUpdater.preApply, nn/updater/BaseMultiLayerUpdater.java:322-380 (per layer gradient view)
        gn = self.model.get("grad_norm")
        if not gn:
            return
        thr = float(self.model.get("grad_norm_threshold", 1.0))
        spans = {}
        for n, key, off, shp, pk in self.layout:
            lo, hi = spans.get(n, (off, off))
            spans[n] = (min(lo, off), max(hi, off + int(np.prod(shp))))
        for n, (lo, hi) in spans.items():
            g = self.grads[lo:hi]
            l2 = float(np.sqrt((g.astype(np.float64) ** 2).sum()))
            if gn == "RenormalizeL2PerLayer":
                g /= np.float32(l2)
            elif gn == "ClipL2PerLayer":
                if l2 > thr:
                    g *= np.float32(thr / l2)
            elif gn == "ClipElementWiseAbsoluteValue":
                np.clip(g, -thr, thr, out=g)
            else:
                raise NotImplementedError(gn)

    def update(self, B):
        m = self.model
        self.grad_normalize()
        div = float(B) if m["minibatch"] else 1.0
        main, bias = self._upd(m["updater"]), self._upd(m.get("bias_updater", m["updater"]))
        ada = m["updater"]["kind"] == "adadelta"
        for n, key, off, shp, pk in self.layout:
            if pk == "bn_stat":
                continue
            cnt = int(np.prod(shp))
            kind, args = bias if pk == "b" else main
            l1, l2 = (m["l1"], m["l2"]) if pk == "w" else (0.0, 0.0)
            st = self.state[2 * off:2 * off + 2 * cnt] if ada else self.state[off:off + cnt]
            o.updater_apply(self.params[off:off + cnt], self.grads[off:off + cnt], st, kind, l1=l1, l2=l2, batch_div=div, **args)

    def fit(self, x, labels):
        B = x.shape[0]
        self.forward(x, labels, True)
        self.backward()
        s = self.score(B)
        self.update(B)
        return s
