"""Vertex-by-vertex parity of one training step of the CUDA path against the CPU oracle (TEST INFRASTRUCTURE).

A free-running comparison of two whole steps measures the CONDITIONING of the network as much as the kernels: one
TF32 rounding flip in an early activation of an untrained 50-layer net moves single gradient blocks by 10 % on either
implementation. This module removes the amplification instead of loosening the tolerance: after one fit() on the GPU
it takes the tensors the GPU actually produced (every stored activation, every vertex' final delta, saved batch-norm
statistics, the flat gradient, the pre- and post-update parameters and updater state) and checks EVERY vertex on its
own -- the oracle restatement of that vertex (ConvolutionLayer / SubsamplingLayer / BatchNormalization / ... exactly as
OracleNet runs them) is fed the GPU's inputs of that vertex and must reproduce the GPU's outputs of that vertex:

  forward   out[n]      = layer_n(out[inputs of n])                                  for every vertex n
  backward  delta[n]    = act'_n . sum over consumers c of  backprop_c(delta[c])     (ComputationGraph.java:2044-2063)
            grad blocks = dW, db, dgamma, dbeta of n from (input of n, delta[n])
  update    params'     = updater(params, grads, state)  block by block             (BaseMultiLayerUpdater.java:223-309)
  score     (loss + L1/L2) / B

Since both sides see identical inputs at every vertex, what is left is (a) fp32 summation order and (b), in TF32 mode,
the ONE rounding to nearest TF32 each producer applies when it stores a tensor that a convolution will consume
(DL4JB200_OPT_TF32_ACTIVATIONS); the filter rounding and the operand truncation are restated exactly (net_oracle.tf32_*).
Hence the bound per element is  |gpu - oracle| <= stages * 2^-11 * magnitude + fp_tol * max|oracle|,  with `stages` the
number of roundings on the GPU's path (0 in FP32 mode and for gradient blocks, which are never rounded) -- i.e. gradient
and post-update parameter blocks are held to fp32-summation accuracy in the very arithmetic bench.py times.
ReLU gates recomputed from a batch-norm output that is within round-off of zero may fall either way; those elements are
counted, must be rare, and are compared separately.
"""
import numpy as np

from oracle import oracle as o
from oracle.net_oracle import OracleNet, tf32_round, tf32_trunc

_ACT = {"identity": o.ACT_IDENTITY, "relu": o.ACT_RELU, "sigmoid": o.ACT_SIGMOID, "tanh": o.ACT_TANH}
_MODE = {"Strict": o.STRICT, "Truncate": o.TRUNCATE, "Same": o.SAME}
_POOL = {"MAX": o.POOL_MAX, "AVG": o.POOL_AVG, "PNORM": o.POOL_PNORM}
HALF_ULP = 2.0 ** -11  # round-to-nearest into a 10-bit mantissa


class Report:
    def __init__(self):
        self.rows, self.flips, self.flip_candidates = [], 0, 0

    def check(self, what, name, got, ref, stages, fp_tol, mag=None, skip=None, floor=None):
        """|got - ref| <= stages*2^-11*mag + fp_tol*max|ref| (+ floor) elementwise; records the worst |got-ref| / max|ref|.
        floor: absolute fp32 round-off allowance for sums that cancel to ~0 (e.g. the bias gradient of a convolution that
        feeds a batch norm is analytically zero): 2e-6 * sum|terms| -- the bias gradient rides in the filter-gradient GEMM
        (delta^T x ones), whose fp32 accumulator truncates; measured up to 1.2e-6 * sum|terms| over 4e5 pixels."""
        got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
        assert got.shape == ref.shape, (what, name, got.shape, ref.shape)
        scale = max(float(np.abs(ref).max()), 1e-30)
        if floor is not None and float(np.max(floor)) * 100 > scale:
            what += ".cancelling"  # the true value is round-off of a cancelling sum: judged on sum|terms|, listed apart
            scale = max(scale, float(np.max(floor)) * 100)
        err = np.abs(got - ref)
        bound = stages * HALF_ULP * 1.002 * (np.abs(ref) if mag is None else np.asarray(mag, np.float64)) + fp_tol * scale
        if floor is not None:
            bound = bound + np.asarray(floor, np.float64)
        bad = err > bound
        if skip is not None:
            bad &= ~skip
        worst = float(err.max() / scale)
        self.rows.append((what, name, worst, int(bad.sum()), got.size))
        assert not bad.any(), (what, name, "elements over bound:", int(bad.sum()), "of", got.size, "worst rel err", worst,
                               "worst excess", float((err - bound).max() / scale))
        return worst

    def worst(self, what):
        r = [x for x in self.rows if x[0] == what]
        return max(r, key=lambda x: x[2]) if r else None

    def summary(self):
        kinds = sorted({r[0] for r in self.rows})
        return {k: (self.worst(k)[1], self.worst(k)[2]) for k in kinds}


def _nchw(net, t, name):
    """device tensor laid out like out[name] -> NCHW / [B, n] host array."""
    s = net.shapes[name]
    a = t.cpu().numpy()
    if len(s) == 3:
        return np.ascontiguousarray(a[..., :s[0]].transpose(0, 3, 1, 2))
    if a.ndim == 4:  # an aliased flatten of a 1x1 map: the buffer is the [B,1,1,cs] NHWC tensor itself
        a = a.reshape(a.shape[0], -1)
    return a[:, :s[0]].copy()


def _tap_views(o, shape, taps, v):
    """yields (tap index, slices into the zero-padded input) for every tap of a pooling window (no dilation)."""
    (kH, kW), (sH, sW) = v["kernel"], v["stride"]
    Ho, Wo, pT, pL = o.conv_geometry(shape[2], shape[3], v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]])
    assert taps.shape[2:] == (Ho, Wo)
    Hp, Wp = max(pT + shape[2], (Ho - 1) * sH + kH), max(pL + shape[3], (Wo - 1) * sW + kW)
    views = [(kh * kW + kw, (slice(None), slice(None), slice(kh, kh + sH * (Ho - 1) + 1, sH), slice(kw, kw + sW * (Wo - 1) + 1, sW)))
             for kh in range(kH) for kw in range(kW)]
    return views, (Hp, Wp, pT, pL)


def _gather_taps(o, a, taps, v):
    """value of the window tap `taps` names (kh * kW + kw), per pooled element."""
    views, (Hp, Wp, pT, pL) = _tap_views(o, a.shape, taps, v)
    ap = np.full(a.shape[:2] + (Hp, Wp), -np.inf, np.float32)
    ap[:, :, pT:pT + a.shape[2], pL:pL + a.shape[3]] = a
    out = np.zeros(taps.shape, np.float32)
    for t, sl in views:
        np.copyto(out, ap[sl], where=taps == t)
    return out


def _scatter_taps(o, d, taps, v, in_shape):
    """SubsamplingLayer.backpropGradient for MAX with given taps: each epsilon goes to the input cell its tap names."""
    views, (Hp, Wp, pT, pL) = _tap_views(o, in_shape, taps, v)
    dxp = np.zeros(tuple(in_shape[:2]) + (Hp, Wp), np.float64)
    for t, sl in views:
        dxp[sl] += np.where(taps == t, d, 0)
    return dxp[:, :, pT:pT + in_shape[2], pL:pL + in_shape[3]]


def replay_step(net, model, x, lab, masks=None, fp_tol=1e-4, grad_tol=1e-4, param_tol=1e-5, max_flip_frac=2e-4,
                forward_only=False):
    """Runs ONE net.fit(x, lab) and checks it vertex by vertex (see module docstring). Returns a Report.
    forward_only: a training-mode forward pass only (SimpleCNN has no loss layer and cannot be fitted)."""
    tf = bool(net.round_tf32)
    q = tf32_round if tf else (lambda a: np.ascontiguousarray(a, np.float32))
    t_ = tf32_trunc if tf else (lambda a: np.ascontiguousarray(a, np.float32))
    st = 1 if tf else 0
    B = x.shape[0]
    rep = Report()
    vmap = {v["name"]: v for v in model["vertices"]}

    if masks:
        net.set_dropout_masks(masks)
    params_pre = net.params.cpu().numpy().copy()
    state_pre = net.state.cpu().numpy().copy()
    net.trace = {}
    if forward_only:
        net.set_batch(x, lab)
        net.forward(True)
        net.sync()
        score = None
    else:
        score = net.fit(x, lab)
        net.sync()
        score = float(score.cpu()[0])
    trace, net.trace = net.trace, None
    grads = net.grads.cpu().numpy()
    params_post = net.params.cpu().numpy()
    loss_gpu = float(net.loss.cpu()[0])

    def P(n, key, arr=params_pre):
        off, cnt, shp, _ = net.pidx[(n, key)]
        return arr[off:off + cnt].reshape(shp)

    # ---------------------------------------------------------------- tensors the GPU produced
    A = {}  # activations, NCHW host

    def act_of(u):
        if u in A:
            return A[u]
        v = vmap[u]
        if v["op"] == "input":
            A[u] = q(x)
        elif v["op"] == "zeropad":
            t, b, l, r = v["pad"]
            A[u] = np.pad(act_of(v["in"][0]), ((0, 0), (0, 0), (t, b), (l, r)))
        elif v["op"] == "flatten" and net.out.get(u, 0) is None:
            a = act_of(v["in"][0])
            A[u] = a.reshape(a.shape[0], -1)
        elif u in net.fused_bn:
            raise KeyError(u)  # never materialised
        else:
            A[u] = net.output_host(u)
        return A[u]

    def lin_of(v):
        """the layer's input as the GPU saw it (dropped copy when the layer has dropout)."""
        n = v["name"]
        if n in net.drop:
            if getattr(net, "stem", None) and net.stem[0] == n:  # row-packed first layer: [B, H, pitch, 4], image at column x_off
                Cc, Hh, Ww = net.shapes[v["in"][0]]
                a = net.drop[n].cpu().numpy()
                return np.ascontiguousarray(a[:, :, net.stem[2]:net.stem[2] + Ww, :Cc].transpose(0, 3, 1, 2))
            return _nchw(net, net.drop[n], v["in"][0])
        return act_of(v["in"][0])

    zref = {}      # oracle pre-activations (for sigmoid / tanh backprop)
    pooled_bn = {}  # BN vertex folded into its max-pooling consumer -> the oracle's unrounded normalised (+activated) tensor
    taps = {}      # such a pooling vertex -> the window taps the GPU saved (routes the backward pass)
    bn_y = {}      # oracle BN output before the fused activation, from the GPU's saved statistics (float64)

    # ---------------------------------------------------------------- forward, vertex by vertex
    for v in model["vertices"]:
        n, op = v["name"], v["op"]
        if op in ("input", "zeropad"):
            continue
        if op == "conv":
            a = lin_of(v)
            if n in net.drop:
                m = masks[n].reshape(act_of(v["in"][0]).shape).astype(np.float32)
                rep.check("fwd.dropout", n, a, act_of(v["in"][0]) * m * np.float32(1.0 / v["dropout"]), st, fp_tol)
            z = o.conv_fwd_im2col(t_(a), q(P(n, "W")), P(n, "b"), stride=v["stride"], pad=v["pad"], dilation=v["dilation"],
                                  mode=_MODE[v["mode"]])
            zref[n] = z
            ref = o.act_fwd(z, _ACT[v["act"]]) if v["act"] != "identity" else z
            rep.check("fwd.conv", n, act_of(n), ref, st if v["act"] != "identity" else 0, fp_tol)
        elif op == "bn":
            xin = act_of(v["in"][0])
            gm, gv = P(n, "mean").copy(), P(n, "var").copy()
            y, mean, std = o.bn_fwd(xin, P(n, "gamma"), P(n, "beta"), gm, gv, v["decay"], v["eps"], True, model["minibatch"])
            sm, ss = (t.cpu().numpy()[:mean.size] for t in net.aux[n])
            rep.check("fwd.bn_mean", n, sm, mean, 0, fp_tol)
            rep.check("fwd.bn_std", n, ss, std, 0, fp_tol)
            rep.check("fwd.bn_running_mean", n, P(n, "mean", params_post), gm, 0, fp_tol)
            rep.check("fwd.bn_running_var", n, P(n, "var", params_post), gv, 0, fp_tol)
            bn_y[n] = y
            if n in net.fused_bn:
                continue  # checked with the Add it is folded into
            if n in getattr(net, "fused_pool_active", ()):
                # folded into the max-pooling layer that consumes it (never written): the pooling branch below checks the
                # GPU's pooled output and saved taps against the oracle's pooling of the oracle's normalised tensor
                ref = o.act_fwd(y, _ACT[v["act"]]) if v["act"] != "identity" else y
                pooled_bn[n] = np.asarray(ref, np.float32)
                A[n] = q(ref)
                continue
            ref = o.act_fwd(y, _ACT[v["act"]]) if v["act"] != "identity" else y
            rep.check("fwd.bn", n, act_of(n), ref, st, fp_tol)
        elif op == "pool" and v["in"][0] in pooled_bn:
            # BatchNormalization + ReLU + MAX pooling in one pass: rounding is monotone, so the stored maximum is the rounded
            # maximum of the unrounded normalised window -- same allowance as a stored BatchNormalization output. The saved
            # tap must hold that maximum (a different tap is only acceptable where two taps agree to within that allowance).
            a = pooled_bn[v["in"][0]]
            ref = o.pool_fwd(a, v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]], _POOL[v["type"]])
            rep.check("fwd.pool", n, act_of(n), ref, st, fp_tol)
            taps[n] = _nchw(net, net.aux[n], n)
            # two taps that round to the same stored value tie (the first wins): up to one full ulp = two roundings apart
            rep.check("fwd.pool_tap", n, _gather_taps(o, a, taps[n], v), ref, 2 * st, fp_tol)
        elif op == "pool":
            a = lin_of(v)
            ref = o.pool_fwd(a, v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]], _POOL[v["type"]])
            rep.check("fwd.pool", n, act_of(n), ref, st if v["type"] != "MAX" else 0, 0.0 if v["type"] == "MAX" else fp_tol)
        elif op == "add":
            a = bn_y[v["in"][0]] if v["in"][0] in net.fused_bn else act_of(v["in"][0])
            b = bn_y[v["in"][1]] if v["in"][1] in net.fused_bn else act_of(v["in"][1])
            s = (a + b).astype(np.float32)
            zref[n] = s
            rep.check("fwd.add", n, act_of(n), o.act_fwd(s, _ACT[v["act"]]) if v["act"] != "identity" else s, st, fp_tol)
        elif op == "act":
            zref[n] = act_of(v["in"][0])
            rep.check("fwd.act", n, act_of(n), o.act_fwd(zref[n], _ACT[v["act"]]), st, fp_tol)
        elif op == "flatten":
            a = act_of(v["in"][0])
            rep.check("fwd.flatten", n, act_of(n), a.reshape(a.shape[0], -1), 0, 0.0)
        elif op in ("dense", "output"):
            a = lin_of(v)
            z = (t_(a).astype(np.float64) @ q(P(n, "W")).T.astype(np.float64) + P(n, "b")[None, :]).astype(np.float32)
            zref[n] = z
            if op == "dense":
                ref = o.act_fwd(z, _ACT[v["act"]]) if v["act"] != "identity" else z
                rep.check("fwd.dense", n, act_of(n), ref, st if v["act"] != "identity" else 0, fp_tol)
            else:
                rep.check("fwd.output", n, act_of(n), z, 0, fp_tol)
                loss, _, delta = o.softmax_nll(act_of(n), lab)
                assert abs(loss_gpu - loss) <= 1e-5 * max(1.0, abs(loss)), (loss_gpu, loss)
                out_delta = delta
        else:
            raise NotImplementedError(op)

    if forward_only:
        return rep

    # ---------------------------------------------------------------- backward, vertex by vertex
    D = {n: _nchw(net, t, n) for n, t in trace.items()}
    exp, mag = {}, {}  # expected epsilon sums per vertex (+ the magnitude that bounds the roundings on the way)

    def give(u, e, m=None):
        while vmap[u]["op"] == "flatten" and net.out.get(u, 0) is None:  # alias: straight through to its input
            src = vmap[u]["in"][0]
            e = e.reshape(act_of(src).shape)
            m = None if m is None else m.reshape(e.shape)
            u = src
        if vmap[u]["op"] in ("input", "zeropad"):
            return
        e = np.asarray(e, np.float64)
        m = np.abs(e) if m is None else np.asarray(m, np.float64)
        if u in exp:
            exp[u] = exp[u] + e
            mag[u] = mag[u] + m + np.abs(exp[u])  # one more stored (rounded) partial sum
        else:
            exp[u], mag[u] = e, m

    gref = np.zeros(net.P, np.float64)   # the oracle's gradient blocks (each from the GPU's inputs of that vertex)
    gfloor = np.zeros(net.P, np.float64)  # round-off allowance for blocks that are cancelling sums
    gtol = np.full(net.P, grad_tol)

    def put(n, key, val, tol=grad_tol, floor=None):
        off, cnt, shp, _ = net.pidx[(n, key)]
        gref[off:off + cnt] = np.asarray(val, np.float64).reshape(-1)
        gtol[off:off + cnt] = tol
        if floor is not None:
            gfloor[off:off + cnt] = np.asarray(floor, np.float64).reshape(-1)

    amb = {}  # vertex -> elements whose incoming epsilon passed a ReLU gate that is within round-off of zero
    for v in reversed(model["vertices"]):
        n, op = v["name"], v["op"]
        if op in ("input", "zeropad") or (op == "flatten" and net.out.get(n, 0) is None):
            continue
        # --- this vertex' delta against the oracle's fan-in sum of its consumers' contributions
        if op == "output":
            rep.check("bwd.delta", n, D[n], out_delta, st, fp_tol)
        else:
            assert n in exp, f"no epsilon reached {n}"
            e = exp[n]
            if op in ("conv", "dense", "add", "act") and v["act"] != "identity":
                if v["act"] == "relu":
                    gate = act_of(n) > 0
                    e, m = e * gate, mag[n] * gate
                else:
                    # reference semantics: afn.backprop(z, eps). The CUDA path takes the derivative from the STORED output
                    # a (sigmoid' = a(1-a), tanh' = 1-a^2); in TF32 mode a carries one rounding, which moves the derivative
                    # by at most 2 * 2^-11 in absolute terms -- allowed for on top of the magnitudes already collected
                    g = o.act_bwd(zref[n], np.ones_like(zref[n]), _ACT[v["act"]]).astype(np.float64)
                    e, m = e * g, mag[n] * np.abs(g) + 2.0 * np.abs(e)
            else:
                m = mag[n]
            # every contribution and every partial sum was stored once, rounded: `m` adds up exactly those magnitudes
            rep.check("bwd.delta", n, D[n], e, st, fp_tol, mag=m, skip=amb.get(n))
            del exp[n], mag[n]
        d = D[n]
        # --- its gradient blocks and its contributions to its inputs, from the GPU's own delta
        if op == "conv":
            dw, db, dx = o.conv_bwd_im2col(t_(lin_of(v)), q(P(n, "W")), t_(d), stride=v["stride"], pad=v["pad"],
                                           dilation=v["dilation"], mode=_MODE[v["mode"]], acc64=True)
            db = d.sum(axis=(0, 2, 3), dtype=np.float64)
            # the tensor core's fp32 accumulator truncates: over reductions of > 2.6e5 pixels (the stem) the filter gradient
            # reaches 1.2e-4 of the block max, so those get 3e-4; everything shorter is held to grad_tol
            put(n, "W", dw, grad_tol if d.shape[0] * d.shape[2] * d.shape[3] < (1 << 18) else 3 * grad_tol)
            put(n, "b", db, floor=2e-6 * np.abs(d).sum(axis=(0, 2, 3), dtype=np.float64))
            src = v["in"][0]
            if vmap[src]["op"] == "zeropad":
                t, b, l, r = vmap[src]["pad"]
                dx = dx[:, :, t:dx.shape[2] - b, l:dx.shape[3] - r]
                src = vmap[src]["in"][0]
            give(src, dx)
        elif op in ("dense", "output"):
            a = lin_of(v)
            dwf, db, dx = t_(d).T.astype(np.float64) @ t_(a).astype(np.float64), d.sum(0, dtype=np.float64), t_(d) @ q(P(n, "W"))
            put(n, "W", dwf)
            put(n, "b", db, floor=2e-6 * np.abs(d).sum(0, dtype=np.float64))
            give(v["in"][0], dx)
        elif op == "bn":
            u = v["in"][0]
            xin = act_of(u)
            sm, ss = (t.cpu().numpy()[:xin.shape[1]] for t in net.aux[n])
            gam = P(n, "gamma")
            dd, ambiguous = d, None
            if v["act"] == "relu":
                # the GPU recomputes the gate as (x-mean)*gamma/std + beta > 0 in fp32; elements within round-off of zero may
                # fall either way: gate them as the GPU's dx says (dx == the ungated formula there or not), count them
                y64 = (xin.astype(np.float64) - sm[None, :, None, None]) * (gam / ss.astype(np.float64))[None, :, None, None] \
                    + P(n, "beta").astype(np.float64)[None, :, None, None]
                ambiguous = np.abs(y64) <= 1e-5 * np.abs(y64).max()
                rep.flip_candidates += int(ambiguous.sum())
                dd = np.where(y64 > 0, d, 0).astype(np.float32)
            dx, dg, dbt = o.bn_bwd(xin, dd, gam, sm, ss)
            put(n, "gamma", dg, grad_tol * 2)
            put(n, "beta", dbt, grad_tol * 2)
            assert not P(n, "mean", grads).any() and not P(n, "var", grads).any()  # BatchNormalization.java:205-208
            give(u, dx)
            if ambiguous is not None and ambiguous.any():
                assert ambiguous.mean() <= max_flip_frac, (n, float(ambiguous.mean()))
                amb.setdefault(u, np.zeros(xin.shape, bool))
                amb[u] |= ambiguous
        elif op == "pool" and n in taps:
            # its input was never materialised: route by the taps the GPU saved (validated in the forward pass above)
            give(v["in"][0], _scatter_taps(o, d, taps[n], v, act_of(v["in"][0]).shape))
        elif op == "pool":
            dx = o.pool_bwd(lin_of(v), d, v["kernel"], v["stride"], v["pad"], (1, 1), _MODE[v["mode"]], _POOL[v["type"]])
            give(v["in"][0], dx)
        elif op in ("add", "act"):
            for u in v["in"]:
                give(u, d)
        elif op == "flatten":
            give(v["in"][0], d.reshape(act_of(v["in"][0]).shape))
        else:
            raise NotImplementedError(op)

    # ---------------------------------------------------------------- gradient blocks (after the layer-wise
    # GradientNormalization, which the step applies in place before the updater: BaseMultiLayerUpdater.java:322-380)
    gn = OracleNet(model, net.layout, net.P, params_pre)
    gn.grads = gref.astype(np.float32)
    if model.get("grad_norm"):
        spans = {}
        for name, key, off, shp, pk in net.layout:
            lo, hi = spans.get(name, (off, off))
            spans[name] = (min(lo, off), max(hi, off + int(np.prod(shp))))
        for name, (lo, hi) in spans.items():  # the allowances scale with the block
            if model["grad_norm"] == "RenormalizeL2PerLayer":
                gfloor[lo:hi] /= max(float(np.sqrt((gref[lo:hi] ** 2).sum())), 1e-30)
        gn.grad_normalize()
    kind = {"W": "W", "b": "b", "gamma": "bn_gamma", "beta": "bn_beta"}
    for name, key, off, shp, pk in net.layout:
        if pk == "bn_stat":
            continue
        cnt = int(np.prod(shp))
        op = vmap[name]["op"]
        what = "grad." + ("conv_" if op == "conv" else "dense_" if op in ("dense", "output") else "") + kind[key]
        fl = gfloor[off:off + cnt]
        rep.check(what, name, grads[off:off + cnt], gn.grads[off:off + cnt], 0, float(gtol[off]), floor=fl if fl.any() else None)

    # ---------------------------------------------------------------- updater and score, block by block
    ref = OracleNet(dict(model, grad_norm=None), net.layout, net.P, params_pre)
    ref.grads[:] = grads
    ref.state[:] = state_pre
    ref.loss = loss_gpu
    s_ref = ref.score(B)
    assert abs(score - s_ref) <= 1e-5 * max(1.0, abs(s_ref)), (score, s_ref)
    ref.update(B)
    for name, key, off, shp, pk in net.layout:
        if pk == "bn_stat":
            continue  # running statistics were checked in the forward pass
        cnt = int(np.prod(shp))
        rep.check("update.params", f"{name}.{key}", params_post[off:off + cnt], ref.params[off:off + cnt], 0, param_tol)
    rep.check("update.state", "state", net.state.cpu().numpy(), ref.state, 0, param_tol)
    return rep
