"""CPU: numerical gradient checks of the oracle's backward passes -- the reference's own way of pinning backprop
(deeplearning4j-core/src/test/java/org/deeplearning4j/gradientcheck/CNNGradientCheckTest.java, BNGradientCheckTest.java,
LRNGradientCheckTests.java, all through GradientCheckUtil: central differences against the analytic gradient, max relative
error 1e-3). The reference holds no literal expected values for gradients, so this is what stands behind every `oracle.*_bwd`
the GPU parity tests compare with: for a scalar loss L = sum(forward(x, params) * r) with a fixed random r, the analytic
gradient is backward(r); it is compared with (L(p + h) - L(p - h)) / 2h on a random sample of coordinates of every input.
The oracle computes in float32, so h is 1e-2 on O(1) values and L is accumulated in float64: the round-off of one forward is
~1e-7 * sum|z r| / h, the truncation error h^2/6 * f''' -- both far below the 2e-3 allowed here."""
import numpy as np
import pytest

H_STEP = 1e-2
MAX_REL = 2e-3
SAMPLES = 12


def _loss(z, r):
    return float(np.sum(z.astype(np.float64) * r.astype(np.float64)))


def _check(name, forward, params, analytic, r, rng, h=H_STEP, tol=MAX_REL):
    """forward(*params) -> z; analytic[i] = dL/dparams[i] (None: not checked)."""
    for i, (p, g) in enumerate(zip(params, analytic)):
        if g is None:
            continue
        assert g.shape == p.shape, (name, i, g.shape, p.shape)
        scale = max(float(np.abs(g).max()), 1e-6)
        flat = p.reshape(-1)
        for j in rng.choice(flat.size, size=min(SAMPLES, flat.size), replace=False):
            old = flat[j]
            flat[j] = old + h
            lp = _loss(forward(*params), r)
            flat[j] = old - h
            lm = _loss(forward(*params), r)
            flat[j] = old
            num = (lp - lm) / (2 * h)
            assert abs(num - float(g.reshape(-1)[j])) <= tol * scale, (name, "input", i, "coord", int(j), num, float(g.reshape(-1)[j]), scale)


@pytest.mark.parametrize("case", [((2, 3, 7, 8), 4, (3, 3), (1, 1), (1, 1), (1, 1), 1), ((2, 4, 9, 9), 5, (3, 3), (2, 2), (0, 0), (1, 1), 2),
                                  ((1, 2, 9, 9), 3, (3, 3), (1, 1), (0, 0), (2, 2), 2), ((2, 3, 8, 8), 6, (1, 1), (2, 2), (0, 0), (1, 1), 1),
                                  ((2, 2, 11, 10), 3, (5, 3), (2, 1), (2, 1), (1, 1), 1)])
def test_convolution_backward(oracle, case):
    shape, cout, k, s, p, d, mode = case
    rng = np.random.default_rng(1)
    x = rng.standard_normal(shape).astype(np.float32)
    w = (rng.standard_normal((cout, shape[1]) + k) * 0.5).astype(np.float32)
    b = rng.standard_normal(cout).astype(np.float32)
    fwd = lambda x_, w_, b_: oracle.conv_fwd(x_, w_, b_, stride=s, pad=p, dilation=d, mode=mode)
    r = rng.standard_normal(fwd(x, w, b).shape).astype(np.float32)
    dw, db, dx = oracle.conv_bwd(x, w, r, stride=s, pad=p, dilation=d, mode=mode)
    _check("conv", fwd, [x, w, b], [dx, dw, db], r, rng)
    dw2, db2, dx2 = oracle.conv_bwd_im2col(x, w, r, stride=s, pad=p, dilation=d, mode=mode)   # the reference's im2col + gemm pipeline
    np.testing.assert_allclose(dw2, dw, rtol=0, atol=1e-4 * np.abs(dw).max())
    np.testing.assert_allclose(dx2, dx, rtol=0, atol=1e-4 * np.abs(dx).max())


@pytest.mark.parametrize("ptype", ["MAX", "AVG", "PNORM"])
@pytest.mark.parametrize("case", [((2, 3, 8, 8), (2, 2), (2, 2), (0, 0), 1), ((2, 2, 9, 11), (3, 3), (2, 2), (1, 1), 1), ((1, 3, 7, 7), (3, 2), (1, 2), (0, 0), 2)])
def test_subsampling_backward(oracle, case, ptype):
    shape, k, s, p, mode = case
    rng = np.random.default_rng(2)
    # distinct values with gaps larger than 2h: a max-pooling window must keep its winner under the perturbation
    x = (rng.permutation(int(np.prod(shape))).reshape(shape) * 0.05 - 2.0).astype(np.float32)
    if ptype == "PNORM":
        x = np.abs(x) + 0.5  # |x|^p is not differentiable at 0
    pt = {"MAX": oracle.POOL_MAX, "AVG": oracle.POOL_AVG, "PNORM": oracle.POOL_PNORM}[ptype]
    fwd = lambda x_: oracle.pool_fwd(x_, k, s, p, (1, 1), mode, pt, 2)
    r = rng.standard_normal(fwd(x).shape).astype(np.float32)
    dx = oracle.pool_bwd(x, r, k, s, p, (1, 1), mode, pt, 2)
    _check("pool-" + ptype, fwd, [x], [dx], r, rng)


@pytest.mark.parametrize("shape", [(4, 3, 5, 6), (6, 8, 3, 3), (16, 5, 1, 1)])
def test_batch_normalization_backward(oracle, shape):
    rng = np.random.default_rng(3)
    x = (rng.standard_normal(shape) * 1.5 + 0.3).astype(np.float32)
    gamma = (1 + 0.2 * rng.standard_normal(shape[1])).astype(np.float32)
    beta = (0.2 * rng.standard_normal(shape[1])).astype(np.float32)

    def fwd(x_, g_, b_):
        return oracle.bn_fwd(x_, g_, b_, np.zeros(shape[1], np.float32), np.ones(shape[1], np.float32), 0.9, 1e-5, True)[0]
    _, mean, std = oracle.bn_fwd(x, gamma, beta, np.zeros(shape[1], np.float32), np.ones(shape[1], np.float32), 0.9, 1e-5, True)
    r = rng.standard_normal(shape).astype(np.float32)
    dx, dg, db = oracle.bn_bwd(x, r, gamma, mean, std)
    # (the batch statistics move with x: the analytic dx includes d mean / d var, BatchNormalization.java:172-196)
    _check("bn", fwd, [x, gamma, beta], [dx, dg, db], r, rng, tol=4e-3)


def test_local_response_normalization_backward(oracle):
    rng = np.random.default_rng(4)
    x = (rng.standard_normal((2, 7, 4, 5)) * 2.0).astype(np.float32)
    kw = dict(k=2.0, n=5.0, alpha=1e-2, beta=0.75)   # a large alpha so that the normaliser actually depends on x
    fwd = lambda x_: oracle.lrn_fwd(x_, **kw)
    r = rng.standard_normal(x.shape).astype(np.float32)
    _check("lrn", fwd, [x], [oracle.lrn_bwd(x, r, **kw)], r, rng)


@pytest.mark.parametrize("kind", ["relu", "sigmoid", "tanh"])
def test_activation_backward(oracle, kind):
    rng = np.random.default_rng(5)
    z = rng.standard_normal((3, 17)).astype(np.float32)
    z[np.abs(z) < 0.05] += 0.2   # keep ReLU away from its kink
    code = {"relu": oracle.ACT_RELU, "sigmoid": oracle.ACT_SIGMOID, "tanh": oracle.ACT_TANH}[kind]
    fwd = lambda z_: oracle.act_fwd(z_, code)
    r = rng.standard_normal(z.shape).astype(np.float32)
    _check("act-" + kind, fwd, [z], [oracle.act_bwd(z, r, code)], r, rng)


def test_softmax_negative_log_likelihood_delta(oracle):
    """delta = dL/dz of the summed loss (BaseOutputLayer: p - labels for softmax + MCXENT)."""
    rng = np.random.default_rng(6)
    z = rng.standard_normal((5, 7)).astype(np.float32)
    lab = np.zeros((5, 7), np.float32)
    lab[np.arange(5), rng.integers(0, 7, 5)] = 1.0
    _, _, delta = oracle.softmax_nll(z, lab)
    flat = z.reshape(-1)
    for j in rng.choice(flat.size, size=SAMPLES, replace=False):
        old = flat[j]
        flat[j] = old + H_STEP
        lp = oracle.softmax_nll(z, lab)[0]
        flat[j] = old - H_STEP
        lm = oracle.softmax_nll(z, lab)[0]
        flat[j] = old
        assert abs((lp - lm) / (2 * H_STEP) - float(delta.reshape(-1)[j])) <= MAX_REL, (int(j),)


@pytest.mark.parametrize("case", [((2, 3, 5, 6), 4, (3, 3), (2, 2), (0, 0)), ((2, 4, 4, 4), 3, (2, 2), (2, 2), (0, 0)), ((1, 2, 6, 5), 3, (3, 3), (1, 1), (1, 1))])
def test_deconvolution2d_backward(oracle, case):
    shape, nout, k, s, p = case
    rng = np.random.default_rng(7)
    x = rng.standard_normal(shape).astype(np.float32)
    w = (rng.standard_normal((shape[1], nout) + k) * 0.5).astype(np.float32)
    b = rng.standard_normal(nout).astype(np.float32)
    fwd = lambda x_, w_, b_: oracle.deconv2d_fwd(x_, w_, b_, stride=s, pad=p)
    r = rng.standard_normal(fwd(x, w, b).shape).astype(np.float32)
    dw, db, dx = oracle.deconv2d_bwd(x, w, r, stride=s, pad=p)
    _check("deconv2d", fwd, [x, w, b], [dx, dw, db], r, rng)


@pytest.mark.parametrize("case", [((2, 3, 7, 7), 2, 5, (3, 3), (1, 1), (1, 1), 1), ((2, 4, 8, 9), 1, 6, (3, 3), (2, 2), (0, 0), 2)])
def test_separable_convolution2d_backward(oracle, case):
    shape, mult, nout, k, s, p, mode = case
    rng = np.random.default_rng(8)
    x = rng.standard_normal(shape).astype(np.float32)
    wd = (rng.standard_normal((mult, shape[1]) + k) * 0.5).astype(np.float32)
    wp = (rng.standard_normal((nout, shape[1] * mult, 1, 1)) * 0.5).astype(np.float32)
    b = rng.standard_normal(nout).astype(np.float32)
    fwd = lambda x_, wd_, wp_, b_: oracle.sepconv2d_fwd(x_, wd_, wp_, b_, stride=s, pad=p, mode=mode)[0]
    r = rng.standard_normal(fwd(x, wd, wp, b).shape).astype(np.float32)
    dwd, dwp, db, dx = oracle.sepconv2d_bwd(x, wd, wp, r, stride=s, pad=p, mode=mode)
    _check("sepconv2d", fwd, [x, wd, wp, b], [dx, dwd, dwp, db], r, rng)


def test_whole_network_gradient(oracle):
    """The network-level oracle (oracle/net_oracle.py: ComputationGraph forward / calcBackpropGradients over a residual graph
    with BatchNormalization, MAX pooling, strided 1x1 and SAME 3x3 convolutions, Add fan-in, softmax output) against central
    differences of its own summed loss, parameter by parameter -- the whole-network form of the reference's gradient checks.
    The flat gradient is d(sum of per-example losses)/d(params) before the updater divides by the minibatch size; the
    running-statistics slots of BatchNormalization carry no gradient (BatchNormalization.java:205-208)."""
    from deeplearning4j_b200 import zoo
    from oracle.net_oracle import OracleNet
    from test_gpu_network import _tiny_resnet
    m = _tiny_resnet()
    B = 6
    P, layout = zoo.param_layout(m)
    p0 = zoo.init_params(m)
    rng = np.random.default_rng(9)
    x = rng.random((B,) + tuple(m["input"]), dtype=np.float32)
    lab = np.zeros((B, 10), np.float32)
    lab[np.arange(B), rng.integers(0, 10, B)] = 1.0
    net = OracleNet(m, layout, P, p0)
    net.forward(x, lab, True)
    net.backward()
    g = net.grads.astype(np.float64).copy()

    def loss_at(params):
        n2 = OracleNet(m, layout, P, params)
        n2.forward(x, lab, True)
        return float(n2.loss)

    # ReLU and max-pooling make the loss piecewise smooth: a step that carries some activation across a kink leaves an O(h)
    # error, float32 round-off of the loss (~1e-6) an O(1/h) one. Each coordinate is therefore tried with a shrinking step
    # (the reference's checks avoid the issue by using smooth activations and double precision, which this float32
    # restatement does not offer) and must match for one of them; a coordinate that sits within 1e-4 of a kink cannot,
    # so up to 3 % of the sample may miss.
    checked, missed = 0, []
    for name, key, off, shp, pk in layout:
        cnt = int(np.prod(shp))
        if pk == "bn_stat":
            assert not g[off:off + cnt].any(), (name, key)
            continue
        scale = max(float(np.abs(g[off:off + cnt]).max()), 1e-3)
        for j in rng.choice(cnt, size=min(3, cnt), replace=False):
            best = np.inf
            for h in (2e-3, 1e-3, 5e-4, 2.5e-4):
                p = p0.copy()
                p[off + j] += h
                lp = loss_at(p)
                p[off + j] -= 2 * h
                lm = loss_at(p)
                best = min(best, abs((lp - lm) / (2 * h) - g[off + j]))
                if best <= 1e-2 * scale + 5e-3:
                    break
            checked += 1
            if best > 1e-2 * scale + 5e-3:
                missed.append((name, key, int(j), float(g[off + j]), float(best), scale))
    assert checked > 40
    assert len(missed) <= 0.03 * checked, missed


def test_whole_network_gradient_smooth_variant(oracle):
    """The same residual graph with tanh instead of ReLU and average instead of max pooling -- a smooth loss, which is how the
    reference's gradient checks are set up -- held to 2e-3 of each block's largest gradient with a single step size."""
    import copy
    from deeplearning4j_b200 import zoo
    from oracle.net_oracle import OracleNet
    from test_gpu_network import _tiny_resnet
    m = copy.deepcopy(_tiny_resnet())
    for v in m["vertices"]:
        if v.get("act") == "relu":
            v["act"] = "tanh"
        if v["op"] == "pool":
            v["type"] = "AVG"
    B = 6
    P, layout = zoo.param_layout(m)
    p0 = zoo.init_params(m)
    rng = np.random.default_rng(9)
    x = rng.random((B,) + tuple(m["input"]), dtype=np.float32)
    lab = np.zeros((B, 10), np.float32)
    lab[np.arange(B), rng.integers(0, 10, B)] = 1.0
    net = OracleNet(m, layout, P, p0)
    net.forward(x, lab, True)
    net.backward()
    g = net.grads.astype(np.float64).copy()

    def loss_at(params):
        n2 = OracleNet(m, layout, P, params)
        n2.forward(x, lab, True)
        return float(n2.loss)

    h, checked = 2e-3, 0
    for name, key, off, shp, pk in layout:
        cnt = int(np.prod(shp))
        if pk == "bn_stat":
            continue
        scale = max(float(np.abs(g[off:off + cnt]).max()), 1e-3)
        for j in rng.choice(cnt, size=min(3, cnt), replace=False):
            p = p0.copy()
            p[off + j] += h
            lp = loss_at(p)
            p[off + j] -= 2 * h
            lm = loss_at(p)
            assert abs((lp - lm) / (2 * h) - g[off + j]) <= 2e-3 * scale + 2e-3, (name, key, int(j), (lp - lm) / (2 * h), g[off + j], scale)
            checked += 1
    assert checked > 40


def _gradcheck_net(m, B, K, masks_seed=None, h=2e-3, tol=2e-3, per_block=3, only=None):
    from deeplearning4j_b200 import zoo
    from oracle.net_oracle import OracleNet
    P, layout = zoo.param_layout(m)
    p0 = zoo.init_params(m)
    rng = np.random.default_rng(21)
    x = rng.random((B,) + tuple(m["input"]), dtype=np.float32)
    lab = np.zeros((B, K), np.float32)
    lab[np.arange(B), rng.integers(0, K, B)] = 1.0
    masks = None
    if masks_seed is not None:  # inverted dropout with fixed masks (inputs of the step, as in the GPU tests)
        from oracle import oracle as o
        shapes = zoo.infer_shapes(m, o.out_size)
        mr = np.random.default_rng(masks_seed)
        masks = {v["name"]: (mr.random((B,) + tuple(shapes[v["in"][0]])) < v["dropout"]).astype(np.float32)
                 for v in m["vertices"] if v.get("dropout")}

    def make(params):
        n = OracleNet(m, layout, P, params)
        n.masks = masks
        n.forward(x, lab, True)
        return n
    net = make(p0)
    net.backward()
    g = net.grads.astype(np.float64).copy()
    # Directional derivatives along random +-1 directions of each parameter block: the signal g.s grows with the block size
    # while the float32 round-off of the loss (~1e-5 here) does not -- single coordinates of the early layers, whose
    # gradients are ~1e-2 behind two saturating activations and dropout, would drown in it.
    checked = 0
    for name, key, off, shp, pk in layout:
        cnt = int(np.prod(shp))
        if pk == "bn_stat" or (only is not None and name not in only):
            continue
        gb = g[off:off + cnt]
        scale = max(float(np.linalg.norm(gb)), 1e-3)
        for _ in range(per_block):
            sgn = rng.choice(np.array([-1.0, 1.0], np.float32), size=cnt)
            p = p0.copy()
            p[off:off + cnt] += np.float32(h) * sgn
            lp = float(make(p).loss)
            p = p0.copy()
            p[off:off + cnt] -= np.float32(h) * sgn
            lm = float(make(p).loss)
            num, ana = (lp - lm) / (2 * h), float(gb @ sgn.astype(np.float64))
            assert abs(num - ana) <= tol * scale + 1e-3, (name, key, num, ana, scale)
            checked += 1
    return checked


def test_sequential_net_with_dense_layers_and_dropout_gradient(oracle):
    """MultiLayerNetwork-style chain (conv -> pool -> conv -> pool -> flatten -> dense -> dense -> softmax) with smooth
    activations, SAME / strided windows, a CnnToFeedForward flatten and the zoo AlexNet's inverted dropout on every layer
    input (mask-fed)."""
    from deeplearning4j_b200 import zoo
    v = [{"name": "input", "op": "input", "in": []},
         zoo._conv("c1", "input", 6, (5, 5), (2, 2), (2, 2), act="tanh"),
         zoo._pool("p1", "c1", "AVG", (3, 3), (2, 2), (1, 1)),
         zoo._conv("c2", "p1", 8, (3, 3), mode="Same", act="sigmoid"),
         zoo._pool("p2", "c2", "PNORM", (2, 2), (2, 2)),
         {"name": "flat", "op": "flatten", "in": ["p2"]},
         {"name": "f1", "op": "dense", "in": ["flat"], "nOut": 12, "act": "tanh"},
         {"name": "f2", "op": "dense", "in": ["f1"], "nOut": 9, "act": "sigmoid"},
         {"name": "output", "op": "output", "in": ["f2"], "nOut": 5}]
    for vv in v[1:]:
        if vv["op"] in ("conv", "pool", "dense", "output"):
            vv["dropout"] = 0.5
    m = {"name": "seq", "input": (2, 17, 19), "vertices": v, "weight_init": ("normal", 0.4),
         "updater": {"kind": "nesterovs", "lr": 1e-2, "momentum": 0.9}, "l1": 0.0, "l2": 0.0, "minibatch": True, "grad_norm": None,
         "seed": 5}
    assert _gradcheck_net(m, 5, 5, masks_seed=None, h=2e-3, tol=1e-2) >= 20
    # With dropout the reference replaces a layer's input by x * mask / p (AbstractLayer.java:360-374) but has NO dropout
    # backprop step: the epsilon a layer hands down is not multiplied by the mask, so below the topmost dropout the
    # "gradient" is not the derivative of the loss -- by the reference's construction, which the oracle restates. Only the
    # output layer's blocks (computed from its dropped input and its own delta) are true derivatives and checked here.
    assert _gradcheck_net(m, 5, 5, masks_seed=77, h=2e-3, tol=1e-2, only=("output",)) >= 6
