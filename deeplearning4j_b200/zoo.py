"""The BASELINE configs as plain-data graphs (layer lists restated from deeplearning4j-zoo).

A model is a dict: {"name", "input": (C,H,W), "vertices": [...], "updater": {...}, "bias_updater", "l1", "l2",
"minibatch", "grad_norm"}; a vertex is {"name", "op", "in": [vertex names], ...op fields}. Ops:
  conv  (nOut, kernel, stride, pad, mode, act, dilation)   ZOO ConvolutionLayer
  bn    (decay, eps, act)  -- act is the FOLLOWING ActivationLayer folded in (fusion; same arithmetic)
  pool  (type, kernel, stride, pad, mode)
  act   (act)     add (act: following ActivationLayer folded in)     zeropad (pad t,b,l,r)
  flatten   CnnToFeedForwardPreProcessor ([B,C,H,W] 'c' -> [B, C*H*W])
  dense (nOut, act)       output (nOut; softmax + NLL/MCXENT)
Parameter order in the flat vector follows vertex order; per layer: conv = bias | W[c_out,c_in,kH,kW] 'c'
(ConvolutionParamInitializer.java:116-123,144-155), bn = gamma | beta | mean | var
(BatchNormalizationParamInitializer.java:80-113), dense/output = W[nIn,nOut] 'f' | bias (DefaultParamInitializer).
"""
import numpy as np

RELU, IDENTITY = "relu", "identity"


def _conv(name, inp, nOut, k, s=(1, 1), p=(0, 0), mode="Truncate", act=IDENTITY, d=(1, 1)):
    return {"name": name, "op": "conv", "in": [inp], "nOut": nOut, "kernel": tuple(k), "stride": tuple(s),
            "pad": tuple(p), "mode": mode, "act": act, "dilation": tuple(d)}


def _bn(name, inp, act=IDENTITY):
    # BatchNormalization() defaults: decay 0.9, eps 1e-5 (nn/conf/layers/BatchNormalization.java:34-35)
    return {"name": name, "op": "bn", "in": [inp], "decay": 0.9, "eps": 1e-5, "act": act}


def _pool(name, inp, ptype, k, s, p=(0, 0), mode="Truncate"):
    return {"name": name, "op": "pool", "in": [inp], "type": ptype, "kernel": tuple(k), "stride": tuple(s),
            "pad": tuple(p), "mode": mode}


def lenet(num_labels=10, input_shape=(1, 28, 28)):
    """ZOO/model/LeNet.java:84-110. SAME mode, XAVIER, AdaDelta (rho .95, eps 1e-6), MCXENT."""
    v = [{"name": "input", "op": "input", "in": []},
         _conv("cnn1", "input", 20, (5, 5), mode="Same", act=RELU),
         _pool("maxpool1", "cnn1", "MAX", (2, 2), (2, 2), mode="Same"),
         _conv("cnn2", "maxpool1", 50, (5, 5), mode="Same", act=RELU),
         _pool("maxpool2", "cnn2", "MAX", (2, 2), (2, 2), mode="Same"),
         {"name": "flat", "op": "flatten", "in": ["maxpool2"]},
         {"name": "ffn1", "op": "dense", "in": ["flat"], "nOut": 500, "act": RELU},
         {"name": "output", "op": "output", "in": ["ffn1"], "nOut": num_labels}]
    return {"name": "lenet", "input": tuple(input_shape), "vertices": v, "weight_init": ("xavier",),
            "updater": {"kind": "adadelta", "rho": 0.95, "eps": 1e-6}, "l1": 0.0, "l2": 0.0, "minibatch": True,
            "grad_norm": None, "seed": 123}


def vgg16(num_labels=1000, input_shape=(3, 224, 224)):
    """ZOO/model/VGG16.java:92-150: 13 conv3x3 p1 + ReLU, 5 max2/2, FC layers commented out in the reference:
    softmax output straight from 25088. Nesterovs defaults (lr 0.1, momentum 0.9), XAVIER, Truncate."""
    v = [{"name": "input", "op": "input", "in": []}]
    prev, i = "input", 0
    for block, (n, c) in enumerate([(2, 64), (2, 128), (3, 256), (3, 512), (3, 512)]):
        for _ in range(n):
            v.append(_conv(f"conv{i}", prev, c, (3, 3), (1, 1), (1, 1), act=RELU))
            prev, i = f"conv{i}", i + 1
        v.append(_pool(f"pool{block}", prev, "MAX", (2, 2), (2, 2)))
        prev = f"pool{block}"
    v.append({"name": "flat", "op": "flatten", "in": [prev]})
    v.append({"name": "output", "op": "output", "in": ["flat"], "nOut": num_labels})
    return {"name": "vgg16", "input": tuple(input_shape), "vertices": v, "weight_init": ("xavier",),
            "updater": {"kind": "nesterovs", "lr": 0.1, "momentum": 0.9}, "l1": 0.0, "l2": 0.0, "minibatch": True,
            "grad_norm": None, "seed": 123}


def alexnet(num_labels=1000, input_shape=(3, 224, 224), dropout=None):
    """ZOO/model/AlexNet.java:84-130. The zoo config's global dropOut(0.5) puts inverted dropout (retain 0.5) on the INPUT
    of every layer, conv1 and the pooling layers included (AbstractLayer.java:360-374); pass dropout=0.5 to get it --
    the masks are then inputs of the step (Network.set_dropout_masks / OracleNet.masks), because nd4j's DropOutInverted
    random stream is not available outside nd4j. Default: off."""
    v = [{"name": "input", "op": "input", "in": []},
         _conv("cnn1", "input", 64, (11, 11), (4, 4), (2, 2), act=RELU),
         _pool("maxpool1", "cnn1", "MAX", (3, 3), (2, 2), (1, 1)),
         _conv("cnn2", "maxpool1", 192, (5, 5), (2, 2), (2, 2), act=RELU),
         _pool("maxpool2", "cnn2", "MAX", (3, 3), (2, 2), mode="Same"),
         _conv("cnn3", "maxpool2", 384, (3, 3), (1, 1), (1, 1), mode="Same", act=RELU),
         _conv("cnn4", "cnn3", 256, (3, 3), (1, 1), (1, 1), mode="Same", act=RELU),
         _conv("cnn5", "cnn4", 256, (3, 3), (1, 1), (1, 1), mode="Same", act=RELU),
         _pool("maxpool3", "cnn5", "MAX", (3, 3), (7, 7), mode="Same"),
         {"name": "flat", "op": "flatten", "in": ["maxpool3"]},
         {"name": "ffn1", "op": "dense", "in": ["flat"], "nOut": 4096, "act": RELU},
         {"name": "ffn2", "op": "dense", "in": ["ffn1"], "nOut": 4096, "act": RELU},
         {"name": "output", "op": "output", "in": ["ffn2"], "nOut": num_labels}]
    if dropout:
        for vv in v[1:]:
            if vv["op"] in ("conv", "pool", "dense", "output"):
                vv["dropout"] = float(dropout)
    return {"name": "alexnet", "input": tuple(input_shape), "vertices": v, "weight_init": ("normal", 0.01),
            "updater": {"kind": "nesterovs", "lr": 1e-2, "momentum": 0.9},
            "bias_updater": {"kind": "nesterovs", "lr": 2e-2, "momentum": 0.9},
            "l1": 0.0, "l2": 5e-4, "minibatch": False, "grad_norm": "RenormalizeL2PerLayer", "seed": 123}


def simplecnn(num_labels=1000, input_shape=(3, 48, 48), dropout=None):
    """ZOO/model/SimpleCNN.java:67-131 (zoo default input 3x48x48): SAME mode, WeightInit.RELU, four blocks of
    [conv k, BN, conv k, BN, ReLU, avg 2x2/2, DropoutLayer(0.5)] with k = 7, 5, 3, 3 and 16, 32, 64, 128 maps, then conv3 256,
    BN, conv3 -> numLabels and a GlobalPoolingLayer(AVG) (= average pooling over the whole remaining map). The reference
    config has no loss layer, so fit() cannot train it (MultiLayerNetwork.java:1405-1408): a forward-only graph, ending at
    the pooled class scores that the final softmax ActivationLayer would normalise. A DropoutLayer is dropout on the next
    layer's input: dropout=0.5 puts it there (mask-fed, see alexnet); default off."""
    v = [{"name": "input", "op": "input", "in": []}]
    prev = "input"
    hw = list(input_shape[1:])
    for blk, (k, c) in enumerate([(7, 16), (5, 32), (3, 64), (3, 128)], 1):
        for j in (1, 2):
            v.append(_conv(f"b{blk}_conv{j}", prev, c, (k, k), mode="Same"))
            if dropout and blk > 1 and j == 1:
                v[-1]["dropout"] = float(dropout)
            v.append(_bn(f"b{blk}_bn{j}", f"b{blk}_conv{j}", RELU if j == 2 else IDENTITY))  # ActivationLayer(ReLU) folded
            prev = f"b{blk}_bn{j}"
        v.append(_pool(f"b{blk}_pool", prev, "AVG", (2, 2), (2, 2), mode="Same"))
        prev = f"b{blk}_pool"
        hw = [(hw[0] + 1) // 2, (hw[1] + 1) // 2]
    v.append(_conv("b5_conv1", prev, 256, (3, 3), mode="Same"))
    if dropout:
        v[-1]["dropout"] = float(dropout)
    v.append(_bn("b5_bn1", "b5_conv1"))
    v.append(_conv("b5_conv2", "b5_bn1", num_labels, (3, 3), mode="Same"))
    v.append(_pool("global_avg", "b5_conv2", "AVG", tuple(hw), (1, 1)))
    v.append({"name": "scores", "op": "flatten", "in": ["global_avg"]})
    return {"name": "simplecnn", "input": tuple(input_shape), "vertices": v, "weight_init": ("relu",),
            "updater": {"kind": "adadelta", "rho": 0.95, "eps": 1e-6}, "l1": 0.0, "l2": 0.0, "minibatch": True,
            "grad_norm": None, "seed": 123}


def resnet50(num_labels=1000, input_shape=(3, 224, 224)):
    """ZOO/model/ResNet50.java:89-225 (DL4J's variant: stage-2 conv block has stride 2 so stages run at
    28/14/7/4 px; the 'avgpool' is a MAX 3x3 s2). ActivationLayers following BN / Add are folded into them.
    RmsProp(0.1, 0.96, 1e-3), N(0, 0.5) init, l1 1e-7, l2 5e-5, Truncate; 3x3 convs SAME."""
    v = [{"name": "input", "op": "input", "in": []},
         {"name": "stem-zero", "op": "zeropad", "in": ["input"], "pad": (3, 3, 3, 3)},
         _conv("stem-cnn1", "stem-zero", 64, (7, 7), (2, 2)),
         _bn("stem-batch1", "stem-cnn1", RELU),
         _pool("stem-maxpool1", "stem-batch1", "MAX", (3, 3), (2, 2))]

    def identity_block(stage, block, f, inp):  # :89-123
        cn, bn, sh = f"res{stage}{block}_branch", f"bn{stage}{block}_branch", f"short{stage}{block}_branch"
        v.extend([_conv(cn + "2a", inp, f[0], (1, 1)), _bn(bn + "2a", cn + "2a", RELU),
                  _conv(cn + "2b", bn + "2a", f[1], (3, 3), mode="Same"), _bn(bn + "2b", cn + "2b", RELU),
                  _conv(cn + "2c", bn + "2b", f[2], (1, 1)), _bn(bn + "2c", cn + "2c"),
                  {"name": sh, "op": "add", "in": [bn + "2c", inp], "act": RELU}])
        return sh

    def conv_block(stage, block, f, inp, stride=(2, 2)):  # :130-169
        cn, bn, sh = f"res{stage}{block}_branch", f"bn{stage}{block}_branch", f"short{stage}{block}_branch"
        v.extend([_conv(cn + "2a", inp, f[0], (1, 1), stride), _bn(bn + "2a", cn + "2a", RELU),
                  _conv(cn + "2b", bn + "2a", f[1], (3, 3), mode="Same"), _bn(bn + "2b", cn + "2b", RELU),
                  _conv(cn + "2c", bn + "2b", f[2], (1, 1)), _bn(bn + "2c", cn + "2c"),
                  _conv(cn + "1", inp, f[2], (1, 1), stride), _bn(bn + "1", cn + "1"),
                  {"name": sh, "op": "add", "in": [bn + "2c", bn + "1"], "act": RELU}])
        return sh

    x = conv_block(2, "a", (64, 64, 256), "stem-maxpool1", (2, 2))  # :194 stride 2, unlike the paper
    for b in "bc":
        x = identity_block(2, b, (64, 64, 256), x)
    x = conv_block(3, "a", (128, 128, 512), x)
    for b in "bcd":
        x = identity_block(3, b, (128, 128, 512), x)
    x = conv_block(4, "a", (256, 256, 1024), x)
    for b in "bcdef":
        x = identity_block(4, b, (256, 256, 1024), x)
    x = conv_block(5, "a", (512, 512, 2048), x)
    for b in "bc":
        x = identity_block(5, b, (512, 512, 2048), x)
    v.append(_pool("avgpool", x, "MAX", (3, 3), (2, 2)))  # :214-216: default PoolingType is MAX
    v.append({"name": "flat", "op": "flatten", "in": ["avgpool"]})
    v.append({"name": "output", "op": "output", "in": ["flat"], "nOut": num_labels})
    return {"name": "resnet50", "input": tuple(input_shape), "vertices": v, "weight_init": ("normal", 0.5),
            "updater": {"kind": "rmsprop", "lr": 0.1, "decay": 0.96, "eps": 1e-3}, "l1": 1e-7, "l2": 5e-5,
            "minibatch": True, "grad_norm": None, "seed": 123}


MODELS = {"lenet": lenet, "vgg16": vgg16, "alexnet": alexnet, "simplecnn": simplecnn, "resnet50": resnet50}


# ---------------------------------------------------------------------------------------------------
def _out_hw(h, w, k, s, p, d, mode, out_size=None):
    if out_size is None:
        from . import _lib
        out_size = _lib.out_size
    m = {"Strict": 0, "Truncate": 1, "Same": 2}[mode]
    return out_size(h, k[0], s[0], p[0], d[0], m)[0], out_size(w, k[1], s[1], p[1], d[1], m)[0]


def infer_shapes(model, out_size=None):
    """-> {vertex name: (C,H,W) or (n,)}; ConvolutionUtils.getOutputSize rules via the C ABI's out_size (callers that
    must not load the CUDA library -- the CPU reference arm of bench.py -- pass their own `out_size(in,k,s,p,d,mode)`)."""
    shapes = {}
    for v in model["vertices"]:
        op = v["op"]
        if op == "input":
            shapes[v["name"]] = tuple(model["input"])
            continue
        s = shapes[v["in"][0]]
        if op == "conv":
            h, w = _out_hw(s[1], s[2], v["kernel"], v["stride"], v["pad"], v["dilation"], v["mode"], out_size)
            shapes[v["name"]] = (v["nOut"], h, w)
        elif op == "pool":
            h, w = _out_hw(s[1], s[2], v["kernel"], v["stride"], v["pad"], (1, 1), v["mode"], out_size)
            shapes[v["name"]] = (s[0], h, w)
        elif op == "zeropad":
            t, b, l, r = v["pad"]
            shapes[v["name"]] = (s[0], s[1] + t + b, s[2] + l + r)
        elif op == "flatten":
            shapes[v["name"]] = (int(np.prod(s)),)
        elif op in ("dense", "output"):
            shapes[v["name"]] = (v["nOut"],)
        else:  # bn, act, add
            shapes[v["name"]] = s
    return shapes


def param_layout(model, out_size=None):
    """-> (total P, [(vertex name, key, offset, shape, kind)]) in flat-vector order (MultiLayerNetwork.java:561-676).
    kind: 'w' (regularised weight), 'b' (bias), 'bn_gb' (gamma/beta), 'bn_stat' (running mean/var, NoOp updater)."""
    shapes = infer_shapes(model, out_size)
    out, off = [], 0
    for v in model["vertices"]:
        op = v["op"]
        if op == "conv":
            cin = shapes[v["in"][0]][0]
            entries = [("b", (v["nOut"],), "b"), ("W", (v["nOut"], cin) + v["kernel"], "w")]
        elif op == "bn":
            c = shapes[v["name"]][0]
            entries = [("gamma", (c,), "bn_gb"), ("beta", (c,), "bn_gb"), ("mean", (c,), "bn_stat"), ("var", (c,), "bn_stat")]
        elif op in ("dense", "output"):
            nin = shapes[v["in"][0]][0]
            entries = [("W", (v["nOut"], nin), "w"), ("b", (v["nOut"],), "b")]  # W stored [nOut][nIn] == 'f' [nIn,nOut]
        else:
            continue
        for key, shp, kind in entries:
            out.append((v["name"], key, off, shp, kind))
            off += int(np.prod(shp))
    return off, out


def init_params(model, seed=None, out_size=None):
    """Host-generated initial flat parameter vector (WeightInitUtil.java:80-150 formulas with our own seeded
    generator -- nd4j's RNG stream is unavailable). conv fanIn = Cin*kH*kW,
    fanOut = Cout*kH*kW/(sH*sW) (ConvolutionParamInitializer.java:188-189)."""
    rng = np.random.default_rng(model.get("seed", 123) if seed is None else seed)
    P, layout = param_layout(model, out_size)
    flat = np.zeros(P, np.float32)
    vmap = {v["name"]: v for v in model["vertices"]}
    wi = model["weight_init"]
    for name, key, off, shp, kind in layout:
        n = int(np.prod(shp))
        v = vmap[name]
        if kind == "w":
            if v["op"] == "conv":
                fan_in = shp[1] * shp[2] * shp[3]
                fan_out = shp[0] * shp[2] * shp[3] / (v["stride"][0] * v["stride"][1])
            else:
                fan_in, fan_out = shp[1], shp[0]
            if wi[0] == "xavier":
                vals = rng.standard_normal(n) * np.sqrt(2.0 / (fan_in + fan_out))
            elif wi[0] == "relu":  # WeightInit.RELU: N(0, 2 / fanIn) (WeightInitUtil.java)
                vals = rng.standard_normal(n) * np.sqrt(2.0 / fan_in)
            else:
                vals = rng.standard_normal(n) * wi[1]
            flat[off:off + n] = vals.astype(np.float32)
        elif key in ("gamma", "var"):
            flat[off:off + n] = 1.0
    return flat
