"""-m gpu: network-level parity of the arithmetic bench.py times (TF32 tensor cores, TF32-representable activations)
at BASELINE.json's image sizes -- every vertex of one training step checked against the CPU oracle on the GPU's own
inputs (tests/step_replay.py explains why this, and not a looser free-running comparison, is the meaningful bar).

Bounds asserted per element: forward activations and deltas |gpu - oracle| <= 2^-11 |value| (the one TF32 rounding a
producer applies on store) + 1e-4 max|oracle|; every gradient block (dW, db, dgamma, dbeta) <= 1e-4 of the block max
(2e-4 for the BN parameter gradients, 3e-4 for filter gradients reduced over more than 2^18 pixels -- the tensor core's
fp32 accumulator truncates), no rounding allowance; every post-update parameter block and the updater state
<= 1e-5 of the block max; score and loss <= 1e-5. The reference's own helper-vs-builtin bound is 1e-4
(deeplearning4j-cuda/src/test/java/org/deeplearning4j/convolution/TestConvolution.java:145,151)."""
import numpy as np
import pytest

from step_replay import replay_step
from test_gpu_network import _data, _tiny_resnet

pytestmark = pytest.mark.gpu


def _masks(model, shapes, B, seed=99):
    rng = np.random.default_rng(seed)
    m = {}
    for v in model["vertices"]:
        if v.get("dropout"):
            m[v["name"]] = (rng.random((B,) + tuple(shapes[v["in"][0]])) < v["dropout"]).astype(np.float32)
    return m or None


def _replay(model, B, K, algo_name, steps=1, forward_only=False, **kw):
    from deeplearning4j_b200 import _lib, network, zoo
    algo = _lib.ALGO_FP32 if algo_name == "fp32" else _lib.ALGO_TC_TF32
    x, lab = _data(model, B, K)
    net = network.Network(model, B, algo=algo, params=zoo.init_params(model))
    try:
        masks = _masks(model, net.shapes, B)
        for _ in range(steps):  # later steps start from the GPU's own updated parameters and updater state
            rep = replay_step(net, model, x, lab, masks=masks, forward_only=forward_only, **kw)
        print(model["name"], B, algo_name, rep.summary())
        return rep
    finally:
        net.close()


@pytest.mark.parametrize("algo_name", ["fp32", "tf32"])
def test_replay_tiny_resnet(algo_name):
    _replay(_tiny_resnet(), 8, 10, algo_name, steps=2)


@pytest.mark.parametrize("algo_name", ["fp32", "tf32"])
def test_replay_lenet_baseline_batch(algo_name):
    from deeplearning4j_b200 import zoo
    _replay(zoo.lenet(), 64, 10, algo_name, steps=2)


def test_replay_sigmoid_tanh_activations():
    """Convolution / dense layers with sigmoid and tanh: the activation derivative is taken from the stored output."""
    from deeplearning4j_b200 import zoo
    m = zoo.lenet()
    for v in m["vertices"]:
        if v["name"] == "cnn1":
            v["act"] = "tanh"
        if v["name"] in ("cnn2", "ffn1"):
            v["act"] = "sigmoid"
    for algo in ("fp32", "tf32"):
        _replay(m, 16, 10, algo, steps=1)


def test_replay_vgg16_224_tf32():
    from deeplearning4j_b200 import zoo
    _replay(zoo.vgg16(), 4, 1000, "tf32")


def test_replay_alexnet_224_tf32_with_dropout():
    from deeplearning4j_b200 import zoo
    _replay(zoo.alexnet(dropout=0.5), 8, 1000, "tf32", steps=2)


def test_replay_resnet50_224_batch32_tf32():
    from deeplearning4j_b200 import zoo
    _replay(zoo.resnet50(), 32, 1000, "tf32")


def test_replay_simplecnn_224_forward_tf32():
    from deeplearning4j_b200 import zoo
    _replay(zoo.simplecnn(input_shape=(3, 224, 224)), 4, 1000, "tf32", forward_only=True)
