"""Host logic without a GPU: the zoo graphs reproduce the reference configurations' parameter counts (SURVEY.md §8 a11:
the flat [1,P] vector is the unit of the updater and of the all-reduce) and shape chains (ConvolutionUtils size rules
through the C ABI's dl4jb200_out_size, which needs no device)."""
import numpy as np

from deeplearning4j_b200 import zoo


def test_parameter_counts_match_reference_models():
    want = {"resnet50": 25_636_712, "vgg16": 39_803_688, "alexnet": 24_400_680}
    for name, p in want.items():
        P, layout = zoo.param_layout(zoo.MODELS[name]())
        assert P == p, (name, P)
        offs = [off for _, _, off, _, _ in layout]
        assert offs == sorted(offs) and offs[0] == 0  # contiguous views in layer order
    P, _ = zoo.param_layout(zoo.lenet(num_labels=10, input_shape=(1, 28, 28)))
    assert P == 1_256_080


def test_flat_layout_order_inside_a_layer():
    # conv: bias first then W (ConvolutionParamInitializer.java:116-123); BN: gamma | beta | mean | var
    _, layout = zoo.param_layout(zoo.resnet50())
    by_layer = {}
    for name, key, off, shp, pk in layout:
        by_layer.setdefault(name, []).append((off, key))
    assert [k for _, k in sorted(by_layer["res2a_branch2a"])] == ["b", "W"]
    assert [k for _, k in sorted(by_layer["bn2a_branch2a"])] == ["gamma", "beta", "mean", "var"]


def test_shape_chains():
    s = zoo.infer_shapes(zoo.resnet50())
    assert s["stem-cnn1"] == (64, 112, 112) and s["stem-maxpool1"] == (64, 55, 55)  # max 3x3/2, no padding
    assert s["res2a_branch2a"][1:] == (28, 28) and s["res3a_branch2a"][1:] == (14, 14)
    assert s["res4a_branch2a"][1:] == (7, 7) and s["res5a_branch2a"][1:] == (4, 4)   # DL4J's variant: half-resolution stages
    a = zoo.infer_shapes(zoo.alexnet())
    assert a["cnn1"] == (64, 55, 55) and a["maxpool1"][1:] == (28, 28) and a["cnn2"][1:] == (14, 14)
    assert a["maxpool2"][1:] == (7, 7) and a["maxpool3"][1:] == (1, 1)               # k3 s7 SAME: negative padding
    c = zoo.infer_shapes(zoo.simplecnn(num_labels=7))
    assert c["b4_pool"] == (128, 3, 3) and c["global_avg"] == (7, 1, 1) and tuple(c["scores"]) == (7,)


def test_weight_init_scales():
    m = zoo.simplecnn(num_labels=7)
    P, layout = zoo.param_layout(m)
    p = zoo.init_params(m)
    for name, key, off, shp, pk in layout:
        if name == "b3_conv2" and key == "W":
            w = p[off:off + int(np.prod(shp))]
            assert abs(float(w.std()) / np.sqrt(2.0 / (shp[1] * shp[2] * shp[3])) - 1.0) < 0.05  # WeightInit.RELU: N(0, 2/fanIn)
