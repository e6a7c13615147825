"""Extracts the literal known-answer vectors of the reference's own unit tests into tests/golden/*.json.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
The JSON files are committed; tests never read /root/reference.
"""
import json
import os
import re

REF = "/root/reference/deeplearning4j-core/src/test/java/org/deeplearning4j"
OUT = os.path.dirname(os.path.abspath(__file__))


def _doubles_after(src, anchor, nth=0):
    """The nth `new double[] {...}` literal after the first occurrence of `anchor`."""
    pos = src.index(anchor)
    lits = re.finditer(r"new double\[\]\s*\{([^}]*)\}", src[pos:])
    for i, m in enumerate(lits):
        if i == nth:
            return [float(t) for t in m.group(1).replace("\n", " ").split(",") if t.strip()]
    raise KeyError(anchor)


def main():
    g = {}
    # --- SubsamplingLayerTest.java:47-63, 65-81, 85-113, 115-133, data :161-165
    s = open(f"{REF}/nn/layers/convolution/SubsamplingLayerTest.java").read()
    g["subsampling"] = {
        "source": "deeplearning4j-core/.../nn/layers/convolution/SubsamplingLayerTest.java:47-133,161-165",
        "kernel": [2, 2], "stride": [2, 2], "pad": [0, 0],
        "input_shape": [1, 2, 4, 4],
        "input": _doubles_after(s, "public INDArray getContainedData()"),
        "max_out": _doubles_after(s, "public void testSubSampleMaxActivate"),
        "avg_out": _doubles_after(s, "public void testSubSampleMeanActivate"),
        "max_bwd_eps": _doubles_after(s, "public void testSubSampleLayerMaxBackprop", 0),
        "max_bwd_out": _doubles_after(s, "public void testSubSampleLayerMaxBackprop", 1),
        "avg_bwd_eps": _doubles_after(s, "public void testSubSampleLayerAvgBackprop", 0),
        "avg_bwd_out": _doubles_after(s, "public void testSubSampleLayerAvgBackprop", 1),
        "out_shape": [1, 2, 2, 2],
    }
    # --- ConvolutionLayerTest.java:264-278 (+ config :317-339)
    s = open(f"{REF}/nn/layers/convolution/ConvolutionLayerTest.java").read()
    g["conv_contained"] = {
        "source": "deeplearning4j-core/.../nn/layers/convolution/ConvolutionLayerTest.java:264-278,317-345",
        "kernel": [2, 2], "stride": [2, 2], "pad": [0, 0], "activation": "sigmoid",
        "w_shape": [2, 1, 2, 2], "w": _doubles_after(s, "public Layer getContainedConfig()", 0),
        "b": _doubles_after(s, "public Layer getContainedConfig()", 1),
        "input_shape": [1, 1, 8, 8], "input": _doubles_after(s, "public INDArray getContainedData()"),
        "out_shape": [1, 2, 4, 4], "out": _doubles_after(s, "public void testActivateResultsContained"),
        "col_shape": [1, 1, 2, 2, 4, 4], "col": _doubles_after(s, "public INDArray getContainedCol()"),
    }
    # --- ConvolutionLayerTest.java:468-546 im2col row/column order (input = arange(36) as [2,2,3,3])
    pos = s.index("public void testCnnIm2ColReshaping")
    rows = re.findall(r"exp2d\.putRow\((\d+), Nd4j\.create\(new double\[\] \{([^}]*)\}", s[pos:])
    g["im2col_order"] = {
        "source": "deeplearning4j-core/.../nn/layers/convolution/ConvolutionLayerTest.java:468-521 (+getInput :446-466)",
        "input_shape": [2, 2, 3, 3], "kernel": [2, 2], "stride": [1, 1], "pad": [0, 0],
        "exp2d": [[float(t) for t in r[1].split(",")] for r in rows[:8]],
    }
    # --- LocalResponseTest.java:36-117
    s = open(f"{REF}/nn/layers/normalization/LocalResponseTest.java").read()
    g["lrn"] = {
        "source": "deeplearning4j-core/.../nn/layers/normalization/LocalResponseTest.java:36-117",
        "shape": [2, 7, 3, 2], "k": 2.0, "n": 5.0, "alpha": 1e-4, "beta": 0.75,
        "x": _doubles_after(s, "private INDArray x ="),
        "y": _doubles_after(s, "private INDArray activationsExpected"),
        "eps": _doubles_after(s, "private INDArray epsilon"),
        "dx": _doubles_after(s, "private INDArray newEpsilonExpected"),
        "dx_asserted_indices": [8, 20], "dx_tol": 1e-4,
    }
    for k, v in g.items():
        with open(os.path.join(OUT, k + ".json"), "w") as f:
            json.dump(v, f, indent=0)
        print(k, {kk: (len(vv) if isinstance(vv, list) else vv) for kk, vv in v.items() if kk != "source"})


if __name__ == "__main__":
    main()
