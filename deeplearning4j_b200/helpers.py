"""Host-side mirror of DL4J's CNN helper SPI over the C ABI (include/dl4jb200.h).

The reference's layers look their helper up by reflection under fixed class names
(`Class.forName("org.deeplearning4j.nn.layers.convolution.CudnnConvolutionHelper")`,
deeplearning4j-nn/.../nn/layers/convolution/ConvolutionLayer.java:64-95 and the three siblings), so a
drop-in must define classes with exactly those names and method signatures. This container has no JVM
(SURVEY.md §0.3): the Java sources live in java/  and this module is the same thin
marshalling layer written in Python -- same class names, method names, argument order and error behaviour
(return None == Java `null` => the layer runs its built-in path; exceptions propagate) -- so the parity
tests read like deeplearning4j-cuda/src/test/.../TestConvolution.java.

Arguments are numpy float32 arrays in DL4J's layout (NCHW, 'c' order) living in HOST memory, exactly as
INDArrays do under nd4j-native; every call stages them to the device, runs the hand-written kernels and
copies the results back. Gradient "views" are written in place.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (ACT_IDENTITY, ACT_LEAKYRELU, ACT_RELU, ACT_SIGMOID, ACT_TANH, ALGO_FP32, ALGO_TC_TF32, NCHW, NHWC,
                   POOL_AVG, POOL_MAX, POOL_PNORM, POOL_SUM, ConvDesc, HelperNotSupported)

CONVOLUTION_MODES = {"Strict": _lib.MODE_STRICT, "Truncate": _lib.MODE_TRUNCATE, "Same": _lib.MODE_SAME}
POOLING_TYPES = {"MAX": POOL_MAX, "AVG": POOL_AVG, "SUM": POOL_SUM, "PNORM": POOL_PNORM}
# IActivation.toString() values the helper handles (CudnnConvolutionHelper.java:527-558 handles the first four)
ACTIVATIONS = {"identity": ACT_IDENTITY, "relu": ACT_RELU, "sigmoid": ACT_SIGMOID, "tanh": ACT_TANH,
               "leakyrelu": ACT_LEAKYRELU}


class Context:
    """One per worker thread / device (ParallelWrapper.java:125-141 pins a worker thread to a device)."""

    _default = {}

    def __init__(self, device=0, stream=None):
        h = C.c_void_p()
        _lib.call("dl4jb200_ctx_create", device, stream, C.byref(h))
        self.handle = h
        self.device = device
        self._pool = {}

    @classmethod
    def default(cls, device=0):
        if device not in cls._default:
            cls._default[device] = cls(device)
        return cls._default[device]

    def close(self):
        if self.handle:
            for ptr, _ in self._pool.values():
                _lib.call("dl4jb200_free", self.handle, ptr)
            self._pool.clear()
            _lib.call("dl4jb200_ctx_destroy", self.handle)
            self.handle = None

    def sync(self):
        _lib.call("dl4jb200_ctx_sync", self.handle)

    def malloc(self, nbytes):
        p = C.c_void_p()
        _lib.call("dl4jb200_malloc", self.handle, nbytes, C.byref(p))
        return p

    def free(self, ptr):
        _lib.call("dl4jb200_free", self.handle, ptr)

    def scratch(self, key, nbytes):
        """Device staging buffer cached by key (the mirror of one INDArray argument slot)."""
        cur = self._pool.get(key)
        if cur is None or cur[1] < nbytes:
            if cur is not None:
                self.free(cur[0])
            cur = (self.malloc(nbytes), nbytes)
            self._pool[key] = cur
        return cur[0]

    def upload(self, key, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float32)
        p = self.scratch(key, max(arr.nbytes, 16))
        _lib.call("dl4jb200_memcpy_h2d", self.handle, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes)
        return p

    def download(self, ptr, shape, dtype=np.float32, out=None):
        if out is None:
            out = np.empty(shape, dtype)
        assert out.flags.c_contiguous
        _lib.call("dl4jb200_memcpy_d2h", self.handle, out.ctypes.data_as(C.c_void_p), ptr, out.nbytes)
        return out


def _write_view(view, values):
    """INDArray.assign into a gradient view (any strides), in place."""
    view[...] = values.reshape(view.shape)


def _is_f32(*arrs):
    return all(a is None or a.dtype == np.float32 for a in arrs)


def conv_desc(x_shape, c_out, kernel, strides, pad, dilation, convolution_mode):
    """ConvolutionUtils.getOutputSize + SAME-mode padding for one call (ConvolutionLayer.java:140-150,314-325)."""
    B, Cin, H, W = x_shape
    mode = CONVOLUTION_MODES[convolution_mode] if isinstance(convolution_mode, str) else convolution_mode
    Ho, pT, _ = _lib.out_size(H, kernel[0], strides[0], pad[0], dilation[0], mode)
    Wo, pL, _ = _lib.out_size(W, kernel[1], strides[1], pad[1], dilation[1], mode)
    return ConvDesc(B, Cin, H, W, c_out, kernel[0], kernel[1], strides[0], strides[1], pT, pL, dilation[0], dilation[1],
                    Ho, Wo)


class _BaseHelper:
    def __init__(self, ctx=None):
        self.ctx = ctx or Context.default()

    def _supported(self):
        return True  # constructing the Context already verified an sm_100 device and the loaded library


class CudnnConvolutionHelper(_BaseHelper):
    """ConvolutionHelper (nn/layers/convolution/ConvolutionHelper.java:35-47)."""

    def __init__(self, ctx=None, algo=ALGO_TC_TF32):
        super().__init__(ctx)
        self.algo = algo

    def checkSupported(self):
        return self._supported()

    def preOutput(self, input, weights, bias, kernel, strides, pad, mode=None, fwdAlgo=None,
                  convolutionMode="Truncate", dilation=(1, 1)):
        if not _is_f32(input, weights, bias):
            return None  # DOUBLE (the reference's -Ddtype=double test runs) falls back to the built-in path
        c = self.ctx
        d = conv_desc(input.shape, weights.shape[0], kernel, strides, pad, dilation, convolutionMode)
        px = c.upload("conv.x", input)
        pw = c.upload("conv.w", weights)
        pb = c.upload("conv.b", bias.reshape(-1)) if bias is not None else None
        zshape = (d.batch, d.c_out, d.out_h, d.out_w)
        pz = c.scratch("conv.z", int(np.prod(zshape)) * 4)
        _lib.call("dl4jb200_conv_fwd", c.handle, C.byref(d), px, NCHW, 0, pw, pb, pz, NCHW, 0, ACT_IDENTITY, self.algo)
        return c.download(pz, zshape)

    def activate(self, z, afn):
        kind = ACTIVATIONS.get(str(afn).lower())
        if kind is None or not _is_f32(z):
            return None
        c = self.ctx
        pz = c.upload("conv.act", z)
        _lib.call("dl4jb200_act_fwd", c.handle, pz, pz, z.size, kind)
        out = c.download(pz, z.shape)
        z[...] = out  # in place on z, as the cuDNN helper does (ConvolutionLayer.java:456-461)
        return z

    def backpropGradient(self, input, weights, delta, kernel, strides, pad, biasGradView, weightGradView, afn=None,
                         mode=None, bwdFilterAlgo=None, bwdDataAlgo=None, convolutionMode="Truncate", dilation=(1, 1)):
        if not _is_f32(input, weights, delta):
            return None
        c = self.ctx
        d = conv_desc(input.shape, weights.shape[0], kernel, strides, pad, dilation, convolutionMode)
        assert tuple(delta.shape) == (d.batch, d.c_out, d.out_h, d.out_w), "delta shape"
        px = c.upload("conv.x", input)
        pw = c.upload("conv.w", weights)
        pd = c.upload("conv.delta", delta)
        pdw = c.scratch("conv.dw", weights.size * 4)
        pdb = c.scratch("conv.db", d.c_out * 4)
        pdx = c.scratch("conv.dx", input.size * 4)
        _lib.call("dl4jb200_conv_bwd", c.handle, C.byref(d), px, NCHW, 0, pw, pd, NCHW, 0, pdw, pdb, pdx, NCHW, 0,
                  self.algo)
        _write_view(weightGradView, c.download(pdw, weights.shape))
        if biasGradView is not None:
            _write_view(biasGradView, c.download(pdb, (d.c_out,)))
        eps_next = c.download(pdx, input.shape)
        # Gradient.setGradientFor(WEIGHT_KEY, weightGradView, 'c') (CudnnConvolutionHelper.java:314-316)
        return {"W": weightGradView, "b": biasGradView}, eps_next


class CudnnSubsamplingHelper(_BaseHelper):
    """SubsamplingHelper (nn/layers/convolution/subsampling/SubsamplingHelper.java:31-39)."""

    def __init__(self, ctx=None):
        super().__init__(ctx)
        self._argmax = None  # kept between activate and backpropGradient like the cuDNN helper's `reduced`
        self._reduced = None

    def checkSupported(self):
        return self._supported()

    def activate(self, input, training, kernel, strides, pad, poolingType, convolutionMode="Truncate", dilation=(1, 1),
                 pnorm=2):
        if not _is_f32(input):
            return None
        c = self.ctx
        pt = POOLING_TYPES[poolingType] if isinstance(poolingType, str) else poolingType
        d = conv_desc(input.shape, input.shape[1], kernel, strides, pad, dilation, convolutionMode)
        px = c.upload("pool.x", input)
        yshape = (d.batch, d.c_in, d.out_h, d.out_w)
        n = int(np.prod(yshape))
        py = c.scratch("pool.y", n * 4)
        pa = c.scratch("pool.argmax", n * 4) if pt == POOL_MAX else None
        _lib.call("dl4jb200_pool_fwd", c.handle, C.byref(d), px, py, pa, NCHW, 0, 0, pt, pnorm)
        self._argmax = (pa, yshape) if (training and pa is not None) else None
        self._reduced = (py, yshape)
        return c.download(py, yshape)

    def argmax(self):
        """int32 window positions (kh*kW+kw) of the last MAX forward; bit-exact tie-break check."""
        pa, shape = self._argmax
        return self.ctx.download(pa, shape, np.int32)

    def backpropGradient(self, input, epsilon, kernel, strides, pad, poolingType, convolutionMode="Truncate",
                         dilation=(1, 1), pnorm=2, pnorm_eps=1e-8):
        if not _is_f32(input, epsilon):
            return None
        c = self.ctx
        pt = POOLING_TYPES[poolingType] if isinstance(poolingType, str) else poolingType
        d = conv_desc(input.shape, input.shape[1], kernel, strides, pad, dilation, convolutionMode)
        assert tuple(epsilon.shape) == (d.batch, d.c_in, d.out_h, d.out_w), "epsilon shape"
        px = c.upload("pool.x", input)
        pe = c.upload("pool.eps", epsilon)
        pdx = c.scratch("pool.dx", input.size * 4)
        py = None
        if pt == POOL_PNORM:
            py = c.scratch("pool.y", epsilon.size * 4)
            _lib.call("dl4jb200_pool_fwd", c.handle, C.byref(d), px, py, None, NCHW, 0, 0, pt, pnorm)
        # argmax is recomputed from x on the device, exactly as the reference's im2col + IsMax does
        # (SubsamplingLayer.java:205-218): the helper call is stateless with respect to the forward pass.
        try:
            _lib.call("dl4jb200_pool_bwd", c.handle, C.byref(d), px, py, pe, None, pdx, NCHW, 0, 0, pt, pnorm,
                      float(pnorm_eps))
        except HelperNotSupported:
            return None
        return {}, c.download(pdx, input.shape)


class CudnnBatchNormalizationHelper(_BaseHelper):
    """BatchNormalizationHelper (nn/layers/normalization/BatchNormalizationHelper.java:29-37)."""

    def __init__(self, ctx=None):
        super().__init__(ctx)
        self._saved = None  # (mean, std) device buffers of the last training forward (meanCache / varCache)

    def checkSupported(self, eps=1e-5):
        return self._supported()  # no cuDNN minimum-eps restriction (CudnnBatchNormalizationHelper.java:100-107)

    def preOutput(self, x, training, shape, gamma, beta, mean, var, decay, eps, minibatch=True):
        if not _is_f32(x, gamma, beta, mean, var) or x.ndim != 4:
            return None
        c = self.ctx
        B, Cc, H, W = x.shape
        px = c.upload("bn.x", x)
        pg = c.upload("bn.gamma", gamma.reshape(-1)) if gamma is not None else None
        pb = c.upload("bn.beta", beta.reshape(-1)) if beta is not None else None
        pm = c.upload("bn.gmean", mean.reshape(-1))
        pv = c.upload("bn.gvar", var.reshape(-1))
        py = c.scratch("bn.y", x.size * 4)
        psm = c.scratch("bn.save_mean", Cc * 4)
        pss = c.scratch("bn.save_std", Cc * 4)
        _lib.call("dl4jb200_bn_fwd", c.handle, px, py, pg, pb, pm, pv, psm, pss, B, Cc, H, W, NCHW, 0, float(decay),
                  float(eps), int(bool(training)), int(bool(minibatch)), ACT_IDENTITY, None)
        if training:  # running statistics are parameters, updated in place in the flat vector
            _write_view(mean, c.download(pm, (Cc,)))
            _write_view(var, c.download(pv, (Cc,)))
            self._saved = (psm, pss, Cc)
        return c.download(py, x.shape)

    def backpropGradient(self, input, epsilon, shape, gamma, dGammaView, dBetaView, eps):
        if not _is_f32(input, epsilon, gamma) or input.ndim != 4:
            return None
        if self._saved is None:
            raise RuntimeError("backpropGradient before a training preOutput")
        c = self.ctx
        B, Cc, H, W = input.shape
        psm, pss, _ = self._saved
        px = c.upload("bn.x", input)
        pe = c.upload("bn.eps", epsilon)
        pg = c.upload("bn.gamma", gamma.reshape(-1)) if gamma is not None else None
        pdx = c.scratch("bn.dx", input.size * 4)
        pdg = c.scratch("bn.dgamma", Cc * 4)
        pdb = c.scratch("bn.dbeta", Cc * 4)
        _lib.call("dl4jb200_bn_bwd", c.handle, px, pe, pg, None, psm, pss, ACT_IDENTITY, pdx, pdg, pdb, B, Cc, H, W, NCHW, 0)
        _write_view(dGammaView, c.download(pdg, (Cc,)))
        _write_view(dBetaView, c.download(pdb, (Cc,)))
        return {"gamma": dGammaView, "beta": dBetaView}, c.download(pdx, input.shape)


class CudnnLocalResponseNormalizationHelper(_BaseHelper):
    """LocalResponseNormalizationHelper (nn/layers/normalization/LocalResponseNormalizationHelper.java:29-36)."""

    def checkSupported(self, k=2.0, n=5.0, alpha=1e-4, beta=0.75):
        return self._supported() and n >= 1  # no cuDNN n/k range limits (CudnnLocalResponseNormalizationHelper.java:100-121)

    def activate(self, x, training, k, n, alpha, beta):
        if not _is_f32(x):
            return None
        c = self.ctx
        B, Cc, H, W = x.shape
        px = c.upload("lrn.x", x)
        py = c.scratch("lrn.y", x.size * 4)
        _lib.call("dl4jb200_lrn_fwd", c.handle, px, py, B, Cc, H, W, NCHW, 0, float(k), float(n), float(alpha), float(beta))
        return c.download(py, x.shape)

    def backpropGradient(self, input, epsilon, k, n, alpha, beta):
        if not _is_f32(input, epsilon):
            return None
        c = self.ctx
        B, Cc, H, W = input.shape
        px = c.upload("lrn.x", input)
        pe = c.upload("lrn.eps", epsilon)
        pdx = c.scratch("lrn.dx", input.size * 4)
        _lib.call("dl4jb200_lrn_bwd", c.handle, px, pe, pdx, B, Cc, H, W, NCHW, 0, float(k), float(n), float(alpha),
                  float(beta))
        return {}, c.download(pdx, input.shape)
