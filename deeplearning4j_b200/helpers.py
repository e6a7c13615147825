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
INDArrays do under nd4j-native. Activations and epsilons a helper RETURNS are `ResidentArray`s: the INDArray
equivalent whose current contents live in the context's host-mirror registry on the device (dl4jb200_buf_acquire
et al.). Passed to the next helper call they are used where they are -- nothing crosses PCIe -- and they come back to
host memory only when host code reads them (np.asarray, indexing, arithmetic), which is what nd4j-cuda's allocator
gives the reference's cuDNN helper (CudnnConvolutionHelper.java:442-447). A conv -> batch-norm -> pooling chain
therefore uploads the minibatch and the parameters once and downloads only the final result. Plain numpy inputs are
host-owned arrays of unknown freshness and are uploaded on every call; parameter gradient "views" are written in
place on the host (the parameter boundary). Context(resident=False) restores copy-in / copy-out for every call.
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

    def __init__(self, device=0, stream=None, resident=True):
        h = C.c_void_p()
        _lib.call("dl4jb200_ctx_create", device, stream, C.byref(h))
        self.handle = h
        self.device = device
        self.resident = resident
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

    def ain(self, key, arr, access=_lib.BUF_READ):
        """Device address of an activation / epsilon operand: a ResidentArray is used where it is (uploaded only if host
        code wrote it since), a plain numpy array is host memory of unknown freshness and is staged through slot `key`."""
        if isinstance(arr, ResidentArray) and arr.ctx is self:
            return arr.device_ptr(access)
        return self.upload(key, np.asarray(arr))

    def aout(self, shape, dtype=np.float32):
        """A new result array, device-resident: (array to return, device address to write)."""
        r = ResidentArray(self, shape, dtype)
        return r, r.dev

    def result(self, r):
        """What the helper hands back: the resident array, or (Context(resident=False)) its host copy right away."""
        return r if self.resident else np.array(r.numpy())

    def transfer_bytes(self):
        """(host->device, device->host) bytes this context has moved so far."""
        a, b = C.c_ulonglong(), C.c_ulonglong()
        _lib.call("dl4jb200_ctx_transfer_bytes", self.handle, C.byref(a), C.byref(b))
        return a.value, b.value

    def download(self, ptr, shape, dtype=np.float32, out=None):
        if out is None:
            out = np.empty(shape, dtype)
        assert out.flags.c_contiguous
        _lib.call("dl4jb200_memcpy_d2h", self.handle, out.ctypes.data_as(C.c_void_p), ptr, out.nbytes)
        return out


class ResidentArray(np.lib.mixins.NDArrayOperatorsMixin):
    """A helper result whose current contents are on the device (state DEVICE_NEWER in the context's registry, keyed by
    the address of its host buffer exactly like an INDArray's). Host code that reads it -- np.asarray, indexing,
    arithmetic, any ndarray attribute -- triggers dl4jb200_buf_sync_to_host once; helper calls use the device copy."""

    __array_priority__ = 100

    def __init__(self, ctx, shape, dtype=np.float32, host=None):
        self.ctx, self.shape, self.dtype = ctx, tuple(int(v) for v in shape), np.dtype(dtype)
        p = C.c_void_p()
        if host is None:  # a helper result: born on the device
            self._host = np.empty(self.shape, self.dtype)  # untouched pages cost nothing until the host really reads
            self._key = C.c_void_p(self._host.ctypes.data)
            _lib.call("dl4jb200_buf_acquire", ctx.handle, self._key, max(self._host.nbytes, 16), _lib.BUF_WRITE, C.byref(p))
        else:             # host data adopted into the registry (state HOST_NEWER: the first kernel that reads it uploads)
            self._host = host
            self._key = C.c_void_p(self._host.ctypes.data)
            _lib.call("dl4jb200_buf_register", ctx.handle, self._key, max(self._host.nbytes, 16), C.byref(p))
            _lib.call("dl4jb200_buf_mark_host_dirty", ctx.handle, self._key)
        self.dev = p

    @classmethod
    def from_host(cls, ctx, arr):
        """Registers a host array (the minibatch, say) so that it is uploaded once and then reused by every helper call
        that reads it, until host code declares a write with mark_host_dirty()."""
        a = np.ascontiguousarray(arr, dtype=np.float32)
        return cls(ctx, a.shape, a.dtype, host=a)

    ndim = property(lambda self: len(self.shape))
    size = property(lambda self: int(np.prod(self.shape)))
    nbytes = property(lambda self: self.size * self.dtype.itemsize)

    def device_ptr(self, access=_lib.BUF_READ):
        """Device address for a kernel that reads (default) and / or writes this array."""
        p = C.c_void_p()
        _lib.call("dl4jb200_buf_acquire", self.ctx.handle, self._key, max(self._host.nbytes, 16), access, C.byref(p))
        self.dev = p
        return p

    def state(self):
        st = C.c_int()
        _lib.call("dl4jb200_buf_state", self.ctx.handle, self._key, C.byref(st))
        return st.value

    def numpy(self):
        """The host view, made current first (a device -> host copy only if the device copy is newer)."""
        _lib.call("dl4jb200_buf_sync_to_host", self.ctx.handle, self._key)
        return self._host

    def mark_host_dirty(self):
        """Host code wrote through numpy(): the next helper call uploads."""
        _lib.call("dl4jb200_buf_mark_host_dirty", self.ctx.handle, self._key)

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a if dtype is None else a.astype(dtype, copy=False)

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        inputs = tuple(x.numpy() if isinstance(x, ResidentArray) else x for x in inputs)
        if "out" in kwargs:
            kwargs["out"] = tuple(x.numpy() if isinstance(x, ResidentArray) else x for x in kwargs["out"])
        return getattr(ufunc, method)(*inputs, **kwargs)

    def __getitem__(self, idx):
        return self.numpy()[idx]

    def __len__(self):
        return self.shape[0]

    def __getattr__(self, name):  # anything else an ndarray offers (reshape, max, ravel, ...) acts on the host view
        if name.startswith("_") or name in ("ctx", "shape", "dtype", "dev"):
            raise AttributeError(name)
        return getattr(self.numpy(), name)

    def __del__(self):
        try:
            if self.ctx.handle:
                _lib.call("dl4jb200_buf_release", self.ctx.handle, self._key)
        except Exception:  # noqa: BLE001 -- interpreter shutdown
            pass


def _write_view(view, values):
    """INDArray.assign into a gradient view (any strides), in place."""
    view[...] = values.reshape(view.shape)


def _is_f32(*arrs):
    return all(a is None or a.dtype == np.float32 for a in arrs)


def _host(a):
    """numpy view of a parameter-sized operand (always staged from host memory)."""
    return a.numpy() if isinstance(a, ResidentArray) else a


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
        px = c.ain("conv.x", input)
        pw = c.upload("conv.w", weights)
        pb = c.upload("conv.b", bias.reshape(-1)) if bias is not None else None
        z, pz = c.aout((d.batch, d.c_out, d.out_h, d.out_w))
        _lib.call("dl4jb200_conv_fwd", c.handle, C.byref(d), px, NCHW, 0, pw, pb, pz, NCHW, 0, ACT_IDENTITY, self.algo)
        return c.result(z)

    def activate(self, z, afn):
        kind = ACTIVATIONS.get(str(afn).lower())
        if kind is None or not _is_f32(z):
            return None
        c = self.ctx
        if isinstance(z, ResidentArray):  # in place on the device copy, as the cuDNN helper does (ConvolutionLayer.java:456-461)
            pz = z.device_ptr(_lib.BUF_READ | _lib.BUF_WRITE)
            _lib.call("dl4jb200_act_fwd", c.handle, pz, pz, z.size, kind)
            return z
        pz = c.upload("conv.act", z)
        _lib.call("dl4jb200_act_fwd", c.handle, pz, pz, z.size, kind)
        out = c.download(pz, z.shape)
        z[...] = out  # in place on z
        return z

    def backpropGradient(self, input, weights, delta, kernel, strides, pad, biasGradView, weightGradView, afn=None,
                         mode=None, bwdFilterAlgo=None, bwdDataAlgo=None, convolutionMode="Truncate", dilation=(1, 1)):
        if not _is_f32(input, weights, delta):
            return None
        c = self.ctx
        d = conv_desc(input.shape, weights.shape[0], kernel, strides, pad, dilation, convolutionMode)
        assert tuple(delta.shape) == (d.batch, d.c_out, d.out_h, d.out_w), "delta shape"
        px = c.ain("conv.x", input)
        pw = c.upload("conv.w", weights)
        pd = c.ain("conv.delta", delta)
        pdw = c.scratch("conv.dw", weights.size * 4)
        pdb = c.scratch("conv.db", d.c_out * 4)
        eps_next, pdx = c.aout(input.shape)
        _lib.call("dl4jb200_conv_bwd", c.handle, C.byref(d), px, NCHW, 0, pw, pd, NCHW, 0, pdw, pdb, pdx, NCHW, 0,
                  self.algo)
        # the gradient views belong to the network's flat gradient vector on the host: the parameter boundary
        _write_view(weightGradView, c.download(pdw, weights.shape))
        if biasGradView is not None:
            _write_view(biasGradView, c.download(pdb, (d.c_out,)))
        # Gradient.setGradientFor(WEIGHT_KEY, weightGradView, 'c') (CudnnConvolutionHelper.java:314-316)
        return {"W": weightGradView, "b": biasGradView}, c.result(eps_next)


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
        px = c.ain("pool.x", input)
        yshape = (d.batch, d.c_in, d.out_h, d.out_w)
        n = int(np.prod(yshape))
        y, py = c.aout(yshape)
        pa = c.scratch("pool.argmax", n * 4) if pt == POOL_MAX else None
        _lib.call("dl4jb200_pool_fwd", c.handle, C.byref(d), px, py, pa, NCHW, 0, 0, pt, pnorm)
        self._argmax = (pa, yshape) if (training and pa is not None) else None
        self._reduced = y
        return c.result(y)

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
        px = c.ain("pool.x", input)
        pe = c.ain("pool.eps", epsilon)
        dx, pdx = c.aout(input.shape)
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
        return {}, c.result(dx)


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
        px = c.ain("bn.x", x)
        pg = c.upload("bn.gamma", gamma.reshape(-1)) if gamma is not None else None
        pb = c.upload("bn.beta", beta.reshape(-1)) if beta is not None else None
        pm = c.upload("bn.gmean", mean.reshape(-1))
        pv = c.upload("bn.gvar", var.reshape(-1))
        y, py = c.aout(x.shape)
        psm = c.scratch("bn.save_mean", Cc * 4)
        pss = c.scratch("bn.save_std", Cc * 4)
        _lib.call("dl4jb200_bn_fwd", c.handle, px, py, pg, pb, pm, pv, psm, pss, B, Cc, H, W, NCHW, 0, float(decay),
                  float(eps), int(bool(training)), int(bool(minibatch)), ACT_IDENTITY, None)
        if training:  # running statistics are parameters, updated in place in the flat vector
            _write_view(mean, c.download(pm, (Cc,)))
            _write_view(var, c.download(pv, (Cc,)))
            self._saved = (psm, pss, Cc)
        return c.result(y)

    def backpropGradient(self, input, epsilon, shape, gamma, dGammaView, dBetaView, eps):
        if not _is_f32(input, epsilon, gamma) or input.ndim != 4:
            return None
        if self._saved is None:
            raise RuntimeError("backpropGradient before a training preOutput")
        c = self.ctx
        B, Cc, H, W = input.shape
        psm, pss, _ = self._saved
        px = c.ain("bn.x", input)
        pe = c.ain("bn.eps", epsilon)
        pg = c.upload("bn.gamma", gamma.reshape(-1)) if gamma is not None else None
        dx, pdx = c.aout(input.shape)
        pdg = c.scratch("bn.dgamma", Cc * 4)
        pdb = c.scratch("bn.dbeta", Cc * 4)
        _lib.call("dl4jb200_bn_bwd", c.handle, px, pe, pg, None, psm, pss, ACT_IDENTITY, pdx, pdg, pdb, B, Cc, H, W, NCHW, 0)
        _write_view(dGammaView, c.download(pdg, (Cc,)))
        _write_view(dBetaView, c.download(pdb, (Cc,)))
        return {"gamma": dGammaView, "beta": dBetaView}, c.result(dx)


class CudnnLocalResponseNormalizationHelper(_BaseHelper):
    """LocalResponseNormalizationHelper (nn/layers/normalization/LocalResponseNormalizationHelper.java:29-36)."""

    def checkSupported(self, k=2.0, n=5.0, alpha=1e-4, beta=0.75):
        return self._supported() and n >= 1  # no cuDNN n/k range limits (CudnnLocalResponseNormalizationHelper.java:100-121)

    def activate(self, x, training, k, n, alpha, beta):
        if not _is_f32(x):
            return None
        c = self.ctx
        B, Cc, H, W = x.shape
        px = c.ain("lrn.x", x)
        y, py = c.aout(x.shape)
        _lib.call("dl4jb200_lrn_fwd", c.handle, px, py, B, Cc, H, W, NCHW, 0, float(k), float(n), float(alpha), float(beta))
        return c.result(y)

    def backpropGradient(self, input, epsilon, k, n, alpha, beta):
        if not _is_f32(input, epsilon):
            return None
        c = self.ctx
        B, Cc, H, W = input.shape
        px = c.ain("lrn.x", input)
        pe = c.ain("lrn.eps", epsilon)
        dx, pdx = c.aout(input.shape)
        _lib.call("dl4jb200_lrn_bwd", c.handle, px, pe, pdx, B, Cc, H, W, NCHW, 0, float(k), float(n), float(alpha),
                  float(beta))
        return {}, c.result(dx)
