"""ctypes front-end of the CPU parity oracle (oracle/dl4j_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, bench.py's cpu_baseline / --impl reference leg and
__graft_entry__.smoke(), never by the product package. All arrays are numpy float32, NCHW, C-contiguous.
Reference citations are in dl4j_oracle.c next to each routine.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdl4joracle.so")

STRICT, TRUNCATE, SAME = 0, 1, 2
POOL_MAX, POOL_AVG, POOL_SUM, POOL_PNORM = 0, 1, 2, 3
ACT_IDENTITY, ACT_RELU, ACT_SIGMOID, ACT_TANH, ACT_LEAKYRELU = 0, 1, 2, 3, 4
UPD_SGD, UPD_NESTEROVS, UPD_RMSPROP, UPD_ADADELTA, UPD_NOOP = 0, 1, 2, 3, 4


def build(force=False):
    src = os.path.join(_HERE, "dl4j_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.o_softmax_nll.restype = C.c_double
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class DL4JInvalidInputException(ValueError):
    pass


class DL4JInvalidConfigException(ValueError):
    pass


def out_size(inp, k, s, p, d=1, mode=TRUNCATE):
    """-> (out, pad_lo, pad_hi); raises like ConvolutionUtils.validateShapes."""
    o, lo, hi = C.c_int(), C.c_int(), C.c_int()
    rc = lib().o_out_size(inp, k, s, p, d, mode, C.byref(o), C.byref(lo), C.byref(hi))
    if rc == 1:
        raise DL4JInvalidInputException("kernel vs input+2*padding")
    if rc == 2:
        raise DL4JInvalidConfigException("ConvolutionMode.Strict: not divisible")
    return o.value, lo.value, hi.value


def conv_geometry(H, W, kernel, stride, pad, dilation=(1, 1), mode=TRUNCATE):
    Ho, pT, _ = out_size(H, kernel[0], stride[0], pad[0], dilation[0], mode)
    Wo, pL, _ = out_size(W, kernel[1], stride[1], pad[1], dilation[1], mode)
    return Ho, Wo, pT, pL


def im2col(x, kernel, stride, pad, dilation=(1, 1), mode=TRUNCATE, padval=0.0):
    x = _f(x)
    B, Cc, H, W = x.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, kernel, stride, pad, dilation, mode)
    col = np.empty((B, Ho, Wo, Cc, kernel[0], kernel[1]), np.float32)
    lib().o_im2col(_p(x), _p(col), B, Cc, H, W, kernel[0], kernel[1], stride[0], stride[1], pT, pL,
                   dilation[0], dilation[1], Ho, Wo, C.c_float(padval))
    return col


def col2im(col, H, W, stride, pad_tl, dilation=(1, 1)):
    col = _f(col)
    B, Ho, Wo, Cc, kH, kW = col.shape
    dx = np.empty((B, Cc, H, W), np.float32)
    lib().o_col2im(_p(col), _p(dx), B, Cc, H, W, kH, kW, stride[0], stride[1], pad_tl[0], pad_tl[1],
                   dilation[0], dilation[1], Ho, Wo)
    return dx


def conv_fwd(x, w, b, kernel=None, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    x, w = _f(x), _f(w)
    b = None if b is None else _f(b).reshape(-1)
    B, Cin, H, W = x.shape
    Cout, _, kH, kW = w.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, (kH, kW), stride, pad, dilation, mode)
    z = np.empty((B, Cout, Ho, Wo), np.float32)
    lib().o_conv_fwd(_p(x), _p(w), _p(b), _p(z), B, Cin, H, W, Cout, kH, kW, stride[0], stride[1], pT, pL,
                     dilation[0], dilation[1], Ho, Wo)
    return z


def conv_fwd_im2col(x, w, b, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    """The literal reference pipeline (ConvolutionLayer.java:385-417): im2col -> sgemm (BLAS, fp32) ->
    +bias -> permute. This is the form timed as the CPU baseline."""
    x, w = _f(x), _f(w)
    B, Cin, H, W = x.shape
    Cout, _, kH, kW = w.shape
    col = im2col(x, (kH, kW), stride, pad, dilation, mode)
    _, Ho, Wo = col.shape[:3]
    z2d = col.reshape(B * Ho * Wo, Cin * kH * kW) @ w.reshape(Cout, -1).T
    if b is not None:
        z2d += _f(b).reshape(1, -1)
    z = np.empty((B, Cout, Ho, Wo), np.float32)
    lib().o_rows_to_nchw(_p(np.ascontiguousarray(z2d)), _p(z), B, Cout, Ho, Wo)
    return z


def conv_bwd(x, w, delta, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE, need_dx=True):
    x, w, delta = _f(x), _f(w), _f(delta)
    B, Cin, H, W = x.shape
    Cout, _, kH, kW = w.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, (kH, kW), stride, pad, dilation, mode)
    assert delta.shape == (B, Cout, Ho, Wo)
    dw = np.empty_like(w)
    db = np.empty((Cout,), np.float32)
    dx = np.empty_like(x) if need_dx else None
    lib().o_conv_bwd(_p(x), _p(w), _p(delta), _p(dw), _p(db), _p(dx), B, Cin, H, W, Cout, kH, kW, stride[0],
                     stride[1], pT, pL, dilation[0], dilation[1], Ho, Wo)
    return dw, db, dx


def conv_bwd_im2col(x, w, delta, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE, acc64=False):
    """Literal reference pipeline (ConvolutionLayer.java:202-257): im2col, two sgemms, col2im. acc64: run the
    filter-gradient GEMM (reduction over B*Ho*Wo, up to millions of terms) in double, for use as a parity oracle."""
    x, w, delta = _f(x), _f(w), _f(delta)
    B, Cin, H, W = x.shape
    Cout, _, kH, kW = w.shape
    col = im2col(x, (kH, kW), stride, pad, dilation, mode)
    _, Ho, Wo = col.shape[:3]
    _, _, pT, pL = conv_geometry(H, W, (kH, kW), stride, pad, dilation, mode)
    d2 = np.empty((B * Ho * Wo, Cout), np.float32)
    lib().o_nchw_to_rows(_p(delta), _p(d2), B, Cout, Ho, Wo)
    col2d = col.reshape(B * Ho * Wo, Cin * kH * kW)
    if acc64:
        dw = np.zeros((Cout, col2d.shape[1]), np.float64)
        step = max(1, (1 << 26) // max(col2d.shape[1], Cout))  # bounded double-precision temporaries
        for r0 in range(0, col2d.shape[0], step):
            dw += d2[r0:r0 + step].T.astype(np.float64) @ col2d[r0:r0 + step].astype(np.float64)
        dw = dw.astype(np.float32).reshape(w.shape)
    else:
        dw = (d2.T @ col2d).reshape(w.shape)
    db = d2.sum(0, dtype=np.float64).astype(np.float32)
    eps2d = d2 @ w.reshape(Cout, -1)
    dx = col2im(eps2d.reshape(B, Ho, Wo, Cin, kH, kW), H, W, stride, (pT, pL), dilation)
    return dw, db, dx


def pool_fwd(x, kernel, stride, pad=(0, 0), dilation=(1, 1), mode=TRUNCATE, ptype=POOL_MAX, pnorm=2,
             want_argmax=False):
    x = _f(x)
    B, Cc, H, W = x.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, kernel, stride, pad, dilation, mode)
    y = np.empty((B, Cc, Ho, Wo), np.float32)
    am = np.empty((B, Cc, Ho, Wo), np.int32) if want_argmax else None
    lib().o_pool_fwd(_p(x), _p(y), _p(am), B, Cc, H, W, kernel[0], kernel[1], stride[0], stride[1], pT, pL,
                     dilation[0], dilation[1], Ho, Wo, ptype, pnorm)
    return (y, am) if want_argmax else y


def pool_bwd(x, eps, kernel, stride, pad=(0, 0), dilation=(1, 1), mode=TRUNCATE, ptype=POOL_MAX, pnorm=2,
             pnorm_eps=1e-8):
    x, eps = _f(x), _f(eps)
    B, Cc, H, W = x.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, kernel, stride, pad, dilation, mode)
    assert eps.shape == (B, Cc, Ho, Wo)
    dx = np.empty_like(x)
    rc = lib().o_pool_bwd(_p(x), _p(eps), _p(dx), B, Cc, H, W, kernel[0], kernel[1], stride[0], stride[1], pT,
                          pL, dilation[0], dilation[1], Ho, Wo, ptype, pnorm, C.c_double(pnorm_eps))
    if rc:
        raise RuntimeError("SUM pooling backprop unsupported (SubsamplingLayer.java:251-253)")
    return dx


def bn_fwd(x, gamma, beta, gmean, gvar, decay=0.9, eps=1e-5, training=True, minibatch=True):
    """gmean / gvar are updated IN PLACE when training. -> (y, save_mean, save_std)"""
    x = _f(x)
    B, Cc, H, W = x.shape
    assert gmean.dtype == np.float32 and gvar.dtype == np.float32
    y = np.empty_like(x)
    sm, ss = np.empty(Cc, np.float32), np.empty(Cc, np.float32)
    lib().o_bn_fwd(_p(x), _p(y), _p(None if gamma is None else _f(gamma)), _p(None if beta is None else _f(beta)),
                   _p(gmean), _p(gvar), _p(sm), _p(ss), B, Cc, H, W, C.c_double(decay), C.c_double(eps),
                   int(training), int(minibatch))
    return y, sm, ss


def bn_bwd(x, eps_in, gamma, mean, std):
    x, eps_in = _f(x), _f(eps_in)
    B, Cc, H, W = x.shape
    dx = np.empty_like(x)
    dg, db = np.empty(Cc, np.float32), np.empty(Cc, np.float32)
    lib().o_bn_bwd(_p(x), _p(eps_in), _p(None if gamma is None else _f(gamma)), _p(_f(mean)), _p(_f(std)), _p(dx),
                   _p(dg), _p(db), B, Cc, H, W)
    return dx, dg, db


def lrn_fwd(x, k=2.0, n=5.0, alpha=1e-4, beta=0.75):
    x = _f(x)
    B, Cc, H, W = x.shape
    y = np.empty_like(x)
    lib().o_lrn_fwd(_p(x), _p(y), None, None, B, Cc, H, W, C.c_double(k), C.c_double(n), C.c_double(alpha),
                    C.c_double(beta))
    return y


def lrn_bwd(x, eps_in, k=2.0, n=5.0, alpha=1e-4, beta=0.75):
    x, eps_in = _f(x), _f(eps_in)
    B, Cc, H, W = x.shape
    dx = np.empty_like(x)
    lib().o_lrn_bwd(_p(x), _p(eps_in), _p(dx), B, Cc, H, W, C.c_double(k), C.c_double(n), C.c_double(alpha),
                    C.c_double(beta))
    return dx


def act_fwd(z, kind):
    z = _f(z)
    a = np.empty_like(z)
    lib().o_act_fwd(_p(z), _p(a), C.c_size_t(z.size), kind)
    return a


def act_bwd(z, eps_in, kind):
    z, eps_in = _f(z), _f(eps_in)
    d = np.empty_like(z)
    lib().o_act_bwd(_p(z), _p(eps_in), _p(d), C.c_size_t(z.size), kind)
    return d


def softmax_nll(z, labels):
    """-> (loss_sum, p, delta)"""
    z, labels = _f(z), _f(labels)
    B, K = z.shape
    p, d = np.empty_like(z), np.empty_like(z)
    loss = lib().o_softmax_nll(_p(z), _p(labels), _p(p), _p(d), B, K)
    return loss, p, d


def updater_apply(w, g, state, kind, lr=0.0, h1=0.0, h2=0.0, l1=0.0, l2=0.0, batch_div=1.0):
    """In place on w and state (float32, contiguous)."""
    assert w.dtype == np.float32 and w.flags.c_contiguous
    g = _f(g)
    lib().o_updater_apply(_p(w), _p(g), _p(state), C.c_size_t(w.size), kind, C.c_double(lr), C.c_double(h1),
                          C.c_double(h2), C.c_double(l1), C.c_double(l2), C.c_double(batch_div))


def dense_fwd(x, wf, b):
    """wf: [nOut, nIn] array holding DL4J's 'f'-order [nIn,nOut] weight (W_f[i + j*nIn])."""
    x, wf = _f(x), _f(wf)
    B, nIn = x.shape
    nOut = wf.shape[0]
    z = np.empty((B, nOut), np.float32)
    lib().o_dense_fwd(_p(x), _p(wf), _p(None if b is None else _f(b)), _p(z), B, nIn, nOut)
    return z


def dense_bwd(x, wf, delta, need_dx=True):
    x, wf, delta = _f(x), _f(wf), _f(delta)
    B, nIn = x.shape
    nOut = wf.shape[0]
    dwf, db = np.empty_like(wf), np.empty(nOut, np.float32)
    dx = np.empty_like(x) if need_dx else None
    lib().o_dense_bwd(_p(x), _p(wf), _p(delta), _p(dwf), _p(db), _p(dx), B, nIn, nOut)
    return dwf, db, dx


def deconv_out_size(inp, k, s, p, d=1, mode=TRUNCATE):
    """ConvolutionUtils.getDeconvolutionOutputSize (util/ConvolutionUtils.java:63-83); Same mode is refused by the layer's
    forward pass (Deconvolution2DLayer.java:147-149)."""
    ke = k + (k - 1) * (d - 1)
    if mode == SAME:
        return s * inp
    return s * (inp - 1) + ke - 2 * p


def deconv2d_fwd(x, w, b, stride=(1, 1), pad=(0, 0), dilation=(1, 1)):
    """w: [nIn, nOut, kH, kW]."""
    x, w = _f(x), _f(w)
    B, nIn, H, W = x.shape
    _, nOut, kH, kW = w.shape
    Ho, Wo = deconv_out_size(H, kH, stride[0], pad[0], dilation[0]), deconv_out_size(W, kW, stride[1], pad[1], dilation[1])
    z = np.empty((B, nOut, Ho, Wo), np.float32)
    lib().o_deconv2d_fwd(_p(x), _p(w), _p(None if b is None else _f(b).reshape(-1)), _p(z), B, nIn, H, W, nOut, kH, kW, stride[0],
                         stride[1], pad[0], pad[1], dilation[0], dilation[1], Ho, Wo)
    return z


def deconv2d_bwd(x, w, delta, stride=(1, 1), pad=(0, 0), dilation=(1, 1)):
    x, w, delta = _f(x), _f(w), _f(delta)
    B, nIn, H, W = x.shape
    _, nOut, kH, kW = w.shape
    Ho, Wo = delta.shape[2:]
    dw, db, dx = np.empty_like(w), np.empty(nOut, np.float32), np.empty_like(x)
    lib().o_deconv2d_bwd(_p(x), _p(w), _p(delta), _p(dw), _p(db), _p(dx), B, nIn, H, W, nOut, kH, kW, stride[0], stride[1], pad[0],
                         pad[1], dilation[0], dilation[1], Ho, Wo)
    return dw, db, dx


def depthwise_fwd(x, wd, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    """wd: [depthMultiplier, nIn, kH, kW] -> u [B, nIn*mult, Ho, Wo] (channel ci*mult + m)."""
    x, wd = _f(x), _f(wd)
    B, Cc, H, W = x.shape
    mult, _, kH, kW = wd.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, (kH, kW), stride, pad, dilation, mode)
    u = np.empty((B, Cc * mult, Ho, Wo), np.float32)
    lib().o_depthwise_fwd(_p(x), _p(wd), _p(u), B, Cc, H, W, mult, kH, kW, stride[0], stride[1], pT, pL, dilation[0], dilation[1], Ho, Wo)
    return u


def depthwise_bwd(x, wd, du, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    x, wd, du = _f(x), _f(wd), _f(du)
    B, Cc, H, W = x.shape
    mult, _, kH, kW = wd.shape
    Ho, Wo, pT, pL = conv_geometry(H, W, (kH, kW), stride, pad, dilation, mode)
    dwd, dx = np.empty_like(wd), np.empty_like(x)
    lib().o_depthwise_bwd(_p(x), _p(wd), _p(du), _p(dwd), _p(dx), B, Cc, H, W, mult, kH, kW, stride[0], stride[1], pT, pL,
                          dilation[0], dilation[1], Ho, Wo)
    return dwd, dx


def sepconv2d_fwd(x, wd, wp, b, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    """SeparableConvolution2DLayer.preOutput: depth-wise then point-wise (wp: [nOut, nIn*mult, 1, 1]) + bias."""
    u = depthwise_fwd(x, wd, stride, pad, dilation, mode)
    return conv_fwd(u, wp, b), u


def sepconv2d_bwd(x, wd, wp, delta, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=TRUNCATE):
    """-> (dWd, dWp, db, dx)."""
    u = depthwise_fwd(x, wd, stride, pad, dilation, mode)
    dwp, db, du = conv_bwd(u, wp, delta)
    dwd, dx = depthwise_bwd(x, wd, du, stride, pad, dilation, mode)
    return dwd, dwp, db, dx
