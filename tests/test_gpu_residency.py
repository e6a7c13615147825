"""-m gpu: residency through the helper SPI. A convolution -> activation -> batch-norm -> max-pooling chain of helper calls
(the sequence MultiLayerNetwork.feedForward issues, MultiLayerNetwork.java:984-1021) and its backward chain must move
across the host link exactly: the minibatch once, the parameters, the epsilon that enters from the host side, the parameter
gradients and running statistics (the parameter boundary) -- and the final results when host code reads them. Counted with
dl4jb200_ctx_transfer_bytes; results checked against the oracle. This is what the reference's cuDNN helper gets from
nd4j-cuda's allocator (CudnnConvolutionHelper.java:442-447)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_helper_chain_keeps_activations_on_the_device(oracle):
    from deeplearning4j_b200 import _lib, helpers as H
    o = oracle
    ctx = H.Context(0)
    try:
        rng = np.random.default_rng(5)
        B, Cin, Hh, W, Cout = 8, 16, 20, 20, 32
        x = rng.random((B, Cin, Hh, W), dtype=np.float32)
        w = (rng.standard_normal((Cout, Cin, 3, 3)) * 0.2).astype(np.float32)
        b = rng.standard_normal(Cout).astype(np.float32)
        g, bt = (rng.random((1, Cout)) + 0.5).astype(np.float32), rng.standard_normal((1, Cout)).astype(np.float32)
        gm, gv = np.zeros((1, Cout), np.float32), np.ones((1, Cout), np.float32)
        conv = H.CudnnConvolutionHelper(ctx, algo=H.ALGO_FP32)
        bn, pool = H.CudnnBatchNormalizationHelper(ctx), H.CudnnSubsamplingHelper(ctx)
        xr = H.ResidentArray.from_host(ctx, x)
        assert xr.state() == _lib.BUF_HOST_NEWER
        h0, d0 = ctx.transfer_bytes()
        # ---- forward chain
        z = conv.preOutput(xr, w, b, (3, 3), (1, 1), (0, 0), None, None, "Same", (1, 1))
        a = conv.activate(z, "relu")
        y = bn.preOutput(a, True, list(a.shape), g, bt, gm, gv, 0.9, 1e-5)
        p = pool.activate(y, True, (2, 2), (2, 2), (0, 0), "MAX", "Truncate", (1, 1))
        h1, d1 = ctx.transfer_bytes()
        assert isinstance(p, H.ResidentArray) and a is z
        assert h1 - h0 == x.nbytes + w.nbytes + b.nbytes + 4 * Cout * 4, "forward: minibatch once + parameters, nothing else"
        assert d1 - d0 == 2 * Cout * 4, "forward: only the running statistics come back (they are parameters)"
        assert xr.state() == _lib.BUF_SYNCED and all(t.state() == _lib.BUF_DEVICE_NEWER for t in (z, y, p))
        # ---- backward chain; the epsilon enters from the host side (an output layer's delta, say)
        eps = rng.standard_normal(p.shape).astype(np.float32)
        _, e_y = pool.backpropGradient(y, eps, (2, 2), (2, 2), (0, 0), "MAX", "Truncate", (1, 1))
        dg, dbt = np.empty((1, Cout), np.float32), np.empty((1, Cout), np.float32)
        _, e_a = bn.backpropGradient(a, e_y, list(a.shape), g, dg, dbt, 1e-5)
        dw, db = np.empty_like(w), np.empty((1, Cout), np.float32)
        relu_mask = None
        h2, d2 = ctx.transfer_bytes()
        assert h2 - h1 == eps.nbytes + g.nbytes, "backward so far: the incoming epsilon and gamma"
        assert d2 - d1 == 2 * Cout * 4, "backward so far: dgamma and dbeta"
        # the layer applies afn.backprop on the host in the reference (ConvolutionLayer.java:171): host READS a and e_a here
        a_host, e_host = np.asarray(a), np.asarray(e_a)
        h3, d3 = ctx.transfer_bytes()
        assert d3 - d2 == a.nbytes + e_a.nbytes and h3 == h2, "host read: exactly the two arrays asked for"
        assert np.asarray(a) is a.numpy() and ctx.transfer_bytes() == (h3, d3), "a second read moves nothing"
        delta = (e_host * (a_host > 0)).astype(np.float32)
        _, e_x = conv.backpropGradient(xr, w, delta, (3, 3), (1, 1), (0, 0), db, dw, "identity", None, None, None, "Same", (1, 1))
        h4, d4 = ctx.transfer_bytes()
        assert h4 - h3 == delta.nbytes + w.nbytes, "conv backward: delta (host-made) + weights; x is still resident"
        assert d4 - d3 == dw.nbytes + Cout * 4, "conv backward: the parameter gradients"
        # ---- host writes a resident array: the next helper that reads it uploads exactly that array
        zh = z.numpy()
        zh *= 0.5
        z.mark_host_dirty()
        assert z.state() == _lib.BUF_HOST_NEWER
        h5, _ = ctx.transfer_bytes()
        pool.activate(z, False, (2, 2), (2, 2), (0, 0), "MAX", "Truncate", (1, 1))
        assert ctx.transfer_bytes()[0] - h5 == z.nbytes
        # ---- and the numbers are the reference's
        zr = o.conv_fwd(x, w, b, stride=(1, 1), pad=(0, 0), mode=o.SAME)
        ar = np.maximum(zr, 0)
        gm2, gv2 = np.zeros(Cout, np.float32), np.ones(Cout, np.float32)
        yr, mean, std = o.bn_fwd(ar, g.ravel(), bt.ravel(), gm2, gv2, 0.9, 1e-5, True, True)
        pr = o.pool_fwd(yr, (2, 2), (2, 2))
        assert np.abs(np.asarray(p) - pr).max() <= 1e-4 * np.abs(pr).max()
        assert np.abs(gm.ravel() - gm2).max() <= 1e-5 and np.abs(gv.ravel() - gv2).max() <= 1e-5
        eyr = o.pool_bwd(yr, eps, (2, 2), (2, 2))
        ear, dgr, dbr = o.bn_bwd(ar, eyr, g.ravel(), mean, std)
        assert np.abs(e_host - ear).max() <= 1e-4 * np.abs(ear).max()
        assert np.abs(dg.ravel() - dgr).max() <= 1e-4 * np.abs(dgr).max()
        dwr, dbr2, dxr = o.conv_bwd(x, w, (ear * (ar > 0)).astype(np.float32), stride=(1, 1), pad=(0, 0), mode=o.SAME)
        assert np.abs(dw - dwr).max() <= 2e-4 * np.abs(dwr).max()
        assert np.abs(np.asarray(e_x) - dxr).max() <= 2e-4 * np.abs(dxr).max()
        # ---- copy-in / copy-out mode still answers with plain host arrays
        ctx2 = H.Context(0, resident=False)
        z2 = H.CudnnConvolutionHelper(ctx2, algo=H.ALGO_FP32).preOutput(x, w, b, (3, 3), (1, 1), (0, 0), None, None, "Same", (1, 1))
        assert isinstance(z2, np.ndarray) and np.abs(z2 - zr).max() <= 1e-4 * np.abs(zr).max()
        ctx2.close()
    finally:
        ctx.close()
