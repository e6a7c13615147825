/*
 * dl4j_oracle.c -- TEST INFRASTRUCTURE ONLY (parity oracle + CPU baseline).
 *
 * A plain-C restatement of the arithmetic that Deeplearning4j's CNN layers
 * perform on the nd4j-native CPU path. Nothing under deeplearning4j_b200/
 * may include, link, import or call this file; only tests/, bench.py's
 * cpu_baseline / --impl reference leg and __graft_entry__.smoke() use it,
 * and only as the checker.
 *
 * The arithmetic itself lives in an un-vendored dependency of the reference
 * (org.nd4j:nd4j-native 1.0.0-SNAPSHOT, pom.xml:58; libnd4j C++), so each
 * routine follows the reference's *call sites* (cited per function, paths
 * relative to /root/reference/deeplearning4j-nn/src/main/java/org/deeplearning4j/)
 * and is pinned by the literal vectors in the reference's unit tests
 * (tests/golden/, see tests/test_oracle_golden.py).
 *
 * Conventions: all tensors are dense NCHW ('c' order) float32; every
 * reduction accumulates in double, so the oracle is the "exact" side of a
 * tolerance comparison. Where the reference runs im2col + gemm + col2im the
 * oracle offers both that literal pipeline (o_im2col / o_col2im, gemm done by
 * the caller with BLAS) and a direct-loop form; tests assert they agree.
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define IDX4(b, c, h, w, C, H, W) ((((size_t)(b) * (C) + (c)) * (H) + (h)) * (W) + (w))

/* ---- util/ConvolutionUtils.java:96-131 (getOutputSize), :213-221 (effectiveKernelSize),
 *      :243-266 (getSameModeTopLeftPadding / BottomRightPadding), :133-211 (validateShapes).
 * mode: 0 Strict, 1 Truncate, 2 Same. Returns 0 ok, 1 invalid input (kernel vs input),
 * 2 invalid config (Strict not divisible). Java int division truncates toward zero, as C does. */
int o_out_size(int in, int k, int s, int p, int d, int mode, int *out, int *pad_lo, int *pad_hi)
{
    int ke = (d == 1) ? k : k + (k - 1) * (d - 1);
    if (mode != 2 && (ke <= 0 || ke > in + 2 * p)) return 1;
    if (mode == 0 && (in - ke + 2 * p) % s != 0) return 2;
    if (mode == 2) {
        int o = (int)ceil(in / (double)s);
        *out = o;
        *pad_lo = ((o - 1) * s + ke - in) / 2;
        *pad_hi = ((o - 1) * s + ke - in + 1) / 2;
    } else {
        *out = (in - ke + 2 * p) / s + 1;
        *pad_lo = p;
        *pad_hi = p;
    }
    return 0;
}

/* ---- Convolution.im2col as called at nn/layers/convolution/ConvolutionLayer.java:385-388
 * (forward) and :213-216 (backward): col is allocated [B,Ho,Wo,C,kH,kW] 'c' and handed to
 * im2col permuted to [B,C,kH,kW,Ho,Wo]; so element (b,oh,ow,c,kh,kw) of the 'c' buffer holds
 * x[b,c,oh*sH-pT+kh*dH, ow*sW-pL+kw*dW] or padval when out of range. Row/column order of the
 * 2d reshape is pinned by ConvolutionLayerTest.java:468-546. padval is 0 for convolution and
 * -FLT_MAX for max-pool backprop (SubsamplingLayer.java:208-215, 270-281). */
void o_im2col(const float *x, float *col, int B, int C, int H, int W, int kH, int kW, int sH, int sW,
              int pT, int pL, int dH, int dW, int Ho, int Wo, float padval)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int oh = 0; oh < Ho; oh++)
            for (int ow = 0; ow < Wo; ow++) {
                float *dst = col + (((size_t)b * Ho + oh) * Wo + ow) * ((size_t)C * kH * kW);
                for (int c = 0; c < C; c++)
                    for (int kh = 0; kh < kH; kh++) {
                        int ih = oh * sH - pT + kh * dH;
                        for (int kw = 0; kw < kW; kw++) {
                            int iw = ow * sW - pL + kw * dW;
                            float v = padval;
                            if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = x[IDX4(b, c, ih, iw, C, H, W)];
                            *dst++ = v;
                        }
                    }
            }
}

/* ---- Convolution.col2im as called at ConvolutionLayer.java:245-253 and SubsamplingLayer.java:263:
 * scatter-add of col [B,Ho,Wo,C,kH,kW] back onto x-shaped [B,C,H,W]; cells falling in the padding
 * are dropped. (The reference's result buffer is [C,B,H,W] permuted; values are identical.) */
void o_col2im(const float *col, float *dx, int B, int C, int H, int W, int kH, int kW, int sH, int sW,
              int pT, int pL, int dH, int dW, int Ho, int Wo)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++) {
            float *plane = dx + IDX4(b, c, 0, 0, C, H, W);
            for (size_t i = 0; i < (size_t)H * W; i++) plane[i] = 0.f;
            for (int oh = 0; oh < Ho; oh++)
                for (int ow = 0; ow < Wo; ow++) {
                    const float *src = col + ((((size_t)b * Ho + oh) * Wo + ow) * C + c) * ((size_t)kH * kW);
                    for (int kh = 0; kh < kH; kh++) {
                        int ih = oh * sH - pT + kh * dH;
                        if (ih < 0 || ih >= H) continue;
                        for (int kw = 0; kw < kW; kw++) {
                            int iw = ow * sW - pL + kw * dW;
                            if (iw < 0 || iw >= W) continue;
                            plane[(size_t)ih * W + iw] += src[kh * kW + kw];
                        }
                    }
                }
        }
}

/* [B,Ho,Wo,Cout] 'c' (the z2d / delta2d row order) <-> NCHW. ConvolutionLayer.java:416-417, 202-206. */
void o_rows_to_nchw(const float *z2d, float *z, int B, int C, int Ho, int Wo)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++)
            for (int i = 0; i < Ho * Wo; i++)
                z[((size_t)b * C + c) * Ho * Wo + i] = z2d[((size_t)b * Ho * Wo + i) * C + c];
}
void o_nchw_to_rows(const float *z, float *z2d, int B, int C, int Ho, int Wo)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int i = 0; i < Ho * Wo; i++)
            for (int c = 0; c < C; c++)
                z2d[((size_t)b * Ho * Wo + i) * C + c] = z[((size_t)b * C + c) * Ho * Wo + i];
}

/* ---- ConvolutionLayer.preOutput, ConvolutionLayer.java:290-429 (CPU branch :385-417), direct form:
 * z[b,co,oh,ow] = bias[co] + sum_{ci,kh,kw} x[b,ci,oh*sH-pT+kh*dH, ow*sW-pL+kw*dW] * W[co,ci,kh,kw];
 * cross-correlation, zero padding, W 'c' order [Cout,Cin,kH,kW]. bias may be NULL (hasBias=false). */
void o_conv_fwd(const float *x, const float *w, const float *bias, float *z, int B, int Cin, int H, int W,
                int Cout, int kH, int kW, int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int co = 0; co < Cout; co++)
            for (int oh = 0; oh < Ho; oh++)
                for (int ow = 0; ow < Wo; ow++) {
                    double acc = bias ? (double)bias[co] : 0.0;
                    for (int ci = 0; ci < Cin; ci++)
                        for (int kh = 0; kh < kH; kh++) {
                            int ih = oh * sH - pT + kh * dH;
                            if (ih < 0 || ih >= H) continue;
                            for (int kw = 0; kw < kW; kw++) {
                                int iw = ow * sW - pL + kw * dW;
                                if (iw < 0 || iw >= W) continue;
                                acc += (double)x[IDX4(b, ci, ih, iw, Cin, H, W)] *
                                       (double)w[(((size_t)co * Cin + ci) * kH + kh) * kW + kw];
                            }
                        }
                    z[IDX4(b, co, oh, ow, Cout, Ho, Wo)] = (float)acc;
                }
}

/* ---- ConvolutionLayer.backpropGradient, ConvolutionLayer.java:131-265 (CPU branch :202-257), direct
 * form. delta = dL/dz (already multiplied by act'(z), :171). dW (:224, gemm beta=0 => overwritten),
 * db = delta2d.sum(1) (:257), dx = col2im(W2d^T . delta2d) (:232-253). Sums over the batch, unscaled. */
void o_conv_bwd(const float *x, const float *w, const float *delta, float *dw, float *db, float *dx, int B,
                int Cin, int H, int W, int Cout, int kH, int kW, int sH, int sW, int pT, int pL, int dH,
                int dW, int Ho, int Wo)
{
    if (dw) {
#pragma omp parallel for collapse(2) schedule(static)
        for (int co = 0; co < Cout; co++)
            for (int ci = 0; ci < Cin; ci++)
                for (int kh = 0; kh < kH; kh++)
                    for (int kw = 0; kw < kW; kw++) {
                        double acc = 0.0;
                        for (int b = 0; b < B; b++)
                            for (int oh = 0; oh < Ho; oh++) {
                                int ih = oh * sH - pT + kh * dH;
                                if (ih < 0 || ih >= H) continue;
                                for (int ow = 0; ow < Wo; ow++) {
                                    int iw = ow * sW - pL + kw * dW;
                                    if (iw < 0 || iw >= W) continue;
                                    acc += (double)delta[IDX4(b, co, oh, ow, Cout, Ho, Wo)] *
                                           (double)x[IDX4(b, ci, ih, iw, Cin, H, W)];
                                }
                            }
                        dw[(((size_t)co * Cin + ci) * kH + kh) * kW + kw] = (float)acc;
                    }
    }
    if (db) {
#pragma omp parallel for schedule(static)
        for (int co = 0; co < Cout; co++) {
            double acc = 0.0;
            for (int b = 0; b < B; b++)
                for (int i = 0; i < Ho * Wo; i++) acc += (double)delta[((size_t)b * Cout + co) * Ho * Wo + i];
            db[co] = (float)acc;
        }
    }
    if (dx) {
#pragma omp parallel for collapse(2) schedule(static)
        for (int b = 0; b < B; b++)
            for (int ci = 0; ci < Cin; ci++)
                for (int h = 0; h < H; h++)
                    for (int ww = 0; ww < W; ww++) {
                        double acc = 0.0;
                        for (int kh = 0; kh < kH; kh++) {
                            int t = h + pT - kh * dH;
                            if (t < 0 || t % sH) continue;
                            int oh = t / sH;
                            if (oh >= Ho) continue;
                            for (int kw = 0; kw < kW; kw++) {
                                int u = ww + pL - kw * dW;
                                if (u < 0 || u % sW) continue;
                                int ow = u / sW;
                                if (ow >= Wo) continue;
                                for (int co = 0; co < Cout; co++)
                                    acc += (double)delta[IDX4(b, co, oh, ow, Cout, Ho, Wo)] *
                                           (double)w[(((size_t)co * Cin + ci) * kH + kh) * kW + kw];
                            }
                        }
                        dx[IDX4(b, ci, h, ww, Cin, H, W)] = (float)acc;
                    }
    }
}

/* ---- SubsamplingLayer.activate, nn/layers/convolution/subsampling/SubsamplingLayer.java:285-363
 * (LegacyPooling2D, :339-360). type: 0 MAX, 1 AVG, 2 SUM, 3 PNORM.
 * MAX ignores padding cells (they act as -FLT_MAX); AVG counts padding as 0 and always divides by
 * kH*kW (extra=1.0, :347-350). argmax (optional, MAX only) = kh*kW+kw of the FIRST maximum in
 * row-major window order, the tie-break pinned by SubsamplingLayerTest.java:85-101. */
void o_pool_fwd(const float *x, float *y, int32_t *argmax, int B, int C, int H, int W, int kH, int kW,
                int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo, int type, int pnorm)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++)
            for (int oh = 0; oh < Ho; oh++)
                for (int ow = 0; ow < Wo; ow++) {
                    double acc = 0.0;
                    float best = -FLT_MAX;
                    int besti = 0;
                    for (int kh = 0; kh < kH; kh++)
                        for (int kw = 0; kw < kW; kw++) {
                            int ih = oh * sH - pT + kh * dH, iw = ow * sW - pL + kw * dW;
                            int in = (ih >= 0 && ih < H && iw >= 0 && iw < W);
                            float v = in ? x[IDX4(b, c, ih, iw, C, H, W)] : (type == 0 ? -FLT_MAX : 0.f);
                            if (type == 0) {
                                if (v > best) { best = v; besti = kh * kW + kw; }
                            } else if (type == 3) {
                                acc += pow(fabs((double)v), (double)pnorm);
                            } else {
                                acc += v;
                            }
                        }
                    size_t o = IDX4(b, c, oh, ow, C, Ho, Wo);
                    if (type == 0) { y[o] = best; if (argmax) argmax[o] = besti; }
                    else if (type == 1) y[o] = (float)(acc / (double)(kH * kW));
                    else if (type == 2) y[o] = (float)acc;
                    else y[o] = (float)pow(acc, 1.0 / (double)pnorm);
                }
}

/* ---- SubsamplingLayer.backpropGradient, SubsamplingLayer.java:112-268.
 * MAX (:205-218, 263): im2col with -FLT_MAX padding, IsMax along the window (first max), times eps,
 * col2im (overlapping windows accumulate). AVG (:223, 263-266): eps broadcast into every window cell,
 * col2im, then / (kH*kW). PNORM (:225-250): dx += eps * x|x|^(p-2) / max(y^(p-1), eps_p). SUM throws
 * (:251-253) -> returns 1. */
int o_pool_bwd(const float *x, const float *eps, float *dx, int B, int C, int H, int W, int kH, int kW,
               int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo, int type, int pnorm,
               double pnorm_eps)
{
    if (type == 2) return 1;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++) {
            float *plane = dx + IDX4(b, c, 0, 0, C, H, W);
            const float *xp = x + IDX4(b, c, 0, 0, C, H, W);
            double *acc = (double *)calloc((size_t)H * W, sizeof(double));
            for (int oh = 0; oh < Ho; oh++)
                for (int ow = 0; ow < Wo; ow++) {
                    double e = eps[IDX4(b, c, oh, ow, C, Ho, Wo)];
                    if (type == 0) {
                        float best = -FLT_MAX;
                        int bh = -1, bw = -1, first = 1;
                        for (int kh = 0; kh < kH; kh++)
                            for (int kw = 0; kw < kW; kw++) {
                                int ih = oh * sH - pT + kh * dH, iw = ow * sW - pL + kw * dW;
                                int in = (ih >= 0 && ih < H && iw >= 0 && iw < W);
                                float v = in ? xp[(size_t)ih * W + iw] : -FLT_MAX;
                                if (first || v > best) { best = v; bh = in ? ih : -1; bw = iw; first = 0; }
                            }
                        if (bh >= 0) acc[(size_t)bh * W + bw] += e;
                    } else if (type == 1) {
                        for (int kh = 0; kh < kH; kh++)
                            for (int kw = 0; kw < kW; kw++) {
                                int ih = oh * sH - pT + kh * dH, iw = ow * sW - pL + kw * dW;
                                if (ih >= 0 && ih < H && iw >= 0 && iw < W)
                                    acc[(size_t)ih * W + iw] += e / (double)(kH * kW);
                            }
                    } else {
                        double s = 0.0;
                        for (int kh = 0; kh < kH; kh++)
                            for (int kw = 0; kw < kW; kw++) {
                                int ih = oh * sH - pT + kh * dH, iw = ow * sW - pL + kw * dW;
                                if (ih >= 0 && ih < H && iw >= 0 && iw < W)
                                    s += pow(fabs((double)xp[(size_t)ih * W + iw]), (double)pnorm);
                            }
                        double yv = pow(s, 1.0 / (double)pnorm);
                        double denom = pow(yv, (double)(pnorm - 1));
                        if (denom < pnorm_eps) denom = pnorm_eps;
                        for (int kh = 0; kh < kH; kh++)
                            for (int kw = 0; kw < kW; kw++) {
                                int ih = oh * sH - pT + kh * dH, iw = ow * sW - pL + kw * dW;
                                if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
                                    double xv = xp[(size_t)ih * W + iw];
                                    acc[(size_t)ih * W + iw] += e * xv * pow(fabs(xv), (double)(pnorm - 2)) / denom;
                                }
                            }
                    }
                }
            for (size_t i = 0; i < (size_t)H * W; i++) plane[i] = (float)acc[i];
            free(acc);
        }
    return 0;
}

/* ---- BatchNormalization.preOutput, nn/layers/normalization/BatchNormalization.java:239-395, rank 4.
 * training!=0: mean/var over (B,H,W) per channel, var biased (:262-266); v = var+eps stored IN PLACE
 * (:273); std = sqrt(v) (:279); y = gamma*(x-mean)/std + beta (:330-355); running stats
 * gMean = decay*gMean + (1-decay)*mean, gVar = decay*gVar + (1-decay)*v  -- v INCLUDES eps (:377-384);
 * minibatch==0 => plain assignment (:386-390). training==0: std = sqrt(gVar), NO eps (:274-279).
 * save_mean / save_std (optional) receive the per-channel mean and std the layer keeps for backward. */
void o_bn_fwd(const float *x, float *y, const float *gamma, const float *beta, float *gmean, float *gvar,
              float *save_mean, float *save_std, int B, int C, int H, int W, double decay, double eps,
              int training, int minibatch)
{
    size_t HW = (size_t)H * W;
    double N = (double)B * (double)HW;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < C; c++) {
        double mean, std;
        if (training) {
            double s = 0.0;
            for (int b = 0; b < B; b++) {
                const float *p = x + ((size_t)b * C + c) * HW;
                for (size_t i = 0; i < HW; i++) s += p[i];
            }
            mean = s / N;
            double q = 0.0;
            for (int b = 0; b < B; b++) {
                const float *p = x + ((size_t)b * C + c) * HW;
                for (size_t i = 0; i < HW; i++) { double d = p[i] - mean; q += d * d; }
            }
            double v = q / N + eps;
            std = sqrt(v);
            if (minibatch) {
                gmean[c] = (float)(decay * gmean[c] + (1.0 - decay) * mean);
                gvar[c] = (float)(decay * gvar[c] + (1.0 - decay) * v);
            } else {
                gmean[c] = (float)mean;
                gvar[c] = (float)v;
            }
        } else {
            mean = gmean[c];
            std = sqrt((double)gvar[c]);
        }
        if (save_mean) save_mean[c] = (float)mean;
        if (save_std) save_std[c] = (float)std;
        double g = gamma ? gamma[c] : 1.0, bt = beta ? beta[c] : 0.0;
        for (int b = 0; b < B; b++) {
            const float *p = x + ((size_t)b * C + c) * HW;
            float *o = y + ((size_t)b * C + c) * HW;
            for (size_t i = 0; i < HW; i++) o[i] = (float)(g * ((p[i] - mean) / std) + bt);
        }
    }
}

/* ---- BatchNormalization.backpropGradient, BatchNormalization.java:93-219 (rank-4 branch :171-210).
 * mean/std are the values of the preceding training forward (xMu, xHat, std are layer state, :46-48).
 * Keeps the (sum xMu) term of dMu (:191) exactly as written. */
void o_bn_bwd(const float *x, const float *eps_in, const float *gamma, const float *mean_in, const float *std_in,
              float *dx, float *dgamma, float *dbeta, int B, int C, int H, int W)
{
    size_t HW = (size_t)H * W;
    double N = (double)B * (double)HW;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < C; c++) {
        double mean = mean_in[c], std = std_in[c], g = gamma ? gamma[c] : 1.0;
        double sum_e = 0, sum_exhat = 0, sum_dxhat_xmu = 0, sum_dxhat = 0, sum_xmu = 0;
        for (int b = 0; b < B; b++) {
            const float *p = x + ((size_t)b * C + c) * HW;
            const float *e = eps_in + ((size_t)b * C + c) * HW;
            for (size_t i = 0; i < HW; i++) {
                double xmu = p[i] - mean, xhat = xmu / std, dxhat = e[i] * g;
                sum_e += e[i];
                sum_exhat += e[i] * xhat;
                sum_dxhat_xmu += dxhat * xmu;
                sum_dxhat += dxhat;
                sum_xmu += xmu;
            }
        }
        double dvar = sum_dxhat_xmu * (-0.5) * pow(std, -3.0);
        double dmu = -sum_dxhat / std + sum_xmu * (-2.0 / N) * dvar;
        for (int b = 0; b < B; b++) {
            const float *p = x + ((size_t)b * C + c) * HW;
            const float *e = eps_in + ((size_t)b * C + c) * HW;
            float *o = dx + ((size_t)b * C + c) * HW;
            for (size_t i = 0; i < HW; i++) {
                double xmu = p[i] - mean, dxhat = e[i] * g;
                o[i] = (float)(dxhat / std + xmu * dvar * (2.0 / N) + dmu / N);
            }
        }
        if (dgamma) dgamma[c] = (float)sum_exhat;
        if (dbeta) dbeta[c] = (float)sum_e;
    }
}

/* ---- LocalResponseNormalization.activate, nn/layers/normalization/LocalResponseNormalization.java:150-190.
 * halfN = (int)n/2 (:155); sumPart_c = sum_{j=max(0,c-h)}^{min(C-1,c+h)} x_j^2; unitScale = k+alpha*sumPart
 * (alpha is NOT divided by n); scale = unitScale^-beta; y = x*scale. Pinned by LocalResponseTest.java:36-60
 * and the scalar loop :153-194. scale_out/unit_out optional (kept by the layer for backward). */
void o_lrn_fwd(const float *x, float *y, float *unit_out, float *scale_out, int B, int C, int H, int W, double k,
               double n, double alpha, double beta)
{
    int h = (int)n / 2;
    size_t HW = (size_t)H * W;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++) {
            int j0 = c - h < 0 ? 0 : c - h, j1 = c + h > C - 1 ? C - 1 : c + h;
            for (size_t i = 0; i < HW; i++) {
                double s = 0.0;
                for (int j = j0; j <= j1; j++) { double v = x[((size_t)b * C + j) * HW + i]; s += v * v; }
                double u = k + alpha * s, sc = pow(u, -beta);
                size_t o = ((size_t)b * C + c) * HW + i;
                y[o] = (float)(x[o] * sc);
                if (unit_out) unit_out[o] = (float)u;
                if (scale_out) scale_out[o] = (float)sc;
            }
        }
}

/* ---- LocalResponseNormalization.backpropGradient, LocalResponseNormalization.java:117-147:
 * dx = eps*scale - 2*alpha*beta * x * (sum_window y_j*eps_j) / unitScale. LocalResponseTest.java:62-117. */
void o_lrn_bwd(const float *x, const float *eps_in, float *dx, int B, int C, int H, int W, double k, double n,
               double alpha, double beta)
{
    int h = (int)n / 2;
    size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; b++) {
        double *u = (double *)malloc(sizeof(double) * C), *r = (double *)malloc(sizeof(double) * C);
        for (size_t i = 0; i < HW; i++) {
            for (int c = 0; c < C; c++) {
                int j0 = c - h < 0 ? 0 : c - h, j1 = c + h > C - 1 ? C - 1 : c + h;
                double s = 0.0;
                for (int j = j0; j <= j1; j++) { double v = x[((size_t)b * C + j) * HW + i]; s += v * v; }
                u[c] = k + alpha * s;
                size_t o = ((size_t)b * C + c) * HW + i;
                r[c] = (double)x[o] * pow(u[c], -beta) * (double)eps_in[o]; /* y * eps */
            }
            for (int c = 0; c < C; c++) {
                int j0 = c - h < 0 ? 0 : c - h, j1 = c + h > C - 1 ? C - 1 : c + h;
                double R = 0.0;
                for (int j = j0; j <= j1; j++) R += r[j];
                size_t o = ((size_t)b * C + c) * HW + i;
                dx[o] = (float)((double)eps_in[o] * pow(u[c], -beta) - 2.0 * alpha * beta * (double)x[o] * R / u[c]);
            }
        }
        free(u);
        free(r);
    }
}

/* ---- IActivation.getActivation / backprop as used at ConvolutionLayer.java:463 / :171 and by
 * ActivationLayer. kind: 0 IDENTITY, 1 RELU, 2 SIGMOID, 3 TANH, 4 LEAKYRELU(0.01). In place allowed. */
static double act_f(double z, int kind)
{
    switch (kind) {
    case 1: return z > 0 ? z : 0;
    case 2: return 1.0 / (1.0 + exp(-z));
    case 3: return tanh(z);
    case 4: return z > 0 ? z : 0.01 * z;
    default: return z;
    }
}
static double act_df(double z, int kind)
{
    switch (kind) {
    case 1: return z > 0 ? 1.0 : 0.0;
    case 2: { double s = 1.0 / (1.0 + exp(-z)); return s * (1 - s); }
    case 3: { double t = tanh(z); return 1 - t * t; }
    case 4: return z > 0 ? 1.0 : 0.01;
    default: return 1.0;
    }
}
void o_act_fwd(const float *z, float *a, size_t n, int kind)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) a[i] = (float)act_f(z[i], kind);
}
void o_act_bwd(const float *z, const float *eps_in, float *delta, size_t n, int kind)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) delta[i] = (float)(act_df(z[i], kind) * (double)eps_in[i]);
}

/* ---- BaseOutputLayer with softmax + NEGATIVELOGLIKELIHOOD / MCXENT (identical for one-hot labels),
 * nn/layers/BaseOutputLayer.java:82-100,147-190: p = softmax(z) row-wise; loss_sum = -sum labels*log(p);
 * delta = p - labels (gradient wrt pre-output, summed not averaged). Returns loss_sum. */
double o_softmax_nll(const float *z, const float *labels, float *p_out, float *delta, int B, int K)
{
    double loss = 0.0;
    for (int b = 0; b < B; b++) {
        const float *zr = z + (size_t)b * K;
        double m = zr[0];
        for (int k = 1; k < K; k++) if (zr[k] > m) m = zr[k];
        double s = 0.0;
        for (int k = 0; k < K; k++) s += exp(zr[k] - m);
        for (int k = 0; k < K; k++) {
            double p = exp(zr[k] - m) / s, l = labels[(size_t)b * K + k];
            if (p_out) p_out[(size_t)b * K + k] = (float)p;
            if (delta) delta[(size_t)b * K + k] = (float)(p - l);
            if (l != 0.0) loss -= l * log(p > 1e-300 ? p : 1e-300);
        }
    }
    return loss;
}

/* ---- Updaters (formulas pinned by deeplearning4j-core/src/test/java/org/deeplearning4j/nn/updater/
 * TestUpdaters.java:62-131 AdaDelta, :386-428 Nesterovs, :431-478 RmsProp); order of operations per
 * nn/updater/UpdaterBlock.java:115-193 and BaseMultiLayerUpdater.java:223-309:
 *   u = Updater(g) ; u += l2*w ; u += l1*sign(w) ; u /= batch (if miniBatch) ; w -= u.
 * kind: 0 SGD(lr), 1 NESTEROVS(lr, momentum=h1; state v), 2 RMSPROP(lr, decay=h1, eps=h2; state c),
 * 3 ADADELTA(rho=h1, eps=h2; state msg|msdx, 2n floats), 4 NOOP (u = 0: BN running stats). */
void o_updater_apply(float *w, const float *g, float *state, size_t n, int kind, double lr, double h1, double h2,
                     double l1, double l2, double batch_div)
{
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) {
        double gi = g[i], u;
        switch (kind) {
        case 1: {
            double v = state[i], vn = h1 * v - lr * gi;
            u = h1 * v - (1.0 + h1) * vn;
            state[i] = (float)vn;
        } break;
        case 2: {
            double c = h1 * state[i] + (1.0 - h1) * gi * gi;
            state[i] = (float)c;
            u = lr * gi / sqrt(c + h2);
        } break;
        case 3: {
            double msg = h1 * state[i] + (1.0 - h1) * gi * gi;
            double msdx = state[n + i];
            u = gi * sqrt(msdx + h2) / sqrt(msg + h2);
            state[i] = (float)msg;
            state[n + i] = (float)(h1 * msdx + (1.0 - h1) * u * u);
        } break;
        case 4: u = 0.0; break;
        default: u = lr * gi;
        }
        double wi = w[i];
        if (l2 > 0) u += l2 * wi;
        if (l1 > 0) u += l1 * (wi > 0 ? 1.0 : (wi < 0 ? -1.0 : 0.0));
        u /= batch_div;
        w[i] = (float)(wi - u);
    }
}

/* Dense pre-output z = x.W + b with W [nIn,nOut] 'f' order (nn/layers/BaseLayer.java:315-344;
 * DefaultParamInitializer reshape 'f'), i.e. W_f[i + j*nIn]. Direct double-accumulate form. */
void o_dense_fwd(const float *x, const float *wf, const float *b, float *z, int B, int nIn, int nOut)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int r = 0; r < B; r++)
        for (int j = 0; j < nOut; j++) {
            double acc = b ? b[j] : 0.0;
            for (int i = 0; i < nIn; i++) acc += (double)x[(size_t)r * nIn + i] * (double)wf[(size_t)j * nIn + i];
            z[(size_t)r * nOut + j] = (float)acc;
        }
}
/* BaseLayer.backpropGradient, BaseLayer.java:71-102: dW = x^T.delta ('f' view), db = sum_rows delta,
 * eps_next = delta.W^T. */
void o_dense_bwd(const float *x, const float *wf, const float *delta, float *dwf, float *db, float *dx, int B,
                 int nIn, int nOut)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int j = 0; j < nOut; j++)
        for (int i = 0; i < nIn; i++) {
            double acc = 0.0;
            for (int r = 0; r < B; r++) acc += (double)x[(size_t)r * nIn + i] * (double)delta[(size_t)r * nOut + j];
            dwf[(size_t)j * nIn + i] = (float)acc;
        }
    if (db)
        for (int j = 0; j < nOut; j++) {
            double acc = 0.0;
            for (int r = 0; r < B; r++) acc += delta[(size_t)r * nOut + j];
            db[j] = (float)acc;
        }
    if (dx) {
#pragma omp parallel for collapse(2) schedule(static)
        for (int r = 0; r < B; r++)
            for (int i = 0; i < nIn; i++) {
                double acc = 0.0;
                for (int j = 0; j < nOut; j++) acc += (double)delta[(size_t)r * nOut + j] * (double)wf[(size_t)j * nIn + i];
                dx[(size_t)r * nIn + i] = (float)acc;
            }
    }
}

/* ---- Deconvolution2DLayer (nn/layers/convolution/Deconvolution2DLayer.java:145-233 forward, :48-138 backward). The layer
 * delegates the arithmetic to libnd4j's "deconv2d" / "deconv2d_bp" custom ops (un-vendored; restated here from the
 * published definition of a transposed convolution that those ops implement): W is [nIn, nOut, kH, kW] 'c'
 * (params/DeconvolutionParamInitializer.java:75-80), output size s*(in-1) + kEff - 2p (util/ConvolutionUtils.java:79-80),
 *   z[b,co,oh,ow] = bias[co] + sum_{ci,kh,kw} x[b,ci,ih,iw] * W[ci,co,kh,kw],  oh = ih*sH - pT + kh*dH (likewise ow)
 * PARITY UNPINNED for these two ops: the reference holds no literal vectors for them (gradient checks only,
 * gradientcheck/CNNGradientCheckTest.java); tests pin forward against backward by a numerical gradient check. */
void o_deconv2d_fwd(const float *x, const float *w, const float *bias, float *z, int B, int nIn, int H, int W, int nOut,
                    int kH, int kW, int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo)
{
    size_t zn = (size_t)B * nOut * Ho * Wo;
    double *acc = (double *)calloc(zn, sizeof(double));
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; b++)
        for (int ci = 0; ci < nIn; ci++)
            for (int ih = 0; ih < H; ih++)
                for (int iw = 0; iw < W; iw++) {
                    double xv = x[(((size_t)b * nIn + ci) * H + ih) * W + iw];
                    for (int co = 0; co < nOut; co++)
                        for (int kh = 0; kh < kH; kh++) {
                            int oh = ih * sH - pT + kh * dH;
                            if (oh < 0 || oh >= Ho) continue;
                            for (int kw = 0; kw < kW; kw++) {
                                int ow = iw * sW - pL + kw * dW;
                                if (ow < 0 || ow >= Wo) continue;
                                acc[(((size_t)b * nOut + co) * Ho + oh) * Wo + ow] += xv * w[(((size_t)ci * nOut + co) * kH + kh) * kW + kw];
                            }
                        }
                }
    for (size_t i = 0; i < zn; i++) {
        int co = (int)((i / ((size_t)Ho * Wo)) % nOut);
        z[i] = (float)(acc[i] + (bias ? (double)bias[co] : 0.0));
    }
    free(acc);
}

/* dW[ci,co,kh,kw] = sum x[b,ci,ih,iw] * delta[b,co,oh,ow]; db[co] = sum delta; dx[b,ci,ih,iw] = sum delta[b,co,oh,ow] * W[ci,co,kh,kw] */
void o_deconv2d_bwd(const float *x, const float *w, const float *delta, float *dw, float *db, float *dx, int B, int nIn,
                    int H, int W, int nOut, int kH, int kW, int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo)
{
#pragma omp parallel for schedule(static)
    for (int ci = 0; ci < nIn; ci++)
        for (int co = 0; co < nOut; co++)
            for (int kh = 0; kh < kH; kh++)
                for (int kw = 0; kw < kW; kw++) {
                    double a = 0.0;
                    for (int b = 0; b < B; b++)
                        for (int ih = 0; ih < H; ih++) {
                            int oh = ih * sH - pT + kh * dH;
                            if (oh < 0 || oh >= Ho) continue;
                            for (int iw = 0; iw < W; iw++) {
                                int ow = iw * sW - pL + kw * dW;
                                if (ow < 0 || ow >= Wo) continue;
                                a += (double)x[(((size_t)b * nIn + ci) * H + ih) * W + iw] *
                                     (double)delta[(((size_t)b * nOut + co) * Ho + oh) * Wo + ow];
                            }
                        }
                    dw[(((size_t)ci * nOut + co) * kH + kh) * kW + kw] = (float)a;
                }
    if (db)
        for (int co = 0; co < nOut; co++) {
            double a = 0.0;
            for (int b = 0; b < B; b++)
                for (int i = 0; i < Ho * Wo; i++) a += delta[((size_t)b * nOut + co) * Ho * Wo + i];
            db[co] = (float)a;
        }
    if (dx) {
#pragma omp parallel for schedule(static)
        for (int b = 0; b < B; b++)
            for (int ci = 0; ci < nIn; ci++)
                for (int ih = 0; ih < H; ih++)
                    for (int iw = 0; iw < W; iw++) {
                        double a = 0.0;
                        for (int co = 0; co < nOut; co++)
                            for (int kh = 0; kh < kH; kh++) {
                                int oh = ih * sH - pT + kh * dH;
                                if (oh < 0 || oh >= Ho) continue;
                                for (int kw = 0; kw < kW; kw++) {
                                    int ow = iw * sW - pL + kw * dW;
                                    if (ow < 0 || ow >= Wo) continue;
                                    a += (double)delta[(((size_t)b * nOut + co) * Ho + oh) * Wo + ow] *
                                         (double)w[(((size_t)ci * nOut + co) * kH + kh) * kW + kw];
                                }
                            }
                        dx[(((size_t)b * nIn + ci) * H + ih) * W + iw] = (float)a;
                    }
    }
}

/* ---- SeparableConvolution2DLayer (nn/layers/convolution/SeparableConvolution2DLayer.java:160-246 forward, :60-152
 * backward; libnd4j "sconv2d" / "sconv2d_bp", un-vendored, restated from their definition): depth-wise filter
 * Wd [depthMultiplier, nIn, kH, kW], point-wise filter Wp [nOut, nIn*depthMultiplier, 1, 1]
 * (params/SeparableConvolutionParamInitializer.java); intermediate channel index = ci * depthMultiplier + m.
 *   u[b, ci*mult+m, oh, ow] = sum_{kh,kw} x[b,ci, oh*sH-pT+kh*dH, ow*sW-pL+kw*dW] * Wd[m,ci,kh,kw]
 *   z[b, co, oh, ow]        = bias[co] + sum_j u[b,j,oh,ow] * Wp[co,j]                                  (unpinned, see above) */
void o_depthwise_fwd(const float *x, const float *wd, float *u, int B, int C, int H, int W, int mult, int kH, int kW, int sH,
                     int sW, int pT, int pL, int dH, int dW, int Ho, int Wo)
{
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; b++)
        for (int ci = 0; ci < C; ci++)
            for (int m = 0; m < mult; m++)
                for (int oh = 0; oh < Ho; oh++)
                    for (int ow = 0; ow < Wo; ow++) {
                        double a = 0.0;
                        for (int kh = 0; kh < kH; kh++) {
                            int ih = oh * sH - pT + kh * dH;
                            if (ih < 0 || ih >= H) continue;
                            for (int kw = 0; kw < kW; kw++) {
                                int iw = ow * sW - pL + kw * dW;
                                if (iw < 0 || iw >= W) continue;
                                a += (double)x[(((size_t)b * C + ci) * H + ih) * W + iw] * (double)wd[(((size_t)m * C + ci) * kH + kh) * kW + kw];
                            }
                        }
                        u[(((size_t)b * C * mult + ci * mult + m) * Ho + oh) * Wo + ow] = (float)a;
                    }
}

/* du: epsilon w.r.t. u. dwd[m,ci,kh,kw] = sum x * du; dx[b,ci,ih,iw] = sum_m,kh,kw du[b,ci*mult+m,oh,ow] * Wd[m,ci,kh,kw] */
void o_depthwise_bwd(const float *x, const float *wd, const float *du, float *dwd, float *dx, int B, int C, int H, int W,
                     int mult, int kH, int kW, int sH, int sW, int pT, int pL, int dH, int dW, int Ho, int Wo)
{
#pragma omp parallel for schedule(static)
    for (int ci = 0; ci < C; ci++)
        for (int m = 0; m < mult; m++)
            for (int kh = 0; kh < kH; kh++)
                for (int kw = 0; kw < kW; kw++) {
                    double a = 0.0;
                    for (int b = 0; b < B; b++)
                        for (int oh = 0; oh < Ho; oh++) {
                            int ih = oh * sH - pT + kh * dH;
                            if (ih < 0 || ih >= H) continue;
                            for (int ow = 0; ow < Wo; ow++) {
                                int iw = ow * sW - pL + kw * dW;
                                if (iw < 0 || iw >= W) continue;
                                a += (double)x[(((size_t)b * C + ci) * H + ih) * W + iw] *
                                     (double)du[(((size_t)b * C * mult + ci * mult + m) * Ho + oh) * Wo + ow];
                            }
                        }
                    dwd[(((size_t)m * C + ci) * kH + kh) * kW + kw] = (float)a;
                }
    if (dx) {
        size_t n = (size_t)B * C * H * W;
        double *acc = (double *)calloc(n, sizeof(double));
#pragma omp parallel for schedule(static)
        for (int b = 0; b < B; b++)
            for (int ci = 0; ci < C; ci++)
                for (int m = 0; m < mult; m++)
                    for (int oh = 0; oh < Ho; oh++)
                        for (int ow = 0; ow < Wo; ow++) {
                            double g = du[(((size_t)b * C * mult + ci * mult + m) * Ho + oh) * Wo + ow];
                            for (int kh = 0; kh < kH; kh++) {
                                int ih = oh * sH - pT + kh * dH;
                                if (ih < 0 || ih >= H) continue;
                                for (int kw = 0; kw < kW; kw++) {
                                    int iw = ow * sW - pL + kw * dW;
                                    if (iw < 0 || iw >= W) continue;
                                    acc[(((size_t)b * C + ci) * H + ih) * W + iw] += g * (double)wd[(((size_t)m * C + ci) * kH + kh) * kW + kw];
                                }
                            }
                        }
        for (size_t i = 0; i < n; i++) dx[i] = (float)acc[i];
        free(acc);
    }
}
