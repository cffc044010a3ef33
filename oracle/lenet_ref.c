/*
 * Plain-C restatement of the examples/mnist training step.  TEST INFRASTRUCTURE ONLY:
 * built into oracle/liblenet_ref.so and called from tests/ (and from
 * __graft_entry__.smoke()) as the checker for the CUDA path; the product never links
 * or loads it.
 *
 * The reference's arithmetic lives in PyTorch (third party, not under /root/reference;
 * image pin pytorch/pytorch:1.0-cuda10.0-cudnn7-runtime, examples/mnist/Dockerfile:1).
 * Each function restates the published definition of the ATen op the reference calls
 * at the cited line of /root/reference/examples/mnist/mnist.py.  Every dot product is
 * accumulated in double and rounded to float once, so this file is the "exact" side
 * of an fp32 comparison; it is itself pinned against tests/golden/ (outputs of the
 * reference run in the build container, see oracle/gen_golden.py) by tests/test_oracle.py.
 *
 * Layouts are the reference's: NCHW activations, [out,in,kh,kw] conv weights,
 * [out,in] linear weights, int64 targets.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* conv5x5 (valid, stride 1) + bias + ReLU + 2x2/2 max-pool: mnist.py:26-27 and :28-29.
 * out[b,oc,py,px] = max over the window of relu(conv); idx = window slot (dy*2+dx) of
 * the first maximum in scan order, which is where max_pool2d's backward routes to. */
static void conv5_relu_pool(const float *in, const float *w, const float *bias, int B, int IC, int H,
                            int OC, float *out, uint8_t *idx) {
    const int OH = H - 4, PH = OH / 2;
    for (int b = 0; b < B; ++b)
        for (int oc = 0; oc < OC; ++oc)
            for (int py = 0; py < PH; ++py)
                for (int px = 0; px < PH; ++px) {
                    float best = 0.f;
                    int bi = 0;
                    for (int q = 0; q < 4; ++q) {
                        const int oy = 2 * py + (q >> 1), ox = 2 * px + (q & 1);
                        double acc = bias[oc];
                        for (int ic = 0; ic < IC; ++ic)
                            for (int ky = 0; ky < 5; ++ky)
                                for (int kx = 0; kx < 5; ++kx)
                                    acc += (double)w[((oc * IC + ic) * 5 + ky) * 5 + kx] *
                                           (double)in[((b * IC + ic) * H + oy + ky) * H + ox + kx];
                        float v = (float)acc;
                        v = v > 0.f ? v : 0.f; /* F.relu */
                        if (q == 0 || v > best) { best = v; bi = q; }
                    }
                    const int o = ((b * OC + oc) * PH + py) * PH + px;
                    out[o] = best;
                    idx[o] = (uint8_t)bi;
                }
}

/* y = x W^T + b, optional ReLU: mnist.py:31 (fc1 + relu) and :32 (fc2). */
static void linear(const float *x, const float *w, const float *bias, int B, int K, int N, int relu, float *y) {
    for (int b = 0; b < B; ++b)
        for (int n = 0; n < N; ++n) {
            double acc = bias[n];
            for (int k = 0; k < K; ++k) acc += (double)x[b * K + k] * (double)w[n * K + k];
            float v = (float)acc;
            y[b * N + n] = (relu && v < 0.f) ? 0.f : v;
        }
}

/* Net.forward, mnist.py:25-33.  Writes every saved activation the backward needs. */
void lenet_forward(const float *c1w, const float *c1b, const float *c2w, const float *c2b, const float *f1w,
                   const float *f1b, const float *f2w, const float *f2b, const float *x, int B, float *p1,
                   uint8_t *i1, float *p2, uint8_t *i2, float *h, float *logp) {
    conv5_relu_pool(x, c1w, c1b, B, 1, 28, 20, p1, i1);
    conv5_relu_pool(p1, c2w, c2b, B, 20, 12, 50, p2, i2);
    linear(p2, f1w, f1b, B, 800, 500, 1, h); /* x.view(-1, 800): NCHW flatten is the identity */
    float *z = (float *)malloc(sizeof(float) * B * 10);
    linear(h, f2w, f2b, B, 500, 10, 0, z);
    for (int b = 0; b < B; ++b) { /* F.log_softmax(dim=1), mnist.py:33 */
        float m = z[b * 10];
        for (int j = 1; j < 10; ++j) m = z[b * 10 + j] > m ? z[b * 10 + j] : m;
        double s = 0.0;
        for (int j = 0; j < 10; ++j) s += exp((double)z[b * 10 + j] - (double)m);
        const double lse = (double)m + log(s);
        for (int j = 0; j < 10; ++j) logp[b * 10 + j] = (float)((double)z[b * 10 + j] - lse);
    }
    free(z);
}

/* F.nll_loss(reduction='mean'), mnist.py:41. */
double lenet_nll_mean(const float *logp, const int64_t *target, int B) {
    double s = 0.0;
    for (int b = 0; b < B; ++b) s -= (double)logp[b * 10 + target[b]];
    return s / B;
}

/* loss.backward(), mnist.py:42: the reverse of lenet_forward + nll mean.  Gradients are
 * written (not accumulated) into the eight g* arrays, shaped like the parameters. */
void lenet_backward(const float *c2w, const float *f1w, const float *f2w, const float *x, const int64_t *target,
                    int B, const float *p1, const uint8_t *i1, const float *p2, const uint8_t *i2, const float *h,
                    const float *logp, float *gc1w, float *gc1b, float *gc2w, float *gc2b, float *gf1w,
                    float *gf1b, float *gf2w, float *gf2b) {
    float *dz2 = (float *)malloc(sizeof(float) * B * 10);
    float *dz1 = (float *)malloc(sizeof(float) * B * 500);
    float *dp2 = (float *)malloc(sizeof(float) * B * 800);
    float *dc2 = (float *)calloc((size_t)B * 50 * 64, sizeof(float));
    float *dp1 = (float *)malloc(sizeof(float) * B * 2880);
    float *dc1 = (float *)calloc((size_t)B * 20 * 576, sizeof(float));

    /* nll_loss + log_softmax backward: (softmax - onehot) / B */
    for (int b = 0; b < B; ++b)
        for (int j = 0; j < 10; ++j)
            dz2[b * 10 + j] = (float)((exp((double)logp[b * 10 + j]) - (j == target[b] ? 1.0 : 0.0)) / B);
    /* fc2: wgrad, bgrad, dgrad (+ ReLU mask of fc1's output) */
    for (int j = 0; j < 10; ++j) {
        double sb = 0.0;
        for (int b = 0; b < B; ++b) sb += dz2[b * 10 + j];
        gf2b[j] = (float)sb;
        for (int k = 0; k < 500; ++k) {
            double s = 0.0;
            for (int b = 0; b < B; ++b) s += (double)dz2[b * 10 + j] * (double)h[b * 500 + k];
            gf2w[j * 500 + k] = (float)s;
        }
    }
    for (int b = 0; b < B; ++b)
        for (int k = 0; k < 500; ++k) {
            double s = 0.0;
            for (int j = 0; j < 10; ++j) s += (double)dz2[b * 10 + j] * (double)f2w[j * 500 + k];
            dz1[b * 500 + k] = h[b * 500 + k] > 0.f ? (float)s : 0.f;
        }
    /* fc1 */
    for (int n = 0; n < 500; ++n) {
        double sb = 0.0;
        for (int b = 0; b < B; ++b) sb += dz1[b * 500 + n];
        gf1b[n] = (float)sb;
        for (int k = 0; k < 800; ++k) {
            double s = 0.0;
            for (int b = 0; b < B; ++b) s += (double)dz1[b * 500 + n] * (double)p2[b * 800 + k];
            gf1w[n * 800 + k] = (float)s;
        }
    }
    for (int b = 0; b < B; ++b)
        for (int k = 0; k < 800; ++k) {
            double s = 0.0;
            for (int n = 0; n < 500; ++n) s += (double)dz1[b * 500 + n] * (double)f1w[n * 800 + k];
            dp2[b * 800 + k] = (float)s;
        }
    /* pool2 + ReLU backward: route to the arg-max slot, kill where the ReLU output was 0 */
    for (int b = 0; b < B; ++b)
        for (int oc = 0; oc < 50; ++oc)
            for (int p = 0; p < 16; ++p) {
                const int o = (b * 50 + oc) * 16 + p;
                if (p2[o] > 0.f) {
                    const int oy = 2 * (p / 4) + (i2[o] >> 1), ox = 2 * (p % 4) + (i2[o] & 1);
                    dc2[(b * 50 + oc) * 64 + oy * 8 + ox] = dp2[o];
                }
            }
    /* conv2: wgrad, bgrad, dgrad */
    for (int oc = 0; oc < 50; ++oc) {
        double sb = 0.0;
        for (int b = 0; b < B; ++b)
            for (int p = 0; p < 64; ++p) sb += dc2[(b * 50 + oc) * 64 + p];
        gc2b[oc] = (float)sb;
        for (int ic = 0; ic < 20; ++ic)
            for (int ky = 0; ky < 5; ++ky)
                for (int kx = 0; kx < 5; ++kx) {
                    double s = 0.0;
                    for (int b = 0; b < B; ++b)
                        for (int oy = 0; oy < 8; ++oy)
                            for (int ox = 0; ox < 8; ++ox) {
                                const float g = dc2[(b * 50 + oc) * 64 + oy * 8 + ox];
                                if (g != 0.f) s += (double)g * (double)p1[((b * 20 + ic) * 12 + oy + ky) * 12 + ox + kx];
                            }
                    gc2w[((oc * 20 + ic) * 5 + ky) * 5 + kx] = (float)s;
                }
    }
    for (int b = 0; b < B; ++b)
        for (int ic = 0; ic < 20; ++ic)
            for (int y = 0; y < 12; ++y)
                for (int xx = 0; xx < 12; ++xx) {
                    double s = 0.0;
                    for (int oc = 0; oc < 50; ++oc)
                        for (int ky = 0; ky < 5; ++ky)
                            for (int kx = 0; kx < 5; ++kx) {
                                const int oy = y - ky, ox = xx - kx;
                                if (oy < 0 || oy >= 8 || ox < 0 || ox >= 8) continue;
                                const float g = dc2[(b * 50 + oc) * 64 + oy * 8 + ox];
                                if (g != 0.f) s += (double)g * (double)c2w[((oc * 20 + ic) * 5 + ky) * 5 + kx];
                            }
                    dp1[((b * 20 + ic) * 12 + y) * 12 + xx] = (float)s;
                }
    /* pool1 + ReLU backward, then conv1 wgrad/bgrad (the input needs no gradient) */
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < 20; ++c)
            for (int p = 0; p < 144; ++p) {
                const int o = (b * 20 + c) * 144 + p;
                if (p1[o] > 0.f) {
                    const int oy = 2 * (p / 12) + (i1[o] >> 1), ox = 2 * (p % 12) + (i1[o] & 1);
                    dc1[(b * 20 + c) * 576 + oy * 24 + ox] = dp1[o];
                }
            }
    for (int c = 0; c < 20; ++c) {
        double sb = 0.0;
        for (int b = 0; b < B; ++b)
            for (int p = 0; p < 576; ++p) sb += dc1[(b * 20 + c) * 576 + p];
        gc1b[c] = (float)sb;
        for (int ky = 0; ky < 5; ++ky)
            for (int kx = 0; kx < 5; ++kx) {
                double s = 0.0;
                for (int b = 0; b < B; ++b)
                    for (int oy = 0; oy < 24; ++oy)
                        for (int ox = 0; ox < 24; ++ox) {
                            const float g = dc1[(b * 20 + c) * 576 + oy * 24 + ox];
                            if (g != 0.f) s += (double)g * (double)x[(b * 28 + oy + ky) * 28 + ox + kx];
                        }
                gc1w[(c * 5 + ky) * 5 + kx] = (float)s;
            }
    }
    free(dz2); free(dz1); free(dp2); free(dc2); free(dp1); free(dc1);
}

/* optim.SGD(lr, momentum).step(), mnist.py:140,43, with DDP's 1/world_size (mnist.py:135-138)
 * as grad_scale.  Rounds exactly like the fp32 ATen kernels do: buf*mu and +g are two
 * roundings; p + (-lr)*buf is one fused multiply-add (ATen's add(alpha) kernel). */
void lenet_sgd_momentum(float *p, const float *g, float *buf, long n, float lr, float mu, float grad_scale,
                        int first_step) {
    const float nlr = -lr;
    for (long i = 0; i < n; ++i) {
        const float gi = g[i] * grad_scale;
        float bi;
        if (first_step) bi = gi;
        else { bi = buf[i] * mu; bi = bi + gi; }
        buf[i] = bi;
        p[i] = fmaf(bi, nlr, p[i]);
    }
}

/* The worker side of examples/smoke-dist/dist_sendrecv.py:32: torch.mul(inp, inp). */
void smoke_square(const float *in, float *out, long n) {
    for (long i = 0; i < n; ++i) out[i] = in[i] * in[i];
}
