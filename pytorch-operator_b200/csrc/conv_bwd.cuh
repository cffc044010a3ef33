// Backward kernels of the two conv + ReLU + max-pool stages (autograd of mnist.py:26-29,
// run by loss.backward() at mnist.py:42).
#pragma once
#include "common.cuh"

namespace b200mnist {

constexpr int kMaxDgSplit = 8;  // fc1 data-gradient split-K upper bound (dp2 partials)

// Gradient w.r.t. the conv2 output map of one image, built in shared memory from the
// gradient of the pooled map: max-pool backward routes each value to the arg-max slot of
// its 2x2 window, ReLU backward kills it where the (post-ReLU) maximum was 0.
//   D_s : [50][ld] floats, already zero-filled; dp2 may be the sum of `nsplit` partials.
// All global loads of a thread (<= 4 elements x (nsplit + 2)) are issued before any is used.
//   PAIRED: rows hold the channel pair {r, r+25} interleaved -- D_s[r*ld + 2*pos + (oc >= 25)] -- so a
//   128-bit load yields two ready-made (d[oc], d[oc+25]) operand pairs for FFMA2.
// Split in two so the global loads can be issued early (before a zero-fill / during the previous
// image's math) and only the shared-memory stores sit behind the barrier.
template <int NTHREADS>
struct DConv2Regs {
    static constexpr int PER = (kP2Elems + NTHREADS - 1) / NTHREADS;
    float g[PER], pv[PER];
    int q[PER];
    __device__ __forceinline__ void load(const float *__restrict__ dp2, int nsplit, size_t split_stride,
                                         const float *__restrict__ p2, const uint8_t *__restrict__ i2, size_t b,
                                         int tid) {
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int e = tid + j * NTHREADS;
            g[j] = 0.f; pv[j] = 0.f; q[j] = 0;
            if (e < kP2Elems) {
                const size_t o = b * kP2Elems + e;
                pv[j] = __ldcg(p2 + o);
                q[j] = i2[o];
                g[j] = sum_partials<kMaxDgSplit>(dp2 + o, nsplit, split_stride);
            }
        }
    }
    template <bool PAIRED>
    __device__ __forceinline__ void store(float *D_s, int ld, int tid) const {
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int e = tid + j * NTHREADS;
            if (e < kP2Elems && pv[j] > 0.f) {
                const int oc = e >> 4, p = e & 15;
                const int oy = 2 * (p >> 2) + (q[j] >> 1), ox = 2 * (p & 3) + (q[j] & 1);
                if (PAIRED) D_s[(oc >= 25 ? oc - 25 : oc) * ld + 2 * (oy * 8 + ox) + (oc >= 25 ? 1 : 0)] = g[j];
                else D_s[oc * ld + oy * 8 + ox] = g[j];
            }
        }
    }
};

// ---------------------------------------------------------------------------------
// conv2 data gradient as GEMM + col2im, per image and per tile of 4 input channels:
//   T[(ic,ky,kx), (oy,ox)] = sum_oc W[oc,(ic,ky,kx)] * dconv2[oc,(oy,ox)]    (100 x 64, K = 50)
//   dp1[ic,y,x]            = sum_{ky,kx} T[(ic,ky,kx), (y-ky, x-kx)]           (valid terms only)
// This does the 1.6 M MAC/image a scatter formulation needs (a gather "full correlation"
// over the zero-padded map would do 3.6 M), with all accumulation in registers and a
// fixed summation order.  The weight tile is loaded once per CTA and reused across the
// images the CTA loops over.
// ---------------------------------------------------------------------------------
constexpr int kDgIcPerCta = 4;
constexpr int kConvBwdThreads = 256;
constexpr int kDgLdT = 68;
constexpr int kDgSmemFloats = 50 * 100 + 50 * 64 + 100 * kDgLdT;  // 15,000 floats = 60,000 B

__device__ __forceinline__ void conv2_dgrad_tile(float *smem, const float *__restrict__ dp2, int nsplit,
                                                 size_t split_stride, const float *__restrict__ p2,
                                                 const uint8_t *__restrict__ i2, const float *__restrict__ w,
                                                 float *__restrict__ dp1, int B, int vx, int gx, int vy) {
    float *W_s = smem;                // [50 oc][100 m]
    float *D_s = W_s + 50 * 100;      // [50 oc][64 n]
    float *T_s = D_s + 50 * 64;       // [100 m][68]
    const int tid = threadIdx.x, ic0 = vy * kDgIcPerCta;
    {
        float4 wv[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const int i = tid + j * kConvBwdThreads;
            if (i < 50 * 25) {
                const int oc = i / 25, m4 = (i % 25) * 4;
                wv[j] = __ldg(reinterpret_cast<const float4 *>(w + (size_t)oc * 500 + ic0 * 25 + m4));
            }
        }
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const int i = tid + j * kConvBwdThreads;
            if (i < 50 * 25) {
                const int oc = i / 25, m4 = (i % 25) * 4;
                *reinterpret_cast<float4 *>(&W_s[oc * 100 + m4]) = wv[j];
            }
        }
    }
    const int tm = tid >> 3, tn = tid & 7;  // tm < 25 active: rows 4*tm..+3; cols 4*tn..+3 and 32+4*tn..+3
    DConv2Regs<kConvBwdThreads> dr;
    if (vx < B) dr.load(dp2, nsplit, split_stride, p2, i2, (size_t)vx, tid);
    for (int b = vx; b < B; b += gx) {
        for (int i = tid; i < 50 * 64 / 4; i += kConvBwdThreads)
            reinterpret_cast<float4 *>(D_s)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();  // also orders W_s (first trip) and the previous trip's T_s readers
        dr.template store<false>(D_s, 64, tid);
        if (b + gx < B) dr.load(dp2, nsplit, split_stride, p2, i2, (size_t)(b + gx), tid);  // next image, in flight
        __syncthreads();
        if (tm < 25) {
            float2 acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = make_float2(0.f, 0.f);
#pragma unroll 5
            for (int oc = 0; oc < 50; ++oc) {
                const float4 a = *reinterpret_cast<const float4 *>(&W_s[oc * 100 + tm * 4]);
                const float4 b0 = *reinterpret_cast<const float4 *>(&D_s[oc * 64 + tn * 4]);
                const float4 b1 = *reinterpret_cast<const float4 *>(&D_s[oc * 64 + 32 + tn * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w};
                const float2 bv[4] = {make_float2(b0.x, b0.y), make_float2(b0.z, b0.w), make_float2(b1.x, b1.y),
                                      make_float2(b1.z, b1.w)};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 aa = make_float2(av[i], av[i]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = __ffma2_rn(aa, bv[j], acc[i][j]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float *row = &T_s[(tm * 4 + i) * kDgLdT];
                *reinterpret_cast<float4 *>(row + tn * 4) = make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
                *reinterpret_cast<float4 *>(row + 32 + tn * 4) =
                    make_float4(acc[i][2].x, acc[i][2].y, acc[i][3].x, acc[i][3].y);
            }
        }
        __syncthreads();
        for (int o = tid; o < kDgIcPerCta * 144; o += kConvBwdThreads) {
            const int icl = o / 144, r = o - icl * 144, y = r / 12, x = r - y * 12;
            float s = 0.f;
#pragma unroll
            for (int ky = 0; ky < 5; ++ky) {
                const int oy = y - ky;
                if (oy < 0 || oy >= 8) continue;
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    const int ox = x - kx;
                    if (ox < 0 || ox >= 8) continue;
                    s += T_s[(icl * 25 + ky * 5 + kx) * kDgLdT + oy * 8 + ox];
                }
            }
            dp1[(size_t)b * kP1Elems + (ic0 + icl) * 144 + r] = s;
        }
    }
}

// ---------------------------------------------------------------------------------
// conv2 weight/bias gradient, dense, register resident:
//   gW[oc,ic,ky,kx] = sum_b sum_{oy,ox} dconv2[b,oc,oy,ox] * p1[b,ic,oy+ky,ox+kx]
// CTA = tile of 4 input channels x a slice of the batch; 8 warps = 4 channels x 2 images in
// flight; lane = output channel pair {lane, lane+25} (25 active lanes), so each thread keeps
// 2x25 accumulators.  Per conv row it loads a 5x12 strip of p1 (warp-uniform broadcast) and
// 2x8 gradient values (row stride 68 floats: conflict-free LDS.128) for 400 FMAs -- about 21
// FMA per shared-memory load.  Each CTA writes its partial [25,056 + 64]; the partials of the
// batch slices are summed in a fixed order by grad_finalize / the fused SGD kernel.
// ---------------------------------------------------------------------------------
constexpr int kWgIcPerCta = 4;
constexpr int kWgLdD = 132;  // floats per channel-pair row: 2 x 64 interleaved + 4 pad (conflict-free LDS.128)
constexpr int kC2PartStride = 25088;  // 25,000 weights @0, 50 biases @25,024
constexpr int kWgSmemFloats = 2 * kWgIcPerCta * 144 + 2 * 25 * kWgLdD;  // 7,752 floats

__device__ __forceinline__ void conv2_wgrad_tile(float *smem, const float *__restrict__ dp2, int nsplit,
                                                 size_t split_stride, const float *__restrict__ p2,
                                                 const uint8_t *__restrict__ i2, const float *__restrict__ p1,
                                                 float *__restrict__ part, int B, int vx, int vy, int gy) {
    float *P_s = smem;                                // [2][4 ic][144]
    float *D_s = smem + 2 * kWgIcPerCta * 144;        // [2][25 pairs][132] (later: the 50 x 100 output tile)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, icl = warp & 3, par = warp >> 2;
    const int ic0 = vx * kWgIcPerCta;
    const bool active = lane < 25;
    float2 acc[5][5], accb = make_float2(0.f, 0.f);  // .x: channel `lane`, .y: channel `lane + 25`
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = 0; j < 5; ++j) acc[i][j] = make_float2(0.f, 0.f);

    for (int bb = vy * 2; bb < B; bb += gy * 2) {
        __syncthreads();  // previous trip's readers are done
        {   // stage the two p1 tiles (global load first, so it overlaps the zero-fill below)
            float4 pv[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int i = tid + j * kConvBwdThreads;
                pv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < 2 * kWgIcPerCta * 36) {
                    const int pr = i / (kWgIcPerCta * 36), r = i - pr * (kWgIcPerCta * 36), b = bb + pr;
                    if (b < B) pv[j] = __ldg(reinterpret_cast<const float4 *>(p1 + (size_t)b * kP1Elems + ic0 * 144) + r);
                }
            }
            DConv2Regs<kConvBwdThreads> d0, d1;
            d0.load(dp2, nsplit, split_stride, p2, i2, (size_t)bb, tid);
            if (bb + 1 < B) d1.load(dp2, nsplit, split_stride, p2, i2, (size_t)bb + 1, tid);
            for (int i = tid; i < 2 * 25 * kWgLdD / 4; i += kConvBwdThreads)
                reinterpret_cast<float4 *>(D_s)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int i = tid + j * kConvBwdThreads;
                if (i < 2 * kWgIcPerCta * 36) reinterpret_cast<float4 *>(P_s)[i] = pv[j];
            }
            __syncthreads();
            d0.template store<true>(D_s, kWgLdD, tid);
            if (bb + 1 < B) d1.template store<true>(D_s + 25 * kWgLdD, kWgLdD, tid);
        }
        __syncthreads();
        if (active && bb + par < B) {
            const float *ps = P_s + (par * kWgIcPerCta + icl) * 144;
            const float *dp = D_s + (par * 25 + lane) * kWgLdD;
#pragma unroll 1
            for (int oy = 0; oy < 8; ++oy) {
                float strip[5][12];
#pragma unroll
                for (int r = 0; r < 5; ++r)
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        const float4 v = *reinterpret_cast<const float4 *>(ps + (oy + r) * 12 + 4 * j);
                        strip[r][4 * j] = v.x; strip[r][4 * j + 1] = v.y;
                        strip[r][4 * j + 2] = v.z; strip[r][4 * j + 3] = v.w;
                    }
                float2 d[8];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 v = *reinterpret_cast<const float4 *>(dp + oy * 16 + 4 * j);
                    d[2 * j] = make_float2(v.x, v.y);
                    d[2 * j + 1] = make_float2(v.z, v.w);
                }
#pragma unroll
                for (int ox = 0; ox < 8; ++ox) {
                    if (icl == 0) { accb.x += d[ox].x; accb.y += d[ox].y; }
#pragma unroll
                    for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                        for (int kx = 0; kx < 5; ++kx) {
                            const float sv = strip[ky][ox + kx];
                            acc[ky][kx] = __ffma2_rn(d[ox], make_float2(sv, sv), acc[ky][kx]);
                        }
                }
            }
        }
    }
    // ---- combine the two image-parity halves, then write the 50 x 100 tile coalesced ----
    __syncthreads();
    float *O_s = D_s;  // [50 oc][100]  (+ 64 bias slots at 5000)
    if (par == 1 && active) {
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                O_s[lane * 100 + icl * 25 + ky * 5 + kx] = acc[ky][kx].x;
                O_s[(lane + 25) * 100 + icl * 25 + ky * 5 + kx] = acc[ky][kx].y;
            }
        if (icl == 0) { O_s[5000 + lane] = accb.x; O_s[5000 + lane + 25] = accb.y; }
    }
    __syncthreads();
    if (par == 0 && active) {
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                O_s[lane * 100 + icl * 25 + ky * 5 + kx] += acc[ky][kx].x;
                O_s[(lane + 25) * 100 + icl * 25 + ky * 5 + kx] += acc[ky][kx].y;
            }
        if (icl == 0) { O_s[5000 + lane] += accb.x; O_s[5000 + lane + 25] += accb.y; }
    }
    __syncthreads();
    float *po = part + (size_t)vy * kC2PartStride;
    for (int i = tid; i < 5000; i += kConvBwdThreads) {
        const int oc = i / 100, m = i - oc * 100;
        po[oc * 500 + ic0 * 25 + m] = O_s[i];
    }
    if (vx == 0 && tid < 50) po[25024 + tid] = O_s[5000 + tid];
}

// One launch for both halves of the conv2 backward (they only share their inputs):
// CTAs [0, n_wg) are weight-gradient tiles (5 channel tiles x wg_gy batch slices), the rest
// data-gradient tiles (dg_gx image slots x 5 channel tiles).  The weight-gradient CTAs come
// first because they are the longer ones.
struct Conv2BwdArgs {
    const float *dp2, *p2, *p1, *w;
    const uint8_t *i2;
    float *dp1, *part;
    size_t split_stride;
    int nsplit, B, wg_gy, n_wg, dg_gx;
    EarlySignal sig;   // fused DDP step: tell the peers that this rank's fc1 gradient is staged (n = 0: nothing)
};
constexpr int kConv2BwdSmemBytes = kDgSmemFloats * 4;  // 60,000 B dynamic (covers both roles)

__global__ void __launch_bounds__(kConvBwdThreads) conv2_bwd_kernel(const __grid_constant__ Conv2BwdArgs a) {

    pdl_wait();
    send_early_signal(a.sig);
    extern __shared__ __align__(16) float smem[];
    int b = blockIdx.x;
    if (b < a.n_wg) {
        conv2_wgrad_tile(smem, a.dp2, a.nsplit, a.split_stride, a.p2, a.i2, a.p1, a.part, a.B, b % 5, b / 5, a.wg_gy);
    } else {
        b -= a.n_wg;
        conv2_dgrad_tile(smem, a.dp2, a.nsplit, a.split_stride, a.p2, a.i2, a.w, a.dp1, a.B, b % a.dg_gx, a.dg_gx,
                         b / a.dg_gx);
    }
}

// ---------------------------------------------------------------------------------
// conv1 weight/bias gradient (the input image needs no gradient).  The conv1 output
// gradient has one non-zero per 2x2 window (the pool arg-max; none where ReLU was inactive),
// so the sum runs over the 144 pooled windows, not the 576 conv outputs.  warp = channel,
// lane = window (five windows per lane): the lane keeps its window's (gradient, input offset)
// in registers and accumulates all 25 taps -- 25 independent FMA chains fed by one LDS each,
// no list building, no dependent shared-memory round trips -- plus the bias gradient; the
// 26 per-lane sums are reduced across the warp once, after the CTA's last image.  The next
// image's pixels and gradient windows are fetched while the current one is accumulated.
// grid = (batch slices, 2 channel halves); each CTA writes a partial [512 + 64].
// ---------------------------------------------------------------------------------
constexpr int kC1wThreads = 320;    // 10 channels x 32 lanes
constexpr int kC1PartStride = 576;  // 500 weights @0, 20 biases @512

// bx / nbx / by: this CTA's batch-slice index, the number of slices, and its channel half (the plain kernel passes
// its block / grid indices; the fused DDP kernel of peer_exchange.cuh shares its grid with another role)
__device__ __forceinline__ void conv1_wgrad_body(const float *__restrict__ x, const float *__restrict__ dp1,
                                                 const float *__restrict__ p1, const uint8_t *__restrict__ i1,
                                                 float *__restrict__ part, int B, const int bx, const int nbx, const int by) {
    // the image is staged as two column-parity planes, xs[c & 1][row][c >> 1]: lanes are consecutive windows, whose
    // pixels are two columns apart -- in a plane they are neighbours, which halves the bank span of every load
    // (ncu on the single-plane layout: 62 % of the shared-memory wavefronts were bank-conflict replays)
    __shared__ __align__(16) float xs[kXElems];
    const int tid = threadIdx.x, wid = tid >> 5, c = by * 10 + wid, lane = tid & 31;
    float acc[25], accb = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) acc[k] = 0.f;
    // per window: gradient and the plane offsets of the arg-max pixel for even kernel columns (oa) and odd ones (ob)
    float gl[5], gl_n[5];
    int oa[5], ob[5], oa_n[5], ob_n[5];
    float4 xv = make_float4(0.f, 0.f, 0.f, 0.f), xv_n = xv;
    auto fetch = [&](int b, float (&g)[5], int (&pa)[5], int (&pb)[5], float4 &xq) {
        const size_t base = ((size_t)b * 20 + c) * 144;
        if (tid < kXElems / 4) xq = __ldg(reinterpret_cast<const float4 *>(x + (size_t)b * kXElems) + tid);
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const int wl = j * 32 + lane;
            g[j] = 0.f;
            pa[j] = 0;
            pb[j] = 0;
            if (wl < 144) {
                const float pv = __ldcg(p1 + base + wl);
                const float gv = __ldcg(dp1 + base + wl);
                const int q = i1[base + wl];
                g[j] = pv > 0.f ? gv : 0.f;
                const int dx = q & 1, at = (2 * (wl / 12) + (q >> 1)) * 14 + wl % 12;   // pixel column 2 wx + dx
                pa[j] = dx * 392 + at;               // column 2 wx + dx + kx, kx even: plane dx, offset + kx / 2
                pb[j] = (dx ^ 1) * 392 + at + dx;    // kx odd: the other plane, offset + (kx - 1) / 2 (+ 1 if dx)
            }
        }
    };
    if (bx < B) fetch(bx, gl_n, oa_n, ob_n, xv_n);
    for (int b = bx; b < B; b += nbx) {
#pragma unroll
        for (int j = 0; j < 5; ++j) { gl[j] = gl_n[j]; oa[j] = oa_n[j]; ob[j] = ob_n[j]; }
        xv = xv_n;
        if (b + nbx < B) fetch(b + nbx, gl_n, oa_n, ob_n, xv_n);
        __syncthreads();  // previous image's readers of xs are done
        if (tid < kXElems / 4) {   // pixels 4 tid .. 4 tid + 3 of one row: columns col, col + 2 -> plane 0; col + 1, col + 3 -> plane 1
            const int row = tid / 7, half = (tid - row * 7) * 2;
            *reinterpret_cast<float2 *>(&xs[row * 14 + half]) = make_float2(xv.x, xv.z);
            *reinterpret_cast<float2 *>(&xs[392 + row * 14 + half]) = make_float2(xv.y, xv.w);
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const float g = gl[j];
            const float *pe = xs + oa[j], *po = xs + ob[j];
            accb += g;
#pragma unroll
            for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    const float v = (kx & 1) ? po[ky * 14 + (kx - 1) / 2] : pe[ky * 14 + kx / 2];
                    acc[ky * 5 + kx] = fmaf(g, v, acc[ky * 5 + kx]);
                }
        }
    }
    float mine = 0.f;   // lane k < 25 ends up with tap k, lane 25 with the bias gradient
#pragma unroll
    for (int k = 0; k < 25; ++k) {
        const float s = warp_sum(acc[k]);
        if (lane == k) mine = s;
    }
    {
        const float s = warp_sum(accb);
        if (lane == 25) mine = s;
    }
    float *po = part + (size_t)bx * kC1PartStride;
    if (lane < 25) po[c * 25 + lane] = mine;
    else if (lane == 25) po[512 + c] = mine;
}

__global__ void __launch_bounds__(kC1wThreads)
conv1_wgrad_kernel(const float *__restrict__ x, const float *__restrict__ dp1, const float *__restrict__ p1,
                   const uint8_t *__restrict__ i1, float *__restrict__ part, int B) {
    pdl_wait();
    conv1_wgrad_body(x, dp1, p1, i1, part, B, blockIdx.x, gridDim.x, blockIdx.y);
}

}  // namespace b200mnist
