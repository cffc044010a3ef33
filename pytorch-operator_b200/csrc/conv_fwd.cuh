// Forward kernels for the two conv5x5 + ReLU + 2x2 max-pool stages (mnist.py:26-29).
#pragma once
#include "common.cuh"

namespace b200mnist {

// ---------------------------------------------------------------------------------
// conv1: 1 -> 20 channels on 28x28 (K = 25 per output: not a contraction worth a tensor
// core).  One CTA = one image x 5 output channels; the 28x28 image is staged once in
// shared memory, each thread owns one pooled pixel: its 6x6 input patch sits in registers
// and is reused for the 4 conv outputs of the window and for all 5 channels (weights are
// warp-uniform shared-memory broadcasts).  ReLU, max and arg-max are taken in registers,
// so the 24x24 conv map never exists in memory.
// ---------------------------------------------------------------------------------
constexpr int kC1ChPerCta = 5;
constexpr int kC1Threads = 160;

__global__ void __launch_bounds__(kC1Threads)
conv1_relu_pool_fwd_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                           float *__restrict__ out, uint8_t *__restrict__ idx, int B) {
    __shared__ __align__(16) float xs[kXElems];
    __shared__ float ws[kC1ChPerCta * 25];
    __shared__ float bs[kC1ChPerCta];
    const int b = blockIdx.x, c0 = blockIdx.y * kC1ChPerCta, t = threadIdx.x;
    const float4 *xg = reinterpret_cast<const float4 *>(x + (size_t)b * kXElems);
    for (int i = t; i < kXElems / 4; i += kC1Threads) reinterpret_cast<float4 *>(xs)[i] = __ldg(xg + i);
    if (t < kC1ChPerCta * 25) ws[t] = __ldg(w + c0 * 25 + t);
    if (t < kC1ChPerCta) bs[t] = __ldg(bias + c0 + t);
    __syncthreads();
    if (t >= 144) return;
    const int py = t / 12, px = t % 12;
    float patch[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float2 v = *reinterpret_cast<const float2 *>(&xs[(2 * py + r) * 28 + 2 * px + 2 * j]);
            patch[r][2 * j] = v.x;
            patch[r][2 * j + 1] = v.y;
        }
#pragma unroll
    for (int c = 0; c < kC1ChPerCta; ++c) {
        float a0 = bs[c], a1 = a0, a2 = a0, a3 = a0;
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const float wv = ws[c * 25 + ky * 5 + kx];
                a0 = fmaf(wv, patch[ky][kx], a0);
                a1 = fmaf(wv, patch[ky][kx + 1], a1);
                a2 = fmaf(wv, patch[ky + 1][kx], a2);
                a3 = fmaf(wv, patch[ky + 1][kx + 1], a3);
            }
        float best = a0;
        int bi = 0;
        if (a1 > best) { best = a1; bi = 1; }
        if (a2 > best) { best = a2; bi = 2; }
        if (a3 > best) { best = a3; bi = 3; }
        const size_t o = ((size_t)b * 20 + c0 + c) * 144 + t;
        out[o] = fmaxf(best, 0.f);
        idx[o] = (uint8_t)bi;
    }
}

// ---------------------------------------------------------------------------------
// Register-tiled 5x5 direct convolution micro-kernel shared by conv2 forward:
// one thread accumulates a 2x2 block of outputs for 10 output channels
// (40 accumulators) over input channels [c_begin, c_end).  Per input channel it loads
// the 6x6 patch once (18 LDS.64) and, per tap, the 10 channel weights as a warp-uniform
// broadcast (2 LDS.128 + 1 LDS.64): 1000 FMA per 93 shared-memory loads.
//   in_s : this lane's image, planes of `plane` floats, rows of IW floats
//   w_s  : [cin][25 taps][12] (10 output channels + 2 pad), 16-byte aligned rows
// ---------------------------------------------------------------------------------
template <int IW>
__device__ __forceinline__ void conv5_tile_accum(const float *in_s, int plane, const float *w_s, int c_begin,
                                                 int c_end, int r0, int c0, float (&acc)[10][4]) {
    for (int c = c_begin; c < c_end; ++c) {
        const float *pl = in_s + c * plane + r0 * IW + c0;
        float patch[6][6];
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const float2 v = *reinterpret_cast<const float2 *>(pl + r * IW + 2 * j);
                patch[r][2 * j] = v.x;
                patch[r][2 * j + 1] = v.y;
            }
        const float *wc = w_s + c * 25 * 12;
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const float *wp = wc + (ky * 5 + kx) * 12;
                const float4 w0 = *reinterpret_cast<const float4 *>(wp);
                const float4 w1 = *reinterpret_cast<const float4 *>(wp + 4);
                const float2 w2 = *reinterpret_cast<const float2 *>(wp + 8);
                const float wv[10] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y};
#pragma unroll
                for (int o = 0; o < 10; ++o) {
                    acc[o][0] = fmaf(wv[o], patch[ky][kx], acc[o][0]);
                    acc[o][1] = fmaf(wv[o], patch[ky][kx + 1], acc[o][1]);
                    acc[o][2] = fmaf(wv[o], patch[ky + 1][kx], acc[o][2]);
                    acc[o][3] = fmaf(wv[o], patch[ky + 1][kx + 1], acc[o][3]);
                }
            }
    }
}

// ---------------------------------------------------------------------------------
// conv2: 20 -> 50 channels on 12x12, K = 500 per output (70% of the forward MACs).
// One CTA = 2 images x 10 output channels.  The 4 warps split the 20 input channels
// (5 each); a lane is (image, pooled pixel), so all lanes of a warp share the weight
// address (broadcast) and hit distinct banks for the activation patch.  The 4 partial
// sums are combined through shared memory (aliasing the image tile), then bias, ReLU,
// max and arg-max are applied and only the pooled 4x4 map is written.
// ---------------------------------------------------------------------------------
constexpr int kC2ImgPerCta = 2;
constexpr int kC2OcPerCta = 10;
constexpr int kC2Threads = 128;

__global__ void __launch_bounds__(kC2Threads)
conv2_relu_pool_fwd_kernel(const float *__restrict__ in, const float *__restrict__ w, const float *__restrict__ bias,
                           float *__restrict__ out, uint8_t *__restrict__ idx, int B) {
    __shared__ __align__(16) float in_s[kC2ImgPerCta * kP1Elems];  // 23,040 B; later the 4x40x32 reduce buffer
    __shared__ __align__(16) float w_s[500 * 12];                  // 24,000 B
    const int t = threadIdx.x, b0 = blockIdx.x * kC2ImgPerCta, oc0 = blockIdx.y * kC2OcPerCta;
    {   // stage both operands with every global load in flight before the first store
        constexpr int kImgVec = kC2ImgPerCta * kP1Elems / 4;        // 1440 float4
        constexpr int kWVec = kC2OcPerCta * 500 / 4;                // 1250 float4 (10 contiguous rows of W)
        constexpr int kImgPer = (kImgVec + kC2Threads - 1) / kC2Threads, kWPer = (kWVec + kC2Threads - 1) / kC2Threads;
        float4 iv[kImgPer], wv[kWPer];
#pragma unroll
        for (int j = 0; j < kImgPer; ++j) {
            const int i = t + j * kC2Threads;
            iv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < kImgVec) {
                const int img = i / (kP1Elems / 4), b = b0 + img;
                if (b < B) iv[j] = __ldg(reinterpret_cast<const float4 *>(in + (size_t)b * kP1Elems) + (i - img * (kP1Elems / 4)));
            }
        }
        const float4 *wg = reinterpret_cast<const float4 *>(w + (size_t)oc0 * 500);
#pragma unroll
        for (int j = 0; j < kWPer; ++j) {
            const int i = t + j * kC2Threads;
            wv[j] = i < kWVec ? __ldg(wg + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int j = 0; j < kImgPer; ++j) {
            const int i = t + j * kC2Threads;
            if (i < kImgVec) reinterpret_cast<float4 *>(in_s)[i] = iv[j];
        }
#pragma unroll
        for (int j = 0; j < kWPer; ++j) {
            const int i = t + j * kC2Threads;
            if (i < kWVec) {  // element e = 4*i of the [10][500] block -> w_s[(ic*25+k)][oc]
                const float v[4] = {wv[j].x, wv[j].y, wv[j].z, wv[j].w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int e = 4 * i + u, oc = e / 500, r = e - oc * 500;
                    w_s[r * 12 + oc] = v[u];
                }
            }
        }
    }
    __syncthreads();
    const int warp = t >> 5, lane = t & 31, img = lane >> 4, pos = lane & 15, py = pos >> 2, px = pos & 3;
    float acc[10][4];
#pragma unroll
    for (int o = 0; o < 10; ++o)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[o][q] = 0.f;
    conv5_tile_accum<12>(in_s + img * kP1Elems, 144, w_s, warp * 5, warp * 5 + 5, 2 * py, 2 * px, acc);
    __syncthreads();  // all patches read; in_s becomes the reduce buffer
    float *red = in_s;
#pragma unroll
    for (int o = 0; o < 10; ++o)
#pragma unroll
        for (int q = 0; q < 4; ++q) red[(warp * 40 + o * 4 + q) * 32 + lane] = acc[o][q];
    __syncthreads();
    const int b = b0 + img;
    for (int o = warp; o < 10; o += 4) {
        const float bv = __ldg(bias + oc0 + o);
        float v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float s = red[(0 * 40 + o * 4 + q) * 32 + lane];
            s += red[(1 * 40 + o * 4 + q) * 32 + lane];
            s += red[(2 * 40 + o * 4 + q) * 32 + lane];
            s += red[(3 * 40 + o * 4 + q) * 32 + lane];
            v[q] = s + bv;
        }
        float best = v[0];
        int bi = 0;
        if (v[1] > best) { best = v[1]; bi = 1; }
        if (v[2] > best) { best = v[2]; bi = 2; }
        if (v[3] > best) { best = v[3]; bi = 3; }
        if (b < B) {
            const size_t oidx = ((size_t)b * 50 + oc0 + o) * 16 + pos;
            out[oidx] = fmaxf(best, 0.f);
            idx[oidx] = (uint8_t)bi;
        }
    }
}

}  // namespace b200mnist
