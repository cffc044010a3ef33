// Forward kernels for the two conv5x5 + ReLU + 2x2 max-pool stages (mnist.py:26-29).
// All multiply-adds are issued as packed FFMA2 (fma.rn.f32x2, new on sm_100): two fp32 FMAs per
// lane per instruction, IEEE-identical to scalar FFMA, pairing two OUTPUT CHANNELS so that the
// weight pair comes straight out of a 64/128-bit shared-memory load.
#pragma once
#include "common.cuh"

namespace b200mnist {

// ---------------------------------------------------------------------------------
// conv1: 1 -> 20 channels on 28x28 (K = 25 per output: not a contraction worth a tensor
// core).  One CTA = one image x groups of 4 output channels (gridDim.y = 5: one group per CTA,
// the small-batch shape that fills the GPU; gridDim.y = 1: all five groups in turn, which at
// large batch amortises the image staging and the patch load); the 28x28 image is staged once in
// shared memory, each thread owns one pooled pixel: its 6x6 input patch sits in registers
// and is reused for the 4 conv outputs of the window and for all channels (weights are
// warp-uniform LDS.128 broadcasts of [tap][4 channels]).  ReLU, max and arg-max are taken in
// registers, so the 24x24 conv map never exists in memory.
// ---------------------------------------------------------------------------------
constexpr int kC1ChPerCta = 4;
constexpr int kC1Threads = 160;

// x_copy (ring variant only): the image is also stored there, for the later kernels of the step
template <bool COPY>
__device__ __forceinline__ void conv1_fwd_body(const float *__restrict__ x, const float *__restrict__ w,
                                               const float *__restrict__ bias, float *__restrict__ out,
                                               uint8_t *__restrict__ idx, int B, float *__restrict__ x_copy) {
    __shared__ __align__(16) float xs[kXElems];
    __shared__ __align__(16) float ws[5][25 * kC1ChPerCta];
    __shared__ float bs[20];
    const int b = blockIdx.x, t = threadIdx.x;
    const float4 *xg = reinterpret_cast<const float4 *>(x + (size_t)b * kXElems);
    float4 xv0 = __ldg(xg + t), xv1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t + kC1Threads < kXElems / 4) xv1 = __ldg(xg + t + kC1Threads);
    if (COPY && blockIdx.y == 0) {
        float4 *xc = reinterpret_cast<float4 *>(x_copy + (size_t)b * kXElems);
        xc[t] = xv0;
        if (t + kC1Threads < kXElems / 4) xc[t + kC1Threads] = xv1;
    }
    for (int cg = blockIdx.y; cg < 5; cg += gridDim.y) {   // this CTA's channel groups: [tap][4 channels] each
        if (t < 25 * kC1ChPerCta) ws[cg][(t % 25) * kC1ChPerCta + t / 25] = __ldg(w + cg * kC1ChPerCta * 25 + t);
        if (t < kC1ChPerCta) bs[cg * kC1ChPerCta + t] = __ldg(bias + cg * kC1ChPerCta + t);
    }
    reinterpret_cast<float4 *>(xs)[t] = xv0;
    if (t + kC1Threads < kXElems / 4) reinterpret_cast<float4 *>(xs)[t + kC1Threads] = xv1;
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
    for (int cg = blockIdx.y; cg < 5; cg += gridDim.y) {
    const int c0 = cg * kC1ChPerCta;
    float2 acc[2][4];  // [channel pair][window slot]
#pragma unroll
    for (int cp = 0; cp < 2; ++cp)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[cp][q] = make_float2(bs[c0 + 2 * cp], bs[c0 + 2 * cp + 1]);
#pragma unroll
    for (int ky = 0; ky < 5; ++ky)
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
            const float4 w4 = *reinterpret_cast<const float4 *>(&ws[cg][(ky * 5 + kx) * kC1ChPerCta]);
            const float2 wp[2] = {make_float2(w4.x, w4.y), make_float2(w4.z, w4.w)};
            const float2 p00 = make_float2(patch[ky][kx], patch[ky][kx]);
            const float2 p01 = make_float2(patch[ky][kx + 1], patch[ky][kx + 1]);
            const float2 p10 = make_float2(patch[ky + 1][kx], patch[ky + 1][kx]);
            const float2 p11 = make_float2(patch[ky + 1][kx + 1], patch[ky + 1][kx + 1]);
#pragma unroll
            for (int cp = 0; cp < 2; ++cp) {
                acc[cp][0] = __ffma2_rn(wp[cp], p00, acc[cp][0]);
                acc[cp][1] = __ffma2_rn(wp[cp], p01, acc[cp][1]);
                acc[cp][2] = __ffma2_rn(wp[cp], p10, acc[cp][2]);
                acc[cp][3] = __ffma2_rn(wp[cp], p11, acc[cp][3]);
            }
        }
#pragma unroll
    for (int c = 0; c < kC1ChPerCta; ++c) {
        const int cp = c >> 1;
        const float a0 = (c & 1) ? acc[cp][0].y : acc[cp][0].x, a1 = (c & 1) ? acc[cp][1].y : acc[cp][1].x;
        const float a2 = (c & 1) ? acc[cp][2].y : acc[cp][2].x, a3 = (c & 1) ? acc[cp][3].y : acc[cp][3].x;
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
}

__global__ void __launch_bounds__(kC1Threads)
conv1_relu_pool_fwd_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                           float *__restrict__ out, uint8_t *__restrict__ idx, int B) {
    pdl_wait();
    conv1_fwd_body<false>(x, w, bias, out, idx, B, nullptr);
}

// The same from a resident ring (RingRef): the cursor is read after the previous step, whose finish_step advanced
// it, has completed; CTA (b, 0) also forwards target b.
__global__ void __launch_bounds__(kC1Threads)
conv1_relu_pool_fwd_ring_kernel(const RingRef r, const float *__restrict__ w, const float *__restrict__ bias,
                                float *__restrict__ out, uint8_t *__restrict__ idx, int B) {
    pdl_wait();
    uint2 st;   // {cursor, steps done}: one 8-byte load, straight from L2
    asm volatile("ld.global.cg.v2.u32 {%0, %1}, [%2];" : "=r"(st.x), "=r"(st.y) : "l"(r.state));
    unsigned seen = 0;   // host-fed ring: issued right behind the state load, so both are in flight together
    if (r.landed && threadIdx.x == 0) seen = ld_acquire_sys(r.landed);
    const unsigned slot = st.x;
    if (r.landed) {   // the batch of this step may still be on its way (normally it landed long ago)
        if (threadIdx.x == 0) {
            SpinClock clk;
            while ((int)(seen - (st.y + 1)) < 0) {
                clk.tick();
                seen = ld_acquire_sys(r.landed);
            }
        }
        __syncthreads();
    }
    if (blockIdx.y == 0 && threadIdx.x == 0) r.t_copy[blockIdx.x] = r.target[(long long)slot * r.t_stride + blockIdx.x];
    conv1_fwd_body<true>(r.x + (long long)slot * r.x_stride, w, bias, out, idx, B, r.x_copy);
}

// ---------------------------------------------------------------------------------
// Register-tiled 5x5 direct convolution micro-kernel of conv2 forward: one thread
// accumulates a 2x2 block of outputs for 10 output channels (20 packed accumulators) over
// input channels [c_begin, c_end).  Per input channel it loads the 6x6 patch once
// (18 LDS.64) and, per tap, the 5 channel-pair weights as a warp-uniform broadcast
// (2 LDS.128 + 1 LDS.64) feeding 20 FFMA2: 500 FFMA2 (= 1000 MAC) per 93 shared-memory loads.
//   in_s : this lane's image, planes of `plane` floats, rows of IW floats
//   w_s  : [cin*25 taps][12] (10 output channels + 2 pad), 16-byte aligned rows
// ---------------------------------------------------------------------------------
template <int IW>
__device__ __forceinline__ void conv5_tile_accum(const float *in_s, int plane, const float *w_s, int c_begin,
                                                 int c_end, int r0, int c0, float2 (&acc)[5][4]) {
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
                const float2 wv[5] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y),
                                      make_float2(w1.z, w1.w), w2};
                const float2 p00 = make_float2(patch[ky][kx], patch[ky][kx]);
                const float2 p01 = make_float2(patch[ky][kx + 1], patch[ky][kx + 1]);
                const float2 p10 = make_float2(patch[ky + 1][kx], patch[ky + 1][kx]);
                const float2 p11 = make_float2(patch[ky + 1][kx + 1], patch[ky + 1][kx + 1]);
#pragma unroll
                for (int o = 0; o < 5; ++o) {
                    acc[o][0] = __ffma2_rn(wv[o], p00, acc[o][0]);
                    acc[o][1] = __ffma2_rn(wv[o], p01, acc[o][1]);
                    acc[o][2] = __ffma2_rn(wv[o], p10, acc[o][2]);
                    acc[o][3] = __ffma2_rn(wv[o], p11, acc[o][3]);
                }
            }
    }
}

// ---------------------------------------------------------------------------------
// conv2: 20 -> 50 channels on 12x12, K = 500 per output (70% of the forward MACs).
// One CTA = 2 images x 10 output channels; the 4 warps split the 20 input channels (5 each, one
// warp per SM sub-partition);
// a lane is (image, pooled pixel), so all lanes of a warp share the weight address
// (broadcast) and hit distinct banks for the activation patch.  The 5 partial sums are
// combined through shared memory (aliasing the staged operands), then bias, ReLU, max and
// arg-max are applied and only the pooled 4x4 map is written.
// ---------------------------------------------------------------------------------
constexpr int kC2ImgPerCta = 2;
constexpr int kC2OcPerCta = 10;
constexpr int kC2Warps = 4;
constexpr int kC2Threads = kC2Warps * 32;

__global__ void __launch_bounds__(kC2Threads)
conv2_relu_pool_fwd_kernel(const float *__restrict__ in, const float *__restrict__ w, const float *__restrict__ bias,
                           float *__restrict__ out, uint8_t *__restrict__ idx, int B) {

    pdl_wait();
    __shared__ __align__(16) float smem_c2[kC2ImgPerCta * kP1Elems + 500 * 12];  // 23,040 + 24,000 B
    float *in_s = smem_c2;                            // [2][20][144]
    float *w_s = smem_c2 + kC2ImgPerCta * kP1Elems;   // [500][12]
    const int t = threadIdx.x, b0 = blockIdx.x * kC2ImgPerCta, oc0 = blockIdx.y * kC2OcPerCta;
    {   // stage both operands with every global load in flight before the first store
        constexpr int kImgVec = kC2ImgPerCta * kP1Elems / 4;                    // 1440 float4
        constexpr int kImgPer = (kImgVec + kC2Threads - 1) / kC2Threads;        // 9
        constexpr int kTapPer = (500 + kC2Threads - 1) / kC2Threads;            // 4
        float4 iv[kImgPer];
#pragma unroll
        for (int j = 0; j < kImgPer; ++j) {
            const int i = t + j * kC2Threads;
            iv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < kImgVec) {
                const int img = i / (kP1Elems / 4), b = b0 + img;
                if (b < B) iv[j] = __ldg(reinterpret_cast<const float4 *>(in + (size_t)b * kP1Elems) + (i - img * (kP1Elems / 4)));
            }
        }
        // weights: a thread owns taps r = t + j*160 and gathers their 10 output channels (each load
        // is coalesced across the warp), then writes each tap row with three 16-byte stores.
        float wv[kTapPer][kC2OcPerCta];
#pragma unroll
        for (int j = 0; j < kTapPer; ++j) {
            const int r = t + j * kC2Threads;
#pragma unroll
            for (int o = 0; o < kC2OcPerCta; ++o) wv[j][o] = r < 500 ? __ldg(w + (size_t)(oc0 + o) * 500 + r) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < kImgPer; ++j) {
            const int i = t + j * kC2Threads;
            if (i < kImgVec) reinterpret_cast<float4 *>(in_s)[i] = iv[j];
        }
#pragma unroll
        for (int j = 0; j < kTapPer; ++j) {
            const int r = t + j * kC2Threads;
            if (r < 500) {
                float4 *row = reinterpret_cast<float4 *>(w_s + r * 12);
                row[0] = make_float4(wv[j][0], wv[j][1], wv[j][2], wv[j][3]);
                row[1] = make_float4(wv[j][4], wv[j][5], wv[j][6], wv[j][7]);
                row[2] = make_float4(wv[j][8], wv[j][9], 0.f, 0.f);
            }
        }
    }
    __syncthreads();
    const int warp = t >> 5, lane = t & 31;
    const int img = lane >> 4, pos = lane & 15, py = pos >> 2, px = pos & 3;
    float2 acc[5][4];
#pragma unroll
    for (int o = 0; o < 5; ++o)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[o][q] = make_float2(0.f, 0.f);
    conv5_tile_accum<12>(in_s + img * kP1Elems, 144, w_s, warp * 5, warp * 5 + 5, 2 * py, 2 * px, acc);
    __syncthreads();  // all operands consumed; the buffer becomes the reduce area
    float2 *red = reinterpret_cast<float2 *>(smem_c2);  // [warp][pair o][slot q][32 lanes] float2 = 25,600 B
#pragma unroll
    for (int o = 0; o < 5; ++o)
#pragma unroll
        for (int q = 0; q < 4; ++q) red[((warp * 5 + o) * 4 + q) * 32 + lane] = acc[o][q];
    __syncthreads();
    // warp w finishes channel pair w (and warp 0 also pair 4): channels 2*pr and 2*pr+1
    const int b = b0 + img;
    for (int pr = warp; pr < 5; pr += kC2Warps) {
    float2 v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float2 s = red[((0 * 5 + pr) * 4 + q) * 32 + lane];
#pragma unroll
        for (int g = 1; g < kC2Warps; ++g) {
            const float2 r = red[((g * 5 + pr) * 4 + q) * 32 + lane];
            s.x += r.x;
            s.y += r.y;
        }
        v[q] = s;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int oc = 2 * pr + h;
        const float bv = __ldg(bias + oc0 + oc);
        const float c0v = (h ? v[0].y : v[0].x) + bv, c1v = (h ? v[1].y : v[1].x) + bv;
        const float c2v = (h ? v[2].y : v[2].x) + bv, c3v = (h ? v[3].y : v[3].x) + bv;
        float best = c0v;
        int bi = 0;
        if (c1v > best) { best = c1v; bi = 1; }
        if (c2v > best) { best = c2v; bi = 2; }
        if (c3v > best) { best = c3v; bi = 3; }
        if (b < B) {
            const size_t oidx = ((size_t)b * 50 + oc0 + oc) * 16 + pos;
            out[oidx] = fmaxf(best, 0.f);
            idx[oidx] = (uint8_t)bi;
        }
    }
    }
}

}  // namespace b200mnist
