// Gradient finalisation, fused SGD+momentum (mnist.py:140,43) and the smoke-dist square.
#pragma once
#include "common.cuh"

namespace b200mnist {

// Sum of the partial buffers of the segment that covers flat index i (fixed order), or
// `fallback` when no segment covers it (the tensor was written to the bucket directly).
__device__ __forceinline__ float reduce_partials(const GradSegments &segs, int i, float fallback) {
#pragma unroll
    for (int s = 0; s < 6; ++s) {
        if (s < segs.count) {
            const GradSegment &sg = segs.seg[s];
            const int r = i - sg.off;
            if (r >= 0 && r < sg.n) {
                return sum_partials<kMaxGradSplit>(sg.part + r, sg.nsplit, (size_t)sg.stride);
            }
        }
    }
    return fallback;
}

// Writes the summed partials into the flat gradient bucket (the tensors that come out of
// batch-split kernels); everything else in the bucket is already final.
__global__ void __launch_bounds__(256) grad_finalize_kernel(float *__restrict__ grads, const GradSegments segs) {
    // segments cover [0, kOffF1W) (conv tensors) and [kOffF2W, kFlatElems) (fc2): skip fc1.
    int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= kOffF1W) i += kOffF2W - kOffF1W;
    if (i >= kFlatElems) return;
    const float cur = 0.f;
    bool covered = false;
#pragma unroll
    for (int s = 0; s < 6; ++s)
        if (s < segs.count && i >= segs.seg[s].off && i < segs.seg[s].off + segs.seg[s].n) covered = true;
    if (covered) grads[i] = reduce_partials(segs, i, cur);
}

// out[i] = sum_s part[s*stride + i]: the same reduction for a stand-alone tensor.
__global__ void __launch_bounds__(256)
reduce_partials_to_kernel(float *__restrict__ out, const float *__restrict__ part, int n, int nsplit, int stride) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    out[i] = sum_partials<kMaxGradSplit>(part + i, nsplit, (size_t)stride);
}

// One SGD+momentum step over a flat fp32 range, 4 elements per thread.  Rounding matches
// ATen's CPU kernels bit for bit: buf*mu and +g are separate roundings (mul_ then add_),
// p + (-lr)*buf is a single fused multiply-add (add_ with alpha).
template <bool FUSED_FINALIZE>
__global__ void __launch_bounds__(256)
sgd_momentum_kernel(float *__restrict__ p, float *__restrict__ g, float *__restrict__ m, int n, float neg_lr,
                    float mu, float grad_scale, int first_step, const GradSegments segs) {
    const int i4 = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (i4 >= n) return;
    if (i4 + 3 < n) {
        float4 gv = *reinterpret_cast<const float4 *>(g + i4);
        if (FUSED_FINALIZE) {
            gv.x = reduce_partials(segs, i4, gv.x);
            gv.y = reduce_partials(segs, i4 + 1, gv.y);
            gv.z = reduce_partials(segs, i4 + 2, gv.z);
            gv.w = reduce_partials(segs, i4 + 3, gv.w);
            *reinterpret_cast<float4 *>(g + i4) = gv;  // keep the bucket inspectable (p.grad)
        }
        float4 pv = *reinterpret_cast<const float4 *>(p + i4);
        float4 mv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!first_step) mv = *reinterpret_cast<const float4 *>(m + i4);
        float gs[4] = {gv.x, gv.y, gv.z, gv.w}, ms[4] = {mv.x, mv.y, mv.z, mv.w}, ps[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gj = __fmul_rn(gs[j], grad_scale);
            ms[j] = first_step ? gj : __fadd_rn(__fmul_rn(ms[j], mu), gj);
            ps[j] = __fmaf_rn(ms[j], neg_lr, ps[j]);
        }
        *reinterpret_cast<float4 *>(m + i4) = make_float4(ms[0], ms[1], ms[2], ms[3]);
        *reinterpret_cast<float4 *>(p + i4) = make_float4(ps[0], ps[1], ps[2], ps[3]);
    } else {
        for (int i = i4; i < n; ++i) {
            float gi = g[i];
            if (FUSED_FINALIZE) { gi = reduce_partials(segs, i, gi); g[i] = gi; }
            gi = __fmul_rn(gi, grad_scale);
            const float mi = first_step ? gi : __fadd_rn(__fmul_rn(m[i], mu), gi);
            m[i] = mi;
            p[i] = __fmaf_rn(mi, neg_lr, p[i]);
        }
    }
}

// Holds the stream busy for `ns` nanoseconds so that the launches that follow queue up behind
// it (profiling: removes host launch gaps from event-to-event kernel timings).
__global__ void delay_kernel(unsigned long long ns) {
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    } while (t - t0 < ns);
}

__global__ void __launch_bounds__(256) square_kernel(const float *__restrict__ in, float *__restrict__ out, long long n) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) out[i] = __fmul_rn(in[i], in[i]);
}

}  // namespace b200mnist
