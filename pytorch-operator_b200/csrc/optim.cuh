// Gradient finalisation, fused SGD+momentum (mnist.py:140,43) and the smoke-dist square.
#pragma once
#include "common.cuh"
#include "conv_bwd.cuh"

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
                float v = sum_partials<32>(sg.part + r, min(sg.nsplit, 32), (size_t)sg.stride);
                for (int s0 = 32; s0 < sg.nsplit; s0 += 32)
                    v += sum_partials<32>(sg.part + (size_t)s0 * sg.stride + r, min(sg.nsplit - s0, 32), (size_t)sg.stride);
                return v;
            }
        }
    }
    return fallback;
}

// Deterministic sum of the per-row losses written by head_kernel (and of the per-row "correct"
// flags when want_correct), by the first 256 threads of one CTA: strided partial sums, then a fixed tree.
__device__ __forceinline__ void reduce_rows_block(const float *rows, int B, float scale, int want_correct,
                                                  float *stats_out) {
    __shared__ float red[256], red2[256];
    const int t = threadIdx.x;
    float s = 0.f, c = 0.f;
    if (t < 256) {   // (a larger CTA may call this: its extra threads only take part in the barriers)
        for (int r = t; r < B; r += 256) {
            s += __ldcg(rows + r);
            if (want_correct) c += __ldcg(rows + B + r);
        }
        red[t] = s;
        red2[t] = c;
    }
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (t < o) { red[t] += red[t + o]; red2[t] += red2[t + o]; }
        __syncthreads();
    }
    if (t == 0) {
        stats_out[0] = red[0] * scale;
        if (want_correct) stats_out[1] = red2[0];
    }
}

struct LossReduce {   // rows == nullptr: nothing to do
    const float *rows;
    float *out;
    int B;
    float scale;
    // resident-input steps (RingRef): the block that reduces the loss also closes the step on the ring
    unsigned *state;                    // [0] cursor -> next slot, [1] steps done + 1
    unsigned slots;
    unsigned long long *report_host;    // host-fed rings: report_host[slot] (PINNED HOST memory) = {step number, loss bits}
};
// End of the step (the SGD kernel's last block, or one CTA of the conv1 weight-gradient launch on a host-fed ring, or
// the last block of the fused DDP tail): the scalar loss, then the ring bookkeeping.  The report to the host is ONE
// 8-byte store (sequence number in the high word, the loss in the low one): the host sees both or neither, so no
// system-scope fence has to order a "done" flag behind the loss at the very end of the step.
__device__ __forceinline__ void finish_step(const LossReduce &lr) {
    if (lr.rows) reduce_rows_block(lr.rows, lr.B, lr.scale, 0, lr.out);
    if (lr.state && threadIdx.x == 0) {
        const unsigned slot = lr.state[0], done = lr.state[1] + 1;
        lr.state[0] = slot + 1 >= lr.slots ? 0u : slot + 1;
        lr.state[1] = done;
        if (lr.report_host) {
            const float loss = lr.rows ? *reinterpret_cast<volatile float *>(lr.out) : 0.f;   // written by this thread above
            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(lr.report_host + slot),
                         "l"(((unsigned long long)done << 32) | (unsigned long long)__float_as_uint(loss)) : "memory");
        }
    }
}

// conv1's weight gradient with one extra grid column whose first CTA closes the step on a HOST-FED ring early
// (finish_step: loss, ring bookkeeping, report to the host).  Everything finish_step needs exists since the loss head
// ran, and only the step's first kernel reads the ring, so nothing has to wait for the SGD kernel -- while the report,
// a store into host memory, would otherwise sit at the very end of the step, where the kernel boundary waits for it to
// be acknowledged across PCIe.
__global__ void __launch_bounds__(kC1wThreads)
conv1_wgrad_finish_kernel(const float *__restrict__ x, const float *__restrict__ dp1, const float *__restrict__ p1,
                          const uint8_t *__restrict__ i1, float *__restrict__ part, int B, const LossReduce lr) {
    pdl_wait();
    if (blockIdx.x == gridDim.x - 1) {
        if (blockIdx.y == 0) finish_step(lr);
        return;
    }
    conv1_wgrad_body(x, dp1, p1, i1, part, B, blockIdx.x, gridDim.x - 1, blockIdx.y);
}

// Writes the summed partials into the flat gradient bucket (the tensors that come out of
// batch-split kernels); everything else in the bucket is already final.
__global__ void __launch_bounds__(256)
grad_finalize_kernel(float *__restrict__ grads, const GradSegments segs, const LossReduce lr) {

    pdl_wait();
    // segments cover [0, kOffF1W) (conv tensors) and [kOffF2W, kFlatElems) (fc2): skip fc1.
    if (blockIdx.x == gridDim.x - 1) {  // last block: the scalar loss from the per-row losses
        if (lr.rows) reduce_rows_block(lr.rows, lr.B, lr.scale, 0, lr.out);
        return;
    }
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
    float v = sum_partials<32>(part + i, min(nsplit, 32), (size_t)stride);
    for (int s0 = 32; s0 < nsplit; s0 += 32)
        v += sum_partials<32>(part + (size_t)s0 * stride + i, min(nsplit - s0, 32), (size_t)stride);
    out[i] = v;
}

// One SGD+momentum step over a flat fp32 range, 4 elements per thread (b200mnist_sgd_momentum: the NCCL
// DDP path and the autograd-bridge loop).  Rounding matches ATen's CPU kernels bit for bit: buf*mu and +g are
// separate roundings (mul_ then add_), p + (-lr)*buf is a single fused multiply-add (add_ with alpha).
__global__ void __launch_bounds__(256)
sgd_momentum_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, int n, float neg_lr,
                    float mu, float grad_scale, int first_step) {
    const int i4 = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (i4 >= n) return;
    if (i4 + 3 < n) {
        const float4 gv = *reinterpret_cast<const float4 *>(g + i4);
        const float4 pv = *reinterpret_cast<const float4 *>(p + i4);
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
            const float gi = __fmul_rn(g[i], grad_scale);
            const float mi = first_step ? gi : __fadd_rn(__fmul_rn(m[i], mu), gi);
            m[i] = mi;
            p[i] = __fmaf_rn(mi, neg_lr, p[i]);
        }
    }
}

// The whole end of a single-process training step in one launch (mnist.py:43 + the tail of :42):
//   blocks [0, kFusedCovBlocks)   : the ~31 k bucket elements whose gradients arrive as batch-split
//                                   partials (conv1, conv2, fc2): one element per thread, all partial
//                                   loads in flight at once, summed in a fixed order, then the update;
//   next kFusedFc1Blocks blocks   : fc1 (93 % of the bucket, gradient already final): float4 per thread;
//   last block                    : the scalar loss from the per-row losses.
constexpr int kFusedCovElems = kOffF1W + (kFlatElems - kOffF2W);           // 30,784
constexpr int kFusedCovBlocks = (kFusedCovElems + 255) / 256;              // 121
constexpr int kFusedFc1Blocks = ((kOffF2W - kOffF1W) / 4 + 255) / 256;     // 392

__device__ __forceinline__ void sgd_update1(float *p, float *m, int i, float gi, float neg_lr, float mu, float scale,
                                            int first_step) {
    gi = __fmul_rn(gi, scale);
    const float mi = first_step ? gi : __fadd_rn(__fmul_rn(m[i], mu), gi);
    m[i] = mi;
    p[i] = __fmaf_rn(mi, neg_lr, p[i]);
}

__global__ void __launch_bounds__(256)
sgd_fused_kernel(float *__restrict__ p, float *__restrict__ g, float *__restrict__ m, float neg_lr, float mu,
                 float grad_scale, int first_step, const GradSegments segs, const LossReduce lr) {

    pdl_wait();
    const int blk = blockIdx.x, t = threadIdx.x;
    if (blk < kFusedCovBlocks) {
        int i = blk * 256 + t;
        if (i >= kFusedCovElems) return;
        if (i >= kOffF1W) i += kOffF2W - kOffF1W;
        const float pv = p[i];                       // issued before the partial loads complete
        const float mv = first_step ? 0.f : m[i];
        const float gi = reduce_partials(segs, i, g[i]);
        g[i] = gi;                                   // keep the bucket inspectable (p.grad)
        const float gs = __fmul_rn(gi, grad_scale);
        const float mi = first_step ? gs : __fadd_rn(__fmul_rn(mv, mu), gs);
        m[i] = mi;
        p[i] = __fmaf_rn(mi, neg_lr, pv);
    } else if (blk < kFusedCovBlocks + kFusedFc1Blocks) {
        const int i4 = kOffF1W + ((blk - kFusedCovBlocks) * 256 + t) * 4;
        if (i4 >= kOffF2W) return;
        const float4 gv = *reinterpret_cast<const float4 *>(g + i4);
        const float4 pv = *reinterpret_cast<const float4 *>(p + i4);
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
        finish_step(lr);
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

// Register-resident FFMA2 loop: the fp32 roofline the SIMT kernels are measured against (bench.py).  Each thread
// keeps 16 independent float2 accumulators (32 FMA chains), `iters` passes of 16 packed FMAs, no memory traffic;
// 2 x 2 x 16 x iters FLOP per thread.  The result is stored so nothing is optimised away.
__global__ void __launch_bounds__(256) fma_peak_kernel(float *out, int iters, float a, float b) {
    float2 acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = make_float2(a + (float)j, b - (float)j + (float)threadIdx.x);
    const float2 x = make_float2(a, b), y = make_float2(b, a);
#pragma unroll 8
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = __ffma2_rn(acc[j], x, y);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += acc[j].x + acc[j].y;
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

// Input pipeline (mnist.py:118-125): gather a batch of 28x28 uint8 images by index and apply ToTensor +
// Normalize -- x/255, then (x - mean)/std, each an IEEE fp32 rounding like the reference's CPU ops -- plus the
// matching labels.  One CTA per image, 196 threads x 4 pixels (32-bit load, 128-bit store).
__global__ void __launch_bounds__(196)
gather_normalize_u8_kernel(const uint8_t *__restrict__ images, const long long *__restrict__ labels,
                           const long long *__restrict__ indices, int n_images, float mean, float stdv,
                           float *__restrict__ data, long long *__restrict__ target) {
    const int b = blockIdx.x, t = threadIdx.x;
    long long src = indices[b];
    if (src < 0 || src >= n_images) src = 0;  // never read out of bounds
    const uchar4 px = *reinterpret_cast<const uchar4 *>(images + (size_t)src * 784 + 4 * t);
    float4 v;
    v.x = __fdiv_rn(__fsub_rn(__fdiv_rn((float)px.x, 255.f), mean), stdv);
    v.y = __fdiv_rn(__fsub_rn(__fdiv_rn((float)px.y, 255.f), mean), stdv);
    v.z = __fdiv_rn(__fsub_rn(__fdiv_rn((float)px.z, 255.f), mean), stdv);
    v.w = __fdiv_rn(__fsub_rn(__fdiv_rn((float)px.w, 255.f), mean), stdv);
    *reinterpret_cast<float4 *>(data + (size_t)b * 784 + 4 * t) = v;
    if (t == 0) target[b] = labels[src];
}

__global__ void __launch_bounds__(256) square_kernel(const float *__restrict__ in, float *__restrict__ out, long long n) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) out[i] = __fmul_rn(in[i], in[i]);
}

}  // namespace b200mnist
