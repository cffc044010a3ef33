// Shared definitions for the sm_100a kernels of libb200mnist.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200mnist {

// Flat parameter bucket: element offsets of the eight tensors (256-byte aligned).
constexpr int kNumTensors = 8;
constexpr int kOffC1W = 0, kOffC1B = 512, kOffC2W = 576, kOffC2B = 25600, kOffF1W = 25664, kOffF1B = 425664,
              kOffF2W = 426176, kOffF2B = 431232, kFlatElems = 431296;
constexpr int kTensorOff[kNumTensors] = {kOffC1W, kOffC1B, kOffC2W, kOffC2B, kOffF1W, kOffF1B, kOffF2W, kOffF2B};
constexpr int kTensorElems[kNumTensors] = {500, 20, 25000, 50, 400000, 500, 5000, 10};

// Per-image activation sizes (floats).
constexpr int kXElems = 784, kP1Elems = 2880, kP2Elems = 800, kHElems = 500, kClasses = 10;

// fc1's tcgen05 GEMMs read hi / lo copies of their operands (tc_gemm.cuh): the copy of p2 [B,800] has this leading
// dimension -- column 800 of the hi copy holds 1 (the bias-gradient column of the weight-gradient GEMM), the rest 0.
constexpr int kP2SplitLd = 832;
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

// Programmatic dependent launch, device side.  pdl_wait(): block until the preceding grid on the stream has
// completed and its writes are visible (returns at once for a normal launch).  pdl_trigger(): allow the next
// grid to be scheduled as soon as every CTA of this grid has got this far.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Wall-clock bound for device-side waits (mbarrier polls, cross-GPU flags): a wait that has not been satisfied
// after kSpinTimeoutNs traps -- the launch fails with an error instead of hanging the GPU.
constexpr unsigned long long kSpinTimeoutNs = 10ull * 1000 * 1000 * 1000;
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// system-scope release store / acquire load: flags that another GPU polls over NVLink (peer_exchange.cuh)
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// An early "ready" signal a step kernel can carry for the fused DDP exchange: thread t < n of CTA 0 stores
// *epoch_src + 1 to dst[t] (a flag word in rank t's peer window; nullptr = skip) once the kernel's inputs are complete.
struct EarlySignal {
    unsigned *dst[8];
    const unsigned *epoch_src;
    int n;
};
__device__ __forceinline__ void send_early_signal(const EarlySignal &sig) {
    const int t = threadIdx.x;
    if (blockIdx.x == 0 && blockIdx.y == 0 && t < sig.n && sig.dst[t])
        st_release_sys(sig.dst[t], *reinterpret_cast<const volatile unsigned *>(sig.epoch_src) + 1);
}

// "Resident input": a ring of batches that already sits in device memory plus a device-side cursor, so that a
// CUDA-graph replay of the step needs no copy and no host-side pointer: conv1's forward kernel reads slot *cursor
// (x + slot * x_stride floats, target + slot * t_stride elements), leaves copies in the step's ordinary input buffers
// (x_copy / t_copy: what the later kernels of the step read) and the end of the step advances the cursor (finish_step:
// the SGD kernel's last block, or one CTA of the conv1 weight-gradient launch on a host-fed ring).
// Host-fed rings (a copy stream refills the slots while the steps run): `landed` holds the sequence number of the
// newest batch that has arrived (written behind each copy, in copy-stream order) and state[1] counts the steps that
// have consumed the ring, so step n (1-based) waits until landed >= n -- on the device, which is what lets the
// compute stream carry nothing but the graph launches.  nullptr for a ring that is simply resident.
struct RingRef {
    const float *x;
    const long long *target;
    long long x_stride, t_stride;
    float *x_copy;
    long long *t_copy;
    unsigned *state;          // [0] cursor: the slot the next step reads; [1] steps done (8-byte aligned pair)
    const unsigned *landed;
    unsigned slots;
};
struct SpinClock {   // wall-clock bound for a spin loop: a producer that never arrives must not hang this GPU
    unsigned long long t0 = 0;
    unsigned n = 0;
    __device__ __forceinline__ void tick() {
        if ((++n & 1023u) == 0u) {
            const unsigned long long now = global_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > kSpinTimeoutNs) __trap();
        }
    }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum of up to MAX partial values p[0], p[stride], ..., in a fixed order.  All MAX loads are
// issued before the first add (a serial "v += p[s*stride]" walk costs one full memory latency
// per partial: at 8-148 partials that dominated the v1 step).  Missing partials read as +0,
// which leaves the sum bit-identical.
template <int MAX>
__device__ __forceinline__ float sum_partials(const float *p, int nsplit, size_t stride) {
    float v[MAX];
#pragma unroll
    for (int s = 0; s < MAX; ++s) v[s] = (s < nsplit) ? __ldcg(p + (size_t)s * stride) : 0.f;
    float acc = v[0];
#pragma unroll
    for (int s = 1; s < MAX; ++s) acc += v[s];
    return acc;
}
constexpr int kMaxGradSplit = 296;  // upper bound of every batch-split weight-gradient kernel (2 CTAs x 148 SMs)

// One gradient tensor whose value is the sum of `nsplit` partial buffers
// (part + s*stride, s < nsplit), landing at flat offset `off` .. off+n.
struct GradSegment {
    int off, n, nsplit, stride;
    const float *part;
};
struct GradSegments {
    GradSegment seg[6];
    int count;
};

}  // namespace b200mnist
