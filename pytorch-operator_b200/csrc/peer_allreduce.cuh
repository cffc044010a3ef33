// DDP gradient averaging + SGD as ONE kernel over NVLink peer memory (replaces, for the fused training step,
// DDP's all_reduce of mnist.py:135-138 followed by optimizer.step() of mnist.py:43).
//
// Every rank keeps its step gradient in a symmetric-memory bucket that all peers have mapped.  The kernel
//   1. tells the peers "my gradient is ready" and waits until all of theirs are (one flag per peer, written
//      with st.release.sys into the peer's signal pad, polled locally with ld.acquire.sys);
//   2. for its share of the bucket reads the value from every rank's bucket -- P2P loads over NVLink 5 /
//      NVSwitch -- sums them in RANK ORDER (so all ranks compute bit-identical sums and the replicas never
//      drift), writes the sum to the local gradient bucket (p.grad) and applies the momentum-SGD update
//      with 1/world_size folded in as grad_scale;
//   3. the last CTA to finish tells the peers "I am done reading" and waits for the same from them, so that
//      when the kernel completes nobody is still reading this rank's bucket and the next backward may
//      overwrite it.
// The exchange is one-shot (each rank reads (ws-1) x 1.7 MB): at this message size the all-reduce is latency
// bound, and a one-shot read has a single barrier + a single NVLink round trip on the critical path.
// Flags carry a monotonically increasing epoch kept in device memory, so the kernel is CUDA-graph replayable
// with constant arguments.  All spins are bounded by wall-clock time (10 s, then trap instead of hanging the GPU).
#pragma once
#include "common.cuh"

namespace b200mnist {

constexpr int kMaxPeers = 8;

struct PeerArgs {
    const float *grads[kMaxPeers];   // every rank's symmetric gradient bucket, as mapped in THIS process
    unsigned *signals[kMaxPeers];    // every rank's signal pad (>= 64 u32), as mapped in THIS process
    int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// wait until *p has reached `want` (epochs only grow; signed difference is wrap-safe)
__device__ __forceinline__ void spin_until(const unsigned *p, unsigned want) {
    unsigned long long t0 = 0;
    for (unsigned spin = 0;; ++spin) {
        if ((int)(ld_acquire_sys(p) - want) >= 0) return;
        if ((spin & 1023u) == 1023u) {   // a peer that never arrives (crashed rank) must not hang this GPU
            const unsigned long long now = global_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > kSpinTimeoutNs) __trap();
        }
    }
}

// signal pad layout (u32): [0..7] "gradient ready" flags by source rank, [8..15] "done reading" flags.
__global__ void __launch_bounds__(256)
allreduce_sgd_kernel(const PeerArgs pa, float *__restrict__ p, float *__restrict__ m, float *__restrict__ gsum,
                     int lo, int hole_lo, int hole_hi, int n, float neg_lr, float mu, float grad_scale, int first_step,
                     unsigned *epoch_ctr, unsigned *ticket) {
    __shared__ bool is_last;
    const unsigned epoch = *reinterpret_cast<volatile unsigned *>(epoch_ctr) + 1;  // bumped by the last CTA only
    const int t = threadIdx.x;
    unsigned *my_pad = pa.signals[pa.rank];
    // one float4 per thread (a single NVLink round trip per CTA) over [lo, n) minus the hole [hole_lo, hole_hi)
    int i4 = lo + (blockIdx.x * 256 + t) * 4;
    if (i4 >= hole_lo) i4 += hole_hi - hole_lo;
    const bool mine = i4 < n;
    // local operands first: their latency overlaps the barrier
    float4 pv = make_float4(0.f, 0.f, 0.f, 0.f), mv = pv, own = pv;
    if (mine) {
        pv = *reinterpret_cast<const float4 *>(p + i4);
        if (!first_step) mv = *reinterpret_cast<const float4 *>(m + i4);
        own = *reinterpret_cast<const float4 *>(pa.grads[pa.rank] + i4);
    }
    // ---- 1. ready barrier ----
    if (blockIdx.x == 0 && t < pa.world && t != pa.rank) st_release_sys(pa.signals[t] + pa.rank, epoch);
    if (t < pa.world && t != pa.rank) spin_until(my_pad + t, epoch);
    __syncthreads();
    // ---- 2. reduce in rank order + SGD ----
    if (mine) {
        float4 v[kMaxPeers];
#pragma unroll
        for (int r = 0; r < kMaxPeers; ++r)
            if (r < pa.world) v[r] = (r == pa.rank) ? own : *reinterpret_cast<const float4 *>(pa.grads[r] + i4);
        float4 s = v[0];
#pragma unroll
        for (int r = 1; r < kMaxPeers; ++r)
            if (r < pa.world) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
        *reinterpret_cast<float4 *>(gsum + i4) = s;
        float gs[4] = {s.x, s.y, s.z, s.w}, ms[4] = {mv.x, mv.y, mv.z, mv.w}, ps[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gj = __fmul_rn(gs[j], grad_scale);
            ms[j] = first_step ? gj : __fadd_rn(__fmul_rn(ms[j], mu), gj);
            ps[j] = __fmaf_rn(ms[j], neg_lr, ps[j]);
        }
        *reinterpret_cast<float4 *>(m + i4) = make_float4(ms[0], ms[1], ms[2], ms[3]);
        *reinterpret_cast<float4 *>(p + i4) = make_float4(ps[0], ps[1], ps[2], ps[3]);
    }
    // ---- 3. done barrier, by the last CTA ----
    __syncthreads();
    if (t == 0) {
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    if (t < pa.world && t != pa.rank) {
        st_release_sys(pa.signals[t] + 8 + pa.rank, epoch);
        spin_until(my_pad + 8 + t, epoch);
    }
    __syncthreads();
    if (t == 0) {
        *ticket = 0u;
        *epoch_ctr = epoch;
    }
}

}  // namespace b200mnist
