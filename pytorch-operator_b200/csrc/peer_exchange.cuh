// DDP gradient averaging + SGD over NVLink peer memory (replaces, for the fused training step, DDP's all_reduce
// of mnist.py:135-138 followed by optimizer.step() of mnist.py:43).
//
// Every rank owns one PEER WINDOW -- a single device allocation that every other rank has mapped (CUDA IPC):
//
//   [ flag page 4096 B | parameter bucket | gradient-sum bucket (p.grad) | staging bucket 0 | staging bucket 1 ]
//
// The backward kernels of step s write this rank's gradient into staging bucket (s & 1).  Two kernels exchange it:
//
//   peer_fc1_kernel  (two-shot, fc1 = 93 % of the bucket; runs on a side stream WHILE the conv backward computes)
//     rank r owns the r-th slice of fc1's range.  It waits until every rank's fc1 gradient is staged, reads its
//     slice from all `world` staging buckets (P2P loads over NVLink 5 / NVSwitch), sums in RANK ORDER, applies
//     momentum-SGD with 1/world folded in (the slice's momentum lives on the owner only), and stores the updated
//     parameters and the gradient sum into EVERY rank's window (P2P stores).  One owner per element: replicas stay
//     bit-identical by construction, and a rank moves 3 x 1/world of the range instead of (world-1) x the range.
//     64-thread CTAs with at most 64 registers, so they fit on the SMs beside the conv backward's CTAs instead of
//     taking slots from its single wave.
//   peer_rest_kernel (one-shot, the ~31 k remaining elements; the tail of the step)
//     waits until every rank's remaining gradient is staged AND every rank's fc1 slice has landed in this rank's
//     window, reads all staging buckets, sums in rank order (bit-identical on every rank) and applies the update.
//
// All peer traffic is plain ld.global / st.global (no TMA, no bulk copies).  Flags are monotonically increasing
// epochs (one per step) kept in the flag page, so both kernels replay from a CUDA graph with constant arguments.
// There is no "done reading" barrier: a staging bucket is rewritten two steps later, and a rank can only get there
// after every peer has passed the barriers of the step in between.  All spins are bounded by wall-clock time
// (10 s, then trap instead of hanging the GPU).
#pragma once
#include "common.cuh"

namespace b200mnist {

constexpr int kMaxPeers = 8;
constexpr size_t kPeerFlagBytes = 4096;
constexpr size_t kBucketBytes = (size_t)kFlatElems * 4;
constexpr size_t kPeerWindowBytes = kPeerFlagBytes + 4 * kBucketBytes;
// flag page, u32 indices.  Written by peers: one slot per source rank, one 128-byte line per kind.
constexpr int kFlagFc1Ready = 0, kFlagFc1Pushed = 32, kFlagRestReady = 64;
// local words: step epochs and CTA tickets of the two kernels; a word the host uses to verify the mapping
constexpr int kFlagEpochFc1 = 128, kFlagTicketFc1 = 129, kFlagEpochRest = 160, kFlagTicketRest = 161, kFlagProbe = 192;

struct PeerWin {
    char *base[kMaxPeers];   // every rank's window, as mapped in THIS process; base[rank] is the local one
    int rank, world;
};
__device__ __forceinline__ unsigned *win_flags(char *b) { return reinterpret_cast<unsigned *>(b); }
__device__ __forceinline__ float *win_params(char *b) { return reinterpret_cast<float *>(b + kPeerFlagBytes); }
__device__ __forceinline__ float *win_gsum(char *b) { return reinterpret_cast<float *>(b + kPeerFlagBytes + kBucketBytes); }
__device__ __forceinline__ float *win_staging(char *b, int parity) {
    return reinterpret_cast<float *>(b + kPeerFlagBytes + (size_t)(2 + parity) * kBucketBytes);
}

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
// L2-only 128-bit load (never served from this SM's L1: the data may have been written by another GPU)
__device__ __forceinline__ float4 ld_cg4(const float *p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st4(float *p, float4 v) { *reinterpret_cast<float4 *>(p) = v; }

// g: summed gradient; returns the new parameter, updates mv in place -- same roundings as sgd_momentum_kernel
__device__ __forceinline__ float4 sgd4(float4 g, float4 &mv, float4 pv, float neg_lr, float mu, float scale, int first_step) {
    float gs[4] = {g.x, g.y, g.z, g.w}, ms[4] = {mv.x, mv.y, mv.z, mv.w}, ps[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float gj = __fmul_rn(gs[j], scale);
        ms[j] = first_step ? gj : __fadd_rn(__fmul_rn(ms[j], mu), gj);
        ps[j] = __fmaf_rn(ms[j], neg_lr, ps[j]);
    }
    mv = make_float4(ms[0], ms[1], ms[2], ms[3]);
    return make_float4(ps[0], ps[1], ps[2], ps[3]);
}

// Last CTA of the grid (ticket) runs `tail`; every CTA's earlier writes are ordered before it at system scope.
template <typename F>
__device__ __forceinline__ void last_cta_then(unsigned *ticket, F tail) {
    __shared__ bool is_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!is_last) return;
    __threadfence_system();
    tail();
}

// set-up check: out[r] = the probe word of rank r's window, read through this process's mapping (the path the
// exchange kernels use)
__global__ void peer_probe_kernel(const __grid_constant__ PeerWin pw, unsigned *out) {
    const int t = threadIdx.x;
    if (t < pw.world) out[t] = ld_acquire_sys(win_flags(pw.base[t]) + kFlagProbe);
}

constexpr int kPeerFc1Threads = 64;
constexpr int kFc1Vec4 = (kOffF2W - kOffF1W) / 4;   // fc1.weight + fc1.bias (+ pad): 100,128 float4

// (64, 16): 16 CTAs x 64 threads per SM caps the kernel at 64 registers; with that ceiling ptxas issues all
// WORLD + 2 loads of an element before the first add (one NVLink round trip per element, not three)
template <int WORLD>
__global__ void __launch_bounds__(kPeerFc1Threads, 16)
peer_fc1_kernel(const __grid_constant__ PeerWin pw, float *__restrict__ m, int parity, float neg_lr, float mu,
                float scale, int first_step) {
    const int t = threadIdx.x, rank = pw.rank;
    unsigned *flags = win_flags(pw.base[rank]);
    const unsigned epoch = *reinterpret_cast<volatile unsigned *>(flags + kFlagEpochFc1) + 1;  // bumped by the last CTA only
    // ---- "my fc1 gradient is staged" -> every peer; wait for theirs ----
    if (blockIdx.x == 0 && t < WORLD && t != rank) {
        __threadfence_system();
        st_release_sys(win_flags(pw.base[t]) + kFlagFc1Ready + rank, epoch);
    }
    if (t < WORLD && t != rank) spin_until(flags + kFlagFc1Ready + t, epoch);
    __syncthreads();
    // ---- my slice: reduce in rank order, update, push parameters + gradient sum to every window ----
    constexpr int per = (kFc1Vec4 + WORLD - 1) / WORLD;
    const int lo = rank * per, hi = min(kFc1Vec4, lo + per);
    float *p_loc = win_params(pw.base[rank]);
    for (int v = lo + blockIdx.x * kPeerFc1Threads + t; v < hi; v += gridDim.x * kPeerFc1Threads) {
        const int i = kOffF1W + 4 * v;
        float4 g[WORLD];
#pragma unroll
        for (int r = 0; r < WORLD; ++r) g[r] = ld_cg4(win_staging(pw.base[r], parity) + i);
        const float4 pv = ld_cg4(p_loc + i);
        float4 mv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!first_step) mv = ld_cg4(m + i);
        float4 s = g[0];
#pragma unroll
        for (int r = 1; r < WORLD; ++r) { s.x += g[r].x; s.y += g[r].y; s.z += g[r].z; s.w += g[r].w; }
        const float4 pn = sgd4(s, mv, pv, neg_lr, mu, scale, first_step);
        st4(m + i, mv);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            st4(win_params(pw.base[r]) + i, pn);
            st4(win_gsum(pw.base[r]) + i, s);
        }
    }
    // ---- "my slice has landed everywhere" -> every peer (consumed by their peer_rest_kernel) ----
    last_cta_then(flags + kFlagTicketFc1, [&] {
        if (t < WORLD && t != rank) st_release_sys(win_flags(pw.base[t]) + kFlagFc1Pushed + rank, epoch);
        __syncthreads();
        if (t == 0) {
            flags[kFlagTicketFc1] = 0u;
            flags[kFlagEpochFc1] = epoch;
        }
    });
}

constexpr int kRestVec4 = (kOffF1W + (kFlatElems - kOffF2W)) / 4;   // conv1, conv2, fc2 (+ pads): 7,696 float4
constexpr int kPeerRestThreads = 128;

// (128, 1): no register ceiling, so all WORLD loads of an element are in flight together (this kernel is the tail
// of the step: its latency is one barrier + one NVLink round trip)
template <int WORLD>
__global__ void __launch_bounds__(kPeerRestThreads, 1)
peer_rest_kernel(const __grid_constant__ PeerWin pw, float *__restrict__ m, int parity, float neg_lr, float mu,
                 float scale, int first_step) {
    const int t = threadIdx.x, rank = pw.rank;
    unsigned *flags = win_flags(pw.base[rank]);
    float *p = win_params(pw.base[rank]);
    const int v = blockIdx.x * kPeerRestThreads + t;
    int i = 4 * v;
    if (i >= kOffF1W) i += kOffF2W - kOffF1W;
    const bool mine = v < kRestVec4;
    // local operands that do not depend on the preceding kernel: their latency overlaps its tail and the barrier
    float4 pv = make_float4(0.f, 0.f, 0.f, 0.f), mv = pv;
    if (mine) {
        pv = *reinterpret_cast<const float4 *>(p + i);
        if (!first_step) mv = *reinterpret_cast<const float4 *>(m + i);
    }
    pdl_wait();   // the gradient finalisation that precedes this kernel on the stream has completed
    const unsigned epoch = *reinterpret_cast<volatile unsigned *>(flags + kFlagEpochRest) + 1;
    if (blockIdx.x == 0 && t < WORLD && t != rank) {
        __threadfence_system();
        st_release_sys(win_flags(pw.base[t]) + kFlagRestReady + rank, epoch);
    }
    if (t < WORLD && t != rank) {
        spin_until(flags + kFlagRestReady + t, epoch);
        // every peer's fc1 slice of THIS step is in my parameter bucket once this kernel (= the step) completes;
        // peer_fc1_kernel and this kernel run once per step each, so their epochs agree
        spin_until(flags + kFlagFc1Pushed + t, epoch);
    }
    __syncthreads();
    if (mine) {
        float4 g[WORLD];
#pragma unroll
        for (int r = 0; r < WORLD; ++r) g[r] = ld_cg4(win_staging(pw.base[r], parity) + i);
        float4 s = g[0];
#pragma unroll
        for (int r = 1; r < WORLD; ++r) { s.x += g[r].x; s.y += g[r].y; s.z += g[r].z; s.w += g[r].w; }
        st4(win_gsum(pw.base[rank]) + i, s);
        const float4 pn = sgd4(s, mv, pv, neg_lr, mu, scale, first_step);
        st4(m + i, mv);
        st4(p + i, pn);
    }
    last_cta_then(flags + kFlagTicketRest, [&] {
        if (t == 0) {
            flags[kFlagTicketRest] = 0u;
            flags[kFlagEpochRest] = epoch;
        }
    });
}

}  // namespace b200mnist
