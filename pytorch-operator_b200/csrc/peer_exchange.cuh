// DDP gradient averaging + SGD over NVLink peer memory, inside the training step's own kernels (replaces DDP's
// all_reduce of mnist.py:135-138 followed by optimizer.step() of mnist.py:43).
//
// Every rank owns one PEER WINDOW -- a single device allocation that every other rank has mapped (CUDA IPC):
//
//   [ flag page 4096 B | parameter bucket | gradient-sum bucket (p.grad) | staging bucket 0 | staging bucket 1 |
//     receive area: 2 parities x 8 source ranks x 30,784 (value, epoch) pairs ]
//
// The step runs on ONE stream; the exchange has no kernel of its own:
//
//   fc1 (93 % of the bucket): two-shot, in the launch of conv1's weight-gradient kernel.
//     The fc backward of step s wrote this rank's fc1 gradient into staging bucket (s & 1).  conv1's weight gradient
//     needs only 2 x min(B, 148) CTAs, so the same launch carries extra CTAs (scheduled first) that exchange fc1:
//     rank r owns the r-th slice of fc1's range; it waits until every rank's gradient is staged, reads its slice
//     from all `world` staging buckets (P2P loads over NVLink 5 / NVSwitch), sums in RANK ORDER, applies
//     momentum-SGD with 1/world folded in (the slice's momentum lives on the owner only) and stores the new
//     parameters and the gradient sum into EVERY rank's window (P2P stores).  One owner per element: replicas stay
//     bit-identical by construction, and a rank moves 3/world of the range instead of (world-1) x the range.
//   conv1, conv2, fc2 (the remaining ~31 k elements): one-shot, in the step's last kernel (peer_tail_kernel).
//     Each thread sums the batch-split partials of one element, PUSHES {value, epoch} as one 8-byte store into every
//     peer's receive area and keeps its value in a register; then it polls its own receive slots until each carries
//     this step's epoch (the data is its own arrival flag: no fence, no separate flag, one NVLink one-way trip on
//     the critical path), adds the values in rank order (bit-identical on every rank) and applies the update.
//
// Measured on 2 GPUs (tools/exchange_probe.py): a system-scope fence issued by every warp of a CTA costs ~12 us
// (the fences of an SM queue up), which is why nothing on the critical path fences: the slice stores are ordered
// before the "landed" flag by the kernel boundary (the flag is raised by the NEXT kernel of the stream), and the
// tail needs no ordering at all.
// All peer traffic is plain ld.global / st.global.  Epochs (one per step, monotonically increasing) live in the
// flag page, so the kernels replay from a CUDA graph with constant arguments: the conv1 kernel's exchange CTAs use
// epoch_a + 1 and store it to epoch_b; the tail kernel uses epoch_b and stores it to epoch_a -- each counter is
// written while no kernel that reads it is running, so no CTA ticket is needed.  A staging bucket / receive slot is
// rewritten two steps later, and a rank only gets there after every peer has passed the barrier of the step in
// between.  All spins are bounded by wall-clock time (10 s, then trap instead of hanging the GPU).
#pragma once
#include "common.cuh"
#include "conv_bwd.cuh"
#include "optim.cuh"

namespace b200mnist {

constexpr int kMaxPeers = 8;
constexpr size_t kPeerFlagBytes = 4096;
constexpr size_t kBucketBytes = (size_t)kFlatElems * 4;
constexpr size_t kRecvSlotBytes = ((size_t)kFusedCovElems * 8 + 255) / 256 * 256;   // (value, epoch) pairs of one source
constexpr size_t kPeerWindowBytes = kPeerFlagBytes + 4 * kBucketBytes + 2 * kMaxPeers * kRecvSlotBytes;
// flag page, u32 indices.  Written by peers: one slot per source rank, one 128-byte line per kind.
constexpr int kFlagFc1Ready = 0, kFlagFc1Pushed = 32;
// local words: the two epoch counters; a word the host uses to verify the mapping
constexpr int kFlagEpochA = 128, kFlagEpochB = 160, kFlagProbe = 192;
// telemetry (local, u64 at byte 2048 + 8*k): globaltimer stamps of the latest step -- 0 exchange CTAs start, 1 their
// barrier passed, 2 exchange CTA 0 done, 3 tail kernel start, 4 its pushes issued, 5 its sum complete (block 0),
// 6 the previous step's slot 5 (5 - 6 = the step period).  Read by ddp.exchange_telemetry().
constexpr int kStampByteOffset = 2048;
__device__ __forceinline__ void stamp(unsigned *flags, int slot) {
    reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(flags) + kStampByteOffset)[slot] = global_ns();
}

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
__device__ __forceinline__ unsigned long long *win_recv(char *b, int parity, int src) {
    return reinterpret_cast<unsigned long long *>(b + kPeerFlagBytes + 4 * kBucketBytes +
                                                  (size_t)(parity * kMaxPeers + src) * kRecvSlotBytes);
}

// one naturally aligned 8-byte access: single-copy atomic, so a reader sees the value and its epoch together
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// wait until *p has reached `want` (epochs only grow; signed difference is wrap-safe)
__device__ __forceinline__ void spin_until(const unsigned *p, unsigned want) {
    SpinClock clk;
    while ((int)(ld_acquire_sys(p) - want) < 0) clk.tick();
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

// set-up check: out[r] = the probe word of rank r's window, read through this process's mapping (the path the
// exchange kernels use)
__global__ void peer_probe_kernel(const __grid_constant__ PeerWin pw, unsigned *out) {
    const int t = threadIdx.x;
    if (t < pw.world) out[t] = ld_acquire_sys(win_flags(pw.base[t]) + kFlagProbe);
}

constexpr int kFc1Vec4 = (kOffF2W - kOffF1W) / 4;   // fc1.weight + fc1.bias (+ pad): 100,128 float4

// The fc1 slice exchange as run by exchange CTA `cta` of `ncta`.
template <int WORLD>
__device__ __forceinline__ void peer_fc1_slice(const PeerWin &pw, float *__restrict__ m, int parity, float neg_lr,
                                               float mu, float scale, int first_step, const int cta, const int ncta,
                                               const int presignaled) {
    const int t = threadIdx.x, rank = pw.rank, nthr = blockDim.x;
    unsigned *flags = win_flags(pw.base[rank]);
    const unsigned epoch = *reinterpret_cast<volatile unsigned *>(flags + kFlagEpochA) + 1;   // written by the tail kernel
    if (cta == 0 && t == 0) {
        stamp(flags, 0);
        flags[kFlagEpochB] = epoch;   // read by this step's tail kernel, which starts after this kernel has completed
    }
    // ---- "my fc1 gradient is staged" -> every peer; wait for theirs.  (The gradient was written by an earlier
    //      kernel of this stream, so it is ordered before this release store.  At small batch the conv2 backward
    //      kernel has already sent this flag ~20 us ago -- EarlySignal, `presignaled` -- so that nobody waits here
    //      for a release fence and a flag's flight time.)
    if (!presignaled && cta == 0 && t < WORLD && t != rank)
        st_release_sys(win_flags(pw.base[t]) + kFlagFc1Ready + rank, epoch);
    if (t < WORLD && t != rank) spin_until(flags + kFlagFc1Ready + t, epoch);
    __syncthreads();
    if (cta == 0 && t == 0) stamp(flags, 1);
    // ---- my slice: reduce in rank order, update, push parameters + gradient sum to every window ----
    constexpr int per = (kFc1Vec4 + WORLD - 1) / WORLD;
    const int lo = rank * per, hi = min(kFc1Vec4, lo + per);
    float *p_loc = win_params(pw.base[rank]);
    for (int v = lo + cta * nthr + t; v < hi; v += ncta * nthr) {
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
    // no fence, no flag here: the tail kernel, which starts when this grid has completed, tells the peers that the
    // slice has landed
    if (cta == 0 && t == 0) stamp(flags, 2);
}

// conv1's weight gradient + the fc1 slice exchange in one launch: grid (nx + batch slices, 2); CTAs with
// blockIdx.x < nx exchange, the others compute the weight gradient (conv1_wgrad_body).
template <int WORLD>
__global__ void __launch_bounds__(kC1wThreads)
conv1_wgrad_peer_kernel(const float *__restrict__ x, const float *__restrict__ dp1, const float *__restrict__ p1,
                        const uint8_t *__restrict__ i1, float *__restrict__ part, int B,
                        const __grid_constant__ PeerWin pw, float *__restrict__ m, int parity, float neg_lr, float mu,
                        float scale, int first_step, int nx, int presignaled, const LossReduce early, int n_early) {
    pdl_wait();
    if (n_early && blockIdx.x == gridDim.x - 1) {   // host-fed ring: one more grid column closes the step on the ring early
        if (blockIdx.y == 0) finish_step(early);    // (see conv1_wgrad_finish_kernel, optim.cuh)
        return;
    }
    if ((int)blockIdx.x < nx)
        peer_fc1_slice<WORLD>(pw, m, parity, neg_lr, mu, scale, first_step, blockIdx.y * nx + blockIdx.x, 2 * nx,
                              presignaled);
    else
        conv1_wgrad_body(x, dp1, p1, i1, part, B, blockIdx.x - nx, gridDim.x - nx - n_early, blockIdx.y);
}

// The end of the step (replaces grad_finalize + all_reduce + optimizer.step()): blocks [0, kFusedCovBlocks) own one
// conv1 / conv2 / fc2 element per thread; the last block reduces the per-row losses, raises this rank's "fc1 slice
// has landed" flag on every peer and closes the step's epoch.
template <int WORLD>
__global__ void __launch_bounds__(256)
peer_tail_kernel(const __grid_constant__ PeerWin pw, float *__restrict__ m, int parity, float neg_lr, float mu,
                 float scale, int first_step, const __grid_constant__ GradSegments segs, const LossReduce lr) {
    const int t = threadIdx.x, rank = pw.rank, blk = blockIdx.x;
    unsigned *flags = win_flags(pw.base[rank]);
    float *p = win_params(pw.base[rank]);
    const int ci = blk * 256 + t;          // compact index over the non-fc1 elements
    int i = ci;
    if (i >= kOffF1W) i += kOffF2W - kOffF1W;
    const bool mine = blk < kFusedCovBlocks && ci < kFusedCovElems;
    float pv = 0.f, mv = 0.f;
    if (mine) {   // written by the previous step: no need to wait for this step's kernels
        pv = p[i];
        if (!first_step) mv = m[i];
    }
    pdl_wait();   // conv1's weight gradient and this rank's fc1 slice exchange have completed
    const unsigned epoch = *reinterpret_cast<volatile unsigned *>(flags + kFlagEpochB);
    if (blk == kFusedCovBlocks) {
        // the preceding grid has completed, so its stores into the peers' windows are ordered before these flags
        if (t < WORLD && t != rank) st_release_sys(win_flags(pw.base[t]) + kFlagFc1Pushed + rank, epoch);
        if (t == 0) flags[kFlagEpochA] = epoch;   // read by the next step's exchange CTAs
        finish_step(lr);
        // the step is over when every peer's fc1 slice has landed in this rank's window as well
        if (t < WORLD && t != rank) spin_until(flags + kFlagFc1Pushed + t, epoch);
        return;
    }
    if (blk == 0 && t == 0) stamp(flags, 3);
    float gi = 0.f;
    if (mine) {
        gi = reduce_partials(segs, i, 0.f);
        const unsigned long long packed = ((unsigned long long)epoch << 32) | (unsigned long long)__float_as_uint(gi);
#pragma unroll
        for (int r = 0; r < WORLD; ++r)
            if (r != rank) st_relaxed_sys_u64(win_recv(pw.base[r], parity, rank) + ci, packed);
    }
    if (blk == 0 && t == 0) stamp(flags, 4);
    if (mine) {
        float v[WORLD];
        SpinClock clk;
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            v[r] = gi;
            if (r != rank) {
                const unsigned long long *slot = win_recv(pw.base[rank], parity, r) + ci;
                unsigned long long got = ld_relaxed_sys_u64(slot);
                while ((unsigned)(got >> 32) != epoch) {
                    clk.tick();
                    got = ld_relaxed_sys_u64(slot);
                }
                v[r] = __uint_as_float((unsigned)got);
            }
        }
        float s = v[0];
#pragma unroll
        for (int r = 1; r < WORLD; ++r) s += v[r];
        win_gsum(pw.base[rank])[i] = s;
        const float gs = __fmul_rn(s, scale);
        const float mi = first_step ? gs : __fadd_rn(__fmul_rn(mv, mu), gs);
        m[i] = mi;
        p[i] = __fmaf_rn(mi, neg_lr, pv);
    }
    if (blk == 0 && t == 0) {
        unsigned long long *st = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(flags) + kStampByteOffset);
        st[6] = st[5];
        st[5] = global_ns();
    }
}

}  // namespace b200mnist
