// 5th-generation tensor-core GEMMs for the fc1 layer at large batch (BASELINE config 5: B = 4096 / rank):
//   forward : h   [B,500]   = relu(p2 [B,800] . W1[500,800]^T + b1)      A K-major,  B K-major
//   dgrad   : dp2 [B,800]   = dz1[B,500] . W1[500,800]                   A K-major,  B MN-major
//   wgrad   : gW1 [500,800] = dz1[B,500]^T . p2[B,800], gb1 = colsum(dz1) A MN-major, B MN-major
// written directly against the sm_100a programming model: TMA (cp.async.bulk.tensor) stages 128-byte
// swizzled K-major tiles, one elected thread issues tcgen05.mma kind::tf32 with the accumulator in
// TMEM, tcgen05.commit signals mbarriers, tcgen05.ld brings the tile back for the epilogue.
//
// Precision: a single TF32 pass (10-bit mantissa) misses the 1e-5 parity bound by two orders of
// magnitude, so the kernel does the error-compensated 3xTF32 product: every fp32 operand x is used as
// hi = x with the low 13 mantissa bits cleared (exactly a TF32 number, whatever rounding the MMA applies
// to its inputs) and lo = x - hi (exact in fp32);  D += A_lo.B_hi + A_hi.B_lo + A_hi.B_hi per k-slice.
// The operands arrive PRE-SPLIT: a small elementwise kernel (split_rows_kernel, one pass per operand per
// step) writes hi and lo copies to the workspace and TMA loads the four tiles of a stage directly -- the
// first version split every landed tile inside the k-loop with 8 extra warps, and ncu put their
// fence.proxy.async + barrier traffic, not the MMAs, on the critical path (2-3x the MMA time).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2-5 = TMEM drain + epilogue (TMEM lane quarter = warp_id % 4).
// Long reductions (wgrad: K = batch): the tensor core's fp32 accumulation error grows with the number
// of chained MMAs, so every kDrain k-blocks the TMEM tile is folded into fp32 registers (round-to-nearest
// adds) by the epilogue warps while the MMA continues into the second TMEM accumulator.
// Every mbarrier wait is bounded: a barrier that never completes traps instead of hanging the GPU.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace b200mnist {
namespace tc {

// CTA tile; kBK fp32 = one 128-byte swizzle row.  With hi and lo copies of both operands a 128 x 64 tile pulled
// 48 KB from L2 per k-block for 0.39 us of MMA time -- ~8.8 TB/s over all SMs, i.e. L2-bandwidth bound; 128 x 128
// needs 64 KB for twice the work.
constexpr int kBM = 128, kBN = 128, kBK = 32;
constexpr int kStages = 3;
constexpr int kThreads = 192;           // warp 0: TMA, warp 1: MMA issuer, warps 2-5: TMEM drain + epilogue
constexpr int kTileA = kBM * kBK * 4, kTileB = kBN * kBK * 4;            // 16,384 B each
constexpr int kStageBytes = 2 * kTileA + 2 * kTileB;                      // hi + lo of both = 65,536 B
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
static_assert(2 * kStages * 8 + 4 * 8 + 8 <= 256, "barrier area");
constexpr int kTmemCols = 512;        // two (128 main + 128 small-term) accumulators (ping-pong): all of TMEM
constexpr int kDrain = 8;            // k-blocks (K = 256) accumulated in TMEM before the tile is folded into fp32 registers

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: poll, and trap after 10 s of wall-clock time (a lost barrier must not hang the box).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    unsigned long long t0 = 0;
    for (uint32_t spin = 0;; ++spin) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(bar), "r"(parity)
                     : "memory");
        if (ok) return;
        if ((spin & 1023u) == 1023u) {
            const unsigned long long now = global_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > kSpinTimeoutNs) __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
                 : "memory");
}
// Shared-memory matrix descriptors are passed to the MMA as their two 32-bit words:
//   lo = (start address >> 4) | (LBO >> 4) << 16        hi = (SBO >> 4) | 1 << 14 (version) | layout << 29
// K-major, 128-byte swizzle (layout 2): rows of 128 B = 32 fp32 along K, LBO unused (1), SBO = 8 rows x 128 B.
// MN-major operand of 32-bit elements: the only layout the TF32 MMA accepts is the 128-byte swizzle with a
// 32-BYTE atom (layout 1, TMA CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): rows of 128 B = 32 contiguous M/N elements,
// the four 32-byte units of a row XOR-permuted by (k-row % 4); an atom is 4 k-rows (SBO = 512 B); 32-element
// chunks along M/N are LBO apart.
// K-major without swizzle (layout 0, tc_conv*.cuh): core matrix = 8 rows x 16 B (contiguous 128 B); LBO = distance
// between the two 16-byte K chunks of a k-step, SBO = distance between consecutive groups of 8 rows.
// Instruction descriptor: D = F32, A = B = TF32, M x N tile; a_mn / b_mn = 1 for an MN-major operand.
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N, int a_mn = 0, int b_mn = 0) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// Warp-uniform issue: the whole warp executes the call with identical operands (so they can live in uniform
// registers) and `leader` (from elect_one()) is non-zero in exactly one lane, which issues the instruction.
// Descriptors are passed as their two 32-bit words.
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void umma_tf32_u(uint32_t d_tmem, uint32_t a_lo32, uint32_t a_hi32, uint32_t b_lo32,
                                            uint32_t b_hi32, uint32_t idesc, uint32_t acc, uint32_t leader) {
    const uint64_t ad = ((uint64_t)a_hi32 << 32) | a_lo32, bd = ((uint64_t)b_hi32 << 32) | b_lo32;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
                 "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc), "r"(leader)
                 : "memory");
}
__device__ __forceinline__ void umma_commit_u(uint32_t bar, uint32_t leader) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t"
                 "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
                 ::"r"(bar), "r"(leader)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
                   "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
                   "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr)
                 : "memory");
}

enum Epilogue { kEpiBiasRelu = 0, kEpiStore = 1, kEpiWgrad = 2 };

struct TcGemmArgs {
    const float *bias;   // kEpiBiasRelu: added per n
    float *C;            // [M][ldc]
    float *colsum;       // kEpiWgrad: colsum[m] = sum_k A(m,k), delivered through a column of ones injected at n == N
    int M, N, K, ldc;
};

// The four tensor maps of one GEMM: hi and lo copies of both operands (same shape and layout).
struct TcMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

// A_MN / B_MN: operand is MN-major in global memory (its M / N index is the contiguous one).
// kEpiWgrad: column N of the B operand is a column of ones in the hi copy (zeros in the lo copy), written by the
// split pass, so that column N of the product is colsum(A) -- fc1's bias gradient rides in the GEMM.
// kb0 / kb1: the range of kBK-wide k-blocks this CTA reduces (split-K).  ATOMIC: the tile is ADDED to C / colsum with
// red.global.add.f32 instead of stored -- used with exactly TWO k-ranges per output tile onto a zeroed destination, so
// the result is p0 + p1 whichever CTA arrives first (fp32 addition commutes): still bitwise reproducible.
template <bool A_MN, bool B_MN, int EPI, bool ATOMIC = false>
__device__ __forceinline__ void gemm_tc_tile(const TcMaps &maps, const TcGemmArgs &g, const int tile_x, const int tile_y,
                                             const int kb0, const int kb1) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    // stage s: [A_hi | A_lo | B_hi | B_lo]; barriers after the last stage
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes);
    uint64_t *full = bars, *empty = bars + kStages;
    uint64_t *acc_full = bars + 2 * kStages, *acc_empty = bars + 2 * kStages + 2;   // [2] each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kStages + 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = tile_y * kBM, n0 = tile_x * kBN;
    const int nkb = kb1 - kb0;
    const int ngroups = (nkb + kDrain - 1) / kDrain;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(smem_u32(&full[s]), 1);
            mbar_init(smem_u32(&empty[s]), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(&acc_full[b]), 1);
            mbar_init(smem_u32(&acc_empty[b]), 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer: four tiles per stage (hi and lo of both operands), one transaction barrier =====
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kStages, ph = (kb / kStages) & 1;
                mbar_wait(smem_u32(&empty[s]), ph ^ 1);
                uint8_t *st = smem + s * kStageBytes;
                const uint32_t bar = smem_u32(&full[s]);
                mbar_arrive_expect_tx(bar, kStageBytes);
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const CUtensorMap *ma = half ? &maps.a_lo : &maps.a_hi, *mb = half ? &maps.b_lo : &maps.b_hi;
                    uint8_t *da = st + half * kTileA, *db = st + 2 * kTileA + half * kTileB;
                    if (A_MN) {
#pragma unroll
                        for (int c = 0; c < kBM / 32; ++c)  // [32 k-rows x 32 m] boxes, 4 KB each
                            tma_load_2d(smem_u32(da + c * 4096), ma, m0 + 32 * c, (kb0 + kb) * kBK, bar);
                    } else {
                        tma_load_2d(smem_u32(da), ma, (kb0 + kb) * kBK, m0, bar);
                    }
                    if (B_MN) {
#pragma unroll
                        for (int c = 0; c < kBN / 32; ++c)
                            tma_load_2d(smem_u32(db + c * 4096), mb, n0 + 32 * c, (kb0 + kb) * kBK, bar);
                    } else {
                        tma_load_2d(smem_u32(db), mb, (kb0 + kb) * kBK, n0, bar);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer.  The WHOLE warp runs the loop so that every operand is warp-uniform (descriptors stay in
        // uniform registers, ~3 instructions per MMA instead of a 14-instruction broadcast sequence); one elected
        // lane issues.  Two MMAs per k-slice: B_lo lies right behind B_hi in the stage, so one N = 2*kBN MMA forms
        //   A_hi . [B_hi ; B_lo]^T  -> columns [0,kBN) += hi.hi, columns [kBN,2kBN) += hi.lo
        // with a single read of A_hi, and  A_lo . B_hi^T  adds lo.hi to the second half.  The small terms have their
        // own accumulator (fewer truncating accumulations into the main one); the drain adds the halves. =====
        const uint32_t leader = elect_one();
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_d, 0);
        constexpr uint32_t idesc2 = umma_idesc_tf32(kBM, 2 * kBN, A_MN ? 1 : 0, B_MN ? 1 : 0);
        constexpr uint32_t idesc1 = umma_idesc_tf32(kBM, kBN, A_MN ? 1 : 0, B_MN ? 1 : 0);
        // descriptor words (see the layout notes above): hi = SBO | version | layout, lo = addr | LBO
        constexpr uint32_t kHiK = (1024u >> 4) | (1u << 14) | (2u << 29), kHiMN = (512u >> 4) | (1u << 14) | (1u << 29);
        constexpr uint32_t kLboK = 1u << 16, kLboMN = (4096u >> 4) << 16;
        constexpr uint32_t a_hi32 = A_MN ? kHiMN : kHiK, b_hi32 = B_MN ? kHiMN : kHiK;
        constexpr uint32_t a_lbo = A_MN ? kLboMN : kLboK, b_lbo = B_MN ? kLboMN : kLboK;
        // 8 fp32 along K: 32 B inside the swizzle row (K-major) / one 1024-B atom of 8 k-rows (MN-major), in 16-B units
        constexpr uint32_t a_step = A_MN ? 64 : 2, b_step = B_MN ? 64 : 2;
        const uint32_t base16 = smem_u32(smem) >> 4;
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % kStages, ph = (kb / kStages) & 1;
            const int grp = kb / kDrain, buf = grp & 1, kin = kb % kDrain;
            if (kin == 0 && grp >= 2) {   // this accumulator was last used by group grp-2: wait until it has been folded
                mbar_wait(smem_u32(&acc_empty[buf]), ((grp >> 1) - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const uint32_t tmem_acc = tm + buf * 2 * kBN;
            mbar_wait(smem_u32(&full[s]), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_hi = base16 + s * (kStageBytes >> 4), a_lo = a_hi + (kTileA >> 4), b_hi = a_hi + (2 * kTileA >> 4);
#pragma unroll
            for (int kk = 0; kk < kBK / 8; ++kk) {
                const uint32_t db = (b_hi + kk * b_step) | b_lbo;
                umma_tf32_u(tmem_acc, (a_hi + kk * a_step) | a_lbo, a_hi32, db, b_hi32, idesc2, (kin | kk) ? 1u : 0u, leader);
                umma_tf32_u(tmem_acc + kBN, (a_lo + kk * a_step) | a_lbo, a_hi32, db, b_hi32, idesc1, 1u, leader);
            }
            umma_commit_u(smem_u32(&empty[s]), leader);          // stage free once these MMAs have read it
            if (kin == kDrain - 1 || kb == nkb - 1) umma_commit_u(smem_u32(&acc_full[buf]), leader);  // group complete
        }
    } else {
        // ===== warps 2-5 (128 threads): fold each completed accumulator group into fp32 registers, then the epilogue =====
        const int q = warp & 3;            // a warp may only touch its own TMEM lane quarter
        float acc[kBN];
#pragma unroll
        for (int j = 0; j < kBN; ++j) acc[j] = 0.f;
        for (int grp = 0; grp < ngroups; ++grp) {
            const int buf = grp & 1;
            mbar_wait(smem_u32(&acc_full[buf]), (grp >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int c = 0; c < kBN / 32; ++c) {
                uint32_t r[32], r2[32];
                tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 2 * kBN + c * 32, r);         // hi.hi
                tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 2 * kBN + kBN + c * 32, r2);  // hi.lo + lo.hi
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[c * 32 + j] += __uint_as_float(r[j]) + __uint_as_float(r2[j]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(smem_u32(&acc_empty[buf]));
        }
        // epilogue: registers -> global, 16 bytes at a time (a thread owns a row of the tile: its columns are contiguous;
        // N, ldc and n0 are multiples of 4.  Scalar stores made this loop 26 % of the kernel's stall samples.)
        const int row = m0 + q * 32 + lane;
        if (row < g.M) {
#pragma unroll
            for (int j = 0; j < kBN; j += 4) {
                const int n = n0 + j;
                if (n < g.N) {
                    float4 v = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
                    if (EPI == kEpiBiasRelu) {
                        const float4 b = __ldg(reinterpret_cast<const float4 *>(g.bias + n));
                        v = make_float4(fmaxf(v.x + b.x, 0.f), fmaxf(v.y + b.y, 0.f), fmaxf(v.z + b.z, 0.f), fmaxf(v.w + b.w, 0.f));
                    }
                    float4 *dst = reinterpret_cast<float4 *>(g.C + (size_t)row * g.ldc + n);
                    if (ATOMIC) atomicAdd(dst, v);
                    else *dst = v;
                } else if (EPI == kEpiWgrad && n == g.N) {
                    if (ATOMIC) atomicAdd(&g.colsum[row], acc[j]);
                    else g.colsum[row] = acc[j];
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(kTmemCols) : "memory");
    }
}

template <bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ TcMaps maps, const TcGemmArgs g) {
    pdl_wait();
    gemm_tc_tile<A_MN, B_MN, EPI>(maps, g, blockIdx.x, blockIdx.y, 0, (g.K + kBK - 1) / kBK);
}

// fc1's two backward GEMMs in ONE launch (they only share dz1): the first n_wg CTAs are weight-gradient tiles
// (gW1 = dz1^T . p2, K = B: 52 output tiles whose k-loop is the batch -- each is split into TWO k-halves that add
// their tiles onto the zeroed gradient, 104 CTAs of half the length), the rest data-gradient tiles (dp2 = dz1 . W1,
// K = 500: many short ones) that fill the remaining SMs while the long tiles run.
__global__ void __launch_bounds__(kThreads, 1)
fc1_bwd_tc_kernel(const __grid_constant__ TcMaps wg_maps, const TcGemmArgs wg, const __grid_constant__ TcMaps dg_maps,
                  const TcGemmArgs dg, const int wg_gx, const int n_wg, const int dg_gx) {
    pdl_wait();
    int b = blockIdx.x;
    if (b < n_wg) {
        const int half = b & 1, tile = b >> 1, nkb = (wg.K + kBK - 1) / kBK, mid = (nkb + 1) / 2;
        gemm_tc_tile<true, true, kEpiWgrad, true>(wg_maps, wg, tile % wg_gx, tile / wg_gx, half ? mid : 0, half ? nkb : mid);
    } else {
        b -= n_wg;
        gemm_tc_tile<false, true, kEpiStore>(dg_maps, dg, b % dg_gx, b / dg_gx, 0, (dg.K + kBK - 1) / kBK);
    }
}

// hi / lo copies of a row-major fp32 matrix [rows][cols] (leading dimension ld_in) into two matrices with leading
// dimension ld_out >= cols (a multiple of 4): hi = x with the low 13 mantissa bits cleared, lo = x - hi.  Columns
// [cols, ld_out) are padding: zero, except hi[row][ones_col] = 1 when ones_col >= 0 (the wgrad GEMM's bias column).
// One float4 per thread; cols % 4 == 0.
__global__ void __launch_bounds__(256)
split_rows_kernel(const float *__restrict__ in, int ld_in, float *__restrict__ hi, float *__restrict__ lo, int ld_out,
                  int rows, int cols, int ones_col) {
    pdl_wait();
    const int v4 = ld_out >> 2;
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long)rows * v4) return;
    const int r = (int)(idx / v4), c = (int)(idx - (long long)r * v4) * 4;
    float4 h = make_float4(0.f, 0.f, 0.f, 0.f), l = h;
    if (c < cols) {
        const float4 x = __ldg(reinterpret_cast<const float4 *>(in + (size_t)r * ld_in + c));
        h.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
        h.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
        h.z = __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u);
        h.w = __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u);
        l = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
    } else if (ones_col >= c && ones_col < c + 4) {
        (&h.x)[ones_col - c] = 1.0f;
    }
    *reinterpret_cast<float4 *>(hi + (size_t)r * ld_out + c) = h;
    *reinterpret_cast<float4 *>(lo + (size_t)r * ld_out + c) = l;
}

// ---- host side: tensor maps through the driver entry point (no link-time dependency on libcuda) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// Row-major fp32 matrix [rows][cols] (cols contiguous), box = box_rows x 32 floats, 128-byte swizzle.
inline bool make_map_kmajor(CUtensorMap *m, const float *base, int rows, int cols, int ld, int box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Same matrix viewed as an MN-major operand: box = 32 rows (k) x 32 contiguous elements (m or n), 4 KB.
inline bool make_map_mnmajor(CUtensorMap *m, const float *base, int rows, int cols, int ld) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)kBK};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace tc
}  // namespace b200mnist
