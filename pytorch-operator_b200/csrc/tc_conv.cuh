// conv2 forward (20 -> 50 channels, 5x5, + bias + ReLU + 2x2 max-pool; mnist.py:28-29) on the 5th-generation
// tensor cores as a SHIFTED-WINDOW implicit GEMM -- no im2col matrix is ever built, not even in shared memory:
//
//   D[M = 128 conv positions of an image pair, N = NT output channels]
//       = sum over the 25 kernel offsets (ky,kx) of   A_(ky,kx)[M, K = 20 input channels] . W_(ky,kx)[N, K]^T
//
// The image pair is staged ONCE per tile in shared memory, channel-chunk major:  plane[chunk = ic/4][position
// (y, img, x) of the 12x12 input][ic%4], 16 bytes per position.  That is exactly the tcgen05 K-major NO-SWIZZLE
// canonical layout ((8,m),(4,2)):((16 B, SBO),(4 B, LBO)) with the 8 rows of a core matrix = 8 consecutive x, SBO =
// one (y, img) step = 12 positions = 192 B and LBO = one chunk plane: GEMM row r = (y, img, x) of the 8x8 output
// map reads input position (y+ky, x+kx), so the A operand of kernel offset (ky,kx) is THE SAME staged planes with
// the descriptor's start address advanced by (ky*24 + kx)*16 bytes.  One hi/lo split per input value (3xTF32)
// instead of one per im2col entry (25x fewer), no gather, and the MMA operands are read from shared memory only by
// the tensor core.  K = 20 is two and a half k-steps of 8: the half steps (channels 16..19) of two consecutive taps
// share one MMA -- its two 16-byte K chunks are chunk 4 at tap o and chunk 4 at tap o+1 (LBO = the distance between
// the two windows, resp. weight blocks) -- and the odd tap out pairs with a plane of zeros, so nothing is padded:
// 63 k-steps instead of 75.
// The weights are pre-split (hi/lo) and pre-arranged per (N tile, offset) by conv2_weight_split_kernel; all 25
// offsets of a tile (NT=32: 128 KB) are loaded once per CTA with five bulk async copies and stay resident while a
// persistent CTA walks image pairs: double-buffered A staging (stager warps), MMA issue (one warp, one elected
// lane), and the bias / ReLU / max-pool / arg-max epilogue (TMEM double-buffered) overlap across tiles.
// 3xTF32 with two MMAs per k-step instead of three: in SS mode every MMA re-reads its 4 KB A operand from shared
// memory (32 cycles at 128 B/clk against 16 cycles of math at N=32), so the B operand stacks [W_hi ; W_lo] along N:
//   MMA 1: A_hi . [W_hi ; W_lo]^T -> columns [0,32) += hi.hi, columns [32,64) += hi.lo      (N = 64)
//   MMA 2: A_lo . W_hi^T          -> columns [32,64) += lo.hi                                (N = 32)
// The small terms thus have their own accumulator, which also keeps the main one to 75 truncating accumulations
// (the tensor core adds into fp32 TMEM with round-toward-zero); the epilogue adds the two halves.
#pragma once
#include "tc_gemm.cuh"

namespace b200mnist {
namespace tc {

constexpr int kWnThreads = 320;   // warp 0: weight loader, warp 1: MMA issuer, warps 2-5: epilogue, warps 6-9: stagers
constexpr int kWnStagers = 128, kWnEpilogue = 128;
constexpr int kWnChunks = 5;                              // 20 input channels = 5 chunks of 4
constexpr int kWnPlane = 2 * 144 * 16;                    // one chunk of an image pair: 288 positions x 16 B = 4,608 B
constexpr int kWnABytes = 2 * kWnChunks * kWnPlane;       // hi planes then lo planes: 46,080 B
constexpr int kWnWGroups = 5;                             // weights arrive in 5 bulk copies of 5 offsets each
template <int NT> __host__ __device__ constexpr int wn_wchunk() { return 2 * NT * 16; }   // [W_hi rows ; W_lo rows] of one chunk
template <int NT> __host__ __device__ constexpr int wn_wblock() { return kWnChunks * wn_wchunk<NT>(); }  // one kernel offset
template <int NT> __host__ __device__ constexpr int wn_wbytes() { return 25 * wn_wblock<NT>(); }
template <int NT> __host__ __device__ constexpr int wn_smem_bytes() {
    return 2 * kWnABytes + kWnPlane /*zeros*/ + wn_wbytes<NT>() + wn_wchunk<NT>() /*zeros*/ + 256 /*barriers*/ + 128 /*align*/;
}
constexpr int kWnScratchFloats = 2 * wn_wbytes<32>() / 4;   // split weights of both N tiles

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void split_tf32(const float (&v)[4], uint4 &hi, uint4 &lo) {
    hi.x = __float_as_uint(v[0]) & 0xFFFFE000u; hi.y = __float_as_uint(v[1]) & 0xFFFFE000u;
    hi.z = __float_as_uint(v[2]) & 0xFFFFE000u; hi.w = __float_as_uint(v[3]) & 0xFFFFE000u;
    lo.x = __float_as_uint(v[0] - __uint_as_float(hi.x)); lo.y = __float_as_uint(v[1] - __uint_as_float(hi.y));
    lo.z = __float_as_uint(v[2] - __uint_as_float(hi.z)); lo.w = __float_as_uint(v[3] - __uint_as_float(hi.w));
}

// conv2.weight [50][20][5][5] -> w2s[N tile][offset ky*5+kx][chunk][hi|lo][oc % NT][ic % 4], the shared-memory image
// of the stacked B operands (rows >= 50 are zero).  One CTA per kernel offset.
template <int NT>
__device__ __forceinline__ void conv2_weight_split_block(const float *__restrict__ w, float *__restrict__ w2s, int o) {
    for (int i = threadIdx.x; i < (64 / NT) * kWnChunks * NT; i += blockDim.x) {
        const int tile = i / (kWnChunks * NT), rem = i - tile * (kWnChunks * NT), chunk = rem / NT, ol = rem - chunk * NT;
        const int oc = tile * NT + ol;
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = oc < 50 ? __ldg(w + (oc * 20 + chunk * 4 + e) * 25 + o) : 0.f;
        uint4 hi, lo;
        split_tf32(v, hi, lo);
        float *dst = w2s + (size_t)tile * (wn_wbytes<NT>() / 4) + (size_t)o * (wn_wblock<NT>() / 4) + (chunk * 2 * NT + ol) * 4;
        *reinterpret_cast<uint4 *>(dst) = hi;
        *reinterpret_cast<uint4 *>(dst + NT * 4) = lo;
    }
}
template <int NT>
__global__ void __launch_bounds__(256) conv2_weight_split_kernel(const float *__restrict__ w, float *__restrict__ w2s) {
    pdl_wait();
    conv2_weight_split_block<NT>(w, w2s, blockIdx.x);
}

template <int NT>
__global__ void __launch_bounds__(kWnThreads, 1)
conv2_fwd_win_kernel(const float *__restrict__ in, const float *__restrict__ w2s, const float *__restrict__ bias,
                     float *__restrict__ out, uint8_t *__restrict__ idx, int B, float *__restrict__ out_hi,
                     float *__restrict__ out_lo) {
    // out_hi / out_lo (optional): hi / lo copies of the pooled output with leading dimension kP2SplitLd, the A operand
    // of fc1's tensor-core GEMMs -- written here so that no separate split pass has to re-read p2
    static_assert(NT == 32, "epilogue reads two 32-column TMEM slabs per tile");
    constexpr int kWBlock = wn_wblock<NT>();
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *a_buf = smem;                                  // [2][hi|lo][chunk][288 positions][4]
    uint8_t *a_zero = a_buf + 2 * kWnABytes;                // one plane of zeros (chunk "5" of every k-step 2)
    uint8_t *w_buf = a_zero + kWnPlane;                     // [25 offsets][chunk][hi|lo][NT][4]
    uint8_t *w_zero = w_buf + wn_wbytes<NT>();              // [2 NT][4] zeros
    uint64_t *bars = reinterpret_cast<uint64_t *>(w_zero + wn_wchunk<NT>());
    uint64_t *w_full = bars, *a_full = bars + kWnWGroups, *a_empty = a_full + 2, *acc_full = a_empty + 2,
             *acc_empty = acc_full + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int npairs = (B + 1) >> 1, n0 = blockIdx.y * NT;

    if (threadIdx.x == 0) {
        for (int g = 0; g < kWnWGroups; ++g) mbar_init(smem_u32(&w_full[g]), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&a_full[s]), kWnStagers);
            mbar_init(smem_u32(&a_empty[s]), 1);
            mbar_init(smem_u32(&acc_full[s]), 1);
            mbar_init(smem_u32(&acc_empty[s]), kWnEpilogue);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(4 * NT)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = threadIdx.x; i < kWnPlane / 16; i += kWnThreads) reinterpret_cast<uint4 *>(a_zero)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x < 2 * NT) reinterpret_cast<uint4 *>(w_zero)[threadIdx.x] = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;
    pdl_wait();   // everything above overlaps the tail of the producer of `in` / `w2s`

    if (warp == 0) {
        // ===== weight loader: the CTA's 25 x (hi, lo) blocks, once =====
        if (lane == 0) {
            constexpr uint32_t kGroupBytes = 5 * kWBlock;
            const uint8_t *src = reinterpret_cast<const uint8_t *>(w2s) + (size_t)blockIdx.y * wn_wbytes<NT>();
            for (int g = 0; g < kWnWGroups; ++g) {
                const uint32_t bar = smem_u32(&w_full[g]);
                mbar_arrive_expect_tx(bar, kGroupBytes);
                bulk_g2s(smem_u32(w_buf + g * kGroupBytes), src + (size_t)g * kGroupBytes, kGroupBytes, bar);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: 63 k-steps (25 taps x 2 + 13 shared half steps) x 2 MMAs.  The WHOLE warp runs the loop so that every operand
        // is warp-uniform (descriptors live in uniform registers, ~3 instructions per MMA); one elected lane issues.
        const uint32_t leader = elect_one();
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_d, 0);
        constexpr uint32_t idesc64 = umma_idesc_tf32(128, 2 * NT), idesc32 = umma_idesc_tf32(128, NT);
        constexpr uint32_t kP = kWnPlane >> 4, kWC = wn_wchunk<NT>() >> 4, kWB = kWBlock >> 4;   // 16-byte units
        constexpr uint32_t kAHi = (192u >> 4) | (1u << 14), kBHi = (128u >> 4) | (1u << 14);      // SBO | version
        const uint32_t a0 = smem_u32(a_buf) >> 4, az = smem_u32(a_zero) >> 4;
        const uint32_t w0 = smem_u32(w_buf) >> 4, wz = smem_u32(w_zero) >> 4;
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&a_full[buf]), ph);
            mbar_wait(smem_u32(&acc_empty[buf]), ph ^ 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_hi = a0 + buf * (kWnABytes >> 4), a_lo = a_hi + kWnChunks * kP;
            const uint32_t d = tm + buf * 2 * NT;
#pragma unroll
            for (int o = 0; o < 25; ++o) {
                if (it == 0 && o % 5 == 0) mbar_wait(smem_u32(&w_full[o / 5]), 0);
                const uint32_t shift = (uint32_t)((o / 5) * 24 + (o % 5));
                const uint32_t b = w0 + o * kWB;
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {      // channels 0..7 and 8..15 of this tap
                    const uint32_t ah = a_hi + ks * 2 * kP, al = a_lo + ks * 2 * kP, bc = b + ks * 2 * kWC;
                    const uint32_t db = bc | (kWC << 16);
                    umma_tf32_u(d, (ah + shift) | (kP << 16), kAHi, db, kBHi, idesc64, (o | ks) ? 1u : 0u, leader);   // hi.hi | hi.lo
                    umma_tf32_u(d + NT, (al + shift) | (kP << 16), kAHi, db, kBHi, idesc32, 1u, leader);              // lo.hi
                }
                // channels 16..19 are half a k-step: pair this tap's chunk 4 with the previous tap's (the second
                // 16-byte K chunk may lie anywhere: LBO = distance between the two taps' windows / weight blocks);
                // the odd tap out (24) pairs with the plane / block of zeros
                if ((o & 1) || o == 24) {
                    const int o0 = o == 24 ? 24 : o - 1;
                    const uint32_t sh0 = (uint32_t)((o0 / 5) * 24 + (o0 % 5));
                    const uint32_t ah = a_hi + 4 * kP + sh0, al = a_lo + 4 * kP + sh0, bc = w0 + o0 * kWB + 4 * kWC;
                    const uint32_t lbo_ah = o == 24 ? az - (a_hi + 4 * kP) : shift - sh0;
                    const uint32_t lbo_al = o == 24 ? az - (a_lo + 4 * kP) : shift - sh0;
                    const uint32_t db = bc | ((o == 24 ? wz - bc : kWB) << 16);
                    umma_tf32_u(d, ah | (lbo_ah << 16), kAHi, db, kBHi, idesc64, 1u, leader);
                    umma_tf32_u(d + NT, al | (lbo_al << 16), kAHi, db, kBHi, idesc32, 1u, leader);
                }
            }
            umma_commit_u(smem_u32(&a_empty[buf]), leader);
            umma_commit_u(smem_u32(&acc_full[buf]), leader);
        }
    } else if (warp >= 6) {
        // ===== stagers: image pair -> hi / lo chunk planes (one split per input value) =====
        const int t = threadIdx.x - 192;
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&a_empty[buf]), ph ^ 1);   // the MMAs that read this buffer two tiles ago are done
            uint8_t *ab = a_buf + buf * kWnABytes;
#pragma unroll
            for (int j0 = 0; j0 < 12; j0 += 4) {
                float v[4][4];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int i = t + (j0 + jj) * kWnStagers;   // (chunk, img, pos)
                    const int chunk = i / 288, rem = i - chunk * 288, img = rem >= 144 ? 1 : 0, pos = rem - img * 144;
                    const int b = pair * 2 + img;
                    const bool ok = i < kWnChunks * 288 && b < B;
                    const float *src = in + (size_t)(ok ? b : 0) * kP1Elems + (ok ? chunk * 4 * 144 + pos : 0);
#pragma unroll
                    for (int e = 0; e < 4; ++e) v[jj][e] = ok ? __ldg(src + e * 144) : 0.f;
                }
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int i = t + (j0 + jj) * kWnStagers;
                    if (i < kWnChunks * 288) {
                        const int chunk = i / 288, rem = i - chunk * 288, img = rem >= 144 ? 1 : 0, pos = rem - img * 144;
                        const int y = pos / 12, x = pos - y * 12;
                        uint4 hi, lo;
                        split_tf32(v[jj], hi, lo);
                        uint8_t *dst = ab + chunk * kWnPlane + ((y * 2 + img) * 12 + x) * 16;
                        *reinterpret_cast<uint4 *>(dst) = hi;
                        *reinterpret_cast<uint4 *>(dst + kWnChunks * kWnPlane) = lo;
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&a_full[buf]));
        }
    } else {
        // ===== epilogue (warps 2-5): bias, 2x2 max-pool + arg-max (partners are lanes ^1, ^16, ^17), ReLU =====
        const int q = warp & 3;                      // TMEM lane quarter this warp may read
        const int img = (lane >> 3) & 1, x = lane & 7;
        const int win = q * 4 + (x >> 1);            // row = q*32 + lane -> y = 2q + (lane >> 4): pooled row q
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&acc_full[buf]), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            uint32_t r[32], r2[32];
            tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 2 * NT, r);        // hi.hi
            tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 2 * NT + NT, r2);  // hi.lo + lo.hi
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(smem_u32(&acc_empty[buf]));   // the accumulator may be overwritten by tile it + 2
            const int b = pair * 2 + img;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int oc = n0 + j;               // warp-uniform
                if (oc >= 50) break;
                const float v0 = (__uint_as_float(r[j]) + __uint_as_float(r2[j])) + __ldg(bias + oc);
                const float v1 = __shfl_xor_sync(0xffffffffu, v0, 1);
                const float v2 = __shfl_xor_sync(0xffffffffu, v0, 16);
                const float v3 = __shfl_xor_sync(0xffffffffu, v0, 17);
                if ((lane & 17) == 0 && b < B) {
                    float best = v0;
                    int bi = 0;
                    if (v1 > best) { best = v1; bi = 1; }
                    if (v2 > best) { best = v2; bi = 2; }
                    if (v3 > best) { best = v3; bi = 3; }
                    const size_t o = ((size_t)b * 50 + oc) * 16 + win;
                    const float pv = fmaxf(best, 0.f);
                    out[o] = pv;
                    idx[o] = (uint8_t)bi;
                    if (out_hi) {
                        const size_t f = (size_t)b * kP2SplitLd + oc * 16 + win;
                        const float hi = tf32_hi(pv);
                        out_hi[f] = hi;
                        out_lo[f] = pv - hi;
                        if (oc == 0 && win == 0) {   // bias-gradient column (columns 801.. only feed discarded outputs)
                            out_hi[(size_t)b * kP2SplitLd + 800] = 1.0f;
                            out_lo[(size_t)b * kP2SplitLd + 800] = 0.0f;
                        }
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(4 * NT) : "memory");
    }
}

}  // namespace tc
}  // namespace b200mnist
