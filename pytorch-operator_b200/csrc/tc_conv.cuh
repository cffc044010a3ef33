// conv2 forward (20 -> 50 channels, 5x5, + bias + ReLU + 2x2 max-pool; mnist.py:28-29) as an implicit GEMM on
// the 5th-generation tensor cores:
//   D[M = conv positions of 2 images (128), N = output channels (50 -> 64)] = A[M, K = (ic,ky,kx) 500 -> 512] . W2[N, K]^T
// Same tcgen05 / TMEM / mbarrier machinery and 3xTF32 split as tc_gemm.cuh; what is new is the A operand: it is
// never materialised in global memory.  Four "builder" warps gather the im2col rows straight from the two images
// staged in shared memory (a 512-entry offset table turns k into ic*144 + ky*12 + kx), split each value into
// hi/lo and write both K-major 128B-swizzled tiles in place -- the swizzle makes the 16-byte stores of the eight
// lanes of a phase bank-conflict free.  The weight tile arrives by TMA (out-of-range rows/columns are zero filled)
// through a 6-deep ring that runs ahead of the A stages and is split by the same warps.  Rows are ordered (image, pool window, window slot), so the 2x2 max-pool and its
// arg-max are three warp shuffles per channel in the epilogue, and only the pooled 4x4 map is written.
#pragma once
#include "tc_gemm.cuh"

namespace b200mnist {
namespace tc {

constexpr int kCvStages = 3;                 // A_hi/A_lo/B_hi/B_lo stages consumed by the MMA
constexpr int kCvRing = 6;                   // raw weight tiles in flight (TMA runs this far ahead of the builders)
constexpr int kCvBuilders = 256;             // 8 builder warps: two threads per GEMM row
constexpr int kCvThreads = 64 + kCvBuilders;
constexpr int kCvImgBytes = 2 * kP1Elems * 4;                      // 23,040 B: two staged images
template <int NT> __host__ __device__ constexpr int cv_stage_bytes() { return 2 * kTileA + 2 * NT * kBK * 4; }
template <int NT> __host__ __device__ constexpr int cv_smem_bytes() {
    return kCvStages * cv_stage_bytes<NT>() + kCvRing * NT * kBK * 4 + kCvImgBytes + 2048 /*table*/ + 1024 + 256;
}

template <int NT>  // N tile: 64 (all 50 channels in one CTA) or 32 (two CTAs per image pair: small batches)
__global__ void __launch_bounds__(kCvThreads, 1)
conv2_fwd_tc_kernel(const __grid_constant__ CUtensorMap map_w, const float *__restrict__ in,
                    const float *__restrict__ bias, float *__restrict__ out, uint8_t *__restrict__ idx, int B) {
    constexpr int kTileBn = NT * kBK * 4;
    constexpr int kStage = cv_stage_bytes<NT>();
    constexpr int kNkb = 16;  // K = 500 padded to 512
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *ring = smem + kCvStages * kStage;                                            // [kCvRing][NT x 32 fp32] raw W tiles
    float *img_s = reinterpret_cast<float *>(ring + kCvRing * kTileBn);                   // [2][20][144]
    int *koff = reinterpret_cast<int *>(reinterpret_cast<uint8_t *>(img_s) + kCvImgBytes);  // [512]
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(koff) + 2048);
    uint64_t *w_full = bars, *w_empty = bars + kCvRing, *built = bars + 2 * kCvRing, *empty = built + kCvStages;
    uint64_t *acc_full = empty + kCvStages;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_full + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b0 = blockIdx.x * 2, n0 = blockIdx.y * NT;

    if (threadIdx.x == 0) {
        for (int r = 0; r < kCvRing; ++r) {
            mbar_init(smem_u32(&w_full[r]), 1);
            mbar_init(smem_u32(&w_empty[r]), kCvBuilders);
        }
        for (int s = 0; s < kCvStages; ++s) {
            mbar_init(smem_u32(&built[s]), kCvBuilders);
            mbar_init(smem_u32(&empty[s]), 1);
        }
        mbar_init(smem_u32(acc_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(NT)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp >= 2) {
        // stage the two images and the k -> offset table
        const int t = threadIdx.x - 64;
        float4 iv[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int i = t + j * kCvBuilders;
            iv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < 2 * kP1Elems / 4) {
                const int img = i / (kP1Elems / 4), b = b0 + img;
                if (b < B) iv[j] = __ldg(reinterpret_cast<const float4 *>(in + (size_t)b * kP1Elems) + (i - img * (kP1Elems / 4)));
            }
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int i = t + j * kCvBuilders;
            if (i < 2 * kP1Elems / 4) reinterpret_cast<float4 *>(img_s)[i] = iv[j];
        }
        for (int k = t; k < 512; k += kCvBuilders) {
            const int ic = k / 25, r = k - ic * 25, ky = r / 5, kx = r - ky * 5;
            koff[k] = k < 500 ? ic * 144 + ky * 12 + kx : -1;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer: raw weight tiles, kCvRing deep (independent of the A stages) =====
        if (lane == 0) {
            for (int kb = 0; kb < kNkb; ++kb) {
                const int r = kb % kCvRing, ph = (kb / kCvRing) & 1;
                mbar_wait(smem_u32(&w_empty[r]), ph ^ 1);
                const uint32_t bar = smem_u32(&w_full[r]);
                mbar_arrive_expect_tx(bar, kTileBn);
                tma_load_2d(smem_u32(ring + r * kTileBn), &map_w, kb * kBK, n0, bar);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_tf32(kBM, NT);
            for (int kb = 0; kb < kNkb; ++kb) {
                const int s = kb % kCvStages, ph = (kb / kCvStages) & 1;
                mbar_wait(smem_u32(&built[s]), ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_hi = smem_u32(smem + s * kStage), a_lo = a_hi + kTileA;
                const uint32_t b_hi = a_hi + 2 * kTileA, b_lo = b_hi + kTileBn;
#pragma unroll
                for (int kk = 0; kk < kBK / 8; ++kk) {
                    const uint32_t off = kk * 32;
                    const uint64_t dah = umma_desc_k_sw128(a_hi + off), dal = umma_desc_k_sw128(a_lo + off);
                    const uint64_t dbh = umma_desc_k_sw128(b_hi + off), dbl = umma_desc_k_sw128(b_lo + off);
                    umma_tf32(tmem_d, dal, dbh, idesc, (kb | kk) ? 1u : 0u);
                    umma_tf32(tmem_d, dah, dbl, idesc, 1u);
                    umma_tf32(tmem_d, dah, dbh, idesc, 1u);
                }
                umma_commit(smem_u32(&empty[s]));
            }
            umma_commit(smem_u32(acc_full));
        }
    } else {
        // ===== builders (warps 2-9): im2col rows -> A_hi / A_lo, split the weight tile; then epilogue =====
        const int bt = threadIdx.x - 64;
        const int t = bt & 127, half = bt >> 7;       // GEMM row, and which 4 of the row's 8 16-byte units
        const int img = t >> 6, win = (t >> 2) & 15, q = t & 3;
        const int oy = 2 * (win >> 2) + (q >> 1), ox = 2 * (win & 3) + (q & 1);
        const float *src = img_s + img * kP1Elems + oy * 12 + ox;
        const uint32_t row_off = (uint32_t)(t >> 3) * 1024 + (uint32_t)(t & 7) * 128;
        for (int kb = 0; kb < kNkb; ++kb) {
            const int s = kb % kCvStages, ph = (kb / kCvStages) & 1;
            mbar_wait(smem_u32(&empty[s]), ph ^ 1);   // the MMAs that read this stage last time have completed
            uint8_t *st = smem + s * kStage;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int j4 = half * 4 + jj;
                const int4 ko = *reinterpret_cast<const int4 *>(koff + kb * kBK + j4 * 4);
                const float v0 = ko.x >= 0 ? src[ko.x] : 0.f, v1 = ko.y >= 0 ? src[ko.y] : 0.f;
                const float v2 = ko.z >= 0 ? src[ko.z] : 0.f, v3 = ko.w >= 0 ? src[ko.w] : 0.f;
                uint4 hi, lo;
                hi.x = __float_as_uint(v0) & 0xFFFFE000u; hi.y = __float_as_uint(v1) & 0xFFFFE000u;
                hi.z = __float_as_uint(v2) & 0xFFFFE000u; hi.w = __float_as_uint(v3) & 0xFFFFE000u;
                lo.x = __float_as_uint(v0 - __uint_as_float(hi.x)); lo.y = __float_as_uint(v1 - __uint_as_float(hi.y));
                lo.z = __float_as_uint(v2 - __uint_as_float(hi.z)); lo.w = __float_as_uint(v3 - __uint_as_float(hi.w));
                const uint32_t o = row_off + (uint32_t)((j4 ^ (t & 7)) << 4);
                *reinterpret_cast<uint4 *>(st + o) = hi;
                *reinterpret_cast<uint4 *>(st + kTileA + o) = lo;
            }
            // weight tile: raw (TMA ring) -> hi / lo in this stage, same offsets (the swizzled layout is preserved)
            const int r = kb % kCvRing, rph = (kb / kCvRing) & 1;
            mbar_wait(smem_u32(&w_full[r]), rph);
            const uint4 *braw = reinterpret_cast<const uint4 *>(ring + r * kTileBn);
            uint4 *bh = reinterpret_cast<uint4 *>(st + 2 * kTileA), *bl = reinterpret_cast<uint4 *>(st + 2 * kTileA + kTileBn);
#pragma unroll
            for (int j = 0; j < kTileBn / 16 / kCvBuilders; ++j) {
                const uint4 v = braw[bt + j * kCvBuilders];
                uint4 hi = make_uint4(v.x & 0xFFFFE000u, v.y & 0xFFFFE000u, v.z & 0xFFFFE000u, v.w & 0xFFFFE000u), lo;
                lo.x = __float_as_uint(__uint_as_float(v.x) - __uint_as_float(hi.x));
                lo.y = __float_as_uint(__uint_as_float(v.y) - __uint_as_float(hi.y));
                lo.z = __float_as_uint(__uint_as_float(v.z) - __uint_as_float(hi.z));
                lo.w = __float_as_uint(__uint_as_float(v.w) - __uint_as_float(hi.w));
                bh[bt + j * kCvBuilders] = hi;
                bl[bt + j * kCvBuilders] = lo;
            }
            mbar_arrive(smem_u32(&w_empty[r]));       // ring slot may be refilled
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&built[s]));
        }
        // ---- epilogue: bias, 2x2 max-pool with arg-max (4 consecutive rows = one window), ReLU ----
        // warp w owns TMEM lane quarter w % 4; the two warps of a quarter take 32 columns each
        const int c = (warp - 2) >> 2;
        if (c * 32 < NT) {
            mbar_wait(smem_u32(acc_full), 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int qt = warp & 3;
            const int row = qt * 32 + lane;
            const int e_img = row >> 6, e_win = (row >> 2) & 15, b = b0 + e_img;
            uint32_t r[32];
            tmem_ld32(tmem_d + ((uint32_t)(qt * 32) << 16) + c * 32, r);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int oc = n0 + c * 32 + j;       // warp-uniform
                if (oc >= 50) break;
                const float v0 = __uint_as_float(r[j]) + __ldg(bias + oc);
                const float v1 = __shfl_down_sync(0xffffffffu, v0, 1);
                const float v2 = __shfl_down_sync(0xffffffffu, v0, 2);
                const float v3 = __shfl_down_sync(0xffffffffu, v0, 3);
                if ((lane & 3) == 0 && b < B) {
                    float best = v0;
                    int bi = 0;
                    if (v1 > best) { best = v1; bi = 1; }
                    if (v2 > best) { best = v2; bi = 2; }
                    if (v3 > best) { best = v3; bi = 3; }
                    const size_t o = ((size_t)b * 50 + oc) * 16 + e_win;
                    out[o] = fmaxf(best, 0.f);
                    idx[o] = (uint8_t)bi;
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(NT) : "memory");
    }
}

}  // namespace tc
}  // namespace b200mnist
