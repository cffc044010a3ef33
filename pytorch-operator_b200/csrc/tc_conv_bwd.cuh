// conv2 data gradient (autograd of mnist.py:28-29, run by loss.backward() at mnist.py:42) on the 5th-generation
// tensor cores: the same GEMM + col2im factorisation as conv2_dgrad_tile (conv_bwd.cuh),
//   T[(img,oy,ox), (ic,ky,kx)] = sum_oc dconv2[(img,oy,ox), oc] * W[oc, (ic,ky,kx)]       M = 128, N = 500, K = 50
//   dp1[img, ic, y, x]         = sum_{ky,kx} T[(img, y-ky, x-kx), (ic,ky,kx)]               (valid terms only)
// with the GEMM as tcgen05.mma kind::tf32 (3xTF32: B stacks [W_hi ; W_lo] along N, see tc_conv.cuh) and the col2im
// as the epilogue.  A CTA owns one N tile = 5 input channels x 25 taps (125 columns) and walks image pairs:
//   stagers  (4 warps): dconv2 of the pair = max-pool/ReLU backward of dp2, scattered as hi/lo into K-major
//                       no-swizzle chunk planes [oc/4][row][oc%4]; only the <= 1600 non-zeros are touched per tile
//                       (the previous tile's are zeroed again by the thread that wrote them), double-buffered;
//   MMA      (1 warp) : 7 k-steps x 2 MMAs (N = 256 and N = 128) into a double-buffered TMEM accumulator;
//   epilogue (4 warps): thread = GEMM row (img,oy,ox); for each column (ic,ky,kx) all 32 lanes of a warp add into
//                       DISTINCT cells (oy+ky, ox+kx) of the warp's private [5 ic][8 rows][12] patch -- no atomics,
//                       fixed order.  Cells of DIFFERENT taps alias across lanes, so the columns are ordered
//                       tap-major (n = tap*5 + ic): the five channels of a tap are independent read-modify-writes,
//                       and a __syncwarp() orders one tap against the next.  Then the two warps of an image are
//                       summed (rows 4..7 overlap) and the 12x12 maps stored coalesced.
#pragma once
#include "conv_bwd.cuh"
#include "tc_conv.cuh"

namespace b200mnist {
namespace tc {

constexpr int kDgThreads = 320;           // warp 0: weight loader, warp 1: MMA issuer, warps 2-5: epilogue, warps 6-9: stagers
constexpr int kDgChunks = 14;             // K = 50 output channels -> 56 = 7 k-steps of 8 = 14 chunks of 4
constexpr int kDgAPlane = 128 * 16;       // one chunk of A: 128 rows x 16 B
constexpr int kDgABytes = 2 * kDgChunks * kDgAPlane;       // hi planes then lo planes: 57,344 B
constexpr int kDgWPlane = 256 * 16;       // one chunk of B: [W_hi 128 rows ; W_lo 128 rows] x 16 B
constexpr int kDgWBytes = kDgChunks * kDgWPlane;           // 57,344 B per N tile
constexpr int kDgPatchLd = 24;            // floats per patch row (12 used): rows 0..3 of a warp hit distinct banks
constexpr int kDgPatch = 5 * 8 * kDgPatchLd;               // one warp's [5 ic][8 rows][24] patch, 960 floats
constexpr int kDgSmemBytes = 2 * kDgABytes + kDgWBytes + 2 * 4 * kDgPatch * 4 + 256 + 128;
constexpr int kDgScratchFloats = 4 * kDgWBytes / 4;        // split weights of the 4 N tiles

// conv2.weight [50][500] -> wd[N tile t][chunk c][hi rows 0..127 | lo rows 128..255][oc % 4]: row n = tap*5 + icl of
// tile t is column (5t + icl)*25 + tap of the [oc][(ic,ky,kx)] matrix (n >= 125 and oc >= 50 are zero).
// One CTA per (tile, chunk).
__global__ void __launch_bounds__(128) conv2_dgrad_weight_split_kernel(const float *__restrict__ w, float *__restrict__ wd) {
    pdl_wait();
    const int t = blockIdx.x / kDgChunks, c = blockIdx.x - t * kDgChunks, n = threadIdx.x;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int oc = c * 4 + e;
        v[e] = (n < 125 && oc < 50) ? __ldg(w + oc * 500 + (t * 5 + n % 5) * 25 + n / 5) : 0.f;
    }
    uint4 hi, lo;
    split_tf32(v, hi, lo);
    float *dst = wd + (size_t)t * (kDgWBytes / 4) + (size_t)c * (kDgWPlane / 4) + n * 4;
    *reinterpret_cast<uint4 *>(dst) = hi;
    *reinterpret_cast<uint4 *>(dst + 128 * 4) = lo;
}

struct Conv2DgradTcArgs {
    const float *dp2, *p2, *wd;
    const uint8_t *i2;
    float *dp1;
    size_t split_stride;
    int nsplit, B;
};

__global__ void __launch_bounds__(kDgThreads, 1) conv2_dgrad_tc_kernel(const Conv2DgradTcArgs g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *a_buf = smem;                                  // [2][hi|lo][chunk][128 rows][4]
    uint8_t *w_buf = a_buf + 2 * kDgABytes;                 // [chunk][hi 128 | lo 128 rows][4]
    float *patch = reinterpret_cast<float *>(w_buf + kDgWBytes);   // [2][4 warps][5][8][24]
    uint64_t *bars = reinterpret_cast<uint64_t *>(patch + 2 * 4 * kDgPatch);
    uint64_t *w_full = bars, *a_full = bars + 1, *a_empty = a_full + 2, *acc_full = a_empty + 2, *acc_empty = acc_full + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int npairs = (g.B + 1) >> 1, tile = blockIdx.y;

    if (threadIdx.x == 0) {
        mbar_init(smem_u32(w_full), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&a_full[s]), 128);
            mbar_init(smem_u32(&a_empty[s]), 1);
            mbar_init(smem_u32(&acc_full[s]), 1);
            mbar_init(smem_u32(&acc_empty[s]), 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // both A buffers and both patch sets start out zero; from then on every writer restores the zeros it disturbed
    for (int i = threadIdx.x; i < 2 * kDgABytes / 16; i += kDgThreads) reinterpret_cast<uint4 *>(a_buf)[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < 2 * 4 * kDgPatch / 4; i += kDgThreads) reinterpret_cast<uint4 *>(patch)[i] = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            const uint32_t bar = smem_u32(w_full);
            mbar_arrive_expect_tx(bar, kDgWBytes);
            bulk_g2s(smem_u32(w_buf), reinterpret_cast<const uint8_t *>(g.wd) + (size_t)tile * kDgWBytes, kDgWBytes, bar);
        }
    } else if (warp == 1) {
        // ===== MMA issuer (whole warp, one elected lane): 7 k-steps x { A_hi.[W_hi;W_lo]^T (N=256), A_lo.W_hi^T (N=128) } =====
        const uint32_t leader = elect_one();
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_d, 0);
        constexpr uint32_t idesc256 = umma_idesc_tf32(128, 256), idesc128 = umma_idesc_tf32(128, 128);
        constexpr uint32_t kHi = (128u >> 4) | (1u << 14);                    // SBO = 128 B | descriptor version
        constexpr uint32_t kAP = kDgAPlane >> 4, kWP = kDgWPlane >> 4;       // 16-byte units
        const uint32_t a0 = smem_u32(a_buf) >> 4, w0 = smem_u32(w_buf) >> 4;
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            if (it == 0) mbar_wait(smem_u32(w_full), 0);
            mbar_wait(smem_u32(&a_full[buf]), ph);
            mbar_wait(smem_u32(&acc_empty[buf]), ph ^ 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_hi = a0 + buf * (kDgABytes >> 4), a_lo = a_hi + kDgChunks * kAP;
            const uint32_t d = tm + buf * 256;
#pragma unroll
            for (int ks = 0; ks < kDgChunks / 2; ++ks) {
                const uint32_t db = (w0 + ks * 2 * kWP) | (kWP << 16);
                umma_tf32_u(d, (a_hi + ks * 2 * kAP) | (kAP << 16), kHi, db, kHi, idesc256, ks ? 1u : 0u, leader);
                umma_tf32_u(d + 128, (a_lo + ks * 2 * kAP) | (kAP << 16), kHi, db, kHi, idesc128, 1u, leader);
            }
            umma_commit_u(smem_u32(&a_empty[buf]), leader);
            umma_commit_u(smem_u32(&acc_full[buf]), leader);
        }
    } else if (warp >= 6) {
        // ===== stagers: max-pool + ReLU backward of dp2 -> scattered hi/lo operand planes =====
        const int t = threadIdx.x - 192;
        constexpr int PER = DConv2Regs<128>::PER;   // 7 elements per image per thread
        DConv2Regs<128> d0, d1;
        int prev[2][2 * PER];                        // byte offsets written two tiles ago into this buffer (-1: none)
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < 2 * PER; ++j) prev[s][j] = -1;
        int it = 0;
        {
            const int b = blockIdx.x * 2;
            if (b < g.B) d0.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)b, t);
            if (b + 1 < g.B) d1.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)b + 1, t);
        }
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&a_empty[buf]), ph ^ 1);
            uint8_t *ab = a_buf + buf * kDgABytes;
            auto scatter = [&](const DConv2Regs<128> &dr, int img, bool have, int (&old)[2 * PER]) {
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    const int slot = img * PER + j;
                    if (old[slot] >= 0) {            // restore the zero this thread disturbed last time
                        *reinterpret_cast<float *>(ab + old[slot]) = 0.f;
                        *reinterpret_cast<float *>(ab + kDgChunks * kDgAPlane + old[slot]) = 0.f;
                        old[slot] = -1;
                    }
                    const int e = t + j * 128;
                    if (have && e < kP2Elems && dr.pv[j] > 0.f) {
                        const int oc = e >> 4, p = e & 15;
                        const int oy = 2 * (p >> 2) + (dr.q[j] >> 1), ox = 2 * (p & 3) + (dr.q[j] & 1);
                        const int row = img * 64 + oy * 8 + ox;
                        const int off = (oc >> 2) * kDgAPlane + row * 16 + (oc & 3) * 4;
                        const float v = dr.g[j];
                        const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
                        *reinterpret_cast<float *>(ab + off) = hi;
                        *reinterpret_cast<float *>(ab + kDgChunks * kDgAPlane + off) = v - hi;
                        old[slot] = off;
                    }
                }
            };
            const int b = pair * 2;
            if (buf == 0) { scatter(d0, 0, b < g.B, prev[0]); scatter(d1, 1, b + 1 < g.B, prev[0]); }
            else          { scatter(d0, 0, b < g.B, prev[1]); scatter(d1, 1, b + 1 < g.B, prev[1]); }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&a_full[buf]));
            const int nb = (pair + gridDim.x) * 2;   // next tile's inputs: in flight while this tile is multiplied
            if (nb < g.B) d0.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)nb, t);
            if (nb + 1 < g.B) d1.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)nb + 1, t);
        }
    } else {
        // ===== epilogue (warps 2-5): col2im into per-warp patches, then combine + store =====
        const int q = warp & 3;                       // TMEM lane quarter: rows 32q..32q+31 = image q>>1, oy 4(q&1)..+3
        const int et = q * 32 + lane;                 // 0..127
        const int cell = (lane >> 3) * kDgPatchLd + (lane & 7);
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            float *mine = patch + (buf * 4 + q) * kDgPatch + cell;
            mbar_wait(smem_u32(&acc_full[buf]), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int s = 0; s < 5; ++s) {
                // 6 taps x 5 channels = 30 columns per slab; the last slab re-reads columns 96..127 for taps 24
                constexpr int kSlabTaps = 6;
                const int c0 = s < 4 ? s * kSlabTaps * 5 : 96, skip = s < 4 ? 0 : 24;
                uint32_t r[32], r2[32];
                tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 256 + c0, r);
                tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * 256 + 128 + c0, r2);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (s == 4) {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(smem_u32(&acc_empty[buf]));
                }
#pragma unroll
                for (int og = 0; og < kSlabTaps; ++og) {
                    const int o = s * kSlabTaps + og;  // tap (ky, kx)
                    if (o < 25) {
                        const int ky = o / 5, kx = o - ky * 5;
                        float cur[5];
#pragma unroll
                        for (int icl = 0; icl < 5; ++icl) cur[icl] = mine[icl * 8 * kDgPatchLd + ky * kDgPatchLd + kx];
#pragma unroll
                        for (int icl = 0; icl < 5; ++icl) {
                            const int jj = skip + og * 5 + icl;
                            mine[icl * 8 * kDgPatchLd + ky * kDgPatchLd + kx] =
                                cur[icl] + (__uint_as_float(r[jj]) + __uint_as_float(r2[jj]));
                        }
                        __syncwarp();   // the next tap's cells overlap this tap's in other lanes
                    }
                }
            }
            asm volatile("bar.sync 2, 128;" ::: "memory");   // the four patches of this tile are complete
            float *pb = patch + buf * 4 * kDgPatch;
#pragma unroll
            for (int k = 0; k < 12; ++k) {
                const int o = et + k * 128;           // (img, icl, pos)
                if (o < 2 * 5 * 144) {
                    const int img = o / 720, rem = o - img * 720, icl = rem / 144, pos = rem - icl * 144;
                    const int y = pos / 12, x = pos - y * 12;
                    float v = 0.f;
                    if (y < 8) {
                        float *c = pb + (2 * img) * kDgPatch + icl * 8 * kDgPatchLd + y * kDgPatchLd + x;
                        v = *c;
                        *c = 0.f;
                    }
                    if (y >= 4) {
                        float *c = pb + (2 * img + 1) * kDgPatch + icl * 8 * kDgPatchLd + (y - 4) * kDgPatchLd + x;
                        v += *c;
                        *c = 0.f;
                    }
                    const int b = pair * 2 + img;
                    if (b < g.B) g.dp1[(size_t)b * kP1Elems + (tile * 5 + icl) * 144 + pos] = v;
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512) : "memory");
    }
}


// ---------------------------------------------------------------------------------------------------------------
// conv2 weight / bias gradient on the tensor cores:
//   gW[oc, ic, ky, kx] = sum_b sum_{oy,ox} dconv2[b, oc, oy, ox] * p1[b, ic, oy+ky, ox+kx]
// as, per kernel column kx (one CTA group each) and kernel row ky (one TMEM accumulator each),
//   D_ky[M = 128: dconv2_hi rows 0..63 | dconv2_lo rows 64..127, N = 48: p1_hi 0..23 | p1_lo 24..47]
//       += A_oy[M, K = 8 positions (oy, ox = 0..7)] . B_(oy+ky)[N, K = 8 positions (oy+ky, kx + 0..7)]^T
// -- ONE MMA forms all four hi/lo products, so 3xTF32 costs a single pass.  Both operands are K-major (positions
// contiguous), no swizzle: A is dconv2 in its natural [oc][position] order (scattered from dp2 as in the data
// gradient), B is the image shifted left by kx, [y][ox/4][n][ox%4], so that the 16-byte chunks stay aligned for
// every ky -- which is why the CTAs are split by kx.  Row n = 20 of B is all ones: column 20 of D_0 is the bias
// gradient.  A CTA = (batch slice, kx) walks its images with double-buffered staging; every kWgDrain images the
// five accumulators are folded into fp32 shared-memory sums (thread = TMEM lane = private cells; ping-pong TMEM),
// which bounds the tensor core's truncating accumulation; at the end the hi and lo row halves are added and the
// slice's partial [25,056 + 64] is written for grad_finalize / the fused SGD kernel (fixed order, no atomics).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kWgThreads = 320;            // warp 0: idle, warp 1: MMA issuer, warps 2-5: drain, warps 6-9: stagers
constexpr int kWgMaxSlices = 29;           // 29 slices x 5 kx groups = 145 CTAs (one wave), <= 32 partials
constexpr int kWgDrain = 8;                // images (= 64 accumulations per accumulator) between folds
constexpr int kWgAPlane = 128 * 16;        // 4 positions x 128 rows
constexpr int kWgABytes = 16 * kWgAPlane;  // 64 positions: 32,768 B
constexpr int kWgBChunk = 48 * 16;         // 4 positions x 48 rows
constexpr int kWgBBytes = 12 * 2 * kWgBChunk;   // 12 input rows x 2 chunks: 18,432 B
constexpr int kWgAccFloats = 5 * 20 * 64 + 64;  // [ky][ic][oc] + bias[oc]
constexpr int kWgSmemBytes = 2 * (kWgABytes + kWgBBytes) + 2 * kWgAccFloats * 4 + 256 + 128;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
}

struct Conv2WgradTcArgs {
    const float *dp2, *p2, *p1;
    const uint8_t *i2;
    float *part;            // [nslices][kC2PartStride]
    size_t split_stride;
    int nsplit, B, nslices;
};

__global__ void __launch_bounds__(kWgThreads, 1) conv2_wgrad_tc_kernel(const Conv2WgradTcArgs g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *a_buf = smem;                                   // [2][16 chunks][128 rows][4 positions]
    uint8_t *b_buf = a_buf + 2 * kWgABytes;                  // [2][12 y][2 chunks][48 rows][4 positions]
    float *acc_s = reinterpret_cast<float *>(b_buf + 2 * kWgBBytes);   // [2 row halves][ky][ic][oc] (+ bias)
    uint64_t *bars = reinterpret_cast<uint64_t *>(acc_s + 2 * kWgAccFloats);
    uint64_t *a_full = bars, *a_empty = bars + 2, *acc_full = bars + 4, *acc_empty = bars + 6;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slice = blockIdx.x, kx = blockIdx.y;
    const int nimg = slice < g.B ? (g.B - slice + g.nslices - 1) / g.nslices : 0;   // images slice, slice+nslices, ...
    const int ngroups = (nimg + kWgDrain - 1) / kWgDrain;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&a_full[s]), 128);
            mbar_init(smem_u32(&a_empty[s]), 1);
            mbar_init(smem_u32(&acc_full[s]), 1);
            mbar_init(smem_u32(&acc_empty[s]), 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // A starts out zero (the stagers restore what they disturb); B's constant rows: n = 20 is all ones (bias), the
    // other padding rows 21..23 and 44..47 are zero; the fp32 sums start at zero
    for (int i = threadIdx.x; i < 2 * kWgABytes / 16; i += kWgThreads) reinterpret_cast<uint4 *>(a_buf)[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < 2 * 24 * 8; i += kWgThreads) {   // (buffer, y, chunk) x 8 constant rows
        const int blk = i >> 3, r = i & 7, n = r < 4 ? 20 + r : 40 + r;
        const float v = n == 20 ? 1.0f : 0.f;
        *reinterpret_cast<float4 *>(b_buf + blk * kWgBChunk + n * 16) = make_float4(v, v, v, v);
    }
    for (int i = threadIdx.x; i < 2 * kWgAccFloats; i += kWgThreads) acc_s[i] = 0.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;
    pdl_wait();

    if (warp == 1) {
        // ===== MMA issuer (whole warp, one elected lane): per image 5 ky x 8 oy MMAs of 128 x 48 x 8 =====
        const uint32_t leader = elect_one();
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_d, 0);
        constexpr uint32_t idesc = umma_idesc_tf32(128, 48);
        constexpr uint32_t kHi = (128u >> 4) | (1u << 14);                       // SBO = 128 B | descriptor version
        constexpr uint32_t kAP = kWgAPlane >> 4, kBC = kWgBChunk >> 4;          // 16-byte units
        const uint32_t a0 = smem_u32(a_buf) >> 4, b0 = smem_u32(b_buf) >> 4;
        for (int i = 0; i < nimg; ++i) {
            const int buf = i & 1, ph = (i >> 1) & 1;
            const int grp = i / kWgDrain, ab = grp & 1, first = (i % kWgDrain) == 0;
            if (first) mbar_wait(smem_u32(&acc_empty[ab]), ((grp >> 1) & 1) ^ 1);   // folded two groups ago
            mbar_wait(smem_u32(&a_full[buf]), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a = a0 + buf * (kWgABytes >> 4), b = b0 + buf * (kWgBBytes >> 4);
            const uint32_t d = tm + ab * 240;
#pragma unroll
            for (int oy = 0; oy < 8; ++oy) {
                const uint32_t da = (a + oy * 2 * kAP) | (kAP << 16);
#pragma unroll
                for (int ky = 0; ky < 5; ++ky) {
                    const uint32_t db = (b + (oy + ky) * 2 * kBC) | (kBC << 16);
                    umma_tf32_u(d + ky * 48, da, kHi, db, kHi, idesc, (first && oy == 0) ? 0u : 1u, leader);
                }
            }
            umma_commit_u(smem_u32(&a_empty[buf]), leader);
            if (i % kWgDrain == kWgDrain - 1 || i == nimg - 1) umma_commit_u(smem_u32(&acc_full[ab]), leader);
        }
    } else if (warp >= 6) {
        // ===== stagers: A = dconv2 (hi rows 0..63, lo rows 64..127), B = p1 shifted by kx (hi rows 0..19, lo 24..43) =====
        const int t = threadIdx.x - 192;
        constexpr int PER = DConv2Regs<128>::PER;
        DConv2Regs<128> dr;
        int prev[2][PER];
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < PER; ++j) prev[s][j] = -1;
        float pv[4][4];      // this thread's (up to 4) B items of the next image
        auto load_b = [&](int b) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int it = t + j * 128;          // (y, chunk, ic)
                if (it < 12 * 2 * 20) {
                    const int y = it / 40, r = it - y * 40, c = r / 20, ic = r - c * 20;
                    const float *src = g.p1 + (size_t)b * kP1Elems + ic * 144 + y * 12 + c * 4 + kx;
#pragma unroll
                    for (int e = 0; e < 4; ++e) pv[j][e] = __ldg(src + e);
                }
            }
        };
        if (nimg > 0) {
            dr.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)slice, t);
            load_b(slice);
        }
        for (int i = 0; i < nimg; ++i) {
            const int buf = i & 1, ph = (i >> 1) & 1;
            mbar_wait(smem_u32(&a_empty[buf]), ph ^ 1);
            uint8_t *ab = a_buf + buf * kWgABytes, *bb = b_buf + buf * kWgBBytes;
            auto scatter = [&](int (&old)[PER]) {
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    if (old[j] >= 0) {               // restore the zeros this thread disturbed two images ago
                        *reinterpret_cast<float *>(ab + old[j]) = 0.f;
                        *reinterpret_cast<float *>(ab + old[j] + 64 * 16) = 0.f;
                        old[j] = -1;
                    }
                    const int e = t + j * 128;
                    if (e < kP2Elems && dr.pv[j] > 0.f) {
                        const int oc = e >> 4, p = e & 15;
                        const int oy = 2 * (p >> 2) + (dr.q[j] >> 1), ox = 2 * (p & 3) + (dr.q[j] & 1);
                        const int off = (oy * 2 + (ox >> 2)) * kWgAPlane + oc * 16 + (ox & 3) * 4;
                        const float v = dr.g[j];
                        const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
                        *reinterpret_cast<float *>(ab + off) = hi;
                        *reinterpret_cast<float *>(ab + off + 64 * 16) = v - hi;
                        old[j] = off;
                    }
                }
            };
            if (buf == 0) scatter(prev[0]); else scatter(prev[1]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int it = t + j * 128;
                if (it < 12 * 2 * 20) {
                    const int y = it / 40, r = it - y * 40, c = r / 20, ic = r - c * 20;
                    uint4 hi, lo;
                    split_tf32(pv[j], hi, lo);
                    uint8_t *dst = bb + (y * 2 + c) * kWgBChunk + ic * 16;
                    *reinterpret_cast<uint4 *>(dst) = hi;
                    *reinterpret_cast<uint4 *>(dst + 24 * 16) = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&a_full[buf]));
            if (i + 1 < nimg) {                      // next image: in flight while this one is multiplied
                const int nb = slice + (i + 1) * g.nslices;
                dr.load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)nb, t);
                load_b(nb);
            }
        }
    } else if (warp >= 2) {
        // ===== drain (warps 2-5): TMEM -> fp32 sums in shared memory; thread = TMEM lane = row of D =====
        const int q = warp & 3, row = q * 32 + lane, half = row >> 6, oc = row & 63;
        float *mine = acc_s + half * kWgAccFloats + oc;       // [ky][ic][oc]: lanes are consecutive oc
        for (int grp = 0; grp < ngroups; ++grp) {
            const int ab = grp & 1;
            mbar_wait(smem_u32(&acc_full[ab]), (grp >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ky = 0; ky < 5; ++ky) {
                uint32_t r[32], r2[16];
                tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + ab * 240 + ky * 48, r);        // n = 0..31
                tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + ab * 240 + ky * 48 + 32, r2);  // n = 32..47
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (ky == 4) {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(smem_u32(&acc_empty[ab]));
                }
#pragma unroll
                for (int ic = 0; ic < 20; ++ic) {     // x p1_hi (n = ic) + x p1_lo (n = 24 + ic)
                    const float lo = __uint_as_float(24 + ic < 32 ? r[24 + ic] : r2[24 + ic - 32]);
                    mine[(ky * 20 + ic) * 64] += __uint_as_float(r[ic]) + lo;
                }
                if (ky == 0) mine[5 * 20 * 64] += __uint_as_float(r[20]);   // ones row: bias gradient
            }
        }
        asm volatile("bar.sync 2, 128;" ::: "memory");
        float *po = g.part + (size_t)slice * kC2PartStride;
        for (int i = row; i < 5 * 20 * 50; i += 128) {        // (ky, ic, oc), oc fastest; hi rows + lo rows
            const int o = i % 50, r = i / 50, ic = r % 20, ky = r / 20;
            const int si = (ky * 20 + ic) * 64 + o;
            po[o * 500 + ic * 25 + ky * 5 + kx] = acc_s[si] + acc_s[kWgAccFloats + si];
        }
        if (kx == 0 && row < 50) po[25024 + row] = acc_s[5 * 20 * 64 + row] + acc_s[kWgAccFloats + 5 * 20 * 64 + row];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512) : "memory");
    }
}

}  // namespace tc
}  // namespace b200mnist
