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
//   epilogue (4 warps): col2im, one input channel (25 columns) at a time.  Every thread parks its row of T in
//                       shared memory ([tap][row], stride 132: conflict-free both ways), one named barrier, then
//                       thread (img, y, kx) -- 24 output rows x 5 kernel columns = 120 of the 128 threads -- sums
//                       over the valid kernel rows ky the eight T values (oy = y-ky, ox = 0..7) of its column kx
//                       with 128-bit loads: every T value is touched exactly once (a per-pixel gather would issue
//                       2.25x as many, mostly predicated-off, loads).  The five kx partial rows of an output row
//                       meet in a 160-byte warp-private strip (x = ox + kx) and are stored.  Fixed summation order,
//                       no atomics.  The park buffer is double-buffered so one barrier per channel suffices.
#pragma once
#include "conv_bwd.cuh"
#include "tc_conv.cuh"

namespace b200mnist {
namespace tc {

constexpr int kDgThreads = 448;           // warp 0: weight loader, 1: MMA issuer, 2-5 and 10-13: two epilogue groups, 6-9: stagers
constexpr int kDgChunks = 14;             // K = 50 output channels -> 56 = 7 k-steps of 8 = 14 chunks of 4
constexpr int kDgAPlane = 128 * 16;       // one chunk of A: 128 rows x 16 B
constexpr int kDgABytes = 2 * kDgChunks * kDgAPlane;       // hi planes then lo planes: 57,344 B
constexpr int kDgWPlane = 256 * 16;       // one chunk of B: [W_hi 128 rows ; W_lo 128 rows] x 16 B
constexpr int kDgWBytes = kDgChunks * kDgWPlane;           // 57,344 B per N tile
constexpr int kDgParkLd = 132;            // floats per tap: 128 rows + 4 (spreads the kx lanes over the banks)
constexpr int kDgPark = 25 * kDgParkLd;   // one input channel of T: [tap][row]
constexpr int kDgStrip = 2 * 4 * 6 * 5 * 8;   // [group][warp][output row of the warp][kx][ox] partial rows
constexpr int kDgSmemBytes = 2 * kDgABytes + kDgWBytes + (2 * kDgPark + kDgStrip) * 4 + 256 + 128;
constexpr int kDgScratchFloats = 4 * kDgWBytes / 4;        // split weights of the 4 N tiles

// Four consecutive pooled pixels (one pooled row of one channel) of the incoming gradient with what max-pool and
// ReLU backward need: 3 vector loads per group instead of 12+ scalar ones.  Group gi of an image = elements 4gi..4gi+3
// of its [50][4][4] map; 200 groups per image, two per stager thread.  dp2 may be `nsplit` partial sums (summed in
// ascending order, like sum_partials).
struct DcGroup {
    float4 g, pv;
    uint32_t q;     // 4 arg-max slots, one per byte
    __device__ __forceinline__ void load(const float *__restrict__ dp2, int nsplit, size_t split_stride,
                                         const float *__restrict__ p2, const uint8_t *__restrict__ i2, size_t b, int gi) {
        const size_t o = b * kP2Elems + (size_t)gi * 4;
        pv = __ldcg(reinterpret_cast<const float4 *>(p2 + o));
        q = __ldg(reinterpret_cast<const uint32_t *>(i2 + o));
        g = __ldcg(reinterpret_cast<const float4 *>(dp2 + o));
        for (int s = 1; s < nsplit; ++s) {
            const float4 c = __ldcg(reinterpret_cast<const float4 *>(dp2 + (size_t)s * split_stride + o));
            g.x += c.x; g.y += c.y; g.z += c.z; g.w += c.w;
        }
    }
    // element px of the group: gradient (0 where ReLU was inactive) and its position in the 8x8 conv2 map
    __device__ __forceinline__ bool elem(int px, int py, float &v, int &oy, int &ox) const {
        const float gv[4] = {g.x, g.y, g.z, g.w}, pp[4] = {pv.x, pv.y, pv.z, pv.w};
        const int qq = (q >> (8 * px)) & 0xFF;
        v = gv[px];
        oy = 2 * py + (qq >> 1);
        ox = 2 * px + (qq & 1);
        return pp[px] > 0.f;
    }
};
constexpr int kDcGroups = kP2Elems / 4;   // 200 per image

// conv2.weight [50][500] -> wd[N tile t][chunk c][hi rows 0..127 | lo rows 128..255][oc % 4]: row n of tile t is
// column t*125 + n of the [oc][(ic,ky,kx)] matrix (n >= 125 and oc >= 50 are zero).  One CTA per (tile, chunk).
__device__ __forceinline__ void conv2_dgrad_weight_split_block(const float *__restrict__ w, float *__restrict__ wd, int blk) {
    const int t = blk / kDgChunks, c = blk - t * kDgChunks, n = threadIdx.x;
    if (n >= 128) return;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int oc = c * 4 + e;
        v[e] = (n < 125 && oc < 50) ? __ldg(w + oc * 500 + t * 125 + n) : 0.f;
    }
    uint4 hi, lo;
    split_tf32(v, hi, lo);
    float *dst = wd + (size_t)t * (kDgWBytes / 4) + (size_t)c * (kDgWPlane / 4) + n * 4;
    *reinterpret_cast<uint4 *>(dst) = hi;
    *reinterpret_cast<uint4 *>(dst + 128 * 4) = lo;
}
__global__ void __launch_bounds__(128) conv2_dgrad_weight_split_kernel(const float *__restrict__ w, float *__restrict__ wd) {
    pdl_wait();
    conv2_dgrad_weight_split_block(w, wd, blockIdx.x);
}
// Both operand copies of conv2.weight in one launch (training step: the weights do not change between the forward
// and the backward pass): CTAs [0, 25) build the forward taps, the other 4 x kDgChunks the data-gradient tiles.
// Blocks beyond those (present when w1 != nullptr) write the hi / lo copies of fc1.weight for fc1's GEMMs, one float4 per thread.
__global__ void __launch_bounds__(256) conv2_weight_split_both_kernel(const float *__restrict__ w, float *__restrict__ w2s,
                                                                      float *__restrict__ wd, const float *__restrict__ w1,
                                                                      float *__restrict__ w1_hi, float *__restrict__ w1_lo) {
    pdl_wait();
    if (blockIdx.x < 25) conv2_weight_split_block<32>(w, w2s, blockIdx.x);
    else if (blockIdx.x < 25 + 4 * kDgChunks) conv2_dgrad_weight_split_block(w, wd, blockIdx.x - 25);
    else {
        const int i = ((blockIdx.x - 25 - 4 * kDgChunks) * 256 + threadIdx.x) * 4;
        if (i < 500 * 800) {
            const float4 x = __ldg(reinterpret_cast<const float4 *>(w1 + i));
            const float4 h = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
            *reinterpret_cast<float4 *>(w1_hi + i) = h;
            *reinterpret_cast<float4 *>(w1_lo + i) = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
        }
    }
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
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
    float *park = reinterpret_cast<float *>(w_buf + kDgWBytes);    // [2 groups][25 taps][132]
    float *strip = park + 2 * kDgPark;                             // [2 groups][4 warps][6][5][8]
    uint64_t *bars = reinterpret_cast<uint64_t *>(strip + kDgStrip);
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
            mbar_init(smem_u32(&acc_empty[s]), 256);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // both A buffers start out zero; from then on the stagers restore the zeros they disturb
    for (int i = threadIdx.x; i < 2 * kDgABytes / 16; i += kDgThreads) reinterpret_cast<uint4 *>(a_buf)[i] = make_uint4(0, 0, 0, 0);
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
    } else if (warp >= 6 && warp < 10) {
        // ===== stagers: max-pool + ReLU backward of dp2 -> scattered hi/lo operand planes =====
        const int t = threadIdx.x - 192;
        DcGroup d[2][2];                             // [image of the pair][group t, group t + 128]
        int prev[2][16];                             // byte offsets written two tiles ago into this buffer (-1: none)
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < 16; ++j) prev[s][j] = -1;
        auto load_pair = [&](int b) {
#pragma unroll
            for (int img = 0; img < 2; ++img)
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    if (b + img < g.B && t + j * 128 < kDcGroups)
                        d[img][j].load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)(b + img), t + j * 128);
        };
        load_pair(blockIdx.x * 2);
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&a_empty[buf]), ph ^ 1);
            uint8_t *ab = a_buf + buf * kDgABytes;
            auto scatter = [&](int (&old)[16]) {
#pragma unroll
                for (int img = 0; img < 2; ++img)
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int gi = t + j * 128, oc = gi >> 2, py = gi & 3;
                        const bool have = pair * 2 + img < g.B && gi < kDcGroups;
#pragma unroll
                        for (int px = 0; px < 4; ++px) {
                            const int slot = (img * 2 + j) * 4 + px;
                            if (old[slot] >= 0) {    // restore the zero this thread disturbed last time
                                *reinterpret_cast<float *>(ab + old[slot]) = 0.f;
                                *reinterpret_cast<float *>(ab + kDgChunks * kDgAPlane + old[slot]) = 0.f;
                                old[slot] = -1;
                            }
                            float v;
                            int oy, ox;
                            if (have && d[img][j].elem(px, py, v, oy, ox)) {
                                const int row = img * 64 + oy * 8 + ox;
                                const int off = (oc >> 2) * kDgAPlane + row * 16 + (oc & 3) * 4;
                                const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
                                *reinterpret_cast<float *>(ab + off) = hi;
                                *reinterpret_cast<float *>(ab + kDgChunks * kDgAPlane + off) = v - hi;
                                old[slot] = off;
                            }
                        }
                    }
            };
            if (buf == 0) scatter(prev[0]); else scatter(prev[1]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&a_full[buf]));
            load_pair((pair + gridDim.x) * 2);       // next tile's inputs: in flight while this tile is multiplied
        }
    } else {
        // ===== epilogue: two groups of four warps (2-5: channels 0, 2, 4; 10-13: channels 1, 3), each with its own
        // park buffer, strips and named barrier, so that two channels are in flight (the per-channel chain TMEM load ->
        // park -> barrier -> 128-bit loads -> strip -> store is latency bound) =====
        const int eg = warp >= 10 ? 1 : 0;
        const int q = warp & 3;                       // TMEM lane quarter: rows 32q..32q+31
        const int row = q * 32 + lane;                // GEMM row (img, oy, ox) = 64 img + 8 oy + ox
        const int gl = lane / 5, kx = lane - gl * 5;  // reduce role: output row G = 6q + gl (lanes 30, 31 idle), column kx
        const int G = q * 6 + gl, img = G >= 12 ? 1 : 0, y = G - img * 12;
        const bool reducer = lane < 30;
        float *pb = park + eg * kDgPark;
        float *my_strip = strip + ((eg * 4 + q) * 6 + (reducer ? gl : 0)) * 40;
        int it = 0;
        for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
            const int buf = it & 1, ph = (it >> 1) & 1;
            mbar_wait(smem_u32(&acc_full[buf]), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int b = pair * 2 + img;
#pragma unroll 1
            for (int icl = eg; icl < 5; icl += 2) {
                // the channel's 25 columns as two overlapping 16-column loads (0..15 and 9..24) per accumulator half
                uint32_t ra[16], rb[16], sa[16], sb[16];
                const uint32_t tc0 = tmem_d + ((uint32_t)(q * 32) << 16) + buf * 256 + icl * 25;
                tmem_ld16(tc0, ra);
                tmem_ld16(tc0 + 9, rb);
                tmem_ld16(tc0 + 128, sa);
                tmem_ld16(tc0 + 128 + 9, sb);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (icl + 2 >= 5) {                   // this group's last read of the accumulator
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(smem_u32(&acc_empty[buf]));
                }
#pragma unroll
                for (int tap = 0; tap < 25; ++tap) {
                    const float v = tap < 16 ? __uint_as_float(ra[tap]) + __uint_as_float(sa[tap])
                                             : __uint_as_float(rb[tap - 9]) + __uint_as_float(sb[tap - 9]);
                    pb[tap * kDgParkLd + row] = v;
                }
                asm volatile("bar.sync %0, 128;" ::"r"(2 + eg) : "memory");   // this channel's T is parked
                if (reducer) {
                    float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;   // partial row: x = kx + 0..7
#pragma unroll
                    for (int ky = 0; ky < 5; ++ky) {
                        const int oy = y - ky;
                        if ((unsigned)oy < 8u) {
                            const float4 *src = reinterpret_cast<const float4 *>(pb + (ky * 5 + kx) * kDgParkLd + img * 64 + oy * 8);
                            const float4 u = src[0], w4 = src[1];
                            lo.x += u.x; lo.y += u.y; lo.z += u.z; lo.w += u.w;
                            hi.x += w4.x; hi.y += w4.y; hi.z += w4.z; hi.w += w4.w;
                        }
                    }
                    reinterpret_cast<float4 *>(my_strip + kx * 8)[0] = lo;
                    reinterpret_cast<float4 *>(my_strip + kx * 8)[1] = hi;
                }
                asm volatile("bar.sync %0, 128;" ::"r"(2 + eg) : "memory");   // park buffer free again; strips written
                if (reducer && b < g.B) {
                    float *dst = g.dp1 + (size_t)b * kP1Elems + (tile * 5 + icl) * 144 + y * 12;
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const int x = kx + 5 * k;
                        if (x < 12) {
                            float v = 0.f;
#pragma unroll
                            for (int k2 = 0; k2 < 5; ++k2)
                                if ((unsigned)(x - k2) < 8u) v += my_strip[k2 * 8 + x - k2];
                            dst[x] = v;
                        }
                    }
                }
                __syncwarp();
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
// as, per kernel column kx (one CTA group each), ONE MMA per output row oy for all five kernel rows ky:
//   D[M = 128: dconv2_hi rows 0..63 | dconv2_lo rows 64..127,  N = 240: ky x (p1_hi 0..23 | p1_lo 24..47)]
//       += A_oy[M, K = 8 positions (oy, ox = 0..7)] . B_oy[N, K = 8 positions (oy+ky, kx + 0..7)]^T
// -- the MMA forms all four hi/lo products (3xTF32 in a single pass) of five kernel rows at once, 128 x 240 x 8:
// math-bound (120 cycles against 92 of shared-memory operand reads).  Both operands are K-major (positions
// contiguous), no swizzle: A is dconv2 in its natural [oc][position] order (scattered from dp2 as in the data
// gradient); B is the image shifted left by kx, [ox/4][y][n][ox%4]: the 16-byte chunks stay aligned for every ky
// (which is why the CTAs are split by kx) and the 48 rows of input row y+1 follow those of row y, so the five
// windows oy+0..4 are simply 240 consecutive operand rows.  Row n = 20 of B is all ones: column 20 of D is the
// bias gradient.  A CTA = (batch slice, kx) walks its images with double-buffered staging; every kWgDrain images the
// five accumulators are folded into fp32 shared-memory sums (thread = TMEM lane = private cells; ping-pong TMEM),
// which bounds the tensor core's truncating accumulation; at the end the hi and lo row halves are added and the
// slice's partial [25,056 + 64] is written for grad_finalize / the fused SGD kernel (fixed order, no atomics).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kWgThreads = 320;            // warp 0: idle, warp 1: MMA issuer, warps 2-5: drain, warps 6-9: stagers
constexpr int kWgMaxSlices = 29;           // 29 slices x 5 kx groups = 145 CTAs (one wave), <= 32 partials
constexpr int kWgDrain = 8;                // images (= 64 accumulations per accumulator) between folds
constexpr int kWgAPlane = 128 * 16;        // 4 positions x 128 rows
constexpr int kWgABytes = 16 * kWgAPlane;  // 64 positions: 32,768 B
constexpr int kWgBChunk = 48 * 16;         // 4 positions x 48 rows (one input row y)
constexpr int kWgBBytes = 2 * 12 * kWgBChunk;   // 2 chunks x 12 input rows: 18,432 B
constexpr int kWgAccLd = 65;                    // floats per (ky, ic) row of the fp32 sums: 64 oc + 1 (conflict-free both ways)
constexpr int kWgAccFloats = 5 * 20 * kWgAccLd + 64 + 4;  // [ky][ic][oc] + bias[oc] (8-byte multiple for the barriers behind it)
constexpr int kWgSmemBytes = 2 * (kWgABytes + kWgBBytes) + 2 * kWgAccFloats * 4 + 256 + 128;

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
    uint8_t *b_buf = a_buf + 2 * kWgABytes;                  // [2][2 chunks][12 y][48 rows][4 positions]
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
    for (int i = threadIdx.x; i < 2 * 24 * 8; i += kWgThreads) {   // (buffer, chunk, y) x 8 constant rows
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
        // ===== MMA issuer (whole warp, one elected lane): per image 8 MMAs (one per oy) of 128 x 240 x 8 =====
        const uint32_t leader = elect_one();
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_d, 0);
        constexpr uint32_t idesc = umma_idesc_tf32(128, 240);
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
            for (int oy = 0; oy < 8; ++oy)   // B rows: input rows oy..oy+4 x 48; its second chunk is 12 input rows on
                umma_tf32_u(d, (a + oy * 2 * kAP) | (kAP << 16), kHi, (b + oy * kBC) | ((12 * kBC) << 16), kHi, idesc,
                            (first && oy == 0) ? 0u : 1u, leader);
            umma_commit_u(smem_u32(&a_empty[buf]), leader);
            if (i % kWgDrain == kWgDrain - 1 || i == nimg - 1) umma_commit_u(smem_u32(&acc_full[ab]), leader);
        }
    } else if (warp >= 6) {
        // ===== stagers: A = dconv2 (hi rows 0..63, lo rows 64..127), B = p1 shifted by kx (hi rows 0..19, lo 24..43) =====
        const int t = threadIdx.x - 192;
        DcGroup d[2];                                // groups t and t + 128 of the image
        int prev[2][8];
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < 8; ++j) prev[s][j] = -1;
        auto load_a = [&](int b) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
                if (t + j * 128 < kDcGroups) d[j].load(g.dp2, g.nsplit, g.split_stride, g.p2, g.i2, (size_t)b, t + j * 128);
        };
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
            load_a(slice);
            load_b(slice);
        }
        for (int i = 0; i < nimg; ++i) {
            const int buf = i & 1, ph = (i >> 1) & 1;
            mbar_wait(smem_u32(&a_empty[buf]), ph ^ 1);
            uint8_t *ab = a_buf + buf * kWgABytes, *bb = b_buf + buf * kWgBBytes;
            auto scatter = [&](int (&old)[8]) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int gi = t + j * 128, oc = gi >> 2, py = gi & 3;
#pragma unroll
                    for (int px = 0; px < 4; ++px) {
                        const int slot = j * 4 + px;
                        if (old[slot] >= 0) {        // restore the zeros this thread disturbed two images ago
                            *reinterpret_cast<float *>(ab + old[slot]) = 0.f;
                            *reinterpret_cast<float *>(ab + old[slot] + 64 * 16) = 0.f;
                            old[slot] = -1;
                        }
                        float v;
                        int oy, ox;
                        if (gi < kDcGroups && d[j].elem(px, py, v, oy, ox)) {
                            const int off = (oy * 2 + (ox >> 2)) * kWgAPlane + oc * 16 + (ox & 3) * 4;
                            const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
                            *reinterpret_cast<float *>(ab + off) = hi;
                            *reinterpret_cast<float *>(ab + off + 64 * 16) = v - hi;
                            old[slot] = off;
                        }
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
                    uint8_t *dst = bb + (c * 12 + y) * kWgBChunk + ic * 16;
                    *reinterpret_cast<uint4 *>(dst) = hi;
                    *reinterpret_cast<uint4 *>(dst + 24 * 16) = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(smem_u32(&a_full[buf]));
            if (i + 1 < nimg) {                      // next image: in flight while this one is multiplied
                const int nb = slice + (i + 1) * g.nslices;
                load_a(nb);
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
                    mine[(ky * 20 + ic) * kWgAccLd] += __uint_as_float(r[ic]) + lo;
                }
                if (ky == 0) mine[5 * 20 * kWgAccLd] += __uint_as_float(r[20]);   // ones row: bias gradient
            }
        }
        asm volatile("bar.sync 2, 128;" ::: "memory");
        float *po = g.part + (size_t)slice * kC2PartStride;
        for (int i = row; i < 5 * 20 * 50; i += 128) {        // (oc, ic, ky), ky fastest: consecutive threads write 20 bytes apart
            const int o = i / 100, r = i - o * 100, ic = r / 5, ky = r - ic * 5;
            const int si = (ky * 20 + ic) * kWgAccLd + o;
            po[o * 500 + ic * 25 + ky * 5 + kx] = acc_s[si] + acc_s[kWgAccFloats + si];   // hi rows + lo rows
        }
        if (kx == 0 && row < 50) po[25024 + row] = acc_s[5 * 20 * kWgAccLd + row] + acc_s[kWgAccFloats + 5 * 20 * kWgAccLd + row];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512) : "memory");
    }
}

}  // namespace tc
}  // namespace b200mnist
