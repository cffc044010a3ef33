// Kernels for the two linear layers, the log-softmax/NLL head and their backward
// (mnist.py:31-33,41-42).
#pragma once
#include "common.cuh"

namespace b200mnist {

// ---------------------------------------------------------------------------------
// fp32 SIMT GEMM  C[M,N] = sum_k A(m,k) * B(k,n), shared-memory tiled, register blocked,
// optional split-K across blockIdx.z (partials are reduced by the consumer kernel, so
// a split costs no extra launch).  fp32 FMA keeps the 1e-5 parity bound that a single
// TF32 pass (10-bit mantissa) cannot.  The next k-tile is fetched into registers while
// the current one is multiplied (one exposed global latency per CTA, not one per tile).
//   A_KCONTIG : A(m,k) = A[m*lda + k]  (else A[k*lda + m])
//   B_KCONTIG : B(k,n) = B[n*ldb + k]  (else B[k*ldb + n])
// Requirements: leading dimensions and the contiguous extents are multiples of 4,
// pointers 16-byte aligned (true for 800, 500 and the 256-byte aligned buckets).
// ---------------------------------------------------------------------------------
constexpr int kGemmBK = 16;

struct GemmArgs {
    const float *A;
    const float *B;
    float *C;             // [z][M][ldc] when gridDim.z > 1
    const float *bias;    // optional, added per n (only when gridDim.z == 1)
    float *rowsum;        // optional: rowsum[m] = sum_k A(m,k) (written by blockIdx.x == 0, gridDim.z == 1)
    int M, N, K, lda, ldb, ldc, kchunk, relu;
};

template <int ROWS, int NT, bool KCONTIG>
struct TileLoader {
    // One operand tile of ROWS (m or n) x kGemmBK (k) floats, moved as float4.
    static constexpr int kVec = ROWS * kGemmBK / 4;
    static constexpr int kPerThread = (kVec + NT - 1) / NT;
    float4 r[kPerThread];

    __device__ __forceinline__ void load(const float *base, int ld, int row0, int nrows, int k0, int kend, int tid) {
#pragma unroll
        for (int j = 0; j < kPerThread; ++j) {
            const int i = tid + j * NT;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < kVec) {
                if (KCONTIG) {
                    const int row = i / (kGemmBK / 4), k4 = (i % (kGemmBK / 4)) * 4;
                    if (row0 + row < nrows && k0 + k4 < kend)
                        v = __ldg(reinterpret_cast<const float4 *>(base + (size_t)(row0 + row) * ld + k0 + k4));
                } else {
                    const int k = i / (ROWS / 4), r4 = (i % (ROWS / 4)) * 4;
                    if (k0 + k < kend && row0 + r4 < nrows)
                        v = __ldg(reinterpret_cast<const float4 *>(base + (size_t)(k0 + k) * ld + row0 + r4));
                }
            }
            r[j] = v;
        }
    }
    __device__ __forceinline__ void store(float *S, int lds, int tid) const {  // S[k][row], row stride lds
#pragma unroll
        for (int j = 0; j < kPerThread; ++j) {
            const int i = tid + j * NT;
            if (i < kVec) {
                if (KCONTIG) {
                    const int row = i / (kGemmBK / 4), k4 = (i % (kGemmBK / 4)) * 4;
                    S[(k4 + 0) * lds + row] = r[j].x;
                    S[(k4 + 1) * lds + row] = r[j].y;
                    S[(k4 + 2) * lds + row] = r[j].z;
                    S[(k4 + 3) * lds + row] = r[j].w;
                } else {
                    const int k = i / (ROWS / 4), r4 = (i % (ROWS / 4)) * 4;
                    *reinterpret_cast<float4 *>(&S[k * lds + r4]) = r[j];
                }
            }
        }
    }
};

template <int BM, int BN, int TM, int TN, bool A_KCONTIG, bool B_KCONTIG>
__device__ __forceinline__ void gemm_tile(const GemmArgs &g, int bx, int by, int bz, float *As, float *Bs) {
    constexpr int NT = (BM / TM) * (BN / TN);
    constexpr int LDA_S = BM + 4, LDB_S = BN + 4;
    const int tid = threadIdx.x, tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int m0 = by * BM, n0 = bx * BN;
    const int kbeg = bz * g.kchunk, kend = min(g.K, kbeg + g.kchunk);
    static_assert(TN % 2 == 0, "accumulators are packed in pairs along n (FFMA2)");
    float2 acc[TM][TN / 2];
    float rs[TM];
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        rs[i] = 0.f;
#pragma unroll
        for (int j = 0; j < TN / 2; ++j) acc[i][j] = make_float2(0.f, 0.f);
    }
    const bool want_rs = (g.rowsum != nullptr) && bx == 0 && tx == 0;
    TileLoader<BM, NT, A_KCONTIG> la;
    TileLoader<BN, NT, B_KCONTIG> lb;
    if (tid < NT) {
        la.load(g.A, g.lda, m0, g.M, kbeg, kend, tid);
        lb.load(g.B, g.ldb, n0, g.N, kbeg, kend, tid);
        la.store(As, LDA_S, tid);
        lb.store(Bs, LDB_S, tid);
    }
    __syncthreads();
    for (int k0 = kbeg; k0 < kend; k0 += kGemmBK) {
        const bool has_next = k0 + kGemmBK < kend;
        if (has_next && tid < NT) {  // in flight while this tile is multiplied
            la.load(g.A, g.lda, m0, g.M, k0 + kGemmBK, kend, tid);
            lb.load(g.B, g.ldb, n0, g.N, k0 + kGemmBK, kend, tid);
        }
        if (tid < NT) {
#pragma unroll
            for (int kk = 0; kk < kGemmBK; ++kk) {
                float a[TM];
                float2 b[TN / 2];
#pragma unroll
                for (int i = 0; i < TM; ++i) a[i] = As[kk * LDA_S + ty * TM + i];
#pragma unroll
                for (int j = 0; j < TN / 2; ++j) b[j] = *reinterpret_cast<const float2 *>(&Bs[kk * LDB_S + tx * TN + 2 * j]);
#pragma unroll
                for (int i = 0; i < TM; ++i) {
                    if (want_rs) rs[i] += a[i];
                    const float2 aa = make_float2(a[i], a[i]);
#pragma unroll
                    for (int j = 0; j < TN / 2; ++j) acc[i][j] = __ffma2_rn(aa, b[j], acc[i][j]);  // 2 FMAs / lane
                }
            }
        }
        __syncthreads();
        if (has_next) {
            if (tid < NT) {
                la.store(As, LDA_S, tid);
                lb.store(Bs, LDB_S, tid);
            }
            __syncthreads();
        }
    }
    if (tid >= NT) return;
    float *C = g.C + (size_t)bz * g.M * g.ldc;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + ty * TM + i;
        if (m >= g.M) continue;
        if (want_rs) g.rowsum[m] = rs[i];
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n >= g.N) continue;
            float v = (j & 1) ? acc[i][j >> 1].y : acc[i][j >> 1].x;
            if (g.bias) v += __ldg(g.bias + n);
            if (g.relu) v = fmaxf(v, 0.f);
            C[(size_t)m * g.ldc + n] = v;
        }
    }
}

template <int BM, int BN, int TM, int TN, bool A_KCONTIG, bool B_KCONTIG>
__global__ void __launch_bounds__((BM / TM) * (BN / TN)) gemm_kernel(const GemmArgs g) {

    pdl_wait();
    __shared__ __align__(16) float As[kGemmBK * (BM + 4)];
    __shared__ __align__(16) float Bs[kGemmBK * (BN + 4)];
    gemm_tile<BM, BN, TM, TN, A_KCONTIG, B_KCONTIG>(g, blockIdx.x, blockIdx.y, blockIdx.z, As, Bs);
}

// ---------------------------------------------------------------------------------
// fc2 weight/bias gradient: gW2[j,k] = sum_b dz2[b,j] h[b,k], gb2[j] = sum_b dz2[b,j].
// Thread = one k (coalesced over h rows), 10 accumulators; the batch is split over CTAs
// (vy), each split writes its own partial [5056 + 64] (summed by the fused SGD /
// grad_finalize).  Rows are fetched 16 at a time with all loads issued up front.
// ---------------------------------------------------------------------------------
constexpr int kFc2wThreads = 128;
constexpr int kFc2PartStride = 5120;  // 5000 weights @0, 10 biases @5056

__device__ __forceinline__ void fc2_wgrad_tile(const float *__restrict__ dz2, const float *__restrict__ h,
                                               float *__restrict__ part, int B, int rows_per_split, int vx, int vy,
                                               float *ds /* 16*10 floats of shared memory */) {
    const int t = threadIdx.x, k = vx * kFc2wThreads + t;
    const int bbeg = vy * rows_per_split, bend = min(B, bbeg + rows_per_split);
    float acc[kClasses], accb = 0.f;
#pragma unroll
    for (int j = 0; j < kClasses; ++j) acc[j] = 0.f;
    const bool mine = t < kFc2wThreads && k < kHElems;
    for (int c0 = bbeg; c0 < bend; c0 += 16) {
        const int nb = min(16, bend - c0);
        float hv[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) hv[r] = (mine && r < nb) ? h[(size_t)(c0 + r) * kHElems + k] : 0.f;
        __syncthreads();
        for (int i = t; i < nb * kClasses; i += blockDim.x) ds[i] = dz2[(size_t)c0 * kClasses + i];
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r < nb) {
#pragma unroll
                for (int j = 0; j < kClasses; ++j) acc[j] = fmaf(ds[r * kClasses + j], hv[r], acc[j]);
                if (vx == 0 && t < kClasses) accb += ds[r * kClasses + t];
            }
        }
    }
    float *p = part + (size_t)vy * kFc2PartStride;
    if (mine) {
#pragma unroll
        for (int j = 0; j < kClasses; ++j) p[j * kHElems + k] = acc[j];
    }
    if (vx == 0 && t < kClasses) p[5056 + t] = accb;
}

__global__ void __launch_bounds__(kFc2wThreads)
fc2_wgrad_kernel(const float *__restrict__ dz2, const float *__restrict__ h, float *__restrict__ part, int B,
                 int rows_per_split) {
    __shared__ float ds[16 * kClasses];
    fc2_wgrad_tile(dz2, h, part, B, rows_per_split, blockIdx.x, blockIdx.y, ds);
}

// ---------------------------------------------------------------------------------
// One launch for the whole backward of the two linear layers (three independent pieces
// that would otherwise be three dependent launches): CTAs [0, n_wg) compute fc1's weight
// gradient tiles (64x64, + bias gradient as row sums), the next n_dg CTAs fc1's data
// gradient tiles (64x32, split-K), the last n_f2 CTAs fc2's weight/bias gradient.
// ---------------------------------------------------------------------------------
struct FcBwdArgs {
    GemmArgs wg, dg;
    int wg_gx, wg_n;          // fc1 wgrad: grid.x, CTA count
    int dg_gx, dg_gy, dg_n;   // fc1 dgrad: grid.x, grid.y, CTA count
    const float *dz2, *h;
    float *fc2_part;
    int B, fc2_rows, fc2_gx;
};

__global__ void __launch_bounds__(256) fc_bwd_kernel(const FcBwdArgs a) {

    pdl_wait();
    __shared__ __align__(16) float As[kGemmBK * (64 + 4)];
    __shared__ __align__(16) float Bs[kGemmBK * (64 + 4)];
    int b = blockIdx.x;
    if (b < a.wg_n) {
        gemm_tile<64, 64, 4, 4, false, false>(a.wg, b % a.wg_gx, b / a.wg_gx, 0, As, Bs);
        return;
    }
    b -= a.wg_n;
    if (b < a.dg_n) {
        const int per_z = a.dg_gx * a.dg_gy;
        const int z = b / per_z, r = b - z * per_z;
        gemm_tile<64, 32, 2, 4, true, false>(a.dg, r % a.dg_gx, r / a.dg_gx, z, As, Bs);
        return;
    }
    b -= a.dg_n;
    fc2_wgrad_tile(a.dz2, a.h, a.fc2_part, a.B, a.fc2_rows, b % a.fc2_gx, b / a.fc2_gx, As);
}

// ---------------------------------------------------------------------------------
// Head: (optional) finish fc1 from its split-K partials + bias + ReLU, then fc2,
// log_softmax, NLL and the first backward steps, one CTA per batch row:
//   h      = relu(sum_s part[s][b][:] + b1)           (nsplit > 0; also written to h_out)
//   z      = h W2^T + b2 ; logp = z - logsumexp(z)
//   rows[b]= -logp[target_b] ;  dz2 = (exp(logp) - onehot) * inv_B ;  dz1 = (dz2 W2) * (h > 0)
// W2 (20 KB) is staged in shared memory with all loads in flight at once; the 10 dot products
// are thread-partial sums + warp shuffles.  The last CTA to finish (threadfence + ticket)
// reduces rows[] in a fixed order, so the scalar loss is deterministic without a second launch.
//   stats_out[0] = sum(rows) * loss_scale ; stats_out[1] = #(argmax == target) if want_correct
// ---------------------------------------------------------------------------------
constexpr int kHeadThreads = 256;
constexpr int kMaxFc1Split = 10;

struct HeadArgs {
    const float *h_in;      // [B,500] when nsplit == 0
    const float *part;      // [nsplit][B][512] fc1 partial sums when nsplit > 0
    const float *b1;
    float *h_out;           // [B,500], written when nsplit > 0
    const float *w2, *b2;
    const int64_t *target;
    float *logp;            // optional [B,10]
    float *dz2;             // optional [B,10]
    float *dz1;             // optional [B,500]
    float *rows;            // [2*B] scratch: per-row nll, per-row correct
    unsigned *ticket;       // zero-initialised once; self-resetting
    float *stats_out;
    int B, nsplit, want_correct;
    int defer_reduce;       // 1: only write rows[]; a later kernel (reduce_rows_block) produces stats_out
    float inv_B, loss_scale;
};

// What the head does on top when the step's fc1 backward runs on the tensor cores (tc_gemm.cuh): hi / lo copies of
// dz1 (the GEMMs' operand) and clearing of the split-K weight-gradient destination (zero_n float4 at zero_buf).
struct HeadTcExtra {
    float *dz1_hi, *dz1_lo;
    float4 *zero_buf;
    int zero_n;
};

template <bool TC_BWD>
__device__ __forceinline__ void head_body(const HeadArgs &a, const HeadTcExtra &x) {

    pdl_wait();
    __shared__ __align__(16) float w2s[kClasses * kHElems];  // 20,000 B
    __shared__ float wred[kHeadThreads / 32][kClasses];
    __shared__ float zs[16];
    __shared__ float dzs[16];
    __shared__ float red[kHeadThreads];
    __shared__ float red2[kHeadThreads];
    __shared__ bool is_last;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    if (TC_BWD && x.zero_buf)
        for (int i = blockIdx.x * kHeadThreads + t; i < x.zero_n; i += gridDim.x * kHeadThreads)
            x.zero_buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    // ---- all global loads of the first row up front: W2 tile and this thread's two h elements ----
    float4 wv[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int i = t + j * kHeadThreads;
        wv[j] = i < kClasses * kHElems / 4 ? __ldg(reinterpret_cast<const float4 *>(a.w2) + i)
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    auto load_row = [&](int b, float (&hv)[2]) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = t + j * kHeadThreads;
            float v = 0.f;
            if (k < kHElems) {
                if (a.nsplit > 0) {
                    v = sum_partials<kMaxFc1Split>(a.part + (size_t)b * 512 + k, a.nsplit, (size_t)a.B * 512);
                    v = fmaxf(v + __ldg(a.b1 + k), 0.f);
                    a.h_out[(size_t)b * kHElems + k] = v;
                } else {
                    v = a.h_in[(size_t)b * kHElems + k];
                }
            }
            hv[j] = v;
        }
    };
    float hv[2], hv_next[2] = {0.f, 0.f};
    load_row(blockIdx.x, hv_next);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int i = t + j * kHeadThreads;
        if (i < kClasses * kHElems / 4) reinterpret_cast<float4 *>(w2s)[i] = wv[j];
    }
    __syncthreads();
    // One row per CTA at the reference batch; at large batch the grid is capped (launch_head) and a CTA walks several
    // rows with W2 staged once (20 KB per row would otherwise be re-read from L2 4096 times), next row's h in flight.
    for (int b = blockIdx.x; b < a.B; b += gridDim.x) {
    hv[0] = hv_next[0];
    hv[1] = hv_next[1];
    if (b + (int)gridDim.x < a.B) load_row(b + gridDim.x, hv_next);
    // ---- z = h W2^T + b2 ----
    float acc[kClasses];
#pragma unroll
    for (int c = 0; c < kClasses; ++c) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = t + j * kHeadThreads;
            if (k < kHElems) s = fmaf(hv[j], w2s[c * kHElems + k], s);
        }
        acc[c] = warp_sum(s);
    }
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < kClasses; ++c) wred[warp][c] = acc[c];
    }
    __syncthreads();
    if (t < kClasses) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kHeadThreads / 32; ++w) s += wred[w][t];
        zs[t] = s + __ldg(a.b2 + t);
    }
    __syncthreads();
    if (warp == 0) {
        // log_softmax + NLL + dz2 on 10 lanes of one warp (shuffle reductions)
        const float z = lane < kClasses ? zs[lane] : -INFINITY;
        float m = z;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
        const unsigned at_max = __ballot_sync(0xffffffffu, lane < kClasses && z == m);
        const int am = __ffs(at_max) - 1;  // first maximum, like torch.max
        const float e = lane < kClasses ? expf(z - m) : 0.f;
        const float se = warp_sum(e);
        const float lse = m + logf(se);
        const long long tg = a.target[b];
        const bool ok = tg >= 0 && tg < kClasses;
        if (lane < kClasses) {
            const float lp = z - lse;
            const bool hit = ok && lane == (int)tg;
            if (a.logp) a.logp[(size_t)b * kClasses + lane] = lp;
            const float d = (expf(lp) - (hit ? 1.f : 0.f)) * a.inv_B;
            dzs[lane] = d;
            if (a.dz2) a.dz2[(size_t)b * kClasses + lane] = d;
            if (hit) a.rows[b] = -lp;
        }
        if (lane == 0) {
            if (!ok) a.rows[b] = __int_as_float(0x7fc00000);  // NaN for an out-of-range target
            if (a.want_correct) a.rows[a.B + b] = (ok && am == (int)tg) ? 1.f : 0.f;
        }
    }
    __syncthreads();
    if (a.dz1) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = t + j * kHeadThreads;
            if (k < kHElems) {
                float s = 0.f;
#pragma unroll
                for (int c = 0; c < kClasses; ++c) s = fmaf(dzs[c], w2s[c * kHElems + k], s);
                const float d = hv[j] > 0.f ? s : 0.f;
                a.dz1[(size_t)b * kHElems + k] = d;
                if (TC_BWD && x.dz1_hi) {
                    const float hi = tf32_hi(d);
                    x.dz1_hi[(size_t)b * kHElems + k] = hi;
                    x.dz1_lo[(size_t)b * kHElems + k] = d - hi;
                }
            }
        }
    }
    }   // rows of this CTA
    if (t == 0) {
        is_last = false;
        if (!a.defer_reduce) {   // this CTA's rows[] entries are written: take a ticket
            __threadfence();
            is_last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (!is_last) return;
    // ---- deterministic last-CTA reduction of the per-row losses ----
    __threadfence();
    float s = 0.f, c = 0.f;
    for (int r = t; r < a.B; r += kHeadThreads) {
        s += __ldcg(a.rows + r);
        if (a.want_correct) c += __ldcg(a.rows + a.B + r);
    }
    red[t] = s;
    red2[t] = c;
    __syncthreads();
    for (int o = kHeadThreads / 2; o > 0; o >>= 1) {
        if (t < o) { red[t] += red[t + o]; red2[t] += red2[t + o]; }
        __syncthreads();
    }
    if (t == 0) {
        a.stats_out[0] = red[0] * a.loss_scale;
        if (a.want_correct) a.stats_out[1] = red2[0];
        *a.ticket = 0u;
    }
}

__global__ void __launch_bounds__(kHeadThreads) head_kernel(const HeadArgs a) { head_body<false>(a, HeadTcExtra{}); }
__global__ void __launch_bounds__(kHeadThreads) head_tc_kernel(const HeadArgs a, const HeadTcExtra x) { head_body<true>(a, x); }

// ---------------------------------------------------------------------------------
// General backward of log_softmax + fc2 for an arbitrary upstream gradient dlogp (the
// autograd bridge; the fused step computes the NLL case inside head_kernel):
//   dz2 = dlogp - exp(logp) * sum_j dlogp ;  dz1 = (dz2 W2) * (h > 0).   One CTA per row.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
head_bwd_kernel(const float *__restrict__ logp, const float *__restrict__ dlogp, const float *__restrict__ h,
                const float *__restrict__ w2, float *__restrict__ dz2, float *__restrict__ dz1) {
    __shared__ float dzs[16];
    const int b = blockIdx.x, t = threadIdx.x;
    if (t == 0) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < kClasses; ++j) s += dlogp[(size_t)b * kClasses + j];
#pragma unroll
        for (int j = 0; j < kClasses; ++j) {
            const float d = dlogp[(size_t)b * kClasses + j] - expf(logp[(size_t)b * kClasses + j]) * s;
            dzs[j] = d;
            dz2[(size_t)b * kClasses + j] = d;
        }
    }
    __syncthreads();
    for (int k = t; k < kHElems; k += 128) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < kClasses; ++j) s = fmaf(dzs[j], __ldg(w2 + j * kHElems + k), s);
        dz1[(size_t)b * kHElems + k] = h[(size_t)b * kHElems + k] > 0.f ? s : 0.f;
    }
}

}  // namespace b200mnist
