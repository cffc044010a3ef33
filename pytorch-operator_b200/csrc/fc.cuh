// Kernels for the two linear layers, the log-softmax/NLL head and their backward
// (mnist.py:31-33,41-42).
#pragma once
#include "common.cuh"

namespace b200mnist {

// ---------------------------------------------------------------------------------
// fp32 SIMT GEMM  C[M,N] = sum_k A(m,k) * B(k,n), shared-memory tiled, register blocked,
// optional split-K across blockIdx.z (partials are reduced by the consumer kernel, so
// a split costs no extra launch).  fp32 FMA keeps the 1e-5 parity bound that a single
// TF32 pass (10-bit mantissa) cannot.
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

template <int BM, int BN, int TM, int TN, bool A_KCONTIG, bool B_KCONTIG>
__global__ void __launch_bounds__((BM / TM) * (BN / TN)) gemm_kernel(const GemmArgs g) {
    constexpr int NT = (BM / TM) * (BN / TN);
    constexpr int LDA_S = BM + 4, LDB_S = BN + 4;
    __shared__ __align__(16) float As[kGemmBK * LDA_S];
    __shared__ __align__(16) float Bs[kGemmBK * LDB_S];
    const int tid = threadIdx.x, tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * g.kchunk, kend = min(g.K, kbeg + g.kchunk);
    float acc[TM][TN];
    float rs[TM];
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        rs[i] = 0.f;
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    }
    const bool want_rs = (g.rowsum != nullptr) && blockIdx.x == 0 && tx == 0;

    for (int k0 = kbeg; k0 < kend; k0 += kGemmBK) {
        // ---- stage A tile (BM x BK) as As[k][m] ----
        if (A_KCONTIG) {
            for (int i = tid; i < BM * (kGemmBK / 4); i += NT) {
                const int m = i / (kGemmBK / 4), k4 = (i % (kGemmBK / 4)) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (m0 + m < g.M && k0 + k4 < kend)
                    v = __ldg(reinterpret_cast<const float4 *>(g.A + (size_t)(m0 + m) * g.lda + k0 + k4));
                As[(k4 + 0) * LDA_S + m] = v.x;
                As[(k4 + 1) * LDA_S + m] = v.y;
                As[(k4 + 2) * LDA_S + m] = v.z;
                As[(k4 + 3) * LDA_S + m] = v.w;
            }
        } else {
            for (int i = tid; i < kGemmBK * (BM / 4); i += NT) {
                const int k = i / (BM / 4), m4 = (i % (BM / 4)) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k0 + k < kend && m0 + m4 < g.M)
                    v = __ldg(reinterpret_cast<const float4 *>(g.A + (size_t)(k0 + k) * g.lda + m0 + m4));
                *reinterpret_cast<float4 *>(&As[k * LDA_S + m4]) = v;
            }
        }
        // ---- stage B tile (BK x BN) as Bs[k][n] ----
        if (B_KCONTIG) {
            for (int i = tid; i < BN * (kGemmBK / 4); i += NT) {
                const int n = i / (kGemmBK / 4), k4 = (i % (kGemmBK / 4)) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n0 + n < g.N && k0 + k4 < kend)
                    v = __ldg(reinterpret_cast<const float4 *>(g.B + (size_t)(n0 + n) * g.ldb + k0 + k4));
                Bs[(k4 + 0) * LDB_S + n] = v.x;
                Bs[(k4 + 1) * LDB_S + n] = v.y;
                Bs[(k4 + 2) * LDB_S + n] = v.z;
                Bs[(k4 + 3) * LDB_S + n] = v.w;
            }
        } else {
            for (int i = tid; i < kGemmBK * (BN / 4); i += NT) {
                const int k = i / (BN / 4), n4 = (i % (BN / 4)) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k0 + k < kend && n0 + n4 < g.N)
                    v = __ldg(reinterpret_cast<const float4 *>(g.B + (size_t)(k0 + k) * g.ldb + n0 + n4));
                *reinterpret_cast<float4 *>(&Bs[k * LDB_S + n4]) = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < kGemmBK; ++kk) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[kk * LDA_S + ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[kk * LDB_S + tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i) {
                if (want_rs) rs[i] += a[i];
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
    float *C = g.C + (size_t)blockIdx.z * g.M * g.ldc;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + ty * TM + i;
        if (m >= g.M) continue;
        if (want_rs) g.rowsum[m] = rs[i];
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n >= g.N) continue;
            float v = acc[i][j];
            if (g.bias) v += __ldg(g.bias + n);
            if (g.relu) v = fmaxf(v, 0.f);
            C[(size_t)m * g.ldc + n] = v;
        }
    }
}

// ---------------------------------------------------------------------------------
// Head: (optional) finish fc1 from its split-K partials + bias + ReLU, then fc2,
// log_softmax, NLL and the first backward steps, one CTA per batch row:
//   h      = relu(sum_s part[s][b][:] + b1)           (nsplit > 0; also written to h_out)
//   z      = h W2^T + b2 ; logp = z - logsumexp(z)      (warp-shuffle dot products)
//   rows[b]= -logp[target_b]  (mode 0/1)  ;  dz2 = (exp(logp) - onehot) * inv_B
//   dz1    = (dz2 W2) * (h > 0)
// The last CTA to finish (threadfence + ticket) reduces rows[] in a fixed order, so the
// scalar loss is deterministic and needs no second launch.
//   stats_out[0] = sum(rows) * loss_scale ; stats_out[1] = #(argmax == target) if want_correct
// ---------------------------------------------------------------------------------
constexpr int kHeadThreads = 128;

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
    float inv_B, loss_scale;
};

__global__ void __launch_bounds__(kHeadThreads) head_kernel(const HeadArgs a) {
    __shared__ float hs[512];
    __shared__ float zs[16];
    __shared__ float dzs[16];
    __shared__ float red[kHeadThreads];
    __shared__ float red2[kHeadThreads];
    __shared__ bool is_last;
    const int b = blockIdx.x, t = threadIdx.x, warp = t >> 5, lane = t & 31;
    for (int k = t; k < kHElems; k += kHeadThreads) {
        float v;
        if (a.nsplit > 0) {
            v = __ldg(a.b1 + k);
            for (int s = 0; s < a.nsplit; ++s) v += a.part[((size_t)s * a.B + b) * 512 + k];
            v = fmaxf(v, 0.f);
            a.h_out[(size_t)b * kHElems + k] = v;
        } else {
            v = a.h_in[(size_t)b * kHElems + k];
        }
        hs[k] = v;
    }
    __syncthreads();
    for (int j = warp; j < kClasses; j += kHeadThreads / 32) {
        float s = 0.f;
        for (int k = lane; k < kHElems; k += 32) s = fmaf(hs[k], __ldg(a.w2 + j * kHElems + k), s);
        s = warp_sum(s);
        if (lane == 0) zs[j] = s + __ldg(a.b2 + j);
    }
    __syncthreads();
    if (t == 0) {
        float m = zs[0];
        int am = 0;
#pragma unroll
        for (int j = 1; j < kClasses; ++j)
            if (zs[j] > m) { m = zs[j]; am = j; }
        float se = 0.f;
#pragma unroll
        for (int j = 0; j < kClasses; ++j) se += expf(zs[j] - m);
        const float lse = m + logf(se);
        const long long tg = a.target[b];
        const bool ok = tg >= 0 && tg < kClasses;
        float nll = __int_as_float(0x7fc00000);  // NaN for an out-of-range target
#pragma unroll
        for (int j = 0; j < kClasses; ++j) {
            const float lp = zs[j] - lse;
            if (a.logp) a.logp[(size_t)b * kClasses + j] = lp;
            if (ok && j == (int)tg) nll = -lp;
            const float d = (expf(lp) - ((ok && j == (int)tg) ? 1.f : 0.f)) * a.inv_B;
            dzs[j] = d;
            if (a.dz2) a.dz2[(size_t)b * kClasses + j] = d;
        }
        a.rows[b] = nll;
        if (a.want_correct) a.rows[a.B + b] = (ok && am == (int)tg) ? 1.f : 0.f;
    }
    __syncthreads();
    if (a.dz1) {
        for (int k = t; k < kHElems; k += kHeadThreads) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < kClasses; ++j) s = fmaf(dzs[j], __ldg(a.w2 + j * kHElems + k), s);
            a.dz1[(size_t)b * kHElems + k] = hs[k] > 0.f ? s : 0.f;
        }
    }
    // ---- deterministic last-CTA reduction of the per-row losses ----
    if (t == 0) {
        __threadfence();
        const unsigned tk = atomicAdd(a.ticket, 1u);
        is_last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
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

// ---------------------------------------------------------------------------------
// General backward of log_softmax + fc2 for an arbitrary upstream gradient dlogp (the
// autograd bridge; the fused step computes the NLL case inside head_kernel):
//   dz2 = dlogp - exp(logp) * sum_j dlogp ;  dz1 = (dz2 W2) * (h > 0).   One CTA per row.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kHeadThreads)
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
    for (int k = t; k < kHElems; k += kHeadThreads) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < kClasses; ++j) s = fmaf(dzs[j], __ldg(w2 + j * kHElems + k), s);
        dz1[(size_t)b * kHElems + k] = h[(size_t)b * kHElems + k] > 0.f ? s : 0.f;
    }
}

// ---------------------------------------------------------------------------------
// fc2 weight/bias gradient: gW2[j,k] = sum_b dz2[b,j] h[b,k], gb2[j] = sum_b dz2[b,j].
// Thread = one k (coalesced over h rows), 10 accumulators; blockIdx.y splits the batch,
// each split writes its own partial [5056 + 64] (reduced by grad_finalize / fused SGD).
// ---------------------------------------------------------------------------------
constexpr int kFc2wThreads = 128;
constexpr int kFc2PartStride = 5120;  // 5000 weights @0, 10 biases @5056

__global__ void __launch_bounds__(kFc2wThreads)
fc2_wgrad_kernel(const float *__restrict__ dz2, const float *__restrict__ h, float *__restrict__ part, int B,
                 int rows_per_split) {
    __shared__ float ds[64 * kClasses];
    const int t = threadIdx.x, k = blockIdx.x * kFc2wThreads + t;
    const int bbeg = blockIdx.y * rows_per_split, bend = min(B, bbeg + rows_per_split);
    float acc[kClasses], accb = 0.f;
#pragma unroll
    for (int j = 0; j < kClasses; ++j) acc[j] = 0.f;
    for (int c0 = bbeg; c0 < bend; c0 += 64) {
        const int nb = min(64, bend - c0);
        __syncthreads();
        for (int i = t; i < nb * kClasses; i += kFc2wThreads) ds[i] = dz2[(size_t)c0 * kClasses + i];
        __syncthreads();
        for (int r = 0; r < nb; ++r) {
            const float hv = k < kHElems ? h[(size_t)(c0 + r) * kHElems + k] : 0.f;
#pragma unroll
            for (int j = 0; j < kClasses; ++j) acc[j] = fmaf(ds[r * kClasses + j], hv, acc[j]);
            if (blockIdx.x == 0 && t < kClasses) accb += ds[r * kClasses + t];
        }
    }
    float *p = part + (size_t)blockIdx.y * kFc2PartStride;
    if (k < kHElems) {
#pragma unroll
        for (int j = 0; j < kClasses; ++j) p[j * kHElems + k] = acc[j];
    }
    if (blockIdx.x == 0 && t < kClasses) p[5056 + t] = accb;
}

}  // namespace b200mnist
