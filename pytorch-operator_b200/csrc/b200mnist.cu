// libb200mnist.so -- C ABI over the sm_100a kernels (declared in include/b200mnist.h).
// Host code here only validates arguments, carves the caller's workspace and enqueues
// kernels on the caller's stream: no allocation, no synchronisation, no retained state.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/b200mnist.h"
#include "common.cuh"
#include "conv_bwd.cuh"
#include "conv_fwd.cuh"
#include "fc.cuh"
#include "optim.cuh"
#include "peer_exchange.cuh"
#include "tc_gemm.cuh"
#include "tc_conv_bwd.cuh"

using namespace b200mnist;

namespace {

thread_local std::string g_last_error;

int fail_arg(int code, const char *msg) {
    g_last_error = msg;
    return -10000 - code;
}
int fail_cuda(cudaError_t e, const char *where) {
    g_last_error = std::string(where) + ": " + cudaGetErrorString(e);
    return -(int)e;
}
// ---- instrumentation: launch counter + optional per-kernel CUDA-event timing ----
constexpr int kProfMax = 64;
struct Profiler {
    bool enabled = false;
    int count = 0;
    cudaEvent_t ev[kProfMax + 1] = {};
    bool created = false;
    const char *names[kProfMax] = {};
};
Profiler g_prof;
long long g_launches = 0;

inline void note_launch(const char *where, cudaStream_t s) {
    ++g_launches;
    if (g_prof.enabled && g_prof.count < kProfMax) {
        g_prof.names[g_prof.count] = where;
        cudaEventRecord(g_prof.ev[++g_prof.count], s);
    }
}
#define CHECK_LAUNCH(where, stream)                           \
    do {                                                      \
        cudaError_t e_ = cudaGetLastError();                  \
        if (e_ != cudaSuccess) return fail_cuda(e_, where);   \
        note_launch(where, stream);                           \
    } while (0)

// Launch with programmatic dependent launch (PDL): the grid may be scheduled while its predecessor on the
// stream is still draining; the kernel itself blocks in pdl_wait() until the predecessor has completed and
// flushed.  Hides the ~2 us launch-to-launch gap of each of the step's dependent kernels, also under CUDA-graph
// replay (stream capture turns the attribute into programmatic graph edges).
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define CHECK_PDL(call, where, stream)                        \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return fail_cuda(e_, where);   \
        note_launch(where, stream);                           \
    } while (0)

// Opt a kernel in to more than 48 KB of dynamic shared memory, once per device (function attributes are per device).
template <typename K>
cudaError_t ensure_dynamic_smem(K kernel, int bytes, unsigned long long *done_mask) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64 && (*done_mask >> dev) & 1ull) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev >= 0 && dev < 64) *done_mask |= 1ull << dev;
    return e;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- how each batch-dependent kernel is split (a pure function of B) ----
struct Plan {
    int fc1_split, fc1_kchunk;  // fc1 forward split-K
    int dg_split, dg_kchunk;    // fc1 data-gradient split-K
    int fc2w_split, fc2w_rows;  // fc2 weight-gradient batch split
    int c2w_split;              // conv2 weight-gradient batch split (= number of partials)
    bool c2w_tc, c2dg_tc;       // conv2 weight / data gradient on the tensor cores (tc_conv_bwd.cuh)
    bool c2_presplit;           // training step: the forward's weight-split launch also builds the data-gradient operand
    int c1w_split;              // conv1 weight-gradient batch split
    int c2dg_gx;                // conv2 data-gradient CTAs per channel tile
};
bool tc_path_enabled(const char *env, int B);
Plan make_plan(int B) {
    Plan p;
    const int gy = ceil_div(B, 64);
    p.fc1_split = clampi(148 / (16 * gy), 1, 10);  // 16 n-tiles x gy m-tiles x split <= one wave of 148 SMs
    p.fc1_kchunk = ceil_div(ceil_div(800, p.fc1_split), kGemmBK) * kGemmBK;
    p.fc1_split = ceil_div(800, p.fc1_kchunk);
    p.dg_split = clampi(ceil_div(100, 25 * gy), 1, 8);
    p.dg_kchunk = ceil_div(ceil_div(500, p.dg_split), kGemmBK) * kGemmBK;
    p.dg_split = ceil_div(500, p.dg_kchunk);
    p.fc2w_split = clampi(ceil_div(B, 16), 1, 32);
    p.fc2w_rows = ceil_div(B, p.fc2w_split);
    p.fc2w_split = ceil_div(B, p.fc2w_rows);
    p.c2w_split = clampi(ceil_div(B, 2), 1, 32);
    // small B: one image per CTA; large B: 148 x 2 channel halves = 296 CTAs = exactly one wave at 2 CTAs/SM
    p.c1w_split = clampi(B <= 64 ? B : B / 2, 1, 148);
    // conv2 backward runs 2 CTAs/SM: 5*c2w_split weight-gradient CTAs + 5*c2dg_gx data-gradient CTAs <= 296
    p.c2dg_gx = clampi(B, 1, (296 - 5 * p.c2w_split) / 5);
    // tuning overrides (development only): B200MNIST_PLAN="fc1_split,dg_split,c2w_split,c2dg_gx" (0 = keep)
    static int ov[4] = {-1, 0, 0, 0};
    if (ov[0] < 0) {
        ov[0] = 0;
        if (const char *e = getenv("B200MNIST_PLAN")) sscanf(e, "%d,%d,%d,%d", &ov[0], &ov[1], &ov[2], &ov[3]);
    }
    if (ov[0] > 0) {   // clamped to what the workspace regions and sum_partials<> are sized for
        p.fc1_kchunk = ceil_div(ceil_div(800, clampi(ov[0], 1, 10)), kGemmBK) * kGemmBK;
        p.fc1_split = ceil_div(800, p.fc1_kchunk);
    }
    if (ov[1] > 0) {
        p.dg_kchunk = ceil_div(ceil_div(500, clampi(ov[1], 1, 8)), kGemmBK) * kGemmBK;
        p.dg_split = ceil_div(500, p.dg_kchunk);
    }
    if (ov[2] > 0) p.c2w_split = clampi(ov[2], 1, 32);
    if (ov[3] > 0) p.c2dg_gx = clampi(ov[3], 1, B);
    // conv2 backward on the tensor cores: auto from kTcConvMinBatch up; B200MNIST_CONV2_WGRAD_TC / _DGRAD_TC = 1 / 0 force
    p.c2w_tc = tc_path_enabled("B200MNIST_CONV2_WGRAD_TC", B);
    p.c2dg_tc = tc_path_enabled("B200MNIST_CONV2_DGRAD_TC", B);
    if (p.c2w_tc) p.c2w_split = B < tc::kWgMaxSlices ? B : tc::kWgMaxSlices;
    p.c2_presplit = p.c2dg_tc && tc_path_enabled("B200MNIST_CONV2_TC", B);
    return p;
}

// ---- workspace carving (all regions 256-byte aligned) ----
struct Workspace {
    unsigned *ticket;
    float *stats;  // [0] loss / nll sum, [1] correct
    float *p1, *p2, *h, *logp, *dz2, *dz1, *dp1, *rows;
    uint8_t *i1, *i2;
    float *fc1_part, *dp2_part, *fc2w_part, *c2w_part, *c1w_part;
    float *w2s;  // conv2.weight split into hi/lo tensor-core operands (tc_conv.cuh)
    float *wd;   // the same for the data-gradient GEMM (tc_conv_bwd.cuh)
    float *fc1ops;  // hi/lo copies of fc1's GEMM operands (Fc1Ops)
    size_t bytes;
};
// hi / lo copies of the operands of fc1's three tcgen05 GEMMs (tc_gemm.cuh), written by split_rows_kernel once per step:
// p2 [B][832] (800 features, column 800 of the hi copy = 1 for the bias gradient, rest zero), dz1 [B][500], W1 [500][800].
constexpr int kP2Ld = kP2SplitLd;
constexpr int kFc1TcFloor = 128;   // fc1's tensor-core GEMMs (and their operand copies) exist from this batch up
struct Fc1Ops {
    float *p2_hi, *p2_lo, *dz1_hi, *dz1_lo, *w1_hi, *w1_lo;
};
inline size_t fc1ops_bytes(int B) {
    return align256((size_t)B * kP2Ld * 4) * 2 + align256((size_t)B * kHElems * 4) * 2 + align256((size_t)500 * 800 * 4) * 2;
}
inline Fc1Ops fc1ops_carve(void *base, int B) {
    char *p = static_cast<char *>(base);
    Fc1Ops o;
    o.p2_hi = reinterpret_cast<float *>(p); p += align256((size_t)B * kP2Ld * 4);
    o.p2_lo = reinterpret_cast<float *>(p); p += align256((size_t)B * kP2Ld * 4);
    o.dz1_hi = reinterpret_cast<float *>(p); p += align256((size_t)B * kHElems * 4);
    o.dz1_lo = reinterpret_cast<float *>(p); p += align256((size_t)B * kHElems * 4);
    o.w1_hi = reinterpret_cast<float *>(p); p += align256((size_t)500 * 800 * 4);
    o.w1_lo = reinterpret_cast<float *>(p);
    return o;
}
Workspace carve(void *base, int B) {
    // Partial-sum regions are sized for the worst split of any batch <= B.
    Workspace w;
    size_t off = 0;
    auto take = [&](size_t nbytes) {
        void *p = base ? static_cast<char *>(base) + off : nullptr;
        off += align256(nbytes);
        return p;
    };
    const size_t Bz = (size_t)B;
    w.ticket = static_cast<unsigned *>(take(256));
    w.stats = static_cast<float *>(take(256));
    w.p1 = static_cast<float *>(take(Bz * kP1Elems * 4));
    w.i1 = static_cast<uint8_t *>(take(Bz * kP1Elems));
    w.p2 = static_cast<float *>(take(Bz * kP2Elems * 4));
    w.i2 = static_cast<uint8_t *>(take(Bz * kP2Elems));
    w.h = static_cast<float *>(take(Bz * kHElems * 4));
    w.logp = static_cast<float *>(take(Bz * kClasses * 4));
    w.dz2 = static_cast<float *>(take(Bz * kClasses * 4));
    w.dz1 = static_cast<float *>(take(Bz * kHElems * 4));
    w.dp1 = static_cast<float *>(take(Bz * kP1Elems * 4));
    w.rows = static_cast<float *>(take(Bz * 2 * 4));
    w.fc1_part = static_cast<float *>(take(10 * Bz * 512 * 4));
    w.dp2_part = static_cast<float *>(take(8 * Bz * kP2Elems * 4));
    w.fc2w_part = static_cast<float *>(take((size_t)32 * kFc2PartStride * 4));
    w.c2w_part = static_cast<float *>(take((size_t)32 * kC2PartStride * 4));
    w.c1w_part = static_cast<float *>(take((size_t)kMaxGradSplit * kC1PartStride * 4));
    w.w2s = static_cast<float *>(take((size_t)tc::kWnScratchFloats * 4));
    w.wd = static_cast<float *>(take((size_t)tc::kDgScratchFloats * 4));
    w.fc1ops = static_cast<float *>(take(B >= kFc1TcFloor ? fc1ops_bytes(B) : 0));   // never used below that batch
    w.bytes = off;
    return w;
}

int check_step_args(const void *params, const void *x, const void *target, int B, void *workspace,
                    int64_t workspace_bytes) {
    if (!params || !x || !target || !workspace) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(params) || !aligned16(x) || (reinterpret_cast<uintptr_t>(workspace) & 255u))
        return fail_arg(3, "params/x must be 16-byte aligned and workspace 256-byte aligned");
    if ((int64_t)carve(nullptr, B).bytes > workspace_bytes) return fail_arg(4, "workspace too small for this batch");
    return 0;
}

// ---- launches of the individual stages ----
int launch_conv1_fwd(const float *x, const float *w, const float *b, float *out, uint8_t *idx, int B, cudaStream_t s,
                     const RingRef *ring = nullptr) {
    // small batches: one CTA per (image, 4-channel group) fills the GPU; large ones: one CTA per image, all groups
    const dim3 grid(B, B >= 1024 ? 1 : 20 / kC1ChPerCta);
    if (ring)   // resident input: reads the ring slot under the cursor and fills x / target for the rest of the step
        CHECK_PDL(launch_pdl(conv1_relu_pool_fwd_ring_kernel, grid, dim3(kC1Threads), 0, s, *ring, w, b, out, idx, B),
                  "conv1_relu_pool_fwd", s);
    else
        CHECK_PDL(launch_pdl(conv1_relu_pool_fwd_kernel, grid, dim3(kC1Threads), 0, s, x, w, b, out, idx, B),
                  "conv1_relu_pool_fwd", s);
    return 0;
}
// Validates a caller's ring description (include/b200mnist.h) and turns it into the kernels' RingRef; x / target are the
// step's ordinary input buffers, which conv1's forward kernel fills from the ring.
int make_ring(const b200mnist_ring_t *ring, const float *x, const int64_t *target, int B, RingRef *out) {
    if (!ring->x || !ring->target || !ring->state) return fail_arg(1, "null pointer in the ring description");
    if (ring->slots < 1) return fail_arg(2, "the ring needs at least one slot");
    if (!aligned16(ring->x) || ((ring->x_stride * 4) & 15) || ring->x_stride < (int64_t)B * kXElems || ring->t_stride < B ||
        (reinterpret_cast<uintptr_t>(ring->state) & 7u))
        return fail_arg(3, "ring slots must be 16-byte aligned and at least one batch apart, `state` 8-byte aligned");
    out->x = ring->x;
    out->target = reinterpret_cast<const long long *>(ring->target);
    out->x_stride = ring->x_stride;
    out->t_stride = ring->t_stride;
    out->x_copy = const_cast<float *>(x);
    out->t_copy = reinterpret_cast<long long *>(const_cast<int64_t *>(target));
    out->state = ring->state;
    out->landed = ring->landed;
    out->slots = ring->slots;
    return 0;
}
LossReduce ring_loss_reduce(const float *rows, float *loss_out, int B, const b200mnist_ring_t *ring) {
    LossReduce l{rows, loss_out, B, 1.0f / (float)B, nullptr, 0u, nullptr};
    if (ring) {
        l.state = ring->state;
        l.slots = ring->slots;
        l.report_host = reinterpret_cast<unsigned long long *>(ring->report_host);
    }
    return l;
}
const char *kTmapErr = "cuTensorMapEncodeTiled failed or is unavailable (driver too old?)";
// conv2 forward on the tensor cores (tcgen05 3xTF32, shifted-window implicit GEMM; tc_conv.cuh).  `w2s` receives the
// hi/lo-split, operand-ordered copy of conv2.weight (tc::kWnScratchFloats floats), rebuilt on every call.
// fc1ops (optional): the kernel also writes the hi / lo copies of its output for fc1's tensor-core GEMMs; with wd_also
// the weight-split launch also writes the hi / lo copies of fc1.weight (w1).
int launch_conv2_fwd_tc(const float *in, const float *w, const float *b, float *out, uint8_t *idx, int B, float *w2s,
                        cudaStream_t s, float *wd_also = nullptr, const Fc1Ops *fc1ops = nullptr, const float *w1 = nullptr) {
    constexpr int NT = 32;
    static unsigned long long attr_mask = 0;
    cudaError_t ea = ensure_dynamic_smem(tc::conv2_fwd_win_kernel<NT>, tc::wn_smem_bytes<NT>(), &attr_mask);
    if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(conv2_fwd_win)");
    if (wd_also) {
        const bool with_w1 = fc1ops && w1;
        CHECK_PDL(launch_pdl(tc::conv2_weight_split_both_kernel, dim3(25 + 4 * tc::kDgChunks + (with_w1 ? ceil_div(500 * 800, 1024) : 0)),
                             dim3(256), 0, s, w, w2s, wd_also, with_w1 ? w1 : nullptr, with_w1 ? fc1ops->w1_hi : nullptr,
                             with_w1 ? fc1ops->w1_lo : nullptr),
                  "conv2 weight split (hi/lo; forward + dgrad operands)", s);
    } else
        CHECK_PDL(launch_pdl(tc::conv2_weight_split_kernel<NT>, dim3(25), dim3(256), 0, s, w, w2s), "conv2 weight split (hi/lo)", s);
    const int npairs = (B + 1) / 2;
    CHECK_PDL(launch_pdl(tc::conv2_fwd_win_kernel<NT>, dim3(npairs < 74 ? npairs : 74, 64 / NT), dim3(tc::kWnThreads),
                         (size_t)tc::wn_smem_bytes<NT>(), s, in, (const float *)w2s, b, out, idx, B,
                         fc1ops ? fc1ops->p2_hi : nullptr, fc1ops ? fc1ops->p2_lo : nullptr),
              "conv2_relu_pool_fwd (tcgen05 3xTF32 shifted-window GEMM)", s);
    return 0;
}
// Which conv2 kernels run inside the fused step.  Measured step time, tcgen05 conv kernels off / on (tools/
// sweep_tc_threshold.sh): B=96 84.3 / 88.0 us, B=128 94.3 / 90.2 us, B=192 129 / 115 us, B=256 163 / 134 us, B=512 294 / 223 us; at the
// reference batch (64) they lose (their fixed costs -- TMEM allocation, zero-filled operand planes, 60-130 KB of weights
// per CTA -- exceed the whole FFMA2 kernel).  So the conv kernels switch on at kTcConvMinBatch; the fc1 GEMMs need
// M = B large enough to fill 128-row tiles and switch at kTcMinBatch (tools/sweep_fc1_tc.sh, step time SIMT / tcgen05
// fc1: B=256 131 / 137 us, B=384 178 / 165 us, B=512 213 / 182 us, B=768 295 / 239 us).
// B200MNIST_CONV2_TC / _DGRAD_TC / _WGRAD_TC = 1 / 0 force each conv kernel on / off (tests, profiling).
constexpr int kTcMinBatch = 384;
constexpr int kTcConvMinBatch = 128;
int fc1_tc_min_batch() {   // B200MNIST_FC1_TC_MIN overrides kTcMinBatch (threshold sweeps)
    const char *e = getenv("B200MNIST_FC1_TC_MIN");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? (v < kFc1TcFloor ? kFc1TcFloor : v) : kTcMinBatch;
}
bool tc_path_enabled(const char *env, int B) {   // read at every (uncaptured) launch: a few hundred ns, lets tests toggle it
    const char *e = getenv(env);
    return (!e || !e[0]) ? B >= kTcConvMinBatch : e[0] == '1';
}
bool conv2_tc_enabled(int B) { return tc_path_enabled("B200MNIST_CONV2_TC", B); }
int launch_conv2_fwd_simt(const float *in, const float *w, const float *b, float *out, uint8_t *idx, int B,
                          cudaStream_t s);
int launch_conv2_fwd(const float *in, const float *w, const float *b, float *out, uint8_t *idx, int B, float *w2s,
                     cudaStream_t s, float *wd_also = nullptr, const Fc1Ops *fc1ops = nullptr, const float *w1 = nullptr) {
    if (w2s && conv2_tc_enabled(B)) return launch_conv2_fwd_tc(in, w, b, out, idx, B, w2s, s, wd_also, fc1ops, w1);
    return launch_conv2_fwd_simt(in, w, b, out, idx, B, s);
}
int launch_conv2_fwd_simt(const float *in, const float *w, const float *b, float *out, uint8_t *idx, int B,
                          cudaStream_t s) {
    CHECK_PDL(launch_pdl(conv2_relu_pool_fwd_kernel, dim3(ceil_div(B, kC2ImgPerCta), 50 / kC2OcPerCta), dim3(kC2Threads),
                         0, s, in, w, b, out, idx, B),
              "conv2_relu_pool_fwd", s);
    return 0;
}
// ---- fc1 on the tensor cores (tcgen05 3xTF32, TMA-fed, operands pre-split into hi / lo copies): large batches ----
int launch_split(const float *in, int ld_in, float *hi, float *lo, int ld_out, int rows, int cols, int ones_col,
                 const char *what, cudaStream_t s) {
    const long long n4 = (long long)rows * (ld_out / 4);
    CHECK_PDL(launch_pdl(tc::split_rows_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, s, in, ld_in, hi, lo, ld_out,
                         rows, cols, ones_col),
              what, s);
    return 0;
}
// maps of a K-major pair (rows x cols, box_rows rows per tile) / an MN-major pair
bool maps_kmajor(CUtensorMap *hi, CUtensorMap *lo, const float *h, const float *l, int rows, int cols, int ld, int box_rows) {
    return tc::make_map_kmajor(hi, h, rows, cols, ld, box_rows) && tc::make_map_kmajor(lo, l, rows, cols, ld, box_rows);
}
bool maps_mnmajor(CUtensorMap *hi, CUtensorMap *lo, const float *h, const float *l, int rows, int cols, int ld) {
    return tc::make_map_mnmajor(hi, h, rows, cols, ld) && tc::make_map_mnmajor(lo, l, rows, cols, ld);
}
// h = relu(p2 . W1^T + b1): splits p2 (with the bias-gradient column for the backward) and W1, then the GEMM
// p2_ready / w1_ready: the producer of that operand has already written its hi / lo copies into `o` (the fused step:
// conv2's tensor-core forward writes p2's, the conv2 weight-split launch W1's); otherwise a split pass runs here.
int launch_fc1_fwd_tc(const float *in, const float *w, const float *bias, float *h, int B, const Fc1Ops &o, cudaStream_t s,
                      bool p2_ready = false, bool w1_ready = false) {
    int rc;
    if (!p2_ready && (rc = launch_split(in, 800, o.p2_hi, o.p2_lo, kP2Ld, B, 800, 800, "fc1 operand split (p2 hi/lo)", s)))
        return rc;
    if (!w1_ready && (rc = launch_split(w, 800, o.w1_hi, o.w1_lo, 800, 500, 800, -1, "fc1 operand split (W1 hi/lo)", s)))
        return rc;
    tc::TcMaps m;
    if (!maps_kmajor(&m.a_hi, &m.a_lo, o.p2_hi, o.p2_lo, B, 800, kP2Ld, tc::kBM) ||
        !maps_kmajor(&m.b_hi, &m.b_lo, o.w1_hi, o.w1_lo, 500, 800, 800, tc::kBN))
        return fail_arg(8, kTmapErr);
    static unsigned long long attr_mask = 0;
    cudaError_t ea = ensure_dynamic_smem(tc::gemm_tc_kernel<false, false, tc::kEpiBiasRelu>, tc::kSmemBytes, &attr_mask);
    if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(gemm_tc)");
    const tc::TcGemmArgs a{bias, h, nullptr, B, 500, 800, 500};
    CHECK_PDL(launch_pdl(tc::gemm_tc_kernel<false, false, tc::kEpiBiasRelu>, dim3(ceil_div(500, tc::kBN), ceil_div(B, tc::kBM)),
                         dim3(tc::kThreads), (size_t)tc::kSmemBytes, s, m, a),
              "fc1 forward gemm (tcgen05 3xTF32)", s);
    return 0;
}
// dp2[B,800] = dz1[B,500] . W1[500,800];  gW1[500,800] = dz1[B,500]^T . p2[B,800], gb1[500] = column sums of dz1:
// both backward GEMMs of fc1 in one launch (tc::fc1_bwd_tc_kernel).  fwd_ops_ready: the hi / lo copies of p2 and W1
// are the ones the forward of this step left in `o`; otherwise they are made here.
int launch_fc1_bwd_tc(const float *dz1, const float *p2, const float *w1, float *gw1, float *gb1, float *dp2, int B,
                      const Fc1Ops &o, bool fwd_ops_ready, bool dz1_ready, bool grads_zeroed, cudaStream_t s) {
    int rc;
    if (!grads_zeroed) {   // the weight-gradient GEMM adds its two k-halves onto gW1 / gb1 (contiguous in the bucket or not)
        cudaError_t e = cudaMemsetAsync(gw1, 0, (size_t)500 * 800 * 4, s);
        if (e == cudaSuccess) e = cudaMemsetAsync(gb1, 0, (size_t)500 * 4, s);
        if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(fc1 gradients)");
    }
    if (!fwd_ops_ready) {
        if ((rc = launch_split(p2, 800, o.p2_hi, o.p2_lo, kP2Ld, B, 800, 800, "fc1 operand split (p2 hi/lo)", s))) return rc;
        if ((rc = launch_split(w1, 800, o.w1_hi, o.w1_lo, 800, 500, 800, -1, "fc1 operand split (W1 hi/lo)", s))) return rc;
    }
    if (!dz1_ready && (rc = launch_split(dz1, 500, o.dz1_hi, o.dz1_lo, 500, B, 500, -1, "fc1 operand split (dz1 hi/lo)", s)))
        return rc;
    tc::TcMaps wm, dm;
    if (!maps_mnmajor(&wm.a_hi, &wm.a_lo, o.dz1_hi, o.dz1_lo, B, 500, 500) ||
        !maps_mnmajor(&wm.b_hi, &wm.b_lo, o.p2_hi, o.p2_lo, B, 801, kP2Ld) ||   // + the column of ones
        !maps_kmajor(&dm.a_hi, &dm.a_lo, o.dz1_hi, o.dz1_lo, B, 500, 500, tc::kBM) ||
        !maps_mnmajor(&dm.b_hi, &dm.b_lo, o.w1_hi, o.w1_lo, 500, 800, 800))
        return fail_arg(8, kTmapErr);
    static unsigned long long attr_mask = 0;
    cudaError_t ea = ensure_dynamic_smem(tc::fc1_bwd_tc_kernel, tc::kSmemBytes, &attr_mask);
    if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(fc1_bwd_tc)");
    const tc::TcGemmArgs wg{nullptr, gw1, gb1, 500, 800, B, 800}, dg{nullptr, dp2, nullptr, B, 800, 500, 800};
    const int wg_gx = ceil_div(800 + 1, tc::kBN), n_wg = 2 * wg_gx * ceil_div(500, tc::kBM);   // + the bias-gradient column; x 2 k-halves
    const int dg_gx = ceil_div(800, tc::kBN), n_dg = dg_gx * ceil_div(B, tc::kBM);
    CHECK_PDL(launch_pdl(tc::fc1_bwd_tc_kernel, dim3(n_wg + n_dg), dim3(tc::kThreads), (size_t)tc::kSmemBytes, s, wm, wg, dm, dg,
                         wg_gx, n_wg, dg_gx),
              "fc1 wgrad + dgrad gemms (tcgen05 3xTF32, one launch)", s);
    return 0;
}
// fc1 forward: returns the number of split-K partials left in `part` (0 = `h` is final).
int launch_fc1_fwd(const float *in, const float *w, const float *bias, float *h, float *part, int B, const Plan &pl,
                   const Fc1Ops &ops, cudaStream_t s, int *nsplit_out, bool p2_ready = false, bool w1_ready = false) {
    if (B >= fc1_tc_min_batch()) {
        *nsplit_out = 0;
        return launch_fc1_fwd_tc(in, w, bias, h, B, ops, s, p2_ready, w1_ready);
    }
    GemmArgs g{};
    g.A = in; g.B = w; g.M = B; g.N = 500; g.K = 800; g.lda = 800; g.ldb = 800; g.kchunk = pl.fc1_kchunk;
    if (pl.fc1_split == 1) {
        g.C = h; g.ldc = 500; g.bias = bias; g.relu = 1;
    } else {
        g.C = part; g.ldc = 512;
    }
    CHECK_PDL(launch_pdl(gemm_kernel<64, 32, 2, 4, true, true>, dim3(ceil_div(500, 32), ceil_div(B, 64), pl.fc1_split),
                         dim3(256), 0, s, g),
              "fc1 forward gemm", s);
    *nsplit_out = pl.fc1_split == 1 ? 0 : pl.fc1_split;
    return 0;
}
int launch_head(const HeadArgs &a, cudaStream_t s, const HeadTcExtra *x = nullptr) {
    // one row per CTA up to 4 CTAs on every SM, then several rows per CTA (W2 staged once per CTA)
    const dim3 grid(a.B < 592 ? a.B : 592);
    if (x)
        CHECK_PDL(launch_pdl(head_tc_kernel, grid, dim3(kHeadThreads), 0, s, a, *x), "fc2/log_softmax/nll head", s);
    else
        CHECK_PDL(launch_pdl(head_kernel, grid, dim3(kHeadThreads), 0, s, a), "fc2/log_softmax/nll head", s);
    return 0;
}
int launch_fc_bwd(const float *dz2, const float *dz1, const float *h, const float *p2, const float *w1, int B,
                  float *gw1, float *gb1, float *dp2_part, float *fc2w_part, const Plan &pl, const Fc1Ops &ops,
                  bool fwd_ops_ready, bool dz1_ready, bool grads_zeroed, cudaStream_t s) {
    const bool use_tc = B >= fc1_tc_min_batch() && pl.dg_split == 1;  // large batch: both fc1 GEMMs on the tensor cores
    FcBwdArgs a{};
    GemmArgs &g = a.wg;  // gW1[n,k] = sum_b dz1[b,n] p2[b,k] ; gb1[n] = sum_b dz1[b,n]
    g.A = dz1; g.B = p2; g.C = gw1; g.rowsum = gb1;
    g.M = 500; g.N = 800; g.K = B; g.lda = 500; g.ldb = 800; g.ldc = 800; g.kchunk = B;
    a.wg_gx = ceil_div(800, 64);
    a.wg_n = a.wg_gx * ceil_div(500, 64);
    GemmArgs &d = a.dg;  // dp2[b,k] = sum_n dz1[b,n] W1[n,k]
    d.A = dz1; d.B = w1; d.C = dp2_part;
    d.M = B; d.N = 800; d.K = 500; d.lda = 500; d.ldb = 800; d.ldc = 800; d.kchunk = pl.dg_kchunk;
    a.dg_gx = ceil_div(800, 32);
    a.dg_gy = ceil_div(B, 64);
    a.dg_n = a.dg_gx * a.dg_gy * pl.dg_split;
    a.dz2 = dz2; a.h = h; a.fc2_part = fc2w_part; a.B = B; a.fc2_rows = pl.fc2w_rows;
    a.fc2_gx = ceil_div(500, kFc2wThreads);
    if (use_tc) {
        a.wg_n = 0;
        a.dg_n = 0;  // the SIMT launch keeps only the (tiny) fc2 weight-gradient role
        int rc = launch_fc1_bwd_tc(dz1, p2, w1, gw1, gb1, dp2_part, B, ops, fwd_ops_ready, dz1_ready, grads_zeroed, s);
        if (rc) return rc;
    }
    CHECK_PDL(launch_pdl(fc_bwd_kernel, dim3(a.wg_n + a.dg_n + a.fc2_gx * pl.fc2w_split), dim3(256), 0, s, a),
              "fc backward (fc1 wgrad + fc1 dgrad + fc2 wgrad)", s);
    return 0;
}
int launch_conv2_bwd(const float *dp2, int nsplit, size_t split_stride, const float *p2, const uint8_t *i2,
                     const float *p1, const float *w, int B, float *dp1, float *c2w_part, float *wd, const Plan &pl,
                     cudaStream_t s, bool wd_ready = false, const EarlySignal *sig = nullptr, bool *sig_sent = nullptr) {
    static unsigned long long attr_mask = 0;
    cudaError_t ea = ensure_dynamic_smem(conv2_bwd_kernel, kConv2BwdSmemBytes, &attr_mask);
    if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(conv2_bwd)");
    Conv2BwdArgs a{};
    a.dp2 = dp2; a.p2 = p2; a.p1 = p1; a.w = w; a.i2 = i2; a.dp1 = dp1; a.part = c2w_part;
    a.split_stride = split_stride; a.nsplit = nsplit; a.B = B;
    a.wg_gy = pl.c2w_split; a.n_wg = 5 * pl.c2w_split; a.dg_gx = pl.c2dg_gx;
    if (sig) a.sig = *sig;
    const bool dg_tc = pl.c2dg_tc && wd != nullptr;
    if (sig_sent) *sig_sent = sig && !(dg_tc && pl.c2w_tc);   // the SIMT kernel (which carries the signal) runs at all
    if (!dg_tc && !pl.c2w_tc) {
        CHECK_PDL(launch_pdl(conv2_bwd_kernel, dim3(a.n_wg + 5 * a.dg_gx), dim3(kConvBwdThreads), kConv2BwdSmemBytes, s, a),
                  "conv2 backward (wgrad + dgrad)", s);
        return 0;
    }
    if (dg_tc) {
        static unsigned long long tc_mask = 0;
        ea = ensure_dynamic_smem(tc::conv2_dgrad_tc_kernel, tc::kDgSmemBytes, &tc_mask);
        if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(conv2_dgrad_tc)");
        if (!wd_ready)   // (the fused training step builds it together with the forward operand)
            CHECK_PDL(launch_pdl(tc::conv2_dgrad_weight_split_kernel, dim3(4 * tc::kDgChunks), dim3(128), 0, s, w, wd),
                      "conv2 dgrad weight split (hi/lo)", s);
        tc::Conv2DgradTcArgs t{};
        t.dp2 = dp2; t.p2 = p2; t.wd = wd; t.i2 = i2; t.dp1 = dp1; t.split_stride = split_stride; t.nsplit = nsplit; t.B = B;
        const int npairs = (B + 1) / 2;
        CHECK_PDL(launch_pdl(tc::conv2_dgrad_tc_kernel, dim3(npairs < 37 ? npairs : 37, 4), dim3(tc::kDgThreads),
                             (size_t)tc::kDgSmemBytes, s, t),
                  "conv2 dgrad (tcgen05 3xTF32 GEMM + col2im)", s);
    }
    if (pl.c2w_tc) {
        static unsigned long long tc_mask = 0;
        ea = ensure_dynamic_smem(tc::conv2_wgrad_tc_kernel, tc::kWgSmemBytes, &tc_mask);
        if (ea != cudaSuccess) return fail_cuda(ea, "cudaFuncSetAttribute(conv2_wgrad_tc)");
        tc::Conv2WgradTcArgs t{};
        t.dp2 = dp2; t.p2 = p2; t.p1 = p1; t.i2 = i2; t.part = c2w_part; t.split_stride = split_stride; t.nsplit = nsplit;
        t.B = B; t.nslices = pl.c2w_split;
        CHECK_PDL(launch_pdl(tc::conv2_wgrad_tc_kernel, dim3(pl.c2w_split, 5), dim3(tc::kWgThreads), (size_t)tc::kWgSmemBytes,
                             s, t),
                  "conv2 wgrad (tcgen05 3xTF32, all four hi/lo products per MMA)", s);
    }
    if (!dg_tc || !pl.c2w_tc) {   // the remaining half on the fp32 SIMT kernel
        if (dg_tc) a.dg_gx = 0;
        if (pl.c2w_tc) a.n_wg = 0;
        CHECK_PDL(launch_pdl(conv2_bwd_kernel, dim3(a.n_wg + 5 * a.dg_gx), dim3(kConvBwdThreads), kConv2BwdSmemBytes, s, a),
                  dg_tc ? "conv2 wgrad" : "conv2 dgrad", s);
    }
    return 0;
}
// What the fused DDP step adds to the backward: the peer windows and the SGD constants of this step.
struct PeerStep {
    PeerWin pw;
    float *m;
    int parity;
    float neg_lr, mu, scale;
    int first;
};
template <int WORLD>
int launch_conv1_bwd_peer(const float *dp1, const float *p1, const uint8_t *i1, const float *x, int B, float *c1w_part,
                          const Plan &pl, const PeerStep &ps, bool presignaled, cudaStream_t s, const LossReduce *early) {
    // exchange CTAs: one float4 of this rank's fc1 slice per thread, an even count split over the two grid rows
    constexpr int per = (kFc1Vec4 + WORLD - 1) / WORLD;
    const int nx = ceil_div(ceil_div(per, kC1wThreads), 2);
    CHECK_PDL(launch_pdl(conv1_wgrad_peer_kernel<WORLD>, dim3(nx + pl.c1w_split + (early ? 1 : 0), 2), dim3(kC1wThreads), 0, s,
                         x, dp1, p1, i1, c1w_part, B, ps.pw, ps.m, ps.parity, ps.neg_lr, ps.mu, ps.scale, ps.first, nx,
                         presignaled ? 1 : 0, early ? *early : LossReduce{}, early ? 1 : 0),
              "conv1 wgrad + fc1 slice exchange (reduce + sgd + push over NVLink)", s);
    return 0;
}
int launch_conv1_bwd(const float *dp1, const float *p1, const uint8_t *i1, const float *x, int B, float *c1w_part,
                     const Plan &pl, cudaStream_t s, const PeerStep *ps = nullptr, bool presignaled = false,
                     const LossReduce *early_finish = nullptr) {
    if (ps) {
        switch (ps->pw.world) {
        case 2: return launch_conv1_bwd_peer<2>(dp1, p1, i1, x, B, c1w_part, pl, *ps, presignaled, s, early_finish);
        case 4: return launch_conv1_bwd_peer<4>(dp1, p1, i1, x, B, c1w_part, pl, *ps, presignaled, s, early_finish);
        default: return launch_conv1_bwd_peer<8>(dp1, p1, i1, x, B, c1w_part, pl, *ps, presignaled, s, early_finish);
        }
    }
    if (early_finish)   // host-fed ring: one more grid column closes the step on the ring (loss, cursor, report to the host)
        CHECK_PDL(launch_pdl(conv1_wgrad_finish_kernel, dim3(pl.c1w_split + 1, 2), dim3(kC1wThreads), 0, s, x, dp1, p1, i1,
                             c1w_part, B, *early_finish),
                  "conv1 wgrad", s);
    else
        CHECK_PDL(launch_pdl(conv1_wgrad_kernel, dim3(pl.c1w_split, 2), dim3(kC1wThreads), 0, s, x, dp1, p1, i1, c1w_part, B),
                  "conv1 wgrad", s);
    return 0;
}
template <int WORLD>
int launch_peer_tail(const PeerStep &ps, const GradSegments &segs, const LossReduce &lrd, cudaStream_t s) {
    CHECK_PDL(launch_pdl(peer_tail_kernel<WORLD>, dim3(kFusedCovBlocks + 1), dim3(256), 0, s, ps.pw, ps.m, ps.parity, ps.neg_lr,
                         ps.mu, ps.scale, ps.first, segs, lrd),
              "step tail: gradient reduce + push + rank-ordered sum + sgd (NVLink P2P)", s);
    return 0;
}

GradSegment seg(int off, int n, const float *part, int nsplit, int stride) {
    GradSegment g;
    g.off = off; g.n = n; g.part = part; g.nsplit = nsplit; g.stride = stride;
    return g;
}
GradSegments step_segments(const Workspace &w, const Plan &pl) {
    GradSegments s{};
    s.seg[0] = seg(kOffC1W, 500, w.c1w_part, pl.c1w_split, kC1PartStride);
    s.seg[1] = seg(kOffC1B, 20, w.c1w_part + 512, pl.c1w_split, kC1PartStride);
    s.seg[2] = seg(kOffC2W, 25000, w.c2w_part, pl.c2w_split, kC2PartStride);
    s.seg[3] = seg(kOffC2B, 50, w.c2w_part + 25024, pl.c2w_split, kC2PartStride);
    s.seg[4] = seg(kOffF2W, 5000, w.fc2w_part, pl.fc2w_split, kFc2PartStride);
    s.seg[5] = seg(kOffF2B, 10, w.fc2w_part + 5056, pl.fc2w_split, kFc2PartStride);
    s.count = 6;
    return s;
}
int launch_finalize(float *grads, const GradSegments &segs, const LossReduce &lr, cudaStream_t s) {
    const int n = kFlatElems - (kOffF2W - kOffF1W);
    CHECK_PDL(launch_pdl(grad_finalize_kernel, dim3(ceil_div(n, 256) + 1), dim3(256), 0, s, grads, segs, lr),
              "grad finalize", s);
    return 0;
}
// out[i] = sum of the partials (used by the per-layer entry points)
int reduce_to(float *out, int n, const float *part, int nsplit, int stride, cudaStream_t s) {
    reduce_partials_to_kernel<<<ceil_div(n, 256), 256, 0, s>>>(out, part, n, nsplit, stride);
    CHECK_LAUNCH("partial reduce", s);
    return 0;
}
int launch_sgd(float *p, float *g, float *m, int64_t n, float lr, float mu, float scale, int first, cudaStream_t s) {
    const int blocks = (int)((n + 1023) / 1024);
    sgd_momentum_kernel<<<blocks, 256, 0, s>>>(p, g, m, (int)n, -lr, mu, scale, first);
    CHECK_LAUNCH("sgd momentum", s);
    return 0;
}
int launch_sgd_fused(float *p, float *g, float *m, float lr, float mu, float scale, int first,
                     const GradSegments &segs, const LossReduce &lrd, cudaStream_t s) {
    CHECK_PDL(launch_pdl(sgd_fused_kernel, dim3(kFusedCovBlocks + kFusedFc1Blocks + 1), dim3(256), 0, s, p, g, m, -lr, mu,
                         scale, first, segs, lrd),
              "sgd momentum", s);
    return 0;
}

// forward (+ optional backward up to the partial gradients)
int run_fwd_bwd(const float *P, float *G, const float *x, const int64_t *target, int B, const Workspace &w,
                const Plan &pl, bool backward, float *logp_out, float *stats_out, float loss_scale, int want_correct,
                cudaStream_t s, const PeerStep *ps = nullptr, const RingRef *ring = nullptr,
                const LossReduce *early_finish = nullptr) {
    int rc;
    if ((rc = launch_conv1_fwd(x, P + kOffC1W, P + kOffC1B, w.p1, w.i1, B, s, ring))) return rc;
    // fc1 on the tensor cores reads hi / lo copies of its operands: their producers write them on the way
    const Fc1Ops ops = fc1ops_carve(w.fc1ops, B);
    const bool fc1_tc = B >= fc1_tc_min_batch(), c2_tc = w.w2s && conv2_tc_enabled(B);
    const bool p2_ready = fc1_tc && c2_tc, w1_ready = p2_ready && backward && pl.c2_presplit;
    const bool fc1_bwd_tc = fc1_tc && backward && pl.dg_split == 1;
    if ((rc = launch_conv2_fwd(w.p1, P + kOffC2W, P + kOffC2B, w.p2, w.i2, B, w.w2s, s,
                               backward && pl.c2_presplit ? w.wd : nullptr, p2_ready ? &ops : nullptr, P + kOffF1W)))
        return rc;
    int nsplit = 0;
    if ((rc = launch_fc1_fwd(w.p2, P + kOffF1W, P + kOffF1B, w.h, w.fc1_part, B, pl, ops, s, &nsplit, p2_ready, w1_ready)))
        return rc;
    HeadArgs a{};
    a.h_in = w.h; a.part = w.fc1_part; a.b1 = P + kOffF1B; a.h_out = w.h;
    a.w2 = P + kOffF2W; a.b2 = P + kOffF2B; a.target = target;
    a.logp = logp_out ? logp_out : w.logp;
    a.dz2 = backward ? w.dz2 : nullptr; a.dz1 = backward ? w.dz1 : nullptr;
    // fc1 backward on the tensor cores: dz1's hi / lo copies, and fc1.weight + fc1.bias gradients (contiguous in the
    // bucket) start from zero (see fc1_bwd_tc_kernel)
    const HeadTcExtra hx{ops.dz1_hi, ops.dz1_lo, reinterpret_cast<float4 *>(G + kOffF1W), (kOffF2W - kOffF1W) / 4};
    a.rows = w.rows; a.ticket = w.ticket; a.stats_out = stats_out;
    a.defer_reduce = backward ? 1 : 0;  // training: a later kernel of the step reduces the per-row losses (finish_step)
    a.B = B; a.nsplit = nsplit; a.want_correct = want_correct; a.inv_B = 1.0f / (float)B; a.loss_scale = loss_scale;
    if ((rc = launch_head(a, s, fc1_bwd_tc ? &hx : nullptr))) return rc;
    if (!backward) return 0;
    if ((rc = launch_fc_bwd(w.dz2, w.dz1, w.h, w.p2, P + kOffF1W, B, G + kOffF1W, G + kOffF1B, w.dp2_part,
                            w.fc2w_part, pl, ops, fc1_tc, fc1_bwd_tc, fc1_bwd_tc, s)))
        return rc;
    EarlySignal sig{};
    bool sig_sent = false;
    if (ps) {   // the conv2 backward's CTA 0 tells the peers that fc1's gradient is staged (fc backward has completed)
        for (int r = 0; r < ps->pw.world; ++r)
            sig.dst[r] = r == ps->pw.rank ? nullptr
                                          : reinterpret_cast<unsigned *>(ps->pw.base[r]) + kFlagFc1Ready + ps->pw.rank;
        sig.epoch_src = reinterpret_cast<const unsigned *>(ps->pw.base[ps->pw.rank]) + kFlagEpochA;
        sig.n = ps->pw.world;
    }
    if ((rc = launch_conv2_bwd(w.dp2_part, pl.dg_split, (size_t)B * kP2Elems, w.p2, w.i2, w.p1, P + kOffC2W, B, w.dp1,
                               w.c2w_part, w.wd, pl, s, pl.c2_presplit, ps ? &sig : nullptr, &sig_sent)))
        return rc;
    return launch_conv1_bwd(w.dp1, w.p1, w.i1, x, B, w.c1w_part, pl, s, ps, sig_sent, early_finish);
}

}  // namespace

extern "C" {

int64_t b200mnist_flat_elems(void) { return kFlatElems; }
int64_t b200mnist_tensor_offset(int i) { return (i < 0 || i >= kNumTensors) ? -1 : kTensorOff[i]; }
int64_t b200mnist_tensor_elems(int i) { return (i < 0 || i >= kNumTensors) ? -1 : kTensorElems[i]; }
const char *b200mnist_last_error(void) { return g_last_error.c_str(); }
int b200mnist_version(void) { return 100; }

int64_t b200mnist_workspace_bytes(int max_batch) {
    if (max_batch < 1 || max_batch > 65536) return -1;
    return (int64_t)carve(nullptr, max_batch).bytes;
}

int64_t b200mnist_workspace_offset(int max_batch, int which) {
    if (max_batch < 1 || max_batch > 65536) return -1;
    const Workspace w = carve(reinterpret_cast<void *>(uintptr_t(256)), max_batch);   // non-null base: offsets + 256
    const void *p[] = {w.p1, w.i1, w.p2, w.i2, w.h, w.logp};
    if (which < 0 || which >= 6) return -1;
    return (int64_t)(reinterpret_cast<uintptr_t>(p[which]) - 256);
}

int b200mnist_forward_backward(const float *params, float *grads, const float *x, const int64_t *target, int B,
                               void *workspace, int64_t workspace_bytes, float *loss_out, float *logp_out,
                               void *stream) {
    int rc = check_step_args(params, x, target, B, workspace, workspace_bytes);
    if (rc) return rc;
    if (!grads || !loss_out || !aligned16(grads)) return fail_arg(5, "grads/loss_out null or grads unaligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const Plan pl = make_plan(B);
    const Workspace w = carve(workspace, B);
    if ((rc = run_fwd_bwd(params, grads, x, target, B, w, pl, true, logp_out, loss_out, 1.0f / (float)B, 0, s)))
        return rc;
    LossReduce lrd{w.rows, loss_out, B, 1.0f / (float)B};
    return launch_finalize(grads, step_segments(w, pl), lrd, s);
}

}  // extern "C"
namespace {
int train_step_impl(float *params, float *grads, float *momentum, const float *x, const int64_t *target, int B,
                    void *workspace, int64_t workspace_bytes, float lr, float mu, int first_step, float *loss_out,
                    const b200mnist_ring_t *ring, void *stream) {
    int rc = check_step_args(params, x, target, B, workspace, workspace_bytes);
    if (rc) return rc;
    if (!grads || !momentum || !loss_out || !aligned16(grads) || !aligned16(momentum))
        return fail_arg(5, "grads/momentum/loss_out null or unaligned");
    RingRef rr{};
    if (ring && (rc = make_ring(ring, x, target, B, &rr))) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const Plan pl = make_plan(B);
    const Workspace w = carve(workspace, B);
    // a host-fed ring is closed by the conv1 weight-gradient launch (conv1_wgrad_finish_kernel); otherwise the SGD
    // kernel's last block does it
    const LossReduce lrd = ring_loss_reduce(w.rows, loss_out, B, ring);
    const bool early = ring && ring->report_host;
    if ((rc = run_fwd_bwd(params, grads, x, target, B, w, pl, true, nullptr, loss_out, 1.0f / (float)B, 0, s, nullptr,
                          ring ? &rr : nullptr, early ? &lrd : nullptr)))
        return rc;
    const GradSegments segs = step_segments(w, pl);
    return launch_sgd_fused(params, grads, momentum, lr, mu, 1.0f, first_step, segs, early ? LossReduce{} : lrd, s);
}
}  // namespace
extern "C" {

int b200mnist_train_step(float *params, float *grads, float *momentum, const float *x, const int64_t *target, int B,
                         void *workspace, int64_t workspace_bytes, float lr, float mu, int first_step,
                         float *loss_out, void *stream) {
    return train_step_impl(params, grads, momentum, x, target, B, workspace, workspace_bytes, lr, mu, first_step, loss_out,
                           nullptr, stream);
}

int b200mnist_train_step_ring(float *params, float *grads, float *momentum, const b200mnist_ring_t *ring_host, float *x,
                              int64_t *target, int B, void *workspace, int64_t workspace_bytes, float lr, float mu,
                              int first_step, float *loss_out, void *stream) {
    if (!ring_host) return fail_arg(1, "null ring description");
    return train_step_impl(params, grads, momentum, x, target, B, workspace, workspace_bytes, lr, mu, first_step, loss_out,
                           ring_host, stream);
}

int b200mnist_forward_eval(const float *params, const float *x, const int64_t *target, int B, void *workspace,
                           int64_t workspace_bytes, float *logp_out, float *stats_out, void *stream) {
    int rc = check_step_args(params, x, target, B, workspace, workspace_bytes);
    if (rc) return rc;
    if (!stats_out) return fail_arg(5, "stats_out is null");
    const Plan pl = make_plan(B);
    const Workspace w = carve(workspace, B);
    return run_fwd_bwd(params, nullptr, x, target, B, w, pl, false, logp_out, stats_out, 1.0f, 1,
                       static_cast<cudaStream_t>(stream));
}

int b200mnist_sgd_momentum(float *params, const float *grads, float *momentum, int64_t n, float lr, float mu,
                           float grad_scale, int first_step, void *stream) {
    if (!params || !grads || !momentum) return fail_arg(1, "null pointer argument");
    if (n < 0 || n > (int64_t)1 << 30) return fail_arg(2, "n out of range");
    if (n == 0) return 0;
    if (!aligned16(params) || !aligned16(grads) || !aligned16(momentum))
        return fail_arg(3, "params/grads/momentum must be 16-byte aligned");
    return launch_sgd(params, const_cast<float *>(grads), momentum, n, lr, mu, grad_scale, first_step,
                      static_cast<cudaStream_t>(stream));
}

/* ---- peer windows: set-up calls (these allocate / map; they are not stream work) ---- */
int64_t b200mnist_peer_window_bytes(void) { return (int64_t)kPeerWindowBytes; }

int b200mnist_peer_window_create(void **window_out_host, void *ipc_handle_out_host) {
    if (!window_out_host || !ipc_handle_out_host) return fail_arg(1, "null pointer argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == B200MNIST_IPC_HANDLE_BYTES, "IPC handle size");
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, kPeerWindowBytes);
    if (e != cudaSuccess) return fail_cuda(e, "cudaMalloc(peer window)");
    e = cudaMemset(p, 0, kPeerWindowBytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return fail_cuda(e, "cudaIpcGetMemHandle(peer window)");
    }
    memcpy(ipc_handle_out_host, &h, sizeof(h));
    *window_out_host = p;
    return 0;
}
int b200mnist_peer_window_open(const void *ipc_handle_host, void **window_out_host) {
    if (!ipc_handle_host || !window_out_host) return fail_arg(1, "null pointer argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle_host, sizeof(h));
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail_cuda(e, "cudaIpcOpenMemHandle(peer window)");
    }
    *window_out_host = p;
    return 0;
}
int b200mnist_peer_window_close(void *window) {
    if (!window) return 0;
    cudaError_t e = cudaIpcCloseMemHandle(window);
    if (e != cudaSuccess) return fail_cuda(e, "cudaIpcCloseMemHandle");
    return 0;
}
int b200mnist_peer_window_destroy(void *window) {
    if (!window) return 0;
    cudaError_t e = cudaFree(window);
    if (e != cudaSuccess) return fail_cuda(e, "cudaFree(peer window)");
    return 0;
}
int b200mnist_peer_window_probe(void *const *windows_host, int world_size, uint32_t *out, void *stream) {
    if (!windows_host || !out) return fail_arg(1, "null pointer argument");
    if (world_size < 1 || world_size > kMaxPeers) return fail_arg(2, "world_size must be in [1, 8]");
    PeerWin pw{};
    for (int r = 0; r < world_size; ++r) {
        if (!windows_host[r]) return fail_arg(1, "null peer window");
        pw.base[r] = static_cast<char *>(windows_host[r]);
    }
    pw.world = world_size;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    peer_probe_kernel<<<1, 32, 0, s>>>(pw, out);
    CHECK_LAUNCH("peer window probe", s);
    return 0;
}

}  // extern "C"
namespace {
int peer_train_step_impl(void *const *windows_host, int rank, int world_size, int parity, float *momentum,
                         const float *x, const int64_t *target, int B, void *workspace, int64_t workspace_bytes,
                         float lr, float mu, int first_step, float *loss_out, const b200mnist_ring_t *ring, void *stream) {
    if (!windows_host || !momentum || !loss_out) return fail_arg(1, "null pointer argument");
    if ((world_size != 2 && world_size != 4 && world_size != 8) || rank < 0 || rank >= world_size)
        return fail_arg(2, "world_size must be 2, 4 or 8 and rank in [0, world_size)");
    if (parity != 0 && parity != 1) return fail_arg(2, "parity must be 0 or 1");
    if (!aligned16(momentum)) return fail_arg(3, "momentum must be 16-byte aligned");
    PeerStep ps{};
    for (int r = 0; r < world_size; ++r) {
        if (!windows_host[r] || (reinterpret_cast<uintptr_t>(windows_host[r]) & 255u))
            return fail_arg(3, "peer windows must be non-null and 256-byte aligned");
        ps.pw.base[r] = static_cast<char *>(windows_host[r]);
    }
    ps.pw.rank = rank;
    ps.pw.world = world_size;
    ps.m = momentum;
    ps.parity = parity;
    ps.neg_lr = -lr;
    ps.mu = mu;
    ps.scale = 1.0f / (float)world_size;
    ps.first = first_step;
    char *own = ps.pw.base[rank];
    const float *params = reinterpret_cast<const float *>(own + kPeerFlagBytes);
    float *staging = reinterpret_cast<float *>(own + kPeerFlagBytes + (size_t)(2 + parity) * kBucketBytes);
    int rc = check_step_args(params, x, target, B, workspace, workspace_bytes);
    if (rc) return rc;
    RingRef rr{};
    if (ring && (rc = make_ring(ring, x, target, B, &rr))) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const Plan pl = make_plan(B);
    const Workspace w = carve(workspace, B);
    // forward, loss head, fc backward (fc1's gradient lands in this step's staging bucket), conv2 backward, then conv1's
    // weight gradient with the fc1 slice exchange in the same launch (which also closes a host-fed ring, see
    // conv1_wgrad_finish_kernel)
    const LossReduce lrd = ring_loss_reduce(w.rows, loss_out, B, ring);
    const bool early = ring && ring->report_host;
    if ((rc = run_fwd_bwd(params, staging, x, target, B, w, pl, true, nullptr, loss_out, 1.0f / (float)B, 0, s, &ps,
                          ring ? &rr : nullptr, early ? &lrd : nullptr)))
        return rc;
    const GradSegments segs = step_segments(w, pl);
    const LossReduce tail_lrd = early ? LossReduce{} : lrd;
    switch (world_size) {
    case 2: return launch_peer_tail<2>(ps, segs, tail_lrd, s);
    case 4: return launch_peer_tail<4>(ps, segs, tail_lrd, s);
    default: return launch_peer_tail<8>(ps, segs, tail_lrd, s);
    }
}
}  // namespace
extern "C" {

int b200mnist_peer_train_step(void *const *windows_host, int rank, int world_size, int parity, float *momentum,
                              const float *x, const int64_t *target, int B, void *workspace, int64_t workspace_bytes,
                              float lr, float mu, int first_step, float *loss_out, void *stream) {
    return peer_train_step_impl(windows_host, rank, world_size, parity, momentum, x, target, B, workspace, workspace_bytes,
                                lr, mu, first_step, loss_out, nullptr, stream);
}

int b200mnist_peer_train_step_ring(void *const *windows_host, int rank, int world_size, int parity, float *momentum,
                                   const b200mnist_ring_t *ring_host, float *x, int64_t *target, int B, void *workspace,
                                   int64_t workspace_bytes, float lr, float mu, int first_step, float *loss_out,
                                   void *stream) {
    if (!ring_host) return fail_arg(1, "null ring description");
    return peer_train_step_impl(windows_host, rank, world_size, parity, momentum, x, target, B, workspace, workspace_bytes,
                                lr, mu, first_step, loss_out, ring_host, stream);
}

int b200mnist_conv1_relu_pool_fwd(const float *x, const float *w, const float *bias, float *out, uint8_t *idx, int B,
                                  void *stream) {
    if (!x || !w || !bias || !out || !idx) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(x)) return fail_arg(3, "x must be 16-byte aligned");
    return launch_conv1_fwd(x, w, bias, out, idx, B, static_cast<cudaStream_t>(stream));
}

int b200mnist_conv2_relu_pool_fwd(const float *in, const float *w, const float *bias, float *out, uint8_t *idx, int B,
                                  void *stream) {
    if (!in || !w || !bias || !out || !idx) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(in)) return fail_arg(3, "in must be 16-byte aligned");
    return launch_conv2_fwd(in, w, bias, out, idx, B, nullptr, static_cast<cudaStream_t>(stream));
}

int64_t b200mnist_conv2_fwd_tc_scratch_bytes(void) { return (int64_t)tc::kWnScratchFloats * 4; }

int b200mnist_conv2_relu_pool_fwd_tc(const float *in, const float *w, const float *bias, float *out, uint8_t *idx,
                                     int B, void *scratch, void *stream) {
    if (!in || !w || !bias || !out || !idx || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(in) || !aligned16(scratch)) return fail_arg(3, "in/scratch must be 16-byte aligned");
    return launch_conv2_fwd_tc(in, w, bias, out, idx, B, static_cast<float *>(scratch), static_cast<cudaStream_t>(stream));
}

// split-K partials (SIMT path) or the hi / lo operand copies (tensor-core path): whichever is larger
int64_t b200mnist_fc1_scratch_bytes(int B) {
    if (B < 1) return -1;
    const int64_t part = (int64_t)10 * B * 512 * 4, ops = (int64_t)fc1ops_bytes(B);
    return part > ops ? part : ops;
}

int b200mnist_fc1_relu_fwd(const float *in, const float *w, const float *bias, float *out, int B, void *scratch,
                           void *stream) {
    if (!in || !w || !bias || !out || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(in) || !aligned16(w) || !aligned16(scratch)) return fail_arg(3, "in/w/scratch must be 16-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Plan pl = make_plan(B);
    pl.fc1_split = 1;  // standalone call: finish in the GEMM epilogue, `scratch` stays unused
    pl.fc1_kchunk = 800;
    int nsplit = 0;
    if (reinterpret_cast<uintptr_t>(scratch) & 255u) return fail_arg(3, "scratch must be 256-byte aligned");
    return launch_fc1_fwd(in, w, bias, out, static_cast<float *>(scratch), B, pl, fc1ops_carve(scratch, B), s, &nsplit);
}

int b200mnist_fc1_relu_fwd_tc(const float *in, const float *w, const float *bias, float *out, int B, void *scratch,
                              void *stream) {
    if (!in || !w || !bias || !out || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(in) || !aligned16(w) || (reinterpret_cast<uintptr_t>(scratch) & 255u))
        return fail_arg(3, "in/w must be 16-byte aligned and scratch 256-byte aligned");
    return launch_fc1_fwd_tc(in, w, bias, out, B, fc1ops_carve(scratch, B), static_cast<cudaStream_t>(stream));
}

int b200mnist_fc2_logsoftmax_nll(const float *h, const float *w, const float *bias, const int64_t *target, int B,
                                 float *logp, float *loss_out, float *dz2, float *dz1, void *scratch, void *stream) {
    if (!h || !w || !bias || !target || !loss_out || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    HeadArgs a{};
    a.h_in = h; a.w2 = w; a.b2 = bias; a.target = target; a.logp = logp; a.dz2 = dz2; a.dz1 = dz1;
    a.ticket = static_cast<unsigned *>(scratch);
    a.rows = static_cast<float *>(scratch) + 64;
    a.stats_out = loss_out;
    a.B = B; a.nsplit = 0; a.want_correct = 0; a.inv_B = 1.0f / (float)B; a.loss_scale = 1.0f / (float)B;
    return launch_head(a, static_cast<cudaStream_t>(stream));
}

/* ---- allocation-free variant of b200mnist_fc2_logsoftmax_nll ------------------------
 * Some hosts that embed this library cannot hand it a scratch buffer: a plugin that is
 * called from inside another runtime's callback is not allowed to call cudaMalloc there
 * (cudaMalloc may synchronise the whole device, which stalls the runtime's other streams),
 * and a process whose device memory has been handed in full to someone else's pool has
 * nothing left to allocate from.  For them the (B + 64)-float scratch of the head kernel
 * lives in a fixed array inside the library image instead.  Every GPU that loads the library
 * carries these 16,640 bytes until the process exits, whether the entry point is used or
 * not, which is why the limit is kept small (B <= B200MNIST_STATIC_MAX_BATCH = 4096). */
constexpr int kStaticMaxBatch = 4096;
static __device__ __align__(256) float g_head_scratch[64 + kStaticMaxBatch];

namespace {
constexpr int kMaxDevices = 64;
float *g_head_scratch_addr[kMaxDevices] = {};  // device address of g_head_scratch, per device ordinal
}  // namespace

int b200mnist_fc2_logsoftmax_nll_static(const float *h, const float *w, const float *bias, const int64_t *target,
                                        int B, float *logp, float *loss_out, float *dz2, float *dz1, void *stream) {
    if (!h || !w || !bias || !target || !loss_out) return fail_arg(1, "null pointer argument");
    if (B < 1) return fail_arg(2, "batch size must be at least 1");
    if (B > kStaticMaxBatch)
        return fail_arg(6, "batch size exceeds B200MNIST_STATIC_MAX_BATCH (4096): the built-in scratch array is "
                           "fixed-size; use b200mnist_fc2_logsoftmax_nll with a caller-provided scratch buffer");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return fail_arg(7, "device ordinal out of range for the static scratch table");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    e = cudaStreamIsCapturing(s, &cap);
    if (e != cudaSuccess) return fail_cuda(e, "cudaStreamIsCapturing");
    if (!g_head_scratch_addr[dev] || cap == cudaStreamCaptureStatusNone) {
        // Set-up for this device, with this device current.  In host code the name
        // g_head_scratch is only a handle for the symbol, not an address a kernel can use:
        // the real per-device address has to be asked for.  The answer is remembered per
        // device ordinal for calls made while the stream is being captured; outside a
        // capture it is asked for again on every call (cheap), so an address remembered
        // before a cudaDeviceReset / context re-creation is not trusted afterwards.
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_head_scratch);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_head_scratch)");
        g_head_scratch_addr[dev] = static_cast<float *>(p);
    }
    float *scratch = g_head_scratch_addr[dev];
    // The array keeps whatever the previous call left in it.  Only the ticket word at its head
    // has to start from zero (the per-row slots are fully rewritten before they are read).
    e = cudaMemsetAsync(scratch, 0, 256, s);
    if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(static scratch)");
    HeadArgs a{};
    a.h_in = h; a.w2 = w; a.b2 = bias; a.target = target; a.logp = logp; a.dz2 = dz2; a.dz1 = dz1;
    a.ticket = reinterpret_cast<unsigned *>(scratch);
    a.rows = scratch + 64;
    a.stats_out = loss_out;
    a.B = B; a.nsplit = 0; a.want_correct = 0; a.inv_B = 1.0f / (float)B; a.loss_scale = 1.0f / (float)B;
    return launch_head(a, s);
}

int b200mnist_head_bwd(const float *logp, const float *dlogp, const float *h, const float *w, int B, float *dz2,
                       float *dz1, void *stream) {
    if (!logp || !dlogp || !h || !w || !dz2 || !dz1) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    head_bwd_kernel<<<B, 128, 0, s>>>(logp, dlogp, h, w, dz2, dz1);
    CHECK_LAUNCH("head backward", s);
    return 0;
}

int64_t b200mnist_fc_bwd_scratch_bytes(int B) {
    if (B < 1) return -1;   // [dp2 split-K partials | fc2 weight-gradient partials | hi / lo operand copies]
    return (int64_t)align256((size_t)8 * B * kP2Elems * 4) + (int64_t)align256((size_t)32 * kFc2PartStride * 4) +
           (int64_t)fc1ops_bytes(B);
}

int b200mnist_fc_bwd(const float *dz2, const float *dz1, const float *h, const float *p2, const float *w1, int B,
                     float *gw2, float *gb2, float *gw1, float *gb1, float *dp2, void *scratch, void *stream) {
    if (!dz2 || !dz1 || !h || !p2 || !w1 || !gw2 || !gb2 || !gw1 || !gb1 || !dp2 || !scratch)
        return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(dz1) || !aligned16(p2) || !aligned16(w1) || !aligned16(scratch))
        return fail_arg(3, "dz1/p2/w1/scratch must be 16-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Plan pl = make_plan(B);
    pl.dg_split = 1;  // standalone call: dp2 is written whole
    pl.dg_kchunk = 512;
    if (reinterpret_cast<uintptr_t>(scratch) & 255u) return fail_arg(3, "scratch must be 256-byte aligned");
    char *sc = static_cast<char *>(scratch) + align256((size_t)8 * B * kP2Elems * 4);
    float *fc2w_part = reinterpret_cast<float *>(sc);
    const Fc1Ops ops = fc1ops_carve(sc + align256((size_t)32 * kFc2PartStride * 4), B);
    int rc = launch_fc_bwd(dz2, dz1, h, p2, w1, B, gw1, gb1, dp2, fc2w_part, pl, ops, false, false, false, s);
    if (rc) return rc;
    if ((rc = reduce_to(gw2, 5000, fc2w_part, pl.fc2w_split, kFc2PartStride, s))) return rc;
    return reduce_to(gb2, 10, fc2w_part + 5056, pl.fc2w_split, kFc2PartStride, s);
}

int64_t b200mnist_conv2_bwd_scratch_bytes(int B) {
    return B < 1 ? -1 : (int64_t)32 * kC2PartStride * 4 + (int64_t)tc::kDgScratchFloats * 4;
}

int b200mnist_conv2_bwd(const float *dp2, const float *p2, const uint8_t *idx2, const float *p1, const float *w, int B,
                        float *gw, float *gb, float *dp1, void *scratch, void *stream) {
    if (!dp2 || !p2 || !idx2 || !p1 || !w || !gw || !gb || !dp1 || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(p1) || !aligned16(w) || !aligned16(scratch) || !aligned16(dp2) || !aligned16(p2) ||
        (reinterpret_cast<uintptr_t>(idx2) & 3u))
        return fail_arg(3, "dp2/p2/p1/w/scratch must be 16-byte aligned and idx2 4-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const Plan pl = make_plan(B);
    float *part = static_cast<float *>(scratch);
    int rc = launch_conv2_bwd(dp2, 1, 0, p2, idx2, p1, w, B, dp1, part, part + (size_t)32 * kC2PartStride, pl, s);
    if (rc) return rc;
    return reduce_to(gw, 25000, part, pl.c2w_split, kC2PartStride, s) ||
           reduce_to(gb, 50, part + 25024, pl.c2w_split, kC2PartStride, s);
}

int64_t b200mnist_conv1_bwd_scratch_bytes(int B) { return B < 1 ? -1 : (int64_t)kMaxGradSplit * kC1PartStride * 4; }

int b200mnist_conv1_bwd(const float *dp1, const float *p1, const uint8_t *idx1, const float *x, int B, float *gw,
                        float *gb, void *scratch, void *stream) {
    if (!dp1 || !p1 || !idx1 || !x || !gw || !gb || !scratch) return fail_arg(1, "null pointer argument");
    if (B < 1 || B > 65536) return fail_arg(2, "batch size must be in [1, 65536]");
    if (!aligned16(x)) return fail_arg(3, "x must be 16-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const Plan pl = make_plan(B);
    float *part = static_cast<float *>(scratch);
    int rc = launch_conv1_bwd(dp1, p1, idx1, x, B, part, pl, s);
    if (rc) return rc;
    return reduce_to(gw, 500, part, pl.c1w_split, kC1PartStride, s) ||
           reduce_to(gb, 20, part + 512, pl.c1w_split, kC1PartStride, s);
}

int b200mnist_gather_normalize_u8(const uint8_t *images, const int64_t *labels, const int64_t *indices, int n,
                                  int n_images, float mean, float stdv, float *data, int64_t *target, void *stream) {
    if (n == 0) return 0;
    if (!images || !labels || !indices || !data || !target) return fail_arg(1, "null pointer argument");
    if (n < 0 || n > 65536 || n_images < 1) return fail_arg(2, "n must be in [0, 65536] and n_images >= 1");
    if ((reinterpret_cast<uintptr_t>(images) & 3u) || !aligned16(data))
        return fail_arg(3, "images must be 4-byte aligned and data 16-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gather_normalize_u8_kernel<<<n, 196, 0, s>>>(images, reinterpret_cast<const long long *>(labels),
                                                reinterpret_cast<const long long *>(indices), n_images, mean, stdv, data,
                                                reinterpret_cast<long long *>(target));
    CHECK_LAUNCH("gather + normalize (input pipeline)", s);
    return 0;
}

int b200mnist_square(const float *in, float *out, int64_t n, void *stream) {
    if (n < 0) return fail_arg(2, "n out of range");
    if (n == 0) return 0;
    if (!in || !out) return fail_arg(1, "null pointer argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    square_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, out, (long long)n);
    CHECK_LAUNCH("square", s);
    return 0;
}

/* ---- host-side plumbing for the double-buffered input pipeline (mnist.py:38, data.to(device)) ----
 * One call = the four stream operations of StepPipeline.submit, without Python in between:
 *   copy stream waits until the slot's previous step has finished (ev_slot_free), H2D copy of the record,
 *   ev_copied recorded on the copy stream, compute stream waits for ev_copied. */
int b200mnist_stage_inputs(void *dst_device, const void *src_host_pinned, int64_t bytes, void *copy_stream,
                           void *compute_stream, void *ev_slot_free, void *ev_copied) {
    if (!dst_device || !src_host_pinned || !ev_slot_free || !ev_copied || bytes <= 0)
        return fail_arg(1, "null pointer argument");
    cudaStream_t cs = static_cast<cudaStream_t>(copy_stream), ms = static_cast<cudaStream_t>(compute_stream);
    cudaError_t e = cudaStreamWaitEvent(cs, static_cast<cudaEvent_t>(ev_slot_free), 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dst_device, src_host_pinned, (size_t)bytes, cudaMemcpyHostToDevice, cs);
    if (e == cudaSuccess) e = cudaEventRecord(static_cast<cudaEvent_t>(ev_copied), cs);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ms, static_cast<cudaEvent_t>(ev_copied), 0);
    if (e != cudaSuccess) return fail_cuda(e, "stage_inputs");
    return 0;
}
int b200mnist_stage_record(void *dst_device, const void *src, int64_t bytes, void *copy_stream, uint32_t *ready_word,
                           uint32_t seq) {
    if (!dst_device || !src || bytes <= 0) return fail_arg(1, "null pointer argument");
    cudaStream_t cs = static_cast<cudaStream_t>(copy_stream);
    cudaError_t e = cudaMemcpyAsync(dst_device, src, (size_t)bytes, cudaMemcpyDefault, cs);
    if (e != cudaSuccess) return fail_cuda(e, "stage_record: cudaMemcpyAsync");
    if (!ready_word) return 0;
    typedef CUresult (*WriteValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
    static WriteValue32Fn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return fail_arg(8, "cuStreamWriteValue32 is unavailable (driver too old?)");
        fn = reinterpret_cast<WriteValue32Fn>(p);
    }
    const CUresult r = fn(reinterpret_cast<CUstream>(cs), reinterpret_cast<CUdeviceptr>(ready_word), seq, 0);
    if (r != CUDA_SUCCESS) return fail_arg(8, "cuStreamWriteValue32 failed");
    return 0;
}
int b200mnist_event_record(void *event, void *stream) {
    cudaError_t e = cudaEventRecord(static_cast<cudaEvent_t>(event), static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail_cuda(e, "cudaEventRecord");
    return 0;
}

/* ---- instrumentation (no counterpart in the reference) ---- */
long long b200mnist_launch_count(void) { return g_launches; }

int b200mnist_profile_begin(void *stream, int delay_us) {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (delay_us > 0) delay_kernel<<<1, 1, 0, s>>>((unsigned long long)delay_us * 1000ull);
    if (!g_prof.created) {
        for (int i = 0; i <= kProfMax; ++i) {
            cudaError_t e = cudaEventCreate(&g_prof.ev[i]);
            if (e != cudaSuccess) return fail_cuda(e, "cudaEventCreate");
        }
        g_prof.created = true;
    }
    g_prof.count = 0;
    g_prof.enabled = true;
    cudaError_t e = cudaEventRecord(g_prof.ev[0], s);
    if (e != cudaSuccess) return fail_cuda(e, "cudaEventRecord");
    return 0;
}

int b200mnist_delay(void *stream, int delay_us) {
    if (delay_us <= 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    delay_kernel<<<1, 1, 0, s>>>((unsigned long long)delay_us * 1000ull);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "delay kernel");
    return 0;
}

int64_t b200mnist_fma_peak(float *out, int blocks, int iters, void *stream) {
    if (!out || blocks < 1 || iters < 1) return fail_arg(1, "out null or blocks/iters < 1");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    fma_peak_kernel<<<blocks, 256, 0, s>>>(out, iters, 1.0000001f, 0.9999999f);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "fma peak kernel");
    return (int64_t)blocks * 256 * 64 * iters;
}

int b200mnist_profile_end(float *ms_out_host, const char **names_out_host, int cap, void *stream) {
    g_prof.enabled = false;
    cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail_cuda(e, "cudaStreamSynchronize");
    int n = g_prof.count < cap ? g_prof.count : cap;
    for (int i = 0; i < n; ++i) {
        e = cudaEventElapsedTime(&ms_out_host[i], g_prof.ev[i], g_prof.ev[i + 1]);
        if (e != cudaSuccess) return fail_cuda(e, "cudaEventElapsedTime");
        names_out_host[i] = g_prof.names[i];
    }
    return n;
}

}  // extern "C"
