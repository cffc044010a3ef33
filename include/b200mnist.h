/*
 * b200mnist.h -- C ABI of libb200mnist.so: the examples/mnist training step and the
 * examples/smoke-dist worker op as hand-written sm_100a CUDA kernels.
 *
 * The reference (jiaqianjing/pytorch-operator) has no native code: its payloads reach
 * native kernels only through PyTorch.  Each entry point below therefore cites the
 * line of the reference PAYLOAD whose library call it replaces
 * (examples/mnist/mnist.py, examples/smoke-dist/dist_sendrecv.py under /root/reference).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; all tensors are
 *     dense fp32 in the reference's layouts (NCHW activations, [out,in,kh,kw] conv
 *     weights, [out,in] linear weights), targets are int64 as torch hands them over;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); every
 *     call only enqueues work on it -- no allocation, no synchronisation, no retained
 *     pointers -- so calls are legal inside CUDA-graph stream capture;
 *   - return value: 0 on success, negative on failure (-cudaError_t for CUDA errors,
 *     -10000-k for argument error k); b200mnist_last_error() gives the text of the last
 *     failure on the calling thread.  Nothing throws across this boundary;
 *   - the caller owns all memory.  Workspaces are caller-allocated, sized by
 *     b200mnist_workspace_bytes(), 256-byte aligned, and must be zero-filled ONCE before
 *     their first use (they hold self-resetting counters);
 *   - one host thread per device context at a time (the reference runs one Python
 *     thread per process/pod).
 */
#ifndef B200MNIST_H
#define B200MNIST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- flat parameter bucket ------------------------------------------------------
 * The eight tensors of Net (mnist.py:17-23) live in ONE flat fp32 buffer so that
 * the DDP gradient exchange (mnist.py:135-138) is a single collective and the SGD
 * update (mnist.py:140,43) a single kernel.  Each tensor starts on a 256-byte
 * boundary; pad elements stay zero.  Index: 0 conv1.weight 1 conv1.bias 2 conv2.weight
 * 3 conv2.bias 4 fc1.weight 5 fc1.bias 6 fc2.weight 7 fc2.bias. */
#define B200MNIST_NUM_TENSORS 8
#define B200MNIST_PARAM_ELEMS 431080 /* real parameters */
#define B200MNIST_FLAT_ELEMS 431296  /* with alignment padding */
int64_t b200mnist_flat_elems(void);
int64_t b200mnist_tensor_offset(int index); /* element offset in the flat buffer, -1 if bad index */
int64_t b200mnist_tensor_elems(int index);  /* element count, -1 if bad index */

const char *b200mnist_last_error(void);
int b200mnist_version(void);

/* Bytes of workspace the step entry points need for batches up to `max_batch`. */
int64_t b200mnist_workspace_bytes(int max_batch);

/* Where a step entry point leaves its saved activations inside the workspace (byte offset from its start, for a
 * workspace sized for `max_batch`; -1 on a bad argument): which = 0 pooled conv1 output p1 [B,20,12,12] fp32,
 * 1 its arg-max slots [B,20,12,12] uint8, 2 pooled conv2 output p2 [B,50,4,4] fp32 (= the [B,800] view of
 * mnist.py:30), 3 its arg-max slots uint8, 4 fc1 output h [B,500] fp32, 5 log-probabilities [B,10] fp32.
 * What autograd would have saved for mnist.py:42; the parity tests read them to teacher-force the oracle. */
int64_t b200mnist_workspace_offset(int max_batch, int which);

/* ---- whole-step entry points ----------------------------------------------------- */

/* Replaces, for one batch, mnist.py:39-42: zero_grad(); output = model(data);
 * loss = F.nll_loss(output, target); loss.backward().
 *   params, grads : flat buckets (B200MNIST_FLAT_ELEMS floats); grads is overwritten
 *   x [B,1,28,28], target int64[B]
 *   loss_out : 1 float, the mean NLL of this batch
 *   logp_out : optional [B,10] log-probabilities (may be NULL)
 * An out-of-range target makes loss_out NaN (torch raises there). */
int b200mnist_forward_backward(const float *params, float *grads, const float *x, const int64_t *target, int B,
                               void *workspace, int64_t workspace_bytes, float *loss_out, float *logp_out,
                               void *stream);

/* Same as b200mnist_forward_backward followed by b200mnist_sgd_momentum in one call,
 * with the gradient reduction fused into the update (single-process training,
 * mnist.py:39-43 when WORLD_SIZE == 1). */
int b200mnist_train_step(float *params, float *grads, float *momentum, const float *x, const int64_t *target, int B,
                         void *workspace, int64_t workspace_bytes, float lr, float mu, int first_step,
                         float *loss_out, void *stream);

/* "Resident input": the batches of many steps already sit in device memory (a ring of slots) and the step picks its
 * own -- for training loops that replay one CUDA graph per step with no copy and no host-side pointer in between
 * (replaces the DataLoader + `data.to(device)` of mnist.py:37-38 when the data fits in HBM).
 *   slot s : images at x + s * x_stride floats ([B,1,28,28] fp32, 16-byte aligned), targets at target + s * t_stride
 *   state  : two uint32 in DEVICE memory (8-byte aligned, zero-filled once): [0] the cursor -- the slot the next step
 *            reads -- and [1] the number of steps that have consumed the ring.  conv1's forward kernel reads them, one of
 *            the step's last two kernels sets them to (cursor + 1) % slots and steps + 1.  The host may rewrite the cursor
 *            between steps (stream-ordered).
 * b200mnist_train_step_ring is b200mnist_train_step reading its batch from the ring: `x` and `target` are now OUTPUT
 * buffers of the step ([B,1,28,28] fp32 and int64[B]): the first kernel leaves the slot's batch there for the rest
 * of the step (and for the caller to inspect).  ring_host is read during the call only.
 * Host-fed ring (optional members, NULL for a ring that is simply resident): a copy stream refills the slots while
 * the steps run, and the compute stream carries nothing but the step launches --
 *   landed      : DEVICE uint32: sequence number (1, 2, ...) of the newest batch that has arrived in the ring, written
 *                 BEHIND each copy in copy-stream order (b200mnist_stage_record).  Step n waits on the device until
 *                 landed >= n.
 *   report_host : uint64[slots] in PINNED HOST memory: step n stores (n << 32 | bits of the loss) into
 *                 report_host[slot of step n] as one 8-byte word: when the high half reads n, the loss of step n is
 *                 final (the low half) and the slot may be refilled.  The store leaves with the step's second-to-last
 *                 kernel (the parameter update completes right behind it, in stream order), and the cursor / steps
 *                 pair of `state` advances there too. */
typedef struct b200mnist_ring {
    const float *x;
    const int64_t *target;
    int64_t x_stride, t_stride;
    uint32_t slots;
    uint32_t *state;
    const uint32_t *landed;
    uint64_t *report_host;
} b200mnist_ring_t;
int b200mnist_train_step_ring(float *params, float *grads, float *momentum, const b200mnist_ring_t *ring_host, float *x,
                              int64_t *target, int B, void *workspace, int64_t workspace_bytes, float lr, float mu,
                              int first_step, float *loss_out, void *stream);

/* Replaces model(data) under torch.no_grad() plus the per-batch reductions of test(),
 * mnist.py:56-61: logp_out [B,10] (may be NULL); stats_out[0] = sum of NLL
 * (F.nll_loss(reduction='sum')), stats_out[1] = number of rows whose argmax equals target. */
int b200mnist_forward_eval(const float *params, const float *x, const int64_t *target, int B, void *workspace,
                           int64_t workspace_bytes, float *logp_out, float *stats_out, void *stream);

/* Replaces optimizer.step() of optim.SGD(lr, momentum) (mnist.py:140,43) fused with the
 * 1/world_size of DDP's gradient averaging (mnist.py:135-138) as grad_scale:
 *   g = grads*grad_scale; buf = first_step ? g : mu*buf + g; p = p - lr*buf
 * over n contiguous floats, rounded exactly like the ATen CPU kernels (bit-exact). */
int b200mnist_sgd_momentum(float *params, const float *grads, float *momentum, int64_t n, float lr, float mu,
                           float grad_scale, int first_step, void *stream);

/* ---- DDP exchange over NVLink peer memory ---------------------------------------------
 * Replaces DDP's gradient all_reduce (mnist.py:135-138) AND optimizer.step() (mnist.py:43) for ranks on one
 * NVLink-connected box, one process per GPU.  Every rank owns a PEER WINDOW: one device allocation
 *   [ flag page B200MNIST_PEER_FLAG_BYTES | parameter bucket | gradient-sum bucket | staging bucket 0 | staging 1 |
 *     receive area ]
 * (buckets of B200MNIST_FLAT_ELEMS floats; b200mnist_peer_window_bytes() in all) that the other ranks map through CUDA IPC.  The model's parameters and
 * p.grad live in the window; the backward of step s writes its gradient into staging bucket (s & 1).
 *
 * Set-up (these allocate / map device memory and synchronise; call them once, outside any capture):
 *   _create : cudaMalloc + zero-fill of a window on the current device; writes its address and a
 *             B200MNIST_IPC_HANDLE_BYTES-byte handle (HOST buffers) to send to the other ranks' processes;
 *   _open   : map another process's window from its handle (works for a GPU this process does not enumerate,
 *             e.g. under CUDA_VISIBLE_DEVICES=<one GPU>, as long as the two GPUs can reach each other);
 *   _close / _destroy : undo _open / _create;
 *   _probe  : set-up check (one tiny kernel): out[r] (DEVICE, world_size uint32) = word
 *             B200MNIST_PEER_PROBE_WORD of rank r's flag page, read through this process's mappings -- the path
 *             the exchange kernels use. */
#define B200MNIST_PEER_FLAG_BYTES 4096
#define B200MNIST_IPC_HANDLE_BYTES 64
#define B200MNIST_PEER_PROBE_WORD 192
int64_t b200mnist_peer_window_bytes(void);
int b200mnist_peer_window_create(void **window_out_host, void *ipc_handle_out_host);
int b200mnist_peer_window_open(const void *ipc_handle_host, void **window_out_host);
int b200mnist_peer_window_close(void *window);
int b200mnist_peer_window_destroy(void *window);
int b200mnist_peer_window_probe(void *const *windows_host, int world_size, uint32_t *out, void *stream);
/* The whole data-parallel training step of mnist.py:39-43 under DDP (mnist.py:135-138) for this rank, on ONE stream:
 * forward, loss, backward, gradient averaging over all ranks and the SGD update -- b200mnist_train_step with the
 * exchange inside its own kernels.
 *   windows_host : HOST array of world_size window addresses as mapped in THIS process ([rank] = the local one)
 *   parity       : which staging bucket / receive slots this step uses (step & 1, the same on every rank)
 *   momentum     : this rank's momentum bucket (local memory)
 * fc1.weight + fc1.bias (93 % of the bucket) are exchanged two-shot by extra CTAs in the launch of conv1's
 * weight-gradient kernel: this rank reduces its 1/world_size slice over all ranks' staging buckets in rank order,
 * applies g = sum/world_size; buf = first_step ? g : mu*buf + g; p = p - lr*buf, and stores the new parameters and
 * the sum into every rank's window (fc1's momentum is therefore kept on the slice's owner only).  The step's last
 * kernel sums the batch-split partial gradients of conv1 / conv2 / fc2, pushes each {value, step epoch} into every
 * peer's receive area, waits for the peers' values, adds them in rank order and applies the same update; it also
 * waits until every peer's fc1 slice has landed here.  Parameters and p.grad (the SUM over ranks) are the window's
 * buckets.  x, target, B, workspace, loss_out as for b200mnist_train_step.  Enqueues only (no allocation,
 * capturable); every rank must call it once per step; world_size in {2, 4, 8}. */
int b200mnist_peer_train_step(void *const *windows_host, int rank, int world_size, int parity, float *momentum,
                              const float *x, const int64_t *target, int B, void *workspace, int64_t workspace_bytes,
                              float lr, float mu, int first_step, float *loss_out, void *stream);
/* The same step reading its batch from a resident ring (see b200mnist_train_step_ring; every rank has its own ring). */
int b200mnist_peer_train_step_ring(void *const *windows_host, int rank, int world_size, int parity, float *momentum,
                                   const b200mnist_ring_t *ring_host, float *x, int64_t *target, int B, void *workspace,
                                   int64_t workspace_bytes, float lr, float mu, int first_step, float *loss_out,
                                   void *stream);
/* ---- per-layer entry points (used by the autograd bridge and the unit tests) ------- */

/* F.max_pool2d(F.relu(conv1(x)), 2, 2), mnist.py:26-27.  x [B,1,28,28], w [20,1,5,5],
 * out [B,20,12,12]; idx [B,20,12,12] uint8 = slot (dy*2+dx) of the window maximum. */
int b200mnist_conv1_relu_pool_fwd(const float *x, const float *w, const float *bias, float *out, uint8_t *idx,
                                  int B, void *stream);
/* F.max_pool2d(F.relu(conv2(x)), 2, 2), mnist.py:28-29.  in [B,20,12,12], w [50,20,5,5],
 * out [B,50,4,4] (= the [B,800] view of mnist.py:30). */
int b200mnist_conv2_relu_pool_fwd(const float *in, const float *w, const float *bias, float *out, uint8_t *idx,
                                  int B, void *stream);
/* The same conv2 + ReLU + max-pool (mnist.py:28-29) on the tensor cores: a shifted-window implicit GEMM (the image
 * pair is staged once in shared memory; each of the 25 kernel offsets is an MMA whose operand descriptor points at
 * the same planes shifted by (ky,kx)), tcgen05.mma kind::tf32 with the 3xTF32 split, pooling in the epilogue.
 * `scratch` (b200mnist_conv2_fwd_tc_scratch_bytes() bytes, 16-byte aligned) receives the hi/lo-split weights.
 * b200mnist_conv2_relu_pool_fwd and the fused step take this path from B >= 128 (measured break-even);
 * B200MNIST_CONV2_TC = 1 / 0 in the environment forces it on / off. */
int64_t b200mnist_conv2_fwd_tc_scratch_bytes(void);
int b200mnist_conv2_relu_pool_fwd_tc(const float *in, const float *w, const float *bias, float *out, uint8_t *idx,
                                     int B, void *scratch, void *stream);
/* F.relu(fc1(x)), mnist.py:31.  in [B,800], w [500,800], out [B,500].  `scratch` must hold
 * b200mnist_fc1_scratch_bytes(B) bytes, 256-byte aligned (split-K partial sums on the fp32 path; hi / lo copies of the
 * operands on the tensor-core path). */
int64_t b200mnist_fc1_scratch_bytes(int B);
int b200mnist_fc1_relu_fwd(const float *in, const float *w, const float *bias, float *out, int B, void *scratch,
                           void *stream);
/* The same F.relu(fc1(x)) (mnist.py:31) forced onto the tensor-core path: both operands are first written to `scratch`
 * as hi / lo copies (the 3xTF32 error-compensated split, fp32-level accuracy), then a TMA-fed tcgen05.mma kind::tf32
 * GEMM with its accumulator in TMEM.  b200mnist_fc1_relu_fwd and the fused step pick it automatically for B >= 384
 * (B200MNIST_FC1_TC_MIN in the environment moves the threshold); this entry point exists so it can be tested at any B.
 * Same `scratch` as b200mnist_fc1_relu_fwd. */
int b200mnist_fc1_relu_fwd_tc(const float *in, const float *w, const float *bias, float *out, int B, void *scratch,
                              void *stream);
/* F.log_softmax(fc2(h), dim=1) and F.nll_loss (mean), mnist.py:32-33,41, plus the start of
 * loss.backward() (mnist.py:42): dz2 = (softmax - onehot)/B and dz1 = (dz2 . W2) * (h > 0).
 * h [B,500]; logp [B,10]; dz2 [B,10]; dz1 [B,500] (NULL to skip the backward part);
 * loss_out 1 float; scratch: (B + 64) floats, zero-filled once. */
int b200mnist_fc2_logsoftmax_nll(const float *h, const float *w, const float *bias, const int64_t *target, int B,
                                 float *logp, float *loss_out, float *dz2, float *dz1, void *scratch,
                                 void *stream);
/* No counterpart in the reference: b200mnist_fc2_logsoftmax_nll for callers that cannot supply a
 * scratch buffer because they may not allocate device memory where they run -- e.g. a plugin
 * invoked from inside another runtime's callback (cudaMalloc is off limits there: it can
 * synchronise the whole device), or a process whose device memory is entirely owned by a
 * foreign pool.  Same arguments and results as b200mnist_fc2_logsoftmax_nll minus `scratch`;
 * the scratch is a fixed 16,640-byte array inside the library image (resident on every GPU
 * that loads the library, for the life of the process), hence the limit
 * B <= B200MNIST_STATIC_MAX_BATCH; larger B returns -10006 with an explanatory message.
 * The call allocates nothing (no cudaMalloc/cudaFree, no allocating library routine).
 * The FIRST call on each device looks up the array's device address (a per-device set-up
 * that is not stream work): make that first call before starting any stream capture.  The
 * address is looked up again on every call made outside a capture; a call made while the stream
 * is capturing uses the address remembered from the last such look-up on that device ordinal, so
 * after cudaDeviceReset (or any re-creation of the device's context) make one uncaptured call
 * before capturing again -- and a graph captured before the reset must not be replayed.
 * All calls on a device share the one array: two calls whose kernels may overlap on the GPU
 * (different streams, no ordering between them) race on it -- serialise them. */
#define B200MNIST_STATIC_MAX_BATCH 4096
int b200mnist_fc2_logsoftmax_nll_static(const float *h, const float *w, const float *bias, const int64_t *target,
                                        int B, float *logp, float *loss_out, float *dz2, float *dz1, void *stream);
/* Backward of log_softmax + fc2 for an arbitrary upstream gradient (autograd of mnist.py:32-33
 * when the loss is not the fused NLL): dz2 = dlogp - exp(logp)*sum_j(dlogp), dz1 = (dz2 . W2)*(h>0). */
int b200mnist_head_bwd(const float *logp, const float *dlogp, const float *h, const float *w, int B, float *dz2,
                       float *dz1, void *stream);
/* Backward of the two linear layers, mnist.py:42: gW2 [10,500], gb2 [10], gW1 [500,800],
 * gb1 [500], dp2 [B,800] = dz1 . W1.  scratch: b200mnist_fc_bwd_scratch_bytes(B), 256-byte aligned. */
int64_t b200mnist_fc_bwd_scratch_bytes(int B);
int b200mnist_fc_bwd(const float *dz2, const float *dz1, const float *h, const float *p2, const float *w1, int B,
                     float *gw2, float *gb2, float *gw1, float *gb1, float *dp2, void *scratch, void *stream);
/* Backward of conv2+ReLU+pool, mnist.py:42: given dp2 [B,50,4,4] (gradient of the pooled
 * output), the saved p2/idx2/p1 and conv2.weight, writes gw [50,20,5,5], gb [50] and
 * dp1 [B,20,12,12].  scratch: b200mnist_conv2_bwd_scratch_bytes(B).  dp2, p2, p1, w and scratch must be
 * 16-byte aligned, idx2 4-byte aligned (vector loads).  From B >= 128 the data and the weight gradient run on the
 * tensor cores (tcgen05, 3xTF32); B200MNIST_CONV2_DGRAD_TC / B200MNIST_CONV2_WGRAD_TC = 1 / 0 in the environment
 * force either kernel on / off. */
int64_t b200mnist_conv2_bwd_scratch_bytes(int B);
int b200mnist_conv2_bwd(const float *dp2, const float *p2, const uint8_t *idx2, const float *p1, const float *w,
                        int B, float *gw, float *gb, float *dp1, void *scratch, void *stream);
/* Backward of conv1+ReLU+pool, mnist.py:42 (weights only: the input needs no gradient).
 * gw [20,1,5,5], gb [20].  scratch: b200mnist_conv1_bwd_scratch_bytes(B). */
int64_t b200mnist_conv1_bwd_scratch_bytes(int B);
int b200mnist_conv1_bwd(const float *dp1, const float *p1, const uint8_t *idx1, const float *x, int B, float *gw,
                        float *gb, void *scratch, void *stream);

/* Replaces the DataLoader of mnist.py:118-125 for a dataset resident on the GPU: for each of the n indices, copy
 * image[index] (28x28 uint8, images is [n_images,784]) to data [n,1,28,28] as ((x/255) - mean)/std -- ToTensor +
 * Normalize with the reference's fp32 roundings, bit-identical pixels -- and labels[index] to target [n]. */
int b200mnist_gather_normalize_u8(const uint8_t *images, const int64_t *labels, const int64_t *indices, int n,
                                  int n_images, float mean, float stdv, float *data, int64_t *target, void *stream);

/* The worker's op in examples/smoke-dist/dist_sendrecv.py:32, torch.mul(inp, inp). */
int b200mnist_square(const float *in, float *out, int64_t n, void *stream);

/* Host plumbing for `data.to(device)` (mnist.py:38) in a double-buffered pipeline: on `copy_stream` wait for
 * ev_slot_free (cudaEvent_t: the step that last used this input slot is done), copy `bytes` from PINNED host
 * memory to the device slot, record ev_copied; make `compute_stream` wait for ev_copied.  Enqueues only.
 * b200mnist_event_record is cudaEventRecord(event, stream) (to mark a slot's step as done). */
int b200mnist_stage_inputs(void *dst_device, const void *src_host_pinned, int64_t bytes, void *copy_stream,
                           void *compute_stream, void *ev_slot_free, void *ev_copied);
int b200mnist_event_record(void *event, void *stream);
/* Refill of a host-fed ring slot (see b200mnist_ring_t): cudaMemcpyAsync of `bytes` from pinned host (or device) memory
 * to dst_device on copy_stream and, if landed_word is not NULL, a stream-ordered 32-bit write of `seq` to that DEVICE
 * word behind it (cuStreamWriteValue32): no kernel, no event, nothing on the compute stream. */
int b200mnist_stage_record(void *dst_device, const void *src, int64_t bytes, void *copy_stream, uint32_t *landed_word,
                           uint32_t seq);

/* ---- instrumentation for bench.py (no counterpart in the reference) ----------------
 * b200mnist_launch_count: kernels this library has launched in this process so far.
 * b200mnist_profile_begin/end: while enabled, a CUDA event is recorded on `stream` after
 * every kernel the library launches; _end synchronises the stream and returns the number n
 * of kernels seen, their durations (ms, event to event) and their names (static strings)
 * in the HOST arrays.  delay_us > 0 first parks the stream for that long so the following
 * launches queue up (no host launch gaps inside the timings).  Not for use under capture. */
long long b200mnist_launch_count(void);
/* b200mnist_delay: parks `stream` for delay_us microseconds (one spinning thread) so that launches issued behind it
 * queue up.  b200mnist_fma_peak: a register-resident packed-FMA (fma.rn.f32x2) loop on `blocks` CTAs of 256 threads,
 * no memory traffic; `out` needs blocks*256 floats; returns the FLOPs the launch performs (blocks*256*64*iters), or a
 * negative error.  Timed by tools/measure_peaks.py to get the fp32 FMA roofline of this GPU. */
int b200mnist_delay(void *stream, int delay_us);
int64_t b200mnist_fma_peak(float *out, int blocks, int iters, void *stream);
int b200mnist_profile_begin(void *stream, int delay_us);
int b200mnist_profile_end(float *ms_out_host, const char **names_out_host, int cap, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200MNIST_H */
