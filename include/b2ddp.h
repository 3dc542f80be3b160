/*
 * b2ddp.h — C ABI of libb2ddp.so: the B200 (sm_100a) kernels for the data-parallel
 * hot path of valohai/distributed-llms-example (label-masked cross-entropy, AdamW
 * step, gradient mean all-reduce over NVLink peer memory).
 *
 * Conventions
 *   - every function returns 0 (B2_OK) or a negative b2_status; a human-readable
 *     message for the calling thread's last failure is at b2_last_error();
 *   - all data pointers are DEVICE pointers owned by the caller unless the comment
 *     says "host"; nothing here allocates device memory except b2_comm_create;
 *   - every launch is asynchronous on the cudaStream_t passed as `stream`
 *     (a void* here so that the header needs no CUDA include);
 *   - no torch types cross this boundary; the Python adapters pass
 *     tensor.data_ptr() and torch.cuda.current_stream().cuda_stream.
 *
 * The reference has no native interface of its own (three Python scripts), so
 * each entry point cites the reference call site whose library kernel it replaces.
 * SP = site-packages of the image (torch 2.11.0, transformers 5.5.0).
 */
#ifndef B2DDP_H
#define B2DDP_H

#include <stddef.h>
#include <stdint.h>
#include <sys/types.h> /* ssize_t */

#ifdef __cplusplus
extern "C" {
#endif

#define B2DDP_ABI_VERSION 1

typedef void* b2_stream_t; /* cudaStream_t */

enum b2_dtype { B2_F32 = 0, B2_BF16 = 1 };

enum b2_status {
  B2_OK = 0,
  B2_ERR_INVALID = -1,   /* bad argument (null pointer, size, alignment, dtype) */
  B2_ERR_CUDA = -2,      /* a CUDA runtime call failed; message has the cudaError */
  B2_ERR_STATE = -3,     /* call order (peers not opened, context destroyed...) */
  B2_ERR_TIMEOUT = -4,   /* a peer did not arrive at a cross-GPU barrier */
  B2_ERR_TOO_LARGE = -5  /* request exceeds a stated fixed limit */
};

/* Message of the last failing call made by this host thread ("" if none). */
const char* b2_last_error(void);
/* B2DDP_ABI_VERSION the library was built with. */
int b2_abi_version(void);

/* ------------------------------------------------------------------------- *
 * (c) label-masked cross-entropy, mean over rows whose label != ignore_index
 *
 * Replaces CrossEntropyLoss(ignore_index=-100)(lm_logits.view(-1, V), labels.view(-1))
 * at SP/transformers/models/t5/modeling_t5.py:1112-1117, reached from
 * train-accelerator.py:220-221, train-task.py:289-290 and train-torchrun.py:140
 * (forward), train-accelerator.py:222 / train-task.py:292 (backward).
 * ------------------------------------------------------------------------- */

/* Bytes of scratch b2_ce_forward needs for (n_rows, vocab): split-row partials
 * and the arrival counter of its last-block reduction. The counter must be zero
 * before the first call (cudaMemset the workspace once); the kernel leaves it zero. */
size_t b2_ce_workspace_bytes(int64_t n_rows, int64_t vocab);

/* logits [n_rows, vocab] row-major, row pitch `row_stride` elements, dtype f32|bf16.
 * labels [n_rows] int64. Outputs: row_lse[n_rows] fp32 = logsumexp of each row
 * (written for ignored rows too), loss_out[1] fp32 = sum_valid(lse - z[y]) / n_valid
 * (NaN when n_valid == 0, like torch), nvalid_out[1] fp32 = n_valid.
 * A label outside [0, vocab) that is not ignore_index poisons loss_out with NaN
 * (torch raises a device assert there). Deterministic: fixed reduction order. */
int b2_ce_forward(const void* logits, const int64_t* labels, int64_t n_rows,
                  int64_t vocab, int64_t row_stride, int64_t ignore_index, int dtype,
                  float* row_lse, float* loss_out, float* nvalid_out,
                  void* workspace, size_t workspace_bytes, b2_stream_t stream);

/* No counterpart in the reference: an addition for hosts that cannot allocate device
 * memory. Same computation and outputs as b2_ce_forward, but the scratch is a fixed array
 * inside the library image instead of a caller workspace, so this path performs no
 * cudaMalloc/cudaFree and takes no workspace argument. Meant for embedders whose runtime
 * owns all device memory after start-up (an allocation would fail or is not permitted) and
 * for thin bindings with nowhere to keep a workspace alive between calls.
 *   - at most B2_CE_NOALLOC_MAX_ROWS rows per call; more returns B2_ERR_TOO_LARGE;
 *   - the first call on each device resolves the array's device address (with that device
 *     current); make that first call before any stream capture starts;
 *   - one array per device shared by every call: calls whose kernels may overlap on the GPU
 *     (unordered streams) race on it and must be serialised by the caller;
 *   - the array costs about 1.03 MiB on every GPU that loads the library, until exit. */
#define B2_CE_NOALLOC_MAX_ROWS 8192
int b2_ce_forward_noalloc(const void* logits, const int64_t* labels, int64_t n_rows,
                          int64_t vocab, int64_t row_stride, int64_t ignore_index, int dtype,
                          float* row_lse, float* loss_out, float* nvalid_out,
                          b2_stream_t stream);

/* dlogits[i, j] = (exp(z_ij - lse_i) - [j == y_i]) * grad_out / n_valid for valid
 * rows, 0 for ignored rows, written in the logits dtype with pitch `row_stride`.
 * grad_out and nvalid are device scalars (no host sync). dlogits must not alias logits. */
int b2_ce_backward(const void* logits, const int64_t* labels, const float* row_lse,
                   const float* grad_out, const float* nvalid, int64_t n_rows,
                   int64_t vocab, int64_t row_stride, int64_t ignore_index, int dtype,
                   void* dlogits, b2_stream_t stream);

/* ------------------------------------------------------------------------- *
 * (c') fused lm_head + cross-entropy on the tensor cores
 *
 * Replaces, for T5-family models, the pair
 *     lm_logits = self.lm_head(sequence_output)          (SP/transformers/models/t5/modeling_t5.py:1110)
 *     loss = CrossEntropyLoss(ignore_index=-100)(lm_logits.view(-1, V), labels.view(-1))   (:1112-1117)
 * reached from train-accelerator.py:220-221, train-task.py:289-290 and train-torchrun.py:140: the
 * [n_rows, vocab] logits are produced tile by tile by tcgen05.mma (bf16 x bf16 -> fp32 in TMEM,
 * operands staged by TMA) and consumed by the cross-entropy epilogue straight out of TMEM, so
 * they never reach HBM in the forward pass.
 *   hidden [n_rows, d_model] bf16 row-major (the lm_head input, already scaled by d_model^-0.5
 *          where the model ties its embeddings), weight [vocab, d_model] bf16 row-major
 *          (nn.Linear layout), labels [n_rows] int64; d_model a multiple of 64; bases 16-byte aligned.
 *   round_bf16 != 0: each logit is rounded to bf16 before the softmax - bit-for-bit what the
 *          unfused path computes from its bf16 logits tensor; 0 keeps the fp32 accumulator.
 * Outputs as b2_ce_forward: row_lse[n_rows], loss_out[1] (NaN when no row is valid),
 * nvalid_out[1]. logits_out: NULL, or a bf16 [n_rows, logits_stride] buffer (pitch >= vocab,
 * multiple of 8, base 16-byte aligned; the up to 7 padding columns between vocab and the next
 * multiple of 8 are written as zeros) that receives the logits once, on their way out of
 * tensor memory (TMA store from the epilogue) - the forward still never reads them back, and the
 * backward can then be b2_ce_backward on that buffer, which is the fastest pairing on B200
 * (re-reading 2 bytes per logit costs less than recomputing 2*d_model flop per logit).
 * workspace: b2_lmce_workspace_bytes(n_rows, vocab) bytes, 16-byte aligned, its first 16 bytes
 * zero before the first call (the kernel leaves them zero). Deterministic.
 * ------------------------------------------------------------------------- */
size_t b2_lmce_workspace_bytes(int64_t n_rows, int64_t vocab);
int b2_lmce_forward(const void* hidden, const void* weight, const int64_t* labels, int64_t n_rows,
                    int64_t vocab, int64_t d_model, int64_t ignore_index, int round_bf16,
                    float* row_lse, float* loss_out, float* nvalid_out, void* logits_out,
                    int64_t logits_stride, void* workspace, size_t workspace_bytes,
                    b2_stream_t stream);
/* Memory-lean backward of the same pair (loss.backward() at train-accelerator.py:222 /
 * train-task.py:292) for a forward that kept no logits: the
 * logits tile is recomputed by the same tcgen05 main loop and the epilogue writes
 *   dlogits[i, j] = (exp(z_ij - lse_i) - [j == y_i]) * grad_out / n_valid      (0 on ignored rows)
 * in bf16 with row pitch `row_stride` elements (>= vocab, multiple of 8). The caller forms
 * d hidden = dlogits @ weight and d weight = dlogits^T @ hidden with library GEMMs. */
int b2_lmce_backward_dlogits(const void* hidden, const void* weight, const int64_t* labels,
                             const float* row_lse, const float* grad_out, const float* nvalid,
                             int64_t n_rows, int64_t vocab, int64_t d_model, int64_t ignore_index,
                             int round_bf16, void* dlogits, int64_t row_stride, b2_stream_t stream);

/* ------------------------------------------------------------------------- *
 * (b) AdamW step over many tensors in one launch
 *
 * Replaces torch.optim.AdamW.step (SP/torch/optim/adam.py:215-272 -> _single_tensor_adam
 * :347-550 / torch._fused_adamw_) called at train-accelerator.py:224, train-task.py:295
 * and by HF Trainer (SP/transformers/trainer.py:1760) for train-torchrun.py:140.
 * Per element, all fp32:
 *   g' = g * (*grad_scale)            (only when grad_scale != NULL: clip fold-in)
 *   p  = p * (1 - lr*wd)
 *   m  = m + (1-beta1) * (g' - m)
 *   v  = v*beta2 + (1-beta2) * g'*g'
 *   p  = p - (lr/bc1) * m / (sqrt(v)/sqrt(bc2) + eps)
 * ------------------------------------------------------------------------- */

typedef struct b2_adamw_hyper {
  double lr;           /* group["lr"] as the host holds it (re-read every step: LambdaLR) */
  double beta1;
  double beta2;
  double eps;
  double weight_decay;
  int64_t step;        /* torch's state["step"] AFTER this step's increment (>= 1) */
} b2_adamw_hyper;
/* The library derives, in double like adam.py:529-534, then rounds to fp32:
 *   decay = 1 - lr*wd, w1 = 1 - beta1, 1 - beta2, step_size = lr / (1 - beta1^step),
 *   bias_correction2_sqrt = sqrt(1 - beta2^step). */

#define B2_ADAMW_MAX_HYPERS 16
#define B2_ADAMW_CHUNK 8192 /* elements per work item; fixed so plans are portable */

typedef struct b2_adamw_tensor {
  void* p;             /* fp32 parameter, updated in place */
  const void* g;       /* fp32 gradient */
  void* m;             /* fp32 exp_avg, updated in place */
  void* v;             /* fp32 exp_avg_sq, updated in place */
  int64_t numel;
  int32_t hyper;       /* index into the hypers[] passed to b2_adamw_step */
  int32_t chunk_begin; /* filled by b2_adamw_plan: first work item of this tensor */
} b2_adamw_tensor;

/* Host-side: fills chunk_begin of tensors_host[0..n) and returns the total number of
 * work items (ceil(numel / B2_ADAMW_CHUNK) summed), or a negative b2_status. */
int64_t b2_adamw_plan(b2_adamw_tensor* tensors_host, int32_t n_tensors);

/* tensors_dev: the planned table copied to device memory by the caller.
 * hypers_host: host array of n_hypers (<= B2_ADAMW_MAX_HYPERS) entries, passed to the
 * kernel by value (no device copy). grad_scale: device fp32 scalar or NULL. */
int b2_adamw_step(const b2_adamw_tensor* tensors_dev, int32_t n_tensors,
                  int64_t total_chunks, const b2_adamw_hyper* hypers_host,
                  int32_t n_hypers, const float* grad_scale, b2_stream_t stream);

/* One flat partition (p, g, m, v of `numel` fp32 each); no table in device memory. */
int b2_adamw_step_flat(void* p, const void* g, void* m, void* v, int64_t numel,
                       const b2_adamw_hyper* hyper_host, const float* grad_scale,
                       b2_stream_t stream);

/* Global L2 norm of all gradients in the table + clip coefficient, one read-only launch
 * (SURVEY.md 8f row 1). Replaces torch.nn.utils.clip_grad_norm_ as HF Trainer calls it
 * (SP/transformers/trainer.py:1753-1757, max_grad_norm=1.0) between sync and step for
 * train-torchrun.py:140. norm_out[0] = ||g||_2, coef_out[0] = min(1, max_norm/(norm+1e-6));
 * pass coef_out as grad_scale to b2_adamw_step instead of rewriting the gradients.
 * tensors_dev is the same planned table b2_adamw_step takes. workspace:
 * b2_grad_norm_workspace_bytes() bytes, 16-byte aligned, first 16 bytes zero before first use. */
size_t b2_grad_norm_workspace_bytes(void);
int b2_grad_norm_clip(const b2_adamw_tensor* tensors_dev, int32_t n_tensors, int64_t total_chunks,
                      float max_norm, void* workspace, size_t workspace_bytes, float* norm_out,
                      float* coef_out, b2_stream_t stream);

/* ------------------------------------------------------------------------- *
 * (a) gradient mean all-reduce over NVLink peer memory
 *
 * Replaces the Reducer's NCCL bucket all-reduce (SP/torch/nn/parallel/distributed.py
 * :1255-1284 -> default_comm_hooks.hpp:36-51) driven from train-accelerator.py:189,222
 * and train-torchrun.py:127,140, and the per-tensor dist.all_reduce + "/= size" of
 * train-task.py:65-69. One process per GPU; the caller moves the 64-byte handles
 * between processes (torch.distributed all_gather in the Python adapter).
 * ------------------------------------------------------------------------- */

typedef struct b2_comm b2_comm;

#define B2_IPC_HANDLE_BYTES 64
#define B2_MAX_WORLD 8

enum b2_algo { B2_ALGO_AUTO = 0, B2_ALGO_ONE_SHOT = 1, B2_ALGO_TWO_SHOT = 2,
               B2_ALGO_TWO_SHOT_INPLACE = 3, /* registered buffers: pull variant (emulation / A-B) */
               B2_ALGO_PUSH_INPLACE = 4      /* registered buffers: one-pass reduce + push (default) */ };

/* Allocates this rank's symmetric staging buffer (2 halves of staging_bytes, used
 * alternately) and its flag pad on the CURRENT device. world in {1,2,4,8}. */
int b2_comm_create(int rank, int world, size_t staging_bytes, b2_comm** out);
/* Writes this rank's B2_IPC_HANDLE_BYTES handle (host memory). */
int b2_comm_get_handle(b2_comm* c, void* handle_out);
/* handles: world * B2_IPC_HANDLE_BYTES host bytes in rank order (own slot ignored). */
int b2_comm_open_peers(b2_comm* c, const void* handles);
/* Largest numel one b2_allreduce call accepts for a wire dtype. */
int64_t b2_comm_max_numel(b2_comm* c, int wire_dtype);
/* buf: fp32 [numel], reduced in place to scale * sum_r buf_r, summed in rank order
 * 0..W-1 on every rank (bit-identical across ranks and across the two algorithms).
 * wire_dtype: B2_F32, or B2_BF16 to move bf16 over NVLink (half the bytes; changes
 * numerics): result = bf16(sum in fp32 of bf16(scale*buf_r)), still stored as fp32.
 * numel of any size: calls larger than the staging half are chunked internally.
 * All ranks must call with the same numel/wire/algo in the same order. */
int b2_allreduce(b2_comm* c, void* buf, int64_t numel, int wire_dtype, float scale,
                 int algo, b2_stream_t stream);
/* Zero-copy path for buffers that live across steps (DDP buckets after the Reducer's rebuild;
 * SURVEY.md 8f row 2). Each rank exports the cudaMalloc allocation that contains its buffer
 * (b2_comm_export: 64-byte handle + byte offset of ptr inside that allocation), the host
 * all-gathers both, and b2_comm_register maps the peers' buffers. b2_allreduce_registered then
 * works IN PLACE on the registered buffers with no staging copy: the owner of a slice reads it
 * from every rank's buffer, adds in rank order and pushes the result into every rank's buffer
 * in the same pass (reads inbound, writes outbound on NVLink), between a one-CTA gate kernel and
 * a final barrier. fp32 on the wire, same rank-ordered sum (bit-identical to b2_allreduce). The buffer must stay
 * allocated until b2_comm_unregister; memory from CUDA's virtual-memory allocators
 * (expandable segments) cannot be exported and has to use b2_allreduce. */
int b2_comm_export(b2_comm* c, const void* ptr, void* handle_out, uint64_t* offset_out);
int b2_comm_register(b2_comm* c, void* local_ptr, size_t bytes, const void* handles,
                     const uint64_t* offsets, int* reg_id_out);
int b2_comm_unregister(b2_comm* c, int reg_id);
int b2_allreduce_registered(b2_comm* c, int reg_id, int64_t numel, float scale,
                            b2_stream_t stream);
/* In-switch reduction (NVLink SHARP / NVLS): the same mean all-reduce as
 * b2_allreduce, with the sum formed inside the NVSwitch (multimem.ld_reduce / multimem.st) so
 * that each NVLink port carries ~S bytes per direction instead of 2(W-1)/W*S. Results are
 * bit-identical on every rank; against the rank-ordered sum they agree to fp32 rounding (the
 * switch's add order), the same standard as NCCL's ring/tree/NVLS orders. World 2, 4 or 8.
 * Set-up, after b2_comm_open_peers, in this order on every rank (host barriers between steps):
 *   rank 0: b2_nvls_create(bytes) -> POSIX fd + rounded size; the host passes the fd to the
 *           other ranks (SCM_RIGHTS over a Unix socket), which call b2_nvls_import(fd, rounded);
 *   all:    b2_nvls_add_device; <barrier>; b2_nvls_bind; <barrier>.
 * b2_nvls_supported: 1 when the current device can join a multicast object. */
int b2_nvls_supported(int* supported_out);
int b2_nvls_create(b2_comm* c, size_t bytes, int* fd_out, size_t* rounded_out);
int b2_nvls_import(b2_comm* c, int fd, size_t rounded_bytes);
int b2_nvls_add_device(b2_comm* c);
int b2_nvls_bind(b2_comm* c);
int b2_allreduce_nvls(b2_comm* c, void* buf, int64_t numel, float scale, b2_stream_t stream);
/* Zero-copy NVLS: buffers ALLOCATED inside the multicast-bound region (same offset on every
 * rank) are reduced in place by the switch with no copy at all: a one-CTA gate kernel, then one
 * pass  v = multimem.ld_reduce [mc+off+i]; multimem.st [mc+off+i], v*scale  over the slice this
 * rank owns, then one barrier. b2_nvls_region reports the region (unicast base address, bytes,
 * and the offset where the pool starts: the bytes before it are b2_allreduce_nvls's staging);
 * b2_nvls_pool_attach makes it the backing store of b2_pool_alloc / b2_pool_free, which have the
 * alloc_fn / free_fn signatures of torch.cuda.memory.CUDAPluggableAllocator (a bump allocator
 * handing torch's caching allocator whole segments), so that a torch MemPool active while the
 * DDP Reducer rebuilds its buckets places the buckets in the region. */
int b2_nvls_region(b2_comm* c, uint64_t* base_out, uint64_t* bytes_out, uint64_t* pool_offset_out);
int b2_nvls_pool_attach(b2_comm* c);
void* b2_pool_alloc(ssize_t size, int device, void* stream);
void b2_pool_free(void* ptr, ssize_t size, int device, void* stream);
int b2_allreduce_nvls_inplace(b2_comm* c, uint64_t offset_bytes, int64_t numel, float scale,
                              b2_stream_t stream);
/* A peer that does not reach a cross-GPU barrier within B2_AR_TIMEOUT_MS (default 600 000 ms,
 * the scale of NCCL's watchdog) is an error, never silently wrong gradients: the rank that gave
 * up raises a sticky status word on EVERY rank, and from then on every all-reduce of this
 * communicator on every rank fills the bucket it was given with NaN instead of reducing it
 * (NCCL would block or abort where the reference calls it: train-task.py:68, the Reducer's
 * all-reduce behind train-accelerator.py:222 / train-torchrun.py:140).
 * b2_comm_status: synchronising check (copies the device word). b2_comm_poll: the same answer
 * from one plain host load of a mapped pinned word the kernels mirror the status into - cheap
 * enough to call every step (the Python adapters do, in the DDP hook, average_gradients and
 * AdamW.step). b2_comm_status_word: host address of that word (non-zero = timed out), for
 * callers that poll from their own C++ (csrc_torch/ddp_hook.cpp). */
int b2_comm_status(b2_comm* c);
int b2_comm_poll(b2_comm* c);
/* Launch shape of the zero-copy all-reduce kernels (defaults from B2_AR_BLOCKS,
 * B2_AR_PACKS_PER_THREAD, B2_AR_UNROLL): at most max_blocks CTAs (<= 128), one more CTA per
 * packs_per_thread 16-byte packs per thread, unroll 0 = tuned default / 1 = more / 2 = fewer
 * loads in flight per thread (in-switch kernel: 0 and 2 = 2 loads, 1 = 8, 3 = 4). Negative = leave
 * unchanged. Same call on every rank. */
int b2_comm_tune(b2_comm* c, int max_blocks, int packs_per_thread, int unroll);
const void* b2_comm_status_word(b2_comm* c);
/* Sum of squares folded into the all-reduce: replaces the separate read of
 * every gradient that torch.nn.utils.clip_grad_norm_ makes between sync and step in HF Trainer
 * (SP/transformers/trainer.py:1753-1757, max_grad_norm=1.0; train-torchrun.py:115-128,140). Once
 * enabled, every all-reduce of this communicator (staging, registered, NVLS) also accumulates
 * sum(x*x) over the reduced values it produced - the slice owner has them in registers - into a
 * double kept in the communicator: per-CTA partials in a fixed order, exchanged with the data,
 * so the total is bit-identical on every rank and run to run. b2_comm_sumsq_numel: elements
 * folded since the last b2_comm_sumsq_clip / b2_comm_sumsq_enable (host-side count; the caller
 * checks it covers exactly the gradients it is about to clip). b2_comm_sumsq_clip: one-thread
 * kernel on `stream` (which must be ordered after the all-reduces): norm_out[0] = sqrt(total),
 * coef_out[0] = min(1, max_norm / (norm + 1e-6)), total <- 0. */
int b2_comm_sumsq_enable(b2_comm* c, int on);
int64_t b2_comm_sumsq_numel(b2_comm* c);
int b2_comm_sumsq_clip(b2_comm* c, float max_norm, float* norm_out, float* coef_out,
                       b2_stream_t stream);
int b2_comm_destroy(b2_comm* c);

/* Test hook for machines with fewer GPUs than ranks: the same device code run as
 * ONE launch over `world` virtual ranks whose buffers all live on the current
 * device (no cross-launch waiting). bufs_host: host array of `world` device
 * pointers, each fp32 [numel], all reduced in place. staging: device scratch of
 * b2_allreduce_emulated_scratch_bytes(world, numel) bytes, 256-byte aligned. */
size_t b2_allreduce_emulated_scratch_bytes(int world, int64_t numel);
int b2_allreduce_emulated(void* const* bufs_host, int world, int64_t numel,
                          int wire_dtype, float scale, int algo, void* staging,
                          b2_stream_t stream);
/* Same, with the sum-of-squares fold switched on: sumsq_out (device, double[world]) receives
 * each virtual rank's total (all equal). */
int b2_allreduce_emulated_sumsq(void* const* bufs_host, int world, int64_t numel,
                                int wire_dtype, float scale, int algo, void* staging,
                                double* sumsq_out, b2_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B2DDP_H */
