/*
 * ddp_b200.h — C ABI of libddp_b200.so: the B200 (sm_100a) gradient-sync +
 * clip + AdamW hot path of a DistributedDataParallel fine-tuning step.
 *
 * Every entry point is `extern "C"`, takes plain pointers / sizes / scalars and
 * an explicit cudaStream_t (as void*), allocates nothing and never throws.
 * Return value: 0 = OK, >0 = cudaError_t, <0 = DDPB200_E_* below.
 *
 * "REF/"   = the upstream repo (naman1618/Distributed-Inference-with-PyTorch-and-LiteLLM)
 * "TORCH/" = torch 2.11.0 site-packages, the un-vendored dependency in which the
 *            reference's DDP / clip / AdamW arithmetic actually lives
 *            (REF/requirements.txt:2, REF/README.md:7).
 *
 * Vocabulary: a *segment* is one parameter tensor's slot (its grad, its bucket
 * view, its optimizer state); a *bucket* is a flat buffer holding the grads of
 * several segments back to back with no padding (TORCH reducer.hpp:347-406);
 * a *chunk* is the unit of work one CTA streams through registers.
 */
#ifndef DDP_B200_H
#define DDP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPB200_ABI_VERSION 2

#if defined(__GNUC__)
#define DDPB200_API __attribute__((visibility("default")))
#else
#define DDPB200_API
#endif

/* element types of a stream of values in HBM */
#define DDPB200_F32  0
#define DDPB200_BF16 1

/* library error codes (negative); positive codes are cudaError_t */
#define DDPB200_E_INVALID      (-1)  /* null pointer, negative size, bad enum        */
#define DDPB200_E_UNSUPPORTED  (-2)  /* dtype combination not compiled               */
#define DDPB200_E_TOO_LARGE    (-3)  /* size beyond a stated limit                   */

/* flags for ddpb200_fused_clip_adamw */
#define DDPB200_ADAMW_ROUND_CLIPPED_GRAD 1u /* g = round_to_grad_dtype(g*coef) first, as
                                               TORCH/nn/utils/clip_grad.py:174 leaves it  */
#define DDPB200_ADAMW_STORE_GRAD         2u /* write the unscaled+clipped grad back        */
#define DDPB200_ADAMW_MAXIMIZE           4u /* TORCH/optim/adam.py maximize=True           */

/* rounding of the grad-norm scalars (ddpb200_l2norm_finalize) */
#define DDPB200_NORM_EXACT_F32   0 /* total = sqrt(sum_t sumsq_t), all fp32                 */
#define DDPB200_NORM_TORCH_BF16  1 /* per-tensor norm, total, total+1e-6, coef and clamp each
                                      rounded to bf16 as torch does for bf16 grads
                                      (TORCH/nn/utils/clip_grad.py:96,106-108,165-169)      */

/* where a segment table lives */
#define DDPB200_TABLE_HOST   0 /* host memory: copied into the launch parameters (no H2D copy, no
                                  lifetime to manage) — for pointers that change every step    */
#define DDPB200_TABLE_DEVICE 1 /* device memory: built once, reused every step                  */

/* One segment (parameter tensor) of a multi-tensor launch, 48 bytes. A table of segments lives
 * on the device (built once, reused every step) or on the host (see DDPB200_TABLE_*). What
 * ptr[0..3] mean is stated per entry point. chunk_begin is the exclusive prefix sum of
 * ceil(numel/chunk_elems) over the table (ddpb200_plan_chunks fills it).
 * A table may be sliced: pass table+first, the slice's n and chunk count; kernels rebase on
 * the first chunk_begin they see. */
typedef struct ddpb200_seg {
  void*   ptr[4];
  int64_t numel;
  int64_t chunk_begin;
} ddpb200_seg;

/* ---- library --------------------------------------------------------------- */
DDPB200_API int         ddpb200_abi_version(void);
DDPB200_API const char* ddpb200_error_string(int code);
/* SM count of the current device (grid sizing); <0 on error. */
DDPB200_API int         ddpb200_sm_count(void);

/* ---- a-1  bucket planner (host only, integer, bit-exact) -------------------
 * Replaces dist._compute_bucket_assignment_by_size
 * (TORCH/include/torch/csrc/distributed/c10d/reducer.hpp:586-598, called at
 *  TORCH/nn/parallel/distributed.py:1237-1244; BucketAccumulator reducer.hpp:39-43).
 * Greedy: tensors are appended per (dtype,device) key; a bucket closes when its byte
 * size reaches the current limit, then that key advances to the next limit; leftovers
 * are appended; a tensor with expect_sparse[idx]!=0 gets a bucket of its own (limit 0).
 * tensor_indices==NULL -> indices are 0..n-1 and buckets are finally sorted by their
 * smallest index; otherwise tensors arrive in grad-ready order and keep it.
 * Outputs: out_indices[n] (tensor indices, bucket after bucket),
 *          out_bucket_start[n+1] (bucket b owns out_indices[start[b]..start[b+1])),
 *          out_bucket_limit[n], *out_n_buckets. */
DDPB200_API int ddpb200_bucket_assign(const int64_t* numel, const int32_t* elem_size,
                          const int64_t* key, const uint8_t* expect_sparse,
                          const int64_t* tensor_indices, int64_t n,
                          const uint64_t* size_limits, int64_t n_limits,
                          int64_t* out_indices, int64_t* out_bucket_start,
                          uint64_t* out_bucket_limit, int64_t* out_n_buckets);

/* Element offsets of each tensor inside its bucket: running sum of numel, no padding
 * (Bucket.offsets/lengths, TORCH reducer.hpp:372-375). out_offset[i] is for out_indices[i];
 * out_bucket_numel[b] is the bucket length in elements. */
DDPB200_API int ddpb200_bucket_layout(const int64_t* numel, const int64_t* indices,
                          const int64_t* bucket_start, int64_t n_buckets,
                          int64_t* out_offset, int64_t* out_bucket_numel);

/* ---- segment-table helper (host only) --------------------------------------
 * Fills seg[i].chunk_begin for a table whose numel fields are set; returns the total
 * chunk count in *out_total_chunks. chunk_elems must be a positive multiple of 2048. */
DDPB200_API int ddpb200_plan_chunks(ddpb200_seg* host_table, int64_t n, int64_t chunk_elems,
                        int64_t* out_total_chunks);
/* A chunk size that gives every CTA of the persistent grid several chunks. */
DDPB200_API int64_t ddpb200_suggest_chunk_elems(int64_t total_numel, int32_t sm_count);

/* ---- a-2 / a-4  pack, unpack (+scale, +cast, + optional sum of squares) ------
 * Replaces the per-parameter `bucket_view = grad * (1/world_size)` of
 * Reducer::mark_variable_ready_dense (TORCH reducer.hpp:276) and the per-parameter
 * `grad.copy_(bucket_view)` of finalize_bucket_dense/copy_bucket_to_grad
 * (reducer.hpp:328-334) by ONE launch per bucket.
 * ptr[0] = source, ptr[1] = destination. dst[i] = round_dst(fp32(src[i]) * scale);
 * scale==1.0f is an exact copy/cast. If sumsq_out != NULL, the fp32 sum of squares of the
 * values as stored in dst is written to sumsq_out[t] per segment (deterministic order;
 * needs chunk_partials[total_chunks] floats and a zeroed uint32 *ticket, which the kernel
 * leaves zero again). dst may alias src (in-place scale of a bucket view). */
DDPB200_API int ddpb200_pack_scale_cast(const ddpb200_seg* table, int32_t table_loc, int64_t n, int64_t total_chunks,
                            int64_t chunk_elems, int32_t src_dtype, int32_t dst_dtype,
                            float scale, float* sumsq_out, float* chunk_partials,
                            uint32_t* ticket, void* stream);

/* ---- a-5  clip_grad_norm_ ----------------------------------------------------
 * Replaces torch._foreach_norm + stack + vector_norm (TORCH/nn/utils/clip_grad.py:96,
 * 106-108) with one pass: ptr[1] = grad (read only). Writes per-segment fp32 sums of
 * squares to sumsq_out[n]. Deterministic (fixed reduction tree; no float atomics). */
DDPB200_API int ddpb200_l2norm_sumsq(const ddpb200_seg* table, int32_t table_loc, int64_t n, int64_t total_chunks,
                         int64_t chunk_elems, int32_t dtype, float* sumsq_out,
                         float* chunk_partials, uint32_t* ticket, void* stream);

/* Same result as ddpb200_l2norm_sumsq for a caller that has no scratch memory to give and no way
 * to get any. (No counterpart in the reference or in torch: an addition of this library.)
 * Who that is: a host binding that only ever sees its own input and output buffers (a cgo / JNI /
 * ctypes caller without a device allocator behind it); a caller inside a region where device
 * allocation is off limits because cudaMalloc synchronises the whole device (the inner loop of a
 * latency-critical step); a process whose device memory has all been handed to a pool that the
 * call site cannot reach.
 * The per-chunk partial sums and the ticket counter live in an array built into the library
 * image: DDPB200_STATIC_SCRATCH_MAX_CHUNKS floats + 256 bytes = 1 MiB + 256 B on every GPU that
 * loads the library, held until the process exits. total_chunks beyond that limit is refused with
 * DDPB200_E_TOO_LARGE ("size exceeds the stated limit"): use a larger chunk_elems, or the entry
 * point above with your own scratch. The function calls no cudaMalloc/cudaFree and nothing that
 * allocates; inputs (table, ptr[1] = grad) and the output sumsq_out[n] are still the caller's.
 * The first call on each device looks the array's address up (cudaGetSymbolAddress, with that
 * device current) and keeps it; make that first call before any stream capture begins.
 * ALL calls on a device share the one array: two calls whose kernels can overlap on the GPU
 * (different streams without an ordering between them) race on it and give wrong sums. Calls on
 * one stream, or ordered by events, are safe. The array keeps the previous call's contents; the
 * ticket is cleared at the start of every call, the partials are overwritten before use. */
#define DDPB200_STATIC_SCRATCH_MAX_CHUNKS 262144
DDPB200_API int ddpb200_l2norm_sumsq_static(const ddpb200_seg* table, int32_t table_loc, int64_t n,
                                            int64_t total_chunks, int64_t chunk_elems, int32_t dtype,
                                            float* sumsq_out, void* stream);

/* total norm + clip coefficient, on device, no host sync
 * (TORCH/nn/utils/clip_grad.py:106-108,165,169). out[0] = total norm (pre-clip),
 * out[1] = min(max_norm/(total+1e-6), 1). mode = DDPB200_NORM_*. */
DDPB200_API int ddpb200_l2norm_finalize(const float* sumsq, int64_t n, int32_t mode, float max_norm,
                            float* out, void* stream);

/* in-place g *= coef[0] for callers that want torch's clip side effect on p.grad
 * (TORCH/nn/utils/clip_grad.py:174). ptr[1] = grad. */
DDPB200_API int ddpb200_scale_by_device_scalar(const ddpb200_seg* table, int32_t table_loc, int64_t n,
                                   int64_t total_chunks, int64_t chunk_elems, int32_t dtype,
                                   const float* coef, void* stream);

/* ---- a-3 + a-5  bucket allreduce fused with the norm pass, over NVLink peer memory -------------
 * Replaces ProcessGroupNCCL::allreduce on the buckets (c10d Reducer all_reduce_bucket,
 * TORCH/include/torch/csrc/distributed/c10d/reducer.hpp:524) followed by torch._foreach_norm
 * (TORCH/nn/utils/clip_grad.py:96): SUM over ranks of the (pre-scaled) bucket arena, in place on every
 * rank, and the per-chunk sums of squares of the result, in one pass.
 * The arena must be symmetric memory: every rank's copy mapped into this process (arena[p]), same
 * layout on all ranks; arena_multicast is its NVSwitch multicast address or NULL. dev_table is this
 * rank's DEVICE table with ptr[1] = bucket views inside arena[rank]. Chunk c is reduced by rank
 * c % world and written to all ranks; its sum of squares is written to partials[p][c] on all ranks.
 * use_multicast != 0: multimem.ld_reduce / multimem.st through the switch ((1+1/n)*S NVLink bytes per
 * GPU and direction); 0: 256-bit peer loads and stores (2(n-1)/n*S), fp32 sum in rank order.
 * scale folds the pack pass's 1/world_size in for gradients that already live in the arena
 * (gradient_as_bucket_view: the in-place `bucket_view.div_(world_size)` of reducer.hpp:276 is then not
 * launched at all): peer mode rounds fp32(g)*scale to the bucket dtype per rank before the sum — the
 * bits a separate pack would have stored, for any world size; multicast mode scales the switch's sum,
 * which is the same number only for a power-of-two scale and outside subnormals. scale = 1 otherwise.
 * The CALLER orders the ranks: a barrier after every rank has packed and before this call, and a
 * barrier after it before anyone reads the arena or the partials (then ddpb200_fold_chunk_partials
 * gives sumsq_out[n], bitwise identical on every rank). */
#define DDPB200_MAX_PEERS 8
typedef struct ddpb200_peer_map {
  void*   arena[DDPB200_MAX_PEERS];
  float*  partials[DDPB200_MAX_PEERS];
  void*   arena_multicast;
  int32_t rank;
  int32_t world;
} ddpb200_peer_map;
DDPB200_API int ddpb200_allreduce_sumsq_chunks(const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks,
                                               int64_t chunk_elems, int32_t dtype,
                                               const ddpb200_peer_map* peers, int32_t use_multicast,
                                               float scale, void* stream);
DDPB200_API int ddpb200_fold_chunk_partials(const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks,
                                            const float* chunk_partials, float* sumsq_out, void* stream);

/* The reduce-scatter half of the above, for the sharded update: the owner of a chunk
 * ((absolute chunk id) % world) stores the reduced chunk only in its OWN arena; the per-chunk sums of
 * squares still go to every rank. Same arguments and barriers as ddpb200_allreduce_sumsq_chunks. */
DDPB200_API int ddpb200_reduce_scatter_sumsq_chunks(const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks,
                                                    int64_t chunk_elems, int32_t dtype,
                                                    const ddpb200_peer_map* peers, int32_t use_multicast,
                                                    float scale, void* stream);

/* ---- §8e/§8f-4  sharded update: AdamW on the owned chunks + all-gather of the parameters -----------
 * ZeRO-1-style extension of a-6 (no counterpart in the reference; the replicated path above is the
 * drop-in): after ddpb200_reduce_scatter_sumsq_chunks and the norm finalize, each rank updates only
 * the chunks it owns — gradients, parameters and its 1/world slice of exp_avg / exp_avg_sq read
 * from LOCAL memory, arithmetic exactly that of ddpb200_fused_clip_adamw — and stores the new
 * parameters into every rank's parameter arena (param_peers: symmetric memory with the SAME layout
 * as the gradient arena; multicast store if use_multicast). dev_table: this rank's device table with
 * ptr[1] = gradient views inside grad_arena_local. Chunk c's state lives at element offset
 * (c / world) * chunk_elems of the shards (chunk ids absolute, as in the table). The caller puts a
 * barrier after this call before any rank reads parameters. Parameters after a step are
 * bit-identical to the replicated path's. */
DDPB200_API int ddpb200_sharded_adamw_allgather(
    const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks, int64_t chunk_elems, int32_t dtype,
    const void* grad_arena_local, const ddpb200_peer_map* param_peers, void* exp_avg_shard,
    void* exp_avg_sq_shard, double lr, const float* lr_ptr, double beta1, double beta2, double eps,
    double weight_decay, int64_t step, const float* clip_coef, uint32_t flags, int32_t use_multicast,
    void* stream);

/* ---- a-6  fused unscale + clip + AdamW ---------------------------------------
 * Replaces torch.optim.AdamW.step (TORCH/optim/adamw.py:20-49, TORCH/optim/adam.py:
 * 215-272; element math of TORCH/include/ATen/native/cuda/fused_adam_utils.cuh:41-93,
 * ADAM_MODE::ADAMW, amsgrad=False) and folds in the `g *= clip_coef` pass of
 * clip_grad.py:174 and GradScaler's unscale. One pass: reads p,g,m,v, writes p,m,v.
 * ptr[0]=param ptr[1]=grad ptr[2]=exp_avg ptr[3]=exp_avg_sq.
 *   g  = fp32(grad); if grad_scale: g /= *grad_scale; if clip_coef: g *= *clip_coef
 *        (rounded to the grad dtype first under DDPB200_ADAMW_ROUND_CLIPPED_GRAD)
 *   p -= (lr*wd)*p;  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g
 *   p -= (lr/bc1) * m / (sqrt(v)/sqrt(bc2) + eps),  bc = 1 - beta^step (step >= 1)
 * lr_ptr (device, optional) overrides lr. found_inf (device, optional): if *found_inf==1
 * the launch is a no-op. step_ptr (device fp32 scalar, optional) overrides step: the bias corrections are
 * then taken on the device from *step_ptr (same formula), which is how the GradScaler flow runs without a
 * host sync — the caller adds 1 to it before the launch and subtracts *found_inf after, as
 * TORCH/optim/adam.py's fused path does with its device-resident state_steps. clip_coef / grad_scale are device pointers so that no host sync
 * sits between the norm and the update. */
DDPB200_API int ddpb200_fused_clip_adamw(const ddpb200_seg* table, int32_t table_loc, int64_t n, int64_t total_chunks,
                             int64_t chunk_elems, int32_t param_dtype, int32_t grad_dtype,
                             int32_t state_dtype, double lr, const float* lr_ptr,
                             double beta1, double beta2, double eps, double weight_decay,
                             int64_t step, const float* step_ptr, const float* clip_coef,
                             const float* grad_scale, const float* found_inf, uint32_t flags, void* stream);

/* The two fp32 scalars the kernel derives from `step`: out[0] = 1 - b1^step,
 * out[1] = sqrt(1 - b2^step), with b = beta rounded to fp32 (as the recurrences use it), the
 * power taken in double on the host and rounded once
 * (TORCH/include/ATen/native/cuda/fused_adam_utils.cuh:131-150). */
DDPB200_API int ddpb200_adamw_bias_corrections(double beta1, double beta2, int64_t step, float* out);

/* ---- self-checks of the hand-written arithmetic (diagnostics; no counterpart in the reference) -----
 * The fast modes of ddpb200_fused_clip_adamw write sqrt.rn and div.rn out as MUFU + packed f32x2 steps
 * (csrc/adamw_math.cuh) and divide by sqrt(bc2) with a reciprocal + residual (csrc/common.cuh). These
 * run the written-out forms beside the IEEE intrinsics on the device and count differing results:
 * out3[0] = number of mismatches, out3[1..2] = the first offending input (zero out3 before the call).
 *   _sqrt          every float bit pattern in [first_bits, first_bits+count) as v (stay below 2^20 =
 *                  0x49800000, the box); with c > 0 also the whole denominator sqrt(v)/c + eps
 *   _div           count pseudo-random (p, a, b) triples spread over the box, p - a/b
 *   _div_by_const  every significand of x in [1,2) for each of the n constants (device array) */
DDPB200_API int ddpb200_selfcheck_sqrt(uint32_t first_bits, uint64_t count, float c, float eps,
                                       unsigned long long* out3, void* stream);
DDPB200_API int ddpb200_selfcheck_div(uint64_t seed, uint64_t count, unsigned long long* out3, void* stream);
DDPB200_API int ddpb200_selfcheck_div_by_const(const float* constants, int32_t n, unsigned long long* out3,
                                               void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDP_B200_H */
