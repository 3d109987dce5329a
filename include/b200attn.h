/*
 * b200attn.h -- C ABI of libb200attn.so: the B200 (sm_100a) paged-attention decode backend.
 *
 * This is the drop-in boundary "b2" of SURVEY.md section 8(b).  The reference
 * (ThakiCloud/vllm-backend-public) has no FFI for this path -- it launches an external vLLM
 * server process (reference benchmark-vllm/vllm_manager.py:371-374,744-773) and never sees
 * the kernels -- so every entry point below cites (a) the vLLM operation an engine behind that
 * launch contract performs and (b) the reference symbol that parameterises it.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / pybind types.
 *   - the caller owns every device and host buffer, including the workspace; the library
 *     allocates nothing on the device and keeps no device-side state.
 *   - every kernel entry point only enqueues work on `stream` (a cudaStream_t passed as void*)
 *     and returns without synchronising.
 *   - return value: 0 (B2_OK) or a negative B2_E* code; the message for the calling thread's last
 *     failure is available from b2_last_error().  Nothing throws or aborts across this boundary.
 *   - device pointers must be 16-byte aligned; head_dim must be 128 (Llama-3 geometry).
 *
 * Tensor layouts (element strides unless noted)
 *   q, out        [B, Hq, d]                 bf16   (row b at q + b*q_stride)
 *   k_new, v_new  [T, Hkv, d]                bf16   (token t at k_new + t*k_token_stride)
 *   k_cache       [num_blocks, Hkv, bs, d]   bf16 or e4m3 bytes ("HND": one (block, head) tile of
 *   v_cache                                  bs x d elements is contiguous)
 *   block_tables  [B, max_nb] int32          physical block id of each logical block
 *   seq_lens      [B] int32                  tokens that attend (including the one just written)
 *   slot_mapping  [T] int64                  block_id*bs + offset, or < 0 to skip the token
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 1

#if defined(__GNUC__)
#define B2_API __attribute__((visibility("default")))
#else
#define B2_API
#endif

/* error codes */
#define B2_OK             0
#define B2_EINVAL        -1   /* bad argument (null pointer, misalignment, non-positive size)   */
#define B2_EUNSUPPORTED  -2   /* shape / dtype outside what the sm_100a kernels are built for     */
#define B2_ECUDA         -3   /* CUDA runtime / driver error at launch or descriptor encoding     */
#define B2_ENOMEM        -4   /* block pool exhausted (block manager) or workspace too small      */
#define B2_ENOTFOUND     -5   /* unknown sequence id (block manager)                              */
#define B2_EEXISTS       -6   /* sequence id already present (block manager)                      */

/* KV-cache element types: VLLMConfig.dtype / additional_args["kv-cache-dtype"]
 * (reference benchmark-vllm/models.py:14,23-28 -> vllm_templates.py:57,82-87). */
#define B2_KV_BF16      0
#define B2_KV_FP8_E4M3  1

B2_API int         b2_version(void);          /* returns B2_ABI_VERSION */
B2_API const char* b2_last_error(void);       /* thread-local, never NULL */

/* ---------------------------------------------------------------------------------------------
 * KV write.  Replaces vLLM `reshape_and_cache` (csrc/cache_kernels.cu) inside the engine the
 * reference launches (vllm_manager.py:371-374); page size = VLLMConfig.block_size
 * (benchmark-vllm/models.py:10 -> --block-size, vllm_templates.py:54).
 *   for t in [0,T): s = slot_mapping[t]; if s < 0 skip;
 *     k_cache[s / bs][h][s % bs][:] = k_new[t][h][:]   (same for v)
 *   B2_KV_FP8_E4M3 stores sat_rne_e4m3(x / *k_scale): clamp to +-448, round to nearest even.
 * k_scale / v_scale: DEVICE pointers to one float each (may be NULL for B2_KV_BF16).
 * cache_block_stride / cache_head_stride are in elements of the cache dtype.
 * ------------------------------------------------------------------------------------------- */
B2_API int b2_reshape_and_cache(const void* k_new, const void* v_new,
                         void* k_cache, void* v_cache,
                         const int64_t* slot_mapping,
                         int T, int Hkv, int d, int bs,
                         int64_t k_token_stride, int64_t v_token_stride,
                         int64_t cache_block_stride, int64_t cache_head_stride,
                         int kv_dtype, const float* k_scale, const float* v_scale,
                         void* stream);

/* Decode-step form of the KV write for engines that keep block_tables RESIDENT on the device: token t is the
 * newest token of batch row t (position seq_lens[t] - 1), so besides writing K/V the call records the block its
 * slot names, block_tables[t][(seq_lens[t] - 1) / bs] = slot_mapping[t] / bs (rows with slot < 0 are skipped).
 * Rewritten every step (idempotent), so a step's host->device traffic is seq_lens + slot_mapping (O(B) words)
 * instead of the whole [B, max_nb] table; a row is uploaded in full only when a new sequence takes it.
 * block_tables: DEVICE int32 [B, max_nb]; seq_lens: DEVICE int32 [B].  Other arguments as b2_reshape_and_cache. */
B2_API int b2_reshape_and_cache_decode(const void* k_new, const void* v_new, void* k_cache, void* v_cache,
                                const int64_t* slot_mapping, int32_t* block_tables, const int32_t* seq_lens,
                                int B, int Hkv, int d, int bs, int max_nb,
                                int64_t k_token_stride, int64_t v_token_stride,
                                int64_t cache_block_stride, int64_t cache_head_stride,
                                int kv_dtype, const float* k_scale, const float* v_scale, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Paged-attention decode.  Replaces vLLM `paged_attention_v1/_v2` (csrc/attention/) for the decode
 * step of the engine behind the reference's launch contract; B <= VLLMConfig.max_num_seqs
 * (models.py:9), max_nb = ceil(max_model_len / block_size) (models.py:15,10), Hq/Hkv are the
 * per-GPU head counts after tensor_parallel_size sharding (models.py:11).
 *   out[b][hq] = softmax_j(scale * <q[b][hq], K[b][j][hq / G]>) . V[b][j][hq / G],  j < seq_lens[b]
 *   G = Hq / Hkv; fp32 softmax and accumulation; out rounded to bf16 once; seq_len 0 -> zeros.
 * num_splits: 0 = choose automatically, 1 = single pass ("v1"), >1 = split-K over the KV length
 * with a log-sum-exp merge ("v2").  The workspace (device memory, 16-byte aligned) must hold
 * b2_paged_decode_workspace_bytes(...) bytes for the same B/Hq/d/max_nb/bs/num_splits; it is
 * scratch (no state survives the call) and may be NULL when that size is 0.
 * ------------------------------------------------------------------------------------------- */
/* The returned size serves every batch size 1..B of the geometry (the automatic schedule cuts smaller batches
 * into finer chunks, so the need of one batch size alone is not monotone in B): size the workspace once for
 * max_num_seqs and never reallocate it under captured CUDA graphs. */
B2_API size_t b2_paged_decode_workspace_bytes(int B, int Hq, int Hkv, int d, int bs, int max_nb,
                                       int num_splits);

/* the split count b2_paged_decode will actually use for num_splits = 0 */
B2_API int b2_paged_decode_auto_splits(int B, int Hq, int Hkv, int d, int bs, int max_nb, int kv_dtype);

/* Host-side schedule choice for callers that hold seq_lens on the host (the scheduler always does; it is the
 * role `max_seq_len` plays for vLLM's paged_attention_v1/v2 choice).  Returns the num_splits to pass to
 * b2_paged_decode: 1 when the single-pass grid (one CTA per (sequence, kv head)) keeps the machine balanced
 * for these lengths, 0 (dynamic persistent schedule) otherwise; negative on bad arguments.  Pure host code. */
B2_API int b2_paged_decode_plan_splits(const int32_t* seq_lens_host, int B, int Hq, int Hkv, int d, int bs,
                                int kv_dtype);

B2_API int b2_paged_decode(void* out, const void* q,
                    const void* k_cache, const void* v_cache,
                    const int32_t* block_tables, const int32_t* seq_lens,
                    int B, int Hq, int Hkv, int d, int bs, int max_nb,
                    int64_t num_blocks,
                    int64_t q_stride, int64_t out_stride,
                    int64_t cache_block_stride, int64_t cache_head_stride,
                    float scale, int kv_dtype,
                    const float* k_scale, const float* v_scale,
                    void* workspace, size_t workspace_bytes, int num_splits,
                    void* stream);

/* The decode step of one layer in ONE call: the KV write of the step's new token per row, then the attention over the
 * rows including it.  Replaces the pair vLLM issues per layer and decode step -- `reshape_and_cache` followed by
 * `paged_attention_v1/_v2` -- inside the engine the reference launches (vllm_manager.py:371-374); sizes as above
 * (models.py:9,10,11,15).  Result and cache contents are bit-identical to b2_reshape_and_cache(k_new, v_new, ...,
 * slot_mapping, T = B, ...) followed by b2_paged_decode(...) with the same arguments.
 *   k_new / v_new: bf16 [B, Hkv, d] rows (token strides in elements); slot_mapping[b] = the slot of row b's new token,
 *   i.e. position seq_lens[b] - 1 of its block table (seq_lens COUNT the new token); slot < 0 skips the write.
 *   The slot must belong to row b alone among the rows of the call (true of every engine: shared prefix blocks are full).
 * With num_splits = 1 (the single-pass grid; b2_paged_decode_plan_splits picks it for long uniform batches) and
 * Hq / Hkv <= 16 it is ONE kernel launch: the CTA that owns (row, kv head) stores the new K/V row itself and issues
 * the TMA copy of the stage that holds it only once the store is visible to the async proxy (fence.proxy.async) -- the step has no second launch
 * and no kernel boundary.  b2_paged_decode_append_fused tells (1 / 0; negative on bad arguments).  Every other schedule
 * cuts a row over several CTAs: the call then launches the KV write in front of the decode grid (same results). */
B2_API int b2_paged_decode_append_fused(int Hq, int Hkv, int num_splits);
B2_API int b2_paged_decode_append(void* out, const void* q,
                           const void* k_new, const void* v_new, const int64_t* slot_mapping,
                           int64_t k_token_stride, int64_t v_token_stride,
                           void* k_cache, void* v_cache,
                           const int32_t* block_tables, const int32_t* seq_lens,
                           int B, int Hq, int Hkv, int d, int bs, int max_nb,
                           int64_t num_blocks,
                           int64_t q_stride, int64_t out_stride,
                           int64_t cache_block_stride, int64_t cache_head_stride,
                           float scale, int kv_dtype,
                           const float* k_scale, const float* v_scale,
                           void* workspace, size_t workspace_bytes, int num_splits,
                           void* stream);

/* b2_paged_decode for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc.  No counterpart in the reference.
 * Same operation, arguments and results as b2_paged_decode, minus the workspace: the split-K partials and the claim
 * counter live in a file-scope device array of B2_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the
 * library holds it until exit).  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
 * A call whose schedule needs more scratch than that (large B x max_nb; b2_paged_decode_workspace_bytes tells) returns
 * B2_ENOMEM with the sizes in b2_last_error -- use b2_paged_decode with a caller-owned workspace then.
 * Every call on a GPU shares the one array: calls must be ordered on the GPU (same stream, or event-ordered); calls
 * whose kernels overlap race on it. */
#define B2_STATIC_SCRATCH_BYTES (32u << 20)
B2_API int b2_paged_decode_static(void* out, const void* q,
                           const void* k_cache, const void* v_cache,
                           const int32_t* block_tables, const int32_t* seq_lens,
                           int B, int Hq, int Hkv, int d, int bs, int max_nb,
                           int64_t num_blocks,
                           int64_t q_stride, int64_t out_stride,
                           int64_t cache_block_stride, int64_t cache_head_stride,
                           float scale, int kv_dtype,
                           const float* k_scale, const float* v_scale,
                           int num_splits, void* stream);

/* Per-engine decode handle: the same operation as b2_paged_decode with everything that does not change between
 * steps fixed at creation -- cache pointers / strides (their two TMA descriptors are encoded once, a driver call
 * each), geometry, scale, kv dtype and scales, the workspace.  A call then is argument checks + kernel launches:
 * no driver call, no getenv, and two launches instead of three for the dynamic schedule (create zeroes the
 * workspace header on `stream`; every call leaves it zeroed).  The handle is host memory only; the caller still owns
 * caches and workspace, which must outlive it and must not be shared with other plans or b2_paged_decode calls
 * in flight.  workspace_bytes >= b2_paged_decode_workspace_bytes(max_B, ..., 0); the first 256 bytes stay reserved
 * for the claim counter, so a call with num_splits > 1 needs b2_paged_decode_workspace_bytes(B, ..., num_splits) + 256
 * (B2_ENOMEM otherwise).  NULL on bad arguments
 * (b2_last_error).  Single-writer, like every handle of this library.  (vLLM analogue: the attention metadata
 * builder + `paged_attention_v1/_v2` behind reference benchmark-vllm/vllm_manager.py:371-374.) */
typedef struct b2_decode_plan b2_decode_plan;
B2_API b2_decode_plan* b2_decode_plan_create(const void* k_cache, const void* v_cache, int max_B, int Hq, int Hkv,
                                      int d, int bs, int max_nb, int64_t num_blocks,
                                      int64_t cache_block_stride, int64_t cache_head_stride,
                                      float scale, int kv_dtype, const float* k_scale, const float* v_scale,
                                      void* workspace, size_t workspace_bytes, void* stream);
B2_API void b2_decode_plan_destroy(b2_decode_plan* plan);
B2_API int b2_decode_plan_run(b2_decode_plan* plan, void* out, const void* q,
                       const int32_t* block_tables, const int32_t* seq_lens, int B,
                       int64_t q_stride, int64_t out_stride, int num_splits, void* stream);
/* The whole decode step of an engine that keeps its block tables resident on the device, in one call:
 * b2_reshape_and_cache_decode (KV write of the step's new token per row + its table entry) followed by
 * b2_decode_plan_run, with the same results in `out`, in the caches the plan was created over (they are WRITTEN here)
 * and in block_tables.  Same vLLM pair as b2_paged_decode_append (`reshape_and_cache` + `paged_attention_v1/_v2`,
 * reference vllm_manager.py:371-374).  One launch when b2_paged_decode_append_fused(Hq, Hkv, num_splits) is 1: the CTA
 * that owns (row, kv head) stores the new K/V row and the table entry, and takes the id of the row's last block from
 * the slot instead of from the entry it is writing.  k_new / v_new / q may be pinned host memory (read in place). */
B2_API int b2_decode_plan_run_append(b2_decode_plan* plan, void* out, const void* q,
                              const void* k_new, const void* v_new, const int64_t* slot_mapping,
                              int64_t k_token_stride, int64_t v_token_stride,
                              int32_t* block_tables, const int32_t* seq_lens, int B,
                              int64_t q_stride, int64_t out_stride, int num_splits, void* stream);

/* Diagnostic / A-B variant kept for the ncu comparison DESIGN.md cites: the same contract as
 * b2_paged_decode, computed by a plain CUDA-core kernel (no TMA, no tensor cores).  Not used by
 * the engine. */
B2_API int b2_paged_decode_simt(void* out, const void* q,
                         const void* k_cache, const void* v_cache,
                         const int32_t* block_tables, const int32_t* seq_lens,
                         int B, int Hq, int Hkv, int d, int bs, int max_nb,
                         int64_t q_stride, int64_t out_stride,
                         int64_t cache_block_stride, int64_t cache_head_stride,
                         float scale, int kv_dtype,
                         const float* k_scale, const float* v_scale,
                         void* stream);

/* ---------------------------------------------------------------------------------------------
 * Prompt-chunk (prefill) attention over the paged cache.  Replaces vLLM's
 * chunked-prefill / prefix attention (flash_attn_varlen_func with a block table, `context_attention_fwd`) in the
 * engine behind the reference's launch contract; chunk sizes follow the scheduler's token budget, table rows
 * max_nb = ceil(max_model_len / block_size) (reference benchmark-vllm/models.py:15,10).
 *   chunk c = token rows [cu_q[c], cu_q[c+1]) of q / out; its tokens sit at positions ctx_start[c] + i of their
 *   sequence, whose blocks are block_tables[c][:].  The chunk's own K/V must already be in the cache
 *   (b2_rope_and_cache / b2_reshape_and_cache run first).
 *   out[t][hq] = softmax_{j <= ctx_start[c] + i}(scale * <q[t][hq], K[c][j][hq / G]>) . V[c][j][hq / G]
 *   fp32 softmax / accumulation, softmax weights rounded to bf16 for the P.V product, out rounded to bf16 once.
 * T = cu_q[n_chunks] (host knowledge: it sizes the grid).  cu_q, ctx_start, block_tables: DEVICE int32.
 * kv_dtype: B2_KV_BF16 only (B2_EUNSUPPORTED otherwise).  Over an fp8-e4m3 cache the caller first gathers the
 * blocks the chunks can see into a bf16 scratch cache with b2_dequant_kv_blocks (below) and passes that scratch
 * with its own block table and the cache's scales: the conversion moves 3 bytes per cache element once, the
 * attention then re-reads bf16.
 * k_scale / v_scale: optional DEVICE fp32 scalars (NULL = 1): the cache holds K / k_scale and V / v_scale;
 * k_scale multiplies the softmax scale, v_scale the output -- exactly the fp8 dequantisation of b2_paged_decode.
 * ------------------------------------------------------------------------------------------- */
B2_API int b2_prefill_attention(void* out, const void* q, const void* k_cache, const void* v_cache,
                         const int32_t* block_tables, const int32_t* cu_q, const int32_t* ctx_start,
                         int n_chunks, int T, int Hq, int Hkv, int d, int bs, int max_nb,
                         int64_t num_blocks, int64_t q_stride, int64_t out_stride,
                         int64_t cache_block_stride, int64_t cache_head_stride,
                         float scale, int kv_dtype, const float* k_scale, const float* v_scale,
                         void* stream);

/* fp8-e4m3 -> bf16 gather of whole cache blocks (the prefill side of VLLMConfig additional_args `kv-cache-dtype`,
 * reference benchmark-vllm/models.py:23-28): out block i = float(e4m3 block src_blocks[i]) for i < n, which is
 * exact in bf16; the per-tensor scales are applied by b2_prefill_attention (k_scale / v_scale arguments).
 * k_out / v_out: bf16 [n][Hkv][bs][d], contiguous, caller-owned.  src_blocks: DEVICE int32 [n]. */
B2_API int b2_dequant_kv_blocks(void* k_out, void* v_out, const void* k_cache, const void* v_cache,
                         const int32_t* src_blocks, int64_t n, int Hkv, int d, int bs,
                         int64_t cache_block_stride, int64_t cache_head_stride, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Element-wise ops of one decoder layer around the attention path: what the
 * launched engine runs between its GEMMs.  bf16 activations (VLLMConfig.dtype, reference
 * benchmark-vllm/models.py:14), fp32 math, one rounding on the way out; strides in elements.
 *
 * b2_rms_norm        residual == NULL: out = bf16(w * x * rsqrt(mean(x^2) + eps))
 *                    else: h = bf16(x + residual) is written to residual, then normalised (vLLM
 *                    fused_add_rms_norm).  out may alias x.
 * b2_silu_and_mul    out[t][i] = silu(gate_up[t][i]) * gate_up[t][I + i]          (vLLM silu_and_mul)
 * b2_rope_and_cache  qkv [T, (Hq + 2 Hkv) * d] = q | k | v per token.  Neox rotary (pairs i, i + d/2) with
 *                    cos_sin fp32 [max_pos, d] (cos[0..d/2) | sin[0..d/2) per position) applied to q and k IN
 *                    PLACE: x1' = bf16(x1*cos - x2*sin), x2' = bf16(x2*cos + x1*sin) (products and sum each
 *                    rounded in fp32, no fma); then k', v go to the paged cache through slot_mapping exactly as
 *                    b2_reshape_and_cache does (slot < 0: rotated in place, not cached).
 *                    (vLLM rotary_embedding + reshape_and_cache in one pass over qkv.)
 * b2_sample          tokens[b] = argmax_i logits[b][i] when temperature == NULL or temperature[b] <= 0 (lowest
 *                    index among equal maxima); else argmax_i (logits[b][i] / T_b + g_i), g_i = -log(-log(u_i)),
 *                    u_i = (hash24(seeds[b], i) + 0.5) / 2^24 with the splitmix64 finaliser (Gumbel-max = one
 *                    draw from softmax(logits / T)).  logits bf16 [B, V], row stride given.
 * ------------------------------------------------------------------------------------------- */
B2_API int b2_rms_norm(void* out, const void* x, void* residual, const void* weight, int T, int H,
                int64_t x_stride, int64_t out_stride, int64_t res_stride, float eps, void* stream);
B2_API int b2_silu_and_mul(void* out, const void* gate_up, int T, int I, int64_t in_stride, int64_t out_stride,
                    void* stream);
B2_API int b2_rope_and_cache(void* qkv, const int32_t* positions, const float* cos_sin, int max_pos,
                      void* k_cache, void* v_cache, const int64_t* slot_mapping, int T, int Hq, int Hkv,
                      int d, int bs, int64_t token_stride, int64_t cache_block_stride, int64_t cache_head_stride,
                      int kv_dtype, const float* k_scale, const float* v_scale, void* stream);
/* b2_rope_and_cache with write_k = 0: the rotated K rows go to the cache only, not back into qkv (an engine that
 * reads K only from the cache saves 2 of 26 KiB per token); q is always rotated in place.  write_k = 1 is
 * b2_rope_and_cache. */
B2_API int b2_rope_and_cache_ex(void* qkv, const int32_t* positions, const float* cos_sin, int max_pos,
                      void* k_cache, void* v_cache, const int64_t* slot_mapping, int T, int Hq, int Hkv,
                      int d, int bs, int64_t token_stride, int64_t cache_block_stride, int64_t cache_head_stride,
                      int kv_dtype, const float* k_scale, const float* v_scale, int write_k, void* stream);
/* The projections of a DECODE step (VLLMConfig.max_num_seqs rows, reference benchmark-vllm/models.py:9):
 * out[M, N] = x[M, K] . W[N, K]^T, bf16 in, fp32 accumulate, one rounding.  M <= 256, K % 64 == 0, N % 4 == 0; ldx / ldo
 * are row strides in elements.  A weight-streaming tcgen05 kernel (swap-AB, split-K over a thread-block cluster with the
 * partial sums reduced through distributed shared memory; csrc/gemm_tc5.cu) for the regime where a library GEMM leaves
 * most of the HBM bandwidth unused; prompt chunks (M > 256) stay on the library.  vLLM: the ColumnParallelLinear /
 * RowParallelLinear forward of a decode step. */
B2_API int b2_skinny_gemm(void* out, const void* x, const void* w, int M, int N, int K, int64_t ldx, int64_t ldo,
                          void* stream);
B2_API int b2_sample(int64_t* tokens, const void* logits, int B, int V, int64_t stride, const float* temperature,
              const int64_t* seeds, void* stream);
/* b2_sample with the OpenAI / vLLM sampling filters (top_k int32 [B], top_p fp32 [B]; either may be NULL): for rows
 * with temperature > 0 the draw is restricted to the candidate set {logit >= t}: t = the top_k-th largest logit
 * (ties kept; top_k <= 0 or >= V: no limit), then raised to the smallest set of top logits whose softmax(logits / T)
 * mass reaches top_p of the top-k mass (ties kept; top_p >= 1: no limit) -- vLLM's order, top-k first.  The draw is the
 * same Gumbel-max with the same hash, so top_k = 0, top_p = 1 is b2_sample bit for bit.  Sort-free: bisection over the
 * 16-bit key space of bf16 (16 passes over the L2-resident row per filter). */
B2_API int b2_sample_filtered(int64_t* tokens, const void* logits, int B, int V, int64_t stride,
                              const float* temperature, const int64_t* seeds, const int32_t* top_k, const float* top_p,
                              void* stream);

/* ---------------------------------------------------------------------------------------------
 * Block manager (host only): the page allocator that produces block_tables / seq_lens /
 * slot_mapping each step.  Pool size follows VLLMConfig.gpu_memory_utilization and block_size
 * (models.py:8,10); per-sequence capacity follows max_model_len (models.py:15).
 * A handle is single-writer (the scheduler thread).  All output pointers are HOST memory.
 * ------------------------------------------------------------------------------------------- */
typedef struct b2_block_manager b2_block_manager;

B2_API b2_block_manager* b2_bm_create(int64_t num_blocks, int bs);          /* NULL on bad args */
B2_API void    b2_bm_destroy(b2_block_manager* bm);
B2_API int64_t b2_bm_num_free(const b2_block_manager* bm);
B2_API int64_t b2_bm_num_seqs(const b2_block_manager* bm);
B2_API int64_t b2_bm_seq_len(const b2_block_manager* bm, int64_t seq_id);   /* < 0: B2_ENOTFOUND */
/* can `n_tokens` more tokens be appended to seq_id (or a new sequence if it is unknown)? */
B2_API int     b2_bm_can_append(const b2_block_manager* bm, int64_t seq_id, int64_t n_tokens);
/* register seq_id with room for n_tokens tokens, length 0 */
B2_API int     b2_bm_alloc_seq(b2_block_manager* bm, int64_t seq_id, int64_t n_tokens);
/* grow seq_id by n_new tokens; writes their n_new slots (block*bs + offset) to slots_out */
B2_API int     b2_bm_append(b2_block_manager* bm, int64_t seq_id, int64_t n_new, int64_t* slots_out);
B2_API int     b2_bm_free_seq(b2_block_manager* bm, int64_t seq_id);
/* block_tables_out[B][max_nb] (unused entries 0) and seq_lens_out[B] for the given sequences */
B2_API int     b2_bm_fill_tables(const b2_block_manager* bm, const int64_t* seq_ids, int B,
                          int32_t* block_tables_out, int max_nb, int32_t* seq_lens_out);
/* one decode step for a batch: append one token to every sequence, then fill the three arrays */
B2_API int     b2_bm_plan_decode(b2_block_manager* bm, const int64_t* seq_ids, int B,
                          int32_t* block_tables_out, int max_nb, int32_t* seq_lens_out,
                          int64_t* slot_mapping_out);

/* ---------------------------------------------------------------------------------------------
 * Continuous-batching scheduler (host only): the caller of the hot path.
 * It owns a block manager and turns a stream of requests into per-step plans under the limits of the
 * reference's launch contract: max_num_seqs (models.py:9), max_model_len (:15), block_size (:10), the block
 * pool sized from gpu_memory_utilization (:8), plus a per-step token budget for chunked prefill.
 *
 * Policy (vLLM-style, decode first): every running sequence decodes one token per step; when the pool cannot
 * hold the step, the most recently admitted running sequence is preempted (its blocks are freed and it goes back
 * to the head of the waiting queue to be recomputed, prompt + tokens generated so far); then waiting requests are
 * admitted first-come-first-served and prefilled in chunks within the remaining token budget.
 *
 * A request holds P prompt tokens and may generate up to max_new_tokens.  After its last prefill chunk the model
 * samples output token #1; each decode step writes the previously sampled token into the cache (one slot),
 * attends over seq_len tokens and samples the next one.  The request finishes by itself when it has sampled
 * max_new_tokens or reached max_model_len; b2_sched_finish ends it earlier (EOS / abort).
 * b2_sched_step assumes the previous plan was executed.  All plan pointers are host memory owned by the
 * scheduler, valid until the next b2_sched_step / b2_sched_destroy.  Single-writer, like the block manager.
 * ------------------------------------------------------------------------------------------- */
typedef struct b2_scheduler b2_scheduler;

typedef struct b2_step_plan {
    /* decode part: one new token for each of n_decode running sequences */
    int            n_decode;
    int            max_nb;              /* row length of the block tables */
    const int64_t* decode_ids;          /* [n_decode] request ids, batch row order */
    const int32_t* block_tables;        /* [n_decode, max_nb] */
    const int32_t* seq_lens;            /* [n_decode] tokens that attend, including the one written this step */
    const int64_t* slot_mapping;        /* [n_decode] slot of the token written this step */
    /* prefill part: n_prefill chunks, n_prefill_tokens tokens in total, chunk after chunk */
    int            n_prefill;
    int            n_prefill_tokens;
    const int64_t* prefill_ids;         /* [n_prefill] */
    const int32_t* prefill_start;       /* [n_prefill] position of the chunk's first token in its sequence */
    const int32_t* prefill_len;         /* [n_prefill] tokens in the chunk */
    const uint8_t* prefill_last;        /* [n_prefill] 1 = the prompt is complete after this chunk: sample token #1 */
    const int64_t* prefill_slots;       /* [n_prefill_tokens] */
    const int32_t* prefill_block_tables;/* [n_prefill, max_nb] tables covering start + len tokens */
    /* bookkeeping the engine must mirror */
    int            n_preempted;
    const int64_t* preempted_ids;       /* recompute later: their KV is gone */
    int            n_finished;
    const int64_t* finished_ids;        /* finished by length after the PREVIOUS plan; their blocks are free again */
} b2_step_plan;

B2_API b2_scheduler* b2_sched_create(int64_t num_blocks, int bs, int max_num_seqs, int max_model_len,
                              int max_num_batched_tokens);                         /* NULL on bad args */
B2_API void    b2_sched_destroy(b2_scheduler* s);
B2_API int     b2_sched_add_request(b2_scheduler* s, int64_t req_id, int prompt_len, int max_new_tokens);
B2_API int     b2_sched_finish(b2_scheduler* s, int64_t req_id);                          /* EOS / abort */
B2_API int     b2_sched_step(b2_scheduler* s, b2_step_plan* plan);
B2_API int64_t b2_sched_num_waiting(const b2_scheduler* s);
B2_API int64_t b2_sched_num_running(const b2_scheduler* s);
B2_API int64_t b2_sched_num_free_blocks(const b2_scheduler* s);
B2_API int64_t b2_sched_generated(const b2_scheduler* s, int64_t req_id);   /* tokens sampled so far; < 0 unknown */

/* ---------------------------------------------------------------------------------------------
 * Tensor-parallel exchange over NVLink peer memory: what `--tensor-parallel-size N` (VLLMConfig.tensor_parallel_size,
 * reference benchmark-vllm/models.py:11 -> vllm_templates.py:55,226; resources x tp :459-461) adds to a decoder
 * layer.  Heads and MLP columns are sharded over tp GPUs (one process per GPU); o_proj / down_proj leave a PARTIAL
 * [T, hidden] sum on every rank.  vLLM runs an all-reduce (NCCL or its custom all-reduce -- which the reference
 * switches off with --disable-custom-all-reduce, vllm_templates.py:62) followed by fused_add_rms_norm; here one kernel
 * does both over peer-mapped buffers (csrc/comm.cu).
 *
 * b2_comm_alloc / open / close / free   a cudaMalloc region plus its 64-byte CUDA IPC handle (the ONE place the
 *     library allocates device memory: the region must be exportable).  Every rank allocates one region of
 *     b2_comm_region_bytes(...) bytes, exchanges handles out of band, opens the peers' handles.
 * b2_comm_create   handle over the tp regions (regions[rank] = this rank's own pointer), tp in {1, 2, 4, 8}.
 *     max_rows x H = the largest partial sum that takes the one-shot path; gather_rows x gather_cols = the [T, V]
 *     logits buffer.
 * b2_comm_input(comm, slot)   device pointer of this rank's partial-sum buffer `slot` (0 after o_proj, 1 after
 *     down_proj), bf16 [max_rows, H] contiguous: the GEMM writes its output there.
 * b2_comm_allreduce_rms_norm   h = bf16(sum_r partial_r (+ residual)) -> residual (when given);
 *     out = weight ? bf16(weight * h * rsqrt(mean(h^2) + eps)) : h.  The sum runs in rank order in fp32, so every
 *     rank computes the same bits.  Every rank must call it with the same (slot, T) in the same order, and
 *     alternate the slots (that is what makes buffer reuse safe, see comm.cu).
 * b2_comm_push_allreduce_rms_norm   the same result by the PUSH protocol (the default in b200attn.tp): `partial`
 *     (bf16 [T, H], row stride given) is ordinary device memory; every rank stores its rows into landing buffers on
 *     every peer, where a slot still holding the sentinel (bf16 -0.0 in all lanes) has not arrived -- no barrier, one
 *     NVLink crossing on the critical path.  A partial sum equal to -0.0 travels as +0.0.
 * b2_comm_allgather_cols   gathered[t][rank*cols + c] = local[t][c] on every rank (b2_comm_gathered: bf16
 *     [gather_rows, tp*cols]); the vocab-parallel lm_head.
 * Both only enqueue on `stream`; they are CUDA-graph capturable (the barrier epoch is device state).  A peer that
 * never arrives traps the waiting kernel after ~10 s (B2_ECUDA at the next sync) instead of hanging the GPU.
 * ------------------------------------------------------------------------------------------- */
#define B2_COMM_HANDLE_BYTES 64
#define B2_COMM_MAX_RANKS 8
B2_API int b2_comm_alloc(size_t bytes, void** ptr, void* handle /* B2_COMM_HANDLE_BYTES out */);
B2_API int b2_comm_open(const void* handle, void** ptr);
B2_API int b2_comm_close(void* ptr);
B2_API int b2_comm_free(void* ptr);
B2_API size_t b2_comm_region_bytes(int tp, int max_rows, int H, int gather_rows, int gather_cols);
B2_API void* b2_comm_create(int rank, int tp, void* const* regions, size_t region_bytes, int max_rows, int H,
                            int gather_rows, int gather_cols);
B2_API void b2_comm_destroy(void* comm);
B2_API void* b2_comm_input(void* comm, int slot);
B2_API void* b2_comm_gathered(void* comm);
/* diagnostic (B2_COMM_DEBUG=1 at b2_comm_create): globaltimer ns of CTA 0 in the last push exchange:
 * kernel start, own rows pushed, first vector complete from every peer, row done.  Synchronises the device. */
B2_API int b2_comm_debug_stamps(void* comm, uint64_t* out4);
B2_API int b2_comm_allreduce_rms_norm(void* comm, int slot, void* out, void* residual, const void* weight, int T,
                                      int64_t out_stride, int64_t res_stride, float eps, void* stream);
B2_API int b2_comm_push_allreduce_rms_norm(void* comm, const void* partial, int64_t partial_stride, void* out,
                                           void* residual, const void* weight, int T, int64_t out_stride,
                                           int64_t res_stride, float eps, void* stream);
B2_API int b2_comm_allgather_cols(void* comm, const void* local, int64_t local_stride, int T, int cols, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
