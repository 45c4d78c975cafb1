/*
 * b200attn.h -- C ABI of the B200 (sm_100a) paged-KV attention library.
 *
 * This is the drop-in boundary of the project: a plain `extern "C"` surface (raw
 * device pointers, sizes, strides, a cudaStream_t) with no torch types.  The
 * Python `B200AttnBackend` (sglang_custom_b200/backend.py) binds it with ctypes;
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Every entry point below replaces one piece of the reference's attention hot
 * path (paths relative to the reference root, python/sglang/srt/...):
 *
 *   b200attn_kv_write            mem_cache/memory_pool.py:420-458  MHATokenToKVPool.set_kv_buffer
 *   b200attn_decode_plan         layers/attention/triton_backend.py:160-226,875-924 (init_forward_metadata,
 *                                get_num_kv_splits_triton) -- device-only, CUDA-graph safe
 *   b200attn_decode              layers/attention/torch_native_backend.py:112-180,226-267 (oracle),
 *                                triton_ops/decode_attention.py:677-728, sgl-kernel/csrc/cpu/decode.cpp:1375
 *   b200attn_decode_partial      stage 1 of the above (triton_ops/decode_attention.py:555-609)
 *   b200attn_decode_merge        triton_ops/decode_attention.py:492-552,611-674, decode.cpp:812-860
 *   b200attn_merge_state         layers/attention/merge_state.py:26-46,
 *                                sgl-kernel/csrc/attention/merge_attn_states.cu:32-106
 *   b200attn_extend              layers/attention/torch_native_backend.py:27-110,182-224 (oracle),
 *                                triton_ops/extend_attention.py:306-438, sgl-kernel/csrc/cpu/extend.cpp:579
 *   b200attn_decode_multiq       cached-prefix part of TARGET_VERIFY batches (triton_backend.py:227-264)
 *   b200attn_decode_static_ws    (no counterpart in the reference) decode for hosts that cannot allocate a workspace
 *   b200attn_decode_paged, b200attn_decode_multiq_paged, b200attn_extend_paged, b200attn_extend_split
 *                                the same three for paged pools (mem_cache/allocator.py:444-447 slot layout)
 *   b200attn_decode_plan_window  layers/attention/triton_backend.py:927-955 (update_sliding_window_buffer)
 *   b200attn_extend_masked       triton_ops/extend_attention.py:158-175,236-247 (sliding window, custom mask)
 *   b200attn_check_loc           mem_cache/memory_pool.py:457-458 (`buf[loc] = x`: which duplicates are legal)
 *   b200attn_flatten_kv_indices  layers/attention/utils.py:5-41 (create_flashinfer_kv_indices_triton)
 *   b200attn_gather_kv           torch_native_backend.py:92-94,155-157 (k_cache[per_req_tokens] page gather)
 *   b200attn_move_kv             mem_cache/memory_pool.py:460-468,1150-1184 (move_kv_cache, all layers in one launch)
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless the name ends in `_host`.
 *   - The library never allocates or frees device memory and never synchronises
 *     the host with the device; every kernel is enqueued on `stream`.
 *     All calls are CUDA-graph-capture safe (b200attn_decode_static_ws states its own rules and makes no such claim).
 *   - Return value: 0 on success, a negative B200ATTN_ERR_* code otherwise; the
 *     message is available (per thread) from b200attn_last_error().
 *   - Strides are in ELEMENTS of the tensor's own dtype unless stated.
 *   - Index tensors `req_pool_indices`/`seq_lens` may be int32 (CUDA-graph
 *     buffers, cuda_graph_runner.py:273-276) or int64 (eager,
 *     schedule_batch.py:1125-1133): `idx_is_i64` says which.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_VERSION 220 /* 0.2.2: + extend_split (KV splits of the tcgen05 extend kernel); 0.2.1: + *_paged (TMA box loads) */

/* element types */
#define B200ATTN_BF16 0
#define B200ATTN_F16 1
#define B200ATTN_FP8_E4M3 2 /* torch.float8_e4m3fn, stored as uint8 in the pool */
#define B200ATTN_F32 3

/* error codes */
#define B200ATTN_OK 0
#define B200ATTN_ERR_INVALID (-1)     /* bad argument (shape, alignment, null) */
#define B200ATTN_ERR_UNSUPPORTED (-2) /* valid in the reference, not built here */
#define B200ATTN_ERR_CUDA (-3)        /* a CUDA runtime call failed */
#define B200ATTN_ERR_WORKSPACE (-4)   /* workspace / plan buffer too small */

/* how the FP8 scale is applied before the e4m3 cast (both are torch behaviours):
 *   DIV: bf16(x / s)      -- torch CPU `div_`, and CUDA `div_` by a 0-dim tensor
 *   RCP: bf16(x * (1/s))  -- torch CUDA `div_` by a python float (BinaryDivTrueKernel.cu) */
#define B200ATTN_SCALE_DIV 0
#define B200ATTN_SCALE_RCP 1

typedef void* b200attn_stream_t; /* cudaStream_t */

int b200attn_version(void);
const char* b200attn_last_error(void);

/* Device facts the host side needs for sizing (no allocation). */
int b200attn_device_info(int* sm_count, int* cc_major, int* cc_minor, int* max_smem_optin);

/* ---------------------------------------------------------------- KV write
 * pool[loc[t], h, :] = cast(k[t, h, :])  for t in [0,T), h in [0,Hkv)   (same for v)
 * Replaces MHATokenToKVPool.set_kv_buffer (memory_pool.py:420-458).
 *   k, v            [T, Hkv, D] in_dtype (bf16/f16); token stride k_tok_stride, head stride k_head_stride
 *   loc             [T] int64, values in [0, P)
 *   k_pool, v_pool  [P, Hkv, D] pool_dtype; token stride pool_tok_stride (elements), head stride D
 *   pool_dtype == in_dtype      : bit copy
 *   pool_dtype == FP8_E4M3      : x -> bf16(x / scale) -> e4m3 (RNE, torch overflow rule)
 *   k_scale/v_scale: if *_scale_dev != NULL it points at n_scales (1 or Hkv) fp32
 *   values on the device (a 0-dim Parameter, or a per-head vector); otherwise the
 *   host scalar *_scale_host is used; a host scalar of 0 means "no scaling" (None).
 *   writeback_scaled != 0 reproduces the reference's in-place `cache_k.div_(scale)`
 *   side effect on the caller's k/v tensors (memory_pool.py:437-440).
 * Duplicate `loc` entries: the reference is last-writer-wins on the CPU (`buf[loc] = x`,
 * memory_pool.py:457-458) and unspecified on CUDA (index_put_ without accumulate); here the
 * order is unspecified too.  SGLang only ever produces duplicates inside the dummy page
 * [0, page_size) (padded CUDA-graph rows, cuda_graph_runner.py:686-695), whose content is
 * never read unmasked; b200attn_check_loc below detects any other duplicate. */
int b200attn_kv_write(const void* k, const void* v, int64_t k_tok_stride, int64_t k_head_stride,
                      int64_t v_tok_stride, int64_t v_head_stride, const int64_t* loc, int T, int Hkv,
                      int D, void* k_pool, void* v_pool, int64_t pool_tok_stride, int in_dtype,
                      int pool_dtype, float k_scale_host, float v_scale_host, const float* k_scale_dev,
                      const float* v_scale_dev, int n_scales, int scale_mode, int writeback_scaled,
                      b200attn_stream_t stream);

/* ---------------------------------------------------------------- fused rotary embedding + KV write
 * The two steps the reference runs right before attention, in one launch:
 *   q, k = rotary_emb(positions, q, k)      layers/rotary_embedding.py:138-166 (forward_native), :218-244 (forward_cuda ->
 *                                           sgl-kernel apply_rope_with_cos_sin_cache_inplace, csrc/elementwise/rope.cu:22-89)
 *   set_kv_buffer(layer, loc, k, v)         mem_cache/memory_pool.py:420-458
 * q [T,Hq,D] and k [T,Hkv,D] are rotated IN PLACE (head stride D, token strides given); the rotated k and v are cast
 * and scattered into the pool with exactly the rules of b200attn_kv_write.  loc == NULL: rotate only, no pool write.
 *   positions      [T] int64 or int32
 *   cos_sin_cache  [max_position, rotary_dim] = cos | sin halves (rotary_embedding.py:125-136); fp32 (what the
 *                  reference keeps on CUDA) or the tensor dtype (what it keeps on CPU, :101-102)
 *   is_neox        != 0: x1 | x2 are the two halves of the rotary part; 0: GPT-J even / odd interleave (:57-67)
 *   math_mode      ROPE_FP32: fp32 products, one rounding (the CUDA op); ROPE_NATIVE: forward_native's op-by-op
 *                  rounding in the tensor dtype -- bit-exact with the reference's CPU path
 * D and rotary_dim must be multiples of 16. */
#define B200ATTN_ROPE_FP32 0
#define B200ATTN_ROPE_NATIVE 1
int b200attn_rope_kv_write(void* q, int64_t q_tok_stride, void* k, int64_t k_tok_stride, const void* v,
                           int64_t v_tok_stride, const void* positions, int pos_is_i64,
                           const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox, int T,
                           int Hq, int Hkv, int D, int dtype, const int64_t* loc, void* k_pool, void* v_pool,
                           int64_t pool_tok_stride, int pool_dtype, float k_scale_host, float v_scale_host,
                           const float* k_scale_dev, const float* v_scale_dev, int n_scales, int scale_mode,
                           int math_mode, b200attn_stream_t stream);

/* ---------------------------------------------------------------- decode
 * Split-KV plan, computed on the device from seq_lens (no host sync, graph safe).
 *
 * The decode kernel is persistent: n_ctas CTAs (a multiple of the SM count).  Every
 * request's KV range is measured in 16-token units; the flattened unit space of the
 * whole batch is cut into n_ctas equal shares of W = max(min_units_per_cta,
 * ceil(total_units / n_ctas)) units, so ragged batches stay balanced.  A request
 * that falls entirely inside one share is finished by that CTA; a request cut
 * across shares produces one fp32 partial per share, combined by the LSE merge.
 *
 * plan layout (int32): [0] total_units [1] W [2] bs [3] n_ctas [4] sliding_window_size (-1: none) [5..7] reserved,
 *   then cum[bs+1] (exclusive prefix sum of units), then first_req[n_ctas].
 */
size_t b200attn_decode_plan_bytes(int max_bs, int n_ctas);
int b200attn_decode_plan(const void* seq_lens, int idx_is_i64, int bs, int n_ctas,
                         int min_units_per_cta, int32_t* plan, size_t plan_bytes,
                         b200attn_stream_t stream);

/* Plan for sliding-window layers (layer.sliding_window_size > -1): request i contributes only its last
 * min(seq_lens[i], sliding_window_size + 1) tokens, i.e. what update_sliding_window_buffer
 * (triton_backend.py:927-955) flattens into window_kv_indices for TritonAttnBackend.forward_decode
 * (triton_backend.py:711-716).  The window is stored in plan[4]; b200attn_decode / _partial read it from the
 * plan, so a model that mixes local and global layers keeps two plan buffers and nothing else changes. */
int b200attn_decode_plan_window(const void* seq_lens, int idx_is_i64, int bs, int n_ctas,
                                int min_units_per_cta, int sliding_window_size, int32_t* plan,
                                size_t plan_bytes, b200attn_stream_t stream);

/* fp32 workspace for split partials: 2 * n_ctas * Hq * (Dv + 1) floats */
size_t b200attn_decode_workspace_bytes(int n_ctas, int Hq, int Dv);

/* Paged decode attention (one query token per request) + LSE merge of the splits.
 *   q        [bs, Hq, D] bf16/f16, token stride q_tok_stride (head stride D)
 *   out      [bs, Hq, D] same dtype as q, token stride out_tok_stride
 *   k_pool, v_pool [P, Hkv, D] kv_dtype (bf16/f16 == q dtype, or FP8_E4M3), token stride pool_tok_stride
 *   req_to_token   [max_reqs, table_stride] int32 : token slot of (request row, position)
 *   req_pool_indices [bs], seq_lens [bs]  (int32 or int64)
 *   out[i,h] = softmax(sm_scale * k_scale * q[i,h] . K[idx_i, h/G]^T) . V[idx_i, h/G] * v_scale,
 *   idx_i = req_to_token[req_pool_indices[i], 0:seq_lens[i]]   (torch_native_backend.py:147-177)
 *   logit_cap > 0 applies cap*tanh(x/cap) to the scaled logits (decode_attention.py:117-118).
 *   plan must have been produced by b200attn_decode_plan for the same seq_lens, bs and n_ctas.
 *   If lse_out != NULL, the natural-log LSE of every (request, head) is stored there (fp32 [bs,Hq]). */
int b200attn_decode(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                    const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                    const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                    const void* seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype,
                    int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale,
                    const int32_t* plan, int n_ctas, void* workspace, size_t workspace_bytes,
                    float* lse_out, b200attn_stream_t stream);

/* The two stages of b200attn_decode, separately launchable (the reference's Triton backend exposes its
 * stage 1 / stage 2 the same way: decode_attention.py:555-609 `_decode_att_m_fwd`, :611-674
 * `_decode_softmax_reducev_fwd`).  b200attn_decode == partial followed by merge on the same stream.
 * b200attn_decode_partial takes exactly the arguments of b200attn_decode. */
int b200attn_decode_partial(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                            const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                            const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                            const void* seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype,
                            int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale,
                            const int32_t* plan, int n_ctas, void* workspace, size_t workspace_bytes,
                            float* lse_out, b200attn_stream_t stream);
int b200attn_decode_merge(void* out, int64_t out_tok_stride, int bs, int Hq, int D, int q_dtype,
                          const int32_t* plan, int n_ctas, void* workspace, size_t workspace_bytes,
                          float* lse_out, b200attn_stream_t stream);

/* nq query tokens per request over the same seq_lens[i] keys, no mask among them: the cached-prefix part of a
 * speculative-decoding verify batch (TARGET_VERIFY: every draft token sees the whole prefix, triton_backend.py:227-264,
 * triton_ops/extend_attention.py stage 1 with SKIP_PREFIX_CUSTOM_MASK).  The nq * (Hq / Hkv) <= 16 rows of a kv head
 * share one pass over K/V.  Arguments as b200attn_decode except
 *   q   [bs, nq, Hq, D]: q_req_stride between requests, q_tok_stride between the nq tokens of a request (elements)
 *   out [bs, nq, Hq, D]: out_req_stride / out_tok_stride likewise
 *   lse_out (optional) fp32, element (i, t, h) at i * lse_req_stride + t * Hq + h
 *   workspace: b200attn_decode_workspace_bytes(n_ctas, Hq * nq, D)
 * The draft-token part (custom mask) is b200attn_extend_masked with
 * sliding_window_size = B200ATTN_WINDOW_NEW_TOKENS_ONLY, combined by b200attn_merge_state. */
int b200attn_decode_multiq(const void* q, int64_t q_req_stride, int64_t q_tok_stride, void* out,
                           int64_t out_req_stride, int64_t out_tok_stride, const void* k_pool, const void* v_pool,
                           int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride,
                           const void* req_pool_indices, const void* seq_lens, int idx_is_i64, int bs, int nq, int Hq,
                           int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale, float logit_cap, float k_scale,
                           float v_scale, const int32_t* plan, int n_ctas, void* workspace, size_t workspace_bytes,
                           float* lse_out, int64_t lse_req_stride, b200attn_stream_t stream);

/* Paged pools (page_size > 1).  SGLang's paged allocator hands out pages of page_size consecutive slots
 * (mem_cache/allocator.py:444-447: `out_pages[:, None] * page_size + arange(page_size)`) and position p of a request
 * sits at req_to_token[r, p] = page(p / page_size) * page_size + p % page_size, so the 16 rows of every 16-token unit
 * [16 j, 16 j + 16) of a request are 16 CONSECUTIVE pool rows when page_size % 16 == 0.  These entry points take the
 * page size; with page_size % 16 == 0 and D = 128 the producers read ONE table entry per unit and fetch the unit's K
 * and V rows as bulk-tensor (TMA, cp.async.bulk.tensor with 128-byte swizzle) boxes instead of 16 row gathers.  Any
 * other page_size runs the row-gather kernel of b200attn_decode; results are identical either way.
 *   stage: 0 = b200attn_decode (stage 1 + merge), 1 = b200attn_decode_partial (stage 1 only)
 * Contract: the consecutive-slot layout above (the caller asserts it by passing page_size > 1), and a plan from
 * b200attn_decode_plan (a sliding-window plan shifts requests off their page boundary: pass page_size = 1 there). */
int b200attn_decode_paged(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                          const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                          const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                          const void* seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype,
                          int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale,
                          const int32_t* plan, int n_ctas, void* workspace, size_t workspace_bytes,
                          float* lse_out, int page_size, int stage, b200attn_stream_t stream);
int b200attn_decode_multiq_paged(const void* q, int64_t q_req_stride, int64_t q_tok_stride, void* out,
                                 int64_t out_req_stride, int64_t out_tok_stride, const void* k_pool,
                                 const void* v_pool, int64_t pool_tok_stride, const int32_t* req_to_token,
                                 int64_t table_stride, const void* req_pool_indices, const void* seq_lens,
                                 int idx_is_i64, int bs, int nq, int Hq, int Hkv, int D, int q_dtype, int kv_dtype,
                                 float sm_scale, float logit_cap, float k_scale, float v_scale, const int32_t* plan,
                                 int n_ctas, void* workspace, size_t workspace_bytes, float* lse_out,
                                 int64_t lse_req_stride, int page_size, b200attn_stream_t stream);

/* Decode with the split-KV workspace built into the library -- this one has NO counterpart in the reference.
 *
 * For hosts that can hand over the tensors of a forward pass but cannot produce one more device buffer: a cgo / JNI /
 * ctypes binding that does not link the CUDA runtime, or a plugin inside a server (SGLang included) that has already
 * given all free HBM to the KV pool and must not cudaMalloc in its step loop (the call can fail for lack of memory, and
 * it stalls every stream on the device while it runs).  Everything else is b200attn_decode_paged with stage = 0:
 * same arguments, same kernels, same results bit for bit; pass page_size = 1 for b200attn_decode's behaviour.
 *
 *   - The fp32 partials go into a B200ATTN_DECODE_STATIC_WS_BYTES array that is part of the library's device image.
 *     Each GPU the library's kernels are loaded on holds that much memory until the process exits.
 *   - Limit: b200attn_decode_workspace_bytes(n_ctas, Hq, D) must not exceed b200attn_decode_static_ws_capacity()
 *     (10 MiB: 148 CTAs x 64 query heads x D 128 fits); larger problems return B200ATTN_ERR_WORKSPACE and the message
 *     in b200attn_last_error() says how much was asked for.  Use fewer CTAs, or b200attn_decode with your own buffer.
 *   - The call allocates nothing, on the host's CUDA heap or anywhere else, and frees nothing.
 *   - The FIRST call on each device resolves the array's device address (cudaGetSymbolAddress, with that device
 *     current -- the device `stream` belongs to must be the current one, as for every entry point here).  Make that
 *     first call during start-up, before any stream capture begins on that device.
 *   - All calls on a device share the one array and find the previous call's partials in it.  Nothing needs to be
 *     zero at the start, so nothing is cleared; but two calls that can overlap on the GPU (different streams without
 *     an event between them, or two host threads on two streams) overwrite each other's partials.  Calls on ONE
 *     stream are ordered and safe.  b200attn_decode with separate workspaces is the way to run decodes concurrently. */
#define B200ATTN_DECODE_STATIC_WS_BYTES (10u << 20)
size_t b200attn_decode_static_ws_capacity(void);
int b200attn_decode_static_ws(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                              const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                              const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                              const void* seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype,
                              int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale,
                              const int32_t* plan, int n_ctas, float* lse_out, int page_size,
                              b200attn_stream_t stream);

/* ---------------------------------------------------------------- 2-way LSE merge
 * out = (e^{la-m} a + e^{lb-m} b) / (e^{la-m} + e^{lb-m}),  lse = log(sum) + m, m = max(la, lb);
 * +inf LSE is treated as -inf (merge_attn_states.cu:47-50).
 *   a, b, out [T, H, D] dtype (bf16/f16/f32), contiguous; lse_* [T, H] fp32; out_lse may be NULL */
int b200attn_merge_state(const void* a, const float* lse_a, const void* b, const float* lse_b,
                         void* out, float* out_lse, int T, int H, int D, int dtype,
                         b200attn_stream_t stream);

/* ---------------------------------------------------------------- extend / prefill
 * Causal attention of E_i new query tokens per request over P_i cached prefix
 * tokens plus the E_i new tokens, all read from the paged pool through
 * req_to_token (the new tokens must already be written: b200attn_kv_write first,
 * exactly as torch_native_backend.py:196-199 does).
 *   q   [T, Hq, D], out [T, Hq, D]  (T = sum E_i), token strides given
 *   extend_seq_lens [bs] int32 (E_i), extend_prefix_lens [bs] int32 (P_i),
 *   extend_start_loc [bs] int32 (row of request i's first query in q/out)
 *   seq_lens[i] must equal P_i + E_i.
 *   key j of request i is visible to its query t iff j <= P_i + t (causal != 0), else all keys.
 *   max_extend_len: upper bound of E_i known on the host (grid sizing, no device sync).
 *   lse_out: optional fp32 [T, Hq] natural-log LSE of every (query, head), or NULL. */
int b200attn_extend(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                    const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                    const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                    const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens,
                    const int32_t* extend_prefix_lens, const int32_t* extend_start_loc, int bs,
                    int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,
                    float logit_cap, float k_scale, float v_scale, int causal, int impl, float* lse_out,
                    b200attn_stream_t stream);

/* Extend with the two extra predicates of TritonAttnBackend.forward_extend (triton_backend.py:658-684 ->
 * triton_ops/extend_attention.py:42-303): a sliding window and a speculative-decoding custom mask
 * (TARGET_VERIFY batches, triton_backend.py:227-264).  All other arguments as b200attn_extend.
 *   sliding_window_size  -1: none.  B200ATTN_WINDOW_NEW_TOKENS_ONLY (-2): no cached-prefix key at all -- the new-token
 *                        part of a verify batch whose prefix part comes from b200attn_decode_multiq.  W >= 0: of the P_i cached-prefix keys only the last Wp = min(P_i, W + 1) are read
 *                        (window_kv_indices, triton_backend.py:299-309,927-955) and, for W > 0, the key at offset jw
 *                        in [0, Wp) of that range is visible to query t (0-based inside the request's new tokens)
 *                        iff t <= jw + W (extend_attention.py:168-174).  New-token keys keep the plain causal rule:
 *                        the reference applies no window there, and neither does this kernel.
 *   custom_mask          NULL: none.  uint8/bool, request i's block starts at mask_indptr[i] (int64 [bs+1], bytes)
 *                        and is [E_i, L_i] row-major with L_i = P_i + E_i.  A new-token key j >= P_i is visible to
 *                        query t iff mask[t, j] != 0 and (causal == 0 or j <= P_i + t); cached-prefix keys consult
 *                        the mask only when skip_prefix_custom_mask == 0 (the reference backend always passes the
 *                        default skip = True, extend_attention.py:322).  EAGLE tree masks are subsets of the causal
 *                        mask, for which this is exactly the reference's predicate.
 *   A query whose keys are all masked produces 0 (the reference's online softmax produces NaN there).
 * impl as in b200attn_extend: a custom mask is served by the register-MMA kernel only (verify batches are
 * num_draft_tokens rows per request); a sliding window alone also by the tcgen05 kernel (AUTO picks it for
 * max_extend_len >= 128, D == 128, causal, logit_cap == 0). */
int b200attn_extend_masked(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                           const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                           const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                           const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens,
                           const int32_t* extend_prefix_lens, const int32_t* extend_start_loc, int bs,
                           int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,
                           float logit_cap, float k_scale, float v_scale, int causal, int impl, int sliding_window_size,
                           const uint8_t* custom_mask, const int64_t* mask_indptr, int skip_prefix_custom_mask,
                           float* lse_out, b200attn_stream_t stream);

/* b200attn_extend_masked for a paged pool (contract as b200attn_decode_paged; sliding_window_size = -1 and
 * custom_mask = NULL give plain b200attn_extend).  With page_size % 16 == 0 and a 16-bit pool the tcgen05 kernel's
 * producer reads one table entry per 16-key page and fetches the page as two bulk-tensor (TMA) boxes of
 * 16 rows x 128 B whose 128-byte swizzle is the UMMA operand layout; FP8 pools keep the register-expanding
 * producers, the register-MMA kernel keeps its row gathers.  Results are identical to page_size = 1. */
int b200attn_extend_paged(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                           const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                           const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                           const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens,
                           const int32_t* extend_prefix_lens, const int32_t* extend_start_loc, int bs,
                           int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,
                           float logit_cap, float k_scale, float v_scale, int causal, int impl, int sliding_window_size,
                           const uint8_t* custom_mask, const int64_t* mask_indptr, int skip_prefix_custom_mask,
                           float* lse_out, int page_size, b200attn_stream_t stream);

/* b200attn_extend_paged with the K/V range of every (request, head pair, 128 q rows) item cut into kv_splits shares
 * that run as separate CTAs of the tcgen05 kernel; each writes a normalised partial + LSE into `workspace`, a second
 * kernel merges them into out / lse_out.  For batches whose items do not fill the GPU -- a chunked-prefill row over a
 * long radix prefix (triton_backend.py:632-685 with extend_prefix_lens >> extend_seq_lens), head-partitioned ranks
 * (configs/model_config.py:342-349) -- where one item per CTA leaves SMs idle for the length of the longest row.
 * Plain batches only: no sliding window, no custom mask; D == 128, logit_cap == 0 (the tcgen05 kernel); otherwise
 * B200ATTN_ERR_UNSUPPORTED.  T = total new tokens (rows of q / out).  kv_splits == 1 is b200attn_extend_paged.
 * Partials are stored in the q dtype, so results differ from the unsplit call by one extra rounding (inside the
 * parity bound, not bit-identical). */
size_t b200attn_extend_split_workspace_bytes(int T, int Hq, int D, int kv_splits);
int b200attn_extend_split(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                          const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                          const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
                          const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens,
                          const int32_t* extend_prefix_lens, const int32_t* extend_start_loc, int bs,
                          int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,
                          float k_scale, float v_scale, int causal, float* lse_out, int page_size, int T, int kv_splits,
                          void* workspace, size_t workspace_bytes, b200attn_stream_t stream);
#define B200ATTN_WINDOW_NEW_TOKENS_ONLY (-2)
#define B200ATTN_EXTEND_AUTO 0
#define B200ATTN_EXTEND_TCGEN05 1 /* tcgen05/TMEM kernel (D == 128, bf16/f16 KV) */
#define B200ATTN_EXTEND_MMA 2     /* register-MMA kernel (short / ragged rows, FP8 KV, D == 64) */
/* (3 was EXTEND_TCGEN05_STEP64, a 64-key-step variant measured slower at C3 -- 1105 vs 1164 TFLOP/s -- and removed
 *  in 0.2.0; the result is kept in DESIGN.md) */

/* Validation of a KV-write slot list (debug aid, O(T^2) -- not for the hot path): after the call
 *   flags[0] = number of pairs t < t' with loc[t] == loc[t'] >= dummy_rows  (a write race: the reference's
 *              last-writer-wins, memory_pool.py:457-458, is not reproduced for them)
 *   flags[1] = number of entries outside [0, pool_rows)
 * flags: 2 int32 on the device.  B200AttnBackend runs it before every KV write when B200ATTN_CHECK_LOC=1. */
int b200attn_check_loc(const int64_t* loc, int T, int64_t dummy_rows, int64_t pool_rows, int32_t* flags,
                       b200attn_stream_t stream);

/* ---------------------------------------------------------------- page-table utilities
 * kv_indices[kv_indptr[i] + j] = req_to_token[req_pool_indices[i], kv_start[i] + j], j < lens[i]
 * (layers/attention/utils.py:5-41).  kv_start may be NULL (0). lens/kv_start/req_pool_indices share idx_is_i64;
 * kv_indptr is int32. */
int b200attn_flatten_kv_indices(const int32_t* req_to_token, int64_t table_stride,
                                const void* req_pool_indices, const void* lens, const void* kv_start,
                                int idx_is_i64, const int32_t* kv_indptr, int32_t* kv_indices, int bs,
                                b200attn_stream_t stream);

/* buf[tgt_loc[i]] = buf[src_loc[i]] for i in [0,n), in EVERY buffer of the pool, one launch.
 * Replaces MHATokenToKVPool.move_kv_cache / copy_all_layer_kv_cache (mem_cache/memory_pool.py:460-468,1150-1184), which
 * EAGLE calls after a verify step with page_size > 1 and topk > 1 to pull the accepted tokens' K/V together
 * (speculative/eagle_utils.py:575-583).
 *   data_ptrs     [n_buffers] uint64, DEVICE array of buffer base addresses (pool.data_ptrs: all k_buffer then v_buffer)
 *   data_strides  [n_buffers] int64, DEVICE array of row sizes in BYTES (pool.data_strides)
 *   max_stride_bytes  host copy of max(data_strides) (sizes the grid; no device read-back)
 *   tgt_loc, src_loc  [n] int64 or int32 row numbers
 * In-place rule of the reference: every source row is read as it was BEFORE the call, also when it is another pair's
 * target.  Duplicate targets are unspecified (as there).  Rows / bases that are not 16-byte multiples take a byte path.
 * Up to 12 800 rows per call (shared-memory staging); more returns B200ATTN_ERR_UNSUPPORTED. */
int b200attn_move_kv(const uint64_t* data_ptrs, const int64_t* data_strides, int n_buffers,
                     int64_t max_stride_bytes, const void* tgt_loc, const void* src_loc, int idx_is_i64, int n,
                     b200attn_stream_t stream);

/* dst[j, :, :] = pool[indices[j], :, :]  -- the page gather `k_cache[per_req_tokens]`, bit exact.
 *   row_bytes = Hkv * D * sizeof(elem) (multiple of 16), pool_tok_stride_bytes likewise */
int b200attn_gather_kv(const void* pool, int64_t pool_tok_stride_bytes, const int32_t* indices,
                       int64_t n, int64_t row_bytes, void* dst, b200attn_stream_t stream);

/* ---------------------------------------------------------------- HiCache KV page transfer
 * Copies n_items items of item_bytes bytes between two KV pools, for layers [start_layer, start_layer + num_layers):
 *     dst(layer, dst_indices[i]) = src(layer, src_indices[i])        for K and (when src_v != NULL) V
 * Replaces sgl-kernel/csrc/kvcacheio/transfer.cu:61-430 (transfer_kv_per_layer, _per_layer_pf_lf, _all_layer,
 * _all_layer_lf_pf and their _mla forms -- pass src_v = dst_v = NULL for the single-buffer MLA forms).
 * Either side may be pinned (mapped) host memory: the kernel then reads / writes it directly over the host link.
 * Layouts (how the address of (layer, index) is formed, transfer.cu:25-59):
 *   LAYER_FIRST  base + layer * layout_dim + index * item_bytes        (layout_dim = bytes of one layer; a per-layer
 *                                                                       call passes the layer's own base, layer 0)
 *   PAGE_FIRST   base + index * layout_dim + layer * item_bytes        (layout_dim = bytes of one index across layers)
 *   LAYER_TABLE  ((uint8_t*)table[layer]) + index * item_bytes         (`base` is a DEVICE array of uint64 pointers)
 * src_indices / dst_indices: device int64 [n_items].  item_bytes must be a multiple of 8 (transfer.cu:142).
 * block_quota > 0 caps the grid at that many CTAs (the reference's knob to leave SMs to a concurrent forward);
 * 0 sizes the grid to the device. */
#define B200ATTN_LAYOUT_LAYER_FIRST 0
#define B200ATTN_LAYOUT_PAGE_FIRST 1
#define B200ATTN_LAYOUT_LAYER_TABLE 2
int b200attn_transfer_kv(const void* src_k, void* dst_k, const void* src_v, void* dst_v, int src_layout,
                         int dst_layout, int64_t src_layout_dim, int64_t dst_layout_dim,
                         const int64_t* src_indices, const int64_t* dst_indices, int64_t n_items,
                         int64_t item_bytes, int start_layer, int num_layers, int block_quota,
                         b200attn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
