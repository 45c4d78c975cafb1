// b200attn_extend: argument checks and the choice between the tcgen05 and the register-MMA kernel.
// Reference entry points this replaces: TritonAttnBackend.forward_extend
// (python/sglang/srt/layers/attention/triton_backend.py:632-685) -> extend_attention_fwd
// (triton_ops/extend_attention.py:306-438); oracle torch_native_backend.py:27-110,182-224.
#include <stdlib.h>

#include "extend.cuh"

using namespace b200attn;

static int tc_launch(const ExtendParams& p, bool q_bf16, int kvb, cudaStream_t st) {
  return extend_tc_launch(p, q_bf16, kvb, st);
}

static int extend_impl(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride, const void* k_pool,
                       const void* v_pool, int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride,
                       const void* req_pool_indices, const void* seq_lens, int idx_is_i64,
                       const int32_t* extend_seq_lens, const int32_t* extend_prefix_lens,
                       const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv, int D,
                       int q_dtype, int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale,
                       int causal, int impl, float* lse_out, int window, const uint8_t* custom_mask,
                       const int64_t* mask_indptr, int skip_prefix_mask, int page_size, b200attn_stream_t stream,
                       int kv_splits = 1, int64_t split_out_stride = 0, int64_t split_lse_stride = 0) {
  B2_REQUIRE(page_size >= 1, B200ATTN_ERR_INVALID, "extend: page_size=%d", page_size);
  B2_REQUIRE(bs >= 0 && Hq > 0 && Hkv > 0 && max_extend_len >= 0, B200ATTN_ERR_INVALID,
             "extend: bad sizes bs=%d Hq=%d Hkv=%d max_extend_len=%d", bs, Hq, Hkv, max_extend_len);
  if (bs == 0 || max_extend_len == 0) return B200ATTN_OK;
  B2_REQUIRE(q && out && k_pool && v_pool && req_to_token && req_pool_indices && seq_lens && extend_seq_lens &&
                 extend_prefix_lens && extend_start_loc,
             B200ATTN_ERR_INVALID, "extend: null pointer");
  B2_REQUIRE(Hq % Hkv == 0, B200ATTN_ERR_INVALID, "extend: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
  B2_REQUIRE(D == 64 || D == 128, B200ATTN_ERR_UNSUPPORTED, "extend: head dim %d (64/128 only)", D);
  B2_REQUIRE(q_dtype == B200ATTN_BF16 || q_dtype == B200ATTN_F16, B200ATTN_ERR_UNSUPPORTED, "extend: q dtype %d", q_dtype);
  B2_REQUIRE(kv_dtype == q_dtype || kv_dtype == B200ATTN_FP8_E4M3, B200ATTN_ERR_UNSUPPORTED,
             "extend: kv dtype %d with q dtype %d", kv_dtype, q_dtype);
  const int kvb = elem_size(kv_dtype);
  B2_REQUIRE(aligned16(q) && aligned16(out) && aligned16(k_pool) && aligned16(v_pool) && q_tok_stride % 8 == 0 &&
                 out_tok_stride % 8 == 0 && (pool_tok_stride * kvb) % 16 == 0,
             B200ATTN_ERR_INVALID, "extend: tensors must be 16-byte aligned");
  B2_REQUIRE(pool_tok_stride >= static_cast<int64_t>(Hkv) * D, B200ATTN_ERR_INVALID, "extend: pool_tok_stride < Hkv*D");
  B2_REQUIRE(k_scale > 0.f && v_scale > 0.f, B200ATTN_ERR_INVALID, "extend: k_scale/v_scale must be > 0");
  B2_REQUIRE(impl == B200ATTN_EXTEND_AUTO || impl == B200ATTN_EXTEND_TCGEN05 || impl == B200ATTN_EXTEND_MMA,
             B200ATTN_ERR_INVALID, "extend: impl %d", impl);

  ExtendParams p;
  p.q = q; p.out = out; p.k_pool = k_pool; p.v_pool = v_pool;
  p.req_to_token = req_to_token; p.req_pool_indices = req_pool_indices; p.seq_lens = seq_lens;
  p.extend_seq_lens = extend_seq_lens; p.extend_prefix_lens = extend_prefix_lens;
  p.extend_start_loc = extend_start_loc;
  p.lse_out = lse_out;
  p.q_tok_stride = q_tok_stride; p.out_tok_stride = out_tok_stride;
  p.pool_tok_stride = pool_tok_stride; p.table_stride = table_stride;
  p.idx_is_i64 = idx_is_i64;
  p.bs = bs; p.Hq = Hq; p.Hkv = Hkv; p.G = Hq / Hkv; p.max_extend_len = max_extend_len;
  p.causal = causal;
  p.scale_raw = sm_scale * k_scale;
  p.scale_log2 = sm_scale * k_scale * kLog2e;
  p.logit_cap = logit_cap;
  p.v_scale = v_scale;
  p.window = window; p.custom_mask = custom_mask; p.mask_indptr = mask_indptr; p.skip_prefix_mask = skip_prefix_mask;
  p.page_size = page_size;
  p.kv_splits = kv_splits; p.split_out_stride = split_out_stride; p.split_lse_stride = split_lse_stride;
  {
    // dev knobs, read once: the exp2 split of the tcgen05 kernel and the phase-timing buffer of its kProf variant
    static const int emu = [] { const char* e = getenv("B200ATTN_EXP_EMU"); return e ? atoi(e) : 0; }();
    p.exp_emu = emu;
    static const int order = [] { const char* e = getenv("B200ATTN_GRID_ORDER"); return e ? atoi(e) : 1; }();
    p.grid_order = order;
    static unsigned long long* const prof = [] { const char* e = getenv("B200ATTN_PROF_PTR"); return e ? reinterpret_cast<unsigned long long*>(strtoull(e, nullptr, 0)) : nullptr; }();
    p.prof = prof;
  }

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool q_bf16 = q_dtype == B200ATTN_BF16;
  const bool tc_ok = extend_tc_supported(p, D, kvb);
  if (window == B200ATTN_WINDOW_NEW_TOKENS_ONLY) {
    B2_REQUIRE(impl != B200ATTN_EXTEND_TCGEN05, B200ATTN_ERR_UNSUPPORTED, "extend: new-tokens-only runs on the register-MMA kernel");
    return extend_mma_masked_launch(p, D, kvb, q_bf16, st);
  }
  if (window >= 0 || custom_mask != nullptr) {
    // the custom mask lives in the register-MMA kernel only (verify batches are a few rows per request); the
    // sliding window is a per-row key range, which the tcgen05 kernel handles too
    B2_REQUIRE(custom_mask == nullptr || mask_indptr != nullptr, B200ATTN_ERR_INVALID,
               "extend: custom_mask without mask_indptr");
    const bool tc_win = tc_ok && custom_mask == nullptr && causal != 0;
    if (impl == B200ATTN_EXTEND_TCGEN05) {
      B2_REQUIRE(tc_win, B200ATTN_ERR_UNSUPPORTED,
                 "extend: the tcgen05 kernel takes a sliding window (D == 128, causal, logit_cap == 0) but no custom mask");
      return tc_launch(p, q_bf16, kvb, st);
    }
    if (impl == B200ATTN_EXTEND_AUTO && tc_win && max_extend_len >= 128) return tc_launch(p, q_bf16, kvb, st);
    return extend_mma_masked_launch(p, D, kvb, q_bf16, st);
  }
  if (impl == B200ATTN_EXTEND_TCGEN05) {
    B2_REQUIRE(tc_ok, B200ATTN_ERR_UNSUPPORTED,
               "extend: the tcgen05 kernel needs D == 128 and logit_cap == 0 (D=%d kv bytes=%d cap=%g)", D,
               kvb, static_cast<double>(logit_cap));
    return tc_launch(p, q_bf16, kvb, st);
  }
  // AUTO: dense prefill tiles go to the tensor-memory kernel; short rows (decode rows of a MIXED
  // batch, tiny chunks) are not worth a 128-row tile.
  if (impl == B200ATTN_EXTEND_AUTO && tc_ok && max_extend_len >= 128) return tc_launch(p, q_bf16, kvb, st);
  return extend_mma_launch(p, D, kvb, q_bf16, st);
}

extern "C" B2_EXPORT int b200attn_extend(const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                                         const void* k_pool, const void* v_pool, int64_t pool_tok_stride,
                                         const int32_t* req_to_token, int64_t table_stride,
                                         const void* req_pool_indices, const void* seq_lens, int idx_is_i64,
                                         const int32_t* extend_seq_lens, const int32_t* extend_prefix_lens,
                                         const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq,
                                         int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale, float logit_cap,
                                         float k_scale, float v_scale, int causal, int impl, float* lse_out,
                                         b200attn_stream_t stream) {
  return extend_impl(q, q_tok_stride, out, out_tok_stride, k_pool, v_pool, pool_tok_stride, req_to_token,
                     table_stride, req_pool_indices, seq_lens, idx_is_i64, extend_seq_lens, extend_prefix_lens,
                     extend_start_loc, bs, max_extend_len, Hq, Hkv, D, q_dtype, kv_dtype, sm_scale, logit_cap,
                     k_scale, v_scale, causal, impl, lse_out, -1, nullptr, nullptr, 1, 1, stream);
}

// Sliding-window layers and speculative-decoding verify batches: TritonAttnBackend.forward_extend with
// window_kv_indices / custom_mask (triton_backend.py:227-264,299-309,658-684; kernel predicate
// triton_ops/extend_attention.py:158-175,236-247).
extern "C" B2_EXPORT int b200attn_extend_masked(
    const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride, const void* k_pool, const void* v_pool,
    int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
    const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens, const int32_t* extend_prefix_lens,
    const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype,
    float sm_scale, float logit_cap, float k_scale, float v_scale, int causal, int impl, int sliding_window_size,
    const uint8_t* custom_mask, const int64_t* mask_indptr, int skip_prefix_custom_mask, float* lse_out,
    b200attn_stream_t stream) {
  B2_REQUIRE(sliding_window_size >= B200ATTN_WINDOW_NEW_TOKENS_ONLY, B200ATTN_ERR_INVALID,
             "extend_masked: sliding_window_size=%d", sliding_window_size);
  return extend_impl(q, q_tok_stride, out, out_tok_stride, k_pool, v_pool, pool_tok_stride, req_to_token,
                     table_stride, req_pool_indices, seq_lens, idx_is_i64, extend_seq_lens, extend_prefix_lens,
                     extend_start_loc, bs, max_extend_len, Hq, Hkv, D, q_dtype, kv_dtype, sm_scale, logit_cap,
                     k_scale, v_scale, causal, impl, lse_out, sliding_window_size, custom_mask,
                     mask_indptr, skip_prefix_custom_mask, 1, stream);
}

// b200attn_extend_masked for a paged pool (slot layout of mem_cache/allocator.py:444-447, see b200attn_decode_paged):
// with page_size % 16 == 0 and a 16-bit pool the tcgen05 kernel's K/V tiles arrive as bulk-tensor (TMA) boxes.
extern "C" B2_EXPORT int b200attn_extend_paged(
    const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride, const void* k_pool, const void* v_pool,
    int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
    const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens, const int32_t* extend_prefix_lens,
    const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype,
    float sm_scale, float logit_cap, float k_scale, float v_scale, int causal, int impl, int sliding_window_size,
    const uint8_t* custom_mask, const int64_t* mask_indptr, int skip_prefix_custom_mask, float* lse_out,
    int page_size, b200attn_stream_t stream) {
  B2_REQUIRE(sliding_window_size >= B200ATTN_WINDOW_NEW_TOKENS_ONLY, B200ATTN_ERR_INVALID,
             "extend_paged: sliding_window_size=%d", sliding_window_size);
  return extend_impl(q, q_tok_stride, out, out_tok_stride, k_pool, v_pool, pool_tok_stride, req_to_token,
                     table_stride, req_pool_indices, seq_lens, idx_is_i64, extend_seq_lens, extend_prefix_lens,
                     extend_start_loc, bs, max_extend_len, Hq, Hkv, D, q_dtype, kv_dtype, sm_scale, logit_cap,
                     k_scale, v_scale, causal, impl, lse_out, sliding_window_size, custom_mask,
                     mask_indptr, skip_prefix_custom_mask, page_size, stream);
}

// KV splits of the tcgen05 kernel + n-way LSE merge (see include/b200attn.h)
extern "C" B2_EXPORT size_t b200attn_extend_split_workspace_bytes(int T, int Hq, int D, int kv_splits) {
  if (T <= 0 || Hq <= 0 || D <= 0 || kv_splits <= 1) return 0;
  const size_t rows = static_cast<size_t>(T) * Hq;
  return static_cast<size_t>(kv_splits) * rows * (static_cast<size_t>(D) * 2 + sizeof(float));
}

extern "C" B2_EXPORT int b200attn_extend_split(
    const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride, const void* k_pool, const void* v_pool,
    int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
    const void* seq_lens, int idx_is_i64, const int32_t* extend_seq_lens, const int32_t* extend_prefix_lens,
    const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv, int D, int q_dtype, int kv_dtype,
    float sm_scale, float k_scale, float v_scale, int causal, float* lse_out, int page_size, int T, int kv_splits,
    void* workspace, size_t workspace_bytes, b200attn_stream_t stream) {
  B2_REQUIRE(kv_splits >= 1 && kv_splits <= 64, B200ATTN_ERR_INVALID, "extend_split: kv_splits=%d", kv_splits);
  if (kv_splits == 1)
    return extend_impl(q, q_tok_stride, out, out_tok_stride, k_pool, v_pool, pool_tok_stride, req_to_token,
                       table_stride, req_pool_indices, seq_lens, idx_is_i64, extend_seq_lens, extend_prefix_lens,
                       extend_start_loc, bs, max_extend_len, Hq, Hkv, D, q_dtype, kv_dtype, sm_scale, 0.f, k_scale,
                       v_scale, causal, B200ATTN_EXTEND_AUTO, lse_out, -1, nullptr, nullptr, 1, page_size, stream);
  if (bs == 0 || max_extend_len == 0 || T == 0) return B200ATTN_OK;
  B2_REQUIRE(T > 0 && D == 128, B200ATTN_ERR_UNSUPPORTED, "extend_split: T=%d D=%d (the tcgen05 kernel needs D == 128)", T, D);
  B2_REQUIRE(workspace != nullptr && aligned16(workspace) &&
                 workspace_bytes >= b200attn_extend_split_workspace_bytes(T, Hq, D, kv_splits),
             B200ATTN_ERR_WORKSPACE, "extend_split: workspace %zu < %zu bytes", workspace_bytes,
             b200attn_extend_split_workspace_bytes(T, Hq, D, kv_splits));
  const int64_t so = static_cast<int64_t>(T) * Hq * D, sl = static_cast<int64_t>(T) * Hq;
  float* ws_lse = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + static_cast<size_t>(kv_splits) * so * 2);
  // stage 1: the partials, written where `out` / `lse_out` would go (token stride Hq * D)
  const int rc = extend_impl(q, q_tok_stride, workspace, static_cast<int64_t>(Hq) * D, k_pool, v_pool, pool_tok_stride,
                             req_to_token, table_stride, req_pool_indices, seq_lens, idx_is_i64, extend_seq_lens,
                             extend_prefix_lens, extend_start_loc, bs, max_extend_len, Hq, Hkv, D, q_dtype, kv_dtype,
                             sm_scale, 0.f, k_scale, v_scale, causal, B200ATTN_EXTEND_TCGEN05, ws_lse, -1, nullptr, nullptr,
                             1, page_size, stream, kv_splits, so, sl);
  if (rc != B200ATTN_OK) return rc;
  return extend_split_merge_launch(workspace, ws_lse, kv_splits, so, sl, T, Hq, D, q_dtype == B200ATTN_BF16, out,
                                   out_tok_stride, lse_out, static_cast<cudaStream_t>(stream));
}

