// Shared declarations of the extend / prefill kernels.
#pragma once
#include "common.cuh"

namespace b200attn {

struct ExtendParams {
  const void* q;
  void* out;
  const void* k_pool;
  const void* v_pool;
  const int32_t* req_to_token;
  const void* req_pool_indices;
  const void* seq_lens;
  const int32_t* extend_seq_lens;
  const int32_t* extend_prefix_lens;
  const int32_t* extend_start_loc;
  float* lse_out;  // optional fp32 [T, Hq]
  int64_t q_tok_stride, out_tok_stride, pool_tok_stride, table_stride;
  int idx_is_i64;
  int bs, Hq, Hkv, G, max_extend_len;
  int causal;
  float scale_log2;  // sm_scale * k_scale * log2(e)
  float scale_raw;   // sm_scale * k_scale
  float logit_cap;
  float v_scale;
  unsigned long long* prof;  // B200ATTN_PROF_PTR: device buffer for clock64 phase timings of CTA (0,0,0) (dev only)
  int grid_order;  // tcgen05 kernel: 0 = (head pair, request, q tile), 1 = (head pair, q tile, request)  (B200ATTN_GRID_ORDER)
  // ---- masked variant (extend_mma.cu, kMasked): sliding window and speculative-decoding custom mask
  // window >= 0: only the last min(P, window+1) cached-prefix keys are read (update_sliding_window_buffer,
  // triton_backend.py:927-955) and, for window > 0, a prefix key at offset jw inside that range is visible to
  // query t iff t <= jw + window (extend_attention.py:168-174); new-token keys stay causal (no window mask there).
  int window;
  // custom_mask != NULL: uint8/bool, request i's block starts at mask_indptr[i] and is [E_i, L_i] row-major;
  // new-token keys take mask (and causal), cached-prefix keys take it only when skip_prefix_mask == 0
  // (extend_attention.py:158-167,236-247)
  const uint8_t* custom_mask;
  const int64_t* mask_indptr;
  int skip_prefix_mask;
  int page_size;  // allocator page size (b200attn_extend_paged); % 16 == 0 with a 16-bit pool: bulk-tensor (TMA) K/V loads
  // KV splits of the tcgen05 kernel (b200attn_extend_split): out / lse_out then point at the partial buffers
  // [split][token][Hq][D] (q dtype) and [split][token][Hq] (fp32), the strides are in elements
  int kv_splits;
  int64_t split_out_stride, split_lse_stride;
  int exp_emu;  // tcgen05 kernel: of every 4 element pairs, this many take the FMA-pipe exp2 (B200ATTN_EXP_EMU, default 0: measured best at C3)
};

// register-MMA kernel (extend_mma.cu): any shape the library accepts
int extend_mma_launch(const ExtendParams& p, int D, int kvb, bool q_bf16, cudaStream_t st);
// same kernel compiled with the window / custom-mask predicate (p.window >= 0 or p.custom_mask != NULL)
int extend_mma_masked_launch(const ExtendParams& p, int D, int kvb, bool q_bf16, cudaStream_t st);

// tcgen05 / TMEM kernel (extend_tc.cu): D == 128, 16-bit or FP8-E4M3 KV, logit_cap == 0
bool extend_tc_supported(const ExtendParams& p, int D, int kvb);
int extend_tc_launch(const ExtendParams& p, bool q_bf16, int kvb, cudaStream_t st);
// LSE merge of the kv_splits partials written by the tcgen05 kernel into the caller's out / lse_out
int extend_split_merge_launch(const void* ws_o, const float* ws_lse, int kv_splits, int64_t split_out_stride,
                              int64_t split_lse_stride, int T, int Hq, int D, bool bf16, void* out, int64_t out_tok_stride,
                              float* lse_out, cudaStream_t st);

}  // namespace b200attn
