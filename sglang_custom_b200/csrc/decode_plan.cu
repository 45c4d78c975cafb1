// Device-side split-KV plan for decode: no host sync, CUDA-graph safe.
//
// Replaces the per-forward metadata the reference builds in init_forward_metadata
// (layers/attention/triton_backend.py:160-226: cumsum(seq_lens), a flattened kv_indices copy,
// and the num_kv_splits heuristic triton_backend.py:875-924).  Here the page table is read in
// place by the decode kernel, so the plan is only bs + n_ctas + 9 integers:
// an exclusive prefix sum of 16-token units and, for every persistent CTA, the request its
// equal share of the flattened unit space starts in.
#include "common.cuh"
#include "decode_plan.cuh"

namespace b200attn {

__global__ void __launch_bounds__(1024) decode_plan_kernel(const void* seq_lens, int is_i64, int bs,
                                                           int n_ctas, int min_units, int window,
                                                           int32_t* plan) {
  __shared__ int warp_tot[32];
  __shared__ int running_sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int32_t* cum = plan + kPlanHeader;
  int32_t* first_req = plan + kPlanHeader + bs + 1;
  if (tid == 0) {
    running_sh = 0;
    cum[0] = 0;
  }
  __syncthreads();
  for (int base = 0; base < bs; base += 1024) {
    const int i = base + tid;
    int u = 0;
    if (i < bs) {
      int64_t L = load_index(seq_lens, is_i64, i);
      if (window >= 0 && L > window + 1) L = window + 1;  // last min(L, window + 1) tokens (triton_backend.py:936-939)
      u = L > 0 ? static_cast<int>((L + kUnit - 1) / kUnit) : 0;
    }
    int incl = u;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += y;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = warp_tot[lane];
#pragma unroll
      for (int s = 1; s < 32; s <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, s);
        if (lane >= s) w += y;
      }
      warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = running_sh + (warp > 0 ? warp_tot[warp - 1] : 0);
    if (i < bs) cum[i + 1] = before + incl;
    __syncthreads();
    if (tid == 0) running_sh += warp_tot[31];
    __syncthreads();
  }
  const int total = running_sh;
  int W = (total + n_ctas - 1) / n_ctas;
  if (W < min_units) W = min_units;
  if (W < 1) W = 1;
  if (tid == 0) {
    plan[kPlanTotalUnits] = total;
    plan[kPlanUnitsPerCta] = W;
    plan[kPlanBs] = bs;
    plan[kPlanNumCtas] = n_ctas;
    plan[kPlanWindow] = window;
  }
  // first request of every CTA: the largest i with cum[i] <= c*W  (cum was written by this block)
  for (int c = tid; c < n_ctas; c += 1024) {
    const int64_t u = static_cast<int64_t>(c) * W;
    int lo = 0, hi = bs;  // invariant: cum[lo] <= u, answer in [lo, hi)
    if (u >= total) {
      lo = bs > 0 ? bs - 1 : 0;
    } else {
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cum[mid] <= u) lo = mid; else hi = mid;
      }
    }
    first_req[c] = lo;
  }
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT size_t b200attn_decode_plan_bytes(int max_bs, int n_ctas) {
  if (max_bs < 0 || n_ctas <= 0) return 0;
  return sizeof(int32_t) * (static_cast<size_t>(kPlanHeader) + max_bs + 1 + n_ctas);
}

static int plan_impl(const void* seq_lens, int idx_is_i64, int bs, int n_ctas, int min_units_per_cta, int window,
                     int32_t* plan, size_t plan_bytes, b200attn_stream_t stream) {
  B2_REQUIRE(bs >= 0 && n_ctas > 0, B200ATTN_ERR_INVALID, "decode_plan: bs=%d n_ctas=%d", bs, n_ctas);
  B2_REQUIRE(window >= -1, B200ATTN_ERR_INVALID, "decode_plan: sliding_window_size=%d", window);
  B2_REQUIRE(plan != nullptr && (bs == 0 || seq_lens != nullptr), B200ATTN_ERR_INVALID, "decode_plan: null pointer");
  B2_REQUIRE(plan_bytes >= b200attn_decode_plan_bytes(bs, n_ctas), B200ATTN_ERR_WORKSPACE,
             "decode_plan: plan buffer %zu < %zu bytes", plan_bytes, b200attn_decode_plan_bytes(bs, n_ctas));
  decode_plan_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(seq_lens, idx_is_i64, bs, n_ctas,
                                                                        min_units_per_cta, window, plan);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

extern "C" B2_EXPORT int b200attn_decode_plan(const void* seq_lens, int idx_is_i64, int bs, int n_ctas,
                                    int min_units_per_cta, int32_t* plan, size_t plan_bytes,
                                    b200attn_stream_t stream) {
  return plan_impl(seq_lens, idx_is_i64, bs, n_ctas, min_units_per_cta, -1, plan, plan_bytes, stream);
}

// Plan for sliding-window layers: request i contributes its last min(L_i, sliding_window_size + 1) tokens
// (update_sliding_window_buffer, triton_backend.py:927-955).  The window is recorded in the plan header and
// b200attn_decode reads it from there, so one plan buffer per window size is all a mixed local/global model needs.
extern "C" B2_EXPORT int b200attn_decode_plan_window(const void* seq_lens, int idx_is_i64, int bs, int n_ctas,
                                                     int min_units_per_cta, int sliding_window_size, int32_t* plan,
                                                     size_t plan_bytes, b200attn_stream_t stream) {
  return plan_impl(seq_lens, idx_is_i64, bs, n_ctas, min_units_per_cta, sliding_window_size, plan, plan_bytes, stream);
}
