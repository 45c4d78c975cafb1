// 2-way LSE merge of partial attention states (prefix (+) suffix).
// Reference: python/sglang/srt/layers/attention/merge_state.py:26-46 (dispatcher),
// sgl-kernel/csrc/attention/merge_attn_states.cu:32-106 (CUDA), triton_ops/merge_state.py:8-63,
// pure-torch statement sgl-kernel/tests/test_merge_state_v2.py:100-133.
// HBM-bound elementwise pass: 16-byte loads, one thread per 16 bytes of a (token, head) row.
#include "common.cuh"
#include "extend.cuh"

namespace b200attn {

template <typename T> struct Vec16;
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ float get(const uint4& v, int i) {
    const uint32_t w = (&v.x)[i >> 1];
    return __uint_as_float((i & 1) ? (w & 0xffff0000u) : (w << 16));
  }
  static __device__ void set(uint4& v, int i, float lo, float hi) { (&v.x)[i >> 1] = pack2<true>(lo, hi); }
};
template <> struct Vec16<__half> {
  static constexpr int N = 8;
  static __device__ float get(const uint4& v, int i) {
    const uint32_t w = (&v.x)[i >> 1];
    return __half2float(__ushort_as_half(static_cast<uint16_t>((i & 1) ? (w >> 16) : (w & 0xffff))));
  }
  static __device__ void set(uint4& v, int i, float lo, float hi) { (&v.x)[i >> 1] = pack2<false>(lo, hi); }
};
template <> struct Vec16<float> {
  static constexpr int N = 4;
  static __device__ float get(const uint4& v, int i) { return __uint_as_float((&v.x)[i]); }
  static __device__ void set(uint4& v, int i, float lo, float hi) {
    (&v.x)[i] = __float_as_uint(lo);
    (&v.x)[i + 1] = __float_as_uint(hi);
  }
};

template <typename T>
// `out` may alias `a` or `b` (the verify route merges in place, backend.py): every thread loads its own 16 bytes of a
// and b before it stores them, and none of the three pointers is __restrict__, so the in-place call is well defined.
__global__ void __launch_bounds__(256) merge_state_kernel(const T* a, const float* __restrict__ lse_a, const T* b,
                                                          const float* __restrict__ lse_b, T* out,
                                                          float* __restrict__ out_lse, int64_t rows, int D) {
  using V = Vec16<T>;
  const int vecs = D / V::N;
  const int64_t total = rows * vecs;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / vecs;
    const int c = static_cast<int>(i % vecs);
    float la = lse_a[r], lb = lse_b[r];
    if (la == INFINITY) la = -INFINITY;
    if (lb == INFINITY) lb = -INFINITY;
    const float m = fmaxf(la, lb);
    // both -inf: the reference computes exp(nan); define the result as 0 / -inf instead
    const float ea = (m == -INFINITY) ? 0.f : expf(la - m);
    const float eb = (m == -INFINITY) ? 0.f : expf(lb - m);
    const float se = ea + eb;
    const float wa = se > 0.f ? ea / se : 0.f, wb = se > 0.f ? eb / se : 0.f;
    const uint4 va = reinterpret_cast<const uint4*>(a)[i];
    const uint4 vb = reinterpret_cast<const uint4*>(b)[i];
    uint4 vo;
#pragma unroll
    for (int e = 0; e < V::N; e += 2) {
      const float x0 = V::get(va, e) * wa + V::get(vb, e) * wb;
      const float x1 = V::get(va, e + 1) * wa + V::get(vb, e + 1) * wb;
      V::set(vo, e, x0, x1);
    }
    reinterpret_cast<uint4*>(out)[i] = vo;
    if (out_lse != nullptr && c == 0) out_lse[r] = se > 0.f ? logf(se) + m : -INFINITY;
  }
}


// n-way LSE merge of the partials the tcgen05 extend kernel writes with kv_splits > 1 (b200attn_extend_split):
// out[t, h, :] = sum_s w_s o_s[t, h, :] / sum_s w_s,  w_s = exp(lse_s[t, h] - max_s lse_s[t, h]);  a split that saw no key
// has lse = -inf and its (unwritten) values are never read.  One thread per 16 bytes of a (token, head) row.
template <typename T>
__global__ void __launch_bounds__(256)
extend_split_merge_kernel(const T* __restrict__ ws_o, const float* __restrict__ ws_lse, int S, int64_t so, int64_t sl,
                          int64_t rows /* T * Hq */, int Hq, int D, T* __restrict__ out, int64_t out_tok_stride,
                          float* __restrict__ lse_out) {
  using V = Vec16<T>;
  const int vecs = D / V::N;
  const int64_t total = rows * vecs;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / vecs;
    const int c = static_cast<int>(i % vecs);
    float m = -INFINITY;
    for (int s = 0; s < S; ++s) m = fmaxf(m, ws_lse[s * sl + r]);
    float acc[V::N];
#pragma unroll
    for (int e = 0; e < V::N; ++e) acc[e] = 0.f;
    float se = 0.f;
    for (int s = 0; s < S; ++s) {
      const float l = ws_lse[s * sl + r];
      if (l == -INFINITY) continue;
      const float w = expf(l - m);
      se += w;
      const uint4 v = *reinterpret_cast<const uint4*>(ws_o + s * so + r * D + c * V::N);
#pragma unroll
      for (int e = 0; e < V::N; ++e) acc[e] += w * V::get(v, e);
    }
    const float inv = se > 0.f ? 1.f / se : 0.f;
    uint4 vo;
#pragma unroll
    for (int e = 0; e < V::N; e += 2) V::set(vo, e, acc[e] * inv, acc[e + 1] * inv);
    const int64_t tok = r / Hq, h = r % Hq;
    *reinterpret_cast<uint4*>(out + tok * out_tok_stride + h * D + c * V::N) = vo;
    if (lse_out != nullptr && c == 0) lse_out[r] = se > 0.f ? m + logf(se) : -INFINITY;
  }
}

int extend_split_merge_launch(const void* ws_o, const float* ws_lse, int kv_splits, int64_t split_out_stride,
                              int64_t split_lse_stride, int T, int Hq, int D, bool bf16, void* out, int64_t out_tok_stride,
                              float* lse_out, cudaStream_t st) {
  const int64_t rows = static_cast<int64_t>(T) * Hq;
  const int64_t total = rows * (D / 8);
  const int blocks = static_cast<int>(total < 148 * 8 * 256 ? (total + 255) / 256 : 148 * 8);
  if (blocks == 0) return B200ATTN_OK;
  if (bf16)
    extend_split_merge_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(ws_o), ws_lse, kv_splits, split_out_stride, split_lse_stride, rows, Hq, D,
        reinterpret_cast<__nv_bfloat16*>(out), out_tok_stride, lse_out);
  else
    extend_split_merge_kernel<__half><<<blocks, 256, 0, st>>>(
        reinterpret_cast<const __half*>(ws_o), ws_lse, kv_splits, split_out_stride, split_lse_stride, rows, Hq, D,
        reinterpret_cast<__half*>(out), out_tok_stride, lse_out);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_merge_state(const void* a, const float* lse_a, const void* b, const float* lse_b,
                                    void* out, float* out_lse, int T, int H, int D, int dtype,
                                    b200attn_stream_t stream) {
  B2_REQUIRE(T >= 0 && H > 0 && D > 0, B200ATTN_ERR_INVALID, "merge_state: bad sizes");
  if (T == 0) return B200ATTN_OK;
  B2_REQUIRE(a && b && lse_a && lse_b && out, B200ATTN_ERR_INVALID, "merge_state: null pointer");
  const int es = elem_size(dtype);
  B2_REQUIRE(dtype == B200ATTN_BF16 || dtype == B200ATTN_F16 || dtype == B200ATTN_F32, B200ATTN_ERR_UNSUPPORTED,
             "merge_state: dtype %d", dtype);
  B2_REQUIRE((D * es) % 16 == 0, B200ATTN_ERR_UNSUPPORTED, "merge_state: head dim %d needs 16-byte rows", D);
  B2_REQUIRE(aligned16(a) && aligned16(b) && aligned16(out), B200ATTN_ERR_INVALID, "merge_state: alignment");
  const int64_t rows = static_cast<int64_t>(T) * H;
  const int64_t total = rows * (D * es / 16);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned g = static_cast<unsigned>(blocks);
  if (dtype == B200ATTN_BF16)
    merge_state_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)a, lse_a, (const __nv_bfloat16*)b, lse_b,
                                                         (__nv_bfloat16*)out, out_lse, rows, D);
  else if (dtype == B200ATTN_F16)
    merge_state_kernel<__half><<<g, 256, 0, st>>>((const __half*)a, lse_a, (const __half*)b, lse_b, (__half*)out,
                                                  out_lse, rows, D);
  else
    merge_state_kernel<float><<<g, 256, 0, st>>>((const float*)a, lse_a, (const float*)b, lse_b, (float*)out, out_lse,
                                                 rows, D);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}
