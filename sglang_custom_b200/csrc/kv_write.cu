// KV-pool scatter write, replacing MHATokenToKVPool.set_kv_buffer
// (reference: python/sglang/srt/mem_cache/memory_pool.py:420-458).
//
// HBM-bound byte mover: one thread owns 8 consecutive elements (16 B of bf16/f16 input)
// of one (token, head) row, K and V handled by the same launch (blockIdx.y selects).
// Algorithmic bytes per token: Hkv * D * (2 * in_size + 2 * pool_size) + 8 (loc).
#include "common.cuh"
#include "kv_quant.cuh"

namespace b200attn {

struct KvWriteParams {
  const void* src[2];  // k, v
  void* pool[2];
  int64_t tok_stride[2];   // elements
  int64_t head_stride[2];  // elements
  const int64_t* loc;
  int64_t pool_tok_stride;  // elements
  const float* scale_dev[2];
  float scale_host[2];
  int n_scales;
  int scale_mode;
  int writeback;
  int T, Hkv, D;
};

// kMode 0: same-dtype bit copy. kMode 1: 16-bit float -> e4m3 with optional scale.
template <bool kBf16, int kMode>
__global__ void __launch_bounds__(256) kv_write_kernel(const KvWriteParams p) {
  // (ternaries, not p.src[blockIdx.y]: a runtime index into a kernel-parameter array forces a local-memory copy)
  const bool is_v = blockIdx.y != 0;
  const int chunks_per_row = p.D >> 3;
  const int64_t total = static_cast<int64_t>(p.T) * p.Hkv * chunks_per_row;
  const uint16_t* __restrict__ src = reinterpret_cast<const uint16_t*>(is_v ? p.src[1] : p.src[0]);
  void* const pool = is_v ? p.pool[1] : p.pool[0];
  const int64_t tok_stride = is_v ? p.tok_stride[1] : p.tok_stride[0];
  const int64_t head_stride = is_v ? p.head_stride[1] : p.head_stride[0];
  const float* const scale_dev = is_v ? p.scale_dev[1] : p.scale_dev[0];
  const float scale_host = is_v ? p.scale_host[1] : p.scale_host[0];
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c < total;
       c += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int ch = static_cast<int>(c % chunks_per_row);
    const int64_t th = c / chunks_per_row;
    const int h = static_cast<int>(th % p.Hkv);
    const int64_t t = th / p.Hkv;
    const int64_t slot = p.loc[t];
    const uint16_t* s = src + t * tok_stride + h * head_stride + ch * 8;
    const uint4 in = *reinterpret_cast<const uint4*>(s);
    const int64_t dst_elem = slot * p.pool_tok_stride + static_cast<int64_t>(h) * p.D + ch * 8;
    if constexpr (kMode == 0) {
      *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(pool) + dst_elem) = in;
    } else {
      float scale = scale_host;
      if (scale_dev != nullptr) scale = scale_dev[p.n_scales > 1 ? h : 0];
      uint4 back;
      const uint2 q8 = quantize_chunk_e4m3<kBf16>(in, scale, p.scale_mode, back);
      *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(pool) + dst_elem) = q8;
      if (p.writeback && scale != 0.0f) *reinterpret_cast<uint4*>(const_cast<uint16_t*>(s)) = back;
    }
  }
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_kv_write(const void* k, const void* v, int64_t k_tok_stride, int64_t k_head_stride,
                                 int64_t v_tok_stride, int64_t v_head_stride, const int64_t* loc, int T,
                                 int Hkv, int D, void* k_pool, void* v_pool, int64_t pool_tok_stride,
                                 int in_dtype, int pool_dtype, float k_scale_host, float v_scale_host,
                                 const float* k_scale_dev, const float* v_scale_dev, int n_scales,
                                 int scale_mode, int writeback_scaled, b200attn_stream_t stream) {
  B2_REQUIRE(T >= 0 && Hkv > 0 && D > 0, B200ATTN_ERR_INVALID, "kv_write: bad sizes T=%d Hkv=%d D=%d", T, Hkv, D);
  if (T == 0) return B200ATTN_OK;
  B2_REQUIRE(k && v && loc && k_pool && v_pool, B200ATTN_ERR_INVALID, "kv_write: null pointer");
  B2_REQUIRE(in_dtype == B200ATTN_BF16 || in_dtype == B200ATTN_F16, B200ATTN_ERR_UNSUPPORTED,
             "kv_write: input dtype %d (bf16/f16 only)", in_dtype);
  B2_REQUIRE(pool_dtype == in_dtype || pool_dtype == B200ATTN_FP8_E4M3, B200ATTN_ERR_UNSUPPORTED,
             "kv_write: pool dtype %d with input dtype %d", pool_dtype, in_dtype);
  B2_REQUIRE(D % 8 == 0, B200ATTN_ERR_INVALID, "kv_write: D=%d must be a multiple of 8", D);
  B2_REQUIRE(k_tok_stride % 8 == 0 && k_head_stride % 8 == 0 && v_tok_stride % 8 == 0 &&
                 v_head_stride % 8 == 0 && aligned16(k) && aligned16(v),
             B200ATTN_ERR_INVALID, "kv_write: k/v rows must be 16-byte aligned");
  const int pool_es = elem_size(pool_dtype);
  B2_REQUIRE((pool_tok_stride * pool_es) % (8 * pool_es) == 0 && aligned16(k_pool) && aligned16(v_pool),
             B200ATTN_ERR_INVALID, "kv_write: pool rows must be aligned");
  B2_REQUIRE(pool_tok_stride >= static_cast<int64_t>(Hkv) * D, B200ATTN_ERR_INVALID,
             "kv_write: pool_tok_stride < Hkv*D");
  B2_REQUIRE(n_scales == 1 || n_scales == Hkv || (!k_scale_dev && !v_scale_dev), B200ATTN_ERR_INVALID,
             "kv_write: n_scales=%d must be 1 or Hkv=%d", n_scales, Hkv);
  B2_REQUIRE(scale_mode == B200ATTN_SCALE_DIV || scale_mode == B200ATTN_SCALE_RCP, B200ATTN_ERR_INVALID,
             "kv_write: scale_mode %d", scale_mode);

  KvWriteParams p;
  p.src[0] = k; p.src[1] = v;
  p.pool[0] = k_pool; p.pool[1] = v_pool;
  p.tok_stride[0] = k_tok_stride; p.tok_stride[1] = v_tok_stride;
  p.head_stride[0] = k_head_stride; p.head_stride[1] = v_head_stride;
  p.loc = loc;
  p.pool_tok_stride = pool_tok_stride;
  p.scale_dev[0] = k_scale_dev; p.scale_dev[1] = v_scale_dev;
  p.scale_host[0] = k_scale_host; p.scale_host[1] = v_scale_host;
  p.n_scales = n_scales;
  p.scale_mode = scale_mode;
  p.writeback = writeback_scaled;
  p.T = T; p.Hkv = Hkv; p.D = D;

  const int64_t total = static_cast<int64_t>(T) * Hkv * (D / 8);
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;  // grid-stride beyond 16 CTAs per SM
  if (blocks > cap) blocks = cap;
  dim3 grid(static_cast<unsigned>(blocks), 2, 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool bf16 = in_dtype == B200ATTN_BF16;
  if (pool_dtype == in_dtype) {
    if (bf16) kv_write_kernel<true, 0><<<grid, threads, 0, st>>>(p);
    else kv_write_kernel<false, 0><<<grid, threads, 0, st>>>(p);
  } else {
    if (bf16) kv_write_kernel<true, 1><<<grid, threads, 0, st>>>(p);
    else kv_write_kernel<false, 1><<<grid, threads, 0, st>>>(p);
  }
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}
