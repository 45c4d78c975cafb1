// Fused rotary embedding + KV-pool write: the two steps the reference runs right before the
// attention kernels, in one launch.
//   rotary_emb(positions, q, k)   python/sglang/srt/layers/rotary_embedding.py:138-166 (forward_native, the CPU
//                                 semantics), :218-244 (forward_cuda -> sgl-kernel apply_rope_with_cos_sin_cache_inplace,
//                                 sgl-kernel/csrc/elementwise/rope.cu:22-89); caller models/llama.py:188-191
//   set_kv_buffer(layer, loc, k, v)   python/sglang/srt/mem_cache/memory_pool.py:420-458
// q and k are rotated in place (as the reference's CUDA op does); the rotated k and v are cast and scattered into the
// pool from registers, which saves one read of k and the launches of the separate KV write.
//
// HBM-bound elementwise kernel.  A work item is (token, head slot, unit): head slots are the Hq query heads, the Hkv
// key heads and the Hkv value heads of the token; a unit is two 16-byte chunks of the head row --
//   rotary part, NeoX style  : chunk u and its partner u + rotary_dim/16 (x1 | x2 halves, rotary_embedding.py:57-58)
//   rotary part, GPT-J style : chunks 2u and 2u+1, each rotated inside itself (even/odd pairs, :60-61)
//   pass-through part / V    : two consecutive chunks, copied
// Math modes: FP32 = cos/sin as stored, one rounding at the end (the CUDA op); NATIVE = forward_native's op-by-op
// rounding in the tensor dtype (cos/sin cast to it first, each product and the sum rounded), bit-exact with the
// reference on CPU.
#include "common.cuh"
#include "kv_quant.cuh"

namespace b200attn {

struct RopeParams {
  void* q;
  void* k;
  const void* v;
  int64_t q_tok_stride, k_tok_stride, v_tok_stride;
  const void* positions;
  int pos_is_i64;
  const void* cache;  // [max_pos, rotary_dim]: cos | sin
  int rotary_dim;
  int T, Hq, Hkv, D;
  const int64_t* loc;  // nullptr: rope only
  void* pool[2];
  int64_t pool_tok_stride;
  const float* scale_dev[2];
  float scale_host[2];
  int n_scales, scale_mode;
};

template <bool kBf16>
__device__ __forceinline__ float rnd(float f) {  // round to the tensor dtype and back
  return in_to_f32<kBf16>(f32_to_in<kBf16>(f));
}

// 8 cos (or sin) values starting at element e of a cache row
template <int kCacheDt /*0 bf16, 1 f16, 3 f32*/>
__device__ __forceinline__ void load8(const void* row, int e, float (&out)[8]) {
  if constexpr (kCacheDt == B200ATTN_F32) {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(row) + e);
    const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(row) + e + 4);
    out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w; out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
  } else {
    const uint4 w = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(row) + e);
    const uint32_t r[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
      out[i] = in_to_f32<kCacheDt == B200ATTN_BF16>(static_cast<uint16_t>(r[i >> 1] >> ((i & 1) * 16)));
  }
}
template <int kCacheDt>
__device__ __forceinline__ void load4(const void* row, int e, float (&out)[4]) {
  if constexpr (kCacheDt == B200ATTN_F32) {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(row) + e);
    out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w;
  } else {
    const uint2 w = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(row) + e);
    const uint32_t r[2] = {w.x, w.y};
#pragma unroll
    for (int i = 0; i < 4; ++i)
      out[i] = in_to_f32<kCacheDt == B200ATTN_BF16>(static_cast<uint16_t>(r[i >> 1] >> ((i & 1) * 16)));
  }
}

template <bool kBf16>
__device__ __forceinline__ void unpack8(const uint4 w, float (&x)[8]) {
  const uint32_t r[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = in_to_f32<kBf16>(static_cast<uint16_t>(r[i >> 1] >> ((i & 1) * 16)));
}
template <bool kBf16>
__device__ __forceinline__ uint4 pack8(const float (&x)[8]) {
  uint32_t r[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    r[i] = static_cast<uint32_t>(f32_to_in<kBf16>(x[2 * i])) | (static_cast<uint32_t>(f32_to_in<kBf16>(x[2 * i + 1])) << 16);
  return make_uint4(r[0], r[1], r[2], r[3]);
}

// o1 = x1 cos - x2 sin, o2 = x2 cos + x1 sin  (rotary_embedding.py:62-63)
template <bool kBf16, bool kNative>
__device__ __forceinline__ void rotate(float x1, float x2, float c, float s, float& o1, float& o2) {
  if constexpr (kNative) {
    c = rnd<kBf16>(c); s = rnd<kBf16>(s);  // cos.to(x.dtype), sin.to(x.dtype)
    o1 = rnd<kBf16>(__fsub_rn(rnd<kBf16>(__fmul_rn(x1, c)), rnd<kBf16>(__fmul_rn(x2, s))));
    o2 = rnd<kBf16>(__fadd_rn(rnd<kBf16>(__fmul_rn(x2, c)), rnd<kBf16>(__fmul_rn(x1, s))));
  } else {
    o1 = __fsub_rn(__fmul_rn(x1, c), __fmul_rn(x2, s));
    o2 = __fadd_rn(__fmul_rn(x2, c), __fmul_rn(x1, s));
  }
}

// kPool: 0 no pool write, 1 same-dtype pool, 2 e4m3 pool
template <bool kBf16, int kCacheDt, bool kNeox, bool kNative, int kPool>
__global__ void __launch_bounds__(256) rope_kv_write_kernel(const RopeParams p) {
  const int units = p.D >> 4;             // units per head row
  const int rot_units = p.rotary_dim >> 4;
  const int slots = p.Hq + 2 * p.Hkv;
  const int64_t total = static_cast<int64_t>(p.T) * slots * units;
  const int cache_es = (kCacheDt == B200ATTN_F32) ? 4 : 2;
  for (int64_t it = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; it < total;
       it += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int u = static_cast<int>(it % units);
    const int64_t th = it / units;
    const int hs = static_cast<int>(th % slots);
    const int64_t t = th / slots;
    const int which = hs < p.Hq ? 0 : (hs < p.Hq + p.Hkv ? 1 : 2);  // q, k, v
    if (kPool == 0 && which == 2) continue;
    const int h = which == 0 ? hs : (which == 1 ? hs - p.Hq : hs - p.Hq - p.Hkv);
    uint16_t* row = which == 0 ? reinterpret_cast<uint16_t*>(p.q) + t * p.q_tok_stride + static_cast<int64_t>(h) * p.D
                  : which == 1 ? reinterpret_cast<uint16_t*>(p.k) + t * p.k_tok_stride + static_cast<int64_t>(h) * p.D
                               : const_cast<uint16_t*>(reinterpret_cast<const uint16_t*>(p.v)) + t * p.v_tok_stride +
                                     static_cast<int64_t>(h) * p.D;
    int ca, cb;  // the two chunks (8 elements each) of this unit
    const bool rot = (which != 2) && (u < rot_units);
    if (rot) {
      if (kNeox) { ca = u; cb = u + rot_units; } else { ca = 2 * u; cb = 2 * u + 1; }
    } else {
      // pass-through chunks follow the rotary part; V rows are all pass-through
      const int base = (which == 2) ? 0 : 2 * rot_units;
      const int uu = (which == 2) ? u : u - rot_units;
      ca = base + 2 * uu; cb = ca + 1;
    }
    uint4 wa = *reinterpret_cast<const uint4*>(row + ca * 8);
    uint4 wb = *reinterpret_cast<const uint4*>(row + cb * 8);
    if (rot) {
      const int64_t pos = load_index(p.positions, p.pos_is_i64, static_cast<int>(t));
      const uint8_t* crow = reinterpret_cast<const uint8_t*>(p.cache) + pos * p.rotary_dim * cache_es;
      const int half = p.rotary_dim >> 1;
      float xa[8], xb[8];
      unpack8<kBf16>(wa, xa);
      unpack8<kBf16>(wb, xb);
      if constexpr (kNeox) {
        float c[8], s[8];
        load8<kCacheDt>(crow, ca * 8, c);
        load8<kCacheDt>(crow, half + ca * 8, s);
#pragma unroll
        for (int i = 0; i < 8; ++i) rotate<kBf16, kNative>(xa[i], xb[i], c[i], s[i], xa[i], xb[i]);
      } else {
        float c[4], s[4];
        load4<kCacheDt>(crow, ca * 4, c);
        load4<kCacheDt>(crow, half + ca * 4, s);
#pragma unroll
        for (int i = 0; i < 4; ++i) rotate<kBf16, kNative>(xa[2 * i], xa[2 * i + 1], c[i], s[i], xa[2 * i], xa[2 * i + 1]);
        load4<kCacheDt>(crow, cb * 4, c);
        load4<kCacheDt>(crow, half + cb * 4, s);
#pragma unroll
        for (int i = 0; i < 4; ++i) rotate<kBf16, kNative>(xb[2 * i], xb[2 * i + 1], c[i], s[i], xb[2 * i], xb[2 * i + 1]);
      }
      wa = pack8<kBf16>(xa);
      wb = pack8<kBf16>(xb);
      *reinterpret_cast<uint4*>(row + ca * 8) = wa;
      *reinterpret_cast<uint4*>(row + cb * 8) = wb;
    }
    if constexpr (kPool != 0) {
      if (which != 0) {
        const int kv = which - 1;
        const int64_t slot = p.loc[t];
        const int64_t dst = slot * p.pool_tok_stride + static_cast<int64_t>(h) * p.D;
        if constexpr (kPool == 1) {
          uint16_t* pool = reinterpret_cast<uint16_t*>(p.pool[kv]) + dst;
          *reinterpret_cast<uint4*>(pool + ca * 8) = wa;
          *reinterpret_cast<uint4*>(pool + cb * 8) = wb;
        } else {
          float scale = p.scale_host[kv];
          if (p.scale_dev[kv] != nullptr) scale = p.scale_dev[kv][p.n_scales > 1 ? h : 0];
          uint8_t* pool = reinterpret_cast<uint8_t*>(p.pool[kv]) + dst;
          uint4 back;
          *reinterpret_cast<uint2*>(pool + ca * 8) = quantize_chunk_e4m3<kBf16>(wa, scale, p.scale_mode, back);
          *reinterpret_cast<uint2*>(pool + cb * 8) = quantize_chunk_e4m3<kBf16>(wb, scale, p.scale_mode, back);
        }
      }
    }
  }
}

namespace {
template <bool kBf16, int kCacheDt, bool kNeox, bool kNative>
void launch_pool(const RopeParams& p, int pool_mode, dim3 grid, cudaStream_t st) {
  if (pool_mode == 0) rope_kv_write_kernel<kBf16, kCacheDt, kNeox, kNative, 0><<<grid, 256, 0, st>>>(p);
  else if (pool_mode == 1) rope_kv_write_kernel<kBf16, kCacheDt, kNeox, kNative, 1><<<grid, 256, 0, st>>>(p);
  else rope_kv_write_kernel<kBf16, kCacheDt, kNeox, kNative, 2><<<grid, 256, 0, st>>>(p);
}
template <bool kBf16, int kCacheDt>
void launch_style(const RopeParams& p, bool neox, bool native, int pool_mode, dim3 grid, cudaStream_t st) {
  if (neox) {
    if (native) launch_pool<kBf16, kCacheDt, true, true>(p, pool_mode, grid, st);
    else launch_pool<kBf16, kCacheDt, true, false>(p, pool_mode, grid, st);
  } else {
    if (native) launch_pool<kBf16, kCacheDt, false, true>(p, pool_mode, grid, st);
    else launch_pool<kBf16, kCacheDt, false, false>(p, pool_mode, grid, st);
  }
}
}  // namespace

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_rope_kv_write(void* q, int64_t q_tok_stride, void* k, int64_t k_tok_stride, const void* v,
                                                int64_t v_tok_stride, const void* positions, int pos_is_i64,
                                                const void* cos_sin_cache, int cache_dtype, int rotary_dim, int is_neox,
                                                int T, int Hq, int Hkv, int D, int dtype, const int64_t* loc,
                                                void* k_pool, void* v_pool, int64_t pool_tok_stride, int pool_dtype,
                                                float k_scale_host, float v_scale_host, const float* k_scale_dev,
                                                const float* v_scale_dev, int n_scales, int scale_mode, int math_mode,
                                                b200attn_stream_t stream) {
  B2_REQUIRE(T >= 0 && Hq > 0 && Hkv > 0 && D > 0, B200ATTN_ERR_INVALID, "rope_kv_write: bad sizes T=%d Hq=%d Hkv=%d D=%d",
             T, Hq, Hkv, D);
  if (T == 0) return B200ATTN_OK;
  B2_REQUIRE(q && k && positions && cos_sin_cache, B200ATTN_ERR_INVALID, "rope_kv_write: null pointer");
  B2_REQUIRE(dtype == B200ATTN_BF16 || dtype == B200ATTN_F16, B200ATTN_ERR_UNSUPPORTED, "rope_kv_write: dtype %d", dtype);
  B2_REQUIRE(cache_dtype == B200ATTN_F32 || cache_dtype == dtype, B200ATTN_ERR_UNSUPPORTED,
             "rope_kv_write: cos_sin_cache dtype %d (fp32 or the tensor dtype)", cache_dtype);
  B2_REQUIRE(D % 16 == 0 && rotary_dim % 16 == 0 && rotary_dim > 0 && rotary_dim <= D, B200ATTN_ERR_UNSUPPORTED,
             "rope_kv_write: head dim %d / rotary dim %d must be multiples of 16", D, rotary_dim);
  B2_REQUIRE(aligned16(q) && aligned16(k) && aligned16(cos_sin_cache) && q_tok_stride % 8 == 0 && k_tok_stride % 8 == 0,
             B200ATTN_ERR_INVALID, "rope_kv_write: q/k rows and the cache must be 16-byte aligned");
  B2_REQUIRE(math_mode == B200ATTN_ROPE_FP32 || math_mode == B200ATTN_ROPE_NATIVE, B200ATTN_ERR_INVALID,
             "rope_kv_write: math_mode %d", math_mode);
  int pool_mode = 0;
  if (loc != nullptr) {
    B2_REQUIRE(v && k_pool && v_pool, B200ATTN_ERR_INVALID, "rope_kv_write: null v / pool pointer");
    B2_REQUIRE(pool_dtype == dtype || pool_dtype == B200ATTN_FP8_E4M3, B200ATTN_ERR_UNSUPPORTED,
               "rope_kv_write: pool dtype %d with tensor dtype %d", pool_dtype, dtype);
    B2_REQUIRE(aligned16(v) && v_tok_stride % 8 == 0 && aligned16(k_pool) && aligned16(v_pool) &&
                   (pool_tok_stride * elem_size(pool_dtype)) % 8 == 0,
               B200ATTN_ERR_INVALID, "rope_kv_write: v rows and pool rows must be aligned");
    B2_REQUIRE(pool_tok_stride >= static_cast<int64_t>(Hkv) * D, B200ATTN_ERR_INVALID, "rope_kv_write: pool_tok_stride < Hkv*D");
    B2_REQUIRE(n_scales == 1 || n_scales == Hkv || (!k_scale_dev && !v_scale_dev), B200ATTN_ERR_INVALID,
               "rope_kv_write: n_scales=%d must be 1 or Hkv=%d", n_scales, Hkv);
    B2_REQUIRE(scale_mode == B200ATTN_SCALE_DIV || scale_mode == B200ATTN_SCALE_RCP, B200ATTN_ERR_INVALID,
               "rope_kv_write: scale_mode %d", scale_mode);
    pool_mode = (pool_dtype == dtype) ? 1 : 2;
  }
  RopeParams p;
  p.q = q; p.k = k; p.v = v;
  p.q_tok_stride = q_tok_stride; p.k_tok_stride = k_tok_stride; p.v_tok_stride = v_tok_stride;
  p.positions = positions; p.pos_is_i64 = pos_is_i64;
  p.cache = cos_sin_cache; p.rotary_dim = rotary_dim;
  p.T = T; p.Hq = Hq; p.Hkv = Hkv; p.D = D;
  p.loc = loc;
  p.pool[0] = k_pool; p.pool[1] = v_pool;
  p.pool_tok_stride = pool_tok_stride;
  p.scale_dev[0] = k_scale_dev; p.scale_dev[1] = v_scale_dev;
  p.scale_host[0] = k_scale_host; p.scale_host[1] = v_scale_host;
  p.n_scales = n_scales; p.scale_mode = scale_mode;

  const int64_t total = static_cast<int64_t>(T) * (Hq + 2 * Hkv) * (D / 16);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  dim3 grid(static_cast<unsigned>(blocks), 1, 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool neox = is_neox != 0, native = math_mode == B200ATTN_ROPE_NATIVE;
  if (dtype == B200ATTN_BF16) {
    if (cache_dtype == B200ATTN_F32) launch_style<true, B200ATTN_F32>(p, neox, native, pool_mode, grid, st);
    else launch_style<true, B200ATTN_BF16>(p, neox, native, pool_mode, grid, st);
  } else {
    if (cache_dtype == B200ATTN_F32) launch_style<false, B200ATTN_F32>(p, neox, native, pool_mode, grid, st);
    else launch_style<false, B200ATTN_F16>(p, neox, native, pool_mode, grid, st);
  }
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}
