// One 16-byte chunk (8 bf16/f16 values) on its way into the KV pool: the cast rules of
// MHATokenToKVPool.set_kv_buffer (python/sglang/srt/mem_cache/memory_pool.py:436-446), shared by the KV-write
// kernel and the fused RoPE + KV-write kernel.
#pragma once
#include "common.cuh"

namespace b200attn {

template <bool kBf16>
__device__ __forceinline__ float in_to_f32(uint16_t b) {
  if constexpr (kBf16) {
    return __uint_as_float(static_cast<uint32_t>(b) << 16);
  } else {
    return __half2float(__ushort_as_half(b));
  }
}
template <bool kBf16>
__device__ __forceinline__ uint16_t f32_to_in(float f) {
  if constexpr (kBf16) {
    return __bfloat16_as_ushort(__float2bfloat16_rn(f));
  } else {
    return __half_as_ushort(__float2half_rn(f));
  }
}

// 8 values -> 8 e4m3 bytes.  scale == 0 means "no scaling" (the reference's k_scale is None).  torch's
// `cache_k.div_(scale)` happens in the tensor's own 16-bit dtype, so the quotient is rounded to that dtype before the
// e4m3 cast; `back` receives those rounded 16-bit values (the reference's in-place side effect on cache_k).
template <bool kBf16>
__device__ __forceinline__ uint2 quantize_chunk_e4m3(const uint4 in, float scale, int scale_mode, uint4& back) {
  const bool scaled = (scale != 0.0f);
  const float inv = (scale_mode == B200ATTN_SCALE_RCP && scaled) ? __fdiv_rn(1.0f, scale) : 0.0f;
  const uint32_t w[4] = {in.x, in.y, in.z, in.w};
  uint32_t bk[4];
  uint32_t out8[2] = {0u, 0u};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint16_t b = static_cast<uint16_t>(w[i >> 1] >> ((i & 1) * 16));
    float f = in_to_f32<kBf16>(b);
    uint16_t rb = b;
    if (scaled) {
      f = (scale_mode == B200ATTN_SCALE_RCP) ? __fmul_rn(f, inv) : __fdiv_rn(f, scale);
      rb = f32_to_in<kBf16>(f);
      f = in_to_f32<kBf16>(rb);
    }
    if (i & 1) {
      bk[i >> 1] |= static_cast<uint32_t>(rb) << 16;
    } else {
      bk[i >> 1] = rb;
    }
    out8[i >> 2] |= static_cast<uint32_t>(f32_to_e4m3_torch(f)) << ((i & 3) * 8);
  }
  back = make_uint4(bk[0], bk[1], bk[2], bk[3]);
  return make_uint2(out8[0], out8[1]);
}

}  // namespace b200attn
