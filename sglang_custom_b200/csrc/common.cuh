// Shared device/host helpers for the b200attn kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200attn.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "b200attn is written for sm_100a only"
#endif

#define B2_EXPORT __attribute__((visibility("default")))

namespace b200attn {

// ------------------------------------------------------------------ host side error plumbing
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define B2_REQUIRE(cond, code, ...)  \
  do {                               \
    if (!(cond)) {                   \
      ::b200attn::set_error(__VA_ARGS__); \
      return (code);                 \
    }                                \
  } while (0)

#define B2_CUDA(expr)                                                  \
  do {                                                                 \
    cudaError_t e__ = (expr);                                          \
    if (e__ != cudaSuccess) return ::b200attn::cuda_fail(e__, #expr);  \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int elem_size(int dtype) {
  switch (dtype) {
    case B200ATTN_BF16:
    case B200ATTN_F16: return 2;
    case B200ATTN_FP8_E4M3: return 1;
    case B200ATTN_F32: return 4;
    default: return 0;
  }
}
int sm_count();
int max_smem_optin();

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// ------------------------------------------------------------------ device helpers
#ifdef __CUDACC__

__device__ __forceinline__ int64_t load_index(const void* p, int is_i64, int i) {
  return is_i64 ? reinterpret_cast<const int64_t*>(p)[i]
                : static_cast<int64_t>(reinterpret_cast<const int32_t*>(p)[i]);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte async global->shared copy (LDGSTS). src_bytes == 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes = 16) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

// D(16x8,f32) += A(16x16) * B(16x8), bf16 or f16 operands
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

template <bool kBf16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (kBf16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// ---------------------------------------------------------------- FP8 E4M3 (torch.float8_e4m3fn)
// fp32 -> e4m3fn with the semantics of torch's cast (c10/util/Float8_e4m3fn.h as observed
// from torch: RNE; |x| that rounds to >= 480, inf and NaN all give S.1111.111; min subnormal 2^-9).
__device__ __forceinline__ uint8_t f32_to_e4m3_torch(float f) {
  uint32_t u = __float_as_uint(f);
  const uint32_t sign = (u >> 24) & 0x80u;
  u &= 0x7fffffffu;
  uint32_t r;
  if (u >= 0x43F00000u) {  // >= 480, inf, NaN
    r = 0x7Fu;
  } else if (u < 0x3C800000u) {  // < 2^-6: lands on the subnormal grid k * 2^-9, k in 0..8
    r = static_cast<uint32_t>(__float2int_rn(__uint_as_float(u) * 512.0f));
  } else {
    const uint32_t lsb = (u >> 20) & 1u;
    u += 0x7FFFFu + lsb;  // round to nearest even at bit 20
    r = (u - (120u << 23)) >> 20;
  }
  return static_cast<uint8_t>(r | sign);
}

// two packed e4m3 (low 16 bits of x) -> packed f16x2 (exact)
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint16_t x) {
  uint32_t r;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;\n" : "=r"(r) : "h"(x));
  return r;
}

// 16-byte streaming global load (read-only path, no L1 allocation)
__device__ __forceinline__ uint4 ld_nc_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

#endif  // __CUDACC__
}  // namespace b200attn
