// Host-side tensor-map encoding for the paged-tile K/V loads (see tma.cuh).
#include "tma.cuh"

#include <mutex>

#include "common.cuh"

namespace b200attn {

namespace {
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn encode_fn() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeFn>(p);
    (void)cudaGetLastError();
  });
  return fn;
}
}  // namespace

int encode_pool_tmap(CUtensorMap* out, const void* pool, int elem_bytes, int64_t pool_row_bytes, int row_segs,
                     int box_rows, int box_segs) {
  EncodeFn fn = encode_fn();
  B2_REQUIRE(fn != nullptr, B200ATTN_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  B2_REQUIRE(elem_bytes == 1 || elem_bytes == 2, B200ATTN_ERR_INVALID, "tensor map: element size %d", elem_bytes);
  B2_REQUIRE(pool_row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(pool) & 15u) == 0, B200ATTN_ERR_INVALID,
             "tensor map: the pool and its row stride must be 16-byte aligned");
  B2_REQUIRE(box_rows >= 1 && box_rows <= 256 && box_segs >= 1 && box_segs <= 256 && box_segs <= row_segs,
             B200ATTN_ERR_INVALID, "tensor map: box %d rows x %d segments", box_rows, box_segs);
  const cuuint32_t inner = static_cast<cuuint32_t>(128 / elem_bytes);
  // The row count is not part of the C ABI (the kernels only ever address slots taken from req_to_token), so the map
  // is given the largest extent the encoder accepts for this stride; no access relies on out-of-bounds filling.
  cuuint64_t rows = (static_cast<cuuint64_t>(1) << 40) / static_cast<cuuint64_t>(pool_row_bytes);
  if (rows > (static_cast<cuuint64_t>(1) << 31)) rows = static_cast<cuuint64_t>(1) << 31;
  const cuuint64_t dims[3] = {inner, rows, static_cast<cuuint64_t>(row_segs)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(pool_row_bytes), 128};  // bytes, dims 1 and 2
  const cuuint32_t box[3] = {inner, static_cast<cuuint32_t>(box_rows), static_cast<cuuint32_t>(box_segs)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult rc = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 3,
                         const_cast<void*>(pool), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(rc == CUDA_SUCCESS, B200ATTN_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %llu, stride %lld)",
             static_cast<int>(rc), static_cast<unsigned long long>(rows), static_cast<long long>(pool_row_bytes));
  return B200ATTN_OK;
}

}  // namespace b200attn
