// Page-table utilities: flatten (kv_indices) and the bit-exact page gather.
// References: python/sglang/srt/layers/attention/utils.py:5-41 (create_flashinfer_kv_indices_triton,
// tested by test/srt/test_create_kvindices.py:18-71) and the torch_native gather
// `k_cache[per_req_tokens]` (layers/attention/torch_native_backend.py:92-94,155-157).
#include "common.cuh"

namespace b200attn {

// one CTA per (request, 2048-entry slab): coalesced int32 copy of a row slice of the table
__global__ void __launch_bounds__(256) flatten_kernel(const int32_t* __restrict__ table, int64_t table_stride,
                                                      const void* rpi, const void* lens, const void* kv_start,
                                                      int is_i64, const int32_t* __restrict__ kv_indptr,
                                                      int32_t* __restrict__ kv_indices) {
  const int i = blockIdx.x;
  const int64_t row = load_index(rpi, is_i64, i);
  const int64_t start = kv_start ? load_index(kv_start, is_i64, i) : 0;
  const int64_t len = load_index(lens, is_i64, i);
  const int64_t out0 = kv_indptr[i];
  const int32_t* src = table + row * table_stride + start;
  for (int64_t j = static_cast<int64_t>(blockIdx.y) * blockDim.x + threadIdx.x; j < len;
       j += static_cast<int64_t>(gridDim.y) * blockDim.x)
    kv_indices[out0 + j] = src[j];
}

// dst[j] = pool[indices[j]] rows of row_bytes; one thread per 16 bytes
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint8_t* __restrict__ pool, int64_t tok_stride_bytes,
                                                          const int32_t* __restrict__ indices, int64_t n,
                                                          int chunks_per_row, uint8_t* __restrict__ dst) {
  const int64_t total = n * chunks_per_row;
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c < total;
       c += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t j = c / chunks_per_row;
    const int ch = static_cast<int>(c % chunks_per_row);
    const int64_t slot = indices[j];
    const uint4 v = *reinterpret_cast<const uint4*>(pool + slot * tok_stride_bytes + ch * 16);
    *reinterpret_cast<uint4*>(dst + c * 16) = v;
  }
}

// Debug check of a KV-write slot list: flags[0] = number of pairs t < t' with loc[t] == loc[t'] >= dummy_rows,
// flags[1] = number of entries outside [0, pool_rows).  One thread per entry scanning the entries after it: a
// validation call (O(T^2 / 2) compares of L1/L2-resident data), never on the hot path.
__global__ void __launch_bounds__(256) check_loc_kernel(const int64_t* __restrict__ loc, int T, int64_t dummy_rows,
                                                        int64_t pool_rows, int32_t* __restrict__ flags) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int64_t x = loc[t];
  if (x < 0 || x >= pool_rows) {
    atomicAdd(flags + 1, 1);
    return;
  }
  if (x < dummy_rows) return;  // the dummy page soaks up padded rows: duplicates there are the contract
  int dup = 0;
  for (int u = t + 1; u < T; ++u) dup += (loc[u] == x);
  if (dup) atomicAdd(flags, dup);
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_check_loc(const int64_t* loc, int T, int64_t dummy_rows, int64_t pool_rows,
                                            int32_t* flags, b200attn_stream_t stream) {
  B2_REQUIRE(T >= 0 && dummy_rows >= 0 && pool_rows > 0, B200ATTN_ERR_INVALID, "check_loc: bad sizes");
  B2_REQUIRE(flags != nullptr && (T == 0 || loc != nullptr), B200ATTN_ERR_INVALID, "check_loc: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  B2_CUDA(cudaMemsetAsync(flags, 0, 2 * sizeof(int32_t), st));
  if (T == 0) return B200ATTN_OK;
  check_loc_kernel<<<(T + 255) / 256, 256, 0, st>>>(loc, T, dummy_rows, pool_rows, flags);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

extern "C" B2_EXPORT int b200attn_flatten_kv_indices(const int32_t* req_to_token, int64_t table_stride,
                                           const void* req_pool_indices, const void* lens, const void* kv_start,
                                           int idx_is_i64, const int32_t* kv_indptr, int32_t* kv_indices, int bs,
                                           b200attn_stream_t stream) {
  B2_REQUIRE(bs >= 0, B200ATTN_ERR_INVALID, "flatten_kv_indices: bs=%d", bs);
  if (bs == 0) return B200ATTN_OK;
  B2_REQUIRE(req_to_token && req_pool_indices && lens && kv_indptr && kv_indices, B200ATTN_ERR_INVALID,
             "flatten_kv_indices: null pointer");
  dim3 grid(bs, 8, 1);
  flatten_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(req_to_token, table_stride, req_pool_indices,
                                                                      lens, kv_start, idx_is_i64, kv_indptr,
                                                                      kv_indices);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

extern "C" B2_EXPORT int b200attn_gather_kv(const void* pool, int64_t pool_tok_stride_bytes, const int32_t* indices, int64_t n,
                                  int64_t row_bytes, void* dst, b200attn_stream_t stream) {
  B2_REQUIRE(n >= 0 && row_bytes > 0, B200ATTN_ERR_INVALID, "gather_kv: bad sizes");
  if (n == 0) return B200ATTN_OK;
  B2_REQUIRE(pool && indices && dst, B200ATTN_ERR_INVALID, "gather_kv: null pointer");
  B2_REQUIRE(row_bytes % 16 == 0 && pool_tok_stride_bytes % 16 == 0 && aligned16(pool) && aligned16(dst),
             B200ATTN_ERR_INVALID, "gather_kv: rows must be 16-byte multiples and aligned");
  const int cpr = static_cast<int>(row_bytes / 16);
  const int64_t total = n * cpr;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 32;
  if (blocks > cap) blocks = cap;
  gather_rows_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint8_t*>(pool), pool_tok_stride_bytes, indices, n, cpr,
      reinterpret_cast<uint8_t*>(dst));
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}
