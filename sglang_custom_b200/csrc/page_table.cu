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

// In-place row move inside every K / V buffer of the pool: buf[tgt[i]] = buf[src[i]] for all i, all buffers, one launch.
// Reference: MHATokenToKVPool.move_kv_cache + copy_all_layer_kv_cache (mem_cache/memory_pool.py:460-468,1150-1184;
// caller speculative/eagle_utils.py:575-583).  That kernel loads a 128-byte column block of ALL source rows and only
// then stores it ("this copy is an inplace operation", :1170-1171), i.e. every read sees the pool as it was before the
// call even when a row is both a source and a target.  The same rule is kept here with more parallelism: a CTA owns one
// column slice of one buffer for all n rows, stages the n source slices in shared memory, and stores after a barrier.
// Distinct CTAs touch disjoint bytes, so there is no ordering between them to get wrong.
__global__ void __launch_bounds__(256) move_kv_kernel(const uint64_t* __restrict__ data_ptrs,
                                                      const int64_t* __restrict__ strides, const void* tgt_loc,
                                                      const void* src_loc, int is_i64, int n, int col_bytes) {
  extern __shared__ uint4 stage[];
  uint8_t* base = reinterpret_cast<uint8_t*>(data_ptrs[blockIdx.y]);
  const int64_t stride = strides[blockIdx.y];
  const int64_t col0 = static_cast<int64_t>(blockIdx.x) * col_bytes;
  if (col0 >= stride) return;
  const int lanes = col_bytes / 16;
  const int total = n * lanes;
  const bool vec = ((reinterpret_cast<uintptr_t>(base) | static_cast<uintptr_t>(stride)) & 15u) == 0;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int64_t off = col0 + (idx % lanes) * 16;
    if (off >= stride) continue;
    const uint8_t* p = base + load_index(src_loc, is_i64, idx / lanes) * stride + off;
    if (vec) {
      stage[idx] = *reinterpret_cast<const uint4*>(p);
    } else {
      uint8_t* s = reinterpret_cast<uint8_t*>(stage + idx);
      const int m = static_cast<int>(stride - off < 16 ? stride - off : 16);
      for (int b = 0; b < m; ++b) s[b] = p[b];
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int64_t off = col0 + (idx % lanes) * 16;
    if (off >= stride) continue;
    uint8_t* p = base + load_index(tgt_loc, is_i64, idx / lanes) * stride + off;
    if (vec) {
      *reinterpret_cast<uint4*>(p) = stage[idx];
    } else {
      const uint8_t* s = reinterpret_cast<const uint8_t*>(stage + idx);
      const int m = static_cast<int>(stride - off < 16 ? stride - off : 16);
      for (int b = 0; b < m; ++b) p[b] = s[b];
    }
  }
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_move_kv(const uint64_t* data_ptrs, const int64_t* data_strides, int n_buffers,
                                          int64_t max_stride_bytes, const void* tgt_loc, const void* src_loc,
                                          int idx_is_i64, int n, b200attn_stream_t stream) {
  B2_REQUIRE(n >= 0 && n_buffers >= 0 && max_stride_bytes > 0, B200ATTN_ERR_INVALID,
             "move_kv: bad sizes n=%d n_buffers=%d max_stride_bytes=%lld", n, n_buffers,
             static_cast<long long>(max_stride_bytes));
  if (n == 0 || n_buffers == 0) return B200ATTN_OK;
  B2_REQUIRE(data_ptrs && data_strides && tgt_loc && src_loc, B200ATTN_ERR_INVALID, "move_kv: null pointer");
  B2_REQUIRE(n_buffers <= 65535, B200ATTN_ERR_UNSUPPORTED, "move_kv: %d buffers > 65535", n_buffers);
  // widest column slice whose n staged rows fit the shared memory of a CTA (128 B = four full sectors per row)
  constexpr int64_t kSmemBudget = 200 * 1024;
  int col = 128;
  while (col > 16 && static_cast<int64_t>(n) * col > kSmemBudget) col >>= 1;
  B2_REQUIRE(static_cast<int64_t>(n) * col <= kSmemBudget, B200ATTN_ERR_UNSUPPORTED,
             "move_kv: %d rows do not fit one CTA's staging (max %lld); split the call", n,
             static_cast<long long>(kSmemBudget / 16));
  const size_t smem = static_cast<size_t>(n) * col;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(move_kv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(kSmemBudget)));
    configured = true;
  }
  dim3 grid(static_cast<unsigned>((max_stride_bytes + col - 1) / col), n_buffers, 1);
  move_kv_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(data_ptrs, data_strides, tgt_loc, src_loc,
                                                                         idx_is_i64, n, col);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

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
