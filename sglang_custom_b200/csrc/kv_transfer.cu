// HiCache KV page transfer: copy token/page items between two KV pools -- device <-> pinned host
// (mapped, read or written by the kernel over PCIe / C2C) or device <-> device -- for one layer or all layers, between
// the layer-first, page-first and per-layer-pointer-table layouts.
// Reference: sgl-kernel/csrc/kvcacheio/transfer.cu:9-175 (transfer_item_warp, the three offset functions,
// transfer_kernel_impl + launcher) and the entry points :177-430; callers mem_cache/memory_pool_host.py and
// mem_cache/memory_pool.py:350-389.
//
// Byte mover, bound by HBM (device <-> device) or by the host link (~50-60 GB/s per direction on PCIe 5).  Design:
// the whole job is flattened into 16-byte chunks -- chunk g belongs to unit g / chunks_per_item, a unit being one
// (item, layer, K|V) triple -- so consecutive threads always touch consecutive addresses whatever the item size
// (the reference gives one warp a whole item with 8-byte accesses, which leaves lanes idle for items < 256 B and
// keeps 256 B in flight per warp).  Every thread keeps eight independent chunks in flight (all loads, then all
// stores); the grid is capped at `block_quota` CTAs when the caller wants to leave SMs to a concurrent forward
// (the reference's parameter of the same name), else sized to the device.
#include "common.cuh"

namespace b200attn {

struct TransferParams {
  const uint8_t* src[2];
  uint8_t* dst[2];
  const int64_t* src_idx;
  const int64_t* dst_idx;
  int64_t n_items, item_bytes;
  int64_t src_dim, dst_dim;
  int src_layout, dst_layout;
  int start_layer, num_layers;
  int n_bufs;  // 2 = K and V, 1 = a single buffer (MLA)
};

__device__ __forceinline__ const uint8_t* item_addr(const uint8_t* base, int layout, int64_t dim, int layer, int64_t index,
                                                    int64_t item_bytes) {
  if (layout == B200ATTN_LAYOUT_LAYER_FIRST) return base + layer * dim + index * item_bytes;
  if (layout == B200ATTN_LAYOUT_PAGE_FIRST) return base + index * dim + layer * item_bytes;
  // per-layer base pointers (the pool keeps one tensor per layer: memory_pool.py:231-255)
  return reinterpret_cast<const uint8_t*>(reinterpret_cast<const uint64_t*>(base)[layer]) + index * item_bytes;
}

template <typename V>
__device__ __forceinline__ V ld_stream(const V* p);
template <>
__device__ __forceinline__ uint4 ld_stream<uint4>(const uint4* p) { return ld_nc_16(p); }
template <>
__device__ __forceinline__ uint2 ld_stream<uint2>(const uint2* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];\n" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(uint4* p, uint4 v) {
  asm volatile("st.global.cg.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream(uint2* p, uint2 v) {
  asm volatile("st.global.cg.v2.u32 [%0], {%1, %2};\n" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// 1024 threads x 8 chunks x 16 B = 128 KiB in flight per CTA: the reference's default quota of two CTAs is then enough
// to fill a PCIe 5 x16 link (measured: 2 CTAs of 256 threads x 4 chunks reach 4.6 GB/s host -> device, the link does 51)
constexpr int kTransferThreads = 1024;
constexpr int kTransferInFlight = 8;

template <typename V>
__global__ void __launch_bounds__(kTransferThreads) transfer_kernel(const TransferParams p) {
  constexpr int kInFlight = kTransferInFlight;
  const int64_t cpi = p.item_bytes / static_cast<int64_t>(sizeof(V));  // chunks per item
  const int64_t units_per_item = static_cast<int64_t>(p.num_layers) * p.n_bufs;
  const int64_t total = p.n_items * units_per_item * cpi;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t g0 = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; g0 < total; g0 += kInFlight * stride) {
    V val[kInFlight];
    V* out[kInFlight];
#pragma unroll
    for (int u = 0; u < kInFlight; ++u) {
      const int64_t g = g0 + u * stride;
      out[u] = nullptr;
      if (g < total) {
        const int64_t unit = g / cpi, c = g % cpi;
        const int64_t item = unit / units_per_item;
        const int r = static_cast<int>(unit % units_per_item);
        const int layer = p.start_layer + r / p.n_bufs, b = r % p.n_bufs;
        // (ternaries, not p.src[b]: a runtime index into a kernel-parameter array forces a local-memory copy of it)
        const uint8_t* s = item_addr(b ? p.src[1] : p.src[0], p.src_layout, p.src_dim, layer, p.src_idx[item], p.item_bytes);
        const uint8_t* d = item_addr(b ? p.dst[1] : p.dst[0], p.dst_layout, p.dst_dim, layer, p.dst_idx[item], p.item_bytes);
        val[u] = ld_stream<V>(reinterpret_cast<const V*>(s) + c);
        out[u] = const_cast<V*>(reinterpret_cast<const V*>(d)) + c;
      }
    }
#pragma unroll
    for (int u = 0; u < kInFlight; ++u)
      if (out[u] != nullptr) st_stream(out[u], val[u]);
  }
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_transfer_kv(const void* src_k, void* dst_k, const void* src_v, void* dst_v, int src_layout,
                                              int dst_layout, int64_t src_layout_dim, int64_t dst_layout_dim,
                                              const int64_t* src_indices, const int64_t* dst_indices, int64_t n_items,
                                              int64_t item_bytes, int start_layer, int num_layers, int block_quota,
                                              b200attn_stream_t stream) {
  B2_REQUIRE(n_items >= 0 && num_layers >= 0 && start_layer >= 0, B200ATTN_ERR_INVALID,
             "transfer_kv: bad sizes n_items=%lld layers=[%d,+%d)", static_cast<long long>(n_items), start_layer, num_layers);
  if (n_items == 0 || num_layers == 0) return B200ATTN_OK;
  B2_REQUIRE(src_k && dst_k && src_indices && dst_indices, B200ATTN_ERR_INVALID, "transfer_kv: null pointer");
  B2_REQUIRE((src_v == nullptr) == (dst_v == nullptr), B200ATTN_ERR_INVALID, "transfer_kv: src_v and dst_v must both be given or both be null");
  // the reference's check: "Item byte size must be divisible by 8" (transfer.cu:142)
  B2_REQUIRE(item_bytes > 0 && item_bytes % 8 == 0, B200ATTN_ERR_INVALID, "transfer_kv: item byte size %lld must be divisible by 8",
             static_cast<long long>(item_bytes));
  for (int l : {src_layout, dst_layout})
    B2_REQUIRE(l == B200ATTN_LAYOUT_LAYER_FIRST || l == B200ATTN_LAYOUT_PAGE_FIRST || l == B200ATTN_LAYOUT_LAYER_TABLE,
               B200ATTN_ERR_INVALID, "transfer_kv: layout %d", l);
  B2_REQUIRE(src_layout_dim % 8 == 0 && dst_layout_dim % 8 == 0, B200ATTN_ERR_INVALID, "transfer_kv: layout dims must be multiples of 8 bytes");

  TransferParams p;
  p.src[0] = static_cast<const uint8_t*>(src_k); p.src[1] = static_cast<const uint8_t*>(src_v);
  p.dst[0] = static_cast<uint8_t*>(dst_k); p.dst[1] = static_cast<uint8_t*>(dst_v);
  p.src_idx = src_indices; p.dst_idx = dst_indices;
  p.n_items = n_items; p.item_bytes = item_bytes;
  p.src_dim = src_layout_dim; p.dst_dim = dst_layout_dim;
  p.src_layout = src_layout; p.dst_layout = dst_layout;
  p.start_layer = start_layer; p.num_layers = num_layers;
  p.n_bufs = src_v ? 2 : 1;

  // 16-byte accesses when every address the kernel forms is 16-byte aligned; pointer tables cannot be checked on the
  // host (they live on the device), so those keep the reference's 8-byte granularity unless the caller's item size and
  // layout dims are multiples of 16 and the table entries are tensor bases (torch allocations are >= 256-byte aligned)
  auto ok16 = [&](const void* base, int layout, int64_t dim) {
    return (layout == B200ATTN_LAYOUT_LAYER_TABLE) || (aligned16(base) && dim % 16 == 0);
  };
  const bool v16 = item_bytes % 16 == 0 && ok16(src_k, src_layout, src_layout_dim) && ok16(dst_k, dst_layout, dst_layout_dim) &&
                   (!src_v || (ok16(src_v, src_layout, src_layout_dim) && ok16(dst_v, dst_layout, dst_layout_dim)));
  const int64_t chunk = v16 ? 16 : 8;
  const int64_t total = n_items * num_layers * p.n_bufs * (item_bytes / chunk);
  const int64_t per_cta = static_cast<int64_t>(kTransferThreads) * kTransferInFlight;
  int64_t blocks = (total + per_cta - 1) / per_cta;
  const int64_t cap = block_quota > 0 ? block_quota : static_cast<int64_t>(sm_count()) * 2;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (v16) transfer_kernel<uint4><<<static_cast<unsigned>(blocks), kTransferThreads, 0, st>>>(p);
  else transfer_kernel<uint2><<<static_cast<unsigned>(blocks), kTransferThreads, 0, st>>>(p);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}
