// Tensor maps for the paged-tile (page_size >= 16) K/V loads: host-side encoding through the driver entry point
// (no libcuda link: cudaGetDriverEntryPoint) and the device-side cp.async.bulk.tensor wrappers.
//
// With page_size > 1 the slots of a page are consecutive (mem_cache/allocator.py:444-447 in the reference:
// `out_pages[:, None] * page_size + arange(page_size)`), and a request's position p sits at
// req_to_token[r, p] = page(p / page_size) * page_size + p % page_size -- so the 16 rows of a 16-token unit are ONE
// box of the pool tensor [rows][Hkv * D]: one table read and one bulk-tensor copy instead of 16 row gathers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200attn {

// Pool viewed as a 3-D tensor for 128-byte-swizzled boxes:
//   dim0 = the 128 bytes of one swizzle row (128 / elem_bytes elements),
//   dim1 = pool rows (token slots), stride pool_row_bytes,
//   dim2 = 128-byte segments of a row (Hkv * D * elem_bytes / 128), stride 128 bytes.
// A box [128 B, box_rows, box_segs] lands in shared memory as [seg][row][128 B] with the XOR swizzle keyed on the row,
// which is what ldmatrix wants (8 rows of one 16-byte column hit 8 different bank groups) and, for box_segs = 1, what
// the UMMA K-major / MN-major SWIZZLE_128B descriptors of the extend kernel expect.
// Returns 0 on success.  The encoding is pure host work (no stream, no sync): safe under CUDA-graph capture.
int encode_pool_tmap(CUtensorMap* out, const void* pool, int elem_bytes, int64_t pool_row_bytes, int row_segs,
                     int box_rows, int box_segs);

#ifdef __CUDACC__
namespace ptx {

// 3-D tiled load, completion on an mbarrier (complete_tx::bytes)
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tmap, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cta.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
          "r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}

}  // namespace ptx
#endif

}  // namespace b200attn
