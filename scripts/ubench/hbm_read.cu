// Read-bandwidth ceilings of the access patterns the decode kernel has to live with (dev tool, not product).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/hbm_read scripts/ubench/hbm_read.cu
// Patterns over a 4 GiB buffer (far beyond the 126 MB L2), read-only, 16-byte loads, result xor-reduced:
//   stream        : fully coalesced grid-stride read
//   rows R bytes  : rows of R contiguous bytes at RANDOM row indices (a permutation), one warp-wide run of
//                   32 x 16 B per 512 B; R = 2048 is a bf16 token row of 8 heads x 128 (C2), R = 128 an FP8 row of
//                   one head (C4 at TP 8), R = 256 / 512 / 1024 the shapes in between
// For each pattern: 148 x {1,2,4} CTAs of 512 threads, 8 independent loads in flight per thread.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
#include <random>

__device__ __forceinline__ uint4 ldnc(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// row_bytes == 0: streaming.  Otherwise chunk c (16 B) of logical position lives in row perm[pos / cpr], chunk pos % cpr.
template <int UNROLL>
__global__ void read_kernel(const uint4* __restrict__ buf, const int* __restrict__ perm, long long n_chunks, int cpr,
                            unsigned* out) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  unsigned acc = 0;
  for (; i + (UNROLL - 1) * stride < n_chunks; i += UNROLL * stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long pos = i + u * stride;
      const long long src = cpr ? static_cast<long long>(__ldg(perm + pos / cpr)) * cpr + pos % cpr : pos;
      v[u] = ldnc(buf + src);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) out[0] = acc;  // never true in practice; keeps the loads alive
}

int main() {
  const long long bytes = 4ll << 30;
  const long long n_chunks = bytes / 16;
  uint4* buf;
  unsigned* out;
  cudaMalloc(&buf, bytes);
  cudaMalloc(&out, 4);
  cudaMemset(buf, 1, bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int row_sizes[] = {0, 2048, 1024, 512, 256, 128};
  for (int rb : row_sizes) {
    int* perm = nullptr;
    int cpr = rb / 16;
    if (rb) {
      const long long rows = bytes / rb;
      std::vector<int> h(rows);
      for (long long r = 0; r < rows; ++r) h[r] = static_cast<int>(r);
      std::mt19937_64 g(1234);
      std::shuffle(h.begin(), h.end(), g);
      cudaMalloc(&perm, rows * 4);
      cudaMemcpy(perm, h.data(), rows * 4, cudaMemcpyHostToDevice);
    }
    for (int mult = 1; mult <= 4; mult *= 2) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        read_kernel<8><<<148 * mult, 512>>>(buf, perm, n_chunks, cpr, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0) best = std::min(best, ms);
      }
      if (rb)
        printf("rows of %4d B at random: %d CTAs/SM  %7.3f ms  %7.1f GB/s\n", rb, mult, best, bytes / best / 1e6);
      else
        printf("stream (coalesced)     : %d CTAs/SM  %7.3f ms  %7.1f GB/s\n", mult, best, bytes / best / 1e6);
    }
    if (perm) cudaFree(perm);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
