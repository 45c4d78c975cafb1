// Cluster-launch-control probe (sm_100a): a grid of N CTAs, each CTA that starts keeps cancelling not-yet-started CTAs
// of the same grid and does their work (here: mark the cancelled block's linear id).  Checks that every id is
// processed exactly once, counts how many CTAs really started and how many items each took.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o clc clc.cu && ./clc
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__global__ void __launch_bounds__(128, 1) clc_probe(int* marks, int* started, int* taken, int spin) {
  extern __shared__ __align__(16) uint8_t dyn[];  // forces 1 CTA / SM through its size
  __shared__ __align__(16) uint4 resp;
  __shared__ __align__(8) uint64_t bar;
  __shared__ int next[4];
  const uint32_t bar_a = smem_u32(&bar), resp_a = smem_u32(&resp);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    atomicAdd(started, 1);
  }
  __syncthreads();
  int bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z, n = 0;
  uint32_t phase = 0;
  while (true) {
    if (threadIdx.x == 0) {
      atomicAdd(&marks[(bz * gridDim.y + by) * gridDim.x + bx], 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 16;\n" ::"r"(bar_a) : "memory");
      asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];\n" ::"r"(resp_a), "r"(bar_a) : "memory");
    }
    // "work"
    long long t0 = clock64();
    while (clock64() - t0 < spin) {}
    ++n;
    if (threadIdx.x == 0) {
      uint32_t done = 0;
      while (!done) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(bar_a), "r"(phase) : "memory");
      }
      uint32_t ok, x, y, z;
      asm volatile(
          "{\n.reg .pred p;\n.reg .b128 r;\nld.shared.b128 r, [%4];\n"
          "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\nselp.u32 %0, 1, 0, p;\n"
          "mov.u32 %1, 0; mov.u32 %2, 0; mov.u32 %3, 0;\n"
          "@p clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%1, %2, %3, _}, r;\n}\n"
          : "=r"(ok), "=r"(x), "=r"(y), "=r"(z) : "r"(resp_a) : "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      next[0] = ok; next[1] = x; next[2] = y; next[3] = z;
    }
    phase ^= 1;
    __syncthreads();
    const int ok = next[0];
    bx = next[1]; by = next[2]; bz = next[3];
    __syncthreads();
    if (!ok) break;
  }
  if (threadIdx.x == 0) atomicMax(taken, n);
}

int main() {
  const dim3 grid(16, 32, 8);
  const int N = grid.x * grid.y * grid.z;
  int *marks, *started, *taken;
  cudaMalloc(&marks, N * 4); cudaMalloc(&started, 4); cudaMalloc(&taken, 4);
  cudaFuncSetAttribute(clc_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int spin : {2000, 60000}) {
    cudaMemset(marks, 0, N * 4); cudaMemset(started, 0, 4); cudaMemset(taken, 0, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    clc_probe<<<grid, 128, 200 * 1024>>>(marks, started, taken, spin);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    std::vector<int> h(N); int hs = 0, ht = 0;
    cudaMemcpy(h.data(), marks, N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&hs, started, 4, cudaMemcpyDeviceToHost); cudaMemcpy(&ht, taken, 4, cudaMemcpyDeviceToHost);
    int bad = 0; for (int v : h) bad += (v != 1);
    printf("spin %d: err=%s items=%d bad=%d ctas_started=%d max_items_per_cta=%d time=%.3f ms (ideal %.3f ms at 1.9 GHz)\n", spin,
           cudaGetErrorString(err), N, bad, hs, ht, ms, (double)N / 148 * spin / 1.9e6);
  }
  return 0;
}
