// Micro-benchmark: tcgen05.mma issue cost and execution rate on B200 (dev tool, not product).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sglang_custom_b200/csrc -o scripts/ubench/umma scripts/ubench/umma.cu
// One CTA per SM; `issuers` warps each issue `count` MMAs (M = 128, N = template, K = 16) back to back from one
// elected lane, commit, and wait.  Reports cycles per MMA as seen by the issuing thread (issue) and until the
// commit barrier fires (exec), for SS (A, B from shared memory) and TS (A from tensor memory) operands.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "common.cuh"
#include "tc_ptx.cuh"

using namespace b200attn;

template <int N, bool kTS>
__global__ void __launch_bounds__(128, 1) k(long long* out, int count, int issuers, int bg_copy) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem + 160 * 1024;
  const uint32_t slot = bars + 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) ptx::tmem_alloc<512>(slot);
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) ptx::mbar_init(bars + 8 * i, 1);
    ptx::fence_mbar_init();
  }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0;
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(slot));
  long long t_issue = 0, t_exec = 0;
  if (warp < issuers) {
    const bool leader = ptx::elect_one();
    constexpr uint32_t idesc = ptx::instr_desc_f16(1, 128, N, 0);
    const uint64_t ad = ptx::smem_desc_sw128(smem + warp * 32768, 16, 1024);
    const uint64_t bd = ptx::smem_desc_sw128(smem + 65536 + warp * 32768, 16, 1024);
    const uint32_t d = tmem_base + warp * 256;       // accumulator columns of this issuer
    const uint32_t a_t = tmem_base + warp * 256 + 128;
    __syncwarp();
    const long long t0 = clock64();
    if (leader) {
      for (int i = 0; i < count; ++i) {
        const uint64_t off = static_cast<uint64_t>((i & 3) * 2);
        if (kTS) ptx::umma_ts(d, a_t + (i & 3) * 8, bd + off, idesc, 1u);
        else ptx::umma_ss(d, ad + off, bd + off, idesc, 1u);
      }
      ptx::umma_commit(bars + 8 * warp);
    }
    __syncwarp();
    const long long t1 = clock64();
    ptx::mbar_wait(bars + 8 * warp, 0);
    const long long t2 = clock64();
    t_issue = t1 - t0; t_exec = t2 - t0;
  } else if (bg_copy && warp == 3) {
    // background shared-memory write traffic (what the K/V producers do): 16-byte stores, 64 KiB per `count/32` MMAs
    uint4 z = make_uint4(1, 2, 3, 4);
    for (int rep = 0; rep < bg_copy; ++rep)
      for (int i = lane; i < 4096; i += 32) reinterpret_cast<uint4*>(smem_raw + 96 * 1024)[i] = z;
  }
  __syncthreads();
  if (blockIdx.x == 0 && lane == 0 && warp < issuers) { out[2 * warp] = t_issue; out[2 * warp + 1] = t_exec; }
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem_base); }
}

template <int N, bool kTS>
void run(int issuers, int bg) {
  long long* out; cudaMalloc(&out, 64);
  auto kern = k<N, kTS>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int count = 512;
  for (int r = 0; r < 2; ++r) kern<<<148, 128, 200 * 1024>>>(out, count, issuers, bg);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[8]; cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost);
  printf("M128 N%-3d %s issuers=%d bg=%d: issue %6.1f cyc/MMA, exec %6.1f cyc/MMA (nominal %d)%s\n", N, kTS ? "TS" : "SS", issuers, bg,
         double(h[0]) / count, double(h[1]) / count / 1.0, 128 * N / 256, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(out);
}

int main() {
  run<64, false>(1, 0); run<128, false>(1, 0); run<256, false>(1, 0);
  run<64, true>(1, 0); run<128, true>(1, 0); run<256, true>(1, 0);
  run<64, false>(2, 0); run<128, false>(2, 0);
  run<64, true>(2, 0); run<128, true>(2, 0);
  // background 16-byte shared-memory stores from one warp: bg x 64 KiB during 512 MMAs
  run<128, false>(1, 16); run<128, false>(1, 32); run<128, false>(1, 64); run<128, false>(1, 128);
  run<128, true>(1, 32); run<128, true>(1, 64); run<64, false>(1, 32);
  return 0;
}
