#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k(const uint8_t* in, uint32_t* out) {
  __shared__ __align__(128) uint8_t sm[2][16][16];
  int lane = threadIdx.x;
  for (int i = lane; i < 512; i += 32) ((uint8_t*)sm)[i] = in[i];
  __syncwarp();
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(&sm[lane >> 4][lane & 15][0]);
  uint32_t r0, r1, r2, r3;
  asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];\n" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
  out[lane * 4 + 0] = r0; out[lane * 4 + 1] = r1; out[lane * 4 + 2] = r2; out[lane * 4 + 3] = r3;
}
__global__ void mm(float* out, int iters) {
  uint32_t a[4] = {threadIdx.x, 2, 3, 4}, b0 = 5, b1 = 6;
  float c[8][4] = {};
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  float s = 0; for (int j = 0; j < 8; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() { return 0; }
