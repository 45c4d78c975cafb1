// Micro-benchmark of the CUDA-core pipes the softmax warps of the extend kernel live on (dev tool, not product).
// Prints cycles per warp-instruction per SMSP for 1, 2 and 4 warps per SMSP.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipes scripts/ubench/pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define REP 16   // independent chains
#define ITERS 256

template <int OP>
__global__ void k(float* out, long long* cyc, float a, float b) {
  float x[REP], y[REP];
#pragma unroll
  for (int i = 0; i < REP; ++i) { x[i] = a + i + threadIdx.x; y[i] = b * i + threadIdx.x; }
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < REP; i += 2) {
      if (OP == 0) {  // FFMA reg,reg,reg (2 per pair)
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y[i]), "f"(y[i + 1]));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 1) {  // FFMA2 pair,pair,pair
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {%3, %2};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 2) {  // FFMA2 pair, pair, imm
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0f3F000000, 0f3F000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 3) {  // FADD2 pair, pair
        asm volatile("{.reg .b64 ra, rb; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3};"
                     "add.rn.f32x2 ra, ra, rb; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 4) {  // FADD2 pair, imm
        asm volatile("{.reg .b64 ra, rb; mov.b64 ra, {%0, %1}; mov.b64 rb, {0f4B400000, 0f4B400000};"
                     "add.rn.f32x2 ra, ra, rb; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]));
      } else if (OP == 5) {  // MUFU.EX2 x2
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i + 1]));
      } else if (OP == 6) {  // FMNMX3
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y[i]), "f"(y[i + 1]));
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 7) {  // FMNMX x2
        asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i]));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i + 1]) : "f"(y[i + 1]));
      } else if (OP == 8) {  // F2FP pack (1 per pair) 
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x[i]), "f"(x[i + 1]));
        x[i] = __uint_as_float(r);
      } else if (OP == 9) {  // FFMA reg, reg, imm-ish: x = x * y + 0.5
        asm volatile("fma.rn.f32 %0, %0, %1, 0f3F000000;" : "+f"(x[i]) : "f"(y[i]));
        asm volatile("fma.rn.f32 %0, %0, %1, 0f3F000000;" : "+f"(x[i + 1]) : "f"(y[i + 1]));
      } else if (OP == 10) {  // mix: 2 MUFU + 1 FFMA2 per pair
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i + 1]));
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0f3F000000, 0f3F000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(y[i]), "+f"(y[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 11) {  // IMAD shift-add x2
        int r0 = __float_as_int(x[i]), r1 = __float_as_int(x[i + 1]);
        asm volatile("mad.lo.s32 %0, %1, 8388608, %0;" : "+r"(r0) : "r"(__float_as_int(y[i])));
        asm volatile("mad.lo.s32 %0, %1, 8388608, %0;" : "+r"(r1) : "r"(__float_as_int(y[i + 1])));
        x[i] = __int_as_float(r0); x[i + 1] = __int_as_float(r1);
      } else if (OP == 12) {  // mix: FFMA2 (imm) + FMNMX x2 (fma pipe + alu pipe)
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0f3F000000, 0f3F000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(y[i + 1]));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(y[i + 1]) : "f"(y[i]));
      } else if (OP == 13) {  // FMUL2 pair, pair
        asm volatile("{.reg .b64 ra, rb; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3};"
                     "mul.rn.f32x2 ra, ra, rb; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
      } else if (OP == 15) {  // MUFU.EX2 bf16x2 (1 per pair)
        uint32_t r = __float_as_uint(x[i]);
        asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r));
        x[i] = __uint_as_float(r);
      } else if (OP == 16) {  // MUFU.EX2 f16x2 (1 per pair)
        uint32_t r = __float_as_uint(x[i]);
        asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r));
        x[i] = __uint_as_float(r);
      } else if (OP == 17) {  // mix: FFMA2(imm) + F2FP
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0f3F000000, 0f3F000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y[i]), "f"(y[i + 1]));
        y[i] = __uint_as_float(r);
      } else if (OP == 18) {  // mix: 2 FMNMX + F2FP
        asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i + 1]));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i + 1]) : "f"(y[i]));
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y[i]), "f"(y[i + 1]));
        y[i] = __uint_as_float(r);
      } else if (OP == 19) {  // the emu-0 softmax mix per pair: FFMA2 + 2 MUFU + FADD2 + F2FP + FMNMX3
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0fBF000000, 0fBF000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i + 1]));
        asm volatile("{.reg .b64 ra, rb; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3};"
                     "add.rn.f32x2 ra, ra, rb; mov.b64 {%0, %1}, ra;}" : "+f"(y[i]), "+f"(y[i + 1]) : "f"(x[i]), "f"(x[i + 1]));
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x[i]), "f"(x[i + 1]));
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(y[i]) : "f"(x[i]), "f"(__uint_as_float(r)));
      } else if (OP == 20) {  // bf16x2-MUFU softmax mix per pair: FFMA2 + F2FP + MUFU.bf16x2 + 2 unpack + FADD2 + FMNMX3
        asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3}; mov.b64 rc, {0fBF000000, 0fBF000000};"
                     "fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0, %1}, ra;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(y[i]), "f"(y[i + 1]));
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x[i + 1]), "f"(x[i]));
        asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r));
        float lo, hi;
        asm volatile("shl.b32 %0, %1, 16;" : "=f"(lo) : "r"(r));
        asm volatile("and.b32 %0, %1, 0xffff0000;" : "=f"(hi) : "r"(r));
        asm volatile("{.reg .b64 ra, rb; mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3};"
                     "add.rn.f32x2 ra, ra, rb; mov.b64 {%0, %1}, ra;}" : "+f"(y[i]), "+f"(y[i + 1]) : "f"(lo), "f"(hi));
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(lo), "f"(hi));
      } else if (OP == 21) {  // F2FP e4m3x2 -> f16x2, x2
        uint32_t r0 = __float_as_uint(x[i]), r1 = __float_as_uint(x[i + 1]);
        asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %0; cvt.rn.f16x2.e4m3x2 %0, lo;}" : "+r"(r0));
        asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %0; cvt.rn.f16x2.e4m3x2 %0, lo;}" : "+r"(r1));
        x[i] = __uint_as_float(r0); x[i + 1] = __uint_as_float(r1);
      } else if (OP == 22) {  // PRMT x2
        uint32_t r0 = __float_as_uint(x[i]), r1 = __float_as_uint(x[i + 1]);
        asm volatile("prmt.b32 %0, %0, %1, 0x1404;" : "+r"(r0) : "r"(__float_as_uint(y[i])));
        asm volatile("prmt.b32 %0, %0, %1, 0x3424;" : "+r"(r1) : "r"(__float_as_uint(y[i + 1])));
        x[i] = __uint_as_float(r0); x[i + 1] = __uint_as_float(r1);
      } else if (OP == 23) {  // LOP3 x2
        uint32_t r0 = __float_as_uint(x[i]), r1 = __float_as_uint(x[i + 1]);
        asm volatile("lop3.b32 %0, %0, %1, 0x80808080, 0xE4;" : "+r"(r0) : "r"(__float_as_uint(y[i])));
        asm volatile("lop3.b32 %0, %0, %1, 0x80808080, 0xE4;" : "+r"(r1) : "r"(__float_as_uint(y[i + 1])));
        x[i] = __uint_as_float(r0); x[i + 1] = __uint_as_float(r1);
      } else if (OP == 24) {  // mix: 2 MUFU + 2 F2FP e4m3 unpack (same pipe?)
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i + 1]));
        uint32_t r0 = __float_as_uint(y[i]), r1 = __float_as_uint(y[i + 1]);
        asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %0; cvt.rn.f16x2.e4m3x2 %0, lo;}" : "+r"(r0));
        asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %0; cvt.rn.f16x2.e4m3x2 %0, lo;}" : "+r"(r1));
        y[i] = __uint_as_float(r0); y[i + 1] = __uint_as_float(r1);
      } else if (OP == 25) {  // SHF x2
        uint32_t r0 = __float_as_uint(x[i]), r1 = __float_as_uint(x[i + 1]);
        asm volatile("shr.b32 %0, %0, 1;" : "+r"(r0));
        asm volatile("shl.b32 %0, %0, 7;" : "+r"(r1));
        x[i] = __uint_as_float(r0); x[i + 1] = __uint_as_float(r1);
      } else if (OP == 14) {  // FADD reg, reg x2
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i]));
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i + 1]) : "f"(y[i + 1]));
      }
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < REP; ++i) s += x[i] + y[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int OP>
void run(const char* name, int instr_per_pair) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  printf("%-34s", name);
  for (int wps = 1; wps <= 4; wps *= 2) {
    k<OP><<<148, 128 * wps>>>(out, cyc, 1.0f, 0.001f);
    k<OP><<<148, 128 * wps>>>(out, cyc, 1.0f, 0.001f);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double per = double(c) / (double(ITERS) * (REP / 2) * instr_per_pair * wps);
    printf("  %dw/SMSP: %6.2f cyc/instr", wps, per);
  }
  printf("   (%d instr per element pair)\n", instr_per_pair);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("FFMA r,r,r", 2);
  run<9>("FFMA r,r,imm", 2);
  run<1>("FFMA2 rr,rr,rr", 1);
  run<2>("FFMA2 rr,rr,imm", 1);
  run<3>("FADD2 rr,rr", 1);
  run<4>("FADD2 rr,imm", 1);
  run<13>("FMUL2 rr,rr", 1);
  run<14>("FADD r,r", 2);
  run<5>("MUFU.EX2", 2);
  run<6>("FMNMX3", 2);
  run<7>("FMNMX", 2);
  run<8>("F2FP.BF16.PACK", 1);
  run<11>("IMAD shl-add", 2);
  run<15>("MUFU.EX2 bf16x2", 1);
  run<16>("MUFU.EX2 f16x2", 1);
  run<17>("mix FFMA2(imm) + F2FP", 2);
  run<18>("mix 2 FMNMX + F2FP", 3);
  run<19>("softmax mix emu0 (6 instr/pair)", 6);
  run<20>("softmax mix bf16x2 MUFU (7/pair)", 7);
  run<21>("F2FP e4m3x2->f16x2", 2);
  run<22>("PRMT", 2);
  run<23>("LOP3", 2);
  run<25>("SHF", 2);
  run<24>("mix 2 MUFU + 2 F2FP.e4m3", 4);
  run<10>("mix 2 MUFU + 1 FFMA2(imm)", 3);
  run<12>("mix 1 FFMA2(imm) + 2 FMNMX", 3);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
