// Throughput of MUFU.TANH / MUFU.EX2 / MUFU.RCP per SM per clock (micro-benchmark, not part of the library).
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(float* out, int iters) {
  float a[8];
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(a[i]));
      if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (OP == 2) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
      if (OP == 5) {
        unsigned r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a[i]), "f"(a[(i + 1) & 7]));
        a[i] = __uint_as_float(r | 0x3f000000u);
      }
      if (OP == 6) {
        asm volatile("shfl.sync.idx.b32 %0, %0, 3, 31, 0xffffffff;" : "+f"(a[i]));
      }
      if (OP == 4) {
        unsigned long long& p = *reinterpret_cast<unsigned long long*>(&a[i & ~1]);
        asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(p));
      }
    }
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char* name) {
  float* out;
  cudaMalloc(&out, 148 * 2 * 1024 * 4);
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  k<OP><<<148 * 2, 1024>>>(out, iters);
  cudaEventRecord(e0);
  k<OP><<<148 * 2, 1024>>>(out, iters);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double ops = 148.0 * 2 * 1024 * 8.0 * iters;
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%s: %.3f ms, %.2f Gop/s, %.2f ops/clk/SM at %d MHz nominal\n", name, ms, ops / ms * 1e-6,
         ops / (ms * 1e-3) / 148.0 / (clk * 1e3), clk / 1000);
  cudaFree(out);
}
int main() {
  run<0>("tanh.approx");
  run<1>("ex2.approx");
  run<2>("rcp.approx");
  run<3>("fma");
  run<4>("fma.f32x2 (instructions; x2 for FMAs)");
  run<5>("cvt.rn.bf16x2.f32 (+1 LOP3)");
  run<6>("shfl.idx");
  return 0;
}
