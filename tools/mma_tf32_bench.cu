// developer aid: throughput of legacy mma.sync.m16n8k8 tf32 on this GPU (FMA/clk/SM), to size the 3xTF32 generic path
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters) {
  float c[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 + 4, b1 = a0 + 5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  if (s == 123.f) out[0] = s;
}
int main() {
  float* d; cudaMalloc(&d, 4);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  for (int warps : {4, 8, 16, 32}) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    k<<<sms, warps * 32>>>(d, 100);
    cudaEventRecord(e0); k<<<sms, warps * 32>>>(d, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double mmas = (double)sms * warps * iters * 8;
    double fma = mmas * 16 * 8 * 8;
    printf("warps/SM %2d: %.1f TFLOP/s tf32 (mma.sync m16n8k8), %.1f FMA/clk/SM at 1.9 GHz, %.2f cycles per mma per SMSP\n", warps,
           2 * fma / (ms * 1e-3) / 1e12, fma / (ms * 1e-3) / sms / 1.9e9, (ms * 1e-3) * 1.9e9 / (iters * 8.0 * warps / 4));
  }
  return 0;
}
