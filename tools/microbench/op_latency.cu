// Dependent-chain latency (cycles per op) of the operations the per-env scalar phase uses, one warp, B200.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o op_latency op_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double c_tab[136];

template <int OP> __global__ void chain(double* out, long long* cyc, int n, double seed, int idx0) {
  double x = seed + threadIdx.x * 1e-9, y = seed * 0.5;
  float f = (float)seed + threadIdx.x * 1e-6f;
  int idx = idx0;
  extern __shared__ double sh[];
  for (int i = threadIdx.x; i < 136; i += blockDim.x) sh[i] = 1.0 + i * 1e-9;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    if (OP == 0) x = fma(x, 1.0000001, y);                       // DFMA
    if (OP == 1) x = x + y;                                      // DADD
    if (OP == 2) x = sqrt(x + 1.0);                              // double sqrt
    if (OP == 3) x = y / (x + 1.0);                              // double div
    if (OP == 4) { f = (float)x; x = (double)f + y; }            // F2F both ways + DADD
    if (OP == 5) f = fmaf(f, 1.0000001f, 0.5f);                  // FFMA
    if (OP == 6) f = rsqrtf(f + 1.0f);                           // MUFU.RSQ + FADD
    if (OP == 7) { x = c_tab[idx]; idx = (idx + (int)x) % 136; } // LDC indexed, dependent
    if (OP == 8) { x = sh[idx]; idx = (idx + (int)x) % 136; }    // LDS dependent
    if (OP == 9) f = __expf(f) * 0.5f;                           // ex2 path
    if (OP == 10) f = sqrtf(f + 1.0f);                           // IEEE sqrtf
    if (OP == 11) f = 1.0f / (f + 1.0f);                         // IEEE div
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  out[threadIdx.x] = x + f + idx;
}

int main() {
  double h[136];
  for (int i = 0; i < 136; ++i) h[i] = 1.0 + i * 1e-9;
  cudaMemcpyToSymbol(c_tab, h, sizeof(h));
  double* out; long long* cyc;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
  const char* names[] = {"DFMA", "DADD", "dsqrt(+DADD)", "ddiv(+DADD)", "F2F.64->32->64 + DADD", "FFMA", "rsqrtf+FADD",
                         "LDC indexed (+F2I,mod)", "LDS (+F2I,mod)", "__expf*0.5", "sqrtf+FADD", "1/(f+1)"};
  const int n = 4096;
#define RUN(OP, TH)                                                                   \
  { for (int r = 0; r < 2; ++r) chain<OP><<<1, TH, 136 * 8>>>(out, cyc, n, 1.5, 3);   \
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);                      \
    printf("%-26s threads=%4d  %7.1f cycles/iter\n", names[OP], TH, (double)c / n); }
  RUN(0, 32) RUN(1, 32) RUN(2, 32) RUN(3, 32) RUN(4, 32) RUN(5, 32) RUN(6, 32) RUN(7, 32) RUN(8, 32) RUN(9, 32) RUN(10, 32) RUN(11, 32)
  // throughput view: 8 warps on one SM (2 per scheduler)
  RUN(0, 256) RUN(2, 256) RUN(3, 256) RUN(5, 256) RUN(6, 256)
  RUN(0, 1024) RUN(5, 1024) RUN(6, 1024)
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
