// Cold vs warm instruction fetch on B200.  One warp per SM runs NB blocks of 2048 unique straight-line FFMAs
// (32 KB of SASS each) `passes` times.  Prints cycles per block and pass: (a) after a 256 MB L2 flush, (b) relaunched
// right away.  Tells: cost of streaming cold code, i-cache capacity (warm pass slow beyond it), retention across launches.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ long long tick(float& a) {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t), "+f"(a));
  return t;
}

template <int SEED> __device__ __forceinline__ void block(float& a, float& b, float& c, float& d) {
#pragma unroll
  for (int i = 0; i < 512; ++i) {
    a = fmaf(a, 1.0f + (SEED * 512 + i) * 1e-7f, 0.5f);
    b = fmaf(b, 1.0f + (SEED * 512 + i) * 2e-7f, 0.25f);
    c = fmaf(c, 1.0f + (SEED * 512 + i) * 3e-7f, 0.125f);
    d = fmaf(d, 1.0f + (SEED * 512 + i) * 4e-7f, 0.0625f);
  }
}

__global__ void k(float* out, long long* cyc, int passes, int nb) {
  float a = threadIdx.x, b = 1, c = 2, d = 3;
  for (int p = 0; p < passes; ++p) {
    long long t[9];
    for (int i = 0; i < 9; ++i) t[i] = 0;
    t[0] = tick(a);
    if (nb > 0) { block<0>(a, b, c, d); a += b + c + d; } t[1] = tick(a);
    if (nb > 1) { block<1>(a, b, c, d); a += b + c + d; } t[2] = tick(a);
    if (nb > 2) { block<2>(a, b, c, d); a += b + c + d; } t[3] = tick(a);
    if (nb > 3) { block<3>(a, b, c, d); a += b + c + d; } t[4] = tick(a);
    if (nb > 4) { block<4>(a, b, c, d); a += b + c + d; } t[5] = tick(a);
    if (nb > 5) { block<5>(a, b, c, d); a += b + c + d; } t[6] = tick(a);
    if (nb > 6) { block<6>(a, b, c, d); a += b + c + d; } t[7] = tick(a);
    if (nb > 7) { block<7>(a, b, c, d); a += b + c + d; } t[8] = tick(a);
    if (threadIdx.x == 0 && blockIdx.x == 0) for (int i = 0; i < 8; ++i) cyc[p * 8 + i] = t[i + 1] - t[i];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
}

__global__ void flush(float* p, size_t n) { for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = 0.f; }

int main() {
  float* out; long long* cyc; float* big;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 24 * 8); cudaMalloc(&big, 256u << 20);
  for (int nb = 1; nb <= 8; nb *= 2) {
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 0) flush<<<1184, 256>>>(big, (256u << 20) / 4);
      k<<<148, 32>>>(out, cyc, 3, nb);
      long long h[24]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      printf("nb=%d (%3d KB) %s\n", nb, nb * 32, mode ? "relaunched right away" : "after a 256 MB L2 flush");
      for (int p = 0; p < 3; ++p) { printf("  pass %d:", p); for (int i = 0; i < nb; ++i) printf(" %6lld", h[p * 8 + i]); printf("\n"); }
    }
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
