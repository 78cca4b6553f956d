// Is instruction fetch from the SM-level cache a per-SM bandwidth limit?  W warps of one CTA per SM walk 4 blocks of
// 2048 unique FFMAs (128 KB of SASS in total, warm), each warp starting at a different block.  Prints cycles per
// block per warp-count: flat => latency-bound prefetch, growing with W => shared fetch bandwidth.
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
__global__ void k(float* out, long long* cyc, int rounds, int same) {
  float a = threadIdx.x, b = 1, c = 2, d = 3;
  const int w = threadIdx.x >> 5;
  long long t0 = 0, t1 = 0;
  for (int r = 0; r < rounds; ++r) {
    if (r == 1) t0 = tick(a);
    for (int i = 0; i < 4; ++i) {
      const int blk = same ? i : ((i + w) & 3);
      if (blk == 0) block<0>(a, b, c, d);
      else if (blk == 1) block<1>(a, b, c, d);
      else if (blk == 2) block<2>(a, b, c, d);
      else block<3>(a, b, c, d);
    }
  }
  t1 = tick(a);
  if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) cyc[w] = (t1 - t0) / (4 * (rounds - 1));
  out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
}
int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 32 * 8);
  for (int same = 0; same < 2; ++same)
    for (int W = 1; W <= 16; W *= 2) {
      k<<<148, 32 * W>>>(out, cyc, 5, same);
      long long h[32]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < W; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("%s  warps/SM=%2d  cycles per 2048-instruction block: warp0 %6lld  max %6lld  => %.2f warp-instr/cycle/SM\n",
             same ? "same block order " : "staggered blocks ", W, h[0], mx, 2048.0 * W / mx);
    }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
