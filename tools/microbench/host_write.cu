// SM-issued writes to pinned host memory over PCIe: what does the shape of the pieces cost?
// One "env" owns a 3168-byte image ((12, 66) fp32); a step writes the 792 bytes of its 3 live rows, minus -- on the pinned
// host path of libchs -- the 16-byte aligned interior of each row's dead column range.  Every mode below issues bulk
// (TMA) shared->global stores from one lane per env, 16 envs per CTA, like step_kernel does.
//   mode 0  the pieces libchs writes today (4 per env, 64 + 176 + 176 + 136 = 552 bytes, no alignment beyond 16)
//   mode 1  the 3 live rows whole (792 bytes, base = e * 3168: 32 mod 64 for odd e)
//   mode 2  mode 1 widened to 64-byte lines  (the bytes added are padding rows: constant zeros under the contract)
//   mode 3  mode 1 widened to 128-byte lines
//   mode 4  mode 1 widened to 256-byte lines
//   mode 5  mode 0 with every piece widened to 64-byte lines (overlapping pieces merged)
//   mode 6  a dense (E, 792) destination (what a compact (D, L) observation contract writes)
//   mode 7  mode 6 with the copy engine instead (cudaMemcpyAsync device->host of E * 792 bytes), for the link's own rate
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o host_write host_write.cu ; run: ./host_write [envs]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, unsigned bytes) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(sa), "r"(bytes) : "memory");
}

constexpr int G = 16, IMG = 3168, ROW = 264, LIVE = 792, DEAD_LO = 56, DEAD_HI = 136;   // bytes

__global__ void __launch_bounds__(256) writer(char* dst, int E, int mode, unsigned long long* bytes_out) {
  extern __shared__ __align__(256) char sm[];          // G slots of 2048 bytes, content irrelevant (zeros)
  for (int i = threadIdx.x; i < G * 2048 / 4; i += blockDim.x) reinterpret_cast<int*>(sm)[i] = 0;
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const int t = threadIdx.x;
  const int ow = t - (blockDim.x - 32);
  unsigned long long n = 0;
  if (ow >= 0 && ow < G) {
    const int e = blockIdx.x * G + ow;
    if (e < E) {
      const char* src = sm + ow * 2048;
      const long long base = (mode == 6) ? (long long)e * 800 : (long long)e * IMG;   // 800: the next multiple of 16
      auto put = [&](long long lo, long long hi) {       // [lo, hi) absolute byte range, 16-byte aligned
        if (lo < 0) lo = 0;
        bulk_store(dst + lo, src + (lo & 255), (unsigned)(hi - lo));
        n += hi - lo;
      };
      if (mode == 0 || mode == 5) {
        long long plo[4], phi[4];
        int np = 0;
        long long cur = 0;
        for (int d = 0; d < 3; ++d) {
          const long long a = (d * ROW + DEAD_LO + 15) & ~15, b = (d * ROW + DEAD_HI) & ~15;
          plo[np] = base + cur; phi[np] = base + a; ++np;
          cur = b;
        }
        plo[np] = base + cur; phi[np] = base + LIVE; ++np;
        if (mode == 5) {
          for (int i = 0; i < np; ++i) { plo[i] &= ~63ll; phi[i] = (phi[i] + 63) & ~63ll; }
          int m = 0;
          for (int i = 1; i < np; ++i) { if (plo[i] <= phi[m]) phi[m] = phi[i]; else { ++m; plo[m] = plo[i]; phi[m] = phi[i]; } }
          np = m + 1;
        }
        for (int i = 0; i < np; ++i) put(plo[i], phi[i]);
      } else {
        const long long al = mode == 2 ? 64 : mode == 3 ? 128 : mode == 4 ? 256 : 16;
        put(base & ~(al - 1), (base + (mode == 6 ? 800 : LIVE) + al - 1) & ~(al - 1));
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  }
  if (bytes_out && n) atomicAdd(bytes_out, n);
}

int main(int argc, char** argv) {
  const int E = argc > 1 ? atoi(argv[1]) : 4096;
  const int reps = 200;
  char* host;
  CK(cudaHostAlloc(&host, (size_t)E * IMG + 4096, cudaHostAllocDefault));
  char* dev_dst;                       // the same patterns into HBM, for comparison
  CK(cudaMalloc(&dev_dst, (size_t)E * IMG + 4096));
  unsigned long long* cnt;
  CK(cudaMalloc(&cnt, 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = (E + G - 1) / G, smem = G * 2048;
  printf("E = %d envs, %d CTAs x 256 threads, %d launches per mode, one lane per env issues the bulk stores\n", E, grid, reps);
  for (int where = 0; where < 2; ++where) {
    char* dst = where == 0 ? host : dev_dst;
    for (int mode = 0; mode <= 7; ++mode) {
      if (mode == 7) {
        if (where == 1) continue;
        for (int i = 0; i < 10; ++i) CK(cudaMemcpyAsync(host, dev_dst, (size_t)E * LIVE, cudaMemcpyDeviceToHost));
        CK(cudaEventRecord(e0));
        for (int i = 0; i < reps; ++i) CK(cudaMemcpyAsync(host, dev_dst, (size_t)E * LIVE, cudaMemcpyDeviceToHost));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("host  mode 7 (copy engine, dense %d B/env): %7.2f us per copy, %6.1f GB/s\n", LIVE, ms * 1e3 / reps, (double)E * LIVE * reps / ms / 1e6);
        continue;
      }
      CK(cudaMemset(cnt, 0, 8));
      writer<<<grid, 256, smem>>>(dst, E, mode, cnt);
      CK(cudaDeviceSynchronize());
      unsigned long long nb; CK(cudaMemcpy(&nb, cnt, 8, cudaMemcpyDeviceToHost));
      for (int i = 0; i < 10; ++i) writer<<<grid, 256, smem>>>(dst, E, mode, nullptr);
      CK(cudaEventRecord(e0));
      for (int i = 0; i < reps; ++i) writer<<<grid, 256, smem>>>(dst, E, mode, nullptr);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      printf("%s mode %d: %6.1f B/env, %7.2f us per launch, %6.1f GB/s\n", where == 0 ? "host " : "hbm  ", mode, (double)nb / E, ms * 1e3 / reps,
             (double)nb * reps / ms / 1e6);
    }
  }
  return 0;
}
