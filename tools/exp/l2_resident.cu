// Micro-benchmark: how much of a buffer that is streamed once per iteration can be KEPT in the
// 126 MB L2 across iterations when its head is loaded with an evict-last priority and the rest with
// evict-first?  (The CG loop re-reads the same matrix every iteration; a resident head is DRAM
// traffic that never happens again.)  Not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_resident l2_resident.cu && ./l2_resident
// Each CTA owns a contiguous slice of the buffer; the first `frac` of every slice is the resident
// part (so the CTAs stay balanced).  Prints effective GB/s of `reps` back-to-back sweeps.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e = (x);                                                       \
    if (e != cudaSuccess) {                                                    \
      printf("%s: %s\n", #x, cudaGetErrorString(e));                           \
      exit(1);                                                                 \
    }                                                                          \
  } while (0)

struct __align__(32) d4 { double x, y, z, w; };

__device__ __forceinline__ d4 ld_first(const double* p) {
  d4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ d4 ld_last(const double* p) {
  d4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_last.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ d4 ld_normal(const double* p) {
  d4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}

// MODE 0: everything evict_first.  1: head evict_last, tail evict_first.  2: head evict_normal
// (default policy), tail evict_first.  3: everything default policy.
template <int T, int MODE>
__global__ void __launch_bounds__(T, 4) sweep(const double* __restrict__ a, long long n4, long long head4,
                                              int rev, double* out) {
  // slice of this CTA, in units of 4 doubles
  const long long s0 = n4 * blockIdx.x / gridDim.x, s1 = n4 * (blockIdx.x + 1) / gridDim.x;
  const long long h1 = s0 + head4;  // resident part of the slice: [s0, h1)
  double acc = 0.0;
  const long long len = s1 - s0;
#pragma unroll 4
  for (long long k = threadIdx.x; k < len; k += T) {
    const long long i = rev ? s1 - 1 - k : s0 + k;
    d4 v;
    if (MODE == 0) v = ld_first(a + 4 * i);
    else if (MODE == 3) v = ld_normal(a + 4 * i);
    else if (i < h1) v = (MODE == 1) ? ld_last(a + 4 * i) : ld_normal(a + 4 * i);
    else v = ld_first(a + 4 * i);
    acc += v.x + v.y + v.z + v.w;
  }
  if (acc == 1.2345e300) out[0] = acc;
}

template <int MODE>
static double run(const double* a, long long n4, long long head_bytes, int grid, int reps, int alt, double* out) {
  const long long head4 = head_bytes / 32 / grid;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 4; ++i) sweep<256, MODE><<<grid, 256>>>(a, n4, head4, alt ? (i & 1) : 0, out);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < reps; ++i) sweep<256, MODE><<<grid, 256>>>(a, n4, head4, alt ? (i & 1) : 0, out);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)n4 * 32.0 * reps / (ms * 1e-3) / 1e9;
}

int main(int argc, char** argv) {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  int l2 = 0, maxpersist = 0;
  CK(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0));
  CK(cudaDeviceGetAttribute(&maxpersist, cudaDevAttrMaxPersistingL2CacheSize, 0));
  printf("SMs %d  L2 %d MB  max persisting L2 %d MB\n", sms, l2 >> 20, maxpersist >> 20);
  const int grid = sms * 4;
  double* out;
  CK(cudaMalloc(&out, 64));
  const long long sizes_mb[] = {217, 1074, 8600};
  for (int pass = 0; pass < 2; ++pass) {
    if (pass == 1) {
      CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)maxpersist));
      size_t got = 0;
      CK(cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize));
      printf("--- with cudaLimitPersistingL2CacheSize = %zu MB\n", got >> 20);
    }
    for (long long mb : sizes_mb) {
      const long long bytes = mb * 1000000ll / 32 * 32;
      double* a;
      CK(cudaMalloc(&a, bytes));
      CK(cudaMemset(a, 0, bytes));
      const long long n4 = bytes / 32;
      const int reps = mb > 2000 ? 6 : 30;
      printf("buffer %lld MB: all-evict_first %.0f GB/s | all-default %.0f | default, alternating direction %.0f\n", mb,
             run<0>(a, n4, 0, grid, reps, 0, out), run<3>(a, n4, 0, grid, reps, 0, out),
             run<3>(a, n4, 0, grid, reps, 1, out));
      const long long heads[] = {16, 32, 48, 64, 80, 96, 112};
      for (long long h : heads) {
        if (h * 1000000ll > bytes) continue;
        const double g1 = run<1>(a, n4, h * 1000000ll, grid, reps, 0, out);
        const double g2 = run<2>(a, n4, h * 1000000ll, grid, reps, 0, out);
        // ideal if the head never touched DRAM again and the tail streamed at the all-evict_first rate
        printf("  head %3lld MB: evict_last head %.0f GB/s | default-policy head %.0f GB/s\n", h, g1, g2);
      }
      CK(cudaFree(a));
    }
  }
  return 0;
}
