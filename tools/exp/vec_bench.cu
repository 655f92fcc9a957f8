// Micro-benchmark for the K2 (x += a p; r -= a s; sum r^2) access pattern: which launch shape and
// access width gets closest to the HBM copy peak.  Not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o vec_bench vec_bench.cu && ./vec_bench [n]
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

struct __align__(32) d4 { double x, y, z, w; };

__device__ __forceinline__ d4 ld4(const double* p) {
  d4 v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ d4 ld4s(const double* p) {  // streamed: no L1 allocation, evict-first
  d4 v;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st4(double* p, d4 v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}
__device__ __forceinline__ void st4s(double* p, d4 v) {
  asm volatile("st.global.L2::evict_first.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}

// MODE 0: grid-stride, 128-bit (the library's shape).  1: grid-stride 256-bit.  2: 256-bit + stream hints.
// 3: contiguous slab per CTA, 256-bit + hints.
template <int T, int U, int MODE>
__global__ void __launch_bounds__(T) k2(double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
                                        const double* __restrict__ s, long long n, double alpha, double* out) {
  double acc = 0.0;
  if (MODE == 0) {
    const long long n2 = n / 2, stride = (long long)gridDim.x * T;
    for (long long i = (long long)blockIdx.x * T + threadIdx.x; i < n2; i += stride * U) {
      double2 xv[U], rv[U], pv[U], sv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = i + u * stride;
        if (j < n2) {
          xv[u] = *(const double2*)(x + 2 * j); pv[u] = *(const double2*)(p + 2 * j);
          rv[u] = *(const double2*)(r + 2 * j); sv[u] = *(const double2*)(s + 2 * j);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = i + u * stride;
        if (j < n2) {
          xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x)); xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
          rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x)); rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
          *(double2*)(x + 2 * j) = xv[u]; *(double2*)(r + 2 * j) = rv[u];
          acc = fma(rv[u].x, rv[u].x, acc); acc = fma(rv[u].y, rv[u].y, acc);
        }
      }
    }
  } else {
    const long long n4 = n / 4;
    long long i0, i1, step;
    if (MODE == 3) {
      const long long per = (n4 + gridDim.x - 1) / gridDim.x;
      i0 = per * blockIdx.x + threadIdx.x; i1 = min(n4, per * (blockIdx.x + 1)); step = T;
    } else {
      i0 = (long long)blockIdx.x * T + threadIdx.x; i1 = n4; step = (long long)gridDim.x * T;
    }
    for (long long i = i0; i < i1; i += step * U) {
      d4 xv[U], rv[U], pv[U], sv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = i + u * step;
        if (j < i1) {
          if (MODE == 1) { xv[u] = ld4(x + 4 * j); pv[u] = ld4(p + 4 * j); rv[u] = ld4(r + 4 * j); sv[u] = ld4(s + 4 * j); }
          else { xv[u] = ld4s(x + 4 * j); pv[u] = ld4s(p + 4 * j); rv[u] = ld4s(r + 4 * j); sv[u] = ld4s(s + 4 * j); }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = i + u * step;
        if (j < i1) {
          xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x)); xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
          xv[u].z = __dadd_rn(xv[u].z, __dmul_rn(alpha, pv[u].z)); xv[u].w = __dadd_rn(xv[u].w, __dmul_rn(alpha, pv[u].w));
          rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x)); rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
          rv[u].z = __dsub_rn(rv[u].z, __dmul_rn(alpha, sv[u].z)); rv[u].w = __dsub_rn(rv[u].w, __dmul_rn(alpha, sv[u].w));
          if (MODE == 1) { st4(x + 4 * j, xv[u]); st4(r + 4 * j, rv[u]); } else { st4s(x + 4 * j, xv[u]); st4s(r + 4 * j, rv[u]); }
          acc = fma(rv[u].x, rv[u].x, acc); acc = fma(rv[u].y, rv[u].y, acc);
          acc = fma(rv[u].z, rv[u].z, acc); acc = fma(rv[u].w, rv[u].w, acc);
        }
      }
    }
  }
  // cheap stand-in for the grid reduction
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc == 12345.678) out[0] = acc;
}

template <int T, int U, int MODE>
static void run(const char* name, int grid, double* x, double* r, double* p, double* s, long long n, double* out) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) k2<T, U, MODE><<<grid, T>>>(x, r, p, s, n, 1e-9, out);
  cudaEventRecord(a);
  const int reps = 20;
  for (int i = 0; i < reps; ++i) k2<T, U, MODE><<<grid, T>>>(x, r, p, s, n, 1e-9, out);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  cudaError_t e = cudaGetLastError();
  printf("%-44s grid %5d  %8.1f us  %7.0f GB/s  %s\n", name, grid, ms / reps * 1e3, 48.0 * n / (ms / reps * 1e-3) / 1e9,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char** argv) {
  const long long n = argc > 1 ? atoll(argv[1]) : 512ll * 512 * 512;
  double *x, *r, *p, *s, *out;
  cudaMalloc(&x, n * 8); cudaMalloc(&r, n * 8); cudaMalloc(&p, n * 8); cudaMalloc(&s, n * 8); cudaMalloc(&out, 64);
  cudaMemset(x, 0, n * 8); cudaMemset(r, 0, n * 8); cudaMemset(p, 0, n * 8); cudaMemset(s, 0, n * 8);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("n = %lld (%.2f GB per launch)\n", n, 48.0 * n / 1e9);
  run<512, 4, 0>("library shape: 128-bit grid-stride T512 U4", sms * 4, x, r, p, s, n, out);
  run<512, 2, 1>("256-bit grid-stride T512 U2", sms * 4, x, r, p, s, n, out);
  run<512, 2, 2>("256-bit + stream hints T512 U2", sms * 4, x, r, p, s, n, out);
  run<256, 2, 2>("256-bit + stream hints T256 U2 x8/SM", sms * 8, x, r, p, s, n, out);
  run<256, 4, 2>("256-bit + stream hints T256 U4 x4/SM", sms * 4, x, r, p, s, n, out);
  run<512, 2, 3>("slab per CTA, 256-bit + hints T512 U2", sms * 2, x, r, p, s, n, out);
  run<256, 2, 3>("slab per CTA, 256-bit + hints T256 U2 x4/SM", sms * 4, x, r, p, s, n, out);
  run<256, 4, 3>("slab per CTA, 256-bit + hints T256 U4 x2/SM", sms * 2, x, r, p, s, n, out);
  run<1024, 1, 2>("256-bit + stream hints T1024 U1 x2/SM", sms * 2, x, r, p, s, n, out);
  run<256, 1, 2>("256-bit + stream hints T256 U1 x16/SM", sms * 16, x, r, p, s, n, out);
  return 0;
}
