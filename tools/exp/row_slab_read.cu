// Micro-benchmark: does HBM deliver its streaming rate when a warp reads its data as 16 interleaved
// row streams of 256 + 128 bytes per step (the long-row SpMV's access pattern: 16 rows of a CSR block
// in lockstep, 32 entries of each per step) instead of one contiguous run?  Not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o row_slab_read row_slab_read.cu && ./row_slab_read
// Buffer: `rows` rows of `len` doubles (values) and `len` ints (indices), back to back like CSR.
// MODE 0: a warp owns 16 consecutive rows and reads 32 entries of each per step (one 8-byte and one
//         4-byte load per lane and entry, ld.global.cs).   MODE 1: the same bytes read contiguously.
// MODE 2: pairs -- lanes 0-15 take row 2i, lanes 16-31 row 2i+1, a 128-bit + a 64-bit load per lane.
// Results on one B200 (profiles/r02_ncu_longrow.md): 2274 / 2982 / 5826 GB/s at 2 CTAs per SM.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                     \
  do {                                                            \
    cudaError_t e = (x);                                          \
    if (e != cudaSuccess) {                                       \
      printf("%s: %s\n", #x, cudaGetErrorString(e));              \
      exit(1);                                                    \
    }                                                             \
  } while (0)

template <int MODE>
__global__ void __launch_bounds__(256) k(const double* __restrict__ v, const int* __restrict__ c, int rows, int len,
                                        double* out) {
  const int lane = threadIdx.x & 31;
  const int W = gridDim.x * 8, gw = blockIdx.x * 8 + (threadIdx.x >> 5);
  double acc = 0.0;
  for (int blk = gw; blk < rows / 16; blk += W) {
    const long long base = (long long)blk * 16 * len;
    if (MODE == 0) {
      for (int off = 0; off < len; off += 32) {
        double vv[16];
        int cc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int kk = off + lane;
          vv[i] = kk < len ? __ldcs(v + base + (long long)i * len + kk) : 0.0;
          cc[i] = kk < len ? __ldcs(c + base + (long long)i * len + kk) : 0;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) acc += vv[i] + (double)cc[i];
      }
    } else if (MODE == 1) {
      const long long tot = 16LL * len;
      for (long long off = 0; off < tot; off += 512) {
        double vv[16];
        int cc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const long long kk = off + i * 32 + lane;
          vv[i] = kk < tot ? __ldcs(v + base + kk) : 0.0;
          cc[i] = kk < tot ? __ldcs(c + base + kk) : 0;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) acc += vv[i] + (double)cc[i];
      }
    } else {
      const int hw = lane >> 4, q2 = (lane & 15) * 2;
      for (int off = 0; off < len; off += 32) {
        double2 vv[8];
        int2 cc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int kk = off + q2;
          const long long e = (base + (long long)(2 * i + hw) * len + kk) & ~1LL;
          vv[i] = kk < len ? __ldcs(reinterpret_cast<const double2*>(v + e)) : make_double2(0.0, 0.0);
          cc[i] = kk < len ? __ldcs(reinterpret_cast<const int2*>(c + e)) : make_int2(0, 0);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) acc += vv[i].x + vv[i].y + (double)(cc[i].x + cc[i].y);
      }
    }
  }
  if (acc == 1.2345e300) out[0] = acc;
}

template <int MODE>
static double run(const double* v, const int* c, int rows, int len, int grid, double* out) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) k<MODE><<<grid, 256>>>(v, c, rows, len, out);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 20; ++i) k<MODE><<<grid, 256>>>(v, c, rows, len, out);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return 12.0 * rows * (double)len * 20 / (ms * 1e-3) / 1e9;
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  double* out;
  CK(cudaMalloc(&out, 64));
  const int len = 344;
  for (int rows : {52272, 209088}) {
    double* v;
    int* c;
    CK(cudaMalloc(&v, (size_t)rows * len * 8 + 1024));
    CK(cudaMalloc(&c, (size_t)rows * len * 4 + 1024));
    CK(cudaMemset(v, 0, (size_t)rows * len * 8 + 1024));
    CK(cudaMemset(c, 0, (size_t)rows * len * 4 + 1024));
    for (int bps : {2, 4})
      printf("len %d rows %d (%.0f MB) %d CTAs/SM: row-interleaved %.0f | contiguous %.0f | pairs %.0f GB/s\n", len, rows,
             12.0 * rows * len / 1e6, bps, run<0>(v, c, rows, len, sms * bps, out), run<1>(v, c, rows, len, sms * bps, out),
             run<2>(v, c, rows, len, sms * bps, out));
    CK(cudaFree(v));
    CK(cudaFree(c));
  }
  return 0;
}
