// banded.cu -- diagonal (DIA) storage for banded matrices with arbitrary coefficients.
//
// Not in the reference, whose only layout is CSR (mv_ops.h:17-23).  Opt-in, reported beside the CSR
// contract figures, never instead of them (like the row-pattern format of pattern.cu, which needs
// constant coefficients and refuses these matrices).
//
// A CSR block whose entries all lie in a band lo <= column - row <= hi that is mostly full -- the
// reference's own course matrix is of this kind: n = 52269, 345 entries per row, cg.c:231-232,260
// -- is re-stored as D = hi - lo + 1 diagonals of n values each, diagonal-major:
//     dia[d * ld + i] = A(i, i + lo + d)        (0.0 where CSR stores nothing or the column is outside)
// The matvec then streams 8 bytes per entry instead of 12, needs no index loads and no gathers:
// lane i of a warp owns row i of a block of 32 rows; for every diagonal the warp reads 256
// contiguous bytes of values and 256 contiguous bytes of the vector (p[i + lo + d], L1/L2 hits: p is
// small against the matrix).  U = 16 diagonals are in flight per thread before the first product.
// Row sums are formed in ascending d = ascending column order with unfused multiply and add, and a
// padding entry contributes +-0.0 to a partial sum that is never -0.0 (it starts at +0.0), i.e.
// nothing: s = A p is bit-identical to the CSR kernels' and to the reference's (for finite p; the
// reference itself multiplies the zeros of its dense-expanded row, mv_ops.c:187-197).
#include <stdio.h>

#include <string>

#include "device_utils.cuh"
#include "devmem.h"
#include "kernels.h"

namespace cgb {

namespace {

// lo / hi of (gather index of the entry) - (gather index of the row's diagonal) over all entries
__global__ void k_dia_span(const int* __restrict__ rowptr, const int* __restrict__ cols, int nrows,
                           long long off, int* __restrict__ out2) {
  const int stride = gridDim.x * blockDim.x;
  int mn = 0x7fffffff, mx = -0x7fffffff - 1;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < nrows; r += stride) {
    const int a = rowptr[r], b = rowptr[r + 1];
    if (b > a) {  // columns ascend within a row
      mn = min(mn, (int)((long long)cols[a] - off - r));
      mx = max(mx, (int)((long long)cols[b - 1] - off - r));
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out2 + 0, mn);
    atomicMax(out2 + 1, mx);
  }
}

__global__ void k_dia_fill(const int* __restrict__ rowptr, const int* __restrict__ cols,
                           const double* __restrict__ vals, int nrows, long long off, int lo,
                           long long ld, double* __restrict__ dia) {
  // a warp per row: consecutive entries of a row go to consecutive diagonals
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < nrows; r += warps) {
    const int a = rowptr[r], b = rowptr[r + 1];
    for (int k = a + lane; k < b; k += 32) {
      const long long d = (long long)cols[k] - off - r - lo;
      dia[d * ld + r] = vals[k];
    }
  }
}

template <int U, int THREADS>
__global__ void __launch_bounds__(THREADS)
k_spmv_dia(const double* __restrict__ dia, const long long ld, const int D, const int lo,
           const int nrows, const long long plen, const long long poff,
           const double* __restrict__ pg, const double* __restrict__ pown, double* __restrict__ s,
           double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
           const int mode, const __grid_constant__ PeerComm cm) {
  const int fin = mode & MV_FIN_MASK;
  if (fin != FIN_NONE && S->done) return;
  const uint64_t pol_keep = policy_evict_last();
  constexpr int WARPS = THREADS / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nblk = (nrows + 31) >> 5;
  double dot = 0.0, dot2 = 0.0;
#pragma unroll 1
  for (int blk = blockIdx.x * WARPS + warp; blk < nblk; blk += gridDim.x * WARPS) {
    const int row = blk * 32 + lane;
    const bool valid = row < nrows;
    const double* vp = dia + (valid ? row : 0);
    const long long p0 = poff + row + lo;  // gather index of this row's first diagonal
    double acc = 0.0;
    // U diagonals in flight per thread, then their products in ascending order.  (Measured on the
    // course-sized matrix: a second batch in flight, values double-buffered in registers, is SLOWER --
    // 37.0 against 32.3 us, 149 registers --, and so are 256-thread CTAs (39.0); 32- and 64-thread
    // CTAs are level with 128 (32.9 / 31.9).  At 145 MB per product the launch is ramp and tail
    // around ~22 us of streaming, like the CSR long-row kernel at this size.  16 rows per warp -- the
    // half-warps taking alternate diagonals, products handed over by shuffle, same sums -- doubles the
    // warps and the loads in flight and is slower still: 41.4 us with 16 loads per lane, 44.1 with 8.)
#pragma unroll 1
    for (int d = 0; d < D; d += U) {
      double v[U], pv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int dd = d + u;
        const long long idx = p0 + dd;
        const bool ok = valid & (dd < D);
        v[u] = ok ? __ldcs(vp + (long long)dd * ld) : 0.0;
        pv[u] = ld_keep_if(pg + (ok ? idx : 0), pol_keep, ok & (idx >= 0) & (idx < plen));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc = __dadd_rn(acc, __dmul_rn(v[u], pv[u]));  // (+-0.0 for padding)
    }
    if (valid) {
      s[row] = acc;
      if (fin != FIN_NONE) {
        const double g = pown[row];
        dot += g * acc;
        if (mode & MV_DUAL) dot2 += g * g;
      }
    }
  }
  if (fin == FIN_NONE) return;
  double v2[2] = {dot, dot2};
  bool last;
  if (mode & MV_DUAL) {
    last = grid_reduce<THREADS, 2>(v2, partials, ticket);
  } else {
    double v1[1] = {dot};
    last = grid_reduce<THREADS, 1>(v1, partials, ticket);
    v2[0] = v1[0];
    v2[1] = 0.0;
  }
  if (last) finish_block<THREADS>(S, cm, fin, v2[0], v2[1]);
}

}  // namespace

void free_banded(DiaDev* B) {
  dev_free(B->dia);
  *B = DiaDev();
}

// 1 = built, 0 = the matrix does not qualify (why says why; stays on CSR), -1 = CUDA failure.
// `off` = gather index of local row 0's diagonal (System::off).
int build_banded(const CsrDev& A, long long off, DiaDev* out, cudaStream_t st, std::string* why) {
  if (A.nrows <= 0 || A.nnz <= 0) {
    *why = "empty block";
    return 0;
  }
  int* d2 = nullptr;
  if (DEV_ALLOC(&d2, 2 * sizeof(int)) != cudaSuccess) return -1;
  const int init[2] = {0x7fffffff, -0x7fffffff - 1};
  int h2[2] = {0, -1};
  cudaError_t e = cudaMemcpyAsync(d2, init, sizeof(init), cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    k_dia_span<<<sm_count() * 4, 256, 0, st>>>(A.rowptr, A.cols, A.nrows, off, d2);
    add_launches(1);
    e = cudaMemcpyAsync(h2, d2, sizeof(h2), cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  dev_free(d2);
  if (e != cudaSuccess) return -1;
  const bool ok = h2[0] <= h2[1];
  if (!ok) {
    *why = "no stored entries";
    return 0;
  }
  const long long D = (long long)h2[1] - h2[0] + 1;
  const double fill = (double)A.nnz / ((double)D * (double)A.nrows);
  // worth it only when the band is mostly full: 8 D bytes per row against 12 per stored entry
  if (D > 8192 || fill < 0.70) {
    char msg[160];
    snprintf(msg, sizeof msg, "band of %lld diagonals is %.0f %% full (needs >= 70 %% and <= 8192 diagonals)", D,
             100.0 * fill);
    *why = msg;
    return 0;
  }
  DiaDev B;
  B.D = (int)D;
  B.lo = h2[0];
  B.ld = ((long long)A.nrows + 31) / 32 * 32;
  const size_t bytes = (size_t)D * (size_t)B.ld * sizeof(double);
  if (DEV_ALLOC(&B.dia, bytes) != cudaSuccess) {
    cudaGetLastError();
    *why = "not enough device memory for the diagonals";
    return 0;
  }
  if (cudaMemsetAsync(B.dia, 0, bytes, st) != cudaSuccess) return -1;
  k_dia_fill<<<sm_count() * 8, 256, 0, st>>>(A.rowptr, A.cols, A.vals, A.nrows, off, B.lo, B.ld, B.dia);
  add_launches(1);
  if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
  *out = B;
  return 1;
}

cudaError_t launch_spmv_dia(const DiaDev& B, int nrows, const double* pg, long long plen, long long poff,
                            const double* pown, double* s, double* partials, unsigned int* ticket,
                            CgState* S, int mode, const PeerComm& cm, cudaStream_t st) {
  if (nrows <= 0) return cudaSuccess;
  // CTAs of 4 warps, one 32-row block per warp (a persistent stride for very tall blocks)
  constexpr int T = 128;
  const int nblk = (nrows + 31) / 32;
  int grid = (nblk + T / 32 - 1) / (T / 32);
  if (grid > 4096) grid = 4096;  // partials per reduction (device_utils.cuh: MAX_GRID)
  k_spmv_dia<16, T><<<grid, T, 0, st>>>(B.dia, B.ld, B.D, B.lo, nrows, plen, poff, pg, pown, s, partials, ticket,
                                        S, mode, cm);
  add_launches(1);
  return cudaGetLastError();
}

}  // namespace cgb
