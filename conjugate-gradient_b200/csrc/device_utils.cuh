// device_utils.cuh -- sm_100a building blocks shared by every kernel of the CG path:
// cache-hinted global loads/stores, mbarrier + 1-D bulk-copy (TMA) wrappers, and the
// deterministic block/grid reductions that replace the reference's sequential dot
// (reference mv_ops.c:117-132).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "state.h"

namespace cgb {

// ---------------------------------------------------------------------------------------------
// L2 policies and hinted accesses
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

// Gathered operand (the search direction): read-only path, keep in L2.
__device__ __forceinline__ double ld_keep(const double* ptr, uint64_t pol) {
  double v;
  asm("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(ptr), "l"(pol));
  return v;
}

// Streamed-once operand, 128 bit: bypass L1 allocation, evict-first in L2.
__device__ __forceinline__ double2 ld_stream2(const double* ptr, uint64_t pol) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;"
               : "=d"(v.x), "=d"(v.y)
               : "l"(ptr), "l"(pol));
  return v;
}

// Streamed-once operand, 256 bit (sm_100 LDG.256), evict-first in L2.  32-byte aligned.
struct __align__(32) double4_t {
  double x, y, z, w;
};
__device__ __forceinline__ double4_t ld_stream4(const double* ptr) {
  uint64_t a, b, c, d;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
               : "=l"(a), "=l"(b), "=l"(c), "=l"(d)
               : "l"(ptr));
  double4_t v;
  v.x = __longlong_as_double((long long)a);
  v.y = __longlong_as_double((long long)b);
  v.z = __longlong_as_double((long long)c);
  v.w = __longlong_as_double((long long)d);
  return v;
}

// ---------------------------------------------------------------------------------------------
// mbarrier + bulk async copy (global -> shared, completes on an mbarrier)
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)),
               "r"(bytes)
               : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_addr(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// bytes must be a multiple of 16; src and dst 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_addr(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar)), "l"(pol)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// Deterministic reductions
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the block; the result is valid in thread 0.  Fixed association order for a fixed
// block size.  `sm` holds at least THREADS/32 doubles.  Safe to call repeatedly.
template <int THREADS>
__device__ __forceinline__ double block_sum(double v, double* sm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect sm against a previous use
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) t += sm[w];
  }
  return t;
}

// Single-pass grid reduction: every block publishes one partial, takes a ticket, and the block
// that draws the last ticket sums all partials in index order (fixed order for a fixed grid, so
// results are run-to-run reproducible; no floating-point atomics).  Returns true in thread 0 of
// that last block only, with the grand total in `total`; the ticket is reset for the next launch.
template <int THREADS>
__device__ __forceinline__ bool grid_sum(double v, double* __restrict__ partials,
                                         unsigned int* __restrict__ ticket, double& total,
                                         const double* __restrict__ extra = nullptr,
                                         int n_extra = 0) {
  __shared__ double sm[THREADS / 32];
  __shared__ int is_last;
  double t = block_sum<THREADS>(v, sm);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = t;
    __threadfence();
    unsigned int tk = atomicAdd(ticket, 1u);
    is_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
  double acc = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += THREADS) acc += __ldcg(partials + i);
  // partials left by a preceding launch of the same reduction (split matvec), same fixed order
  for (int i = threadIdx.x; i < n_extra; i += THREADS) acc += __ldcg(extra + i);
  total = block_sum<THREADS>(acc, sm);
  if (threadIdx.x == 0) {
    *ticket = 0u;
    return true;
  }
  return false;
}

// ---------------------------------------------------------------------------------------------
// Scalar bookkeeping of the loop (reference cg.c:113,125-134), run by exactly one thread per
// iteration: either the last block of the fused kernel (one GPU) or a one-thread kernel after the
// cross-GPU sum.
// ---------------------------------------------------------------------------------------------

// After p.s is known (reference cg.c:113).  Breakdown policy: the reference divides regardless
// and poisons x with NaN once r is exactly zero; here the loop stops before the update.
__device__ __forceinline__ void finish_ps(CgState* S, double ps) {
  S->ps = ps;
  if (S->rr == 0.0) {
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  } else if (!(ps > 0.0) || isinf(ps)) {
    S->done = 1;
    S->status = CGB200_ST_BREAKDOWN;
  } else {
    S->alpha = S->rr / ps;
  }
}

// After the new r.r is known (reference cg.c:125-129).  `iter` counts x updates done.
__device__ __forceinline__ void finish_rr(CgState* S, double rr_new) {
  const int k = S->iter;  // the reference's k at the break test
  S->iter = k + 1;
  if (S->thresh > 0.0 && rr_new <= S->thresh) {
    S->done = 1;
    S->status = CGB200_ST_CONVERGED;
  } else if (k >= S->max_iter) {
    S->done = 1;
    S->status = CGB200_ST_MAXITER;
  } else {
    S->beta = rr_new / S->rr;
  }
  S->rr = rr_new;
}

// After b.b is known (loop entry, reference cg.c:104-108: x = 0, r = p = b).
__device__ __forceinline__ void finish_init(CgState* S, double bb) {
  S->bb = bb;
  S->rr = bb;
  S->thresh = S->tol * S->tol * bb;
  S->iter = 0;
  S->alpha = 0.0;
  S->beta = 0.0;
  S->ps = 0.0;
  S->done = 0;
  S->status = CGB200_ST_RUNNING;
  if (bb == 0.0) {  // b = 0: x = 0 is the solution; the reference would return NaN
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  }
}

}  // namespace cgb
