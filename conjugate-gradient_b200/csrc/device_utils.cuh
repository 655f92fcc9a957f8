// device_utils.cuh -- sm_100a building blocks shared by every kernel of the CG path:
// cache-hinted global loads/stores, mbarrier + 1-D bulk-copy (TMA) wrappers, and the
// deterministic block/grid reductions that replace the reference's sequential dot
// (reference mv_ops.c:117-132).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "comm.h"
#include "kernels.h"
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

// Same, predicated inside the instruction (0.0 when !ok): no branch around the load, so a row's
// gathers stay in one basic block and can all be in flight together.
__device__ __forceinline__ double ld_keep_if(const double* ptr, uint64_t pol, bool ok) {
  double v;
  asm("{\n"
      ".reg .pred p;\n"
      "setp.ne.s32 p, %3, 0;\n"
      "mov.f64 %0, 0d0000000000000000;\n"
      "@p ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;\n"
      "}\n"
      : "=d"(v)
      : "l"(ptr), "l"(pol), "r"((int)ok));
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

// Operand that other CTAs (or peer GPUs) rewrite between two reads of the SAME launch (the search
// direction inside the persistent dense solve kernel): 256 bit, served by L2, never by this SM's L1,
// which is not coherent and is only invalidated at launch boundaries.
__device__ __forceinline__ double4_t ld_l2_4(const double* ptr) {
  double4_t v;
  asm volatile("ld.global.cg.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(ptr));
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

// Asks L2 to fetch [src, src + bytes) (multiple of 16, 16-byte aligned); nothing is written.
__device__ __forceinline__ void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
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

// Single-pass grid reduction of NV values at once: every block publishes one partial per value,
// takes a ticket, and the block that draws the last ticket sums all partials in index order
// (fixed order for a fixed grid, so results are run-to-run reproducible; no floating-point
// atomics).  Returns a block-uniform "this is the last block"; the grand totals are left in v[]
// of thread 0 of that block, and the ticket is reset for the next launch.  Value k's partials live
// at partials + k * PART_STRIDE; `extra` (same layout) holds n_extra partials left by a preceding
// launch of the same reduction (split matvec), folded in in the same fixed order.
constexpr int PART_STRIDE = 8192;
constexpr int MAX_GRID = 4096;

template <int THREADS, int NV>
__device__ __forceinline__ bool grid_reduce(double (&v)[NV], double* __restrict__ partials,
                                            unsigned int* __restrict__ ticket,
                                            const double* __restrict__ extra = nullptr,
                                            int n_extra = 0, int nparts = -1) {
  // nparts >= 0: only CTAs 0 .. nparts-1 call this (the others have no share in the sum and go
  // straight to the barrier that follows): nparts tickets, nparts partials
  const int np = nparts < 0 ? (int)gridDim.x : nparts;
  __shared__ double sm[THREADS / 32];
  __shared__ int is_last;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const double t = block_sum<THREADS>(v[k], sm);
    if (threadIdx.x == 0) partials[k * PART_STRIDE + blockIdx.x] = t;
  }
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int tk = atomicAdd(ticket, 1u);
    is_last = (tk == (unsigned int)np - 1);
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double acc = 0.0;
    for (int i = threadIdx.x; i < np; i += THREADS)
      acc += __ldcg(partials + k * PART_STRIDE + i);
    for (int i = threadIdx.x; i < n_extra; i += THREADS) acc += __ldcg(extra + k * PART_STRIDE + i);
    v[k] = block_sum<THREADS>(acc, sm);
  }
  if (threadIdx.x == 0) *ticket = 0u;
  return true;
}

// One-value convenience form used by the standalone operators: true in thread 0 of the last block.
template <int THREADS>
__device__ __forceinline__ bool grid_sum(double v, double* __restrict__ partials,
                                         unsigned int* __restrict__ ticket, double& total) {
  double a[1] = {v};
  const bool last = grid_reduce<THREADS, 1>(a, partials, ticket);
  total = a[0];
  return last && threadIdx.x == 0;
}

// ---------------------------------------------------------------------------------------------
// Peer-memory exchange (comm.h): system-scope accesses, the in-kernel cross-GPU sum, the
// gather-vector push and the wait on its flags.
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(double* p, double v) {
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
// Gathered operand in a boundary tile: the halo was written by a peer while this kernel may
// already have been running, so go to L2 (the coherence point for peer writes), never L1.
__device__ __forceinline__ double ld_l2(const double* p) {
  double v;
  asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double ld_l2_if(const double* p, bool ok) {
  double v;
  asm volatile("{\n"
               ".reg .pred p;\n"
               "setp.ne.s32 p, %2, 0;\n"
               "mov.f64 %0, 0d0000000000000000;\n"
               "@p ld.relaxed.gpu.global.f64 %0, [%1];\n"
               "}\n"
               : "=d"(v)
               : "l"(p), "r"((int)ok));
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr unsigned long long PEER_TIMEOUT_NS = 4000000000ull;  // a hung peer must not hang this GPU

// Spins until *flag satisfies (exact ? == want : >= want).  false on time-out.
__device__ __forceinline__ bool spin_until(const unsigned long long* flag, unsigned long long want,
                                           bool exact) {
  unsigned long long t0 = 0;
  for (unsigned int it = 0;; ++it) {
    const unsigned long long f = ld_acquire_sys(flag);
    if (exact ? f == want : f >= want) return true;
    if ((it & 255u) == 255u) {
      const unsigned long long now = global_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > PEER_TIMEOUT_NS) return false;
    }
  }
}

__device__ __forceinline__ void comm_failed(CgState* S) {
  S->done = 1;
  S->status = CGB200_ST_COMM;
}

// Sum of (t0, t1) over all ranks, formed inside the calling kernel.  Called by every thread of
// ONE block per GPU (the last block of a fused reduction) with the local totals valid in thread
// 0; thread g < world stores this rank's pair into rank g's mailbox and then collects rank g's
// pair from the local mailbox (flagged words, comm.h: RedSlot); thread 0 adds the G pairs in rank
// order, so every rank obtains the same bits and takes the same branches.  Slots are indexed by the reduction number modulo 4: a
// peer can be at most one reduction ahead (it needs this rank's next contribution to go further).
__device__ __forceinline__ void st_ll2(unsigned long long* p, unsigned long long a,
                                       unsigned long long b) {
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_ll2(const unsigned long long* p, unsigned long long& a,
                                       unsigned long long& b) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}

template <int THREADS>
__device__ __forceinline__ bool peer_allreduce(const PeerComm& cm, CgState* S, double& t0,
                                               double& t1) {
  static_assert(THREADS >= MAX_RANKS, "block too small");
  __shared__ double sh[2 * MAX_RANKS + 2];
  __shared__ unsigned long long sh_epoch;
  __shared__ int sh_fail;
  if (threadIdx.x == 0) {
    const unsigned long long e = cm.local->red_epoch + 1;
    cm.local->red_epoch = e;
    sh_epoch = e;
    sh[2 * MAX_RANKS] = t0;
    sh[2 * MAX_RANKS + 1] = t1;
    sh_fail = 0;
  }
  __syncthreads();
  const unsigned long long e = sh_epoch;
  const int slot = (int)(e & 3ull);
  if ((int)threadIdx.x < cm.world) {
    const int g = threadIdx.x;
    const unsigned long long flag = (e % 0xffffffffull + 1ull) << 32;  // never 0 (fresh memory)
    const unsigned long long a = (unsigned long long)__double_as_longlong(sh[2 * MAX_RANKS]);
    const unsigned long long b = (unsigned long long)__double_as_longlong(sh[2 * MAX_RANKS + 1]);
    RedSlot* dst = &cm.mbox[g]->red[slot][cm.rank];
    st_ll2(&dst->w[0], flag | (a & 0xffffffffull), flag | (a >> 32));
    st_ll2(&dst->w[2], flag | (b & 0xffffffffull), flag | (b >> 32));
    const RedSlot* src = &cm.mbox[cm.rank]->red[slot][g];
    unsigned long long w0, w1, w2, w3, tstart = 0;
    bool ok = false;
    for (unsigned int it = 0;; ++it) {
      ld_ll2(&src->w[0], w0, w1);
      ld_ll2(&src->w[2], w2, w3);
      const unsigned long long hi = 0xffffffff00000000ull;
      if ((w0 & hi) == flag && (w1 & hi) == flag && (w2 & hi) == flag && (w3 & hi) == flag) {
        ok = true;
        break;
      }
      if ((it & 255u) == 255u) {
        const unsigned long long now = global_ns();
        if (tstart == 0) tstart = now;
        else if (now - tstart > PEER_TIMEOUT_NS) break;
      }
    }
    if (ok) {
      sh[2 * g] = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
      sh[2 * g + 1] = __longlong_as_double((long long)((w2 & 0xffffffffull) | (w3 << 32)));
    } else {
      sh_fail = 1;
    }
  }
  __syncthreads();
  const bool ok = sh_fail == 0;
  if (threadIdx.x == 0) {
    if (ok) {
      double a = 0.0, b = 0.0;
      for (int g = 0; g < cm.world; ++g) {
        a += sh[2 * g];
        b += sh[2 * g + 1];
      }
      t0 = a;
      t1 = b;
    } else {
      comm_failed(S);
    }
  }
  return ok;
}

// Element j of the owned block of the gather vector has just been computed: store it into the
// halo slots of every peer that gathers it.
__device__ __forceinline__ void push_elem(const PeerComm& cm, long long j, double v) {
  for (int t = 0; t < cm.npush; ++t)
    if (j >= cm.push[t].begin && j < cm.push[t].end) cm.push[t].dst[j - cm.push[t].begin] = v;
}

// Called by every thread of every CTA of a kernel that used push_elem(), after its last store:
// the CTA that draws the last ticket knows all remote stores of the grid have been fenced and
// raises this rank's flag in each target's mailbox.
__device__ __forceinline__ void push_finish(const PeerComm& cm) {
  if (cm.npush <= 0) return;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&cm.local->push_ticket, 1u);
    if (tk == gridDim.x - 1) {
      cm.local->push_ticket = 0u;
      __threadfence_system();
      const unsigned long long e = cm.local->push_epoch + 1;
      cm.local->push_epoch = e;
      for (int t = 0; t < cm.npush; ++t)
        st_release_sys(&cm.mbox[cm.push[t].peer]->halo_flag[cm.rank], e);
    }
  }
}

// One thread: wait until every source this rank gathers from has completed as many pushes as
// this rank has (the ranks push in lockstep, once before every matvec).
__device__ __forceinline__ void wait_halo(const PeerComm& cm, CgState* S) {
  if (cm.nwait <= 0) return;
  const unsigned long long e = *(volatile unsigned long long*)&cm.local->push_epoch;
  for (int i = 0; i < cm.nwait; ++i)
    if (!spin_until(&cm.mbox[cm.rank]->halo_flag[cm.wait_src[i]], e, false)) {
      comm_failed(S);
      return;
    }
}

// ---------------------------------------------------------------------------------------------
// Grid barrier of the one-kernel iteration (co-resident grid, cooperative launch): an epoch in
// device memory.  The CTA that closed a fused reduction releases; everyone else waits.
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int ld_volatile_int(const int* p) {
  return *(const volatile int*)p;
}
__device__ __forceinline__ double ld_volatile_f64(const double* p) {
  return *(const volatile double*)p;
}

// Every thread of the releasing CTA calls this after its last store of the phase.
template <int THREADS>
__device__ __forceinline__ void grid_release(unsigned long long* epoch, unsigned long long e) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    st_release_gpu(epoch, e);
  }
}

constexpr unsigned long long GRID_TIMEOUT_NS = 12000000000ull;  // > PEER_TIMEOUT_NS

// Every thread of every CTA calls this; false (block-uniform) if the grid never got there.
__device__ __forceinline__ bool grid_wait(const unsigned long long* epoch, unsigned long long e,
                                          CgState* S) {
  __shared__ int sh_ok;
  if (threadIdx.x == 0) {
    int ok = 1;
    unsigned long long t0 = 0;
    for (unsigned int it = 0; ld_acquire_gpu(epoch) < e; ++it) {
      if ((it & 1023u) == 1023u) {
        const unsigned long long now = global_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > GRID_TIMEOUT_NS) {
          ok = 0;
          comm_failed(S);
          break;
        }
      }
    }
    sh_ok = ok;
  }
  __syncthreads();
  const bool ok = sh_ok != 0;
  __syncthreads();  // sh_ok may be rewritten by the next barrier
  return ok;
}

// ---------------------------------------------------------------------------------------------
// Scalar bookkeeping of the loop (reference cg.c:113,125-134), run by exactly one thread per
// iteration: the last block of the fused kernel (after the in-kernel cross-GPU sum when several
// GPUs take part), or a one-thread kernel after an NCCL sum.
// ---------------------------------------------------------------------------------------------

// After p.s is known (reference cg.c:113).  Breakdown policy: the reference divides regardless
// and poisons x with NaN once r is exactly zero; here the loop stops before the update -- but only
// where the division has no finite result (r.r == 0, p.s == 0, p.s not finite).  A NEGATIVE p.s is
// let through exactly as the reference lets it: on a negative-definite system (a Laplacian
// assembled with the other sign) the recurrence produces the iterates of (-A, -b).
__device__ __forceinline__ void finish_ps(CgState* S, double ps) {
  S->ps = ps;
  if (S->rr == 0.0) {
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  } else if (ps == 0.0 || !isfinite(ps)) {
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
  S->sr_first = 1;
  if (bb == 0.0) {  // b = 0: x = 0 is the solution; the reference would return NaN
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  }
}

// Single-reduction variant (Chronopoulos-Gear): after w = A r, with rw = r.w and rr = r.r from the
// same fused reduction.  The first call (loop entry, r = b) only forms alpha_0 = rr / rw; later
// calls close the x update that preceded them: stop tests as in finish_rr, then
//   beta = rr / rr_old,   alpha = rr / (rw - beta * rr / alpha_old).
__device__ __forceinline__ void finish_sr(CgState* S, double rw, double rr) {
  S->ps = rw;
  if (S->sr_first) {
    S->sr_first = 0;
    if (rr == 0.0) {
      S->done = 1;
      S->status = CGB200_ST_EXACT;
    } else if (rw == 0.0 || !isfinite(rw)) {
      S->done = 1;
      S->status = CGB200_ST_BREAKDOWN;
    } else {
      S->alpha = rr / rw;
      S->beta = 0.0;
    }
    S->rr = rr;
    return;
  }
  const int k = S->iter;
  S->iter = k + 1;
  if (S->thresh > 0.0 && rr <= S->thresh) {
    S->done = 1;
    S->status = CGB200_ST_CONVERGED;
  } else if (k >= S->max_iter) {
    S->done = 1;
    S->status = CGB200_ST_MAXITER;
  } else if (rr == 0.0) {
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  } else {
    const double beta = rr / S->rr;
    const double denom = rw - beta * rr / S->alpha;
    if (denom == 0.0 || !isfinite(denom)) {
      S->done = 1;
      S->status = CGB200_ST_BREAKDOWN;
    } else {
      S->beta = beta;
      S->alpha = rr / denom;
    }
  }
  S->rr = rr;
}

// General loop entry (extensions: initial guess x0, Jacobi preconditioner M = diag(A)).  With a
// preconditioner the recurrence runs on r.z (z = M^-1 r), which takes the place of r.r in S->rr, and
// the stop test keeps looking at the true r.r, carried in S->aux[0].
//   FIN_BB     b.b known: threshold and a clean state (nothing decided yet)
//   FIN_INIT2  r0 = b - A x0 formed: r0.z0 and r0.r0
//   FIN_RZ     after the x/r update: r.z and r.r
__device__ __forceinline__ void finish_bb(CgState* S, double bb) {
  S->bb = bb;
  S->thresh = S->tol * S->tol * bb;
  S->iter = 0;
  S->alpha = 0.0;
  S->beta = 0.0;
  S->ps = 0.0;
  S->done = 0;
  S->status = CGB200_ST_RUNNING;
  S->sr_first = 1;
}

__device__ __forceinline__ void finish_init2(CgState* S, double rz, double rr) {
  S->rr = rz;
  S->aux[0] = rr;
  if (rr == 0.0) {  // x0 solves the system exactly
    S->done = 1;
    S->status = CGB200_ST_EXACT;
  } else if (S->thresh > 0.0 && rr <= S->thresh) {
    S->done = 1;
    S->status = CGB200_ST_CONVERGED;
  }
}

__device__ __forceinline__ void finish_rz(CgState* S, double rz_new, double rr_new) {
  const int k = S->iter;
  S->iter = k + 1;
  if (S->thresh > 0.0 && rr_new <= S->thresh) {
    S->done = 1;
    S->status = CGB200_ST_CONVERGED;
  } else if (k >= S->max_iter) {
    S->done = 1;
    S->status = CGB200_ST_MAXITER;
  } else {
    S->beta = rz_new / S->rr;
  }
  S->rr = rz_new;
  S->aux[0] = rr_new;
}

__device__ __forceinline__ void apply_finish(CgState* S, int fin, double t0, double t1) {
  if (fin == FIN_INIT) finish_init(S, t0);
  else if (fin == FIN_PS) finish_ps(S, t0);
  else if (fin == FIN_RR) finish_rr(S, t0);
  else if (fin == FIN_SR) finish_sr(S, t0, t1);
  else if (fin == FIN_BB) finish_bb(S, t0);
  else if (fin == FIN_INIT2) finish_init2(S, t0, t1);
  else if (fin == FIN_RZ) finish_rz(S, t0, t1);
}

// Block-wide epilogue of a fused reduction, run by the last block only (totals in thread 0).
template <int THREADS>
__device__ __forceinline__ void finish_block(CgState* S, const PeerComm& cm, int fin, double t0,
                                             double t1) {
  if (fin == FIN_LOCAL) {  // NCCL mode / kernel timing: leave the local sums for the caller
    if (threadIdx.x == 0) {
      S->local[0] = t0;
      S->local[1] = t1;
    }
    return;
  }
  bool ok = true;
  if (cm.world > 1) ok = peer_allreduce<THREADS>(cm, S, t0, t1);
  if (ok && threadIdx.x == 0) apply_finish(S, fin, t0, t1);
}

}  // namespace cgb
