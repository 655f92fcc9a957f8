// kernels.cu -- hand-written sm_100a kernels of the conjugate-gradient path.
//
// Reference functions replaced (all of mv_ops.c's arithmetic and the body of cg.c's loop):
//   k_spmv_stream     mv_mult + mat_get_row on CSR     mv_ops.c:160-201, 99-113   (fused with p.s)
//   k_gemv_rows       mv_mult on full rows (nnz==n^2)  mv_ops.c:160-201           (fused with p.s)
//   k_update_xr       sv_mult, vec_add, sv_mult, vec_sub, dot_product   cg.c:115-123,129
//   k_update_p        sv_mult, vec_add                 cg.c:131-132
//   k_init            new_mv_struct_with_size, 2x mv_deep_copy, dot     cg.c:104-108
//   k_dot/k_scale/k_add/k_sub   the standalone operators of mv_ops.h    mv_ops.c:117-259
//
// All kernels are HBM-bound streaming kernels; no tensor cores (GEMV/SpMV is not a dense
// contraction).  Grids are persistent multiples of the SM count; reductions are deterministic.
// Arithmetic in the vector updates and the CSR row sums is deliberately unfused (separate
// multiply and add, __dmul_rn/__dadd_rn) because the reference is built without FMA contraction
// (Makefile:2, x86-64 -g): s = A p is then bit-identical to the reference and only the dot
// products differ (by summation order).
#include "kernels.h"

#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "device_utils.cuh"

namespace cgb {

static int g_launches = 0;
long long launch_count() { return g_launches; }
void add_launches(long long n) { g_launches += (int)n; }

// ---------------------------------------------------------------------------------------------
// K1-sparse: CSR "stream" SpMV fused with p.s
//
// A tile is R consecutive rows (one consumer thread per row).  The tile's slice of values[],
// col_indices[] and row_ptr[] is contiguous in memory, so a dedicated producer warp moves it into
// shared memory with three 1-D bulk async copies (TMA, UBLKCP) that complete on a "full"
// mbarrier; STAGES tiles are in flight per CTA, so outstanding bytes cost no registers.  Each
// consumer thread walks its own row out of shared memory in ascending column order (the
// reference's order, mv_ops.c:106-110,192-194) and gathers p[col] through the read-only path with
// an L2 evict-last policy, while the matrix streams with evict-first so it does not push p out of
// L2.  A consumer warp releases a stage by arriving on its "empty" mbarrier; there is no
// CTA-wide barrier in the steady state, warps drift freely across tiles.  Tiles are dealt
// round-robin to a persistent grid so that all CTAs sweep the matrix as one front and the
// gathered window of p (two grid planes for a 7-point stencil) stays L2-resident: HBM sees p once.
//
// Rows longer than the stage capacity are handled by splitting the tile's nnz span in chunks of
// CAP entries; a row simply keeps accumulating across chunks (still ascending order).
// ---------------------------------------------------------------------------------------------

template <int R, int CAP, int STAGES>
struct SpmvSmem {
  static constexpr int RPN = R + 4;  // row_ptr entries copied per tile (R+1 rounded up to 16 B)
  static constexpr size_t vals_bytes = (size_t)STAGES * CAP * 8;
  static constexpr size_t cols_bytes = (size_t)STAGES * CAP * 4;
  static constexpr size_t rp_bytes = (size_t)STAGES * RPN * 4;
  static constexpr size_t bar_bytes = (size_t)STAGES * 16;
  static constexpr size_t total = vals_bytes + cols_bytes + rp_bytes + bar_bytes + 16;
  // CTAs per SM that shared memory and the thread count allow.  Told to the compiler through
  // __launch_bounds__: ptxas otherwise budgets registers for the occupancy the block size alone
  // would permit (3 CTAs of 544 threads -> 40 registers) and serialises the row's gathers to fit.
  static constexpr int by_smem = (int)((227u * 1024u) / (total + 2048u));
  static constexpr int by_threads = 2048 / (R + 32);
  static constexpr int min_ctas = by_smem < by_threads ? (by_smem < 1 ? 1 : by_smem) : by_threads;
};

// Up to three runs of tiles, visited in this order.  With several GPUs run 0 holds the interior
// tiles (columns inside the owned block) and runs 1-2 the boundary tiles that gather halo values.
__device__ __forceinline__ int vtile_to_tile(const TileRange& tr, int v, const bool rev = false) {
  const int n0 = tr.e[0] - tr.b[0];
  if (v < n0) return rev ? tr.e[0] - 1 - v : tr.b[0] + v;
  v -= n0;
  const int n1 = tr.e[1] - tr.b[1];
  if (v < n1) return rev ? tr.e[1] - 1 - v : tr.b[1] + v;
  v -= n1;
  return rev ? tr.e[2] - 1 - v : tr.b[2] + v;
}

// One row's share of a staged chunk: entries [k, k1) of the stage, ascending column order.
// COHERENT = the gathered values may have been written by a peer GPU during this kernel.
//
// The gathers are the only long-latency operations of the consumer, so the body is kept free of
// branches: the staged column indices and values are read unconditionally (reading a few entries
// past the row stays inside the CTA's shared memory) and the gather is predicated inside the
// instruction.  All (up to 8) gathers of a row are then issued back to back before the first
// product is formed; with a branch around each load ptxas emitted load, multiply, add per entry,
// i.e. one or two gathers in flight per thread.  The sum itself is still taken entry by entry in
// ascending column order with unfused multiply and add (mv_ops.c:106-110,192-194).
template <bool COHERENT>
__device__ __forceinline__ double walk_row(const double* __restrict__ v_s,
                                           const int* __restrict__ c_s, int k, const int k1,
                                           const double* __restrict__ pg, const uint64_t pol,
                                           double acc) {
  while (k < k1) {
    int c[8];
    double v[8], pv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) c[j] = c_s[k + j];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      pv[j] = COHERENT ? ld_l2_if(pg + c[j], (k + j) < k1) : ld_keep_if(pg + c[j], pol, (k + j) < k1);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = v_s[k + j];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const double t = __dadd_rn(acc, __dmul_rn(v[j], pv[j]));
      acc = (k + j) < k1 ? t : acc;
    }
    k += 8;
  }
  return acc;
}

template <int R, int CAP, int STAGES>
__global__ void __launch_bounds__(R + 32, SpmvSmem<R, CAP, STAGES>::min_ctas)
k_spmv_stream(const double* __restrict__ vals, const int* __restrict__ cols,
              const int* __restrict__ rowptr, int nrows, const double* __restrict__ pg,
              const double* __restrict__ pown, double* __restrict__ s, const TileRange tr,
              const int wait_from, const long long lead, const long long plen,
              double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
              const int mode, const int hint, const __grid_constant__ PeerComm cm) {
  using SM = SpmvSmem<R, CAP, STAGES>;
  constexpr int THREADS = R + 32;
  constexpr int CWARPS = R / 32;
  const int fin = mode & MV_FIN_MASK;
  const bool rev = (mode & MV_REV) != 0;
  if (fin != FIN_NONE && S->done) return;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm_vals = reinterpret_cast<double*>(smem_raw);
  int* sm_cols = reinterpret_cast<int*>(smem_raw + SM::vals_bytes);
  int* sm_rp = reinterpret_cast<int*>(smem_raw + SM::vals_bytes + SM::cols_bytes);
  uint64_t* full =
      reinterpret_cast<uint64_t*>(smem_raw + SM::vals_bytes + SM::cols_bytes + SM::rp_bytes);
  uint64_t* empty = full + STAGES;

  const int tid = threadIdx.x;
  const int nvt = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);

  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], CWARPS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  double dot = 0.0, dot2 = 0.0;

  if (tid >= R) {
    // ------------------------------ producer warp ------------------------------
    if (tid == R) {
      const uint64_t pol_stream = policy_evict_first();
      int item = 0;
      bool waited = false;
      // bounds of the next tile are fetched one tile ahead so their latency is never exposed
      int vt = blockIdx.x;
      int nlo = 0, nhi = 0;
      if (vt < nvt) {
        const int t = vtile_to_tile(tr, vt, rev);
        nlo = __ldg(rowptr + t * R);
        nhi = __ldg(rowptr + min(t * R + R, nrows));
      }
#pragma unroll 1
      for (; vt < nvt; vt += gridDim.x) {
        if (vt >= wait_from && !waited) {
          // first boundary tile of this CTA: the peers' rows of the gather vector must have
          // landed.  The consumers are ordered behind this through the stage's full barrier.
          wait_halo(cm, S);
          waited = true;
        }
        const int t = vtile_to_tile(tr, vt, rev);
        const int lo = nlo & ~3, hi = (nhi + 3) & ~3;
        const int vt2 = vt + gridDim.x;
        if (vt2 < nvt) {
          const int t2 = vtile_to_tile(tr, vt2, rev);
          nlo = __ldg(rowptr + t2 * R);
          nhi = __ldg(rowptr + min(t2 * R + R, nrows));
          if (lead >= 0) {
            // banded matrix: the only gathered values of the next tile that nobody has touched yet
            // are the R entries at its upper band edge; have L2 fetch them one tile ahead so the
            // consumers' first touch is an L2 hit instead of a DRAM round trip per row
            const long long i0 = ((long long)t2 * R + lead) & ~1ll;
            const long long cnt = min((long long)R + 2, plen - i0) & ~1ll;
            if (cnt > 0) bulk_prefetch_l2(pg + i0, (uint32_t)cnt * 8u);
          }
        }
        const int nchunks = max(1, (hi - lo + CAP - 1) / CAP);
#pragma unroll 1
        for (int chunk = 0; chunk < nchunks; ++chunk, ++item) {
          const int stage = item % STAGES;
          if (item >= STAGES) mbar_wait(&empty[stage], ((uint32_t)(item / STAGES) & 1u) ^ 1u);
          const int c0 = lo + chunk * CAP;
          const int cnt = min(CAP, hi - c0);
          const uint32_t bytes = (uint32_t)cnt * 12u + (chunk == 0 ? (uint32_t)SM::RPN * 4u : 0u);
          mbar_expect_tx(&full[stage], bytes);
          if (cnt > 0) {
            bulk_g2s(sm_vals + (size_t)stage * CAP, vals + c0, (uint32_t)cnt * 8u, &full[stage],
                     pol_stream);
            bulk_g2s(sm_cols + (size_t)stage * CAP, cols + c0, (uint32_t)cnt * 4u, &full[stage],
                     pol_stream);
          }
          if (chunk == 0)
            bulk_g2s(sm_rp + (size_t)stage * SM::RPN, rowptr + (size_t)t * R,
                     (uint32_t)SM::RPN * 4u, &full[stage], pol_stream);
        }
      }
    }
  } else {
    // ------------------------------ consumer warps ------------------------------
    const uint64_t pol_keep = hint ? policy_evict_last() : policy_evict_normal();
    const int lane = tid & 31;
    int item = 0;
#pragma unroll 1
    for (int vt = blockIdx.x; vt < nvt; vt += gridDim.x) {
      const int t = vtile_to_tile(tr, vt, rev);
      const int row = t * R + tid;
      const int rows_valid = min(R, nrows - t * R);
      const bool boundary = vt >= wait_from;
      int rs = 0, re = 0, lo = 0, nchunks = 1;
      double acc = 0.0;
#pragma unroll 1
      for (int chunk = 0; chunk < nchunks; ++chunk, ++item) {
        const int stage = item % STAGES;
        mbar_wait(&full[stage], (uint32_t)(item / STAGES) & 1u);
        const double* v_s = sm_vals + (size_t)stage * CAP;
        const int* c_s = sm_cols + (size_t)stage * CAP;
        if (chunk == 0) {
          const int* rp_s = sm_rp + (size_t)stage * SM::RPN;
          lo = rp_s[0] & ~3;
          const int hi = (rp_s[rows_valid] + 3) & ~3;
          nchunks = max(1, (hi - lo + CAP - 1) / CAP);
          if (tid < rows_valid) {
            rs = rp_s[tid];
            re = rp_s[tid + 1];
          }
        }
        const int c0 = lo + chunk * CAP;
        const int k = max(rs - c0, 0);
        const int k1 = min(re - c0, CAP);
        if (boundary)
          acc = walk_row<true>(v_s, c_s, k, k1, pg, pol_keep, acc);
        else
          acc = walk_row<false>(v_s, c_s, k, k1, pg, pol_keep, acc);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[stage]);  // this warp is done with the stage
      }
      if (tid < rows_valid) {
        s[row] = acc;
        if (fin != FIN_NONE) {
          const double g = pown[row];
          dot += g * acc;
          if (mode & MV_DUAL) dot2 += g * g;
        }
      }
    }
  }

  if (fin == FIN_NONE) return;
  const int nv = (mode & MV_DUAL) ? 2 : 1;
  if (mode & MV_PARTIAL) {
    // first half of a split launch (interior rows): leave the partials for the finishing launch
    // (the caller passes partials + MV_EXTRA_OFFSET)
    __shared__ double smp[THREADS / 32];
    const double t = block_sum<THREADS>(dot, smp);
    if (tid == 0) partials[blockIdx.x] = t;
    if (nv == 2) {
      const double t2 = block_sum<THREADS>(dot2, smp);
      if (tid == 0) partials[PART_STRIDE + blockIdx.x] = t2;
    }
    return;
  }
  // `partials` may carry extra slots written by a preceding MV_PARTIAL launch
  double v[2] = {dot, dot2};
  bool last;
  if (nv == 2) {
    last = grid_reduce<THREADS, 2>(v, partials, ticket, partials + MV_EXTRA_OFFSET, mode >> 8);
  } else {
    double v1[1] = {dot};
    last = grid_reduce<THREADS, 1>(v1, partials, ticket, partials + MV_EXTRA_OFFSET, mode >> 8);
    v[0] = v1[0];
    v[1] = 0.0;
  }
  if (last) finish_block<THREADS>(S, cm, fin, v[0], v[1]);
}

// ---------------------------------------------------------------------------------------------
// K1-sparse for long rows: a group of G lanes per row (CSR blocks whose mean row length is tens
// to hundreds of entries -- banded and finite-element matrices, the course matrix of the
// reference with ~345 entries per row, cg.c:231-232,260).
//
// The stream kernel above gives a row to one thread, which is right for stencil rows (all of a
// row's gathers in flight at once) and wrong for long ones: a 512-row tile of 345-entry rows is
// 50 pipeline stages, each thread takes ~7 entries per stage, and the launch is bound by the
// stage hand-over latency (0.35 ms per product on n = 52269, ten times the HBM time).  Here the G
// lanes of a group load G consecutive entries of their row at a time (values, column indices and
// the gathered operands of UNR such steps are in flight together), each forms its product, and
// the products are then added in ascending column order by a chain of G shuffles and adds that
// every lane of the group executes alike: s[i] has the bits of the reference's left-to-right sum
// (mv_ops.c:106-110,192-194), as in the stream kernel.  A warp serves 32/G rows, so the chain's
// instructions are shared by that many rows.  CTAs own contiguous row ranges; one value per row
// is stored by the group's first lane, which also carries the fused p.s (and p.p) partial sums.
// Single GPU or NCCL mode only (no halo wait in this form).
// ---------------------------------------------------------------------------------------------

template <int THREADS, int G, int UNR>
__global__ void __launch_bounds__(THREADS)
k_spmv_rowgroup(const double* __restrict__ vals, const int* __restrict__ cols,
                const int* __restrict__ rowptr, const int nrows, const double* __restrict__ pg,
                const double* __restrict__ pown, double* __restrict__ s,
                double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
                const int mode, const __grid_constant__ PeerComm cm) {
  constexpr int GROUPS = THREADS / G;  // rows in flight per CTA
  const int fin = mode & MV_FIN_MASK;
  const bool rev = (mode & MV_REV) != 0;
  if (fin != FIN_NONE && S->done) return;
  const uint64_t pol_keep = policy_evict_last();
  const int gl = threadIdx.x & (G - 1);  // lane within the group
  const int grp = threadIdx.x / G;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1)));
  const int cta = rev ? (int)gridDim.x - 1 - (int)blockIdx.x : (int)blockIdx.x;
  const int r0 = (int)((long long)nrows * cta / gridDim.x);
  const int r1 = (int)((long long)nrows * (cta + 1) / gridDim.x);
  double dot = 0.0, dot2 = 0.0;
#pragma unroll 1
  for (int row = r0 + grp; row < r1; row += GROUPS) {
    const int rs = __ldg(rowptr + row), re = __ldg(rowptr + row + 1);
    double acc = 0.0;
#pragma unroll 1
    for (int k0 = rs; k0 < re; k0 += G * UNR) {
      double t[UNR];
      int c[UNR];
      double v[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int k = k0 + u * G + gl;
        const bool ok = k < re;
        c[u] = ok ? __ldcs(cols + k) : 0;
        v[u] = ok ? __ldcs(vals + k) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) t[u] = ld_keep_if(pg + c[u], pol_keep, k0 + u * G + gl < re);
#pragma unroll
      for (int u = 0; u < UNR; ++u) t[u] = __dmul_rn(v[u], t[u]);
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int kb = k0 + u * G;
#pragma unroll
        for (int j = 0; j < G; ++j) {
          const double tj = __shfl_sync(gmask, t[u], j, G);
          const double a2 = __dadd_rn(acc, tj);
          acc = (kb + j < re) ? a2 : acc;
        }
      }
    }
    if (gl == 0) {
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

// ---------------------------------------------------------------------------------------------
// K1-sparse for long rows, second form: a warp per block of 32 rows, transposed through shared
// memory.  The shuffle chain of k_spmv_rowgroup is bound by the shuffle unit for wide groups and
// by the number of memory lines a load instruction touches for narrow ones.  Here the warp loads
// 32 consecutive entries of ONE row per instruction (256 B of values, 128 B of indices: fully
// coalesced), for each of its 32 rows in turn, and leaves the PRODUCTS in a 32 x 33 tile of
// shared memory; then lane i adds row i's 32 products in ascending column order straight out of
// the tile (conflict-free 64-bit reads), all 32 chains side by side.  Same bits as the
// reference's left-to-right row sum; about a quarter of an instruction per entry.  The indices
// and values of the NEXT 32-column slab are fetched into registers before the chain of the
// current one runs, so the DRAM latency of a slab hides behind the additions of the previous one.
// ---------------------------------------------------------------------------------------------

// Round 2 (ncu --set full on the shipped <256,16> form, profiles/r02_ncu_longrow.md): DRAM traffic equals
// the algorithmic bytes, DRAM at 54 % of its peak, 16 warps per SM (119 registers), 1.1 warp
// instructions per stored entry; top stalls long scoreboard (the gathers behind the slab's loads),
// barrier (the launch tail: 3267 blocks over 2368 warps) and wait.  The kernel streams at 0.84 of
// the HBM peak on a matrix four times the size (869 MB) and 0.68 on the course-sized one (217 MB,
// 49 us per product): what is missing there is a fixed ~20 us of ramp and tail, not bandwidth.
// Tried against it and dropped, each measured slower on the course-sized matrix: equal runs of
// stored entries per warp with 8-row units (0.53), an L2-resident head of the matrix (no gain: the
// launch is not DRAM-bound), a ring of 16-byte asynchronous copies with gather-ahead (0.45), a
// thread per row fed by per-row bulk copies (0.31), blocks drawn from an atomic counter with a
// predicate-free path for full slabs (0.64), paired 128-bit loads covering two rows per
// instruction (0.61), three CTAs per SM at 80 registers (0.34, spills), the next slab's indices
// requested right behind the gathers instead of after the products (0.65).
template <int THREADS, int ROWS, bool COHERENT>
__global__ void __launch_bounds__(THREADS)
k_spmv_rowblock(const double* __restrict__ vals, const int* __restrict__ cols,
                const int* __restrict__ rowptr, const int nrows, const double* __restrict__ pg,
                const double* __restrict__ pown, double* __restrict__ s,
                double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
                const int mode, const __grid_constant__ PeerComm cm) {
  // COHERENT (several GPUs, peer-memory mode): part of the gather vector is stored by the peers'
  // direction updates while this kernel may already be running, so every CTA first waits for
  // their completion flags and the gathers bypass L1.  (The stream kernel overlaps that wait with
  // its interior tiles; here the whole launch sits behind it.)
  // ROWS rows per warp and work unit (lane i < ROWS owns row i of the unit): fewer rows mean
  // fewer registers for the prefetched slab, hence more warps per SM, and finer units -- both
  // matter while a unit's time is set by latencies rather than by its bytes
  constexpr int WARPS = THREADS / 32;
  constexpr int LD = 33;  // tile row stride in doubles: lane i reads bank pair 2i + 2j (mod 32)
  __shared__ double tile[WARPS][ROWS * LD];
  const int fin = mode & MV_FIN_MASK;
  const bool rev = (mode & MV_REV) != 0;
  if (fin != FIN_NONE && S->done) return;
  const uint64_t pol_keep = policy_evict_last();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* const tw = tile[warp];
  if (COHERENT) {
    if (threadIdx.x == 0) wait_halo(cm, S);
    __syncthreads();
  }
  const int nblk = (nrows + ROWS - 1) / ROWS;
  const int cta = rev ? (int)gridDim.x - 1 - (int)blockIdx.x : (int)blockIdx.x;
  const int b0 = (int)((long long)nblk * cta / gridDim.x);
  const int b1 = (int)((long long)nblk * (cta + 1) / gridDim.x);
  double dot = 0.0, dot2 = 0.0;
#pragma unroll 1
  for (int blk = b0 + warp; blk < b1; blk += WARPS) {
    const int row = blk * ROWS + lane;
    const bool rv = lane < ROWS && row < nrows;
    const int rs = rv ? __ldg(rowptr + row) : 0;
    const int re = rv ? __ldg(rowptr + row + 1) : 0;
    const int maxlen = __reduce_max_sync(0xffffffffu, re - rs);
    double acc = 0.0;
    int c[ROWS];
    double v[ROWS];
    unsigned ok = 0u;  // bit i: this lane holds an entry of row i in the slab in registers
    // slab 0 into registers
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
      const int k = __shfl_sync(0xffffffffu, rs, i) + lane;
      const bool o = k < __shfl_sync(0xffffffffu, re, i);
      ok |= (o ? 1u : 0u) << i;
      c[i] = o ? __ldcs(cols + k) : 0;
      v[i] = o ? __ldcs(vals + k) : 0.0;
    }
#pragma unroll 1
    for (int off = 0; off < maxlen; off += 32) {
      // products of the slab in registers -> tile
      double t[ROWS];
#pragma unroll
      for (int i = 0; i < ROWS; ++i)
        t[i] = COHERENT ? ld_l2_if(pg + c[i], (ok >> i) & 1u) : ld_keep_if(pg + c[i], pol_keep, (ok >> i) & 1u);
#pragma unroll
      for (int i = 0; i < ROWS; ++i) tw[i * LD + lane] = __dmul_rn(v[i], t[i]);
      __syncwarp();
      // next slab's indices and values on their way while this slab is added up
      ok = 0u;
      if (off + 32 < maxlen) {
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
          const int k = __shfl_sync(0xffffffffu, rs, i) + off + 32 + lane;
          const bool o = k < __shfl_sync(0xffffffffu, re, i);
          ok |= (o ? 1u : 0u) << i;
          c[i] = o ? __ldcs(cols + k) : 0;
          v[i] = o ? __ldcs(vals + k) : 0.0;
        }
      }
      if (lane < ROWS) {
        const int left = re - rs - off;  // entries of this lane's row that are in the tile
        const double* mine = tw + lane * LD;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const double a2 = __dadd_rn(acc, mine[j]);
          acc = j < left ? a2 : acc;
        }
      }
      __syncwarp();
    }
    if (rv) {
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

// ---------------------------------------------------------------------------------------------
// K1-dense: row-block GEMV over values[] only, fused with p.s
//
// When every CSR row is full (nnz == n*n, row_ptr[i] == i*n) values[] is a row-major dense
// matrix and col_indices carries no information: stream 8 B/entry instead of 12.  A CTA owns
// RB rows at a time; its threads stride across the columns with 256-bit loads, so each p chunk
// (L1/L2 resident) is reused for RB rows of A and RB independent wide loads are in flight per
// thread.  Row groups are dealt round-robin to a persistent grid.
// ---------------------------------------------------------------------------------------------

template <int THREADS, int RB, int VEC>
__global__ void __launch_bounds__(THREADS)
k_gemv_rows(const double* __restrict__ A, long long ld, int nrows, int ncols,
            const double* __restrict__ pg, const double* __restrict__ pown,
            double* __restrict__ s, double* __restrict__ partials,
            unsigned int* __restrict__ ticket, CgState* S, const int mode, const int wait,
            const __grid_constant__ PeerComm cm) {
  const int fin = mode & MV_FIN_MASK;
  if (fin != FIN_NONE && S->done) return;
  __shared__ double sm_red[RB][THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (wait) {  // the peers' slices of the gather vector must have landed (all-gather mode)
    if (tid == 0) wait_halo(cm, S);
    __syncthreads();
  }
  const uint64_t pol_stream = policy_evict_first();
  const int ngroups = (nrows + RB - 1) / RB;
  double dot = 0.0, dot2 = 0.0;

#pragma unroll 1
  for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
    const int row0 = g * RB;
    const double* rowp[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) rowp[r] = A + (long long)min(row0 + r, nrows - 1) * ld;
    double acc[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = 0.0;

    if constexpr (VEC == 4) {
#pragma unroll 2
      for (int c = tid * 4; c < ncols; c += THREADS * 4) {
        const double4_t pv = *reinterpret_cast<const double4_t*>(pg + c);
        double4_t a[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) a[r] = ld_stream4(rowp[r] + c);
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          acc[r] = fma(a[r].x, pv.x, acc[r]);
          acc[r] = fma(a[r].y, pv.y, acc[r]);
          acc[r] = fma(a[r].z, pv.z, acc[r]);
          acc[r] = fma(a[r].w, pv.w, acc[r]);
        }
      }
    } else if constexpr (VEC == 2) {
#pragma unroll 2
      for (int c = tid * 2; c < ncols; c += THREADS * 2) {
        const double2 pv = *reinterpret_cast<const double2*>(pg + c);
        double2 a[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) a[r] = ld_stream2(rowp[r] + c, pol_stream);
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          acc[r] = fma(a[r].x, pv.x, acc[r]);
          acc[r] = fma(a[r].y, pv.y, acc[r]);
        }
      }
    } else {
#pragma unroll 4
      for (int c = tid; c < ncols; c += THREADS) {
        const double pv = pg[c];
#pragma unroll
        for (int r = 0; r < RB; ++r) acc[r] = fma(__ldg(rowp[r] + c), pv, acc[r]);
      }
    }

#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = warp_sum(acc[r]);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) sm_red[r][warp] = acc[r];
    }
    __syncthreads();
    if (tid < RB && row0 + tid < nrows) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < THREADS / 32; ++w) t += sm_red[tid][w];
      s[row0 + tid] = t;
      if (fin != FIN_NONE) {
        const double g = pown[row0 + tid];
        dot += g * t;
        if (mode & MV_DUAL) dot2 += g * g;
      }
    }
  }

  if (fin == FIN_NONE) return;
  double v[2] = {dot, dot2};
  bool last;
  if (mode & MV_DUAL) {
    last = grid_reduce<THREADS, 2>(v, partials, ticket);
  } else {
    double v1[1] = {dot};
    last = grid_reduce<THREADS, 1>(v1, partials, ticket);
    v[0] = v1[0];
    v[1] = 0.0;
  }
  if (last) finish_block<THREADS>(S, cm, fin, v[0], v[1]);
}

// ---------------------------------------------------------------------------------------------
// K1-dense, balanced form (the default whenever 256-bit loads are possible).
//
// k_gemv_rows deals whole row groups to CTAs, so the slowest CTA owns ceil(groups / grid) of them:
// 3 % idle tail on one GPU (8192 groups over 444 CTAs), 30 % on an eighth of the matrix.  Here the
// work is cut into units of RB rows x CW columns and CTA c owns the contiguous unit range
// [U c / grid, U (c+1) / grid): every CTA streams the same number of bytes (+- one unit).  Row
// groups that straddle two CTAs are finished by the last CTA of the grid: each CTA leaves at most
// two records {group, RB partial row sums} (its first and its last group when incomplete), and the
// CTA that draws the last ticket adds the records of a group in CTA order -- fixed order, so s and
// p.s stay run-to-run reproducible -- before it closes the fused reduction.
// ---------------------------------------------------------------------------------------------

constexpr int GEMV_REC_OFFSET = 2 * PART_STRIDE;  // records live behind the reduction partials
constexpr int GEMV_CNT_OFFSET = 3 * PART_STRIDE;  // per-group arrival counters (unsigned, zero at rest)
constexpr int GEMV_GDOT_OFFSET = 1024;            // dot-product shares of the shared groups: the
                                                  // `extra` partials of grid_reduce, slot = owner CTA
constexpr int GEMV_MAX_GRID = 819;                // PART_STRIDE / (2 * (RB + 1)) records, < 1024 slots

// Unit u belongs to the CTA with the largest c such that U c / G <= u.
__device__ __forceinline__ int gemv_cta_of(long long u, long long U, int G) {
  return (int)(((u + 1) * G - 1) / U);
}

// Cold tail of gemv_units_stream, kept out of line so that its index arithmetic (64-bit divisions)
// costs the streaming loop no registers.  Called by every thread of the CTA after the records
// have been written; gid0/gid1 are the group numbers of this CTA's records (nrec of them).
template <int THREADS, int RB>
__device__ __noinline__ void gemv_units_shared_groups(int nrows, int ncols, const int ngroups,
                                                      const double* __restrict__ pown,
                                                      double* __restrict__ s,
                                                      double* __restrict__ partials,
                                                      const bool dots, const int nrec,
                                                      const double gid0, const double gid1) {
  constexpr int CW = THREADS * 4;
  constexpr int REC = RB + 1;
  __shared__ int sm_fin[2][3];  // per record: finishing?, first CTA, last CTA of the group
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = (int)gridDim.x;
  const int cpg = (ncols + CW - 1) / CW;
  const long long U = (long long)ngroups * cpg;
  const long long ubeg = U * blockIdx.x / G;
  const long long u1 = U * (blockIdx.x + 1) / G;
  const double* recs = partials + GEMV_REC_OFFSET;
  unsigned int* cnt = reinterpret_cast<unsigned int*>(partials + GEMV_CNT_OFFSET);
  double* gd = partials + GEMV_GDOT_OFFSET;
  if (tid == 0) {
    // this CTA's own slot stays empty unless the group cut by its upper boundary starts in it
    if ((int)blockIdx.x < G - 1) {
      const long long gcut = u1 / cpg;
      if (u1 % cpg == 0 || ubeg > gcut * cpg) {
        gd[blockIdx.x] = 0.0;
        gd[PART_STRIDE + blockIdx.x] = 0.0;
      }
    }
    __threadfence();
    for (int i = 0; i < 2; ++i) {
      int fin = 0, kf = 0, kl = 0;
      if (i < nrec) {
        const long long g0 = (long long)(i == 0 ? gid0 : gid1) * cpg;
        kf = gemv_cta_of(g0, U, G);
        kl = gemv_cta_of(g0 + cpg - 1, U, G);
        const unsigned int old = atomicAdd(&cnt[kf], 1u);
        if (old == (unsigned int)(kl - kf)) {  // every contributor has arrived
          cnt[kf] = 0u;
          fin = 1;
        }
      }
      sm_fin[i][0] = fin;
      sm_fin[i][1] = kf;
      sm_fin[i][2] = kl;
    }
    __threadfence();
  }
  __syncthreads();
#pragma unroll 1
  for (int i = 0; i < 2; ++i) {
    if (!sm_fin[i][0]) continue;  // block-uniform
    if (warp == 0) {
      const int kf = sm_fin[i][1], kl = sm_fin[i][2];
      const long long g = (long long)(i == 0 ? gid0 : gid1);
      const int row = (int)g * RB + lane;
      double t = 0.0;
      if (lane < RB) {
#pragma unroll 4
        for (int k = kf; k <= kl; ++k) {
          // CTA k's record of group g: its second record iff g is not its first group and its
          // first group was incomplete too (it starts inside a group)
          const long long bk = U * k / G;
          const int slot = (bk / cpg != g && bk % cpg != 0) ? 1 : 0;
          t += __ldcg(recs + ((size_t)k * 2 + slot) * REC + 1 + lane);
        }
        if (row < nrows) s[row] = t;
      }
      if (dots) {
        double d1 = 0.0, d2 = 0.0;
        if (lane < RB && row < nrows) {
          const double gv = __ldcg(pown + row);
          d1 = gv * t;
          d2 = gv * gv;
        }
        double a1 = 0.0, a2 = 0.0;
#pragma unroll
        for (int r = 0; r < RB; ++r) {  // row order
          a1 += __shfl_sync(0xffffffffu, d1, r);
          a2 += __shfl_sync(0xffffffffu, d2, r);
        }
        if (lane == 0) {
          gd[kf] = a1;
          gd[PART_STRIDE + kf] = a2;
        }
      }
    }
  }
  // (the stores above are ordered before this CTA's ticket by the fence inside grid_reduce, which
  // follows a block-wide barrier)
}


// The unit loop of one CTA: streams its unit range, writes the rows it owns whole, leaves its (at
// most two) incomplete row groups as records behind the reduction partials, and returns its share
// of the fused dot products.  A shared group is finished by whichever of its CTAs arrives last
// (one arrival counter per group): it adds the group's records in CTA order -- fixed order, s stays
// run-to-run reproducible -- writes the rows and leaves the group's share of the dot products in
// the slot of the group's first CTA, which the closing reduction folds in in slot order.  So the
// grid's last CTA has nothing left to finish and the fix-ups hide behind the stream of the others.
template <int THREADS, int RB>
__device__ __forceinline__ void gemv_units_stream(const double* __restrict__ A, long long ld,
                                                  int nrows, int ncols, const int ngroups,
                                                  const double* __restrict__ pg,
                                                  const double* __restrict__ pown,
                                                  double* __restrict__ s,
                                                  double* __restrict__ partials, const bool dots,
                                                  double& dot, double& dot2) {
  constexpr int CW = THREADS * 4;  // columns per unit: one 256-bit load per thread
  constexpr int REC = RB + 1;
  static_assert(RB <= 32, "row group wider than a warp");
  __shared__ double sm_red[RB][THREADS / 32];
  __shared__ double sm_rec[2][REC];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 2) sm_rec[tid][0] = -1.0;
  __syncthreads();

  const int G = (int)gridDim.x;
  const int cpg = (ncols + CW - 1) / CW;          // units per row group
  const long long U = (long long)ngroups * cpg;
  const long long ubeg = U * blockIdx.x / G;
  const long long u1 = U * (blockIdx.x + 1) / G;
  long long u = ubeg;
  int nrec = 0;

#pragma unroll 1
  while (u < u1) {
    const int g = (int)(u / cpg), ka = (int)(u % cpg);
    const int kb = (int)min((long long)cpg, ka + (u1 - u));
    const int row0 = g * RB;
    const double* rowp[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) rowp[r] = A + (long long)min(row0 + r, nrows - 1) * ld;
    double acc[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = 0.0;
    const int cend = min(kb * CW, ncols);
#pragma unroll 2
    for (int c = ka * CW + tid * 4; c < cend; c += CW) {
      const double4_t pv = ld_l2_4(pg + c);
      double4_t a[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) a[r] = ld_stream4(rowp[r] + c);
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        acc[r] = fma(a[r].x, pv.x, acc[r]);
        acc[r] = fma(a[r].y, pv.y, acc[r]);
        acc[r] = fma(a[r].z, pv.z, acc[r]);
        acc[r] = fma(a[r].w, pv.w, acc[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = warp_sum(acc[r]);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) sm_red[r][warp] = acc[r];
    }
    __syncthreads();
    const bool whole = (ka == 0 && kb == cpg);
    if (tid < RB) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < THREADS / 32; ++w) t += sm_red[tid][w];
      if (whole) {
        if (row0 + tid < nrows) {
          s[row0 + tid] = t;
          if (dots) {
            const double gv = __ldcg(pown + row0 + tid);
            dot += gv * t;
            dot2 += gv * gv;
          }
        }
      } else {
        sm_rec[nrec][1 + tid] = t;
        if (tid == 0) sm_rec[nrec][0] = (double)g;
      }
    }
    if (!whole) ++nrec;  // at most the first and the last group of a CTA are incomplete
    u += kb - ka;
  }
  __syncthreads();
  double* recs = partials + GEMV_REC_OFFSET;
  if (tid < 2 * REC) recs[(size_t)blockIdx.x * 2 * REC + tid] = (&sm_rec[0][0])[tid];

  __syncthreads();  // the records are written (block-wide) before thread 0 fences and arrives
  gemv_units_shared_groups<THREADS, RB>(nrows, ncols, ngroups, pown, s, partials, dots, nrec,
                                        sm_rec[0][0], sm_rec[1][0]);
}

template <int THREADS, int RB>
__global__ void __launch_bounds__(THREADS, 4)
k_gemv_units(const double* __restrict__ A, long long ld, int nrows, int ncols,
             const double* __restrict__ pg, const double* __restrict__ pown,
             double* __restrict__ s, double* __restrict__ partials,
             unsigned int* __restrict__ ticket, CgState* S, const int mode, const int wait,
             const __grid_constant__ PeerComm cm) {
  const int fin = mode & MV_FIN_MASK;
  if (fin != FIN_NONE && S->done) return;
  if (wait) {  // the peers' slices of the gather vector must have landed (all-gather mode)
    if (threadIdx.x == 0) wait_halo(cm, S);
  }
  double dot = 0.0, dot2 = 0.0;
  gemv_units_stream<THREADS, RB>(A, ld, nrows, ncols, (nrows + RB - 1) / RB, pg, pown, s, partials,
                                 fin != FIN_NONE, dot, dot2);
  if (fin == FIN_NONE) return;
  double v[2] = {dot, dot2};
  if (!grid_reduce<THREADS, 2>(v, partials, ticket, partials + GEMV_GDOT_OFFSET,
                               (int)gridDim.x - 1))
    return;
  finish_block<THREADS>(S, cm, fin, v[0], (mode & MV_DUAL) ? v[1] : 0.0);
}

// ---------------------------------------------------------------------------------------------
// Dynamic tail of the dense matvec (one-kernel iteration only).
//
// A static equal split leaves the fastest SMs idle at the end of the launch: per-CTA stream times
// differ by +-1 % SYSTEMATICALLY (the same SMs are slow in every launch and every process,
// profiles/r01c_dense_cta_times_n32768.json), 46 us between the first and the last CTA on the
// 8.6 GB matrix.  So the last few per cent of the row groups are not dealt statically: they form
// a pool of chunks (one row group x GEMV_PCH units of columns) that CTAs draw from an atomic
// counter once their static share is done.  Who computes a chunk varies from run to run; what is
// computed does not: every chunk leaves its RB partial row sums and its share of p.s in slots
// that belong to the chunk, and both are later added in chunk order -- results stay run-to-run
// bit-reproducible.  The pool rows' s is formed in phase B from the chunk records.
//
// pool buffer (doubles): [cpc x P x RB chunk records][P * cpc chunk dot shares][counter]
// ---------------------------------------------------------------------------------------------

constexpr int GEMV_PCH = 2;  // units of columns per pool chunk (64 KB of matrix)

__host__ __device__ __forceinline__ int gemv_chunks_per_group(int ncols, int threads) {
  const int cpg = (ncols + threads * 4 - 1) / (threads * 4);
  return (cpg + GEMV_PCH - 1) / GEMV_PCH;
}

template <int THREADS, int RB>
__device__ __forceinline__ void gemv_pool_phase(const double* __restrict__ A, long long ld,
                                                int nrows, int ncols, const int ngs, const int P,
                                                const double* __restrict__ pg,
                                                const double* __restrict__ pown,
                                                double* __restrict__ pool) {
  constexpr int CW = THREADS * 4;
  __shared__ double sm_pred[RB][THREADS / 32];
  __shared__ unsigned int sm_next;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cpc = gemv_chunks_per_group(ncols, THREADS);
  const unsigned int nchunks = (unsigned int)P * cpc;
  double* rec = pool;
  double* cdot = pool + (size_t)cpc * P * RB;
  unsigned int* ctr = reinterpret_cast<unsigned int*>(cdot + nchunks);
  if (tid == 0) sm_next = atomicAdd(ctr, 1u);
  __syncthreads();
  unsigned int q = sm_next;
#pragma unroll 1
  while (q < nchunks) {
    __syncthreads();  // everyone holds q: sm_next and sm_pred may be rewritten
    unsigned int nxt = 0;
    if (tid == 0) nxt = atomicAdd(ctr, 1u);  // the next draw travels while this chunk streams
    const int j = (int)(q / cpc), k = (int)(q % cpc);
    const int row0 = (ngs + j) * RB;
    const double* rowp[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) rowp[r] = A + (long long)min(row0 + r, nrows - 1) * ld;
    double gv = 0.0;
    if (tid < RB && row0 + tid < nrows) gv = __ldcg(pown + row0 + tid);
    double acc[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = 0.0;
    const int cend = min((k + 1) * GEMV_PCH * CW, ncols);
#pragma unroll 2
    for (int c = k * GEMV_PCH * CW + tid * 4; c < cend; c += CW) {
      const double4_t pv = ld_l2_4(pg + c);
      double4_t a[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) a[r] = ld_stream4(rowp[r] + c);
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        acc[r] = fma(a[r].x, pv.x, acc[r]);
        acc[r] = fma(a[r].y, pv.y, acc[r]);
        acc[r] = fma(a[r].z, pv.z, acc[r]);
        acc[r] = fma(a[r].w, pv.w, acc[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = warp_sum(acc[r]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) sm_pred[r][warp] = acc[r];
    }
    if (tid == 0) sm_next = nxt;
    __syncthreads();
    if (warp == 0) {
      double t = 0.0;
      if (lane < RB) {
#pragma unroll
        for (int w = 0; w < THREADS / 32; ++w) t += sm_pred[lane][w];
        rec[((size_t)k * P + j) * RB + lane] = t;
      }
      const double d = (lane < RB && row0 + lane < nrows) ? gv * t : 0.0;
      double a1 = 0.0;
#pragma unroll
      for (int r = 0; r < RB; ++r) a1 += __shfl_sync(0xffffffffu, d, r);  // row order
      if (lane == 0) cdot[q] = a1;
    }
    q = sm_next;
  }
}

// ---------------------------------------------------------------------------------------------
// Dense path: the whole SOLVE in one launch (classic recurrence, cg.c:104-135).
//
// On the dense path the vectors are n doubles against n^2 of matrix: K2 and K3 are pure latency
// (5-9 us each plus two launch boundaries), which is 2 % of an iteration on one GPU and 25 % on
// an eighth of the matrix.  The persistent GEMV grid is co-resident (cooperative launch), so the
// loop continues behind grid barriers instead of launches -- loop entry and every iteration:
//   phase I  (init != 0) x = 0, r = p = b, b.b; p pushed to the peers   -> last CTA: [sum] state
//   per iteration, up to `iters` of them:
//   phase A  s = A p, p.s (unit-range GEMV as above)         -> last CTA: [cross-GPU sum] alpha
//   barrier  (the CTA that drew the last ticket releases the grid)
//            while the reductions and vector phases run HBM is idle: every CTA asks L2 to fetch
//            the head of its matrix range for the NEXT iteration (the matrix does not depend on p)
//   phase B  x += alpha p ; r -= alpha s ; r.r  (each CTA a slice) -> last CTA: [sum] beta, stop
//   barrier
//   phase C  p = r + beta p, pushed into the peers' gather vectors
//   barrier  (arrival count: the last CTA raises the peers' flags and releases the grid)
// The barriers are epochs in CgState (grid_epoch): a CTA reads the epoch when it starts -- nobody
// can have released yet, a release needs every CTA's arrival -- and waits for epoch + 1, + 2, ...
// A stop (fixed count reached, tolerance met, breakdown) is seen by every CTA behind the barrier
// that follows the decision, and the launch ends: a tolerance solve needs no host in the loop.
// `timing` != 0 leaves the scalars alone (alpha = beta = 0 state of system_time_kernel).
// (Tried and dropped: the two vector phases run by the one CTA that closes the matvec's reduction,
// one grid barrier per iteration instead of three.  Slower at every block size -- 54.1 against 52.7 us
// per iteration at 4096 rows, 358 against 327 us at 16384: one CTA against memory latency loses to
// 592 CTAs behind two barriers.)
// ---------------------------------------------------------------------------------------------

struct DenseSolveArgs {
  int iters;        // iterations this launch may run (it stops earlier when the loop is done)
  int init;         // run the loop entry first (a new solve); 0 = continue the solve in CgState
  int max_iter;     // init only: reference semantics, stop after the update at k == max_iter
  double tol;       // init only
  int wait;         // the peers' slices of p take part (all-gather by push)
  int timing;
  int pool_groups;
  int prefetch_bytes;  // matrix bytes per CTA asked into L2 ahead of the next iteration (0 = none)
};

// Every thread of every CTA calls this after its last store of a phase: the CTA that arrives last
// (optionally after raising the peers' halo flags: the pushes of this rank are complete) releases
// epoch e.  Replaces a launch boundary between p's update and the next matvec.
template <int THREADS>
__device__ __forceinline__ void grid_arrive(CgState* S, const PeerComm& cm, const bool pushed,
                                            const unsigned long long e, const int nparts) {
  if (pushed) __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int tk = atomicAdd(&S->ticket[3], 1u);
    if (tk == (unsigned int)nparts - 1) {  // only the CTAs that own rows arrive
      S->ticket[3] = 0u;
      __threadfence();
      if (pushed) {
        __threadfence_system();
        const unsigned long long pe = cm.local->push_epoch + 1;
        cm.local->push_epoch = pe;
        for (int t = 0; t < cm.npush; ++t)
          st_release_sys(&cm.mbox[cm.push[t].peer]->halo_flag[cm.rank], pe);
      }
      st_release_gpu(&S->grid_epoch, e);
    }
  }
}

// `trace` (normally null): cycle counts of the phases, accumulated by thread 0 of the CTA that
// closes each phase, relative to that CTA's own start of the iteration (tools/dense_trace.py).
template <int THREADS, int RB>
__global__ void __launch_bounds__(THREADS, 4)
k_cg_dense_solve(const double* __restrict__ A, long long ld, int nrows, int ncols,
                 const double* pg, double* pown, double* __restrict__ x, double* __restrict__ r,
                 double* s, const double* __restrict__ b, double* __restrict__ partials, CgState* S,
                 unsigned long long* __restrict__ trace, double* __restrict__ pool,
                 const DenseSolveArgs a, const __grid_constant__ PeerComm cm) {
  const int tid = threadIdx.x;
  if (!a.init && S->done) return;
  __shared__ unsigned long long sh_e0, sh_t0;
  long long c0 = 0;
  if (tid == 0) sh_e0 = ld_acquire_gpu(&S->grid_epoch);
  __syncthreads();
  unsigned long long e = sh_e0;
  const long long gstride = (long long)gridDim.x * THREADS;
  const bool pushing = cm.npush > 0;
  const int pool_groups = a.pool_groups;
  const int ngs = (nrows + RB - 1) / RB - pool_groups;  // statically dealt row groups
  // The vector phases hand row i to thread i of the grid: only the first `nact` CTAs own rows (16 of
  // 592 on an eighth of the n = 32768 system).  Only they take part in those phases' reductions and
  // arrival barrier -- the others would each add a zero partial, a fence and a ticket to the
  // critical path -- and everybody meets at the epoch wait.
  const int nact = (int)min((long long)gridDim.x, ((long long)nrows + THREADS - 1) / THREADS);
  const bool active = (int)blockIdx.x < nact;

  // ---- phase I: loop entry (cg.c:104-108) -----------------------------------------------------
  if (a.init) {
    double acc = 0.0;
    for (long long i = (long long)blockIdx.x * THREADS + tid; i < nrows; i += gstride) {
      const double v = b[i];
      x[i] = 0.0;
      r[i] = v;
      pown[i] = v;
      if (pushing) push_elem(cm, i, v);
      acc = fma(v, v, acc);
    }
    if (active) {
      double v[1] = {acc};
      if (grid_reduce<THREADS, 1>(v, partials, &S->ticket[2], nullptr, 0, nact)) {
        if (tid == 0) {  // a fresh solve: the scalars of the previous one are void
          S->tol = a.tol > 0.0 ? a.tol : 0.0;
          S->max_iter = a.max_iter;
          S->variant = 0;
        }
        finish_block<THREADS>(S, cm, FIN_INIT, v[0], 0.0);
      }
      // (the reduction's partials are consumed before anyone can pass the barrier below)
      grid_arrive<THREADS>(S, cm, pushing, e + 1, nact);
    }
    if (!grid_wait(&S->grid_epoch, e + 1, S)) return;
    e += 1;
    if (ld_volatile_int(&S->done)) return;
  }

#pragma unroll 1
  for (int it = 0; it < a.iters; ++it) {
    if (tid == 0) {
      if (trace) {
        c0 = clock64();
        sh_t0 = global_ns();
      }
      if (a.wait) wait_halo(cm, S);  // the peers' slices of p have landed (all-gather by push)
      if (trace && blockIdx.x == 0) atomicAdd(trace + 0, (unsigned long long)(clock64() - c0));
    }
    // (the __syncthreads at the top of gemv_units_stream orders the wait before the block's loads)

    // ---- phase A -----------------------------------------------------------------------------
    double dot = 0.0, dot2 = 0.0;
    gemv_units_stream<THREADS, RB>(A, ld, nrows, ncols, ngs, pg, pown, s, partials, true, dot, dot2);
    if (pool_groups > 0)
      gemv_pool_phase<THREADS, RB>(A, ld, nrows, ncols, ngs, pool_groups, pg, pown, pool);
    if (trace && tid == 0) {  // spread of the CTAs' start and stream-end times (ns, global timer)
      atomicMin(trace + 8, sh_t0);
      atomicMax(trace + 9, sh_t0);
      const unsigned long long t1 = global_ns();
      atomicMin(trace + 10, t1);
      atomicMax(trace + 11, t1);
      atomicAdd(trace + 12, t1 - sh_t0);
      trace[16 + blockIdx.x] += t1 - sh_t0;  // per-CTA sum of GEMV durations (this CTA is the only writer)
      unsigned int smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      trace[16 + 1024 + blockIdx.x] = smid;
    }
    {
      double v[1] = {dot};
      if (grid_reduce<THREADS, 1>(v, partials, &S->ticket[0], partials + GEMV_GDOT_OFFSET,
                                  (int)gridDim.x - 1)) {
        if (trace && tid == 0) {
          atomicAdd(trace + 1, (unsigned long long)(clock64() - c0));
          // per-launch spreads (ns): CTA starts, ends of the CTAs' GEMV shares, first start -> first end
          volatile unsigned long long* tv = trace;
          const unsigned long long s0 = tv[8], s1 = tv[9], f0 = tv[10], f1 = tv[11];
          tv[13] += s1 - s0;
          tv[14] += f1 - f0;
          tv[15] += f0 - s0;
          tv[2] += f1 - s0;  // first start -> last end
          tv[8] = ~0ull;
          tv[10] = ~0ull;
          tv[9] = 0ull;
          tv[11] = 0ull;
        }
        if (pool_groups > 0) {  // the pool's shares of p.s, in chunk order
          const int cpc = gemv_chunks_per_group(ncols, THREADS);
          const int nchunks = pool_groups * cpc;
          const double* cdot = pool + (size_t)cpc * pool_groups * RB;
          double cd = 0.0;
          for (int i = tid; i < nchunks; i += THREADS) cd += __ldcg(cdot + i);
          __shared__ double sm_cd[THREADS / 32];
          cd = block_sum<THREADS>(cd, sm_cd);
          if (tid == 0) {
            v[0] += cd;
            *reinterpret_cast<unsigned int*>(pool + (size_t)cpc * pool_groups * RB + nchunks) = 0u;
          }
        }
        if (!a.timing) finish_block<THREADS>(S, cm, FIN_PS, v[0], 0.0);
        if (trace && tid == 0) {
          atomicAdd(trace + 3, (unsigned long long)(clock64() - c0));
          atomicAdd(trace + 7, 1ull);
        }
        grid_release<THREADS>(&S->grid_epoch, e + 1);
      }
    }
    if (!grid_wait(&S->grid_epoch, e + 1, S)) return;
    if (ld_volatile_int(&S->done)) return;

    // HBM has nothing to do until the next matvec: have L2 fetch the head of this CTA's matrix
    // range (the first row group it will stream) -- evict-first loads then find it there
    if (a.prefetch_bytes > 0 && it + 1 < a.iters && tid < RB) {
      constexpr int CW = THREADS * 4;
      const int cpg = (ncols + CW - 1) / CW;
      const long long U = (long long)ngs * cpg;
      const long long ubeg = U * blockIdx.x / gridDim.x, uend = U * (blockIdx.x + 1) / gridDim.x;
      const int g0 = (int)(ubeg / cpg), ka = (int)(ubeg % cpg);
      const int row = g0 * RB + tid;
      long long cols_ahead = min((long long)(cpg - ka), uend - ubeg) * CW;
      cols_ahead = min(cols_ahead, (long long)ncols - (long long)ka * CW);
      long long bytes = min(cols_ahead * 8, (long long)(a.prefetch_bytes / RB)) & ~15ll;
      if (row < nrows && bytes > 0)
        bulk_prefetch_l2(A + (long long)row * ld + (long long)ka * CW, (uint32_t)bytes);
    }

    // ---- phase B -----------------------------------------------------------------------------
    const double alpha = ld_volatile_f64(&S->alpha);
    const long long pool_row0 = pool_groups > 0 ? (long long)ngs * RB : (long long)nrows;
    double acc = 0.0;
    for (long long i = (long long)blockIdx.x * THREADS + tid; i < nrows; i += gstride) {
      const double pv = pown[i];
      double sv;
      if (i >= pool_row0) {  // a pool row: its chunks' partial sums, in chunk order
        const int cpc = gemv_chunks_per_group(ncols, THREADS);
        const size_t jr = (size_t)(i - pool_row0);  // == j * RB + r
        sv = 0.0;
        for (int k = 0; k < cpc; ++k) sv += __ldcg(pool + (size_t)k * pool_groups * RB + jr);
        s[i] = sv;
      } else {
        sv = __ldcg(s + i);  // written by other CTAs of this launch: read at L2
      }
      const double xn = __dadd_rn(x[i], __dmul_rn(alpha, pv));
      const double rn = __dsub_rn(r[i], __dmul_rn(alpha, sv));
      x[i] = xn;
      r[i] = rn;
      acc = fma(rn, rn, acc);
    }
    if (active) {
      double v[1] = {acc};
      if (grid_reduce<THREADS, 1>(v, partials, &S->ticket[1], nullptr, 0, nact)) {
        if (trace && tid == 0) atomicAdd(trace + 4, (unsigned long long)(clock64() - c0));
        if (!a.timing) finish_block<THREADS>(S, cm, FIN_RR, v[0], 0.0);
        if (trace && tid == 0) atomicAdd(trace + 5, (unsigned long long)(clock64() - c0));
        grid_release<THREADS>(&S->grid_epoch, e + 2);
      }
    }
    if (!grid_wait(&S->grid_epoch, e + 2, S)) return;
    if (ld_volatile_int(&S->done)) return;

    // ---- phase C -----------------------------------------------------------------------------
    const double beta = ld_volatile_f64(&S->beta);
    for (long long i = (long long)blockIdx.x * THREADS + tid; i < nrows; i += gstride) {
      const double pn = __dadd_rn(__ldcg(r + i), __dmul_rn(beta, pown[i]));
      pown[i] = pn;
      if (pushing) push_elem(cm, i, pn);
    }
    if (active) grid_arrive<THREADS>(S, cm, pushing, e + 3, nact);
    if (it + 1 == a.iters) {  // nothing of this launch reads p any more: the next launch will
      if (trace && tid == 0 && blockIdx.x == 0) atomicAdd(trace + 6, (unsigned long long)(clock64() - c0));
      return;
    }
    if (!grid_wait(&S->grid_epoch, e + 3, S)) return;
    if (trace && tid == 0 && blockIdx.x == 0) atomicAdd(trace + 6, (unsigned long long)(clock64() - c0));
    e += 3;
  }
}

// ---------------------------------------------------------------------------------------------
// Streaming vector kernels.  128-bit accesses, UNROLL independent loads per array in flight.
// The tail (n odd) is handled by one thread.
// ---------------------------------------------------------------------------------------------

constexpr int VT = 512;   // threads per CTA for the vector kernels
constexpr int VU = 4;     // unroll

// number of double2 elements the vector path may cover: all of them when every pointer is 16-byte
// aligned, none otherwise (the scalar loop then covers the whole range)
__device__ __forceinline__ long long vec_pairs(long long n, const void* a, const void* b,
                                               const void* c, const void* d) {
  const uintptr_t m = (uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d;
  return (m & 15) ? 0 : (n >> 1);
}

// number of 4-element groups the 256-bit path may cover: all of them when every pointer is 32-byte
// aligned, none otherwise.  256-bit accesses move 3 % more bytes per second than 128-bit ones on
// the mixed read/write vector kernels (tools/exp/vec_bench.cu: 6740 vs 6567 GB/s on 512^3).
__device__ __forceinline__ long long vec_quads(long long n, const void* a, const void* b,
                                               const void* c, const void* d) {
  const uintptr_t m = (uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d;
  return (m & 31) ? 0 : (n >> 2);
}
__device__ __forceinline__ double4_t ld4(const double* p) {
  double4_t v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st4(double* p, const double4_t v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z),
               "d"(v.w)
               : "memory");
}
constexpr int VU4 = 2;  // unroll of the 256-bit loops (same bytes in flight as VU x 128 bit)

__device__ __forceinline__ double2 ld2(const double* p) {
  return *reinterpret_cast<const double2*>(p);
}
__device__ __forceinline__ void st2(double* p, double2 v) { *reinterpret_cast<double2*>(p) = v; }

// K2: x += alpha p ; r -= alpha s ; partial r.r           (cg.c:115-123 and the dot of cg.c:129)
__global__ void __launch_bounds__(VT)
k_update_xr(double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
            const double* __restrict__ s, long long n, double* __restrict__ partials,
            unsigned int* __restrict__ ticket, CgState* S, const int fin, const int rev,
            const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  const double alpha = S->alpha;
  const long long n4 = vec_quads(n, x, r, p, s);
  const long long n2 = n4 > 0 ? 2 * n4 : vec_pairs(n, x, r, p, s);
  const long long stride = (long long)gridDim.x * VT;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n4; i += stride * VU4) {
    double4_t xv[VU4], rv[VU4], pv[VU4], sv[VU4];
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        xv[u] = ld4(x + 4 * j);
        pv[u] = ld4(p + 4 * j);
        rv[u] = ld4(r + 4 * j);
        sv[u] = ld4(s + 4 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x));
        xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
        xv[u].z = __dadd_rn(xv[u].z, __dmul_rn(alpha, pv[u].z));
        xv[u].w = __dadd_rn(xv[u].w, __dmul_rn(alpha, pv[u].w));
        rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x));
        rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
        rv[u].z = __dsub_rn(rv[u].z, __dmul_rn(alpha, sv[u].z));
        rv[u].w = __dsub_rn(rv[u].w, __dmul_rn(alpha, sv[u].w));
        st4(x + 4 * j, xv[u]);
        st4(r + 4 * j, rv[u]);
        acc = fma(rv[u].x, rv[u].x, acc);
        acc = fma(rv[u].y, rv[u].y, acc);
        acc = fma(rv[u].z, rv[u].z, acc);
        acc = fma(rv[u].w, rv[u].w, acc);
      }
    }
  }
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; n4 == 0 && i < n2; i += stride * VU) {
    double2 xv[VU], rv[VU], pv[VU], sv[VU];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        xv[u] = ld2(x + 2 * j);
        pv[u] = ld2(p + 2 * j);
        rv[u] = ld2(r + 2 * j);
        sv[u] = ld2(s + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x));
        xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
        rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x));
        rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
        st2(x + 2 * j, xv[u]);
        st2(r + 2 * j, rv[u]);
        acc = fma(rv[u].x, rv[u].x, acc);
        acc = fma(rv[u].y, rv[u].y, acc);
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    const double xn = __dadd_rn(x[j], __dmul_rn(alpha, p[j]));
    const double rn = __dsub_rn(r[j], __dmul_rn(alpha, s[j]));
    x[j] = xn;
    r[j] = rn;
    acc = fma(rn, rn, acc);
  }
  double v[1] = {acc};
  if (grid_reduce<VT, 1>(v, partials, ticket)) finish_block<VT>(S, cm, fin, v[0], 0.0);
}

// K3: p = r + beta p, and the rows of p a peer GPU gathers go straight into its halo slots
// (cg.c:131-132)
__global__ void __launch_bounds__(VT)
k_update_p(double* __restrict__ p, const double* __restrict__ r, long long n, const CgState* S,
           const int rev, const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  const double beta = S->beta;
  const bool pushing = cm.npush > 0;
  const long long n4 = vec_quads(n, p, r, nullptr, nullptr);
  const long long n2 = n4 > 0 ? 2 * n4 : vec_pairs(n, p, r, nullptr, nullptr);
  const long long stride = (long long)gridDim.x * VT;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n4; i += stride * VU4) {
    double4_t rv[VU4], pv[VU4];
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        rv[u] = ld4(r + 4 * j);
        pv[u] = ld4(p + 4 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        pv[u].x = __dadd_rn(rv[u].x, __dmul_rn(beta, pv[u].x));
        pv[u].y = __dadd_rn(rv[u].y, __dmul_rn(beta, pv[u].y));
        pv[u].z = __dadd_rn(rv[u].z, __dmul_rn(beta, pv[u].z));
        pv[u].w = __dadd_rn(rv[u].w, __dmul_rn(beta, pv[u].w));
        st4(p + 4 * j, pv[u]);
        if (pushing) {
          push_elem(cm, 4 * j, pv[u].x);
          push_elem(cm, 4 * j + 1, pv[u].y);
          push_elem(cm, 4 * j + 2, pv[u].z);
          push_elem(cm, 4 * j + 3, pv[u].w);
        }
      }
    }
  }
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; n4 == 0 && i < n2; i += stride * VU) {
    double2 rv[VU], pv[VU];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        rv[u] = ld2(r + 2 * j);
        pv[u] = ld2(p + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        pv[u].x = __dadd_rn(rv[u].x, __dmul_rn(beta, pv[u].x));
        pv[u].y = __dadd_rn(rv[u].y, __dmul_rn(beta, pv[u].y));
        st2(p + 2 * j, pv[u]);
        if (pushing) {
          push_elem(cm, 2 * j, pv[u].x);
          push_elem(cm, 2 * j + 1, pv[u].y);
        }
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    const double pn = __dadd_rn(r[j], __dmul_rn(beta, p[j]));
    p[j] = pn;
    if (pushing) push_elem(cm, j, pn);
  }
  push_finish(cm);
}

// K2 + K3 in one launch for systems whose vectors one wave of CTAs covers (single GPU, 32-byte
// aligned vectors, at most 8 elements per thread).  On such systems the two vector kernels move a
// few hundred KB and cost 5-9 us each -- launch, ramp and drain; the matvec of the reference's
// course-sized matrix takes 49 us.  Here the grid does K2, meets at a grid barrier behind the
// r.r reduction (the CTA that closes it forms beta, cg.c:129, and releases the epoch) and goes on
// with K3 on the r and p it still holds in registers: one launch and two L2 round trips less per
// iteration.  Same element-to-thread mapping and summation order as k_update_xr, so r.r -- and
// with it every iterate -- has the bits of the three-kernel loop.  Launched co-operatively (all
// CTAs resident: the barrier cannot starve).
__global__ void __launch_bounds__(VT)
k_update_xrp(double* __restrict__ x, double* __restrict__ r, double* __restrict__ p,
             const double* __restrict__ s, long long n, double* __restrict__ partials,
             unsigned int* __restrict__ ticket, CgState* S, const int rev,
             const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  __shared__ unsigned long long sh_e0;
  if (threadIdx.x == 0) sh_e0 = ld_acquire_gpu(&S->grid_epoch);  // nobody can have released yet
  __syncthreads();
  const unsigned long long e = sh_e0 + 1;
  const double alpha = S->alpha;
  const long long n4 = n >> 2;  // (the launcher checked the alignment)
  const long long stride = (long long)gridDim.x * VT;
  const long long i = (long long)blockIdx.x * VT + threadIdx.x;
  double acc = 0.0;
  double4_t rv[VU4], pv[VU4];
  {
    double4_t xv[VU4], sv[VU4];
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        xv[u] = ld4(x + 4 * j);
        pv[u] = ld4(p + 4 * j);
        rv[u] = ld4(r + 4 * j);
        sv[u] = ld4(s + 4 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU4; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n4 - 1 - jj : jj;
      if (jj < n4) {
        xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x));
        xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
        xv[u].z = __dadd_rn(xv[u].z, __dmul_rn(alpha, pv[u].z));
        xv[u].w = __dadd_rn(xv[u].w, __dmul_rn(alpha, pv[u].w));
        rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x));
        rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
        rv[u].z = __dsub_rn(rv[u].z, __dmul_rn(alpha, sv[u].z));
        rv[u].w = __dsub_rn(rv[u].w, __dmul_rn(alpha, sv[u].w));
        st4(x + 4 * j, xv[u]);
        st4(r + 4 * j, rv[u]);
        acc = fma(rv[u].x, rv[u].x, acc);
        acc = fma(rv[u].y, rv[u].y, acc);
        acc = fma(rv[u].z, rv[u].z, acc);
        acc = fma(rv[u].w, rv[u].w, acc);
      }
    }
  }
  // the last n mod 4 elements: one each for the first threads of the grid
  const long long jt = 4 * n4 + i;
  double rt = 0.0, pt = 0.0;
  if (jt < n) {
    pt = p[jt];
    const double xn = __dadd_rn(x[jt], __dmul_rn(alpha, pt));
    rt = __dsub_rn(r[jt], __dmul_rn(alpha, s[jt]));
    x[jt] = xn;
    r[jt] = rt;
    acc = fma(rt, rt, acc);
  }
  double v[1] = {acc};
  if (grid_reduce<VT, 1>(v, partials, ticket)) {
    finish_block<VT>(S, cm, FIN_RR, v[0], 0.0);
    grid_release<VT>(&S->grid_epoch, e);
  }
  if (!grid_wait(&S->grid_epoch, e, S)) return;
  if (ld_volatile_int(&S->done)) return;
  const double beta = ld_volatile_f64(&S->beta);
#pragma unroll
  for (int u = 0; u < VU4; ++u) {
    const long long jj = i + u * stride;
    const long long j = rev ? n4 - 1 - jj : jj;
    if (jj < n4) {
      pv[u].x = __dadd_rn(rv[u].x, __dmul_rn(beta, pv[u].x));
      pv[u].y = __dadd_rn(rv[u].y, __dmul_rn(beta, pv[u].y));
      pv[u].z = __dadd_rn(rv[u].z, __dmul_rn(beta, pv[u].z));
      pv[u].w = __dadd_rn(rv[u].w, __dmul_rn(beta, pv[u].w));
      st4(p + 4 * j, pv[u]);
    }
  }
  if (jt < n) p[jt] = __dadd_rn(rt, __dmul_rn(beta, pt));
}

// Single-reduction variant, the whole vector update in one pass (g = residual, held in the gather
// vector; d = search direction; s = A d; w = A g):
//   d = g + beta d ; s = w + beta s ; x += alpha d ; g -= alpha s ; push the new g to the peers
__global__ void __launch_bounds__(VT)
k_update_sr(double* __restrict__ x, double* __restrict__ g, double* __restrict__ d,
            double* __restrict__ s, const double* __restrict__ w, long long n, const CgState* S,
            const int rev, const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  const double alpha = S->alpha, beta = S->beta;
  const bool pushing = cm.npush > 0;
  const uintptr_t al = (uintptr_t)x | (uintptr_t)g | (uintptr_t)d | (uintptr_t)s | (uintptr_t)w;
  const long long n2 = (al & 15) ? 0 : (n >> 1);
  const long long stride = (long long)gridDim.x * VT;
  constexpr int U = 2;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n2; i += stride * U) {
    double2 xv[U], gv[U], dv[U], sv[U], wv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n2 - 1 - jj : jj;
      if (jj < n2) {
        gv[u] = ld2(g + 2 * j);
        dv[u] = ld2(d + 2 * j);
        wv[u] = ld2(w + 2 * j);
        sv[u] = ld2(s + 2 * j);
        xv[u] = ld2(x + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long jj = i + u * stride;
      const long long j = rev ? n2 - 1 - jj : jj;
      if (jj < n2) {
        dv[u].x = __dadd_rn(gv[u].x, __dmul_rn(beta, dv[u].x));
        dv[u].y = __dadd_rn(gv[u].y, __dmul_rn(beta, dv[u].y));
        sv[u].x = __dadd_rn(wv[u].x, __dmul_rn(beta, sv[u].x));
        sv[u].y = __dadd_rn(wv[u].y, __dmul_rn(beta, sv[u].y));
        xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, dv[u].x));
        xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, dv[u].y));
        gv[u].x = __dsub_rn(gv[u].x, __dmul_rn(alpha, sv[u].x));
        gv[u].y = __dsub_rn(gv[u].y, __dmul_rn(alpha, sv[u].y));
        st2(d + 2 * j, dv[u]);
        st2(s + 2 * j, sv[u]);
        st2(x + 2 * j, xv[u]);
        st2(g + 2 * j, gv[u]);
        if (pushing) {
          push_elem(cm, 2 * j, gv[u].x);
          push_elem(cm, 2 * j + 1, gv[u].y);
        }
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    const double dn = __dadd_rn(g[j], __dmul_rn(beta, d[j]));
    const double sn = __dadd_rn(w[j], __dmul_rn(beta, s[j]));
    const double xn = __dadd_rn(x[j], __dmul_rn(alpha, dn));
    const double gn = __dsub_rn(g[j], __dmul_rn(alpha, sn));
    d[j] = dn;
    s[j] = sn;
    x[j] = xn;
    g[j] = gn;
    if (pushing) push_elem(cm, j, gn);
  }
  push_finish(cm);
}

// Loop entry: x = 0, r = b, p = b, partial b.b (cg.c:104-108); p is pushed to the peers.
// Single-reduction variant (sr): the gather vector p holds the residual b, the direction buffer r
// and s start at zero (the first update then forms d = g, s = w with beta = 0).
__global__ void __launch_bounds__(VT)
k_init(double* __restrict__ x, double* __restrict__ r, double* __restrict__ p,
       double* __restrict__ s, const double* __restrict__ b, long long n, const int sr,
       double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S, const int fin,
       const __grid_constant__ PeerComm cm) {
  const long long stride = (long long)gridDim.x * VT;
  const bool pushing = cm.npush > 0;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n; i += stride) {
    const double v = b[i];
    x[i] = 0.0;
    r[i] = sr ? 0.0 : v;
    if (sr) s[i] = 0.0;
    p[i] = v;
    if (pushing) push_elem(cm, i, v);
    acc = fma(v, v, acc);
  }
  push_finish(cm);
  double v[1] = {acc};
  if (grid_reduce<VT, 1>(v, partials, ticket)) finish_block<VT>(S, cm, fin, v[0], 0.0);
}

// ---------------------------------------------------------------------------------------------
// General loop entry and the Jacobi-preconditioned vector kernels.  Extensions:
// the reference always starts from x = 0 (cg.c:104) and has no preconditioner.  Same unfused
// roundings as the reference's operators; z = dinv .* r is never stored.
// ---------------------------------------------------------------------------------------------

// b.b (threshold of the stop test); x = 0 unless the caller supplied an initial guess
__global__ void __launch_bounds__(VT)
k_init_bb(double* __restrict__ x, const double* __restrict__ b, long long n, const int zero_x,
          double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
          const int fin, const __grid_constant__ PeerComm cm) {
  const long long stride = (long long)gridDim.x * VT;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n; i += stride) {
    const double v = b[i];
    if (zero_x) x[i] = 0.0;
    acc = fma(v, v, acc);
  }
  double v[1] = {acc};
  if (grid_reduce<VT, 1>(v, partials, ticket)) finish_block<VT>(S, cm, fin, v[0], 0.0);
}

// r = b - s (s = A x0; s == nullptr: r = b), p = z = dinv .* r (dinv == nullptr: z = r), p pushed
// to the peers, partial r.z and r.r
__global__ void __launch_bounds__(VT)
k_init_r(double* __restrict__ r, double* __restrict__ p, const double* __restrict__ b,
         const double* __restrict__ s, const double* __restrict__ dinv, long long n,
         double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
         const int fin, const __grid_constant__ PeerComm cm) {
  const long long stride = (long long)gridDim.x * VT;
  const bool pushing = cm.npush > 0;
  double arz = 0.0, arr = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n; i += stride) {
    const double rv = s ? __dsub_rn(b[i], s[i]) : b[i];
    const double zv = dinv ? __dmul_rn(dinv[i], rv) : rv;
    r[i] = rv;
    p[i] = zv;
    if (pushing) push_elem(cm, i, zv);
    arz = fma(rv, zv, arz);
    arr = fma(rv, rv, arr);
  }
  push_finish(cm);
  double v[2] = {arz, arr};
  if (grid_reduce<VT, 2>(v, partials, ticket)) finish_block<VT>(S, cm, fin, v[0], v[1]);
}

// K2 with the preconditioner: x += alpha p ; r -= alpha s ; partial r.(dinv .* r) and r.r
__global__ void __launch_bounds__(VT)
k_update_xr_pc(double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
               const double* __restrict__ s, const double* __restrict__ dinv, long long n,
               double* __restrict__ partials, unsigned int* __restrict__ ticket, CgState* S,
               const int fin, const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  const double alpha = S->alpha;
  const uintptr_t al = (uintptr_t)x | (uintptr_t)r | (uintptr_t)p | (uintptr_t)s | (uintptr_t)dinv;
  const long long n2 = (al & 15) ? 0 : (n >> 1);
  const long long stride = (long long)gridDim.x * VT;
  constexpr int U = 2;
  double arz = 0.0, arr = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n2; i += stride * U) {
    double2 xv[U], rv[U], pv[U], sv[U], dv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        xv[u] = ld2(x + 2 * j);
        pv[u] = ld2(p + 2 * j);
        rv[u] = ld2(r + 2 * j);
        sv[u] = ld2(s + 2 * j);
        dv[u] = ld2(dinv + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        xv[u].x = __dadd_rn(xv[u].x, __dmul_rn(alpha, pv[u].x));
        xv[u].y = __dadd_rn(xv[u].y, __dmul_rn(alpha, pv[u].y));
        rv[u].x = __dsub_rn(rv[u].x, __dmul_rn(alpha, sv[u].x));
        rv[u].y = __dsub_rn(rv[u].y, __dmul_rn(alpha, sv[u].y));
        st2(x + 2 * j, xv[u]);
        st2(r + 2 * j, rv[u]);
        arz = fma(rv[u].x, __dmul_rn(dv[u].x, rv[u].x), arz);
        arz = fma(rv[u].y, __dmul_rn(dv[u].y, rv[u].y), arz);
        arr = fma(rv[u].x, rv[u].x, arr);
        arr = fma(rv[u].y, rv[u].y, arr);
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    const double xn = __dadd_rn(x[j], __dmul_rn(alpha, p[j]));
    const double rn = __dsub_rn(r[j], __dmul_rn(alpha, s[j]));
    x[j] = xn;
    r[j] = rn;
    arz = fma(rn, __dmul_rn(dinv[j], rn), arz);
    arr = fma(rn, rn, arr);
  }
  double v[2] = {arz, arr};
  if (grid_reduce<VT, 2>(v, partials, ticket)) finish_block<VT>(S, cm, fin, v[0], v[1]);
}

// K3 with the preconditioner: p = dinv .* r + beta p, pushed to the peers
__global__ void __launch_bounds__(VT)
k_update_p_pc(double* __restrict__ p, const double* __restrict__ r, const double* __restrict__ dinv,
              long long n, const CgState* S, const __grid_constant__ PeerComm cm) {
  if (S->done) return;
  const double beta = S->beta;
  const bool pushing = cm.npush > 0;
  const long long n2 = vec_pairs(n, p, r, dinv, nullptr);
  const long long stride = (long long)gridDim.x * VT;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n2; i += stride * VU) {
    double2 rv[VU], pv[VU], dv[VU];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        rv[u] = ld2(r + 2 * j);
        pv[u] = ld2(p + 2 * j);
        dv[u] = ld2(dinv + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        pv[u].x = __dadd_rn(__dmul_rn(dv[u].x, rv[u].x), __dmul_rn(beta, pv[u].x));
        pv[u].y = __dadd_rn(__dmul_rn(dv[u].y, rv[u].y), __dmul_rn(beta, pv[u].y));
        st2(p + 2 * j, pv[u]);
        if (pushing) {
          push_elem(cm, 2 * j, pv[u].x);
          push_elem(cm, 2 * j + 1, pv[u].y);
        }
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    const double pn = __dadd_rn(__dmul_rn(dinv[j], r[j]), __dmul_rn(beta, p[j]));
    p[j] = pn;
    if (pushing) push_elem(cm, j, pn);
  }
  push_finish(cm);
}

// dinv[i] = 1 / A_ii for the owned rows.  CSR: the entry whose gather index is diag_off + i
// (columns ascend within a row: binary search).  bad counts rows without a positive finite diagonal.
__global__ void k_diag_inv_csr(const int* __restrict__ rowptr, const int* __restrict__ cols,
                               const double* __restrict__ vals, int nrows, long long diag_off,
                               double* __restrict__ dinv, int* __restrict__ bad) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) {
    int lo = rowptr[i], hi = rowptr[i + 1];
    const long long want = diag_off + i;
    double d = 0.0;
    while (lo < hi) {
      const int mid = lo + ((hi - lo) >> 1);
      const long long c = cols[mid];
      if (c == want) {
        d = vals[mid];
        break;
      }
      if (c < want) lo = mid + 1;
      else hi = mid;
    }
    const bool okd = d > 0.0 && !isinf(d);
    dinv[i] = okd ? 1.0 / d : 0.0;
    if (!okd) atomicAdd(bad, 1);
  }
}

__global__ void k_diag_inv_dense(const double* __restrict__ vals, long long ld, int nrows,
                                 long long row0, double* __restrict__ dinv, int* __restrict__ bad) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) {
    const double d = vals[(size_t)i * ld + row0 + i];
    const bool okd = d > 0.0 && !isinf(d);
    dinv[i] = okd ? 1.0 / d : 0.0;
    if (!okd) atomicAdd(bad, 1);
  }
}

// Stand-alone push of the owned block's boundary rows (operator-API matvecs, residual checks):
// the loop itself never launches this, its vector kernels push as they write.
__global__ void __launch_bounds__(VT)
k_push(const double* __restrict__ p, const __grid_constant__ PeerComm cm) {
  const long long stride = (long long)gridDim.x * VT;
  for (int t = 0; t < cm.npush; ++t) {
    const long long b0 = cm.push[t].begin, cnt = cm.push[t].end - b0;
    double* dst = cm.push[t].dst;
    for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < cnt; i += stride)
      dst[i] = p[b0 + i];
  }
  push_finish(cm);
}

// One-thread bookkeeping kernel used after an NCCL sum of S->local (NCCL mode only).
__global__ void k_finish(CgState* S, int which) {
  if (which != FIN_INIT && which != FIN_BB && S->done) return;
  apply_finish(S, which, S->local[0], S->local[1]);
}

// Standalone operators of mv_ops.h ------------------------------------------------------------

// dot_product (mv_ops.c:117-132): result left in *out (device)
__global__ void __launch_bounds__(VT)
k_dot(const double* __restrict__ a, const double* __restrict__ b, long long n,
      double* __restrict__ partials, unsigned int* __restrict__ ticket, double* __restrict__ out) {
  const long long n2 = vec_pairs(n, a, b, nullptr, nullptr);
  const long long stride = (long long)gridDim.x * VT;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n2; i += stride * VU) {
    double2 av[VU], bv[VU];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        av[u] = ld2(a + 2 * j);
        bv[u] = ld2(b + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        acc = fma(av[u].x, bv[u].x, acc);
        acc = fma(av[u].y, bv[u].y, acc);
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride)
    acc = fma(a[j], b[j], acc);
  double total;
  if (grid_sum<VT>(acc, partials, ticket, total)) *out = total;
}

// op: 0 = scale (out = sca*a; mv_ops.c:134-158), 1 = add (203-230), 2 = sub (232-259);
// 3 / 4 = a +/- fl(sca*b): a scale whose only reader is the add/sub that follows it (the operator
// API's resident mode defers the scale, abi.cu) -- the same two roundings in one pass;
// 5 / 6 = the same with fl(sca*b) stored to out2 too (the add/sub overwrites the scale's source).
// out may alias a or b (element-wise, each element read before it is written).
template <int OP>
__global__ void __launch_bounds__(VT)
k_ewise(double sca, const double* a, const double* b, double* out, long long n, double* out2) {
  const long long n2 = vec_pairs(n, a, OP != 0 ? b : nullptr, out, OP >= 5 ? out2 : nullptr);
  const long long stride = (long long)gridDim.x * VT;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n2; i += stride * VU) {
    double2 av[VU], bv[VU];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        av[u] = ld2(a + 2 * j);
        if (OP != 0) bv[u] = ld2(b + 2 * j);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long j = i + u * stride;
      if (j < n2) {
        double2 o;
        if (OP == 0) {
          o.x = __dmul_rn(sca, av[u].x);
          o.y = __dmul_rn(sca, av[u].y);
        } else if (OP == 1) {
          o.x = __dadd_rn(av[u].x, bv[u].x);
          o.y = __dadd_rn(av[u].y, bv[u].y);
        } else if (OP == 2) {
          o.x = __dsub_rn(av[u].x, bv[u].x);
          o.y = __dsub_rn(av[u].y, bv[u].y);
        } else if (OP == 3) {  // a + fl(sca * b): sv_mult then vec_add, same two roundings
          o.x = __dadd_rn(av[u].x, __dmul_rn(sca, bv[u].x));
          o.y = __dadd_rn(av[u].y, __dmul_rn(sca, bv[u].y));
        } else if (OP == 4) {  // a - fl(sca * b)
          o.x = __dsub_rn(av[u].x, __dmul_rn(sca, bv[u].x));
          o.y = __dsub_rn(av[u].y, __dmul_rn(sca, bv[u].y));
        } else {               // 5 / 6: as 3 / 4, and the scaled vector is stored as well (out2)
          double2 t;
          t.x = __dmul_rn(sca, bv[u].x);
          t.y = __dmul_rn(sca, bv[u].y);
          st2(out2 + 2 * j, t);
          o.x = OP == 5 ? __dadd_rn(av[u].x, t.x) : __dsub_rn(av[u].x, t.x);
          o.y = OP == 5 ? __dadd_rn(av[u].y, t.y) : __dsub_rn(av[u].y, t.y);
        }
        st2(out + 2 * j, o);
      }
    }
  }
  for (long long j = 2 * n2 + (long long)blockIdx.x * VT + threadIdx.x; j < n; j += stride) {
    if (OP >= 5) {
      const double t = __dmul_rn(sca, b[j]);
      const double av = a[j];
      out2[j] = t;
      out[j] = OP == 5 ? __dadd_rn(av, t) : __dsub_rn(av, t);
    } else {
      out[j] = OP == 0   ? __dmul_rn(sca, a[j])
               : OP == 1 ? __dadd_rn(a[j], b[j])
               : OP == 2 ? __dsub_rn(a[j], b[j])
               : OP == 3 ? __dadd_rn(a[j], __dmul_rn(sca, b[j]))
                         : __dsub_rn(a[j], __dmul_rn(sca, b[j]));
    }
  }
}

// r = b - s and partial sums of r.r and b.b (independent true-residual check)
__global__ void __launch_bounds__(VT)
k_resid_norms(const double* __restrict__ b, const double* __restrict__ s, long long n,
              double* __restrict__ partials, unsigned int* __restrict__ ticket,
              double* __restrict__ out2) {
  const long long stride = (long long)gridDim.x * VT;
  double num = 0.0, den = 0.0;
  for (long long i = (long long)blockIdx.x * VT + threadIdx.x; i < n; i += stride) {
    const double d = b[i] - s[i];
    num = fma(d, d, num);
    den = fma(b[i], b[i], den);
  }
  double total;
  if (grid_sum<VT>(num, partials, ticket, total)) out2[0] = total;
  if (grid_sum<VT>(den, partials + gridDim.x, ticket + 1, total)) out2[1] = total;
}

// ---------------------------------------------------------------------------------------------
// Matrix analysis / generators (set-up, not on the iteration loop)
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ double uniform_pm_half(unsigned long long seed, unsigned long long ctr) {
  const unsigned long long h = splitmix64(splitmix64(seed) ^ (ctr * 0xD1342543DE82EF95ull + 1ull));
  return (double)(h >> 11) * (1.0 / 9007199254740992.0) - 0.5;
}

__global__ void k_gen_rhs(double* b, long long row0, long long n, unsigned long long seed) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    b[i] = uniform_pm_half(seed, (unsigned long long)(row0 + i));
}

// Stored entries in rows [0, row) of the Dirichlet 5/7-point operator on nx*ny*nz, natural order.
__device__ __forceinline__ long long lap_nnz_before(long long row, long long nx, long long ny,
                                                    long long nz) {
  const long long pl = nx * ny;
  const long long L = row / nx, xx = row % nx;
  const long long P = row / pl, q = row % pl;
  long long missing = 0;
  missing += L + (xx > 0 ? 1 : 0);                                   // x == 0      (no -x)
  missing += L;                                                      // x == nx-1   (no +x)
  missing += P * nx + (q < nx ? q : nx);                             // y == 0
  missing += P * nx + (q > (ny - 1) * nx ? q - (ny - 1) * nx : 0);   // y == ny-1
  missing += row < pl ? row : pl;                                    // z == 0
  missing += row > (nz - 1) * pl ? row - (nz - 1) * pl : 0;          // z == nz-1
  return 7 * row - missing;
}

// Local rows [row0, row0+nloc) of the operator; columns are written already shifted by
// -col_shift (the local gather index space).
__global__ void k_gen_lap(int nx, int ny, int nz, long long row0, int nloc, long long col_shift,
                          int* __restrict__ rowptr, int* __restrict__ cols,
                          double* __restrict__ vals) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long pl = (long long)nx * ny;
  const double diag = nz == 1 ? 4.0 : 6.0;
  const long long base = lap_nnz_before(row0, nx, ny, nz);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i <= nloc; i += stride) {
    const long long row = row0 + i;
    long long k = lap_nnz_before(row, nx, ny, nz) - base;
    rowptr[i] = (int)k;
    if (i == nloc) break;
    const int xx = (int)(row % nx), yy = (int)((row / nx) % ny), zz = (int)(row / pl);
    if (zz > 0) { cols[k] = (int)(row - pl - col_shift); vals[k++] = -1.0; }
    if (yy > 0) { cols[k] = (int)(row - nx - col_shift); vals[k++] = -1.0; }
    if (xx > 0) { cols[k] = (int)(row - 1 - col_shift); vals[k++] = -1.0; }
    cols[k] = (int)(row - col_shift); vals[k++] = diag;
    if (xx < nx - 1) { cols[k] = (int)(row + 1 - col_shift); vals[k++] = -1.0; }
    if (yy < ny - 1) { cols[k] = (int)(row + nx - col_shift); vals[k++] = -1.0; }
    if (zz < nz - 1) { cols[k] = (int)(row + pl - col_shift); vals[k++] = -1.0; }
  }
}

// Rows [row0, row0+nloc) of the dense SPD test matrix, values only (row-major, ld = n).
__global__ void k_gen_dense(int n, int row0, int nloc, unsigned long long seed,
                            double* __restrict__ vals) {
  const long long total = (long long)nloc * n;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const long long i = row0 + e / n, j = e % n;
    const long long lo = i < j ? i : j, hi = i < j ? j : i;
    vals[e] = (i == j) ? 0.5 * n + 1.0
                       : uniform_pm_half(seed, (unsigned long long)lo * (unsigned long long)n +
                                                   (unsigned long long)hi);
  }
}

// min / max column index over all stored entries (halo discovery) and the "every row is full"
// test (row_ptr[i] == i*ncols).  out[0] = min col, out[1] = max col, out[2] = #rows not full,
// out[3] = #decreasing steps of row_ptr (a valid CSR has none).
__global__ void k_analyze(const int* __restrict__ rowptr, const int* __restrict__ cols, int nrows,
                          long long nnz, int ncols, int* __restrict__ out) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  int mn = 0x7fffffff, mx = -0x7fffffff - 1, bad = 0, dec = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride)
    if (rowptr[i + 1] < rowptr[i]) ++dec;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride) {
    const int c = cols[i];
    mn = min(mn, c);
    mx = max(mx, c);
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i <= nrows; i += stride)
    if ((long long)rowptr[i] != i * (long long)ncols) ++bad;
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    bad += __shfl_xor_sync(0xffffffffu, bad, o);
    dec += __shfl_xor_sync(0xffffffffu, dec, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out + 0, mn);
    atomicMax(out + 1, mx);
    if (bad) atomicAdd(out + 2, bad);
    if (dec) atomicAdd(out + 3, dec);
  }
}

// Largest (gather index of a row's last entry) - (local row) over all rows: the offset of the upper
// band edge in the gather vector (columns are ascending within a row).
__global__ void k_upper_band(const int* __restrict__ rowptr, const int* __restrict__ cols, int nrows,
                             int* __restrict__ out) {
  const int stride = gridDim.x * blockDim.x;
  int mx = -0x7fffffff - 1;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < nrows; r += stride) {
    const int a = rowptr[r], b = rowptr[r + 1];
    if (b > a) mx = max(mx, cols[b - 1] - r);
  }
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, mx);
}

__global__ void k_shift_cols(int* __restrict__ cols, long long nnz, int shift) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride)
    cols[i] -= shift;
}

// ---------------------------------------------------------------------------------------------
// Launch wrappers
// ---------------------------------------------------------------------------------------------

static int g_sms = 0;
int sm_count() {
  if (!g_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms <= 0) g_sms = 148;
  }
  return g_sms;
}

static inline int vec_grid(long long n) {
  const long long want = (n / 2 + (long long)VT * VU - 1) / ((long long)VT * VU);
  const long long cap = (long long)sm_count() * 4;
  long long g = want < cap ? want : cap;
  return (int)(g < 1 ? 1 : g);
}

int max_partials() { return PART_STRIDE; }
int partials_doubles() { return 3 * PART_STRIDE + 1024; }  // two reduction values, GEMV records, group counters
size_t partials_zero_offset() { return (size_t)GEMV_CNT_OFFSET; }  // doubles from here on start as zero

// SpMV tile configurations.  One is chosen per process (CGB200_SPMV_CFG, default 0); DESIGN.md
// records the sweep that picked the default.
struct SpmvCfg {
  int R, CAP, STAGES;
  size_t smem;
  const void* fn;
  void (*launch)(int grid, cudaStream_t st, const CsrDev& A, const double* pg, const double* pown,
                 double* s, const TileRange& tr, int wait_from, long long lead, long long plen,
                 double* partials, unsigned int* ticket, CgState* S, int mode, int hint,
                 const PeerComm& cm);
};

template <int R, int CAP, int STAGES>
static void spmv_launch_t(int grid, cudaStream_t st, const CsrDev& A, const double* pg,
                          const double* pown, double* s, const TileRange& tr, int wait_from,
                          long long lead, long long plen, double* partials, unsigned int* ticket,
                          CgState* S, int mode, int hint, const PeerComm& cm) {
  k_spmv_stream<R, CAP, STAGES><<<grid, R + 32, SpmvSmem<R, CAP, STAGES>::total, st>>>(
      A.vals, A.cols, A.rowptr, A.nrows, pg, pown, s, tr, wait_from, lead, plen, partials, ticket,
      S, mode, hint, cm);
}

#define SPMV_CFG(R, CAP, ST)                                                        \
  SpmvCfg {                                                                         \
    R, CAP, ST, SpmvSmem<R, CAP, ST>::total, (const void*)k_spmv_stream<R, CAP, ST>, \
        spmv_launch_t<R, CAP, ST>                                                   \
  }

static const SpmvCfg g_spmv_cfgs[] = {
    SPMV_CFG(512, 3584, 2),  // 0  default: 2 CTAs/SM x 16 consumer warps, double-buffered, 7 nnz/row
    SPMV_CFG(256, 2048, 2),  // 1
    SPMV_CFG(256, 2048, 4),  // 2
    SPMV_CFG(128, 1024, 4),  // 3
    SPMV_CFG(128, 1024, 3),  // 4
    SPMV_CFG(256, 2048, 3),  // 5
    SPMV_CFG(512, 4096, 3),  // 6
    SPMV_CFG(512, 4096, 2),  // 7
    SPMV_CFG(384, 2688, 2),  // 8
    SPMV_CFG(608, 4256, 2),  // 9
    SPMV_CFG(480, 3840, 1),  // 10
    SPMV_CFG(992, 7936, 1),  // 11
    SPMV_CFG(576, 4032, 2),  // 12
};

static int g_spmv_cfg = -1, g_spmv_bps = 0, g_spmv_hint = 1;

static const SpmvCfg& spmv_cfg() {
  if (g_spmv_cfg < 0) {
    const char* e = getenv("CGB200_SPMV_CFG");
    int i = e ? atoi(e) : 0;
    const int ncfg = (int)(sizeof(g_spmv_cfgs) / sizeof(g_spmv_cfgs[0]));
    if (i < 0 || i >= ncfg) i = 0;
    g_spmv_cfg = i;
    const char* h = getenv("CGB200_SPMV_HINT");
    if (h) g_spmv_hint = atoi(h);
    const SpmvCfg& c = g_spmv_cfgs[i];
    cudaFuncSetAttribute(c.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_spmv_bps, c.fn, c.R + 32, c.smem);
    if (g_spmv_bps < 1) g_spmv_bps = 1;
    const char* b = getenv("CGB200_SPMV_BPS");
    if (b && atoi(b) > 0 && atoi(b) < g_spmv_bps) g_spmv_bps = atoi(b);
  }
  return g_spmv_cfgs[g_spmv_cfg];
}

int spmv_rows_per_tile() { return spmv_cfg().R; }

int spmv_grid_for(int ntiles) {
  spmv_cfg();
  int grid = sm_count() * g_spmv_bps;
  if (grid > MAX_GRID) grid = MAX_GRID;
  return grid > ntiles ? ntiles : grid;
}

cudaError_t launch_spmv(const CsrDev& A, const double* pg, const double* pown, double* s,
                        const TileRange& tr, int wait_from, double* partials,
                        unsigned int* ticket, CgState* S, int mode, const PeerComm& cm,
                        cudaStream_t st) {
  const int nvt = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);
  if (nvt <= 0) return cudaSuccess;
  const SpmvCfg& c = spmv_cfg();
  c.launch(spmv_grid_for(nvt), st, A, pg, pown, s, tr, wait_from, A.lead, A.plen, partials, ticket,
           S, mode, g_spmv_hint, cm);
  ++g_launches;
  return cudaGetLastError();
}

// Mean row length from which a single-GPU CSR block takes a long-row kernel instead of the stream
// kernel (CGB200_SPMV_ROWGROUP_MIN; 0 = never).  Measured on B200 (tests/rowgroup_sweep.py,
// profiles/r01f_longrow_sweep.log), matvec time stream / group(2 lanes) / row block(16 rows, 256 thr):
//   33 entries per row, n = 1e6:    0.075 / 0.125 / 0.120 ms     (stream at 0.85 of the HBM peak)
//   65 entries per row, n = 4e5:    0.080 / 0.097 / 0.078 ms
//   345 entries per row, n = 52269: 0.333 / 0.061 / 0.049 ms     (row block at 0.67 of the HBM peak)
// The stream kernel's per-stage hand-over dominates once a 512-row tile needs tens of stages.
int spmv_rowgroup_min() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CGB200_SPMV_ROWGROUP_MIN");
    v = e ? atoi(e) : 64;
    if (v < 0) v = 0;
  }
  return v;
}

template <int T, int G, int UNR>
static cudaError_t launch_spmv_rowgroup_cfg(const CsrDev& A, const double* pg, const double* pown,
                                            double* s, double* partials, unsigned int* ticket,
                                            CgState* S, int mode, const PeerComm& cm,
                                            cudaStream_t st) {
  static int bps = 0;
  if (!bps) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_rowgroup<T, G, UNR>, T, 0);
    if (bps < 1) bps = 1;
  }
  const int rows_per_pass = T / G;
  int grid = sm_count() * bps;
  const int need = (A.nrows + rows_per_pass - 1) / rows_per_pass;
  if (grid > need) grid = need;
  if (grid > MAX_GRID) grid = MAX_GRID;
  k_spmv_rowgroup<T, G, UNR><<<grid, T, 0, st>>>(A.vals, A.cols, A.rowptr, A.nrows, pg, pown, s,
                                                 partials, ticket, S, mode, cm);
  ++g_launches;
  return cudaGetLastError();
}

template <int T, int ROWS, bool COH = false>
static cudaError_t launch_spmv_rowblock_cfg(const CsrDev& A, const double* pg, const double* pown,
                                            double* s, double* partials, unsigned int* ticket,
                                            CgState* S, int mode, const PeerComm& cm,
                                            cudaStream_t st) {
  static int bps = 0;
  if (!bps) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_rowblock<T, ROWS, COH>, T, 0);
    if (bps < 1) bps = 1;
  }
  const int nblk = (A.nrows + ROWS - 1) / ROWS;
  int grid = sm_count() * bps;
  const int need = (nblk + T / 32 - 1) / (T / 32);
  if (grid > need) grid = need;
  if (grid > MAX_GRID) grid = MAX_GRID;
  k_spmv_rowblock<T, ROWS, COH><<<grid, T, 0, st>>>(A.vals, A.cols, A.rowptr, A.nrows, pg, pown, s,
                                                    partials, ticket, S, mode, cm);
  ++g_launches;
  return cudaGetLastError();
}

// The default long-row kernel behind the peers' completion flags (several GPUs, peer-memory mode).
cudaError_t launch_spmv_longrow_coherent(const CsrDev& A, const double* pg, const double* pown,
                                         double* s, double* partials, unsigned int* ticket,
                                         CgState* S, int mode, const PeerComm& cm, cudaStream_t st) {
  if (A.nrows <= 0) return cudaSuccess;
  return launch_spmv_rowblock_cfg<256, 16, true>(A, pg, pown, s, partials, ticket, S, mode, cm, st);
}

cudaError_t launch_spmv_rowgroup(const CsrDev& A, const double* pg, const double* pown, double* s,
                                 double* partials, unsigned int* ticket, CgState* S, int mode,
                                 const PeerComm& cm, cudaStream_t st) {
  if (A.nrows <= 0) return cudaSuccess;
  // CGB200_SPMV_LONGROW: group = the shuffle-chain kernel; block, block16, block8 (optionally
  // followed by "x256": threads per CTA) = the shared-memory transposition with 32 / 16 / 8 rows
  // per warp and work unit
  static int form = -1;
  if (form < 0) {
    const char* e = getenv("CGB200_SPMV_LONGROW");
    form = 1016;  // block16x256 (tests/rowgroup_sweep.py)
    if (e && !strcmp(e, "group")) form = 0;
    if (e && !strncmp(e, "block", 5)) {
      const int rows = atoi(e + 5) > 0 ? atoi(e + 5) : 32;
      form = (rows == 8 ? 8 : rows == 16 ? 16 : 32) + (strstr(e, "x256") ? 1000 : 0);
    }
  }
#define RB_CASE(F, T, R) \
  if (form == F) return launch_spmv_rowblock_cfg<T, R>(A, pg, pown, s, partials, ticket, S, mode, cm, st)
  RB_CASE(32, 128, 32);
  RB_CASE(16, 128, 16);
  RB_CASE(8, 128, 8);
  RB_CASE(1016, 256, 16);
  RB_CASE(1008, 256, 8);
#undef RB_CASE
  // lanes per row, entries per lane in flight, threads per CTA (CGB200_SPMV_ROWGROUP_G = "G,UNR,T"):
  // fewer lanes share the addition chain between more rows of a warp, more lanes make the loads
  // of a row wider and give a small matrix more threads
  static int cfg = -1;
  if (cfg < 0) {
    int g = 2, u = 8, t = 256;
    const char* e = getenv("CGB200_SPMV_ROWGROUP_G");
    if (e) sscanf(e, "%d,%d,%d", &g, &u, &t);
    cfg = g * 10000 + u * 1000 + t;
  }
#define RG_CASE(G, U, T)   \
  case G * 10000 + U * 1000 + T: \
    return launch_spmv_rowgroup_cfg<T, G, U>(A, pg, pown, s, partials, ticket, S, mode, cm, st)
  switch (cfg) {
    RG_CASE(2, 8, 128);
    RG_CASE(4, 8, 256);
    RG_CASE(8, 4, 256);
    RG_CASE(32, 1, 256);
    default: break;
  }
#undef RG_CASE
  return launch_spmv_rowgroup_cfg<256, 2, 8>(A, pg, pown, s, partials, ticket, S, mode, cm, st);
}

constexpr int GV_T = 256;
constexpr int GV_RB = 4;

template <int VEC>
static cudaError_t launch_gemv_vec(const double* A, long long ld, int nrows, int ncols,
                                   const double* pg, const double* pown, double* s,
                                   double* partials, unsigned int* ticket, CgState* S, int mode,
                                   int wait, const PeerComm& cm, cudaStream_t st) {
  static int bps = 0;
  if (!bps) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_gemv_rows<GV_T, GV_RB, VEC>, GV_T, 0);
    if (bps < 1) bps = 1;
  }
  const int ngroups = (nrows + GV_RB - 1) / GV_RB;
  int grid = sm_count() * bps;
  if (grid > ngroups) grid = ngroups;
  if (grid > MAX_GRID) grid = MAX_GRID;
  k_gemv_rows<GV_T, GV_RB, VEC><<<grid, GV_T, 0, st>>>(A, ld, nrows, ncols, pg, pown, s, partials,
                                                      ticket, S, mode, wait, cm);
  ++g_launches;
  return cudaGetLastError();
}

static cudaError_t launch_gemv_units(const double* A, long long ld, int nrows, int ncols,
                                     const double* pg, const double* pown, double* s,
                                     double* partials, unsigned int* ticket, CgState* S, int mode,
                                     int wait, const PeerComm& cm, cudaStream_t st) {
  static int bps = 0;
  if (!bps) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_gemv_units<GV_T, GV_RB>, GV_T, 0);
    if (bps < 1) bps = 1;
    const char* e = getenv("CGB200_GEMV_BPS");  // fewer CTAs per SM (experiments)
    if (e && atoi(e) > 0 && atoi(e) < bps) bps = atoi(e);
  }
  const long long units =
      (long long)((nrows + GV_RB - 1) / GV_RB) * ((ncols + GV_T * 4 - 1) / (GV_T * 4));
  long long grid = (long long)sm_count() * bps;
  if (grid > units) grid = units;
  const long long cap = GEMV_MAX_GRID;  // room for two records per CTA
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  k_gemv_units<GV_T, GV_RB><<<(int)grid, GV_T, 0, st>>>(A, ld, nrows, ncols, pg, pown, s, partials,
                                                       ticket, S, mode, wait, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_gemv(const double* A, long long ld, int nrows, int ncols, const double* pg,
                        const double* pown, double* s, double* partials, unsigned int* ticket,
                        CgState* S, int mode, int wait, const PeerComm& cm, cudaStream_t st) {
  static int force_rows = -1;
  if (force_rows < 0) {
    const char* e = getenv("CGB200_GEMV_ROWS");  // 1 = whole row groups per CTA (comparison runs)
    force_rows = (e && atoi(e) > 0) ? 1 : 0;
  }
  const bool al32 = ((uintptr_t)A % 32 == 0) && ((uintptr_t)pg % 32 == 0);
  const bool al16 = ((uintptr_t)A % 16 == 0) && ((uintptr_t)pg % 16 == 0);
  if (ld % 4 == 0 && ncols % 4 == 0 && al32) {
    if (!force_rows)
      return launch_gemv_units(A, ld, nrows, ncols, pg, pown, s, partials, ticket, S, mode, wait,
                               cm, st);
    return launch_gemv_vec<4>(A, ld, nrows, ncols, pg, pown, s, partials, ticket, S, mode, wait,
                              cm, st);
  }
  if (ld % 2 == 0 && ncols % 2 == 0 && al16)
    return launch_gemv_vec<2>(A, ld, nrows, ncols, pg, pown, s, partials, ticket, S, mode, wait,
                              cm, st);
  return launch_gemv_vec<1>(A, ld, nrows, ncols, pg, pown, s, partials, ticket, S, mode, wait, cm,
                            st);
}

// The one-kernel iteration needs the unit-range GEMV's alignment (256-bit loads) and a grid that is
// co-resident by construction: cooperative launch, so an impossible grid is an error, never a hang.
static int g_dense_fused = -1;
void set_dense_iter_enabled(int on) { g_dense_fused = on ? 1 : 0; }
bool dense_iter_enabled() {
  if (g_dense_fused < 0) {
    const char* e = getenv("CGB200_DENSE_FUSED");  // 0 = three kernels per iteration (comparison runs)
    g_dense_fused = (e && atoi(e) == 0) ? 0 : 1;
  }
  return g_dense_fused != 0;
}
bool dense_iter_supported(const double* A, long long ld, int ncols, const double* pg) {
  if (!dense_iter_enabled()) return false;
  return ld % 4 == 0 && ncols % 4 == 0 && ((uintptr_t)A % 32 == 0) && ((uintptr_t)pg % 32 == 0);
}

// Row groups left to the dynamic pool of the one-kernel iteration: about 3 % of the matrix plus
// 40 MB (the end-of-launch spread has a fixed and a proportional part), at most a quarter, none
// for small blocks.  CGB200_GEMV_POOL_PCT overrides the percentage (0 = static split only).
int dense_pool_groups(int nrows, int ncols) {
  static int pct = -2;
  if (pct == -2) {
    const char* e = getenv("CGB200_GEMV_POOL_PCT");
    pct = e ? atoi(e) : -1;
  }
  const int ngroups = (nrows + GV_RB - 1) / GV_RB;
  const double bytes = 8.0 * (double)nrows * (double)ncols;
  if (bytes < 64e6 || ngroups < 4 * sm_count()) return 0;
  double f = pct >= 0 ? pct / 100.0 : 0.03 + 40e6 / bytes;
  if (f > 0.25) f = 0.25;
  int P = (int)(f * ngroups);
  if (P > 8192) P = 8192;
  return P < 0 ? 0 : P;
}

size_t dense_pool_doubles(int nrows, int ncols) {
  const int P = dense_pool_groups(nrows, ncols);
  if (P <= 0) return 0;
  const size_t cpc = (size_t)gemv_chunks_per_group(ncols, GV_T);
  return cpc * P * GV_RB + cpc * P + 8;
}

// CGB200_DENSE_PREFETCH_KB: matrix bytes per CTA asked into L2 ahead of the next iteration's matvec.
static int dense_prefetch_bytes() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CGB200_DENSE_PREFETCH_KB");
    v = (e ? atoi(e) : 96) * 1024;
    if (v < 0) v = 0;
  }
  return v;
}

cudaError_t launch_cg_dense_solve(const double* A, long long ld, int nrows, int ncols,
                                  const double* pg, double* pown, double* x, double* r, double* s,
                                  const double* b, double* partials, CgState* S, int iters,
                                  int init, int max_iter, double tol, int wait, int timing,
                                  unsigned long long* trace, double* pool, const PeerComm& cm,
                                  cudaStream_t st) {
  static int bps = 0;
  if (!bps) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_cg_dense_solve<GV_T, GV_RB>, GV_T, 0);
    if (bps < 1) bps = 1;
    const char* e = getenv("CGB200_GEMV_BPS");  // fewer CTAs per SM (experiments)
    if (e && atoi(e) > 0 && atoi(e) < bps) bps = atoi(e);
  }
  const int P = pool ? dense_pool_groups(nrows, ncols) : 0;
  const long long units =
      (long long)((nrows + GV_RB - 1) / GV_RB - P) * ((ncols + GV_T * 4 - 1) / (GV_T * 4));
  long long grid = (long long)sm_count() * bps;
  if (grid > units) grid = units;
  const long long cap = GEMV_MAX_GRID;  // room for two records per CTA
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  DenseSolveArgs a;
  a.iters = iters;
  a.init = init;
  a.max_iter = max_iter;
  a.tol = tol;
  a.wait = wait;
  a.timing = timing;
  a.pool_groups = P;
  a.prefetch_bytes = dense_prefetch_bytes();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(GV_T);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, k_cg_dense_solve<GV_T, GV_RB>, A, ld, nrows, ncols,
                                           pg, pown, x, r, s, b, partials, S, trace, pool, a, cm);
  ++g_launches;
  return e != cudaSuccess ? e : cudaGetLastError();
}

cudaError_t launch_update_xr(double* x, double* r, const double* p, const double* s, long long n,
                             double* partials, unsigned int* ticket, CgState* S, int fin, int rev,
                             const PeerComm& cm, cudaStream_t st) {
  k_update_xr<<<vec_grid(n), VT, 0, st>>>(x, r, p, s, n, partials, ticket, S, fin, rev, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_update_p(double* p, const double* r, long long n, const CgState* S, int rev,
                            const PeerComm& cm, cudaStream_t st) {
  k_update_p<<<vec_grid(n), VT, 0, st>>>(p, r, n, S, rev, cm);
  ++g_launches;
  return cudaGetLastError();
}

// K2 + K3 fused (k_update_xrp) when one resident wave of CTAs covers the vectors; false = not this system
static int g_fuse_vec = -1;  // -1: CGB200_FUSE_VEC not looked at yet
void set_fused_vector_update(int on) { g_fuse_vec = on ? 1 : 0; }

bool update_xrp_applies(const double* x, const double* r, const double* p, const double* s, long long n) {
  static int limit = -1;
  if (g_fuse_vec < 0) {
    const char* e = getenv("CGB200_FUSE_VEC");
    g_fuse_vec = (e && atoi(e) == 0) ? 0 : 1;
  }
  const int on = g_fuse_vec;
  if (limit < 0) {
    int bps = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_update_xrp, VT, 0) != cudaSuccess) bps = 0;
    limit = bps * sm_count();
  }
  if (!on || n < 1) return false;
  if ((((uintptr_t)x | (uintptr_t)r | (uintptr_t)p | (uintptr_t)s) & 31) != 0) return false;
  const long long want = (n / 2 + (long long)VT * VU - 1) / ((long long)VT * VU);
  // one pass: every thread holds its elements (at most 2 quads + 1) across the barrier.  Worth it up to
  // half a wave of CTAs (measured, it/s fused vs three kernels: 32x32 grid 87.4k vs 81.0k, course-sized
  // band 17.46k vs 17.03k, 512x512 grid 59.3k vs 58.8k; at 512 000 unknowns 47.4k vs 48.4k: not there)
  return want <= limit && want <= (long long)sm_count() / 2;
}

cudaError_t launch_update_xrp(double* x, double* r, double* p, const double* s, long long n,
                              double* partials, unsigned int* ticket, CgState* S, int rev,
                              const PeerComm& cm, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)vec_grid(n));
  cfg.blockDim = dim3(VT);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, k_update_xrp, x, r, p, s, n, partials, ticket, S, rev, cm);
  ++g_launches;
  return e != cudaSuccess ? e : cudaGetLastError();
}

cudaError_t launch_update_sr(double* x, double* g, double* d, double* s, const double* w,
                             long long n, const CgState* S, int rev, const PeerComm& cm,
                             cudaStream_t st) {
  k_update_sr<<<vec_grid(n), VT, 0, st>>>(x, g, d, s, w, n, S, rev, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_init(double* x, double* r, double* p, double* s, const double* b, long long n,
                        int sr, double* partials, unsigned int* ticket, CgState* S, int fin,
                        const PeerComm& cm, cudaStream_t st) {
  k_init<<<vec_grid(n), VT, 0, st>>>(x, r, p, s, b, n, sr, partials, ticket, S, fin, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_push(const double* p, const PeerComm& cm, cudaStream_t st) {
  if (cm.npush <= 0) return cudaSuccess;
  long long most = 1;
  for (int t = 0; t < cm.npush; ++t)
    most = std::max(most, cm.push[t].end - cm.push[t].begin);
  k_push<<<vec_grid(2 * most), VT, 0, st>>>(p, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_finish(CgState* S, int which, cudaStream_t st) {
  k_finish<<<1, 1, 0, st>>>(S, which);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_init_bb(double* x, const double* b, long long n, int zero_x, double* partials,
                           unsigned int* ticket, CgState* S, int fin, const PeerComm& cm,
                           cudaStream_t st) {
  k_init_bb<<<vec_grid(n), VT, 0, st>>>(x, b, n, zero_x, partials, ticket, S, fin, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_init_r(double* r, double* p, const double* b, const double* s,
                          const double* dinv, long long n, double* partials, unsigned int* ticket,
                          CgState* S, int fin, const PeerComm& cm, cudaStream_t st) {
  k_init_r<<<vec_grid(n), VT, 0, st>>>(r, p, b, s, dinv, n, partials, ticket, S, fin, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_update_xr_pc(double* x, double* r, const double* p, const double* s,
                                const double* dinv, long long n, double* partials,
                                unsigned int* ticket, CgState* S, int fin, const PeerComm& cm,
                                cudaStream_t st) {
  k_update_xr_pc<<<vec_grid(n), VT, 0, st>>>(x, r, p, s, dinv, n, partials, ticket, S, fin, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_update_p_pc(double* p, const double* r, const double* dinv, long long n,
                               const CgState* S, const PeerComm& cm, cudaStream_t st) {
  k_update_p_pc<<<vec_grid(n), VT, 0, st>>>(p, r, dinv, n, S, cm);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_diag_inv(const CsrDev& A, bool dense, long long ld, long long diag_off,
                            double* dinv, int* bad, cudaStream_t st) {
  if (dense)
    k_diag_inv_dense<<<sm_count() * 4, 256, 0, st>>>(A.vals, ld, A.nrows, diag_off, dinv, bad);
  else
    k_diag_inv_csr<<<sm_count() * 4, 256, 0, st>>>(A.rowptr, A.cols, A.vals, A.nrows, diag_off,
                                                   dinv, bad);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_dot(const double* a, const double* b, long long n, double* partials,
                       unsigned int* ticket, double* out, cudaStream_t st) {
  k_dot<<<vec_grid(n), VT, 0, st>>>(a, b, n, partials, ticket, out);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_ewise(int op, double sca, const double* a, const double* b, double* out,
                         long long n, cudaStream_t st, double* out2) {
  const int g = vec_grid(n);
  if (op == 0)
    k_ewise<0><<<g, VT, 0, st>>>(sca, a, b, out, n, nullptr);
  else if (op == 1)
    k_ewise<1><<<g, VT, 0, st>>>(sca, a, b, out, n, nullptr);
  else if (op == 2)
    k_ewise<2><<<g, VT, 0, st>>>(sca, a, b, out, n, nullptr);
  else if (op == 3)
    k_ewise<3><<<g, VT, 0, st>>>(sca, a, b, out, n, nullptr);
  else if (op == 4)
    k_ewise<4><<<g, VT, 0, st>>>(sca, a, b, out, n, nullptr);
  else if (op == 5)
    k_ewise<5><<<g, VT, 0, st>>>(sca, a, b, out, n, out2);
  else
    k_ewise<6><<<g, VT, 0, st>>>(sca, a, b, out, n, out2);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_resid_norms(const double* b, const double* s, long long n, double* partials,
                               unsigned int* ticket, double* out2, cudaStream_t st) {
  int g = vec_grid(n);
  if (g > max_partials() / 2) g = max_partials() / 2;
  k_resid_norms<<<g, VT, 0, st>>>(b, s, n, partials, ticket, out2);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_gen_rhs(double* b, long long row0, long long n, unsigned long long seed,
                           cudaStream_t st) {
  k_gen_rhs<<<sm_count() * 4, 256, 0, st>>>(b, row0, n, seed);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_gen_lap(int nx, int ny, int nz, long long row0, int nloc, long long col_shift,
                           int* rowptr, int* cols, double* vals, cudaStream_t st) {
  k_gen_lap<<<sm_count() * 8, 256, 0, st>>>(nx, ny, nz, row0, nloc, col_shift, rowptr, cols, vals);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_gen_dense(int n, int row0, int nloc, unsigned long long seed, double* vals,
                             cudaStream_t st) {
  k_gen_dense<<<sm_count() * 8, 256, 0, st>>>(n, row0, nloc, seed, vals);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_analyze(const int* rowptr, const int* cols, int nrows, long long nnz, int ncols,
                           int* out3, cudaStream_t st) {
  static const int init[4] = {0x7fffffff, -0x7fffffff - 1, 0, 0};
  cudaError_t e = cudaMemcpyAsync(out3, init, sizeof(init), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  k_analyze<<<sm_count() * 4, 256, 0, st>>>(rowptr, cols, nrows, nnz, ncols, out3);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_upper_band(const int* rowptr, const int* cols, int nrows, int* out1,
                              cudaStream_t st) {
  static const int init = -0x7fffffff - 1;
  cudaError_t e = cudaMemcpyAsync(out1, &init, sizeof(int), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  k_upper_band<<<sm_count() * 4, 256, 0, st>>>(rowptr, cols, nrows, out1);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_shift_cols(int* cols, long long nnz, int shift, cudaStream_t st) {
  if (shift == 0) return cudaSuccess;
  k_shift_cols<<<sm_count() * 4, 256, 0, st>>>(cols, nnz, shift);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace cgb
