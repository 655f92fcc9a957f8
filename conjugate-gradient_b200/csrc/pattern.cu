// pattern.cu -- row-pattern dictionary format for CSR matrices with few distinct rows, and its SpMV.
//
// SURVEY 8(f) N4 ("alternative matrix formats behind the same API"): not in the reference, whose
// only layout is CSR (mv_ops.h:17-23).  A constant-coefficient stencil matrix has only a handful
// of distinct rows once a row is written as {(column - row, value)}: 27 for the 7-point Laplacian
// with Dirichlet boundaries, 9 for the 5-point one, 2h+1+2h for a Toeplitz band.  Such a matrix is
// stored as ONE BYTE per row (the row's pattern number) plus a dictionary that lives in shared
// memory: the matvec then streams 1 byte per row instead of 12 per entry (84 per row for the
// 7-point stencil), and the iteration's HBM traffic drops from 12 nnz + 92 n to 89 n bytes.
//
// The arithmetic is unchanged: a row is still summed entry by entry in ascending column order
// with unfused multiply and add (mv_ops.c:106-110,192-194), from the same values, so s = A p is
// bit-identical to the CSR kernel and to the reference.  Detection is exact: rows are hashed into
// a small table to find the candidate patterns, then every row is compared entry by entry (column
// offsets and value bits) with the dictionary entry it was assigned; any mismatch, more than 255
// patterns or a row longer than 32 entries leaves the system on the CSR path.
//
// Opt-in (cgb200_system_use_pattern_format): the contract figures of DESIGN.md section 4 are
// quoted on the CSR stream kernel; numbers measured in this format are reported separately.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "device_utils.cuh"
#include "kernels.h"
#include "solver.h"

namespace cgb {

constexpr int PAT_MAX_TYPES = 255;
constexpr int PAT_MAX_LEN = 32;
constexpr int PAT_TABLE = 2048;  // hash slots (power of two)
constexpr int PAT_PROBES = 64;

struct PatSlot {
  unsigned long long hash;  // 0 = empty
  int row;                  // a row that has this pattern
  int pad;
};

__device__ __forceinline__ unsigned long long pat_mix(unsigned long long h, unsigned long long v) {
  h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
  h *= 0xbf58476d1ce4e5b9ull;
  h ^= h >> 29;
  return h;
}

__device__ __forceinline__ unsigned long long pat_hash(const int* __restrict__ cols,
                                                       const double* __restrict__ vals, int a,
                                                       int b, int row) {
  unsigned long long h = pat_mix(0x243f6a8885a308d3ull, (unsigned long long)(b - a));
  for (int j = a; j < b; ++j) {
    h = pat_mix(h, (unsigned long long)(unsigned int)(cols[j] - row));
    h = pat_mix(h, (unsigned long long)__double_as_longlong(vals[j]));
  }
  return h ? h : 1ull;
}

// Pass 1: every row's pattern hash goes into the table; flags[0] = too many / too long,
// flags[1] = number of distinct hashes.
__global__ void k_pat_collect(const int* __restrict__ rowptr, const int* __restrict__ cols,
                              const double* __restrict__ vals, int nrows, PatSlot* table,
                              int* flags) {
  const int stride = gridDim.x * blockDim.x;
  for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += stride) {
    if (*(volatile int*)&flags[0]) return;
    const int a = rowptr[row], b = rowptr[row + 1];
    if (b - a > PAT_MAX_LEN) {
      flags[0] = 1;
      return;
    }
    const unsigned long long h = pat_hash(cols, vals, a, b, row);
    int slot = (int)(h & (PAT_TABLE - 1));
    bool placed = false;
    for (int probe = 0; probe < PAT_PROBES && !placed; ++probe) {
      unsigned long long cur = *(volatile unsigned long long*)&table[slot].hash;
      if (cur == 0ull) {
        cur = atomicCAS(&table[slot].hash, 0ull, h);
        if (cur == 0ull) {
          table[slot].row = row;
          if (atomicAdd(&flags[1], 1) >= PAT_MAX_TYPES) flags[0] = 1;
          placed = true;
          break;
        }
      }
      if (cur == h) placed = true;
      slot = (slot + 1) & (PAT_TABLE - 1);
    }
    if (!placed) flags[0] = 1;
  }
}

// Pass 2: assign each row its pattern number and verify the row against the dictionary exactly.
__global__ void k_pat_assign(const int* __restrict__ rowptr, const int* __restrict__ cols,
                             const double* __restrict__ vals, int nrows,
                             const PatSlot* __restrict__ table,
                             const unsigned char* __restrict__ slot_id,
                             const int* __restrict__ len, const int* __restrict__ rel,
                             const double* __restrict__ val, int lpad,
                             unsigned char* __restrict__ type, int* flags) {
  const int stride = gridDim.x * blockDim.x;
  for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += stride) {
    const int a = rowptr[row], b = rowptr[row + 1];
    const unsigned long long h = pat_hash(cols, vals, a, b, row);
    int slot = (int)(h & (PAT_TABLE - 1));
    int id = -1;
    for (int probe = 0; probe < PAT_PROBES; ++probe) {
      const unsigned long long cur = table[slot].hash;
      if (cur == h) {
        id = slot_id[slot];
        break;
      }
      if (cur == 0ull) break;
      slot = (slot + 1) & (PAT_TABLE - 1);
    }
    bool ok = id >= 0 && len[id] == b - a;
    for (int j = a; ok && j < b; ++j) {
      ok = rel[id * lpad + (j - a)] == cols[j] - row &&
           __double_as_longlong(val[id * lpad + (j - a)]) == __double_as_longlong(vals[j]);
    }
    if (!ok) {
      flags[0] = 1;
      return;
    }
    type[row] = (unsigned char)id;
  }
}

void free_pattern(PatDev* P) {
  cudaFree(P->type);
  cudaFree(P->len);
  cudaFree(P->rel);
  cudaFree(P->val);
  *P = PatDev();
}

#define PAT_OK(call)                                                                  \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      set_error(std::string("pattern format: ") + #call + ": " + cudaGetErrorString(e__)); \
      rc = -1;                                                                        \
      goto done;                                                                      \
    }                                                                                 \
  } while (0)

// 1 = built, 0 = the matrix does not compress (reason in *why), -1 = CUDA failure.
int build_pattern(const CsrDev& A, PatDev* out, cudaStream_t st, std::string* why) {
  int rc = 0;
  PatSlot* d_table = nullptr;
  int* d_flags = nullptr;
  unsigned char* d_slot_id = nullptr;
  PatDev P;
  std::vector<PatSlot> table(PAT_TABLE);
  std::vector<PatSlot> used;
  std::vector<unsigned char> slot_id(PAT_TABLE, 0);
  std::vector<int> len, rel;
  std::vector<double> val;
  int flags[2] = {0, 0};
  int lpad = 8;
  const int nrows = A.nrows;
  const int grid = sm_count() * 8;
  if (nrows <= 0 || !A.cols || !A.rowptr) {
    *why = "no CSR arrays";
    return 0;
  }
  PAT_OK(cudaMalloc(&d_table, sizeof(PatSlot) * PAT_TABLE));
  PAT_OK(cudaMalloc(&d_flags, 2 * sizeof(int)));
  PAT_OK(cudaMemsetAsync(d_table, 0, sizeof(PatSlot) * PAT_TABLE, st));
  PAT_OK(cudaMemsetAsync(d_flags, 0, 2 * sizeof(int), st));
  k_pat_collect<<<grid, 256, 0, st>>>(A.rowptr, A.cols, A.vals, nrows, d_table, d_flags);
  add_launches(1);
  PAT_OK(cudaGetLastError());
  PAT_OK(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
  PAT_OK(cudaMemcpyAsync(table.data(), d_table, sizeof(PatSlot) * PAT_TABLE, cudaMemcpyDeviceToHost, st));
  PAT_OK(cudaStreamSynchronize(st));
  if (flags[0]) {
    *why = "more than 255 distinct rows, or a row longer than 32 entries";
    goto done;
  }
  for (int i = 0; i < PAT_TABLE; ++i)
    if (table[i].hash) {
      used.push_back(table[i]);
      used.back().pad = i;  // remember the slot
    }
  std::sort(used.begin(), used.end(),
            [](const PatSlot& a, const PatSlot& b) { return a.hash < b.hash; });
  {
    // the dictionary: one representative row per pattern, read back from the device
    std::vector<std::vector<int>> rels(used.size());
    std::vector<std::vector<double>> vals(used.size());
    int maxlen = 0;
    for (size_t t = 0; t < used.size(); ++t) {
      int ab[2];
      PAT_OK(cudaMemcpy(ab, A.rowptr + used[t].row, 2 * sizeof(int), cudaMemcpyDeviceToHost));
      const int L = ab[1] - ab[0];
      rels[t].resize(L);
      vals[t].resize(L);
      if (L > 0) {
        PAT_OK(cudaMemcpy(rels[t].data(), A.cols + ab[0], L * sizeof(int), cudaMemcpyDeviceToHost));
        PAT_OK(cudaMemcpy(vals[t].data(), A.vals + ab[0], L * sizeof(double), cudaMemcpyDeviceToHost));
      }
      for (int j = 0; j < L; ++j) rels[t][j] -= used[t].row;
      maxlen = std::max(maxlen, L);
      slot_id[used[t].pad] = (unsigned char)t;
    }
    lpad = std::max(8, (maxlen + 7) / 8 * 8);
    const size_t T = used.size();
    len.assign(T, 0);
    rel.assign(T * lpad, 0);
    val.assign(T * lpad, 0.0);
    for (size_t t = 0; t < T; ++t) {
      len[t] = (int)rels[t].size();
      for (int j = 0; j < len[t]; ++j) {
        rel[t * lpad + j] = rels[t][j];
        val[t * lpad + j] = vals[t][j];
      }
    }
  }
  P.ntypes = (int)used.size();
  P.lpad = lpad;
  PAT_OK(cudaMalloc(&P.type, (size_t)nrows + 64));
  PAT_OK(cudaMalloc(&P.len, sizeof(int) * std::max(1, P.ntypes)));
  PAT_OK(cudaMalloc(&P.rel, sizeof(int) * std::max<size_t>(1, rel.size())));
  PAT_OK(cudaMalloc(&P.val, sizeof(double) * std::max<size_t>(1, val.size())));
  PAT_OK(cudaMalloc(&d_slot_id, PAT_TABLE));
  PAT_OK(cudaMemcpyAsync(P.len, len.data(), sizeof(int) * len.size(), cudaMemcpyHostToDevice, st));
  PAT_OK(cudaMemcpyAsync(P.rel, rel.data(), sizeof(int) * rel.size(), cudaMemcpyHostToDevice, st));
  PAT_OK(cudaMemcpyAsync(P.val, val.data(), sizeof(double) * val.size(), cudaMemcpyHostToDevice, st));
  PAT_OK(cudaMemcpyAsync(d_slot_id, slot_id.data(), PAT_TABLE, cudaMemcpyHostToDevice, st));
  PAT_OK(cudaMemsetAsync(d_flags, 0, 2 * sizeof(int), st));
  k_pat_assign<<<grid, 256, 0, st>>>(A.rowptr, A.cols, A.vals, nrows, d_table, d_slot_id, P.len,
                                     P.rel, P.val, lpad, P.type, d_flags);
  add_launches(1);
  PAT_OK(cudaGetLastError());
  PAT_OK(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
  PAT_OK(cudaStreamSynchronize(st));
  if (flags[0]) {  // two different rows shared a hash: stay on CSR rather than guess
    *why = "row verification against the dictionary failed (hash collision)";
    goto done;
  }
  *out = P;
  P = PatDev();
  rc = 1;
done:
  cudaFree(d_table);
  cudaFree(d_flags);
  cudaFree(d_slot_id);
  free_pattern(&P);
  return rc;
}

// ---------------------------------------------------------------------------------------------
// The matvec.  One thread per row, R rows per tile (the stream kernel's tile, so the interior /
// boundary classification of the multi-GPU layer applies unchanged), tiles dealt round-robin to a
// persistent grid so the chip sweeps the vector as one front.  A warp reads 32 consecutive
// pattern bytes and, entry by entry, 32 consecutive elements of p: every gather is a coalesced
// 256-byte access, mostly served by L1/L2 (HBM sees p once).
// ---------------------------------------------------------------------------------------------

template <bool COHERENT>
__device__ __forceinline__ double pattern_row(const double* __restrict__ val_s,
                                              const int* __restrict__ rel_s, const int L,
                                              const double* __restrict__ prow, const uint64_t pol) {
  double acc = 0.0;
  for (int k = 0; k < L; k += 8) {
    int c[8];
    double v[8], pv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) c[j] = rel_s[k + j];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      pv[j] = COHERENT ? ld_l2_if(prow + c[j], (k + j) < L) : ld_keep_if(prow + c[j], pol, (k + j) < L);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = val_s[k + j];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const double t = __dadd_rn(acc, __dmul_rn(v[j], pv[j]));
      acc = (k + j) < L ? t : acc;
    }
  }
  return acc;
}

__device__ __forceinline__ int pat_vtile_to_tile(const TileRange& tr, int v) {
  const int n0 = tr.e[0] - tr.b[0];
  if (v < n0) return tr.b[0] + v;
  v -= n0;
  const int n1 = tr.e[1] - tr.b[1];
  return v < n1 ? tr.b[1] + v : tr.b[2] + (v - n1);
}

template <int R>
__global__ void __launch_bounds__(R, 2)
k_spmv_pattern(const unsigned char* __restrict__ type, const int* __restrict__ len_g,
               const int* __restrict__ rel_g, const double* __restrict__ val_g, const int ntypes,
               const int lpad, const int nrows, const double* __restrict__ pg,
               const double* __restrict__ pown, double* __restrict__ s, const TileRange tr,
               const int wait_from, double* __restrict__ partials,
               unsigned int* __restrict__ ticket, CgState* S, const int mode, const int hint,
               const __grid_constant__ PeerComm cm) {
  const int fin = mode & MV_FIN_MASK;
  if (fin != FIN_NONE && S->done) return;
  extern __shared__ __align__(16) unsigned char pat_smem[];
  double* val_s = reinterpret_cast<double*>(pat_smem);
  int* rel_s = reinterpret_cast<int*>(val_s + (size_t)ntypes * lpad);
  int* len_s = rel_s + (size_t)ntypes * lpad;
  const int tid = threadIdx.x;
  for (int i = tid; i < ntypes * lpad; i += R) {
    val_s[i] = val_g[i];
    rel_s[i] = rel_g[i];
  }
  for (int i = tid; i < ntypes; i += R) len_s[i] = len_g[i];
  __syncthreads();

  const uint64_t pol_keep = hint ? policy_evict_last() : policy_evict_normal();
  const int nvt = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);
  double dot = 0.0, dot2 = 0.0;
  bool waited = false;
#pragma unroll 1
  for (int vt = blockIdx.x; vt < nvt; vt += gridDim.x) {
    const bool boundary = vt >= wait_from;
    if (boundary && !waited) {  // block-uniform: the peers' rows of the gather vector have landed
      if (tid == 0) wait_halo(cm, S);
      __syncthreads();
      waited = true;
    }
    const int row = pat_vtile_to_tile(tr, vt) * R + tid;
    if (row < nrows) {
      const int ty = type[row];
      const int L = len_s[ty];
      const double* prow = pg + row;
      const double acc = boundary
                             ? pattern_row<true>(val_s + ty * lpad, rel_s + ty * lpad, L, prow, pol_keep)
                             : pattern_row<false>(val_s + ty * lpad, rel_s + ty * lpad, L, prow, pol_keep);
      s[row] = acc;
      if (fin != FIN_NONE) {
        const double g = pown[row];
        dot += g * acc;
        if (mode & MV_DUAL) dot2 += g * g;
      }
    }
  }

  if (fin == FIN_NONE) return;
  const int nv = (mode & MV_DUAL) ? 2 : 1;
  if (mode & MV_PARTIAL) {  // first half of a split launch: leave the partials for the second
    __shared__ double smp[R / 32];
    const double t = block_sum<R>(dot, smp);
    if (tid == 0) partials[blockIdx.x] = t;
    if (nv == 2) {
      const double t2 = block_sum<R>(dot2, smp);
      if (tid == 0) partials[PART_STRIDE + blockIdx.x] = t2;
    }
    return;
  }
  double v[2] = {dot, dot2};
  bool last;
  if (nv == 2) {
    last = grid_reduce<R, 2>(v, partials, ticket, partials + MV_EXTRA_OFFSET, mode >> 8);
  } else {
    double v1[1] = {dot};
    last = grid_reduce<R, 1>(v1, partials, ticket, partials + MV_EXTRA_OFFSET, mode >> 8);
    v[0] = v1[0];
    v[1] = 0.0;
  }
  if (last) finish_block<R>(S, cm, fin, v[0], v[1]);
}

constexpr int PAT_R = 512;  // must equal the stream kernel's rows per tile (tile classification)

static size_t pat_smem_bytes(const PatDev& P) {
  return (size_t)P.ntypes * P.lpad * 12 + (size_t)P.ntypes * 4 + 16;
}

int pattern_grid_for(const PatDev& P, int ntiles) {
  static int bps = 0;
  static size_t bps_smem = ~(size_t)0;
  const size_t smem = pat_smem_bytes(P);
  if (!bps || bps_smem != smem) {
    cudaFuncSetAttribute(k_spmv_pattern<PAT_R>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)std::max<size_t>(smem, 1024));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_pattern<PAT_R>, PAT_R, smem);
    if (bps < 1) bps = 1;
    bps_smem = smem;
  }
  int grid = sm_count() * bps;
  if (grid > MAX_GRID) grid = MAX_GRID;
  return grid > ntiles ? ntiles : grid;
}

cudaError_t launch_spmv_pattern(const PatDev& P, int nrows, const double* pg, const double* pown,
                                double* s, const TileRange& tr, int wait_from, double* partials,
                                unsigned int* ticket, CgState* S, int mode, const PeerComm& cm,
                                cudaStream_t st) {
  const int nvt = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);
  if (nvt <= 0) return cudaSuccess;
  if (spmv_rows_per_tile() != PAT_R) return cudaErrorInvalidConfiguration;
  k_spmv_pattern<PAT_R><<<pattern_grid_for(P, nvt), PAT_R, pat_smem_bytes(P), st>>>(
      P.type, P.len, P.rel, P.val, P.ntypes, P.lpad, nrows, pg, pown, s, tr, wait_from, partials,
      ticket, S, mode, 1, cm);
  add_launches(1);
  return cudaGetLastError();
}

}  // namespace cgb
