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
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "devmem.h"
#include "device_utils.cuh"
#include "kernels.h"
#include "solver.h"

namespace cgb {

constexpr int PAT_MAX_TYPES = 255;
constexpr int PAT_MAX_LEN = 32;
constexpr int PAT_TABLE = 2048;  // hash slots (power of two)
constexpr int PAT_PROBES = 64;

constexpr int PT_R = 512;   // rows per tile = consumer threads (the stream kernel's tile)
constexpr int PT_PB = 4;    // consecutive tiles staged together
constexpr long long PT_MAX_STAGE_BYTES = 96 * 1024;  // gather-vector windows of one stage

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
  dev_free(P->type);
  dev_free(P->len);
  dev_free(P->rel);
  dev_free(P->val);
  dev_free(P->soff);
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
  PAT_OK(DEV_ALLOC(&d_table, sizeof(PatSlot) * PAT_TABLE));
  PAT_OK(DEV_ALLOC(&d_flags, 2 * sizeof(int)));
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
    P.maxlen = maxlen;
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
  PAT_OK(DEV_ALLOC(&P.type, (size_t)nrows + 64));
  PAT_OK(DEV_ALLOC(&P.len, sizeof(int) * std::max(1, P.ntypes)));
  PAT_OK(DEV_ALLOC(&P.rel, sizeof(int) * std::max<size_t>(1, rel.size())));
  PAT_OK(DEV_ALLOC(&P.val, sizeof(double) * std::max<size_t>(1, val.size())));
  PAT_OK(DEV_ALLOC(&d_slot_id, PAT_TABLE));
  {
    // staged form: cluster the column offsets of all patterns into windows
    std::vector<int> offs;
    for (int t = 0; t < P.ntypes; ++t)
      for (int j = 0; j < len[t]; ++j) offs.push_back(rel[(size_t)t * lpad + j]);
    std::sort(offs.begin(), offs.end());
    offs.erase(std::unique(offs.begin(), offs.end()), offs.end());
    const int BS = PT_PB * PT_R;
    int nwin = 0, lo[PAT_MAX_WIN], hi[PAT_MAX_WIN];
    bool fits = !offs.empty();
    for (size_t i = 0; i < offs.size() && fits; ++i) {
      if (nwin > 0 && (long long)offs[i] - hi[nwin - 1] <= BS / 2) {
        hi[nwin - 1] = offs[i];
      } else if (nwin < PAT_MAX_WIN) {
        lo[nwin] = offs[i] & ~1;  // even (also for negative offsets): 16-byte aligned copies
        hi[nwin] = offs[i];
        ++nwin;
      } else {
        fits = false;
      }
    }
    long long elems = 0;
    for (int w = 0; w < nwin && fits; ++w) {
      P.wlo[w] = lo[w];
      P.wspan[w] = hi[w] - lo[w];
      P.wbase[w] = (int)elems;
      elems += ((long long)BS + P.wspan[w] + 1 + 1) & ~1ll;
      if (elems * 8 > PT_MAX_STAGE_BYTES) fits = false;
    }
    if (fits) {
      std::vector<int> soff((size_t)std::max(1, P.ntypes) * lpad, 0);
      for (int t = 0; t < P.ntypes; ++t)
        for (int j = 0; j < len[t]; ++j) {
          const int d = rel[(size_t)t * lpad + j];
          int w = 0;
          while (w + 1 < nwin && d >= lo[w + 1]) ++w;
          soff[(size_t)t * lpad + j] = P.wbase[w] + d - lo[w];
        }
      PAT_OK(DEV_ALLOC(&P.soff, sizeof(int) * soff.size()));
      PAT_OK(cudaMemcpy(P.soff, soff.data(), sizeof(int) * soff.size(), cudaMemcpyHostToDevice));
      P.nwin = nwin;
      P.stage_elems = (int)elems;
      P.staged = true;
    }
  }
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
  dev_free(d_table);
  dev_free(d_flags);
  dev_free(d_slot_id);
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

__device__ __forceinline__ int pat_vtile_to_tile(const TileRange& tr, int v, const bool rev) {
  const int n0 = tr.e[0] - tr.b[0];
  if (v < n0) return rev ? tr.e[0] - 1 - v : tr.b[0] + v;
  v -= n0;
  const int n1 = tr.e[1] - tr.b[1];
  if (v < n1) return rev ? tr.e[1] - 1 - v : tr.b[1] + v;
  v -= n1;
  return rev ? tr.e[2] - 1 - v : tr.b[2] + v;
}

// Two rows of at most 8 entries each, all 16 gathers issued before the first product is formed.
template <bool COHERENT>
__device__ __forceinline__ void pattern_rows2(const double* __restrict__ val_s,
                                              const int* __restrict__ rel_s, const int offA,
                                              const int offB, const int LA, const int LB,
                                              const double* __restrict__ prowA,
                                              const double* __restrict__ prowB, const uint64_t pol,
                                              double& accA, double& accB) {
  int cA[8], cB[8];
  double pvA[8], pvB[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    cA[j] = rel_s[offA + j];
    cB[j] = rel_s[offB + j];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    pvA[j] = COHERENT ? ld_l2_if(prowA + cA[j], j < LA) : ld_keep_if(prowA + cA[j], pol, j < LA);
    pvB[j] = COHERENT ? ld_l2_if(prowB + cB[j], j < LB) : ld_keep_if(prowB + cB[j], pol, j < LB);
  }
  double a = 0.0, b = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const double ta = __dadd_rn(a, __dmul_rn(val_s[offA + j], pvA[j]));
    const double tb = __dadd_rn(b, __dmul_rn(val_s[offB + j], pvB[j]));
    a = j < LA ? ta : a;
    b = j < LB ? tb : b;
  }
  accA = a;
  accB = b;
}

// Tiles are handed out in blocks of PAT_B consecutive tiles: rows i +- (a few rows) and, for grid
// stencils, i +- (one grid line) of a tile are then rows of the same CTA's previous / next tile
// and hit in L1; only the far neighbours (i +- one plane) come from L2.  The pattern bytes of a
// whole block are fetched before its first gather, and two tiles are in flight per thread.
constexpr int PAT_B = 8;

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
  const bool rev = (mode & MV_REV) != 0;
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
  const bool dots = fin != FIN_NONE;
  double dot = 0.0, dot2 = 0.0;
  bool waited = false;
#pragma unroll 1
  for (int v0 = blockIdx.x * PAT_B; v0 < nvt; v0 += gridDim.x * PAT_B) {
    const int nb = min(PAT_B, nvt - v0);
    const bool boundary = v0 + nb - 1 >= wait_from;
    if (boundary && !waited) {  // block-uniform: the peers' rows of the gather vector have landed
      if (tid == 0) wait_halo(cm, S);
      __syncthreads();
      waited = true;
    }
    int row[PAT_B], ty[PAT_B];
#pragma unroll
    for (int j = 0; j < PAT_B; ++j) {
      row[j] = j < nb ? pat_vtile_to_tile(tr, v0 + j, rev) * R + tid : nrows;
      ty[j] = row[j] < nrows ? (int)type[row[j]] : -1;
    }
    if (lpad == 8) {
#pragma unroll
      for (int j = 0; j < PAT_B; j += 2) {
        const int LA = ty[j] >= 0 ? len_s[ty[j]] : 0, LB = ty[j + 1] >= 0 ? len_s[ty[j + 1]] : 0;
        const int offA = max(ty[j], 0) * 8, offB = max(ty[j + 1], 0) * 8;
        const double* pa = pg + min(row[j], nrows - 1);
        const double* pb = pg + min(row[j + 1], nrows - 1);
        double accA, accB;
        if (boundary)
          pattern_rows2<true>(val_s, rel_s, offA, offB, LA, LB, pa, pb, pol_keep, accA, accB);
        else
          pattern_rows2<false>(val_s, rel_s, offA, offB, LA, LB, pa, pb, pol_keep, accA, accB);
        if (ty[j] >= 0) {
          s[row[j]] = accA;
          if (dots) {
            const double g = pown[row[j]];
            dot += g * accA;
            dot2 += g * g;
          }
        }
        if (ty[j + 1] >= 0) {
          s[row[j + 1]] = accB;
          if (dots) {
            const double g = pown[row[j + 1]];
            dot += g * accB;
            dot2 += g * g;
          }
        }
      }
    } else {
#pragma unroll 1
      for (int j = 0; j < PAT_B; ++j) {
        int tyj = -1, rowj = 0;
#pragma unroll
        for (int q = 0; q < PAT_B; ++q)  // (register arrays: select, do not index)
          if (q == j) {
            tyj = ty[q];
            rowj = row[q];
          }
        if (tyj < 0) continue;
        const int L = len_s[tyj];
        const double* prow = pg + rowj;
        const double acc =
            boundary ? pattern_row<true>(val_s + tyj * lpad, rel_s + tyj * lpad, L, prow, pol_keep)
                     : pattern_row<false>(val_s + tyj * lpad, rel_s + tyj * lpad, L, prow, pol_keep);
        s[rowj] = acc;
        if (dots) {
          const double g = pown[rowj];
          dot += g * acc;
          dot2 += g * g;
        }
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

// ---------------------------------------------------------------------------------------------
// Staged form.  The register-gather kernel above is bound by how many loads an SM can keep in
// flight (ncu: long-scoreboard stalls, 25 % of DRAM bandwidth).  Here a producer thread copies,
// for a block of PT_PB consecutive tiles, the few contiguous windows of the gather vector that
// the block's rows can touch (one per cluster of column offsets) and the block's pattern bytes
// into shared memory with bulk async copies (TMA) on an mbarrier ring; the consumers read their
// operands from shared memory.  Outstanding bytes cost no registers and no L1 miss slots.
// ---------------------------------------------------------------------------------------------

struct PatWin {
  int nwin;
  int lo[PAT_MAX_WIN], span[PAT_MAX_WIN], base[PAT_MAX_WIN];
  int stage_elems;
};

// Blocks never straddle two runs of the tile range, so a block's rows are consecutive.
__device__ __forceinline__ bool pt_decode(const TileRange& tr, int g, const bool rev, int& tile0,
                                          int& nb, int& vfirst) {
  int pre = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int lenk = tr.e[k] - tr.b[k];
    const int nblk = (lenk + PT_PB - 1) / PT_PB;
    if (g < nblk) {
      if (rev) {  // the run's blocks from its end: block g ends PT_PB * g tiles before the run's end
        const int hi = tr.e[k] - g * PT_PB;
        tile0 = max(tr.b[k], hi - PT_PB);
        nb = hi - tile0;
      } else {
        tile0 = tr.b[k] + g * PT_PB;
        nb = min(PT_PB, tr.e[k] - tile0);
      }
      vfirst = pre + g * PT_PB;
      return true;
    }
    g -= nblk;
    pre += lenk;
  }
  return false;
}

static int pt_num_blocks(const TileRange& tr) {
  int n = 0;
  for (int k = 0; k < 3; ++k) n += (tr.e[k] - tr.b[k] + PT_PB - 1) / PT_PB;
  return n;
}

// NE = operands read per row when every pattern has at most 8 entries (3, 5, 7 or 8: the longest
// pattern rounded up) -- a padding entry then costs no shared-memory read.
template <int R, int NE>
__global__ void __launch_bounds__(R + 32, 1)
k_spmv_pattern_tma(const unsigned char* __restrict__ type, const int* __restrict__ len_g,
                   const int* __restrict__ soff_g, const double* __restrict__ val_g,
                   const int ntypes, const int lpad, const PatWin win, const int stages,
                   const int nrows, const long long plen, const double* __restrict__ pg,
                   const double* __restrict__ pown, double* __restrict__ s, const TileRange tr,
                   const int wait_from, double* __restrict__ partials,
                   unsigned int* __restrict__ ticket, CgState* S, const int mode,
                   const __grid_constant__ PeerComm cm) {
  constexpr int THREADS = R + 32;
  constexpr int CWARPS = R / 32;
  constexpr int TYB = PT_PB * R;  // pattern bytes per stage
  const int fin = mode & MV_FIN_MASK;
  const bool rev = (mode & MV_REV) != 0;
  if (fin != FIN_NONE && S->done) return;

  extern __shared__ __align__(128) unsigned char pt_smem[];
  double* p_area = reinterpret_cast<double*>(pt_smem);
  unsigned char* ty_area = pt_smem + (size_t)stages * win.stage_elems * 8;
  double* val_s = reinterpret_cast<double*>(ty_area + (size_t)stages * TYB);
  int* soff_s = reinterpret_cast<int*>(val_s + (size_t)ntypes * lpad);
  int* len_s = soff_s + (size_t)ntypes * lpad;
  uint64_t* full = reinterpret_cast<uint64_t*>(
      (reinterpret_cast<uintptr_t>(len_s + ntypes) + 7) & ~(uintptr_t)7);
  uint64_t* empty = full + 4;

  const int tid = threadIdx.x;
  for (int i = tid; i < ntypes * lpad; i += THREADS) {
    val_s[i] = val_g[i];
    soff_s[i] = soff_g[i];
  }
  for (int i = tid; i < ntypes; i += THREADS) len_s[i] = len_g[i];
  if (tid == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], CWARPS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  double dot = 0.0, dot2 = 0.0;
  if (tid >= R) {
    // ------------------------------ producer ------------------------------
    if (tid == R) {
      const uint64_t pol_p = policy_evict_last();      // p is gathered again by the neighbouring blocks
      const uint64_t pol_ty = policy_evict_first();    // pattern bytes stream once
      const long long plen2 = (plen + 1) & ~1ll;       // the gather vector is padded past plen
      bool waited = false;
      int item = 0;
      int tile0 = 0, nb = 0, vfirst = 0;
#pragma unroll 1
      for (int g = blockIdx.x; pt_decode(tr, g, rev, tile0, nb, vfirst); g += gridDim.x, ++item) {
        if (vfirst + nb - 1 >= wait_from && !waited) {
          // first block with boundary tiles: the peers' rows of the gather vector must have landed,
          // and the bulk copies (async proxy) must be ordered behind that acquire
          wait_halo(cm, S);
          asm volatile("fence.proxy.async;" ::: "memory");
          waited = true;
        }
        const int stage = item % stages;
        if (item >= stages) mbar_wait(&empty[stage], ((uint32_t)(item / stages) & 1u) ^ 1u);
        const long long r0 = (long long)tile0 * R;
        const int rows_in = (int)min((long long)nb * R, (long long)nrows - r0);
        const uint32_t ty_bytes = (uint32_t)((rows_in + 15) & ~15);
        long long cs[PAT_MAX_WIN], ce[PAT_MAX_WIN];
        uint32_t total = ty_bytes;
#pragma unroll
        for (int w = 0; w < PAT_MAX_WIN; ++w) {
          cs[w] = ce[w] = 0;
          if (w < win.nwin) {
            const long long gs = r0 + win.lo[w];
            const long long ge = (r0 + (long long)nb * R + win.lo[w] + win.span[w] + 1 + 1) & ~1ll;
            cs[w] = max(gs, 0ll);
            ce[w] = min(ge, plen2);
            if (ce[w] > cs[w]) total += (uint32_t)(ce[w] - cs[w]) * 8u;
          }
        }
        mbar_expect_tx(&full[stage], total);
        bulk_g2s(ty_area + (size_t)stage * TYB, type + r0, ty_bytes, &full[stage], pol_ty);
        double* pst = p_area + (size_t)stage * win.stage_elems;
#pragma unroll
        for (int w = 0; w < PAT_MAX_WIN; ++w)
          if (w < win.nwin && ce[w] > cs[w])
            bulk_g2s(pst + win.base[w] + (cs[w] - (r0 + win.lo[w])), pg + cs[w],
                     (uint32_t)(ce[w] - cs[w]) * 8u, &full[stage], pol_p);
      }
    }
  } else {
    // ------------------------------ consumers ------------------------------
    const bool dots = fin != FIN_NONE;
    const int lane = tid & 31;
    int item = 0;
    int tile0 = 0, nb = 0, vfirst = 0;
#pragma unroll 1
    for (int g = blockIdx.x; pt_decode(tr, g, rev, tile0, nb, vfirst); g += gridDim.x, ++item) {
      const int stage = item % stages;
      mbar_wait(&full[stage], (uint32_t)(item / stages) & 1u);
      const double* pw = p_area + (size_t)stage * win.stage_elems;
      const unsigned char* ty_s = ty_area + (size_t)stage * TYB;
      const long long r0 = (long long)tile0 * R;
      if (lpad == 8) {
        double acc[PT_PB];
        bool valid[PT_PB];
#pragma unroll
        for (int j = 0; j < PT_PB; ++j) {
          const int q = j * R + tid;
          valid[j] = j < nb && r0 + q < nrows;
          const int ty = valid[j] ? (int)ty_s[q] : 0;
          const int L = valid[j] ? len_s[ty] : 0;
          const int base = ty * 8;
          // the shared-memory pipe is what bounds this kernel (ncu: l1tex 83 %): dictionary entries
          // come as 128-bit broadcast reads
          const int4 so0 = *reinterpret_cast<const int4*>(soff_s + base);
          const int4 so1 = *reinterpret_cast<const int4*>(soff_s + base + 4);
          const int so[8] = {so0.x, so0.y, so0.z, so0.w, so1.x, so1.y, so1.z, so1.w};
          double pv[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) pv[e] = e < NE ? pw[so[e] + q] : 0.0;
          double vv[8];
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const double2 t2 = *reinterpret_cast<const double2*>(val_s + base + e);
            vv[e] = t2.x;
            vv[e + 1] = t2.y;
          }
          double a = 0.0;
#pragma unroll
          for (int e = 0; e < NE; ++e) {
            const double t = __dadd_rn(a, __dmul_rn(vv[e], pv[e]));
            a = e < L ? t : a;
          }
          acc[j] = a;
        }
#pragma unroll
        for (int j = 0; j < PT_PB; ++j)
          if (valid[j]) {
            const long long row = r0 + j * R + tid;
            s[row] = acc[j];
            if (dots) {
              const double gv = pown[row];
              dot += gv * acc[j];
              dot2 += gv * gv;
            }
          }
      } else {
#pragma unroll 1
        for (int j = 0; j < nb; ++j) {
          const int q = j * R + tid;
          const long long row = r0 + q;
          if (row >= nrows) break;
          const int ty = ty_s[q];
          const int L = len_s[ty];
          const int base = ty * lpad;
          double a = 0.0;
          for (int e = 0; e < L; ++e)
            a = __dadd_rn(a, __dmul_rn(val_s[base + e], pw[soff_s[base + e] + q]));
          s[row] = a;
          if (dots) {
            const double gv = pown[row];
            dot += gv * a;
            dot2 += gv * gv;
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);  // this warp is done with the stage
    }
  }

  if (fin == FIN_NONE) return;
  const int nv = (mode & MV_DUAL) ? 2 : 1;
  if (mode & MV_PARTIAL) {  // first half of a split launch: leave the partials for the second
    __shared__ double smp[THREADS / 32];
    const double t = block_sum<THREADS>(dot, smp);
    if (tid == 0) partials[blockIdx.x] = t;
    if (nv == 2) {
      const double t2 = block_sum<THREADS>(dot2, smp);
      if (tid == 0) partials[PART_STRIDE + blockIdx.x] = t2;
    }
    return;
  }
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

static size_t pat_smem_bytes(const PatDev& P) {
  return (size_t)P.ntypes * P.lpad * 12 + (size_t)P.ntypes * 4 + 16;
}

// Stages of the staged kernel that fit one CTA per SM (0 = use the register-gather kernel).
static int pt_stages(const PatDev& P, size_t* smem_out) {
  static int off = -1;
  if (off < 0) {
    const char* e = getenv("CGB200_PATTERN_TMA");  // 0 = register-gather kernel (comparison runs)
    off = (e && atoi(e) == 0) ? 1 : 0;
  }
  if (!P.staged || off) return 0;
  const size_t fixed = (size_t)P.ntypes * P.lpad * 12 + (size_t)P.ntypes * 4 + 8 + 64 + 128;
  const size_t per_stage = (size_t)P.stage_elems * 8 + (size_t)PT_PB * PT_R;
  const size_t budget = 200 * 1024;
  if (fixed + 2 * per_stage > budget) return 0;
  int st = (int)((budget - fixed) / per_stage);
  if (st > 4) st = 4;
  if (smem_out) *smem_out = fixed + (size_t)st * per_stage;
  return st;
}

int pattern_grid_for(const PatDev& P, const TileRange& tr) {
  const int ntiles = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);
  size_t smem = 0;
  if (pt_stages(P, &smem) > 0) {
    static size_t set_for = 0;
    if (set_for < smem) {
      cudaFuncSetAttribute(k_spmv_pattern_tma<PT_R, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(k_spmv_pattern_tma<PT_R, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(k_spmv_pattern_tma<PT_R, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(k_spmv_pattern_tma<PT_R, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      set_for = smem;
    }
    int grid = sm_count();
    const int nblk = pt_num_blocks(tr);
    return grid > nblk ? nblk : grid;
  }
  static int bps = 0;
  static size_t bps_smem = ~(size_t)0;
  smem = pat_smem_bytes(P);
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
                                double* s, const TileRange& tr, int wait_from, long long plen,
                                double* partials, unsigned int* ticket, CgState* S, int mode,
                                const PeerComm& cm, cudaStream_t st) {
  const int nvt = (tr.e[0] - tr.b[0]) + (tr.e[1] - tr.b[1]) + (tr.e[2] - tr.b[2]);
  if (nvt <= 0) return cudaSuccess;
  if (spmv_rows_per_tile() != PAT_R) return cudaErrorInvalidConfiguration;
  const int grid = pattern_grid_for(P, tr);
  size_t smem = 0;
  const int stages = pt_stages(P, &smem);
  if (stages > 0 && ((uintptr_t)pg % 16 == 0)) {
    PatWin w;
    w.nwin = P.nwin;
    for (int i = 0; i < PAT_MAX_WIN; ++i) {
      w.lo[i] = P.wlo[i];
      w.span[i] = P.wspan[i];
      w.base[i] = P.wbase[i];
    }
    w.stage_elems = P.stage_elems;
#define PT_LAUNCH(NE)                                                                          \
  k_spmv_pattern_tma<PT_R, NE><<<grid, PT_R + 32, smem, st>>>(                                 \
      P.type, P.len, P.soff, P.val, P.ntypes, P.lpad, w, stages, nrows, plen, pg, pown, s, tr, \
      wait_from, partials, ticket, S, mode, cm)
    if (P.maxlen <= 3) PT_LAUNCH(3);
    else if (P.maxlen <= 5) PT_LAUNCH(5);
    else if (P.maxlen <= 7) PT_LAUNCH(7);
    else PT_LAUNCH(8);
#undef PT_LAUNCH
  } else if (stages > 0) {
    return cudaErrorMisalignedAddress;  // the gather vector is always 16-byte aligned in this library
  } else {
    k_spmv_pattern<PAT_R><<<grid, PAT_R, pat_smem_bytes(P), st>>>(
        P.type, P.len, P.rel, P.val, P.ntypes, P.lpad, nrows, pg, pown, s, tr, wait_from, partials,
        ticket, S, mode, 1, cm);
  }
  add_launches(1);
  return cudaGetLastError();
}

}  // namespace cgb
