// kernels.h -- host-side launch interface of the sm_100a kernels (kernels.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "comm.h"
#include "state.h"

namespace cgb {

// Device-resident CSR row block.  Arrays are padded so the bulk copies of the stream kernel may
// over-read: vals/cols by >= 16 entries, rowptr by >= rows_per_tile + 8 entries.
struct CsrDev {
  double* vals = nullptr;
  int* cols = nullptr;   // indices into the gather vector (already shifted to local space)
  int* rowptr = nullptr;
  int nrows = 0;
  long long nnz = 0;
  // banded matrices: gather index of the upper band edge = local row + lead (-1: no prefetch);
  // plen = length of the gather vector
  long long lead = -1, plen = 0;
};

// Row-pattern dictionary format (pattern.cu): one byte per row naming the row's pattern
// {(column - row, value)}; the dictionary (ntypes x lpad entries, lpad a multiple of 8) is copied
// into shared memory by the kernel.  Built only for matrices with at most 255 distinct rows.
constexpr int PAT_MAX_WIN = 4;
struct PatDev {
  unsigned char* type = nullptr;  // nrows
  int* len = nullptr;             // ntypes
  int* rel = nullptr;             // ntypes * lpad, column (gather space) minus local row
  double* val = nullptr;          // ntypes * lpad
  int ntypes = 0, lpad = 0, maxlen = 0;
  // staged form (k_spmv_pattern_tma): the column offsets of all patterns fall into at most
  // PAT_MAX_WIN clusters; for a block of consecutive rows each cluster is one contiguous window of
  // the gather vector, copied into shared memory by bulk async copies.  soff = position of an
  // entry's operand inside the staged windows, relative to the row's position in the block.
  bool staged = false;
  int* soff = nullptr;            // ntypes * lpad
  int nwin = 0;
  int wlo[PAT_MAX_WIN] = {0, 0, 0, 0};    // lowest column offset of the window (even)
  int wspan[PAT_MAX_WIN] = {0, 0, 0, 0};  // highest minus lowest offset
  int wbase[PAT_MAX_WIN] = {0, 0, 0, 0};  // first element of the window inside a stage
  int stage_elems = 0;            // doubles of gather vector per stage
};

// Diagonal (DIA) storage of a banded block (banded.cu): dia[d * ld + i] = A(i, i + lo + d), D diagonals.
struct DiaDev {
  double* dia = nullptr;
  int D = 0, lo = 0;
  long long ld = 0;
};

// `mode` of the matvec kernels: a finish kind (what the last CTA does with the fused dot
// products) in the low bits plus flags.
//   FIN_NONE   s = A p only (mv_mult of the operator API)
//   FIN_PS     + p.s   -> alpha = rr / ps                      (classic recurrence, cg.c:113)
//   FIN_SR     + r.w and r.r (MV_DUAL) -> alpha, beta          (single-reduction variant)
//   FIN_LOCAL  + the sums are left in S->local for the caller (NCCL mode, kernel timing)
// FIN_INIT / FIN_RR are the finish kinds of the vector kernels; FIN_BB / FIN_INIT2 / FIN_RZ those
// of the general loop entry (initial guess, Jacobi preconditioner: extensions, not in cg.c).
enum : int {
  FIN_NONE = 0, FIN_INIT = 1, FIN_PS = 2, FIN_RR = 3, FIN_SR = 4, FIN_LOCAL = 5,
  FIN_BB = 6, FIN_INIT2 = 7, FIN_RZ = 8
};
constexpr int MV_FIN_MASK = 0x0f;
constexpr int MV_PARTIAL = 0x10;  // publish per-CTA partials only (first half of a split launch)
constexpr int MV_DUAL = 0x20;     // also reduce the square of the gathered vector's owned block
constexpr int MV_REV = 0x40;      // visit each run of tiles from its last tile to its first (sweep direction)
// a finishing launch adds `extra` partials left at partials[MV_EXTRA_OFFSET ..] by a preceding
// MV_PARTIAL launch: mode |= extra << 8
constexpr int MV_EXTRA_OFFSET = 4096;
constexpr int NO_WAIT = 0x7fffffff;

// Up to three runs of tiles [b[i], e[i]), visited in this order by the stream SpMV.
struct TileRange {
  int b[3], e[3];
};

int sm_count();
int max_partials();
int partials_doubles();
size_t partials_zero_offset();  // the tail of a partials buffer (group counters) must start zeroed
int spmv_rows_per_tile();
int spmv_grid_for(int ntiles);
long long launch_count();
void add_launches(long long n);

cudaError_t launch_spmv(const CsrDev& A, const double* pg, const double* pown, double* s,
                        const TileRange& tr, int wait_from, double* partials,
                        unsigned int* ticket, CgState* S, int mode, const PeerComm& cm,
                        cudaStream_t st);
int pattern_grid_for(const PatDev& P, const TileRange& tr);
cudaError_t launch_spmv_pattern(const PatDev& P, int nrows, const double* pg, const double* pown,
                                double* s, const TileRange& tr, int wait_from, long long plen,
                                double* partials, unsigned int* ticket, CgState* S, int mode,
                                const PeerComm& cm, cudaStream_t st);
void free_pattern(PatDev* P);
void free_banded(DiaDev* B);
cudaError_t launch_spmv_dia(const DiaDev& B, int nrows, const double* pg, long long plen, long long poff,
                            const double* pown, double* s, double* partials, unsigned int* ticket,
                            CgState* S, int mode, const PeerComm& cm, cudaStream_t st);
cudaError_t launch_gemv(const double* A, long long ld, int nrows, int ncols, const double* pg,
                        const double* pown, double* s, double* partials, unsigned int* ticket,
                        CgState* S, int mode, int wait, const PeerComm& cm, cudaStream_t st);
// dense path: the whole classic solve in one cooperative launch (kernels.cu: k_cg_dense_solve)
void set_dense_iter_enabled(int on);
bool dense_iter_enabled();
bool dense_iter_supported(const double* A, long long ld, int ncols, const double* pg);
// One cooperative launch runs the loop entry (init != 0) and up to `iters` iterations; it ends by
// itself when the loop stops.  max_iter / tol are read at the loop entry only.
cudaError_t launch_cg_dense_solve(const double* A, long long ld, int nrows, int ncols,
                                  const double* pg, double* pown, double* x, double* r, double* s,
                                  const double* b, double* partials, CgState* S, int iters,
                                  int init, int max_iter, double tol, int wait, int timing,
                                  unsigned long long* trace, double* pool, const PeerComm& cm,
                                  cudaStream_t st);
int dense_pool_groups(int nrows, int ncols);       // row groups dealt dynamically (0 = none)
size_t dense_pool_doubles(int nrows, int ncols);   // size of the zero-initialised pool buffer
// rev != 0: the vector kernels sweep from the last element to the first (see solver.cu: sweep_dir)
cudaError_t launch_update_xr(double* x, double* r, const double* p, const double* s, long long n,
                             double* partials, unsigned int* ticket, CgState* S, int fin, int rev,
                             const PeerComm& cm, cudaStream_t st);
// K2 and K3 in one co-operative launch for small single-GPU systems (kernels.cu: k_update_xrp)
void set_fused_vector_update(int on);  // A/B switch (CGB200_FUSE_VEC), for systems solved afterwards
bool update_xrp_applies(const double* x, const double* r, const double* p, const double* s, long long n);
cudaError_t launch_update_xrp(double* x, double* r, double* p, const double* s, long long n,
                              double* partials, unsigned int* ticket, CgState* S, int rev,
                              const PeerComm& cm, cudaStream_t st);
cudaError_t launch_update_p(double* p, const double* r, long long n, const CgState* S, int rev,
                            const PeerComm& cm, cudaStream_t st);
cudaError_t launch_update_sr(double* x, double* g, double* d, double* s, const double* w,
                             long long n, const CgState* S, int rev, const PeerComm& cm,
                             cudaStream_t st);
cudaError_t launch_init(double* x, double* r, double* p, double* s, const double* b, long long n,
                        int sr, double* partials, unsigned int* ticket, CgState* S, int fin,
                        const PeerComm& cm, cudaStream_t st);
// general loop entry (initial guess) and the Jacobi-preconditioned vector kernels (extensions)
cudaError_t launch_init_bb(double* x, const double* b, long long n, int zero_x, double* partials,
                           unsigned int* ticket, CgState* S, int fin, const PeerComm& cm,
                           cudaStream_t st);
cudaError_t launch_init_r(double* r, double* p, const double* b, const double* s,
                          const double* dinv, long long n, double* partials, unsigned int* ticket,
                          CgState* S, int fin, const PeerComm& cm, cudaStream_t st);
cudaError_t launch_update_xr_pc(double* x, double* r, const double* p, const double* s,
                                const double* dinv, long long n, double* partials,
                                unsigned int* ticket, CgState* S, int fin, const PeerComm& cm,
                                cudaStream_t st);
cudaError_t launch_update_p_pc(double* p, const double* r, const double* dinv, long long n,
                               const CgState* S, const PeerComm& cm, cudaStream_t st);
// dinv = 1 / diag of the owned rows; diag_off = gather index (CSR) or global column (dense) of
// row 0's diagonal; *bad = rows without a positive finite diagonal
cudaError_t launch_diag_inv(const CsrDev& A, bool dense, long long ld, long long diag_off,
                            double* dinv, int* bad, cudaStream_t st);
// CSR matvec with a group of lanes per row, for blocks of long rows on one GPU (bit-identical to
// launch_spmv); spmv_rowgroup_min() is the mean row length from which solver.cu chooses it.
cudaError_t launch_spmv_rowgroup(const CsrDev& A, const double* pg, const double* pown, double* s,
                                 double* partials, unsigned int* ticket, CgState* S, int mode,
                                 const PeerComm& cm, cudaStream_t st);
int spmv_rowgroup_min();
// Several GPUs in peer-memory mode: the default long-row kernel, every CTA first waiting for the
// peers' halo flags, gathers bypassing L1 (the other forms are single-GPU / NCCL-mode only).
cudaError_t launch_spmv_longrow_coherent(const CsrDev& A, const double* pg, const double* pown,
                                         double* s, double* partials, unsigned int* ticket,
                                         CgState* S, int mode, const PeerComm& cm, cudaStream_t st);
cudaError_t launch_push(const double* p, const PeerComm& cm, cudaStream_t st);
cudaError_t launch_finish(CgState* S, int which, cudaStream_t st);
cudaError_t launch_dot(const double* a, const double* b, long long n, double* partials,
                       unsigned int* ticket, double* out, cudaStream_t st);
cudaError_t launch_ewise(int op, double sca, const double* a, const double* b, double* out,
                         long long n, cudaStream_t st, double* out2 = nullptr);
cudaError_t launch_resid_norms(const double* b, const double* s, long long n, double* partials,
                               unsigned int* ticket, double* out2, cudaStream_t st);
cudaError_t launch_gen_rhs(double* b, long long row0, long long n, unsigned long long seed,
                           cudaStream_t st);
cudaError_t launch_gen_lap(int nx, int ny, int nz, long long row0, int nloc, long long col_shift,
                           int* rowptr, int* cols, double* vals, cudaStream_t st);
cudaError_t launch_gen_dense(int n, int row0, int nloc, unsigned long long seed, double* vals,
                             cudaStream_t st);
cudaError_t launch_analyze(const int* rowptr, const int* cols, int nrows, long long nnz, int ncols,
                           int* out3, cudaStream_t st);
cudaError_t launch_upper_band(const int* rowptr, const int* cols, int nrows, int* out1,
                              cudaStream_t st);
cudaError_t launch_shift_cols(int* cols, long long nnz, int shift, cudaStream_t st);

}  // namespace cgb
