// kernels.h -- host-side launch interface of the sm_100a kernels (kernels.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

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
};

// `mode` of the matvec kernels
enum : int {
  SPMV_PLAIN = 0,         // s = A p only (mv_mult of the operator API)
  SPMV_FINISH = 1,        // + p.s, grid-reduced, finish_ps() on this GPU
  SPMV_FINISH_LOCAL = 2,  // + p.s, grid-reduced into S->local (a cross-GPU sum follows)
  SPMV_PARTIAL_ONLY = 3,  // + p.s, per-CTA partials only (first half of a split launch)
};
// a finishing launch adds `extra` partials left at partials[SPMV_EXTRA_OFFSET ..] by a preceding
// SPMV_PARTIAL_ONLY launch: mode = base | (extra << 8)
constexpr int SPMV_EXTRA_OFFSET = 4096;

enum : int { FIN_INIT = 0, FIN_PS = 1, FIN_RR = 2 };

int sm_count();
int max_partials();
int spmv_rows_per_tile();
int spmv_grid_for(int ntiles);
long long launch_count();
void add_launches(long long n);

cudaError_t launch_spmv(const CsrDev& A, const double* pg, const double* pown, double* s,
                        int tile_begin0, int tile_end0, int tile_begin1, int tile_end1,
                        double* partials, unsigned int* ticket, CgState* S, int mode,
                        cudaStream_t st);
cudaError_t launch_gemv(const double* A, long long ld, int nrows, int ncols, const double* pg,
                        const double* pown, double* s, double* partials, unsigned int* ticket,
                        CgState* S, int mode, cudaStream_t st);
cudaError_t launch_update_xr(double* x, double* r, const double* p, const double* s, long long n,
                             double* partials, unsigned int* ticket, CgState* S, int finish_local,
                             cudaStream_t st);
cudaError_t launch_update_p(double* p, const double* r, long long n, const CgState* S,
                            cudaStream_t st);
cudaError_t launch_init(double* x, double* r, double* p, const double* b, long long n,
                        double* partials, unsigned int* ticket, CgState* S, int finish_local,
                        cudaStream_t st);
cudaError_t launch_finish(CgState* S, int which, cudaStream_t st);
cudaError_t launch_dot(const double* a, const double* b, long long n, double* partials,
                       unsigned int* ticket, double* out, cudaStream_t st);
cudaError_t launch_ewise(int op, double sca, const double* a, const double* b, double* out,
                         long long n, cudaStream_t st);
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
cudaError_t launch_shift_cols(int* cols, long long nnz, int shift, cudaStream_t st);

}  // namespace cgb
