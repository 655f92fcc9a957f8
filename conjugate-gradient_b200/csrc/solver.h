// solver.h -- host-side objects of libcgb200: the process context, device-resident systems and
// the solve entry points used by the C ABI (abi.cu).
#pragma once

#include <cuda_runtime.h>
#include <math.h>

#include <map>
#include <string>
#include <vector>

#include "comm.h"
#include "kernels.h"
#include "partition.h"
#include "state.h"

namespace cgb {

struct Ctx {
  bool ready = false;
  int device = 0;
  cudaStream_t stream = nullptr;       // every kernel of the path runs here
  cudaStream_t comm_stream = nullptr;  // halo exchange, overlapped with interior tiles
  double* partials = nullptr;          // grid-reduction partials for the standalone operators
  CgState* d_state = nullptr;          // scratch state for the standalone operators
  CgState* h_state = nullptr;          // pinned: [0],[1] poll slots, [2] upload, [3] final
  void* d_scratch = nullptr;           // 64 B
  void* h_scratch = nullptr;           // 64 B pinned
  cudaEvent_t poll_ev[2] = {nullptr, nullptr};
  cudaEvent_t t0 = nullptr, t1 = nullptr, t2 = nullptr, t3 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // multi-GPU
  int rank = 0, world = 1;
  void* comm = nullptr;  // ncclComm_t
  // how the GPUs talk inside the loop: 1 = peer memory over NVLink (kernels store into each
  // other's buffers, comm.h), 0 = NCCL calls between the kernels.  CGB200_COMM=p2p|nccl.
  int comm_mode = 1;
  bool team_primary = false;  // this process is rank 0 of a GPU team it started itself (team.cu)
};

Ctx& ctx();
int ensure_init(int device);
void shutdown();
void set_error(const std::string& m);
const char* last_error();

int dist_unique_id(void* id128);
int dist_init(int rank, int world, const void* id128);
int dist_finalize();
int dist_barrier();
const char* comm_mode_name();

// One row block [row0, row0 + nloc) of the system and everything the loop needs for it.
struct System {
  long long n_global = 0;
  long long row0 = 0;
  int nloc = 0;
  bool dense = false;  // values[] is a row-major nloc x n_global block, no index arrays
  CsrDev A;
  PatDev pat;               // row-pattern dictionary form of A (pattern.cu), built on request
  bool pat_active = false;  // the matvec runs k_spmv_pattern instead of the CSR stream kernel
  DiaDev dia;               // diagonal form of a banded A (banded.cu), built on request
  bool dia_active = false;  // the matvec runs k_spmv_dia
  // GPU team (team.cu): the matrix arrays and b / x live in an allocation that the primary process
  // made on this device and exported; this system only borrows them
  bool ext_matrix = false, ext_bx = false;

  // gather space of the matvec: p_ext = [ lower halo | owned rows | upper halo ] (neighbour mode)
  // or the whole global vector (single rank, all-gather mode)
  double* p_ext = nullptr;
  double* p_own = nullptr;  // p_ext + off
  long long off = 0, ext_len = 0;
  int h_lo = 0, h_hi = 0;
  bool allgather = false;
  int world = 1, rank = 0;
  std::vector<long long> part_row0;
  std::vector<int> part_nloc;
  int send_lo = 0, send_hi = 0;  // rows sent to rank-1 / rank+1 in neighbour mode

  // tiles [tb_lo, tb_hi) touch owned columns only; the rest wait for the halo
  int ntiles = 0, tb_lo = 0, tb_hi = 0;

  double *b = nullptr, *x = nullptr, *r = nullptr, *s = nullptr;
  double* w = nullptr;  // single-reduction variant only: w = A r (allocated on first use)
  double* dinv = nullptr;  // 1 / diag(A) of the owned rows (Jacobi preconditioner; built on first use)
  bool pc_loop = false;    // the solve in progress runs the preconditioned vector kernels
  CgState* state = nullptr;
  double* partials = nullptr;
  double* gemv_pool = nullptr;          // dense path: chunk records of the matvec's dynamic tail
  unsigned long long* trace = nullptr;  // phase trace of the one-kernel iteration (diagnostics)
  std::map<int, cudaGraphExec_t> graphs;  // 4 * iterations per graph + 2 * one-kernel + variant -> graph

  // peer-memory exchange (comm.h).  comm_buf = [Mailbox | CommLocal | p_ext] is the one
  // allocation exported through CUDA IPC; peer_base[g] is rank g's comm_buf mapped here.
  bool p2p = false;
  void* comm_buf = nullptr;
  void* peer_base[MAX_RANKS] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  PeerComm cm = {};      // active description (world > 1 only when p2p)
  PeerComm cm_off = {};  // disabled description (single GPU, NCCL mode)

  double matrix_bytes() const;  // bytes of A the matvec streams per launch in the current format
  double iter_bytes() const;  // contract bytes per iteration: A + 88 n
  double k1_bytes() const;    // share of the matvec kernel: A + 16 n
  double comm_bytes_per_iter() const;  // bytes this rank sends to peers per classic iteration
  ~System();
};

struct SolveOpts {
  int max_iter = 0;
  double tol = 0.0;
  int variant = 0;
  int check_every = 0;
  int flags = 0;  // CGB200_OPT_JACOBI (1) | CGB200_OPT_X0 (2)
};

struct SolveResult {
  int iterations;
  int status;
  double rr, bb, relres, seconds, bytes_per_iter;
  int path, world;
  long long kernel_launches;
};

// Systems created while a SoloScope is alive are single-rank whatever world the process is in.
struct SoloScope {
  SoloScope();
  ~SoloScope();
};

System* system_lap(int nx, int ny, int nz, int z0, int nzl, unsigned long long bseed);
System* system_dense(int n, int row0, int nrows, unsigned long long aseed, unsigned long long bseed);
// Copies of large arrays between ordinary (pageable) host memory and the CURRENT device, staged
// through page-locked buffers by a few host threads (solver.cu).  In that case st is synchronised
// first and the data is in place on return; otherwise they are a cudaMemcpyAsync on st.
cudaError_t upload_host(void* dst_dev, const void* src_host, size_t bytes, cudaStream_t st);
cudaError_t download_host(void* dst_host, const void* src_dev, size_t bytes, cudaStream_t st);
void stage_release();  // frees the staging buffer (shutdown)

System* system_from_csr(long long n_global, long long row0, int nrows, const int* row_ptr,
                        const int* cols, const double* vals, const double* b);
// A row block whose arrays are already on this device (padded like alloc_matrix pads them; rowptr
// local from 0, cols global; cols / rowptr null for a full-row block): borrowed, never freed here.
System* system_from_device(long long n_global, long long row0, int nrows, long long nnz, bool dense,
                           double* vals, int* cols, int* rowptr, double* b, double* x);
int classify_tiles(System* S);
bool dense_one_kernel(const System* S);  // classic iterations of S run as one cooperative launch
int system_solve(System* S, const SolveOpts& o, SolveResult* res);
int system_solve_async(System* S, const SolveOpts& o, long long* ticket_out);
int system_solve_wait(long long ticket, SolveResult* res);
// 1 = S now runs the pattern matvec, 0 = not compressible (stays on CSR; reason in last_error), -1 = failure
int system_use_pattern(System* S, int on);
// the same for the diagonal (banded) format; single rank only
int system_use_banded(System* S, int on);
int build_banded(const CsrDev& A, long long off, DiaDev* out, cudaStream_t st, std::string* why);
int build_pattern(const CsrDev& A, PatDev* out, cudaStream_t st, std::string* why);
int system_trace(System* S, int on, unsigned long long* out16);
int system_set_rhs(System* S, const double* b_host);
int system_gen_rhs(System* S, unsigned long long seed);
int system_get_x(System* S, double* x_host);
int system_set_x(System* S, const double* x_host);  // initial guess for a CGB200_OPT_X0 solve
int system_matvec(System* S, const double* v_host, double* s_host);
int system_matvec_dev(System* S, const double* v_dev, double* s_host_or_null);
int system_matvec_into(System* S, const double* v_dev, double* out_dev);
int system_true_relres(System* S, double* relres);
int system_time_kernel(System* S, int which, int reps, double* ms_out);
int system_download(System* S, double* vals, int* cols, int* row_ptr);
int system_get_b(System* S, double* b_host);
int timer_start();
int timer_stop(double* ms_out);

}  // namespace cgb
