/* cgb200.h -- extensions of libcgb200.so beyond the reference's mv_ops.h / conj_grad surface.
 *
 * None of these symbols is required by a drop-in caller of mv_ops.h.  They expose what the
 * reference does not have: a tolerance stop, device-resident systems (so large matrices never
 * pass through host structs), row-block partitioning across GPUs, and measurement hooks.
 * Plain C ABI: pointers and sizes only.
 */
#ifndef CGB200_H
#define CGB200_H

#include <stdint.h>

#include "mv_ops.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status of a solve ------------------------------------------------------------------- */
#define CGB200_RUNNING 0
#define CGB200_MAXITER 1   /* ran the fixed count (the reference's only exit, cg.c:125-127) */
#define CGB200_CONVERGED 2 /* r.r <= tol^2 * b.b */
#define CGB200_BREAKDOWN 3 /* p.Ap zero or not finite: stopped, last finite x kept (a negative p.Ap
                              is let through as the reference lets it, cg.c:113) */
#define CGB200_EXACT 4     /* r.r == 0: stopped where the reference would compute 0/0 = NaN */
#define CGB200_COMM_TIMEOUT 5 /* multi-GPU: a peer-memory exchange timed out (the call returns -1) */

/* ---- recurrence variants ------------------------------------------------------------------ */
#define CGB200_CLASSIC 0          /* Hestenes-Stiefel, the reference's recurrence (default) */
/* Chronopoulos-Gear recurrence: w = A r is formed instead of A p, r.r and r.w come out of ONE
 * fused reduction inside the matvec kernel, and the four vector updates are one pass.  Same x in
 * exact arithmetic, two kernels and one (cross-GPU) reduction per iteration instead of three and
 * two; costs one more matvec at loop entry.  Same iteration-count semantics as the classic loop. */
#define CGB200_SINGLE_REDUCTION 1

/* ---- solve options beyond the reference (cgb200_opts.flags; classic recurrence only) ---------
 * The reference always starts from x = 0 (cg.c:104-108) and has no preconditioner; both options
 * are additive and leave the default path untouched.
 * CGB200_OPT_JACOBI  diagonal (Jacobi) preconditioner M = diag(A): z = M^-1 r is formed inside the
 *                    two vector kernels (never stored), alpha = r.z / p.Ap, beta = r.z / (r.z)_prev,
 *                    p = z + beta p; the stop test still looks at the true r.r.  Costs 16 n more
 *                    bytes per iteration.  Refused (-1) when a row has no positive finite diagonal.
 * CGB200_OPT_X0      start from an initial guess: r0 = b - A x0 (one more matvec at loop entry).
 *                    cgb200_solve: *x on entry is the guess (not modified, not freed; *x is replaced
 *                    by a freshly allocated vector as in conj_grad).  cgb200_system_solve: the guess
 *                    is the x the system holds -- the previous solve's, or cgb200_system_set_x. */
#define CGB200_OPT_JACOBI 1
#define CGB200_OPT_X0 2

/* ---- operator-API residency modes --------------------------------------------------------- */
#define CGB200_EAGER 0    /* every op uploads its inputs and downloads its output (default) */
#define CGB200_RESIDENT 1 /* outputs stay on the device until cgb200_to_host()/print/get_row */

typedef struct cgb200_opts {
  int max_iter;    /* reference semantics: M performs M+1 matvecs and x/r updates */
  double tol;      /* stop when ||r||/||b|| <= tol; 0 = fixed count (reference behaviour) */
  int variant;     /* CGB200_CLASSIC | CGB200_SINGLE_REDUCTION */
  int check_every; /* iterations per graph launch / host poll of the stop flag; 0 = default */
  int flags;       /* 0 (the reference's loop) or CGB200_OPT_JACOBI | CGB200_OPT_X0 */
} cgb200_opts;

typedef struct cgb200_result {
  int iterations; /* x updates performed (matvecs) */
  int status;     /* CGB200_MAXITER ... */
  double rr;      /* final recursive r.r */
  double bb;      /* b.b */
  double relres;  /* sqrt(rr/bb) */
  double seconds; /* device time of the loop (CUDA events on the solve stream) */
  double bytes_per_iter; /* contract bytes this rank moves per iteration */
  int path;       /* 0 = CSR stream SpMV, 1 = dense row-block GEMV */
  int world;      /* ranks that took part */
  long long kernel_launches; /* kernels launched by this solve (graph nodes included) */
} cgb200_result;

/* ---- process-wide context ----------------------------------------------------------------- */

/* Binds the library to a CUDA device (default: LOCAL_RANK, else 0).  Optional: the first call
 * that needs the device initialises lazily.  0 / -1. */
int cgb200_init(int device);
void cgb200_shutdown(void);
/* Message of the last failure on the calling process ("" if none). */
const char *cgb200_last_error(void);
/* Kernels launched by the library so far in this process (graph launches count their nodes). */
long long cgb200_launch_count(void);
/* 1 when a CUDA device is usable, 0 otherwise (never touches the oracle or any CPU path). */
int cgb200_device_available(void);
/* Guard mode (debug aid; environment variable CGB200_GUARD=1, read at the first allocation): every
 * device buffer of the library sits between two 4 KiB guard bands and is filled, guards and
 * payload, with 0xFF -- an out-of-bounds store damages a guard, an out-of-bounds or
 * uninitialised load yields NaN / -1 and breaks parity.  Returns -1 when guard mode is off, else
 * the number of buffers whose guards were found damaged so far (freed ones included;
 * cgb200_last_error() and stderr name the allocation site).  Synchronises the device. */
long long cgb200_guard_check(void);
/* Positive control: stores one byte just outside a scratch buffer on either side and reports
 * whether the guards caught both (1), missed one (0); -1 when guard mode is off. */
int cgb200_guard_selftest(void);

/* ---- operator-API residency (struct address -> device mirror table) ------------------------ */
/* CGB200_EAGER (default): every operator call uploads its inputs and downloads its output --
 * correct for any caller.  CGB200_RESIDENT: outputs stay on the device (host arrays go stale)
 * until cgb200_to_host / print_sparse / mat_get_row / mv_deep_copy of a matrix ask for them;
 * mv_deep_copy of a device-resident vector copies device to device.  The environment variable
 * CGB200_MODE=resident selects resident mode at start-up for callers that cannot be edited (the
 * reference's own cg.c loop: tests/test_gpu_dropin.py). */
int cgb200_set_mode(int mode);
/* Device copies of caller matrices (mv_mult, conj_grad, cgb200_solve on host structs).  The
 * reference reads the caller's arrays on every call (mv_ops.c:187-197), so a caller may edit
 * values[] in place between two calls; the device copy is re-validated on every call:
 *   CGB200_CACHE_TRACKED (default) the kernel is asked whether the arrays were WRITTEN since the
 *                        upload (userfaultfd asynchronous write-protection + PAGEMAP_SCAN, Linux
 *                        >= 6.7: the page-aligned interior of each array of >= 1 MiB; their partial
 *                        first / last pages and smaller arrays are hashed): exact for every write
 *                        that goes through the CPU, ~1 ms per 10 GB instead of a pass over the
 *                        arrays.  Written, or the kernel cannot tell (no support, page-locked
 *                        memory -- device DMA bypasses the page tables --, any error): the full
 *                        hash below decides.  The caller pays one minor fault per 4 KiB page it
 *                        rewrites.  CGB200_WRITE_TRACKING=0 disables the tracking.  mv_mult and
 *                        the solves ask the question on a host thread while the device already
 *                        works on the copy it has, and look at the answer before any result leaves
 *                        the library: a write means that result is dropped and the call repeats
 *                        after the ordinary validation.  Two structs over the same arrays share
 *                        one watch; the arrays must not be written WHILE a call on them runs (the
 *                        reference reads them during the call too).
 *   CGB200_CACHE_STRICT  hash of every byte of values / col_indices / row_ptr on every call,
 *                        several host threads: any in-place edit is seen and causes a fresh upload;
 *   CGB200_CACHE_SAMPLED 96 samples + the array addresses (constant time): notices a rebuilt
 *                        matrix, can miss a small in-place edit -- call cgb200_invalidate;
 *   CGB200_CACHE_OFF     every call uploads.
 * Environment: CGB200_MATRIX_CACHE=tracked|strict|sampled|off.  A matrix handed over explicitly
 * with cgb200_to_device(A), and every matrix while in CGB200_RESIDENT mode, is pinned: the caller
 * has declared the device copy current until cgb200_invalidate(A) / free_mv_struct(A), and only the
 * sampled check runs.  cgb200_cache_stats: {uploads, full hashes, calls settled by write tracking
 * alone, calls where tracking reported a write or could not tell, tracking available (0/1)}. */
#define CGB200_CACHE_STRICT 0
#define CGB200_CACHE_SAMPLED 1
#define CGB200_CACHE_OFF 2
#define CGB200_CACHE_TRACKED 3
int cgb200_cache_stats(long long *out5);
/* 1 if the kernel's write tracking is usable in this process (probed and self-tested on first use;
 * host only, needs no GPU), 0 if CGB200_CACHE_TRACKED will hash like CGB200_CACHE_STRICT. */
int cgb200_write_tracking_available(void);
/* The mechanism itself, for tests and diagnostics (host only, no GPU; not needed by callers).
 * [p, p + bytes) is ordinary anonymous memory; only its page-aligned interior is covered.
 * op 0: start watching (0 / -1); op 1: was any page written since the marks were last cleared?
 * (1 / 0 / -1 = cannot tell, e.g. not watched); op 2: clear the marks (0 / -1); op 3: stop watching. */
int cgb200_write_tracking_op(const void *p, size_t bytes, int op);
int cgb200_set_matrix_cache(int policy);
/* The content hash CGB200_CACHE_STRICT compares (host only, needs no device; 0 for NULL): every
 * byte of values / col_indices / row_ptr (a vector: its values).  Independent of the number of
 * hashing threads (CGB200_HASH_THREADS). */
unsigned long long cgb200_content_hash(const struct __mv_sparse *A);
int cgb200_to_device(struct __mv_sparse *obj); /* upload now; the mirror becomes valid (matrices: pinned) */
int cgb200_to_host(struct __mv_sparse *obj);   /* bring values back if the device copy is newer */
int cgb200_invalidate(struct __mv_sparse *obj); /* caller changed host arrays in place */
void cgb200_free_all(struct __mv_sparse *obj);  /* struct + its host arrays + device mirror */

/* ---- several GPUs behind conj_grad / cgb200_solve ------------------------------------------- */
/* The reference's entry point takes host structs and knows nothing about GPUs (cg.c:24).  With
 * gpus > 1 -- or the environment variable CGB200_GPUS=N ("all": every visible GPU), for callers
 * that cannot be edited, e.g. the reference's own cg.c relinked -- conj_grad and cgb200_solve cut
 * the caller's matrix into N contiguous row blocks of equal stored entries, one per GPU, and run
 * the same loop over all of them: the calling process becomes rank 0 of a team of N processes (it
 * starts N-1 `cgb200_worker` processes, found next to the library or through CGB200_WORKER, which
 * stay up until cgb200_shutdown or exit), puts every block on its GPU itself and copies b in and x
 * out of every GPU per call.  The matrix cache (above) applies unchanged: an unchanged matrix stays
 * on its GPUs.  Systems with fewer than 64 rows per GPU, and every call when the team cannot be
 * started (fewer GPUs visible, worker missing: one line on stderr), run on one GPU.  mv_mult and
 * the vector operators always run on the calling process's GPU.  0 / -1 (gpus outside 1..8). */
int cgb200_set_gpus(int gpus);
int cgb200_gpus_in_use(void); /* ranks of the running team, 1 without one */
/* Body of the cgb200_worker executable (driver/cgb200_worker.c); not for users. */
int cgb200_worker_main(int argc, char **argv);

/* ---- solve on host structs with options ---------------------------------------------------- */
/* conj_grad(M, A, b, &x) == cgb200_solve with {max_iter = M, tol = 0, CGB200_CLASSIC}. */
int cgb200_solve(struct __mv_sparse *A, struct __mv_sparse *b, struct __mv_sparse **x,
                 const cgb200_opts *opts, cgb200_result *res);

/* ---- device-resident systems (one row block of A, b, x and the work vectors) -------------- */
/* Threading: the library has one context per process (one device, one solve stream).  Every entry
 * point of this header and of mv_ops.h may be called from any thread; calls are serialised by one
 * process-wide lock (the reference's operators are stateless, mv_ops.c has no globals -- here the
 * state is the device).  cgb200_solve_wait holds the lock while it waits. */
typedef struct cgb200_system cgb200_system;

/* Rows [z0*nx*ny, (z0+nzl)*nx*ny) of the Dirichlet 5-point (nz == 1) or 7-point Laplacian on an
 * nx*ny*nz grid, natural ordering, generated on the device as CSR; b_i = U(-1/2,1/2)(bseed, i). */
cgb200_system *cgb200_system_lap(int nx, int ny, int nz, int z0, int nzl, uint64_t bseed);
/* Rows [row0, row0+nrows) of the dense SPD test matrix, values only. */
cgb200_system *cgb200_system_dense(int n, int row0, int nrows, uint64_t aseed, uint64_t bseed);
/* Rows [row0, row0+nrows) of a CSR matrix held in host arrays.  row_ptr[nrows+1] are this block's
 * row offsets INTO cols[] / vals[]: either a local array that starts at 0 with the block's own
 * cols / vals, or a slice of the global row_ptr (row_ptr = global_row_ptr + row0, starting at any
 * offset) with the GLOBAL cols / vals arrays -- the block's entries are then
 * cols[row_ptr[0] .. row_ptr[nrows]).  Column indices are global, 0-based and ascending within a
 * row; b holds the block's nrows right-hand-side entries (may be NULL: set later).  The arrays are
 * checked on upload (column range, non-decreasing row_ptr): a malformed matrix is refused with
 * NULL and a message.  A full-row matrix (row_ptr[i] == i*n_global) takes the dense path. */
cgb200_system *cgb200_system_from_csr(long long n_global, long long row0, int nrows,
                                      const int *row_ptr, const int *cols, const double *vals,
                                      const double *b);
void cgb200_system_free(cgb200_system *sys);
int cgb200_system_set_rhs(cgb200_system *sys, const double *b_host);   /* H2D, nrows doubles */
int cgb200_system_gen_rhs(cgb200_system *sys, uint64_t bseed);         /* on the device */
int cgb200_system_get_x(cgb200_system *sys, double *x_host);           /* D2H, nrows doubles */
int cgb200_system_set_x(cgb200_system *sys, const double *x_host);     /* H2D: guess for CGB200_OPT_X0 */
int cgb200_system_solve(cgb200_system *sys, const cgb200_opts *opts, cgb200_result *res);
/* Queues a fixed-count solve (opts->tol must be 0: the reference's semantics) behind whatever is
 * already queued on the library's stream and returns at once with a ticket; right-hand sides
 * set with cgb200_system_gen_rhs are queued the same way, so many solves run back to back without
 * the GPU waiting for the host.  At most 8 tickets may be outstanding.  cgb200_solve_wait blocks
 * until that solve has finished and fills the result.  x stays on the device: read it with
 * cgb200_system_get_x after the wait of the LAST queued solve on that system. */
int cgb200_system_solve_async(cgb200_system *sys, const cgb200_opts *opts, long long *ticket);
int cgb200_solve_wait(long long ticket, cgb200_result *res);
/* s = A v for a host vector v[n_local] (single rank) -> s_host; exercises the loop's matvec
 * kernel alone (mode PLAIN). */
int cgb200_system_matvec(cgb200_system *sys, const double *v_host, double *s_host);
/* ||b - A x|| / ||b|| recomputed from scratch on the device (all ranks get the global value). */
int cgb200_system_true_relres(cgb200_system *sys, double *relres);
/* Average device time (ms, CUDA events on the launching stream) of `reps` back-to-back launches
 * of one kernel of the loop on this system's buffers.  which: 1 = matvec+dot (K1),
 * 2 = x/r update + r.r (K2), 3 = direction update (K3); single-reduction variant: 4 = the fused
 * vector update, 5 = matvec with both dot products; 6 = the dense path's persistent solve kernel
 * (only where cgb200_system_one_kernel_iteration() is 1): launches of 5 iterations each, and the
 * figure returned is the time per ITERATION (launch duration / 5). */
int cgb200_system_time_kernel(cgb200_system *sys, int which, int reps, double *ms);
/* 1 if a classic solve of this system is ONE cooperative launch (dense row block, one GPU or
 * peer-memory mode): loop entry and every iteration -- matvec, both reductions, both vector
 * updates -- behind grid barriers; the launch ends by itself when the loop stops. */
int cgb200_system_one_kernel_iteration(cgb200_system *sys);
/* Process-wide switch for the above (default on; CGB200_DENSE_FUSED=0 also turns it off): with 0 the
 * dense path runs the same three kernels per iteration as the CSR path.  For comparison runs. */
void cgb200_set_one_kernel_iteration(int on);
/* Small single-GPU CSR systems (vectors covered by one resident wave of CTAs, about 600 000 unknowns)
 * run the two vector kernels of an iteration as ONE co-operative launch with a grid barrier behind
 * the r.r reduction: 2 launches per iteration instead of 3, same bits.  Process-wide switch (default
 * on; CGB200_FUSE_VEC=0 also turns it off), for systems solved afterwards.  For comparison runs. */
void cgb200_set_fused_vector_update(int on);
/* Row-pattern dictionary format (not in the reference, whose only layout is CSR,
 * mv_ops.h:17-23).  A CSR row block with at most 255 distinct rows -- a row being the list of
 * (column - row, value) pairs, as in constant-coefficient stencil and Toeplitz band matrices -- is
 * re-encoded as one byte per row plus a dictionary, and the matvec of the loop streams that byte
 * instead of 12 bytes per entry.  Row sums are formed from the same values in the same ascending
 * column order with unfused multiply and add: s = A p is bit-identical to the CSR kernel's.
 * on != 0: build (first time) and switch; returns 1 if the system now runs the pattern matvec,
 * 0 if the matrix does not qualify (it stays on CSR; cgb200_last_error() says why), -1 on a CUDA
 * failure.  on == 0 switches back to CSR.  With several ranks each rank decides for its own rows.
 * The environment variable CGB200_FORMAT=pattern makes every CSR system try the format when it is
 * created (callers that cannot be edited; `cg --pattern`). */
int cgb200_system_use_pattern_format(cgb200_system *sys, int on);
/* Diagonal (DIA) storage for banded matrices with arbitrary coefficients (not in the
 * reference).  A CSR block whose entries lie in a band lo <= column - row <= hi that is at least
 * 70 % full (the reference's course matrix: 345 entries per row, cg.c:231-232,260) is re-stored as
 * hi - lo + 1 diagonals of one value per row; the matvec of the loop then streams 8 bytes per entry
 * instead of 12, without index loads or gathers.  Row sums are formed in ascending column order
 * with unfused multiply and add and padding contributes exact zeros: s = A p is bit-identical to
 * the CSR kernels' for finite p.  Same return convention as cgb200_system_use_pattern_format (1 = now
 * active, 0 = does not qualify and stays on CSR, -1 = CUDA failure; on == 0 switches back).  One
 * GPU only (a multi-rank block answers 0).  CGB200_FORMAT=banded tries it on every CSR system at
 * creation.  Reported beside the CSR contract figures, never instead of them. */
int cgb200_system_use_banded_format(cgb200_system *sys, int on);
/* Diagnostics of the one-kernel iteration: SM-cycle counts of its phases, summed over launches by
 * the CTA that closes each phase, relative to that CTA's start.  out16 (may be NULL) receives the
 * counters accumulated so far: [0] wait for the peers' slices of p, [1] GEMV done (last ticket),
 * [2] (ns) first CTA start to last CTA's end of the GEMV, [3] alpha known (after the cross-GPU sum), [4] x/r update done,
 * [5] beta known, [6] p updated and pushed, [7] launches traced; global-timer ns over all CTAs
 * [12] sum of the CTAs' GEMV durations, and per-launch spreads summed over launches: [13] latest
 * minus earliest CTA start, [14] latest minus earliest end of a CTA's GEMV share, [15] earliest
 * start to earliest end ([8..11] are scratch); [16 + c] sum of CTA c's GEMV durations (ns),
 * [16 + 1024 + c] the SM CTA c ran on.  out16 holds 16 + 2048 words.
 * on != 0 (re)starts tracing, on == 0 stops it; either way the system's captured graphs are
 * rebuilt. */
int cgb200_system_trace(cgb200_system *sys, int on, unsigned long long *out16);
/* Contract bytes per iteration for this rank and the share of the matvec kernel. */
int cgb200_system_bytes(cgb200_system *sys, double *iter_bytes, double *k1_bytes);
/* Bytes this rank sends to its peers over NVLink per classic iteration: its share of the
 * search-direction exchange (halo rows, or its whole block to every peer for a dense system) plus
 * the two scalar sums.  0 on one GPU. */
double cgb200_system_comm_bytes(cgb200_system *sys);
/* Copies the matrix block to host arrays (any may be NULL): vals[nnz], cols[nnz] (global column
 * indices), row_ptr[rows+1] (local, from 0).  Lets a device-generated system be handed to code
 * that wants the reference's host structs (the cg driver's --gen, bench.py's end-to-end leg). */
int cgb200_system_download(cgb200_system *sys, double *vals, int *cols, int *row_ptr);
int cgb200_system_get_b(cgb200_system *sys, double *b_host);
/* One CUDA-event pair on the library's solve stream: start, run any calls, stop -> elapsed ms. */
int cgb200_timer_start(void);
int cgb200_timer_stop(double *ms);
long long cgb200_system_rows(cgb200_system *sys);
long long cgb200_system_nnz(cgb200_system *sys);

/* ---- an entry point that allocates nothing ---------------------------------------------------- */
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  *result_dev = sum_i a_dev[i] * b_dev[i] for DEVICE vectors of n doubles, on the current
 * device, enqueued on `stream` (a cudaStream_t; NULL = the default stream) and not waited for.  The
 * reduction's scratch (one partial sum per 2^18 elements and a ticket) lives in a file-scope static
 * __device__ array of 8.5 KB that every GPU which loads the library holds until exit, hence the
 * limit: n <= CGB200_DOT_STATIC_MAX_N, larger n is refused with -1 and a message.  Neither this call
 * nor anything it calls allocates or frees device memory, and it does not create the library's
 * context.  Every call on a device shares the one array: calls that overlap on the same GPU (two
 * streams, two threads) race -- serialise them.  The result has the same bits on every GPU and for
 * every grid size (partials belong to chunks of the vectors and are added in chunk order); it
 * agrees with dot_product to the summation-order bound 4 n eps sum|a_i b_i|.  0 / -1. */
#define CGB200_DOT_STATIC_MAX_N 268435456LL /* 2^28 */
int cgb200_dot_static(const double *a_dev, const double *b_dev, long long n, double *result_dev,
                      void *stream);

/* ---- systems on disk (host only) ------------------------------------------ */
/* Replaces read_input_file (reference cg.c:146-218): fills caller-created A and b (new_mv_struct)
 * from the reference's 4-line comma-separated text format -- line 0 column indices, 1 row
 * pointers, 2 matrix values, 3 right-hand side; n = row pointers - 1 (cg.c:204) -- or from the
 * binary sidecar cgb200_write_system(..., 1) produces (recognised by its magic).  The file is
 * mapped and parsed by several threads (CGB200_IO_THREADS, default: all cores up to 32) into
 * malloc'ed arrays of the exact size, which the caller owns like the reference's.  Every value has
 * the bits the reference's strtol/strtod give it.  Unlike the reference (whose return code main
 * ignores, cg.c:65, and whose reader spins forever on a file without a final newline): 0 / -1 with
 * cgb200_last_error(); a missing final newline, CR LF and blanks around tokens are accepted; an
 * empty or non-numeric token, a missing line or inconsistent counts are errors.  Needs no GPU. */
int cgb200_read_system(const char *path, struct __mv_sparse *A, struct __mv_sparse *b);
/* Writes A (CSR) and b in the text format above (binary == 0: integers in decimal, reals as the
 * shortest decimal that reads back to the same bits, so the reference's reader reproduces the
 * arrays exactly) or as the binary sidecar (binary != 0).  0 / -1. */
int cgb200_write_system(const char *path, const struct __mv_sparse *A, const struct __mv_sparse *b,
                        int binary);

/* ---- multi-GPU: one process per GPU, NCCL over NVLink -------------------------------------- */
/* Rank 0 obtains a 128-byte NCCL unique id and ships it to the other ranks by any means
 * (torch.distributed broadcast in bench.py, a pipe in the cg driver). */
/* Host-only planning of the row-block decomposition (csrc/partition.h).  info4 holds, per rank,
 * {row0, nloc, mincol, maxcol} (global column span of that rank's rows).  out10 receives
 * {allgather, h_lo, h_hi, off, ext_len, col_shift, send_lo, send_hi, recv_lo, recv_hi}.
 * Needs no device. */
int cgb200_plan_partition(int world, int rank, long long n_global, const long long *info4,
                          int dense, long long *out10);
/* Peer-memory mode: which of its rows rank `rank` stores into which peer's gather vector, and
 * whose completion flags it waits for (csrc/partition.h plan_push; host-only, needs no device).
 * out[42]: {npush, 8 x {peer, begin, end, dst_index}, nwait, 8 x wait_src}; unused entries -1/0. */
int cgb200_plan_push(int world, int rank, long long n_global, const long long *info4, int dense,
                     long long *out42);
int cgb200_dist_unique_id(void *id128);
int cgb200_dist_init(int rank, int world, const void *id128);
int cgb200_dist_finalize(void);
int cgb200_dist_rank(void);
int cgb200_dist_world(void);
/* All ranks have finished the work queued so far (set-up / tear-down ordering). */
int cgb200_dist_barrier(void);
/* How the GPUs exchange data inside the loop: "single" (one rank), "p2p" (default: kernels store
 * halo rows and partial sums straight into peer memory over NVLink, CUDA IPC mappings) or "nccl"
 * (NCCL send/recv, all-gather and all-reduce between the kernels; CGB200_COMM=nccl, and the
 * automatic fallback when IPC mapping is not possible -- see cgb200_system_comm_is_p2p). */
const char *cgb200_dist_comm_mode(void);
int cgb200_system_comm_is_p2p(cgb200_system *sys);

#ifdef __cplusplus
}
#endif

#endif /* CGB200_H */
