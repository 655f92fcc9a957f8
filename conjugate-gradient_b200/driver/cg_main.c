/* cg_main.c -- the `cg` command of the reference, on libcgb200.so.
 *
 * Keeps the reference's command line (cg.c:42-85):
 *
 *     cg <input-data> <max-iterations> [suppress-output]
 *
 * reads the reference's 4-line comma-separated text format (line 0 column indices, 1 row
 * pointers, 2 matrix values, 3 right-hand side; cg.c:173-201), calls conj_grad() and prints what
 * the reference prints ("CG took approx N seconds", then print_sparse(x)).  Differences, all
 * deliberate: a file that cannot be opened or parsed is an error (the reference ignores the
 * reader's return code and iterates on empty structs, cg.c:65); a missing final newline and tokens
 * of any length are accepted, and the file is parsed by several threads (cgb200_read_system) (the reference never terminates / overruns its 64-byte token buffer);
 * `suppress-output` starting with 'y' really suppresses the vector (the reference parses it and
 * then ignores it, cg.c:56-57,78).
 *
 * Extensions (none needed by a reference user):
 *     --tol T            stop when ||r||/||b|| <= T (default 0: the reference's fixed count)
 *     --check-every M    iterations per graph launch between stop-flag polls
 *     --gen dense:N | lap2d:N | lap3d:N    generate the system on the device instead of reading it
 *     --gpus G           row-block partition over G GPUs (one process per GPU, NCCL)
 *     --json             one machine-readable line: iterations, relres, seconds, it/s, GB/s
 *     --variant classic|sr   the reference's recurrence (default) or the single-reduction one
 *     --banded           store a mostly full band as diagonals: 8 bytes per entry, no indices (one GPU)
 *     --pattern          store stencil-like matrices as one byte per row (row-pattern dictionary
 *                        format; same arithmetic, fewer bytes) when the matrix qualifies
 *     --jacobi           diagonal (Jacobi) preconditioner (classic recurrence only)
 *     --dump FILE        write the system in the reference's text format before solving (a --gen
 *                        system becomes an input file the unmodified reference can read);
 *     --dump-binary FILE the same as the binary sidecar cgb200_read_system also accepts
 */
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include "cgb200.h"

/* Reads the 4-line text format (or the binary sidecar) into A and b through the library's reader
 * (csrc/sysio.cpp, the replacement of the reference's read_input_file, cg.c:146-218).
 * 0 on success, -1 with a message on stderr. */
static int read_system(const char *path, struct __mv_sparse *A, struct __mv_sparse *b) {
  if (cgb200_read_system(path, A, b) != 0) {
    fprintf(stderr, "Error: %s\n", cgb200_last_error());
    return -1;
  }
  return 0;
}

/* ---- options ------------------------------------------------------------------------------ */
typedef struct {
  const char *input;
  int max_iter;
  int suppress;
  double tol;
  int check_every;
  const char *gen;
  const char *dump;   /* write the system (text format, or binary sidecar) before solving */
  int dump_binary;
  int gpus;
  int json;
  int variant;
  int pattern;
  int banded;
  int jacobi;
} options;

static void usage(const char *p) {
  fprintf(stderr,
          "Usage: %s <input-data> <max-iterations> [suppress-output]\n"
          "       %s --gen dense:N|lap2d:N|lap3d:N <max-iterations> [--gpus G]\n"
          "       options: --tol T  --check-every M  --json  --variant classic|sr  --pattern  --banded  --jacobi\n"
          "                --dump FILE | --dump-binary FILE   (write the system before solving)\n", p, p);
}

static int parse(int argc, char **argv, options *o) {
  memset(o, 0, sizeof(*o));
  o->max_iter = -1;
  o->gpus = 1;
  int npos = 0;
  for (int i = 1; i < argc; ++i) {
    const char *a = argv[i];
    if (!strcmp(a, "--tol") && i + 1 < argc) o->tol = strtod(argv[++i], NULL);
    else if (!strcmp(a, "--check-every") && i + 1 < argc) o->check_every = atoi(argv[++i]);
    else if (!strcmp(a, "--gen") && i + 1 < argc) o->gen = argv[++i];
    else if (!strcmp(a, "--dump") && i + 1 < argc) o->dump = argv[++i];
    else if (!strcmp(a, "--dump-binary") && i + 1 < argc) { o->dump = argv[++i]; o->dump_binary = 1; }
    else if (!strcmp(a, "--gpus") && i + 1 < argc) o->gpus = atoi(argv[++i]);
    else if (!strcmp(a, "--json")) o->json = 1;
    else if (!strcmp(a, "--pattern")) o->pattern = 1;
    else if (!strcmp(a, "--banded")) o->banded = 1;
    else if (!strcmp(a, "--jacobi")) o->jacobi = 1;
    else if (!strcmp(a, "--variant") && i + 1 < argc) {
      const char *v = argv[++i];
      if (!strcmp(v, "sr") || !strcmp(v, "single-reduction")) o->variant = CGB200_SINGLE_REDUCTION;
      else if (!strcmp(v, "classic")) o->variant = CGB200_CLASSIC;
      else { fprintf(stderr, "unknown variant %s\n", v); return -1; }
    }
    else if (a[0] == '-' && a[1] == '-') { fprintf(stderr, "unknown option %s\n", a); return -1; }
    else {
      if (o->gen && npos == 0) npos = 1; /* no input file with --gen */
      if (npos == 0) o->input = a;
      else if (npos == 1) o->max_iter = (int)strtol(a, NULL, 10);
      else if (npos == 2) o->suppress = a[0] == 'y';
      ++npos;
    }
  }
  if ((!o->input && !o->gen) || o->max_iter < 0 || o->gpus < 1) return -1;
  if (o->pattern) setenv("CGB200_FORMAT", "pattern", 1); /* read by the library when a system is created */
  if (o->banded) setenv("CGB200_FORMAT", "banded", 1);   /* diagonal storage for mostly full bands (one GPU) */
  return 0;
}

static void print_json(const cgb200_result *r, double wall) {
  const double its = r->seconds > 0 ? r->iterations / r->seconds : 0.0;
  printf("{\"iterations\": %d, \"status\": %d, \"relres\": %.6e, \"device_seconds\": %.6f, \"wall_seconds\": %.6f, "
         "\"it_per_s\": %.3f, \"bytes_per_iter_per_rank\": %.0f, \"gbs_per_gpu\": %.1f, \"path\": \"%s\", \"gpus\": %d, "
         "\"kernel_launches\": %lld}\n",
         r->iterations, r->status, r->relres, r->seconds, wall, its, r->bytes_per_iter,
         r->seconds > 0 ? r->bytes_per_iter * r->iterations / r->seconds / 1e9 : 0.0,
         r->path ? "dense row-block GEMV" : "CSR stream SpMV", r->world, r->kernel_launches);
}

/* ---- device-generated / distributed run --------------------------------------------------- */
static int run_rank(const options *o, int rank, int world, const unsigned char *id) {
  if (cgb200_init(rank) != 0) { fprintf(stderr, "cg: %s\n", cgb200_last_error()); return 1; }
  if (world > 1 && cgb200_dist_init(rank, world, id) != 0) {
    fprintf(stderr, "cg[%d]: %s\n", rank, cgb200_last_error());
    return 1;
  }
  cgb200_system *sys = NULL;
  struct __mv_sparse A = {0}, b = {0};
  if (o->gen) {
    char kind[16] = {0};
    int N = 0;
    if (sscanf(o->gen, "%15[^:]:%d", kind, &N) != 2 || N < 1) { fprintf(stderr, "bad --gen %s\n", o->gen); return 1; }
    if (!strcmp(kind, "dense")) {
      int nl = N / world, r0 = rank * nl;
      if (rank == world - 1) nl = N - r0;
      sys = cgb200_system_dense(N, r0, nl, 1, 7);
    } else if (!strcmp(kind, "lap2d")) {
      if (world > 1) { /* y-slabs of the 2D grid are z-slabs of an N x 1 x N grid?  keep it simple */
        fprintf(stderr, "--gen lap2d is single-GPU; use lap3d for --gpus\n");
        return 1;
      }
      sys = cgb200_system_lap(N, N, 1, 0, 1, 7);
    } else if (!strcmp(kind, "lap3d")) {
      int nzl = N / world, z0 = rank * nzl;
      if (rank == world - 1) nzl = N - z0;
      sys = cgb200_system_lap(N, N, N, z0, nzl, 7);
    } else { fprintf(stderr, "bad --gen %s\n", o->gen); return 1; }
  } else {
    if (read_system(o->input, &A, &b) != 0) return 1;
    int nl = A.size / world, r0 = rank * nl;
    if (rank == world - 1) nl = A.size - r0;
    sys = cgb200_system_from_csr(A.size, r0, nl, A.row_ptr + r0, A.col_indices, A.values, b.values + r0);
  }
  if (!sys) { fprintf(stderr, "cg[%d]: %s\n", rank, cgb200_last_error()); return 1; }
  if (o->dump) {
    if (world > 1) { fprintf(stderr, "--dump is single-GPU\n"); return 1; }
    if (o->gen) { /* bring the generated block back as the reference's host structs */
      const long long n = cgb200_system_rows(sys), nnz = cgb200_system_nnz(sys);
      A.size = (int)n; A.nnz = (int)nnz;
      A.values = malloc(sizeof(double) * (nnz > 0 ? nnz : 1));
      A.col_indices = malloc(sizeof(int) * (nnz > 0 ? nnz : 1));
      A.row_ptr = malloc(sizeof(int) * (n + 1));
      b.size = b.nnz = (int)n;
      b.values = malloc(sizeof(double) * (n > 0 ? n : 1));
      if (!A.values || !A.col_indices || !A.row_ptr || !b.values ||
          cgb200_system_download(sys, A.values, A.col_indices, A.row_ptr) != 0 ||
          cgb200_system_get_b(sys, b.values) != 0) {
        fprintf(stderr, "cg: --dump: %s\n", cgb200_last_error());
        return 1;
      }
    }
    if (cgb200_write_system(o->dump, &A, &b, o->dump_binary) != 0) {
      fprintf(stderr, "cg: --dump: %s\n", cgb200_last_error());
      return 1;
    }
  }
  cgb200_opts opts = {o->max_iter, o->tol, o->variant, o->check_every, o->jacobi ? CGB200_OPT_JACOBI : 0};
  cgb200_result res;
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  if (cgb200_system_solve(sys, &opts, &res) != 0) { fprintf(stderr, "cg[%d]: %s\n", rank, cgb200_last_error()); return 1; }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  const double wall = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
  double tr = 0.0;
  if (cgb200_system_true_relres(sys, &tr) != 0) { fprintf(stderr, "cg[%d]: %s\n", rank, cgb200_last_error()); return 1; }
  if (rank == 0) {
    printf("CG took approx %d seconds\n", (int)wall);
    printf("iterations %d  status %d  relres %.3e  true relres %.3e  %.1f it/s\n", res.iterations, res.status,
           res.relres, tr, res.seconds > 0 ? res.iterations / res.seconds : 0.0);
    if (o->json) print_json(&res, wall);
  }
  if (!o->suppress && world == 1) {
    const long long n = cgb200_system_rows(sys);
    struct __mv_sparse *x = new_mv_struct_with_size((int)n);
    cgb200_system_get_x(sys, x->values);
    print_sparse(x);
    cgb200_free_all(x);
  }
  cgb200_system_free(sys);
  if (world > 1) cgb200_dist_finalize();
  cgb200_shutdown();
  return 0;
}

int main(int argc, char **argv) {
  options o;
  if (parse(argc, argv, &o) != 0) { usage(argv[0]); return -1; }

  if (o.gen || o.gpus > 1) {
    if (o.gpus == 1) return run_rank(&o, 0, 1, NULL);
    /* one process per GPU; rank 0 creates the NCCL id and hands it to the others through pipes.
     * CUDA is never initialised in this parent (it would not survive the fork). */
    int (*pipes)[2] = malloc(sizeof(int[2]) * o.gpus);
    for (int r = 1; r < o.gpus; ++r) if (pipe(pipes[r]) != 0) { perror("pipe"); return 1; }
    pid_t *pids = malloc(sizeof(pid_t) * o.gpus);
    for (int r = 0; r < o.gpus; ++r) {
      pids[r] = fork();
      if (pids[r] == 0) {
        unsigned char id[128];
        if (r == 0) {
          if (cgb200_dist_unique_id(id) != 0) { fprintf(stderr, "cg: %s\n", cgb200_last_error()); _exit(1); }
          for (int k = 1; k < o.gpus; ++k) if (write(pipes[k][1], id, 128) != 128) _exit(1);
        } else {
          if (read(pipes[r][0], id, 128) != 128) _exit(1);
        }
        const int rrc = run_rank(&o, r, o.gpus, id);
        fflush(stdout); /* _exit does not: rank 0's report would be lost when stdout is a pipe */
        fflush(stderr);
        _exit(rrc);
      }
    }
    int rc = 0;
    for (int r = 0; r < o.gpus; ++r) {
      int st = 0;
      waitpid(pids[r], &st, 0);
      if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = 1;
    }
    return rc;
  }

  /* ---- the reference's flow (cg.c:59-84) ---- */
  struct __mv_sparse *mat_A = new_mv_struct();
  struct __mv_sparse *vec_b = new_mv_struct();
  struct __mv_sparse *vec_x = NULL;
  if (read_system(o.input, mat_A, vec_b) != 0) return 1;
  if (o.dump && cgb200_write_system(o.dump, mat_A, vec_b, o.dump_binary) != 0) {
    fprintf(stderr, "cg: --dump: %s\n", cgb200_last_error());
    return 1;
  }

  time_t start = time(NULL);
  int rc;
  cgb200_result res;
  memset(&res, 0, sizeof(res));
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  if (o.tol > 0.0 || o.json || o.check_every > 0 || o.jacobi || o.variant != CGB200_CLASSIC) {
    cgb200_opts opts = {o.max_iter, o.tol, o.variant, o.check_every, o.jacobi ? CGB200_OPT_JACOBI : 0};
    rc = cgb200_solve(mat_A, vec_b, &vec_x, &opts, &res);
  } else {
    rc = conj_grad(o.max_iter, mat_A, vec_b, &vec_x);
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  time_t end = time(NULL);
  if (rc != 0) {
    fprintf(stderr, "cg: solve failed: %s\n", cgb200_last_error());
    return 1;
  }
  printf("CG took approx %d seconds\n", (int)(end - start));
  if (o.json) print_json(&res, (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec));
  if (!o.suppress) print_sparse(vec_x);

  cgb200_free_all(vec_x);
  cgb200_free_all(mat_A);
  cgb200_free_all(vec_b);
  cgb200_shutdown();
  return 0;
}
