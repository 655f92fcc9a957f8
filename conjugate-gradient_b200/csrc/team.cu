// team.cu -- several GPUs behind conj_grad(max_iter, A, b, &x): see team.h.
//
// Protocol between the primary (rank 0, the caller's process) and its workers: fixed-size
// messages over a pipe pair per worker; bulk data never travels through them -- the primary puts
// every block on its GPU itself and the worker maps the device buffers (CUDA IPC).
#include "team.h"

#include <dlfcn.h>
#include <errno.h>
#include <signal.h>
#include <spawn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mv_ops.h"

extern char** environ;

namespace cgb {

namespace {

enum : int { CMD_CREATE = 1, CMD_SOLVE = 2, CMD_FREE = 3, CMD_QUIT = 4 };

struct TeamMsg {
  int cmd;
  int ok;            // reply: 1 = done, 0 = failed (err says why)
  long long n_global, row0, nnz;
  int nrows, dense;
  int handle;        // which of the worker's systems (create assigns, solve / free name it)
  size_t off_vals, off_cols, off_rowptr, off_b, off_x, bytes;
  cudaIpcMemHandle_t mem;
  SolveOpts o;
  SolveResult r;
  char err[200];
};

int write_all(int fd, const void* p, size_t n) {
  const char* c = (const char*)p;
  while (n > 0) {
    const ssize_t k = write(fd, c, n);
    if (k < 0) {
      if (errno == EINTR) continue;
      return -1;
    }
    c += k;
    n -= (size_t)k;
  }
  return 0;
}

int read_all(int fd, void* p, size_t n) {
  char* c = (char*)p;
  while (n > 0) {
    const ssize_t k = read(fd, c, n);
    if (k < 0) {
      if (errno == EINTR) continue;
      return -1;
    }
    if (k == 0) return -1;  // the other side is gone
    c += k;
    n -= (size_t)k;
  }
  return 0;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
void clear_msg(TeamMsg& m) { memset((void*)&m, 0, sizeof(TeamMsg)); }

struct Worker {
  pid_t pid = -1;
  int to = -1, from = -1;  // pipe ends of the primary
  int device = -1;
};

int g_requested = -1;   // -1: CGB200_GPUS not looked at yet
bool g_active = false;
bool g_broken = false;  // a worker died or a collective failed: the team is not used again
std::vector<Worker> g_workers;  // index = rank - 1

void team_fail(const std::string& why) {
  g_broken = true;
  set_error("GPU team: " + why);
}

std::string worker_path() {
  if (const char* e = getenv("CGB200_WORKER")) return e;
  Dl_info info;
  if (dladdr((const void*)&team_fail, &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    const size_t k = p.rfind('/');
    return (k == std::string::npos ? std::string(".") : p.substr(0, k)) + "/cgb200_worker";
  }
  return "cgb200_worker";
}

}  // namespace

struct TeamSystem {
  System* mine = nullptr;             // rank 0's block, in this process
  long long n = 0;
  bool dense = false;
  std::vector<long long> row0;        // per rank
  std::vector<int> nrows;
  std::vector<void*> base;            // per rank >= 1: the allocation on that rank's device
  std::vector<size_t> off_b, off_x;
  std::vector<int> handle;            // per rank >= 1: the worker's name for its system
};

int team_requested() {
  if (g_requested < 0) {
    const char* e = getenv("CGB200_GPUS");
    int n = e ? atoi(e) : 1;
    if (e && strcmp(e, "all") == 0) {
      if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        n = 1;
      }
    }
    if (n > MAX_RANKS) n = MAX_RANKS;
    g_requested = n < 1 ? 1 : n;
  }
  return g_requested;
}

void team_request(int gpus) {
  if (gpus > MAX_RANKS) gpus = MAX_RANKS;
  g_requested = gpus < 1 ? 1 : gpus;
}

bool team_active() { return g_active && !g_broken; }

// Spawns the workers, joins them and this process into one NCCL world.  The workers take the
// devices after this process's own, in order.
int team_start() {
  if (g_active) return g_broken ? -1 : 0;
  const int n = team_requested();
  if (n <= 1 || g_broken) return -1;
  if (ensure_init(-1)) return -1;
  Ctx& c = ctx();
  if (c.world > 1) {
    set_error("GPU team: this process already belongs to a multi-rank world (CGB200_GPUS is for single-process callers)");
    g_requested = 1;
    return -1;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < n) {
    cudaGetLastError();
    set_error("GPU team: CGB200_GPUS asks for " + std::to_string(n) + " GPUs, " + std::to_string(ndev) + " visible");
    g_requested = 1;  // do not try again on every call
    return -1;
  }
  const std::string exe = worker_path();
  if (access(exe.c_str(), X_OK) != 0) {
    set_error("GPU team: worker executable not found: " + exe + " (built by make; CGB200_WORKER overrides the path)");
    g_requested = 1;
    return -1;
  }
  unsigned char id[128];
  if (dist_unique_id(id)) {
    g_requested = 1;
    return -1;
  }
  signal(SIGPIPE, SIG_IGN);  // a dead worker is reported by write(), not by killing the caller
  g_workers.assign(n - 1, Worker());
  for (int r = 1; r < n; ++r) {
    int down[2], up[2];  // primary -> worker, worker -> primary
    if (pipe(down) != 0 || pipe(up) != 0) {
      team_fail("pipe() failed");
      return -1;
    }
    Worker& w = g_workers[r - 1];
    w.device = (c.device + r) % ndev;
    char a_rank[16], a_world[16], a_dev[16], a_in[16], a_out[16];
    snprintf(a_rank, sizeof a_rank, "%d", r);
    snprintf(a_world, sizeof a_world, "%d", n);
    snprintf(a_dev, sizeof a_dev, "%d", w.device);
    snprintf(a_in, sizeof a_in, "%d", down[0]);
    snprintf(a_out, sizeof a_out, "%d", up[1]);
    char* argv[] = {(char*)exe.c_str(), a_rank, a_world, a_dev, a_in, a_out, nullptr};
    posix_spawn_file_actions_t fa;
    posix_spawn_file_actions_init(&fa);
    posix_spawn_file_actions_addclose(&fa, down[1]);
    posix_spawn_file_actions_addclose(&fa, up[0]);
    for (int k = 0; k < r - 1; ++k) {  // earlier workers' pipe ends are none of this one's business
      posix_spawn_file_actions_addclose(&fa, g_workers[k].to);
      posix_spawn_file_actions_addclose(&fa, g_workers[k].from);
    }
    const int rc = posix_spawn(&w.pid, exe.c_str(), &fa, nullptr, argv, environ);
    posix_spawn_file_actions_destroy(&fa);
    close(down[0]);
    close(up[1]);
    if (rc != 0) {
      team_fail(std::string("cannot start ") + exe + ": " + strerror(rc));
      return -1;
    }
    w.to = down[1];
    w.from = up[0];
    if (write_all(w.to, id, sizeof id)) {
      team_fail("worker " + std::to_string(r) + " does not answer");
      return -1;
    }
  }
  // two handshakes: "my device is up" before anybody enters the NCCL rendezvous (a worker that
  // cannot start must not leave the others waiting in it), "I have joined" after it
  for (int pass = 0; pass < 2; ++pass) {
    if (pass == 1 && dist_init(0, n, id)) {  // returns once every worker has joined
      g_broken = true;
      return -1;
    }
    for (int r = 1; r < n; ++r) {
      TeamMsg m;
      clear_msg(m);
      if (read_all(g_workers[r - 1].from, &m, sizeof m) || !m.ok) {
        m.err[sizeof m.err - 1] = 0;
        team_fail("worker " + std::to_string(r) + " failed to start: " + (m.err[0] ? m.err : "no answer"));
        return -1;
      }
    }
  }
  c.team_primary = true;
  g_active = true;
  return 0;
}

void team_stop() {
  if (g_workers.empty()) return;
  for (Worker& w : g_workers) {
    if (w.to >= 0) {
      TeamMsg m;
      clear_msg(m);
      m.cmd = CMD_QUIT;
      write_all(w.to, &m, sizeof m);
    }
  }
  if (g_active) dist_finalize();  // (each side destroys its own communicator)
  for (Worker& w : g_workers) {
    if (w.to >= 0) close(w.to);
    if (w.from >= 0) close(w.from);
    if (w.pid > 0) {
      int st = 0;
      waitpid(w.pid, &st, 0);
    }
  }
  g_workers.clear();
  ctx().team_primary = false;
  g_active = false;
  g_broken = false;
}

// One request to every worker, then (after `between`, the primary's own share of the collective)
// one reply from each.  Replies are left in `msgs`.
template <typename F>
static int team_round(std::vector<TeamMsg>& msgs, F between) {
  for (size_t k = 0; k < g_workers.size(); ++k)
    if (write_all(g_workers[k].to, &msgs[k], sizeof(TeamMsg))) {
      team_fail("worker " + std::to_string(k + 1) + " is gone");
      return -1;
    }
  const int mine = between();
  const std::string my_err = mine != 0 ? last_error() : "";
  int rc = mine;
  for (size_t k = 0; k < g_workers.size(); ++k) {
    if (read_all(g_workers[k].from, &msgs[k], sizeof(TeamMsg))) {
      team_fail("worker " + std::to_string(k + 1) + " died");
      return -1;
    }
    if (!msgs[k].ok) {
      msgs[k].err[sizeof msgs[k].err - 1] = 0;
      team_fail("worker " + std::to_string(k + 1) + ": " + msgs[k].err);
      rc = -1;
    }
  }
  if (mine != 0) team_fail("rank 0: " + my_err);  // the ranks no longer agree on what exists
  return rc;
}

// ---- the caller's CSR matrix, checked on the host before anything is sent anywhere ------------
// (a block refused by one rank only would leave the others waiting in a collective)
static int validate_csr_host(const struct __mv_sparse* A, bool* full) {
  const long long n = A->size, nnz = A->nnz;
  const int* rp = A->row_ptr;
  if (n < 1 || nnz < 0 || rp[0] != 0 || rp[n] != nnz) {
    set_error("GPU team: malformed CSR (row_ptr[0] != 0 or row_ptr[size] != nnz)");
    return -1;
  }
  bool isfull = nnz == n * n;
  for (long long i = 0; i < n; ++i) {
    if (rp[i + 1] < rp[i]) {
      set_error("GPU team: malformed CSR (row_ptr decreases)");
      return -1;
    }
    if (isfull && (long long)rp[i] != i * n) isfull = false;
  }
  *full = isfull;
  if (isfull) return 0;
  const int* ci = A->col_indices;
  size_t nt = std::thread::hardware_concurrency();
  if (nt > 16) nt = 16;
  if (nt < 1 || nnz < (1 << 20)) nt = 1;
  std::vector<int> bad(nt, 0);
  auto work = [&](size_t t) {
    const long long k0 = nnz * (long long)t / (long long)nt, k1 = nnz * (long long)(t + 1) / (long long)nt;
    int b = 0;
    for (long long k = k0; k < k1; ++k) b |= (ci[k] < 0) | (ci[k] >= n);
    bad[t] = b;
  };
  std::vector<std::thread> th;
  for (size_t t = 1; t < nt; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  for (int b : bad)
    if (b) {
      set_error("GPU team: malformed CSR (column index out of range; indices are 0-based)");
      return -1;
    }
  return 0;
}

TeamSystem* team_create(const struct __mv_sparse* A) {
  if (!team_active()) return nullptr;
  Ctx& c = ctx();
  const int G = c.world;
  bool full = false;
  if (validate_csr_host(A, &full)) return nullptr;
  const long long n = A->size, nnz = A->nnz;
  const int* rp = A->row_ptr;
  if (n < G) {
    set_error("GPU team: fewer rows than GPUs");
    return nullptr;
  }
  TeamSystem* T = new TeamSystem();
  T->n = n;
  T->dense = full;
  T->row0.assign(G, 0);
  T->nrows.assign(G, 0);
  T->base.assign(G, nullptr);
  T->off_b.assign(G, 0);
  T->off_x.assign(G, 0);
  T->handle.assign(G, -1);
  {  // contiguous row blocks of equal stored entries (at least one row each)
    std::vector<long long> cut(G + 1, 0);
    cut[G] = n;
    for (int g = 1; g < G; ++g) {
      const long long target = nnz * g / G;
      long long lo = 0, hi = n;
      while (lo < hi) {
        const long long mid = (lo + hi) / 2;
        if ((long long)rp[mid] < target) lo = mid + 1;
        else hi = mid;
      }
      cut[g] = lo;
    }
    for (int g = 1; g < G; ++g) cut[g] = std::max(cut[g], cut[g - 1] + 1);
    for (int g = G - 1; g >= 1; --g) cut[g] = std::min(cut[g], cut[g + 1] - 1);
    for (int g = 0; g < G; ++g) {
      T->row0[g] = cut[g];
      T->nrows[g] = (int)(cut[g + 1] - cut[g]);
    }
  }
  // every worker's block goes onto its device from here: one allocation [vals | cols | rowptr | b | x],
  // padded as solver.cu pads its own (the bulk copies of the stream kernel may over-read)
  std::vector<TeamMsg> msgs(G - 1);
  bool ok = true;
  for (int g = 1; g < G && ok; ++g) {
    const long long r0 = T->row0[g];
    const int nr = T->nrows[g];
    const long long k0 = rp[r0], bn = (long long)rp[r0 + nr] - k0;
    TeamMsg& m = msgs[g - 1];
    clear_msg(m);
    m.cmd = CMD_CREATE;
    m.n_global = n;
    m.row0 = r0;
    m.nrows = nr;
    m.nnz = bn;
    m.dense = full ? 1 : 0;
    size_t off = 0;
    m.off_vals = off;
    off = align_up(off + ((size_t)bn + 64) * sizeof(double), 256);
    m.off_cols = off;
    if (!full) off = align_up(off + ((size_t)bn + 64) * sizeof(int), 256);
    m.off_rowptr = off;
    const size_t rpad = (size_t)nr + 1 + 1024 + 64;
    if (!full) off = align_up(off + rpad * sizeof(int), 256);
    m.off_b = off;
    off = align_up(off + ((size_t)nr + 16) * sizeof(double), 256);
    m.off_x = off;
    off = align_up(off + ((size_t)nr + 16) * sizeof(double), 256);
    m.bytes = align_up(off, 2u << 20);  // never a sub-allocation of a block shared with other buffers
    T->off_b[g] = m.off_b;
    T->off_x[g] = m.off_x;
    char* base = nullptr;
    ok = cudaSetDevice(g_workers[g - 1].device) == cudaSuccess && cudaMalloc(&base, m.bytes) == cudaSuccess &&
         cudaMemset(base, 0, m.bytes) == cudaSuccess;
    if (ok) {
      T->base[g] = base;
      ok = upload_host(base + m.off_vals, A->values + k0, sizeof(double) * bn, nullptr) == cudaSuccess;
      if (ok && !full) {
        ok = upload_host(base + m.off_cols, A->col_indices + k0, sizeof(int) * bn, nullptr) == cudaSuccess;
        std::vector<int> lrp(rpad, (int)bn);  // local offsets; the padding repeats the end offset
        for (int i = 0; i <= nr; ++i) lrp[i] = (int)((long long)rp[r0 + i] - k0);
        ok = ok && cudaMemcpy(base + m.off_rowptr, lrp.data(), sizeof(int) * rpad, cudaMemcpyHostToDevice) == cudaSuccess;
      }
      ok = ok && cudaDeviceSynchronize() == cudaSuccess && cudaIpcGetMemHandle(&m.mem, base) == cudaSuccess;
    }
  }
  const std::string cuda_err = ok ? "" : cudaGetErrorString(cudaGetLastError());
  cudaSetDevice(c.device);
  if (!ok) {
    set_error("GPU team: putting a block on its GPU failed: " + cuda_err);
    for (int g = 1; g < G; ++g)
      if (T->base[g]) {
        cudaSetDevice(g_workers[g - 1].device);
        cudaFree(T->base[g]);
      }
    cudaSetDevice(c.device);
    delete T;
    return nullptr;
  }
  // collective: every rank builds its system (plan, halo lists, peer mappings) at the same time
  const int rc = team_round(msgs, [&]() -> int {
    T->mine = system_from_csr(n, 0, T->nrows[0], rp, A->col_indices, A->values, nullptr);
    return T->mine ? 0 : -1;
  });
  if (rc != 0) {  // the team is marked broken: nothing of it is used, or released collectively, again
    delete T;
    return nullptr;
  }
  for (int g = 1; g < G; ++g) T->handle[g] = msgs[g - 1].handle;
  return T;
}

void team_destroy(TeamSystem* T) {
  if (!T) return;
  Ctx& c = ctx();
  if (team_active()) {
    std::vector<TeamMsg> msgs(g_workers.size());
    for (size_t k = 0; k < msgs.size(); ++k) {
      clear_msg(msgs[k]);
      msgs[k].cmd = CMD_FREE;
      msgs[k].handle = T->handle[k + 1];
    }
    team_round(msgs, [&]() -> int {
      cudaStreamSynchronize(c.stream);
      delete T->mine;  // collective with the workers' own delete (peer mappings are closed together)
      T->mine = nullptr;
      return 0;
    });
    for (size_t g = 1; g < T->base.size(); ++g)
      if (T->base[g]) {
        cudaSetDevice(g_workers[g - 1].device);
        cudaFree(T->base[g]);
      }
    cudaSetDevice(c.device);
  }
  // (a broken team's blocks go with its processes)
  delete T;
}

int team_solve(TeamSystem* T, const double* b_host, const double* x0_host, double* x_host,
               const SolveOpts& o, SolveResult* res) {
  if (!T || !team_active()) {
    set_error("GPU team: not running");
    return -1;
  }
  // what any rank would refuse on its own is refused HERE, before the request goes out: a rank that
  // backs out of a collective solve alone leaves the others waiting for it
  if ((o.variant != 0 && o.variant != 1) || ((o.flags & 3) != 0 && o.variant != 0) || (o.flags & ~3) != 0) {
    set_error("GPU team: unknown variant / flags, or CGB200_OPT_JACOBI / CGB200_OPT_X0 with the single-reduction variant");
    return -1;
  }
  if (((o.flags & 2) != 0) != (x0_host != nullptr)) {
    set_error("GPU team: CGB200_OPT_X0 and the initial guess must come together");
    return -1;
  }
  Ctx& c = ctx();
  const int G = c.world;
  bool ok = true;
  // every rank's slice of b (and of the guess) onto its GPU, with that GPU current (the staged copy
  // of a large slice from pageable memory runs on streams of the current device, solver.cu)
  for (int g = 1; g < G && ok; ++g) {
    char* base = (char*)T->base[g];
    const size_t bytes = sizeof(double) * (size_t)T->nrows[g];
    ok = cudaSetDevice(g_workers[g - 1].device) == cudaSuccess &&
         upload_host(base + T->off_b[g], b_host + T->row0[g], bytes, nullptr) == cudaSuccess;
    if (ok && x0_host) ok = upload_host(base + T->off_x[g], x0_host + T->row0[g], bytes, nullptr) == cudaSuccess;
    if (ok) ok = cudaStreamSynchronize(nullptr) == cudaSuccess;
  }
  cudaSetDevice(c.device);
  if (ok) ok = system_set_rhs(T->mine, b_host) == 0 && (!x0_host || system_set_x(T->mine, x0_host) == 0);
  if (!ok) {
    set_error(std::string("GPU team: copying the right-hand side in failed: ") + cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  std::vector<TeamMsg> msgs(G - 1);
  for (int g = 1; g < G; ++g) {
    clear_msg(msgs[g - 1]);
    msgs[g - 1].cmd = CMD_SOLVE;
    msgs[g - 1].handle = T->handle[g];
    msgs[g - 1].o = o;
  }
  if (team_round(msgs, [&]() -> int { return system_solve(T->mine, o, res); }) != 0) return -1;
  if (system_get_x(T->mine, x_host)) return -1;
  for (int g = 1; g < G; ++g) {
    const bool got = cudaSetDevice(g_workers[g - 1].device) == cudaSuccess &&
                     download_host(x_host + T->row0[g], (char*)T->base[g] + T->off_x[g],
                                   sizeof(double) * T->nrows[g], nullptr) == cudaSuccess &&
                     cudaStreamSynchronize(nullptr) == cudaSuccess;
    cudaSetDevice(c.device);
    if (!got) {
      set_error(std::string("GPU team: copying x out failed: ") + cudaGetErrorString(cudaGetLastError()));
      return -1;
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// The worker: rank r of the team, on its own GPU, serving the primary's requests until told to quit.
// argv: rank world device fd_in fd_out
// ---------------------------------------------------------------------------------------------
int worker_main(int argc, char** argv) {
  if (argc < 6) {
    fprintf(stderr, "cgb200_worker is started by libcgb200.so (CGB200_GPUS=N); it takes no user arguments\n");
    return 2;
  }
  const int rank = atoi(argv[1]), world = atoi(argv[2]), device = atoi(argv[3]);
  const int fin = atoi(argv[4]), fout = atoi(argv[5]);
  unsigned char id[128];
  if (read_all(fin, id, sizeof id)) return 1;
  TeamMsg m;
  auto reply = [&](bool ok) {
    m.ok = ok ? 1 : 0;
    if (!ok) snprintf(m.err, sizeof m.err, "%s", last_error());
    return write_all(fout, &m, sizeof m);
  };
  clear_msg(m);
  const bool up = ensure_init(device) == 0;
  if (reply(up) || !up) return 1;
  const bool joined = dist_init(rank, world, id) == 0;
  if (reply(joined) || !joined) return 1;
  struct Held {
    System* sys;
    void* base;
  };
  std::map<int, Held> held;
  int next_handle = 1;
  for (;;) {
    if (read_all(fin, &m, sizeof m)) _exit(0);  // the primary is gone: so are we (no collective tear-down)
    if (m.cmd == CMD_QUIT) break;
    if (m.cmd == CMD_CREATE) {
      void* base = nullptr;
      System* S = nullptr;
      if (cudaIpcOpenMemHandle(&base, m.mem, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        set_error(std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(cudaGetLastError()));
      } else {
        char* p = (char*)base;
        S = system_from_device(m.n_global, m.row0, m.nrows, m.nnz, m.dense != 0, (double*)(p + m.off_vals),
                               (int*)(p + m.off_cols), (int*)(p + m.off_rowptr), (double*)(p + m.off_b),
                               (double*)(p + m.off_x));
      }
      if (S) {
        m.handle = next_handle++;
        held[m.handle] = Held{S, base};
      }
      if (reply(S != nullptr)) break;
    } else if (m.cmd == CMD_SOLVE) {
      auto it = held.find(m.handle);
      bool ok = it != held.end();
      if (!ok) set_error("unknown system handle");
      else ok = system_solve(it->second.sys, m.o, &m.r) == 0;
      if (reply(ok)) break;
    } else if (m.cmd == CMD_FREE) {
      auto it = held.find(m.handle);
      if (it != held.end()) {
        cudaStreamSynchronize(ctx().stream);
        delete it->second.sys;
        cudaIpcCloseMemHandle(it->second.base);
        held.erase(it);
      }
      if (reply(true)) break;
    } else {
      set_error("unknown command");
      if (reply(false)) break;
    }
  }
  dist_finalize();
  shutdown();
  return 0;
}

}  // namespace cgb
