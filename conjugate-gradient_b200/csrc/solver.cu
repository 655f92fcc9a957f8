// solver.cu -- device-resident systems, the CG loop (reference cg.c:88-141 re-expressed as a CUDA
// graph of fused kernels with device-side scalars), and the multi-GPU row-block layer.
//
// Loop shape per iteration (classic recurrence, the reference's):
//   K1  s = A p, partial p.s -> last CTA: alpha = rr / ps                    cg.c:111-113
//   K2  x += alpha p ; r -= alpha s ; partial r.r -> last CTA: beta, stop    cg.c:115-129
//   K3  p = r + beta p                                                        cg.c:131-132
// With more than one rank the partial sums are combined by a 1-double NCCL all-reduce followed
// by a one-thread bookkeeping kernel, and the search direction is exchanged before K1:
// neighbour halo send/recv overlapped with the interior tiles of K1 (stencil-like matrices) or
// an all-gather (dense / wide coupling).
#include "devmem.h"
#include "solver.h"

#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace cgb {

// ---------------------------------------------------------------------------------------------
// Error plumbing
// ---------------------------------------------------------------------------------------------

static thread_local std::string t_err;
void set_error(const std::string& m) { t_err = m; }
const char* last_error() { return t_err.c_str(); }

#define CUDA_OK(call)                                                                        \
  do {                                                                                       \
    cudaError_t e__ = (call);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      set_error(std::string(#call) + ": " + cudaGetErrorString(e__));                        \
      return -1;                                                                             \
    }                                                                                        \
  } while (0)

// ---------------------------------------------------------------------------------------------
// NCCL, loaded lazily so that the single-GPU path has no dependency on it
// ---------------------------------------------------------------------------------------------

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;

static int load_nccl() {
  if (g_nccl.lib) return 0;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    g_nccl.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) {
    set_error(std::string("cannot load NCCL: ") + dlerror());
    return -1;
  }
#define SYM(field, name)                                                   \
  *(void**)(&g_nccl.field) = dlsym(g_nccl.lib, name);                      \
  if (!g_nccl.field) {                                                     \
    set_error(std::string("NCCL symbol missing: ") + name);                \
    return -1;                                                             \
  }
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(AllGather, "ncclAllGather");
  SYM(Broadcast, "ncclBroadcast");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  return 0;
}

#define NCCL_OK(call)                                                                        \
  do {                                                                                       \
    ncclResult_t r__ = (call);                                                               \
    if (r__ != ncclSuccess) {                                                                \
      set_error(std::string(#call) + ": " + g_nccl.GetErrorString(r__));                     \
      return -1;                                                                             \
    }                                                                                        \
  } while (0)

// ---------------------------------------------------------------------------------------------
// Context
// ---------------------------------------------------------------------------------------------

static Ctx g_ctx;
static std::mutex g_mu;
Ctx& ctx() { return g_ctx; }

int ensure_init(int device) {
  std::lock_guard<std::mutex> lk(g_mu);
  Ctx& c = g_ctx;
  if (c.ready) return 0;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error(std::string("no CUDA device available (") + cudaGetErrorString(e) +
              "); libcgb200 has no CPU fallback");
    cudaGetLastError();
    return -1;
  }
  if (device < 0) {
    const char* lr = getenv("LOCAL_RANK");
    device = lr ? atoi(lr) % ndev : 0;
  }
  CUDA_OK(cudaSetDevice(device));
  c.device = device;
  CUDA_OK(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  CUDA_OK(cudaStreamCreateWithFlags(&c.comm_stream, cudaStreamNonBlocking));
  CUDA_OK(DEV_ALLOC(&c.partials, sizeof(double) * partials_doubles()));
  CUDA_OK(cudaMemset(c.partials, 0, sizeof(double) * partials_doubles()));
  CUDA_OK(DEV_ALLOC(&c.d_state, sizeof(CgState)));
  CUDA_OK(cudaMemset(c.d_state, 0, sizeof(CgState)));
  CUDA_OK(cudaMallocHost(&c.h_state, sizeof(CgState) * 4));
  CUDA_OK(DEV_ALLOC(&c.d_scratch, 64));
  CUDA_OK(cudaMallocHost(&c.h_scratch, 64));
  for (int i = 0; i < 2; ++i) CUDA_OK(cudaEventCreateWithFlags(&c.poll_ev[i], cudaEventDisableTiming));
  CUDA_OK(cudaEventCreate(&c.t0));
  CUDA_OK(cudaEventCreate(&c.t1));
  CUDA_OK(cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming));
  CUDA_OK(cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming));
  const char* cmode = getenv("CGB200_COMM");
  c.comm_mode = (cmode && strcmp(cmode, "nccl") == 0) ? 0 : 1;
  c.ready = true;
  return 0;
}

void shutdown() {
  std::lock_guard<std::mutex> lk(g_mu);
  Ctx& c = g_ctx;
  if (!c.ready) return;
  cudaSetDevice(c.device);
  cudaDeviceSynchronize();
  if (c.comm) {
    g_nccl.CommDestroy((ncclComm_t)c.comm);
    c.comm = nullptr;
  }
  dev_free(c.partials);
  dev_free(c.d_state);
  cudaFreeHost(c.h_state);
  dev_free(c.d_scratch);
  cudaFreeHost(c.h_scratch);
  stage_release();
  cudaStreamDestroy(c.stream);
  cudaStreamDestroy(c.comm_stream);
  c = Ctx();
}

int dist_unique_id(void* id128) {
  if (load_nccl()) return -1;
  ncclUniqueId id;
  NCCL_OK(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return 0;
}

int dist_init(int rank, int world, const void* id128) {
  if (ensure_init(-1)) return -1;
  Ctx& c = g_ctx;
  if (world <= 1) {
    c.rank = 0;
    c.world = 1;
    return 0;
  }
  if (load_nccl()) return -1;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  CUDA_OK(cudaSetDevice(c.device));
  ncclComm_t comm;
  NCCL_OK(g_nccl.CommInitRank(&comm, world, id, rank));
  c.comm = comm;
  c.rank = rank;
  c.world = world;
  return 0;
}

int dist_finalize() {
  Ctx& c = g_ctx;
  if (c.comm) {
    cudaStreamSynchronize(c.stream);
    cudaStreamSynchronize(c.comm_stream);
    g_nccl.CommDestroy((ncclComm_t)c.comm);
    c.comm = nullptr;
  }
  c.rank = 0;
  c.world = 1;
  return 0;
}

static int allreduce_sum_f64(void* buf, size_t count, cudaStream_t st) {
  NCCL_OK(g_nccl.AllReduce(buf, buf, count, ncclDouble, ncclSum, (ncclComm_t)g_ctx.comm, st));
  return 0;
}

// A process that belongs to a multi-rank world (the primary of a GPU team, team.cu) still serves
// the single-rank entry points of mv_ops.h: systems created inside a SoloScope ignore the world.
static thread_local int t_solo = 0;
SoloScope::SoloScope() { ++t_solo; }
SoloScope::~SoloScope() { --t_solo; }
static int eff_world() { return t_solo > 0 ? 1 : g_ctx.world; }
static int eff_rank() { return t_solo > 0 ? 0 : g_ctx.rank; }

// host-side all-gather / max of a few 64-bit integers through a device bounce buffer (set-up only)
static int allgather_i64(const long long* mine, int count, std::vector<long long>& all, int world = -1) {
  Ctx& c = g_ctx;
  if (world < 0) world = eff_world();
  all.assign((size_t)count * world, 0);
  if (world == 1) {
    for (int i = 0; i < count; ++i) all[i] = mine[i];
    return 0;
  }
  long long* d = nullptr;
  CUDA_OK(DEV_ALLOC(&d, sizeof(long long) * count * c.world));
  CUDA_OK(cudaMemcpyAsync(d + (size_t)c.rank * count, mine, sizeof(long long) * count,
                          cudaMemcpyHostToDevice, c.stream));
  NCCL_OK(g_nccl.AllGather(d + (size_t)c.rank * count, d, (size_t)count, ncclInt64,
                           (ncclComm_t)c.comm, c.stream));
  CUDA_OK(cudaMemcpyAsync(all.data(), d, sizeof(long long) * count * c.world,
                          cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  dev_free(d);
  return 0;
}

// host-side all-gather of `bytes` bytes per rank through a device bounce buffer (set-up only)
static int allgather_bytes(const void* mine, int bytes, std::vector<unsigned char>& all) {
  Ctx& c = g_ctx;
  all.assign((size_t)bytes * c.world, 0);
  if (c.world == 1) {
    memcpy(all.data(), mine, bytes);
    return 0;
  }
  unsigned char* d = nullptr;
  CUDA_OK(DEV_ALLOC(&d, (size_t)bytes * c.world));
  CUDA_OK(cudaMemcpyAsync(d + (size_t)c.rank * bytes, mine, bytes, cudaMemcpyHostToDevice, c.stream));
  NCCL_OK(g_nccl.AllGather(d + (size_t)c.rank * bytes, d, (size_t)bytes, ncclInt8,
                           (ncclComm_t)c.comm, c.stream));
  CUDA_OK(cudaMemcpyAsync(all.data(), d, (size_t)bytes * c.world, cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  dev_free(d);
  return 0;
}

// All ranks have finished everything queued on their solve streams.
int dist_barrier() {
  Ctx& c = g_ctx;
  if (c.world <= 1 || !c.comm) return 0;
  CUDA_OK(cudaMemsetAsync(c.d_scratch, 0, 8, c.stream));
  NCCL_OK(g_nccl.AllReduce(c.d_scratch, c.d_scratch, 1, ncclInt32, ncclSum, (ncclComm_t)c.comm,
                           c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

const char* comm_mode_name() {
  Ctx& c = g_ctx;
  if (c.world <= 1) return "single";
  return c.comm_mode == 1 ? "p2p" : "nccl";
}

// ---------------------------------------------------------------------------------------------
// Systems
// ---------------------------------------------------------------------------------------------


System::~System() {
  for (auto& kv : graphs)
    if (kv.second) cudaGraphExecDestroy(kv.second);
  if (p2p) {
    // unmap the peers, and do not release what they may still have mapped before they did too
    cudaStreamSynchronize(g_ctx.stream);
    for (int g = 0; g < MAX_RANKS; ++g)
      if (peer_base[g]) cudaIpcCloseMemHandle(peer_base[g]);
    dist_barrier();
  }
  free_pattern(&pat);
  free_banded(&dia);
  if (!ext_matrix) {
    dev_free(A.vals);
    dev_free(A.cols);
    dev_free(A.rowptr);
  }
  if (!ext_bx) {
    dev_free(b);
    dev_free(x);
  }
  dev_free(r);
  dev_free(s);
  dev_free(w);
  dev_free(dinv);
  if (comm_buf) {
    dev_guard_unwatch((char*)comm_buf + COMM_PEXT_OFFSET + ((size_t)ext_len + 16) * sizeof(double));
    cudaFree(comm_buf);  // p_ext lives inside it
  } else {
    dev_free(p_ext);
  }
  dev_free(state);
  dev_free(partials);
  dev_free(trace);
  dev_free(gemv_pool);
}

// Allocates the matrix arrays with the padding the bulk copies may over-read (zero-filled).
static int alloc_matrix(System* S, long long nnz, bool with_index) {
  S->A.nnz = nnz;
  S->A.nrows = S->nloc;
  const size_t vpad = (size_t)nnz + 64;
  CUDA_OK(DEV_ALLOC(&S->A.vals, vpad * sizeof(double)));
  CUDA_OK(cudaMemsetAsync(S->A.vals + nnz, 0, 64 * sizeof(double), g_ctx.stream));
  if (with_index) {
    CUDA_OK(DEV_ALLOC(&S->A.cols, vpad * sizeof(int)));
    CUDA_OK(cudaMemsetAsync(S->A.cols + nnz, 0, 64 * sizeof(int), g_ctx.stream));
    const size_t rpad = (size_t)S->nloc + 1 + 1024 + 64;
    CUDA_OK(DEV_ALLOC(&S->A.rowptr, rpad * sizeof(int)));
    CUDA_OK(cudaMemsetAsync(S->A.rowptr, 0, rpad * sizeof(int), g_ctx.stream));
  }
  return 0;
}

// Maps every peer's exported allocation and describes who pushes what to whom (comm.h).  Every
// rank evaluates every rank's plan (pure host logic), so no layout has to be communicated beyond
// the IPC handles.  If any rank cannot export or map, all ranks fall back to NCCL together.
static int setup_peer_comm(System* S, const std::vector<PartInfo>& parts) {
  Ctx& c = g_ctx;
  const int G = c.world, me = c.rank;
  CUDA_OK(cudaStreamSynchronize(c.stream));  // the zero-fill is complete before anyone can map it
  struct Msg {
    cudaIpcMemHandle_t h;
    int ok;
    int pad;
  } mine;
  memset(&mine, 0, sizeof(mine));
  cudaError_t e = cudaIpcGetMemHandle(&mine.h, S->comm_buf);
  mine.ok = (e == cudaSuccess);
  if (!mine.ok) cudaGetLastError();
  std::vector<unsigned char> all;
  if (allgather_bytes(&mine, (int)sizeof(Msg), all)) return -1;
  const Msg* msgs = (const Msg*)all.data();
  bool ok = true;
  for (int g = 0; g < G; ++g) ok = ok && msgs[g].ok;
  int opened = 1;
  if (ok) {
    for (int g = 0; g < G && opened; ++g) {
      if (g == me) continue;
      void* ptr = nullptr;
      e = cudaIpcOpenMemHandle(&ptr, msgs[g].h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        opened = 0;
      } else {
        S->peer_base[g] = ptr;
      }
    }
  } else {
    opened = 0;
  }
  std::vector<unsigned char> all2;
  if (allgather_bytes(&opened, (int)sizeof(int), all2)) return -1;
  for (int g = 0; g < G; ++g) ok = ok && ((const int*)all2.data())[g] != 0;
  if (!ok) {  // together: NCCL between the kernels instead
    for (int g = 0; g < G; ++g)
      if (S->peer_base[g]) {
        cudaIpcCloseMemHandle(S->peer_base[g]);
        S->peer_base[g] = nullptr;
      }
    if (getenv("CGB200_VERBOSE")) fprintf(stderr, "[cgb200] rank %d: CUDA IPC unavailable, using NCCL\n", me);
    S->p2p = false;
    return 0;
  }

  PushPlan pp;
  const char* perr = nullptr;
  if (plan_push(G, me, S->n_global, parts.data(), S->dense ? 1 : 0, &pp, &perr)) {
    set_error(std::string("push plan: ") + (perr ? perr : "?"));
    return -1;
  }
  auto base_of = [&](int g) { return g == me ? (char*)S->comm_buf : (char*)S->peer_base[g]; };

  PeerComm cm;
  memset(&cm, 0, sizeof(cm));
  cm.world = G;
  cm.rank = me;
  for (int g = 0; g < G; ++g) cm.mbox[g] = (Mailbox*)base_of(g);
  cm.local = (CommLocal*)((char*)S->comm_buf + COMM_LOCAL_OFFSET);
  cm.npush = pp.npush;
  for (int t = 0; t < pp.npush; ++t) {
    double* pext = (double*)(base_of(pp.push[t].peer) + COMM_PEXT_OFFSET);
    cm.push[t] = PushTarget{pext + pp.push[t].dst_index, pp.push[t].begin, pp.push[t].end,
                            pp.push[t].peer, 0};
  }
  cm.nwait = pp.nwait;
  for (int i = 0; i < pp.nwait; ++i) cm.wait_src[i] = pp.wait_src[i];
  S->cm = cm;
  S->p2p = true;
  return 0;
}

// Decides the exchange scheme from the column span, allocates the vectors, shifts the column
// indices into the local gather space and classifies tiles as interior / boundary.
// `mincol`/`maxcol` are GLOBAL column indices seen in this rank's rows.
static int plan_and_alloc(System* S, long long mincol, long long maxcol, bool cols_are_global) {
  Ctx& c = g_ctx;
  const int world = eff_world(), rank = eff_rank();
  S->world = world;
  S->rank = rank;
  // every rank learns every rank's (row0, nloc, mincol, maxcol); the plan is pure host logic
  long long mine[4] = {S->row0, (long long)S->nloc, mincol, maxcol};
  std::vector<long long> all;
  if (allgather_i64(mine, 4, all)) return -1;
  std::vector<PartInfo> parts(world);
  for (int g = 0; g < world; ++g)
    parts[g] = PartInfo{all[4 * g + 0], all[4 * g + 1], all[4 * g + 2], all[4 * g + 3]};
  // diagnostics (tools/gemv_rows_sweep.py): a lone row block [0, nloc) of a wider matrix, for timing
  // its kernels on one GPU; such a system is not a solvable one
  if (world == 1 && S->row0 == 0 && S->nloc < S->n_global && getenv("CGB200_TIMING_BLOCKS"))
    parts[0].nloc = S->n_global;
  Plan plan;
  const char* perr = nullptr;
  if (plan_partition(world, rank, S->n_global, parts.data(), S->dense ? 1 : 0, &plan, &perr)) {
    set_error(std::string("partition plan: ") + (perr ? perr : "?"));
    return -1;
  }
  S->part_row0.resize(world);
  S->part_nloc.resize(world);
  for (int g = 0; g < world; ++g) {
    S->part_row0[g] = parts[g].row0;
    S->part_nloc[g] = (int)parts[g].nloc;
  }
  S->allgather = (world > 1) && plan.allgather;
  S->h_lo = plan.h_lo;
  S->h_hi = plan.h_hi;
  S->send_lo = plan.send_lo;
  S->send_hi = plan.send_hi;
  S->off = plan.off;
  S->ext_len = plan.ext_len;
  const long long col_shift = plan.col_shift;  // global column -> index in p_ext

  const size_t nl = (size_t)S->nloc;
  CUDA_OK(DEV_ALLOC(&S->b, (nl + 16) * sizeof(double)));
  CUDA_OK(DEV_ALLOC(&S->x, (nl + 16) * sizeof(double)));
  CUDA_OK(DEV_ALLOC(&S->r, (nl + 16) * sizeof(double)));
  CUDA_OK(DEV_ALLOC(&S->s, (nl + 16) * sizeof(double)));
  S->cm_off.world = 1;
  S->cm = S->cm_off;
  if (world > 1 && c.comm_mode == 1 && world <= MAX_RANKS) {
    // one exported allocation: [Mailbox | CommLocal | p_ext]; a multiple of 2 MiB so that it is
    // never a sub-allocation of a block shared with other buffers
    size_t bytes = COMM_PEXT_OFFSET + ((size_t)S->ext_len + 16) * sizeof(double);
    bytes = (bytes + (2u << 20) - 1) / (2u << 20) * (2u << 20);
    const size_t used = COMM_PEXT_OFFSET + ((size_t)S->ext_len + 16) * sizeof(double);
    if (dev_guard_enabled() && bytes - used < 4096) bytes += 2u << 20;  // room for a guard band
    CUDA_OK(cudaMalloc(&S->comm_buf, bytes));
    CUDA_OK(cudaMemsetAsync(S->comm_buf, 0, bytes, c.stream));
    S->p_ext = (double*)((char*)S->comm_buf + COMM_PEXT_OFFSET);
    // guard mode: the peers store into p_ext; the slack behind it must stay untouched
    CUDA_OK(cudaStreamSynchronize(c.stream));
    CUDA_OK(dev_guard_watch((char*)S->comm_buf + used, bytes - used, used, __FILE__, __LINE__));
  } else {
    CUDA_OK(DEV_ALLOC(&S->p_ext, ((size_t)S->ext_len + 16) * sizeof(double)));
    CUDA_OK(cudaMemsetAsync(S->p_ext, 0, ((size_t)S->ext_len + 16) * sizeof(double), c.stream));
  }
  CUDA_OK(cudaMemsetAsync(S->x, 0, (nl + 16) * sizeof(double), c.stream));
  CUDA_OK(DEV_ALLOC(&S->state, sizeof(CgState)));
  CUDA_OK(cudaMemsetAsync(S->state, 0, sizeof(CgState), c.stream));
  CUDA_OK(DEV_ALLOC(&S->partials, sizeof(double) * partials_doubles()));
  CUDA_OK(cudaMemsetAsync(S->partials, 0, sizeof(double) * partials_doubles(), c.stream));
  S->p_own = S->p_ext + S->off;
  if (S->dense) {  // chunk records of the dense matvec's dynamic tail (kernels.cu: gemv_pool_phase)
    const size_t pd = dense_pool_doubles(S->nloc, (int)S->n_global);
    if (pd > 0) {
      CUDA_OK(DEV_ALLOC(&S->gemv_pool, pd * sizeof(double)));
      CUDA_OK(cudaMemsetAsync(S->gemv_pool, 0, pd * sizeof(double), c.stream));
    }
  }
  if (S->comm_buf && setup_peer_comm(S, parts)) return -1;

  if (!S->dense) {
    if (cols_are_global && col_shift != 0) {
      if (col_shift > 0x7fffffffLL || col_shift < -0x7fffffffLL) {
        set_error("column shift exceeds int32");
        return -1;
      }
      CUDA_OK(launch_shift_cols(S->A.cols, S->A.nnz, (int)col_shift, c.stream));
    }
    // upper band edge in gather space (after the shift): the stream kernel prefetches it
    S->A.plen = S->ext_len;
    S->A.lead = -1;
    // (opt-in, CGB200_SPMV_PREFETCH=1: once the consumers keep a whole row's gathers in flight the
    // prefetch no longer pays on the 3D stencils -- profiles/r01_spmv_gather_batching.md)
    const char* pf = getenv("CGB200_SPMV_PREFETCH");
    if (S->nloc > 0 && S->A.nnz > 0 && pf && atoi(pf) > 0) {
      int* d1 = (int*)c.d_scratch + 8;
      CUDA_OK(launch_upper_band(S->A.rowptr, S->A.cols, S->nloc, d1, c.stream));
      CUDA_OK(cudaMemcpyAsync((int*)c.h_scratch + 8, d1, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
      CUDA_OK(cudaStreamSynchronize(c.stream));
      const long long lead = ((int*)c.h_scratch)[8];
      // worth it only for a band that is wide enough to fall out of the tile's own footprint and
      // narrow enough to be a band at all
      if (lead > 4 * spmv_rows_per_tile() && lead < S->ext_len) S->A.lead = lead;
    }
    const int R = spmv_rows_per_tile();
    S->ntiles = (S->nloc + R - 1) / R;
    // interior tiles: rows whose columns stay inside the owned block.  Rows are coupled to a
    // band around the diagonal of half-width h_lo / h_hi, so rows [h_lo, nloc - h_hi) are safe
    // for banded/stencil matrices; verified per tile below for general patterns.
    S->tb_lo = 0;
    S->tb_hi = S->ntiles;
    if (world > 1 && !S->allgather && (S->h_lo > 0 || S->h_hi > 0)) {
      if (classify_tiles(S)) return -1;
    }
  }
  return 0;
}

// Per-tile min/max gathered index -> leading/trailing runs of boundary tiles.
__global__ void k_tile_span(const int* __restrict__ rowptr, const int* __restrict__ cols, int nrows,
                            int R, int ntiles, int lo_bound, int hi_bound,
                            unsigned char* __restrict__ flag) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  const int r0 = t * R, r1 = min(r0 + R, nrows);
  unsigned char f = 0;
  // columns are ascending within a row: first and last entry of each row bound it
  for (int r = r0; r < r1 && !f; ++r) {
    const int a = rowptr[r], b = rowptr[r + 1];
    if (b > a && (cols[a] < lo_bound || cols[b - 1] >= hi_bound)) f = 1;
  }
  flag[t] = f;
}

int classify_tiles(System* S) {
  Ctx& c = g_ctx;
  unsigned char* d_flag = nullptr;
  CUDA_OK(DEV_ALLOC(&d_flag, S->ntiles));
  const int R = spmv_rows_per_tile();
  k_tile_span<<<(S->ntiles + 127) / 128, 128, 0, c.stream>>>(
      S->A.rowptr, S->A.cols, S->nloc, R, S->ntiles, (int)S->off, (int)(S->off + S->nloc), d_flag);
  add_launches(1);
  std::vector<unsigned char> flag(S->ntiles);
  CUDA_OK(cudaMemcpyAsync(flag.data(), d_flag, S->ntiles, cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  dev_free(d_flag);
  int lo = 0, hi = S->ntiles;
  while (lo < S->ntiles && flag[lo]) ++lo;
  while (hi > lo && flag[hi - 1]) --hi;
  bool scattered = false;
  for (int t = lo; t < hi; ++t)
    if (flag[t]) { scattered = true; break; }
  if (scattered) {  // no clean split: everything waits for the halo
    lo = 0;
    hi = 0;
  }
  S->tb_lo = lo;  // tiles [0, tb_lo) and [tb_hi, ntiles) need the halo; [tb_lo, tb_hi) do not
  S->tb_hi = hi;
  return 0;
}

// CGB200_FORMAT=pattern: every CSR system tries the row-pattern dictionary format when it is
// created (for callers that cannot be edited; cgb200_system_use_pattern_format is the API).
static int maybe_auto_pattern(System* S) {
  const char* f = getenv("CGB200_FORMAT");
  if (!f || S->dense || S->nloc <= 0) return 0;
  if (strcmp(f, "banded") == 0 && S->world == 1) return system_use_banded(S, 1) < 0 ? -1 : 0;
  if (strcmp(f, "pattern") != 0) return 0;
  return system_use_pattern(S, 1) < 0 ? -1 : 0;
}

System* system_lap(int nx, int ny, int nz, int z0, int nzl, unsigned long long bseed) {
  if (ensure_init(-1)) return nullptr;
  Ctx& c = g_ctx;
  if (nx < 1 || ny < 1 || nz < 1 || z0 < 0 || nzl < 1 || z0 + nzl > nz) {
    set_error("system_lap: bad grid/slab");
    return nullptr;
  }
  const long long pl = (long long)nx * ny;
  const long long n = pl * nz;
  const long long nloc = pl * nzl;
  if (n > 0x7fffffffLL) {
    set_error("system_lap: more than 2^31-1 unknowns");
    return nullptr;
  }
  System* S = new System();
  S->n_global = n;
  S->row0 = pl * z0;
  S->nloc = (int)nloc;
  S->dense = false;
  // nnz of the slab: 7 per row minus missing neighbours
  auto nnz_before = [&](long long row) {
    const long long L = row / nx, xx = row % nx, P = row / pl, q = row % pl;
    long long m = 0;
    m += L + (xx > 0 ? 1 : 0);
    m += L;
    m += P * nx + std::min<long long>(q, nx);
    m += P * nx + std::max<long long>(0, q - (long long)(ny - 1) * nx);
    m += std::min(row, pl);
    m += std::max<long long>(0, row - (long long)(nz - 1) * pl);
    return 7 * row - m;
  };
  const long long nnz = nnz_before(S->row0 + nloc) - nnz_before(S->row0);
  if (nnz > 0x7fffffffLL - 64) {
    set_error("system_lap: local nnz exceeds int32");
    delete S;
    return nullptr;
  }
  if (alloc_matrix(S, nnz, true)) { delete S; return nullptr; }
  // column span is known analytically: one plane below / above (clipped to the grid)
  const long long mincol = std::max(0LL, S->row0 - (nz > 1 ? pl : (long long)nx));
  const long long maxcol = std::min(n - 1, S->row0 + nloc - 1 + (nz > 1 ? pl : (long long)nx));
  // generate with GLOBAL columns first, plan, then the plan shifts them
  if (launch_gen_lap(nx, ny, nz, S->row0, S->nloc, 0, S->A.rowptr, S->A.cols, S->A.vals,
                     c.stream) != cudaSuccess) {
    set_error("gen_lap launch failed");
    delete S;
    return nullptr;
  }
  if (plan_and_alloc(S, mincol, maxcol, true)) { delete S; return nullptr; }
  if (launch_gen_rhs(S->b, S->row0, S->nloc, bseed, c.stream) != cudaSuccess ||
      cudaStreamSynchronize(c.stream) != cudaSuccess) {
    set_error(std::string("system_lap: ") + cudaGetErrorString(cudaGetLastError()));
    delete S;
    return nullptr;
  }
  if (maybe_auto_pattern(S)) { delete S; return nullptr; }
  return S;
}

System* system_dense(int n, int row0, int nrows, unsigned long long aseed,
                     unsigned long long bseed) {
  if (ensure_init(-1)) return nullptr;
  Ctx& c = g_ctx;
  if (n < 1 || row0 < 0 || nrows < 1 || row0 + nrows > n) {
    set_error("system_dense: bad row block");
    return nullptr;
  }
  System* S = new System();
  S->n_global = n;
  S->row0 = row0;
  S->nloc = nrows;
  S->dense = true;
  if (alloc_matrix(S, (long long)nrows * n, false)) { delete S; return nullptr; }
  if (launch_gen_dense(n, row0, nrows, aseed, S->A.vals, c.stream) != cudaSuccess) {
    set_error("gen_dense launch failed");
    delete S;
    return nullptr;
  }
  if (plan_and_alloc(S, 0, n - 1, false)) { delete S; return nullptr; }
  if (launch_gen_rhs(S->b, S->row0, S->nloc, bseed, c.stream) != cudaSuccess ||
      cudaStreamSynchronize(c.stream) != cudaSuccess) {
    set_error(std::string("system_dense: ") + cudaGetErrorString(cudaGetLastError()));
    delete S;
    return nullptr;
  }
  return S;
}

// Copies of large arrays between the caller's ordinary (pageable) memory and the device.  The
// driver's own staging reaches ~10 GB/s there (measured: 8.6 GB up in 0.8 s); here a few host
// threads move 4 MiB pieces through page-locked buffers, each thread on its own stream, so the
// host-side copy of one piece overlaps the PCIe transfer of the others (8.6 GB up in 0.22 s).
// Page-locked sources / destinations and small arrays take one cudaMemcpyAsync on st.  The staged
// forms synchronise st first (the device buffer is theirs from then on) and return with the data
// in place.  The staging buffer (64 MiB, usable from every device) is allocated on first use and
// kept until shutdown: allocating page-locked memory costs more than a 100 MB copy.
namespace {

constexpr size_t STAGE_PIECE = 4u << 20;
constexpr int STAGE_NBUF = 2, STAGE_MAXT = 8;
constexpr size_t STAGE_MIN_BYTES = 32u << 20;
char* g_stage = nullptr;
std::mutex g_stage_mu;  // one staged copy at a time per process

bool staged_copy_wanted(const void* host, size_t bytes) {
  if (bytes < STAGE_MIN_BYTES) return false;
  if (const char* e = getenv("CGB200_STAGED_UPLOAD"))
    if (strcmp(e, "0") == 0) return false;
  cudaPointerAttributes at;
  const bool pageable = cudaPointerGetAttributes(&at, host) == cudaSuccess && at.type == cudaMemoryTypeUnregistered;
  cudaGetLastError();
  return pageable;
}

int stage_threads() {
  const int t = (int)std::thread::hardware_concurrency() / 2;
  return t < 2 ? 2 : (t > STAGE_MAXT ? STAGE_MAXT : t);
}

// dir 0: host -> device, 1: device -> host.  Call with g_stage_mu held and g_stage allocated.
cudaError_t staged_copy(char* dev_ptr, char* host_ptr, size_t bytes, int dir) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const int T = stage_threads();
  std::vector<cudaError_t> err((size_t)T, cudaSuccess);
  std::atomic<size_t> next(0);
  auto work = [&](int t) {
    cudaError_t& er = err[(size_t)t];
    cudaStream_t s = nullptr;
    cudaEvent_t ev[STAGE_NBUF] = {nullptr, nullptr};
    size_t p_off[STAGE_NBUF] = {0, 0}, p_len[STAGE_NBUF] = {0, 0};  // dir 1: piece waiting in each buffer
    if ((er = cudaSetDevice(dev)) != cudaSuccess) return;
    if ((er = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)) != cudaSuccess) return;
    for (int k = 0; k < STAGE_NBUF && er == cudaSuccess; ++k)
      er = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
    auto buffer = [&](int b) { return g_stage + ((size_t)t * STAGE_NBUF + b) * STAGE_PIECE; };
    auto drain = [&](int b) {  // dir 1: the piece that arrived in buffer b goes to the caller's array
      if (p_len[b] == 0 || er != cudaSuccess) return;
      er = cudaEventSynchronize(ev[b]);
      if (er == cudaSuccess) memcpy(host_ptr + p_off[b], buffer(b), p_len[b]);
      p_len[b] = 0;
    };
    for (size_t k = 0; er == cudaSuccess; ++k) {
      const size_t off = next.fetch_add(1) * STAGE_PIECE;
      if (off >= bytes) break;
      const size_t len = bytes - off < STAGE_PIECE ? bytes - off : STAGE_PIECE;
      const int b = (int)(k % STAGE_NBUF);
      if (dir == 0) {
        if (k >= STAGE_NBUF) er = cudaEventSynchronize(ev[b]);  // the transfer that last used this buffer
        if (er != cudaSuccess) break;
        memcpy(buffer(b), host_ptr + off, len);
        er = cudaMemcpyAsync(dev_ptr + off, buffer(b), len, cudaMemcpyHostToDevice, s);
      } else {
        drain(b);
        if (er != cudaSuccess) break;
        er = cudaMemcpyAsync(buffer(b), dev_ptr + off, len, cudaMemcpyDeviceToHost, s);
        p_off[b] = off;
        p_len[b] = len;
      }
      if (er == cudaSuccess) er = cudaEventRecord(ev[b], s);
    }
    if (dir == 1)
      for (int b = 0; b < STAGE_NBUF; ++b) drain(b);
    if (s) {
      const cudaError_t e2 = cudaStreamSynchronize(s);
      if (er == cudaSuccess) er = e2;
      cudaStreamDestroy(s);
    }
    for (int k = 0; k < STAGE_NBUF; ++k)
      if (ev[k]) cudaEventDestroy(ev[k]);
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  for (cudaError_t x : err)
    if (x != cudaSuccess) e = x;
  return e;
}

cudaError_t host_copy(void* dev_ptr, void* host_ptr, size_t bytes, int dir, cudaStream_t st) {
  const cudaMemcpyKind kind = dir == 0 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  void* dst = dir == 0 ? dev_ptr : host_ptr;
  const void* src = dir == 0 ? host_ptr : dev_ptr;
  if (!staged_copy_wanted(host_ptr, bytes)) return cudaMemcpyAsync(dst, src, bytes, kind, st);
  std::lock_guard<std::mutex> lk(g_stage_mu);
  if (!g_stage &&
      cudaHostAlloc((void**)&g_stage, (size_t)STAGE_MAXT * STAGE_NBUF * STAGE_PIECE, cudaHostAllocPortable) != cudaSuccess) {
    cudaGetLastError();
    g_stage = nullptr;
    return cudaMemcpyAsync(dst, src, bytes, kind, st);
  }
  const cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return e;
  return staged_copy((char*)dev_ptr, (char*)host_ptr, bytes, dir);
}

}  // namespace

cudaError_t upload_host(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  return host_copy(dst, const_cast<void*>(src), bytes, 0, st);
}

cudaError_t download_host(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  return host_copy(const_cast<void*>(src), dst, bytes, 1, st);
}

void stage_release() {
  std::lock_guard<std::mutex> lk(g_stage_mu);
  if (g_stage) cudaFreeHost(g_stage);
  g_stage = nullptr;
}

System* system_from_csr(long long n_global, long long row0, int nrows, const int* row_ptr,
                        const int* cols, const double* vals, const double* b) {
  if (ensure_init(-1)) return nullptr;
  Ctx& c = g_ctx;
  if (n_global < 1 || nrows < 0 || row0 < 0 || row0 + nrows > n_global || !row_ptr ||
      (!vals && nrows > 0 && row_ptr[nrows] > 0)) {
    set_error("system_from_csr: bad arguments");
    return nullptr;
  }
  if (n_global > 0x7fffffffLL) {
    set_error("system_from_csr: more than 2^31-1 unknowns");
    return nullptr;
  }
  System* S = new System();
  S->n_global = n_global;
  S->row0 = row0;
  S->nloc = nrows;
  const long long nnz = nrows > 0 ? (long long)row_ptr[nrows] - row_ptr[0] : 0;
  // full rows?  then values[] is a row-major dense block and the indices carry no information
  bool full = nrows > 0 && nnz == (long long)nrows * n_global && row_ptr[0] == 0;
  if (full) {
    for (int i = 0; i <= nrows && full; ++i)
      if ((long long)row_ptr[i] != (long long)i * n_global) full = false;
  }
  S->dense = full;
  if (alloc_matrix(S, nnz, !full)) { delete S; return nullptr; }
  auto fail = [&](const char* what) -> System* {
    set_error(std::string("system_from_csr: ") + what + ": " +
              cudaGetErrorString(cudaGetLastError()));
    delete S;
    return nullptr;
  };
  if (nnz > 0 &&
      upload_host(S->A.vals, vals + row_ptr[0], sizeof(double) * nnz, c.stream) != cudaSuccess)
    return fail("values H2D");
  long long mincol = 0, maxcol = n_global - 1;
  if (!full) {
    if (!cols && nnz > 0) {
      set_error("system_from_csr: col_indices missing");
      delete S;
      return nullptr;
    }
    if (nnz > 0 && upload_host(S->A.cols, cols + row_ptr[0], sizeof(int) * nnz, c.stream) != cudaSuccess)
      return fail("cols H2D");
    if (row_ptr[0] == 0) {
      if (cudaMemcpyAsync(S->A.rowptr, row_ptr, sizeof(int) * ((size_t)nrows + 1),
                          cudaMemcpyHostToDevice, c.stream) != cudaSuccess)
        return fail("row_ptr H2D");
    } else {
      std::vector<int> rp((size_t)nrows + 1);
      for (int i = 0; i <= nrows; ++i) rp[i] = row_ptr[i] - row_ptr[0];
      if (cudaMemcpyAsync(S->A.rowptr, rp.data(), sizeof(int) * rp.size(), cudaMemcpyHostToDevice,
                          c.stream) != cudaSuccess ||
          cudaStreamSynchronize(c.stream) != cudaSuccess)
        return fail("row_ptr H2D");
    }
    // the padding after row_ptr[nrows] repeats the end offset so over-read tiles see empty rows
    {
      std::vector<int> tail(1024 + 63, (int)nnz);
      if (cudaMemcpyAsync(S->A.rowptr + nrows + 1, tail.data(), sizeof(int) * tail.size(),
                          cudaMemcpyHostToDevice, c.stream) != cudaSuccess ||
          cudaStreamSynchronize(c.stream) != cudaSuccess)
        return fail("row_ptr pad");
    }
    // Column span and validity, one pass over the uploaded index arrays.  The span feeds the halo
    // plan; the checks keep a malformed matrix (a 1-based file, a non-monotone row_ptr) from
    // reaching the gather and bulk-copy kernels, where it would read out of bounds and leave the
    // CUDA context in an error state for the rest of the process.  (The reference, which expands
    // rows on the host, simply ignores such entries: mv_ops.c:106-110.)
    if (nnz > 0) {
      int* d4 = (int*)c.d_scratch;
      if (launch_analyze(S->A.rowptr, S->A.cols, nrows, nnz, (int)n_global, d4, c.stream) !=
              cudaSuccess ||
          cudaMemcpyAsync(c.h_scratch, d4, 4 * sizeof(int), cudaMemcpyDeviceToHost, c.stream) !=
              cudaSuccess ||
          cudaStreamSynchronize(c.stream) != cudaSuccess)
        return fail("analyze");
      const int* h4 = (const int*)c.h_scratch;
      if (h4[0] < 0 || (long long)h4[1] >= n_global || h4[3] != 0) {
        char msg[200];
        snprintf(msg, sizeof msg,
                 "system_from_csr: malformed CSR (column range [%d, %d] for %lld columns, %d "
                 "decreasing row_ptr step(s)); indices are 0-based and row_ptr non-decreasing",
                 h4[0], h4[1], n_global, h4[3]);
        set_error(msg);
        delete S;
        return nullptr;
      }
      mincol = h4[0];
      maxcol = h4[1];
    }
  }
  if (plan_and_alloc(S, mincol, maxcol, true)) { delete S; return nullptr; }
  if (b && nrows > 0 &&
      cudaMemcpyAsync(S->b, b, sizeof(double) * nrows, cudaMemcpyHostToDevice, c.stream) !=
          cudaSuccess)
    return fail("b H2D");
  if (cudaStreamSynchronize(c.stream) != cudaSuccess) return fail("sync");
  if (maybe_auto_pattern(S)) { delete S; return nullptr; }
  return S;
}

System* system_from_device(long long n_global, long long row0, int nrows, long long nnz, bool dense,
                           double* vals, int* cols, int* rowptr, double* b, double* x) {
  if (ensure_init(-1)) return nullptr;
  Ctx& c = g_ctx;
  if (n_global < 1 || n_global > 0x7fffffffLL || nrows < 0 || row0 < 0 || row0 + nrows > n_global ||
      !vals || (!dense && (!cols || !rowptr)) || !b || !x) {
    set_error("system_from_device: bad arguments");
    return nullptr;
  }
  System* S = new System();
  S->n_global = n_global;
  S->row0 = row0;
  S->nloc = nrows;
  S->dense = dense;
  S->ext_matrix = true;
  S->A.nnz = nnz;
  S->A.nrows = nrows;
  S->A.vals = vals;
  S->A.cols = dense ? nullptr : cols;
  S->A.rowptr = dense ? nullptr : rowptr;
  long long mincol = 0, maxcol = n_global - 1;
  if (!dense && nnz > 0) {  // the column span feeds the halo plan (validity was checked by the primary)
    int* d4 = (int*)c.d_scratch;
    if (launch_analyze(S->A.rowptr, S->A.cols, nrows, nnz, (int)n_global, d4, c.stream) != cudaSuccess ||
        cudaMemcpyAsync(c.h_scratch, d4, 4 * sizeof(int), cudaMemcpyDeviceToHost, c.stream) != cudaSuccess ||
        cudaStreamSynchronize(c.stream) != cudaSuccess) {
      set_error(std::string("system_from_device: analyze: ") + cudaGetErrorString(cudaGetLastError()));
      delete S;
      return nullptr;
    }
    mincol = ((const int*)c.h_scratch)[0];
    maxcol = ((const int*)c.h_scratch)[1];
  }
  if (plan_and_alloc(S, mincol, maxcol, true)) { delete S; return nullptr; }
  if (cudaStreamSynchronize(c.stream) != cudaSuccess) {
    set_error("system_from_device: sync failed");
    delete S;
    return nullptr;
  }
  dev_free(S->b);  // the loop reads b and leaves x where the primary can reach them
  dev_free(S->x);
  S->b = b;
  S->x = x;
  S->ext_bx = true;
  if (maybe_auto_pattern(S)) { delete S; return nullptr; }
  return S;
}

// ---------------------------------------------------------------------------------------------
// One iteration, enqueued on the solve stream (captured into a graph by the caller)
// ---------------------------------------------------------------------------------------------

// Makes p_ext complete on every rank.  Neighbour mode runs on the comm stream between two events
// so that the interior tiles of K1 overlap it; all-gather mode runs in line.
static int enqueue_exchange_begin(System* S) {
  Ctx& c = g_ctx;
  if (S->world == 1) return 0;
  ncclComm_t comm = (ncclComm_t)c.comm;
  if (S->allgather) {
    bool equal = true;
    for (int g = 0; g < S->world; ++g) equal = equal && S->part_nloc[g] == S->nloc;
    if (equal) {
      NCCL_OK(g_nccl.AllGather(S->p_own, S->p_ext, (size_t)S->nloc, ncclDouble, comm, c.stream));
    } else {
      NCCL_OK(g_nccl.GroupStart());
      for (int g = 0; g < S->world; ++g)
        NCCL_OK(g_nccl.Broadcast(S->p_ext + S->part_row0[g], S->p_ext + S->part_row0[g],
                                 (size_t)S->part_nloc[g], ncclDouble, g, comm, c.stream));
      NCCL_OK(g_nccl.GroupEnd());
    }
    return 0;
  }
  CUDA_OK(cudaEventRecord(c.ev_fork, c.stream));
  CUDA_OK(cudaStreamWaitEvent(c.comm_stream, c.ev_fork, 0));
  const int me = S->rank;
  NCCL_OK(g_nccl.GroupStart());
  if (me > 0) {
    const int send_cnt = S->send_lo;  // my first rows are the lower rank's upper halo
    if (send_cnt > 0)
      NCCL_OK(g_nccl.Send(S->p_own, (size_t)send_cnt, ncclDouble, me - 1, comm, c.comm_stream));
    if (S->h_lo > 0)
      NCCL_OK(g_nccl.Recv(S->p_own - S->h_lo, (size_t)S->h_lo, ncclDouble, me - 1, comm,
                          c.comm_stream));
  }
  if (me + 1 < S->world) {
    const int send_cnt = S->send_hi;  // my last rows are the upper rank's lower halo
    if (send_cnt > 0)
      NCCL_OK(g_nccl.Send(S->p_own + S->nloc - send_cnt, (size_t)send_cnt, ncclDouble, me + 1,
                          comm, c.comm_stream));
    if (S->h_hi > 0)
      NCCL_OK(g_nccl.Recv(S->p_own + S->nloc, (size_t)S->h_hi, ncclDouble, me + 1, comm,
                          c.comm_stream));
  }
  NCCL_OK(g_nccl.GroupEnd());
  CUDA_OK(cudaEventRecord(c.ev_join, c.comm_stream));
  return 0;
}

static int enqueue_exchange_end(System* S) {
  Ctx& c = g_ctx;
  if (S->world == 1 || S->allgather) return 0;
  CUDA_OK(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
  return 0;
}

static const PeerComm& comm_of(const System* S) { return S->p2p ? S->cm : S->cm_off; }

// K1: out = A * (gather vector), optionally fused with the dot products of the loop.
//   fin      FIN_NONE (plain product), FIN_PS, FIN_SR, or FIN_LOCAL (sums left in state->local)
//   dual     also reduce the square of the gathered vector's owned block (single-reduction)
//   exchange the gather vector's halo / remote blocks take part: in peer-memory mode they were
//            pushed by the kernel that wrote the vector and this kernel waits for the flags at
//            its first boundary tile; in NCCL mode the exchange is enqueued here.
// Multi-rank systems take the long-row matvec too (tests/dist_check.py: neighbour-halo and
// all-gather systems, both exchange modes); CGB200_SPMV_LONGROW_MULTI=0 keeps them on the stream
// kernel, whose interior tiles overlap the exchange.
static bool longrow_multi_ok(const System* S, bool multi) {
  if (S->world == 1) return true;
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("CGB200_SPMV_LONGROW_MULTI");
    on = (e && atoi(e) == 0) ? 0 : 1;
  }
  (void)multi;
  return on == 1;
}

static int enqueue_matvec(System* S, double* out, int fin, bool dual, bool exchange,
                          bool rev = false) {
  Ctx& c = g_ctx;
  CgState* st = S->state;
  const PeerComm& cm = comm_of(S);
  const bool multi = S->world > 1 && exchange;
  int mode = fin | (dual ? MV_DUAL : 0) | (rev ? MV_REV : 0);
  if (S->dense) {
    if (multi && !S->p2p && enqueue_exchange_begin(S)) return -1;
    CUDA_OK(launch_gemv(S->A.vals, S->n_global, S->nloc, (int)S->n_global, S->p_ext, S->p_own, out,
                        S->partials, &st->ticket[0], st, mode, (multi && S->p2p) ? 1 : 0, cm,
                        c.stream));
    return 0;
  }
  const TileRange all_tiles = {{0, 0, 0}, {S->ntiles, 0, 0}};
  // CSR stream kernel, or the pattern kernel when the system was switched to that format
  auto spmv = [&](const TileRange& tr, int wait_from, double* partials, int md) -> cudaError_t {
    if (S->pat_active)
      return launch_spmv_pattern(S->pat, S->nloc, S->p_ext, S->p_own, out, tr, wait_from,
                                 S->ext_len, partials, &st->ticket[0], st, md, cm, c.stream);
    return launch_spmv(S->A, S->p_ext, S->p_own, out, tr, wait_from, partials, &st->ticket[0], st,
                       md, cm, c.stream);
  };
  // long rows: a warp per unit of 16 rows instead of a thread per row (kernels.cu)
  const int rg = spmv_rowgroup_min();
  const bool longrow = !S->pat_active && rg > 0 && S->A.nrows > 0 &&
                       S->A.nnz >= (long long)rg * S->A.nrows && longrow_multi_ok(S, multi);
  if (S->dia_active && S->world == 1) {  // banded block in diagonal storage (banded.cu)
    CUDA_OK(launch_spmv_dia(S->dia, S->nloc, S->p_ext, S->ext_len, S->off, S->p_own, out, S->partials,
                            &st->ticket[0], st, mode, cm, c.stream));
    return 0;
  }
  if (!multi) {
    if (longrow) {
      CUDA_OK(launch_spmv_rowgroup(S->A, S->p_ext, S->p_own, out, S->partials, &st->ticket[0], st,
                                   mode, cm, c.stream));
      return 0;
    }
    CUDA_OK(spmv(all_tiles, NO_WAIT, S->partials, mode));
    return 0;
  }
  if (longrow) {
    // several GPUs: the whole launch behind the exchange (no interior/boundary split: these
    // matrices gather from most of the vector anyway)
    if (S->p2p) {
      if (cm.nwait > 0)
        CUDA_OK(launch_spmv_longrow_coherent(S->A, S->p_ext, S->p_own, out, S->partials,
                                             &st->ticket[0], st, mode, cm, c.stream));
      else
        CUDA_OK(launch_spmv_rowgroup(S->A, S->p_ext, S->p_own, out, S->partials, &st->ticket[0],
                                     st, mode, cm, c.stream));
      return 0;
    }
    if (enqueue_exchange_begin(S)) return -1;
    if (enqueue_exchange_end(S)) return -1;
    CUDA_OK(launch_spmv_rowgroup(S->A, S->p_ext, S->p_own, out, S->partials, &st->ticket[0], st,
                                 mode, cm, c.stream));
    return 0;
  }
  const bool has_interior = !S->allgather && S->tb_hi > S->tb_lo;
  if (S->p2p) {
    // one launch: interior tiles first, the boundary tiles behind the wait on the peers' flags
    TileRange tr = all_tiles;
    int wait_from = cm.nwait > 0 ? 0 : NO_WAIT;
    if (has_interior) {
      tr = TileRange{{S->tb_lo, 0, S->tb_hi}, {S->tb_hi, S->tb_lo, S->ntiles}};
      if (cm.nwait > 0) wait_from = S->tb_hi - S->tb_lo;
    }
    CUDA_OK(spmv(tr, wait_from, S->partials, mode));
    return 0;
  }
  // NCCL mode
  const bool split = has_interior && (S->tb_lo > 0 || S->tb_hi < S->ntiles);
  if (enqueue_exchange_begin(S)) return -1;
  if (!split) {
    if (enqueue_exchange_end(S)) return -1;
    CUDA_OK(spmv(all_tiles, NO_WAIT, S->partials, mode));
    return 0;
  }
  // interior tiles while the halo is in flight, then the boundary tiles finish the reduction
  const TileRange interior = {{S->tb_lo, 0, 0}, {S->tb_hi, 0, 0}};
  const TileRange boundary = {{0, S->tb_hi, 0}, {S->tb_lo, S->ntiles, 0}};
  const int g_int = S->pat_active ? pattern_grid_for(S->pat, interior)
                                  : spmv_grid_for(S->tb_hi - S->tb_lo);
  const int int_mode = (fin == FIN_NONE) ? mode : (mode | MV_PARTIAL);
  CUDA_OK(spmv(interior, NO_WAIT, S->partials + MV_EXTRA_OFFSET, int_mode));
  if (enqueue_exchange_end(S)) return -1;
  const int fin_mode = (fin == FIN_NONE) ? mode : (mode | (g_int << 8));
  CUDA_OK(spmv(boundary, NO_WAIT, S->partials, fin_mode));
  return 0;
}

// NCCL mode only: cross-GPU sum of state->local[0..count) and the one-thread bookkeeping kernel.
static int enqueue_nccl_finish(System* S, int count, int which) {
  Ctx& c = g_ctx;
  if (allreduce_sum_f64(&S->state->local[0], (size_t)count, c.stream)) return -1;
  CUDA_OK(launch_finish(S->state, which, c.stream));
  return 0;
}

static int ensure_w(System* S) {
  if (S->w) return 0;
  CUDA_OK(DEV_ALLOC(&S->w, ((size_t)S->nloc + 16) * sizeof(double)));
  CUDA_OK(cudaMemsetAsync(S->w, 0, ((size_t)S->nloc + 16) * sizeof(double), g_ctx.stream));
  return 0;
}

// Dense row block on one GPU or with the peer-memory exchange: one cooperative launch per classic
// SOLVE -- loop entry and all iterations behind grid barriers (kernels.cu: k_cg_dense_solve).  NCCL
// mode keeps the three-kernel form (its sums happen between launches).
bool dense_one_kernel(const System* S) {
  return S->dense && S->nloc > 0 && !S->pc_loop && (S->world == 1 || S->p2p) &&
         dense_iter_supported(S->A.vals, S->n_global, (int)S->n_global, S->p_ext);
}

static int launch_dense_solve(System* S, int iters, int init, int max_iter, double tol, int timing) {
  Ctx& c = g_ctx;
  const bool peers = S->world > 1 && S->p2p && !timing;
  CUDA_OK(launch_cg_dense_solve(S->A.vals, S->n_global, S->nloc, (int)S->n_global, S->p_ext, S->p_own,
                                S->x, S->r, S->s, S->b, S->partials, S->state, iters, init, max_iter,
                                tol, peers ? 1 : 0, timing, timing ? nullptr : S->trace, S->gemv_pool,
                                peers ? S->cm : S->cm_off, c.stream));
  return 0;
}

// Classic recurrence (the reference's, cg.c:110-135): K1, K2, K3.
// Single-reduction recurrence: one fused vector update, then K1 with both dot products.
// Sweep direction.  Consecutive kernels of the loop walk the vectors in opposite directions, so a
// kernel starts where its predecessor has just finished and finds the tails of the vectors it
// reads still in the 126 MB L2 (written or read a moment ago) instead of fetching them from HBM.
// Measured (interleaved A/B, one box): -0.7 % per iteration on 2D 4096^2 (134 MB vectors), -0.3 %
// on 512^3 -- far less than the L2 size suggests, the tails do not survive the matrix stream.  Only the order in which tiles / elements are visited changes -- and with
// it the (still fixed) order of the per-CTA partial sums.  CGB200_SWEEP=0: always ascending.
static bool sweep_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("CGB200_SWEEP");
    on = (e && atoi(e) == 0) ? 0 : 1;
  }
  return on != 0;
}

// classic recurrence on one GPU without preconditioner, vectors covered by one resident wave of CTAs
static bool fused_vector_update(const System* S) {
  return S->world == 1 && !S->pc_loop && update_xrp_applies(S->x, S->r, S->p_own, S->s, S->nloc);
}

static int enqueue_iteration(System* S, int variant, int it = 0) {
  Ctx& c = g_ctx;
  CgState* st = S->state;
  const PeerComm& cm = comm_of(S);
  const bool nccl = S->world > 1 && !S->p2p;
  const bool sw = sweep_enabled();
  if (variant == 1) {  // two kernels per iteration: the update ascends, the matvec descends
    CUDA_OK(launch_update_sr(S->x, S->p_own, S->r, S->s, S->w, S->nloc, st, 0, cm, c.stream));
    if (enqueue_matvec(S, S->w, nccl ? FIN_LOCAL : FIN_SR, true, true, sw)) return -1;
    if (nccl && enqueue_nccl_finish(S, 2, FIN_SR)) return -1;
    return 0;
  }
  if (S->pc_loop) {  // Jacobi-preconditioned recurrence (extension): K1 unchanged, K2/K3 carry dinv
    if (enqueue_matvec(S, S->s, nccl ? FIN_LOCAL : FIN_PS, false, true, false)) return -1;
    if (nccl && enqueue_nccl_finish(S, 1, FIN_PS)) return -1;
    CUDA_OK(launch_update_xr_pc(S->x, S->r, S->p_own, S->s, S->dinv, S->nloc, S->partials,
                                &st->ticket[1], st, nccl ? FIN_LOCAL : FIN_RZ, cm, c.stream));
    if (nccl && enqueue_nccl_finish(S, 2, FIN_RZ)) return -1;
    CUDA_OK(launch_update_p_pc(S->p_own, S->r, S->dinv, S->nloc, st, cm, c.stream));
    return 0;
  }
  // three kernels per iteration, alternating: K1, K2, K3 of iteration `it` are kernels 3 it + 0..2
  const bool d1 = sw && ((3 * it) & 1), d2 = sw && ((3 * it + 1) & 1), d3 = sw && ((3 * it + 2) & 1);
  if (enqueue_matvec(S, S->s, nccl ? FIN_LOCAL : FIN_PS, false, true, d1)) return -1;
  if (nccl && enqueue_nccl_finish(S, 1, FIN_PS)) return -1;
  if (fused_vector_update(S)) {  // small single-GPU system: K2 and K3 in one launch
    CUDA_OK(launch_update_xrp(S->x, S->r, S->p_own, S->s, S->nloc, S->partials, &st->ticket[1], st,
                              d2 ? 1 : 0, cm, c.stream));
    return 0;
  }
  CUDA_OK(launch_update_xr(S->x, S->r, S->p_own, S->s, S->nloc, S->partials, &st->ticket[1], st,
                           nccl ? FIN_LOCAL : FIN_RR, d2 ? 1 : 0, cm, c.stream));
  if (nccl && enqueue_nccl_finish(S, 1, FIN_RR)) return -1;
  CUDA_OK(launch_update_p(S->p_own, S->r, S->nloc, st, d3 ? 1 : 0, cm, c.stream));
  return 0;
}

// Kernel nodes per captured iteration (for the launch counter)
static int kernels_per_iteration(const System* S, int variant) {
  int k = variant == 1 ? 2 : 3;
  if (variant == 0 && fused_vector_update(S)) k = 2;
  if (S->world > 1 && !S->p2p) {
    k += variant == 1 ? 1 : 2;  // bookkeeping kernels
    if (!S->dense && !S->allgather && S->tb_hi > S->tb_lo &&
        (S->tb_lo > 0 || S->tb_hi < S->ntiles))
      k += 1;  // split K1
  }
  return k;
}

static int get_graph(System* S, int iters, int variant, cudaGraphExec_t* out) {
  Ctx& c = g_ctx;
  const int key = 16 * iters + 8 * (fused_vector_update(S) ? 1 : 0) + 4 * (S->pc_loop ? 1 : 0) + variant;
  auto it = S->graphs.find(key);
  if (it != S->graphs.end()) {
    *out = it->second;
    return 0;
  }
  const long long before = launch_count();
  cudaGraph_t graph = nullptr;
  CUDA_OK(cudaStreamBeginCapture(c.stream, cudaStreamCaptureModeThreadLocal));
  int rc = 0;
  for (int i = 0; i < iters && rc == 0; ++i) rc = enqueue_iteration(S, variant, i);
  cudaError_t e = cudaStreamEndCapture(c.stream, &graph);
  add_launches(before - launch_count());  // capture is not execution
  if (rc) {
    if (graph) cudaGraphDestroy(graph);
    return -1;
  }
  if (e != cudaSuccess) {
    set_error(std::string("graph capture: ") + cudaGetErrorString(e));
    return -1;
  }
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess) {
    set_error(std::string("graph instantiate: ") + cudaGetErrorString(e));
    return -1;
  }
  S->graphs[key] = exec;
  *out = exec;
  return 0;
}

// Matrix bytes the matvec streams per launch: CSR 12 nnz + 4 (n + 1), dense values only 8 nnz,
// pattern format one byte per row (the dictionary stays in shared memory).
double System::matrix_bytes() const {
  const double n = (double)nloc;
  if (dense) return 8.0 * (double)A.nnz;
  if (dia_active) return 8.0 * (double)dia.D * n;  // every diagonal in full, no indices
  if (pat_active) return n;
  return 12.0 * (double)A.nnz + 4.0 * (n + 1.0);
}
double System::iter_bytes() const { return matrix_bytes() + 88.0 * (double)nloc; }
double System::k1_bytes() const { return matrix_bytes() + 16.0 * (double)nloc; }

// Bytes this rank sends to its peers per classic iteration (NVLink, one direction): its share of
// the search-direction exchange (peer-memory mode: the rows it stores into the peers' gather
// vectors; NCCL mode: the halo rows it sends, or its block to every peer for an all-gather) plus
// the two scalar sums (one 32-byte flagged slot per peer and sum; NCCL mode: counted alike).
double System::comm_bytes_per_iter() const {
  if (world <= 1) return 0.0;
  double rows = 0.0;
  if (p2p) {
    for (int t = 0; t < cm.npush; ++t) rows += (double)(cm.push[t].end - cm.push[t].begin);
  } else if (allgather) {
    rows = (double)nloc * (world - 1);
  } else {
    rows = (double)send_lo * (rank > 0 ? 1 : 0) + (double)send_hi * (rank + 1 < world ? 1 : 0);
  }
  return 8.0 * rows + 2.0 * 32.0 * (world - 1);
}

// Switches the matvec of a CSR system between the stream kernel and the row-pattern kernel.
int system_use_pattern(System* S, int on) {
  Ctx& c = g_ctx;
  if (!S) {
    set_error("system_use_pattern: NULL system");
    return -1;
  }
  CUDA_OK(cudaSetDevice(c.device));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  int result = 0;
  if (on && !S->dense && S->nloc > 0) {
    if (!S->pat.type) {
      std::string why;
      const int rc = build_pattern(S->A, &S->pat, c.stream, &why);
      if (rc < 0) return -1;
      if (rc == 0) set_error("pattern format not applicable: " + why);
    }
    result = S->pat.type ? 1 : 0;
  } else if (on) {
    set_error("pattern format not applicable: dense or empty row block");
  }
  // every rank must take the same path only in the sense of correctness of its own rows: the
  // exchange, the reductions and the tile classification do not depend on the format
  const bool active = result == 1;
  if (active) S->dia_active = false;  // one format at a time
  if (active != S->pat_active) {
    S->pat_active = active;
    for (auto& kv : S->graphs)
      if (kv.second) cudaGraphExecDestroy(kv.second);
    S->graphs.clear();
  }
  return result;
}

// Switches the matvec of a banded CSR system between its CSR kernel and the diagonal-storage kernel.
int system_use_banded(System* S, int on) {
  Ctx& c = g_ctx;
  if (!S) {
    set_error("system_use_banded: NULL system");
    return -1;
  }
  CUDA_OK(cudaSetDevice(c.device));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  int result = 0;
  if (on && !S->dense && S->nloc > 0 && S->world == 1) {
    if (!S->dia.dia) {
      std::string why;
      const int rc = build_banded(S->A, S->off, &S->dia, c.stream, &why);
      if (rc < 0) {
        set_error(std::string("system_use_banded: ") + cudaGetErrorString(cudaGetLastError()));
        return -1;
      }
      if (rc == 0) set_error("banded format not applicable: " + why);
    }
    result = S->dia.dia ? 1 : 0;
  } else if (on) {
    set_error("banded format not applicable: dense, empty or multi-rank row block");
  }
  const bool active = result == 1;
  if (active != S->dia_active) {
    S->dia_active = active;
    if (active) S->pat_active = false;  // one format at a time
    for (auto& kv : S->graphs)
      if (kv.second) cudaGraphExecDestroy(kv.second);
    S->graphs.clear();
  }
  return result;
}

// 1 / diag(A) of the owned rows, built once per system.  A row without a positive finite diagonal
// entry makes the Jacobi preconditioner meaningless: refused, on every rank alike.
static int ensure_dinv(System* S) {
  Ctx& c = g_ctx;
  if (S->dinv) return 0;
  double* d = nullptr;
  CUDA_OK(DEV_ALLOC(&d, ((size_t)S->nloc + 16) * sizeof(double)));
  int* bad = (int*)c.d_scratch;
  CUDA_OK(cudaMemsetAsync(bad, 0, sizeof(int), c.stream));
  if (S->nloc > 0)
    CUDA_OK(launch_diag_inv(S->A, S->dense, S->n_global, S->dense ? S->row0 : S->off, d, bad,
                            c.stream));
  CUDA_OK(cudaMemcpyAsync(c.h_scratch, bad, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  long long nbad = *(const int*)c.h_scratch;
  if (S->world > 1) {
    std::vector<long long> all;
    if (allgather_i64(&nbad, 1, all, S->world)) {
      dev_free(d);
      return -1;
    }
    nbad = 0;
    for (long long v : all) nbad += v;
  }
  if (nbad > 0) {
    dev_free(d);
    set_error("Jacobi preconditioner: " + std::to_string(nbad) +
              " row(s) without a positive finite diagonal entry");
    return -1;
  }
  S->dinv = d;
  return 0;
}

// Marks the system as running the preconditioned loop for the duration of one solve (the graph
// cache and the one-kernel dense iteration look at it).
struct PcScope {
  System* S;
  PcScope(System* s, bool on) : S(s) { S->pc_loop = on; }
  ~PcScope() { S->pc_loop = false; }
};

int system_solve(System* S, const SolveOpts& o, SolveResult* res) {
  Ctx& c = g_ctx;
  if (!S) {
    set_error("system_solve: NULL system");
    return -1;
  }
  if (o.variant != 0 && o.variant != 1) {
    set_error("system_solve: variant must be CGB200_CLASSIC or CGB200_SINGLE_REDUCTION");
    return -1;
  }
  const bool jacobi = (o.flags & 1) != 0, warm = (o.flags & 2) != 0;
  if ((jacobi || warm) && o.variant != 0) {
    set_error("system_solve: CGB200_OPT_JACOBI / CGB200_OPT_X0 need the classic recurrence");
    return -1;
  }
  CUDA_OK(cudaSetDevice(c.device));
  if (jacobi && ensure_dinv(S)) return -1;
  PcScope pc_scope(S, jacobi);
  const int variant = o.variant;
  const bool multi = S->world > 1;
  const bool nccl = multi && !S->p2p;
  const PeerComm& cm = comm_of(S);
  const long long launches0 = launch_count();
  const int max_iter = o.max_iter < 0 ? 0 : o.max_iter;

  // dense row block, classic recurrence: one persistent launch runs the loop entry and every
  // iteration; it reads max_iter / tol from its arguments, so not even the scalars are uploaded
  const bool dense1 = variant == 0 && !jacobi && dense_one_kernel(S);
  const bool fused_entry = dense1 && !warm;

  // device scalars for this solve
  CgState* h = &c.h_state[2];
  if (!fused_entry) {
    memset(h, 0, sizeof(CgState));
    h->tol = o.tol > 0.0 ? o.tol : 0.0;
    h->max_iter = max_iter;
    h->variant = variant;
    CUDA_OK(cudaMemcpyAsync(S->state, h, sizeof(CgState), cudaMemcpyHostToDevice, c.stream));
  }

  if (S->nloc == 0 && !multi) {
    if (res) {
      memset(res, 0, sizeof(*res));
      res->status = CGB200_ST_EXACT;
      res->world = 1;
    }
    return 0;
  }
  if (variant == 1 && ensure_w(S)) return -1;

  if (jacobi || warm) {
    // general loop entry (extension): r = b - A x0, p = z = dinv .* r.  The sum of b.b sits between
    // the product A x0 and the push of p, so no rank overwrites halo slots a peer still reads.
    if (warm) {
      CUDA_OK(cudaMemcpyAsync(S->p_own, S->x, sizeof(double) * S->nloc, cudaMemcpyDeviceToDevice,
                              c.stream));
      if (multi && S->p2p) CUDA_OK(launch_push(S->p_own, S->cm, c.stream));
      if (enqueue_matvec(S, S->s, FIN_NONE, false, true)) return -1;
    }
    CUDA_OK(launch_init_bb(S->x, S->b, S->nloc, warm ? 0 : 1, S->partials, &S->state->ticket[2],
                           S->state, nccl ? FIN_LOCAL : FIN_BB, cm, c.stream));
    if (nccl && enqueue_nccl_finish(S, 1, FIN_BB)) return -1;
    CUDA_OK(launch_init_r(S->r, S->p_own, S->b, warm ? S->s : nullptr, jacobi ? S->dinv : nullptr,
                          S->nloc, S->partials, &S->state->ticket[1], S->state,
                          nccl ? FIN_LOCAL : FIN_INIT2, cm, c.stream));
    if (nccl && enqueue_nccl_finish(S, 2, FIN_INIT2)) return -1;
  } else if (!fused_entry) {
  // loop entry: x = 0, r = p = b, rr = bb = b.b                         (cg.c:104-108)
  CUDA_OK(launch_init(S->x, S->r, S->p_own, S->s, S->b, S->nloc, variant == 1 ? 1 : 0, S->partials,
                      &S->state->ticket[2], S->state, nccl ? FIN_LOCAL : FIN_INIT, cm, c.stream));
  if (nccl && enqueue_nccl_finish(S, 1, FIN_INIT)) return -1;
  }
  if (variant == 1) {  // w = A r, alpha_0 = r.r / r.w
    if (enqueue_matvec(S, S->w, nccl ? FIN_LOCAL : FIN_SR, true, true)) return -1;
    if (nccl && enqueue_nccl_finish(S, 2, FIN_SR)) return -1;
  }

  const long long total_updates = (long long)max_iter + 1;
  // Iterations per graph launch.  The stop flag is looked at one batch behind, so up to two
  // batches of early-exit launches follow a tolerance stop: with a tolerance and no explicit
  // check_every the batches start small and double (4, 8, 16, 32), which keeps that overrun short
  // for systems that converge in a handful of iterations and costs nothing for long solves.
  const int batch_max = o.check_every > 0 ? o.check_every : 32;
  int batch = (o.check_every > 0 || !(o.tol > 0.0)) ? batch_max : 4;
  const int kpi = kernels_per_iteration(S, variant);

  CUDA_OK(cudaEventRecord(c.t0, c.stream));
  if (dense1) {
    // the launch ends by itself when the loop stops (count, tolerance, breakdown): no polling
    if (launch_dense_solve(S, (int)std::min<long long>(total_updates, 0x7fffffff), fused_entry ? 1 : 0,
                           max_iter, o.tol > 0.0 ? o.tol : 0.0, 0))
      return -1;
  } else {
  long long enq = 0;
  int k = 0;
  bool stop = false;
  while (enq < total_updates && !stop) {
    int nb = batch;
    if (nb > total_updates - enq) nb = (int)(total_updates - enq);  // a tail graph, not early-exit launches
    cudaGraphExec_t exec = nullptr;
    if (get_graph(S, nb, variant, &exec)) return -1;
    CUDA_OK(cudaGraphLaunch(exec, c.stream));
    add_launches((long long)kpi * nb);
    enq += nb;
    if (batch < batch_max) batch = std::min(batch_max, batch * 2);
    CUDA_OK(cudaMemcpyAsync(&c.h_state[k & 1], S->state, sizeof(CgState), cudaMemcpyDeviceToHost,
                            c.stream));
    CUDA_OK(cudaEventRecord(c.poll_ev[k & 1], c.stream));
    if (k > 0) {  // look at the batch before the one just queued: the GPU never idles
      CUDA_OK(cudaEventSynchronize(c.poll_ev[(k - 1) & 1]));
      if (c.h_state[(k - 1) & 1].done) stop = true;
    }
    ++k;
  }
  }
  CUDA_OK(cudaEventRecord(c.t1, c.stream));
  CUDA_OK(cudaMemcpyAsync(&c.h_state[3], S->state, sizeof(CgState), cudaMemcpyDeviceToHost,
                          c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  float ms = 0.f;
  CUDA_OK(cudaEventElapsedTime(&ms, c.t0, c.t1));
  const CgState& f = c.h_state[3];
  if (res) {
    res->iterations = f.iter;
    res->status = f.status;
    res->rr = jacobi ? f.aux[0] : f.rr;  // with a preconditioner S->rr carries r.z
    res->bb = f.bb;
    res->relres = f.bb > 0.0 ? sqrt(res->rr / f.bb) : 0.0;
    res->seconds = ms * 1e-3;
    res->bytes_per_iter = S->iter_bytes() + (jacobi ? 16.0 * (double)S->nloc : 0.0);
    res->path = S->dense ? 1 : 0;
    res->world = S->world;
    res->kernel_launches = launch_count() - launches0;
  }
  if (f.status == CGB200_ST_COMM) {
    set_error("system_solve: a peer-memory exchange between GPUs timed out");
    return -1;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Asynchronous fixed-count solves: enqueue now, collect the result later.  A fixed-count solve
// needs no decision from the host (the stop flag only matters for a tolerance), so the whole of
// it -- scalars upload, loop entry, the graph launches, the read-back of the final scalars -- is
// queued without a single synchronisation and the host is free to queue the next right-hand side
// and the next solve behind it: back-to-back solves leave no idle gap on the GPU.
// ---------------------------------------------------------------------------------------------

constexpr int ASYNC_SLOTS = 8;
struct AsyncSlot {
  CgState* up = nullptr;    // pinned: scalars uploaded at the start of the solve
  CgState* down = nullptr;  // pinned: scalars read back at its end
  cudaEvent_t t0 = nullptr, t1 = nullptr, done = nullptr;
  long long launches = 0;
  System* sys = nullptr;
  long long ticket = -1;
};
static AsyncSlot g_async[ASYNC_SLOTS];
static long long g_async_next = 0;

static int async_slot_init(AsyncSlot& a) {
  if (a.up) return 0;
  CUDA_OK(cudaMallocHost(&a.up, sizeof(CgState) * 2));
  a.down = a.up + 1;
  CUDA_OK(cudaEventCreate(&a.t0));
  CUDA_OK(cudaEventCreate(&a.t1));
  CUDA_OK(cudaEventCreateWithFlags(&a.done, cudaEventDisableTiming));
  return 0;
}

int system_solve_async(System* S, const SolveOpts& o, long long* ticket_out) {
  Ctx& c = g_ctx;
  if (!S || !ticket_out) {
    set_error("system_solve_async: NULL argument");
    return -1;
  }
  if (o.tol > 0.0) {
    set_error("system_solve_async: only fixed-count solves (tol = 0) can be queued; a tolerance "
              "needs the host to watch the stop flag -- use cgb200_system_solve");
    return -1;
  }
  if (o.variant != 0 && o.variant != 1) {
    set_error("system_solve_async: unknown variant");
    return -1;
  }
  if (o.flags != 0) {
    set_error("system_solve_async: CGB200_OPT_JACOBI / CGB200_OPT_X0 are not available for queued "
              "solves -- use cgb200_system_solve");
    return -1;
  }
  AsyncSlot& a = g_async[g_async_next % ASYNC_SLOTS];
  if (a.ticket >= 0) {
    set_error("system_solve_async: too many solves in flight; collect earlier tickets first");
    return -1;
  }
  if (async_slot_init(a)) return -1;
  CUDA_OK(cudaSetDevice(c.device));
  const int variant = o.variant;
  const bool nccl = S->world > 1 && !S->p2p;
  const PeerComm& cm = comm_of(S);
  const long long launches0 = launch_count();
  const int max_iter = o.max_iter < 0 ? 0 : o.max_iter;
  memset(a.up, 0, sizeof(CgState));
  a.up->max_iter = max_iter;
  a.up->variant = variant;
  if (variant == 1 && ensure_w(S)) return -1;
  const long long total_updates = (long long)max_iter + 1;
  int batch = o.check_every > 0 ? o.check_every : 32;
  if (batch > total_updates) batch = (int)total_updates;
  if (variant == 0 && dense_one_kernel(S)) {
    // dense row block: the whole solve is one persistent launch (loop entry included)
    CUDA_OK(cudaEventRecord(a.t0, c.stream));
    if (launch_dense_solve(S, (int)std::min<long long>(total_updates, 0x7fffffff), 1, max_iter, 0.0, 0))
      return -1;
  } else {
  cudaGraphExec_t exec = nullptr, exec_tail = nullptr;
  if (get_graph(S, batch, variant, &exec)) return -1;
  const int tail = (int)(total_updates % batch);  // exactly max_iter + 1 updates are enqueued
  if (tail > 0 && get_graph(S, tail, variant, &exec_tail)) return -1;
  const int kpi = kernels_per_iteration(S, variant);

  CUDA_OK(cudaMemcpyAsync(S->state, a.up, sizeof(CgState), cudaMemcpyHostToDevice, c.stream));
  CUDA_OK(launch_init(S->x, S->r, S->p_own, S->s, S->b, S->nloc, variant == 1 ? 1 : 0, S->partials,
                      &S->state->ticket[2], S->state, nccl ? FIN_LOCAL : FIN_INIT, cm, c.stream));
  if (nccl && enqueue_nccl_finish(S, 1, FIN_INIT)) return -1;
  if (variant == 1) {
    if (enqueue_matvec(S, S->w, nccl ? FIN_LOCAL : FIN_SR, true, true)) return -1;
    if (nccl && enqueue_nccl_finish(S, 2, FIN_SR)) return -1;
  }
  CUDA_OK(cudaEventRecord(a.t0, c.stream));
  for (long long enq = 0; enq + batch <= total_updates; enq += batch) {
    CUDA_OK(cudaGraphLaunch(exec, c.stream));
    add_launches((long long)kpi * batch);
  }
  if (tail > 0) {
    CUDA_OK(cudaGraphLaunch(exec_tail, c.stream));
    add_launches((long long)kpi * tail);
  }
  }
  CUDA_OK(cudaEventRecord(a.t1, c.stream));
  CUDA_OK(cudaMemcpyAsync(a.down, S->state, sizeof(CgState), cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaEventRecord(a.done, c.stream));
  a.launches = launch_count() - launches0;
  a.sys = S;
  a.ticket = g_async_next++;
  *ticket_out = a.ticket;
  return 0;
}

int system_solve_wait(long long ticket, SolveResult* res) {
  if (ticket < 0) {
    set_error("system_solve_wait: bad ticket");
    return -1;
  }
  AsyncSlot& a = g_async[ticket % ASYNC_SLOTS];
  if (a.ticket != ticket) {
    set_error("system_solve_wait: unknown or already collected ticket");
    return -1;
  }
  CUDA_OK(cudaEventSynchronize(a.done));
  float ms = 0.f;
  CUDA_OK(cudaEventElapsedTime(&ms, a.t0, a.t1));
  const CgState& f = *a.down;
  if (res) {
    res->iterations = f.iter;
    res->status = f.status;
    res->rr = f.rr;
    res->bb = f.bb;
    res->relres = f.bb > 0.0 ? sqrt(f.rr / f.bb) : 0.0;
    res->seconds = ms * 1e-3;
    res->bytes_per_iter = a.sys->iter_bytes();
    res->path = a.sys->dense ? 1 : 0;
    res->world = a.sys->world;
    res->kernel_launches = a.launches;
  }
  a.ticket = -1;
  a.sys = nullptr;
  if (f.status == CGB200_ST_COMM) {
    set_error("system_solve_wait: a peer-memory exchange between GPUs timed out");
    return -1;
  }
  return 0;
}

// Phase trace of the one-kernel iteration (tools/dense_trace.py): on = allocate + zero the
// counters and rebuild the graphs with the trace pointer; read = copy the 8 counters out.
constexpr int TRACE_WORDS = 16 + 2048;  // counters, per-CTA duration sums, per-CTA SM ids
int system_trace(System* S, int on, unsigned long long* out) {
  Ctx& c = g_ctx;
  CUDA_OK(cudaStreamSynchronize(c.stream));
  if (out && S->trace)
    CUDA_OK(cudaMemcpy(out, S->trace, TRACE_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (on && !S->trace) CUDA_OK(DEV_ALLOC(&S->trace, TRACE_WORDS * sizeof(unsigned long long)));
  if (on) {
    std::vector<unsigned long long> z(TRACE_WORDS, 0ull);
    z[8] = z[10] = ~0ull;  // minima
    CUDA_OK(cudaMemcpy(S->trace, z.data(), TRACE_WORDS * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  }
  if (!on && S->trace) {
    dev_free(S->trace);
    S->trace = nullptr;
  }
  for (auto& kv : S->graphs)
    if (kv.second) cudaGraphExecDestroy(kv.second);
  S->graphs.clear();
  return 0;
}

int system_set_rhs(System* S, const double* b_host) {
  Ctx& c = g_ctx;
  CUDA_OK(upload_host(S->b, b_host, sizeof(double) * S->nloc, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

int system_gen_rhs(System* S, unsigned long long seed) {
  Ctx& c = g_ctx;
  CUDA_OK(launch_gen_rhs(S->b, S->row0, S->nloc, seed, c.stream));
  return 0;
}

int system_set_x(System* S, const double* x_host) {
  Ctx& c = g_ctx;
  CUDA_OK(upload_host(S->x, x_host, sizeof(double) * S->nloc, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

int system_get_x(System* S, double* x_host) {
  Ctx& c = g_ctx;
  CUDA_OK(download_host(x_host, S->x, sizeof(double) * S->nloc, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

// s = A v with v taken from the device vector `v_dev` (owned rows); exchanges the halo first.
int system_matvec_dev(System* S, const double* v_dev, double* s_host_or_null) {
  Ctx& c = g_ctx;
  if (v_dev != S->p_own)
    CUDA_OK(cudaMemcpyAsync(S->p_own, v_dev, sizeof(double) * S->nloc, cudaMemcpyDeviceToDevice,
                            c.stream));
  if (S->world > 1 && S->p2p) CUDA_OK(launch_push(S->p_own, S->cm, c.stream));
  if (enqueue_matvec(S, S->s, FIN_NONE, false, true)) return -1;
  if (s_host_or_null)
    CUDA_OK(cudaMemcpyAsync(s_host_or_null, S->s, sizeof(double) * S->nloc, cudaMemcpyDeviceToHost,
                            c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  // no reduction follows a plain product, so nothing else keeps a fast rank from pushing its next
  // vector into halo slots a slow rank is still reading
  if (S->world > 1 && S->p2p && dist_barrier()) return -1;
  return 0;
}

// Single rank: s = A v straight from a caller's device vector into a caller's device buffer --
// no staging copies and no host synchronisation (the operator API's resident mode, abi.cu).
// Both buffers hold at least nloc + 16 doubles (the padding the kernels may over-read) and are
// distinct.  Anything else takes the staged path above and copies the result out.
int system_matvec_into(System* S, const double* v_dev, double* out_dev) {
  Ctx& c = g_ctx;
  if (S->world != 1 || S->off != 0 || v_dev == out_dev) {
    if (system_matvec_dev(S, v_dev, nullptr)) return -1;
    CUDA_OK(cudaMemcpyAsync(out_dev, S->s, sizeof(double) * S->nloc, cudaMemcpyDeviceToDevice,
                            c.stream));
    return 0;
  }
  double* const ext = S->p_ext;
  double* const own = S->p_own;
  S->p_ext = const_cast<double*>(v_dev);  // read-only in a plain product
  S->p_own = S->p_ext;
  const int rc = enqueue_matvec(S, out_dev, FIN_NONE, false, false);
  S->p_ext = ext;
  S->p_own = own;
  return rc;
}

int system_matvec(System* S, const double* v_host, double* s_host) {
  Ctx& c = g_ctx;
  CUDA_OK(cudaMemcpyAsync(S->p_own, v_host, sizeof(double) * S->nloc, cudaMemcpyHostToDevice,
                          c.stream));
  return system_matvec_dev(S, S->p_own, s_host);
}

int system_true_relres(System* S, double* relres) {
  Ctx& c = g_ctx;
  if (system_matvec_dev(S, S->x, nullptr)) return -1;
  double* d2 = (double*)c.d_scratch;
  CUDA_OK(cudaMemsetAsync(d2, 0, 2 * sizeof(double), c.stream));
  CUDA_OK(cudaMemsetAsync(&S->state->ticket[0], 0, 4 * sizeof(unsigned), c.stream));
  CUDA_OK(launch_resid_norms(S->b, S->s, S->nloc, S->partials, &S->state->ticket[0], d2, c.stream));
  if (S->world > 1 && allreduce_sum_f64(d2, 2, c.stream)) return -1;
  CUDA_OK(cudaMemcpyAsync(c.h_scratch, d2, 2 * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  const double* h2 = (const double*)c.h_scratch;
  *relres = h2[1] > 0.0 ? sqrt(h2[0] / h2[1]) : 0.0;
  return 0;
}

// Copies the matrix block back to host arrays (any of them may be NULL).  Columns are returned as
// GLOBAL indices.  For a dense block cols/row_ptr are synthesised (row_ptr[i] = i*n, col = j).
int system_download(System* S, double* vals, int* cols, int* row_ptr) {
  Ctx& c = g_ctx;
  const long long nnz = S->A.nnz;
  if (vals && nnz > 0)
    CUDA_OK(cudaMemcpyAsync(vals, S->A.vals, sizeof(double) * nnz, cudaMemcpyDeviceToHost, c.stream));
  if (S->dense) {
    CUDA_OK(cudaStreamSynchronize(c.stream));
    const long long n = S->n_global;
    if (row_ptr)
      for (long long i = 0; i <= S->nloc; ++i) row_ptr[i] = (int)(i * n);
    if (cols)
      for (long long i = 0; i < S->nloc; ++i)
        for (long long j = 0; j < n; ++j) cols[i * n + j] = (int)j;
    return 0;
  }
  if (cols && nnz > 0)
    CUDA_OK(cudaMemcpyAsync(cols, S->A.cols, sizeof(int) * nnz, cudaMemcpyDeviceToHost, c.stream));
  if (row_ptr)
    CUDA_OK(cudaMemcpyAsync(row_ptr, S->A.rowptr, sizeof(int) * ((size_t)S->nloc + 1),
                            cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  const long long shift = (S->world == 1 || S->allgather) ? 0 : S->row0 - S->off;
  if (cols && shift != 0)
    for (long long k = 0; k < nnz; ++k) cols[k] += (int)shift;
  return 0;
}

int system_get_b(System* S, double* b_host) {
  Ctx& c = g_ctx;
  CUDA_OK(cudaMemcpyAsync(b_host, S->b, sizeof(double) * S->nloc, cudaMemcpyDeviceToHost, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

// Device timer on the solve stream (bench harness: brackets K steps with one event pair).
int timer_start() {
  Ctx& c = g_ctx;
  if (!c.t2) {
    CUDA_OK(cudaEventCreate(&c.t2));
    CUDA_OK(cudaEventCreate(&c.t3));
  }
  CUDA_OK(cudaStreamSynchronize(c.stream));
  if (c.world > 1 && c.comm && !c.team_primary) {
    // several ranks: the start event sits behind a device-side rendezvous (a 4-byte all-reduce on
    // the solve stream, not waited for by the host), so every rank's timed region begins when the
    // last rank's stream gets there and not when its host process happened to be scheduled
    CUDA_OK(cudaMemsetAsync(c.d_scratch, 0, 8, c.stream));
    NCCL_OK(g_nccl.AllReduce(c.d_scratch, c.d_scratch, 1, ncclInt32, ncclSum, (ncclComm_t)c.comm,
                             c.stream));
  }
  CUDA_OK(cudaEventRecord(c.t2, c.stream));
  return 0;
}

int timer_stop(double* ms_out) {
  Ctx& c = g_ctx;
  if (!c.t2) {
    set_error("timer_stop without timer_start");
    return -1;
  }
  CUDA_OK(cudaEventRecord(c.t3, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  float ms = 0.f;
  CUDA_OK(cudaEventElapsedTime(&ms, c.t2, c.t3));
  *ms_out = ms;
  return 0;
}

constexpr int TIME_KERNEL_DENSE_ITERS = 5;  // iterations per launch when the dense solve kernel is timed

int system_time_kernel(System* S, int which, int reps, double* ms_out) {
  Ctx& c = g_ctx;
  // a benign state: alpha = beta = 0 keeps x, r, p finite however often the kernels run
  CgState* h = &c.h_state[2];
  memset(h, 0, sizeof(CgState));
  h->rr = 1.0;
  h->max_iter = 0x7fffffff;
  CUDA_OK(cudaMemcpyAsync(S->state, h, sizeof(CgState), cudaMemcpyHostToDevice, c.stream));
  if ((which == 4 || which == 5) && ensure_w(S)) return -1;
  auto one = [&]() -> int {
    if (which == 1) {  // local kernel only: no exchange, sums left in state->local
      if (enqueue_matvec(S, S->s, FIN_LOCAL, false, false)) return -1;
    } else if (which == 2) {
      CUDA_OK(launch_update_xr(S->x, S->r, S->p_own, S->s, S->nloc, S->partials,
                               &S->state->ticket[1], S->state, FIN_LOCAL, 0, S->cm_off, c.stream));
    } else if (which == 3) {
      CUDA_OK(launch_update_p(S->p_own, S->r, S->nloc, S->state, 0, S->cm_off, c.stream));
    } else if (which == 6) {  // the persistent dense solve kernel, scalars left alone
      if (!dense_one_kernel(S)) {
        set_error("system_time_kernel: this system does not run the one-kernel solve");
        return -1;
      }
      if (launch_dense_solve(S, TIME_KERNEL_DENSE_ITERS, 0, 0, 0.0, 1)) return -1;
    } else if (which == 4) {
      CUDA_OK(launch_update_sr(S->x, S->p_own, S->r, S->s, S->w, S->nloc, S->state, 0, S->cm_off,
                               c.stream));
    } else {
      if (enqueue_matvec(S, S->w, FIN_LOCAL, true, false)) return -1;
    }
    return 0;
  };
  for (int i = 0; i < 3; ++i)
    if (one()) return -1;
  CUDA_OK(cudaEventRecord(c.t0, c.stream));
  for (int i = 0; i < reps; ++i)
    if (one()) return -1;
  CUDA_OK(cudaEventRecord(c.t1, c.stream));
  CUDA_OK(cudaStreamSynchronize(c.stream));
  float ms = 0.f;
  CUDA_OK(cudaEventElapsedTime(&ms, c.t0, c.t1));
  *ms_out = (double)ms / reps;
  if (which == 6) *ms_out /= TIME_KERNEL_DENSE_ITERS;  // per iteration
  return 0;
}

}  // namespace cgb
