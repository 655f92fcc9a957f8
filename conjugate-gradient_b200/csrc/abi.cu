// abi.cu -- the C ABI of libcgb200.so.
//
// Part 1: the reference's operator API (include/mv_ops.h; reference mv_ops.h:26-42) and its
//         solver entry point conj_grad (reference cg.c:24,88), symbol for symbol.  Host structs
//         in, host structs out; the arithmetic runs in the kernels of kernels.cu.
// Part 2: the cgb200_* extensions (include/cgb200.h).
//
// There is no CPU implementation of any arithmetic in this file: without a usable CUDA device
// every arithmetic entry point fails (-1 / -1.0) and cgb200_last_error() says so.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <map>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include "devmem.h"
#include "../../include/cgb200.h"
#include "solver.h"
#include "team.h"
#include "writewatch.h"

using namespace cgb;

namespace {

#define CU_OK(call)                                                              \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      set_error(std::string(#call) + ": " + cudaGetErrorString(e__));            \
      return -1;                                                                 \
    }                                                                            \
  } while (0)

std::mutex g_tab_mu;
// doubles behind every device vector of the operator API: the matvec kernels may over-read their
// gather vector by up to a vector load (solver.cu pads the systems' own vectors the same way)
constexpr size_t VEC_PAD = 16;
// CGB200_MODE=resident starts the operator API in resident mode: the way to run an unmodified
// caller of mv_ops.h (the reference's own cg.c loop) without a transfer per operator call
static int initial_mode() {
  const char* e = getenv("CGB200_MODE");
  return (e && strcmp(e, "resident") == 0) ? CGB200_RESIDENT : CGB200_EAGER;
}
int g_mode = initial_mode();

// ---- device mirrors of caller structs (keyed by struct address) -----------------------------

struct MatrixEntry {
  System* sys = nullptr;       // single-GPU copy (mv_mult, and solves without a GPU team)
  TeamSystem* team = nullptr;  // row blocks over the GPU team (solves with CGB200_GPUS > 1)
  const double* values = nullptr;
  const int* cols = nullptr;
  const int* rowptr = nullptr;
  int size = 0, nnz = 0;
  uint64_t fp = 0;       // sampled fingerprint (cheap; every call)
  uint64_t content = 0;  // hash of every byte of the three arrays (strict policy)
  bool has_content = false;
  bool pinned = false;   // the caller declared the device copy current (cgb200_to_device / resident mode)
  // tracked policy: the page-aligned interiors of the three arrays are watched for writes
  // (writewatch.cpp); their partial first / last pages are hashed instead (edge)
  bool watched = false;
  uint64_t seen[3] = {0, 0, 0};  // epoch of each watched range at this entry's last validation
  uint64_t edge = 0;
};

// what the cache did, for tests and the bench line (cgb200_cache_stats)
long long g_stat_uploads = 0, g_stat_hashes = 0, g_stat_tracked_clean = 0, g_stat_tracked_dirty = 0;

struct VectorEntry {
  double* d = nullptr;
  int cap = 0;
  int n = 0;
  // host array the mirror was last validated against: a struct that was released with plain free()
  // (the reference's free_mv_struct is just that, mv_ops.c:39-42) and whose address malloc hands
  // out again comes back with another values pointer -- the stale mirror is then not trusted
  const double* host_ptr = nullptr;
  bool dev_valid = false;   // device copy holds the current values
  bool host_valid = false;  // host array holds the current values
  // resident mode: v.v of the current device values, once a dot_product(v, v) has computed it.
  // The reference's loop asks for r.r twice per residual (cg.c:129 then cg.c:113) and once more
  // for the copy it keeps (r_prev, cg.c:120,129): the kernel is deterministic, so handing back the
  // stored value is bit-identical to recomputing it and saves two launches and two host
  // synchronisations per iteration.  Cleared by every write; carried over by mv_deep_copy.
  bool sd_valid = false;
  double sd = 0.0;
  // resident mode: a deferred sv_mult.  The mirror's values are lazy_s * (the values of lazy_src's
  // mirror) and have not been computed: the reference's loop reads each scaled vector exactly once,
  // in the vec_add / vec_sub that follows (cg.c:115-118, 122-123, 131-132), and that operator then
  // forms a +/- fl(s*b) itself in one pass -- same two roundings, one pass over memory less.  Any
  // other reader computes the values first (materialize), and so does anything about to change
  // or release the source (before_write), so a deferred vector never observes a changed source.
  bool lazy = false;
  double lazy_s = 0.0;
  const void* lazy_src = nullptr;
};

std::unordered_map<const void*, MatrixEntry> g_mats;
std::unordered_map<const void*, VectorEntry> g_vecs;

// eager-mode scratch vectors
double* g_scratch[3] = {nullptr, nullptr, nullptr};
size_t g_scratch_cap[3] = {0, 0, 0};

int scratch(int slot, size_t n, double** out) {
  if (g_scratch_cap[slot] < n + VEC_PAD) {
    if (g_scratch[slot]) dev_free(g_scratch[slot]);
    g_scratch[slot] = nullptr;
    g_scratch_cap[slot] = 0;
    CU_OK(DEV_ALLOC(&g_scratch[slot], (n + VEC_PAD) * sizeof(double)));
    g_scratch_cap[slot] = n + VEC_PAD;
  }
  *out = g_scratch[slot];
  return 0;
}

// ---- is the device copy of a caller's matrix still the caller's matrix? ------------------------
//
// The reference reads the caller's arrays on every call (mv_ops.c:187-197), so an unmodified
// caller may edit values[] in place between two mv_mult / conj_grad calls and expects the new
// product.  The device copy is therefore re-validated on EVERY call that takes a matrix:
//   strict  (default)  a hash of every byte of values[], col_indices[], row_ptr[] (several host
//                      threads; one pass over host memory, an order of magnitude cheaper than
//                      the upload from pageable memory it saves).  A change confined to one
//                      8-byte word is always seen (every step of the hash is a bijection of its
//                      state); anything else with probability 1 - 2^-64.
//   sampled            96 samples + the array addresses: constant time, sees a rebuilt matrix,
//                      misses a small in-place edit -- cgb200_invalidate() is then the contract.
//   off                no cache: every call uploads.
// CGB200_MATRIX_CACHE=strict|sampled|off, cgb200_set_matrix_cache().  A matrix the caller has
// handed over explicitly -- cgb200_to_device(A) -- and every matrix in resident mode is PINNED:
// the caller has stated that the device copy is current until cgb200_invalidate / free, and only
// the sampled check runs (a rebuilt matrix behind the same struct is still noticed).
int g_cache_policy = -1;  // CGB200_CACHE_*
int cache_policy() {
  if (g_cache_policy < 0) {
    const char* e = getenv("CGB200_MATRIX_CACHE");
    g_cache_policy = CGB200_CACHE_TRACKED;
    if (e && strcmp(e, "strict") == 0) g_cache_policy = CGB200_CACHE_STRICT;
    if (e && strcmp(e, "sampled") == 0) g_cache_policy = CGB200_CACHE_SAMPLED;
    if (e && strcmp(e, "off") == 0) g_cache_policy = CGB200_CACHE_OFF;
  }
  return g_cache_policy;
}

uint64_t mix(uint64_t h, uint64_t v) {
  h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
  return h;
}
uint64_t fingerprint(const struct __mv_sparse* A) {
  uint64_t h = 1469598103934665603ull;
  const long long nnz = A->nnz, n = A->size;
  const long long step_v = nnz > 64 ? nnz / 64 : 1;
  for (long long k = 0; k < nnz; k += step_v) {
    uint64_t bits;
    memcpy(&bits, &A->values[k], 8);
    h = mix(h, bits);
    if (A->col_indices) h = mix(h, (uint64_t)(uint32_t)A->col_indices[k]);
  }
  if (nnz > 0) {
    uint64_t bits;
    memcpy(&bits, &A->values[nnz - 1], 8);
    h = mix(h, bits);
  }
  const long long step_r = n > 32 ? n / 32 : 1;
  for (long long i = 0; i <= n; i += step_r) h = mix(h, (uint64_t)(uint32_t)A->row_ptr[i]);
  h = mix(h, (uint64_t)(uint32_t)A->row_ptr[n]);
  return h;
}

// Hash of one byte range: four independent lanes of (h ^ word) * odd constant, so the loop is
// bound by memory, not by the multiplier's latency.
uint64_t hash_range(const unsigned char* p, size_t bytes) {
  const uint64_t P0 = 0x9E3779B97F4A7C15ull, P1 = 0xC2B2AE3D27D4EB4Full, P2 = 0x165667B19E3779F9ull,
                 P3 = 0xD6E8FEB86659FD93ull;
  uint64_t h0 = 0x243F6A8885A308D3ull, h1 = 0x13198A2E03707344ull, h2 = 0xA4093822299F31D0ull,
           h3 = 0x082EFA98EC4E6C89ull;
  size_t i = 0;
  for (; i + 32 <= bytes; i += 32) {
    uint64_t w[4];
    memcpy(w, p + i, 32);
    h0 = (h0 ^ w[0]) * P0;
    h1 = (h1 ^ w[1]) * P1;
    h2 = (h2 ^ w[2]) * P2;
    h3 = (h3 ^ w[3]) * P3;
  }
  for (; i < bytes; i += 8) {
    uint64_t w = 0;
    memcpy(&w, p + i, bytes - i < 8 ? bytes - i : 8);
    h0 = (h0 ^ w) * P0;
    h0 = (h0 << 29) | (h0 >> 35);
  }
  uint64_t h = h0;
  h = (h ^ ((h1 << 17) | (h1 >> 47))) * P1;
  h = (h ^ ((h2 << 31) | (h2 >> 33))) * P2;
  h = (h ^ ((h3 << 43) | (h3 >> 21))) * P3;
  return h ^ (h >> 32);
}

// Hash of a whole array, cut into 4 MiB pieces dealt to up to 32 threads; the pieces' hashes are
// chained in order, so the result does not depend on the thread count.
uint64_t hash_array(const void* ptr, size_t bytes, uint64_t seed) {
  if (!ptr || bytes == 0) return seed * 0x9E3779B97F4A7C15ull + 1;
  const size_t PIECE = 4u << 20;
  const size_t npieces = (bytes + PIECE - 1) / PIECE;
  std::vector<uint64_t> hp(npieces);
  const unsigned char* p = (const unsigned char*)ptr;
  auto work = [&](size_t t, size_t nt) {
    for (size_t k = t; k < npieces; k += nt) {
      const size_t off = k * PIECE;
      hp[k] = hash_range(p + off, bytes - off < PIECE ? bytes - off : PIECE);
    }
  };
  size_t nt = std::thread::hardware_concurrency();
  if (const char* e = getenv("CGB200_HASH_THREADS")) nt = (size_t)atoi(e);
  if (nt > 32) nt = 32;
  if (nt > npieces) nt = npieces;
  if (nt <= 1) {
    work(0, 1);
  } else {
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; ++t) th.emplace_back(work, t, nt);
    work(0, nt);
    for (auto& x : th) x.join();
  }
  uint64_t h = seed ^ (uint64_t)bytes;
  for (size_t k = 0; k < npieces; ++k) h = (h ^ hp[k]) * 0xD6E8FEB86659FD93ull + 0x9E3779B97F4A7C15ull;
  return h;
}

uint64_t content_hash(const struct __mv_sparse* A) {
  const size_t nnz = A->nnz > 0 ? (size_t)A->nnz : 0;
  uint64_t h = hash_array(A->values, nnz * sizeof(double), 1);
  h = hash_array(A->col_indices, A->col_indices ? nnz * sizeof(int) : 0, h);
  h = hash_array(A->row_ptr, ((size_t)A->size + 1) * sizeof(int), h);
  return h;
}

bool is_vector(const struct __mv_sparse* v) { return v->col_indices == NULL && v->row_ptr == NULL; }

// Device system for a host CSR struct, uploaded on first use and whenever the struct's arrays no
// longer hold what was uploaded (policy above).  pin: the caller hands the matrix over explicitly.
// The three host arrays of a matrix struct.
struct HostArr {
  const void* p;
  size_t bytes;
};
void host_arrays(const struct __mv_sparse* A, HostArr out[3]) {
  const size_t nnz = A->nnz > 0 ? (size_t)A->nnz : 0;
  out[0] = HostArr{A->values, nnz * sizeof(double)};
  out[1] = HostArr{A->col_indices, A->col_indices ? nnz * sizeof(int) : 0};
  out[2] = HostArr{A->row_ptr, ((size_t)A->size + 1) * sizeof(int)};
}
void host_arrays(const MatrixEntry& e, HostArr out[3]) {
  const size_t nnz = e.nnz > 0 ? (size_t)e.nnz : 0;
  out[0] = HostArr{e.values, nnz * sizeof(double)};
  out[1] = HostArr{e.cols, e.cols ? nnz * sizeof(int) : 0};
  out[2] = HostArr{e.rowptr, ((size_t)e.size + 1) * sizeof(int)};
}

constexpr size_t WATCH_MIN_BYTES = 1u << 20;  // smaller arrays are hashed: cheaper than a page-table walk is worth

// Hash of what write tracking does not cover: the partial first and last pages of the watched
// arrays, and the whole of the arrays too small to watch.  A few KB to a few MB per call.
uint64_t edge_hash(const HostArr a[3]) {
  const size_t pg = ww_page();
  uint64_t h = 7;
  for (int k = 0; k < 3; ++k) {
    const unsigned char* p = (const unsigned char*)a[k].p;
    if (!p || a[k].bytes == 0) continue;
    if (a[k].bytes < WATCH_MIN_BYTES) {
      h = hash_array(p, a[k].bytes, h);
      continue;
    }
    const uintptr_t lo = ((uintptr_t)p + pg - 1) / pg * pg, hi = ((uintptr_t)p + a[k].bytes) / pg * pg;
    h = hash_array(p, lo - (uintptr_t)p, h);
    h = hash_array((const void*)hi, (uintptr_t)p + a[k].bytes - hi, h);
  }
  return h;
}

// Pageable memory only: device DMA into page-locked memory (cudaHostAlloc, cudaHostRegister) does
// not go through the CPU's page tables and would not be seen.
bool plain_pageable(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeUnregistered;
}

// Watched ranges are shared between cache entries: two matrix structs may point at the same arrays,
// and re-arming a range for one of them erases the "written" marks the other has not looked at yet.
// So every range carries an epoch that is bumped whenever its marks are cleared, and an entry trusts
// "no page written" only while the epoch is the one it saw at its own last validation.  Ranges that
// overlap without being identical are not watched (those entries hash).
struct WatchedRange {
  size_t bytes = 0;
  int refs = 0;
  uint64_t epoch = 0;
};
std::map<uintptr_t, WatchedRange> g_ranges;  // by start address; guarded by g_tab_mu like the cache

bool watchable(const HostArr& h) { return h.p && h.bytes >= WATCH_MIN_BYTES; }

bool range_acquire(const HostArr& h, uint64_t* seen) {
  const uintptr_t lo = (uintptr_t)h.p, hi = lo + h.bytes;
  auto nx = g_ranges.lower_bound(lo);
  if (nx != g_ranges.end() && nx->first == lo && nx->second.bytes == h.bytes) {
    ++nx->second.refs;  // marks left by writes before now stay: this entry's first check hashes
    *seen = nx->second.epoch;
    return true;
  }
  if (nx != g_ranges.end() && nx->first < hi) return false;
  if (nx != g_ranges.begin()) {
    auto pv = std::prev(nx);
    if (pv->first + pv->second.bytes > lo) return false;
  }
  if (!plain_pageable(h.p) || !ww_arm(h.p, h.bytes)) return false;
  WatchedRange r;
  r.bytes = h.bytes;
  r.refs = 1;
  g_ranges[lo] = r;
  *seen = 0;
  return true;
}

void range_release(const HostArr& h) {
  auto it = g_ranges.find((uintptr_t)h.p);
  if (it == g_ranges.end() || it->second.bytes != h.bytes) return;
  if (--it->second.refs > 0) return;
  ww_disarm(h.p, h.bytes);
  g_ranges.erase(it);
}

void watch_release(const HostArr a[3]) {
  for (int k = 0; k < 3; ++k)
    if (watchable(a[k])) range_release(a[k]);
}

bool watch_acquire(const HostArr a[3], uint64_t seen[3]) {
  if (!ww_available()) return false;
  bool any = false;
  for (int k = 0; k < 3; ++k) {
    if (!watchable(a[k])) continue;
    if (!range_acquire(a[k], &seen[k])) {
      for (int j = 0; j < k; ++j)
        if (watchable(a[j])) range_release(a[j]);
      return false;
    }
    any = true;
  }
  if (!any) return false;  // (nothing large enough to watch: the caller just hashes, which is cheap then)
  return true;
}

// 0 = no watched array was written since this entry last validated, 1 = one was (or its marks
// were cleared on behalf of another entry), -1 = cannot tell
int watch_state(const HostArr a[3], const uint64_t seen[3]) {
  int out = 0;
  for (int k = 0; k < 3; ++k) {
    if (!watchable(a[k])) continue;
    auto it = g_ranges.find((uintptr_t)a[k].p);
    if (it == g_ranges.end() || it->second.bytes != a[k].bytes) return -1;
    if (!plain_pageable(a[k].p)) return -1;  // page-locked since (cudaHostRegister): DMA writes are invisible
    const int w = ww_written(a[k].p, a[k].bytes);
    if (w < 0) return -1;
    out |= w | (it->second.epoch != seen[k] ? 1 : 0);
  }
  return out;
}

// Before the arrays are re-read for the deciding hash: clear the marks of the ranges that carry
// any (bumping their epoch for the other entries that share them) and note the epochs.  false =
// tracking lost (the entry has released its ranges and hashes from now on).
bool watch_refresh(const HostArr a[3], uint64_t seen[3]) {
  bool ok = true;
  for (int k = 0; k < 3 && ok; ++k) {
    if (!watchable(a[k])) continue;
    auto it = g_ranges.find((uintptr_t)a[k].p);
    if (it == g_ranges.end() || it->second.bytes != a[k].bytes || !plain_pageable(a[k].p)) {
      ok = false;
      break;
    }
    if (ww_written(a[k].p, a[k].bytes) != 0) {
      ++it->second.epoch;
      if (!ww_rearm(a[k].p, a[k].bytes)) ok = false;
    }
    seen[k] = it->second.epoch;
  }
  if (!ok) watch_release(a);
  return ok;
}

void release_entry(MatrixEntry& e) {
  if (e.watched) {
    HostArr a[3];
    host_arrays(e, a);
    watch_release(a);
    e.watched = false;
  }
  delete e.sys;
  e.sys = nullptr;
  if (e.team) team_destroy(e.team);  // collective with the workers
  e.team = nullptr;
}

// The tracked policy's question to the kernel takes 0.1-0.5 ms on a multi-GB matrix (a walk over
// its page tables) -- against a 1.2 ms matvec.  Callers that are about to start device work on the
// matrix hand in a PendingCheck: the question is then asked on a host thread WHILE the device works
// on the copy it already has, and settle() says afterwards whether that copy was the current one.
// If not (the arrays were written, or the kernel cannot tell), nothing has left the library yet:
// the caller validates the ordinary way and repeats the device work.
struct PendingCheck {
  std::thread th;
  MatrixEntry* entry = nullptr;
  int w = 0;
  uint64_t edge = 0;
  bool started = false;
  ~PendingCheck() {
    if (th.joinable()) th.join();
  }
  // true = the device copy that was used is the caller's current matrix
  bool settle() {
    if (!started) return true;
    th.join();
    started = false;
    if (w == 0 && edge == entry->edge) {
      ++g_stat_tracked_clean;
      return true;
    }
    return false;
  }
};

// The cache entry of a caller's matrix with the requested device copy in it: one GPU (want_team
// false) or row blocks over the GPU team.  Validation is the same for both (policy above).
MatrixEntry* matrix_entry(struct __mv_sparse* A, bool pin, bool want_team, PendingCheck* pc = nullptr) {
  if (!A->row_ptr || !A->values || (!A->col_indices && A->nnz > 0)) {
    set_error("matrix struct lacks CSR arrays");
    return nullptr;
  }
  if (ctx().world > 1 && !ctx().team_primary) {
    set_error("host-struct entry points are single-rank; use the cgb200_system_* API when distributed");
    return nullptr;
  }
  const int policy = cache_policy();
  const uint64_t fp = fingerprint(A);
  uint64_t content = 0;
  bool have_content = false;
  MatrixEntry* e = nullptr;
  auto it = g_mats.find(A);
  if (it != g_mats.end()) {
    MatrixEntry& old = it->second;
    bool same = policy != CGB200_CACHE_OFF && old.values == A->values && old.cols == A->col_indices &&
                old.rowptr == A->row_ptr && old.size == A->size && old.nnz == A->nnz && old.fp == fp;
    const bool trusted = old.pinned || g_mode == CGB200_RESIDENT || policy == CGB200_CACHE_SAMPLED;
    if (same && !trusted) {
      bool settled = false;
      if (policy == CGB200_CACHE_TRACKED && old.watched) {
        // nothing written to the watched interiors and the unwatched edges hash alike: unchanged,
        // exactly, without reading the arrays
        HostArr a[3];
        host_arrays(A, a);
        if (pc && (want_team ? old.team != nullptr : old.sys != nullptr)) {  // asked while the device works: PendingCheck
          pc->entry = &old;
          pc->started = true;
          const HostArr a0 = a[0], a1 = a[1], a2 = a[2];
          const uint64_t s0 = old.seen[0], s1 = old.seen[1], s2 = old.seen[2];
          pc->th = std::thread([pc, a0, a1, a2, s0, s1, s2]() {
            const HostArr aa[3] = {a0, a1, a2};
            const uint64_t ss[3] = {s0, s1, s2};
            pc->w = watch_state(aa, ss);
            pc->edge = pc->w == 0 ? edge_hash(aa) : 0;
          });
          return &old;
        }
        const int w = watch_state(a, old.seen);
        if (w == 0 && edge_hash(a) == old.edge) {
          ++g_stat_tracked_clean;
          settled = true;
        } else if (w != 0) {
          // written (perhaps with the same values) or unknown: the hash decides.  The marks are
          // cleared BEFORE the arrays are read, so a write racing with the hash is seen next time.
          ++g_stat_tracked_dirty;
          old.watched = watch_refresh(a, old.seen);
          if (old.watched) old.edge = edge_hash(a);
        }
      }
      if (!settled) {
        content = content_hash(A);
        have_content = true;
        ++g_stat_hashes;
        same = old.has_content && old.content == content;
        if (same && policy == CGB200_CACHE_TRACKED && !old.watched) {  // try to watch from here on
          HostArr a[3];
          host_arrays(A, a);
          old.watched = watch_acquire(a, old.seen);
          if (old.watched) {  // armed after the hash read the arrays: one more look at them settles it
            old.edge = edge_hash(a);
            if (content_hash(A) != content) same = false;
          }
        }
      }
    }
    if (same) {
      e = &old;
      if (pin) e->pinned = true;
    } else {
      release_entry(old);
      g_mats.erase(it);
    }
  }
  std::thread hasher;  // the hash for the next call's comparison is taken while the upload runs
  bool fresh = false;
  const bool trace = getenv("CGB200_TRACE_UPLOAD") != nullptr;  // first-call timing to stderr
  const auto t0 = std::chrono::steady_clock::now();
  auto ms_since = [&](std::chrono::steady_clock::time_point a) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count();
  };
  double t_arm = 0.0, t_up = 0.0;
  if (!e) {
    MatrixEntry ne;
    ne.values = A->values;
    ne.cols = A->col_indices;
    ne.rowptr = A->row_ptr;
    ne.size = A->size;
    ne.nnz = A->nnz;
    ne.fp = fp;
    ne.content = content;
    ne.has_content = have_content;
    ne.pinned = pin;
    e = &(g_mats[A] = ne);
    fresh = true;
    ++g_stat_uploads;
    if (policy == CGB200_CACHE_TRACKED) {
      // armed BEFORE the arrays are read for the upload and the hash: a write that races with them
      // is seen at the next call (and costs a re-upload) instead of being missed
      HostArr a[3];
      host_arrays(A, a);
      e->watched = watch_acquire(a, e->seen);
      if (e->watched) e->edge = edge_hash(a);
      t_arm = ms_since(t0);
    }
    if ((policy == CGB200_CACHE_STRICT || policy == CGB200_CACHE_TRACKED) && !have_content) {
      ++g_stat_hashes;
      hasher = std::thread([&]() { content = content_hash(A); });
    }
  }
  bool ok = true;
  if (want_team && !e->team) {
    e->team = team_create(A);
    ok = e->team != nullptr;
  } else if (!want_team && !e->sys) {
    SoloScope solo;  // (the primary of a GPU team is in a multi-rank world: this copy is not)
    e->sys = system_from_csr(A->size, 0, A->size, A->row_ptr, A->col_indices, A->values, nullptr);
    ok = e->sys != nullptr;
  }
  t_up = ms_since(t0);
  if (hasher.joinable()) {
    hasher.join();
    e->content = content;
    e->has_content = true;
  }
  if (trace && fresh)
    fprintf(stderr, "[cgb200] matrix %lld x %lld, %lld entries: write tracking armed %.1f ms, on the device %.1f ms, "
                    "content hash joined %.1f ms%s\n", (long long)A->size, (long long)A->size, (long long)A->nnz, t_arm,
            t_up, ms_since(t0), e->watched ? " (watched)" : "");
  if (!ok) {
    if (fresh || (!e->sys && !e->team)) {
      release_entry(*e);
      g_mats.erase(A);
    }
    return nullptr;
  }
  return e;
}

// Device system for a host CSR struct, uploaded on first use and whenever the struct's arrays no
// longer hold what was uploaded (policy above).  pin: the caller hands the matrix over explicitly.
System* matrix_system(struct __mv_sparse* A, bool pin = false, PendingCheck* pc = nullptr) {
  MatrixEntry* e = matrix_entry(A, pin, false, pc);
  return e ? e->sys : nullptr;
}

// Several GPUs behind the host-struct solve (team.h): asked for, worth it, and the team is up.
bool want_team(const struct __mv_sparse* A) {
  const int g = team_requested();
  if (g <= 1 || A->size < 64 * g || !A->row_ptr) return false;
  if (ctx().world > 1 && !ctx().team_primary) return false;
  if (team_active()) return true;
  static bool warned = false;
  if (team_start() == 0) return true;
  if (!warned) {
    warned = true;
    fprintf(stderr, "[cgb200] CGB200_GPUS=%d: %s -- solving on one GPU\n", g, last_error());
  }
  return false;
}

std::vector<const void*> g_lazy;  // structs whose mirror is a deferred scale (a handful at most)

void lazy_forget(const void* obj) {
  for (size_t i = 0; i < g_lazy.size(); ++i)
    if (g_lazy[i] == obj) {
      g_lazy[i] = g_lazy.back();
      g_lazy.pop_back();
      return;
    }
}

// Computes the values of a deferred scale into its own buffer.
int materialize(const void* obj) {
  auto it = g_vecs.find(obj);
  if (it == g_vecs.end() || !it->second.lazy) return 0;
  VectorEntry& e = it->second;
  auto is = g_vecs.find(e.lazy_src);
  if (is == g_vecs.end() || !is->second.d || !e.d) {
    set_error("internal: deferred scale lost its source");
    return -1;
  }
  if (e.n > 0) CU_OK(launch_ewise(0, e.lazy_s, is->second.d, nullptr, e.d, e.n, ctx().stream));
  e.lazy = false;
  e.lazy_src = nullptr;
  lazy_forget(obj);
  return 0;
}

// obj's device values are about to change or go away: whoever defers a scale of them computes now.
int before_write(const void* obj) {
  if (g_lazy.empty()) return 0;
  for (size_t i = 0; i < g_lazy.size();) {
    const void* l = g_lazy[i];
    auto it = g_vecs.find(l);
    if (l != obj && it != g_vecs.end() && it->second.lazy && it->second.lazy_src == obj) {
      if (materialize(l)) return -1;  // removes l from g_lazy: look at slot i again
      continue;
    }
    ++i;
  }
  return 0;
}

// Device vectors released by free_mv_struct / cgb200_free_all / a resize are kept for the next
// operator output of the same capacity: a caller written like the reference's loop creates and
// drops two vectors per iteration (cg.c:117,120), and a cudaMalloc + cudaFree pair per vector
// (the free synchronises the device) would cost more than the operators themselves on small
// systems.  Bounded: at most 16 buffers / 4 GiB; flushed on allocation failure and at shutdown.
std::multimap<int, double*> g_vec_pool;  // capacity (doubles) -> buffer
size_t g_vec_pool_bytes = 0;
constexpr size_t VEC_POOL_MAX_BYTES = 4ull << 30;
constexpr size_t VEC_POOL_MAX_COUNT = 16;

void vec_pool_flush() {
  for (auto& kv : g_vec_pool) dev_free(kv.second);
  g_vec_pool.clear();
  g_vec_pool_bytes = 0;
}

void vec_release(double* d, int cap) {
  if (!d) return;
  const size_t bytes = ((size_t)cap + VEC_PAD) * sizeof(double);
  // guard mode frees at once: the guards are checked at the free, and a reused buffer would hold
  // stale values instead of the poison that exposes uninitialised reads
  if (dev_guard_enabled() || g_vec_pool.size() >= VEC_POOL_MAX_COUNT || g_vec_pool_bytes + bytes > VEC_POOL_MAX_BYTES) {
    dev_free(d);
    return;
  }
  g_vec_pool.emplace(cap, d);
  g_vec_pool_bytes += bytes;
}

void drop_mirror(const void* obj) {
  auto im = g_mats.find(obj);
  if (im != g_mats.end()) {
    release_entry(im->second);
    g_mats.erase(im);
  }
  if (!g_lazy.empty()) {
    before_write(obj);  // a deferred scale of these values is computed before they go
    lazy_forget(obj);
  }
  auto iv = g_vecs.find(obj);
  if (iv != g_vecs.end()) {
    vec_release(iv->second.d, iv->second.cap);
    g_vecs.erase(iv);
  }
}

int ensure_vec_capacity(VectorEntry& e, int n) {
  if (e.cap < n || !e.d) {
    if (e.d) vec_release(e.d, e.cap);
    e.d = nullptr;
    e.cap = 0;
    auto it = g_vec_pool.find(n);  // exact capacity only: a pooled buffer never grows a footprint
    if (it != g_vec_pool.end()) {
      e.d = it->second;
      g_vec_pool_bytes -= ((size_t)n + VEC_PAD) * sizeof(double);
      g_vec_pool.erase(it);
    } else {
      cudaError_t err = DEV_ALLOC(&e.d, ((size_t)n + VEC_PAD) * sizeof(double));
      if (err != cudaSuccess && !g_vec_pool.empty()) {
        cudaGetLastError();
        vec_pool_flush();
        err = DEV_ALLOC(&e.d, ((size_t)n + VEC_PAD) * sizeof(double));
      }
      CU_OK(err);
    }
    e.cap = n;
  }
  e.n = n;
  return 0;
}

// Device pointer holding the current values of an input vector.
int dev_input(struct __mv_sparse* v, int slot, const double** out) {
  Ctx& c = ctx();
  const int n = v->size;
  if (g_mode == CGB200_RESIDENT) {
    VectorEntry& e = g_vecs[v];
    if (e.dev_valid && e.n == n && e.host_ptr == v->values) {
      if (e.lazy && materialize(v)) return -1;
      *out = e.d;
      return 0;
    }
    if (before_write(v)) return -1;
    if (e.lazy) {
      e.lazy = false;
      lazy_forget(v);
    }
    if (ensure_vec_capacity(e, n)) return -1;
    if (n > 0)
      CU_OK(upload_host(e.d, v->values, sizeof(double) * n, c.stream));
    e.dev_valid = true;
    e.host_valid = true;
    e.sd_valid = false;
    e.host_ptr = v->values;
    *out = e.d;
    return 0;
  }
  double* d;
  if (scratch(slot, n, &d)) return -1;
  if (n > 0)
    CU_OK(upload_host(d, v->values, sizeof(double) * n, c.stream));
  *out = d;
  return 0;
}

// The reference's allocate-or-reuse rule for outputs (mv_ops.c:141-152 and its twins).
int prepare_out(struct __mv_sparse** out, int size, int nnz) {
  if (*out == NULL) {
    *out = (struct __mv_sparse*)calloc(1, sizeof(struct __mv_sparse));
    if (!*out) return -1;
    (*out)->values = (double*)calloc(size > 0 ? size : 1, sizeof(double));
    (*out)->col_indices = NULL;
    (*out)->row_ptr = NULL;
  } else {
    (*out)->values = (double*)realloc((*out)->values, sizeof(double) * (size > 0 ? size : 1));
  }
  if (!(*out)->values) {
    set_error("host allocation failed");
    return -1;
  }
  (*out)->size = size;
  (*out)->nnz = nnz;
  return 0;
}

// Device pointer an op writes its result to (may coincide with an input mirror: in-place).
int dev_output(struct __mv_sparse* o, int slot, double** out) {
  const int n = o->size;
  if (g_mode == CGB200_RESIDENT) {
    if (before_write(o)) return -1;
    VectorEntry& e = g_vecs[o];
    if (e.lazy) {  // overwritten before anybody read it: the scale is never computed
      e.lazy = false;
      e.lazy_src = nullptr;
      lazy_forget(o);
    }
    if (e.cap < n || !e.d) {
      // growing must keep nothing: the op overwrites all n elements
      if (ensure_vec_capacity(e, n)) return -1;
    }
    e.n = n;
    *out = e.d;
    return 0;
  }
  return scratch(slot, n, out);
}

// Publishes an op's result: eager -> copy to the host array now; resident -> mark the mirror.
int finish_output(struct __mv_sparse* o, double* d) {
  Ctx& c = ctx();
  if (g_mode == CGB200_RESIDENT) {
    VectorEntry& e = g_vecs[o];
    e.dev_valid = true;
    e.host_valid = false;
    e.sd_valid = false;
    e.host_ptr = o->values;
    return 0;
  }
  if (o->size > 0)
    CU_OK(download_host(o->values, d, sizeof(double) * o->size, c.stream));
  CU_OK(cudaStreamSynchronize(c.stream));
  return 0;
}

int sync_to_host(struct __mv_sparse* o) {
  auto it = g_vecs.find(o);
  if (it == g_vecs.end()) return 0;
  VectorEntry& e = it->second;
  if (e.lazy && materialize(o)) return -1;
  if (e.dev_valid && !e.host_valid && e.n == o->size && o->values && e.host_ptr == o->values) {
    Ctx& c = ctx();
    if (e.n > 0)
      CU_OK(download_host(o->values, e.d, sizeof(double) * e.n, c.stream));
    CU_OK(cudaStreamSynchronize(c.stream));
    e.host_valid = true;
  }
  return 0;
}

// Resident mode's two shortcuts (VectorEntry::lazy).  Returns 0 when the call has been dealt with,
// -1 on failure, 1 when the general path should run.
//   sv_mult(s, a, &out), out distinct from a: nothing is launched; out's mirror becomes "s * a".
//   vec_add / vec_sub whose second operand is such a deferred scale: one pass forms
//   a +/- fl(s * src); when the output is the scale's source (p = r + beta p through
//   sres = beta p, cg.c:131-132) the same pass also stores the scaled vector.
int ewise_resident(int op, double sca, struct __mv_sparse* a, struct __mv_sparse* b,
                   struct __mv_sparse** out) {
  Ctx& c = ctx();
  const int n = a->size, nnz = a->nnz;
  if (op == 0) {
    if (*out == a || n <= 0) return 1;
    const double* da = nullptr;
    if (dev_input(a, 0, &da)) return -1;  // a on the device, computed if it was deferred itself
    if (prepare_out(out, n, nnz)) return -1;
    double* dout = nullptr;
    if (dev_output(*out, 2, &dout)) return -1;  // buffer for a later materialize; drops old state
    VectorEntry& e = g_vecs[*out];
    e.dev_valid = true;
    e.host_valid = false;
    e.sd_valid = false;
    e.host_ptr = (*out)->values;
    e.lazy = true;
    e.lazy_s = sca;
    e.lazy_src = a;
    g_lazy.push_back(*out);
    return 0;
  }
  auto ib = g_vecs.find(b);
  if (ib == g_vecs.end() || !ib->second.lazy || !ib->second.dev_valid || ib->second.n != n || b == a)
    return 1;
  const void* src = ib->second.lazy_src;
  const double s = ib->second.lazy_s;
  auto is = g_vecs.find(src);
  if (is == g_vecs.end() || !is->second.d) return 1;
  const double* da = nullptr;
  if (dev_input(a, 0, &da)) return -1;
  ib = g_vecs.find(b);  // (dev_input may have rehashed the table or touched what b defers)
  if (ib == g_vecs.end() || !ib->second.lazy || ib->second.lazy_src != src) return 1;
  if (src == (const void*)*out) {
    // the output is the scale's source (p = r + beta p through sres = beta p, cg.c:131-132): one
    // pass stores the scaled vector where b's values belong and the sum over the source
    VectorEntry& eb = ib->second;
    double* db = eb.d;
    eb.lazy = false;  // its values exist once the kernel below has run
    eb.lazy_src = nullptr;
    lazy_forget(b);
    if (prepare_out(out, n, nnz)) return -1;
    double* dout = nullptr;
    if (dev_output(*out, 2, &dout)) return -1;  // computes whoever else defers a scale of *out
    if (*out == a) da = dout;
    CU_OK(launch_ewise(op == 1 ? 5 : 6, s, da, dout, dout, n, c.stream, db));
    return finish_output(*out, dout);
  }
  if (prepare_out(out, n, nnz)) return -1;
  double* dout = nullptr;
  if (dev_output(*out, 2, &dout)) return -1;  // (*out == b: b stops being deferred, it is overwritten)
  if (*out == a) da = dout;
  const double* dsrc = g_vecs.find(src)->second.d;
  CU_OK(launch_ewise(op == 1 ? 3 : 4, s, da, dsrc, dout, n, c.stream));
  return finish_output(*out, dout);
}

int ewise(int op, double sca, struct __mv_sparse* a, struct __mv_sparse* b, struct __mv_sparse** out) {
  if (!a || (op != 0 && !b) || !out) return -1;
  if (op != 0 && a->size != b->size) return -1;
  if (ensure_init(-1)) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  Ctx& c = ctx();
  const int n = a->size, nnz = a->nnz;
  if (g_mode == CGB200_RESIDENT) {
    const int rc = ewise_resident(op, sca, a, b, out);
    if (rc <= 0) return rc;  // done (0) or failed (-1); 1 = take the general path below
  }
  // inputs first: *out may alias an input whose host array prepare_out() is about to realloc
  const double *da = nullptr, *db = nullptr;
  if (dev_input(a, 0, &da)) return -1;
  if (op != 0 && dev_input(b, 1, &db)) return -1;
  if (g_mode == CGB200_EAGER) CU_OK(cudaStreamSynchronize(c.stream));
  if (prepare_out(out, n, nnz)) return -1;
  double* dout = nullptr;
  if (dev_output(*out, 2, &dout)) return -1;
  if (g_mode == CGB200_RESIDENT) {  // dev_output may have re-allocated an aliased mirror
    if (*out == a) da = dout;
    if (op != 0 && *out == b) db = dout;
  }
  CU_OK(launch_ewise(op, sca, da, db, dout, n, c.stream));
  return finish_output(*out, dout);
}

int solve_host(struct __mv_sparse* A, struct __mv_sparse* b, struct __mv_sparse** x,
               const SolveOpts& o, SolveResult* res) {
  if (!A || !b || !x) {
    set_error("NULL argument");
    return -1;
  }
  if (A->size != b->size) {
    set_error("A and b sizes differ");
    return -1;
  }
  if (ensure_init(-1)) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  Ctx& c = ctx();
  if (want_team(A)) {  // CGB200_GPUS > 1: the same call, the matrix cut into row blocks over the GPU team
    PendingCheck pc;  // (tracked policy: A's arrays are checked while the team already solves)
    MatrixEntry* e = matrix_entry(A, false, true, &pc);
    if (e && e->team) {
      if (sync_to_host(b)) return -1;
      const int n = b->size;
      const double* x0 = nullptr;
      if (o.flags & CGB200_OPT_X0) {
        if (!*x || (*x)->size != n || !(*x)->values) {
          set_error("CGB200_OPT_X0: *x must be a vector of A's size holding the initial guess");
          return -1;
        }
        if (sync_to_host(*x)) return -1;
        x0 = (*x)->values;
      }
      struct __mv_sparse* xv = (struct __mv_sparse*)calloc(1, sizeof(struct __mv_sparse));
      if (!xv) return -1;
      xv->size = n;
      xv->nnz = n;
      xv->values = (double*)calloc(n > 0 ? n : 1, sizeof(double));
      bool ok = xv->values != nullptr && team_solve(e->team, b->values, x0, xv->values, o, res) == 0;
      if (ok && !pc.settle()) {
        // A was written since it was cut and uploaded: that solve is dropped; validate the ordinary
        // way (hash, re-cut and upload if different) and solve again
        e = matrix_entry(A, false, true);
        ok = e && e->team && team_solve(e->team, b->values, x0, xv->values, o, res) == 0;
      }
      if (!ok) {
        free(xv->values);
        free(xv);
        return -1;
      }
      *x = xv;
      return 0;
    }
    if (team_active()) return -1;  // the team is fine, this matrix was refused: the reason is in last_error
    // (the team broke while taking the matrix: fall through to one GPU)
  }
  PendingCheck pc;
  System* sys = matrix_system(A, false, &pc);
  if (!sys) return -1;
  if (sync_to_host(b)) return -1;
  const int n = b->size;
  if (o.flags & CGB200_OPT_X0) {  // extension: *x holds the initial guess (the reference starts at 0)
    if (!*x || (*x)->size != n || !(*x)->values) {
      set_error("CGB200_OPT_X0: *x must be a vector of A's size holding the initial guess");
      return -1;
    }
    if (sync_to_host(*x)) return -1;
  }
  for (int attempt = 0;; ++attempt) {
    if (n > 0) CU_OK(upload_host(sys->b, b->values, sizeof(double) * n, c.stream));
    if ((o.flags & CGB200_OPT_X0) && n > 0) CU_OK(upload_host(sys->x, (*x)->values, sizeof(double) * n, c.stream));
    if (system_solve(sys, o, res)) return -1;
    if (attempt > 0 || pc.settle()) break;
    // the arrays were written since the upload (or the kernel cannot tell): that solve ran on a
    // stale copy and is dropped -- validate the ordinary way (hash, upload if different), again
    sys = matrix_system(A);
    if (!sys) return -1;
  }
  // the reference hands back a freshly allocated vector (cg.c:104,138)
  struct __mv_sparse* xv = (struct __mv_sparse*)calloc(1, sizeof(struct __mv_sparse));
  if (!xv) return -1;
  xv->size = n;
  xv->nnz = n;
  xv->values = (double*)calloc(n > 0 ? n : 1, sizeof(double));
  if (!xv->values) {
    free(xv);
    return -1;
  }
  if (n > 0 && system_get_x(sys, xv->values)) {
    free(xv->values);
    free(xv);
    return -1;
  }
  *x = xv;
  return 0;
}

}  // namespace

// =============================================================================================
// Part 1: the reference surface
// =============================================================================================

extern "C" {

struct __mv_sparse* new_mv_struct(void) {
  return (struct __mv_sparse*)calloc(1, sizeof(struct __mv_sparse));
}

struct __mv_sparse* new_mv_struct_with_size(int size) {
  struct __mv_sparse* v = new_mv_struct();
  if (!v) return NULL;
  v->size = size;
  v->nnz = size;
  v->values = (double*)calloc(size > 0 ? size : 1, sizeof(double));
  return v;
}

void free_mv_struct(struct __mv_sparse* obj) {
  if (!obj) return;
  {
    std::lock_guard<std::mutex> lk(g_tab_mu);
    drop_mirror(obj);
  }
  free(obj);  // arrays stay with the caller, as in the reference (mv_ops.c:41)
}

struct __mv_sparse* mv_deep_copy(struct __mv_sparse* orig) {
  if (!orig) return NULL;
  {
    std::lock_guard<std::mutex> lk(g_tab_mu);
    // resident mode, vector whose current values live on the device (cg.c:117,120 copy x and r
    // every iteration): copy device to device and leave the host array of the copy stale, like
    // every other resident-mode output -- no transfer, no synchronisation
    auto it = g_vecs.find(orig);
    if (it != g_vecs.end() && it->second.lazy && materialize(orig)) return NULL;
    if (g_mode == CGB200_RESIDENT && !orig->col_indices && !orig->row_ptr && it != g_vecs.end() &&
        it->second.dev_valid && !it->second.host_valid && it->second.n == orig->size &&
        it->second.host_ptr == orig->values) {
      struct __mv_sparse* cp = new_mv_struct();
      if (!cp) return NULL;
      cp->size = orig->size;
      cp->nnz = orig->nnz;
      cp->values = (double*)calloc(orig->nnz > 0 ? (size_t)orig->nnz : 1, sizeof(double));
      if (!cp->values) {
        free(cp);
        return NULL;
      }
      const double* src = it->second.d;  // (taken before g_vecs[cp] may rehash the table)
      const int n = it->second.n;
      const bool sd_valid = it->second.sd_valid;
      const double sd = it->second.sd;
      VectorEntry& e = g_vecs[cp];
      if (ensure_vec_capacity(e, n) ||
          (n > 0 && cudaMemcpyAsync(e.d, src, sizeof(double) * n, cudaMemcpyDeviceToDevice,
                                    ctx().stream) != cudaSuccess)) {
        set_error("mv_deep_copy: device copy failed");
        drop_mirror(cp);
        free(cp->values);
        free(cp);
        return NULL;
      }
      e.dev_valid = true;
      e.host_valid = false;
      e.sd_valid = sd_valid;
      e.sd = sd;
      e.host_ptr = cp->values;
      return cp;
    }
    if (sync_to_host(orig)) return NULL;
  }
  struct __mv_sparse* cp = new_mv_struct();
  if (!cp) return NULL;
  cp->size = orig->size;
  cp->nnz = orig->nnz;
  const size_t nv = orig->nnz > 0 ? (size_t)orig->nnz : 1;
  cp->values = (double*)calloc(nv, sizeof(double));
  if (orig->values && orig->nnz > 0) memcpy(cp->values, orig->values, sizeof(double) * orig->nnz);
  if (orig->col_indices) {
    cp->col_indices = (int*)calloc(nv, sizeof(int));
    if (orig->nnz > 0) memcpy(cp->col_indices, orig->col_indices, sizeof(int) * orig->nnz);
  }
  if (orig->row_ptr) {
    cp->row_ptr = (int*)calloc((size_t)orig->size + 1, sizeof(int));
    memcpy(cp->row_ptr, orig->row_ptr, sizeof(int) * ((size_t)orig->size + 1));
  }
  return cp;
}

void print_sparse(struct __mv_sparse* obj) {
  printf("Sparse Object:\n");
  if (!obj) {
    printf("\tObject is NULL\n");
    return;
  }
  {
    std::lock_guard<std::mutex> lk(g_tab_mu);
    sync_to_host(obj);
  }
  printf("\tSize: %d\n", obj->size);
  printf("\tNNZ: %d\n", obj->nnz);
  printf("\tValues: %p\n", (void*)obj->values);
  for (int i = 0; i < obj->nnz; i++) printf("\t%f\n", obj->values[i]);
}

int mat_get_row(struct __mv_sparse* mat_A, int row_id, double* p_row) {
  if (!mat_A || !p_row) return -1;
  if (!mat_A->row_ptr || row_id < 0 || row_id >= mat_A->size) return -1;
  memset(p_row, 0, sizeof(double) * (size_t)mat_A->size);
  for (int k = mat_A->row_ptr[row_id]; k < mat_A->row_ptr[row_id + 1]; ++k) {
    const int c = mat_A->col_indices[k];
    if (c >= 0 && c < mat_A->size) p_row[c] = mat_A->values[k];
  }
  return 0;
}

double dot_product(struct __mv_sparse* vec_a, struct __mv_sparse* vec_b) {
  if (!vec_a || !vec_b) return -1.0;
  if (vec_a->size != vec_b->size) return -1.0;
  if (ensure_init(-1)) return -1.0;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  Ctx& c = ctx();
  const double *da = nullptr, *db = nullptr;
  if (dev_input(vec_a, 0, &da)) return -1.0;
  const bool self = vec_b == vec_a && g_mode == CGB200_RESIDENT;
  if (self) {
    const VectorEntry& e = g_vecs[vec_a];
    if (e.sd_valid) return e.sd;
  }
  if (vec_b == vec_a)
    db = da;
  else if (dev_input(vec_b, 1, &db))
    return -1.0;
  double* dres = (double*)c.d_scratch;
  if (launch_dot(da, db, vec_a->size, c.partials, &c.d_state->ticket[0], dres, c.stream) !=
          cudaSuccess ||
      cudaMemcpyAsync(c.h_scratch, dres, sizeof(double), cudaMemcpyDeviceToHost, c.stream) !=
          cudaSuccess ||
      cudaStreamSynchronize(c.stream) != cudaSuccess) {
    set_error(std::string("dot_product: ") + cudaGetErrorString(cudaGetLastError()));
    return -1.0;
  }
  const double v = *(double*)c.h_scratch;
  if (self) {
    VectorEntry& e = g_vecs[vec_a];
    e.sd = v;
    e.sd_valid = true;
  }
  return v;
}

int sv_mult(double s, struct __mv_sparse* vec_a, struct __mv_sparse** out) {
  return ewise(0, s, vec_a, NULL, out);
}

int vec_add(struct __mv_sparse* vec_a, struct __mv_sparse* vec_b, struct __mv_sparse** out) {
  return ewise(1, 0.0, vec_a, vec_b, out);
}

int vec_sub(struct __mv_sparse* vec_a, struct __mv_sparse* vec_b, struct __mv_sparse** out) {
  return ewise(2, 0.0, vec_a, vec_b, out);
}

int mv_mult(struct __mv_sparse* mat_A, struct __mv_sparse* vec_b, struct __mv_sparse** out) {
  if (!mat_A || !vec_b || !out) return -1;
  if (mat_A->size != vec_b->size) return -1;
  if (ensure_init(-1)) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  Ctx& c = ctx();
  PendingCheck pc;  // (tracked policy: the kernel is asked about A's arrays while the product runs)
  System* sys = matrix_system(mat_A, false, &pc);
  if (!sys) return -1;
  const int n = mat_A->size;
  const double* db = nullptr;
  if (dev_input(vec_b, 0, &db)) return -1;
  if (g_mode == CGB200_EAGER) CU_OK(cudaStreamSynchronize(c.stream));
  if (prepare_out(out, n, vec_b->nnz)) return -1;
  if (n == 0) return 0;
  double* dout = nullptr;
  if (dev_output(*out, 2, &dout)) return -1;
  if (system_matvec_into(sys, db, dout)) return -1;
  if (finish_output(*out, dout)) return -1;
  if (pc.settle()) return 0;
  // A was written since its upload: that product is dropped; validate the ordinary way and redo it
  sys = matrix_system(mat_A);
  if (!sys) return -1;
  if (system_matvec_into(sys, db, dout)) return -1;
  return finish_output(*out, dout);
}

int conj_grad(int max_iter, struct __mv_sparse* mat_A, struct __mv_sparse* vec_b,
              struct __mv_sparse** x) {
  SolveOpts o;
  o.max_iter = max_iter;
  o.tol = 0.0;
  o.variant = CGB200_CLASSIC;
  SolveResult r;
  return solve_host(mat_A, vec_b, x, o, &r);
}

// =============================================================================================
// Part 2: extensions
// =============================================================================================

int cgb200_init(int device) { return ensure_init(device); }

void cgb200_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  for (auto& kv : g_mats) release_entry(kv.second);
  g_mats.clear();
  team_stop();  // the workers leave with the last matrix
  for (auto& kv : g_vecs) dev_free(kv.second.d);
  vec_pool_flush();
  g_vecs.clear();
  g_lazy.clear();
  for (int i = 0; i < 3; ++i) {
    if (g_scratch[i]) dev_free(g_scratch[i]);
    g_scratch[i] = nullptr;
    g_scratch_cap[i] = 0;
  }
  shutdown();
}

const char* cgb200_last_error(void) { return last_error(); }
long long cgb200_launch_count(void) { return launch_count(); }
long long cgb200_guard_check(void) { return dev_guard_check(); }
int cgb200_guard_selftest(void) {
  if (ensure_init(-1)) return -1;
  return dev_guard_selftest();
}

int cgb200_device_available(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n > 0 ? 1 : 0;
}

int cgb200_set_mode(int mode) {
  if (mode != CGB200_EAGER && mode != CGB200_RESIDENT) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (mode == CGB200_EAGER && g_mode == CGB200_RESIDENT) {
    // leaving resident mode: nothing may stay newer on the device than on the host
    for (auto& kv : g_vecs) {
      struct __mv_sparse* o = (struct __mv_sparse*)kv.first;
      if (kv.second.dev_valid && !kv.second.host_valid) sync_to_host(o);
      kv.second.dev_valid = false;
      kv.second.sd_valid = false;
    }
  }
  g_mode = mode;
  return 0;
}

int cgb200_set_matrix_cache(int policy) {
  if (policy != CGB200_CACHE_STRICT && policy != CGB200_CACHE_SAMPLED && policy != CGB200_CACHE_OFF &&
      policy != CGB200_CACHE_TRACKED)
    return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  g_cache_policy = policy;
  return 0;
}

int cgb200_write_tracking_available(void) { return ww_available() ? 1 : 0; }

int cgb200_write_tracking_op(const void* p, size_t bytes, int op) {
  if (!ww_available()) return -1;
  switch (op) {
    case 0: return ww_arm(p, bytes) ? 0 : -1;
    case 1: return ww_written(p, bytes);
    case 2: return ww_rearm(p, bytes) ? 0 : -1;
    case 3: ww_disarm(p, bytes); return 0;
    default: return -1;
  }
}

int cgb200_cache_stats(long long* out5) {
  if (!out5) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  out5[0] = g_stat_uploads;
  out5[1] = g_stat_hashes;
  out5[2] = g_stat_tracked_clean;
  out5[3] = g_stat_tracked_dirty;
  out5[4] = ww_available() ? 1 : 0;
  return 0;
}

unsigned long long cgb200_content_hash(const struct __mv_sparse* A) {
  if (!A || !A->values) return 0ull;
  if (!A->row_ptr) {  // a vector: its values only
    return hash_array(A->values, (A->size > 0 ? (size_t)A->size : 0) * sizeof(double), 1);
  }
  return content_hash(A);
}

int cgb200_set_gpus(int gpus) {
  if (gpus < 1 || gpus > MAX_RANKS) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (gpus != team_requested() && team_active()) {  // another team size: the old team's copies go first
    for (auto& kv : g_mats)
      if (kv.second.team) {
        team_destroy(kv.second.team);
        kv.second.team = nullptr;
      }
    team_stop();
  }
  team_request(gpus);
  return 0;
}

int cgb200_gpus_in_use(void) { return team_active() ? ctx().world : 1; }

int cgb200_worker_main(int argc, char** argv) { return worker_main(argc, argv); }

int cgb200_to_device(struct __mv_sparse* obj) {
  if (!obj) return -1;
  if (ensure_init(-1)) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!is_vector(obj))  // pinned until invalidate / free; with a GPU team the blocks go to their GPUs now
    return matrix_entry(obj, true, want_team(obj)) ? 0 : -1;
  if (before_write(obj)) return -1;
  VectorEntry& e = g_vecs[obj];
  if (e.lazy) {
    e.lazy = false;
    lazy_forget(obj);
  }
  if (ensure_vec_capacity(e, obj->size)) return -1;
  if (obj->size > 0)
    CU_OK(cudaMemcpyAsync(e.d, obj->values, sizeof(double) * obj->size, cudaMemcpyHostToDevice,
                          ctx().stream));
  CU_OK(cudaStreamSynchronize(ctx().stream));
  e.dev_valid = true;
  e.host_valid = true;
  e.sd_valid = false;
  e.host_ptr = obj->values;
  return 0;
}

int cgb200_to_host(struct __mv_sparse* obj) {
  if (!obj) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return sync_to_host(obj);
}

int cgb200_invalidate(struct __mv_sparse* obj) {
  if (!obj) return -1;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  drop_mirror(obj);
  return 0;
}

void cgb200_free_all(struct __mv_sparse* obj) {
  if (!obj) return;
  {
    std::lock_guard<std::mutex> lk(g_tab_mu);
    drop_mirror(obj);
  }
  free(obj->values);
  free(obj->col_indices);
  free(obj->row_ptr);
  free(obj);
}

static void fill_result(const SolveResult& r, cgb200_result* res) {
  if (!res) return;
  res->iterations = r.iterations;
  res->status = r.status;
  res->rr = r.rr;
  res->bb = r.bb;
  res->relres = r.relres;
  res->seconds = r.seconds;
  res->bytes_per_iter = r.bytes_per_iter;
  res->path = r.path;
  res->world = r.world;
  res->kernel_launches = r.kernel_launches;
}

static SolveOpts to_opts(const cgb200_opts* o) {
  SolveOpts s;
  if (o) {
    s.max_iter = o->max_iter;
    s.tol = o->tol;
    s.variant = o->variant;
    s.check_every = o->check_every;
    s.flags = o->flags;
  }
  return s;
}

int cgb200_solve(struct __mv_sparse* A, struct __mv_sparse* b, struct __mv_sparse** x,
                 const cgb200_opts* opts, cgb200_result* res) {
  SolveResult r;
  memset(&r, 0, sizeof(r));
  const int rc = solve_host(A, b, x, to_opts(opts), &r);
  if (rc == 0) fill_result(r, res);
  return rc;
}

cgb200_system* cgb200_system_lap(int nx, int ny, int nz, int z0, int nzl, uint64_t bseed) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return (cgb200_system*)system_lap(nx, ny, nz, z0, nzl, bseed);
}

cgb200_system* cgb200_system_dense(int n, int row0, int nrows, uint64_t aseed, uint64_t bseed) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return (cgb200_system*)system_dense(n, row0, nrows, aseed, bseed);
}

cgb200_system* cgb200_system_from_csr(long long n_global, long long row0, int nrows,
                                      const int* row_ptr, const int* cols, const double* vals,
                                      const double* b) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return (cgb200_system*)system_from_csr(n_global, row0, nrows, row_ptr, cols, vals, b);
}

void cgb200_system_free(cgb200_system* sys) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys) return;
  cudaStreamSynchronize(ctx().stream);
  delete (System*)sys;
}

int cgb200_system_set_rhs(cgb200_system* sys, const double* b_host) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !b_host) return -1;
  return system_set_rhs((System*)sys, b_host);
}

int cgb200_system_gen_rhs(cgb200_system* sys, uint64_t bseed) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys) return -1;
  return system_gen_rhs((System*)sys, bseed);
}

int cgb200_system_set_x(cgb200_system* sys, const double* x_host) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !x_host) return -1;
  return system_set_x((System*)sys, x_host);
}

int cgb200_system_get_x(cgb200_system* sys, double* x_host) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !x_host) return -1;
  return system_get_x((System*)sys, x_host);
}

int cgb200_system_solve(cgb200_system* sys, const cgb200_opts* opts, cgb200_result* res) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  SolveResult r;
  memset(&r, 0, sizeof(r));
  const int rc = system_solve((System*)sys, to_opts(opts), &r);
  if (rc == 0) fill_result(r, res);
  return rc;
}

int cgb200_system_solve_async(cgb200_system* sys, const cgb200_opts* opts, long long* ticket) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return system_solve_async((System*)sys, to_opts(opts), ticket);
}

int cgb200_solve_wait(long long ticket, cgb200_result* res) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  SolveResult r;
  memset(&r, 0, sizeof(r));
  const int rc = system_solve_wait(ticket, &r);
  fill_result(r, res);
  return rc;
}

int cgb200_system_matvec(cgb200_system* sys, const double* v_host, double* s_host) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !v_host || !s_host) return -1;
  return system_matvec((System*)sys, v_host, s_host);
}

int cgb200_system_true_relres(cgb200_system* sys, double* relres) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !relres) return -1;
  return system_true_relres((System*)sys, relres);
}

int cgb200_system_time_kernel(cgb200_system* sys, int which, int reps, double* ms) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !ms || reps < 1 || which < 1 || which > 6) return -1;
  return system_time_kernel((System*)sys, which, reps, ms);
}

int cgb200_system_download(cgb200_system* sys, double* vals, int* cols, int* row_ptr) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys) return -1;
  return system_download((System*)sys, vals, cols, row_ptr);
}

int cgb200_system_get_b(cgb200_system* sys, double* b_host) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys || !b_host) return -1;
  return system_get_b((System*)sys, b_host);
}

int cgb200_timer_start(void) { return ensure_init(-1) ? -1 : timer_start(); }
int cgb200_timer_stop(double* ms) { return ms ? timer_stop(ms) : -1; }

int cgb200_system_bytes(cgb200_system* sys, double* iter_bytes, double* k1_bytes) {
  if (!sys) return -1;
  if (iter_bytes) *iter_bytes = ((System*)sys)->iter_bytes();
  if (k1_bytes) *k1_bytes = ((System*)sys)->k1_bytes();
  return 0;
}

double cgb200_system_comm_bytes(cgb200_system* sys) {
  return sys ? ((System*)sys)->comm_bytes_per_iter() : -1.0;
}

long long cgb200_system_rows(cgb200_system* sys) { return sys ? ((System*)sys)->nloc : -1; }
long long cgb200_system_nnz(cgb200_system* sys) { return sys ? ((System*)sys)->A.nnz : -1; }

int cgb200_plan_partition(int world, int rank, long long n_global, const long long* info4,
                          int dense, long long* out10) {
  if (!info4 || !out10 || world < 1) return -1;
  std::vector<PartInfo> parts(world);
  for (int g = 0; g < world; ++g)
    parts[g] = PartInfo{info4[4 * g], info4[4 * g + 1], info4[4 * g + 2], info4[4 * g + 3]};
  Plan p;
  const char* err = nullptr;
  if (plan_partition(world, rank, n_global, parts.data(), dense, &p, &err)) {
    set_error(err ? err : "plan_partition failed");
    return -1;
  }
  out10[0] = p.allgather;
  out10[1] = p.h_lo;
  out10[2] = p.h_hi;
  out10[3] = p.off;
  out10[4] = p.ext_len;
  out10[5] = p.col_shift;
  out10[6] = p.send_lo;
  out10[7] = p.send_hi;
  out10[8] = p.recv_lo;
  out10[9] = p.recv_hi;
  return 0;
}

int cgb200_plan_push(int world, int rank, long long n_global, const long long* info4, int dense,
                     long long* out) {
  if (!info4 || !out || world < 1 || world > 8) return -1;
  std::vector<PartInfo> parts(world);
  for (int g = 0; g < world; ++g)
    parts[g] = PartInfo{info4[4 * g], info4[4 * g + 1], info4[4 * g + 2], info4[4 * g + 3]};
  PushPlan pp;
  const char* err = nullptr;
  if (plan_push(world, rank, n_global, parts.data(), dense, &pp, &err)) {
    set_error(err ? err : "plan_push failed");
    return -1;
  }
  out[0] = pp.npush;
  for (int t = 0; t < 8; ++t) {
    const bool on = t < pp.npush;
    out[1 + 4 * t + 0] = on ? pp.push[t].peer : -1;
    out[1 + 4 * t + 1] = on ? pp.push[t].begin : 0;
    out[1 + 4 * t + 2] = on ? pp.push[t].end : 0;
    out[1 + 4 * t + 3] = on ? pp.push[t].dst_index : 0;
  }
  out[33] = pp.nwait;
  for (int i = 0; i < 8; ++i) out[34 + i] = i < pp.nwait ? pp.wait_src[i] : -1;
  return 0;
}

int cgb200_dist_unique_id(void* id128) { return id128 ? dist_unique_id(id128) : -1; }
int cgb200_dist_init(int rank, int world, const void* id128) {
  if (world > 1 && !id128) return -1;
  return dist_init(rank, world, id128);
}
int cgb200_dist_finalize(void) { return dist_finalize(); }
int cgb200_dist_rank(void) { return ctx().rank; }
int cgb200_dist_world(void) { return ctx().world; }
int cgb200_dist_barrier(void) { return dist_barrier(); }
const char* cgb200_dist_comm_mode(void) { return comm_mode_name(); }
int cgb200_system_use_pattern_format(cgb200_system* sys, int on) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return system_use_pattern((System*)sys, on);
}

int cgb200_system_use_banded_format(cgb200_system* sys, int on) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  return system_use_banded((System*)sys, on);
}

int cgb200_system_trace(cgb200_system* sys, int on, unsigned long long* out16) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  if (!sys) return -1;
  return system_trace((System*)sys, on, out16);
}

void cgb200_set_one_kernel_iteration(int on) { set_dense_iter_enabled(on); }
void cgb200_set_fused_vector_update(int on) { set_fused_vector_update(on); }

int cgb200_system_one_kernel_iteration(cgb200_system* sys) {
  return sys && dense_one_kernel((System*)sys) ? 1 : 0;
}

int cgb200_system_comm_is_p2p(cgb200_system* sys) { return sys && ((System*)sys)->p2p ? 1 : 0; }

}  // extern "C"
