// devmem.cu -- see devmem.h.
#include "devmem.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "solver.h"

namespace {

constexpr size_t GUARD = 4096;  // bytes on either side; keeps the payload 4 KiB aligned

struct Rec {
  void* base;
  size_t bytes;  // payload
  size_t total;
  const char* file;
  int line;
};

std::mutex g_mu;
std::unordered_map<void*, Rec> g_live;  // payload pointer -> record
std::unordered_map<void*, Rec> g_watch; // extra regions that must stay 0xFF (base = region, total = length)
long long g_damaged = 0;                // allocations found damaged when they were freed
int g_mode = -1;                        // -1 unknown, 0 off, 1 on

bool guard_on() {
  if (g_mode < 0) {
    const char* e = getenv("CGB200_GUARD");
    g_mode = (e && *e && strcmp(e, "0") != 0) ? 1 : 0;
  }
  return g_mode == 1;
}

// bytes of the two guard bands of `r` that no longer hold 0xFF (device must be idle)
size_t damaged_bytes(const Rec& r, size_t* first_off) {
  const size_t tail = r.total - GUARD - r.bytes;
  std::vector<unsigned char> h(GUARD + tail);
  if (cudaMemcpy(h.data(), r.base, GUARD, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
  if (cudaMemcpy(h.data() + GUARD, (char*)r.base + GUARD + r.bytes, tail, cudaMemcpyDeviceToHost) !=
      cudaSuccess)
    return 0;
  size_t bad = 0;
  for (size_t i = 0; i < h.size(); ++i)
    if (h[i] != 0xFF) {
      if (!bad && first_off) *first_off = i < GUARD ? i : r.bytes + i;  // relative to base + 0
      ++bad;
    }
  return bad;
}

// bytes of a watched region that no longer hold 0xFF
size_t damaged_watch(const Rec& r, size_t* first_off) {
  std::vector<unsigned char> h(r.total);
  if (cudaMemcpy(h.data(), r.base, r.total, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
  size_t bad = 0;
  for (size_t i = 0; i < h.size(); ++i)
    if (h[i] != 0xFF) {
      if (!bad && first_off) *first_off = GUARD + r.bytes + i;  // report() subtracts GUARD
      ++bad;
    }
  return bad;
}

void report(const Rec& r, size_t bad, size_t first) {
  char msg[512];
  const long long rel = (long long)first - (long long)GUARD;  // offset from the payload start
  snprintf(msg, sizeof msg,
           "CGB200_GUARD: %zu guard byte(s) overwritten around the %zu-byte buffer allocated at %s:%d "
           "(first at payload offset %lld)",
           bad, r.bytes, r.file, r.line, rel);
  fprintf(stderr, "%s\n", msg);
  cgb::set_error(msg);
}

}  // namespace

bool dev_guard_enabled() {
  std::lock_guard<std::mutex> lk(g_mu);
  return guard_on();
}

cudaError_t dev_alloc(void** p, size_t bytes, const char* file, int line) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!guard_on()) return cudaMalloc(p, bytes);
  const size_t total = GUARD + (bytes + GUARD - 1) / GUARD * GUARD + GUARD;
  void* base = nullptr;
  cudaError_t e = cudaMalloc(&base, total);
  if (e != cudaSuccess) return e;
  // poison everything; the library's streams may be non-blocking, so wait for the fill
  e = cudaMemset(base, 0xFF, total);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(base);
    return e;
  }
  *p = (char*)base + GUARD;
  g_live[*p] = Rec{base, bytes, total, file, line};
  return cudaSuccess;
}

cudaError_t dev_free(void* p) {
  if (!p) return cudaSuccess;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!guard_on()) return cudaFree(p);
  auto it = g_live.find(p);
  if (it == g_live.end()) return cudaFree(p);  // not ours (exported buffers stay on cudaMalloc)
  cudaDeviceSynchronize();
  size_t first = 0;
  const size_t bad = damaged_bytes(it->second, &first);
  if (bad) {
    ++g_damaged;
    report(it->second, bad, first);
  }
  void* base = it->second.base;
  g_live.erase(it);
  return cudaFree(base);
}

long long dev_guard_check() {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!guard_on()) return -1;
  cudaDeviceSynchronize();
  long long n = g_damaged;
  for (auto& kv : g_live) {
    size_t first = 0;
    const size_t bad = damaged_bytes(kv.second, &first);
    if (bad) {
      ++n;
      report(kv.second, bad, first);
    }
  }
  for (auto& kv : g_watch) {
    size_t first = 0;
    const size_t bad = damaged_watch(kv.second, &first);
    if (bad) {
      ++n;
      report(kv.second, bad, first);
    }
  }
  return n;
}

cudaError_t dev_guard_watch(void* region, size_t len, size_t payload_bytes, const char* file, int line) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!guard_on() || !region || len == 0) return cudaSuccess;
  if (len > (64u << 10)) len = 64u << 10;
  cudaError_t e = cudaMemset(region, 0xFF, len);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess) g_watch[region] = Rec{region, payload_bytes, len, file, line};
  return e;
}

void dev_guard_unwatch(void* region) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_watch.find(region);
  if (it == g_watch.end()) return;
  cudaDeviceSynchronize();
  size_t first = 0;
  const size_t bad = damaged_watch(it->second, &first);
  if (bad) {
    ++g_damaged;
    report(it->second, bad, first);
  }
  g_watch.erase(it);
}

// Positive control for the tests: one byte stored past the end of a 1000-byte buffer must be seen
// (and one before its start); the bytes are restored so the buffer frees clean.
// 1 = both detected, 0 = missed, -1 = guard mode off or a CUDA failure.
int dev_guard_selftest() {
  if (!dev_guard_enabled()) return -1;
  unsigned char* d = nullptr;
  if (DEV_ALLOC(&d, 1000) != cudaSuccess) return -1;
  int seen = 0;
  for (long long off : {1000LL, -1LL}) {
    if (cudaMemset(d + off, 0x00, 1) != cudaSuccess) return -1;
    cudaDeviceSynchronize();
    {
      std::lock_guard<std::mutex> lk(g_mu);
      size_t first = 0;
      if (damaged_bytes(g_live[d], &first) == 1 && (long long)first - (long long)GUARD == off) ++seen;
    }
    if (cudaMemset(d + off, 0xFF, 1) != cudaSuccess) return -1;
  }
  dev_free(d);
  return seen == 2 ? 1 : 0;
}
