// writewatch.cpp -- "has the caller written to this array since we looked?" without reading it.
//
// The reference reads the caller's matrix arrays on every call (mv_ops.c:187-197), so a drop-in has
// to notice an in-place edit of values[] between two calls.  Hashing every byte per call is exact
// but costs a pass over host memory (0.1 s for the 12.9 GB of the dense n = 32768 system, against a
// 6 ms solve).  Linux >= 6.7 can answer the question directly: userfaultfd write-protection in
// ASYNC mode marks a page "written" in its page-table entry on the first CPU write (resolved inside
// the kernel, no handler, no signal) and ioctl(PAGEMAP_SCAN) reports such pages -- 0.6 ms per GiB
// for a clean range.  Only the ranges registered here are affected; the caller pays one minor
// fault per 4 KiB page it rewrites.
//
// Exactness: every write that goes through the CPU's page tables is seen, the kernel's own
// copy_to_user (read(2) into the array) included.  Writes that bypass them are NOT: device DMA into
// page-locked memory.  abi.cu therefore never trusts this for arrays that CUDA reports as pinned or
// registered, and falls back to the content hash whenever anything here answers "cannot tell".
// Only the page-aligned INTERIOR of an array is watched (its first and last partial pages may be
// shared with other allocations; abi.cu hashes those few bytes instead).
#include "writewatch.h"

#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/ioctl.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <thread>
#include <vector>

#if defined(__linux__) && __has_include(<linux/userfaultfd.h>) && __has_include(<linux/fs.h>)
#include <linux/fs.h>
#include <linux/userfaultfd.h>
#if defined(PAGEMAP_SCAN) && defined(UFFD_FEATURE_WP_ASYNC)
#define CGB_HAVE_WW 1
#endif
#endif

namespace cgb {

#ifdef CGB_HAVE_WW

namespace {

int g_uffd = -1, g_pm = -1;
int g_state = 0;  // 0 = not tried, 1 = usable, -1 = unavailable
long g_page = 4096;

// any written page in [a, b)?  1 / 0; -1 = the range is not (or no longer) watched
int scan_range(uintptr_t a, uintptr_t b) {
  struct page_region reg[1];
  struct pm_scan_arg arg;
  memset(&arg, 0, sizeof arg);
  arg.size = sizeof arg;
  arg.start = a;
  arg.end = b;
  arg.vec = (uintptr_t)reg;
  arg.vec_len = 1;
  arg.max_pages = 1;  // one hit is enough
  arg.category_mask = PAGE_IS_WRITTEN;
  arg.return_mask = PAGE_IS_WRITTEN;
  arg.flags = PM_SCAN_CHECK_WPASYNC;  // fail instead of guessing when the range is not async-wp registered
  const long r = ioctl(g_pm, PAGEMAP_SCAN, &arg);
  if (r < 0) return -1;
  return r > 0 ? 1 : 0;
}

bool interior(const void* p, size_t bytes, uintptr_t* a, uintptr_t* b) {
  const uintptr_t lo = ((uintptr_t)p + g_page - 1) / g_page * g_page;
  const uintptr_t hi = ((uintptr_t)p + bytes) / g_page * g_page;
  if (hi <= lo) return false;
  *a = lo;
  *b = hi;
  return true;
}

bool self_test() {  // the mechanism does what this file relies on, on this kernel
  const size_t len = 4 * (size_t)g_page;
  char* m = (char*)mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  if (m == MAP_FAILED) return false;
  memset(m, 1, len);
  bool ok = ww_arm(m, len);
  ok = ok && ww_written(m, len) == 0;
  if (ok) {
    volatile char c = m[g_page + 5];  // a read is not a write
    (void)c;
    ok = ww_written(m, len) == 0;
    m[2 * g_page + 7] = 9;
    ok = ok && ww_written(m, len) == 1;
    ok = ok && ww_rearm(m, len) && ww_written(m, len) == 0;
  }
  ww_disarm(m, len);
  munmap(m, len);
  return ok;
}

}  // namespace

bool ww_available() {
  if (g_state != 0) return g_state > 0;
  g_state = -1;
  const char* e = getenv("CGB200_WRITE_TRACKING");
  if (e && strcmp(e, "0") == 0) return false;
  g_page = sysconf(_SC_PAGESIZE);
  g_uffd = (int)syscall(SYS_userfaultfd, O_CLOEXEC | O_NONBLOCK | UFFD_USER_MODE_ONLY);
  if (g_uffd < 0) return false;
  struct uffdio_api api;
  memset(&api, 0, sizeof api);
  api.api = UFFD_API;
  api.features = UFFD_FEATURE_WP_ASYNC | UFFD_FEATURE_WP_UNPOPULATED;
  if (ioctl(g_uffd, UFFDIO_API, &api) != 0 || !(api.features & UFFD_FEATURE_WP_ASYNC)) {
    close(g_uffd);
    g_uffd = -1;
    return false;
  }
  g_pm = open("/proc/self/pagemap", O_RDONLY | O_CLOEXEC);
  if (g_pm < 0) {
    close(g_uffd);
    g_uffd = -1;
    return false;
  }
  g_state = 1;  // (the self-test calls the functions below)
  if (!self_test()) {
    g_state = -1;
    close(g_uffd);
    close(g_pm);
    g_uffd = g_pm = -1;
  }
  return g_state > 0;
}

bool ww_arm(const void* p, size_t bytes) {
  if (g_state <= 0 || !p) return false;
  uintptr_t a, b;
  if (!interior(p, bytes, &a, &b)) return false;
  struct uffdio_register rg;
  memset(&rg, 0, sizeof rg);
  rg.range.start = a;
  rg.range.len = b - a;
  rg.mode = UFFDIO_REGISTER_MODE_WP;
  if (ioctl(g_uffd, UFFDIO_REGISTER, &rg) != 0) return false;  // not anonymous memory, other owner, ...
  return ww_rearm(p, bytes);
}

namespace {

// [a, b) dealt to a few threads in page-aligned pieces (page-table walks of multi-GiB ranges)
template <class F>
void split_range(uintptr_t a, uintptr_t b, F&& f) {
  const size_t len = b - a;
  const size_t CH = 512u << 20;
  size_t nt = (len + CH - 1) / CH;
  if (nt > 8) nt = 8;
  if (nt <= 1) {
    f(0, a, b);
    return;
  }
  std::vector<std::thread> th;
  auto work = [&](size_t t) {
    const uintptr_t c0 = a + (len / nt * t) / g_page * g_page;
    const uintptr_t c1 = t + 1 == nt ? b : a + (len / nt * (t + 1)) / g_page * g_page;
    if (c1 > c0) f(t, c0, c1);
  };
  for (size_t t = 1; t < nt; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
}

}  // namespace

bool ww_rearm(const void* p, size_t bytes) {
  if (g_state <= 0 || !p) return false;
  uintptr_t a, b;
  if (!interior(p, bytes, &a, &b)) return false;
  int bad[8] = {0};
  split_range(a, b, [&](size_t t, uintptr_t c0, uintptr_t c1) {
    struct uffdio_writeprotect wp;
    memset(&wp, 0, sizeof wp);
    wp.range.start = c0;
    wp.range.len = c1 - c0;
    wp.mode = UFFDIO_WRITEPROTECT_MODE_WP;
    if (ioctl(g_uffd, UFFDIO_WRITEPROTECT, &wp) != 0) bad[t] = 1;
  });
  for (int x : bad)
    if (x) return false;
  return true;
}

int ww_written(const void* p, size_t bytes) {
  if (g_state <= 0 || !p) return -1;
  uintptr_t a, b;
  if (!interior(p, bytes, &a, &b)) return -1;
  int res[8] = {0};
  split_range(a, b, [&](size_t t, uintptr_t c0, uintptr_t c1) { res[t] = scan_range(c0, c1); });
  int out = 0;
  for (int r : res) {
    if (r < 0) return -1;
    out |= r;
  }
  return out;
}

void ww_disarm(const void* p, size_t bytes) {
  if (g_state <= 0 || !p) return;
  uintptr_t a, b;
  if (!interior(p, bytes, &a, &b)) return;
  struct uffdio_range r;
  r.start = a;
  r.len = b - a;
  ioctl(g_uffd, UFFDIO_UNREGISTER, &r);
}

size_t ww_page() {
  ww_available();
  return (size_t)g_page;
}

#else  // no kernel headers for it: the callers fall back to hashing

bool ww_available() { return false; }
bool ww_arm(const void*, size_t) { return false; }
bool ww_rearm(const void*, size_t) { return false; }
int ww_written(const void*, size_t) { return -1; }
void ww_disarm(const void*, size_t) {}
size_t ww_page() { return 4096; }

#endif

}  // namespace cgb
