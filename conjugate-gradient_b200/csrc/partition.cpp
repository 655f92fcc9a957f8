// partition.cpp -- see partition.h
#include "partition.h"

#include <algorithm>

namespace cgb {

static inline long long round_up16(long long v) { return (v + 15) / 16 * 16; }

static void halos(const PartInfo& p, long long n_global, int dense, long long* lo, long long* hi) {
  const long long row1 = p.row0 + p.nloc;
  *lo = std::max(0LL, p.row0 - p.mincol);
  *hi = std::max(0LL, p.maxcol - (row1 - 1));
  if (dense) {
    *lo = p.row0;
    *hi = n_global - row1;
  }
  if (p.nloc == 0) *lo = *hi = 0;
}

int plan_partition(int world, int rank, long long n_global, const PartInfo* parts, int dense,
                   Plan* out, const char** err) {
  auto fail = [&](const char* m) {
    if (err) *err = m;
    return -1;
  };
  if (world < 1 || rank < 0 || rank >= world || !parts || !out) return fail("bad arguments");
  long long next = 0;
  for (int g = 0; g < world; ++g) {
    if (parts[g].row0 != next || parts[g].nloc < 0) return fail("row blocks must tile [0, n) in rank order");
    next += parts[g].nloc;
  }
  if (next != n_global) return fail("row blocks do not cover the matrix");

  bool wide = dense != 0;
  for (int g = 0; g < world && !wide; ++g) {
    long long lo, hi;
    halos(parts[g], n_global, dense, &lo, &hi);
    if (lo > 0 && (g == 0 || lo > parts[g - 1].nloc)) wide = true;
    if (hi > 0 && (g == world - 1 || hi > parts[g + 1].nloc)) wide = true;
  }
  const PartInfo& me = parts[rank];
  Plan p = {};
  if (world == 1 || wide) {
    p.allgather = 1;
    p.off = me.row0;
    p.ext_len = n_global;
    p.col_shift = 0;
  } else {
    long long lo, hi;
    halos(me, n_global, dense, &lo, &hi);
    p.allgather = 0;
    p.h_lo = (int)lo;
    p.h_hi = (int)hi;
    p.off = round_up16(lo);  // keeps the owned block 128-byte aligned for the 128-bit kernels
    p.ext_len = p.off + me.nloc + hi;
    p.col_shift = me.row0 - p.off;
    p.recv_lo = p.h_lo;
    p.recv_hi = p.h_hi;
    if (rank > 0) {  // my first rows are the lower rank's upper halo
      long long l2, h2;
      halos(parts[rank - 1], n_global, dense, &l2, &h2);
      p.send_lo = (int)h2;
    }
    if (rank + 1 < world) {  // my last rows are the upper rank's lower halo
      long long l2, h2;
      halos(parts[rank + 1], n_global, dense, &l2, &h2);
      p.send_hi = (int)l2;
    }
  }
  *out = p;
  return 0;
}

int plan_push(int world, int rank, long long n_global, const PartInfo* parts, int dense,
              PushPlan* out, const char** err) {
  if (world > 8) {
    if (err) *err = "peer-memory mode supports at most 8 ranks";
    return -1;
  }
  Plan plans[8] = {};
  for (int g = 0; g < world; ++g)
    if (plan_partition(world, g, n_global, parts, dense, &plans[g], err)) return -1;
  PushPlan pp = {};
  const Plan& me = plans[rank];
  const long long nloc = parts[rank].nloc;
  if (world == 1) {
    *out = pp;
    return 0;
  }
  if (me.allgather) {
    // every peer holds the whole vector: my block sits at my global row offset in each of them
    for (int q = 0; q < world; ++q) {
      if (q == rank) continue;
      pp.push[pp.npush++] = PushItem{q, 0, nloc, parts[rank].row0};
      pp.wait_src[pp.nwait++] = q;
    }
  } else {
    if (rank > 0 && me.send_lo > 0) {  // my first rows are the lower rank's upper halo
      const Plan& q = plans[rank - 1];
      pp.push[pp.npush++] = PushItem{rank - 1, 0, me.send_lo, q.off + parts[rank - 1].nloc};
    }
    if (rank + 1 < world && me.send_hi > 0) {  // my last rows are the upper rank's lower halo
      const Plan& q = plans[rank + 1];
      pp.push[pp.npush++] = PushItem{rank + 1, nloc - me.send_hi, nloc, q.off - q.h_lo};
    }
    if (me.h_lo > 0) pp.wait_src[pp.nwait++] = rank - 1;
    if (me.h_hi > 0) pp.wait_src[pp.nwait++] = rank + 1;
  }
  *out = pp;
  return 0;
}

}  // namespace cgb
