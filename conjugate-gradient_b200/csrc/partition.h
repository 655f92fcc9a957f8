// partition.h -- host-only planning of the row-block decomposition (no CUDA in here, so it is
// unit-testable on a CPU box; tests/test_partition_gloo.py drives it with world_size 2 over gloo).
//
// The reference has no partitioning to match (its Makefile's mpiexec lines launch an MPI-free
// binary); this layer is new.  Rank g owns rows [row0[g], row0[g]+nloc[g]) of A and
// the same slices of x, r, p, s, b.  From the span of column indices each rank touches the plan
// picks how the search direction is exchanged before the matvec:
//   neighbour mode  every rank needs only a halo that lies inside its two adjacent ranks
//                   (stencil / banded matrices): p_ext = [lower halo | owned | upper halo]
//   all-gather mode anything wider (dense): p_ext is the whole vector
#pragma once

namespace cgb {

struct PartInfo {
  long long row0, nloc;      // owned rows
  long long mincol, maxcol;  // smallest / largest GLOBAL column index stored in those rows
};

struct Plan {
  int allgather;        // 1 = gather space is the whole vector (also the single-rank layout)
  int h_lo, h_hi;       // halo rows needed below / above the owned block (0 in all-gather mode)
  long long off;        // index of the first owned row inside p_ext
  long long ext_len;    // length of p_ext
  long long col_shift;  // gather index = global column - col_shift
  int send_lo, send_hi; // rows sent to rank-1 (my first rows) / rank+1 (my last rows)
  int recv_lo, recv_hi; // == h_lo / h_hi in neighbour mode
};

// 0 on success; -1 (and a message in *err when non-null) when the blocks do not tile
// [0, n_global) in rank order.  All ranks compute the same mode from the same `parts`.
int plan_partition(int world, int rank, long long n_global, const PartInfo* parts, int dense,
                   Plan* out, const char** err);

// Peer-memory mode (comm.h): who stores which of its rows where.  Rank `rank` stores rows
// [begin, end) of its owned block at index dst_index.. of rank `peer`'s gather vector p_ext, and
// must see a completion flag from every rank in wait_src before gathering outside its own block.
struct PushItem {
  int peer;
  long long begin, end;   // local row range of the sender
  long long dst_index;    // first index in the receiver's p_ext
};

struct PushPlan {
  int npush;
  PushItem push[8];
  int nwait;
  int wait_src[8];
};

// Derived from plan_partition() of every rank (each rank can evaluate every rank's plan, so the
// layout never has to be communicated).  world <= 8.  0 / -1 like plan_partition.
int plan_push(int world, int rank, long long n_global, const PartInfo* parts, int dense,
              PushPlan* out, const char** err);

}  // namespace cgb
