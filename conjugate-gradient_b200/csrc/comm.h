// comm.h -- peer-memory exchange between the GPUs of one node (host- and device-visible layout).
//
// The reference has no communication at all (its mpiexec targets run an MPI-free binary, SURVEY
// D6).  With one process per GPU, each rank exports ONE device allocation per system through CUDA
// IPC: [ Mailbox | CommLocal | gather vector p_ext ].  Peers map it and the kernels of the loop
// talk to each other directly over NVLink:
//   * the kernel that produces the new gather vector (p = r + beta p, or the residual in the
//     single-reduction variant) also stores the rows a neighbour needs straight into that
//     neighbour's halo slots, and the last CTA raises a flag in the neighbour's mailbox;
//   * the matvec kernel waits for the flags only when it reaches its first boundary tile
//     (interior tiles run first, so the transfer is hidden behind them);
//   * the last CTA of every fused reduction writes its partial sums into every peer's mailbox and
//     sums the G partials it received in rank order (bit-identical on all ranks), so alpha/beta
//     are formed inside the kernel that produced the partials: no NCCL call, no extra launch.
// NCCL stays for set-up (handle exchange, barriers) and as a selectable fallback (CGB200_COMM=nccl).
#pragma once

namespace cgb {

constexpr int MAX_RANKS = 8;

// One rank's contribution to one cross-GPU sum, in flagged words (the scheme NCCL calls LL): each
// 64-bit word carries 32 bits of payload and the 32-bit number of the reduction it belongs to, and
// an aligned 64-bit store is indivisible, so a word whose flag matches is a word whose payload has
// arrived.  No fence on the sending side, no acquire on the receiving side: one NVLink crossing.
struct RedSlot {
  unsigned long long w[4];  // {v0.lo, e}, {v0.hi, e}, {v1.lo, e}, {v1.hi, e}
};

// Lives in the owner's device memory; written by peers (and by the owner itself for its own slot).
struct Mailbox {
  RedSlot red[4][MAX_RANKS];                // [reduction number & 3][source rank]
  unsigned long long halo_flag[MAX_RANKS];  // [source rank] = gather-vector pushes completed by it
};

// Device-local counters, never touched by peers.
struct CommLocal {
  unsigned long long red_epoch;   // cross-GPU reductions completed
  unsigned long long push_epoch;  // gather-vector pushes completed by this rank
  unsigned int push_ticket;
  unsigned int pad;
};

constexpr size_t COMM_LOCAL_OFFSET = 2048;  // CommLocal inside the exported allocation
constexpr size_t COMM_PEXT_OFFSET = 4096;   // p_ext inside the exported allocation

// Rows [begin, end) of the owned block go to dst[0 .. end-begin) in rank `peer`'s memory.
struct PushTarget {
  double* dst;
  long long begin, end;
  int peer;
  int pad;
};

// By-value kernel parameter.  world <= 1 disables everything (single GPU, or NCCL mode).
struct PeerComm {
  int world, rank;
  int npush;  // targets of the gather-vector push
  int nwait;  // sources whose flag the matvec needs before touching its halo
  Mailbox* mbox[MAX_RANKS];  // mbox[rank] is local, the others are peer mappings
  CommLocal* local;
  PushTarget push[MAX_RANKS];
  int wait_src[MAX_RANKS];
};

}  // namespace cgb
