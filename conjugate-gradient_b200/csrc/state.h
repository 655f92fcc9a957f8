// state.h -- the device-resident scalars of the CG loop (host- and device-visible layout).
// One instance lives in device memory per solve; kernels read alpha/beta from it and the
// finishing thread of each reduction updates it, so an iteration never returns to the host
// (reference cg.c:113,129 compute these scalars on the host between operator calls).
#pragma once

#define CGB200_ST_RUNNING 0
#define CGB200_ST_MAXITER 1    /* fixed count reached (the reference's only exit, cg.c:125-127) */
#define CGB200_ST_CONVERGED 2  /* r.r <= tol^2 * b.b (extension) */
#define CGB200_ST_BREAKDOWN 3  /* p.Ap zero or not finite (negative is let through, cg.c:113) */
#define CGB200_ST_EXACT 4      /* r.r == 0 exactly: stopped where the reference would divide 0/0 */
#define CGB200_ST_COMM 5       /* a peer-memory exchange between GPUs timed out */

struct CgState {
  double rr;      // r.r of the current residual
  double ps;      // p.(A p)   (single-reduction variant: r.(A r))
  double alpha;   // rr / ps                      (cg.c:113)
  double beta;    // rr_new / rr                  (cg.c:129)
  double bb;      // b.b
  double tol;     // relative tolerance on ||r||/||b||; 0 = fixed count (reference behaviour)
  double thresh;  // tol^2 * bb
  double local[2];  // NCCL mode: this rank's partial sums awaiting the cross-GPU sum
  int iter;       // x updates performed so far
  int max_iter;   // reference semantics: stop after the update at k == max_iter
  int done;
  int status;
  int variant;    // 0 classic (the reference's recurrence), 1 single-reduction
  int sr_first;   // single-reduction: the next fused reduction is the loop-entry one
  unsigned int ticket[4];  // grid-reduction tickets (one per fused kernel)
  unsigned long long grid_epoch;  // grid barriers passed by the one-kernel iteration (dense path)
  double aux[3];
};
