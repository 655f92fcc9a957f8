// team.h -- several GPUs behind the reference's own entry point.
//
// conj_grad(max_iter, A, b, &x) (reference cg.c:24,88) takes host structs and knows nothing about
// GPUs.  With CGB200_GPUS=N (or cgb200_set_gpus) the calling process becomes rank 0 of a team of N
// processes, one per GPU: it spawns N-1 workers (the cgb200_worker executable next to the library),
// joins them into one NCCL world, cuts the caller's CSR matrix into N contiguous row blocks of
// equal stored entries, puts every block on its GPU itself (cudaMemcpy from the caller's arrays,
// no host-side copy) and hands the device buffers to the block's worker through CUDA IPC.  From
// there on the team runs exactly the multi-rank loop of solver.cu (peer-memory exchange between the
// kernels, NCCL fallback); the primary copies b in and x out of every GPU.  The workers stay up
// between calls, and a matrix that has not changed stays on its GPUs (the matrix cache of abi.cu).
#pragma once

#include "solver.h"

struct __mv_sparse;

namespace cgb {

struct TeamSystem;  // one matrix spread over the team

int team_requested();             // GPUs asked for (CGB200_GPUS / cgb200_set_gpus); <= 1: no team
void team_request(int gpus);
bool team_active();               // the workers are up and this process is rank 0 of their world
int team_start();                 // idempotent; 0 / -1 (then the caller falls back to one GPU)
void team_stop();

TeamSystem* team_create(const struct __mv_sparse* A);   // collective; nullptr on failure
void team_destroy(TeamSystem* T);                        // collective
int team_solve(TeamSystem* T, const double* b_host, const double* x0_host, double* x_host,
               const SolveOpts& o, SolveResult* res);    // collective

int worker_main(int argc, char** argv);  // body of the cgb200_worker executable

}  // namespace cgb
