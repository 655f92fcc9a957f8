// noalloc.cu -- cgb200_dot_static: the dot product for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers may
// not call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid
// it.  The library's own dot_product (abi.cu -> k_dot) reduces through a scratch buffer of per-CTA
// partial sums and a ticket that the context allocates at start-up.  Here the same single-pass
// reduction keeps that scratch in a file-scope static __device__ array: nothing is allocated, by
// this entry point or by anything it calls (it does not even create the library's context).
//
// The array is a fixed 8.5 KB that every GPU which loads the library holds until exit.  Its size
// sets the limit: one partial sum per chunk of CHUNK elements, at most MAX_CHUNKS chunks, so
// n <= CGB200_DOT_STATIC_MAX_N = 2^28 elements (2 GiB per vector); larger n is refused.  Because
// the partials belong to chunks, not to CTAs, and are added in chunk order, the result has the
// same bits whatever the grid size -- and on every GPU model.
//
// Every call shares the one array per device, and it keeps the last call's contents: the ticket is
// cleared by a memset node in front of every launch, the partials are all rewritten before they
// are read.  Two calls that overlap on the same GPU (different streams, or threads) RACE on it:
// callers serialise their calls per device.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/cgb200.h"
#include "solver.h"  // set_error only: no context, no streams, no buffers of the library are touched

namespace {

constexpr int NA_THREADS = 512;
constexpr int NA_CHUNK = 1 << 18;      // elements per partial sum
constexpr int NA_MAX_CHUNKS = 1024;    // 2^28 elements
constexpr int NA_MAX_DEVICES = 64;

struct __align__(256) DotScratch {
  double partial[NA_MAX_CHUNKS];
  unsigned int ticket;                  // CTAs that have finished; zero when a launch starts
  unsigned int pad[63];
};

static __device__ __align__(256) DotScratch g_dot_scratch;

// The array's DEVICE address, once per device.  In host code `g_dot_scratch` is only a stand-in for
// the symbol, not an address a kernel can use: it has to be translated with the device current.
DotScratch* g_scratch_addr[NA_MAX_DEVICES] = {nullptr};

__global__ void __launch_bounds__(NA_THREADS)
k_dot_static(const double* __restrict__ a, const double* __restrict__ b, const long long n,
             const int nchunks, DotScratch* __restrict__ sc, double* __restrict__ out) {
  __shared__ double sm[NA_THREADS / 32];
  __shared__ int is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const long long e0 = (long long)c * NA_CHUNK;
    const long long e1 = e0 + NA_CHUNK < n ? e0 + NA_CHUNK : n;
    double acc = 0.0;
    for (long long i = e0 + threadIdx.x; i < e1; i += NA_THREADS) acc = fma(a[i], b[i], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __syncthreads();  // sm free again
    if (lane == 0) sm[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < NA_THREADS / 32; ++w) t += sm[w];
      sc->partial[c] = t;
    }
  }
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int tk = atomicAdd(&sc->ticket, 1u);
    is_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // the CTA that arrives last adds the chunks' partial sums in chunk order
  double acc = 0.0;
  for (int c = threadIdx.x; c < nchunks; c += NA_THREADS) acc += __ldcg(&sc->partial[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __syncthreads();
  if (lane == 0) sm[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < NA_THREADS / 32; ++w) t += sm[w];
    *out = t;
    sc->ticket = 0u;
  }
}

}  // namespace

extern "C" int cgb200_dot_static(const double* a_dev, const double* b_dev, long long n,
                                 double* result_dev, void* stream) {
  using cgb::set_error;
  if (!result_dev || n < 0 || (n > 0 && (!a_dev || !b_dev))) {  // (empty vectors may be NULL)
    set_error("cgb200_dot_static: NULL pointer or negative length");
    return -1;
  }
  if (n > CGB200_DOT_STATIC_MAX_N) {
    set_error("cgb200_dot_static: n = " + std::to_string(n) + " exceeds CGB200_DOT_STATIC_MAX_N = " +
              std::to_string((long long)CGB200_DOT_STATIC_MAX_N) +
              " (the static scratch array holds one partial sum per 2^18 elements, 1024 of them); "
              "use dot_product, or split the vectors");
    return -1;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) {
    set_error(std::string("cgb200_dot_static: no current CUDA device: ") + cudaGetErrorString(e));
    cudaGetLastError();
    return -1;
  }
  if (!g_scratch_addr[dev]) {  // set-up, once per device, with that device current
    DotScratch* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_dot_scratch);
    if (e != cudaSuccess || !p) {
      set_error(std::string("cgb200_dot_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
      cudaGetLastError();
      return -1;
    }
    g_scratch_addr[dev] = p;
  }
  DotScratch* sc = g_scratch_addr[dev];
  cudaStream_t st = (cudaStream_t)stream;
  // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
  e = cudaMemsetAsync(&sc->ticket, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) {
    set_error(std::string("cgb200_dot_static: clearing the ticket: ") + cudaGetErrorString(e));
    return -1;
  }
  const int nchunks = n > 0 ? (int)((n + NA_CHUNK - 1) / NA_CHUNK) : 1;  // n == 0: one empty chunk -> 0.0
  int grid = nchunks < 592 ? nchunks : 592;
  k_dot_static<<<grid, NA_THREADS, 0, st>>>(a_dev, b_dev, n, nchunks, sc, result_dev);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error(std::string("cgb200_dot_static: launch: ") + cudaGetErrorString(e));
    return -1;
  }
  cgb::add_launches(1);
  return 0;
}
