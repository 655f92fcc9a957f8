// devmem.h -- device allocations of the library, with an optional guard mode.
//
// CGB200_GUARD=1 (read once, at the first allocation) turns every allocation made through
// DEV_ALLOC into [4 KiB guard | payload | guard up to the next 4 KiB], all of it filled with 0xFF:
//   * a kernel that WRITES outside its buffer damages a guard band -- found when the buffer is
//     freed and by cgb200_guard_check(), which names the allocation site;
//   * a kernel that READS outside its buffer, or reads payload nobody wrote, gets NaN doubles /
//     -1 indices, which the parity tests against the oracle then cannot miss.
// It stands in for compute-sanitizer's memcheck/initcheck, which is not available on the GPU pool
// this library is developed on (tests/test_gpu_guard.py re-runs the parity suites in this mode).
// With the variable unset DEV_ALLOC is cudaMalloc and dev_free is cudaFree.
//
// Buffers exported to other processes (cudaIpcGetMemHandle needs the allocation's base address)
// stay on plain cudaMalloc.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

cudaError_t dev_alloc(void** p, size_t bytes, const char* file, int line);
cudaError_t dev_free(void* p);
// -1 when guard mode is off; else the number of allocations whose guard bands were found damaged
// so far (at free time) plus those damaged among the live ones.  Synchronises the device.
long long dev_guard_check();
bool dev_guard_enabled();
int dev_guard_selftest();
// Guard mode only (no-ops otherwise): the slack behind a payload inside a plain cudaMalloc
// allocation (the exported peer-memory buffer) is filled with 0xFF and checked like a guard band.
// `payload_bytes` only labels the report.  At most 64 KiB are watched.
cudaError_t dev_guard_watch(void* region, size_t len, size_t payload_bytes, const char* file, int line);
void dev_guard_unwatch(void* region);

#define DEV_ALLOC(pp, bytes) dev_alloc((void**)(pp), (bytes), __FILE__, __LINE__)
