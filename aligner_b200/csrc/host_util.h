// host_util.h — host-side plumbing shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/aligner_b200.h"

namespace alnr {

extern std::atomic<uint64_t> g_launches;
void set_error_detail(const std::string &s);

// Maps a CUDA failure to ALNR_CUDA_ERROR / ALNR_OUT_OF_MEMORY and records the detail text.
inline int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
  char buf[512];
  std::snprintf(buf, sizeof(buf), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  set_error_detail(buf);
  (void)cudaGetLastError();
  return e == cudaErrorMemoryAllocation ? ALNR_OUT_OF_MEMORY : ALNR_CUDA_ERROR;
}

#define ALNR_CUDA(call)                                                        \
  do {                                                                         \
    cudaError_t _e = (call);                                                   \
    if (_e != cudaSuccess) return ::alnr::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define ALNR_LAUNCHED()                                                        \
  do {                                                                         \
    ::alnr::g_launches.fetch_add(1, std::memory_order_relaxed);                \
    cudaError_t _e = cudaGetLastError();                                       \
    if (_e != cudaSuccess) return ::alnr::cuda_fail(_e, "kernel launch", __FILE__, __LINE__); \
  } while (0)

// Caching device allocator: cudaMalloc/cudaFree cost milliseconds per GB and synchronise the
// device, so buffers released by one call are kept per device and reused by the next
// (alnr_release_cached_memory() returns them to the driver).  Thread-safe.
cudaError_t pool_alloc(void **p, size_t bytes, size_t *granted);
void pool_free(void *p, size_t granted);
void pool_release_all();
size_t pool_cached_bytes(int device);  // bytes of idle blocks the pool holds for `device`

// RAII device buffer backed by the pool.
template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  size_t granted = 0;
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p) pool_free(p, granted);
    p = nullptr;
    n = 0;
    granted = 0;
  }
  cudaError_t alloc(size_t count) {
    release();
    if (count == 0) count = 1;
    void *q = nullptr;
    cudaError_t e = pool_alloc(&q, count * sizeof(T), &granted);
    if (e == cudaSuccess) { p = (T *)q; n = count; } else { p = nullptr; granted = 0; }
    return e;
  }
};

}  // namespace alnr
