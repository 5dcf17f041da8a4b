// pool.cu — per-device caching allocator behind DevBuf (see host_util.h).
#include <map>
#include <mutex>
#include <vector>

#include "host_util.h"

namespace alnr {

namespace {
struct Pool {
  std::mutex mu;
  // device -> (size -> free blocks of that size)
  std::map<int, std::multimap<size_t, void *>> free_blocks;
  std::map<void *, int> owner;  // block -> device it was allocated on
};
Pool &pool() {
  static Pool *p = new Pool();  // leaked on purpose: must outlive static destructors
  return *p;
}
size_t round_up(size_t b) {
  const size_t g = b < (1u << 20) ? 4096 : (2u << 20);
  return (b + g - 1) / g * g;
}
}  // namespace

cudaError_t pool_alloc(void **p, size_t bytes, size_t *granted) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const size_t need = round_up(bytes);
  Pool &P = pool();
  {
    std::lock_guard<std::mutex> g(P.mu);
    auto &fb = P.free_blocks[dev];
    // best fit: an exact-size block if there is one (repeated calls of one shape then always hit),
    // else the smallest larger block unless it is wastefully large
    auto it = fb.lower_bound(need);
    if (it != fb.end() && it->first <= need + need / 4 + (8u << 20)) {
      *p = it->second;
      *granted = it->first;
      fb.erase(it);
      return cudaSuccess;
    }
  }
  e = cudaMalloc(p, need);
  if (e == cudaErrorMemoryAllocation) {  // give cached blocks back and retry once
    (void)cudaGetLastError();
    pool_release_all();
    e = cudaMalloc(p, need);
  }
  if (e != cudaSuccess) return e;
  *granted = need;
  std::lock_guard<std::mutex> g(P.mu);
  P.owner[*p] = dev;
  return cudaSuccess;
}

void pool_free(void *p, size_t granted) {
  if (!p) return;
  Pool &P = pool();
  std::lock_guard<std::mutex> g(P.mu);
  auto it = P.owner.find(p);
  const int dev = it == P.owner.end() ? 0 : it->second;
  P.free_blocks[dev].emplace(granted, p);
}

size_t pool_cached_bytes(int device) {
  Pool &P = pool();
  std::lock_guard<std::mutex> g(P.mu);
  size_t total = 0;
  auto it = P.free_blocks.find(device);
  if (it != P.free_blocks.end())
    for (auto &b : it->second) total += b.first;
  return total;
}

void pool_release_all() {
  Pool &P = pool();
  std::vector<std::pair<int, void *>> blocks;
  {
    std::lock_guard<std::mutex> g(P.mu);
    for (auto &d : P.free_blocks) {
      for (auto &b : d.second) { blocks.emplace_back(d.first, b.second); P.owner.erase(b.second); }
      d.second.clear();
    }
  }
  int cur = 0;
  cudaGetDevice(&cur);
  for (auto &b : blocks) { cudaSetDevice(b.first); cudaFree(b.second); }
  cudaSetDevice(cur);
}

}  // namespace alnr

extern "C" void alnr_release_cached_memory(void) { alnr::pool_release_all(); }
