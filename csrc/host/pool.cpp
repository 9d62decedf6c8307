// Host staging-buffer pool: the C++ counterpart of ResourcePool/ArrayPool
// (HC/resource/ArrayPool.java:73-228): free lists keyed by power-of-two size class,
// release() returns a block to its list, free() gives it back to the OS, clean() drops
// every cached block, stats for log().  Blocks are 4 KiB aligned so they can be
// cudaHostRegister'ed (pinned) by the caller for asynchronous H2D copies.
#include <cstdint>
#include <cstdlib>
#include <map>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace {
struct Pool {
  std::mutex mu;
  std::map<size_t, std::vector<void*>> free_lists;   // size class -> cached blocks
  std::unordered_map<void*, size_t> in_use;          // block -> size class
  size_t cached_bytes = 0, in_use_bytes = 0, total_allocs = 0, pool_hits = 0;
};
Pool& pool() {
  static Pool p;
  return p;
}
size_t round_pow2(size_t n) {
  size_t c = 4096;
  while (c < n) c <<= 1;
  return c;
}
}  // namespace

extern "C" {

void* hh_pool_get(size_t bytes, int approximate) {
  Pool& p = pool();
  const size_t cls = approximate ? round_pow2(bytes) : ((bytes + 4095) / 4096 * 4096);
  std::lock_guard<std::mutex> lock(p.mu);
  ++p.total_allocs;
  auto it = p.free_lists.find(cls);
  void* blk = nullptr;
  if (it != p.free_lists.end() && !it->second.empty()) {
    blk = it->second.back();
    it->second.pop_back();
    p.cached_bytes -= cls;
    ++p.pool_hits;
  } else {
    if (posix_memalign(&blk, 4096, cls ? cls : 4096) != 0) return nullptr;
  }
  p.in_use[blk] = cls;
  p.in_use_bytes += cls;
  return blk;
}

// Returns the block to the pool (ArrayPool.releaseArray). 0 = ok, -1 = unknown block.
int hh_pool_release(void* blk) {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  auto it = p.in_use.find(blk);
  if (it == p.in_use.end()) return -1;
  const size_t cls = it->second;
  p.in_use.erase(it);
  p.in_use_bytes -= cls;
  p.free_lists[cls].push_back(blk);
  p.cached_bytes += cls;
  return 0;
}

// Drops the block entirely (ArrayPool.freeArray).
int hh_pool_free(void* blk) {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  auto it = p.in_use.find(blk);
  if (it == p.in_use.end()) return -1;
  p.in_use_bytes -= it->second;
  p.in_use.erase(it);
  free(blk);
  return 0;
}

// Frees every cached block (ResourcePool.clean).
void hh_pool_clean() {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  for (auto& kv : p.free_lists)
    for (void* b : kv.second) free(b);
  p.free_lists.clear();
  p.cached_bytes = 0;
}

// out[0..4] = cached bytes, in-use bytes, total get() calls, pool hits, size classes (ResourcePool.log).
void hh_pool_stats(unsigned long long* out) {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  out[0] = p.cached_bytes;
  out[1] = p.in_use_bytes;
  out[2] = p.total_allocs;
  out[3] = p.pool_hits;
  out[4] = p.free_lists.size();
}

}  // extern "C"
