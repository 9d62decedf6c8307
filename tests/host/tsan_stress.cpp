// ThreadSanitizer stress of the host runtime (csrc/host/pool.cpp, parse.cpp): the reference relies on
// `synchronized` pools (HC/resource/ArrayPool.java) and has no race detector; here the C++ counterpart is
// hammered from many threads under -fsanitize=thread.  Built and run by tests/test_host_tsan.py.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

extern "C" {
void* hh_pool_get(size_t bytes, int approximate);
int hh_pool_release(void* blk);
int hh_pool_free(void* blk);
void hh_pool_clean();
void hh_pool_stats(unsigned long long* out);
long long hh_parse_doubles(const char* data, long long len, double* out, long long cap, int threads);
}

int main() {
  std::atomic<long long> errors{0};
  std::vector<std::thread> ts;
  for (int t = 0; t < 8; ++t) {
    ts.emplace_back([t, &errors] {
      std::vector<void*> held;
      for (int i = 0; i < 4000; ++i) {
        void* b = hh_pool_get((size_t)(1 + (i * 37 + t * 101) % 70000), i & 1);
        if (!b) { ++errors; continue; }
        std::memset(b, t, 64);
        held.push_back(b);
        if (held.size() > 16) {
          if (hh_pool_release(held.front()) != 0) ++errors;
          held.erase(held.begin());
        }
        if (i % 997 == 0) { unsigned long long st[8] = {0}; hh_pool_stats(st); }
        if (i % 1999 == 0 && t == 0) hh_pool_clean();
      }
      for (size_t k = 0; k < held.size(); ++k)
        if ((k & 1 ? hh_pool_free(held[k]) : hh_pool_release(held[k])) != 0) ++errors;
    });
  }
  // the parser runs its own worker threads; several callers at once
  for (int t = 0; t < 3; ++t) {
    ts.emplace_back([t, &errors] {
      std::string s;
      for (int i = 0; i < 20000; ++i) s += std::to_string(i * 0.5 + t) + ((i % 10 == 9) ? "\n" : ",");
      std::vector<double> out(20000);
      for (int rep = 0; rep < 5; ++rep) {
        long long n = hh_parse_doubles(s.data(), (long long)s.size(), out.data(), (long long)out.size(), 4);
        if (n != 20000 || out[7] != 3.5 + t) ++errors;
      }
    });
  }
  for (auto& th : ts) th.join();
  std::printf("tsan_stress errors=%lld\n", errors.load());
  return errors.load() == 0 ? 0 : 1;
}
