// harp_b200 — sparse graph kernels of the contrib applications (sm_100a).
//
// pagerank_push_kernel: one PageRank propagation over the pages a worker owns — page s hands pr[s] / outdeg(s) to each
// of its targets, a page without out-links adds its rank to the dangling mass that is spread over everybody
// (contrib/src/main/java/edu/iu/pagerank/PageRankMapper.java: the per-iteration loop that fills the global KV table).
// One warp per source page, lanes over its targets, fp64 `red.global.add` into the dense rank vector.
#include "common.cuh"
#include "harp_b200.h"

namespace hb {
namespace graph {

__global__ void __launch_bounds__(256)
pagerank_push_kernel(const long long* __restrict__ ptr, const int* __restrict__ idx, const long long* __restrict__ src_ids,
                     const double* __restrict__ pr, double* __restrict__ out, double* __restrict__ dangling, const int m) {
  const int lane = threadIdx.x & 31;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  double dsum = 0.0;
  for (int i = wid; i < m; i += nw) {
    const long long k0 = ptr[i], k1 = ptr[i + 1];
    const double p = pr[src_ids[i]];
    if (k0 == k1) {
      if (lane == 0) dsum += p;
      continue;
    }
    const double share = p / (double)(k1 - k0);
    for (long long k = k0 + lane; k < k1; k += 32) red_add_f64(out + idx[k], share);
  }
  if (lane == 0 && dsum != 0.0) red_add_f64(dangling, dsum);
}

}  // namespace graph
}  // namespace hb

using namespace hb;

extern "C" {

// out[t] += pr[s] / outdeg(s) for every link s -> t of the m local source pages (CSR ptr / idx, global ids in src_ids);
// *dangling += sum of pr[s] over local pages without out-links.  out and dangling are accumulated into.
int hb_pagerank_push(const long long* ptr, const int* idx, const long long* src_ids, const double* pr, double* out,
                     double* dangling, int m, void* stream) {
  if (m <= 0) return 0;
  long long blocks = ((long long)m * 32 + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  graph::pagerank_push_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(ptr, idx, src_ids, pr, out, dangling, m);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
