// harp_b200 — colour-coding dynamic programme of subgraph counting (sm_100a).
//
// For a rooted sub-template T split at an edge into an active part Ta (holding the root) and a passive part Tp, the
// number of colourful embeddings of T rooted at vertex v that use exactly the colour set C is
//     count_T[v][C] = sum over S1 subset C, |S1| = |Ta| of  count_Ta[v][S1] * ( sum over u in N(v) of count_Tp[u][C \ S1] )
// (ml/java/src/main/java/edu/iu/subgraph/colorcount_HJ.java:861-925, the hot loop of the reference).  Tables are
// COMPRESSED: a sub-template of size s keeps one column per colour set of size s (C(k, s) of them), not 2^k.
// One warp per vertex: the warp sums the passive rows of the vertex's neighbours into shared memory (coalesced row
// reads), keeps the vertex's active row next to it, and then every lane forms the sums of products for its colour
// sets from a precomputed split table (pairs of column indices, shared by all vertices, L1/L2 resident).
#include "common.cuh"
#include "harp_b200.h"

namespace hb {
namespace subgraph {

constexpr int WARPS = 8;

__global__ void __launch_bounds__(WARPS * 32)
combine_kernel(const long long* __restrict__ nbr_ptr, const int* __restrict__ nbr_idx, const float* __restrict__ act,
               const float* __restrict__ pas, const short* __restrict__ split_a, const short* __restrict__ split_p,
               const int na, const int np, const int nC, const int nsplit, float* __restrict__ out, const int n) {
  extern __shared__ __align__(16) float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nap = (na + 3) & ~3, npp = (np + 3) & ~3;        // keep both rows 16-byte aligned
  float* a_s = sm + (size_t)warp * (nap + npp);
  float* p_s = a_s + nap;
  for (int v = blockIdx.x * WARPS + warp; v < n; v += gridDim.x * WARPS) {
    for (int j = lane; j < na; j += 32) a_s[j] = act[(size_t)v * na + j];
    const long long e0 = nbr_ptr[v], e1 = nbr_ptr[v + 1];
    if ((np & 3) == 0) {
      // 16-byte row reads, four neighbours' rows in flight per lane
      const int nq = np >> 2;
      for (int q0 = 0; q0 < nq; q0 += 32) {
        const int q = q0 + lane;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long e = e0; e < e1; e += 4) {
          int u[4];
#pragma unroll
          for (int t = 0; t < 4; ++t) u[t] = (e + t < e1) ? nbr_idx[e + t] : -1;
          float4 r[4];
#pragma unroll
          for (int t = 0; t < 4; ++t)
            r[t] = (u[t] >= 0 && q < nq) ? *reinterpret_cast<const float4*>(pas + (size_t)u[t] * np + q * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int t = 0; t < 4; ++t) { acc.x += r[t].x; acc.y += r[t].y; acc.z += r[t].z; acc.w += r[t].w; }
        }
        if (q < nq) *reinterpret_cast<float4*>(p_s + q * 4) = acc;
      }
    } else {
      for (int j0 = 0; j0 < np; j0 += 32 * 4) {
        // four independent column chunks per lane keep several loads in flight
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (long long e = e0; e < e1; ++e) {
          const float* row = pas + (size_t)nbr_idx[e] * np;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int j = j0 + q * 32 + lane;
            if (j < np) acc[q] += row[j];
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int j = j0 + q * 32 + lane;
          if (j < np) p_s[j] = acc[q];
        }
      }
    }
    __syncwarp();
    for (int c = lane; c < nC; c += 32) {
      const short* sa = split_a + (size_t)c * nsplit;
      const short* sp = split_p + (size_t)c * nsplit;
      float s = 0.f;
      for (int j = 0; j < nsplit; ++j) s = fmaf(a_s[sa[j]], p_s[sp[j]], s);
      out[(size_t)v * nC + c] = s;
    }
    __syncwarp();
  }
}

}  // namespace subgraph
}  // namespace hb

using namespace hb;

extern "C" {

// out[v, c] = sum_j act[v, split_a[c, j]] * (sum over neighbours u of v of pas[u, split_p[c, j]]).
// nbr_ptr [n + 1] (int64), nbr_idx [nnz] = row of `pas` for every neighbour; act [n, na], pas [*, np], out [n, nC];
// split_a / split_p [nC, nsplit] int16.
int hb_subgraph_combine(const long long* nbr_ptr, const int* nbr_idx, const float* act, const float* pas,
                        const short* split_a, const short* split_p, int na, int np, int nC, int nsplit, float* out,
                        int n, void* stream) {
  HB_REQUIRE(na >= 1 && np >= 1 && nC >= 1 && nsplit >= 1 && na < 32768 && np < 32768, "hb_subgraph_combine: bad table widths");
  if (n <= 0) return 0;
  const size_t smem = (size_t)subgraph::WARPS * (((na + 3) & ~3) + ((np + 3) & ~3)) * sizeof(float);
  HB_REQUIRE(smem <= 200 * 1024, "hb_subgraph_combine: rows of %d + %d columns do not fit shared memory", na, np);
  HB_CUDA_OK(cudaFuncSetAttribute(subgraph::combine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long blocks = ((long long)n + subgraph::WARPS - 1) / subgraph::WARPS;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  subgraph::combine_kernel<<<(unsigned)blocks, subgraph::WARPS * 32, smem, (cudaStream_t)stream>>>(
      nbr_ptr, nbr_idx, act, pas, split_a, split_p, na, np, nC, nsplit, out, n);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
