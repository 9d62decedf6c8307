// harp_b200 — WDA-MDS (SMACOF with deterministic annealing): B(X) X and the stress in ONE pass over the distance rows.
//
// The reference computes B(X) X in BCCalcTask and the stress in StressCalcTask, each a sweep over the worker's rows of
// the target-distance matrix (ml/java/src/main/java/edu/iu/wdamds/BCCalcTask.java, StressCalcTask.java:44-101,
// WDAMDSMapper.java:585-725).  On the device both are bandwidth bound on that matrix (4 bytes per pair); formed with
// tensor ops they cost five or six n_local x n temporaries each.  Here one CTA owns a row i: its threads walk the row
// with coalesced loads, recompute d_ij(X) from the (L1/L2 resident) layout, and accumulate
//     b_i   = sum_j r_ij,          r_ij = w_ij * max(delta_ij - shift, 0) / d_ij   (0 where d_ij = 0)
//     bx_i  = sum_j r_ij X_j       ->  (B X)_i = b_i X_i - bx_i
//     s_num = sum_j w_ij (d_ij - dt_ij)^2,   s_den = sum_j w_ij dt_ij^2            (stress = s_num / s_den)
// with w_ij = 1 except w_i,id(i) = 0.  dim <= 8.
#include "common.cuh"
#include "harp_b200.h"

namespace hb {
namespace mds {

constexpr int THREADS = 256;
constexpr int MAX_DIM = 8;

template <int DIMT>          // compile-time bound of the layout dimension (2, 3, 4 or 8): the per-pair loops unroll to it
__global__ void __launch_bounds__(THREADS)
bc_stress_kernel(const float* __restrict__ D, const long long* __restrict__ ids, const float* __restrict__ X,
                 const int m, const int n, const int dim, const float shift, float* __restrict__ BX,
                 double* __restrict__ stress) {
  __shared__ float red[THREADS / 32][MAX_DIM + 3];
  constexpr int MD = DIMT;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = blockIdx.x; i < m; i += gridDim.x) {
    const long long gi = ids[i];
    float xi[MD];
#pragma unroll
    for (int c = 0; c < MD; ++c) xi[c] = c < dim ? X[gi * dim + c] : 0.f;
    float b = 0.f, snum = 0.f, sden = 0.f, bx[MD];
#pragma unroll
    for (int c = 0; c < MD; ++c) bx[c] = 0.f;
    const float* drow = D + (size_t)i * n;
    // four columns per thread and trip: four independent (distance, layout row) loads in flight hide the stream's latency
    for (int j0 = threadIdx.x; j0 < n; j0 += THREADS * 4) {
      float dv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int j = j0 + q * THREADS;
        dv[q] = j < n ? drow[j] : 0.f;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int j = j0 + q * THREADS;
        if (j < n) {
          float xj[MD], d2 = 0.f;
#pragma unroll
          for (int c = 0; c < MD; ++c) {
            xj[c] = c < dim ? X[(size_t)j * dim + c] : 0.f;
            const float t = xi[c] - xj[c];
            d2 = fmaf(t, t, d2);
          }
          const float dist = sqrtf(d2);
          const float w = (j == gi) ? 0.f : 1.f;
          const float dt = fmaxf(dv[q] - shift, 0.f);
          const float r = dist > 1e-12f ? w * dt / dist : 0.f;
          b += r;
#pragma unroll
          for (int c = 0; c < MD; ++c) bx[c] = fmaf(r, xj[c], bx[c]);
          const float e = dist - dt;
          snum = fmaf(w * e, e, snum);
          sden = fmaf(w * dt, dt, sden);
        }
      }
    }
    float vals[MAX_DIM + 3];
#pragma unroll
    for (int c = 0; c < MAX_DIM; ++c) vals[c] = c < MD ? bx[c < MD ? c : 0] : 0.f;
    vals[MAX_DIM] = b; vals[MAX_DIM + 1] = snum; vals[MAX_DIM + 2] = sden;
#pragma unroll
    for (int q = 0; q < MAX_DIM + 3; ++q) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) vals[q] += __shfl_xor_sync(0xffffffffu, vals[q], o);
    }
    __syncthreads();
    if (lane == 0) {
#pragma unroll
      for (int q = 0; q < MAX_DIM + 3; ++q) red[warp][q] = vals[q];
    }
    __syncthreads();
    if (threadIdx.x < MAX_DIM + 3) {
      float s = 0.f;
#pragma unroll
      for (int w2 = 0; w2 < THREADS / 32; ++w2) s += red[w2][threadIdx.x];
      red[0][threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < dim) BX[(size_t)i * dim + threadIdx.x] = red[0][MAX_DIM] * X[gi * dim + threadIdx.x] - red[0][threadIdx.x];
    if (threadIdx.x == 0 && stress != nullptr) {
      atomicAdd(stress, (double)red[0][MAX_DIM + 1]);
      atomicAdd(stress + 1, (double)red[0][MAX_DIM + 2]);
    }
  }
}

}  // namespace mds
}  // namespace hb

using namespace hb;

extern "C" {

// D [m, n] target distances of the m local rows (global ids in ids[m]); X [n, dim] current layout (dim <= 8);
// BX [m, dim] <- rows of B(X) X; stress[0..1] += (numerator, denominator) of the normalised stress (may be NULL).
int hb_mds_bc_stress(const float* D, const long long* ids, const float* X, int m, int n, int dim, float shift, float* BX,
                     double* stress, void* stream) {
  HB_REQUIRE(dim >= 1 && dim <= mds::MAX_DIM, "hb_mds_bc_stress: dim = %d outside 1..%d", dim, mds::MAX_DIM);
  if (m <= 0 || n <= 0) return 0;
  int grid = m < sm_count() * 8 ? m : sm_count() * 8;
  if (dim <= 2) mds::bc_stress_kernel<2><<<grid, mds::THREADS, 0, (cudaStream_t)stream>>>(D, ids, X, m, n, dim, shift, BX, stress);
  else if (dim == 3) mds::bc_stress_kernel<3><<<grid, mds::THREADS, 0, (cudaStream_t)stream>>>(D, ids, X, m, n, dim, shift, BX, stress);
  else if (dim == 4) mds::bc_stress_kernel<4><<<grid, mds::THREADS, 0, (cudaStream_t)stream>>>(D, ids, X, m, n, dim, shift, BX, stress);
  else mds::bc_stress_kernel<8><<<grid, mds::THREADS, 0, (cudaStream_t)stream>>>(D, ids, X, m, n, dim, shift, BX, stress);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
