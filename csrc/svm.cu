// SVM training on the device: the whole SMO loop of one binary C-SVC problem in ONE kernel launch.
//
// Replaces DAAL's svm::training::Batch (Boser method) behind
// ml/daal/.../daal_svm/MultiClassDenseBatch/SVMDaalCollectiveMapper.java:150-199 and the contrib iterative SVM's
// libsvm call (contrib/.../svm/SVMMapper.java).  SMO is a chain of tiny dependent steps (pick the maximal violating
// pair, move two multipliers, patch the gradient with two kernel-matrix rows): run from the host it costs a
// handful of launches and a device->host sync per step; here one CTA keeps iterating by itself:
//   * 1024 threads stride over the n multipliers: f_k = -y_k grad_k, arg-max over the "up" set, arg-min over the
//     "low" set (warp shuffles, then one shared-memory round; ties go to the lowest index, like the host solver
//     csrc/host/smo.cpp, which is this kernel's oracle);
//   * thread 0 clips the step to the box and updates the two multipliers;
//   * all threads patch grad with rows i and j of the (symmetric, precomputed) kernel matrix.
// The kernel matrix itself comes from a GEMM (daal/classify.py::kernel_linear / kernel_rbf).  One CTA is the right
// grain: a one-vs-one multi-class model is k(k-1)/2 independent problems, launched on separate streams.
#include "common.cuh"
#include "../include/harp_b200.h"

#include <math_constants.h>

namespace hb {
namespace svm {

constexpr int SMO_THREADS = 1024;

struct Best { double v; int idx; };

__device__ __forceinline__ Best better_max(Best a, Best b) {   // larger value; on ties the lower index
  return (b.v > a.v || (b.v == a.v && b.idx >= 0 && (a.idx < 0 || b.idx < a.idx))) ? b : a;
}
__device__ __forceinline__ Best better_min(Best a, Best b) {
  return (b.v < a.v || (b.v == a.v && b.idx >= 0 && (a.idx < 0 || b.idx < a.idx))) ? b : a;
}
__device__ __forceinline__ Best shfl_down(Best x, int o) {
  Best r;
  r.v = __shfl_down_sync(0xffffffffu, x.v, o);
  r.idx = __shfl_down_sync(0xffffffffu, x.idx, o);
  return r;
}

__global__ void __launch_bounds__(SMO_THREADS, 1)
smo_kernel(const double* __restrict__ K, const double* __restrict__ y, int n, double C, double tol, int max_iter,
           double* __restrict__ alpha, double* __restrict__ grad, int* __restrict__ iters_out) {
  __shared__ Best s_up[32], s_lo[32];
  __shared__ int s_i, s_j, s_stop;
  __shared__ double s_t;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int k = tid; k < n; k += SMO_THREADS) { alpha[k] = 0.0; grad[k] = -1.0; }
  __syncthreads();
  int it = 0;
  for (; it < max_iter; ++it) {
    Best up = {-CUDART_INF, -1}, lo = {CUDART_INF, -1};
    for (int k = tid; k < n; k += SMO_THREADS) {
      const double yk = y[k], ak = alpha[k];
      const double f = -yk * grad[k];
      const bool in_up = (yk > 0 && ak < C) || (yk < 0 && ak > 0);
      const bool in_lo = (yk > 0 && ak > 0) || (yk < 0 && ak < C);
      if (in_up) up = better_max(up, Best{f, k});
      if (in_lo) lo = better_min(lo, Best{f, k});
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      up = better_max(up, shfl_down(up, o));
      lo = better_min(lo, shfl_down(lo, o));
    }
    if (lane == 0) { s_up[warp] = up; s_lo[warp] = lo; }
    __syncthreads();
    if (warp == 0) {
      up = s_up[lane];
      lo = s_lo[lane];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        up = better_max(up, shfl_down(up, o));
        lo = better_min(lo, shfl_down(lo, o));
      }
      if (lane == 0) {
        const int bi = up.idx, bj = lo.idx;
        int stop = (bi < 0 || bj < 0 || up.v - lo.v < tol) ? 1 : 0;
        if (!stop) {
          double eta = K[(size_t)bi * n + bi] + K[(size_t)bj * n + bj] - 2.0 * K[(size_t)bi * n + bj];
          if (!(eta > 1e-12)) eta = 1e-12;
          const double yi = y[bi], yj = y[bj];
          double t = (up.v - lo.v) / eta;                     // alpha_i += y_i t, alpha_j -= y_j t, t >= 0
          const double ti = yi > 0 ? C - alpha[bi] : alpha[bi];
          const double tj = yj > 0 ? alpha[bj] : C - alpha[bj];
          if (ti < t) t = ti;
          if (tj < t) t = tj;
          alpha[bi] += yi * t;
          alpha[bj] -= yj * t;
          s_t = t;
        }
        s_i = bi; s_j = bj; s_stop = stop;
      }
    }
    __syncthreads();
    if (s_stop) break;
    const double t = s_t;
    const double* Ki = K + (size_t)s_i * n;                   // K is symmetric: row i = column i
    const double* Kj = K + (size_t)s_j * n;
    for (int k = tid; k < n; k += SMO_THREADS) grad[k] += t * y[k] * (Ki[k] - Kj[k]);
    __syncthreads();
  }
  if (tid == 0 && iters_out != nullptr) *iters_out = it;
}

}  // namespace svm
}  // namespace hb

using namespace hb;

extern "C" {

// One binary C-SVC problem: K[n,n] (row-major, symmetric, fp64), y in {-1,+1}; outputs alpha[n], grad[n] and the
// number of SMO steps (device int, may be NULL).  Everything is stream-ordered; nothing is read back.
int hb_smo_binary(const double* K, const double* y, int n, double C, double tol, int max_iter, double* alpha,
                  double* grad, int* iters_out, void* stream) {
  HB_REQUIRE(n >= 1 && K != nullptr && y != nullptr && alpha != nullptr && grad != nullptr, "hb_smo_binary: bad arguments");
  HB_REQUIRE(C > 0 && max_iter >= 0, "hb_smo_binary: need C > 0 and max_iter >= 0");
  svm::smo_kernel<<<1, svm::SMO_THREADS, 0, (cudaStream_t)stream>>>(K, y, n, C, tol, max_iter, alpha, grad, iters_out);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
