// C-SVC dual by SMO with the maximal-violating-pair working-set rule (WSS-1), on a precomputed kernel matrix.
// Host counterpart of csrc/svm.cu::smo_kernel (same steps, same tie-breaking: lowest index wins), used for CPU
// tensors and as the oracle of the device kernel.  The reference runs this through DAAL's svm::training::Batch
// with the Boser method (ml/daal/.../daal_svm/MultiClassDenseBatch/SVMDaalCollectiveMapper.java:150-199).
//
//   minimise  1/2 a^T Q a - e^T a,   Q_ij = y_i y_j K_ij,   0 <= a_i <= C,   y^T a = 0
//
// grad_i = (Q a)_i - 1 is maintained incrementally; f_i = -y_i grad_i.
#include <cmath>
#include <limits>

extern "C" {

// K: n x n row-major, y in {-1,+1}.  alpha and grad are outputs (n doubles each).  Returns the iterations used.
long long hh_smo_binary(const double* K, const double* y, long long n, double C, double tol, long long max_iter,
                        double* alpha, double* grad) {
  for (long long i = 0; i < n; ++i) { alpha[i] = 0.0; grad[i] = -1.0; }
  long long it = 0;
  for (; it < max_iter; ++it) {
    double fi = -std::numeric_limits<double>::infinity(), fj = std::numeric_limits<double>::infinity();
    long long bi = -1, bj = -1;
    for (long long k = 0; k < n; ++k) {
      const double f = -y[k] * grad[k];
      const bool up = (y[k] > 0 && alpha[k] < C) || (y[k] < 0 && alpha[k] > 0);
      const bool lo = (y[k] > 0 && alpha[k] > 0) || (y[k] < 0 && alpha[k] < C);
      if (up && f > fi) { fi = f; bi = k; }
      if (lo && f < fj) { fj = f; bj = k; }
    }
    if (bi < 0 || bj < 0 || fi - fj < tol) break;
    double eta = K[bi * n + bi] + K[bj * n + bj] - 2.0 * K[bi * n + bj];
    if (!(eta > 1e-12)) eta = 1e-12;
    const double yi = y[bi], yj = y[bj];
    double t = (fi - fj) / eta;                                   // alpha_i += y_i t, alpha_j -= y_j t, t >= 0
    const double ti = yi > 0 ? C - alpha[bi] : alpha[bi];
    const double tj = yj > 0 ? alpha[bj] : C - alpha[bj];
    if (ti < t) t = ti;
    if (tj < t) t = tj;
    alpha[bi] += yi * t;
    alpha[bj] -= yj * t;
    const double* Ki = K + bi * n;                                // K is symmetric: row i = column i
    const double* Kj = K + bj * n;
    for (long long k = 0; k < n; ++k) grad[k] += t * y[k] * (Ki[k] - Kj[k]);
  }
  return it;
}

}  // extern "C"
