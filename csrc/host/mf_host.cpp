// Sequential host restatements of the MF-SGD and LDA-CGS passes: the CPU path of ops/mf.py / ops/lda.py and the
// "reference build stand-in" that BASELINE.md times on the host CPUs (the Java reference cannot run here).
//
// hh_sgd_pass   = SGDMPTask.doRun (ml/java/src/main/java/edu/iu/sgd/SGDMPTask.java:46-77) over a bucketed panel, in
//                 the device kernel's own order: Latin-square steps t, row-blocks i, column runs inside block
//                 (i, c(i,t)), ratings inside a run — exactly what ops/mf.py::_cpu_sgd_pass does in Python.
// hh_lda_pass   = LDAMPTask.sample (ml/java/src/main/java/edu/iu/lda/LDAMPTask.java:85-331) with dense rows:
//                 p_k = (n_dk + alpha)(n_wk + beta) / (n_k + V beta), counter-based RNG like the device kernel.
#include <cstdint>
#include <vector>

namespace {
inline int col_block(int i, int t, int B, int C) { return C != B ? (int)(((long long)i * C / B + t) % C) : (i + t) % B; }

inline double uniform01(uint64_t seed, uint64_t ctr) {
  uint64_t x = seed + ctr * 0x9E3779B97F4A7C15ULL;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
  x ^= x >> 27; x *= 0x94D049BB133111EBULL;
  x ^= x >> 31;
  return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}
}  // namespace

extern "C" {

// W [rows, rank], H [cols, rank] row-major, T = double or float selected by `is_double`.  Returns ratings trained.
long long hh_sgd_pass(void* W_, void* H_, int rank, const int* blk_seg_ptr, const int* seg_col, const int* seg_ptr,
                      const int* r_row, const float* r_val, int B, int C, double eps, double lambda, int is_double) {
  long long n = 0;
  std::vector<double> w(rank), h(rank);
  for (int t = 0; t < C; ++t) {
    for (int i = 0; i < B; ++i) {
      const int c = col_block(i, t, B, C);
      for (int s = blk_seg_ptr[(size_t)i * C + c]; s < blk_seg_ptr[(size_t)i * C + c + 1]; ++s) {
        const size_t col = (size_t)(seg_col[s] & 0x7FFFFFFF);
        for (int f = 0; f < rank; ++f)
          h[f] = is_double ? ((double*)H_)[col * rank + f] : (double)((float*)H_)[col * rank + f];
        for (int k = seg_ptr[s]; k < seg_ptr[s + 1]; ++k) {
          const size_t row = (size_t)r_row[k];
          double dot = 0.0;
          for (int f = 0; f < rank; ++f) {
            w[f] = is_double ? ((double*)W_)[row * rank + f] : (double)((float*)W_)[row * rank + f];
            dot += w[f] * h[f];
          }
          const double e = dot - (double)r_val[k];
          for (int f = 0; f < rank; ++f) {
            const double wn = w[f] - eps * (e * h[f] + lambda * w[f]);
            h[f] = h[f] - eps * (e * w[f] + lambda * h[f]);
            if (is_double) ((double*)W_)[row * rank + f] = wn; else ((float*)W_)[row * rank + f] = (float)wn;
          }
          ++n;
        }
        for (int f = 0; f < rank; ++f) {
          if (is_double) ((double*)H_)[col * rank + f] = h[f]; else ((float*)H_)[col * rank + f] = (float)h[f];
        }
      }
    }
  }
  return n;
}

// n_dk [docs, K], n_wk [words, K] int32 (n_dk may be int16 when dk16 != 0); commons[k] = 1 / (n_k + V beta).
long long hh_lda_pass(void* n_dk_, int dk16, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                      const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, double alpha, double beta,
                      unsigned long long seed) {
  long long n = 0;
  std::vector<double> p(K);
  for (int t = 0; t < B; ++t) {
    for (int i = 0; i < B; ++i) {
      const int c = (i + t) % B;
      for (int s = blk_seg_ptr[(size_t)i * B + c]; s < blk_seg_ptr[(size_t)i * B + c + 1]; ++s) {
        int* wrow = n_wk + (size_t)(seg_word[s] & 0x7FFFFFFF) * K;
        for (int tk = seg_ptr[s]; tk < seg_ptr[s + 1]; ++tk) {
          const size_t d = (size_t)tok_doc[tk];
          const int zo = tok_z[tk];
          int16_t* d16 = dk16 ? (int16_t*)n_dk_ + d * K : nullptr;
          int* d32 = dk16 ? nullptr : (int*)n_dk_ + d * K;
          if (dk16) d16[zo] -= 1; else d32[zo] -= 1;
          wrow[zo] -= 1;
          double total = 0.0;
          for (int k = 0; k < K; ++k) {
            const double nd = dk16 ? (double)d16[k] : (double)d32[k];
            // the frozen topic sums still count this token under its old topic, as in the reference iteration
            p[k] = (nd + alpha) * ((double)wrow[k] + beta) * (double)commons[k];
            total += p[k];
          }
          const double u = uniform01(seed, (uint64_t)tk * 2654435761ULL + (uint64_t)i) * total;
          double acc = 0.0;
          int zn = K - 1;
          for (int k = 0; k < K; ++k) { acc += p[k]; if (acc > u) { zn = k; break; } }
          if (dk16) d16[zn] += 1; else d32[zn] += 1;
          wrow[zn] += 1;
          tok_z[tk] = zn;
          ++n;
        }
      }
    }
  }
  return n;
}

}  // extern "C"
