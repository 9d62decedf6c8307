// Multi-threaded ASCII number parser for the loaders (HarpDAALDataSource / MTReader /
// VStore / PointLoadTask equivalents): splits the buffer at separator boundaries, parses the
// chunks in parallel with a hand-rolled strtod-free fast path, then compacts the results.
// Separators: blank, tab, comma, CR, LF, ':' (libsvm / "id : v v v" model files).
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace {

inline bool is_sep(char c) { return c == ' ' || c == '\t' || c == ',' || c == '\n' || c == '\r' || c == ':'; }

// Parses one token [p, e); falls back to strtod for anything unusual (hex, inf, nan, > 18 digits).
inline double parse_token(const char* p, const char* e) {
  const char* s = p;
  bool neg = false;
  if (s < e && (*s == '-' || *s == '+')) { neg = (*s == '-'); ++s; }
  uint64_t mant = 0;
  int digits = 0, frac = 0;
  bool seen_dot = false, ok = s < e;
  for (; s < e; ++s) {
    const char c = *s;
    if (c >= '0' && c <= '9') {
      if (digits < 18) { mant = mant * 10 + uint64_t(c - '0'); ++digits; if (seen_dot) ++frac; }
      else if (!seen_dot) { ok = false; break; }   // too many integer digits: let strtod handle it
    } else if (c == '.' && !seen_dot) {
      seen_dot = true;
    } else {
      break;
    }
  }
  int exp10 = 0;
  if (ok && s < e && (*s == 'e' || *s == 'E')) {
    ++s;
    bool eneg = false;
    if (s < e && (*s == '-' || *s == '+')) { eneg = (*s == '-'); ++s; }
    int ev = 0, ed = 0;
    for (; s < e && *s >= '0' && *s <= '9'; ++s, ++ed) ev = ev * 10 + (*s - '0');
    if (ed == 0) ok = false;
    exp10 = eneg ? -ev : ev;
  }
  if (!ok || s != e || digits == 0) {
    char buf[64];
    size_t n = size_t(e - p) < sizeof(buf) - 1 ? size_t(e - p) : sizeof(buf) - 1;
    memcpy(buf, p, n);
    buf[n] = 0;
    return strtod(buf, nullptr);
  }
  const int e10 = exp10 - frac;
  double v = double(mant);
  static const double P10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                               1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  if (digits <= 15 && e10 >= -22 && e10 <= 22) {
    v = e10 < 0 ? v / P10[-e10] : v * P10[e10];   // exact for <= 15 digits (Clinger fast path)
  } else {
    char buf[64];
    size_t n = size_t(e - p) < sizeof(buf) - 1 ? size_t(e - p) : sizeof(buf) - 1;
    memcpy(buf, p, n);
    buf[n] = 0;
    return strtod(buf, nullptr);
  }
  return neg ? -v : v;
}

void parse_range(const char* b, const char* e, std::vector<double>& out) {
  const char* p = b;
  while (p < e) {
    while (p < e && is_sep(*p)) ++p;
    const char* t = p;
    while (p < e && !is_sep(*p)) ++p;
    if (p > t) out.push_back(parse_token(t, p));
  }
}

}  // namespace

extern "C" {

// Parses every number in data[0, len) into out[0, cap); returns the count, or -1 if cap is too small.
long long hh_parse_doubles(const char* data, long long len, double* out, long long cap, int threads) {
  if (len <= 0) return 0;
  int nt = threads > 0 ? threads : int(std::thread::hardware_concurrency());
  if (nt < 1) nt = 1;
  if (len < (1 << 16)) nt = 1;
  if (nt > 64) nt = 64;
  std::vector<long long> cut(nt + 1);
  cut[0] = 0;
  cut[nt] = len;
  for (int i = 1; i < nt; ++i) {
    long long c = len * i / nt;
    while (c < len && !is_sep(data[c])) ++c;   // never split a token
    cut[i] = c;
  }
  std::vector<std::vector<double>> parts(nt);
  std::vector<std::thread> th;
  for (int i = 0; i < nt; ++i) {
    if (nt == 1) {
      parse_range(data + cut[i], data + cut[i + 1], parts[i]);
    } else {
      th.emplace_back([&, i] { parse_range(data + cut[i], data + cut[i + 1], parts[i]); });
    }
  }
  for (auto& t : th) t.join();
  long long total = 0;
  for (auto& v : parts) total += (long long)v.size();
  if (total > cap) return -1;
  long long off = 0;
  for (auto& v : parts) {
    if (!v.empty()) memcpy(out + off, v.data(), v.size() * sizeof(double));
    off += (long long)v.size();
  }
  return total;
}

// Big-endian IEEE doubles (Java DataOutputStream.writeDouble) -> native float32/float64.
void hh_be_doubles_to_f32(const unsigned char* src, long long n, float* dst) {
  for (long long i = 0; i < n; ++i) {
    uint64_t v;
    memcpy(&v, src + i * 8, 8);
    v = __builtin_bswap64(v);
    double d;
    memcpy(&d, &v, 8);
    dst[i] = float(d);
  }
}

// COO (row, col) -> CSR row offsets by counting sort; returns the permutation that sorts by row
// (stable), so values/cols can be gathered.  (HarpDAALDataSource.COOToCSR, :439-496)
void hh_coo_to_csr(const long long* rows, long long nnz, long long nrows, long long* row_offsets,
                   long long* perm) {
  for (long long r = 0; r <= nrows; ++r) row_offsets[r] = 0;
  for (long long i = 0; i < nnz; ++i) ++row_offsets[rows[i] + 1];
  for (long long r = 0; r < nrows; ++r) row_offsets[r + 1] += row_offsets[r];
  std::vector<long long> cursor(row_offsets, row_offsets + nrows);
  for (long long i = 0; i < nnz; ++i) perm[cursor[rows[i]]++] = i;
}

int hh_version() { return 100; }

}  // extern "C"
