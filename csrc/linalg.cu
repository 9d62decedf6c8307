// harp_b200 — dense linear-algebra kernels behind the DAAL-style algorithms (sm_100a).
//
//   K21  tsqr_kernel        R factor of a tall-skinny fp32 matrix by Householder reflections, one streaming pass
//                           over X (daal_qr / daal_svd / daal_pca-svd step 1 and step 2,
//                           ml/daal/.../daal_svd/SVDDaalCollectiveMapper.java:136-219).
//   K16  potrs_batched      batched f x f Cholesky factor + solve, one warp per system.
//   K22  als_solve_kernel   implicit-ALS step 4: per row, the normal matrix Y'Y + Y_u'(C_u - I)Y_u + lambda I is built
//                           in registers, factored and solved in shared memory by one warp
//                           (ml/daal/.../daal_als/ALSDaalCollectiveMapper.java:354-492).
//
// TSQR design.  A CTA owns a contiguous range of rows and keeps ONE running upper-triangular R (D x D) in shared
// memory.  It streams tiles of 128*RPT rows: every thread holds RPT whole rows of the tile in registers, and the
// stacked matrix [R ; tile] is reduced column by column with the structured reflector v = [R_kk - alpha ; tile[:,k]]
// (rows k+1.. of R are already zero in column k, so the reflector only touches row k of R and the tile).  The dot
// products of column k with columns k..D-1 do not depend on the norm, so the norm and all the projections come out of
// ONE block-wide reduction per column: a transposing butterfly inside the warp (31 shuffles for 32 sums), four
// partials through shared memory, two barriers per column.  While a tile is being reduced the next one is already in
// flight: the rows are fetched with cp.async.bulk into a padded shared buffer (pitch D+4 floats keeps the 128-bit
// row reads conflict-free) that is free as soon as the previous tile sits in registers.  X is read exactly once; the
// per-CTA triangles are stacked and reduced by the same kernel until one is left.
#include "common.cuh"
#include "harp_b200.h"

namespace hb {
namespace tsqr {

constexpr int THREADS = 128;          // two CTAs per SM: one CTA's barrier and reduction latency hides behind the other
constexpr int NW = THREADS / 32;

template <int D>
struct Cfg {
  static constexpr int RPT = D <= 16 ? 8 : (D <= 32 ? 4 : 2);   // rows per thread
  static constexpr int TILE = THREADS * RPT;
  static constexpr int PITCH = D + 4;
  static constexpr int CW = D < 32 ? D : 32;                    // sums per butterfly
  static constexpr size_t smem_bytes() {
    return (size_t)TILE * PITCH * 4 + (size_t)D * D * 4 + (size_t)NW * D * 4 + (size_t)D * 4 + 64;
  }
};

// One column of the reduction.  K is a template parameter so that every register-array index is static.
template <int D, int RPT, int K>
__device__ __forceinline__ void column_step(float (&b)[RPT][D], float* __restrict__ Rm, float* __restrict__ part,
                                            float* __restrict__ wv, const int warp, const int lane, const int tid) {
  constexpr int CW = Cfg<D>::CW;
  constexpr int C_FIRST = (K / CW) * CW;
  // ---- stage 1: per-warp sums of tile[:,K] * tile[:,j] for j >= K ----
#pragma unroll
  for (int c0 = C_FIRST; c0 < D; c0 += CW) {
    float p[CW];
#pragma unroll
    for (int i = 0; i < CW; ++i) {
      const int j = c0 + i;
      float s = 0.f;
      if (j >= K) {
#pragma unroll
        for (int r = 0; r < RPT; ++r) s = fmaf(b[r][K], b[r][j], s);
      }
      p[i] = s;
    }
    int lo = K - c0 > 0 ? K - c0 : 0;              // slots below `lo` are identically zero in every lane
#pragma unroll
    for (int h = CW / 2; h >= 1; h >>= 1) {
      const bool up = (lane & h) != 0;
#pragma unroll
      for (int i = 0; i < h; ++i) {
        if (i + h >= lo) {
          const float keep = up ? p[i + h] : p[i];
          const float send = up ? p[i] : p[i + h];
          p[i] = keep + __shfl_xor_sync(0xffffffffu, send, h);
        }
      }
      lo = lo - h > 0 ? lo - h : 0;
    }
    if (CW < 32) p[0] += __shfl_xor_sync(0xffffffffu, p[0], 16);
    const int j = c0 + (lane & (CW - 1));
    if (lane < CW && j >= K) part[warp * D + j] = p[0];
  }
  __syncthreads();
  // ---- stage 2: thread j finishes column j: new R[K][j] and the coefficient of the row update ----
  if (tid >= K && tid < D) {
    float gk = 0.f, gj = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      gk += part[w * D + K];
      gj += part[w * D + tid];
    }
    const float rkk = Rm[K * D + K];
    float coef = 0.f;
    const float normsq = fmaf(rkk, rkk, gk);
    if (gk > 0.f && normsq > 1e-34f) {             // below that alpha * v0 would leave the normal range
      const float alpha = -copysignf(normsq * rsqrtf(normsq), rkk);
      const float v0 = rkk - alpha;
      const float tau = __frcp_rn(0.5f * fmaf(v0, v0, gk));   // 2 / (v'v): exactly orthogonal for the alpha in use
      if (tid == K) {
        Rm[K * D + K] = alpha;
      } else {
        const float rkj = Rm[K * D + tid];
        const float w = fmaf(v0, rkj, gj);
        Rm[K * D + tid] = rkj - tau * v0 * w;
        coef = tau * w;
      }
    }
    wv[tid] = coef;
  }
  __syncthreads();
  // ---- stage 3: tile[:, j] -= tile[:, K] * coef_j ----
  if (K + 1 < D) {
#pragma unroll
    for (int j4 = ((K + 1) / 4) * 4; j4 < D; j4 += 4) {
      const float4 c = *reinterpret_cast<const float4*>(wv + j4);
      const float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int j = j4 + q;
        if (j > K) {
#pragma unroll
          for (int r = 0; r < RPT; ++r) b[r][j] = fmaf(-b[r][K], cc[q], b[r][j]);
        }
      }
    }
  }
}

template <int D, int RPT, int K>
struct Columns {
  static __device__ __forceinline__ void run(float (&b)[RPT][D], float* Rm, float* part, float* wv, int warp, int lane,
                                             int tid) {
    column_step<D, RPT, K>(b, Rm, part, wv, warp, lane, tid);
    Columns<D, RPT, K + 1>::run(b, Rm, part, wv, warp, lane, tid);
  }
};
template <int D, int RPT>
struct Columns<D, RPT, D> {
  static __device__ __forceinline__ void run(float (&)[RPT][D], float*, float*, float*, int, int, int) {}
};

// X: [n, d] fp32, row pitch `ld` floats.  Every CTA reduces rows [cta*rows_per_cta, ...) and writes its triangle to
// out[cta][D][D] (row-major, zeros below the diagonal, columns >= d zero).  BULK: rows are fetched with cp.async.bulk
// (needs d == D, ld % 4 == 0, 16-byte aligned base); otherwise the tile is staged with plain coalesced loads.
template <int D, bool BULK>
__global__ void __launch_bounds__(THREADS, 2)
tsqr_kernel(const float* __restrict__ X, const long long n, const int d, const long long ld,
            const long long rows_per_cta, float* __restrict__ out) {
  using C = Cfg<D>;
  constexpr int RPT = C::RPT, TILE = C::TILE, PITCH = C::PITCH;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  float* Rm = tile + (size_t)TILE * PITCH;
  float* part = Rm + D * D;
  float* wv = part + NW * D;
  uint64_t* bar = reinterpret_cast<uint64_t*>(wv + D);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long r_begin = (long long)blockIdx.x * rows_per_cta;
  long long r_end = r_begin + rows_per_cta;
  if (r_end > n) r_end = n;
  for (int i = tid; i < D * D; i += THREADS) Rm[i] = 0.f;
  if (BULK && tid == 0) {
    mbar_init(smem_u32(bar), 1);
    fence_barrier_init();
  }
  __syncthreads();

  auto issue = [&](long long row0) {                       // fetch rows [row0, row0 + TILE) ∩ [.., r_end)
    if (BULK) {
      long long left = r_end - row0;
      const int rows = left > TILE ? TILE : (int)left;
      if (tid == 0) mbar_expect_tx(smem_u32(bar), (uint32_t)rows * D * 4u);
#pragma unroll
      for (int r = 0; r < RPT; ++r) {
        const int lr = tid + r * THREADS;
        if (lr < rows) bulk_load_1d(smem_u32(tile + (size_t)lr * PITCH), X + (row0 + lr) * ld, D * 4, smem_u32(bar));
      }
    }
  };

  uint32_t phase = 0;
  if (r_begin < r_end) issue(r_begin);
  for (long long row0 = r_begin; row0 < r_end; row0 += TILE) {
    long long left = r_end - row0;
    const int rows = left > TILE ? TILE : (int)left;
    if (BULK) {
      mbar_wait(smem_u32(bar), phase);
      phase ^= 1;
    } else {
      for (int i = tid; i < rows * D; i += THREADS) {
        const int lr = i / D, c = i - lr * D;
        tile[(size_t)lr * PITCH + c] = c < d ? X[(row0 + lr) * ld + c] : 0.f;
      }
      __syncthreads();
    }
    float b[RPT][D];
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
      const int lr = tid + r * THREADS;
      if (lr < rows) {
#pragma unroll
        for (int c = 0; c < D; c += 4) {
          const float4 v = *reinterpret_cast<const float4*>(tile + (size_t)lr * PITCH + c);
          b[r][c] = v.x; b[r][c + 1] = v.y; b[r][c + 2] = v.z; b[r][c + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int c = 0; c < D; ++c) b[r][c] = 0.f;
      }
    }
    __syncthreads();                                       // the staging buffer is free again
    if (row0 + TILE < r_end) issue(row0 + TILE);
    Columns<D, RPT, 0>::run(b, Rm, part, wv, warp, lane, tid);
  }
  __syncthreads();
  float* o = out + (size_t)blockIdx.x * D * D;
  for (int i = tid; i < D * D; i += THREADS) {
    const int r = i / D, c = i - r * D;
    o[i] = c >= r ? Rm[i] : 0.f;
  }
}

template <int D>
int launch(const float* X, long long n, int d, long long ld, int grid, long long rows_per_cta, float* out,
           cudaStream_t stream) {
  const bool bulk = d == D && ld % 4 == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0;
  const size_t smem = Cfg<D>::smem_bytes();
  if (bulk) {
    HB_CUDA_OK(cudaFuncSetAttribute(tsqr_kernel<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tsqr_kernel<D, true><<<grid, THREADS, smem, stream>>>(X, n, d, ld, rows_per_cta, out);
  } else {
    HB_CUDA_OK(cudaFuncSetAttribute(tsqr_kernel<D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tsqr_kernel<D, false><<<grid, THREADS, smem, stream>>>(X, n, d, ld, rows_per_cta, out);
  }
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int D>
int run(const float* X, long long n, int d, long long ld, float* work, float* R_out, cudaStream_t stream) {
  // level 0 over X, then over the stacked triangles until one CTA is enough
  constexpr int TILE = Cfg<D>::TILE;
  const float* src = X;
  long long rows = n, pitch = ld;
  int cols = d;
  float* buf[2] = {work, work + (size_t)2 * sm_count() * D * D};
  int which = 0;
  for (;;) {
    long long tiles = (rows + TILE - 1) / TILE;
    int grid = src == X ? 2 * sm_count() : (int)((tiles + 3) / 4);
    if (grid > tiles) grid = (int)tiles;
    if (grid < 1) grid = 1;
    long long per = ((tiles + grid - 1) / grid) * TILE;
    grid = (int)((rows + per - 1) / per);
    if (grid < 1) grid = 1;
    float* dst = grid == 1 ? R_out : buf[which];
    if (int rc = launch<D>(src, rows, cols, pitch, grid, per, dst, stream)) return rc;
    if (grid == 1) return 0;
    src = dst;
    rows = (long long)grid * D;
    pitch = D;
    cols = D;
    which ^= 1;
  }
}

}  // namespace tsqr

// ------------------------------------------------------------------------------------------------------------
// Cholesky factor + solve of one f x f system by one warp, matrix in shared memory with pitch F + 1.
// ------------------------------------------------------------------------------------------------------------
namespace chol {

// In place: lower triangle of A becomes L (A = L L').  Returns false if a pivot is not positive.
template <int F>
__device__ __forceinline__ bool factor(float* __restrict__ A, const int f, const int lane) {
  constexpr int P = F + 1;
  bool ok = true;
  for (int k = 0; k < f; ++k) {
    const float akk = A[k * P + k];
    if (!(akk > 0.f)) ok = false;
    const float inv = rsqrtf(fmaxf(akk, 1e-30f));
    __syncwarp();
    for (int i = k + lane; i < f; i += 32) A[i * P + k] = i == k ? akk * inv : A[i * P + k] * inv;
    __syncwarp();
    // column k now holds L[k][k] = sqrt(akk) (akk * inv) and L[i][k] = A[i][k] * inv
    for (int i = k + 1 + lane; i < f; i += 32) {
      const float lik = A[i * P + k];
      for (int j = k + 1; j <= i; ++j) A[i * P + j] = fmaf(-lik, A[j * P + k], A[i * P + j]);
    }
    __syncwarp();
  }
  return ok;
}

// Solves L L' x = b in place (b in shared memory).
template <int F>
__device__ __forceinline__ void solve(const float* __restrict__ A, float* __restrict__ b, const int f, const int lane) {
  constexpr int P = F + 1;
  for (int k = 0; k < f; ++k) {                  // forward: L z = b
    const float zk = b[k] / A[k * P + k];
    __syncwarp();
    if (lane == 0) b[k] = zk;
    for (int i = k + 1 + lane; i < f; i += 32) b[i] = fmaf(-A[i * P + k], zk, b[i]);
    __syncwarp();
  }
  for (int k = f - 1; k >= 0; --k) {             // backward: L' x = z
    const float xk = b[k] / A[k * P + k];
    __syncwarp();
    if (lane == 0) b[k] = xk;
    for (int i = lane; i < k; i += 32) b[i] = fmaf(-A[k * P + i], xk, b[i]);
    __syncwarp();
  }
}

constexpr int WARPS = 4;

// A: [batch][f][f] symmetric positive definite (only the lower triangle is read), B: [batch][f].
// X: [batch][f] solutions; L_out (optional): [batch][f][f] lower factors; info[0] counts failed factorisations.
template <int F>
__global__ void __launch_bounds__(WARPS * 32)
potrs_batched_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ X,
                     float* __restrict__ L_out, const int f, const long long batch, int* __restrict__ info) {
  extern __shared__ __align__(16) float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* As = sm + (size_t)warp * (F * (F + 1) + F);
  float* bs = As + F * (F + 1);
  for (long long m = (long long)blockIdx.x * WARPS + warp; m < batch; m += (long long)gridDim.x * WARPS) {
    const float* a = A + (size_t)m * f * f;
    for (int i = lane; i < f * f; i += 32) {
      const int r = i / f, c = i - r * f;
      As[r * (F + 1) + c] = a[i];
    }
    for (int i = lane; i < f; i += 32) bs[i] = B ? B[(size_t)m * f + i] : 0.f;
    __syncwarp();
    const bool ok = factor<F>(As, f, lane);
    if (!ok && lane == 0) atomicAdd(info, 1);
    if (L_out) {
      float* l = L_out + (size_t)m * f * f;
      for (int i = lane; i < f * f; i += 32) {
        const int r = i / f, c = i - r * f;
        l[i] = c <= r ? As[r * (F + 1) + c] : 0.f;
      }
    }
    if (X) {
      solve<F>(As, bs, f, lane);
      for (int i = lane; i < f; i += 32) X[(size_t)m * f + i] = bs[i];
    }
    __syncwarp();
  }
}

}  // namespace chol

// ------------------------------------------------------------------------------------------------------------
// Implicit ALS, step 4: x_u = (Y'Y + Y_u'(C_u - I)Y_u + lambda I)^-1 Y_u' C_u p_u,  c_ui = 1 + alpha r_ui.
// One warp per row u.  The 32 lanes form a 4 x 8 grid; lane (lr, lc) keeps the (F/4) x (F/8) block of the normal
// matrix in registers and adds one weighted outer product per rating.  Ratings are taken four at a time: the four
// factor rows are fetched with coalesced loads, parked in shared memory, and every lane picks the F/4 + F/8 entries
// it needs.  Rows are claimed dynamically (row lengths are heavy-tailed).
// ------------------------------------------------------------------------------------------------------------
namespace als {

constexpr int WARPS = 8;
constexpr int BATCH = 4;

template <int F>
constexpr size_t warp_smem_floats() { return (size_t)F * (F + 1) + F + BATCH * F; }

template <int F>
__global__ void __launch_bounds__(WARPS * 32)
solve_kernel(const long long* __restrict__ ptr, const int* __restrict__ idx, const float* __restrict__ val,
             const float* __restrict__ Y, const float* __restrict__ YtY, const float lam, const float alpha,
             const int* __restrict__ rows, const int n_rows, const int f, float* __restrict__ Xout,
             int* __restrict__ counter, int* __restrict__ info) {
  constexpr int TR = F / 4, TC = F / 8, Q = (F + 31) / 32, P = F + 1;
  extern __shared__ __align__(16) float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* As = sm + (size_t)warp * warp_smem_floats<F>();
  float* bs = As + F * P;
  float* ys = bs + F;
  const int lr = lane >> 3, lc = lane & 7;
  const int r0 = lr * TR, c0 = lc * TC;
  for (;;) {
    int claim = 0;
    if (lane == 0) claim = atomicAdd(counter, 1);
    claim = __shfl_sync(0xffffffffu, claim, 0);
    if (claim >= n_rows) break;
    const int u = rows ? rows[claim] : claim;
    const long long a = ptr[u], e = ptr[u + 1];
    float acc[TR][TC];
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
      for (int j = 0; j < TC; ++j) {
        const int r = r0 + i, c = c0 + j;
        float v = YtY[r * F + c];
        if (r == c) v += r < f ? lam : 1.f;        // padded coordinates get an identity block
        acc[i][j] = v;
      }
    float bv[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) bv[q] = 0.f;
    for (long long t = a; t < e; t += BATCH) {
      float yl[BATCH][Q], w[BATCH];
#pragma unroll
      for (int s = 0; s < BATCH; ++s) {
        const bool on = t + s < e;
        const int it = on ? idx[t + s] : 0;
        w[s] = on ? alpha * val[t + s] : 0.f;
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int c = lane + 32 * q;
          yl[s][q] = (on && c < F) ? Y[(size_t)it * F + c] : 0.f;
        }
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < BATCH; ++s)
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int c = lane + 32 * q;
          if (c < F) ys[s * F + c] = yl[s][q];
          bv[q] = fmaf(t + s < e ? 1.f + w[s] : 0.f, yl[s][q], bv[q]);
        }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < BATCH; ++s) {
        float yr[TR], yc[TC];
#pragma unroll
        for (int i = 0; i < TR; ++i) yr[i] = ys[s * F + r0 + i] * w[s];
#pragma unroll
        for (int j = 0; j < TC; ++j) yc[j] = ys[s * F + c0 + j];
#pragma unroll
        for (int i = 0; i < TR; ++i)
#pragma unroll
          for (int j = 0; j < TC; ++j) acc[i][j] = fmaf(yr[i], yc[j], acc[i][j]);
      }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
      for (int j = 0; j < TC; ++j) As[(r0 + i) * P + c0 + j] = acc[i][j];
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int c = lane + 32 * q;
      if (c < F) bs[c] = bv[q];
    }
    __syncwarp();
    if (a == e) {                                   // no ratings: the factor row is zero (as the reference leaves it)
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = lane + 32 * q;
        if (c < F) Xout[(size_t)u * F + c] = 0.f;
      }
      continue;
    }
    if (!chol::factor<F>(As, f, lane) && lane == 0) atomicAdd(info, 1);
    chol::solve<F>(As, bs, f, lane);
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int c = lane + 32 * q;
      if (c < F) Xout[(size_t)u * F + c] = c < f ? bs[c] : 0.f;
    }
    __syncwarp();
  }
}

}  // namespace als
}  // namespace hb

// ------------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------------
using namespace hb;

extern "C" {

// Padded column count the TSQR kernel runs a d-column matrix at (0: not supported, use the library route).
int hb_tsqr_pad(int d) { return d <= 0 ? 0 : d <= 16 ? 16 : d <= 32 ? 32 : d <= 64 ? 64 : 0; }

size_t hb_tsqr_workspace_bytes(int d) {
  const int D = hb_tsqr_pad(d);
  return (size_t)4 * sm_count() * D * D * sizeof(float);
}

// R_out: [D, D] fp32 with D = hb_tsqr_pad(d); the leading d x d block is the triangle (signs not normalised).
int hb_tsqr(const float* X, long long n, int d, long long ld, float* work, float* R_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int D = hb_tsqr_pad(d);
  HB_REQUIRE(D > 0, "hb_tsqr: d = %d is outside 1..64", d);
  HB_REQUIRE(n >= 0 && ld >= d, "hb_tsqr: bad shape (n=%lld, d=%d, ld=%lld)", n, d, ld);
  if (n == 0) {
    HB_CUDA_OK(cudaMemsetAsync(R_out, 0, (size_t)D * D * sizeof(float), stream));
    return 0;
  }
  switch (D) {
    case 16: return tsqr::run<16>(X, n, d, ld, work, R_out, stream);
    case 32: return tsqr::run<32>(X, n, d, ld, work, R_out, stream);
    default: return tsqr::run<64>(X, n, d, ld, work, R_out, stream);
  }
}

int hb_chol_pad(int f) { return f <= 0 ? 0 : f <= 16 ? 16 : f <= 32 ? 32 : f <= 64 ? 64 : 0; }

// Batched Cholesky: A [batch, f, f] (lower triangle read), B [batch, f] or NULL, X [batch, f] or NULL,
// L [batch, f, f] or NULL; info[0] is incremented once per matrix with a non-positive pivot.
int hb_potrs_batched(const float* A, const float* B, float* X, float* L, int f, long long batch, int* info,
                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int F = hb_chol_pad(f);
  HB_REQUIRE(F > 0, "hb_potrs_batched: f = %d is outside 1..64", f);
  HB_REQUIRE(X == nullptr || B != nullptr, "hb_potrs_batched: a solve needs right-hand sides");
  if (batch <= 0) return 0;
  long long blocks = (batch + chol::WARPS - 1) / chol::WARPS;
  const long long cap = (long long)sm_count() * 8;
  const int grid = (int)(blocks < cap ? blocks : cap);
#define HB_POTRS(FF)                                                                                             \
  {                                                                                                              \
    const size_t smem = (size_t)chol::WARPS * (FF * (FF + 1) + FF) * sizeof(float);                              \
    HB_CUDA_OK(cudaFuncSetAttribute(chol::potrs_batched_kernel<FF>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                    (int)smem));                                                                 \
    chol::potrs_batched_kernel<FF><<<grid, chol::WARPS * 32, smem, stream>>>(A, B, X, L, f, batch, info);        \
  }
  switch (F) {
    case 16: HB_POTRS(16) break;
    case 32: HB_POTRS(32) break;
    default: HB_POTRS(64) break;
  }
#undef HB_POTRS
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// One half-iteration of implicit ALS for `n_rows` rows (rows[] lists them, NULL = 0..n_rows-1).
// ptr/idx/val: CSR of the ratings of THIS side's rows (ptr indexed by row id); Y: the other side's factors
// [n_other, F] with F = hb_chol_pad(f), columns >= f zero; YtY: [F, F]; Xout: [*, F]; counter/info: device ints.
int hb_als_solve(const long long* ptr, const int* idx, const float* val, const float* Y, const float* YtY, float lam,
                 float alpha, const int* rows, int n_rows, int f, float* Xout, int* counter, int* info,
                 void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int F = hb_chol_pad(f);
  HB_REQUIRE(F > 0, "hb_als_solve: f = %d is outside 1..64", f);
  if (n_rows <= 0) return 0;
  HB_CUDA_OK(cudaMemsetAsync(counter, 0, sizeof(int), stream));
  int grid = sm_count() * (F == 64 ? 1 : 2);
  const int need = (n_rows + als::WARPS - 1) / als::WARPS;
  if (grid > need) grid = need;
#define HB_ALS(FF)                                                                                          \
  {                                                                                                         \
    const size_t smem = (size_t)als::WARPS * als::warp_smem_floats<FF>() * sizeof(float);                   \
    HB_CUDA_OK(cudaFuncSetAttribute(als::solve_kernel<FF>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                    (int)smem));                                                            \
    als::solve_kernel<FF><<<grid, als::WARPS * 32, smem, stream>>>(ptr, idx, val, Y, YtY, lam, alpha, rows, \
                                                                   n_rows, f, Xout, counter, info);         \
  }
  switch (F) {
    case 16: HB_ALS(16) break;
    case 32: HB_ALS(32) break;
    default: HB_ALS(64) break;
  }
#undef HB_ALS
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
