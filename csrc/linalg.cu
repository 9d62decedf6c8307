// harp_b200 — dense linear-algebra kernels behind the DAAL-style algorithms (sm_100a).
//
//   K21  tsqr_kernel        R factor of a tall-skinny fp32 matrix by Householder reflections, one streaming pass
//                           over X (daal_qr / daal_svd / daal_pca-svd step 1 and step 2,
//                           ml/daal/.../daal_svd/SVDDaalCollectiveMapper.java:136-219).
//   K16  potrs_batched      batched f x f Cholesky factor + solve, one warp per system.
//   K22  als_solve_kernel   implicit-ALS step 4: per row, the normal matrix Y'Y + Y_u'(C_u - I)Y_u + lambda I is built
//                           in registers, factored and solved by one warp
//                           (ml/daal/.../daal_als/ALSDaalCollectiveMapper.java:354-492).
//
// TSQR design.  A CTA owns a contiguous range of rows and keeps ONE running upper-triangular R (D x D) in shared
// memory.  It streams tiles; the stacked matrix [R ; tile] is reduced column by column with the structured reflector
// v = [R_kk - alpha ; tile[:,k]] (rows k+1.. of R are already zero in column k, so the reflector only touches row k
// of R and the tile).  The tile lives in registers, laid out so that NO cross-lane reduction is needed: a group of
// D/4 lanes holds 32 rows, lane c of the group owning columns c, c + D/4, c + 2D/4, c + 3D/4 of every one of them.
// Column k is broadcast from its owner lane with one shuffle per row; every lane then forms the dot products of that
// column with its own four columns and applies the update to them.  Per column the CTA exchanges one row of partial
// dot products through shared memory (one barrier); the scalar part of the reflector is recomputed by every lane.
// The dot products do not depend on the norm, so the norm comes out of the same exchange.  Because ownership is
// cyclic, after D/4 columns every lane's first register column is dead and the code moves on to a body that no
// longer touches it: four bodies, each a rolled loop over D/4 columns, ~1000 instructions in total.  (A first
// version kept whole rows per thread with all D columns unrolled; its 370 KB of straight-line code ran at 30 % issue
// utilisation waiting for instruction fetch.)  While a tile is being reduced the next one is in flight: every warp
// fetches its own 16 KB strip with cp.async.bulk into shared memory and signals its own mbarrier.  X is read exactly
// once; the per-CTA triangles are stacked and reduced by the same kernel until one is left.
#include "common.cuh"
#include "harp_b200.h"

namespace hb {
namespace tsqr {

constexpr int WARPS = 4;                     // two CTAs per SM: one CTA's barrier + reflector latency hides behind the other
constexpr int THREADS = WARPS * 32;
constexpr int CPL = 4;                       // columns per lane
constexpr int RW = 32;                       // rows per lane group

template <int D>
struct Cfg {
  static constexpr int LPG = D / CPL;        // lanes per group
  static constexpr int G = 32 / LPG;         // groups per warp
  static constexpr int STRIP = RW * G;       // rows per warp and tile
  static constexpr int TILE = WARPS * STRIP;
  static constexpr size_t smem_bytes() {
    return (size_t)TILE * D * 4 + (size_t)2 * D * D * 4 + (size_t)2 * WARPS * D * 4 + WARPS * 8 + 64;
  }
};

// Columns [Q*D/4, (Q+1)*D/4): register column Q of the owner lanes is eliminated, columns Q..3 are updated.
template <int D, int Q>
__device__ __forceinline__ void phase(float (&b)[RW][CPL], const float* __restrict__ Rcur, float* __restrict__ Rnext,
                                      float* __restrict__ part, int& pbuf, const int warp, const int lane) {
  constexpr int LPG = Cfg<D>::LPG;
  const int grp = lane / LPG, cl = lane - grp * LPG;
#pragma unroll 1
  for (int kk = 0; kk < LPG; ++kk) {
    const int k = Q * LPG + kk;
    const int src = grp * LPG + kk;
    float bk[RW];
#pragma unroll
    for (int i = 0; i < RW; ++i) bk[i] = __shfl_sync(0xffffffffu, b[i][Q], src);
    float s[CPL][2];
#pragma unroll
    for (int q = 0; q < CPL; ++q) s[q][0] = s[q][1] = 0.f;
#pragma unroll
    for (int i = 0; i < RW; ++i)
#pragma unroll
      for (int q = Q; q < CPL; ++q) s[q][i & 1] = fmaf(bk[i], b[i][q], s[q][i & 1]);
    float t[CPL];
#pragma unroll
    for (int q = 0; q < CPL; ++q) t[q] = s[q][0] + s[q][1];
#pragma unroll
    for (int off = LPG; off < 32; off <<= 1)
#pragma unroll
      for (int q = Q; q < CPL; ++q) t[q] += __shfl_xor_sync(0xffffffffu, t[q], off);
    float* mine = part + (size_t)(pbuf * WARPS + warp) * D;
    if (grp == 0) *reinterpret_cast<float4*>(mine + cl * 4) = make_float4(t[0], t[1], t[2], t[3]);
    __syncthreads();
    float g[CPL] = {0.f, 0.f, 0.f, 0.f}, gk = 0.f;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const float* pw = part + (size_t)(pbuf * WARPS + w) * D;
      const float4 v = *reinterpret_cast<const float4*>(pw + cl * 4);
      g[0] += v.x; g[1] += v.y; g[2] += v.z; g[3] += v.w;
      gk += pw[kk * 4 + Q];
    }
    pbuf ^= 1;
    const float rkk = Rcur[k * D + k];
    const float normsq = fmaf(rkk, rkk, gk);
    const bool valid = gk > 0.f && normsq > 1e-34f;   // below that alpha * v0 would leave the normal range
    const float alpha = -copysignf(normsq * rsqrtf(fmaxf(normsq, 1e-34f)), rkk);
    const float v0 = rkk - alpha;
    const float tau = valid ? __frcp_rn(0.5f * fmaf(v0, v0, gk)) : 0.f;   // 2 / (v'v)
    float coef[CPL];
#pragma unroll
    for (int q = Q; q < CPL; ++q) {
      const int j = cl + q * LPG;
      const float rkj = Rcur[k * D + j];
      const float w = fmaf(v0, rkj, g[q]);
      coef[q] = j > k ? tau * w : 0.f;
      if (warp == 0 && grp == 0 && j >= k) Rnext[k * D + j] = !valid ? rkj : (j == k ? alpha : rkj - tau * v0 * w);
    }
#pragma unroll
    for (int i = 0; i < RW; ++i)
#pragma unroll
      for (int q = Q; q < CPL; ++q) b[i][q] = fmaf(-bk[i], coef[q], b[i][q]);
  }
}

// X: [n, d] fp32, row pitch `ld` floats.  Every CTA reduces rows [cta*rows_per_cta, ...) and writes its triangle to
// out[cta][D][D] (row-major, zeros below the diagonal, columns >= d zero).  BULK: rows are fetched with cp.async.bulk
// (needs d == D, ld % 4 == 0, 16-byte aligned base); otherwise the strip is staged with plain loads.
template <int D, bool BULK>
__global__ void __launch_bounds__(THREADS, 2)
tsqr_kernel(const float* __restrict__ X, const long long n, const int d, const long long ld,
            const long long rows_per_cta, float* __restrict__ out) {
  using C = Cfg<D>;
  constexpr int LPG = C::LPG, STRIP = C::STRIP, TILE = C::TILE;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* stage = reinterpret_cast<float*>(smem_raw);
  float* Rm = stage + (size_t)TILE * D;            // two triangles: read one, write the other, swap per tile
  float* part = Rm + 2 * D * D;
  uint64_t* bars = reinterpret_cast<uint64_t*>(part + 2 * WARPS * D);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grp = lane / LPG, cl = lane - grp * LPG;
  const long long r_begin = (long long)blockIdx.x * rows_per_cta;
  long long r_end = r_begin + rows_per_cta;
  if (r_end > n) r_end = n;
  for (int i = tid; i < 2 * D * D; i += THREADS) Rm[i] = 0.f;
  if (BULK && lane == 0) {
    mbar_init(smem_u32(bars + warp), 1);
    fence_barrier_init();
  }
  __syncthreads();
  float* my_stage = stage + (size_t)warp * STRIP * D;
  const uint32_t my_bar = smem_u32(bars + warp);

  auto strip_rows = [&](long long row0) -> int {            // rows of this warp's strip in the tile at row0
    const long long left = r_end - (row0 + (long long)warp * STRIP);
    return left <= 0 ? 0 : (left > STRIP ? STRIP : (int)left);
  };
  auto issue = [&](long long row0) {
    const int rows = strip_rows(row0);
    if (!BULK || rows == 0) return;
    if (lane == 0) mbar_expect_tx(my_bar, (uint32_t)rows * D * 4u);
    const float* src = X + (row0 + (long long)warp * STRIP) * ld;
    for (int r = lane; r < rows; r += 32) bulk_load_1d(smem_u32(my_stage + (size_t)r * D), src + (long long)r * ld, D * 4, my_bar);
  };

  uint32_t phase_bit = 0;
  int cur = 0, pbuf = 0;
  if (r_begin < r_end) issue(r_begin);
  for (long long row0 = r_begin; row0 < r_end; row0 += TILE) {
    const int rows = strip_rows(row0);
    if (BULK) {
      if (rows > 0) {
        mbar_wait(my_bar, phase_bit);
        phase_bit ^= 1;
      }
    } else {
      const float* src = X + (row0 + (long long)warp * STRIP) * ld;
      for (int i = lane; i < rows * D; i += 32) {
        const int r = i / D, c = i - r * D;
        my_stage[i] = c < d ? src[(long long)r * ld + c] : 0.f;
      }
      __syncwarp();
    }
    float b[RW][CPL];
#pragma unroll
    for (int i = 0; i < RW; ++i) {
      const int r = grp * RW + i;
#pragma unroll
      for (int q = 0; q < CPL; ++q) b[i][q] = r < rows ? my_stage[(size_t)r * D + cl + q * LPG] : 0.f;
    }
    __syncwarp();                                          // the strip buffer is free again
    if (row0 + TILE < r_end) issue(row0 + TILE);
    const float* Rc = Rm + cur * D * D;
    float* Rn = Rm + (cur ^ 1) * D * D;
    phase<D, 0>(b, Rc, Rn, part, pbuf, warp, lane);
    phase<D, 1>(b, Rc, Rn, part, pbuf, warp, lane);
    phase<D, 2>(b, Rc, Rn, part, pbuf, warp, lane);
    phase<D, 3>(b, Rc, Rn, part, pbuf, warp, lane);
    cur ^= 1;
  }
  __syncthreads();
  const float* Rc = Rm + cur * D * D;
  float* o = out + (size_t)blockIdx.x * D * D;
  for (int i = tid; i < D * D; i += THREADS) {
    const int r = i / D, c = i - r * D;
    o[i] = c >= r ? Rc[i] : 0.f;
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
    int grid = src == X ? 2 * sm_count() : (int)((tiles + 1) / 2);
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
// Cholesky factor + solve of one F x F system by one warp, matrix in REGISTERS.
//
// The 32 lanes form a 4 x 8 grid; lane (lr, lc) holds the elements with row = 4a + lr and col = 8b + lc (cyclic in
// both directions), so at step k every lane still has live elements and the finished part of the matrix drops out as
// whole register rows / columns.  Column k and row k of the symmetric trailing matrix are fetched with one shuffle
// per register (row k sits in the lanes with lr == k % 4, column k in the lanes with lc == k % 8), the rank-one update
// runs on registers, and the right-hand side rides along as an extra row, which makes the forward substitution part of
// the factorisation.  Steps are grouped in phases of eight columns: inside a phase the register column is static and
// the two candidate register rows are picked with a select, so each phase is a rolled loop of eight steps.
// ------------------------------------------------------------------------------------------------------------
namespace chol {

template <int F, int BQ>
__device__ __forceinline__ void factor_phase(float (&acc)[F / 4][F / 8], float (&bx)[F / 8], const int lr, const int lc,
                                             bool& ok) {
  constexpr int TR = F / 4, TC = F / 8, A0 = 2 * BQ;
#pragma unroll 1
  for (int kk = 0; kk < 8; ++kk) {
    const int k = 8 * BQ + kk;
    const bool hi = kk >= 4;                       // row k sits in register row A0 + hi of the lanes lr == k % 4
    const int own = (k & 3) << 3;
    const float d = __shfl_sync(0xffffffffu, hi ? acc[A0 + 1][BQ] : acc[A0][BQ], own + kk);
    if (!(d > 0.f)) ok = false;
    const float dc = fmaxf(d, 1e-30f);
    float inv = rsqrtf(dc);
    inv = inv * fmaf(-0.5f * dc, inv * inv, 1.5f);   // one Newton step: full fp32 accuracy
    float lj[TC];
#pragma unroll
    for (int b = BQ; b < TC; ++b) lj[b] = __shfl_sync(0xffffffffu, hi ? acc[A0 + 1][b] : acc[A0][b], own + lc) * inv;
    if (lc <= kk) lj[BQ] = 0.f;                    // columns <= k are finished
    float li[TR];
#pragma unroll
    for (int a = A0; a < TR; ++a) li[a] = __shfl_sync(0xffffffffu, acc[a][BQ], (lr << 3) + kk) * inv;
    const float zk = __shfl_sync(0xffffffffu, bx[BQ], (lr << 3) + kk) * inv;
    if (lc == kk) {                                // the owner lanes keep the finished column of L
#pragma unroll
      for (int a = A0; a < TR; ++a) {
        const int row = 4 * a + lr;
        if (row >= k) acc[a][BQ] = row == k ? d * inv : li[a];
      }
      bx[BQ] = zk;
    }
    if (4 * A0 + lr <= k) li[A0] = 0.f;            // rows <= k are finished (only these two register rows can be)
    if (4 * (A0 + 1) + lr <= k) li[A0 + 1] = 0.f;
#pragma unroll
    for (int a = A0; a < TR; ++a)
#pragma unroll
      for (int b = BQ; b < TC; ++b) acc[a][b] = fmaf(-li[a], lj[b], acc[a][b]);
#pragma unroll
    for (int b = BQ; b < TC; ++b) bx[b] = fmaf(-zk, lj[b], bx[b]);
  }
}

template <int F, int BQ>
__device__ __forceinline__ void back_phase(const float (&acc)[F / 4][F / 8], float (&bx)[F / 8], const int lc) {
  constexpr int A0 = 2 * BQ;
#pragma unroll 1
  for (int kk = 7; kk >= 0; --kk) {
    const int k = 8 * BQ + kk;
    const bool hi = kk >= 4;
    const int own = (k & 3) << 3;
    const float lkk = __shfl_sync(0xffffffffu, hi ? acc[A0 + 1][BQ] : acc[A0][BQ], own + kk);
    const float zk = __shfl_sync(0xffffffffu, bx[BQ], kk);
    const float xk = zk / lkk;
    if (lc == kk) bx[BQ] = xk;
#pragma unroll
    for (int b = 0; b <= BQ; ++b) {
      float v = __shfl_sync(0xffffffffu, hi ? acc[A0 + 1][b] : acc[A0][b], own + lc);
      if (b == BQ && lc >= kk) v = 0.f;            // only columns < k
      bx[b] = fmaf(-v, xk, bx[b]);
    }
  }
}

template <int F, int BQ>
struct Phases {
  static __device__ __forceinline__ void factor(float (&acc)[F / 4][F / 8], float (&bx)[F / 8], int lr, int lc, bool& ok) {
    factor_phase<F, BQ>(acc, bx, lr, lc, ok);
    Phases<F, BQ + 1>::factor(acc, bx, lr, lc, ok);
  }
  static __device__ __forceinline__ void back(const float (&acc)[F / 4][F / 8], float (&bx)[F / 8], int lc) {
    Phases<F, BQ + 1>::back(acc, bx, lc);
    back_phase<F, BQ>(acc, bx, lc);
  }
};
template <int F>
struct Phases<F, F / 8> {
  static __device__ __forceinline__ void factor(float (&)[F / 4][F / 8], float (&)[F / 8], int, int, bool&) {}
  static __device__ __forceinline__ void back(const float (&)[F / 4][F / 8], float (&)[F / 8], int) {}
};

// acc: the symmetric matrix (both triangles), bx: the right-hand side replicated over lr.  On return the lower
// triangle of acc is L, bx is the solution (x[8b + lc] in bx[b] of every lane with that lc).
template <int F>
__device__ __forceinline__ bool factor_solve(float (&acc)[F / 4][F / 8], float (&bx)[F / 8], const int lr, const int lc,
                                             const bool want_solution) {
  bool ok = true;
  Phases<F, 0>::factor(acc, bx, lr, lc, ok);
  if (want_solution) Phases<F, 0>::back(acc, bx, lc);
  return ok;
}

constexpr int WARPS = 4;

// A: [batch][f][f] symmetric positive definite, B: [batch][f] or NULL.
// X: [batch][f] solutions; L_out (optional): [batch][f][f] lower factors; info[0] counts failed factorisations.
template <int F>
__global__ void __launch_bounds__(WARPS * 32)
potrs_batched_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ X,
                     float* __restrict__ L_out, const int f, const long long batch, int* __restrict__ info) {
  constexpr int TR = F / 4, TC = F / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int lr = lane >> 3, lc = lane & 7;
  for (long long m = (long long)blockIdx.x * WARPS + warp; m < batch; m += (long long)gridDim.x * WARPS) {
    const float* a_ = A + (size_t)m * f * f;
    float acc[TR][TC], bx[TC];
#pragma unroll
    for (int a = 0; a < TR; ++a)
#pragma unroll
      for (int b = 0; b < TC; ++b) {
        const int r = 4 * a + lr, c = 8 * b + lc;
        acc[a][b] = (r < f && c < f) ? a_[(size_t)r * f + c] : (r == c ? 1.f : 0.f);
      }
#pragma unroll
    for (int b = 0; b < TC; ++b) bx[b] = (B && 8 * b + lc < f) ? B[(size_t)m * f + 8 * b + lc] : 0.f;
    const bool ok = factor_solve<F>(acc, bx, lr, lc, X != nullptr);
    if (!ok && lane == 0) atomicAdd(info, 1);
    if (L_out) {
      float* l = L_out + (size_t)m * f * f;
#pragma unroll
      for (int a = 0; a < TR; ++a)
#pragma unroll
        for (int b = 0; b < TC; ++b) {
          const int r = 4 * a + lr, c = 8 * b + lc;
          if (r < f && c < f) l[(size_t)r * f + c] = c <= r ? acc[a][b] : 0.f;
        }
    }
    if (X && lr == 0) {
#pragma unroll
      for (int b = 0; b < TC; ++b)
        if (8 * b + lc < f) X[(size_t)m * f + 8 * b + lc] = bx[b];
    }
  }
}

}  // namespace chol

// ------------------------------------------------------------------------------------------------------------
// Implicit ALS, step 4: x_u = (Y'Y + Y_u'(C_u - I)Y_u + lambda I)^-1 Y_u' C_u p_u,  c_ui = 1 + alpha r_ui.
// One warp per row u, normal matrix in the register layout of chol:: above: one weighted outer product per rating is
// added to every lane's (F/4) x (F/8) elements.  Ratings are taken four at a time: the four factor rows are fetched
// with coalesced loads and parked in shared memory twice, permuted so that each lane reads the entries of its rows
// and of its columns with 128-bit loads.  Rows are claimed dynamically (row lengths are heavy-tailed).
// ------------------------------------------------------------------------------------------------------------
namespace als {

constexpr int WARPS = 8;
constexpr int BATCH = 4;

template <int F>
constexpr size_t warp_smem_floats() { return (size_t)2 * BATCH * F; }

template <int F>
__global__ void __launch_bounds__(WARPS * 32)
solve_kernel(const long long* __restrict__ ptr, const int* __restrict__ idx, const float* __restrict__ val,
             const float* __restrict__ Y, const float* __restrict__ YtY, const float lam, const float alpha,
             const int* __restrict__ rows, const int n_rows, const int f, float* __restrict__ Xout,
             int* __restrict__ counter, int* __restrict__ info) {
  constexpr int TR = F / 4, TC = F / 8, Q = (F + 31) / 32;
  extern __shared__ __align__(16) float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* ysr = sm + (size_t)warp * warp_smem_floats<F>();   // [BATCH][4][TR]: y[4a + lr] at lr*TR + a
  float* ysc = ysr + BATCH * F;                              // [BATCH][8][TC]: y[8b + lc] at lc*TC + b
  const int lr = lane >> 3, lc = lane & 7;
  for (;;) {
    int claim = 0;
    if (lane == 0) claim = atomicAdd(counter, 1);
    claim = __shfl_sync(0xffffffffu, claim, 0);
    if (claim >= n_rows) break;
    const int u = rows ? rows[claim] : claim;
    const long long a0 = ptr[u], e = ptr[u + 1];
    if (a0 == e) {                                  // no ratings: the factor row is zero (as the reference leaves it)
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = lane + 32 * q;
        if (c < F) Xout[(size_t)u * F + c] = 0.f;
      }
      continue;
    }
    float acc[TR][TC], bx[TC];
#pragma unroll
    for (int a = 0; a < TR; ++a)
#pragma unroll
      for (int b = 0; b < TC; ++b) {
        const int r = 4 * a + lr, c = 8 * b + lc;
        float v = YtY[r * F + c];
        if (r == c) v += r < f ? lam : 1.f;        // padded coordinates get an identity block
        acc[a][b] = v;
      }
#pragma unroll
    for (int b = 0; b < TC; ++b) bx[b] = 0.f;
    for (long long t = a0; t < e; t += BATCH) {
      float yl[BATCH][Q], w[BATCH], cw[BATCH];
#pragma unroll
      for (int s = 0; s < BATCH; ++s) {
        const bool on = t + s < e;
        const int it = on ? idx[t + s] : 0;
        w[s] = on ? alpha * val[t + s] : 0.f;
        cw[s] = on ? 1.f + w[s] : 0.f;
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
          if (c < F) {
            ysr[s * F + (c & 3) * TR + (c >> 2)] = yl[s][q];
            ysc[s * F + (c & 7) * TC + (c >> 3)] = yl[s][q];
          }
        }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < BATCH; ++s) {
        float yr[TR], yc[TC];
#pragma unroll
        for (int a = 0; a < TR; ++a) yr[a] = ysr[s * F + lr * TR + a] * w[s];
#pragma unroll
        for (int b = 0; b < TC; ++b) yc[b] = ysc[s * F + lc * TC + b];
#pragma unroll
        for (int a = 0; a < TR; ++a)
#pragma unroll
          for (int b = 0; b < TC; ++b) acc[a][b] = fmaf(yr[a], yc[b], acc[a][b]);
#pragma unroll
        for (int b = 0; b < TC; ++b) bx[b] = fmaf(cw[s], yc[b], bx[b]);
      }
    }
    if (!chol::factor_solve<F>(acc, bx, lr, lc, true) && lane == 0) atomicAdd(info, 1);
    if (lr == 0) {
#pragma unroll
      for (int b = 0; b < TC; ++b) {
        const int c = 8 * b + lc;
        Xout[(size_t)u * F + c] = c < f ? bx[b] : 0.f;
      }
    }
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
#define HB_POTRS(FF) chol::potrs_batched_kernel<FF><<<grid, chol::WARPS * 32, 0, stream>>>(A, B, X, L, f, batch, info);
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

