// Gram / cross-product kernels for the DAAL-style algorithms (covariance, PCA-correlation,
// linear & ridge regression by normal equations, low-order moments).
//
// Replaces the local step of Intel DAAL's distributed algorithms — covariance
// DistributedStep1Local (ml/daal/.../daal_cov/densedistri/COVDaalCollectiveMapper.java:143-152),
// pca (daal_pca/cordensedistr/PCADaalCollectiveMapper.java:121-127), linear_regression normEq
// (daal_linreg/normaleq/LinRegDaalCollectiveMapper.java:162-167), low_order_moments
// (daal_mom/densedistri/MOMDaalCollectiveMapper.java) — which run MKL SYRK/GEMM on Xeon.
//
// G = XᵀX for row-major X[n, d] (fp32, d % 32 == 0) on the tensor cores:
//   * operands are MN-major: the TMA box {32 features, KC rows, 8 feature-slabs} of a 3-D view
//     of X lands in shared memory exactly in the canonical MN-major tf32 layout
//     (SWIZZLE_128B_BASE32B: [slab][row][32 floats]; LBO = KC*128 B between slabs, SBO = 512 B
//     between 4-row groups; TMA side: SWIZZLE_128B_ATOM_32B),
//     so X is never transposed or copied;
//   * each CTA owns a 256x256 macro-tile of G (both 128x256 TMEM accumulators = all 512
//     columns) for one slab of rows, K-steps of 8 rows, kind::tf32, fp32 accumulate;
//   * only macro-tiles on or above the diagonal are computed (diagonal ones load one operand);
//   * split-K partial tiles are folded into G with 16-byte L2 reductions; a mirror pass fills
//     the lower triangle.
// Column moments (sum, sum of squares, min, max) come from a separate streaming kernel.
#include <stdlib.h>
#include "common.cuh"
#include "../include/harp_b200.h"

#include <math_constants.h>
#include <string.h>

namespace hb {
namespace gram {

constexpr int MT = 256;              // macro-tile edge (features)
constexpr int KC = 16;               // rows per pipeline stage
constexpr int UK = 8;                // rows per tcgen05.mma (tf32)
constexpr int NSTAGE = 6;
constexpr int SLAB = 32;             // features per 128-byte slab
constexpr int OPER_BYTES = MT * KC * 4;          // 16 KB: one operand, one stage
constexpr int STAGE_BYTES = 2 * OPER_BYTES;      // A + B
constexpr int LBO_BYTES = KC * SLAB * 4;         // 2048: next 32-feature slab
constexpr int SBO_BYTES = 4 * SLAB * 4;          // 512: next group of 4 rows (the 32B-base swizzle atom)
constexpr int KSTEP_BYTES = UK * SLAB * 4;       // 1024: 8 rows = one tcgen05.mma K step
constexpr int NSTAT = 8;                 // column-statistics warps
constexpr int NTHREADS = (6 + NSTAT) * 32;   // TMA warp, MMA warp, 4 epilogue warps, the statistics warps
constexpr int STAT_WARP0 = 6;
constexpr int TMEM_COLS = 512;

struct Smem {
  static constexpr int stages = 0;
  static constexpr int bars = NSTAGE * STAGE_BYTES;   // 196608
  static constexpr int tmem_slot = bars + 32 * 8;
  static constexpr int prog = tmem_slot + 8;         // number of pipeline stages the TMA warp has issued
  static constexpr int total = tmem_slot + 16;
};
enum : int { B_FULL = 0, B_EMPTY = NSTAGE, T_FULL = 2 * NSTAGE, T_EMPTY = 2 * NSTAGE + 1 };

// MN-major operand descriptor.  For 32-bit (tf32) MN-major operands the only legal shared-memory
// layout is SWIZZLE_128B_BASE32B (layout type 1): 128-byte rows of 32 features, atoms of 4 rows,
// 32-byte chunks XOR-swizzled by (row % 4) — what TMA produces with SWIZZLE_128B_ATOM_32B.
__device__ __forceinline__ uint64_t desc_mnmajor_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(LBO_BYTES >> 4) << 16;
  d |= static_cast<uint64_t>(SBO_BYTES >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(1) << 61;   // SWIZZLE_128B_BASE32B
  return d;
}

__device__ __forceinline__ float4 ld_nc_v4_hint(const float* p, uint64_t policy) {
  float4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(policy));
  return v;
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar,
                                            int32_t c0, int32_t c1, int32_t c2, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}

// unit -> (macro-tile row ti, col tj >= ti, row slab s)
__device__ __forceinline__ void decode_unit(long long u, int ntri, int nmt, int& ti, int& tj, long long& s) {
  s = u / ntri;
  int t = (int)(u % ntri);
  ti = 0;
  int rowlen = nmt;
  while (t >= rowlen) { t -= rowlen; ++ti; --rowlen; }
  tj = ti + t;
}

// ----------------------------------------------------------------------------
// Fused partial-result allreduce (FUSED = 1): the launch that computes this rank's XᵀX also sums the
// ranks' partial Gram matrices over NVLink peer memory, so every rank leaves the launch holding the global
// result (the reference gathers DAAL partial results on a master and broadcasts the model back,
// HD/data_comm/HarpDAALComm.java:203-246, e.g. daal_cov/densedistri/COVDaalCollectiveMapper.java:143-176).
// State block per rank, same offset in every rank's symmetric slab:  GramCtl (256 B) | G[d*d].
// Tail, run by all CTAs of the launch (one per SM, co-resident):
//   1. local completion counter (all split-K reductions into the local G are issued);
//   2. CTA 0 publishes flag1[rank] = epoch+1 on every peer; every CTA waits for all peers' flag1;
//   3. every rank sums its 1/N row-slice of the upper macro-tile triangle over the peers in rank order
//      (bit-deterministic, identical on all ranks) and stores it into every peer's G;
//   4. second counter; its last CTA publishes flag2; every CTA waits for all peers' flag2;
//   5. lower triangle <- upper triangle, locally.
// ----------------------------------------------------------------------------
struct GramCtl {
  uint32_t done1, done2, pad0[2];
  uint32_t flag1[HB_MAX_RANKS];
  uint32_t flag2[HB_MAX_RANKS];
};
static_assert(sizeof(GramCtl) <= 256, "GramCtl must fit its 256-byte slot");
struct GramFused {
  uint8_t* peer[HB_MAX_RANKS];
  int rank, world;
  uint32_t epoch;
  int pad;
};

// Column statistics produced by the same launch (all four arrays of d entries, initialised by the host wrapper):
// the statistics warps of a CTA re-read, through L2, the rows its TMA warp is streaming for a DIAGONAL macro-tile
// (each 256-column block of X is on exactly one diagonal tile per row slab), so X still leaves HBM once.
struct GramStats {
  double* sum;
  double* sumsq;
  float* mn;
  float* mx;
  int dbg;          // HB_GRAM_STATS_DBG triage switches: 1 no loads, 2 no arithmetic, 8 pace on the MMA pipeline
};

__device__ __forceinline__ void gram_spin(const uint32_t* p, uint32_t target, bool sys, const char* what) {
  const long long t0 = clock64();
  while ((int32_t)((sys ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - target) < 0) {
    if (clock64() - t0 > 20000000000LL) {
      printf("harp_b200: fused Gram tail timed out waiting for %s (cta %d thread %d)\n", what, blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

__device__ __noinline__ void gram_fused_tail(const GramFused& fa, int d) {
  const int tid = threadIdx.x;
  const int world = fa.world;
  const uint32_t target = fa.epoch + 1u;
  uint8_t* me = fa.peer[fa.rank];
  GramCtl* ctl = reinterpret_cast<GramCtl*>(me);
  float* G = reinterpret_cast<float*>(me + 256);
  // 1.
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    atomicAdd(&ctl->done1, 1u);
    gram_spin(&ctl->done1, gridDim.x, false, "local CTAs");
  }
  __syncthreads();
  const size_t gtid = blockIdx.x * (size_t)blockDim.x + tid;
  const size_t gthreads = (size_t)gridDim.x * blockDim.x;
  const int d4 = d >> 2;
  if (world > 1) {
    // 2.
    if (tid < world) {
      if (blockIdx.x == 0) {
        fence_sys();
        st_release_sys(&reinterpret_cast<GramCtl*>(fa.peer[tid])->flag1[fa.rank], target);
      }
      gram_spin(&ctl->flag1[tid], target, true, "peer partial results");
    }
    __syncthreads();
    // 3.
    const int rows_per = (d + world - 1) / world;
    const int r0 = min(d, fa.rank * rows_per), r1 = min(d, r0 + rows_per);
    const size_t total = (size_t)(r1 - r0) * d4;
    constexpr int PB = 4;
    for (size_t i = gtid; i < total; i += gthreads) {
      const int row = r0 + (int)(i / d4);
      const int c4 = (int)(i % d4);
      if (c4 * 4 + 3 < (row / MT) * MT) continue;          // below the computed macro-tile triangle
      const size_t off = 256 + ((size_t)row * d + c4 * 4) * sizeof(float);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int p0 = 0; p0 < world; p0 += PB) {
        float4 v[PB];
#pragma unroll
        for (int pp = 0; pp < PB; ++pp) {
          v[pp] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p0 + pp < world) v[pp] = ld_volatile_f4(fa.peer[p0 + pp] + off);
        }
#pragma unroll
        for (int pp = 0; pp < PB; ++pp) { acc.x += v[pp].x; acc.y += v[pp].y; acc.z += v[pp].z; acc.w += v[pp].w; }
      }
      for (int p = 0; p < world; ++p) *reinterpret_cast<float4*>(fa.peer[(fa.rank + p) % world] + off) = acc;
    }
    fence_sys();
  }
  // 4.
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const uint32_t old = atomicAdd(&ctl->done2, 1u);
    if (world > 1 && old == gridDim.x - 1) {          // last CTA of this rank: all its pushes are fenced
      fence_sys();
      for (int p = 0; p < world; ++p)
        st_release_sys(&reinterpret_cast<GramCtl*>(fa.peer[p])->flag2[fa.rank], target);
    }
    gram_spin(&ctl->done2, gridDim.x, false, "local CTAs (2)");
  }
  if (world > 1 && tid < world) gram_spin(&ctl->flag2[tid], target, true, "peer reduced slices");
  __syncthreads();
  // 5.
  const size_t total = (size_t)d * d;
  for (size_t idx = gtid; idx < total; idx += gthreads) {
    const int r = (int)(idx / d), c = (int)(idx % d);
    if (r > c) G[idx] = ld_volatile_f32(G + (size_t)c * d + r);
  }
}

template <int FUSED>
__global__ void __launch_bounds__(NTHREADS, 1)
gram_tc_kernel(const __grid_constant__ CUtensorMap tmap, long long n, int d, int nmt, int ntri,
               long long rows_per_slab, long long nslabs, float* __restrict__ G, const GramFused fa,
               const float* __restrict__ X, const GramStats stats) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0) { if (threadIdx.x == 0) printf("harp_b200: smem base not 1024-aligned\n"); __trap(); }
  const uint32_t bars = base + Smem::bars;
  auto bar = [&](int i) -> uint32_t { return bars + 8u * i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem_raw + Smem::tmem_slot);
  volatile uint32_t* prog = reinterpret_cast<volatile uint32_t*>(smem_raw + Smem::prog);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long nunits = (long long)ntri * nslabs;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar(B_FULL + s), 1); mbar_init(bar(B_EMPTY + s), 1); }
    mbar_init(bar(T_FULL), 1);
    mbar_init(bar(T_EMPTY), 4);
    *prog = 0u;
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc_1sm(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS); tmem_relinquish_1sm(); }
  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmap);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const uint64_t pol = policy_evict_last();     // a row slab is re-read by every macro-tile of its wave
      uint32_t it = 0;
      for (long long u = blockIdx.x; u < nunits; u += gridDim.x) {
        int ti, tj; long long s;
        decode_unit(u, ntri, nmt, ti, tj, s);
        const long long r0 = s * rows_per_slab;
        long long r1 = r0 + rows_per_slab; if (r1 > n) r1 = n;
        const bool diag = (ti == tj);
        for (long long r = r0; r < r1; r += KC, ++it) {
          const uint32_t st = it % NSTAGE, ph = (it / NSTAGE) & 1;
          mbar_wait(bar(B_EMPTY + st), ph ^ 1);
          mbar_expect_tx(bar(B_FULL + st), diag ? OPER_BYTES : STAGE_BYTES);
          const uint32_t dst = base + Smem::stages + st * STAGE_BYTES;
          tma_load_3d(dst, &tmap, bar(B_FULL + st), 0, (int32_t)r, ti * (MT / SLAB), pol);
          if (!diag) tma_load_3d(dst + OPER_BYTES, &tmap, bar(B_FULL + st), 0, (int32_t)r, tj * (MT / SLAB), pol);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp, one elected lane issues) =====================
    {
      const uint32_t idesc = umma_idesc_tf32(128, MT, 1, 1);   // both operands MN-major
      const uint64_t dhi = desc_mnmajor_sw128(0);
      uint32_t it = 0, gu = 0;
      for (long long u = blockIdx.x; u < nunits; u += gridDim.x, ++gu) {
        int ti, tj; long long s;
        decode_unit(u, ntri, nmt, ti, tj, s);
        const long long r0 = s * rows_per_slab;
        long long r1 = r0 + rows_per_slab; if (r1 > n) r1 = n;
        const bool diag = (ti == tj);
        mbar_wait(bar(T_EMPTY), (gu & 1) ^ 1);
        tc_fence_after();
        uint32_t first = 1;
        for (long long r = r0; r < r1; r += KC, ++it) {
          const uint32_t st = it % NSTAGE, ph = (it / NSTAGE) & 1;
          mbar_wait(bar(B_FULL + st), ph);
          if (lane == 0) *prog = it + 1u;                 // the stage has landed in L2 and smem: statistics warps may read its rows
          tc_fence_after();
          const uint32_t a_addr = base + Smem::stages + st * STAGE_BYTES;
          const uint32_t b_addr = diag ? a_addr : a_addr + OPER_BYTES;
          const uint64_t da0 = dhi | ((a_addr & 0x3FFFFu) >> 4);
          const uint64_t db0 = dhi | ((b_addr & 0x3FFFFu) >> 4);
          long long rem = r1 - r;                       // rows past n are zero-filled by TMA
          int ksteps = (int)((rem + UK - 1) / UK); if (ksteps > KC / UK) ksteps = KC / UK;
          const bool last = (r + KC >= r1);
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < KC / UK; ++j) {
              if (j < ksteps) {
#pragma unroll
                for (int mh = 0; mh < 2; ++mh)            // the two 128-feature halves of the A side
                  umma_tf32_1sm(tmem_base + mh * MT, da0 + (uint64_t)((mh * 4 * LBO_BYTES + j * KSTEP_BYTES) >> 4),
                                db0 + (uint64_t)((j * KSTEP_BYTES) >> 4), idesc, (first && j == 0) ? 0u : 1u);
              }
            }
            umma_commit_1sm(bar(B_EMPTY + st));
            if (last) umma_commit_1sm(bar(T_FULL));
          }
          __syncwarp();
          first = 0;
        }
      }
    }
  } else if (warp >= STAT_WARP0) {
    // ===================== Column statistics of the diagonal tiles' rows =====================
    // Warp sw covers 128 of the tile's 256 columns (sw & 1) and 4 of a stage's 16 rows (sw >> 1), one float4 per lane
    // and row.  The tile's rows pass through L2 about one latency (~700 cycles) after the TMA warp asked for them and
    // a stage lasts ~520 cycles; to hit L2 the warps must keep the TMA warp's pace, so three stages of loads stay in
    // flight per lane (four register buffers) and the per-stage arithmetic is fp32 (four rows), folded into the fp64
    // accumulators once per stage.
    if (stats.sum != nullptr) {
      const int sw = warp - STAT_WARP0;
      const int chalf = sw & 1, rpart = sw >> 1;
      constexpr int RPS = KC / (NSTAT / 2);
      uint32_t it = 0;
      for (long long u = blockIdx.x; u < nunits; u += gridDim.x) {
        int ti, tj; long long s;
        decode_unit(u, ntri, nmt, ti, tj, s);
        const long long r0 = s * rows_per_slab;
        long long r1 = r0 + rows_per_slab; if (r1 > n) r1 = n;
        const uint32_t nst = (uint32_t)((r1 - r0 + KC - 1) / KC);
        const int col = ti * MT + chalf * (MT / 2) + lane * 4;
        if (ti != tj || col >= d) { it += nst; continue; }
        double ds[4], dss[4];
        float lo[4], hi[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) { ds[q] = 0.0; dss[q] = 0.0; lo[q] = CUDART_INF_F; hi[q] = -CUDART_INF_F; }
        const float* base = X + (size_t)col;
        const uint64_t pol = policy_evict_last();            // same L2 priority as the TMA loads of the same lines
        float4 buf[4][RPS];
        auto fetch = [&](float4 (&b)[RPS], uint32_t k) {       // rows of stage k that this warp owns
          if (stats.dbg & 8) while (*prog <= it + k) __nanosleep(32);   // optional: only rows whose stage has landed (slower, see EXPERIMENTS.md)
          const long long rb = r0 + (long long)k * KC + rpart * RPS;
#pragma unroll
          for (int j = 0; j < RPS; ++j) {
            const long long row = rb + j < r1 ? rb + j : r1 - 1;
            if (stats.dbg & 1) b[j] = make_float4(1.f, 2.f, 3.f, (float)k); else
            b[j] = ld_nc_v4_hint(base + (size_t)row * d, pol);
          }
        };
        auto consume = [&](const float4 (&b)[RPS], uint32_t k) {
          const long long rb = r0 + (long long)k * KC + rpart * RPS;
          float fs[4] = {0.f, 0.f, 0.f, 0.f}, fss[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < RPS; ++j) {
            if (stats.dbg & 2) { lo[0] = fminf(lo[0], b[j].x + b[j].y + b[j].z + b[j].w); continue; }
            if (rb + j < r1) {
              const float e[4] = {b[j].x, b[j].y, b[j].z, b[j].w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                fs[q] += e[q]; fss[q] = fmaf(e[q], e[q], fss[q]);
                lo[q] = fminf(lo[q], e[q]); hi[q] = fmaxf(hi[q], e[q]);
              }
            }
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) { ds[q] += (double)fs[q]; dss[q] += (double)fss[q]; }
        };
        fetch(buf[0], 0);
        if (nst > 1) fetch(buf[1], 1);
        if (nst > 2) fetch(buf[2], 2);
        for (uint32_t k = 0; k < nst; k += 4) {
          if (k + 3 < nst) fetch(buf[3], k + 3);
          consume(buf[0], k);
          if (k + 1 < nst) {
            if (k + 4 < nst) fetch(buf[0], k + 4);
            consume(buf[1], k + 1);
          }
          if (k + 2 < nst) {
            if (k + 5 < nst) fetch(buf[1], k + 5);
            consume(buf[2], k + 2);
          }
          if (k + 3 < nst) {
            if (k + 6 < nst) fetch(buf[2], k + 6);
            consume(buf[3], k + 3);
          }
        }
        it += nst;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          atomicAdd(stats.sum + col + q, ds[q]);
          atomicAdd(stats.sumsq + col + q, dss[q]);
          int* imn = reinterpret_cast<int*>(stats.mn + col + q);
          int* imx = reinterpret_cast<int*>(stats.mx + col + q);
          if (lo[q] >= 0.f) atomicMin(imn, __float_as_int(lo[q])); else atomicMax(reinterpret_cast<unsigned int*>(imn), __float_as_uint(lo[q]));
          if (hi[q] >= 0.f) atomicMax(imx, __float_as_int(hi[q])); else atomicMin(reinterpret_cast<unsigned int*>(imx), __float_as_uint(hi[q]));
        }
      }
    }
  } else {
    // ===================== Epilogue: TMEM -> G (split-K reduction in L2) =====================
    const int q = warp & 3;
    uint32_t gu = 0;
    for (long long u = blockIdx.x; u < nunits; u += gridDim.x, ++gu) {
      int ti, tj; long long s;
      decode_unit(u, ntri, nmt, ti, tj, s);
      mbar_wait(bar(T_FULL), gu & 1);
      tc_fence_after();
#pragma unroll 1
      for (int mh = 0; mh < 2; ++mh) {
        const int grow = ti * MT + mh * 128 + q * 32 + lane;           // row of G owned by this thread
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + mh * MT;
#pragma unroll 1
        for (int c = 0; c < MT / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait_regs(v);
#ifdef HB_GRAM_DEBUG
          if (blockIdx.x == 0 && mh == 0 && c == 0 && q == 0 && lane < 2)
            printf("dbg tmem lane %d: %f %f %f %f\n", lane, __uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3]));
#endif
          const int gcol = tj * MT + c * 32;
          if (grow < d && gcol < d) {
            float* dst = G + (size_t)grow * d + gcol;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              red_add_v4(dst + j, make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                              __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(T_EMPTY));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_1sm(tmem_base, TMEM_COLS);
  if (FUSED) gram_fused_tail(fa, d);
}

// Only macro-tiles with tj >= ti were computed: copy the upper triangle onto the lower one.
__global__ void mirror_kernel(float* __restrict__ G, int d) {
  const size_t total = (size_t)d * d;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx / d), c = (int)(idx % d);
    if (r > c) G[idx] = G[(size_t)c * d + r];      // strict lower triangle <- upper (exactly symmetric output)
  }
}

// SIMT fallback (any d): one block per 32x32 tile of G, rows strided across the grid's z.
__global__ void gram_simt_kernel(const float* __restrict__ X, long long n, int d, float* __restrict__ G,
                                 long long rows_per_block) {
  __shared__ float As[32][33], Bs[32][33];
  const int ti = blockIdx.y, tj = blockIdx.x;
  if (tj < ti) return;
  const long long r0 = blockIdx.z * rows_per_block;
  long long r1 = r0 + rows_per_block; if (r1 > n) r1 = n;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8 threads
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (long long r = r0; r < r1; r += 32) {
    for (int k = ty; k < 32; k += 8) {
      const long long row = r + k;
      const int ca = ti * 32 + tx, cb = tj * 32 + tx;
      As[k][tx] = (row < r1 && ca < d) ? X[(size_t)row * d + ca] : 0.f;
      Bs[k][tx] = (row < r1 && cb < d) ? X[(size_t)row * d + cb] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const float b = Bs[k][tx];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(As[k][ty + 8 * i], b, acc[i]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gr = ti * 32 + ty + 8 * i, gc = tj * 32 + tx;
    if (gr < d && gc < d) atomicAdd(G + (size_t)gr * d + gc, acc[i]);
  }
}
__global__ void mirror32_kernel(float* __restrict__ G, int d) {
  const size_t total = (size_t)d * d;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx / d), c = (int)(idx % d);
    if (r / 32 > c / 32) G[idx] = G[(size_t)c * d + r];
  }
}

// Column moments: sum, sum of squares, min, max over rows (DAAL low_order_moments partials).
// out = [4][d] doubles for sum / sumsq, floats reinterpret for min/max kept in separate arrays.
__global__ void moments_kernel(const float* __restrict__ X, long long n, int d, double* __restrict__ sum,
                               double* __restrict__ sumsq, float* __restrict__ mn, float* __restrict__ mx) {
  // block = 256 threads: 32 column-lanes x 8 row-lanes; grid.x tiles columns, grid.y strides rows
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cl;
  double s = 0.0, ss = 0.0;
  float lo = CUDART_INF_F, hi = -CUDART_INF_F;
  if (col < d) {
    for (long long r = blockIdx.y * 8LL + rl; r < n; r += (long long)gridDim.y * 8) {
      const float v = X[(size_t)r * d + col];
      s += v; ss += (double)v * v;
      lo = fminf(lo, v); hi = fmaxf(hi, v);
    }
  }
  __shared__ double sh_s[8][32], sh_ss[8][32];
  __shared__ float sh_lo[8][32], sh_hi[8][32];
  sh_s[rl][cl] = s; sh_ss[rl][cl] = ss; sh_lo[rl][cl] = lo; sh_hi[rl][cl] = hi;
  __syncthreads();
  if (rl == 0 && col < d) {
    for (int i = 1; i < 8; ++i) {
      s += sh_s[i][cl]; ss += sh_ss[i][cl];
      lo = fminf(lo, sh_lo[i][cl]); hi = fmaxf(hi, sh_hi[i][cl]);
    }
    atomicAdd(sum + col, s);
    atomicAdd(sumsq + col, ss);
    // float atomic min/max through the ordered-int trick
    int* imn = reinterpret_cast<int*>(mn + col);
    int* imx = reinterpret_cast<int*>(mx + col);
    if (lo >= 0.f) atomicMin(imn, __float_as_int(lo)); else atomicMax(reinterpret_cast<unsigned int*>(imn), __float_as_uint(lo));
    if (hi >= 0.f) atomicMax(imx, __float_as_int(hi)); else atomicMin(reinterpret_cast<unsigned int*>(imx), __float_as_uint(hi));
  }
}
// Same reduction for d % 4 == 0: a lane owns 4 adjacent columns (one 16-byte load per row), a warp a 128-column
// stripe, four rows are in flight per lane; ~2x the bandwidth of the scalar kernel on wide matrices.
__global__ void __launch_bounds__(256) moments_vec4_kernel(const float* __restrict__ X, long long n, int d,
                                                           double* __restrict__ sum, double* __restrict__ sumsq,
                                                           float* __restrict__ mn, float* __restrict__ mx) {
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cl) * 4;
  double ds[4] = {0, 0, 0, 0}, dss[4] = {0, 0, 0, 0};
  float lo[4], hi[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { lo[q] = CUDART_INF_F; hi[q] = -CUDART_INF_F; }
  if (col < d) {
    const long long stride = (long long)gridDim.y * 8;
    long long r = blockIdx.y * 8LL + rl;
    for (; r + 3 * stride < n; r += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ld_nc_v4(X + (size_t)(r + u * stride) * d + col);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          ds[q] += (double)e[q]; dss[q] = fma((double)e[q], (double)e[q], dss[q]);
          lo[q] = fminf(lo[q], e[q]); hi[q] = fmaxf(hi[q], e[q]);
        }
      }
    }
    for (; r < n; r += stride) {
      const float4 v = ld_nc_v4(X + (size_t)r * d + col);
      const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ds[q] += (double)e[q]; dss[q] = fma((double)e[q], (double)e[q], dss[q]);
        lo[q] = fminf(lo[q], e[q]); hi[q] = fmaxf(hi[q], e[q]);
      }
    }
  }
  __shared__ double sh_s[8][32][4], sh_ss[8][32][4];
  __shared__ float sh_lo[8][32][4], sh_hi[8][32][4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { sh_s[rl][cl][q] = ds[q]; sh_ss[rl][cl][q] = dss[q]; sh_lo[rl][cl][q] = lo[q]; sh_hi[rl][cl][q] = hi[q]; }
  __syncthreads();
  if (rl < 4 && col < d) {       // warp q of the first four finishes column col + q
    const int q = rl;
    double a = 0, b = 0;
    float l = CUDART_INF_F, h = -CUDART_INF_F;
    for (int i = 0; i < 8; ++i) {
      a += sh_s[i][cl][q]; b += sh_ss[i][cl][q];
      l = fminf(l, sh_lo[i][cl][q]); h = fmaxf(h, sh_hi[i][cl][q]);
    }
    atomicAdd(sum + col + q, a);
    atomicAdd(sumsq + col + q, b);
    int* imn = reinterpret_cast<int*>(mn + col + q);
    int* imx = reinterpret_cast<int*>(mx + col + q);
    if (l >= 0.f) atomicMin(imn, __float_as_int(l)); else atomicMax(reinterpret_cast<unsigned int*>(imn), __float_as_uint(l));
    if (h >= 0.f) atomicMax(imx, __float_as_int(h)); else atomicMin(reinterpret_cast<unsigned int*>(imx), __float_as_uint(h));
  }
}
__global__ void moments_init_kernel(int d, double* sum, double* sumsq, float* mn, float* mx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d) { sum[i] = 0.0; sumsq[i] = 0.0; mn[i] = CUDART_INF_F; mx[i] = -CUDART_INF_F; }
}

}  // namespace gram
}  // namespace hb

namespace hb {
namespace gram {

// K14: [Σx ; Yᵀ X] in ONE streaming pass over X (DAAL linear_regression / ridge partials: Xᵀy and the column sums,
// ml/daal/.../daal_linreg/normaleq/LinRegDaalCollectiveMapper.java:162-167).  Thread t of a CTA owns four
// consecutive columns; a CTA walks a slab of rows with coalesced 16-byte loads, the row's few y values come in
// as one broadcast load; accumulation is fp64 per thread (the pass is HBM bound: 8 B of X per 2 (m + 1) flops).
template <int M>
__global__ void __launch_bounds__(256) xty_kernel(const float* __restrict__ X, long long n, int d,
                                                  const float* __restrict__ Y, double* __restrict__ out) {
  const int c4 = blockIdx.x * blockDim.x + threadIdx.x;          // float4 column group
  if (c4 * 4 >= d) return;
  const long long rows_per = (n + gridDim.y - 1) / gridDim.y;
  const long long r0 = blockIdx.y * rows_per, r1 = min(n, r0 + rows_per);
  double acc[M + 1][4];
#pragma unroll
  for (int j = 0; j <= M; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0;
  for (long long r = r0; r < r1; ++r) {
    const float4 x = ld_nc_v4(X + (size_t)r * d + c4 * 4);
    acc[0][0] += x.x; acc[0][1] += x.y; acc[0][2] += x.z; acc[0][3] += x.w;
#pragma unroll
    for (int j = 0; j < M; ++j) {
      const double y = (double)__ldg(Y + (size_t)r * M + j);
      acc[j + 1][0] += y * x.x; acc[j + 1][1] += y * x.y; acc[j + 1][2] += y * x.z; acc[j + 1][3] += y * x.w;
    }
  }
#pragma unroll
  for (int j = 0; j <= M; ++j)
#pragma unroll
    for (int e = 0; e < 4; ++e) red_add_f64(out + (size_t)j * d + c4 * 4 + e, acc[j][e]);
}

// rows of X minus a per-column shift (fp32): the centring step of the two-pass covariance for data whose mean is
// large against its spread, where TF32 products of the raw values would keep almost nothing of the spread.
__global__ void center_rows_kernel(const float* __restrict__ X, long long n, int d, const float* __restrict__ shift,
                                   float* __restrict__ out) {
  const int d4 = d >> 2;
  const long long total = n * d4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % d4);
    const float4 x = ld_nc_v4(X + i * 4);
    const float4 m = __ldg(reinterpret_cast<const float4*>(shift) + c4);
    // rounded to the nearest TF32 value here: the tensor core would otherwise TRUNCATE the low 13 bits, a one-sided error
    // that biases every sum of squares by ~ -5e-4 relative; round-to-nearest errors are unbiased and average out
    auto rna = [](float v) { uint32_t u; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v)); return __uint_as_float(u); };
    *reinterpret_cast<float4*>(out + i * 4) = make_float4(rna(x.x - m.x), rna(x.y - m.y), rna(x.z - m.z), rna(x.w - m.w));
  }
}

}  // namespace gram
}  // namespace hb

using namespace hb;
using namespace hb::gram;

struct GramPlan { CUtensorMap tm; int nmt, ntri, grid; long long rows_per_slab, nslabs; };

// Tensor map + work decomposition of one XᵀX pass (tensor-core path).
static int gram_plan(GramPlan& P, const float* X, long long n, int d) {
  // 3-D view of X: dim0 = 32 floats inside a slab, dim1 = rows, dim2 = slab index
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    HB_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    HB_REQUIRE(q == cudaDriverEntryPointSuccess && p != nullptr, "cuTensorMapEncodeTiled unavailable");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  const long long nn = n > 0 ? n : 1;
  cuuint64_t gdim[3] = {(cuuint64_t)SLAB, (cuuint64_t)nn, (cuuint64_t)(d / SLAB)};
  cuuint64_t gstr[2] = {(cuuint64_t)d * 4, (cuuint64_t)SLAB * 4};
  cuuint32_t box[3] = {SLAB, KC, MT / SLAB};
  cuuint32_t est[3] = {1, 1, 1};
  CUresult r = fn(&P.tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(X), gdim, gstr, box, est,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HB_REQUIRE(r == CUDA_SUCCESS, "hb_gram: cuTensorMapEncodeTiled failed (%d)", (int)r);
  P.nmt = (d + MT - 1) / MT;
  P.ntri = P.nmt * (P.nmt + 1) / 2;
  const int sms = sm_count();
  // row slabs: enough units to balance the grid, each a multiple of KC rows
  long long nslabs = (5LL * sms + P.ntri - 1) / P.ntri;
  long long rows_per_slab = ((nn + nslabs - 1) / nslabs + KC - 1) / KC * KC;
  if (rows_per_slab < 4 * KC) rows_per_slab = 4 * KC;
  nslabs = n > 0 ? (n + rows_per_slab - 1) / rows_per_slab : 0;
  P.rows_per_slab = rows_per_slab;
  P.nslabs = nslabs;
  const long long nunits = nslabs * P.ntri;
  P.grid = sms;
  if ((long long)P.grid > nunits) P.grid = (int)nunits;
  if (P.grid < 1) P.grid = 1;
  return 0;
}

extern "C" {

int hb_gram_tc_supported(int d) { return (d >= 32 && d % 32 == 0 && d <= 16384) ? 1 : 0; }

// G[d,d] (+)= XᵀX.  flags: HB_GRAM_NO_ZERO accumulates, HB_GRAM_FORCE_SIMT uses the fp32 SIMT kernel.
int hb_gram(const float* X, long long n, int d, float* G, int flags, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(n >= 0 && d >= 1, "hb_gram: bad shape n=%lld d=%d", n, d);
  if ((flags & HB_GRAM_NO_ZERO) == 0) HB_CUDA_OK(cudaMemsetAsync(G, 0, (size_t)d * d * sizeof(float), stream));
  if (n == 0) return 0;
  const bool tc = (flags & HB_GRAM_FORCE_SIMT) == 0 && hb_gram_tc_supported(d) &&
                  (reinterpret_cast<uintptr_t>(X) & 15) == 0 && (reinterpret_cast<uintptr_t>(G) & 15) == 0 &&
                  n < (1LL << 31);
  if (tc) {
    GramPlan P;
    if (gram_plan(P, X, n, d) != 0) return -1;
    const int sms = sm_count();
    HB_CUDA_OK(cudaFuncSetAttribute(gram_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::total));
    GramFused nofuse;
    memset(&nofuse, 0, sizeof(nofuse));
    GramStats nostats;
    memset(&nostats, 0, sizeof(nostats));
    gram_tc_kernel<0><<<P.grid, NTHREADS, Smem::total, stream>>>(P.tm, n, d, P.nmt, P.ntri, P.rows_per_slab, P.nslabs, G, nofuse, X, nostats);
    HB_CUDA_OK(cudaGetLastError());
    mirror_kernel<<<sms * 2, 256, 0, stream>>>(G, d);
    HB_CUDA_OK(cudaGetLastError());
  } else {
    const int nt = (d + 31) / 32;
    long long zb = (sm_count() * 4 + (long long)nt * nt - 1) / ((long long)nt * nt);
    if (zb < 1) zb = 1;
    if (zb > 65535) zb = 65535;
    long long rows_per_block = ((n + zb - 1) / zb + 31) / 32 * 32;
    zb = (n + rows_per_block - 1) / rows_per_block;
    dim3 grid(nt, nt, (unsigned)zb);
    gram_simt_kernel<<<grid, 256, 0, stream>>>(X, n, d, G, rows_per_block);
    HB_CUDA_OK(cudaGetLastError());
    mirror32_kernel<<<sm_count() * 2, 256, 0, stream>>>(G, d);
    HB_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

size_t hb_gram_fused_state_bytes(int d) { return 256 + (size_t)d * d * sizeof(float); }

// XᵀX of this rank's rows + allreduce over the ranks in ONE launch; the result is left in every rank's state
// block at byte offset 256 (row-major d x d).  `epoch` = 0, 1, 2, ... per state block, identical on all ranks;
// the block must be zero-filled once before epoch 0 (and all ranks synchronised) so the flags start at 0.
int hb_gram_allreduce_fused_stats(const float* X, long long n, int d, void* const* peer_state, int rank, int world,
                                  unsigned epoch, double* sum, double* sumsq, float* mn, float* mx, void* stream_);
int hb_gram_allreduce_fused(const float* X, long long n, int d, void* const* peer_state, int rank, int world,
                            unsigned epoch, void* stream_) {
  return hb_gram_allreduce_fused_stats(X, n, d, peer_state, rank, world, epoch, nullptr, nullptr, nullptr, nullptr, stream_);
}

// Same launch, and the column statistics of THIS rank's rows (sum, sum of squares in fp64, min, max; d entries each,
// (re)initialised here) come out of it too: X is streamed from HBM once for Gram + moments.  sum == NULL skips them.
int hb_gram_allreduce_fused_stats(const float* X, long long n, int d, void* const* peer_state, int rank, int world,
                                  unsigned epoch, double* sum, double* sumsq, float* mn, float* mx, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(sum == nullptr || (sumsq != nullptr && mn != nullptr && mx != nullptr), "hb_gram_allreduce_fused_stats: all four statistics arrays or none");
  if (sum != nullptr) {
    moments_init_kernel<<<(d + 255) / 256, 256, 0, stream>>>(d, sum, sumsq, mn, mx);
    HB_CUDA_OK(cudaGetLastError());
  }
  GramStats stats;
  stats.sum = sum; stats.sumsq = sumsq; stats.mn = mn; stats.mx = mx;
  stats.dbg = getenv("HB_GRAM_STATS_DBG") ? atoi(getenv("HB_GRAM_STATS_DBG")) : 0;
  HB_REQUIRE(world >= 1 && world <= HB_MAX_RANKS && rank >= 0 && rank < world, "hb_gram_allreduce_fused: bad rank %d / world %d", rank, world);
  HB_REQUIRE(n >= 0 && n < (1LL << 31) && hb_gram_tc_supported(d), "hb_gram_allreduce_fused: shape (n=%lld, d=%d) not supported", n, d);
  GramFused fa;
  memset(&fa, 0, sizeof(fa));
  for (int i = 0; i < world; ++i) {
    HB_REQUIRE(peer_state[i] != nullptr && (reinterpret_cast<uintptr_t>(peer_state[i]) & 255) == 0,
               "hb_gram_allreduce_fused: state pointer of rank %d is null or not 256-byte aligned", i);
    fa.peer[i] = reinterpret_cast<uint8_t*>(peer_state[i]);
  }
  fa.rank = rank; fa.world = world; fa.epoch = epoch;
  uint8_t* me = fa.peer[rank];
  float* G = reinterpret_cast<float*>(me + 256);
  HB_REQUIRE(n == 0 || (X != nullptr && (reinterpret_cast<uintptr_t>(X) & 15) == 0), "hb_gram_allreduce_fused: X must be 16-byte aligned");
  // counters of this launch + the accumulator; peers stopped touching this block when our previous launch ended
  HB_CUDA_OK(cudaMemsetAsync(me, 0, 16, stream));
  HB_CUDA_OK(cudaMemsetAsync(G, 0, (size_t)d * d * sizeof(float), stream));
  GramPlan P;
  if (gram_plan(P, n > 0 ? X : G, n, d) != 0) return -1;
  HB_CUDA_OK(cudaFuncSetAttribute(gram_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::total));
  // the tail spins on counters fed by the other CTAs of the launch: one CTA per SM, cooperative launch
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)P.grid);
  cfg.blockDim = dim3(NTHREADS);
  cfg.dynamicSmemBytes = Smem::total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  HB_CUDA_OK(cudaLaunchKernelEx(&cfg, gram_tc_kernel<1>, P.tm, n, d, P.nmt, P.ntri, P.rows_per_slab, P.nslabs, G, fa, X, stats));
  return 0;
}

// Column sum / sum of squares (double) and min / max (float) over the rows of X[n,d].
int hb_moments(const float* X, long long n, int d, double* sum, double* sumsq, float* mn, float* mx,
               int flags, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(n >= 0 && d >= 1, "hb_moments: bad shape");
  if ((flags & HB_GRAM_NO_ZERO) == 0) {
    moments_init_kernel<<<(d + 255) / 256, 256, 0, stream>>>(d, sum, sumsq, mn, mx);
    HB_CUDA_OK(cudaGetLastError());
  }
  if (n == 0) return 0;
  const bool vec4 = (d % 4 == 0) && (reinterpret_cast<uintptr_t>(X) & 15) == 0;
  const int gx = vec4 ? (d / 4 + 31) / 32 : (d + 31) / 32;
  long long gy = (sm_count() * 8 + gx - 1) / gx;
  if (gy > (n + 7) / 8) gy = (n + 7) / 8;
  if (gy < 1) gy = 1;
  if (gy > 65535) gy = 65535;
  if (vec4) moments_vec4_kernel<<<dim3(gx, (unsigned)gy), 256, 0, stream>>>(X, n, d, sum, sumsq, mn, mx);
  else moments_kernel<<<dim3(gx, (unsigned)gy), 256, 0, stream>>>(X, n, d, sum, sumsq, mn, mx);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// out[(m + 1), d] (double, accumulated into: zero it first) : row 0 += column sums of X, row 1 + j += Σ_r Y[r, j] X[r, :].
// X [n, d] and Y [n, m] row-major fp32, d % 4 == 0, 0 <= m <= 8.
int hb_xty(const float* X, long long n, int d, const float* Y, int m, double* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(n >= 0 && d >= 4 && d % 4 == 0 && m >= 0 && m <= 8, "hb_xty: need d %% 4 == 0 and at most 8 dependent variables (d=%d, m=%d)", d, m);
  HB_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "hb_xty: X must be 16-byte aligned");
  if (n == 0) return 0;
  const int gx = (d / 4 + 255) / 256;
  long long gy = (long long)sm_count() * 8 / gx;
  if (gy > n) gy = n;
  if (gy < 1) gy = 1;
  const dim3 grid((unsigned)gx, (unsigned)gy);
  using namespace hb::gram;
  switch (m) {
    case 0: xty_kernel<0><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 1: xty_kernel<1><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 2: xty_kernel<2><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 3: xty_kernel<3><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 4: xty_kernel<4><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 5: xty_kernel<5><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 6: xty_kernel<6><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    case 7: xty_kernel<7><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
    default: xty_kernel<8><<<grid, 256, 0, stream>>>(X, n, d, Y, out); break;
  }
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// out[r, :] = X[r, :] - shift[:]   (fp32, d % 4 == 0)
int hb_center_rows(const float* X, long long n, int d, const float* shift, float* out, void* stream_) {
  HB_REQUIRE(n >= 0 && d >= 4 && d % 4 == 0, "hb_center_rows: need d %% 4 == 0 (d=%d)", d);
  HB_REQUIRE(((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(shift) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
             "hb_center_rows: pointers must be 16-byte aligned");
  if (n == 0) return 0;
  long long blocks = (n * (d / 4) + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  hb::gram::center_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream_>>>(X, n, d, shift, out);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
