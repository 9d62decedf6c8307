// K-means assignment step for B200 (sm_100a).
//
// Replaces the reference's scalar Java hot loop (CenCalcTask.run,
// ml/java/.../kmeans/regroupallgather/CenCalcTask.java:67-100, and the contrib
// variant KmeansMapper.java:173-215) and DAAL's kmeans.DistributedStep1Local.
//
// Design (B200-first, not a translation).  Three kernels share the building blocks below:
//   assign_tc_pair_kernel<DBG, EXT, FUSED>   the production path: a cluster of two CTAs drives one
//       tcgen05.mma.cta_group::2 kind::tf32 pipeline (M256 x N256 x K8) over fp32 tiles that TMA puts in
//       shared memory (SWIZZLE_128B); the 256x256 fp32 accumulator is double-buffered in TMEM.  Warps 0..7
//       of each CTA are the epilogue (tcgen05.ld, packed arg-max, fold into sums/counts with
//       red.global.add.v4.f32 that resolve in L2), warp 8 the TMA producer, warp 9 (leader CTA) the MMA
//       issuer.  EXT = 1 folds -|c|^2/2 into the GEMM as one extra K = 8 step; FUSED = 1 appends the
//       regroup -> centroid update -> allgather over NVLink peer memory (fused_tail), i.e. a whole Lloyd
//       iteration is ONE launch per GPU.
//   assign_tc_kernel                         the first, single-CTA form (cta_group::1, 4 epilogue warps,
//       |c|^2 added in the epilogue); kept as fallback for tiny inputs and for hb_kmeans_step_static.
//   assign_simt_kernel                       plain fp32 SIMT kernel, the exact-arithmetic fallback for shapes
//       the tensor path does not take (d % 4 != 0 or d > 128).
// In all of them the point tile stays resident in smem while every centroid tile streams past it out of
// L2, and nothing but the final [k x d] sums ever goes back to HBM.
#include "common.cuh"
#include "../include/harp_b200.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <math_constants.h>

#include <mutex>
#include <stdlib.h>

namespace hb {
namespace kmeans {

constexpr int BM = 128;          // points per tile (UMMA M)
constexpr int BN = 256;          // centroids per tile (UMMA N)
constexpr int BK = 32;           // fp32 elements per k-chunk = one 128-byte swizzle row
constexpr int UK = 8;            // K per tcgen05.mma for tf32
constexpr int MAX_KCH = 4;       // d <= 128
constexpr int NSTAGE = 4;        // centroid smem ring depth
constexpr int X_CHUNK_BYTES = BM * BK * 4;   // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 4;   // 32 KB
constexpr int CN_BYTES = BN * 4;             // 1 KB
constexpr int NTHREADS = 192;                // 6 warps
constexpr int TMEM_COLS = 512;

struct SmemLayout {
  // offsets relative to the 1024-aligned base
  static constexpr int x = 0;
  static constexpr int b = x + MAX_KCH * X_CHUNK_BYTES;         // 65536
  static constexpr int cn = b + NSTAGE * B_STAGE_BYTES;         // +131072
  static constexpr int bars = cn + 2 * CN_BYTES;                // +2048
  static constexpr int tmem_slot = bars + 32 * 8;
  static constexpr int total = tmem_slot + 16;
};
constexpr int SMEM_BYTES = SmemLayout::total + 1024;  // + alignment slack

// barrier indices
enum : int {
  BAR_X_FULL = 0,
  BAR_X_EMPTY = 1,
  BAR_B_FULL = 2,                     // NSTAGE
  BAR_B_EMPTY = BAR_B_FULL + NSTAGE,  // NSTAGE
  BAR_T_FULL = BAR_B_EMPTY + NSTAGE,  // 2
  BAR_T_EMPTY = BAR_T_FULL + 2,       // 2
  BAR_CN_FULL = BAR_T_EMPTY + 2,      // 2
  BAR_COUNT = BAR_CN_FULL + 2
};

// ----------------------------------------------------------------------------
// prep: ‖c‖² (padded with +inf to a multiple of BN), zero the accumulators
// ----------------------------------------------------------------------------
// tf32 keeps 10 explicit mantissa bits: clearing the low 13 makes the value exactly representable, so the
// tensor core's own truncation of it is the identity
__device__ __forceinline__ float tf32_trunc(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

__global__ void prep_kernel(const float* __restrict__ C, int k, int d, int kpad,
                            float* __restrict__ cn, float* __restrict__ cn_ext, float* __restrict__ ones_ext,
                            float* __restrict__ sums,
                            float* __restrict__ counts, double* __restrict__ obj, int zero_acc) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int j = warp; j < kpad; j += nwarps) {
    float s = CUDART_INF_F;
    if (j < k) {
      s = 0.f;
      if ((d & 3) == 0) {     // same element order as the fused kernel's tail: both produce identical bits
        const float4* row4 = reinterpret_cast<const float4*>(C + (size_t)j * d);
        for (int c4 = lane; c4 < (d >> 2); c4 += 32) {
          const float4 v = row4[c4];
          s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        }
      } else {
        for (int t = lane; t < d; t += 32) {
          float v = C[(size_t)j * d + t];
          s = fmaf(v, v, s);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    if (lane == 0 && cn != nullptr) cn[j] = s;
    if (cn_ext != nullptr && lane < 8) {
      // row j of the extra GEMM operand: -‖c‖²/2 as an exact sum of three tf32 numbers (hi, mid, lo);
      // with the matching A row [1,1,1,0,...] one more K=8 MMA step turns the accumulator into
      // x·c - ‖c‖²/2, whose arg-MAX is the nearest centroid.  Padding rows get a huge negative bias.
      float v = 0.f;
      if (j >= k) { v = (lane == 0) ? -3.0e38f : 0.f; v = tf32_trunc(v); }
      else {
        const float t0 = -0.5f * s;
        const float hi = tf32_trunc(t0);
        const float mid = tf32_trunc(t0 - hi);
        const float lo = tf32_trunc(t0 - hi - mid);
        v = lane == 0 ? hi : (lane == 1 ? mid : (lane == 2 ? lo : 0.f));
      }
      cn_ext[(size_t)j * 8 + lane] = v;
    }
  }
  if (ones_ext != nullptr) {
    const int tid0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid0 < 128 * 8) ones_ext[tid0] = ((tid0 & 7) < 3) ? 1.0f : 0.0f;
  }
  if (zero_acc) {
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t nth = (size_t)gridDim.x * blockDim.x;
    const size_t total = (size_t)k * d;
    for (size_t i = tid; i < total; i += nth) sums[i] = 0.f;
    for (size_t i = tid; i < (size_t)k; i += nth) counts[i] = 0.f;
    if (tid == 0 && obj != nullptr) *obj = 0.0;
  }
}

// ----------------------------------------------------------------------------
// Split-precision operands (PREC = 1 kernels).  tcgen05 has no fp32-input MMA; TF32 keeps 10 mantissa bits of
// each input, so near-ties between centroids can resolve differently from fp32 arithmetic.  Here every fp32 value
// v (times a power-of-two `scale` in fp16 mode) is stored as TWO 16-bit numbers hi = rn16(v), lo = rn16(v - hi)
// and the distance GEMM is evaluated as hi·hi + lo·hi + hi·lo with kind::f16 MMAs (fp32 accumulate): with fp16
// halves (11 + 11 mantissa bits, values scaled into fp16's range) the dropped lo·lo term is ~2^-22 relative, i.e.
// fp32-faithful products at 1.5x the tensor time of TF32 (kind::f16 runs at twice the kind::tf32 rate); bf16
// halves (8 + 8 bits, no scaling needed) give ~2^-16.
// Row layout: [hi_0 .. hi_{dp-1} | lo_0 .. lo_{dp-1}] 16-bit, dp = d rounded up to 64 — for d = 128 a row is 512
// bytes, the same as the fp32 row, and the hi / lo halves are 128-byte K-chunks for TMA + SWIZZLE_128B.
// ----------------------------------------------------------------------------
__host__ __device__ constexpr int split_dp(int d) { return (d + 63) / 64 * 64; }

__device__ __forceinline__ unsigned short cvt16(float v, int fmt16) {
  if (fmt16 == 0) return __half_as_ushort(__float2half_rn(v));
  return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float back16(unsigned short h, int fmt16) {
  if (fmt16 == 0) return __half2float(__ushort_as_half(h));
  return __bfloat162float(__ushort_as_bfloat16(h));
}
__device__ __forceinline__ void split16(float v, int fmt16, unsigned short& hi, unsigned short& lo) {
  hi = cvt16(v, fmt16);
  lo = cvt16(v - back16(hi, fmt16), fmt16);
}

// rows of X (or C) -> split rows; one thread per 4 consecutive elements
__global__ void split_rows_kernel(const float* __restrict__ X, long long n, int d, float scale, int fmt16,
                                  unsigned short* __restrict__ Xs) {
  const int dp = split_dp(d), q4 = dp >> 2;
  const long long total = n * q4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / q4;
    const int c = (int)(i % q4) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < d) v = *reinterpret_cast<const float4*>(X + (size_t)row * d + c);      // d % 4 == 0
    unsigned short h[4], l[4];
    split16(v.x * scale, fmt16, h[0], l[0]); split16(v.y * scale, fmt16, h[1], l[1]);
    split16(v.z * scale, fmt16, h[2], l[2]); split16(v.w * scale, fmt16, h[3], l[3]);
    unsigned short* dst = Xs + (size_t)row * 2 * dp + c;
    *reinterpret_cast<uint2*>(dst) = make_uint2(h[0] | ((uint32_t)h[1] << 16), h[2] | ((uint32_t)h[3] << 16));
    *reinterpret_cast<uint2*>(dst + dp) = make_uint2(l[0] | ((uint32_t)l[1] << 16), l[2] | ((uint32_t)l[3] << 16));
  }
}

// -(scale^2)‖c‖²/2 as three bf16 terms in a 32-byte operand row (one K = 16 step of kind::f16 with bf16 inputs)
__device__ __forceinline__ void bias_row_bf16(float t0, uint32_t (&w)[2]) {
  const __nv_bfloat16 hi = __float2bfloat16_rn(t0);
  const float r1 = t0 - __bfloat162float(hi);
  const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
  const __nv_bfloat16 lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
  w[0] = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(mid) << 16);
  w[1] = (uint32_t)__bfloat16_as_ushort(lo);
}

// split-mode counterpart of prep_kernel's operand part: Cs rows, bias rows (kpad x 8 words), the ones tile
__global__ void prep_split_kernel(const float* __restrict__ C, int k, int d, int kpad, float scale, int fmt16,
                                  unsigned short* __restrict__ Cs, uint32_t* __restrict__ ext, uint32_t* __restrict__ ones) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int dp = split_dp(d);
  for (int j = warp; j < kpad; j += nwarps) {
    float s = 0.f;
    if (j < k) {
      const float4* row4 = reinterpret_cast<const float4*>(C + (size_t)j * d);
      for (int c4 = lane; c4 < (dp >> 2); c4 += 32) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c4 < (d >> 2)) v = row4[c4];
        s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        unsigned short h[4], l[4];
        split16(v.x * scale, fmt16, h[0], l[0]); split16(v.y * scale, fmt16, h[1], l[1]);
        split16(v.z * scale, fmt16, h[2], l[2]); split16(v.w * scale, fmt16, h[3], l[3]);
        unsigned short* dst = Cs + (size_t)j * 2 * dp + c4 * 4;
        *reinterpret_cast<uint2*>(dst) = make_uint2(h[0] | ((uint32_t)h[1] << 16), h[2] | ((uint32_t)h[3] << 16));
        *reinterpret_cast<uint2*>(dst + dp) = make_uint2(l[0] | ((uint32_t)l[1] << 16), l[2] | ((uint32_t)l[3] << 16));
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    if (lane < 8) {
      uint32_t w[2];
      bias_row_bf16(j < k ? -0.5f * s * scale * scale : -3.0e38f, w);
      ext[(size_t)j * 8 + lane] = lane < 2 ? w[lane] : 0u;
    }
  }
  const int tid0 = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid0 < 128 * 8) ones[tid0] = (tid0 & 7) == 0 ? 0x3F803F80u : ((tid0 & 7) == 1 ? 0x00003F80u : 0u);   // bf16 [1,1,1,0..]
}

// Running arg-min over 32 accumulator columns held in registers.  d = ‖c‖² - 2·x·c.
__device__ __forceinline__ void argmin_chunk(const uint32_t (&v)[32], const float4* __restrict__ cn4,
                                             int cbase, float& best, int& bi) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 cnv = cn4[j >> 2];
    const float d0 = fmaf(-2.f, __uint_as_float(v[j + 0]), cnv.x);
    const float d1 = fmaf(-2.f, __uint_as_float(v[j + 1]), cnv.y);
    const float d2 = fmaf(-2.f, __uint_as_float(v[j + 2]), cnv.z);
    const float d3 = fmaf(-2.f, __uint_as_float(v[j + 3]), cnv.w);
    const float m = fminf(fminf(d0, d1), fminf(d2, d3));
    if (m < best) {
      best = m;
      // lowest index wins ties inside the group
      bi = cbase + j + (d0 == m ? 0 : (d1 == m ? 1 : (d2 == m ? 2 : 3)));
    }
  }
}

// Branch-free arg-min over 32 accumulator columns: the column's index inside the chunk is
// packed into the low mantissa byte of d = ‖c‖² - 2·x·c (a 2^-15 relative perturbation,
// far below the TF32 product error), so one FMNMX tree yields both the minimum and where it
// was.  No data-dependent branch, no serial compare chain.
template <bool kNoLds = false>
__device__ __forceinline__ float packed_min32(const uint32_t (&v)[32], const float4* __restrict__ cn4) {
  float key[32];
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 cnv = kNoLds ? make_float4(1.f, 2.f, 3.f, 4.f) : cn4[j >> 2];
    const float d0 = fmaf(-2.f, __uint_as_float(v[j + 0]), cnv.x);
    const float d1 = fmaf(-2.f, __uint_as_float(v[j + 1]), cnv.y);
    const float d2 = fmaf(-2.f, __uint_as_float(v[j + 2]), cnv.z);
    const float d3 = fmaf(-2.f, __uint_as_float(v[j + 3]), cnv.w);
    // one PRMT each: low mantissa byte <- column index (AND+OR with two immediates costs two LOP3)
    key[j + 0] = __uint_as_float(__byte_perm(__float_as_uint(d0), (uint32_t)(j + 0), 0x3214));
    key[j + 1] = __uint_as_float(__byte_perm(__float_as_uint(d1), (uint32_t)(j + 1), 0x3214));
    key[j + 2] = __uint_as_float(__byte_perm(__float_as_uint(d2), (uint32_t)(j + 2), 0x3214));
    key[j + 3] = __uint_as_float(__byte_perm(__float_as_uint(d3), (uint32_t)(j + 3), 0x3214));
  }
#pragma unroll
  for (int w = 16; w > 0; w >>= 1) {
#pragma unroll
    for (int j = 0; j < w; ++j) key[j] = fminf(key[j], key[j + w]);
  }
  return key[0];
}

// Folds the 32 points owned by this warp (row0 + lane, arg-min index `bi` per lane) into
// sums / counts with 16-byte L2 reductions; returns this lane's share of Σ ‖x - c‖² (exact fp32).
__device__ __forceinline__ float fold_points(const float* __restrict__ X, const float* __restrict__ C,
                                             long long n, int d, long long row0, int bi, int lane,
                                             float* __restrict__ sums, float* __restrict__ counts,
                                             int* __restrict__ assign, bool want_obj,
                                             int i_begin = 0, int i_end = 32) {
  const int nl4 = d >> 2;
  if (assign != nullptr && i_begin == 0 && row0 + lane < n) assign[row0 + lane] = bi;
  float sq = 0.f;
  const bool act = lane < nl4;
  // 8 rows per batch: all loads are issued before the first reduction needs its data, so the
  // batch costs one memory round trip instead of eight.
#pragma unroll 1
  for (int i0 = i_begin; i0 < i_end; i0 += 8) {
    float4 xv[8], cv[8];
    int idx[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      idx[u] = __shfl_sync(0xffffffffu, bi, i0 + u);
      const long long row = row0 + i0 + u;
      xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      cv[u] = xv[u];
      if (act && row < n) {
        xv[u] = __ldcs(reinterpret_cast<const float4*>(X + (size_t)row * d) + lane);
        if (want_obj) cv[u] = __ldg(reinterpret_cast<const float4*>(C + (size_t)idx[u] * d) + lane);
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const long long row = row0 + i0 + u;
      if (row < n) {
        if (act) {
          red_add_v4(sums + (size_t)idx[u] * d + lane * 4, xv[u]);
          if (want_obj) {
            const float e0 = xv[u].x - cv[u].x, e1 = xv[u].y - cv[u].y, e2 = xv[u].z - cv[u].z,
                        e3 = xv[u].w - cv[u].w;
            sq = fmaf(e0, e0, sq); sq = fmaf(e1, e1, sq); sq = fmaf(e2, e2, sq); sq = fmaf(e3, e3, sq);
          }
        }
        if (lane == 0) red_add_f32(counts + idx[u], 1.0f);
      }
    }
  }
  return sq;
}

// ----------------------------------------------------------------------------
// Tensor-core assignment + accumulation kernel
// ----------------------------------------------------------------------------
__global__ void __launch_bounds__(NTHREADS, 1)
assign_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_c,
                 const float* __restrict__ X, const float* __restrict__ C,
                 const float* __restrict__ cn_pad, long long n, int d, int k, int kpad,
                 float* __restrict__ sums, float* __restrict__ counts, int* __restrict__ assign,
                 double* __restrict__ obj) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bars = base + SmemLayout::bars;
  auto bar = [&](int i) -> uint32_t { return bars + 8u * i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(base_ptr + SmemLayout::tmem_slot);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int kch = (d + BK - 1) / BK;
  const int nct = kpad / BN;                                   // centroid tiles
  const long long ntiles = (n + BM - 1) / BM;                  // point tiles
  const uint32_t x_bytes = (uint32_t)kch * X_CHUNK_BYTES;

  if (threadIdx.x == 0) {
    mbar_init(bar(BAR_X_FULL), 1);
    mbar_init(bar(BAR_X_EMPTY), 1);
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(bar(BAR_B_FULL + s), 1);
      mbar_init(bar(BAR_B_EMPTY + s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar(BAR_T_FULL + a), 1);
      mbar_init(bar(BAR_T_EMPTY + a), 4);  // one arrival per epilogue warp
      mbar_init(bar(BAR_CN_FULL + a), 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_1sm(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    tmem_relinquish_1sm();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_c);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const uint64_t pol_keep = policy_evict_last();    // centroids: re-read by every tile
      const uint64_t pol_stream = policy_evict_first(); // points: read once
      uint32_t it = 0;   // B-stage counter
      uint32_t ga = 0;   // centroid-tile counter
      uint32_t gp = 0;   // point-tile counter
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++gp) {
        mbar_wait(bar(BAR_X_EMPTY), (gp & 1) ^ 1);
        mbar_expect_tx(bar(BAR_X_FULL), x_bytes);
        for (int kc = 0; kc < kch; ++kc)
          tma_load_2d_hint(base + SmemLayout::x + kc * X_CHUNK_BYTES, &tmap_x, bar(BAR_X_FULL),
                           kc * BK, (int32_t)(tile * BM), pol_stream);
        for (int ct = 0; ct < nct; ++ct, ++ga) {
          const uint32_t a = ga & 1, aph = (ga >> 1) & 1;
          // the ‖c‖² slot is recycled together with its accumulator
          mbar_wait(bar(BAR_T_EMPTY + a), aph ^ 1);
          mbar_expect_tx(bar(BAR_CN_FULL + a), CN_BYTES);
          bulk_load_1d(base + SmemLayout::cn + a * CN_BYTES, cn_pad + (size_t)ct * BN, CN_BYTES,
                       bar(BAR_CN_FULL + a));
          for (int kc = 0; kc < kch; ++kc, ++it) {
            const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
            mbar_wait(bar(BAR_B_EMPTY + s), ph ^ 1);
            mbar_expect_tx(bar(BAR_B_FULL + s), B_STAGE_BYTES);
            tma_load_2d_hint(base + SmemLayout::b + s * B_STAGE_BYTES, &tmap_c, bar(BAR_B_FULL + s),
                             kc * BK, ct * BN, pol_keep);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp, one elected lane issues) =====================
    {
      const uint32_t idesc = umma_idesc_tf32(BM, BN);
      const uint64_t dhi = umma_desc_kmajor_sw128(0);
      uint32_t it = 0, ga = 0, gp = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++gp) {
        mbar_wait(bar(BAR_X_FULL), gp & 1);
        for (int ct = 0; ct < nct; ++ct, ++ga) {
          const uint32_t a = ga & 1, aph = (ga >> 1) & 1;
          mbar_wait(bar(BAR_T_EMPTY + a), aph ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + a * BN;
          for (int kc = 0; kc < kch; ++kc, ++it) {
            const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
            mbar_wait(bar(BAR_B_FULL + s), ph);
            tc_fence_after();
            const uint64_t da0 = dhi | (((base + SmemLayout::x + kc * X_CHUNK_BYTES) & 0x3FFFFu) >> 4);
            const uint64_t db0 = dhi | (((base + SmemLayout::b + s * B_STAGE_BYTES) & 0x3FFFFu) >> 4);
            int ksteps = (d - kc * BK + UK - 1) / UK;
            if (ksteps > BK / UK) ksteps = BK / UK;
            if (elect_one()) {
#pragma unroll
              for (int j = 0; j < BK / UK; ++j)
                if (j < ksteps)
                  umma_tf32_1sm(tmem_d, da0 + (uint64_t)(j * UK * 4 / 16), db0 + (uint64_t)(j * UK * 4 / 16), idesc,
                                (kc | j) != 0 ? 1u : 0u);
              umma_commit_1sm(bar(BAR_B_EMPTY + s));      // frees the smem stage when the MMAs retire
              if (kc == kch - 1) umma_commit_1sm(bar(BAR_T_FULL + a));   // accumulator ready for the epilogue
            }
            __syncwarp();
          }
        }
        if (elect_one()) umma_commit_1sm(bar(BAR_X_EMPTY));   // point tile no longer read by the tensor core
        __syncwarp();
      }
    }
  } else {
    // ===================== Epilogue warps (4 x 32 rows) =====================
    const int q = warp & 3;                              // TMEM lane quadrant of this warp
    uint32_t ga = 0;
    double obj_local = 0.0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      float best = CUDART_INF_F;
      int bi = 0;
      for (int ct = 0; ct < nct; ++ct, ++ga) {
        const uint32_t a = ga & 1, aph = (ga >> 1) & 1;
        mbar_wait(bar(BAR_CN_FULL + a), aph);
        mbar_wait(bar(BAR_T_FULL + a), aph);
        tc_fence_after();
        const float4* cn4 = reinterpret_cast<const float4*>(base_ptr + SmemLayout::cn + a * CN_BYTES);
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + a * BN;
        // Software-pipelined: the TMEM load of chunk c+1 is in flight while chunk c is reduced.
        // Per group of 4 columns: 4 FFMA + 3 FMNMX + 1 compare; the index search only runs when
        // the group actually improves on the running minimum (~ln(k) times per point).
        uint32_t va[32], vb[32];
        tmem_ld_32x32(taddr, va);
#pragma unroll 1
        for (int c = 0; c < BN / 32; c += 2) {
          tmem_ld_wait_regs(va);
          tmem_ld_32x32(taddr + (c + 1) * 32, vb);
          argmin_chunk(va, cn4 + c * 8, ct * BN + c * 32, best, bi);
          tmem_ld_wait_regs(vb);
          if (c + 2 < BN / 32) tmem_ld_32x32(taddr + (c + 2) * 32, va);
          argmin_chunk(vb, cn4 + (c + 1) * 8, ct * BN + (c + 1) * 32, best, bi);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_T_EMPTY + a));
      }
      // ---- fold the 32 points of this warp into sums / counts (L2 reductions) ----
      const long long row0 = tile * BM + q * 32;
      const float sq = fold_points(X, C, n, d, row0, bi, lane, sums, counts, assign, obj != nullptr);
      obj_local += (double)sq;
    }
    if (obj != nullptr) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) obj_local += __shfl_xor_sync(0xffffffffu, obj_local, o);
      if (lane == 0) red_add_f64(obj, obj_local);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_1sm(tmem_base, TMEM_COLS);
}


// ----------------------------------------------------------------------------
// v2: CTA-pair kernel (tcgen05 cta_group::2).  Two CTAs on one TPC share every
// centroid tile: each loads only its 128-row half of the 256-centroid tile and
// the tensor cores of both SMs read both halves, so L2->SMEM traffic per point
// is halved (the v1 kernel is bound by exactly that traffic: ncu shows 7.7 TB/s
// of L2 reads with the tensor pipe 46 % busy).  UMMA shape M=256 (128 points per
// CTA) x N=256 x K=8.  The point tile is double-buffered so the next tile's TMA
// overlaps the current tile's 40 centroid tiles.
// ----------------------------------------------------------------------------
namespace pair {
constexpr int NTHREADS2 = 320;                   // warps 0..7 epilogue, warp 8 TMA, warp 9 MMA
constexpr int NSTAGE2 = 5;
constexpr int NCN = 8;                           // ‖c‖² ring slots (decoupled from the 2 accumulators)
constexpr int NEXT = 3;                          // ring slots of the 4 KB GEMM-extension operand (EXT kernels)
constexpr int EXT_BYTES = 128 * 8 * 4;           // 128 rows x 8 floats (one tf32 K step), SWIZZLE_32B
constexpr int BH = BN / 2;                        // centroid rows each CTA loads
constexpr int B_HALF_BYTES = BH * BK * 4;         // 16 KB
constexpr int X_TILE_BYTES = MAX_KCH * X_CHUNK_BYTES;  // 64 KB
struct Smem {
  static constexpr int x = 0;                                   // 2 x 64 KB
  static constexpr int b = x + 2 * X_TILE_BYTES;                // NSTAGE2 x 16 KB
  static constexpr int cn = b + NSTAGE2 * B_HALF_BYTES;         // NCN x 1 KB  |  EXT: A_ext + NEXT x B_ext
  static constexpr int aext = cn;
  static constexpr int bext = cn + EXT_BYTES;
  static constexpr int bars = cn + 16384;
  static constexpr int tmem_slot = bars + 48 * 8;
  static constexpr int merge = tmem_slot + 16;                  // 128 x 4 B hand-over between warp halves
  static constexpr int total = merge + 128 * 4;
};
static_assert(Smem::total <= 232448, "pair kernel exceeds the 227 KB shared-memory limit");
enum : int {
  X_FULL = 0,                   // 2, used in the leader CTA
  X_EMPTY = 2,                  // 2, both CTAs (multicast commit)
  B_FULL = 4,                   // NSTAGE2, leader
  B_EMPTY = B_FULL + NSTAGE2,   // NSTAGE2, both
  T_FULL = B_EMPTY + NSTAGE2,   // 2, both
  T_EMPTY = T_FULL + 2,         // 2, leader (8 arrivals: 4 epilogue warps x 2 CTAs)
  CN_FULL = T_EMPTY + 2,        // NCN, local            | EXT: BX_FULL (leader) in the first NEXT slots
  CN_EMPTY = CN_FULL + NCN,     // NCN, local (8 arrivals)| EXT: BX_EMPTY (both)
  AEXT_FULL = CN_EMPTY + NCN,   // EXT: the constant A_ext tiles of both CTAs have landed (leader)
  NBARS = AEXT_FULL + 1
};
static_assert(NCN * CN_BYTES <= 16384 && EXT_BYTES * (1 + NEXT) <= 16384, "cn / ext region is 16 KB");
static_assert(NBARS <= 48, "barrier area too small");
}  // namespace pair

// K-major operand in the SWIZZLE_32B layout (32-byte rows = one tf32 K step, 8-row groups 256 B apart).
__device__ __forceinline__ uint64_t umma_desc_kmajor_sw32(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(256 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(6) << 61;             // SWIZZLE_32B
  return d;
}

// Arg-max counterpart of packed_min32 for the EXT kernels (accumulator already holds x·c - ‖c‖²/2).
__device__ __forceinline__ float packed_max32(const uint32_t (&v)[32]) {
  float key[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) key[j] = __uint_as_float(__byte_perm(v[j], (uint32_t)j, 0x3214));
#pragma unroll
  for (int w = 16; w > 0; w >>= 1) {
#pragma unroll
    for (int j = 0; j < w; ++j) key[j] = fmaxf(key[j], key[j + w]);
  }
  return key[0];
}

// Exact arg-max over 32 accumulator columns for the split-precision kernels: the packed form above overwrites the
// low mantissa byte (a 2^-15 relative perturbation, more than the split products' own error), so here the maximum
// is taken on the untouched values and the column is looked up only when a lane's running best improves — about
// ln(k) times per point, so the branch is cold.
__device__ __forceinline__ void exact_max32(const uint32_t (&v)[32], int cbase, float& bestx, int& bi) {
  float t[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) t[j] = fmaxf(__uint_as_float(v[j]), __uint_as_float(v[j + 16]));
#pragma unroll
  for (int w = 8; w > 0; w >>= 1) {
#pragma unroll
    for (int j = 0; j < w; ++j) t[j] = fmaxf(t[j], t[j + w]);
  }
  const float m = t[0];
  if (m > bestx) {
    bestx = m;
    int idx = 31;
#pragma unroll
    for (int j = 30; j >= 0; --j) idx = (__uint_as_float(v[j]) == m) ? j : idx;      // lowest column wins ties
    bi = cbase + idx;
  }
}

// ----------------------------------------------------------------------------
// Fused iteration (FUSED = 1): the same launch that computes distances / arg-min /
// partial sums also performs Harp's regroup -> centroid update -> allgather
// (ml/java/.../kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185) over
// NVLink peer memory, and prepares the GEMM-extension operand of the next iteration.
//
// Every rank keeps one "fused state" block at the same offset of its symmetric slab:
//   ctl | ones_ext | ext[2] | C[2] | acc[2]          ([2] = ping-pong by iteration parity)
// Iteration e reads C[e&1]/ext[e&1], folds into acc[e&1]; its tail writes C[~e&1]/ext[~e&1] of EVERY
// rank and zeroes the local acc[~e&1] for the next iteration.  Tail, executed by all CTAs (one per SM,
// co-resident):
//   1. local:  every CTA adds 1 to ctl.done1 and spins until all CTAs of this GPU have folded;
//   2. global: CTA 0 publishes flag1[rank] = e+1 on every peer; every CTA waits for all peers' flag1;
//   3. warp-per-row over this rank's 1/N of the centroids: sum the peers' partial sums and counts in
//      rank order (bit-deterministic), divide, push the row and its [-|c|^2/2 hi|mid|lo] operand row
//      into every peer; CTA 0 also adds up the peers' objective terms;
//   4. the last CTA through ctl.done2 publishes flag2[rank] = e+1 on every peer; CTA 0 waits for all
//      peers' flag2 before exiting, so kernel completion on a rank means its new centroids are whole
//      and nobody still reads its acc[e&1].
// ----------------------------------------------------------------------------
struct FusedCtl {
  uint32_t done1, done2, pad0[2];
  uint32_t flag1[HB_MAX_RANKS];
  uint32_t flag2[HB_MAX_RANKS];
  double obj[2];          // this rank's objective term per parity
  double obj_total;       // sum over ranks, written by the tail
  double pad1;
};
static_assert(sizeof(FusedCtl) <= 256, "FusedCtl must fit its 256-byte slot");

struct FusedArgs {
  uint8_t* peer[HB_MAX_RANKS];   // fused-state base of every rank (peer[rank] is local memory)
  int rank, world;
  uint32_t epoch;
  int keep_empty;
  unsigned long long off_ext[2], off_c[2], off_acc[2];
  unsigned long long off_cs[2];  // split-precision centroid rows (PREC = 1)
  float scale;                   // power of two applied to X and C before the 16-bit split (1 for bf16)
  int fmt16;                     // 0 = fp16 halves, 1 = bf16 halves
};

__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t target, bool sys, const char* what) {
  const long long t0 = clock64();
  while ((int32_t)((sys ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - target) < 0) {
    if (clock64() - t0 > 20000000000LL) {
      printf("harp_b200: fused K-means tail timed out waiting for %s (cta %d thread %d)\n", what, blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

template <int PREC>
__device__ __noinline__ void fused_tail(const FusedArgs& fa, int k, int d, uint32_t* smem_flag) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nwarps_cta = blockDim.x >> 5;
  const uint32_t cur = fa.epoch & 1u, nxt = cur ^ 1u, target = fa.epoch + 1u;
  uint8_t* me = fa.peer[fa.rank];
  FusedCtl* ctl = reinterpret_cast<FusedCtl*>(me);
  // 1. all CTAs of this GPU have issued their reductions into acc[cur]
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    atomicAdd(&ctl->done1, 1u);
    spin_until(&ctl->done1, gridDim.x, false, "local CTAs");
  }
  __syncthreads();
  // 2. every rank's partial sums are complete
  if (tid < fa.world) {
    if (blockIdx.x == 0) {
      fence_sys();
      st_release_sys(&reinterpret_cast<FusedCtl*>(fa.peer[tid])->flag1[fa.rank], target);
    }
    spin_until(&ctl->flag1[tid], target, true, "peer partial sums");
  }
  __syncthreads();
  // 3. regroup (reduce this rank's rows over all peers) -> divide -> allgather (push to all peers)
  const int world = fa.world;
  const int rows_per = (k + world - 1) / world;
  const int r0 = min(k, fa.rank * rows_per), r1 = min(k, r0 + rows_per);
  const int gw = blockIdx.x * nwarps_cta + warp, nw = gridDim.x * nwarps_cta;
  const size_t acc_cur = fa.off_acc[cur];
  // 4 rows per warp at a time and up to 4 peers' loads in flight per row, so a batch costs about two NVLink
  // round trips instead of rows x peers; sums are still added in rank order
  const int d4 = d >> 2;
  constexpr int RB = 4, PB = 4;
  for (int rowb = r0 + gw * RB; rowb < r1; rowb += nw * RB) {
    float cnt[RB];
#pragma unroll
    for (int u = 0; u < RB; ++u) cnt[u] = 0.f;
    for (int p0 = 0; p0 < world; p0 += PB) {
      float cv[PB][RB];
#pragma unroll
      for (int pp = 0; pp < PB; ++pp)
#pragma unroll
        for (int u = 0; u < RB; ++u) {
          cv[pp][u] = 0.f;
          if (p0 + pp < world && rowb + u < r1)
            cv[pp][u] = ld_volatile_f32(fa.peer[p0 + pp] + acc_cur + ((size_t)k * d + rowb + u) * sizeof(float));
        }
#pragma unroll
      for (int pp = 0; pp < PB; ++pp)
#pragma unroll
        for (int u = 0; u < RB; ++u) cnt[u] += cv[pp][u];
    }
    float sq[RB];
#pragma unroll
    for (int u = 0; u < RB; ++u) sq[u] = 0.f;
    for (int c4 = lane; c4 < d4; c4 += 32) {
      float4 sacc[RB];
#pragma unroll
      for (int u = 0; u < RB; ++u) sacc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int p0 = 0; p0 < world; p0 += PB) {
        float4 xv[PB][RB];
#pragma unroll
        for (int pp = 0; pp < PB; ++pp)
#pragma unroll
          for (int u = 0; u < RB; ++u) {
            xv[pp][u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p0 + pp < world && rowb + u < r1)
              xv[pp][u] = ld_volatile_f4(fa.peer[p0 + pp] + acc_cur + ((size_t)(rowb + u) * d + c4 * 4) * sizeof(float));
          }
#pragma unroll
        for (int pp = 0; pp < PB; ++pp)
#pragma unroll
          for (int u = 0; u < RB; ++u) {
            sacc[u].x += xv[pp][u].x; sacc[u].y += xv[pp][u].y; sacc[u].z += xv[pp][u].z; sacc[u].w += xv[pp][u].w;
          }
      }
#pragma unroll
      for (int u = 0; u < RB; ++u) {
        if (rowb + u >= r1) continue;
        const size_t eoff = ((size_t)(rowb + u) * d + c4 * 4) * sizeof(float);
        float4 c;
        if (cnt[u] > 0.f) c = make_float4(sacc[u].x / cnt[u], sacc[u].y / cnt[u], sacc[u].z / cnt[u], sacc[u].w / cnt[u]);
        else if (fa.keep_empty) c = *reinterpret_cast<const float4*>(me + fa.off_c[cur] + eoff);
        else c = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < world; ++p)
          *reinterpret_cast<float4*>(fa.peer[(fa.rank + p) % world] + fa.off_c[nxt] + eoff) = c;
        if (PREC == 1) {       // the next iteration's GEMM operand: the row split into 16-bit halves
          const int dp = split_dp(d);
          unsigned short h[4], l[4];
          split16(c.x * fa.scale, fa.fmt16, h[0], l[0]); split16(c.y * fa.scale, fa.fmt16, h[1], l[1]);
          split16(c.z * fa.scale, fa.fmt16, h[2], l[2]); split16(c.w * fa.scale, fa.fmt16, h[3], l[3]);
          const uint2 hv = make_uint2(h[0] | ((uint32_t)h[1] << 16), h[2] | ((uint32_t)h[3] << 16));
          const uint2 lv = make_uint2(l[0] | ((uint32_t)l[1] << 16), l[2] | ((uint32_t)l[3] << 16));
          const size_t soff = ((size_t)(rowb + u) * 2 * dp + c4 * 4) * sizeof(unsigned short);
          for (int p = 0; p < world; ++p) {
            uint8_t* dst = fa.peer[(fa.rank + p) % world] + fa.off_cs[nxt] + soff;
            *reinterpret_cast<uint2*>(dst) = hv;
            *reinterpret_cast<uint2*>(dst + dp * sizeof(unsigned short)) = lv;
          }
        }
        sq[u] = fmaf(c.x, c.x, sq[u]); sq[u] = fmaf(c.y, c.y, sq[u]); sq[u] = fmaf(c.z, c.z, sq[u]); sq[u] = fmaf(c.w, c.w, sq[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < RB; ++u) {
      float sv = sq[u];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sv += __shfl_xor_sync(0xffffffffu, sv, o);
      if (lane < 8 && rowb + u < r1) {
        float v;
        if (PREC == 1) {
          uint32_t w[2];
          bias_row_bf16(-0.5f * sv * fa.scale * fa.scale, w);
          v = __uint_as_float(lane < 2 ? w[lane] : 0u);
        } else {
          const float t0 = -0.5f * sv;
          const float hi = tf32_trunc(t0);
          const float mid = tf32_trunc(t0 - hi);
          const float lo = tf32_trunc(t0 - hi - mid);
          v = lane == 0 ? hi : (lane == 1 ? mid : (lane == 2 ? lo : 0.f));
        }
        for (int p = 0; p < world; ++p)
          *reinterpret_cast<float*>(fa.peer[(fa.rank + p) % world] + fa.off_ext[nxt] + ((size_t)(rowb + u) * 8 + lane) * sizeof(float)) = v;
      }
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    double tot = 0.0;
    for (int p = 0; p < world; ++p) {
      const volatile double* po = &reinterpret_cast<const FusedCtl*>(fa.peer[p])->obj[cur];
      tot += *po;
    }
    ctl->obj_total = tot;
  }
  // accumulators of the NEXT iteration: no peer reads acc[nxt] any more (each acknowledged the end of its
  // previous tail before our previous launch completed); zeroed here rather than at the start of the next
  // launch so the lines are still in L2 when the next launch's reductions arrive
  {
    float4* z = reinterpret_cast<float4*>(me + fa.off_acc[nxt]);
    const size_t total4 = ((size_t)k * d + k + 3) / 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + tid; i < total4; i += (size_t)gridDim.x * blockDim.x)
      z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (blockIdx.x == 0 && tid == 0) ctl->obj[nxt] = 0.0;
  }
  // 4. publish "my pushes are done, I no longer read your acc[cur]"
  fence_sys();
  __syncthreads();
  if (tid == 0) {
    const uint32_t old = atomicAdd(&ctl->done2, 1u);
    const bool last = (old == gridDim.x - 1);
    if (last) {            // every CTA is past its done1 spin: recycle both counters for the next launch
      ctl->done2 = 0u;
      ctl->done1 = 0u;
      fence_sys();
    }
    *smem_flag = last ? 1u : 0u;
  }
  __syncthreads();
  const bool last = (*smem_flag != 0u);
  if (tid < world) {
    if (last) {
      fence_sys();
      st_release_sys(&reinterpret_cast<FusedCtl*>(fa.peer[tid])->flag2[fa.rank], target);
    }
    if (blockIdx.x == 0) spin_until(&ctl->flag2[tid], target, true, "peer centroid pushes");
  }
}

template <int DBG, int EXT, int FUSED = 0, int PREC = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(pair::NTHREADS2, 1)
assign_tc_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_c,
                      const __grid_constant__ CUtensorMap tmap_ones, const __grid_constant__ CUtensorMap tmap_cext,
                      const float* __restrict__ X, const float* __restrict__ C,
                      const float* __restrict__ cn_pad, long long n, int d, int k, int kpad,
                      float* __restrict__ sums, float* __restrict__ counts, int* __restrict__ assign,
                      double* __restrict__ obj, const FusedArgs fa) {
  using namespace pair;
  constexpr int dbg = DBG;
  // Role layout: epilogue warps 0..7 (two per TMEM lane quadrant, which is fixed by warp id % 4), then the
  // TMA producer and the MMA issuer.
  constexpr int TMA_WARP = 8, MMA_WARP = 9;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0) { if (threadIdx.x == 0) printf("harp_b200: smem base not 1024-aligned\n"); __trap(); }
  const uint32_t bars = base + Smem::bars;
  auto bar = [&](int i) -> uint32_t { return bars + 8u * i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem_raw + Smem::tmem_slot);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();           // 0 = leader
  const bool leader = (cta == 0);
  // K-chunks (one 128-byte swizzle row each) of a point tile / per centroid tile: fp32 -> d / 32 chunks of 32
  // floats; split precision -> hi chunks then lo chunks of 64 halves
  const int dh = (d + 63) / 64;
  const int kch = PREC == 1 ? 2 * dh : (d + BK - 1) / BK;
  const int bke = PREC == 1 ? 64 : BK;                      // elements per chunk (TMA coordinate step)
  const int nct = kpad / BN;
  const long long npairs = (n + 2 * BM - 1) / (2 * BM);   // pair-tiles of 256 points
  const long long cluster_id = blockIdx.x >> 1;
  const long long nclusters = gridDim.x >> 1;
  const uint32_t x_bytes = (uint32_t)kch * X_CHUNK_BYTES;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(X_FULL + i), 1);
      mbar_init(bar(X_EMPTY + i), 1);
      mbar_init(bar(T_FULL + i), 1);
      mbar_init(bar(T_EMPTY + i), 16);   // 8 epilogue warps x 2 CTAs
    }
    for (int i = 0; i < NCN; ++i) {
      mbar_init(bar(CN_FULL + i), 1);
      mbar_init(bar(CN_EMPTY + i), EXT ? 1 : 8);
    }
    mbar_init(bar(AEXT_FULL), 1);
    for (int s = 0; s < NSTAGE2; ++s) {
      mbar_init(bar(B_FULL + s), 1);
      mbar_init(bar(B_EMPTY + s), 1);
    }
    fence_barrier_init();
  }
  if (warp == MMA_WARP) {
    tmem_alloc_2sm(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    tmem_relinquish_2sm();
  }
  if (warp == TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_c);
  }
  tc_fence_before();
  cluster_sync_all();            // barriers of both CTAs initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == TMA_WARP) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      const uint64_t pol_keep = policy_evict_last();
      const uint64_t pol_stream = policy_evict_first();
      uint32_t it = 0, ga = 0, gp = 0;
      if (EXT) {      // the constant [1,1,1,0..] operand rows, once per CTA
        if (leader) mbar_expect_tx(bar(AEXT_FULL), 2 * EXT_BYTES);
        tma_load_2d_2sm(base + Smem::aext, &tmap_ones, bar(AEXT_FULL) & kPeerBitMask, 0, 0, pol_keep);
      }
      for (long long tp = cluster_id; tp < npairs; tp += nclusters, ++gp) {
        const uint32_t xb = gp & 1, xph = (gp >> 1) & 1;
        mbar_wait(bar(X_EMPTY + xb), xph ^ 1);
        if (leader) mbar_expect_tx(bar(X_FULL + xb), 2 * x_bytes);
        const int32_t row0 = (int32_t)(tp * 2 * BM + cta * BM);
        for (int kc = 0; kc < kch; ++kc)
          tma_load_2d_2sm(base + Smem::x + xb * X_TILE_BYTES + kc * X_CHUNK_BYTES, &tmap_x,
                          bar(X_FULL + xb) & kPeerBitMask, kc * bke, row0, pol_stream);
        for (int ct = 0; ct < nct; ++ct, ++ga) {
          // ‖c‖² tile -> its own 8-deep ring: with only as many slots as accumulators this
          // prefetch would be gated on the epilogue of tile ct-2 and epilogue + TMA latency
          // would sit on the tensor core's critical path.
          if (!EXT) {
            const uint32_t cs = ga % NCN, cph = (ga / NCN) & 1;
            mbar_wait(bar(CN_EMPTY + cs), cph ^ 1);
            mbar_expect_tx(bar(CN_FULL + cs), CN_BYTES);
            bulk_load_1d(base + Smem::cn + cs * CN_BYTES, cn_pad + (size_t)ct * BN, CN_BYTES,
                         bar(CN_FULL + cs));
          } else {
            // EXT: this CTA's 128 rows of the [-‖c‖²/2 hi|mid|lo] operand for centroid tile ct
            const uint32_t es = ga % NEXT, eph = (ga / NEXT) & 1;
            mbar_wait(bar(CN_EMPTY + es), eph ^ 1);
            if (leader) mbar_expect_tx(bar(CN_FULL + es), 2 * EXT_BYTES);
            tma_load_2d_2sm(base + Smem::bext + es * EXT_BYTES, &tmap_cext, bar(CN_FULL + es) & kPeerBitMask,
                            0, ct * BN + (int)cta * BH, pol_keep);
          }
          for (int kc = 0; kc < kch; ++kc, ++it) {
            const uint32_t s = it % NSTAGE2, ph = (it / NSTAGE2) & 1;
            mbar_wait(bar(B_EMPTY + s), ph ^ 1);
            if (leader) mbar_expect_tx(bar(B_FULL + s), 2 * B_HALF_BYTES);
            tma_load_2d_2sm(base + Smem::b + s * B_HALF_BYTES, &tmap_c, bar(B_FULL + s) & kPeerBitMask,
                            kc * bke, ct * BN + (int)cta * BH, pol_keep);
          }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer (leader CTA) =====================
    // The whole warp walks the loop converged and one elected lane issues: descriptors stay
    // warp-uniform, so the compiler keeps them in uniform registers instead of emitting an
    // ELECT / R2UR.BROADCAST loop per tcgen05 instruction.  The issue latency of a tile sits
    // on the epilogue -> issuer -> tensor pipe -> epilogue chain, so every instruction counts.
    if (leader) {
      const uint32_t idesc = umma_idesc_tf32(2 * BM, BN);
      const uint32_t idesc16 = umma_idesc_16(2 * BM, BN, (uint32_t)fa.fmt16);      // split halves: fp16 or bf16
      const uint32_t idesc_b16 = umma_idesc_16(2 * BM, BN, 1u);                    // bias step: always bf16
      const uint64_t dhi = umma_desc_kmajor_sw128(0);          // address-free part of the descriptor
      uint32_t it = 0, ga = 0, gp = 0;
      long long prof_issue = 0, prof_n = 0;
      if (EXT) mbar_wait(bar(AEXT_FULL), 0);
      for (long long tp = cluster_id; tp < npairs; tp += nclusters, ++gp) {
        const uint32_t xb = gp & 1, xph = (gp >> 1) & 1;
        mbar_wait(bar(X_FULL + xb), xph);
        for (int ct = 0; ct < nct; ++ct, ++ga) {
          const uint32_t a = ga & 1, aph = (ga >> 1) & 1;
          mbar_wait_cluster(bar(T_EMPTY + a), aph ^ 1);
          tc_fence_after();
          long long t0 = 0;
          if (dbg & 16) t0 = clock64();
          const uint32_t tmem_d = tmem_base + a * BN;
          for (int kc = 0; kc < kch; ++kc, ++it) {
            const uint32_t s = it % NSTAGE2, ph = (it / NSTAGE2) & 1;
            mbar_wait(bar(B_FULL + s), ph);
            tc_fence_after();
            const uint64_t db0 = dhi | (((base + Smem::b + s * B_HALF_BYTES) & 0x3FFFFu) >> 4);
            if (PREC == 1) {
              // stage kc holds B chunk kc: hi half (kc < dh) meets A hi AND A lo, lo half meets A hi only:
              // x·c ~ hi·hi + lo·hi + hi·lo, 4 K=16 steps per 64-element chunk
              const int hb = kc < dh ? kc : kc - dh;
              const uint64_t da_hi = dhi | (((base + Smem::x + xb * X_TILE_BYTES + hb * X_CHUNK_BYTES) & 0x3FFFFu) >> 4);
              const uint64_t da_lo = dhi | (((base + Smem::x + xb * X_TILE_BYTES + (dh + hb) * X_CHUNK_BYTES) & 0x3FFFFu) >> 4);
              int ksteps = (d - hb * 64 + 15) / 16;
              if (ksteps > 4) ksteps = 4;
              if (elect_one()) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (j < ksteps) umma_f16_2sm(tmem_d, da_hi + (uint64_t)(2 * j), db0 + (uint64_t)(2 * j), idesc16, (kc | j) != 0 ? 1u : 0u);
                if (kc < dh) {
#pragma unroll
                  for (int j = 0; j < 4; ++j)
                    if (j < ksteps) umma_f16_2sm(tmem_d, da_lo + (uint64_t)(2 * j), db0 + (uint64_t)(2 * j), idesc16, 1u);
                }
                umma_commit_2sm_mc(bar(B_EMPTY + s), 0x3);
              }
              __syncwarp();
              continue;
            }
            const uint64_t da0 = dhi | (((base + Smem::x + xb * X_TILE_BYTES + kc * X_CHUNK_BYTES) & 0x3FFFFu) >> 4);
            int ksteps = (d - kc * BK + UK - 1) / UK;
            if (ksteps > BK / UK) ksteps = BK / UK;
            if (elect_one()) {
#pragma unroll
              for (int j = 0; j < BK / UK; ++j)
                if (j < ksteps)
                  umma_tf32_2sm(tmem_d, da0 + (uint64_t)(j * UK * 4 / 16), db0 + (uint64_t)(j * UK * 4 / 16), idesc,
                                (kc | j) != 0 ? 1u : 0u);
              umma_commit_2sm_mc(bar(B_EMPTY + s), 0x3);     // frees the stage in both CTAs
            }
            __syncwarp();
          }
          if (EXT) {   // one more K=8 step: accumulator += [1,1,1,0..] · [-‖c‖²/2 hi,mid,lo,0..]ᵀ
            const uint32_t es = ga % NEXT, eph = (ga / NEXT) & 1;
            mbar_wait(bar(CN_FULL + es), eph);
            tc_fence_after();
            if (elect_one()) {
              if (PREC == 1)
                umma_f16_2sm(tmem_d, umma_desc_kmajor_sw32(base + Smem::aext),
                             umma_desc_kmajor_sw32(base + Smem::bext + es * EXT_BYTES), idesc_b16, 1u);
              else
              umma_tf32_2sm(tmem_d, umma_desc_kmajor_sw32(base + Smem::aext),
                            umma_desc_kmajor_sw32(base + Smem::bext + es * EXT_BYTES), idesc, 1u);
              umma_commit_2sm_mc(bar(CN_EMPTY + es), 0x3);
              umma_commit_2sm_mc(bar(T_FULL + a), 0x3);        // accumulators ready in both CTAs
            }
          } else if (elect_one()) {
            umma_commit_2sm_mc(bar(T_FULL + a), 0x3);
          }
          __syncwarp();
          if (dbg & 16) { prof_issue += clock64() - t0; ++prof_n; }
        }
        if (elect_one()) umma_commit_2sm_mc(bar(X_EMPTY + xb), 0x3);
        __syncwarp();
      }
      if ((dbg & 16) && lane == 0 && blockIdx.x < 2 && prof_n > 0)
        printf("prof cta %d issuer: tiles %lld issue/tile %lld cycles\n", blockIdx.x, prof_n, prof_issue / prof_n);
    }
  } else {
    // ===================== Epilogue warps (both CTAs) =====================
    // 8 warps: two per TMEM lane quadrant; warp (q, h) reduces columns [h*128, h*128+128) of
    // every accumulator for the quadrant's 32 point rows.
    const int q = warp & 3;
    const int h = warp >> 2;
    constexpr int CH = BN / 32 / 2;          // 4 chunks of 32 columns per warp per tile
    float* merge = reinterpret_cast<float*>(smem_raw + Smem::merge);
    uint32_t ga = 0;
    double obj_local = 0.0;
    long long prof_wait = 0, prof_work = 0, prof_n = 0;
    for (long long tp = cluster_id; tp < npairs; tp += nclusters) {
      // !EXT: best = min of ‖c‖² - 2 x·c;  EXT: best = -(max of x·c - ‖c‖²/2), kept negated so the merge below is shared
      float best = CUDART_INF_F;
      float bestx = -CUDART_INF_F;
      int bi = 0;
      for (int ct = 0; ct < nct; ++ct, ++ga) {
        const uint32_t a = ga & 1, aph = (ga >> 1) & 1;
        const int tbase = ct * BN + h * (BN / 2);
        const uint32_t cs = ga % NCN, cph = (ga / NCN) & 1;
        const float4* cng = reinterpret_cast<const float4*>(smem_raw + Smem::cn + cs * CN_BYTES) + h * (BN / 8);
        long long tw0 = 0;
        if (dbg & 16) tw0 = clock64();
        if (!EXT) mbar_wait(bar(CN_FULL + cs), cph);
        mbar_wait(bar(T_FULL + a), aph);
        if (dbg & 16) { const long long now = clock64(); prof_wait += now - tw0; tw0 = now; }
        tc_fence_after();
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + a * BN + h * (BN / 2);
        const bool skip_math = ((dbg & 32) && !leader) || ((dbg & 64) && leader);
        if (EXT == 2 && !(dbg & 1) && !skip_math) {
          // one TMEM load per warp per tile: all 128 columns of this warp's half at once
          uint32_t v[128];
          tmem_ld_32x128(taddr, v);
          tmem_ld_wait_regs128(v);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(bar(T_EMPTY + a), 0);     // accumulator free as soon as it is in registers
          float key[128];
#pragma unroll
          for (int j = 0; j < 128; ++j) key[j] = __uint_as_float(__byte_perm(v[j], (uint32_t)j, 0x3214));
#pragma unroll
          for (int w = 64; w > 0; w >>= 1) {
#pragma unroll
            for (int j = 0; j < w; ++j) key[j] = fmaxf(key[j], key[j + w]);
          }
          const bool gt = key[0] > bestx;
          bestx = gt ? key[0] : bestx;
          bi = gt ? tbase : bi;
          if (dbg & 16) { prof_work += clock64() - tw0; ++prof_n; }
          continue;
        }
        if (!(dbg & 1) && !skip_math) {
          uint32_t va[32], vb[32];
          tmem_ld_32x32(taddr, va);
#pragma unroll
          for (int c = 0; c < CH; c += 2) {
            tmem_ld_wait_regs(va);
            tmem_ld_32x32(taddr + (c + 1) * 32, vb);
            if (dbg & 2) { best = fminf(best, __uint_as_float(va[0]) + __uint_as_float(va[31])); } else
            if (PREC == 1) {
              exact_max32(va, tbase + c * 32, bestx, bi);
            } else if (EXT) {
              const float m = packed_max32(va);
              const bool gt = m > bestx;
              bestx = gt ? m : bestx;
              bi = gt ? tbase + c * 32 : bi;
            } else {
              const float m = packed_min32<(DBG & 8) != 0>(va, cng + c * 8);
              const bool lt = m < best;
              best = lt ? m : best;
              bi = lt ? tbase + c * 32 : bi;
            }
            tmem_ld_wait_regs(vb);
            if (c + 2 < CH) tmem_ld_32x32(taddr + (c + 2) * 32, va);
            else {
              // last chunk is in registers: hand the accumulator back before doing its math
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(bar(T_EMPTY + a), 0);
            }
            if (dbg & 2) { best = fminf(best, __uint_as_float(vb[0]) + __uint_as_float(vb[31])); } else
            if (PREC == 1) {
              exact_max32(vb, tbase + (c + 1) * 32, bestx, bi);
            } else if (EXT) {
              const float m = packed_max32(vb);
              const bool gt = m > bestx;
              bestx = gt ? m : bestx;
              bi = gt ? tbase + (c + 1) * 32 : bi;
            } else {
              const float m = packed_min32<(DBG & 8) != 0>(vb, cng + (c + 1) * 8);
              const bool lt = m < best;
              best = lt ? m : best;
              bi = lt ? tbase + (c + 1) * 32 : bi;
            }
          }
        }
        if ((dbg & 1) || skip_math) {                     // (debug paths skip the loop that hands it back)
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(bar(T_EMPTY + a), 0);
        }
        if (!EXT) {
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(CN_EMPTY + cs));   // local ‖c‖² slot
        }
        if (dbg & 16) { prof_work += clock64() - tw0; ++prof_n; }
      }
      // unpack the in-chunk column, then merge the two column halves of each row through smem
      if (PREC == 1) best = -bestx;       // exact arg-max: bi is already the full index
      else if (EXT) { bi += (int)(__float_as_uint(bestx) & 255u); best = -bestx; }
      else bi += (int)(__float_as_uint(best) & 255u);
      // (512 B of smem: best goes over first, then the index, then the verdict comes back)
      const int slot = q * 32 + lane;
      if (h == 1) merge[slot] = best;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float obest = 0.f;
      if (h == 0) obest = merge[slot];
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (h == 1) merge[slot] = __int_as_float(bi);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (h == 0) {
        if (obest < best) { best = obest; bi = __float_as_int(merge[slot]); }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (h == 0) merge[slot] = __int_as_float(bi);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (h == 1) bi = __float_as_int(merge[slot]);
      asm volatile("bar.sync 1, 256;" ::: "memory");       // merge[] is rewritten by the next tile
      const long long row0 = tp * 2 * BM + (long long)cta * BM + q * 32;
      if (!(dbg & 4))
        obj_local += fold_points(X, C, n, d, row0, bi, lane, sums, counts, assign, obj != nullptr,
                                 h * 16, h * 16 + 16);
      else if (best == 12345.f && assign) assign[0] = bi;
    }
    if (obj != nullptr) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) obj_local += __shfl_xor_sync(0xffffffffu, obj_local, o);
      if (lane == 0) red_add_f64(obj, obj_local);
    }
    if ((dbg & 16) && lane == 0 && blockIdx.x < 2 && prof_n > 0)
      printf("prof cta %d warp %d: tiles %lld wait/tile %lld work/tile %lld cycles\n", blockIdx.x, warp,
             prof_n, prof_wait / prof_n, prof_work / prof_n);
  }

  tc_fence_before();
  cluster_sync_all();            // nobody leaves while the peer may still touch its smem / TMEM / barriers
  if (warp == MMA_WARP) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  if (FUSED) fused_tail<PREC>(fa, k, d, reinterpret_cast<uint32_t*>(smem_raw + Smem::merge));
}

// ----------------------------------------------------------------------------
// Exact fp32 SIMT fallback: (x-c)² form, one thread per point, centroids staged
// through shared memory.
// ----------------------------------------------------------------------------
__global__ void assign_simt_kernel(const float* __restrict__ X, long long n, int d,
                                   const float* __restrict__ C, int k, int tc,
                                   int* __restrict__ assign, float* __restrict__ bestd) {
  extern __shared__ float cs[];
  const long long row = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  float best = CUDART_INF_F;
  int bi = 0;
  for (int c0 = 0; c0 < k; c0 += tc) {
    const int nc = min(tc, k - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < nc * d; i += blockDim.x) cs[i] = C[(size_t)c0 * d + i];
    __syncthreads();
    if (row < n) {
      const float* x = X + (size_t)row * d;
      for (int c = 0; c < nc; ++c) {
        const float* cc = cs + c * d;
        float acc = 0.f;
        for (int j = 0; j < d; ++j) {
          const float e = x[j] - cc[j];
          acc = fmaf(e, e, acc);
        }
        if (acc < best) { best = acc; bi = c0 + c; }
      }
    }
  }
  if (row < n) {
    assign[row] = bi;
    bestd[row] = best;
  }
}

__global__ void accumulate_simt_kernel(const float* __restrict__ X, long long n, int d,
                                       const int* __restrict__ assign,
                                       const float* __restrict__ bestd, float* __restrict__ sums,
                                       float* __restrict__ counts, double* __restrict__ obj) {
  const int warp_in_block = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long wid = blockIdx.x * (long long)(blockDim.x >> 5) + warp_in_block;
  const long long nw = (long long)gridDim.x * (blockDim.x >> 5);
  double o = 0.0;
  for (long long row = wid; row < n; row += nw) {
    const int idx = assign[row];
    for (int j = lane; j < d; j += 32) atomicAdd(sums + (size_t)idx * d + j, X[(size_t)row * d + j]);
    if (lane == 0) {
      atomicAdd(counts + idx, 1.0f);
      o += (double)bestd[row];
    }
  }
  if (obj != nullptr && lane == 0 && o != 0.0) atomicAdd(obj, o);
}

// ----------------------------------------------------------------------------
// finalize: centroid = sum / count; clusters that attracted nothing collapse to
// the zero vector (ml/java semantics, KMeansCollectiveMapper.java:170-183) or
// keep their previous position (keep_empty != 0).
// ----------------------------------------------------------------------------
__global__ void finalize_kernel(const float* __restrict__ sums, const float* __restrict__ counts,
                                int k, int d, float* __restrict__ C, int keep_empty) {
  const size_t total = (size_t)k * d;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const float cnt = counts[i / d];
    if (cnt > 0.f) C[i] = sums[i] / cnt;
    else if (!keep_empty) C[i] = 0.f;
  }
}

}  // namespace kmeans
}  // namespace hb

using namespace hb;
using namespace hb::kmeans;

extern "C" {

// Scratch layout: [ ‖c‖² : kpad floats | pad to 256 B ][ -‖c‖²/2 as 3 tf32 terms : kpad x 8 floats ][ ones : 128 x 8 ].
// The first part alone (hb_kmeans_min_workspace_bytes) is enough for the single-CTA kernel; a caller that
// provides only that much gets that kernel.
static size_t ws_base_bytes(int k) {
  const size_t kpad = ((size_t)k + BN - 1) / BN * BN;
  return kpad * sizeof(float) + 256;
}
size_t hb_kmeans_min_workspace_bytes(int k, int d) {
  (void)d;
  return ws_base_bytes(k);
}
size_t hb_kmeans_workspace_bytes(int k, int d) {
  const size_t kpad = ((size_t)k + BN - 1) / BN * BN;
  (void)d;
  return ws_base_bytes(k) + kpad * 8 * sizeof(float) + 128 * 8 * sizeof(float);
}

int hb_kmeans_tc_supported(int d, int k) { return (d % 4 == 0 && d >= 4 && d <= BK * MAX_KCH && k >= 1) ? 1 : 0; }

int hb_kmeans_step(const float* X, long long n, int d, const float* C, int k, float* sums,
                   float* counts, int* assign, double* obj, void* workspace, size_t workspace_bytes,
                   int flags, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(n >= 0 && d >= 1 && k >= 1, "hb_kmeans_step: bad shape n=%lld d=%d k=%d", n, d, k);
  const int kpad = (k + BN - 1) / BN * BN;
  HB_REQUIRE(workspace != nullptr && workspace_bytes >= ws_base_bytes(k),
             "hb_kmeans_step: workspace too small (%zu < %zu)", workspace_bytes, ws_base_bytes(k));
  const bool have_ext = workspace_bytes >= hb_kmeans_workspace_bytes(k, d);
  HB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
             "hb_kmeans_step: workspace must be 256-byte aligned");
  float* cn = reinterpret_cast<float*>(workspace);
  float* cn_ext = have_ext ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + ws_base_bytes(k)) : nullptr;
  float* ones_ext = have_ext ? cn_ext + (size_t)kpad * 8 : nullptr;
  const bool force_simt = (flags & HB_KMEANS_FORCE_SIMT) != 0;
  const bool zero_acc = (flags & HB_KMEANS_NO_ZERO) == 0;
  const bool use_tc = !force_simt && hb_kmeans_tc_supported(d, k) &&
                      (reinterpret_cast<uintptr_t>(X) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(sums) & 15) == 0;

  {
    const int threads = 256;
    int blocks = (kpad * 32 + threads - 1) / threads;
    const int cap = sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    prep_kernel<<<blocks, threads, 0, stream>>>(C, k, d, kpad, cn, cn_ext, ones_ext, sums, counts, obj, zero_acc ? 1 : 0);
    HB_CUDA_OK(cudaGetLastError());
  }
  if (n == 0) return 0;

  if (use_tc) {
    CUtensorMap tx, tc;
    if (make_tmap_2d_f32(&tx, X, (uint64_t)n, (uint64_t)d, (uint64_t)d * 4, BM, BK,
                         CU_TENSOR_MAP_SWIZZLE_128B) != 0)
      return -3;
    if (make_tmap_2d_f32(&tc, C, (uint64_t)k, (uint64_t)d, (uint64_t)d * 4, BN, BK,
                         CU_TENSOR_MAP_SWIZZLE_128B) != 0)
      return -3;
    int dev = 0;
    HB_CUDA_OK(cudaGetDevice(&dev));
    const long long ntiles = (n + BM - 1) / BM;
    const bool use_pair = (flags & HB_KMEANS_SINGLE_CTA) == 0 && ntiles >= 2;
    if (use_pair) {
      // CTA-pair kernel: each CTA loads a 128-row half of every centroid tile
      CUtensorMap tch;
      if (make_tmap_2d_f32(&tch, C, (uint64_t)k, (uint64_t)d, (uint64_t)d * 4, BN / 2, BK,
                           CU_TENSOR_MAP_SWIZZLE_128B) != 0)
        return -3;
      const int dbg = (flags >> 8) & 0xff;
      // EXT (default when the workspace holds the extension operands): ‖c‖² folded into the GEMM
      const bool ext = have_ext && (flags & HB_KMEANS_NO_EXT) == 0;
      const bool ext128 = ext && (flags & HB_KMEANS_EXT_X128) != 0;
      CUtensorMap tones = tch, tcext = tch;
      if (ext) {
        if (make_tmap_2d_f32(&tones, ones_ext, 128, 8, 32, 128, 8, CU_TENSOR_MAP_SWIZZLE_32B) != 0) return -3;
        if (make_tmap_2d_f32(&tcext, cn_ext, (uint64_t)kpad, 8, 32, BN / 2, 8, CU_TENSOR_MAP_SWIZZLE_32B) != 0) return -3;
      }
      auto kern = ext128 ? assign_tc_pair_kernel<0, 2> : (ext ? assign_tc_pair_kernel<0, 1> : assign_tc_pair_kernel<0, 0>);
#ifdef HB_DEBUG_KERNELS       // triage variants (masked epilogues, per-role cycle counters): not in the product build
      if (dbg == 1) kern = assign_tc_pair_kernel<1, 0>;
      else if (dbg == 2) kern = assign_tc_pair_kernel<2, 0>;
      else if (dbg == 4) kern = ext ? assign_tc_pair_kernel<4, 1> : assign_tc_pair_kernel<4, 0>;
      else if (dbg == 7) kern = assign_tc_pair_kernel<7, 0>;
      else if (dbg == 36) kern = assign_tc_pair_kernel<36, 1>;
      else if (dbg == 68) kern = assign_tc_pair_kernel<68, 1>;
      else if (dbg == 16) kern = ext128 ? assign_tc_pair_kernel<16, 2> : (ext ? assign_tc_pair_kernel<16, 1> : assign_tc_pair_kernel<16, 0>);
#else
      HB_REQUIRE(dbg == 0, "hb_kmeans_step: debug kernel variant %d needs a build with -DHB_DEBUG_KERNELS", dbg);
#endif
      HB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pair::Smem::total));
      const long long npairs = (n + 2 * BM - 1) / (2 * BM);
      long long nclusters = sm_count() / 2;
      if (const char* e = getenv("HB_KMEANS_MAX_CLUSTERS")) {   // perf-triage knob
        const long long cap = atoll(e);
        if (cap > 0 && cap < nclusters) nclusters = cap;
      }
      if (nclusters > npairs) nclusters = npairs;
      FusedArgs nofuse;
      memset(&nofuse, 0, sizeof(nofuse));
      kern<<<(unsigned)(2 * nclusters), pair::NTHREADS2, pair::Smem::total, stream>>>(
          tx, tch, tones, tcext, X, C, cn, n, d, k, kpad, sums, counts, assign, obj, nofuse);
      HB_CUDA_OK(cudaGetLastError());
    } else {
      // (cheap and idempotent: set on every call rather than cached in an unsynchronised host static)
      HB_CUDA_OK(cudaFuncSetAttribute(assign_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
      int grid = sm_count();
      if ((long long)grid > ntiles) grid = (int)ntiles;
      assign_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, stream>>>(tx, tc, X, C, cn, n, d, k, kpad, sums,
                                                               counts, assign, obj);
      HB_CUDA_OK(cudaGetLastError());
    }
  } else {
    // SIMT path needs an assignment + distance buffer; it lives behind ‖c‖² in the workspace
    // only if the caller sized it with hb_kmeans_simt_workspace_bytes.
    const size_t need = hb_kmeans_workspace_bytes(k, d) + (size_t)n * 8;
    HB_REQUIRE(workspace_bytes >= need,
               "hb_kmeans_step: SIMT path needs %zu workspace bytes (got %zu)", need, workspace_bytes);
    uint8_t* p = reinterpret_cast<uint8_t*>(workspace) + hb_kmeans_workspace_bytes(k, d);
    int* asg = reinterpret_cast<int*>(p);
    float* bd = reinterpret_cast<float*>(p + (size_t)n * 4);
    int tcn = (32 * 1024) / (4 * d);
    if (tcn < 1) tcn = 1;
    if (tcn > 64) tcn = 64;
    if (tcn > k) tcn = k;
    const size_t smem = (size_t)tcn * d * 4;
    HB_REQUIRE(smem <= 48 * 1024, "hb_kmeans_step: d=%d too large for the SIMT fallback", d);
    const int threads = 128;
    const long long blocks = (n + threads - 1) / threads;
    assign_simt_kernel<<<(unsigned)blocks, threads, smem, stream>>>(X, n, d, C, k, tcn, asg, bd);
    HB_CUDA_OK(cudaGetLastError());
    accumulate_simt_kernel<<<sm_count() * 8, 256, 0, stream>>>(X, n, d, asg, bd, sums, counts, obj);
    HB_CUDA_OK(cudaGetLastError());
    if (assign != nullptr)
      HB_CUDA_OK(cudaMemcpyAsync(assign, asg, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

// ---- fused iteration: distance GEMM + arg-min + partial sums + regroup/update/allgather in one launch ----
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
struct FusedLayout { size_t ones, ext[2], c[2], acc[2], cs[2], total; };
static FusedLayout fused_layout(int k, int d) {
  const size_t kpad = ((size_t)k + BN - 1) / BN * BN;
  FusedLayout L;
  size_t o = 256;                                   // FusedCtl
  L.ones = o; o += 128 * 8 * sizeof(float);
  for (int i = 0; i < 2; ++i) { L.ext[i] = o; o = align_up(o + kpad * 8 * sizeof(float), 256); }
  for (int i = 0; i < 2; ++i) { L.c[i] = o; o = align_up(o + (size_t)k * d * sizeof(float), 256); }
  for (int i = 0; i < 2; ++i) { L.acc[i] = o; o = align_up(o + ((size_t)k * d + k + 3) / 4 * 16, 256); }
  for (int i = 0; i < 2; ++i) { L.cs[i] = o; o = align_up(o + (size_t)k * 2 * split_dp(d) * sizeof(unsigned short), 256); }
  L.total = o;
  return L;
}
size_t hb_kmeans_fused_state_bytes(int k, int d) { return fused_layout(k, d).total; }

int hb_kmeans_fused_offsets(int k, int d, unsigned long long* out8) {
  HB_REQUIRE(out8 != nullptr, "hb_kmeans_fused_offsets: null output");
  const FusedLayout L = fused_layout(k, d);
  out8[0] = L.c[0]; out8[1] = L.c[1]; out8[2] = L.acc[0]; out8[3] = L.acc[1];
  out8[4] = offsetof(FusedCtl, obj); out8[5] = offsetof(FusedCtl, obj_total); out8[6] = L.ext[0]; out8[7] = L.ext[1];
  return 0;
}

// Split-precision helpers.  prec: 0 = TF32 products (default), 1 = fp16 halves (values scaled by `scale`, a power of
// two that brings max |x| near 2^14), 2 = bf16 halves (scale 1).
size_t hb_kmeans_split_row_bytes(int d) { return (size_t)2 * split_dp(d) * sizeof(unsigned short); }

int hb_kmeans_split_rows(const float* X, long long n, int d, float scale, int prec, void* Xs, void* stream_) {
  HB_REQUIRE(prec == 1 || prec == 2, "hb_kmeans_split_rows: prec must be 1 (fp16 halves) or 2 (bf16 halves)");
  HB_REQUIRE(d >= 4 && d % 4 == 0 && d <= BK * MAX_KCH, "hb_kmeans_split_rows: need d %% 4 == 0 and d <= 128 (got %d)", d);
  HB_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0 && (reinterpret_cast<uintptr_t>(Xs) & 15) == 0, "hb_kmeans_split_rows: pointers must be 16-byte aligned");
  if (n == 0) return 0;
  long long blocks = (n * (split_dp(d) / 4) + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  split_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream_>>>(X, n, d, scale, prec == 1 ? 0 : 1,
                                                                         reinterpret_cast<unsigned short*>(Xs));
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_kmeans_fused_prep(void* state, int k, int d, int parity, void* stream_);

// (Re)computes the operands of centroid buffer `parity` for precision mode `prec`.
int hb_kmeans_fused_prep2(void* state, int k, int d, int parity, int prec, float scale, void* stream_) {
  if (prec == 0) return hb_kmeans_fused_prep(state, k, d, parity, stream_);
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(state != nullptr && (parity == 0 || parity == 1) && (prec == 1 || prec == 2), "hb_kmeans_fused_prep2: bad arguments");
  const FusedLayout L = fused_layout(k, d);
  const int kpad = (k + BN - 1) / BN * BN;
  uint8_t* st = reinterpret_cast<uint8_t*>(state);
  const int threads = 256;
  int blocks = (kpad * 32 + threads - 1) / threads;
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  prep_split_kernel<<<blocks, threads, 0, stream>>>(reinterpret_cast<const float*>(st + L.c[parity]), k, d, kpad, scale,
                                                    prec == 1 ? 0 : 1, reinterpret_cast<unsigned short*>(st + L.cs[parity]),
                                                    reinterpret_cast<uint32_t*>(st + L.ext[parity]),
                                                    reinterpret_cast<uint32_t*>(st + L.ones));
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// (Re)computes the GEMM-extension operand of centroid buffer `parity` from its current contents.
int hb_kmeans_fused_prep(void* state, int k, int d, int parity, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(state != nullptr && (parity == 0 || parity == 1), "hb_kmeans_fused_prep: bad arguments");
  const FusedLayout L = fused_layout(k, d);
  const int kpad = (k + BN - 1) / BN * BN;
  uint8_t* st = reinterpret_cast<uint8_t*>(state);
  const int threads = 256;
  int blocks = (kpad * 32 + threads - 1) / threads;
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  prep_kernel<<<blocks, threads, 0, stream>>>(reinterpret_cast<const float*>(st + L.c[parity]), k, d, kpad, nullptr,
                                              reinterpret_cast<float*>(st + L.ext[parity]),
                                              reinterpret_cast<float*>(st + L.ones), nullptr, nullptr, nullptr, 0);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// Zeroes the state, loads the initial centroids into buffer 0 and prepares both operand buffers.  Every rank
// must have returned from this (and synchronised with its peers) before any rank launches the first step.
int hb_kmeans_fused_init(void* state, int k, int d, const float* C0, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(state != nullptr && k >= 1 && d >= 4 && d % 4 == 0 && hb_kmeans_tc_supported(d, k),
             "hb_kmeans_fused_init: shape (k=%d, d=%d) not supported by the fused kernel", k, d);
  HB_REQUIRE((reinterpret_cast<uintptr_t>(state) & 255) == 0, "hb_kmeans_fused_init: state must be 256-byte aligned");
  const FusedLayout L = fused_layout(k, d);
  HB_CUDA_OK(cudaMemsetAsync(state, 0, L.total, stream));
  if (C0 != nullptr)
    HB_CUDA_OK(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(state) + L.c[0], C0, (size_t)k * d * sizeof(float),
                               cudaMemcpyDeviceToDevice, stream));
  if (hb_kmeans_fused_prep(state, k, d, 1, stream_) != 0) return -1;   // padding rows of buffer 1
  return hb_kmeans_fused_prep(state, k, d, 0, stream_);
}

int hb_kmeans_fused_step2(const float* X, const void* Xs, long long n, int d, int k, void* const* peer_state, int rank,
                          int world, unsigned epoch, int keep_empty, int* assign, int prec, float scale, void* stream_);

int hb_kmeans_fused_step(const float* X, long long n, int d, int k, void* const* peer_state, int rank, int world,
                         unsigned epoch, int keep_empty, int* assign, void* stream_) {
  return hb_kmeans_fused_step2(X, nullptr, n, d, k, peer_state, rank, world, epoch, keep_empty, assign, 0, 1.0f, stream_);
}

// Xs: the points as split rows (hb_kmeans_split_rows) when prec != 0; X (fp32) is always needed: the partial sums and
// the objective are accumulated from the exact values.
int hb_kmeans_fused_step2(const float* X, const void* Xs, long long n, int d, int k, void* const* peer_state, int rank,
                          int world, unsigned epoch, int keep_empty, int* assign, int prec, float scale, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(prec >= 0 && prec <= 2, "hb_kmeans_fused_step2: prec must be 0 (tf32), 1 (fp16 halves) or 2 (bf16 halves)");
  HB_REQUIRE(prec == 0 || n == 0 || (Xs != nullptr && (reinterpret_cast<uintptr_t>(Xs) & 15) == 0), "hb_kmeans_fused_step2: split rows missing or misaligned");
  HB_REQUIRE(world >= 1 && world <= HB_MAX_RANKS && rank >= 0 && rank < world, "hb_kmeans_fused_step: bad rank %d / world %d", rank, world);
  HB_REQUIRE(n >= 0 && k >= 1 && hb_kmeans_tc_supported(d, k), "hb_kmeans_fused_step: shape (n=%lld, k=%d, d=%d) not supported", n, k, d);
  HB_REQUIRE(n == 0 || (X != nullptr && (reinterpret_cast<uintptr_t>(X) & 15) == 0), "hb_kmeans_fused_step: X must be 16-byte aligned");
  const FusedLayout L = fused_layout(k, d);
  const int kpad = (k + BN - 1) / BN * BN;
  const uint32_t cur = epoch & 1u;
  FusedArgs fa;
  memset(&fa, 0, sizeof(fa));
  for (int i = 0; i < world; ++i) {
    HB_REQUIRE(peer_state[i] != nullptr, "hb_kmeans_fused_step: null state pointer for rank %d", i);
    fa.peer[i] = reinterpret_cast<uint8_t*>(peer_state[i]);
  }
  fa.rank = rank; fa.world = world; fa.epoch = epoch; fa.keep_empty = keep_empty;
  fa.scale = prec == 1 ? scale : 1.0f; fa.fmt16 = prec == 2 ? 1 : 0;
  for (int i = 0; i < 2; ++i) { fa.off_ext[i] = L.ext[i]; fa.off_c[i] = L.c[i]; fa.off_acc[i] = L.acc[i]; fa.off_cs[i] = L.cs[i]; }
  uint8_t* me = fa.peer[rank];
  const float* C = reinterpret_cast<const float*>(me + L.c[cur]);
  float* sums = reinterpret_cast<float*>(me + L.acc[cur]);
  float* counts = sums + (size_t)k * d;
  double* obj = &reinterpret_cast<FusedCtl*>(me)->obj[cur];
  CUtensorMap tx, tch, tones, tcext;
  // a rank without points still takes part in the collective tail; its X map is never dereferenced
  const float* Xm = n > 0 ? X : C;
  const uint64_t nm = n > 0 ? (uint64_t)n : (uint64_t)k;
  if (prec == 0) {
    if (make_tmap_2d_f32(&tx, Xm, nm, (uint64_t)d, (uint64_t)d * 4, BM, BK, CU_TENSOR_MAP_SWIZZLE_128B) != 0) return -3;
    if (make_tmap_2d_f32(&tch, C, (uint64_t)k, (uint64_t)d, (uint64_t)d * 4, BN / 2, BK, CU_TENSOR_MAP_SWIZZLE_128B) != 0) return -3;
  } else {
    // split rows: 2 * dp 16-bit columns, boxes of 64 columns (128 bytes) x 128 rows
    const uint64_t cols = (uint64_t)2 * split_dp(d);
    const void* Xsm = n > 0 ? Xs : (const void*)(me + L.cs[cur]);
    if (make_tmap_2d_u16(&tx, Xsm, nm, cols, cols * 2, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B) != 0) return -3;
    if (make_tmap_2d_u16(&tch, me + L.cs[cur], (uint64_t)k, cols, cols * 2, BN / 2, 64, CU_TENSOR_MAP_SWIZZLE_128B) != 0) return -3;
  }
  if (make_tmap_2d_f32(&tones, me + L.ones, 128, 8, 32, 128, 8, CU_TENSOR_MAP_SWIZZLE_32B) != 0) return -3;
  if (make_tmap_2d_f32(&tcext, me + L.ext[cur], (uint64_t)kpad, 8, 32, BN / 2, 8, CU_TENSOR_MAP_SWIZZLE_32B) != 0) return -3;
  auto kern = prec == 0 ? assign_tc_pair_kernel<0, 1, 1, 0> : assign_tc_pair_kernel<0, 1, 1, 1>;
#ifdef HB_DEBUG_KERNELS
  if (const char* e = getenv("HB_FUSED_DBG")) { if (atoi(e) == 16) kern = assign_tc_pair_kernel<16, 1, 1>; else if (atoi(e) == 4) kern = assign_tc_pair_kernel<4, 1, 1>; }   // triage knob
#endif
  HB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pair::Smem::total));
  const long long npairs = (n + 2 * BM - 1) / (2 * BM);
  long long nclusters = sm_count() / 2;
  if (nclusters > npairs) nclusters = npairs;
  if (nclusters < 1) nclusters = 1;
  // The tail spins on counters fed by the other CTAs of this launch, so all of them must be resident at once:
  // one CTA per SM, never more CTAs than SMs, and a cooperative launch so the driver refuses otherwise.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * nclusters));
  cfg.blockDim = dim3(pair::NTHREADS2);
  cfg.dynamicSmemBytes = pair::Smem::total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  HB_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tx, tch, tones, tcext, Xm, C, (const float*)nullptr, n, d, k, kpad, sums, counts,
                                assign, obj, fa));
  return 0;
}

size_t hb_kmeans_simt_workspace_bytes(long long n, int k, int d) {
  return hb_kmeans_workspace_bytes(k, d) + (size_t)n * 8;
}

int hb_kmeans_finalize(const float* sums, const float* counts, int k, int d, float* C,
                       int keep_empty, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(k >= 1 && d >= 1, "hb_kmeans_finalize: bad shape");
  const size_t total = (size_t)k * d;
  int blocks = (int)((total + 255) / 256);
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  finalize_kernel<<<blocks, 256, 0, stream>>>(sums, counts, k, d, C, keep_empty);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ----------------------------------------------------------------------------
// hb_kmeans_step_static — the same assignment pass, for callers that cannot
// hand this library any scratch memory.
//
// Who that is: a host program that reaches this library through the bare C ABI
// from a runtime that owns no CUDA allocator of its own (a JVM worker through
// JNI/FFI, a Fortran or C service), or a process whose device memory has been
// claimed up front by someone else's pool (an inference server, PyTorch's
// caching allocator configured to take the whole GPU) so that a cudaMalloc from
// here would fail or, worse, synchronise the whole device in the middle of
// their pipeline.  hb_kmeans_step() asks such callers for a workspace they have
// no good way to produce.  This variant keeps the workspace inside the library
// image instead: a fixed array in device global memory that exists, per GPU,
// from the moment the library is loaded on that GPU until the process exits.
//
// Consequences the caller must know:
//   * The array is sized for at most HB_KMEANS_STATIC_MAX_K centroids (64 KiB
//     of ‖c‖² values plus padding); larger k is refused with an error, as is
//     any (d, k) the tensor-core path does not take, because the SIMT fallback
//     needs scratch proportional to the number of points.
//   * All calls on one GPU share that one array.  Each call rewrites the part
//     it uses before reading it (nothing in it has to start at zero and stale
//     contents of an earlier call are harmless), but two calls that overlap in
//     time on the same GPU — different streams, different host threads — race
//     on it.  Serialise them, or use hb_kmeans_step() with separate workspaces.
//   * The first call on each GPU resolves the array's device address
//     (cudaGetSymbolAddress) and sets the kernel's shared-memory attribute.
//     Make that first call during start-up, before any stream capture begins.
//   * Nothing in this path calls cudaMalloc, cudaFree or an allocating library
//     routine.
// ----------------------------------------------------------------------------
namespace hb {
namespace kmeans {

constexpr int STATIC_MAX_K = HB_KMEANS_STATIC_MAX_K;
constexpr size_t STATIC_SCRATCH_FLOATS = (size_t)STATIC_MAX_K + 64;  // ‖c‖² (multiple of 256) + 256 B pad

// Device-resident scratch.  In host code this name is only a handle for the
// runtime's symbol table, never an address a kernel may dereference.
static __device__ __align__(256) float g_static_scratch[STATIC_SCRATCH_FLOATS];

static std::mutex g_static_mu;
static float* g_static_ptr[64] = {nullptr};

// Returns the device address of g_static_scratch on the current device.
static int static_scratch_ptr(float** out) {
  int dev = 0;
  HB_CUDA_OK(cudaGetDevice(&dev));
  HB_REQUIRE(dev >= 0 && dev < 64, "hb_kmeans_step_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    HB_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_static_scratch));
    HB_REQUIRE(p != nullptr, "hb_kmeans_step_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return 0;
}

}  // namespace kmeans
}  // namespace hb

extern "C" int hb_kmeans_step_static(const float* X, long long n, int d, const float* C, int k,
                                     float* sums, float* counts, int* assign, double* obj,
                                     void* stream) {
  HB_REQUIRE(k >= 1 && k <= hb::kmeans::STATIC_MAX_K,
             "hb_kmeans_step_static: k=%d exceeds the built-in scratch (max %d centroids); "
             "use hb_kmeans_step with a caller-provided workspace",
             k, hb::kmeans::STATIC_MAX_K);
  HB_REQUIRE(hb_kmeans_tc_supported(d, k),
             "hb_kmeans_step_static: d=%d is not handled without caller scratch "
             "(needs d %% 4 == 0 and d <= 128); use hb_kmeans_step",
             d);
  // hb_kmeans_step would quietly take its SIMT branch for misaligned pointers and then complain about a workspace this
  // caller never supplied: say what is actually wrong, before anything is queued
  HB_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(sums) & 15) == 0,
             "hb_kmeans_step_static: X, C and sums must be 16-byte aligned (the tensor-core path is the only one "
             "that runs without caller scratch)");
  const size_t bytes = hb::kmeans::STATIC_SCRATCH_FLOATS * sizeof(float);
  HB_REQUIRE(hb_kmeans_min_workspace_bytes(k, d) <= bytes,
             "hb_kmeans_step_static: internal scratch too small for k=%d", k);
  float* scratch = nullptr;
  if (hb::kmeans::static_scratch_ptr(&scratch) != 0) return -1;
  // 64 KiB of scratch holds ‖c‖² only, which is what the single-CTA tcgen05 kernel needs
  return hb_kmeans_step(X, n, d, C, k, sums, counts, assign, obj, scratch, bytes, HB_KMEANS_SINGLE_CTA, stream);
}
