// K-means assignment step for B200 (sm_100a).
//
// Replaces the reference's scalar Java hot loop (CenCalcTask.run,
// ml/java/.../kmeans/regroupallgather/CenCalcTask.java:67-100, and the contrib
// variant KmeansMapper.java:173-215) and DAAL's kmeans.DistributedStep1Local.
//
// Design (B200-first, not a translation):
//   * -2·X·Cᵀ runs on the 5th-gen tensor cores: tcgen05.mma kind::tf32 reading
//     fp32 tiles straight out of shared memory (TMA, SWIZZLE_128B), with the
//     128x256 fp32 accumulator double-buffered in TMEM.
//   * A persistent, warp-specialised CTA per SM: warp 0 = TMA producer,
//     warp 1 = single-thread MMA issuer, warps 2..5 = epilogue.  The point tile
//     (128 x D) stays resident in smem while all centroid tiles stream past it
//     out of L2.
//   * The epilogue pulls distances out of TMEM (tcgen05.ld), adds ‖c‖² from a
//     small smem ring, keeps the running arg-min, and then folds each point
//     into sums[argmin]/counts[argmin] with 16-byte vector reductions
//     (red.global.add.v4.f32) that resolve in L2.  Nothing but the final
//     [k x d] sums ever goes back to HBM.
//   * A plain fp32 SIMT kernel is kept as the exact-arithmetic fallback for
//     shapes the tensor path does not take (d % 4 != 0 or d > 128).
#include "common.cuh"
#include "../include/harp_b200.h"

#include <math_constants.h>

#include <mutex>

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
__global__ void prep_kernel(const float* __restrict__ C, int k, int d, int kpad,
                            float* __restrict__ cn, float* __restrict__ sums,
                            float* __restrict__ counts, double* __restrict__ obj, int zero_acc) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int j = warp; j < kpad; j += nwarps) {
    float s = CUDART_INF_F;
    if (j < k) {
      s = 0.f;
      for (int t = lane; t < d; t += 32) {
        float v = C[(size_t)j * d + t];
        s = fmaf(v, v, s);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    if (lane == 0) cn[j] = s;
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
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_tf32(BM, BN);
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
            const uint32_t a_addr = base + SmemLayout::x + kc * X_CHUNK_BYTES;
            const uint32_t b_addr = base + SmemLayout::b + s * B_STAGE_BYTES;
            int ksteps = (d - kc * BK + UK - 1) / UK;
            if (ksteps > BK / UK) ksteps = BK / UK;
            for (int j = 0; j < ksteps; ++j) {
              const uint64_t da = umma_desc_kmajor_sw128(a_addr + j * UK * 4);
              const uint64_t db = umma_desc_kmajor_sw128(b_addr + j * UK * 4);
              umma_tf32_1sm(tmem_d, da, db, idesc, (kc | j) != 0 ? 1u : 0u);
            }
            umma_commit_1sm(bar(BAR_B_EMPTY + s));      // frees the smem stage when the MMAs retire
          }
          umma_commit_1sm(bar(BAR_T_FULL + a));          // accumulator ready for the epilogue
        }
        umma_commit_1sm(bar(BAR_X_EMPTY));               // point tile no longer read by the tensor core
      }
    }
  } else {
    // ===================== Epilogue warps (4 x 32 rows) =====================
    const int q = warp & 3;                              // TMEM lane quadrant of this warp
    const int nl4 = d >> 2;                              // float4s per row
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
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
          const int cbase = ct * BN + c * 32;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 cnv = cn4[c * 8 + (j >> 2)];
            const float d0 = fmaf(-2.f, __uint_as_float(v[j + 0]), cnv.x);
            const float d1 = fmaf(-2.f, __uint_as_float(v[j + 1]), cnv.y);
            const float d2 = fmaf(-2.f, __uint_as_float(v[j + 2]), cnv.z);
            const float d3 = fmaf(-2.f, __uint_as_float(v[j + 3]), cnv.w);
            if (d0 < best) { best = d0; bi = cbase + j + 0; }
            if (d1 < best) { best = d1; bi = cbase + j + 1; }
            if (d2 < best) { best = d2; bi = cbase + j + 2; }
            if (d3 < best) { best = d3; bi = cbase + j + 3; }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_T_EMPTY + a));
      }
      // ---- fold the 32 points of this warp into sums / counts (L2 reductions) ----
      const long long row0 = tile * BM + q * 32;
      if (assign != nullptr && row0 + lane < n) assign[row0 + lane] = bi;
      float sq = 0.f;
#pragma unroll 4
      for (int i = 0; i < 32; ++i) {
        const long long row = row0 + i;
        const int idx = __shfl_sync(0xffffffffu, bi, i);
        if (row < n) {
          if (lane < nl4) {
            const float4 xv = ld_nc_v4(X + (size_t)row * d + lane * 4);
            red_add_v4(sums + (size_t)idx * d + lane * 4, xv);
            if (obj != nullptr) {
              const float4 cv = *reinterpret_cast<const float4*>(C + (size_t)idx * d + lane * 4);
              const float e0 = xv.x - cv.x, e1 = xv.y - cv.y, e2 = xv.z - cv.z, e3 = xv.w - cv.w;
              sq = fmaf(e0, e0, sq); sq = fmaf(e1, e1, sq); sq = fmaf(e2, e2, sq); sq = fmaf(e3, e3, sq);
            }
          }
          if (lane == 0) atomicAdd(counts + idx, 1.0f);
        }
      }
      obj_local += (double)sq;
    }
    if (obj != nullptr) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) obj_local += __shfl_xor_sync(0xffffffffu, obj_local, o);
      if (lane == 0) atomicAdd(obj, obj_local);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_1sm(tmem_base, TMEM_COLS);
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

size_t hb_kmeans_workspace_bytes(int k, int d) {
  const size_t kpad = ((size_t)k + BN - 1) / BN * BN;
  (void)d;
  return kpad * sizeof(float) + 256;
}

int hb_kmeans_tc_supported(int d, int k) { return (d % 4 == 0 && d >= 4 && d <= BK * MAX_KCH && k >= 1) ? 1 : 0; }

int hb_kmeans_step(const float* X, long long n, int d, const float* C, int k, float* sums,
                   float* counts, int* assign, double* obj, void* workspace, size_t workspace_bytes,
                   int flags, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  HB_REQUIRE(n >= 0 && d >= 1 && k >= 1, "hb_kmeans_step: bad shape n=%lld d=%d k=%d", n, d, k);
  const int kpad = (k + BN - 1) / BN * BN;
  HB_REQUIRE(workspace != nullptr && workspace_bytes >= hb_kmeans_workspace_bytes(k, d),
             "hb_kmeans_step: workspace too small (%zu < %zu)", workspace_bytes,
             hb_kmeans_workspace_bytes(k, d));
  HB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
             "hb_kmeans_step: workspace must be 256-byte aligned");
  float* cn = reinterpret_cast<float*>(workspace);
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
    prep_kernel<<<blocks, threads, 0, stream>>>(C, k, d, kpad, cn, sums, counts, obj, zero_acc ? 1 : 0);
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
    static bool attr_set[64] = {false};
    int dev = 0;
    HB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev]) {
      HB_CUDA_OK(cudaFuncSetAttribute(assign_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      SMEM_BYTES));
      attr_set[dev] = true;
    }
    const long long ntiles = (n + BM - 1) / BM;
    int grid = sm_count();
    if ((long long)grid > ntiles) grid = (int)ntiles;
    assign_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, stream>>>(tx, tc, X, C, cn, n, d, k, kpad, sums,
                                                             counts, assign, obj);
    HB_CUDA_OK(cudaGetLastError());
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
  const size_t bytes = hb::kmeans::STATIC_SCRATCH_FLOATS * sizeof(float);
  HB_REQUIRE(hb_kmeans_workspace_bytes(k, d) <= bytes,
             "hb_kmeans_step_static: internal scratch too small for k=%d", k);
  float* scratch = nullptr;
  if (hb::kmeans::static_scratch_ptr(&scratch) != 0) return -1;
  return hb_kmeans_step(X, n, d, C, k, sums, counts, assign, obj, scratch, bytes, 0, stream);
}
