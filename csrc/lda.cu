// Collapsed Gibbs sampling for LDA.
//
// Replaces LDAMPTask.doRun / sample (ml/java/src/main/java/edu/iu/lda/LDAMPTask.java:85-331), which
// is SparseLDA (s/r/t buckets over sorted sparse rows) run by one Java thread per (doc-split,
// word-split) block under the dymoro Scheduler's row/column exclusivity, with the global topic
// sums frozen for the duration of an iteration (LDAMPCollectiveMapper.java:464-475).
//
// B200 design: tokens of one panel (this worker's documents x the visiting word-topic slice) are
// bucketed B x B and sorted by word inside a block, exactly like the MF-SGD panels; ONE persistent
// cooperative kernel walks the blocks in Latin-square order with ticket hand-over of word-blocks,
// so no two CTAs ever sample for the same documents or the same words concurrently.  A warp owns
// one word run: phi_k = (n_wk + beta) / (n_k + V beta) for its 32 interleaved topic lanes lives
// in registers and is patched by +-1/(n_k + V beta) as counts move; per token the warp reads the
// document's dense topic row with coalesced 128-byte loads, forms p_k = (n_dk + alpha) phi_k,
// scans, draws, and moves one count.  K <= 2048 topics (dense rows: at K ~ 1000 the dense product
// costs one 4 KB row read per token, which is what bounds the kernel; the sparse bucket walk of
// the CPU code would diverge across lanes).  Count updates are integer atomics, so the tables stay
// exactly consistent with the assignments even where two warps of a CTA meet on a document.
#include "common.cuh"
#include "../include/harp_b200.h"

namespace hb {
namespace lda {

constexpr int LDA_THREADS = 512;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ int ldcg_i32(const int* p) {
  int r;
  asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
// counter-based RNG: one 64-bit mix per draw
__device__ __forceinline__ float uniform01(unsigned long long seed, unsigned long long ctr) {
  unsigned long long x = seed + ctr * 0x9E3779B97F4A7C15ULL;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
  x ^= x >> 27; x *= 0x94D049BB133111EBULL;
  x ^= x >> 31;
  return (float)(x >> 40) * (1.0f / 16777216.0f);      // 24 random bits -> [0, 1)
}

struct LdaArgs {
  int* n_dk;                 // [docs_local, K]   document-topic counts
  int* n_wk;                 // [words_slice, K]  word-topic counts of the visiting slice
  const float* commons;      // [K]               1 / (n_k + V beta), frozen for the iteration
  const int* blk_seg_ptr;    // [B*B + 1]
  const int* seg_word;       // [nseg]            slice-local word id (bit 31: run was split)
  const int* seg_ptr;        // [nseg + 1]
  const int* tok_doc;        // [ntok]            worker-local document id
  int* tok_z;                // [ntok]            topic assignment (updated in place)
  unsigned int* col_ticket;  // [B]
  unsigned long long* trained;
  unsigned long long deadline_ns;
  unsigned long long seed;
  unsigned int ticket_base;
  int B, K;
  float alpha, beta;
};

template <int KJ>   // topics per lane: K <= 32 * KJ
__global__ void __launch_bounds__(LDA_THREADS, 1) lda_pass_kernel(LdaArgs a) {
  __shared__ int s_next;
  __shared__ int s_stop;
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x, B = a.B, K = a.K;
  unsigned long long local_trained = 0;
  const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
  __shared__ float s_cm[32 * KJ];     // commons: 1 / (n_k + V beta)
  for (int k = threadIdx.x; k < 32 * KJ; k += blockDim.x) s_cm[k] = (k < K) ? a.commons[k] : 0.f;
  __syncthreads();

  for (int t = 0; t < B; ++t) {
    const int c = (i + t) % B;
    if (threadIdx.x == 0) {
      const unsigned int want = a.ticket_base + (unsigned)t;
      while (ld_acquire_gpu(a.col_ticket + c) != want) { __nanosleep(64); }
      s_next = a.blk_seg_ptr[(size_t)i * B + c];
      s_stop = (t_end != 0 && globaltimer_ns() > t_end) ? 1 : 0;
    }
    __syncthreads();
    const bool expired = (s_stop != 0);
    const int s_end = a.blk_seg_ptr[(size_t)i * B + c + 1];
    if (!expired) {
      for (;;) {
        int seg = 0;
        if (lane == 0) seg = atomicAdd(&s_next, 1);
        seg = __shfl_sync(0xffffffffu, seg, 0);
        if (seg >= s_end) break;
        const int word = a.seg_word[seg] & 0x7FFFFFFF;
        const int k0 = a.seg_ptr[seg], k1 = a.seg_ptr[seg + 1];
        int* wrow = a.n_wk + (size_t)word * K;
        float phi[KJ];
#pragma unroll
        for (int j = 0; j < KJ; ++j) {
          const int k = lane + 32 * j;
          phi[j] = (k < K) ? ((float)ldcg_i32(wrow + k) + a.beta) * s_cm[k] : 0.f;
        }
        for (int tk = k0; tk < k1; ++tk) {
          const int d = a.tok_doc[tk];
          const int zold = a.tok_z[tk];
          int* drow = a.n_dk + (size_t)d * K;
          // take the token out of the distribution (registers only; the tables move once, below)
          {
            const int jo = zold >> 5;
            if ((zold & 31) == lane) {
              const float cz = s_cm[zold];
#pragma unroll
              for (int j = 0; j < KJ; ++j) if (j == jo) phi[j] -= cz;
            }
          }
          // p_k = (n_dk + alpha) * phi_k over this lane's topics
          float p[KJ];
          float sum = 0.f;
#pragma unroll
          for (int j = 0; j < KJ; ++j) {
            const int k = lane + 32 * j;
            int nd = (k < K) ? ldcg_i32(drow + k) : 0;
            if (k == zold) nd = max(nd - 1, 0);        // this token itself
            p[j] = (k < K) ? ((float)nd + a.alpha) * phi[j] : 0.f;
            sum += p[j];
          }
          float cum = sum;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const float v = __shfl_up_sync(0xffffffffu, cum, o);
            if (lane >= o) cum += v;
          }
          const float total = __shfl_sync(0xffffffffu, cum, 31);
          const float u = uniform01(a.seed, (unsigned long long)tk * 2654435761ULL + (unsigned long long)i) * total;
          const unsigned int ball = __ballot_sync(0xffffffffu, cum > u);
          const int L = ball ? (__ffs(ball) - 1) : 31;
          int znew = zold;
          if (lane == L) {
            float acc = cum - sum;
            znew = -1;
#pragma unroll
            for (int j = 0; j < KJ; ++j) {
              acc += p[j];
              if (znew < 0 && acc > u && lane + 32 * j < K) znew = lane + 32 * j;
            }
            if (znew < 0) {          // rounding at the top of the range: last valid topic of this lane
#pragma unroll
              for (int j = KJ - 1; j >= 0; --j) if (znew < 0 && lane + 32 * j < K && p[j] > 0.f) znew = lane + 32 * j;
              if (znew < 0) znew = zold;
            }
          }
          znew = __shfl_sync(0xffffffffu, znew, L);
          // put the token back with its new topic; the count tables only move when the topic changed
          {
            const int jn = znew >> 5;
            if ((znew & 31) == lane) {
              const float cz = s_cm[znew];
#pragma unroll
              for (int j = 0; j < KJ; ++j) if (j == jn) phi[j] += cz;
              if (znew != zold) { atomicAdd(wrow + znew, 1); atomicAdd(drow + znew, 1); }
            }
            if ((zold & 31) == lane && znew != zold) { atomicSub(wrow + zold, 1); atomicSub(drow + zold, 1); }
          }
          if (lane == 0) a.tok_z[tk] = znew;
        }
        if (lane == 0) local_trained += (unsigned long long)(k1 - k0);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.col_ticket + c), "r"(a.ticket_base + (unsigned)t + 1u) : "memory");
    }
  }
  if (lane == 0 && local_trained) atomicAdd(a.trained, local_trained);
}

// topic_sums[k] (+)= Σ_w n_wk[w][k]   (LDAMPCollectiveMapper.java:442-462)
__global__ void topic_sum_kernel(const int* __restrict__ n_wk, long long words, int K, long long* __restrict__ sums) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  long long s = 0;
  for (long long w = blockIdx.y; w < words; w += gridDim.y) s += n_wk[(size_t)w * K + k];
  if (s) atomicAdd(reinterpret_cast<unsigned long long*>(sums + k), (unsigned long long)s);
}

// counts[(idx0[t]) * K + z[t]] += 1 for all tokens (table initialisation from assignments)
__global__ void count_kernel(const int* __restrict__ idx0, const int* __restrict__ z, long long n, int K, int* __restrict__ counts) {
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x)
    atomicAdd(counts + (size_t)idx0[t] * K + z[t], 1);
}

}  // namespace lda
}  // namespace hb

using namespace hb;
using namespace hb::lda;

extern "C" {

int hb_lda_max_blocks(int K) {
  int dev = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  const void* fn = K <= 256 ? (const void*)lda_pass_kernel<8> : K <= 512 ? (const void*)lda_pass_kernel<16>
                 : K <= 1024 ? (const void*)lda_pass_kernel<32> : (const void*)lda_pass_kernel<64>;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, LDA_THREADS, 0) != cudaSuccess) return 0;
  return per_sm * sm_count();
}

int hb_lda_pass(int* n_dk, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, float alpha, float beta,
                unsigned long long seed, unsigned int* col_ticket, unsigned int ticket_base,
                unsigned long long deadline_ns, unsigned long long* trained, void* stream) {
  HB_REQUIRE(B >= 1 && K >= 2 && K <= 2048, "hb_lda_pass: need 2 <= K <= 2048 topics (got %d)", K);
  HB_REQUIRE(B <= hb_lda_max_blocks(K), "hb_lda_pass: B=%d blocks cannot be co-resident (max %d)", B, hb_lda_max_blocks(K));
  LdaArgs a;
  a.n_dk = n_dk; a.n_wk = n_wk; a.commons = commons; a.blk_seg_ptr = blk_seg_ptr; a.seg_word = seg_word;
  a.seg_ptr = seg_ptr; a.tok_doc = tok_doc; a.tok_z = tok_z; a.col_ticket = col_ticket; a.trained = trained;
  a.deadline_ns = deadline_ns; a.seed = seed; a.ticket_base = ticket_base; a.B = B; a.K = K; a.alpha = alpha; a.beta = beta;
  void* args[] = {&a};
  const void* fn = K <= 256 ? (const void*)lda_pass_kernel<8> : K <= 512 ? (const void*)lda_pass_kernel<16>
                 : K <= 1024 ? (const void*)lda_pass_kernel<32> : (const void*)lda_pass_kernel<64>;
  HB_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(B), dim3(LDA_THREADS), args, 0, (cudaStream_t)stream));
  return 0;
}

int hb_lda_topic_sums(const int* n_wk, long long words, int K, long long* sums, void* stream) {
  if (words == 0) return 0;
  int gy = (int)(words < 256 ? words : 256);
  topic_sum_kernel<<<dim3((K + 127) / 128, gy), 128, 0, (cudaStream_t)stream>>>(n_wk, words, K, sums);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_lda_count(const int* idx0, const int* z, long long n, int K, int* counts, void* stream) {
  if (n == 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  count_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(idx0, z, n, K, counts);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
