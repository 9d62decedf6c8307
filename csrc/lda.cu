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
#include "rotate_ring.cuh"
#include "../include/harp_b200.h"

#include <string.h>

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

// Branch-free loads of a lane's topic pair (k, k+1) of a count row with K entries.  Rows start at row*K elements,
// so the pair is naturally aligned for ONE wide load only when K is even (`keven`); odd K takes two narrow loads.
// Everything is predicated, nothing branches: the KP loads of a row must issue back to back, a branch per pair
// would put one memory round trip per pair on the token's critical path.
__device__ __forceinline__ void ld_pair_i32(const int* row, int k, int K, bool keven, int& v0, int& v1) {
  const unsigned pv = (keven && k < K) ? 1u : 0u, p0 = (!keven && k < K) ? 1u : 0u, p1 = (!keven && k + 1 < K) ? 1u : 0u;
  asm volatile(
      "{\n"
      ".reg .pred pv, p0, p1;\n"
      "setp.ne.u32 pv, %3, 0;\n"
      "setp.ne.u32 p0, %4, 0;\n"
      "setp.ne.u32 p1, %5, 0;\n"
      "mov.b32 %0, 0;\n"
      "mov.b32 %1, 0;\n"
      "@pv ld.global.cg.v2.s32 {%0, %1}, [%2];\n"
      "@p0 ld.global.cg.s32 %0, [%2];\n"
      "@p1 ld.global.cg.s32 %1, [%2+4];\n"
      "}\n"
      : "=&r"(v0), "=&r"(v1)
      : "l"(row + k), "r"(pv), "r"(p0), "r"(p1));
}
// 16-bit counts: returns lo | hi << 16
__device__ __forceinline__ unsigned ld_pair_u16(const unsigned short* row, int k, int K, bool keven) {
  const unsigned pv = (keven && k < K) ? 1u : 0u, p0 = (!keven && k < K) ? 1u : 0u, p1 = (!keven && k + 1 < K) ? 1u : 0u;
  unsigned v, lo, hi;
  asm volatile(
      "{\n"
      ".reg .pred pv, p0, p1;\n"
      "setp.ne.u32 pv, %4, 0;\n"
      "setp.ne.u32 p0, %5, 0;\n"
      "setp.ne.u32 p1, %6, 0;\n"
      "mov.b32 %0, 0;\n"
      "mov.b32 %1, 0;\n"
      "mov.b32 %2, 0;\n"
      "@pv ld.global.cg.b32 %0, [%3];\n"
      "@p0 ld.global.cg.u16 %1, [%3];\n"
      "@p1 ld.global.cg.u16 %2, [%3+2];\n"
      "}\n"
      : "=&r"(v), "=&r"(lo), "=&r"(hi)
      : "l"(row + k), "r"(pv), "r"(p0), "r"(p1));
  return v | lo | (hi << 16);
}

// Lane <-> topic mapping: lane l owns the topic PAIRS m = l + 32 jp (topics 2m, 2m+1), so that one 64-bit load
// fetches a lane's two word-topic counts and, with 16-bit document-topic counts (DK16), one 32-bit load fetches
// its two document-topic counts: a K = 1000 document row is 2 KB instead of 4 KB, and the kernel is bound by
// exactly that row traffic (the table does not fit in L2).
__device__ __forceinline__ int topic_owner_lane(int z) { return (z >> 1) & 31; }
__device__ __forceinline__ int topic_slot(int z) { return ((z >> 6) << 1) | (z & 1); }   // index into phi[]

template <int KJ, bool DK16>   // topics per lane: K <= 32 * KJ, KJ even
__global__ void __launch_bounds__(LDA_THREADS, 1) lda_pass_kernel(LdaArgs a) {
  constexpr int KP = KJ / 2;          // topic pairs per lane
  __shared__ int s_next;
  __shared__ int s_stop;
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x, B = a.B, K = a.K;
  const bool keven = (K & 1) == 0;
  unsigned long long local_trained = 0;
  const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
  __shared__ float s_cm[32 * KJ];     // commons: 1 / (n_k + V beta)
  for (int k = threadIdx.x; k < 32 * KJ; k += blockDim.x) s_cm[k] = (k < K) ? a.commons[k] : 0.f;
  __syncthreads();

  auto load_doc_pair16 = [&](const unsigned short* drow, int jp) -> unsigned {
    return ld_pair_u16(drow, 2 * lane + 64 * jp, K, keven);
  };

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
        for (int jp = 0; jp < KP; ++jp) {
          const int k = 2 * lane + 64 * jp;
          int w0, w1;
          ld_pair_i32(wrow, k, K, keven, w0, w1);
          phi[2 * jp] = (k < K) ? ((float)w0 + a.beta) * s_cm[k] : 0.f;
          phi[2 * jp + 1] = (k + 1 < K) ? ((float)w1 + a.beta) * s_cm[k + 1] : 0.f;
        }
        for (int kb = k0; kb < k1; kb += 32) {
          const int nb = min(32, k1 - kb);
          // documents and current topics of up to 32 tokens of the run, one per lane
          const int docs_l = (lane < nb) ? a.tok_doc[kb + lane] : 0;
          const int zs_l = (lane < nb) ? a.tok_z[kb + lane] : 0;
          int znew_l = zs_l;
          // DK16: the next token's document row is fetched while the current token is sampled
          unsigned nxt[DK16 ? KP : 1];
          if (DK16) {
            const int d0 = __shfl_sync(0xffffffffu, docs_l, 0);
            const unsigned short* drow0 = reinterpret_cast<const unsigned short*>(a.n_dk) + (size_t)d0 * K;
#pragma unroll
            for (int jp = 0; jp < KP; ++jp) nxt[jp] = load_doc_pair16(drow0, jp);
          }
          for (int j = 0; j < nb; ++j) {
            const int d = __shfl_sync(0xffffffffu, docs_l, j);
            const int zold = __shfl_sync(0xffffffffu, zs_l, j);
            const int tk = kb + j;
            // take the token out of the distribution (registers only; the tables move once, below)
            if (topic_owner_lane(zold) == lane) {
              const float cz = s_cm[zold];
              const int so = topic_slot(zold);
#pragma unroll
              for (int q = 0; q < KJ; ++q) if (q == so) phi[q] -= cz;
            }
            // p_k = (n_dk + alpha) * phi_k over this lane's topics
            float p[KJ];
            float sum = 0.f;
            if (DK16) {
              unsigned cur[KP];
#pragma unroll
              for (int jp = 0; jp < KP; ++jp) cur[jp] = nxt[jp];
              if (j + 1 < nb) {
                const int dn = __shfl_sync(0xffffffffu, docs_l, j + 1);
                const unsigned short* drown = reinterpret_cast<const unsigned short*>(a.n_dk) + (size_t)dn * K;
#pragma unroll
                for (int jp = 0; jp < KP; ++jp) nxt[jp] = load_doc_pair16(drown, jp);
              }
#pragma unroll
              for (int jp = 0; jp < KP; ++jp) {
                const int k = 2 * lane + 64 * jp;
                int n0 = (int)(cur[jp] & 0xFFFFu), n1 = (int)(cur[jp] >> 16);
                if (k == zold) n0 = max(n0 - 1, 0);          // this token itself
                if (k + 1 == zold) n1 = max(n1 - 1, 0);
                p[2 * jp] = (k < K) ? ((float)n0 + a.alpha) * phi[2 * jp] : 0.f;
                p[2 * jp + 1] = (k + 1 < K) ? ((float)n1 + a.alpha) * phi[2 * jp + 1] : 0.f;
                sum += p[2 * jp] + p[2 * jp + 1];
              }
            } else {
              const int* drow = a.n_dk + (size_t)d * K;
#pragma unroll
              for (int jp = 0; jp < KP; ++jp) {
                const int k = 2 * lane + 64 * jp;
                int n0, n1;
                ld_pair_i32(drow, k, K, keven, n0, n1);
                if (k == zold) n0 = max(n0 - 1, 0);
                if (k + 1 == zold) n1 = max(n1 - 1, 0);
                p[2 * jp] = (k < K) ? ((float)n0 + a.alpha) * phi[2 * jp] : 0.f;
                p[2 * jp + 1] = (k + 1 < K) ? ((float)n1 + a.alpha) * phi[2 * jp + 1] : 0.f;
                sum += p[2 * jp] + p[2 * jp + 1];
              }
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
              for (int q = 0; q < KJ; ++q) {
                const int k = 2 * lane + 64 * (q >> 1) + (q & 1);
                acc += p[q];
                if (znew < 0 && acc > u && k < K) znew = k;
              }
              if (znew < 0) {          // rounding at the top of the range: last valid topic of this lane
#pragma unroll
                for (int q = KJ - 1; q >= 0; --q) {
                  const int k = 2 * lane + 64 * (q >> 1) + (q & 1);
                  if (znew < 0 && k < K && p[q] > 0.f) znew = k;
                }
                if (znew < 0) znew = zold;
              }
            }
            znew = __shfl_sync(0xffffffffu, znew, L);
            // put the token back with its new topic; the count tables only move when the topic changed
            if (topic_owner_lane(znew) == lane) {
              const float cz = s_cm[znew];
              const int sn = topic_slot(znew);
#pragma unroll
              for (int q = 0; q < KJ; ++q) if (q == sn) phi[q] += cz;
            }
            if (znew != zold) {
              if (lane == 0) {
                atomicAdd(wrow + znew, 1);
                atomicSub(wrow + zold, 1);
                if (DK16) {
                  int* d32 = reinterpret_cast<int*>(reinterpret_cast<unsigned short*>(a.n_dk) + (size_t)d * K);
                  if (keven) {
                    atomicAdd(d32 + (znew >> 1), (znew & 1) ? 65536 : 1);
                    atomicSub(d32 + (zold >> 1), (zold & 1) ? 65536 : 1);
                  } else {
                    // rows are only 2-byte aligned when K is odd: address the containing aligned word
                    unsigned short* e_new = reinterpret_cast<unsigned short*>(a.n_dk) + (size_t)d * K + znew;
                    unsigned short* e_old = reinterpret_cast<unsigned short*>(a.n_dk) + (size_t)d * K + zold;
                    const uintptr_t an = reinterpret_cast<uintptr_t>(e_new), ao = reinterpret_cast<uintptr_t>(e_old);
                    atomicAdd(reinterpret_cast<int*>(an & ~(uintptr_t)3), (an & 2) ? 65536 : 1);
                    atomicSub(reinterpret_cast<int*>(ao & ~(uintptr_t)3), (ao & 2) ? 65536 : 1);
                  }
                } else {
                  int* drow = a.n_dk + (size_t)d * K;
                  atomicAdd(drow + znew, 1);
                  atomicSub(drow + zold, 1);
                }
              }
              if (DK16 && j + 1 < nb) {
                // the row fetched ahead is stale if the next token is in the same document: patch it
                const int dn = __shfl_sync(0xffffffffu, docs_l, j + 1);
                if (dn == d) {
                  if (topic_owner_lane(znew) == lane) {
                    const int jn = znew >> 6;
#pragma unroll
                    for (int jp = 0; jp < KP; ++jp) if (jp == jn) nxt[jp] += (znew & 1) ? 65536u : 1u;
                  }
                  if (topic_owner_lane(zold) == lane) {
                    const int jo = zold >> 6;
#pragma unroll
                    for (int jp = 0; jp < KP; ++jp) if (jp == jo) nxt[jp] -= (zold & 1) ? 65536u : 1u;
                  }
                }
              }
            }
            if (lane == j) znew_l = znew;
          }
          if (lane < nb) a.tok_z[kb + lane] = znew_l;
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

// ----------------------------------------------------------------------------
// Shared-memory variant of the sampling pass (the default whenever a document row is a whole number of 16-byte
// chunks: K % 8 == 0 with 16-bit document counts, K % 4 == 0 with 32-bit ones, K <= 1024).
//
// The register kernel above spends ~1100 warp instructions per token, most of them on emulating "element z of
// this warp's phi / document row" with predicated selects over 32 registers.  Here both live in shared memory:
//   * phi_k of the run's word: one 4 KB array per warp, patched by one lane as counts move;
//   * the document-topic row: copied global -> shared with cp.async (16-byte chunks, no registers), double
//     buffered so the next token's row is in flight while the current token is sampled.
// Per token a lane then reads its 16 topic pairs with one 64-bit (phi) and one 32/64-bit (counts) LDS each,
// keeps the running prefix in registers and finds the drawn topic with a compare mask.
// ----------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// warps per CTA: as many as the per-warp buffers (phi + two row buffers) leave room for in 227 KB
template <int KJ, bool DK16> constexpr int lda_smem_warps() {
  return (KJ * 32 * (4 + (DK16 ? 4 : 8)) * 24 + KJ * 32 * 4 <= 200 * 1024) ? 24 : 16;
}

// One word run of a cell, sampled by one warp with its phi vector and two document-row buffers in shared memory.
// `a` provides n_dk, n_wk, seg_word, seg_ptr, tok_doc, tok_z, seed, alpha, beta; `salt` decorrelates the random
// streams of different row-blocks.  Returns the number of tokens of the run.
template <int KJ, bool DK16, typename Core>
__device__ __forceinline__ unsigned long long lda_run_segment_smem(const Core& a, int seg, int lane, int K, int salt,
                                                                    float* phi, uint8_t* rowbuf, const float* s_cm) {
  __builtin_assume(__isGlobal(a.seg_word)); __builtin_assume(__isGlobal(a.seg_ptr)); __builtin_assume(__isGlobal(a.tok_doc));
  __builtin_assume(__isGlobal(a.tok_z)); __builtin_assume(__isGlobal(a.n_dk)); __builtin_assume(__isGlobal(a.n_wk));
  constexpr int KS = 32 * KJ;
  constexpr int KP = KJ / 2;
  constexpr int ROWB = DK16 ? KS * 2 : KS * 4;
  const uint32_t rowbuf_s = smem_u32(rowbuf);
  const int row_bytes = DK16 ? K * 2 : K * 4;
  const int nchunks = row_bytes >> 4;
  const int i = salt;
  auto issue_row = [&](int d, int buf) {
    const uint8_t* src = reinterpret_cast<const uint8_t*>(a.n_dk) + (size_t)d * row_bytes;
    for (int ch = lane; ch < nchunks; ch += 32) cp_async16(rowbuf_s + buf * ROWB + ch * 16, src + ch * 16);
    cp_async_commit();
  };
  auto row_add = [&](int buf, int z, int delta) {    // one lane: counts[z] += delta in the shared copy, clamped at 0
    if (DK16) {
      unsigned short* e = reinterpret_cast<unsigned short*>(rowbuf + buf * ROWB) + z;
      const int v = (int)*e + delta;
      *e = (unsigned short)max(v, 0);
    } else {
      int* e = reinterpret_cast<int*>(rowbuf + buf * ROWB) + z;
      *e = max(*e + delta, 0);
    }
  };
  const int word = a.seg_word[seg] & 0x7FFFFFFF;
  const int k0 = a.seg_ptr[seg], k1 = a.seg_ptr[seg + 1];
  int* wrow = a.n_wk + (size_t)word * K;
  __syncwarp();
#pragma unroll 4
  for (int k = lane; k < KS; k += 32) phi[k] = (k < K) ? ((float)ldcg_i32(wrow + k) + a.beta) * s_cm[k] : 0.f;
  __syncwarp();
  for (int kb = k0; kb < k1; kb += 32) {
    const int nb = min(32, k1 - kb);
    const int docs_l = (lane < nb) ? a.tok_doc[kb + lane] : 0;
    const int zs_l = (lane < nb) ? a.tok_z[kb + lane] : 0;
    int znew_l = zs_l;
    int pend_doc = -1, pend_new = 0, pend_old = 0;      // table moves the row fetched ahead may have missed
    issue_row(__shfl_sync(0xffffffffu, docs_l, 0), 0);
    for (int j = 0; j < nb; ++j) {
      const int d = __shfl_sync(0xffffffffu, docs_l, j);
      const int zold = __shfl_sync(0xffffffffu, zs_l, j);
      const int tk = kb + j;
      const int buf = j & 1;
      if (j + 1 < nb) {
        issue_row(__shfl_sync(0xffffffffu, docs_l, j + 1), buf ^ 1);
        cp_async_wait<1>();
      } else {
        cp_async_wait<0>();
      }
      __syncwarp();
      // take the token out of both factors (shared copies only; the tables move once, below)
      if (lane == 0) {
        if (pend_doc == d) { row_add(buf, pend_new, 1); row_add(buf, pend_old, -1); }
        phi[zold] -= s_cm[zold];
        row_add(buf, zold, -1);
      }
      __syncwarp();
      // p_k = (n_dk + alpha) * phi_k, running prefix over this lane's topics (2m, 2m+1), m = lane + 32 jp
      float cpre[KJ];
      float run = 0.f;
      const float2* phi2 = reinterpret_cast<const float2*>(phi);
#pragma unroll
      for (int jp = 0; jp < KP; ++jp) {
        const int m = lane + 32 * jp;
        const float2 ph = phi2[m];
        float n0, n1;
        if (DK16) {
          const unsigned w = reinterpret_cast<const unsigned*>(rowbuf + buf * ROWB)[m];
          n0 = (float)(w & 0xFFFFu); n1 = (float)(w >> 16);
        } else {
          const int2 w = reinterpret_cast<const int2*>(rowbuf + buf * ROWB)[m];
          n0 = (float)w.x; n1 = (float)w.y;
        }
        run = fmaf(n0 + a.alpha, ph.x, run); cpre[2 * jp] = run;
        run = fmaf(n1 + a.alpha, ph.y, run); cpre[2 * jp + 1] = run;
      }
      float cum = run;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float v = __shfl_up_sync(0xffffffffu, cum, o);
        if (lane >= o) cum += v;
      }
      const float total = __shfl_sync(0xffffffffu, cum, 31);
      const float u = uniform01(a.seed, (unsigned long long)tk * 2654435761ULL + (unsigned long long)i) * total;
      const unsigned int ball = __ballot_sync(0xffffffffu, cum > u);
      const int L = ball ? (__ffs(ball) - 1) : 31;
      const float target = u - (cum - run);                  // position inside this lane's stretch
      unsigned mask = 0u;
#pragma unroll
      for (int q = 0; q < KJ; ++q) mask |= (cpre[q] > target) ? (1u << q) : 0u;
      int zmine = zold;
      if (mask != 0u) {
        const int q = __ffs(mask) - 1;
        const int k = 2 * (lane + 32 * (q >> 1)) + (q & 1);
        if (k < K) zmine = k;
      }
      const int znew = __shfl_sync(0xffffffffu, zmine, L);
      // put the token back with its new topic; the count tables only move when the topic changed
      if (lane == 0) {
        phi[znew] += s_cm[znew];
        if (znew != zold) {
          atomicAdd(wrow + znew, 1);
          atomicSub(wrow + zold, 1);
          if (DK16) {
            int* d32 = reinterpret_cast<int*>(reinterpret_cast<unsigned short*>(a.n_dk) + (size_t)d * K);
            atomicAdd(d32 + (znew >> 1), (znew & 1) ? 65536 : 1);       // rows are 16-byte aligned here
            atomicSub(d32 + (zold >> 1), (zold & 1) ? 65536 : 1);
          } else {
            int* drow = a.n_dk + (size_t)d * K;
            atomicAdd(drow + znew, 1);
            atomicSub(drow + zold, 1);
          }
        }
      }
      // the next token's row was requested before these table moves: remember them in case it is the same document
      pend_doc = (znew != zold) ? d : -1; pend_new = znew; pend_old = zold;
      if (lane == j) znew_l = znew;
    }
    if (lane < nb) a.tok_z[kb + lane] = znew_l;
  }
  return (unsigned long long)(k1 - k0);
}

template <int KJ, bool DK16>
__global__ void __launch_bounds__(32 * lda_smem_warps<KJ, DK16>(), 1) lda_pass_smem_kernel(LdaArgs a) {
  constexpr int KS = 32 * KJ;                       // padded topic count
  constexpr int KP = KJ / 2;                        // topic pairs per lane
  constexpr int NW = lda_smem_warps<KJ, DK16>();
  constexpr int ROWB = DK16 ? KS * 2 : KS * 4;      // bytes of one document-row buffer
  extern __shared__ __align__(16) uint8_t lsm[];
  float* s_cm = reinterpret_cast<float*>(lsm);                                   // [KS]
  float* s_phi_all = s_cm + KS;                                                  // [NW][KS]
  uint8_t* s_row_all = reinterpret_cast<uint8_t*>(s_phi_all + NW * KS);          // [NW][2][ROWB]
  __shared__ int s_next;
  __shared__ int s_stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x, B = a.B, K = a.K;
  unsigned long long local_trained = 0;
  const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
  for (int k = threadIdx.x; k < KS; k += blockDim.x) s_cm[k] = (k < K) ? a.commons[k] : 0.f;
  for (int k = threadIdx.x; k < NW * 2 * ROWB / 4; k += blockDim.x) reinterpret_cast<int*>(s_row_all)[k] = 0;
  float* phi = s_phi_all + warp * KS;
  uint8_t* rowbuf = s_row_all + (size_t)warp * 2 * ROWB;
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
        local_trained += lda_run_segment_smem<KJ, DK16>(a, seg, lane, K, i, phi, rowbuf, s_cm);
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

// ----------------------------------------------------------------------------
// Whole-iteration kernel: N rotation steps x S word-topic slices in ONE cooperative launch; the last warp of
// every CTA is the communication warp of rotate_ring.cuh, pulling the next use of the other slice over NVLink
// while the sampling warps work (LDAMPCollectiveMapper.java:263-300 — `for j < N: for k < numModelSlices:
// getSplitMap(k); schedule(); rotate(k)`).  Workers are clusters of `cs` CTAs exactly as in the MF-SGD epoch
// kernel (csrc/mf.cu): few, fat workers keep the ticket ring short when the panels are small.
// ----------------------------------------------------------------------------
struct LdaPanelDesc {
  const int* blk_seg_ptr;
  const int* seg_word;
  const int* seg_ptr;
  const int* tok_doc;
  int* tok_z;
  unsigned int* col_ticket;
};

struct LdaEpochArgs {
  void* n_dk;
  const float* commons;
  LdaPanelDesc panels[rot::MAX_STEPS];   // kernel-parameter space: every field stays warp-uniform
  int panel_of_step[rot::MAX_STEPS];
  unsigned long long* trained;
  unsigned long long deadline_ns;
  unsigned long long seed;
  unsigned int ticket_epoch;
  int cs, K;
  float alpha, beta;
  rot::RingArgs ring;
};

struct LdaCore {
  int* n_dk; int* n_wk;
  const int* seg_word; const int* seg_ptr; const int* tok_doc; int* tok_z;
  unsigned long long seed; float alpha, beta;
};

__device__ __forceinline__ void lmbar_arrive_remote(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(cta) : "memory");
}
__device__ __forceinline__ void lmbar_arrive_remote_relaxed(uint32_t bar, uint32_t cta) {     // after ONE fence by the caller
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(cta) : "memory");
}
__device__ __forceinline__ void lst_remote_u32(uint32_t addr, uint32_t cta, uint32_t v) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "st.shared::cluster.u32 [ra], %2;\n"
      "}\n" ::"r"(addr), "r"(cta), "r"(v) : "memory");
}
__device__ __forceinline__ void lmbar_wait_acq_cluster(uint32_t bar, uint32_t parity) {
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.b32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) break;
  }
}

__device__ __forceinline__ uint32_t latom_add_remote_u32(uint32_t addr, uint32_t cta, uint32_t v) {
  uint32_t old;
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %1, %2;\n"
      "atom.shared::cluster.add.u32 %0, [ra], %3;\n"
      "}\n" : "=r"(old) : "r"(addr), "r"(cta), "r"(v) : "memory");
  return old;
}

template <int KJ, bool DK16>
__global__ void __launch_bounds__(32 * (lda_smem_warps<KJ, DK16>() + 1), 1)
lda_epoch_kernel(const __grid_constant__ LdaEpochArgs a) {
  constexpr int KS = 32 * KJ;
  constexpr int NW = lda_smem_warps<KJ, DK16>();         // sampling warps; warp NW is the communication warp
  constexpr int ROWB = DK16 ? KS * 2 : KS * 4;
  extern __shared__ __align__(16) uint8_t lsm[];
  float* s_cm = reinterpret_cast<float*>(lsm);
  float* s_phi_all = s_cm + KS;
  uint8_t* s_row_all = reinterpret_cast<uint8_t*>(s_phi_all + NW * KS);
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ unsigned int s_flags;
  __shared__ int s_end;
  __shared__ unsigned int s_next;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cs = (uint32_t)a.cs;
  const uint32_t crank = cs > 1 ? cluster_ctarank() : 0u;
  const int worker = blockIdx.x / cs, B = gridDim.x / cs, K = a.K;
  const uint32_t bar_go = smem_u32(&s_bar[0]), bar_done = smem_u32(&s_bar[1]);
  const bool ring_on = a.ring.world > 0;
  if (threadIdx.x == 0) {
    mbar_init(bar_go, 1);
    mbar_init(bar_done, cs);
    s_flags = 0u;
    fence_barrier_init();
  }
  for (int k = threadIdx.x; k < KS; k += blockDim.x) s_cm[k] = (k < K) ? a.commons[k] : 0.f;
  for (int k = threadIdx.x; k < NW * 2 * ROWB / 4; k += blockDim.x) reinterpret_cast<int*>(s_row_all)[k] = 0;
  if (cs > 1) cluster_sync_all(); else __syncthreads();

  if (warp == NW) {
    if (ring_on) rot::comm_warp_loop(a.ring);
  } else {
    float* phi = s_phi_all + warp * KS;
    uint8_t* rowbuf = s_row_all + (size_t)warp * 2 * ROWB;
    const bool leader = (crank == 0u && threadIdx.x == 0);
    unsigned long long local_trained = 0;
    uint32_t ph_go = 0, ph_done = 0;
    bool dead = false;
    const unsigned int tbase = a.ticket_epoch * (unsigned)B;
    for (int g = 0; g < a.ring.steps && !dead; ++g) {
      const LdaPanelDesc* pdp = &a.panels[a.panel_of_step[g]];
      LdaCore core;
      core.n_dk = reinterpret_cast<int*>(a.n_dk);
      core.n_wk = ring_on ? reinterpret_cast<int*>(rot::slice_ptr(a.ring, g))
                          : reinterpret_cast<int*>(a.ring.peer[0] + a.ring.off_buf[g % a.ring.S][0]);
      core.seg_word = pdp->seg_word; core.seg_ptr = pdp->seg_ptr; core.tok_doc = pdp->tok_doc; core.tok_z = pdp->tok_z;
      core.seed = a.seed + (unsigned long long)g * 0x9E3779B97F4A7C15ULL; core.alpha = a.alpha; core.beta = a.beta;
      const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
      for (int t = 0; t < B; ++t) {
        const int c = (worker + t) % B;
        if (leader) {
          rot::RingCtl* ctl = ring_on ? rot::ctl_of(a.ring, a.ring.rank) : nullptr;
          unsigned int* tick = pdp->col_ticket + c;
          const int* bsp = pdp->blk_seg_ptr + (size_t)worker * B + c;
          const int s0 = bsp[0], s1 = bsp[1];
          unsigned int fl = 0u;
          if (t == 0 && ring_on && !rot::wait_slice(a.ring, g)) fl = 2u;
          if (fl == 0u) {
            const unsigned int want = tbase + (unsigned)t;
            if (ld_acquire_gpu(tick) != want) {
              const unsigned long long t0 = rot::now_ns();
              unsigned spins = 0;
              while (ld_relaxed_gpu(tick) != want) {       // relaxed polls: an acquire load per poll would keep invalidating L1
                __nanosleep(100);
                if ((++spins & 255u) == 0u && ring_on) {
                  if (rot::aborted(ctl)) { fl = 2u; break; }
                  if (a.ring.timeout_ns != 0ULL && rot::now_ns() - t0 > a.ring.timeout_ns) {
                    if (atomicCAS(&ctl->error, 0u, (unsigned)rot::ERR_TICKET) == 0u) ctl->error_detail = (unsigned)g;
                    fl = 2u; break;
                  }
                }
              }
              (void)ld_acquire_gpu(tick);
            }
          }
          if (fl == 0u && t_end != 0ULL && globaltimer_ns() > t_end) fl = 1u;
          s_next = (unsigned)s0;
          if (cs > 1) {
            for (uint32_t q = 0; q < cs; ++q) {
              lst_remote_u32(smem_u32(&s_flags), q, fl);
              lst_remote_u32(smem_u32(&s_end), q, (unsigned)s1);
            }
            asm volatile("fence.acq_rel.cluster;" ::: "memory");      // one fence, then relaxed arrivals (see csrc/mf.cu)
            for (uint32_t q = 0; q < cs; ++q) lmbar_arrive_remote_relaxed(bar_go, q);
          } else {
            s_flags = fl; s_end = s1;
            mbar_arrive(bar_go);
          }
        }
        if (threadIdx.x == 0) { if (cs > 1) lmbar_wait_acq_cluster(bar_go, ph_go); else { while (!mbar_try_wait(bar_go, ph_go)) {} } }
        ph_go ^= 1u;
        asm volatile("bar.sync 2, %0;" ::"n"(32 * NW) : "memory");     // the others park here instead of spinning
        const unsigned int fl = *reinterpret_cast<volatile unsigned int*>(&s_flags);
        if (!(fl & 3u)) {
          const int s1 = *reinterpret_cast<volatile int*>(&s_end);
          for (;;) {
            int seg = 0;
            if (lane == 0) seg = cs > 1 ? (int)latom_add_remote_u32(smem_u32(&s_next), 0, 1u) : (int)atomicAdd(&s_next, 1u);
            seg = __shfl_sync(0xffffffffu, seg, 0);
            if (seg >= s1) break;
            local_trained += lda_run_segment_smem<KJ, DK16>(core, seg, lane, K, worker, phi, rowbuf, s_cm);
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * NW) : "memory");
        if (cs > 1) {
          if (threadIdx.x == 0) { __threadfence(); lmbar_arrive_remote_relaxed(bar_done, 0); }
          if (leader) lmbar_wait_acq_cluster(bar_done, ph_done);
        }
        if (leader) {
          ph_done ^= 1u;
          asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(pdp->col_ticket + c), "r"(tbase + (unsigned)t + 1u) : "memory");
        }
        if (fl & 2u) { dead = true; break; }
      }
      if (threadIdx.x == 0 && ring_on && !dead) rot::step_done(a.ring, g);
    }
    if (lane == 0 && local_trained) atomicAdd(a.trained, local_trained);
  }
  if (cs > 1) cluster_sync_all(); else __syncthreads();
}

template <int KJ, bool DK16> constexpr size_t lda_smem_bytes() {
  return (size_t)(32 * KJ) * 4 * (1 + lda_smem_warps<KJ, DK16>()) +
         (size_t)lda_smem_warps<KJ, DK16>() * 2 * (DK16 ? 32 * KJ * 2 : 32 * KJ * 4);
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

static const void* lda_fn(int K, bool dk16) {
  if (dk16)
    return K <= 256 ? (const void*)lda_pass_kernel<8, true> : K <= 512 ? (const void*)lda_pass_kernel<16, true>
         : K <= 1024 ? (const void*)lda_pass_kernel<32, true> : (const void*)lda_pass_kernel<64, true>;
  return K <= 256 ? (const void*)lda_pass_kernel<8, false> : K <= 512 ? (const void*)lda_pass_kernel<16, false>
       : K <= 1024 ? (const void*)lda_pass_kernel<32, false> : (const void*)lda_pass_kernel<64, false>;
}

int hb_lda_max_blocks(int K) {
  const bool dk16 = false;   // the int32 variant needs at least as many registers as the 16-bit one
  int dev = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  const void* fn = lda_fn(K, dk16);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, LDA_THREADS, 0) != cudaSuccess) return 0;
  if (per_sm > 1) per_sm = 1;       // the shared-memory kernel fits one CTA per SM
  return per_sm * sm_count();
}

static int lda_launch(void* n_dk, bool dk16, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                      const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, float alpha, float beta,
                      unsigned long long seed, unsigned int* col_ticket, unsigned int ticket_base,
                      unsigned long long deadline_ns, unsigned long long* trained, void* stream) {
  HB_REQUIRE(B >= 1 && K >= 2 && K <= 2048, "hb_lda_pass: need 2 <= K <= 2048 topics (got %d)", K);
  HB_REQUIRE(B <= hb_lda_max_blocks(K), "hb_lda_pass: B=%d blocks cannot be co-resident (max %d)", B, hb_lda_max_blocks(K));
  HB_REQUIRE((reinterpret_cast<uintptr_t>(n_dk) & 7) == 0 && (reinterpret_cast<uintptr_t>(n_wk) & 7) == 0,
             "hb_lda_pass: count tables must be 8-byte aligned");
  LdaArgs a;
  a.n_dk = reinterpret_cast<int*>(n_dk); a.n_wk = n_wk; a.commons = commons; a.blk_seg_ptr = blk_seg_ptr; a.seg_word = seg_word;
  a.seg_ptr = seg_ptr; a.tok_doc = tok_doc; a.tok_z = tok_z; a.col_ticket = col_ticket; a.trained = trained;
  a.deadline_ns = deadline_ns; a.seed = seed; a.ticket_base = ticket_base; a.B = B; a.K = K; a.alpha = alpha; a.beta = beta;
  void* args[] = {&a};
  const bool aligned = dk16 ? (K % 8 == 0) : (K % 4 == 0);
  if (aligned && K <= 1024 && getenv("HB_LDA_GENERIC") == nullptr) {
    const void* fn = nullptr;
    size_t smem = 0;
    int threads = LDA_THREADS;
#define HB_LDA_PICK(KJ_)                                                                                     \
    do {                                                                                                      \
      if (dk16) { fn = (const void*)lda_pass_smem_kernel<KJ_, true>; smem = lda_smem_bytes<KJ_, true>(); threads = 32 * lda_smem_warps<KJ_, true>(); }    \
      else { fn = (const void*)lda_pass_smem_kernel<KJ_, false>; smem = lda_smem_bytes<KJ_, false>(); threads = 32 * lda_smem_warps<KJ_, false>(); }       \
    } while (0)
    if (K <= 256) HB_LDA_PICK(8); else if (K <= 512) HB_LDA_PICK(16); else HB_LDA_PICK(32);
#undef HB_LDA_PICK
    HB_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    HB_REQUIRE(B <= sm_count(), "hb_lda_pass: B=%d blocks cannot be co-resident (max %d)", B, sm_count());
    HB_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(B), dim3(threads), args, smem, (cudaStream_t)stream));
    return 0;
  }
  const void* fn = lda_fn(K, dk16);
  HB_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(B), dim3(LDA_THREADS), args, 0, (cudaStream_t)stream));
  return 0;
}

int hb_lda_pass(int* n_dk, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, float alpha, float beta,
                unsigned long long seed, unsigned int* col_ticket, unsigned int ticket_base,
                unsigned long long deadline_ns, unsigned long long* trained, void* stream) {
  return lda_launch(n_dk, false, n_wk, commons, blk_seg_ptr, seg_word, seg_ptr, tok_doc, tok_z, B, K, alpha, beta, seed,
                    col_ticket, ticket_base, deadline_ns, trained, stream);
}

// Same pass with the document-topic table stored as unsigned 16-bit counts (every document must be shorter than
// 65536 tokens): half the row traffic that bounds the kernel, and the next token's row is fetched while the
// current one is sampled.
int hb_lda_pass_dk16(unsigned short* n_dk16, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                     const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, float alpha, float beta,
                     unsigned long long seed, unsigned int* col_ticket, unsigned int ticket_base,
                     unsigned long long deadline_ns, unsigned long long* trained, void* stream) {
  return lda_launch(n_dk16, true, n_wk, commons, blk_seg_ptr, seg_word, seg_ptr, tok_doc, tok_z, B, K, alpha, beta, seed,
                    col_ticket, ticket_base, deadline_ns, trained, stream);
}

// ---- whole-iteration launch (rotation fused); see hb_sgd_epoch for the ring arguments ----
static const void* lda_epoch_fn(int K, bool dk16, size_t* smem, int* threads) {
#define HB_LDA_EPICK(KJ_)                                                                                       \
  do {                                                                                                           \
    if (dk16) { *smem = lda_smem_bytes<KJ_, true>(); *threads = 32 * (lda_smem_warps<KJ_, true>() + 1); return (const void*)lda_epoch_kernel<KJ_, true>; }   \
    *smem = lda_smem_bytes<KJ_, false>(); *threads = 32 * (lda_smem_warps<KJ_, false>() + 1); return (const void*)lda_epoch_kernel<KJ_, false>;              \
  } while (0)
  if (K <= 256) HB_LDA_EPICK(8);
  if (K <= 512) HB_LDA_EPICK(16);
  HB_LDA_EPICK(32);
#undef HB_LDA_EPICK
}

// 1 when the whole-iteration kernel handles (K, table width): rows must be whole 16-byte chunks, K <= 1024.
int hb_lda_epoch_supported(int K, int dk16) { return (K >= 2 && K <= 1024 && (dk16 ? K % 8 == 0 : K % 4 == 0)) ? 1 : 0; }

int hb_lda_epoch_max_workers(int K, int dk16, int cs) {
  if (cs < 1 || cs > 16 || (cs & (cs - 1)) != 0 || !hb_lda_epoch_supported(K, dk16)) return 0;
  size_t smem = 0; int threads = 0;
  const void* fn = lda_epoch_fn(K, dk16 != 0, &smem, &threads);
  if (cs > 8 && cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (cs == 1) return sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(sm_count() / cs * cs));
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int hb_lda_epoch(void* n_dk, int dk16, const float* commons, const void* panels_host, int npanels, const int* panel_of_step_host,
                 int steps, int S, int workers, int cs, int K, float alpha, float beta, unsigned long long seed,
                 unsigned ticket_epoch, unsigned long long deadline_ns, unsigned long long* trained,
                 void* const* peer_state, int my_rank, int world, unsigned long long off_ctl,
                 const unsigned long long* off_buf4, const unsigned long long* bytes2, const int* src_rank_host,
                 unsigned cnt_base, const unsigned* par_base2, unsigned long long timeout_ns, void* stream) {
  HB_REQUIRE(hb_lda_epoch_supported(K, dk16), "hb_lda_epoch: K=%d (dk16=%d) not supported (need K <= 1024, rows of whole 16-byte chunks)", K, dk16);
  HB_REQUIRE(workers >= 1 && workers <= hb_lda_epoch_max_workers(K, dk16, cs),
             "hb_lda_epoch: %d workers of %d CTAs cannot be co-resident (max %d)", workers, cs, hb_lda_epoch_max_workers(K, dk16, cs));
  HB_REQUIRE((reinterpret_cast<uintptr_t>(n_dk) & 15) == 0, "hb_lda_epoch: the document-topic table must be 16-byte aligned");
  LdaEpochArgs a;
  memset(&a, 0, sizeof(a));
  if (rot::fill_ring(a.ring, peer_state, my_rank, world, S, steps, off_ctl, off_buf4, bytes2, src_rank_host, cnt_base,
                     par_base2, timeout_ns) != 0) return -2;
  HB_REQUIRE(npanels >= 1 && npanels <= rot::MAX_STEPS && panels_host != nullptr && panel_of_step_host != nullptr,
             "hb_lda_epoch: need 1..%d panel descriptors (got %d)", rot::MAX_STEPS, npanels);
  memcpy(a.panels, panels_host, (size_t)npanels * sizeof(LdaPanelDesc));
  for (int g = 0; g < steps; ++g) {
    HB_REQUIRE(panel_of_step_host[g] >= 0 && panel_of_step_host[g] < npanels, "hb_lda_epoch: step %d names panel %d of %d", g, panel_of_step_host[g], npanels);
    a.panel_of_step[g] = panel_of_step_host[g];
  }
  a.n_dk = n_dk; a.commons = commons; a.trained = trained; a.deadline_ns = deadline_ns; a.seed = seed;
  a.ticket_epoch = ticket_epoch; a.cs = cs; a.K = K; a.alpha = alpha; a.beta = beta;
  size_t smem = 0; int threads = 0;
  const void* fn = lda_epoch_fn(K, dk16 != 0, &smem, &threads);
  HB_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return rot::launch_epoch(fn, threads, smem, workers, cs, a, stream);
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
