// Collapsed Gibbs sampling for LDA with the SparseLDA decomposition, for any number of topics up to 16 384.
//
// Replaces LDAMPTask.sample (ml/java/src/main/java/edu/iu/lda/LDAMPTask.java:209-254: the s / r / t buckets over
// sparse rows) and the sorted packed rows it walks (SortTask.java:39-58, TopicCountList).  The dense kernels of
// lda.cu form p_k for ALL K topics per token (one 2 KB document row per token at K = 1000, K <= 2048 at most); the
// reference's headline runs use K = 10 000.  Here the conditional
//     p(k)  ~  (n_dk + alpha) (n_wk + beta) c_k ,   c_k = 1 / (n_k + V beta)
// is split the SparseLDA way into
//     alpha beta c_k            smoothing bucket: one constant mass per iteration
//   + alpha n_wk c_k            word bucket: non-zero only where the word has counts
//   + n_dk (n_wk + beta) c_k    document bucket: non-zero only on the document's topics
// and evaluated sparsely WITHOUT sparse count rows (which two warps could not update concurrently):
//   * the document bucket is a sum over the document's TOKENS, sum_j (n_w,z_j + beta) c_z_j, because a token list
//     with repeats is the sparse document row (n_dk copies of k); every rank keeps its documents' current
//     assignments contiguous in `doc_z`, a warp reads the ~100 entries with coalesced loads and gathers the
//     matching word counts; the drawn token's topic is the new topic.  No document-topic table is read at all;
//   * the word bucket lives as per-warp block sums (one per 32 topics) in shared memory, built once per word run
//     and patched as counts move; a draw walks 32 block sums per lane, then one 128-byte row segment;
//   * the smoothing bucket uses the same two-level walk over block sums of c_k shared by the CTA.
// Per token: O(doc length + K / 32) instead of O(K); the word-topic table stays dense int32 and is updated with
// atomics, so the tables remain exactly consistent with the assignments whatever the interleaving.
#include "common.cuh"
#include "rotate_ring.cuh"
#include "../include/harp_b200.h"

#include <string.h>

namespace hb {
namespace ldatl {

constexpr int NW = 24;                       // sampling warps per CTA; warp NW is the communication warp
constexpr int THREADS = 32 * (NW + 1);
constexpr int MAX_K = 16384;

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float uniform01(unsigned long long seed, unsigned long long ctr) {
  unsigned long long x = seed + ctr * 0x9E3779B97F4A7C15ULL;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
  x ^= x >> 27; x *= 0x94D049BB133111EBULL;
  x ^= x >> 31;
  return (float)(x >> 40) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ int ldcg_i32(const int* p) {
  int r;
  asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ float warp_incl_scan(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

struct PanelDesc {
  const int* blk_seg_ptr;
  const int* seg_word;
  const int* seg_ptr;
  const int* tok_doc;
  int* tok_z;
  unsigned int* col_ticket;
  const int* tok_pos;          // position of the token inside doc_z
};

struct Core {
  int* n_wk;
  int* n_wb;                   // [words][nb] exact per-32-topic block counts of n_wk (NULL: not kept)
  const int* seg_word; const int* seg_ptr; const int* tok_doc; int* tok_z; const int* tok_pos;
  int* doc_z; const int* d_ptr;
  unsigned long long seed;
  float alpha, beta, csum;     // csum = sum_k c_k
};

// Two-level inverse CDF over `nb` block sums held in shared memory: returns the block whose running sum first
// exceeds u (last block on rounding) and leaves u relative to that block's start.
__device__ __forceinline__ int pick_block(const float* bs, int nb, float& u, int lane) {
  float run = 0.f;
  for (int b0 = 0; b0 < nb; b0 += 32) {
    const float v = (b0 + lane < nb) ? fmaxf(bs[b0 + lane], 0.f) : 0.f;
    const float inc = warp_incl_scan(v, lane);
    const float tot = __shfl_sync(0xffffffffu, inc, 31);
    if (u < run + tot || b0 + 32 >= nb) {
      const unsigned m = __ballot_sync(0xffffffffu, run + inc > u && b0 + lane < nb);
      int l = m ? (__ffs(m) - 1) : min(31, nb - 1 - b0);
      const float before = __shfl_sync(0xffffffffu, inc - v, l);
      u -= run + before;
      return b0 + l;
    }
    run += tot;
  }
  return nb - 1;
}

// One word run sampled by one warp.  `bs` = this warp's block sums of n_wk c_k for the run's word.
__device__ __forceinline__ unsigned long long run_segment(const Core& a, int seg, int lane, int K, int salt, float* bs,
                                                          const float* s_cm, const float* s_cbs) {
  const int nb = (K + 31) >> 5;
  const int word = a.seg_word[seg] & 0x7FFFFFFF;
  const int k0 = a.seg_ptr[seg], k1 = a.seg_ptr[seg + 1];
  int* wrow = a.n_wk + (size_t)word * K;
  // block sums of the word bucket, one coalesced 128-byte row segment per block.  With the block counts only the
  // blocks that hold any of the word's tokens are read (a word with 100 tokens touches at most 100 of the 313 blocks
  // of a 10 000-topic row; most words far fewer), instead of the whole 40 KB row per run.
  int* wbrow = a.n_wb ? a.n_wb + (size_t)word * nb : nullptr;
  float wsum = 0.f;
  __syncwarp();
  if (wbrow != nullptr) {
    for (int b0 = 0; b0 < nb; b0 += 32) {
      const int b = b0 + lane;
      const int cnt = b < nb ? ldcg_i32(wbrow + b) : 0;
      if (b < nb) bs[b] = 0.f;
      unsigned m = __ballot_sync(0xffffffffu, cnt > 0);
      __syncwarp();
      while (m) {
        const int bb = b0 + __ffs(m) - 1;
        m &= m - 1;
        const int k = bb * 32 + lane;
        float v = (k < K) ? (float)ldcg_i32(wrow + k) * s_cm[k] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) bs[bb] = v;
        wsum += v;
      }
    }
  } else {
    for (int b = 0; b < nb; ++b) {
      const int k = b * 32 + lane;
      float v = (k < K) ? (float)ldcg_i32(wrow + k) * s_cm[k] : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) bs[b] = v;
      wsum += v;
    }
  }
  __syncwarp();
  for (int tk = k0; tk < k1; ++tk) {
    const int d = a.tok_doc[tk], zold = a.tok_z[tk], pos = a.tok_pos[tk];
    const float cz = s_cm[zold];
    // take the token out (shared copies only; the table moves once, below)
    if (lane == 0) bs[zold >> 5] -= cz;
    __syncwarp();
    const float wex = fmaxf(wsum - cz, 0.f);
    // ---- document bucket: sum over the document's other tokens of (n_w,z + beta) c_z ----
    const int p0 = a.d_ptr[d], p1 = a.d_ptr[d + 1];
    float dsum = 0.f;
    // the first KEEP rounds (documents up to 128 tokens) stay in registers so that a draw from this bucket does not
    // gather the word counts a second time
    constexpr int KEEP = 4;
    float vr[KEEP], tr[KEEP];
    int zr[KEEP];
#pragma unroll
    for (int r = 0; r < KEEP; ++r) { vr[r] = 0.f; tr[r] = 0.f; zr[r] = zold; }
#pragma unroll
    for (int r = 0; r < KEEP; ++r) {
      const int j = p0 + r * 32 + lane;
      if (p0 + r * 32 < p1) {
        if (j < p1 && j != pos) {
          zr[r] = a.doc_z[j];
          const int nw = bs[zr[r] >> 5] > 1e-12f ? ldcg_i32(wrow + zr[r]) - (zr[r] == zold ? 1 : 0) : 0;   // empty block: no gather
          vr[r] = ((float)max(nw, 0) + a.beta) * s_cm[zr[r]];
        }
        float t = vr[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        tr[r] = t;
        dsum += t;
      }
    }
    for (int j0 = p0 + KEEP * 32; j0 < p1; j0 += 32) {
      const int j = j0 + lane;
      float v = 0.f;
      if (j < p1 && j != pos) {
        const int z = a.doc_z[j];
        const int nw = bs[z >> 5] > 1e-12f ? ldcg_i32(wrow + z) - (z == zold ? 1 : 0) : 0;
        v = ((float)max(nw, 0) + a.beta) * s_cm[z];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      dsum += v;
    }
    const float amass = a.alpha * (wex + a.beta * a.csum);
    float u = uniform01(a.seed, (unsigned long long)tk * 2654435761ULL + (unsigned long long)salt) * (dsum + amass);
    int znew = zold;
    if (u < dsum) {
      bool found = false;
#pragma unroll
      for (int r = 0; r < KEEP; ++r) {
        if (!found && p0 + r * 32 < p1) {
          const bool last = (p0 + (r + 1) * 32 >= p1);
          if (u < tr[r] || last) {
            const float inc = warp_incl_scan(vr[r], lane);
            const unsigned m = __ballot_sync(0xffffffffu, inc > u && vr[r] > 0.f);
            const unsigned any = __ballot_sync(0xffffffffu, vr[r] > 0.f);
            const int l = m ? (__ffs(m) - 1) : (any ? (31 - __clz(any)) : 0);
            znew = __shfl_sync(0xffffffffu, zr[r], l);
            found = true;
          } else {
            u -= tr[r];
          }
        }
      }
      // longer documents: walk the rest of the list again and stop in the round that contains u
      for (int j0 = p0 + KEEP * 32; j0 < p1 && !found; j0 += 32) {
        const int j = j0 + lane;
        float v = 0.f;
        int z = zold;
        if (j < p1 && j != pos) {
          z = a.doc_z[j];
          const int nw = bs[z >> 5] > 1e-12f ? ldcg_i32(wrow + z) - (z == zold ? 1 : 0) : 0;
          v = ((float)max(nw, 0) + a.beta) * s_cm[z];
        }
        const float inc = warp_incl_scan(v, lane);
        const float tot = __shfl_sync(0xffffffffu, inc, 31);
        if (u < tot || j0 + 32 >= p1) {
          const unsigned m = __ballot_sync(0xffffffffu, inc > u && v > 0.f);
          const unsigned any = __ballot_sync(0xffffffffu, v > 0.f);
          const int l = m ? (__ffs(m) - 1) : (any ? (31 - __clz(any)) : 0);
          znew = __shfl_sync(0xffffffffu, z, l);
          found = true;
        }
        u -= tot;
      }
    } else {
      u = (u - dsum) / a.alpha;                       // position inside  n_wk c_k  |  beta c_k
      const bool word_bucket = u < wex;
      if (!word_bucket) u = (u - wex) / a.beta;
      const int b = pick_block(word_bucket ? bs : s_cbs, nb, u, lane);
      const int k = b * 32 + lane;
      float v = 0.f;
      if (k < K) v = word_bucket ? (float)max(ldcg_i32(wrow + k) - (k == zold ? 1 : 0), 0) * s_cm[k] : s_cm[k];
      const float inc = warp_incl_scan(v, lane);
      const unsigned m = __ballot_sync(0xffffffffu, inc > u && v > 0.f);
      const unsigned any = __ballot_sync(0xffffffffu, v > 0.f);
      if (m | any) znew = b * 32 + (m ? (__ffs(m) - 1) : (31 - __clz(any)));
    }
    // put the token back with its new topic
    if (lane == 0) {
      bs[znew >> 5] += s_cm[znew];
      if (znew != zold) {
        atomicAdd(wrow + znew, 1);
        atomicSub(wrow + zold, 1);
        if (wbrow != nullptr && (znew >> 5) != (zold >> 5)) {
          atomicAdd(wbrow + (znew >> 5), 1);
          atomicSub(wbrow + (zold >> 5), 1);
        }
        a.doc_z[pos] = znew;
        a.tok_z[tk] = znew;
      }
    }
    wsum = wex + s_cm[znew];
    __syncwarp();
  }
  return (unsigned long long)(k1 - k0);
}


struct EpochArgs {
  const float* commons;                  // [K] c_k, frozen for the iteration
  int* doc_z;                            // this rank's tokens in document order: current topic of each
  const int* d_ptr;                      // [docs + 1]
  PanelDesc panels[rot::MAX_STEPS];
  int panel_of_step[rot::MAX_STEPS];
  unsigned long long* trained;
  unsigned long long deadline_ns;
  unsigned long long seed;
  unsigned int ticket_epoch;
  int cs, K;
  float alpha, beta;
  unsigned long long wb_off;             // element offset of the block counts inside a slice (0: none)
  rot::RingArgs ring;
};

__device__ __forceinline__ void arrive_remote_relaxed(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(cta) : "memory");
}
__device__ __forceinline__ void st_remote(uint32_t addr, uint32_t cta, uint32_t v) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "st.shared::cluster.u32 [ra], %2;\n"
      "}\n" ::"r"(addr), "r"(cta), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_remote(uint32_t addr, uint32_t cta, uint32_t v) {
  uint32_t old;
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %1, %2;\n"
      "atom.shared::cluster.add.u32 %0, [ra], %3;\n"
      "}\n" : "=r"(old) : "r"(addr), "r"(cta), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void wait_acq_cluster(uint32_t bar, uint32_t parity) {
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

// Whole iteration in one launch: same skeleton as lda_epoch_kernel (csrc/lda.cu) — workers = clusters of `cs` CTAs
// walking the Latin-square ticket ring, last warp of every CTA = communication warp of rotate_ring.cuh.
__global__ void __launch_bounds__(THREADS, 1) epoch_kernel(const __grid_constant__ EpochArgs a) {
  extern __shared__ __align__(16) uint8_t lsm[];
  const int K = a.K, KS = (K + 31) / 32 * 32, nb = KS / 32;
  float* s_cm = reinterpret_cast<float*>(lsm);           // [KS] c_k
  float* s_cbs = s_cm + KS;                               // [nb] block sums of c_k
  float* s_bs_all = s_cbs + nb;                           // [NW][nb] per-warp block sums of the word bucket
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ unsigned int s_flags;
  __shared__ int s_end;
  __shared__ unsigned int s_next;
  __shared__ float s_csum;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cs = (uint32_t)a.cs;
  const uint32_t crank = cs > 1 ? cluster_ctarank() : 0u;
  const int worker = blockIdx.x / cs, B = gridDim.x / cs;
  const uint32_t bar_go = smem_u32(&s_bar[0]), bar_done = smem_u32(&s_bar[1]);
  const bool ring_on = a.ring.world > 0;
  if (threadIdx.x == 0) {
    mbar_init(bar_go, 1);
    mbar_init(bar_done, cs);
    s_flags = 0u;
    fence_barrier_init();
  }
  for (int k = threadIdx.x; k < KS; k += blockDim.x) s_cm[k] = (k < K) ? a.commons[k] : 0.f;
  __syncthreads();
  for (int b = warp; b < nb; b += blockDim.x >> 5) {
    float v = s_cm[b * 32 + lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_cbs[b] = v;
  }
  __syncthreads();
  if (warp == 0) {
    float v = 0.f;
    for (int b = lane; b < nb; b += 32) v += s_cbs[b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_csum = v;
  }
  if (cs > 1) cluster_sync_all(); else __syncthreads();

  if (warp == NW) {
    if (ring_on) rot::comm_warp_loop(a.ring);
  } else {
    float* bs = s_bs_all + warp * nb;
    const bool leader = (crank == 0u && threadIdx.x == 0);
    unsigned long long local_trained = 0;
    uint32_t ph_go = 0, ph_done = 0;
    bool dead = false;
    const unsigned int tbase = a.ticket_epoch * (unsigned)B;
    for (int g = 0; g < a.ring.steps && !dead; ++g) {
      const PanelDesc* pdp = &a.panels[a.panel_of_step[g]];
      Core core;
      core.n_wk = ring_on ? reinterpret_cast<int*>(rot::slice_ptr(a.ring, g))
                          : reinterpret_cast<int*>(a.ring.peer[0] + a.ring.off_buf[g % a.ring.S][0]);
      core.n_wb = a.wb_off ? core.n_wk + a.wb_off : nullptr;
      core.seg_word = pdp->seg_word; core.seg_ptr = pdp->seg_ptr; core.tok_doc = pdp->tok_doc; core.tok_z = pdp->tok_z;
      core.tok_pos = pdp->tok_pos; core.doc_z = a.doc_z; core.d_ptr = a.d_ptr;
      core.seed = a.seed + (unsigned long long)g * 0x9E3779B97F4A7C15ULL; core.alpha = a.alpha; core.beta = a.beta;
      core.csum = s_csum;
      __builtin_assume(__isGlobal(core.n_wk)); __builtin_assume(__isGlobal(core.seg_word)); __builtin_assume(__isGlobal(core.seg_ptr));
      __builtin_assume(__isGlobal(core.tok_doc)); __builtin_assume(__isGlobal(core.tok_z)); __builtin_assume(__isGlobal(core.tok_pos));
      __builtin_assume(__isGlobal(core.doc_z)); __builtin_assume(__isGlobal(core.d_ptr));
      const unsigned long long t_end = a.deadline_ns ? gtimer() + a.deadline_ns : 0ULL;
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
              while (ld_relaxed_gpu(tick) != want) {
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
          if (fl == 0u && t_end != 0ULL && gtimer() > t_end) fl = 1u;
          s_next = (unsigned)s0;
          if (cs > 1) {
            for (uint32_t q = 0; q < cs; ++q) { st_remote(smem_u32(&s_flags), q, fl); st_remote(smem_u32(&s_end), q, (unsigned)s1); }
            asm volatile("fence.acq_rel.cluster;" ::: "memory");
            for (uint32_t q = 0; q < cs; ++q) arrive_remote_relaxed(bar_go, q);
          } else {
            s_flags = fl; s_end = s1;
            mbar_arrive(bar_go);
          }
        }
        if (threadIdx.x == 0) { if (cs > 1) wait_acq_cluster(bar_go, ph_go); else { while (!mbar_try_wait(bar_go, ph_go)) {} } }
        ph_go ^= 1u;
        asm volatile("bar.sync 2, %0;" ::"n"(32 * NW) : "memory");
        const unsigned int fl = *reinterpret_cast<volatile unsigned int*>(&s_flags);
        if (!(fl & 3u)) {
          const int s1 = *reinterpret_cast<volatile int*>(&s_end);
          for (;;) {
            int seg = 0;
            if (lane == 0) seg = cs > 1 ? (int)atom_add_remote(smem_u32(&s_next), 0, 1u) : (int)atomicAdd(&s_next, 1u);
            seg = __shfl_sync(0xffffffffu, seg, 0);
            if (seg >= s1) break;
            local_trained += run_segment(core, seg, lane, K, worker, bs, s_cm, s_cbs);
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * NW) : "memory");
        if (cs > 1) {
          if (threadIdx.x == 0) { __threadfence(); arrive_remote_relaxed(bar_done, 0); }
          if (leader) wait_acq_cluster(bar_done, ph_done);
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

// K11: sum over a count array of lgamma(count + add), skipping zeros (they contribute lgamma(add), which the caller
// adds in closed form): CalcLikelihoodTask.java:57-77 without float64 copies of the tables.
__global__ void lgamma_sum_kernel(const int* __restrict__ counts, long long n, double add, double* __restrict__ out,
                                  unsigned long long* __restrict__ nnz) {
  double acc = 0.0;
  unsigned long long nz = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = counts[i];
    if (c > 0) { acc += lgamma((double)c + add); ++nz; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    nz += __shfl_xor_sync(0xffffffffu, nz, o);
  }
  if ((threadIdx.x & 31) == 0 && nz) { red_add_f64(out, acc); atomicAdd(nnz, nz); }
}

}  // namespace ldatl
}  // namespace hb

using namespace hb;
using namespace hb::ldatl;

static size_t tl_smem_bytes(int K) {
  const int KS = (K + 31) / 32 * 32, nb = KS / 32;
  return (size_t)(KS + nb * (1 + NW)) * sizeof(float);
}

extern "C" {

int hb_lda_tl_supported(int K) { return (K >= 2 && K <= MAX_K) ? 1 : 0; }

int hb_lda_tl_max_workers(int K, int cs) {
  if (cs < 1 || cs > 16 || (cs & (cs - 1)) != 0 || !hb_lda_tl_supported(K)) return 0;
  if (cs > 8 && cudaFuncSetAttribute(epoch_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 0; }
  const size_t smem = tl_smem_bytes(K);
  if (cudaFuncSetAttribute(epoch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (cs == 1) return sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(sm_count() / cs * cs));
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, epoch_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// One LDA iteration (steps = N rotations x S slices) with the token-list sampler; arguments as hb_lda_epoch, except:
// no document-topic table — doc_z[ntok_local] holds this rank's tokens in document order (current topics),
// d_ptr[docs + 1] the documents' extents, and every panel descriptor has a 7th pointer tok_pos (token -> index
// into doc_z).  K <= 16384.
// Same, with per-32-topic block counts kept next to the word-topic rows: every slice is
// [words x K counts | words x ceil(K/32) block counts], wb_off = words x K (in elements; 0 = no block counts).
int hb_lda_tl_epoch2(int* doc_z, const int* d_ptr, const float* commons, const void* panels_host, int npanels,
                     const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha, float beta,
                     unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                     unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                     unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                     const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                     unsigned long long timeout_ns, unsigned long long wb_off, void* stream);
int hb_lda_tl_epoch(int* doc_z, const int* d_ptr, const float* commons, const void* panels_host, int npanels,
                    const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha, float beta,
                    unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                    unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                    unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                    const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                    unsigned long long timeout_ns, void* stream) {
  return hb_lda_tl_epoch2(doc_z, d_ptr, commons, panels_host, npanels, panel_of_step_host, steps, S, workers, cs, K, alpha,
                          beta, seed, ticket_epoch, deadline_ns, trained, peer_state, my_rank, world, off_ctl, off_buf4,
                          bytes2, src_rank_host, cnt_base, par_base2, timeout_ns, 0ULL, stream);
}

int hb_lda_tl_epoch2(int* doc_z, const int* d_ptr, const float* commons, const void* panels_host, int npanels,
                     const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha, float beta,
                     unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                     unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                     unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                     const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                     unsigned long long timeout_ns, unsigned long long wb_off, void* stream) {
  HB_REQUIRE(hb_lda_tl_supported(K), "hb_lda_tl_epoch: need 2 <= K <= %d topics (got %d)", MAX_K, K);
  HB_REQUIRE(workers >= 1 && workers <= hb_lda_tl_max_workers(K, cs),
             "hb_lda_tl_epoch: %d workers of %d CTAs cannot be co-resident (max %d)", workers, cs, hb_lda_tl_max_workers(K, cs));
  EpochArgs a;
  memset(&a, 0, sizeof(a));
  if (rot::fill_ring(a.ring, peer_state, my_rank, world, S, steps, off_ctl, off_buf4, bytes2, src_rank_host, cnt_base,
                     par_base2, timeout_ns) != 0) return -2;
  HB_REQUIRE(npanels >= 1 && npanels <= rot::MAX_STEPS && panels_host != nullptr && panel_of_step_host != nullptr,
             "hb_lda_tl_epoch: need 1..%d panel descriptors (got %d)", rot::MAX_STEPS, npanels);
  memcpy(a.panels, panels_host, (size_t)npanels * sizeof(PanelDesc));
  for (int g = 0; g < steps; ++g) {
    HB_REQUIRE(panel_of_step_host[g] >= 0 && panel_of_step_host[g] < npanels, "hb_lda_tl_epoch: step %d names panel %d of %d", g, panel_of_step_host[g], npanels);
    a.panel_of_step[g] = panel_of_step_host[g];
  }
  a.commons = commons; a.doc_z = doc_z; a.d_ptr = d_ptr; a.trained = trained; a.deadline_ns = deadline_ns; a.seed = seed;
  a.ticket_epoch = ticket_epoch; a.cs = cs; a.K = K; a.alpha = alpha; a.beta = beta; a.wb_off = wb_off;
  const size_t smem = tl_smem_bytes(K);
  HB_CUDA_OK(cudaFuncSetAttribute(epoch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return rot::launch_epoch((const void*)epoch_kernel, THREADS, smem, workers, cs, a, stream);
}

// *out += sum over counts[i] > 0 of lgamma(counts[i] + add); *nnz += how many were non-zero.
int hb_lgamma_sum(const int* counts, long long n, double add, double* out, unsigned long long* nnz, void* stream) {
  if (n == 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  lgamma_sum_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(counts, n, add, out, nnz);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
