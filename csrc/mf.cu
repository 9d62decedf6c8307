// Matrix-factorisation kernels: SGD block pass, RMSE, CCD++ rank-one update.
//
// Replaces SGDMPTask.doRun (ml/java/src/main/java/edu/iu/sgd/SGDMPTask.java:46-77), the
// row/column-exclusive dynamic block scheduler with its timer (ml/java/.../dymoro/Scheduler.java:
// 95-237), RMSETask (ml/java/.../sgd/RMSETask.java:67-103), the native mf_sgd.Distri kernel of the
// experimental Harp-DAAL build (experimental/.../daal_sgd/SGDDaalCollectiveMapper.java:274-352) and
// CCDMPTask.doRowCCD/doColCCD (ml/java/.../ccd/CCDMPTask.java:67-125).
//
// SGD design.  The ratings of one "panel" (this worker's rows x the visiting H slice) are bucketed
// into a B x B grid of blocks and, inside a block, sorted by column.  ONE persistent cooperative
// kernel runs B CTAs; CTA i owns row-block i for the whole pass and walks the column-blocks in
// Latin-square order c = (i + t) mod B.  A column-block is handed from CTA to CTA through an
// acquire/release ticket in global memory, so no two CTAs ever touch the same rows or the same
// columns at the same time (the invariant the reference's Scheduler maintains with its
// rowCount/colCount bookkeeping) and there is no grid-wide barrier: CTAs drift apart by as much
// as the load imbalance wants.  Inside a CTA a warp takes one column run at a time: h stays in
// registers for the whole run (sequential SGD down the column, exactly the reference's inner
// loop order), each rating costs one 128-bit-per-lane read and write of its w row and a 5-step
// shuffle reduction.  Two warps of a CTA may meet on a row (different columns); that rare
// intra-block collision is resolved last-writer-wins.  A %globaltimer deadline gives the
// reference's timer-bounded partial pass; the number of ratings actually trained is returned.
#include "common.cuh"
#include "rotate_ring.cuh"
#include "../include/harp_b200.h"

#include <cooperative_groups.h>
#include <stdlib.h>
#include <string.h>

namespace hb {
namespace mf {

// Threads per CTA.  Rank <= 128 needs few registers: 1024 threads x 1 CTA per SM beat 512 x 2 (same warps per SM,
// half as many ring steps with 4x larger cells: 17.5 vs 19.8 ms per epoch on the Netflix-shaped bench).
#ifndef HB_SGD_THREADS
#define HB_SGD_THREADS 1024
#endif
__host__ __device__ constexpr int sgd_threads(int RV) { return RV == 1 ? HB_SGD_THREADS : 512; }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 ldcg4(const float* p) {
  float4 r;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stcg4(float* p, float4 v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct SgdArgs {
  float* W;                  // [rows_local, rank]
  float* H;                  // [cols_slice, rank]
  const int* blk_seg_ptr;    // [B*B + 1]  segments (column runs) of block (i, c) = i * B + c
  const int* seg_col;        // [nseg]     column of the run (slice-local index)
  const int* seg_ptr;        // [nseg + 1] ratings of the run
  const int* r_row;          // [nnz]      row (worker-local index)
  const float* r_val;        // [nnz]
  unsigned int* col_ticket;  // [B]        hand-over tickets (monotone across launches)
  unsigned long long* trained;   // [1]    ratings updated (atomicAdd)
  unsigned long long deadline_ns;  // time budget in ns measured from kernel entry; 0 = no deadline
  unsigned int ticket_base;
  int B;                     // CTAs = row-blocks
  int C;                     // column-blocks (ring schedule: C == B)
  int dynamic;               // 0 = Latin-square ticket ring, 1 = dynamic: any free column-block not yet visited
  unsigned int* col_lock;    // [C] dynamic schedule: 0 = free, 1 = held (all 0 between launches)
  int rank;
  float eps, lambda;
};

#ifndef HB_SGD_PF
#define HB_SGD_PF 1     // deeper w prefetch measured slower (the pass is bound by random-row DRAM traffic, not latency)
#endif
// One column run (segment) of a block: h stays in registers, the run's ratings are trained in order.
// Returns the number of ratings of the run (same value in every lane).
// The run's ratings trained in order with h in registers.  `h` (the column's vector) and the first 32 row ids /
// values (`rows0`, `vals0`: lane j holds rating k0 + j) are passed in, so a caller can fetch them for the NEXT run
// while the current one is being trained.
template <int RV, typename Core>
__device__ __forceinline__ unsigned long long sgd_process_run(const Core& a, int col_raw, int k0, int k1, float4 (&h)[RV],
                                                              int rows0, float vals0, int lane, int nl4) {
  constexpr int PF = RV == 1 ? HB_SGD_PF : (RV == 2 ? 2 : 1);     // w rows in flight per warp
    // bit 31 of seg_col marks a sub-run of a long column run that was split so that several warps
    // can work on a popular column at once; such sub-runs merge their h deltas additively
    const bool shared_col = col_raw < 0;
    const int col = col_raw & 0x7FFFFFFF;
    float4 h0[RV];
#pragma unroll
    for (int u = 0; u < RV; ++u) h0[u] = h[u];
    // Row ids and values of the run come in with one coalesced load (lane j holds rating k0 + j); the w rows
    // of the next PF ratings are kept in flight, so a rating costs 1/PF of a memory round trip, not a whole one.
    for (int kb = k0; kb < k1; kb += 32) {
      const int nb = min(32, k1 - kb);
      const int rows_l = (kb == k0) ? rows0 : ((lane < nb) ? a.r_row[kb + lane] : 0);
      const float vals_l = (kb == k0) ? vals0 : ((lane < nb) ? a.r_val[kb + lane] : 0.f);
      // a w row fetched early is stale if one of the PF ratings before it hits the same row (duplicate
      // (row, col) ratings only): those are flagged in bit 31 and re-read at their turn
      unsigned dup = 0;
#pragma unroll
      for (int q = 1; q <= PF; ++q) {
        const int prev = __shfl_up_sync(0xffffffffu, rows_l, q);
        if (lane >= q && lane < nb && prev == rows_l) dup = 0x80000000u;
      }
      const unsigned rows_e = (unsigned)rows_l | dup;
      float4 wq[PF][RV];
#pragma unroll
      for (int q = 0; q < PF; ++q) {
        const int rq = (int)(__shfl_sync(0xffffffffu, rows_e, q) & 0x7FFFFFFFu);
#pragma unroll
        for (int u = 0; u < RV; ++u)
          wq[q][u] = (q < nb && lane + 32 * u < nl4) ? ldcg4(a.W + (size_t)rq * a.rank + (lane + 32 * u) * 4)
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      for (int j0 = 0; j0 < nb; j0 += PF) {
#pragma unroll
        for (int q = 0; q < PF; ++q) {
          const int j = j0 + q;
          if (j < nb) {
            const unsigned re = __shfl_sync(0xffffffffu, rows_e, j);
            const int row = (int)(re & 0x7FFFFFFFu);
            const float v = __shfl_sync(0xffffffffu, vals_l, j);
            float4 w[RV];
#pragma unroll
            for (int u = 0; u < RV; ++u) w[u] = wq[q][u];
            if (re & 0x80000000u) {
#pragma unroll
              for (int u = 0; u < RV; ++u)
                if (lane + 32 * u < nl4) w[u] = ldcg4(a.W + (size_t)row * a.rank + (lane + 32 * u) * 4);
            }
            if (j + PF < nb) {
              const int rn = (int)(__shfl_sync(0xffffffffu, rows_e, j + PF) & 0x7FFFFFFFu);
#pragma unroll
              for (int u = 0; u < RV; ++u)
                if (lane + 32 * u < nl4) wq[q][u] = ldcg4(a.W + (size_t)rn * a.rank + (lane + 32 * u) * 4);
            }
            float dot = 0.f;
#pragma unroll
            for (int u = 0; u < RV; ++u)
              dot += w[u].x * h[u].x + w[u].y * h[u].y + w[u].z * h[u].z + w[u].w * h[u].w;
            const float e = warp_sum(dot) - v;
            const float ee = a.eps * e, el = a.eps * a.lambda;
#pragma unroll
            for (int u = 0; u < RV; ++u) {
              // w -= eps (e h + lambda w);  h -= eps (e w_old + lambda h)     (SGDMPTask.java:60-72)
              float4 wn, hn;
              wn.x = w[u].x - (ee * h[u].x + el * w[u].x); hn.x = h[u].x - (ee * w[u].x + el * h[u].x);
              wn.y = w[u].y - (ee * h[u].y + el * w[u].y); hn.y = h[u].y - (ee * w[u].y + el * h[u].y);
              wn.z = w[u].z - (ee * h[u].z + el * w[u].z); hn.z = h[u].z - (ee * w[u].z + el * h[u].z);
              wn.w = w[u].w - (ee * h[u].w + el * w[u].w); hn.w = h[u].w - (ee * w[u].w + el * h[u].w);
              h[u] = hn;
              if (lane + 32 * u < nl4) stcg4(a.W + (size_t)row * a.rank + (lane + 32 * u) * 4, wn);
            }
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < RV; ++u) {
      if (lane + 32 * u < nl4) {
        float* hp = a.H + (size_t)col * a.rank + (lane + 32 * u) * 4;
        if (shared_col)
          red_add_v4(hp, make_float4(h[u].x - h0[u].x, h[u].y - h0[u].y, h[u].z - h0[u].z, h[u].w - h0[u].w));
        else
          stcg4(hp, h[u]);
      }
    }
    return (unsigned long long)(k1 - k0);
}

// One column run (segment) of a block, fetched and trained.  Returns the number of ratings of the run (same value
// in every lane).
template <int RV, typename Core>
__device__ __forceinline__ unsigned long long sgd_run_segment(const Core& a, int seg, int lane, int nl4) {
  const int col_raw = a.seg_col[seg];
  const int col = col_raw & 0x7FFFFFFF;
  const int k0 = a.seg_ptr[seg], k1 = a.seg_ptr[seg + 1];
  float4 h[RV];
#pragma unroll
  for (int u = 0; u < RV; ++u)
    h[u] = (lane + 32 * u < nl4) ? ldcg4(a.H + (size_t)col * a.rank + (lane + 32 * u) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
  const int rows0 = (k0 + lane < k1) ? a.r_row[k0 + lane] : 0;
  const float vals0 = (k0 + lane < k1) ? a.r_val[k0 + lane] : 0.f;
  return sgd_process_run<RV>(a, col_raw, k0, k1, h, rows0, vals0, lane, nl4);
}

template <int RV>   // float4s per lane: rank <= 128 * RV
__global__ void __launch_bounds__(sgd_threads(RV), (RV == 1 && sgd_threads(RV) <= 512) ? 2 : 1) sgd_pass_kernel(SgdArgs a) {
  __shared__ int s_next;      // next segment of the current block
  __shared__ int s_stop;
  __shared__ int s_c;         // column-block of this step
  __shared__ unsigned s_done[64];   // dynamic schedule: column-blocks this CTA has already visited
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x, B = a.B, C = a.C;
  const int nl4 = a.rank >> 2;
  unsigned long long local_trained = 0;
  bool expired = false;
  const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
  if (threadIdx.x < 64) s_done[threadIdx.x] = 0u;
  __syncthreads();
  int scan_from = (int)(((long long)i * C) / B);     // staggered start: no contention on the first step
  for (int t = 0; t < C; ++t) {
    // ---- acquire a column-block ----
    if (a.dynamic) {
      // The reference's Scheduler hands a worker ANY free column block it has not trained yet
      // (dymoro/Scheduler.java:95-237).  Same here: warp 0 probes 32 candidates at a time and takes the first
      // free one with a CAS, so a slow neighbour never stalls the ring behind it.
      if (warp == 0) {
        int got = -1;
        while (got < 0) {
          for (int base = 0; base < C && got < 0; base += 32) {
            int cand = scan_from + base + lane;
            const bool valid = base + lane < C;
            cand %= C;
            const bool open = valid && !((s_done[cand >> 5] >> (cand & 31)) & 1u) &&
                              ld_acquire_gpu(a.col_lock + cand) == 0u;
            unsigned m = __ballot_sync(0xffffffffu, open);
            while (m != 0u && got < 0) {
              const int l = __ffs(m) - 1;
              m &= m - 1;
              const int cc = __shfl_sync(0xffffffffu, cand, l);
              unsigned old = 1u;
              if (lane == 0) old = atomicCAS(a.col_lock + cc, 0u, 1u);
              old = __shfl_sync(0xffffffffu, old, 0);
              if (old == 0u) got = cc;
            }
          }
          if (got < 0) __nanosleep(200);
        }
        if (lane == 0) {
          __threadfence();
          s_done[got >> 5] |= 1u << (got & 31);
          s_c = got;
          s_next = a.blk_seg_ptr[(size_t)i * C + got];
          s_stop = (t_end != 0 && globaltimer_ns() > t_end) ? 1 : 0;
        }
        scan_from = (got + 1) % C;
      }
    } else if (threadIdx.x == 0) {
      const int cr = (i + t) % B;
      const unsigned int want = a.ticket_base + (unsigned)t;
      while (ld_acquire_gpu(a.col_ticket + cr) != want) { __nanosleep(64); }
      s_c = cr;
      s_next = a.blk_seg_ptr[(size_t)i * C + cr];
      s_stop = (t_end != 0 && globaltimer_ns() > t_end) ? 1 : 0;
    }
    __syncthreads();
    expired = (s_stop != 0);
    const int c = s_c;
    const int s_end = a.blk_seg_ptr[(size_t)i * C + c + 1];
    if (!expired) {
      for (;;) {
        int seg = 0;
        if (lane == 0) seg = atomicAdd(&s_next, 1);
        seg = __shfl_sync(0xffffffffu, seg, 0);
        if (seg >= s_end) break;
        local_trained += sgd_run_segment<RV>(a, seg, lane, nl4);
      }
    }
    // ---- release column-block c to the next CTA in the ring ----
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (a.dynamic)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.col_lock + c), "r"(0u) : "memory");
      else
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.col_ticket + c), "r"(a.ticket_base + (unsigned)t + 1u) : "memory");
    }
  }
  if (lane == 0 && local_trained) atomicAdd(a.trained, local_trained);
}

// ----------------------------------------------------------------------------
// Whole-epoch kernel: N rotation steps x S model slices in ONE cooperative launch, with the model
// rotation done by a communication warp per CTA over NVLink peer memory (rotate_ring.cuh) while the
// other warps train (SGDCollectiveMapper.java:263-277 — `for j < N: for k < numModelSlices:
// getSplitMap(k); schedule(); rotate(k)` — plus Rotator.java:53-67 / RotateTask.java:107-124).
//
// A *worker* (the unit that owns a row-block and is handed column-blocks through the ticket ring) is a
// thread-block cluster of CS CTAs: small panels (many GPUs) want FEW, FAT workers — a pass costs B hand-overs
// plus nnz / (B * warps-per-worker) updates, so with 4 or 8 SMs per worker the ring has 37 or 18 steps
// instead of 148 while all SMs stay busy.  Inside a worker the cell's column runs are dealt to the
// CS x 31 training warps round-robin (runs are stored longest first, so the deal is balanced); the
// hand-over inside the cluster goes through two mbarriers in distributed shared memory, the hand-over
// between workers through the acquire/release column tickets in global memory.
// ----------------------------------------------------------------------------
struct SgdPanelDesc {
  const int* blk_seg_ptr;    // [B*B + 1]
  const int* seg_col;
  const int* seg_ptr;
  const int* r_row;
  const float* r_val;
  unsigned int* col_ticket;  // [B], holds ticket_epoch * B between launches
};

struct SgdEpochArgs {
  float* W;
  SgdPanelDesc panels[rot::MAX_STEPS];   // kernel-parameter space: every field stays warp-uniform
  int panel_of_step[rot::MAX_STEPS];     // [steps] index into panels
  unsigned long long* trained;
  unsigned long long deadline_ns;  // per step, 0 = none
  unsigned int ticket_epoch;       // launches of this kernel so far (every panel is visited once per launch)
  int cs;                          // CTAs per worker (cluster size)
  int C;                           // column-blocks per panel (>= workers; the surplus hides the hand-over latency)
  unsigned long long* trace;       // optional [steps][B][workers][3] %globaltimer stamps: wait start, go, done (tuning aid)
  int rank;
  float eps, lambda;
  rot::RingArgs ring;
};

struct SgdCore {
  float* W; float* H;
  const int* seg_col; const int* seg_ptr; const int* r_row; const float* r_val;
  int rank; float eps, lambda;
};

// CTA sizes: 31 training warps + the communication warp for the one-run-per-warp cell, 16 + 1 for the quad-run cell
// (it needs ~100 registers per thread).  The last warp is the communication warp.
__host__ __device__ constexpr int epoch_threads(bool quad) { return quad ? 544 : 1024; }

__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t cta) {
  // release at cluster scope: the arriving thread's prior writes (and, through the CTA barrier before it,
  // its CTA's) are visible to whoever acquires the barrier phase
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(cta) : "memory");
}
// Relaxed form: the caller has already issued ONE fence for all the arrivals it is about to make (a release
// arrive per destination CTA costs a MEMBAR.ALL.GPU each, and the leader makes `cs` of them per hand-over).
__device__ __forceinline__ void mbar_arrive_remote_relaxed(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar), "r"(cta) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
__device__ __forceinline__ void st_remote_u32(uint32_t addr, uint32_t cta, uint32_t v) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "st.shared::cluster.u32 [ra], %2;\n"
      "}\n" ::"r"(addr), "r"(cta), "r"(v) : "memory");
}
__device__ __forceinline__ void mbar_wait_acq_cluster(uint32_t bar, uint32_t parity) {
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

__device__ __forceinline__ uint32_t atom_add_remote_u32(uint32_t addr, uint32_t cta, uint32_t v) {
  uint32_t old;
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %1, %2;\n"
      "atom.shared::cluster.add.u32 %0, [ra], %3;\n"
      "}\n" : "=r"(old) : "r"(addr), "r"(cta), "r"(v) : "memory");
  return old;
}

// All runs of one cell, claimed one at a time from the worker's shared counter.  Kept out of line so that its
// register allocation is that of the per-panel kernel's hot loop, independent of the epoch loop's live state.
template <int RV>
__device__ __noinline__ unsigned long long sgd_run_cell(const SgdCore core, int s1, bool clustered, uint32_t next_addr,
                                                        int lane, int nl4) {
  // the descriptors travel through the parameter block as plain 64-bit words: tell the compiler they are global
  // addresses, or every rating costs four generic loads
  __builtin_assume(__isGlobal(core.seg_col)); __builtin_assume(__isGlobal(core.seg_ptr));
  __builtin_assume(__isGlobal(core.r_row)); __builtin_assume(__isGlobal(core.r_val));
  __builtin_assume(__isGlobal(core.W)); __builtin_assume(__isGlobal(core.H));
  unsigned long long n = 0;
  for (;;) {
    int seg = 0;
    if (lane == 0) {
      if (clustered) seg = (int)atom_add_remote_u32(next_addr, 0, 1u);
      else asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(seg) : "r"(next_addr) : "memory");
    }
    seg = __shfl_sync(0xffffffffu, seg, 0);
    if (seg >= s1) break;
    n += sgd_run_segment<RV>(core, seg, lane, nl4);
  }
  return n;
}

// ----------------------------------------------------------------------------
// Quad-run cell (rank <= 128): a warp trains FOUR column runs at once, eight lanes per run, sixteen floats of w
// and h per lane.  The one-run-per-warp form above spends 86 issued instructions per rating (ncu: the pass is
// instruction-issue bound, a 5-step shuffle reduction and the address arithmetic for four floats of real work per
// lane); here every issued instruction serves four ratings, the dot product needs three shuffle steps, and a warp
// keeps four w rows in flight instead of one.  Runs are stored longest first, so the four runs a warp claims
// together have nearly the same length.  Semantics per run are unchanged: ratings in order, h in registers,
// split runs merge their h deltas with red.add.
// ----------------------------------------------------------------------------
__device__ __forceinline__ float4 ld_w4(const float* p, bool on) {
  return on ? ldcg4(p) : make_float4(0.f, 0.f, 0.f, 0.f);
}

__device__ __noinline__ unsigned long long sgd_run_cell_quad(const SgdCore core, int s1, bool clustered, uint32_t next_addr,
                                                             int lane) {
  __builtin_assume(__isGlobal(core.seg_col)); __builtin_assume(__isGlobal(core.seg_ptr));
  __builtin_assume(__isGlobal(core.r_row)); __builtin_assume(__isGlobal(core.r_val));
  __builtin_assume(__isGlobal(core.W)); __builtin_assume(__isGlobal(core.H));
  const int grp = lane >> 3, sub = lane & 7;
  const int nl4 = core.rank >> 2;
  bool on[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) on[q] = (q * 8 + sub) < nl4;
  const float ee_scale = core.eps, el = core.eps * core.lambda;
  unsigned long long n = 0;
  for (;;) {
    int base = 0;
    if (lane == 0) {
      if (clustered) base = (int)atom_add_remote_u32(next_addr, 0, 4u);
      else asm volatile("atom.shared.add.u32 %0, [%1], 4;" : "=r"(base) : "r"(next_addr) : "memory");
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= s1) break;
    const int seg = base + grp;
    const bool valid = seg < s1;
    int col_raw = 0, k0 = 0, k1 = 0;
    if (valid) { col_raw = core.seg_col[seg]; k0 = core.seg_ptr[seg]; k1 = core.seg_ptr[seg + 1]; }
    const bool shared_col = col_raw < 0;
    const int col = col_raw & 0x7FFFFFFF;
    const int len = k1 - k0;
    float4 h[4], h0[4];
    float* hrow = core.H + (size_t)col * core.rank + sub * 4;
#pragma unroll
    for (int q = 0; q < 4; ++q) { h[q] = ld_w4(hrow + q * 32, valid && on[q]); h0[q] = h[q]; }
    int maxlen = len;
    maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, 8));
    maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, 16));
    for (int kb = 0; kb < maxlen; kb += 8) {
      const int nb = min(8, len - kb);                    // ratings of MY run in this batch (<= 0: run is over)
      const int nbmax = min(8, maxlen - kb);
      const int rows_l = (sub < nb) ? core.r_row[k0 + kb + sub] : 0;
      const float vals_l = (sub < nb) ? core.r_val[k0 + kb + sub] : 0.f;
      int row = __shfl_sync(0xffffffffu, rows_l, 0, 8);
      float4 wq[4];
      {
        const float* wrow = core.W + (size_t)row * core.rank + sub * 4;
#pragma unroll
        for (int q = 0; q < 4; ++q) wq[q] = ld_w4(wrow + q * 32, nb > 0 && on[q]);
      }
      bool reload = false;
      for (int j = 0; j < nbmax; ++j) {
        const bool act = j < nb;
        const float v = __shfl_sync(0xffffffffu, vals_l, j, 8);
        float* wrow = core.W + (size_t)row * core.rank + sub * 4;
        float4 w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) w[q] = wq[q];
        if (reload) {             // the same row twice in a row (duplicate rating): the copy fetched ahead is stale
#pragma unroll
          for (int q = 0; q < 4; ++q) w[q] = ld_w4(wrow + q * 32, on[q]);
        }
        // fetch the next rating's row while this one is trained
        const int rown = __shfl_sync(0xffffffffu, rows_l, (j + 1) & 7, 8);
        const bool more = (j + 1 < nb);
        reload = more && (rown == row);
        if (more && !reload) {
          const float* wn = core.W + (size_t)rown * core.rank + sub * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q) wq[q] = ld_w4(wn + q * 32, on[q]);
        }
        float dot = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) dot += w[q].x * h[q].x + w[q].y * h[q].y + w[q].z * h[q].z + w[q].w * h[q].w;
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        const float ee = ee_scale * (dot - v);
        if (act) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            // w -= eps (e h + lambda w);  h -= eps (e w_old + lambda h)     (SGDMPTask.java:60-72)
            float4 wn, hn;
            wn.x = w[q].x - (ee * h[q].x + el * w[q].x); hn.x = h[q].x - (ee * w[q].x + el * h[q].x);
            wn.y = w[q].y - (ee * h[q].y + el * w[q].y); hn.y = h[q].y - (ee * w[q].y + el * h[q].y);
            wn.z = w[q].z - (ee * h[q].z + el * w[q].z); hn.z = h[q].z - (ee * w[q].z + el * h[q].z);
            wn.w = w[q].w - (ee * h[q].w + el * w[q].w); hn.w = h[q].w - (ee * w[q].w + el * h[q].w);
            h[q] = hn;
            if (on[q]) stcg4(wrow + q * 32, wn);
          }
        }
        row = rown;
      }
    }
    if (valid) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (on[q]) {
          if (shared_col)
            red_add_v4(hrow + q * 32, make_float4(h[q].x - h0[q].x, h[q].y - h0[q].y, h[q].z - h0[q].z, h[q].w - h0[q].w));
          else
            stcg4(hrow + q * 32, h[q]);
        }
      }
    }
    int tot = (sub == 0) ? len : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    n += (unsigned long long)tot;
  }
  return n;
}

template <int RV, bool QUAD>
__global__ void __launch_bounds__(epoch_threads(QUAD), 1) sgd_epoch_kernel(const __grid_constant__ SgdEpochArgs a) {
  constexpr int EPOCH_THREADS = epoch_threads(QUAD);
  constexpr int EPOCH_COMM_WARP = EPOCH_THREADS / 32 - 1;
  constexpr int EPOCH_TRAIN_THREADS = EPOCH_THREADS - 32;
  __shared__ __align__(8) unsigned long long s_bar[2];    // [0] go (1 arrival), [1] done (cs arrivals; used in the leader)
  __shared__ unsigned int s_flags;                        // bit 0: stop (deadline), bit 1: abort
  __shared__ int s_end;                                   // one past the cell's last column run
  __shared__ unsigned int s_next;                         // next unclaimed run of the cell (the leader CTA's copy is used)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cs = (uint32_t)a.cs;
  const uint32_t crank = cs > 1 ? cluster_ctarank() : 0u;
  const int worker = blockIdx.x / cs, B = gridDim.x / cs, C = a.C;
  const uint32_t bar_go = smem_u32(&s_bar[0]), bar_done = smem_u32(&s_bar[1]);
  const int nl4 = a.rank >> 2;
  const bool ring_on = a.ring.world > 0;
  // without a ring there is nothing to communicate: the last warp trains too
  const int ntrain = ring_on ? EPOCH_TRAIN_THREADS : EPOCH_THREADS;
  if (threadIdx.x == 0) {
    mbar_init(bar_go, 1);
    mbar_init(bar_done, cs);
    s_flags = 0u;
    fence_barrier_init();
  }
  if (cs > 1) cluster_sync_all(); else __syncthreads();

  if (ring_on && warp == EPOCH_COMM_WARP) {
    rot::comm_warp_loop(a.ring);
  } else {
    const bool leader = (crank == 0u && threadIdx.x == 0);
    rot::RingCtl* ctl = ring_on ? rot::ctl_of(a.ring, a.ring.rank) : nullptr;
    unsigned long long local_trained = 0;
    uint32_t ph_go = 0, ph_done = 0;
    bool dead = false;
    const unsigned int tbase = a.ticket_epoch * (unsigned)B;
    for (int g = 0; g < a.ring.steps && !dead; ++g) {
      const SgdPanelDesc* pdp = &a.panels[a.panel_of_step[g]];
      SgdCore core;
      core.W = a.W;
      core.H = ring_on ? reinterpret_cast<float*>(rot::slice_ptr(a.ring, g))
                       : reinterpret_cast<float*>(a.ring.peer[0] + a.ring.off_buf[g % a.ring.S][0]);
      core.seg_col = pdp->seg_col; core.seg_ptr = pdp->seg_ptr; core.r_row = pdp->r_row; core.r_val = pdp->r_val;
      core.rank = a.rank; core.eps = a.eps; core.lambda = a.lambda;
      const unsigned long long t_end = a.deadline_ns ? globaltimer_ns() + a.deadline_ns : 0ULL;
      for (int t = 0; t < C; ++t) {
        const int c = (worker + t) % C;
        if (leader) {
          unsigned int* tick = pdp->col_ticket + c;
          const int* bsp = pdp->blk_seg_ptr + (size_t)worker * C + c;
          const int s0 = bsp[0], s1 = bsp[1];                 // independent of the ticket: issued before the spin
          unsigned int fl = 0u;
          unsigned long long* tr = a.trace ? a.trace + (((size_t)g * C + t) * B + worker) * 27 : nullptr;
          if (tr) tr[0] = rot::now_ns();
          if (t == 0 && ring_on && !rot::wait_slice(a.ring, g)) fl = 2u;
          if (fl == 0u) {
            // column-block c is ours once its previous holder has released it
            // block c is visited at step t' by worker (c - t') mod C when that is < B: our turn comes after the
            // workers worker+1 .. worker+t (mod C) that exist
            const int before = min(t, B - 1 - worker) + max(0, worker + t - C + 1);
            const unsigned int want = tbase + (unsigned)before;
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
          if (tr) tr[1] = rot::now_ns();
          s_next = (unsigned)s0;
          if (cs > 1) {
            for (uint32_t q = 0; q < cs; ++q) {
              st_remote_u32(smem_u32(&s_flags), q, fl);
              st_remote_u32(smem_u32(&s_end), q, (unsigned)s1);
            }
            fence_acq_rel_cluster();         // one fence, then relaxed arrivals: publishes the stores above and what the
            for (uint32_t q = 0; q < cs; ++q) mbar_arrive_remote_relaxed(bar_go, q);      // ticket acquire made visible
          } else {
            s_flags = fl; s_end = s1;
            mbar_arrive(bar_go);
          }
        }
        // One thread per CTA waits for the worker's leader (unbounded: the leader may be waiting for a peer GPU and
        // broadcasts an abort flag if it gives up); everybody else parks at a hardware barrier instead of spinning
        // on the mbarrier, which cost a quarter more issued instructions.
        if (threadIdx.x == 0) { if (cs > 1) mbar_wait_acq_cluster(bar_go, ph_go); else { while (!mbar_try_wait(bar_go, ph_go)) {} } }
        ph_go ^= 1u;
        asm volatile("bar.sync 2, %0;" ::"r"(ntrain) : "memory");
        unsigned long long* trc = (a.trace && threadIdx.x == 0 && crank < 8u) ? a.trace   /* the slot holds 8 CTAs */ + (((size_t)g * C + t) * B + worker) * 27 + 3 + crank * 3 : nullptr;
        if (trc) trc[0] = rot::now_ns();
        const unsigned int fl = *reinterpret_cast<volatile unsigned int*>(&s_flags);
        if (!(fl & 3u)) {
          const int s1 = *reinterpret_cast<volatile int*>(&s_end);
          if (QUAD) local_trained += sgd_run_cell_quad(core, s1, cs > 1, smem_u32(&s_next), lane);
          else local_trained += sgd_run_cell<RV>(core, s1, cs > 1, smem_u32(&s_next), lane, nl4);
        }
        // ---- hand the column-block on ----
        asm volatile("bar.sync 1, %0;" ::"r"(ntrain) : "memory");
        if (trc) trc[1] = rot::now_ns();
        // One fence per CTA (its writes, ordered before thread 0 by the barrier above), then a relaxed arrival at the
        // leader; the leader's ticket store below is the gpu-scope release that hands everything to the next worker.
        // With a single CTA per worker the release store alone does it.
        if (cs > 1) {
          if (threadIdx.x == 0) { __threadfence(); if (trc) trc[2] = rot::now_ns(); mbar_arrive_remote_relaxed(bar_done, 0); }
          if (leader) mbar_wait_acq_cluster(bar_done, ph_done);
        }
        if (leader) {
          ph_done ^= 1u;
          // tickets keep moving even when the step was cut short (deadline / abort) so nobody waits for ever
          if (a.trace) a.trace[(((size_t)g * C + t) * B + worker) * 27 + 2] = rot::now_ns();
          {
            const int before = min(t, B - 1 - worker) + max(0, worker + t - C + 1);
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(pdp->col_ticket + c), "r"(tbase + (unsigned)before + 1u) : "memory");
          }
        }
        if (fl & 2u) { dead = true; break; }
      }
      if (threadIdx.x == 0 && ring_on && !dead) rot::step_done(a.ring, g);
    }
    if (lane == 0 && local_trained) atomicAdd(a.trained, local_trained);
  }
  // nobody leaves while a cluster peer may still arrive on its barriers
  if (cs > 1) cluster_sync_all(); else __syncthreads();
}

// Σ (v - w·h)^2 over a rating list; one warp per rating.
__global__ void rmse_kernel(const float* __restrict__ W, const float* __restrict__ H, const int* __restrict__ rows,
                            const int* __restrict__ cols, const float* __restrict__ vals, long long n, int rank,
                            double* __restrict__ sq_err) {
  const int lane = threadIdx.x & 31;
  const long long wid = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  double acc = 0.0;
  for (long long k = wid; k < n; k += nw) {
    const float* w = W + (size_t)rows[k] * rank;
    const float* h = H + (size_t)cols[k] * rank;
    float dot = 0.f;
    for (int j = lane; j < rank; j += 32) dot = fmaf(w[j], h[j], dot);
    dot = warp_sum(dot);
    const float e = vals[k] - dot;
    acc += (double)e * e;
  }
  if (lane == 0 && acc != 0.0) atomicAdd(sq_err, acc);
}

// CCD++ rank-one subproblem for one side: given residuals r_ij = v_ij - Σ_f w_if h_jf (kept per
// rating), feature f:  for each row i:  w_if <- Σ_j (r_ij + w_if h_jf) h_jf / (λ + Σ_j h_jf²), then
// r_ij updated in place.  CSR over the side being updated; one warp per row.
// (CCDMPTask.doRowCCD / doColCCD, ml/java/.../ccd/CCDMPTask.java:67-125)
__global__ void ccd_update_kernel(const int* __restrict__ ptr, const int* __restrict__ idx,
                                  float* __restrict__ resid, const int* __restrict__ resid_perm,
                                  float* __restrict__ Wf, const float* __restrict__ Hf, int nrows, float lambda) {
  const int lane = threadIdx.x & 31;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  for (int i = wid; i < nrows; i += nw) {
    const int k0 = ptr[i], k1 = ptr[i + 1];
    if (k0 == k1) continue;
    const float wi = Wf[i];
    float up = 0.f, down = 0.f;
    for (int k = k0 + lane; k < k1; k += 32) {
      const float hj = Hf[idx[k]];
      const int rk = resid_perm ? resid_perm[k] : k;
      up = fmaf(resid[rk] + wi * hj, hj, up);
      down = fmaf(hj, hj, down);
    }
    up = warp_sum(up); down = warp_sum(down);
    const float wnew = up / (lambda * (float)(k1 - k0) + down);      // CCDMPTask.java:80: down = lambda * numV + sum h^2
    const float delta = wi - wnew;
    for (int k = k0 + lane; k < k1; k += 32) {
      const int rk = resid_perm ? resid_perm[k] : k;
      resid[rk] += delta * Hf[idx[k]];
    }
    if (lane == 0) Wf[i] = wnew;
  }
}


// One CCD++ coordinate sweep of ONE feature over the rows (or columns) a rank owns, in a single pass over their
// ratings (CCDMPTask.java:67-125: up = sum (m1 + w h) h, down = lambda numV + sum h^2, z* = up / down,
// m1 -= (z* - w) h).  The two orders of the residual (by row for the W sweeps, by column for the H sweeps) are kept
// in step without extra passes: a sweep leaves the OTHER order stale by one rank-one term, and the next sweep over
// that order applies it on the fly:
//     true residual = res[k] - pd_other[idx[k]] * pd_self[i]
// (pd_other = change of the other side's vector in the pending sweep, all ids; pd_self = this side's value of the
// pending feature for the owned id i).  G threads work on one id: a warp for short lists, a whole CTA for long ones.
template <int G, int U>
__global__ void __launch_bounds__(256) ccd_sweep_kernel(const int* __restrict__ ptr, const int* __restrict__ idx,
                                                        float* __restrict__ res, const float* __restrict__ pd_other,
                                                        const float* __restrict__ pd_self, const float* __restrict__ other,
                                                        float* __restrict__ self, float* __restrict__ dself, int n,
                                                        float lambda) {
  __shared__ float sh[2][8];
  const int lg = threadIdx.x % G;
  const int gi = (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) / G);
  const int ng = (int)(((size_t)gridDim.x * blockDim.x) / G);
  for (int i = gi; i < n; i += ng) {
    const int k0 = ptr[i], k1 = ptr[i + 1];
    if (k0 == k1) {
      if (lg == 0 && dself) dself[i] = 0.f;
      continue;
    }
    const float s_old = self[i];
    const float pds = pd_other ? pd_self[i] : 0.f;
    float up = 0.f, down = 0.f;
    // the first U*G ratings of the list stay in registers between the two passes (U independent index -> gather
    // chains in flight per thread); longer lists re-read their tail
    float rc[U], oc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int k = k0 + lg + u * G;
      rc[u] = 0.f; oc[u] = 0.f;
      if (k < k1) {
        const int j = idx[k];
        const float o = other[j];
        float r = res[k];
        if (pd_other) r = fmaf(-pd_other[j], pds, r);
        rc[u] = r; oc[u] = o;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      up = fmaf(fmaf(s_old, oc[u], rc[u]), oc[u], up);
      down = fmaf(oc[u], oc[u], down);
    }
    for (int k = k0 + lg + U * G; k < k1; k += G) {
      const int j = idx[k];
      const float o = other[j];
      float r = res[k];
      if (pd_other) r = fmaf(-pd_other[j], pds, r);
      up = fmaf(fmaf(s_old, o, r), o, up);
      down = fmaf(o, o, down);
    }
    up = warp_sum(up); down = warp_sum(down);
    if (G > 32) {
      const int w = threadIdx.x >> 5;
      __syncthreads();
      if ((threadIdx.x & 31) == 0) { sh[0][w] = up; sh[1][w] = down; }
      __syncthreads();
      up = 0.f; down = 0.f;
#pragma unroll
      for (int q = 0; q < G / 32; ++q) { up += sh[0][q]; down += sh[1][q]; }
    }
    const float s_new = up / (lambda * (float)(k1 - k0) + down);
    const float delta = s_new - s_old;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int k = k0 + lg + u * G;
      if (k < k1) res[k] = fmaf(-delta, oc[u], rc[u]);
    }
    for (int k = k0 + lg + U * G; k < k1; k += G) {
      const int j = idx[k];
      float r = res[k];
      if (pd_other) r = fmaf(-pd_other[j], pds, r);
      res[k] = fmaf(-delta, other[j], r);
    }
    if (lg == 0) {
      self[i] = s_new;
      if (dself) dself[i] = delta;
    }
  }
}

}  // namespace mf
}  // namespace hb

using namespace hb;
using namespace hb::mf;

extern "C" {

// Largest B (CTAs = blocks per side) a single cooperative launch can keep resident.
int hb_sgd_max_blocks(int rank) {
  int dev = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  const void* fn = rank <= 128 ? (const void*)sgd_pass_kernel<1> : rank <= 256 ? (const void*)sgd_pass_kernel<2>
                                                                                 : (const void*)sgd_pass_kernel<4>;
  const int threads = sgd_threads(rank <= 128 ? 1 : (rank <= 256 ? 2 : 4));
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, 0) != cudaSuccess) return 0;
  return per_sm * sm_count();
}

static int sgd_launch(SgdArgs& a, void* stream);

int hb_sgd_pass(float* W, float* H, const int* blk_seg_ptr, const int* seg_col, const int* seg_ptr,
                const int* r_row, const float* r_val, int B, int rank, float eps, float lambda,
                unsigned int* col_ticket, unsigned int ticket_base, unsigned long long deadline_ns,
                unsigned long long* trained, void* stream) {
  SgdArgs a;
  a.W = W; a.H = H; a.blk_seg_ptr = blk_seg_ptr; a.seg_col = seg_col; a.seg_ptr = seg_ptr; a.r_row = r_row;
  a.r_val = r_val; a.col_ticket = col_ticket; a.trained = trained; a.deadline_ns = deadline_ns;
  a.ticket_base = ticket_base; a.B = B; a.C = B; a.dynamic = 0; a.col_lock = nullptr;
  a.rank = rank; a.eps = eps; a.lambda = lambda;
  return sgd_launch(a, stream);
}

// Dynamic schedule: B row-blocks (CTAs) x C column-blocks (C >= B, C <= 2048); blk_seg_ptr is [B*C + 1] and
// col_lock [C] must be all zero (it is again when the pass completes).
int hb_sgd_pass_dynamic(float* W, float* H, const int* blk_seg_ptr, const int* seg_col, const int* seg_ptr,
                        const int* r_row, const float* r_val, int B, int C, int rank, float eps, float lambda,
                        unsigned int* col_lock, unsigned long long deadline_ns, unsigned long long* trained,
                        void* stream) {
  HB_REQUIRE(C >= B && C <= 2048, "hb_sgd_pass_dynamic: need B <= C <= 2048 (B=%d, C=%d)", B, C);
  SgdArgs a;
  a.W = W; a.H = H; a.blk_seg_ptr = blk_seg_ptr; a.seg_col = seg_col; a.seg_ptr = seg_ptr; a.r_row = r_row;
  a.r_val = r_val; a.col_ticket = nullptr; a.trained = trained; a.deadline_ns = deadline_ns;
  a.ticket_base = 0; a.B = B; a.C = C; a.dynamic = 1; a.col_lock = col_lock;
  a.rank = rank; a.eps = eps; a.lambda = lambda;
  return sgd_launch(a, stream);
}

static int sgd_launch(SgdArgs& a, void* stream) {
  const int B = a.B, rank = a.rank;
  HB_REQUIRE(B >= 1 && rank >= 4 && rank % 4 == 0 && rank <= 512, "hb_sgd_pass: need rank %% 4 == 0 and rank <= 512 (got %d)", rank);
  HB_REQUIRE(B <= hb_sgd_max_blocks(rank), "hb_sgd_pass: B=%d blocks cannot be co-resident (max %d)", B, hb_sgd_max_blocks(rank));
  void* args[] = {&a};
  const void* fn = rank <= 128 ? (const void*)sgd_pass_kernel<1> : rank <= 256 ? (const void*)sgd_pass_kernel<2>
                                                                                 : (const void*)sgd_pass_kernel<4>;
  // cooperative launch: all B CTAs are guaranteed resident, which the ticket ring relies on
  const int threads = sgd_threads(rank <= 128 ? 1 : (rank <= 256 ? 2 : 4));
  HB_CUDA_OK(cudaLaunchCooperativeKernel(fn, dim3(B), dim3(threads), args, 0, (cudaStream_t)stream));
  return 0;
}

// ---- whole-epoch launch (rotation fused) ----
static unsigned long long* g_epoch_trace = nullptr;
// Tuning aid: when set, the next hb_sgd_epoch launches write three %globaltimer stamps per (step, hand-over, worker)
// — wait start, go, done — into this device buffer of steps * B * B * 27 uint64 (3 leader stamps + 3 per CTA of the worker: work start, work end, fenced) (NULL switches it off).
int hb_sgd_epoch_trace(unsigned long long* buf) { g_epoch_trace = buf; return 0; }
static bool sgd_use_quad(int rank) {
  static const bool off = getenv("HB_SGD_QUAD") != nullptr && atoi(getenv("HB_SGD_QUAD")) == 0;     // A/B knob
  return rank <= 128 && !off;
}
static const void* sgd_epoch_fn(int rank) {
  if (sgd_use_quad(rank)) return (const void*)sgd_epoch_kernel<1, true>;
  return rank <= 128 ? (const void*)sgd_epoch_kernel<1, false> : rank <= 256 ? (const void*)sgd_epoch_kernel<2, false>
                                                                                : (const void*)sgd_epoch_kernel<4, false>;
}
static int sgd_epoch_threads(int rank) { return epoch_threads(sgd_use_quad(rank)); }

// How many workers (clusters of `cs` CTAs, one CTA per SM) one cooperative launch keeps resident.
int hb_sgd_epoch_max_workers(int rank, int cs) {
  if (cs < 1 || cs > 16 || (cs & (cs - 1)) != 0) return 0;
  if (cs > 8 && cudaFuncSetAttribute(sgd_epoch_fn(rank), cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(sm_count() / cs * cs));
  cfg.blockDim = dim3((unsigned)sgd_epoch_threads(rank));
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cs == 1) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sgd_epoch_fn(rank), sgd_epoch_threads(rank), 0) != cudaSuccess) return 0;
    return per_sm > 0 ? sm_count() : 0;
  }
  if (cudaOccupancyMaxActiveClusters(&n, sgd_epoch_fn(rank), &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int hb_ring_ctl_bytes(void) { return (int)sizeof(rot::RingCtl); }

// One epoch = `steps` panel passes (step g trains panel panel_of_step[g] against slice g % S) in ONE launch.
//   panels: HOST array of `npanels` 6-pointer descriptors {blk_seg_ptr, seg_col, seg_ptr, r_row, r_val, col_ticket};
//   every panel must be bucketed with B = `workers` blocks per side and is visited exactly once per launch
//   (its tickets advance by B), ticket_epoch = number of earlier launches on these panels.
//   world == 0: no ring, peer_state[0] + off_buf4[2k] is slice k.  world >= 1: symmetric ring state of every
//   rank (world == 1 is a loop-back ring onto this rank's own buffers).
int hb_sgd_epoch(float* W, const void* panels_host, int npanels, const int* panel_of_step_host, int steps, int S, int workers,
                 int cs_and_flags /* CTAs per worker */, int rank, float eps, float lambda, unsigned ticket_epoch, unsigned long long deadline_ns,
                 unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                 unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                 const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                 unsigned long long timeout_ns, void* stream) {
  const int cs = cs_and_flags & 0xff;
  const int C = ((cs_and_flags >> 8) & 0xff) ? ((cs_and_flags >> 8) & 0xff) : workers;     // column-blocks per panel
  HB_REQUIRE(C >= workers && C <= 255, "hb_sgd_epoch: %d column-blocks for %d workers", C, workers);
  HB_REQUIRE(rank >= 4 && rank % 4 == 0 && rank <= 512, "hb_sgd_epoch: need rank %% 4 == 0 and rank <= 512 (got %d)", rank);
  HB_REQUIRE(workers >= 1 && workers <= hb_sgd_epoch_max_workers(rank, cs),
             "hb_sgd_epoch: %d workers of %d CTAs cannot be co-resident (max %d)", workers, cs, hb_sgd_epoch_max_workers(rank, cs));
  SgdEpochArgs a;
  memset(&a, 0, sizeof(a));
  if (rot::fill_ring(a.ring, peer_state, my_rank, world, S, steps, off_ctl, off_buf4, bytes2, src_rank_host, cnt_base,
                par_base2, timeout_ns) != 0) return -2;
  HB_REQUIRE(npanels >= 1 && npanels <= rot::MAX_STEPS && panels_host != nullptr && panel_of_step_host != nullptr,
             "hb_sgd_epoch: need 1..%d panel descriptors (got %d)", rot::MAX_STEPS, npanels);
  memcpy(a.panels, panels_host, (size_t)npanels * sizeof(SgdPanelDesc));
  for (int g = 0; g < steps; ++g) {
    HB_REQUIRE(panel_of_step_host[g] >= 0 && panel_of_step_host[g] < npanels, "hb_sgd_epoch: step %d names panel %d of %d", g, panel_of_step_host[g], npanels);
    a.panel_of_step[g] = panel_of_step_host[g];
  }
  a.W = W;
  a.trained = trained; a.deadline_ns = deadline_ns; a.ticket_epoch = ticket_epoch; a.cs = cs; a.C = C; a.trace = g_epoch_trace; a.rank = rank;
  a.eps = eps; a.lambda = lambda;
  return rot::launch_epoch(sgd_epoch_fn(rank), sgd_epoch_threads(rank), 0, workers, cs, a, stream);
}

// Device wall clock (ns) usable as a deadline base for hb_sgd_pass.
__global__ void read_timer_kernel(unsigned long long* out) { *out = globaltimer_ns(); }
int hb_device_time_ns(unsigned long long* out_dev, void* stream) {
  read_timer_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(out_dev);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_mf_sq_err(const float* W, const float* H, const int* rows, const int* cols, const float* vals,
                 long long n, int rank, double* sq_err, void* stream) {
  if (n == 0) return 0;
  long long blocks = (n * 32 + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  rmse_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(W, H, rows, cols, vals, n, rank, sq_err);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_ccd_update(const int* ptr, const int* idx, float* resid, const int* resid_perm, float* Wf,
                  const float* Hf, int nrows, float lambda, void* stream) {
  if (nrows == 0) return 0;
  long long blocks = ((long long)nrows * 32 + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  ccd_update_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(ptr, idx, resid, resid_perm, Wf, Hf, nrows, lambda);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// One CCD++ sweep of one feature over n owned ids (see ccd_sweep_kernel).  pd_other / pd_self may both be NULL (the
// residual order is up to date); dself (the change of every owned entry) may be NULL.  wide != 0: one CTA per id.
int hb_ccd_sweep(const int* ptr, const int* idx, float* res, const float* pd_other, const float* pd_self,
                 const float* other, float* self, float* dself, int n, float lambda, int wide, void* stream) {
  HB_REQUIRE((pd_other == nullptr) == (pd_self == nullptr), "hb_ccd_sweep: pending correction needs both vectors");
  if (n <= 0) return 0;
  const int per_cta = wide ? 1 : 8;
  long long blocks = ((long long)n + per_cta - 1) / per_cta;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (wide) ccd_sweep_kernel<256, 24><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(ptr, idx, res, pd_other, pd_self, other, self, dself, n, lambda);
  else ccd_sweep_kernel<32, 8><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(ptr, idx, res, pd_other, pd_self, other, self, dself, n, lambda);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
