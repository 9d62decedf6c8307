// Model rotation fused into the compute kernel (K5): device-side ring over NVLink peer memory.
//
// Replaces the reference's Rotator / RotateTask pair (ml/java/src/main/java/edu/iu/dymoro/Rotator.java:
// 53-67, RotateTask.java:107-124, :158-210) and the rotate collective underneath it
// (core/harp-collective/.../collective/LocalGlobalSyncCollective.java:710-772).  There a background
// thread ships model slice k to the next worker while the compute threads train on slice 1-k.  Here
// ONE persistent cooperative kernel runs a whole epoch (N rotation steps x S slices): in every CTA the
// last warp is a *communication warp* that pulls the NEXT use of a slice out of the source rank's
// symmetric slab (16-byte loads over NVLink, stripe = 1/gridDim of the slice per CTA) into the local
// ping-pong buffer while the other warps train on the current slice.  No host round trip, no separate
// rotate launch, no grid-wide barrier.
//
// Protocol (all counters monotone, one set per slice k; "use" u = how many times this kernel family
// has trained on slice k so far; every rank does the same uses in the same order):
//   done[k]     local, gpu scope   += 1 by every CTA that finished training use u
//   ready[k]    cross-GPU          = u + 1 once all CTAs of this rank finished use u       (read by consumer)
//   avail[k]    local, gpu scope   = u + 1: CTA 0's comm warp saw the SOURCE's ready >= u + 1 (relay, so
//                                   that only one warp per GPU polls across NVLink)
//   pulled[k]   local, gpu scope   += 1 by every CTA whose stripe of use u + 1 has landed
//   consumed[k] cross-GPU          = u + 1 written INTO THE SOURCE by the consumer when the whole of use u's
//                                   result has been pulled; the source may then overwrite that buffer
// Buffer parity: use u of this launch (local index j) lives in buf[k][par_base[k] ^ (j & 1)] on every rank.
// The source rank of each pull comes from a per-step table, so randomised rotation orders
// (dymoro/RotationUtil.java:58-90) cost nothing extra; a null table is the plain ring (rank - 1).
// Waits are bounded by `timeout_ns` of %globaltimer (0 = unbounded): on expiry the waiter records a code in
// ctl->error and every loop of the kernel drains instead of trapping, so the CUDA context survives
// and the host can raise a normal exception.
#pragma once

#include "common.cuh"
#include "../include/harp_b200.h"

#include <string.h>

namespace hb {
namespace rot {

constexpr int MAX_SLICES = 2;
constexpr int MAX_STEPS = HB_MAX_RANKS * MAX_SLICES;    // steps of one launch: N rotations x S slices

struct RingCtl {                      // one per rank, in the symmetric slab, zeroed once at creation
  uint32_t done[MAX_SLICES];
  uint32_t pulled[MAX_SLICES];
  uint32_t avail[MAX_SLICES];
  uint32_t ready[MAX_SLICES];
  uint32_t consumed[MAX_SLICES];
  uint32_t error;                     // 0 = fine; else HB_RING_ERR_* of the first wait that expired
  uint32_t error_detail;
  uint32_t pad[4];
};
static_assert(sizeof(RingCtl) == 64, "RingCtl is one 64-byte block");

enum : uint32_t {
  ERR_CONSUMED = 1,   // consumer never pulled our previous result
  ERR_READY = 2,      // source never finished the use we want to pull
  ERR_PULLED = 3,     // local comm warps never delivered the slice
  ERR_TICKET = 4,     // a column-block ticket never arrived
  ERR_ABORTED = 5     // another waiter of this GPU gave up first
};

struct RingArgs {
  uint8_t* peer[HB_MAX_RANKS];             // ring state block of every rank (peer[rank] is local)
  unsigned long long off_ctl;              // RingCtl
  unsigned long long off_buf[MAX_SLICES][2];
  unsigned long long bytes[MAX_SLICES];    // bytes of a slice (multiple of 16)
  int src_rank[MAX_STEPS];                 // source rank of the pull issued at step g
  int rank, world, S, steps;
  uint32_t cnt_base;                       // uses of each slice completed by earlier launches
  uint32_t par_base[MAX_SLICES];           // parity of the buffer holding the first use of this launch
  unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ RingCtl* ctl_of(const RingArgs& r, int rk) {
  return reinterpret_cast<RingCtl*>(r.peer[rk] + r.off_ctl);
}

// Spins until *p >= target (wrap-safe).  Returns false if the wait was abandoned: this waiter's time budget
// ran out (it then publishes `code`), or somebody else on this GPU already did.
template <bool SYS>
__device__ __forceinline__ bool wait_ge(const uint32_t* p, uint32_t target, RingCtl* ctl, uint32_t code,
                                        unsigned long long timeout_ns, uint32_t detail = 0) {
  if ((int32_t)((SYS ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - target) >= 0) return true;
  const unsigned long long t0 = now_ns();
  unsigned spins = 0;
  for (;;) {
    // relaxed polls (an acquire load invalidates the SM's L1 each time), one acquire when the value is there
    if ((int32_t)((SYS ? ld_relaxed_sys(p) : ld_relaxed_gpu(p)) - target) >= 0) {
      (void)(SYS ? ld_acquire_sys(p) : ld_acquire_gpu(p));
      return true;
    }
    __nanosleep(SYS ? 200 : 100);
    if ((++spins & 63u) == 0u) {
      if (ld_relaxed_gpu(&ctl->error) != 0u) return false;
      if (timeout_ns != 0ULL && now_ns() - t0 > timeout_ns) {
        if (atomicCAS(&ctl->error, 0u, code) == 0u) ctl->error_detail = detail;
        __threadfence();
        return false;
      }
    }
  }
}

__device__ __forceinline__ bool aborted(const RingCtl* ctl) { return ld_relaxed_gpu(&ctl->error) != 0u; }

// Body of the communication warp of CTA blockIdx.x (all 32 lanes call it).  Walks the launch's steps and, for
// each, pulls this CTA's stripe of the slice's NEXT use from the step's source rank.
__device__ __forceinline__ void comm_warp_loop(const RingArgs& r) {
  if (r.world <= 0) return;
  const int lane = threadIdx.x & 31;
  RingCtl* ctl = ctl_of(r, r.rank);
  const uint32_t G = gridDim.x;
  for (int g = 0; g < r.steps; ++g) {
    const int k = g % r.S, j = g / r.S;
    const uint32_t u = r.cnt_base + (uint32_t)j;
    const int src = r.src_rank[g];
    int ok = 1;
    if (lane == 0) {
      // 1. the buffer we are about to overwrite held use u-1; its consumer must have copied it (which also
      //    implies that our own CTAs finished training on it)
      ok = wait_ge<true>(&ctl->consumed[k], u, ctl, ERR_CONSUMED, r.timeout_ns, (uint32_t)g) ? 1 : 0;
      // 2. the source has finished use u of this slice
      if (ok) {
        if (blockIdx.x == 0) {
          ok = wait_ge<true>(&ctl_of(r, src)->ready[k], u + 1u, ctl, ERR_READY, r.timeout_ns, (uint32_t)g) ? 1 : 0;
          if (ok) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&ctl->avail[k]), "r"(u + 1u) : "memory");
        } else {
          ok = wait_ge<false>(&ctl->avail[k], u + 1u, ctl, ERR_READY, r.timeout_ns, (uint32_t)g) ? 1 : 0;
        }
      }
    }
    ok = __shfl_sync(0xffffffffu, ok, 0);
    if (!ok) return;
    // 3. this CTA's stripe: source's buf[k][use u] -> local buf[k][use u+1]
    const uint32_t par = r.par_base[k] ^ (uint32_t)(j & 1);
    const uint8_t* sbuf = r.peer[src] + r.off_buf[k][par];
    uint8_t* dbuf = r.peer[r.rank] + r.off_buf[k][par ^ 1u];
    const size_t nvec = (size_t)(r.bytes[k] >> 4);
    const size_t per = (nvec + G - 1) / G;
    const size_t v0 = min(nvec, (size_t)blockIdx.x * per), v1 = min(nvec, v0 + per);
    constexpr int UN = 8;                       // 8 x 512 B in flight per warp
    for (size_t i = v0 + lane; i < v1; i += 32 * UN) {
      int4 tmp[UN];
#pragma unroll
      for (int q = 0; q < UN; ++q)
        if (i + (size_t)q * 32 < v1) tmp[q] = ld_volatile_v4(sbuf + (i + (size_t)q * 32) * 16);
#pragma unroll
      for (int q = 0; q < UN; ++q)
        if (i + (size_t)q * 32 < v1) *reinterpret_cast<int4*>(dbuf + (i + (size_t)q * 32) * 16) = tmp[q];
    }
    __syncwarp();
    // 4. publish: locally "stripe landed"; the last CTA also tells the source its buffer is free
    if (lane == 0) {
      __threadfence();
      const uint32_t old = atomicAdd(&ctl->pulled[k], 1u);
      if (old == G * (u + 1u) - 1u) {
        fence_sys();
        st_release_sys(&ctl_of(r, src)->consumed[k], u + 1u);
      }
    }
  }
}

// Called by ONE thread of a worker before it first touches slice k at step g: the slice's current use
// must have landed completely (nothing to wait for on the very first use, which the host placed).
__device__ __forceinline__ bool wait_slice(const RingArgs& r, int g) {
  if (r.world <= 0) return true;
  const int k = g % r.S;
  const uint32_t u = r.cnt_base + (uint32_t)(g / r.S);
  if (u == 0u) return true;
  RingCtl* ctl = ctl_of(r, r.rank);
  return wait_ge<false>(&ctl->pulled[k], gridDim.x * u, ctl, ERR_PULLED, r.timeout_ns, (uint32_t)g);
}

// Called by ONE thread per CTA after the CTA's last write of step g (and a CTA-wide barrier).
__device__ __forceinline__ void step_done(const RingArgs& r, int g) {
  if (r.world <= 0) return;
  const int k = g % r.S;
  const uint32_t u = r.cnt_base + (uint32_t)(g / r.S);
  RingCtl* ctl = ctl_of(r, r.rank);
  fence_sys();
  const uint32_t old = atomicAdd(&ctl->done[k], 1u);
  if (old == gridDim.x * (u + 1u) - 1u) {
    fence_sys();
    st_release_sys(&ctl->ready[k], u + 1u);
  }
}

__device__ __forceinline__ void* slice_ptr(const RingArgs& r, int g) {
  const int k = g % r.S;
  const uint32_t par = r.par_base[k] ^ (uint32_t)((g / r.S) & 1);
  return r.peer[r.rank] + r.off_buf[k][par];
}

// ---- host side: argument marshalling and the cooperative (+ cluster) launch shared by the epoch kernels ----
inline int fill_ring(RingArgs& r, void* const* peer_state, int my_rank, int world, int S, int steps,
                     unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                     const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                     unsigned long long timeout_ns) {
  memset(&r, 0, sizeof(r));
  HB_REQUIRE(S >= 1 && S <= MAX_SLICES && steps >= 0 && steps % S == 0 && steps <= MAX_STEPS, "ring: bad slice count %d / steps %d", S, steps);
  HB_REQUIRE(world >= 0 && world <= HB_MAX_RANKS, "ring: bad world %d", world);
  HB_REQUIRE(world == 0 || (my_rank >= 0 && my_rank < world), "ring: bad rank %d of %d", my_rank, world);
  const int np = world > 0 ? world : 1;
  for (int i = 0; i < np; ++i) {
    HB_REQUIRE(peer_state[i] != nullptr, "ring: null state pointer for rank %d", i);
    r.peer[i] = reinterpret_cast<uint8_t*>(peer_state[i]);
  }
  r.off_ctl = off_ctl;
  for (int k = 0; k < S; ++k) {
    r.off_buf[k][0] = off_buf4[2 * k]; r.off_buf[k][1] = off_buf4[2 * k + 1];
    r.bytes[k] = bytes2[k];
    r.par_base[k] = par_base2 ? (par_base2[k] & 1u) : 0u;
    HB_REQUIRE((r.off_buf[k][0] & 15) == 0 && (r.off_buf[k][1] & 15) == 0 && (r.bytes[k] & 15) == 0,
               "ring: slice buffers must be 16-byte aligned and a multiple of 16 bytes long");
  }
  for (int g = 0; g < steps; ++g) {
    r.src_rank[g] = (world > 0) ? (src_rank_host ? src_rank_host[g] : (my_rank + world - 1) % world) : 0;
    HB_REQUIRE(world == 0 || (r.src_rank[g] >= 0 && r.src_rank[g] < world), "ring: bad source rank %d at step %d", r.src_rank[g], g);
  }
  r.rank = world > 0 ? my_rank : 0; r.world = world; r.S = S; r.steps = steps;
  r.cnt_base = cnt_base; r.timeout_ns = timeout_ns;
  return 0;
}

template <typename Args>
inline int launch_epoch(const void* fn, int threads, size_t smem, int workers, int cs, Args& a, void* stream) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(workers * cs));
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeCooperative;      // every CTA must be resident: the ticket ring and the ring flags spin
  attr[0].val.cooperative = 1;
  int na = 1;
  if (cs > 1) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = (unsigned)cs; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    na = 2;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  void* args[] = {&a};
  HB_CUDA_OK(cudaLaunchKernelExC(&cfg, fn, args));
  return 0;
}

}  // namespace rot
}  // namespace hb
