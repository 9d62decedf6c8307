// In-kernel collectives over NVLink 5 / NVSwitch peer memory.
//
// Harp's collectives move whole serialised tables over TCP in log N or N-1
// hops (HC/collective/AllreduceCollective.java:150-292, AllgatherCollective
// .java:147-213, RegroupCollective.java:154-226, BcastCollective.java:338-386,
// LocalGlobalSyncCollective.java:710-772).  On an NVSwitch box every GPU
// reaches every peer at full bandwidth, so each op here is ONE kernel per rank
// that loads/stores the peers' symmetric slabs directly:
//
//   allreduce      = barrier · every rank reduces its 1/N slice from all peers
//                    (fixed rank order → bit-deterministic) and stores the
//                    result into every peer · barrier
//   reduce_scatter = same, result kept locally only          (dense regroup)
//   allgather      = every rank stores its slice into every peer
//   broadcast      = every non-root pulls from the root
//   rotate         = every rank pulls its predecessor's block (arbitrary map)
//   reduce         = root reduces all peers' buffers
//
// Synchronisation: per-CTA flag words in the slab, monotonically increasing
// tickets (no reset race): st.release.sys / ld.acquire.sys.  All ranks launch
// the same grid; CTA b only ever handshakes with CTA b of the peers, which is
// sufficient because kernel completion implies completion of all CTAs.
#include "common.cuh"
#include "../include/harp_b200.h"

#include <string.h>

namespace hb {
namespace coll {

constexpr int MAX_RANKS = HB_MAX_RANKS;
constexpr int MAX_CTAS = HB_COLL_MAX_CTAS;

struct Peers {
  uint8_t* base[MAX_RANKS];  // peer-mapped slab base of every rank (index = rank)
  unsigned long long timeout_ns;
};

__device__ __forceinline__ uint32_t* flag_ptr(uint8_t* slab, size_t flags_off, int cta, int src) {
  return reinterpret_cast<uint32_t*>(slab + flags_off) + (size_t)cta * MAX_RANKS + src;
}

// All-to-all handshake of CTA blockIdx.x across ranks.  `ticket` must increase
// with every use.  Called by all threads of the CTA.
// The wait is bounded by `Peers::timeout_ns` of %globaltimer (0 = unbounded; the host default is the reference's
// DATA_MAX_WAIT_TIME, 1800 s, HB_WAIT_TIMEOUT_S overrides).  When it expires the CTA records (ticket, peer) in the
// error words behind the flags, returns false and the kernel skips its data phase: the context stays usable and
// the host raises from SymmetricWorkspace.check().
__device__ __forceinline__ uint32_t* error_words(uint8_t* slab, size_t flags_off) {
  return reinterpret_cast<uint32_t*>(slab + flags_off) + (size_t)MAX_CTAS * MAX_RANKS;
}
__device__ __forceinline__ bool cta_barrier_all(const Peers& P, int rank, int world, size_t flags_off,
                                                uint32_t ticket) {
  __syncthreads();
  int bad = 0;
  if (threadIdx.x < world) {
    const int p = threadIdx.x;
    // the release store is cumulative over what the barrier above ordered before it: the CTA's prior (peer) writes are
    // visible to whoever acquires the flag; a separate fence.sys in front of it only added a second MEMBAR.SYS
    st_release_sys(flag_ptr(P.base[p], flags_off, blockIdx.x, rank), ticket);
    const uint32_t* mine = flag_ptr(P.base[rank], flags_off, blockIdx.x, p);
    if ((int32_t)(ld_acquire_sys(mine) - ticket) < 0) {
      unsigned long long t0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      unsigned spins = 0;
      while ((int32_t)(ld_relaxed_sys(mine) - ticket) < 0) {
        __nanosleep(100);
        if ((++spins & 255u) == 0u && P.timeout_ns != 0ULL) {
          unsigned long long t1;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
          if (t1 - t0 > P.timeout_ns) {
            uint32_t* err = error_words(P.base[rank], flags_off);
            if (atomicCAS(err, 0u, ticket | 0x80000000u) == 0u) err[1] = (uint32_t)p;
            bad = 1;
            break;
          }
        }
      }
      (void)ld_acquire_sys(mine);
    }
  }
  return __syncthreads_or(bad) == 0;
}

template <typename T> struct Vec16 { static constexpr int N = 16 / sizeof(T); T v[16 / sizeof(T)]; };

template <typename T> __device__ __forceinline__ Vec16<T> ld16(const void* p) {
  Vec16<T> r;
  *reinterpret_cast<int4*>(&r) = ld_volatile_v4(p);
  return r;
}
template <typename T> __device__ __forceinline__ void st16(void* p, const Vec16<T>& v) {
  *reinterpret_cast<int4*>(p) = *reinterpret_cast<const int4*>(&v);
}

template <int OP, typename T> __device__ __forceinline__ T apply(T a, T b) {
  if (OP == HB_OP_SUM) return a + b;
  if (OP == HB_OP_MINUS) return a - b;            // folded in ascending rank order: v0 - v1 - ... (Operation.MINUS)
  if (OP == HB_OP_MAX) return a > b ? a : b;
  if (OP == HB_OP_MIN) return a < b ? a : b;
  if (OP == HB_OP_PROD) return a * b;
  return a + b;
}

// mode: 0 = allreduce (store to all), 1 = reduce_scatter (store locally),
//       2 = reduce to root (`root` reduces everything)
template <int OP, typename T>
__global__ void __launch_bounds__(512)
reduce_kernel(Peers P, int rank, int world, size_t off, size_t count, size_t flags_off,
              uint32_t ticket, int mode, int root) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
  constexpr int VN = Vec16<T>::N;
  size_t lo, hi;  // element range this rank reduces
  if (mode == 2) {
    lo = 0; hi = (rank == root) ? count : 0;
  } else {
    const size_t per = ((count + world - 1) / world + VN - 1) / VN * VN;
    lo = (size_t)rank * per; hi = lo + per;
    if (lo > count) lo = count;
    if (hi > count) hi = count;
  }
  const size_t nvec = (hi - lo) / VN;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  for (size_t i = tid; i < nvec; i += nth) {
    const size_t byte = off + (lo + i * VN) * sizeof(T);
    Vec16<T> acc = ld16<T>(P.base[0] + byte);
    for (int p = 1; p < world; ++p) {
      Vec16<T> x = ld16<T>(P.base[p] + byte);
#pragma unroll
      for (int e = 0; e < VN; ++e) acc.v[e] = apply<OP, T>(acc.v[e], x.v[e]);
    }
    if (mode == 0) {
      for (int p = 0; p < world; ++p) st16<T>(P.base[(rank + p) % world] + byte, acc);
    } else {
      st16<T>(P.base[rank] + byte, acc);
    }
  }
  // scalar tail of the range
  for (size_t e = lo + nvec * VN + tid; e < hi; e += nth) {
    const size_t byte = off + e * sizeof(T);
    T acc = *reinterpret_cast<const volatile T*>(P.base[0] + byte);
    for (int p = 1; p < world; ++p)
      acc = apply<OP, T>(acc, *reinterpret_cast<const volatile T*>(P.base[p] + byte));
    if (mode == 0) {
      for (int p = 0; p < world; ++p) *reinterpret_cast<T*>(P.base[p] + byte) = acc;
    } else {
      *reinterpret_cast<T*>(P.base[rank] + byte) = acc;
    }
  }
  (void)cta_barrier_all(P, rank, world, flags_off, ticket + 1);
}

// Byte mover used by allgather / broadcast / rotate.
//   mode 0: allgather — push [off + rank*bytes, +bytes) into every peer (same offset)
//   mode 1: broadcast — pull [off, +bytes) from `root`
//   mode 2: rotate    — pull [src_off, +bytes) of rank `root` into [off, +bytes) locally
__global__ void __launch_bounds__(512)
move_kernel(Peers P, int rank, int world, size_t off, size_t src_off, size_t bytes,
            size_t flags_off, uint32_t ticket, int mode, int root) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  const size_t nvec = bytes / 16;
  if (mode == 0) {
    const uint8_t* src = P.base[rank] + off + (size_t)rank * bytes;
    for (size_t i = tid; i < nvec; i += nth) {
      const int4 v = *reinterpret_cast<const int4*>(src + i * 16);
      for (int p = 1; p < world; ++p) {
        const int q = (rank + p) % world;
        *reinterpret_cast<int4*>(P.base[q] + off + (size_t)rank * bytes + i * 16) = v;
      }
    }
    for (size_t b = nvec * 16 + tid; b < bytes; b += nth) {
      const uint8_t v = src[b];
      for (int p = 1; p < world; ++p) P.base[(rank + p) % world][off + (size_t)rank * bytes + b] = v;
    }
  } else {
    const bool active = (mode == 2) || (rank != root);
    if (active) {
      const uint8_t* src = P.base[root] + (mode == 2 ? src_off : off);
      uint8_t* dst = P.base[rank] + off;
      for (size_t i = tid; i < nvec; i += nth)
        *reinterpret_cast<int4*>(dst + i * 16) = ld_volatile_v4(src + i * 16);
      for (size_t b = nvec * 16 + tid; b < bytes; b += nth) dst[b] = src[b];
    }
  }
  (void)cta_barrier_all(P, rank, world, flags_off, ticket + 1);
}

__global__ void barrier_kernel(Peers P, int rank, int world, size_t flags_off, uint32_t ticket) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
}

// ----------------------------------------------------------------------------
// Fused K-means tail = Harp's regroup -> divide -> allgather in ONE kernel
// (MJ/kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185): rank r owns
// centroid rows [r*rows_per, (r+1)*rows_per); it sums those rows of every
// peer's [sums|counts] slab, divides, and stores the new centroid rows (and the
// summed counts) into every peer's C buffer.
// ----------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
kmeans_regroup_allgather_kernel(Peers P, int rank, int world, size_t acc_off, size_t c_off, int k,
                                int d, size_t flags_off, uint32_t ticket, int keep_empty) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
  const int rows_per = (k + world - 1) / world;
  const int r0 = min(k, rank * rows_per), r1 = min(k, r0 + rows_per);
  const int d4 = d >> 2;  // d % 4 == 0 guaranteed by the host
  const size_t counts_off = acc_off + (size_t)k * d * sizeof(float);
  const size_t total = (size_t)(r1 - r0) * d4;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  for (size_t i = tid; i < total; i += nth) {
    const int row = r0 + (int)(i / d4);
    const int c4 = (int)(i % d4);
    const size_t sbyte = acc_off + ((size_t)row * d + c4 * 4) * sizeof(float);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    float cnt = 0.f;
    for (int p = 0; p < world; ++p) {
      const float4 x = ld_volatile_f4(P.base[p] + sbyte);
      s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
      cnt += ld_volatile_f32(P.base[p] + counts_off + (size_t)row * 4);
    }
    const size_t cbyte = c_off + ((size_t)row * d + c4 * 4) * sizeof(float);
    if (cnt > 0.f) {
      const float inv = 1.f / cnt;
      const float4 c = make_float4(s.x * inv, s.y * inv, s.z * inv, s.w * inv);
      for (int p = 0; p < world; ++p) *reinterpret_cast<float4*>(P.base[(rank + p) % world] + cbyte) = c;
    } else if (!keep_empty) {
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int p = 0; p < world; ++p) *reinterpret_cast<float4*>(P.base[(rank + p) % world] + cbyte) = z;
    }
  }
  (void)cta_barrier_all(P, rank, world, flags_off, ticket + 1);
}


// ----------------------------------------------------------------------------
// Row collectives over a dense table [rows x row_elems] that sits at the same offset of every rank's slab.
// Ownership is Harp's default Partitioner: row (= partition id) p belongs to rank p % world
// (partition/Partitioner.java:36-43).
//   mode 0  regroup      owner sums its rows over all ranks, in rank order, in place        (K2, RegroupCollective
//                        .java:154-226 with a SUM combiner = reduce-scatter by ownership)
//   mode 1  allgather    owner stores its rows into every peer                              (AllgatherCollective.java:
//                        147-213 after a regroup: the second half of `aggregate`)
//   mode 2  push         every rank adds ALL rows of its local table into the owners' global tables with
//                        red.global.add over the peer pointer (K18, LocalGlobalSyncCollective.java:456-549); the
//                        owner's global table holds its rows densely: row p at index p / world
//   mode 3  pull         every rank adds the owners' global rows into its local table (gather over peer
//                        pointers; `pull` combines, callers zero first: :564-698)
//   mode 4  pull, bcast  same result, but the owner PUSHES each of its rows into every peer's local table — the
//                        useBcast branch for rows that all workers want (:654-659)
// ----------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void red_add_t(T* p, T v);
template <> __device__ __forceinline__ void red_add_t<float>(float* p, float v) { red_add_f32(p, v); }
template <> __device__ __forceinline__ void red_add_t<double>(double* p, double v) { red_add_f64(p, v); }

template <typename T>
__global__ void __launch_bounds__(512)
rows_kernel(Peers P, int rank, int world, size_t loc_off, size_t glob_off, long long rows, long long row_elems,
            size_t flags_off, uint32_t ticket, int mode) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  if (mode == 0 || mode == 1 || mode == 4) {
    // iterate over the rows this rank owns
    const long long nown = rows > rank ? (rows - rank + world - 1) / world : 0;
    const size_t total = (size_t)nown * row_elems;
    for (size_t i = tid; i < total; i += nth) {
      const long long p = rank + (long long)(i / row_elems) * world;
      const size_t e = i % row_elems;
      if (mode == 0) {
        const size_t byte = loc_off + ((size_t)p * row_elems + e) * sizeof(T);
        T acc = *reinterpret_cast<const volatile T*>(P.base[0] + byte);
        for (int q = 1; q < world; ++q) acc += *reinterpret_cast<const volatile T*>(P.base[q] + byte);
        *reinterpret_cast<T*>(P.base[rank] + byte) = acc;
      } else if (mode == 1) {
        const size_t byte = loc_off + ((size_t)p * row_elems + e) * sizeof(T);
        const T v = *reinterpret_cast<const T*>(P.base[rank] + byte);
        for (int q = 1; q < world; ++q) *reinterpret_cast<T*>(P.base[(rank + q) % world] + byte) = v;
      } else {
        const T v = *reinterpret_cast<const T*>(P.base[rank] + glob_off + ((size_t)(p / world) * row_elems + e) * sizeof(T));
        const size_t byte = loc_off + ((size_t)p * row_elems + e) * sizeof(T);
        for (int q = 0; q < world; ++q) red_add_t<T>(reinterpret_cast<T*>(P.base[(rank + q) % world] + byte), v);
      }
    }
  } else {
    const size_t total = (size_t)rows * row_elems;
    for (size_t i = tid; i < total; i += nth) {
      const long long p = (long long)(i / row_elems);
      const size_t e = i % row_elems;
      const int owner = (int)(p % world);
      T* loc = reinterpret_cast<T*>(P.base[rank] + loc_off + ((size_t)p * row_elems + e) * sizeof(T));
      uint8_t* glob = P.base[owner] + glob_off + ((size_t)(p / world) * row_elems + e) * sizeof(T);
      if (mode == 2) red_add_t<T>(reinterpret_cast<T*>(glob), *loc);
      else *loc += *reinterpret_cast<const volatile T*>(glob);
    }
  }
  (void)cta_barrier_all(P, rank, world, flags_off, ticket + 1);
}


// ----------------------------------------------------------------------------
// NVLS collectives: the slab is also bound to an NVSwitch multicast object; `mc` is this rank's mapping of it.
// A multimem.ld_reduce on an mc address returns the SUM of that address over every rank's slab, computed in the
// switch; a multimem.st writes every rank's slab at once.  One-shot allreduce: rank r reduces its 1/world slice
// with ld_reduce and multicasts the result with st — two NVSwitch traversals, no N-way peer loads
// (AllreduceCollective.java:150-292).  Synchronisation stays on the peer-to-peer flag barrier above.
//   mode 0 allreduce (f32 sum, in place)   mode 1 broadcast from `root`   mode 2 allgather of `world` equal shards
// ----------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add_v4(const void* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_v4(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(512)
nvls_kernel(Peers P, uint8_t* mc, int rank, int world, size_t off, size_t bytes, size_t flags_off, uint32_t ticket,
            int mode, int root) {
  if (!cta_barrier_all(P, rank, world, flags_off, ticket)) return;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  const size_t nvec = bytes >> 4;
  if (mode == 0) {
    const size_t per = (nvec + world - 1) / world;
    const size_t v0 = min(nvec, (size_t)rank * per), v1 = min(nvec, v0 + per);
    for (size_t i = v0 + tid; i < v1; i += nth) {
      const float4 v = multimem_ld_reduce_add_v4(mc + off + i * 16);
      multimem_st_v4(mc + off + i * 16, v);
    }
  } else if (mode == 1) {
    if (rank == root)
      for (size_t i = tid; i < nvec; i += nth)
        multimem_st_v4(mc + off + i * 16, *reinterpret_cast<const float4*>(P.base[rank] + off + i * 16));
  } else {
    const size_t shard = nvec / world;
    for (size_t i = tid; i < shard; i += nth) {
      const size_t b = off + ((size_t)rank * shard + i) * 16;
      multimem_st_v4(mc + b, *reinterpret_cast<const float4*>(P.base[rank] + b));
    }
  }
  // the multimem stores are ordered before the flags of the closing barrier by its bar.sync + release stores
  (void)cta_barrier_all(P, rank, world, flags_off, ticket + 1);
}

}  // namespace coll
}  // namespace hb

using namespace hb;
using namespace hb::coll;

static unsigned long long g_timeout_ns = 1800ULL * 1000000000ULL;      // DATA_MAX_WAIT_TIME of the reference

static int fill_peers(Peers& P, const void* const* peer_ptrs, int rank, int world) {
  P.timeout_ns = g_timeout_ns;
  HB_REQUIRE(world >= 1 && world <= MAX_RANKS, "world=%d exceeds HB_MAX_RANKS=%d", world, MAX_RANKS);
  HB_REQUIRE(rank >= 0 && rank < world, "bad rank %d of %d", rank, world);
  for (int i = 0; i < MAX_RANKS; ++i)
    P.base[i] = (i < world) ? (uint8_t*)const_cast<void*>(peer_ptrs[i]) : nullptr;
  return 0;
}

static int pick_grid(size_t work_items, int threads) {
  size_t g = (work_items + threads - 1) / threads;
  if (g < 1) g = 1;
  size_t cap = (size_t)hb::sm_count();
  if (cap > MAX_CTAS) cap = MAX_CTAS;
  return (int)(g > cap ? cap : g);
}

template <int OP, typename T>
static int launch_reduce(const Peers& P, int rank, int world, size_t off, size_t count,
                         size_t flags_off, uint32_t ticket, int mode, int root, int grid,
                         cudaStream_t s) {
  reduce_kernel<OP, T><<<grid, 512, 0, s>>>(P, rank, world, off, count, flags_off, ticket, mode, root);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
static int dispatch_op(int op, const Peers& P, int rank, int world, size_t off, size_t count,
                       size_t flags_off, uint32_t ticket, int mode, int root, int grid,
                       cudaStream_t s) {
  switch (op) {
    case HB_OP_SUM: return launch_reduce<HB_OP_SUM, T>(P, rank, world, off, count, flags_off, ticket, mode, root, grid, s);
    case HB_OP_MINUS: return launch_reduce<HB_OP_MINUS, T>(P, rank, world, off, count, flags_off, ticket, mode, root, grid, s);
    case HB_OP_MAX: return launch_reduce<HB_OP_MAX, T>(P, rank, world, off, count, flags_off, ticket, mode, root, grid, s);
    case HB_OP_MIN: return launch_reduce<HB_OP_MIN, T>(P, rank, world, off, count, flags_off, ticket, mode, root, grid, s);
    case HB_OP_PROD: return launch_reduce<HB_OP_PROD, T>(P, rank, world, off, count, flags_off, ticket, mode, root, grid, s);
  }
  hb::set_error("unsupported reduction op %d", op);
  return -2;
}

extern "C" {

// flags + 256 bytes whose first two words report an expired wait: (0x80000000 | ticket, peer rank waited for)
size_t hb_coll_flags_bytes(void) { return (size_t)MAX_CTAS * MAX_RANKS * sizeof(uint32_t) + 256; }

// Bound (ns) on every in-kernel wait for a peer; 0 = wait for ever.
int hb_coll_set_timeout_ns(unsigned long long ns) { g_timeout_ns = ns; return 0; }

// NVLS data path (multimem.ld_reduce / multimem.st through `mc_ptr`, this rank's mapping of the multicast object the
// slabs are bound to).  mode 0: in-place f32 sum allreduce of [offset, +bytes); 1: broadcast from root; 2: allgather
// of `world` equal shards.  bytes must be a multiple of 16 (x world for mode 2).
int hb_coll_nvls(const void* const* peer_ptrs, void* mc_ptr, int rank, int world, size_t offset_bytes, size_t bytes, int mode,
                 int root, size_t flags_offset, uint32_t ticket, void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  HB_REQUIRE(mc_ptr != nullptr, "hb_coll_nvls: no multicast mapping");
  HB_REQUIRE(mode >= 0 && mode <= 2 && root >= 0 && root < world, "hb_coll_nvls: bad mode %d / root %d", mode, root);
  HB_REQUIRE((offset_bytes & 15) == 0 && (bytes & 15) == 0 && (mode != 2 || (bytes / 16) % world == 0),
             "hb_coll_nvls: offset and size must be multiples of 16 bytes (and of 16 x world for allgather)");
  const size_t work = mode == 1 ? bytes / 16 : bytes / 16 / world;
  const int grid = pick_grid(work + 1, 512);
  nvls_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(P, (uint8_t*)mc_ptr, rank, world, offset_bytes, bytes, flags_offset,
                                                      ticket, mode, root);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_coll_rows(const void* const* peer_ptrs, int rank, int world, size_t local_offset_bytes, size_t global_offset_bytes,
                 long long rows, long long row_elems, int dtype, int mode, size_t flags_offset, uint32_t ticket,
                 void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  HB_REQUIRE(mode >= 0 && mode <= 4, "hb_coll_rows: bad mode %d", mode);
  HB_REQUIRE(dtype == HB_F32 || dtype == HB_F64, "hb_coll_rows: float32 / float64 tables only (dtype %d)", dtype);
  HB_REQUIRE(rows >= 0 && row_elems >= 1, "hb_coll_rows: bad shape");
  const int grid = pick_grid((size_t)rows * (size_t)row_elems / (mode == 2 || mode == 3 ? 1 : world) + 1, 512);
  if (dtype == HB_F32)
    rows_kernel<float><<<grid, 512, 0, (cudaStream_t)stream>>>(P, rank, world, local_offset_bytes, global_offset_bytes,
                                                               rows, row_elems, flags_offset, ticket, mode);
  else
    rows_kernel<double><<<grid, 512, 0, (cudaStream_t)stream>>>(P, rank, world, local_offset_bytes, global_offset_bytes,
                                                                rows, row_elems, flags_offset, ticket, mode);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_coll_reduce(const void* const* peer_ptrs, int rank, int world, size_t offset_bytes,
                   size_t count, int dtype, int op, int mode, int root, size_t flags_offset,
                   uint32_t ticket, void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  HB_REQUIRE((offset_bytes & 15) == 0, "hb_coll_reduce: offset must be 16-byte aligned");
  HB_REQUIRE(mode >= 0 && mode <= 2, "hb_coll_reduce: bad mode %d", mode);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t esz = (dtype == HB_F64 || dtype == HB_I64) ? 8 : (dtype == HB_I16 ? 2 : (dtype == HB_I8 ? 1 : 4));
  const size_t per_rank = mode == 2 ? count : (count + world - 1) / world;
  const int grid = pick_grid(per_rank * esz / 16 + 1, 512);
  switch (dtype) {
    case HB_F32: return dispatch_op<float>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
    case HB_F64: return dispatch_op<double>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
    case HB_I32: return dispatch_op<int>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
    case HB_I64: return dispatch_op<long long>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
    case HB_I16: return dispatch_op<short>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
    case HB_I8: return dispatch_op<signed char>(op, P, rank, world, offset_bytes, count, flags_offset, ticket, mode, root, grid, s);
  }
  hb::set_error("hb_coll_reduce: unsupported dtype %d", dtype);
  return -2;
}

int hb_coll_move(const void* const* peer_ptrs, int rank, int world, size_t offset_bytes,
                 size_t src_offset_bytes, size_t bytes, int mode, int root, size_t flags_offset,
                 uint32_t ticket, void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  HB_REQUIRE((offset_bytes & 15) == 0 && (src_offset_bytes & 15) == 0,
             "hb_coll_move: offsets must be 16-byte aligned");
  HB_REQUIRE(mode >= 0 && mode <= 2, "hb_coll_move: bad mode %d", mode);
  HB_REQUIRE(root >= 0 && root < world, "hb_coll_move: bad root/source rank %d", root);
  HB_REQUIRE(mode != 0 || (bytes & 15) == 0 || world == 1,
             "hb_coll_move: allgather shard size must be a multiple of 16 bytes");
  const int grid = pick_grid(bytes / 16 + 1, 512);
  move_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(P, rank, world, offset_bytes, src_offset_bytes,
                                                      bytes, flags_offset, ticket, mode, root);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_coll_barrier(const void* const* peer_ptrs, int rank, int world, size_t flags_offset,
                    uint32_t ticket, void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, rank, world, flags_offset, ticket);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

int hb_kmeans_regroup_allgather(const void* const* peer_ptrs, int rank, int world,
                                size_t acc_offset_bytes, size_t c_offset_bytes, int k, int d,
                                int keep_empty, size_t flags_offset, uint32_t ticket, void* stream) {
  Peers P;
  if (fill_peers(P, peer_ptrs, rank, world) != 0) return -2;
  HB_REQUIRE(d % 4 == 0, "hb_kmeans_regroup_allgather: d must be a multiple of 4");
  HB_REQUIRE((acc_offset_bytes & 15) == 0 && (c_offset_bytes & 15) == 0, "offsets must be 16-byte aligned");
  const int rows_per = (k + world - 1) / world;
  const int grid = pick_grid((size_t)rows_per * (d / 4), 512);
  kmeans_regroup_allgather_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(
      P, rank, world, acc_offset_bytes, c_offset_bytes, k, d, flags_offset, ticket, keep_empty);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- legacy CUDA-IPC plumbing (fallback when torch symmetric memory is unavailable) ----
int hb_ipc_export(const void* dev_ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  HB_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr)));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
int hb_ipc_open(const void* handle64, void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  HB_CUDA_OK(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int hb_ipc_close(void* ptr) {
  HB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}
// Allocates (cudaMalloc) a zeroed slab that can be exported with hb_ipc_export.
int hb_slab_alloc(size_t bytes, void** out_ptr) {
  HB_CUDA_OK(cudaMalloc(out_ptr, bytes));
  HB_CUDA_OK(cudaMemset(*out_ptr, 0, bytes));
  return 0;
}
int hb_slab_free(void* ptr) {
  HB_CUDA_OK(cudaFree(ptr));
  return 0;
}

}  // extern "C"
