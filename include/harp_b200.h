/* harp_b200 — C ABI of the native B200 (sm_100a) library.
 *
 * Every entry point takes raw device pointers plus a CUDA stream (passed as
 * void*; NULL = legacy default stream) and returns 0 on success.  On failure
 * it returns a negative code and hb_last_error() describes what went wrong
 * (thread-local string).  Nothing here allocates device memory unless the
 * comment on the function says so: callers own every buffer, including the
 * scratch ("workspace") some kernels need.
 *
 * Reference citations are paths inside chathurawidanage/harp
 * (HC = core/harp-collective/src/main/java/edu/iu/harp,
 *  MJ = ml/java/src/main/java/edu/iu).
 */
#ifndef HARP_B200_H_
#define HARP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------ */
const char* hb_last_error(void);
int hb_version(void);
int hb_device_sm_count(void);
/* Overwrites `bytes` of `buf` (should exceed the 126 MB L2) so the next timed
 * kernel starts from a cold L2.  Benchmark hygiene helper. */
int hb_l2_flush(void* buf, size_t bytes, void* stream);

/* ---- K-means (MJ/kmeans/regroupallgather/CenCalcTask.java:67-100,
 *      contrib/.../kmeans/allreduce/KmeansMapper.java:173-215) ------------- */
#define HB_KMEANS_FORCE_SIMT 1 /* use the exact-fp32 SIMT kernel */
#define HB_KMEANS_NO_ZERO 2    /* accumulate into sums/counts/obj instead of zeroing them */

/* Scratch the tensor-core path needs: ‖c‖² padded to a multiple of 256. */
size_t hb_kmeans_workspace_bytes(int k, int d);
/* Scratch for the SIMT path (adds per-point assignment + distance). */
size_t hb_kmeans_simt_workspace_bytes(long long n, int k, int d);
/* 1 when (d, k) is handled by the tcgen05 path (d % 4 == 0, d <= 128). */
int hb_kmeans_tc_supported(int d, int k);

/* One assignment pass over X[n,d] (row-major fp32) against C[k,d]:
 *   sums[k,d] (+)= Σ points assigned to each centroid, counts[k] (+)= how many,
 *   assign[n] (optional) = arg-min centroid, *obj (optional, double) (+)= Σ min ‖x-c‖².
 * workspace: 256-byte aligned device scratch of hb_kmeans_workspace_bytes(). */
int hb_kmeans_step(const float* X, long long n, int d, const float* C, int k, float* sums,
                   float* counts, int* assign, double* obj, void* workspace,
                   size_t workspace_bytes, int flags, void* stream);

/* C[j,:] = sums[j,:] / counts[j]; empty clusters become the zero vector
 * (MJ/kmeans/regroupallgather/KMeansCollectiveMapper.java:170-183) unless
 * keep_empty != 0, in which case C[j,:] is left untouched. */
int hb_kmeans_finalize(const float* sums, const float* counts, int k, int d, float* C,
                       int keep_empty, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HARP_B200_H_ */
