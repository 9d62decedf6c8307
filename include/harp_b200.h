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
#define HB_KMEANS_SINGLE_CTA 4 /* use the 1-CTA tcgen05 kernel instead of the CTA-pair (cta_group::2) one */
#define HB_KMEANS_NO_EXT 8     /* CTA-pair kernel: add ‖c‖² in the epilogue instead of folding it into the GEMM */
#define HB_KMEANS_EXT_X128 16  /* CTA-pair kernel with folded ‖c‖²: one 128-column TMEM load per warp per tile (slower; kept for study) */

/* Scratch of the default (CTA-pair) tensor-core path: ‖c‖² padded to a multiple of 256, plus the
 * -‖c‖²/2 GEMM-extension operand (3 tf32 terms per centroid) and a constant tile. */
size_t hb_kmeans_workspace_bytes(int k, int d);
/* Smallest workspace hb_kmeans_step accepts (‖c‖² only); with it the single-CTA kernel is used. */
size_t hb_kmeans_min_workspace_bytes(int k, int d);
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

/* ---- Fused K-means iteration: ONE launch per Lloyd iteration, collective included ----
 *
 * The launch that runs the tcgen05 distance GEMM, the arg-min and the partial-sum accumulation also
 * performs the reference's regroup -> centroid update -> allgather
 * (MJ/kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185) over NVLink peer memory and
 * prepares the GEMM operands of the next iteration.  Every rank owns a "fused state" block of
 * hb_kmeans_fused_state_bytes(k, d) bytes, 256-byte aligned, at addresses the peers can load/store
 * (symmetric memory / CUDA IPC); on one GPU any device buffer works.
 *
 *   hb_kmeans_fused_offsets -> out8 = byte offsets inside the block of
 *       { C[0], C[1], acc[0], acc[1], obj terms (2 doubles), obj total (1 double), ext[0], ext[1] };
 *       iteration e reads C[e & 1], writes [sums | counts] to acc[e & 1] and the new centroids to C[~e & 1].
 *   hb_kmeans_fused_init    zeroes the block and loads the initial centroids C0 (device pointer or NULL) into
 *       C[0].  All ranks must have completed it (host-side barrier) before anyone launches the first step.
 *   hb_kmeans_fused_prep    rebuilds the operands of C[parity] after the caller overwrote that buffer.
 *   hb_kmeans_fused_step    one iteration; `epoch` = 0, 1, 2, ... must be the same on every rank.  When the
 *       launch completes on a rank, its C[~epoch & 1] is whole and the objective total (sum over ranks of the
 *       sum of min squared distances to the INPUT centroids) is in the block.  Needs hb_kmeans_tc_supported(d, k);
 *       the grid is one CTA per SM, launched cooperatively because its tail synchronises the CTAs of the launch. */
size_t hb_kmeans_fused_state_bytes(int k, int d);
int hb_kmeans_fused_offsets(int k, int d, unsigned long long* out8);
int hb_kmeans_fused_init(void* state, int k, int d, const float* C0, void* stream);
int hb_kmeans_fused_prep(void* state, int k, int d, int parity, void* stream);
int hb_kmeans_fused_step(const float* X, long long n, int d, int k, void* const* peer_state, int rank, int world,
                         unsigned epoch, int keep_empty, int* assign, void* stream);

/* Same pass as hb_kmeans_step for callers that cannot provide scratch memory: programs that
 * come in through this C ABI from a runtime with no CUDA allocator of its own (JNI/FFI hosts),
 * or whose device memory is already fully claimed by another pool, so a cudaMalloc here would
 * fail or stall their streams.  The scratch lives in a fixed device array inside the library
 * (held on every GPU that loads it until exit), which limits k to HB_KMEANS_STATIC_MAX_K and
 * (d, k) to what hb_kmeans_tc_supported accepts; anything else is refused with an error.
 * All calls on a GPU share that array: calls that overlap in time on one GPU race — serialise
 * them.  Each call rewrites what it uses, so leftovers from an earlier call do not matter.
 * The first call on each GPU looks up the array's device address; make it at start-up, before
 * any stream capture.  Allocates nothing.  (No counterpart in the reference.) */
#define HB_KMEANS_STATIC_MAX_K 16384
int hb_kmeans_step_static(const float* X, long long n, int d, const float* C, int k, float* sums,
                          float* counts, int* assign, double* obj, void* stream);

/* ---- Gram / moments (csrc/gram.cu) — local step of the DAAL-style distributed algorithms
 * (ml/daal/.../daal_cov/densedistri/COVDaalCollectiveMapper.java:143-152,
 *  daal_pca/cordensedistr/PCADaalCollectiveMapper.java:121-127,
 *  daal_linreg/normaleq/LinRegDaalCollectiveMapper.java:162-167, daal_mom/...) ------------- */
#define HB_GRAM_FORCE_SIMT 1
#define HB_GRAM_NO_ZERO 2
/* 1 when d is handled by the tcgen05 path (d % 32 == 0). */
int hb_gram_tc_supported(int d);
/* G[d,d] (+)= X^T X for row-major fp32 X[n,d]; TF32 products, fp32 accumulate; full symmetric output. */
int hb_gram(const float* X, long long n, int d, float* G, int flags, void* stream);
/* X^T X of this rank's rows AND the allreduce of the ranks' partial results in ONE launch (DAAL's
 * "partial result -> master -> broadcast", HD/data_comm/HarpDAALComm.java:203-246, done over NVLink peer memory
 * by the kernel that produced the partial).  Every rank owns a 256-byte-aligned, peer-addressable block of
 * hb_gram_fused_state_bytes(d) bytes, zero-filled once before epoch 0 (then barrier across ranks); the global
 * d x d result is left at byte offset 256 of every rank's block.  epoch = 0, 1, 2, ... identical on all ranks.
 * Needs hb_gram_tc_supported(d) and n < 2^31. */
size_t hb_gram_fused_state_bytes(int d);
int hb_gram_allreduce_fused(const float* X, long long n, int d, void* const* peer_state, int rank, int world,
                            unsigned epoch, void* stream);
/* K14: out[(m+1), d] (double, accumulated: zero first): row 0 += column sums of X, row 1+j += sum_r Y[r,j] X[r,:]; one
 * streaming pass (MD/daal_linreg/normaleq/LinRegDaalCollectiveMapper.java:162-167).  d % 4 == 0, m <= 8. */
int hb_xty(const float* X, long long n, int d, const float* Y, int m, double* out, void* stream);
/* out[r,:] = X[r,:] - shift[:] (centring step of the two-pass covariance; d % 4 == 0). */
int hb_center_rows(const float* X, long long n, int d, const float* shift, float* out, void* stream);
/* Per-column sum / sum of squares (double) and min / max (float) of X[n,d]. */
int hb_moments(const float* X, long long n, int d, double* sum, double* sumsq, float* mn, float* mx,
               int flags, void* stream);

/* ---- Matrix factorisation (csrc/mf.cu) — SGDMPTask.doRun (MJ/sgd/SGDMPTask.java:46-77), the
 * dymoro block Scheduler (MJ/dymoro/Scheduler.java:95-237), RMSETask (MJ/sgd/RMSETask.java:67-103),
 * CCDMPTask (MJ/ccd/CCDMPTask.java:67-125) -------------------------------------------------- */
/* How many blocks per side one cooperative launch can keep resident for this rank. */
int hb_sgd_max_blocks(int rank);
/* One SGD pass over a panel bucketed into B x B blocks (block (i,c) = runs
 * blk_seg_ptr[i*B+c] .. blk_seg_ptr[i*B+c+1]); run s covers ratings seg_ptr[s]..seg_ptr[s+1] of
 * column seg_col[s].  col_ticket[B] must hold `ticket_base` on entry and holds ticket_base + B on
 * exit.  deadline_ns (time budget from kernel entry, 0 = none) bounds the pass as the reference's
 * Timer does; *trained += ratings updated. */
int hb_sgd_pass(float* W, float* H, const int* blk_seg_ptr, const int* seg_col, const int* seg_ptr,
                const int* r_row, const float* r_val, int B, int rank, float eps, float lambda,
                unsigned int* col_ticket, unsigned int ticket_base, unsigned long long deadline_ns,
                unsigned long long* trained, void* stream);
/* Same pass with the reference's dynamic block scheduling (dymoro/Scheduler.java:95-237): B row-blocks (one CTA
 * each) x C >= B column-blocks (C <= 2048); a CTA takes ANY free column-block it has not visited yet.
 * blk_seg_ptr is [B*C + 1]; col_lock[C] must be all zero on entry and is all zero again on exit. */
int hb_sgd_pass_dynamic(float* W, float* H, const int* blk_seg_ptr, const int* seg_col, const int* seg_ptr,
                        const int* r_row, const float* r_val, int B, int C, int rank, float eps, float lambda,
                        unsigned int* col_lock, unsigned long long deadline_ns, unsigned long long* trained,
                        void* stream);
/* Writes the device wall clock (ns) to *out_dev. */
int hb_device_time_ns(unsigned long long* out_dev, void* stream);
/* *sq_err += sum (v - w.h)^2 over a rating list. */
int hb_mf_sq_err(const float* W, const float* H, const int* rows, const int* cols, const float* vals,
                 long long n, int rank, double* sq_err, void* stream);
/* CCD++ rank-one update of one feature column Wf[nrows] against Hf, residuals updated in place. */
int hb_ccd_update(const int* ptr, const int* idx, float* resid, const int* resid_perm, float* Wf,
                  const float* Hf, int nrows, float lambda, void* stream);

/* ---- LDA collapsed Gibbs sampling (csrc/lda.cu) — LDAMPTask.doRun/sample
 * (MJ/lda/LDAMPTask.java:85-331), topic sums (MJ/lda/LDAMPCollectiveMapper.java:442-475) ------- */
int hb_lda_max_blocks(int K);
/* One sampling pass over a token panel bucketed B x B like the SGD panels (documents = rows, words =
 * columns).  n_dk[docs,K], n_wk[words_slice,K] are dense int32 count tables updated in place with
 * atomics, tok_z holds the assignments, commons[k] = 1/(n_k + V beta) is frozen for the pass. */
int hb_lda_pass(int* n_dk, int* n_wk, const float* commons, const int* blk_seg_ptr, const int* seg_word,
                const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K, float alpha, float beta,
                unsigned long long seed, unsigned int* col_ticket, unsigned int ticket_base,
                unsigned long long deadline_ns, unsigned long long* trained, void* stream);
/* Same pass with the document-topic table held as unsigned 16-bit counts (every document shorter than 65536
 * tokens; table and rows 8-byte aligned): the kernel is bound by one document-row read per token, so this halves
 * its traffic, and the next token's row is fetched while the current token is sampled. */
int hb_lda_pass_dk16(unsigned short* n_dk16, int* n_wk, const float* commons, const int* blk_seg_ptr,
                     const int* seg_word, const int* seg_ptr, const int* tok_doc, int* tok_z, int B, int K,
                     float alpha, float beta, unsigned long long seed, unsigned int* col_ticket,
                     unsigned int ticket_base, unsigned long long deadline_ns, unsigned long long* trained,
                     void* stream);
/* sums[k] += sum over words of n_wk[w][k] (int64). */
int hb_lda_topic_sums(const int* n_wk, long long words, int K, long long* sums, void* stream);
/* counts[idx0[t] * K + z[t]] += 1 for every token. */
int hb_lda_count(const int* idx0, const int* z, long long n, int K, int* counts, void* stream);

/* ---- SVM training (csrc/svm.cu) — DAAL svm::training::Batch (Boser) behind
 * MD/daal_svm/MultiClassDenseBatch/SVMDaalCollectiveMapper.java:150-199 ------------------------ */
/* The whole SMO loop (maximal-violating-pair rule) of one binary C-SVC problem in ONE launch of one CTA:
 * K[n,n] row-major symmetric fp64 kernel matrix, y in {-1,+1}; outputs alpha[n], grad[n] (= Q alpha - e) and the
 * number of steps taken (*iters_out, device int, may be NULL).  Stream-ordered, nothing is read back; independent
 * problems (one-vs-one pairs) can run on separate streams. */
int hb_smo_binary(const double* K, const double* y, int n, double C, double tol, int max_iter, double* alpha,
                  double* grad, int* iters_out, void* stream);

/* ---- NVLink peer-memory collectives (csrc/coll.cu) ------------------------
 * Replace the TCP collectives of HC/collective/<Op>Collective.java.  Every rank passes the
 * same table `peer_ptrs[world]` of peer-mapped slab base addresses (index =
 * rank; entry `rank` is the local slab), the same byte offsets into the slab,
 * and the same `ticket`.  Tickets must grow by at least 2 per call on a given
 * flags area (each call consumes `ticket` and `ticket + 1`).  The flags area
 * (hb_coll_flags_bytes() bytes at `flags_offset` inside every slab) must be
 * zeroed once before first use.  All ranks must issue the calls in the same
 * order; waits for a peer are bounded by hb_coll_set_timeout_ns. */
#define HB_MAX_RANKS 16
#define HB_COLL_MAX_CTAS 256
enum { HB_F32 = 0, HB_F64 = 1, HB_I32 = 2, HB_I64 = 3, HB_I16 = 4, HB_I8 = 5 };   /* the six array types of HC/resource */
enum { HB_OP_SUM = 0, HB_OP_MINUS = 1, HB_OP_PROD = 2, HB_OP_MAX = 3, HB_OP_MIN = 4 };

size_t hb_coll_flags_bytes(void);
/* Bound, in ns of %globaltimer, on every in-kernel wait for a peer (default 1800 s like the reference's
 * DATA_MAX_WAIT_TIME, HC/io/Constant.java:35; 0 = unbounded).  An expired wait does not trap: the kernel records
 * (0x80000000 | ticket, peer) in the two words at flags_offset + MAX_CTAS * MAX_RANKS * 4 and skips its data phase. */
int hb_coll_set_timeout_ns(unsigned long long ns);
/* Row collectives on a dense [rows, row_elems] table at the same offset of every slab; row p is owned by rank
 * p % world (HC/partition/Partitioner.java:36-43).  mode 0: regroup = owner sums its rows over all ranks
 * (HC/collective/RegroupCollective.java:154-226); 1: owner stores its rows into every peer (allgather half of
 * aggregate, :252-297); 2: push = red.add of every local row into the owner's global table, whose row for id p is at
 * index p / world (HC/collective/LocalGlobalSyncCollective.java:456-549); 3: pull = local row += owner's global row
 * (:564-698); 4: pull with useBcast = owners add their rows into every peer's local table (:654-659). */
int hb_coll_rows(const void* const* peer_ptrs, int rank, int world, size_t local_offset_bytes,
                 size_t global_offset_bytes, long long rows, long long row_elems, int dtype, int mode,
                 size_t flags_offset, uint32_t ticket, void* stream);
/* mode 0: allreduce in place (HC/collective/AllreduceCollective.java:150-292)
 * mode 1: reduce-scatter — rank r ends with the reduced r-th slice in place
 *         (dense regroup, HC/collective/RegroupCollective.java:154-226)
 * mode 2: reduce to `root` (HC/collective/ReduceCollective.java:150-336) */
int hb_coll_reduce(const void* const* peer_ptrs, int rank, int world, size_t offset_bytes,
                   size_t count, int dtype, int op, int mode, int root, size_t flags_offset,
                   uint32_t ticket, void* stream);
/* mode 0: allgather — shard `rank` of size `bytes` at offset+rank*bytes is pushed to all peers
 *         (HC/collective/AllgatherCollective.java:147-213)
 * mode 1: broadcast of [offset, +bytes) from `root` (HC/collective/BcastCollective.java:338-386)
 * mode 2: rotate — pull [src_offset, +bytes) of rank `root` into local [offset, +bytes)
 *         (HC/collective/LocalGlobalSyncCollective.java:710-772) */
int hb_coll_move(const void* const* peer_ptrs, int rank, int world, size_t offset_bytes,
                 size_t src_offset_bytes, size_t bytes, int mode, int root, size_t flags_offset,
                 uint32_t ticket, void* stream);
/* (HC/collective/Communication.java:61-183) */
int hb_coll_barrier(const void* const* peer_ptrs, int rank, int world, size_t flags_offset,
                    uint32_t ticket, void* stream);
/* Fused regroup -> divide -> allgather of K-means partial sums
 * (MJ/kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185): reads every
 * peer's [sums(k*d) | counts(k)] at acc_offset, writes new centroids at c_offset
 * of every peer. */
int hb_kmeans_regroup_allgather(const void* const* peer_ptrs, int rank, int world,
                                size_t acc_offset_bytes, size_t c_offset_bytes, int k, int d,
                                int keep_empty, size_t flags_offset, uint32_t ticket, void* stream);

/* CUDA-IPC plumbing used when torch symmetric memory is unavailable.
 * hb_slab_alloc is the only function of this library that allocates device memory. */
int hb_ipc_export(const void* dev_ptr, void* handle64);
int hb_ipc_open(const void* handle64, void** out_ptr);
int hb_ipc_close(void* ptr);
int hb_slab_alloc(size_t bytes, void** out_ptr);
int hb_slab_free(void* ptr);

/* ---- split-precision K-means (fp32-faithful tensor-core mode) ----
 * prec 0: TF32; 1: fp16 hi/lo halves of x * scale (scale a power of two); 2: bf16 hi/lo halves.
 * Three kind::f16 MMAs per product (hi*hi + lo*hi + hi*lo), exact two-pass arg-max.
 * (MJ/kmeans/regroupallgather/CenCalcTask.java:67-100) */
size_t hb_kmeans_split_row_bytes(int d);
int hb_kmeans_split_rows(const float* X, long long n, int d, float scale, int prec, void* Xs, void* stream);
int hb_kmeans_fused_prep2(void* state, int k, int d, int parity, int prec, float scale, void* stream);
int hb_kmeans_fused_step2(const float* X, const void* Xs, long long n, int d, int k, void* const* peer_state, int rank,
                          int world, unsigned epoch, int keep_empty, int* assign, int prec, float scale, void* stream);

/* ---- whole-epoch rotation kernels (csrc/rotate_ring.cuh) ----
 * One cooperative launch trains `steps` panels; the model slices rotate inside the kernel: a communication warp per
 * CTA pulls the next slice from the symmetric slab of src_rank[step] while the other warps train
 * (MJ/dymoro/Rotator.java:53-67, RotateTask.java:107-124,158-210).  peer_state, off_ctl, off_buf4, bytes2, src_rank, cnt_base,
 * par_base2 and timeout_ns come from harp_b200.parallel.ring.RingState.ring_args(); world 0 = no ring, 1 = loop-back.
 * Bounded waits record a code in the ring control block (words 10, 11) and drain instead of trapping. */
int hb_ring_ctl_bytes(void);
int hb_sgd_epoch_max_workers(int rank, int cs);
int hb_sgd_epoch_trace(unsigned long long* buf);
int hb_sgd_epoch(float* W, const void* panels_host, int npanels, const int* panel_of_step_host, int steps, int S,
                 int workers, int cs_and_flags, int rank, float eps, float lambda, unsigned ticket_epoch,
                 unsigned long long deadline_ns, unsigned long long* trained, void* const* peer_state, int my_rank,
                 int world, unsigned long long off_ctl, const unsigned long long* off_buf4,
                 const unsigned long long* bytes2, const int* src_rank_host, unsigned cnt_base,
                 const unsigned* par_base2, unsigned long long timeout_ns, void* stream);
int hb_lda_epoch_supported(int K, int dk16);
int hb_lda_epoch_max_workers(int K, int dk16, int cs);
int hb_lda_epoch(void* n_dk, int dk16, const float* commons, const void* panels_host, int npanels,
                 const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha, float beta,
                 unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                 unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                 unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                 const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                 unsigned long long timeout_ns, void* stream);
/* Token-list (SparseLDA-style) sampler, K up to 16384: the doc side is the document's own topic list
 * (MJ/lda/LDAMPTask.java:85-132,209-254). */
int hb_lda_tl_supported(int K);
int hb_lda_tl_max_workers(int K, int cs);
int hb_lda_tl_epoch(int* doc_z, const int* d_ptr, const float* commons, const void* panels_host, int npanels,
                    const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha,
                    float beta, unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                    unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                    unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                    const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                    unsigned long long timeout_ns, void* stream);
/* hb_lda_tl_epoch2: the same with exact per-32-topic block counts kept behind the word-topic rows of every slice
 * ([words x K | words x ceil(K/32)], wb_off = words x K elements): a run reads only the non-empty blocks of its row. */
int hb_lda_tl_epoch2(int* doc_z, const int* d_ptr, const float* commons, const void* panels_host, int npanels,
                     const int* panel_of_step_host, int steps, int S, int workers, int cs, int K, float alpha,
                     float beta, unsigned long long seed, unsigned ticket_epoch, unsigned long long deadline_ns,
                     unsigned long long* trained, void* const* peer_state, int my_rank, int world,
                     unsigned long long off_ctl, const unsigned long long* off_buf4, const unsigned long long* bytes2,
                     const int* src_rank_host, unsigned cnt_base, const unsigned* par_base2,
                     unsigned long long timeout_ns, unsigned long long wb_off, void* stream);
/* sum over nonzero counts of lgamma(count + add), fp64, plus the number of nonzeros
 * (MJ/lda/CalcLikelihoodTask.java:57-77). */
int hb_lgamma_sum(const int* counts, long long n, double add, double* out, unsigned long long* nnz, void* stream);

/* One CCD++ sweep of one feature over n owned ids in a single pass (residual update included; the other residual
 * order is brought up to date on the fly through pd_other / pd_self).  (MJ/ccd/CCDMPTask.java:67-125) */
int hb_ccd_sweep(const int* ptr, const int* idx, float* res, const float* pd_other, const float* pd_self,
                 const float* other, float* self, float* dself, int n, float lambda, int wide, void* stream);

/* Gram + allreduce + column statistics of this rank's rows in one launch (sum == NULL: no statistics). */
int hb_gram_allreduce_fused_stats(const float* X, long long n, int d, void* const* peer_state, int rank, int world,
                                  unsigned epoch, double* sum, double* sumsq, float* mn, float* mx, void* stream);

/* Colour-coding DP step of subgraph counting: out[v, c] = sum_j act[v, split_a[c, j]] * (sum over neighbours u of v of
 * pas[u, split_p[c, j]]); compressed tables (one column per colour set of the sub-template's size).
 * (MJ/subgraph/colorcount_HJ.java:861-925) */
int hb_subgraph_combine(const long long* nbr_ptr, const int* nbr_idx, const float* act, const float* pas,
                        const short* split_a, const short* split_p, int na, int np, int nC, int nsplit, float* out,
                        int n, void* stream);

/* WDA-MDS: rows of B(X) X and the stress numerator / denominator in one pass over the local distance rows
 * (MJ/wdamds/BCCalcTask.java, StressCalcTask.java:44-101); dt = max(D - shift, 0), dim <= 8. */
int hb_mds_bc_stress(const float* D, const long long* ids, const float* X, int m, int n, int dim, float shift, float* BX,
                     double* stress, void* stream);

/* One PageRank propagation over m local source pages (contrib PageRankMapper.java): out[t] += pr[s] / outdeg(s),
 * *dangling += pr of pages without out-links. */
int hb_pagerank_push(const long long* ptr, const int* idx, const long long* src_ids, const double* pr, double* out,
                     double* dangling, int m, void* stream);

/* ---- dense linear algebra (csrc/linalg.cu) ---- */
int hb_tsqr_pad(int d);
size_t hb_tsqr_workspace_bytes(int d);
int hb_tsqr(const float* X, long long n, int d, long long ld, float* work, float* R_out, void* stream);
int hb_chol_pad(int f);
int hb_potrs_batched(const float* A, const float* B, float* X, float* L, int f, long long batch, int* info,
                     void* stream);
int hb_als_solve(const long long* ptr, const int* idx, const float* val, const float* Y, const float* YtY, float lam,
                 float alpha, const int* rows, int n_rows, int f, float* Xout, int* counter, int* info, void* stream);

/* ---- NVLink SHARP (multimem) and the VMM slab behind it (csrc/nvls.cu) ----
 * hb_coll_nvls: mode 0 allreduce (multimem.ld_reduce + multimem.st, f32 sum), 1 broadcast from `root`, 2 allgather.
 * The slab is a cuMemCreate allocation exported as a POSIX fd, mapped by every rank, and bound to a multicast object. */
int hb_coll_nvls(const void* const* peer_ptrs, void* mc_ptr, int rank, int world, size_t offset_bytes, size_t bytes,
                 int mode, int root, size_t flags_offset, uint32_t ticket, void* stream);
int hb_nvls_probe(int world, size_t bytes, char* msg, int msg_len);
int hb_vmm_granularity(int world, size_t* out);
int hb_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out);
int hb_vmm_import(int fd, unsigned long long* handle_out);
int hb_vmm_map(unsigned long long handle, size_t bytes, size_t align, void** ptr_out);
int hb_vmm_unmap(void* ptr, size_t bytes);
int hb_vmm_release(unsigned long long handle);
int hb_mc_create(int world, size_t bytes, unsigned long long* handle_out, int* fd_out);
int hb_mc_add_device(unsigned long long mc_handle);
int hb_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* HARP_B200_H_ */
