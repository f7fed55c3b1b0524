/*
 * lmi_b200.h -- C-ABI of the B200-native LMI query-time hot path (liblmi_b200.so).
 *
 * The reference (TerkaSlan/sisap23-laion-challenge-learned-index) is pure Python and has no
 * FFI of its own; its boundary for this path is three Python methods.  Every entry point
 * below names the reference interface it replaces (file:line under search/li/), and
 * INTEGRATION.md shows the ctypes binding a reference maintainer would add.
 *
 *   reference                                                     -> this library
 *   NeuralNetwork.predict_proba   model.py:214-229                -> lmi_navigate
 *   NeuralNetwork.predict         model.py:201-212                -> lmi_predict
 *   LearnedIndex.search           LearnedIndex.py:22-101          -> lmi_search / lmi_search_device
 *   LearnedIndex.search_single    LearnedIndex.py:103-195         -> lmi_search_rank
 *   groupby('category') + .loc    LearnedIndex.py:143,152,168     -> lmi_store_* (done once, not per rank)
 *
 * Conventions: plain pointers and sizes only.  "host" pointers are ordinary process memory,
 * "dev" pointers are CUDA device memory on the index's device.  All functions return 0 on
 * success or a negative lmi_status; lmi_last_error() gives a thread-local message.  One
 * search may be in flight per index; calls on one index must be serialised by the caller
 * (the reference is single-threaded and not re-entrant either, LearnedIndex.py:67).
 * Results: ids are the reference's 1-based DataFrame labels (search.py:71-72) = id_base + row;
 * never-written slots are (id 0, dist 10000.0) exactly like LearnedIndex.py:137-138.
 */
#ifndef LMI_B200_H
#define LMI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lmi_index lmi_index;

typedef enum {
  LMI_OK = 0,
  LMI_ERR_ARG = -1,      /* bad argument (the reference would raise AssertionError / numpy error) */
  LMI_ERR_STATE = -2,    /* e.g. search before the MLP / store is set (LearnedIndex.py:60) */
  LMI_ERR_CUDA = -3,     /* CUDA runtime / driver failure */
  LMI_ERR_NOMEM = -4
} lmi_status;

/* shard policies for multi-GPU (one process per GPU; see DESIGN.md "multi-GPU") */
#define LMI_SHARD_BUCKETS 0   /* whole buckets per rank, longest-processing-time assignment */
#define LMI_SHARD_STRIPES 1   /* every bucket split in `world` contiguous stripes */

/* rescoring kernels (option "rescore") */
#define LMI_RESCORE_SIMT 0    /* CUDA-core reference kernel (cross-check in tests) */
#define LMI_RESCORE_TC   1    /* tcgen05 kernel, query tile resident in TMEM, objects streamed by TMA, one CTA per tile */
#define LMI_RESCORE_TC_SS 2   /* tcgen05 kernel, both operands streamed through shared memory (d > 768 fallback) */
#define LMI_RESCORE_TC_PAIR 3 /* tcgen05 cta_group::2 kernel: query-tile pairs on 2-CTA clusters, each CTA holds half of
                                 every object tile (default) */

typedef struct {
  /* shape of the last search */
  int32_t nq, n_buckets, k, kprime;
  int64_t n_objects_local;       /* objects held by this index (this rank) */
  int64_t n_work_items;          /* (bucket, query tile, object chunk) items executed */
  int64_t pairs;                 /* sum_b m_b * n_b  (query-object pairs scored, no padding) */
  int64_t bucket_bytes;          /* 2*d*sum_{b: m_b>0} n_b  (fp16 bytes of every probed bucket, once) */
  int64_t n_partial_lists;       /* partial top-k' lists written by the rescoring kernel */
  int64_t n_uncertified;         /* queries whose exactness certificate failed and that were therefore re-derived exactly
                                    (float64 over every probed object) by the fallback kernel */
  double  max_abs_score_err;     /* max |fp32 cosine - fp64 cosine| over refined candidates */
  /* device time of the last search, milliseconds (CUDA events on the search stream) */
  float ms_navigate, ms_plan, ms_rescore, ms_finalize, ms_total;
  /* mean rescoring-kernel duration over the searches issued since the previous lmi_get_stats call
   * (at most the last 256), and how many went into it; option "timing" = 1 */
  float ms_rescore_avg;
  int32_t n_rescore_timed;
  int64_t padded_pairs;          /* sum over work items of 128 x ceil64(objects): pairs the MMA tiles compute, padding included */
  float k3_sm_mhz;               /* SM clock during the last rescoring kernel (clock64 / globaltimer on CTA 0; tcgen05 kernel only) */
  int32_t chunk_objects;         /* objects per work item used by the last search (option "chunk_objects", 0 = automatic) */
} lmi_stats;

const char* lmi_last_error(void);
const char* lmi_version(void);

/* ---- lifetime ------------------------------------------------------------------------ */
/* d_nav: navigation dimension (96, pca96); d: rescoring dimension (768, clip768; multiple
 * of 64, <= 1024); n_categories: leaf buckets C.  Binds to CUDA device `device`. */
int lmi_create(lmi_index** out, int device, int d_nav, int d, int n_categories);
void lmi_destroy(lmi_index* idx);

/* name/value tunables: "rescore" (LMI_RESCORE_*), "chunk_objects" (0 = chosen per search), "kprime_extra",
 * "epi_sets" (CTA-pair kernel: epilogue warp sets, 0 = automatic), "item_order" (CTA-pair work list: 0 = full pairs
 * first, 1 = bucket-major, 2 = experiment: 0 + three-tile buckets as three half pairs), "l2_hint" (CTA-pair kernel,
 * experiment: 1 = a chunk's last reader loads with L2 evict_first), "lockstep" (CTA-pair kernel, experiment: 1 = a half
 * pair follows its chunk's full pair tile by tile; measured slower, default 0), "timing" (1 = record per-stage CUDA events). */
int lmi_set_option(lmi_index* idx, const char* name, int64_t value);

/* multi-GPU: this index holds shard `rank` of `world`.  Call before lmi_store_begin. */
int lmi_set_shard(lmi_index* idx, int rank, int world, int policy);

/* ---- model: replaces the torch Sequential in li.model.Model (model.py:15-83) ------------ */
/* n_layers nn.Linear layers with ReLU between them; W[l] is [out_dims[l], in] row-major
 * fp32 exactly as nn.Linear.weight, b[l] is [out_dims[l]].  in of layer 0 is d_nav and
 * out_dims[n_layers-1] must equal n_categories.  Host pointers, borrowed for the call. */
int lmi_set_mlp(lmi_index* idx, int n_layers, const int32_t* out_dims,
                const float* const* W, const float* const* b);

/* ---- bucket store: replaces groupby('category') + data_search.loc[...] ------------------ */
/* pred_categories: the per-object bucket = output of LearnedIndex.build (LearnedIndex.py:240),
 * int64 [n_total] (host or device, says `on_device`).  Decides the CSR layout (and the shard). */
int lmi_store_begin(lmi_index* idx, const int64_t* pred_categories, int64_t n_total,
                    int64_t id_base, int on_device);
/* append rows [first_row, first_row+rows) of data_search (fp16 bits, row-major [rows, d]);
 * may be called in any order / chunking until every row was seen once. */
int lmi_store_append(lmi_index* idx, const uint16_t* rows_f16, int64_t first_row, int64_t rows,
                     int on_device);
/* same for a strided source: element (r, c) of the chunk lies at base[r * row_stride + c * col_stride] (strides in
 * ELEMENTS).  Row-major rows (col_stride = 1, row_stride >= d) or feature-major columns (row_stride = 1,
 * col_stride >= rows: how pandas lays out the block of a DataFrame built as search.py:83-84 builds data_search, so
 * it is uploaded as it lies instead of being transposed on the host).  Host sources go through two pinned staging
 * buffers, overlapping the host-side gather with the H2D copy. */
int lmi_store_append_strided(lmi_index* idx, const uint16_t* base, int64_t first_row, int64_t rows,
                             int64_t row_stride, int64_t col_stride, int on_device);
int lmi_store_end(lmi_index* idx);
/* on-disk cache of the finished device layout (CSR fp16 rows + inverse norms + ids) of this index / shard.
 * key32: 32 bytes chosen by the caller (a hash of pred_categories and a fingerprint of the data) stored in the
 * file and compared on load; lmi_store_load fails with LMI_ERR_STATE on any mismatch (shape, shard, key). */
int lmi_store_save(lmi_index* idx, const char* path, const uint8_t* key32);
int lmi_store_load(lmi_index* idx, const char* path, const uint8_t* key32);
/* convenience: begin + append(all) + end from host memory */
int lmi_set_buckets(lmi_index* idx, const uint16_t* data_f16, const int64_t* pred_categories,
                    int64_t n_total, int64_t id_base);

/* ---- navigation ------------------------------------------------------------------------ */
/* predict_proba (model.py:214-229): the n_top most probable categories per query, descending.
 * out_classes int64 [nq, n_top]; out_probs fp32 [nq, n_top] or NULL. */
int lmi_navigate(lmi_index* idx, const float* q_nav, int64_t nq, int n_top,
                 int64_t* out_classes, float* out_probs);
/* predict (model.py:201-212): argmax category of every row; out int64 [n]. */
int lmi_predict(lmi_index* idx, const float* x_nav, int64_t n, int64_t* out_category);

/* ---- search ---------------------------------------------------------------------------- */
/* LearnedIndex.search (LearnedIndex.py:22-101).  q_nav fp32 [nq, d_nav]; q_f16 fp16 bits
 * [nq, d]; out_dists float64 [nq, k] ascending; out_ids uint32 [nq, k].  Host buffers;
 * host<->device copies happen inside the call; synchronous on return. */
int lmi_search(lmi_index* idx, const float* q_nav, const uint16_t* q_f16, int64_t nq,
               int n_buckets, int k, double* out_dists, uint32_t* out_ids);
/* same with device-resident inputs/outputs, asynchronous on `stream` (a cudaStream_t). */
int lmi_search_device(lmi_index* idx, const float* q_nav_dev, const uint16_t* q_f16_dev,
                      int64_t nq, int n_buckets, int k, double* out_dists_dev,
                      uint32_t* out_ids_dev, void* stream);
/* LearnedIndex.search_single (LearnedIndex.py:103-195): one rank, caller-supplied category per
 * query (int64 [nq]); threshold (float64 [nq]) or NULL keeps only distances strictly below it
 * (utils.py:23). */
int lmi_search_rank(lmi_index* idx, const uint16_t* q_f16, int64_t nq,
                    const int64_t* category_per_query, int k, const double* threshold,
                    double* out_dists, uint32_t* out_ids);

/* ---- multi-GPU exchange (one NCCL all-gather between the two calls) -- */
/* per-rank partial result: `lmi_partial_bytes(nq,k)` bytes per rank, device memory. */
int64_t lmi_partial_bytes(int64_t nq, int k);
int lmi_search_partial_device(lmi_index* idx, const float* q_nav_dev, const uint16_t* q_f16_dev,
                              int64_t nq, int n_buckets, int k, void* partial_dev, void* stream);
/* The same with the bucket lists supplied by the caller (int32 [nq, n_buckets], device; ids outside [0, C) match no
 * bucket): lets every rank navigate only its slice of the queries (lmi_navigate_device) and all-gather the lists
 * (nq * n_buckets * 4 bytes) instead of navigating all queries on every rank. */
int lmi_navigate_device(lmi_index* idx, const float* q_nav_dev, int64_t nq, int n_top,
                        int32_t* out_classes_dev, void* stream);
int lmi_search_partial_classes_device(lmi_index* idx, const int32_t* classes_dev, const uint16_t* q_f16_dev,
                                      int64_t nq, int n_buckets, int k, void* partial_dev, void* stream);
/* gathered_dev: `world` partials back to back (rank-major). */
int lmi_merge_partials_device(lmi_index* idx, const void* gathered_dev, int world, int64_t nq,
                              int k, double* out_dists_dev, uint32_t* out_ids_dev, void* stream);

/* ---- allocation-free navigation --------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  predict_proba on device buffers:
 * q_nav_dev fp32 [nq, d_nav]; out_classes_dev int64 [nq, n_top] (descending probability, what lmi_navigate returns);
 * out_probs_dev fp32 [nq, n_top] or NULL; asynchronous on `stream`.  The call allocates nothing: the kernel's int32
 * bucket lists go to a 4 MiB device array that is part of the library image (held by every GPU that loads the library
 * until the process exits).  Limit: nq * n_top <= 1,048,576, LMI_ERR_ARG beyond it.  All calls on a device share that
 * array: calls ordered on one stream are safe, calls that overlap on the GPU race.  The first call on a device resolves
 * the array's address (cudaGetSymbolAddress): make it before a stream capture begins. */
int lmi_navigate_device_noalloc(lmi_index* idx, const float* q_nav_dev, int64_t nq, int n_top,
                                int64_t* out_classes_dev, float* out_probs_dev, void* stream);

/* ---- introspection --------------------------------------------------------------------- */
int lmi_get_stats(lmi_index* idx, lmi_stats* out);
/* CSR layout of this shard: bucket_ptr int64 [C+1] (host out). */
int lmi_get_bucket_ptr(lmi_index* idx, int64_t* out_bucket_ptr);
/* name and per-launch device time of the dominant (rescoring) kernel of the last search */
const char* lmi_rescore_kernel_name(lmi_index* idx);

#ifdef __cplusplus
}
#endif
#endif /* LMI_B200_H */
