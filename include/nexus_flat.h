/*
 * nexus_flat.h -- C ABI of the B200-native exact flat-search engine (libnexus_flat.so).
 *
 * This is the drop-in boundary for the ONE hot path of hanhainebula/Nexus that this repository
 * replaces: "score query embeddings against a corpus embedding matrix and return top-k", plus the
 * pool + L2-normalise epilogue that produces those embeddings.  The reference has no FFI of its own
 * (it is pure Python and duck-types a Faiss index object), so every entry point below cites the
 * reference call site whose work it takes over (paths are relative to the reference repository):
 *
 *   nxf_create / nxf_add      faiss.index_factory(D,"Flat",METRIC_INNER_PRODUCT) + .train/.add
 *                             Nexus/evaluation/text_retrieval/utils.py:176,188-189
 *                             faiss.IndexFlatIP / IndexFlatL2 + index_cpu_to_gpu + .add
 *                             Nexus/training/embedder/recommendation/modeling.py:182-195
 *                             NumpyIPIndex.__init__            utils.py:16-18
 *   nxf_search                faiss_index.search(q.astype(f32), k)      utils.py:224
 *                             NumpyIPIndex.search                       utils.py:20-31
 *                             index_flat.search(query_vec, k)           modeling.py:197
 *                             item_index.search(user_emb, topk)         Nexus/inference/embedder/recommendation/base.py:182
 *                             index.search(q[None,:], k=20)             serve/inference_service_text_retrieval.py:70
 *                             matmul + full torch.sort                  Nexus/evaluation/mmeb_v2/eval_embedding.py:359-361
 *   nxf_merge_topk            Faiss IndexShards host-side merge behind
 *                             index_cpu_to_all_gpus(co.shard=True)      utils.py:178-183
 *   nxf_pool_normalize        BaseEmbedder.pooling + F.normalize        Nexus/inference/embedder/text_retrieval/base.py:333-335,355-379
 *                             BaseDecoderEmbedder.pooling               .../decoder/base.py:309-310,330-342
 *                             MultimodalEmbedder._pool_hidden_states    .../multimodal_retrieval/generic.py:104-126
 *                             MMEBEmbeddingModel._pooling_last          Nexus/evaluation/mmeb_v2/models.py:136-151
 *   nxf_reconstruct           index.reconstruct_n (used by write_index-shaped persistence;
 *                             serve/inference_service_text_retrieval.py:35 reads such a file)
 *
 * Conventions
 *   - Plain C: pointers, sizes, ints.  No C++/torch types cross this boundary.
 *   - Every function returns NXF_OK (0) or a negative NXF_E* code; the message for the calling
 *     thread is available from nxf_last_error().  There is NO CPU fallback: without a usable
 *     sm_100 device every compute entry point fails with NXF_ENODEVICE / NXF_ECUDA.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Calls whose
 *     inputs and outputs are all device pointers are asynchronous on `stream`; a call with any
 *     host pointer returns after its host buffers are safe to reuse / filled.
 *   - Result layout follows Faiss: row-major [nq, k]; rows sorted best-first (descending inner
 *     product, ascending squared L2 distance); when fewer than k rows exist the tail is filled with
 *     id -1 and score -FLT_MAX (IP) / +FLT_MAX (L2) -- the padding searcher.py:156 already skips.
 *   - Row ids are insertion order, 0-based, int64, plus the index's id offset (nxf_set_id_offset),
 *     which is how a row shard reports global ids.
 *   - Thread-safety and streams: one handle may be used from several threads and on several streams.
 *     Calls on the SAME handle are serialised on the host (a mutex) AND on the device: the handle's
 *     scratch arena, staging buffers and row storage are shared, so a call issued on a stream other
 *     than the previous call's first makes its stream wait (cudaStreamWaitEvent) for the device work
 *     the previous call enqueued.  Searches on two streams therefore return correct results without
 *     any caller-side synchronisation, but they run one after the other on the device (every search
 *     kernel fills the GPU; the reference is single-threaded at this seam).  Distinct handles share
 *     nothing and overlap freely.
 */
#ifndef NEXUS_FLAT_H_
#define NEXUS_FLAT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define NXF_API
#else
#define NXF_API __attribute__((visibility("default")))
#endif

typedef struct nxf_index nxf_index;

/* error codes */
enum {
  NXF_OK = 0,
  NXF_EINVAL = -1,    /* bad argument (message says which) */
  NXF_ECUDA = -2,     /* a CUDA runtime / driver call failed */
  NXF_ENOMEM = -3,    /* device or host allocation failed */
  NXF_ENODEVICE = -4, /* no CUDA device, or the device is not sm_100 */
  NXF_EUNSUPPORTED = -5
};

/* metric: what "best" means */
enum { NXF_METRIC_IP = 0, NXF_METRIC_L2 = 1 };

/* element types for row data (and for the index's storage) */
enum { NXF_F32 = 0, NXF_F16 = 1, NXF_BF16 = 2 };

/* pooling rules of nxf_pool_normalize */
enum {
  NXF_POOL_CLS = 0,        /* h[:,0]                                   base.py:372-373 */
  NXF_POOL_MEAN = 1,       /* sum(h*mask)/sum(mask), fp32               base.py:374-377 */
  NXF_POOL_LAST = 2,       /* left-padded batch -> h[:,-1], else h[b, sum(mask[b])-1]  decoder/base.py:330-342 */
  NXF_POOL_LAST_ARGMAX = 3 /* h[b, position of the last 1 in mask[b]]   mmeb_v2/models.py:146-151 */
};

/* nxf_search flags */
enum {
  NXF_SEARCH_DEFAULT = 0,
  NXF_SEARCH_FORCE_EXACT = 1 << 0,   /* fp32 FFMA tile kernel (KX) regardless of shape */
  NXF_SEARCH_FORCE_STREAM = 1 << 1,  /* small-batch HBM-streaming kernel (K2); nq <= NXF_STREAM_MAX_NQ */
  NXF_SEARCH_FORCE_TENSOR = 1 << 2,  /* tcgen05 GEMM + fused top-k (K1); needs f16/bf16 storage */
  NXF_SEARCH_NO_RESCORE = 1 << 3,    /* K1 only: skip the fp32-query re-score of the over-fetched candidates */
  NXF_SEARCH_TENSOR_1CTA = 1 << 4,   /* K1 only: cta_group::1 tiles (debug / A-B comparison) */
  NXF_SEARCH_FORCE_TSTREAM = 1 << 5  /* small-batch tcgen05 streaming kernel (K2T); nq <= 128, f16/bf16 storage */
};
#define NXF_STREAM_MAX_NQ 4
#define NXF_TSTREAM_MAX_NQ 128
#define NXF_MAX_K 2048

/* which kernel family served the last search on this handle (nxf_stats.last_path) */
enum { NXF_PATH_NONE = 0, NXF_PATH_STREAM = 1, NXF_PATH_TENSOR = 2, NXF_PATH_EXACT = 3, NXF_PATH_TSTREAM = 4, NXF_PATH_SUBSET = 5 };

typedef struct nxf_stats {
  int64_t kernel_launches; /* number of this library's kernels launched on behalf of the handle */
  int64_t searches;
  int64_t rows_added;
  int32_t last_path;       /* NXF_PATH_* */
  int32_t last_k_fetch;    /* candidates per query the main kernel kept (k or the over-fetch k') */
  int32_t last_chunks;     /* corpus chunks per query tile in the last K1 launch */
  int32_t last_grid;       /* CTAs of the last main kernel */
  int64_t scratch_bytes;   /* device scratch currently held */
  int64_t storage_bytes;   /* device bytes of row storage currently held */
  int64_t main_kernel_count; /* main search kernels timed since nxf_set_profiling(idx, 1) */
  double main_kernel_ms;     /* their summed device time (CUDA events on the launching stream) */
  /* certified search (see nxf_search): queries whose candidate pass was checked, and how many of them were searched again by the
   * segmented tensor-core pass (level 2) and by the exact fp32 kernels (level 3) because their certificate failed */
  int64_t cert_checked;
  int64_t cert_retry_l2;
  int64_t cert_retry_l3;
} nxf_stats;

/* ---- library ---- */
NXF_API const char* nxf_version(void);
NXF_API const char* nxf_last_error(void);
/* number of visible CUDA devices (>= 0) or a negative error code */
NXF_API int nxf_device_count(void);
/* Host arithmetic only (no device needed): the rank m that nxf_search's sampled thresholds use for a sample of
 * `sample_tiles` out of `n_tiles` evenly spread corpus tiles and k kept candidates -- the smallest m (capped at k) with
 * exp(-(m ln(m/lambda) - m + lambda)) <= e^-30, lambda = k * sample_tiles / n_tiles, i.e. a Chernoff bound on
 * P(Poisson(lambda) >= m), which dominates the hypergeometric count of global top-k rows inside the sample.  The m-th best
 * score of the sample is then beaten by >= k rows of the whole corpus except with probability < e^-30 per query; results
 * do not depend on it (counts are verified and a short query is searched again), only speed does.  Replaces nothing in the
 * reference (Faiss keeps a heap per query); exported so that the bound can be tested without a GPU. */
NXF_API int nxf_sample_rank(int k, int64_t sample_tiles, int64_t n_tiles);

/* ---- index lifetime ---- */
/* dim >= 1; metric NXF_METRIC_*; store_dtype NXF_F32/F16/BF16; device = CUDA ordinal.
 * Allocates the handle only; device memory is taken at the first nxf_add / nxf_reserve. */
NXF_API int nxf_create(int dim, int metric, int store_dtype, int device, nxf_index** out);
NXF_API void nxf_destroy(nxf_index* idx);
/* drop all rows (keeps the allocation) */
NXF_API int nxf_reset(nxf_index* idx);
/* make room for n_total rows so later adds do not reallocate */
NXF_API int nxf_reserve(nxf_index* idx, int64_t n_total);
NXF_API int64_t nxf_ntotal(const nxf_index* idx);
NXF_API int nxf_dim(const nxf_index* idx);
NXF_API int nxf_metric(const nxf_index* idx);
NXF_API int nxf_store_dtype(const nxf_index* idx);
NXF_API int nxf_device(const nxf_index* idx);
/* ids reported by nxf_search are (row + offset): a row shard sets its first global row here */
NXF_API int nxf_set_id_offset(nxf_index* idx, int64_t offset);
NXF_API int nxf_get_stats(const nxf_index* idx, nxf_stats* out);
/* enable != 0: bracket every main search kernel (K1/K2/KX) with CUDA events on its stream; the summed
 * durations are folded into nxf_stats by nxf_get_stats (which then synchronises those events).
 * Enabling resets the two counters. */
NXF_API int nxf_set_profiling(nxf_index* idx, int enable);

/* ---- build / add ---- */
/* Append n rows of `dim` elements of type src_dtype, row stride `src_ld` elements (0 = dim).
 * rows is a host pointer (src_on_device = 0) or a device pointer on the index's device.
 * The index converts to its storage type and keeps its own copy; the caller keeps `rows`. */
NXF_API int nxf_add(nxf_index* idx, const void* rows, int64_t n, int src_dtype, int64_t src_ld,
                    int src_on_device, void* stream);

/* Copy rows [first, first+n) back as fp32 (dst_on_device selects the pointer kind). */
NXF_API int nxf_reconstruct(nxf_index* idx, int64_t first, int64_t n, float* dst, int dst_on_device,
                            void* stream);

/* ---- search ---- */
/* q: nq x dim of q_dtype, row stride q_ld elements (0 = dim), host or device (q_on_device).
 * out_scores [nq,k] float32 and out_ids [nq,k] int64, caller-allocated, both host or both device
 * (out_on_device).  1 <= k <= NXF_MAX_K.  flags: NXF_SEARCH_*.
 *
 * Exactness.  The reference arithmetic is plain fp32 (utils.py:21-22,224).  The tensor-core kernels see 16-bit operands, so unless
 * the queries already are of the index's 16-bit storage type a search is CERTIFIED: the candidate pass keeps k' > k rows per query
 * (2k for 16-bit storage: only the query is rounded, every pass score is within 2^-9 |q||c| of the fp32 score; 4k for fp32 storage,
 * which streams a bf16 copy of its rows: 2^-8 |q||c|), the candidates are re-scored with fp32 FMAs on the stored rows, and the
 * result of a query stands only if its k-th re-scored score clears the smallest kept pass score by that bound -- then no row
 * outside the candidates can belong to the top-k.  A query that fails is searched again by a segmented tensor-core pass whose
 * products reproduce the fp32 ones to ~1e-4 |q||c| (same certificate), and, failing that too, by the exact fp32 FMA kernels.
 * The certificate costs one 4-byte read-back per call: a certified search drains `stream` before it returns, device buffers or
 * not.  NXF_SEARCH_NO_RESCORE switches re-score and certificate off (raw tensor-core scores); the counters are in nxf_stats. */
NXF_API int nxf_search(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld,
                       int q_on_device, int k, float* out_scores, int64_t* out_ids, int out_on_device,
                       int flags, void* stream);

/* ---- per-query candidate lists ("local ranking") ----
 * Replaces the loop at Nexus/evaluation/mmeb_v2/eval_embedding.py:384-404 (eval_type != "global"): per query, stack the
 * embeddings of that query's own candidates, q @ cand^T (models.py:186-188), torch.sort(descending); the metrics read the
 * first 10 entries (utils/eval_utils/metrics.py:12).  The candidate embeddings are rows of the index; query i ranks the rows
 * rows[offsets[i] .. offsets[i+1]) (CSR; a row may appear in many lists and more than once in one list).
 * out_scores [nq,k] float32 and out_pos [nq,k] int64: the POSITIONS inside the query's own list (what the reference's
 * `ranked_idx` holds), best first, ties by ascending position (the order of a stable descending sort); columns beyond the
 * list length are padded with -1 and -FLT_MAX (IP) / +FLT_MAX (L2).  Scores are fp32 arithmetic on the stored rows and the
 * fp32 queries (inner product, or squared L2 distance for an L2 index).
 * q, (offsets, rows) and the outputs are host or device pointers as the three *_on_device flags say; total = offsets[nq].
 * A row id outside [0, ntotal) or a malformed offsets array is an error (NXF_EINVAL; the reference raises KeyError on an
 * unknown candidate name): the call synchronises `stream` to report it. */
NXF_API int nxf_search_subsets(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int q_on_device,
                               const int64_t* offsets, const int64_t* rows, int64_t total, int lists_on_device, int k,
                               float* out_scores, int64_t* out_pos, int out_on_device, void* stream);

/* ---- multi-shard merge (device pointers only) ----
 * scores [n_lists, nq, k] float32 and ids [n_lists, nq, k] int64 (global ids, -1 = padding), each
 * list's rows sorted best-first -> out_scores/out_ids [nq, k_out] sorted best-first, k_out <= NXF_MAX_K.
 * `metric` decides the direction.  Used after the NCCL all-gather of per-shard lists, and exposed
 * so that the multi-GPU logic is testable on one GPU. */
NXF_API int nxf_merge_topk(const float* scores, const int64_t* ids, int n_lists, int64_t nq, int k,
                           int k_out, int metric, float* out_scores, int64_t* out_ids, int device,
                           void* stream);

/* ---- row-sharded search: packed per-shard lists (device pointers only) ----
 * What one shard contributes to the multi-GPU search that replaces faiss.index_cpu_to_all_gpus(index, co) with
 * co.shard = True (Nexus/evaluation/text_retrieval/utils.py:178-183; Faiss merges IndexShards results on the host):
 * nxf_search_packed is nxf_search with device queries whose result rows are written as [nq, k] 8-byte entries
 * (float32 score, int32 SHARD-LOCAL row; row -1 = padding, score as in nxf_search) -- one buffer, so the per-shard lists
 * travel in ONE all-gather of 8 bytes per entry (instead of float32 scores + int64 ids = 12 bytes in two collectives).
 * nxf_merge_packed merges n_lists such lists ([n_lists, nq, k], rows best-first) into [nq, k_out] float32 scores and
 * int64 GLOBAL ids: id = row + list_offsets[list] (list_offsets: HOST array of n_lists int64, the first global row of
 * every shard).  n_lists <= 64.  Ties between lists are ordered by ascending global id, like nxf_merge_topk. */
typedef struct nxf_entry {
  float score;
  int32_t row;
} nxf_entry;
NXF_API int nxf_search_packed(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int k, nxf_entry* out,
                              int flags, void* stream);
NXF_API int nxf_merge_packed(const nxf_entry* lists, const int64_t* list_offsets, int n_lists, int64_t nq, int k, int k_out,
                             int metric, float* out_scores, int64_t* out_ids, int device, void* stream);

/* ---- rank of the positive among the top-k (device pointers only) ----
 * Replaces torch.cat([pos_scores, topk_scores]) + stable torch.sort(descending) + `indice[:, :k] == 0` in
 * BaseRetriever.eval_step, Nexus/training/embedder/recommendation/modeling.py:199-205: for every row the positive's
 * position in the descending stable order of [pos, top_1..top_k] is the number of top-k scores strictly greater than it
 * (equal scores sort after the positive, which comes first in the concatenation).  pos_scores [nq] float32,
 * topk_scores [nq, k] float32 (as returned by nxf_search; padding -FLT_MAX counts as a score) ->
 * rank [nq] int32 in [0, k]  and, if pred is not NULL, pred [nq, k] uint8 with pred[b, j] = (j == rank[b]). */
NXF_API int nxf_positive_rank(const float* pos_scores, const float* topk_scores, int64_t nq, int k, int32_t* rank,
                              uint8_t* pred, int device, void* stream);

/* ---- embedder epilogue (device pointers only) ----
 * hidden [B, L, H] of h_dtype (contiguous), mask [B, L] int64 (may be NULL for NXF_POOL_CLS),
 * out [B, H] of out_dtype.  normalize != 0 applies x / max(||x||_2, eps) (F.normalize);
 * eps <= 0 selects the reference's host-numpy form (no clamp: a zero row becomes NaN). */
NXF_API int nxf_pool_normalize(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L,
                               int64_t H, int pooling, int normalize, float eps, void* out, int out_dtype,
                               int device, void* stream);

/* ---- mean pooling without any allocation (device pointers only) ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * nxf_pool_normalize(..., NXF_POOL_MEAN, ...), but the kernel's workspace (per-row control words + fp32 partial sums) is a static
 * __device__ array inside the library instead of a cached cudaMalloc'd buffer: the call allocates and frees NOTHING (no cudaMalloc,
 * cudaFree, stream-ordered or library allocation).  Every GPU that loads the library holds that array (64 KB + 8 MB) until exit.
 * Limits, rejected with NXF_EINVAL / NXF_EUNSUPPORTED and a message: B <= 4096 rows; B * H * 4 bytes <= 8 MB; hidden states
 * 16-byte aligned with H * sizeof(element) a multiple of 16 and at most 512 such vectors per row; mask must not be NULL.
 * All calls share the one array per device and it keeps the last call's contents (the control words are cleared on `stream`
 * before every launch): calls that may overlap on the GPU -- two streams, two threads -- race; order them yourself. */
NXF_API int nxf_pool_mean_normalize_static(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L, int64_t H,
                                           int normalize, float eps, void* out, int out_dtype, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NEXUS_FLAT_H_ */
