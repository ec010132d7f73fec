// params.cuh -- kernel parameter blocks and launcher prototypes shared between api.cu and the
// kernel translation units.
#pragma once

#include "sink.cuh"

namespace nxf {

struct KxParams {
  const float* q;  // [nq, ldq] fp32, device
  int64_t ldq;
  const void* corpus;  // [n, ld] storage type
  int64_t ld;
  const float* half_norm;  // 0.5*|c|^2 per row (L2 metric) or nullptr
  int64_t n;
  int64_t nq;
  int dim;
  int num_qt;
  int chunks;
  int64_t tiles_per_chunk;
  int64_t n_tiles;
  SinkParams sink;
};

// Virtual -> physical order of the corpus tiles (DESIGN.md "sampled thresholds"): the first ns virtual tiles are the physical
// tiles 0, s, 2s, ... -- of those, every s0-th comes first (n0 tiles) -- then all the others in order.  s <= 1: identity.
struct TilePerm {
  int64_t s = 0, ns = 0, s0 = 0, n0 = 0;
  // floor(2^64 / d) + 1 for d = s - 1 and d = s0 - 1 (0 when d <= 1): j / d == hi64(j * m) exactly for every 32-bit j, which keeps
  // the per-tile index arithmetic of the single producer / issuer threads off the integer-division slow path
  uint64_t m1 = 0, m0 = 0;
  void finish() {
    m1 = s - 1 > 1 ? ~uint64_t(0) / static_cast<uint64_t>(s - 1) + 1 : 0;
    m0 = s0 - 1 > 1 ? ~uint64_t(0) / static_cast<uint64_t>(s0 - 1) + 1 : 0;
  }
};
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t perm_div(uint32_t j, uint64_t m) {  // j / d with m = floor(2^64 / d) + 1; m == 0: d == 1
  if (m == 0) return j;
  const uint64_t t = static_cast<uint64_t>(j) * static_cast<uint32_t>(m >> 32) + __umulhi(j, static_cast<uint32_t>(m));
  return static_cast<uint32_t>(t >> 32);
}
__device__ __forceinline__ int64_t perm_tile(const TilePerm& tp, int64_t v64) {
  if (tp.s <= 1) return v64;
  // tile indices fit 32 bits (< 2^31 rows per device)
  const uint32_t v = static_cast<uint32_t>(v64), ns = static_cast<uint32_t>(tp.ns);
  if (v >= ns) {  // the j-th tile that is not a multiple of s
    const uint32_t j = v - ns;
    return static_cast<int64_t>(j + perm_div(j, tp.m1) + 1u);
  }
  uint32_t i = v;  // index among the multiples of s
  if (tp.s0 > 1) {
    const uint32_t s0 = static_cast<uint32_t>(tp.s0), n0 = static_cast<uint32_t>(tp.n0);
    if (v < n0) i = v * s0;
    else { const uint32_t j = v - n0; i = j + perm_div(j, tp.m0) + 1u; }
  }
  return static_cast<int64_t>(i) * tp.s;
}
#endif

struct K1Params {
  int64_t n;    // corpus rows
  int64_t nq;   // query rows
  int kblocks;  // K blocks of 64 elements accumulated per tile: ld / 64, or (segmented) segments * split_kb
  int split_kb; // 0, or the plane width in K blocks of a segmented search (the retry level of the certified search):
                // the queries are two bf16 planes [hi | mid] per row (hi = bf16(x), mid = bf16(x - hi)) and a tile accumulates
  int seg_mode; //   1 (fp32 storage: corpus planes hi = tmap_c, mid = tmap_c2): (q_hi, c_hi), (q_hi, c_mid), (q_mid, c_hi)
                //   2 (16-bit storage, one corpus plane):                       (q_hi, c), (q_mid, c)
  int num_qt;   // query tiles of 128*CG rows (<= number of clusters)
  int epi_sets; // 1 or 2 epilogue warp sets (set s filters accumulator buffer s); 2 when a tile holds too few K blocks to hide one set
  int64_t tile_begin, tile_end;  // (virtual) corpus tiles [begin, end) of this pass (a search is 1-3 passes: sample, sample, rest)
  TilePerm perm;
  const uint32_t* run_flag;  // not NULL: the launch does nothing unless *run_flag != 0 (the fall-back pass after a failed verification)
  // end of every pass: each row's candidates (its sink.k best if it holds more) are appended to the query's arena slice
  int round_tiles;     // corpus tiles per round (the L2-resident window)
  int sync_window;     // soft round barrier: a producer may run this many rounds ahead of the rest (0 = no barrier)
  uint32_t* sync_ctr;  // [1] zeroed per launch: rounds completed, summed over clusters
  const float* half_norm;
  uint32_t* prof;      // debug bit5: 4 x uint64 {prunes, prune cycles, end-merge cycles, warps} summed over epilogue warps
  int est_only;  // estimate-only pass: a slice that beats the row's threshold records its maximum instead of its scores
  int split;  // 1: two sets, <= 2 K blocks per tile: accumulators computed and handed over as 128-column halves (search_tc.cu SPLIT)
  int debug;  // timing experiments (only in builds with -DNXF_K1_DEBUG_BUILD=1): bit0 epilogue skips the selection; bit1 also skips the TMEM loads; bit2 producer issues no TMA;
              // bit3 issuer issues no MMA; bit4 epilogue computes the maxima but never appends
  SinkParams sink;
};

struct K2Params {
  const float* q;  // [nq, ldq] fp32 device
  int64_t ldq;
  const void* corpus;
  int64_t ld;  // elements per stored row (row bytes are a multiple of 128)
  const float* half_norm;
  int64_t n;
  int nq;
  int dim;
  int stages;
  int64_t n_blocks;  // ceil(n / rows-per-stage)
  SinkParams sink;
};

struct K2TParams {
  int64_t n;     // corpus rows
  int nq;        // queries (<= 128)
  int nq_pad;    // nq rounded up to 16 = UMMA N
  int kblocks;   // 64-element K blocks per row (queries and corpus alike): ld / 64
  int stages;    // ring depth
  int tbufs;     // TMEM accumulator buffers
  int tbuf_stride;  // TMEM columns between buffers
  int check_every;  // tiles per epilogue set between two buffer-overflow checks (cap = 2k' + epi_sets*128*check_every)
  int epi_sets;     // 2..4 epilogue warp sets (four warps each, one per SM sub-partition); set s filters tiles s, s + epi_sets, ...
  // a search is 1-3 launches over consecutive tile ranges (DESIGN.md "K2T, threshold seeding"); the per-(CTA, query) lists and
  // their counts stay in global memory between launches
  int64_t tile_begin, tile_end;  // this launch streams (virtual) corpus tiles [tile_begin, tile_end), CTA b takes tile_begin + b + i * grid
  TilePerm perm;
  const uint32_t* run_flag;  // not NULL: the launch does nothing unless *run_flag != 0 (fall-back after a failed verification)
  uint32_t* cnt_g;  // [grid][nq_pad] list lengths carried from launch to launch
  int first_pass;   // lists start empty
  int last_pass;    // lists are reduced to k' and flushed to the arena at the end (else: lengths stored to cnt_g)
  int no_check;     // one tile per CTA (<= 128 entries per list, below cap): skip the overflow checks
  int est_only;     // estimate pass: nothing is appended, the CTA's best score per query goes to cnt_g (as a monotone key) instead
  // threshold fold (bf16, ip): score - tau comes out of the tensor core, see k2t_stream_kernel
  int fold;
  const float* q_f32;         // [nq][ldq] the queries in fp32 (for |q|, the lower bound that stands in for a threshold of -inf)
  int64_t ldq;
  int dim;
  const uint32_t* max_norm2;  // device word: float bits of max |row|^2
  const float* half_norm;
  SinkParams sink;
};

struct MergeArenaParams {
  const Entry* arena;
  const uint32_t* qcnt;
  int64_t arena_stride;
  int k_fetch;  // candidates to keep before the re-score (<= NXF_MAX_K)
  int k_out;    // columns of the output
  int metric;   // NXF_METRIC_*
  int rescore;  // recompute the kept candidates' scores in fp32 from q and the stored rows
  // certificate of the candidate pass (DESIGN.md "certified search"): every row NOT among the kept k' has a pass score <= tau' (the
  // smallest kept pass score), hence a true score <= tau' + eps with eps = eps_rel * |q| * max|c| (+ eps_abs * max|c|); the re-scored
  // top-k is exact iff its k-th true score >= tau' + eps.  Queries that fail are flagged for the retry level.
  int cert;                   // evaluate it (needs rescore)
  int q_round;                // the pass saw the queries rounded to this type (NXF_BF16 / NXF_F16; -1: exactly, e.g. as [hi | mid] planes):
                              // that part of the bound is the ACTUAL |q - round(q)| of the query, not the worst case 2^-9 |q|
  float eps_rel, eps_abs;     // everything else, relative to |q| (row rounding of fp32 storage, accumulation slack) / absolute
  const uint32_t* max_norm2;  // device word: float bits of max over stored rows of |row|^2
  int64_t n_rows;             // rows in the index (a query that kept them all passes)
  uint32_t* cert_fail;        // [nq] 1 = failed
  uint32_t* cert_count;       // [1] number of failed queries (atomicAdd)
  const float* q;  // [nq, ldq] fp32 (needed when rescore or metric == L2)
  int64_t ldq;
  const void* corpus;
  int64_t ld;
  int dim;
  int64_t id_offset;
  float* out_scores;  // [nq, k_out]
  int64_t* out_ids;   // [nq, k_out]
  Entry* out_packed;  // not NULL: write [nq, k_out] (score, shard-local row; -1 = padding) entries instead (nxf_search_packed)
};

struct SubsetParams {
  const float* q;          // [nq, ldq] fp32, device
  int64_t ldq;
  const void* corpus;      // [n_rows, ld] of the storage type
  int64_t ld;
  int dim;
  int64_t n_rows;
  const int64_t* offsets;  // [nq + 1] CSR offsets into rows
  const int64_t* rows;     // [total] index rows of every query's candidates
  int64_t total;
  Entry* scratch;          // [total] (score, position) of lists longer than NXF_MAX_K
  int k_out;
  int metric;
  float* out_scores;       // [nq, k_out]
  int64_t* out_pos;        // [nq, k_out] positions inside the query's own list, -1 = padding
  uint32_t* err;           // bit 0: a row id outside [0, n_rows); bit 1: malformed offsets
  int short_max;           // lists of up to this many entries are served by the warp-per-list kernel (-1: none)
};

constexpr int kMaxPackedLists = 64;
struct MergeListsParams {
  const float* scores;  // [n_lists, nq, k]
  const int64_t* ids;   // [n_lists, nq, k]
  const Entry* packed;  // not NULL: [n_lists, nq, k] (score, list-local row) entries instead of scores / ids; global id = row + offs[list]
  int64_t offs[kMaxPackedLists];
  int n_lists;
  int64_t nq;
  int k;
  int k_out;
  int metric;
  float* out_scores;
  int64_t* out_ids;
};

int launch_kx(int store_dtype, const KxParams& p, int grid, cudaStream_t stream);
// q16: [nq_rows, q_ld] 16-bit queries; corpus: [n, ld] 16-bit rows; corpus2 (seg_mode 1 only): the mid plane, same shape
int launch_k1(int store_dtype, int cta_group, const void* q16, int64_t nq_rows, int64_t q_ld, const void* corpus, const void* corpus2,
              int64_t ld, const K1Params& p, int grid, cudaStream_t stream);
int k1_smem_bytes(int cta_group);
int launch_k2(int store_dtype, const K2Params& p, int nq_t, int rw, int grid, size_t smem, cudaStream_t stream);
size_t k2_smem_bytes(int64_t ld, int elem_size, int nq_t, int rw, int stages);
int launch_k2t(int store_dtype, const void* q16, int64_t nq_rows, int64_t q_ld, const void* corpus, int64_t ld, const K2TParams& p, int grid,
               cudaStream_t stream);
size_t k2t_smem_bytes(int kblocks, int nq_pad, int stages, int fold);
int k2t_plan_stages(int kblocks, int nq_pad, int fold);
int launch_seed_tau_lists(const Entry* cand, const uint32_t* cnt_g, int grid, int nq_pad, int cap, int k_fetch, uint32_t* tau_g, int nq,
                          cudaStream_t stream);
int launch_seed_tau_est(const uint32_t* est_g, int grid, int nq_pad, int m, uint32_t* tau_g, int nq, cudaStream_t stream);
int launch_select_tau(Entry* arena, uint32_t* qcnt, int64_t arena_stride, int m, uint32_t* tau_g, int64_t nq, int reset, cudaStream_t stream);
int launch_verify_counts(uint32_t* qcnt, uint32_t* tau_g, int64_t nq, int need, uint32_t* flag, cudaStream_t stream);
int launch_merge_arena(int store_dtype, const MergeArenaParams& p, int64_t nq, cudaStream_t stream);
int launch_merge_lists(const MergeListsParams& p, cudaStream_t stream);
int launch_subset_search(int store_dtype, const SubsetParams& p, int64_t nq, int64_t min_len, int64_t max_len, cudaStream_t stream);
int launch_positive_rank(const float* pos, const float* topk, int64_t nq, int k, int32_t* rank, uint8_t* pred, cudaStream_t stream);
int launch_convert_rows(const void* src, int src_dtype, int64_t src_ld, void* dst, int dst_dtype, int64_t dst_ld, int64_t n, int dim,
                        cudaStream_t stream);
int launch_split_rows(const float* src, int64_t ld, void* dst, int64_t n, int dst_dtype, cudaStream_t stream);
int launch_mid_rows(const float* src, int64_t ld, void* dst, int64_t n, cudaStream_t stream);
int launch_row_norm_max(const void* rows, int dtype, int64_t ld, int64_t n, int dim, uint32_t* max_norm2, cudaStream_t stream);
int launch_gather_rows(const void* src, int64_t row_bytes, const int64_t* rows, int64_t n, void* dst, cudaStream_t stream);
int launch_scatter_results(const float* src_scores, const int64_t* src_ids, const Entry* src_packed, const int64_t* rows, int64_t n, int k,
                           float* dst_scores, int64_t* dst_ids, Entry* dst_packed, cudaStream_t stream);
int launch_half_norm(const void* rows, int dtype, int64_t ld, int64_t n, int dim, float* out, cudaStream_t stream);
int launch_pool(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize,
                float eps, void* out, int out_dtype, cudaStream_t stream);

}  // namespace nxf
