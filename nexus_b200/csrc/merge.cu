// merge.cu -- final per-query selection + ordering kernels.
//
//   merge_arena_kernel : one CTA per query.  Input: the query's arena slice (unordered candidates
//       flushed by K1/K2/KX).  Block radix-select to k_fetch, optional exact fp32 re-score of the
//       survivors against the stored rows (fixes the query-side rounding of the tensor-core pass and
//       reports true fp32 scores), bitonic sort by (score desc, row asc), write Faiss-shaped rows
//       [k_out] of (float32 score, int64 id) with -1 / -FLT_MAX padding.
//   merge_lists_kernel : one CTA per query.  Input: n_lists sorted per-shard lists (the result of the
//       NCCL all-gather in the row-sharded multi-GPU search, replacing Faiss IndexShards' host merge
//       behind utils.py:178-183).  Same select + sort on (score, global id).
#include <cstdlib>

#include "params.cuh"
#include "block_sort.cuh"

namespace nxf {

// ---- arena merge (+ optional rescore, certificate) -------------------------------------------
// Round 2 (the kernel is what a row shard pays on top of its GEMM and does not shrink with the shard):
//   * slices of <= 2 048 entries are staged in shared memory once; the selection passes run there instead of over global memory;
//   * with a certificate, only candidates that can still reach the top-k are re-scored: k candidates have a pass score >= a_k (the
//     k-th largest pass score), so the true k-th score is >= a_k - eps, and a candidate whose pass score is below a_k - 2 eps cannot
//     beat it -- it is dropped before the gather (C2: ~150 of the k' = 200 rows per query are read);
//   * four candidate rows in flight per warp instead of two;
//   * up to 512 survivors are ordered by counting (rank = number of larger keys; one barrier) instead of a bitonic network.
namespace mg { constexpr int STAGE = 2048; constexpr int RANK_SORT_MAX = 512; }

template <typename T>
__global__ void __launch_bounds__(mg::THREADS) merge_arena_kernel(MergeArenaParams p) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  uint64_t* keys = reinterpret_cast<uint64_t*>(dyn_smem);                       // [MAXK] sort keys; first the selected entries
  Entry* stage = reinterpret_cast<Entry*>(dyn_smem + mg::MAXK * 8);             // [STAGE] the staged slice, later the live candidates
  float* q_s = reinterpret_cast<float*>(dyn_smem + mg::MAXK * 8 + mg::STAGE * 8);  // [ld] when rescore / L2
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh_prefix;
  __shared__ int sh_need, sh_gt, sh_eq, sh_live;
  __shared__ float sh_qnorm, sh_qlo, sh_tau_min, sh_sk;
  __shared__ float red[mg::THREADS / 32], red_lo[mg::THREADS / 32];

  const int64_t qi = blockIdx.x;
  const int tid = threadIdx.x;
  const Entry* __restrict__ gsrc = p.arena + qi * p.arena_stride;
  int64_t m64 = p.qcnt[qi];
  NXF_DEV_CHECK(m64 <= p.arena_stride, "arena slice count beyond its stride");
  if (m64 > p.arena_stride) m64 = p.arena_stride;
  const int M = static_cast<int>(m64);
  const int keep = min(M, p.k_fetch);
  NXF_DEV_CHECK(keep <= mg::MAXK, "k' beyond the merge's shared-memory arrays");
  Entry* ent = reinterpret_cast<Entry*>(keys);  // gather target aliases the key array (8 B each)

  const bool need_q = p.rescore || p.metric == NXF_METRIC_L2;
  if (need_q) {
    float part = 0.f, part_lo = 0.f;
    for (int d = tid; d < static_cast<int>(p.ld); d += mg::THREADS) {  // ldq >= ld and both are zero padded past dim
      const float v = d < p.dim ? p.q[qi * p.ldq + d] : 0.f;
      q_s[d] = v;
      part += v * v;
      if (p.cert && p.q_round >= 0) {   // what the candidate pass lost of this query: q - round(q)
        const float r = p.q_round == NXF_F16 ? __half2float(__float2half_rn(v)) : __bfloat162float(__float2bfloat16_rn(v));
        part_lo = fmaf(v - r, v - r, part_lo);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { part += __shfl_xor_sync(kFullMask, part, o); part_lo += __shfl_xor_sync(kFullMask, part_lo, o); }
    if ((tid & 31) == 0) { red[tid >> 5] = part; red_lo[tid >> 5] = part_lo; }
    __syncthreads();
    if (tid == 0) {
      float s = 0.f, sl = 0.f;
      for (int w = 0; w < mg::THREADS / 32; ++w) { s += red[w]; sl += red_lo[w]; }
      sh_qnorm = s;
      sh_qlo = sl;
    }
  }
  const bool staged = M <= mg::STAGE;
  if (staged) for (int i = tid; i < M; i += mg::THREADS) stage[i] = gsrc[i];
  if (tid == 0) { sh_gt = 0; sh_eq = 0; sh_live = 0; sh_sk = 0.f; }
  __syncthreads();

  // ---- the k' best by pass score -> ent[0, keep) ----
  if (M > keep) {
    if (staged) block_radix_select(M, keep, [&](int i) { return okey(stage[i].score); }, hist, &sh_prefix, &sh_need);
    else block_radix_select(M, keep, [&](int i) { return okey(gsrc[i].score); }, hist, &sh_prefix, &sh_need);
    const uint32_t kth = sh_prefix;
    const int need = sh_need;
    const int n_gt = keep - need;
    for (int i = tid; i < M; i += mg::THREADS) {
      const Entry e = staged ? stage[i] : gsrc[i];
      const uint32_t key = okey(e.score);
      if (key > kth) {
        ent[atomicAdd(&sh_gt, 1)] = e;
      } else if (key == kth) {
        const int t = atomicAdd(&sh_eq, 1);
        if (t < need) ent[n_gt + t] = e;
      }
    }
  } else {
    for (int i = tid; i < M; i += mg::THREADS) ent[i] = staged ? stage[i] : gsrc[i];
  }
  __syncthreads();

  // ---- certificate, first half: tau' = the smallest PASS score among the kept candidates bounds the pass score of every other row ----
  float eps = 0.f;
  if (p.cert) {
    float mn = INFINITY;
    for (int i = tid; i < keep; i += mg::THREADS) mn = fminf(mn, ent[i].score);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(kFullMask, mn, o));
    if ((tid & 31) == 0) red[tid >> 5] = mn;
    __syncthreads();
    if (tid == 0) {
      float m = red[0];
      for (int w = 1; w < mg::THREADS / 32; ++w) m = fminf(m, red[w]);
      sh_tau_min = m;
    }
    eps = (sqrtf(sh_qlo) * 1.001f + p.eps_rel * sqrtf(sh_qnorm) + p.eps_abs) * sqrtf(__uint_as_float(*p.max_norm2)) * 1.0001f;
  }

  // ---- the candidates worth re-scoring -> live[0, n_live) ----
  Entry* live = ent;
  int n_live = keep;
  if (p.rescore && p.cert && keep > p.k_out) {
    block_radix_select(keep, p.k_out, [&](int i) { return okey(ent[i].score); }, hist, &sh_prefix, &sh_need);
    const float floor_s = okey_inv(sh_prefix) - 2.f * eps;     // a_k - 2 eps
    for (int i = tid; i < keep; i += mg::THREADS) {
      const Entry e = ent[i];
      if (e.score >= floor_s) {
        const int at = atomicAdd(&sh_live, 1);
        NXF_DEV_CHECK(at < mg::STAGE, "live candidate list overflow");
        stage[at] = e;
      }
    }
    __syncthreads();
    live = stage;
    n_live = sh_live;
  }

  // ---- exact fp32 re-score: one warp per candidate, four candidates in flight per warp, 16-byte loads ----
  // (rows are ld = 64k elements long and zero padded, q_s holds ld entries: no tail handling)
  if (p.rescore) {
    const T* __restrict__ corpus = reinterpret_cast<const T*>(p.corpus);
    const int warp = tid >> 5, lane = tid & 31;
    const int ld = static_cast<int>(p.ld);
    constexpr int NW = mg::THREADS / 32;
    constexpr int RIF = 4;  // rows in flight per warp
    for (int c0 = warp; c0 < n_live; c0 += RIF * NW) {
      const T* r[RIF];
      bool on[RIF];
      float dot[RIF], nn[RIF];
#pragma unroll
      for (int u = 0; u < RIF; ++u) {
        const int c = c0 + u * NW;
        on[u] = c < n_live;
        r[u] = corpus + static_cast<int64_t>(live[on[u] ? c : c0].row) * p.ld;
        dot[u] = 0.f; nn[u] = 0.f;
      }
      for (int d0 = lane * 8; d0 < ld; d0 += 256) {
        float x[RIF][8];
#pragma unroll
        for (int u = 0; u < RIF; ++u) load8_f32<T>(r[u] + d0, x[u]);
        const float4 q0 = *reinterpret_cast<const float4*>(q_s + d0), q1 = *reinterpret_cast<const float4*>(q_s + d0 + 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int u = 0; u < RIF; ++u)
#pragma unroll
          for (int e = 0; e < 8; ++e) { dot[u] = fmaf(qv[e], x[u][e], dot[u]); nn[u] = fmaf(x[u][e], x[u][e], nn[u]); }
      }
#pragma unroll
      for (int u = 0; u < RIF; ++u) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { dot[u] += __shfl_xor_sync(kFullMask, dot[u], o); nn[u] += __shfl_xor_sync(kFullMask, nn[u], o); }
        if (lane == 0 && on[u]) live[c0 + u * NW].score = (p.metric == NXF_METRIC_L2) ? (dot[u] - 0.5f * nn[u]) : dot[u];
      }
    }
  }
  __syncthreads();

  // ---- order by (score desc, row asc) and write ----
  const float qn = (p.metric == NXF_METRIC_L2) ? sh_qnorm : 0.f;
  auto emit = [&](int j, uint64_t kk) {
    const float sc = sort_key_score(kk);
    const float s = (p.metric == NXF_METRIC_L2) ? fmaxf(qn - 2.f * sc, 0.f) : sc;
    const int32_t row = sort_key_row(kk);
    if (p.out_packed != nullptr) {
      Entry e; e.score = s; e.row = row;                     // shard-local: the shard merge adds the shard's offset
      p.out_packed[qi * p.k_out + j] = e;
    } else {
      p.out_scores[qi * p.k_out + j] = s;
      p.out_ids[qi * p.k_out + j] = static_cast<int64_t>(row) + p.id_offset;
    }
  };
  const int n_out = min(n_live, p.k_out);
  if (n_live <= mg::RANK_SORT_MAX) {
    // counting sort: keys are distinct (rows are), the rank of a key is the number of larger ones
    uint64_t mine[mg::RANK_SORT_MAX / mg::THREADS];
#pragma unroll
    for (int u = 0; u < mg::RANK_SORT_MAX / mg::THREADS; ++u) {
      const int i = tid + u * mg::THREADS;
      mine[u] = 0;
      if (i < n_live) { const Entry e = live[i]; mine[u] = sort_key(e.score, e.row); }
    }
    __syncthreads();   // every thread has read its entries of live[] (which may alias keys[])
#pragma unroll
    for (int u = 0; u < mg::RANK_SORT_MAX / mg::THREADS; ++u) {
      const int i = tid + u * mg::THREADS;
      if (i < n_live) keys[i] = mine[u];
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < mg::RANK_SORT_MAX / mg::THREADS; ++u) {
      const int i = tid + u * mg::THREADS;
      if (i >= n_live) continue;
      int rank = 0;
      for (int j = 0; j < n_live; ++j) rank += (keys[j] > mine[u]) ? 1 : 0;
      if (rank < p.k_out) emit(rank, mine[u]);
      if (rank == p.k_out - 1) sh_sk = sort_key_score(mine[u]);
    }
  } else {
    const int P = next_pow2(max(n_live, 2));
    uint64_t tmp[mg::MAXK / mg::THREADS];
#pragma unroll
    for (int u = 0; u < mg::MAXK / mg::THREADS; ++u) {
      const int i = tid + u * mg::THREADS;
      tmp[u] = 0;  // below every real key
      if (i < n_live) { const Entry e = live[i]; tmp[u] = sort_key(e.score, e.row); }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < mg::MAXK / mg::THREADS; ++u) {
      const int i = tid + u * mg::THREADS;
      if (i < P) keys[i] = tmp[u];
    }
    block_bitonic_desc<false>(keys, nullptr, P);
    for (int j = tid; j < n_out; j += mg::THREADS) emit(j, keys[j]);
    if (tid == 0 && n_live >= p.k_out) sh_sk = sort_key_score(keys[p.k_out - 1]);
  }
  for (int j = n_out + tid; j < p.k_out; j += mg::THREADS) {
    if (p.out_packed != nullptr) {
      Entry e; e.score = (p.metric == NXF_METRIC_L2) ? FLT_MAX : -FLT_MAX; e.row = -1;
      p.out_packed[qi * p.k_out + j] = e;
    } else {
      p.out_scores[qi * p.k_out + j] = (p.metric == NXF_METRIC_L2) ? FLT_MAX : -FLT_MAX;
      p.out_ids[qi * p.k_out + j] = -1;
    }
  }

  // ---- certificate, second half: the re-scored k-th score must clear tau' + eps (params.cuh, MergeArenaParams::cert) ----
  if (p.cert) {
    __syncthreads();
    if (tid == 0) {
      bool pass = static_cast<int64_t>(keep) >= p.n_rows;     // every row of the index was a candidate
      if (!pass && n_live >= p.k_out) pass = sh_sk >= sh_tau_min + eps;
      p.cert_fail[qi] = pass ? 0u : 1u;
      if (!pass) atomicAdd(p.cert_count, 1u);
    }
  }
}

int launch_merge_arena(int store_dtype, const MergeArenaParams& p, int64_t nq, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  const bool need_q = p.rescore || p.metric == NXF_METRIC_L2;
  const size_t smem = static_cast<size_t>(mg::MAXK) * 8 + static_cast<size_t>(mg::STAGE) * 8 + (need_q ? static_cast<size_t>(p.ld) * 4 : 0);
  if (smem > 200 * 1024) { set_error("merge: dim %d too large for the re-score stage", p.dim); return NXF_EUNSUPPORTED; }
#define NXF_LAUNCH_MERGE(T)                                                                                          \
  do {                                                                                                               \
    if (smem > 48 * 1024) {                                                                                          \
      const int rc_ = ensure_dyn_smem(reinterpret_cast<const void*>(merge_arena_kernel<T>), (int)smem);              \
      if (rc_) return rc_;                                                                                           \
    }                                                                                                                \
    merge_arena_kernel<T><<<static_cast<unsigned>(nq), mg::THREADS, smem, stream>>>(p);                              \
  } while (0)
  switch (store_dtype) {
    case NXF_F32: NXF_LAUNCH_MERGE(float); break;
    case NXF_F16: NXF_LAUNCH_MERGE(__half); break;
    case NXF_BF16: NXF_LAUNCH_MERGE(__nv_bfloat16); break;
    default: set_error("merge: bad storage dtype %d", store_dtype); return NXF_EINVAL;
  }
#undef NXF_LAUNCH_MERGE
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}


// ---- threshold seeding -------------------------------------------------------------------------
// K2T keeps one candidate list per (CTA, query); a CTA's own k'-th best only reflects 1/grid of the corpus, so its threshold
// is weak.  Between two launches of a search this kernel looks at ALL lists of a query (one CTA per query): gathers their
// keys into shared memory, selects the k'-th largest and raises tau_g to it.  That is a valid lower bound on the final k'-th
// score (k' held entries reach it), and it reflects every row streamed so far: the next launch rejects all but
// ~k' * (rows of the launch / rows before it) scores per query.  The lists are not modified.
constexpr int SEED_MAX_KEYS = 48 * 1024;  // 192 KB of shared memory; more candidates than that: the threshold is left as it is

__global__ void __launch_bounds__(mg::THREADS) seed_tau_lists_kernel(const Entry* __restrict__ cand, const uint32_t* __restrict__ cnt_g,
                                                                     int grid, int nq_pad, int cap, int k_fetch, uint32_t* tau_g) {
  extern __shared__ uint32_t seed_smem[];
  __shared__ uint32_t hist[256];
  __shared__ int offs[256 + 1];  // grid <= 256 (one CTA per SM)
  uint32_t* keys = seed_smem;
  const int q = blockIdx.x;
  const int tid = threadIdx.x;
  if (tid < 32) {  // exclusive scan of the list lengths by one warp, 8 lists per lane
    int c[8], sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int b = tid * 8 + j;
      c[j] = b < grid ? min(static_cast<int>(cnt_g[static_cast<int64_t>(b) * nq_pad + q]), cap) : 0;
      sum += c[j];
    }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (tid >= o) incl += t;
    }
    int run = incl - sum;
#pragma unroll
    for (int j = 0; j < 8; ++j) { offs[tid * 8 + j] = run; run += c[j]; }
    if (tid == 31) offs[256] = incl;
  }
  __syncthreads();
  const int total = offs[256];
  if (total < k_fetch || total > SEED_MAX_KEYS) return;  // uniform per block
  // gather: warp w copies the keys of lists w, w + 8, ... (independent coalesced loads, no index search)
  const int warp = tid >> 5, lane = tid & 31;
  uint32_t lmin = 0xFFFFFFFFu, lmax = 0u;
  for (int b = warp; b < grid; b += mg::THREADS / 32) {
    const int o = offs[b], c = offs[b + 1] - o;
    const Entry* __restrict__ src = cand + (static_cast<int64_t>(b) * nq_pad + q) * cap;
    for (int i = lane; i < c; i += 32) {
      const uint32_t key = okey(src[i].score);
      keys[o + i] = key;
      lmin = min(lmin, key);
      lmax = max(lmax, key);
    }
  }
  lmin = __reduce_min_sync(0xFFFFFFFFu, lmin);
  lmax = __reduce_max_sync(0xFFFFFFFFu, lmax);
  uint32_t* red = hist;  // [2][8] counts, then min / max
  if (lane == 0) { red[32 + warp] = lmin; red[48 + warp] = lmax; }
  __syncthreads();
  lmin = 0xFFFFFFFFu; lmax = 0u;
#pragma unroll
  for (int w = 0; w < mg::THREADS / 32; ++w) { lmin = min(lmin, red[32 + w]); lmax = max(lmax, red[48 + w]); }
  // bitwise descent below the common prefix of min and max (select.cuh), counts reduced over the block; all threads track T
  uint32_t T = lmax;
  const uint32_t diff = lmin ^ lmax;
  if (diff != 0u) {
    const int top = 31 - __clz(diff);
    T = (top == 31) ? 0u : (lmax & ~((2u << top) - 1u));
    int par = 0;
#pragma unroll 1
    for (int b = top; b >= 0; --b, par ^= 1) {
      const uint32_t c_key = T | (1u << b);
      unsigned c = 0;
      for (int i = tid; i < total; i += mg::THREADS) c += (keys[i] >= c_key) ? 1u : 0u;
      c = __reduce_add_sync(0xFFFFFFFFu, c);
      if (lane == 0) red[par * 8 + warp] = c;
      __syncthreads();  // red[par] complete; red[par ^ 1] of the previous step is no longer read by anyone
      unsigned tot = 0;
#pragma unroll
      for (int w = 0; w < mg::THREADS / 32; ++w) tot += red[par * 8 + w];
      if (tot >= static_cast<unsigned>(k_fetch)) {
        T = c_key;
        if (tot == static_cast<unsigned>(k_fetch)) break;  // uniform: exactly k' keys >= T
      }
    }
  }
  if (tid == 0) atomicMax(&tau_g[q], T);
}

// K2T's estimate pass leaves one key per (CTA, query) in est_g: the best score among the rows that CTA streamed.  The m-th
// largest of those maxima is a lower bound of the m-th best score of the streamed sample (m different rows reach it), hence --
// for m = sample_rank() -- of the k'-th best of the whole corpus with the usual probability; verified after the main pass like
// every sampled threshold.  One warp per query, bitwise descent on the keys (<= 8 per lane).
__global__ void __launch_bounds__(256) seed_tau_est_kernel(const uint32_t* __restrict__ est_g, int grid, int nq_pad, int m, uint32_t* tau_g, int nq) {
  const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (q >= nq) return;  // whole warps
  const int lane = threadIdx.x & 31;
  uint32_t key[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int b = lane + 32 * j;
    key[j] = b < grid ? est_g[static_cast<int64_t>(b) * nq_pad + q] : 0u;
  }
  uint32_t prefix = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t cand = prefix | (1u << bit);
    uint32_t c = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) c += key[j] >= cand ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (c >= static_cast<uint32_t>(m)) prefix = cand;
  }
  // prefix = the m-th largest key (0 when fewer than m CTAs saw a row); -inf and below leave the threshold alone
  if (lane == 0 && prefix > kKeyNegInf) atomicMax(&tau_g[q], prefix);
}

int launch_seed_tau_est(const uint32_t* est_g, int grid, int nq_pad, int m, uint32_t* tau_g, int nq, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  if (grid > 256) { set_error("seed: %d CTAs per query (max 256)", grid); return NXF_EUNSUPPORTED; }
  seed_tau_est_kernel<<<static_cast<unsigned>((nq * 32 + 255) / 256), 256, 0, stream>>>(est_g, grid, nq_pad, m, tau_g, nq);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

int launch_seed_tau_lists(const Entry* cand, const uint32_t* cnt_g, int grid, int nq_pad, int cap, int k_fetch, uint32_t* tau_g, int nq,
                          cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  if (grid > 256) { set_error("seed: %d lists per query (max 256)", grid); return NXF_EUNSUPPORTED; }
  const size_t smem = static_cast<size_t>(std::min<int64_t>(SEED_MAX_KEYS, static_cast<int64_t>(grid) * cap)) * sizeof(uint32_t);
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(seed_tau_lists_kernel), SEED_MAX_KEYS * 4)) return rc;
  seed_tau_lists_kernel<<<static_cast<unsigned>(nq), mg::THREADS, smem, stream>>>(cand, cnt_g, grid, nq_pad, cap, k_fetch, tau_g);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// ---- between two K1 passes: the slice's m-th best becomes the query's threshold ------------------
// One warp per query: the slice (what the passes so far appended) is reduced in place to its m best entries, whose m-th score
// goes to tau_g.  `reset` empties the slice instead (the estimate-only pass: its tiles are streamed again).
__global__ void __launch_bounds__(256) select_tau_kernel(Entry* __restrict__ arena, uint32_t* __restrict__ qcnt, int64_t arena_stride, int m,
                                                         uint32_t* tau_g, int64_t nq, int reset) {
  const int64_t q = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  if (q >= nq) return;  // whole warps
  const unsigned lane = threadIdx.x & 31;
  int n = static_cast<int>(min(static_cast<int64_t>(qcnt[q]), arena_stride));
  __syncwarp();
  if (n > m) {
    const uint32_t kth = warp_prune(arena + q * arena_stride, n, m);
    if (lane == 0) atomicMax(&tau_g[q], kth);
    n = m;
  }
  if (lane == 0) qcnt[q] = reset ? 0u : static_cast<uint32_t>(n);
}

int launch_select_tau(Entry* arena, uint32_t* qcnt, int64_t arena_stride, int m, uint32_t* tau_g, int64_t nq, int reset, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  select_tau_kernel<<<static_cast<unsigned>((nq * 32 + 255) / 256), 256, 0, stream>>>(arena, qcnt, arena_stride, m, tau_g, nq, reset);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// ---- verification of sampled thresholds --------------------------------------------------------
// K1 starts its main pass from a threshold ESTIMATED on a sample (DESIGN.md "sampled thresholds"); the estimate is below the
// true k'-th score except with probability ~e^-30 per query.  A query for which fewer than `need` candidates came back is
// searched again from scratch by the fall-back pass: its slice is emptied and its threshold reset, every other query gets
// +inf (the fall-back pass then appends nothing for it), and *flag tells the fall-back launch whether to run at all.
__global__ void __launch_bounds__(256) verify_counts_kernel(uint32_t* __restrict__ qcnt, uint32_t* __restrict__ tau_g, int64_t nq, int need,
                                                            uint32_t* flag) {
  const int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (q >= nq) return;
  if (qcnt[q] < static_cast<uint32_t>(need)) {
    qcnt[q] = 0u;
    tau_g[q] = 0u;
    atomicOr(flag, 1u);
  } else {
    tau_g[q] = 0xFF800000u;  // okey(+inf)
  }
}

int launch_verify_counts(uint32_t* qcnt, uint32_t* tau_g, int64_t nq, int need, uint32_t* flag, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  verify_counts_kernel<<<static_cast<unsigned>((nq + 255) / 256), 256, 0, stream>>>(qcnt, tau_g, nq, need, flag);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// ---- multi-list merge (multi-GPU shard merge) ---------------------------------------------------
template <bool PACKED>
__global__ void __launch_bounds__(mg::THREADS) merge_lists_kernel(const __grid_constant__ MergeListsParams p) {
  __shared__ uint64_t keys[mg::MAXK];
  __shared__ int64_t pay[mg::MAXK];
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh_prefix;
  __shared__ int sh_need, sh_gt, sh_eq, sh_valid;

  const int64_t qi = blockIdx.x;
  const int tid = threadIdx.x;
  const int M = p.n_lists * p.k;
  const bool l2 = p.metric == NXF_METRIC_L2;
  auto score_at = [&](int i) -> float {
    const int l = i / p.k, j = i - l * p.k;
    if (PACKED) return p.packed[(static_cast<int64_t>(l) * p.nq + qi) * p.k + j].score;
    return p.scores[(static_cast<int64_t>(l) * p.nq + qi) * p.k + j];
  };
  auto id_at = [&](int i) -> int64_t {
    const int l = i / p.k, j = i - l * p.k;
    if (PACKED) {
      const int32_t row = p.packed[(static_cast<int64_t>(l) * p.nq + qi) * p.k + j].row;
      return row < 0 ? -1 : static_cast<int64_t>(row) + p.offs[l];
    }
    return p.ids[(static_cast<int64_t>(l) * p.nq + qi) * p.k + j];
  };
  // key 0 for padding (id < 0) so that it sorts last in either metric
  auto key_at = [&](int i) -> uint32_t {
    if (id_at(i) < 0) return 0u;
    const float s = score_at(i);
    return okey(l2 ? -s : s);
  };
  if (tid == 0) { sh_gt = 0; sh_eq = 0; sh_valid = 0; }
  __syncthreads();
  int local_valid = 0;
  for (int i = tid; i < M; i += mg::THREADS) local_valid += (id_at(i) >= 0);
  if (local_valid) atomicAdd(&sh_valid, local_valid);
  __syncthreads();
  const int keep = min(min(sh_valid, p.k_out), mg::MAXK);

  if (sh_valid > keep) {
    block_radix_select(M, keep, key_at, hist, &sh_prefix, &sh_need);
    const uint32_t kth = sh_prefix;
    const int need = sh_need;
    const int n_gt = keep - need;
    for (int i = tid; i < M; i += mg::THREADS) {
      const uint32_t key = key_at(i);
      if (key == 0u) continue;
      int slot = -1;
      if (key > kth) slot = atomicAdd(&sh_gt, 1);
      else if (key == kth) { const int t = atomicAdd(&sh_eq, 1); if (t < need) slot = n_gt + t; }
      if (slot >= 0) {
        const int64_t id = id_at(i);
        // tie order: smaller global id first (low 32 bits of the key hold ~id, saturated)
        const uint32_t low = static_cast<uint32_t>(0x7FFFFFFF) - static_cast<uint32_t>(id > 0x7FFFFFFE ? 0x7FFFFFFE : id);
        keys[slot] = (static_cast<uint64_t>(key) << 32) | low;
        pay[slot] = id;
      }
    }
  } else {
    for (int i = tid; i < M; i += mg::THREADS) {
      const uint32_t key = key_at(i);
      if (key == 0u) continue;
      const int slot = atomicAdd(&sh_gt, 1);
      const int64_t id = id_at(i);
      const uint32_t low = static_cast<uint32_t>(0x7FFFFFFF) - static_cast<uint32_t>(id > 0x7FFFFFFE ? 0x7FFFFFFE : id);
      keys[slot] = (static_cast<uint64_t>(key) << 32) | low;
      pay[slot] = id;
    }
  }
  __syncthreads();
  const int P = next_pow2(max(keep, 2));
  for (int i = keep + tid; i < P; i += mg::THREADS) { keys[i] = 0; pay[i] = -1; }
  block_bitonic_desc<true>(keys, pay, P);
  for (int j = tid; j < p.k_out; j += mg::THREADS) {
    float s;
    int64_t id;
    if (j < keep) {
      const float sc = okey_inv(static_cast<uint32_t>(keys[j] >> 32));
      s = l2 ? -sc : sc;
      id = pay[j];
    } else {
      s = l2 ? FLT_MAX : -FLT_MAX;
      id = -1;
    }
    p.out_scores[qi * p.k_out + j] = s;
    p.out_ids[qi * p.k_out + j] = id;
  }
}

// The same merge for lists that ARE sorted best-first (every list nxf_search / nxf_search_packed writes is: score descending, id
// ascending among equal scores), without selection or sorting: an entry's place in the merged order is its position in its own list
// plus, for every other list, the number of entries that precede it there -- a binary search.  One CTA per query, the 64-bit keys
// (monotone score key | inverted id) of all lists in shared memory, one barrier.  2 lists x 100: 0.14 ms -> ~20 us for 6 980 queries.
constexpr int kRankMergeMax = 4096;   // entries per query (n_lists * k) served by this kernel; beyond: the select + sort kernel
template <bool PACKED>
__global__ void __launch_bounds__(128) merge_sorted_lists_kernel(const __grid_constant__ MergeListsParams p) {
  extern __shared__ __align__(16) unsigned char ms_smem[];
  const int M = p.n_lists * p.k;
  uint64_t* keys = reinterpret_cast<uint64_t*>(ms_smem);            // [M]
  int64_t* pay = reinterpret_cast<int64_t*>(ms_smem + static_cast<size_t>(M) * 8);  // [M]
  __shared__ int sh_valid;
  const int64_t qi = blockIdx.x;
  const int tid = threadIdx.x;
  const bool l2 = p.metric == NXF_METRIC_L2;
  if (tid == 0) sh_valid = 0;
  __syncthreads();
  int valid = 0;
  for (int i = tid; i < M; i += 128) {
    const int l = i / p.k, j = i - l * p.k;
    const int64_t at = (static_cast<int64_t>(l) * p.nq + qi) * p.k + j;
    float sc;
    int64_t id;
    if (PACKED) {
      const Entry e = p.packed[at];
      sc = e.score;
      id = e.row < 0 ? -1 : static_cast<int64_t>(e.row) + p.offs[l];
    } else {
      sc = p.scores[at];
      id = p.ids[at];
    }
    uint64_t kk = 0;   // padding: below every real key
    if (id >= 0) {
      const uint32_t low = static_cast<uint32_t>(0x7FFFFFFF) - static_cast<uint32_t>(id > 0x7FFFFFFE ? 0x7FFFFFFE : id);
      kk = (static_cast<uint64_t>(okey(l2 ? -sc : sc)) << 32) | low;
      ++valid;
    }
    keys[i] = kk;
    pay[i] = id;
  }
  if (valid) atomicAdd(&sh_valid, valid);
  __syncthreads();
  const int total = sh_valid;
  for (int i = tid; i < M; i += 128) {
    const uint64_t kk = keys[i];
    if (kk == 0) continue;
    const int l = i / p.k, j = i - l * p.k;
    int rank = j;                                   // the entries before it in its own (sorted) list
    for (int o = 0; o < p.n_lists && rank < p.k_out; ++o) {
      if (o == l) continue;
      const uint64_t* lst = keys + o * p.k;
      // number of entries of list o that come first: key greater, or equal and the list has the smaller index
      int lo = 0, hi = p.k;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const uint64_t km = lst[mid];
        if (km > kk || (km == kk && o < l)) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < p.k_out) {
      const float sc = okey_inv(static_cast<uint32_t>(kk >> 32));
      p.out_scores[qi * p.k_out + rank] = l2 ? -sc : sc;
      p.out_ids[qi * p.k_out + rank] = pay[i];
    }
  }
  for (int j = total + tid; j < p.k_out; j += 128) {
    p.out_scores[qi * p.k_out + j] = l2 ? FLT_MAX : -FLT_MAX;
    p.out_ids[qi * p.k_out + j] = -1;
  }
}

int launch_merge_lists(const MergeListsParams& p, cudaStream_t stream) {
  if (p.nq <= 0) return NXF_OK;
  const int64_t M = static_cast<int64_t>(p.n_lists) * p.k;
  static const bool rank_merge = []() { const char* v = getenv("NXF_MERGE_SORTED"); return !(v && *v == '0'); }();
  if (rank_merge && M <= kRankMergeMax) {
    const size_t smem = static_cast<size_t>(M) * 16;
    if (p.packed != nullptr) {
      if (smem > 48 * 1024) { if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(merge_sorted_lists_kernel<true>), static_cast<int>(smem))) return rc; }
      merge_sorted_lists_kernel<true><<<static_cast<unsigned>(p.nq), 128, smem, stream>>>(p);
    } else {
      if (smem > 48 * 1024) { if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(merge_sorted_lists_kernel<false>), static_cast<int>(smem))) return rc; }
      merge_sorted_lists_kernel<false><<<static_cast<unsigned>(p.nq), 128, smem, stream>>>(p);
    }
    NXF_CUDA(cudaGetLastError());
    return NXF_OK;
  }
  if (p.packed != nullptr) merge_lists_kernel<true><<<static_cast<unsigned>(p.nq), mg::THREADS, 0, stream>>>(p);
  else merge_lists_kernel<false><<<static_cast<unsigned>(p.nq), mg::THREADS, 0, stream>>>(p);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// ---- rank of the positive among the top-k (recsys eval_step) --------------------------------------
// One warp per row: rank = #{j : topk[j] > pos}; the reference's cat + stable descending sort puts scores equal to the
// positive after it (modeling.py:199-205).  HBM-bound: nq*k*4 bytes in, nq*(4 + k) out.
__global__ void __launch_bounds__(256) positive_rank_kernel(const float* __restrict__ pos, const float* __restrict__ topk, int64_t nq, int k,
                                                            int32_t* __restrict__ rank, uint8_t* __restrict__ pred) {
  const int64_t b = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  if (b >= nq) return;  // whole warps
  const int lane = threadIdx.x & 31;
  const float p = pos[b];
  const float* row = topk + b * k;
  int c = 0;
  for (int j = lane; j < k; j += 32) c += (row[j] > p) ? 1 : 0;
  c = __reduce_add_sync(0xFFFFFFFFu, c);
  if (lane == 0) rank[b] = c;
  if (pred != nullptr)
    for (int j = lane; j < k; j += 32) pred[b * k + j] = (j == c) ? 1 : 0;
}

int launch_positive_rank(const float* pos, const float* topk, int64_t nq, int k, int32_t* rank, uint8_t* pred, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  positive_rank_kernel<<<static_cast<unsigned>((nq * 32 + 255) / 256), 256, 0, stream>>>(pos, topk, nq, k, rank, pred);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

}  // namespace nxf
