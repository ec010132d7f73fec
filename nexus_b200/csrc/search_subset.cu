// search_subset.cu -- KS: per-query candidate lists ("local ranking").
//
// Replaces the loop at Nexus/evaluation/mmeb_v2/eval_embedding.py:384-404 (eval_type != "global"): for every query the
// reference stacks the embeddings of that query's own candidate names, uploads them, computes q @ cand^T
// (models.py:186-188) and fully sorts the row (torch.sort descending); the metrics then read the first 10 entries
// (utils/eval_utils/metrics.py:12).  Here the candidate embeddings are rows of a FlatIndex (uploaded once), the lists are a
// CSR (offsets [nq + 1], rows [total]) and ONE launch scores every list and returns each query's best k POSITIONS inside
// its own list, best first -- the same numbers the reference's `ranked_idx[:k]` holds.
//
// Bound: HBM (a gather of whole rows).  Algorithmic bytes per launch: total * ld * s (the listed rows, each read once)
// + total * 8 (the row ids) + nq * dim * 4 (queries) + nq * k * 12 (results).
//
// One CTA (256 threads) per query.  The query sits in shared memory as fp32; a row is read with 16-byte loads by LPR lanes
// (8 / 16 / 32 for rows of 64 / 128 / >= 256 elements, so short rows do not idle three quarters of a warp), two rows in
// flight per lane group, fp32 FMAs on the stored values (every product and sum is fp32: no rounding of the query).  Lists
// up to 2048 entries never leave shared memory; longer ones park their (score, position) pairs in a global scratch slice.
// Then block radix select of the k best (when the list is longer than k) and a bitonic sort by (score desc, position asc)
// -- the order of a STABLE descending sort, so ties are deterministic where torch.sort's are not.
#include "params.cuh"
#include "block_sort.cuh"

namespace nxf {

template <typename T, int LPR, bool L2>
__global__ void __launch_bounds__(mg::THREADS) subset_search_kernel(SubsetParams p) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  Entry* ent = reinterpret_cast<Entry*>(dyn_smem);                               // [MAXK] scored entries of a short list
  uint64_t* keys = reinterpret_cast<uint64_t*>(dyn_smem + mg::MAXK * 8);         // [MAXK] the kept entries, then their sort keys
  float* q_s = reinterpret_cast<float*>(dyn_smem + 2 * mg::MAXK * 8);            // [ld]
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh_prefix;
  __shared__ int sh_need, sh_gt, sh_eq;
  __shared__ float sh_qnorm;
  __shared__ float red[mg::THREADS / 32];

  const int64_t qi = blockIdx.x;
  const int tid = threadIdx.x;
  const int ld = static_cast<int>(p.ld);
  constexpr bool l2 = L2;

  int64_t off0 = p.offsets[qi], off1 = p.offsets[qi + 1];
  if (off0 < 0 || off1 < off0 || off1 > p.total || off1 - off0 > (int64_t(1) << 30)) {
    if (tid == 0) atomicOr(p.err, 2u);
    off1 = off0 = 0;
  }
  const int L = static_cast<int>(off1 - off0);
  if (L <= p.short_max) return;   // the warp-per-list kernel serves this query (uniform over the CTA)
  const bool small = L <= mg::MAXK;
  Entry* dst = small ? ent : p.scratch + off0;

  float part = 0.f;
  for (int d = tid; d < ld; d += mg::THREADS) {
    const float v = d < p.dim ? p.q[qi * p.ldq + d] : 0.f;
    q_s[d] = v;
    part += v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(kFullMask, part, o);
  if ((tid & 31) == 0) red[tid >> 5] = part;
  if (tid == 0) { sh_gt = 0; sh_eq = 0; }
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int w = 0; w < mg::THREADS / 32; ++w) s += red[w];
    sh_qnorm = s;
  }

  // ---- score the listed rows ----
  {
    constexpr int GPW = 32 / LPR;                      // lane groups per warp
    constexpr int NG = (mg::THREADS / 32) * GPW;       // lane groups per CTA
    const T* __restrict__ corpus = reinterpret_cast<const T*>(p.corpus);
    const int64_t* __restrict__ rows = p.rows + off0;
    const int lane = tid & 31, sub = lane / LPR, l = lane % LPR;
    const int g = (tid >> 5) * GPW + sub;
    for (int c0 = 0; c0 < L; c0 += 2 * NG) {           // warp-uniform trip count: every lane takes part in the shuffles
      const int ca = c0 + g, cb = c0 + NG + g;
      int64_t ra = ca < L ? rows[ca] : -2, rb = cb < L ? rows[cb] : -2;
      const bool bad_a = ca < L && (ra < 0 || ra >= p.n_rows), bad_b = cb < L && (rb < 0 || rb >= p.n_rows);
      if (bad_a || bad_b) atomicOr(p.err, 1u);
      const bool va = ca < L && !bad_a, vb = cb < L && !bad_b;
      const T* pa = corpus + (va ? ra : 0) * p.ld;
      const T* pb = corpus + (vb ? rb : 0) * p.ld;
      float dot0 = 0.f, nn0 = 0.f, dot1 = 0.f, nn1 = 0.f;
      for (int d0 = l * 8; d0 < ld; d0 += LPR * 8) {
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (va) load8_f32<T>(pa + d0, a);
        if (vb) load8_f32<T>(pb + d0, b);
        const float4 q0 = *reinterpret_cast<const float4*>(q_s + d0), q1 = *reinterpret_cast<const float4*>(q_s + d0 + 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          dot0 = fmaf(qv[e], a[e], dot0);
          dot1 = fmaf(qv[e], b[e], dot1);
          if (L2) { nn0 = fmaf(a[e], a[e], nn0); nn1 = fmaf(b[e], b[e], nn1); }   // IP: half the FMAs (the kernel is close to issue bound)
        }
      }
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) {
        dot0 += __shfl_xor_sync(kFullMask, dot0, o);
        dot1 += __shfl_xor_sync(kFullMask, dot1, o);
        if (L2) { nn0 += __shfl_xor_sync(kFullMask, nn0, o); nn1 += __shfl_xor_sync(kFullMask, nn1, o); }
      }
      if (l == 0) {
        // a row id outside the index is reported through p.err; it ranks behind everything
        if (ca < L) { Entry e; e.score = va ? (l2 ? dot0 - 0.5f * nn0 : dot0) : -FLT_MAX; e.row = ca; dst[ca] = e; }
        if (cb < L) { Entry e; e.score = vb ? (l2 ? dot1 - 0.5f * nn1 : dot1) : -FLT_MAX; e.row = cb; dst[cb] = e; }
      }
    }
  }
  __syncthreads();

  // ---- keep the best min(L, k_out) ----
  // short lists are sorted whole (a bitonic sort of <= 512 keys costs fewer barriers than the four radix passes)
  const int keep = (L <= 512) ? L : min(L, p.k_out);
  Entry* kept = reinterpret_cast<Entry*>(keys);
  const Entry* src = dst;
  if (L > keep) {
    block_radix_select(L, keep, [&](int i) { return okey(src[i].score); }, hist, &sh_prefix, &sh_need);
    const uint32_t kth = sh_prefix;
    const int need = sh_need;
    const int n_gt = keep - need;
    for (int i = tid; i < L; i += mg::THREADS) {
      const Entry e = src[i];
      const uint32_t key = okey(e.score);
      if (key > kth) {
        kept[atomicAdd(&sh_gt, 1)] = e;
      } else if (key == kth) {
        const int t = atomicAdd(&sh_eq, 1);
        if (t < need) kept[n_gt + t] = e;
      }
    }
  } else {
    for (int i = tid; i < L; i += mg::THREADS) kept[i] = src[i];
  }
  __syncthreads();

  // ---- sort by (score desc, position asc) ----
  const int P = next_pow2(max(keep, 2));
  for (int i = tid; i < P; i += mg::THREADS) {   // kept[] aliases keys[] slot for slot: a thread rewrites only the slots it reads
    uint64_t kk = 0;
    if (i < keep) { const Entry e = kept[i]; kk = sort_key(e.score, e.row); }
    keys[i] = kk;
  }
  block_bitonic_desc<false>(keys, nullptr, P);

  const float qn = l2 ? sh_qnorm : 0.f;
  for (int j = tid; j < p.k_out; j += mg::THREADS) {
    float s;
    int64_t pos;
    if (j < keep) {   // keep >= min(L, k_out)
      const uint64_t kk = keys[j];
      const float sc = sort_key_score(kk);
      s = l2 ? fmaxf(qn - 2.f * sc, 0.f) : sc;
      pos = sort_key_row(kk);
    } else {
      s = l2 ? FLT_MAX : -FLT_MAX;
      pos = -1;
    }
    p.out_scores[qi * p.k_out + j] = s;
    p.out_pos[qi * p.k_out + j] = pos;
  }
}

// ---- short lists: one WARP per query --------------------------------------------------------------
// A 256-thread CTA per list is mostly fixed cost on a list of 20 rows (0.44 of the HBM peak on 50 000 x 20 x 768, 0.21 on
// 65 536 x 128 x 64 in round 1).  Lists of up to KS_SHORT_MAX entries are served by one warp each (eight queries per CTA): the query
// in the warp's shared-memory slice, LPR lanes per row and two rows in flight per lane group as above, the (score, position) keys
// in a 2 KB slice, and the order by counting (rank = number of larger keys) -- no block barrier anywhere.
constexpr int KS_SHORT_MAX = 256;
constexpr int KS_WARPS = 8;

template <typename T, int LPR, bool L2>
__global__ void __launch_bounds__(KS_WARPS * 32) subset_search_warp_kernel(SubsetParams p, int64_t nq) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  const int ld = static_cast<int>(p.ld);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q_s = reinterpret_cast<float*>(dyn_smem) + static_cast<size_t>(warp) * ld;                                  // [ld]
  uint64_t* keys = reinterpret_cast<uint64_t*>(dyn_smem + static_cast<size_t>(KS_WARPS) * ld * 4) + warp * KS_SHORT_MAX;  // [KS_SHORT_MAX]
  const T* __restrict__ corpus = reinterpret_cast<const T*>(p.corpus);
  constexpr int GPW = 32 / LPR;
  const int sub = lane / LPR, l = lane % LPR;
  for (int64_t qi = static_cast<int64_t>(blockIdx.x) * KS_WARPS + warp; qi < nq; qi += static_cast<int64_t>(gridDim.x) * KS_WARPS) {
    int64_t off0 = p.offsets[qi], off1 = p.offsets[qi + 1];
    if (off0 < 0 || off1 < off0 || off1 > p.total || off1 - off0 > (int64_t(1) << 30)) {
      if (lane == 0) atomicOr(p.err, 2u);
      off1 = off0 = 0;
    }
    const int L = static_cast<int>(off1 - off0);
    if (L > KS_SHORT_MAX) continue;   // the CTA kernel serves this query
    float part = 0.f;
    for (int d = lane; d < ld; d += 32) {
      const float v = d < p.dim ? p.q[qi * p.ldq + d] : 0.f;
      q_s[d] = v;
      part += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(kFullMask, part, o);
    const float qn = L2 ? part : 0.f;
    __syncwarp();
    const int64_t* __restrict__ rows = p.rows + off0;
    for (int c0 = 0; c0 < L; c0 += 2 * GPW) {   // warp-uniform trip count
      const int ca = c0 + sub, cb = c0 + GPW + sub;
      int64_t ra = ca < L ? rows[ca] : -2, rb = cb < L ? rows[cb] : -2;
      const bool bad_a = ca < L && (ra < 0 || ra >= p.n_rows), bad_b = cb < L && (rb < 0 || rb >= p.n_rows);
      if (bad_a || bad_b) atomicOr(p.err, 1u);
      const bool va = ca < L && !bad_a, vb = cb < L && !bad_b;
      const T* pa = corpus + (va ? ra : 0) * p.ld;
      const T* pb = corpus + (vb ? rb : 0) * p.ld;
      float dot0 = 0.f, nn0 = 0.f, dot1 = 0.f, nn1 = 0.f;
      for (int d0 = l * 8; d0 < ld; d0 += LPR * 8) {
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (va) load8_f32<T>(pa + d0, a);
        if (vb) load8_f32<T>(pb + d0, b);
        const float4 q0 = *reinterpret_cast<const float4*>(q_s + d0), q1 = *reinterpret_cast<const float4*>(q_s + d0 + 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          dot0 = fmaf(qv[e], a[e], dot0);
          dot1 = fmaf(qv[e], b[e], dot1);
          if (L2) { nn0 = fmaf(a[e], a[e], nn0); nn1 = fmaf(b[e], b[e], nn1); }
        }
      }
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) {
        dot0 += __shfl_xor_sync(kFullMask, dot0, o);
        dot1 += __shfl_xor_sync(kFullMask, dot1, o);
        if (L2) { nn0 += __shfl_xor_sync(kFullMask, nn0, o); nn1 += __shfl_xor_sync(kFullMask, nn1, o); }
      }
      if (l == 0) {
        if (ca < L) keys[ca] = sort_key(va ? (L2 ? dot0 - 0.5f * nn0 : dot0) : -FLT_MAX, ca);
        if (cb < L) keys[cb] = sort_key(vb ? (L2 ? dot1 - 0.5f * nn1 : dot1) : -FLT_MAX, cb);
      }
    }
    __syncwarp();
    // order: keys are distinct (positions are), the rank of a key is the number of larger ones
    for (int j = lane; j < L; j += 32) {
      const uint64_t mine = keys[j];
      int rank = 0;
      for (int i = 0; i < L; ++i) rank += (keys[i] > mine) ? 1 : 0;
      if (rank < p.k_out) {
        const float sc = sort_key_score(mine);
        p.out_scores[qi * p.k_out + rank] = L2 ? fmaxf(qn - 2.f * sc, 0.f) : sc;
        p.out_pos[qi * p.k_out + rank] = sort_key_row(mine);
      }
    }
    for (int j = L + lane; j < p.k_out; j += 32) {
      p.out_scores[qi * p.k_out + j] = L2 ? FLT_MAX : -FLT_MAX;
      p.out_pos[qi * p.k_out + j] = -1;
    }
    __syncwarp();   // keys / q_s are reused by this warp's next query
  }
}

// min_len / max_len: shortest / longest list when the caller knows them (host lists), else -1: both kernels are launched and each
// skips the queries of the other (a CTA that finds a short list returns at once).
int launch_subset_search(int store_dtype, const SubsetParams& p_in, int64_t nq, int64_t min_len, int64_t max_len, cudaStream_t stream) {
  if (nq <= 0) return NXF_OK;
  SubsetParams p = p_in;
  static const bool warp_on = []() { const char* v = getenv("NXF_KS_WARP"); return !(v && *v == '0'); }();
  const size_t wsmem = static_cast<size_t>(KS_WARPS) * (static_cast<size_t>(p.ld) * 4 + KS_SHORT_MAX * 8);
  const bool warp_ok = warp_on && wsmem <= 200 * 1024;
  const bool run_warp = warp_ok && (min_len < 0 || min_len <= KS_SHORT_MAX);
  const bool run_cta = !run_warp || max_len < 0 || max_len > KS_SHORT_MAX;
  p.short_max = run_warp ? KS_SHORT_MAX : -1;
  if (run_warp) {
    int dev = 0, sms = 148;
    NXF_CUDA(cudaGetDevice(&dev));
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = static_cast<int>(std::min<int64_t>((nq + KS_WARPS - 1) / KS_WARPS, static_cast<int64_t>(sms) * 8));
#define NXF_LAUNCH_KSW3(T, LPR, L2)                                                                                             \
  do {                                                                                                                          \
    if (wsmem > 48 * 1024) {                                                                                                    \
      const int rc_ = ensure_dyn_smem(reinterpret_cast<const void*>(subset_search_warp_kernel<T, LPR, L2>), (int)wsmem);        \
      if (rc_) return rc_;                                                                                                      \
    }                                                                                                                           \
    subset_search_warp_kernel<T, LPR, L2><<<grid, KS_WARPS * 32, wsmem, stream>>>(p, nq);                                       \
  } while (0)
#define NXF_LAUNCH_KSW2(T, LPR)                                      \
  do {                                                               \
    if (p.metric == NXF_METRIC_L2) NXF_LAUNCH_KSW3(T, LPR, true);    \
    else NXF_LAUNCH_KSW3(T, LPR, false);                             \
  } while (0)
#define NXF_LAUNCH_KSW(T)                          \
  do {                                             \
    if (p.ld >= 256) NXF_LAUNCH_KSW2(T, 32);       \
    else if (p.ld >= 128) NXF_LAUNCH_KSW2(T, 16);  \
    else NXF_LAUNCH_KSW2(T, 8);                    \
  } while (0)
    switch (store_dtype) {
      case NXF_F32: NXF_LAUNCH_KSW(float); break;
      case NXF_F16: NXF_LAUNCH_KSW(__half); break;
      case NXF_BF16: NXF_LAUNCH_KSW(__nv_bfloat16); break;
      default: set_error("search_subsets: bad storage dtype %d", store_dtype); return NXF_EINVAL;
    }
#undef NXF_LAUNCH_KSW
#undef NXF_LAUNCH_KSW2
#undef NXF_LAUNCH_KSW3
    NXF_CUDA(cudaGetLastError());
    if (!run_cta) return NXF_OK;
  }
  const size_t smem = static_cast<size_t>(mg::MAXK) * 16 + static_cast<size_t>(p.ld) * 4;
  if (smem > 200 * 1024) { set_error("search_subsets: dim %d too large", p.dim); return NXF_EUNSUPPORTED; }
#define NXF_LAUNCH_SUBSET3(T, LPR, L2)                                                                                          \
  do {                                                                                                                          \
    if (smem > 48 * 1024) {                                                                                                     \
      const int rc_ = ensure_dyn_smem(reinterpret_cast<const void*>(subset_search_kernel<T, LPR, L2>), (int)smem);              \
      if (rc_) return rc_;                                                                                                      \
    }                                                                                                                           \
    subset_search_kernel<T, LPR, L2><<<static_cast<unsigned>(nq), mg::THREADS, smem, stream>>>(p);                              \
  } while (0)
#define NXF_LAUNCH_SUBSET2(T, LPR)                                      \
  do {                                                                  \
    if (p.metric == NXF_METRIC_L2) NXF_LAUNCH_SUBSET3(T, LPR, true);    \
    else NXF_LAUNCH_SUBSET3(T, LPR, false);                             \
  } while (0)
#define NXF_LAUNCH_SUBSET(T)                          \
  do {                                                \
    if (p.ld >= 256) NXF_LAUNCH_SUBSET2(T, 32);       \
    else if (p.ld >= 128) NXF_LAUNCH_SUBSET2(T, 16);  \
    else NXF_LAUNCH_SUBSET2(T, 8);                    \
  } while (0)
  switch (store_dtype) {
    case NXF_F32: NXF_LAUNCH_SUBSET(float); break;
    case NXF_F16: NXF_LAUNCH_SUBSET(__half); break;
    case NXF_BF16: NXF_LAUNCH_SUBSET(__nv_bfloat16); break;
    default: set_error("search_subsets: bad storage dtype %d", store_dtype); return NXF_EINVAL;
  }
#undef NXF_LAUNCH_SUBSET
#undef NXF_LAUNCH_SUBSET2
#undef NXF_LAUNCH_SUBSET3
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

}  // namespace nxf
