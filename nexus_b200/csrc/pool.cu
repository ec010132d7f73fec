// pool.cu -- K0: fused pool + L2-normalise (+ cast) epilogue over encoder last-hidden-states.
//
// One kernel replaces the chain the reference runs after every encoder forward:
//   pooling()      Nexus/inference/embedder/text_retrieval/base.py:355-379  (cls | mean)
//                  .../decoder/base.py:330-342                              (last token, padding-side rule)
//                  Nexus/evaluation/mmeb_v2/models.py:146-151               (last token by flip/argmax)
//   F.normalize    base.py:334-335 (x / max(||x||, eps)); host-numpy form base.py:672-673 (no eps)
// HBM-bound: algorithmic bytes = B*L*H*sizeof(T) + B*L*8 + B*H*sizeof(O) for mean,
//            B*H*(sizeof(T)+sizeof(O)) + B*L*8 for cls / last.
// Grid: one CTA per batch row; 16-byte vector loads along H; the mean path skips masked positions
// entirely (the reference multiplies them by 0) and splits L over thread groups.
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

namespace nxf {

#ifndef NXF_POOL_CTAS
#define NXF_POOL_CTAS (6 * 3 * 148)
#endif
#ifndef NXF_POOL_WS_CACHE_DEFAULT
#define NXF_POOL_WS_CACHE_DEFAULT 1   // 1: the split kernel's workspace is cached per (device, stream) (NXF_POOL_WS_CACHE overrides)
#endif
#ifndef NXF_POOL_MINB
#define NXF_POOL_MINB 3   // min resident CTAs per SM asked of the split kernel (register cap: 54 -> 40 registers, 2 -> 3 CTAs of 480 threads)
#endif
#ifndef NXF_POOL_UNROLL
#define NXF_POOL_UNROLL 4
#endif
namespace k0 {
constexpr int MEAN_UNROLL = NXF_POOL_UNROLL;
constexpr int THREADS = 256;
constexpr int MAX_GROUPS = 8;
}  // namespace k0

__device__ __forceinline__ float block_sum_f(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
  for (int w = 0; w < k0::THREADS / 32; ++w) s += red[w];
  return s;
}
__device__ __forceinline__ long long block_sum_ll(long long v, long long* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  long long s = 0;
  for (int w = 0; w < k0::THREADS / 32; ++w) s += red[w];
  return s;
}

template <typename T, int VEC> struct VecLoad;
template <typename T> struct VecLoad<T, 1> {
  __device__ static __forceinline__ void load(const T* p, float (&x)[1]) { x[0] = to_f32<T>(*p); }
};
template <> struct VecLoad<float, 4> {
  __device__ static __forceinline__ void load(const float* p, float (&x)[4]) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
  }
};
template <> struct VecLoad<__half, 8> {
  __device__ static __forceinline__ void load(const __half* p, float (&x)[8]) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i])); x[2 * i] = f.x; x[2 * i + 1] = f.y; }
  }
};
template <> struct VecLoad<__nv_bfloat16, 8> {
  __device__ static __forceinline__ void load(const __nv_bfloat16* p, float (&x)[8]) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { x[2 * i] = __uint_as_float(w[i] << 16); x[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u); }
  }
};

template <typename T, typename O, int VEC>
__global__ void __launch_bounds__(k0::THREADS) pool_norm_kernel(const T* __restrict__ hidden, const int64_t* __restrict__ mask,
                                                                int64_t B, int64_t L, int64_t H, int pooling, int normalize,
                                                                float eps, O* __restrict__ out, int groups) {
  extern __shared__ __align__(16) float vals[];  // [groups][H]
  __shared__ float red_f[k0::THREADS / 32];
  __shared__ long long red_ll[k0::THREADS / 32];
  const int tid = threadIdx.x;
  const int64_t b = blockIdx.x;
  const T* hb = hidden + b * L * H;
  const int64_t* mb = mask ? mask + b * L : nullptr;
  const int nvec = static_cast<int>(H / VEC);

  if (pooling == NXF_POOL_MEAN) {
    // thread (g, v): positions l = g, g+groups, ...; vector column v, v+tpg, ...
    const int tpg = k0::THREADS / groups;  // threads per group
    const int g = tid / tpg;
    const int v0 = tid - g * tpg;
    long long wsum = 0;
    for (int64_t l = tid; l < L; l += k0::THREADS) wsum += mb[l];
    const long long den = block_sum_ll(wsum, red_ll);
    for (int v = v0; v < nvec; v += tpg) {
      float acc[VEC];
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] = 0.f;
      if (g < groups) {
#pragma unroll 4
        for (int64_t l = g; l < L; l += groups) {
          const int64_t w = mb[l];
          if (w != 0) {
            float x[VEC];
            VecLoad<T, VEC>::load(hb + l * H + static_cast<int64_t>(v) * VEC, x);
            const float wf = static_cast<float>(w);
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] = fmaf(x[e], wf, acc[e]);
          }
        }
#pragma unroll
        for (int e = 0; e < VEC; ++e) vals[static_cast<int64_t>(g) * H + v * VEC + e] = acc[e];
      }
    }
    __syncthreads();
    const float fden = static_cast<float>(den);
    for (int64_t h = tid; h < H; h += k0::THREADS) {
      float s = 0.f;
      for (int gg = 0; gg < groups; ++gg) s += vals[static_cast<int64_t>(gg) * H + h];
      vals[h] = s / fden;  // 0/0 -> NaN for an empty mask, like the reference
    }
  } else {
    int64_t pos = 0;
    if (pooling == NXF_POOL_LAST) {
      long long lastcol = 0, rowsum = 0;
      for (int64_t bb = tid; bb < B; bb += k0::THREADS) lastcol += mask[bb * L + (L - 1)];
      for (int64_t l = tid; l < L; l += k0::THREADS) rowsum += mb[l];
      lastcol = block_sum_ll(lastcol, red_ll);
      rowsum = block_sum_ll(rowsum, red_ll);
      if (lastcol == B) pos = L - 1;                 // left padding: h[:, -1]      decoder/base.py:336-338
      else { pos = rowsum - 1; if (pos < 0) pos += L; if (pos < 0) pos = 0; if (pos >= L) pos = L - 1; }  // :340-342
    } else if (pooling == NXF_POOL_LAST_ARGMAX) {
      // flip().argmax(): first maximum of the reversed mask = last position holding the row maximum
      long long best = LLONG_MIN, bestpos = 0;
      for (int64_t l = tid; l < L; l += k0::THREADS) {
        const long long w = mb[l];
        if (w > best || (w == best && l > bestpos)) { best = w; bestpos = l; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const long long ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
        const long long op = __shfl_xor_sync(0xFFFFFFFFu, bestpos, o);
        if (ob > best || (ob == best && op > bestpos)) { best = ob; bestpos = op; }
      }
      __shared__ long long sb[k0::THREADS / 32], sp[k0::THREADS / 32];
      if ((tid & 31) == 0) { sb[tid >> 5] = best; sp[tid >> 5] = bestpos; }
      __syncthreads();
      best = sb[0]; bestpos = sp[0];
      for (int w = 1; w < k0::THREADS / 32; ++w)
        if (sb[w] > best || (sb[w] == best && sp[w] > bestpos)) { best = sb[w]; bestpos = sp[w]; }
      pos = bestpos;
    }
    const T* src = hb + pos * H;
    for (int v = tid; v < nvec; v += k0::THREADS) {
      float x[VEC];
      VecLoad<T, VEC>::load(src + static_cast<int64_t>(v) * VEC, x);
#pragma unroll
      for (int e = 0; e < VEC; ++e) vals[v * VEC + e] = x[e];
    }
  }
  __syncthreads();

  float scale = 1.f;
  if (normalize) {
    float ss = 0.f;
    for (int64_t h = tid; h < H; h += k0::THREADS) ss = fmaf(vals[h], vals[h], ss);
    ss = block_sum_f(ss, red_f);
    const float nrm = sqrtf(ss);
    scale = 1.f / (eps > 0.f ? fmaxf(nrm, eps) : nrm);
  }
  O* ob = out + b * H;
  for (int64_t h = tid; h < H; h += k0::THREADS) ob[h] = from_f32<O>(normalize ? vals[h] * scale : vals[h]);
}

// ---- mean pooling at the HBM roofline ------------------------------------------------------------
// One CTA per batch row cannot fill 148 SMs (B is 64..256 where L is long), so L is split over `splits` CTAs per row:
// grid (splits, B), block = G groups x nvec threads (thread (g, v) owns vector column v and positions l0+g, l0+g+G, ...;
// consecutive threads read consecutive 16-byte vectors of one position, masked positions are skipped).  Every CTA
// writes its fp32 partial sums to a workspace [B][splits][H]; the last CTA of a row to arrive (one atomic counter per
// row) adds the partials in a fixed order -- results do not depend on the arrival order --, divides by the mask sum,
// normalises, casts and stores.
template <typename T, typename O, int VEC>
__global__ void __launch_bounds__(512, NXF_POOL_MINB) pool_mean_split_kernel(const T* __restrict__ hidden, const int64_t* __restrict__ mask, int64_t L,
                                                              int64_t H, int normalize, float eps, O* __restrict__ out, int G,
                                                              float* __restrict__ partial, unsigned long long* __restrict__ den_g,
                                                              unsigned int* __restrict__ counter) {
  extern __shared__ __align__(16) float vals[];  // [G][H], reused as [H] by the finalising CTA
  __shared__ float red_f[16];
  __shared__ unsigned int s_last;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int splits = gridDim.x, sp = blockIdx.x;
  const int64_t b = blockIdx.y;
  const int nvec = static_cast<int>(H / VEC);
  const int64_t lc = (L + splits - 1) / splits;
  const int64_t l0 = sp * lc, l1 = min(L, l0 + lc);
  const T* hb = hidden + b * L * H;
  const int64_t* mb = mask + b * L;

  __shared__ int s_lo, s_hi;
  if (tid == 0) { s_lo = INT_MAX; s_hi = -1; }
  __syncthreads();
  unsigned long long wsum = 0;
  int lo = INT_MAX, hi = -1;  // first / last unmasked position of the slice, relative to l0
  for (int64_t l = l0 + tid; l < l1; l += nthr) {
    const int64_t w = mb[l];
    wsum += static_cast<unsigned long long>(w);
    if (w != 0) { lo = min(lo, static_cast<int>(l - l0)); hi = max(hi, static_cast<int>(l - l0)); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xFFFFFFFFu, wsum, o);
  lo = __reduce_min_sync(0xFFFFFFFFu, lo);
  hi = __reduce_max_sync(0xFFFFFFFFu, hi);
  if ((tid & 31) == 0) {
    if (wsum != 0) atomicAdd(&den_g[b], wsum);
    atomicMin(&s_lo, lo);
    atomicMax(&s_hi, hi);
  }
  __syncthreads();
  const int64_t n0 = l0 + s_lo, n1 = (s_hi >= 0) ? l0 + s_hi + 1 : n0;  // empty when the whole slice is masked

  const int g = tid / nvec, v0 = tid - g * nvec;  // threads beyond G*nvec (block rounded up to a warp) idle
  if (g < G) {
    for (int v = v0; v < nvec; v += nvec) {  // block covers all vector columns when nvec <= 512 (H <= 4096 for 16-bit types)
      float acc[VEC];
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] = 0.f;
      const T* col = hb + static_cast<int64_t>(v) * VEC;
      // unconditional loads over [n0, n1) -- the part of the slice between its first and last unmasked position -- so that the
      // unrolled loads are independent of the mask words; a masked position inside that range contributes h * 0 like the
      // reference's h * mask.float()
#pragma unroll k0::MEAN_UNROLL
      for (int64_t l = n0 + g; l < n1; l += G) {
#ifdef NXF_POOL_COND
        const int64_t w = mb[l];
        if (w != 0) {
          float x[VEC];
          VecLoad<T, VEC>::load(col + l * H, x);
          const float wf = static_cast<float>(w);
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[e] = fmaf(x[e], wf, acc[e]);
        }
#else
        float x[VEC];
        VecLoad<T, VEC>::load(col + l * H, x);
        const float wf = static_cast<float>(mb[l]);
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] = fmaf(x[e], wf, acc[e]);
#endif
      }
#pragma unroll
      for (int e = 0; e < VEC; ++e) vals[static_cast<int64_t>(g) * H + v * VEC + e] = acc[e];
    }
  }
  __syncthreads();
  float* my_part = partial + (b * splits + sp) * H;
  for (int64_t h = tid; h < H; h += nthr) {
    float s = 0.f;
    for (int gg = 0; gg < G; ++gg) s += vals[static_cast<int64_t>(gg) * H + h];
    my_part[h] = s;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(&counter[b], 1u) == static_cast<unsigned>(splits - 1)) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const float fden = static_cast<float>(static_cast<long long>(*reinterpret_cast<volatile unsigned long long*>(&den_g[b])));
  float ss = 0.f;
  for (int64_t h = tid; h < H; h += nthr) {
    float s = 0.f;
    for (int j = 0; j < splits; ++j) s += __ldcg(partial + (b * splits + j) * H + h);
    s /= fden;  // 0/0 -> NaN for an empty mask, like the reference
    vals[h] = s;
    ss = fmaf(s, s, ss);
  }
  float scale = 1.f;
  if (normalize) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, o);
    if ((tid & 31) == 0) red_f[tid >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < (nthr + 31) / 32; ++w) tot += red_f[w];
    const float nrm = sqrtf(tot);
    scale = 1.f / (eps > 0.f ? fmaxf(nrm, eps) : nrm);
  }
  O* ob = out + b * H;
  for (int64_t h = tid; h < H; h += nthr) ob[h] = from_f32<O>(normalize ? vals[h] * scale : vals[h]);
  // leave the row's control words as they were found (zero): the workspace is cached per (device, stream) and the next launch on
  // the stream starts without a memset.  Nobody else touches them any more: every CTA of the row has already arrived.
  __syncthreads();  // every thread of this (last) CTA has read den_g[b] by now
  if (tid == 0) { den_g[b] = 0ull; counter[b] = 0u; }
}

// Workspace of the split kernel, cached per (device, stream): [control words of up to 65 535 rows (den u64 | counter u32), zero
// between launches][fp32 partial sums].  Launches on one stream are ordered, so they can share it; other streams get their own.
// (cudaMallocAsync + memset + cudaFreeAsync per call were a third of the host time of a default encode batch's pooling call.)
namespace {
struct PoolWs { unsigned char* p = nullptr; size_t bytes = 0; };
constexpr size_t kPoolCtlBytes = size_t(65536) * 16;
std::mutex g_pool_mu;
std::map<std::pair<int, cudaStream_t>, PoolWs> g_pool_ws;

int pool_workspace(size_t part_bytes, cudaStream_t stream, unsigned char** out) {
  int dev = 0;
  NXF_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_pool_mu);
  PoolWs& w = g_pool_ws[{dev, stream}];
  const size_t want = kPoolCtlBytes + part_bytes;
  if (w.bytes < want) {
    if (w.p) { NXF_CUDA(cudaStreamSynchronize(stream)); cudaFree(w.p); w.p = nullptr; w.bytes = 0; }
    const size_t alloc = (want + (size_t(4) << 20)) & ~((size_t(1) << 20) - 1);
    NXF_CUDA(cudaMalloc(reinterpret_cast<void**>(&w.p), alloc));
    w.bytes = alloc;
    NXF_CUDA(cudaMemsetAsync(w.p, 0, kPoolCtlBytes, stream));
  }
  *out = w.p;
  return NXF_OK;
}
}  // namespace

// ---- mean pooling as a persistent TMA stream (round 2) ---------------------------------------------
// The split kernel above runs ~2 000 short CTAs whose fixed path (mask scan -> first loads -> shared-memory reduce -> fence -> atomic)
// is as long as their streaming time: 0.38-0.46 of the HBM peak on the reference's default encode batch (256 x 512 x 768 f16).
// Here ONE CTA per SM stays resident and a producer warp keeps a ring of shared-memory stages full with cp.async.bulk copies --
// the positions of a row are contiguous in [B, L, H], so a chunk of positions is one contiguous block -- while eight consumer
// warps accumulate from shared memory.  Work = segments (row b, slice sp of `splits`), fetched from a global counter by the
// producer, which also scans the segment's mask (first / last unmasked position, weights) while earlier chunks are in flight: the
// per-segment fixed costs overlap the stream instead of serialising with it.  splits == 1: the CTA finalises the row itself (no
// workspace traffic); splits > 1: partial sums + last-arriver finalisation as in the split kernel (same fixed summation order).
namespace k0s {
constexpr int CONSUMER_WARPS = 8;
constexpr int CONSUMERS = CONSUMER_WARPS * 32;
constexpr int THREADS = CONSUMERS + 32;
constexpr int MAX_STAGES = 8;
constexpr int MAX_CH = 32;  // positions per stage (one producer lane per position)
struct __align__(16) StageMeta {
  int seg;        // segment id, -1 = no more work
  int npos;       // positions in this stage
  int last;       // last stage of its segment
  int pad;
  unsigned long long wsum;  // (last stage) sum of the segment's mask words
  float w[MAX_CH];
};
}  // namespace k0s

template <typename T, typename O, int VEC>
__global__ void __launch_bounds__(k0s::THREADS, 1) pool_mean_stream_kernel(const T* __restrict__ hidden, const int64_t* __restrict__ mask, int64_t B,
                                                                          int64_t L, int64_t H, int normalize, float eps, O* __restrict__ out,
                                                                          int splits, int CH, int stages, int G, float* __restrict__ partial,
                                                                          unsigned long long* __restrict__ den_g, unsigned int* __restrict__ counter,
                                                                          unsigned int* __restrict__ work_ctr, unsigned int* __restrict__ exit_ctr) {
  using namespace k0s;
  extern __shared__ __align__(128) unsigned char sm[];
  const int64_t row_bytes = H * static_cast<int64_t>(sizeof(T));
  const int64_t stage_bytes = (row_bytes * CH + 127) / 128 * 128;
  unsigned char* ring = sm;
  StageMeta* meta = reinterpret_cast<StageMeta*>(ring + stage_bytes * stages);
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + MAX_STAGES);
  uint64_t* empty = full + MAX_STAGES;
  float* vals = reinterpret_cast<float*>(empty + MAX_STAGES);   // [G][H], reused as [H]
  __shared__ float red_f[CONSUMER_WARPS];
  __shared__ unsigned int s_last;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const unsigned lane = tid & 31;
  const int64_t n_seg = B * splits;
  const int64_t lc = (L + splits - 1) / splits;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CONSUMER_WARPS); }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == CONSUMER_WARPS) {
    // ------------------------------- producer warp -------------------------------
    // Work is walked as (segment, sub-block) pairs; a sub-block = up to 32 * MR positions whose mask words the warp holds in
    // registers (lane j: positions base + j + 32 r).  The NEXT pair's mask words are loaded before the current pair's chunks are
    // pushed, so the producer never waits for a mask load between two bulk copies (a first version read the weights of every
    // chunk from global memory right before its copy: ~0.7 us per chunk, twice the chunk's HBM time).
    constexpr int MR = 8;
    const int SB = (32 * MR / CH) * CH;   // positions per sub-block: a multiple of the chunk, so chunks never straddle sub-blocks
    int s = 0;
    uint32_t ph = 0;
    struct Pair { int64_t seg, base, end; };
    // segment ids are fetched ONE SEGMENT AHEAD: the atomic's round trip (~1 us) overlaps the pushes of the current segment
    // instead of stalling the producer once per segment (short rows: as long as the segment's own HBM time)
    auto issue_fetch = [&]() -> unsigned int { return lane == 0 ? atomicAdd(work_ctr, 1u) : 0u; };
    auto take = [&](unsigned int v) -> int64_t { return static_cast<int64_t>(__shfl_sync(0xFFFFFFFFu, v, 0)); };
    auto first_of = [&](int64_t seg) -> Pair {
      Pair pr; pr.seg = seg;
      if (seg >= n_seg) { pr.base = 0; pr.end = 0; return pr; }
      const int64_t b = seg / splits;
      const int64_t l0 = (seg - b * splits) * lc;
      pr.base = l0; pr.end = min(L, l0 + lc);
      return pr;
    };
    auto load_words = [&](const Pair& pr, int64_t (&m)[MR]) {
#pragma unroll
      for (int r = 0; r < MR; ++r) m[r] = 0;
      if (pr.seg >= n_seg) return;
      const int64_t* mb = mask + (pr.seg / splits) * L;
      const int64_t stop = min(pr.end, pr.base + SB);
#pragma unroll
      for (int r = 0; r < MR; ++r) {
        const int64_t l = pr.base + lane + 32 * r;
        if (l < stop) m[r] = mb[l];
      }
    };
    Pair nxt = first_of(take(issue_fetch()));
    unsigned int ahead = issue_fetch();
    int64_t nxt_m[MR];
    load_words(nxt, nxt_m);
    unsigned long long seg_wsum = 0;
    for (;;) {
      const Pair cur = nxt;
      int64_t m[MR];
#pragma unroll
      for (int r = 0; r < MR; ++r) m[r] = nxt_m[r];
      if (cur.seg >= n_seg) {
        mbar_wait(&empty[s], ph ^ 1);
        if (lane == 0) { meta[s].seg = -1; meta[s].npos = 0; meta[s].last = 0; }
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
        break;
      }
      const bool last_sub = cur.base + SB >= cur.end;
      if (last_sub) { nxt = first_of(take(ahead)); ahead = issue_fetch(); }
      else { nxt = cur; nxt.base = cur.base + SB; }
      load_words(nxt, nxt_m);   // in flight while the chunks of `cur` are pushed

      const int64_t b = cur.seg / splits;
      const int64_t stop = min(cur.end, cur.base + SB);
      unsigned long long wsum = 0;
      int lo = INT_MAX, hi = -1;
#pragma unroll
      for (int r = 0; r < MR; ++r) {
        wsum += static_cast<unsigned long long>(m[r]);
        if (m[r] != 0) { lo = min(lo, static_cast<int>(lane) + 32 * r); hi = max(hi, static_cast<int>(lane) + 32 * r); }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xFFFFFFFFu, wsum, o);
      lo = __reduce_min_sync(0xFFFFFFFFu, lo);
      hi = __reduce_max_sync(0xFFFFFFFFu, hi);
      seg_wsum += wsum;
      const unsigned char* hb = reinterpret_cast<const unsigned char*>(hidden + b * L * H);
      // chunks over the unmasked span of the sub-block, aligned to the chunk grid of the sub-block
      if (hi >= 0) {
        for (int c0 = (lo / CH) * CH; c0 <= hi; c0 += CH) {
          const int first = max(c0, lo);
          const int np = static_cast<int>(min(static_cast<int64_t>(min(c0 + CH, hi + 1)), stop - cur.base)) - first;
          // weight of position first + lane: word (first + lane) of the sub-block lives in lane (idx % 32), register idx / 32
          const int idx = first + static_cast<int>(lane);
          float wv = 0.f;
#pragma unroll
          for (int r = 0; r < MR; ++r) {
            const int64_t t = __shfl_sync(0xFFFFFFFFu, m[r], idx & 31);
            if ((idx >> 5) == r) wv = static_cast<float>(t);
          }
          mbar_wait(&empty[s], ph ^ 1);
          NXF_DEV_CHECK(np >= 1 && np <= CH && cur.base + first + np <= cur.end, "K0 chunk outside its segment");
          if (static_cast<int>(lane) < np) meta[s].w[lane] = wv;
          if (lane == 0) { meta[s].seg = static_cast<int>(cur.seg); meta[s].npos = np; meta[s].last = 0; meta[s].wsum = 0ull; }
          __syncwarp();
          if (lane == 0) {
            const uint32_t bytes = static_cast<uint32_t>(np * row_bytes);
            mbar_expect_tx(&full[s], bytes);
            bulk_load_1d(ring + s * stage_bytes, hb + (cur.base + first) * row_bytes, bytes, &full[s]);
          }
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
      if (last_sub) {
        // the segment's terminator: no data, the mask sum of the whole segment
        mbar_wait(&empty[s], ph ^ 1);
        if (lane == 0) { meta[s].seg = static_cast<int>(cur.seg); meta[s].npos = 0; meta[s].last = 1; meta[s].wsum = seg_wsum; }
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
        if (++s == stages) { s = 0; ph ^= 1; }
        seg_wsum = 0;
      }
    }
    // the last producer to leave puts the two scheduling words back to zero for the next launch on this workspace
    if (lane == 0) {
      __threadfence();
      if (atomicAdd(exit_ctr, 1u) == gridDim.x - 1) { *work_ctr = 0u; *exit_ctr = 0u; __threadfence(); }
    }
    return;
  }

  // ------------------------------------ consumers ------------------------------------
  const int nvec = static_cast<int>(H / VEC);
  const int g = tid / nvec, v = tid - g * nvec;
  const bool active = g < G;
  float acc[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) acc[e] = 0.f;
  int s = 0;
  uint32_t ph = 0;
  for (;;) {
    mbar_wait(&full[s], ph);
    const int seg = meta[s].seg;
    const int npos = meta[s].npos;
    const int last = meta[s].last;
    const unsigned long long wsum = meta[s].wsum;
    if (active) {
      const T* st = reinterpret_cast<const T*>(ring + s * stage_bytes) + static_cast<int64_t>(v) * VEC;
#pragma unroll 4
      for (int pp = g; pp < npos; pp += G) {
        float x[VEC];
        VecLoad<T, VEC>::load(st + static_cast<int64_t>(pp) * H, x);
        const float wf = meta[s].w[pp];
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] = fmaf(x[e], wf, acc[e]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
    if (++s == stages) { s = 0; ph ^= 1; }
    if (seg < 0) break;
    if (!last) continue;

    // ---- end of a segment ----
    const int64_t b = seg / splits;
    const int sp = seg - static_cast<int>(b) * splits;
    if (active) {
#pragma unroll
      for (int e = 0; e < VEC; ++e) { vals[static_cast<int64_t>(g) * H + v * VEC + e] = acc[e]; acc[e] = 0.f; }
    }
    asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(CONSUMERS) : "memory");
    float fden;
    if (splits == 1) {
      for (int64_t h = tid; h < H; h += CONSUMERS) {
        float t = 0.f;
        for (int gg = 0; gg < G; ++gg) t += vals[static_cast<int64_t>(gg) * H + h];
        vals[h] = t;   // group 0's slot: only this thread reads and writes column h
      }
      fden = static_cast<float>(static_cast<long long>(wsum));
    } else {
      float* my_part = partial + (b * splits + sp) * H;
      for (int64_t h = tid; h < H; h += CONSUMERS) {
        float t = 0.f;
        for (int gg = 0; gg < G; ++gg) t += vals[static_cast<int64_t>(gg) * H + h];
        my_part[h] = t;
      }
      __threadfence();
      asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(CONSUMERS) : "memory");
      if (tid == 0) {
        if (wsum != 0) atomicAdd(&den_g[b], wsum);
        __threadfence();
        s_last = (atomicAdd(&counter[b], 1u) == static_cast<unsigned>(splits - 1)) ? 1u : 0u;
      }
      asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(CONSUMERS) : "memory");
      if (!s_last) continue;
      __threadfence();
      fden = static_cast<float>(static_cast<long long>(*reinterpret_cast<volatile unsigned long long*>(&den_g[b])));
      for (int64_t h = tid; h < H; h += CONSUMERS) {
        float t = 0.f;
        for (int j = 0; j < splits; ++j) t += __ldcg(partial + (b * splits + j) * H + h);
        vals[h] = t;
      }
    }
    // divide, normalise, cast, store (vals[h] is touched by thread h % CONSUMERS only until the barrier below)
    float ss = 0.f;
    for (int64_t h = tid; h < H; h += CONSUMERS) {
      const float t = vals[h] / fden;  // 0/0 -> NaN for an empty mask, like the reference
      vals[h] = t;
      ss = fmaf(t, t, ss);
    }
    float scale = 1.f;
    if (normalize) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, o);
      if (lane == 0) red_f[warp] = ss;
      asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(CONSUMERS) : "memory");
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < CONSUMER_WARPS; ++w) tot += red_f[w];
      const float nrm = sqrtf(tot);
      scale = 1.f / (eps > 0.f ? fmaxf(nrm, eps) : nrm);
    }
    O* ob = out + b * H;
    for (int64_t h = tid; h < H; h += CONSUMERS) ob[h] = from_f32<O>(normalize ? vals[h] * scale : vals[h]);
    if (splits > 1 && tid == 0) { den_g[b] = 0ull; counter[b] = 0u; }   // leave the row's control words as they were found
    asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(CONSUMERS) : "memory");   // vals / red_f free for the next segment
  }
}

template <typename T, typename O, int VEC>
static int launch_pool_mean_split(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int normalize, float eps,
                                  void* out, cudaStream_t stream) {
  const int nvec = static_cast<int>(H / VEC);
  const int G = std::max(1, 512 / nvec);
  const int threads = std::min(512, (G * nvec + 31) / 32 * 32);
  // ~3 CTAs of this size are resident per SM: aim at >= 6 waves so that the last, partial wave costs little
  int splits = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>((NXF_POOL_CTAS + B - 1) / B, std::max<int64_t>(1, L / 64))));
  const size_t smem = static_cast<size_t>(G) * H * 4;
  const size_t part_bytes = static_cast<size_t>(B) * splits * H * 4;
  auto kern = pool_mean_split_kernel<T, O, VEC>;
  if (smem > 48 * 1024) { if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem))) return rc; }
  const char* ws_env = getenv("NXF_POOL_WS_CACHE");  // read per call
  const bool cached_ws = ws_env ? atoi(ws_env) != 0 : NXF_POOL_WS_CACHE_DEFAULT != 0;
  if (cached_ws) {
    unsigned char* ws = nullptr;
    int rc = pool_workspace(part_bytes, stream, &ws);
    if (rc) return rc;
    unsigned long long* den = reinterpret_cast<unsigned long long*>(ws);
    unsigned int* counter = reinterpret_cast<unsigned int*>(ws + size_t(65536) * 8);
    float* partial = reinterpret_cast<float*>(ws + kPoolCtlBytes);
    kern<<<dim3(static_cast<unsigned>(splits), static_cast<unsigned>(B)), threads, smem, stream>>>(
        (const T*)hidden, mask, L, H, normalize, eps, (O*)out, G, partial, den, counter);
    NXF_CUDA(cudaGetLastError());
    return NXF_OK;
  }
  // stream-ordered allocation per call: partials | den (u64 per row) | counter (u32 per row), control words zeroed by a memset
  const size_t ctl_bytes = static_cast<size_t>(B) * 16;
  unsigned char* ws = nullptr;
  NXF_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&ws), part_bytes + ctl_bytes, stream));
  unsigned long long* den = reinterpret_cast<unsigned long long*>(ws + part_bytes);
  unsigned int* counter = reinterpret_cast<unsigned int*>(ws + part_bytes + static_cast<size_t>(B) * 8);
  cudaError_t e = cudaMemsetAsync(ws + part_bytes, 0, ctl_bytes, stream);
  if (e == cudaSuccess) {
    kern<<<dim3(static_cast<unsigned>(splits), static_cast<unsigned>(B)), threads, smem, stream>>>(
        (const T*)hidden, mask, L, H, normalize, eps, (O*)out, G, reinterpret_cast<float*>(ws), den, counter);
    e = cudaGetLastError();
  }
  cudaFreeAsync(ws, stream);
  NXF_CUDA(e);
  return NXF_OK;
}

// Returns NXF_OK after launching, or 1 when the shape does not fit the stream kernel (the caller falls back to the split kernel).
template <typename T, typename O, int VEC>
static int launch_pool_mean_stream(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int normalize, float eps,
                                   void* out, cudaStream_t stream) {
  using namespace k0s;
  const int nvec = static_cast<int>(H / VEC);
  if (nvec > CONSUMERS || B >= 65535) return 1;
  const int G = std::max(1, CONSUMERS / nvec);
  const int64_t row_bytes = H * static_cast<int64_t>(sizeof(T));
  const int CH = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(MAX_CH, 32768 / row_bytes)));
  const int64_t stage_bytes = (row_bytes * CH + 127) / 128 * 128;
  const int64_t fixed = static_cast<int64_t>(MAX_STAGES) * sizeof(StageMeta) + 2 * MAX_STAGES * 8 + static_cast<int64_t>(G) * H * 4 + 256;
  const int stages = static_cast<int>(std::min<int64_t>(MAX_STAGES, (200 * 1024 - fixed) / stage_bytes));
  if (stages < 3) return 1;
  int dev = 0, sms = 148;
  NXF_CUDA(cudaGetDevice(&dev));
  static int sm_cache[32] = {0};
  if (dev >= 0 && dev < 32) {
    if (sm_cache[dev] == 0) NXF_CUDA(cudaDeviceGetAttribute(&sm_cache[dev], cudaDevAttrMultiProcessorCount, dev));
    sms = sm_cache[dev];
  }
  // segments: at least ~6 per CTA for the dynamic schedule to balance ragged masks, each at least two stages long
  int splits = 1;
  while (B * splits < 6 * static_cast<int64_t>(sms) && L / (2 * splits) >= 2 * CH && splits < 1024) splits *= 2;
  const size_t part_bytes = splits > 1 ? static_cast<size_t>(B) * splits * H * 4 : 0;
  unsigned char* ws = nullptr;
  int rc = pool_workspace(part_bytes, stream, &ws);
  if (rc) return rc;
  unsigned long long* den = reinterpret_cast<unsigned long long*>(ws);
  unsigned int* counter = reinterpret_cast<unsigned int*>(ws + size_t(65536) * 8);
  unsigned int* sched = reinterpret_cast<unsigned int*>(den + 65535);   // row 65 535 never exists: its den word holds the two scheduling counters
  float* partial = reinterpret_cast<float*>(ws + kPoolCtlBytes);
  const size_t smem = static_cast<size_t>(stage_bytes) * stages + fixed;
  auto kern = pool_mean_stream_kernel<T, O, VEC>;
  if (int r2 = ensure_dyn_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem))) return r2;
  const int grid = static_cast<int>(std::min<int64_t>(sms, B * splits));
  kern<<<grid, THREADS, smem, stream>>>((const T*)hidden, mask, B, L, H, normalize, eps, (O*)out, splits, CH, stages, G, partial, den, counter,
                                       sched, sched + 1);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

template <typename T, typename O, int VEC>
static int launch_pool_t(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize, float eps,
                         void* out, cudaStream_t stream) {
  const int nvec = static_cast<int>(H / VEC);
  // mean over vectorisable rows: the split kernel (any B fills the GPU); B <= 65535 is the grid's y limit
  if (pooling == NXF_POOL_MEAN && VEC > 1 && nvec <= 512 && B <= 65535 && static_cast<size_t>(std::max(1, 512 / nvec)) * H * 4 <= 160 * 1024)
  {
    if constexpr (VEC > 1) {
      // the persistent TMA stream first (NXF_POOL_STREAM=0: the split kernel, for A/B timing); it needs the cached workspace
      const char* se = getenv("NXF_POOL_STREAM");
      const char* we = getenv("NXF_POOL_WS_CACHE");
      const bool ws_cached = we ? atoi(we) != 0 : NXF_POOL_WS_CACHE_DEFAULT != 0;
      // Measured (B200, same box, fraction of the HBM peak, stream | split): 3 000 x 128 x 384 f16 0.54 | 0.41, 1 024 x 256 x 1 024
      // bf16 0.82 | -, 64 x 8 192 x 1 536 bf16 0.87 | 0.82, 256 x 512 x 768 f16 0.39 | 0.45.  The stream wins when a CTA owns whole
      // rows (B >= 6 rows per SM: no cross-CTA combine) or rows are long (>= 1 MB: the per-segment combine is rare); in between the
      // end-of-segment round trips (fence + arrival counter) on the consumers' path cost more than the split kernel's short CTAs.
      // (A variant with dedicated finalizer warps was measured too: 0.25 / 0.52 / 0.67 -- slower; profiles/r2_k0_stream_variants.txt.)
      const bool whole_rows = B >= 6 * 148;
      const bool long_rows = L * H * static_cast<int64_t>(sizeof(T)) >= (int64_t(1) << 20);
      const bool force = se && *se == '1';
      if (!(se && *se == '0') && ws_cached && (whole_rows || long_rows || force)) {
        const int rc = launch_pool_mean_stream<T, O, VEC>(hidden, mask, B, L, H, normalize, eps, out, stream);
        if (rc <= 0) return rc;
      }
    }
    return launch_pool_mean_split<T, O, VEC>(hidden, mask, B, L, H, normalize, eps, out, stream);
  }
  int groups = 1;
  if (pooling == NXF_POOL_MEAN) {
    groups = std::max(1, std::min(k0::MAX_GROUPS, k0::THREADS / std::max(1, std::min(nvec, k0::THREADS))));
    while (k0::THREADS % groups) --groups;
    while (groups > 1 && static_cast<size_t>(groups) * H * 4 > 160 * 1024) groups >>= 1;
  }
  const size_t smem = static_cast<size_t>(groups) * H * 4;
  if (smem > 200 * 1024) { set_error("pool_normalize: hidden size %lld too large", (long long)H); return NXF_EUNSUPPORTED; }
  auto kern = pool_norm_kernel<T, O, VEC>;
  if (smem > 48 * 1024) { if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem))) return rc; }
  kern<<<static_cast<unsigned>(B), k0::THREADS, smem, stream>>>((const T*)hidden, mask, B, L, H, pooling, normalize, eps, (O*)out, groups);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

template <typename T, int VEC>
static int launch_pool_o(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize, float eps,
                         void* out, int out_dtype, cudaStream_t stream) {
  switch (out_dtype) {
    case NXF_F32: return launch_pool_t<T, float, VEC>(hidden, mask, B, L, H, pooling, normalize, eps, out, stream);
    case NXF_F16: return launch_pool_t<T, __half, VEC>(hidden, mask, B, L, H, pooling, normalize, eps, out, stream);
    case NXF_BF16: return launch_pool_t<T, __nv_bfloat16, VEC>(hidden, mask, B, L, H, pooling, normalize, eps, out, stream);
  }
  set_error("pool_normalize: bad output dtype %d", out_dtype);
  return NXF_EINVAL;
}

int launch_pool(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize,
                float eps, void* out, int out_dtype, cudaStream_t stream) {
  if (B <= 0) return NXF_OK;
  const bool aligned = (reinterpret_cast<uintptr_t>(hidden) % 16 == 0);
  switch (h_dtype) {
    case NXF_F32:
      if (aligned && H % 4 == 0) return launch_pool_o<float, 4>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
      return launch_pool_o<float, 1>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
    case NXF_F16:
      if (aligned && H % 8 == 0) return launch_pool_o<__half, 8>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
      return launch_pool_o<__half, 1>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
    case NXF_BF16:
      if (aligned && H % 8 == 0) return launch_pool_o<__nv_bfloat16, 8>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
      return launch_pool_o<__nv_bfloat16, 1>(hidden, mask, B, L, H, pooling, normalize, eps, out, out_dtype, stream);
  }
  set_error("pool_normalize: bad hidden dtype %d", h_dtype);
  return NXF_EINVAL;
}

// ---- mean pooling without any allocation: nxf_pool_mean_normalize_static ---------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the split
// mean kernel above with its workspace (per-row control words + fp32 partial sums) in a file-scope __device__ array instead of the
// cached cudaMalloc'd workspace.  Every GPU that loads the library holds the array (kStaticScratchBytes = 64 KB + 8 MB) until the
// process exits.  In HOST code the array's name is a host-side stand-in, not a device address: the address the kernel gets comes
// from cudaGetSymbolAddress, taken once per device with that device current.  All calls share the array and it keeps the last
// call's contents: the control words are cleared on the call's stream before every launch; calls that overlap on the GPU race.
constexpr size_t kStaticCtlBytes = 64 * 1024;               // den (u64) + arrival counter (u32) for up to kStaticMaxRows rows
constexpr size_t kStaticPartBytes = size_t(8) << 20;        // fp32 partial sums [B][splits][H]
constexpr size_t kStaticScratchBytes = kStaticCtlBytes + kStaticPartBytes;
constexpr int64_t kStaticMaxRows = 4096;                    // 4096 * 8 + 4096 * 4 = 48 KB of control words
static __device__ __align__(256) unsigned char g_pool_static_scratch[kStaticScratchBytes];

namespace {
std::mutex g_static_mu;
unsigned char* g_static_ptr[64] = {nullptr};   // device address of the array, per device ordinal

// the device must be current
int static_scratch_address(int device, unsigned char** out) {
  if (device < 0 || device >= 64) { set_error("nxf_pool_mean_normalize_static: device ordinal %d outside [0, 64)", device); return NXF_EINVAL; }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (g_static_ptr[device] == nullptr) {
    void* ptr = nullptr;
    NXF_CUDA(cudaGetSymbolAddress(&ptr, g_pool_static_scratch));
    g_static_ptr[device] = static_cast<unsigned char*>(ptr);
  }
  *out = g_static_ptr[device];
  return NXF_OK;
}

template <typename T, typename O, int VEC>
int launch_pool_mean_static_t(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int normalize, float eps, void* out,
                              unsigned char* scratch, cudaStream_t stream) {
  const int nvec = static_cast<int>(H / VEC);
  if (nvec > 512 || static_cast<size_t>(std::max(1, 512 / nvec)) * H * 4 > 160 * 1024) {
    set_error("nxf_pool_mean_normalize_static: hidden size %lld does not fit the split kernel (at most 512 16-byte vectors per row)", (long long)H);
    return NXF_EUNSUPPORTED;
  }
  const int G = std::max(1, 512 / nvec);
  const int threads = std::min(512, (G * nvec + 31) / 32 * 32);
  // as many slices of L per row as the split kernel would take, halved until the partial sums fit the static array
  int splits = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>((NXF_POOL_CTAS + B - 1) / B, std::max<int64_t>(1, L / 64))));
  while (splits > 1 && static_cast<size_t>(B) * splits * H * 4 > kStaticPartBytes) splits = (splits + 1) / 2;
  if (static_cast<size_t>(B) * splits * H * 4 > kStaticPartBytes) {
    set_error("nxf_pool_mean_normalize_static: B * H * 4 = %lld bytes of partial sums exceed the static scratch (%lld bytes)",
              (long long)(B * H * 4), (long long)kStaticPartBytes);
    return NXF_EINVAL;
  }
  const size_t smem = static_cast<size_t>(G) * H * 4;
  auto kern = pool_mean_split_kernel<T, O, VEC>;
  if (smem > 48 * 1024) { if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem))) return rc; }
  unsigned long long* den = reinterpret_cast<unsigned long long*>(scratch);
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + kStaticMaxRows * 8);
  float* partial = reinterpret_cast<float*>(scratch + kStaticCtlBytes);
  // whatever an earlier (possibly aborted) call left behind: the control words must start at zero
  NXF_CUDA(cudaMemsetAsync(scratch, 0, static_cast<size_t>(kStaticMaxRows) * 12, stream));
  kern<<<dim3(static_cast<unsigned>(splits), static_cast<unsigned>(B)), threads, smem, stream>>>(
      (const T*)hidden, mask, L, H, normalize, eps, (O*)out, G, partial, den, counter);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

template <typename T, int VEC>
int launch_pool_mean_static_o(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int normalize, float eps, void* out,
                              int out_dtype, unsigned char* scratch, cudaStream_t stream) {
  switch (out_dtype) {
    case NXF_F32: return launch_pool_mean_static_t<T, float, VEC>(hidden, mask, B, L, H, normalize, eps, out, scratch, stream);
    case NXF_F16: return launch_pool_mean_static_t<T, __half, VEC>(hidden, mask, B, L, H, normalize, eps, out, scratch, stream);
    case NXF_BF16: return launch_pool_mean_static_t<T, __nv_bfloat16, VEC>(hidden, mask, B, L, H, normalize, eps, out, scratch, stream);
  }
  set_error("nxf_pool_mean_normalize_static: bad output dtype %d", out_dtype);
  return NXF_EINVAL;
}
}  // namespace

}  // namespace nxf

// C ABI (declared in include/nexus_flat.h).  Defined here, not in api.cu: only this translation unit can name the static array.
extern "C" int nxf_pool_mean_normalize_static(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L, int64_t H,
                                              int normalize, float eps, void* out, int out_dtype, int device, void* stream_) {
  using namespace nxf;
  NXF_REQUIRE(hidden != nullptr && out != nullptr && mask != nullptr, "nxf_pool_mean_normalize_static: NULL buffer");
  NXF_REQUIRE(B >= 0 && L >= 1 && H >= 1, "nxf_pool_mean_normalize_static: bad shape B=%lld L=%lld H=%lld", (long long)B, (long long)L, (long long)H);
  NXF_REQUIRE(B <= kStaticMaxRows, "nxf_pool_mean_normalize_static: B=%lld rows exceed the static scratch's maximum of %lld (use nxf_pool_normalize)",
              (long long)B, (long long)kStaticMaxRows);
  if (B == 0) return NXF_OK;
  const bool aligned = reinterpret_cast<uintptr_t>(hidden) % 16 == 0;
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__);
  unsigned char* scratch = nullptr;
  int rc = static_scratch_address(device, &scratch);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (rc == NXF_OK) {
    if (h_dtype == NXF_F32 && aligned && H % 4 == 0)
      rc = launch_pool_mean_static_o<float, 4>(hidden, mask, B, L, H, normalize, eps, out, out_dtype, scratch, stream);
    else if (h_dtype == NXF_F16 && aligned && H % 8 == 0)
      rc = launch_pool_mean_static_o<__half, 8>(hidden, mask, B, L, H, normalize, eps, out, out_dtype, scratch, stream);
    else if (h_dtype == NXF_BF16 && aligned && H % 8 == 0)
      rc = launch_pool_mean_static_o<__nv_bfloat16, 8>(hidden, mask, B, L, H, normalize, eps, out, out_dtype, scratch, stream);
    else {
      set_error("nxf_pool_mean_normalize_static: needs f32/f16/bf16 hidden states, 16-byte aligned, H a multiple of 16 bytes (dtype %d, H %lld)",
                h_dtype, (long long)H);
      rc = NXF_EUNSUPPORTED;
    }
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return rc;
}
