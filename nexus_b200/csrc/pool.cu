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

template <typename T, typename O, int VEC>
static int launch_pool_t(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize, float eps,
                         void* out, cudaStream_t stream) {
  const int nvec = static_cast<int>(H / VEC);
  // mean over vectorisable rows: the split kernel (any B fills the GPU); B <= 65535 is the grid's y limit
  if (pooling == NXF_POOL_MEAN && VEC > 1 && nvec <= 512 && B <= 65535 && static_cast<size_t>(std::max(1, 512 / nvec)) * H * 4 <= 160 * 1024)
    return launch_pool_mean_split<T, O, VEC>(hidden, mask, B, L, H, normalize, eps, out, stream);
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

}  // namespace nxf
