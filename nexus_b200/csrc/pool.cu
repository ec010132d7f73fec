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

#include "common.cuh"

namespace nxf {

namespace k0 {
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

template <typename T, typename O, int VEC>
static int launch_pool_t(const void* hidden, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize, float eps,
                         void* out, cudaStream_t stream) {
  const int nvec = static_cast<int>(H / VEC);
  int groups = 1;
  if (pooling == NXF_POOL_MEAN) {
    groups = std::max(1, std::min(k0::MAX_GROUPS, k0::THREADS / std::max(1, std::min(nvec, k0::THREADS))));
    while (k0::THREADS % groups) --groups;
    while (groups > 1 && static_cast<size_t>(groups) * H * 4 > 160 * 1024) groups >>= 1;
  }
  const size_t smem = static_cast<size_t>(groups) * H * 4;
  if (smem > 200 * 1024) { set_error("pool_normalize: hidden size %lld too large", (long long)H); return NXF_EUNSUPPORTED; }
  auto kern = pool_norm_kernel<T, O, VEC>;
  if (smem > 48 * 1024) NXF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
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
