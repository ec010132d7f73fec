// block_sort.cuh -- block-level (256 threads) selection and ordering helpers shared by merge.cu and search_subset.cu:
// radix select of the k-th largest monotone key, in-place bitonic sort of 64-bit keys in shared memory, 16-byte row loads as fp32.
#pragma once
#include "params.cuh"

namespace nxf {

namespace mg {
constexpr int THREADS = 256;
constexpr int MAXK = NXF_MAX_K;  // 2048
}  // namespace mg

// ---- block-level helpers -------------------------------------------------------------------

// Finds the exact key of the `need`-th largest of n keys produced by key_at(i).  Returns it in
// *kth_out and the number of entries equal to kth that must be kept in *need_out (shared memory).
template <typename KeyAt>
__device__ __forceinline__ void block_radix_select(int n, int need_in, KeyAt key_at, uint32_t* hist, uint32_t* sh_prefix,
                                                   int* sh_need) {
  const int tid = threadIdx.x;
  if (tid == 0) { *sh_prefix = 0; *sh_need = need_in; }
  uint32_t mask = 0;
  __syncthreads();
#pragma unroll 1
  for (int pass = 3; pass >= 0; --pass) {
    const int shift = pass * 8;
    hist[tid] = 0;  // THREADS == 256
    __syncthreads();
    const uint32_t prefix = *sh_prefix;
    const int need = *sh_need;
    // (a warp-aggregated histogram -- __match_any_sync, one atomic per distinct bin per warp -- was tried for the passes in which
    // nearly every key hits the same bin and LOST: 0.16 -> 0.27 ms on a 29 600-entry slice; match.any costs more than the
    // same-address serialisation it removes.  Large slices are avoided at the source instead: sampled thresholds, api.cu)
    for (int i = tid; i < n; i += mg::THREADS) {
      const uint32_t key = key_at(i);
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 0xFF], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      const unsigned lane = tid;
      uint32_t c[8], lane_sum = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { c[j] = hist[255 - (lane * 8 + j)]; lane_sum += c[j]; }
      uint32_t incl = lane_sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(kFullMask, incl, o);
        if (lane >= static_cast<unsigned>(o)) incl += t;
      }
      const uint32_t excl = incl - lane_sum;
      if (excl < static_cast<uint32_t>(need) && static_cast<uint32_t>(need) <= incl) {
        uint32_t run = excl;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (run < static_cast<uint32_t>(need) && static_cast<uint32_t>(need) <= run + c[j]) {
            *sh_prefix = prefix | ((255u - (lane * 8 + j)) << shift);
            *sh_need = need - static_cast<int>(run);
          }
          run += c[j];
        }
      }
    }
    mask |= 0xFFu << shift;
    __syncthreads();
  }
}

// In-place bitonic sort, descending, of P (power of two) 64-bit keys in shared memory, with an
// optional parallel payload array.
template <bool PAYLOAD>
__device__ __forceinline__ void block_bitonic_desc(uint64_t* keys, int64_t* payload, int P) {
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int t = threadIdx.x; t < (P >> 1); t += mg::THREADS) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const uint64_t a = keys[lo], b = keys[hi];
        if ((a < b) == desc) {
          keys[lo] = b; keys[hi] = a;
          if (PAYLOAD) { const int64_t x = payload[lo]; payload[lo] = payload[hi]; payload[hi] = x; }
        }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ int next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// 8 consecutive stored elements (16-byte aligned) as fp32
template <typename T>
__device__ __forceinline__ void load8_f32(const T* __restrict__ p, float (&out)[8]);
template <>
__device__ __forceinline__ void load8_f32<float>(const float* __restrict__ p, float (&out)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w; out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
}
template <>
__device__ __forceinline__ void load8_f32<__nv_bfloat16>(const __nv_bfloat16* __restrict__ p, float (&out)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    out[2 * i] = __uint_as_float(w[i] << 16);
    out[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
template <>
__device__ __forceinline__ void load8_f32<__half>(const __half* __restrict__ p, float (&out)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(h[i]);
    out[2 * i] = f.x;
    out[2 * i + 1] = f.y;
  }
}

}  // namespace nxf
