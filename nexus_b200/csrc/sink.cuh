// sink.cuh -- where the search kernels put the candidates that pass the running threshold.
//
// Global-memory scratch shared by K1 (tcgen05), K2 (stream) and KX (exact fp32):
//   tau_g [nq_pad]              uint32  monotone key of a valid lower bound on each query's final k-th
//                                       score (atomicMax by every owner after a prune; starts at 0)
//   qcnt  [nq_pad]              uint32  number of entries flushed to the query's arena slice
//   arena [nq_pad * stride]     Entry   per-query candidate slices, merged by merge_arena_kernel
//   cand  [owners * cap]        Entry   per-owner private append buffers (owner = epilogue thread in
//                                       K1, (CTA,row) in KX, (warp,query) in K2)
#pragma once

#include "select.cuh"

namespace nxf {

struct SinkParams {
  uint32_t* tau_g;
  uint32_t* qcnt;
  Entry* arena;
  Entry* cand;
  int64_t arena_stride;  // entries per query in the arena
  int cap;               // entries per owner buffer
  int k;                 // entries kept per owner (k or the over-fetch k')
};

// Warp-cooperative flush of one owner's buffer (n <= k entries, already pruned) to the query arena.
// All lanes call with identical arguments.
__device__ __forceinline__ void warp_flush(const SinkParams& sp, const Entry* __restrict__ buf, int n, int64_t query) {
  if (n <= 0) return;
  uint32_t pos = 0;
  if (lane_id() == 0) pos = atomicAdd(&sp.qcnt[query], static_cast<uint32_t>(n));
  pos = __shfl_sync(kFullMask, pos, 0);
  Entry* dst = sp.arena + query * sp.arena_stride + pos;
  NXF_DEV_CHECK(static_cast<int64_t>(pos) + n <= sp.arena_stride, "arena slice overflow in warp_flush");
  if (static_cast<int64_t>(pos) + n > sp.arena_stride) return;  // cannot happen by construction; never write OOB
  for (int i = lane_id(); i < n; i += 32) dst[i] = buf[i];
}

__device__ __forceinline__ float tau_from_key(uint32_t key) { return key <= kKeyNegInf ? -INFINITY : okey_inv(key); }
__device__ __forceinline__ float tau_load(const uint32_t* tau_g, int64_t query) {
  return tau_from_key(*reinterpret_cast<const volatile uint32_t*>(tau_g + query));
}
// The load alone: a warp stalls at the first instruction that USES a loaded register, and tau_load's key -> float conversion is
// that instruction -- a prefetched threshold must stay a raw key until the point of use (K1 on 64-d rows spent 5.5 % of its stall
// samples on the conversion, i.e. on an L2 round trip per tile that the tile's own work was meant to hide).
__device__ __forceinline__ uint32_t tau_load_key(const uint32_t* tau_g, int64_t query) {
  return *reinterpret_cast<const volatile uint32_t*>(tau_g + query);
}

}  // namespace nxf
