// select.cuh -- exact k-selection primitives shared by all search kernels.
//
//   warp_prune   : one warp reduces a candidate buffer in global memory (n > k entries) to its
//                  exact top-k (unsorted, compacted to the front) and returns a threshold T with
//                  exactly-k-or-ties semantics: T <= the k-th key, and every kept entry has key >= T.
//                  Bitwise descent on the monotone 32-bit score key with warp REDUX counting (no shared
//                  memory, no atomics), then one in-place compaction pass.
//   Sink         : per-owner streaming top-k state (threshold tau, count, buffer) used by the
//                  epilogues: "score > tau" filter, append, prune when the buffer is nearly full.
#pragma once

#include "common.cuh"

namespace nxf {

constexpr unsigned kFullMask = 0xFFFFFFFFu;

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m; asm volatile("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m;
}

// Exact selection by bitwise descent on the monotone 32-bit key: T = the largest value with count(key >= T) >= k is the
// k-th largest key.  Bits above the highest bit in which min and max differ are common to all keys and are skipped, so a
// buffer of scores from a narrow band (the usual case next to a threshold) needs ~20 counting steps.  No shared memory,
// no atomics: a counting step is R compares per lane and one warp REDUX.

// Register-resident variant for n <= 32*R: the buffer is read ONCE (R independent coalesced loads per lane, all in flight
// together).  All 32 lanes call with identical (buf, n, k); requires n > k >= 1.  On return buf[0..k) holds the k largest
// entries (ties at the k-th value resolved by buffer position) and the k-th key is returned.
template <int R>
__device__ __noinline__ uint32_t warp_prune_reg(Entry* __restrict__ buf, int n, int k) {
  __builtin_assume(__isGlobal(buf));
  const unsigned lane = lane_id();
  Entry e[R];
  uint32_t key[R];
  uint32_t lmin = 0xFFFFFFFFu, lmax = 0u;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = r * 32 + lane;
    e[r].score = 0.f; e[r].row = -1;
    key[r] = 0u;  // below every counted threshold (thresholds are >= 1)
    if (i < n) {
      e[r] = buf[i];
      key[r] = okey(e[r].score);
      lmin = min(lmin, key[r]);
      lmax = max(lmax, key[r]);
    }
  }
  lmin = __reduce_min_sync(kFullMask, lmin);
  lmax = __reduce_max_sync(kFullMask, lmax);
  uint32_t T = lmax;  // all keys equal: the k-th is that key
  const uint32_t diff = lmin ^ lmax;
  if (diff != 0u) {
    const int top = 31 - __clz(diff);
    T = (top == 31) ? 0u : (lmax & ~((2u << top) - 1u));  // the common prefix
#pragma unroll 1
    for (int b = top; b >= 0; --b) {
      const uint32_t cand = T | (1u << b);
      unsigned c = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) c += (key[r] >= cand) ? 1u : 0u;
      c = __reduce_add_sync(kFullMask, c);
      if (c >= static_cast<unsigned>(k)) {
        T = cand;
        if (c == static_cast<unsigned>(k)) break;  // exactly k keys >= cand: the top-k set is settled, cand is a lower bound of the k-th key
      }
    }
  }
  const uint32_t kth = T;
  unsigned gt = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) gt += (key[r] > kth) ? 1u : 0u;
  gt = __reduce_add_sync(kFullMask, gt);
  const int need = k - static_cast<int>(gt);  // entries equal to the k-th key that are kept
  int out = 0, ties = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool in = r * 32 + static_cast<int>(lane) < n;
    const bool eq = in && key[r] == kth;
    const unsigned eqm = __ballot_sync(kFullMask, eq);
    const bool keep = in && (key[r] > kth || (eq && ties + __popc(eqm & lanemask_lt()) < need));
    const unsigned km = __ballot_sync(kFullMask, keep);
    if (keep) buf[out + __popc(km & lanemask_lt())] = e[r];   // every slot was read into registers before any write
    out += __popc(km);
    ties += __popc(eqm);
  }
  __syncwarp();
  return kth;
}

// Key-resident variant for n <= 32*R with R too large to also hold the entries: only the 32-bit keys live in registers
// during the descent (one coalesced read of the buffer), the compaction pass re-reads the entries.  Same contract.
template <int R>
__device__ __noinline__ uint32_t warp_prune_keys(Entry* __restrict__ buf, int n, int k) {
  __builtin_assume(__isGlobal(buf));
  const unsigned lane = lane_id();
  uint32_t key[R];
  uint32_t lmin = 0xFFFFFFFFu, lmax = 0u;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = r * 32 + lane;
    key[r] = 0u;
    if (i < n) {
      key[r] = okey(buf[i].score);
      lmin = min(lmin, key[r]);
      lmax = max(lmax, key[r]);
    }
  }
  lmin = __reduce_min_sync(kFullMask, lmin);
  lmax = __reduce_max_sync(kFullMask, lmax);
  uint32_t T = lmax;
  const uint32_t diff = lmin ^ lmax;
  if (diff != 0u) {
    const int top = 31 - __clz(diff);
    T = (top == 31) ? 0u : (lmax & ~((2u << top) - 1u));
#pragma unroll 1
    for (int b = top; b >= 0; --b) {
      const uint32_t cand = T | (1u << b);
      unsigned c = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) c += (key[r] >= cand) ? 1u : 0u;
      c = __reduce_add_sync(kFullMask, c);
      if (c >= static_cast<unsigned>(k)) {
        T = cand;
        if (c == static_cast<unsigned>(k)) break;  // exactly k keys >= cand: the top-k set is settled, cand is a lower bound of the k-th key
      }
    }
  }
  const uint32_t kth = T;
  unsigned gt = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) gt += (key[r] > kth) ? 1u : 0u;
  gt = __reduce_add_sync(kFullMask, gt);
  const int need = k - static_cast<int>(gt);
  int out = 0, ties = 0;
#pragma unroll 4
  for (int r = 0; r < R; ++r) {
    const int i = r * 32 + static_cast<int>(lane);
    const bool in = i < n;
    Entry e;
    e.score = 0.f; e.row = -1;
    if (in) e = buf[i];
    const uint32_t kk = in ? okey(e.score) : 0u;
    const bool eq = in && kk == kth;
    const unsigned eqm = __ballot_sync(kFullMask, eq);
    const bool keep = in && (kk > kth || (eq && ties + __popc(eqm & lanemask_lt()) < need));
    const unsigned km = __ballot_sync(kFullMask, keep);
    const int pos = out + __popc(km & lanemask_lt());
    __syncwarp();  // every lane has read its slot of this round before anyone writes (pos <= i)
    if (keep) buf[pos] = e;
    out += __popc(km);
    ties += __popc(eqm);
    __syncwarp();
    if (r * 32 + 32 >= n) break;
  }
  return kth;
}

// Memory-resident variant for larger buffers: two bits per pass over the (L1/L2-resident) buffer, then one in-place
// compaction pass.  Same contract as above.
static __device__ __noinline__ uint32_t warp_prune_mem(Entry* __restrict__ buf, int n, int k) {
  __builtin_assume(__isGlobal(buf));
  const unsigned lane = lane_id();
  uint32_t lmin = 0xFFFFFFFFu, lmax = 0u;
  for (int i = lane; i < n; i += 32) {
    const uint32_t key = okey(buf[i].score);
    lmin = min(lmin, key);
    lmax = max(lmax, key);
  }
  lmin = __reduce_min_sync(kFullMask, lmin);
  lmax = __reduce_max_sync(kFullMask, lmax);
  uint32_t T = lmax;
  const uint32_t diff = lmin ^ lmax;
  if (diff != 0u) {
    const int top = 31 - __clz(diff);
    T = (top == 31) ? 0u : (lmax & ~((2u << top) - 1u));
#pragma unroll 1
    for (int b = top; b >= 0; b -= 2) {
      const uint32_t hi = 1u << b, lo = (b > 0) ? (1u << (b - 1)) : 0u;
      const uint32_t c3 = T | hi | lo, c2 = T | hi, c1 = T | lo;  // b == 0: c3 == c2, c1 == T
      unsigned n3 = 0, n2 = 0, n1 = 0;
      for (int i = lane; i < n; i += 32) {
        const uint32_t key = okey(buf[i].score);
        n3 += (key >= c3) ? 1u : 0u;
        n2 += (key >= c2) ? 1u : 0u;
        n1 += (key >= c1) ? 1u : 0u;
      }
      n3 = __reduce_add_sync(kFullMask, n3);
      n2 = __reduce_add_sync(kFullMask, n2);
      n1 = __reduce_add_sync(kFullMask, n1);
      const unsigned kk = static_cast<unsigned>(k);
      unsigned hit = n3;
      if (n3 >= kk) T = c3;
      else if (n2 >= kk) { T = c2; hit = n2; }
      else if (lo != 0u && n1 >= kk) { T = c1; hit = n1; }
      if (hit == kk) break;  // exactly k keys >= T
    }
  }
  const uint32_t kth = T;
  unsigned gt = 0;
  for (int i = lane; i < n; i += 32) gt += (okey(buf[i].score) > kth) ? 1u : 0u;
  gt = __reduce_add_sync(kFullMask, gt);
  const int need = k - static_cast<int>(gt);
  int out = 0, ties = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    Entry e;
    e.score = 0.f; e.row = -1;
    uint32_t key = 0;
    const bool in = i < n;
    if (in) { e = buf[i]; key = okey(e.score); }
    const bool eq = in && key == kth;
    const unsigned eqm = __ballot_sync(kFullMask, eq);
    const int my_tie = ties + __popc(eqm & lanemask_lt());
    const bool keep = in && (key > kth || (eq && my_tie < need));
    const unsigned km = __ballot_sync(kFullMask, keep);
    const int pos = out + __popc(km & lanemask_lt());
    __syncwarp();  // every lane has read its slot of this round before anyone writes (pos <= i)
    if (keep) buf[pos] = e;
    out += __popc(km);
    ties += __popc(eqm);
    __syncwarp();
  }
  return kth;
}

// Dispatcher.  All 32 lanes call with identical arguments; requires n > k >= 1.
__device__ __forceinline__ uint32_t warp_prune(Entry* __restrict__ buf, int n, int k) {
  if (n <= 32 * 8) return warp_prune_reg<8>(buf, n, k);
  if (n <= 32 * 16) return warp_prune_reg<16>(buf, n, k);
  if (n <= 32 * 32) return warp_prune_reg<32>(buf, n, k);
  if (n <= 32 * 48) return warp_prune_keys<48>(buf, n, k);
  if (n <= 32 * 96) return warp_prune_keys<96>(buf, n, k);
  return warp_prune_mem(buf, n, k);
}

}  // namespace nxf
