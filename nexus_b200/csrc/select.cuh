// select.cuh -- exact k-selection primitives shared by all search kernels.
//
//   warp_prune   : one warp reduces a candidate buffer in global memory (n > k entries) to its
//                  exact top-k (unsorted, compacted to the front) and returns the k-th key.
//                  Radix select over the monotone 32-bit score key: 4 passes of 8 bits with a
//                  256-bin shared-memory histogram, then one in-place compaction pass.
//   Sink         : per-owner streaming top-k state (threshold tau, count, buffer) used by the
//                  epilogues: "score > tau" filter, append, prune when the buffer is nearly full.
#pragma once

#include "common.cuh"

namespace nxf {

constexpr unsigned kFullMask = 0xFFFFFFFFu;

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m; asm volatile("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m;
}

// All 32 lanes call this with identical (buf, n, k, hist).  hist: 256 uint32 of shared memory
// private to the warp.  Requires n > k >= 1.  On return buf[0..k) holds the k largest entries by
// score key (ties at the k-th value resolved by buffer position) and the k-th key is returned.
static __device__ __noinline__ uint32_t warp_prune_global(Entry* __restrict__ buf, int n, int k, uint32_t* hist) {
  __builtin_assume(__isShared(hist));
  __builtin_assume(__isGlobal(buf));
  const unsigned lane = lane_id();
  uint32_t prefix = 0, mask = 0;
  int need = k;  // we want the need-th largest among entries matching (key & mask) == prefix
#pragma unroll 1
  for (int pass = 3; pass >= 0; --pass) {
    const int shift = pass * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) hist[lane * 8 + j] = 0;
    __syncwarp();
    for (int i = lane; i < n; i += 32) {
      uint32_t key = okey(buf[i].score);
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 0xFF], 1u);
    }
    __syncwarp();
    // lane l owns digits [255-8l-7, 255-8l]: descending order across lanes
    uint32_t c[8];
    uint32_t lane_sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = hist[255 - (lane * 8 + j)]; lane_sum += c[j]; }
    uint32_t incl = lane_sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(kFullMask, incl, o);
      if (lane >= static_cast<unsigned>(o)) incl += t;
    }
    const uint32_t excl = incl - lane_sum;
    const bool mine = (excl < static_cast<uint32_t>(need)) && (static_cast<uint32_t>(need) <= incl);
    uint32_t digit = 0, above = 0;
    if (mine) {
      uint32_t run = excl;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (run < static_cast<uint32_t>(need) && static_cast<uint32_t>(need) <= run + c[j]) {
          digit = 255 - (lane * 8 + j);
          above = run;
        }
        run += c[j];
      }
    }
    const unsigned who = __ballot_sync(kFullMask, mine);
    const int src = __ffs(who) - 1;  // exactly one lane when n >= need (guaranteed by n > k)
    digit = __shfl_sync(kFullMask, digit, src);
    above = __shfl_sync(kFullMask, above, src);
    need -= static_cast<int>(above);
    prefix |= digit << shift;
    mask |= 0xFFu << shift;
    __syncwarp();
  }
  const uint32_t kth = prefix;  // exact key of the k-th largest; keep `need` entries equal to it
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


// The digit scan shared by both variants: given the 256-bin histogram of the current digit among the
// entries matching the prefix, find the bin holding the need-th largest and the count above it.
__device__ __forceinline__ void warp_scan_digit(const uint32_t* hist, int need, unsigned lane, uint32_t* digit_out, uint32_t* above_out) {
  uint32_t c[8];
  uint32_t lane_sum = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { c[j] = hist[255 - (lane * 8 + j)]; lane_sum += c[j]; }
  uint32_t incl = lane_sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(kFullMask, incl, o);
    if (lane >= static_cast<unsigned>(o)) incl += t;
  }
  const uint32_t excl = incl - lane_sum;
  const bool mine = (excl < static_cast<uint32_t>(need)) && (static_cast<uint32_t>(need) <= incl);
  uint32_t digit = 0, above = 0;
  if (mine) {
    uint32_t run = excl;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (run < static_cast<uint32_t>(need) && static_cast<uint32_t>(need) <= run + c[j]) { digit = 255 - (lane * 8 + j); above = run; }
      run += c[j];
    }
  }
  const int src = __ffs(__ballot_sync(kFullMask, mine)) - 1;
  *digit_out = __shfl_sync(kFullMask, digit, src);
  *above_out = __shfl_sync(kFullMask, above, src);
}

// Register-resident variant for n <= 32*R: the buffer is read ONCE (R independent coalesced loads per
// lane, all in flight together), the four radix passes and the compaction run out of registers.
template <int R>
__device__ __noinline__ uint32_t warp_prune_reg(Entry* __restrict__ buf, int n, int k, uint32_t* hist) {
  __builtin_assume(__isShared(hist));
  __builtin_assume(__isGlobal(buf));
  const unsigned lane = lane_id();
  Entry e[R];
  uint32_t key[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = r * 32 + lane;
    e[r].score = 0.f; e[r].row = -1;
    if (i < n) e[r] = buf[i];
    key[r] = okey(e[r].score);
  }
  uint32_t prefix = 0, mask = 0;
  int need = k;
#pragma unroll 1
  for (int pass = 3; pass >= 0; --pass) {
    const int shift = pass * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) hist[lane * 8 + j] = 0;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (r * 32 + static_cast<int>(lane) < n && (key[r] & mask) == prefix) atomicAdd(&hist[(key[r] >> shift) & 0xFF], 1u);
    __syncwarp();
    uint32_t digit, above;
    warp_scan_digit(hist, need, lane, &digit, &above);
    need -= static_cast<int>(above);
    prefix |= digit << shift;
    mask |= 0xFFu << shift;
    __syncwarp();
  }
  const uint32_t kth = prefix;
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

// Dispatcher.  All 32 lanes call with identical arguments; requires n > k >= 1.
__device__ __forceinline__ uint32_t warp_prune(Entry* __restrict__ buf, int n, int k, uint32_t* hist) {
  if (n <= 32 * 8) return warp_prune_reg<8>(buf, n, k, hist);
  if (n <= 32 * 16) return warp_prune_reg<16>(buf, n, k, hist);
  if (n <= 32 * 32) return warp_prune_reg<32>(buf, n, k, hist);
  return warp_prune_global(buf, n, k, hist);
}

}  // namespace nxf
