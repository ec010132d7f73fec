// search_stream.cu -- K2: small-batch (nq <= 4) exact search, bound by HBM bandwidth.
//
// This is the shape in which the reference actually calls the index today: one query per request in
// serve/inference_service_text_retrieval.py:70, and (for fp32 indices) the tail chunks of
// utils.py:221-224.  Every query needs one pass over the whole corpus, so the kernel is judged on
// achieved GB/s: algorithmic bytes = N*ld*sizeof(T) per call.
//
// Structure (persistent, one CTA per SM):
//   warp 8      producer: cp.async.bulk (TMA 1-D) of R contiguous corpus rows per stage into a
//               shared-memory ring, completion on an mbarrier (full[s]); waits empty[s] before reuse.
//   warps 0..7  consumers: warp w owns rows [w*RW, (w+1)*RW) of the stage; lanes split each row in
//               16-byte chunks (conflict-free LDS.128), fp32 FMA against the queries held in shared
//               memory, a transposing butterfly reduce leaves one (row, query) score per lane pair,
//               then the streaming top-k sink (threshold filter -> append -> warp prune, select.cuh).
#include "params.cuh"

namespace nxf {

namespace k2 {
constexpr int CONSUMER_WARPS = 8;
constexpr int THREADS = (CONSUMER_WARPS + 1) * 32;
constexpr int MAX_STAGES = 4;
}  // namespace k2

template <typename T> struct Chunk;  // 16 bytes of storage -> E floats
template <> struct Chunk<float> {
  static constexpr int E = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&x)[4]) {
    x[0] = __uint_as_float(v.x); x[1] = __uint_as_float(v.y); x[2] = __uint_as_float(v.z); x[3] = __uint_as_float(v.w);
  }
};
template <> struct Chunk<__nv_bfloat16> {
  static constexpr int E = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&x)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      x[2 * i] = __uint_as_float(w[i] << 16);
      x[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
};
template <> struct Chunk<__half> {
  static constexpr int E = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&x)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      x[2 * i] = f.x; x[2 * i + 1] = f.y;
    }
  }
};

// Butterfly that halves the number of live values at each step while there is more than one:
// afterwards every lane holds the full 32-lane sum of value index `base` (returned).
template <int V>
__device__ __forceinline__ int warp_reduce_transpose(float (&v)[V], unsigned lane) {
  int base = 0;
  int live = V;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    if (live > 1) {
      const int half = live >> 1;
      const bool upper = (lane & o) != 0;
#pragma unroll
      for (int i = 0; i < V / 2; ++i) {
        if (i < half) {
          const float send = upper ? v[i] : v[i + half];
          const float keep = upper ? v[i + half] : v[i];
          v[i] = keep + __shfl_xor_sync(kFullMask, send, o);
        }
      }
      base += upper ? half : 0;
      live = half;
    } else {
      v[0] += __shfl_xor_sync(kFullMask, v[0], o);
    }
  }
  return base;
}
template <>
__device__ __forceinline__ int warp_reduce_transpose<1>(float (&v)[1], unsigned) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v[0] += __shfl_xor_sync(kFullMask, v[0], o);
  return 0;
}

template <typename T, int NQ, int RW>
__global__ void __launch_bounds__(k2::THREADS, 1) k2_stream_kernel(K2Params p) {
  using namespace k2;
  constexpr int E = Chunk<T>::E;
  constexpr int R = CONSUMER_WARPS * RW;  // rows per stage
  constexpr int V = RW * NQ;              // scores per warp per stage
  static_assert(V <= 16, "too many live scores per warp");
  extern __shared__ __align__(128) unsigned char smem[];

  const int64_t row_bytes = p.ld * static_cast<int64_t>(sizeof(T));
  const int64_t stage_bytes = row_bytes * R;
  unsigned char* stage_base = smem;
  float* q_s = reinterpret_cast<float*>(smem + stage_bytes * p.stages);                 // [NQ][ld]
  uint64_t* full = reinterpret_cast<uint64_t*>(q_s + static_cast<int64_t>(NQ) * p.ld);  // [stages]
  uint64_t* empty = full + MAX_STAGES;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const unsigned lane = tid & 31;

  for (int64_t i = tid; i < static_cast<int64_t>(NQ) * p.ld; i += THREADS) {
    const int qi = static_cast<int>(i / p.ld);
    const int d = static_cast<int>(i - qi * p.ld);
    q_s[i] = (qi < p.nq && d < p.dim) ? p.q[qi * p.ldq + d] : 0.f;
  }
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CONSUMER_WARPS); }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == CONSUMER_WARPS) {
    // ------------------------------- producer -------------------------------
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      const unsigned char* src = reinterpret_cast<const unsigned char*>(p.corpus);
      for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
        mbar_wait(&empty[s], ph ^ 1);
        const int64_t r0 = b * R;
        const int64_t rows = (p.n - r0 < R) ? (p.n - r0) : static_cast<int64_t>(R);
        const uint32_t bytes = static_cast<uint32_t>(rows * row_bytes);
        mbar_expect_tx(&full[s], bytes);
        bulk_load_1d(stage_base + s * stage_bytes, src + r0 * row_bytes, bytes, &full[s]);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
    return;
  }

  // --------------------------------- consumers ---------------------------------
  const int cap = p.sink.cap;
  const int k = p.sink.k;
  Entry* my_cand = p.sink.cand + (static_cast<int64_t>(blockIdx.x) * CONSUMER_WARPS + warp) * NQ * cap;
  float tau[NQ];
  int cnt[NQ];
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) { cnt[qi] = 0; tau[qi] = (qi < p.nq) ? tau_load(p.sink.tau_g, qi) : INFINITY; }

  // which (row, query) score this lane holds after the reduce: fixed by the lane id
  int my_base;
  {
    float dummy[V];
#pragma unroll
    for (int i = 0; i < V; ++i) dummy[i] = 0.f;
    my_base = warp_reduce_transpose<V>(dummy, lane);
  }
  const int my_r = my_base / NQ;
  const int my_q = my_base - my_r * NQ;
  const bool owner = (lane & (32 / V - 1)) == 0;  // lanes sharing a score: the lowest one appends

  const int chunks = static_cast<int>(row_bytes >> 4);
  int s = 0;
  uint32_t ph = 0;
  for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
    mbar_wait(&full[s], ph);
    const unsigned char* st = stage_base + s * stage_bytes + static_cast<int64_t>(warp) * RW * row_bytes;
    float acc[V];
#pragma unroll
    for (int i = 0; i < V; ++i) acc[i] = 0.f;
#pragma unroll 2
    for (int ci = lane; ci < chunks; ci += 32) {
      uint4 cv[RW];
#pragma unroll
      for (int r = 0; r < RW; ++r) cv[r] = *reinterpret_cast<const uint4*>(st + r * row_bytes + (static_cast<int64_t>(ci) << 4));
      float qv[NQ][E];
#pragma unroll
      for (int qi = 0; qi < NQ; ++qi) {
        const float4* qp = reinterpret_cast<const float4*>(q_s + qi * p.ld + ci * E);
#pragma unroll
        for (int e4 = 0; e4 < E / 4; ++e4) {
          const float4 f = qp[e4];
          qv[qi][4 * e4] = f.x; qv[qi][4 * e4 + 1] = f.y; qv[qi][4 * e4 + 2] = f.z; qv[qi][4 * e4 + 3] = f.w;
        }
      }
#pragma unroll
      for (int r = 0; r < RW; ++r) {
        float x[E];
        Chunk<T>::unpack(cv[r], x);
#pragma unroll
        for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
          for (int e = 0; e < E; ++e) acc[r * NQ + qi] = fmaf(x[e], qv[qi][e], acc[r * NQ + qi]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);  // stage consumed (all lanes' reads are done)
    if (++s == p.stages) { s = 0; ph ^= 1; }

    warp_reduce_transpose<V>(acc, lane);
    float sc = acc[0];
    const int64_t grow = b * R + warp * RW + my_r;
    float my_tau = tau[0];
#pragma unroll
    for (int qi = 1; qi < NQ; ++qi) if (my_q == qi) my_tau = tau[qi];
    bool pass = owner && grow < p.n;
    if (p.half_norm != nullptr && pass) sc -= p.half_norm[grow];
    pass = pass && sc > my_tau;
    if (__any_sync(kFullMask, pass)) {
#pragma unroll
      for (int qi = 0; qi < NQ; ++qi) {
        const bool mine = pass && my_q == qi;
        const unsigned m = __ballot_sync(kFullMask, mine);
        if (m) {
          Entry* buf = my_cand + static_cast<int64_t>(qi) * cap;
          if (mine) {
            Entry e; e.score = sc; e.row = static_cast<int32_t>(grow);
            NXF_DEV_CHECK(cnt[qi] + __popc(m) <= cap, "K2 per-(warp, query) buffer overflow");
            buf[cnt[qi] + __popc(m & lanemask_lt())] = e;
          }
          cnt[qi] += __popc(m);
          __syncwarp();
          if (cnt[qi] > cap - RW) {  // the next stage could add RW more
            const uint32_t kth = warp_prune(buf, cnt[qi], k);
            cnt[qi] = k;
            if (lane == 0) atomicMax(&p.sink.tau_g[qi], kth);
            tau[qi] = fmaxf(tau[qi], okey_inv(kth));
          }
        }
      }
    }
    // pick up other warps' / CTAs' progress once per stage
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
      if (qi < p.nq) tau[qi] = fmaxf(tau[qi], tau_load(p.sink.tau_g, qi));
  }

  // ---- end: prune to k and flush ----
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    if (qi >= p.nq) continue;
    Entry* buf = my_cand + static_cast<int64_t>(qi) * cap;
    int n = cnt[qi];
    if (n > k) { warp_prune(buf, n, k); n = k; }
    __syncwarp();
    warp_flush(p.sink, buf, n, qi);
  }
}

size_t k2_smem_bytes(int64_t ld, int elem_size, int nq_t, int rw, int stages) {
  const int64_t stage_bytes = ld * elem_size * k2::CONSUMER_WARPS * rw;
  return static_cast<size_t>(stage_bytes * stages + static_cast<int64_t>(nq_t) * ld * 4 + 2 * k2::MAX_STAGES * 8);
}

template <typename T, int NQ, int RW>
static int launch_k2_t(const K2Params& p, int grid, size_t smem, cudaStream_t stream) {
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(k2_stream_kernel<T, NQ, RW>), static_cast<int>(smem))) return rc;
  k2_stream_kernel<T, NQ, RW><<<grid, k2::THREADS, smem, stream>>>(p);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}
template <typename T>
static int launch_k2_d(const K2Params& p, int nq_t, int rw, int grid, size_t smem, cudaStream_t stream) {
#define NXF_K2_CASE(NQ, RW) if (nq_t == NQ && rw == RW) return launch_k2_t<T, NQ, RW>(p, grid, smem, stream)
  NXF_K2_CASE(1, 4); NXF_K2_CASE(1, 2); NXF_K2_CASE(1, 1);
  NXF_K2_CASE(2, 4); NXF_K2_CASE(2, 2); NXF_K2_CASE(2, 1);
  NXF_K2_CASE(4, 4); NXF_K2_CASE(4, 2); NXF_K2_CASE(4, 1);
#undef NXF_K2_CASE
  set_error("launch_k2: unsupported (nq_t=%d, rw=%d)", nq_t, rw);
  return NXF_EINVAL;
}

int launch_k2(int store_dtype, const K2Params& p, int nq_t, int rw, int grid, size_t smem, cudaStream_t stream) {
  switch (store_dtype) {
    case NXF_F32: return launch_k2_d<float>(p, nq_t, rw, grid, smem, stream);
    case NXF_F16: return launch_k2_d<__half>(p, nq_t, rw, grid, smem, stream);
    case NXF_BF16: return launch_k2_d<__nv_bfloat16>(p, nq_t, rw, grid, smem, stream);
  }
  set_error("launch_k2: bad storage dtype %d", store_dtype);
  return NXF_EINVAL;
}

}  // namespace nxf
