// search_exact.cu -- KX: exact fp32 flat search on CUDA cores (FFMA), fused streaming top-k.
//
// Same arithmetic as the reference's own implementation, NumpyIPIndex.search
// (Nexus/evaluation/text_retrieval/utils.py:20-31: fp32 cast, q @ emb.T, top-k), for any storage type:
// stored elements are widened to fp32 and every product/accumulation is an fp32 FMA.  It serves
// fp32-stored indices (BASELINE config 1) and is the on-device ground truth for the tensor-core path.
//
// Work decomposition: unit = (corpus chunk c, query tile qt), qt fastest so that co-running CTAs
// read the same corpus rows (L2 reuse).  A CTA owns 128 query rows for the whole unit: per-row
// threshold/count live in shared memory, candidates go to a CTA-private global buffer, and a row is
// pruned by one warp (bitwise-descent select, select.cuh) when its buffer could overflow during the next tile.
#include "params.cuh"

namespace nxf {

namespace kx {
constexpr int BM = 128;  // queries per CTA tile
constexpr int BN = 128;  // corpus rows per tile
constexpr int BK = 16;
constexpr int THREADS = 256;
constexpr int PAD = 4;
}  // namespace kx

template <typename T>
__global__ void __launch_bounds__(kx::THREADS) kx_search_kernel(KxParams p) {
  using namespace kx;
  __shared__ __align__(16) float As[BK][BM + PAD];
  __shared__ __align__(16) float Bs[BK][BN + PAD];
  __shared__ float tau_s[BM];
  __shared__ int cnt_s[BM];
  __shared__ int any_prune_s;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int tx = tid & 15;  // column group
  const int ty = tid >> 4;  // row group
  const T* __restrict__ corpus = reinterpret_cast<const T*>(p.corpus);
  const int cap = p.sink.cap;
  const int k = p.sink.k;
  Entry* my_cand = p.sink.cand + static_cast<int64_t>(blockIdx.x) * BM * cap;

  const int64_t num_units = static_cast<int64_t>(p.num_qt) * p.chunks;
  for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x) {
    const int qt = static_cast<int>(u % p.num_qt);
    const int64_t c = u / p.num_qt;
    const int64_t t0 = c * p.tiles_per_chunk;
    const int64_t t1 = min(t0 + p.tiles_per_chunk, p.n_tiles);
    const int64_t q0 = static_cast<int64_t>(qt) * BM;

    if (tid < BM) {
      cnt_s[tid] = 0;
      tau_s[tid] = (q0 + tid < p.nq) ? tau_load(p.sink.tau_g, q0 + tid) : INFINITY;  // padded rows accept nothing
    }
    if (tid == 0) any_prune_s = 0;
    __syncthreads();

    for (int64_t t = t0; t < t1; ++t) {
      const int64_t r0 = t * BN;
      float acc[8][8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

      for (int kb = 0; kb < p.dim; kb += BK) {
        // ---- stage A (queries) and B (corpus) slices, transposed to [k][row] ----
        {
          const int row = tid >> 1;        // 0..127
          const int kk0 = (tid & 1) * 8;   // 0 or 8
          const int64_t gq = q0 + row;
          const int64_t gr = r0 + row;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int d = kb + kk0 + j;
            float a = 0.f, b = 0.f;
            if (d < p.dim) {
              if (gq < p.nq) a = p.q[gq * p.ldq + d];
              if (gr < p.n) b = to_f32<T>(corpus[gr * p.ld + d]);
            }
            As[kk0 + j][row] = a;
            Bs[kk0 + j][row] = b;
          }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
          const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
          const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
          const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
          const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
          const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
          const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
      }

      // ---- epilogue: threshold filter + append ----
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int row = (i < 4) ? (ty * 4 + i) : (64 + ty * 4 + (i - 4));
        const float tau = tau_s[row];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int col = (j < 4) ? (tx * 4 + j) : (64 + tx * 4 + (j - 4));
          const int64_t gr = r0 + col;
          float s = acc[i][j];
          if (p.half_norm != nullptr && gr < p.n) s -= p.half_norm[gr];
          if (gr < p.n && s > tau) {
            const int pos = atomicAdd(&cnt_s[row], 1);
            Entry e; e.score = s; e.row = static_cast<int32_t>(gr);
            NXF_DEV_CHECK(pos < cap, "KX per-row candidate buffer overflow");
            my_cand[static_cast<int64_t>(row) * cap + pos] = e;
          }
        }
      }
      __syncthreads();
      // ---- prune rows that could overflow during the next tile (a tile adds <= BN per row) ----
      if (tid < BM && cnt_s[tid] > cap - BN) any_prune_s = 1;
      __syncthreads();
      if (any_prune_s) {
        for (int row = warp; row < BM; row += THREADS / 32) {
          const int n = cnt_s[row];
          if (n > cap - BN && n > k) {
            const uint32_t kth = warp_prune(my_cand + static_cast<int64_t>(row) * cap, n, k);
            if (lane == 0) {
              cnt_s[row] = k;
              tau_s[row] = fmaxf(tau_s[row], okey_inv(kth));
              if (q0 + row < p.nq) atomicMax(&p.sink.tau_g[q0 + row], kth);
            }
          }
        }
        __syncthreads();
        if (tid == 0) any_prune_s = 0;
        // refresh thresholds from other CTAs' progress
        if (tid < BM && q0 + tid < p.nq) tau_s[tid] = fmaxf(tau_s[tid], tau_load(p.sink.tau_g, q0 + tid));
        __syncthreads();
      }
    }

    // ---- unit end: prune to k, publish threshold, flush to the arena ----
    for (int row = warp; row < BM; row += THREADS / 32) {
      int n = cnt_s[row];
      Entry* buf = my_cand + static_cast<int64_t>(row) * cap;
      if (q0 + row >= p.nq) continue;
      if (n > k) {
        const uint32_t kth = warp_prune(buf, n, k);
        if (lane == 0) atomicMax(&p.sink.tau_g[q0 + row], kth);
        n = k;
      }
      __syncwarp();
      warp_flush(p.sink, buf, n, q0 + row);
    }
    __syncthreads();
  }
}

int launch_kx(int store_dtype, const KxParams& p, int grid, cudaStream_t stream) {
  switch (store_dtype) {
    case NXF_F32: kx_search_kernel<float><<<grid, kx::THREADS, 0, stream>>>(p); break;
    case NXF_F16: kx_search_kernel<__half><<<grid, kx::THREADS, 0, stream>>>(p); break;
    case NXF_BF16: kx_search_kernel<__nv_bfloat16><<<grid, kx::THREADS, 0, stream>>>(p); break;
    default: set_error("launch_kx: bad storage dtype %d", store_dtype); return NXF_EINVAL;
  }
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

}  // namespace nxf
