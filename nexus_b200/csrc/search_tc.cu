// search_tc.cu -- K1: exact flat search as a tcgen05 GEMM with the top-k selection fused into the
// epilogue, so the [Q, N] score matrix never exists in HBM.
//
// Replaces, for f16/bf16-stored indices and nq > NXF_STREAM_MAX_NQ, the work behind
// faiss_index.search(...) at Nexus/evaluation/text_retrieval/utils.py:224 (and the sharded fp16 GPU
// variant at utils.py:178-183), modeling.py:197, recommendation/base.py:182 and the
// matmul + full sort at mmeb_v2/eval_embedding.py:359-361.
//
// Tile: 128 query rows per CTA (UMMA M = 128 * CG) x 256 corpus rows (UMMA N) x 64 (one 128-byte
// swizzle atom of K per pipeline stage), fp32 accumulators double-buffered in TMEM (2 x 256 columns).
// Warp roles (256 threads, one CTA per SM):
//   warp 0  TMA producer: queries tile (A) + corpus tile (B) per K-block into a smem ring
//   warp 1  MMA issuer (one lane; with CG = 2 only the pair's leader CTA issues, cta_group::2)
//   warp 2  TMEM allocator / deallocator
//   warps 4-7  epilogue: thread <-> accumulator lane <-> query row.  tcgen05.ld 32 columns at a time,
//           3-input max tree, compare with the row's running threshold tau; the common case is one
//           warp vote and no branch taken.  Survivors are appended to the thread's private candidate
//           buffer (global memory, L2 resident); a buffer that could overflow is reduced to its exact
//           top-k by the whole warp (radix select, select.cuh) and tau is raised to the k-th score,
//           also published to tau_g[query] so every CTA working on the same query benefits.
// Work decomposition: unit = (corpus chunk, query tile), query tile fastest: the CTAs running at any
// moment share the corpus rows they stream (one HBM pass, L2 hits for the rest).
#include "params.cuh"

namespace nxf {

namespace k1 {
constexpr int BLOCK_M = 128;
constexpr int BLOCK_N = 256;
constexpr int BLOCK_K = 64;
constexpr int UMMA_K = 16;
constexpr int THREADS = 256;
constexpr int EPI_WARP0 = 4;
constexpr int EPI_WARPS = 4;
constexpr int TMEM_COLS = 512;
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB

template <int CG> struct Cfg {
  static constexpr int LOAD_N = BLOCK_N / CG;
  static constexpr int B_BYTES = LOAD_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (CG == 1) ? 4 : 6;
  static constexpr int RING_BYTES = STAGES * STAGE_BYTES;  // 192 KB either way
  // ring | full[STAGES] empty[STAGES] tfull[2] tempty[2] | tmem ptr | hist[EPI_WARPS][256]
  static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8;
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + RING_BYTES + BAR_BYTES + 16 + EPI_WARPS * 256 * 4;
};
}  // namespace k1

template <int CG, bool BF16>
__global__ void __launch_bounds__(k1::THREADS, 1)
k1_search_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c, K1Params p) {
  using namespace k1;
  using C = Cfg<CG>;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::RING_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_ptr_s = reinterpret_cast<uint32_t*>(tempty + 2);
  uint32_t* hist = tmem_ptr_s + 4;

  const int warp = __shfl_sync(kFullMask, threadIdx.x >> 5, 0);
  const unsigned lane = lane_id();
  const uint32_t rank = (CG == 2) ? cluster_ctarank() : 0u;
  const int64_t cluster_id = blockIdx.x / CG;
  const int64_t num_clusters = gridDim.x / CG;
  const int64_t num_units = static_cast<int64_t>(p.num_qt) * p.chunks;

  if (CG == 2) cluster_sync_all();  // both CTAs of the pair are resident before cta_group::2 alloc
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full[s], CG); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], CG * EPI_WARPS); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<CG>(tmem_ptr_s, TMEM_COLS);
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_ptr_s);

  auto a_ptr = [&](int s) { return smem + s * C::STAGE_BYTES; };
  auto b_ptr = [&](int s) { return smem + s * C::STAGE_BYTES + A_BYTES; };

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int64_t u = cluster_id; u < num_units; u += num_clusters) {
        const int qt = static_cast<int>(u % p.num_qt);
        const int64_t c = u / p.num_qt;
        const int64_t t0 = p.tile_begin + c * p.tiles_per_chunk;
        const int64_t t1 = min(t0 + p.tiles_per_chunk, p.n_tiles);
        const int32_t q_row0 = static_cast<int32_t>((static_cast<int64_t>(qt) * CG + rank) * BLOCK_M);
        for (int64_t t = t0; t < t1; ++t) {
          const int32_t c_row0 = static_cast<int32_t>(t * BLOCK_N + rank * C::LOAD_N);
          for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&empty[s], ph ^ 1);
            if (CG == 1) {
              mbar_expect_tx(&full[s], C::STAGE_BYTES);
              tma_load_2d(&tmap_q, &full[s], a_ptr(s), kb * BLOCK_K, q_row0, kEvictLast);
              tma_load_2d(&tmap_c, &full[s], b_ptr(s), kb * BLOCK_K, c_row0, kEvictNormal);
            } else {
              tma_load_2d_2sm(&tmap_q, &full[s], a_ptr(s), kb * BLOCK_K, q_row0, kEvictLast);
              tma_load_2d_2sm(&tmap_c, &full[s], b_ptr(s), kb * BLOCK_K, c_row0, kEvictNormal);
              if (rank == 0) mbar_expect_tx(&full[s], 2 * C::STAGE_BYTES);
              else mbar_arrive_cluster(&full[s], 0);
            }
            if (++s == C::STAGES) { s = 0; ph ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = umma_idesc_f16(BF16, BLOCK_M * CG, BLOCK_N);
      int s = 0;
      uint32_t ph = 0;
      uint32_t iter = 0;
      for (int64_t u = cluster_id; u < num_units; u += num_clusters) {
        const int64_t c = u / p.num_qt;
        const int64_t t0 = p.tile_begin + c * p.tiles_per_chunk;
        const int64_t t1 = min(t0 + p.tiles_per_chunk, p.n_tiles);
        for (int64_t t = t0; t < t1; ++t, ++iter) {
          const uint32_t as = iter & 1, aph = (iter >> 1) & 1;
          mbar_wait(&tempty[as], aph ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BLOCK_N;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&full[s], ph);
            tc_fence_after();
            const uint64_t a_desc = umma_desc_sw128(smem_u32(a_ptr(s)));
            const uint64_t b_desc = umma_desc_sw128(smem_u32(b_ptr(s)));
#pragma unroll
            for (int k4 = 0; k4 < BLOCK_K / UMMA_K; ++k4) {
              // advance 16 elements = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
              umma_f16<CG>(d_tmem, a_desc + 2 * k4, b_desc + 2 * k4, idesc, (kb | k4) != 0 ? 1u : 0u);
            }
            umma_commit<CG>(&empty[s]);  // smem stage reusable once these MMAs retire
            if (++s == C::STAGES) { s = 0; ph ^= 1; }
          }
          umma_commit<CG>(&tfull[as]);   // accumulator complete -> epilogue
        }
      }
      // do not tear the barriers down while the last remote arrivals may still be in flight
      if (iter > 0) {
        const uint32_t last = iter - 1;
        mbar_wait(&tempty[last & 1], (last >> 1) & 1);
        if (iter > 1) {
          const uint32_t prev = iter - 2;
          mbar_wait(&tempty[prev & 1], (prev >> 1) & 1);
        }
      }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0) {
    // ============================== epilogue ==============================
    const int ew = warp - EPI_WARP0;  // == warp % 4: the TMEM lane quarter this warp may read
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
    uint32_t* my_hist = hist + ew * 256;
    const int cap = p.sink.cap;
    const int k = p.sink.k;
    Entry* warp_buf0 = p.sink.cand + (static_cast<int64_t>(blockIdx.x) * BLOCK_M + ew * 32) * cap;  // lane 0's buffer
    Entry* my_buf = warp_buf0 + static_cast<int64_t>(lane) * cap;
    const int32_t n_rows = static_cast<int32_t>(p.n);
    uint32_t iter = 0;

    for (int64_t u = cluster_id; u < num_units; u += num_clusters) {
      const int qt = static_cast<int>(u % p.num_qt);
      const int64_t c = u / p.num_qt;
      const int64_t t0 = p.tile_begin + c * p.tiles_per_chunk;
      const int64_t t1 = min(t0 + p.tiles_per_chunk, p.n_tiles);
      const int64_t qrow0 = (static_cast<int64_t>(qt) * CG + rank) * BLOCK_M + ew * 32;  // lane 0's query
      const int64_t qrow = qrow0 + lane;
      const bool valid = qrow < p.nq;
      float tau = valid ? tau_load(p.sink.tau_g, qrow) : INFINITY;  // padded rows accept nothing
      int cnt = 0;
      float tau_next = tau;

      for (int64_t t = t0; t < t1; ++t, ++iter) {
        const uint32_t as = iter & 1, aph = (iter >> 1) & 1;
        tau = fmaxf(tau, tau_next);  // other CTAs' progress on this query (loaded during the previous tile)
        mbar_wait(&tfull[as], aph);
        tc_fence_after();
        const int32_t col0 = static_cast<int32_t>(t * BLOCK_N);
#pragma unroll 1
        for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
          uint32_t vr[32];
          if (!(p.debug & 2)) {
            tmem_ld32(taddr + as * BLOCK_N + ch * 32, vr);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) vr[j] = 0;
          }
          if (ch == BLOCK_N / 32 - 1) {
            // accumulator fully read: hand the TMEM buffer back to the MMA issuer
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CG == 2 && rank != 0) mbar_arrive_cluster(&tempty[as], 0); else mbar_arrive(&tempty[as]);
            }
          }
          if (p.debug & 1) continue;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(vr[j]);
          const int32_t cbase = col0 + ch * 32;
          if (p.half_norm != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] -= p.half_norm[min(cbase + j, n_rows - 1)];
          }
          float m = fmax3(v[0], v[1], v[2]);
#pragma unroll
          for (int j = 3; j < 31; j += 2) m = fmax3(m, v[j], v[j + 1]);
          m = fmaxf(m, v[31]);
          unsigned hm = __ballot_sync(kFullMask, m > tau);
          if (hm) {
            // Warp-cooperative append: a row with a survivor publishes its 32 scores through shared memory, then
            // the 32 lanes test one column each and the survivors are written with one coalesced store.
            float* stage = reinterpret_cast<float*>(my_hist);  // the prune histogram is idle here
            do {
              const int src = __ffs(hm) - 1;
              hm &= hm - 1;
              if (static_cast<int>(lane) == src) {
                float4* d4 = reinterpret_cast<float4*>(stage);
#pragma unroll
                for (int j = 0; j < 8; ++j) d4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
              }
              __syncwarp();
              const float sc = stage[lane];
              const float tau_src = __shfl_sync(kFullMask, tau, src);
              const int cnt_src = __shfl_sync(kFullMask, cnt, src);
              const int32_t col = cbase + static_cast<int32_t>(lane);
              const bool pass = sc > tau_src && col < n_rows;
              const unsigned pm = __ballot_sync(kFullMask, pass);
              if (pass) {
                Entry e; e.score = sc; e.row = col;
                (warp_buf0 + static_cast<int64_t>(src) * cap)[cnt_src + __popc(pm & lanemask_lt())] = e;
              }
              if (static_cast<int>(lane) == src) cnt += __popc(pm);
              __syncwarp();
            } while (hm);
            __syncwarp();
            unsigned need = __ballot_sync(kFullMask, cnt > cap - 32);
            while (need) {
              const int src = __ffs(need) - 1;
              need &= need - 1;
              const int n_src = __shfl_sync(kFullMask, cnt, src);
              const uint32_t kth = warp_prune(warp_buf0 + static_cast<int64_t>(src) * cap, n_src, k, my_hist);
              if (static_cast<int>(lane) == src) {
                cnt = k;
                tau = fmaxf(tau, okey_inv(kth));
                if (valid) atomicMax(&p.sink.tau_g[qrow], kth);
              }
              __syncwarp();
            }
          }
        }
        if (valid) tau_next = tau_load(p.sink.tau_g, qrow);
      }

      // ---- unit end: prune to k, publish, flush the 32 rows of this warp to the arena ----
      __syncwarp();
      unsigned over = __ballot_sync(kFullMask, cnt > k);
      while (over) {
        const int src = __ffs(over) - 1;
        over &= over - 1;
        const int n_src = __shfl_sync(kFullMask, cnt, src);
        const uint32_t kth = warp_prune(warp_buf0 + static_cast<int64_t>(src) * cap, n_src, k, my_hist);
        if (static_cast<int>(lane) == src) {
          cnt = k;
          if (valid) atomicMax(&p.sink.tau_g[qrow], kth);
        }
        __syncwarp();
      }
      if (!valid) cnt = 0;
      uint32_t pos = 0;
      if (cnt > 0) pos = atomicAdd(&p.sink.qcnt[qrow], static_cast<uint32_t>(cnt));
      if (static_cast<int64_t>(pos) + cnt > p.sink.arena_stride) cnt = 0;  // cannot happen by construction
      __syncwarp();
#pragma unroll 4
      for (int src = 0; src < 32; ++src) {
        const int n_src = __shfl_sync(kFullMask, cnt, src);
        const uint32_t pos_src = __shfl_sync(kFullMask, pos, src);
        const Entry* from = warp_buf0 + static_cast<int64_t>(src) * cap;
        Entry* to = p.sink.arena + (qrow0 + src) * p.sink.arena_stride + pos_src;
        for (int i = lane; i < n_src; i += 32) to[i] = from[i];
      }
      __syncwarp();
    }
  }

  // ------------------------------ teardown ------------------------------
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || p == nullptr) return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

// rows x ld 16-bit matrix, box = 64 elements (128 B, one swizzle atom) x box_rows
static int make_tmap(CUtensorMap* m, const void* base, int64_t rows, int64_t ld, int box_rows, bool bf16) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return NXF_ECUDA; }
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(ld), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld ld=%lld)", (int)r, (long long)rows, (long long)ld); return NXF_ECUDA; }
  return NXF_OK;
}

template <int CG, bool BF16>
static int launch_k1_t(const CUtensorMap& tq, const CUtensorMap& tc, const K1Params& p, int grid, cudaStream_t stream) {
  using C = k1::Cfg<CG>;
  auto kern = k1_search_kernel<CG, BF16>;
  NXF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(k1::THREADS);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  NXF_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tc, p));
  return NXF_OK;
}

// q16: [nq_rows, ld] 16-bit queries (device), corpus: [n, ld] 16-bit rows (device)
int launch_k1(int store_dtype, int cta_group, const void* q16, int64_t nq_rows, const void* corpus, int64_t ld, const K1Params& p,
              int grid, cudaStream_t stream) {
  const bool bf16 = store_dtype == NXF_BF16;
  if (store_dtype != NXF_BF16 && store_dtype != NXF_F16) { set_error("tensor-core search needs f16/bf16 storage"); return NXF_EUNSUPPORTED; }
  CUtensorMap tq, tc;
  int rc = make_tmap(&tq, q16, nq_rows, ld, k1::BLOCK_M, bf16);
  if (rc) return rc;
  rc = make_tmap(&tc, corpus, p.n, ld, k1::BLOCK_N / cta_group, bf16);
  if (rc) return rc;
  if (cta_group == 2) return bf16 ? launch_k1_t<2, true>(tq, tc, p, grid, stream) : launch_k1_t<2, false>(tq, tc, p, grid, stream);
  return bf16 ? launch_k1_t<1, true>(tq, tc, p, grid, stream) : launch_k1_t<1, false>(tq, tc, p, grid, stream);
}

int k1_smem_bytes(int cta_group) { return cta_group == 2 ? k1::Cfg<2>::SMEM_BYTES : k1::Cfg<1>::SMEM_BYTES; }

}  // namespace nxf
