// search_stream_tc.cu -- K2T: small-batch (nq <= 128) exact search on tcgen05 with the CORPUS as the M operand.
//
// The shape in which the reference calls search today is a handful of queries against the whole corpus: 32-row
// chunks (Nexus/evaluation/text_retrieval/utils.py:221-224), 128 users per recsys batch
// (Nexus/inference/embedder/recommendation/base.py:182), one query per request
// (serve/inference_service_text_retrieval.py:70).  Every such call is one pass over the corpus, i.e. bound by HBM
// bandwidth: algorithmic bytes = N*ld*2 per launch.  CUDA-core FMAs stop keeping up with HBM beyond one or two
// queries, so the dot products go to the tensor cores -- but, unlike K1, with the roles swapped:
//   A (M = 128)  = 128 corpus rows, streamed by TMA through a shared-memory ring, 16 KB per 64-wide K block
//   B (N = nq')  = the queries (nq rounded up to 16), loaded ONCE and resident in shared memory for the whole launch
//   D in TMEM    = [128 corpus rows x nq'] fp32, several buffers deep
// so shared memory and L2 only ever carry the corpus, once.  The epilogue thread owns a corpus row and compares its
// nq' scores with the per-query thresholds kept in shared memory (v - tau > 0 on a max tree); survivors are appended
// to per-(CTA, query) candidate buffers with a shared-memory atomic counter.  A named barrier with an OR reduction
// once per tile decides (uniformly) whether any buffer must be pruned before the next tile.
#include "params.cuh"

namespace nxf {

// Knock-out builds for locating the stage that paces the stream (tools/build_variant.sh k2t_koN "-DNXF_K2T_KNOCKOUT=N"; results are
// wrong by construction, timing only):  1 = the epilogue only waits for and releases the accumulator buffers;  2 = it also reads
// them (tcgen05.ld) but filters nothing;  3 = as 1, and the issuer issues no MMA (only the commits): TMA + barriers alone;
// 5 = everything runs, but every threshold is +inf: the filter's cost without a single survivor.
#ifndef NXF_K2T_KNOCKOUT
#define NXF_K2T_KNOCKOUT 0
#endif
// who issues inside the producer warp: 1 = the lane elect.sync picks, 0 = lane 0 (same-box A/B with tools/build_variant.sh: no
// difference; the issuer warp's choice follows the template parameter SETS, see the issuer)
#ifndef NXF_K2T_ELECT_PRODUCER
#define NXF_K2T_ELECT_PRODUCER 1
#endif

namespace k2t {
constexpr int BLOCK_M = 128;   // corpus rows per tile
constexpr int BLOCK_K = 64;
constexpr int UMMA_K = 16;
constexpr int EPI_WARP0 = 4;
// Epilogue warp sets (template parameter SETS = 2..4, four warps each, one per SM sub-partition): set s filters tiles s, s + SETS, ...
// Few queries: two sets keep up and leave the issue slots of sub-partitions 0/1 to the producer and the MMA issuer (one query on
// 128-byte rows: 0.218 ms with two sets, 0.244 ms with three); from ~64 queries on the per-tile filter is bound by dependent-issue
// latency and more sets hide it (128 queries: 0.439 ms with two, 0.366 with three, 0.342 with four -- at 102 registers per thread).
constexpr int threads(int sets) { return 128 + 128 * sets; }
constexpr int TMEM_COLS = 512;
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB per stage
constexpr int MAX_STAGES = 12;
constexpr int MAX_TBUF = 8;
constexpr int MAX_NQ = 128;
// shared memory after the 1024-aligned (queries | ring): barriers, tau, counts
constexpr int TAIL_BYTES = (2 * MAX_STAGES + 2 * MAX_TBUF + 2) * 8 + 16 + MAX_NQ * 4 * 2 + 2 * MAX_NQ * 4;  // + tauv[2][MAX_NQ] (FOLD)
constexpr int ONES_BYTES = BLOCK_M * 128;  // FOLD: the constant A' tile (128 rows x one 128-byte swizzle row)
}  // namespace k2t

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait_dep16(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
// named barrier 1 over the epilogue threads with an OR reduction of a predicate
template <int SETS>
__device__ __forceinline__ bool epi_bar_or(bool pred) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %1, 0;\n\t"
      "barrier.cta.red.or.pred.aligned q, 1, %2, p;\n\t"
      "selp.u32 %0, 1, 0, q;\n\t}"
      : "=r"(r) : "r"(static_cast<uint32_t>(pred)), "n"(128 * SETS) : "memory");
  return r != 0;
}
template <int SETS>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("barrier.cta.sync.aligned 1, %0;" ::"n"(128 * SETS) : "memory"); }

// One 16-query slice of a corpus row's scores: d = v - tau on a max tree; the common case is one compare, branch not taken.
// The thresholds of the slice are passed in (loaded by the caller BEFORE it waits for the TMEM load, so that the two latencies
// overlap), and the maximum is a balanced tree of depth 3 -- with two epilogue warps per sub-partition it is dependent-issue
// latency, not issue slots, that bounds this loop at 128 queries.
struct TauSlice { float4 t[4]; };
__device__ __forceinline__ TauSlice k2t_load_tau(const float* __restrict__ tau_s, int ch) {
  TauSlice r;
  const float4* tp = reinterpret_cast<const float4*>(tau_s + ch * 16);  // same address for every lane: broadcast
#pragma unroll
  for (int j4 = 0; j4 < 4; ++j4) r.t[j4] = tp[j4];
  return r;
}
template <bool L2>
__device__ __forceinline__ void k2t_chunk(const uint32_t (&vr)[16], const TauSlice& ts, int ch, float hn, int64_t row, bool row_ok,
                                          int* __restrict__ cnt_s, uint32_t* __restrict__ cand_words, int cap) {
  float v[16], d[16];
#pragma unroll
  for (int j4 = 0; j4 < 4; ++j4) {
    const float4 tq = ts.t[j4];
#pragma unroll
    for (int e = 0; e < 4; ++e) v[4 * j4 + e] = L2 ? __uint_as_float(vr[4 * j4 + e]) - hn : __uint_as_float(vr[4 * j4 + e]);
    d[4 * j4] = v[4 * j4] - tq.x;
    d[4 * j4 + 1] = v[4 * j4 + 1] - tq.y;
    d[4 * j4 + 2] = v[4 * j4 + 2] - tq.z;
    d[4 * j4 + 3] = v[4 * j4 + 3] - tq.w;
  }
  const float m0 = fmax3(d[0], d[1], d[2]), m1 = fmax3(d[3], d[4], d[5]), m2 = fmax3(d[6], d[7], d[8]);
  const float m3 = fmax3(d[9], d[10], d[11]), m4 = fmax3(d[12], d[13], d[14]);
  const float m = fmax3(fmax3(m0, m1, m2), fmaxf(m3, m4), d[15]);
  if (m > 0.f && row_ok) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (d[j] > 0.f) {  // v > tau (tau = +inf for padded columns gives d = -inf / NaN: never true)
        const int qi = ch * 16 + j;
        const int pos = atomicAdd(&cnt_s[qi], 1);
        NXF_DEV_CHECK(pos < cap, "K2T per-(CTA, query) list overflow");
        const uint32_t w = static_cast<uint32_t>(qi * cap + pos);
        cand_words[2 * static_cast<uint64_t>(w)] = __float_as_uint(v[j]);
        cand_words[2 * static_cast<uint64_t>(w) + 1] = static_cast<uint32_t>(row);
      }
    }
  }
}

// FOLD fast filter: the largest of 32 folded scores (11 FMNMX3 + 1 FMNMX, depth 4)
__device__ __forceinline__ float k2t_max32(const uint32_t (&r)[32]) {
  float m[11];
#pragma unroll
  for (int i = 0; i < 10; ++i) m[i] = fmax3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
  m[10] = fmaxf(__uint_as_float(r[30]), __uint_as_float(r[31]));
  const float a = fmax3(m[0], m[1], m[2]), b = fmax3(m[3], m[4], m[5]), c = fmax3(m[6], m[7], m[8]), d = fmaxf(m[9], m[10]);
  return fmaxf(fmax3(a, b, c), d);
}

// FOLD: thread q's query threshold as three bf16 pieces t0 + t1 + t2 <= tau (the last one rounded down, so the folded threshold never
// exceeds the true one: nothing that beats tau is filtered out), written as -t into row q of the B' tile (128-byte swizzled rows:
// 16-byte chunk c of row r sits at chunk position c ^ (r % 8)), and their exact fp32 sum into tauv[q].
__device__ __forceinline__ void k2t_write_fold(unsigned char* __restrict__ tile, float* __restrict__ tauv, int q, float tau, float lower) {
  float t = fmaxf(tau, lower);               // -inf (no threshold yet) -> the Cauchy-Schwarz lower bound of any score of this query
  if (!(t < 3.0e38f)) t = 3.0e38f;           // padded columns (+inf): nothing passes
  const __nv_bfloat16 b0 = __float2bfloat16_rn(t);
  const float r1 = t - __bfloat162float(b0);
  const __nv_bfloat16 b1 = __float2bfloat16_rn(r1);
  const float r2 = r1 - __bfloat162float(b1);
  const __nv_bfloat16 b2 = __float2bfloat16_rd(r2);
  tauv[q] = (__bfloat162float(b0) + __bfloat162float(b1)) + __bfloat162float(b2);   // exact: three disjoint 8-bit pieces
  const uint32_t n0 = static_cast<uint32_t>(__bfloat16_as_ushort(b0)) ^ 0x8000u, n1 = static_cast<uint32_t>(__bfloat16_as_ushort(b1)) ^ 0x8000u;
  const uint32_t n2 = static_cast<uint32_t>(__bfloat16_as_ushort(b2)) ^ 0x8000u;   // sign flipped: the tile holds -t
  uint2 w;
  w.x = n0 | (n1 << 16);
  w.y = n2;
  *reinterpret_cast<uint2*>(tile + (q >> 3) * 1024 + (q & 7) * 128 + ((q & 7) << 4)) = w;   // chunk 0 of row q: one 8-byte store
}

// Estimate pass: nothing is appended; the best score of this CTA's rows per query is kept (as a monotone key) in est_s.
// One REDUX.MAX per query and warp, lane j keeps query j's maximum, 16 shared-memory atomics per slice.
template <bool L2>
__device__ __forceinline__ void k2t_chunk_est(const uint32_t (&vr)[16], int ch, float hn, bool row_ok, uint32_t* __restrict__ est_s,
                                              unsigned lane) {
  uint32_t mine = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float v = L2 ? __uint_as_float(vr[j]) - hn : __uint_as_float(vr[j]);
    const uint32_t m = __reduce_max_sync(kFullMask, row_ok ? okey(v) : 0u);
    if (lane == static_cast<unsigned>(j)) mine = m;
  }
  if (lane < 16) atomicMax(&est_s[ch * 16 + lane], mine);
}

// FOLD (bf16, inner product, short rows; K2TParams::fold): one extra tcgen05.mma per tile, K = 16, of a constant A' tile (columns 0-2
// of every corpus row = 1) against a B' tile holding -tau of every query as three bf16 pieces, so the accumulator ends as
// score - tau and the epilogue's filter needs neither the thresholds nor a subtraction.  The issuer runs up to `tbufs` tiles ahead of
// the epilogue, so B' exists in two versions: the tiles of round R (round = SETS * check_every tiles) use version R & 1, and the
// thresholds the epilogue agrees on at the end of round R are written into version R & 1 -- whose tiles are all consumed by then --
// for round R + 2.  The threshold the filter sees is therefore up to two rounds old (a few more appends, never a wrong result).
template <bool BF16, bool L2, int SETS, bool FOLD>
__global__ void __launch_bounds__(k2t::threads(SETS), 1)
k2t_stream_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c, K2TParams p) {
  using namespace k2t;
  constexpr int EPI_SETS = SETS;
  static_assert(!FOLD || (BF16 && !L2), "the threshold fold is built for bf16 rows and the inner-product metric");
  extern __shared__ unsigned char smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array itself (an integer round trip would make every later access generic)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int q_bytes = p.kblocks * p.nq_pad * 128;          // resident queries: [kblocks][nq_pad rows][128 B], 1024-aligned blocks
  unsigned char* q_smem = smem;
  unsigned char* ring = smem + ((q_bytes + 1023) & ~1023);
  unsigned char* ones_tile = ring + p.stages * A_BYTES;                       // FOLD only: [128][128 B]
  const int tq_bytes = FOLD ? ((p.nq_pad * 128 + 1023) & ~1023) : 0;
  unsigned char* tauq_tile = ones_tile + (FOLD ? ONES_BYTES : 0);             // FOLD only: two versions of [nq_pad][128 B]
  unsigned char* tail = tauq_tile + 2 * tq_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint64_t* tempty = tfull + MAX_TBUF;
  uint64_t* qfull = tempty + MAX_TBUF;
  uint32_t* tmem_ptr_s = reinterpret_cast<uint32_t*>(qfull + 2);
  float* tau_s = reinterpret_cast<float*>(tmem_ptr_s + 4);  // [MAX_NQ]
  int* cnt_s = reinterpret_cast<int*>(tau_s + MAX_NQ);      // [MAX_NQ]
  float* tauv = reinterpret_cast<float*>(cnt_s + MAX_NQ);   // FOLD: [2][MAX_NQ] folded thresholds of the two versions
  float fold_lower = -INFINITY;                             // FOLD, thread q < nq: -(|q| * max|c|) * 1.01, below every score of query q

  if (p.run_flag != nullptr && ld_relaxed_u32(p.run_flag) == 0u) return;  // uniform over the grid: written by an earlier kernel
  const int warp = __shfl_sync(kFullMask, threadIdx.x >> 5, 0);
  const unsigned lane = lane_id();
  const int64_t t_first = p.tile_begin + blockIdx.x;  // this CTA's tiles: t_first, t_first + grid, ... < t_end
  const int64_t t_end = p.tile_end;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < p.tbufs; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 4); }
    mbar_init(qfull, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<1>(tmem_ptr_s, TMEM_COLS);
  if (threadIdx.x < MAX_NQ) {
#if NXF_K2T_KNOCKOUT == 5   // the whole filter runs, nothing ever passes it
    tau_s[threadIdx.x] = INFINITY;
#else
    tau_s[threadIdx.x] = (static_cast<int>(threadIdx.x) < p.nq) ? tau_load(p.sink.tau_g, threadIdx.x) : INFINITY;  // padded columns accept nothing
#endif
    cnt_s[threadIdx.x] = (p.first_pass || static_cast<int>(threadIdx.x) >= p.nq)
                             ? 0 : static_cast<int>(p.cnt_g[static_cast<int64_t>(blockIdx.x) * p.nq_pad + threadIdx.x]);
  }
  if (FOLD) {
    // zero the three tiles, set the ones columns, fold the initial thresholds into both versions
    uint4* z = reinterpret_cast<uint4*>(ones_tile);
    const int n16 = (ONES_BYTES + 2 * tq_bytes) / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    if (threadIdx.x < BLOCK_M) {
      const int r = threadIdx.x;
      uint2 w;
      w.x = 0x3F803F80u;   // bf16 1.0, 1.0
      w.y = 0x00003F80u;   // 1.0, 0
      *reinterpret_cast<uint2*>(ones_tile + (r >> 3) * 1024 + (r & 7) * 128 + ((r & 7) << 4)) = w;
    }
    if (threadIdx.x < MAX_NQ && static_cast<int>(threadIdx.x) < p.nq_pad) {
      const int q = threadIdx.x;
      if (q < p.nq) {
        float ss = 0.f;
        const float4* qr = reinterpret_cast<const float4*>(p.q_f32 + static_cast<int64_t>(q) * p.ldq);   // rows are zero padded to ldq
#pragma unroll 8
        for (int d = 0; d < p.kblocks * 16; ++d) {
          const float4 x = qr[d];
          ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
        }
        fold_lower = -(sqrtf(ss) * sqrtf(__uint_as_float(*p.max_norm2)) * 1.01f + 1e-30f);
      }
      k2t_write_fold(tauq_tile, tauv, q, tau_s[q], fold_lower);
      k2t_write_fold(tauq_tile + tq_bytes, tauv + MAX_NQ, q, tau_s[q], fold_lower);
    }
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_ptr_s);

  if (warp == 0) {
    // ============================== TMA producer ==============================
    // All 32 lanes run the loop and wait on the barriers; lane 0 alone issues.  (With the loop inside `if (lane == 0)` the other 31
    // lanes sit at the closing __syncwarp as a second, divergent instruction stream of the same warp and take a quarter of its issue
    // turns -- measured on 128-byte rows, where this thread and the MMA issuer pace the whole stream.)
    if (lane == 0) {
      mbar_expect_tx(qfull, static_cast<uint32_t>(q_bytes));
      for (int kb = 0; kb < p.kblocks; ++kb) tma_load_2d(&tmap_q, qfull, q_smem + kb * p.nq_pad * 128, kb * BLOCK_K, 0, kEvictLast);
    }
    int s = 0;
    uint32_t ph = 0;
    for (int64_t t = t_first; t < t_end; t += gridDim.x) {
      const int32_t row0 = static_cast<int32_t>(perm_tile(p.perm, t) * BLOCK_M);
      for (int kb = 0; kb < p.kblocks; ++kb) {
        mbar_wait(&empty[s], ph ^ 1);
        if (NXF_K2T_ELECT_PRODUCER ? elect_one() : lane == 0) {
          mbar_expect_tx(&full[s], A_BYTES);
          tma_load_2d(&tmap_c, &full[s], ring + s * A_BYTES, kb * BLOCK_K, row0, kEvictFirst);
        }
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    // The issuing thread paces the whole stream on short rows (one K block per tile: a tile every ~700 clocks at HBM speed), so the
    // per-tile instruction chain is kept short: running counters instead of divisions, descriptors advanced by addition, warp-uniform
    // loop.  (A second issuer warp on alternate tiles and pairs of warp sets sharing a tile were built and measured in round 2 --
    // no gain at 128 queries, and their run-time parameters in these loops cost the 1-32 query kernels 3-5 %: removed again,
    // profiles/r2_k2t_knockout.md.)
    const uint32_t idesc = umma_idesc_f16(BF16, BLOCK_M, p.nq_pad);
    mbar_wait(qfull, 0);
    tc_fence_after();
    int s = 0;
    uint32_t ph = 0;
    uint32_t b = 0, bph = 0;
    const uint64_t a_desc0 = umma_desc_sw128(smem_u32(ring));          // stage s: + s * (A_BYTES >> 4) in the address field
    const uint64_t b_desc0 = umma_desc_sw128(smem_u32(q_smem));        // query block qb: + qb * (nq_pad * 128 >> 4)
    const uint32_t q_blk = static_cast<uint32_t>(p.nq_pad) * 8u;
    const uint64_t ones_desc = umma_desc_sw128(smem_u32(ones_tile));
    const uint64_t tauq_desc0 = umma_desc_sw128(smem_u32(tauq_tile));  // version v: + v * (tq_bytes >> 4)
    const uint32_t fold_round = static_cast<uint32_t>(EPI_SETS) * static_cast<uint32_t>(p.check_every);
    uint32_t in_round = 0, fold_ver = 0;
    for (int64_t t = t_first; t < t_end; t += gridDim.x) {
      mbar_wait(&tempty[b], bph ^ 1);
      const uint32_t d_tmem = tmem_base + b * p.tbuf_stride;
      for (int kb = 0; kb < p.kblocks; ++kb) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        // Who issues: the lane elect.sync picks (kernels with three or four warp sets, i.e. more than 32 queries) or lane 0 (two
        // sets).  Under `lane == 0` nvcc wraps each tcgen05 instruction in an elect-and-retry loop (five dependent instructions each);
        // under elect.sync the MMAs and commits of a tile are consecutive.  Same box, C5 corpus, k = 10, whole search: 64 queries
        // 0.272 -> 0.243 ms and 128 queries 0.342 -> 0.334 ms with elect.sync, but 1 / 8 queries 0.234 / 0.239 -> 0.243 / 0.245 ms
        // (32: equal) -- hence both forms, chosen at compile time with the set count (a run-time choice between two copies in this
        // loop cost the 32-query case 6 %; profiles/r2_k2t_elect_ab.log; the producer's TMA is insensitive to it).
        constexpr bool ELECT_ISSUE = SETS >= 3;
        auto issue_tile_step = [&]() {
          const uint64_t a_desc = a_desc0 + static_cast<uint32_t>(s) * static_cast<uint32_t>(A_BYTES >> 4);
          const uint64_t b_desc = b_desc0 + static_cast<uint32_t>(kb) * q_blk;
#if NXF_K2T_KNOCKOUT != 3
          umma_f16<1>(d_tmem, a_desc, b_desc, idesc, kb != 0 ? 1u : 0u);
#pragma unroll
          for (int k4 = 1; k4 < BLOCK_K / UMMA_K; ++k4) umma_f16<1>(d_tmem, a_desc + 2 * k4, b_desc + 2 * k4, idesc, 1u);
#endif
          umma_commit<1>(&empty[s]);
          if (kb == p.kblocks - 1) {
            if (FOLD) umma_f16<1>(d_tmem, ones_desc, tauq_desc0 + fold_ver * static_cast<uint32_t>(tq_bytes >> 4), idesc, 1u);   // += -tau
            umma_commit<1>(&tfull[b]);
          }
        };
        if constexpr (ELECT_ISSUE) {
          if (elect_one()) issue_tile_step();
        } else {
          if (lane == 0) issue_tile_step();
        }
        __syncwarp();
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
      if (FOLD && ++in_round == fold_round) { in_round = 0; fold_ver ^= 1u; }
      if (++b == static_cast<uint32_t>(p.tbufs)) { b = 0; bph ^= 1; }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0) {
    // ============================== epilogue ==============================
    // Two sets of four warps (one warp of each set per SM sub-partition): set s filters this CTA's tiles s, s + 2, ...  The
    // thresholds, list counters and lists are shared by both sets (shared-memory atomics); the sets meet every 2*check_every
    // tiles at a named barrier, where the lists that could overflow are pruned while nobody appends.
    const int ew8 = warp - EPI_WARP0;  // 0..7
    const int ew = ew8 & 3;            // == warp % 4: the TMEM lane quarter this warp may read
    const int set = ew8 >> 2;
    const int et = ew * 32 + static_cast<int>(lane);   // 0..127: corpus row within the tile
    const int eq = ew8 * 32 + static_cast<int>(lane);  // 0..255: the query this thread looks after (if < nq)
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
    const int cap = p.sink.cap;
    const int k = p.sink.k;
    Entry* cta_cand = p.sink.cand + static_cast<int64_t>(blockIdx.x) * p.nq_pad * cap;  // [nq_pad][cap]
    uint32_t* cand_words = reinterpret_cast<uint32_t*>(cta_cand);
    const int chunks = p.nq_pad / 16;
    const uint32_t round_tiles = static_cast<uint32_t>(EPI_SETS) * static_cast<uint32_t>(p.check_every);  // tiles between two overflow checks / threshold refreshes
    // a list gains <= 128 entries per tile, i.e. <= 128 * round_tiles between two checks, and the counts read at a check are
    // exact (every appender has arrived at the barrier): cap - full_mark = one round of appends
    const int full_mark = cap - static_cast<int>(round_tiles) * BLOCK_M;
    const uint32_t n_it = t_first < t_end ? static_cast<uint32_t>((t_end - t_first + gridDim.x - 1) / gridDim.x) : 0u;
    uint32_t tau_next_key = (eq < p.nq) ? tau_load_key(p.sink.tau_g, eq) : 0u;   // raw key until used (sink.cuh)
    // accumulator buffer and phase of this set's next tile (iter = set, set + 2, ...: round_tiles is even, so the sequence runs on
    // across rounds), kept as running counters: iter % tbufs and iter / tbufs are runtime divisions otherwise
    uint32_t b = static_cast<uint32_t>(set) % static_cast<uint32_t>(p.tbufs), bph = (static_cast<uint32_t>(set) / static_cast<uint32_t>(p.tbufs)) & 1u;

    uint32_t ver = 0;   // FOLD: version of the folded thresholds the tiles of this round were computed with
    for (uint32_t r0 = 0; r0 < n_it; r0 += round_tiles, ver ^= 1u) {
      const uint32_t r1 = min(n_it, r0 + round_tiles);
      const float* tv = tauv + ver * MAX_NQ;
      for (uint32_t iter = r0 + set; iter < r1; iter += EPI_SETS) {  // r0 is a multiple of EPI_SETS: iter % EPI_SETS is the set
        const int64_t t = t_first + static_cast<int64_t>(iter) * gridDim.x;
        const int64_t row = perm_tile(p.perm, t) * BLOCK_M + et;
        const bool row_ok = row < p.n;
        const float hn = (L2 && row_ok) ? p.half_norm[row] : 0.f;
        mbar_wait(&tfull[b], bph);
        tc_fence_after();
        const uint32_t tbase = taddr + b * p.tbuf_stride;
        if constexpr (FOLD) {
          // Folded thresholds: a row has a survivor iff the largest of its nq_pad values is positive, so the common case is one
          // compare per ROW, not per 16-query slice: 32-column loads (the next one in flight while a block is reduced), a max tree
          // per block, one warp vote.  (Knock-out builds, profiles/r2_k2t_knockout.md: with no survivors at all the per-slice filter
          // adds 65 us to the 204 us the stream takes with the filter compiled out, this one 30 us.)
          uint32_t fa[32], fb[32];
          const int ncol = p.nq_pad;   // a multiple of 32 (host), <= 128
          float m0, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
          tmem_ld32(tbase, fa);
          tmem_ld_wait_dep(fa);
          if (ncol > 32) tmem_ld32(tbase + 32, fb);
          m0 = k2t_max32(fa);
          if (ncol > 32) {
            tmem_ld_wait_dep(fb);
            if (ncol > 64) tmem_ld32(tbase + 64, fa);
            m1 = k2t_max32(fb);
          }
          if (ncol > 64) {
            tmem_ld_wait_dep(fa);
            if (ncol > 96) tmem_ld32(tbase + 96, fb);
            m2 = k2t_max32(fa);
          }
          if (ncol > 96) {
            tmem_ld_wait_dep(fb);
            m3 = k2t_max32(fb);
          }
          // which 32-query blocks hold a survivor for some row of this warp (about one warp in eight on the main pass, then
          // usually a single block): only those are read again -- and the buffer goes back to the issuer as soon as the last
          // of them is in registers, before the appends (shared-memory atomic + two global stores per survivor)
          const bool h0 = row_ok && m0 > 0.f, h1 = row_ok && m1 > 0.f, h2 = row_ok && m2 > 0.f, h3 = row_ok && m3 > 0.f;
          uint32_t hb = 0;
          if (__any_sync(kFullMask, h0 || h1 || h2 || h3)) {
            hb = (__any_sync(kFullMask, h0) ? 1u : 0u) | (__any_sync(kFullMask, h1) ? 2u : 0u) | (__any_sync(kFullMask, h2) ? 4u : 0u) |
                 (__any_sync(kFullMask, h3) ? 8u : 0u);
          }
          if (hb == 0u) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[b]);
          }
          while (hb != 0u) {   // warp-uniform
            const int blk = __ffs(hb) - 1;
            hb &= hb - 1u;
            tmem_ld32(tbase + blk * 32, fa);
            tmem_ld_wait_dep(fa);
            if (hb == 0u) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&tempty[b]);
            }
            const bool mine = blk == 0 ? h0 : (blk == 1 ? h1 : (blk == 2 ? h2 : h3));
            if (mine) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float d = __uint_as_float(fa[j]);
                if (d > 0.f) {
                  const int qi = blk * 32 + j;
                  const int pos = atomicAdd(&cnt_s[qi], 1);
                  NXF_DEV_CHECK(pos < cap, "K2T per-(CTA, query) list overflow");
                  const uint32_t w = static_cast<uint32_t>(qi * cap + pos);
                  cand_words[2 * static_cast<uint64_t>(w)] = __float_as_uint(d + tv[qi]);
                  cand_words[2 * static_cast<uint64_t>(w) + 1] = static_cast<uint32_t>(row);
                }
              }
            }
            __syncwarp();
          }
        } else {
        // two register buffers: the tcgen05.ld of slice ch+1 is in flight while slice ch is filtered
        uint32_t va[16], vb[16];
#if NXF_K2T_KNOCKOUT == 1 || NXF_K2T_KNOCKOUT == 3
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[b]);
        for (int ch = chunks; ch < chunks; ch += 2) {
#else
        tmem_ld16(tbase, va);
#pragma unroll 1
        for (int ch = 0; ch < chunks; ch += 2) {
#endif
          const TauSlice ta = k2t_load_tau(tau_s, ch);
          tmem_ld_wait_dep16(va);
          const bool more = ch + 1 < chunks;
          if (more) tmem_ld16(tbase + (ch + 1) * 16, vb);
          else {  // accumulator fully read: hand the buffer back to the MMA issuer
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[b]);
          }
#if NXF_K2T_KNOCKOUT == 0 || NXF_K2T_KNOCKOUT == 5
          if (p.est_only) k2t_chunk_est<L2>(va, ch, hn, row_ok, reinterpret_cast<uint32_t*>(cnt_s), lane);
          else k2t_chunk<L2>(va, ta, ch, hn, row, row_ok, cnt_s, cand_words, cap);
#endif
          if (more) {
            const TauSlice tb = k2t_load_tau(tau_s, ch + 1);
            tmem_ld_wait_dep16(vb);
            if (ch + 2 < chunks) tmem_ld16(tbase + (ch + 2) * 16, va);
            else {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&tempty[b]);
            }
#if NXF_K2T_KNOCKOUT == 0 || NXF_K2T_KNOCKOUT == 5
            if (p.est_only) k2t_chunk_est<L2>(vb, ch + 1, hn, row_ok, reinterpret_cast<uint32_t*>(cnt_s), lane);
            else k2t_chunk<L2>(vb, tb, ch + 1, hn, row, row_ok, cnt_s, cand_words, cap);
#endif
          }
        }
        }  // !FOLD
        b += EPI_SETS;
        while (b >= static_cast<uint32_t>(p.tbufs)) { b -= static_cast<uint32_t>(p.tbufs); bph ^= 1u; }
      }
      if (p.no_check) continue;
      // ---- end of a round: could any query's list overflow during the next round? ----
      const bool any_full = epi_bar_or<SETS>(eq < p.nq && cnt_s[eq] > full_mark);
      if (any_full) {
        for (int qi = ew8; qi < p.nq; qi += 4 * EPI_SETS) {  // warp w prunes queries w, w+8, ...
          const int n_q = cnt_s[qi];
          if (n_q > full_mark && n_q > k) {
            const uint32_t kth = warp_prune(cta_cand + static_cast<int64_t>(qi) * cap, n_q, k);
            if (lane == 0) {
              cnt_s[qi] = k;
              tau_s[qi] = fmaxf(tau_s[qi], okey_inv(kth));
              atomicMax(&p.sink.tau_g[qi], kth);
            }
            __syncwarp();
          }
        }
        epi_bar_sync<SETS>();
      }
      // other CTAs' progress: thread eq looks after query eq (value loaded one round ago)
      if (eq < p.nq) {
        const float tau_next = tau_from_key(tau_next_key);
        if (tau_next > tau_s[eq]) tau_s[eq] = tau_next;
        tau_next_key = tau_load_key(p.sink.tau_g, eq);
        if (FOLD) {
          // this round's version is free (every tile computed with it has been consumed): it becomes the thresholds of round R + 2
          k2t_write_fold(tauq_tile + ver * tq_bytes, tauv + ver * MAX_NQ, eq, tau_s[eq], fold_lower);
          fence_proxy_async();
        }
      }
    }

    // ---- end of the launch: the last one reduces every list to k' and flushes it to the arena; the others only record the
    // list lengths (the lists stay where they are for the seed kernel and the next launch) ----
    epi_bar_sync<SETS>();
    if (!p.last_pass) {
      if (eq < p.nq) p.cnt_g[static_cast<int64_t>(blockIdx.x) * p.nq_pad + eq] = static_cast<uint32_t>(cnt_s[eq]);
    } else {
      for (int qi = ew8; qi < p.nq; qi += 4 * EPI_SETS) {
        int n_q = cnt_s[qi];
        Entry* buf = cta_cand + static_cast<int64_t>(qi) * cap;
        if (n_q > k) {
          const uint32_t kth = warp_prune(buf, n_q, k);
          if (lane == 0) atomicMax(&p.sink.tau_g[qi], kth);
          n_q = k;
        }
        __syncwarp();
        warp_flush(p.sink, buf, n_q, qi);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<1>(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int make_tmap_2d_sw128(CUtensorMap* m, const void* base, int64_t rows, int64_t ld, int box_rows, bool bf16);  // search_tc.cu

size_t k2t_smem_bytes(int kblocks, int nq_pad, int stages, int fold) {
  const size_t q_bytes = (static_cast<size_t>(kblocks) * nq_pad * 128 + 1023) & ~size_t(1023);
  const size_t fold_bytes = fold ? k2t::ONES_BYTES + 2 * ((static_cast<size_t>(nq_pad) * 128 + 1023) & ~size_t(1023)) : 0;
  return 1024 + q_bytes + static_cast<size_t>(stages) * k2t::A_BYTES + fold_bytes + k2t::TAIL_BYTES;
}

// Picks the ring depth; returns 0 when the resident queries leave room for fewer than 4 stages (caller uses K1 instead).
int k2t_plan_stages(int kblocks, int nq_pad, int fold) {
  for (int st = k2t::MAX_STAGES; st >= 4; --st)
    if (k2t_smem_bytes(kblocks, nq_pad, st, fold) <= 227 * 1024) return st;
  return 0;
}

int launch_k2t(int store_dtype, const void* q16, int64_t nq_rows, int64_t q_ld, const void* corpus, int64_t ld, const K2TParams& p, int grid,
               cudaStream_t stream) {
  const bool bf16 = store_dtype == NXF_BF16;
  if (store_dtype != NXF_BF16 && store_dtype != NXF_F16) { set_error("tensor stream search needs f16/bf16 storage"); return NXF_EUNSUPPORTED; }
  CUtensorMap tq, tc;
  int rc = make_tmap_2d_sw128(&tq, q16, nq_rows, q_ld, p.nq_pad, bf16);
  if (rc) return rc;
  rc = make_tmap_2d_sw128(&tc, corpus, p.n, ld, k2t::BLOCK_M, bf16);
  if (rc) return rc;
  const bool l2 = p.half_norm != nullptr;
  const bool fold = p.fold && !p.est_only;   // the estimate pass has no thresholds to fold (same shared-memory budget either way)
  if (p.fold && (!bf16 || l2 || !p.q_f32 || !p.max_norm2)) { set_error("k2t: the threshold fold needs bf16 rows, the ip metric, q_f32 and max_norm2"); return NXF_EINVAL; }
  const size_t smem = k2t_smem_bytes(p.kblocks, p.nq_pad, p.stages, p.fold);
#define NXF_K2T_LAUNCH4(B, L, S, F)                                                                                                   \
  do {                                                                                                                                \
    rc = ensure_dyn_smem(reinterpret_cast<const void*>(k2t_stream_kernel<B, L, S, F>), static_cast<int>(smem));                       \
    if (rc) return rc;                                                                                                                \
    k2t_stream_kernel<B, L, S, F><<<grid, k2t::threads(S), smem, stream>>>(tq, tc, p);                                                \
  } while (0)
#define NXF_K2T_LAUNCH3(B, L, S) NXF_K2T_LAUNCH4(B, L, S, false)
#define NXF_K2T_LAUNCH(B, L)                     \
  do {                                           \
    if (p.epi_sets == 2) NXF_K2T_LAUNCH3(B, L, 2);      \
    else if (p.epi_sets == 3) NXF_K2T_LAUNCH3(B, L, 3); \
    else NXF_K2T_LAUNCH3(B, L, 4);                      \
  } while (0)
  if (p.epi_sets < 2 || p.epi_sets > 4) { set_error("k2t: %d epilogue sets (2..4)", p.epi_sets); return NXF_EINVAL; }
  if (fold) {
    if (p.epi_sets == 2) NXF_K2T_LAUNCH4(true, false, 2, true);
    else if (p.epi_sets == 3) NXF_K2T_LAUNCH4(true, false, 3, true);
    else NXF_K2T_LAUNCH4(true, false, 4, true);
  } else if (bf16 && l2) NXF_K2T_LAUNCH(true, true);
  else if (bf16) NXF_K2T_LAUNCH(true, false);
  else if (l2) NXF_K2T_LAUNCH(false, true);
  else NXF_K2T_LAUNCH(false, false);
#undef NXF_K2T_LAUNCH
#undef NXF_K2T_LAUNCH3
#undef NXF_K2T_LAUNCH4
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

}  // namespace nxf
