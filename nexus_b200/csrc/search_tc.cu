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
//           compare and no branch taken.  Survivors are appended to the thread's private candidate
//           buffer (global memory); a buffer that could overflow is reduced to its exact top-k by the
//           whole warp (bitwise-descent select, select.cuh) and tau is raised to the k-th score, also
//           published to tau_g[query] so every CTA working on the same query benefits.  At the end of a
//           launch every row appends what it holds to the query's arena slice.
// A search is several launches of this kernel over a sample-first permutation of the corpus tiles (estimate-only pass,
// two sample passes, the main pass, a fall-back pass that normally returns at once): DESIGN.md "sampled thresholds";
// the schedule lives in api.cu, select_tau_kernel / verify_counts_kernel in merge.cu run between the launches.
// Work decomposition ("interleaved rounds"): the corpus is cut into rounds of R tiles sized to stay
// L2-resident (tens of MB).  Inside a round the num_qt x R tile-units are laid on a line, query tile major, and cut
// into P equal pieces, one per cluster -- so every cluster works on the SAME <= 2 query tiles for the whole launch
// (their thresholds, counters and candidate buffers never move) while all clusters consume the same round of the
// corpus at different offsets: each corpus tile comes from HBM once and from L2 for the other query tiles.  A soft
// (timeout-guarded) round barrier between the producers keeps the clusters within one round of each other.
#include <mutex>

#include "params.cuh"

namespace nxf {

// Timing experiments (producer without loads, issuer without MMAs, epilogue without selection, prune counters; K1Params::debug) are
// compiled in only with -DNXF_K1_DEBUG_BUILD=1 (NXF_EXTRA_NVCC_FLAGS, tools/k1_variants.sh): the shipped kernel carries none of
// their tests in its hot loops (~1 % of stall samples each on three of them).
#ifndef NXF_K1_DEBUG_BUILD
#define NXF_K1_DEBUG_BUILD 0
#endif
#define NXF_K1_DBG(p_debug, bits) (NXF_K1_DEBUG_BUILD != 0 && ((p_debug) & (bits)) != 0)

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
  // ring | full[STAGES] empty[STAGES] tfull[4] tempty[4] | tmem ptr   (SPLIT kernels use all four of each, the others two)
  static constexpr int BAR_BYTES = (2 * STAGES + 8) * 8;
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + RING_BYTES + BAR_BYTES + 16;
};
}  // namespace k1


// ---------------------------------------------------------------------------------------------
// epilogue helpers
// ---------------------------------------------------------------------------------------------
struct EpiState {
  float tau;  // this row's running threshold (monotone)
  int cnt;    // entries in this row's private candidate buffer
};

__device__ __forceinline__ uint32_t ld_cg_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ Entry ld_cg_entry(const Entry* p) {
  uint32_t a, b;
  asm volatile("ld.global.cg.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
  Entry e;
  e.score = __uint_as_float(a);
  e.row = static_cast<int32_t>(b);
  return e;
}

// One 32-column slice of the accumulator: this thread's row x 32 corpus rows [cbase, cbase+32).
// Common case: 18 max ops and one branch not taken.  A row whose slice maximum beats its threshold scans the
// 8-column group(s) that did and appends the survivors to its PRIVATE buffer (no cross-lane work, bounded time: the
// accumulator is held while this runs).  Buffers are pruned between tiles (phase B in the kernel), never here.
// ADJ = false is the hot instantiation: the accumulator registers are used in place (no copies).  ADJ = true (L2 metric
// or the ragged last corpus tile) first subtracts 0.5*|c|^2 and masks the rows past the end of the corpus.
template <bool ADJ>
__device__ __forceinline__ void epi_chunk(const uint32_t (&vr)[32], int32_t cbase, int32_t n_rows,
                                          const float* __restrict__ half_norm, EpiState& st, uint32_t* __restrict__ cand_words,
                                          uint32_t my_first, int debug, int est_only) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(vr[j]);
  if (ADJ) {
    if (half_norm != nullptr) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] -= half_norm[min(cbase + j, n_rows - 1)];
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) if (cbase + j >= n_rows) v[j] = -INFINITY;
  }
  float g[4];
#pragma unroll
  for (int q4 = 0; q4 < 4; ++q4) {
    float a = fmax3(v[8 * q4], v[8 * q4 + 1], v[8 * q4 + 2]);
    a = fmax3(a, v[8 * q4 + 3], v[8 * q4 + 4]);
    a = fmax3(a, v[8 * q4 + 5], v[8 * q4 + 6]);
    g[q4] = fmaxf(a, v[8 * q4 + 7]);
  }
  const float m = fmaxf(fmax3(g[0], g[1], g[2]), g[3]);
  if (m > st.tau && !NXF_K1_DBG(debug, 16)) {
    const float tau = st.tau;
    // entry index in the candidate arena (uniform base + 32-bit index: one IMAD.WIDE per address); score and row go out
    // as two 32-bit stores so that no register pairing is forced on the accumulator values of the common path
    uint32_t w = my_first + static_cast<uint32_t>(st.cnt);
    if (est_only) {
      // estimate-only pass: the slice maximum stands for the slice (its rank statistics bound the scores' from below)
      cand_words[2 * static_cast<uint64_t>(w)] = __float_as_uint(m);
      cand_words[2 * static_cast<uint64_t>(w) + 1] = static_cast<uint32_t>(cbase);
      st.cnt += 1;
      return;
    }
#pragma unroll
    for (int q4 = 0; q4 < 4; ++q4) {
      if (g[q4] > tau) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float sc = v[8 * q4 + j];
          if (sc > tau) {
            cand_words[2 * static_cast<uint64_t>(w)] = __float_as_uint(sc);
            cand_words[2 * static_cast<uint64_t>(w) + 1] = static_cast<uint32_t>(cbase + 8 * q4 + j);
            ++w;
          }
        }
      }
    }
    st.cnt = static_cast<int>(w - my_first);
  }
}
// (the bound of these thread-private appends -- a row gains <= 256 entries per tile and is pruned once fewer than 256 slots are
// left -- is asserted after every tile in the kernel: NXF_DEV_CHECK(cur.cnt <= cap))

// Round schedule shared by the three warp roles.  Line positions [a, b) of cluster j in a round of R tiles cover
// query tiles qt = pos / R at corpus offsets pos % R.
struct RoundPiece {
  int64_t R, a, b;
};
__device__ __forceinline__ RoundPiece round_piece(const K1Params& p, int64_t r0, int64_t cluster_id, int64_t num_clusters) {
  RoundPiece rp;
  rp.R = min(static_cast<int64_t>(p.round_tiles), p.tile_end - r0);
  rp.a = cluster_id * p.num_qt * rp.R / num_clusters;
  rp.b = (cluster_id + 1) * p.num_qt * rp.R / num_clusters;
  return rp;
}

// SETS = epilogue warp sets.  One set (four warps, one per SM sub-partition) filters a 256 x 256 tile in ~3 600 clocks, bound by
// dependent-issue latency; the MMAs of a tile take 512 clocks per K block, so from 7 K blocks on (D >= 448) one set hides behind the
// tensor pipe.  Shorter rows get a second set on the other TMEM buffer (set s filters the tiles with iter % 2 == s), with its own
// per-row thresholds, counters and candidate buffers -- to the rest of the search it is one more owner of the query.  Rows of one or
// two K blocks get four groups on split accumulators (SETS = 4 with SPLIT, below): 640 threads, 96 registers each.
// SPLIT (rows of one or two K blocks, two sets): each 256-column accumulator is computed and handed over as two 128-column halves
// with their own full / empty barriers (two N = 128 MMAs per K step instead of one N = 256).  With one buffer per set, a set that
// hands its buffer back then waits for the whole MMA of its next tile (19 % of the epilogue's samples on 64-d rows); with halves, the
// MMA of the next tile's first half runs while the set is still filtering this tile's second half.  Under cta_group::2 each CTA holds
// half of the tile's corpus rows, so half h of the accumulator is corpus rows [64 h, 64 h + 64) of BOTH CTAs' slices: columns
// 0-63 = rows 64 h .. of the tile, columns 64-127 = rows 128 + 64 h .. (cta_group::1: rows 128 h .. 128 h + 127).
template <int CG, bool BF16, int SETS, bool SPLIT>
__global__ void __launch_bounds__(128 + 128 * SETS, 1)
k1_search_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c,
                 const __grid_constant__ CUtensorMap tmap_c2, K1Params p) {
  using namespace k1;
  using C = Cfg<CG>;
  extern __shared__ unsigned char smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array itself (an integer round trip would make every later access generic)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::RING_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 4;
  uint32_t* tmem_ptr_s = reinterpret_cast<uint32_t*>(tempty + 4);

  if (p.run_flag != nullptr && ld_relaxed_u32(p.run_flag) == 0u) return;  // uniform over the grid: written by an earlier kernel
  const int warp = __shfl_sync(kFullMask, threadIdx.x >> 5, 0);
  const unsigned lane = lane_id();
  const uint32_t rank = (CG == 2) ? cluster_ctarank() : 0u;
  const int64_t cluster_id = blockIdx.x / CG;
  const int64_t num_clusters = gridDim.x / CG;

  if (CG == 2) cluster_sync_all();  // both CTAs of the pair are resident before cta_group::2 alloc
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_c);
    if (p.seg_mode == 1) tma_prefetch_desc(&tmap_c2);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full[s], CG); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 4; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], CG * EPI_WARPS); }
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
    // The whole warp runs the loops and waits on the barriers; one elected lane issues.  With the loop inside `if (lane == 0)` nvcc
    // wraps every TMA / tcgen05 instruction (uniform-datapath operands in a divergent region) in an elect-and-retry loop of five
    // dependent instructions; under elect.sync it emits them back to back -- on 128-byte rows, where one K block is a whole tile,
    // that chain paced the stream (profiles/r2_k2t_knockout.md).
    {
      int s = 0;
      uint32_t ph = 0;
      int64_t ri = 0;
      for (int64_t r0 = p.tile_begin; r0 < p.tile_end; r0 += p.round_tiles, ++ri) {
        if (p.sync_window > 0 && rank == 0 && ri >= p.sync_window) {
          // soft round barrier: wait until (on average) every cluster finished issuing round ri - window; bounded spin, so a
          // cluster that is not resident (foreign kernels on some SMs) costs lockstep, never progress
          const uint32_t want = static_cast<uint32_t>(num_clusters * (ri - p.sync_window + 1));
          for (int spin = 0; spin < 20000 && __shfl_sync(kFullMask, ld_relaxed_u32(p.sync_ctr), 0) < want; ++spin) __nanosleep(100);
        }
        const RoundPiece rp = round_piece(p, r0, cluster_id, num_clusters);
        for (int64_t pos = rp.a; pos < rp.b;) {
          const int64_t qt = pos / rp.R;
          const int64_t lo = pos - qt * rp.R, hi = min(rp.b - qt * rp.R, rp.R);
          const int32_t q_row0 = static_cast<int32_t>((qt * CG + rank) * BLOCK_M);
          for (int64_t t = r0 + lo; t < r0 + hi; ++t) {
            const int32_t c_row0 = static_cast<int32_t>(perm_tile(p.perm, t) * BLOCK_N + rank * C::LOAD_N);
            for (int kb = 0; kb < p.kblocks; ++kb) {
              // column (in elements) of this K block in the query / corpus rows.  Segmented search (K1Params::seg_mode): the query
              // planes are [hi | mid] in one row; mode 1 walks (q_hi, c_hi), (q_hi, c_mid), (q_mid, c_hi) with the corpus mid plane behind
              // its own tensor map, mode 2 walks (q_hi, c), (q_mid, c)
              int32_t q_col = kb * BLOCK_K, c_col = kb * BLOCK_K;
              const CUtensorMap* cm = &tmap_c;
              if (p.split_kb > 0) {
                const int seg = kb / p.split_kb, w = kb - seg * p.split_kb;
                c_col = w * BLOCK_K;
                if (p.seg_mode == 1) {
                  q_col = ((seg == 2 ? p.split_kb : 0) + w) * BLOCK_K;
                  if (seg == 1) cm = &tmap_c2;
                } else {
                  q_col = (seg * p.split_kb + w) * BLOCK_K;
                }
              }
              mbar_wait(&empty[s], ph ^ 1);
              if (elect_one()) {
                if (NXF_K1_DBG(p.debug, 4)) {  // pipeline timing only: no loads, the MMAs run on whatever the ring holds
                  if (CG == 1 || rank == 0) mbar_arrive(&full[s]); else mbar_arrive_cluster(&full[s], 0);
                } else if (CG == 1) {
                  mbar_expect_tx(&full[s], C::STAGE_BYTES);
                  tma_load_2d(&tmap_q, &full[s], a_ptr(s), q_col, q_row0, kEvictLast);
                  tma_load_2d(cm, &full[s], b_ptr(s), c_col, c_row0, kEvictNormal);
                } else {
                  tma_load_2d_2sm(&tmap_q, &full[s], a_ptr(s), q_col, q_row0, kEvictLast);
                  tma_load_2d_2sm(cm, &full[s], b_ptr(s), c_col, c_row0, kEvictNormal);
                  if (rank == 0) mbar_expect_tx(&full[s], 2 * C::STAGE_BYTES);
                  else mbar_arrive_cluster(&full[s], 0);
                }
              }
              __syncwarp();
              if (++s == C::STAGES) { s = 0; ph ^= 1; }
            }
          }
          pos = qt * rp.R + hi;
        }
        if (p.sync_window > 0 && rank == 0 && lane == 0) red_relaxed_add_u32(p.sync_ctr, 1u);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    if (rank == 0) {   // all 32 lanes: warp-uniform loops, one elected lane issues (see the producer)
      constexpr uint32_t idesc = umma_idesc_f16(BF16, BLOCK_M * CG, BLOCK_N);
      constexpr uint32_t idesc_half = umma_idesc_f16(BF16, BLOCK_M * CG, BLOCK_N / 2);
      int s = 0;
      uint32_t ph = 0;
      uint32_t iter = 0;
      for (int64_t r0 = p.tile_begin; r0 < p.tile_end; r0 += p.round_tiles) {
        const RoundPiece rp = round_piece(p, r0, cluster_id, num_clusters);
        for (int64_t i = rp.a; i < rp.b; ++i, ++iter) {  // one accumulator tile per line position, whatever its query tile
          const uint32_t as = iter & 1, aph = (iter >> 1) & 1;
          if constexpr (SPLIT) {
            // kblocks <= 2 <= STAGES (host): all K blocks of the tile are resident at once, so the ring positions are walked twice,
            // once per half; the stages go back to the producer with the second half's MMAs
            const int s0 = s;
            const uint32_t ph0 = ph;
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
              mbar_wait(&tempty[as * 2 + h], aph ^ 1);
              tc_fence_after();
              const uint32_t d_tmem = tmem_base + as * BLOCK_N + h * (BLOCK_N / 2);
              s = s0; ph = ph0;
              for (int kb = 0; kb < p.kblocks; ++kb) {
                mbar_wait(&full[s], ph);
                tc_fence_after();
                const uint64_t a_desc = umma_desc_sw128(smem_u32(a_ptr(s)));
                // this CTA's rows of the half: LOAD_N / 2 rows of 128 bytes further into its B tile
                const uint64_t b_desc = umma_desc_sw128(smem_u32(b_ptr(s))) + static_cast<uint32_t>(h) * ((C::LOAD_N / 2) * 128 >> 4);
                if (elect_one()) {
#pragma unroll
                  for (int k4 = 0; k4 < BLOCK_K / UMMA_K; ++k4)
                    umma_f16<CG>(d_tmem, a_desc + 2 * k4, b_desc + 2 * k4, idesc_half, (kb | k4) != 0 ? 1u : 0u);
                  if (h == 1) umma_commit<CG>(&empty[s]);
                  if (kb == p.kblocks - 1) umma_commit<CG>(&tfull[as * 2 + h]);
                }
                __syncwarp();
                if (++s == C::STAGES) { s = 0; ph ^= 1; }
              }
            }
            continue;
          }
          mbar_wait(&tempty[as], aph ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BLOCK_N;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&full[s], ph);
            tc_fence_after();
            const uint64_t a_desc = umma_desc_sw128(smem_u32(a_ptr(s)));
            const uint64_t b_desc = umma_desc_sw128(smem_u32(b_ptr(s)));
            if (elect_one()) {
              if (!NXF_K1_DBG(p.debug, 8)) {
#pragma unroll
                for (int k4 = 0; k4 < BLOCK_K / UMMA_K; ++k4) {
                  // advance 16 elements = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
                  umma_f16<CG>(d_tmem, a_desc + 2 * k4, b_desc + 2 * k4, idesc, (kb | k4) != 0 ? 1u : 0u);
                }
              }
              umma_commit<CG>(&empty[s]);  // smem stage reusable once these MMAs retire
              if (kb == p.kblocks - 1) umma_commit<CG>(&tfull[as]);   // accumulator complete -> epilogue
            }
            __syncwarp();
            if (++s == C::STAGES) { s = 0; ph ^= 1; }
          }
        }
      }
      // do not tear the barriers down while the last remote arrivals may still be in flight
      if (iter > 0) {
        const uint32_t last = iter - 1;
        if (SPLIT) { mbar_wait(&tempty[(last & 1) * 2], (last >> 1) & 1); mbar_wait(&tempty[(last & 1) * 2 + 1], (last >> 1) & 1); }
        else mbar_wait(&tempty[last & 1], (last >> 1) & 1);
        if (iter > 1) {
          const uint32_t prev = iter - 2;
          if (SPLIT) { mbar_wait(&tempty[(prev & 1) * 2], (prev >> 1) & 1); mbar_wait(&tempty[(prev & 1) * 2 + 1], (prev >> 1) & 1); }
          else mbar_wait(&tempty[prev & 1], (prev >> 1) & 1);
        }
      }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0) {
    // ============================== epilogue ==============================
    const int ew = (warp - EPI_WARP0) & 3;  // == warp % 4: the TMEM lane quarter this warp may read
    // grp = owner index of this warp's rows (its own thresholds, counters, candidate buffers).  SETS = 4 (SPLIT only): the two groups of
    // a buffer take one 128-column half each -- group (set, half) waits for, reads and hands back exactly that half.
    static_assert(SETS != 4 || SPLIT, "four groups need the split accumulators");
    const uint32_t grp = static_cast<uint32_t>(warp - EPI_WARP0) >> 2;  // 0 .. SETS-1
    const uint32_t set = SETS == 4 ? (grp & 1u) : grp;                   // the accumulator buffer (tile parity) it filters
    const uint32_t my_half = SETS == 4 ? (grp >> 1) : 0u;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
    const int cap = p.sink.cap;
    const int k = p.sink.k;
    const int32_t n_rows = static_cast<int32_t>(p.n);
    uint32_t* cand_words = reinterpret_cast<uint32_t*>(p.sink.cand);
    // this cluster's (at most two) query tiles: slot 0 = qt_base, slot 1 = qt_base + 1
    const int64_t qt_base = cluster_id * p.num_qt / num_clusters;
    auto slot_buf0 = [&](int slot) {  // lane 0's candidate buffer of this warp for that slot
      return p.sink.cand + (((static_cast<int64_t>(blockIdx.x) * 2 + slot) * SETS + grp) * BLOCK_M + ew * 32) * cap;
    };
    auto slot_qrow0 = [&](int slot) { return ((qt_base + slot) * CG + rank) * BLOCK_M + ew * 32; };
    auto slot_valid = [&](int slot) { return qt_base + slot < p.num_qt && slot_qrow0(slot) + lane < p.nq; };
    EpiState cur, alt;  // per-row state of the active slot / the other slot (swapped when the piece crosses a query tile)
    cur.cnt = 0; alt.cnt = 0;
    cur.tau = slot_valid(0) ? tau_load(p.sink.tau_g, slot_qrow0(0) + lane) : INFINITY;  // padded rows accept nothing
    alt.tau = slot_valid(1) ? tau_load(p.sink.tau_g, slot_qrow0(1) + lane) : INFINITY;
    int cur_slot = 0;
    uint32_t iter = 0;

    for (int64_t r0 = p.tile_begin; r0 < p.tile_end; r0 += p.round_tiles) {
      const RoundPiece rp = round_piece(p, r0, cluster_id, num_clusters);
      for (int64_t pos = rp.a; pos < rp.b;) {
        const int64_t qt = pos / rp.R;
        const int64_t lo = pos - qt * rp.R, hi = min(rp.b - qt * rp.R, rp.R);
        pos = qt * rp.R + hi;
        const int slot = static_cast<int>(qt - qt_base);
        if (slot != cur_slot) {
          const EpiState tmp = cur; cur = alt; alt = tmp;
          cur_slot = slot;
        }
        Entry* warp_buf0 = slot_buf0(slot);
        const uint32_t my_first = static_cast<uint32_t>((warp_buf0 - p.sink.cand) + static_cast<int64_t>(lane) * cap);
        const int64_t qrow = slot_qrow0(slot) + lane;
        const bool valid = qrow < p.nq;
        uint32_t tau_next_key = valid ? tau_load_key(p.sink.tau_g, qrow) : 0u;   // raw key: converted where it is used (sink.cuh)

        for (int64_t t = r0 + lo; t < r0 + hi; ++t, ++iter) {
          if (SETS >= 2 && (iter & 1u) != set) continue;  // the other set's accumulator buffer
          const uint32_t as = iter & 1, aph = (iter >> 1) & 1;
          // SPLIT: barrier pair of this buffer's halves; the second half is waited for just before its first load
          uint64_t* const tf = SPLIT ? &tfull[as * 2 + my_half] : &tfull[as];
          uint64_t* const te = SPLIT ? &tempty[as * 2 + my_half] : &tempty[as];
          mbar_wait(tf, aph);
          tc_fence_after();
          const int32_t col0 = static_cast<int32_t>(perm_tile(p.perm, t) * BLOCK_N);
          // corpus row (within the tile) of the first column of 32-column slice ch -- see SPLIT above the kernel
          auto slice_row = [&](int ch) -> int32_t {
            if (!SPLIT || CG == 1) return ch * 32;
            const int h = ch >> 2, j = ch & 3;
            return (j < 2) ? 64 * h + 32 * j : 128 + 64 * h + 32 * (j - 2);
          };
          auto hand_back = [&](uint64_t* bar) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2 && rank != 0) mbar_arrive_cluster(bar, 0); else mbar_arrive(bar); }
          };
          // slow-path epilogue: the last corpus tile (rows past n are TMA zero fill) or the L2 metric
          const bool adj = (col0 + BLOCK_N > n_rows) || (p.half_norm != nullptr);
          const uint32_t tbase = taddr + as * BLOCK_N;
          if (NXF_K1_DBG(p.debug, 2)) {
            // pipeline timing only: hand the accumulator back without reading it
            hand_back(te);
            if (SPLIT && SETS != 4) { mbar_wait(tf + 1, aph); tc_fence_after(); hand_back(te + 1); }
            continue;
          }
          // Phase A (holds the accumulator): two register buffers, the tcgen05.ld of slice ch+1 is in flight while slice
          // ch is filtered
          uint32_t va[32];
          if constexpr (SETS == 4) {
            // 640 threads: 102 registers each leave no room for two slices in flight; twenty resident warps hide the load instead
            const int c0 = static_cast<int>(my_half) * (BLOCK_N / 64);
#pragma unroll 1
            for (int ch = c0; ch < c0 + BLOCK_N / 64; ++ch) {
              tmem_ld32(tbase + ch * 32, va);
              tmem_ld_wait_dep(va);
              if (ch + 1 == c0 + BLOCK_N / 64) hand_back(te);
              if (adj) epi_chunk<true>(va, col0 + slice_row(ch), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
              else epi_chunk<false>(va, col0 + slice_row(ch), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
            }
          } else {
          uint32_t vb[32];
          tmem_ld32(tbase, va);
#pragma unroll 1
          for (int ch = 0; ch < BLOCK_N / 32; ch += 2) {
            tmem_ld_wait_dep(va);
            tmem_ld32(tbase + (ch + 1) * 32, vb);
            if (!NXF_K1_DBG(p.debug, 1)) {
              if (adj) epi_chunk<true>(va, col0 + slice_row(ch), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
              else epi_chunk<false>(va, col0 + slice_row(ch), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
            }
            tmem_ld_wait_dep(vb);
            if (SPLIT && ch + 2 == BLOCK_N / 64) {
              // first half fully read: back to the issuer (which starts the next tile's first half at once); second half ready?
              hand_back(te);
              mbar_wait(tf + 1, aph);
              tc_fence_after();
            }
            if (ch + 2 < BLOCK_N / 32) {
              tmem_ld32(tbase + (ch + 2) * 32, va);
            } else {
              // accumulator fully read: hand the TMEM buffer (SPLIT: its second half) back to the MMA issuer
              hand_back(SPLIT ? te + 1 : te);
            }
            if (!NXF_K1_DBG(p.debug, 1)) {
              if (adj) epi_chunk<true>(vb, col0 + slice_row(ch + 1), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
              else epi_chunk<false>(vb, col0 + slice_row(ch + 1), n_rows, p.half_norm, cur, cand_words, my_first, p.debug, p.est_only);
            }
          }
          }  // SETS != 4
          NXF_DEV_CHECK(cur.cnt <= cap, "K1 private candidate buffer overflow");
          // other owners' progress on this query: loaded during the previous tile, folded in only now so that its latency
          // hides behind phase A
          cur.tau = fmaxf(cur.tau, tau_from_key(tau_next_key));
          // Phase B (the accumulator is already back with the MMA issuer): rows whose buffer could overflow during the next
          // tile (<= 256 appends) are reduced to their exact top-k' by the whole warp; tau rises to the k'-th score and is
          // published for the other owners of the query
          __syncwarp();  // this tile's appends (thread-private global stores) are visible to the warp
          unsigned need = __ballot_sync(kFullMask, cur.cnt > cap - BLOCK_N);
          const bool prof = NXF_K1_DBG(p.debug, 32) && need;  // timing experiment: prune count / cycles per warp into the counter words
          const long long pc0 = prof ? clock64() : 0;
          const int pn = __popc(need);
          if (need && !NXF_K1_DBG(p.debug, 64) && cur.cnt > cap - BLOCK_N) {
            // the buffers were written tiles ago and may have left L2: every needy row pulls its own lines in now, so the serial
            // warp-cooperative prunes below find them there
            const char* b = reinterpret_cast<const char*>(warp_buf0 + static_cast<int64_t>(lane) * cap);
            for (int o = 0; o < cur.cnt * 8; o += 128) prefetch_l2(b + o);
          }
          while (need) {
            const int src = __ffs(need) - 1;
            need &= need - 1;
            const int n_src = __shfl_sync(kFullMask, cur.cnt, src);
            const uint32_t kth = warp_prune(warp_buf0 + static_cast<int64_t>(src) * cap, n_src, k);
            if (static_cast<int>(lane) == src) {
              cur.cnt = k;
              cur.tau = fmaxf(cur.tau, okey_inv(kth));
              if (valid) atomicMax(&p.sink.tau_g[qrow], kth);
            }
            __syncwarp();
          }
          if (prof && lane == 0) {
            atomicAdd(reinterpret_cast<unsigned long long*>(p.prof), static_cast<unsigned long long>(pn));
            atomicAdd(reinterpret_cast<unsigned long long*>(p.prof) + 1, static_cast<unsigned long long>(clock64() - pc0));
          }
          if (valid) tau_next_key = tau_load_key(p.sink.tau_g, qrow);
        }
      }
    }

    // ---- end of the launch (pass): every row's candidates (its sink.k best if it holds more) go to the query's arena slice ----
    __syncwarp();
    const long long mc0 = NXF_K1_DBG(p.debug, 32) ? clock64() : 0;
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
      if (qt_base + slot >= p.num_qt) break;
      int my_cnt = (slot == cur_slot) ? cur.cnt : alt.cnt;
      Entry* warp_buf0 = slot_buf0(slot);
      const int64_t my_q = slot_qrow0(slot) + lane;
      if (my_q >= p.nq) my_cnt = 0;
      // Everything a row holds goes out (<= cap entries; the slice is sized for it, api.cu k1_arena_stride).  Round 1 reduced rows
      // holding more than sink.k entries to their k best first -- a serial warp-cooperative select per row, up to 64 of them per
      // warp: ~0.13 ms at the end of EVERY pass, sample passes included (their rows all hold more than the sample rank m), which
      // was most of what a 1/8 shard paid on top of its GEMM.  select_tau_kernel / merge_arena_kernel select over the union anyway.
      // one atomic per row, all 32 in flight together; then the warp copies the rows one after the other
      uint32_t my_pos = 0;
      if (my_cnt > 0) my_pos = atomicAdd(&p.sink.qcnt[my_q], static_cast<uint32_t>(my_cnt));
      NXF_DEV_CHECK(static_cast<int64_t>(my_pos) + my_cnt <= p.sink.arena_stride, "K1 arena slice overflow");
      if (static_cast<int64_t>(my_pos) + my_cnt > p.sink.arena_stride) my_cnt = 0;  // cannot happen by construction; never write OOB
      unsigned todo = __ballot_sync(kFullMask, my_cnt > 0);
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int n = __shfl_sync(kFullMask, my_cnt, src);
        const uint32_t pos = __shfl_sync(kFullMask, my_pos, src);
        const Entry* buf = warp_buf0 + static_cast<int64_t>(src) * cap;
        Entry* dst = p.sink.arena + (slot_qrow0(slot) + src) * p.sink.arena_stride + pos;
        for (int i = lane; i < n; i += 32) dst[i] = buf[i];
      }
    }
    if (NXF_K1_DBG(p.debug, 32) && lane == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof) + 2, static_cast<unsigned long long>(clock64() - mc0));
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof) + 3, 1ull);
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
// The same few maps are asked for again and again (five K1 launches per search over one corpus and one query buffer; the serve
// path repeats the identical search shape per request): a small cache keyed by everything the encoding depends on.
namespace {
struct TmapSlot { const void* base; int64_t rows, ld; int box_rows; bool bf16, valid; CUtensorMap map; };
std::mutex g_tmap_mu;
TmapSlot g_tmap_cache[16];
unsigned g_tmap_next = 0;
}  // namespace

static int encode_tmap_2d_sw128(CUtensorMap* m, const void* base, int64_t rows, int64_t ld, int box_rows, bool bf16);

int make_tmap_2d_sw128(CUtensorMap* m, const void* base, int64_t rows, int64_t ld, int box_rows, bool bf16) {
  std::lock_guard<std::mutex> lk(g_tmap_mu);
  for (const TmapSlot& s : g_tmap_cache)
    if (s.valid && s.base == base && s.rows == rows && s.ld == ld && s.box_rows == box_rows && s.bf16 == bf16) { *m = s.map; return NXF_OK; }
  const int rc = encode_tmap_2d_sw128(m, base, rows, ld, box_rows, bf16);
  if (rc) return rc;
  TmapSlot& s = g_tmap_cache[g_tmap_next++ % 16];
  s.base = base; s.rows = rows; s.ld = ld; s.box_rows = box_rows; s.bf16 = bf16; s.map = *m; s.valid = true;
  return NXF_OK;
}

static int encode_tmap_2d_sw128(CUtensorMap* m, const void* base, int64_t rows, int64_t ld, int box_rows, bool bf16) {
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

template <int CG, bool BF16, int SETS, bool SPLIT = false>
static int launch_k1_t(const CUtensorMap& tq, const CUtensorMap& tc, const CUtensorMap& tc2, const K1Params& p, int grid, cudaStream_t stream) {
  using C = k1::Cfg<CG>;
  auto kern = k1_search_kernel<CG, BF16, SETS, SPLIT>;
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), C::SMEM_BYTES)) return rc;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(128 + 128 * SETS);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  NXF_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, tc, tc2, p));
  return NXF_OK;
}

// q16: [nq_rows, q_ld] 16-bit queries (device), corpus (and corpus2, the mid plane of a mode-1 segmented search): [n, ld] 16-bit rows
int launch_k1(int store_dtype, int cta_group, const void* q16, int64_t nq_rows, int64_t q_ld, const void* corpus, const void* corpus2,
              int64_t ld, const K1Params& p, int grid, cudaStream_t stream) {
  const bool bf16 = store_dtype == NXF_BF16;
  if (store_dtype != NXF_BF16 && store_dtype != NXF_F16) { set_error("tensor-core search needs f16/bf16 storage"); return NXF_EUNSUPPORTED; }
  if (p.seg_mode == 1 && corpus2 == nullptr) { set_error("launch_k1: segmented search over fp32 storage needs the mid plane"); return NXF_EINVAL; }
  CUtensorMap tq, tc, tc2;
  int rc = make_tmap_2d_sw128(&tq, q16, nq_rows, q_ld, k1::BLOCK_M, bf16);
  if (rc) return rc;
  rc = make_tmap_2d_sw128(&tc, corpus, p.n, ld, k1::BLOCK_N / cta_group, bf16);
  if (rc) return rc;
  if (p.seg_mode == 1) {
    rc = make_tmap_2d_sw128(&tc2, corpus2, p.n, ld, k1::BLOCK_N / cta_group, bf16);
    if (rc) return rc;
  } else {
    tc2 = tc;
  }
  if (p.split) {
    if ((p.epi_sets != 2 && p.epi_sets != 4) || p.kblocks > 2) { set_error("k1: split accumulators need two or four epilogue sets and <= 2 K blocks per tile"); return NXF_EINVAL; }
    if (p.epi_sets == 4) {
      if (cta_group == 2) return bf16 ? launch_k1_t<2, true, 4, true>(tq, tc, tc2, p, grid, stream) : launch_k1_t<2, false, 4, true>(tq, tc, tc2, p, grid, stream);
      return bf16 ? launch_k1_t<1, true, 4, true>(tq, tc, tc2, p, grid, stream) : launch_k1_t<1, false, 4, true>(tq, tc, tc2, p, grid, stream);
    }
    if (cta_group == 2) return bf16 ? launch_k1_t<2, true, 2, true>(tq, tc, tc2, p, grid, stream) : launch_k1_t<2, false, 2, true>(tq, tc, tc2, p, grid, stream);
    return bf16 ? launch_k1_t<1, true, 2, true>(tq, tc, tc2, p, grid, stream) : launch_k1_t<1, false, 2, true>(tq, tc, tc2, p, grid, stream);
  }
  if (p.epi_sets == 2) {
    if (cta_group == 2) return bf16 ? launch_k1_t<2, true, 2>(tq, tc, tc2, p, grid, stream) : launch_k1_t<2, false, 2>(tq, tc, tc2, p, grid, stream);
    return bf16 ? launch_k1_t<1, true, 2>(tq, tc, tc2, p, grid, stream) : launch_k1_t<1, false, 2>(tq, tc, tc2, p, grid, stream);
  }
  if (cta_group == 2) return bf16 ? launch_k1_t<2, true, 1>(tq, tc, tc2, p, grid, stream) : launch_k1_t<2, false, 1>(tq, tc, tc2, p, grid, stream);
  return bf16 ? launch_k1_t<1, true, 1>(tq, tc, tc2, p, grid, stream) : launch_k1_t<1, false, 1>(tq, tc, tc2, p, grid, stream);
}

int k1_smem_bytes(int cta_group) { return cta_group == 2 ? k1::Cfg<2>::SMEM_BYTES : k1::Cfg<1>::SMEM_BYTES; }

}  // namespace nxf
