// api.cu -- the C ABI of libnexus_flat.so (see include/nexus_flat.h): index handle, add / convert /
// shard upload, search dispatch (K2 stream | K1 tcgen05 | KX exact), merge, pool+normalise.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <unordered_map>
#include <thread>
#include <vector>

#include "params.cuh"

namespace nxf {

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d in `%s`", static_cast<int>(e), cudaGetErrorString(e), file, line, what);
  if (e == cudaErrorMemoryAllocation) return NXF_ENOMEM;
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver || e == cudaErrorInvalidDevice) return NXF_ENODEVICE;
  return NXF_ECUDA;
}

int ensure_dyn_smem(const void* kernel, int bytes) {
  struct Lim { int v[32]; };
  static std::mutex mu;
  static std::unordered_map<const void*, Lim> seen;
  int dev = 0;
  NXF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 32) { NXF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); return NXF_OK; }
  std::lock_guard<std::mutex> lk(mu);
  Lim& l = seen[kernel];  // value-initialised to zeros
  if (l.v[dev] >= bytes) return NXF_OK;
  NXF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  l.v[dev] = bytes;
  return NXF_OK;
}

}  // namespace nxf

using namespace nxf;

// ------------------------------------------------------------------------------------------
// the handle
// ------------------------------------------------------------------------------------------
struct nxf_index {
  int dim = 0, metric = 0, store_dtype = 0, device = 0;
  int64_t ld = 0;          // stored row stride in elements (dim rounded up to 64, zero padded)
  int64_t n = 0, capacity = 0;
  void* data = nullptr;    // [capacity, ld] of store_dtype
  float* half_norm = nullptr;  // [capacity] 0.5*|row|^2 (L2 only)
  // fp32 storage only: the bf16 operand planes of the tensor-core path (DESIGN.md "certified search"), built on first use
  void* hi = nullptr;          // [hi_cap + 256, ld] bf16(x): the candidate pass streams this plane alone (2 bytes per element)
  void* mid = nullptr;         // [mid_cap + 256, ld] bf16(x - hi): built only when a query fails its certificate (retry level)
  int64_t hi_cap = 0, hi_rows = 0, mid_cap = 0, mid_rows = 0;  // rows allocated / rows converted so far
  bool hi_failed = false, mid_failed = false;  // the plane did not fit in device memory (until the storage changes)
  uint32_t* max_norm2 = nullptr;  // device word: float bits of max |stored row|^2, raised by every add (the certificate's |c| bound)
  int64_t id_offset = 0;
  int sm_count = 0;
  int cc_major = 0;
  // scratch arena (grown on demand, reused across searches)
  unsigned char* scratch = nullptr;
  size_t scratch_bytes = 0;
  // pinned staging for host <-> device traffic
  unsigned char* pinned = nullptr;
  size_t pinned_bytes = 0;
  nxf_stats stats{};
  bool profiling = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;  // pending (start, stop) pairs
  // Stream ordering between calls.  The scratch arena, the staging buffers and (add / reconstruct) the row storage are shared by
  // every call on the handle, and device-in / device-out calls return before their kernels ran: each call records `last_use` on its
  // stream when it has enqueued its work, and a call on a DIFFERENT stream first makes that stream wait for it (UseScope below).
  cudaEvent_t last_use = nullptr;
  cudaStream_t last_stream = nullptr;
  bool used = false;
  std::mutex mu;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = (cudaSetDevice(dev) == cudaSuccess);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// RAII: on entry the call's stream waits for the previous call on this handle if that ran on another stream; on exit (every
// return path) the event is re-recorded behind whatever this call enqueued.  Constructed under idx->mu with the device current.
struct UseScope {
  nxf_index* idx;
  cudaStream_t stream;
  UseScope(nxf_index* i, cudaStream_t s) : idx(i), stream(s) {
    if (idx->used && idx->last_stream != stream && idx->last_use) cudaStreamWaitEvent(stream, idx->last_use, 0);
  }
  ~UseScope() {
    if (!idx->last_use && cudaEventCreateWithFlags(&idx->last_use, cudaEventDisableTiming) != cudaSuccess) { idx->last_use = nullptr; return; }
    if (cudaEventRecord(idx->last_use, stream) == cudaSuccess) { idx->last_stream = stream; idx->used = true; }
  }
};

int check_dtype(int dt, const char* what) {
  if (dt != NXF_F32 && dt != NXF_F16 && dt != NXF_BF16) {
    set_error("%s: unknown dtype %d (NXF_F32=0, NXF_F16=1, NXF_BF16=2)", what, dt);
    return NXF_EINVAL;
  }
  return NXF_OK;
}

int ensure_scratch(nxf_index* idx, size_t bytes) {
  if (bytes <= idx->scratch_bytes) return NXF_OK;
  if (idx->scratch) { cudaFree(idx->scratch); idx->scratch = nullptr; idx->scratch_bytes = 0; }
  const size_t want = (bytes + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
  NXF_CUDA(cudaMalloc(reinterpret_cast<void**>(&idx->scratch), want));
  idx->scratch_bytes = want;
  idx->stats.scratch_bytes = static_cast<int64_t>(want);
  return NXF_OK;
}

int ensure_pinned(nxf_index* idx, size_t bytes) {
  if (bytes <= idx->pinned_bytes) return NXF_OK;
  if (idx->pinned) { cudaFreeHost(idx->pinned); idx->pinned = nullptr; idx->pinned_bytes = 0; }
  NXF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&idx->pinned), bytes));
  idx->pinned_bytes = bytes;
  return NXF_OK;
}

int ensure_capacity(nxf_index* idx, int64_t want, cudaStream_t stream) {
  if (want <= idx->capacity) return NXF_OK;
  int64_t cap = std::max<int64_t>(want, idx->capacity + idx->capacity / 2);
  cap = round_up64(cap, 256);
  const size_t esz = dtype_size(idx->store_dtype);
  void* nd = nullptr;
  // +256 rows of slack so that tile-granular TMA boxes never leave the allocation
  NXF_CUDA(cudaMalloc(&nd, static_cast<size_t>(cap + 256) * idx->ld * esz));
  float* nh = nullptr;
  if (idx->metric == NXF_METRIC_L2) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&nh), static_cast<size_t>(cap + 256) * sizeof(float));
    if (e != cudaSuccess) { cudaFree(nd); return cuda_fail(e, "cudaMalloc(half_norm)", __FILE__, __LINE__); }
  }
  if (idx->n > 0) {
    NXF_CUDA(cudaMemcpyAsync(nd, idx->data, static_cast<size_t>(idx->n) * idx->ld * esz, cudaMemcpyDeviceToDevice, stream));
    if (nh) NXF_CUDA(cudaMemcpyAsync(nh, idx->half_norm, static_cast<size_t>(idx->n) * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    NXF_CUDA(cudaStreamSynchronize(stream));
  }
  if (idx->data) cudaFree(idx->data);
  if (idx->half_norm) cudaFree(idx->half_norm);
  if (idx->hi) cudaFree(idx->hi);
  if (idx->mid) cudaFree(idx->mid);
  idx->hi = idx->mid = nullptr;
  idx->hi_cap = idx->hi_rows = idx->mid_cap = idx->mid_rows = 0;
  idx->hi_failed = idx->mid_failed = false;
  idx->data = nd;
  idx->half_norm = nh;
  idx->capacity = cap;
  idx->stats.storage_bytes = static_cast<int64_t>((cap + 256) * idx->ld * esz);
  return NXF_OK;
}

// fp32 storage: bring a bf16 operand plane (which = 0: hi = bf16(x); 1: mid = bf16(x - hi)) up to date with the stored rows.  Returns
// NXF_ENOMEM (and remembers it) when the plane does not fit; the caller then stays on / falls back to the FFMA kernels.
int ensure_plane(nxf_index* idx, int which, cudaStream_t stream) {
  void*& plane = which ? idx->mid : idx->hi;
  int64_t& cap = which ? idx->mid_cap : idx->hi_cap;
  int64_t& rows = which ? idx->mid_rows : idx->hi_rows;
  bool& failed = which ? idx->mid_failed : idx->hi_failed;
  if (failed) return NXF_ENOMEM;
  if (cap < idx->capacity) {
    if (plane) { cudaFree(plane); plane = nullptr; cap = 0; rows = 0; }
    const size_t bytes = static_cast<size_t>(idx->capacity + 256) * idx->ld * sizeof(uint16_t);
    cudaError_t e = cudaMalloc(&plane, bytes);
    if (e != cudaSuccess) {
      cudaGetLastError();
      plane = nullptr;
      failed = true;
      return NXF_ENOMEM;
    }
    cap = idx->capacity;
    idx->stats.storage_bytes += static_cast<int64_t>(bytes);
  }
  if (rows < idx->n) {
    const float* src = static_cast<const float*>(idx->data) + static_cast<size_t>(rows) * idx->ld;
    uint16_t* dst = static_cast<uint16_t*>(plane) + static_cast<size_t>(rows) * idx->ld;
    int rc = which ? launch_mid_rows(src, idx->ld, dst, idx->n - rows, stream)
                   : launch_convert_rows(src, NXF_F32, idx->ld, dst, NXF_BF16, idx->ld, idx->n - rows, static_cast<int>(idx->ld), stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
    rows = idx->n;
  }
  return NXF_OK;
}

// choose the number of corpus chunks so that (query tiles x chunks) fills `grid` CTAs evenly
void plan_units(int num_qt, int64_t n_tiles, int grid, int* chunks_out, int64_t* tpc_out) {
  if (n_tiles <= 0) { *chunks_out = 1; *tpc_out = 1; return; }
  const int64_t max_chunks = std::max<int64_t>(1, std::min<int64_t>(n_tiles, 4 * static_cast<int64_t>(grid)));
  int64_t best_c = 1;
  double best_eff = -1.0;
  for (int64_t c = 1; c <= max_chunks; ++c) {
    const int64_t tpc = ceil_div64(n_tiles, c);
    const int64_t cc = ceil_div64(n_tiles, tpc);
    const int64_t units = cc * num_qt;
    const int64_t waves = ceil_div64(units, grid);
    // cost model: every wave costs tpc tile-times (+1 for per-unit flush overhead)
    const double cost = static_cast<double>(waves) * static_cast<double>(tpc + 1);
    const double ideal = static_cast<double>(n_tiles) * num_qt / grid;
    const double eff = ideal / cost;
    if (eff > best_eff + 1e-9) { best_eff = eff; best_c = cc; }
    if (eff >= 0.985) break;
  }
  const int64_t tpc = ceil_div64(n_tiles, best_c);
  *tpc_out = tpc;
  *chunks_out = static_cast<int>(ceil_div64(n_tiles, tpc));
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Host rows -> pinned staging with several threads: one thread copies at 8-12 GB/s, a Gen5 x16 link moves ~55 GB/s, so the
// add of a host corpus (utils.py:169,189: N*D*4 bytes) is bound by this copy unless it is split (NXF_ADD_THREADS, default 8).
void parallel_memcpy(unsigned char* dst, const unsigned char* src, size_t bytes, int threads) {
  if (threads <= 1 || bytes < (size_t(8) << 20)) { std::memcpy(dst, src, bytes); return; }
  const size_t part = (bytes / threads + 4095) & ~size_t(4095);
  std::vector<std::thread> th;
  for (int t = 1; t < threads; ++t) {
    const size_t o = static_cast<size_t>(t) * part;
    if (o >= bytes) break;
    th.emplace_back([=] { std::memcpy(dst + o, src + o, std::min(part, bytes - o)); });
  }
  std::memcpy(dst, src, std::min(part, bytes));
  for (auto& x : th) x.join();
}

// K1's arena slice of a query (a tile's line segment is cut by pieces of num_qt / clusters of its length: owners_max clusters own
// a piece of one query tile)
int64_t k1_arena_stride(int64_t clusters, int64_t num_qt, int k_fetch, int epi_sets, int cap) {
  const int64_t owners_max = std::min<int64_t>(clusters, (clusters + num_qt - 1) / std::max<int64_t>(num_qt, 1) + 1);
  // every epilogue set of an owning cluster flushes whatever its row holds at the end of a pass (<= cap entries, no prune first),
  // on top of the list the earlier passes left (<= k')
  return owners_max * epi_sets * cap + k_fetch;
}

// Rank m in a sample of `s_tiles` of `n_tiles` corpus tiles whose score is, except with probability < e^-30 per query, exceeded by
// at least k rows of the whole corpus: the number of global top-k rows that fall into the sample is hypergeometric, bounded by
// the Poisson(k * s / n) upper tail P(X >= m) <= exp(-(m ln(m / lambda) - m + lambda)).
int sample_rank(int k, int64_t s_tiles, int64_t n_tiles) {
  const double lambda = static_cast<double>(k) * static_cast<double>(s_tiles) / static_cast<double>(std::max<int64_t>(n_tiles, 1));
  int m = static_cast<int>(lambda) + 1;
  while (m < k && m * std::log(m / lambda) - m + lambda < 30.0) ++m;
  return m;
}

int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  if (!v || !*v) return dflt;
  return std::atoi(v);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// library
// ------------------------------------------------------------------------------------------
extern "C" {

const char* nxf_version(void) { return "nexus_flat 0.1.0 (sm_100a; tcgen05/TMA)"; }
const char* nxf_last_error(void) { return g_err; }

int nxf_sample_rank(int k, int64_t sample_tiles, int64_t n_tiles) {
  if (k < 1 || sample_tiles < 1 || n_tiles < sample_tiles) return NXF_EINVAL;
  return std::min(k, sample_rank(k, sample_tiles, n_tiles));
}

int nxf_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__);
  }
  return n;
}

int nxf_create(int dim, int metric, int store_dtype, int device, nxf_index** out) {
  NXF_REQUIRE(out != nullptr, "nxf_create: out is NULL");
  *out = nullptr;
  NXF_REQUIRE(dim >= 1 && dim <= 16384, "nxf_create: dim %d out of range [1, 16384]", dim);
  NXF_REQUIRE(metric == NXF_METRIC_IP || metric == NXF_METRIC_L2, "nxf_create: unknown metric %d", metric);
  int rc = check_dtype(store_dtype, "nxf_create(store_dtype)");
  if (rc) return rc;
  int ndev = nxf_device_count();
  if (ndev < 0) return ndev;
  if (ndev == 0) { set_error("nxf_create: no CUDA device is visible; this engine has no CPU fallback"); return NXF_ENODEVICE; }
  NXF_REQUIRE(device >= 0 && device < ndev, "nxf_create: device %d out of range (visible devices: %d)", device, ndev);
  cudaDeviceProp prop;
  NXF_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("nxf_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
    return NXF_ENODEVICE;
  }
  nxf_index* idx = new (std::nothrow) nxf_index();
  if (!idx) { set_error("nxf_create: out of host memory"); return NXF_ENOMEM; }
  idx->dim = dim;
  idx->metric = metric;
  idx->store_dtype = store_dtype;
  idx->device = device;
  idx->ld = round_up64(dim, 64);
  idx->sm_count = prop.multiProcessorCount;
  idx->cc_major = prop.major;
  *out = idx;
  return NXF_OK;
}

void nxf_destroy(nxf_index* idx) {
  if (!idx) return;
  {
    DeviceGuard g(idx->device);
    if (idx->data) cudaFree(idx->data);
    if (idx->half_norm) cudaFree(idx->half_norm);
    if (idx->hi) cudaFree(idx->hi);
    if (idx->mid) cudaFree(idx->mid);
    if (idx->max_norm2) cudaFree(idx->max_norm2);
    if (idx->scratch) cudaFree(idx->scratch);
    if (idx->pinned) cudaFreeHost(idx->pinned);
    if (idx->last_use) cudaEventDestroy(idx->last_use);
    for (auto& pr : idx->prof_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  }
  delete idx;
}

int nxf_reset(nxf_index* idx) {
  NXF_REQUIRE(idx != nullptr, "nxf_reset: NULL index");
  std::lock_guard<std::mutex> lk(idx->mu);
  idx->n = 0;
  idx->hi_rows = idx->mid_rows = 0;   // max_norm2 stays: a stale (larger) bound only makes the certificate more conservative
  return NXF_OK;
}

int nxf_reserve(nxf_index* idx, int64_t n_total) {
  NXF_REQUIRE(idx != nullptr, "nxf_reserve: NULL index");
  NXF_REQUIRE(n_total >= 0 && n_total < (int64_t(1) << 31) - 512, "nxf_reserve: %lld rows exceed the per-device limit", (long long)n_total);
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  UseScope use(idx, nullptr);
  return ensure_capacity(idx, n_total, nullptr);
}

int64_t nxf_ntotal(const nxf_index* idx) { return idx ? idx->n : -1; }
int nxf_dim(const nxf_index* idx) { return idx ? idx->dim : -1; }
int nxf_metric(const nxf_index* idx) { return idx ? idx->metric : -1; }
int nxf_store_dtype(const nxf_index* idx) { return idx ? idx->store_dtype : -1; }
int nxf_device(const nxf_index* idx) { return idx ? idx->device : -1; }

int nxf_set_id_offset(nxf_index* idx, int64_t offset) {
  NXF_REQUIRE(idx != nullptr, "nxf_set_id_offset: NULL index");
  idx->id_offset = offset;
  return NXF_OK;
}

static void fold_profile_events(nxf_index* idx) {
  for (auto& pr : idx->prof_events) {
    float ms = 0.f;
    if (cudaEventSynchronize(pr.second) == cudaSuccess && cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) {
      idx->stats.main_kernel_ms += ms;
      idx->stats.main_kernel_count++;
    }
    cudaEventDestroy(pr.first);
    cudaEventDestroy(pr.second);
  }
  idx->prof_events.clear();
}

int nxf_set_profiling(nxf_index* idx, int enable) {
  NXF_REQUIRE(idx != nullptr, "nxf_set_profiling: NULL index");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  fold_profile_events(idx);
  idx->profiling = enable != 0;
  if (enable) { idx->stats.main_kernel_ms = 0.0; idx->stats.main_kernel_count = 0; }
  return NXF_OK;
}

int nxf_get_stats(const nxf_index* idx_c, nxf_stats* out) {
  nxf_index* idx = const_cast<nxf_index*>(idx_c);
  NXF_REQUIRE(idx != nullptr && out != nullptr, "nxf_get_stats: NULL argument");
  if (!idx->prof_events.empty()) {
    std::lock_guard<std::mutex> lk(idx->mu);
    DeviceGuard g(idx->device);
    fold_profile_events(idx);
  }
  *out = idx->stats;
  return NXF_OK;
}

// ------------------------------------------------------------------------------------------
// add
// ------------------------------------------------------------------------------------------
int nxf_add(nxf_index* idx, const void* rows, int64_t n, int src_dtype, int64_t src_ld, int src_on_device, void* stream_) {
  NXF_REQUIRE(idx != nullptr, "nxf_add: NULL index");
  NXF_REQUIRE(n >= 0, "nxf_add: negative row count");
  if (n == 0) return NXF_OK;
  NXF_REQUIRE(rows != nullptr, "nxf_add: rows is NULL");
  int rc = check_dtype(src_dtype, "nxf_add(src_dtype)");
  if (rc) return rc;
  if (src_ld == 0) src_ld = idx->dim;
  NXF_REQUIRE(src_ld >= idx->dim, "nxf_add: row stride %lld smaller than dim %d", (long long)src_ld, idx->dim);
  std::lock_guard<std::mutex> lk(idx->mu);
  NXF_REQUIRE(idx->n + n < (int64_t(1) << 31) - 512, "nxf_add: %lld rows exceed the per-device limit (shard across devices)",
              (long long)(idx->n + n));
  DeviceGuard g(idx->device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  UseScope use(idx, stream);
  rc = ensure_capacity(idx, idx->n + n, stream);
  if (rc) return rc;
  const size_t esz = dtype_size(idx->store_dtype);
  const size_t ssz = dtype_size(src_dtype);
  unsigned char* dst = static_cast<unsigned char*>(idx->data) + static_cast<size_t>(idx->n) * idx->ld * esz;

  if (src_on_device) {
    rc = launch_convert_rows(rows, src_dtype, src_ld, dst, idx->store_dtype, idx->ld, n, idx->dim, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
  } else {
    // host rows: double-buffered pinned staging -> async H2D -> convert/re-pitch on the device
    const size_t row_bytes = static_cast<size_t>(src_ld) * ssz;
    const size_t chunk_rows = std::max<size_t>(1, std::min<size_t>(static_cast<size_t>(n), (size_t(64) << 20) / row_bytes));
    const size_t chunk_bytes = chunk_rows * row_bytes;
    rc = ensure_pinned(idx, 2 * chunk_bytes);
    if (rc) return rc;
    rc = ensure_scratch(idx, 2 * chunk_bytes);
    if (rc) return rc;
    cudaEvent_t ev[2];
    NXF_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    NXF_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    const unsigned char* src = static_cast<const unsigned char*>(rows);
    int b = 0;
    bool used[2] = {false, false};
    const int add_threads = std::max(1, std::min(env_int("NXF_ADD_THREADS", 8), static_cast<int>(std::thread::hardware_concurrency())));
    for (int64_t r0 = 0; r0 < n; r0 += static_cast<int64_t>(chunk_rows), b ^= 1) {
      const int64_t rn = std::min<int64_t>(static_cast<int64_t>(chunk_rows), n - r0);
      if (used[b]) NXF_CUDA(cudaEventSynchronize(ev[b]));  // staging buffer b free again
      parallel_memcpy(idx->pinned + b * chunk_bytes, src + static_cast<size_t>(r0) * row_bytes, static_cast<size_t>(rn) * row_bytes, add_threads);
      NXF_CUDA(cudaMemcpyAsync(idx->scratch + b * chunk_bytes, idx->pinned + b * chunk_bytes, static_cast<size_t>(rn) * row_bytes,
                               cudaMemcpyHostToDevice, stream));
      rc = launch_convert_rows(idx->scratch + b * chunk_bytes, src_dtype, src_ld, dst + static_cast<size_t>(r0) * idx->ld * esz,
                               idx->store_dtype, idx->ld, rn, idx->dim, stream);
      if (rc) return rc;
      idx->stats.kernel_launches++;
      NXF_CUDA(cudaEventRecord(ev[b], stream));
      used[b] = true;
    }
    NXF_CUDA(cudaStreamSynchronize(stream));
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
  }
  if (idx->metric == NXF_METRIC_L2) {
    rc = launch_half_norm(dst, idx->store_dtype, idx->ld, n, idx->dim, idx->half_norm + idx->n, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
  }
  if (!idx->max_norm2) {
    NXF_CUDA(cudaMalloc(reinterpret_cast<void**>(&idx->max_norm2), 256));
    NXF_CUDA(cudaMemsetAsync(idx->max_norm2, 0, 256, stream));
  }
  rc = launch_row_norm_max(dst, idx->store_dtype, idx->ld, n, idx->dim, idx->max_norm2, stream);
  if (rc) return rc;
  idx->stats.kernel_launches++;
  idx->n += n;
  idx->stats.rows_added += n;
  return NXF_OK;
}

int nxf_reconstruct(nxf_index* idx, int64_t first, int64_t n, float* dst, int dst_on_device, void* stream_) {
  NXF_REQUIRE(idx != nullptr && dst != nullptr, "nxf_reconstruct: NULL argument");
  std::lock_guard<std::mutex> lk(idx->mu);
  NXF_REQUIRE(first >= 0 && n >= 0 && first + n <= idx->n, "nxf_reconstruct: rows [%lld, %lld) outside [0, %lld)", (long long)first,
              (long long)(first + n), (long long)idx->n);
  if (n == 0) return NXF_OK;
  DeviceGuard g(idx->device);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  UseScope use(idx, stream);
  const size_t esz = dtype_size(idx->store_dtype);
  const unsigned char* src = static_cast<const unsigned char*>(idx->data) + static_cast<size_t>(first) * idx->ld * esz;
  if (dst_on_device) {
    int rc = launch_convert_rows(src, idx->store_dtype, idx->ld, dst, NXF_F32, idx->dim, n, idx->dim, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
    return NXF_OK;
  }
  const int64_t chunk = std::max<int64_t>(1, (int64_t(64) << 20) / (static_cast<int64_t>(idx->dim) * 4));
  int rc = ensure_scratch(idx, static_cast<size_t>(std::min(chunk, n)) * idx->dim * 4);
  if (rc) return rc;
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t rn = std::min(chunk, n - r0);
    rc = launch_convert_rows(src + static_cast<size_t>(r0) * idx->ld * esz, idx->store_dtype, idx->ld, idx->scratch, NXF_F32, idx->dim, rn,
                             idx->dim, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
    NXF_CUDA(cudaMemcpyAsync(dst + r0 * idx->dim, idx->scratch, static_cast<size_t>(rn) * idx->dim * 4, cudaMemcpyDeviceToHost, stream));
    NXF_CUDA(cudaStreamSynchronize(stream));
  }
  return NXF_OK;
}

}  // extern "C" (search helpers below are C++)

// ------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------
namespace {

struct Carver {
  unsigned char* base;
  size_t off = 0;
  explicit Carver(unsigned char* b) : base(b) {}
  template <typename T> T* take(size_t count) {
    off = align_up(off, 256);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
};

struct SearchPlan {
  int path = NXF_PATH_NONE;
  int cta_group = 1;
  int k_eff = 0;     // min(k, n)
  int k_fetch = 0;   // per-owner keep count in the main kernel
  int rescore = 0;
  int use_hi = 0;    // fp32 storage through its bf16 hi plane (K1 / K2T)
  int seg_mode = 0;  // K1 retry level: 1 = fp32 storage, three segments over (hi, mid); 2 = 16-bit storage, two query planes
  int level = 1;     // certified search: 1 candidate pass, 2 segmented retry, 3 exact kernels (no certificate needed)
  int cert = 0;      // evaluate the certificate in the arena merge
  float eps_rel = 0.f, eps_abs = 0.f;
  int q_round = -1;  // type the candidate pass rounds the queries to (its share of the bound is measured per query), -1: none
  int seed = 0;      // K2T: threshold-seeding pass over the first tile of every CTA
  int cap = 0;
  int epi_sets = 2;
  int grid = 0;
  int num_qt = 0, chunks = 1;
  int64_t tiles_per_chunk = 1, n_tiles = 0;
  // K1 interleaved rounds
  int round_tiles = 1, sync_window = 0;
  int64_t arena_stride = 0;
  int64_t owners = 0;
  // K2
  int nq_t = 1, rw = 4, stages = 2;
  size_t k2_smem = 0;
  int64_t batch = 0;  // queries per launch
};

// internal flag bits (above the public NXF_SEARCH_* ones): which level of the certified search this call is
constexpr int kFlagLevel2 = 1 << 20;
constexpr int kFlagLevel3 = 1 << 21;
constexpr int kForceMask = NXF_SEARCH_FORCE_EXACT | NXF_SEARCH_FORCE_STREAM | NXF_SEARCH_FORCE_TENSOR | NXF_SEARCH_FORCE_TSTREAM;

// fp32 accumulation of the tensor cores over `kblocks` K blocks of 64 elements: 4 MMAs per block, two roundings each (the sum of
// the 16 products, then the add into the accumulator), each taken as <= 1 ulp (2^-23, doubled for a truncating adder) of a
// partial sum that never exceeds sum|q_i c_i| <= |q||c|
// (+8 of the constant: K2T's threshold fold -- one more MMA, the 2^-24-relative residual of the three-piece threshold and the
// d + tau reconstruction of a survivor's score)
float acc_slack(int kblocks) { return static_cast<float>(kblocks * 8 + 24) * 2.384185791015625e-07f; }

// K2T's threshold fold (search_stream_tc.cu), an experiment kept behind NXF_K2T_FOLD=1: short rows, many queries.  Measured on the C5
// corpus (profiles/r2_k2t_knockout.md): 128 queries k = 10, main launch 298 -> 282 us, but +9 us on the sample launch and 368 -> 453 us at
// k = 500 (every survivor re-reads its 32-query block); 64 queries and fewer: slower.  Not the default.
bool k2t_use_fold(const nxf_index* idx, int nq_pad) {
  if (env_int("NXF_K2T_FOLD", 0) == 0) return false;
  if (idx->metric != NXF_METRIC_IP) return false;
  if (idx->store_dtype != NXF_BF16 && idx->store_dtype != NXF_F32) return false;   // fp32 rows: K2T streams their bf16 plane
  return idx->ld / 64 <= env_int("NXF_K2T_FOLD_MAX_KB", 2) && nq_pad >= env_int("NXF_K2T_FOLD_MIN_NQ", 32);
}

}  // namespace

static int plan_search(const nxf_index* idx, int64_t nq, int q_dtype, int k, int flags, bool hi_ok, SearchPlan* sp) {
  SearchPlan& P = *sp;
  const bool store16 = idx->store_dtype != NXF_F32;
  const bool tens = store16 || hi_ok;  // K1 / K2T can run: 16-bit rows, or fp32 rows with their bf16 hi plane
  P.k_eff = static_cast<int>(std::min<int64_t>(k, idx->n));
  P.level = (flags & kFlagLevel3) ? 3 : ((flags & kFlagLevel2) ? 2 : 1);
  const int64_t row_bytes = idx->ld * dtype_size(idx->store_dtype);
  int path;
  if (P.level == 3) path = (nq <= 16) ? NXF_PATH_STREAM : NXF_PATH_EXACT;   // exact fp32 FMAs on the stored values
  else if (P.level == 2) path = NXF_PATH_TENSOR;                             // segmented K1
  else if (flags & NXF_SEARCH_FORCE_EXACT) path = NXF_PATH_EXACT;
  else if (flags & NXF_SEARCH_FORCE_STREAM) path = NXF_PATH_STREAM;
  else if (flags & NXF_SEARCH_FORCE_TENSOR) path = NXF_PATH_TENSOR;
  else if (flags & NXF_SEARCH_FORCE_TSTREAM) path = NXF_PATH_TSTREAM;
  // small batches stream the corpus once through the tensor cores (K2T); round 2: also a single query on long rows (same GB/s as
  // K2 there, and a tenth of its arena to merge) and fp32 storage, which streams its 2-byte hi plane instead of the 4-byte rows
  // ... except 65-128 queries on 128-byte rows of a large corpus: there K2T is bound by its per-tile filter (one tile = 16 KB of stream),
  // and K1 with four epilogue groups on split accumulators is faster (C5 corpus, same box: 128 queries k = 10 0.332 -> 0.287 ms,
  // k = 500 0.429 -> 0.334 ms; 96 queries 0.305 -> 0.284 / 0.391 -> 0.326; at 64 queries K2T wins, 0.243 vs 0.296)
  else if (tens && idx->ld == 64 && nq > env_int("NXF_TSTREAM_SHORT_MAX_NQ", 64) && nq <= NXF_TSTREAM_MAX_NQ &&
           idx->n >= env_int("NXF_TSTREAM_SHORT_MIN_ROWS", 2000000) && (store16 || hi_ok)) path = NXF_PATH_TENSOR;
  else if (store16 && nq <= NXF_TSTREAM_MAX_NQ && nq >= env_int("NXF_TSTREAM_MIN_NQ", 1)) path = NXF_PATH_TSTREAM;
  else if (hi_ok && nq <= NXF_TSTREAM_MAX_NQ && nq >= env_int("NXF_TSTREAM_MIN_NQ_F32", 1)) path = NXF_PATH_TSTREAM;
  else if (nq <= NXF_STREAM_MAX_NQ) path = NXF_PATH_STREAM;
  else path = tens ? NXF_PATH_TENSOR : NXF_PATH_EXACT;

  if (path == NXF_PATH_TSTREAM) {
    // K2T keeps the (16-padded) queries resident in shared memory; if that leaves fewer than 4 ring stages fall back to the kernels
    // that do not need it
    int nq_pad = static_cast<int>(round_up64(std::min<int64_t>(nq, NXF_TSTREAM_MAX_NQ), 16));
    const bool fold = k2t_use_fold(idx, nq_pad);
    if (fold) nq_pad = (nq_pad + 31) & ~31;   // the folded filter reads the accumulator 32 columns at a time
    const int st = tens ? k2t_plan_stages(static_cast<int>(idx->ld / 64), nq_pad, fold) : 0;
    if (st == 0) {
      if (flags & NXF_SEARCH_FORCE_TSTREAM) {
        set_error("nxf_search: the tensor stream kernel needs queries that fit in shared memory (nq=%lld, dim=%d) and, for f32 storage, the bf16 plane",
                  (long long)nq, idx->dim);
        return NXF_EUNSUPPORTED;
      }
      path = nq <= NXF_STREAM_MAX_NQ ? NXF_PATH_STREAM : (tens ? NXF_PATH_TENSOR : NXF_PATH_EXACT);
    } else {
      P.stages = st;
      P.nq_t = nq_pad;
    }
  }
  if (path == NXF_PATH_STREAM) {
    // rows per consumer warp per stage: keep a stage <= 48 KB and at least two stages in flight
    int rw = 4;
    while (rw > 1 && row_bytes * 8 * rw > 48 * 1024) rw >>= 1;
    const int nq_t = nq >= 3 ? 4 : (nq == 2 ? 2 : 1);
    int stages = 4;
    while (stages > 1 && k2_smem_bytes(idx->ld, dtype_size(idx->store_dtype), nq_t, rw, stages) > 200 * 1024) --stages;
    if (stages < 2) {
      if (flags & NXF_SEARCH_FORCE_STREAM) { set_error("nxf_search: rows of %lld bytes are too long for the stream kernel", (long long)row_bytes); return NXF_EUNSUPPORTED; }
      path = (P.level == 3 || !tens) ? NXF_PATH_EXACT : NXF_PATH_TENSOR;
    } else {
      P.rw = rw; P.nq_t = nq_t; P.stages = stages;
      P.k2_smem = k2_smem_bytes(idx->ld, dtype_size(idx->store_dtype), nq_t, rw, stages);
    }
  }
  if (path == NXF_PATH_TENSOR && !tens) {
    set_error("nxf_search: the tensor-core path over f32 storage needs its bf16 plane, which did not fit in device memory (or NXF_F32_TENSOR=0)");
    return NXF_EUNSUPPORTED;
  }
  P.path = path;
  const bool tpath = path == NXF_PATH_TENSOR || path == NXF_PATH_TSTREAM;
  P.use_hi = (tpath && !store16) ? 1 : 0;
  P.k_fetch = P.k_eff;
  P.rescore = 0;
  // ---- certified search: the MMA sees queries (and, for fp32 storage, rows) rounded to 16 bits.  Unless the caller's
  // queries already are of the storage type, over-fetch k' candidates, re-score them in fp32 and certify the result ----
  const bool q_exact = store16 && q_dtype == idx->store_dtype;
  if (tpath && !(flags & NXF_SEARCH_NO_RESCORE) && !q_exact) {
    P.rescore = 1;
    const int kb = static_cast<int>(idx->ld / 64);
    int64_t kf;
    if (P.level == 2) {
      // segmented retry: the pass score is within ~1e-4 |q||c| of the fp32 score, a quarter more candidates is plenty
      kf = static_cast<int64_t>(P.k_eff) + std::max<int64_t>(16, (P.k_eff + 3) / 4);
      P.seg_mode = store16 ? 2 : 1;
      if (path != NXF_PATH_TENSOR) P.seg_mode = 0;
      P.eps_rel = (store16 ? 7.62939453125e-06f : 3.f * 7.62939453125e-06f) + acc_slack(kb * (store16 ? 2 : 3));   // 2^-17, 3 * 2^-17
    } else if (store16) {
      // query rounding only: |q - q16| <= 2^-9 |q| -> every pass score within 2^-9 |q||c| of the fp32 score; k' = 2k keeps the gap
      // between ranks k and k' above that on the BASELINE configs (C2: 0.0069 against 0.0020)
      kf = std::max<int64_t>(static_cast<int64_t>(std::max(1, env_int("NXF_KFETCH_MULT16", 2))) * P.k_eff, P.k_eff + 16);
      P.q_round = idx->store_dtype;          // the merge measures |q - round(q)| itself (<= 2^-9 |q|, ~0.6 of that on average)
      P.eps_rel = acc_slack(kb);
    } else {
      // fp32 rows through their bf16 hi plane: both operands rounded, (2^-8 + 2^-16) |q||c|; k' = 4k
      kf = std::max<int64_t>(static_cast<int64_t>(std::max(1, env_int("NXF_KFETCH_MULT32", 4))) * P.k_eff, P.k_eff + 64);
      P.q_round = NXF_BF16;                  // query part measured; the rows' part: |round(q) . (c - bf16(c))| <= (1 + 2^-9) |q| 2^-9 |c|
      P.eps_rel = 0.001953125f * 1.002f + acc_slack(kb);
    }
    if (idx->store_dtype == NXF_F16) P.eps_abs = std::sqrt(static_cast<float>(idx->dim)) * 5.9604644775390625e-08f;  // f16 subnormals: 2^-24 per element
    kf = (kf + 7) / 8 * 8;
    P.k_fetch = static_cast<int>(std::max<int64_t>(P.k_eff, std::min<int64_t>(std::min<int64_t>(kf, NXF_MAX_K), idx->n)));
    P.cert = env_int("NXF_CERT", 1) != 0 ? 1 : 0;
    if (!P.cert && P.level == 1) {
      // NXF_CERT=0: round 1's fixed over-fetch k' = k + max(16, k/4), no certificate, no retry (kept for A/B timing)
      int64_t k1 = static_cast<int64_t>(P.k_eff) + std::max<int64_t>(16, (P.k_eff + 3) / 4);
      k1 = (k1 + 7) / 8 * 8;
      P.k_fetch = static_cast<int>(std::max<int64_t>(P.k_eff, std::min<int64_t>(std::min<int64_t>(k1, NXF_MAX_K), idx->n)));
    }
  }
  if (path == NXF_PATH_STREAM) {
    const int64_t R = 8 * P.rw;
    const int64_t n_blocks = ceil_div64(idx->n, R);
    P.grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(idx->sm_count, n_blocks)));
    P.cap = std::max(2 * P.k_eff, P.k_eff + 32);
    P.owners = static_cast<int64_t>(P.grid) * 8 * P.nq_t;
    P.arena_stride = static_cast<int64_t>(P.grid) * 8 * P.k_eff;
    P.batch = P.nq_t;
  } else if (path == NXF_PATH_TSTREAM) {
    P.n_tiles = ceil_div64(idx->n, 128);
    P.grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(idx->sm_count, P.n_tiles)));
    // overflow checks every `ce` tiles: short rows (few K blocks per tile) cannot afford a barrier per tile
    P.rw = std::max(1, std::min(8, 16 / std::max<int>(1, static_cast<int>(idx->ld / 64))));
    // epilogue warp sets: two while the filter keeps up, three / four once it is latency bound (search_stream_tc.cu); NXF_K2T_SETS=2..4 overrides
    {
      const int64_t nqp = round_up64(nq, 16);
      P.epi_sets = std::max(2, std::min(4, env_int("NXF_K2T_SETS", nqp <= 32 ? 2 : (nqp <= 96 ? 3 : 4))));
    }
    P.cap = 2 * P.k_fetch + P.epi_sets * 128 * P.rw;   // pruned once more than 2k' entries are held (one round of appends on top)
    P.owners = static_cast<int64_t>(P.grid) * P.nq_t;
    // threshold seeding (DESIGN.md "K2T"): sample passes first, the sample's m-th best into tau_g.  With the estimate-only first pass
    // the sample launches cost ~25 us: on 128-byte rows they pay from nq * k' ~ 200 on (C5 corpus, 16 queries k = 10: 0.309 -> 0.238 ms;
    // one query k = 500: 0.280 -> 0.237 ms; one query k = 10: 0.218 -> 0.232 ms, left alone); long rows amortise the appends over more
    // stream bytes and keep the old threshold (C2, 2-4 queries: -1 % with the launches).  NXF_K2T_SEED=0/1 overrides
    {
      const int seed_env = env_int("NXF_K2T_SEED", -1);
      const int64_t seed_from = idx->ld * 2 < 512 ? 192 : 768;
      // round 2: also whenever the unseeded search would leave one list of k' per CTA for the arena merge to reduce (grid * k' entries
      // per query: one query, k' = 200 on C2 = 29 600 entries, 0.16-0.27 ms in a single merge CTA -- more than the sample launches cost)
      const bool big_arena = static_cast<int64_t>(P.grid) * P.k_fetch >= env_int("NXF_K2T_SEED_ARENA", 4096);
      P.seed = seed_env >= 0 ? seed_env != 0 : (nq * static_cast<int64_t>(P.k_fetch) >= seed_from || big_arena);
    }
    P.arena_stride = static_cast<int64_t>(P.grid) * P.k_fetch;
    P.batch = std::min<int64_t>(nq, NXF_TSTREAM_MAX_NQ);
  } else if (path == NXF_PATH_EXACT) {
    P.n_tiles = ceil_div64(idx->n, 128);
    P.num_qt = static_cast<int>(ceil_div64(nq, 128));
    const int max_grid = 2 * idx->sm_count;
    plan_units(P.num_qt, P.n_tiles, max_grid, &P.chunks, &P.tiles_per_chunk);
    P.grid = static_cast<int>(std::min<int64_t>(max_grid, static_cast<int64_t>(P.num_qt) * P.chunks));
    P.cap = std::max(2 * P.k_eff, P.k_eff + 128);
    P.owners = static_cast<int64_t>(P.grid) * 128;
    P.arena_stride = static_cast<int64_t>(P.chunks) * P.k_eff;
    P.batch = nq;
  } else {
    // one 128-row query tile: a CTA pair (M = 256) would spend half its MMAs on padding
    int cg = env_int("NXF_K1_CTA_GROUP", 0);
    if (cg != 1 && cg != 2) cg = (nq <= 128) ? 1 : 2;
    if (flags & NXF_SEARCH_TENSOR_1CTA) cg = 1;
    P.cta_group = cg;
    const int clusters = idx->sm_count / cg;
    // one launch serves at most one query tile per cluster (every cluster keeps <= 2 query tiles' state for the whole launch)
    P.batch = std::min<int64_t>(nq, static_cast<int64_t>(clusters) * 128 * cg);
    P.n_tiles = ceil_div64(idx->n, 256);
    P.num_qt = static_cast<int>(ceil_div64(P.batch, 128 * cg));
    const int64_t units = static_cast<int64_t>(P.num_qt) * P.n_tiles;
    const int used = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(clusters, units)));
    P.grid = used * cg;
    // round = the corpus window all clusters share through L2: a multiple of used/gcd(used, num_qt) tiles makes the P pieces of
    // a round exactly equal; sized to NXF_K1_ROUND_MB (default 32 MB)
    {
      int a = used, b = P.num_qt;
      while (b) { const int t = a % b; a = b; b = t; }
      const int64_t unit_tiles = used / a;
      const int64_t tile_bytes = 256 * idx->ld * 2;
      const int64_t target = static_cast<int64_t>(std::max(1, env_int("NXF_K1_ROUND_MB", 32))) << 20;
      const int64_t m = std::max<int64_t>(1, (target + unit_tiles * tile_bytes / 2) / (unit_tiles * tile_bytes));
      P.round_tiles = static_cast<int>(std::min<int64_t>(unit_tiles * m, std::max<int64_t>(P.n_tiles, 1)));
      P.sync_window = env_int("NXF_K1_SYNC_WINDOW", 1);
      if (P.num_qt == 1) {
        // one query tile: nothing is shared between clusters, so every cluster streams one contiguous chunk, unsynchronised
        P.round_tiles = static_cast<int>(std::max<int64_t>(P.n_tiles, 1));
        P.sync_window = 0;
      }
    }
    P.chunks = static_cast<int>(ceil_div64(std::max<int64_t>(P.n_tiles, 1), P.round_tiles));  // reported as "chunks": rounds per launch
    // pruned between tiles once fewer than 256 (one tile of appends) slots are left
    P.cap = std::max(2, env_int("NXF_K1_CAP_MULT", 2)) * P.k_fetch + 256;
    // a second epilogue warp set when a tile holds too few K blocks for one set to hide behind the MMAs (search_tc.cu); NXF_K1_SETS=1/2
    // overrides.  Measured (4 M rows, 16 384 queries, same box, one -> two sets): D = 64 +65 %, 128 +35 %, 256 +7 %, 320 +3 %, 384-512 +2 %,
    // 768 and 1536 +-0
    {
      const int kb_tile = static_cast<int>(idx->ld / 64) * (P.seg_mode == 1 ? 3 : (P.seg_mode == 2 ? 2 : 1));
      // small searches (C1: 21 tile-units per cluster) lose 7 % to the second set's fixed costs (twice the lists to flush and merge)
      const int64_t units_per_cluster = static_cast<int64_t>(P.num_qt) * P.n_tiles / std::max(used, 1);
      // rows of one or two K blocks: four groups on split accumulators (C5 full batch, same box: kernel 31.9 -> 27.8 ms; D = 128 40.0 -> 37.2 ms)
      const int want = env_int("NXF_K1_SETS", units_per_cluster >= 128 ? (kb_tile <= 2 ? 4 : (kb_tile <= 8 ? 2 : 1)) : 1);
      // four groups exist only with the split accumulators (rows of <= 2 K blocks, search_tc.cu)
      P.epi_sets = (want >= 4 && kb_tile <= 2) ? 4 : (want >= 2 ? 2 : 1);
    }
    P.owners = static_cast<int64_t>(P.grid) * 128 * 2 * P.epi_sets;  // two query-tile slots per CTA, one buffer per (slot, set, row)
    // arena slice of a query: the list left by the sample passes (<= k' entries) + up to cap entries from every (cluster, set) that
    // owns a piece of the query tile in a pass
    P.arena_stride = k1_arena_stride(used, P.num_qt, P.k_fetch, P.epi_sets, P.cap);
  }
  return NXF_OK;
}

// One level of the (certified) search over queries [0, nq).  Called with idx->mu held and the index's device current.
// out_packed != NULL (device pointer): rows are written as 8-byte (score, shard-local row) entries instead of the two arrays.
static int search_locked(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int q_on_device, int k, float* out_scores,
                         int64_t* out_ids, Entry* out_packed, int out_on_device, int flags, cudaStream_t stream) {
  int rc;
  const bool l2 = idx->metric == NXF_METRIC_L2;
  const int64_t total = nq * k;

  // ---- empty index: Faiss fills with -1 / neutral scores ----
  if (idx->n == 0) {
    const float fill = l2 ? FLT_MAX : -FLT_MAX;
    if (!out_on_device) {
      for (int64_t i = 0; i < total; ++i) { out_scores[i] = fill; out_ids[i] = -1; }
      return NXF_OK;
    }
    // device outputs: run the arena merge over an all-empty arena (qcnt = 0)
  }

  // fp32 storage goes to the tensor cores through its bf16 hi plane, built here on first use (NXF_F32_TENSOR=0 keeps it on the FFMA
  // kernels); the mid plane is only built when the retry level needs it
  bool hi_ok = false;
  if (idx->store_dtype == NXF_F32 && idx->n > 0 && !(flags & (NXF_SEARCH_FORCE_EXACT | NXF_SEARCH_FORCE_STREAM | kFlagLevel3)) &&
      ((flags & (NXF_SEARCH_FORCE_TENSOR | NXF_SEARCH_FORCE_TSTREAM | kFlagLevel2)) || env_int("NXF_F32_TENSOR", 1) != 0)) {
    rc = ensure_plane(idx, 0, stream);
    if (rc == NXF_OK && (flags & kFlagLevel2)) rc = ensure_plane(idx, 1, stream);
    if (rc == NXF_OK) hi_ok = true;
    else if (rc != NXF_ENOMEM) return rc;
    else if (flags & kFlagLevel2) flags = (flags & ~kFlagLevel2) | kFlagLevel3;   // no room for the mid plane: straight to the exact kernels
  }

  // plan; if the per-query arena forces several launches, re-plan for the launch size
  SearchPlan P;
  int64_t batch = nq;
  for (int it = 0; it < 8; ++it) {
    P = SearchPlan();
    rc = plan_search(idx, batch, q_dtype, k, flags, hi_ok, &P);
    if (rc) return rc;
    if (idx->n == 0) break;
    if (P.path == NXF_PATH_STREAM) { batch = std::min<int64_t>(nq, P.batch); break; }
    if (P.path == NXF_PATH_TSTREAM) { batch = std::min<int64_t>(nq, P.batch); break; }
    if (P.path == NXF_PATH_TENSOR && P.batch < batch) { batch = P.batch; continue; }  // re-plan for the per-launch query count
    const int64_t budget = (int64_t(6) << 30) << (it >= 6 ? 1 : 0);  // last tries: accept a larger arena
    const int64_t per_q = std::max<int64_t>(P.arena_stride * 8, 1);
    const int64_t gran = 128 * P.cta_group;
    const int64_t max_q = std::max<int64_t>(gran, budget / per_q / gran * gran);
    if (batch <= max_q || it == 7) break;
    batch = max_q;
  }
  if (idx->n == 0) { P.path = NXF_PATH_NONE; batch = nq; P.arena_stride = 1; P.owners = 0; P.cap = 1; P.k_fetch = 0; P.k_eff = 0; P.rescore = 0; P.cert = 0; }
  const bool multi_batch = batch < nq;

  // K1: the last launch of a multi-batch search holds fewer query tiles, so more clusters share a tile: its slices are longer
  int64_t rem_arena = 0, rem_stride = 0;
  if (P.path == NXF_PATH_TENSOR && multi_batch && nq % batch != 0) {
    const int64_t rem = nq % batch;
    const int64_t nqt = ceil_div64(rem, 128 * P.cta_group);
    const int64_t cl = std::max<int64_t>(1, std::min<int64_t>(P.grid / P.cta_group, nqt * P.n_tiles));
    rem_stride = k1_arena_stride(cl, nqt, P.k_fetch, P.epi_sets, P.cap);
    rem_arena = rem * rem_stride;
  }

  // operands of the tensor paths: 16-bit rows as stored, or the bf16 planes of fp32 rows
  const bool tpath = P.path == NXF_PATH_TENSOR || P.path == NXF_PATH_TSTREAM;
  const void* c_rows = P.use_hi ? idx->hi : idx->data;
  const void* c_rows2 = P.seg_mode == 1 ? idx->mid : nullptr;
  const int c_dt = P.use_hi ? NXF_BF16 : idx->store_dtype;
  const int64_t q16_ld = P.seg_mode ? 2 * idx->ld : idx->ld;

  // ---- carve scratch (two passes: size, then pointers) ----
  const size_t qsz = dtype_size(q_dtype);
  unsigned char* base = nullptr;
  float* q_f32 = nullptr; void* q16 = nullptr; void* q_src = nullptr; uint32_t* tau_g = nullptr; uint32_t* qcnt = nullptr;
  Entry* arena = nullptr; Entry* cand = nullptr; float* o_scores = nullptr; int64_t* o_ids = nullptr; uint32_t* k2t_cnt = nullptr;
  uint32_t* cert_w = nullptr;   // [nq] failure flags of the certificate, then one counter word
  const int64_t nq_pad = round_up64(batch, 256);
  for (int pass = 0; pass < 2; ++pass) {
    Carver cv(base);
    q_src = q_on_device ? nullptr : cv.take<unsigned char>(static_cast<size_t>(batch) * q_ld * qsz);
    q_f32 = cv.take<float>(static_cast<size_t>(nq_pad) * idx->ld);
    q16 = tpath ? cv.take<uint16_t>(static_cast<size_t>(nq_pad) * q16_ld) : nullptr;
    // tau | qcnt | 64 control words: [0..5] K1 round counters (one per pass), [8] verification flag, [16..23] debug cycle counters
    tau_g = cv.take<uint32_t>(static_cast<size_t>(2 * nq_pad + 64));
    qcnt = tau_g ? tau_g + nq_pad : nullptr;
    arena = cv.take<Entry>(static_cast<size_t>(std::max<int64_t>(batch * P.arena_stride, rem_arena)));
    cand = cv.take<Entry>(static_cast<size_t>(std::max<int64_t>(P.owners, 1)) * P.cap);
    k2t_cnt = P.path == NXF_PATH_TSTREAM ? cv.take<uint32_t>(static_cast<size_t>(std::max<int64_t>(P.owners, 1))) : nullptr;
    cert_w = P.cert ? cv.take<uint32_t>(static_cast<size_t>(nq) + 64) : nullptr;
    if (!out_on_device) {
      o_scores = cv.take<float>(static_cast<size_t>(batch) * k);
      o_ids = cv.take<int64_t>(static_cast<size_t>(batch) * k);
    }
    if (pass == 0) {
      rc = ensure_scratch(idx, cv.off + 256);
      if (rc) return rc;
      base = idx->scratch;
    }
  }
  if (P.cert) NXF_CUDA(cudaMemsetAsync(cert_w + nq, 0, 64 * sizeof(uint32_t), stream));

  for (int64_t b0 = 0; b0 < nq; b0 += batch) {
    const int64_t bn = std::min(batch, nq - b0);
    const unsigned char* qb = static_cast<const unsigned char*>(q) + static_cast<size_t>(b0) * q_ld * qsz;
    const void* q_dev = qb;
    if (!q_on_device) {
      NXF_CUDA(cudaMemcpyAsync(q_src, qb, static_cast<size_t>(bn) * q_ld * qsz, cudaMemcpyHostToDevice, stream));
      q_dev = q_src;
    }
    // queries -> fp32 [bn, ld] (zero padded); 16-bit copy (or [hi | mid] planes) for the MMA
    rc = launch_convert_rows(q_dev, q_dtype, q_ld, q_f32, NXF_F32, idx->ld, bn, idx->dim, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;
    if (tpath && P.seg_mode) {
      rc = launch_split_rows(q_f32, idx->ld, q16, bn, c_dt, stream);
      if (rc) return rc;
      idx->stats.kernel_launches++;
    } else if (tpath) {
      rc = launch_convert_rows(q_dev, q_dtype, q_ld, q16, c_dt, idx->ld, bn, idx->dim, stream);
      if (rc) return rc;
      idx->stats.kernel_launches++;
    }
    NXF_CUDA(cudaMemsetAsync(tau_g, 0, static_cast<size_t>(2 * nq_pad + 64) * sizeof(uint32_t), stream));

    cudaEvent_t pe0 = nullptr, pe1 = nullptr;
    if (idx->profiling && P.path != NXF_PATH_NONE) {
      NXF_CUDA(cudaEventCreate(&pe0));
      NXF_CUDA(cudaEventCreate(&pe1));
      NXF_CUDA(cudaEventRecord(pe0, stream));
    }
    const int64_t arena_stride = (P.path == NXF_PATH_TENSOR && bn < batch && rem_stride > 0) ? rem_stride : P.arena_stride;
    SinkParams sink;
    sink.tau_g = tau_g; sink.qcnt = qcnt; sink.arena = arena; sink.cand = cand;
    sink.arena_stride = arena_stride; sink.cap = P.cap; sink.k = P.k_fetch;

    if (P.path == NXF_PATH_STREAM) {
      // K2 handles up to nq_t queries per launch; a forced stream search of more queries loops
      for (int64_t s0 = 0; s0 < bn; s0 += P.nq_t) {
        const int sn = static_cast<int>(std::min<int64_t>(P.nq_t, bn - s0));
        K2Params kp;
        kp.q = q_f32 + s0 * idx->ld; kp.ldq = idx->ld; kp.corpus = idx->data; kp.ld = idx->ld;
        kp.half_norm = l2 ? idx->half_norm : nullptr; kp.n = idx->n; kp.nq = sn; kp.dim = idx->dim; kp.stages = P.stages;
        kp.n_blocks = ceil_div64(idx->n, 8 * P.rw);
        kp.sink = sink;
        kp.sink.tau_g = tau_g + s0; kp.sink.qcnt = qcnt + s0; kp.sink.arena = arena + s0 * P.arena_stride;
        rc = launch_k2(idx->store_dtype, kp, P.nq_t, P.rw, P.grid, P.k2_smem, stream);
        if (rc) return rc;
        idx->stats.kernel_launches++;
      }
    } else if (P.path == NXF_PATH_TSTREAM) {
      K2TParams kp;
      kp.n = idx->n; kp.nq = static_cast<int>(bn); kp.nq_pad = static_cast<int>(round_up64(bn, 16)); kp.kblocks = static_cast<int>(idx->ld / 64);
      // bf16 rows (or the bf16 plane of fp32 rows), ip, short rows, many queries: the threshold comparison moves into the tensor core
      // (and a round of at least 2 * tbufs tiles: the issuer must never be two threshold versions ahead of the epilogue)
      kp.fold = (k2t_use_fold(idx, kp.nq_pad) && c_dt == NXF_BF16 && P.epi_sets * P.rw >= 16) ? 1 : 0;
      if (kp.fold) kp.nq_pad = (kp.nq_pad + 31) & ~31;   // 32-column accumulator loads (P.nq_t was planned the same way)
      kp.q_f32 = q_f32; kp.ldq = idx->ld; kp.dim = idx->dim; kp.max_norm2 = idx->max_norm2;
      kp.stages = k2t_plan_stages(kp.kblocks, kp.nq_pad, kp.fold);
      kp.tbuf_stride = kp.nq_pad <= 32 ? 32 : (kp.nq_pad <= 64 ? 64 : 128);
      kp.tbufs = std::min(8, 512 / kp.tbuf_stride);
      kp.half_norm = l2 ? idx->half_norm : nullptr;
      kp.check_every = P.rw;
      kp.epi_sets = P.epi_sets;
      kp.sink = sink;
      kp.sink.cap = P.cap;
      const int64_t c_ld = idx->ld;
      // Sampled thresholds, as in K1: up to two short launches over evenly spread sample tiles first (about one tile
      // per CTA, then about 16 more); after each, the seed kernel raises tau_g to the sample's m-th best over ALL lists of the
      // query (m from sample_rank(): at least k' rows of the corpus beat it except with probability ~e^-30).  The counts are
      // verified after the last launch; a query that came back short is searched again from -inf by the fall-back launch.
      kp.perm = TilePerm();
      kp.run_flag = nullptr;
      kp.cnt_g = k2t_cnt;
      int64_t bounds[4] = {0, 0, 0, P.n_tiles};
      int n_pass = 0;
      if (P.seed) {
        const int64_t g = P.grid;
        if (P.n_tiles >= 68 * g) {
          kp.perm.s = P.n_tiles / (17 * g); kp.perm.ns = ceil_div64(P.n_tiles, kp.perm.s);
          kp.perm.s0 = ceil_div64(kp.perm.ns, g); kp.perm.n0 = ceil_div64(kp.perm.ns, kp.perm.s0);
          bounds[++n_pass] = kp.perm.n0;
          bounds[++n_pass] = kp.perm.ns;
        } else if (P.n_tiles >= 4 * g) {
          kp.perm.s = P.n_tiles / g; kp.perm.ns = ceil_div64(P.n_tiles, kp.perm.s);
          kp.perm.s0 = 1; kp.perm.n0 = kp.perm.ns;
          bounds[++n_pass] = kp.perm.ns;
        }
      }
      kp.perm.finish();
      bounds[++n_pass] = P.n_tiles;
      kp.est_only = 0;
      // With two sample passes the first (about one tile per CTA) is estimate-only: the CTA's best score per query, the m-th
      // largest of which over the CTAs seeds the threshold; its tiles are streamed again by the second pass, which starts from
      // empty lists.  (Appending everything from -inf instead filled 128 lists per CTA, cost a 19 k-key selection per query
      // in between and a prune of every list at the start of the next launch: 140 us of a 550 us search at 128 queries.)
      const bool est_first = n_pass == 3 && env_int("NXF_K2T_EST", 1) != 0;
      for (int ps = 0; ps + 1 < n_pass; ++ps) {
        const bool est = est_first && ps == 0;
        kp.tile_begin = (est_first && ps == 1) ? 0 : bounds[ps]; kp.tile_end = bounds[ps + 1];
        kp.first_pass = ps == 0 || (est_first && ps == 1); kp.last_pass = 0; kp.est_only = est ? 1 : 0;
        kp.no_check = (est || kp.tile_end - kp.tile_begin <= P.grid) ? 1 : 0;
        rc = launch_k2t(c_dt, q16, bn, q16_ld, c_rows, c_ld, kp, P.grid, stream);
        if (rc) return rc;
        kp.est_only = 0;
        const int m = std::min(P.k_fetch, sample_rank(P.k_fetch, bounds[ps + 1], P.n_tiles));
        if (est) {
          // the m-th largest of min(grid, tiles) per-CTA maxima; more than that many is not available: leave the threshold alone
          if (m <= std::min<int64_t>(P.grid, bounds[1])) rc = launch_seed_tau_est(k2t_cnt, P.grid, kp.nq_pad, m, tau_g, static_cast<int>(bn), stream);
        } else {
          rc = launch_seed_tau_lists(cand, k2t_cnt, P.grid, kp.nq_pad, P.cap, m, tau_g, static_cast<int>(bn), stream);
        }
        if (rc) return rc;
        idx->stats.kernel_launches += 2;
      }
      kp.tile_begin = bounds[n_pass - 1]; kp.tile_end = P.n_tiles;
      kp.first_pass = n_pass == 1; kp.last_pass = 1; kp.no_check = 0;
      if (n_pass > 1) {
        rc = launch_k2t(c_dt, q16, bn, q16_ld, c_rows, c_ld, kp, P.grid, stream);
        if (rc) return rc;
        uint32_t* flag = tau_g + 2 * nq_pad + 8;
        rc = launch_verify_counts(qcnt, tau_g, bn, static_cast<int>(std::min<int64_t>(P.k_fetch, idx->n)), flag, stream);
        if (rc) return rc;
        idx->stats.kernel_launches += 2;
        // the fall-back launch: everything again from empty lists, for the queries the verification reset (all others: tau = +inf)
        kp.tile_begin = 0; kp.first_pass = 1; kp.run_flag = flag;
      }
      rc = launch_k2t(c_dt, q16, bn, q16_ld, c_rows, c_ld, kp, P.grid, stream);
      if (rc) return rc;
      idx->stats.kernel_launches++;
    } else if (P.path == NXF_PATH_EXACT) {
      KxParams kp;
      kp.q = q_f32; kp.ldq = idx->ld; kp.corpus = idx->data; kp.ld = idx->ld; kp.half_norm = l2 ? idx->half_norm : nullptr;
      kp.n = idx->n; kp.nq = bn; kp.dim = idx->dim;
      kp.num_qt = static_cast<int>(ceil_div64(bn, 128)); kp.chunks = P.chunks; kp.tiles_per_chunk = P.tiles_per_chunk; kp.n_tiles = P.n_tiles;
      kp.sink = sink;
      rc = launch_kx(idx->store_dtype, kp, P.grid, stream);
      if (rc) return rc;
      idx->stats.kernel_launches++;
    } else if (P.path == NXF_PATH_TENSOR) {
      K1Params kp;
      kp.n = idx->n; kp.nq = bn; kp.kblocks = static_cast<int>(idx->ld / 64);
      kp.split_kb = 0; kp.seg_mode = P.seg_mode;
      if (P.seg_mode) { kp.split_kb = kp.kblocks; kp.kblocks *= (P.seg_mode == 1 ? 3 : 2); }
      kp.num_qt = static_cast<int>(ceil_div64(bn, 128 * P.cta_group));
      kp.epi_sets = P.epi_sets;
      // short rows: the accumulator in two halves, so that a set never waits for the whole MMA of its next tile
      // (NXF_K1_SPLIT=0: whole accumulators; =2: required -- the call fails if the plan does not allow it, so that a test knows it ran)
      const int split_env = env_int("NXF_K1_SPLIT", 1);
      kp.split = (P.epi_sets >= 2 && kp.kblocks <= 2 && (split_env != 0 || P.epi_sets == 4)) ? 1 : 0;
      NXF_REQUIRE(split_env != 2 || kp.split, "nxf_search: NXF_K1_SPLIT=2 but the plan has %d epilogue set(s) and %d K blocks per tile",
                  P.epi_sets, kp.kblocks);
      kp.half_norm = l2 ? idx->half_norm : nullptr;
      kp.debug = env_int("NXF_K1_DEBUG", 0);
      kp.est_only = 0;
      kp.sink = sink;
      kp.round_tiles = P.round_tiles; kp.sync_window = P.sync_window;
      // the last launch of a multi-batch search may hold fewer query tiles than planned: never more clusters than tile-units
      int grid = P.grid;
      {
        const int64_t units = static_cast<int64_t>(kp.num_qt) * P.n_tiles;
        if (units < grid / P.cta_group) grid = static_cast<int>(std::max<int64_t>(units, 1)) * P.cta_group;
      }
      NXF_REQUIRE(kp.num_qt <= grid / P.cta_group || P.n_tiles == 0, "nxf_search: internal: %d query tiles for %d clusters", kp.num_qt,
                  grid / P.cta_group);
      // Sampled thresholds: two short passes over an evenly spread sample of the corpus tiles (1/256, then 1/16 of
      // them) leave in tau_g the score of a rank chosen so that -- except with probability ~e^-30 per query -- at least k' rows of
      // the whole corpus beat it; the main pass over the other tiles then keeps ~3 k' candidates per query instead of
      // k' ln(N / k') per owner.  The counts are verified afterwards; a query that came back short is searched again.
      kp.perm = TilePerm();
      kp.run_flag = nullptr;
      kp.prof = tau_g + 2 * nq_pad + 16;
      struct Pass { int64_t begin, end; int k; int est_only; };
      Pass passes[4];
      int n_pass = 0;
      {
        const int est_env = env_int("NXF_K1_EST", -1);
        const int64_t s1 = std::max(2, env_int("NXF_K1_EST_S1", 16)), s0 = std::max(2, env_int("NXF_K1_EST_S0", 16));
        const bool est = (est_env >= 0 ? est_env != 0 : true) && P.n_tiles >= 4 * s1 * s0 && kp.debug == 0;
        if (est) {
          kp.perm.s = s1; kp.perm.ns = ceil_div64(P.n_tiles, s1);
          kp.perm.s0 = s0; kp.perm.n0 = ceil_div64(kp.perm.ns, s0);
          kp.perm.finish();
          // estimate-only pass: a few tiles per cluster (so that no row fills its buffer), only slice maxima are recorded; it spares
          // the first sample pass the cost of starting from -inf (every score appended, every row pruned after its first tile)
          const int64_t clusters = grid / P.cta_group;
          const int64_t n_e = std::max<int64_t>(2, std::min<int64_t>(std::min<int64_t>(16, kp.perm.n0), 7 * clusters / std::max(kp.num_qt, 1)));
          passes[n_pass++] = {0, n_e, std::min(P.k_fetch, sample_rank(P.k_fetch, n_e, P.n_tiles)), 1};
          passes[n_pass++] = {0, kp.perm.n0, std::min(P.k_fetch, sample_rank(P.k_fetch, kp.perm.n0, P.n_tiles)), 0};
          passes[n_pass++] = {kp.perm.n0, kp.perm.ns, std::min(P.k_fetch, sample_rank(P.k_fetch, kp.perm.ns, P.n_tiles)), 0};
          passes[n_pass++] = {kp.perm.ns, P.n_tiles, P.k_fetch, 0};
        } else {
          passes[n_pass++] = {0, P.n_tiles, P.k_fetch, 0};
        }
      }
      for (int ps = 0; ps < n_pass; ++ps) {
        kp.tile_begin = passes[ps].begin; kp.tile_end = passes[ps].end;
        kp.sink.k = passes[ps].k;
        kp.est_only = passes[ps].est_only;
        kp.sync_ctr = tau_g + 2 * nq_pad + ps;   // one round counter per pass (all zeroed above)
        rc = launch_k1(c_dt, P.cta_group, q16, bn, q16_ld, c_rows, c_rows2, idx->ld, kp, grid, stream);
        if (rc) return rc;
        idx->stats.kernel_launches++;
        if (ps + 1 < n_pass) {
          rc = launch_select_tau(arena, qcnt, arena_stride, passes[ps].k, tau_g, bn, passes[ps].est_only, stream);
          if (rc) return rc;
          idx->stats.kernel_launches++;
        }
      }
      kp.est_only = 0;
      if (n_pass > 1) {
        // verification + fall-back: the launch below returns at once unless some query holds fewer than min(k', n) candidates
        uint32_t* flag = tau_g + 2 * nq_pad + 8;
        rc = launch_verify_counts(qcnt, tau_g, bn, static_cast<int>(std::min<int64_t>(P.k_fetch, idx->n)), flag, stream);
        if (rc) return rc;
        kp.tile_begin = 0; kp.tile_end = P.n_tiles; kp.sink.k = P.k_fetch;
        kp.run_flag = flag;
        kp.sync_ctr = tau_g + 2 * nq_pad + 5;
        rc = launch_k1(c_dt, P.cta_group, q16, bn, q16_ld, c_rows, c_rows2, idx->ld, kp, grid, stream);
        if (rc) return rc;
        idx->stats.kernel_launches += 2;
      }
    }

    if (P.path == NXF_PATH_TENSOR && (env_int("NXF_K1_DEBUG", 0) & 32)) {
      unsigned long long h[4] = {0, 0, 0, 0};
      NXF_CUDA(cudaMemcpyAsync(h, tau_g + 2 * nq_pad + 16, sizeof(h), cudaMemcpyDeviceToHost, stream));
      NXF_CUDA(cudaStreamSynchronize(stream));
      fprintf(stderr, "[nxf K1 prof] prunes=%llu prune_cycles/warp=%.0f end_merge_cycles/warp=%.0f cycles/prune=%.0f (epilogue warps %llu)\n", h[0],
              h[3] ? double(h[1]) / h[3] : 0.0, h[3] ? double(h[2]) / h[3] : 0.0, h[0] ? double(h[1]) / h[0] : 0.0, h[3]);
    }
    if (pe0) {
      NXF_CUDA(cudaEventRecord(pe1, stream));
      idx->prof_events.emplace_back(pe0, pe1);
    }
    MergeArenaParams mp;
    mp.arena = arena; mp.qcnt = qcnt; mp.arena_stride = arena_stride; mp.k_fetch = P.k_fetch; mp.k_out = k; mp.metric = idx->metric;
    mp.rescore = P.rescore; mp.q = q_f32; mp.ldq = idx->ld; mp.corpus = idx->data; mp.ld = idx->ld; mp.dim = idx->dim;
    mp.id_offset = idx->id_offset;
    mp.cert = P.cert; mp.q_round = P.q_round; mp.eps_rel = P.eps_rel; mp.eps_abs = P.eps_abs; mp.max_norm2 = idx->max_norm2; mp.n_rows = idx->n;
    mp.cert_fail = P.cert ? cert_w + b0 : nullptr; mp.cert_count = P.cert ? cert_w + nq : nullptr;
    mp.out_scores = out_on_device ? (out_scores ? out_scores + b0 * k : nullptr) : o_scores;
    mp.out_ids = out_on_device ? (out_ids ? out_ids + b0 * k : nullptr) : o_ids;
    mp.out_packed = out_packed ? out_packed + b0 * k : nullptr;
    rc = launch_merge_arena(idx->store_dtype, mp, bn, stream);
    if (rc) return rc;
    idx->stats.kernel_launches++;

    if (!out_on_device) {
      NXF_CUDA(cudaMemcpyAsync(out_scores + b0 * k, o_scores, static_cast<size_t>(bn) * k * sizeof(float), cudaMemcpyDeviceToHost, stream));
      NXF_CUDA(cudaMemcpyAsync(out_ids + b0 * k, o_ids, static_cast<size_t>(bn) * k * sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
    }
    if (!out_on_device || !q_on_device || multi_batch) NXF_CUDA(cudaStreamSynchronize(stream));
  }
  idx->stats.searches++;
  idx->stats.last_path = P.path;
  idx->stats.last_k_fetch = P.k_fetch;
  idx->stats.last_chunks = P.chunks;
  idx->stats.last_grid = P.grid;
  if (!P.cert) return NXF_OK;

  // ---- certificate: how many queries could not be certified?  (4 bytes back; the stream is drained here, which is what makes a
  // certified device-in / device-out search synchronous on its stream) ----
  uint32_t n_fail = 0;
  NXF_CUDA(cudaMemcpyAsync(&n_fail, cert_w + nq, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  NXF_CUDA(cudaStreamSynchronize(stream));
  idx->stats.cert_checked += nq;
  if (n_fail == 0) return NXF_OK;

  // ---- widen and retry: the failed queries go through the next level (segmented K1, then the exact fp32 kernels) ----
  std::vector<uint32_t> h_fail(static_cast<size_t>(nq));
  NXF_CUDA(cudaMemcpy(h_fail.data(), cert_w, static_cast<size_t>(nq) * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  std::vector<int64_t> rows;
  rows.reserve(n_fail);
  for (int64_t i = 0; i < nq; ++i) if (h_fail[static_cast<size_t>(i)]) rows.push_back(i);
  const int64_t nf = static_cast<int64_t>(rows.size());
  if (nf == 0) return NXF_OK;
  const int next_level = P.level == 1 ? 2 : 3;
  if (next_level == 2) idx->stats.cert_retry_l2 += nf; else idx->stats.cert_retry_l3 += nf;
  const int next_flags = (flags & ~(kForceMask | kFlagLevel2 | kFlagLevel3 | NXF_SEARCH_TENSOR_1CTA)) | (next_level == 2 ? kFlagLevel2 : kFlagLevel3);
  const nxf_stats keep_stats = idx->stats;
  const size_t row_bytes_q = static_cast<size_t>(q_ld) * qsz;
  if (!q_on_device && !out_on_device) {
    // host in, host out: compact on the host
    std::vector<unsigned char> hq(static_cast<size_t>(nf) * row_bytes_q);
    for (int64_t i = 0; i < nf; ++i) std::memcpy(hq.data() + static_cast<size_t>(i) * row_bytes_q, static_cast<const unsigned char*>(q) + static_cast<size_t>(rows[i]) * row_bytes_q, row_bytes_q);
    std::vector<float> hs(static_cast<size_t>(nf) * k);
    std::vector<int64_t> hi(static_cast<size_t>(nf) * k);
    rc = search_locked(idx, hq.data(), nf, q_dtype, q_ld, 0, k, hs.data(), hi.data(), nullptr, 0, next_flags, stream);
    if (rc) return rc;
    for (int64_t i = 0; i < nf; ++i) {
      std::memcpy(out_scores + rows[i] * k, hs.data() + i * k, static_cast<size_t>(k) * sizeof(float));
      std::memcpy(out_ids + rows[i] * k, hi.data() + i * k, static_cast<size_t>(k) * sizeof(int64_t));
    }
  } else {
    // any device side: compacted queries and the retry's results live in temporary device buffers (a rare path: plain cudaMalloc)
    unsigned char* d_q = nullptr; int64_t* d_rows = nullptr; float* d_s = nullptr; int64_t* d_i = nullptr; Entry* d_p = nullptr;
    auto cleanup = [&]() { cudaFree(d_q); cudaFree(d_rows); cudaFree(d_s); cudaFree(d_i); cudaFree(d_p); };
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&d_q), static_cast<size_t>(nf) * row_bytes_q);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&d_rows), static_cast<size_t>(nf) * sizeof(int64_t));
    if (e == cudaSuccess && out_packed) e = cudaMalloc(reinterpret_cast<void**>(&d_p), static_cast<size_t>(nf) * k * sizeof(Entry));
    if (e == cudaSuccess && !out_packed) e = cudaMalloc(reinterpret_cast<void**>(&d_s), static_cast<size_t>(nf) * k * sizeof(float));
    if (e == cudaSuccess && !out_packed) e = cudaMalloc(reinterpret_cast<void**>(&d_i), static_cast<size_t>(nf) * k * sizeof(int64_t));
    if (e != cudaSuccess) { cleanup(); return cuda_fail(e, "cudaMalloc(retry buffers)", __FILE__, __LINE__); }
    e = cudaMemcpyAsync(d_rows, rows.data(), static_cast<size_t>(nf) * sizeof(int64_t), cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) { cleanup(); return cuda_fail(e, "cudaMemcpyAsync(retry rows)", __FILE__, __LINE__); }
    if (q_on_device) {
      rc = launch_gather_rows(q, static_cast<int64_t>(row_bytes_q), d_rows, nf, d_q, stream);
    } else {
      std::vector<unsigned char> hq(static_cast<size_t>(nf) * row_bytes_q);
      for (int64_t i = 0; i < nf; ++i) std::memcpy(hq.data() + static_cast<size_t>(i) * row_bytes_q, static_cast<const unsigned char*>(q) + static_cast<size_t>(rows[i]) * row_bytes_q, row_bytes_q);
      e = cudaMemcpy(d_q, hq.data(), hq.size(), cudaMemcpyHostToDevice);
      rc = e == cudaSuccess ? NXF_OK : cuda_fail(e, "cudaMemcpy(retry queries)", __FILE__, __LINE__);
    }
    if (rc == NXF_OK) rc = search_locked(idx, d_q, nf, q_dtype, q_ld, 1, k, d_s, d_i, d_p, 1, next_flags, stream);
    if (rc == NXF_OK) {
      if (out_on_device) {
        rc = launch_scatter_results(d_s, d_i, d_p, d_rows, nf, k, out_scores, out_ids, out_packed, stream);
      } else {
        std::vector<float> hs(static_cast<size_t>(nf) * k);
        std::vector<int64_t> hi(static_cast<size_t>(nf) * k);
        e = cudaMemcpyAsync(hs.data(), d_s, hs.size() * sizeof(float), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(hi.data(), d_i, hi.size() * sizeof(int64_t), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "retry results", __FILE__, __LINE__);
        else for (int64_t i = 0; i < nf; ++i) {
          std::memcpy(out_scores + rows[i] * k, hs.data() + i * k, static_cast<size_t>(k) * sizeof(float));
          std::memcpy(out_ids + rows[i] * k, hi.data() + i * k, static_cast<size_t>(k) * sizeof(int64_t));
        }
      }
    }
    cudaStreamSynchronize(stream);
    cleanup();
    if (rc) return rc;
  }
  // the primary level describes the search in the stats; the retry only adds its counters
  idx->stats.last_path = keep_stats.last_path; idx->stats.last_k_fetch = keep_stats.last_k_fetch;
  idx->stats.last_chunks = keep_stats.last_chunks; idx->stats.last_grid = keep_stats.last_grid;
  idx->stats.searches = keep_stats.searches;
  return NXF_OK;
}

static int search_impl(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int q_on_device, int k, float* out_scores,
                       int64_t* out_ids, Entry* out_packed, int out_on_device, int flags, void* stream_) {
  NXF_REQUIRE(idx != nullptr, "nxf_search: NULL index");
  NXF_REQUIRE(nq >= 0, "nxf_search: negative query count");
  NXF_REQUIRE(k >= 1 && k <= NXF_MAX_K, "nxf_search: k=%d outside [1, %d]", k, NXF_MAX_K);
  if (nq == 0) return NXF_OK;
  NXF_REQUIRE(q != nullptr && ((out_scores != nullptr && out_ids != nullptr) || out_packed != nullptr), "nxf_search: NULL buffer");
  int rc = check_dtype(q_dtype, "nxf_search(q_dtype)");
  if (rc) return rc;
  if (q_ld == 0) q_ld = idx->dim;
  NXF_REQUIRE(q_ld >= idx->dim, "nxf_search: query stride %lld smaller than dim %d", (long long)q_ld, idx->dim);
  NXF_REQUIRE((flags & ~(kForceMask | NXF_SEARCH_NO_RESCORE | NXF_SEARCH_TENSOR_1CTA)) == 0, "nxf_search: unknown flag bits 0x%x", flags);
  const int nforce = !!(flags & NXF_SEARCH_FORCE_EXACT) + !!(flags & NXF_SEARCH_FORCE_STREAM) + !!(flags & NXF_SEARCH_FORCE_TENSOR) +
                     !!(flags & NXF_SEARCH_FORCE_TSTREAM);
  NXF_REQUIRE(nforce <= 1, "nxf_search: more than one NXF_SEARCH_FORCE_* flag");

  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  UseScope use(idx, stream);
  return search_locked(idx, q, nq, q_dtype, q_ld, q_on_device, k, out_scores, out_ids, out_packed, out_on_device, flags, stream);
}

extern "C" {

int nxf_search(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int q_on_device, int k, float* out_scores,
               int64_t* out_ids, int out_on_device, int flags, void* stream_) {
  if (nq > 0) NXF_REQUIRE(out_scores != nullptr && out_ids != nullptr, "nxf_search: NULL buffer");
  return search_impl(idx, q, nq, q_dtype, q_ld, q_on_device, k, out_scores, out_ids, nullptr, out_on_device, flags, stream_);
}

int nxf_search_packed(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int k, nxf_entry* out, int flags, void* stream_) {
  if (nq > 0) NXF_REQUIRE(out != nullptr, "nxf_search_packed: NULL buffer");
  static_assert(sizeof(nxf_entry) == sizeof(Entry), "packed entry layout");
  return search_impl(idx, q, nq, q_dtype, q_ld, 1, k, nullptr, nullptr, reinterpret_cast<Entry*>(out), 1, flags, stream_);
}

// ------------------------------------------------------------------------------------------
// per-query candidate lists (KS, search_subset.cu)
// ------------------------------------------------------------------------------------------
int nxf_search_subsets(nxf_index* idx, const void* q, int64_t nq, int q_dtype, int64_t q_ld, int q_on_device, const int64_t* offsets,
                       const int64_t* rows, int64_t total, int lists_on_device, int k, float* out_scores, int64_t* out_pos,
                       int out_on_device, void* stream_) {
  NXF_REQUIRE(idx != nullptr, "nxf_search_subsets: NULL index");
  NXF_REQUIRE(nq >= 0 && total >= 0, "nxf_search_subsets: negative count (nq=%lld total=%lld)", (long long)nq, (long long)total);
  NXF_REQUIRE(nq < (int64_t(1) << 31), "nxf_search_subsets: nq=%lld too large for one call", (long long)nq);
  NXF_REQUIRE(k >= 1 && k <= NXF_MAX_K, "nxf_search_subsets: k=%d outside [1, %d]", k, NXF_MAX_K);
  if (nq == 0) return NXF_OK;
  NXF_REQUIRE(q != nullptr && offsets != nullptr && out_scores != nullptr && out_pos != nullptr, "nxf_search_subsets: NULL buffer");
  NXF_REQUIRE(rows != nullptr || total == 0, "nxf_search_subsets: NULL row list");
  int rc = check_dtype(q_dtype, "nxf_search_subsets(q_dtype)");
  if (rc) return rc;
  if (q_ld == 0) q_ld = idx->dim;
  NXF_REQUIRE(q_ld >= idx->dim, "nxf_search_subsets: query stride %lld smaller than dim %d", (long long)q_ld, idx->dim);
  if (!lists_on_device) {
    NXF_REQUIRE(offsets[0] == 0 && offsets[nq] == total, "nxf_search_subsets: offsets must run from 0 to total=%lld (got %lld..%lld)",
                (long long)total, (long long)offsets[0], (long long)offsets[nq]);
  }

  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  UseScope use(idx, stream);
  const size_t qsz = dtype_size(q_dtype);

  unsigned char* base = nullptr;
  unsigned char* q_src = nullptr; float* q_f32 = nullptr; int64_t* d_off = nullptr; int64_t* d_rows = nullptr; Entry* scr = nullptr;
  float* o_scores = nullptr; int64_t* o_pos = nullptr; uint32_t* err = nullptr;
  for (int pass = 0; pass < 2; ++pass) {
    Carver cv(base);
    err = cv.take<uint32_t>(64);
    q_src = q_on_device ? nullptr : cv.take<unsigned char>(static_cast<size_t>(nq) * q_ld * qsz);
    q_f32 = cv.take<float>(static_cast<size_t>(nq) * idx->ld);
    if (!lists_on_device) {
      d_off = cv.take<int64_t>(static_cast<size_t>(nq) + 1);
      d_rows = cv.take<int64_t>(static_cast<size_t>(std::max<int64_t>(total, 1)));
    }
    scr = cv.take<Entry>(static_cast<size_t>(std::max<int64_t>(total, 1)));
    if (!out_on_device) {
      o_scores = cv.take<float>(static_cast<size_t>(nq) * k);
      o_pos = cv.take<int64_t>(static_cast<size_t>(nq) * k);
    }
    if (pass == 0) {
      rc = ensure_scratch(idx, cv.off + 256);
      if (rc) return rc;
      base = idx->scratch;
    }
  }
  NXF_CUDA(cudaMemsetAsync(err, 0, sizeof(uint32_t), stream));
  const void* q_dev = q;
  if (!q_on_device) {
    NXF_CUDA(cudaMemcpyAsync(q_src, q, static_cast<size_t>(nq) * q_ld * qsz, cudaMemcpyHostToDevice, stream));
    q_dev = q_src;
  }
  if (!lists_on_device) {
    NXF_CUDA(cudaMemcpyAsync(d_off, offsets, (static_cast<size_t>(nq) + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
    if (total > 0) NXF_CUDA(cudaMemcpyAsync(d_rows, rows, static_cast<size_t>(total) * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
  }
  rc = launch_convert_rows(q_dev, q_dtype, q_ld, q_f32, NXF_F32, idx->ld, nq, idx->dim, stream);
  if (rc) return rc;
  idx->stats.kernel_launches++;

  SubsetParams sp;
  sp.q = q_f32; sp.ldq = idx->ld; sp.corpus = idx->data; sp.ld = idx->ld; sp.dim = idx->dim; sp.n_rows = idx->n;
  sp.offsets = lists_on_device ? offsets : d_off; sp.rows = lists_on_device ? rows : d_rows; sp.total = total; sp.scratch = scr;
  sp.k_out = k; sp.metric = idx->metric;
  sp.out_scores = out_on_device ? out_scores : o_scores; sp.out_pos = out_on_device ? out_pos : o_pos; sp.err = err;
  cudaEvent_t pe0 = nullptr, pe1 = nullptr;
  if (idx->profiling) {
    NXF_CUDA(cudaEventCreate(&pe0));
    NXF_CUDA(cudaEventCreate(&pe1));
    NXF_CUDA(cudaEventRecord(pe0, stream));
  }
  int64_t min_len = -1, max_len = -1;   // known for host lists: lets the launcher skip the kernel (CTA / warp per list) nobody needs
  if (!lists_on_device) {
    min_len = INT64_MAX; max_len = 0;
    for (int64_t i = 0; i < nq; ++i) {
      const int64_t len = offsets[i + 1] - offsets[i];
      min_len = std::min(min_len, len); max_len = std::max(max_len, len);
    }
    if (min_len < 0) { min_len = -1; max_len = -1; }   // malformed: both kernels run and report it
  }
  sp.short_max = -1;
  rc = launch_subset_search(idx->store_dtype, sp, nq, min_len, max_len, stream);
  if (rc) return rc;
  idx->stats.kernel_launches++;
  if (pe0) {
    NXF_CUDA(cudaEventRecord(pe1, stream));
    idx->prof_events.emplace_back(pe0, pe1);
  }
  if (!out_on_device) {
    NXF_CUDA(cudaMemcpyAsync(out_scores, o_scores, static_cast<size_t>(nq) * k * sizeof(float), cudaMemcpyDeviceToHost, stream));
    NXF_CUDA(cudaMemcpyAsync(out_pos, o_pos, static_cast<size_t>(nq) * k * sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  }
  uint32_t h_err = 0;
  NXF_CUDA(cudaMemcpyAsync(&h_err, err, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  NXF_CUDA(cudaStreamSynchronize(stream));
  idx->stats.searches++;
  idx->stats.last_path = NXF_PATH_SUBSET;
  idx->stats.last_k_fetch = k;
  idx->stats.last_grid = static_cast<int32_t>(std::min<int64_t>(nq, INT32_MAX));
  NXF_REQUIRE(!(h_err & 2u), "nxf_search_subsets: offsets are not a non-decreasing CSR inside [0, total=%lld]", (long long)total);
  NXF_REQUIRE(!(h_err & 1u), "nxf_search_subsets: a candidate row id lies outside [0, ntotal=%lld)", (long long)idx->n);
  return NXF_OK;
}

// ------------------------------------------------------------------------------------------
// merge + pool (device pointers only)
// ------------------------------------------------------------------------------------------
int nxf_merge_topk(const float* scores, const int64_t* ids, int n_lists, int64_t nq, int k, int k_out, int metric, float* out_scores,
                   int64_t* out_ids, int device, void* stream_) {
  NXF_REQUIRE(scores && ids && out_scores && out_ids, "nxf_merge_topk: NULL buffer");
  NXF_REQUIRE(n_lists >= 1 && n_lists <= 4096, "nxf_merge_topk: n_lists=%d outside [1, 4096]", n_lists);
  NXF_REQUIRE(k >= 1 && k_out >= 1 && k_out <= NXF_MAX_K, "nxf_merge_topk: k=%d k_out=%d (max %d)", k, k_out, NXF_MAX_K);
  NXF_REQUIRE(static_cast<int64_t>(n_lists) * k < (int64_t(1) << 30), "nxf_merge_topk: too many candidates per query");
  NXF_REQUIRE(metric == NXF_METRIC_IP || metric == NXF_METRIC_L2, "nxf_merge_topk: unknown metric %d", metric);
  NXF_REQUIRE(nq >= 0, "nxf_merge_topk: negative nq");
  DeviceGuard g(device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  MergeListsParams p;
  p.scores = scores; p.ids = ids; p.packed = nullptr; p.n_lists = n_lists; p.nq = nq; p.k = k; p.k_out = k_out; p.metric = metric;
  p.out_scores = out_scores; p.out_ids = out_ids;
  return launch_merge_lists(p, static_cast<cudaStream_t>(stream_));
}

int nxf_merge_packed(const nxf_entry* lists, const int64_t* list_offsets, int n_lists, int64_t nq, int k, int k_out, int metric,
                     float* out_scores, int64_t* out_ids, int device, void* stream_) {
  NXF_REQUIRE(lists && list_offsets && out_scores && out_ids, "nxf_merge_packed: NULL buffer");
  NXF_REQUIRE(n_lists >= 1 && n_lists <= kMaxPackedLists, "nxf_merge_packed: n_lists=%d outside [1, %d]", n_lists, kMaxPackedLists);
  NXF_REQUIRE(k >= 1 && k_out >= 1 && k_out <= NXF_MAX_K, "nxf_merge_packed: k=%d k_out=%d (max %d)", k, k_out, NXF_MAX_K);
  NXF_REQUIRE(metric == NXF_METRIC_IP || metric == NXF_METRIC_L2, "nxf_merge_packed: unknown metric %d", metric);
  NXF_REQUIRE(nq >= 0, "nxf_merge_packed: negative nq");
  DeviceGuard g(device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  MergeListsParams p;
  p.scores = nullptr; p.ids = nullptr; p.packed = reinterpret_cast<const Entry*>(lists);
  for (int l = 0; l < kMaxPackedLists; ++l) p.offs[l] = l < n_lists ? list_offsets[l] : 0;
  p.n_lists = n_lists; p.nq = nq; p.k = k; p.k_out = k_out; p.metric = metric;
  p.out_scores = out_scores; p.out_ids = out_ids;
  return launch_merge_lists(p, static_cast<cudaStream_t>(stream_));
}

int nxf_positive_rank(const float* pos_scores, const float* topk_scores, int64_t nq, int k, int32_t* rank, uint8_t* pred, int device,
                      void* stream_) {
  NXF_REQUIRE(pos_scores && topk_scores && rank, "nxf_positive_rank: NULL buffer");
  NXF_REQUIRE(nq >= 0 && k >= 1, "nxf_positive_rank: bad shape nq=%lld k=%d", (long long)nq, k);
  DeviceGuard g(device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  return launch_positive_rank(pos_scores, topk_scores, nq, k, rank, pred, static_cast<cudaStream_t>(stream_));
}

int nxf_pool_normalize(const void* hidden, int h_dtype, const int64_t* mask, int64_t B, int64_t L, int64_t H, int pooling, int normalize,
                       float eps, void* out, int out_dtype, int device, void* stream_) {
  NXF_REQUIRE(hidden != nullptr && out != nullptr, "nxf_pool_normalize: NULL buffer");
  NXF_REQUIRE(B >= 0 && L >= 1 && H >= 1, "nxf_pool_normalize: bad shape B=%lld L=%lld H=%lld", (long long)B, (long long)L, (long long)H);
  // base.py:378-379: anything but the known rules raises NotImplementedError in the reference
  NXF_REQUIRE(pooling >= NXF_POOL_CLS && pooling <= NXF_POOL_LAST_ARGMAX, "nxf_pool_normalize: pooling method %d not implemented", pooling);
  NXF_REQUIRE(pooling == NXF_POOL_CLS || mask != nullptr, "nxf_pool_normalize: this pooling rule needs the attention mask");
  int rc = check_dtype(h_dtype, "nxf_pool_normalize(h_dtype)");
  if (rc) return rc;
  rc = check_dtype(out_dtype, "nxf_pool_normalize(out_dtype)");
  if (rc) return rc;
  DeviceGuard g(device);
  if (!g.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
  return launch_pool(hidden, h_dtype, mask, B, L, H, pooling, normalize, eps, out, out_dtype, static_cast<cudaStream_t>(stream_));
}

}  // extern "C"
