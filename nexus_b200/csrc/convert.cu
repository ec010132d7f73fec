// convert.cu -- row conversion kernels used by nxf_add / nxf_search / nxf_reconstruct:
// any of {f32, f16, bf16} -> any of them, re-pitching rows to the index's padded row stride and
// zero-filling the padding, plus 0.5*|row|^2 for the L2 metric (computed from the STORED values).
#include <algorithm>

#include "params.cuh"

namespace nxf {

template <typename S, typename D>
__global__ void __launch_bounds__(256) convert_rows_kernel(const S* __restrict__ src, int64_t src_ld, D* __restrict__ dst,
                                                           int64_t dst_ld, int64_t n, int dim) {
  const int64_t total = n * dst_ld;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / dst_ld;
    const int d = static_cast<int>(i - r * dst_ld);
    const float v = d < dim ? to_f32<S>(src[r * src_ld + d]) : 0.f;
    dst[i] = from_f32<D>(v);
  }
}

template <typename S>
static int convert_dispatch_dst(const void* src, int64_t src_ld, void* dst, int dst_dtype, int64_t dst_ld, int64_t n, int dim,
                                cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int64_t total = n * dst_ld;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(total, 256), 148 * 16));
  switch (dst_dtype) {
    case NXF_F32: convert_rows_kernel<S, float><<<grid, 256, 0, stream>>>((const S*)src, src_ld, (float*)dst, dst_ld, n, dim); break;
    case NXF_F16: convert_rows_kernel<S, __half><<<grid, 256, 0, stream>>>((const S*)src, src_ld, (__half*)dst, dst_ld, n, dim); break;
    case NXF_BF16: convert_rows_kernel<S, __nv_bfloat16><<<grid, 256, 0, stream>>>((const S*)src, src_ld, (__nv_bfloat16*)dst, dst_ld, n, dim); break;
    default: set_error("convert: bad destination dtype %d", dst_dtype); return NXF_EINVAL;
  }
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

int launch_convert_rows(const void* src, int src_dtype, int64_t src_ld, void* dst, int dst_dtype, int64_t dst_ld, int64_t n, int dim,
                        cudaStream_t stream) {
  switch (src_dtype) {
    case NXF_F32: return convert_dispatch_dst<float>(src, src_ld, dst, dst_dtype, dst_ld, n, dim, stream);
    case NXF_F16: return convert_dispatch_dst<__half>(src, src_ld, dst, dst_dtype, dst_ld, n, dim, stream);
    case NXF_BF16: return convert_dispatch_dst<__nv_bfloat16>(src, src_ld, dst, dst_dtype, dst_ld, n, dim, stream);
  }
  set_error("convert: bad source dtype %d", src_dtype);
  return NXF_EINVAL;
}

// fp32 rows -> two 16-bit planes per row, [hi(ld) | mid(ld)]: hi = T(x), mid = T(x - hi), so x = hi + mid up to 2^-17 |x| (bf16; less for
// f16).  The QUERY operand of the segmented retry level of the certified search (K1Params::seg_mode): with the corpus planes of
// fp32 storage the three products q_hi*c_hi + q_hi*c_mid + q_mid*c_hi reproduce an fp32 inner product to ~2^-16 per term; with
// 16-bit storage q_hi*c + q_mid*c removes the query rounding.
template <typename T>
__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ src, int64_t ld, T* __restrict__ dst, int64_t n) {
  const int64_t total = n * ld;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / ld;
    const int64_t d = i - r * ld;
    const float x = src[i];
    const T hi = from_f32<T>(x);
    const T mid = from_f32<T>(x - to_f32<T>(hi));
    dst[r * 2 * ld + d] = hi;
    dst[r * 2 * ld + ld + d] = mid;
  }
}

int launch_split_rows(const float* src, int64_t ld, void* dst, int64_t n, int dst_dtype, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n * ld, 256), 148 * 16));
  if (dst_dtype == NXF_BF16) split_rows_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(src, ld, static_cast<__nv_bfloat16*>(dst), n);
  else if (dst_dtype == NXF_F16) split_rows_kernel<__half><<<grid, 256, 0, stream>>>(src, ld, static_cast<__half*>(dst), n);
  else { set_error("split_rows: bad plane dtype %d", dst_dtype); return NXF_EINVAL; }
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// fp32 rows -> the mid plane alone (pitch ld): mid = bf16(x - bf16(x)).  The hi plane of fp32 storage is a plain bf16 conversion.
__global__ void __launch_bounds__(256) mid_rows_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t total) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float x = src[i];
    dst[i] = __float2bfloat16_rn(x - __bfloat162float(__float2bfloat16_rn(x)));
  }
}

int launch_mid_rows(const float* src, int64_t ld, void* dst, int64_t n, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n * ld, 256), 148 * 16));
  mid_rows_kernel<<<grid, 256, 0, stream>>>(src, static_cast<__nv_bfloat16*>(dst), n * ld);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// max over rows of |row|^2 (from the STORED values), kept as float bits in one device word: non-negative floats order like their
// bit patterns, so atomicMax on uint32 is exact.  The certificate of the candidate pass bounds every row's score error by it.
template <typename T>
__global__ void __launch_bounds__(256) row_norm_max_kernel(const T* __restrict__ rows, int64_t ld, int64_t n, int dim, uint32_t* __restrict__ out) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  float best = 0.f;
  for (int64_t r = warp; r < n; r += nwarps) {
    float s = 0.f;
    for (int d = lane; d < dim; d += 32) { const float v = to_f32<T>(rows[r * ld + d]); s = fmaf(v, v, s); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    best = fmaxf(best, s);
  }
  if (lane == 0 && best > 0.f) atomicMax(out, __float_as_uint(best));
}

int launch_row_norm_max(const void* rows, int dtype, int64_t ld, int64_t n, int dim, uint32_t* max_norm2, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n, 8), 148 * 8));
  switch (dtype) {
    case NXF_F32: row_norm_max_kernel<float><<<grid, 256, 0, stream>>>((const float*)rows, ld, n, dim, max_norm2); break;
    case NXF_F16: row_norm_max_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)rows, ld, n, dim, max_norm2); break;
    case NXF_BF16: row_norm_max_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)rows, ld, n, dim, max_norm2); break;
    default: set_error("row_norm_max: bad dtype %d", dtype); return NXF_EINVAL;
  }
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// dst[i] = src row rows[i] (row_bytes each, a multiple of 2): compacts the queries that failed the certificate for the retry level
__global__ void __launch_bounds__(256) gather_rows_kernel(const unsigned char* __restrict__ src, int64_t row_bytes, const int64_t* __restrict__ rows,
                                                          int64_t n, unsigned char* __restrict__ dst) {
  const int64_t total = n * (row_bytes >> 1);
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / (row_bytes >> 1), o = i - r * (row_bytes >> 1);
    reinterpret_cast<uint16_t*>(dst)[i] = reinterpret_cast<const uint16_t*>(src + rows[r] * row_bytes)[o];
  }
}

int launch_gather_rows(const void* src, int64_t row_bytes, const int64_t* rows, int64_t n, void* dst, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n * (row_bytes >> 1), 256), 148 * 16));
  gather_rows_kernel<<<grid, 256, 0, stream>>>(static_cast<const unsigned char*>(src), row_bytes, rows, n, static_cast<unsigned char*>(dst));
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

// result rows of the retry level -> their places in the caller's output (row i of src goes to row rows[i] of dst)
__global__ void __launch_bounds__(256) scatter_results_kernel(const float* __restrict__ ss, const int64_t* __restrict__ si, const Entry* __restrict__ sp,
                                                              const int64_t* __restrict__ rows, int64_t n, int k, float* __restrict__ ds,
                                                              int64_t* __restrict__ di, Entry* __restrict__ dp) {
  const int64_t total = n * k;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / k, j = i - r * k;
    const int64_t o = rows[r] * k + j;
    if (dp != nullptr) dp[o] = sp[i];
    else { ds[o] = ss[i]; di[o] = si[i]; }
  }
}

int launch_scatter_results(const float* src_scores, const int64_t* src_ids, const Entry* src_packed, const int64_t* rows, int64_t n, int k,
                           float* dst_scores, int64_t* dst_ids, Entry* dst_packed, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n * k, 256), 148 * 16));
  scatter_results_kernel<<<grid, 256, 0, stream>>>(src_scores, src_ids, src_packed, rows, n, k, dst_scores, dst_ids, dst_packed);
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

template <typename T>
__global__ void __launch_bounds__(256) half_norm_kernel(const T* __restrict__ rows, int64_t ld, int64_t n, int dim, float* __restrict__ out) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    float s = 0.f;
    for (int d = lane; d < dim; d += 32) { const float v = to_f32<T>(rows[r * ld + d]); s = fmaf(v, v, s); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    if (lane == 0) out[r] = 0.5f * s;
  }
}

int launch_half_norm(const void* rows, int dtype, int64_t ld, int64_t n, int dim, float* out, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n, 8), 148 * 8));
  switch (dtype) {
    case NXF_F32: half_norm_kernel<float><<<grid, 256, 0, stream>>>((const float*)rows, ld, n, dim, out); break;
    case NXF_F16: half_norm_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)rows, ld, n, dim, out); break;
    case NXF_BF16: half_norm_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)rows, ld, n, dim, out); break;
    default: set_error("half_norm: bad dtype %d", dtype); return NXF_EINVAL;
  }
  NXF_CUDA(cudaGetLastError());
  return NXF_OK;
}

}  // namespace nxf
