// convert.cu -- row conversion kernels used by nxf_add / nxf_search / nxf_reconstruct:
// any of {f32, f16, bf16} -> any of them, re-pitching rows to the index's padded row stride and
// zero-filling the padding, plus 0.5*|row|^2 for the L2 metric (computed from the STORED values).
#include <algorithm>

#include "common.cuh"

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

// fp32 rows -> two bf16 planes per row, [hi(ld) | mid(ld)]: hi = bf16(x), mid = bf16(x - hi).  x = hi + mid up to 2^-17 |x|, so the three
// bf16 products q_hi*c_hi + q_hi*c_mid + q_mid*c_hi reproduce an fp32 inner product to ~2^-16 per term -- the operand layout of the
// tensor-core path over fp32 storage (DESIGN.md "fp32 on tcgen05").
__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ src, int64_t ld, __nv_bfloat16* __restrict__ dst, int64_t n) {
  const int64_t total = n * ld;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / ld;
    const int64_t d = i - r * ld;
    const float x = src[i];
    const __nv_bfloat16 hi = __float2bfloat16_rn(x);
    const __nv_bfloat16 mid = __float2bfloat16_rn(x - __bfloat162float(hi));
    dst[r * 2 * ld + d] = hi;
    dst[r * 2 * ld + ld + d] = mid;
  }
}

int launch_split_rows(const float* src, int64_t ld, void* dst, int64_t n, cudaStream_t stream) {
  if (n <= 0) return NXF_OK;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div64(n * ld, 256), 148 * 16));
  split_rows_kernel<<<grid, 256, 0, stream>>>(src, ld, static_cast<__nv_bfloat16*>(dst), n);
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
