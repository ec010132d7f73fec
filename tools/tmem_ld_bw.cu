// tmem_ld_bw.cu -- how fast can an SM read its tensor memory?  (B200, sm_100a)
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_ld_bw tools/tmem_ld_bw.cu && /tmp/tmem_ld_bw
//
// Why it matters here: K1 on 64-d rows (config 5, full batch) and K2T on 128-byte rows with 128 queries produce one fp32 score per
// 2 FLOP-pairs of input -- every score has to come out of TMEM through tcgen05.ld before it can be compared with a threshold.  Both
// kernels level off at ~14.5 scores per clock per SM whatever the number of epilogue warps; this program measures the ceiling of
// that path alone: W warps per CTA (one CTA per SM) do nothing but tcgen05.ld.32x32b.xN + tcgen05.wait::ld in a loop.
// Output: bytes per clock per SM and fp32 values per clock per SM, per (warps, columns per load, loads in flight).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

template <int X>
__device__ __forceinline__ void ld(uint32_t taddr, uint32_t& sink) {
  if constexpr (X == 32) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    sink ^= r[0] ^ r[31];
  } else {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    sink ^= r[0] ^ r[15];
  }
}

// two loads in flight before one wait
template <int X>
__device__ __forceinline__ void ld2(uint32_t t0, uint32_t t1, uint32_t& sink) {
  static_assert(X == 32, "x32 only");
  uint32_t a[32], b[32];
#define LD32(r, t)                                                                                                             \
  asm volatile(                                                                                                                \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                                \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                                \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                                \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), \
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),   \
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),   \
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                                                                     \
      : "r"(t) : "memory")
  LD32(a, t0);
  LD32(b, t1);
#undef LD32
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  sink ^= a[0] ^ a[31] ^ b[0] ^ b[31];
}

template <int X, int INFLIGHT>
__global__ void __launch_bounds__(512, 1) tmem_read_kernel(int iters, unsigned long long* cycles, uint32_t* out) {
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = *reinterpret_cast<volatile uint32_t*>(&tmem_ptr) + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  uint32_t sink = 0;
  __syncthreads();
  const long long c0 = clock64();
  for (int i = 0; i < iters; ++i) {
    // walk the 512 columns: 16 (x32) or 32 (x16) loads per sweep
    if constexpr (INFLIGHT == 2) {
#pragma unroll 4
      for (int c = 0; c < 512; c += 2 * X) ld2<X>(base + c, base + c + X, sink);
    } else {
#pragma unroll 4
      for (int c = 0; c < 512; c += X) ld<X>(base + c, sink);
    }
  }
  __syncthreads();
  const long long c1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = static_cast<unsigned long long>(c1 - c0);
  if (sink == 0x12345678u) out[0] = sink;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*reinterpret_cast<volatile uint32_t*>(&tmem_ptr)), "r"(512) : "memory");
}

template <int X, int INFLIGHT>
static void run(int warps, int sms, unsigned long long* d_cyc, uint32_t* d_out) {
  const int iters = 2000;
  tmem_read_kernel<X, INFLIGHT><<<sms, warps * 32>>>(10, d_cyc, d_out);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  tmem_read_kernel<X, INFLIGHT><<<sms, warps * 32>>>(iters, d_cyc, d_out);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("warps %d x%d: CUDA error %s\n", warps, X, cudaGetErrorString(e)); return; }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long h[256];
  cudaMemcpy(h, d_cyc, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost);
  double cyc = 0;
  for (int i = 0; i < sms; ++i) cyc += static_cast<double>(h[i]);
  cyc /= sms;
  const double bytes_per_sm = static_cast<double>(iters) * 512.0 * 32 * 4 * warps;   // every warp reads its 32 lanes x 512 columns x 4 B per sweep
  printf("{\"warps\": %2d, \"cols_per_ld\": %d, \"lds_in_flight\": %d, \"bytes_per_clk_per_sm\": %.1f, \"fp32_per_clk_per_sm\": %.2f, "
         "\"avg_cycles\": %.0f, \"ms\": %.3f, \"GHz\": %.3f}\n",
         warps, X, INFLIGHT, bytes_per_sm / cyc, bytes_per_sm / cyc / 4.0, cyc, ms, cyc / (ms * 1e6));
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount;
  printf("device: %s, %d SMs\n", prop.name, sms);
  unsigned long long* d_cyc; uint32_t* d_out;
  cudaMalloc(&d_cyc, 256 * sizeof(unsigned long long));
  cudaMalloc(&d_out, 64);
  for (int warps : {4, 8, 12, 16}) {
    run<32, 1>(warps, sms, d_cyc, d_out);
    run<32, 2>(warps, sms, d_cyc, d_out);
    run<16, 1>(warps, sms, d_cyc, d_out);
  }
  return 0;
}
