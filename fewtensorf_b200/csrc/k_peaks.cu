// k_peaks.cu — measurement aids, built into their OWN small library (libftrf_peaks.so; only bench.py and
// scripts/measure_peaks.py load it): the two roofline denominators MEASURED_PEAKS.json does not carry
// (SURVEY.md section 8d "Roofline per kernel").
//   ftrf_peak_l2_read : every thread streams a device buffer that fits L2 (8-40 MB) `iters` times with 16-byte
//                       L2-cached loads (ld.global.cg: no L1), 8 loads in flight per thread -> L2 read bandwidth
//   ftrf_peak_l2_red  : the same sweep with red.global.add.v4.f32 (what the gather backward kernels issue)
//                       -> L2 reduction bandwidth (bytes of gradient reduced per second)
//   ftrf_peak_l1_read : every CTA re-reads a 32 KB slice with ld.global.ca -> aggregate L1 bandwidth (the ceiling of
//                       gathers whose texel reuse is served by L1)
//   ftrf_peak_tf32    : one CTA per SM, one thread issuing back-to-back tcgen05.mma.kind::tf32 M=128 N=256 K=8 from
//                       shared-memory operands (no-swizzle K-major, or SWIZZLE_128B) -> dense tf32 tensor peak
// The host times them with CUDA events; nothing here is on the product path.
#include <cuda_runtime.h>
#include <stdint.h>
#include "tc.cuh"

namespace ftrf {

__global__ void __launch_bounds__(512)
l2_read_kernel(const float4* __restrict__ buf, size_t n_vec, int iters, float* __restrict__ sink) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    size_t i = t;
    for (; i + 7 * stride < n_vec; i += 8 * stride) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(buf + i + u * stride);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += (v[u].x + v[u].y) + (v[u].z + v[u].w);
    }
    for (; i < n_vec; i += stride) { float4 v = __ldcg(buf + i); acc += (v.x + v.y) + (v.z + v.w); }
  }
  if (acc == 1234.5678f) *sink = acc;      // keeps the loads alive
}

__global__ void __launch_bounds__(512)
l2_red_kernel(float* __restrict__ buf, size_t n_vec, int iters) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < iters; ++it)
    for (size_t i = t; i < n_vec; i += stride)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(buf + 4 * i), "f"(1.f), "f"(0.5f), "f"(0.25f), "f"(0.125f) : "memory");
}

// every CTA re-reads its own 32 KB slice (L1-resident after the first pass) with 16-byte ld.global.ca loads
__global__ void __launch_bounds__(512)
l1_read_kernel(const float4* __restrict__ buf, int iters, float* __restrict__ sink) {
  const float4* mine = buf + (size_t)blockIdx.x * 2048;       // 2048 x 16 B = 32 KB
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    float4 v[4];
    const int t = (threadIdx.x + it * 32) & 511;            // a different warp-sized run of the slice every pass
#pragma unroll
    for (int u = 0; u < 4; ++u)
      asm volatile("ld.global.ca.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                   : "l"(mine + t + u * 512) : "memory");
#pragma unroll
    for (int u = 0; u < 4; ++u) acc += (v[u].x + v[u].y) + (v[u].z + v[u].w);
  }
  if (acc == 1234.5678f) *sink = acc;
}

// MODE 0: both operands K-major no-swizzle (the layout the product kernels use); 1: both SWIZZLE_128B K-major
template <int MODE>
__global__ void __launch_bounds__(128, 1) tf32_peak_kernel(int iters) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  float* sm = reinterpret_cast<float*>(smem_raw);
  for (int i = threadIdx.x; i < 40960; i += blockDim.x) sm[i] = 0.f;
  if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (threadIdx.x < 32) tc::tmem_alloc(&slot, 512);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = tc::idesc_tf32(128, 256);
    const uint32_t a0 = tc::smem_u32(sm), b0 = a0 + 32768;
    uint64_t da[4], db[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (MODE == 0) {
        da[s] = tc::smem_desc(a0 + s * 256, 128, 1024);
        db[s] = tc::smem_desc(b0 + s * 256, 128, 1024);
      } else {
        da[s] = tc::smem_desc(a0 + s * 32, 16, 1024) | (2ull << 61);
        db[s] = tc::smem_desc(b0 + s * 32, 16, 1024) | (2ull << 61);
      }
    }
    for (int i = 0; i < iters; i += 4) {
#pragma unroll
      for (int s = 0; s < 4; ++s) tc::mma_tf32_ss(tmem, da[s], db[s], idesc, true);
    }
    tc::mma_commit(&bar);
    tc::mbar_wait(&bar, 0);
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

}  // namespace ftrf

static int last_error() { return cudaGetLastError() == cudaSuccess ? 0 : -3; }

extern "C" int ftrf_peak_l2_read(const void* buf, size_t bytes, int iters, int ctas, void* sink, void* stream) {
  if (!buf || bytes < 16 || iters < 1 || ctas < 1 || !sink) return -1;
  ftrf::l2_read_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>((const float4*)buf, bytes / 16, iters, (float*)sink);
  return last_error();
}

extern "C" int ftrf_peak_l2_red(void* buf, size_t bytes, int iters, int ctas, void* stream) {
  if (!buf || bytes < 16 || iters < 1 || ctas < 1) return -1;
  ftrf::l2_red_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>((float*)buf, bytes / 16, iters);
  return last_error();
}

// bytes of one launch = ctas * iters * 32768; buf needs ctas * 32 KB
extern "C" int ftrf_peak_l1_read(const void* buf, int iters, int ctas, void* sink, void* stream) {
  if (!buf || iters < 1 || ctas < 1 || !sink) return -1;
  ftrf::l1_read_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>((const float4*)buf, iters, (float*)sink);
  return last_error();
}

// flops of one launch = ctas * iters * 2 * 128 * 256 * 8
extern "C" int ftrf_peak_tf32(int mode, int iters, int ctas, void* stream) {
  if (iters < 4 || ctas < 1) return -1;
  if (mode == 0) {
    cudaFuncSetAttribute(ftrf::tf32_peak_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 163840);
    ftrf::tf32_peak_kernel<0><<<ctas, 128, 163840, (cudaStream_t)stream>>>(iters);
  } else {
    cudaFuncSetAttribute(ftrf::tf32_peak_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 163840);
    ftrf::tf32_peak_kernel<1><<<ctas, 128, 163840, (cudaStream_t)stream>>>(iters);
  }
  return last_error();
}
