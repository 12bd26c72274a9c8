// k_mma_probe.cu — micro-probe (debug aid, not on the product path): issue rate of back-to-back
// tcgen05.mma.kind::tf32 dispatches for a few operand placements, one CTA per SM, one issuing thread.
#include "common.cuh"
#include "tc.cuh"

namespace ftrf {

__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}

// MODE 0: SS N=128 K-major no-swizzle   1: SS N=256   2: TS N=128   3: TS N=256   4: SS N=128, B MN-major
//      5: SS N=64   6: SS N=128 both SWIZZLE_128B K-major   7: SS N=256 SW128   8: TS N=128 B SW128
// Descriptors are built before the timed loop; the loop body is 4 dispatches (the 4 k-steps of a chunk).
template <int MODE>
__global__ void __launch_bounds__(128, 1) mma_probe_kernel(int iters, long long* out) {
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
    constexpr int N = (MODE == 1 || MODE == 3 || MODE == 7) ? 256 : (MODE == 5 ? 64 : 128);
    uint32_t idesc = tc::idesc_tf32(128, N);
    if (MODE == 4) idesc |= (1u << 16);
    const uint32_t a0 = tc::smem_u32(sm), b0 = a0 + 32768;
    uint64_t da[4], db[4];
    uint32_t ta[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      da[s] = tc::smem_desc(a0 + s * 256, 128, 1024);
      db[s] = tc::smem_desc(b0 + s * 256, 128, 1024);
      if (MODE == 4) db[s] = tc::smem_desc(b0 + s * 4096, 4096, 128);
      if (MODE == 6 || MODE == 7) {
        da[s] = tc::smem_desc(a0 + s * 32, 16, 1024) | (2ull << 61);
        db[s] = tc::smem_desc(b0 + s * 32, 16, 1024) | (2ull << 61);
      }
      if (MODE == 8) db[s] = tc::smem_desc(b0 + s * 32, 16, 1024) | (2ull << 61);
      ta[s] = tmem + 256 + s * 8;
    }
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 4) {
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (MODE == 2 || MODE == 3 || MODE == 8) mma_tf32_ts(tmem, ta[s], db[s], idesc, true);
        else tc::mma_tf32_ss(tmem, da[s], db[s], idesc, true);
      }
    }
    long long t1 = clock64();
    tc::mma_commit(&bar);
    tc::mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

// Layout probe: A region filled with (16-byte unit index) or (element index within the unit); B is a
// K-major identity, so D[m][k] (k < 8) returns whichever shared-memory float the tensor core used as A(m,k).
__global__ void __launch_bounds__(128, 1) layout_probe_kernel(int a_mn, uint32_t lbo, uint32_t sbo, int within, float* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  float* A = reinterpret_cast<float*>(smem_raw);            // 64 KB region
  float* B = A + 16384;                                     // K-major identity [128 x 8]: units (n/8)*2*8... LBO 128, SBO 256
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) A[i] = within ? (float)(i & 3) : (float)(i >> 2);
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) B[i] = 0.f;
  __syncthreads();
  if (threadIdx.x < 8) {      // B[n][k] = 1 for n == k < 8 : K-major no-swizzle, rows 0-7 in core matrix 0 (k 0-3) and LBO later (k 4-7)
    int n = threadIdx.x, k = n;
    B[(k >> 2) * 32 + n * 4 + (k & 3)] = 1.f;               // unit (k/4) at LBO = 128 B = 32 floats; row n at 16 B
  }
  if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (threadIdx.x < 32) tc::tmem_alloc(&slot, 512);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    uint32_t idesc = tc::idesc_tf32(128, 128) | (a_mn ? (1u << 15) : 0u);
    uint64_t da = tc::smem_desc(tc::smem_u32(A), lbo, sbo);
    uint64_t db = tc::smem_desc(tc::smem_u32(B), 128, 256);
    tc::mma_tf32_ss(tmem, da, db, idesc, false);
    tc::mma_commit(&bar);
    tc::mbar_wait(&bar, 0);
  }
  __syncthreads();
  tc::fence_after_sync();
  {
    float acc[32];
    const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
    tc::tmem_ld32(tmem + (((uint32_t)(q * 32)) << 16), acc);
    tc::tmem_wait_ld();
    for (int k = 0; k < 8; ++k) out[(q * 32 + lane) * 8 + k] = acc[k];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

}  // namespace ftrf

extern "C" int ftrf_debug_layout_probe(int a_mn, unsigned lbo, unsigned sbo, int within, float* out_device, void* stream) {
  cudaFuncSetAttribute(ftrf::layout_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304);
  ftrf::layout_probe_kernel<<<1, 128, 98304, (cudaStream_t)stream>>>(a_mn, lbo, sbo, within, out_device);
  return ftrf::check_launch("layout_probe");
}

extern "C" int ftrf_debug_mma_probe(int mode, int iters, int ctas, long long* out_device, void* stream) {
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 163840);
    kern<<<ctas, 128, 163840, (cudaStream_t)stream>>>(iters, out_device);
  };
  switch (mode) {
    case 0: go(ftrf::mma_probe_kernel<0>); break;
    case 1: go(ftrf::mma_probe_kernel<1>); break;
    case 2: go(ftrf::mma_probe_kernel<2>); break;
    case 3: go(ftrf::mma_probe_kernel<3>); break;
    case 4: go(ftrf::mma_probe_kernel<4>); break;
    case 5: go(ftrf::mma_probe_kernel<5>); break;
    case 6: go(ftrf::mma_probe_kernel<6>); break;
    case 7: go(ftrf::mma_probe_kernel<7>); break;
    default: go(ftrf::mma_probe_kernel<8>); break;
  }
  return ftrf::check_launch("mma_probe");
}
