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

// Issue-side cost of what a fused kernel's MMA thread does around its dispatches: groups of 12 TS-form MMAs (one operand
// chunk of the 3xTF32 scheme) with, per group, `commits` tcgen05.commit, `waits` waits on an mbarrier whose phase has
// already completed, and optionally a tcgen05.fence::after_thread_sync.  out[0] = clocks for `groups` groups (issue),
// out[1] = until the last MMA has completed.
__global__ void __launch_bounds__(640, 1) issue_probe_kernel(int groups, int commits, int waits, int fence, int per_group, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t bar, done_bar, sink_bar[4];
  __shared__ uint32_t slot[2];
  float* sm = reinterpret_cast<float*>(smem_raw);
  for (int i = threadIdx.x; i < 40960; i += blockDim.x) sm[i] = 0.f;
  if (threadIdx.x == 0) {
    slot[1] = 7u;
    tc::mbar_init(&bar, 1);
    tc::mbar_init(&done_bar, 1);
    for (int i = 0; i < 4; ++i) tc::mbar_init(sink_bar + i, 1);
    tc::fence_barrier_init();
  }
  if (threadIdx.x < 32) tc::tmem_alloc(slot, 512);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = slot[0];
  if (threadIdx.x >= 128) {                      // noise warps beside the issuer: what slows its dispatches?
    const int noise = fence >> 4, w = threadIdx.x >> 5;
    const uint32_t t_lane = ((uint32_t)((w & 3) * 32)) << 16;
    float x = (float)threadIdx.x, y = 1.0001f;
    volatile uint32_t* stop = slot + 1;
    while (*stop == 7u) {
      if (noise == 1) {
#pragma unroll
        for (int i = 0; i < 64; ++i) { x = fmaf(x, y, 0.5f); y = fmaf(y, 1.0001f, x * 1e-9f); }
      } else if (noise == 2) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          tc::tmem_st4(tmem + t_lane + 448 + (uint32_t)(i * 4), x, y, x, y);
          tc::tmem_st4(tmem + t_lane + 480 + (uint32_t)(i * 4), x, y, x, y);
        }
        tc::tmem_wait_st();
        float v[8];
        tc::tmem_ld8(tmem + t_lane + 448, v);
        tc::tmem_wait_ld();
        x += v[0];
      } else if (noise == 3) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float q;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(q) : "r"(tc::smem_u32(sm + 36864 + ((threadIdx.x * 4 + i * 64) & 1023))) : "memory");
          x += q;
        }
      } else {
        __nanosleep(200);
      }
    }
    if (x == 1234.5f) out[2] = (long long)x;
  }
  if (threadIdx.x == 32 && commits >= 8) {       // a second issuing thread (own warp) on its own accumulator / barriers
    commits -= 8;
    while (!tc::mbar_try_wait(&bar, 0)) {}
    const uint32_t idesc = tc::idesc_tf32(128, 128);
    const uint32_t b_lo = tc::desc_lo(tc::smem_u32(sm) + 65536, 128), d_hi = tc::desc_hi(1024);
    for (int g = 0; g < groups; ++g) {
      for (int w = 0; w < waits; ++w) tc::mbar_wait(&bar, 0);
      if (fence) tc::fence_after_sync();
      const uint32_t a = tmem + 384 + (uint32_t)(g & 1) * 64, d = tmem + 128;
      for (int s = 0; s < per_group / 3; ++s) {
        tc::mma_tf32_ts_lo(d, a + 32 + (s & 3) * 8, b_lo + (s & 3) * 16, d_hi, idesc, 1u);
        tc::mma_tf32_ts_lo(d, a + (s & 3) * 8, b_lo + 1024 + (s & 3) * 16, d_hi, idesc, 1u);
        tc::mma_tf32_ts_lo(d, a + (s & 3) * 8, b_lo + (s & 3) * 16, d_hi, idesc, 1u);
      }
      for (int k = 0; k < commits; ++k) tc::mma_commit(sink_bar + 2 + k);
    }
    tc::mma_commit(&done_bar);
  }
  if (threadIdx.x == 0) {
    const bool two = commits >= 8;
    if (two) commits -= 8;
    tc::mbar_arrive(&bar);                       // phase 0 of `bar` is complete: waits on parity 0 return at once
    const uint32_t idesc = tc::idesc_tf32(128, 128);
    const uint32_t b_lo = tc::desc_lo(tc::smem_u32(sm) + 32768, 128), d_hi = tc::desc_hi(1024);
    long long t0 = clock64();
    for (int g = 0; g < groups; ++g) {
      for (int w = 0; w < waits; ++w) {
        if (((fence >> 1) & 3) == 0) tc::mbar_wait(&bar, 0);
        else if (((fence >> 1) & 3) == 1) {
          uint32_t ok;
          do {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(tc::smem_u32(&bar)), "r"(0) : "memory");
          } while (!ok);
        } else {
          uint32_t v;
          do {
            asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(tc::smem_u32(slot) + 4) : "memory");
          } while (v == 0xffffffffu);
        }
      }
      if (fence & 1) tc::fence_after_sync();
      const uint32_t a = tmem + 256 + (uint32_t)(g & 1) * 64, d = tmem;
      for (int s = 0; s < per_group / 3; ++s) {
        tc::mma_tf32_ts_lo(d, a + 32 + (s & 3) * 8, b_lo + (s & 3) * 16, d_hi, idesc, 1u);
        tc::mma_tf32_ts_lo(d, a + (s & 3) * 8, b_lo + 1024 + (s & 3) * 16, d_hi, idesc, 1u);
        tc::mma_tf32_ts_lo(d, a + (s & 3) * 8, b_lo + (s & 3) * 16, d_hi, idesc, 1u);
      }
      for (int k = 0; k < commits; ++k) tc::mma_commit(sink_bar + k);
    }
    long long t1 = clock64();
    tc::mma_commit(&done_bar);
    tc::mbar_wait(&done_bar, 0);
    if (two) tc::mbar_wait(&done_bar, 1);
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    *reinterpret_cast<volatile uint32_t*>(slot + 1) = 0u;      // stops the noise warps
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

// TS-form check: C[128 x 128] = A[128 x 32] B[128 x 32]^T with A written to TMEM by tcgen05.st (thread = row, one
// 32-bit column per k) and B in shared memory (K-major no-swizzle).  Inputs are expected to be tf32-representable.
__global__ void __launch_bounds__(128, 1) ts_gemm_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ Cout) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  float* sB = reinterpret_cast<float*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int e = tid; e < 128 * 8; e += 128) {
    const int n = e >> 3, ku = e & 7;
    *reinterpret_cast<float4*>(sB + tc::canon_off(n, ku, 8)) = *reinterpret_cast<const float4*>(B + n * 32 + ku * 4);
  }
  if (tid == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc(&slot, 256);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = slot;
  const uint32_t t_lane = ((uint32_t)(warp * 32)) << 16;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const float4 v = *reinterpret_cast<const float4*>(A + tid * 32 + u * 4);
    tc::tmem_st4(tmem + t_lane + 128 + u * 4, v.x, v.y, v.z, v.w);
  }
  tc::tmem_wait_st();
  tc::fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    tc::fence_after_sync();
    const uint32_t idesc = tc::idesc_tf32(128, 128);
    const uint32_t b_lo = tc::desc_lo(tc::smem_u32(sB), 128), d_hi = tc::desc_hi(1024);
#pragma unroll
    for (int s = 0; s < 4; ++s) tc::mma_tf32_ts_lo(tmem, tmem + 128 + s * 8, b_lo + s * 16, d_hi, idesc, s ? 1u : 0u);
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  for (int j = 0; j < 4; ++j) {
    float acc[32];
    tc::tmem_ld32(tmem + t_lane + j * 32, acc);
    tc::tmem_wait_ld();
    for (int i = 0; i < 32; ++i) Cout[tid * 128 + j * 32 + i] = acc[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace ftrf

extern "C" int ftrf_debug_issue_probe(int groups, int commits, int waits, int fence, int per_group, int ctas, long long* out_device,
                                      void* stream) {
  if (commits < 0 || (commits & 7) > 2 || per_group % 3) return -1;
  cudaFuncSetAttribute(ftrf::issue_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 163840);
  ftrf::issue_probe_kernel<<<ctas, (fence >> 4) ? 128 + 512 : 128, 163840, (cudaStream_t)stream>>>(groups, commits, waits, fence, per_group, out_device);
  return ftrf::check_launch("issue_probe");
}

extern "C" int ftrf_test_ts_gemm(const float* A, const float* B, float* C, void* stream) {
  cudaFuncSetAttribute(ftrf::ts_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  ftrf::ts_gemm_kernel<<<1, 128, 32768, (cudaStream_t)stream>>>(A, B, C);
  return ftrf::check_launch("ts_gemm");
}

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
