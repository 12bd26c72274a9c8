// k_wgrad3x.cu — weight gradients of the shading MLP on the tensor cores:
//   dW[o][k] += sum_s G[s][o] * A[s][k]        (G = dL/d pre-activation [M][128], A = layer input [M][K])
// i.e. D[128 x Np] = G^T A with the SAMPLE index as the MMA K dimension.  Both operands are row-major
// over samples in global memory, so the producer threads transpose while filling shared memory: a
// thread gathers 4 consecutive samples of one column (coalesced across the warp) into one 16-byte
// K-major unit.  hi/lo tf32 split as in k_gemm3x.cu (3 MMAs per k-step).  The accumulator stays in
// TMEM over every chunk the CTA owns and is flushed once with atomics; the bias gradient rides along
// as a virtual all-ones column of A.
#include "gemm3x.cuh"
#include "tc.cuh"

namespace ftrf {

constexpr int W_KC = 32;     // samples per chunk (4 k-steps of 8)

__global__ void __launch_bounds__(256, 1)
wgrad3x_kernel(const Wgrad3xArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const int ncol = a.K + (a.dbias ? 1 : 0), Np = g_np(ncol);
  const int gsz = 128 * W_KC, asz = Np * W_KC;                // floats per hi (or lo) tile
  float* base = reinterpret_cast<float*>(smem_raw);           // [2 buffers][Ghi, Glo, Ahi, Alo]
  const int bufsz = 2 * gsz + 2 * asz;
  uint64_t* bars = reinterpret_cast<uint64_t*>(base + 2 * bufsz);
  uint64_t* bar_free = bars;
  uint64_t* bar_acc = bars + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t M = a.count ? min((int64_t)*a.count, a.cap) : a.cap;
  const int64_t nchunks = (M + W_KC - 1) / W_KC;
  if ((int64_t)blockIdx.x >= nchunks) return;

  if (tid == 0) {
    tc::mbar_init(bar_free + 0, 1);
    tc::mbar_init(bar_free + 1, 1);
    tc::mbar_init(bar_acc, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t idesc = tc::idesc_tf32(128, Np);
  const uint32_t lbo = 128, sbo = (W_KC / 4) * 128;

  constexpr int GU = 128 * (W_KC / 4) / 256;     // G units per thread (4)
  constexpr int AU = 6;                          // A units per thread (Np <= 192)
  const int a_units = Np * (W_KC / 4);
  auto load_chunk = [&](int64_t ch, float gv[GU][4], float av[AU][4]) {
    const int64_t s0 = ch * W_KC;
#pragma unroll
    for (int u = 0; u < GU; ++u) {
      int e = tid + u * 256, o = e & 127, ku = e >> 7;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int64_t s = s0 + ku * 4 + i;
        gv[u][i] = s < M ? a.G[s * a.ldg + o] : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < AU; ++u) {
      int e = tid + u * 256;
      if (e < a_units) {
        int n = e % Np, ku = e / Np;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          int64_t s = s0 + ku * 4 + i;
          av[u][i] = (s < M) ? (n < a.K ? a.A[s * a.lda + n] : (n == a.K && a.dbias ? 1.f : 0.f)) : 0.f;
        }
      }
    }
  };
  // Two register sets alternate (loop unrolled by 2, so they are static registers): the loads of the
  // chunk two iterations ahead are issued right after the current one is written to shared memory.
  auto body = [&](int64_t ch, uint32_t it, float (&gv)[GU][4], float (&av)[AU][4]) {
    if (ch >= nchunks) return;
    const int buf = it & 1;
    float* Ghi = base + buf * bufsz;
    float* Glo = Ghi + gsz;
    float* Ahi = Glo + gsz;
    float* Alo = Ahi + asz;
    if (it >= 2) tc::mbar_wait(bar_free + buf, ((it >> 1) - 1) & 1);
#pragma unroll
    for (int u = 0; u < GU; ++u) {               // G^T: rows o, K = samples
      int e = tid + u * 256, o = e & 127, ku = e >> 7;
      float4 h, l;
      tc::split_tf32(gv[u][0], h.x, l.x); tc::split_tf32(gv[u][1], h.y, l.y);
      tc::split_tf32(gv[u][2], h.z, l.z); tc::split_tf32(gv[u][3], h.w, l.w);
      int off = tc::canon_off(o, ku, W_KC / 4);
      *reinterpret_cast<float4*>(Ghi + off) = h;
      *reinterpret_cast<float4*>(Glo + off) = l;
    }
#pragma unroll
    for (int u = 0; u < AU; ++u) {               // A^T: rows k (+ the virtual ones column), K = samples
      int e = tid + u * 256;
      if (e < a_units) {
        int n = e % Np, ku = e / Np;
        float4 h, l;
        tc::split_tf32(av[u][0], h.x, l.x); tc::split_tf32(av[u][1], h.y, l.y);
        tc::split_tf32(av[u][2], h.z, l.z); tc::split_tf32(av[u][3], h.w, l.w);
        int off = tc::canon_off(n, ku, W_KC / 4);
        *reinterpret_cast<float4*>(Ahi + off) = h;
        *reinterpret_cast<float4*>(Alo + off) = l;
      }
    }
    if (ch + 2 * (int64_t)gridDim.x < nchunks) load_chunk(ch + 2 * (int64_t)gridDim.x, gv, av);
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc::fence_after_sync();
      const uint32_t gh = tc::smem_u32(Ghi), gl = tc::smem_u32(Glo), ah = tc::smem_u32(Ahi), al = tc::smem_u32(Alo);
#pragma unroll
      for (int s = 0; s < W_KC / 8; ++s) {
        uint32_t ko = s * 256;
        uint64_t dgh = tc::smem_desc(gh + ko, lbo, sbo), dgl = tc::smem_desc(gl + ko, lbo, sbo);
        uint64_t dah = tc::smem_desc(ah + ko, lbo, sbo), dal = tc::smem_desc(al + ko, lbo, sbo);
        tc::mma_tf32_ss(tmem, dgl, dah, idesc, (it | (uint32_t)s) != 0);
        tc::mma_tf32_ss(tmem, dgh, dal, idesc, true);
        tc::mma_tf32_ss(tmem, dgh, dah, idesc, true);
      }
      tc::mma_commit(bar_free + buf);
    }
  };
  float g0[GU][4], a0[AU][4], g1[GU][4], a1[AU][4];
  load_chunk(blockIdx.x, g0, a0);
  if ((int64_t)blockIdx.x + gridDim.x < nchunks) load_chunk((int64_t)blockIdx.x + gridDim.x, g1, a1);
  uint32_t it = 0;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += 2 * (int64_t)gridDim.x, it += 2) {
    body(ch, it, g0, a0);
    body(ch + gridDim.x, it + 1, g1, a1);
  }
  if (tid == 0) tc::mma_commit(bar_acc);
  tc::mbar_wait(bar_acc, 0);
  tc::fence_after_sync();
  {
    const int q = warp & 3, o = q * 32 + lane;
    for (int cc = warp >> 2; cc * 32 < Np; cc += 2) {
      float acc[32];
      tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(cc * 32), acc);
      tc::tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        int col = cc * 32 + j;
        float v = acc[j];
        if (col < a.K) {
          int dst = a.use_perm ? a.perm[col] : col;
          if (v != 0.f && dst >= 0) atomicAdd(a.dW + (size_t)o * a.ldw + dst, v);
        }
        else if (col == a.K && a.dbias) { if (v != 0.f) atomicAdd(a.dbias + o, v); }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

int launch_wgrad3x(const Wgrad3xArgs& a, cudaStream_t st) {
  int Np = g_np(a.K + (a.dbias ? 1 : 0));
  if (a.K < 1 || Np > 192 || a.cap < 0) { set_error("wgrad3x: bad shape K=%d", a.K); return FTRF_E_ARG; }
  size_t smem = (size_t)2 * (2 * 128 * W_KC + 2 * Np * W_KC) * 4 + 64;
  if (smem > 227 * 1024) { set_error("wgrad3x: K=%d needs %zu B of shared memory", a.K, smem); return FTRF_E_ARG; }
  if (a.cap == 0) return FTRF_OK;
  cudaFuncSetAttribute(wgrad3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t chunks = (a.cap + W_KC - 1) / W_KC;
  int grid = (int)(chunks < FTRF_SMS ? chunks : FTRF_SMS);
  wgrad3x_kernel<<<grid, 256, smem, st>>>(a);
  return check_launch("wgrad3x");
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_test_wgrad3x(const float* G, int ldg, const float* A, int lda, int K, int64_t M, float* dW, int ldw,
                                 float* dbias, void* stream) {
  ftrf::Wgrad3xArgs a{};
  a.G = G; a.ldg = ldg; a.A = A; a.lda = lda; a.K = K; a.count = nullptr; a.cap = M; a.dW = dW; a.ldw = ldw; a.dbias = dbias;
  return ftrf::launch_wgrad3x(a, (cudaStream_t)stream);
}
#endif
