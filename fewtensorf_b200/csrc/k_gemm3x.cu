// k_gemm3x.cu — the dense contraction of the shading MLP on the 5th-gen tensor cores.
//   out[M][N] = epilogue( A[M][K] * B[N][K]^T )        fp32 in, fp32 out
// computed as 3xTF32: A = Ah + Al, B = Bh + Bl (tf32 pairs), D = Ah*Bh + Ah*Bl + Al*Bh accumulated in
// fp32 in TMEM by tcgen05.mma.kind::tf32 — ~2^-21 relative per product, i.e. fp32-grade, which is what
// the 1e-4 parity budget on rgb / gradients needs (plain TF32 is ~1e-3).
//
// Persistent CTAs (one per SM).  B (the layer's weight matrix, <= 160 x 152) is split once per CTA
// into hi/lo and stays resident in shared memory in the UMMA canonical K-major no-swizzle layout;
// A streams through in 128-row x 32-column chunks: all threads LDG -> split -> STS (double buffered),
// one elected thread issues the MMAs asynchronously, tcgen05.commit releases the chunk buffer through
// an mbarrier, and after the last chunk the accumulator tile is read back with tcgen05.ld for the
// fused epilogue (bias, ReLU, ReLU-mask of a saved activation).
#include "common.cuh"
#include "gemm3x.cuh"
#include "tc.cuh"

namespace ftrf {


constexpr int G_KC = 32;                 // K columns per streamed A chunk
constexpr int G_TM = 128;                // rows per tile (UMMA M)
constexpr int G_ABUF = G_TM * G_KC;      // floats per hi (or lo) chunk buffer



__global__ void __launch_bounds__(256, 1)
gemm3x_kernel(const Gemm3xArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const int Np = g_np(a.N), Kp = g_kp(a.K), KUB = Kp / 4;
  float* Bhi = reinterpret_cast<float*>(smem_raw);
  float* Blo = Bhi + (size_t)Np * Kp;
  float* Abuf = Blo + (size_t)Np * Kp;                       // [2 buffers][hi, lo][G_ABUF]
  uint64_t* bars = reinterpret_cast<uint64_t*>(Abuf + 4 * G_ABUF);
  uint64_t* bar_free = bars;                                  // [2]
  uint64_t* bar_acc = bars + 2;                               // [2] accumulator stages
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  float* s_bias = reinterpret_cast<float*>(bars + 6);        // [Np]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t M = a.count ? min((int64_t)*a.count, a.cap) : a.cap;
  if ((int64_t)blockIdx.x * G_TM >= M) return;

  if (tid == 0) {
    tc::mbar_init(bar_free + 0, 1);
    tc::mbar_init(bar_free + 1, 1);
    tc::mbar_init(bar_acc + 0, 1);
    tc::mbar_init(bar_acc + 1, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  for (int e = tid; e < Np; e += blockDim.x) s_bias[e] = (a.bias && e < a.N) ? __ldg(a.bias + e) : 0.f;
  // resident B: split into hi / lo, canonical layout.  Lanes run along the contiguous global axis.
  const bool k_contig = a.sk == 1;
  const int dbg = a.variant;     // bring-up switches: 2 no MMA, 4 no epilogue, 8 no A loads, 16 no B fill
#pragma unroll 4
  for (int e = tid; e < ((dbg & 16) ? 0 : Np * KUB); e += blockDim.x) {
    int n = k_contig ? e / KUB : e % Np, ku = k_contig ? e % KUB : e / Np;
    float hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int k = ku * 4 + j;
      float v = (n < a.N && k < a.K) ? __ldg(a.B + (int64_t)n * a.sn + (int64_t)k * a.sk) : 0.f;
      tc::split_tf32(v, hi[j], lo[j]);
    }
    int off = tc::canon_off(n, ku, KUB);
    *reinterpret_cast<float4*>(Bhi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(Blo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  const uint32_t a_lbo = 128, a_sbo = (G_KC / 4) * 128, b_lbo = 128, b_sbo = (uint32_t)KUB * 128;
  const uint32_t idesc = tc::idesc_tf32(G_TM, Np);
  const int nchunks = (Kp + G_KC - 1) / G_KC;
  const int64_t ntiles = (M + G_TM - 1) / G_TM;
  const int my_tiles = (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const int total = my_tiles * nchunks;

  // this thread's 4 (row, k-unit) slots of a chunk: 8 rows x 64 B per warp item
  int slot_r[4], slot_ku[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int id = warp * 4 + j;
    slot_r[j] = (id >> 1) * 8 + (lane & 7);
    slot_ku[j] = (id & 1) * 4 + (lane >> 3);
  }
  auto load_chunk = [&](int step, float4 v[4]) {
    int t = step / nchunks, c = step - t * nchunks;
    int64_t m0 = ((int64_t)blockIdx.x + (int64_t)t * gridDim.x) * G_TM;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int64_t row = m0 + slot_r[j];
      int col = c * G_KC + slot_ku[j] * 4;
      v[j] = (row < M && col < a.ka && !(dbg & 8)) ? ld4(a.A + row * a.lda + col) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  // TMEM -> registers -> global for one finished tile (accumulator stage `stg`)
  auto epilogue = [&](int t) {
    const int stg = t & 1;
    tc::mbar_wait(bar_acc + stg, (t >> 1) & 1);
    tc::fence_after_sync();
    const int64_t m0 = ((int64_t)blockIdx.x + (int64_t)t * gridDim.x) * G_TM;
    const int q = warp & 3;
    const int64_t row = m0 + q * 32 + lane;
    for (int cc = warp >> 2; cc * 32 < ((dbg & 4) ? 0 : Np); cc += 2) {
      float acc[32];
      tc::tmem_ld32(tmem + (uint32_t)(stg * 256) + ((uint32_t)(q * 32) << 16) + (uint32_t)(cc * 32), acc);
      tc::tmem_wait_ld();
      if (row < M) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          int col = cc * 32 + j;
          if (col >= a.N) break;
          float o[4], mk[4] = {1.f, 1.f, 1.f, 1.f};
          if (a.mask) {
            if (col + 3 < a.N && (a.ldm & 3) == 0) {
              float4 m4 = ld4(a.mask + row * a.ldm + col);
              mk[0] = m4.x; mk[1] = m4.y; mk[2] = m4.z; mk[3] = m4.w;
            } else
              for (int e = 0; e < 4 && col + e < a.N; ++e) mk[e] = __ldg(a.mask + row * a.ldm + col + e);
          }
          float* dst = a.out + row * a.ldo + col;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float x = acc[j + e];
            if (a.accumulate && col + e < a.N) x += dst[e];
            x += s_bias[col + e];
            if (a.relu) x = fmaxf(x, 0.f);
            if (!(mk[e] > 0.f)) x = 0.f;
            o[e] = x;
          }
          if (col + 3 < a.N && (a.ldo & 3) == 0) *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
          else
            for (int e = 0; e < 4 && col + e < a.N; ++e) dst[e] = o[e];
        }
      }
    }
    tc::fence_before_sync();     // ordered before the next __syncthreads / MMA that reuses this stage
  };

  // One pipeline step = one A chunk.  Three register sets rotate (the loop is unrolled by 3 so the
  // sets are static registers): the LDGs of step s+3 are issued as soon as step s has been written
  // to shared memory, i.e. global latency is covered by ~3 steps of MMA time.
  auto body = [&](int step, float4 (&v)[4]) {
    if (step >= total) return;
    const int t = step / nchunks, c = step - t * nchunks;
    const int buf = step & 1;
    float* Ahi = Abuf + buf * 2 * G_ABUF;
    float* Alo = Ahi + G_ABUF;
    if (step >= 2) tc::mbar_wait(bar_free + buf, ((step >> 1) - 1) & 1);   // MMAs that read this buffer are done
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float4 h, l;
      tc::split_tf32(v[j].x, h.x, l.x); tc::split_tf32(v[j].y, h.y, l.y);
      tc::split_tf32(v[j].z, h.z, l.z); tc::split_tf32(v[j].w, h.w, l.w);
      int off = tc::canon_off(slot_r[j], slot_ku[j], G_KC / 4);
      *reinterpret_cast<float4*>(Ahi + off) = h;
      *reinterpret_cast<float4*>(Alo + off) = l;
    }
    if (step + 3 < total) load_chunk(step + 3, v);
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {              // one thread issues this chunk's MMAs (async); commit frees the buffer
      tc::fence_after_sync();
      const uint32_t acc_t = tmem + (uint32_t)((t & 1) * 256);
      const int ksteps = min(G_KC, Kp - c * G_KC) / 8;
      const uint32_t ah = tc::smem_u32(Ahi), al = tc::smem_u32(Alo);
      const uint32_t bh = tc::smem_u32(Bhi) + (uint32_t)(c * (G_KC / 4)) * 128;
      const uint32_t bl = tc::smem_u32(Blo) + (uint32_t)(c * (G_KC / 4)) * 128;
      for (int s = 0; s < ((dbg & 2) ? 0 : ksteps); ++s) {
        uint32_t ko = s * 256;      // 8 tf32 = two 16-byte units along K
        uint64_t dah = tc::smem_desc(ah + ko, a_lbo, a_sbo), dal = tc::smem_desc(al + ko, a_lbo, a_sbo);
        uint64_t dbh = tc::smem_desc(bh + ko, b_lbo, b_sbo), dbl = tc::smem_desc(bl + ko, b_lbo, b_sbo);
        tc::mma_tf32_ss(acc_t, dal, dbh, idesc, (c | s) != 0);
        tc::mma_tf32_ss(acc_t, dah, dbl, idesc, true);
        tc::mma_tf32_ss(acc_t, dah, dbh, idesc, true);
      }
      tc::mma_commit(bar_free + buf);
      if (c == nchunks - 1) tc::mma_commit(bar_acc + (t & 1));
    }
    // the previous tile's epilogue runs while this tile's first MMAs execute
    if (c == 0 && t > 0) epilogue(t - 1);
  };
  float4 r0[4], r1[4], r2[4];
  load_chunk(0, r0);
  if (1 < total) load_chunk(1, r1);
  if (2 < total) load_chunk(2, r2);
  for (int step = 0; step < total; step += 3) {
    body(step, r0);
    body(step + 1, r1);
    body(step + 2, r2);
  }
  epilogue(my_tiles - 1);
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

size_t gemm3x_smem(int N, int K) {
  return (size_t)2 * g_np(N) * g_kp(K) * 4 + (size_t)4 * G_ABUF * 4 + 64 + (size_t)g_np(N) * 4;
}

int launch_gemm3x(const Gemm3xArgs& a, cudaStream_t st) {
  if (a.N < 1 || a.N > 256 || a.K < 1 || (a.ka & 3) || (a.lda & 3) || a.cap < 0) { set_error("gemm3x: bad shape N=%d K=%d ka=%d lda=%d", a.N, a.K, a.ka, a.lda); return FTRF_E_ARG; }
  size_t smem = gemm3x_smem(a.N, a.K);
  if (smem > 227 * 1024) { set_error("gemm3x: N=%d K=%d needs %zu B of shared memory", a.N, a.K, smem); return FTRF_E_ARG; }
  if (a.cap == 0) return FTRF_OK;
  cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t tiles = (a.cap + G_TM - 1) / G_TM;
  int grid = (int)(tiles < FTRF_SMS ? tiles : FTRF_SMS);
  gemm3x_kernel<<<grid, 256, smem, st>>>(a);
  return check_launch("gemm3x");
}

}  // namespace ftrf

// test / bring-up entry (declared in include/fewtensorf.h)
#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_test_gemm3x(const float* A, int lda, int ka, int64_t M, const float* B, int64_t sn, int64_t sk, int N,
                                int K, const float* bias, const float* mask, int ldm, int relu, float* out, int ldo,
                                int variant, void* stream) {
  ftrf::Gemm3xArgs a{};
  a.A = A; a.lda = lda; a.ka = ka; a.count = nullptr; a.cap = M;
  a.B = B; a.sn = sn; a.sk = sk; a.N = N; a.K = K;
  a.bias = bias; a.mask = mask; a.ldm = ldm; a.relu = relu; a.out = out; a.ldo = ldo; a.variant = variant;
  return ftrf::launch_gemm3x(a, (cudaStream_t)stream);
}
#endif
