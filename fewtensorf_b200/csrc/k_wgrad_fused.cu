// k_wgrad_fused.cu — both weight gradients of the shading MLP in ONE persistent tcgen05 kernel:
//   dW2[o][i] += sum_s g_h2[s][o] * h1[s][i]                      (128 x 128)
//   dW1[o][k] += sum_s g_h1[s][o] * X[s][k],  db1[o] += sum_s g_h1[s][o]   (128 x kp, bias as a ones column)
// (autograd of models/mlp.py:34-39's Linear layers.)  The SAMPLE index is the MMA K dimension, so both
// operands of both products are transposed views of the row-major activations.  (UMMA "MN-major"
// operands would avoid the transposition, but kind::tf32 with the no-swizzle layout returns zeros for a
// transposed operand on this part — scripts/layout_probe.py — so the tiles are K-major and the producer
// threads transpose: a thread gathers 4 consecutive samples of ONE column with four warp-coalesced
// 4-byte loads into one 16-byte K-major unit.)
//   shared tile [col/8][s/4][col%8][s%4]   <=>   UMMA K-major no-swizzle, LBO = 128 B, SBO = 512 B.
// 8 producer warps stream 16-sample chunks (LDG -> tf32 hi/lo split -> STS.128, register double
// buffering), warp 8 issues the 12 MMAs per chunk (3xTF32) into two TMEM accumulators that stay
// resident for the whole kernel; the epilogue adds them into a zeroed internal-order scratch with
// red.global.add.v4.f32, and a small kernel un-permutes dW1 into the parameter gradient.
#include "fused.cuh"
#include "gemm3x.cuh"
#include "tc.cuh"

namespace ftrf {

FTRF_SWITCH(int, g_wgrad_variant, 0);

constexpr int WG_KS = 16;                 // samples per chunk (2 k-steps of 8)
constexpr int WG_STAGES = 3;
constexpr int WG_NGX_MAX = 48;            // 16-byte column groups of the X operand: data + ones column, padded to a multiple of 4
constexpr int WG_G_FLOATS = WG_KS * 128;  // one hi (or lo) tile of a 128-wide operand
constexpr int WG_PROD_WARPS = 16;
constexpr int WG_PROD = WG_PROD_WARPS * 32;
constexpr int WG_THREADS = WG_PROD + 32;


// Predicated read-only 4-byte load that stays where it is written: `volatile` keeps it ordered with the
// mbarrier waits / arrivals around it, so the prefetch of a later chunk really is issued before the
// current chunk is processed (plain loads get sunk towards their use under register pressure).
__device__ __forceinline__ float ldg_early(const float* p, bool ok) {
  float v;
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "setp.ne.b32 q, %2, 0;\n\t"
      "mov.f32 %0, 0f00000000;\n\t"
      "@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"((int)ok));
  return v;
}

__device__ __forceinline__ void red4g(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_fused_kernel(const WgradFusedArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  float* ring = reinterpret_cast<float*>(smem_raw);                    // [STAGES][G2h,G2l,H1h,H1l,G1h,G1l,Xh,Xl]
  const int WG_NGX = a.ngt;
  const int WG_X_FLOATS = WG_KS * WG_NGX * 4;
  const int WG_STAGE_FLOATS = 2 * (3 * WG_G_FLOATS + WG_X_FLOATS);      // G2, H1, G1, X, each hi + lo
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + (size_t)WG_STAGES * WG_STAGE_FLOATS);
  uint64_t* full = bars;                      // [STAGES]  8 producer warps
  uint64_t* empty = bars + WG_STAGES;         // [STAGES]  tcgen05.commit
  uint64_t* done = bars + 2 * WG_STAGES;      // accumulators final
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float fold_dw3[3] = {0.f, 0.f, 0.f}, fold_db2 = 0.f, fold_db3[3] = {0.f, 0.f, 0.f};     // output-layer gradients (fold3)
  const int64_t M = a.count ? min((int64_t)*a.count, a.cap) : a.cap;
  const int64_t nchunks = (M + WG_KS - 1) / WG_KS;
  if ((int64_t)blockIdx.x >= nchunks) return;
  const int my_chunks = (int)((nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x);

  if (tid == 0) {
    for (int i = 0; i < WG_STAGES; ++i) { tc::mbar_init(full + i, WG_PROD_WARPS); tc::mbar_init(empty + i, 1); }
    tc::mbar_init(done, 1);
    tc::fence_barrier_init();
  }
  if (warp == WG_PROD_WARPS) tc::tmem_alloc(tmem_slot, 512);
  // constant part of every X stage: the ones column (bias gradient) and the zero padding columns
  for (int e = tid; e < WG_STAGES * 2 * (WG_NGX * 4 - a.kp) * 4; e += WG_THREADS) {
    int ku = e & 3, r = e >> 2;
    int col = a.kp + r % (WG_NGX * 4 - a.kp); r /= (WG_NGX * 4 - a.kp);
    int hl = r & 1, stg = r >> 1;
    float* X = ring + (size_t)stg * WG_STAGE_FLOATS + 6 * WG_G_FLOATS + hl * WG_X_FLOATS;
    float one = (col == a.kp && hl == 0) ? 1.f : 0.f;
    *reinterpret_cast<float4*>(X + tc::canon_off(col, ku, 4)) = make_float4(one, one, one, one);
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;            // dW2: columns 0-127, dW1: columns 128-...

  if (warp < WG_PROD_WARPS) {
    // ============================================================ producers
    // unit (col, ku) = samples 4ku..4ku+3 of one column; lanes run over col (coalesced 4-byte loads).
    // Two register sets rotate, so the loads of chunk j+2 are in flight while chunks j, j+1 are written (17 warps
    // leave 96 registers per thread: a third set spills once the output-layer gradients live here too).
    float v[2][20];
    const int xu = a.kp * 4;                              // X units per chunk
    const int gcol = tid & 127, gku = tid >> 7;           // this thread's unit of every 128-wide tile
    float w3c[3] = {0.f, 0.f, 0.f};
    float (&dw3)[3] = fold_dw3, &db2 = fold_db2, (&db3)[3] = fold_db3;
    if (a.fold3) {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) w3c[ch] = __ldg(a.w3 + ch * 128 + gcol);
    }
    auto load = [&](int j, float (&r)[20]) {
      const int64_t s0 = ((int64_t)blockIdx.x + (int64_t)j * gridDim.x) * WG_KS;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t s = s0 + gku * 4 + i;
        const bool ok = s < M;
        r[i] = ldg_early(a.g2 + s * 128 + gcol, ok);
        r[4 + i] = ldg_early(a.h1 + s * 128 + gcol, ok);
        r[8 + i] = ldg_early(a.g1 + s * 128 + gcol, ok);
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {                       // X: kp columns x 4 units
        const int id = tid + WG_PROD * q;
        const int ku = id / a.kp, col = id - ku * a.kp;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int64_t s = s0 + ku * 4 + i;
          r[12 + q * 4 + i] = ldg_early(a.x + s * a.ldx + col, id < xu && s < M);
        }
      }
    };
    auto put = [&](float* hi, float* lo, int off, const float* x) {
      float4 h, l;
      tc::split_tf32(x[0], h.x, l.x); tc::split_tf32(x[1], h.y, l.y);
      tc::split_tf32(x[2], h.z, l.z); tc::split_tf32(x[3], h.w, l.w);
      *reinterpret_cast<float4*>(hi + off) = h;
      *reinterpret_cast<float4*>(lo + off) = l;
    };
    auto store = [&](int j, float (&r)[20]) {
      const int stg = j % WG_STAGES;
      if (a.fold3) {
        // r[0..3] holds h2 of this thread's column for four samples: rebuild g_h2 and collect the output layer's gradients
        const int64_t s0 = ((int64_t)blockIdx.x + (int64_t)j * gridDim.x) * WG_KS + gku * 4;
        if (j + 1 < my_chunks) {                        // the next chunk's go values: into L1 while this chunk is written
          const float4* nx = a.go4 + ((int64_t)blockIdx.x + (int64_t)(j + 1) * gridDim.x) * WG_KS + gku * 4;
          asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 g = s0 + i < M ? __ldg(a.go4 + s0 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
          const float h = r[i];
          dw3[0] = fmaf(g.x, h, dw3[0]); dw3[1] = fmaf(g.y, h, dw3[1]); dw3[2] = fmaf(g.z, h, dw3[2]);
          const float gh = h > 0.f ? g.x * w3c[0] + g.y * w3c[1] + g.z * w3c[2] : 0.f;
          db2 += gh;
          if (gcol == 0) { db3[0] += g.x; db3[1] += g.y; db3[2] += g.z; }
          r[i] = gh;
        }
      }
      tc::mbar_wait(empty + stg, ((j / WG_STAGES) & 1) ^ 1);
      float* S = ring + (size_t)stg * WG_STAGE_FLOATS;
      const int off = tc::canon_off(gcol, gku, 4);
      put(S, S + WG_G_FLOATS, off, r);
      put(S + 2 * WG_G_FLOATS, S + 3 * WG_G_FLOATS, off, r + 4);
      put(S + 4 * WG_G_FLOATS, S + 5 * WG_G_FLOATS, off, r + 8);
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int id = tid + WG_PROD * q;
        const int ku = id / a.kp, col = id - ku * a.kp;
        if (id < xu) put(S + 6 * WG_G_FLOATS, S + 6 * WG_G_FLOATS + WG_X_FLOATS, tc::canon_off(col, ku, 4), r + 12 + q * 4);
      }
      tc::fence_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(full + stg);
    };
    load(0, v[0]);
    if (1 < my_chunks) load(1, v[1]);
    for (int j = 0; j < my_chunks; j += 2) {
      store(j, v[0]);
      if (j + 2 < my_chunks) load(j + 2, v[0]);
      if (j + 1 < my_chunks) {
        store(j + 1, v[1]);
        if (j + 3 < my_chunks) load(j + 3, v[1]);
      }
    }
  } else if (lane == 0) {
    // ============================================================ MMA issuer
    const uint32_t idesc2 = tc::idesc_tf32(128, 128), idesc1 = tc::idesc_tf32(128, WG_NGX * 4);
    const uint32_t hi_w = tc::desc_hi(512);                                  // SBO: 8-row group stride (4 units of 128 B)
    const uint32_t lo_g = tc::desc_lo(tc::smem_u32(ring), 128);              // LBO: next 16-byte unit along K
    const uint32_t lo_x = tc::desc_lo(tc::smem_u32(ring + 6 * WG_G_FLOATS), 128);
    const uint32_t GU = WG_G_FLOATS * 4 / 16, XU = WG_X_FLOATS * 4 / 16, SU = WG_STAGE_FLOATS * 4 / 16;
    for (int j = 0; j < my_chunks; ++j) {
      const int stg = j % WG_STAGES;
      tc::mbar_wait(full + stg, (j / WG_STAGES) & 1);
      tc::fence_after_sync();
      const uint32_t g = lo_g + (uint32_t)stg * SU, x = lo_x + (uint32_t)stg * SU;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        const uint32_t kk = ks * 16;                       // 8 samples = two 16-byte units = 256 B
        const uint32_t acc = (j | ks) ? 1u : 0u;
        // dW2 = g_h2^T h1 :  A = G2 (hi at +0, lo at +GU), B = H1 (hi +2GU, lo +3GU)
        tc::mma_tf32_ss_lo(tmem, g + GU + kk, g + 2 * GU + kk, hi_w, idesc2, acc);
        tc::mma_tf32_ss_lo(tmem, g + kk, g + 3 * GU + kk, hi_w, idesc2, 1u);
        tc::mma_tf32_ss_lo(tmem, g + kk, g + 2 * GU + kk, hi_w, idesc2, 1u);
        // dW1 = g_h1^T X   :  A = G1 (hi +4GU, lo +5GU), B = X (hi +0, lo +XU)
        tc::mma_tf32_ss_lo(tmem + 128, g + 5 * GU + kk, x + kk, hi_w, idesc1, acc);
        tc::mma_tf32_ss_lo(tmem + 128, g + 4 * GU + kk, x + XU + kk, hi_w, idesc1, 1u);
        tc::mma_tf32_ss_lo(tmem + 128, g + 4 * GU + kk, x + kk, hi_w, idesc1, 1u);
      }
      tc::mma_commit(empty + stg);
    }
    tc::mma_commit(done);
  }
  // ============================================================ epilogue: TMEM -> red.global.add.v4
  float* part = ring;                                   // [4 gku][128 col][4]: free once `done` has fired
  if (warp < WG_PROD_WARPS) {
    tc::mbar_wait(done, 0);
    tc::fence_after_sync();
    if (a.fold3) {
      const int gcol = tid & 127, gku = tid >> 7;
      *reinterpret_cast<float4*>(part + (gku * 128 + gcol) * 4) = make_float4(fold_dw3[0], fold_dw3[1], fold_dw3[2], fold_db2);
      if (gcol == 0) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
          if (fold_db3[ch] != 0.f) atomicAdd(a.db3 + ch, fold_db3[ch]);
      }
    }
    const int q = warp & 3, o = q * 32 + lane;
    const uint32_t t_lane = ((uint32_t)(q * 32)) << 16;
    const int ncb = 4 + (WG_NGX * 4 + 31) / 32;
    for (int cb = warp >> 2; cb < ncb; cb += WG_PROD_WARPS / 4) {         // 4 column blocks of dW2, then those of dW1
      float acc[32];
      tc::tmem_ld32(tmem + t_lane + (uint32_t)(cb * 32), acc);
      tc::tmem_wait_ld();
      float* dst = cb < 4 ? a.dw2 + (size_t)o * 128 + cb * 32 : a.dw1i + (size_t)o * (WG_NGX * 4) + (cb - 4) * 32;
      const int nu = cb < 4 ? 8 : min(8, WG_NGX - (cb - 4) * 8);
      if (cb >= 4 || a.dw2_vec) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (u < nu) red4g(dst + 4 * u, acc[4 * u], acc[4 * u + 1], acc[4 * u + 2], acc[4 * u + 3]);
      } else {
#pragma unroll
        for (int u = 0; u < 32; ++u) atomicAdd(dst + u, acc[u]);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (a.fold3 && tid < 128) {                           // sum the four sample-quarter partials of every hidden unit
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 p = *reinterpret_cast<const float4*>(part + (q * 128 + tid) * 4);
      s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
    }
    if (s.x != 0.f) atomicAdd(a.dw3 + tid, s.x);
    if (s.y != 0.f) atomicAdd(a.dw3 + 128 + tid, s.y);
    if (s.z != 0.f) atomicAdd(a.dw3 + 256 + tid, s.z);
    if (s.w != 0.f) atomicAdd(a.db2 + tid, s.w);
  }
  if (warp == WG_PROD_WARPS) tc::tmem_dealloc(tmem, 512);
}

// dW1 (internal column order, column kp = bias gradient) -> parameter gradients
__global__ void wgrad_unpermute_kernel(const float* __restrict__ dw1i, FusedSpec fs, int ld, float* __restrict__ g_w1, int ldw,
                                       float* __restrict__ g_b1) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 128 * ld) return;
  const int o = e / ld, k = e - o * ld;
  const float v = dw1i[e];
  if (k < fs.kp) {
    const int dst = fs.orig[k];
    if (dst >= 0 && v != 0.f) atomicAdd(g_w1 + (size_t)o * ldw + dst, v);
  } else if (k == fs.kp && g_b1 && v != 0.f) {
    atomicAdd(g_b1 + o, v);
  }
}

size_t wgrad_fused_scratch_bytes() { return (size_t)128 * WG_NGX_MAX * 4 * sizeof(float); }

int launch_wgrad_fused(const WgradFusedArgs& a_in, const FusedSpec& fs, float* g_w1, int ldw, float* g_b1, cudaStream_t st,
                       const FtrfCall* call) {
  WgradFusedArgs a = a_in;
  a.ngt = (a.kp / 4 + 1 + 3) & ~3;
  a.dw2_vec = ((uintptr_t)a.dw2 & 15) == 0;
  a.variant = g_wgrad_variant;
  if (a.kp < 4 || a.ngt > WG_NGX_MAX || (a.kp & 3) || (a.ldx & 3) || a.cap < 0) { set_error("wgrad_fused: bad shape kp=%d ldx=%d", a.kp, a.ldx); return FTRF_E_ARG; }
  if (a.cap == 0) return FTRF_OK;
  if (!(call && prep_ahead_take_scratch(*call)) &&
      cudaMemsetAsync(a.dw1i, 0, wgrad_fused_scratch_bytes(), st) != cudaSuccess) return check_launch("wgrad scratch memset");
  const size_t smem = (size_t)WG_STAGES * 2 * (3 * WG_G_FLOATS + WG_KS * a.ngt * 4) * 4 + 16 * 8;
  cudaFuncSetAttribute(wgrad_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t chunks = (a.cap + WG_KS - 1) / WG_KS;
  int grid = (int)(chunks < FTRF_SMS ? chunks : FTRF_SMS);
  wgrad_fused_kernel<<<grid, WG_THREADS, smem, st>>>(a);
  if (int e = check_launch("wgrad_fused")) return e;
  // nothing later in the backward reads g_w1 / g_b1: off the serial chain, joined at the end of the call
  cudaStream_t tail = call ? side_tail_begin(*call, st) : nullptr;
  wgrad_unpermute_kernel<<<(128 * a.ngt * 4 + 255) / 256, 256, 0, tail ? tail : st>>>(a.dw1i, fs, a.ngt * 4, g_w1, ldw, g_b1);
  if (tail) side_tail_end(*call);
  return check_launch("wgrad_unpermute");
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_debug_wgrad_variant(int v) { ftrf::g_wgrad_variant = v; return 0; }

// unit-test entry: identity column order
extern "C" int ftrf_test_wgrad_fused(const float* g2, const float* h1, const float* g1, const float* x, int ldx, int kp,
                                     int64_t M, float* dw2, float* dw1, int ldw, float* db1, float* scratch, void* stream) {
  ftrf::WgradFusedArgs a{};
  a.g2 = g2; a.h1 = h1; a.g1 = g1; a.x = x; a.ldx = ldx; a.kp = kp; a.count = nullptr; a.cap = M; a.dw2 = dw2;
  a.dw1i = scratch;
  ftrf::FusedSpec fs{};
  fs.kp = kp;
  for (int k = 0; k < kp; ++k) fs.orig[k] = (int16_t)(k < ldw ? k : -1);
  return ftrf::launch_wgrad_fused(a, fs, dw1, ldw, db1, (cudaStream_t)stream, nullptr);
}
#endif
