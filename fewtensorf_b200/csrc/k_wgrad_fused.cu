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
constexpr int WG_G8 = ACT16_G8;                     // floats of one 8-column group of a raw stage (16 rows x 8 + padding)
constexpr int WG_RAW_OPS = 2, WG_RAW_STAGES = 2;   // RAW build: operand stages (69.6 KB each) and raw stages (34.6 KB each); 1 + 4 measured slower
constexpr int WG_L2_AHEAD = 0;                     // chunks the bulk-copy warp prefetches into L2 ahead of its copies
constexpr int WG_THREADS = WG_PROD + 64;   // + the MMA-issuing warp and the bulk-copy warp
FTRF_SWITCH(int, g_wgrad_raw, 1);          // ftrf_debug_set_wgrad_raw (test build): 0 = producers load from global memory themselves


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

// RAW = true (shipped): the activation rows of a chunk — 16 consecutive samples, i.e. one contiguous block of each array —
// are brought into a two-stage shared-memory ring by cp.async.bulk (warp 17), and the producers transpose from THERE.
// With RAW = false the producers load from global memory themselves (four 4-byte loads per K-major unit, two register sets
// in flight): ncu showed them stalled on those loads 45 % of the time (L1 hit rate 4 %: every byte is used once) with DRAM at
// 29 % — memory-level parallelism bounded by registers.  Bulk copies need none: two stages keep 69 KB per SM in flight.
template <bool RAW>
__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_fused_kernel(const WgradFusedArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int NSTG = RAW ? WG_RAW_OPS : WG_STAGES, RS = WG_RAW_STAGES;
  float* ring = reinterpret_cast<float*>(smem_raw);                    // [NSTG][G2h,G2l,H1h,H1l,G1h,G1l,Xh,Xl]
  const int WG_NGX = a.ngt;
  const int WG_X_FLOATS = WG_KS * WG_NGX * 4;
  const int WG_STAGE_FLOATS = 2 * (3 * WG_G_FLOATS + WG_X_FLOATS);      // G2, H1, G1, X, each hi + lo
  // raw stage: h2 | h1 | g_h1 | X | go4.  Block layout of the saved activations (perm): every 8-column group of the 16 rows is
  // 512 contiguous bytes, copied to a 544-byte slot (WG_G8 floats) so that the four groups a warp reads sit in different banks
  const int RAW_A = 16 * WG_G8, RAW_X = 3 * RAW_A, RAW_GO = RAW_X + ((a.ldx + 7) / 8) * WG_G8, RAW_FLOATS = RAW_GO + WG_KS * 4;
  const bool perm = a.count != nullptr && a.count[WS_ACT_LAYOUT - 1] != 0;        // (count = ws.counters + 1)
  float* raw = ring + (size_t)NSTG * WG_STAGE_FLOATS;                  // [RS][RAW_FLOATS] (RAW only)
  uint64_t* bars = reinterpret_cast<uint64_t*>(raw + (RAW ? RS * (size_t)RAW_FLOATS : 0));
  uint64_t* full = bars;                      // [NSTG]  16 producer warps
  uint64_t* empty = bars + NSTG;              // [NSTG]  tcgen05.commit
  uint64_t* done = bars + 2 * NSTG;           // accumulators final
  uint64_t* raw_full = done + 1;              // [RS]  expect_tx + bulk copies
  uint64_t* raw_empty = raw_full + RS;        // [RS]  16 producer warps have read the stage
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(raw_empty + RS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float fold_dw3[3] = {0.f, 0.f, 0.f}, fold_db2 = 0.f, fold_db3[3] = {0.f, 0.f, 0.f};     // output-layer gradients (fold3)
  const int64_t M = a.count ? min((int64_t)*a.count, a.cap) : a.cap;
  const int64_t nchunks = (M + WG_KS - 1) / WG_KS;
  if ((int64_t)blockIdx.x >= nchunks) return;
  const int my_chunks = (int)((nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x);

  if (tid == 0) {
    for (int i = 0; i < NSTG; ++i) { tc::mbar_init(full + i, WG_PROD_WARPS); tc::mbar_init(empty + i, 1); }
    tc::mbar_init(done, 1);
    for (int i = 0; i < RS; ++i) { tc::mbar_init(raw_full + i, 1); tc::mbar_init(raw_empty + i, WG_PROD_WARPS); }
    tc::fence_barrier_init();
  }
  if (warp == WG_PROD_WARPS) tc::tmem_alloc(tmem_slot, 512);
  // constant part of every X stage: the ones column (bias gradient) and the zero padding columns
  for (int e = tid; e < NSTG * 2 * (WG_NGX * 4 - a.kp) * 4; e += WG_THREADS) {
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
    int xdst[2];                                          // operand offsets of this thread's X units (-1: none)
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int id = tid + WG_PROD * q;
      const int ku = id / a.kp, col = id - ku * a.kp;
      xdst[q] = id < xu ? tc::canon_off(col, ku, 4) : -1;
    }
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
        const size_t o = perm ? act16_off(s, gcol, 128) : (size_t)s * 128 + gcol;
        r[i] = ldg_early(a.g2 + o, ok);
        r[4 + i] = ldg_early(a.h1 + o, ok);
        r[8 + i] = ldg_early(a.g1 + o, ok);
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {                       // X: kp columns x 4 units
        const int id = tid + WG_PROD * q;
        const int ku = id / a.kp, col = id - ku * a.kp;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int64_t s = s0 + ku * 4 + i;
          r[12 + q * 4 + i] = ldg_early(a.x + (perm ? act16_off(s, col, a.ldx) : (size_t)s * a.ldx + col), id < xu && s < M);
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
    auto store = [&](int j, float (&r)[20], const float4* gor) {      // gor: this thread's four go4 records (RAW), else nullptr
      const int stg = j % NSTG;
      if (a.fold3) {
        // r[0..3] holds h2 of this thread's column for four samples: rebuild g_h2 and collect the output layer's gradients
        const int64_t s0 = ((int64_t)blockIdx.x + (int64_t)j * gridDim.x) * WG_KS + gku * 4;
        if (!RAW && j + 1 < my_chunks) {                // the next chunk's go values: into L1 while this chunk is written
          const float4* nx = a.go4 + ((int64_t)blockIdx.x + (int64_t)(j + 1) * gridDim.x) * WG_KS + gku * 4;
          asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 g = RAW ? gor[i] : (s0 + i < M ? __ldg(a.go4 + s0 + i) : make_float4(0.f, 0.f, 0.f, 0.f));
          const float h = r[i];
          dw3[0] = fmaf(g.x, h, dw3[0]); dw3[1] = fmaf(g.y, h, dw3[1]); dw3[2] = fmaf(g.z, h, dw3[2]);
          const float gh = h > 0.f ? g.x * w3c[0] + g.y * w3c[1] + g.z * w3c[2] : 0.f;
          db2 += gh;
          if (gcol == 0) { db3[0] += g.x; db3[1] += g.y; db3[2] += g.z; }
          r[i] = gh;
        }
      }
      tc::mbar_wait(empty + stg, ((j / NSTG) & 1) ^ 1);
      float* S = ring + (size_t)stg * WG_STAGE_FLOATS;
      const int off = tc::canon_off(gcol, gku, 4);
      put(S, S + WG_G_FLOATS, off, r);
      put(S + 2 * WG_G_FLOATS, S + 3 * WG_G_FLOATS, off, r + 4);
      put(S + 4 * WG_G_FLOATS, S + 5 * WG_G_FLOATS, off, r + 8);
#pragma unroll
      for (int q = 0; q < 2; ++q)
        if (xdst[q] >= 0) put(S + 6 * WG_G_FLOATS, S + 6 * WG_G_FLOATS + WG_X_FLOATS, xdst[q], r + 12 + q * 4);
      tc::fence_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(full + stg);
    };
    if constexpr (RAW) {
      // per-thread constants of the X part (a division by kp per chunk showed up as the hottest integer work)
      int xsrc[2], xrow[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int id = tid + WG_PROD * q;
        const int ku = id / a.kp, col = id - ku * a.kp;
        xrow[q] = id < xu ? ku * 4 : WG_KS;             // first sample row of the unit (WG_KS: no unit)
        xsrc[q] = RAW_X + (perm ? (col >> 3) * WG_G8 + ku * 4 * 8 + (col & 7) : ku * 4 * a.ldx + col);
      }
      const int gsrc = perm ? (gcol >> 3) * WG_G8 + gku * 4 * 8 + (gcol & 7) : gku * 4 * 128 + gcol;
      const int rstep = perm ? 8 : 128, xstep = perm ? 8 : a.ldx;      // floats between consecutive rows
      for (int j = 0; j < my_chunks; ++j) {
        const int rs = j % RS;
        const float* R = raw + (size_t)rs * RAW_FLOATS;
        const int64_t s0 = ((int64_t)blockIdx.x + (int64_t)j * gridDim.x) * WG_KS;
        const int nrow = (int)min((int64_t)WG_KS, M - s0);         // rows past the end were not copied: stale shared memory
        float (&r)[20] = v[0];
        float4 gor[4];
        tc::mbar_wait(raw_full + rs, (j / RS) & 1);
        if (nrow == WG_KS) {                              // every chunk but the last: no per-row masking
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            r[i] = R[gsrc + i * rstep];
            r[4 + i] = R[RAW_A + gsrc + i * rstep];
            r[8 + i] = R[2 * RAW_A + gsrc + i * rstep];
            gor[i] = a.fold3 ? *reinterpret_cast<const float4*>(R + RAW_GO + (gku * 4 + i) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int q = 0; q < 2; ++q) {
#pragma unroll
            for (int i = 0; i < 4; ++i) r[12 + q * 4 + i] = xrow[q] < WG_KS ? R[xsrc[q] + i * xstep] : 0.f;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int sl = gku * 4 + i;
            const bool ok = sl < nrow;
            r[i] = ok ? R[gsrc + i * rstep] : 0.f;
            r[4 + i] = ok ? R[RAW_A + gsrc + i * rstep] : 0.f;
            r[8 + i] = ok ? R[2 * RAW_A + gsrc + i * rstep] : 0.f;
            gor[i] = (ok && a.fold3) ? *reinterpret_cast<const float4*>(R + RAW_GO + sl * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int q = 0; q < 2; ++q) {                     // X: kp columns x 4 units
#pragma unroll
            for (int i = 0; i < 4; ++i) r[12 + q * 4 + i] = xrow[q] + i < nrow ? R[xsrc[q] + i * xstep] : 0.f;
          }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(raw_empty + rs);   // (release: this warp's reads of the stage are done)
        store(j, r, gor);
      }
    } else {
      load(0, v[0]);
      if (1 < my_chunks) load(1, v[1]);
      for (int j = 0; j < my_chunks; j += 2) {
        store(j, v[0], nullptr);
        if (j + 2 < my_chunks) load(j + 2, v[0]);
        if (j + 1 < my_chunks) {
          store(j + 1, v[1], nullptr);
          if (j + 3 < my_chunks) load(j + 3, v[1]);
        }
      }
    }
  } else if (warp == WG_PROD_WARPS + 1) {
    // ============================================================ bulk-copy warp: the raw rows of chunk j -> raw stage j % 2
    if (RAW && lane == 0) {
      // the shared-memory ring holds two chunks, i.e. one copy in flight while the other stage is read: far too little
      // for DRAM latency.  cp.async.bulk.prefetch.L2 needs no destination: chunks up to WG_L2_AHEAD ahead are pulled
      // into L2, and the copies into shared memory then pay L2 latency only.
      auto prefetch = [&](int jj) {
        if (jj >= my_chunks) return;
        const int64_t p0 = ((int64_t)blockIdx.x + (int64_t)jj * gridDim.x) * WG_KS;
        const uint32_t prow = (uint32_t)min((int64_t)WG_KS, M - p0);
        auto pf = [](const void* ptr, uint32_t bytes) {
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
        };
        pf(a.g2 + p0 * 128, prow * 512u);
        pf(a.h1 + p0 * 128, prow * 512u);
        pf(a.g1 + p0 * 128, prow * 512u);
        pf(a.x + p0 * a.ldx, prow * (uint32_t)(a.ldx * 4));
      };
      for (int jj = RS; jj < RS + WG_L2_AHEAD; ++jj) prefetch(jj);
      for (int j = 0; j < my_chunks; ++j) {
        prefetch(j + RS + WG_L2_AHEAD);
        const int rs = j % RS;
        float* R = raw + (size_t)rs * RAW_FLOATS;
        const int64_t s0 = ((int64_t)blockIdx.x + (int64_t)j * gridDim.x) * WG_KS;
        const uint32_t rows = (uint32_t)min((int64_t)WG_KS, M - s0);
        tc::mbar_wait(raw_empty + rs, ((j / RS) & 1) ^ 1);
        if (perm) {
          // whole 16-row blocks (the arrays are padded to them), padding included: the stage has the global layout
          const uint32_t hb = (uint32_t)act16_block_floats(128) * 4, xb = (uint32_t)act16_block_floats(a.ldx) * 4;
          const int64_t blk = s0 >> 4;
          tc::mbar_expect_tx(raw_full + rs, 3 * hb + xb + (a.fold3 ? rows * 16u : 0u));
          tc::bulk_g2s(R, a.g2 + blk * (hb / 4), hb, raw_full + rs);
          tc::bulk_g2s(R + RAW_A, a.h1 + blk * (hb / 4), hb, raw_full + rs);
          tc::bulk_g2s(R + 2 * RAW_A, a.g1 + blk * (hb / 4), hb, raw_full + rs);
          tc::bulk_g2s(R + RAW_X, a.x + blk * (xb / 4), xb, raw_full + rs);
        } else {
          tc::mbar_expect_tx(raw_full + rs, rows * (uint32_t)(3 * 512 + a.ldx * 4 + (a.fold3 ? 16 : 0)));
          tc::bulk_g2s(R, a.g2 + s0 * 128, rows * 512u, raw_full + rs);
          tc::bulk_g2s(R + RAW_A, a.h1 + s0 * 128, rows * 512u, raw_full + rs);
          tc::bulk_g2s(R + 2 * RAW_A, a.g1 + s0 * 128, rows * 512u, raw_full + rs);
          tc::bulk_g2s(R + RAW_X, a.x + s0 * a.ldx, rows * (uint32_t)(a.ldx * 4), raw_full + rs);
        }
        if (a.fold3) tc::bulk_g2s(R + RAW_GO, a.go4 + s0, rows * 16u, raw_full + rs);
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
      const int stg = j % NSTG;
      tc::mbar_wait(full + stg, (j / NSTG) & 1);
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
  const size_t stage_b = (size_t)2 * (3 * WG_G_FLOATS + WG_KS * a.ngt * 4) * 4;
  const size_t raw_b = (size_t)(3 * 16 * WG_G8 + ((a.ldx + 7) / 8) * WG_G8 + WG_KS * 4) * 4;
  const bool use_raw = g_wgrad_raw && WG_RAW_OPS * stage_b + WG_RAW_STAGES * raw_b + 256 <= 227 * 1024 &&
                       (((uintptr_t)a.g2 | (uintptr_t)a.h1 | (uintptr_t)a.g1 | (uintptr_t)a.x | (uintptr_t)a.go4) & 15) == 0;
  const size_t smem = use_raw ? WG_RAW_OPS * stage_b + WG_RAW_STAGES * raw_b + 24 * 8 : (size_t)WG_STAGES * stage_b + 24 * 8;
  int64_t chunks = (a.cap + WG_KS - 1) / WG_KS;
  int grid = (int)(chunks < FTRF_SMS ? chunks : FTRF_SMS);
  if (use_raw) {
    cudaFuncSetAttribute(wgrad_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    wgrad_fused_kernel<true><<<grid, WG_THREADS, smem, st>>>(a);
  } else {
    cudaFuncSetAttribute(wgrad_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    wgrad_fused_kernel<false><<<grid, WG_THREADS, smem, st>>>(a);
  }
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
extern "C" int ftrf_debug_set_wgrad_raw(int on) { ftrf::g_wgrad_raw = on; return 0; }

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
