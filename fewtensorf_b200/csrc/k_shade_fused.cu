// k_shade_fused.cu — stage 3 forward as ONE persistent, warp-specialised tcgen05 kernel:
//   positional encoding (+ FreeNeRF encoding masks) -> Linear+ReLU -> Linear+ReLU -> Linear -> sigmoid
// (models/mlp.py:8-13, :41-69, :88-107, :127-155) for a pair of 128-sample tiles at a time, without
// sending X / h1 / h2 through HBM (they are written out only when saved for backward).
//
//   warps 0-7  (256 threads, one per sample row; warps 0-3 = sub-tile a, 4-7 = sub-tile b):
//              build the encoded input rows chunk by chunk (32 columns) straight into the UMMA
//              operand ring (tf32 hi / lo split in registers), later turn the layer-1 accumulator
//              (tcgen05.ld) into the layer-2 operand chunks (bias + ReLU + split), and finish with
//              the layer-2 epilogue: bias + ReLU, the 128->3 layer on CUDA cores, sigmoid.
//   warp 8     one elected lane issues tcgen05.mma.kind::tf32 (3xTF32: Al*Bh + Ah*Bl + Ah*Bh, fp32
//              accumulate in TMEM, fp32-grade accuracy) for every (activation chunk, weight chunk).
//   warp 9     one elected lane streams the pre-split weight chunks (32 KB each, prepared once per
//              call by prep_weights_kernel) from L2 into a shared-memory ring with cp.async.bulk
//              (TMA), completion on mbarriers.  Each weight chunk serves both sub-tiles.
// Roles meet only through mbarriers: a_full / a_empty (operand ring), w_full / w_empty (weight
// ring), acc1_full / acc2_full (tcgen05.commit).  TMEM: acc1[a,b] columns 0-255, acc2[a,b] 256-511.
#include "fused.cuh"
#include "tc.cuh"

namespace ftrf {

constexpr int FZ_A_STAGES = 2;
constexpr int FZ_W_STAGES = 2;
constexpr int FZ_CHUNK_FLOATS = 128 * 32;              // one hi (or lo) [128 x 32] chunk
constexpr int FZ_THREADS = 352;                        // 8 compute warps, two MMA-issuing warps (one per sub-tile), the weight streamer
constexpr int FZ_PAIR_ROWS = 256;

// ------------------------------------------------------------------------------------------ prep
// Weight images (see fused.cuh).  One launch per forward call: the weights change every step.
__global__ void __launch_bounds__(256)
prep_weights_kernel(const FtrfField f, const FusedSpec fs, float* __restrict__ img, float* __restrict__ w1p, int in_dim) {
  const int nchunks = fs.n1 + 4;
  const int total = nchunks * 128 * 8;                  // (chunk, row n, 16-byte unit ku)
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int ku = e & 7, n = (e >> 3) & 127, ch = e >> 10;
    float hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v;
      if (ch < fs.n1) {
        int k = ch * 32 + ku * 4 + j;
        int o = k < fs.kp ? fs.orig[k] : -1;
        v = o >= 0 ? __ldg(f.w1 + (size_t)n * in_dim + o) : 0.f;
      } else {
        v = __ldg(f.w2 + (size_t)n * FC + (ch - fs.n1) * 32 + ku * 4 + j);
      }
      tc::split_tf32(v, hi[j], lo[j]);
    }
    float* base = img + (size_t)ch * (2 * FZ_CHUNK_FLOATS);
    const int off = tc::canon_off(n, ku, 8);
    *reinterpret_cast<float4*>(base + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(base + FZ_CHUNK_FLOATS + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
  if (w1p) {     // plain fp32 W1 in internal column order, [128][kp] (backward: g_X = g_h1 W1)
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < FC * fs.kp; e += gridDim.x * blockDim.x) {
      int n = e / fs.kp, k = e - n * fs.kp, o = fs.orig[k];
      w1p[e] = o >= 0 ? __ldg(f.w1 + (size_t)n * in_dim + o) : 0.f;
    }
  }
}

// ------------------------------------------------------------------------------------------ PTX bits
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(tc::smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(tc::smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void split4(const float v[4], float4& h, float4& l) {
  tc::split_tf32(v[0], h.x, l.x); tc::split_tf32(v[1], h.y, l.y);
  tc::split_tf32(v[2], h.z, l.z); tc::split_tf32(v[3], h.w, l.w);
}

// sin / cos with an explicit two-constant Cody-Waite reduction to [-pi, pi] followed by the SFU
// approximations (abs error ~2^-21 there): ~12 instructions instead of sincosf()'s ~50, which made the
// input build and the encoding chain rule the longest phases of the fused kernels.  Absolute error stays
// ~1e-6 for |x| up to ~1e4 (the PE arguments are app features times 2^f).
__device__ __forceinline__ void fast_sincos(float x, float* s, float* c) {
  const float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(-k, 6.2831854820251465f, x);          // 2 pi, high part (fp32)
  r = fmaf(-k, -1.7484556000744883e-7f, r);            // 2 pi, low part
  __sincosf(r, s, c);
}

// one full 32-byte sector per store instruction (sm_100: 256-bit global stores)
__device__ __forceinline__ void st_global_v8(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
               "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}

// Bit i = (v[i] > 0) for 32 values that are >= +0 (outputs of a ReLU): as unsigned integers they are non-zero exactly when
// positive, so `v + 0xffffffff` carries out exactly then, and `m = m + m + carry` shifts the bit in — two integer
// instructions per value (the dgrad kernel reads these 4 bytes instead of 128 bytes of activations).
__device__ __forceinline__ uint32_t relu_sign_bits(const float v[32]) {
  uint32_t p[4] = {0u, 0u, 0u, 0u};          // four independent carry chains of eight values each
#pragma unroll
  for (int i = 7; i >= 0; --i) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t junk;
      asm("{\n\tadd.cc.u32 %1, %2, 0xffffffff;\n\taddc.u32 %0, %0, %0;\n\t}" : "+r"(p[q]), "=r"(junk) : "r"(__float_as_uint(v[q * 8 + i])));
    }
  }
  return p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24);
}

struct SPair { int meta; float scale, ma, mb; };     // meta = kind | sa << 8 | sb << 16

// Compile-time copy of make_fused_spec()'s column recipe for the shipped configurations: with every
// loop of the input build unrolled, the per-row source values stay in registers (the generic path
// indexes them dynamically, i.e. through local memory, which the 200 KB of operand rings leave
// almost no L1 for).
struct PairInfo { int kind, sa, sb, fexp; };
template <int MODE, int D_, int FPOS, int FVIEW, int FFEA>
struct StaticSpec {
  static constexpr bool is_static = true;
  static constexpr int D = D_;
  static constexpr int ncopy = D_ + 3, ncp = (ncopy + 1) / 2;
  static constexpr int nb = MODE == FTRF_SHADE_MLP ? 3 : 2;
  __host__ __device__ static constexpr int blk_src(int b) {
    return MODE == FTRF_SHADE_MLP_FEA ? (b == 0 ? SRC_FEA : SRC_VIEW)
         : MODE == FTRF_SHADE_MLP_PE ? (b == 0 ? SRC_POS : SRC_VIEW)
         : (b == 0 ? SRC_POS : (b == 1 ? SRC_VIEW : SRC_FEA));
  }
  __host__ __device__ static constexpr int blk_f(int b) { return blk_src(b) == SRC_FEA ? FFEA : (blk_src(b) == SRC_VIEW ? FVIEW : FPOS); }
  __host__ __device__ static constexpr int blk_c(int b) { return blk_src(b) == SRC_FEA ? D_ : 3; }
  __host__ __device__ static constexpr int blk_s0(int b) { return blk_src(b) == SRC_FEA ? 0 : (blk_src(b) == SRC_VIEW ? D_ : D_ + 3); }
  __host__ __device__ static constexpr int npairs() {
    int n = ncp;
    for (int b = 0; b < nb; ++b) n += blk_f(b) * blk_c(b);
    return (n + 3) & ~3;
  }
  static constexpr int kp = npairs() * 2;
  static constexpr int n1 = (kp + 31) / 32;
  __host__ __device__ static constexpr PairInfo pair(int P) {
    if (P < ncp) return (2 * P + 1 < ncopy) ? PairInfo{PK_COPY2, 2 * P, 2 * P + 1, 0} : PairInfo{PK_COPY1, 2 * P, 0, 0};
    int q = P - ncp;
    for (int b = 0; b < nb; ++b) {
      const int n = blk_f(b) * blk_c(b);
      if (q < n) return PairInfo{PK_TRIG, blk_s0(b) + q / blk_f(b), 0, q % blk_f(b)};
      q -= n;
    }
    return PairInfo{PK_ZERO, 0, 0, 0};
  }
};
struct DynamicSpec { static constexpr bool is_static = false; };

// optional phase timeline of CTA 0 (debug aid, scripts/fused_timeline.py): dbg[role*512 + n]
FTRF_SWITCH(long long*, g_fused_dbg, nullptr);
FTRF_SWITCH(long long*, g_dgrad_dbg, nullptr);
FTRF_SWITCH(int, g_act_perm, 1);           // ftrf_debug_set_act_perm (test build): 0 = saved activations stay row-major
FTRF_SWITCH(int, g_dgrad_masks, 1);        // ftrf_debug_set_dgrad_masks (test build): 0 = the dgrad kernel reads h1 / h2 back
#define FZ_T(role, n) do { if (dbg && blockIdx.x == 0 && (n) < 512) dbg[(role) * 512 + (n)] = clock64(); } while (0)


// ------------------------------------------------------------------------------------------ kernel
template <class SS>
__global__ void __launch_bounds__(FZ_THREADS, 1)
shade_fused_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const FusedSpec fs, const float* __restrict__ wimg, long long* dbg,
                       const int perm) {       // perm: save X / h1 / h2 in the 16-row block layout (act16_off), else row-major
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  float* a_ring = reinterpret_cast<float*>(smem_raw);                               // [A_STAGES][2 subs][hi, lo][4096]
  float* w_ring = a_ring + (size_t)FZ_A_STAGES * 4 * FZ_CHUNK_FLOATS;               // [W_STAGES][hi, lo][4096]
  float* s_b1 = w_ring + (size_t)FZ_W_STAGES * 2 * FZ_CHUNK_FLOATS;                 // [128]
  float* s_b2 = s_b1 + FC;                                                          // [128]
  float* s_w3 = s_b2 + FC;                                                          // [3][128]
  SPair* s_pair = reinterpret_cast<SPair*>(s_w3 + 3 * FC);                          // [kp/2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_pair + FZ_MAX_K / 2);
  uint64_t* a_full = bars;                       // [A_STAGES]  8 warp arrivals
  uint64_t* a_empty = a_full + FZ_A_STAGES;      // [A_STAGES]  tcgen05.commit
  uint64_t* w_full = a_empty + FZ_A_STAGES;      // [W_STAGES]  expect_tx + bulk copy
  uint64_t* w_empty = w_full + FZ_W_STAGES;      // [W_STAGES]  tcgen05.commit
  uint64_t* acc_full = w_empty + FZ_W_STAGES;    // [2]         layer 1, layer 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if (blockIdx.x == 0 && tid == 0 && c.save_for_backward) ws.counters[WS_ACT_LAYOUT] = perm;     // the layout travels with the data
  const int64_t npairs = (n + FZ_PAIR_ROWS - 1) / FZ_PAIR_ROWS;
  if ((int64_t)blockIdx.x >= npairs) return;
  const int my_pairs = (int)((npairs - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const int n1 = fs.n1, nck = n1 + 4;

  if (tid == 0) {
    for (int i = 0; i < FZ_A_STAGES; ++i) { tc::mbar_init(a_full + i, 8); tc::mbar_init(a_empty + i, 2); }
    for (int i = 0; i < FZ_W_STAGES; ++i) { tc::mbar_init(w_full + i, 1); tc::mbar_init(w_empty + i, 2); }
    tc::mbar_init(acc_full + 0, 2);                // one commit per issuing thread
    tc::mbar_init(acc_full + 1, 2);
    tc::fence_barrier_init();
  }
  if (warp == 8) tc::tmem_alloc(tmem_slot, 512);
  for (int e = tid; e < FC; e += FZ_THREADS) { s_b1[e] = __ldg(f.b1 + e); s_b2[e] = __ldg(f.b2 + e); }
  for (int e = tid; e < 3 * FC; e += FZ_THREADS) s_w3[e] = __ldg(f.w3 + e);
  for (int e = tid; e < fs.kp / 2; e += FZ_THREADS) {
    const PairRec r = fs.pair[e];
    SPair p;
    p.meta = (int)r.kind | ((int)r.sa << 8) | ((int)r.sb << 16);
    p.scale = (float)(1 << r.fexp);
    p.ma = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.ma) : 1.f;
    p.mb = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.mb) : 1.f;
    s_pair[e] = p;
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 8) {
    // =============================================================== compute warps (one row per thread)
    const int sub = warp >> 2, q = warp & 3, r_in = q * 32 + lane;
    const uint32_t t_lane = ((uint32_t)(q * 32)) << 16;
    const uint32_t t_acc1 = tmem + t_lane + (uint32_t)(sub * 128);
    const uint32_t t_acc2 = tmem + t_lane + (uint32_t)(256 + sub * 128);
    const float b3[3] = {__ldg(f.b3), __ldg(f.b3 + 1), __ldg(f.b3 + 2)};
    const bool save = c.save_for_backward != 0;
    const int D = fs.D;
    const bool tl = warp == 0 && lane == 0;
    uint32_t ac = 0;                                   // operand-ring chunk counter
    auto pair_row = [&](int it) { return ((int64_t)blockIdx.x + (int64_t)it * gridDim.x) * FZ_PAIR_ROWS + sub * 128 + r_in; };
    // sources of one row: [feat(D) | view(3) | pos(3)]
    auto load_src = [&](int64_t row, float* dst) {
      const bool live = row < n;
      if constexpr (SS::is_static) {
#pragma unroll
        for (int k = 0; k < SS::D + 6; ++k) dst[k] = 0.f;
        if (live) {
          const float* fp = ws.a_feat + row * SS::D;
#pragma unroll
          for (int k = 0; k < SS::D; ++k) dst[k] = fp[k];
          float v[3], p[3];
          shade_geometry(f, c, ws, row, v, p);
#pragma unroll
          for (int k = 0; k < 3; ++k) { dst[SS::D + k] = v[k]; dst[SS::D + 3 + k] = p[k]; }
        }
      } else {
        if (live) {
          const float* fp = ws.a_feat + row * D;
          for (int k = 0; k < D; ++k) dst[k] = fp[k];
          float v[3], p[3];
          shade_geometry(f, c, ws, row, v, p);
#pragma unroll
          for (int k = 0; k < 3; ++k) { dst[D + k] = v[k]; dst[D + 3 + k] = p[k]; }
        } else {
          for (int k = 0; k < D + 6; ++k) dst[k] = 0.f;
        }
      }
    };
    float src[FZ_MAX_SRC];                             // sources of the pair whose input rows are being built
    float da_s = 0.f, da_c = 1.f;                      // running (sin, cos) of the current source at the current octave

    // one 32-column chunk of the encoded input rows -> operand ring (+ X saved for backward)
    auto x_chunk = [&](int ch, int64_t row) {
      const bool live = row < n;
      const int stg = ac % FZ_A_STAGES;
      tc::mbar_wait(a_empty + stg, ((ac / FZ_A_STAGES) & 1) ^ 1);
      float* Ahi = a_ring + ((size_t)stg * 2 + sub) * (2 * FZ_CHUNK_FLOATS);
      float* Alo = Ahi + FZ_CHUNK_FLOATS;
      float xs[4];                                        // the even unit of a pair, until its odd neighbour is ready
      auto emit = [&](int u, const float v[4]) {
        float4 h, l;
        split4(v, h, l);
        const int off = tc::canon_off(r_in, u, 8);
        *reinterpret_cast<float4*>(Ahi + off) = h;
        *reinterpret_cast<float4*>(Alo + off) = l;
        // X saved for the weight-gradient kernel: two units per store = one full 32-byte sector (kp is a multiple of 8;
        // 16-byte stores left half-written sectors behind: 15 us of the kernel went into them)
        if (u & 1) {
          if (save && live) {
            const float w[8] = {xs[0], xs[1], xs[2], xs[3], v[0], v[1], v[2], v[3]};
            st_global_v8(ws.a_x + (perm ? act16_off(row, ch * 32 + (u - 1) * 4, ws.ldx) : (size_t)row * ws.ldx + ch * 32 + (u - 1) * 4), w);
          }
        } else {
          xs[0] = v[0]; xs[1] = v[1]; xs[2] = v[2]; xs[3] = v[3];
        }
      };
      if constexpr (SS::is_static) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (ch * 32 + u * 4 < SS::kp) {
            float v[4];
#pragma unroll
            for (int p2 = 0; p2 < 2; ++p2) {
              const int P = (ch * 32 + u * 4) / 2 + p2;
              const PairInfo pi = SS::pair(P);
              float x0 = 0.f, x1 = 0.f;
              if (pi.kind == PK_TRIG) {
                const SPair pr = s_pair[P];
                // octaves 0, 4, 8, ...: sincos of the scaled argument; the others: double-angle step from the octave
                // below (the pairs of one channel are consecutive, frequency-minor).  Each step doubles the absolute
                // error, so at most three are chained (pos_pe 8 of configs/free_tensorf_human.txt would chain seven).
                if ((pi.fexp & 3) == 0) fast_sincos(src[pi.sa] * (float)(1 << pi.fexp), &da_s, &da_c);
                else { const float s2 = 2.f * da_s * da_c; da_c = fmaf(-2.f * da_s, da_s, 1.f); da_s = s2; }
                x0 = da_s * pr.ma;
                x1 = da_c * pr.mb;
              } else if (pi.kind == PK_COPY2) {
                x0 = src[pi.sa];
                x1 = src[pi.sb];
              } else if (pi.kind == PK_COPY1) {
                x0 = src[pi.sa];
              }
              v[2 * p2] = x0;
              v[2 * p2 + 1] = x1;
            }
            emit(u, v);
          }
        }
      } else {
        const int units = min(8, (fs.kp - ch * 32) / 4);
        for (int u = 0; u < units; ++u) {
          float v[4];
#pragma unroll
          for (int p2 = 0; p2 < 2; ++p2) {
            const SPair pr = s_pair[(ch * 32 + u * 4) / 2 + p2];
            const int kind = pr.meta & 0xff, sa = (pr.meta >> 8) & 0xff, sb = (pr.meta >> 16) & 0xff;
            float x0 = 0.f, x1 = 0.f;
            if (kind == PK_TRIG) {
              if (((__float_as_int(pr.scale) >> 23) & 3) == 3) fast_sincos(src[sa] * pr.scale, &da_s, &da_c);   // 2^f with f % 4 == 0
              else { const float s2 = 2.f * da_s * da_c; da_c = fmaf(-2.f * da_s, da_s, 1.f); da_s = s2; }
              x0 = da_s * pr.ma;
              x1 = da_c * pr.mb;
            } else if (kind == PK_COPY2) {
              x0 = src[sa];
              x1 = src[sb];
            } else if (kind == PK_COPY1) {
              x0 = src[sa];
            }
            v[2 * p2] = x0;
            v[2 * p2 + 1] = x1;
          }
          emit(u, v);
        }
      }
      tc::fence_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(a_full + stg);
      ++ac;
    };
    // Schedule per pair `it` (the operand ring keeps production order = MMA consumption order):
    //   tail chunks of X(it) | wait acc1 | h1 chunks (layer-2 operand) | HEAD chunks of X(it+1) | wait acc2 | epilogue(it)
    // so the tensor pipe already works on the next pair's first layer while this pair's epilogue runs.
    constexpr int HEAD = FZ_A_STAGES;
    load_src(pair_row(0), src);
    if constexpr (SS::is_static) {
#pragma unroll
      for (int ch = 0; ch < (HEAD < SS::n1 ? HEAD : SS::n1); ++ch) x_chunk(ch, pair_row(0));
    } else {
      for (int ch = 0; ch < min(HEAD, n1); ++ch) x_chunk(ch, pair_row(0));
    }
    for (int it = 0; it < my_pairs; ++it) {
      const int64_t row = pair_row(it), row_next = pair_row(it + 1);
      const bool live = row < n, more = it + 1 < my_pairs;
      if (tl) FZ_T(0, it * 8 + 0);
      if constexpr (SS::is_static) {
#pragma unroll
        for (int ch = HEAD; ch < SS::n1; ++ch) x_chunk(ch, row);
      } else {
        for (int ch = HEAD; ch < n1; ++ch) x_chunk(ch, row);
      }
      if (tl) FZ_T(0, it * 8 + 1);
      if (more) load_src(row_next, src);               // in flight during the layer-1 epilogue below
      if (tl) FZ_T(0, it * 8 + 2);
      // ---- layer-1 accumulator -> layer-2 operand chunks (bias + ReLU + split)
      tc::mbar_wait(acc_full + 0, it & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(0, it * 8 + 3);
      // ReLU sign bits of h1 (words 0..3) / h2 (4..7) for the dgrad kernel, word-major [8][cap_app]: a warp's store of one
      // word is 128 contiguous bytes
      uint32_t* const sgn = reinterpret_cast<uint32_t*>(ws.a_gx) + row;
      for (int j = 0; j < 4; ++j, ++ac) {
        const int stg = ac % FZ_A_STAGES;
        float acc[32];
        tc::tmem_ld32(t_acc1 + (uint32_t)(j * 32), acc);
        tc::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = fmaxf(acc[i] + s_b1[j * 32 + i], 0.f);
        if (save && live) {
#pragma unroll
          for (int u = 0; u < 4; ++u) st_global_v8(ws.a_h1 + (perm ? act16_off(row, j * 32 + u * 8, FC) : (size_t)row * FC + j * 32 + u * 8), acc + 8 * u);
          sgn[(size_t)j * c.cap_app] = relu_sign_bits(acc);
        }
        tc::mbar_wait(a_empty + stg, ((ac / FZ_A_STAGES) & 1) ^ 1);
        float* Ahi = a_ring + ((size_t)stg * 2 + sub) * (2 * FZ_CHUNK_FLOATS);
        float* Alo = Ahi + FZ_CHUNK_FLOATS;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float4 h, l;
          split4(acc + 4 * u, h, l);
          const int off = tc::canon_off(r_in, u, 8);
          *reinterpret_cast<float4*>(Ahi + off) = h;
          *reinterpret_cast<float4*>(Alo + off) = l;
        }
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(a_full + stg);
      }
      if (tl) FZ_T(0, it * 8 + 4);
      // ---- head chunks of the next pair's input rows (acc1 is free: every warp has read it)
      if (more) {
        if constexpr (SS::is_static) {
#pragma unroll
          for (int ch = 0; ch < (HEAD < SS::n1 ? HEAD : SS::n1); ++ch) x_chunk(ch, row_next);
        } else {
          for (int ch = 0; ch < min(HEAD, n1); ++ch) x_chunk(ch, row_next);
        }
      }
      // ---- layer-2 epilogue: bias + ReLU, layer 3 (128 -> 3) on CUDA cores, sigmoid
      tc::mbar_wait(acc_full + 1, it & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(0, it * 8 + 5);
      float o3[3] = {b3[0], b3[1], b3[2]};
      for (int j = 0; j < 4; ++j) {
        float acc[32];
        tc::tmem_ld32(t_acc2 + (uint32_t)(j * 32), acc);
        tc::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = fmaxf(acc[i] + s_b2[j * 32 + i], 0.f);
        if (save && live) {
#pragma unroll
          for (int u = 0; u < 4; ++u) st_global_v8(ws.a_h2 + (perm ? act16_off(row, j * 32 + u * 8, FC) : (size_t)row * FC + j * 32 + u * 8), acc + 8 * u);
          sgn[(size_t)(4 + j) * c.cap_app] = relu_sign_bits(acc);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const float4* wr = reinterpret_cast<const float4*>(s_w3 + k * FC + j * 32);
          float sacc = 0.f;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            float4 w4 = wr[u];
            sacc = fmaf(acc[4 * u], w4.x, sacc); sacc = fmaf(acc[4 * u + 1], w4.y, sacc);
            sacc = fmaf(acc[4 * u + 2], w4.z, sacc); sacc = fmaf(acc[4 * u + 3], w4.w, sacc);
          }
          o3[k] += sacc;
        }
      }
      if (tl) FZ_T(0, it * 8 + 6);
      tc::fence_before_sync();     // accumulator reads are ordered before the a_full arrivals that follow
      if (live) {
#pragma unroll
        for (int k = 0; k < 3; ++k) ws.a_rgb[row * 3 + k] = 1.f / (1.f + expf(-o3[k]));
      }
    }
  } else if (warp == 8 || warp == 9) {
    // =============================================================== MMA issuers: warp 8 sub-tile a, warp 9 sub-tile b.
    // What an issuing thread does between dispatches is not hidden behind them (two waits + three commits per chunk cost
    // as much as 10 of its 12 MMAs, scripts/issue_probe.py), but two threads interleave on the tensor pipe: each one's
    // overhead runs behind the other's products.  A stage is free when both have committed (barrier counts 2).
    if (lane == 0) {
      const int sub = warp - 8;
      const uint32_t idesc = tc::idesc_tf32(128, 128);
      const uint32_t a_lo0 = tc::desc_lo(tc::smem_u32(a_ring), 128), w_lo0 = tc::desc_lo(tc::smem_u32(w_ring), 128);
      const uint32_t d_hi = tc::desc_hi(1024);
      uint32_t ac = 0, wc = 0;
      for (int it = 0; it < my_pairs; ++it) {
        for (int ch = 0; ch < nck; ++ch, ++ac, ++wc) {
          const int as = ac % FZ_A_STAGES, wsg = wc % FZ_W_STAGES;
          const bool l1 = ch < n1;
          const int ksteps = l1 ? min(32, fs.kp - ch * 32) / 8 : 4;
          if (sub == 0) FZ_T(1, wc * 3 + 0);
          tc::mbar_wait(w_full + wsg, (wc / FZ_W_STAGES) & 1);
          if (sub == 0) FZ_T(1, wc * 3 + 1);
          tc::mbar_wait(a_full + as, (ac / FZ_A_STAGES) & 1);
          if (sub == 0) FZ_T(1, wc * 3 + 2);
          tc::fence_after_sync();
          // descriptor low words: address >> 4 (+ LBO field); a k-step advances by 256 B = 16 units
          const uint32_t w_lo = w_lo0 + (uint32_t)wsg * (2 * FZ_CHUNK_FLOATS * 4 / 16);
          const uint32_t a_lo = a_lo0 + (uint32_t)as * (4 * FZ_CHUNK_FLOATS * 4 / 16);
          const uint32_t d0 = tmem + (uint32_t)(l1 ? 0 : 256);
          const uint32_t fresh = (l1 ? ch == 0 : ch == n1) ? 0u : 1u;
          constexpr uint32_t LO = FZ_CHUNK_FLOATS * 4 / 16;      // hi -> lo half of a chunk, in 16-byte units
          {
            const uint32_t ah = a_lo + (uint32_t)sub * (2 * LO), d = d0 + (uint32_t)(sub * 128);
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              if (s < ksteps) {
                tc::mma_tf32_ss_lo(d, ah + LO + s * 16, w_lo + s * 16, d_hi, idesc, s == 0 ? fresh : 1u);
                tc::mma_tf32_ss_lo(d, ah + s * 16, w_lo + LO + s * 16, d_hi, idesc, 1u);
                tc::mma_tf32_ss_lo(d, ah + s * 16, w_lo + s * 16, d_hi, idesc, 1u);
              }
            }
          }
          tc::mma_commit(a_empty + as);
          tc::mma_commit(w_empty + wsg);
          if (ch == n1 - 1) tc::mma_commit(acc_full + 0);
          if (ch == nck - 1) tc::mma_commit(acc_full + 1);
        }
      }
    }
  } else {
    // =============================================================== weight-chunk streamer (TMA)
    if (lane == 0) {
      uint32_t wc = 0;
      for (int it = 0; it < my_pairs; ++it)
        for (int ch = 0; ch < nck; ++ch, ++wc) {
          const int wsg = wc % FZ_W_STAGES;
          tc::mbar_wait(w_empty + wsg, ((wc / FZ_W_STAGES) & 1) ^ 1);
          FZ_T(2, wc);
          mbar_expect_tx(w_full + wsg, (uint32_t)FZ_CHUNK_BYTES);
          bulk_g2s(w_ring + (size_t)wsg * 2 * FZ_CHUNK_FLOATS, wimg + (size_t)ch * 2 * FZ_CHUNK_FLOATS, (uint32_t)FZ_CHUNK_BYTES,
                   w_full + wsg);
        }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 8) tc::tmem_dealloc(tmem, 512);
}

#ifdef FTRF_TEST_BUILD
// ================================================================================================
// TEST BUILD ONLY — an experiment that did NOT beat the kernel above and is kept for the measurements it carries
// (scripts/issue_probe.py, scripts/fused_timeline.py): 104 us per training step against 89 us,
// 11.4 ms against 10.6 ms per MLP_Fea frame.  What it taught: (i) tcgen05.st / tcgen05.ld traffic of the compute warps
// slows concurrent MMAs by ~40 % (issue probe with noise warps), so moving the operand from shared memory to tensor
// memory only trades one contention for another; (ii) with 256 TMEM columns left beside the two accumulators the operand
// ring is two stages deep, and every producer step pays a commit -> wait -> tcgen05.wait::st -> arrive round trip of
// ~1 k clocks that such a shallow ring cannot hide.
// Forward, second generation ("TS"): the activation operand of both layers lives in TENSOR MEMORY, written there by
// the compute warps with tcgen05.st (thread = sample row, one 32-bit column per k, tf32 hi | lo), and the MMAs take
// it from there (tcgen05.mma with A in TMEM); shared memory only carries the streamed weight chunks.
// What shapes it (scripts/issue_probe.py, measured): whatever the MMA-issuing thread does between dispatches is NOT
// hidden behind them — a successful barrier wait costs it ~225 clk (any shared-memory round trip does), the first
// commit of a batch ~125, further ones ~50 — so 12-MMA operand chunks with two waits and three commits each ran at
// 113 clk per MMA (the first-generation kernel above).  Here an operand GROUP is 64 k-columns (24 MMAs), the issuer
// waits ONCE per group (the weight streamer forwards the weights' arrival to the operand barrier) and commits twice.
//   warps 0-15  thread = (sample row, slice t): warp % 4 = TMEM lane quadrant (32 rows of the 128-row tile), warp / 4 = t.
//               All sixteen warps build every operand group together (slice t = its columns 16 t .. 16 t + 15): X groups
//               for layer 1, bias + ReLU + split of the layer-1 accumulator for layer 2.  The layer-2 epilogue (bias +
//               ReLU, partial 128 -> 3 product) goes by 32-column slice; partials meet in shared memory.
//   warp 16     issues every MMA (3xTF32), in ring order X groups, h1 groups per tile
//   warp 17     streams the weight chunks (cp.async.bulk, 3 stages x 64 KB) and arrives on a group's barrier for them
// TMEM: acc1 columns 0-127, acc2 128-255, operand ring 2 stages x (64 hi + 64 lo) columns at 256.
// A warp runs  epi1(i) -> X(i+1) -> epi2(i)  per tile; with the in-order issue that is also what protects the
// accumulators (L1 of tile i+1 follows every warp's reads of acc1(i), L2 of tile i+1 every warp's epi2(i)).
constexpr int FT_CWARPS = 16;
constexpr int FT_THREADS = (FT_CWARPS + 2) * 32;
constexpr int FT_R_STAGES = 2;                     // operand ring (TMEM), 128 columns per stage
constexpr int FT_W_STAGES = 3;                     // weight ring (shared memory), 64 KB per stage
constexpr int FT_GK = 64;                          // k-columns of an operand group
constexpr int FT_TILE = 128;
constexpr uint32_t FT_ACC1 = 0, FT_ACC2 = 128, FT_RING = 256;
int g_fwd_ts = 0;                                  // ftrf_debug_set_fwd_ts: 1 = this kernel instead of the SS kernel above

template <int V> struct IC { static constexpr int value = V; };

template <class SS>
__global__ void __launch_bounds__(FT_THREADS, 1)
shade_fwd_ts_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const FusedSpec fs, const float* __restrict__ wimg, long long* dbg) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  float* w_ring = reinterpret_cast<float*>(smem_raw);                               // [W_STAGES][2 chunks][hi, lo][4096]
  float* s_b1 = w_ring + (size_t)FT_W_STAGES * 4 * FZ_CHUNK_FLOATS;                 // [128]
  float* s_b2 = s_b1 + FC;                                                          // [128]
  float* s_w3 = s_b2 + FC;                                                          // [3][128]
  float4* s_part = reinterpret_cast<float4*>(s_w3 + 3 * FC);                        // [2][4 slices][128] partial layer-3 sums
  SPair* s_pair = reinterpret_cast<SPair*>(s_part + 2 * 4 * FT_TILE);               // [kp/2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_pair + FZ_MAX_K / 2);
  uint64_t* o_full = bars;                       // [R_STAGES]  16 warp arrivals + the weight streamer's
  uint64_t* o_empty = o_full + FT_R_STAGES;      // [R_STAGES]  tcgen05.commit
  uint64_t* w_full = o_empty + FT_R_STAGES;      // [W_STAGES]  expect_tx + bulk copies
  uint64_t* w_empty = w_full + FT_W_STAGES;      // [W_STAGES]  tcgen05.commit
  uint64_t* acc_full = w_empty + FT_W_STAGES;    // [2]         layer 1, layer 2
  uint64_t* part_full = acc_full + 2;            // [1]         16 warp arrivals: the partial layer-3 sums of a tile are in s_part
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(part_full + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int64_t ntiles = (n + FT_TILE - 1) / FT_TILE;
  if ((int64_t)blockIdx.x >= ntiles) return;
  const int my_tiles = (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const int n1 = fs.n1, nxg = (fs.kp + FT_GK - 1) / FT_GK, ng = nxg + 2;        // W1 chunks; X groups, groups per tile

  if (tid == 0) {
    for (int i = 0; i < FT_R_STAGES; ++i) { tc::mbar_init(o_full + i, FT_CWARPS + 1); tc::mbar_init(o_empty + i, 1); }
    for (int i = 0; i < FT_W_STAGES; ++i) { tc::mbar_init(w_full + i, 1); tc::mbar_init(w_empty + i, 1); }
    tc::mbar_init(acc_full + 0, 1);
    tc::mbar_init(acc_full + 1, 1);
    tc::mbar_init(part_full, FT_CWARPS);
    tc::fence_barrier_init();
  }
  if (warp == FT_CWARPS) tc::tmem_alloc(tmem_slot, 512);
  for (int e = tid; e < FC; e += FT_THREADS) { s_b1[e] = __ldg(f.b1 + e); s_b2[e] = __ldg(f.b2 + e); }
  for (int e = tid; e < 3 * FC; e += FT_THREADS) s_w3[e] = __ldg(f.w3 + e);
  for (int e = tid; e < fs.kp / 2; e += FT_THREADS) {
    const PairRec r = fs.pair[e];
    SPair p;
    p.meta = (int)r.kind | ((int)r.sa << 8) | ((int)r.sb << 16);
    p.scale = (float)(1 << r.fexp);
    p.ma = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.ma) : 1.f;
    p.mb = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.mb) : 1.f;
    s_pair[e] = p;
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < FT_CWARPS) {
    // =============================================================== compute warps: thread = (sample row, slice t)
    const int team = warp >> 2, q = warp & 3, r_in = q * 32 + lane;
    const uint32_t t_lane = ((uint32_t)(q * 32)) << 16;
    const uint32_t t_acc1 = tmem + t_lane + FT_ACC1 + (uint32_t)(team * 16);    // + 64 G: this thread's columns of h1 group G
    const uint32_t t_acc2 = tmem + t_lane + FT_ACC2 + (uint32_t)(team * 32);
    const uint32_t t_ring = tmem + t_lane + FT_RING;
    const float b3[3] = {__ldg(f.b3), __ldg(f.b3 + 1), __ldg(f.b3 + 2)};
    const bool save = c.save_for_backward != 0;
    const int D = fs.D;
    const bool tl = warp == 0 && lane == 0;
    auto tile_row = [&](int it) { return ((int64_t)blockIdx.x + (int64_t)it * gridDim.x) * FT_TILE + r_in; };
    // sources of one row: [feat(D) | view(3) | pos(3)]
    auto load_src = [&](int64_t row, float* dst) {
      const bool live = row < n;
      if constexpr (SS::is_static) {
#pragma unroll
        for (int k = 0; k < SS::D + 6; ++k) dst[k] = 0.f;
        if (live) {
          const float* fp = ws.a_feat + row * SS::D;
#pragma unroll
          for (int k = 0; k < SS::D; ++k) dst[k] = fp[k];
          float v[3], p[3];
          shade_geometry(f, c, ws, row, v, p);
#pragma unroll
          for (int k = 0; k < 3; ++k) { dst[SS::D + k] = v[k]; dst[SS::D + 3 + k] = p[k]; }
        }
      } else {
        if (live) {
          const float* fp = ws.a_feat + row * D;
          for (int k = 0; k < D; ++k) dst[k] = fp[k];
          float v[3], p[3];
          shade_geometry(f, c, ws, row, v, p);
#pragma unroll
          for (int k = 0; k < 3; ++k) { dst[D + k] = v[k]; dst[D + 3 + k] = p[k]; }
        } else {
          for (int k = 0; k < D + 6; ++k) dst[k] = 0.f;
        }
      }
    };
    float src[FZ_MAX_SRC];
    float da_s = 0.f, da_c = 1.f;                      // running (sin, cos) of the current source at the current octave

    // stage of ring use `g`: wait until the MMAs that read its previous content have completed
    auto ring_acquire = [&](uint32_t g) -> uint32_t {
      const uint32_t stg = g % FT_R_STAGES;
      tc::mbar_wait(o_empty + stg, ((g / FT_R_STAGES) & 1) ^ 1);
      tc::fence_after_sync();
      return t_ring + stg * (2 * FT_GK);
    };
    auto ring_publish = [&](uint32_t g) {
      tc::tmem_wait_st();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(o_full + (g % FT_R_STAGES));
    };
    auto put4 = [&](uint32_t base, int u, const float v[4]) {      // unit u (4 columns) of a group: hi | lo
      float4 h, l;
      split4(v, h, l);
      tc::tmem_st4(base + (uint32_t)(u * 4), h.x, h.y, h.z, h.w);
      tc::tmem_st4(base + FT_GK + (uint32_t)(u * 4), l.x, l.y, l.z, l.w);
    };

    // this thread's 16 columns (units 4 t .. 4 t + 3) of X group G -> operand ring (+ X saved for backward).  The first trig
    // pair of the part takes a fresh sincos: the octave below it is another thread's.
    auto x_part = [&](auto gc, auto tmc, int gd, int64_t row, uint32_t g) {
      constexpr int G = decltype(gc)::value, TM = decltype(tmc)::value;
      const int grp = SS::is_static ? G : gd, tm = SS::is_static ? TM : team;
      const bool live = row < n;
      const uint32_t base = ring_acquire(g);
      auto emit = [&](int u, int col, const float v[4]) {
        put4(base, u, v);
        if (save && live) *reinterpret_cast<float4*>(ws.a_x + row * ws.ldx + col) = make_float4(v[0], v[1], v[2], v[3]);
      };
      if constexpr (SS::is_static) {
#pragma unroll
        for (int uu = 0; uu < 4; ++uu) {
          constexpr int C0 = G * FT_GK + TM * 16;        // first column of the part
          if (C0 + uu * 4 < SS::kp) {
            float v[4];
#pragma unroll
            for (int p2 = 0; p2 < 2; ++p2) {
              const int P = (C0 + uu * 4) / 2 + p2;
              const PairInfo pi = SS::pair(P);
              float x0 = 0.f, x1 = 0.f;
              if (pi.kind == PK_TRIG) {
                const SPair pr = s_pair[P];
                if ((pi.fexp & 3) == 0 || P == C0 / 2) fast_sincos(src[pi.sa] * (float)(1 << pi.fexp), &da_s, &da_c);
                else { const float s2 = 2.f * da_s * da_c; da_c = fmaf(-2.f * da_s, da_s, 1.f); da_s = s2; }
                x0 = da_s * pr.ma;
                x1 = da_c * pr.mb;
              } else if (pi.kind == PK_COPY2) {
                x0 = src[pi.sa];
                x1 = src[pi.sb];
              } else if (pi.kind == PK_COPY1) {
                x0 = src[pi.sa];
              }
              v[2 * p2] = x0;
              v[2 * p2 + 1] = x1;
            }
            emit(4 * TM + uu, C0 + uu * 4, v);
          }
        }
      } else {
        for (int uu = 0; uu < 4; ++uu) {
          const int col = grp * FT_GK + tm * 16 + uu * 4;
          if (col < fs.kp) {                             // warp-uniform
            float v[4];
#pragma unroll
            for (int p2 = 0; p2 < 2; ++p2) {
              const SPair pr = s_pair[col / 2 + p2];
              const int kind = pr.meta & 0xff, sa = (pr.meta >> 8) & 0xff, sb = (pr.meta >> 16) & 0xff;
              float x0 = 0.f, x1 = 0.f;
              if (kind == PK_TRIG) {
                if (((__float_as_int(pr.scale) >> 23) & 3) == 3 || (uu == 0 && p2 == 0)) fast_sincos(src[sa] * pr.scale, &da_s, &da_c);
                else { const float s2 = 2.f * da_s * da_c; da_c = fmaf(-2.f * da_s, da_s, 1.f); da_s = s2; }
                x0 = da_s * pr.ma;
                x1 = da_c * pr.mb;
              } else if (kind == PK_COPY2) {
                x0 = src[sa];
                x1 = src[sb];
              } else if (kind == PK_COPY1) {
                x0 = src[sa];
              }
              v[2 * p2] = x0;
              v[2 * p2 + 1] = x1;
            }
            emit(4 * tm + uu, col, v);
          }
        }
      }
      ring_publish(g);
    };
    // every X group of one tile (this thread's slice of each); g0 = ring use of group 0
    auto build_team = [&](auto tmc, int64_t row, uint32_t g0) {
      if constexpr (SS::is_static) {
        constexpr int NXG = (SS::kp + FT_GK - 1) / FT_GK;
        if constexpr (0 < NXG) x_part(IC<0>{}, tmc, 0, row, g0 + 0);
        if constexpr (1 < NXG) x_part(IC<1>{}, tmc, 1, row, g0 + 1);
        if constexpr (2 < NXG) x_part(IC<2>{}, tmc, 2, row, g0 + 2);
      }
    };
    auto build_x = [&](int64_t row, uint32_t g0) {
      if constexpr (SS::is_static) {
        switch (team) {
          case 0: build_team(IC<0>{}, row, g0); break;
          case 1: build_team(IC<1>{}, row, g0); break;
          case 2: build_team(IC<2>{}, row, g0); break;
          default: build_team(IC<3>{}, row, g0); break;
        }
      } else {
        for (int k = 0; k < nxg; ++k) x_part(IC<0>{}, IC<0>{}, k, row, g0 + (uint32_t)k);
      }
    };

    load_src(tile_row(0), src);
    build_x(tile_row(0), 0u);
    for (int it = 0; it < my_tiles; ++it) {
      const int64_t row = tile_row(it), row_next = tile_row(it + 1);
      const bool live = row < n, more = it + 1 < my_tiles;
      const uint32_t g0 = (uint32_t)it * (uint32_t)ng;
      if (tl) FZ_T(0, it * 8 + 0);
      if (more) load_src(row_next, src);               // in flight during the layer-1 epilogue below
      // ---- layer-1 accumulator -> the two layer-2 operand groups (bias + ReLU + split), this thread's 16 columns of each
      tc::mbar_wait(acc_full + 0, it & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(0, it * 8 + 1);
      {
        float acc[32];
#pragma unroll
        for (int j = 0; j < 4; ++j) tc::tmem_ld8(t_acc1 + (uint32_t)((j >> 1) * FT_GK + (j & 1) * 8), acc + 8 * j);
        tc::tmem_wait_ld();                            // every read of acc1 precedes this warp's first X publish below
#pragma unroll
        for (int G = 0; G < 2; ++G) {
          const int c0 = G * FT_GK + team * 16;
#pragma unroll
          for (int i = 0; i < 16; ++i) acc[16 * G + i] = fmaxf(acc[16 * G + i] + s_b1[c0 + i], 0.f);
          if (save && live) {
            st_global_v8(ws.a_h1 + row * FC + c0, acc + 16 * G);
            st_global_v8(ws.a_h1 + row * FC + c0 + 8, acc + 16 * G + 8);
          }
          const uint32_t g = g0 + (uint32_t)(nxg + G);
          const uint32_t base = ring_acquire(g);
#pragma unroll
          for (int uu = 0; uu < 4; ++uu) put4(base, 4 * team + uu, acc + 16 * G + 4 * uu);
          ring_publish(g);
        }
      }
      if (tl) FZ_T(0, it * 8 + 2);
      // ---- the next tile's input rows
      if (more) build_x(row_next, g0 + (uint32_t)ng);
      if (tl) FZ_T(0, it * 8 + 3);
      // ---- layer-2 accumulator slice: bias + ReLU, partial 128 -> 3 product
      tc::mbar_wait(acc_full + 1, it & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(0, it * 8 + 4);
      float o3[3] = {0.f, 0.f, 0.f};
      {
        float acc[32];
        tc::tmem_ld32(t_acc2, acc);
        tc::tmem_wait_ld();
        tc::fence_before_sync();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = fmaxf(acc[i] + s_b2[team * 32 + i], 0.f);
        if (save && live) {
#pragma unroll
          for (int u = 0; u < 4; ++u) st_global_v8(ws.a_h2 + row * FC + team * 32 + u * 8, acc + 8 * u);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const float4* wr = reinterpret_cast<const float4*>(s_w3 + k * FC + team * 32);
          float sacc = 0.f;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            float4 w4 = wr[u];
            sacc = fmaf(acc[4 * u], w4.x, sacc); sacc = fmaf(acc[4 * u + 1], w4.y, sacc);
            sacc = fmaf(acc[4 * u + 2], w4.z, sacc); sacc = fmaf(acc[4 * u + 3], w4.w, sacc);
          }
          o3[k] = sacc;
        }
      }
      float4* part = s_part + (size_t)(it & 1) * 4 * FT_TILE;
      part[team * FT_TILE + r_in] = make_float4(o3[0], o3[1], o3[2], 0.f);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(part_full);       // (release: the partial sums above are visible to the waiter)
      if (team == (it & 3)) tc::mbar_wait(part_full, it & 1);
      if (team == (it & 3) && live) {                  // the finishing slice rotates; nobody else waits
        const float4 p0 = part[r_in], p1 = part[FT_TILE + r_in], p2 = part[2 * FT_TILE + r_in], p3 = part[3 * FT_TILE + r_in];
        const float o0 = b3[0] + ((p0.x + p1.x) + (p2.x + p3.x));
        const float o1 = b3[1] + ((p0.y + p1.y) + (p2.y + p3.y));
        const float o2 = b3[2] + ((p0.z + p1.z) + (p2.z + p3.z));
        ws.a_rgb[row * 3 + 0] = 1.f / (1.f + expf(-o0));
        ws.a_rgb[row * 3 + 1] = 1.f / (1.f + expf(-o1));
        ws.a_rgb[row * 3 + 2] = 1.f / (1.f + expf(-o2));
      }
      if (tl) FZ_T(0, it * 8 + 5);
    }
  } else if (warp == FT_CWARPS) {
    // =============================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_tf32(128, 128);
      const uint32_t w_lo0 = tc::desc_lo(tc::smem_u32(w_ring), 128);
      const uint32_t d_hi = tc::desc_hi(1024);
      uint32_t g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        for (int k = 0; k < ng; ++k, ++g) {
          const uint32_t as = g % FT_R_STAGES, wsg = g % FT_W_STAGES;
          const bool l1 = k < nxg;
          const int ksteps = l1 ? min(FT_GK, fs.kp - k * FT_GK) / 8 : FT_GK / 8;
          FZ_T(1, g * 2 + 0);
          tc::mbar_wait(o_full + as, (g / FT_R_STAGES) & 1);       // operands AND weights (the streamer arrives for them)
          FZ_T(1, g * 2 + 1);
          tc::fence_after_sync();
          const uint32_t w_lo = w_lo0 + wsg * (uint32_t)(4 * FZ_CHUNK_FLOATS * 4 / 16);
          const uint32_t a_hi = tmem + FT_RING + as * (2 * FT_GK), a_lo = a_hi + FT_GK;
          const uint32_t d = tmem + (l1 ? FT_ACC1 : FT_ACC2);
          const uint32_t fresh = (k == 0 || k == nxg) ? 0u : 1u;
          constexpr uint32_t LO = FZ_CHUNK_FLOATS * 4 / 16;      // hi -> lo half of a weight chunk, in 16-byte units
#pragma unroll
          for (int s = 0; s < FT_GK / 8; ++s) {
            if (s < ksteps) {
              const uint32_t wb = w_lo + (uint32_t)(s >> 2) * (2 * LO) + (uint32_t)(s & 3) * 16;   // chunk s / 4 of the stage
              tc::mma_tf32_ts_lo(d, a_lo + s * 8, wb, d_hi, idesc, s == 0 ? fresh : 1u);
              tc::mma_tf32_ts_lo(d, a_hi + s * 8, wb + LO, d_hi, idesc, 1u);
              tc::mma_tf32_ts_lo(d, a_hi + s * 8, wb, d_hi, idesc, 1u);
            }
          }
          tc::mma_commit(o_empty + as);
          tc::mma_commit(w_empty + wsg);
          if (k == nxg - 1) tc::mma_commit(acc_full + 0);
          if (k == ng - 1) tc::mma_commit(acc_full + 1);
        }
      }
    }
  } else {
    // =============================================================== weight streamer (TMA): group g's chunks into stage g % 3,
    // up to three groups ahead; arrives on the group's operand barrier once its weights have landed
    if (lane == 0) {
      const uint32_t total = (uint32_t)my_tiles * (uint32_t)ng;
      uint32_t issued = 0;
      for (uint32_t g = 0; g < total; ++g) {
        while (issued < total && issued < g + FT_W_STAGES) {
          const uint32_t wsg = issued % FT_W_STAGES;
          const int k = (int)(issued % (uint32_t)ng);
          const int c0 = k < nxg ? 2 * k : n1 + 2 * (k - nxg);                   // first weight chunk of the group
          const int nc = k < nxg ? min(2, n1 - 2 * k) : 2;
          tc::mbar_wait(w_empty + wsg, ((issued / FT_W_STAGES) & 1) ^ 1);
          mbar_expect_tx(w_full + wsg, (uint32_t)(nc * FZ_CHUNK_BYTES));
          for (int j = 0; j < nc; ++j)
            bulk_g2s(w_ring + ((size_t)wsg * 2 + j) * 2 * FZ_CHUNK_FLOATS, wimg + (size_t)(c0 + j) * 2 * FZ_CHUNK_FLOATS,
                     (uint32_t)FZ_CHUNK_BYTES, w_full + wsg);
          ++issued;
        }
        tc::mbar_wait(w_full + (g % FT_W_STAGES), (g / FT_W_STAGES) & 1);
        if (g >= FT_R_STAGES) tc::mbar_wait(o_full + (g % FT_R_STAGES), ((g - FT_R_STAGES) / FT_R_STAGES) & 1);   // its previous phase is over
        tc::mbar_arrive(o_full + (g % FT_R_STAGES));
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == FT_CWARPS) tc::tmem_dealloc(tmem, 512);
}

static size_t fwd_ts_smem() {
  return (size_t)FT_W_STAGES * 4 * FZ_CHUNK_FLOATS * 4 + (size_t)(2 * FC + 3 * FC) * 4 + (size_t)2 * 4 * FT_TILE * sizeof(float4) +
         (size_t)(FZ_MAX_K / 2) * sizeof(SPair) + 32 * 8 + 16;
}

#endif  // FTRF_TEST_BUILD

// ================================================================================================
// Backward input-gradient chain as ONE persistent kernel:
//   go = g_rgb * sigmoid'  ->  g_h2 = relu'(h2) * (go W3)  ->  g_h1 = relu'(h1) * (g_h2 W2)  ->  g_X = g_h1 W1  ->  g_feat
// (chain rule through the encoding) per 128-sample tile.  g_h2 is formed in registers from the saved h2 (it is never
// written: the output-layer backward used to be a kernel of its own that wrote it and this one read it back, 105 MB
// per step); go (3 floats per sample) is left in the workspace for the weight-gradient kernel, which rebuilds g_h2
// the same way and also owns dW3 / db3 / db2; g_h1 is written out for that kernel; g_X never leaves the SM.
// Roles: two groups of 8 compute warps work on alternate tiles (one group's loads / epilogue overlap
// the other group's products), thread = (sample row, half of the columns); warp 16 issues the MMAs tile
// by tile; warp 17 streams the weights by TMA: per tile 8 sub-chunks (K = 16) of W2^T [128 x 16] then 8
// of W1^T [160 x 16] (internal column order).  Each group owns a 2-stage operand ring, so a group never
// runs more than one barrier phase ahead of the MMA warp.  TMEM: acc1 columns 0-127 (shared: a tile's
// acc1 is drained before its second product ends), acc2 of group g at 128 + 160 g.
constexpr int BZ_GROUPS = 2, BZ_A_STAGES = 2, BZ_W_STAGES = 2;
constexpr int BZ_W1_ROWS = 160;                                  // kp (152) padded to a multiple of 16
constexpr int BZ_SUB_K = 16;                                     // k extent of a streamed weight sub-chunk
constexpr int BZ_W2_SUB_FLOATS = 2 * 128 * BZ_SUB_K;             // hi + lo
constexpr int BZ_W1_SUB_FLOATS = 2 * BZ_W1_ROWS * BZ_SUB_K;
constexpr int BZ_WSTAGE_FLOATS = 2 * BZ_W1_SUB_FLOATS;            // a stage = the two sub-chunks (K = 32) one operand chunk needs:
                                                                 // one barrier wait and one commit of the issuing thread per chunk
                                                                 // instead of two (each wait costs it ~225 clk, scripts/issue_probe.py)
constexpr int BZ_THREADS = (BZ_GROUPS * 8 + 2) * 32;

__global__ void __launch_bounds__(256)
prep_weights_bwd_kernel(const FtrfField f, const FusedSpec fs, float* __restrict__ img, int in_dim) {
  // sub-chunks 0-7: B[n = i][k = o] = W2[o][i], o in [16 sc, 16 sc + 16);  8-15: B[n = kint][k = o] = W1[o][orig(kint)]
  const int per2 = 128 * 4, per1 = BZ_W1_ROWS * 4;
  const int total = 8 * per2 + 8 * per1;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const bool second = e >= 8 * per2;
    const int ee = second ? e - 8 * per2 : e, per = second ? per1 : per2;
    const int sc = ee / per, r = ee - sc * per, ku = r & 3, nrow = r >> 2;
    float hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = sc * BZ_SUB_K + ku * 4 + j;
      float v;
      if (!second) v = __ldg(f.w2 + (size_t)o * FC + nrow);
      else {
        const int oc = nrow < fs.kp ? fs.orig[nrow] : -1;
        v = oc >= 0 ? __ldg(f.w1 + (size_t)o * in_dim + oc) : 0.f;
      }
      tc::split_tf32(v, hi[j], lo[j]);
    }
    float* base = img + (second ? (size_t)8 * BZ_W2_SUB_FLOATS + (size_t)sc * BZ_W1_SUB_FLOATS : (size_t)sc * BZ_W2_SUB_FLOATS);
    const int half = (second ? BZ_W1_SUB_FLOATS : BZ_W2_SUB_FLOATS) / 2;
    const int off = tc::canon_off(nrow, ku, 4);
    *reinterpret_cast<float4*>(base + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(base + half + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
}

__device__ __forceinline__ void ld_global_v8(const float* p, float* v) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}

// which 32-column chunks of g_X hold a column that column-half H turns into g_feat (channels split 14 / rest)
constexpr int BZ_SPLIT = 14;
template <class SS>
__host__ __device__ constexpr bool bz_owned(int sa, int H) { return sa < SS::D && (sa < BZ_SPLIT ? 0 : 1) == H; }
template <class SS>
__host__ __device__ constexpr uint32_t bz_chunk_mask(int H) {
  uint32_t m = 0;
  for (int P = 0; 2 * P < SS::kp; ++P) {
    const PairInfo pi = SS::pair(P);
    bool use = false;
    if (pi.kind == PK_COPY2) use = bz_owned<SS>(pi.sa, H) || bz_owned<SS>(pi.sb, H);
    else if (pi.kind == PK_COPY1 || pi.kind == PK_TRIG) use = bz_owned<SS>(pi.sa, H);
    if (use) m |= 1u << (P / 16);
  }
  return m;
}

template <class SS, int H>
__device__ __forceinline__ void bz_epilogue(uint32_t t_acc2, const SPair* s_pair, const float* __restrict__ feat_row,
                                            float* __restrict__ gfeat_row, bool live) {
  constexpr int C0 = H == 0 ? 0 : BZ_SPLIT, C1 = H == 0 ? (SS::D < BZ_SPLIT ? SS::D : BZ_SPLIT) : SS::D;
  constexpr int NC = C1 - C0 > 0 ? C1 - C0 : 1;
  constexpr uint32_t CM = bz_chunk_mask<SS>(H);
  float gf[NC], ft[NC];
#pragma unroll
  for (int k = 0; k < NC; ++k) { gf[k] = 0.f; ft[k] = (live && C0 + k < C1) ? __ldg(feat_row + C0 + k) : 0.f; }
  float da_s = 0.f, da_c = 1.f;
#pragma unroll
  for (int cb = 0; cb < SS::n1; ++cb) {
    if (!((CM >> cb) & 1u)) continue;
    float gx[32];
    tc::tmem_ld32(t_acc2 + (uint32_t)(cb * 32), gx);
    tc::tmem_wait_ld();
#pragma unroll
    for (int pp = 0; pp < 16; ++pp) {
      const int P = cb * 16 + pp;
      if (2 * P >= SS::kp) continue;
      const PairInfo pi = SS::pair(P);
      if (pi.kind == PK_COPY2 || pi.kind == PK_COPY1) {
        if (bz_owned<SS>(pi.sa, H)) gf[pi.sa - C0] += gx[2 * pp];
        if (pi.kind == PK_COPY2 && bz_owned<SS>(pi.sb, H)) gf[pi.sb - C0] += gx[2 * pp + 1];
      } else if (pi.kind == PK_TRIG && bz_owned<SS>(pi.sa, H)) {
        const SPair pr = s_pair[P];
        if ((pi.fexp & 3) == 0) fast_sincos(ft[pi.sa - C0] * (float)(1 << pi.fexp), &da_s, &da_c);
        else { const float s2 = 2.f * da_s * da_c; da_c = fmaf(-2.f * da_s, da_s, 1.f); da_s = s2; }
        // d/dfeat of (sin(a) ms, cos(a) mc), a = feat 2^f
        gf[pi.sa - C0] += (float)(1 << pi.fexp) * (da_c * pr.ma * gx[2 * pp] - da_s * pr.mb * gx[2 * pp + 1]);
      }
    }
  }
  tc::fence_before_sync();
  if (live) {
#pragma unroll
    for (int k = 0; k < NC; ++k)
      if (C0 + k < C1) gfeat_row[C0 + k] = gf[k];
  }
}

template <class SS>
__global__ void __launch_bounds__(BZ_THREADS, 1)
shade_fused_dgrad_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const FusedSpec fs, const float* __restrict__ wimg, long long* dbg,
                         const int masks) {     // masks: the forward left the ReLU sign bits of h1 / h2 in ws.a_gx (32 B per sample)
  static_assert(SS::is_static && SS::kp <= BZ_W1_ROWS, "dgrad kernel is built for the static specs only");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  float* a_ring = reinterpret_cast<float*>(smem_raw);                               // [group][A_STAGES][hi, lo][4096]
  float* w_ring = a_ring + (size_t)BZ_GROUPS * BZ_A_STAGES * 2 * FZ_CHUNK_FLOATS;   // [W_STAGES][BZ_WSTAGE_FLOATS]
  SPair* s_pair = reinterpret_cast<SPair*>(w_ring + (size_t)BZ_W_STAGES * BZ_WSTAGE_FLOATS);
  float* s_w3 = reinterpret_cast<float*>(s_pair + FZ_MAX_K / 2);                    // [3][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_w3 + 3 * FC);
  uint64_t* a_full = bars;                                   // [group][A_STAGES]  4 warp arrivals (the column half that owns the chunk)
  uint64_t* a_empty = a_full + BZ_GROUPS * BZ_A_STAGES;
  uint64_t* w_full = a_empty + BZ_GROUPS * BZ_A_STAGES;      // [W_STAGES]
  uint64_t* w_empty = w_full + BZ_W_STAGES;
  uint64_t* acc_full = w_empty + BZ_W_STAGES;                // [group][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + BZ_GROUPS * 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int64_t ntiles = (n + 127) / 128;
  if ((int64_t)blockIdx.x >= ntiles) return;
  const int my_tiles = (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const int perm = ws.counters[WS_ACT_LAYOUT];          // layout the forward saved in: g_h1 goes out the same way

  if (tid == 0) {
    for (int i = 0; i < BZ_GROUPS * BZ_A_STAGES; ++i) { tc::mbar_init(a_full + i, 4); tc::mbar_init(a_empty + i, 1); }
    for (int i = 0; i < BZ_W_STAGES; ++i) { tc::mbar_init(w_full + i, 1); tc::mbar_init(w_empty + i, 1); }
    for (int i = 0; i < BZ_GROUPS * 2; ++i) tc::mbar_init(acc_full + i, 1);
    tc::fence_barrier_init();
  }
  if (warp == 16) tc::tmem_alloc(tmem_slot, 512);
  for (int e = tid; e < 3 * FC; e += BZ_THREADS) s_w3[e] = __ldg(f.w3 + e);
  for (int e = tid; e < fs.kp / 2; e += BZ_THREADS) {
    const PairRec r = fs.pair[e];
    SPair p;
    p.meta = (int)r.kind | ((int)r.sa << 8) | ((int)r.sb << 16);
    p.scale = (float)(1 << r.fexp);
    p.ma = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.ma) : 1.f;
    p.mb = r.kind == PK_TRIG ? enc_mask_at(f, r.msrc, r.mb) : 1.f;
    s_pair[e] = p;
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 16) {
    // =============================================================== compute warps
    const int g = warp >> 3, half = (warp >> 2) & 1, q = warp & 3, r_in = q * 32 + lane;
    const uint32_t t_lane = ((uint32_t)(q * 32)) << 16;
    float* my_ring = a_ring + (size_t)g * BZ_A_STAGES * 2 * FZ_CHUNK_FLOATS;
    uint64_t* my_full = a_full + g * BZ_A_STAGES;
    uint64_t* my_empty = a_empty + g * BZ_A_STAGES;
    auto put_chunk = [&](uint32_t ac, const float* vals) {          // 32 values of this row -> operand ring stage
      const int stg = ac % BZ_A_STAGES;
      tc::mbar_wait(my_empty + stg, ((ac / BZ_A_STAGES) & 1) ^ 1);
      float* Ahi = my_ring + (size_t)stg * 2 * FZ_CHUNK_FLOATS;
      float* Alo = Ahi + FZ_CHUNK_FLOATS;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float4 h, l;
        split4(vals + 4 * u, h, l);
        const int off = tc::canon_off(r_in, u, 8);
        *reinterpret_cast<float4*>(Ahi + off) = h;
        *reinterpret_cast<float4*>(Alo + off) = l;
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(my_full + stg);
    };
    const bool tl = lane == 0 && q == 0 && half == 0;
    for (int li = 0; g + 2 * li < my_tiles; ++li) {
      const int it = g + 2 * li;
      const int64_t row = ((int64_t)blockIdx.x + (int64_t)it * gridDim.x) * 128 + r_in;
      const bool live = row < n;
      const uint32_t ac0 = (uint32_t)li * 8;
      if (tl) FZ_T(g, li * 8 + 0);
      // ---- g_h2 chunks (half, half + 2) -> operand ring (A of g_h2 W2); all loads in flight at once.
      // g_h2 = relu'(h2) * (go W3) with go = dL/d rgb * rgb (1 - rgb) (sigmoid backward, models/mlp.py:66-67)
      uint32_t hm[2] = {0u, 0u};                            // relu'(h1) of my two chunks as bit masks
      {
        float v[64];
        if (live) {
          if (masks) {                                      // sign bits instead of the values: v > 0 <=> bit set
            const uint32_t* mk = reinterpret_cast<const uint32_t*>(ws.a_gx) + row;      // word-major [8][cap_app]
            hm[0] = __ldg(mk + (size_t)half * c.cap_app);
            hm[1] = __ldg(mk + (size_t)(half + 2) * c.cap_app);
            const uint32_t b0 = __ldg(mk + (size_t)(4 + half) * c.cap_app), b1 = __ldg(mk + (size_t)(6 + half) * c.cap_app);
#pragma unroll
            for (int i = 0; i < 32; ++i) { v[i] = (float)((b0 >> i) & 1u); v[32 + i] = (float)((b1 >> i) & 1u); }
          } else {
#pragma unroll
            for (int jj = 0; jj < 2; ++jj)
#pragma unroll
              for (int u = 0; u < 4; ++u) ld_global_v8(ws.a_h2 + row * FC + (jj * 2 + half) * 32 + u * 8, v + jj * 32 + 8 * u);
          }
          float go[3];
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            const float r = ws.a_rgb[row * 3 + ch];
            go[ch] = ws.a_grgb[row * 3 + ch] * r * (1.f - r);
          }
          if (half == 0) *reinterpret_cast<float4*>(ws.a_gh2 + row * 4) = make_float4(go[0], go[1], go[2], 0.f);   // go4, for the wgrad kernel
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            const float* w3 = s_w3 + (jj * 2 + half) * 32;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float g = go[0] * w3[i] + go[1] * w3[FC + i] + go[2] * w3[2 * FC + i];
              v[jj * 32 + i] = v[jj * 32 + i] > 0.f ? g : 0.f;
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 64; ++i) v[i] = 0.f;
        }
        put_chunk(ac0 + half, v);
        put_chunk(ac0 + 2 + half, v + 32);
      }
      if (tl) FZ_T(g, li * 8 + 1);
      // ---- relu'(h1) of my two chunks as bit masks, fetched while the first product runs (unless the forward left them)
      if (live && !masks) {
        float h1[64];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
          for (int u = 0; u < 4; ++u) ld_global_v8(ws.a_h1 + row * FC + (jj * 2 + half) * 32 + u * 8, h1 + jj * 32 + 8 * u);
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
          for (int i = 0; i < 32; ++i) hm[jj] |= (h1[jj * 32 + i] > 0.f ? 1u : 0u) << i;
      }
      // ---- g_h1 = relu'(h1) * acc1: saved for the weight-gradient kernel and fed to the second product
      tc::mbar_wait(acc_full + g * 2 + 0, li & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(g, li * 8 + 2);
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = jj * 2 + half;
        float acc[32];
        tc::tmem_ld32(tmem + t_lane + (uint32_t)(j * 32), acc);
        tc::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = ((hm[jj] >> i) & 1u) ? acc[i] : 0.f;
        if (live) {
#pragma unroll
          for (int u = 0; u < 4; ++u) st_global_v8(ws.a_gh1 + (perm ? act16_off(row, j * 32 + u * 8, FC) : (size_t)row * FC + j * 32 + u * 8), acc + 8 * u);
        }
        put_chunk(ac0 + 4 + j, acc);
      }
      if (tl) FZ_T(g, li * 8 + 3);
      // ---- g_X (acc2) -> g_feat: copy column + chain rule through sin / cos, channels split between the halves
      tc::mbar_wait(acc_full + g * 2 + 1, li & 1);
      tc::fence_after_sync();
      if (tl) FZ_T(g, li * 8 + 4);
      const uint32_t t_acc2 = tmem + t_lane + (uint32_t)(128 + BZ_W1_ROWS * g);
      if (half == 0) bz_epilogue<SS, 0>(t_acc2, s_pair, ws.a_feat + row * SS::D, ws.a_gfeat + row * SS::D, live);
      else bz_epilogue<SS, 1>(t_acc2, s_pair, ws.a_feat + row * SS::D, ws.a_gfeat + row * SS::D, live);
      if (tl) FZ_T(g, li * 8 + 5);
    }
  } else if (warp == 16) {
    // =============================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc2 = tc::idesc_tf32(128, 128), idesc1 = tc::idesc_tf32(128, BZ_W1_ROWS);
      const uint32_t a_lo0 = tc::desc_lo(tc::smem_u32(a_ring), 128), w_lo0 = tc::desc_lo(tc::smem_u32(w_ring), 128);
      const uint32_t a_hi = tc::desc_hi(1024), w_hi = tc::desc_hi(512);
      constexpr uint32_t ALO = FZ_CHUNK_FLOATS * 4 / 16;
      uint32_t wc = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int g = it & 1, li = it >> 1;
        for (int ch = 0; ch < 8; ++ch) {
          const uint32_t ac = (uint32_t)li * 8 + ch;
          const int as = ac % BZ_A_STAGES;
          const bool first = ch < 4;
          const uint32_t ah = a_lo0 + (uint32_t)(g * BZ_A_STAGES + as) * (2 * ALO);
          const uint32_t wlo = (uint32_t)((first ? BZ_W2_SUB_FLOATS : BZ_W1_SUB_FLOATS) / 2 * 4 / 16);
          const uint32_t d = tmem + (first ? 0u : (uint32_t)(128 + BZ_W1_ROWS * g)), idesc = first ? idesc2 : idesc1;
          FZ_T(2, (it * 8 + ch) * 2 + 0);
          tc::mbar_wait(a_full + g * BZ_A_STAGES + as, (ac / BZ_A_STAGES) & 1);
          FZ_T(2, (it * 8 + ch) * 2 + 1);
          const int wsg = wc % BZ_W_STAGES;
          tc::mbar_wait(w_full + wsg, (wc / BZ_W_STAGES) & 1);
          tc::fence_after_sync();
#pragma unroll
          for (int sub = 0; sub < 2; ++sub) {
            const uint32_t wh = w_lo0 + (uint32_t)wsg * (BZ_WSTAGE_FLOATS * 4 / 16) + (uint32_t)sub * (2 * wlo);
#pragma unroll
            for (int s = 0; s < 2; ++s) {
              const uint32_t ao = (uint32_t)(sub * 2 + s) * 16, wo = (uint32_t)s * 16;
              const uint32_t fresh = ((ch & 3) == 0 && sub == 0 && s == 0) ? 0u : 1u;
              tc::mma_tf32_ss_lo2(d, ah + ALO + ao, a_hi, wh + wo, w_hi, idesc, fresh);
              tc::mma_tf32_ss_lo2(d, ah + ao, a_hi, wh + wlo + wo, w_hi, idesc, 1u);
              tc::mma_tf32_ss_lo2(d, ah + ao, a_hi, wh + wo, w_hi, idesc, 1u);
            }
          }
          tc::mma_commit(w_empty + wsg);
          ++wc;
          tc::mma_commit(a_empty + g * BZ_A_STAGES + as);
          if (ch == 3) tc::mma_commit(acc_full + g * 2 + 0);
          if (ch == 7) tc::mma_commit(acc_full + g * 2 + 1);
        }
      }
    }
  } else {
    // =============================================================== weight sub-chunk streamer (TMA)
    if (lane == 0) {
      uint32_t wc = 0;
      for (int it = 0; it < my_tiles; ++it)
        for (int sc = 0; sc < 16; sc += 2, ++wc) {          // sub-chunks sc, sc + 1 are adjacent in the image: one copy
          const int wsg = wc % BZ_W_STAGES;
          const bool first = sc < 8;
          const uint32_t bytes = (uint32_t)(first ? BZ_W2_SUB_FLOATS : BZ_W1_SUB_FLOATS) * 4 * 2;
          const float* src = first ? wimg + (size_t)sc * BZ_W2_SUB_FLOATS
                                   : wimg + (size_t)8 * BZ_W2_SUB_FLOATS + (size_t)(sc - 8) * BZ_W1_SUB_FLOATS;
          tc::mbar_wait(w_empty + wsg, ((wc / BZ_W_STAGES) & 1) ^ 1);
          mbar_expect_tx(w_full + wsg, bytes);
          bulk_g2s(w_ring + (size_t)wsg * BZ_WSTAGE_FLOATS, src, bytes, w_full + wsg);
        }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 16) tc::tmem_dealloc(tmem, 512);
}

size_t fused_bwd_wimg_floats() { return (size_t)8 * BZ_W2_SUB_FLOATS + (size_t)8 * BZ_W1_SUB_FLOATS; }

// g_h2 (from layer3_bwd) -> g_h1 (saved) and g_feat.  *handled = false: configuration not covered.
// Which fused dgrad build covers this field (-1: none, the per-layer GEMMs run instead).
static int dgrad_variant(const FtrfField& f, const FusedSpec& fs) {
  using SFea = StaticSpec<FTRF_SHADE_MLP_FEA, 27, 2, 2, 2>;
  using SPe = StaticSpec<FTRF_SHADE_MLP_PE, 27, 2, 2, 2>;
  using SFew = StaticSpec<FTRF_SHADE_MLP, 18, 8, 2, 2>;      // configs/free_tensorf_human.txt:46-54: 153 inputs
  const bool pe2 = f.app_dim == 27 && f.view_pe == 2;
  if (f.shading == FTRF_SHADE_MLP_FEA && pe2 && f.fea_pe == 2 && fs.kp == SFea::kp) return 0;
  if (f.shading == FTRF_SHADE_MLP_PE && pe2 && f.pos_pe == 2 && fs.kp == SPe::kp) return 1;
  if (f.shading == FTRF_SHADE_MLP && f.app_dim == 18 && f.pos_pe == 8 && f.view_pe == 2 && f.fea_pe == 2 && fs.kp == SFew::kp) return 2;
  return -1;                                     // (MLP 27/2/2/2: kp = 168 > 160 rows of the W1^T chunk; generic path)
}

// The weight images only depend on the parameters, not on anything the step computes, so ftrf_forward_ex builds
// them on a side stream at the very start of the call (beside march / density) instead of in front of the fused
// forward and dgrad kernels, where the two ~7 us latency-bound launches sat on the step's serial chain.
// What a call has pending (images built ahead, side-stream work to join) is recorded in the caller's FtrfCall
// (lib_flags / lib_ctx, see call_state()): the backward may run on another host thread than the forward (autograd
// engine), and a split backward's two halves may run on different threads, without any process-wide state.
// Each host thread only CACHES a side stream and its events here (created on first use, never freed: a context
// recorded in a call must outlive the thread that made it).
struct SideCtx {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, done = nullptr;            // forward: weight images
  cudaEvent_t tail_fork = nullptr, tail_done = nullptr;  // backward: un-permute of dW1
  cudaEvent_t scan_fork = nullptr, scan_done = nullptr;  // backward: reverse scan of the integration stage
  int dev = -1;
};
static thread_local SideCtx* t_side = nullptr;

// The calling thread's side-stream context for the current device (nullptr: none, callers run in line).
static SideCtx* ensure_side() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (t_side) return t_side->dev == dev ? t_side : nullptr;     // (one process per GPU; a thread that switches devices runs in line)
  SideCtx* s = new SideCtx();
  s->dev = dev;
  bool ok = cudaStreamCreateWithFlags(&s->side, cudaStreamNonBlocking) == cudaSuccess;
  cudaEvent_t* evs[6] = {&s->fork, &s->done, &s->tail_fork, &s->tail_done, &s->scan_fork, &s->scan_done};
  for (int i = 0; ok && i < 6; ++i) ok = cudaEventCreateWithFlags(evs[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
    delete s;
    return nullptr;
  }
  t_side = s;
  return s;
}

// A tail that nothing later in the backward reads (the un-permute of the dW1 scratch into g_w1 / g_b1) runs on the
// side stream beside the appearance and density backward kernels: side_tail_begin() orders the side stream after
// everything enqueued on `st` so far and returns it (nullptr: no side stream, run in line); side_tail_end() records
// its completion in the call; side_tail_join() makes `st` wait for it (end of the backward call).
cudaStream_t side_tail_begin(const FtrfCall& c, cudaStream_t st) {
  SideCtx* s = ensure_side();
  if (!s) return nullptr;
  CallState& cs = call_state(c);
  if ((cs.flags & (CALL_TAIL_PENDING | CALL_SCAN_PENDING)) && cs.ctx != s) return nullptr;   // another thread's work pending: in line
  if (cudaEventRecord(s->tail_fork, st) != cudaSuccess || cudaStreamWaitEvent(s->side, s->tail_fork, 0) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return s->side;
}
void side_tail_end(const FtrfCall& c) {
  SideCtx* s = t_side;
  CallState& cs = call_state(c);
  if (s && cudaEventRecord(s->tail_done, s->side) == cudaSuccess) { cs.flags |= CALL_TAIL_PENDING; cs.ctx = s; }
}
int side_tail_join(const FtrfCall& c, cudaStream_t st) {
  CallState& cs = call_state(c);
  if (!(cs.flags & CALL_TAIL_PENDING)) return FTRF_OK;
  cs.flags &= ~(uint64_t)CALL_TAIL_PENDING;
  if (cudaStreamWaitEvent(st, static_cast<SideCtx*>(cs.ctx)->tail_done, 0) != cudaSuccess) return check_launch("side tail join");
  return FTRF_OK;
}

// Same for the reverse scan of the integration backward (only the density backward reads its sigma gradients).
cudaStream_t side_scan_begin(const FtrfCall& c, cudaStream_t st) {
  SideCtx* s = ensure_side();
  if (!s) return nullptr;
  CallState& cs = call_state(c);
  if ((cs.flags & (CALL_TAIL_PENDING | CALL_SCAN_PENDING)) && cs.ctx != s) return nullptr;
  if (cudaEventRecord(s->scan_fork, st) != cudaSuccess || cudaStreamWaitEvent(s->side, s->scan_fork, 0) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return s->side;
}
void side_scan_end(const FtrfCall& c) {
  SideCtx* s = t_side;
  CallState& cs = call_state(c);
  if (s && cudaEventRecord(s->scan_done, s->side) == cudaSuccess) { cs.flags |= CALL_SCAN_PENDING; cs.ctx = s; }
}
int side_scan_join(const FtrfCall& c, cudaStream_t st) {
  CallState& cs = call_state(c);
  if (!(cs.flags & CALL_SCAN_PENDING)) return FTRF_OK;
  cs.flags &= ~(uint64_t)CALL_SCAN_PENDING;
  if (cudaStreamWaitEvent(st, static_cast<SideCtx*>(cs.ctx)->scan_done, 0) != cudaSuccess) return check_launch("side scan join");
  return FTRF_OK;
}

bool prep_ahead_take_scratch(const FtrfCall& c) {
  CallState& cs = call_state(c);
  const bool had = (cs.flags & CALL_SCRATCH_READY) != 0;
  cs.flags &= ~(uint64_t)CALL_SCRATCH_READY;
  return had;
}

// phase 0 (before the call's first kernel is launched): mark the fork point on the caller's stream; phase 1 (after
// that launch, so the host work below does not delay it): build the images on the side stream behind the mark.
int shade_prep_ahead(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st, int phase) {
  CallState& cs = call_state(c);
  if (phase == 0) {
    cs.flags = 0;
    cs.ctx = nullptr;
  } else if (!(cs.flags & CALL_FORKED)) {
    return FTRF_OK;
  }
  if (f.shading == FTRF_SHADE_SH || c.cap_app == 0 || !ws.wimg) return FTRF_OK;
  InSpec sp = make_spec(f);
  FusedSpec fs;
  if (sp.in_dim != f.in_dim || !make_fused_spec(f, sp, fs) || fs.kp > ws.ldx) return FTRF_OK;
  SideCtx* s = ensure_side();
  if (!s) return FTRF_OK;                        // no side stream: the launchers build the images in line
  // after everything already on the caller's stream (the previous call's kernels still read the images)
  if (phase == 0) {
    if (cudaEventRecord(s->fork, st) != cudaSuccess) return check_launch("prep_ahead fork");
    cs.flags |= CALL_FORKED;
    cs.ctx = s;
    return FTRF_OK;
  }
  cs.flags &= ~(uint64_t)CALL_FORKED;
  if (cudaStreamWaitEvent(s->side, s->fork, 0) != cudaSuccess) return check_launch("prep_ahead fork wait");
  float* img = ws.wimg;
  float* w1p = c.save_for_backward ? img + fused_w1p_offset_floats() : nullptr;
  prep_weights_kernel<<<72, 128, 0, s->side>>>(f, fs, img, w1p, sp.in_dim);
  if (int e = check_launch("prep_weights")) return e;
  cs.flags |= CALL_IMG_FWD_PENDING;
  if (c.save_for_backward && dgrad_variant(f, fs) >= 0) {
    float* bimg = ws.wimg + fused_bwd_offset_floats();
    prep_weights_bwd_kernel<<<72, 128, 0, s->side>>>(f, fs, bimg, sp.in_dim);
    if (int e = check_launch("prep_weights_bwd")) return e;
    // the weight-gradient kernel's zeroed dW1 scratch (read by nothing before it): same side stream
    float* scratch = ws.wimg + fused_scratch_offset_floats();
    if (cudaMemsetAsync(scratch, 0, wgrad_fused_scratch_bytes(), s->side) != cudaSuccess) return check_launch("wgrad scratch memset");
    cs.flags |= CALL_IMG_BWD_READY | CALL_SCRATCH_READY;
  }
  if (cudaEventRecord(s->done, s->side) != cudaSuccess) return check_launch("prep_ahead done");
  return FTRF_OK;
}

bool shade_dgrad_is_fused(const FtrfField& f, const FtrfCall& c, const Ws& ws) {
  InSpec sp = make_spec(f);
  FusedSpec fs;
  return make_fused_spec(f, sp, fs) && fs.kp <= ws.ldx && ws.wimg && dgrad_variant(f, fs) >= 0;
}

int launch_shade_dgrad_fused(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st, bool* handled) {
  *handled = false;
  InSpec sp = make_spec(f);
  FusedSpec fs;
  if (!make_fused_spec(f, sp, fs) || fs.kp > ws.ldx || !ws.wimg) return FTRF_OK;
  using SFea = StaticSpec<FTRF_SHADE_MLP_FEA, 27, 2, 2, 2>;
  using SPe = StaticSpec<FTRF_SHADE_MLP_PE, 27, 2, 2, 2>;
  using SFew = StaticSpec<FTRF_SHADE_MLP, 18, 8, 2, 2>;
  const int which = dgrad_variant(f, fs);
  if (which < 0) return FTRF_OK;
  *handled = true;
  if (c.cap_app == 0) return FTRF_OK;
  float* img = ws.wimg + fused_bwd_offset_floats();
  CallState& cs = call_state(c);
  const int masks = (cs.flags & CALL_MASKS_READY) && g_dgrad_masks ? 1 : 0;     // this call's fused forward wrote the sign masks
  cs.flags &= ~(uint64_t)CALL_MASKS_READY;
  if (cs.flags & CALL_IMG_BWD_READY) {
    cs.flags &= ~(uint64_t)CALL_IMG_BWD_READY;      // built by this call's forward on the side stream, which the forward joined
  } else {
    prep_weights_bwd_kernel<<<72, 128, 0, st>>>(f, fs, img, sp.in_dim);
    if (int e = check_launch("prep_weights_bwd")) return e;
  }
  const size_t smem = ((size_t)BZ_GROUPS * BZ_A_STAGES * 2 * FZ_CHUNK_FLOATS + (size_t)BZ_W_STAGES * BZ_WSTAGE_FLOATS) * 4 +
                      (size_t)(FZ_MAX_K / 2) * sizeof(SPair) + (size_t)3 * FC * 4 + 24 * 8 + 16;
  int64_t tiles = (c.cap_app + 127) / 128;
  int grid = (int)(tiles < FTRF_SMS ? tiles : FTRF_SMS);
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, BZ_THREADS, smem, st>>>(f, c, ws, fs, img, g_dgrad_dbg, masks);
  };
  if (which == 0) go(shade_fused_dgrad_kernel<SFea>);
  else if (which == 1) go(shade_fused_dgrad_kernel<SPe>);
  else go(shade_fused_dgrad_kernel<SFew>);
  return check_launch("shade_fused_dgrad");
}

static size_t fused_fwd_smem() {
  return (size_t)FZ_A_STAGES * 4 * FZ_CHUNK_FLOATS * 4 + (size_t)FZ_W_STAGES * 2 * FZ_CHUNK_FLOATS * 4 +
         (size_t)(2 * FC + 3 * FC) * 4 + (size_t)(FZ_MAX_K / 2) * sizeof(SPair) + 16 * 8 + 16;
}

// Forward of the MLP shading modes.  *handled = false when this configuration is not covered.
int launch_shade_fused(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st, bool* handled) {
  *handled = false;
  InSpec sp = make_spec(f);
  FusedSpec fs;
  if (!make_fused_spec(f, sp, fs) || fs.kp > ws.ldx || !ws.wimg) return FTRF_OK;
  *handled = true;
  if (c.cap_app == 0) return FTRF_OK;
  float* img = ws.wimg;
  float* w1p = c.save_for_backward ? img + fused_w1p_offset_floats() : nullptr;
  CallState& cs = call_state(c);
  if (cs.flags & CALL_IMG_FWD_PENDING) {
    cs.flags &= ~(uint64_t)CALL_IMG_FWD_PENDING;   // built at the start of this call on the side stream: join
    if (cudaStreamWaitEvent(st, static_cast<SideCtx*>(cs.ctx)->done, 0) != cudaSuccess) return check_launch("prep_ahead join");
  } else {
    cs.flags &= ~(uint64_t)(CALL_IMG_BWD_READY | CALL_SCRATCH_READY);   // (stage-level call: nothing was built ahead)
    prep_weights_kernel<<<72, 128, 0, st>>>(f, fs, img, w1p, sp.in_dim);
    if (int e = check_launch("prep_weights")) return e;
  }
  const size_t smem = fused_fwd_smem();
  // the block layout of the saved activations is only understood by the fused backward kernels: both must cover this field
  const int perm = (g_act_perm && c.save_for_backward && shade_dgrad_is_fused(f, c, ws)) ? 1 : 0;
  int64_t pairs = (c.cap_app + FZ_PAIR_ROWS - 1) / FZ_PAIR_ROWS;
  int grid = (int)(pairs < FTRF_SMS ? pairs : FTRF_SMS);
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, FZ_THREADS, smem, st>>>(f, c, ws, fs, img, g_fused_dbg, perm);
  };
  // shipped configurations (app_dim 27, pe 2/2/2 — configs/*.txt, config.yaml) get the unrolled build
  using SFea = StaticSpec<FTRF_SHADE_MLP_FEA, 27, 2, 2, 2>;
  using SPe = StaticSpec<FTRF_SHADE_MLP_PE, 27, 2, 2, 2>;
  using SMlp = StaticSpec<FTRF_SHADE_MLP, 27, 2, 2, 2>;
  using SFew = StaticSpec<FTRF_SHADE_MLP, 18, 8, 2, 2>;      // configs/free_tensorf_human.txt:46-54
  const bool pe2 = f.app_dim == 27 && f.view_pe == 2;
#ifdef FTRF_TEST_BUILD
  if (g_fwd_ts) {                                  // the TMEM-operand experiment (see shade_fwd_ts_kernel)
    const int64_t tiles = (c.cap_app + FT_TILE - 1) / FT_TILE;
    const int grid_ts = (int)(tiles < FTRF_SMS ? tiles : FTRF_SMS);
    auto go_ts = [&](auto kern) {
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd_ts_smem());
      kern<<<grid_ts, FT_THREADS, fwd_ts_smem(), st>>>(f, c, ws, fs, img, g_fused_dbg);
    };
    if (f.shading == FTRF_SHADE_MLP_FEA && pe2 && f.fea_pe == 2 && fs.kp == SFea::kp) go_ts(shade_fwd_ts_kernel<SFea>);
    else if (f.shading == FTRF_SHADE_MLP_PE && pe2 && f.pos_pe == 2 && fs.kp == SPe::kp) go_ts(shade_fwd_ts_kernel<SPe>);
    else if (f.shading == FTRF_SHADE_MLP && pe2 && f.pos_pe == 2 && f.fea_pe == 2 && fs.kp == SMlp::kp) go_ts(shade_fwd_ts_kernel<SMlp>);
    else if (f.shading == FTRF_SHADE_MLP && f.app_dim == 18 && f.pos_pe == 8 && f.view_pe == 2 && f.fea_pe == 2 && fs.kp == SFew::kp) go_ts(shade_fwd_ts_kernel<SFew>);
    else go_ts(shade_fwd_ts_kernel<DynamicSpec>);
    return check_launch("shade_fwd_ts");
  }
#endif
  if (c.save_for_backward && ws.a_gx) cs.flags |= CALL_MASKS_READY;      // the kernels below leave the ReLU sign masks in ws.a_gx
  if (f.shading == FTRF_SHADE_MLP_FEA && pe2 && f.fea_pe == 2 && fs.kp == SFea::kp) go(shade_fused_fwd_kernel<SFea>);
  else if (f.shading == FTRF_SHADE_MLP_PE && pe2 && f.pos_pe == 2 && fs.kp == SPe::kp) go(shade_fused_fwd_kernel<SPe>);
  else if (f.shading == FTRF_SHADE_MLP && pe2 && f.pos_pe == 2 && f.fea_pe == 2 && fs.kp == SMlp::kp) go(shade_fused_fwd_kernel<SMlp>);
  else if (f.shading == FTRF_SHADE_MLP && f.app_dim == 18 && f.pos_pe == 8 && f.view_pe == 2 && f.fea_pe == 2 && fs.kp == SFew::kp) go(shade_fused_fwd_kernel<SFew>);
  else go(shade_fused_fwd_kernel<DynamicSpec>);
  return check_launch("shade_fused_fwd");
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_debug_set_act_perm(int on) {
  ftrf::g_act_perm = on;
  return FTRF_OK;
}

extern "C" int ftrf_debug_set_dgrad_masks(int on) {
  ftrf::g_dgrad_masks = on;
  return FTRF_OK;
}

extern "C" int ftrf_debug_set_fwd_ts(int on) {
  ftrf::g_fwd_ts = on;
  return FTRF_OK;
}

extern "C" int ftrf_debug_dgrad_timeline(void* device_buffer_3x512_i64) {
  ftrf::g_dgrad_dbg = (long long*)device_buffer_3x512_i64;
  return FTRF_OK;
}

extern "C" int ftrf_debug_fused_timeline(void* device_buffer_3x512_i64) {
  ftrf::g_fused_dbg = (long long*)device_buffer_3x512_i64;
  return FTRF_OK;
}
#endif
