// k_appearance.cu — stage 2b: VM / CP appearance gather + basis_mat projection on the packed
// shaded-sample list, and its backward.  Replaces TensorVMSplit/TensorCP.compute_appfeature
// (models/tensoRF.py:230-263, :388-415; basis_mat :149/:338).
//
// A CTA of 256 threads owns a tile of 64 shaded samples, 4 lanes per sample (same channel-chunk
// split as k_density.cu).  The plane*line products (144 for VM 48x3, n_comp for CP) go to shared
// memory, then every 4-lane group does its sample's app_dim x NP mat-vec against the basis matrix
// held in shared memory.  Algorithmic bytes per shaded sample (VM 48x3): 3*48*(4+2)*4 = 3456 B
// forward, 6912 B backward (re-read + gradient reduction).
#include "common.cuh"
#include "tc.cuh"

namespace ftrf {

constexpr int AGS = 4;          // lanes per sample
constexpr int ATILE = 64;       // samples per CTA tile
constexpr int LDP = ATILE + 4;  // row stride of the transposed product tile

__host__ __device__ inline int app_np(const FtrfField& f) {
  return f.model == FTRF_MODEL_VM ? f.app.n_comp[0] + f.app.n_comp[1] + f.app.n_comp[2] : f.app.n_comp[0];
}

__device__ __forceinline__ void app_point(const FtrfField& f, const FtrfCall& c, const Ws& ws, int64_t s, float xn[3]) {
  if (ws.direct_xn) {          // points-in call (ftrf_app_feature)
#pragma unroll
    for (int k = 0; k < 3; ++k) xn[k] = __ldg(ws.direct_xn + 3 * s + k);
    return;
  }
  int slot = ws.a_slot[s];
  Ray ray = load_ray_ws(f, c, ws, ws.s_ray[slot]);
  float p[3];
  sample_pt(ray, sample_z(f, c, ray, ws.s_idx[slot]), p);
  normalize_pt(f, p, xn);
}

// products of this lane's channel chunks -> P row (shared)
__device__ __forceinline__ void app_products(const FtrfField& f, const float xn[3], int sub, float* prow) {
  const FtrfFactorSet& a = f.app;
  if (f.model == FTRF_MODEL_VM) {
    int base = 0;
    const Taps3 T3 = make_taps3(a, xn);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int C = a.n_comp[k];
      const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
      for (int ch = sub * 4; ch < C; ch += AGS * 4) {
        float4 p = plane4(a.plane[k], a.plane_w[k], C, tx, ty, ch);
        float4 l = line4(a.line[k], C, tl, ch);
        if (a.mask[k]) {                                   // tensoRF.py:250, :257
          float4 m = mask4(a.mask[k], a.mask_stride[k], ch);
          p = mul4(p, m);
          l = mul4(l, m);
        }
        *reinterpret_cast<float4*>(prow + base + ch) = mul4(p, l);
      }
      base += C;
    }
  } else {
    const int C = a.n_comp[0];
    Tap t0 = make_tap(xn[vecm(0)], a.line_len[0]);
    Tap t1 = make_tap(xn[vecm(1)], a.line_len[1]);
    Tap t2 = make_tap(xn[vecm(2)], a.line_len[2]);
    for (int ch = sub * 4; ch < C; ch += AGS * 4) {
      float4 v = mul4(mul4(line4(a.line[0], C, t0, ch), line4(a.line[1], C, t1, ch)), line4(a.line[2], C, t2, ch));
      if (a.mask[0]) v = mul4(v, mask4(a.mask[0], a.mask_stride[0], ch));   // tensoRF.py:415
      *reinterpret_cast<float4*>(prow + ch) = v;
    }
  }
}

// products of this lane's channel chunks -> column `sl` of the TRANSPOSED tile PT[k][LDP]
__device__ __forceinline__ void app_products_t(const FtrfField& f, const float xn[3], int sub, float* PT, int sl) {
  const FtrfFactorSet& a = f.app;
  auto put = [&](int k, float4 v) {
    float* p = PT + (size_t)k * LDP + sl;
    p[0] = v.x; p[LDP] = v.y; p[2 * LDP] = v.z; p[3 * LDP] = v.w;
  };
  if (f.model == FTRF_MODEL_VM) {
    int base = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int C = a.n_comp[k];
      Tap tx = make_tap(xn[mat0(k)], a.plane_w[k]);
      Tap ty = make_tap(xn[mat1(k)], a.plane_h[k]);
      Tap tl = make_tap(xn[vecm(k)], a.line_len[k]);
      for (int ch = sub * 4; ch < C; ch += AGS * 4) {
        float4 p = plane4(a.plane[k], a.plane_w[k], C, tx, ty, ch);
        float4 l = line4(a.line[k], C, tl, ch);
        if (a.mask[k]) {                                   // tensoRF.py:250, :257
          float4 m = mask4(a.mask[k], a.mask_stride[k], ch);
          p = mul4(p, m);
          l = mul4(l, m);
        }
        put(base + ch, mul4(p, l));
      }
      base += C;
    }
  } else {
    const int C = a.n_comp[0];
    Tap t0 = make_tap(xn[vecm(0)], a.line_len[0]);
    Tap t1 = make_tap(xn[vecm(1)], a.line_len[1]);
    Tap t2 = make_tap(xn[vecm(2)], a.line_len[2]);
    for (int ch = sub * 4; ch < C; ch += AGS * 4) {
      float4 v = mul4(mul4(line4(a.line[0], C, t0, ch), line4(a.line[1], C, t1, ch)), line4(a.line[2], C, t2, ch));
      if (a.mask[0]) v = mul4(v, mask4(a.mask[0], a.mask_stride[0], ch));   // tensoRF.py:415
      put(ch, v);
    }
  }
}

// Forward.  smem: WT [NP][32] (basis_mat transposed, outputs padded to 32) | PT [NP][LDP] (products,
// transposed).  Phase 1: 4 lanes per sample gather the plane x line products.  Phase 2: the
// [64 x NP] x [NP x app_dim] projection as a register-tiled mini GEMM — warp w owns outputs 4w..4w+3,
// a thread owns 4 samples and half of K (the two halves meet in one shuffle), so every pair of
// 16-byte shared-memory loads feeds 16 FMAs.
__global__ void __launch_bounds__(256)
app_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ __align__(16) float smem[];
  const int NP = app_np(f), D = f.app_dim;
  float* WT = smem;                          // [NP][32]
  float* PT = WT + (size_t)NP * 32;          // [NP][LDP]
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  for (int e = threadIdx.x; e < NP * 32; e += blockDim.x) {
    int k = e >> 5, o = e & 31;
    WT[e] = o < D ? __ldg(f.basis + (size_t)o * NP + k) : 0.f;
  }
  const int sl = threadIdx.x / AGS, sub = threadIdx.x % AGS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int khalf = lane & 1, sq = lane >> 1;            // phase 2: K half, sample quad
  const int kh = (NP / 2 + 3) & ~3;                      // first half of K (multiple of 4)
  for (int64_t t0 = (int64_t)blockIdx.x * ATILE; t0 < n; t0 += (int64_t)gridDim.x * ATILE) {
    const int64_t s = t0 + sl;
    __syncthreads();                     // WT visible / previous tile's PT consumed
    if (s < n) {
      float xn[3];
      app_point(f, c, ws, s, xn);
      app_products_t(f, xn, sub, PT, sl);
    } else {
      for (int k = sub; k < NP; k += AGS) PT[(size_t)k * LDP + sl] = 0.f;
    }
    __syncthreads();
    if (warp * 4 < D) {
      float acc[4][4] = {};
      const int k0 = khalf ? kh : 0, k1 = khalf ? NP : kh;
      const float* pp = PT + sq * 4;
      const float* wp = WT + warp * 4;
#pragma unroll 4
      for (int k = k0; k < k1; ++k) {
        float4 pv = *reinterpret_cast<const float4*>(pp + (size_t)k * LDP);
        float4 wv = *reinterpret_cast<const float4*>(wp + (size_t)k * 32);
        const float ps[4] = {pv.x, pv.y, pv.z, pv.w}, wo[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ps[i], wo[j], acc[i][j]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += __shfl_xor_sync(FULL_MASK, acc[i][j], 1);
      // lane khalf = 0 writes samples 0,1 of the quad, khalf = 1 samples 2,3
#pragma unroll
      for (int ii = 0; ii < 2; ++ii) {
        const int64_t so = t0 + sq * 4 + khalf * 2 + ii;
        if (so < n)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (warp * 4 + j < D) ws.a_feat[so * D + warp * 4 + j] = khalf ? acc[2 + ii][j] : acc[ii][j];
      }
    }
  }
}

// Forward on the tensor cores (VM, every n_comp a multiple of 8 and <= 64, app_dim <= 32).
// A CTA owns tiles of 128 shaded samples.  Per tile: 128 threads compute the three taps of their sample
// once into shared memory; then, per factor pair k, all 256 threads gather plane x line products —
// consecutive lanes take consecutive 16-byte channel chunks of one texel, so a warp reads whole 192-byte
// texels — split them into tf32 hi / lo and store them as a UMMA K-major operand tile [128 x C_k]; one
// thread issues the 3xTF32 products against the pre-split basis matrix image [32 x C_k]
// (tcgen05.mma M=128 N=32 K=8, accumulating over the three pairs in 32 TMEM columns); the epilogue reads
// the accumulator rows back and stores app_dim features per sample.  Two CTAs per SM overlap one CTA's
// gathers with the other's products.
constexpr int TCT = 128;
constexpr int TC_THREADS = 256 + 64;   // 256 workers + two warps computing the next tile's taps
constexpr int TC_LBO_F = 36;      // operand tile: 144 B between K-adjacent core matrices (128 B + 16 B skew: the 12 lanes that
                                  // write the chunks of one sample row then hit different banks)
FTRF_SWITCH(int, g_app_fwd_tc, 1);           // bring-up switch (ftrf_debug_set_app_tc)

__global__ void __launch_bounds__(TC_THREADS, 2)
app_fwd_tc_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int CM) {
  extern __shared__ __align__(1024) uint8_t app_raw[];
  const FtrfFactorSet& a = f.app;
  const int NP = a.n_comp[0] + a.n_comp[1] + a.n_comp[2], D = f.app_dim;
  float* Bimg = reinterpret_cast<float*>(app_raw);             // per pair k: hi [32 x C_k] then lo [32 x C_k]
  float* Phi = Bimg + (size_t)2 * 32 * NP;                      // [128 x CM]
  float* Plo = Phi + (size_t)(TCT / 8) * (CM / 4) * TC_LBO_F;
  float4* TAP0 = reinterpret_cast<float4*>(Plo + (size_t)(TCT / 8) * (CM / 4) * TC_LBO_F);   // 2 x [128][3] (i0, i1, w0, w1) per axis
  uint64_t* bar = reinterpret_cast<uint64_t*>(TAP0 + 2 * TCT * 3);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if ((int64_t)blockIdx.x * TCT >= n) return;

  {   // basis matrix image (outputs padded to 32 rows of zeros)
    int base = 0, kb = 0;
    for (int k = 0; k < 3; ++k) {
      const int Ck = a.n_comp[k], nq = Ck >> 2;
      for (int e = tid; e < 32 * nq; e += TC_THREADS) {
        const int o = e / nq, ku = e - o * nq;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f), h, l;
        if (o < D) w = ld4(f.basis + (size_t)o * NP + kb + ku * 4);
        const float v[4] = {w.x, w.y, w.z, w.w};
        tc::split_tf32(v[0], h.x, l.x); tc::split_tf32(v[1], h.y, l.y); tc::split_tf32(v[2], h.z, l.z); tc::split_tf32(v[3], h.w, l.w);
        const int off = tc::canon_off(o, ku, nq);
        *reinterpret_cast<float4*>(Bimg + base + off) = h;
        *reinterpret_cast<float4*>(Bimg + base + 32 * Ck + off) = l;
      }
      base += 2 * 32 * Ck;
      kb += Ck;
    }
  }
  // Taps of a tile: slot -> ray -> position is three dependent global round trips.  Two extra warps compute them
  // one tile ahead (two samples per thread) while the 256 workers gather; nobody waits for the chain any more.
  auto make_taps = [&](int64_t tbase, float4* dst) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int r = (tid - 256) * 2 + u;
      const int64_t s = tbase + r;
      float4 tp[3];
#pragma unroll
      for (int j = 0; j < 3; ++j) tp[j] = make_float4(0.f, 0.f, 0.f, 0.f);      // index 0, weight 0
      if (s < n) {
        float xn[3];
        app_point(f, c, ws, s, xn);
        const Taps3 T = make_taps3(a, xn);
#pragma unroll
        for (int j = 0; j < 3; ++j) tp[j] = make_float4(__int_as_float(T.a[j].i0), __int_as_float(T.a[j].i1), T.a[j].w0, T.a[j].w1);
      }
#pragma unroll
      for (int j = 0; j < 3; ++j) dst[r * 3 + j] = tp[j];
    }
  };
  if (tid >= 256) make_taps((int64_t)blockIdx.x * TCT, TAP0);
  if (tid == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc(tslot, 32);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tslot;
  const uint32_t idesc = tc::idesc_tf32(128, 32);
  uint32_t ph = 0;
  int par = 0;                      // tap buffer of the current tile

  for (int64_t t0 = (int64_t)blockIdx.x * TCT; t0 < n; t0 += (int64_t)gridDim.x * TCT) {
    const float4* TAP = TAP0 + (size_t)par * (TCT * 3);
    if (tid >= 256) {
      const int64_t tn = t0 + (int64_t)gridDim.x * TCT;
      if (tn < n) make_taps(tn, TAP0 + (size_t)(par ^ 1) * (TCT * 3));
    }
    int bbase = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int Ck = a.n_comp[k], nq = Ck >> 2, W = a.plane_w[k];
      if (k > 0) {                      // the previous pair's products have consumed the operand tile
        tc::mbar_wait(bar, ph);
        ph ^= 1;
        tc::fence_after_sync();
      }
      // two items per iteration, all twelve gathers issued before the first use
      for (int item = tid; item < TCT * nq && tid < 256; item += 512) {
        float4 pv[2], lv[2];
        int off[2];
        bool on[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int it2 = item + u * 256;
          on[u] = it2 < TCT * nq;
          const int sl = on[u] ? it2 / nq : 0, cq = on[u] ? it2 - sl * nq : 0, ch = cq * 4;
          const float4 ax = TAP[sl * 3 + mat0(k)], ay = TAP[sl * 3 + mat1(k)], al = TAP[sl * 3 + vecm(k)];
          const Tap tx = {__float_as_int(ax.x), __float_as_int(ax.y), ax.z, ax.w};
          const Tap ty = {__float_as_int(ay.x), __float_as_int(ay.y), ay.z, ay.w};
          const Tap tl = {__float_as_int(al.x), __float_as_int(al.y), al.z, al.w};
          pv[u] = plane4(a.plane[k], W, Ck, tx, ty, ch);
          lv[u] = line4(a.line[k], Ck, tl, ch);
          if (a.mask[k]) {                                   // (plane m)(line m), tensoRF.py:250, :257
            const float4 m = mask4(a.mask[k], a.mask_stride[k], ch);
            pv[u] = mul4(pv[u], m);
            lv[u] = mul4(lv[u], m);
          }
          off[u] = ((sl >> 3) * nq + cq) * TC_LBO_F + (sl & 7) * 4;
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (!on[u]) continue;
          const float4 v4 = mul4(pv[u], lv[u]);
          float4 h, l;
          const float v[4] = {v4.x, v4.y, v4.z, v4.w};
          tc::split_tf32(v[0], h.x, l.x); tc::split_tf32(v[1], h.y, l.y); tc::split_tf32(v[2], h.z, l.z); tc::split_tf32(v[3], h.w, l.w);
          *reinterpret_cast<float4*>(Phi + off[u]) = h;
          *reinterpret_cast<float4*>(Plo + off[u]) = l;
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      __syncthreads();
      if (tid == 0) {
        tc::fence_after_sync();
        const uint32_t hi_w = tc::desc_hi((uint32_t)nq * 128u), hi_a = tc::desc_hi((uint32_t)nq * TC_LBO_F * 4u);
        const uint32_t ah = tc::desc_lo(tc::smem_u32(Phi), TC_LBO_F * 4), al_ = tc::desc_lo(tc::smem_u32(Plo), TC_LBO_F * 4);
        const uint32_t bh = tc::desc_lo(tc::smem_u32(Bimg + bbase), 128), bl = tc::desc_lo(tc::smem_u32(Bimg + bbase + 32 * Ck), 128);
        for (int ks = 0; ks < (Ck >> 3); ++ks) {
          const uint32_t o = (uint32_t)ks * 16u;          // two core-matrix columns per k-step
          const uint32_t oa = (uint32_t)ks * (2u * TC_LBO_F * 4u / 16u);
          tc::mma_tf32_ss_lo2(tmem, al_ + oa, hi_a, bh + o, hi_w, idesc, (k == 0 && ks == 0) ? 0u : 1u);
          tc::mma_tf32_ss_lo2(tmem, ah + oa, hi_a, bl + o, hi_w, idesc, 1u);
          tc::mma_tf32_ss_lo2(tmem, ah + oa, hi_a, bh + o, hi_w, idesc, 1u);
        }
        tc::mma_commit(bar);
      }
      bbase += 2 * 32 * Ck;
    }
    tc::mbar_wait(bar, ph);
    ph ^= 1;
    tc::fence_after_sync();
    // accumulator rows -> shared (the operand tile is free: its last products have completed) -> one
    // contiguous block of app_dim * 128 floats in global memory, written with full 16-byte lanes
    float* stage = Phi;
    if (warp < 4) {
      float acc[32];
      tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), acc);
      tc::tmem_wait_ld();
#pragma unroll
      for (int o = 0; o < 32; ++o)
        if (o < D) stage[tid * D + o] = acc[o];
    }
    tc::fence_before_sync();
    __syncthreads();
    {
      const int rows = (int)min((int64_t)TCT, n - t0);
      const int total = rows * D;                       // floats; the block starts 16-byte aligned (t0 % 128 == 0)
      float* dst = ws.a_feat + t0 * D;
      if ((((size_t)dst) & 15) == 0) {
        for (int e = tid * 4; e < total && tid < 256; e += 256 * 4) {
          if (e + 4 <= total) *reinterpret_cast<float4*>(dst + e) = *reinterpret_cast<const float4*>(stage + e);
          else for (int j = e; j < total; ++j) dst[j] = stage[j];
        }
      } else {
        for (int e = tid; e < total && tid < 256; e += 256) dst[e] = stage[e];
      }
    }
    tc::fence_before_sync();
    __syncthreads();                    // accumulator read, operand tile free, the next tile's taps written
    par ^= 1;
  }
  if (warp == 0) tc::tmem_dealloc(tmem, 32);
}


// Warp-specialised build of the same forward (one CTA per SM).  In the kernel above a CTA alternates gather ->
// __syncthreads -> products -> wait (the operand tile is single-buffered) three times per tile and then drains the
// accumulator behind two more block-wide barriers; two CTAs per SM overlap that only coarsely (ncu: L1 at 45 %,
// warps active 28 %).  Here the roles never wait for each other inside a tile:
//   warps 0-11   gather workers: plane x line products of pair k -> operand buffer p % 2 (tf32 hi | lo, K-major)
//   warps 12-13  taps of the NEXT tile (slot -> ray -> position, three dependent global round trips)
//   warp 14      one lane issues the 3xTF32 products of a filled buffer into accumulator tile % 2 (32 TMEM columns each)
//   warps 15-18  epilogue: accumulator rows -> shared staging -> one contiguous block of app_dim * 128 floats
// all meeting through mbarriers only (operand full / empty, taps full / empty, accumulator full / empty).
constexpr int AW_WORKERS = 384, AW_TAP0 = 384, AW_MMA_WARP = 14, AW_EPI_WARP0 = 15;
constexpr int AW_THREADS = (12 + 2 + 1 + 4) * 32;
FTRF_SWITCH(int, g_app_fwd_ws, 1);           // ftrf_debug_set_app_ws (test build): 0 = app_fwd_tc_kernel above

// CP = true: TensorCP (models/tensoRF.py:388-415): one product of three line samples per component, the C components
// cut into chunks of CM (<= 64) that play the role of VM's three factor pairs.
template <bool CP>
__global__ void __launch_bounds__(AW_THREADS, 1)
app_fwd_ws_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int CM) {
  extern __shared__ __align__(1024) uint8_t app_raw[];
  const FtrfFactorSet& a = f.app;
  const int NP = CP ? a.n_comp[0] : a.n_comp[0] + a.n_comp[1] + a.n_comp[2], D = f.app_dim;
  const int nch = CP ? a.n_comp[0] / CM : 3;                        // operand chunks per tile
  const size_t PSZ = (size_t)(TCT / 8) * (CM / 4) * TC_LBO_F;      // floats of one hi (or lo) operand tile [128 x CM]
  float* Bimg = reinterpret_cast<float*>(app_raw);                 // per pair k: hi [32 x C_k] then lo [32 x C_k]
  float* P = Bimg + (size_t)2 * 32 * NP;                           // [2 buffers][hi, lo][PSZ]
  float4* TAP0 = reinterpret_cast<float4*>(P + 4 * PSZ);           // 2 x [128][3] (i0, i1, w0, w1) per axis
  float* stage = reinterpret_cast<float*>(TAP0 + 2 * TCT * 3);     // [128][D]
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage + TCT * 32);
  uint64_t* p_full = bars;            // [2] 12 warp arrivals
  uint64_t* p_empty = bars + 2;       // [2] tcgen05.commit
  uint64_t* tap_full = bars + 4;      // [2] 2 warp arrivals
  uint64_t* tap_empty = bars + 6;     // [2] 12 warp arrivals
  uint64_t* acc_full = bars + 8;      // [2] tcgen05.commit
  uint64_t* acc_empty = bars + 10;    // [2] 4 warp arrivals
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 12);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int64_t ntiles = (n + TCT - 1) / TCT;
  if ((int64_t)blockIdx.x >= ntiles) return;
  const int my_tiles = (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  auto tile_base = [&](int i) { return ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * TCT; };

  {   // basis matrix image (outputs padded to 32 rows of zeros)
    int base = 0, kb = 0;
    for (int k = 0; k < nch; ++k) {
      const int Ck = CP ? CM : a.n_comp[k], nq = Ck >> 2;
      for (int e = tid; e < 32 * nq; e += AW_THREADS) {
        const int o = e / nq, ku = e - o * nq;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f), h, l;
        if (o < D) w = ld4(f.basis + (size_t)o * NP + kb + ku * 4);
        const float v[4] = {w.x, w.y, w.z, w.w};
        tc::split_tf32(v[0], h.x, l.x); tc::split_tf32(v[1], h.y, l.y); tc::split_tf32(v[2], h.z, l.z); tc::split_tf32(v[3], h.w, l.w);
        const int off = tc::canon_off(o, ku, nq);
        *reinterpret_cast<float4*>(Bimg + base + off) = h;
        *reinterpret_cast<float4*>(Bimg + base + 32 * Ck + off) = l;
      }
      base += 2 * 32 * Ck;
      kb += Ck;
    }
  }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(p_full + i, 12); tc::mbar_init(p_empty + i, 1);
      tc::mbar_init(tap_full + i, 2); tc::mbar_init(tap_empty + i, 12);
      tc::mbar_init(acc_full + i, 1); tc::mbar_init(acc_empty + i, 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == AW_MMA_WARP) tc::tmem_alloc(tslot, 64);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tslot;

  if (tid < AW_WORKERS) {
    // =============================================================== gather workers
    for (int i = 0; i < my_tiles; ++i) {
      const float4* TAP = TAP0 + (size_t)(i & 1) * (TCT * 3);
      tc::mbar_wait(tap_full + (i & 1), (i >> 1) & 1);
      for (int k = 0; k < nch; ++k) {
        const int Ck = CP ? CM : a.n_comp[k], nq = Ck >> 2, W = CP ? 0 : a.plane_w[k];
        const uint32_t p = (uint32_t)i * (uint32_t)nch + (uint32_t)k, b = p & 1u;
        float* Phi = P + (size_t)b * 2 * PSZ;
        float* Plo = Phi + PSZ;
        tc::mbar_wait(p_empty + b, ((p >> 1) & 1u) ^ 1u);
        // two items per iteration, all twelve gathers issued before the first use
        for (int item = tid; item < TCT * nq; item += 2 * AW_WORKERS) {
          float4 pv[2], lv[2], mv[2];
          int off[2];
          bool on[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int it2 = item + u * AW_WORKERS;
            on[u] = it2 < TCT * nq;
            const int sl = on[u] ? it2 / nq : 0, cq = on[u] ? it2 - sl * nq : 0, ch = cq * 4;
            const float4 ax = TAP[sl * 3 + (CP ? 0 : mat0(k))], ay = TAP[sl * 3 + (CP ? 1 : mat1(k))], al = TAP[sl * 3 + (CP ? 2 : vecm(k))];
            const Tap tx = {__float_as_int(ax.x), __float_as_int(ax.y), ax.z, ax.w};
            const Tap ty = {__float_as_int(ay.x), __float_as_int(ay.y), ay.z, ay.w};
            const Tap tl = {__float_as_int(al.x), __float_as_int(al.y), al.z, al.w};
            mv[u] = make_float4(1.f, 1.f, 1.f, 1.f);
            if (CP) {                                            // TAP[sl][j] is line j's tap: (l0 l1) l2, then the mask
              const int cg = k * CM + ch;                        // component index in the lines
              pv[u] = mul4(line4(a.line[0], NP, tx, cg), line4(a.line[1], NP, ty, cg));
              lv[u] = line4(a.line[2], NP, tl, cg);
              if (a.mask[0]) mv[u] = mask4(a.mask[0], a.mask_stride[0], cg);      // tensoRF.py:415
            } else {
              pv[u] = plane4(a.plane[k], W, Ck, tx, ty, ch);
              lv[u] = line4(a.line[k], Ck, tl, ch);
              if (a.mask[k]) {                                   // (plane m)(line m), tensoRF.py:250, :257
                const float4 m = mask4(a.mask[k], a.mask_stride[k], ch);
                pv[u] = mul4(pv[u], m);
                lv[u] = mul4(lv[u], m);
              }
            }
            off[u] = ((sl >> 3) * nq + cq) * TC_LBO_F + (sl & 7) * 4;
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            if (!on[u]) continue;
            float4 v4 = mul4(pv[u], lv[u]);
            if (CP) v4 = mul4(v4, mv[u]);
            float4 h, l;
            const float v[4] = {v4.x, v4.y, v4.z, v4.w};
            tc::split_tf32(v[0], h.x, l.x); tc::split_tf32(v[1], h.y, l.y); tc::split_tf32(v[2], h.z, l.z); tc::split_tf32(v[3], h.w, l.w);
            *reinterpret_cast<float4*>(Phi + off[u]) = h;
            *reinterpret_cast<float4*>(Plo + off[u]) = l;
          }
        }
        tc::fence_async_smem();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(p_full + b);
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tap_empty + (i & 1));
    }
  } else if (warp < AW_MMA_WARP) {
    // =============================================================== taps, one tile ahead of the workers
    for (int i = 0; i < my_tiles; ++i) {
      float4* dst = TAP0 + (size_t)(i & 1) * (TCT * 3);
      const int64_t tbase = tile_base(i);
      tc::mbar_wait(tap_empty + (i & 1), ((i >> 1) & 1) ^ 1);
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int r = (tid - AW_TAP0) * 2 + u;
        const int64_t s = tbase + r;
        float4 tp[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) tp[j] = make_float4(0.f, 0.f, 0.f, 0.f);      // index 0, weight 0
        if (s < n) {
          float xn[3];
          app_point(f, c, ws, s, xn);
          Taps3 T;
          if (CP) {
#pragma unroll
            for (int j = 0; j < 3; ++j) T.a[j] = make_tap(xn[vecm(j)], a.line_len[j]);
          } else {
            T = make_taps3(a, xn);
          }
#pragma unroll
          for (int j = 0; j < 3; ++j) tp[j] = make_float4(__int_as_float(T.a[j].i0), __int_as_float(T.a[j].i1), T.a[j].w0, T.a[j].w1);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) dst[r * 3 + j] = tp[j];
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tap_full + (i & 1));
    }
  } else if (warp == AW_MMA_WARP) {
    // =============================================================== products
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_tf32(128, 32);
      for (int i = 0; i < my_tiles; ++i) {
        const uint32_t ab = (uint32_t)i & 1u, d = tmem + ab * 32u;
        tc::mbar_wait(acc_empty + ab, (((uint32_t)i >> 1) & 1u) ^ 1u);
        int bbase = 0;
        for (int k = 0; k < nch; ++k) {
          const int Ck = CP ? CM : a.n_comp[k], nq = Ck >> 2;
          const uint32_t p = (uint32_t)i * (uint32_t)nch + (uint32_t)k, b = p & 1u;
          const float* Phi = P + (size_t)b * 2 * PSZ;
          tc::mbar_wait(p_full + b, (p >> 1) & 1u);
          tc::fence_after_sync();
          const uint32_t hi_w = tc::desc_hi((uint32_t)nq * 128u), hi_a = tc::desc_hi((uint32_t)nq * TC_LBO_F * 4u);
          const uint32_t ah = tc::desc_lo(tc::smem_u32(Phi), TC_LBO_F * 4), al_ = tc::desc_lo(tc::smem_u32(Phi + PSZ), TC_LBO_F * 4);
          const uint32_t bh = tc::desc_lo(tc::smem_u32(Bimg + bbase), 128), bl = tc::desc_lo(tc::smem_u32(Bimg + bbase + 32 * Ck), 128);
          for (int ks = 0; ks < (Ck >> 3); ++ks) {
            const uint32_t o = (uint32_t)ks * 16u;          // two core-matrix columns per k-step
            const uint32_t oa = (uint32_t)ks * (2u * TC_LBO_F * 4u / 16u);
            tc::mma_tf32_ss_lo2(d, al_ + oa, hi_a, bh + o, hi_w, idesc, (k == 0 && ks == 0) ? 0u : 1u);
            tc::mma_tf32_ss_lo2(d, ah + oa, hi_a, bl + o, hi_w, idesc, 1u);
            tc::mma_tf32_ss_lo2(d, ah + oa, hi_a, bh + o, hi_w, idesc, 1u);
          }
          tc::mma_commit(p_empty + b);
          if (k == nch - 1) tc::mma_commit(acc_full + ab);
          bbase += 2 * 32 * Ck;
        }
      }
    }
  } else {
    // =============================================================== epilogue: accumulator rows -> staging -> global
    const int q = warp & 3, row = q * 32 + lane, et = tid - AW_EPI_WARP0 * 32;
    for (int i = 0; i < my_tiles; ++i) {
      const uint32_t ab = (uint32_t)i & 1u;
      const int64_t t0 = tile_base(i);
      tc::mbar_wait(acc_full + ab, ((uint32_t)i >> 1) & 1u);
      tc::fence_after_sync();
      float acc[32];
      tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + ab * 32u, acc);
      tc::tmem_wait_ld();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(acc_empty + ab);
#pragma unroll
      for (int o = 0; o < 32; ++o)
        if (o < D) stage[row * D + o] = acc[o];
      asm volatile("bar.sync 2, 128;" ::: "memory");
      {
        const int rows = (int)min((int64_t)TCT, n - t0);
        const int total = rows * D;                       // floats; the block starts 16-byte aligned (t0 % 128 == 0)
        float* dst = ws.a_feat + t0 * D;
        if ((((size_t)dst) & 15) == 0) {
          for (int e = et * 4; e < total; e += 128 * 4) {
            if (e + 4 <= total) *reinterpret_cast<float4*>(dst + e) = *reinterpret_cast<const float4*>(stage + e);
            else for (int j = e; j < total; ++j) dst[j] = stage[j];
          }
        } else {
          for (int e = et; e < total; e += 128) dst[e] = stage[e];
        }
      }
      asm volatile("bar.sync 2, 128;" ::: "memory");      // the staging block is free for the next tile
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == AW_MMA_WARP) tc::tmem_dealloc(tmem, 64);
}

// Backward: dWb += G^T P (register tiles, flushed once per CTA), g_prod = Wb^T g_feat, scatter.
constexpr int BW_TT = 2;        // (4 outputs x 4 comps) register tiles per thread
FTRF_SWITCH(int, g_app_bwd_onegather, 1);    // VM: single-gather kernel (app_bwd_vm_kernel); 0 = two-gather kernel below
FTRF_SWITCH(int, g_app_bwd_runs, -1);        // samples per run-merging segment of the VM scatter; 0 = one reduction set per sample, -1 = auto

__global__ void __launch_bounds__(512, 1)
app_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int arun) {
  extern __shared__ __align__(16) float smem[];
  const int NP = app_np(f), NPs = NP + 4, D = f.app_dim, Dp = (D + 3) & ~3;
  float* Wb = smem;
  float* P = Wb + (size_t)D * NPs;
  float* G = P + (size_t)ATILE * NPs;
  float* GT = G + (size_t)ATILE * Dp;          // [Dp][LDP]  g_feat transposed (samples contiguous)
  float* XN = GT + (size_t)Dp * LDP;           // [ATILE][3]  normalised sample positions
  const FtrfFactorSet& a = f.app;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  for (int e = threadIdx.x; e < D * NP; e += blockDim.x) Wb[(e / NP) * NPs + (e % NP)] = __ldg(f.basis + e);
  const int sl = threadIdx.x / AGS, sub = threadIdx.x % AGS;
  const int n_og = Dp / 4, n_kg = NP / 4, n_tt = n_og * n_kg;
  float dW[BW_TT][16];
#pragma unroll
  for (int t = 0; t < BW_TT; ++t)
#pragma unroll
    for (int i = 0; i < 16; ++i) dW[t][i] = 0.f;

  for (int64_t t0 = (int64_t)blockIdx.x * ATILE; t0 < n; t0 += (int64_t)gridDim.x * ATILE) {
    const int64_t s = t0 + sl;
    __syncthreads();
    float xn[3] = {0.f, 0.f, 0.f};
    if (sl >= ATILE) {
    } else if (s < n) {
      app_point(f, c, ws, s, xn);
      if (sub < 3) XN[sl * 3 + sub] = sub == 0 ? xn[0] : (sub == 1 ? xn[1] : xn[2]);
      app_products(f, xn, sub, P + (size_t)sl * NPs);
      for (int o = sub; o < Dp; o += AGS) {
        float g = o < D ? ws.a_gfeat[s * D + o] : 0.f;
        G[sl * Dp + o] = g;
        GT[o * LDP + sl] = g;
      }
    } else {
      if (sub < 3) XN[sl * 3 + sub] = 0.f;
      for (int k = sub * 4; k < NP; k += AGS * 4) *reinterpret_cast<float4*>(P + (size_t)sl * NPs + k) = make_float4(0, 0, 0, 0);
      for (int o = sub; o < Dp; o += AGS) { G[sl * Dp + o] = 0.f; GT[o * LDP + sl] = 0.f; }
    }
    __syncthreads();
    // (a) basis gradient: 4x4 register tiles over (output group, component group)
#pragma unroll
    for (int t = 0; t < BW_TT; ++t) {
      int tt = threadIdx.x + t * blockDim.x;
      if (tt < n_tt) {
        int og = tt / n_kg, kg = tt % n_kg;
        for (int r = 0; r < ATILE; ++r) {
          float4 gv = *reinterpret_cast<const float4*>(G + r * Dp + og * 4);
          float4 pv = *reinterpret_cast<const float4*>(P + (size_t)r * NPs + kg * 4);
          const float gg[4] = {gv.x, gv.y, gv.z, gv.w}, pp[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dW[t][i * 4 + j] = fmaf(gg[i], pp[j], dW[t][i * 4 + j]);
        }
      }
    }
    // (b) gradient w.r.t. the products, GQ[64 x NP] = G[64 x D] Wb[D x NP], as a register-tiled mini
    // GEMM (4 samples x 4 components per thread) written over the product tile
    __syncthreads();                       // every dW tile has consumed P
    for (int id = threadIdx.x; id < n_kg * (ATILE / 4); id += blockDim.x) {
      const int kg = id % n_kg, sq = id / n_kg;
      float acc[4][4] = {};
      for (int o = 0; o < D; ++o) {
        float4 wv = *reinterpret_cast<const float4*>(Wb + (size_t)o * NPs + kg * 4);
        float4 gv = *reinterpret_cast<const float4*>(GT + o * LDP + sq * 4);
        const float ww[4] = {wv.x, wv.y, wv.z, wv.w}, gs[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(gs[i], ww[j], acc[i][j]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(P + (size_t)(sq * 4 + i) * NPs + kg * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // scatter into plane / line gradients
    if (arun > 0 && f.model == FTRF_MODEL_VM) {
      // run merging (see k_density.cu): a thread owns one 4-channel chunk of one factor and walks `arun`
      // consecutive samples of the tile, reducing into global memory only when the cell changes
      const int nq = NP / 4, nsegs = (ATILE + arun - 1) / arun;
      const int tile_n = (int)min((int64_t)ATILE, n - t0);
      for (int item = threadIdx.x; item < nsegs * nq; item += blockDim.x) {
        const int seg = item / nq, q = item - seg * nq;
        int k = 0, base = 0;
        while (k < 2 && 4 * q >= base + a.n_comp[k]) { base += a.n_comp[k]; ++k; }
        const int ch = 4 * q - base, C = a.n_comp[k], W = a.plane_w[k], H = a.plane_h[k], L = a.line_len[k];
        const int m0 = mat0(k), m1 = mat1(k), mv = vecm(k);
        float4 m2 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (a.mask[k]) { const float4 m = mask4(a.mask[k], a.mask_stride[k], ch); m2 = mul4(m, m); }
        PlaneRun pr; LineRun lr;
        pr.kx = -1; pr.ky = -1; lr.kl = -1;
        const int i1 = min(tile_n, (seg + 1) * arun);
        for (int i = seg * arun; i < i1; ++i) {
          const Tap tx = make_tap(XN[i * 3 + m0], W);
          const Tap ty = make_tap(XN[i * 3 + m1], H);
          const Tap tl = make_tap(XN[i * 3 + mv], L);
          const float4 gq = mul4(*reinterpret_cast<const float4*>(P + (size_t)i * NPs + 4 * q), m2);
          const float4 pv = plane4(a.plane[k], W, C, tx, ty, ch);
          const float4 lv = line4(a.line[k], C, tl, ch);
          const float4 gp = mul4(gq, lv), gl = mul4(gq, pv);
          const int kx = tx.i0 | (tx.i1 << 16), ky = ty.i0 | (ty.i1 << 16), kl = tl.i0 | (tl.i1 << 16);
          if (kx != pr.kx || ky != pr.ky) {
            flush_plane(pr, a.g_plane[k], W, C, ch);
            pr.kx = kx; pr.ky = ky;
            pr.a00 = z4(); pr.a10 = z4(); pr.a01 = z4(); pr.a11 = z4();
          }
          pr.a00 = fma4(tx.w0 * ty.w0, gp, pr.a00);
          pr.a10 = fma4(tx.w1 * ty.w0, gp, pr.a10);
          pr.a01 = fma4(tx.w0 * ty.w1, gp, pr.a01);
          pr.a11 = fma4(tx.w1 * ty.w1, gp, pr.a11);
          if (kl != lr.kl) {
            flush_line(lr, a.g_line[k], C, ch);
            lr.kl = kl;
            lr.a0 = z4(); lr.a1 = z4();
          }
          lr.a0 = fma4(tl.w0, gl, lr.a0);
          lr.a1 = fma4(tl.w1, gl, lr.a1);
        }
        flush_plane(pr, a.g_plane[k], W, C, ch);
        flush_line(lr, a.g_line[k], C, ch);
      }
    } else if (arun > 0) {
      // TensorCP, run merging over the three lines: every sample of a step lands in the same few hundred rows of each
      // line, so per-sample reductions are bound by same-address contention; a thread owns one 4-channel chunk and
      // walks `arun` consecutive samples, reducing a line's running gradient only when its cell changes
      const int C = a.n_comp[0], nq = C / 4, nsegs = (ATILE + arun - 1) / arun;
      const int tile_n = (int)min((int64_t)ATILE, n - t0);
      for (int item = threadIdx.x; item < nsegs * nq; item += blockDim.x) {
        const int seg = item / nq, q = item - seg * nq, ch = 4 * q;
        float4 m1 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (a.mask[0]) m1 = mask4(a.mask[0], a.mask_stride[0], ch);          // tensoRF.py:415 (once)
        LineRun lr[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) lr[k].kl = -1;
        const int i1 = min(tile_n, (seg + 1) * arun);
        for (int i = seg * arun; i < i1; ++i) {
          Tap t[3];
          float4 lv[3];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            t[k] = make_tap(XN[i * 3 + vecm(k)], a.line_len[k]);
            lv[k] = line4(a.line[k], C, t[k], ch);
          }
          const float4 gq = mul4(*reinterpret_cast<const float4*>(P + (size_t)i * NPs + ch), m1);
          const float4 gk[3] = {mul4(gq, mul4(lv[1], lv[2])), mul4(gq, mul4(lv[0], lv[2])), mul4(gq, mul4(lv[0], lv[1]))};
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const int kl = t[k].i0 | (t[k].i1 << 16);
            if (kl != lr[k].kl) {
              flush_line(lr[k], a.g_line[k], C, ch);
              lr[k].kl = kl;
              lr[k].a0 = z4(); lr[k].a1 = z4();
            }
            lr[k].a0 = fma4(t[k].w0, gk[k], lr[k].a0);
            lr[k].a1 = fma4(t[k].w1, gk[k], lr[k].a1);
          }
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) flush_line(lr[k], a.g_line[k], C, ch);
      }
    } else if (sl < ATILE && s < n) {
      const float* gqrow = P + (size_t)sl * NPs;
      if (f.model == FTRF_MODEL_VM) {
        int base = 0;
        const Taps3 T3 = make_taps3(a, xn);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int C = a.n_comp[k], W = a.plane_w[k];
          const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
          for (int ch = sub * 4; ch < C; ch += AGS * 4) {
            float4 gq = *reinterpret_cast<const float4*>(gqrow + base + ch);
            if (a.mask[k]) {
              float4 m = mask4(a.mask[k], a.mask_stride[k], ch);
              gq = mul4(gq, mul4(m, m));
            }
            float4 pv = plane4(a.plane[k], W, C, tx, ty, ch);
            float4 lv = line4(a.line[k], C, tl, ch);
            float4 gp = mul4(gq, lv), gl = mul4(gq, pv);
            float* gp0 = a.g_plane[k] + ((size_t)ty.i0 * W) * C + ch;
            float* gp1 = a.g_plane[k] + ((size_t)ty.i1 * W) * C + ch;
            float w;
            if ((w = tx.w0 * ty.w0) != 0.f) red4(gp0 + (size_t)tx.i0 * C, scale4(w, gp));
            if ((w = tx.w1 * ty.w0) != 0.f) red4(gp0 + (size_t)tx.i1 * C, scale4(w, gp));
            if ((w = tx.w0 * ty.w1) != 0.f) red4(gp1 + (size_t)tx.i0 * C, scale4(w, gp));
            if ((w = tx.w1 * ty.w1) != 0.f) red4(gp1 + (size_t)tx.i1 * C, scale4(w, gp));
            if (tl.w0 != 0.f) red4(a.g_line[k] + (size_t)tl.i0 * C + ch, scale4(tl.w0, gl));
            if (tl.w1 != 0.f) red4(a.g_line[k] + (size_t)tl.i1 * C + ch, scale4(tl.w1, gl));
          }
          base += C;
        }
      } else {
        const int C = a.n_comp[0];
        Tap t[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) t[k] = make_tap(xn[vecm(k)], a.line_len[k]);
        for (int ch = sub * 4; ch < C; ch += AGS * 4) {
          float4 gq = *reinterpret_cast<const float4*>(gqrow + ch);
          if (a.mask[0]) gq = mul4(gq, mask4(a.mask[0], a.mask_stride[0], ch));
          float4 l0 = line4(a.line[0], C, t[0], ch), l1 = line4(a.line[1], C, t[1], ch), l2 = line4(a.line[2], C, t[2], ch);
          float4 gk[3] = {mul4(gq, mul4(l1, l2)), mul4(gq, mul4(l0, l2)), mul4(gq, mul4(l0, l1))};
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            if (t[k].w0 != 0.f) red4(a.g_line[k] + (size_t)t[k].i0 * C + ch, scale4(t[k].w0, gk[k]));
            if (t[k].w1 != 0.f) red4(a.g_line[k] + (size_t)t[k].i1 * C + ch, scale4(t[k].w1, gk[k]));
          }
        }
      }
    }
  }
  // flush the basis gradient tiles
#pragma unroll
  for (int t = 0; t < BW_TT; ++t) {
    int tt = threadIdx.x + t * blockDim.x;
    if (tt < n_tt) {
      int og = tt / n_kg, kg = tt % n_kg;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int o = og * 4 + i;
        if (o < D)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float v = dW[t][i * 4 + j];
            if (v != 0.f) atomicAdd(f.g_basis + (size_t)o * NP + kg * 4 + j, v);
          }
      }
    }
  }
}

// VM backward, one gather per sample.  Order per 64-sample tile: g_feat rows and the three taps of every
// sample go to shared memory; GQ = G Wb (gradient w.r.t. the plane x line products) is formed first; then
// ONE pass over the factor texels both rebuilds the products (for dWb += G^T P afterwards) and turns GQ
// into plane / line gradients with run merging.  Work items of that pass are (segment of `arun` samples,
// 16-byte channel chunk), consecutive lanes on consecutive chunks of the same texel.
__global__ void __launch_bounds__(256, 2)
app_bwd_vm_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int arun) {
  extern __shared__ __align__(16) float smem[];
  const FtrfFactorSet& a = f.app;
  const int NP = app_np(f), NPs = NP + 4, D = f.app_dim, Dp = (D + 3) & ~3;
  float* Wb = smem;
  float* P = Wb + (size_t)D * NPs;
  float* Q = P + (size_t)ATILE * NPs;
  float* G = Q + (size_t)ATILE * NPs;
  float* GT = G + (size_t)ATILE * Dp;          // [Dp][LDP]  g_feat transposed (samples contiguous)
  float4* TAP = reinterpret_cast<float4*>(GT + (size_t)Dp * LDP);   // [ATILE][3]
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  for (int e = threadIdx.x; e < D * NP; e += blockDim.x) Wb[(e / NP) * NPs + (e % NP)] = __ldg(f.basis + e);
  const int sl = threadIdx.x / AGS, sub = threadIdx.x % AGS;
  const int n_og = Dp / 4, n_kg = NP / 4, n_tt = n_og * n_kg;
  const int nsegs = (ATILE + arun - 1) / arun;
  float dW[BW_TT][16];
#pragma unroll
  for (int t = 0; t < BW_TT; ++t)
#pragma unroll
    for (int i = 0; i < 16; ++i) dW[t][i] = 0.f;

  for (int64_t t0 = (int64_t)blockIdx.x * ATILE; t0 < n; t0 += (int64_t)gridDim.x * ATILE) {
    const int64_t s = t0 + sl;
    const int tile_n = (int)min((int64_t)ATILE, n - t0);
    __syncthreads();                               // previous tile's dWb pass has consumed P and G
    if (s < n) {
      for (int o = sub; o < Dp; o += AGS) {
        const float g = o < D ? ws.a_gfeat[s * D + o] : 0.f;
        G[sl * Dp + o] = g;
        GT[o * LDP + sl] = g;
      }
    } else {
      for (int o = sub; o < Dp; o += AGS) { G[sl * Dp + o] = 0.f; GT[o * LDP + sl] = 0.f; }
      for (int k = sub * 4; k < NP; k += AGS * 4) *reinterpret_cast<float4*>(P + (size_t)sl * NPs + k) = make_float4(0, 0, 0, 0);
    }
    if (threadIdx.x < ATILE) {
      const int64_t st = t0 + threadIdx.x;
      float4 tp[3];
#pragma unroll
      for (int j = 0; j < 3; ++j) tp[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (st < n) {
        float xn[3];
        app_point(f, c, ws, st, xn);
        const Taps3 T = make_taps3(a, xn);
#pragma unroll
        for (int j = 0; j < 3; ++j) tp[j] = make_float4(__int_as_float(T.a[j].i0), __int_as_float(T.a[j].i1), T.a[j].w0, T.a[j].w1);
      }
#pragma unroll
      for (int j = 0; j < 3; ++j) TAP[threadIdx.x * 3 + j] = tp[j];
    }
    __syncthreads();
    // GQ[64 x NP] = G[64 x D] Wb[D x NP]: register-tiled mini GEMM (4 samples x 4 components per thread)
    for (int id = threadIdx.x; id < n_kg * (ATILE / 4); id += blockDim.x) {
      const int kg = id % n_kg, sq = id / n_kg;
      float acc[4][4] = {};
      for (int o = 0; o < D; ++o) {
        const float4 wv = *reinterpret_cast<const float4*>(Wb + (size_t)o * NPs + kg * 4);
        const float4 gv = *reinterpret_cast<const float4*>(GT + o * LDP + sq * 4);
        const float ww[4] = {wv.x, wv.y, wv.z, wv.w}, gs[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(gs[i], ww[j], acc[i][j]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(Q + (size_t)(sq * 4 + i) * NPs + kg * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // gather once: products for dWb, plane / line gradients with run merging
    for (int item = threadIdx.x; item < nsegs * n_kg; item += blockDim.x) {
      const int seg = item / n_kg, q = item - seg * n_kg;
      int k = 0, base = 0;
      while (k < 2 && 4 * q >= base + a.n_comp[k]) { base += a.n_comp[k]; ++k; }
      const int ch = 4 * q - base, C = a.n_comp[k], W = a.plane_w[k];
      const int m0 = mat0(k), m1 = mat1(k), mv = vecm(k);
      float4 m2 = make_float4(1.f, 1.f, 1.f, 1.f);
      if (a.mask[k]) { const float4 m = mask4(a.mask[k], a.mask_stride[k], ch); m2 = mul4(m, m); }
      const float* pl = a.plane[k] + ch;
      const float* ln = a.line[k] + ch;
      PlaneRun pr; LineRun lr;
      pr.kx = -1; pr.ky = -1; lr.kl = -1;
      const int i1 = min(tile_n, (seg + 1) * arun);
      for (int i = seg * arun; i < i1; ++i) {
        const float4 ax = TAP[i * 3 + m0], ay = TAP[i * 3 + m1], al = TAP[i * 3 + mv];
        const int x0 = __float_as_int(ax.x), x1 = __float_as_int(ax.y), y0 = __float_as_int(ay.x), y1 = __float_as_int(ay.y);
        const int l0 = __float_as_int(al.x), l1 = __float_as_int(al.y);
        const float* r0 = pl + ((size_t)y0 * W) * C;
        const float* r1 = pl + ((size_t)y1 * W) * C;
        const float4 c00 = ld4(r0 + (size_t)x0 * C), c10 = ld4(r0 + (size_t)x1 * C);
        const float4 c01 = ld4(r1 + (size_t)x0 * C), c11 = ld4(r1 + (size_t)x1 * C);
        const float4 e0 = ld4(ln + (size_t)l0 * C), e1 = ld4(ln + (size_t)l1 * C);
        const float w00 = ax.z * ay.z, w10 = ax.w * ay.z, w01 = ax.z * ay.w, w11 = ax.w * ay.w;
        float4 pv = z4(), lv = z4();
        pv = fma4(w00, c00, pv); pv = fma4(w10, c10, pv); pv = fma4(w01, c01, pv); pv = fma4(w11, c11, pv);
        lv = fma4(al.z, e0, lv); lv = fma4(al.w, e1, lv);
        *reinterpret_cast<float4*>(P + (size_t)i * NPs + 4 * q) = mul4(mul4(pv, lv), m2);
        const float4 gq = mul4(*reinterpret_cast<const float4*>(Q + (size_t)i * NPs + 4 * q), m2);
        const float4 gp = mul4(gq, lv), gl = mul4(gq, pv);
        const int kx = x0 | (x1 << 16), ky = y0 | (y1 << 16), kl = l0 | (l1 << 16);
        if (kx != pr.kx || ky != pr.ky) {
          flush_plane(pr, a.g_plane[k], W, C, ch);
          pr.kx = kx; pr.ky = ky;
          pr.a00 = z4(); pr.a10 = z4(); pr.a01 = z4(); pr.a11 = z4();
        }
        pr.a00 = fma4(w00, gp, pr.a00);
        pr.a10 = fma4(w10, gp, pr.a10);
        pr.a01 = fma4(w01, gp, pr.a01);
        pr.a11 = fma4(w11, gp, pr.a11);
        if (kl != lr.kl) {
          flush_line(lr, a.g_line[k], C, ch);
          lr.kl = kl;
          lr.a0 = z4(); lr.a1 = z4();
        }
        lr.a0 = fma4(al.z, gl, lr.a0);
        lr.a1 = fma4(al.w, gl, lr.a1);
      }
      flush_plane(pr, a.g_plane[k], W, C, ch);
      flush_line(lr, a.g_line[k], C, ch);
    }
    __syncthreads();
    // basis gradient: 4x4 register tiles over (output group, component group)
#pragma unroll
    for (int t = 0; t < BW_TT; ++t) {
      const int tt = threadIdx.x + t * blockDim.x;
      if (tt < n_tt) {
        const int og = tt / n_kg, kg = tt % n_kg;
        for (int r = 0; r < ATILE; ++r) {
          const float4 gv = *reinterpret_cast<const float4*>(G + r * Dp + og * 4);
          const float4 pv = *reinterpret_cast<const float4*>(P + (size_t)r * NPs + kg * 4);
          const float gg[4] = {gv.x, gv.y, gv.z, gv.w}, pp[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dW[t][i * 4 + j] = fmaf(gg[i], pp[j], dW[t][i * 4 + j]);
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < BW_TT; ++t) {
    const int tt = threadIdx.x + t * blockDim.x;
    if (tt < n_tt) {
      const int og = tt / n_kg, kg = tt % n_kg;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int o = og * 4 + i;
        if (o < D)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float v = dW[t][i * 4 + j];
            if (v != 0.f) atomicAdd(f.g_basis + (size_t)o * NP + kg * 4 + j, v);
          }
      }
    }
  }
}

static int app_check(const FtrfField& f, size_t smem) {
  int NP = app_np(f), Dp = (f.app_dim + 3) & ~3;
  if (NP % 4 || (Dp / 4) * (NP / 4) > BW_TT * 256) { set_error("appearance: unsupported component count %d", NP); return FTRF_E_ARG; }
  if (smem > 220 * 1024) { set_error("appearance: shared memory %zu B too large", smem); return FTRF_E_ARG; }
  return FTRF_OK;
}

int launch_appearance(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (f.app_dim > 32) { set_error("appearance: app_dim %d > 32 unsupported", f.app_dim); return FTRF_E_ARG; }
  const FtrfFactorSet& a = f.app;
  const int cm = f.model == FTRF_MODEL_VM ? (a.n_comp[0] > a.n_comp[1] ? (a.n_comp[0] > a.n_comp[2] ? a.n_comp[0] : a.n_comp[2]) : (a.n_comp[1] > a.n_comp[2] ? a.n_comp[1] : a.n_comp[2])) : 0;
  if (g_app_fwd_tc && f.model == FTRF_MODEL_VM && a.n_comp[0] % 8 == 0 && a.n_comp[1] % 8 == 0 && a.n_comp[2] % 8 == 0 && cm <= 64) {
    if (g_app_fwd_ws) {
      const size_t smw = ((size_t)2 * 32 * app_np(f) + (size_t)4 * (TCT / 8) * (cm / 4) * TC_LBO_F + (size_t)TCT * 32) * sizeof(float) +
                         (size_t)2 * TCT * 3 * sizeof(float4) + 16 * 8;
      if (smw <= 227 * 1024) {
        cudaFuncSetAttribute(app_fwd_ws_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smw);
        app_fwd_ws_kernel<false><<<FTRF_SMS, AW_THREADS, smw, st>>>(f, c, ws, cm);
        return check_launch("app_fwd_ws");
      }
    }
    const size_t sm = ((size_t)2 * 32 * app_np(f) + (size_t)2 * (TCT / 8) * (cm / 4) * TC_LBO_F) * sizeof(float) + (size_t)2 * TCT * 3 * sizeof(float4) + 32;
    cudaFuncSetAttribute(app_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    app_fwd_tc_kernel<<<FTRF_SMS * 2, TC_THREADS, sm, st>>>(f, c, ws, cm);
    return check_launch("app_fwd_tc");
  }
  if (g_app_fwd_ws && f.model == FTRF_MODEL_CP && a.n_comp[0] % 8 == 0) {
    // TensorCP on the tensor cores: components in chunks of the largest multiple of 8 that divides them (288 -> 6 x 48)
    const int C = a.n_comp[0];
    int ck = 0;
    for (int d = 64; d >= 8 && !ck; d -= 8) if (C % d == 0) ck = d;
    const size_t smw = ((size_t)2 * 32 * C + (size_t)4 * (TCT / 8) * (ck / 4) * TC_LBO_F + (size_t)TCT * 32) * sizeof(float) +
                       (size_t)2 * TCT * 3 * sizeof(float4) + 16 * 8;
    if (ck && smw <= 227 * 1024) {
      cudaFuncSetAttribute(app_fwd_ws_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smw);
      app_fwd_ws_kernel<true><<<FTRF_SMS, AW_THREADS, smw, st>>>(f, c, ws, ck);
      return check_launch("app_fwd_ws_cp");
    }
  }
  size_t smem = ((size_t)app_np(f) * 32 + (size_t)app_np(f) * LDP) * sizeof(float);
  if (int e = app_check(f, smem)) return e;
  cudaFuncSetAttribute(app_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  app_fwd_kernel<<<FTRF_SMS * 3, 256, smem, st>>>(f, c, ws);
  return check_launch("app_fwd");
}

// ------------------------------------------------------------------------------------------------------------
// Appearance backward with both small contractions on the tensor cores (VM).  app_bwd_vm_kernel above spends most
// of its instructions in two CUDA-core mini GEMMs over shared memory (GQ = G Wb and dWb += G^T P); here
//   GQ [64 x NP]  = G [64 x 32] Wb [32 x NP]           one tcgen05 batch per tile (M = 128 with rows 64.. zero)
//   dWb^T [NP x 32] += P^T [NP x 64] G [64 x 32]       accumulated in TMEM over all tiles of the CTA, two M = 128
//                                                       tiles of component rows (0..127 and NP-128..NP-1)
// both 3xTF32 (operands split into tf32 hi / lo, three products), fp32 accumulation: fp32-grade like everything else.
// The gather / run-merging scatter pass in between is the one of app_bwd_vm_kernel; it reads GQ from shared memory
// (copied out of TMEM) and writes the products straight into the K-major operand tile of the second contraction
// (four consecutive samples of one component per 16-byte store; the row-group stride is skewed by 16 bytes so the
// eight lanes of a store phase hit different bank groups).  One CTA of 576 threads per SM.
constexpr int BT = 64;                 // samples per tile
constexpr int BT_WORK = 576;            // threads of the gather / scatter pass: one (4-sample segment, 4-component group) each
constexpr int BT_THREADS = BT_WORK + 64;   // + two warps that compute the NEXT tile's taps meanwhile
constexpr int BT_ISSUER = BT_WORK + 32;  // the thread that issues the MMA batches: lane 0 of the second extra warp, so the
                                         // dWb batch of a tile is issued while the 576 workers already stage the next one
constexpr int BT_RUN = 4;              // samples per scatter segment = one 16-byte K unit of the product operand
constexpr int BT_PT_SBO_F = 16 * 32 + 4;   // floats between 8-row groups of the product operand (2048 B + 16 B skew)
FTRF_SWITCH(int, g_app_bwd_tc, 1);                  // ftrf_debug_set_app_bwd_tc

struct AppBwdTcLayout { int np, rows_pt, qcols, tmem_cols; size_t wbi, g, gt, pt, q, tap, gf, bars, total; };
static __host__ __device__ inline AppBwdTcLayout app_bwd_tc_layout(int NP) {
  AppBwdTcLayout L;
  L.np = NP;
  L.rows_pt = NP > 128 ? NP : 128;
  L.qcols = (NP + 31) & ~31;
  int need = L.qcols + 64, cols = 32;
  while (cols < need) cols <<= 1;
  L.tmem_cols = cols;
  size_t off = 0;
  L.wbi = off; off += (size_t)2 * NP * 32;                        // Wb image  [NP rows][K 32] hi, lo
  L.g = off; off += (size_t)2 * 128 * 32;                          // G         [128 rows][K 32] hi, lo (rows 64.. zero)
  L.gt = off; off += (size_t)2 * 2 * 32 * BT;                      // G^T       [32 rows][K 64] hi, lo, two tile buffers
  L.pt = off; off += (size_t)2 * (L.rows_pt / 8) * BT_PT_SBO_F;    // P^T       [rows_pt][K 64] hi, lo, skewed
  L.q = off; off += (size_t)BT * (NP + 4);                         // GQ fp32   [64][NP + 4]
  L.tap = off; off += (size_t)2 * BT * 3 * 4;                      // taps      [64][3] float4, two tile buffers
  L.gf = off; off += (size_t)BT * 32;                              // next tile's g_feat block (fp32, prefetched)
  L.bars = off; off += 8;
  L.total = off * sizeof(float);
  return L;
}

// CP = true: TensorCP (models/tensoRF.py:388-415).  The C components are cut into NB blocks of NP = C / NB (<= 144, what
// the operand tiles of one CTA hold); CTA j works on block j % NB for the tiles j / NB, j / NB + grid / NB, ... — its basis
// image, its product operand and its dWb accumulator only ever see that block, so nothing is reloaded per tile.  The gather /
// scatter pass multiplies three line samples instead of a plane and a line sample and merges runs on three lines.
template <bool CP>
__global__ void __launch_bounds__(BT_THREADS, 1)
app_bwd_vm_tc_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int NB) {
  extern __shared__ __align__(1024) float bsm[];
  const FtrfFactorSet& a = f.app;
  const int NPT = app_np(f), NP = NPT / NB, cb0 = CP ? ((int)blockIdx.x % NB) * NP : 0;     // all components; this CTA's block
  const int NPs = NP + 4, D = f.app_dim, n_kg = NP / 4;
  const int64_t tile0 = (int64_t)(blockIdx.x / NB) * BT, tile_step = (int64_t)(gridDim.x / NB) * BT;
  const AppBwdTcLayout L = app_bwd_tc_layout(NP);
  float* WbH = bsm + L.wbi; float* WbL = WbH + (size_t)NP * 32;
  float* GH = bsm + L.g;    float* GL = GH + 128 * 32;
  float* GT0 = bsm + L.gt;                                         // buffer b: hi at GT0 + b * 2 * 32 * BT, lo 32 * BT after it
  const int pt_half = (L.rows_pt / 8) * BT_PT_SBO_F;
  float* PTH = bsm + L.pt;  float* PTL = PTH + pt_half;
  float* Q = bsm + L.q;
  float4* TAP0 = reinterpret_cast<float4*>(bsm + L.tap);
  float* GF = bsm + L.gf;                                          // [64 * D] contiguous copy of the tile's g_feat rows
  uint64_t* bars = reinterpret_cast<uint64_t*>(bsm + L.bars);      // [0] GQ batch done, [1] dWb batch done
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if (tile0 >= n) return;

  // resident operands: Wb image (row = component, k = output o, zero for o >= D); zero G rows 64..127 and P^T
  for (int e = tid; e < NP * 8; e += BT_THREADS) {
    const int nrow = e >> 3, ku = e & 7;
    float h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = ku * 4 + j;
      tc::split_tf32(o < D ? __ldg(f.basis + (size_t)o * NPT + cb0 + nrow) : 0.f, h[j], l[j]);
    }
    const int off = tc::canon_off(nrow, ku, 8);
    *reinterpret_cast<float4*>(WbH + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(WbL + off) = make_float4(l[0], l[1], l[2], l[3]);
  }
  for (int e = tid * 4; e < 2 * 128 * 32; e += BT_THREADS * 4) *reinterpret_cast<float4*>(GH + e) = z4();
  for (int e = tid * 4; e < 2 * pt_half; e += BT_THREADS * 4) *reinterpret_cast<float4*>(PTH + e) = z4();
  if (tid == 0) { tc::mbar_init(bars, 1); tc::mbar_init(bars + 1, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc(tslot, (uint32_t)L.tmem_cols);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tslot;
  const uint32_t idesc_q = tc::idesc_tf32(128, NP), idesc_w = tc::idesc_tf32(128, 32);
  const uint32_t tq = tmem, tw0 = tmem + (uint32_t)L.qcols, tw1 = tw0 + 32u;
  const bool two_m = NP > 128;
  uint32_t ph_q = 0, ph_w = 0;
  bool first = true;
  int gtb = 0;
  // taps of one tile (slot -> ray -> position: three dependent global round trips) — computed one tile ahead by
  // the two extra warps while the others run the gather / scatter pass
  auto make_taps = [&](int64_t tbase, float4* dst, int r) {
    const int64_t st = tbase + r;
    float4 tp[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) tp[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (st < n) {
      float xn[3];
      app_point(f, c, ws, st, xn);
      Taps3 T;
      if (CP) {                                           // line j samples axis vecm(j)
#pragma unroll
        for (int j = 0; j < 3; ++j) T.a[j] = make_tap(xn[vecm(j)], a.line_len[j]);
      } else {
        T = make_taps3(a, xn);
      }
#pragma unroll
      for (int j = 0; j < 3; ++j) tp[j] = make_float4(__int_as_float(T.a[j].i0), __int_as_float(T.a[j].i1), T.a[j].w0, T.a[j].w1);
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) dst[r * 3 + j] = tp[j];
    // the tile's g_feat rows are one contiguous block: coalesced copy into shared memory
    const int64_t lim = min((int64_t)BT, n - tbase) * D;
    for (int e = r; e < BT * D; e += 64) GF[e] = e < lim ? __ldg(ws.a_gfeat + tbase * D + e) : 0.f;
  };
  if (tid >= BT_WORK) make_taps(tile0, TAP0, tid - BT_WORK);
  __syncthreads();                        // the first tile's staging reads the g_feat block

  for (int64_t t0 = tile0; t0 < n; t0 += tile_step) {
    const int tile_n = (int)min((int64_t)BT, n - t0);
    // ---- stage g_feat as both operands, and the taps.  The previous tile's dWb products may still be running:
    // they read the other G^T buffer and P^T, which is not written before the wait below.
    float* GTH = GT0 + (size_t)gtb * (2 * 32 * BT);           // gtb = parity of this tile: selects its G^T and tap buffers
    float* GTL = GTH + 32 * BT;
    for (int e = tid; e < BT * 8; e += BT_THREADS) {
      const int sl = e >> 3, ku = e & 7;
      float h[4], l[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int o = ku * 4 + j;
        const float g = o < D ? GF[sl * D + o] : 0.f;          // (rows past the end were stored as zeros)
        tc::split_tf32(g, h[j], l[j]);
        const int offt = tc::canon_off(o, sl >> 2, BT / 4) + (sl & 3);
        GTH[offt] = h[j];
        GTL[offt] = l[j];
      }
      const int off = tc::canon_off(sl, ku, 8);
      *reinterpret_cast<float4*>(GH + off) = make_float4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<float4*>(GL + off) = make_float4(l[0], l[1], l[2], l[3]);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    // ---- GQ = G Wb
    if (tid == BT_ISSUER) {
      tc::fence_after_sync();
      const uint32_t hi = tc::desc_hi(8u * 128u);
      const uint32_t ah = tc::desc_lo(tc::smem_u32(GH), 128), al = tc::desc_lo(tc::smem_u32(GL), 128);
      const uint32_t bh = tc::desc_lo(tc::smem_u32(WbH), 128), bl = tc::desc_lo(tc::smem_u32(WbL), 128);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t o = (uint32_t)ks * 16u;
        tc::mma_tf32_ss_lo(tq, al + o, bh + o, hi, idesc_q, ks ? 1u : 0u);
        tc::mma_tf32_ss_lo(tq, ah + o, bl + o, hi, idesc_q, 1u);
        tc::mma_tf32_ss_lo(tq, ah + o, bh + o, hi, idesc_q, 1u);
      }
      tc::mma_commit(bars);
    }
    tc::mbar_wait(bars, ph_q);
    ph_q ^= 1;
    if (!first) {                         // issued before this tile's GQ batch by the same thread: complete as well
      tc::mbar_wait(bars + 1, ph_w);
      ph_w ^= 1;
    }
    tc::fence_after_sync();
    if ((warp & 3) < 2) {                 // rows 0..63 live in TMEM lanes 0..63: warps 0,1 (mod 4) can read them
      const int q = warp & 3, row = q * 32 + lane;
      for (int cc = warp >> 2; cc * 32 < NP; cc += (BT_THREADS / 32 + 3) / 4) {
        float acc[32];
        tc::tmem_ld32(tq + ((uint32_t)(q * 32) << 16) + (uint32_t)(cc * 32), acc);
        tc::tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          if (cc * 32 + j < NP)
            *reinterpret_cast<float4*>(Q + (size_t)row * NPs + cc * 32 + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
      }
    }
    tc::fence_before_sync();
    __syncthreads();
    // ---- gather once: products into the dWb operand, plane / line gradients with run merging
    const float4* TAP = TAP0 + (size_t)gtb * (BT * 3);
    if (tid >= BT_WORK) {
      const int64_t tn = t0 + tile_step;
      if (tn < n) make_taps(tn, TAP0 + (size_t)(gtb ^ 1) * (BT * 3), tid - BT_WORK);
    }
    for (int item = tid; item < (BT / BT_RUN) * n_kg && tid < BT_WORK; item += BT_WORK) {
      const int seg = item / n_kg, q = item - seg * n_kg;
      float pq[BT_RUN][4];
      if (CP) {
        const int ch = cb0 + 4 * q, C = NPT;              // component index in the lines
        float4 m1 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (a.mask[0]) m1 = mask4(a.mask[0], a.mask_stride[0], ch);          // tensoRF.py:415 (once)
        const float* ln0 = a.line[0] + ch;
        const float* ln1 = a.line[1] + ch;
        const float* ln2 = a.line[2] + ch;
        LineRun lr[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) lr[j].kl = -1;
#pragma unroll
        for (int u = 0; u < BT_RUN; ++u) {
          const int i = seg * BT_RUN + u;
          pq[u][0] = pq[u][1] = pq[u][2] = pq[u][3] = 0.f;
          if (i >= tile_n) continue;
          const float4 t0 = TAP[i * 3 + 0], t1 = TAP[i * 3 + 1], t2 = TAP[i * 3 + 2];
          const int a0 = __float_as_int(t0.x), a1 = __float_as_int(t0.y), b0 = __float_as_int(t1.x), b1 = __float_as_int(t1.y);
          const int c0 = __float_as_int(t2.x), c1 = __float_as_int(t2.y);
          const float4 e00 = ld4(ln0 + (size_t)a0 * C), e01 = ld4(ln0 + (size_t)a1 * C);
          const float4 e10 = ld4(ln1 + (size_t)b0 * C), e11 = ld4(ln1 + (size_t)b1 * C);
          const float4 e20 = ld4(ln2 + (size_t)c0 * C), e21 = ld4(ln2 + (size_t)c1 * C);
          float4 l0 = z4(), l1 = z4(), l2 = z4();
          l0 = fma4(t0.z, e00, l0); l0 = fma4(t0.w, e01, l0);
          l1 = fma4(t1.z, e10, l1); l1 = fma4(t1.w, e11, l1);
          l2 = fma4(t2.z, e20, l2); l2 = fma4(t2.w, e21, l2);
          const float4 prod = mul4(mul4(mul4(l0, l1), l2), m1);
          pq[u][0] = prod.x; pq[u][1] = prod.y; pq[u][2] = prod.z; pq[u][3] = prod.w;
          const float4 gq = mul4(*reinterpret_cast<const float4*>(Q + (size_t)i * NPs + 4 * q), m1);
          const float4 gk[3] = {mul4(gq, mul4(l1, l2)), mul4(gq, mul4(l0, l2)), mul4(gq, mul4(l0, l1))};
          const int key[3] = {a0 | (a1 << 16), b0 | (b1 << 16), c0 | (c1 << 16)};
          const float w0[3] = {t0.z, t1.z, t2.z}, w1[3] = {t0.w, t1.w, t2.w};
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            if (key[j] != lr[j].kl) {
              flush_line(lr[j], a.g_line[j], C, ch);
              lr[j].kl = key[j];
              lr[j].a0 = z4(); lr[j].a1 = z4();
            }
            lr[j].a0 = fma4(w0[j], gk[j], lr[j].a0);
            lr[j].a1 = fma4(w1[j], gk[j], lr[j].a1);
          }
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) flush_line(lr[j], a.g_line[j], C, ch);
      } else {
        int k = 0, base = 0;
        while (k < 2 && 4 * q >= base + a.n_comp[k]) { base += a.n_comp[k]; ++k; }
        const int ch = 4 * q - base, C = a.n_comp[k], W = a.plane_w[k];
        const int m0 = mat0(k), m1 = mat1(k), mv = vecm(k);
        float4 m2 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (a.mask[k]) { const float4 m = mask4(a.mask[k], a.mask_stride[k], ch); m2 = mul4(m, m); }
        const float* pl = a.plane[k] + ch;
        const float* ln = a.line[k] + ch;
        PlaneRun pr; LineRun lr;
        pr.kx = -1; pr.ky = -1; lr.kl = -1;
  #pragma unroll
        for (int u = 0; u < BT_RUN; ++u) {
          const int i = seg * BT_RUN + u;
          pq[u][0] = pq[u][1] = pq[u][2] = pq[u][3] = 0.f;
          if (i >= tile_n) continue;
          const float4 ax = TAP[i * 3 + m0], ay = TAP[i * 3 + m1], al = TAP[i * 3 + mv];
          const int x0 = __float_as_int(ax.x), x1 = __float_as_int(ax.y), y0 = __float_as_int(ay.x), y1 = __float_as_int(ay.y);
          const int l0 = __float_as_int(al.x), l1 = __float_as_int(al.y);
          const float* r0 = pl + ((size_t)y0 * W) * C;
          const float* r1 = pl + ((size_t)y1 * W) * C;
          const float4 c00 = ld4(r0 + (size_t)x0 * C), c10 = ld4(r0 + (size_t)x1 * C);
          const float4 c01 = ld4(r1 + (size_t)x0 * C), c11 = ld4(r1 + (size_t)x1 * C);
          const float4 e0 = ld4(ln + (size_t)l0 * C), e1 = ld4(ln + (size_t)l1 * C);
          const float w00 = ax.z * ay.z, w10 = ax.w * ay.z, w01 = ax.z * ay.w, w11 = ax.w * ay.w;
          float4 pv = z4(), lv = z4();
          pv = fma4(w00, c00, pv); pv = fma4(w10, c10, pv); pv = fma4(w01, c01, pv); pv = fma4(w11, c11, pv);
          lv = fma4(al.z, e0, lv); lv = fma4(al.w, e1, lv);
          const float4 prod = mul4(mul4(pv, lv), m2);
          pq[u][0] = prod.x; pq[u][1] = prod.y; pq[u][2] = prod.z; pq[u][3] = prod.w;
          const float4 gq = mul4(*reinterpret_cast<const float4*>(Q + (size_t)i * NPs + 4 * q), m2);
          const float4 gp = mul4(gq, lv), gl = mul4(gq, pv);
          const int kx = x0 | (x1 << 16), ky = y0 | (y1 << 16), kl = l0 | (l1 << 16);
          if (kx != pr.kx || ky != pr.ky) {
            flush_plane(pr, a.g_plane[k], W, C, ch);
            pr.kx = kx; pr.ky = ky;
            pr.a00 = z4(); pr.a10 = z4(); pr.a01 = z4(); pr.a11 = z4();
          }
          pr.a00 = fma4(w00, gp, pr.a00);
          pr.a10 = fma4(w10, gp, pr.a10);
          pr.a01 = fma4(w01, gp, pr.a01);
          pr.a11 = fma4(w11, gp, pr.a11);
          if (kl != lr.kl) {
            flush_line(lr, a.g_line[k], C, ch);
            lr.kl = kl;
            lr.a0 = z4(); lr.a1 = z4();
          }
          lr.a0 = fma4(al.z, gl, lr.a0);
          lr.a1 = fma4(al.w, gl, lr.a1);
        }
        flush_plane(pr, a.g_plane[k], W, C, ch);
        flush_line(lr, a.g_line[k], C, ch);
      }
      // rows 4q..4q+3 (components), K unit `seg` (samples 4 seg .. 4 seg + 3)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int row = 4 * q + j;
        float h[4], l[4];
#pragma unroll
        for (int u = 0; u < BT_RUN; ++u) tc::split_tf32(pq[u][j], h[u], l[u]);
        const int off = (row >> 3) * BT_PT_SBO_F + seg * 32 + (row & 7) * 4;
        *reinterpret_cast<float4*>(PTH + off) = make_float4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<float4*>(PTL + off) = make_float4(l[0], l[1], l[2], l[3]);
      }
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    // ---- dWb^T += P^T G
    if (tid == BT_ISSUER) {
      tc::fence_after_sync();
      const uint32_t hi_a = tc::desc_hi((uint32_t)BT_PT_SBO_F * 4u), hi_b = tc::desc_hi((uint32_t)(BT / 4) * 128u);
      const uint32_t bh = tc::desc_lo(tc::smem_u32(GTH), 128), bl = tc::desc_lo(tc::smem_u32(GTL), 128);
      for (int mt = 0; mt < (two_m ? 2 : 1); ++mt) {
        const int row0 = mt ? NP - 128 : 0;
        const uint32_t ah = tc::desc_lo(tc::smem_u32(PTH + (row0 >> 3) * BT_PT_SBO_F), 128);
        const uint32_t al = tc::desc_lo(tc::smem_u32(PTL + (row0 >> 3) * BT_PT_SBO_F), 128);
        const uint32_t td = mt ? tw1 : tw0;
#pragma unroll
        for (int ks = 0; ks < BT / 8; ++ks) {
          const uint32_t o = (uint32_t)ks * 16u;
          tc::mma_tf32_ss_lo2(td, al + o, hi_a, bh + o, hi_b, idesc_w, (first && ks == 0) ? 0u : 1u);
          tc::mma_tf32_ss_lo2(td, ah + o, hi_a, bl + o, hi_b, idesc_w, 1u);
          tc::mma_tf32_ss_lo2(td, ah + o, hi_a, bh + o, hi_b, idesc_w, 1u);
        }
      }
      tc::mma_commit(bars + 1);
    }
    first = false;
    gtb ^= 1;
  }
  // ---- flush the basis gradient: TMEM row = component, column = output
  tc::mbar_wait(bars + 1, ph_w);
  tc::fence_after_sync();
  if (warp < 8) {
    const int mt = warp >> 2, q = warp & 3;
    if (mt == 0 || two_m) {
      const int m = q * 32 + lane, comp = mt ? NP - 128 + m : m;
      float acc[32];
      tc::tmem_ld32((mt ? tw1 : tw0) + ((uint32_t)(q * 32) << 16), acc);
      tc::tmem_wait_ld();
      const bool mine = mt ? comp >= 128 : comp < NP;      // the second tile repeats rows NP-128 .. 127
      if (mine)
#pragma unroll
        for (int o = 0; o < 32; ++o)
          if (o < D && acc[o] != 0.f) atomicAdd(f.g_basis + (size_t)o * NPT + cb0 + comp, acc[o]);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, (uint32_t)L.tmem_cols);
}

int launch_appearance_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  int NPs = app_np(f) + 4, Dp = (f.app_dim + 3) & ~3;
  size_t smem = ((size_t)f.app_dim * NPs + (size_t)ATILE * NPs + (size_t)ATILE * Dp + (size_t)Dp * LDP + (size_t)ATILE * 3) * sizeof(float);
  if (int e = app_check(f, smem)) return e;
  cudaFuncSetAttribute(app_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  // run-merging scatter: (ATILE / run) * NP / 4 work items per tile; auto = the shortest run whose items fit
  // one pass of the CTA (48x3 components: run 10, 252 items)
  // TensorCP (288 components): 512 threads, so that the scatter's (segment, chunk) items still fit one pass
  const int threads = f.model == FTRF_MODEL_CP ? 512 : 256;
  int arun = g_app_bwd_runs;
  if (arun < 0) {
    arun = 4;
    while (arun < ATILE && ((ATILE + arun - 1) / arun) * (app_np(f) / 4) > threads) ++arun;
  }
  if (f.model == FTRF_MODEL_VM && g_app_bwd_tc && arun != 0) {
    const int NP = app_np(f);
    const AppBwdTcLayout L = app_bwd_tc_layout(NP);
    if (NP % 16 == 0 && NP >= 16 && NP <= 256 && f.app_dim <= 32 && L.total + 1024 <= 227 * 1024 && L.tmem_cols <= 512) {
      cudaFuncSetAttribute(app_bwd_vm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
      app_bwd_vm_tc_kernel<false><<<FTRF_SMS, BT_THREADS, L.total, st>>>(f, c, ws, 1);
      return check_launch("app_bwd_vm_tc");
    }
  }
  if (f.model == FTRF_MODEL_CP && g_app_bwd_tc && arun != 0) {
    // component blocks of at most 144 (the operand tiles of one CTA), as few as divide C into multiples of 16; the grid is
    // rounded down to a multiple of the block count (288 components: 2 blocks x 74 CTAs)
    const int C = app_np(f);
    int nb = 0;
    for (int t = 1; t <= 8 && !nb; ++t)
      if (C % t == 0 && (C / t) % 16 == 0 && C / t <= 144 && C / t >= 16) nb = t;
    if (nb && f.app_dim <= 32) {
      const AppBwdTcLayout L = app_bwd_tc_layout(C / nb);
      if (L.total + 1024 <= 227 * 1024 && L.tmem_cols <= 512) {
        cudaFuncSetAttribute(app_bwd_vm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
        app_bwd_vm_tc_kernel<true><<<(FTRF_SMS / nb) * nb, BT_THREADS, L.total, st>>>(f, c, ws, nb);
        return check_launch("app_bwd_cp_tc");
      }
    }
  }
  if (f.model == FTRF_MODEL_VM && arun > 0 && g_app_bwd_onegather) {
    const size_t sm2 = smem + ((size_t)ATILE * NPs + (size_t)ATILE * 3 * 4) * sizeof(float);
    cudaFuncSetAttribute(app_bwd_vm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
    app_bwd_vm_kernel<<<FTRF_SMS * 2, 256, sm2, st>>>(f, c, ws, arun);
    return check_launch("app_bwd_vm");
  }
  app_bwd_kernel<<<FTRF_SMS * 2, threads, smem, st>>>(f, c, ws, arun);
  return check_launch("app_bwd");
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_debug_set_app_tc(int on) {
  ftrf::g_app_fwd_tc = on;
  return 0;
}

extern "C" int ftrf_debug_set_app_ws(int on) {
  ftrf::g_app_fwd_ws = on;
  return 0;
}

extern "C" int ftrf_debug_set_app_bwd_tc(int on) {
  ftrf::g_app_bwd_tc = on;
  return 0;
}

extern "C" int ftrf_debug_set_app_runs(int runs) {
  if (runs == -2) { ftrf::g_app_bwd_onegather = 0; runs = -1; }
  else ftrf::g_app_bwd_onegather = 1;
  ftrf::g_app_bwd_runs = runs;
  return 0;
}
#endif
