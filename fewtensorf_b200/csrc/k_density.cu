// k_density.cu — stage 2a: VM / CP density gather + sigma activation on the packed sample list,
// and its scatter-add backward.  Replaces TensorVMSplit/TensorCP.compute_densityfeature
// (models/tensoRF.py:207-227, :358-386) and feature2density (models/tensorBase.py:291-295).
//
// Layout: factors are channel-last, so the C channels of one texel are one contiguous 4C-byte
// vector.  A group of GS=4 lanes owns one sample; lane `sub` handles float4 channel chunks
// sub, sub+4, ...  (C=16: one 16-byte load per lane per corner, the 4 lanes cover one 64-byte
// texel).  Consecutive packed samples of a ray are half a texel apart, so neighbouring groups hit
// the same or adjacent texels: the working set is L1/L2 resident.
// Algorithmic bytes per sample (VM, C=16x3): 3*16*(4+2)*4 = 1152 B forward;
// backward re-reads them and reduces 1152 B of gradient (2304 B).
#include "common.cuh"

namespace ftrf {

constexpr int GS = 4;   // lanes per sample
FTRF_SWITCH(int, g_density_bwd_runs, 12);  // longest segment a lane group of the run-merging backward walks; 0 = one reduction set per sample (ftrf_debug_set_runs)

__device__ __forceinline__ float group_sum(float v) {
  v += __shfl_xor_sync(FULL_MASK, v, 1);
  v += __shfl_xor_sync(FULL_MASK, v, 2);
  return v;
}

// VM density feature of this lane's channel chunks from ready-made taps
__device__ __forceinline__ float density_partial_vm(const FtrfField& f, const Taps3& T3, int sub) {
  const FtrfFactorSet& d = f.den;
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int C = d.n_comp[k];
    const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
    for (int c = sub * 4; c < C; c += GS * 4) {
      float4 p = plane4(d.plane[k], d.plane_w[k], C, tx, ty, c);
      float4 l = line4(d.line[k], C, tl, c);
      if (d.mask[k]) {                       // (plane*m)*(line*m), tensoRF.py:224-225
        float4 m = mask4(d.mask[k], d.mask_stride[k], c);
        p = mul4(p, m);
        l = mul4(l, m);
      }
      acc += sum4(mul4(p, l));
    }
  }
  return acc;
}

// The three taps of a sample, one axis per lane of its 4-lane group, exchanged by shuffles (every lane of
// the warp must call this).
__device__ __forceinline__ Taps3 taps3_group(const FtrfFactorSet& d, const float xn[3], int sub) {
  const int a = sub < 3 ? sub : 0;
  const int size = a == 0 ? d.plane_w[0] : (a == 1 ? d.plane_h[0] : d.line_len[0]);
  const Tap mine = make_tap(a == 0 ? xn[0] : (a == 1 ? xn[1] : xn[2]), size);
  const int base = lane_id() & ~(GS - 1);
  Taps3 T;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    T.a[k].i0 = __shfl_sync(FULL_MASK, mine.i0, base + k);
    T.a[k].i1 = __shfl_sync(FULL_MASK, mine.i1, base + k);
    T.a[k].w0 = __shfl_sync(FULL_MASK, mine.w0, base + k);
    T.a[k].w1 = __shfl_sync(FULL_MASK, mine.w1, base + k);
  }
  return T;
}

// Partial (this lane's channel chunks) density feature of one normalised point.
__device__ __forceinline__ float density_partial(const FtrfField& f, const float xn[3], int sub) {
  const FtrfFactorSet& d = f.den;
  float acc = 0.f;
  if (f.model == FTRF_MODEL_VM) {
    const Taps3 T3 = make_taps3(d, xn);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int C = d.n_comp[k];
      const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
      for (int c = sub * 4; c < C; c += GS * 4) {
        float4 p = plane4(d.plane[k], d.plane_w[k], C, tx, ty, c);
        float4 l = line4(d.line[k], C, tl, c);
        if (d.mask[k]) {                       // (plane*m)*(line*m), tensoRF.py:224-225
          float4 m = mask4(d.mask[k], d.mask_stride[k], c);
          p = mul4(p, m);
          l = mul4(l, m);
        }
        acc += sum4(mul4(p, l));
      }
    }
  } else {
    const int C = d.n_comp[0];
    Tap t0 = make_tap(xn[vecm(0)], d.line_len[0]);
    Tap t1 = make_tap(xn[vecm(1)], d.line_len[1]);
    Tap t2 = make_tap(xn[vecm(2)], d.line_len[2]);
    for (int c = sub * 4; c < C; c += GS * 4) {
      float4 v = mul4(mul4(line4(d.line[0], C, t0, c), line4(d.line[1], C, t1, c)), line4(d.line[2], C, t2, c));
      if (d.mask[0]) v = mul4(v, mask4(d.mask[0], d.mask_stride[0], c));     // tensoRF.py:384
      acc += sum4(v);
    }
  }
  return acc;
}

__device__ __forceinline__ float activate(const FtrfField& f, float feat) {
  return f.act == FTRF_ACT_SOFTPLUS ? softplus_f(feat + f.density_shift) : fmaxf(feat, 0.f);
}

__global__ void __launch_bounds__(256)
density_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  // a sample list that overflowed has unwritten holes (rays that did not fit): the host repeats the
  // call with a larger list, nothing may be read from this one
  if (ws.counters[2] & 1) return;
  const int64_t n = min((int64_t)ws.counters[0], c.cap_samples);
  const int sub = threadIdx.x & (GS - 1);
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  // all lanes of a warp run the same trip count so the group shuffles stay converged
  const int64_t n_pad = (n + 7) & ~(int64_t)7;
  // the list entry and the ray of the NEXT iteration are fetched while this one gathers: without that a
  // sample is three dependent global round trips (entry -> ray -> texels) with nothing else in flight
  const bool vm = f.model == FTRF_MODEL_VM;
  Ray ray_n = {};
  int idx_n = 0;
  if (g0 < n) { ray_n = load_ray_ws(f, c, ws, ws.s_ray[g0]); idx_n = ws.s_idx[g0]; }
  for (int64_t s = g0; s < n_pad; s += gstride) {
    const Ray ray = ray_n;
    const int idx = idx_n;
    const int64_t s2 = s + gstride;
    int r2 = 0;
    if (s2 < n) { r2 = ws.s_ray[s2]; idx_n = ws.s_idx[s2]; }
    float part = 0.f;
    float xn[3] = {0.f, 0.f, 0.f};
    if (s < n) {
      float p[3];
      sample_pt(ray, sample_z(f, c, ray, idx), p);
      normalize_pt(f, p, xn);
    }
    if (vm) {
      const Taps3 T = taps3_group(f.den, xn, sub);       // one axis per lane, shared by shuffles
      if (s < n) part = density_partial_vm(f, T, sub);
    } else if (s < n) {
      part = density_partial(f, xn, sub);
    }
    if (s2 < n) ray_n = load_ray_ws(f, c, ws, r2);
    float feat = group_sum(part);
    if (s < n && sub == 0) ws.s_sigma[s] = activate(f, feat);
  }
}

// sigma at arbitrary world-space points (compute_alpha's inner part, tensorBase.py:298-313)
__global__ void __launch_bounds__(256)
density_at_kernel(const FtrfField f, const float* __restrict__ xyz, int64_t n, float* __restrict__ out) {
  const int sub = threadIdx.x & (GS - 1);
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  const int64_t n_pad = (n + 7) & ~(int64_t)7;
  for (int64_t s = g0; s < n_pad; s += gstride) {
    float part = 0.f;
    bool keep = false;
    if (s < n) {
      float p[3] = {__ldg(xyz + 3 * s), __ldg(xyz + 3 * s + 1), __ldg(xyz + 3 * s + 2)}, xn[3];
      keep = f.has_alpha_mask ? alpha_mask_keep(f, p) : true;
      if (keep) {
        normalize_pt(f, p, xn);
        part = density_partial(f, xn, sub);
      }
    }
    float feat = group_sum(part);
    if (s < n && sub == 0) out[s] = keep ? activate(f, feat) : 0.f;
  }
}

// compute_densityfeature on points (models/tensoRF.py:207-227, :358-386): normalised coordinates in, the raw
// sigma feature out (no alphaMask, no activation) — the module method, on the gather code of the hot path
__global__ void __launch_bounds__(256)
density_feature_kernel(const FtrfField f, const float* __restrict__ xn_in, int64_t n, float* __restrict__ out) {
  const int sub = threadIdx.x & (GS - 1);
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  const int64_t n_pad = (n + 7) & ~(int64_t)7;
  for (int64_t s = g0; s < n_pad; s += gstride) {
    float part = 0.f;
    if (s < n) {
      const float xn[3] = {__ldg(xn_in + 3 * s), __ldg(xn_in + 3 * s + 1), __ldg(xn_in + 3 * s + 2)};
      part = density_partial(f, xn, sub);
    }
    const float feat = group_sum(part);
    if (s < n && sub == 0) out[s] = feat;
  }
}

// Occupancy lattice: getDenseAlpha + compute_alpha + the `>= alphaMask_thres` test of updateAlphaMask
// (tensorBase.py:215-230, :298-318, :245) without materialising the [X,Y,Z,3] point tensor or the
// float alpha lattice.  A warp owns one 32-texel word of the (Z,Y,X)-ordered bit field: four rounds of
// eight lattice points (four lanes per point, as in the sample kernels), each round's verdicts
// collected by one ballot.  Point k-coordinate = aabb0*(1-s) + aabb1*s with s = linspace(0,1,G)[i]
// (host table `lin` = the three linspaces back to back), every operation rounded separately as in
// torch.  alpha = clamp(1 - exp(-sigma*length), 0, 1); raw bit = alpha >= thres (the 3^3 max-pool
// that follows is a dilation of these bits because max(.) >= t <=> any(. >= t)).
__global__ void __launch_bounds__(256)
dense_alpha_bits_kernel(const FtrfField f, const float* __restrict__ lin, const int X, const int Y, const int Z,
                        const float length, const float thres, uint32_t* __restrict__ raw, float* __restrict__ alpha_out) {
  const int sub = threadIdx.x & (GS - 1), grp = lane_id() >> 2;
  const int xw = (X + 31) >> 5;
  const int64_t words = (int64_t)xw * Y * Z;
  const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t wstride = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = w0; w < words; w += wstride) {
    const int xb = (int)(w % xw) << 5;
    const int y = (int)((w / xw) % Y), z = (int)(w / ((int64_t)xw * Y));
    const float sy = __ldg(lin + X + y), sz = __ldg(lin + X + Y + z);
    uint32_t word = 0u;
#pragma unroll 1
    for (int j = 0; j < 4; ++j) {
      const int x = xb + j * 8 + grp;
      float part = 0.f;
      bool keep = false;
      if (x < X) {
        const float s[3] = {__ldg(lin + x), sy, sz};
        float p[3], xn[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
          p[k] = __fadd_rn(__fmul_rn(f.aabb[k], __fsub_rn(1.f, s[k])), __fmul_rn(f.aabb[3 + k], s[k]));
        keep = f.has_alpha_mask ? alpha_mask_keep(f, p) : true;
        if (keep) {
          normalize_pt(f, p, xn);
          part = density_partial(f, xn, sub);
        }
      }
      const float feat = group_sum(part);
      const float sigma = keep ? activate(f, feat) : 0.f;
      float a = __fsub_rn(1.f, expf(__fmul_rn(-sigma, length)));
      a = fminf(fmaxf(a, 0.f), 1.f);
      if (alpha_out && sub == 0 && x < X) alpha_out[((size_t)z * Y + y) * X + x] = a;
      const uint32_t b = __ballot_sync(FULL_MASK, sub == 0 && x < X && a >= thres);
#pragma unroll
      for (int g = 0; g < 8; ++g) word |= ((b >> (4 * g)) & 1u) << (j * 8 + g);
    }
    if (lane_id() == 0) raw[w] = word;
  }
}

// ------------------------------------------------------------------------------------ backward
// d sigma_feature arrives in ws.s_gsig.  Each lane scatters its own channel chunks with
// red.global.add.v4.f32 (one 16-byte reduction per corner per chunk).
__global__ void __launch_bounds__(256)
density_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  if (ws.counters[2] & 1) return;
  const int64_t n = min((int64_t)ws.counters[0], c.cap_samples);
  const int sub = threadIdx.x & (GS - 1);
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  const FtrfFactorSet& d = f.den;
  for (int64_t s = g0; s < n; s += gstride) {
    const float g = ws.s_gsig[s];
    if (g == 0.f) continue;
    Ray ray = load_ray_ws(f, c, ws, ws.s_ray[s]);
    float p[3], xn[3];
    sample_pt(ray, sample_z(f, c, ray, ws.s_idx[s]), p);
    normalize_pt(f, p, xn);
    if (f.model == FTRF_MODEL_VM) {
      const Taps3 T3 = make_taps3(d, xn);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int C = d.n_comp[k], W = d.plane_w[k];
        const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
        for (int ch = sub * 4; ch < C; ch += GS * 4) {
          float4 pv = plane4(d.plane[k], W, C, tx, ty, ch);
          float4 lv = line4(d.line[k], C, tl, ch);
          float4 gm = make_float4(g, g, g, g);
          if (d.mask[k]) {
            float4 m = mask4(d.mask[k], d.mask_stride[k], ch);
            gm = mul4(gm, mul4(m, m));
          }
          float4 gp = mul4(gm, lv);      // d/d plane sample
          float4 gl = mul4(gm, pv);      // d/d line sample
          float* gp0 = d.g_plane[k] + ((size_t)ty.i0 * W) * C + ch;
          float* gp1 = d.g_plane[k] + ((size_t)ty.i1 * W) * C + ch;
          float w;
          if ((w = tx.w0 * ty.w0) != 0.f) red4(gp0 + (size_t)tx.i0 * C, scale4(w, gp));
          if ((w = tx.w1 * ty.w0) != 0.f) red4(gp0 + (size_t)tx.i1 * C, scale4(w, gp));
          if ((w = tx.w0 * ty.w1) != 0.f) red4(gp1 + (size_t)tx.i0 * C, scale4(w, gp));
          if ((w = tx.w1 * ty.w1) != 0.f) red4(gp1 + (size_t)tx.i1 * C, scale4(w, gp));
          if (tl.w0 != 0.f) red4(d.g_line[k] + (size_t)tl.i0 * C + ch, scale4(tl.w0, gl));
          if (tl.w1 != 0.f) red4(d.g_line[k] + (size_t)tl.i1 * C + ch, scale4(tl.w1, gl));
        }
      }
    } else {
      const int C = d.n_comp[0];
      Tap t[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) t[k] = make_tap(xn[vecm(k)], d.line_len[k]);
      for (int ch = sub * 4; ch < C; ch += GS * 4) {
        float4 l0 = line4(d.line[0], C, t[0], ch), l1 = line4(d.line[1], C, t[1], ch),
               l2 = line4(d.line[2], C, t[2], ch);
        float4 gm = make_float4(g, g, g, g);
        if (d.mask[0]) gm = mul4(gm, mask4(d.mask[0], d.mask_stride[0], ch));
        float4 gk[3] = {mul4(gm, mul4(l1, l2)), mul4(gm, mul4(l0, l2)), mul4(gm, mul4(l0, l1))};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (t[k].w0 != 0.f) red4(d.g_line[k] + (size_t)t[k].i0 * C + ch, scale4(t[k].w0, gk[k]));
          if (t[k].w1 != 0.f) red4(d.g_line[k] + (size_t)t[k].i1 * C + ch, scale4(t[k].w1, gk[k]));
        }
      }
    }
  }
}

// VM backward with run merging.  A lane group walks up to 12 consecutive entries of the packed list (the
// samples of one ray are consecutive and half a texel apart), keeps the gradient of the plane cell
// and of the line cell it is in in registers, and issues the reductions only when the cell changes:
// ~2x fewer red.global operations than one set per sample (the kernel is bound by their issue rate).

__global__ void __launch_bounds__(256, 2)
density_bwd_vm_runs_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int DRUN_MAX) {
  if (ws.counters[2] & 1) return;
  const int64_t n = min((int64_t)ws.counters[0], c.cap_samples);
  const int sub = threadIdx.x & (GS - 1);
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  // every lane group gets the same number of equally long segments (the grid is one resident wave)
  const int64_t per = (n + gstride - 1) / gstride;
  const int64_t rounds = max((int64_t)1, (per + DRUN_MAX - 1) / DRUN_MAX);
  const int64_t DRUN = max((int64_t)1, (per + rounds - 1) / rounds);
  const int64_t nseg = (n + DRUN - 1) / DRUN;
  const FtrfFactorSet& d = f.den;
  const int cmax = max(d.n_comp[0], max(d.n_comp[1], d.n_comp[2]));
  for (int64_t seg = g0; seg < nseg; seg += gstride) {
    const int64_t s0 = seg * DRUN, s1 = min(n, s0 + DRUN);
    for (int ch = sub * 4; ch < cmax; ch += GS * 4) {
      PlaneRun pr[3];
      LineRun lr[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) { pr[k].kx = -1; pr[k].ky = -1; lr[k].kl = -1; }
      for (int64_t s = s0; s < s1; ++s) {
        const float g = ws.s_gsig[s];
        if (g == 0.f) continue;
        Ray ray = load_ray_ws(f, c, ws, ws.s_ray[s]);
        float p[3], xn[3];
        sample_pt(ray, sample_z(f, c, ray, ws.s_idx[s]), p);
        normalize_pt(f, p, xn);
        const Taps3 T3 = make_taps3(d, xn);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int C = d.n_comp[k], W = d.plane_w[k];
          if (ch >= C) continue;
          const Tap &tx = T3.a[mat0(k)], &ty = T3.a[mat1(k)], &tl = T3.a[vecm(k)];
          const float4 pv = plane4(d.plane[k], W, C, tx, ty, ch);
          const float4 lv = line4(d.line[k], C, tl, ch);
          float4 gm = make_float4(g, g, g, g);
          if (d.mask[k]) {
            const float4 m = mask4(d.mask[k], d.mask_stride[k], ch);
            gm = mul4(gm, mul4(m, m));
          }
          const float4 gp = mul4(gm, lv);      // d/d plane sample
          const float4 gl = mul4(gm, pv);      // d/d line sample
          const int kx = tx.i0 | (tx.i1 << 16), ky = ty.i0 | (ty.i1 << 16), kl = tl.i0 | (tl.i1 << 16);
          if (kx != pr[k].kx || ky != pr[k].ky) {
            flush_plane(pr[k], d.g_plane[k], W, C, ch);
            pr[k].kx = kx; pr[k].ky = ky;
            pr[k].a00 = z4(); pr[k].a10 = z4(); pr[k].a01 = z4(); pr[k].a11 = z4();
          }
          pr[k].a00 = fma4(tx.w0 * ty.w0, gp, pr[k].a00);
          pr[k].a10 = fma4(tx.w1 * ty.w0, gp, pr[k].a10);
          pr[k].a01 = fma4(tx.w0 * ty.w1, gp, pr[k].a01);
          pr[k].a11 = fma4(tx.w1 * ty.w1, gp, pr[k].a11);
          if (kl != lr[k].kl) {
            flush_line(lr[k], d.g_line[k], C, ch);
            lr[k].kl = kl;
            lr[k].a0 = z4(); lr[k].a1 = z4();
          }
          lr[k].a0 = fma4(tl.w0, gl, lr[k].a0);
          lr[k].a1 = fma4(tl.w1, gl, lr[k].a1);
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        flush_plane(pr[k], d.g_plane[k], d.plane_w[k], d.n_comp[k], ch);
        flush_line(lr[k], d.g_line[k], d.n_comp[k], ch);
      }
    }
  }
}

// Same scatter with the per-sample work hoisted: the four lanes of a group first compute the taps of the
// segment's samples (three samples each) into shared memory, then walk the segment once per factor pair
// with only that pair's accumulators live.  ~70 registers instead of 128: three CTAs per SM instead of two,
// and no lane recomputes what its neighbours compute.
#ifndef FTRF_DSEG
#define FTRF_DSEG 12
#endif
#ifndef FTRF_DBW_CTAS
#define FTRF_DBW_CTAS 2      // two CTAs of <= 128 registers (no spills) beat three of 85: 94 -> 85 us (scripts/variant_sweep.sh)
#endif
constexpr int DSEG = FTRF_DSEG;          // longest segment (samples) a lane group walks
constexpr int DBW_CTAS = FTRF_DBW_CTAS;  // resident CTAs per SM the grid is sized for

__global__ void __launch_bounds__(256, DBW_CTAS)
density_bwd_vm_staged_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ __align__(16) uint8_t dbw_smem[];
  float4 (*s_tap)[DSEG][3] = reinterpret_cast<float4 (*)[DSEG][3]>(dbw_smem);
  float (*s_g)[DSEG] = reinterpret_cast<float (*)[DSEG]>(dbw_smem + sizeof(float4) * (256 / GS) * DSEG * 3);
  if (ws.counters[2] & 1) return;
  const int64_t n = min((int64_t)ws.counters[0], c.cap_samples);
  const int sub = threadIdx.x & (GS - 1), grp = threadIdx.x / GS;
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  // every lane group gets the same number of equally long segments (the grid is one resident wave)
  const int64_t per = (n + gstride - 1) / gstride;
  const int64_t rounds = max((int64_t)1, (per + DSEG - 1) / DSEG);
  const int DRUN = (int)max((int64_t)1, (per + rounds - 1) / rounds);
  const int64_t nseg = (n + DRUN - 1) / DRUN;
  const FtrfFactorSet& d = f.den;
  const int cmax = max(d.n_comp[0], max(d.n_comp[1], d.n_comp[2]));
  const int64_t seg_pad = (nseg + 7) & ~(int64_t)7;        // whole warps take the same trips (__syncwarp below)
  for (int64_t seg = g0; seg < seg_pad; seg += gstride) {
    const int64_t s0 = seg * DRUN;
    const int len = seg < nseg ? (int)min((int64_t)DRUN, n - s0) : 0;
    __syncwarp();
    for (int j = sub; j < len; j += GS) {
      const int64_t s = s0 + j;
      const float g = ws.s_gsig[s];
      s_g[grp][j] = g;
      if (g != 0.f) {
        const Ray ray = load_ray_ws(f, c, ws, ws.s_ray[s]);
        float p[3], xn[3];
        sample_pt(ray, sample_z(f, c, ray, ws.s_idx[s]), p);
        normalize_pt(f, p, xn);
        const Taps3 T = make_taps3(d, xn);
#pragma unroll
        for (int a = 0; a < 3; ++a)
          s_tap[grp][j][a] = make_float4(__int_as_float(T.a[a].i0), __int_as_float(T.a[a].i1), T.a[a].w0, T.a[a].w1);
      }
    }
    __syncwarp();
    for (int ch = sub * 4; ch < cmax; ch += GS * 4) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int C = d.n_comp[k], W = d.plane_w[k];
        if (ch >= C) continue;
        float4 m2 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (d.mask[k]) { const float4 m = mask4(d.mask[k], d.mask_stride[k], ch); m2 = mul4(m, m); }
        const float* pl = d.plane[k] + ch;
        const float* ln = d.line[k] + ch;
        PlaneRun pr; LineRun lr;
        pr.kx = -1; pr.ky = -1; lr.kl = -1;
        for (int j = 0; j < len; ++j) {
          const float g = s_g[grp][j];
          if (g == 0.f) continue;
          const float4 ax = s_tap[grp][j][mat0(k)], ay = s_tap[grp][j][mat1(k)], al = s_tap[grp][j][vecm(k)];
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
          const float4 gm = scale4(g, m2);
          const float4 gp = mul4(gm, lv), gl = mul4(gm, pv);
          const int kx = x0 | (x1 << 16), ky = y0 | (y1 << 16), kl = l0 | (l1 << 16);
          if (kx != pr.kx || ky != pr.ky) {
            flush_plane(pr, d.g_plane[k], W, C, ch);
            pr.kx = kx; pr.ky = ky;
            pr.a00 = z4(); pr.a10 = z4(); pr.a01 = z4(); pr.a11 = z4();
          }
          pr.a00 = fma4(w00, gp, pr.a00);
          pr.a10 = fma4(w10, gp, pr.a10);
          pr.a01 = fma4(w01, gp, pr.a01);
          pr.a11 = fma4(w11, gp, pr.a11);
          if (kl != lr.kl) {
            flush_line(lr, d.g_line[k], C, ch);
            lr.kl = kl;
            lr.a0 = z4(); lr.a1 = z4();
          }
          lr.a0 = fma4(al.z, gl, lr.a0);
          lr.a1 = fma4(al.w, gl, lr.a1);
        }
        flush_plane(pr, d.g_plane[k], W, C, ch);
        flush_line(lr, d.g_line[k], C, ch);
      }
    }
  }
}

// TensorCP: the same staged walk over the three lines (models/tensoRF.py:358-386: feature = sum_c prod_k line_k).  Every
// sample of a step lands in the same few hundred line rows, so one reduction set per sample (density_bwd_kernel) is
// bound by same-address contention in L2; consecutive samples of a ray share their line cells on all three axes, and
// a lane group that keeps the running gradient of its current cell per line in registers issues a third of them.
__global__ void __launch_bounds__(256, DBW_CTAS)
density_bwd_cp_staged_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ __align__(16) uint8_t dbw_smem[];
  float4 (*s_tap)[DSEG][3] = reinterpret_cast<float4 (*)[DSEG][3]>(dbw_smem);
  float (*s_g)[DSEG] = reinterpret_cast<float (*)[DSEG]>(dbw_smem + sizeof(float4) * (256 / GS) * DSEG * 3);
  if (ws.counters[2] & 1) return;
  const int64_t n = min((int64_t)ws.counters[0], c.cap_samples);
  const int sub = threadIdx.x & (GS - 1), grp = threadIdx.x / GS;
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const int64_t gstride = (int64_t)gridDim.x * blockDim.x / GS;
  const int64_t per = (n + gstride - 1) / gstride;
  const int64_t rounds = max((int64_t)1, (per + DSEG - 1) / DSEG);
  const int DRUN = (int)max((int64_t)1, (per + rounds - 1) / rounds);
  const int64_t nseg = (n + DRUN - 1) / DRUN;
  const FtrfFactorSet& d = f.den;
  const int C = d.n_comp[0];
  const int64_t seg_pad = (nseg + 7) & ~(int64_t)7;        // whole warps take the same trips (__syncwarp below)
  for (int64_t seg = g0; seg < seg_pad; seg += gstride) {
    const int64_t s0 = seg * DRUN;
    const int len = seg < nseg ? (int)min((int64_t)DRUN, n - s0) : 0;
    __syncwarp();
    for (int j = sub; j < len; j += GS) {
      const int64_t s = s0 + j;
      const float g = ws.s_gsig[s];
      s_g[grp][j] = g;
      if (g != 0.f) {
        const Ray ray = load_ray_ws(f, c, ws, ws.s_ray[s]);
        float p[3], xn[3];
        sample_pt(ray, sample_z(f, c, ray, ws.s_idx[s]), p);
        normalize_pt(f, p, xn);
#pragma unroll
        for (int k = 0; k < 3; ++k) {                      // tap of line k (axis vecMode[k])
          const Tap t = make_tap(xn[vecm(k)], d.line_len[k]);
          s_tap[grp][j][k] = make_float4(__int_as_float(t.i0), __int_as_float(t.i1), t.w0, t.w1);
        }
      }
    }
    __syncwarp();
    for (int ch = sub * 4; ch < C; ch += GS * 4) {
      float4 m1 = make_float4(1.f, 1.f, 1.f, 1.f);
      if (d.mask[0]) m1 = mask4(d.mask[0], d.mask_stride[0], ch);          // tensoRF.py:384 (once, not squared)
      LineRun lr[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) lr[k].kl = -1;
      for (int j = 0; j < len; ++j) {
        const float g = s_g[grp][j];
        if (g == 0.f) continue;
        float4 t[3], lv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          t[k] = s_tap[grp][j][k];
          const float* ln = d.line[k] + ch;
          lv[k] = fma4(t[k].w, ld4(ln + (size_t)__float_as_int(t[k].y) * C), scale4(t[k].z, ld4(ln + (size_t)__float_as_int(t[k].x) * C)));
        }
        const float4 gm = scale4(g, m1);
        const float4 gk[3] = {mul4(gm, mul4(lv[1], lv[2])), mul4(gm, mul4(lv[0], lv[2])), mul4(gm, mul4(lv[0], lv[1]))};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int kl = __float_as_int(t[k].x) | (__float_as_int(t[k].y) << 16);
          if (kl != lr[k].kl) {
            flush_line(lr[k], d.g_line[k], C, ch);
            lr[k].kl = kl;
            lr[k].a0 = z4(); lr[k].a1 = z4();
          }
          lr[k].a0 = fma4(t[k].z, gk[k], lr[k].a0);
          lr[k].a1 = fma4(t[k].w, gk[k], lr[k].a1);
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) flush_line(lr[k], d.g_line[k], C, ch);
    }
  }
}

static int sample_grid(int64_t n_hint) {
  int64_t blocks = (n_hint * GS + 255) / 256;
  int64_t cap = (int64_t)FTRF_SMS * 8;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

int launch_density(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  // the sample count lives on the device: persistent grid sized for the SM count, grid-stride inside
  density_fwd_kernel<<<FTRF_SMS * 8, 256, 0, st>>>(f, c, ws);
  return check_launch("density_fwd");
}

int launch_density_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (f.model == FTRF_MODEL_VM && g_density_bwd_runs == 12) {
    constexpr size_t smem = (sizeof(float4) * 3 + sizeof(float)) * (256 / GS) * DSEG;
    cudaFuncSetAttribute(density_bwd_vm_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    density_bwd_vm_staged_kernel<<<FTRF_SMS * DBW_CTAS, 256, smem, st>>>(f, c, ws);
  }
  else if (f.model == FTRF_MODEL_VM && g_density_bwd_runs) density_bwd_vm_runs_kernel<<<FTRF_SMS * 2, 256, 0, st>>>(f, c, ws, g_density_bwd_runs);
  else if (f.model == FTRF_MODEL_CP && g_density_bwd_runs) {
    constexpr size_t smem = (sizeof(float4) * 3 + sizeof(float)) * (256 / GS) * DSEG;
    cudaFuncSetAttribute(density_bwd_cp_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    density_bwd_cp_staged_kernel<<<FTRF_SMS * DBW_CTAS, 256, smem, st>>>(f, c, ws);
  }
  else density_bwd_kernel<<<FTRF_SMS * 8, 256, 0, st>>>(f, c, ws);
  return check_launch("density_bwd");
}

int launch_dense_alpha_bits(const FtrfField& f, const float* lin, int X, int Y, int Z, float length, float thres,
                            uint32_t* raw, float* alpha_out, cudaStream_t st) {
  const int64_t words = (int64_t)((X + 31) / 32) * Y * Z;
  const int64_t blocks = (words + 7) / 8;                       // 8 warps (words) per CTA
  const int64_t cap = (int64_t)FTRF_SMS * 8 * 4;
  dense_alpha_bits_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(f, lin, X, Y, Z, length, thres, raw, alpha_out);
  return check_launch("dense_alpha_bits");
}

int launch_density_feature(const FtrfField& f, const float* xn, int64_t n, float* out, cudaStream_t st) {
  if (n == 0) return FTRF_OK;
  density_feature_kernel<<<sample_grid(n), 256, 0, st>>>(f, xn, n, out);
  return check_launch("density_feature");
}

int launch_density_at(const FtrfField& f, const float* xyz, int64_t n, float* out, cudaStream_t st) {
  if (n == 0) return FTRF_OK;
  density_at_kernel<<<sample_grid(n), 256, 0, st>>>(f, xyz, n, out);
  return check_launch("density_at");
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_debug_set_runs(int density_runs) {
  ftrf::g_density_bwd_runs = density_runs;
  return 0;
}
#endif
