// common.cuh — shared device helpers and the carved workspace view.
// sm_100a only.  Arithmetic that decides WHICH samples exist (stage 1) uses explicitly
// rounded intrinsics (__fmul_rn/__fadd_rn/...) so nvcc cannot contract it into FMAs: the
// reference evaluates those expressions as separate torch ops (models/tensorBase.py:193-206).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/fewtensorf.h"

#define FTRF_SMS 148

// Bring-up / cross-check switches exist only in the test build (libfewtensorf_b200_test.so, -DFTRF_TEST_BUILD);
// the product library compiles them to constants and carries no process-global mutable state for them.
#ifdef FTRF_TEST_BUILD
#define FTRF_SWITCH(type, name, value) type name = value
#else
#define FTRF_SWITCH(type, name, value) static constexpr type name = value
#endif
#define FULL_MASK 0xffffffffu

namespace ftrf {

// Typed view of the workspace buffer (see FTRF_WS_* in fewtensorf.h).
struct Ws {
  int32_t* counters;
  float* tmin;
  int32_t* ray_off;
  int32_t* ray_cnt;
  float* acc;
  float* rgb_pre;
  int32_t* s_ray;
  int32_t* s_idx;
  float* s_sigma;
  float* s_w;
  float* s_T;
  int32_t* s_app;
  float* s_gsig;
  int32_t* a_slot;
  float* a_feat;
  float* a_rgb;
  float* a_h1;
  float* a_h2;
  float* a_grgb;
  float* a_gh1;
  float* a_gh2;
  float* a_gfeat;
  float* a_x;      // [cap_app][ldx]  MLP input rows (tensor-core path)
  float* a_gx;     // [cap_app][ldx]  dL/dX
  float* wimg;     // weight images of the fused MLP kernels (fused.cuh)
  int ldx;
  // points-in calls (ftrf_app_feature / ftrf_shade_points): the shaded "list" is the caller's point array itself
  const float* direct_xn;     // [n][3] normalised coordinates, or nullptr (per-ray path: slot -> ray -> position)
  const float* direct_view;   // [n][3] view directions
};

// library-owned hand-off bits of FtrfCall::lib_flags
enum : uint64_t {
  CALL_IMG_FWD_PENDING = 1,   // forward weight image being built on the side stream, join before the fused forward
  CALL_IMG_BWD_READY = 2,     // this call's forward built the dgrad weight image
  CALL_SCRATCH_READY = 4,     // ... and zeroed the wgrad scratch
  CALL_TAIL_PENDING = 8,      // un-permute of dW1 running on the side stream
  CALL_SCAN_PENDING = 16,     // reverse scan of the integration backward running on the side stream
  CALL_FORKED = 32,           // phase 0 of shade_prep_ahead recorded its fork event
  CALL_MASKS_READY = 64,      // the fused forward left the ReLU sign masks of h1 / h2 (32 B per shaded sample) in ws.a_gx
};
struct CallState { uint64_t flags; void* ctx; };
inline CallState& call_state(const FtrfCall& c) {
  return *reinterpret_cast<CallState*>(const_cast<uint64_t*>(&c.lib_flags));
}

size_t carve(const FtrfField& f, const FtrfCall& c, int64_t* off, void* base, Ws* ws);
void set_error(const char* fmt, ...);
int check_launch(const char* what);
void prof_begin(int stage, cudaStream_t st);   // no-ops unless ftrf_profile_enable(1)
void prof_end(int stage, cudaStream_t st);

// launchers (one per stage; defined in the k_*.cu files)
int launch_march(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_density(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_integrate(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_appearance(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_shade(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_composite(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_integrate_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_grgb(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_integrate_bwd_sigma(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_shade_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t, int part = 0);
int launch_appearance_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_density_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_density_at(const FtrfField&, const float* xyz, int64_t n, float* out, cudaStream_t);
int launch_density_feature(const FtrfField&, const float* xn, int64_t n, float* out, cudaStream_t);
bool prof_enabled();
int launch_dense_alpha_bits(const FtrfField&, const float* lin, int X, int Y, int Z, float length, float thres,
                            uint32_t* raw, float* alpha_out, cudaStream_t);

#ifdef __CUDACC__

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// ---------------------------------------------------------------- stage-1 exact arithmetic
struct Ray {
  float o[3], d[3];
  float tmin;
  float u;      // per-ray jitter (train, non-ndc) else 0
};

// t_min of models/tensorBase.py:193-196 — IEEE division, min, max over xyz, clamp(near, far).
__device__ __forceinline__ float ray_tmin(const FtrfField& f, const float o[3], const float d[3]) {
  float t = -INFINITY;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float v = (d[k] == 0.f) ? 1e-6f : d[k];
    float ra = __fdiv_rn(__fsub_rn(f.aabb[3 + k], o[k]), v);
    float rb = __fdiv_rn(__fsub_rn(f.aabb[k], o[k]), v);
    t = fmaxf(t, fminf(ra, rb));
  }
  return fminf(fmaxf(t, f.near_), f.far_);
}

// z of sample i: t_min + stepSize * (i [+ u])   (tensorBase.py:198-203); ndc: host z table (:181-183).
__device__ __forceinline__ float sample_z(const FtrfField& f, const FtrfCall& c, const Ray& r, int i) {
  if (c.ndc) return __ldg(c.jitter + i);
  float rng = (float)i;
  if (c.is_train) rng = __fadd_rn(rng, r.u);
  return __fadd_rn(r.tmin, __fmul_rn(f.step_size, rng));
}

__device__ __forceinline__ void sample_pt(const Ray& r, float z, float p[3]) {
#pragma unroll
  for (int k = 0; k < 3; ++k) p[k] = __fadd_rn(r.o[k], __fmul_rn(r.d[k], z));   // tensorBase.py:205
}

__device__ __forceinline__ bool in_aabb(const FtrfField& f, const float p[3]) {        // tensorBase.py:206
  bool out = false;
#pragma unroll
  for (int k = 0; k < 3; ++k) out |= (f.aabb[k] > p[k]) | (p[k] > f.aabb[3 + k]);
  return !out;
}

// normalize_coord, tensorBase.py:130-131: (x - aabb0) * invaabbSize - 1
__device__ __forceinline__ void normalize_pt(const FtrfField& f, const float p[3], float xn[3]) {
#pragma unroll
  for (int k = 0; k < 3; ++k)
    xn[k] = __fsub_rn(__fmul_rn(__fsub_rn(p[k], f.aabb[k]), f.inv_aabb[k]), 1.f);
}

// grid_sample(align_corners=True) pixel coordinate: ((c + 1) / 2) * (size - 1)
__device__ __forceinline__ float unnorm(float c, int size) {
  return __fmul_rn(__fmul_rn(__fadd_rn(c, 1.f), 0.5f), (float)(size - 1));
}

__device__ __forceinline__ Ray load_ray(const FtrfField& f, const FtrfCall& c, int64_t r) {
  Ray ray;
  const float* p = c.rays + r * c.ray_cols;
#pragma unroll
  for (int k = 0; k < 3; ++k) { ray.o[k] = __ldg(p + k); ray.d[k] = __ldg(p + 3 + k); }
  ray.tmin = c.ndc ? 0.f : ray_tmin(f, ray.o, ray.d);
  ray.u = (c.is_train && !c.ndc) ? __ldg(c.jitter + r) : 0.f;
  return ray;
}

// Same ray with t_min read back from the workspace (the march kernel stored the value it computed, so this
// is bit-identical and saves the six IEEE divisions per sample in every kernel downstream of stage 1).
__device__ __forceinline__ Ray load_ray_ws(const FtrfField& f, const FtrfCall& c, const Ws& ws, int64_t r) {
  Ray ray;
  const float* p = c.rays + r * c.ray_cols;
#pragma unroll
  for (int k = 0; k < 3; ++k) { ray.o[k] = __ldg(p + k); ray.d[k] = __ldg(p + 3 + k); }
  ray.tmin = c.ndc ? 0.f : ws.tmin[r];
  ray.u = (c.is_train && !c.ndc) ? __ldg(c.jitter + r) : 0.f;
  return ray;
}

// alphaMask keep decision — equivalent to `grid_sample(volume, c, align_corners=True) > 0`
// (tensorBase.py:41-48, :350-351) for a {0,1} volume: some in-bounds corner holds 1 with all three
// 1-D weights > 0.  The lower-corner weight (x0+1-fx) is always > 0; the upper one (fx-x0) is 0
// exactly when fx is an integer.
__device__ __forceinline__ bool am_bit(const FtrfField& f, int x, int y, int z) {
  int xw = (f.am_dim[0] + 31) >> 5;
  uint32_t w = __ldg(f.am_bits + ((size_t)z * f.am_dim[1] + y) * xw + (x >> 5));
  return (w >> (x & 31)) & 1u;
}

__device__ __forceinline__ bool alpha_mask_keep(const FtrfField& f, const float p[3]) {
  float fx[3];
  int i0[3];
  bool up[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float c = __fsub_rn(__fmul_rn(__fsub_rn(p[k], f.am_aabb0[k]), f.am_inv[k]), 1.f);
    fx[k] = unnorm(c, f.am_dim[k]);
    float fl = floorf(fx[k]);
    // guard the int conversion; anything this far out has no in-bounds corner
    if (!(fl >= -2.f && fl <= 1e9f)) return false;
    i0[k] = (int)fl;
    up[k] = __fsub_rn(fx[k], fl) > 0.f;
  }
  if (f.am_coarse && i0[0] >= 0 && i0[1] >= 0 && i0[2] >= 0 && i0[0] < f.am_dim[0] && i0[1] < f.am_dim[1] && i0[2] < f.am_dim[2]) {
    // both interpolation corners along every axis lie in [8c, 8c+8]: covered by the coarse bit of block c
    const int CX = (f.am_dim[0] + 7) >> 3, CY = (f.am_dim[1] + 7) >> 3, cxw = (CX + 31) >> 5;
    const int cx = i0[0] >> 3, cy = i0[1] >> 3, cz = i0[2] >> 3;
    uint32_t w = __ldg(f.am_coarse + ((size_t)cz * CY + cy) * cxw + (cx >> 5));
    if (!((w >> (cx & 31)) & 1u)) return false;
  }
#pragma unroll
  for (int dz = 0; dz < 2; ++dz) {
    int z = i0[2] + dz;
    if (z < 0 || z >= f.am_dim[2] || (dz && !up[2])) continue;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      int y = i0[1] + dy;
      if (y < 0 || y >= f.am_dim[1] || (dy && !up[1])) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        int x = i0[0] + dx;
        if (x < 0 || x >= f.am_dim[0] || (dx && !up[0])) continue;
        if (am_bit(f, x, y, z)) return true;
      }
    }
  }
  return false;
}

// ---------------------------------------------------------------- interpolation taps
// 1-D linear taps of grid_sample(align_corners=True, padding zeros) along an axis of `size` texels.
struct Tap {
  int i0, i1;     // clamped indices (always addressable)
  float w0, w1;   // weights, zeroed when the tap is out of bounds
};

__device__ __forceinline__ Tap make_tap(float c, int size) {
  float fx = unnorm(c, size);
  float fl = floorf(fx);
  Tap t;
  float w1 = fx - fl, w0 = (fl + 1.f) - fx;
  fl = fminf(fmaxf(fl, -2.f), (float)size);     // keeps the int conversion defined
  int i = (int)fl;
  bool ok0 = (i >= 0) & (i < size), ok1 = (i + 1 >= 0) & (i + 1 < size);
  t.i0 = min(max(i, 0), size - 1);
  t.i1 = min(max(i + 1, 0), size - 1);
  t.w0 = ok0 ? w0 : 0.f;
  t.w1 = ok1 ? w1 : 0.f;
  return t;
}

// The three planes and three lines of a VM factor set sample each axis at ONE resolution (plane k is
// gridSize[mat1(k)] x gridSize[mat0(k)], line k is gridSize[vecm(k)]; tensoRF.py:186-198; checked by
// the ABI), so a sample has three distinct taps, not nine.
struct Taps3 { Tap a[3]; };
__device__ __forceinline__ Taps3 make_taps3(const FtrfFactorSet& s, const float xn[3]) {
  Taps3 t;
  t.a[0] = make_tap(xn[0], s.plane_w[0]);
  t.a[1] = make_tap(xn[1], s.plane_h[0]);
  t.a[2] = make_tap(xn[2], s.line_len[0]);
  return t;
}

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float4 fma4(float w, float4 v, float4 a) {
  a.x = fmaf(w, v.x, a.x); a.y = fmaf(w, v.y, a.y); a.z = fmaf(w, v.z, a.z); a.w = fmaf(w, v.w, a.w);
  return a;
}
__device__ __forceinline__ float4 mul4(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 scale4(float s, float4 a) { return make_float4(s * a.x, s * a.y, s * a.z, s * a.w); }
__device__ __forceinline__ float sum4(float4 a) { return (a.x + a.y) + (a.z + a.w); }

// decomposition-mask value for 4 consecutive channels (stride 0 = scalar)
__device__ __forceinline__ float4 mask4(const float* m, int stride, int c) {
  if (stride == 0) { float v = __ldg(m); return make_float4(v, v, v, v); }
  return ld4(m + c);
}

// bilinear plane sample for 4 channels starting at c (channel-last [H][W][C])
__device__ __forceinline__ float4 plane4(const float* pl, int W, int C, const Tap& tx, const Tap& ty, int c) {
  const float* r0 = pl + ((size_t)ty.i0 * W) * C + c;
  const float* r1 = pl + ((size_t)ty.i1 * W) * C + c;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  a = fma4(tx.w0 * ty.w0, ld4(r0 + (size_t)tx.i0 * C), a);
  a = fma4(tx.w1 * ty.w0, ld4(r0 + (size_t)tx.i1 * C), a);
  a = fma4(tx.w0 * ty.w1, ld4(r1 + (size_t)tx.i0 * C), a);
  a = fma4(tx.w1 * ty.w1, ld4(r1 + (size_t)tx.i1 * C), a);
  return a;
}

__device__ __forceinline__ float4 line4(const float* ln, int C, const Tap& t, int c) {
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  a = fma4(t.w0, ld4(ln + (size_t)t.i0 * C + c), a);
  a = fma4(t.w1, ld4(ln + (size_t)t.i1 * C + c), a);
  return a;
}

// vectorised reduction into global memory (sm_90+: red.global.add.v4.f32)
__device__ __forceinline__ void red4(float* p, float4 v) {
  // no "memory" clobber: the gradient buffers are never read back inside a kernel, and the clobber would
  // pin the gathers of the next sample behind the reductions of this one
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
}

// run merging of the gather backward kernels: gradient of the plane / line cell a lane is walking through,
// reduced into global memory when the cell changes
struct PlaneRun { int kx, ky; float4 a00, a10, a01, a11; };   // k = i0 | i1 << 16 of the taps
struct LineRun { int kl; float4 a0, a1; };

__device__ __forceinline__ float4 z4() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void flush_plane(const PlaneRun& r, float* gp, int W, int C, int ch) {
  if (r.kx < 0) return;
  const int x0 = r.kx & 0xffff, x1 = r.kx >> 16, y0 = r.ky & 0xffff, y1 = r.ky >> 16;
  float* r0 = gp + ((size_t)y0 * W) * C + ch;
  float* r1 = gp + ((size_t)y1 * W) * C + ch;
  red4(r0 + (size_t)x0 * C, r.a00);
  red4(r0 + (size_t)x1 * C, r.a10);
  red4(r1 + (size_t)x0 * C, r.a01);
  red4(r1 + (size_t)x1 * C, r.a11);
}
__device__ __forceinline__ void flush_line(const LineRun& r, float* gl, int C, int ch) {
  if (r.kl < 0) return;
  red4(gl + (size_t)(r.kl & 0xffff) * C + ch, r.a0);
  red4(gl + (size_t)(r.kl >> 16) * C + ch, r.a1);
}

__device__ __forceinline__ float softplus_f(float x) {          // F.softplus, beta 1, threshold 20
  return x > 20.f ? x : log1pf(expf(x));
}

// matMode = [[0,1],[0,2],[1,2]], vecMode = [2,1,0]   (tensorBase.py:60-61)
__host__ __device__ __forceinline__ int mat0(int k) { return k == 2 ? 1 : 0; }
__host__ __device__ __forceinline__ int mat1(int k) { return k == 0 ? 1 : 2; }
__host__ __device__ __forceinline__ int vecm(int k) { return 2 - k; }

#endif  // __CUDACC__
}  // namespace ftrf
