// shade_common.cuh — MLP input specification (concat order + positional encoding + FreeNeRF
// encoding masks) shared by the forward and backward shading kernels, and the CUDA-core tile GEMM.
//   MLP_Fea  models/mlp.py:41-69   [fea, view, PE(fea)*m_fea, PE(view)*m_view]
//   MLP_PE   models/mlp.py:88-107  [fea, view, PE(pos)*m_pos, PE(view)*m_view]
//   MLP      models/mlp.py:127-155 [fea, view, PE(pos)*m_pos, PE(view)*m_view, PE(fea)*m_fea]
//   PE       models/mlp.py:8-13    channel-major / frequency-minor, [sin | cos]
#pragma once
#include "common.cuh"

namespace ftrf {

constexpr int ST = 64;        // shaded samples per CTA tile
constexpr int FC = 128;       // featureC
constexpr int HS = FC + 4;    // row stride of hidden-activation tiles in shared memory

enum { SRC_POS = 0, SRC_VIEW = 1, SRC_FEA = 2 };

struct PeBlock {
  int src, F, C, col0;
};

struct InSpec {
  int D;          // app_dim
  int nblk;
  PeBlock b[3];
  int in_dim;     // true MLP input width
  int kp;         // in_dim rounded up to a multiple of 4 (padding columns are zero)
  int need_pos;
};

inline InSpec make_spec(const FtrfField& f) {
  InSpec s{};
  s.D = f.app_dim;
  int col = f.app_dim + 3;
  auto add = [&](int src, int F, int C) {
    if (F <= 0) return;
    s.b[s.nblk++] = PeBlock{src, F, C, col};
    col += 2 * F * C;
    if (src == SRC_POS) s.need_pos = 1;
  };
  if (f.shading == FTRF_SHADE_MLP_FEA) { add(SRC_FEA, f.fea_pe, f.app_dim); add(SRC_VIEW, f.view_pe, 3); }
  else if (f.shading == FTRF_SHADE_MLP_PE) { add(SRC_POS, f.pos_pe, 3); add(SRC_VIEW, f.view_pe, 3); }
  else if (f.shading == FTRF_SHADE_MLP) { add(SRC_POS, f.pos_pe, 3); add(SRC_VIEW, f.view_pe, 3); add(SRC_FEA, f.fea_pe, f.app_dim); }
  s.in_dim = col;
  s.kp = (col + 3) & ~3;
  return s;
}

#ifdef __CUDACC__

// view direction fed to the shading module: raw d, or d/|d| for ndc rays (tensorBase.py:337-347)
__device__ __forceinline__ void shade_geometry(const FtrfField& f, const FtrfCall& c, const Ws& ws, int64_t s,
                                               float view[3], float pos[3]) {
  if (ws.direct_xn) {          // points-in call (ftrf_shade_points): the caller's arrays are the list
#pragma unroll
    for (int k = 0; k < 3; ++k) { pos[k] = __ldg(ws.direct_xn + 3 * s + k); view[k] = __ldg(ws.direct_view + 3 * s + k); }
    return;
  }
  int slot = ws.a_slot[s];
  Ray ray = load_ray_ws(f, c, ws, ws.s_ray[slot]);
  float p[3];
  sample_pt(ray, sample_z(f, c, ray, ws.s_idx[slot]), p);
  normalize_pt(f, p, pos);
  float inv = 1.f;
  if (c.ndc) inv = sqrtf(ray.d[0] * ray.d[0] + ray.d[1] * ray.d[1] + ray.d[2] * ray.d[2]);
#pragma unroll
  for (int k = 0; k < 3; ++k) view[k] = c.ndc ? ray.d[k] / inv : ray.d[k];
}

__device__ __forceinline__ float enc_mask_at(const FtrfField& f, int src, int e) {
  const float* m = f.enc_mask[src];
  if (!m) return 1.f;
  return f.enc_mask_len[src] == 1 ? __ldg(m) : __ldg(m + e);
}

// value of MLP input column `col` for one sample
__device__ __forceinline__ float in_col(const FtrfField& f, const InSpec& sp, int col, const float* feat,
                                        const float* view, const float* pos) {
  if (col < sp.D) return feat[col];
  if (col < sp.D + 3) return view[col - sp.D];
  for (int b = 0; b < sp.nblk; ++b) {
    const PeBlock& pb = sp.b[b];
    int e = col - pb.col0, half = pb.F * pb.C;
    if (e < 2 * half) {
      bool is_cos = e >= half;
      int q = is_cos ? e - half : e;
      int ch = q / pb.F, fq = q - ch * pb.F;
      const float* src = pb.src == SRC_POS ? pos : (pb.src == SRC_VIEW ? view : feat);
      float arg = src[ch] * (float)(1 << fq);
      float v = is_cos ? cosf(arg) : sinf(arg);
      return v * enc_mask_at(f, pb.src, e);
    }
  }
  return 0.f;   // padding column
}

// acc[4][8] += A[64][K] (row-major, lda) x B[K][128] (k-major, ldb).  Thread (ty,tx) of a 16x16
// layout owns samples ty*4..+3 and outputs {tx*4..+3} U {64+tx*4..+3}.  K % 4 == 0.
template <bool B_GLOBAL>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                                          int K, float acc[4][8], int ty, int tx) {
  const float* a0 = A + (size_t)(ty * 4) * lda;
  for (int k = 0; k < K; k += 4) {
    float a[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float4 v = *reinterpret_cast<const float4*>(a0 + (size_t)i * lda + k);
      a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float* brow = B + (size_t)(k + kk) * ldb;
      float4 b0, b1;
      if (B_GLOBAL) { b0 = ld4(brow + tx * 4); b1 = ld4(brow + 64 + tx * 4); }
      else { b0 = *reinterpret_cast<const float4*>(brow + tx * 4); b1 = *reinterpret_cast<const float4*>(brow + 64 + tx * 4); }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float av = a[i][kk];
        acc[i][0] = fmaf(av, b0.x, acc[i][0]); acc[i][1] = fmaf(av, b0.y, acc[i][1]);
        acc[i][2] = fmaf(av, b0.z, acc[i][2]); acc[i][3] = fmaf(av, b0.w, acc[i][3]);
        acc[i][4] = fmaf(av, b1.x, acc[i][4]); acc[i][5] = fmaf(av, b1.y, acc[i][5]);
        acc[i][6] = fmaf(av, b1.z, acc[i][6]); acc[i][7] = fmaf(av, b1.w, acc[i][7]);
      }
    }
  }
}

// Stage feat/view/pos of a tile into shared memory and expand the MLP input rows X[ST][lda].
// `stage` needs ST*(D+6) floats.  Must be called by all 256 threads; ends with __syncthreads().
__device__ __forceinline__ void build_x_tile(const FtrfField& f, const FtrfCall& c, const Ws& ws, const InSpec& sp,
                                             int64_t t0, int64_t n, float* stage, float* X, int lda) {
  const int D = sp.D;
  float* s_feat = stage;
  float* s_view = stage + ST * D;
  float* s_pos = s_view + ST * 3;
  for (int e = threadIdx.x; e < ST * D; e += blockDim.x) {
    int64_t s = t0 + e / D;
    s_feat[e] = s < n ? ws.a_feat[s * D + (e % D)] : 0.f;
  }
  if (threadIdx.x < ST) {
    int64_t s = t0 + threadIdx.x;
    float v[3] = {0.f, 0.f, 0.f}, p[3] = {0.f, 0.f, 0.f};
    if (s < n) shade_geometry(f, c, ws, s, v, p);
#pragma unroll
    for (int k = 0; k < 3; ++k) { s_view[threadIdx.x * 3 + k] = v[k]; s_pos[threadIdx.x * 3 + k] = p[k]; }
  }
  __syncthreads();
  const int kp = sp.kp;
  for (int e = threadIdx.x; e < ST * kp; e += blockDim.x) {
    int s = e / kp, col = e - s * kp;
    X[(size_t)s * lda + col] = in_col(f, sp, col, s_feat + s * D, s_view + s * 3, s_pos + s * 3);
  }
  __syncthreads();
}

#endif
}  // namespace ftrf
