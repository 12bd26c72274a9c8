// k_shade.cu — stage 3 forward: positional encoding (+ FreeNeRF encoding mask) -> 3-layer MLP ->
// sigmoid, or SH degree-2 shading.  Replaces MLPRender_Fea / MLPRender_PE / MLPRender
// (models/mlp.py:27-155) and SHRender + eval_sh_bases(2,.) (models/mlp.py:15-19, models/sh.py:87-112).
//
// v1 data path (fp32 CUDA cores, exact fp32 products so the 1e-4 parity budget holds): persistent
// CTAs, both weight matrices resident in shared memory (k-major), 64-sample tiles, 4x8 register
// tiles per thread.  FLOPs per shaded sample: 2*(in*128 + 128*128 + 128*3) = 71 936 (MLP_Fea).
#include "shade_common.cuh"

namespace ftrf {

__global__ void __launch_bounds__(256, 1)
shade_mlp_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp) {
  extern __shared__ __align__(16) float smem[];
  const int KP = sp.kp, KA = KP + 4;
  float* W1T = smem;                         // [KP][128]
  float* W2T = W1T + (size_t)KP * FC;        // [128][128]
  float* W3s = W2T + FC * FC;                // [3][128]
  float* b1s = W3s + 3 * FC;
  float* b2s = b1s + FC;
  float* b3s = b2s + FC;                     // [4]
  float* X = b3s + 4;                        // [ST][KA]  (later: H2 [ST][HS])
  float* H = X + (size_t)ST * (KA > HS ? KA : HS);   // [ST][HS]  (first: staging of feat/view/pos)

  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if ((int64_t)blockIdx.x * ST >= n) return;

  for (int e = threadIdx.x; e < KP * FC; e += blockDim.x) {
    int o = e % FC, k = e / FC;
    W1T[e] = k < sp.in_dim ? __ldg(f.w1 + (size_t)o * sp.in_dim + k) : 0.f;
  }
  for (int e = threadIdx.x; e < FC * FC; e += blockDim.x) {
    int o = e % FC, k = e / FC;
    W2T[e] = __ldg(f.w2 + (size_t)o * FC + k);
  }
  for (int e = threadIdx.x; e < 3 * FC; e += blockDim.x) W3s[e] = __ldg(f.w3 + e);
  if (threadIdx.x < FC) { b1s[threadIdx.x] = __ldg(f.b1 + threadIdx.x); b2s[threadIdx.x] = __ldg(f.b2 + threadIdx.x); }
  if (threadIdx.x < 3) b3s[threadIdx.x] = __ldg(f.b3 + threadIdx.x);

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int64_t t0 = (int64_t)blockIdx.x * ST; t0 < n; t0 += (int64_t)gridDim.x * ST) {
    __syncthreads();
    build_x_tile(f, c, ws, sp, t0, n, H, X, KA);

    float acc[4][8];
    // ---- layer 1
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = b1s[(j < 4 ? 0 : 60) + tx * 4 + j];
    tile_gemm<false>(X, KA, W1T, FC, KP, acc, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int s = ty * 4 + i;
      float4 v0 = make_float4(fmaxf(acc[i][0], 0.f), fmaxf(acc[i][1], 0.f), fmaxf(acc[i][2], 0.f), fmaxf(acc[i][3], 0.f));
      float4 v1 = make_float4(fmaxf(acc[i][4], 0.f), fmaxf(acc[i][5], 0.f), fmaxf(acc[i][6], 0.f), fmaxf(acc[i][7], 0.f));
      *reinterpret_cast<float4*>(H + s * HS + tx * 4) = v0;
      *reinterpret_cast<float4*>(H + s * HS + 64 + tx * 4) = v1;
      if (c.save_for_backward && t0 + s < n) {
        float* g = ws.a_h1 + (size_t)(t0 + s) * FC;
        *reinterpret_cast<float4*>(g + tx * 4) = v0;
        *reinterpret_cast<float4*>(g + 64 + tx * 4) = v1;
      }
    }
    __syncthreads();
    // ---- layer 2 (output tile overwrites X)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = b2s[(j < 4 ? 0 : 60) + tx * 4 + j];
    tile_gemm<false>(H, HS, W2T, FC, FC, acc, ty, tx);
    float* H2 = X;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int s = ty * 4 + i;
      float4 v0 = make_float4(fmaxf(acc[i][0], 0.f), fmaxf(acc[i][1], 0.f), fmaxf(acc[i][2], 0.f), fmaxf(acc[i][3], 0.f));
      float4 v1 = make_float4(fmaxf(acc[i][4], 0.f), fmaxf(acc[i][5], 0.f), fmaxf(acc[i][6], 0.f), fmaxf(acc[i][7], 0.f));
      *reinterpret_cast<float4*>(H2 + s * HS + tx * 4) = v0;
      *reinterpret_cast<float4*>(H2 + s * HS + 64 + tx * 4) = v1;
      if (c.save_for_backward && t0 + s < n) {
        float* g = ws.a_h2 + (size_t)(t0 + s) * FC;
        *reinterpret_cast<float4*>(g + tx * 4) = v0;
        *reinterpret_cast<float4*>(g + 64 + tx * 4) = v1;
      }
    }
    __syncthreads();
    // ---- layer 3 + sigmoid
    if (threadIdx.x < 3 * ST) {
      int s = threadIdx.x & (ST - 1), ch = threadIdx.x / ST;
      float a = b3s[ch];
      const float* h = H2 + s * HS;
      const float* w = W3s + ch * FC;
      for (int k = 0; k < FC; k += 4) {
        float4 hv = *reinterpret_cast<const float4*>(h + k);
        float4 wv = *reinterpret_cast<const float4*>(w + k);
        a = fmaf(hv.x, wv.x, a); a = fmaf(hv.y, wv.y, a); a = fmaf(hv.z, wv.z, a); a = fmaf(hv.w, wv.w, a);
      }
      if (t0 + s < n) ws.a_rgb[(size_t)(t0 + s) * 3 + ch] = 1.f / (1.f + expf(-a));
    }
  }
}

// ------------------------------------------------------------------------------------------ SH
__device__ __forceinline__ void sh_basis2(const float v[3], float b[9]) {   // models/sh.py:98-112
  const float x = v[0], y = v[1], z = v[2];
  const float xx = x * x, yy = y * y, zz = z * z;
  b[0] = 0.28209479177387814f;
  b[1] = -0.4886025119029199f * y;
  b[2] = 0.4886025119029199f * z;
  b[3] = -0.4886025119029199f * x;
  b[4] = 1.0925484305920792f * (x * y);
  b[5] = -1.0925484305920792f * (y * z);
  b[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
  b[7] = -1.0925484305920792f * (x * z);
  b[8] = 0.5462742152960396f * (xx - yy);
}

__global__ void __launch_bounds__(256)
shade_sh_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (int64_t)gridDim.x * blockDim.x) {
    float view[3], pos[3], b[9];
    shade_geometry(f, c, ws, s, view, pos);
    sh_basis2(view, b);
    const float* ft = ws.a_feat + s * f.app_dim;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      float a = 0.f;
#pragma unroll
      for (int j = 0; j < 9; ++j) a = fmaf(b[j], ft[ch * 9 + j], a);
      ws.a_rgb[s * 3 + ch] = fmaxf(a + 0.5f, 0.f);          // models/mlp.py:18
    }
  }
}

__global__ void __launch_bounds__(256)
shade_sh_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (int64_t)gridDim.x * blockDim.x) {
    float view[3], pos[3], b[9];
    shade_geometry(f, c, ws, s, view, pos);
    sh_basis2(view, b);
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      float g = ws.a_rgb[s * 3 + ch] > 0.f ? ws.a_grgb[s * 3 + ch] : 0.f;
#pragma unroll
      for (int j = 0; j < 9; ++j) ws.a_gfeat[s * f.app_dim + ch * 9 + j] = g * b[j];
    }
  }
}

size_t shade_fwd_smem(const InSpec& sp) {
  int ka = sp.kp + 4 > HS ? sp.kp + 4 : HS;
  size_t fl = (size_t)sp.kp * FC + FC * FC + 3 * FC + 2 * FC + 4 + (size_t)ST * ka + (size_t)ST * HS;
  return fl * sizeof(float);
}

// cuda_core: the fp32 CUDA-core kernels keep W1 and W2 whole in shared memory, which bounds the input width;
// the tcgen05 back ends split a wide first layer instead (k_shade_tc.cu).
int shade_check(const FtrfField& f, const InSpec& sp, bool cuda_core) {
  if (f.feature_c != FC) { set_error("shade: featureC must be %d (got %d)", FC, f.feature_c); return FTRF_E_ARG; }
  if (sp.in_dim != f.in_dim) { set_error("shade: in_dim %d does not match the shading mode (%d)", f.in_dim, sp.in_dim); return FTRF_E_ARG; }
  if (sp.D + 6 > HS) { set_error("shade: unsupported input width %d", sp.in_dim); return FTRF_E_ARG; }
  if (cuda_core && shade_fwd_smem(sp) > 227 * 1024) {
    set_error("shade: input width %d needs too much shared memory on the CUDA-core back end (use the tensor-core one)", sp.in_dim);
    return FTRF_E_ARG;
  }
  return FTRF_OK;
}

int launch_shade_sh(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (f.app_dim != 27) { set_error("SH shading needs app_dim 27"); return FTRF_E_ARG; }
  shade_sh_fwd_kernel<<<FTRF_SMS * 4, 256, 0, st>>>(f, c, ws);
  return check_launch("shade_sh_fwd");
}

int launch_shade_v1(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  InSpec sp = make_spec(f);
  if (int e = shade_check(f, sp, true)) return e;
  size_t smem = shade_fwd_smem(sp);
  cudaFuncSetAttribute(shade_mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  shade_mlp_fwd_kernel<<<FTRF_SMS, 256, smem, st>>>(f, c, ws, sp);
  return check_launch("shade_mlp_fwd");
}

int launch_shade_sh_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  shade_sh_bwd_kernel<<<FTRF_SMS * 4, 256, 0, st>>>(f, c, ws);
  return check_launch("shade_sh_bwd");
}

}  // namespace ftrf
