// k_shade_sh.cu — SH degree-2 shading (SHRender + eval_sh_bases(2,.), models/mlp.py:15-19, models/sh.py:87-112),
// forward and backward, and the argument check shared by the MLP shading back ends.
#include "shade_common.cuh"

namespace ftrf {

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

// Width limits of the shading MLP kernels (the tcgen05 back ends split a wide first layer, k_shade_tc.cu).
int shade_check(const FtrfField& f, const InSpec& sp) {
  if (f.feature_c != FC) { set_error("shade: featureC must be %d (got %d)", FC, f.feature_c); return FTRF_E_ARG; }
  if (sp.in_dim != f.in_dim) { set_error("shade: in_dim %d does not match the shading mode (%d)", f.in_dim, sp.in_dim); return FTRF_E_ARG; }
  if (sp.D + 6 > HS) { set_error("shade: unsupported input width %d", sp.in_dim); return FTRF_E_ARG; }
  return FTRF_OK;
}

int launch_shade_sh(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (f.app_dim != 27) { set_error("SH shading needs app_dim 27"); return FTRF_E_ARG; }
  shade_sh_fwd_kernel<<<FTRF_SMS * 4, 256, 0, st>>>(f, c, ws);
  return check_launch("shade_sh_fwd");
}

int launch_shade_sh_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  shade_sh_bwd_kernel<<<FTRF_SMS * 4, 256, 0, st>>>(f, c, ws);
  return check_launch("shade_sh_bwd");
}

}  // namespace ftrf
