// k_raygen.cu — camera rays generated on the device: a render call then needs a
// 3x4 pose and the intrinsics instead of 15 MB of host rays per 800x800 frame (renderer.py:18 copies every
// chunk host -> device).  Follows dataLoader/ray_utils.py:24-41 (get_ray_directions: pixel centre + 0.5,
// (i - cx) / fx, (j - cy) / fy, 1), :44-63 (the blender-convention variant: y and z negated),
// dataLoader/blender.py:59 (directions normalised once per dataset) and ray_utils.py:67-85 (get_rays:
// d_world = d_cam @ R^T, origin = c2w[:, 3] for every pixel).
#include "common.cuh"

namespace ftrf {

struct RayGenArgs {
  int W, row0, row_step;
  int64_t n;                 // rays to write = n_rows * W
  float fx, fy, cx, cy;
  float r[9], o[3];
  int normalize, blender;
  float* out;                // [n][6]
};

__global__ void __launch_bounds__(256)
raygen_kernel(const RayGenArgs a) {
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < a.n; t += (int64_t)gridDim.x * blockDim.x) {
    const int lr = (int)(t / a.W), col = (int)(t - (int64_t)lr * a.W);
    const int row = a.row0 + lr * a.row_step;
    const float i = (float)col + 0.5f, j = (float)row + 0.5f;
    float dx = __fdiv_rn(i - a.cx, a.fx), dy = __fdiv_rn(j - a.cy, a.fy), dz = 1.f;
    if (a.blender) { dy = -dy; dz = -1.f; }
    if (a.normalize) {
      const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
      dx = __fdiv_rn(dx, n); dy = __fdiv_rn(dy, n); dz = __fdiv_rn(dz, n);
    }
    float* o = a.out + t * 6;
    o[0] = a.o[0]; o[1] = a.o[1]; o[2] = a.o[2];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      o[3 + k] = __fadd_rn(__fadd_rn(__fmul_rn(dx, a.r[3 * k]), __fmul_rn(dy, a.r[3 * k + 1])), __fmul_rn(dz, a.r[3 * k + 2]));
  }
}

}  // namespace ftrf

using namespace ftrf;

extern "C" int ftrf_generate_rays(const FtrfCamera* cam, int32_t row0, int32_t row_step, int32_t n_rows, float* rays_out,
                                  void* stream) {
  if (!cam || !rays_out || cam->H <= 0 || cam->W <= 0 || row_step <= 0 || row0 < 0 || n_rows < 0 ||
      (n_rows > 0 && row0 + (int64_t)(n_rows - 1) * row_step >= cam->H) || cam->fx == 0.f || cam->fy == 0.f) {
    set_error("generate_rays: bad argument (H=%d W=%d rows %d:+%d x%d)", cam ? cam->H : -1, cam ? cam->W : -1, row0, row_step, n_rows);
    return FTRF_E_ARG;
  }
  if (n_rows == 0) return FTRF_OK;
  RayGenArgs a{};
  a.W = cam->W; a.row0 = row0; a.row_step = row_step; a.n = (int64_t)n_rows * cam->W;
  a.fx = cam->fx; a.fy = cam->fy; a.cx = cam->cx; a.cy = cam->cy;
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) a.r[3 * r + c] = cam->c2w[4 * r + c];
    a.o[r] = cam->c2w[4 * r + 3];
  }
  a.normalize = cam->normalize; a.blender = cam->blender; a.out = rays_out;
  int64_t blocks = (a.n + 255) / 256;
  const int64_t cap = (int64_t)FTRF_SMS * 8;
  raygen_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(a);
  return check_launch("generate_rays");
}
