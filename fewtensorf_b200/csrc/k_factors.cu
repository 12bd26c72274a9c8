// k_factors.cu — factor maintenance between training stages (the half that is not a
// regulariser on planes): everything works on the channel-last factor layout of the hot path ([H][W][C] planes,
// [L][C] lines), so upsample / shrink produce parameters the gather kernels read without a re-layout pass.
//   ftrf_resample_factor   F.interpolate(mode='bilinear', align_corners=True) of one plane or line
//                          (models/tensoRF.py:267-288 upsample_volume_grid, :418-435 for TensorCP)
//   ftrf_crop_factor       the index crop of shrink (models/tensoRF.py:290-327, :437-466)
//   ftrf_ortho_reg         vector_comp_diffs (models/tensoRF.py:175-187): mean |off-diagonal| of each line's C x C Gram
//                          matrix, value and gradient
#include "common.cuh"

namespace ftrf {

// dst[oy][ox][:] = bilinear(src) with ATen's align_corners arithmetic: scale = (in - 1) / (out - 1) (0 when out == 1),
// real = scale * o, i0 = floor(real) clamped, lambda1 = real - i0, neighbour offset 0 at the last texel.
__global__ void __launch_bounds__(256)
resample_kernel(const float* __restrict__ src, int C, int H, int W, float* __restrict__ dst, int H2, int W2, float sy, float sx) {
  const int cq = C >> 2;
  const int64_t total = (int64_t)H2 * W2 * cq;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int q = (int)(e % cq);
    const int64_t pix = e / cq;
    const int ox = (int)(pix % W2), oy = (int)(pix / W2);
    const float ry = sy * (float)oy, rx = sx * (float)ox;
    const int y0 = min((int)ry, H - 1), x0 = min((int)rx, W - 1);
    const int y1 = y0 + (y0 < H - 1 ? 1 : 0), x1 = x0 + (x0 < W - 1 ? 1 : 0);
    const float ly1 = ry - (float)y0, lx1 = rx - (float)x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
    const float4 v00 = ld4(src + ((size_t)y0 * W + x0) * C + 4 * q), v01 = ld4(src + ((size_t)y0 * W + x1) * C + 4 * q);
    const float4 v10 = ld4(src + ((size_t)y1 * W + x0) * C + 4 * q), v11 = ld4(src + ((size_t)y1 * W + x1) * C + 4 * q);
    float4 o;
    o.x = ly0 * (lx0 * v00.x + lx1 * v01.x) + ly1 * (lx0 * v10.x + lx1 * v11.x);
    o.y = ly0 * (lx0 * v00.y + lx1 * v01.y) + ly1 * (lx0 * v10.y + lx1 * v11.y);
    o.z = ly0 * (lx0 * v00.z + lx1 * v01.z) + ly1 * (lx0 * v10.z + lx1 * v11.z);
    o.w = ly0 * (lx0 * v00.w + lx1 * v01.w) + ly1 * (lx0 * v10.w + lx1 * v11.w);
    *reinterpret_cast<float4*>(dst + (size_t)pix * C + 4 * q) = o;
  }
}

// One CTA per (row i of the Gram matrix, line).  Phase 1: thread j forms G[i][j] = sum_l M[i][l] M[j][l] (memory is
// [L][C]: the loads of a warp are one contiguous run per l), adds |G[i][j]| (j != i) to the loss and leaves
// sign(G[i][j]) in shared memory.  Phase 2 (gradient wanted): g[l][i] += 2 w / (C (C - 1)) * sum_j sign_ij M[j][l] — the
// Gram matrix is symmetric, so row i and column i of sign contribute the same amount.
__global__ void __launch_bounds__(256)
ortho_kernel(const float* __restrict__ ln, float* __restrict__ g, int C, int L, float weight, const float* __restrict__ dev_scale,
             double* __restrict__ loss_out) {
  extern __shared__ float sgn[];          // [C]
  __shared__ float part[8];
  const int i = blockIdx.x;
  float mine = 0.f;
  for (int j = threadIdx.x; j < C; j += blockDim.x) {
    float acc = 0.f;
    for (int l = 0; l < L; ++l) acc = fmaf(__ldg(ln + (size_t)l * C + i), __ldg(ln + (size_t)l * C + j), acc);
    const bool off = j != i;
    sgn[j] = off ? (acc > 0.f ? 1.f : (acc < 0.f ? -1.f : 0.f)) : 0.f;
    if (off) mine += fabsf(acc);
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(FULL_MASK, mine, d);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = mine;
  __syncthreads();
  const float inv = 1.f / ((float)C * (float)(C - 1));
  if (threadIdx.x == 0 && loss_out) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += part[w];
    atomicAdd(loss_out, (double)(s * inv));
  }
  if (!g) return;
  const float scale = 2.f * weight * inv * (dev_scale ? __ldg(dev_scale) : 1.f);
  for (int l = threadIdx.x; l < L; l += blockDim.x) {
    const float* row = ln + (size_t)l * C;
    float acc = 0.f;
    for (int j = 0; j < C; ++j) acc = fmaf(sgn[j], __ldg(row + j), acc);
    g[(size_t)l * C + i] += scale * acc;          // element (l, i) belongs to this CTA alone
  }
}

}  // namespace ftrf

using namespace ftrf;

extern "C" int ftrf_resample_factor(const float* src, int32_t C, int32_t H, int32_t W, float* dst, int32_t H2, int32_t W2,
                                    void* stream) {
  if (!src || !dst || C < 4 || (C & 3) || H < 1 || W < 1 || H2 < 1 || W2 < 1) { set_error("resample_factor: bad argument"); return FTRF_E_ARG; }
  const float sy = H2 > 1 ? (float)(H - 1) / (float)(H2 - 1) : 0.f, sx = W2 > 1 ? (float)(W - 1) / (float)(W2 - 1) : 0.f;
  const int64_t total = (int64_t)H2 * W2 * (C >> 2);
  const int64_t blocks = (total + 255) / 256;
  const int grid = (int)(blocks < (int64_t)FTRF_SMS * 16 ? blocks : (int64_t)FTRF_SMS * 16);
  resample_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, C, H, W, dst, H2, W2, sy, sx);
  return check_launch("resample_factor");
}

extern "C" int ftrf_crop_factor(const float* src, int32_t C, int32_t H, int32_t W, int32_t y0, int32_t y1, int32_t x0, int32_t x1,
                                float* dst, void* stream) {
  if (!src || !dst || C < 1 || y0 < 0 || x0 < 0 || y1 > H || x1 > W || y1 <= y0 || x1 <= x0) { set_error("crop_factor: bad box [%d,%d)x[%d,%d) of %dx%d", y0, y1, x0, x1, H, W); return FTRF_E_ARG; }
  const size_t row = (size_t)(x1 - x0) * C * sizeof(float);
  if (cudaMemcpy2DAsync(dst, row, src + ((size_t)y0 * W + x0) * C, (size_t)W * C * sizeof(float), row, (size_t)(y1 - y0),
                        cudaMemcpyDeviceToDevice, (cudaStream_t)stream) != cudaSuccess)
    return check_launch("crop_factor");
  return FTRF_OK;
}

extern "C" int ftrf_ortho_reg(const float* const* lines, float* const* g_lines, const int32_t* n_comp, const int32_t* line_len,
                              int n_lines, float weight, const float* dev_scale, double* loss_out, void* stream) {
  if (!lines || !n_comp || !line_len || n_lines < 1) { set_error("ortho_reg: bad argument"); return FTRF_E_ARG; }
  for (int k = 0; k < n_lines; ++k) {
    if (!lines[k] || n_comp[k] < 2 || line_len[k] < 1) { set_error("ortho_reg: line %d needs at least 2 components", k); return FTRF_E_ARG; }
    ortho_kernel<<<n_comp[k], 256, (size_t)n_comp[k] * sizeof(float), (cudaStream_t)stream>>>(
        lines[k], g_lines ? g_lines[k] : nullptr, n_comp[k], line_len[k], weight, dev_scale, loss_out);
    if (int e = check_launch("ortho_reg")) return e;
  }
  return FTRF_OK;
}
