// k_occupancy.cu — the occupancy side of the hot path: what produces the alphaMask
// the march kernel culls with, and the ray pre-filter that consults it.
//   updateAlphaMask   models/tensorBase.py:233-256   -> dense_alpha_bits_kernel (k_density.cu) + mask_dilate_kernel
//   filtering_rays    models/tensorBase.py:259-288   -> filter_rays_kernel
// Everything stays bit-packed (32 x-texels per word, (Z,Y,X) order): a 256^3 rebuild reads the factors
// once, writes 2 MiB of raw bits and 2 MiB of dilated bits, and never holds the 201 MB point lattice or
// the 67 MB float volumes of the reference formulation.
#include "common.cuh"
#include <climits>

namespace ftrf {

// stats: int32[8] = {min x, min y, min z, max x, max y, max z, count lo, count hi}
__global__ void mask_stats_init_kernel(int32_t* stats) {
  if (threadIdx.x < 3) stats[threadIdx.x] = INT_MAX;
  else if (threadIdx.x < 6) stats[threadIdx.x] = -1;
  else if (threadIdx.x < 8) stats[threadIdx.x] = 0;
}

// F.max_pool3d(kernel 3, stride 1, padding 1) + the two threshold assignments (tensorBase.py:243-246) on
// the thresholded bits: out = OR of the 27 neighbours (padding never wins a max), one thread per word.
// Also the reductions updateAlphaMask needs afterwards: index bounding box of the set texels (the
// lattice points are monotone in the index, so xyz_min / xyz_max are the points at these indices,
// :250-253) and their count (:257).
__global__ void __launch_bounds__(256)
mask_dilate_kernel(const uint32_t* __restrict__ raw, const int X, const int Y, const int Z, uint32_t* __restrict__ bits,
                   int32_t* __restrict__ stats) {
  const int xw = (X + 31) >> 5;
  const int64_t words = (int64_t)xw * Y * Z;
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {-1, -1, -1};
  unsigned cnt = 0;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    const int xi = (int)(w % xw), y = (int)((w / xw) % Y), z = (int)(w / ((int64_t)xw * Y));
    uint32_t out = 0u;
    for (int zz = max(z - 1, 0); zz <= min(z + 1, Z - 1); ++zz)
      for (int yy = max(y - 1, 0); yy <= min(y + 1, Y - 1); ++yy) {
        const uint32_t* row = raw + ((size_t)zz * Y + yy) * xw;
        const uint32_t c = __ldg(row + xi);
        const uint32_t l = xi > 0 ? __ldg(row + xi - 1) : 0u, r = xi + 1 < xw ? __ldg(row + xi + 1) : 0u;
        out |= c | (c << 1) | (c >> 1) | (l >> 31) | (r << 31);
      }
    if (xi == xw - 1 && (X & 31)) out &= (1u << (X & 31)) - 1u;
    bits[w] = out;
    if (out) {
      mn[0] = min(mn[0], xi * 32 + __ffs(out) - 1);
      mx[0] = max(mx[0], xi * 32 + 31 - __clz(out));
      mn[1] = min(mn[1], y); mx[1] = max(mx[1], y);
      mn[2] = min(mn[2], z); mx[2] = max(mx[2], z);
      cnt += __popc(out);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    mn[k] = __reduce_min_sync(FULL_MASK, mn[k]);
    mx[k] = __reduce_max_sync(FULL_MASK, mx[k]);
  }
  cnt = __reduce_add_sync(FULL_MASK, cnt);
  if (lane_id() == 0 && cnt) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { atomicMin(stats + k, mn[k]); atomicMax(stats + 3 + k, mx[k]); }
    atomicAdd(reinterpret_cast<unsigned long long*>(stats + 6), (unsigned long long)cnt);
  }
}

// bit field -> the float (Z,Y,X) volume of {0,1} the reference's AlphaGridMask exposes (tensorBase.py:31-39)
__global__ void __launch_bounds__(256)
mask_unpack_kernel(const uint32_t* __restrict__ bits, const int X, const int Y, const int Z, float* __restrict__ vol) {
  const int xw = (X + 31) >> 5;
  const int64_t n = (int64_t)X * Y * Z;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % X);
    const int64_t row = i / X;
    vol[i] = (float)((__ldg(bits + row * xw + (x >> 5)) >> (x & 31)) & 1u);
  }
}

// filtering_rays, alphaMask branch (tensorBase.py:285-286): keep a ray iff any of its n_samples
// un-jittered march points (sample_ray(is_train=False), :189-208 — t_min clamped to [near, far], NO box
// test: points outside the box still interpolate against the zero padding) reads alpha > 0.  Warp per
// ray, 32 samples per ballot, stops at the first hit.
__global__ void __launch_bounds__(256)
filter_rays_kernel(const FtrfField f, const float* __restrict__ rays, const int ray_cols, const int64_t n, const int n_samples,
                   uint8_t* __restrict__ keep) {
  const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t wstride = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int lane = lane_id();
  for (int64_t r = w0; r < n; r += wstride) {
    Ray ray;
    const float* p = rays + r * ray_cols;
#pragma unroll
    for (int k = 0; k < 3; ++k) { ray.o[k] = __ldg(p + k); ray.d[k] = __ldg(p + 3 + k); }
    ray.tmin = ray_tmin(f, ray.o, ray.d);
    bool hit = false;
    for (int i0 = 0; i0 < n_samples && !hit; i0 += 32) {
      const int i = i0 + lane;
      bool k = false;
      if (i < n_samples) {
        float pt[3];
        sample_pt(ray, __fadd_rn(ray.tmin, __fmul_rn(f.step_size, (float)i)), pt);
        k = alpha_mask_keep(f, pt);
      }
      hit = __any_sync(FULL_MASK, k);
    }
    if (lane == 0) keep[r] = hit ? 1 : 0;
  }
}

// filtering_rays, bbox_only branch (tensorBase.py:277-283): slab test without the near/far clamp.
__global__ void __launch_bounds__(256)
filter_rays_bbox_kernel(const FtrfField f, const float* __restrict__ rays, const int ray_cols, const int64_t n,
                        uint8_t* __restrict__ keep) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
    const float* p = rays + r * ray_cols;
    float tmin = -INFINITY, tmax = INFINITY;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float o = __ldg(p + k), d = __ldg(p + 3 + k);
      const float v = (d == 0.f) ? 1e-6f : d;
      const float ra = __fdiv_rn(__fsub_rn(f.aabb[3 + k], o), v), rb = __fdiv_rn(__fsub_rn(f.aabb[k], o), v);
      tmin = fmaxf(tmin, fminf(ra, rb));
      tmax = fminf(tmax, fmaxf(ra, rb));
    }
    keep[r] = tmax > tmin ? 1 : 0;
  }
}

static int grid_for(int64_t threads) {
  int64_t blocks = (threads + 255) / 256;
  const int64_t cap = (int64_t)FTRF_SMS * 8 * 4;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

}  // namespace ftrf

extern "C" int ftrf_dense_alpha_bits(const FtrfField* f, const float* lin, int32_t x, int32_t y, int32_t z, float length,
                                     float thres, uint32_t* raw_bits, float* alpha_out, void* stream) {
  if (!f || !lin || !raw_bits || x <= 0 || y <= 0 || z <= 0) { ftrf::set_error("dense_alpha_bits: bad argument"); return FTRF_E_ARG; }
  return ftrf::launch_dense_alpha_bits(*f, lin, x, y, z, length, thres, raw_bits, alpha_out, (cudaStream_t)stream);
}

extern "C" int ftrf_dilate_alpha_mask(const uint32_t* raw_bits, int32_t x, int32_t y, int32_t z, uint32_t* bits,
                                      int32_t* stats, void* stream) {
  if (!raw_bits || !bits || !stats || raw_bits == bits || x <= 0 || y <= 0 || z <= 0) {
    ftrf::set_error("dilate_alpha_mask: bad argument");
    return FTRF_E_ARG;
  }
  const int64_t words = (int64_t)ftrf_alpha_mask_words(x, y, z);
  ftrf::mask_stats_init_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(stats);
  if (int e = ftrf::check_launch("mask_stats_init")) return e;
  ftrf::mask_dilate_kernel<<<ftrf::grid_for(words), 256, 0, (cudaStream_t)stream>>>(raw_bits, x, y, z, bits, stats);
  return ftrf::check_launch("dilate_alpha_mask");
}

extern "C" int ftrf_unpack_alpha_mask(const uint32_t* bits, int32_t x, int32_t y, int32_t z, float* volume, void* stream) {
  if (!bits || !volume || x <= 0 || y <= 0 || z <= 0) { ftrf::set_error("unpack_alpha_mask: bad argument"); return FTRF_E_ARG; }
  ftrf::mask_unpack_kernel<<<ftrf::grid_for((int64_t)x * y * z), 256, 0, (cudaStream_t)stream>>>(bits, x, y, z, volume);
  return ftrf::check_launch("unpack_alpha_mask");
}

extern "C" int ftrf_filter_rays(const FtrfField* f, const float* rays, int32_t ray_cols, int64_t n, int32_t n_samples,
                                int32_t bbox_only, uint8_t* keep, void* stream) {
  if (!f || !keep || n < 0 || (n > 0 && !rays) || ray_cols < 6 || (!bbox_only && n_samples < 0)) {
    ftrf::set_error("filter_rays: bad argument");
    return FTRF_E_ARG;
  }
  if (!bbox_only && !f->has_alpha_mask) { ftrf::set_error("filter_rays: the field has no alphaMask (use bbox_only)"); return FTRF_E_ARG; }
  if (n == 0) return FTRF_OK;
  if (bbox_only)
    ftrf::filter_rays_bbox_kernel<<<ftrf::grid_for(n), 256, 0, (cudaStream_t)stream>>>(*f, rays, ray_cols, n, keep);
  else
    ftrf::filter_rays_kernel<<<ftrf::grid_for(n * 32), 256, 0, (cudaStream_t)stream>>>(*f, rays, ray_cols, n, n_samples, keep);
  return ftrf::check_launch("filter_rays");
}
