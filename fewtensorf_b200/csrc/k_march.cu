// k_march.cu — stage 1: AABB ray sampling + alphaMask occupancy culling + warp-ballot compaction.
// Replaces TensorBase.sample_ray / sample_ray_ndc and the alphaMask cull of forward
// (models/tensorBase.py:178-208, :349-354).  The [N,S,3] point tensor is never materialised:
// one warp owns one ray, lanes stride over the sample index i, each 32-wide step is decided
// with one __ballot_sync, and survivors are written as a packed, per-ray contiguous, in-order
// list (ray id, i).  Bit-exact with the reference by construction (see common.cuh).
#include "common.cuh"

namespace ftrf {

constexpr int MARCH_WARPS = 8;

__global__ void __launch_bounds__(MARCH_WARPS * 32)
march_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ uint32_t s_ballots[];          // [MARCH_WARPS][chunks]
  const int S = c.n_samples;
  const int chunks = (S + 31) >> 5;
  const int warp = threadIdx.x >> 5, lane = lane_id();
  uint32_t* bal = s_ballots + warp * chunks;
  const int64_t ray0 = (int64_t)blockIdx.x * MARCH_WARPS + warp;
  const int64_t stride = (int64_t)gridDim.x * MARCH_WARPS;

  for (int64_t r = ray0; r < c.n_rays; r += stride) {
    Ray ray = load_ray(f, c, r);
    int count = 0;
    for (int ch = 0; ch < chunks; ++ch) {
      int i = (ch << 5) + lane;
      bool keep = false;
      if (i < S) {
        float p[3];
        sample_pt(ray, sample_z(f, c, ray, i), p);
        keep = in_aabb(f, p);
        if (keep && f.has_alpha_mask) keep = alpha_mask_keep(f, p);
      }
      uint32_t b = __ballot_sync(FULL_MASK, keep);
      if (lane == 0) bal[ch] = b;
      count += __popc(b);
    }
    // one allocation per ray on the global cursor
    int base = 0;
    if (lane == 0) {
      base = count ? atomicAdd(ws.counters + 0, count) : 0;
      bool fits = (int64_t)base + count <= c.cap_samples;
      if (!fits) { atomicOr(ws.counters + 2, 1); }
      ws.tmin[r] = ray.tmin;
      ws.ray_off[r] = base;
      ws.ray_cnt[r] = fits ? count : 0;
      if (!fits) count = 0;
    }
    base = __shfl_sync(FULL_MASK, base, 0);
    count = __shfl_sync(FULL_MASK, count, 0);
    __syncwarp();
    if (count) {
      int run = base;
      for (int ch = 0; ch < chunks; ++ch) {
        uint32_t b = bal[ch];
        if ((b >> lane) & 1u) {
          int pos = run + __popc(b & ((1u << lane) - 1u));
          ws.s_ray[pos] = (int32_t)r;
          ws.s_idx[pos] = (ch << 5) + lane;
        }
        run += __popc(b);
      }
    }
    __syncwarp();
  }
}

int launch_march(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  int chunks = (c.n_samples + 31) / 32;
  size_t smem = (size_t)MARCH_WARPS * chunks * sizeof(uint32_t);
  if (smem > 200 * 1024) { set_error("march: n_samples %d too large", c.n_samples); return FTRF_E_ARG; }
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(march_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t blocks = (c.n_rays + MARCH_WARPS - 1) / MARCH_WARPS;
  int64_t cap = (int64_t)FTRF_SMS * 8;             // 8 resident CTAs/SM worth of warps, then grid-stride
  int grid = (int)(blocks < cap ? blocks : cap);
  march_kernel<<<grid, MARCH_WARPS * 32, smem, st>>>(f, c, ws);
  return check_launch("march");
}

// ---------------------------------------------------------------------------------------------
// AlphaGridMask volume (float {0,1}, (Z,Y,X)) -> bit field, 32 x-texels per word
// (models/tensorBase.py:30-39; the reference samples the float volume with a 3-D grid_sample).
__global__ void pack_mask_kernel(const float* __restrict__ vol, int X, int Y, int Z, uint32_t* __restrict__ bits) {
  int xw = (X + 31) >> 5;
  int64_t words = (int64_t)xw * Y * Z;
  int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // one warp per word
  int lane = threadIdx.x & 31;
  int64_t wstride = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = gw; w < words; w += wstride) {
    int64_t row = w / xw;
    int x = (int)(w % xw) * 32 + lane;
    bool on = x < X && vol[row * X + x] > 0.5f;
    uint32_t b = __ballot_sync(FULL_MASK, on);
    if (lane == 0) bits[w] = b;
  }
}

// Coarse occupancy: one bit per 8x8x8 block of texels, set iff any texel of the block OR of its +1
// neighbours in x / y / z is set.  A sample whose lower interpolation corner lies in a block with a
// clear bit cannot touch a set texel, so the march kernel rejects it with one L1-resident load
// instead of eight dependent ones; samples in blocks with a set bit take the exact test.
__global__ void coarse_mask_kernel(const uint32_t* __restrict__ bits, int X, int Y, int Z, uint32_t* __restrict__ coarse) {
  const int xw = (X + 31) >> 5;
  const int CX = (X + 7) >> 3, CY = (Y + 7) >> 3, CZ = (Z + 7) >> 3, cxw = (CX + 31) >> 5;
  const int64_t words = (int64_t)cxw * CY * CZ;
  const int lane = threadIdx.x & 31;
  const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, wstride = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = gw; w < words; w += wstride) {
    const int cw = (int)(w % cxw), cy = (int)((w / cxw) % CY), cz = (int)(w / ((int64_t)cxw * CY));
    const int cx = cw * 32 + lane;
    bool on = false;
    if (cx < CX) {
      const int x0 = cx * 8, x1 = min(x0 + 8, X - 1);
      for (int z = cz * 8; z <= min(cz * 8 + 8, Z - 1) && !on; ++z)
        for (int y = cy * 8; y <= min(cy * 8 + 8, Y - 1) && !on; ++y) {
          const uint32_t* row = bits + ((size_t)z * Y + y) * xw;
          for (int x = x0; x <= x1; ++x)
            if ((row[x >> 5] >> (x & 31)) & 1u) { on = true; break; }
        }
    }
    uint32_t b = __ballot_sync(FULL_MASK, on);
    if (lane == 0) coarse[w] = b;
  }
}

}  // namespace ftrf

extern "C" size_t ftrf_alpha_mask_coarse_words(int32_t x, int32_t y, int32_t z) {
  return (size_t)(((x + 7) / 8 + 31) / 32) * (size_t)((y + 7) / 8) * (size_t)((z + 7) / 8);
}

extern "C" int ftrf_build_coarse_mask(const uint32_t* bits, int32_t x, int32_t y, int32_t z, uint32_t* coarse, void* stream) {
  if (!bits || !coarse || x <= 0 || y <= 0 || z <= 0) { ftrf::set_error("build_coarse_mask: bad argument"); return FTRF_E_ARG; }
  int64_t words = (int64_t)ftrf_alpha_mask_coarse_words(x, y, z);
  int64_t blocks = (words * 32 + 255) / 256;
  int grid = (int)(blocks < FTRF_SMS * 16 ? blocks : FTRF_SMS * 16);
  ftrf::coarse_mask_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(bits, x, y, z, coarse);
  return ftrf::check_launch("build_coarse_mask");
}

extern "C" size_t ftrf_alpha_mask_words(int32_t x, int32_t y, int32_t z) {
  return (size_t)((x + 31) / 32) * (size_t)y * (size_t)z;
}

extern "C" int ftrf_pack_alpha_mask(const float* volume, int32_t x, int32_t y, int32_t z, uint32_t* bits,
                                    void* stream) {
  if (!volume || !bits || x <= 0 || y <= 0 || z <= 0) { ftrf::set_error("pack_alpha_mask: bad argument"); return FTRF_E_ARG; }
  int64_t words = (int64_t)ftrf_alpha_mask_words(x, y, z);
  int64_t blocks = (words * 32 + 255) / 256;
  int grid = (int)(blocks < FTRF_SMS * 16 ? blocks : FTRF_SMS * 16);
  ftrf::pack_mask_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(volume, x, y, z, bits);
  return ftrf::check_launch("pack_alpha_mask");
}
