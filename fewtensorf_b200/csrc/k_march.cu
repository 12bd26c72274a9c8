// k_march.cu — stage 1: AABB ray sampling + alphaMask occupancy culling + warp-ballot compaction.
// Replaces TensorBase.sample_ray / sample_ray_ndc and the alphaMask cull of forward
// (models/tensorBase.py:178-208, :349-354).  The [N,S,3] point tensor is never materialised:
// one warp owns one ray, lanes stride over the sample index i, each 32-wide step is decided
// with one __ballot_sync, and survivors are written as a packed, per-ray contiguous, in-order
// list (ray id, i).  Bit-exact with the reference by construction (see common.cuh).
#include "common.cuh"

namespace ftrf {

constexpr int MARCH_WARPS = 8;
constexpr int MARCH_RPW = 4;          // rays a warp marches per list allocation

// Can any sample with index in [i0, i1] touch a set alphaMask texel?  Conservative: the span's bounding
// box in (approximate) texel coordinates, widened by a texel, against the 32^3-block occupancy level
// (one word per (z, y) block row, bit = x block).  `la + lb * i` is the texel coordinate of sample i.
__device__ __forceinline__ bool span_may_hit(const FtrfField& f, const uint32_t* __restrict__ sup, const float la[3], const float lb[3],
                                             int i0, int i1) {
  int lo[3], hi[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float fa = fmaf(lb[k], (float)i0, la[k]), fb = fmaf(lb[k], (float)i1, la[k]);
    const float mn = fminf(fa, fb) - 1.f, mx = fmaxf(fa, fb) + 1.f;
    if (!(mx >= -1.f) || !(mn <= (float)f.am_dim[k])) return false;      // no lower corner in [-1, dim - 1]
    lo[k] = max(__float2int_rd(mn), 0) >> 5;
    hi[k] = min(__float2int_rd(mx), f.am_dim[k] - 1) >> 5;
    if (hi[k] - lo[k] > 1) return true;                                  // longer than a block: not decided here
  }
  const int SY = (f.am_dim[1] + 31) >> 5;
  const uint32_t xmask = (2u << hi[0]) - (1u << lo[0]);
  const uint32_t any = __ldg(sup + lo[2] * SY + lo[1]) | __ldg(sup + lo[2] * SY + hi[1]) |
                       __ldg(sup + hi[2] * SY + lo[1]) | __ldg(sup + hi[2] * SY + hi[1]);
  return (any & xmask) != 0u;
}

// The cursor of the packed list is ONE address: an atomicAdd per ray serialises in L2, so a CTA marches
// MARCH_WARPS * rpw rays, scans their counts in shared memory and allocates once.
// Work per ray: (1) one lane per 32-sample chunk decides from the coarse occupancy whether the chunk can
// hold a kept sample at all (most of a ray is empty space); (2) flagged chunks get the per-sample test:
// exact box test, coarse block of the sample by approximate arithmetic, and only in occupied blocks the
// exact, reference-ordered alphaMask test.  Keep decisions are the reference's, bit for bit.
__global__ void __launch_bounds__(MARCH_WARPS * 32, 4)
march_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const int rpw) {
  extern __shared__ uint32_t s_ballots[];          // [MARCH_WARPS][rpw][chunks]
  __shared__ int s_cnt[MARCH_WARPS * MARCH_RPW];
  __shared__ int s_off[MARCH_WARPS * MARCH_RPW];
  const int S = c.n_samples;
  const int chunks = (S + 31) >> 5;
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int RPC = MARCH_WARPS * rpw;               // rays per CTA step (rpw <= MARCH_RPW; 1 for small batches)
  const int64_t steps = (c.n_rays + (int64_t)gridDim.x * RPC - 1) / ((int64_t)gridDim.x * RPC);
  const bool coarse_fast = f.has_alpha_mask && f.am_coarse && !c.ndc;
  const int unit = (chunks + 31) >> 5;             // chunks per flag: one lane decides `unit` consecutive chunks
  const int CY = (f.am_dim[1] + 7) >> 3, cxw = (((f.am_dim[0] + 7) >> 3) + 31) >> 5;
  const bool use_flags = coarse_fast && f.am_dim[0] <= 1024;
  const uint32_t* sup = f.am_coarse + (size_t)cxw * CY * ((f.am_dim[2] + 7) >> 3);   // 32^3 level follows the 8^3 level

  for (int64_t it = 0; it < steps; ++it) {
    const int64_t r0 = (it * gridDim.x + blockIdx.x) * RPC + warp * rpw;
    int used[MARCH_RPW], count[MARCH_RPW];
    float tmin[MARCH_RPW];
#pragma unroll
    for (int q = 0; q < MARCH_RPW; ++q) {
      const int64_t r = r0 + q;
      uint32_t* bal = s_ballots + (size_t)(warp * rpw + q) * chunks;
      used[q] = 0; count[q] = 0; tmin[q] = 0.f;
      if (q >= rpw || r >= c.n_rays) continue;
      const Ray ray = load_ray(f, c, r);
      tmin[q] = ray.tmin;
      // alphaMask texel coordinate as a linear function of the sample index (approximate: it only selects
      // coarse blocks, which tolerate +-1 texel)
      float la[3], lb[3];
      bool finite = true;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float G = f.am_inv[k] * 0.5f * (float)(f.am_dim[k] - 1);
        la[k] = G * (ray.o[k] - f.am_aabb0[k] + ray.d[k] * (ray.tmin + f.step_size * ray.u));
        lb[k] = G * ray.d[k] * f.step_size;
        finite &= fabsf(la[k]) < 1e9f && fabsf(lb[k]) < 1e5f;
      }
      uint32_t flags = 0xffffffffu;
      if (use_flags && finite) {
        const int c0 = lane * unit;
        const bool fl = c0 < chunks && span_may_hit(f, sup, la, lb, c0 * 32, (c0 + unit) * 32 - 1);
        flags = __ballot_sync(FULL_MASK, fl);
      }
      for (int ch = lane; ch < chunks; ch += 32) bal[ch] = 0u;
      __syncwarp();
      bool done = false;
      while (flags && !done) {
        const int u = __ffs(flags) - 1;
        flags &= flags - 1;
        for (int ch = u * unit; ch < min((u + 1) * unit, chunks) && !done; ++ch) {
          const int i = (ch << 5) + lane;
          bool keep = false, gone = true;
          if (i < S) {
            float p[3];
            sample_pt(ray, sample_z(f, c, ray, i), p);
            keep = in_aabb(f, p);
            // left the box through a face the ray moves away from: z_i and o + d z_i are monotone in i (every
            // operation is a correctly rounded monotone function), so every later sample is outside too
            gone = false;
#pragma unroll
            for (int k = 0; k < 3; ++k)
              gone |= (ray.d[k] > 0.f && p[k] > f.aabb[3 + k]) || (ray.d[k] < 0.f && p[k] < f.aabb[k]);
            if (keep && f.has_alpha_mask) {
              bool maybe = true;
              if (coarse_fast && finite) {
                const float fi = (float)i;
                const int ix = __float2int_rd(fmaf(lb[0], fi, la[0])), iy = __float2int_rd(fmaf(lb[1], fi, la[1])),
                          iz = __float2int_rd(fmaf(lb[2], fi, la[2]));
                if ((unsigned)ix < (unsigned)f.am_dim[0] && (unsigned)iy < (unsigned)f.am_dim[1] && (unsigned)iz < (unsigned)f.am_dim[2]) {
                  const uint32_t wd = __ldg(f.am_coarse + ((size_t)(iz >> 3) * CY + (iy >> 3)) * cxw + (ix >> 8));
                  maybe = (wd >> ((ix >> 3) & 31)) & 1u;
                }
              }
              keep = maybe && alpha_mask_keep(f, p);
            }
          }
          const uint32_t b = __ballot_sync(FULL_MASK, keep);
          if (lane == 0) bal[ch] = b;
          count[q] += __popc(b);
          used[q] = ch + 1;
          if (!c.ndc && __all_sync(FULL_MASK, gone)) done = true;
        }
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int q = 0; q < MARCH_RPW; ++q)
        if (q < rpw) s_cnt[warp * rpw + q] = count[q];
    }
    __syncthreads();
    if (warp == 0) {                                 // exclusive scan of the RPC (<= 32) counts + one allocation
      const int v = lane < RPC ? s_cnt[lane] : 0;
      int inc = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(FULL_MASK, inc, d);
        if (lane >= d) inc += t;
      }
      int base = 0;
      if (lane == 31 && inc) base = atomicAdd(ws.counters + 0, inc);
      base = __shfl_sync(FULL_MASK, base, 31);
      if (lane < RPC) s_off[lane] = base + inc - v;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < MARCH_RPW; ++q) {
      const int64_t r = r0 + q;
      if (q >= rpw || r >= c.n_rays) continue;
      const uint32_t* bal = s_ballots + (size_t)(warp * rpw + q) * chunks;
      const int base = s_off[warp * rpw + q];
      const bool fits = (int64_t)base + count[q] <= c.cap_samples;
      if (lane == 0) {
        if (!fits) atomicOr(ws.counters + 2, 1);
        ws.tmin[r] = tmin[q];
        ws.ray_off[r] = count[q] ? base : 0;
        ws.ray_cnt[r] = fits ? count[q] : 0;
      }
      if (fits && count[q]) {
        int pos0 = base;
        for (int ch = 0; ch < used[q]; ++ch) {
          const uint32_t b = bal[ch];
          if ((b >> lane) & 1u) {
            const int pos = pos0 + __popc(b & ((1u << lane) - 1u));
            ws.s_ray[pos] = (int32_t)r;
            ws.s_idx[pos] = (ch << 5) + lane;
          }
          pos0 += __popc(b);
        }
      }
    }
    __syncthreads();                                 // s_cnt / s_off / ballots are reused by the next step
  }
}

int launch_march(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  int chunks = (c.n_samples + 31) / 32;
  int64_t cap = (int64_t)FTRF_SMS * 8;             // 8 resident CTAs/SM worth of warps, then grid-stride
  const int rpw = c.n_rays <= cap * MARCH_WARPS ? 1 : MARCH_RPW;
  size_t smem = (size_t)MARCH_WARPS * rpw * chunks * sizeof(uint32_t);
  if (smem > 200 * 1024) { set_error("march: n_samples %d too large", c.n_samples); return FTRF_E_ARG; }
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(march_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t blocks = (c.n_rays + MARCH_WARPS * rpw - 1) / (MARCH_WARPS * rpw);
  int grid = (int)(blocks < cap ? blocks : cap);
  march_kernel<<<grid, MARCH_WARPS * 32, smem, st>>>(f, c, ws, rpw);
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
    // grid_sample(volume) > 0 on a NON-NEGATIVE volume <=> some touched texel is > 0 (the wrapper rejects negatives)
    bool on = x < X && vol[row * X + x] > 0.f;
    uint32_t b = __ballot_sync(FULL_MASK, on);
    if (lane == 0) bits[w] = b;
  }
}

// Coarse occupancy: one bit per 8x8x8 block of texels, set iff any texel in [8c - 1, 8c + 9]^3 is set:
// the block, the +1 upper interpolation corner, and one more texel on either side.  A sample whose
// lower interpolation corner i0 lies in a block with a clear bit cannot touch a set texel, and the
// extra ring makes that true even when i0 is only known to +-1 (the march kernel locates the block
// with three fused multiply-adds per sample instead of the exact, reference-ordered arithmetic, and
// runs the exact test only where the bit is set).
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
      const int x0 = max(cx * 8 - 1, 0), x1 = min(cx * 8 + 9, X - 1);
      for (int z = max(cz * 8 - 1, 0); z <= min(cz * 8 + 9, Z - 1) && !on; ++z)
        for (int y = max(cy * 8 - 1, 0); y <= min(cy * 8 + 9, Y - 1) && !on; ++y) {
          const uint32_t* row = bits + ((size_t)z * Y + y) * xw;
          for (int x = x0; x <= x1; ++x)
            if ((row[x >> 5] >> (x & 31)) & 1u) { on = true; break; }
        }
    }
    uint32_t b = __ballot_sync(FULL_MASK, on);
    if (lane == 0) coarse[w] = b;
  }
}

// 32^3 level from the 8^3 level: bit x of row (z, y) = OR of the 4 x 4 x 4 coarse bits of that block
// (for x > 1024 texels the row word cannot hold the bits and the march kernel does not use the level).
__global__ void super_mask_kernel(const uint32_t* __restrict__ coarse, int X, int Y, int Z, uint32_t* __restrict__ sup) {
  const int CX = (X + 7) >> 3, CY = (Y + 7) >> 3, CZ = (Z + 7) >> 3, cxw = (CX + 31) >> 5;
  const int SY = (Y + 31) >> 5, SZ = (Z + 31) >> 5;
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= SY * SZ) return;
  const int sy = row % SY, sz = row / SY;
  uint32_t out = 0u;
  for (int cz = sz * 4; cz < min(sz * 4 + 4, CZ); ++cz)
    for (int cy = sy * 4; cy < min(sy * 4 + 4, CY); ++cy)
      for (int cx = 0; cx < min(CX, 128); ++cx)
        if ((coarse[((size_t)cz * CY + cy) * cxw + (cx >> 5)] >> (cx & 31)) & 1u) out |= 1u << (cx >> 2);
  sup[row] = out;
}

}  // namespace ftrf

static size_t coarse8_words(int32_t x, int32_t y, int32_t z) {
  return (size_t)(((x + 7) / 8 + 31) / 32) * (size_t)((y + 7) / 8) * (size_t)((z + 7) / 8);
}

// 8^3 level followed by the 32^3 level (one word per (z, y) block row; x <= 1024 texels)
extern "C" size_t ftrf_alpha_mask_coarse_words(int32_t x, int32_t y, int32_t z) {
  return coarse8_words(x, y, z) + (size_t)((y + 31) / 32) * (size_t)((z + 31) / 32);
}

extern "C" int ftrf_build_coarse_mask(const uint32_t* bits, int32_t x, int32_t y, int32_t z, uint32_t* coarse, void* stream) {
  if (!bits || !coarse || x <= 0 || y <= 0 || z <= 0) { ftrf::set_error("build_coarse_mask: bad argument"); return FTRF_E_ARG; }
  int64_t words = (int64_t)coarse8_words(x, y, z);
  int64_t blocks = (words * 32 + 255) / 256;
  int grid = (int)(blocks < FTRF_SMS * 16 ? blocks : FTRF_SMS * 16);
  ftrf::coarse_mask_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(bits, x, y, z, coarse);
  if (int e = ftrf::check_launch("build_coarse_mask")) return e;
  const int rows = ((y + 31) / 32) * ((z + 31) / 32);
  ftrf::super_mask_kernel<<<(rows + 127) / 128, 128, 0, (cudaStream_t)stream>>>(coarse, x, y, z, coarse + words);
  return ftrf::check_launch("build_super_mask");
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
