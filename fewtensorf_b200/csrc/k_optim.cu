// k_optim.cu — the two dense passes over the factor grids that sit either side of the per-ray hot path
// inside one training iteration:
//   * regularisers: TVLoss of the VM planes (loss.py:120-141 through models/tensoRF.py:195-205) and the L1
//     of the density factors (models/tensoRF.py:189-193, :468-472), value AND gradient in ONE pass, the
//     gradient added straight into the flat gradient buffer the ray kernels reduce into;
//   * torch.optim.Adam over every parameter (train.py:272-273, :374-376) as ONE launch over a segment
//     table instead of torch's per-tensor / multi-pass foreach chain.
// Both are pure HBM streams: 12 B per plane element (x read once, g read + written; the four neighbours come
// out of L1/L2) and 28 B per parameter (p, g, m, v in; p, m, v out).
#include "common.cuh"

namespace ftrf {

// ------------------------------------------------------------------------------------ regularisers
struct RegPlane {
  const float* x; float* g;
  int H, W, C;
  float ch, cw;        // loss scale of one squared vertical / horizontal difference (already x 2 x 1e-2 / count)
  float l1;            // 1 / numel
};
struct RegArgs {
  RegPlane p[3];
  float w_tv, w_l1;            // weights of the two losses in the gradient
  const float* dev_scale;      // optional device scalar multiplying both weights (autograd's upstream gradient)
  double* loss;                // [2]: += TV loss, += L1 loss (unweighted, as the reference methods return them)
};

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) - (v < 0.f); }

// Channel-last plane [H][W][C]: one thread owns the 4 channels of one texel.  Differences are counted once
// per pair (towards h+1 / w+1); the gradient of texel t collects both pairs it belongs to on each axis.
// Persistent grid over tiles of T.rows image rows x 256 float4 of a row, the three planes taken as ONE tile
// space (a 16-channel 300^2 plane alone quantises badly).  A thread walks DOWN its column of the tile with
// the rows above / below in registers, so each plane element comes in from L2 once (+ 2/rows halo) rather
// than three times — the per-SM L2 port, not HBM, was the limit of the flat version; left / right neighbours
// are the neighbouring threads' centre loads (L1).  Few blocks also keep the two double atomics per block
// that publish the loss values few (same-address atomics serialise).
constexpr int REG_THREADS = 256;
struct RegTiles { uint32_t first[4]; uint32_t chunks[3]; int rows; };     // tile prefix per plane, column chunks per row, rows per tile

__global__ void __launch_bounds__(REG_THREADS, 6)
reg_planes_kernel(const RegArgs a, const RegTiles T) {
  float tv = 0.f, l1 = 0.f;
  float scale = 1.f;
  if (a.dev_scale) scale = __ldg(a.dev_scale);
  for (uint32_t tile = blockIdx.x; tile < T.first[3]; tile += gridDim.x) {
    const int k = tile < T.first[1] ? 0 : (tile < T.first[2] ? 1 : 2);
    const RegPlane& P = a.p[k];
    const uint32_t lt = tile - T.first[k];
    const uint32_t rg = lt / T.chunks[k], cc = lt - rg * T.chunks[k];
    const uint32_t c4n = P.C >> 2;
    const uint32_t rowq = (uint32_t)P.W * c4n;          // float4 per image row
    const uint32_t col = cc * REG_THREADS + threadIdx.x; // float4 index inside the row
    if (col >= rowq) continue;
    const int w = (int)(col / c4n);
    const int h0 = (int)rg * T.rows;
    const int h1 = h0 + T.rows < P.H ? h0 + T.rows : P.H;
    const float gtv = a.w_tv * scale, gl1 = a.w_l1 * scale * P.l1;
    const bool want_g = P.g != nullptr;
    const bool left = w > 0 && want_g, right = w + 1 < P.W;
    const float ch = 2.f * gtv * P.ch, cw = 2.f * gtv * P.cw;
    const float4* X = reinterpret_cast<const float4*>(P.x) + (size_t)h0 * rowq + col;
    float4* G = reinterpret_cast<float4*>(P.g) + (size_t)h0 * rowq + col;
    float4 x = __ldg(X);
    float4 yu = (h0 > 0 && want_g) ? __ldg(X - rowq) : x;     // a missing neighbour is replaced by x (difference 0)
#pragma unroll 2
    for (int h = h0; h < h1; ++h, X += rowq, G += rowq) {
      const float4 yd = h + 1 < P.H ? __ldg(X + rowq) : x;
      const float4 yr = right ? __ldg(X + c4n) : x;
      const float4 yl = left ? __ldg(X - c4n) : x;
      const float4 dd = make_float4(x.x - yd.x, x.y - yd.y, x.z - yd.z, x.w - yd.w);
      const float4 dr = make_float4(x.x - yr.x, x.y - yr.y, x.z - yr.z, x.w - yr.w);
      tv = fmaf(P.ch, dd.x * dd.x + dd.y * dd.y + dd.z * dd.z + dd.w * dd.w, tv);
      tv = fmaf(P.cw, dr.x * dr.x + dr.y * dr.y + dr.z * dr.z + dr.w * dr.w, tv);
      l1 = fmaf(fabsf(x.x) + fabsf(x.y) + fabsf(x.z) + fabsf(x.w), P.l1, l1);
      if (want_g) {
        // d/dx of c (x - y)^2 is 2 c (x - y)
        float4 o = *G;
        o.x += ch * (dd.x + (x.x - yu.x)) + cw * (dr.x + (x.x - yl.x)) + gl1 * sgn(x.x);
        o.y += ch * (dd.y + (x.y - yu.y)) + cw * (dr.y + (x.y - yl.y)) + gl1 * sgn(x.y);
        o.z += ch * (dd.z + (x.z - yu.z)) + cw * (dr.z + (x.z - yl.z)) + gl1 * sgn(x.z);
        o.w += ch * (dd.w + (x.w - yu.w)) + cw * (dr.w + (x.w - yl.w)) + gl1 * sgn(x.w);
        *G = o;
      }
      yu = x;
      x = yd;
    }
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    tv += __shfl_xor_sync(FULL_MASK, tv, d);
    l1 += __shfl_xor_sync(FULL_MASK, l1, d);
  }
  __shared__ float s_tv[REG_THREADS / 32], s_l1[REG_THREADS / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_tv[warp] = tv; s_l1[warp] = l1; }
  __syncthreads();
  if (threadIdx.x == 0 && a.loss) {
    double t = 0.0, l = 0.0;
    for (int i = 0; i < REG_THREADS / 32; ++i) { t += s_tv[i]; l += s_l1[i]; }
    atomicAdd(a.loss + 0, t);
    atomicAdd(a.loss + 1, l);
  }
}

// L1 of the three lines [L][C] (tiny): one element per thread, blockIdx.y = line (a block-per-line loop made
// ~20 dependent global round trips and took longer than the 52 MB plane pass).
struct RegLines { const float* x[3]; float* g[3]; int n[3]; float w_l1; const float* dev_scale; double* loss; };
__global__ void __launch_bounds__(256)
reg_lines_kernel(const RegLines a) {
  const int k = blockIdx.y;
  const float* x = a.x[k];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = x != nullptr && i < a.n[k];
  const float inv = 1.f / (float)(a.n[k] > 0 ? a.n[k] : 1);
  float l1 = 0.f;
  if (live) {
    const float scale = a.dev_scale ? __ldg(a.dev_scale) : 1.f;
    const float v = __ldg(x + i);
    l1 = fabsf(v);
    if (a.g[k]) a.g[k][i] += a.w_l1 * scale * inv * sgn(v);
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) l1 += __shfl_xor_sync(FULL_MASK, l1, d);
  __shared__ float s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = l1;
  __syncthreads();
  if (threadIdx.x == 0 && a.loss && x != nullptr && blockIdx.x * blockDim.x < a.n[k]) {
    double l = 0.0;
    for (int j = 0; j < 8; ++j) l += s[j];
    atomicAdd(a.loss + 1, l * (double)inv);
  }
}

int launch_reg_factors(const FtrfFactorSet& fs, int model, float w_tv, float w_l1, const float* dev_scale,
                       double* loss, int with_lines, cudaStream_t st) {
  RegArgs a{};
  a.w_tv = w_tv; a.w_l1 = w_l1; a.dev_scale = dev_scale; a.loss = loss;
  if (model == FTRF_MODEL_VM) {
    for (int k = 0; k < 3; ++k) {
      RegPlane& P = a.p[k];
      P.x = fs.plane[k]; P.g = fs.g_plane[k];
      P.H = fs.plane_h[k]; P.W = fs.plane_w[k]; P.C = fs.n_comp[k];
      // TVLoss (loss.py:124-138): 2 (h_tv / count_h + w_tv / count_w) / batch, and the caller's 1e-2 (tensoRF.py:198)
      const double cnt_h = (double)P.C * (P.H - 1) * P.W, cnt_w = (double)P.C * P.H * (P.W - 1);
      if (P.H < 2 || P.W < 2) { set_error("reg_factors: plane %d is %dx%d, TVLoss needs >= 2x2", k, P.H, P.W); return FTRF_E_ARG; }
      P.ch = (float)(2.0 * 1e-2 / cnt_h);
      P.cw = (float)(2.0 * 1e-2 / cnt_w);
      P.l1 = (float)(1.0 / ((double)P.C * P.H * P.W));
      const int64_t n4 = (int64_t)P.H * P.W * (P.C >> 2);
      if (n4 >= (int64_t)1 << 31) { set_error("reg_factors: plane %d too large", k); return FTRF_E_ARG; }
    }
    static int per_sm = 0;
    if (!per_sm) {
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, reg_planes_kernel, REG_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
    }
    // rows per tile: the walk costs rows + 2 (halo) row loads per tile, and the tiles should fill whole waves
    RegTiles T{};
    double best = 1e30;
    for (int rows = 4; rows <= 32; ++rows) {
      uint32_t tiles = 0;
      for (int k = 0; k < 3; ++k)
        tiles += (uint32_t)((a.p[k].W * (a.p[k].C >> 2) + REG_THREADS - 1) / REG_THREADS) * (uint32_t)((a.p[k].H + rows - 1) / rows);
      const double waves = (double)((tiles + FTRF_SMS * per_sm - 1) / (FTRF_SMS * per_sm));
      const double cost = waves * (rows + 2);
      if (cost < best) { best = cost; T.rows = rows; }
    }
    for (int k = 0; k < 3; ++k) {
      T.chunks[k] = (uint32_t)((a.p[k].W * (a.p[k].C >> 2) + REG_THREADS - 1) / REG_THREADS);
      T.first[k + 1] = T.first[k] + T.chunks[k] * (uint32_t)((a.p[k].H + T.rows - 1) / T.rows);
    }
    const uint32_t cap = (uint32_t)FTRF_SMS * per_sm;
    reg_planes_kernel<<<T.first[3] < cap ? T.first[3] : cap, REG_THREADS, 0, st>>>(a, T);
    if (int e = check_launch("reg_planes")) return e;
  }
  if (with_lines) {
    RegLines l{};
    l.w_l1 = w_l1; l.dev_scale = dev_scale; l.loss = loss;
    for (int k = 0; k < 3; ++k) { l.x[k] = fs.line[k]; l.g[k] = fs.g_line[k]; l.n[k] = fs.line_len[k] * fs.n_comp[k]; }
    int most = 1;
    for (int k = 0; k < 3; ++k) most = l.n[k] > most ? l.n[k] : most;
    reg_lines_kernel<<<dim3((most + 255) / 256, 3), 256, 0, st>>>(l);
    if (int e = check_launch("reg_lines")) return e;
  }
  return FTRF_OK;
}

// ------------------------------------------------------------------------------------ Adam
constexpr int ADAM_MAX_SEGS = 48;
constexpr int ADAM_TILE = 4096;            // elements per block
struct AdamSeg { float* p; const float* g; float* m; float* v; int64_t n; float step_size; int first_block; };
struct AdamArgs {
  AdamSeg s[ADAM_MAX_SEGS];
  int n_segs;
  float beta1, beta2, eps, inv_bc2_sqrt;
  const float* guard;        // optional device float: a non-zero value turns the whole launch into a no-op
};

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, const AdamArgs& a, float step_size) {
  m = fmaf(1.f - a.beta1, g - m, m);                       // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(a.beta2, v, (1.f - a.beta2) * g * g);            // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = sqrtf(v) * a.inv_bc2_sqrt + a.eps;
  p -= step_size * (m / denom);
}

__global__ void __launch_bounds__(256)
adam_kernel(const __grid_constant__ AdamArgs a) {
  if (a.guard && __ldg(a.guard) != 0.f) return;
  int k = 0;
  while (k + 1 < a.n_segs && (int)blockIdx.x >= a.s[k + 1].first_block) ++k;
  const AdamSeg& S = a.s[k];
  const int64_t lo = (int64_t)(blockIdx.x - S.first_block) * ADAM_TILE;
  const int64_t hi = lo + ADAM_TILE < S.n ? lo + ADAM_TILE : S.n;
  const bool vec = ((((uintptr_t)S.p | (uintptr_t)S.g | (uintptr_t)S.m | (uintptr_t)S.v) & 15) == 0);
  if (vec) {
    const int64_t hi4 = lo + ((hi - lo) & ~(int64_t)3);
    for (int64_t i = lo + threadIdx.x * 4; i < hi4; i += 256 * 4) {
      float4 p = *reinterpret_cast<float4*>(S.p + i);
      const float4 g = *reinterpret_cast<const float4*>(S.g + i);
      float4 m = *reinterpret_cast<float4*>(S.m + i);
      float4 v = *reinterpret_cast<float4*>(S.v + i);
      adam1(p.x, g.x, m.x, v.x, a, S.step_size);
      adam1(p.y, g.y, m.y, v.y, a, S.step_size);
      adam1(p.z, g.z, m.z, v.z, a, S.step_size);
      adam1(p.w, g.w, m.w, v.w, a, S.step_size);
      *reinterpret_cast<float4*>(S.p + i) = p;
      *reinterpret_cast<float4*>(S.m + i) = m;
      *reinterpret_cast<float4*>(S.v + i) = v;
    }
    for (int64_t i = hi4 + threadIdx.x; i < hi; i += 256) adam1(S.p[i], S.g[i], S.m[i], S.v[i], a, S.step_size);
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += 256) adam1(S.p[i], S.g[i], S.m[i], S.v[i], a, S.step_size);
  }
}

}  // namespace ftrf

using namespace ftrf;

extern "C" int ftrf_reg_factors(const FtrfFactorSet* fs, int model, float w_tv, float w_l1, const float* dev_scale,
                                double* loss_out, int with_lines, void* stream) {
  if (!fs) { set_error("reg_factors: null factor set"); return FTRF_E_ARG; }
  for (int k = 0; k < 3; ++k) {
    if (fs->n_comp[k] <= 0 || (fs->n_comp[k] & 3)) { set_error("reg_factors: n_comp[%d]=%d must be a positive multiple of 4", k, fs->n_comp[k]); return FTRF_E_ARG; }
    if (model == FTRF_MODEL_VM && !fs->plane[k]) { set_error("reg_factors: plane %d is NULL", k); return FTRF_E_ARG; }
    if (with_lines && !fs->line[k]) { set_error("reg_factors: line %d is NULL", k); return FTRF_E_ARG; }
  }
  if (model != FTRF_MODEL_VM && w_tv != 0.f) {
    // models/tensoRF.py:474-484 applies TVLoss to (1,C,L,1) lines: count_w = 0 and the value is 0/0 upstream
    set_error("reg_factors: TV of CP lines is NaN in the reference (loss.py:129, count_w = 0); not provided");
    return FTRF_E_ARG;
  }
  return launch_reg_factors(*fs, model, w_tv, w_l1, dev_scale, loss_out, with_lines, (cudaStream_t)stream);
}

extern "C" int ftrf_adam_step_guarded(const FtrfAdamSegment* segs, int n_segs, float beta1, float beta2, float eps, int step,
                                      const float* skip_if_nonzero, void* stream) {
  if (!segs || n_segs <= 0 || step < 1) { set_error("adam_step: bad argument"); return FTRF_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  // torch/optim/adam.py _single_tensor_adam: bias corrections on the host in double
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  for (int s0 = 0; s0 < n_segs; s0 += ADAM_MAX_SEGS) {
    AdamArgs a{};
    a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
    a.guard = skip_if_nonzero;
    int blocks = 0;
    const int n = n_segs - s0 < ADAM_MAX_SEGS ? n_segs - s0 : ADAM_MAX_SEGS;
    int used = 0;
    for (int i = 0; i < n; ++i) {
      const FtrfAdamSegment& in = segs[s0 + i];
      if (in.n < 0 || (in.n > 0 && (!in.param || !in.grad || !in.exp_avg || !in.exp_avg_sq))) {
        set_error("adam_step: segment %d has a NULL pointer", s0 + i);
        return FTRF_E_ARG;
      }
      if (in.n == 0) continue;
      AdamSeg& o = a.s[used++];
      o.p = in.param; o.g = in.grad; o.m = in.exp_avg; o.v = in.exp_avg_sq; o.n = in.n;
      o.step_size = (float)((double)in.lr / bc1);
      o.first_block = blocks;
      blocks += (int)((in.n + ADAM_TILE - 1) / ADAM_TILE);
    }
    a.n_segs = used;
    if (!used) continue;
    adam_kernel<<<blocks, 256, 0, st>>>(a);
    if (int e = check_launch("adam_step")) return e;
  }
  return FTRF_OK;
}

extern "C" int ftrf_adam_step(const FtrfAdamSegment* segs, int n_segs, float beta1, float beta2, float eps, int step,
                              void* stream) {
  return ftrf_adam_step_guarded(segs, n_segs, beta1, beta2, eps, step, nullptr, stream);
}
