// k_shade_tc.cu — stage 3 on the tensor cores: the shading MLP as tcgen05 3xTF32 contractions
// (k_gemm3x.cu / k_wgrad3x.cu) plus the thin CUDA-core kernels around them:
//   build_x      positional encoding + FreeNeRF encoding mask -> X[V_a][ldx]   (models/mlp.py:8-13, :41-64)
//   layer3_fwd   rgb = sigmoid(h2 W3^T + b3)                                    (models/mlp.py:66-67)
//   layer3_bwd   sigmoid' , g_h2 = relu'(h2) * (go W3), dW3, db3
//   pe_bwd       g_feat from g_X (chain rule through sin/cos and the mask)
// Forward:  build_x -> gemm(X,W1)+b1,ReLU -> gemm(h1,W2)+b2,ReLU -> layer3
// Backward: layer3_bwd -> gemm(g_h2,W2^T)*relu'(h1) -> gemm(g_h1,W1^T) -> pe_bwd ; wgrad(g_h2,h1), wgrad(g_h1,X)
#include "shade_common.cuh"
#include "gemm3x.cuh"
#include "fused.cuh"

namespace ftrf {

// ---------------------------------------------------------------------------------------- build_x
// Per-column recipe, built once per CTA: where the value comes from and what is applied to it.
struct XCol {
  int src;       // offset into the staged [feat(D) | view(3) | pos(3)] row, or -1 for a padding column
  int mode;      // 0 copy, 1 sin, 2 cos
  float scale;   // 2^f
  float mask;    // FreeNeRF encoding-mask value of this column (1 if none)
};

__global__ void __launch_bounds__(256)
build_x_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp, int ldx) {
  extern __shared__ __align__(16) float stage[];      // ST * (D + 6) floats, then XCol[ldx]
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int D = sp.D, RW = D + 6;
  XCol* cols = reinterpret_cast<XCol*>(stage + ST * RW);
  for (int col = threadIdx.x; col < ldx; col += blockDim.x) {
    XCol x{-1, 0, 1.f, 1.f};
    if (col < D) x.src = col;
    else if (col < D + 3) x.src = col;                  // view follows feat in the staged row
    else
      for (int b = 0; b < sp.nblk; ++b) {
        const PeBlock& pb = sp.b[b];
        int e = col - pb.col0, half = pb.F * pb.C;
        if (e >= 0 && e < 2 * half) {
          int q = e >= half ? e - half : e;
          int ch = q / pb.F, fq = q - ch * pb.F;
          x.src = (pb.src == SRC_FEA ? 0 : (pb.src == SRC_VIEW ? D : D + 3)) + ch;
          x.mode = e >= half ? 2 : 1;
          x.scale = (float)(1 << fq);
          x.mask = enc_mask_at(f, pb.src, e);
        }
      }
    cols[col] = x;
  }
  for (int64_t t0 = (int64_t)blockIdx.x * ST; t0 < n; t0 += (int64_t)gridDim.x * ST) {
    __syncthreads();
    for (int e = threadIdx.x; e < ST * D; e += blockDim.x) {
      int s = e / D, ch = e - s * D;
      stage[s * RW + ch] = t0 + s < n ? ws.a_feat[(t0 + s) * D + ch] : 0.f;
    }
    if (threadIdx.x < ST) {
      int64_t s = t0 + threadIdx.x;
      float v[3] = {0.f, 0.f, 0.f}, p[3] = {0.f, 0.f, 0.f};
      if (s < n) shade_geometry(f, c, ws, s, v, p);
#pragma unroll
      for (int k = 0; k < 3; ++k) { stage[threadIdx.x * RW + D + k] = v[k]; stage[threadIdx.x * RW + D + 3 + k] = p[k]; }
    }
    __syncthreads();
    const int rows = (int)min((int64_t)ST, n - t0);
    for (int e = threadIdx.x; e < rows * ldx; e += blockDim.x) {
      int s = e / ldx, col = e - s * ldx;
      XCol x = cols[col];
      float v = 0.f;
      if (x.src >= 0) {
        v = stage[s * RW + x.src];
        if (x.mode) {
          float arg = v * x.scale;
          v = (x.mode == 1 ? sinf(arg) : cosf(arg)) * x.mask;
        }
      }
      ws.a_x[(size_t)t0 * ldx + e] = v;
    }
  }
}

// ---------------------------------------------------------------------------------------- layer 3
// one warp per shaded sample: lane owns hidden units 4*lane .. 4*lane+3
__global__ void __launch_bounds__(256)
layer3_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int lane = threadIdx.x & 31;
  float4 w[3];
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) w[ch] = ld4(f.w3 + ch * FC + lane * 4);
  const float b0 = __ldg(f.b3), b1 = __ldg(f.b3 + 1), b2 = __ldg(f.b3 + 2);
  const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t s = wid; s < n; s += nw) {
    float4 h = *reinterpret_cast<const float4*>(ws.a_h2 + s * FC + lane * 4);
    float a[3];
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) a[ch] = h.x * w[ch].x + h.y * w[ch].y + h.z * w[ch].z + h.w * w[ch].w;
#pragma unroll
    for (int d = 16; d; d >>= 1)
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) a[ch] += __shfl_xor_sync(FULL_MASK, a[ch], d);
    if (lane < 3) {
      float v = lane == 0 ? a[0] + b0 : (lane == 1 ? a[1] + b1 : a[2] + b2);
      ws.a_rgb[s * 3 + lane] = 1.f / (1.f + expf(-v));
    }
  }
}

// g_h2, dW3, db3 (+ db2 = column sums of g_h2).  Per-lane partial sums over the warp's samples,
// then a block reduction in shared memory and one atomic flush per CTA.
__global__ void __launch_bounds__(256)
layer3_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  __shared__ float red[8][16 * 32];       // [warp][12 dW3 + 4 db2 per lane]
  __shared__ float red3[8][3];
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 w[3];
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) w[ch] = ld4(f.w3 + ch * FC + lane * 4);
  float dw[3][4] = {}, db2[4] = {}, db3[3] = {};
  const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // four rows per iteration: their loads are all issued before the first use (one 512-byte row in flight
  // per warp left the kernel latency-bound at 2 TB/s)
  constexpr int U = 4;
  for (int64_t s0 = wid; s0 < n; s0 += nw * U) {
    float4 h[U];
    float r[U][3], gr[U][3];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t s = s0 + (int64_t)u * nw;
      if (s < n) {
        h[u] = *reinterpret_cast<const float4*>(ws.a_h2 + s * FC + lane * 4);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) { r[u][ch] = ws.a_rgb[s * 3 + ch]; gr[u][ch] = ws.a_grgb[s * 3 + ch]; }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t s = s0 + (int64_t)u * nw;
      if (s >= n) continue;
      float go[3];
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        go[ch] = gr[u][ch] * r[u][ch] * (1.f - r[u][ch]);
        db3[ch] += go[ch];
      }
      const float hh[4] = {h[u].x, h[u].y, h[u].z, h[u].w};
      const float ww[3][4] = {{w[0].x, w[0].y, w[0].z, w[0].w}, {w[1].x, w[1].y, w[1].z, w[1].w}, {w[2].x, w[2].y, w[2].z, w[2].w}};
      float g[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[j] = hh[j] > 0.f ? go[0] * ww[0][j] + go[1] * ww[1][j] + go[2] * ww[2][j] : 0.f;
        db2[j] += g[j];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) dw[ch][j] = fmaf(go[ch], hh[j], dw[ch][j]);
      }
      *reinterpret_cast<float4*>(ws.a_gh2 + s * FC + lane * 4) = make_float4(g[0], g[1], g[2], g[3]);
    }
  }
#pragma unroll
  for (int ch = 0; ch < 3; ++ch)
#pragma unroll
    for (int j = 0; j < 4; ++j) red[warp][(ch * 4 + j) * 32 + lane] = dw[ch][j];
#pragma unroll
  for (int j = 0; j < 4; ++j) red[warp][(12 + j) * 32 + lane] = db2[j];
  if (lane == 0) { red3[warp][0] = db3[0]; red3[warp][1] = db3[1]; red3[warp][2] = db3[2]; }   // same for all lanes
  __syncthreads();
  for (int e = threadIdx.x; e < 16 * 32; e += blockDim.x) {
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) v += red[q][e];
    int item = e >> 5, ln = e & 31;          // item < 12: dW3[ch][4*ln + j]; else db2[4*ln + j]
    if (v != 0.f) {
      if (item < 12) atomicAdd(f.g_w3 + (item >> 2) * FC + ln * 4 + (item & 3), v);
      else atomicAdd(f.g_b2 + ln * 4 + (item - 12), v);
    }
  }
  if (threadIdx.x < 3) {
    float v = 0.f;
    for (int q = 0; q < 8; ++q) v += red3[q][threadIdx.x];
    if (v != 0.f) atomicAdd(f.g_b3 + threadIdx.x, v);
  }
}

// ---------------------------------------------------------------------------------------- pe_bwd
// g_feat[ch] = gX[ch] + sum over PE(fea) blocks of 2^f * (cos(a) m_s gX[sin col] - sin(a) m_c gX[cos col])
__global__ void __launch_bounds__(256)
pe_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int D = sp.D, ldx = ws.ldx;
  const int64_t total = n * D;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t s = e / D;
    int ch = (int)(e - s * D);
    const float* gx = ws.a_gx + s * ldx;
    float g = gx[ch];
    for (int b = 0; b < sp.nblk; ++b) {
      const PeBlock& pb = sp.b[b];
      if (pb.src != SRC_FEA) continue;
      float v = ws.a_feat[e];
      int half = pb.F * pb.C;
      for (int fq = 0; fq < pb.F; ++fq) {
        float sc = (float)(1 << fq), sn, cs;
        sincosf(v * sc, &sn, &cs);
        int q = ch * pb.F + fq;
        float ms = enc_mask_at(f, SRC_FEA, q), mc = enc_mask_at(f, SRC_FEA, half + q);
        g += sc * (cs * ms * gx[pb.col0 + q] - sn * mc * gx[pb.col0 + half + q]);
      }
    }
    ws.a_gfeat[e] = g;
  }
}

// same chain rule for X rows in the fused kernels' internal column order (fused.cuh)
__global__ void __launch_bounds__(256)
pe_bwd_internal_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp, int icol0, int icol1, int icol2) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  const int D = sp.D, ldx = ws.ldx;
  const int icol[3] = {icol0, icol1, icol2};
  const int64_t total = n * D;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t s = e / D;
    int ch = (int)(e - s * D);
    const float* gx = ws.a_gx + s * ldx;
    float g = gx[ch];
    for (int b = 0; b < sp.nblk; ++b) {
      const PeBlock& pb = sp.b[b];
      if (pb.src != SRC_FEA) continue;
      float v = ws.a_feat[e];
      int half = pb.F * pb.C;
      for (int fq = 0; fq < pb.F; ++fq) {
        float sc = (float)(1 << fq), sn, cs;
        sincosf(v * sc, &sn, &cs);
        int q = ch * pb.F + fq;
        float ms = enc_mask_at(f, SRC_FEA, q), mc = enc_mask_at(f, SRC_FEA, half + q);
        g += sc * (cs * ms * gx[icol[b] + 2 * q] - sn * mc * gx[icol[b] + 2 * q + 1]);
      }
    }
    ws.a_gfeat[e] = g;
  }
}

// ---------------------------------------------------------------------------------------- drivers
int shade_check(const FtrfField& f, const InSpec& sp);
int launch_shade_v1(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_shade_bwd_v1(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_shade_sh(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);
int launch_shade_sh_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);

// 2 = fused tcgen05 kernels (per-layer tcgen05 GEMMs for the shapes they do not cover); the test build can also
// force 1 = one tcgen05 GEMM per layer, 0 = fp32 CUDA cores (k_shade.cu / k_shade_bwd.cu, cross-checks only)
FTRF_SWITCH(int, g_shade_path, 2);

static Gemm3xArgs gemm_args(const Ws& ws, const FtrfCall& c) {
  Gemm3xArgs g{};
  g.count = ws.counters + 1;
  g.cap = c.cap_app;
  return g;
}

size_t gemm3x_smem(int N, int K);
constexpr size_t SMEM_MAX = 227 * 1024;

// A layer whose resident weight operand does not fit shared memory (first layer of a wide input: MLP shading
// with pos_pe 8 is 198 columns, configs/free_tensorf_human.txt) runs as two launches.
//   split along K (forward): the second launch adds the first one's partial sums in its epilogue
static int launch_gemm3x_ksplit(Gemm3xArgs g, cudaStream_t st) {
  if (gemm3x_smem(g.N, g.K) <= SMEM_MAX) return launch_gemm3x(g, st);
  const int k1 = ((g.K / 2) + 31) & ~31;             // whole A chunks in the first half
  Gemm3xArgs a = g, b = g;
  a.K = k1; a.ka = k1; a.bias = nullptr; a.relu = 0; a.mask = nullptr;
  if (int e = launch_gemm3x(a, st)) return e;
  b.A = g.A + k1; b.ka = g.ka - k1; b.B = g.B + (int64_t)k1 * g.sk; b.K = g.K - k1; b.accumulate = 1;
  return launch_gemm3x(b, st);
}
//   split along N (dgrad into the input columns): two independent column ranges of the output
static int launch_gemm3x_nsplit(Gemm3xArgs g, cudaStream_t st) {
  if (gemm3x_smem(g.N, g.K) <= SMEM_MAX) return launch_gemm3x(g, st);
  const int n1 = ((g.N / 2) + 15) & ~15;
  Gemm3xArgs a = g, b = g;
  a.N = n1;
  if (int e = launch_gemm3x(a, st)) return e;
  b.B = g.B + (int64_t)n1 * g.sn; b.N = g.N - n1; b.out = g.out + n1;
  if (g.bias) b.bias = g.bias + n1;
  if (g.mask) b.mask = g.mask + n1;
  return launch_gemm3x(b, st);
}
//   weight gradient of a wide layer: two column ranges of dW (the bias gradient rides with the first)
static int launch_wgrad3x_split(Wgrad3xArgs w, cudaStream_t st) {
  if (g_np(w.K + (w.dbias ? 1 : 0)) <= 192) return launch_wgrad3x(w, st);
  const int k1 = ((w.K / 2) + 15) & ~15;
  Wgrad3xArgs a = w, b = w;
  a.K = k1;
  if (w.use_perm) return launch_wgrad3x(w, st);      // (the permuted internal layout never gets this wide)
  if (int e = launch_wgrad3x(a, st)) return e;
  b.A = w.A + k1; b.K = w.K - k1; b.dW = w.dW + k1; b.dbias = nullptr;
  return launch_wgrad3x(b, st);
}

bool shade_path_is_fused() { return g_shade_path == 2; }

int launch_shade(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (f.shading == FTRF_SHADE_SH) return launch_shade_sh(f, c, ws, st);
#ifdef FTRF_TEST_BUILD
  if (!g_shade_path) return launch_shade_v1(f, c, ws, st);
#endif
  InSpec sp = make_spec(f);
  if (int e = shade_check(f, sp)) return e;
  if (c.cap_app == 0) return FTRF_OK;
  if (g_shade_path == 2) {
    bool handled = false;
    if (int e = launch_shade_fused(f, c, ws, st, &handled)) return e;
    if (handled) return FTRF_OK;
  }
  size_t smem = (size_t)ST * (sp.D + 6) * sizeof(float) + (size_t)ws.ldx * 16;
  build_x_kernel<<<FTRF_SMS * 4, 256, smem, st>>>(f, c, ws, sp, ws.ldx);
  if (int e = check_launch("build_x")) return e;
  Gemm3xArgs g = gemm_args(ws, c);
  g.A = ws.a_x; g.lda = ws.ldx; g.ka = ws.ldx; g.B = f.w1; g.sn = sp.in_dim; g.sk = 1; g.N = FC; g.K = sp.in_dim;
  g.bias = f.b1; g.relu = 1; g.out = ws.a_h1; g.ldo = FC;
  if (int e = launch_gemm3x_ksplit(g, st)) return e;
  g.A = ws.a_h1; g.lda = FC; g.ka = FC; g.B = f.w2; g.sn = FC; g.sk = 1; g.N = FC; g.K = FC;
  g.bias = f.b2; g.out = ws.a_h2;
  if (int e = launch_gemm3x(g, st)) return e;
  layer3_fwd_kernel<<<FTRF_SMS * 8, 256, 0, st>>>(f, c, ws);
  return check_launch("layer3_fwd");
}

// part 0: everything.  part 1: down to g_feat (what the appearance backward waits for).  part 2: the MLP
// weight gradients, which nothing downstream reads.  Only the fused path splits; the other back ends do
// all their work in part 1.
int launch_shade_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st, int part) {
  if (f.shading == FTRF_SHADE_SH) return part == 2 ? FTRF_OK : launch_shade_sh_bwd(f, c, ws, st);
#ifdef FTRF_TEST_BUILD
  if (!g_shade_path) return part == 2 ? FTRF_OK : launch_shade_bwd_v1(f, c, ws, st);
#endif
  InSpec sp = make_spec(f);
  if (int e = shade_check(f, sp)) return e;
  if (c.cap_app == 0) return FTRF_OK;
  FusedSpec fs;
  const bool internal = g_shade_path == 2 && make_fused_spec(f, sp, fs) && fs.kp <= ws.ldx && ws.wimg;   // X saved by the fused forward
  const float* w1p = ws.wimg ? ws.wimg + fused_w1p_offset_floats() : nullptr;
  if (part == 2) {
    if (!internal) return FTRF_OK;
    prof_begin(9, st);
    WgradFusedArgs wa{};
    wa.g2 = ws.a_gh2; wa.h1 = ws.a_h1; wa.g1 = ws.a_gh1; wa.x = ws.a_x; wa.ldx = ws.ldx; wa.kp = fs.kp;
    wa.count = ws.counters + 1; wa.cap = c.cap_app; wa.dw2 = f.g_w2;
    if (shade_dgrad_is_fused(f, c, ws)) {     // that kernel formed g_h2 itself and left go4 where g_h2 used to go
      wa.fold3 = 1;
      wa.g2 = ws.a_h2;
      wa.go4 = reinterpret_cast<const float4*>(ws.a_gh2);
      wa.w3 = f.w3; wa.dw3 = f.g_w3; wa.db3 = f.g_b3; wa.db2 = f.g_b2;
    }
    wa.dw1i = ws.wimg + fused_scratch_offset_floats();
    int e = launch_wgrad_fused(wa, fs, f.g_w1, sp.in_dim, f.g_b1, st, &c);
    prof_end(9, st);
    return e;
  }
  // stage 7: output-layer backward as a kernel of its own — only where the fused dgrad kernel does not run (it forms
  // g_h2 in its prologue and the fused wgrad kernel takes dW3 / db3 / db2); stage 8: dgrad chain down to g_feat
  const bool fold3 = internal && shade_dgrad_is_fused(f, c, ws);
  if (!fold3) {
    prof_begin(7, st);
    layer3_bwd_kernel<<<FTRF_SMS * 2, 256, 0, st>>>(f, c, ws);
    if (int e = check_launch("layer3_bwd")) return e;
    prof_end(7, st);
  }
  prof_begin(8, st);
  bool fused_dgrad = false;
  if (internal) {
    if (int e = launch_shade_dgrad_fused(f, c, ws, st, &fused_dgrad)) return e;
  }
  Gemm3xArgs g = gemm_args(ws, c);
  if (!fused_dgrad) {
    // g_h1 = relu'(h1) * (g_h2 W2):  B(n=i,k=o) = W2[o][i]
    g.A = ws.a_gh2; g.lda = FC; g.ka = FC; g.B = f.w2; g.sn = 1; g.sk = FC; g.N = FC; g.K = FC;
    g.mask = ws.a_h1; g.ldm = FC; g.out = ws.a_gh1; g.ldo = FC;
    if (int e = launch_gemm3x(g, st)) return e;
  }
  if (!fused_dgrad) {
    g = gemm_args(ws, c);                   // g_X = g_h1 W1:  B(n=j,k=o) = W1[o][j]
    g.A = ws.a_gh1; g.lda = FC; g.ka = FC; g.B = f.w1; g.sn = 1; g.sk = sp.in_dim; g.N = sp.in_dim; g.K = FC;
    if (internal) { g.B = w1p; g.sk = fs.kp; g.N = fs.kp; }
    g.out = ws.a_gx; g.ldo = ws.ldx;
    if (int e = launch_gemm3x_nsplit(g, st)) return e;
    if (internal) pe_bwd_internal_kernel<<<FTRF_SMS * 8, 256, 0, st>>>(f, c, ws, sp, fs.blk_icol0[0], fs.blk_icol0[1], fs.blk_icol0[2]);
    else pe_bwd_kernel<<<FTRF_SMS * 8, 256, 0, st>>>(f, c, ws, sp);
    if (int e = check_launch("pe_bwd")) return e;
  }
  prof_end(8, st);
  if (internal) return part == 1 ? FTRF_OK : launch_shade_bwd(f, c, ws, st, 2);   // both weight gradients in one kernel (k_wgrad_fused.cu)
  prof_begin(9, st);
  Wgrad3xArgs w{};
  w.count = ws.counters + 1; w.cap = c.cap_app;
  w.G = ws.a_gh2; w.ldg = FC; w.A = ws.a_h1; w.lda = FC; w.K = FC; w.dW = f.g_w2; w.ldw = FC; w.dbias = nullptr;  // db2 done in layer3_bwd
  if (int e = launch_wgrad3x(w, st)) return e;
  prof_end(9, st); prof_begin(10, st);
  w.G = ws.a_gh1; w.A = ws.a_x; w.lda = ws.ldx; w.K = sp.in_dim; w.dW = f.g_w1; w.ldw = sp.in_dim; w.dbias = f.g_b1;
  if (internal) {
    w.K = fs.kp; w.use_perm = 1;
    for (int k = 0; k < fs.kp; ++k) w.perm[k] = fs.orig[k];
  }
  int e = launch_wgrad3x_split(w, st);
  prof_end(10, st);
  return e;
}

}  // namespace ftrf

#ifdef FTRF_TEST_BUILD
extern "C" int ftrf_set_shade_path(int tensor_cores) {
  ftrf::g_shade_path = tensor_cores < 0 ? 0 : (tensor_cores > 2 ? 2 : tensor_cores);
  return FTRF_OK;
}
#endif
