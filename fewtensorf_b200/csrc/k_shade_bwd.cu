// k_shade_bwd.cu — stage 3 backward (v1, fp32 CUDA cores): the autograd replay of the shading MLP
// (models/mlp.py:66-69 and friends; train.py:374-376) split into three kernels:
//   shade_bwd_hidden : d rgb -> sigmoid' -> layer 3 -> ReLU -> layer 2 -> ReLU ; writes g_h2, g_h1,
//                      accumulates dW3, db3, db2, db1
//   shade_bwd_input  : g_h1 -> layer 1 -> positional-encoding backward -> g_feat
//   shade_wgrad<L>   : dW2 = g_h2^T h1,  dW1 = g_h1^T X  (X rebuilt from feat, never stored)
// Weight gradients are reduced in registers over all tiles a CTA owns and flushed once.
#include "shade_common.cuh"

namespace ftrf {

template <int NB>   // NB 64-column halves: acc[4][4*NB] += A[64][K] x B[K][64*NB]
__device__ __forceinline__ void tile_gemm_n(const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                                            int K, float (*acc)[4 * NB], int ty, int tx) {
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
#pragma unroll
      for (int h = 0; h < NB; ++h) {
        float4 b = *reinterpret_cast<const float4*>(brow + h * 64 + tx * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float av = a[i][kk];
          acc[i][h * 4 + 0] = fmaf(av, b.x, acc[i][h * 4 + 0]); acc[i][h * 4 + 1] = fmaf(av, b.y, acc[i][h * 4 + 1]);
          acc[i][h * 4 + 2] = fmaf(av, b.z, acc[i][h * 4 + 2]); acc[i][h * 4 + 3] = fmaf(av, b.w, acc[i][h * 4 + 3]);
        }
      }
    }
  }
}

// load a [ST][128] row-major global tile into shared memory (row stride HS); rows >= n are zero
__device__ __forceinline__ void load_tile128(const float* __restrict__ g, int64_t t0, int64_t n, float* sm) {
  for (int e = threadIdx.x; e < ST * (FC / 4); e += blockDim.x) {
    int s = e / (FC / 4), q = e % (FC / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t0 + s < n) v = *reinterpret_cast<const float4*>(g + (size_t)(t0 + s) * FC + q * 4);
    *reinterpret_cast<float4*>(sm + s * HS + q * 4) = v;
  }
}

__global__ void __launch_bounds__(256, 1)
shade_bwd_hidden_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ __align__(16) float smem[];
  float* W2 = smem;                 // [o][i] natural layout = k-major for g_h1 = g_h2 W2
  float* W3s = W2 + FC * FC;        // [3][128]
  float* T2 = W3s + 3 * FC;         // [ST][HS]  h2, then g_h2
  float* T1 = T2 + ST * HS;         // [ST][HS]  h1, then g_h1
  float* go = T1 + ST * HS;         // [ST][4]
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if ((int64_t)blockIdx.x * ST >= n) return;
  for (int e = threadIdx.x; e < FC * FC / 4; e += blockDim.x)
    reinterpret_cast<float4*>(W2)[e] = ld4(f.w2 + (size_t)e * 4);
  for (int e = threadIdx.x; e < 3 * FC; e += blockDim.x) W3s[e] = __ldg(f.w3 + e);

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float dw3[2] = {0.f, 0.f}, db3 = 0.f, db2 = 0.f, db1 = 0.f;

  for (int64_t t0 = (int64_t)blockIdx.x * ST; t0 < n; t0 += (int64_t)gridDim.x * ST) {
    __syncthreads();
    load_tile128(ws.a_h2, t0, n, T2);
    load_tile128(ws.a_h1, t0, n, T1);
    if (threadIdx.x < 3 * ST) {
      int s = threadIdx.x / 3, ch = threadIdx.x % 3;
      float v = 0.f;
      if (t0 + s < n) {
        float r = ws.a_rgb[(size_t)(t0 + s) * 3 + ch];
        v = ws.a_grgb[(size_t)(t0 + s) * 3 + ch] * r * (1.f - r);      // sigmoid backward
      }
      go[s * 4 + ch] = v;
    }
    __syncthreads();
    // dW3 / db3
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int e = threadIdx.x + q * 256;
      if (e < 3 * FC) {
        int ch = e / FC, o = e % FC;
        float a = 0.f;
        for (int s = 0; s < ST; ++s) a = fmaf(go[s * 4 + ch], T2[s * HS + o], a);
        dw3[q] += a;
      }
    }
    if (threadIdx.x < 3) {
      float a = 0.f;
      for (int s = 0; s < ST; ++s) a += go[s * 4 + threadIdx.x];
      db3 += a;
    }
    __syncthreads();
    // g_h2 = (h2 > 0) * go W3      (in place)
    for (int e = threadIdx.x; e < ST * FC; e += blockDim.x) {
      int s = e / FC, o = e % FC;
      float h = T2[s * HS + o];
      float g = h > 0.f ? go[s * 4] * W3s[o] + go[s * 4 + 1] * W3s[FC + o] + go[s * 4 + 2] * W3s[2 * FC + o] : 0.f;
      T2[s * HS + o] = g;
      if (t0 + s < n) ws.a_gh2[(size_t)(t0 + s) * FC + o] = g;
    }
    __syncthreads();
    if (threadIdx.x < FC) {
      float a = 0.f;
      for (int s = 0; s < ST; ++s) a += T2[s * HS + threadIdx.x];
      db2 += a;
    }
    // g_h1 = (h1 > 0) * g_h2 W2    (in place over the h1 tile)
    float acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    tile_gemm<false>(T2, HS, W2, FC, FC, acc, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int s = ty * 4 + i;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float* p = T1 + s * HS + h * 64 + tx * 4;
        float4 hv = *reinterpret_cast<float4*>(p);
        float4 g = make_float4(hv.x > 0.f ? acc[i][h * 4] : 0.f, hv.y > 0.f ? acc[i][h * 4 + 1] : 0.f,
                               hv.z > 0.f ? acc[i][h * 4 + 2] : 0.f, hv.w > 0.f ? acc[i][h * 4 + 3] : 0.f);
        *reinterpret_cast<float4*>(p) = g;
        if (t0 + s < n) *reinterpret_cast<float4*>(ws.a_gh1 + (size_t)(t0 + s) * FC + h * 64 + tx * 4) = g;
      }
    }
    __syncthreads();
    if (threadIdx.x < FC) {
      float a = 0.f;
      for (int s = 0; s < ST; ++s) a += T1[s * HS + threadIdx.x];
      db1 += a;
    }
  }
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    int e = threadIdx.x + q * 256;
    if (e < 3 * FC) atomicAdd(f.g_w3 + e, dw3[q]);
  }
  if (threadIdx.x < 3) atomicAdd(f.g_b3 + threadIdx.x, db3);
  if (threadIdx.x < FC) { atomicAdd(f.g_b2 + threadIdx.x, db2); atomicAdd(f.g_b1 + threadIdx.x, db1); }
}

// g_X = g_h1 W1, then PE backward into g_feat.  W1 natural [o][in] is k-major for this product.
__global__ void __launch_bounds__(256, 1)
shade_bwd_input_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp) {
  extern __shared__ __align__(16) float smem[];
  const int LDW = 192, KA = LDW + 4;
  float* W1 = smem;                  // [128][192] (columns >= in_dim are zero)
  float* G1 = W1 + FC * LDW;         // [ST][HS]
  float* GX = G1 + ST * HS;          // [ST][KA]
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if ((int64_t)blockIdx.x * ST >= n) return;
  for (int e = threadIdx.x; e < FC * LDW; e += blockDim.x) {
    int o = e / LDW, k = e % LDW;
    W1[e] = k < sp.in_dim ? __ldg(f.w1 + (size_t)o * sp.in_dim + k) : 0.f;
  }
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4, D = sp.D;
  for (int64_t t0 = (int64_t)blockIdx.x * ST; t0 < n; t0 += (int64_t)gridDim.x * ST) {
    __syncthreads();
    load_tile128(ws.a_gh1, t0, n, G1);
    __syncthreads();
    {
      float acc[4][8];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      tile_gemm_n<2>(G1, HS, W1, LDW, FC, acc, ty, tx);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          *reinterpret_cast<float4*>(GX + (ty * 4 + i) * KA + h * 64 + tx * 4) =
              make_float4(acc[i][h * 4], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
    }
    if (sp.kp > 128) {
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
      tile_gemm_n<1>(G1, HS, W1 + 128, LDW, FC, acc, ty, tx);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(GX + (ty * 4 + i) * KA + 128 + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // PE backward: g_feat[ch] = gX[ch] + sum over PE(fea) blocks
    for (int e = threadIdx.x; e < ST * D; e += blockDim.x) {
      int s = e / D, ch = e % D;
      if (t0 + s >= n) continue;
      const float* gx = GX + s * KA;
      float g = gx[ch];
      for (int b = 0; b < sp.nblk; ++b) {
        const PeBlock& pb = sp.b[b];
        if (pb.src != SRC_FEA) continue;
        float v = ws.a_feat[(size_t)(t0 + s) * D + ch];
        int half = pb.F * pb.C;
        for (int fq = 0; fq < pb.F; ++fq) {
          float sc = (float)(1 << fq), sn, cs;
          sincosf(v * sc, &sn, &cs);
          int q = ch * pb.F + fq;
          float ms = enc_mask_at(f, SRC_FEA, q), mc = enc_mask_at(f, SRC_FEA, half + q);
          g += sc * (cs * ms * gx[pb.col0 + q] - sn * mc * gx[pb.col0 + half + q]);
        }
      }
      ws.a_gfeat[(size_t)(t0 + s) * D + ch] = g;
    }
  }
}

// dW[o][k] += sum_s G[s][o] A[s][k].  LAYER 1: A = X (rebuilt), K = in_dim;  LAYER 2: A = h1, K = 128.
template <int LAYER>
__global__ void __launch_bounds__(256, 1)
shade_wgrad_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const InSpec sp) {
  constexpr int NJ = LAYER == 1 ? 3 : 2;            // 64-column chunks of K handled per thread
  extern __shared__ __align__(16) float smem[];
  const int K = LAYER == 1 ? sp.in_dim : FC;
  const int KA = LAYER == 1 ? 196 : HS;
  float* G = smem;                 // [ST][HS]
  float* A = G + ST * HS;          // [ST][KA]
  float* stage = A + ST * KA;      // layer 1 only: ST*(D+6)
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if ((int64_t)blockIdx.x * ST >= n) return;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[8][4 * NJ];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4 * NJ; ++j) acc[i][j] = 0.f;
  if (LAYER == 1) {
    for (int e = threadIdx.x; e < ST * KA; e += blockDim.x) A[e] = 0.f;   // columns >= kp stay zero
  }
  for (int64_t t0 = (int64_t)blockIdx.x * ST; t0 < n; t0 += (int64_t)gridDim.x * ST) {
    __syncthreads();
    load_tile128(LAYER == 1 ? ws.a_gh1 : ws.a_gh2, t0, n, G);
    if (LAYER == 1) build_x_tile(f, c, ws, sp, t0, n, stage, A, KA);
    else load_tile128(ws.a_h1, t0, n, A);
    __syncthreads();
    for (int s = 0; s < ST; ++s) {
      float4 g0 = *reinterpret_cast<const float4*>(G + s * HS + ty * 8);
      float4 g1 = *reinterpret_cast<const float4*>(G + s * HS + ty * 8 + 4);
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float4 a = *reinterpret_cast<const float4*>(A + s * KA + j * 64 + tx * 4);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          acc[i][j * 4 + 0] = fmaf(g[i], a.x, acc[i][j * 4 + 0]); acc[i][j * 4 + 1] = fmaf(g[i], a.y, acc[i][j * 4 + 1]);
          acc[i][j * 4 + 2] = fmaf(g[i], a.z, acc[i][j * 4 + 2]); acc[i][j * 4 + 3] = fmaf(g[i], a.w, acc[i][j * 4 + 3]);
        }
      }
    }
  }
  float* dst = LAYER == 1 ? f.g_w1 : f.g_w2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int o = ty * 8 + i;
#pragma unroll
    for (int j = 0; j < NJ; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        int k = j * 64 + tx * 4 + q;
        if (k < K) atomicAdd(dst + (size_t)o * K + k, acc[i][j * 4 + q]);
      }
  }
}

int shade_check_v1(const FtrfField& f, const InSpec& sp);
int launch_shade_sh_bwd(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t);

int launch_shade_bwd_v1(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  InSpec sp = make_spec(f);
  if (int e = shade_check_v1(f, sp)) return e;
  if (sp.kp > 192) { set_error("shade backward: input width %d > 192 unsupported", sp.in_dim); return FTRF_E_ARG; }
  size_t sm_h = (size_t)(FC * FC + 3 * FC + 2 * ST * HS + ST * 4) * sizeof(float);
  size_t sm_i = (size_t)(FC * 192 + ST * HS + ST * 196) * sizeof(float);
  size_t sm_w1 = (size_t)(ST * HS + ST * 196 + ST * (sp.D + 6)) * sizeof(float);
  size_t sm_w2 = (size_t)(ST * HS + ST * HS) * sizeof(float);
  cudaFuncSetAttribute(shade_bwd_hidden_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_h);
  cudaFuncSetAttribute(shade_bwd_input_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_i);
  cudaFuncSetAttribute(shade_wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_w1);
  cudaFuncSetAttribute(shade_wgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_w2);
  prof_begin(7, st);
  shade_bwd_hidden_kernel<<<FTRF_SMS, 256, sm_h, st>>>(f, c, ws);
  if (int e = check_launch("shade_bwd_hidden")) return e;
  prof_end(7, st); prof_begin(8, st);
  shade_bwd_input_kernel<<<FTRF_SMS, 256, sm_i, st>>>(f, c, ws, sp);
  if (int e = check_launch("shade_bwd_input")) return e;
  prof_end(8, st); prof_begin(9, st);
  shade_wgrad_kernel<2><<<FTRF_SMS, 256, sm_w2, st>>>(f, c, ws, sp);
  if (int e = check_launch("shade_wgrad2")) return e;
  prof_end(9, st); prof_begin(10, st);
  shade_wgrad_kernel<1><<<FTRF_SMS, 256, sm_w1, st>>>(f, c, ws, sp);
  int e = check_launch("shade_wgrad1");
  prof_end(10, st);
  return e;
}

}  // namespace ftrf
