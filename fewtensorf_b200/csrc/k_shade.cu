// k_shade.cu — TEST BUILD ONLY (libfewtensorf_b200_test.so): the first, fp32 CUDA-core forward of the shading MLP
// (models/mlp.py:27-155), kept as an independent cross-check of the tcgen05 kernels the product runs.
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

size_t shade_fwd_smem(const InSpec& sp) {
  int ka = sp.kp + 4 > HS ? sp.kp + 4 : HS;
  size_t fl = (size_t)sp.kp * FC + FC * FC + 3 * FC + 2 * FC + 4 + (size_t)ST * ka + (size_t)ST * HS;
  return fl * sizeof(float);
}

int shade_check(const FtrfField& f, const InSpec& sp);
// the CUDA-core kernels keep W1 and W2 whole in shared memory, which bounds the input width
int shade_check_v1(const FtrfField& f, const InSpec& sp) {
  if (int e = shade_check(f, sp)) return e;
  if (shade_fwd_smem(sp) > 227 * 1024) {
    set_error("shade: input width %d needs too much shared memory on the CUDA-core back end (use the tensor-core one)", sp.in_dim);
    return FTRF_E_ARG;
  }
  return FTRF_OK;
}

int launch_shade_v1(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  InSpec sp = make_spec(f);
  if (int e = shade_check_v1(f, sp)) return e;
  size_t smem = shade_fwd_smem(sp);
  cudaFuncSetAttribute(shade_mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  shade_mlp_fwd_kernel<<<FTRF_SMS, 256, smem, st>>>(f, c, ws, sp);
  return check_launch("shade_mlp_fwd");
}

}  // namespace ftrf
