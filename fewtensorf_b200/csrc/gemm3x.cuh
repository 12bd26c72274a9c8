// gemm3x.cuh — argument blocks of the tcgen05 3xTF32 contractions (k_gemm3x.cu, k_wgrad3x.cu).
#pragma once
#include "common.cuh"

namespace ftrf {

// out[M][N] = epi(A[M][K] * B[N][K]^T)
struct Gemm3xArgs {
  const float* A; int lda; int ka;       // A[m][k], k < ka valid (ka % 4 == 0), others read as 0
  const int32_t* count; int64_t cap;     // M = count ? min(*count, cap) : cap
  const float* B; int64_t sn, sk;        // B element (n,k) at B[n*sn + k*sk]
  int N, K;
  const float* bias;                     // [N] or null
  const float* mask; int ldm;            // out = mask[m][n] > 0 ? acc : 0   (ReLU backward) or null
  int relu;                              // out = max(acc, 0)
  float* out; int ldo;
  int variant;                           // debug: bit0 swaps the LBO/SBO roles
  int accumulate;                        // acc += out[m][n] before bias / ReLU / mask (second half of a K-split layer)
};

// dW[o][k] += sum_s G[s][o] * A[s][k]   (o < 128, k < K);  dbias[o] += sum_s G[s][o] if dbias
struct Wgrad3xArgs {
  const float* G; int ldg;               // [M][128]
  const float* A; int lda; int K;        // [M][lda], columns k < K used
  const int32_t* count; int64_t cap;
  float* dW; int ldw;                    // [128][ldw]
  float* dbias;                          // [128] or null
  int use_perm;                          // dW column of input column k is perm[k] (-1: dropped)
  int16_t perm[192];
};

__host__ __device__ inline int g_np(int N) { return (N + 15) & ~15; }
__host__ __device__ inline int g_kp(int K) { return (K + 7) & ~7; }

int launch_gemm3x(const Gemm3xArgs& a, cudaStream_t st);
int launch_wgrad3x(const Wgrad3xArgs& a, cudaStream_t st);

}  // namespace ftrf
