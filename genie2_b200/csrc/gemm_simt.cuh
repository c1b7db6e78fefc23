// Generic fp32 CUDA-core GEMM for the small / odd-shaped products of the path
// (embedding tables, IPA projections, structure-net MLPs: M = B*N <= ~1k rows).
//   C[m,n] = act( sum_k A[m,k] * W[n,k] + bias[n] ) + R[m,n]
// A [M,K] (lda), W [N,K] (ldw) -- torch.nn.Linear layout, so no host-side transposes.
// 64x64x16 tiles, 256 threads, 4x4 outputs per thread, bounds-checked on every edge.
#pragma once
#include "common.cuh"

template <int ACT>  // 0 none, 1 relu
__global__ void __launch_bounds__(256) sgemm_nt_kernel(const float* __restrict__ A, int lda,
                                                       const float* __restrict__ W, int ldw,
                                                       const float* __restrict__ bias,
                                                       const float* __restrict__ R, int ldr,
                                                       float* __restrict__ C, int ldc, int M, int N, int K) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float As[BK][BM + 4];
  __shared__ float Ws[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // loader mapping: 64 rows x 16 k; thread -> (row = tid/4, k4 = (tid%4)*4)
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  for (int k0 = 0; k0 < K; k0 += BK) {
    {
      const int m = m0 + lr;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + lk + q;
        As[lk + q][lr] = (m < M && k < K) ? A[(size_t)m * lda + k] : 0.f;
      }
      const int n = n0 + lr;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + lk + q;
        Ws[lk + q][lr] = (n < N && k < K) ? W[(size_t)n * ldw + k] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? bias[n] : 0.f);
      if (ACT == 1) v = fmaxf(v, 0.f);
      if (R) v += R[(size_t)m * ldr + n];
      C[(size_t)m * ldc + n] = v;
    }
  }
}

static inline void sgemm_nt(cudaStream_t st, const float* A, int lda, const float* W, int ldw, const float* bias,
                            const float* R, int ldr, float* C, int ldc, int M, int N, int K, bool relu) {
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  if (relu)
    sgemm_nt_kernel<1><<<grid, 256, 0, st>>>(A, lda, W, ldw, bias, R, ldr, C, ldc, M, N, K);
  else
    sgemm_nt_kernel<0><<<grid, 256, 0, st>>>(A, lda, W, ldw, bias, R, ldr, C, ldc, M, N, K);
}
