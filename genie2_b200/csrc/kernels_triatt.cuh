// K8 triangle attention (modules/triangular_attention.py:67-130, primitives.py:222-281), fp32 CUDA
// cores.  Off in the base configuration (genie/config.py:62); specialised for 4 heads x 32.
// Pipeline per module (api.cu run_triatt): LayerNorm rows -> one GEMM for q|k|v|g (512 cols) and one for
// the triangle bias (4 cols) -> triatt_core_kernel (this file) -> output GEMM with residual.
// The ending-node variant is the same computation on the transposed pair map
// (triangular_attention.py:84-86,127-128); it is done by index mapping, no transposed copy exists:
//   pix(i, j) = starting ? (i, j) : (j, i)     bias(j, k) read at starting ? (j, k) : (k, j).
#pragma once
#include "common.cuh"

#define TA_H 4
#define TA_C 32

// One CTA per (row i of the possibly transposed map, sample b); thread-per-query online softmax over
// the keys of that row, K / V of one head in shared memory.
//   qkvg [B*N*N][512]: q | k | v | sigmoid-input g, each [H][C];  tbias [B*N*N][4];  og [B*N*N][128]
__global__ void __launch_bounds__(256) triatt_core_kernel(const float* __restrict__ qkvg, const float* __restrict__ tbias,
                                                          const int* __restrict__ rmask, float* __restrict__ og, int N,
                                                          int starting) {
  extern __shared__ __align__(16) float ta_sm[];
  float* Ks = ta_sm;                    // [N][32]
  float* Vs = ta_sm + (size_t)N * TA_C; // [N][32]
  float* mb = Vs + (size_t)N * TA_C;    // [N] mask bias of the keys
  const int b = blockIdx.y, i = blockIdx.x, tid = threadIdx.x;
  const size_t sample = (size_t)b * N * N;
  auto pix = [&](int r, int c) -> size_t { return sample + (starting ? (size_t)r * N + c : (size_t)c * N + r); };
  const int mi = rmask[b * N + i];
  for (int k = tid; k < N; k += 256) mb[k] = 1e9f * ((float)(mi * rmask[b * N + k]) - 1.f);
  const float scale = rsqrtf((float)TA_C);
  for (int h = 0; h < TA_H; ++h) {
    __syncthreads();
    for (int e = tid; e < N * (TA_C / 4); e += 256) {
      const int k = e >> 3, c4 = (e & 7) * 4;
      const float* row = qkvg + pix(i, k) * 512 + h * TA_C + c4;
      st4(Ks + k * TA_C + c4, ld4(row + 128));
      st4(Vs + k * TA_C + c4, ld4(row + 256));
    }
    __syncthreads();
    for (int j = tid; j < N; j += 256) {
      const size_t pj = pix(i, j);
      float q[TA_C], acc[TA_C];
#pragma unroll
      for (int c = 0; c < TA_C; c += 4) {
        const float4 t = ld4(qkvg + pj * 512 + h * TA_C + c);
        q[c] = t.x; q[c + 1] = t.y; q[c + 2] = t.z; q[c + 3] = t.w;
        acc[c] = acc[c + 1] = acc[c + 2] = acc[c + 3] = 0.f;
      }
      float m = -3.0e38f, l = 0.f;
      for (int k = 0; k < N; ++k) {
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < TA_C; c += 4) {
          const float4 kk = ld4(Ks + k * TA_C + c);
          s = fmaf(q[c], kk.x, s); s = fmaf(q[c + 1], kk.y, s); s = fmaf(q[c + 2], kk.z, s); s = fmaf(q[c + 3], kk.w, s);
        }
        const size_t pb = sample + (starting ? (size_t)j * N + k : (size_t)k * N + j);
        s = s * scale + mb[k] + tbias[pb * TA_H + h];
        const float mn = fmaxf(m, s), corr = expf(m - mn), pe = expf(s - mn);
        l = l * corr + pe;
        m = mn;
#pragma unroll
        for (int c = 0; c < TA_C; c += 4) {
          const float4 vv = ld4(Vs + k * TA_C + c);
          acc[c] = fmaf(pe, vv.x, acc[c] * corr); acc[c + 1] = fmaf(pe, vv.y, acc[c + 1] * corr);
          acc[c + 2] = fmaf(pe, vv.z, acc[c + 2] * corr); acc[c + 3] = fmaf(pe, vv.w, acc[c + 3] * corr);
        }
      }
      const float inv = 1.f / l;
#pragma unroll
      for (int c = 0; c < TA_C; c += 4) {
        const float4 gg = ld4(qkvg + pj * 512 + 384 + h * TA_C + c);
        float4 o;
        o.x = acc[c] * inv * sigmoidf_(gg.x); o.y = acc[c + 1] * inv * sigmoidf_(gg.y);
        o.z = acc[c + 2] * inv * sigmoidf_(gg.z); o.w = acc[c + 3] * inv * sigmoidf_(gg.w);
        st4(og + pj * 128 + h * TA_C + c, o);
      }
    }
  }
}

static inline size_t triatt_smem_bytes(int N) { return ((size_t)N * TA_C * 2 + N) * sizeof(float); }
static inline void triatt_set_attributes() {
  cudaFuncSetAttribute(triatt_core_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}
