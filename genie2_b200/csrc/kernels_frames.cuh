// K1 frenet_frames, K11 ddpm_step (+frames), backbone update / frame compose.
#pragma once
#include "common.cuh"

// ---------------------------------------------------------------------------------
// Frenet-Serret frames of one sample, executed by one CTA (utils/geo_utils.py:21-85).
//   x      : smem or global coords of this sample [N,3]
//   chain  : [N], len = sum(mask)
//   raw    : smem scratch [N,9]   (the "rots_" array before the fix-up loops)
//   out    : global [N,9]
// tbn[k] has COLUMNS t_{k+1}, b_k, n_k (geo_utils.py:60) and lands at residue k+1
// (geo_utils.py:67).  The two sequential fix-up loops (:69-77) are evaluated in parallel:
// the start loop only ever reads entries it has not yet written (index j+1 > j), so
// start[j] = raw[j+1]; the end loop reads index j-1 (python wrap at j = 0), which may
// itself have been rewritten by the end loop when j-1 is also a chain end (1-residue
// chains) -- resolved by walking back.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void frenet_tbn(const float* xm, const float* x0, const float* xp, float* r, float eps) {
  // residue k+1 from x_k (xm), x_{k+1} (x0), x_{k+2} (xp): t_k = unit(x0 - xm), t_{k+1} = unit(xp - x0)
  float t0[3], t1[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { t0[c] = x0[c] - xm[c]; t1[c] = xp[c] - x0[c]; }
  float n0 = sqrtf(eps + t0[0] * t0[0] + t0[1] * t0[1] + t0[2] * t0[2]);
  float n1 = sqrtf(eps + t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2]);
#pragma unroll
  for (int c = 0; c < 3; ++c) { t0[c] = t0[c] / n0; t1[c] = t1[c] / n1; }
  float b[3] = {t0[1] * t1[2] - t0[2] * t1[1], t0[2] * t1[0] - t0[0] * t1[2], t0[0] * t1[1] - t0[1] * t1[0]};
  float nb = sqrtf(eps + b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
#pragma unroll
  for (int c = 0; c < 3; ++c) b[c] = b[c] / nb;
  float n[3] = {b[1] * t1[2] - b[2] * t1[1], b[2] * t1[0] - b[0] * t1[2], b[0] * t1[1] - b[1] * t1[0]};
#pragma unroll
  for (int c = 0; c < 3; ++c) { r[c * 3 + 0] = t1[c]; r[c * 3 + 1] = b[c]; r[c * 3 + 2] = n[c]; }
}

__device__ __forceinline__ bool chain_start(const int* ch, int j) { return j == 0 || ch[j] != ch[j - 1]; }
__device__ __forceinline__ bool chain_end(const int* ch, int j, int len) { return j == len - 1 || ch[j] != ch[j + 1]; }

// value of rots_[q] after the START loop
__device__ __forceinline__ void frenet_after_start(const float* raw, const int* ch, int len, int N, int q, float* r) {
  int src = q;
  if (q < len && chain_start(ch, q)) src = q + 1;   // reads raw (unmodified) entry j+1
  if (src >= N) src = N - 1;                          // reference would raise IndexError here
#pragma unroll
  for (int e = 0; e < 9; ++e) r[e] = raw[src * 9 + e];
}

__device__ void frenet_block(const float* x, const int* ch, int len, int N, float* raw, float* out) {
  const float eps = 1e-10f;
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    float r[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (j >= 1 && j < len - 1) frenet_tbn(x + (j - 1) * 3, x + j * 3, x + (j + 1) * 3, r, eps);
#pragma unroll
    for (int e = 0; e < 9; ++e) raw[j * 9 + e] = r[e];
  }
  __syncthreads();
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    float r[9];
    if (j < len && chain_end(ch, j, len)) {
      // end loop: rots_[j] = rots_[j-1] (current value).  Walk back over consecutive chain ends.
      int q = j - 1;
      if (q < 0) q += N;                               // python negative index
      while (q < j && q >= 1 && q < len && chain_end(ch, q, len)) q -= 1;
      if (q == 0 && q < j && q < len && chain_end(ch, 0, len)) q = N - 1;   // rots_[0] <- rots_[-1]
      frenet_after_start(raw, ch, len, N, q, r);
    } else {
      frenet_after_start(raw, ch, len, N, j, r);
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) out[j * 9 + e] = r[e];
  }
}

__global__ void frenet_kernel(const float* __restrict__ trans, const int* __restrict__ chain,
                              const int* __restrict__ mask, int N, float* __restrict__ rots) {
  extern __shared__ float sm[];
  float* xs = sm;                 // [N*3]
  float* raw = sm + N * 3;        // [N*9]
  int* chs = reinterpret_cast<int*>(raw + N * 9);
  __shared__ int s_len;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) s_len = 0;
  __syncthreads();
  int loc = 0;
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    loc += mask[b * N + j];
    chs[j] = chain[b * N + j];
#pragma unroll
    for (int c = 0; c < 3; ++c) xs[j * 3 + c] = trans[((size_t)b * N + j) * 3 + c];
  }
  atomicAdd(&s_len, loc);
  __syncthreads();
  frenet_block(xs, chs, s_len, N, raw, rots + (size_t)b * N * 9);
}

// ---------------------------------------------------------------------------------
// K11: one reverse step (sampler/base.py:249-272) fused with the frames of x_{t-1}
// (:275-279).  One CTA per sample.  Schedule tables indexed by the 1-based step.
// noise: the N(0,1) draw (the host's torch.randn buffer or the injected tape, indexed
// tape[step]); nullptr -> drawn in-kernel with Philox keyed by (seed, step, element).
// step is read from *step_dev when step_dev != nullptr (CUDA-graph replay).
// ---------------------------------------------------------------------------------
__global__ void ddpm_step_kernel(float* __restrict__ trans, float* __restrict__ rots, const float* __restrict__ z,
                                 const float* __restrict__ noise, const int* __restrict__ chain,
                                 const int* __restrict__ mask, const float* __restrict__ alphas,
                                 const float* __restrict__ sqrt_alphas, const float* __restrict__ sqrt_omac,
                                 const float* __restrict__ sqrt_betas, int step_arg, const int* __restrict__ step_dev,
                                 size_t noise_step_stride, float scale, uint64_t seed, int N) {
  extern __shared__ float sm[];
  float* xs = sm;
  float* raw = sm + N * 3;
  int* chs = reinterpret_cast<int*>(raw + N * 9);
  __shared__ int s_len;
  const int b = blockIdx.x;
  const int step = step_dev ? *step_dev : step_arg;
  if (threadIdx.x == 0) s_len = 0;
  __syncthreads();
  const float alpha = alphas[step];
  const float w_z = (1.f - alpha) / sqrt_omac[step];
  const float inv_sa = 1.f / sqrt_alphas[step];
  const float sigma = sqrt_betas[step];
  const float* nz = noise ? noise + (size_t)step * noise_step_stride : nullptr;
  int loc = 0;
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    const int m = mask[b * N + j];
    loc += m;
    chs[j] = chain[b * N + j];
    const float mf = (float)m;
    float gen[3] = {0.f, 0.f, 0.f};
    if (!nz && step != 1) normal3(seed, (uint32_t)step, (uint32_t)(b * N + j), gen);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const size_t o = ((size_t)b * N + j) * 3 + c;
      float mean = inv_sa * (trans[o] - w_z * z[o]);
      mean = mean * mf;
      float v = mean;
      if (step != 1) {
        const float ts = scale * sigma;          // params['scale'] * trans_sigma
        v = (mean + ts * (nz ? nz[o] : gen[c])) * mf;
      }
      xs[j * 3 + c] = v;
      trans[o] = v;
    }
  }
  atomicAdd(&s_len, loc);
  __syncthreads();
  frenet_block(xs, chs, s_len, N, raw, rots + (size_t)b * N * 9);
}

// ---------------------------------------------------------------------------------
// BackboneUpdate + T.compose (modules/backbone_update.py:40-66, utils/affine_utils.py:109-116)
// one warp per residue: 6 dot products of length c_s, quaternion -> rotation, compose.
// ---------------------------------------------------------------------------------
__global__ void backbone_update_kernel(const float* __restrict__ s, const float* __restrict__ W /*[6,c_s]*/,
                                       const float* __restrict__ bias, float* __restrict__ rots,
                                       float* __restrict__ trans, int rows, int c_s) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  float acc[6] = {0, 0, 0, 0, 0, 0};
  const float* sr = s + (size_t)warp * c_s;
  for (int k = lane; k < c_s; k += 32) {
    const float v = sr[k];
#pragma unroll
    for (int o = 0; o < 6; ++o) acc[o] = fmaf(v, W[o * c_s + k], acc[o]);
  }
#pragma unroll
  for (int o = 0; o < 6; ++o) acc[o] = warp_sum(acc[o]) + bias[o];
  if (lane == 0) {
    const float denom = sqrtf(acc[0] * acc[0] + acc[1] * acc[1] + acc[2] * acc[2] + 1.f);
    const float a = 1.f / denom, b = acc[0] / denom, c = acc[1] / denom, d = acc[2] / denom;
    float ru[9], r1[9], rn[9];
    quat_to_rot(a, b, c, d, ru);
    float* R = rots + (size_t)warp * 9;
    float* T = trans + (size_t)warp * 3;
#pragma unroll
    for (int e = 0; e < 9; ++e) r1[e] = R[e];
    mat3_mul(r1, ru, rn);
    const float tx = acc[3], ty = acc[4], tz = acc[5];
    const float nt0 = r1[0] * tx + r1[1] * ty + r1[2] * tz + T[0];
    const float nt1 = r1[3] * tx + r1[4] * ty + r1[5] * tz + T[1];
    const float nt2 = r1[6] * tx + r1[7] * ty + r1[8] * tz + T[2];
#pragma unroll
    for (int e = 0; e < 9; ++e) R[e] = rn[e];
    T[0] = nt0; T[1] = nt1; T[2] = nt2;
  }
}
