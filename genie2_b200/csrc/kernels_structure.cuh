// Structure network kernels: IPA (K9) and row LayerNorm; the GEMMs use gemm_simt.cuh.
// Specialised for the base IPA geometry H = 12 heads, C = 16, Pq = 4, Pv = 8, c_p = 128.
#pragma once
#include "common.cuh"

#include "ipa_dims.h"

// LayerNorm over the last dim C of [rows, C]; one warp per row; optional second output copy.
__global__ void layernorm_rows_kernel(const float* __restrict__ in, float* __restrict__ out, float* __restrict__ out2,
                                      const float* __restrict__ gamma, const float* __restrict__ beta, int rows, int C) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* r = in + (size_t)warp * C;
  float sum = 0.f;
  for (int k = lane; k < C; k += 32) sum += r[k];
  const float mean = warp_sum(sum) / (float)C;
  float d2 = 0.f;
  for (int k = lane; k < C; k += 32) { const float d = r[k] - mean; d2 = fmaf(d, d, d2); }
  const float rstd = rsqrtf(warp_sum(d2) / (float)C + 1e-5f);
  for (int k = lane; k < C; k += 32) {
    const float v = (r[k] - mean) * rstd * gamma[k] + beta[k];
    out[(size_t)warp * C + k] = v;
    if (out2) out2[(size_t)warp * C + k] = v;
  }
}

// Map the projected points into the global frame in place: T.apply (invariant_point_attention.py:149,162).
// One thread per (row, point); points are stored xyz-major in the projection row.
__global__ void ipa_points_to_global_kernel(float* __restrict__ qkv, const float* __restrict__ rots,
                                            const float* __restrict__ trans, int rows) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int NP = IPA_H * IPA_PQ + IPA_H * (IPA_PQ + IPA_PV);   // 48 + 144
  if (idx >= rows * NP) return;
  const int row = idx / NP, pt = idx % NP;
  float* base;
  int stride;
  if (pt < IPA_H * IPA_PQ) { base = qkv + (size_t)row * IPA_QKV + IPA_OFF_QP + pt; stride = IPA_H * IPA_PQ; }
  else { base = qkv + (size_t)row * IPA_QKV + IPA_OFF_KVP + (pt - IPA_H * IPA_PQ); stride = IPA_H * (IPA_PQ + IPA_PV); }
  const float x = base[0], y = base[stride], z = base[2 * stride];
  const float* R = rots + (size_t)row * 9;
  const float* T = trans + (size_t)row * 3;
  base[0] = R[0] * x + R[1] * y + R[2] * z + T[0];
  base[stride] = R[3] * x + R[4] * y + R[5] * z + T[1];
  base[2 * stride] = R[6] * x + R[7] * y + R[8] * z + T[2];
}

// ---------------------------------------------------------------------------------
// K9b ipa_attention (invariant_point_attention.py:181-259).  One CTA per query row (b,i),
// 256 threads.  Pass 1: logits[h][j] for every key j (pair bias from one coalesced read of
// z[b,i,j,:], scalar q.k, point distances, mask); softmax over j in shared memory; pass 2:
// o, o_pt, o_pair as probability-weighted sums (second read of the z row comes from L2).
//   WbT [128][12] pair-bias weight (c-major), bb [12], hw [12] = softplus(head_weights) * sqrt(1/54)
//   cat_out [B*N, 2112]
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ipa_attention_kernel(const float* __restrict__ qkv, const float* __restrict__ z,
                                                            const float* __restrict__ rots,
                                                            const float* __restrict__ trans,
                                                            const int* __restrict__ mask, const float* __restrict__ Wb,
                                                            const float* __restrict__ bb, const float* __restrict__ hw,
                                                            float* __restrict__ cat_out, int N, int row0) {
  extern __shared__ __align__(16) float sm[];
  float* logit = sm;                          // [12][N]
  float* qs = sm + IPA_H * N;                 // q (192) + q_pts (144)
  float* red = qs + 192 + 144;                // o_pt accumulators [288] for the inverse-frame stage
  const int b = blockIdx.y, i = blockIdx.x + row0, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const size_t row_i = (size_t)b * N + i;
  for (int k = tid; k < 192; k += 256) qs[k] = qkv[row_i * IPA_QKV + k];
  for (int k = tid; k < 144; k += 256) qs[192 + k] = qkv[row_i * IPA_QKV + IPA_OFF_QP + k];
  // lane's slice of the pair-bias weight: channels lane*4..+3 for 12 heads
  float wb[IPA_H][4];
#pragma unroll
  for (int h = 0; h < IPA_H; ++h)
#pragma unroll
    for (int e = 0; e < 4; ++e) wb[h][e] = Wb[h * IPA_CP + lane * 4 + e];
  const int mi = mask[row_i];
  __syncthreads();
  const float* zrow = z + row_i * (size_t)N * IPA_CP;
  const float sc_qk = sqrtf(1.f / (3.f * IPA_C)), sc_b = sqrtf(1.f / 3.f);
  for (int j = warp; j < N; j += 8) {
    const float4 zv = ld4(zrow + (size_t)j * IPA_CP + lane * 4);
    float pb[IPA_H];
#pragma unroll
    for (int h = 0; h < IPA_H; ++h) pb[h] = zv.x * wb[h][0] + zv.y * wb[h][1] + zv.z * wb[h][2] + zv.w * wb[h][3];
#pragma unroll
    for (int h = 0; h < IPA_H; ++h) pb[h] = warp_sum(pb[h]);
    if (lane < IPA_H) {
      const int h = lane;
      float bias = 0.f;
#pragma unroll
      for (int hh = 0; hh < IPA_H; ++hh) if (hh == h) bias = pb[hh];
      const size_t row_j = (size_t)b * N + j;
      const float* kj = qkv + row_j * IPA_QKV + IPA_OFF_KV + h * 2 * IPA_C;
      float qk = 0.f;
#pragma unroll
      for (int c = 0; c < IPA_C; ++c) qk = fmaf(qs[h * IPA_C + c], kj[c], qk);
      const float* kp = qkv + row_j * IPA_QKV + IPA_OFF_KVP;
      float d2 = 0.f;
#pragma unroll
      for (int pp = 0; pp < IPA_PQ; ++pp)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float d = qs[192 + c * (IPA_H * IPA_PQ) + h * IPA_PQ + pp] - kp[c * (IPA_H * 12) + h * 12 + pp];
          d2 = fmaf(d, d, d2);
        }
      const float sq = (float)(mi * mask[row_j]);
      logit[h * N + j] = qk * sc_qk + sc_b * (bias + bb[h]) + d2 * hw[h] * (-0.5f) + 1e5f * (sq - 1.f);
    }
  }
  __syncthreads();
  // softmax over j per head
  for (int h = warp; h < IPA_H; h += 8) {
    float mx = -3.0e38f;
    for (int j = lane; j < N; j += 32) mx = fmaxf(mx, logit[h * N + j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < N; j += 32) { const float e = expf(logit[h * N + j] - mx); logit[h * N + j] = e; sum += e; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int j = lane; j < N; j += 32) logit[h * N + j] *= inv;
  }
  __syncthreads();
  float* cat = cat_out + row_i * IPA_CAT;
  // o_pair[h][c] = sum_j a[h][j] z[j][c]: thread (c = tid & 127, head group = tid >> 7 -> 6 heads)
  {
    const int c = tid & 127, h0 = (tid >> 7) * 6;
    float acc[6] = {0, 0, 0, 0, 0, 0};
    for (int j = 0; j < N; ++j) {
      const float zv = zrow[(size_t)j * IPA_CP + c];
#pragma unroll
      for (int e = 0; e < 6; ++e) acc[e] = fmaf(logit[(h0 + e) * N + j], zv, acc[e]);
    }
#pragma unroll
    for (int e = 0; e < 6; ++e) cat[192 + 4 * 96 + (h0 + e) * IPA_CP + c] = acc[e];
  }
  // o[h][c] (192 outputs) and o_pt[h][p][xyz] (288 outputs, global frame)
  for (int o = tid; o < 192 + 288; o += 256) {
    int h, off;
    if (o < 192) { h = o / IPA_C; off = IPA_OFF_KV + h * 2 * IPA_C + IPA_C + (o % IPA_C); }
    else { const int e = o - 192; const int c = e / 96, r = e % 96; h = r / IPA_PV; off = IPA_OFF_KVP + c * 144 + h * 12 + IPA_PQ + (r % IPA_PV); }
    float acc = 0.f;
    const float* src = qkv + (size_t)b * N * IPA_QKV + off;
    for (int j = 0; j < N; ++j) acc = fmaf(logit[h * N + j], src[(size_t)j * IPA_QKV], acc);
    if (o < 192) cat[o] = acc; else red[o - 192] = acc;   // red[c*96 + h*8 + p]
  }
  __syncthreads();
  if (tid < 96) {          // inverse frame + norm (invariant_point_attention.py:227-240)
    const float* R = rots + row_i * 9;
    const float* T = trans + row_i * 3;
    const float x = red[tid] - T[0], y = red[96 + tid] - T[1], zc = red[192 + tid] - T[2];
    const float lx = R[0] * x + R[3] * y + R[6] * zc;      // R^T (v - t)
    const float ly = R[1] * x + R[4] * y + R[7] * zc;
    const float lz = R[2] * x + R[5] * y + R[8] * zc;
    cat[192 + tid] = lx;
    cat[192 + 96 + tid] = ly;
    cat[192 + 192 + tid] = lz;
    cat[192 + 288 + tid] = sqrtf(lx * lx + ly * ly + lz * lz + 1e-8f);
  }
}
