// fp32 CUDA-core (FFMA) implementation of the pair stack -- the exact-parity path
// (GB_PREC_FP32) and the on-device reference the tcgen05 kernels are validated against.
//   K4 trimul_proj      LN_in -> a, b (gated, masked, channel-major) and the output gate g
//   K5 trimul_contract  per-channel N x N x N contraction (outgoing / incoming)
//   K6 trimul_out       LN_out -> linear_z -> * g -> p +=
//   K7 pair_transition  LN -> 128->512 -> ReLU -> 512->128 -> mask -> p = (p + .) * mask
// All specialised for c_p = c_hidden_mul = 128.  A CTA owns 128 consecutive pixels
// (rows of the [B*N*N, 128] view of p); 256 threads, each an 8x8 register tile split into
// 4+4 rows / 4+4 cols so shared-memory reads are conflict-free 128-bit loads.
#pragma once
#include "common.cuh"

#define PS_LD 132                      // smem leading dim of the [k][m] activation tile
#define PS_A_FLOATS (128 * PS_LD)
#define PS_W_FLOATS (128 * 128)

__device__ __forceinline__ int ps_row(int ty, int i) { return (i < 4) ? ty * 4 + i : 64 + ty * 4 + (i - 4); }
__device__ __forceinline__ int ps_col(int tx, int j) { return (j < 4) ? tx * 4 + j : 64 + tx * 4 + (j - 4); }

// acc[8][8] += As[k][rows] * Ws[k][cols] over k = 0..127
__device__ __forceinline__ void ps_tile_mma(const float* __restrict__ As, const float* __restrict__ Ws,
                                            float (&acc)[8][8], int tx, int ty) {
#pragma unroll 4
  for (int k = 0; k < 128; ++k) {
    const float4 a0 = ld4(As + k * PS_LD + ty * 4), a1 = ld4(As + k * PS_LD + 64 + ty * 4);
    const float4 w0 = ld4(Ws + k * 128 + tx * 4), w1 = ld4(Ws + k * 128 + 64 + tx * 4);
    const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
  }
}

__device__ __forceinline__ void ps_zero(float (&acc)[8][8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
}

// copy one packed [128 k][128 n] weight tile into smem
__device__ __forceinline__ void ps_load_w(float* Ws, const float* __restrict__ Wg) {
  for (int i = threadIdx.x; i < PS_W_FLOATS / 4; i += 256) st4(Ws + i * 4, ld4(Wg + i * 4));
}

// LayerNorm(128) of 128 pixel-major rows -> As[k][m] (normalised, affine applied).
// One warp per row; lane holds channels lane + 32 q.  Rows >= M are zero-filled.
__device__ __forceinline__ void ps_ln_rows_to_smem(float* As, const float* __restrict__ src, size_t row0, size_t M,
                                                   const float* __restrict__ gamma, const float* __restrict__ beta) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float g[4], be[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { g[q] = gamma[lane + 32 * q]; be[q] = beta[lane + 32 * q]; }
  for (int m = warp; m < 128; m += 8) {
    const size_t row = row0 + m;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (row < M) {
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = src[row * 128 + lane + 32 * q];
    }
    const float mean = warp_sum(v[0] + v[1] + v[2] + v[3]) * (1.f / 128.f);
    float d2 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) { const float d = v[q] - mean; d2 = fmaf(d, d, d2); }
    const float rstd = rsqrtf(warp_sum(d2) * (1.f / 128.f) + 1e-5f);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      As[(lane + 32 * q) * PS_LD + m] = (row < M) ? (v[q] - mean) * rstd * g[q] + be[q] : 0.f;
  }
}

__device__ __forceinline__ float ps_pair_mask(const int* __restrict__ rmask, size_t row, int N) {
  const size_t NN = (size_t)N * N;
  const size_t b = row / NN, rem = row % NN;
  const int i = (int)(rem / N), j = (int)(rem % N);
  return (float)(rmask[b * N + i] * rmask[b * N + j]);
}

// ---------------------------------------------------------------------------------
// K4 trimul_proj (triangular_multiplicative_update.py:97-103,107)
//   Wpk : [5][128 k][128 n]  tiles 0,1 -> a (cols 0..63 = a_p rows c0.., cols 64..127 = a_g rows c0..,
//         c0 = 64 * tile), tiles 2,3 -> b likewise, tile 4 -> linear_g (128 cols)
//   bpk : [5][128]
//   a_out, b_out : channel-major [B][128][N*N]     g_out : pixel-major [M][128]
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) trimul_proj_simt_kernel(
    const float* __restrict__ p, size_t M, int N, const int* __restrict__ rmask, const float* __restrict__ ln_g,
    const float* __restrict__ ln_b, const float* __restrict__ Wpk, const float* __restrict__ bpk,
    float* __restrict__ a_out, float* __restrict__ b_out, float* __restrict__ g_out) {
  extern __shared__ __align__(16) float smem[];
  float* As = smem;
  float* Ws = smem + PS_A_FLOATS;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const size_t row0 = (size_t)blockIdx.x * 128;
  const size_t NN = (size_t)N * N;
  ps_ln_rows_to_smem(As, p, row0, M, ln_g, ln_b);
  float mk[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const size_t row = row0 + ps_row(ty, i);
    mk[i] = (row < M) ? ps_pair_mask(rmask, row, N) : 0.f;
  }
  for (int t = 0; t < 5; ++t) {
    __syncthreads();                      // As ready (t=0) / previous staging reads done
    ps_load_w(Ws, Wpk + (size_t)t * PS_W_FLOATS);
    __syncthreads();
    float acc[8][8];
    ps_zero(acc);
    ps_tile_mma(As, Ws, acc, tx, ty);
    const float* bias = bpk + t * 128;
    if (t == 4) {                         // output gate, pixel-major
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const size_t row = row0 + ps_row(ty, i);
        if (row >= M) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int c = ps_col(tx, h * 4);
          float4 o;
          o.x = sigmoidf_(acc[i][h * 4 + 0] + bias[c + 0]);
          o.y = sigmoidf_(acc[i][h * 4 + 1] + bias[c + 1]);
          o.z = sigmoidf_(acc[i][h * 4 + 2] + bias[c + 2]);
          o.w = sigmoidf_(acc[i][h * 4 + 3] + bias[c + 3]);
          st4(g_out + row * 128 + c, o);
        }
      }
      continue;
    }
    __syncthreads();                      // all warps done reading Ws -> reuse as staging [64 ch][PS_LD]
    float* St = Ws;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cl = tx * 4 + j;          // local channel 0..63
      const float bp = bias[cl], bg = bias[64 + cl];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float4 o;
        o.x = (acc[h * 4 + 0][j] + bp) * sigmoidf_(acc[h * 4 + 0][4 + j] + bg) * mk[h * 4 + 0];
        o.y = (acc[h * 4 + 1][j] + bp) * sigmoidf_(acc[h * 4 + 1][4 + j] + bg) * mk[h * 4 + 1];
        o.z = (acc[h * 4 + 2][j] + bp) * sigmoidf_(acc[h * 4 + 2][4 + j] + bg) * mk[h * 4 + 2];
        o.w = (acc[h * 4 + 3][j] + bp) * sigmoidf_(acc[h * 4 + 3][4 + j] + bg) * mk[h * 4 + 3];
        st4(St + cl * PS_LD + h * 64 + ty * 4, o);
      }
    }
    __syncthreads();
    float* dst = (t < 2) ? a_out : b_out;
    const int c0 = (t & 1) * 64;
    const int warp = tid >> 5, lane = tid & 31;
    for (int cl = warp; cl < 64; cl += 8) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int m = lane + 32 * e;
        const size_t row = row0 + m;
        if (row < M) {
          const size_t b = row / NN, rem = row % NN;
          dst[(b * 128 + c0 + cl) * NN + rem] = St[cl * PS_LD + m];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// K5 trimul_contract (triangular_multiplicative_update.py:57-82), channel-major operands.
//   outgoing: x[i,j] = sum_k a[i,k] b[j,k]      incoming: x[i,j] = sum_k a[k,i] b[k,j]
// grid (ceil(N/64), ceil(N/64), B*128); 64x64x16 tiles, 256 threads, 4x4 per thread.
// ---------------------------------------------------------------------------------
template <bool OUTGOING>
__global__ void __launch_bounds__(256) trimul_contract_simt_kernel(const float* __restrict__ a,
                                                                   const float* __restrict__ b,
                                                                   float* __restrict__ x, int N) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const size_t plane = (size_t)blockIdx.z * N * N;
  const float* A = a + plane;
  const float* Bm = b + plane;
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < N; k0 += 16) {
    if (OUTGOING) {
      const int r = tid >> 2, kk = (tid & 3) * 4;       // row r of the tile, 4 consecutive k
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + kk + q;
        As[kk + q][r] = (i0 + r < N && k < N) ? A[(size_t)(i0 + r) * N + k] : 0.f;
        Bs[kk + q][r] = (j0 + r < N && k < N) ? Bm[(size_t)(j0 + r) * N + k] : 0.f;
      }
    } else {
      const int kk = tid >> 4, c4 = (tid & 15) * 4;     // k row kk, 4 consecutive columns
      const int k = k0 + kk;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        As[kk][c4 + q] = (k < N && i0 + c4 + q < N) ? A[(size_t)k * N + i0 + c4 + q] : 0.f;
        Bs[kk][c4 + q] = (k < N && j0 + c4 + q < N) ? Bm[(size_t)k * N + j0 + c4 + q] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float4 av4 = ld4(&As[k][ty * 4]), bv4 = ld4(&Bs[k][tx * 4]);
      const float av[4] = {av4.x, av4.y, av4.z, av4.w}, bv[4] = {bv4.x, bv4.y, bv4.z, bv4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ii = i0 + ty * 4 + i;
    if (ii >= N) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int jj = j0 + tx * 4 + j;
      if (jj < N) x[plane + (size_t)ii * N + jj] = acc[i][j];
    }
  }
}

// ---------------------------------------------------------------------------------
// K6 trimul_out (triangular_multiplicative_update.py:105-108 + residual pair_transform_net.py:111-112)
//   x channel-major [B][128][N*N]; p += (linear_z(LN_out(x)) ) * g
//   WzT [128 k][128 n], bz [128]
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) trimul_out_simt_kernel(
    const float* __restrict__ x, const float* __restrict__ g, float* __restrict__ p, size_t M, int N,
    const float* __restrict__ ln_g, const float* __restrict__ ln_b, const float* __restrict__ WzT,
    const float* __restrict__ bz) {
  extern __shared__ __align__(16) float smem[];
  float* As = smem;
  float* Ws = smem + PS_A_FLOATS;
  __shared__ float s_mean[128], s_rstd[128];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const size_t row0 = (size_t)blockIdx.x * 128;
  const size_t NN = (size_t)N * N;
  // coalesced channel-major load: warp w takes channels w, w+8, ...; lanes along pixels
  const int warp = tid >> 5, lane = tid & 31;
  for (int c = warp; c < 128; c += 8) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int m = lane + 32 * e;
      const size_t row = row0 + m;
      float v = 0.f;
      if (row < M) {
        const size_t b = row / NN, rem = row % NN;
        v = x[(b * 128 + c) * NN + rem];
      }
      As[c * PS_LD + m] = v;
    }
  }
  ps_load_w(Ws, WzT);
  __syncthreads();
  // LN over channels per pixel: 2 threads per pixel, 64 channels each
  {
    const int m = tid >> 1, half = tid & 1;
    float sum = 0.f;
    for (int c = half * 64; c < half * 64 + 64; ++c) sum += As[c * PS_LD + m];
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    const float mean = sum * (1.f / 128.f);
    float d2 = 0.f;
    for (int c = half * 64; c < half * 64 + 64; ++c) { const float d = As[c * PS_LD + m] - mean; d2 = fmaf(d, d, d2); }
    d2 += __shfl_xor_sync(0xffffffffu, d2, 1);
    if (half == 0) { s_mean[m] = mean; s_rstd[m] = rsqrtf(d2 * (1.f / 128.f) + 1e-5f); }
  }
  __syncthreads();
  for (int idx = tid; idx < 128 * 128; idx += 256) {
    const int c = idx >> 7, m = idx & 127;
    As[c * PS_LD + m] = (As[c * PS_LD + m] - s_mean[m]) * s_rstd[m] * ln_g[c] + ln_b[c];
  }
  __syncthreads();
  float acc[8][8];
  ps_zero(acc);
  ps_tile_mma(As, Ws, acc, tx, ty);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const size_t row = row0 + ps_row(ty, i);
    if (row >= M) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int c = ps_col(tx, h * 4);
      const float4 gv = ld4(g + row * 128 + c);
      float4 pv = ld4(p + row * 128 + c);
      pv.x += (acc[i][h * 4 + 0] + bz[c + 0]) * gv.x;
      pv.y += (acc[i][h * 4 + 1] + bz[c + 1]) * gv.y;
      pv.z += (acc[i][h * 4 + 2] + bz[c + 2]) * gv.z;
      pv.w += (acc[i][h * 4 + 3] + bz[c + 3]) * gv.w;
      st4(p + row * 128 + c, pv);
    }
  }
}

// ---------------------------------------------------------------------------------
// K7 pair_transition (pair_transition.py:48-87) + residual + layer-final mask
// (pair_transform_net.py:116-117):  p = (p + (W2 relu(W1 LN(p) + b1) + b2) * mask) * mask
//   W1T [4][128 k][128 n] (hidden tile t = columns 128t..128t+127), b1 [512]
//   W2T [4][128 k'][128 n] (k' = hidden index within tile t), b2 [128]
// The 512-wide hidden activation never leaves shared memory.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) pair_transition_simt_kernel(
    float* __restrict__ p, size_t M, int N, const int* __restrict__ rmask, const float* __restrict__ ln_g,
    const float* __restrict__ ln_b, const float* __restrict__ W1T, const float* __restrict__ b1,
    const float* __restrict__ W2T, const float* __restrict__ b2, int n_hidden_tiles) {
  extern __shared__ __align__(16) float smem[];
  float* As = smem;
  float* Ws = smem + PS_A_FLOATS;
  float* Hs = Ws + PS_W_FLOATS;            // [128 k'][PS_LD]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const size_t row0 = (size_t)blockIdx.x * 128;
  ps_ln_rows_to_smem(As, p, row0, M, ln_g, ln_b);
  float out[8][8];
  ps_zero(out);
  for (int t = 0; t < n_hidden_tiles; ++t) {
    __syncthreads();                       // As ready / previous GEMM2 finished with Ws, Hs
    ps_load_w(Ws, W1T + (size_t)t * PS_W_FLOATS);
    __syncthreads();
    float acc[8][8];
    ps_zero(acc);
    ps_tile_mma(As, Ws, acc, tx, ty);
    // hidden = relu(acc + b1) -> Hs[k' = col][m = row]
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int col = ps_col(tx, j);
      const float bb = b1[t * 128 + col];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float4 o;
        o.x = fmaxf(acc[h * 4 + 0][j] + bb, 0.f);
        o.y = fmaxf(acc[h * 4 + 1][j] + bb, 0.f);
        o.z = fmaxf(acc[h * 4 + 2][j] + bb, 0.f);
        o.w = fmaxf(acc[h * 4 + 3][j] + bb, 0.f);
        st4(Hs + col * PS_LD + h * 64 + ty * 4, o);
      }
    }
    __syncthreads();                       // Hs complete, everyone done with Ws (W1 tile)
    ps_load_w(Ws, W2T + (size_t)t * PS_W_FLOATS);
    __syncthreads();
    ps_tile_mma(Hs, Ws, out, tx, ty);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const size_t row = row0 + ps_row(ty, i);
    if (row >= M) continue;
    const float mk = ps_pair_mask(rmask, row, N);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int c = ps_col(tx, h * 4);
      float4 pv = ld4(p + row * 128 + c);
      pv.x = (pv.x + (out[i][h * 4 + 0] + b2[c + 0]) * mk) * mk;
      pv.y = (pv.y + (out[i][h * 4 + 1] + b2[c + 1]) * mk) * mk;
      pv.z = (pv.z + (out[i][h * 4 + 2] + b2[c + 2]) * mk) * mk;
      pv.w = (pv.w + (out[i][h * 4 + 3] + b2[c + 3]) * mk) * mk;
      st4(p + row * 128 + c, pv);
    }
  }
}
