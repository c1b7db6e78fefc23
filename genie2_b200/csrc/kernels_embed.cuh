// K2 single_embed and K3 pair_embed (SingleFeatureNet / PairFeatureNet).
#pragma once
#include "common.cuh"

// ---------------------------------------------------------------------------------
// sinusoidal_encoding (utils/encoding.py:5-25) with the reference's fp32 op order:
//   idx d (0-based): k = d + 1; even d -> cos(v*pi / N^(2(k-1)/D)); odd d -> sin(v*pi / N^(2k/D))
// enc[r, d] for the integer values v[r].
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float sinus_enc(int v, float Nf, int D, int d) {
  const int k = d + 1;
  const float pi = 3.14159265358979323846f;
  if ((d & 1) == 0) {
    const float e = (float)(2 * (k - 1)) / (float)D;
    const float div = powf(Nf, e);
    return cosf((float)v * pi / div);
  } else {
    const float e = (float)(2 * k) / (float)D;
    const float div = powf(Nf, e);
    return sinf((float)v * pi / div);
  }
}

// time features for every timestep t = 0..T: X[t, 0..D)
__global__ void time_encoding_kernel(float* __restrict__ X, int T1, int n_timestep, int D) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T1 * D) return;
  const int t = i / D, d = i % D;
  X[i] = sinus_enc(t, (float)n_timestep, D, d);
}

// Step-invariant single features (single_feature_net.py:103-142 minus the timestep block):
// X[r, :] = [pos(c_pos) | chain(c_chain) | aatype*fsm (20) | fsm | fsm | interface]   (c_pos+c_chain+23)
__global__ void single_static_features_kernel(float* __restrict__ X, const int* __restrict__ residue_index,
                                              const int* __restrict__ chain_index, const int* __restrict__ aatype,
                                              const uint8_t* __restrict__ fsm, const uint8_t* __restrict__ interface,
                                              int rows, int c_pos, int c_chain, int max_n_res, int max_n_chain) {
  const int F = c_pos + c_chain + 23;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * F) return;
  const int r = i / F, d = i % F;
  float v;
  if (d < c_pos) v = sinus_enc(residue_index[r], (float)max_n_res, c_pos, d);
  else if (d < c_pos + c_chain) v = sinus_enc(chain_index[r], (float)max_n_chain, c_chain, d - c_pos);
  else {
    const int e = d - c_pos - c_chain;
    const float f = fsm[r] ? 1.f : 0.f;
    if (e < 20) v = (float)aatype[r * 20 + e] * f;
    else if (e < 22) v = f;
    else v = interface[r] ? 1.f : 0.f;
  }
  X[i] = v;
}

// s[b,n,:] = (s_static[b,n,:] + time_table[t_b,:]) * residue_mask[b,n]
__global__ void single_combine_kernel(float* __restrict__ s, const float* __restrict__ s_static,
                                      const float* __restrict__ time_table, const int* __restrict__ timesteps,
                                      const int* __restrict__ mask, int B, int N, int c_s) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * N * c_s) return;
  const int c = (int)(i % c_s);
  const size_t r = i / c_s;
  const int b = (int)(r / N);
  s[i] = (s_static[i] + time_table[(size_t)timesteps[b] * c_s + c]) * (float)mask[r];
}

// ---------------------------------------------------------------------------------
// K3 pair_embed (pair_feature_net.py:117-160).  One CTA per (b, i) row, 128 threads = c_p
// channels; pixels j processed in tiles of 32.  Per pixel:
//   p = ( Pi[b,i] + Pj[b,j] + W_relpos[:, d_ij] + same_chain * W_relpos[:, 66]
//         + W_t[:, 0:37] . softbins(d_ij) m_ij + W_t[:, 37:41] . quat(R_j R_i) m_ij
//         + (W_t[:,41] + W_t[:,42]) fst_ij
//         + [ W_m[:, 0:37] . softbins(dmotif_ij) fs_i fs_j fst_ij + (W_m[:,37]+W_m[:,38]) fst_ij ] ) * m_ij
// m_ij = residue_mask_i * residue_mask_j, fst = fixed_structure_mask, fs = fixed_sequence_mask.
// Weights arrive transposed [in][128] so the per-channel reads are coalesced.
// ---------------------------------------------------------------------------------
#define PE_TJ 128   // keys per tile = block size: every thread builds the features of one key
#define PE_NBIN_MAX 40

template <int NBIN>
__device__ __forceinline__ void soft_bins(float d, float dmin, float dstep, float* out /*smem, stride PE_TJ*/, float scale) {
  constexpr int nbin = NBIN;
  // softmax_k( -4 |d - (dmin + k dstep)| ), pair_feature_net.py:246-263
  float mx = -1e30f;
  for (int k = 0; k < nbin; ++k) mx = fmaxf(mx, -4.f * fabsf(d - (dmin + (float)k * dstep)));
  float sum = 0.f;
  for (int k = 0; k < nbin; ++k) {
    const float e = expf(-4.f * fabsf(d - (dmin + (float)k * dstep)) - mx);
    out[k * PE_TJ] = e;
    sum += e;
  }
  const float inv = scale / sum;
  for (int k = 0; k < nbin; ++k) out[k * PE_TJ] *= inv;
}

template <int NBIN>
__global__ void __launch_bounds__(128, 4) pair_embed_kernel(
    float* __restrict__ p, const float* __restrict__ PiPj /*[B*N, 256]: Pi | Pj*/, const float* __restrict__ rots,
    const float* __restrict__ trans, const int* __restrict__ residue_index, const int* __restrict__ chain_index,
    const int* __restrict__ mask, const uint8_t* __restrict__ fsm, const uint8_t* __restrict__ fst,
    const float* __restrict__ motif_pos, const float* __restrict__ WrelT /*[2k+3][128]*/,
    const float* __restrict__ WtT /*[nbin+6][128]*/, const float* __restrict__ WmT /*[nbin+2][128]*/,
    const float* __restrict__ quat_follow, int N, int relpos_k, float dmin, float dstep,
    int has_motif, int row0 /* first row of a row-sharded launch (grid.x = rows) */) {
  constexpr int nbin = NBIN;
  // feature-major tiles: the per-channel dot products read four keys per 16-byte broadcast load
  __shared__ __align__(16) float feat[NBIN + 8][PE_TJ];   // bins | quat(4) | fst | relidx | samechain | pm
  __shared__ __align__(16) float mfeat[NBIN][PE_TJ];      // motif bins (already masked)
  __shared__ float Ri[9], xi[3], mi3[3];
  const int b = blockIdx.y, i = blockIdx.x + row0, c = threadIdx.x;
  const size_t row_i = (size_t)b * N + i;
  if (c < 9) Ri[c] = rots[row_i * 9 + c];
  if (c < 3) { xi[c] = trans[row_i * 3 + c]; mi3[c] = motif_pos[row_i * 3 + c]; }
  const int ri = residue_index[row_i], ci = chain_index[row_i], mi = mask[row_i];
  const int fsi = fsm[row_i];
  // per-thread weight columns
  float2 wt2[NBIN + 4];      // each weight twice: the second operand of the packed FMAs
#pragma unroll
  for (int k = 0; k < nbin + 4; ++k) { const float w = WtT[k * 128 + c]; wt2[k] = make_float2(w, w); }
  const float wt_fst = WtT[(nbin + 4) * 128 + c] + WtT[(nbin + 5) * 128 + c];
  const float wm_fst = WmT[nbin * 128 + c] + WmT[(nbin + 1) * 128 + c];
  const float w_same = WrelT[(2 * relpos_k + 2) * 128 + c];
  const float pi_c = PiPj[row_i * 256 + c];
  __syncthreads();

  for (int j0 = 0; j0 < N; j0 += PE_TJ) {
    if (j0 + c < N) {
      const int j = j0 + c;
      const size_t row_j = (size_t)b * N + j;
      const float pm = (float)(mi * mask[row_j]);
      const float dx = xi[0] - trans[row_j * 3 + 0], dy = xi[1] - trans[row_j * 3 + 1], dz = xi[2] - trans[row_j * 3 + 2];
      const float d = sqrtf(1e-10f + dx * dx + dy * dy + dz * dz);
      soft_bins<NBIN>(d, dmin, dstep, &feat[0][c], pm);
      float Rj[9], R[9], q[4];
#pragma unroll
      for (int e = 0; e < 9; ++e) Rj[e] = rots[row_j * 9 + e];
      mat3_mul(Rj, Ri, R);                       // r_ij = R_j . R_i  (pair_feature_net.py:288-291)
      rot_to_quat_canonical(R, q);
      if (quat_follow) {
        const float* qf = quat_follow + (row_i * N + j) * 4;
        if (q[0] * qf[0] + q[1] * qf[1] + q[2] * qf[2] + q[3] * qf[3] < 0.f) { q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2]; q[3] = -q[3]; }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) feat[nbin + e][c] = q[e] * pm;
      const int fs = fst[row_i * N + j];
      feat[nbin + 4][c] = (float)fs;
      const int same = (chain_index[row_j] == ci);
      int dr = ri - residue_index[row_j] + relpos_k;
      dr = dr < 0 ? 0 : (dr > 2 * relpos_k ? 2 * relpos_k : dr);
      feat[nbin + 5][c] = __int_as_float(same ? dr : 2 * relpos_k + 1);
      feat[nbin + 6][c] = (float)same;
      feat[nbin + 7][c] = pm;
      if (has_motif) {
        if (fs) {
          const float ex = mi3[0] - motif_pos[row_j * 3 + 0], ey = mi3[1] - motif_pos[row_j * 3 + 1], ez = mi3[2] - motif_pos[row_j * 3 + 2];
          const float dm = sqrtf(1e-10f + ex * ex + ey * ey + ez * ez);
          soft_bins<NBIN>(dm, dmin, dstep, &mfeat[0][c], (float)(fsi * (int)fsm[row_j]));
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < nbin + 8; ++k) feat[k][c] = 0.f;   // keys past the end: relidx 0, fst 0, never stored
    }
    __syncthreads();
    const int jn = min(PE_TJ, N - j0);
    for (int jj = 0; jj < jn; jj += 8) {
      // global operands of the eight pixels first: their latency hides behind the 41 x 8 FMAs below
      const float4 rela = *reinterpret_cast<const float4*>(&feat[nbin + 5][jj]), relb = *reinterpret_cast<const float4*>(&feat[nbin + 5][jj + 4]);
      const float relv[8] = {rela.x, rela.y, rela.z, rela.w, relb.x, relb.y, relb.z, relb.w};
      float pj[8], wr[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int jc = min(j0 + jj + u, N - 1);
        pj[u] = PiPj[((size_t)b * N + jc) * 256 + 128 + c];
        wr[u] = WrelT[__float_as_int(relv[u]) * 128 + c];
      }
      // packed fp32x2 FMAs (two pixels per instruction, each lane an exact IEEE fma in the same k order as before)
      float2 t[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
      for (int k = 0; k < nbin + 4; ++k) {
        const float4 f = *reinterpret_cast<const float4*>(&feat[k][jj]), g = *reinterpret_cast<const float4*>(&feat[k][jj + 4]);
        t[0] = __ffma2_rn(wt2[k], make_float2(f.x, f.y), t[0]);
        t[1] = __ffma2_rn(wt2[k], make_float2(f.z, f.w), t[1]);
        t[2] = __ffma2_rn(wt2[k], make_float2(g.x, g.y), t[2]);
        t[3] = __ffma2_rn(wt2[k], make_float2(g.z, g.w), t[3]);
      }
      const float tv[8] = {t[0].x, t[0].y, t[1].x, t[1].y, t[2].x, t[2].y, t[3].x, t[3].y};
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (jj + u >= jn) break;
        const float fsv = feat[nbin + 4][jj + u], samev = feat[nbin + 6][jj + u], pmv = feat[nbin + 7][jj + u];
        float acc = pi_c + pj[u];
        acc += wr[u];
        acc = fmaf(samev, w_same, acc);
        acc += tv[u];
        acc = fmaf(fsv, wt_fst, acc);
        if (has_motif && fsv != 0.f) {
          float m = 0.f;
          for (int k = 0; k < nbin; ++k) m = fmaf(WmT[k * 128 + c], mfeat[k][jj + u], m);
          acc += m + wm_fst;
        }
        p[(row_i * N + j0 + jj + u) * 128 + c] = acc * pmv;
      }
    }
    __syncthreads();
  }
}
