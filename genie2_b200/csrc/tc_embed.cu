// K3 on tensor cores (GB_PREC_BF16): PairFeatureNet.forward (pair_feature_net.py:117-160) with the 41-term
// (distance bins + quaternion) and the 37-term (motif bins) dot products per pixel and channel as one tcgen05 GEMM.
//
// The CUDA-core kernel (kernels_embed.cuh, kept for the fp32 parity mode) is bound by shared memory -> register
// bandwidth: every LDS.128 of the broadcast feature tile costs 4 cycles and feeds only 4 (8 packed) FMAs.  Here one
// CTA owns one (b, i) row; per tile of 128 keys j
//   threads 0..127    distance soft bins of key j            -> bf16 row j of the A image, k = 8..44
//   threads 128..255  quaternion of R_j R_i, side features   -> k = 0..3 of the same row; motif bins -> second k-block
//   thread 0          D[j][c] = A[j][:] . Wf[c][:]  (M = 128 keys, N = 128 channels, K = 64 or 128)
//   all warps         TMEM -> swizzled fp32 staging -> per key row, lanes over channels:
//                     p = (Pi + Pj + W_rel[d_ij] + same W_rel[66] + D + fst (w_fst [+ wm_fst])) * m_ij   (coalesced rows)
#include <vector>

#include "common.cuh"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

#define PEK_NBIN 37
#define PEK_THREADS 256
// A 2 x 16K | W 2 x 16K | staging 64K | side 4 x 512 | barriers
#define PEK_OFF_W 32768
#define PEK_OFF_S 65536
#define PEK_OFF_SIDE (65536 + 65536)
#define PEK_OFF_BAR (PEK_OFF_SIDE + 2048)
#define PEK_SMEM (PEK_OFF_BAR + 64 + 1024)
// without motif conditioning the second k-blocks are not needed: A 16K | W 16K | staging | side | barriers
#define PEK_SMEM_NOMOTIF (PEK_SMEM - 32768)

// softmax_k(-4 |d - (dmin + k dstep)|) * scale for k < 37 as five 16-byte chunks of bf16 (k = 0..39, the last three zero)
__device__ __forceinline__ void soft_bins_bf16(float d, float dmin, float dstep, float scale, uint4 (&ch)[5]) {
  float e[40];
  float mx = -1e30f;
#pragma unroll
  for (int k = 0; k < PEK_NBIN; ++k) {
    e[k] = -4.f * fabsf(d - (dmin + (float)k * dstep));
    mx = fmaxf(mx, e[k]);
  }
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < PEK_NBIN; ++k) {
    e[k] = __expf(e[k] - mx);
    sum += e[k];
  }
  const float inv = scale / sum;
#pragma unroll
  for (int k = PEK_NBIN; k < 40; ++k) e[k] = 0.f;
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    ch[q].x = pack_bf16(e[8 * q + 0] * inv, e[8 * q + 1] * inv);
    ch[q].y = pack_bf16(e[8 * q + 2] * inv, e[8 * q + 3] * inv);
    ch[q].z = pack_bf16(e[8 * q + 4] * inv, e[8 * q + 5] * inv);
    ch[q].w = pack_bf16(e[8 * q + 6] * inv, e[8 * q + 7] * inv);
  }
}

__device__ __forceinline__ void put_chunk(uint8_t* img, uint32_t row, uint32_t q, uint4 v) {
  *reinterpret_cast<uint4*>(img + row * 128u + ((q ^ (row & 7u)) << 4)) = v;
}

template <bool MOTIF>
__global__ void __launch_bounds__(PEK_THREADS, MOTIF ? 1 : 2) pair_embed_tc_kernel(
    float* __restrict__ p, const float* __restrict__ PiPj /*[B*N, 256]: Pi | Pj*/, const float* __restrict__ rots,
    const float* __restrict__ trans, const int* __restrict__ residue_index, const int* __restrict__ chain_index,
    const int* __restrict__ mask, const uint8_t* __restrict__ fsm, const uint8_t* __restrict__ fst,
    const float* __restrict__ motif_pos, const float* __restrict__ WrelT /*[2k+3][128]*/, const bf16* __restrict__ Wimg,
    const float* __restrict__ wfst /*[2][128]: template, motif*/, const float* __restrict__ quat_follow, int N, int relpos_k,
    float dmin, float dstep, int row0) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  constexpr uint32_t KB = MOTIF ? 2 : 1, SHR = MOTIF ? 0 : 32768;      // the no-motif layout drops the second k-blocks
  uint8_t* A = sm;
  uint8_t* W = sm + (MOTIF ? PEK_OFF_W : 16384);
  uint8_t* S = sm + PEK_OFF_S - SHR;
  float* side = reinterpret_cast<float*>(sm + PEK_OFF_SIDE - SHR);      // [4][128]: relidx (int) | same | fst | pm
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(sm + PEK_OFF_BAR - SHR);
  uint64_t* bar_mma = bar_w + 1;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_w + 2);
  __shared__ float Ri[9], xi[3], mi3[3];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, i = blockIdx.x + row0;
  const size_t row_i = (size_t)b * N + i;
  if (tid == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
    mbar_expect_tx(bar_w, KB * 16384u);
    bulk_g2s(W, Wimg, KB * 16384u, bar_w);
  }
  if (warp == 1) tmem_alloc(tslot, 128);
  if (tid < 9) Ri[tid] = rots[row_i * 9 + tid];
  if (tid < 3) { xi[tid] = trans[row_i * 3 + tid]; mi3[tid] = motif_pos[row_i * 3 + tid]; }
  const int ri = residue_index[row_i], ci = chain_index[row_i], mi = mask[row_i], fsi = fsm[row_i];
  // per-lane constants of the combine phase: channels 4 lane .. 4 lane + 3
  const float4 pi4 = ld4(PiPj + row_i * 256 + lane * 4);
  const float4 wsame4 = ld4(WrelT + (size_t)(2 * relpos_k + 2) * 128 + lane * 4);
  float4 wfst4 = ld4(wfst + lane * 4);
  if (MOTIF) { const float4 m4 = ld4(wfst + 128 + lane * 4); wfst4.x += m4.x; wfst4.y += m4.y; wfst4.z += m4.z; wfst4.w += m4.w; }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);

  uint32_t tile = 0;
  for (int j0 = 0; j0 < N; j0 += 128, ++tile) {
    // ---- features of key j = j0 + (tid & 127)
    {
      const uint32_t r = tid & 127;
      const int j = j0 + (int)r;
      const bool live = j < N;
      const size_t row_j = (size_t)b * N + (live ? j : 0);
      const float pm = live ? (float)(mi * mask[row_j]) : 0.f;
      if (tid < 128) {
        uint4 ch[5];
        if (live) {
          const float dx = xi[0] - trans[row_j * 3 + 0], dy = xi[1] - trans[row_j * 3 + 1], dz = xi[2] - trans[row_j * 3 + 2];
          soft_bins_bf16(sqrtf(1e-10f + dx * dx + dy * dy + dz * dz), dmin, dstep, pm, ch);
        } else {
#pragma unroll
          for (int q = 0; q < 5; ++q) ch[q] = zero4;
        }
#pragma unroll
        for (int q = 0; q < 5; ++q) put_chunk(A, r, 1 + q, ch[q]);
      } else {
        uint4 c0 = zero4;
        int fs = 0, relidx = 0, same = 0;
        if (live) {
          float Rj[9], R[9], q[4];
#pragma unroll
          for (int e = 0; e < 9; ++e) Rj[e] = rots[row_j * 9 + e];
          mat3_mul(Rj, Ri, R);                       // r_ij = R_j . R_i  (pair_feature_net.py:288-291)
          rot_to_quat_canonical(R, q);
          if (quat_follow) {
            const float* qf = quat_follow + (row_i * N + j) * 4;
            if (q[0] * qf[0] + q[1] * qf[1] + q[2] * qf[2] + q[3] * qf[3] < 0.f) { q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2]; q[3] = -q[3]; }
          }
          c0.x = pack_bf16(q[0] * pm, q[1] * pm);
          c0.y = pack_bf16(q[2] * pm, q[3] * pm);
          fs = fst[row_i * N + j];
          same = chain_index[row_j] == ci;
          int dr = ri - residue_index[row_j] + relpos_k;
          dr = dr < 0 ? 0 : (dr > 2 * relpos_k ? 2 * relpos_k : dr);
          relidx = same ? dr : 2 * relpos_k + 1;
        }
        put_chunk(A, r, 0, c0);
        put_chunk(A, r, 6, zero4);
        put_chunk(A, r, 7, zero4);
        side[r] = __int_as_float(relidx);
        side[128 + r] = (float)same;
        side[256 + r] = (float)fs;
        side[384 + r] = pm;
        if (MOTIF) {
          uint4 ch[5];
          if (live && fs) {
            const float ex = mi3[0] - motif_pos[row_j * 3 + 0], ey = mi3[1] - motif_pos[row_j * 3 + 1], ez = mi3[2] - motif_pos[row_j * 3 + 2];
            soft_bins_bf16(sqrtf(1e-10f + ex * ex + ey * ey + ez * ez), dmin, dstep, (float)(fsi * (int)fsm[row_j]), ch);
          } else {
#pragma unroll
            for (int q = 0; q < 5; ++q) ch[q] = zero4;
          }
#pragma unroll
          for (int q = 0; q < 5; ++q) put_chunk(A + 16384, r, q, ch[q]);
#pragma unroll
          for (int q = 5; q < 8; ++q) put_chunk(A + 16384, r, q, zero4);
        }
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (tid == 0) {
      mbar_wait(bar_w, 0);
      tc_fence_after();
      umma_tile(tmem, smem_u32(A), smem_u32(W), 128, KB, false);
      umma_commit(bar_mma);
    }
    __syncwarp();
    mbar_wait(bar_mma, tile & 1);
    tc_fence_after();
    // ---- TMEM -> staging: thread = key row (lane quarter warp & 3), warps 0-3 channels 0..63, warps 4-7 channels 64..127
    {
      const uint32_t q = warp & 3, m = q * 32 + lane, hcol = (warp >> 2) * 64;
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const uint32_t col0 = hcol + cc * 32;
        uint32_t v[32];
        tmem_ld32(tmem + ((q * 32u) << 16) + col0, v);
        tmem_ld_wait();
#pragma unroll
        for (int k4 = 0; k4 < 8; ++k4)
          *reinterpret_cast<float4*>(S + m * 512u + ((((col0 >> 2) + k4) ^ (m & 31u)) << 4)) =
              make_float4(__uint_as_float(v[4 * k4]), __uint_as_float(v[4 * k4 + 1]), __uint_as_float(v[4 * k4 + 2]), __uint_as_float(v[4 * k4 + 3]));
      }
    }
    tc_fence_before();
    __syncthreads();
    // ---- combine: a warp owns 16 consecutive keys, a lane 4 channels; 4 keys in flight
    {
      const int jn = min(128, N - j0);
      for (int r0 = warp * 16; r0 < warp * 16 + 16 && r0 < jn; r0 += 4) {
        float4 pj[4], wr[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = min(r0 + u, jn - 1);
          pj[u] = ld4(PiPj + ((size_t)b * N + j0 + r) * 256 + 128 + lane * 4);
          wr[u] = ld4(WrelT + (size_t)__float_as_int(side[r]) * 128 + lane * 4);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = r0 + u;
          if (r >= jn) break;
          const float4 t = *reinterpret_cast<const float4*>(S + (uint32_t)r * 512u + ((((uint32_t)lane) ^ ((uint32_t)r & 31u)) << 4));
          const float same = side[128 + r], fs = side[256 + r], pm = side[384 + r];
          float4 o;
          o.x = (pi4.x + pj[u].x + wr[u].x + same * wsame4.x + t.x + fs * wfst4.x) * pm;
          o.y = (pi4.y + pj[u].y + wr[u].y + same * wsame4.y + t.y + fs * wfst4.y) * pm;
          o.z = (pi4.z + pj[u].z + wr[u].z + same * wsame4.z + t.z + fs * wfst4.z) * pm;
          o.w = (pi4.w + pj[u].w + wr[u].w + same * wsame4.w + t.w + fs * wfst4.w) * pm;
          st4(p + (row_i * N + j0 + r) * 128 + lane * 4, o);
        }
      }
    }
    __syncthreads();      // side arrays, staging and A are rewritten by the next tile
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// Wf image [2 kb][128 channels][64 k]: block 0: k = 0..3 quaternion (W_t[:, 37:41]), k = 8..44 distance bins (W_t[:, 0:37]);
// block 1: k = 0..36 motif bins (W_m[:, 0:37]).  wfst = W_t[:,41] + W_t[:,42] | W_m[:,37] + W_m[:,38].
int tc_pack_pair_embed(gb_ctx* c, const std::vector<float>& wt, const std::vector<float>& wm, int nbin) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  s->Wfeat_img = nullptr;
  if (nbin != PEK_NBIN) return GB_OK;      // other bin counts keep the CUDA-core kernel
  const int ct = nbin + 6, cm = nbin + 2;
  std::vector<uint16_t> img(2 * 128 * 64);
  pack_image(img.data(), 128, 128, [&](int r, int k) -> float {
    if (k < 4) return wt[(size_t)r * ct + nbin + k];
    if (k >= 8 && k < 8 + nbin) return wt[(size_t)r * ct + (k - 8)];
    if (k >= 64 && k < 64 + nbin) return wm[(size_t)r * cm + (k - 64)];
    return 0.f;
  });
  std::vector<float> wf(256);
  for (int r = 0; r < 128; ++r) {
    wf[r] = wt[(size_t)r * ct + nbin + 4] + wt[(size_t)r * ct + nbin + 5];
    wf[128 + r] = wm[(size_t)r * cm + nbin] + wm[(size_t)r * cm + nbin + 1];
  }
  s->Wfeat_img = (bf16*)upload_bytes(c, img.data(), img.size() * 2);
  s->wfst = (float*)upload_bytes(c, wf.data(), wf.size() * 4);
  if (!s->Wfeat_img || !s->wfst) return gb_fail(c, GB_ERR_CUDA, "tc_pack_pair_embed: device allocation failed");
  cudaError_t e;
  if ((e = cudaFuncSetAttribute(pair_embed_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PEK_SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(pair_embed_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PEK_SMEM_NOMOTIF)) != cudaSuccess)
    return gb_fail(c, GB_ERR_CUDA, "tc_pack_pair_embed: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  return GB_OK;
}

bool tc_pair_embed_supported(const gb_ctx* c) {
  const TcState* s = static_cast<const TcState*>(c->tc_state);
  return s && s->Wfeat_img && c->d.template_dist_n_bin == PEK_NBIN && c->d.c_p == 128;
}

int tc_pair_embed(gb_ctx* c, const float* quat_follow, int rows, int row0, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const gb_dims& d = c->d;
  const dim3 grid(rows, c->B);
  if (c->has_motif)
    pair_embed_tc_kernel<true><<<grid, PEK_THREADS, PEK_SMEM, st>>>(c->p, c->pipj, c->rots_w, c->trans_w, c->residue_index, c->chain_index,
        c->mask, c->fsm, c->fst, c->motif_pos, c->WrelT, s->Wfeat_img, s->wfst, quat_follow, c->N, d.relpos_k, d.template_dist_min,
        d.template_dist_step, row0);
  else
    pair_embed_tc_kernel<false><<<grid, PEK_THREADS, PEK_SMEM_NOMOTIF, st>>>(c->p, c->pipj, c->rots_w, c->trans_w, c->residue_index,
        c->chain_index, c->mask, c->fsm, c->fst, c->motif_pos, c->WrelT, s->Wfeat_img, s->wfst, quat_follow, c->N, d.relpos_k,
        d.template_dist_min, d.template_dist_step, row0);
  GB_LAUNCH_CHECK(c, "pair_embed_tc");
  return GB_OK;
}
