// K4 trimul_proj_tc (triangular_multiplicative_update.py:97-103,107), persistent and warp-specialised.
//
//   a = (W_ap z + b_ap) * sigmoid(W_ag z + b_ag) * mask      z = LayerNorm_in(p)
//   b = likewise                                              g = sigmoid(W_g z + b_g)
//
// One CTA per SM walks the pixel tiles (2 rows x 64 columns of one sample's pair map) round robin.
//   warps 8..15  producers: 64 KB of p rows in flight per SM in registers, LayerNorm statistics with
//                 8 lanes per row, normalised rows (gamma / beta are folded into the weights on the
//                 host) written as the bf16 K-major operand image X[buf] (two buffers)
//   warp 16      MMA issue (converged warp, elected lane, descriptors in uniform registers).  The a / b
//                 weights are the A operand and live in TENSOR MEMORY for the whole kernel (columns
//                 0..255, loaded once per CTA): TS-mode tcgen05.mma reads only the 64 pixel rows of X
//                 from shared memory per instruction (measured 46 cycles per 128x64x16 MMA against 72 in
//                 SS mode, scripts/micro/mma_rate.cu).  Five "units" per tile go through a ring of two
//                 128-column accumulator slots (columns 256..511):
//                   a0 a1 b0 b1 : D[channel][64 pixels of tile row 0/1], proj in columns 0..63, gate 64..127
//                   g           : D[pixel][channel], SS mode, W_g image in shared memory
//   warps 0..7   epilogue: drain a slot, release it at once, gate, mask, store: a thread owns one channel
//                 and 32 pixels = 64 contiguous bytes of that channel's operand image row, written as two
//                 32-byte (full sector) stores
// Weights are scaled by 1/2 on the host so that sigmoid(y) = (tanh(y/2) + 1)/2 costs one MUFU and one
// packed FMA:  a = u * t + u  with u = (W_ap z + b_ap)/2,  t = tanh((W_ag z + b_ag)/2).
// Output formats: a / b [B][128 ch][Np/64 kb][Np rows][64] tile images (incoming update: transposed
// pixels, see tc_pair.cu), g [B*N*N][128] bf16.
#include <vector>

#include "common.cuh"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

#define K4_THREADS 544
#define K4_OFF_X 32768u
#define K4_OFF_BAR 98304u
#define K4_SMEM (98304 + 128 + 1408 + 1024)   // barriers | masks 800 B + g bias 512 B (+ pad) | alignment slack

__device__ __forceinline__ float tanh_approx(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(x));
  return t;
}

// D[128 ch][64 px] = Wrows[128][128] . Xrows[64][128]^T ; both operands are [2 kb][rows][64] images
__device__ __forceinline__ void k4_mma_T64(uint32_t tmem_d, uint32_t w_rows, uint32_t x_rows) {
  const uint32_t idesc = umma_idesc(128, 64);
#pragma unroll
  for (int kb = 0; kb < 2; ++kb)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      umma(tmem_d, umma_desc(w_rows + kb * (256 * 128) + k * 32), umma_desc(x_rows + kb * 16384 + k * 32), idesc, (kb | k) ? 1u : 0u);
}

__global__ void __launch_bounds__(K4_THREADS, 1) trimul_proj_tc_kernel(
    const float* __restrict__ p, int N, int Np, int ntiles, const int* __restrict__ rmask, const bf16* __restrict__ Wimg,
    const float* __restrict__ bias, bf16* __restrict__ a_out, bf16* __restrict__ b_out, bf16* __restrict__ g_out, int incoming,
    const TileRange tr, const ImgLayout La, const ImgLayout Lb) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* W = sm;                       // g weight image 32K (a / b weights live in tensor memory)
  uint8_t* X = sm + K4_OFF_X;            // 2 x 32K
  uint64_t* full_x = reinterpret_cast<uint64_t*>(sm + K4_OFF_BAR);
  uint64_t* empty_x = full_x + 2;
  uint64_t* slot_full = full_x + 4;
  uint64_t* slot_free = full_x + 8;
  uint64_t* bar_w = full_x + 12;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(full_x + 14);
  float* cmask = reinterpret_cast<float*>(sm + K4_OFF_BAR + 128);     // [3][64] residue mask of the tile's columns
  float* rowm = cmask + 192;                                          // [3][2]  ... of its two rows
  float* cfull = cmask + 198;                                         // [3][2]  1 if all 32 columns of a half are unmasked
  float* gbias = cmask + 204;                                         // [128]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(full_x + i, 8); mbar_init(empty_x + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(slot_full + i, 1); mbar_init(slot_free + i, 8); }
    mbar_init(bar_w, 1);
    fence_mbar_init();
  }
  if (warp == 16) tmem_alloc(tslot, 512);
  if (tid < 128) gbias[tid] = bias[512 + tid];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  if (warp < 8) {
    // a_p | a_g | b_p | b_g rows -> TMEM columns [0,256): lane = output channel, 64 columns = K 128 bf16 in pairs
    const uint32_t q = warp & 3;
#pragma unroll 1
    for (int bi = 0; bi < 2; ++bi) {
      const int blk = (warp >> 2) * 2 + bi;
      const uint4* src = reinterpret_cast<const uint4*>(Wimg + 16384 + ((size_t)blk * 128 + q * 32 + lane) * 128);
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t v[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint4 t = src[hf * 8 + i];
          v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
        tmem_st32(tmem + ((q * 32u) << 16) + blk * 64 + hf * 32, v);
      }
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();            // weights are in tensor memory; from here on p is read and the images are written
  pdl_launch();
  const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 16) {
    // ------------------------------------------------------------------ weights + MMA issue (converged warp)
    if (lane == 0) {
      mbar_expect_tx(bar_w, 32768u);
      bulk_g2s(W, Wimg, 32768u, bar_w);
    }
    __syncwarp();
    mbar_wait(bar_w, 0);
    const uint32_t wb = smem_u32(W), tm = warp_uniform(tmem);
    uint32_t U = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const uint32_t buf = it & 1, xb = smem_u32(X) + buf * 32768u;
      mbar_wait(full_x + buf, (it >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (uint32_t u = 0; u < 5; ++u, ++U) {
        const uint32_t slot = U & 1u, d = tm + 256u + slot * 128u;
        mbar_wait(slot_free + slot, ((U >> 1) & 1u) ^ 1u);
        tc_fence_after();
        if (elect_one()) {
          if (u < 4) {
            const uint32_t wt = tm + (u >> 1) * 128u, xr = xb + (u & 1u) * 8192u;
#pragma unroll
            for (uint32_t ks = 0; ks < 8; ++ks) {
              const uint64_t bd = umma_desc(xr + (ks >> 2) * 16384u + (ks & 3u) * 32u);
              umma_ts(d, wt + ks * 8u, bd, umma_idesc(128, 64), ks);
              umma_ts(d + 64u, wt + 64u + ks * 8u, bd, umma_idesc(128, 64), ks);
            }
          } else {
            umma_tile(d, xb, wb, 128, 2, false);
          }
          umma_commit(slot_full + slot);
          if (u == 4) umma_commit(empty_x + buf);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ producers
    const int pw = warp - 8, ptid = tid - 256;
    const uint32_t rsub = lane >> 3, l8 = lane & 7;
    for (int it = 0; it < my_tiles; ++it) {
      const PixTile T = pix_tile(blockIdx.x + it * gridDim.x, tr);
      const uint32_t buf = it & 1;
      const size_t sample = (size_t)T.b * N * N;
      float4 v[4][4];
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) {
        const uint32_t m = 16 * pw + gi + 4 * rsub;
        const uint32_t row = 2 * T.t2 + (m >> 6), col = 64 * T.jb + (m & 63);
        const bool ok = row < (uint32_t)N && col < (uint32_t)N;
        const float* src = p + (sample + (incoming ? (size_t)col * N + row : (size_t)row * N + col)) * 128 + l8 * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) v[gi][k] = ok ? ld4(src + k * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float mk = 0.f;
      if (ptid < 66) {          // 64 column masks, then the 2 row masks
        const uint32_t r = ptid < 64 ? 64 * T.jb + ptid : 2 * T.t2 + (ptid - 64);
        if (r < (uint32_t)N) mk = (float)rmask[T.b * N + r];
      }
      mbar_wait(empty_x + buf, ((it >> 1) & 1) ^ 1);
      if (ptid < 64) {
        cmask[(it % 3) * 64 + ptid] = mk;
        const unsigned on = __ballot_sync(0xffffffffu, mk != 0.f);
        if (lane == 0) cfull[(it % 3) * 2 + pw] = on == 0xffffffffu ? 1.f : 0.f;
      } else if (ptid < 66) rowm[(it % 3) * 2 + (ptid - 64)] = mk;
      uint8_t* Xb = X + buf * 32768u;
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) {
        const uint32_t m = 16 * pw + gi + 4 * rsub;
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) s += (v[gi][k].x + v[gi][k].y) + (v[gi][k].z + v[gi][k].w);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        const float mean = s * (1.f / 128.f);
        float ss = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          v[gi][k].x -= mean; v[gi][k].y -= mean; v[gi][k].z -= mean; v[gi][k].w -= mean;
          ss = fmaf(v[gi][k].x, v[gi][k].x, ss); ss = fmaf(v[gi][k].y, v[gi][k].y, ss);
          ss = fmaf(v[gi][k].z, v[gi][k].z, ss); ss = fmaf(v[gi][k].w, v[gi][k].w, ss);
        }
        ss += __shfl_xor_sync(0xffffffffu, ss, 1);
        ss += __shfl_xor_sync(0xffffffffu, ss, 2);
        ss += __shfl_xor_sync(0xffffffffu, ss, 4);
        const float rstd = rsqrtf(ss * (1.f / 128.f) + 1e-5f);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint2 o;
          o.x = pack_bf16(v[gi][k].x * rstd, v[gi][k].y * rstd);
          o.y = pack_bf16(v[gi][k].z * rstd, v[gi][k].w * rstd);
          const uint32_t cik = (k & 1) * 4 + (l8 >> 1);
          *reinterpret_cast<uint2*>(Xb + (k >> 1) * 16384u + m * 128u + ((cik ^ (m & 7u)) << 4) + (l8 & 1u) * 8u) = o;
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_x + buf);
    }
  } else {
    // ------------------------------------------------------------------ epilogue
    const uint32_t q = warp & 3, c2 = warp >> 2;
    const uint32_t t_lane = tmem + ((q * 32u) << 16);
    const uint32_t ch = q * 32 + lane;
    const float2 b_ap = make_float2(bias[ch], bias[ch]), b_ag = make_float2(bias[128 + ch], bias[128 + ch]);
    const float2 b_bp = make_float2(bias[256 + ch], bias[256 + ch]), b_bg = make_float2(bias[384 + ch], bias[384 + ch]);
    uint32_t U = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const PixTile T = pix_tile(blockIdx.x + it * gridDim.x, tr);
      const float* cm = cmask + (it % 3) * 64 + c2 * 32;
      uint8_t* const dst_a = reinterpret_cast<uint8_t*>(a_out + img_off(La, T.b, ch, T.jb, 2 * T.t2));
      uint8_t* const dst_b = reinterpret_cast<uint8_t*>(b_out + img_off(Lb, T.b, ch, T.jb, 2 * T.t2));
      mbar_wait(full_x + (it & 1), (it >> 1) & 1);      // orders the producers' mask writes before our reads
      const bool cols_on = cfull[(it % 3) * 2 + c2] != 0.f;
      const float rm0 = rowm[(it % 3) * 2], rm1 = rowm[(it % 3) * 2 + 1];
#pragma unroll
      for (uint32_t u = 0; u < 4; ++u, ++U) {
        const uint32_t slot = U & 1u, hh = u & 1u, orow = 2 * T.t2 + hh;
        const float2 bp = (u < 2) ? b_ap : b_bp, bg = (u < 2) ? b_ag : b_bg;
        uint8_t* dst = ((u < 2) ? dst_a : dst_b) + hh * 128;
        const bool row_in = orow < (uint32_t)N, row_on = (hh ? rm1 : rm0) != 0.f;
        mbar_wait(slot_full + slot, (U >> 1) & 1u);
        tc_fence_after();
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t vp[16], vg[16];
          tmem_ld16(t_lane + 256u + slot * 128u + c2 * 32u + hf * 16u, vp);
          tmem_ld16(t_lane + 256u + slot * 128u + 64u + c2 * 32u + hf * 16u, vg);
          tmem_ld_wait();
          if (hf == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(slot_free + slot);
          }
          if (!row_in) continue;
          uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
          if (row_on) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              float2 uu = __fadd2_rn(make_float2(__uint_as_float(vp[2 * e]), __uint_as_float(vp[2 * e + 1])), bp);
              if (!cols_on) uu = __fmul2_rn(uu, *reinterpret_cast<const float2*>(cm + hf * 16 + 2 * e));
              const float2 tg = __fadd2_rn(make_float2(__uint_as_float(vg[2 * e]), __uint_as_float(vg[2 * e + 1])), bg);
              const float2 o = __ffma2_rn(uu, make_float2(tanh_approx(tg.x), tanh_approx(tg.y)), uu);
              w[e] = pack_bf16(o.x, o.y);
            }
          }
          // the two 16-byte chunks of this half share one 32-byte sector; the swizzle may swap them
          const bool swap = orow & 1u;
          const uint4 w0 = make_uint4(w[0], w[1], w[2], w[3]), w1 = make_uint4(w[4], w[5], w[6], w[7]);
          st_global_32B(dst + ((((uint32_t)(c2 * 4 + hf * 2)) ^ (orow & 6u)) << 4), swap ? w1 : w0, swap ? w0 : w1);
        }
      }
      {   // unit g: this thread = pixel m, channels [64 c2, 64 c2 + 64)
        const uint32_t slot = U & 1u;
        const uint32_t m = q * 32 + lane;
        const uint32_t row = 2 * T.t2 + (m >> 6), col = 64 * T.jb + (m & 63);
        const bool valid = row < (uint32_t)N && col < (uint32_t)N;
        const size_t px = (size_t)T.b * N * N + (incoming ? (size_t)col * N + row : (size_t)row * N + col);
        uint4* dst = reinterpret_cast<uint4*>(g_out + px * 128 + c2 * 64);
        mbar_wait(slot_full + slot, (U >> 1) & 1u);
        tc_fence_after();
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t v[32];
          tmem_ld32(t_lane + 256u + slot * 128u + c2 * 64u + hf * 32u, v);
          tmem_ld_wait();
          if (hf == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(slot_free + slot);
          }
          if (!valid) continue;
          const float4* bg4 = reinterpret_cast<const float4*>(gbias + c2 * 64 + hf * 32);
#pragma unroll
          for (int k8 = 0; k8 < 2; ++k8) {
            uint4 o[2];
#pragma unroll
            for (int k4 = 0; k4 < 2; ++k4) {
              const uint32_t* vv = v + 16 * k8 + 8 * k4;
              const float4 b0 = bg4[4 * k8 + 2 * k4], b1 = bg4[4 * k8 + 2 * k4 + 1];
              o[k4].x = pack_bf16(fmaf(0.5f, tanh_approx(__uint_as_float(vv[0]) + b0.x), 0.5f), fmaf(0.5f, tanh_approx(__uint_as_float(vv[1]) + b0.y), 0.5f));
              o[k4].y = pack_bf16(fmaf(0.5f, tanh_approx(__uint_as_float(vv[2]) + b0.z), 0.5f), fmaf(0.5f, tanh_approx(__uint_as_float(vv[3]) + b0.w), 0.5f));
              o[k4].z = pack_bf16(fmaf(0.5f, tanh_approx(__uint_as_float(vv[4]) + b1.x), 0.5f), fmaf(0.5f, tanh_approx(__uint_as_float(vv[5]) + b1.y), 0.5f));
              o[k4].w = pack_bf16(fmaf(0.5f, tanh_approx(__uint_as_float(vv[6]) + b1.z), 0.5f), fmaf(0.5f, tanh_approx(__uint_as_float(vv[7]) + b1.w), 0.5f));
            }
            st_global_32B(dst + hf * 4 + k8 * 2, o[0], o[1]);
          }
        }
        ++U;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 16) {
    __syncwarp();
    tmem_dealloc(tmem, 512);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int tc_proj_init(gb_ctx* c) {
  cudaError_t e = cudaFuncSetAttribute(trimul_proj_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K4_SMEM);
  if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "tc_proj_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  TcState* s = static_cast<TcState*>(c->tc_state);
  if (cudaDeviceGetAttribute(&s->num_sms, cudaDevAttrMultiProcessorCount, c->device) != cudaSuccess || s->num_sms < 1)
    return gb_fail(c, GB_ERR_CUDA, "tc_proj_init: cannot read the SM count");
  return tc_proj_x3_init(c);
}

// Weight images with LayerNorm_in's gamma folded in and everything scaled by 1/2 (exact in bf16):
//   W' = W * gamma / 2 ;  bias' = (b + W beta) / 2       order: a_p | a_g | b_p | b_g | g
int tc_pack_trimul_proj(gb_ctx* c, TriMulW& w, const std::vector<float>* Wm[5], const std::vector<float>* bm[5],
                        const std::vector<float>& ln_g, const std::vector<float>& ln_b) {
  // [g: K-major operand image 32 KB][a_p | a_g | b_p | b_g: row-major bf16 [128][128], loaded into tensor memory]
  std::vector<uint16_t> img(5 * 128 * 128);
  auto wv = [&](int t, int r, int k) { return 0.5f * (*Wm[t])[(size_t)r * 128 + k] * ln_g[k]; };
  pack_image(img.data(), 128, 128, [&](int r, int k) { return wv(4, r, k); });
  for (int t = 0; t < 4; ++t)
    for (int r = 0; r < 128; ++r)
      for (int k = 0; k < 128; ++k) img[16384 + ((size_t)t * 128 + r) * 128 + k] = f2bf(wv(t, r, k));
  w.tc_Wproj = upload_bytes(c, img.data(), img.size() * 2);
  std::vector<float> bias(5 * 128);
  for (int t = 0; t < 5; ++t)
    for (int r = 0; r < 128; ++r) {
      double acc = (*bm[t])[r];
      for (int k = 0; k < 128; ++k) acc += (double)(*Wm[t])[(size_t)r * 128 + k] * ln_b[k];
      bias[t * 128 + r] = (float)(0.5 * acc);
    }
  w.tc_bproj = (float*)upload_bytes(c, bias.data(), bias.size() * 4);
  if (!w.tc_Wproj || !w.tc_bproj) return gb_fail(c, GB_ERR_CUDA, "tc_pack_trimul_proj: device allocation failed");
  return tc_pack_trimul_proj_x3(c, w, Wm, bm, ln_g, ln_b);
}

int tc_launch_trimul_proj(gb_ctx* c, const TriMulW& w, bool outgoing, const float* p, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int B = c->B, N = c->N, Np = s->Np, W = c->shard_world;
  TileRange tr = full_tile_range(N, Np);
  ImgLayout La = img_layout(B, Np, 1, 1), Lb = La;
  if (W > 1) {
    // this rank's rows of p: image rows [r0, r0 + RB) (outgoing) or k-blocks [r0 / 64, ...) of every image row (incoming)
    const uint32_t RB = (uint32_t)N / W, r0 = c->shard_rank * RB;
    // blocks (channel group d, me) of both images, cut by rows (outgoing) or by k-blocks (incoming): tc_common.cuh ImgLayout
    if (outgoing) { tr = TileRange{r0 >> 1, RB >> 1, 0u, (uint32_t)Np >> 6}; La = Lb = img_layout(B, Np, W, 1, W); }
    else { tr = TileRange{0u, (uint32_t)(N + 1) >> 1, r0 >> 6, RB >> 6}; La = Lb = img_layout(B, Np, 1, W, W); }
  }
  const int tiles = B * (int)(tr.nt2 * tr.njb);
  const int grid = tiles < s->num_sms ? tiles : s->num_sms;
  // plain stream order for this one: launched with the programmatic-serialization attribute it runs 16 us SLOWER per
  // launch (measured, also with the wait at the very top of the kernel); its pdl_wait / pdl_launch are then no-ops
  // for itself but pdl_launch still lets the contraction kernel start early
  trimul_proj_tc_kernel<<<grid, K4_THREADS, K4_SMEM, st>>>(p, N, Np, tiles, c->mask, (const bf16*)w.tc_Wproj, w.tc_bproj, s->a, s->b,
                                                           s->g, outgoing ? 0 : 1, tr, La, Lb);
  GB_LAUNCH_CHECK(c, "trimul_proj_tc");
  return GB_OK;
}
