// K4 for GB_PREC_BF16X3 (triangular_multiplicative_update.py:97-103,107): the persistent warp-specialised kernel of tc_proj.cu
// with SPLIT operands.  z = LayerNorm_in(p) is written as two operand images X_hi = bf16(z), X_lo = bf16(z - X_hi); the
// weights are carried as W_hi | W_lo; every product is the three passes  W_hi X_hi + W_hi X_lo + W_lo X_hi  into the same fp32
// tensor-memory accumulator (~2^-16 per product instead of 2^-9).
//
// Tensor memory cannot hold hi and lo of all four a / b weight matrices (2 x 256 columns) next to the accumulators, so the
// kernel is launched TWICE per update, each launch with ONE gated projection resident in tensor memory:
//   half 0:  a = (W_ap z + b_ap) sigmoid(W_ag z + b_ag) mask   and   g = sigmoid(W_g z + b_g)   (units a0 a1 g per tile)
//   half 1:  b = likewise                                                                        (units b0 b1 per tile)
// Tensor memory: [0,64) W_p hi | [64,128) W_g hi | [128,192) W_p lo | [192,256) W_g lo | two 128-column accumulator slots.
// Shared memory: W_g hi 32K | W_g lo 32K | two X buffers of (hi 32K | lo 32K).
// Outputs: a / b as hi | lo tile-image pairs (the lo plane `lo_ab` elements behind the hi plane, same ImgLayout), g fp32
// [B*N*N][128].  The gate uses ex2 / rcp (1e-6) instead of the 2^-11 tanh.approx of the bf16 kernel; -log2(e) is folded into
// the gate weights and biases when they are packed, so sigmoid(y) = rcp(1 + ex2(y')) is two MUFUs and one add.
#include <vector>

#include "common.cuh"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

// 1 / (1 + 2^y): the logistic function of a pre-activation that was scaled by -log2(e) at pack time
__device__ __forceinline__ float sigmoid_ex2(float y) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(y));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
  return r;
}

#define K4S_THREADS 544
#define K4S_OFF_X 65536u
#define K4S_OFF_BAR 196608u
#define K4S_SMEM (196608 + 128 + 1408 + 1024)

__global__ void __launch_bounds__(K4S_THREADS, 1) trimul_proj_x3_kernel(
    const float* __restrict__ p, int N, int Np, int ntiles, const int* __restrict__ rmask, const bf16* __restrict__ Wimg_hi,
    const bf16* __restrict__ Wimg_lo, const float* __restrict__ bias, bf16* __restrict__ out, size_t lo_ab, float* __restrict__ g_out,
    int incoming, int half, const TileRange tr, const ImgLayout Lo) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* W = sm;                       // g weight images: hi 32K | lo 32K
  uint8_t* X = sm + K4S_OFF_X;           // 2 x (hi 32K | lo 32K)
  uint64_t* full_x = reinterpret_cast<uint64_t*>(sm + K4S_OFF_BAR);
  uint64_t* empty_x = full_x + 2;
  uint64_t* slot_full = full_x + 4;
  uint64_t* slot_free = full_x + 8;
  uint64_t* bar_w = full_x + 12;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(full_x + 14);
  float* cmask = reinterpret_cast<float*>(sm + K4S_OFF_BAR + 128);    // [3][64] residue mask of the tile's columns
  float* rowm = cmask + 192;                                          // [3][2]  ... of its two rows
  float* cfull = cmask + 198;                                         // [3][2]  1 if all 32 columns of a half are unmasked
  float* gbias = cmask + 204;                                         // [128]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t NU = half == 0 ? 3u : 2u;                            // units per tile

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
    // this half's proj | gate rows, hi (warps 0-3) and lo (warps 4-7) -> TMEM columns [0,256): lane = output channel,
    // 64 columns = K 128 bf16 in pairs
    const uint32_t q = warp & 3;
    const bf16* Wsrc = (warp >> 2) ? Wimg_lo : Wimg_hi;
#pragma unroll 1
    for (int bi = 0; bi < 2; ++bi) {
      const int blk = (warp >> 2) * 2 + bi;            // TMEM block: p hi, g hi, p lo, g lo
      const uint4* src = reinterpret_cast<const uint4*>(Wsrc + 16384 + ((size_t)(half * 2 + bi) * 128 + q * 32 + lane) * 128);
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
    if (half == 0) {
      if (lane == 0) {
        mbar_expect_tx(bar_w, 65536u);
        bulk_g2s(W, Wimg_hi, 32768u, bar_w);
        bulk_g2s(W + 32768, Wimg_lo, 32768u, bar_w);
      }
      __syncwarp();
      mbar_wait(bar_w, 0);
    }
    const uint32_t wb = smem_u32(W), tm = warp_uniform(tmem);
    uint32_t U = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const uint32_t buf = it & 1, xb = smem_u32(X) + buf * 65536u;
      mbar_wait(full_x + buf, (it >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (uint32_t u = 0; u < NU; ++u, ++U) {
        const uint32_t slot = U & 1u, d = tm + 256u + slot * 128u;
        mbar_wait(slot_free + slot, ((U >> 1) & 1u) ^ 1u);
        tc_fence_after();
        if (elect_one()) {
          if (u < 2) {
            const uint32_t xr = xb + u * 8192u;
#pragma unroll
            for (uint32_t ks = 0; ks < 8; ++ks) {
              const uint32_t ko = (ks >> 2) * 16384u + (ks & 3u) * 32u;
              const uint64_t bh = umma_desc(xr + ko), bl = umma_desc(xr + 32768u + ko);
              umma_ts(d, tm + ks * 8u, bh, umma_idesc(128, 64), ks);
              umma_ts(d, tm + ks * 8u, bl, umma_idesc(128, 64), 1u);
              umma_ts(d, tm + 128u + ks * 8u, bh, umma_idesc(128, 64), 1u);
              umma_ts(d + 64u, tm + 64u + ks * 8u, bh, umma_idesc(128, 64), ks);
              umma_ts(d + 64u, tm + 64u + ks * 8u, bl, umma_idesc(128, 64), 1u);
              umma_ts(d + 64u, tm + 192u + ks * 8u, bh, umma_idesc(128, 64), 1u);
            }
          } else {
            umma_tile(d, xb, wb, 128, 2, false);
            umma_tile(d, xb + 32768u, wb, 128, 2, true);
            umma_tile(d, xb, wb + 32768u, 128, 2, true);
          }
          umma_commit(slot_full + slot);
          if (u == NU - 1) umma_commit(empty_x + buf);
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
      uint8_t* Xb = X + buf * 65536u;
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
          uint2 oh, ol;
          split_bf16x2(v[gi][k].x * rstd, v[gi][k].y * rstd, oh.x, ol.x);
          split_bf16x2(v[gi][k].z * rstd, v[gi][k].w * rstd, oh.y, ol.y);
          const uint32_t cik = (k & 1) * 4 + (l8 >> 1);
          uint8_t* dst = Xb + (k >> 1) * 16384u + m * 128u + ((cik ^ (m & 7u)) << 4) + (l8 & 1u) * 8u;
          *reinterpret_cast<uint2*>(dst) = oh;
          *reinterpret_cast<uint2*>(dst + 32768u) = ol;
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
    const float b_p = bias[half * 256 + ch], b_g = bias[half * 256 + 128 + ch];
    uint32_t U = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const PixTile T = pix_tile(blockIdx.x + it * gridDim.x, tr);
      const float* cm = cmask + (it % 3) * 64 + c2 * 32;
      uint8_t* const dst0 = reinterpret_cast<uint8_t*>(out + img_off(Lo, T.b, ch, T.jb, 2 * T.t2));
      mbar_wait(full_x + (it & 1), (it >> 1) & 1);      // orders the producers' mask writes before our reads
      const bool cols_on = cfull[(it % 3) * 2 + c2] != 0.f;
      const float rm0 = rowm[(it % 3) * 2], rm1 = rowm[(it % 3) * 2 + 1];
#pragma unroll
      for (uint32_t u = 0; u < 2; ++u, ++U) {
        const uint32_t slot = U & 1u, orow = 2 * T.t2 + u;
        uint8_t* dst = dst0 + u * 128;
        const bool row_in = orow < (uint32_t)N, row_on = (u ? rm1 : rm0) != 0.f;
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
          uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u}, wl[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
          if (row_on) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              float u0 = __uint_as_float(vp[2 * e]) + b_p, u1 = __uint_as_float(vp[2 * e + 1]) + b_p;
              if (!cols_on) { u0 *= cm[hf * 16 + 2 * e]; u1 *= cm[hf * 16 + 2 * e + 1]; }
              const float o0 = u0 * sigmoid_ex2(__uint_as_float(vg[2 * e]) + b_g), o1 = u1 * sigmoid_ex2(__uint_as_float(vg[2 * e + 1]) + b_g);
              split_bf16x2(o0, o1, w[e], wl[e]);
            }
          }
          // the two 16-byte chunks of this half share one 32-byte sector; the swizzle may swap them
          const bool swap = orow & 1u;
          uint8_t* d32 = dst + ((((uint32_t)(c2 * 4 + hf * 2)) ^ (orow & 6u)) << 4);
          const uint4 w0 = make_uint4(w[0], w[1], w[2], w[3]), w1 = make_uint4(w[4], w[5], w[6], w[7]);
          st_global_32B(d32, swap ? w1 : w0, swap ? w0 : w1);
          const uint4 l0 = make_uint4(wl[0], wl[1], wl[2], wl[3]), l1 = make_uint4(wl[4], wl[5], wl[6], wl[7]);
          st_global_32B(d32 + lo_ab * 2, swap ? l1 : l0, swap ? l0 : l1);
        }
      }
      if (half == 0) {   // unit g: this thread = pixel m, channels [64 c2, 64 c2 + 64), fp32
        const uint32_t slot = U & 1u;
        const uint32_t m = q * 32 + lane;
        const uint32_t row = 2 * T.t2 + (m >> 6), col = 64 * T.jb + (m & 63);
        const bool valid = row < (uint32_t)N && col < (uint32_t)N;
        const size_t px = (size_t)T.b * N * N + (incoming ? (size_t)col * N + row : (size_t)row * N + col);
        float* dst = g_out + px * 128 + c2 * 64;
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
          const float* bg = gbias + c2 * 64 + hf * 32;
#pragma unroll
          for (int k8 = 0; k8 < 4; ++k8) {
            uint32_t o[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = __float_as_uint(sigmoid_ex2(__uint_as_float(v[8 * k8 + e]) + bg[8 * k8 + e]));
            st_global_32B(dst + hf * 32 + k8 * 8, make_uint4(o[0], o[1], o[2], o[3]), make_uint4(o[4], o[5], o[6], o[7]));
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
int tc_proj_x3_init(gb_ctx* c) {
  cudaError_t e = cudaFuncSetAttribute(trimul_proj_x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K4S_SMEM);
  if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "tc_proj_x3_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  return GB_OK;
}

// hi and lo weight images with LayerNorm_in's gamma folded in; the three gates (a_g, b_g, g) also carry -log2(e), so that the
// kernel's sigmoid is rcp(1 + ex2(.)):
//   W' = sc * W * gamma = W'_hi + W'_lo ;  bias' = sc * (b + W beta)         order: a_p | a_g | b_p | b_g | g
int tc_pack_trimul_proj_x3(gb_ctx* c, TriMulW& w, const std::vector<float>* Wm[5], const std::vector<float>* bm[5],
                           const std::vector<float>& ln_g, const std::vector<float>& ln_b) {
  std::vector<uint16_t> hi(5 * 128 * 128), lo(5 * 128 * 128);
  const float NLOG2E = -1.4426950408889634f;
  auto gate = [](int t) { return t == 1 || t == 3 || t == 4; };
  auto wv = [&](int t, int r, int k) { return (gate(t) ? NLOG2E : 1.f) * ((*Wm[t])[(size_t)r * 128 + k] * ln_g[k]); };
  pack_image(hi.data(), 128, 128, [&](int r, int k) { return wv(4, r, k); });
  pack_image(lo.data(), 128, 128, [&](int r, int k) { return bf_lo(wv(4, r, k)); });
  for (int t = 0; t < 4; ++t)
    for (int r = 0; r < 128; ++r)
      for (int k = 0; k < 128; ++k) {
        hi[16384 + ((size_t)t * 128 + r) * 128 + k] = f2bf(wv(t, r, k));
        lo[16384 + ((size_t)t * 128 + r) * 128 + k] = f2bf(bf_lo(wv(t, r, k)));
      }
  w.x3_Wproj_hi = upload_bytes(c, hi.data(), hi.size() * 2);
  w.x3_Wproj_lo = upload_bytes(c, lo.data(), lo.size() * 2);
  std::vector<float> bias(5 * 128);
  for (int t = 0; t < 5; ++t)
    for (int r = 0; r < 128; ++r) {
      double acc = (*bm[t])[r];
      for (int k = 0; k < 128; ++k) acc += (double)(*Wm[t])[(size_t)r * 128 + k] * ln_b[k];
      bias[t * 128 + r] = (float)(gate(t) ? (double)NLOG2E * acc : acc);
    }
  w.x3_bproj = (float*)upload_bytes(c, bias.data(), bias.size() * 4);
  if (!w.x3_Wproj_hi || !w.x3_Wproj_lo || !w.x3_bproj) return gb_fail(c, GB_ERR_CUDA, "tc_pack_trimul_proj_x3: device allocation failed");
  return GB_OK;
}

// two launches: half 0 writes the a images and g, half 1 the b images
int tc_launch_trimul_proj_x3(gb_ctx* c, const TriMulW& w, bool outgoing, const float* p, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int B = c->B, N = c->N, Np = s->Np;
  if (c->shard_world > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding is built for the bf16 pair stack only");
  const TileRange tr = full_tile_range(N, Np);
  const ImgLayout L = img_layout(B, Np, 1, 1);
  const int tiles = B * (int)(tr.nt2 * tr.njb);
  const int grid = tiles < s->num_sms ? tiles : s->num_sms;
  for (int half = 0; half < 2; ++half) {
    trimul_proj_x3_kernel<<<grid, K4S_THREADS, K4S_SMEM, st>>>(p, N, Np, tiles, c->mask, (const bf16*)w.x3_Wproj_hi, (const bf16*)w.x3_Wproj_lo,
                                                               w.x3_bproj, half ? s->b : s->a, s->lo_plane, c->g, outgoing ? 0 : 1, half, tr, L);
    GB_LAUNCH_CHECK(c, "trimul_proj_x3");
  }
  return GB_OK;
}
