// K7 pair_transition_tc (modules/pair_transition.py:48-87 + pair_transform_net.py:116-117), persistent and
// warp-specialised:   p = (p + (W2 relu(W1 LN(p) + b1) + b2) * mask) * mask
//
// One CTA per SM walks PAIRS of 128-pixel tiles (A, B), so every 32 KB weight block streamed from L2 is
// used for 256 pixels.  The hidden layer is processed in chunks of 128 channels:
//   G1(t,c)  D1[128 px][128 hid]  = X_t . W1_c^T              SS mode (X, W1_c in shared memory)
//   E1(t,c)  H_t = bf16(relu(D1 + b1_c))                       TMEM -> registers -> TMEM (never leaves the SM)
//   G2(t,c)  D2_t[128 px][128 out] += H_t . W2_c^T             TS mode (A = H_t from tensor memory)
//   E2(t)    p = (p + (D2_t + b2) mask) mask                   thread = pixel, 32-byte global accesses
// Tensor memory (512 columns): D1 128 | H_A 64 | H_B 64 | D2_A 128 | D2_B 128.
// The MMA warp issues  G1(A,c) G2(A,c-1) G1(B,c) G2(B,c-1)  so that the tensor pipe always has work
// while the epilogue warps turn D1 into H.  LayerNorm's gamma / beta are folded into W1 / b1 on the host.
//   warps 0..7   E1 / E2        warps 8..15  LayerNorm producers (registers -> bf16 operand image ring)
//   warp 16      MMA issue      warp 17      weight-block loader (cp.async.bulk ring of three 32 KB slots)
#include <vector>

#include "common.cuh"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

#define K7_THREADS 576
#define K7_OFF_W 98304u
#define K7_OFF_BAR 196608u
#define K7_SMEM (196608 + 256 + 4096 + 512 + 1024 + 16384 + 1024)   // X ring | W ring | barriers | b1 | b2 | masks | E2 staging | slack

// position of a weight block in the per-iteration stream  W1_0, W1_1, W2_0, W1_2, W2_1, ..., W2_{C-1}
__device__ __forceinline__ uint32_t k7_pos_w1(uint32_t c) { return c == 0 ? 0u : 2u * c - 1u; }
__device__ __forceinline__ uint32_t k7_pos_w2(uint32_t c, uint32_t C) { return c == C - 1 ? 2u * C - 1u : 2u * c + 2u; }

__device__ __forceinline__ uint32_t relu_pack_bf16(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

__global__ void __launch_bounds__(K7_THREADS, 1) pair_transition_tc_kernel(
    float* __restrict__ p, uint32_t M, uint32_t px0, int N, uint32_t npairs, const int* __restrict__ rmask, const bf16* __restrict__ W1img,
    const float* __restrict__ b1, const bf16* __restrict__ W2img, const float* __restrict__ b2, uint32_t C) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* XR = sm;                      // 3 x 32 KB operand images of LayerNorm'ed pixel tiles
  uint8_t* WR = sm + K7_OFF_W;           // 3 x 32 KB weight blocks
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + K7_OFF_BAR);
  uint64_t *x_full = bars, *x_free = bars + 3, *w_full = bars + 6, *w_free = bars + 9;
  uint64_t *d1_full = bars + 12, *d1_free = bars + 13, *h_full = bars + 14, *h_free = bars + 16;
  uint64_t *d2_full = bars + 18, *d2_free = bars + 20;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 22);
  float* b1s = reinterpret_cast<float*>(sm + K7_OFF_BAR + 256);
  float* b2s = b1s + 1024;
  float* maskS = b2s + 128;                                   // [2][128] pair masks of the current tile pair
  uint8_t* stg = sm + K7_OFF_BAR + 256 + 4096 + 512 + 1024;   // 16 KB staging slab of E2
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int i = 0; i < 3; ++i) { mbar_init(x_full + i, 8); mbar_init(x_free + i, 1); mbar_init(w_full + i, 1); mbar_init(w_free + i, 1); }
    mbar_init(d1_full, 1); mbar_init(d1_free, 8);
    for (int i = 0; i < 2; ++i) { mbar_init(h_full + i, 8); mbar_init(h_free + i, 1); mbar_init(d2_full + i, 1); mbar_init(d2_free + i, 8); }
    fence_mbar_init();
  }
  if (warp == 16) tmem_alloc(tslot, 512);
  for (uint32_t i = tid; i < C * 128; i += K7_THREADS) b1s[i] = b1[i];
  if (tid < 128) b2s[tid] = b2[tid];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  const uint32_t iters = (npairs - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const uint32_t T_D1 = 0, T_H = 128, T_D2 = 256;
  if (warp != 17) {      // the weight loader touches nothing the previous kernel wrote: it starts streaming at once
    pdl_wait();
    pdl_launch();
  }

  if (warp == 17) {
    // ------------------------------------------------------------------ weight-block loader
    if (lane == 0) {
      uint32_t J = 0;
      for (uint32_t it = 0; it < iters; ++it)
        for (uint32_t pos = 0; pos < 2 * C; ++pos, ++J) {
          // invert the position map: pos 0 -> W1_0; odd pos 2c-1 -> W1_c; even pos 2c+2 -> W2_c; last -> W2_{C-1}
          const bf16* src;
          if (pos == 0) src = W1img;
          else if (pos == 2 * C - 1) src = W2img + (size_t)(C - 1) * 16384;
          else if (pos & 1u) src = W1img + (size_t)((pos + 1) >> 1) * 16384;
          else src = W2img + (size_t)((pos - 2) >> 1) * 16384;
          const uint32_t slot = J % 3u;
          mbar_wait(w_free + slot, ((J / 3u) & 1u) ^ 1u);
          mbar_expect_tx(w_full + slot, 32768u);
          bulk_g2s(WR + slot * 32768u, src, 32768u, w_full + slot);
        }
    }
    __syncwarp();
  } else if (warp == 16) {
    // ------------------------------------------------------------------ MMA issue (converged warp, elected lane)
    const uint32_t tm = warp_uniform(tmem), xr = smem_u32(XR), wr = smem_u32(WR);
    uint32_t n1 = 0;                                  // G1 units issued so far (D1 phase)
    for (uint32_t it = 0; it < iters; ++it) {
      const uint32_t lt0 = 2 * it;                    // local tile counters of A / B
      auto G1 = [&](uint32_t t, uint32_t c) {
        const uint32_t lt = lt0 + t, xs = lt % 3u, J = it * 2 * C + k7_pos_w1(c), ws = J % 3u;
        if (c == 0) { mbar_wait(x_full + xs, (lt / 3u) & 1u); }
        if (t == 0) mbar_wait(w_full + ws, (J / 3u) & 1u);
        mbar_wait(d1_free, (n1 & 1u) ^ 1u);
        tc_fence_after();
        if (elect_one()) {
          umma_tile(tm + T_D1, xr + xs * 32768u, wr + ws * 32768u, 128, 2, false);
          umma_commit(d1_full);
          if (t == 1) umma_commit(w_free + ws);
          if (c == C - 1) umma_commit(x_free + xs);
        }
        __syncwarp();
        ++n1;
      };
      auto G2 = [&](uint32_t t, uint32_t c) {
        const uint32_t J = it * 2 * C + k7_pos_w2(c, C), ws = J % 3u, k = it * C + c;
        if (t == 0) mbar_wait(w_full + ws, (J / 3u) & 1u);
        if (c == 0) mbar_wait(d2_free + t, (it & 1u) ^ 1u);
        mbar_wait(h_full + t, k & 1u);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t wb = wr + ws * 32768u;
#pragma unroll
          for (uint32_t ks = 0; ks < 8; ++ks)
            umma_ts(tm + T_D2 + t * 128u, tm + T_H + t * 64u + ks * 8u, umma_desc(wb + (ks >> 2) * 16384u + (ks & 3u) * 32u),
                    umma_idesc(128, 128), (c | ks) ? 1u : 0u);
          umma_commit(h_free + t);
          if (t == 1) umma_commit(w_free + ws);
          if (c == C - 1) umma_commit(d2_full + t);
        }
        __syncwarp();
      };
      G1(0, 0);
      G1(1, 0);
      for (uint32_t c = 1; c < C; ++c) {
        G1(0, c);
        G2(0, c - 1);
        G1(1, c);
        G2(1, c - 1);
      }
      G2(0, C - 1);
      G2(1, C - 1);
    }
  }
  else if (warp >= 8) {
    // ------------------------------------------------------------------ LayerNorm producers, then E2 of the finished pair
    const uint32_t pw = warp - 8, rsub = lane >> 3, l8 = lane & 7;
    const uint32_t q = pw & 3, c2 = pw >> 2;
    const uint32_t t_lane = tmem + ((q * 32u) << 16);
    auto pair_mask = [&](uint32_t row) -> float {
      if (row >= M) return 0.f;
      row += px0;                                   // p points at pixel px0 of the map (row-sharded launch)
      const uint32_t NN = (uint32_t)N * N, b = row / NN, rem = row - b * NN, i = rem / N, j = rem - i * N;
      return (float)(rmask[b * N + i] * rmask[b * N + j]);
    };
    // LayerNorm of one row group (4 rows per warp instruction, 8 lanes per row) -> bf16 operand image
    auto row0_of = [&](uint32_t lt) -> uint32_t { return (2 * (blockIdx.x + (lt >> 1) * gridDim.x) + (lt & 1u)) * 128u; };
    auto load_group = [&](uint32_t lt, int gi, float4 (&v)[4]) {
      const uint32_t row = row0_of(lt) + 16 * pw + gi + 4 * rsub;
      const float* src = p + (size_t)row * 128 + l8 * 4;
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = row < M ? ld4(src + k * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    auto norm_group = [&](uint32_t lt, int gi, float4 (&v)[4]) {
      uint8_t* Xb = XR + (lt % 3u) * 32768u;
      const uint32_t m = 16 * pw + gi + 4 * rsub;
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      const float mean = s * (1.f / 128.f);
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        v[k].x -= mean; v[k].y -= mean; v[k].z -= mean; v[k].w -= mean;
        ss = fmaf(v[k].x, v[k].x, ss); ss = fmaf(v[k].y, v[k].y, ss);
        ss = fmaf(v[k].z, v[k].z, ss); ss = fmaf(v[k].w, v[k].w, ss);
      }
      ss += __shfl_xor_sync(0xffffffffu, ss, 1);
      ss += __shfl_xor_sync(0xffffffffu, ss, 2);
      ss += __shfl_xor_sync(0xffffffffu, ss, 4);
      const float rstd = rsqrtf(ss * (1.f / 128.f) + 1e-5f);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint2 o;
        o.x = pack_bf16(v[k].x * rstd, v[k].y * rstd);
        o.y = pack_bf16(v[k].z * rstd, v[k].w * rstd);
        const uint32_t cik = (k & 1) * 4 + (l8 >> 1);
        *reinterpret_cast<uint2*>(Xb + (k >> 1) * 16384u + m * 128u + ((cik ^ (m & 7u)) << 4) + (l8 & 1u) * 8u) = o;
      }
    };
    auto produce = [&](uint32_t lt) {
      float4 v[4][4];
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) load_group(lt, gi, v[gi]);
      mbar_wait(x_free + lt % 3u, ((lt / 3u) & 1u) ^ 1u);
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) norm_group(lt, gi, v[gi]);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(x_full + lt % 3u);
    };
    // E2(t): p = (p + (D2_t + b2) mask) mask.  The accumulator goes through a swizzled 16 KB staging slab
    // (128 pixels x 32 channels) so that the read-modify-write of p is done with fully coalesced 128-byte
    // row segments; thread-per-row 32-byte accesses cost one L2 request per lane and were 4x slower.
    auto E2 = [&](uint32_t t, uint32_t it) {
      const uint32_t row0 = (2 * (blockIdx.x + it * gridDim.x) + t) * 128u;
      const uint32_t m1 = q * 32 + lane;                 // phase 1: this thread's pixel row (TMEM lane)
      const uint32_t r4 = lane >> 3, ch = lane & 7;      // phase 2: row within a group of four, 16-byte chunk
      mbar_wait(d2_full + t, it & 1u);
      tc_fence_after();
      auto load_res = [&](uint32_t slab, float4 (&pv)[4]) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t row = row0 + pw * 16 + g * 4 + r4;
          pv[g] = row < M ? ld4(p + (size_t)row * 128 + slab * 32 + ch * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      };
      float4 pva[4], pvb[4];
      load_res(0, pva);
      auto do_slab = [&](uint32_t slab, float4 (&pv)[4], float4 (&pnext)[4]) {
        uint32_t v[16];
        tmem_ld16(t_lane + T_D2 + t * 128u + slab * 32u + c2 * 16u, v);
        if (slab < 3) load_res(slab + 1, pnext);       // residual rows of the next slab, one slab ahead
        tmem_ld_wait();
        if (slab == 3) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(d2_free + t);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(stg + m1 * 128u + (((c2 * 4 + j) ^ (m1 & 7u)) << 4)) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float4 bv = *reinterpret_cast<const float4*>(b2s + slab * 32 + ch * 4);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t m = pw * 16 + g * 4 + r4, row = row0 + m;
          const float4 a = *reinterpret_cast<const float4*>(stg + m * 128u + ((ch ^ (m & 7u)) << 4));
          const float mk = maskS[t * 128 + m];
          float4 o;
          o.x = (pv[g].x + (a.x + bv.x) * mk) * mk;
          o.y = (pv[g].y + (a.y + bv.y) * mk) * mk;
          o.z = (pv[g].z + (a.z + bv.z) * mk) * mk;
          o.w = (pv[g].w + (a.w + bv.w) * mk) * mk;
          if (row < M) st4(p + (size_t)row * 128 + slab * 32 + ch * 4, o);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
      };
      do_slab(0, pva, pvb);
      do_slab(1, pvb, pva);
      do_slab(2, pva, pvb);
      do_slab(3, pvb, pva);
    };
    if (iters > 0) { produce(0); produce(1); }
    for (uint32_t it = 0; it < iters; ++it) {
      // pair masks of this iteration's 256 pixels (read in E2 after its first named barrier)
      maskS[pw * 32 + lane] = pair_mask(2 * (blockIdx.x + it * gridDim.x) * 128u + pw * 32u + lane);
      if (it + 1 < iters) { produce(2 * it + 2); produce(2 * it + 3); }
      E2(0, it);
      E2(1, it);
    }
  } else {
    // ------------------------------------------------------------------ E1 / E2
    const uint32_t q = warp & 3, c2 = warp >> 2;
    const uint32_t t_lane = tmem + ((q * 32u) << 16);
    uint32_t n1 = 0;
    for (uint32_t it = 0; it < iters; ++it) {
      for (uint32_t c = 0; c < C; ++c)
#pragma unroll 1
        for (uint32_t t = 0; t < 2; ++t, ++n1) {
          // E1(t,c): this thread = pixel lane, hidden columns [64 c2, 64 c2 + 64) of the chunk
          const uint32_t k = it * C + c;
          mbar_wait(d1_full, n1 & 1u);
          tc_fence_after();
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t v[32], hp[16];
            tmem_ld32(t_lane + T_D1 + c2 * 64u + hf * 32u, v);
            tmem_ld_wait();
            if (hf == 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(d1_free);
            }
            const float2* bb = reinterpret_cast<const float2*>(b1s + c * 128u + c2 * 64u + hf * 32u);
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const float2 s = __fadd2_rn(make_float2(__uint_as_float(v[2 * e]), __uint_as_float(v[2 * e + 1])), bb[e]);
              hp[e] = relu_pack_bf16(s.x, s.y);
            }
            if (hf == 0) {              // H_t is still being read by G2(t, c-1) until h_free completes
              mbar_wait(h_free + t, (k & 1u) ^ 1u);
              tc_fence_after();
            }
            tmem_st16(t_lane + T_H + t * 64u + c2 * 32u + hf * 16u, hp);
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(h_full + t);
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
int tc_transition_init(gb_ctx* c) {
  cudaError_t e = cudaFuncSetAttribute(pair_transition_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K7_SMEM);
  if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "tc_transition_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  return tc_transition_x3_init(c);
}

// W1' = W1 * gamma (LayerNorm folded), b1' = b1 + W1 beta;  chunk images [chunk][2 kb][128 rows][64]
int tc_pack_transition(gb_ctx* c, TransitionW& w, const std::vector<float>& w1, const std::vector<float>& b1,
                       const std::vector<float>& w2, const std::vector<float>& ln_g, const std::vector<float>& ln_b) {
  const int nh = c->d.pair_transition_n, HID = 128 * nh;
  std::vector<uint16_t> i1((size_t)nh * 128 * 128), i2((size_t)nh * 128 * 128);
  for (int t = 0; t < nh; ++t) {
    pack_image(i1.data() + (size_t)t * 16384, 128, 128, [&](int r, int k) { return w1[(size_t)(t * 128 + r) * 128 + k] * ln_g[k]; });
    pack_image(i2.data() + (size_t)t * 16384, 128, 128, [&](int r, int k) { return w2[(size_t)r * HID + t * 128 + k]; });
  }
  std::vector<float> bf(HID);
  for (int n = 0; n < HID; ++n) {
    double acc = b1[n];
    for (int k = 0; k < 128; ++k) acc += (double)w1[(size_t)n * 128 + k] * ln_b[k];
    bf[n] = (float)acc;
  }
  w.tc_W1 = upload_bytes(c, i1.data(), i1.size() * 2);
  w.tc_W2 = upload_bytes(c, i2.data(), i2.size() * 2);
  w.tc_b1 = (float*)upload_bytes(c, bf.data(), bf.size() * 4);
  if (!w.tc_W1 || !w.tc_W2 || !w.tc_b1) return gb_fail(c, GB_ERR_CUDA, "tc_pack_transition: device allocation failed");
  return tc_pack_transition_x3(c, w, w1, w2, ln_g);
}

int tc_pair_transition(gb_ctx* c, const TransitionW& w, float* p, cudaStream_t st) {
  if (c->precision == GB_PREC_BF16X3) return tc_pair_transition_x3(c, w, p, st);
  size_t M = (size_t)c->B * c->N * c->N, px0 = 0;
  if (M >= (1ull << 31)) return gb_fail(c, GB_ERR_UNSUPPORTED, "pair map too large");
  if (c->shard_world > 1) {                         // B == 1: this rank's rows are one contiguous pixel range
    M /= c->shard_world;
    px0 = M * c->shard_rank;
  }
  TcState* s = static_cast<TcState*>(c->tc_state);
  const uint32_t npairs = (uint32_t)((M + 255) / 256);
  const uint32_t grid = npairs < (uint32_t)s->num_sms ? npairs : (uint32_t)s->num_sms;
  launch_pdl_cls(4, pair_transition_tc_kernel, grid, K7_THREADS, K7_SMEM, st, p + px0 * 128, (uint32_t)M, (uint32_t)px0, c->N, npairs, c->mask, (const bf16*)w.tc_W1, w.tc_b1,
                                                               (const bf16*)w.tc_W2, w.b2, (uint32_t)c->d.pair_transition_n);
  GB_LAUNCH_CHECK(c, "pair_transition_tc");
  return GB_OK;
}
