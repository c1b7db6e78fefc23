// tcgen05 / TMEM implementation of the structure network for GB_PREC_BF16
// (structure_net.py:76-116, invariant_point_attention.py:100-260, structure_transition.py:34-70).
// Activations that feed a GEMM are kept as bf16 operand tile images [K/64][Rp rows][64]
// (tc_common.cuh) written by the producing kernel's epilogue, so every GEMM moves both operands
// with cp.async.bulk and never converts in its main loop.
//   gemm_img_tc_kernel      C = A_img . W_img^T (+bias)(+residual)(relu) -> fp32 and/or bf16 image
//   ln_img_kernel           row LayerNorm (or plain copy) -> fp32 and/or bf16 image
//   ipa_prep_tc_kernel      once per eval: pair bias of all layers (one 128->96 GEMM over p) and
//                           the bf16 copy of p laid out as MN-major operand images per query row
//   ipa_attention_tc_kernel G query rows per CTA: logits / softmax / o / o_pt on CUDA cores,
//                           o_pair = a . z on tensor cores (z^T as MN-major A, probabilities as B)
#include <type_traits>
#include <vector>

#include "common.cuh"
#include "ipa_dims.h"
#include "shard.h"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

// ---------------------------------------------------------------------------------------------
// gemm_img_tc_kernel: one CTA = 128 rows x 64 output columns; warp 0 streams k-blocks of A (16 KB)
// and W (8 KB) through an 8-stage mbarrier ring, warp 1 issues the MMAs, all warps run the epilogue.
// ---------------------------------------------------------------------------------------------
// NS = 2 (GB_PREC_BF16X3): A, W (and the image output O) are hi | lo split components, component c `A_lo` / `W_lo` / `O_lo`
// elements behind component c - 1; a stage holds NS x [A_c 16K | W_c 8K] and every k step issues the passes of SplitPasses.
#define GI_STAGE_BYTES (16384 + 8192)
template <int NS> struct GI {
  // GB_PREC_BF16X3: the k-blocks are spread round robin over NACC tensor-memory accumulators that the epilogue adds in fp32
  // (round to nearest).  tcgen05.mma adds into its fp32 accumulator without rounding to nearest, so a long chain drifts
  // (1e-5 relative on K = 2112, above the 2^-17 of the hi | lo operands); short chains + an IEEE sum keep the split mode's
  // accuracy where its operands put it.
  static constexpr int NACC = NS == 1 ? 1 : 4;
  static constexpr int STAGES = NS == 1 ? 8 : 4;
  static constexpr uint32_t STAGE = NS * GI_STAGE_BYTES;
  static constexpr uint32_t SMEM = STAGES * STAGE + 1024 + 256;
};

struct GemmImgArgs {
  const bf16* A;      // [nkb][Rp][64]
  const bf16* W;      // [nkb][Nw][64]
  const float* bias;  // [Nw] or null
  const float* res;   // fp32 residual [R][ldr] or null
  float* C;           // fp32 out [R][ldc] or null
  bf16* O;            // bf16 image out [Nw/64][Rp][64] or null
  int Rp, Nw, nkb, R, ldc, ldr, relu;
  int mblk0;          // first 128-row block of this launch (row-sharded launches compute their own rows only)
  int kvsplit;        // > 0: IPA projection scatter, value = N (keys per sample): columns < IPA_QW go to rows [R][IPA_QW],
                      // the rest to key-major planes [B][IPA_NPLANE][N][4] behind them (see ipa_dims.h)
  size_t A_lo, W_lo, O_lo;   // NS > 1 only
  int ksplit;         // 3: blockIdx.z takes k-blocks [z nkb / 3, (z + 1) nkb / 3) and writes its partial sums to C / C1 / C2
                      // (no bias / residual / image: the consumer adds the partials in a fixed order)
  float *C1, *C2;
};

template <int NS>
__global__ void __launch_bounds__(256, 1) gemm_img_tc_kernel(const GemmImgArgs g) {
  constexpr int GI_STAGES = GI<NS>::STAGES;
  constexpr uint32_t STAGE = GI<NS>::STAGE;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + GI_STAGES * STAGE);
  uint64_t* empty = full + GI_STAGES;
  uint64_t* done = empty + GI_STAGES;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(done + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t n0 = blockIdx.x * 64u, m0 = (g.mblk0 + blockIdx.y) * 128u;
  const int nkb = g.ksplit > 1 ? g.nkb / g.ksplit : g.nkb, kb0 = g.ksplit > 1 ? (int)blockIdx.z * nkb : 0;   // this CTA's k-blocks
  if (tid == 0) {
    for (int s = 0; s < GI_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  constexpr int NACC = GI<NS>::NACC;
  if (warp == 2) tmem_alloc(tslot, 64 * NACC);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  if (warp == 0) {
    if (lane == 0) {
      // the weight halves of the first ring of stages do not depend on the previous kernel: in flight before pdl_wait
      const int npre = nkb < GI_STAGES ? nkb : GI_STAGES;
      for (int kb = 0; kb < npre; ++kb) {
        mbar_expect_tx(full + kb, STAGE);
        const bf16* wsrc = g.W + ((size_t)(kb0 + kb) * g.Nw + n0) * 64;
#pragma unroll
        for (int c = 0; c < NS; ++c) bulk_g2s(sm + kb * STAGE + c * GI_STAGE_BYTES + 16384, wsrc + c * g.W_lo, 8192, full + kb);
      }
      pdl_wait();
      pdl_launch();
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % GI_STAGES;
        uint8_t* st = sm + s * STAGE;
        if (kb >= GI_STAGES) {
          mbar_wait(empty + s, ((kb / GI_STAGES) - 1) & 1);
          mbar_expect_tx(full + s, STAGE);
          const bf16* wsrc = g.W + ((size_t)(kb0 + kb) * g.Nw + n0) * 64;
#pragma unroll
          for (int c = 0; c < NS; ++c) bulk_g2s(st + c * GI_STAGE_BYTES + 16384, wsrc + c * g.W_lo, 8192, full + s);
        }
        const bf16* asrc = g.A + ((size_t)(kb0 + kb) * g.Rp + m0) * 64;
#pragma unroll
        for (int c = 0; c < NS; ++c) bulk_g2s(st + c * GI_STAGE_BYTES, asrc + c * g.A_lo, 16384, full + s);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc(128, 64), tm = warp_uniform(tmem);
    for (int kb = 0; kb < nkb; ++kb) {
      const int s = kb % GI_STAGES;
      mbar_wait(full + s, (kb / GI_STAGES) & 1);
      tc_fence_after();
      const uint32_t sa = smem_u32(sm + s * STAGE), sb = sa + 16384;
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int ps = 0; ps < SplitPasses<NS>::N; ++ps)
            umma(tm + (uint32_t)(kb % NACC) * 64u, umma_desc(sa + SplitPasses<NS>::a(ps) * GI_STAGE_BYTES + k * 32),
                 umma_desc(sb + SplitPasses<NS>::b(ps) * GI_STAGE_BYTES + k * 32), idesc, ((kb / NACC) | k | ps) ? 1u : 0u);
        umma_commit(empty + s);
        if (kb == nkb - 1) umma_commit(done);
      }
      __syncwarp();
    }
  }
  pdl_wait();        // (the residual read and the output writes of the epilogue; the loader thread has already waited)
  mbar_wait(done, 0);
  tc_fence_after();
  __syncwarp();
  const uint32_t q = warp & 3, h = warp >> 2, m = m0 + q * 32 + lane, c0 = n0 + h * 32;
  uint32_t v[32];
  tmem_ld32(tmem + ((q * 32u) << 16) + h * 32, v);
  tmem_ld_wait();
#pragma unroll
  for (int a = 1; a < NACC; ++a) {
    if (a < nkb) {                       // (warp-uniform) accumulators no k-block went to hold nothing
      uint32_t w[32];
      tmem_ld32(tmem + ((q * 32u) << 16) + a * 64 + h * 32, w);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) + __uint_as_float(w[e]));
    }
  }
  if (m < (uint32_t)g.R) {
    float o[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      float t = __uint_as_float(v[e]) + (g.bias ? g.bias[c0 + e] : 0.f);
      if (g.relu) t = fmaxf(t, 0.f);
      o[e] = t;
    }
    if (g.res) {
      const float4* r4 = reinterpret_cast<const float4*>(g.res + (size_t)m * g.ldr + c0);
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float4 r = r4[e]; o[4 * e] += r.x; o[4 * e + 1] += r.y; o[4 * e + 2] += r.z; o[4 * e + 3] += r.w; }
    }
    if (g.C && g.kvsplit) {
      const uint32_t bb = m / (uint32_t)g.kvsplit, jj = m - bb * (uint32_t)g.kvsplit;
      float* planes = g.C + (size_t)g.R * IPA_QW;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const uint32_t col = c0 + 4 * e;
        float* dst = col < IPA_QW ? g.C + (size_t)m * IPA_QW + col
                                  : planes + (((size_t)bb * IPA_NPLANE + ((col - IPA_QW) >> 2)) * g.kvsplit + jj) * 4;
        *reinterpret_cast<float4*>(dst) = make_float4(o[4 * e], o[4 * e + 1], o[4 * e + 2], o[4 * e + 3]);
      }
    } else if (g.C) {
      float* cbase = g.ksplit > 1 && blockIdx.z == 1 ? g.C1 : g.ksplit > 1 && blockIdx.z == 2 ? g.C2 : g.C;
      float4* c4 = reinterpret_cast<float4*>(cbase + (size_t)m * g.ldc + c0);
#pragma unroll
      for (int e = 0; e < 8; ++e) c4[e] = make_float4(o[4 * e], o[4 * e + 1], o[4 * e + 2], o[4 * e + 3]);
    }
    if (g.O) {
      uint8_t* row = reinterpret_cast<uint8_t*>(g.O + ((size_t)blockIdx.x * g.Rp + m) * 64);
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t c0[NS], c1[NS], c2[NS], c3[NS];
        split_bf16x2_n<NS>(o[8 * ch + 0], o[8 * ch + 1], c0);
        split_bf16x2_n<NS>(o[8 * ch + 2], o[8 * ch + 3], c1);
        split_bf16x2_n<NS>(o[8 * ch + 4], o[8 * ch + 5], c2);
        split_bf16x2_n<NS>(o[8 * ch + 6], o[8 * ch + 7], c3);
#pragma unroll
        for (int c = 0; c < NS; ++c)
          *reinterpret_cast<uint4*>(row + (size_t)c * g.O_lo * 2 + (((h * 4 + ch) ^ (m & 7u)) << 4)) = make_uint4(c0[c], c1[c], c2[c], c3[c]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 64 * NACC);
}

// Row-sharded runs: the split-K partial sums of this rank's rows are added here, in ln_img_kernel's order
// ((((p0 + p1) + p2) + bias) + residual), so that the all-gathered rows are bit-identical to the one-GPU result.
__global__ void sum_partials_kernel(float* __restrict__ p0, const float* __restrict__ p1, const float* __restrict__ p2,
                                    const float* __restrict__ bias, const float* __restrict__ res, int row0, int rows, int C) {
  pdl_wait();
  pdl_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x, per_row = C >> 2;
  if (i >= rows * per_row) return;
  const int c = (i % per_row) * 4;
  const size_t o = (size_t)(row0 + i / per_row) * C + c;
  float4 v = ld4(p0 + o);
  const float4 a = ld4(p1 + o), b = ld4(p2 + o), bi = ld4(bias + c), re = ld4(res + o);
  v.x = (((v.x + a.x) + b.x) + bi.x) + re.x; v.y = (((v.y + a.y) + b.y) + bi.y) + re.y;
  v.z = (((v.z + a.z) + b.z) + bi.z) + re.z; v.w = (((v.w + a.w) + b.w) + bi.w) + re.w;
  st4(p0 + o, v);
}

// ---------------------------------------------------------------------------------------------
// ln_img_kernel: one warp per row of [R, C] (C % 128 == 0, C <= 512): optional LayerNorm, then any
// of: fp32 out, second fp32 out (the 'states' list), bf16 operand image [C/64][Rp][64].
// ---------------------------------------------------------------------------------------------
__global__ void ln_img_kernel(const float* __restrict__ in, float* __restrict__ out, float* __restrict__ out2,
                              bf16* __restrict__ img, const float* __restrict__ gamma, const float* __restrict__ beta,
                              int R, int C, int Rp, int normalize, const float* __restrict__ Wbb,
                              const float* __restrict__ bbb, float* __restrict__ rots,
                              float* __restrict__ trans, size_t img_lo, int ns /* split components of the image, img_lo elements apart */,
                              const float* __restrict__ in2, const float* __restrict__ in3 /* split-K partial sums, or null */,
                              const float* __restrict__ addb /*[C] bias*/, const float* __restrict__ addr /*[R][C] residual*/) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  pdl_wait();
  pdl_launch();
  if (row >= R) return;
  const int nq = C >> 7;
  float4 v[4];
  float sum = 0.f;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    v[q] = q < nq ? ld4(in + (size_t)row * C + (q * 32 + lane) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (in2 && q < nq) {      // the input is a split-K GEMM: (((p0 + p1) + p2) + bias) + residual, always in this order
      const size_t o = (size_t)row * C + (q * 32 + lane) * 4;
      const float4 a = ld4(in2 + o), bq = ld4(in3 + o), bi = ld4(addb + (q * 32 + lane) * 4), re = ld4(addr + o);
      v[q].x = (((v[q].x + a.x) + bq.x) + bi.x) + re.x; v[q].y = (((v[q].y + a.y) + bq.y) + bi.y) + re.y;
      v[q].z = (((v[q].z + a.z) + bq.z) + bi.z) + re.z; v[q].w = (((v[q].w + a.w) + bq.w) + bi.w) + re.w;
    }
    sum += v[q].x + v[q].y + v[q].z + v[q].w;
  }
  if (normalize) {
    const float mean = warp_sum(sum) / (float)C;
    float d2 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (q >= nq) break;
      v[q].x -= mean; v[q].y -= mean; v[q].z -= mean; v[q].w -= mean;
      d2 += v[q].x * v[q].x + v[q].y * v[q].y + v[q].z * v[q].z + v[q].w * v[q].w;
    }
    const float rstd = rsqrtf(warp_sum(d2) / (float)C + 1e-5f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (q >= nq) break;
      const float4 g4 = ld4(gamma + (q * 32 + lane) * 4), b4 = ld4(beta + (q * 32 + lane) * 4);
      v[q].x = v[q].x * rstd * g4.x + b4.x; v[q].y = v[q].y * rstd * g4.y + b4.y;
      v[q].z = v[q].z * rstd * g4.z + b4.z; v[q].w = v[q].w * rstd * g4.w + b4.w;
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (q >= nq) break;
    const int c = (q * 32 + lane) * 4;
    if (out) st4(out + (size_t)row * C + c, v[q]);
    if (out2) st4(out2 + (size_t)row * C + c, v[q]);
    if (img) {
      uint2 o, ol;
      split_bf16x2(v[q].x, v[q].y, o.x, ol.x);
      split_bf16x2(v[q].z, v[q].w, o.y, ol.y);
      uint8_t* dst = reinterpret_cast<uint8_t*>(img + ((size_t)(c >> 6) * Rp + row) * 64) +
                     (((uint32_t)((c & 63) >> 3) ^ (row & 7u)) << 4) + ((c & 7) >> 2) * 8;
      *reinterpret_cast<uint2*>(dst) = o;
      if (ns > 1) *reinterpret_cast<uint2*>(dst + img_lo * 2) = ol;
    }
  }
  if (Wbb) {
    // BackboneUpdate of this row on the row just written (backbone_update.py:40-66, affine_utils.py:109-116): the warp
    // already holds it in registers, so the separate kernel (one more pass over s and one more launch) is folded in
    float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (q >= nq) break;
      const int c = (q * 32 + lane) * 4;
#pragma unroll
      for (int o = 0; o < 6; ++o) {
        const float4 w4 = ld4(Wbb + o * C + c);
        acc[o] = fmaf(v[q].x, w4.x, acc[o]); acc[o] = fmaf(v[q].y, w4.y, acc[o]);
        acc[o] = fmaf(v[q].z, w4.z, acc[o]); acc[o] = fmaf(v[q].w, w4.w, acc[o]);
      }
    }
#pragma unroll
    for (int o = 0; o < 6; ++o) acc[o] = warp_sum(acc[o]) + bbb[o];
    if (lane == 0) {
      const float denom = sqrtf(acc[0] * acc[0] + acc[1] * acc[1] + acc[2] * acc[2] + 1.f);
      float ru[9], r1[9], rn[9];
      quat_to_rot(1.f / denom, acc[0] / denom, acc[1] / denom, acc[2] / denom, ru);
      float* Rm = rots + (size_t)row * 9;
      float* Tm = trans + (size_t)row * 3;
#pragma unroll
      for (int e = 0; e < 9; ++e) r1[e] = Rm[e];
      mat3_mul(r1, ru, rn);
      const float tx = acc[3], ty = acc[4], tz = acc[5];
      const float nt0 = r1[0] * tx + r1[1] * ty + r1[2] * tz + Tm[0];
      const float nt1 = r1[3] * tx + r1[4] * ty + r1[5] * tz + Tm[1];
      const float nt2 = r1[6] * tx + r1[7] * ty + r1[8] * tz + Tm[2];
#pragma unroll
      for (int e = 0; e < 9; ++e) Rm[e] = rn[e];
      Tm[0] = nt0; Tm[1] = nt1; Tm[2] = nt2;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// split_img_kernel: fp32 rows [R, C] -> three bf16 operand images hi | lo | hi (hi = bf16(x), lo = bf16(x - hi)) in
// three consecutive [C/64][Rp][64] buffers.  Against a weight image [W_hi | W_hi | W_lo] one GEMM of K = 3 C computes
// x_hi W_hi + x_lo W_hi + x_hi W_lo: the product to ~2^-16 relative, from bf16 tensor-core passes.  One warp per row.
// ---------------------------------------------------------------------------------------------
__global__ void split_img_kernel(const float* __restrict__ in, bf16* __restrict__ img, int R, int C, int Rp, size_t stride /*elements between the three images*/) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  pdl_wait();
  pdl_launch();
  if (row >= R) return;
  for (int c = lane * 4; c < C; c += 128) {
    const float4 v = ld4(in + (size_t)row * C + c);
    uint2 hi, lo;
    hi.x = pack_bf16(v.x, v.y);
    hi.y = pack_bf16(v.z, v.w);
    lo.x = pack_bf16(v.x - __uint_as_float(hi.x << 16), v.y - __uint_as_float(hi.x & 0xffff0000u));
    lo.y = pack_bf16(v.z - __uint_as_float(hi.y << 16), v.w - __uint_as_float(hi.y & 0xffff0000u));
    uint8_t* dst = reinterpret_cast<uint8_t*>(img + ((size_t)(c >> 6) * Rp + row) * 64) +
                   (((uint32_t)((c & 63) >> 3) ^ ((uint32_t)row & 7u)) << 4) + ((c & 7) >> 2) * 8;
    *reinterpret_cast<uint2*>(dst) = hi;
    *reinterpret_cast<uint2*>(dst + stride * 2) = lo;
    *reinterpret_cast<uint2*>(dst + stride * 4) = hi;
  }
}

// ---------------------------------------------------------------------------------------------
// ipa_prep_tc_kernel (once per denoiser eval, after the pair stack).  p does not change across the
// structure layers, so for all layers l at once:
//   pbias[l][b][i][h][j] = sqrt(1/3) * (linear_b_l(z_ij)[h])      (invariant_point_attention.py:184-190)
// (the mask term 1e5 (m_i m_j - 1) of :206-212 is added by the attention kernel, last, as the reference does: on a masked
//  query row every logit sits at -1e5 where fp32 has 7 fractional bits, so the ORDER of the additions decides the result)
//   z_img[b][i][mb][j][64]  bf16 copy of z_ij[64 mb ..], chunk-swizzled by j & 7: for a fixed query
//                           row i this is the MN-major (channel-contiguous) A operand of o_pair.
// One CTA = 128 pixels (2 rows x 64 cols); Wb image [2 kb][96 rows = l*12+h][64]; bias96 pre-scaled.
// ---------------------------------------------------------------------------------------------
// NS = 2 (GB_PREC_BF16X3): A, the Wb image (components directly behind each other) and z_img (`z_lo` elements apart) are
// hi | lo split components and the pair-bias GEMM runs the passes of SplitPasses<NS>; two CTAs per SM instead of three.
template <int NS> struct PrepSmem {
  static constexpr uint32_t A = NS * 32768, W = NS * 24576, TOTAL = A + W + 1024 + 64;
};

template <int NS>
__global__ void __launch_bounds__(256, NS == 1 ? 3 : 2) ipa_prep_tc_kernel(const float* __restrict__ p, int B, int N, int Np,
                                                             const bf16* __restrict__ Wimg,
                                                             const float* __restrict__ bias96, float* __restrict__ pbias,
                                                             bf16* __restrict__ z_img, int ncol, const TileRange tr, size_t z_lo) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* A = sm;
  uint8_t* W = sm + PrepSmem<NS>::A;
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(sm + PrepSmem<NS>::A + PrepSmem<NS>::W);
  uint64_t* bar_mma = bar_w + 1;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_w + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const PixTile T = pix_tile(blockIdx.x, tr);
  if (tid == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
    mbar_expect_tx(bar_w, PrepSmem<NS>::W);
    bulk_g2s(W, Wimg, PrepSmem<NS>::W, bar_w);
  }
  if (warp == 1) tmem_alloc(tslot, 128);
  pdl_wait();            // p comes from the last pair-transition kernel
  {
    const uint32_t kb = lane >> 4, chunk = (lane & 15) >> 1, half = lane & 1;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      float4 v[4];
      bool ok[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t m = warp * 16 + it * 4 + u, i = 2 * T.t2 + (m >> 6), j = 64 * T.jb + (m & 63);
        ok[u] = i < (uint32_t)N && j < (uint32_t)N;
        v[u] = ok[u] ? ld4(p + (((size_t)T.b * N + i) * N + j) * 128 + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t m = warp * 16 + it * 4 + u, i = 2 * T.t2 + (m >> 6), j = 64 * T.jb + (m & 63);
        uint32_t cx[NS], cy[NS];
        split_bf16x2_n<NS>(v[u].x, v[u].y, cx);
        split_bf16x2_n<NS>(v[u].z, v[u].w, cy);
        const uint32_t inrow = ((chunk ^ (m & 7u)) << 4) + half * 8u;
#pragma unroll
        for (int c = 0; c < NS; ++c) *reinterpret_cast<uint2*>(A + c * 32768u + kb * 16384u + m * 128u + inrow) = make_uint2(cx[c], cy[c]);
        if (ok[u]) {
          uint8_t* zr = reinterpret_cast<uint8_t*>(z_img) + ((((size_t)T.b * N + i) * 2 + kb) * Np + j) * 128;
#pragma unroll
          for (int c = 0; c < NS; ++c) *reinterpret_cast<uint2*>(zr + (size_t)c * z_lo * 2 + inrow) = make_uint2(cx[c], cy[c]);
        }
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  pdl_launch();
  if (tid == 0) {
    mbar_wait(bar_w, 0);
    tc_fence_after();
#pragma unroll
    for (int ps = 0; ps < SplitPasses<NS>::N; ++ps)
      umma_tile(tmem, smem_u32(A) + SplitPasses<NS>::a(ps) * 32768u, smem_u32(W) + SplitPasses<NS>::b(ps) * 24576u, 96, 2, ps > 0);
    umma_commit(bar_mma);
  }
  __syncwarp();
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  const uint32_t q = warp & 3, h = warp >> 2, m = q * 32 + lane;
  const uint32_t i = 2 * T.t2 + (m >> 6), j = 64 * T.jb + (m & 63);
  const bool valid = i < (uint32_t)N && j < (uint32_t)N;
  // warps 0-3: column chunks 0 and 2; warps 4-7: chunk 1.  The chunk index is a compile-time constant inside the lambda so
  // that layer / head of every column (n / 12, n % 12) fold into immediates: with a run-time chunk the divisions and the
  // 64-bit address arithmetic of this loop were 60 % of the kernel's instructions.
  {
    float* const pb0 = pbias + (((size_t)T.b * N + i) * 12) * Np + j;
    const size_t lstride = (size_t)B * N * 12 * Np;
    auto chunk = [&](auto cc_) {
      constexpr int cc = decltype(cc_)::value;
      uint32_t v[32];
      tmem_ld32(tmem + ((q * 32u) << 16) + cc * 32, v);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int n = cc * 32 + e, l = n / 12, hh = n - l * 12;
          if (n < ncol) pb0[(size_t)l * lstride + (size_t)hh * Np] = __uint_as_float(v[e]) + __ldg(bias96 + n);
        }
      }
    };
    if (h == 0) {
      chunk(std::integral_constant<int, 0>{});
      chunk(std::integral_constant<int, 2>{});
    } else {
      chunk(std::integral_constant<int, 1>{});
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------------------------
// ipa_points_planes_kernel: T.apply on the projected points (invariant_point_attention.py:149,162) in the
// Q-row / key-plane layout of ipa_dims.h.  One thread per (row, head): 4 q points, 4 k points, 8 v points.
// ---------------------------------------------------------------------------------------------
__global__ void ipa_points_planes_kernel(float* __restrict__ qkv, const float* __restrict__ rots, const float* __restrict__ trans,
                                         int R, int N) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_wait();
  pdl_launch();
  if (idx >= R * IPA_H) return;
  const int row = idx / IPA_H, h = idx - row * IPA_H, b = row / N, j = row - b * N;
  const float* Rm = rots + (size_t)row * 9;
  const float* T = trans + (size_t)row * 3;
  const float r0 = Rm[0], r1 = Rm[1], r2 = Rm[2], r3 = Rm[3], r4 = Rm[4], r5 = Rm[5], r6 = Rm[6], r7 = Rm[7], r8 = Rm[8];
  const float t0 = T[0], t1 = T[1], t2 = T[2];
  auto apply4 = [&](float4& x, float4& y, float4& z) {
    const float4 a = x, bq = y, cz = z;
    x = make_float4(r0 * a.x + r1 * bq.x + r2 * cz.x + t0, r0 * a.y + r1 * bq.y + r2 * cz.y + t0, r0 * a.z + r1 * bq.z + r2 * cz.z + t0,
                    r0 * a.w + r1 * bq.w + r2 * cz.w + t0);
    y = make_float4(r3 * a.x + r4 * bq.x + r5 * cz.x + t1, r3 * a.y + r4 * bq.y + r5 * cz.y + t1, r3 * a.z + r4 * bq.z + r5 * cz.z + t1,
                    r3 * a.w + r4 * bq.w + r5 * cz.w + t1);
    z = make_float4(r6 * a.x + r7 * bq.x + r8 * cz.x + t2, r6 * a.y + r7 * bq.y + r8 * cz.y + t2, r6 * a.z + r7 * bq.z + r8 * cz.z + t2,
                    r6 * a.w + r7 * bq.w + r8 * cz.w + t2);
  };
  {   // q points: Q row, 192 + c*48 + h*4 + p
    float4* q = reinterpret_cast<float4*>(qkv + (size_t)row * IPA_QW + 192 + h * 4);
    float4 x = q[0], y = q[12], z = q[24];
    apply4(x, y, z);
    q[0] = x; q[12] = y; q[24] = z;
  }
  float4* pl = reinterpret_cast<float4*>(qkv + (size_t)R * IPA_QW) + ((size_t)b * IPA_NPLANE + h * IPA_PLANES_PER_HEAD) * N + j;
  {   // k points: planes e = 4, 5, 6
    float4 x = pl[(size_t)4 * N], y = pl[(size_t)5 * N], z = pl[(size_t)6 * N];
    apply4(x, y, z);
    pl[(size_t)4 * N] = x; pl[(size_t)5 * N] = y; pl[(size_t)6 * N] = z;
  }
#pragma unroll
  for (int pg = 0; pg < 2; ++pg) {   // v points: x planes 11, 12; y 13, 14; z 15, 16
    float4 x = pl[(size_t)(11 + pg) * N], y = pl[(size_t)(13 + pg) * N], z = pl[(size_t)(15 + pg) * N];
    apply4(x, y, z);
    pl[(size_t)(11 + pg) * N] = x; pl[(size_t)(13 + pg) * N] = y; pl[(size_t)(15 + pg) * N] = z;
  }
}

// ---------------------------------------------------------------------------------------------
// ipa_attention_tc_kernel (invariant_point_attention.py:181-259), Np <= 1024.
// One CTA = G consecutive query rows i0.. of sample b (IA_THREADS threads):
//   1. logits: thread j owns key j; per head it loads k / k_pts of its key once (registers) and scores
//      it against the G queries (q, q_pts broadcast from shared memory); + precomputed pair bias.
//   2. softmax over j per (head, query) in shared memory; probabilities are kept fp32 (for o, o_pt)
//      and written as bf16 B-operand images a_img[g] = [Np/64 kb][16 rows = head][64 j].
//   3. warp 0 streams z_img of each query row in 64-key chunks (2 x 8 KB, cp.async.bulk, 4-deep ring),
//      warp 1 issues D_g[c][h] += zT[c][j] . a_g[h][j]  (M = 128 channels, N = 16, MN-major A),
//      warps 2-7 meanwhile accumulate o and o_pt on CUDA cores.
//   4. epilogue: inverse frame + norms for o_pt; everything is written as bf16 straight into the
//      operand image of the output projection (cat_img, K = 2112: o | o_pt.x | .y | .z | |o_pt| | o_pair).
// ---------------------------------------------------------------------------------------------
#define IA_ZRING 2
#define IA_THREADS 768   // the CUDA-core phases run at ~0.2 IPC per warp (dependent FMA chains): occupancy, not bandwidth, bounds them
// Queries per CTA (template parameter G): every CTA streams the whole K / V of its sample from L2, so that traffic
// per query is 1 / G; the fp32 logits [12][Np][G] and the probability images must fit in shared memory:
// G = 8 up to 256 keys, 4 up to 512, 2 up to 1024.
// ns = 2 (GB_PREC_BF16X3): ring slots hold the hi | lo components of a z chunk (2 x 16 KB), the probability images exist as
// a hi and a lo set
struct IaSmem { uint32_t aimg, P, qs, red, fr, bar, total; };
__host__ __device__ inline IaSmem ia_smem(int G, int Np, int ns = 1) {
  IaSmem o;
  o.aimg = IA_ZRING * ns * 16384;
  o.P = o.aimg + ns * G * (Np >> 6) * 2048;
  o.qs = o.P + 12 * Np * G * 4;
  o.red = o.qs + G * 12 * 28 * 4;
  o.fr = o.red + G * 480 * 4;
  o.bar = o.fr + G * 12 * 4;
  o.total = o.bar + 128 + 1024;
  return o;
}
// Short chains and small batches: a CTA's fixed cost (prologue, 120 reductions, epilogue) dominates, so take the smallest
// G that still fits all CTAs in one wave (B * ceil(N / G) <= SMs); otherwise the largest the shared memory allows.
static inline int ia_queries_per_cta(int Np, int B = 0, int N = 0, int num_sms = 0, int ns = 1) {
  int gmax = Np <= 256 ? 8 : Np <= 512 ? 4 : 2;
  if (ns == 2) gmax = gmax > 2 ? gmax / 2 : 0;     // hi | lo probability images: half the queries; Np > 512 does not fit
  if (gmax == 0) return 0;
  for (int g = 2; g < gmax && num_sms > 0; g *= 2)
    if (B * ((N + g - 1) / g) <= num_sms) return g;
  return gmax;
}

// G consecutive floats of the logit / probability array
template <int G> __device__ __forceinline__ void p_load(const float* pp, float (&t)[G]) {
  if constexpr (G % 4 == 0) {
#pragma unroll
    for (int g4 = 0; g4 < G / 4; ++g4) {
      const float4 v = *reinterpret_cast<const float4*>(pp + 4 * g4);
      t[4 * g4] = v.x; t[4 * g4 + 1] = v.y; t[4 * g4 + 2] = v.z; t[4 * g4 + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int g2 = 0; g2 < G / 2; ++g2) {
      const float2 v = *reinterpret_cast<const float2*>(pp + 2 * g2);
      t[2 * g2] = v.x; t[2 * g2 + 1] = v.y;
    }
  }
}
template <int G> __device__ __forceinline__ void p_store(float* pp, const float (&t)[G]) {
  if constexpr (G % 4 == 0) {
#pragma unroll
    for (int g4 = 0; g4 < G / 4; ++g4) *reinterpret_cast<float4*>(pp + 4 * g4) = make_float4(t[4 * g4], t[4 * g4 + 1], t[4 * g4 + 2], t[4 * g4 + 3]);
  } else {
#pragma unroll
    for (int g2 = 0; g2 < G / 2; ++g2) *reinterpret_cast<float2*>(pp + 2 * g2) = make_float2(t[2 * g2], t[2 * g2 + 1]);
  }
}

// 2-byte store of element k of row r of a [.][Rp][64] operand image
template <int NS = 1>
__device__ __forceinline__ void img_store(bf16* img, int Rp, uint32_t r, uint32_t k, float v, size_t lo = 0) {
  uint8_t* dst = reinterpret_cast<uint8_t*>(img + ((size_t)(k >> 6) * Rp + r) * 64) + ((((k & 63) >> 3) ^ (r & 7u)) << 4) + ((k & 7) << 1);
#pragma unroll
  for (int c = 0; c < NS; ++c) {
    const bf16 hc = __float2bfloat16_rn(v);
    *reinterpret_cast<bf16*>(dst + (size_t)c * lo * 2) = hc;
    v -= __bfloat162float(hc);
  }
}

template <int G, int NS>
__global__ void __launch_bounds__(IA_THREADS, 1) ipa_attention_tc_kernel(
    const float* __restrict__ qkv, const float* __restrict__ pbias, const bf16* __restrict__ z_img,
    const float* __restrict__ rots, const float* __restrict__ trans, const float* __restrict__ hw,
    bf16* __restrict__ cat_img, int N, int Np, int Rp, int row0, const int* __restrict__ rmask, size_t z_lo, size_t cat_lo) {
  constexpr uint32_t ZSLOT = NS * 16384u;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  const IaSmem L = ia_smem(G, Np, NS);
  uint8_t* zring = sm;
  uint8_t* aimg = sm + L.aimg;                                 // [g][Np/64 kb][16 rows = head][64 j]
  float* P = reinterpret_cast<float*>(sm + L.P);               // [12 h][Np j][G]
  float* qs = reinterpret_cast<float*>(sm + L.qs);             // [g][h][28]: q*sc (16) | q_pts c*4+p (12)
  float* oacc = reinterpret_cast<float*>(sm + L.red);          // [g][h*40 + e]: o (e < 16) | o_pt global frame (16 + c*8 + p)
  float* fr = reinterpret_cast<float*>(sm + L.fr);             // [g][12] R (9) t (3)
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + L.bar);
  uint64_t* empty = full + IA_ZRING;
  uint64_t* done = empty + IA_ZRING;                           // [G]
  uint32_t* tslot = reinterpret_cast<uint32_t*>(done + G);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, i0 = row0 + blockIdx.x * G;        // row0: first query row of a row-sharded launch
  const int ng = min(G, N - i0), nch = Np >> 6;
  const uint32_t aimg_g = (uint32_t)nch * 2048u;               // bytes of one query's probability image
  constexpr uint32_t TCOLS = G * 16 < 32 ? 32 : G * 16;
  const size_t rowb = (size_t)b * N;

  if (tid == 0) {
    for (int s = 0; s < IA_ZRING; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    for (int g = 0; g < G; ++g) mbar_init(done + g, 1);
    fence_mbar_init();
    // prefetch the first ring of z chunks: independent of everything this kernel computes
    const int tot = ng * nch;
    for (int it = 0; it < min(tot, IA_ZRING); ++it) {
      const int g = it / nch, ch = it - g * nch;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(z_img) + ((rowb + i0 + g) * 2 * Np + ch * 64) * 128;
      mbar_expect_tx(full + it, ZSLOT);
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        bulk_g2s(zring + it * ZSLOT + c * 16384, src + (size_t)c * z_lo * 2, 8192, full + it);
        bulk_g2s(zring + it * ZSLOT + c * 16384 + 8192, src + (size_t)c * z_lo * 2 + (size_t)Np * 128, 8192, full + it);
      }
    }
  }
  if (warp == 2) tmem_alloc(tslot, TCOLS);
  if (warp == 3 && lane < ng) {
    // L2 prefetch of what this CTA will stream later, all of it written by ipa_prep several launches back: the z rows of
    // its queries (2 k-blocks x Np x 128 B per query and component; phase 3 was bound by their HBM latency through the
    // two-slot ring) and the pair-bias rows (12 x Np floats per query, read in the logits phase)
    const size_t row = rowb + i0 + lane;
#pragma unroll
    for (int c = 0; c < NS; ++c) {
      const uint8_t* zr = reinterpret_cast<const uint8_t*>(z_img) + (size_t)c * z_lo * 2 + row * 2 * Np * 128;
      for (uint32_t off = 0; off < 2u * Np * 128u; off += 32768u) l2_prefetch(zr + off, min(32768u, 2u * Np * 128u - off));
    }
    l2_prefetch(pbias + row * 12 * Np, 12u * Np * 4u);
  }
  pdl_wait();          // the z prefetch above reads what ipa_prep wrote (several launches back); q / k / v come from the previous kernel
  // q, q_pts of the G queries; a_img cleared (rows 12..15 and padded keys must read as zero)
  const float sc_qk = sqrtf(1.f / (3.f * IPA_C));
  for (int e = tid; e < G * 12 * 28; e += IA_THREADS) {
    const int g = e / 336, r = e - g * 336, h = r / 28, x = r - h * 28;
    float v = 0.f;
    if (g < ng) {
      const float* row = qkv + (rowb + i0 + g) * IPA_QW;
      v = x < 16 ? row[h * 16 + x] * sc_qk : row[192 + ((x - 16) >> 2) * 48 + h * 4 + ((x - 16) & 3)];
    }
    qs[e] = v;
  }
  for (int e = tid; e < (int)(NS * G * aimg_g / 16); e += IA_THREADS) reinterpret_cast<uint4*>(aimg)[e] = make_uint4(0, 0, 0, 0);
  for (int e = tid; e < G * 480; e += IA_THREADS) oacc[e] = 0.f;
  if (tid < G * 12) {
    const int g = tid / 12, x = tid - g * 12;
    fr[tid] = g < ng ? (x < 9 ? rots[(rowb + i0 + g) * 9 + x] : trans[(rowb + i0 + g) * 3 + x - 9]) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  pdl_launch();        // after the tensor-memory allocation: a dependent CTA on this SM can no longer get in its way

  // ---- 1. logits: thread (j mod 256, head group); keys in blocks of 256
  {
    constexpr int HPG = 12 / (IA_THREADS / 256);   // heads per thread group
    const int hg = tid >> 8;
    float mq[G];                                  // residue masks of the G queries
#pragma unroll
    for (int g = 0; g < G; ++g) mq[g] = g < ng ? (float)rmask[rowb + i0 + g] : 0.f;
    for (int j = tid & 255; j < Np; j += 256) {
      const bool kv = j < N;
      const float mk = kv ? (float)rmask[rowb + j] : 0.f;
      const float4* kpl = reinterpret_cast<const float4*>(qkv + (size_t)gridDim.y * N * IPA_QW) + (size_t)b * IPA_NPLANE * N + (kv ? j : 0);
#pragma unroll
      for (int hh = 0; hh < HPG; ++hh) {       // unrolled: the K planes and pair-bias words of later heads are in flight early
        const int h = hg * HPG + hh;
        float out[G];
#pragma unroll
        for (int g = 0; g < G; ++g) out[g] = -3.0e38f;
        if (kv) {
          // the G pair-bias words of this (head, key) first, all in flight together
          float pbv[G];
#pragma unroll
          for (int g = 0; g < G; ++g) pbv[g] = g < ng ? __ldg(pbias + ((rowb + i0 + g) * 12 + h) * Np + j) : 0.f;
          // packed fp32x2 arithmetic: kk[0..7] = k (two components per register pair), kk[8..13] = -k_pts
          float2 kk[14];
#pragma unroll
          for (int e = 0; e < 7; ++e) {       // planes 0..3 = k, 4..6 = k_pts x / y / z: consecutive keys, consecutive 16-byte words
            const float4 t = kpl[(size_t)(h * IPA_PLANES_PER_HEAD + e) * N];
            kk[2 * e] = e < 4 ? make_float2(t.x, t.y) : make_float2(-t.x, -t.y);
            kk[2 * e + 1] = e < 4 ? make_float2(t.z, t.w) : make_float2(-t.z, -t.w);
          }
          const float gam = -0.5f * hw[h];
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const float4* q4 = reinterpret_cast<const float4*>(qs + (g * 12 + h) * 28);
            float2 qk = make_float2(0.f, 0.f), d2 = make_float2(0.f, 0.f);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float4 t = q4[e];
              qk = __ffma2_rn(make_float2(t.x, t.y), kk[2 * e], qk);
              qk = __ffma2_rn(make_float2(t.z, t.w), kk[2 * e + 1], qk);
            }
#pragma unroll
            for (int e = 4; e < 7; ++e) {
              const float4 t = q4[e];
              const float2 da = __fadd2_rn(make_float2(t.x, t.y), kk[2 * e]), db = __fadd2_rn(make_float2(t.z, t.w), kk[2 * e + 1]);
              d2 = __ffma2_rn(da, da, d2);
              d2 = __ffma2_rn(db, db, d2);
            }
            const float pb = pbv[g];
            // the reference's order (invariant_point_attention.py:181-212): q.k, + pair bias, + point term, + mask term LAST
            out[g] = (((qk.x + qk.y) + pb) + gam * (d2.x + d2.y)) + 1e5f * (mq[g] * mk - 1.f);
          }
        }
        p_store<G>(P + ((size_t)h * Np + j) * G, out);
      }
    }
  }
  __syncthreads();
  // ---- 2. softmax over j; one warp per head
  for (int h = warp; h < 12; h += IA_THREADS / 32) {
    float mx[G], sum[G], t[G];
#pragma unroll
    for (int g = 0; g < G; ++g) { mx[g] = -3.0e38f; sum[g] = 0.f; }
    for (int j = lane; j < Np; j += 32) {
      p_load<G>(P + ((size_t)h * Np + j) * G, t);
#pragma unroll
      for (int g = 0; g < G; ++g) mx[g] = fmaxf(mx[g], t[g]);
    }
#pragma unroll
    for (int g = 0; g < G; ++g) mx[g] = warp_max(mx[g]);
    for (int j = lane; j < Np; j += 32) {
      float* pp = P + ((size_t)h * Np + j) * G;
      p_load<G>(pp, t);
#pragma unroll
      for (int g = 0; g < G; ++g) { t[g] = __expf(t[g] - mx[g]); sum[g] += t[g]; }
      p_store<G>(pp, t);
    }
#pragma unroll
    for (int g = 0; g < G; ++g) sum[g] = 1.f / warp_sum(sum[g]);
    for (int j = lane; j < Np; j += 32) {
      float* pp = P + ((size_t)h * Np + j) * G;
      const uint32_t off = (j >> 6) * 2048u + h * 128u + (((((uint32_t)j & 63) >> 3) ^ ((uint32_t)h & 7u)) << 4) + ((j & 7) << 1);
      p_load<G>(pp, t);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        t[g] *= sum[g];
        float rest = t[g];
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          const bf16 hc = __float2bfloat16_rn(rest);
          *reinterpret_cast<bf16*>(aimg + (c * G + g) * aimg_g + off) = hc;
          rest -= __bfloat162float(hc);
        }
      }
      p_store<G>(pp, t);
    }
  }
  fence_proxy_async();
  __syncthreads();
  // ---- 3. role split
  if (warp == 0) {
    if (lane == 0) {
      const int tot = ng * nch;
      for (int it = IA_ZRING; it < tot; ++it) {
        const int s = it % IA_ZRING, g = it / nch, ch = it - g * nch;
        mbar_wait(empty + s, ((it / IA_ZRING) - 1) & 1);
        const uint8_t* src = reinterpret_cast<const uint8_t*>(z_img) + ((rowb + i0 + g) * 2 * Np + ch * 64) * 128;
        mbar_expect_tx(full + s, ZSLOT);
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          bulk_g2s(zring + s * ZSLOT + c * 16384, src + (size_t)c * z_lo * 2, 8192, full + s);
          bulk_g2s(zring + s * ZSLOT + c * 16384 + 8192, src + (size_t)c * z_lo * 2 + (size_t)Np * 128, 8192, full + s);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(128, 16) | (1u << 15);     // A is MN-major
      int it = 0;
      for (int g = 0; g < ng; ++g) {
        for (int ch = 0; ch < nch; ++ch, ++it) {
          const int s = it % IA_ZRING;
          mbar_wait(full + s, (it / IA_ZRING) & 1);
          tc_fence_after();
          const uint32_t za = smem_u32(zring + s * ZSLOT), ab = smem_u32(aimg + g * aimg_g + ch * 2048);
#pragma unroll
          for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int ps = 0; ps < SplitPasses<NS>::N; ++ps)
              umma(tmem + g * 16, umma_desc_mn(za + SplitPasses<NS>::a(ps) * 16384 + k * 2048, 8192),
                   umma_desc(ab + SplitPasses<NS>::b(ps) * G * aimg_g + k * 32), idesc, (ch | k | ps) ? 1u : 0u);
          umma_commit(empty + s);
        }
        umma_commit(done + g);
      }
    }
    __syncwarp();
  } else {
    // o[h][16] and o_pt[h][8][3] (global frame): 120 float4 groups (10 value planes per head).  One warp
    // per group, lanes over keys (coalesced plane reads, conflict-free P reads), butterfly reduction.
    const float4* vpl = reinterpret_cast<const float4*>(qkv + (size_t)gridDim.y * N * IPA_QW) + (size_t)b * IPA_NPLANE * N;
    for (int t = warp - 2; t < 120; t += IA_THREADS / 32 - 2) {       // one value plane per pass, 8 key loads in flight together
      const int h = t / 10, ev = t - h * 10;
      const float4* src = vpl + (size_t)(h * IPA_PLANES_PER_HEAD + 7 + ev) * N;
      float2 acc[G / 2][4];        // {query 2 gp, query 2 gp + 1} x value component e
#pragma unroll
      for (int gp = 0; gp < G / 2; ++gp) acc[gp][0] = acc[gp][1] = acc[gp][2] = acc[gp][3] = make_float2(0.f, 0.f);
      for (int j0 = 0; j0 < N; j0 += 256) {
        float4 vv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          vv[u] = j < N ? src[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          if (j < N) {
            float av[G];
            p_load<G>(P + ((size_t)h * Np + j) * G, av);
            const float4 v = vv[u];
            const float2 vx = make_float2(v.x, v.x), vy = make_float2(v.y, v.y), vz = make_float2(v.z, v.z), vw = make_float2(v.w, v.w);
#pragma unroll
            for (int gp = 0; gp < G / 2; ++gp) {        // two queries per packed FMA: their probabilities are adjacent in P
              const float2 a2 = make_float2(av[2 * gp], av[2 * gp + 1]);
              acc[gp][0] = __ffma2_rn(a2, vx, acc[gp][0]); acc[gp][1] = __ffma2_rn(a2, vy, acc[gp][1]);
              acc[gp][2] = __ffma2_rn(a2, vz, acc[gp][2]); acc[gp][3] = __ffma2_rn(a2, vw, acc[gp][3]);
            }
          }
        }
      }
      // 4 G per-lane partial sums -> lane v ends up with the warp total of value v = 4 g + e: at every step a lane keeps
      // one half of its values and trades the other half with its partner (4 G - 1 shuffles instead of 5 per value)
      {
        constexpr int NV = 4 * G;
        float a[NV];
#pragma unroll
        for (int gp = 0; gp < G / 2; ++gp)
#pragma unroll
          for (int e = 0; e < 4; ++e) { a[8 * gp + e] = acc[gp][e].x; a[8 * gp + 4 + e] = acc[gp][e].y; }
        // lane distances 16 .. NV first (plain butterfly), then the trading steps NV / 2 .. 1: every value is summed over
        // the same tree (16, 8, 4, 2, 1) whatever G is, so the result does not depend on the queries-per-CTA choice
#pragma unroll
        for (int o = 16; o >= NV; o >>= 1)
#pragma unroll
          for (int i = 0; i < NV; ++i) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
#pragma unroll
        for (int sft = NV / 2; sft >= 1; sft >>= 1) {
          const bool up = (lane & sft) != 0;
#pragma unroll
          for (int i = 0; i < sft; ++i) {
            const float keep = up ? a[i + sft] : a[i], send = up ? a[i] : a[i + sft];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, sft);
          }
        }
        if (lane < NV) oacc[(lane >> 2) * 480 + h * 40 + ev * 4 + (lane & 3)] = a[0];
      }
    }
  }
  __syncthreads();
  // ---- 4. inverse frame + norm (invariant_point_attention.py:227-240)
  for (int e = tid; e < ng * 192; e += IA_THREADS) {
    const int g = e / 192, x = e - g * 192;
    img_store<NS>(cat_img, Rp, (uint32_t)(rowb + i0 + g), x, oacc[g * 480 + (x >> 4) * 40 + (x & 15)], cat_lo);
  }
  for (int e = tid; e < ng * 96; e += IA_THREADS) {
    const int g = e / 96, pt = e - g * 96;
    const float* R = fr + g * 12;
    const float* op = oacc + g * 480 + (pt >> 3) * 40 + 16 + (pt & 7);
    const float x = op[0] - R[9], y = op[8] - R[10], zc = op[16] - R[11];
    const float lx = R[0] * x + R[3] * y + R[6] * zc, ly = R[1] * x + R[4] * y + R[7] * zc, lz = R[2] * x + R[5] * y + R[8] * zc;
    const uint32_t r = (uint32_t)(rowb + i0 + g);
    img_store<NS>(cat_img, Rp, r, 192 + pt, lx, cat_lo);
    img_store<NS>(cat_img, Rp, r, 288 + pt, ly, cat_lo);
    img_store<NS>(cat_img, Rp, r, 384 + pt, lz, cat_lo);
    img_store<NS>(cat_img, Rp, r, 480 + pt, sqrtf(lx * lx + ly * ly + lz * lz + 1e-8f), cat_lo);
  }
  // o_pair: TMEM lane = channel c, columns 16 g + h
  {
    const uint32_t q = warp & 3, c = q * 32 + lane;
    for (int g = (warp >> 2); g < ng; g += IA_THREADS / 128) {
      mbar_wait(done + g, 0);
      tc_fence_after();
      __syncwarp();
      uint32_t v[16];
      tmem_ld16(tmem + ((q * 32u) << 16) + g * 16, v);
      tmem_ld_wait();
      const uint32_t r = (uint32_t)(rowb + i0 + g);
#pragma unroll
      for (int h = 0; h < 12; ++h) img_store<NS>(cat_img, Rp, r, 576 + h * 128 + c, __uint_as_float(v[h]), cat_lo);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TCOLS);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

int tc_struct_init(gb_ctx* c) {
  cudaError_t e;
  if ((e = cudaFuncSetAttribute(gemm_img_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GI<1>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(gemm_img_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GI<2>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_prep_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, PrepSmem<1>::TOTAL)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_prep_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, PrepSmem<2>::TOTAL)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_attention_tc_kernel<8, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ia_smem(8, 256).total)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_attention_tc_kernel<4, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ia_smem(4, 512).total)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_attention_tc_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ia_smem(2, 1024).total)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_attention_tc_kernel<4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ia_smem(4, 256, 2).total)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(ipa_attention_tc_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ia_smem(2, 512, 2).total)) != cudaSuccess)
    return gb_fail(c, GB_ERR_CUDA, "tc_struct_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  return GB_OK;
}

bool tc_struct_supported(const gb_ctx* c, int N) {
  const int cs = c->d.c_s;
  return cs % 128 == 0 && cs <= 512 && N <= 1024 && c->d.n_structure_layer * IPA_H <= 96;
}

// bytes of the block of bf16 images [s | ln1 | h1 | h2 | cat | z]; GB_PREC_BF16X3 keeps the lo component of every image in a
// second block of the same layout behind it
static size_t struct_img_block(const gb_ctx* c, int B, int N) {
  const size_t Np = (size_t)((N + 63) / 64) * 64, Rp = ((size_t)B * N + 127) / 128 * 128;
  const size_t kcs = c->d.c_s / 64;
  return 4 * al256(kcs * Rp * 128) + al256((size_t)(IPA_CAT / 64) * Rp * 128) + al256((size_t)B * N * 2 * Np * 128);
}

size_t tc_struct_workspace_bytes(const gb_ctx* c, int B, int N) {
  if (!tc_struct_supported(c, N)) return 0;
  const size_t Np = (size_t)((N + 63) / 64) * 64;
  return 2 * struct_img_block(c, B, N) + al256((size_t)c->d.n_structure_layer * B * N * IPA_H * Np * 4) + 1024;
}

int tc_struct_bind(gb_ctx* c, void* base, size_t bytes, int B, int N) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  s->Rp = 0;
  if (!tc_struct_supported(c, N)) return GB_OK;
  if (bytes < tc_struct_workspace_bytes(c, B, N) - 1024) return gb_fail(c, GB_ERR_INVALID, "tc_struct_bind: too small");
  const size_t Rp = ((size_t)B * N + 127) / 128 * 128;
  const size_t kcs = c->d.c_s / 64, act = al256(kcs * Rp * 128), cat = al256((size_t)(IPA_CAT / 64) * Rp * 128);
  const size_t blk = struct_img_block(c, B, N);
  char* q = static_cast<char*>(base);
  s->s_img = reinterpret_cast<bf16*>(q);
  s->ln1_img = reinterpret_cast<bf16*>(q + act);
  s->h1_img = reinterpret_cast<bf16*>(q + 2 * act);
  s->h2_img = reinterpret_cast<bf16*>(q + 3 * act);
  s->cat_img = reinterpret_cast<bf16*>(q + 4 * act);
  s->z_img = reinterpret_cast<bf16*>(q + 4 * act + cat);
  s->img_lo = blk / 2;                      // elements from any of the images above to its lo twin
  s->pbias = reinterpret_cast<float*>(q + 2 * blk);
  s->Rp = (int)Rp;
  // padded rows of the activation images and padded keys of z_img feed MMAs: keep them finite (zero)
  if (cudaMemset(q, 0, 2 * blk) != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "tc_struct_bind: memset failed");
  return GB_OK;
}

// rows x K matrix at(r, k) -> two operand images [K/64][rows][64]: hi | lo split components (GB_PREC_BF16X3 reads both, the
// bf16 mode the first)
template <typename F>
static bf16* pack_split2(gb_ctx* c, int rows, int K, F at) {
  std::vector<uint16_t> img((size_t)2 * rows * K);
  pack_image(img.data(), rows, K, at);
  pack_image(img.data() + (size_t)rows * K, rows, K, [&](int r, int k) { return bf_lo(at(r, k)); });
  return (bf16*)upload_bytes(c, img.data(), img.size() * 2);
}
// W [N][K] fp32 (nn.Linear layout)
static bf16* pack_linear(gb_ctx* c, const std::vector<float>& w, int N, int K) {
  return pack_split2(c, N, K, [&](int r, int k) { return w[(size_t)r * K + k]; });
}

int tc_pack_struct_layer(gb_ctx* c, int l, const std::vector<float>& wqkv, const std::vector<float>& bqkv, const std::vector<float>& wout,
                         const std::vector<const std::vector<float>*>& w123) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int cs = c->d.c_s;
  if (cs % 64 != 0) return GB_OK;
  if ((int)s->st.size() <= l) s->st.resize(l + 1);
  TcStructW& w = s->st[l];
  w = TcStructW();
  {
    // output columns reordered: q | q_pts | key-major planes (h, e) of 4 floats (ipa_dims.h)
    w.Wqkv = pack_split2(c, IPA_QKV, cs, [&](int r, int k) { return wqkv[(size_t)ipa_old_col(r) * cs + k]; });
    std::vector<float> bp(IPA_QKV);
    for (int r = 0; r < IPA_QKV; ++r) bp[r] = bqkv[ipa_old_col(r)];
    w.bqkv = (float*)upload_bytes(c, bp.data(), bp.size() * 4);
  }
  w.Wout = pack_linear(c, wout, cs, IPA_CAT);
  for (size_t t = 0; t + 2 < w123.size(); t += 3) {
    w.W1.push_back(pack_linear(c, *w123[t], cs, cs));
    w.W2.push_back(pack_linear(c, *w123[t + 1], cs, cs));
    w.W3.push_back(pack_linear(c, *w123[t + 2], cs, cs));
  }
  if (!w.Wqkv || !w.Wout || !w.bqkv) return gb_fail(c, GB_ERR_CUDA, "tc_pack_struct_layer: device allocation failed");
  return GB_OK;
}

// linear_b of every structure layer stacked into one 128 -> 96 projection, sqrt(1/3) folded in; hi | lo images
int tc_pack_ipa_bias(gb_ctx* c, const std::vector<const std::vector<float>*>& wb, const std::vector<const std::vector<float>*>& bb) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int L = (int)wb.size();
  if (L * IPA_H > 96) return GB_OK;
  const float sc = sqrtf(1.f / 3.f);
  auto at = [&](int r, int k) { return r < L * IPA_H ? sc * (*wb[r / IPA_H])[(size_t)(r % IPA_H) * 128 + k] : 0.f; };
  std::vector<float> b96(96, 0.f);
  for (int r = 0; r < L * IPA_H; ++r) b96[r] = sc * (*bb[r / IPA_H])[r % IPA_H];
  s->Wb_img = pack_split2(c, 96, 128, at);
  s->bias96 = (float*)upload_bytes(c, b96.data(), b96.size() * 4);
  if (!s->Wb_img || !s->bias96) return gb_fail(c, GB_ERR_CUDA, "tc_pack_ipa_bias: device allocation failed");
  return GB_OK;
}

// x3: A / O are images with their lo twins `img_lo` elements behind, W is hi | lo images of K * Nw elements each
static int gemm_img(gb_ctx* c, cudaStream_t st, const bf16* A, int K, const bf16* W, int Nw, const float* bias, const float* res,
                    int ldr, float* C, int ldc, bf16* O, bool relu, int kvsplit = 0, int mblk0 = 0, int nmblk = 0, bool x3 = false,
                    int ksplit = 1, float* C1 = nullptr, float* C2 = nullptr) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  GemmImgArgs g;
  g.mblk0 = mblk0;
  g.A = A; g.W = W; g.bias = bias; g.res = res; g.C = C; g.O = O;
  g.Rp = s->Rp; g.Nw = Nw; g.nkb = K / 64; g.R = c->B * c->N; g.ldc = ldc; g.ldr = ldr; g.relu = relu ? 1 : 0; g.kvsplit = kvsplit;
  g.A_lo = g.O_lo = s->img_lo;
  g.W_lo = (size_t)K * Nw;
  g.ksplit = ksplit; g.C1 = C1; g.C2 = C2;
  if (ksplit > 1 && (ksplit != 3 || !C1 || !C2 || bias || res || O || kvsplit || (K / 64) % 3)) return gb_fail(c, GB_ERR_INVALID, "gemm_img: bad split-K request");
  const dim3 grid(Nw / 64, nmblk ? nmblk : s->Rp / 128, ksplit);
  if (x3) launch_pdl(gemm_img_tc_kernel<2>, grid, 256, GI<2>::SMEM, st, g);
  else launch_pdl(gemm_img_tc_kernel<1>, grid, 256, GI<1>::SMEM, st, g);
  GB_LAUNCH_CHECK(c, "gemm_img_tc");
  return GB_OK;
}

// PairFeatureNet's two projections of s (pair_feature_net.py:117-121): pipj[R][256] = s . [W_i | W_j]^T as one tensor-core
// GEMM with split-bf16 operands (these two terms carry most of the embedding; plain bf16 operands cost 25 % more error in
// p and pushed the L = 1024 z error past its gate).  Also leaves the bf16 operand image of s in s_img, which is what the structure net starts from (s does not
// change through the pair stack), so tc_structure_net skips its own conversion.
int tc_pack_pipj(gb_ctx* c, const std::vector<float>& wij /*[256][c_s]*/) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int cs = c->d.c_s;
  s->Wpipj_img = nullptr;
  if (cs % 64 != 0) return GB_OK;
  // [W_hi | W_hi | W_lo] along K (see split_img_kernel)
  std::vector<uint16_t> img((size_t)256 * 3 * cs);
  pack_image(img.data(), 256, 3 * cs, [&](int r, int k) -> float {
    const float w = wij[(size_t)r * cs + k % cs];
    return k < 2 * cs ? w : bf_lo(w);
  });
  s->Wpipj_img = (bf16*)upload_bytes(c, img.data(), img.size() * 2);
  if (!s->Wpipj_img) return gb_fail(c, GB_ERR_CUDA, "tc_pack_pipj: device allocation failed");
  return GB_OK;
}

bool tc_pipj_supported(const gb_ctx* c) {
  const TcState* s = static_cast<const TcState*>(c->tc_state);
  return s && s->Rp && s->Wpipj_img && s->s_img && (size_t)(s->ln1_img - s->s_img) == (size_t)(c->d.c_s / 64) * s->Rp * 64 &&
         (size_t)(s->h1_img - s->ln1_img) == (size_t)(c->d.c_s / 64) * s->Rp * 64;
}

int tc_pair_pipj(gb_ctx* c, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int R = c->B * c->N, cs = c->d.c_s;
  // s_img | ln1_img | h1_img are consecutive buffers of one image each and the last two are free at this point of the eval
  const size_t stride = (size_t)(s->ln1_img - s->s_img);
  launch_pdl(split_img_kernel, (unsigned)((R * 32 + 255) / 256), 256, 0, st, (const float*)c->s, s->s_img, R, cs, s->Rp, stride);
  GB_LAUNCH_CHECK(c, "s_split_img");
  s->s_img_valid = c->precision == GB_PREC_BF16;      // (the split mode also needs the lo twin of s_img: tc_structure_net makes both)
  return gemm_img(c, st, s->s_img, 3 * cs, s->Wpipj_img, 256, nullptr, nullptr, 0, c->pipj, 256, nullptr, false);
}

static void launch_ln(gb_ctx* c, cudaStream_t st, const float* in, float* out, float* out2, bf16* img, const float* gamma, const float* beta,
                      int normalize, const StructLayerW* bb, bool x3, const float* in2 = nullptr, const float* in3 = nullptr,
                      const float* addb = nullptr, const float* addr = nullptr) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int R = c->B * c->N;
  launch_pdl(ln_img_kernel, (unsigned)((R * 32 + 255) / 256), 256, 0, st, in, out, out2, img, gamma, beta, R, c->d.c_s, s->Rp, normalize,
             bb ? (const float*)bb->Wbb : (const float*)nullptr, bb ? (const float*)bb->bbb : (const float*)nullptr,
             bb ? c->rots_w : (float*)nullptr, bb ? c->trans_w : (float*)nullptr, s->img_lo, x3 ? 2 : 1, in2, in3, addb, addr);
}

// StructureNet.forward (structure_net.py:189-243) on c->s / c->p / c->rots_w / c->trans_w, tensor-core modes.
int tc_structure_net(gb_ctx* c, float* states_out, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const gb_dims& d = c->d;
  const int B = c->B, N = c->N, R = B * N, cs = d.c_s, Np = s->Np, Rp = s->Rp;
  const bool x3 = c->precision == GB_PREC_BF16X3;
  if (!Rp || (int)s->st.size() < d.n_structure_layer || !s->Wb_img) return gb_fail(c, GB_ERR_INVALID, "tc_structure_net: not bound");
  const size_t Rcs = (size_t)R * cs;
  int rc;
  // row-sharded (B == 1): pair bias, z images, attention and the output projection for this rank's query rows only;
  // the [N, c_s] result of every layer is all-gathered and the O(N) rest runs replicated
  const int W = c->shard_world, rows = N / W, row0 = c->shard_rank * rows;
  if (x3 && W > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding is built for the bf16 mode only");
  if (x3 && Np > 512) return gb_fail(c, GB_ERR_UNSUPPORTED, "bf16x3 structure net: N = %d > 512 (the hi | lo probability images do not fit in shared memory); use bf16 or fp32", N);
  TileRange tr = full_tile_range(N, Np);
  if (W > 1) { tr.nt2 = (uint32_t)rows >> 1; tr.t2_0 = (uint32_t)row0 >> 1; }
  const unsigned ptiles = (unsigned)B * tr.nt2 * tr.njb;
  ph_mark(st, "ipa_prep");
  if (x3) launch_pdl_cls(5, ipa_prep_tc_kernel<2>, ptiles, 256, PrepSmem<2>::TOTAL, st, (const float*)c->p, B, N, Np, (const bf16*)s->Wb_img,
                         (const float*)s->bias96, s->pbias, s->z_img, d.n_structure_layer * IPA_H, tr, s->img_lo);
  else launch_pdl_cls(5, ipa_prep_tc_kernel<1>, ptiles, 256, PrepSmem<1>::TOTAL, st, (const float*)c->p, B, N, Np, (const bf16*)s->Wb_img,
                      (const float*)s->bias96, s->pbias, s->z_img, d.n_structure_layer * IPA_H, tr, (size_t)0);
  GB_LAUNCH_CHECK(c, "ipa_prep_tc");
  if (!s->s_img_valid) {      // already made by tc_pair_pipj of this eval (bf16 mode)
    launch_ln(c, st, c->s, nullptr, nullptr, s->s_img, nullptr, nullptr, 0, nullptr, x3);
    GB_LAUNCH_CHECK(c, "s_to_img");
  }
  s->s_img_valid = false;
  int state_idx = 1;
  for (int blk = 0; blk < d.n_structure_block; ++blk)
    for (int l = 0; l < d.n_structure_layer; ++l) {
      const StructLayerW& w = c->st[l];
      const TcStructW& t = s->st[l];
      ph_mark(st, "qkv + points");
      if ((rc = gemm_img(c, st, s->s_img, cs, t.Wqkv, IPA_QKV, t.bqkv, nullptr, 0, c->qkv, IPA_QKV, nullptr, false, N, 0, 0, x3))) return rc;
      launch_pdl(ipa_points_planes_kernel, (R * IPA_H + 255) / 256, 256, 0, st, c->qkv, c->rots_w, c->trans_w, R, N);
      GB_LAUNCH_CHECK(c, "ipa_points_planes");
      ph_mark(st, "ipa_attention");
      {
        const int nq = W > 1 ? rows : N, G = ia_queries_per_cta(Np, B, nq, s->num_sms, x3 ? 2 : 1);
        const dim3 grid((nq + G - 1) / G, B);
        const float* pb = s->pbias + (size_t)l * B * N * IPA_H * Np;
        const size_t smem = ia_smem(G, Np, x3 ? 2 : 1).total, zl = x3 ? s->img_lo : 0;
#define IA_LAUNCH(GG, SP) launch_pdl(ipa_attention_tc_kernel<GG, SP>, grid, IA_THREADS, smem, st, (const float*)c->qkv, pb, (const bf16*)s->z_img, \
                                     (const float*)c->rots_w, (const float*)c->trans_w, (const float*)w.hw, s->cat_img, N, Np, Rp, row0, (const int*)c->mask, zl, zl)
        if (x3) { if (G == 4) IA_LAUNCH(4, 2); else IA_LAUNCH(2, 2); }
        else if (G == 8) IA_LAUNCH(8, 1);
        else if (G == 4) IA_LAUNCH(4, 1);
        else IA_LAUNCH(2, 1);
#undef IA_LAUNCH
      }
      GB_LAUNCH_CHECK(c, "ipa_attention_tc");
      ph_mark(st, "out_proj");
      // output projection, K = 2112 = 33 k-blocks: split three ways over grid.z (144 CTAs instead of 48, a third of the serial
      // MMA chain each; also a third of the accumulation length inside the tensor core); the partial sums land in sA / h1 / h2
      // and the LayerNorm kernel adds them, the bias and the residual in a fixed order
      const bool split3 = (IPA_CAT / 64) % 3 == 0, ks3 = W == 1 && split3;
      float* ln_in = c->sA;
      if (W > 1) {
        // this rank's query rows only, same split and same order of additions as on one GPU; then every rank pulls the other
        // ranks' rows out of their memory (peer-mapped, shard.cu).  Two exchange buffers alternate from layer to layer: a
        // rank rewrites its rows two layers later, after a barrier that every rank reaches only once its pull has finished.
        float* X = c->sX[(blk * d.n_structure_layer + l) & 1];
        ln_in = X;
        if (split3) {
          if ((rc = gemm_img(c, st, s->cat_img, IPA_CAT, t.Wout, cs, nullptr, nullptr, 0, X, cs, nullptr, false, 0, row0 / 128, rows / 128, false, 3, c->h1, c->h2))) return rc;
          launch_pdl(sum_partials_kernel, (unsigned)((rows * (cs / 4) + 255) / 256), 256, 0, st, X, (const float*)c->h1, (const float*)c->h2,
                     (const float*)w.bout, (const float*)c->s, row0, rows, cs);
          GB_LAUNCH_CHECK(c, "sum_partials");
        } else if ((rc = gemm_img(c, st, s->cat_img, IPA_CAT, t.Wout, cs, w.bout, c->s, cs, X, cs, nullptr, false, 0, row0 / 128, rows / 128))) return rc;
        ph_mark(st, "barrier");
        if ((rc = shard_barrier(c, st))) return rc;
        ph_mark(st, "pull rows");
        if ((rc = shard_pull_gather(c, X, (size_t)rows * cs * sizeof(float), st))) return rc;
        ph_mark(st, "out_proj");
      }
      if (W == 1 && !ks3 && (rc = gemm_img(c, st, s->cat_img, IPA_CAT, t.Wout, cs, w.bout, c->s, cs, c->sA, cs, nullptr, false, 0, 0, 0, x3))) return rc;
      if (ks3) {
        if ((rc = gemm_img(c, st, s->cat_img, IPA_CAT, t.Wout, cs, nullptr, nullptr, 0, c->sA, cs, nullptr, false, 0, 0, 0, x3, 3, c->h1, c->h2))) return rc;
        launch_ln(c, st, c->sA, c->sB, nullptr, s->ln1_img, w.ln1_g, w.ln1_b, 1, nullptr, x3, c->h1, c->h2, w.bout, c->s);
      } else
      launch_ln(c, st, ln_in, c->sB, nullptr, s->ln1_img, w.ln1_g, w.ln1_b, 1, nullptr, x3);
      GB_LAUNCH_CHECK(c, "ipa_layer_norm");
      ph_mark(st, "transition + bb");
      float* cur = c->sB;
      float* other = c->sA;
      for (int tl = 0; tl < d.n_structure_transition_layer; ++tl) {
        if (tl > 0) {
          launch_ln(c, st, cur, nullptr, nullptr, s->ln1_img, nullptr, nullptr, 0, nullptr, x3);
          GB_LAUNCH_CHECK(c, "to_img");
        }
        if ((rc = gemm_img(c, st, s->ln1_img, cs, t.W1[tl], cs, w.b1[tl], nullptr, 0, nullptr, 0, s->h1_img, true, 0, 0, 0, x3))) return rc;
        if ((rc = gemm_img(c, st, s->h1_img, cs, t.W2[tl], cs, w.b2[tl], nullptr, 0, nullptr, 0, s->h2_img, true, 0, 0, 0, x3))) return rc;
        if ((rc = gemm_img(c, st, s->h2_img, cs, t.W3[tl], cs, w.b3[tl], cur, cs, other, cs, nullptr, false, 0, 0, 0, x3))) return rc;
        float* tmp = cur; cur = other; other = tmp;
      }
      float* st_out = states_out ? states_out + (size_t)state_idx * Rcs : nullptr;
      launch_ln(c, st, cur, c->s, st_out, s->s_img, w.ln2_g, w.ln2_b, 1, &w, x3);
      GB_LAUNCH_CHECK(c, "transition_layer_norm_backbone_update");
      ++state_idx;
    }
  return GB_OK;
}
