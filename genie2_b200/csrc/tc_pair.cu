// tcgen05 / TMEM implementation of the pair stack (GB_PREC_BF16): bf16 operands, fp32
// accumulation in tensor memory, fp32 LayerNorm statistics, fp32 residual stream p.
//   K4 trimul_proj_tc      (tc_proj.cu) LN_in -> gated / masked a, b (bf16 tile images, the operand
//                          format of K5) and the output gate g (bf16)
//   K5 trimul_contract_tc  per (sample, channel): X = A . B^T, operands streamed with cp.async.bulk
//   K6 trimul_out_tc       LN_out over channels -> linear_z -> * g -> p +=
//   K7 pair_transition_tc  (tc_transition.cu) LN -> 128->512 -> ReLU -> 512->128, hidden never leaves the SM
// This file: shared helpers, K5, K6, workspace and dispatch.
// Reference: modules/triangular_multiplicative_update.py:57-110, modules/pair_transition.py:48-87.
#include <string.h>

#include <vector>

#include "common.cuh"
#include "shard.h"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;



// ---------------------------------------------------------------------------------------------
// shared helpers
// ---------------------------------------------------------------------------------------------
// LayerNorm(128) of up to 128 pixel rows -> bf16 A operand image [2 kb][128 rows][64].
// Warp w normalises rows 16w..16w+15, eight at a time (eight 512-byte coalesced loads in flight);
// lane l holds channels 4l..4l+3.  src_row(m) returns the fp32 row pointer or nullptr (zero row).
template <typename RowFn>
__device__ __forceinline__ void ln_rows_to_image(uint8_t* A, RowFn src_row, const float* __restrict__ gamma,
                                                 const float* __restrict__ beta) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float4 g4 = ld4(gamma + lane * 4), b4 = ld4(beta + lane * 4);
  const uint32_t kb = lane >> 4, chunk = (lane & 15) >> 1, half = lane & 1;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const float* r = src_row(warp * 16 + it * 8 + u);
      v[u] = r ? ld4(r + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint32_t m = warp * 16 + it * 8 + u;
      const float mean = warp_sum(v[u].x + v[u].y + v[u].z + v[u].w) * (1.f / 128.f);
      const float dx = v[u].x - mean, dy = v[u].y - mean, dz = v[u].z - mean, dw = v[u].w - mean;
      const float rstd = rsqrtf(warp_sum(dx * dx + dy * dy + dz * dz + dw * dw) * (1.f / 128.f) + 1e-5f);
      uint2 o;
      o.x = pack_bf16(dx * rstd * g4.x + b4.x, dy * rstd * g4.y + b4.y);
      o.y = pack_bf16(dz * rstd * g4.z + b4.z, dw * rstd * g4.w + b4.w);
      *reinterpret_cast<uint2*>(A + kb * 16384u + m * 128u + ((chunk ^ (m & 7u)) << 4) + half * 8u) = o;
    }
  }
}

// Same for the eight rows m0 .. m0 + 7 of this warp only (the caller chooses the row range).
template <typename RowFn>
__device__ __forceinline__ void ln_rows8_to_image(uint8_t* A, uint32_t m0, RowFn src_row, const float* __restrict__ gamma,
                                                  const float* __restrict__ beta) {
  const int lane = threadIdx.x & 31;
  const float4 g4 = ld4(gamma + lane * 4), b4 = ld4(beta + lane * 4);
  const uint32_t kb = lane >> 4, chunk = (lane & 15) >> 1, half = lane & 1;
  float4 v[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const float* r = src_row(m0 + u);
    v[u] = r ? ld4(r + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const uint32_t m = m0 + u;
    const float mean = warp_sum(v[u].x + v[u].y + v[u].z + v[u].w) * (1.f / 128.f);
    const float dx = v[u].x - mean, dy = v[u].y - mean, dz = v[u].z - mean, dw = v[u].w - mean;
    const float rstd = rsqrtf(warp_sum(dx * dx + dy * dy + dz * dz + dw * dw) * (1.f / 128.f) + 1e-5f);
    uint2 o;
    o.x = pack_bf16(dx * rstd * g4.x + b4.x, dy * rstd * g4.y + b4.y);
    o.y = pack_bf16(dz * rstd * g4.z + b4.z, dw * rstd * g4.w + b4.w);
    *reinterpret_cast<uint2*>(A + kb * 16384u + m * 128u + ((chunk ^ (m & 7u)) << 4) + half * 8u) = o;
  }
}

// fp32 staging tile [128 rows][128 cols] with 16-byte chunks XOR-swizzled by the row, so that a
// thread-per-row writer (TMEM epilogue) and a warp-per-row reader are both bank-conflict free.
__device__ __forceinline__ float4* stage_ptr(uint8_t* S, uint32_t m, uint32_t chunk) {
  return reinterpret_cast<float4*>(S + m * 512u + ((chunk ^ (m & 31u)) << 4));
}
// this thread's 32 TMEM columns [col0, col0+32) (+ bias) -> staging row m
__device__ __forceinline__ void tmem_to_stage(uint8_t* S, uint32_t taddr, uint32_t m, uint32_t col0,
                                              const float* __restrict__ bias) {
  uint32_t v[32];
  tmem_ld32(taddr, v);
  tmem_ld_wait();
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    float4 o;
    o.x = __uint_as_float(v[4 * q + 0]) + bias[col0 + 4 * q + 0];
    o.y = __uint_as_float(v[4 * q + 1]) + bias[col0 + 4 * q + 1];
    o.z = __uint_as_float(v[4 * q + 2]) + bias[col0 + 4 * q + 2];
    o.w = __uint_as_float(v[4 * q + 3]) + bias[col0 + 4 * q + 3];
    *stage_ptr(S, m, (col0 >> 2) + q) = o;
  }
}

// K7 (pair transition) lives in tc_transition.cu.
// ---------------------------------------------------------------------------------------------
// Pixel tiles of the triangle kernels: a CTA owns 128 pixels = 2 consecutive rows x 64 consecutive
// columns of one sample's N x N pair map; tile-local pixel m = 64 * ii + jj.
//   tile -> (b, t2, jb);  row = 2 * t2 + ii,  col = 64 * jb + jj
// K4 reads pixel (row, col) for the outgoing update and pixel (col, row) for the incoming one, so
// that in both cases it emits A[row][k = col] (and B likewise) K-major and K5 is a plain A . B^T:
//   outgoing x_ij = sum_k a_ik b_jk ;  incoming x_ij = sum_k a_ki b_kj = sum_k aT_ik bT_jk.
// ---------------------------------------------------------------------------------------------
// K4 (gated projections) lives in tc_proj.cu.
// ---------------------------------------------------------------------------------------------
// K5 trimul_contract_tc (triangular_multiplicative_update.py:57-82 with the operand orientation
// fixed by K4): per (sample b, channel c): X[i][j] = sum_k A[i][k] B[j][k].
// One CTA per (b, c, 128-row block, column block of NJ <= 256): warp 0 streams 64-wide k-blocks
// of A (128 rows, 16 KB) and B (NJ rows) into a ring with cp.async.bulk, warp 1 issues the MMAs
// into a 128 x NJ fp32 TMEM accumulator, all 8 warps run the epilogue (bf16 x, channel-major).
// ---------------------------------------------------------------------------------------------
#define K5_STAGES 2
#define K5_STAGE_BYTES (16384 + 32768)
#define K5_SMEM (K5_STAGES * K5_STAGE_BYTES + 1024 + 128)

__global__ void __launch_bounds__(256, 2) trimul_contract_tc_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b,
                                                                    bf16* __restrict__ x, int N, int Np, int NJ,
                                                                    const ImgLayout La, const ImgLayout Lb, int ib0, int nib,
                                                                    int ch0, int nchs, const XLayout Lx) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + K5_STAGES * K5_STAGE_BYTES);
  uint64_t* empty = full + K5_STAGES;
  uint64_t* done = empty + K5_STAGES;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(done + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nkb = Np >> 6, njb = (Np + NJ - 1) / NJ;
  uint32_t blk = blockIdx.x;
  const uint32_t jb = blk % njb; blk /= njb;
  const uint32_t ib = ib0 + blk % nib; blk /= nib;  // blk = b * nchs + c; a sharded launch owns channels [ch0, ch0 + nchs)
  const uint32_t nj = min((uint32_t)NJ, (uint32_t)Np - jb * NJ);   // columns of this block (multiple of 64)
  const uint32_t bb = blk / (uint32_t)nchs, cch = (uint32_t)ch0 + blk % (uint32_t)nchs;
  const uint32_t bpiece = (Lb.RL % nj == 0) ? nj : 128u;            // B rows that are contiguous in the (possibly row-split) image
  const uint32_t a_rows = min(128u, (uint32_t)Np - ib * 128u);      // rows present in the plane (64 or 128)

  if (tid == 0) {
    for (int s = 0; s < K5_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tslot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  pdl_wait();
  pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % K5_STAGES;
        if (kb >= K5_STAGES) mbar_wait(empty + s, ((kb / K5_STAGES) - 1) & 1);
        uint8_t* st = sm + s * K5_STAGE_BYTES;
        mbar_expect_tx(full + s, (a_rows + nj) * 128u);
        bulk_g2s(st, a + img_off(La, bb, cch, kb, ib * 128u), a_rows * 128u, full + s);
        for (uint32_t r = 0; r < nj; r += bpiece)
          bulk_g2s(st + 16384 + r * 128u, b + img_off(Lb, bb, cch, kb, jb * NJ + r), min(bpiece, nj - r) * 128u, full + s);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc(128, nj), tm = warp_uniform(tmem);
    for (int kb = 0; kb < nkb; ++kb) {
      const int s = kb % K5_STAGES;
      mbar_wait(full + s, (kb / K5_STAGES) & 1);
      tc_fence_after();
      const uint32_t sa = smem_u32(sm + s * K5_STAGE_BYTES), sb = sa + 16384;
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma(tm, umma_desc(sa + k * 32), umma_desc(sb + k * 32), idesc, (kb | k) ? 1u : 0u);
        umma_commit(empty + s);
        if (kb == nkb - 1) umma_commit(done);
      }
      __syncwarp();
    }
  }
  mbar_wait(done, 0);
  tc_fence_after();
  __syncwarp();
  // epilogue: thread = output row i (TMEM lane); warps 0-3 take even 32-column groups, 4-7 odd ones
  const uint32_t q = warp & 3, h = warp >> 2, i = ib * 128 + q * 32 + lane;
  const uint32_t t_lane = tmem + ((q * 32u) << 16);
  // x tile images: [b][t2 = i / 2][jb64][ii = i & 1][channel][64 j], 16-byte chunks swizzled by channel & 7:
  // the 32 KB block of one (t2, jb64) pixel tile is exactly K6's MN-major A operand.
  for (uint32_t c0 = h * 32; c0 < nj; c0 += 64) {
    uint32_t v[32];
    tmem_ld32(t_lane + c0, v);
    tmem_ld_wait();
    if (i < (uint32_t)N) {
      const uint32_t j = jb * NJ + c0, jb64 = j >> 6, q0 = (j & 63) >> 3;
      uint8_t* dst = reinterpret_cast<uint8_t*>(x + x_off(Lx, bb, i, jb64, cch));
      const uint32_t sw = cch & 7u;
#pragma unroll
      for (int kp = 0; kp < 2; ++kp) {     // two 16-byte chunks = one 32-byte sector per store
        uint4 o[2];
#pragma unroll
        for (int k4 = 0; k4 < 2; ++k4) {
          const uint32_t* vv = v + 16 * kp + 8 * k4;
          o[k4].x = pack_bf16(__uint_as_float(vv[0]), __uint_as_float(vv[1]));
          o[k4].y = pack_bf16(__uint_as_float(vv[2]), __uint_as_float(vv[3]));
          o[k4].z = pack_bf16(__uint_as_float(vv[4]), __uint_as_float(vv[5]));
          o[k4].w = pack_bf16(__uint_as_float(vv[6]), __uint_as_float(vv[7]));
        }
        const bool swap = sw & 1u;
        st_global_32B(dst + (((q0 + 2 * kp) ^ (sw & 6u)) << 4), swap ? o[1] : o[0], swap ? o[0] : o[1]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 256);
}

// ---------------------------------------------------------------------------------------------
// K5, persistent form: one CTA per SM walks the (b, channel, row block, column block) units round robin.  Warp 0 streams
// k-blocks through a 4-stage ring that runs ahead across unit boundaries, warp 1 issues the MMAs into one of TWO 256-column
// TMEM accumulators, warps 2-9 drain the other one -- load, MMA and drain of consecutive units overlap instead of
// relying on a second resident CTA.  Same operands, same k order, same stores as the kernel above: bit-identical x.
// ---------------------------------------------------------------------------------------------
// SPLIT (GB_PREC_BF16X3): a, b and x are hi | lo image pairs (lo plane `lo_ab` / `lo_x` elements behind the hi plane).  A
// stage holds [A_hi 16K | B_hi 32K | A_lo 16K | B_lo 32K] (two stages instead of four) and every k step issues the three
// passes A_hi B_hi + A_lo B_hi + A_hi B_lo into the same fp32 accumulator; x is written as bf16(x) and bf16(x - bf16(x)).
#define K5P_THREADS 320
template <bool SPLIT> struct K5P {
  static constexpr uint32_t STAGES = SPLIT ? 2 : 4;
  static constexpr uint32_t STAGE_BYTES = SPLIT ? 2 * K5_STAGE_BYTES : K5_STAGE_BYTES;
  static constexpr uint32_t SMEM = STAGES * STAGE_BYTES + 1024 + 128;
};

template <bool SPLIT>
__global__ void __launch_bounds__(K5P_THREADS, 1) trimul_contract_pers_tc_kernel(
    const bf16* __restrict__ a, const bf16* __restrict__ b, bf16* __restrict__ x, int N, int Np, int NJ, const ImgLayout La,
    const ImgLayout Lb, int ib0, int nib, uint32_t nunits, size_t lo_ab, size_t lo_x, int ch0, int nchs, const XLayout Lx) {
  constexpr uint32_t K5P_STAGES = K5P<SPLIT>::STAGES, STAGE_BYTES = K5P<SPLIT>::STAGE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + K5P_STAGES * STAGE_BYTES);
  uint64_t* empty = full + K5P_STAGES;
  uint64_t* acc_full = empty + K5P_STAGES;      // [2] MMAs of a unit complete
  uint64_t* acc_empty = acc_full + 2;           // [2] the 8 epilogue warps have read the accumulator
  uint32_t* tslot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t nkb = (uint32_t)Np >> 6, njb = (uint32_t)(Np + NJ - 1) / NJ;

  if (tid == 0) {
    for (int s = 0; s < K5P_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(acc_full + s, 1); mbar_init(acc_empty + s, 8); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  pdl_wait();
  pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (uint32_t u = blockIdx.x; u < nunits; u += gridDim.x) {
        uint32_t blk = u;
        const uint32_t jb = blk % njb; blk /= njb;
        const uint32_t ib = ib0 + blk % nib; blk /= nib;
        const uint32_t nj = min((uint32_t)NJ, (uint32_t)Np - jb * NJ);
        const uint32_t bb = blk / (uint32_t)nchs, cch = (uint32_t)ch0 + blk % (uint32_t)nchs;
        const uint32_t bpiece = (Lb.RL % nj == 0) ? nj : 128u;
        const uint32_t a_rows = min(128u, (uint32_t)Np - ib * 128u);
        for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
          const uint32_t s = it % K5P_STAGES;
          if (it >= K5P_STAGES) mbar_wait(empty + s, ((it / K5P_STAGES) - 1) & 1);
          uint8_t* st = sm + s * STAGE_BYTES;
          mbar_expect_tx(full + s, (SPLIT ? 2u : 1u) * (a_rows + nj) * 128u);
          const size_t ao = img_off(La, bb, cch, kb, ib * 128u);
          bulk_g2s(st, a + ao, a_rows * 128u, full + s);
          if (SPLIT) bulk_g2s(st + K5_STAGE_BYTES, a + lo_ab + ao, a_rows * 128u, full + s);
          for (uint32_t r = 0; r < nj; r += bpiece) {
            const size_t bo = img_off(Lb, bb, cch, kb, jb * NJ + r);
            bulk_g2s(st + 16384 + r * 128u, b + bo, min(bpiece, nj - r) * 128u, full + s);
            if (SPLIT) bulk_g2s(st + K5_STAGE_BYTES + 16384 + r * 128u, b + lo_ab + bo, min(bpiece, nj - r) * 128u, full + s);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    const uint32_t tm = warp_uniform(tmem);
    uint32_t it = 0, n = 0;
    for (uint32_t u = blockIdx.x; u < nunits; u += gridDim.x, ++n) {
      const uint32_t jb = u % njb;
      const uint32_t nj = min((uint32_t)NJ, (uint32_t)Np - jb * NJ);
      const uint32_t idesc = umma_idesc(128, nj), slot = n & 1u, acc = tm + slot * 256u;
      if (n >= 2) {
        mbar_wait(acc_empty + slot, ((n >> 1) - 1) & 1);
        tc_fence_after();
      }
      for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
        const uint32_t s = it % K5P_STAGES;
        mbar_wait(full + s, (it / K5P_STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(sm + s * STAGE_BYTES), sb = sa + 16384;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            umma(acc, umma_desc(sa + k * 32), umma_desc(sb + k * 32), idesc, (kb | k) ? 1u : 0u);
            if (SPLIT) {
              umma(acc, umma_desc(sa + K5_STAGE_BYTES + k * 32), umma_desc(sb + k * 32), idesc, 1u);
              umma(acc, umma_desc(sa + k * 32), umma_desc(sb + K5_STAGE_BYTES + k * 32), idesc, 1u);
            }
          }
          umma_commit(empty + s);
          if (kb == nkb - 1) umma_commit(acc_full + slot);
        }
        __syncwarp();
      }
    }
  } else {
    // epilogue warps 2..9: thread = output row (TMEM lane quarter = warp & 3); warps 2-5 take even 32-column groups, 6-9 odd ones
    const uint32_t q = warp & 3, h = (uint32_t)(warp - 2) >> 2;
    uint32_t n = 0;
    for (uint32_t u = blockIdx.x; u < nunits; u += gridDim.x, ++n) {
      uint32_t blk = u;
      const uint32_t jb = blk % njb; blk /= njb;
      const uint32_t ib = ib0 + blk % nib; blk /= nib;
      const uint32_t nj = min((uint32_t)NJ, (uint32_t)Np - jb * NJ);
      const uint32_t bb = blk / (uint32_t)nchs, cch = (uint32_t)ch0 + blk % (uint32_t)nchs, slot = n & 1u;
      const uint32_t i = ib * 128 + q * 32 + lane;
      const uint32_t t_lane = tmem + ((q * 32u) << 16) + slot * 256u;
      mbar_wait(acc_full + slot, (n >> 1) & 1);
      tc_fence_after();
      __syncwarp();
      for (uint32_t c0 = h * 32; c0 < nj; c0 += 64) {
        uint32_t v[32];
        tmem_ld32(t_lane + c0, v);
        tmem_ld_wait();
        if (i < (uint32_t)N) {
          const uint32_t j = jb * NJ + c0, jb64 = j >> 6, q0 = (j & 63) >> 3;
          uint8_t* dst = reinterpret_cast<uint8_t*>(x + x_off(Lx, bb, i, jb64, cch));
          const uint32_t sw = cch & 7u;
#pragma unroll
          for (int kp = 0; kp < 2; ++kp) {
            uint4 o[2], ol[2];
#pragma unroll
            for (int k4 = 0; k4 < 2; ++k4) {
              const uint32_t* vv = v + 16 * kp + 8 * k4;
              if (SPLIT) {
                split_bf16x2(__uint_as_float(vv[0]), __uint_as_float(vv[1]), o[k4].x, ol[k4].x);
                split_bf16x2(__uint_as_float(vv[2]), __uint_as_float(vv[3]), o[k4].y, ol[k4].y);
                split_bf16x2(__uint_as_float(vv[4]), __uint_as_float(vv[5]), o[k4].z, ol[k4].z);
                split_bf16x2(__uint_as_float(vv[6]), __uint_as_float(vv[7]), o[k4].w, ol[k4].w);
              } else {
                o[k4].x = pack_bf16(__uint_as_float(vv[0]), __uint_as_float(vv[1]));
                o[k4].y = pack_bf16(__uint_as_float(vv[2]), __uint_as_float(vv[3]));
                o[k4].z = pack_bf16(__uint_as_float(vv[4]), __uint_as_float(vv[5]));
                o[k4].w = pack_bf16(__uint_as_float(vv[6]), __uint_as_float(vv[7]));
              }
            }
            const bool swap = sw & 1u;
            uint8_t* d32 = dst + (((q0 + 2 * kp) ^ (sw & 6u)) << 4);
            st_global_32B(d32, swap ? o[1] : o[0], swap ? o[0] : o[1]);
            if (SPLIT) st_global_32B(d32 + lo_x * 2, swap ? ol[1] : ol[0], swap ? ol[0] : ol[1]);
          }
        }
      }
      // every lane's TMEM reads have completed (tmem_ld_wait): hand the accumulator back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + slot);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
// K6 trimul_out_tc (triangular_multiplicative_update.py:105-108 + residual, pair_transform_net.py:111-112)
//   p[pixel] += (W_z LN_out(x[:, pixel]) + b_z) * g[pixel]
// The LayerNorm over channels is applied AFTER the GEMM:  with W' = W_z * gamma (bf16),
//   u[n] = sum_c W'[n][c],  v[n] = sum_c W_z[n][c] beta[c] + b_z[n]:
//   W_z LN(x) + b_z = rstd * (W' x - mean * u) + v
// so the raw bf16 x tile is the MMA operand as it lies in memory (one 32 KB cp.async.bulk, MN-major A:
// pixels contiguous, channels along K) and no thread touches x except to take its per-pixel mean / rstd
// from shared memory while the tensor core runs.  u is summed from the bf16-rounded W' so the mean
// term cancels exactly.  Shared: X 32K | W' 32K (then the 64K fp32 staging tile); 128 TMEM columns.
// ---------------------------------------------------------------------------------------------
#define K6_SMEM (65536 + 1024 + 64 + 5120)

__global__ void __launch_bounds__(256, 3) trimul_out_tc_kernel(
    const PeerImg px, const bf16* __restrict__ g, float* __restrict__ p, int N, int Np,
    const bf16* __restrict__ Wzimg, const float* __restrict__ uvec, const float* __restrict__ vvec, const TileRange tr, const XLayout Lx) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* X = sm;
  uint8_t* W = sm + 32768;
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(sm + 65536);
  uint64_t* bar_mma = bar_w + 1;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_w + 2);
  float* red = reinterpret_cast<float*>(sm + 65536 + 64);     // [8][128] partials, then mean [128], rstd [128]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const PixTile T = pix_tile(blockIdx.x, tr);
  if (tid == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
    mbar_expect_tx(bar_w, 65536);
    bulk_g2s(W, Wzimg, 32768, bar_w);
  }
  if (warp == 1) tmem_alloc(tslot, 128);
  {
    // the tile's p and g rows are only read at the very end (after the GEMM, the statistics and the TMEM drain), with few
    // bytes in flight per warp: ask L2 for them now.  Safe before the dependency wait: L2 is the point of coherence.
    if (tid >= 32 && tid < 36) {
      const uint32_t ii = (tid >> 1) & 1, row = 2 * T.t2 + ii, col = 64 * T.jb;
      if (row < (uint32_t)N && col < (uint32_t)N) {
        const uint32_t npx = min(64u, (uint32_t)N - col);
        const size_t px = ((size_t)T.b * N + row) * N + col;
        if (tid & 1) l2_prefetch(g + px * 128, npx * 256u);
        else l2_prefetch(p + px * 128, npx * 512u);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  pdl_wait();            // x comes from the contraction kernel
  pdl_launch();
  if (tid == 0) {
    // one 32 KB tile; channel-sharded runs fetch it as W x 2 pieces [i & 1][CG channels][64], each from the rank that
    // contracted that channel group (peer memory over NVLink)
    if (Lx.W == 1) bulk_g2s(X, px.p[0] + x_off(Lx, T.b, 2 * T.t2, T.jb, 0), 32768, bar_w);
    else {
      // the W - 1 remote pieces first, starting on another owner per rank and per tile (the requests of all ranks spread over
      // all memories instead of arriving at rank 0 first), the local piece last
      const uint32_t me = (2u * T.t2) / Lx.RL;
      for (uint32_t k = 0; k + 1u < Lx.W; ++k) {
        const uint32_t cg = (me + 1u + (k + blockIdx.x) % (Lx.W - 1u)) % Lx.W;
        for (uint32_t ii = 0; ii < 2; ++ii)
          bulk_g2s(X + ii * 16384u + cg * Lx.CG * 128u, peer_ptr(px, cg) + x_off(Lx, T.b, 2 * T.t2 + ii, T.jb, cg * Lx.CG), Lx.CG * 128u, bar_w);
      }
      for (uint32_t ii = 0; ii < 2; ++ii)
        bulk_g2s(X + ii * 16384u + me * Lx.CG * 128u, peer_ptr(px, me) + x_off(Lx, T.b, 2 * T.t2 + ii, T.jb, me * Lx.CG), Lx.CG * 128u, bar_w);
    }
  }
  if (tid == 0) {
    mbar_wait(bar_w, 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc(128, 128) | (1u << 15);      // A (x tile) is MN-major
#pragma unroll
    for (int k = 0; k < 8; ++k)
      umma(tmem, umma_desc_mn(smem_u32(X) + k * 2048, 16384), umma_desc(smem_u32(W) + (k >> 2) * 16384 + (k & 3) * 32), idesc, k ? 1u : 0u);
    umma_commit(bar_mma);
  }
  __syncwarp();
  // per-pixel statistics over the 128 channels, from the same shared tile (two passes, fp32).  A thread owns one 16-byte
  // chunk column (8 consecutive pixels) of 8 channel rows: 8 conflict-free LDS.128 bring its 64 values into registers;
  // partial sums meet through a shuffle (lane ^ 16 = the neighbouring channel group) and an [8][128] shared array.
  mbar_wait(bar_w, 0);
  {
    const uint32_t pc = tid & 15, cg = tid >> 4;                  // pixel chunk (8 pixels), channel group (8 channels)
    const uint32_t q8 = pc & 7;
    const uint8_t* base = X + (pc >> 3) * 16384u + cg * 1024u;    // channel row c = 8 cg + k holds chunk q8 at position q8 ^ k
    uint4 w[8];
#pragma unroll
    for (uint32_t k = 0; k < 8; ++k) w[k] = *reinterpret_cast<const uint4*>(base + k * 128u + ((q8 ^ k) << 4));
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t ww[4] = {w[k].x, w[k].y, w[k].z, w[k].w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        acc[2 * t] += __uint_as_float(ww[t] << 16);
        acc[2 * t + 1] += __uint_as_float(ww[t] & 0xffff0000u);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
    float* part = red + (cg >> 1) * 128 + pc * 8;                 // red: [8][128] partials | mean [128] | rstd [128]
    if ((lane & 16) == 0) {
      st4(part, make_float4(acc[0], acc[1], acc[2], acc[3]));
      st4(part + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
    }
    __syncthreads();
    if (tid < 128) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) t += red[r * 128 + tid];
      red[1024 + tid] = t * (1.f / 128.f);
    }
    __syncthreads();
    float mean[8];
    {
      const float4 m0 = ld4(red + 1024 + pc * 8), m1 = ld4(red + 1024 + pc * 8 + 4);
      mean[0] = m0.x; mean[1] = m0.y; mean[2] = m0.z; mean[3] = m0.w; mean[4] = m1.x; mean[5] = m1.y; mean[6] = m1.z; mean[7] = m1.w;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t ww[4] = {w[k].x, w[k].y, w[k].z, w[k].w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float d0 = __uint_as_float(ww[t] << 16) - mean[2 * t], d1 = __uint_as_float(ww[t] & 0xffff0000u) - mean[2 * t + 1];
        acc[2 * t] = fmaf(d0, d0, acc[2 * t]);
        acc[2 * t + 1] = fmaf(d1, d1, acc[2 * t + 1]);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
    if ((lane & 16) == 0) {                                       // the partials of pass 1 were consumed before the last barrier
      st4(part, make_float4(acc[0], acc[1], acc[2], acc[3]));
      st4(part + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
    }
    __syncthreads();
    if (tid < 128) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) t += red[r * 128 + tid];
      red[1152 + tid] = rsqrtf(t * (1.f / 128.f) + 1e-5f);
    }
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  __syncthreads();                                               // stats visible; X / W dead -> staging
  const uint32_t q = warp & 3, h = warp >> 2, m = q * 32 + lane;
  const uint32_t t_lane = tmem + ((q * 32u) << 16);
  uint8_t* S = sm;
  {
    const float mean = red[1024 + m], rstd = red[1152 + m];
#pragma unroll 1
    for (int cc = 0; cc < 2; ++cc) {
      const uint32_t col0 = h * 64 + cc * 32;
      uint32_t v[32];
      tmem_ld32(t_lane + col0, v);
      tmem_ld_wait();
#pragma unroll
      for (int k4 = 0; k4 < 8; ++k4) {
        const float4 u4 = ld4(uvec + col0 + 4 * k4), v4 = ld4(vvec + col0 + 4 * k4);
        float4 o;
        o.x = rstd * (__uint_as_float(v[4 * k4 + 0]) - mean * u4.x) + v4.x;
        o.y = rstd * (__uint_as_float(v[4 * k4 + 1]) - mean * u4.y) + v4.y;
        o.z = rstd * (__uint_as_float(v[4 * k4 + 2]) - mean * u4.z) + v4.z;
        o.w = rstd * (__uint_as_float(v[4 * k4 + 3]) - mean * u4.w) + v4.w;
        *stage_ptr(S, m, (col0 >> 2) + k4) = o;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  for (int r0 = warp * 16; r0 < warp * 16 + 16; r0 += 4) {     // 4 pixels in flight per warp
    float4 pv[4];
    uint2 gv[4];
    size_t px[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t r = r0 + u, row = 2 * T.t2 + (r >> 6), col = 64 * T.jb + (r & 63);
      ok[u] = row < (uint32_t)N && col < (uint32_t)N;
      px[u] = ((size_t)T.b * N + row) * N + col;
      if (ok[u]) {
        const uint32_t cl = lane ^ (r & 31);
        pv[u] = ld4(p + px[u] * 128 + cl * 4);
        gv[u] = *reinterpret_cast<const uint2*>(g + px[u] * 128 + cl * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t r = r0 + u;
      if (!ok[u]) continue;
      const float4 v = *reinterpret_cast<const float4*>(S + r * 512u + (lane << 4));
      const float2 g01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv[u].x));
      const float2 g23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv[u].y));
      float4 o = pv[u];
      o.x += v.x * g01.x;
      o.y += v.y * g01.y;
      o.z += v.z * g23.x;
      o.w += v.w * g23.y;
      st4(p + px[u] * 128 + ((lane ^ (r & 31)) << 2), o);
    }
  }
  if (warp == 1) tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------------------------
// K6, GB_PREC_BF16X3: the x tile arrives as a hi | lo image pair (two 32 KB bulk copies), the folded weight W' = W_z * gamma
// as hi and lo images; D = x_hi W'_hi + x_lo W'_hi + x_hi W'_lo.  W'_lo takes W'_hi's place in shared memory once the first
// two passes have completed (96 KB per CTA keeps two CTAs per SM); a ninth warp does the copies and the MMA issue so the
// eight statistics warps never wait for it.  Statistics are taken from hi + lo (exact in fp32); u = row sums of
// W'_hi + W'_lo; g is fp32.  Same staging / coalesced read-modify-write of p as the bf16 kernel.
// ---------------------------------------------------------------------------------------------
#define K6S_THREADS 288
#define K6S_SMEM (98304 + 1024 + 64 + 5120)

// per-pixel sum over the 128 channels of (hi + lo - mean)^POW of this thread's (8-pixel chunk, 8-channel group)
template <bool SQ>
__device__ __forceinline__ void k6s_partial(const uint8_t* base, uint32_t q8, const float (&mean)[8], float (&acc)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
  for (uint32_t k = 0; k < 8; ++k) {
    const uint4 wh = *reinterpret_cast<const uint4*>(base + k * 128u + ((q8 ^ k) << 4));
    const uint4 wl = *reinterpret_cast<const uint4*>(base + 32768u + k * 128u + ((q8 ^ k) << 4));
    const uint32_t hh[4] = {wh.x, wh.y, wh.z, wh.w}, ll[4] = {wl.x, wl.y, wl.z, wl.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 a = unpack_bf16x2(hh[t]), b = unpack_bf16x2(ll[t]);
      if (SQ) {
        const float d0 = (a.x + b.x) - mean[2 * t], d1 = (a.y + b.y) - mean[2 * t + 1];
        acc[2 * t] = fmaf(d0, d0, acc[2 * t]);
        acc[2 * t + 1] = fmaf(d1, d1, acc[2 * t + 1]);
      } else {
        acc[2 * t] += a.x + b.x;
        acc[2 * t + 1] += a.y + b.y;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
}

__global__ void __launch_bounds__(K6S_THREADS, 2) trimul_out_x3_kernel(
    const bf16* __restrict__ x, size_t lo_x, const float* __restrict__ g, float* __restrict__ p, int N, int Np,
    const bf16* __restrict__ Wz_hi, const bf16* __restrict__ Wz_lo, const float* __restrict__ uvec, const float* __restrict__ vvec,
    const TileRange tr) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* X = sm;                       // hi 32K | lo 32K
  uint8_t* W = sm + 65536;
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(sm + 98304);
  uint64_t* bar_w2 = bar_w + 1;
  uint64_t* bar_mma1 = bar_w + 2;
  uint64_t* bar_mma = bar_w + 3;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_w + 4);
  float* red = reinterpret_cast<float*>(sm + 98304 + 64);     // [8][128] partials, then mean [128], rstd [128]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const PixTile T = pix_tile(blockIdx.x, tr);
  if (tid == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_w2, 1);
    mbar_init(bar_mma1, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tslot, 128);
  if (tid >= 32 && tid < 36) {           // L2 prefetch of the tile's p and g rows (read last, few bytes in flight per warp)
    const uint32_t ii = (tid >> 1) & 1, row = 2 * T.t2 + ii, col = 64 * T.jb;
    if (row < (uint32_t)N && col < (uint32_t)N) {
      const uint32_t npx = min(64u, (uint32_t)N - col);
      const size_t px = ((size_t)T.b * N + row) * N + col;
      l2_prefetch((tid & 1 ? g : p) + px * 128, npx * 512u);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  if (warp == 8) {
    if (lane == 0) {
      mbar_expect_tx(bar_w, 98304);
      bulk_g2s(W, Wz_hi, 32768, bar_w);
    }
    pdl_wait();            // x comes from the contraction kernel
    pdl_launch();
    if (lane == 0) {
      const size_t off = (((size_t)T.b * (Np >> 1) + T.t2) * (Np >> 6) + T.jb) * 16384;   // (the split mode is not sharded)
      bulk_g2s(X, x + off, 32768, bar_w);
      bulk_g2s(X + 32768, x + lo_x + off, 32768, bar_w);
    }
    __syncwarp();
    mbar_wait(bar_w, 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc(128, 128) | (1u << 15);      // A (x tile) is MN-major
    const uint32_t xs = smem_u32(X), ws = smem_u32(W), tm = warp_uniform(tmem);
    if (elect_one()) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma(tm, umma_desc_mn(xs + k * 2048, 16384), umma_desc(ws + (k >> 2) * 16384 + (k & 3) * 32), idesc, k ? 1u : 0u);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma(tm, umma_desc_mn(xs + 32768 + k * 2048, 16384), umma_desc(ws + (k >> 2) * 16384 + (k & 3) * 32), idesc, 1u);
      umma_commit(bar_mma1);
    }
    __syncwarp();
    mbar_wait(bar_mma1, 0);          // W'_hi has been read: W'_lo takes its place
    if (lane == 0) {
      mbar_expect_tx(bar_w2, 32768);
      bulk_g2s(W, Wz_lo, 32768, bar_w2);
    }
    __syncwarp();
    mbar_wait(bar_w2, 0);
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma(tm, umma_desc_mn(xs + k * 2048, 16384), umma_desc(ws + (k >> 2) * 16384 + (k & 3) * 32), idesc, 1u);
      umma_commit(bar_mma);
    }
    __syncwarp();
  } else {
    pdl_wait();
    pdl_launch();
    // per-pixel statistics over the 128 channels of hi + lo, two passes; thread = (8-pixel chunk, 8-channel group)
    mbar_wait(bar_w, 0);
    const uint32_t pc = tid & 15, cg = tid >> 4, q8 = pc & 7;
    const uint8_t* base = X + (pc >> 3) * 16384u + cg * 1024u;
    float acc[8], mean[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    k6s_partial<false>(base, q8, mean, acc);
    float* part = red + (cg >> 1) * 128 + pc * 8;
    if ((lane & 16) == 0) {
      st4(part, make_float4(acc[0], acc[1], acc[2], acc[3]));
      st4(part + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (tid < 128) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) t += red[r * 128 + tid];
      red[1024 + tid] = t * (1.f / 128.f);
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    {
      const float4 m0 = ld4(red + 1024 + pc * 8), m1 = ld4(red + 1024 + pc * 8 + 4);
      mean[0] = m0.x; mean[1] = m0.y; mean[2] = m0.z; mean[3] = m0.w; mean[4] = m1.x; mean[5] = m1.y; mean[6] = m1.z; mean[7] = m1.w;
    }
    k6s_partial<true>(base, q8, mean, acc);
    if ((lane & 16) == 0) {
      st4(part, make_float4(acc[0], acc[1], acc[2], acc[3]));
      st4(part + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (tid < 128) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) t += red[r * 128 + tid];
      red[1152 + tid] = rsqrtf(t * (1.f / 128.f) + 1e-5f);
    }
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  __syncthreads();                                               // stats visible; X / W dead -> staging
  uint8_t* S = sm;
  if (warp < 8) {
    const uint32_t q = warp & 3, h = warp >> 2, m = q * 32 + lane;
    const uint32_t t_lane = tmem + ((q * 32u) << 16);
    const float mean = red[1024 + m], rstd = red[1152 + m];
#pragma unroll 1
    for (int cc = 0; cc < 2; ++cc) {
      const uint32_t col0 = h * 64 + cc * 32;
      uint32_t v[32];
      tmem_ld32(t_lane + col0, v);
      tmem_ld_wait();
#pragma unroll
      for (int k4 = 0; k4 < 8; ++k4) {
        const float4 u4 = ld4(uvec + col0 + 4 * k4), v4 = ld4(vvec + col0 + 4 * k4);
        float4 o;
        o.x = rstd * (__uint_as_float(v[4 * k4 + 0]) - mean * u4.x) + v4.x;
        o.y = rstd * (__uint_as_float(v[4 * k4 + 1]) - mean * u4.y) + v4.y;
        o.z = rstd * (__uint_as_float(v[4 * k4 + 2]) - mean * u4.z) + v4.z;
        o.w = rstd * (__uint_as_float(v[4 * k4 + 3]) - mean * u4.w) + v4.w;
        *stage_ptr(S, m, (col0 >> 2) + k4) = o;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp < 8) {
    for (int r0 = warp * 16; r0 < warp * 16 + 16; r0 += 4) {     // 4 pixels in flight per warp
      float4 pv[4], gv[4];
      size_t px[4];
      bool ok[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t r = r0 + u, row = 2 * T.t2 + (r >> 6), col = 64 * T.jb + (r & 63);
        ok[u] = row < (uint32_t)N && col < (uint32_t)N;
        px[u] = ((size_t)T.b * N + row) * N + col;
        if (ok[u]) {
          const uint32_t cl = lane ^ (r & 31);
          pv[u] = ld4(p + px[u] * 128 + cl * 4);
          gv[u] = ld4(g + px[u] * 128 + cl * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t r = r0 + u;
        if (!ok[u]) continue;
        const float4 v = *reinterpret_cast<const float4*>(S + r * 512u + (lane << 4));
        float4 o = pv[u];
        o.x = fmaf(v.x, gv[u].x, o.x);
        o.y = fmaf(v.y, gv[u].y, o.y);
        o.z = fmaf(v.z, gv[u].z, o.z);
        o.w = fmaf(v.w, gv[u].w, o.w);
        st4(p + px[u] * 128 + ((lane ^ (r & 31)) << 2), o);
      }
    }
  }
  if (warp == 1) tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int tc_init(gb_ctx* c) {
  cudaError_t e;
  if ((e = cudaFuncSetAttribute(trimul_contract_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K5_SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(trimul_contract_pers_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, K5P<false>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(trimul_contract_pers_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, K5P<true>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(trimul_out_x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K6S_SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(trimul_out_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K6_SMEM)) != cudaSuccess)
    return gb_fail(c, GB_ERR_CUDA, "tc_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  c->tc_state = new TcState();
  int rc = tc_proj_init(c);
  if (rc == GB_OK) rc = tc_transition_init(c);
  if (rc == GB_OK) rc = tc_triatt_init(c);
  return rc != GB_OK ? rc : tc_struct_init(c);
}

void tc_destroy(gb_ctx* c) {
  delete static_cast<TcState*>(c->tc_state);
  c->tc_state = nullptr;
}

int tc_pack_trimul(gb_ctx* c, TriMulW& w, const std::vector<float>& ap, const std::vector<float>& ag,
                   const std::vector<float>& bp, const std::vector<float>& bg, const std::vector<float>& lg,
                   const std::vector<float>& lz, const std::vector<float>& apb, const std::vector<float>& agb,
                   const std::vector<float>& bpb, const std::vector<float>& bgb, const std::vector<float>& lgb,
                   const std::vector<float>& lzb, const std::vector<float>& lni_g, const std::vector<float>& lni_b,
                   const std::vector<float>& lno_g, const std::vector<float>& lno_b) {
  {
    const std::vector<float>* Wm[5] = {&ap, &ag, &bp, &bg, &lg};
    const std::vector<float>* bm[5] = {&apb, &agb, &bpb, &bgb, &lgb};
    const int rc = tc_pack_trimul_proj(c, w, Wm, bm, lni_g, lni_b);
    if (rc != GB_OK) return rc;
  }
  // K6 folds LayerNorm_out into the GEMM: W' = W_z * gamma, u = row sums of bf16(W'), v = W_z beta + b_z
  std::vector<uint16_t> wz(128 * 128);
  pack_image(wz.data(), 128, 128, [&](int r, int k) { return lz[(size_t)r * 128 + k] * lno_g[k]; });
  w.tc_Wz = upload_bytes(c, wz.data(), wz.size() * 2);
  std::vector<float> uv(256);
  for (int n = 0; n < 128; ++n) {
    double u = 0.0, v = lzb[n];
    for (int k = 0; k < 128; ++k) {
      const uint32_t bits = (uint32_t)f2bf(lz[(size_t)n * 128 + k] * lno_g[k]) << 16;
      float wr;
      memcpy(&wr, &bits, 4);
      u += wr;
      v += (double)lz[(size_t)n * 128 + k] * lno_b[k];
    }
    uv[n] = (float)u;
    uv[128 + n] = (float)v;
  }
  w.tc_uv = (float*)upload_bytes(c, uv.data(), uv.size() * 4);
  if (!w.tc_Wz || !w.tc_uv) return gb_fail(c, GB_ERR_CUDA, "tc_pack_trimul: device allocation failed");
  // GB_PREC_BF16X3: W' as hi | lo images, u = row sums of W'_hi + W'_lo (what the three passes multiply the mean with)
  std::vector<uint16_t> wl(128 * 128);
  pack_image(wl.data(), 128, 128, [&](int r, int k) { return bf_lo(lz[(size_t)r * 128 + k] * lno_g[k]); });
  w.x3_Wz_hi = w.tc_Wz;
  w.x3_Wz_lo = upload_bytes(c, wl.data(), wl.size() * 2);
  for (int n = 0; n < 128; ++n) {
    double u = 0.0;
    for (int k = 0; k < 128; ++k) {
      const float wf = lz[(size_t)n * 128 + k] * lno_g[k];
      u += (double)bf_round(wf) + (double)bf_round(bf_lo(wf));
    }
    uv[n] = (float)u;
  }
  w.x3_uv = (float*)upload_bytes(c, uv.data(), uv.size() * 4);
  if (!w.x3_Wz_lo || !w.x3_uv) return gb_fail(c, GB_ERR_CUDA, "tc_pack_trimul: device allocation failed");
  return GB_OK;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

size_t tc_workspace_bytes(const gb_ctx* c, int B, int N) {
  const size_t Np = (size_t)((N + 63) / 64) * 64;
  const size_t plane = (size_t)B * 128 * Np * Np * 2;
  const size_t gb = align256((size_t)B * N * N * 128 * 2);
  // triangle attention (tc_triatt.cu) keeps Q / K / V^T / gate in planes of its own: the a / b / x planes rely on their
  // k-padding staying zero from one triangle update to the next
  const size_t ta = c->d.include_tri_att ? 2 * (3 * align256(plane) + gb) : 0;     // hi planes, then their lo twins (bf16x3)
  return 6 * align256(plane) + gb + ta + 1024 + tc_struct_workspace_bytes(c, B, N);
}

int tc_bind_workspace(gb_ctx* c, void* base, size_t bytes, int B, int N) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const size_t Np = (size_t)((N + 63) / 64) * 64;
  const size_t plane = align256((size_t)B * 128 * Np * Np * 2);
  const size_t gbytes = align256((size_t)B * N * N * 128 * 2);
  if (bytes < 6 * plane + gbytes) return gb_fail(c, GB_ERR_INVALID, "tc_bind_workspace: too small");
  char* q = static_cast<char*>(base);
  s->Np = (int)Np;
  s->a = reinterpret_cast<bf16*>(q);
  s->b = reinterpret_cast<bf16*>(q + plane);
  s->x = reinterpret_cast<bf16*>(q + 2 * plane);
  s->g = reinterpret_cast<bf16*>(q + 3 * plane);
  s->ab_bytes = 2 * plane;
  // [a | b | x | g (bf16)] [a_lo | b_lo | x_lo]: the lo planes of GB_PREC_BF16X3 sit at one common offset behind their hi planes
  s->lo_plane = (3 * plane + gbytes) / 2;
  // the k-padding (N <= k < Np) of the operand images is never written by K4 and must read as zero
  if (cudaMemset(q, 0, 3 * plane) != cudaSuccess || cudaMemset(q + 3 * plane + gbytes, 0, 3 * plane) != cudaSuccess)
    return gb_fail(c, GB_ERR_CUDA, "tc_bind_workspace: memset failed");
  size_t used = 6 * plane + gbytes;
  s->ta_q = s->ta_k = s->ta_vt = s->ta_g = nullptr;
  if (c->d.include_tri_att) {
    if (bytes < used + 2 * (3 * plane + gbytes)) return gb_fail(c, GB_ERR_INVALID, "tc_bind_workspace: too small");
    s->ta_q = reinterpret_cast<bf16*>(q + used);
    s->ta_k = reinterpret_cast<bf16*>(q + used + plane);
    s->ta_vt = reinterpret_cast<bf16*>(q + used + 2 * plane);
    s->ta_g = reinterpret_cast<bf16*>(q + used + 3 * plane);
    s->ta_lo = (3 * plane + gbytes) / 2;
    used += 2 * (3 * plane + gbytes);
  }
  return tc_struct_bind(c, q + used, bytes - used, B, N);
}

int tc_trimul_stage(gb_ctx* c, const TriMulW& w, bool outgoing, int stage, float* p, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int B = c->B, N = c->N, Np = s->Np, W = c->shard_world;
  const bool x3 = c->precision == GB_PREC_BF16X3;
  if (x3 && W > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding is built for the bf16 pair stack only");
  // Row-sharded (W > 1, B == 1): K4 runs on this rank's rows of p and writes blocks (d, me) of the operand images, in its own
  // memory, for every channel group d; the contraction is sharded by CHANNEL (x^c = a^c b^cT is independent per channel):
  // rank d takes its 128 / W channels for all rows and all k.  It first pulls blocks (d, s) of a and b out of rank s's memory
  // (peer-mapped over NVLink, shard.cu: one copy kernel, bulk copies both ways, no staging buffers, no rendezvous) because
  // K5 reads every operand block 4 - 8 times and peer reads bypass the local L2; K5 then runs on local memory and writes x
  // blocks (s, d) locally; K6 of rank s fetches its rows of x straight from every channel owner (each tile is read once).
  // Two device-side barriers per update (operands ready / x ready) replace every collective; per rank and update
  // 3 (W - 1) / W^2 of an image cross NVLink (a, b, x) against (W - 1) / W for an all-gather of b.  Write-after-read across
  // updates is covered by the same barriers: K4 of update n + 1 follows this rank's "x ready" barrier of update n, which every
  // rank reaches only after its pull and K5 of update n; K5 of update n + 1 follows "operands ready", which follows every
  // rank's K6 of update n.
  const ImgLayout Lab = W > 1 ? (outgoing ? img_layout(B, Np, W, 1, W) : img_layout(B, Np, 1, W, W)) : img_layout(B, Np, 1, 1);
  const XLayout Lx = x_layout(B, Np, W);
  PeerImg px;
  for (int r = 0; r < 8; ++r) {
    const char* base = W > 1 ? (const char*)c->peer_ws[r < W ? r : 0] : (const char*)c->ws_alloc;
    if (W > 1 && !base) return gb_fail(c, GB_ERR_INVALID, "row sharding: peer workspaces are not mapped");
    px.p[r] = (const bf16*)(base + ((const char*)s->x - (const char*)c->ws_alloc));
  }
  int rc;
  if (stage == 0 || stage == 1) {
    ph_mark(st, "K4 proj");
    rc = x3 ? tc_launch_trimul_proj_x3(c, w, outgoing, p, st) : tc_launch_trimul_proj(c, w, outgoing, p, st);
    if (rc != GB_OK) return rc;
    if (W > 1) {
      ph_mark(st, "barrier");
      if ((rc = shard_barrier(c, st)) != GB_OK) return rc;                 // operands ready on every rank
      ph_mark(st, "pull a, b");
      if ((rc = shard_pull_blocks(c, s->a, s->b, Lab.blk * 2, st)) != GB_OK) return rc;
    }
  }
  if (stage == 0 || stage == 2) {
    ph_mark(st, "K5 contract");
    const int NJ = 256;
    const int nib = (Np + 127) / 128, ib0 = 0;
    const int nchs = 128 / W, ch0 = W > 1 ? c->shard_rank * nchs : 0;     // sharded: this rank's channel group, every row block
    const unsigned cgrid = (unsigned)B * nchs * nib * ((Np + NJ - 1) / NJ);
    static const int k5pers = [] { const char* e = getenv("GB_K5_PERS"); return e ? atoi(e) : 1; }();
    if (k5pers || x3) {
      const unsigned pgrid = cgrid < (unsigned)s->num_sms ? cgrid : (unsigned)s->num_sms;
      if (x3) launch_pdl_cls(2, trimul_contract_pers_tc_kernel<true>, pgrid, K5P_THREADS, K5P<true>::SMEM, st, s->a, s->b, s->x, N, Np, NJ, Lab, Lab, ib0, nib, (uint32_t)cgrid, s->lo_plane, s->lo_plane, ch0, nchs, Lx);
      else launch_pdl_cls(2, trimul_contract_pers_tc_kernel<false>, pgrid, K5P_THREADS, K5P<false>::SMEM, st, s->a, s->b, s->x, N, Np, NJ, Lab, Lab, ib0, nib, (uint32_t)cgrid, (size_t)0, (size_t)0, ch0, nchs, Lx);
    } else
    launch_pdl_cls(2, trimul_contract_tc_kernel, cgrid, 256, K5_SMEM, st, s->a, s->b, s->x, N, Np, NJ, Lab, Lab, ib0, nib, ch0, nchs, Lx);
    GB_LAUNCH_CHECK(c, "trimul_contract_tc");
    if (W > 1) ph_mark(st, "barrier");
    if (W > 1 && (rc = shard_barrier(c, st)) != GB_OK) return rc;        // x ready on every rank
  }
  if (stage == 0 || stage == 3) {
    ph_mark(st, "K6 out");
    TileRange tr = full_tile_range(N, Np);
    if (W > 1) { tr.nt2 = (uint32_t)(N / W) >> 1; tr.t2_0 = c->shard_rank * tr.nt2; }
    const unsigned tiles = (unsigned)B * tr.nt2 * tr.njb;
    if (x3) launch_pdl_cls(3, trimul_out_x3_kernel, tiles, K6S_THREADS, K6S_SMEM, st, s->x, s->lo_plane, c->g, p, N, Np, (const bf16*)w.x3_Wz_hi, (const bf16*)w.x3_Wz_lo, w.x3_uv, w.x3_uv + 128, tr);
    else launch_pdl_cls(3, trimul_out_tc_kernel, tiles, 256, K6_SMEM, st, px, s->g, p, N, Np, (const bf16*)w.tc_Wz, w.tc_uv, w.tc_uv + 128, tr, Lx);
    GB_LAUNCH_CHECK(c, "trimul_out_tc");
  }
  return GB_OK;
}

int tc_trimul(gb_ctx* c, const TriMulW& w, bool outgoing, float* p, cudaStream_t st) { return tc_trimul_stage(c, w, outgoing, 0, p, st); }

extern "C" int gb_dbg_tc_copy(gb_ctx* c, int which, void* dst, int64_t bytes) {
  if (!c || !c->tc_state || !c->features_set) return GB_ERR_INVALID;
  TcState* s = static_cast<TcState*>(c->tc_state);
  if (!dst) return s->Np;
  const size_t plane = (size_t)c->B * 128 * s->Np * s->Np * 2, gb = (size_t)c->B * c->N * c->N * 128 * 2;
  // 0..3: a, b, x, g (bf16); 4..6: lo planes of a, b, x (GB_PREC_BF16X3); 7: fp32 g of that mode
  const void* src = which == 0 ? (void*)s->a : which == 1 ? (void*)s->b : which == 2 ? (void*)s->x : which == 3 ? (void*)s->g
                  : which == 4 ? (void*)(s->a + s->lo_plane) : which == 5 ? (void*)(s->b + s->lo_plane)
                  : which == 6 ? (void*)(s->x + s->lo_plane) : which == 7 ? (void*)c->g : nullptr;
  const size_t have = which == 3 ? gb : which == 7 ? 2 * gb : plane;
  if (!src || bytes < 0 || (size_t)bytes > have) return gb_fail(c, GB_ERR_INVALID, "gb_dbg_tc_copy: bad arguments");
  GB_CUDA(c, cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice));
  return s->Np;
}
