// K8 on tensor cores (GB_PREC_BF16, and GB_PREC_BF16X3 with split hi | lo operands: template parameter NS = 2): triangle
// attention around the starting / ending node
// (modules/triangular_attention.py:67-130, modules/primitives.py:222-281; 4 heads x 32).
//   p += linear_o( softmax_k( q_ij . k_ik / sqrt(32) + 1e9 (mask_ik - 1) + b_jk ) v_ik * sigmoid(g_ij) )
// "Row" i and positions j, k live in the ATTENTION frame: for the starting node that is the pair map itself, for the
// ending node its transpose (triangular_attention.py:84-86,127-128); the transpose is an index mapping when p is read and
// written, pix(i, j) = starting ? (i, j) : (j, i); every intermediate buffer is laid out in the attention frame.
//
//   triatt_proj_tc_kernel   one CTA = 128 consecutive positions j of one row (b, i): LayerNorm -> bf16 operand image X;
//                           four 128 x 128 x 128 tcgen05 GEMMs into the 512 tensor-memory columns:
//                             D_q [px][ch] | D_k [px][ch] | D_g [px][ch] | D_vT [ch][px] = W_v X^T (so V lands key-contiguous)
//                           outputs: Q, K row images [b][i][2 kb][Np rows][64] (1/sqrt(32) folded into W_q), V^T images
//                           [b][i][Np/64][128 ch][64 keys], gate sigmoid(.) bf16 [b][i][j][128], and the triangle bias
//                           b_jk = linear(x_jk) [b][4][Np/4][j][4] in fp32 on CUDA cores from the normalised row (4 dot products).
//   triatt_core_tc_kernel   one CTA = 128 queries j of one row (b, i), all four heads, keys in blocks of 256:
//                             S = Q_h K_h^T            SS MMA, M = 128, N <= 256, K = 32 (two k steps inside a 64-wide image)
//                             softmax                  256 threads: thread = (query row, half of the keys); bias and mask
//                                                      added in the reference's order; running max / sum (flash style)
//                             O_h += P V_h             TS MMA: P (bf16) from tensor memory, V^T rows of head h from shared
//                           then og = O / l * gate as an operand image, D = og W_o^T, p[pix] += D + b_o.
//   Tensor memory: S 256 | P 128 | O 4 x 32 columns.
// NS = 2 (bf16x3): every operand is a hi | lo pair (the lo plane `lo` elements behind the hi plane), every product the three
// passes hi hi + lo hi + hi lo.  Projections: the hi | lo weights of one matrix at a time stream through two 64 KB slots.  Core:
// key blocks of 128 instead of 256 (K / V^T hi | lo fill the same 64 KB each), P as hi and lo halves of the same 128 columns,
// the output-projection weights take the K buffer's place after the last block.
#include <vector>

#include "common.cuh"
#include "tc_api.h"
#include "tc_common.cuh"
#include "tc_host.h"
#include "tc_state.h"

using namespace tc;

#define TA_HEADS 4
#define TA_CH 32

__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// 64 fp32 values of one thread -> one 128-byte row of an operand image (8 chunks of 8 bf16, chunk c at position c ^ (r & 7))
__device__ __forceinline__ void store_image_row(uint8_t* row, uint32_t r, const uint32_t (&v)[64]) {
  const bool swap = r & 1u;
#pragma unroll
  for (int cp = 0; cp < 4; ++cp) {     // chunk pairs share a 32-byte sector
    uint4 o[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const uint32_t* s = v + (2 * cp + c) * 8;
      o[c].x = pack_bf16(__uint_as_float(s[0]), __uint_as_float(s[1]));
      o[c].y = pack_bf16(__uint_as_float(s[2]), __uint_as_float(s[3]));
      o[c].z = pack_bf16(__uint_as_float(s[4]), __uint_as_float(s[5]));
      o[c].w = pack_bf16(__uint_as_float(s[6]), __uint_as_float(s[7]));
    }
    st_global_32B(row + (((2u * cp) ^ (r & 6u)) << 4), swap ? o[1] : o[0], swap ? o[0] : o[1]);
  }
}

// the same for the NS components of a split row (component c `lo_bytes` behind component c - 1)
template <int NS>
__device__ __forceinline__ void store_image_row_split(uint8_t* row, size_t lo_bytes, uint32_t r, const uint32_t (&v)[64]) {
  if constexpr (NS == 1) {
    store_image_row(row, r, v);
  } else {
    const bool swap = r & 1u;
#pragma unroll
    for (int cp = 0; cp < 4; ++cp) {
      uint32_t h[8], l[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) split_bf16x2(__uint_as_float(v[cp * 16 + 2 * e]), __uint_as_float(v[cp * 16 + 2 * e + 1]), h[e], l[e]);
      uint8_t* d = row + (((2u * cp) ^ (r & 6u)) << 4);
      const uint4 h0 = make_uint4(h[0], h[1], h[2], h[3]), h1 = make_uint4(h[4], h[5], h[6], h[7]);
      const uint4 l0 = make_uint4(l[0], l[1], l[2], l[3]), l1 = make_uint4(l[4], l[5], l[6], l[7]);
      st_global_32B(d, swap ? h1 : h0, swap ? h0 : h1);
      st_global_32B(d + lo_bytes, swap ? l1 : l0, swap ? l0 : l1);
    }
  }
}

// D[128 x 128] (+)= A . B^T over K = 128 with NS-way split operands: A / B images hi at the base, lo 32 KB behind
template <int NS>
__device__ __forceinline__ void umma_tile_split(uint32_t d, uint32_t a, uint32_t b, bool acc) {
  umma_tile(d, a, b, 128, 2, acc);
  if constexpr (NS == 2) {
    umma_tile(d, a + 32768u, b, 128, 2, true);
    umma_tile(d, a, b + 32768u, 128, 2, true);
  }
}

// ---------------------------------------------------------------------------------------------
// projections
// ---------------------------------------------------------------------------------------------
// shared memory: X (NS x 32 KB) | weights (NS = 1: four resident 32 KB images; NS = 2: two 64 KB slots, hi | lo of one matrix)
// | triangle-bias staging [4][128] | barriers
template <int NS> struct TAP {
  static constexpr uint32_t X = NS * 32768u, W = 131072u, SMEM = X + W + 2048 + 128 + 1024;
};

template <int NS>
__global__ void __launch_bounds__(256, 1) triatt_proj_tc_kernel(
    const float* __restrict__ p, int N, int Np, const float* __restrict__ ln_g, const float* __restrict__ ln_b,
    const bf16* __restrict__ Wimg /* NS = 1: q (scaled) | k | v | g; NS = 2: q_hi q_lo k_hi k_lo v_hi v_lo g_hi g_lo */,
    const float* __restrict__ Wb /*[4][128]*/, const float* __restrict__ bg, bf16* __restrict__ Qimg, bf16* __restrict__ Kimg,
    bf16* __restrict__ VTimg, bf16* __restrict__ G, float* __restrict__ tb, int starting, size_t lo /* elements to the lo planes */) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* X = sm;
  uint8_t* W = sm + TAP<NS>::X;
  float* tbs = reinterpret_cast<float*>(sm + TAP<NS>::X + TAP<NS>::W);    // [4][128]
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(sm + TAP<NS>::X + TAP<NS>::W + 2048);   // [2] weights (slot) loaded
  uint64_t* bar_f = bar_w + 2;                                                        // [2] slot read by its MMAs
  uint64_t* bar_mma = bar_w + 4;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_w + 5);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int jt = blockIdx.x, i = blockIdx.y, b = blockIdx.z;
  const size_t sample = (size_t)b * N * N;
  if (tid == 0) {
    for (int k = 0; k < 2; ++k) { mbar_init(bar_w + k, 1); mbar_init(bar_f + k, 1); }
    mbar_init(bar_mma, 1);
    fence_mbar_init();
    // weights: nothing the previous kernel writes
    if (NS == 1) {
      mbar_expect_tx(bar_w, 4 * 32768);
      bulk_g2s(W, Wimg, 4 * 32768, bar_w);
    } else {
      mbar_expect_tx(bar_w, 65536);
      bulk_g2s(W, Wimg, 65536, bar_w);                         // q hi | lo
      mbar_expect_tx(bar_w + 1, 65536);
      bulk_g2s(W + 65536, Wimg + 2 * 16384, 65536, bar_w + 1); // k hi | lo
    }
  }
  if (warp == 1) tmem_alloc(tslot, 512);
  pdl_wait();
  {
    // LayerNorm of the tile's rows -> bf16 image(s); the triangle bias from the normalised fp32 row.  Positions >= N
    // (padding up to Np) become zero rows: q = k = v = 0 there.
    const float4 g4 = ld4(ln_g + lane * 4), b4 = ld4(ln_b + lane * 4);
    float4 wb[TA_HEADS];
#pragma unroll
    for (int h = 0; h < TA_HEADS; ++h) wb[h] = ld4(Wb + h * 128 + lane * 4);
    const uint32_t kb = lane >> 4, chunk = (lane & 15) >> 1, half = lane & 1;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int m = warp * 16 + it * 8 + u, j = jt * 128 + m;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < N) v[u] = ld4(p + (sample + (starting ? (size_t)i * N + j : (size_t)j * N + i)) * 128 + lane * 4);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t m = warp * 16 + it * 8 + u;
        const bool ok = jt * 128 + (int)m < N;
        const float mean = warp_sum(v[u].x + v[u].y + v[u].z + v[u].w) * (1.f / 128.f);
        const float dx = v[u].x - mean, dy = v[u].y - mean, dz = v[u].z - mean, dw = v[u].w - mean;
        const float rstd = rsqrtf(warp_sum(dx * dx + dy * dy + dz * dz + dw * dw) * (1.f / 128.f) + 1e-5f);
        float4 y = make_float4(dx * rstd * g4.x + b4.x, dy * rstd * g4.y + b4.y, dz * rstd * g4.z + b4.z, dw * rstd * g4.w + b4.w);
        if (!ok) y = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t cx[NS], cy[NS];
        split_bf16x2_n<NS>(y.x, y.y, cx);
        split_bf16x2_n<NS>(y.z, y.w, cy);
#pragma unroll
        for (int c = 0; c < NS; ++c)
          *reinterpret_cast<uint2*>(X + c * 32768u + kb * 16384u + m * 128u + ((chunk ^ (m & 7u)) << 4) + half * 8u) = make_uint2(cx[c], cy[c]);
#pragma unroll
        for (int h = 0; h < TA_HEADS; ++h) {
          const float t = warp_sum(y.x * wb[h].x + y.y * wb[h].y + y.z * wb[h].z + y.w * wb[h].w);
          if (lane == h) tbs[h * 128 + m] = t;
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
  if (warp == 0) {
    const uint32_t xs = smem_u32(X), ws = smem_u32(W), tm = warp_uniform(tmem);
    if constexpr (NS == 1) {
      mbar_wait(bar_w, 0);
      tc_fence_after();
      if (elect_one()) {
        umma_tile(tm, xs, ws, 128, 2, false);                    // D_q  [px][ch]
        umma_tile(tm + 128u, xs, ws + 32768u, 128, 2, false);    // D_k
        umma_tile(tm + 256u, xs, ws + 3 * 32768u, 128, 2, false);// D_g
        umma_tile(tm + 384u, ws + 2 * 32768u, xs, 128, 2, false);// D_vT [ch][px]
        umma_commit(bar_mma);
      }
      __syncwarp();
    } else {
      // matrices q, k through slots 0, 1; then g (slot 0) and v (slot 1) once the first two have been read
      mbar_wait(bar_w, 0);
      tc_fence_after();
      if (elect_one()) { umma_tile_split<NS>(tm, xs, ws, false); umma_commit(bar_f); }
      __syncwarp();
      mbar_wait(bar_w + 1, 0);
      tc_fence_after();
      if (elect_one()) { umma_tile_split<NS>(tm + 128u, xs, ws + 65536u, false); umma_commit(bar_f + 1); }
      __syncwarp();
      mbar_wait(bar_f, 0);
      if (lane == 0) {
        mbar_expect_tx(bar_w, 65536);
        bulk_g2s(W, Wimg + 6 * 16384, 65536, bar_w);             // g hi | lo
      }
      __syncwarp();
      mbar_wait(bar_f + 1, 0);
      if (lane == 0) {
        mbar_expect_tx(bar_w + 1, 65536);
        bulk_g2s(W + 65536, Wimg + 4 * 16384, 65536, bar_w + 1); // v hi | lo
      }
      __syncwarp();
      mbar_wait(bar_w, 1);
      tc_fence_after();
      if (elect_one()) umma_tile_split<NS>(tm + 256u, xs, ws, false);
      __syncwarp();
      mbar_wait(bar_w + 1, 1);
      tc_fence_after();
      if (elect_one()) { umma_tile_split<NS>(tm + 384u, ws + 65536u, xs, false); umma_commit(bar_mma); }
      __syncwarp();
    }
  }
  // triangle bias rows (coalesced over the positions of the tile)
  for (int e = tid; e < TA_HEADS * 128; e += 256) {
    const int h = e >> 7, m = e & 127, j = jt * 128 + m;
    // layout [b][h][Np / 4 key quads][N query rows][4]: the core's threads (lane = query row) read one 16-byte quad each,
    // 512 contiguous bytes per warp load (a row-major [j][k] layout costs 32 lines per load instruction)
    if (j < Np) tb[((((size_t)b * TA_HEADS + h) * (Np >> 2) + (j >> 2)) * N + i) * 4 + (j & 3)] = tbs[e];
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  const uint32_t q = warp & 3, hf = warp >> 2, m = q * 32 + lane;
  const uint32_t t_lane = tmem + ((q * 32u) << 16);
  const int j = jt * 128 + (int)m;
  const size_t rowi = (size_t)b * N + i, lob = lo * 2;
  {
    uint32_t v[64];
    // q, k: this thread = position j, channels [64 hf, 64 hf + 64) = k-block hf of the row image
    tmem_ld32(t_lane + hf * 64u, *reinterpret_cast<uint32_t(*)[32]>(v));
    tmem_ld32(t_lane + hf * 64u + 32u, *reinterpret_cast<uint32_t(*)[32]>(v + 32));
    tmem_ld_wait();
    if (j < Np) store_image_row_split<NS>(reinterpret_cast<uint8_t*>(Qimg + ((rowi * 2 + hf) * Np + j) * 64), lob, (uint32_t)j, v);
    tmem_ld32(t_lane + 128u + hf * 64u, *reinterpret_cast<uint32_t(*)[32]>(v));
    tmem_ld32(t_lane + 128u + hf * 64u + 32u, *reinterpret_cast<uint32_t(*)[32]>(v + 32));
    tmem_ld_wait();
    if (j < Np) store_image_row_split<NS>(reinterpret_cast<uint8_t*>(Kimg + ((rowi * 2 + hf) * Np + j) * 64), lob, (uint32_t)j, v);
    // gate: sigmoid(W_g x + b_g), plain row-major bf16 [b][i][j][128] (hi | lo planes for NS = 2)
    tmem_ld32(t_lane + 256u + hf * 64u, *reinterpret_cast<uint32_t(*)[32]>(v));
    tmem_ld32(t_lane + 256u + hf * 64u + 32u, *reinterpret_cast<uint32_t(*)[32]>(v + 32));
    tmem_ld_wait();
    if (j < N) {
      uint8_t* dst = reinterpret_cast<uint8_t*>(G + (rowi * N + j) * 128 + hf * 64);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[8], ol[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int ch = c * 16 + 2 * e;
          const float x0 = __uint_as_float(v[ch]) + bg[hf * 64 + ch], x1 = __uint_as_float(v[ch + 1]) + bg[hf * 64 + ch + 1];
          if constexpr (NS == 1) o[e] = pack_bf16(fast_sigmoid(x0), fast_sigmoid(x1));
          else split_bf16x2(sigmoid_acc(x0), sigmoid_acc(x1), o[e], ol[e]);
        }
        st_global_32B(dst + c * 32, make_uint4(o[0], o[1], o[2], o[3]), make_uint4(o[4], o[5], o[6], o[7]));
        if constexpr (NS == 2) st_global_32B(dst + c * 32 + lob, make_uint4(ol[0], ol[1], ol[2], ol[3]), make_uint4(ol[4], ol[5], ol[6], ol[7]));
      }
    }
    // v^T: this thread = channel m, keys [64 hf, 64 hf + 64) of the tile = key block 2 jt + hf
    tmem_ld32(t_lane + 384u + hf * 64u, *reinterpret_cast<uint32_t(*)[32]>(v));
    tmem_ld32(t_lane + 384u + hf * 64u + 32u, *reinterpret_cast<uint32_t(*)[32]>(v + 32));
    tmem_ld_wait();
    const uint32_t kb64 = 2 * jt + hf;
    if (kb64 < (uint32_t)(Np >> 6))
      store_image_row_split<NS>(reinterpret_cast<uint8_t*>(VTimg + (((rowi * (Np >> 6)) + kb64) * 128 + m) * 64), lob, m, v);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
// attention core + output projection + residual
// ---------------------------------------------------------------------------------------------
// shared memory: Q tile (NS x 32 KB) | K block 64 KB | V^T block 64 KB | W_o 32 KB (NS = 1; NS = 2 loads hi | lo over the K block at
// the end) | reduction scratch, mask bias, barriers.  Keys per block: 256 (NS = 1) or 128 (NS = 2: hi | lo fill the same 64 KB).
template <int NS> struct TAC {
  static constexpr uint32_t KB = NS == 1 ? 256u : 128u;
  static constexpr uint32_t Q = NS * 32768u, OFF_K = Q, OFF_VT = Q + 65536u, OFF_WO = Q + 2 * 65536u;
  static constexpr uint32_t OFF_MISC = OFF_WO + (NS == 1 ? 32768u : 0u), SMEM = OFF_MISC + 2048 /*red*/ + 1024 /*mbs*/ + 128 + 1024;
};
#define LOG2E 1.4426950408889634f

template <int NS>
__global__ void __launch_bounds__(256, 1) triatt_core_tc_kernel(
    const bf16* __restrict__ Qimg, const bf16* __restrict__ Kimg, const bf16* __restrict__ VTimg, const bf16* __restrict__ G,
    const float* __restrict__ tb, const int* __restrict__ rmask, const bf16* __restrict__ Woimg /* NS = 2: hi | lo */, const float* __restrict__ bo,
    float* __restrict__ p, int N, int Np, int starting, size_t lo /* elements to the lo planes of Q / K / V^T / G */) {
  constexpr uint32_t KB = TAC<NS>::KB;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint8_t* Qs = sm;                       // [2 kb][128 rows][64]; later the og image
  uint8_t* Ks = sm + TAC<NS>::OFF_K;      // NS x [2 kb][nk rows][64] (component c at c * 32768)
  uint8_t* VTs = sm + TAC<NS>::OFF_VT;    // NS x [nk/64][128 ch][64 keys]
  uint8_t* Wos = NS == 1 ? sm + TAC<NS>::OFF_WO : Ks;
  float* red = reinterpret_cast<float*>(sm + TAC<NS>::OFF_MISC);      // [2 halves][128 rows] (+ [2][128] for the sums)
  float* mbs = red + 512;                                               // [256] mask bias of the keys of the block
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(sm + TAC<NS>::OFF_MISC + 3072);
  uint64_t *bar_ld = bar_q + 1, *bar_s = bar_q + 2, *bar_o = bar_q + 3, *bar_w = bar_q + 4;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_q + 5);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qt = blockIdx.x, i = blockIdx.y, b = blockIdx.z;
  const size_t rowi = (size_t)b * N + i;
  const int qrows = min(128, Np - qt * 128);
  if (tid == 0) {
    mbar_init(bar_q, 1); mbar_init(bar_ld, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1); mbar_init(bar_w, 1);
    fence_mbar_init();
    if (NS == 1) {
      mbar_expect_tx(bar_w, 32768);
      bulk_g2s(Wos, Woimg, 32768, bar_w);
    }
  }
  if (warp == 1) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  const uint32_t T_S = 0, T_P = 256, T_O = 384;
  pdl_wait();
  pdl_launch();
  if (tid == 0) {
    mbar_expect_tx(bar_q, NS * 2u * qrows * 128u);
    for (int c = 0; c < NS; ++c)
      for (int kb = 0; kb < 2; ++kb)
        bulk_g2s(Qs + c * 32768 + kb * 16384, Qimg + c * lo + ((rowi * 2 + kb) * Np + qt * 128) * 64, qrows * 128u, bar_q);
  }
  const uint32_t q = warp & 3, hf = warp >> 2, m = q * 32 + lane;
  const uint32_t t_lane = tmem + ((q * 32u) << 16);
  const int j = qt * 128 + (int)m, jc = j < N ? j : N - 1;     // padded query rows read a valid bias row (never written)
  const int mi = rmask[b * N + i];
  float mrun[TA_HEADS], lrun[TA_HEADS];
#pragma unroll
  for (int h = 0; h < TA_HEADS; ++h) { mrun[h] = -3.0e38f; lrun[h] = 0.f; }
  const int nblk = (Np + (int)KB - 1) / (int)KB;
  const uint32_t tm = warp_uniform(tmem), qs = smem_u32(Qs), ks = smem_u32(Ks), vts = smem_u32(VTs);
  uint32_t ns = 0;                     // S commits so far (phase of bar_s)

  auto issue_S = [&](int h, uint32_t nk) {      // warp 0, converged
    tc_fence_after();
    if (elect_one()) {
      const uint32_t kb = h >> 1, off = (h & 1) * 64u;
#pragma unroll
      for (uint32_t ps = 0; ps < (NS == 1 ? 1u : 3u); ++ps) {      // hi hi | lo hi | hi lo
        const uint32_t qa = qs + (ps == 1 ? 32768u : 0u), ka = ks + (ps == 2 ? 32768u : 0u);
#pragma unroll
        for (uint32_t s2 = 0; s2 < 2; ++s2)
          umma(tm + T_S, umma_desc(qa + kb * 16384u + off + s2 * 32u), umma_desc(ka + kb * (nk * 128u) + off + s2 * 32u), umma_idesc(128, nk), ps | s2);
      }
      umma_commit(bar_s);
    }
    __syncwarp();
  };

  for (int blk = 0; blk < nblk; ++blk) {
    const int k0 = blk * (int)KB;
    const uint32_t nk = (uint32_t)min((int)KB, Np - k0), hw = nk >> 1;  // keys of this block; per half
    if (blk > 0) {
      mbar_wait(bar_o, (blk - 1) & 1);    // the P V MMAs of the previous block have read K / V^T and updated O
      tc_fence_after();
      __syncthreads();
    }
    if (tid == 0) {
      mbar_expect_tx(bar_ld, NS * (2u * nk * 128u + nk * 256u));
      for (int c = 0; c < NS; ++c) {
        for (int kb = 0; kb < 2; ++kb)
          bulk_g2s(Ks + c * 32768 + kb * (nk * 128u), Kimg + c * lo + ((rowi * 2 + kb) * Np + k0) * 64, nk * 128u, bar_ld);
        bulk_g2s(VTs + c * 32768, VTimg + c * lo + ((rowi * (Np >> 6)) + (k0 >> 6)) * 8192, nk * 256u, bar_ld);
      }
    }
    if (tid < (int)nk) {
      const int k = k0 + tid;
      mbs[tid] = k < N ? 1e9f * ((float)(mi * rmask[b * N + k]) - 1.f) : -3.0e38f;
    }
    if (blk == 0) mbar_wait(bar_q, 0);
    mbar_wait(bar_ld, blk & 1);
    __syncthreads();
    if (warp == 0) issue_S(0, nk);
    for (int h = 0; h < TA_HEADS; ++h) {
      mbar_wait(bar_s, ns & 1);          // S_h complete (and every MMA issued before it: P V of the previous head)
      ++ns;
      tc_fence_after();
      // ---- pass 1: s = q.k + mask bias + triangle bias (the reference's order), written back; row maximum
      const float* tbr = tb + ((((size_t)b * TA_HEADS + h) * (Np >> 2) + ((k0 + hf * hw) >> 2)) * N + jc) * 4;   // + key quad * 4 N
      float mx = -3.0e38f;
      for (uint32_t c = 0; c < hw; c += 32) {
        uint32_t v[32];
        tmem_ld32(t_lane + T_S + hf * hw + c, v);
        float4 t4[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) t4[e] = ld4(tbr + (size_t)((c >> 2) + e) * N * 4);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float4 mb4 = *reinterpret_cast<const float4*>(mbs + hf * hw + c + 4 * e);
          const float s0 = (__uint_as_float(v[4 * e]) + mb4.x) + t4[e].x, s1 = (__uint_as_float(v[4 * e + 1]) + mb4.y) + t4[e].y;
          const float s2 = (__uint_as_float(v[4 * e + 2]) + mb4.z) + t4[e].z, s3 = (__uint_as_float(v[4 * e + 3]) + mb4.w) + t4[e].w;
          mx = fmaxf(fmaxf(mx, fmaxf(s0, s1)), fmaxf(s2, s3));
          v[4 * e] = __float_as_uint(s0); v[4 * e + 1] = __float_as_uint(s1); v[4 * e + 2] = __float_as_uint(s2); v[4 * e + 3] = __float_as_uint(s3);
        }
        tmem_st32(t_lane + T_S + hf * hw + c, v);
      }
      tmem_st_wait();
      red[hf * 128 + m] = mx;
      __syncthreads();
      const float mnew = fmaxf(mrun[h], fmaxf(red[m], red[128 + m]));
      const float corr = ex2_approx((mrun[h] - mnew) * LOG2E);
      mrun[h] = mnew;
      // ---- pass 2: probabilities (bf16) into tensor memory, row sums; O_h rescaled when the maximum moved
      const float nm = -mnew * LOG2E;
      float sum = 0.f;
      for (uint32_t c = 0; c < hw; c += 32) {
        uint32_t v[32], pk[16], pl[16];
        tmem_ld32(t_lane + T_S + hf * hw + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float p0 = ex2_approx(fmaf(__uint_as_float(v[2 * e]), LOG2E, nm)), p1 = ex2_approx(fmaf(__uint_as_float(v[2 * e + 1]), LOG2E, nm));
          if constexpr (NS == 1) {
            pk[e] = pack_bf16(p0, p1);
            // sum what the tensor core will multiply with: the bf16-rounded probabilities
            const float2 r = unpack_bf16x2(pk[e]);
            sum += r.x + r.y;
          } else {
            split_bf16x2(p0, p1, pk[e], pl[e]);
            sum += p0 + p1;
          }
        }
        tmem_st16(t_lane + T_P + ((hf * hw + c) >> 1), pk);
        if constexpr (NS == 2) tmem_st16(t_lane + T_P + 64u + ((hf * hw + c) >> 1), pl);     // lo halves behind the 64 hi columns
      }
      lrun[h] = lrun[h] * corr + sum;
      if (blk > 0 && hf == 0) {
        uint32_t o[32];
        tmem_ld32(t_lane + T_O + h * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * corr);
        tmem_st32(t_lane + T_O + h * 32, o);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncthreads();
      if (warp == 0) {
        tc_fence_after();
        if (elect_one()) {
          // O_h (+)= P . V_h : A = P from tensor memory (16 keys = 8 columns per step), B = rows [32 h, 32 h + 32) of the V^T blocks
          for (uint32_t kb64 = 0; kb64 < (nk >> 6); ++kb64)
#pragma unroll
            for (uint32_t s4 = 0; s4 < 4; ++s4)
#pragma unroll
              for (uint32_t ps = 0; ps < (NS == 1 ? 1u : 3u); ++ps)      // P_hi V_hi | P_lo V_hi | P_hi V_lo
                umma_ts(tm + T_O + h * 32u, tm + T_P + (ps == 1 ? 64u : 0u) + kb64 * 32u + s4 * 8u,
                        umma_desc(vts + (ps == 2 ? 32768u : 0u) + kb64 * 16384u + h * 4096u + s4 * 32u), umma_idesc(128, 32),
                        (blk | (int)kb64 | (int)s4 | (int)ps) ? 1u : 0u);
          if (h == TA_HEADS - 1) umma_commit(bar_o);
        }
        __syncwarp();
        if (h + 1 < TA_HEADS) issue_S(h + 1, nk);
      }
    }
  }
  mbar_wait(bar_o, (nblk - 1) & 1);
  tc_fence_after();
  // ---- og = O / l * gate -> operand image (in place of the Q tile); this thread: row m, heads 2 hf and 2 hf + 1
  // exchange the partial row sums of the two key halves, one head at a time through red[2][128]
  float inv[TA_HEADS];
#pragma unroll
  for (int h = 0; h < TA_HEADS; ++h) {
    red[hf * 128 + m] = lrun[h];
    __syncthreads();
    inv[h] = 1.f / (red[m] + red[128 + m]);
    __syncthreads();
  }
  if (NS == 2 && tid == 0) {      // every MMA that read the K block has completed: the output-projection weights take its place
    mbar_expect_tx(bar_w, 65536);
    bulk_g2s(Wos, Woimg, 65536, bar_w);
  }
  {
    const uint16_t* grow = reinterpret_cast<const uint16_t*>(G) + (rowi * N + (j < N ? j : 0)) * 128;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int h = 2 * hf + hh;
      uint32_t o[32];
      tmem_ld32(t_lane + T_O + h * 32, o);
      uint4 g4[4], gl4[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        g4[e] = *reinterpret_cast<const uint4*>(grow + h * 32 + e * 8);
        if constexpr (NS == 2) gl4[e] = *reinterpret_cast<const uint4*>(grow + lo + h * 32 + e * 8);
      }
      tmem_ld_wait();
      const uint32_t* gw = reinterpret_cast<const uint32_t*>(g4);
      const uint32_t* gwl = reinterpret_cast<const uint32_t*>(gl4);
#pragma unroll
      for (int c = 0; c < 4; ++c) {      // 4 chunks of 8 channels
        uint32_t w[4], wl[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 gg = unpack_bf16x2(gw[c * 4 + e]);
          if constexpr (NS == 2) {
            const float2 gq = unpack_bf16x2(gwl[c * 4 + e]);
            gg.x += gq.x;
            gg.y += gq.y;
          }
          const float y0 = __uint_as_float(o[c * 8 + 2 * e]) * inv[h] * gg.x, y1 = __uint_as_float(o[c * 8 + 2 * e + 1]) * inv[h] * gg.y;
          if constexpr (NS == 1) w[e] = pack_bf16(y0, y1);
          else split_bf16x2(y0, y1, w[e], wl[e]);
        }
        const uint32_t chunk = (uint32_t)(h & 1) * 4u + c;       // channel h*32 + c*8 -> k-block h >> 1, chunk within the 64
        uint8_t* dst = Qs + (h >> 1) * 16384u + m * 128u + ((chunk ^ (m & 7u)) << 4);
        *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
        if constexpr (NS == 2) *reinterpret_cast<uint4*>(dst + 32768u) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
      }
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    mbar_wait(bar_w, 0);
    tc_fence_after();
    if (elect_one()) {
      umma_tile_split<NS>(tm + T_S, qs, smem_u32(Wos), false);
      umma_commit(bar_s);
    }
    __syncwarp();
  }
  mbar_wait(bar_s, ns & 1);
  tc_fence_after();
  {
    // D + b_o -> fp32 staging tile in the (now idle) K buffer, 16-byte chunks XOR-swizzled by the row so that the
    // thread-per-row writer and the warp-per-row reader are both conflict free; then p[pix] += . with one coalesced
    // 512-byte row per warp instruction (the ending node's rows are N pixels apart, each still 512 contiguous bytes)
    uint8_t* S = NS == 1 ? Ks : VTs;      // (NS = 2: the K buffer holds W_o)
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      tmem_ld32(t_lane + T_S + hf * 64 + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float4 b4 = ld4(bo + hf * 64 + c * 32 + 4 * e);
        const uint32_t chunk = (uint32_t)(hf * 16 + c * 8 + e);
        *reinterpret_cast<float4*>(S + m * 512u + ((chunk ^ (m & 31u)) << 4)) =
            make_float4(__uint_as_float(v[4 * e]) + b4.x, __uint_as_float(v[4 * e + 1]) + b4.y, __uint_as_float(v[4 * e + 2]) + b4.z,
                        __uint_as_float(v[4 * e + 3]) + b4.w);
      }
    }
    __syncthreads();
    for (int r0 = warp * 16; r0 < warp * 16 + 16; r0 += 4) {
      float4 pv[4];
      float* dst[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int jr = qt * 128 + r0 + u;
        dst[u] = nullptr;
        if (jr < N) {
          dst[u] = p + ((size_t)b * N * N + (starting ? (size_t)i * N + jr : (size_t)jr * N + i)) * 128 + lane * 4;
          pv[u] = ld4(dst[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (!dst[u]) continue;
        const uint32_t r = (uint32_t)(r0 + u);
        const float4 d = *reinterpret_cast<const float4*>(S + r * 512u + (((uint32_t)lane ^ (r & 31u)) << 4));
        st4(dst[u], make_float4(pv[u].x + d.x, pv[u].y + d.y, pv[u].z + d.z, pv[u].w + d.w));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int tc_triatt_init(gb_ctx* c) {
  cudaError_t e;
  if ((e = cudaFuncSetAttribute(triatt_proj_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAP<1>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(triatt_proj_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAP<2>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(triatt_core_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAC<1>::SMEM)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(triatt_core_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAC<2>::SMEM)) != cudaSuccess)
    return gb_fail(c, GB_ERR_CUDA, "tc_triatt_init: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
  return GB_OK;
}

// wq, wk, wv, wg [128 = 4 heads x 32][128 in]; wo [128 out][128 = 4 heads x 32].  bf16: q (scaled) | k | v | g images and the
// W_o image; bf16x3: hi | lo image pairs q_hi q_lo k_hi k_lo v_hi v_lo g_hi g_lo and W_o hi | lo.
int tc_pack_triatt(gb_ctx* c, TriAttW& w, const std::vector<float>& wq, const std::vector<float>& wk, const std::vector<float>& wv,
                   const std::vector<float>& wg, const std::vector<float>& wo) {
  std::vector<uint16_t> img(4 * 16384), io(16384), img3(8 * 16384), io3(2 * 16384);
  const float sc = 1.f / sqrtf((float)TA_CH);
  const std::vector<float>* src[4] = {&wq, &wk, &wv, &wg};
  for (int t = 0; t < 4; ++t) {
    auto at = [&](int r, int k) { return (t == 0 ? sc : 1.f) * (*src[t])[(size_t)r * 128 + k]; };
    pack_image(img.data() + (size_t)t * 16384, 128, 128, at);
    pack_image(img3.data() + (size_t)(2 * t) * 16384, 128, 128, at);
    pack_image(img3.data() + (size_t)(2 * t + 1) * 16384, 128, 128, [&](int r, int k) { return bf_lo(at(r, k)); });
  }
  auto ato = [&](int r, int k) { return wo[(size_t)r * 128 + k]; };
  pack_image(io.data(), 128, 128, ato);
  pack_image(io3.data(), 128, 128, ato);
  pack_image(io3.data() + 16384, 128, 128, [&](int r, int k) { return bf_lo(ato(r, k)); });
  w.tc_Wqkvg = upload_bytes(c, img.data(), img.size() * 2);
  w.tc_Wo = upload_bytes(c, io.data(), io.size() * 2);
  w.x3_Wqkvg = upload_bytes(c, img3.data(), img3.size() * 2);
  w.x3_Wo = upload_bytes(c, io3.data(), io3.size() * 2);
  if (!w.tc_Wqkvg || !w.tc_Wo || !w.x3_Wqkvg || !w.x3_Wo) return gb_fail(c, GB_ERR_CUDA, "tc_pack_triatt: device allocation failed");
  return GB_OK;
}

// p += TriangleAttention(p) on tensor cores (bf16 or bf16x3).  Q / K / V^T / gate live in planes of their own
// (TcState::ta_*, their lo twins ta_lo elements behind), the triangle bias in c->ta_bias ([B][4][Np/4][N][4] fp32).
int tc_tri_attention(gb_ctx* c, const TriAttW& w, bool starting, float* p, cudaStream_t st) {
  TcState* s = static_cast<TcState*>(c->tc_state);
  const int B = c->B, N = c->N, Np = s->Np;
  const bool x3 = c->precision == GB_PREC_BF16X3;
  if (c->shard_world > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "triangle attention is not row-sharded");
  if (!w.tc_Wqkvg || !w.tc_Wo || !w.x3_Wqkvg || !w.x3_Wo || !c->ta_bias || !s->ta_q)
    return gb_fail(c, GB_ERR_INVALID, "tc_tri_attention: weights not packed / workspace not bound");
  const dim3 grid((Np + 127) / 128, N, B);
  if (x3) {
    launch_pdl(triatt_proj_tc_kernel<2>, grid, 256, TAP<2>::SMEM, st, (const float*)p, N, Np, (const float*)w.ln_g, (const float*)w.ln_b,
               (const bf16*)w.x3_Wqkvg, (const float*)w.Wbias, (const float*)w.bg, s->ta_q, s->ta_k, s->ta_vt, s->ta_g, c->ta_bias, starting ? 1 : 0, s->ta_lo);
    GB_LAUNCH_CHECK(c, "triatt_proj_tc");
    launch_pdl(triatt_core_tc_kernel<2>, grid, 256, TAC<2>::SMEM, st, (const bf16*)s->ta_q, (const bf16*)s->ta_k, (const bf16*)s->ta_vt, (const bf16*)s->ta_g,
               (const float*)c->ta_bias, (const int*)c->mask, (const bf16*)w.x3_Wo, (const float*)w.bo, p, N, Np, starting ? 1 : 0, s->ta_lo);
  } else {
    launch_pdl(triatt_proj_tc_kernel<1>, grid, 256, TAP<1>::SMEM, st, (const float*)p, N, Np, (const float*)w.ln_g, (const float*)w.ln_b,
               (const bf16*)w.tc_Wqkvg, (const float*)w.Wbias, (const float*)w.bg, s->ta_q, s->ta_k, s->ta_vt, s->ta_g, c->ta_bias, starting ? 1 : 0, (size_t)0);
    GB_LAUNCH_CHECK(c, "triatt_proj_tc");
    launch_pdl(triatt_core_tc_kernel<1>, grid, 256, TAC<1>::SMEM, st, (const bf16*)s->ta_q, (const bf16*)s->ta_k, (const bf16*)s->ta_vt, (const bf16*)s->ta_g,
               (const float*)c->ta_bias, (const int*)c->mask, (const bf16*)w.tc_Wo, (const float*)w.bo, p, N, Np, starting ? 1 : 0, (size_t)0);
  }
  GB_LAUNCH_CHECK(c, "triatt_core_tc");
  return GB_OK;
}
