// sm_100a building blocks for the tensor-core kernels: mbarrier, 1-D bulk async copies
// (cp.async.bulk), TMEM allocation, tcgen05.mma / commit / ld and the shared-memory operand
// format they all agree on.
//
// Operand format ("SW128 K-major tile image"): a matrix of R rows (R % 8 == 0) by 64 bf16 of
// K is stored as R rows of 128 bytes; the 16-byte chunk q (8 consecutive k) of row r sits at
// chunk position q ^ (r & 7).  Eight rows form one 1024-byte swizzle atom (SBO = 1024).  A
// K = 128 operand is two such images back to back.  The same byte image is used in global
// memory for the packed weights and for the a / b operands of the triangle contraction, so
// a tile moves global -> shared with one cp.async.bulk and needs no tensor map.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte offset of element (r, k) inside one [R][64] bf16 tile image, k in [0, 64)
__host__ __device__ __forceinline__ uint32_t sw128_off(uint32_t r, uint32_t k) {
  return r * 128u + ((((k >> 3) ^ (r & 7u)) << 4) | ((k & 7u) << 1));
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok;
}
// try_wait with a suspend-time hint: the thread sleeps in hardware until the phase completes or `ns` nanoseconds pass (it
// wakes at once on completion), so a waiting warp issues one instruction per wake-up instead of spinning.
__device__ __forceinline__ uint32_t mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
      : "memory");
  return ok;
}
// Bounded wait: a protocol bug traps (launch error reported to the caller) instead of hanging the GPU.  The wait suspends
// in hardware (10 us hint per try) and only looks at the clock every 64th try after the first 64: in round 1 the CS2R /
// ISETP / BRA of a clock64() polling loop were a quarter of all instructions the persistent kernels issued, taken from
// the warps that had work.  Time-out ~2 s.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t tries = 0;
  long long t0 = 0;
  while (!mbar_try_wait_hint(bar, parity, 10000u)) {
    ++tries;
    if (tries == 64u) t0 = clock64();
    else if (tries > 64u && (tries & 63u) == 0u && clock64() - t0 > 4000000000LL) __trap();
  }
}

// ------------------------------------------------------------------ async copies / proxies
// global -> shared, bytes % 16 == 0, both addresses 16-byte aligned; completes on `bar` (complete_tx)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global bulk copy (bulk async-group completion): commit, then wait until all but `Pending` groups have finished
// READING shared memory (the slot may be refilled) or have finished altogether
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int Pending> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(Pending) : "memory"); }
template <int Pending> __device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(Pending) : "memory"); }
// ask L2 for [src, src + bytes) (bytes % 16 == 0, src 16-byte aligned); no completion, no destination
__device__ __forceinline__ void l2_prefetch(const void* src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
// make generic-proxy shared-memory writes visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------ TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t* slot_smem, uint32_t ncols) {   // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {      // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 consecutive fp32 columns of this thread's TMEM lane (lane = 32 * (warp % 4) + laneid)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// registers -> 32 consecutive columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
      "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
      "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// One lane of a converged warp.  tcgen05.mma / commit are issued as  `if (elect_one()) ...`  from
// warp-uniform control flow with warp-uniform operands: inside a lane-divergent branch the compiler
// cannot keep the descriptors in uniform registers and wraps every UTCHMMA in an R2UR waterfall loop
// (~100 issue cycles per MMA instead of ~10).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t warp_uniform(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }

// ------------------------------------------------------------------ tcgen05.mma (kind::f16, bf16 x bf16 -> fp32)
// shared-memory matrix descriptor of a SW128 K-major tile image starting at `saddr` (1024-byte
// aligned tile base + optional k advance of 32 bytes per 16 bf16)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) /* LBO (unused for swizzled K-major) */
         | (64ull << 32)                                      /* SBO = 1024 B between 8-row groups */
         | (1ull << 46)                                       /* descriptor version (sm_100)       */
         | (2ull << 61);                                      /* SWIZZLE_128B                      */
}
// MN-major A descriptor: 64-channel blocks LBO bytes apart, 8-key groups 1024 B apart, SWIZZLE_128B
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr, uint32_t lbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D fp32, A/B bf16, both K-major, M x N (or bit 15 for an MN-major A)
__host__ __device__ constexpr uint32_t umma_idesc(uint32_t M, uint32_t N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// A operand from tensor memory (128 lanes = rows, 8 columns = 16 bf16 of K packed in pairs), B from shared memory
__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on `bar` once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// D[128 x N] (+)= A[128 x 64*nkb] . B[N x 64*nkb]^T ; A image at a_base (nkb images of 128 rows),
// B image at b_base (nkb images of N rows).  One thread.
__device__ __forceinline__ void umma_tile(uint32_t tmem_d, uint32_t a_base, uint32_t b_base, uint32_t N, int nkb,
                                          bool accumulate_first) {
  const uint32_t idesc = umma_idesc(128, N);
  uint32_t acc = accumulate_first ? 1u : 0u;
  for (int kb = 0; kb < nkb; ++kb) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      umma(tmem_d, umma_desc(a_base + kb * (128 * 128) + k * 32), umma_desc(b_base + kb * (N * 128) + k * 32), idesc, acc);
      acc = 1u;
    }
  }
}

// sigmoid(x) = 0.5 tanh(x / 2) + 0.5 with the one-instruction MUFU tanh (max rel. error 2^-11, below
// the 2^-9 round-off of the bf16 value it feeds)
__device__ __forceinline__ float fast_sigmoid(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return fmaf(0.5f, t, 0.5f);
}

// one 32-byte (full sector) global store; p 32-byte aligned.  Two 16-byte stores from different
// instructions reach L2 as two partial-sector writes and halve the effective store bandwidth.
__device__ __forceinline__ void st_global_32B(void* p, uint4 lo, uint4 hi) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w),
               "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w)
               : "memory");
}

__device__ __forceinline__ void ld_global_32B(const float* p, float4& lo, float4& hi) {
  asm volatile("ld.global.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
               : "l"(p)
               : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Split of fp32 values into bf16 hi + bf16 lo (GB_PREC_BF16X3): hi = bf16(v), lo = bf16(v - hi); v - hi is exact in fp32, so
// hi + lo carries 16 mantissa bits and  x w ~ x_hi w_hi + x_lo w_hi + x_hi w_lo  to ~2^-16 of |x w|.
__device__ __forceinline__ void split_bf16x2(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = pack_bf16(a, b);
  lo = pack_bf16(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xffff0000u));
}
// NS-way split (NS = 1, 2): component c = bf16 of what the components before it left over
template <int NS>
__device__ __forceinline__ void split_bf16x2_n(float a, float b, uint32_t (&c)[NS]) {
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    c[i] = pack_bf16(a, b);
    a -= __uint_as_float(c[i] << 16);
    b -= __uint_as_float(c[i] & 0xffff0000u);
  }
}
// operand components (A, B) of pass i of an NS-way split product, largest terms first.  The structure-net kernels that use
// this table (gemm_img, ipa_prep, ipa_attention) also run the lo x lo pass: they are latency-bound, the pass is nearly free
// (+38 us per eval) and, with the accumulation chains cut short (GI<2>::NACC), it took the 1000-step trajectory from 0.128 to
// 0.10 A.  The pair-stack kernels (K4 - K7) keep three passes: a fourth there cost 120 us and did not help the trajectory.
// A third component (six passes) was tried in round 1 and bought nothing while the fp32 accumulation of tcgen05.mma itself
// left ~1e-5 on a K = 2112 product.
template <int NS> struct SplitPasses;
template <> struct SplitPasses<1> { static constexpr int N = 1; __device__ static constexpr int a(int) { return 0; } __device__ static constexpr int b(int) { return 0; } };
template <> struct SplitPasses<2> { static constexpr int N = 4; __device__ static constexpr int a(int i) { return (i == 1 || i == 3) ? 1 : 0; } __device__ static constexpr int b(int i) { return (i == 2 || i == 3) ? 1 : 0; } };
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) { return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u)); }
// sigmoid to ~1e-6 relative (ex2.approx + rcp.approx): the split mode cannot use the 2^-11 tanh.approx of the bf16 mode
__device__ __forceinline__ float sigmoid_acc(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// ------------------------------------------------------------------ kernel-side conveniences
// dynamic shared memory base rounded up to 1024 bytes (SWIZZLE_128B atoms)
__device__ __forceinline__ uint8_t* smem_base_1024(uint8_t* raw) {
  const uint32_t a = smem_u32(raw);
  return raw + (((a + 1023u) & ~1023u) - a);
}


// Pixel tile of the triangle / IPA-prep kernels: 128 pixels = 2 consecutive rows x 64 consecutive
// columns of one sample's N x N pair map; tile-local pixel m = 64 * ii + jj,
// row = 2 * t2 + ii, col = 64 * jb + jj.
// Programmatic dependent launch.  Every kernel of the chain runs its prologue (barriers, tensor-memory allocation,
// weight prefetch: nothing the previous kernel writes), then pdl_wait() -- the previous grid has completed and its
// writes are visible -- then pdl_launch(), which lets the NEXT kernel start its own prologue while this one computes.
// Wait-then-launch keeps at most one waiting grid resident and makes everything older than the previous grid
// complete before a kernel starts (so a prologue may read data produced two or more launches back).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

struct PixTile { uint32_t b, t2, jb; };
__device__ __forceinline__ PixTile pix_tile(uint32_t blk, int N, int Np) {
  const uint32_t njb = Np >> 6, nt2 = (N + 1) >> 1;
  PixTile t;
  t.jb = blk % njb;
  const uint32_t r = blk / njb;
  t.t2 = r % nt2;
  t.b = r / nt2;
  return t;
}
// Restricted tile set of a row-sharded launch: row pairs [t2_0, t2_0 + nt2) x column blocks [jb_0, jb_0 + njb).
struct TileRange { uint32_t t2_0, nt2, jb_0, njb; };
__device__ __forceinline__ PixTile pix_tile(uint32_t blk, const TileRange& tr) {
  PixTile t;
  t.jb = tr.jb_0 + blk % tr.njb;
  const uint32_t r = blk / tr.njb;
  t.t2 = tr.t2_0 + r % tr.nt2;
  t.b = r / tr.nt2;
  return t;
}
static inline TileRange full_tile_range(int N, int Np) { return TileRange{0u, (uint32_t)(N + 1) >> 1, 0u, (uint32_t)Np >> 6}; }

// Addressing of the a / b operand tile images (elements).  One GPU: [b][ch][kb][Np rows][64].  Row-sharded runs cut the
// image into contiguous blocks so that an in-place exchange assembles it without packing kernels: by the producing
// rank's rows (outgoing update, RL = rows per rank) or k-blocks (incoming update, KBL = k-blocks per rank), and by
// channel groups of CG = 128 / world channels, the unit the channel-sharded contraction consumes:
//   [channel group][row block][k block] [b][CG][KBL][RL][64]
// Rank s writes blocks (d, s) for every channel group d; the all-to-all moves block (d, s) to rank d, which then holds
// its CG channels for all rows and all k.  CG = 128, one block: the plain image.
struct ImgLayout {
  uint32_t RL, KBL, nkblk, CG, nsrc;
  size_t blk;
};
static inline ImgLayout img_layout(int B, int Np, int row_split, int k_split, int cg_split = 1) {
  ImgLayout L;
  L.RL = (uint32_t)Np / row_split;
  L.KBL = (uint32_t)(Np >> 6) / k_split;
  L.nkblk = k_split;
  L.CG = 128u / cg_split;
  L.nsrc = (uint32_t)(row_split * k_split);
  L.blk = (size_t)B * L.CG * L.KBL * L.RL * 64;
  return L;
}
__device__ __forceinline__ size_t img_off(const ImgLayout& L, uint32_t b, uint32_t ch, uint32_t kb, uint32_t row) {
  const uint32_t rb = row / L.RL, kq = kb / L.KBL, cg = ch / L.CG;
  return ((size_t)cg * L.nsrc + rb * L.nkblk + kq) * L.blk +
         ((((size_t)b * L.CG + (ch - cg * L.CG)) * L.KBL + (kb - kq * L.KBL)) * L.RL + (row - rb * L.RL)) * 64;
}

// Row-sharded runs read blocks other ranks produced straight from those ranks' memory (peer-mapped over NVLink, shard.cu):
// p[r] = rank r's copy of the buffer; one GPU: p[0].
struct PeerImg { const __nv_bfloat16* p[8]; };
// p[r] with r known only at run time, through static indices: indexing the kernel parameter dynamically makes the compiler
// copy the array to local memory at kernel entry, for every thread
__device__ __forceinline__ const __nv_bfloat16* peer_ptr(const PeerImg& im, uint32_t r) {
  const __nv_bfloat16* q = im.p[0];
#pragma unroll
  for (uint32_t k = 1; k < 8; ++k) q = (r == k) ? im.p[k] : q;
  return q;
}

// Addressing of the contraction result x (elements of the 128-byte row of 64 columns j of channel ch, image row i, column
// block jb64).  One GPU: pixel-tile images [b][i / 2][jb64][i & 1][128 ch][64] (the 32 KB of one tile are K6's MN-major A
// operand).  Channel-sharded: [row block s][channel group d][b][local i / 2][jb64][i & 1][CG][64]: rank d (owner of
// channel group d) writes blocks (s, d) for every row block s, the all-to-all returns block (s, d) to rank s.
struct XLayout {
  uint32_t RL, CG, W, njb;     // rows per rank (Np when unsharded), channels per group, world, Np / 64
  size_t blk;
};
static inline XLayout x_layout(int B, int Np, int world) {
  XLayout L;
  L.RL = (uint32_t)Np / world;
  L.CG = 128u / world;
  L.W = (uint32_t)world;
  L.njb = (uint32_t)Np >> 6;
  L.blk = (size_t)B * (L.RL >> 1) * L.njb * 2 * L.CG * 64;
  return L;
}
__device__ __forceinline__ size_t x_off(const XLayout& L, uint32_t b, uint32_t i, uint32_t jb64, uint32_t ch) {
  const uint32_t rb = i / L.RL, il = i - rb * L.RL, cg = ch / L.CG;
  return ((size_t)rb * L.W + cg) * L.blk + (((((size_t)b * (L.RL >> 1) + (il >> 1)) * L.njb + jb64) * 2 + (il & 1)) * L.CG + (ch - cg * L.CG)) * 64;
}


}  // namespace tc
