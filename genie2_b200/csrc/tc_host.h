// Host-side helpers of the tensor-core modules: bf16 rounding and operand tile-image packing.
#pragma once
#include <stdlib.h>
#include <string.h>

#include "ctx.h"
#include "tc_common.cuh"

static uint16_t f2bf(float f) {   // round to nearest even
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

static float bf_round(float f) {
  const uint32_t b = (uint32_t)f2bf(f) << 16;
  float r;
  memcpy(&r, &b, 4);
  return r;
}
// the part of f a bf16 cannot hold: f = bf_round(f) + bf_lo(f) to 16 mantissa bits (GB_PREC_BF16X3 weight images)
static float bf_lo(float f) { return f - bf_round(f); }

// rows x K (K % 64 == 0) fp32 matrix given by at(r, k) -> tile images [K/64][rows][64]
template <typename F>
static void pack_image(uint16_t* dst, int rows, int K, F at) {
  for (int kb = 0; kb < K / 64; ++kb)
    for (int r = 0; r < rows; ++r)
      for (int k = 0; k < 64; ++k)
        dst[(size_t)kb * rows * 64 + tc::sw128_off(r, k) / 2] =
            f2bf(at(r, kb * 64 + k));
}

static void* upload_bytes(gb_ctx* c, const void* src, size_t bytes) {
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr;
  if (cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(p); return nullptr; }
  c->weight_allocs.push_back(p);
  return p;
}


// Launch with the programmatic-stream-serialization attribute: the kernel may start while its predecessor in the
// stream is still running and synchronises with it through tc::pdl_wait().  Only for kernels that call pdl_wait()
// before touching anything the predecessor writes or reads.  GB_NO_PDL=1 turns the attribute off (plain stream order).
template <typename... KArgs, typename... Args>
static inline void launch_pdl_cls(int cls, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  static const int mask = [] { const char* e = getenv("GB_PDL_MASK"); return e ? atoi(e) : 0xff; }();
  const bool off = !((mask >> cls) & 1);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = off ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);   // errors surface through cudaGetLastError in GB_LAUNCH_CHECK
}

template <typename... KArgs, typename... Args>
static inline void launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  launch_pdl_cls(0, kern, grid, block, smem, st, static_cast<Args&&>(args)...);
}
