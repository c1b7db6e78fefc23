// Host-side helpers of the tensor-core modules: bf16 rounding and operand tile-image packing.
#pragma once
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

