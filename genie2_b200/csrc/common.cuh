// Shared device helpers for the genie2_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GB_WARP 32

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// 3x3 helpers, row-major r[9]
__device__ __forceinline__ void mat3_mul(const float* a, const float* b, float* c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      c[i * 3 + j] = a[i * 3 + 0] * b[0 * 3 + j] + a[i * 3 + 1] * b[1 * 3 + j] + a[i * 3 + 2] * b[2 * 3 + j];
}

// Unit quaternion (w,x,y,z) of a rotation matrix: Shepperd's branch on the largest of the
// four squared components, sign convention w >= 0.  Same eigenvector (up to sign) as the
// reference's eigh of the 4x4 K matrix (utils/affine_utils.py:336-355).
__device__ __forceinline__ void rot_to_quat_canonical(const float* r, float* q) {
  const float xx = r[0], xy = r[1], xz = r[2], yx = r[3], yy = r[4], yz = r[5], zx = r[6], zy = r[7], zz = r[8];
  const float d0 = 1.f + xx + yy + zz, d1 = 1.f + xx - yy - zz, d2 = 1.f - xx + yy - zz, d3 = 1.f - xx - yy + zz;
  float dm = d0;
  int idx = 0;
  if (d1 > dm) { dm = d1; idx = 1; }
  if (d2 > dm) { dm = d2; idx = 2; }
  if (d3 > dm) { dm = d3; idx = 3; }
  dm = fmaxf(dm, 1e-30f);
  const float s = 0.5f / sqrtf(dm);
  float w, x, y, z;
  if (idx == 0)      { w = dm;      x = zy - yz; y = xz - zx; z = yx - xy; }
  else if (idx == 1) { w = zy - yz; x = dm;      y = xy + yx; z = xz + zx; }
  else if (idx == 2) { w = xz - zx; x = xy + yx; y = dm;      z = yz + zy; }
  else               { w = yx - xy; x = xz + zx; y = yz + zy; z = dm; }
  const float sg = (w < 0.f) ? -s : s;
  q[0] = w * sg; q[1] = x * sg; q[2] = y * sg; q[3] = z * sg;
}

// (a,b,c,d) -> R, the table at utils/affine_utils.py:299-334 written out.
__device__ __forceinline__ void quat_to_rot(float a, float b, float c, float d, float* r) {
  r[0] = a * a + b * b - c * c - d * d; r[1] = 2.f * (b * c - a * d);         r[2] = 2.f * (b * d + a * c);
  r[3] = 2.f * (b * c + a * d);         r[4] = a * a - b * b + c * c - d * d; r[5] = 2.f * (c * d - a * b);
  r[6] = 2.f * (b * d - a * c);         r[7] = 2.f * (c * d + a * b);         r[8] = a * a - b * b - c * c + d * d;
}

// Philox4x32-10 counter RNG (Salmon et al. 2011) + Box-Muller: the in-kernel N(0,1) source
// of the reverse step when the caller supplies no noise buffer.
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t* out) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// three independent N(0,1) draws keyed by (seed, step, element)
__device__ __forceinline__ void normal3(uint64_t seed, uint32_t step, uint32_t elem, float* n) {
  uint32_t r[4];
  philox4x32_10(elem, step, 0x47454e49u, 0x45320000u, (uint32_t)seed, (uint32_t)(seed >> 32), r);
  const float u0 = ((float)r[0] + 0.5f) * 2.3283064365386963e-10f, u1 = ((float)r[1] + 0.5f) * 2.3283064365386963e-10f;
  const float u2 = ((float)r[2] + 0.5f) * 2.3283064365386963e-10f, u3 = ((float)r[3] + 0.5f) * 2.3283064365386963e-10f;
  const float ra = sqrtf(-2.f * logf(u0)), rb = sqrtf(-2.f * logf(u2));
  float s, c;
  sincosf(6.283185307179586f * u1, &s, &c);
  n[0] = ra * c; n[1] = ra * s;
  n[2] = rb * cosf(6.283185307179586f * u3);
}
