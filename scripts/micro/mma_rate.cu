// Microbenchmark: issue rate of tcgen05.mma (kind::f16, M=128) for SS (A, B in shared memory) and
// TS (A in tensor memory) operand modes and N = 64 / 128 / 256.  One CTA per SM, one issuing warp.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I genie2_b200/csrc scripts/micro/mma_rate.cu -o /tmp/mma_rate
#include <stdio.h>
#include "tc_common.cuh"
using namespace tc;

__global__ void __launch_bounds__(128, 1) mma_rate(int mode, int N, int reps, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = smem_base_1024(raw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 196608);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar + 1);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 196608 / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(tslot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = warp_uniform(*tslot);
  if (warp == 0) {
    const uint32_t a = smem_u32(sm), b = smem_u32(sm) + 65536;
    const uint32_t idesc = umma_idesc(128, N);
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (elect_one()) {
#pragma unroll
        for (uint32_t ks = 0; ks < 8; ++ks) {
          const uint64_t bd = umma_desc(b + (ks >> 2) * (uint32_t)(N * 128) + (ks & 3u) * 32u);
          if (mode == 0) umma(tm + 256, umma_desc(a + (ks >> 2) * 16384u + (ks & 3u) * 32u), bd, idesc, 1);
          else umma_ts(tm + 256, tm + ks * 8u, bd, idesc, 1);
        }
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar);
    __syncwarp();
    mbar_wait(bar, 0);
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int grid : {1, 148})
    for (int mode = 0; mode < 2; ++mode)
      for (int N : {64, 128, 256}) {
        const int reps = 500;
        mma_rate<<<grid, 128, 198 * 1024>>>(mode, N, reps, d);
        long long h = 0;
        cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("grid %3d %s N=%3d : %.1f cycles per MMA (128 x %d x 16)  [%s]\n", grid, mode ? "TS" : "SS", N, (double)h / (reps * 8), N,
               cudaGetErrorString(e));
      }
  return 0;
}
