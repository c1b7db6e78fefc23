// Microbenchmark: how fast can one B200 move a block out of / into a peer's memory over NVLink, by transport?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/p2p_bw scripts/p2p_bw.cu && scripts/p2p_bw [MB]
// One process, two GPUs with peer access enabled (the same NVLink path the row-sharded runs take through CUDA IPC).
// Each transport is timed one way (GPU 0 only) and both ways at once (GPU 0 and GPU 1 doing the same thing to each other,
// which is what the sharded exchange does).  Prints GB/s per direction.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// bulk copies both ways through a ring of ST chunks of CH bytes, loads ST - 2 ahead, one issuing thread per CTA
template <uint32_t CH, uint32_t ST>
__global__ void __launch_bounds__(32) tma_copy(const char* src, char* dst, size_t bytes) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + ST * CH);
  if (threadIdx.x) return;
  for (uint32_t i = 0; i < ST; ++i) mbar_init(full + i, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const size_t total = bytes / CH, first = blockIdx.x, step = gridDim.x;
  const size_t mine = first < total ? (total - first + step - 1) / step : 0;
  auto issue = [&](size_t i) {
    const uint32_t slot = (uint32_t)(i % ST);
    mbar_expect(full + slot, CH);
    g2s(sm + slot * CH, src + (first + i * step) * CH, CH, full + slot);
  };
  for (size_t i = 0; i < ST - 2 && i < mine; ++i) issue(i);
  for (size_t i = 0; i < mine; ++i) {
    if (i + ST - 2 < mine) {
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      issue(i + ST - 2);
    }
    const uint32_t slot = (uint32_t)(i % ST);
    mbar_wait(full + slot, (uint32_t)(i / ST) & 1u);
    s2g(dst + (first + i * step) * CH, sm + slot * CH, CH);
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// plain vector loads and stores, U x 16 bytes in flight per thread
template <int U>
__global__ void __launch_bounds__(512) ldg_copy(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < n16; i += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) dst[i + u * stride] = v[u];
  }
  for (; i < n16; i += stride) dst[i] = src[i];
}

struct Dev {
  int id;
  char *mine, *scratch;
  cudaStream_t st;
  cudaEvent_t e0, e1;
};

template <class F>
static void run(const char* name, Dev* d, size_t bytes, F launch) {
  for (int both = 0; both < 2; ++both) {
    const int nd = both ? 2 : 1;
    float best[2] = {1e30f, 1e30f};
    for (int rep = 0; rep < 6; ++rep) {
      for (int k = 0; k < nd; ++k) { CK(cudaSetDevice(d[k].id)); CK(cudaEventRecord(d[k].e0, d[k].st)); }
      for (int k = 0; k < nd; ++k) { CK(cudaSetDevice(d[k].id)); launch(d[k], d[1 - k]); CK(cudaEventRecord(d[k].e1, d[k].st)); }
      for (int k = 0; k < nd; ++k) {
        CK(cudaSetDevice(d[k].id));
        CK(cudaStreamSynchronize(d[k].st));
        float ms;
        CK(cudaEventElapsedTime(&ms, d[k].e0, d[k].e1));
        if (rep && ms < best[k]) best[k] = ms;
      }
    }
    if (both) printf("%-44s both ways: %7.1f / %7.1f GB/s per direction\n", name, bytes / best[0] / 1e6, bytes / best[1] / 1e6);
    else printf("%-44s one way  : %7.1f GB/s\n", name, bytes / best[0] / 1e6);
  }
}

int main(int argc, char** argv) {
  const size_t bytes = (size_t)(argc > 1 ? atoi(argv[1]) : 128) << 20;
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
  Dev d[2];
  for (int k = 0; k < 2; ++k) {
    d[k].id = k;
    CK(cudaSetDevice(k));
    CK(cudaDeviceEnablePeerAccess(1 - k, 0));
    CK(cudaMalloc(&d[k].mine, bytes));
    CK(cudaMalloc(&d[k].scratch, bytes));
    CK(cudaMemset(d[k].mine, k + 1, bytes));
    CK(cudaStreamCreate(&d[k].st));
    CK(cudaEventCreate(&d[k].e0));
    CK(cudaEventCreate(&d[k].e1));
    CK(cudaFuncSetAttribute(tma_copy<16384, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 16384 + 128));
    CK(cudaFuncSetAttribute(tma_copy<8192, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, 10 * 8192 + 128));
    CK(cudaFuncSetAttribute(tma_copy<32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768 + 128));
    CK(cudaFuncSetAttribute(tma_copy<4096, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * 4096 + 128));
  }
  printf("block of %zu MB between GPU 0 and GPU 1\n", bytes >> 20);
  const size_t n16 = bytes / 16;
  run("local copy, ldg x4 (HBM reference)", d, bytes, [&](Dev& me, Dev&) { ldg_copy<4><<<148 * 4, 512, 0, me.st>>>((const uint4*)me.mine, (uint4*)me.scratch, n16); });
  run("cudaMemcpyAsync pull (copy engine)", d, bytes, [&](Dev& me, Dev& peer) { CK(cudaMemcpyAsync(me.scratch, peer.mine, bytes, cudaMemcpyDeviceToDevice, me.st)); });
  run("cudaMemcpyAsync push (copy engine)", d, bytes, [&](Dev& me, Dev& peer) { CK(cudaMemcpyAsync(peer.scratch, me.mine, bytes, cudaMemcpyDeviceToDevice, me.st)); });
  run("pull, ldg x4, 592 CTAs x 512", d, bytes, [&](Dev& me, Dev& peer) { ldg_copy<4><<<148 * 4, 512, 0, me.st>>>((const uint4*)peer.mine, (uint4*)me.scratch, n16); });
  run("pull, ldg x8, 592 CTAs x 512", d, bytes, [&](Dev& me, Dev& peer) { ldg_copy<8><<<148 * 4, 512, 0, me.st>>>((const uint4*)peer.mine, (uint4*)me.scratch, n16); });
  run("pull, ldg x8, 148 CTAs x 512", d, bytes, [&](Dev& me, Dev& peer) { ldg_copy<8><<<148, 512, 0, me.st>>>((const uint4*)peer.mine, (uint4*)me.scratch, n16); });
  run("push, ldg local + stg peer x4, 592 CTAs", d, bytes, [&](Dev& me, Dev& peer) { ldg_copy<4><<<148 * 4, 512, 0, me.st>>>((const uint4*)me.mine, (uint4*)peer.scratch, n16); });
  run("pull, bulk 6 x 16 KB, 296 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<16384, 6><<<296, 32, 6 * 16384 + 128, me.st>>>(peer.mine, me.scratch, bytes); });
  run("pull, bulk 6 x 32 KB, 148 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<32768, 6><<<148, 32, 6 * 32768 + 128, me.st>>>(peer.mine, me.scratch, bytes); });
  run("pull, bulk 10 x 8 KB, 148 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<8192, 10><<<148, 32, 10 * 8192 + 128, me.st>>>(peer.mine, me.scratch, bytes); });
  run("pull, bulk 16 x 4 KB, 444 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<4096, 16><<<444, 32, 16 * 4096 + 128, me.st>>>(peer.mine, me.scratch, bytes); });
  run("push, bulk 6 x 16 KB, 296 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<16384, 6><<<296, 32, 6 * 16384 + 128, me.st>>>(me.mine, peer.scratch, bytes); });
  run("push, bulk 10 x 8 KB, 148 CTAs", d, bytes, [&](Dev& me, Dev& peer) { tma_copy<8192, 10><<<148, 32, 10 * 8192 + 128, me.st>>>(me.mine, peer.scratch, bytes); });
  // correctness spot check of the last pull variant
  CK(cudaSetDevice(0));
  tma_copy<8192, 10><<<148, 32, 10 * 8192 + 128, d[0].st>>>(d[1].mine, d[0].scratch, bytes);
  unsigned char h[2];
  CK(cudaMemcpy(h, d[0].scratch + bytes - 2, 2, cudaMemcpyDeviceToHost));
  printf("check: %d %d (expect 2 2)\n", h[0], h[1]);
  return 0;
}
