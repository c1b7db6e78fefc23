// Row sharding of one long sample (B = 1) over `world` GPUs of one box: NCCL communicator owned by the
// context, in-place exchanges enqueued on the caller's stream between the kernels of gb_denoise.
//
// What is sharded (rows [rank * N / world, (rank + 1) * N / world) of the pair map p on each rank):
//   pair embedding, K4 / K6 of both triangle updates, the pair transition, the IPA attention and its
//   output projection.  The contraction K5 is sharded by CHANNEL instead (x^c = a^c b^cT is independent
//   per channel): rank d contracts channels [d * 128 / world, (d + 1) * 128 / world) for all rows.
//   What is exchanged:
//     each triangle update   all-to-all of the a and b operand images (one NCCL group: every rank ends up
//                            with its channel group for all rows and all k), all-to-all of x back to the
//                            row owners: 3 (W - 1) / W^2 of an image (N^2 * 128 * 2 bytes) received per
//                            rank, against (W - 1) / W .. (W^2 - 1) / W^2 for the all-gather of b (+ the
//                            all-to-all of a) that row-sharding the contraction needs
//     each IPA layer         all-gather of the [N, c_s] fp32 result (1.5 MB at N = 1024)
//   p itself never moves (only when the caller asks for aux.p).  The operand images and x are laid out
//   as W x W grids of contiguous blocks (tc_common.cuh, ImgLayout / XLayout): block (dest, source) is
//   written in place by the source rank and travels as is, so there are no packing kernels.  The same
//   kernels run on the same tiles in the same k order as on one GPU, so the sharded result is
//   bit-identical to the single-GPU one.
//
// libnccl is resolved at run time (dlopen) so that single-GPU users need no NCCL at all; inside a
// torch process this picks up the libnccl.so.2 torch already loaded.
// Reference: there is no counterpart (the reference runs one sample per GPU, multiprocessor.py:73-100).
#include "shard.h"
#include "tc_common.cuh"

#include <dlfcn.h>
#include <string.h>

namespace {
using namespace tc;

// the slice of the NCCL ABI used here (nccl.h, stable since 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclUint8 = 1 };

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

NcclApi g_nccl;

const char* load_nccl() {
  if (g_nccl.handle) return nullptr;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return "libnccl.so.2 not found (dlopen)";
  NcclApi a;
  a.handle = h;
  a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
  a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
  a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
  a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
  a.GetVersion = (decltype(a.GetVersion))dlsym(h, "ncclGetVersion");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.GetErrorString) return "libnccl: missing symbols";
  g_nccl = a;
  return nullptr;
}

constexpr size_t FLAG_BYTES = 2u << 20;     // an allocation of its own (IPC handles share whole allocations)

// All-gather of n bytes per rank between the hosts, through the communicator (setup only).
int exchange_bytes(gb_ctx* c, const void* mine, void* all, size_t n) {
  const int W = c->shard_world;
  char* d = nullptr;
  GB_CUDA(c, cudaMalloc((void**)&d, n * W));
  cudaError_t e = cudaMemcpy(d + (size_t)c->shard_rank * n, mine, n, cudaMemcpyHostToDevice);
  ncclResult_t r = 0;
  if (e == cudaSuccess) r = g_nccl.AllGather(d + (size_t)c->shard_rank * n, d, n, ncclUint8, (ncclComm_t)c->shard_comm, (cudaStream_t)0);
  if (e == cudaSuccess && r == 0) e = cudaStreamSynchronize((cudaStream_t)0);
  if (e == cudaSuccess && r == 0) e = cudaMemcpy(all, d, n * W, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (r != 0) return gb_fail(c, GB_ERR_CUDA, "handle exchange (ncclAllGather): %s", g_nccl.GetErrorString(r));
  if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "handle exchange: %s", cudaGetErrorString(e));
  return GB_OK;
}

// Map the allocation `local` of every rank into this process: out[r] = rank r's, out[me] = local.  Collective.
int map_peers(gb_ctx* c, void* local, void** out) {
  const int W = c->shard_world, me = c->shard_rank;
  cudaIpcMemHandle_t mine, all[8];
  GB_CUDA(c, cudaIpcGetMemHandle(&mine, local));
  const int rc = exchange_bytes(c, &mine, all, sizeof(mine));
  if (rc != GB_OK) return rc;
  for (int r = 0; r < W; ++r) {
    out[r] = local;
    if (r == me) continue;
    const cudaError_t e = cudaIpcOpenMemHandle(&out[r], all[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int q = 0; q < r; ++q) if (q != me) cudaIpcCloseMemHandle(out[q]);
      return gb_fail(c, GB_ERR_CUDA, "cudaIpcOpenMemHandle (rank %d's memory): %s -- row sharding needs peer access between the GPUs of one box",
                     r, cudaGetErrorString(e));
    }
  }
  return GB_OK;
}

// One warp; lane r signals rank r (its flags[me] = epoch, release at system scope: after this GPU's earlier writes) and waits
// for rank r's signal (own flags[r] >= epoch, acquire).  A rank that never arrives traps after ~10 s instead of hanging the GPU.
__global__ void peer_barrier_kernel(gb_peer_flags pf, int W, int me) {
  __shared__ unsigned epoch_s;
  unsigned* mine = pf.f[me];
  if (threadIdx.x == 0) { epoch_s = mine[8] + 1u; mine[8] = epoch_s; }
  __syncthreads();
  const unsigned epoch = epoch_s;
  const int r = threadIdx.x;
  if (r >= W || r == me) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pf.f[r] + me), "r"(epoch) : "memory");
  const long long t0 = clock64();
  unsigned v;
  for (;;) {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + r) : "memory");
    if ((int)(v - epoch) >= 0) break;
    if (clock64() - t0 > 20000000000LL) __trap();
  }
}

// Pull n blocks of `bytes` each out of peer memory into local memory: peer -> shared -> local, bulk copies both ways, one
// issuing thread per CTA, a ring of PULL_STAGES chunks of PULL_CHUNK bytes per CTA with the loads four chunks ahead (296 CTAs x
// 64 KB = 19 MB in flight against the ~2 MB that NVLink's bandwidth-delay product asks for).  Chunks are dealt round-robin over the blocks so that every
// peer link is busy from the start.
constexpr uint32_t PULL_CHUNK = 16384, PULL_STAGES = 6;
constexpr size_t PULL_SMEM = PULL_STAGES * PULL_CHUNK + 1024 + 128;
struct PullJob {
  const char* src[14];
  char* dst[14];
  int n;
  size_t bytes;
};

__global__ void __launch_bounds__(32, 1) peer_pull_kernel(const PullJob job) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_base_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + PULL_STAGES * PULL_CHUNK);
  if (threadIdx.x != 0) return;
  for (uint32_t i = 0; i < PULL_STAGES; ++i) mbar_init(full + i, 1);
  fence_mbar_init();
  const size_t per = job.bytes / PULL_CHUNK, total = per * (size_t)job.n;
  const size_t first = blockIdx.x, step = gridDim.x;
  const size_t mine = first < total ? (total - first + step - 1) / step : 0;
  auto issue = [&](size_t i) {
    const size_t ch = first + i * step;
    const uint32_t blk = (uint32_t)(ch % (size_t)job.n), slot = (uint32_t)(i % PULL_STAGES);
    mbar_expect_tx(full + slot, PULL_CHUNK);
    bulk_g2s(sm + slot * PULL_CHUNK, job.src[blk] + (ch / (size_t)job.n) * PULL_CHUNK, PULL_CHUNK, full + slot);
  };
  // loads run PULL_STAGES - 2 chunks ahead: the slot refilled at step i held chunk i - 2, whose store has been read by then
  for (size_t i = 0; i < PULL_STAGES - 2 && i < mine; ++i) issue(i);
  for (size_t i = 0; i < mine; ++i) {
    const size_t nxt = i + PULL_STAGES - 2;
    if (nxt < mine) {
      bulk_wait_read<1>();
      issue(nxt);
    }
    const uint32_t slot = (uint32_t)(i % PULL_STAGES);
    mbar_wait(full + slot, (uint32_t)(i / PULL_STAGES) & 1u);
    const size_t ch = first + i * step;
    bulk_s2g(job.dst[ch % (size_t)job.n] + (ch / (size_t)job.n) * PULL_CHUNK, sm + slot * PULL_CHUNK, PULL_CHUNK);
  }
  bulk_wait_all<0>();
}

}  // namespace

extern "C" int gb_shard_unique_id(void* out128) {
  if (!out128) return gb_fail(nullptr, GB_ERR_INVALID, "gb_shard_unique_id: NULL pointer");
  if (const char* e = load_nccl()) return gb_fail(nullptr, GB_ERR_MISSING, "gb_shard_unique_id: %s", e);
  ncclUniqueId id;
  const ncclResult_t r = g_nccl.GetUniqueId(&id);
  if (r != 0) return gb_fail(nullptr, GB_ERR_CUDA, "ncclGetUniqueId: %s", g_nccl.GetErrorString(r));
  memcpy(out128, &id, sizeof(id));
  return GB_OK;
}

extern "C" int gb_shard_init(gb_ctx* c, int world, int rank, const void* id128) {
  if (!c) return GB_ERR_INVALID;
  if (world < 1 || world > 8 || rank < 0 || rank >= world) return gb_fail(c, GB_ERR_INVALID, "gb_shard_init: bad rank %d of %d (one box, at most 8 GPUs)", rank, world);
  if (c->features_set) return gb_fail(c, GB_ERR_INVALID, "gb_shard_init: call before gb_set_features");
  shard_destroy(c);
  if (world == 1) return GB_OK;
  if (!id128) return gb_fail(c, GB_ERR_INVALID, "gb_shard_init: NULL unique id");
  if (const char* e = load_nccl()) return gb_fail(c, GB_ERR_MISSING, "gb_shard_init: %s", e);
  GB_CUDA(c, cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclComm_t comm = nullptr;
  const ncclResult_t r = g_nccl.CommInitRank(&comm, world, id, rank);
  if (r != 0) return gb_fail(c, GB_ERR_CUDA, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
  c->shard_comm = comm;
  c->shard_world = world;
  c->shard_rank = rank;
  // barrier flags: flags[r] = last epoch rank r announced to this rank, [8] = this rank's own epoch counter
  GB_CUDA(c, cudaMalloc(&c->flag_alloc, FLAG_BYTES));
  GB_CUDA(c, cudaMemset(c->flag_alloc, 0, FLAG_BYTES));
  void* peers[8];
  const int rc = map_peers(c, c->flag_alloc, peers);
  if (rc != GB_OK) return rc;
  for (int r = 0; r < world; ++r) c->peer_flags[r] = (unsigned*)peers[r];
  return GB_OK;
}

extern "C" int gb_shard_info(const gb_ctx* c, int* world, int* rank) {
  if (!c) return GB_ERR_INVALID;
  if (world) *world = c->shard_world;
  if (rank) *rank = c->shard_rank;
  return GB_OK;
}

int shard_map_workspace(gb_ctx* c) {
  if (c->shard_world <= 1) return GB_OK;
  if (!c->shard_comm || !c->ws_alloc) return gb_fail(c, GB_ERR_INVALID, "shard_map_workspace: no communicator / workspace");
  return map_peers(c, c->ws_alloc, c->peer_ws);
}

int shard_unmap_workspace(gb_ctx* c, bool barrier) {
  const int W = c->shard_world;
  for (int r = 0; r < W; ++r) {
    if (c->peer_ws[r] && r != c->shard_rank) cudaIpcCloseMemHandle(c->peer_ws[r]);
    c->peer_ws[r] = nullptr;
  }
  if (barrier && W > 1 && c->shard_comm) {     // no rank frees its workspace before every rank has let go of it
    char one = 1, all[8];
    return exchange_bytes(c, &one, all, 1);
  }
  return GB_OK;
}

static int launch_pull(gb_ctx* c, const PullJob& job, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    GB_CUDA(c, cudaFuncSetAttribute(peer_pull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PULL_SMEM));
    attr_set = true;
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
  const size_t chunks = job.bytes / PULL_CHUNK * (size_t)job.n;
  const unsigned grid = (unsigned)(chunks < (size_t)(2 * sms) ? chunks : (size_t)(2 * sms));
  peer_pull_kernel<<<grid, 32, PULL_SMEM, st>>>(job);
  GB_LAUNCH_CHECK(c, "peer_pull");
  return GB_OK;
}

// Blocks (me, s), s != me, of two W x W grids of blocks (block (d, s) at base + (d * W + s) * block_bytes, written by rank s,
// wanted by rank d) are copied from rank s's memory into
// this rank's: after it this rank holds row `me` of both grids.  The caller has put a shard_barrier behind the producers.
int shard_pull_blocks(gb_ctx* c, void* base, void* base2, size_t block_bytes, cudaStream_t st) {
  const int W = c->shard_world, me = c->shard_rank;
  if (W <= 1) return GB_OK;
  if (block_bytes % PULL_CHUNK) return gb_fail(c, GB_ERR_INVALID, "shard_pull_blocks: block of %zu bytes", block_bytes);
  PullJob job;
  job.n = 0;
  job.bytes = block_bytes;
  for (void* b : {base, base2}) {
    if (!b) continue;
    const size_t off = (const char*)b - (const char*)c->ws_alloc;
    for (int k = 1; k < W; ++k) {
      const int peer = (me + k) % W;        // every rank starts on another peer
      if (!c->peer_ws[peer]) return gb_fail(c, GB_ERR_INVALID, "shard_pull_blocks: peer workspaces are not mapped");
      job.src[job.n] = (const char*)c->peer_ws[peer] + off + ((size_t)me * W + peer) * block_bytes;
      job.dst[job.n] = (char*)b + ((size_t)me * W + peer) * block_bytes;
      ++job.n;
    }
  }
  return launch_pull(c, job, st);
}

// In-place all-gather by pulling: rank s's piece lies at base + s * piece_bytes in every rank's workspace.
int shard_pull_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st) {
  const int W = c->shard_world, me = c->shard_rank;
  if (W <= 1) return GB_OK;
  if (piece_bytes % PULL_CHUNK) return gb_fail(c, GB_ERR_INVALID, "shard_pull_gather: piece of %zu bytes", piece_bytes);
  PullJob job;
  job.n = 0;
  job.bytes = piece_bytes;
  const size_t off = (const char*)base - (const char*)c->ws_alloc;
  for (int k = 1; k < W; ++k) {
    const int peer = (me + k) % W;
    if (!c->peer_ws[peer]) return gb_fail(c, GB_ERR_INVALID, "shard_pull_gather: peer workspaces are not mapped");
    job.src[job.n] = (const char*)c->peer_ws[peer] + off + (size_t)peer * piece_bytes;
    job.dst[job.n] = (char*)base + (size_t)peer * piece_bytes;
    ++job.n;
  }
  return launch_pull(c, job, st);
}

int shard_barrier(gb_ctx* c, cudaStream_t st) {
  const int W = c->shard_world;
  if (W <= 1) return GB_OK;
  gb_peer_flags pf;
  for (int r = 0; r < 8; ++r) pf.f[r] = c->peer_flags[r < W ? r : 0];
  if (!pf.f[0]) return gb_fail(c, GB_ERR_INVALID, "shard_barrier: flags not mapped");
  peer_barrier_kernel<<<1, 32, 0, st>>>(pf, W, c->shard_rank);
  GB_LAUNCH_CHECK(c, "peer_barrier");
  c->peer_barriers++;
  return GB_OK;
}

void shard_destroy(gb_ctx* c) {
  shard_unmap_workspace(c, false);
  for (int r = 0; r < c->shard_world; ++r) {
    if (c->peer_flags[r] && r != c->shard_rank) cudaIpcCloseMemHandle(c->peer_flags[r]);
    c->peer_flags[r] = nullptr;
  }
  if (c->flag_alloc) cudaFree(c->flag_alloc);
  c->flag_alloc = nullptr;
  if (c->shard_comm && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)c->shard_comm);
  c->shard_comm = nullptr;
  c->shard_world = 1;
  c->shard_rank = 0;
}

int shard_all_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st) {
  if (c->shard_world <= 1) return GB_OK;
  if (!c->shard_comm) return gb_fail(c, GB_ERR_INVALID, "shard_all_gather: no communicator");
  const ncclResult_t r = g_nccl.AllGather((const char*)base + (size_t)c->shard_rank * piece_bytes, base, piece_bytes, ncclUint8,
                                          (ncclComm_t)c->shard_comm, st);
  if (r != 0) return gb_fail(c, GB_ERR_CUDA, "ncclAllGather: %s", g_nccl.GetErrorString(r));
  c->collectives++;
  return GB_OK;
}
