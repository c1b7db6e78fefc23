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

#include <dlfcn.h>
#include <string.h>

namespace {

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
  ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
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
  a.Send = (decltype(a.Send))dlsym(h, "ncclSend");
  a.Recv = (decltype(a.Recv))dlsym(h, "ncclRecv");
  a.GroupStart = (decltype(a.GroupStart))dlsym(h, "ncclGroupStart");
  a.GroupEnd = (decltype(a.GroupEnd))dlsym(h, "ncclGroupEnd");
  a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
  a.GetVersion = (decltype(a.GetVersion))dlsym(h, "ncclGetVersion");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.GetErrorString || !a.Send || !a.Recv || !a.GroupStart ||
      !a.GroupEnd) return "libnccl: missing symbols";
  g_nccl = a;
  return nullptr;
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
  if (world < 1 || rank < 0 || rank >= world) return gb_fail(c, GB_ERR_INVALID, "gb_shard_init: bad rank %d of %d", rank, world);
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
  return GB_OK;
}

extern "C" int gb_shard_info(const gb_ctx* c, int* world, int* rank) {
  if (!c) return GB_ERR_INVALID;
  if (world) *world = c->shard_world;
  if (rank) *rank = c->shard_rank;
  return GB_OK;
}

void shard_destroy(gb_ctx* c) {
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

int shard_all_to_all(gb_ctx* c, void* base, size_t block_bytes, cudaStream_t st, void* base2) {
  const int W = c->shard_world, me = c->shard_rank;
  if (W <= 1) return GB_OK;
  if (!c->shard_comm) return gb_fail(c, GB_ERR_INVALID, "shard_all_to_all: no communicator");
  ncclComm_t comm = (ncclComm_t)c->shard_comm;
  char* bases[2] = {(char*)base, (char*)base2};
  ncclResult_t r = g_nccl.GroupStart();
  for (int i = 0; i < 2; ++i) {
    char* b = bases[i];
    if (!b) continue;
    for (int peer = 0; peer < W && r == 0; ++peer) {
      if (peer == me) continue;
      r = g_nccl.Send(b + ((size_t)peer * W + me) * block_bytes, block_bytes, ncclUint8, peer, comm, st);
      if (r == 0) r = g_nccl.Recv(b + ((size_t)me * W + peer) * block_bytes, block_bytes, ncclUint8, peer, comm, st);
    }
  }
  const ncclResult_t e = g_nccl.GroupEnd();
  if (r == 0) r = e;
  if (r != 0) return gb_fail(c, GB_ERR_CUDA, "all-to-all (ncclSend / ncclRecv): %s", g_nccl.GetErrorString(r));
  c->collectives++;
  return GB_OK;
}
