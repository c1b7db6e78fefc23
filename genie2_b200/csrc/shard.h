// Row sharding of one long sample across the GPUs of a box: communicator, peer-mapped memory,
// device-side barriers and the copy kernels that pull the peers' blocks.
#pragma once
#include "ctx.h"

struct gb_peer_flags { unsigned* f[8]; };

// In-place NCCL all-gather on the caller's stream: rank r's piece lies at base + r * piece_bytes (setup and the optional
// aux.p output only; nothing on the path of an eval).
int shard_all_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st);
// Peer-mapped workspaces (CUDA IPC, one process per GPU of one box): after gb_set_features allocated the workspace every
// rank maps every other rank's (collective); kernels then read their operands straight from the producer's memory over
// NVLink.  shard_unmap_workspace (collective: it ends with a host barrier) must run before the workspace is freed.
int shard_map_workspace(gb_ctx* c);
int shard_unmap_workspace(gb_ctx* c, bool barrier);
// Device-side barrier of all ranks on the caller's stream (flags in peer-mapped memory, system-scope release / acquire):
// everything the ranks wrote before it is visible to peer reads after it.  No host involvement, no NCCL kernel.
int shard_barrier(gb_ctx* c, cudaStream_t st);
// Copy blocks (me, s), s != me, of one or two world x world grids of blocks (block (d, s) at base + (d * world + s) *
// block_bytes: written by rank s, wanted by rank d) out of rank s's workspace into this rank's over NVLink (one kernel).  Put a shard_barrier between the producers and this.
int shard_pull_blocks(gb_ctx* c, void* base, void* base2, size_t block_bytes, cudaStream_t st);
// In-place all-gather by pulling (rank s's piece at base + s * piece_bytes of every rank's workspace); same rule.
int shard_pull_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st);
void shard_destroy(gb_ctx* c);
