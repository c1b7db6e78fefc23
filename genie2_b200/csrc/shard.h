// Row sharding of one long sample across the GPUs of a box: NCCL plumbing.
#pragma once
#include "ctx.h"

// In-place all-gather on the caller's stream: rank r's piece lies at base + r * piece_bytes.
int shard_all_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st);
// In-place all-to-all on a world x world grid of blocks: block (d, s) at base + (d * world + s) * block_bytes is sent by
// rank s (which wrote column s) to the same position on rank d (which ends up with the whole row d).  base2: a second
// buffer of the same layout exchanged in the same NCCL group (one kernel, one rendezvous).
int shard_all_to_all(gb_ctx* c, void* base, size_t block_bytes, cudaStream_t st, void* base2 = nullptr);
void shard_destroy(gb_ctx* c);
