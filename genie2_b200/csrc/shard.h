// Row sharding of one long sample across the GPUs of a box: NCCL plumbing.
#pragma once
#include "ctx.h"

// In-place all-gather on the caller's stream: rank r's piece lies at base + r * piece_bytes.
int shard_all_gather(gb_ctx* c, void* base, size_t piece_bytes, cudaStream_t st);
void shard_destroy(gb_ctx* c);
