// Internal context of the genie2_b200 library (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/genie2_b200.h"

struct HostTensor {
  std::vector<float> v;
  std::vector<int64_t> shape;
};

struct TriMulW {
  float *ln_in_g, *ln_in_b, *ln_out_g, *ln_out_b;
  float *Wpk, *bpk;   // [5][128][128], [5][128]   (fp32 SIMT packing, see kernels_pair_simt.cuh)
  float *WzT, *bz;    // [128 k][128 n], [128]
  // tensor-core packings (K-major [n][k] rows, i.e. nn.Linear layout), filled by tc_pack
  void *tc_Wproj, *tc_Wz;
  float *tc_bproj, *tc_uv;   // tc_uv: [128] u | [128] v of the LayerNorm-folded output projection
  // GB_PREC_BF16X3: hi | lo images of the same (unscaled) weights, see tc_proj_x3.cu / tc_pair.cu
  void *x3_Wproj_hi, *x3_Wproj_lo, *x3_Wz_hi, *x3_Wz_lo;
  float *x3_bproj, *x3_uv;
};
struct TransitionW {
  float *ln_g, *ln_b, *W1T, *b1, *W2T, *b2;
  void *tc_W1, *tc_W2;   // bf16 chunk images; W1 carries the LayerNorm gamma
  float* tc_b1;          // b1 + W1 beta
  void *x3_W1, *x3_W2;   // GB_PREC_BF16X3: 64-wide hidden chunks, each block hi 16K | lo 16K (tc_transition_x3.cu)
};
struct TriAttW {
  float *ln_g, *ln_b, *Wbias /*[H][128]*/, *Wqkvg /*[4*H*C][128]: q|k|v|g*/, *bg, *Wo /*[128][H*C]*/, *bo, *bqkvg /*[4*H*C]: 0|0|0|bg*/;
  void *tc_Wqkvg = nullptr, *tc_Wo = nullptr;   // bf16 operand images of the tensor-core kernels (tc_triatt.cu)
  void *x3_Wqkvg = nullptr, *x3_Wo = nullptr;   // hi | lo image pairs (GB_PREC_BF16X3)
};
struct PairLayerW {
  TriMulW mul_out, mul_in;
  TriAttW att_start, att_end;
  TransitionW tr;
};
struct StructLayerW {
  float *Wqkv, *bqkv;           // [1152][c_s], [1152]
  float *Wb, *bb, *hw;          // [12][128], [12], [12]
  float *Wout, *bout;           // [c_s][2112]
  float *ln1_g, *ln1_b;
  std::vector<float*> W1, b1, W2, b2, W3, b3;
  float *ln2_g, *ln2_b;
  float *Wbb, *bbb;             // [6][c_s], [6]
};

struct gb_ctx {
  gb_dims d;
  int device = 0;
  int precision = GB_PREC_FP32;
  mutable std::string err;
  int64_t launches = 0;

  std::unordered_map<std::string, HostTensor> host;   // state dict as loaded
  std::vector<void*> weight_allocs;                    // device buffers owned by the weights
  bool weights_loaded = false, schedule_set = false, features_set = false;

  // packed weights
  float *W_single = nullptr;      // [c_s][855] raw
  float *time_table = nullptr;    // [T+1][c_s]
  float *W_pipj = nullptr;        // [2*c_p][c_s]
  float *WrelT = nullptr, *WtT = nullptr, *WmT = nullptr;
  std::vector<PairLayerW> pair;
  std::vector<StructLayerW> st;

  // schedule tables [T+1]
  float *alphas = nullptr, *sqrt_alphas = nullptr, *sqrt_omac = nullptr, *sqrt_betas = nullptr;   // alphas owns the block

  // bound features
  int B = 0, N = 0, has_motif = 0;
  void* feat_alloc = nullptr;
  int *residue_index = nullptr, *chain_index = nullptr, *mask = nullptr, *aatype = nullptr;
  uint8_t *fsm = nullptr, *fst = nullptr, *iface = nullptr;
  float *motif_pos = nullptr, *s_static = nullptr;

  // workspace
  void* ws_alloc = nullptr;
  size_t ws_bytes = 0;
  size_t feat_cap = 0, ws_cap = 0;   // bytes behind feat_alloc / ws_alloc (grow-only across gb_set_features calls)
  float *s = nullptr, *sA = nullptr, *sB = nullptr, *h1 = nullptr, *h2 = nullptr, *pipj = nullptr;
  float* sX[2] = {nullptr, nullptr};
  float *p = nullptr, *a = nullptr, *b = nullptr, *x = nullptr, *g = nullptr;
  float *ta_x = nullptr, *ta_bias = nullptr;   // triangle attention: LN(p) / gated o [M][128], bias [M][4]
  float *qkv = nullptr, *cat = nullptr, *rots_w = nullptr, *trans_w = nullptr, *z = nullptr;
  float *trans_io = nullptr, *rots_io = nullptr;   // gb_sample_host state
  int* t_dev = nullptr;                            // [B] timesteps
  // gb_sample_loop: one reverse step captured as a CUDA graph and replayed; the step index lives on the device
  int* step_dev = nullptr;                         // [1] step of the next replay, decremented by the graph itself
  void* step_graph = nullptr;                      // cudaGraphExec_t
  int64_t step_launches = 0;                       // kernels in one captured step
  void* cap_stream = nullptr;                      // private stream the capture runs on
  struct StepKey {
    const void *trans = nullptr, *rots = nullptr, *tape = nullptr;
    float scale = 0.f;
    uint64_t seed = 0, epoch = 0;
    int precision = -1;
    bool operator==(const StepKey& o) const {
      return trans == o.trans && rots == o.rots && tape == o.tape && scale == o.scale && seed == o.seed && epoch == o.epoch && precision == o.precision;
    }
  } step_key, last_key;                            // key of the captured graph / of the previous gb_sample_loop call
  uint64_t epoch = 0;                              // bumped by everything that invalidates a captured step (features, weights, schedule, sharding)
  void* tape_alloc = nullptr;
  size_t tape_bytes = 0;
  // row sharding of one long sample over `shard_world` GPUs (shard.cu); 1 = off
  int shard_world = 1, shard_rank = 0;
  void* shard_comm = nullptr;   // ncclComm_t
  // peer-mapped memory of the other ranks (CUDA IPC; shard.cu): peer_ws[r] = rank r's workspace allocation as seen from this
  // GPU (own entry = ws_alloc), peer_flags[r] = rank r's barrier flags (own entry = flag_alloc)
  void* peer_ws[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  unsigned* peer_flags[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  void* flag_alloc = nullptr;
  long long peer_barriers = 0;
  int64_t collectives = 0;      // NCCL collectives enqueued since creation
  // tensor-core scratch (bf16/tf32 operand copies), owned by the tc module
  void* tc_state = nullptr;
  // device address of the file-scope static scratch of gb_k_ipa_static (api.cu), resolved once in gb_create
  float* static_scratch = nullptr;
};

// GB_PHASE_TIMING=1: gb_denoise prints, per call, the device time between consecutive marks summed by name (stderr; the
// eval is synchronised for it -- a diagnosis switch, not a mode).  ph_mark closes the running phase and opens `name`.
void ph_mark(cudaStream_t st, const char* name);
void ph_report(const gb_ctx* ctx, cudaStream_t st);

// set ctx->err (printf style) and return code
int gb_fail(const gb_ctx* ctx, int code, const char* fmt, ...);

#define GB_CUDA(ctx, call)                                                                          \
  do {                                                                                              \
    cudaError_t e_ = (call);                                                                        \
    if (e_ != cudaSuccess) return gb_fail(ctx, GB_ERR_CUDA, "%s failed: %s (%s:%d)", #call,        \
                                          cudaGetErrorString(e_), __FILE__, __LINE__);             \
  } while (0)

#define GB_LAUNCH_CHECK(ctx, name)                                                                  \
  do {                                                                                              \
    (ctx)->launches++;                                                                              \
    cudaError_t e_ = cudaGetLastError();                                                            \
    if (e_ != cudaSuccess) return gb_fail(ctx, GB_ERR_CUDA, "launch of %s failed: %s", name,       \
                                          cudaGetErrorString(e_));                                 \
  } while (0)
