/*
 * genie2_b200 -- C ABI of the B200-native Genie 2 reverse-diffusion sampling hot path.
 *
 * This is the drop-in boundary: a flat extern "C" surface (plain pointers and sizes,
 * no torch / C++ types) that a reference-side binding (ctypes)
 * calls instead of the PyTorch-eager code paths cited next to each entry point.
 * Citations are file:line into the reference tree (marvinli00/genie2).
 *
 * Conventions
 *   - every function returns 0 on success or a negative gb_status; the message is
 *     available from gb_last_error().  Unsupported dimensions, NULL pointers and
 *     CUDA failures are errors, never a fallback (there is no CPU path).
 *   - all tensors are row-major, contiguous, fp32 unless stated, in the reference's
 *     own layouts: trans [B,N,3], rots [B,N,3,3], s [B,N,c_s], p [B,N,N,c_p].
 *   - pointers are DEVICE pointers unless the parameter is documented as "host or
 *     device" (copied with cudaMemcpyDefault) or the function name ends in _host.
 *   - a gb_ctx is bound to one device and is not thread-safe; work is enqueued on the
 *     caller's stream (a cudaStream_t passed as void*), asynchronously.
 *   - the context owns only packed weights and its workspace; results are written
 *     into caller-owned buffers.  Nothing is allocated after gb_set_features().
 */
#ifndef GENIE2_B200_H_
#define GENIE2_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GB_VERSION 200

typedef struct gb_ctx gb_ctx;

typedef enum gb_status {
  GB_OK = 0,
  GB_ERR_INVALID = -1,      /* bad argument / NULL pointer / wrong call order */
  GB_ERR_UNSUPPORTED = -2,  /* dimension or option outside the compiled specialisation */
  GB_ERR_CUDA = -3,         /* a CUDA runtime call or kernel failed */
  GB_ERR_MISSING = -4       /* a required weight tensor was not supplied */
} gb_status;

/* Arithmetic of the GEMM-shaped work (pair stack, IPA projections / o_pair, structure transitions).  LayerNorm
 * statistics, softmax, gating, residual adds, frames and the posterior step are fp32 in every mode, and the residual
 * stream p is stored in fp32 in every mode.
 *   FP32    fp32 FFMA on CUDA cores: the exact-parity path (1e-4 on every tensor Denoiser.forward returns).
 *   BF16X3  tcgen05.mma kind::f16 with SPLIT operands: every operand v is carried as hi = bf16(v), lo = bf16(v - hi)
 *           and every product is the three passes hi*hi + lo*hi + hi*lo accumulated in fp32 tensor memory (~2^-16
 *           per product); intermediates between kernels (a, b, x, H, activations) are stored as hi | lo image pairs
 *           or fp32.  This is the tensor-core mode that meets the 1e-3 per-eval / 0.1 A trajectory gates.
 *   BF16    tcgen05.mma kind::f16, plain bf16 operands and bf16 intermediates: fastest, ~1.6e-2 per eval.
 * GB_PREC_TF32 is the former name of slot 1 (declared, never built, in revision 100); it selects BF16X3. */
typedef enum gb_precision {
  GB_PREC_FP32 = 0,
  GB_PREC_BF16X3 = 1,
  GB_PREC_TF32 = 1,
  GB_PREC_BF16 = 2
} gb_precision;

/* Model dimensions: the keyword arguments of Denoiser.__init__ (genie/model/model.py:20-61)
 * as built at genie/diffusion/ddpm.py:26-31 from genie/config.py:41-80. */
typedef struct gb_dims {
  int32_t c_s, c_p, n_timestep;
  float rescale;
  int32_t c_pos_emb, c_chain_emb, c_timestep_emb, max_n_res, max_n_chain;
  int32_t relpos_k;
  float template_dist_min, template_dist_step;
  int32_t template_dist_n_bin;
  int32_t n_pair_transform_layer, include_mul_update, include_tri_att;
  int32_t c_hidden_mul, c_hidden_tri_att, n_head_tri, pair_transition_n;
  int32_t n_structure_layer, n_structure_block;
  int32_t c_hidden_ipa, n_head_ipa, n_qk_point, n_v_point;
  int32_t n_structure_transition_layer;
} gb_dims;

/* One named fp32 tensor of Denoiser.state_dict() (names exactly as the reference's,
 * e.g. "pair_transform_net.net.0.tri_mul_out.linear_a_p.weight").  data: host or device. */
typedef struct gb_tensor {
  const char* name;
  const float* data;
  int32_t ndim;
  int64_t shape[4];
} gb_tensor;

/* The batched, padded feature dictionary (genie/utils/feat_utils.py:304-321), the keys
 * the forward pass reads.  All pointers host or device; copied into the context. */
typedef struct gb_features {
  int32_t B, N;
  const int32_t* residue_index;        /* [B,N]    */
  const int32_t* chain_index;          /* [B,N]    */
  const int32_t* residue_mask;         /* [B,N]    */
  const uint8_t* fixed_sequence_mask;  /* [B,N]    bool */
  const uint8_t* fixed_structure_mask; /* [B,N,N]  bool */
  const uint8_t* interface_mask;       /* [B,N]    bool */
  const int32_t* aatype;               /* [B,N,20] */
  const float* atom_positions;         /* [B,N,3]  motif coordinates */
} gb_features;

/* Optional extra outputs / test-only inputs of gb_denoise; any member may be NULL.
 * These are the other keys Denoiser.forward returns (model.py:186-192). */
typedef struct gb_aux {
  float* s;                /* [B,N,c_s]                                   */
  float* p;                /* [B,N,N,c_p]                                 */
  float* states;           /* [1+n_layer*n_block,B,N,c_s]                 */
  float* out_rots;         /* [B,N,3,3] final frames ('ts')               */
  float* out_trans;        /* [B,N,3]                                     */
  const float* quat_follow;/* [B,N,N,4] TEST ONLY: copy the quaternion sign of this field
                              (the reference's eigh sign is LAPACK-defined,
                              affine_utils.py:336-355); NULL = canonical w >= 0 */
} gb_aux;

int gb_version(void);
const char* gb_last_error(const gb_ctx* ctx); /* ctx may be NULL: last error of a failed gb_create */

/* Denoiser.__init__ (model.py:20-123). */
int gb_create(gb_ctx** out, const gb_dims* dims, int device);
void gb_destroy(gb_ctx* ctx);

/* load_state_dict: copies and repacks (gate|proj concatenation, transposes, the
 * timestep-embedding table W_time . enc(t) for t = 0..n_timestep). */
int gb_load_weights(gb_ctx* ctx, const gb_tensor* tensors, int n);

/* DDPM.setup_schedule tables the sampler reads (ddpm.py:36-66; base.py:249-250,270);
 * each of length n_timestep + 1, host or device. */
int gb_set_schedule(gb_ctx* ctx, const float* alphas, const float* sqrt_alphas,
                    const float* sqrt_one_minus_alphas_cumprod, const float* sqrt_betas);

int gb_set_precision(gb_ctx* ctx, int precision);

/* Binds a feature batch (base.py:216-224), sizes the workspace for (B,N) and computes
 * the step-invariant parts of the embeddings. */
int gb_set_features(gb_ctx* ctx, const gb_features* f);

/* compute_frenet_frames (utils/geo_utils.py:21-85) on the bound batch. */
int gb_frenet(gb_ctx* ctx, const float* trans, float* rots_out, void* stream);

/* Denoiser.forward (model.py:125-192): timesteps int32 [B] (device); z_out [B,N,3]. */
int gb_denoise(gb_ctx* ctx, const float* rots, const float* trans, const int32_t* timesteps,
               float* z_out, const gb_aux* aux, void* stream);

/* One reverse step, body of BaseSampler._sample (sampler/base.py:249-282): posterior mean,
 * scaled noise (none at step 1), masks, then the Frenet frames of x_{t-1}.
 * noise [B,N,3] N(0,1) draw (device) or NULL to draw it in-kernel (Philox, seed/step keyed). */
int gb_p_sample(gb_ctx* ctx, float* trans_inout, float* rots_inout, const float* z, int step,
                float scale, const float* noise, uint64_t seed, void* stream);

/* steps t_start, t_start-1, ..., t_end (inclusive) of the loop at base.py:236-282,
 * no host synchronisation inside.  noise_tape NULL or [n_timestep+1,B,N,3] with
 * tape[t] the draw used at step t.  Three or more steps, or a repeated call with the same
 * arguments, replay ONE captured step as a CUDA graph on `stream` (the step index lives in
 * device memory); results are bit-identical to launching kernel by kernel (GB_STEP_GRAPH=0). */
int gb_sample_loop(gb_ctx* ctx, float* trans_inout, float* rots_inout, int t_start, int t_end,
                   float scale, const float* noise_tape, uint64_t seed, void* stream);

/* End to end with HOST buffers (what BaseSampler._sample does around the loop,
 * base.py:226-233,285): uploads x_T [B,N,3] (and the optional host tape), builds the
 * frames, runs steps t_start..t_end, downloads x_{t_end-1} into x_out; synchronises. */
int gb_sample_host(gb_ctx* ctx, const float* x_T_host, float* x_out_host, int t_start, int t_end,
                   float scale, const float* noise_tape_host, uint64_t seed);

/* kernels launched by this context since creation / bytes of workspace held */
int64_t gb_launch_count(const gb_ctx* ctx);
int64_t gb_workspace_bytes(const gb_ctx* ctx);

/* ---- row sharding of ONE long sample (B = 1) over the GPUs of a box (BASELINE config C5) ----
 * No reference counterpart: the reference farms whole samples, one process per GPU
 * (genie/utils/multiprocessor.py:73-100), and a length-1024 sample runs on a single device.
 * One process per GPU; every rank creates its own context on its own device, loads the same
 * weights, binds the same features and makes the same calls.  Rank r owns rows
 * [r N / world, (r+1) N / world) of the pair representation and channels [r 128 / world, ...) of the
 * triangle contraction; the ranks read each other's operand blocks and per-layer IPA rows straight
 * from peer-mapped memory over NVLink (CUDA IPC, set up by gb_set_features; device-side barriers and
 * copy kernels on the caller's stream; NCCL only carries the setup), and every rank returns the
 * full z.  The result is bit-identical to the single-GPU one.
 *   gb_shard_unique_id: rank 0 fills 128 bytes (an ncclUniqueId) that the caller distributes
 *                       to the other ranks by any means (torch.distributed, a file, MPI).
 *   gb_shard_init:      collective; before gb_set_features.  world == 1 turns sharding off.
 * Requirements: bf16 precision, B == 1, N a multiple of 128 * world, world <= 8 GPUs of one box with
 * peer access, no triangle attention; gb_set_features is collective on a sharded context.
 * libnccl.so.2 is resolved with dlopen on first use (GB_ERR_MISSING if absent). */
int gb_shard_unique_id(void* out128);
int gb_shard_init(gb_ctx* ctx, int world, int rank, const void* id128);
int gb_shard_info(const gb_ctx* ctx, int* world, int* rank);
int64_t gb_collective_count(const gb_ctx* ctx);   /* NCCL collectives enqueued by this context */
int64_t gb_peer_barrier_count(const gb_ctx* ctx); /* device-side barriers of the ranks (flags in peer-mapped memory) enqueued */

/* ---- per-kernel entry points (unit parity tests; same kernels the forward uses) ---- */

/* geo_utils.py:21-85 without a context. */
int gb_k_frenet(const float* trans, const int32_t* chain_index, const int32_t* residue_mask,
                int B, int N, float* rots_out, void* stream);

/* TriangleMultiplicativeUpdate.forward (triangular_multiplicative_update.py:84-110) of pair
 * layer `layer`; outgoing != 0 selects tri_mul_out.  p [B,N,N,c_p] is updated in place
 * (p += update), i.e. including the residual add at pair_transform_net.py:111-112. */
int gb_k_trimul(gb_ctx* ctx, int layer, int outgoing, float* p, void* stream);

/* One kernel of the tensor-core (bf16) triangle update, for per-kernel timing and tests: stage 1 = LayerNorm +
 * gated projections (triangular_multiplicative_update.py:97-103,107), 2 = the per-channel contraction
 * (:57-82), 3 = LayerNorm_out + linear_z + gate + residual (:105-108).  Stages 2, 3 consume what the
 * previous stage of the same context left in the workspace. */
int gb_k_trimul_stage(gb_ctx* ctx, int layer, int outgoing, int stage, float* p, void* stream);

/* PairTransition.forward + residual + the layer-final mask (pair_transition.py:48-87,
 * pair_transform_net.py:116-117): p = (p + transition(p)) * mask, in place. */
int gb_k_pair_transition(gb_ctx* ctx, int layer, float* p, void* stream);

/* TriangleAttention{Starting,Ending}Node.forward + residual (triangular_attention.py:67-130). */
int gb_k_tri_attention(gb_ctx* ctx, int layer, int starting, float* p, void* stream);

/* InvariantPointAttention.forward (invariant_point_attention.py:100-260) of structure layer
 * `layer`: out [B,N,c_s] = ipa(s, p, T(rots,trans), residue_mask) (no residual). */
int gb_k_ipa(gb_ctx* ctx, int layer, const float* s, const float* p, const float* rots,
             const float* trans, float* out, void* stream);

/* gb_k_ipa for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same arguments, same kernels and the
 * same result as gb_k_ipa (fp32 CUDA cores), but its scratch (projection rows [B*N][1152] + concat rows [B*N][2112]) is a
 * file-scope static __device__ array of the library instead of the context's workspace, and the call performs no
 * cudaMalloc / cudaFree and calls no allocating routine.  Limit: B * N <= GB_K_IPA_STATIC_MAX_ROWS of the bound feature
 * batch (larger batches return GB_ERR_UNSUPPORTED); the array is 13,369,344 bytes on every GPU that loads the library,
 * held until the process exits.  ALL calls, from every context on a device, share that one array and it keeps the last
 * call's contents (nothing in it needs to start at zero: every element read is written earlier in the same call): calls
 * that can overlap on the GPU (different streams, different contexts on one device) RACE -- serialise them.  Not
 * available on a row-sharded context. */
#define GB_K_IPA_STATIC_MAX_ROWS 1024
int gb_k_ipa_static(gb_ctx* ctx, int layer, const float* s, const float* p, const float* rots,
                    const float* trans, float* out, void* stream);

/* TEST ONLY: copy one bf16 intermediate of the last tensor-core triangle update into dst (device):
 * which = 0/1: a / b operand tile images [B][128][Np/64][Np][64]; 2: x tile images [B][Np/2][Np/64][2][128][64];
 * 3: g [B*N*N][128].  Returns the padded length Np (> 0) or a negative gb_status; dst may be NULL
 * to query Np only.  Lets unit tests check each tcgen05 kernel separately. */
int gb_dbg_tc_copy(gb_ctx* ctx, int which, void* dst, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* GENIE2_B200_H_ */
