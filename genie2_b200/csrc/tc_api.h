// Interface between the orchestration (api.cu) and the tcgen05 tensor-core module (tc_*.cu).
#pragma once
#include <vector>

#include "ctx.h"

int tc_init(gb_ctx* c);
void tc_destroy(gb_ctx* c);
int tc_pack_trimul(gb_ctx* c, TriMulW& w, const std::vector<float>& ap, const std::vector<float>& ag,
                   const std::vector<float>& bp, const std::vector<float>& bg, const std::vector<float>& lg,
                   const std::vector<float>& lz, const std::vector<float>& apb, const std::vector<float>& agb,
                   const std::vector<float>& bpb, const std::vector<float>& bgb, const std::vector<float>& lgb,
                   const std::vector<float>& lzb, const std::vector<float>& lni_g, const std::vector<float>& lni_b,
                   const std::vector<float>& lno_g, const std::vector<float>& lno_b);
// K4 (tc_proj.cu): folded / scaled weight images, SM count, launch
int tc_proj_init(gb_ctx* c);
int tc_pack_trimul_proj(gb_ctx* c, TriMulW& w, const std::vector<float>* Wm[5], const std::vector<float>* bm[5],
                        const std::vector<float>& ln_g, const std::vector<float>& ln_b);
int tc_launch_trimul_proj(gb_ctx* c, const TriMulW& w, bool outgoing, const float* p, cudaStream_t st);
// GB_PREC_BF16X3 (split-bf16, three passes per product): tc_proj_x3.cu, tc_transition_x3.cu
int tc_proj_x3_init(gb_ctx* c);
int tc_pack_trimul_proj_x3(gb_ctx* c, TriMulW& w, const std::vector<float>* Wm[5], const std::vector<float>* bm[5],
                           const std::vector<float>& ln_g, const std::vector<float>& ln_b);
int tc_launch_trimul_proj_x3(gb_ctx* c, const TriMulW& w, bool outgoing, const float* p, cudaStream_t st);
int tc_transition_x3_init(gb_ctx* c);
int tc_pack_transition_x3(gb_ctx* c, TransitionW& w, const std::vector<float>& w1, const std::vector<float>& w2,
                          const std::vector<float>& ln_g);
int tc_pair_transition_x3(gb_ctx* c, const TransitionW& w, float* p, cudaStream_t st);
int tc_transition_init(gb_ctx* c);
int tc_pack_transition(gb_ctx* c, TransitionW& w, const std::vector<float>& w1, const std::vector<float>& b1,
                       const std::vector<float>& w2, const std::vector<float>& ln_g, const std::vector<float>& ln_b);
// workspace the tensor-core path needs for a (B,N) batch, carved by gb_set_features
size_t tc_workspace_bytes(const gb_ctx* c, int B, int N);
int tc_bind_workspace(gb_ctx* c, void* base, size_t bytes, int B, int N);
// p += TriangleMultiplicativeUpdate(p) with tcgen05 kernels (precision = c->precision)
int tc_trimul(gb_ctx* c, const TriMulW& w, bool outgoing, float* p, cudaStream_t st);
// one kernel of the update only: stage 1 = K4 projections, 2 = K5 contraction, 3 = K6 output (0 = all)
int tc_trimul_stage(gb_ctx* c, const TriMulW& w, bool outgoing, int stage, float* p, cudaStream_t st);
// p = (p + PairTransition(p)) * mask
int tc_pair_transition(gb_ctx* c, const TransitionW& w, float* p, cudaStream_t st);

// ---- triangle attention on tensor cores (tc_triatt.cu, GB_PREC_BF16)
int tc_triatt_init(gb_ctx* c);
int tc_pack_triatt(gb_ctx* c, TriAttW& w, const std::vector<float>& wq, const std::vector<float>& wk, const std::vector<float>& wv,
                   const std::vector<float>& wg, const std::vector<float>& wo);
int tc_tri_attention(gb_ctx* c, const TriAttW& w, bool starting, float* p, cudaStream_t st);

// ---- pair embedding on tensor cores (tc_embed.cu); wt [128][nbin + 6], wm [128][nbin + 2] as in the state dict
int tc_pack_pair_embed(gb_ctx* c, const std::vector<float>& wt, const std::vector<float>& wm, int nbin);
bool tc_pair_embed_supported(const gb_ctx* c);
int tc_pair_embed(gb_ctx* c, const float* quat_follow, int rows, int row0, cudaStream_t st);

// Pi | Pj = s . [W_i | W_j]^T on tensor cores (tc_struct.cu); leaves the operand image of s for the structure net
int tc_pack_pipj(gb_ctx* c, const std::vector<float>& wij);
bool tc_pipj_supported(const gb_ctx* c);
int tc_pair_pipj(gb_ctx* c, cudaStream_t st);

// ---- structure network on tensor cores (tc_struct.cu)
int tc_struct_init(gb_ctx* c);
bool tc_struct_supported(const gb_ctx* c, int N);
size_t tc_struct_workspace_bytes(const gb_ctx* c, int B, int N);
int tc_struct_bind(gb_ctx* c, void* base, size_t bytes, int B, int N);
int tc_pack_struct_layer(gb_ctx* c, int l, const std::vector<float>& wqkv, const std::vector<float>& bqkv, const std::vector<float>& wout,
                         const std::vector<const std::vector<float>*>& w123);
int tc_pack_ipa_bias(gb_ctx* c, const std::vector<const std::vector<float>*>& wb, const std::vector<const std::vector<float>*>& bb);
// StructureNet.forward on c->s / c->p / c->rots_w / c->trans_w (bf16 mode); states_out may be NULL
int tc_structure_net(gb_ctx* c, float* states_out, cudaStream_t st);
// SIMT kernels of api.cu the tensor-core structure net reuses
int simt_points_to_global(gb_ctx* c, cudaStream_t st);
int simt_backbone_update(gb_ctx* c, const StructLayerW& w, cudaStream_t st);
