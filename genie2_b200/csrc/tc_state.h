// State of the tensor-core modules (tc_pair.cu, tc_struct.cu), hung off gb_ctx::tc_state.
#pragma once
#include <cuda_bf16.h>

#include <vector>

typedef __nv_bfloat16 bf16;

struct TcStructW {                 // per structure layer, bf16 operand images [K/64][N][64]
  bf16 *Wqkv = nullptr, *Wout = nullptr;   // Wqkv rows (and bqkv) in the plane order of ipa_dims.h
  float* bqkv = nullptr;
  std::vector<bf16*> W1, W2, W3;
};

struct TcState {
  int Np = 0;                        // N rounded up to a multiple of 64
  bf16 *a = nullptr, *b = nullptr;   // [B][128][Np/64][Np][64] tile images
  bf16 *x = nullptr;                 // [B][128][Np][Np]
  bf16 *g = nullptr;                 // [B*N*N][128]
  size_t ab_bytes = 0;
  bf16 *ta_q = nullptr, *ta_k = nullptr, *ta_vt = nullptr, *ta_g = nullptr;   // triangle attention (tc_triatt.cu), only with include_tri_att
  size_t ta_lo = 0;                  // elements from any of them to its lo twin (GB_PREC_BF16X3)
  size_t lo_plane = 0;               // GB_PREC_BF16X3: elements from a / b / x (the hi planes) to their lo planes
  int num_sms = 0;
  // structure net
  int Rp = 0;                        // B*N rounded up to a multiple of 128
  bf16 *s_img = nullptr, *cat_img = nullptr, *ln1_img = nullptr, *h1_img = nullptr, *h2_img = nullptr;
  bf16* z_img = nullptr;             // [B][N][2][Np][64]
  size_t img_lo = 0;                 // GB_PREC_BF16X3: elements from any image above to its lo twin
  float* pbias = nullptr;            // [L][B][N][12][Np]
  bf16* Wpipj_img = nullptr;         // [c_s/64][256][64]: [W_i | W_j] of the pair feature net
  bool s_img_valid = false;          // s_img holds the image of the current single features (set by tc_pair_pipj)
  bf16* Wfeat_img = nullptr;         // pair embedding: [2 kb][128 ch][64 k] (tc_embed.cu)
  float* wfst = nullptr;             // [2][128]
  bf16* Wb_img = nullptr;            // [2][96][64]  rows l*12+h, scaled by sqrt(1/3)
  float* bias96 = nullptr;
  std::vector<TcStructW> st;
};
