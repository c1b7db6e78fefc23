// IPA geometry of the base configuration and the layout of the fused projection / concat rows.
#pragma once
#define IPA_H 12
#define IPA_C 16
#define IPA_PQ 4
#define IPA_PV 8
#define IPA_CP 128
// layout of one row of the fused projection buffer qkv [B*N, 1152]:
//   [0,192) q (h,c) | [192,576) kv (h, k16|v16) | [576,720) q_pts (xyz-major: c*48 + h*4 + p)
//   | [720,1152) kv_pts (xyz-major: c*144 + h*12 + p, p<4 = k points, p>=4 = v points)
#define IPA_QKV 1152
#define IPA_OFF_KV 192
#define IPA_OFF_QP 576
#define IPA_OFF_KVP 720
// concat row: o(192) | o_pt.x(96) | o_pt.y(96) | o_pt.z(96) | |o_pt|(96) | o_pair(12*128)
#define IPA_CAT (IPA_H * (IPA_C + IPA_PV * 4 + IPA_CP))

// Tensor-core path: the same 1152 outputs per row, columns reordered by the weight packer so that the
// projection GEMM can scatter them into
//   Q rows  [R][IPA_QW]:  q (h*16 + c) | q_pts (192 + c*48 + h*4 + p)
//   planes  [B][IPA_NPLANE][N][4] (key-major, one float4 per key): plane = h*17 + e
//     e 0..3  k[4e..4e+3]        e 4..6   k_pts coordinate e-4, points 0..3
//     e 7..10 v[4(e-7)..]        e 11..16 v_pts coordinate (e-11)/2, points 4*((e-11)%2) .. +3
// Consecutive keys are consecutive 16-byte words, so thread-per-key / lane-per-key reads are coalesced.
#define IPA_QW 336
#define IPA_NPLANE 204
#define IPA_PLANES_PER_HEAD 17
static inline int ipa_old_col(int n) {   // new output column -> column of the reference-ordered row (layout above)
  if (n < 192) return n;
  if (n < IPA_QW) return IPA_OFF_QP + (n - 192);
  const int plane = (n - IPA_QW) >> 2, comp = (n - IPA_QW) & 3, h = plane / IPA_PLANES_PER_HEAD, e = plane % IPA_PLANES_PER_HEAD;
  if (e < 4) return IPA_OFF_KV + h * 32 + 4 * e + comp;
  if (e < 7) return IPA_OFF_KVP + (e - 4) * 144 + h * 12 + comp;
  if (e < 11) return IPA_OFF_KV + h * 32 + 16 + 4 * (e - 7) + comp;
  return IPA_OFF_KVP + ((e - 11) >> 1) * 144 + h * 12 + 4 + ((e - 11) & 1) * 4 + comp;
}
