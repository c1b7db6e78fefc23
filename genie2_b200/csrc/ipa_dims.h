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
