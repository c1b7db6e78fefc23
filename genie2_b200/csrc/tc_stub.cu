// Placeholder for the tcgen05 tensor-core module until tc_pair.cu lands.
#include "tc_api.h"
int tc_init(gb_ctx*) { return GB_OK; }
void tc_destroy(gb_ctx*) {}
int tc_pack_trimul(gb_ctx*, TriMulW&, const std::vector<float>&, const std::vector<float>&, const std::vector<float>&,
                   const std::vector<float>&, const std::vector<float>&, const std::vector<float>&, const std::vector<float>&,
                   const std::vector<float>&, const std::vector<float>&, const std::vector<float>&, const std::vector<float>&) { return GB_OK; }
int tc_pack_transition(gb_ctx*, TransitionW&, const std::vector<float>&, const std::vector<float>&) { return GB_OK; }
size_t tc_workspace_bytes(const gb_ctx*, int, int) { return 0; }
int tc_bind_workspace(gb_ctx*, void*, size_t) { return GB_OK; }
int tc_trimul(gb_ctx* c, const TriMulW&, bool, float*, cudaStream_t) { return gb_fail(c, GB_ERR_UNSUPPORTED, "tensor-core path not built"); }
int tc_pair_transition(gb_ctx* c, const TransitionW&, float*, cudaStream_t) { return gb_fail(c, GB_ERR_UNSUPPORTED, "tensor-core path not built"); }
