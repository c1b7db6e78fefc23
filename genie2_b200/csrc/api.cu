// C ABI of genie2_b200 (include/genie2_b200.h): context, weight packing, feature binding and
// the orchestration of Denoiser.forward / the reverse-diffusion step as a fixed kernel sequence.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cmath>

#include "ctx.h"
#include "gemm_simt.cuh"
#include "kernels_embed.cuh"
#include "kernels_frames.cuh"
#include "kernels_pair_simt.cuh"
#include "kernels_structure.cuh"
#include "kernels_triatt.cuh"
#include "shard.h"
#include "tc_api.h"

static std::string g_create_err;

// Selects the context's device for the duration of an entry point and restores the caller's current device afterwards,
// so that two contexts on different GPUs can be driven from one host thread.
struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) changed = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
};

int gb_fail(const gb_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf; else g_create_err = buf;
  return code;
}

extern "C" int gb_version(void) { return GB_VERSION; }
extern "C" const char* gb_last_error(const gb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
extern "C" int64_t gb_launch_count(const gb_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int64_t gb_workspace_bytes(const gb_ctx* ctx) { return ctx ? (int64_t)ctx->ws_bytes : 0; }
namespace {
struct PhaseMarks {
  bool on = false;
  std::vector<std::pair<const char*, cudaEvent_t>> marks;
  PhaseMarks() { const char* e = getenv("GB_PHASE_TIMING"); on = e && e[0] == '1'; }
} g_phases;
}  // namespace

void ph_mark(cudaStream_t st, const char* name) {
  if (!g_phases.on) return;
  cudaEvent_t ev;
  if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  g_phases.marks.emplace_back(name, ev);
}

void ph_report(const gb_ctx* ctx, cudaStream_t st) {
  if (!g_phases.on) return;
  ph_mark(st, "end");
  cudaStreamSynchronize(st);
  std::vector<std::pair<std::string, std::pair<float, int>>> rows;
  float total = 0.f;
  for (size_t i = 0; i + 1 < g_phases.marks.size(); ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, g_phases.marks[i].second, g_phases.marks[i + 1].second);
    total += ms;
    bool found = false;
    for (auto& r : rows) if (r.first == g_phases.marks[i].first) { r.second.first += ms; r.second.second++; found = true; break; }
    if (!found) rows.push_back({g_phases.marks[i].first, {ms, 1}});
  }
  fprintf(stderr, "[gb phases] rank %d of %d, N = %d: %.3f ms\n", ctx->shard_rank, ctx->shard_world, ctx->N, total);
  for (auto& r : rows) fprintf(stderr, "[gb phases]   %-16s %8.3f ms  x%d\n", r.first.c_str(), r.second.first, r.second.second);
  for (auto& m : g_phases.marks) cudaEventDestroy(m.second);
  g_phases.marks.clear();
}

extern "C" int64_t gb_collective_count(const gb_ctx* ctx) { return ctx ? ctx->collectives : 0; }
extern "C" int64_t gb_peer_barrier_count(const gb_ctx* ctx) { return ctx ? ctx->peer_barriers : 0; }

// Scratch of gb_k_ipa_static (the entry point that allocates nothing): the fused projection rows [rows][1152] followed by
// the concat rows [rows][2112], rows <= GB_K_IPA_STATIC_MAX_ROWS: 13,369,344 bytes that every GPU which loads this library
// holds until the process exits.  In HOST code this name is only a stand-in for the symbol, not a device address: the
// address is taken once per context with cudaGetSymbolAddress in gb_create (the context's device current) and only that
// pointer is ever passed to a kernel.
static __device__ __align__(256) float g_ipa_static_scratch[(size_t)GB_K_IPA_STATIC_MAX_ROWS * (IPA_QKV + IPA_CAT)];

static const int kSmemPair = (PS_A_FLOATS + PS_W_FLOATS) * 4;
static const int kSmemTrans = (PS_A_FLOATS + PS_W_FLOATS + PS_A_FLOATS) * 4;

extern "C" int gb_create(gb_ctx** out, const gb_dims* d, int device) {
  if (!out || !d) return gb_fail(nullptr, GB_ERR_INVALID, "gb_create: NULL argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return gb_fail(nullptr, GB_ERR_CUDA, "gb_create: no CUDA device (this library has no CPU path)");
  if (device < 0 || device >= ndev) return gb_fail(nullptr, GB_ERR_INVALID, "gb_create: bad device %d", device);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return gb_fail(nullptr, GB_ERR_CUDA, "gb_create: cudaGetDeviceProperties failed");
  if (prop.major != 10)
    return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "gb_create: device %d is sm_%d%d; this build targets sm_100a (B200) only",
                   device, prop.major, prop.minor);
  // compiled specialisations (SURVEY.md section 5 "Config / flags")
  if (d->c_p != 128 || d->c_hidden_mul != 128) return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "c_p and c_hidden_mul must be 128");
  if (d->n_head_ipa != IPA_H || d->c_hidden_ipa != IPA_C || d->n_qk_point != IPA_PQ || d->n_v_point != IPA_PV)
    return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "IPA geometry must be H=12, C=16, Pq=4, Pv=8");
  if (d->template_dist_n_bin != 37) return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "template_dist_n_bin must be 37");
  if (d->pair_transition_n < 1 || d->pair_transition_n > 8) return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "pair_transition_n out of range");
  if (d->include_tri_att && (d->n_head_tri != TA_H || d->c_hidden_tri_att != TA_C))
    return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "triangle attention must be 4 heads x 32");
  if (d->c_s % 4 != 0 || d->c_s > 1024) return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "c_s must be a multiple of 4, <= 1024");
  if (d->rescale == 0.f) return gb_fail(nullptr, GB_ERR_INVALID, "rescale must be non-zero");
  DeviceGuard dg(device);
  gb_ctx* c = new gb_ctx();
  c->d = *d;
  c->device = device;
  // opt in to > 48 KB dynamic shared memory
  cudaFuncSetAttribute(trimul_proj_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemPair);
  cudaFuncSetAttribute(trimul_out_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemPair);
  cudaFuncSetAttribute(pair_transition_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTrans);
  cudaFuncSetAttribute(ipa_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  triatt_set_attributes();
  cudaFuncSetAttribute(frenet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(ddpm_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (cudaGetLastError() != cudaSuccess) { delete c; return gb_fail(nullptr, GB_ERR_CUDA, "cudaFuncSetAttribute failed"); }
  {
    void* sp = nullptr;     // device `device` is current here (DeviceGuard above)
    const cudaError_t se = cudaGetSymbolAddress(&sp, g_ipa_static_scratch);
    if (se != cudaSuccess || !sp) { delete c; return gb_fail(nullptr, GB_ERR_CUDA, "gb_create: cudaGetSymbolAddress(static scratch) failed: %s", cudaGetErrorString(se)); }
    c->static_scratch = static_cast<float*>(sp);
  }
  int rc = tc_init(c);
  if (rc != GB_OK) { g_create_err = c->err; delete c; return rc; }
  *out = c;
  return GB_OK;
}

static void drop_step_graph(gb_ctx* c) {
  if (c->step_graph) cudaGraphExecDestroy((cudaGraphExec_t)c->step_graph);
  c->step_graph = nullptr;
  c->step_key = gb_ctx::StepKey();
  ++c->epoch;
}

static void free_features(gb_ctx* c) {
  drop_step_graph(c);
  if (c->feat_alloc) cudaFree(c->feat_alloc);
  if (c->ws_alloc) cudaFree(c->ws_alloc);
  c->feat_alloc = c->ws_alloc = nullptr;
  c->feat_cap = c->ws_cap = 0;
  c->features_set = false;
}

extern "C" void gb_destroy(gb_ctx* c) {
  if (!c) return;
  DeviceGuard dg(c->device);
  shard_destroy(c);
  tc_destroy(c);
  free_features(c);
  if (c->cap_stream) cudaStreamDestroy((cudaStream_t)c->cap_stream);
  if (c->tape_alloc) cudaFree(c->tape_alloc);
  if (c->alphas) cudaFree(c->alphas);
  for (void* p : c->weight_allocs) cudaFree(p);
  delete c;
}

// ------------------------------------------------------------------ weight helpers
static float* upload(gb_ctx* c, const std::vector<float>& v) {
  void* p = nullptr;
  if (cudaMalloc(&p, v.size() * sizeof(float)) != cudaSuccess) return nullptr;
  if (cudaMemcpy(p, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(p); return nullptr; }
  c->weight_allocs.push_back(p);
  return (float*)p;
}

struct Getter {
  gb_ctx* c;
  bool ok = true;
  std::string missing;
  const std::vector<float>& get(const std::string& name, std::initializer_list<int64_t> shape) {
    static std::vector<float> empty;
    auto it = c->host.find(name);
    if (it == c->host.end()) { if (ok) missing = name; ok = false; return empty; }
    size_t n = 1, k = 0;
    bool same = it->second.shape.size() == shape.size();
    for (int64_t s : shape) { n *= (size_t)s; if (same && it->second.shape[k] != s) same = false; ++k; }
    if (!same || it->second.v.size() != n) { if (ok) missing = name + " (shape mismatch)"; ok = false; return empty; }
    return it->second.v;
  }
};

// dst[k][n] = src[n][k]   src [rows n][cols k]
static std::vector<float> transpose(const std::vector<float>& src, int rows, int cols) {
  std::vector<float> d((size_t)rows * cols);
  if (src.empty()) return d;
  for (int n = 0; n < rows; ++n)
    for (int k = 0; k < cols; ++k) d[(size_t)k * rows + n] = src[(size_t)n * cols + k];
  return d;
}
static std::vector<float> concat(std::initializer_list<const std::vector<float>*> parts) {
  std::vector<float> d;
  for (auto* p : parts) d.insert(d.end(), p->begin(), p->end());
  return d;
}

static int pack_trimul(gb_ctx* c, Getter& G, const std::string& pre, TriMulW& w) {
  const auto& ap = G.get(pre + ".linear_a_p.weight", {128, 128});
  const auto& ag = G.get(pre + ".linear_a_g.weight", {128, 128});
  const auto& bp = G.get(pre + ".linear_b_p.weight", {128, 128});
  const auto& bg = G.get(pre + ".linear_b_g.weight", {128, 128});
  const auto& lg = G.get(pre + ".linear_g.weight", {128, 128});
  const auto& lz = G.get(pre + ".linear_z.weight", {128, 128});
  const auto& apb = G.get(pre + ".linear_a_p.bias", {128});
  const auto& agb = G.get(pre + ".linear_a_g.bias", {128});
  const auto& bpb = G.get(pre + ".linear_b_p.bias", {128});
  const auto& bgb = G.get(pre + ".linear_b_g.bias", {128});
  const auto& lgb = G.get(pre + ".linear_g.bias", {128});
  const auto& lzb = G.get(pre + ".linear_z.bias", {128});
  const auto& lig = G.get(pre + ".layer_norm_in.weight", {128});
  const auto& lib = G.get(pre + ".layer_norm_in.bias", {128});
  const auto& log_ = G.get(pre + ".layer_norm_out.weight", {128});
  const auto& lob = G.get(pre + ".layer_norm_out.bias", {128});
  if (!G.ok) return GB_ERR_MISSING;
  std::vector<float> Wpk(5 * 128 * 128), bpk(5 * 128);
  for (int t = 0; t < 5; ++t)
    for (int n = 0; n < 128; ++n) {
      const std::vector<float>* W;
      const std::vector<float>* bsrc;
      int row;
      if (t == 4) { W = &lg; bsrc = &lgb; row = n; }
      else {
        const int c0 = (t & 1) * 64;
        const bool gate = n >= 64;
        row = c0 + (n & 63);
        W = (t < 2) ? (gate ? &ag : &ap) : (gate ? &bg : &bp);
        bsrc = (t < 2) ? (gate ? &agb : &apb) : (gate ? &bgb : &bpb);
      }
      bpk[t * 128 + n] = (*bsrc)[row];
      for (int k = 0; k < 128; ++k) Wpk[((size_t)t * 128 + k) * 128 + n] = (*W)[(size_t)row * 128 + k];
    }
  w.Wpk = upload(c, Wpk); w.bpk = upload(c, bpk);
  w.WzT = upload(c, transpose(lz, 128, 128)); w.bz = upload(c, lzb);
  w.ln_in_g = upload(c, lig); w.ln_in_b = upload(c, lib);
  w.ln_out_g = upload(c, log_); w.ln_out_b = upload(c, lob);
  return tc_pack_trimul(c, w, ap, ag, bp, bg, lg, lz, apb, agb, bpb, bgb, lgb, lzb, lig, lib, log_, lob);
}

static int pack_transition(gb_ctx* c, Getter& G, const std::string& pre, TransitionW& w) {
  const int nh = c->d.pair_transition_n;
  const auto& w1 = G.get(pre + ".linear_1.weight", {128 * nh, 128});
  const auto& b1 = G.get(pre + ".linear_1.bias", {128 * nh});
  const auto& w2 = G.get(pre + ".linear_2.weight", {128, 128 * nh});
  const auto& b2 = G.get(pre + ".linear_2.bias", {128});
  const auto& lg = G.get(pre + ".layer_norm.weight", {128});
  const auto& lb = G.get(pre + ".layer_norm.bias", {128});
  if (!G.ok) return GB_ERR_MISSING;
  std::vector<float> W1T((size_t)nh * 128 * 128), W2T((size_t)nh * 128 * 128);
  for (int t = 0; t < nh; ++t)
    for (int k = 0; k < 128; ++k)
      for (int n = 0; n < 128; ++n) {
        W1T[((size_t)t * 128 + k) * 128 + n] = w1[((size_t)t * 128 + n) * 128 + k];        // hidden col n of tile t
        W2T[((size_t)t * 128 + k) * 128 + n] = w2[(size_t)n * (128 * nh) + t * 128 + k];    // k = hidden idx in tile
      }
  w.W1T = upload(c, W1T); w.b1 = upload(c, b1); w.W2T = upload(c, W2T); w.b2 = upload(c, b2);
  w.ln_g = upload(c, lg); w.ln_b = upload(c, lb);
  return tc_pack_transition(c, w, w1, b1, w2, lg, lb);
}

static int pack_triatt(gb_ctx* c, Getter& G, const std::string& pre, TriAttW& w) {
  const int hc = TA_H * TA_C;
  const auto& lg = G.get(pre + ".layer_norm.weight", {128});
  const auto& lb = G.get(pre + ".layer_norm.bias", {128});
  const auto& wl = G.get(pre + ".linear.weight", {TA_H, 128});
  const auto& wq = G.get(pre + ".mha.linear_q.weight", {hc, 128});
  const auto& wk = G.get(pre + ".mha.linear_k.weight", {hc, 128});
  const auto& wv = G.get(pre + ".mha.linear_v.weight", {hc, 128});
  const auto& wg = G.get(pre + ".mha.linear_g.weight", {hc, 128});
  const auto& bg = G.get(pre + ".mha.linear_g.bias", {hc});
  const auto& wo = G.get(pre + ".mha.linear_o.weight", {128, hc});
  const auto& bo = G.get(pre + ".mha.linear_o.bias", {128});
  if (!G.ok) return GB_ERR_MISSING;
  w.ln_g = upload(c, lg); w.ln_b = upload(c, lb); w.Wbias = upload(c, wl);
  w.Wqkvg = upload(c, concat({&wq, &wk, &wv, &wg})); w.bg = upload(c, bg);
  std::vector<float> bq(4 * hc, 0.f);   // q, k, v have no bias (primitives.py:243-251)
  std::copy(bg.begin(), bg.end(), bq.begin() + 3 * hc);
  w.bqkvg = upload(c, bq);
  w.Wo = upload(c, wo); w.bo = upload(c, bo);
  return tc_pack_triatt(c, w, wq, wk, wv, wg, wo);
}

extern "C" int gb_load_weights(gb_ctx* c, const gb_tensor* tensors, int n) {
  if (!c || !tensors || n <= 0) return gb_fail(c, GB_ERR_INVALID, "gb_load_weights: bad arguments");
  DeviceGuard dg(c->device);
  for (void* p : c->weight_allocs) cudaFree(p);
  c->weight_allocs.clear();
  c->host.clear();
  c->weights_loaded = false;
  // s_static (gb_set_features) is W_single . static features of the weights loaded at that time: a re-bind is required
  c->features_set = false;
  drop_step_graph(c);
  for (int i = 0; i < n; ++i) {
    const gb_tensor& t = tensors[i];
    if (!t.name || !t.data || t.ndim < 1 || t.ndim > 4) return gb_fail(c, GB_ERR_INVALID, "gb_load_weights: tensor %d malformed", i);
    std::string name = t.name;
    if (name.compare(0, 6, "model.") == 0) name = name.substr(6);   // Lightning prefix (ddpm.py:26)
    HostTensor h;
    size_t cnt = 1;
    for (int k = 0; k < t.ndim; ++k) { h.shape.push_back(t.shape[k]); cnt *= (size_t)t.shape[k]; }
    h.v.resize(cnt);
    GB_CUDA(c, cudaMemcpy(h.v.data(), t.data, cnt * sizeof(float), cudaMemcpyDefault));
    c->host[name] = std::move(h);
  }
  const gb_dims& d = c->d;
  Getter G{c};
  const int c_s = d.c_s, F = d.c_pos_emb + d.c_chain_emb + d.c_timestep_emb + 23;
  // ---- single feature net: raw weight + timestep table W_time . enc(t)
  const auto& ws = G.get("single_feature_net.linear.weight", {c_s, F});
  const auto& wi = G.get("pair_feature_net.linear_s_p_i.weight", {128, c_s});
  const auto& wj = G.get("pair_feature_net.linear_s_p_j.weight", {128, c_s});
  const auto& wr = G.get("pair_feature_net.linear_relpos.weight", {128, 2 * d.relpos_k + 3});
  const auto& wt = G.get("pair_feature_net.linear_template.weight", {128, d.template_dist_n_bin + 6});
  const auto& wm = G.get("pair_feature_net.linear_motif_template.weight", {128, d.template_dist_n_bin + 2});
  if (!G.ok) return gb_fail(c, GB_ERR_MISSING, "gb_load_weights: missing tensor %s", G.missing.c_str());
  c->W_single = upload(c, ws);
  c->W_pipj = upload(c, concat({&wi, &wj}));
  {
    const int rc = tc_pack_pipj(c, concat({&wi, &wj}));
    if (rc != GB_OK) return rc;
  }
  c->WrelT = upload(c, transpose(wr, 128, 2 * d.relpos_k + 3));
  c->WtT = upload(c, transpose(wt, 128, d.template_dist_n_bin + 6));
  c->WmT = upload(c, transpose(wm, 128, d.template_dist_n_bin + 2));
  {
    const int rc = tc_pack_pair_embed(c, wt, wm, d.template_dist_n_bin);
    if (rc != GB_OK) return rc;
  }
  {
    const int T1 = d.n_timestep + 1, D = d.c_timestep_emb;
    float* X = nullptr;
    GB_CUDA(c, cudaMalloc(&X, (size_t)T1 * D * sizeof(float)));
    void* tt = nullptr;
    if (cudaMalloc(&tt, (size_t)T1 * c_s * sizeof(float)) != cudaSuccess) { cudaFree(X); return gb_fail(c, GB_ERR_CUDA, "cudaMalloc time table"); }
    c->weight_allocs.push_back(tt);
    c->time_table = (float*)tt;
    time_encoding_kernel<<<(T1 * D + 255) / 256, 256>>>(X, T1, d.n_timestep, D);
    sgemm_nt(0, X, D, c->W_single + d.c_pos_emb + d.c_chain_emb, F, nullptr, nullptr, 0, c->time_table, c_s, T1, c_s, D, false);
    cudaError_t e = cudaDeviceSynchronize();
    cudaFree(X);
    if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "time table kernels failed: %s", cudaGetErrorString(e));
  }
  // ---- pair transform layers
  c->pair.assign(d.n_pair_transform_layer, PairLayerW());
  for (int l = 0; l < d.n_pair_transform_layer; ++l) {
    const std::string pre = "pair_transform_net.net." + std::to_string(l);
    int rc = GB_OK;
    if (d.include_mul_update) {
      rc = pack_trimul(c, G, pre + ".tri_mul_out", c->pair[l].mul_out);
      if (rc == GB_OK) rc = pack_trimul(c, G, pre + ".tri_mul_in", c->pair[l].mul_in);
    }
    if (rc == GB_OK && d.include_tri_att) {
      rc = pack_triatt(c, G, pre + ".tri_att_start", c->pair[l].att_start);
      if (rc == GB_OK) rc = pack_triatt(c, G, pre + ".tri_att_end", c->pair[l].att_end);
    }
    if (rc == GB_OK) rc = pack_transition(c, G, pre + ".pair_transition", c->pair[l].tr);
    if (rc == GB_ERR_MISSING) return gb_fail(c, rc, "gb_load_weights: missing tensor %s", G.missing.c_str());
    if (rc != GB_OK) return rc;
  }
  // ---- structure layers
  c->st.assign(d.n_structure_layer, StructLayerW());
  const int H = IPA_H, hc = IPA_H * IPA_C;
  std::vector<const std::vector<float>*> wb_all, bb_all;
  for (int l = 0; l < d.n_structure_layer; ++l) {
    const std::string pre = "structure_net.net." + std::to_string(l);
    StructLayerW& w = c->st[l];
    const auto& wq = G.get(pre + ".ipa.linear_q.weight", {hc, c_s});
    const auto& wkv = G.get(pre + ".ipa.linear_kv.weight", {2 * hc, c_s});
    const auto& wqp = G.get(pre + ".ipa.linear_q_points.weight", {H * IPA_PQ * 3, c_s});
    const auto& wkvp = G.get(pre + ".ipa.linear_kv_points.weight", {H * (IPA_PQ + IPA_PV) * 3, c_s});
    const auto& bq = G.get(pre + ".ipa.linear_q.bias", {hc});
    const auto& bkv = G.get(pre + ".ipa.linear_kv.bias", {2 * hc});
    const auto& bqp = G.get(pre + ".ipa.linear_q_points.bias", {H * IPA_PQ * 3});
    const auto& bkvp = G.get(pre + ".ipa.linear_kv_points.bias", {H * (IPA_PQ + IPA_PV) * 3});
    const auto& wb = G.get(pre + ".ipa.linear_b.weight", {H, 128});
    const auto& bb = G.get(pre + ".ipa.linear_b.bias", {H});
    const auto& hwt = G.get(pre + ".ipa.head_weights", {H});
    const auto& wo = G.get(pre + ".ipa.linear_out.weight", {c_s, IPA_CAT});
    const auto& bo = G.get(pre + ".ipa.linear_out.bias", {c_s});
    const auto& l1g = G.get(pre + ".ipa_layer_norm.weight", {c_s});
    const auto& l1b = G.get(pre + ".ipa_layer_norm.bias", {c_s});
    const auto& l2g = G.get(pre + ".transition.layer_norm.weight", {c_s});
    const auto& l2b = G.get(pre + ".transition.layer_norm.bias", {c_s});
    const auto& wbb = G.get(pre + ".bb_update.linear.weight", {6, c_s});
    const auto& bbb = G.get(pre + ".bb_update.linear.bias", {6});
    if (!G.ok) return gb_fail(c, GB_ERR_MISSING, "gb_load_weights: missing tensor %s", G.missing.c_str());
    w.Wqkv = upload(c, concat({&wq, &wkv, &wqp, &wkvp}));
    w.bqkv = upload(c, concat({&bq, &bkv, &bqp, &bkvp}));
    w.Wb = upload(c, wb); w.bb = upload(c, bb);
    std::vector<float> hw(H);
    for (int h = 0; h < H; ++h) {
      const float xh = hwt[h];
      const float sp = (xh > 20.f) ? xh : log1pf(expf(xh));                       // nn.Softplus
      hw[h] = sp * sqrtf(1.f / (3.f * ((float)IPA_PQ * 9.f / 2.f)));             // invariant_point_attention.py:197-200
    }
    w.hw = upload(c, hw);
    w.Wout = upload(c, wo); w.bout = upload(c, bo);
    w.ln1_g = upload(c, l1g); w.ln1_b = upload(c, l1b); w.ln2_g = upload(c, l2g); w.ln2_b = upload(c, l2b);
    w.Wbb = upload(c, wbb); w.bbb = upload(c, bbb);
    std::vector<const std::vector<float>*> w123;
    wb_all.push_back(&wb); bb_all.push_back(&bb);
    for (int t = 0; t < d.n_structure_transition_layer; ++t) {
      const std::string q = pre + ".transition.layers." + std::to_string(t);
      const auto& a1 = G.get(q + ".linear_1.weight", {c_s, c_s});
      const auto& c1 = G.get(q + ".linear_1.bias", {c_s});
      const auto& a2 = G.get(q + ".linear_2.weight", {c_s, c_s});
      const auto& c2 = G.get(q + ".linear_2.bias", {c_s});
      const auto& a3 = G.get(q + ".linear_3.weight", {c_s, c_s});
      const auto& c3 = G.get(q + ".linear_3.bias", {c_s});
      if (!G.ok) return gb_fail(c, GB_ERR_MISSING, "gb_load_weights: missing tensor %s", G.missing.c_str());
      w.W1.push_back(upload(c, a1)); w.b1.push_back(upload(c, c1));
      w.W2.push_back(upload(c, a2)); w.b2.push_back(upload(c, c2));
      w.W3.push_back(upload(c, a3)); w.b3.push_back(upload(c, c3));
      w123.push_back(&a1); w123.push_back(&a2); w123.push_back(&a3);
    }
    int rc = tc_pack_struct_layer(c, l, concat({&wq, &wkv, &wqp, &wkvp}), concat({&bq, &bkv, &bqp, &bkvp}), wo, w123);
    if (rc != GB_OK) return rc;
  }
  {
    int rc = tc_pack_ipa_bias(c, wb_all, bb_all);
    if (rc != GB_OK) return rc;
  }
  for (void* p : c->weight_allocs) if (!p) return gb_fail(c, GB_ERR_CUDA, "gb_load_weights: device allocation failed");
  c->host.clear();
  c->weights_loaded = true;
  ++c->epoch;
  return GB_OK;
}

extern "C" int gb_set_schedule(gb_ctx* c, const float* alphas, const float* sqrt_alphas, const float* sqrt_omac,
                               const float* sqrt_betas) {
  if (!c || !alphas || !sqrt_alphas || !sqrt_omac || !sqrt_betas) return gb_fail(c, GB_ERR_INVALID, "gb_set_schedule: NULL argument");
  DeviceGuard dg(c->device);
  const size_t n = (size_t)c->d.n_timestep + 1;
  if (!c->alphas) {
    float* base = nullptr;
    GB_CUDA(c, cudaMalloc(&base, 4 * n * sizeof(float)));
    c->alphas = base; c->sqrt_alphas = base + n; c->sqrt_omac = base + 2 * n; c->sqrt_betas = base + 3 * n;
  }
  GB_CUDA(c, cudaMemcpy(c->alphas, alphas, n * sizeof(float), cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->sqrt_alphas, sqrt_alphas, n * sizeof(float), cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->sqrt_omac, sqrt_omac, n * sizeof(float), cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->sqrt_betas, sqrt_betas, n * sizeof(float), cudaMemcpyDefault));
  c->schedule_set = true;
  return GB_OK;
}

extern "C" int gb_set_precision(gb_ctx* c, int precision) {
  if (!c) return GB_ERR_INVALID;
  if (precision != GB_PREC_FP32 && precision != GB_PREC_BF16X3 && precision != GB_PREC_BF16)
    return gb_fail(c, GB_ERR_INVALID, "gb_set_precision: unknown precision %d", precision);
  c->precision = precision;
  return GB_OK;
}

// ------------------------------------------------------------------ features + workspace
struct Carver {
  char* base;
  size_t off = 0;
  template <typename T> T* take(size_t n) {
    off = (off + 255) & ~(size_t)255;
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

static void carve_features(gb_ctx* c, Carver& k, int B, int N) {
  const size_t R = (size_t)B * N;
  c->residue_index = k.take<int>(R);
  c->chain_index = k.take<int>(R);
  c->mask = k.take<int>(R);
  c->aatype = k.take<int>(R * 20);
  c->fsm = k.take<uint8_t>(R);
  c->iface = k.take<uint8_t>(R);
  c->fst = k.take<uint8_t>(R * N);
  c->motif_pos = k.take<float>(R * 3);
  c->s_static = k.take<float>(R * c->d.c_s);
}

static void carve_workspace(gb_ctx* c, Carver& k, int B, int N) {
  const size_t R = (size_t)B * N, M = R * N, cs = c->d.c_s;
  c->s = k.take<float>(R * cs);
  c->sA = k.take<float>(R * cs);
  c->sB = k.take<float>(R * cs);
  c->h1 = k.take<float>(R * cs);
  c->h2 = k.take<float>(R * cs);
  c->sX[0] = k.take<float>(R * cs);      // row-sharded structure net: the rows the ranks exchange, alternating per layer
  c->sX[1] = k.take<float>(R * cs);
  c->pipj = k.take<float>(R * 256);
  c->qkv = k.take<float>(R * IPA_QKV);
  c->cat = k.take<float>(R * IPA_CAT);
  c->rots_w = k.take<float>(R * 9);
  c->trans_w = k.take<float>(R * 3);
  c->z = k.take<float>(R * 3);
  c->trans_io = k.take<float>(R * 3);
  c->rots_io = k.take<float>(R * 9);
  c->t_dev = k.take<int>(B);
  c->step_dev = k.take<int>(1);
  c->p = k.take<float>(M * 128);
  c->a = k.take<float>(M * 128);
  c->b = k.take<float>(M * 128);
  c->x = k.take<float>(M * 128);
  c->g = k.take<float>(M * 128);
  if (c->d.include_tri_att) {   // a..g double as the [M][512] q|k|v|g buffer of K8
    c->ta_x = k.take<float>(M * 128);
    c->ta_bias = k.take<float>((size_t)B * N * (((size_t)N + 63) / 64 * 64) * TA_H);   // [B][4][N][Np] (tensor-core kernels) >= [M][4]
  }
}

extern "C" int gb_set_features(gb_ctx* c, const gb_features* f) {
  if (!c || !f) return gb_fail(c, GB_ERR_INVALID, "gb_set_features: NULL argument");
  if (!c->weights_loaded) return gb_fail(c, GB_ERR_INVALID, "gb_set_features: call gb_load_weights first");
  if (f->B < 1 || f->N < 3) return gb_fail(c, GB_ERR_INVALID, "gb_set_features: need B >= 1 and N >= 3 (got %d, %d)", f->B, f->N);
  if (!f->residue_index || !f->chain_index || !f->residue_mask || !f->fixed_sequence_mask || !f->fixed_structure_mask ||
      !f->interface_mask || !f->aatype || !f->atom_positions)
    return gb_fail(c, GB_ERR_INVALID, "gb_set_features: NULL feature pointer");
  if ((size_t)f->N * 13 * 4 + 64 > 200 * 1024) return gb_fail(c, GB_ERR_UNSUPPORTED, "N = %d too large for the frame kernels", f->N);
  if (c->shard_world > 1 && (f->B != 1 || f->N % (128 * c->shard_world) != 0))
    return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding over %d GPUs needs B == 1 and N a multiple of %d (got B = %d, N = %d)",
                   c->shard_world, 128 * c->shard_world, f->B, f->N);
  DeviceGuard dg(c->device);
  const int B = f->B, N = f->N;
  if (!c->features_set || B != c->B || N != c->N) {
    GB_CUDA(c, cudaDeviceSynchronize());
    Carver kf{nullptr}, kw{nullptr};
    carve_features(c, kf, B, N);
    carve_workspace(c, kw, B, N);
    const size_t tcb = tc_workspace_bytes(c, B, N);
    // grow-only: a batch that fits in what an earlier, larger batch allocated is re-carved in place (a length sweep that
    // runs its largest batches first allocates once)
    if (kf.off + 256 > c->feat_cap || kw.off + 512 + tcb > c->ws_cap) {
      // row-sharded: the peers read this workspace over NVLink (shard.cu); they let go of the old one first (collective)
      if (c->shard_world > 1 && c->ws_alloc) { const int urc = shard_unmap_workspace(c, true); if (urc != GB_OK) return urc; }
      free_features(c);
      GB_CUDA(c, cudaMalloc(&c->feat_alloc, kf.off + 256));
      cudaError_t e = cudaMalloc(&c->ws_alloc, kw.off + 512 + tcb);
      if (e != cudaSuccess) {
        cudaFree(c->feat_alloc); c->feat_alloc = nullptr;
        return gb_fail(c, GB_ERR_CUDA, "workspace allocation of %zu bytes failed: %s", kw.off + tcb, cudaGetErrorString(e));
      }
      c->feat_cap = kf.off + 256;
      c->ws_cap = kw.off + 512 + tcb;
      if (c->shard_world > 1) { const int mrc = shard_map_workspace(c); if (mrc != GB_OK) return mrc; }
    } else {
      drop_step_graph(c);
      c->features_set = false;
    }
    c->ws_bytes = kf.off + kw.off + tcb + 768;
    // GB_DEBUG_POISON=1: fill the fresh workspace with NaN patterns so that any read of a never-written location shows
    // up in the parity tests (cudaMalloc does not promise zeroed memory)
    if (const char* env = getenv("GB_DEBUG_POISON"))
      if (env[0] == '1') {
        GB_CUDA(c, cudaMemset(c->ws_alloc, 0xFF, kw.off + 512 + tcb));
        GB_CUDA(c, cudaMemset(c->feat_alloc, 0xFF, kf.off + 256));
      }
    Carver kf2{(char*)c->feat_alloc}, kw2{(char*)c->ws_alloc};
    carve_features(c, kf2, B, N);
    carve_workspace(c, kw2, B, N);
    kw2.off = (kw2.off + 255) & ~(size_t)255;
    int rc = tc_bind_workspace(c, (char*)c->ws_alloc + kw2.off, tcb, B, N);
    if (rc != GB_OK) return rc;
    c->B = B; c->N = N;
  }
  const size_t R = (size_t)B * N;
  GB_CUDA(c, cudaMemcpy(c->residue_index, f->residue_index, R * 4, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->chain_index, f->chain_index, R * 4, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->mask, f->residue_mask, R * 4, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->aatype, f->aatype, R * 20 * 4, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->fsm, f->fixed_sequence_mask, R, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->iface, f->interface_mask, R, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->fst, f->fixed_structure_mask, R * N, cudaMemcpyDefault));
  GB_CUDA(c, cudaMemcpy(c->motif_pos, f->atom_positions, R * 3 * 4, cudaMemcpyDefault));
  // has_motif: any fixed_structure_mask bit set (host check on a copy; once per batch)
  {
    std::vector<uint8_t> h(R * N);
    GB_CUDA(c, cudaMemcpy(h.data(), c->fst, R * N, cudaMemcpyDeviceToHost));
    c->has_motif = 0;
    for (uint8_t v : h) if (v) { c->has_motif = 1; break; }
  }
  // step-invariant part of SingleFeatureNet: s_static = W[:, non-time columns] . features
  const gb_dims& d = c->d;
  const int Fs = d.c_pos_emb + d.c_chain_emb + 23, F = Fs + d.c_timestep_emb;
  float* X = nullptr;
  GB_CUDA(c, cudaMalloc(&X, R * Fs * sizeof(float)));
  single_static_features_kernel<<<(unsigned)((R * Fs + 255) / 256), 256>>>(X, c->residue_index, c->chain_index, c->aatype, c->fsm,
                                                                            c->iface, (int)R, d.c_pos_emb, d.c_chain_emb,
                                                                            d.max_n_res, d.max_n_chain);
  // two GEMMs over the two non-time column blocks of W [c_s][F]: [0, c_pos+c_chain) and [F-23, F)
  const int K1 = d.c_pos_emb + d.c_chain_emb;
  sgemm_nt(0, X, Fs, c->W_single, F, nullptr, nullptr, 0, c->s_static, d.c_s, (int)R, d.c_s, K1, false);
  sgemm_nt(0, X + K1, Fs, c->W_single + K1 + d.c_timestep_emb, F, nullptr, c->s_static, d.c_s, c->s_static, d.c_s, (int)R,
           d.c_s, 23, false);
  cudaError_t e = cudaDeviceSynchronize();
  cudaFree(X);
  if (e != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "static embedding kernels failed: %s", cudaGetErrorString(e));
  c->launches += 3;
  c->features_set = true;
  ++c->epoch;
  return GB_OK;
}

// ------------------------------------------------------------------ small orchestration kernels
__global__ void init_frames_kernel(const float* __restrict__ rots, const float* __restrict__ trans, float* rots_w,
                                   float* trans_w, float rescale, int rows) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows * 9) rots_w[i] = rots[i];
  if (i < rows * 3) trans_w[i] = trans[i] * rescale;
}
// z = trans - trans_w * (1/rescale)  (model.py:184-187); optional copies of the final frames
__global__ void finish_kernel(const float* __restrict__ trans, const float* __restrict__ trans_w,
                              const float* __restrict__ rots_w, float* __restrict__ z, float* out_trans,
                              float* out_rots, float inv_rescale, int rows) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows * 3) {
    const float t = trans_w[i] * inv_rescale;
    z[i] = trans[i] - t;
    if (out_trans) out_trans[i] = t;
  }
  if (out_rots && i < rows * 9) out_rots[i] = rots_w[i];
}
__global__ void fill_int_kernel(int* p, int v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// graph-replayed reverse step: timesteps[b] = *step at the start of a step, *step -= 1 at its end
__global__ void step_begin_kernel(int* __restrict__ t_dev, const int* __restrict__ step, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) t_dev[i] = *step;
}
__global__ void step_end_kernel(int* step) { *step -= 1; }

static int check_ready(gb_ctx* c, const char* who) {
  if (!c) return GB_ERR_INVALID;
  if (!c->weights_loaded) return gb_fail(c, GB_ERR_INVALID, "%s: weights not loaded", who);
  if (!c->features_set) return gb_fail(c, GB_ERR_INVALID, "%s: features not bound (gb_set_features)", who);
  return GB_OK;
}

static int x3diag() { static const int v = [] { const char* e = getenv("GB_X3_DIAG"); return e ? atoi(e) : 0; }(); return v; }
static int run_trimul(gb_ctx* c, const TriMulW& w, bool outgoing, float* p, cudaStream_t st) {
  if (c->precision != GB_PREC_FP32 && !((x3diag() & 2) && c->precision == GB_PREC_BF16X3)) return tc_trimul(c, w, outgoing, p, st);
  const int B = c->B, N = c->N;
  const size_t M = (size_t)B * N * N;
  const unsigned tiles = (unsigned)((M + 127) / 128);
  trimul_proj_simt_kernel<<<tiles, 256, kSmemPair, st>>>(p, M, N, c->mask, w.ln_in_g, w.ln_in_b, w.Wpk, w.bpk, c->a, c->b, c->g);
  GB_LAUNCH_CHECK(c, "trimul_proj_simt");
  dim3 grid((N + 63) / 64, (N + 63) / 64, B * 128);
  if (outgoing) trimul_contract_simt_kernel<true><<<grid, 256, 0, st>>>(c->a, c->b, c->x, N);
  else trimul_contract_simt_kernel<false><<<grid, 256, 0, st>>>(c->a, c->b, c->x, N);
  GB_LAUNCH_CHECK(c, "trimul_contract_simt");
  trimul_out_simt_kernel<<<tiles, 256, kSmemPair, st>>>(c->x, c->g, p, M, N, w.ln_out_g, w.ln_out_b, w.WzT, w.bz);
  GB_LAUNCH_CHECK(c, "trimul_out_simt");
  return GB_OK;
}

// K8: p += TriangleAttention(p) (pair_transform_net.py:109-119).  bf16 and bf16x3: tcgen05 kernels (tc_triatt.cu, the latter
// with split hi | lo operands and three passes per product); fp32: the CUDA-core kernels below.
static int run_triatt(gb_ctx* c, const TriAttW& w, bool starting, float* p, cudaStream_t st) {
  if (c->precision != GB_PREC_FP32) return tc_tri_attention(c, w, starting, p, st);
  const int B = c->B, N = c->N;
  const size_t M = (size_t)B * N * N;
  if (M > (size_t)INT32_MAX) return gb_fail(c, GB_ERR_UNSUPPORTED, "triangle attention: B*N*N too large");
  if (triatt_smem_bytes(N) > 200 * 1024) return gb_fail(c, GB_ERR_UNSUPPORTED, "triangle attention: N = %d too large", N);
  float* qkvg = c->a;   // a | b | x | g are contiguous: [M][512]
  layernorm_rows_kernel<<<(unsigned)((M * 32 + 255) / 256), 256, 0, st>>>(p, c->ta_x, nullptr, w.ln_g, w.ln_b, (int)M, 128);
  GB_LAUNCH_CHECK(c, "triatt_layer_norm");
  sgemm_nt(st, c->ta_x, 128, w.Wqkvg, 128, w.bqkvg, nullptr, 0, qkvg, 512, (int)M, 512, 128, false);
  GB_LAUNCH_CHECK(c, "triatt_qkvg");
  sgemm_nt(st, c->ta_x, 128, w.Wbias, 128, nullptr, nullptr, 0, c->ta_bias, TA_H, (int)M, TA_H, 128, false);
  GB_LAUNCH_CHECK(c, "triatt_bias");
  triatt_core_kernel<<<dim3(N, B), 256, triatt_smem_bytes(N), st>>>(qkvg, c->ta_bias, c->mask, c->ta_x, N, starting ? 1 : 0);
  GB_LAUNCH_CHECK(c, "triatt_core");
  sgemm_nt(st, c->ta_x, 128, w.Wo, 128, w.bo, p, 128, p, 128, (int)M, 128, 128, false);
  GB_LAUNCH_CHECK(c, "triatt_out");
  return GB_OK;
}

static int run_transition(gb_ctx* c, const TransitionW& w, float* p, cudaStream_t st) {
  if (c->precision != GB_PREC_FP32 && !((x3diag() & 4) && c->precision == GB_PREC_BF16X3)) return tc_pair_transition(c, w, p, st);
  const size_t M = (size_t)c->B * c->N * c->N;
  const unsigned tiles = (unsigned)((M + 127) / 128);
  pair_transition_simt_kernel<<<tiles, 256, kSmemTrans, st>>>(p, M, c->N, c->mask, w.ln_g, w.ln_b, w.W1T, w.b1, w.W2T, w.b2,
                                                              c->d.pair_transition_n);
  GB_LAUNCH_CHECK(c, "pair_transition_simt");
  return GB_OK;
}

static int run_ipa(gb_ctx* c, const StructLayerW& w, const float* s, const float* p, const float* rots, const float* trans,
                   float* out, const float* residual, cudaStream_t st) {
  const int B = c->B, N = c->N, R = B * N, cs = c->d.c_s;
  sgemm_nt(st, s, cs, w.Wqkv, cs, w.bqkv, nullptr, 0, c->qkv, IPA_QKV, R, IPA_QKV, cs, false);
  GB_LAUNCH_CHECK(c, "ipa_proj");
  const int NP = IPA_H * IPA_PQ + IPA_H * (IPA_PQ + IPA_PV);
  ipa_points_to_global_kernel<<<(R * NP + 255) / 256, 256, 0, st>>>(c->qkv, rots, trans, R);
  GB_LAUNCH_CHECK(c, "ipa_points_to_global");
  const size_t smem = ((size_t)IPA_H * N + 192 + 144 + 288) * sizeof(float);
  if (smem > 200 * 1024) return gb_fail(c, GB_ERR_UNSUPPORTED, "N = %d too large for ipa_attention", N);
  // row-sharded (B == 1): attention and output projection for this rank's query rows, then the [N, c_s] result is
  // all-gathered; everything else in the structure net is O(N) and runs replicated
  const int W = c->shard_world, rows = N / W, row0 = c->shard_rank * rows;
  ipa_attention_kernel<<<dim3(rows, B), 256, smem, st>>>(c->qkv, p, rots, trans, c->mask, w.Wb, w.bb, w.hw, c->cat, N, row0);
  GB_LAUNCH_CHECK(c, "ipa_attention");
  sgemm_nt(st, c->cat + (size_t)row0 * IPA_CAT, IPA_CAT, w.Wout, IPA_CAT, w.bout, residual ? residual + (size_t)row0 * cs : nullptr, cs,
           out + (size_t)row0 * cs, cs, W > 1 ? rows : R, cs, IPA_CAT, false);
  GB_LAUNCH_CHECK(c, "ipa_out");
  if (W > 1) return shard_all_gather(c, out, (size_t)rows * cs * sizeof(float), st);
  return GB_OK;
}

int simt_points_to_global(gb_ctx* c, cudaStream_t st) {
  const int R = c->B * c->N, NP = IPA_H * IPA_PQ + IPA_H * (IPA_PQ + IPA_PV);
  ipa_points_to_global_kernel<<<(R * NP + 255) / 256, 256, 0, st>>>(c->qkv, c->rots_w, c->trans_w, R);
  GB_LAUNCH_CHECK(c, "ipa_points_to_global");
  return GB_OK;
}
int simt_backbone_update(gb_ctx* c, const StructLayerW& w, cudaStream_t st) {
  const int R = c->B * c->N;
  backbone_update_kernel<<<(R * 32 + 255) / 256, 256, 0, st>>>(c->s, w.Wbb, w.bbb, c->rots_w, c->trans_w, R, c->d.c_s);
  GB_LAUNCH_CHECK(c, "backbone_update");
  return GB_OK;
}

extern "C" int gb_frenet(gb_ctx* c, const float* trans, float* rots_out, void* stream) {
  int rc = check_ready(c, "gb_frenet");
  if (rc) return rc;
  if (!trans || !rots_out) return gb_fail(c, GB_ERR_INVALID, "gb_frenet: NULL pointer");
  DeviceGuard dg(c->device);
  const size_t smem = (size_t)c->N * 13 * 4;
  frenet_kernel<<<c->B, 256, smem, (cudaStream_t)stream>>>(trans, c->chain_index, c->mask, c->N, rots_out);
  GB_LAUNCH_CHECK(c, "frenet");
  return GB_OK;
}

extern "C" int gb_k_frenet(const float* trans, const int32_t* chain_index, const int32_t* residue_mask, int B, int N,
                           float* rots_out, void* stream) {
  if (!trans || !chain_index || !residue_mask || !rots_out || B < 1 || N < 3) return gb_fail(nullptr, GB_ERR_INVALID, "gb_k_frenet: bad arguments");
  const size_t smem = (size_t)N * 13 * 4;
  if (smem > 200 * 1024) return gb_fail(nullptr, GB_ERR_UNSUPPORTED, "gb_k_frenet: N too large");
  cudaFuncSetAttribute(frenet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  frenet_kernel<<<B, 256, smem, (cudaStream_t)stream>>>(trans, chain_index, residue_mask, N, rots_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return gb_fail(nullptr, GB_ERR_CUDA, "frenet launch failed: %s", cudaGetErrorString(e));
  return GB_OK;
}

extern "C" int gb_denoise(gb_ctx* c, const float* rots, const float* trans, const int32_t* timesteps, float* z_out,
                          const gb_aux* aux, void* stream) {
  int rc = check_ready(c, "gb_denoise");
  if (rc) return rc;
  if (!rots || !trans || !timesteps || !z_out) return gb_fail(c, GB_ERR_INVALID, "gb_denoise: NULL pointer");
  if (c->shard_world > 1 && c->precision != GB_PREC_BF16)
    return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding is built for the tensor-core (bf16) pair stack only");
  DeviceGuard dg(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  const gb_dims& d = c->d;
  const int B = c->B, N = c->N, R = B * N, cs = d.c_s;
  const size_t Rcs = (size_t)R * cs;
  ph_mark(st, "embed");
  init_frames_kernel<<<(R * 9 + 255) / 256, 256, 0, st>>>(rots, trans, c->rots_w, c->trans_w, d.rescale, R);
  GB_LAUNCH_CHECK(c, "init_frames");
  // SingleFeatureNet
  single_combine_kernel<<<(unsigned)((Rcs + 255) / 256), 256, 0, st>>>(c->s, c->s_static, c->time_table, timesteps, c->mask, B, N, cs);
  GB_LAUNCH_CHECK(c, "single_combine");
  if (aux && aux->s) GB_CUDA(c, cudaMemcpyAsync(aux->s, c->s, Rcs * 4, cudaMemcpyDeviceToDevice, st));
  if (aux && aux->states) GB_CUDA(c, cudaMemcpyAsync(aux->states, c->s, Rcs * 4, cudaMemcpyDeviceToDevice, st));
  // PairFeatureNet.  Tensor-core modes never drop to the CUDA-core kernels silently: a shape the tensor-core structure
  // path does not cover is an error (use GB_PREC_FP32).  GB_PIPJ_TC=0 / GB_PE_TC=0 are explicit debugging switches.
  const bool tcmode = c->precision != GB_PREC_FP32;
  if (tcmode && !(tc_struct_supported(c, N) && tc_pipj_supported(c)))
    return gb_fail(c, GB_ERR_UNSUPPORTED, "tensor-core modes need c_s %% 128 == 0, c_s <= 512, N <= 1024 and n_structure_layer <= 8 (got c_s = %d, N = %d); use fp32",
                   cs, N);
  static const int pipj_tc = [] { const char* e = getenv("GB_PIPJ_TC"); return e ? atoi(e) : 1; }();
  if (pipj_tc && tcmode) {
    if ((rc = tc_pair_pipj(c, st))) return rc;
  } else {
    sgemm_nt(st, c->s, cs, c->W_pipj, cs, nullptr, nullptr, 0, c->pipj, 256, R, 256, cs, false);
    GB_LAUNCH_CHECK(c, "pair_pipj");
  }
  const int W = c->shard_world, rows_l = N / W, row0 = c->shard_rank * rows_l;   // W == 1: all rows
  static const int pe_tc = [] { const char* e = getenv("GB_PE_TC"); return e ? atoi(e) : 1; }();
  if (c->precision == GB_PREC_BF16 && !tc_pair_embed_supported(c))
    return gb_fail(c, GB_ERR_UNSUPPORTED, "bf16 pair embedding: packed feature weights missing (template_dist_n_bin must be 37)");
  if (pe_tc && c->precision == GB_PREC_BF16) {
    if ((rc = tc_pair_embed(c, aux ? aux->quat_follow : nullptr, rows_l, row0, st))) return rc;
  } else {
    // fp32 and bf16x3: the 41- / 78-term feature dot on CUDA cores in fp32 (one launch per eval)
  pair_embed_kernel<37><<<dim3(rows_l, B), 128, 0, st>>>(c->p, c->pipj, c->rots_w, c->trans_w, c->residue_index, c->chain_index, c->mask,
                                                    c->fsm, c->fst, c->motif_pos, c->WrelT, c->WtT, c->WmT,
                                                    aux ? aux->quat_follow : nullptr, N, d.relpos_k, d.template_dist_min,
                                                    d.template_dist_step, c->has_motif, row0);
  GB_LAUNCH_CHECK(c, "pair_embed");
  }
  // PairTransformNet
  for (int l = 0; l < d.n_pair_transform_layer; ++l) {
    if (d.include_mul_update) {
      if ((rc = run_trimul(c, c->pair[l].mul_out, true, c->p, st))) return rc;
      if ((rc = run_trimul(c, c->pair[l].mul_in, false, c->p, st))) return rc;
    }
    if (d.include_tri_att) {
      if (W > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "row sharding: triangle attention is not sharded in this revision");
      if ((rc = run_triatt(c, c->pair[l].att_start, true, c->p, st))) return rc;
      if ((rc = run_triatt(c, c->pair[l].att_end, false, c->p, st))) return rc;
    }
    ph_mark(st, "K7 transition");
    if ((rc = run_transition(c, c->pair[l].tr, c->p, st))) return rc;
  }
  ph_mark(st, "barrier");
  // row-sharded: no rank leaves the pair stack while a peer's last K6 may still be reading its x (its memory stays valid for
  // the peers for as long as work of this eval is in flight anywhere)
  if (W > 1 && d.include_mul_update && (rc = shard_barrier(c, st))) return rc;
  if (aux && aux->p) {
    if (W > 1 && (rc = shard_all_gather(c, c->p, (size_t)rows_l * N * 128 * 4, st))) return rc;   // every rank returns the whole map
    GB_CUDA(c, cudaMemcpyAsync(aux->p, c->p, (size_t)R * N * 128 * 4, cudaMemcpyDeviceToDevice, st));
  }
  // StructureNet
  float* s = c->s;
  int state_idx = 1;
  if (tcmode && !((x3diag() & 1) && c->precision == GB_PREC_BF16X3)) {
    if ((rc = tc_structure_net(c, aux ? aux->states : nullptr, st))) return rc;
  } else
  for (int blk = 0; blk < d.n_structure_block; ++blk)
    for (int l = 0; l < d.n_structure_layer; ++l) {
      const StructLayerW& w = c->st[l];
      if ((rc = run_ipa(c, w, s, c->p, c->rots_w, c->trans_w, c->sA, s, st))) return rc;           // sA = s + ipa
      layernorm_rows_kernel<<<(R * 32 + 255) / 256, 256, 0, st>>>(c->sA, c->sB, nullptr, w.ln1_g, w.ln1_b, R, cs);
      GB_LAUNCH_CHECK(c, "ipa_layer_norm");
      float* cur = c->sB;
      float* other = c->sA;
      for (int t = 0; t < d.n_structure_transition_layer; ++t) {
        sgemm_nt(st, cur, cs, w.W1[t], cs, w.b1[t], nullptr, 0, c->h1, cs, R, cs, cs, true);
        GB_LAUNCH_CHECK(c, "st_linear_1");
        sgemm_nt(st, c->h1, cs, w.W2[t], cs, w.b2[t], nullptr, 0, c->h2, cs, R, cs, cs, true);
        GB_LAUNCH_CHECK(c, "st_linear_2");
        sgemm_nt(st, c->h2, cs, w.W3[t], cs, w.b3[t], cur, cs, other, cs, R, cs, cs, false);
        GB_LAUNCH_CHECK(c, "st_linear_3");
        float* tmp = cur; cur = other; other = tmp;
      }
      float* st_out = (aux && aux->states) ? aux->states + (size_t)state_idx * Rcs : nullptr;
      layernorm_rows_kernel<<<(R * 32 + 255) / 256, 256, 0, st>>>(cur, c->s, st_out, w.ln2_g, w.ln2_b, R, cs);
      GB_LAUNCH_CHECK(c, "transition_layer_norm");
      ++state_idx;
      s = c->s;
      backbone_update_kernel<<<(R * 32 + 255) / 256, 256, 0, st>>>(s, w.Wbb, w.bbb, c->rots_w, c->trans_w, R, cs);
      GB_LAUNCH_CHECK(c, "backbone_update");
    }
  finish_kernel<<<(R * 9 + 255) / 256, 256, 0, st>>>(trans, c->trans_w, c->rots_w, z_out, aux ? aux->out_trans : nullptr,
                                                     aux ? aux->out_rots : nullptr, 1.f / d.rescale, R);
  GB_LAUNCH_CHECK(c, "finish");
  ph_report(c, st);
  return GB_OK;
}

static int launch_ddpm(gb_ctx* c, float* trans, float* rots, const float* z, int step, float scale, const float* noise,
                       size_t noise_step_stride, uint64_t seed, cudaStream_t st, const int* step_dev = nullptr) {
  const size_t smem = (size_t)c->N * 13 * 4;
  ddpm_step_kernel<<<c->B, 256, smem, st>>>(trans, rots, z, noise, c->chain_index, c->mask, c->alphas, c->sqrt_alphas,
                                            c->sqrt_omac, c->sqrt_betas, step, step_dev, noise_step_stride, scale, seed, c->N);
  GB_LAUNCH_CHECK(c, "ddpm_step");
  return GB_OK;
}

extern "C" int gb_p_sample(gb_ctx* c, float* trans_inout, float* rots_inout, const float* z, int step, float scale,
                           const float* noise, uint64_t seed, void* stream) {
  int rc = check_ready(c, "gb_p_sample");
  if (rc) return rc;
  if (!c->schedule_set) return gb_fail(c, GB_ERR_INVALID, "gb_p_sample: schedule not set");
  if (!trans_inout || !rots_inout || !z) return gb_fail(c, GB_ERR_INVALID, "gb_p_sample: NULL pointer");
  if (step < 1 || step > c->d.n_timestep) return gb_fail(c, GB_ERR_INVALID, "gb_p_sample: step %d outside [1, %d]", step, c->d.n_timestep);
  DeviceGuard dg(c->device);
  return launch_ddpm(c, trans_inout, rots_inout, z, step, scale, noise, 0, seed, (cudaStream_t)stream);
}

extern "C" int gb_sample_loop(gb_ctx* c, float* trans_inout, float* rots_inout, int t_start, int t_end, float scale,
                              const float* noise_tape, uint64_t seed, void* stream) {
  int rc = check_ready(c, "gb_sample_loop");
  if (rc) return rc;
  if (!c->schedule_set) return gb_fail(c, GB_ERR_INVALID, "gb_sample_loop: schedule not set");
  if (!trans_inout || !rots_inout) return gb_fail(c, GB_ERR_INVALID, "gb_sample_loop: NULL pointer");
  if (t_end < 1 || t_start > c->d.n_timestep || t_start < t_end)
    return gb_fail(c, GB_ERR_INVALID, "gb_sample_loop: bad step range [%d, %d]", t_start, t_end);
  DeviceGuard dg(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t stride = (size_t)c->B * c->N * 3;
  // Three or more steps on one GPU: capture ONE step (its ~105 launches, programmatic-dependent-launch edges included) as
  // a CUDA graph whose kernels read the step index from device memory, and replay it.  Short chains at L <= 128 are
  // otherwise bound by the host's launch rate, not by the GPU.  The capture runs on a private stream (the caller's may be
  // the legacy default stream, which cannot capture); replays go to the caller's stream.  GB_STEP_GRAPH=0 keeps the plain
  // launch loop (also used for row-sharded runs: NCCL calls stay outside graphs here).
  static const int graph_mode = [] { const char* e = getenv("GB_STEP_GRAPH"); return e ? atoi(e) : 1; }();
  // Short calls (one step per call, as the host-buffer entry point gb_sample_host is typically driven) use the graph
  // from the second call with the same arguments on: a capture costs about as much as launching the step once.
  gb_ctx::StepKey key;
  key.trans = trans_inout; key.rots = rots_inout; key.tape = noise_tape; key.scale = scale; key.seed = seed;
  key.epoch = c->epoch; key.precision = c->precision;
  const bool repeated = key == c->last_key;
  c->last_key = key;
  if (graph_mode && c->shard_world == 1 && (t_start - t_end + 1 >= 3 || repeated)) {
    if (!c->step_graph || !(key == c->step_key)) {
      if (c->step_graph) cudaGraphExecDestroy((cudaGraphExec_t)c->step_graph);
      c->step_graph = nullptr;
      const int64_t l0 = c->launches;
      if (!c->cap_stream) {
        cudaStream_t ns = nullptr;
        GB_CUDA(c, cudaStreamCreateWithFlags(&ns, cudaStreamNonBlocking));
        c->cap_stream = ns;
      }
      cudaStream_t cs = (cudaStream_t)c->cap_stream;
      GB_CUDA(c, cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
      step_begin_kernel<<<(c->B + 255) / 256, 256, 0, cs>>>(c->t_dev, c->step_dev, c->B);
      rc = cudaGetLastError() == cudaSuccess ? GB_OK : gb_fail(c, GB_ERR_CUDA, "step_begin launch failed during capture");
      c->launches++;
      if (rc == GB_OK) rc = gb_denoise(c, rots_inout, trans_inout, c->t_dev, c->z, nullptr, cs);
      if (rc == GB_OK) rc = launch_ddpm(c, trans_inout, rots_inout, c->z, 0, scale, noise_tape, noise_tape ? stride : 0, seed, cs, c->step_dev);
      if (rc == GB_OK) {
        step_end_kernel<<<1, 1, 0, cs>>>(c->step_dev);
        c->launches++;
      }
      cudaGraph_t graph = nullptr;
      const cudaError_t ce = cudaStreamEndCapture(cs, &graph);
      c->step_launches = c->launches - l0;
      c->launches = l0;                                  // nothing ran yet: replays are counted below
      if (rc != GB_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
      if (ce != cudaSuccess || !graph) return gb_fail(c, GB_ERR_CUDA, "gb_sample_loop: stream capture failed: %s", cudaGetErrorString(ce));
      cudaGraphExec_t exec = nullptr;
      const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
      cudaGraphDestroy(graph);
      if (ie != cudaSuccess) return gb_fail(c, GB_ERR_CUDA, "gb_sample_loop: cudaGraphInstantiate failed: %s", cudaGetErrorString(ie));
      c->step_graph = exec;
      c->step_key = key;
    }
    fill_int_kernel<<<1, 32, 0, st>>>(c->step_dev, t_start, 1);
    GB_LAUNCH_CHECK(c, "set_step");
    for (int step = t_start; step >= t_end; --step) {
      GB_CUDA(c, cudaGraphLaunch((cudaGraphExec_t)c->step_graph, st));
      c->launches += c->step_launches;
    }
    return GB_OK;
  }
  for (int step = t_start; step >= t_end; --step) {
    fill_int_kernel<<<(c->B + 255) / 256, 256, 0, st>>>(c->t_dev, step, c->B);
    GB_LAUNCH_CHECK(c, "fill_timesteps");
    if ((rc = gb_denoise(c, rots_inout, trans_inout, c->t_dev, c->z, nullptr, stream))) return rc;
    if ((rc = launch_ddpm(c, trans_inout, rots_inout, c->z, step, scale, noise_tape, noise_tape ? stride : 0, seed, st))) return rc;
  }
  return GB_OK;
}

extern "C" int gb_sample_host(gb_ctx* c, const float* x_T_host, float* x_out_host, int t_start, int t_end, float scale,
                              const float* noise_tape_host, uint64_t seed) {
  int rc = check_ready(c, "gb_sample_host");
  if (rc) return rc;
  if (!x_T_host || !x_out_host) return gb_fail(c, GB_ERR_INVALID, "gb_sample_host: NULL pointer");
  if (!c->schedule_set) return gb_fail(c, GB_ERR_INVALID, "gb_sample_host: schedule not set");
  if (t_end < 1 || t_start > c->d.n_timestep || t_start < t_end)      // before any copy sized from the range
    return gb_fail(c, GB_ERR_INVALID, "gb_sample_host: bad step range [%d, %d]", t_start, t_end);
  DeviceGuard dg(c->device);
  const size_t stride = (size_t)c->B * c->N * 3;
  cudaStream_t st = 0;
  GB_CUDA(c, cudaMemcpyAsync(c->trans_io, x_T_host, stride * 4, cudaMemcpyHostToDevice, st));
  const float* tape = nullptr;
  if (noise_tape_host) {
    const size_t need = ((size_t)c->d.n_timestep + 1) * stride * 4;
    if (need > c->tape_bytes) {
      if (c->tape_alloc) cudaFree(c->tape_alloc);
      c->tape_alloc = nullptr; c->tape_bytes = 0;
      GB_CUDA(c, cudaMalloc(&c->tape_alloc, need));
      c->tape_bytes = need;
    }
    // only the slices the loop reads: steps t_end..t_start
    const size_t lo = (size_t)t_end * stride, cnt = (size_t)(t_start - t_end + 1) * stride;
    GB_CUDA(c, cudaMemcpyAsync((float*)c->tape_alloc + lo, noise_tape_host + lo, cnt * 4, cudaMemcpyHostToDevice, st));
    tape = (const float*)c->tape_alloc;
  }
  if ((rc = gb_frenet(c, c->trans_io, c->rots_io, st))) return rc;
  if ((rc = gb_sample_loop(c, c->trans_io, c->rots_io, t_start, t_end, scale, tape, seed, st))) return rc;
  GB_CUDA(c, cudaMemcpyAsync(x_out_host, c->trans_io, stride * 4, cudaMemcpyDeviceToHost, st));
  GB_CUDA(c, cudaStreamSynchronize(st));
  return GB_OK;
}

// ------------------------------------------------------------------ per-kernel entry points
extern "C" int gb_k_trimul(gb_ctx* c, int layer, int outgoing, float* p, void* stream) {
  int rc = check_ready(c, "gb_k_trimul");
  if (rc) return rc;
  DeviceGuard dg(c->device);
  if (!p || layer < 0 || layer >= c->d.n_pair_transform_layer || !c->d.include_mul_update) return gb_fail(c, GB_ERR_INVALID, "gb_k_trimul: bad arguments");
  return run_trimul(c, outgoing ? c->pair[layer].mul_out : c->pair[layer].mul_in, outgoing != 0, p, (cudaStream_t)stream);
}

extern "C" int gb_k_trimul_stage(gb_ctx* c, int layer, int outgoing, int stage, float* p, void* stream) {
  int rc = check_ready(c, "gb_k_trimul_stage");
  if (rc) return rc;
  DeviceGuard dg(c->device);
  if (!p || layer < 0 || layer >= c->d.n_pair_transform_layer || !c->d.include_mul_update || stage < 1 || stage > 3)
    return gb_fail(c, GB_ERR_INVALID, "gb_k_trimul_stage: bad arguments");
  if (c->precision == GB_PREC_FP32) return gb_fail(c, GB_ERR_UNSUPPORTED, "gb_k_trimul_stage: tensor-core modes (bf16, bf16x3) only");
  return tc_trimul_stage(c, outgoing ? c->pair[layer].mul_out : c->pair[layer].mul_in, outgoing != 0, stage, p, (cudaStream_t)stream);
}

extern "C" int gb_k_pair_transition(gb_ctx* c, int layer, float* p, void* stream) {
  int rc = check_ready(c, "gb_k_pair_transition");
  if (rc) return rc;
  DeviceGuard dg(c->device);
  if (!p || layer < 0 || layer >= c->d.n_pair_transform_layer) return gb_fail(c, GB_ERR_INVALID, "gb_k_pair_transition: bad arguments");
  return run_transition(c, c->pair[layer].tr, p, (cudaStream_t)stream);
}

extern "C" int gb_k_tri_attention(gb_ctx* c, int layer, int starting, float* p, void* stream) {
  int rc = check_ready(c, "gb_k_tri_attention");
  if (rc) return rc;
  DeviceGuard dg(c->device);
  if (!c->d.include_tri_att) return gb_fail(c, GB_ERR_INVALID, "gb_k_tri_attention: include_tri_att is off in this context");
  if (!p || layer < 0 || layer >= c->d.n_pair_transform_layer) return gb_fail(c, GB_ERR_INVALID, "gb_k_tri_attention: bad arguments");
  return run_triatt(c, starting ? c->pair[layer].att_start : c->pair[layer].att_end, starting != 0, p, (cudaStream_t)stream);
}

extern "C" int gb_k_ipa(gb_ctx* c, int layer, const float* s, const float* p, const float* rots, const float* trans,
                        float* out, void* stream) {
  int rc = check_ready(c, "gb_k_ipa");
  if (rc) return rc;
  DeviceGuard dg(c->device);
  if (!s || !p || !rots || !trans || !out || layer < 0 || layer >= c->d.n_structure_layer) return gb_fail(c, GB_ERR_INVALID, "gb_k_ipa: bad arguments");
  return run_ipa(c, c->st[layer], s, p, rots, trans, out, nullptr, (cudaStream_t)stream);
}

// gb_k_ipa without any allocation and without the context's workspace as scratch: the same kernels as run_ipa (fp32 CUDA
// cores, one GPU) on the library's static scratch array.  Nothing in the scratch has to start at zero: the projection GEMM
// writes all 1152 columns of every row before they are read, the attention kernel all 2112 columns of the concat row.
extern "C" int gb_k_ipa_static(gb_ctx* c, int layer, const float* s, const float* p, const float* rots, const float* trans,
                               float* out, void* stream) {
  int rc = check_ready(c, "gb_k_ipa_static");
  if (rc) return rc;
  if (!s || !p || !rots || !trans || !out || layer < 0 || layer >= c->d.n_structure_layer) return gb_fail(c, GB_ERR_INVALID, "gb_k_ipa_static: bad arguments");
  const int B = c->B, N = c->N, R = B * N, cs = c->d.c_s;
  if (R > GB_K_IPA_STATIC_MAX_ROWS)
    return gb_fail(c, GB_ERR_UNSUPPORTED, "gb_k_ipa_static: B * N = %d rows exceed the static scratch (GB_K_IPA_STATIC_MAX_ROWS = %d); use gb_k_ipa",
                   R, GB_K_IPA_STATIC_MAX_ROWS);
  if (c->shard_world > 1) return gb_fail(c, GB_ERR_UNSUPPORTED, "gb_k_ipa_static: not available on a row-sharded context");
  if (!c->static_scratch) return gb_fail(c, GB_ERR_INVALID, "gb_k_ipa_static: static scratch address not resolved");
  const size_t smem = ((size_t)IPA_H * N + 192 + 144 + 288) * sizeof(float);
  if (smem > 200 * 1024) return gb_fail(c, GB_ERR_UNSUPPORTED, "gb_k_ipa_static: N = %d too large for ipa_attention", N);
  DeviceGuard dg(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  const StructLayerW& w = c->st[layer];
  float* qkv = c->static_scratch;                                   // [R][1152]
  float* cat = c->static_scratch + (size_t)R * IPA_QKV;             // [R][2112]
  sgemm_nt(st, s, cs, w.Wqkv, cs, w.bqkv, nullptr, 0, qkv, IPA_QKV, R, IPA_QKV, cs, false);
  GB_LAUNCH_CHECK(c, "ipa_proj(static)");
  const int NP = IPA_H * IPA_PQ + IPA_H * (IPA_PQ + IPA_PV);
  ipa_points_to_global_kernel<<<(R * NP + 255) / 256, 256, 0, st>>>(qkv, rots, trans, R);
  GB_LAUNCH_CHECK(c, "ipa_points_to_global(static)");
  ipa_attention_kernel<<<dim3(N, B), 256, smem, st>>>(qkv, p, rots, trans, c->mask, w.Wb, w.bb, w.hw, cat, N, 0);
  GB_LAUNCH_CHECK(c, "ipa_attention(static)");
  sgemm_nt(st, cat, IPA_CAT, w.Wout, IPA_CAT, w.bout, nullptr, 0, out, cs, R, cs, IPA_CAT, false);
  GB_LAUNCH_CHECK(c, "ipa_out(static)");
  return GB_OK;
}
