// C ABI (include/lutldpc.h) over the decoder object.  No torch types, no exceptions across the ABI.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "decoder.hpp"
#include "design.hpp"

using namespace lutldpc;

static thread_local std::string g_create_error;

#define CKH(call, where)                                   \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) return h->cuda_fail(e_, where); \
  } while (0)

extern "C" {

int lutldpc_abi_version(void) { return LUTLDPC_ABI_VERSION; }

int lutldpc_device_count(void) {
  int n = 0;
  return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}
int lutldpc_set_device(int device) {
  if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; return LUTLDPC_ERR_CUDA; }
  return LUTLDPC_OK;
}

const char* lutldpc_last_error(const lutldpc_decoder* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

static int finish_create(lutldpc_decoder* h, lutldpc_decoder** out) {
  int rc = h->init_device();
  if (rc != LUTLDPC_OK) {
    g_create_error = h->err;
    delete h;
    return rc;
  }
  *out = h;
  return LUTLDPC_OK;
}

int lutldpc_create_from_bundle(const char* bundle_path, lutldpc_decoder** out) {
  if (!bundle_path || !out) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  *out = nullptr;
  lutldpc_decoder* h = new (std::nothrow) lutldpc_decoder();
  if (!h) { g_create_error = "out of memory"; return LUTLDPC_ERR_ARG; }
  std::string err;
  if (!load_decoder_file(bundle_path, h->m, err)) {
    g_create_error = err;
    delete h;
    return LUTLDPC_ERR_MODEL;
  }
  return finish_create(h, out);
}

int lutldpc_create(const lutldpc_desc* d, lutldpc_decoder** out) {
  if (!d || !out) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  *out = nullptr;
  if (!d->dv_vec || !d->dc_vec || !d->cn_msg_idx || !d->Nq_Msg || !d->reuse_vec || !d->qb_Cha || !d->var_tree_string ||
      d->nvar <= 0 || d->nchk <= 0 || d->num_edges <= 0 || d->max_iters <= 0 || d->n_qb_Msg < 0 ||
      (d->n_qb_Msg > 0 && !d->qb_Msg)) {
    g_create_error = "lutldpc_desc: missing array or non-positive size";
    return LUTLDPC_ERR_ARG;
  }
  lutldpc_decoder* h = new (std::nothrow) lutldpc_decoder();
  if (!h) { g_create_error = "out of memory"; return LUTLDPC_ERR_ARG; }
  DecoderModel& m = h->m;
  m.nvar = d->nvar; m.nchk = d->nchk; m.nchk_lin_indep = d->nchk_lin_indep; m.num_edges = d->num_edges;
  m.max_iters = d->max_iters; m.Nq_Cha = d->Nq_Cha; m.minLUT = d->min_lut != 0;
  m.dv.assign(d->dv_vec, d->dv_vec + d->nvar);
  m.dc.assign(d->dc_vec, d->dc_vec + d->nchk);
  m.cn_msg_idx.assign(d->cn_msg_idx, d->cn_msg_idx + d->num_edges);
  m.Nq_Msg.assign(d->Nq_Msg, d->Nq_Msg + d->max_iters);
  m.reuse.assign(d->reuse_vec, d->reuse_vec + d->max_iters);
  std::string err;
  if (d->Nq_Cha < 2 || d->Nq_Cha > 16) err = "Nq_Cha out of range";
  if (err.empty()) {
    if (d->Nq_Cha_2_Nq_Msg_map) m.cha2msg.assign(d->Nq_Cha_2_Nq_Msg_map, d->Nq_Cha_2_Nq_Msg_map + d->Nq_Cha);
    m.qb_Cha.assign(d->qb_Cha, d->qb_Cha + d->Nq_Cha - 1);
    if (d->n_qb_Msg) m.qb_Msg.assign(d->qb_Msg, d->qb_Msg + d->n_qb_Msg);
    if (!parse_tree_array(d->var_tree_string, d->var_tree_len, m.var_trees, err)) err = "var trees: " + err;
    m.var_tree_text.assign(d->var_tree_string, d->var_tree_len);
    if (d->chk_tree_string && d->chk_tree_len) m.chk_tree_text.assign(d->chk_tree_string, d->chk_tree_len);
  }
  if (err.empty() && d->chk_tree_string && d->chk_tree_len) {
    std::string e2;
    if (!parse_tree_array(d->chk_tree_string, d->chk_tree_len, m.chk_trees, e2) && !m.minLUT) err = "chk trees: " + e2;
  }
  if (err.empty() && d->gen_parity && d->gen_rows > 0 && d->gen_cols > 0) {
    m.gen_rows = d->gen_rows; m.gen_cols = d->gen_cols;
    m.gen.assign(d->gen_parity, d->gen_parity + (size_t)d->gen_rows * d->gen_cols);
  }
  if (err.empty()) finalize_model(m, err);
  if (!err.empty()) {
    g_create_error = err;
    delete h;
    return LUTLDPC_ERR_MODEL;
  }
  return finish_create(h, out);
}

// ------------------------------------------------------------------------------------ LUT design (host only)
int lutldpc_design_luts(const lutldpc_design_params* p, lutldpc_design_result* out) {
  if (!p || !out) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  memset(out, 0, sizeof *out);
  if (!p->dv_vec || !p->dc_vec || !p->Nq_Msg || p->nvar <= 0 || p->nchk <= 0 || p->max_iters <= 0) {
    g_create_error = "lutldpc_design_params: missing array or non-positive size";
    return LUTLDPC_ERR_ARG;
  }
  design::Params dp;
  dp.dv.assign(p->dv_vec, p->dv_vec + p->nvar);
  dp.dc.assign(p->dc_vec, p->dc_vec + p->nchk);
  dp.max_iters = p->max_iters; dp.Nq_Cha = p->Nq_Cha;
  dp.Nq_Msg.assign(p->Nq_Msg, p->Nq_Msg + p->max_iters);
  if (p->reuse_vec) dp.reuse_vec.assign(p->reuse_vec, p->reuse_vec + p->max_iters);
  dp.min_lut = p->min_lut != 0; dp.sigma2 = p->sigma2;
  if (p->tree_method) dp.tree_method = p->tree_method;
  design::Result r;
  std::string err;
  if (!design::design_luts(dp, r, err)) { g_create_error = err; return LUTLDPC_ERR_MODEL; }
  auto dup = [](const void* src, size_t bytes) -> void* {
    void* d = malloc(bytes ? bytes : 1);
    if (d && bytes) memcpy(d, src, bytes);
    return d;
  };
  out->n_qb_Cha = (int32_t)r.qb_Cha.size(); out->n_qb_Msg = (int32_t)r.qb_Msg.size(); out->n_map = (int32_t)r.cha2msg.size();
  out->qb_Cha = (double*)dup(r.qb_Cha.data(), r.qb_Cha.size() * 8);
  out->qb_Msg = (double*)dup(r.qb_Msg.data(), r.qb_Msg.size() * 8);
  std::vector<int32_t> m32(r.cha2msg.begin(), r.cha2msg.end());
  out->Nq_Cha_2_Nq_Msg_map = (int32_t*)dup(m32.data(), m32.size() * 4);
  out->var_tree_len = r.var_tree_string.size(); out->chk_tree_len = r.chk_tree_string.size();
  out->var_tree_string = (char*)dup(r.var_tree_string.data(), r.var_tree_string.size());
  out->chk_tree_string = (char*)dup(r.chk_tree_string.data(), r.chk_tree_string.size());
  if (!out->qb_Cha || !out->qb_Msg || !out->Nq_Cha_2_Nq_Msg_map || !out->var_tree_string || !out->chk_tree_string) {
    lutldpc_design_free(out);
    g_create_error = "out of memory";
    return LUTLDPC_ERR_ARG;
  }
  return LUTLDPC_OK;
}

void lutldpc_design_free(lutldpc_design_result* r) {
  if (!r) return;
  free(r->qb_Cha); free(r->qb_Msg); free(r->Nq_Cha_2_Nq_Msg_map); free(r->var_tree_string); free(r->chk_tree_string);
  memset(r, 0, sizeof *r);
}

// ------------------------------------------------------------------------------------ graphs and generators
int lutldpc_alist_graph(const char* alist_path, int rank_mode, lutldpc_graph* out) {
  if (!alist_path || !out) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  memset(out, 0, sizeof *out);
  ParityRows H;
  std::string err;
  if (!load_alist_file(alist_path, H, err)) { g_create_error = err; return LUTLDPC_ERR_MODEL; }
  FlatGraph g;
  flatten_graph(H, g);
  int rank = -1;
  if (rank_mode == 1) rank = gf2_rank_host(H);
  else if (rank_mode == 2) {
    Gf2Result r;
    if (!gf2_systematic_device(H, r, false, nullptr, err)) { g_create_error = err; return LUTLDPC_ERR_CUDA; }
    rank = r.rank;
  } else if (rank_mode != 0) { g_create_error = "rank_mode must be 0 (none), 1 (host) or 2 (device)"; return LUTLDPC_ERR_ARG; }
  out->nvar = g.nvar; out->nchk = g.nchk; out->num_edges = g.num_edges; out->nchk_lin_indep = rank;
  out->dv_vec = (int32_t*)malloc(sizeof(int32_t) * std::max(1, g.nvar));
  out->dc_vec = (int32_t*)malloc(sizeof(int32_t) * std::max(1, g.nchk));
  out->cn_msg_idx = (int32_t*)malloc(sizeof(int32_t) * std::max(1, g.num_edges));
  if (!out->dv_vec || !out->dc_vec || !out->cn_msg_idx) { lutldpc_graph_free(out); g_create_error = "out of memory"; return LUTLDPC_ERR_ARG; }
  memcpy(out->dv_vec, g.dv.data(), sizeof(int32_t) * g.nvar);
  memcpy(out->dc_vec, g.dc.data(), sizeof(int32_t) * g.nchk);
  memcpy(out->cn_msg_idx, g.cn_msg_idx.data(), sizeof(int32_t) * g.num_edges);
  return LUTLDPC_OK;
}

void lutldpc_graph_free(lutldpc_graph* g) {
  if (!g) return;
  free(g->dv_vec); free(g->dc_vec); free(g->cn_msg_idx);
  memset(g, 0, sizeof *g);
}

int lutldpc_create_systematic(const lutldpc_decoder* src, lutldpc_decoder** out, int32_t* col_perm) {
  if (!src || !out) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  *out = nullptr;
  if (cudaSetDevice(src->device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; return LUTLDPC_ERR_CUDA; }
  ParityRows H, Hp;
  std::string err;
  const DecoderModel& sm = src->m;
  if (!rows_from_graph(sm.nvar, sm.nchk, sm.dv, sm.dc, sm.cn_msg_idx, H, err)) { g_create_error = err; return LUTLDPC_ERR_MODEL; }
  Gf2Result r;
  if (!gf2_systematic_device(H, r, true, nullptr, err)) { g_create_error = err; return LUTLDPC_ERR_CUDA; }
  if (r.rank == 0 || r.rank >= sm.nvar) {
    if (r.d_P) cudaFree(r.d_P);
    g_create_error = "parity-check matrix has no information columns";
    return LUTLDPC_ERR_MODEL;
  }
  lutldpc_decoder* h = new (std::nothrow) lutldpc_decoder();
  if (!h) { cudaFree(r.d_P); g_create_error = "out of memory"; return LUTLDPC_ERR_ARG; }
  h->m = sm;
  permute_columns(H, r.col_perm, Hp);
  FlatGraph g;
  flatten_graph(Hp, g);
  h->m.dv = g.dv; h->m.dc = g.dc; h->m.cn_msg_idx = g.cn_msg_idx;
  h->m.nchk_lin_indep = r.rank;
  h->m.gen.clear(); h->m.gen_rows = h->m.gen_cols = 0;
  if (!finalize_model(h->m, err)) { cudaFree(r.d_P); delete h; g_create_error = err; return LUTLDPC_ERR_MODEL; }
  h->d_gen_bits = r.d_P;
  h->gen_bits_wp = r.P_words_per_row; h->gen_bits_rows = r.rank; h->gen_bits_cols = sm.nvar - r.rank;
  int rc = finish_create(h, out);
  if (rc != LUTLDPC_OK) return rc;
  (*out)->psc = src->psc; (*out)->pisc = src->pisc; (*out)->init_mode = src->init_mode;
  if (col_perm) memcpy(col_perm, r.col_perm.data(), sizeof(int32_t) * sm.nvar);
  return LUTLDPC_OK;
}

int lutldpc_save_code(const lutldpc_decoder* h, const char* path) {
  if (!h || !path) return LUTLDPC_ERR_ARG;
  std::string err;
  if (!save_it_codec_file(path, h->m, err)) {
    const_cast<lutldpc_decoder*>(h)->err = err;
    return LUTLDPC_ERR_STATE;
  }
  return LUTLDPC_OK;
}

int lutldpc_convert_codec(const char* in_path, const char* out_it_path) {
  if (!in_path || !out_it_path) { g_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  DecoderModel m;
  std::string err;
  if (!load_decoder_file(in_path, m, err)) { g_create_error = err; return LUTLDPC_ERR_MODEL; }
  if (!save_it_codec_file(out_it_path, m, err)) { g_create_error = err; return LUTLDPC_ERR_STATE; }
  return LUTLDPC_OK;
}

void lutldpc_destroy(lutldpc_decoder* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  delete h;
}

int lutldpc_get_info(const lutldpc_decoder* h, lutldpc_info* info) {
  if (!h || !info) return LUTLDPC_ERR_ARG;
  const DecoderModel& m = h->m;
  memset(info, 0, sizeof *info);
  info->nvar = m.nvar; info->nchk = m.nchk; info->nchk_lin_indep = m.nchk_lin_indep; info->ninfo = m.ninfo();
  info->num_edges = m.num_edges; info->max_iters = h->max_iters; info->Nq_Cha = m.Nq_Cha; info->min_lut = m.minLUT;
  info->has_gen = h->has_generator(); info->n_degree_classes = (int)h->vn_classes.size(); info->rate = m.rate();
  int nvc = (int)h->vn_classes.size();
  for (int c = 0; c < nvc; ++c) {  // first iteration slot is representative (shapes rarely change per slot)
    info->lookups_per_frame_iter += h->var_progs[c].n_ops * h->vn_classes[c].n;
    info->lookups_per_frame_iter_naive += h->var_progs[c].n_naive * h->vn_classes[c].n;
  }
  for (int t = 0; t + 1 < h->n_iter_slots; ++t)
    for (int c = 0; c < nvc; ++c) {
      if (!h->vn_classes[c].n) continue;
      ++info->n_var_programs;
      info->n_fast_var_programs += h->var_progs[(size_t)t * nvc + c].fast_dv != 0;
    }
  info->onchip = h->onchip_ok ? 1 : 0;
  info->onchip_smem_bytes = h->onchip_fits ? (int32_t)h->oc.smem_bytes : 0;
  return LUTLDPC_OK;
}

int lutldpc_set_onchip(lutldpc_decoder* h, int on) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (on && !h->onchip_fits)
    return h->fail(LUTLDPC_ERR_STATE, "the decoder does not qualify for the on-chip regime (LUT check nodes, or the state "
                                      "of a word column does not fit shared memory twice per SM)");
  h->onchip_ok = on != 0;
  return LUTLDPC_OK;
}

int lutldpc_set_exit_conditions(lutldpc_decoder* h, int max_iters, int psc, int pisc) {
  if (!h) return LUTLDPC_ERR_ARG;
  // the decision trees belong to the last designed iteration (dec_update_lut(.., max_iters-1, ..),
  // src/LDPC_Code_LUT.cpp:342), so the iteration count is fixed by the design
  if (max_iters != h->m.max_iters)
    return h->fail(LUTLDPC_ERR_ARG, "max_iters must equal the number of designed iterations (" +
                                        std::to_string(h->m.max_iters) + ")");
  h->max_iters = max_iters; h->psc = psc != 0; h->pisc = pisc != 0;
  return LUTLDPC_OK;
}

int lutldpc_set_initial_message_mode(lutldpc_decoder* h, int mode) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (mode != LUTLDPC_INIT_CONT && mode != LUTLDPC_INIT_QCHA) return h->fail(LUTLDPC_ERR_ARG, "unknown initial message mode");
  if (mode == LUTLDPC_INIT_QCHA && h->m.cha2msg.empty()) return h->fail(LUTLDPC_ERR_STATE, "decoder has no Nq_Cha_2_Nq_Msg_map");
  if (mode == LUTLDPC_INIT_CONT && (int)h->m.qb_Msg.size() != h->m.Nq_Msg[0] - 1)
    return h->fail(LUTLDPC_ERR_STATE, "decoder has no message quantiser qb_Msg for continuous inputs");
  h->init_mode = mode;
  return LUTLDPC_OK;
}

int lutldpc_set_stream(lutldpc_decoder* h, void* cuda_stream) {
  if (!h) return LUTLDPC_ERR_ARG;
  h->stream = (cudaStream_t)cuda_stream;
  return LUTLDPC_OK;
}

int lutldpc_set_profiling(lutldpc_decoder* h, int on) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (h->profiling && !on) h->flush_profile();
  h->profiling = on != 0;
  if (on) {
    h->evn = 0;
    for (int i = 0; i < LUTLDPC_PROFILE_KINDS; ++i) { h->prof_ms[i] = 0; h->prof_n[i] = 0; }
  }
  return LUTLDPC_OK;
}

int lutldpc_get_profile(lutldpc_decoder* h, double* ms, int64_t* launches, int32_t* class_degree, int32_t* class_nodes) {
  if (!h || !ms || !launches) return LUTLDPC_ERR_ARG;
  int rc = h->flush_profile();
  if (rc) return rc;
  for (int i = 0; i < LUTLDPC_PROFILE_KINDS; ++i) { ms[i] = h->prof_ms[i]; launches[i] = h->prof_n[i]; }
  for (int c = 0; c < 16; ++c) {
    bool have = c < (int)h->vn_classes.size();
    if (class_degree) class_degree[c] = have ? h->vn_classes[c].degree : 0;
    if (class_nodes) class_nodes[c] = have ? h->vn_classes[c].n : 0;
  }
  return LUTLDPC_OK;
}

int lutldpc_set_poll_interval(lutldpc_decoder* h, int every) {
  if (!h) return LUTLDPC_ERR_ARG;
  h->poll_every = every < 0 ? 0 : every;
  return LUTLDPC_OK;
}

// ------------------------------------------------------------------------------------ decode
static int check_decode_args(lutldpc_decoder* h, int n_frames, const void* a, const void* b2, const void* c) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (n_frames < 0) return h->fail(LUTLDPC_ERR_ARG, "n_frames < 0");
  if (n_frames > 0 && (!a || !b2 || !c)) return h->fail(LUTLDPC_ERR_ARG, "NULL buffer");
  return LUTLDPC_OK;
}

int lutldpc_decode_quantized_device(lutldpc_decoder* h, int n_frames, const uint8_t* d_cha, const uint8_t* d_msg,
                                    uint8_t* d_out_bits, int32_t* d_out_iters) {
  int rc = check_decode_args(h, n_frames, d_cha, d_msg, d_out_bits);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  const int N = h->m.nvar, maxb = h->max_batch_frames();
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    int nf = std::min(maxb, n_frames - f0);
    rc = h->run_batch(nf, d_cha + (size_t)f0 * N, d_msg + (size_t)f0 * N, nullptr);
    if (rc) return rc;
    CKH(launch_unpack(h->g, h->b, d_out_bits + (size_t)f0 * N, h->stream), "unpack");
    h->launched(KIND_UNPACK);
    if (d_out_iters)
      CKH(cudaMemcpyAsync(d_out_iters + f0, h->b.iters, (size_t)nf * 4, cudaMemcpyDeviceToDevice, h->stream), "iters copy");
  }
  return LUTLDPC_OK;
}

// Host-buffer decode.  The call is cut into chunks and pipelined over three streams: while chunk c is decoded
// on the handle's stream, the inputs of chunk c+1 travel host->device and the results of chunk c-1
// device->host on two copy streams (PCIe is full duplex); device staging is double-buffered.  With pinned
// host buffers the copies are asynchronous and hide behind the decoder; with pageable buffers the runtime
// stages them synchronously and the call degrades to copy-then-decode, with the same results.
static int host_chunk_frames(lutldpc_decoder* h, int n_frames) {
  const int maxb = h->max_batch_frames();
  if (const char* e = getenv("LUTLDPC_HOST_CHUNK")) {
    int v = atoi(e);
    if (v > 0) return std::min(maxb, std::max(256, (v + 255) / 256 * 256));
  }
  // ~384 MB of labels per chunk: large enough for the kernels to run at batch efficiency (N=64800: 6144 frames),
  // small enough that the first copy in and the last copy out -- the only ones not hidden -- stay short
  int c = (int)std::min<long long>(maxb, ((3LL << 27) / std::max(1, h->m.nvar) + 512) / 1024 * 1024);
  (void)n_frames;
  return std::min(maxb, std::max(2048, c));
}

// Chunk sizes of one host-buffer call.  Only the first copy in and the last copy out are exposed, and the
// decoder runs closer to its batch efficiency on large chunks: long calls therefore start with a small chunk,
// grow by 2x per chunk (the copy of chunk c+1 must fit under the decode of chunk c: ~2.4 us per frame over
// PCIe against >= 6 us per frame of decoding for N=64800), run chunks of twice the nominal size in the middle
// and shrink again at the end.  Short calls keep the uniform nominal chunks.
static std::vector<int> host_chunk_plan(lutldpc_decoder* h, int n_frames, bool single) {
  std::vector<int> plan;
  if (single) { plan.push_back(n_frames); return plan; }
  int C = host_chunk_frames(h, n_frames);
  const int maxb = h->max_batch_frames();
  if (const char* e = getenv("LUTLDPC_HOST_RAMP_NOMINAL")) {  // tests: ramped plan on a small nominal chunk
    const int v = atoi(e);
    if (v >= 1024) C = std::min(maxb / 2, v / 256 * 256);
  }
  auto r256 = [](long long v) { return (int)std::max<long long>(256, (v + 128) / 256 * 256); };
  const int up[2] = {r256(C / 2), C};
  const int down[2] = {C, r256(C / 2)};
  const long long ramps = (long long)up[0] + up[1] + down[0] + down[1];
  const bool ramped = !getenv("LUTLDPC_HOST_CHUNK") && n_frames >= ramps && 2 * C <= maxb;
  if (!ramped) {
    for (int f = 0; f < n_frames; f += C) plan.push_back(std::min(C, n_frames - f));
    return plan;
  }
  for (int u : up) plan.push_back(u);
  long long mid = n_frames - ramps;
  if (mid > 0) {
    const int n_mid = (int)((mid + 2 * C - 1) / (2 * C));
    const int per = r256((mid + n_mid - 1) / n_mid);
    while (mid > 0) {
      const int t = (int)std::min<long long>(per, mid);
      plan.push_back(t);
      mid -= t;
    }
  }
  for (int d : down) plan.push_back(d);
  return plan;
}

// packed: `cha` holds both labels of a symbol in one byte (msg == nullptr) and out_bits is bit-packed, NB bytes per frame
static int decode_host(lutldpc_decoder* h, int n_frames, const uint8_t* cha, const uint8_t* msg, const double* llr,
                       uint8_t* out_bits, int32_t* out_iters, TraceSink* tr, bool packed = false) {
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  if (n_frames == 0) return LUTLDPC_OK;
  const int N = h->m.nvar;
  const size_t out_per_frame = packed ? (size_t)(N + 7) / 8 : (size_t)N;
  const std::vector<int> plan = host_chunk_plan(h, n_frames, tr != nullptr);
  const int n_chunks = (int)plan.size();
  std::vector<int> first(n_chunks + 1, 0);
  int chunk = 0;  // largest chunk: capacity of the staging buffers
  for (int c = 0; c < n_chunks; ++c) {
    first[c + 1] = first[c] + plan[c];
    chunk = std::max(chunk, plan[c]);
  }
  cudaStream_t st = h->stream;
  HostPipe& hp = h->pipe;
  if (!hp.s_in) {
    CKH(cudaStreamCreateWithFlags(&hp.s_in, cudaStreamNonBlocking), "stream");
    CKH(cudaStreamCreateWithFlags(&hp.s_out, cudaStreamNonBlocking), "stream");
    for (int i = 0; i < 2; ++i) {
      CKH(cudaEventCreateWithFlags(&hp.ev_in[i], cudaEventDisableTiming), "event");
      CKH(cudaEventCreateWithFlags(&hp.ev_used[i], cudaEventDisableTiming), "event");
      CKH(cudaEventCreateWithFlags(&hp.ev_dec[i], cudaEventDisableTiming), "event");
      CKH(cudaEventCreateWithFlags(&hp.ev_out[i], cudaEventDisableTiming), "event");
    }
  }
  const size_t cap = (size_t)std::min(chunk, n_frames) * N;
  for (int i = 0; i < (n_chunks > 1 ? 2 : 1); ++i) {
    CKH(hp.lab_cha[i].ensure(cap), "alloc labels");
    if (!packed) CKH(hp.lab_msg[i].ensure(cap), "alloc labels");
    CKH(hp.out[i].ensure(cap), "alloc out");
    CKH(hp.its[i].ensure((size_t)std::min(chunk, n_frames)), "alloc iters");
    if (llr) CKH(hp.llr[i].ensure(cap), "alloc llr");
  }
  // inputs of the previous call on the handle's stream may still be in flight: order the first copy after them
  CKH(cudaEventRecord(hp.ev_used[0], st), "event");
  CKH(cudaEventRecord(hp.ev_used[1], st), "event");
  auto h2d = [&](int c) -> int {
    const int slot = c & 1, f0 = first[c], nf = plan[c];
    const size_t cnt = (size_t)nf * N;
    CKH(cudaStreamWaitEvent(hp.s_in, hp.ev_used[slot], 0), "wait");  // staging slot consumed by chunk c-2
    if (llr) {
      CKH(cudaMemcpyAsync(hp.llr[slot].p, llr + (size_t)f0 * N, cnt * 8, cudaMemcpyHostToDevice, hp.s_in), "H2D llr");
    } else {
      CKH(cudaMemcpyAsync(hp.lab_cha[slot].p, cha + (size_t)f0 * N, cnt, cudaMemcpyHostToDevice, hp.s_in), "H2D cha");
      if (!packed)
        CKH(cudaMemcpyAsync(hp.lab_msg[slot].p, msg + (size_t)f0 * N, cnt, cudaMemcpyHostToDevice, hp.s_in), "H2D msg");
    }
    CKH(cudaEventRecord(hp.ev_in[slot], hp.s_in), "event");
    return LUTLDPC_OK;
  };
  int rc = h2d(0);
  if (rc) return rc;
  for (int c = 0; c < n_chunks; ++c) {
    const int slot = c & 1, f0 = first[c], nf = plan[c];
    const size_t cnt = (size_t)nf * N;
    if (c + 1 < n_chunks && (rc = h2d(c + 1))) return rc;
    CKH(cudaStreamWaitEvent(st, hp.ev_in[slot], 0), "wait");
    if (llr)
      CKH(launch_quantize(hp.llr[slot].p, cnt, h->d_qb_cha.p, (int)h->m.qb_Cha.size(), h->d_qb_msg.p,
                          (int)h->m.qb_Msg.size(), h->d_cha2msg.p, h->init_mode == LUTLDPC_INIT_QCHA,
                          hp.lab_cha[slot].p, hp.lab_msg[slot].p, st), "quantize");
    static const bool trace_host = getenv("LUTLDPC_HOST_TRACE") != nullptr;
    const auto t_a = std::chrono::steady_clock::now();
    rc = h->run_batch(nf, hp.lab_cha[slot].p, packed ? nullptr : hp.lab_msg[slot].p, tr);
    if (rc) return rc;
    if (trace_host)
      fprintf(stderr, "host chunk %d frames %d run_batch %.3f ms\n", c, nf,
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_a).count());
    CKH(cudaEventRecord(hp.ev_used[slot], st), "event");
    if (c >= 2) CKH(cudaStreamWaitEvent(st, hp.ev_out[slot], 0), "wait");  // results of chunk c-2 have left
    if (packed) CKH(launch_unpack_bits(h->g, h->b, hp.out[slot].p, st), "unpack_bits");
    else CKH(launch_unpack(h->g, h->b, hp.out[slot].p, st), "unpack");
    h->launched(KIND_UNPACK);
    CKH(cudaMemcpyAsync(hp.its[slot].p, h->b.iters, (size_t)nf * 4, cudaMemcpyDeviceToDevice, st), "iters copy");
    CKH(cudaEventRecord(hp.ev_dec[slot], st), "event");
    CKH(cudaStreamWaitEvent(hp.s_out, hp.ev_dec[slot], 0), "wait");
    CKH(cudaMemcpyAsync(out_bits + (size_t)f0 * out_per_frame, hp.out[slot].p, (size_t)nf * out_per_frame,
                        cudaMemcpyDeviceToHost, hp.s_out), "D2H bits");
    if (out_iters)
      CKH(cudaMemcpyAsync(out_iters + f0, hp.its[slot].p, (size_t)nf * 4, cudaMemcpyDeviceToHost, hp.s_out), "D2H iters");
    CKH(cudaEventRecord(hp.ev_out[slot], hp.s_out), "event");
  }
  CKH(cudaStreamSynchronize(hp.s_out), "sync");
  CKH(cudaStreamSynchronize(st), "sync");
  return LUTLDPC_OK;
}

int lutldpc_decode_quantized(lutldpc_decoder* h, int n_frames, const uint8_t* cha, const uint8_t* msg,
                             uint8_t* out_bits, int32_t* out_iters) {
  int rc = check_decode_args(h, n_frames, cha, msg, out_bits);
  if (rc) return rc;
  return decode_host(h, n_frames, cha, msg, nullptr, out_bits, out_iters, nullptr);
}

int lutldpc_decode_packed(lutldpc_decoder* h, int n_frames, const uint8_t* cha_msg, uint8_t* out_bits_packed,
                          int32_t* out_iters) {
  int rc = check_decode_args(h, n_frames, cha_msg, cha_msg, out_bits_packed);
  if (rc) return rc;
  return decode_host(h, n_frames, cha_msg, nullptr, nullptr, out_bits_packed, out_iters, nullptr, true);
}

int lutldpc_decode_packed_device(lutldpc_decoder* h, int n_frames, const uint8_t* d_cha_msg, uint8_t* d_out_bits_packed,
                                 int32_t* d_out_iters) {
  int rc = check_decode_args(h, n_frames, d_cha_msg, d_cha_msg, d_out_bits_packed);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  const int N = h->m.nvar, NB = (N + 7) / 8, maxb = h->max_batch_frames();
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    int nf = std::min(maxb, n_frames - f0);
    rc = h->run_batch(nf, d_cha_msg + (size_t)f0 * N, nullptr, nullptr);
    if (rc) return rc;
    CKH(launch_unpack_bits(h->g, h->b, d_out_bits_packed + (size_t)f0 * NB, h->stream), "unpack_bits");
    h->launched(KIND_UNPACK);
    if (d_out_iters)
      CKH(cudaMemcpyAsync(d_out_iters + f0, h->b.iters, (size_t)nf * 4, cudaMemcpyDeviceToDevice, h->stream), "iters copy");
  }
  return LUTLDPC_OK;
}

int lutldpc_decode_quantized_traced(lutldpc_decoder* h, int n_frames, const uint8_t* cha, const uint8_t* msg,
                                    uint8_t* out_bits, int32_t* out_iters, int trace_frame, uint8_t* trace_msgs,
                                    int32_t* trace_kind, int32_t* trace_iter, int max_rec, int32_t* n_rec) {
  int rc = check_decode_args(h, n_frames, cha, msg, out_bits);
  if (rc) return rc;
  if (trace_frame < 0 || trace_frame >= n_frames || !trace_msgs || !trace_kind || !trace_iter || !n_rec || max_rec < 0)
    return h->fail(LUTLDPC_ERR_ARG, "bad trace arguments");
  if (n_frames > h->max_batch_frames()) return h->fail(LUTLDPC_ERR_ARG, "traced decode is limited to one batch");
  TraceSink tr;
  tr.frame = trace_frame; tr.max_rec = max_rec; tr.msgs = trace_msgs; tr.kind = trace_kind; tr.iter = trace_iter;
  rc = decode_host(h, n_frames, cha, msg, nullptr, out_bits, out_iters, &tr);
  *n_rec = tr.n_rec;
  return rc;
}

int lutldpc_decode_llr(lutldpc_decoder* h, int n_frames, const double* llr, uint8_t* out_bits, int32_t* out_iters) {
  int rc = check_decode_args(h, n_frames, llr, llr, out_bits);
  if (rc) return rc;
  if (h->init_mode == LUTLDPC_INIT_CONT && (int)h->m.qb_Msg.size() != h->m.Nq_Msg[0] - 1)
    return h->fail(LUTLDPC_ERR_STATE, "decoder has no message quantiser qb_Msg for continuous inputs");
  return decode_host(h, n_frames, nullptr, nullptr, llr, out_bits, out_iters, nullptr);
}

int lutldpc_decode_llr_device(lutldpc_decoder* h, int n_frames, const double* d_llr, uint8_t* d_out_bits,
                              int32_t* d_out_iters) {
  int rc = check_decode_args(h, n_frames, d_llr, d_llr, d_out_bits);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  const int N = h->m.nvar, maxb = h->max_batch_frames();
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    int nf = std::min(maxb, n_frames - f0);
    size_t cnt = (size_t)nf * N;
    CKH(h->buf_lab_cha.ensure(cnt), "alloc labels");
    CKH(h->buf_lab_msg.ensure(cnt), "alloc labels");
    CKH(launch_quantize(d_llr + (size_t)f0 * N, cnt, h->d_qb_cha.p, (int)h->m.qb_Cha.size(), h->d_qb_msg.p,
                        (int)h->m.qb_Msg.size(), h->d_cha2msg.p, h->init_mode == LUTLDPC_INIT_QCHA, h->buf_lab_cha.p,
                        h->buf_lab_msg.p, h->stream), "quantize");
    rc = h->run_batch(nf, h->buf_lab_cha.p, h->buf_lab_msg.p, nullptr);
    if (rc) return rc;
    CKH(launch_unpack(h->g, h->b, d_out_bits + (size_t)f0 * N, h->stream), "unpack");
    h->launched(KIND_UNPACK);
    if (d_out_iters)
      CKH(cudaMemcpyAsync(d_out_iters + f0, h->b.iters, (size_t)nf * 4, cudaMemcpyDeviceToDevice, h->stream), "iters copy");
  }
  return LUTLDPC_OK;
}

int lutldpc_quantize(lutldpc_decoder* h, int n_frames, const double* llr, uint8_t* cha, uint8_t* msg) {
  int rc = check_decode_args(h, n_frames, llr, cha, msg);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  size_t cnt = (size_t)n_frames * h->m.nvar;
  if (!cnt) return LUTLDPC_OK;
  cudaStream_t st = h->stream;
  CKH(h->buf_llr.ensure(cnt), "alloc llr");
  CKH(h->buf_lab_cha.ensure(cnt), "alloc labels");
  CKH(h->buf_lab_msg.ensure(cnt), "alloc labels");
  CKH(cudaMemcpyAsync(h->buf_llr.p, llr, cnt * 8, cudaMemcpyHostToDevice, st), "H2D llr");
  CKH(launch_quantize(h->buf_llr.p, cnt, h->d_qb_cha.p, (int)h->m.qb_Cha.size(), h->d_qb_msg.p, (int)h->m.qb_Msg.size(),
                      h->d_cha2msg.p, h->init_mode == LUTLDPC_INIT_QCHA, h->buf_lab_cha.p, h->buf_lab_msg.p, st), "quantize");
  CKH(cudaMemcpyAsync(cha, h->buf_lab_cha.p, cnt, cudaMemcpyDeviceToHost, st), "D2H");
  CKH(cudaMemcpyAsync(msg, h->buf_lab_msg.p, cnt, cudaMemcpyDeviceToHost, st), "D2H");
  CKH(cudaStreamSynchronize(st), "sync");
  return LUTLDPC_OK;
}

int lutldpc_quantize_device(lutldpc_decoder* h, int n_frames, const double* d_llr, uint8_t* d_cha, uint8_t* d_msg) {
  int rc = check_decode_args(h, n_frames, d_llr, d_cha, d_msg);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  size_t cnt = (size_t)n_frames * h->m.nvar;
  if (!cnt) return LUTLDPC_OK;
  CKH(launch_quantize(d_llr, cnt, h->d_qb_cha.p, (int)h->m.qb_Cha.size(), h->d_qb_msg.p, (int)h->m.qb_Msg.size(),
                      h->d_cha2msg.p, h->init_mode == LUTLDPC_INIT_QCHA, d_cha, d_msg, h->stream), "quantize");
  return LUTLDPC_OK;
}

int lutldpc_syndrome_check(lutldpc_decoder* h, int n_frames, const uint8_t* bits, uint8_t* ok) {
  int rc = check_decode_args(h, n_frames, bits, bits, ok);
  if (rc) return rc;
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  const int N = h->m.nvar, maxb = h->max_batch_frames();
  cudaStream_t st = h->stream;
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    int nf = std::min(maxb, n_frames - f0);
    size_t cnt = (size_t)nf * N;
    if ((rc = h->ensure_batch(nf))) return rc;
    CKH(h->buf_lab_cha.ensure(cnt), "alloc");
    CKH(cudaMemcpyAsync(h->buf_lab_cha.p, bits + (size_t)f0 * N, cnt, cudaMemcpyHostToDevice, st), "H2D bits");
    CKH(launch_reset_flags(h->b, st), "reset_flags");
    CKH(launch_pack_bits(h->buf_lab_cha.p, h->g, h->b, st), "pack_bits");
    CKH(launch_syndrome(h->g, h->b, st), "syndrome");
    std::vector<uint32_t> sf(h->b.W);
    CKH(cudaMemcpyAsync(sf.data(), h->b.syn_fail, sf.size() * 4, cudaMemcpyDeviceToHost, st), "D2H flags");
    CKH(cudaStreamSynchronize(st), "sync");
    for (int f = 0; f < nf; ++f) ok[f0 + f] = !((sf[f >> 3] >> (4 * (f & 7))) & 1);
  }
  return LUTLDPC_OK;
}

int lutldpc_last_call_stats(const lutldpc_decoder* h, double* device_ms, int64_t* kernel_launches,
                            int64_t* frame_iterations) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (!h->have_stats) return LUTLDPC_ERR_STATE;
  if (cudaEventSynchronize(h->ev1) != cudaSuccess) return LUTLDPC_ERR_CUDA;
  float ms = 0;
  if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return LUTLDPC_ERR_CUDA;
  if (device_ms) *device_ms = ms;
  if (kernel_launches) *kernel_launches = h->launches;
  if (frame_iterations && h->stat_frame_iters >= 0) {
    *frame_iterations = h->stat_frame_iters;
  } else if (frame_iterations) {
    std::vector<int32_t> it((size_t)h->last_frames);
    if (cudaMemcpy(it.data(), h->b.iters, it.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return LUTLDPC_ERR_CUDA;
    int64_t s = 0;
    for (int32_t v : it) s += v < 0 ? -v : v;
    *frame_iterations = s;
  }
  return LUTLDPC_OK;
}

int lutldpc_set_repack(lutldpc_decoder* h, int on, double alpha) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (on && !(alpha > 0.0 && alpha <= 1.0)) return h->fail(LUTLDPC_ERR_ARG, "repack alpha must be in (0, 1]");
  h->repack_on = on != 0;
  if (on) h->repack_alpha = alpha;
  return LUTLDPC_OK;
}

int lutldpc_get_repack_count(const lutldpc_decoder* h, int64_t* repacks) {
  if (!h || !repacks) return LUTLDPC_ERR_ARG;
  *repacks = h->repacks;
  return LUTLDPC_OK;
}

}  // extern "C"
