// Decoder object: device set-up and the pass schedule of one batch.  See decoder.hpp.
#include "decoder.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

using namespace lutldpc;

#define CK(call, where)                         \
  do {                                          \
    cudaError_t e_ = (call);                    \
    if (e_ != cudaSuccess) return cuda_fail(e_, where); \
  } while (0)

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s && *s ? atoi(s) : dflt;
}

lutldpc_decoder::~lutldpc_decoder() {
  if (d_gen_bits) cudaFree(d_gen_bits);
  if (h_active) cudaFreeHost(h_active);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  if (ev_sim0) cudaEventDestroy(ev_sim0);
  if (ev_sim1) cudaEventDestroy(ev_sim1);
  if (ev_lag) cudaEventDestroy(ev_lag);
  for (cudaEvent_t e : evpool) cudaEventDestroy(e);
}

int lutldpc_decoder::cuda_fail(cudaError_t e, const char* where) {
  err = std::string("CUDA error in ") + where + ": " + cudaGetErrorString(e);
  return LUTLDPC_ERR_CUDA;
}

int lutldpc_decoder::init_device() {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(LUTLDPC_ERR_CUDA, "no CUDA device available (this library has no CPU fallback)");
  CK(cudaGetDevice(&device), "cudaGetDevice");
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
  if (prop.major < 10)
    return fail(LUTLDPC_ERR_CUDA, std::string("device ") + prop.name + " is not sm_100 (kernels are built for sm_100a only)");
  CK(configure_kernels(), "configure_kernels");
  max_iters = m.max_iters;

  CK(d_vn_edge0.upload(m.vn_edge0), "upload vn_edge0");
  CK(d_dv.upload(m.dv), "upload dv");
  CK(d_cn_start.upload(m.cn_start), "upload cn_start");
  CK(d_cn_msg_idx.upload(m.cn_msg_idx), "upload cn_msg_idx");
  CK(d_chk_equ_idx.upload(m.chk_equ_idx), "upload chk_equ_idx");
  CK(d_qb_cha.upload(m.qb_Cha), "upload qb_Cha");
  CK(d_qb_msg.upload(m.qb_Msg), "upload qb_Msg");
  CK(d_cha2msg.upload(m.cha2msg), "upload cha2msg");
  if (!m.gen.empty()) CK(d_gen.upload(m.gen), "upload generator");
  g.vn_edge0 = d_vn_edge0.p; g.dv = d_dv.p; g.cn_start = d_cn_start.p; g.cn_msg_idx = d_cn_msg_idx.p;
  g.chk_equ_idx = d_chk_equ_idx.p; g.nvar = m.nvar; g.nchk = m.nchk; g.num_edges = m.num_edges;
  g.max_dc = *std::max_element(m.dc.begin(), m.dc.end());

  // degree classes (set_trees: one tree per degree slot, src/LDPC_Code_LUT.cpp:128-139)
  int nvc = m.var_trees.n_deg;
  vn_classes = std::vector<NodeClass>(nvc);
  {
    std::vector<std::vector<int32_t>> lists(nvc);
    for (int v = 0; v < m.nvar; ++v) lists[m.var_tree_idx_degree[v]].push_back(v);
    for (int c = 0; c < nvc; ++c) {
      vn_classes[c].degree = m.var_trees.at(0, c).num_leaves;
      vn_classes[c].n = (int)lists[c].size();
      CK(vn_classes[c].list.upload(lists[c]), "upload vn class list");
      std::vector<int2> pr(lists[c].size());
      for (size_t i = 0; i < pr.size(); ++i) pr[i] = make_int2(lists[c][i], m.vn_edge0[lists[c][i]]);
      CK(vn_classes[c].pairs.upload(pr), "upload vn class pairs");
    }
  }
  if (m.minLUT) {  // min-sum: checks grouped by exact degree, edge rows stored back to back
    std::map<int, std::vector<int32_t>> rows;
    bool ok = true;
    for (int c = 0; c < m.nchk; ++c) {
      ok &= cn_exact_supported(m.dc[c]);
      std::vector<int32_t>& r = rows[m.dc[c]];
      r.insert(r.end(), m.cn_msg_idx.begin() + m.cn_start[c], m.cn_msg_idx.begin() + m.cn_start[c + 1]);
    }
    const char* e = getenv("LUTLDPC_CN_GENERIC");
    use_cn_exact = ok && !(e && e[0] == '1');
    if (use_cn_exact) {
      cn_exact = std::vector<NodeClass>(rows.size());
      // launch order of the degree classes: descending degree by default.  Measured on N=64800 (dc 7/8/9) the
      // order 9, 8, 7 takes 0.70 ms per check pass of 16384 frames against 0.76 ms for 7, 8, 9: the launch that
      // follows the variable pass competes with the write-back of its dirty lines, and the larger class pays more.
      const bool rev = env_int("LUTLDPC_CN_ASCENDING", 0) == 0;
      int i = rev ? (int)rows.size() - 1 : 0;
      for (auto& kv : rows) {
        cn_exact[i].degree = kv.first;
        cn_exact[i].n = (int)(kv.second.size() / kv.first);
        CK(cn_exact[i].list.upload(kv.second), "upload cn rows");
        i += rev ? -1 : 1;
      }
    }
  }
  int ncc = m.minLUT ? 0 : m.chk_trees.n_deg;
  cn_classes = std::vector<NodeClass>(ncc);
  if (ncc) {
    std::vector<std::vector<int32_t>> lists(ncc);
    for (int c = 0; c < m.nchk; ++c) lists[m.chk_tree_idx_degree[c]].push_back(c);
    for (int c = 0; c < ncc; ++c) {
      cn_classes[c].degree = m.chk_trees.at(0, c).num_leaves + 1;
      cn_classes[c].n = (int)lists[c].size();
      CK(cn_classes[c].list.upload(lists[c]), "upload cn class list");
    }
  }

  if (const char* e = getenv("LUTLDPC_GENERIC_VN")) use_fast_vn = !(e[0] == '1');
  repack_on = env_int("LUTLDPC_REPACK", 1) != 0;
  if (const char* e = getenv("LUTLDPC_REPACK_ALPHA")) repack_alpha = atof(e);
  // compile every (iteration slot, degree class) tree into a node program
  n_iter_slots = m.var_trees.n_iter;
  std::vector<int> slot_first_iter(n_iter_slots, -1);
  for (int ii = 0; ii < m.max_iters; ++ii) {
    int t = m.var_tree_idx_iter[ii];
    if (slot_first_iter[t] < 0) slot_first_iter[t] = ii;
    else if (m.Nq_Msg[ii] != m.Nq_Msg[slot_first_iter[t]])
      return fail(LUTLDPC_ERR_MODEL, "iterations that reuse a LUT slot must share the message resolution");
  }
  std::vector<uint8_t> blob;
  auto add_prog = [&](const NodeProgram& p, ProgRef& r) {
    std::vector<uint8_t> img = p.image();
    r.off = blob.size();
    r.bytes = (uint32_t)img.size();
    r.n_slots = (uint32_t)p.n_slots;
    r.n_in = (uint32_t)p.n_in;
    r.n_ops = (int64_t)p.ops.size();
    r.n_naive = p.n_lookups_naive;
    r.tables_off = reinterpret_cast<const DevProgHeader*>(img.data())->tables_off;
    r.q_out = p.q_out;
    blob.insert(blob.end(), img.begin(), img.end());
  };
  var_progs.assign((size_t)n_iter_slots * nvc, ProgRef());
  for (int t = 0; t < n_iter_slots; ++t) {
    bool dec = t == n_iter_slots - 1;  // the last slot holds the decision trees
    int ii = slot_first_iter[t];
    for (int c = 0; c < nvc; ++c) {
      if (vn_classes[c].n == 0) continue;
      const Tree& tr = m.var_trees.at(t, c);
      NodeProgram p;
      std::string perr;
      if (!compile_program(tr, vn_classes[c].degree, dec ? PROG_DEC : PROG_VAR, m.Nq_Msg[ii], m.Nq_Cha, p, perr))
        return fail(LUTLDPC_ERR_MODEL, "var tree (slot " + std::to_string(t) + ", degree " +
                                           std::to_string(vn_classes[c].degree) + "): " + perr);
      // every iteration that uses this slot must be followed by the alphabet the slot's root emits: the next check
      // pass reads the codes with code_bits(Nq_Msg[ii+1])
      if (!dec)
        for (int i2 = 0; i2 + 1 < m.max_iters; ++i2)
          if (m.var_tree_idx_iter[i2] == t && tr.nodes[0].K != m.Nq_Msg[i2 + 1])
            return fail(LUTLDPC_ERR_MODEL, "var tree output alphabet does not match Nq_Msg of the next iteration");
      if (dec && tr.nodes[0].K != 2) return fail(LUTLDPC_ERR_MODEL, "decision tree root must be binary");
      add_prog(p, var_progs[(size_t)t * nvc + c]);
      if (!dec && use_fast_vn && fast_vn_matches(p)) var_progs[(size_t)t * nvc + c].fast_dv = vn_classes[c].degree;
    }
  }
  {  // fused VN launch plan: classes whose program is the specialised one in every iteration slot
    const char* e = getenv("LUTLDPC_VN_UNFUSED");
    bool want = !(e && e[0] == '1');
    std::vector<std::pair<double, int4>> tiles;
    for (int c = 0; c < nvc && want; ++c) {
      if (!vn_classes[c].n) continue;
      bool all_fast = true;
      for (int t = 0; t + 1 < n_iter_slots; ++t) all_fast &= var_progs[(size_t)t * nvc + c].fast_dv != 0;
      // only the low degrees share a launch: their code fits 40 registers (12 CTAs per SM); a high-degree class in
      // the same kernel would drag every class down to its register count, so it keeps its own k_vn_fast<D> launch
      if (!all_fast || n_iter_slots < 2 || fused_classes.size() >= 8 || vn_classes[c].degree > 8) continue;
      int j = (int)fused_classes.size();
      fused_classes.push_back(c);
      fused_max_dv = std::max(fused_max_dv, vn_classes[c].degree);
      int ops = (int)var_progs[c].n_ops;
      int per = std::max(1, env_int("LUTLDPC_VN_TILE_OPS", 256) / std::max(1, ops));
      int nt = (vn_classes[c].n + per - 1) / per;
      for (int i = 0; i < nt; ++i)
        tiles.push_back({(i + 0.5) / nt, make_int4(j, i * per, std::min(per, vn_classes[c].n - i * per), 0)});
    }
    // a single class keeps its own k_vn_fast<D> launch (measured on the (3,6) N=1000 code: 3.54 Gbit/s against 3.45
    // through the fused kernel); LUTLDPC_VN_FUSE_MIN=1 sends it through the fused kernel anyway
    if (fused_classes.empty() || env_int("LUTLDPC_VN_FUSE_MIN", 2) > (int)fused_classes.size()) fused_classes.clear();
    else {
      std::stable_sort(tiles.begin(), tiles.end(), [](const std::pair<double, int4>& a, const std::pair<double, int4>& b2) {
        return a.first < b2.first;
      });
      std::vector<int4> flat;
      for (auto& t : tiles) flat.push_back(t.second);
      n_fused_tiles = (int)flat.size();
      CK(d_fused_tiles.upload(flat), "upload fused tiles");
    }
  }
  chk_progs.assign((size_t)n_iter_slots * ncc, ProgRef());
  for (int t = 0; t < n_iter_slots && ncc; ++t) {
    int ii = slot_first_iter[t];
    for (int c = 0; c < ncc; ++c) {
      if (cn_classes[c].n == 0) continue;
      NodeProgram p;
      std::string perr;
      if (!compile_program(m.chk_trees.at(t, c), cn_classes[c].degree, PROG_CHK, m.Nq_Msg[ii], m.Nq_Cha, p, perr))
        return fail(LUTLDPC_ERR_MODEL, "chk tree (slot " + std::to_string(t) + ", degree " +
                                           std::to_string(cn_classes[c].degree) + "): " + perr);
      if (m.chk_trees.at(t, c).nodes[0].K != m.Nq_Msg[ii])
        return fail(LUTLDPC_ERR_MODEL, "chk tree output alphabet does not match Nq_Msg");
      add_prog(p, chk_progs[(size_t)t * ncc + c]);
    }
  }
  for (auto& r : var_progs)
    if (r.bytes + (size_t)r.n_slots * kNodeThreads * 4 > 200 * 1024)
      return fail(LUTLDPC_ERR_MODEL, "a node program (LUTs + value slots) exceeds the 200 KiB shared-memory budget");
  for (auto& r : chk_progs)
    if (r.bytes + (size_t)r.n_slots * kNodeThreads * 4 > 200 * 1024)
      return fail(LUTLDPC_ERR_MODEL, "a node program (LUTs + value slots) exceeds the 200 KiB shared-memory budget");
  CK(d_prog_blob.upload(blob), "upload node programs");
  {
    int rc = init_onchip();
    if (rc) return rc;
  }
  CK(cudaHostAlloc((void**)&h_active, 8 * sizeof(uint32_t), cudaHostAllocMapped), "cudaHostAlloc");
  memset(h_active, 0, 8 * sizeof(uint32_t));
  CK(cudaHostGetDevicePointer((void**)&d_active_mirror, h_active, 0), "cudaHostGetDevicePointer");
  CK(cudaEventCreateWithFlags(&ev_lag, cudaEventDisableTiming), "cudaEventCreate");
  lag_poll = env_int("LUTLDPC_LAG_POLL", 1) != 0;
  tail_lag_frames = env_int("LUTLDPC_TAIL_LAG", 1024);
  CK(cudaEventCreate(&ev0), "cudaEventCreate");
  CK(cudaEventCreate(&ev1), "cudaEventCreate");
  return LUTLDPC_OK;
}

// profiling mode: one event after every launch; launch i took ev[i] - ev[i-1]
void lutldpc_decoder::launched(int kind) {
  ++launches;
  if (!profiling) return;
  if (evn == (int)evpool.size()) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    evpool.push_back(e);
    evkind.push_back(0);
  }
  evkind[evn] = kind;
  cudaEventRecord(evpool[evn], stream);
  ++evn;
}

int lutldpc_decoder::flush_profile() {
  if (evn > 0) {
    CK(cudaEventSynchronize(evpool[evn - 1]), "profile sync");
    static const bool dump = env_int("LUTLDPC_PROFILE_DUMP", 0) != 0;  // per-launch timeline on stderr
    for (int i = 1; i < evn; ++i) {
      if (evkind[i] < 0) continue;
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, evpool[i - 1], evpool[i]), "profile elapsed");
      prof_ms[evkind[i]] += ms;
      prof_n[evkind[i]] += 1;
      if (dump) fprintf(stderr, "launch %d kind %d ms %.4f\n", i, evkind[i], ms);
    }
  }
  evn = 0;
  return LUTLDPC_OK;
}

int lutldpc_decoder::max_batch_frames() const {
  // bound the resident batch to ~16 GiB of message state
  double per_frame = m.num_edges / 2.0 + m.nvar / 2.0 + m.nvar / 8.0 + 8;
  long f = (long)(16.0 * 1024 * 1024 * 1024 / per_frame);
  f = std::min(f, 65536L);
  // the hot kernels index msgs / cha / hard with 32-bit element offsets: rows * W must stay below 2^32
  const long rows = std::max<long>(std::max(m.num_edges, m.nvar), 1);
  f = std::min(f, ((1L << 32) - 1) / rows * 8 - 256);
  f = std::max(256L, f / 256 * 256);
  return (int)f;
}

int lutldpc_decoder::ensure_batch(int n_frames, bool onchip) {
  int W = ((n_frames + 255) / 256) * 32;
  if (onchip) CK(buf_msg0.ensure((size_t)m.nvar * W), "alloc msg0");  // the messages never leave shared memory
  else CK(buf_msgs.ensure((size_t)m.num_edges * W + 4), "alloc msgs");  // + slack: the repack gather over-reads 2 words
  CK(buf_cha.ensure((size_t)m.nvar * W + 4), "alloc cha");
  CK(buf_hard.ensure((size_t)m.nvar * W), "alloc hard");
  CK(buf_flags.ensure((size_t)4 * W + 32), "alloc flags");
  CK(buf_cols.ensure((size_t)W + W / 4), "alloc work lists");
  CK(buf_iters.ensure((size_t)8 * W), "alloc iters");
  b.msgs = onchip ? nullptr : buf_msgs.p; b.msg0 = onchip ? buf_msg0.p : nullptr;
  b.cha = buf_cha.p; b.hard = buf_hard.p;
  b.done = buf_flags.p; b.vn_fail = buf_flags.p + W; b.cn_fail = buf_flags.p + 2 * W; b.syn_fail = buf_flags.p + 3 * W;
  b.n_active = buf_flags.p + 4 * W;
  b.iters = buf_iters.p;
  b.cols = buf_cols.p; b.grp4 = buf_cols.p + W;
  b.host_cols = W; b.host_grp4 = W / 4;  // upper bounds until the first poll
  b.mirror = d_active_mirror;
  b.W = W; b.F = n_frames;
  return LUTLDPC_OK;
}

// LDPC_Code_LUT::lut_decode for a whole batch (src/LDPC_Code_LUT.cpp:259-353; control flow: SURVEY A.6).
int lutldpc_decoder::begin_batch(int n_frames, bool traced) {
  batch_onchip = onchip_ok && !traced;
  int rc = ensure_batch(n_frames, batch_onchip);
  if (rc) return rc;
  launches = 0;
  if (profiling && (rc = flush_profile())) return rc;
  CK(cudaEventRecord(ev0, stream), "event");
  launched(-1);
  --launches;
  CK(launch_reset_flags(b, stream), "reset_flags"); launched(KIND_RESET);
  return LUTLDPC_OK;
}

int lutldpc_decoder::repack_into(BatchBuffers& cur, int slot, int n_live) {
  RepackCtx& c = rctx[slot];
  const int W = std::max(32, ((n_live + 255) / 256) * 32);
  CK(c.msgs.ensure((size_t)m.num_edges * W + 4), "alloc repack msgs");
  CK(c.cha.ensure((size_t)m.nvar * W + 4), "alloc repack cha");
  CK(c.hard.ensure((size_t)m.nvar * W), "alloc repack hard");
  CK(c.flags.ensure((size_t)4 * W + 32), "alloc repack flags");
  CK(c.cols.ensure((size_t)W + W / 4), "alloc repack lists");
  CK(c.iters.ensure((size_t)8 * W), "alloc repack iters");
  CK(c.fid.ensure((size_t)8 * W), "alloc repack frame ids");
  CK(buf_rsrc.ensure((size_t)8 * W), "alloc repack plan");
  BatchBuffers n;
  n.msgs = c.msgs.p; n.cha = c.cha.p; n.hard = c.hard.p;
  n.done = c.flags.p; n.vn_fail = c.flags.p + W; n.cn_fail = c.flags.p + 2 * W; n.syn_fail = c.flags.p + 3 * W;
  n.n_active = c.flags.p + 4 * W;
  n.iters = c.iters.p; n.fid = c.fid.p; n.mirror = d_active_mirror;
  n.cols = c.cols.p; n.grp4 = c.cols.p + W;
  n.host_cols = (n_live + 7) / 8; n.host_grp4 = (n.host_cols + 3) / 4;
  n.W = W; n.F = n_live;
  // frames that finished in a repacked layout leave it now: their results go back to the original one
  if (cur.fid) {
    int rc = retire_from(cur);
    if (rc) return rc;
  }
  CK(launch_repack(g, cur, n, buf_rsrc.p, stream), "repack"); launched(KIND_REPACK);
  launches += 3;  // plan + msgs rows + cha rows + memset
  cur = n;
  ++repacks;
  return LUTLDPC_OK;
}

int lutldpc_decoder::retire_from(const BatchBuffers& cur) {
  if (!cur.fid) return LUTLDPC_OK;
  CK(buf_ow_first.ensure((size_t)2 * b.W + 1), "alloc retire plan");
  CK(buf_ow_mask.ensure((size_t)2 * b.W), "alloc retire plan");
  CK(launch_retire(g, cur, b, buf_ow_first.p, buf_ow_mask.p, stream), "retire");
  launched(KIND_RETIRE);
  launches += 1;
  return LUTLDPC_OK;
}

int lutldpc_decoder::run_batch(int n_frames, const uint8_t* d_cha_lab, const uint8_t* d_msg_lab, TraceSink* tr) {
  int rc = begin_batch(n_frames, tr != nullptr);
  if (rc) return rc;
  CK(launch_pack(d_cha_lab, d_msg_lab, g, b, m.Nq_Cha, m.Nq_Msg[0], stream), "pack"); launched(KIND_PACK);
  return run_decode(tr);
}

int lutldpc_decoder::run_decode(TraceSink* tr) { return batch_onchip ? run_onchip() : run_passes(tr); }

// The whole lut_decode of the batch in one persistent kernel (onchip.cu): cha / msg0 / done in, hard / iters out.
int lutldpc_decoder::run_onchip() {
  OnchipParams P = oc;
  P.W = b.W;
  P.cha = b.cha; P.msg0 = b.msg0; P.hard = b.hard; P.iters = b.iters; P.done = b.done;
  P.psc = psc; P.pisc = pisc;
  CK(launch_onchip(P, oc_max_fast_dv, oc_ctas, stream), "onchip");
  launched(KIND_ONCHIP);
  CK(cudaEventRecord(ev1, stream), "event");
  have_stats = true;
  last_frames = b.F;
  stat_frame_iters = -1;
  return LUTLDPC_OK;
}

// Set-up of the on-chip regime.  A decoder qualifies when its checks are min-sum and the state of one word column
// (E message words with the bank skew, N channel words, N hard bytes) is small enough that a CTA holding a few columns,
// one slot of program images and the interpreter's value slots fits shared memory at least twice per SM.
int lutldpc_decoder::init_onchip() {
  onchip_ok = onchip_fits = false;
  if (!m.minLUT) return LUTLDPC_OK;
  const int nvc = (int)vn_classes.size();
  std::vector<int> cls;
  for (int c = 0; c < nvc; ++c)
    if (vn_classes[c].n) cls.push_back(c);
  if (cls.empty() || (int)cls.size() > kOnchipMaxClasses) return LUTLDPC_OK;
  std::map<int, std::vector<int32_t>> rows;
  for (int c = 0; c < m.nchk; ++c) {
    std::vector<int32_t>& r = rows[m.dc[c]];
    r.insert(r.end(), m.cn_msg_idx.begin() + m.cn_start[c], m.cn_msg_idx.begin() + m.cn_start[c + 1]);
  }
  if ((int)rows.size() > kOnchipMaxClasses) return LUTLDPC_OK;
  auto al = [](size_t x, size_t a) { return (x + a - 1) / a * a; };
  const size_t E = (size_t)m.num_edges, N = (size_t)m.nvar;
  const size_t msgs_bytes = al((E + E / 32 + 1) * 4, 128), cha_bytes = al(N * 4, 128), hard_bytes = al(N, 128);
  oc_state_bytes = msgs_bytes + cha_bytes + hard_bytes;
  if (oc_state_bytes > (size_t)env_int("LUTLDPC_ONCHIP_MAX_STATE", 100 * 1024)) return LUTLDPC_OK;
  const int dec_slot = n_iter_slots - 1;
  // program images: class k of every iteration slot at the same offset, the decision slot laid out on its own
  memset(&oc, 0, sizeof oc);
  std::vector<uint32_t> it_sz(cls.size(), 0), dec_sz(cls.size(), 0);
  int n_slots_max = 0;
  bool any_interp_vn = false;
  oc_max_fast_dv = 0;
  for (size_t k = 0; k < cls.size(); ++k) {
    const int c = cls[k];
    bool fast = n_iter_slots >= 2;
    for (int t = 0; t < dec_slot; ++t) {
      const ProgRef& r = var_progs[(size_t)t * nvc + c];
      it_sz[k] = std::max(it_sz[k], r.bytes);
      fast &= r.fast_dv != 0;
    }
    for (int t = 0; t < dec_slot && !fast; ++t) n_slots_max = std::max(n_slots_max, (int)var_progs[(size_t)t * nvc + c].n_slots);
    any_interp_vn |= !fast;
    const ProgRef& rd = var_progs[(size_t)dec_slot * nvc + c];
    dec_sz[k] = rd.bytes;
    n_slots_max = std::max(n_slots_max, (int)rd.n_slots);
    oc.vc[k].pairs = vn_classes[c].pairs.p;
    oc.vc[k].n = vn_classes[c].n;
    oc.vc[k].dv = vn_classes[c].degree;
    oc.vc[k].fast = fast ? 1 : 0;
    if (fast) oc_max_fast_dv = std::max(oc_max_fast_dv, vn_classes[c].degree);
  }
  uint32_t off = 0, doff = 0;
  for (size_t k = 0; k < cls.size(); ++k) {
    oc.vc[k].img_off = off; off += (uint32_t)al(it_sz[k], 128);
    oc.vc[k].dec_off = doff; doff += (uint32_t)al(dec_sz[k], 128);
  }
  oc.n_vc = (int)cls.size();
  oc.slot_bytes = off; oc.dec_bytes = doff;
  oc.slot_stride = std::max(off, doff);
  // interpreted variable passes need value slots for every thread; the (rare) decision pass runs on two warps' worth
  oc.slot_threads = any_interp_vn ? kOnchipThreads : 64;
  const size_t tab_bytes = oc.slot_stride, slots_bytes = al((size_t)n_slots_max * oc.slot_threads * 4, 128);
  // columns per CTA: as many as leave room for about four CTAs per SM (LUTLDPC_ONCHIP_COLS overrides)
  const size_t fixed = tab_bytes + slots_bytes;
  CK(configure_onchip(), "configure onchip kernel");
  const int by_regs = std::max(1, onchip_ctas_per_sm(oc_max_fast_dv, 0));  // CTAs per SM the register file allows
  const size_t target = (size_t)kOnchipMaxSmem / std::max(1, std::min(by_regs, env_int("LUTLDPC_ONCHIP_CTAS", 4)));
  int C = target > fixed ? (int)((target - fixed) / oc_state_bytes) : 0;
  C = std::max(1, std::min(kOnchipMaxCols, C));
  if (int v = env_int("LUTLDPC_ONCHIP_COLS", 0)) C = std::max(1, std::min(kOnchipMaxCols, v));
  while (C > 1 && fixed + (size_t)C * oc_state_bytes > (size_t)kOnchipMaxSmem) --C;
  oc.C = C;
  oc.col_bytes = (uint32_t)oc_state_bytes;
  oc.off_cha = (uint32_t)msgs_bytes;
  oc.off_hard = (uint32_t)(msgs_bytes + cha_bytes);
  oc.off_tab = (uint32_t)((size_t)C * oc_state_bytes);
  oc.off_slots = (uint32_t)(oc.off_tab + tab_bytes);
  oc.smem_bytes = (uint32_t)(oc.off_slots + slots_bytes);
  if (oc.smem_bytes > (uint32_t)kOnchipMaxSmem) return LUTLDPC_OK;
  oc_ctas_per_sm = onchip_ctas_per_sm(oc_max_fast_dv, oc.smem_bytes);
  if (oc_ctas_per_sm < 1 || oc_ctas_per_sm * C < 2) return LUTLDPC_OK;
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device), "device attribute");
  oc_ctas = sms * oc_ctas_per_sm;
  // upload: slot images, iteration tables, check rows, work items
  std::vector<uint8_t> host_blob((size_t)d_prog_blob.n);
  CK(cudaMemcpy(host_blob.data(), d_prog_blob.p, host_blob.size(), cudaMemcpyDeviceToHost), "read back programs");
  std::vector<uint8_t> images((size_t)n_iter_slots * oc.slot_stride, 0);
  for (int t = 0; t < n_iter_slots; ++t)
    for (size_t k = 0; k < cls.size(); ++k) {
      const ProgRef& r = var_progs[(size_t)t * nvc + cls[k]];
      const uint32_t at = t == dec_slot ? oc.vc[k].dec_off : oc.vc[k].img_off;
      memcpy(images.data() + (size_t)t * oc.slot_stride + at, host_blob.data() + r.off, r.bytes);
    }
  CK(d_oc_images.upload(images), "upload slot images");
  std::vector<int32_t> q(m.max_iters);
  for (int i = 0; i < m.max_iters; ++i) q[i] = code_bits(m.Nq_Msg[i]);
  CK(d_oc_iter_slot.upload(m.var_tree_idx_iter), "upload iteration slots");
  CK(d_oc_iter_q.upload(q), "upload iteration widths");
  CK(d_oc_queue.ensure(4), "alloc queue");
  oc_cn = std::vector<NodeClass>(rows.size());
  int i = (int)rows.size() - 1;
  for (auto& kv : rows) {  // descending degree
    oc_cn[i].degree = kv.first;
    oc_cn[i].n = (int)(kv.second.size() / kv.first);
    CK(oc_cn[i].list.upload(kv.second), "upload cn rows");
    --i;
  }
  for (size_t k = 0; k < oc_cn.size(); ++k) oc.cc[k] = OnchipCn{oc_cn[k].list.p, oc_cn[k].n, oc_cn[k].degree};
  oc.n_cc = (int)oc_cn.size();
  // work items: 32 nodes of a class, the most expensive class first (its units are handed out first)
  std::vector<std::pair<int64_t, int4>> vi, ci;
  for (size_t k = 0; k < cls.size(); ++k) {
    const int64_t cost = std::max<int64_t>(1, var_progs[cls[k]].n_ops);
    for (int f = 0; f < oc.vc[k].n; f += 32) vi.push_back({cost, make_int4((int)k, f, std::min(32, oc.vc[k].n - f), 0)});
  }
  for (size_t k = 0; k < oc_cn.size(); ++k)
    for (int f = 0; f < oc_cn[k].n; f += 32) ci.push_back({oc_cn[k].degree, make_int4((int)k, f, std::min(32, oc_cn[k].n - f), 0)});
  auto by_cost = [](const std::pair<int64_t, int4>& a, const std::pair<int64_t, int4>& b2) { return a.first > b2.first; };
  std::stable_sort(vi.begin(), vi.end(), by_cost);
  std::stable_sort(ci.begin(), ci.end(), by_cost);
  std::vector<int4> flat;
  for (auto& x : vi) flat.push_back(x.second);
  oc.n_vn_items = (int)flat.size();
  CK(d_oc_vn_items.upload(flat), "upload vn items");
  flat.clear();
  for (auto& x : ci) flat.push_back(x.second);
  oc.n_cn_items = (int)flat.size();
  CK(d_oc_cn_items.upload(flat), "upload cn items");
  oc.vn_items = d_oc_vn_items.p; oc.cn_items = d_oc_cn_items.p;
  oc.slot_images = d_oc_images.p;
  oc.iter_slot = d_oc_iter_slot.p; oc.iter_q = d_oc_iter_q.p;
  oc.max_iters = m.max_iters; oc.dec_slot = dec_slot; oc.q_cha = code_bits(m.Nq_Cha);
  oc.nvar = m.nvar; oc.nchk = m.nchk; oc.num_edges = m.num_edges;
  oc.vn_edge0 = d_vn_edge0.p; oc.cn_start = d_cn_start.p; oc.chk_equ_idx = d_chk_equ_idx.p;
  oc.queue = d_oc_queue.p;
  onchip_fits = true;
  // Default: on chip only where it was measured faster than the HBM regime -- compile-time variable programs of degree
  // <= 8 for every class (the (3,6) N=1000 decoder: 4.9 against 4.45 Gbit/s).  Interpreted trees (file templates,
  // 2- and 3-bit messages) and high degrees (113-233 registers per thread: one CTA per SM) run 0.4-0.6x as fast on
  // chip as through HBM; they take the regime only on request (lutldpc_set_onchip, LUTLDPC_ONCHIP=2).
  const int mode = env_int("LUTLDPC_ONCHIP", 1);
  onchip_ok = mode == 2 || (mode == 1 && !any_interp_vn && oc_max_fast_dv <= 8);
  return LUTLDPC_OK;
}

// Expects cha / msgs / hard of the current batch to be initialised (pack or the channel kernel).
int lutldpc_decoder::run_passes(TraceSink* tr) {
  int rc = 0;
  const int n_frames = b.F;
  BatchBuffers cur = b;  // working layout: the batch itself until the first repack
  int rslot = 0;
  int n_live_known = n_frames;  // unfinished frames at the last waiting poll
  const int nvc = (int)vn_classes.size(), ncc = (int)cn_classes.size();
  const uint8_t* blob = d_prog_blob.p;
  cudaStream_t st = stream;
  auto record = [&](int kind, int it, int Nq) -> int {
    if (!tr || !tr->live || tr->n_rec >= tr->max_rec) return 0;
    CK(buf_trace.ensure((size_t)m.num_edges), "alloc trace");
    CK(launch_extract_frame(cur, m.num_edges, tr->frame, Nq, buf_trace.p, st), "extract_frame");
    CK(cudaMemcpyAsync(tr->msgs + (size_t)tr->n_rec * m.num_edges, buf_trace.p, m.num_edges, cudaMemcpyDeviceToHost, st),
       "trace copy");
    CK(cudaStreamSynchronize(st), "trace sync");
    tr->kind[tr->n_rec] = kind;
    tr->iter[tr->n_rec] = it;
    ++tr->n_rec;
    return 0;
  };
  auto traced_frame_done = [&](bool& done) -> int {  // after a finalize
    int32_t it = 0;
    uint32_t dw = 0;
    CK(cudaMemcpyAsync(&dw, cur.done + (tr->frame >> 3), 4, cudaMemcpyDeviceToHost, st), "trace done copy");
    CK(cudaMemcpyAsync(&it, cur.iters + tr->frame, 4, cudaMemcpyDeviceToHost, st), "trace iters copy");
    CK(cudaStreamSynchronize(st), "trace sync");
    done = (dw >> (4 * (tr->frame & 7))) & 1;
    return 0;
  };
  // counters of one finalize, as they arrive on the host: [0] unfinished frames, [2] live words, [3] live groups.
  // They only ever shrink, so a copy that is one iteration old still bounds every later launch.
  auto process_counts = [&](const uint32_t* hc, bool& all_done) -> int {
    all_done = hc[0] == 0;
    cur.host_cols = (int)hc[2];
    cur.host_grp4 = (int)hc[3];
    // frame repacking: too many slots of the live words belong to frames that have already returned
    const int n_live = (int)hc[0];
    n_live_known = n_live;
    const int dense_cols = (n_live + 7) / 8;
    if (repack_on && !tr && n_live > 0 && (double)n_live < repack_alpha * 8.0 * cur.host_cols &&
        (std::max(32, ((n_live + 255) / 256) * 32) < cur.W || 2 * dense_cols <= cur.host_cols)) {
      int rc2 = repack_into(cur, rslot, n_live);
      if (rc2) return rc2;
      rslot ^= 1;
    }
    return 0;
  };
  auto poll = [&](bool& all_done) -> int {  // waiting poll: exact counts, the stream drains
    all_done = false;
    if (poll_every <= 0) return 0;
    // the finalize kernel wrote the counters into host-mapped memory: no copy (a D2H memcpy on this stream would
    // queue behind the result transfer of the previous chunk on the copy engine and stall the whole schedule)
    CK(cudaStreamSynchronize(st), "poll sync");
    return process_counts(h_active, all_done);
  };
  bool all_done = false;
  bool lagging = lag_poll && !tr && poll_every == 1, lag_inflight = false;
  bool lag_skip = false;
  const int tail_lag = lag_poll && !tr && poll_every == 1 ? tail_lag_frames : -1;
  if (tr) tr->live = true;
  if (pisc) {  // :277
    CK(launch_syndrome(g, cur, st), "syndrome"); launched(KIND_SYNDROME);
    CK(launch_finalize(FIN_INIT, 0, cur, st), "finalize"); launched(KIND_FINALIZE);
    if (tr) { bool d = false; if ((rc = traced_frame_done(d))) return rc; if (d) tr->live = false; }
    if ((rc = poll(all_done))) return rc;
  }
  if ((rc = record(LUTLDPC_TRACE_INIT, -1, m.Nq_Msg[0]))) return rc;
  int polled = 0;
  for (int ii = 0; ii < max_iters && !all_done; ++ii) {
    const int t = m.var_tree_idx_iter[ii];
    const int want = psc && ii > 0;
    // check-node pass (:304-310); its sign parities complete the early-exit test of VN pass ii-1
    if (m.minLUT && use_cn_exact) {
      for (auto& cl : cn_exact) {
        CK(launch_cn_exact(cl.degree, cl.list.p, cl.n, cur, code_bits(m.Nq_Msg[ii]), want, st), "cn_exact");
        launched(KIND_CN);
      }
    } else if (m.minLUT) {
      CK(launch_cn_minsum(g, cur, code_bits(m.Nq_Msg[ii]), want, st), "cn_minsum"); launched(KIND_CN);
    } else {
      for (int c = 0; c < ncc; ++c) {
        const ProgRef& r = chk_progs[(size_t)t * ncc + c];
        NodeLaunch nl{blob + r.off, r.bytes, r.n_slots, r.n_in, cn_classes[c].list.p, nullptr, cn_classes[c].n};
        CK(launch_node(NODE_CHK, nl, g, cur, want, st), "chk_lut"); launched(KIND_CHK0 + std::min(c, 15));
      }
    }
    if ((rc = record(LUTLDPC_TRACE_CN, ii, m.Nq_Msg[ii]))) return rc;
    if (want) {  // psc: VN pass ii-1 unanimous and syndrome zero -> return ii (:327)
      CK(launch_finalize(FIN_ITER, ii, cur, st), "finalize"); launched(KIND_FINALIZE);
      if (tr && tr->live) {
        bool d = false;
        if ((rc = traced_frame_done(d))) return rc;
        if (d) {  // the reference returned before dumping VN(ii-1) and never ran CN(ii)
          tr->n_rec = std::max(0, tr->n_rec - 2);
          tr->live = false;
        }
      }
      if (poll_every > 0 && ++polled % poll_every == 0) {
        // Lagged poll: the counters are copied without waiting and looked at one iteration later, so the stream
        // never drains.  Used (a) while no frame of the batch has returned yet -- there is nothing to decide;
        // the first copy that shows a returned frame switches to the waiting poll -- and (b) once few frames are
        // left (tail_lag): those iterations are short, a drained stream would cost as much as the passes, and
        // counts that are one iteration old still bound the grids and the repacked layout.
        const bool tail = !lagging && n_live_known <= tail_lag;
        bool wait = !lagging && !tail;
        if (lagging || tail) {
          if (lag_inflight) {
            CK(cudaEventSynchronize(ev_lag), "poll sync");
            lag_inflight = false;
            if (tail) {
              // the mirror may still hold the counts of the layout that was just left: skip one reading after a repack
              if (lag_skip) lag_skip = false;
              else {
                const int64_t r0 = repacks;
                if ((rc = process_counts(h_active, all_done))) return rc;
                if (all_done) break;
                lag_skip = repacks != r0;
              }
            } else if (h_active[0] != (uint32_t)n_live_known) {
              lagging = false;
              wait = true;
            }
          }
          if (!wait) {
            CK(cudaEventRecord(ev_lag, st), "poll event");  // the counters arrive through the mapped mirror
            lag_inflight = true;
          }
        }
        if (wait) {
          if ((rc = poll(all_done))) return rc;
          if (all_done) break;
        }
      }
    }
    if (ii != max_iters - 1) {  // variable-node pass (:320-325)
      if (!fused_classes.empty()) {
        FusedVnParams P;
        memset(&P, 0, sizeof P);
        for (size_t j = 0; j < fused_classes.size(); ++j) {
          const int c = fused_classes[j];
          const ProgRef& r = var_progs[(size_t)t * nvc + c];
          P.c[j] = FusedVnClass{blob + r.off + r.tables_off, vn_classes[c].pairs.p, r.fast_dv, r.q_out};
        }
        CK(launch_vn_fused(P, fused_max_dv, d_fused_tiles.p, n_fused_tiles, g, cur, psc, st), "var_fused");
        launched(KIND_VN0 + 15);
      }
      for (int c = 0; c < nvc; ++c) {
        if (!vn_classes[c].n) continue;
        if (std::find(fused_classes.begin(), fused_classes.end(), c) != fused_classes.end()) continue;
        const ProgRef& r = var_progs[(size_t)t * nvc + c];
        NodeLaunch nl{blob + r.off, r.bytes, r.n_slots, r.n_in, vn_classes[c].list.p, vn_classes[c].pairs.p, vn_classes[c].n};
        if (r.fast_dv) CK(launch_vn_fast(r.fast_dv, blob + r.off + r.tables_off, r.q_out, nl, g, cur, psc, st), "var_fast");
        else CK(launch_node(NODE_VAR, nl, g, cur, psc, st), "var_lut");
        launched(KIND_VN0 + std::min(c, 15));
      }
      if ((rc = record(LUTLDPC_TRACE_VN, ii, m.Nq_Msg[ii + 1]))) return rc;
    } else if ((rc = record(LUTLDPC_TRACE_VN, ii, m.Nq_Msg[ii]))) {  // the reference dumps here too (:331)
      return rc;
    }
  }
  if (!all_done) {  // decision + final syndrome (:340-349)
    const int t = m.var_tree_idx_iter[max_iters - 1];
    for (int c = 0; c < nvc; ++c) {
      if (!vn_classes[c].n) continue;
      const ProgRef& r = var_progs[(size_t)t * nvc + c];
      NodeLaunch nl{blob + r.off, r.bytes, r.n_slots, r.n_in, vn_classes[c].list.p, vn_classes[c].pairs.p, vn_classes[c].n};
      CK(launch_node(NODE_DEC, nl, g, cur, 0, st), "dec_lut"); launched(KIND_DEC0 + std::min(c, 15));
    }
    CK(launch_syndrome(g, cur, st), "syndrome"); launched(KIND_SYNDROME);
    CK(launch_finalize(FIN_LAST, max_iters, cur, st), "finalize"); launched(KIND_FINALIZE);
  }
  if ((rc = retire_from(cur))) return rc;
  CK(cudaEventRecord(ev1, st), "event");
  have_stats = true;
  last_frames = n_frames;
  stat_frame_iters = -1;
  return LUTLDPC_OK;
}
