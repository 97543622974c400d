// See model.hpp.  Host-only C++ (no CUDA in this file).
#include "model.hpp"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>

namespace lutldpc {

// ------------------------------------------------------------------ tree text format
namespace {
struct Cursor {
  const char* p;
  const char* end;
  bool next(int& v) {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p;
    if (p >= end) return false;
    bool neg = false;
    if (*p == '-') { neg = true; ++p; }
    long acc = 0;
    int nd = 0;
    while (p < end && *p >= '0' && *p <= '9') { acc = acc * 10 + (*p - '0'); ++p; ++nd; }
    if (!nd) return false;
    v = (int)(neg ? -acc : acc);
    return true;
  }
};

// One node record: "<#children>\n<type> <|Q|> <K>\n[Q ...]" followed by the children, pre-order
// (format documented in the reference's trees/README.md:87-96).
int parse_node(Cursor& c, Tree& t, int depth) {
  int nchild, type, qlen, K;
  if (depth > 4096) return -1;
  if (!c.next(nchild) || !c.next(type) || !c.next(qlen) || !c.next(K)) return -1;
  if (nchild < 0 || qlen < 0 || type < 0 || type > 3) return -1;
  int me = (int)t.nodes.size();
  t.nodes.emplace_back();
  t.nodes[me].type = type;
  t.nodes[me].K = K;
  t.nodes[me].Q.resize(qlen);
  for (int i = 0; i < qlen; ++i)
    if (!c.next(t.nodes[me].Q[i])) return -1;
  for (int i = 0; i < nchild; ++i) {
    int ch = parse_node(c, t, depth + 1);
    if (ch < 0) return -1;
    t.nodes[me].child.push_back(ch);
  }
  return me;
}
}  // namespace

bool parse_tree_array(const char* txt, size_t len, TreeArray& out, std::string& err) {
  Cursor c{txt, txt + len};
  out = TreeArray();
  int ni;
  if (!c.next(ni) || ni < 0) { err = "tree array: missing iteration-slot count"; return false; }
  for (int ii = 0; ii < ni; ++ii) {
    int nd;
    if (!c.next(nd) || nd < 0) { err = "tree array: missing degree-slot count"; return false; }
    if (ii == 0) out.n_deg = nd;
    else if (nd != out.n_deg) { err = "tree array: ragged degree slots"; return false; }
    for (int dd = 0; dd < nd; ++dd) {
      Tree t;
      if (!c.next(t.tree_type) || !c.next(t.num_leaves)) { err = "tree array: short tree header"; return false; }
      if (parse_node(c, t, 0) != 0) { err = "tree array: malformed node record"; return false; }
      out.t.push_back(std::move(t));
    }
  }
  out.n_iter = ni;
  return true;
}

// ------------------------------------------------------------------ bundle file
namespace {
struct Reader {
  FILE* f;
  bool ok = true;
  void raw(void* dst, size_t n) { if (ok && n && fread(dst, 1, n, f) != n) ok = false; }
  template <class T> void vec(std::vector<T>& v, size_t n) { v.resize(n); raw(v.data(), n * sizeof(T)); }
};
}  // namespace

// Layout: oracle/ref_driver.cpp write_bundle (the dump of a reference LDPC_Code_LUT object).
bool load_bundle_file(const char* path, DecoderModel& m, std::string& err) {
  FILE* f = fopen(path, "rb");
  if (!f) { err = std::string("cannot open bundle ") + path; return false; }
  Reader r{f};
  char magic[8];
  int32_t h[10];
  r.raw(magic, 8);
  r.raw(h, sizeof h);
  if (!r.ok || memcmp(magic, "LUTLDPC1", 8) != 0) { fclose(f); err = "bundle: bad magic/header"; return false; }
  m = DecoderModel();
  m.nvar = h[0]; m.nchk = h[1]; m.nchk_lin_indep = h[2]; m.num_edges = h[3];
  m.max_iters = h[4]; m.Nq_Cha = h[5]; m.minLUT = h[6] != 0;
  int n_qb_msg = h[7], has_gen = h[8];
  if (m.nvar <= 0 || m.nchk <= 0 || m.num_edges <= 0 || m.max_iters <= 0 || m.Nq_Cha < 2 || n_qb_msg < 0 ||
      m.nvar > (1 << 26) || m.num_edges > (1 << 28) || m.max_iters > (1 << 16) || m.Nq_Cha > 16 || n_qb_msg > 64) {
    fclose(f); err = "bundle: header out of range"; return false;
  }
  r.vec(m.dv, m.nvar); r.vec(m.dc, m.nchk); r.vec(m.cn_msg_idx, m.num_edges);
  r.vec(m.Nq_Msg, m.max_iters); r.vec(m.reuse, m.max_iters); r.vec(m.cha2msg, m.Nq_Cha);
  r.vec(m.qb_Cha, m.Nq_Cha - 1); r.vec(m.qb_Msg, n_qb_msg);
  for (int k = 0; k < 2 && r.ok; ++k) {
    int32_t len = 0;
    r.raw(&len, 4);
    if (!r.ok || len < 0) { r.ok = false; break; }
    std::vector<char> txt((size_t)len + 1, 0);
    r.raw(txt.data(), (size_t)len);
    if (!r.ok) break;
    TreeArray ta;
    std::string perr;
    bool parsed = parse_tree_array(txt.data(), (size_t)len, ta, perr);
    if (k == 0) {
      if (!parsed) { fclose(f); err = "bundle: var trees: " + perr; return false; }
      m.var_trees = std::move(ta);
    } else if (parsed) {
      m.chk_trees = std::move(ta);
    }
  }
  if (r.ok && has_gen) {
    int32_t rc[2];
    r.raw(rc, 8);
    if (r.ok && rc[0] >= 0 && rc[1] >= 0) {
      m.gen_rows = rc[0]; m.gen_cols = rc[1];
      r.vec(m.gen, (size_t)rc[0] * rc[1]);
    } else r.ok = false;
  }
  fclose(f);
  if (!r.ok) { err = "bundle: truncated file"; return false; }
  return finalize_model(m, err);
}

bool finalize_model(DecoderModel& m, std::string& err) {
  char buf[256];
  if ((int)m.dv.size() != m.nvar || (int)m.dc.size() != m.nchk || (int)m.cn_msg_idx.size() != m.num_edges ||
      (int)m.Nq_Msg.size() != m.max_iters || (int)m.reuse.size() != m.max_iters) {
    err = "model: array sizes do not match the header"; return false;
  }
  long sdv = 0, sdc = 0;
  for (int d : m.dv) { if (d < 1) { err = "model: variable node of degree < 1"; return false; } sdv += d; }
  for (int d : m.dc) {
    // a degree-1 check makes the reference emit an out-of-range magnitude (min2 keeps its sentinel,
    // src/LDPC_Code_LUT.cpp:362-389); refuse such graphs instead of reproducing undefined labels
    if (d < 2) { err = "model: check node of degree < 2 is not decodable by the LUT decoder"; return false; }
    sdc += d;
  }
  if (sdv != m.num_edges || sdc != m.num_edges) { err = "model: degree sums != num_edges"; return false; }
  for (int e : m.cn_msg_idx) if (e < 0 || e >= m.num_edges) { err = "model: cn_msg_idx out of range"; return false; }
  if (m.Nq_Cha < 2 || m.Nq_Cha > 16 || (m.Nq_Cha & 1)) { err = "model: Nq_Cha must be even and <= 16"; return false; }
  for (int q : m.Nq_Msg) if (q < 2 || q > 16 || (q & 1)) { err = "model: Nq_Msg must be even and <= 16"; return false; }
  if ((int)m.qb_Cha.size() != m.Nq_Cha - 1) { err = "model: qb_Cha must hold Nq_Cha-1 thresholds"; return false; }
  if (!m.cha2msg.empty() && (int)m.cha2msg.size() != m.Nq_Cha) { err = "model: Nq_Cha_2_Nq_Msg_map size"; return false; }
  // set_trees, src/LDPC_Code_LUT.cpp:120-160
  if (m.reuse[0] != 0 || m.reuse[m.max_iters - 1] != 0) { err = "model: reuse_vec must start and end with 0"; return false; }
  m.var_tree_idx_iter.resize(m.max_iters);
  int acc = 0;
  for (int i = 0; i < m.max_iters; ++i) { acc += (m.reuse[i] == 0); m.var_tree_idx_iter[i] = acc - 1; }
  if (m.var_trees.n_iter != acc) {
    snprintf(buf, sizeof buf, "model: %d var-tree iteration slots, reuse_vec needs %d", m.var_trees.n_iter, acc);
    err = buf; return false;
  }
  m.var_tree_idx_degree.assign(m.nvar, -1);
  for (int v = 0; v < m.nvar; ++v) {
    for (int dd = 0; dd < m.var_trees.n_deg; ++dd)
      if (m.var_trees.at(0, dd).num_leaves == m.dv[v]) { m.var_tree_idx_degree[v] = dd; break; }
    if (m.var_tree_idx_degree[v] < 0) {
      snprintf(buf, sizeof buf, "model: no variable-node tree with %d leaves", m.dv[v]);
      err = buf; return false;
    }
  }
  if (!m.minLUT) {
    // chk_tree_idx_iter == var_tree_idx_iter, src/LDPC_Code_LUT.cpp:146
    if (m.chk_trees.n_iter < acc) { err = "model: min_lut=false but check-node trees are missing"; return false; }
    m.chk_tree_idx_degree.assign(m.nchk, -1);
    for (int c = 0; c < m.nchk; ++c) {
      for (int dd = 0; dd < m.chk_trees.n_deg; ++dd)
        if (m.chk_trees.at(0, dd).num_leaves + 1 == m.dc[c]) { m.chk_tree_idx_degree[c] = dd; break; }
      if (m.chk_tree_idx_degree[c] < 0) { err = "model: no check-node tree for a check degree"; return false; }
    }
  }
  m.vn_edge0.resize(m.nvar + 1);
  m.vn_edge0[0] = 0;
  for (int v = 0; v < m.nvar; ++v) m.vn_edge0[v + 1] = m.vn_edge0[v] + m.dv[v];
  m.cn_start.resize(m.nchk + 1);
  m.cn_start[0] = 0;
  for (int c = 0; c < m.nchk; ++c) m.cn_start[c + 1] = m.cn_start[c] + m.dc[c];
  // chk_equ_idx (src/LDPC_Code_LUT.cpp:531-535): the VN owning each listed edge
  std::vector<int32_t> edge_vn(m.num_edges);
  for (int v = 0; v < m.nvar; ++v)
    for (int e = m.vn_edge0[v]; e < m.vn_edge0[v + 1]; ++e) edge_vn[e] = v;
  m.chk_equ_idx.resize(m.num_edges);
  std::vector<uint8_t> seen(m.num_edges, 0);
  for (int i = 0; i < m.num_edges; ++i) {
    if (seen[m.cn_msg_idx[i]]) { err = "model: cn_msg_idx lists an edge twice"; return false; }
    seen[m.cn_msg_idx[i]] = 1;
    m.chk_equ_idx[i] = edge_vn[m.cn_msg_idx[i]];
  }
  if (m.nchk_lin_indep < 0 || m.nchk_lin_indep > m.nvar) { err = "model: nchk_lin_indep out of range"; return false; }
  return true;
}

// ------------------------------------------------------------------ storage code
int code_bits(int Nq) {
  int q = 1;
  while ((1 << q) < Nq) ++q;
  return q;
}
int label_to_code(int m, int Nq) {
  int nz = Nq / 2, q = code_bits(Nq);
  return m < nz ? ((1 << (q - 1)) | (nz - 1 - m)) : (m - nz);
}
int code_to_label(int c, int Nq) {
  int nz = Nq / 2, q = code_bits(Nq);
  int sign = (c >> (q - 1)) & 1, mag = c & ((1 << (q - 1)) - 1);
  if (mag >= nz || c >= (1 << q)) return -1;
  return sign ? nz - 1 - mag : nz + mag;
}
