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
      m.var_tree_text.assign(txt.data(), (size_t)len);
    } else {
      if (parsed) m.chk_trees = std::move(ta);
      m.chk_tree_text.assign(txt.data(), (size_t)len);
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

// ---------------------------------------------------------------------------------------------
// The reference's own codec file: IT++ it_file (format version 3) holding the variables that
// LDPC_Code_LUT::save_code writes (src/LDPC_Code_LUT.cpp:643-697) and load_code reads (:568-641).
// Layout (itpp/itpp/base/itfile.h:98-125, itfile.cpp): "IT++" 0x03, then blocks of
//   u64 hdr_bytes, u64 data_bytes, u64 block_bytes, name\0 type\0 desc\0, data
// little endian; int32 = 4 bytes, bool = 1 byte, ivec/dvec/bvec/string = u64 count + elements
// (int32 / f64 / 1 byte / char).  chk_equ_idx (ivecArray) is not needed: it is re-derived from
// cn_msg_idx like in decoder_parameterization (:515-535).
// ---------------------------------------------------------------------------------------------
namespace {
struct ItVar {
  std::string type;
  const uint8_t* p = nullptr;
  uint64_t n = 0;
};
struct ItFile {
  std::vector<uint8_t> buf;
  std::vector<std::pair<std::string, ItVar>> vars;  // file order; names may repeat (the generator appends its own)
  const ItVar* find(const char* name, const char* type) const {
    for (auto& kv : vars)
      if (kv.first == name && kv.second.type == type) return &kv.second;
    return nullptr;
  }
  bool i32(const char* name, int32_t& v) const {
    const ItVar* x = find(name, "int32");
    if (!x || x->n != 4) return false;
    memcpy(&v, x->p, 4);
    return true;
  }
  bool flag(const char* name, bool& v) const {
    const ItVar* x = find(name, "bool");
    if (!x || x->n != 1) return false;
    v = x->p[0] != 0;
    return true;
  }
  template <class T>
  bool arr(const char* name, const char* type, size_t elem, std::vector<T>& out) const {
    const ItVar* x = find(name, type);
    uint64_t cnt = 0;
    if (!x || x->n < 8) return false;
    memcpy(&cnt, x->p, 8);
    if (cnt > (1ull << 31) || x->n != 8 + cnt * elem) return false;
    out.resize((size_t)cnt);
    for (size_t i = 0; i < (size_t)cnt; ++i) {
      if (elem == 4) { int32_t v; memcpy(&v, x->p + 8 + 4 * i, 4); out[i] = (T)v; }
      else if (elem == 8) { double v; memcpy(&v, x->p + 8 + 8 * i, 8); out[i] = (T)v; }
      else out[i] = (T)x->p[8 + i];
    }
    return true;
  }
};

bool read_it_file(const char* path, ItFile& f, std::string& err) {
  FILE* fp = fopen(path, "rb");
  if (!fp) { err = std::string("cannot open codec file ") + path; return false; }
  fseek(fp, 0, SEEK_END);
  long sz = ftell(fp);
  fseek(fp, 0, SEEK_SET);
  if (sz < 5) { fclose(fp); err = "codec file: too short"; return false; }
  f.buf.resize((size_t)sz);
  size_t got = fread(f.buf.data(), 1, (size_t)sz, fp);
  fclose(fp);
  if (got != (size_t)sz) { err = "codec file: read error"; return false; }
  if (memcmp(f.buf.data(), "IT++", 4) != 0 || f.buf[4] != 3) { err = "codec file: not an IT++ file of format version 3"; return false; }
  size_t off = 5;
  while (off + 24 <= f.buf.size()) {
    uint64_t h[3];
    memcpy(h, f.buf.data() + off, 24);
    if (h[2] < 24 || h[0] < 27 || h[2] > f.buf.size() - off || h[0] > h[2] || h[1] > h[2] - h[0]) { err = "codec file: corrupt block header"; return false; }
    const char* s = reinterpret_cast<const char*>(f.buf.data() + off + 24);
    const char* end = reinterpret_cast<const char*>(f.buf.data() + off + h[0]);
    std::string str[3];
    for (int k = 0; k < 3; ++k) {
      const void* z = memchr(s, 0, (size_t)(end - s));
      if (!z) { err = "codec file: corrupt block strings"; return false; }
      str[k].assign(s, (const char*)z);
      s = (const char*)z + 1;
    }
    ItVar v;
    v.type = str[1];
    v.p = f.buf.data() + off + h[0];
    v.n = h[1];
    if (!v.type.empty()) f.vars.push_back({str[0], v});  // type "" marks a deleted variable
    off += (size_t)h[2];
  }
  return true;
}
}  // namespace

bool load_it_codec_file(const char* path, DecoderModel& m, std::string& err) {
  ItFile f;
  if (!read_it_file(path, f, err)) return false;
  int32_t ver = 0, nvar = 0, nchk = 0, nli = 0, max_iters = 0, nq_cha = 0;
  bool h_def = false, luts_def = false, min_lut = false;
  if (!f.i32("Fileversion", ver) || ver != 1) { err = "codec file: unsupported Fileversion (src/LDPC_Code_LUT.cpp:35)"; return false; }
  if (!f.flag("H_defined", h_def) || !f.flag("LUTs_defined", luts_def) || !h_def || !luts_def) {
    err = "codec file: no parity-check matrix or no LUTs stored";
    return false;
  }
  m = DecoderModel();
  std::vector<uint8_t> reuse;
  std::vector<char> vt, ct;
  bool ok = f.i32("nvar", nvar) && f.i32("nchk", nchk) && f.i32("nchk_lin_indep", nli) && f.i32("max_iters", max_iters) &&
            f.i32("Nq_Cha", nq_cha) && f.flag("minLUT", min_lut) && f.arr("dv_vec", "ivec", 4, m.dv) &&
            f.arr("dc_vec", "ivec", 4, m.dc) && f.arr("cn_msg_idx", "ivec", 4, m.cn_msg_idx) &&
            f.arr("Nq_Msg", "ivec", 4, m.Nq_Msg) && f.arr("qb_Cha", "dvec", 8, m.qb_Cha) && f.arr("reuse_vec", "bvec", 1, reuse) &&
            f.arr("var_tree_string", "string", 1, vt);
  if (!ok) { err = "codec file: a variable of LDPC_Code_LUT::save_code is missing or has the wrong type"; return false; }
  f.arr("qb_Msg", "dvec", 8, m.qb_Msg);
  f.arr("Nq_Cha_2_Nq_Msg_map", "ivec", 4, m.cha2msg);
  f.arr("chk_tree_string", "string", 1, ct);
  m.nvar = nvar; m.nchk = nchk; m.nchk_lin_indep = nli; m.max_iters = max_iters; m.Nq_Cha = nq_cha; m.minLUT = min_lut;
  m.num_edges = (int)m.cn_msg_idx.size();
  m.reuse.assign(reuse.begin(), reuse.end());
  if (m.nvar <= 0 || m.nchk <= 0 || m.num_edges <= 0 || m.max_iters <= 0 || m.nvar > (1 << 26) || m.num_edges > (1 << 28)) {
    err = "codec file: header out of range";
    return false;
  }
  std::string perr;
  if (!parse_tree_array(vt.data(), vt.size(), m.var_trees, perr)) { err = "codec file: var trees: " + perr; return false; }
  m.var_tree_text.assign(vt.begin(), vt.end());
  m.chk_tree_text.assign(ct.begin(), ct.end());
  if (!ct.empty()) {
    TreeArray ta;
    if (parse_tree_array(ct.data(), ct.size(), ta, perr)) m.chk_trees = std::move(ta);
    else if (!m.minLUT) { err = "codec file: chk trees: " + perr; return false; }
  }
  // generator appended by LDPC_Generator_Systematic::save (itpp/itpp/comm/ldpc.cpp:1002-1009): GF2mat "G" =
  // u64 nrows, ncols, nwords + rows of bytes, bit k of byte j = column 8j+k (itpp/itpp/base/gf2mat.cpp:1066-1082)
  if (const ItVar* g = f.find("G", "GF2mat")) {
    uint64_t d[3] = {0, 0, 0};
    if (g->n >= 24) memcpy(d, g->p, 24);
    if (g->n >= 24 && d[0] > 0 && d[1] > 0 && d[2] == (d[1] + 7) / 8 && g->n == 24 + d[0] * d[2] && d[0] * d[1] < (1ull << 33)) {
      m.gen_rows = (int)d[0]; m.gen_cols = (int)d[1];
      m.gen.assign((size_t)d[0] * d[1], 0);
      for (uint64_t r = 0; r < d[0]; ++r)
        for (uint64_t c = 0; c < d[1]; ++c)
          m.gen[(size_t)(r * d[1] + c)] = (g->p[24 + r * d[2] + (c >> 3)] >> (c & 7)) & 1;
    }
  }
  return finalize_model(m, err);
}

bool save_it_codec_file(const char* path, const DecoderModel& m, std::string& err) {
  if (m.var_tree_text.empty()) { err = "save_code: the decoder holds no tree text"; return false; }
  FILE* fp = fopen(path, "wb");
  if (!fp) { err = std::string("save_code: cannot open ") + path; return false; }
  bool ok = fwrite("IT++\x03", 1, 5, fp) == 5;
  auto block = [&](const char* name, const char* type, const std::string& data) {
    const uint64_t hdr = 24 + strlen(name) + 1 + strlen(type) + 1 + 1, db = data.size(), blk = hdr + db;
    ok = ok && fwrite(&hdr, 8, 1, fp) == 1 && fwrite(&db, 8, 1, fp) == 1 && fwrite(&blk, 8, 1, fp) == 1;
    ok = ok && fwrite(name, 1, strlen(name) + 1, fp) == strlen(name) + 1 && fwrite(type, 1, strlen(type) + 1, fp) == strlen(type) + 1;
    ok = ok && fputc(0, fp) == 0;  // empty description
    ok = ok && (data.empty() || fwrite(data.data(), 1, data.size(), fp) == data.size());
  };
  auto i32 = [&](const char* name, int32_t v) { block(name, "int32", std::string(reinterpret_cast<const char*>(&v), 4)); };
  auto flag = [&](const char* name, bool v) { block(name, "bool", std::string(1, v ? '\1' : '\0')); };
  auto counted = [](uint64_t n) { return std::string(reinterpret_cast<const char*>(&n), 8); };
  auto ivec = [&](const char* name, const std::vector<int32_t>& v) {
    std::string d = counted(v.size());
    d.append(reinterpret_cast<const char*>(v.data()), 4 * v.size());
    block(name, "ivec", d);
  };
  auto dvec = [&](const char* name, const std::vector<double>& v) {
    std::string d = counted(v.size());
    d.append(reinterpret_cast<const char*>(v.data()), 8 * v.size());
    block(name, "dvec", d);
  };
  i32("Fileversion", 1);  // LUT_LDPC_binary_file_version, src/LDPC_Code_LUT.cpp:35
  flag("H_defined", true);
  flag("G_defined", false);
  flag("LUTs_defined", true);
  i32("nvar", m.nvar);
  i32("nchk", m.nchk);
  i32("nchk_lin_indep", m.nchk_lin_indep);
  ivec("dv_vec", m.dv);
  ivec("dc_vec", m.dc);
  {  // chk_equ_idx: Array<ivec>, the sorted variable indices of every check (:531-535)
    std::string d = counted((uint64_t)m.nchk);
    for (int c = 0; c < m.nchk; ++c) {
      const int a = m.cn_start[c], b = m.cn_start[c + 1];
      d += counted((uint64_t)(b - a));
      d.append(reinterpret_cast<const char*>(m.chk_equ_idx.data() + a), 4 * (size_t)(b - a));
    }
    block("chk_equ_idx", "ivecArray", d);
  }
  ivec("cn_msg_idx", m.cn_msg_idx);
  i32("Nq_Cha", m.Nq_Cha);
  ivec("Nq_Msg", m.Nq_Msg);
  ivec("Nq_Cha_2_Nq_Msg_map", m.cha2msg);
  dvec("qb_Cha", m.qb_Cha);
  dvec("qb_Msg", m.qb_Msg);
  {
    std::string d = counted(m.reuse.size());
    for (int32_t r : m.reuse) d.push_back(r ? '\1' : '\0');
    block("reuse_vec", "bvec", d);
  }
  flag("minLUT", m.minLUT);
  i32("output_verbosity", 0);
  i32("max_iters", m.max_iters);
  block("var_tree_string", "string", counted(m.var_tree_text.size()) + m.var_tree_text);
  block("chk_tree_string", "string", counted(m.chk_tree_text.size()) + m.chk_tree_text);
  ok = (fclose(fp) == 0) && ok;
  if (!ok) err = std::string("save_code: write error on ") + path;
  return ok;
}

bool load_bp_codec_file(const char* path, BpCodecFile& o, std::string& err) {
  ItFile f;
  if (!read_it_file(path, f, err)) return false;
  int32_t ver = 0;
  bool h_def = false, g_def = false;
  if (!f.i32("Fileversion", ver) || ver != 3) { err = "LDPC_Code::load_code(): Unsupported file format"; return false; }  // ldpc.cpp:42,1258-1260
  if (!f.flag("H_defined", h_def) || !h_def) { err = "bp codec file: no parity-check matrix stored"; return false; }
  f.flag("G_defined", g_def);
  o = BpCodecFile();
  int32_t nvar = 0, nchk = 0, nli = 0;
  bool ok = f.i32("nvar", nvar) && f.i32("ncheck", nchk) && f.i32("ncheck_lin_indep", nli) && f.arr("C", "ivec", 4, o.C) &&
            f.arr("V", "ivec", 4, o.V) && f.arr("sumX1", "ivec", 4, o.sumX1) && f.arr("sumX2", "ivec", 4, o.sumX2) &&
            f.arr("iind", "ivec", 4, o.iind) && f.arr("jind", "ivec", 4, o.jind);
  if (!ok || nvar <= 0 || nchk <= 0 || (int)o.sumX1.size() != nvar || (int)o.sumX2.size() != nchk || o.V.size() % (size_t)nchk ||
      o.C.size() % (size_t)nvar) {
    err = "bp codec file: a variable of LDPC_Code::save_code is missing or has the wrong size";
    return false;
  }
  o.nvar = nvar; o.ncheck = nchk; o.ncheck_lin_indep = nli;
  if (g_def) {
    const ItVar* g = f.find("G", "GF2mat");
    uint64_t d[3] = {0, 0, 0};
    if (g && g->n >= 24) memcpy(d, g->p, 24);
    if (!g || d[0] == 0 || d[1] == 0 || d[2] != (d[1] + 7) / 8 || g->n != 24 + d[0] * d[2] || d[0] + d[1] != (uint64_t)nvar) {
      err = "bp codec file: generator block missing or inconsistent";
      return false;
    }
    o.gen_rows = (int)d[0]; o.gen_cols = (int)d[1];
    o.gen_bytes.assign(g->p + 24, g->p + 24 + d[0] * d[2]);
  }
  return true;
}

bool save_bp_codec_file(const char* path, const BpCodecFile& m, std::string& err) {
  FILE* fp = fopen(path, "wb");
  if (!fp) { err = std::string("save_code: cannot open ") + path; return false; }
  bool ok = fwrite("IT++\x03", 1, 5, fp) == 5;
  auto block = [&](const char* name, const char* type, const std::string& data) {
    const uint64_t hdr = 24 + strlen(name) + 1 + strlen(type) + 1 + 1, db = data.size(), blk = hdr + db;
    ok = ok && fwrite(&hdr, 8, 1, fp) == 1 && fwrite(&db, 8, 1, fp) == 1 && fwrite(&blk, 8, 1, fp) == 1;
    ok = ok && fwrite(name, 1, strlen(name) + 1, fp) == strlen(name) + 1 && fwrite(type, 1, strlen(type) + 1, fp) == strlen(type) + 1;
    ok = ok && fputc(0, fp) == 0;
    ok = ok && (data.empty() || fwrite(data.data(), 1, data.size(), fp) == data.size());
  };
  auto i32 = [&](const char* name, int32_t v) { block(name, "int32", std::string(reinterpret_cast<const char*>(&v), 4)); };
  auto flag = [&](const char* name, bool v) { block(name, "bool", std::string(1, v ? '\1' : '\0')); };
  auto counted = [](uint64_t n) { return std::string(reinterpret_cast<const char*>(&n), 8); };
  auto ivec = [&](const char* name, const std::vector<int32_t>& v) {
    std::string d = counted(v.size());
    d.append(reinterpret_cast<const char*>(v.data()), 4 * v.size());
    block(name, "ivec", d);
  };
  i32("Fileversion", 3);  // LDPC_binary_file_version, itpp/itpp/comm/ldpc.cpp:42
  flag("H_defined", true);
  flag("G_defined", m.gen_rows > 0);
  i32("nvar", m.nvar);
  i32("ncheck", m.ncheck);
  i32("ncheck_lin_indep", m.ncheck_lin_indep);
  ivec("C", m.C);
  ivec("V", m.V);
  ivec("sumX1", m.sumX1);
  ivec("sumX2", m.sumX2);
  ivec("iind", m.iind);
  ivec("jind", m.jind);
  if (m.gen_rows > 0) {  // LDPC_Generator_Systematic::save (:1002-1009); its "Fileversion" is the same 3 and stays in place
    block("G_type", "string", counted(10) + "systematic");
    const uint64_t d[3] = {(uint64_t)m.gen_rows, (uint64_t)m.gen_cols, (uint64_t)(m.gen_cols + 7) / 8};
    std::string g(reinterpret_cast<const char*>(d), 24);
    g.append(reinterpret_cast<const char*>(m.gen_bytes.data()), m.gen_bytes.size());
    block("G", "GF2mat", g);
  }
  ok = (fclose(fp) == 0) && ok;
  if (!ok) err = std::string("save_code: write error on ") + path;
  return ok;
}

// bundle ("LUTLDPC1", oracle/ref_driver.cpp) or the reference's codec file ("IT++"), by magic
bool load_decoder_file(const char* path, DecoderModel& m, std::string& err) {
  FILE* fp = fopen(path, "rb");
  if (!fp) { err = std::string("cannot open ") + path; return false; }
  char magic[4] = {0, 0, 0, 0};
  size_t got = fread(magic, 1, 4, fp);
  fclose(fp);
  if (got == 4 && memcmp(magic, "IT++", 4) == 0) return load_it_codec_file(path, m, err);
  return load_bundle_file(path, m, err);
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

// ------------------------------------------------------------------ node-program compiler
namespace {

struct RawOp {
  int node;
  std::vector<int> in;  // value ids
  int out;              // value id
};

struct Compiler {
  const Tree& t;
  ProgMode mode;
  std::vector<int> val_A;  // label alphabet of every value id
  std::map<std::vector<int>, int> memo;
  std::vector<RawOp> ops;
  int naive = 0;
  std::string err;

  bool is_leaf(const TreeNode& nd) const {
    // var/dec trees: a leaf of either type consumes the next input (src/LUT_Tree.cpp:403-407);
    // check trees: only MSG leaves exist (src/LUT_Tree.cpp:421-425)
    return mode == PROG_CHK ? nd.type == NODE_MSG : (nd.type == NODE_MSG || nd.type == NODE_CHA);
  }
  int eval(int n, const std::vector<int>& seq, size_t& pos) {
    const TreeNode& nd = t.nodes[n];
    if (is_leaf(nd)) {
      if (pos >= seq.size()) { err = "tree consumes more inputs than the node degree provides"; return -1; }
      return seq[pos++];
    }
    if (nd.child.empty()) { err = "inner tree node without children"; return -1; }
    if ((int)nd.child.size() > kMaxOpInputs) { err = "tree node with more than 16 inputs"; return -1; }
    std::vector<int> key;
    key.push_back(n);
    for (int ch : nd.child) {
      int v = eval(ch, seq, pos);
      if (v < 0) return -1;
      key.push_back(v);
    }
    ++naive;
    auto it = memo.find(key);
    if (it != memo.end()) return it->second;
    int id = (int)val_A.size();
    val_A.push_back(nd.K);
    RawOp op;
    op.node = n;
    op.in.assign(key.begin() + 1, key.end());
    op.out = id;
    ops.push_back(op);
    memo[key] = id;
    return id;
  }
};

struct BuiltTable {
  std::vector<uint8_t> bytes;
  std::vector<int> q, sh;
  int idx_bits = 0, q_out = 0;
  bool sym = false;
};

// Re-indexes one node's Q table from the reference's mixed-radix label domain
// (label = sum base_k*child_k, base_{k+1} = base_k*K_k, src/LUT_Tree.cpp:409-417) into the
// storage-code domain (index = concatenated sign/magnitude codes).
bool build_var_table(const Tree& t, int n, const std::vector<int>& A, BuiltTable& bt, std::string& err) {
  const TreeNode& nd = t.nodes[n];
  int nin = (int)A.size(), B = 0;
  bt.q.resize(nin); bt.sh.resize(nin);
  for (int k = 0; k < nin; ++k) { bt.q[k] = code_bits(A[k]); bt.sh[k] = B; B += bt.q[k]; }
  if (B > 17) { err = "LUT node index wider than 17 bits is not supported"; return false; }
  if (nd.K < 2 || nd.K > 16) { err = "LUT node output alphabet out of range"; return false; }
  bt.q_out = code_bits(nd.K);
  int qlen = (int)nd.Q.size();
  size_t full = (size_t)1 << B;
  std::vector<int> T(full, -1);
  for (size_t idx = 0; idx < full; ++idx) {
    long label = 0, base = 1;
    bool valid = true;
    for (int k = 0; k < nin; ++k) {
      int code = (int)((idx >> bt.sh[k]) & ((1u << bt.q[k]) - 1));
      int m = code_to_label(code, A[k]);
      if (m < 0) { valid = false; break; }
      label += base * m;
      base *= t.nodes[nd.child[k]].K;
    }
    if (!valid) continue;
    if (label >= 2L * qlen) {
      err = "a tree leaf is fed a wider alphabet than its LUT was designed for (label outside the table)";
      return false;
    }
    int r = label < qlen ? nd.Q[label] : nd.K - 1 - nd.Q[2 * qlen - 1 - label];  // src/LUT_Tree.cpp:414-417
    if (r < 0 || r >= nd.K) { err = "LUT entry outside the node's output alphabet"; return false; }
    T[idx] = label_to_code(r, nd.K);
  }
  // symmetry: flipping every input's sign flips the output's sign
  size_t F = 0;
  for (int k = 0; k < nin; ++k) F |= (size_t)1 << (bt.sh[k] + bt.q[k] - 1);
  int Hout = 1 << (bt.q_out - 1);
  bool sym = true;
  for (size_t idx = 0; idx < full && sym; ++idx) {
    int a = T[idx], b = T[idx ^ F];
    if ((a < 0) != (b < 0) || (a >= 0 && b != (a ^ Hout))) sym = false;
  }
  bt.sym = sym;
  bt.idx_bits = sym ? B - 1 : B;
  if (bt.idx_bits > 16) { err = "LUT node index wider than 16 bits is not supported"; return false; }
  size_t n_ent = (size_t)1 << bt.idx_bits;
  bt.bytes.resize(n_ent);
  for (size_t idx = 0; idx < n_ent; ++idx) bt.bytes[idx] = (uint8_t)(T[idx] < 0 ? 0 : T[idx]);
  return true;
}

// Check-tree node (src/LUT_Tree.cpp:427-444): index from the magnitudes, entry = value for parity 1.
bool build_chk_table(const Tree& t, int n, const std::vector<int>& A, BuiltTable& bt, std::string& err) {
  const TreeNode& nd = t.nodes[n];
  int nin = (int)A.size(), B = 0;
  bt.q.resize(nin); bt.sh.resize(nin);
  for (int k = 0; k < nin; ++k) { bt.q[k] = code_bits(A[k]); bt.sh[k] = B; B += bt.q[k] - 1; }
  if (B > 16) { err = "check LUT node index wider than 16 bits is not supported"; return false; }
  if (nd.K < 2 || nd.K > 16) { err = "LUT node output alphabet out of range"; return false; }
  bt.q_out = code_bits(nd.K);
  bt.idx_bits = B;
  bt.sym = false;
  int qlen = (int)nd.Q.size();
  size_t n_ent = (size_t)1 << B;
  bt.bytes.assign(n_ent, 0);
  for (size_t idx = 0; idx < n_ent; ++idx) {
    long label = 0, base = 1;
    bool valid = true;
    for (int k = 0; k < nin; ++k) {
      int mag = (int)((idx >> bt.sh[k]) & ((1u << (bt.q[k] - 1)) - 1));
      if (mag >= A[k] / 2) { valid = false; break; }
      label += base * mag;
      base *= t.nodes[nd.child[k]].K / 2;
    }
    if (!valid) continue;
    if (label >= qlen) { err = "check LUT label outside the table"; return false; }
    int r = nd.Q[label];
    if (r < 0 || r >= nd.K) { err = "LUT entry outside the node's output alphabet"; return false; }
    bt.bytes[idx] = (uint8_t)label_to_code(r, nd.K);
  }
  return true;
}

}  // namespace

bool compile_program(const Tree& t, int deg, ProgMode mode, int Nq_in, int Nq_cha, NodeProgram& out,
                     std::string& err) {
  out = NodeProgram();
  Compiler c{t, mode};
  bool has_cha = mode != PROG_CHK;
  int n_in = deg;
  for (int k = 0; k < n_in; ++k) c.val_A.push_back(Nq_in);
  if (has_cha) c.val_A.push_back(Nq_cha);
  int n_inputs = n_in + (has_cha ? 1 : 0);
  int n_out = mode == PROG_DEC ? 1 : deg;
  std::vector<int> outs(n_out);
  for (int ii = 0; ii < n_out; ++ii) {
    std::vector<int> seq;
    for (int k = 0; k < n_in; ++k)
      if (mode == PROG_DEC || k != ii) seq.push_back(k);
    if (has_cha) seq.push_back(n_in);
    size_t pos = 0;
    int v = c.eval(0, seq, pos);
    if (v < 0) { err = c.err; return false; }
    if (pos != seq.size()) { err = "tree leaf count does not match the node degree"; return false; }
    outs[ii] = v;
  }
  // tables (shared between ops of the same node fed the same alphabets)
  std::map<std::vector<int>, std::pair<uint32_t, BuiltTable>> tables;
  out.ops.resize(c.ops.size());
  for (size_t i = 0; i < c.ops.size(); ++i) {
    const RawOp& ro = c.ops[i];
    std::vector<int> key{ro.node}, A;
    for (int v : ro.in) { A.push_back(c.val_A[v]); key.push_back(c.val_A[v]); }
    auto it = tables.find(key);
    if (it == tables.end()) {
      BuiltTable bt;
      bool ok = mode == PROG_CHK ? build_chk_table(t, ro.node, A, bt, err) : build_var_table(t, ro.node, A, bt, err);
      if (!ok) return false;
      size_t off = (out.tables.size() + 127) & ~(size_t)127;  // a 128-entry table spans each bank exactly once
      out.tables.resize(off + bt.bytes.size(), 0);
      memcpy(out.tables.data() + off, bt.bytes.data(), bt.bytes.size());
      it = tables.emplace(key, std::make_pair((uint32_t)off, std::move(bt))).first;
    }
    const BuiltTable& bt = it->second.second;
    DevOp& d = out.ops[i];
    memset(&d, 0, sizeof d);
    d.table_off = it->second.first;
    d.n_in = (uint8_t)ro.in.size();
    d.kind = bt.idx_bits <= 8 ? OP_NARROW : OP_WIDE;
    d.q_out = (uint8_t)bt.q_out;
    d.sym = bt.sym;
    d.chk = mode == PROG_CHK;
    d.idx_bits = (uint8_t)bt.idx_bits;
    for (size_t k = 0; k < ro.in.size(); ++k) { d.q[k] = (uint8_t)bt.q[k]; d.sh[k] = (uint8_t)bt.sh[k]; }
  }
  // slot allocation: inputs own slots [0, n_inputs); outputs stay live to the end
  int n_vals = (int)c.val_A.size();
  std::vector<int> last_use(n_vals, -1), slot(n_vals, -1);
  for (size_t i = 0; i < c.ops.size(); ++i)
    for (int v : c.ops[i].in) last_use[v] = (int)i;
  for (int v : outs) last_use[v] = 1 << 30;
  for (int v = 0; v < n_inputs; ++v) slot[v] = v;
  std::vector<int> free_slots;
  int n_slots = n_inputs;
  for (size_t i = 0; i < c.ops.size(); ++i) {
    const RawOp& ro = c.ops[i];
    // the destination may not alias an input of the same op (the kernel reads all inputs first, but
    // keeping them distinct costs nothing)
    int s;
    if (!free_slots.empty()) { s = free_slots.back(); free_slots.pop_back(); }
    else s = n_slots++;
    slot[ro.out] = s;
    for (size_t k = 0; k < ro.in.size(); ++k) {
      int v = ro.in[k];
      bool dup = false;
      for (size_t j = 0; j < k; ++j) dup |= ro.in[j] == v;
      if (!dup && v >= n_inputs && last_use[v] == (int)i) free_slots.push_back(slot[v]);
    }
    if (last_use[ro.out] < 0) free_slots.push_back(s);  // dead value (cannot happen with memoised eval)
  }
  if (n_slots > 255) { err = "node program needs more than 255 value slots"; return false; }
  for (size_t i = 0; i < c.ops.size(); ++i) {
    out.ops[i].out_slot = (uint8_t)slot[c.ops[i].out];
    for (size_t k = 0; k < c.ops[i].in.size(); ++k) out.ops[i].in_slot[k] = (uint8_t)slot[c.ops[i].in[k]];
  }
  out.out_slot.resize(n_out);
  for (int ii = 0; ii < n_out; ++ii) out.out_slot[ii] = (uint8_t)slot[outs[ii]];
  for (const RawOp& ro : c.ops) {
    out.raw_node.push_back(ro.node);
    out.raw_in.push_back(ro.in);
  }
  out.raw_out = outs;
  out.n_in = n_in;
  out.n_out = n_out;
  out.n_slots = n_slots;
  out.has_cha = has_cha;
  out.n_lookups_naive = c.naive;
  out.q_in = code_bits(Nq_in);
  out.q_out = code_bits(t.nodes[0].K);
  return true;
}

std::vector<uint8_t> NodeProgram::image() const {
  auto al = [](size_t x) { return (x + 127) & ~(size_t)127; };
  DevProgHeader h;
  memset(&h, 0, sizeof h);
  h.n_ops = (uint32_t)ops.size();
  h.n_slots = (uint32_t)n_slots;
  h.n_in = (uint32_t)n_in;
  h.n_out = (uint32_t)n_out;
  h.ops_off = (uint32_t)sizeof(DevProgHeader);
  h.outs_off = (uint32_t)(h.ops_off + ops.size() * sizeof(DevOp));
  h.tables_off = (uint32_t)al(h.outs_off + out_slot.size());
  h.table_bytes = (uint32_t)tables.size();
  h.total_bytes = (uint32_t)al(h.tables_off + tables.size());
  h.has_cha = has_cha;
  h.q_in = (uint32_t)q_in;
  h.q_out = (uint32_t)q_out;
  std::vector<uint8_t> img(h.total_bytes, 0);
  memcpy(img.data(), &h, sizeof h);
  if (!ops.empty()) memcpy(img.data() + h.ops_off, ops.data(), ops.size() * sizeof(DevOp));
  if (!out_slot.empty()) memcpy(img.data() + h.outs_off, out_slot.data(), out_slot.size());
  if (!tables.empty()) memcpy(img.data() + h.tables_off, tables.data(), tables.size());
  return img;
}

}  // namespace lutldpc
