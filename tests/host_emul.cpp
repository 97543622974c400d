// TEST INFRASTRUCTURE ONLY.  Runs the per-thread arithmetic the CUDA kernels execute
// (lut_ldpc_b200/csrc/core.cuh: SWAR min-sum, node-program evaluation) and the host-side tree
// compiler (model.cpp) on the CPU, thread by thread, with the same pass schedule as
// lutldpc_decoder::run_passes.  It lets the CPU test-suite check the packed-nibble algebra and the
// compiled node programs against the oracle without a GPU.  It is never linked into the product
// library and is not a fallback: tests/test_host_emul.py builds it into tests/_build/.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../lut_ldpc_b200/csrc/core.cuh"
#include "../lut_ldpc_b200/csrc/model.hpp"

using namespace lutldpc;

namespace {
struct VecVals {
  std::vector<uint32_t> v;
  uint32_t get(int s) const { return v[s]; }
  void set(int s, uint32_t x) { v[s] = x; }
};
struct Prog {
  NodeProgram p;
  bool ok = false;
};
struct Emul {
  DecoderModel m;
  std::vector<Prog> var_progs, chk_progs;  // [slot][class]
  std::vector<std::vector<int>> vn_lists, cn_lists;
  std::vector<int> vn_deg, cn_deg;
  std::string err;
};

uint32_t lab2code(int m, int Nq) { return (uint32_t)label_to_code(m, Nq); }

void run_node(const NodeProgram& p, VecVals& vals) {
  run_program((uint32_t)p.ops.size(), p.ops.data(), p.tables.data(), vals);
}
}  // namespace

extern "C" {

void* emul_create(const char* bundle, char* err, int errlen) {
  Emul* e = new Emul();
  std::string s;
  auto failed = [&](const std::string& msg) -> void* {
    snprintf(err, errlen, "%s", msg.c_str());
    delete e;
    return nullptr;
  };
  if (!load_decoder_file(bundle, e->m, s)) return failed(s);
  DecoderModel& m = e->m;
  int nvc = m.var_trees.n_deg, ncc = m.minLUT ? 0 : m.chk_trees.n_deg, ns = m.var_trees.n_iter;
  e->vn_lists.resize(nvc); e->cn_lists.resize(ncc);
  for (int v = 0; v < m.nvar; ++v) e->vn_lists[m.var_tree_idx_degree[v]].push_back(v);
  for (int c = 0; c < m.nchk && ncc; ++c) e->cn_lists[m.chk_tree_idx_degree[c]].push_back(c);
  for (int c = 0; c < nvc; ++c) e->vn_deg.push_back(m.var_trees.at(0, c).num_leaves);
  for (int c = 0; c < ncc; ++c) e->cn_deg.push_back(m.chk_trees.at(0, c).num_leaves + 1);
  std::vector<int> first(ns, -1);
  for (int ii = 0; ii < m.max_iters; ++ii)
    if (first[m.var_tree_idx_iter[ii]] < 0) first[m.var_tree_idx_iter[ii]] = ii;
  e->var_progs.resize((size_t)ns * nvc);
  e->chk_progs.resize((size_t)ns * ncc);
  for (int t = 0; t < ns; ++t) {
    for (int c = 0; c < nvc; ++c) {
      if (e->vn_lists[c].empty()) continue;
      Prog& pr = e->var_progs[(size_t)t * nvc + c];
      if (!compile_program(m.var_trees.at(t, c), e->vn_deg[c], t == ns - 1 ? PROG_DEC : PROG_VAR, m.Nq_Msg[first[t]],
                           m.Nq_Cha, pr.p, s))
        return failed("var tree: " + s);
      pr.ok = true;
    }
    for (int c = 0; c < ncc; ++c) {
      if (e->cn_lists[c].empty()) continue;
      Prog& pr = e->chk_progs[(size_t)t * ncc + c];
      if (!compile_program(m.chk_trees.at(t, c), e->cn_deg[c], PROG_CHK, m.Nq_Msg[first[t]], m.Nq_Cha, pr.p, s))
        return failed("chk tree: " + s);
      pr.ok = true;
    }
  }
  return e;
}

void emul_destroy(void* h) { delete (Emul*)h; }

// stats[0] = ops per VN pass (compiled), stats[1] = node evaluations of the reference's recursion
void emul_stats(void* h, int64_t* stats) {
  Emul* e = (Emul*)h;
  int nvc = e->m.var_trees.n_deg;
  stats[0] = stats[1] = 0;
  for (int c = 0; c < nvc; ++c) {
    if (!e->var_progs[c].ok) continue;
    stats[0] += (int64_t)e->var_progs[c].p.ops.size() * e->vn_lists[c].size();
    stats[1] += (int64_t)e->var_progs[c].p.n_lookups_naive * e->vn_lists[c].size();
  }
}

// The guarded min-sum variant (vectorised CN kernel) against the plain one on random code words.
// Returns the number of mismatches.
int emul_cng_selftest(int trials, uint32_t seed) {
  auto rnd = [&]() { seed = seed * 1664525u + 1013904223u; return seed; };
  int bad = 0;
  for (int t = 0; t < trials; ++t) {
    int q = 1 + (int)(rnd() >> 8) % 4, dc = 2 + (int)(rnd() >> 8) % 14;
    QMask qm = make_qmask(q);
    uint32_t keep = (kNibOnes << q) - kNibOnes;  // only q-bit codes occur
    uint32_t w[16];
    for (int k = 0; k < dc; ++k) {
      uint32_t a = rnd(), b = rnd();
      w[k] = ((a << 16) ^ (b >> 4) ^ (t % 5 == 0 ? 0 : a)) & keep;
      if (t % 7 == 0) w[k] &= w[0];  // provoke ties
    }
    CnAcc a; CnAccG g;
    cn_init(a, w[0], w[1], qm); cng_init(g, w[0], w[1], qm);
    for (int k = 2; k < dc; ++k) { cn_accum(a, w[k], qm); cng_accum(g, w[k], qm); }
    for (int k = 0; k < dc; ++k) bad += cn_emit(a, w[k], qm) != cng_emit(g, w[k], qm);
    bad += ((a.par ^ g.par) & qm.H) != 0;
  }
  return bad;
}

// Same contract as lutldpc_decode_quantized_traced.
int emul_decode(void* h, int F, const uint8_t* cha_lab, const uint8_t* msg_lab, int psc, int pisc, uint8_t* out_bits,
                int32_t* iters, int trace_frame, uint8_t* tmsgs, int32_t* tkind, int32_t* titer, int max_rec,
                int32_t* n_rec_out) {
  Emul* e = (Emul*)h;
  const DecoderModel& m = e->m;
  const int N = m.nvar, E = m.num_edges, W = (F + 7) / 8, max_iters = m.max_iters;
  const int nvc = m.var_trees.n_deg, ncc = m.minLUT ? 0 : m.chk_trees.n_deg;
  std::vector<uint32_t> msgs((size_t)E * W), cha((size_t)N * W), done(W), vn_fail(W, 0), cn_fail(W, 0), syn_fail(W, 0);
  std::vector<uint8_t> hard((size_t)N * W);
  const int qc = code_bits(m.Nq_Cha);
  for (int w = 0; w < W; ++w) {
    uint32_t d = 0;
    for (int j = 0; j < 8; ++j) {
      if (8 * w + j >= F) d |= 1u << (4 * j);
      else iters[8 * w + j] = 0;
    }
    done[w] = d;
  }
  for (int v = 0; v < N; ++v)
    for (int w = 0; w < W; ++w) {
      uint32_t cw = 0, mw = 0;
      for (int j = 0; j < 8; ++j) {
        int f = 8 * w + j;
        if (f >= F) break;
        cw |= lab2code(cha_lab[(size_t)f * N + v], m.Nq_Cha) << (4 * j);
        mw |= lab2code(msg_lab[(size_t)f * N + v], m.Nq_Msg[0]) << (4 * j);
      }
      cha[(size_t)v * W + w] = cw;
      hard[(size_t)v * W + w] = (uint8_t)nib_bits_to_byte(cw, qc - 1);
      for (int ed = m.vn_edge0[v]; ed < m.vn_edge0[v + 1]; ++ed) msgs[(size_t)ed * W + w] = mw;
    }
  int n_rec = 0;
  bool live = trace_frame >= 0;
  auto record = [&](int kind, int it, int Nq) {
    if (!live || n_rec >= max_rec) return;
    int q = code_bits(Nq);
    (void)q;
    for (int ed = 0; ed < E; ++ed) {
      uint32_t c = (msgs[(size_t)ed * W + (trace_frame >> 3)] >> (4 * (trace_frame & 7))) & 0xF;
      tmsgs[(size_t)n_rec * E + ed] = (uint8_t)code_to_label((int)c, Nq);
    }
    tkind[n_rec] = kind; titer[n_rec] = it; ++n_rec;
  };
  auto frame_done = [&](int f) { return (done[f >> 3] >> (4 * (f & 7))) & 1; };
  auto syndrome = [&]() {
    for (int c = 0; c < m.nchk; ++c)
      for (int w = 0; w < W; ++w) {
        if (done[w] == kNibOnes) continue;
        uint32_t x = 0;
        for (int k = m.cn_start[c]; k < m.cn_start[c + 1]; ++k) x ^= hard[(size_t)m.chk_equ_idx[k] * W + w];
        if (x) syn_fail[w] |= byte_to_nib_bits(x);
      }
  };
  auto finalize = [&](int mode, int value) {
    uint32_t active = 0;
    for (int w = 0; w < W; ++w) {
      uint32_t d = done[w], conv;
      if (mode == 0) { conv = ~d & ~syn_fail[w] & kNibOnes; syn_fail[w] = 0; }
      else if (mode == 1) { conv = ~d & ~(vn_fail[w] | cn_fail[w]) & kNibOnes; vn_fail[w] = cn_fail[w] = 0; }
      else conv = ~d & kNibOnes;
      uint32_t sf = mode == 2 ? syn_fail[w] : 0;
      for (int j = 0; j < 8; ++j)
        if ((conv >> (4 * j)) & 1) iters[8 * w + j] = ((sf >> (4 * j)) & 1) ? -value : value;
      if (mode == 2) syn_fail[w] = 0;
      d |= conv;
      done[w] = d;
      active += __builtin_popcount(~d & kNibOnes);
    }
    return active;
  };
  auto node_pass = [&](int mode, const NodeProgram& p, const std::vector<int>& list, int want_flags) {
    VecVals vals;
    vals.v.resize(p.n_slots);
    const int so = p.q_out - 1, si = p.q_in - 1;
    for (int node : list)
      for (int w = 0; w < W; ++w) {
        uint32_t dn = done[w];
        if (dn == kNibOnes) continue;
        uint32_t done_byte = nib_bits_to_byte(dn, 0), par = 0;
        int s = mode == 2 ? m.cn_start[node] : m.vn_edge0[node];
        for (int k = 0; k < p.n_in; ++k) {
          size_t row = mode == 2 ? (size_t)m.cn_msg_idx[s + k] : (size_t)(s + k);
          uint32_t x = msgs[row * W + w];
          vals.set(k, x);
          par ^= x;
        }
        if (mode != 2) vals.set(p.n_in, cha[(size_t)node * W + w]);
        run_node(p, vals);
        if (mode == 0) {
          uint32_t A = 0xFFFFFFFFu, O = 0, first = 0;
          for (int ii = 0; ii < p.n_out; ++ii) {
            uint32_t r = vals.get(p.out_slot[ii]);
            msgs[(size_t)(s + ii) * W + w] = r;
            A &= r; O |= r;
            if (ii == 0) first = r;
          }
          if (want_flags) {
            vn_fail[w] |= ((A ^ O) >> so) & kNibOnes;
            uint8_t& hp = hard[(size_t)node * W + w];
            hp = (uint8_t)((hp & done_byte) | (nib_bits_to_byte(first, so) & ~done_byte));
          }
        } else if (mode == 2) {
          for (int ii = 0; ii < p.n_out; ++ii) msgs[(size_t)m.cn_msg_idx[s + ii] * W + w] = vals.get(p.out_slot[ii]);
          if (want_flags) cn_fail[w] |= (par >> si) & kNibOnes;
        } else {
          uint8_t& hp = hard[(size_t)node * W + w];
          hp = (uint8_t)((hp & done_byte) | (nib_bits_to_byte(vals.get(p.out_slot[0]), so) & ~done_byte));
        }
      }
  };
  uint32_t active = (uint32_t)F;
  if (pisc) {
    syndrome();
    active = finalize(0, 0);
    if (live && frame_done(trace_frame)) live = false;
  }
  record(0, -1, m.Nq_Msg[0]);
  for (int ii = 0; ii < max_iters && active; ++ii) {
    const int t = m.var_tree_idx_iter[ii];
    const int want = psc && ii > 0;
    if (m.minLUT) {
      QMask qm = make_qmask(code_bits(m.Nq_Msg[ii]));
      for (int c = 0; c < m.nchk; ++c)
        for (int w = 0; w < W; ++w) {
          if (done[w] == kNibOnes) continue;
          int s = m.cn_start[c], dc = m.cn_start[c + 1] - s;
          CnAcc a;
          cn_init(a, msgs[(size_t)m.cn_msg_idx[s] * W + w], msgs[(size_t)m.cn_msg_idx[s + 1] * W + w], qm);
          for (int k = 2; k < dc; ++k) cn_accum(a, msgs[(size_t)m.cn_msg_idx[s + k] * W + w], qm);
          for (int k = 0; k < dc; ++k) {
            size_t o = (size_t)m.cn_msg_idx[s + k] * W + w;
            msgs[o] = cn_emit(a, msgs[o], qm);
          }
          if (want) cn_fail[w] |= (a.par >> qm.s) & kNibOnes;
        }
    } else {
      for (int c = 0; c < ncc; ++c)
        if (e->chk_progs[(size_t)t * ncc + c].ok) node_pass(2, e->chk_progs[(size_t)t * ncc + c].p, e->cn_lists[c], want);
    }
    record(1, ii, m.Nq_Msg[ii]);
    if (want) {
      active = finalize(1, ii);
      if (live && frame_done(trace_frame)) { n_rec = n_rec >= 2 ? n_rec - 2 : 0; live = false; }
      if (!active) break;
    }
    if (ii != max_iters - 1) {
      for (int c = 0; c < nvc; ++c)
        if (e->var_progs[(size_t)t * nvc + c].ok) node_pass(0, e->var_progs[(size_t)t * nvc + c].p, e->vn_lists[c], psc);
      record(2, ii, m.Nq_Msg[ii + 1]);
    } else {
      record(2, ii, m.Nq_Msg[ii]);
    }
  }
  if (active) {
    const int t = m.var_tree_idx_iter[max_iters - 1];
    for (int c = 0; c < nvc; ++c)
      if (e->var_progs[(size_t)t * nvc + c].ok) node_pass(1, e->var_progs[(size_t)t * nvc + c].p, e->vn_lists[c], 0);
    syndrome();
    finalize(2, max_iters);
  }
  for (int f = 0; f < F; ++f)
    for (int v = 0; v < N; ++v) out_bits[(size_t)f * N + v] = (hard[(size_t)v * W + (f >> 3)] >> (f & 7)) & 1;
  if (n_rec_out) *n_rec_out = n_rec;
  return 0;
}
}
