// LUT design on the host: the step of LDPC_BER_Sim_LUT::load that turns (code, design sigma, resolutions) into the
// quantiser thresholds and the LUT trees of an LDPC_Code_LUT.  Host-only C++, double precision.
//
// What this restates (paths relative to the reference tree):
//   LDPC_Code_LUT::design_luts                 src/LDPC_Code_LUT.cpp:699-746
//   get_lut_tree_templates                     src/LDPC_DE.cpp:1124-1290     (auto_bin_balanced and file templates)
//   LDPC_DE_LUT::evolve(save_luts = true)      src/LDPC_DE.cpp:198-326
//   set_channel_pmf / get_quant_bound          src/LDPC_DE.cpp:400-412, 561-601
//   chk_update_irr / var_update_irr            src/LDPC_DE.cpp:414-558       (min-sum or LUT checks, joint_root strategy)
//   joint_root_irr_lut_design / level_lut_tree_update   src/LDPC_DE.cpp:1345-1466
//   chk_update_minsum(pmf)                     src/LDPC_DE.cpp:1061-1089
//   LUT_Tree::update / var_update / chk_update src/LUT_Tree.cpp:114-129, 686-766
//   quant_mi_sym, sym_llr_sort_unique, product pmfs, get_gaussian_pmf   src/common.cpp:30-70, 140-149, 193-369
//   get_empirical_ensemble                     src/LDPC_Ensemble.cpp:391-423
//   tree text format                           src/LUT_Tree.cpp:488-517, 847-864
//
// The tables must equal the reference's entry for entry (the decoder is bit-exact only with identical LUTs), and the
// information-optimal quantiser breaks ties between candidate boundaries by floating-point comparison: every sum,
// product and division below is therefore done in the reference's order (IT++'s sum() adds left to right, vec / scalar
// divides element by element, ...).  tests/test_design_host.py compares the result with the reference-written golden
// bundles byte for byte.
#include "design.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <limits>
#include <map>
#include <sstream>

namespace lutldpc {
namespace design {

typedef std::vector<double> vec;
typedef std::vector<int> ivec;

namespace {

double vsum(const vec& v) {  // itpp::sum: left to right
  double m = 0;
  for (double x : v) m += x;
  return m;
}
vec vdiv(const vec& v, double t) {  // itpp: v / t, element by element
  vec r(v.size());
  for (size_t i = 0; i < v.size(); ++i) r[i] = v[i] / t;
  return r;
}
vec flip(const vec& x) {
  vec y(x.size());
  for (size_t i = 0; i < x.size(); ++i) y[i] = x[x.size() - 1 - i];
  return y;
}
vec kron(const vec& x, const vec& y) {  // src/common.cpp:180-190
  vec z(x.size() * y.size());
  for (size_t a = 0; a < x.size(); ++a)
    for (size_t b = 0; b < y.size(); ++b) z[a * y.size() + b] = x[a] * y[b];
  return z;
}
double qfunc(double x) { return 0.5 * ::erfc(x / 1.41421356237310); }  // itpp/itpp/base/math/error.cpp:217-220

// src/common.cpp:140-149
vec gaussian_pmf(double mu, double sig, int N, double delta) {
  vec pmf(N);
  pmf[0] = 1 - qfunc(((-N / 2.0 + 1) * delta - mu) / sig);
  for (int nn = 1; nn < N - 1; ++nn)
    pmf[nn] = qfunc(((nn - N / 2.0) * delta - mu) / sig) - qfunc(((nn + 1 - N / 2.0) * delta - mu) / sig);
  pmf[N - 1] = qfunc(((N / 2.0 - 1) * delta - mu) / sig);
  return vdiv(pmf, vsum(pmf));
}

double x_log2_y(double x, double y) { return x == 0 ? 0.0 : x * std::log2(y); }  // src/common.cpp:162-167

// src/common.cpp:333-369: sort the labels by LLR, merge exactly equal LLRs (llr_delta = 0)
vec sym_llr_sort_unique(const vec& p_in, ivec& idx_in, ivec& idx_sorted) {
  const int M_in = (int)p_in.size();
  vec llr(M_in);
  for (int i = 0; i < M_in; ++i) llr[i] = std::log(p_in[i]) - std::log(p_in[M_in - 1 - i]);
  // ascending LLR, equal LLRs by ascending index (the reference sorts, then re-sorts every run of equal values)
  idx_in.resize(M_in);
  for (int i = 0; i < M_in; ++i) idx_in[i] = i;
  std::sort(idx_in.begin(), idx_in.end(), [&](int a, int b) { return llr[a] < llr[b] || (llr[a] == llr[b] && a < b); });
  idx_sorted.assign(M_in / 2, 0);
  double dupl = llr[idx_in[0]];
  int dupl_idx = 0, num_dupl = 0;
  for (int mm = 1; mm < M_in / 2; ++mm) {
    if (std::abs(llr[idx_in[mm]] - dupl) <= 0.0) ++num_dupl;
    else ++dupl_idx;
    idx_sorted[mm] = dupl_idx;
    dupl = llr[idx_in[mm]];
  }
  const int mx = *std::max_element(idx_sorted.begin(), idx_sorted.end());
  ivec second(M_in / 2);
  for (int mm = 0; mm < M_in / 2; ++mm) second[mm] = 2 * mx + 1 - idx_sorted[M_in / 2 - 1 - mm];
  idx_sorted.insert(idx_sorted.end(), second.begin(), second.end());
  const int M = M_in - 2 * num_dupl;
  vec p_sorted(M, 0.0);
  for (int mm = 0; mm < M_in; ++mm) p_sorted[idx_sorted[mm]] += p_in[idx_in[mm]];
  return p_sorted;
}

// src/common.cpp:230-331: information-optimal symmetric quantiser of a binary-input pmf to Nq labels (dynamic program)
bool quant_mi_sym(vec& p_out, ivec& Q_out, const vec& p_in, int Nq, bool sorted, std::string& err) {
  const int K = Nq, M_in = (int)p_in.size();
  if (M_in % 2 || K % 2 || M_in == 0) { err = "quant_mi_sym: input pmf and label count must be even"; return false; }
  vec p_sorted;
  int M;
  ivec idx_in, idx_sorted;
  if (!sorted) {
    p_sorted = sym_llr_sort_unique(p_in, idx_in, idx_sorted);
    M = (int)p_sorted.size();
  } else {
    idx_in.resize(M_in);
    for (int i = 0; i < M_in; ++i) idx_in[i] = i;
    idx_sorted = idx_in;
    p_sorted = p_in;
    M = M_in;
  }
  if (K >= M) {  // trivial case
    Q_out.assign(M_in, 0);
    int outlabel = 0;
    for (int mm = 0; mm < M_in / 2; ++mm) {
      if (idx_sorted[mm] > outlabel) ++outlabel;
      Q_out[idx_in[M_in - 1 - mm]] = K - 1 - outlabel;
      Q_out[idx_in[mm]] = outlabel;
    }
    p_out.assign(K, 0.0);
    for (int mm = 0; mm < M_in; ++mm) p_out[Q_out[mm]] += p_in[mm];
    return true;
  }
  const int H = M / 2, Kh = K / 2;
  // partial mutual information of merging the sorted labels ap..a (and their mirror images) into one output label
  std::vector<double> g((size_t)H * H, 0.0);
  for (int ap = 0; ap < H; ++ap) {
    double p_plus = 0, p_minus = 0;
    for (int a = ap; a < H; ++a) {
      p_plus += p_sorted[H + a];
      p_minus += p_sorted[H - 1 - a];
      g[(size_t)ap * H + a] = x_log2_y(p_plus, 2 * p_plus / (p_plus + p_minus)) + x_log2_y(p_minus, 2 * p_minus / (p_plus + p_minus));
    }
  }
  std::vector<double> S((size_t)H * Kh, 0.0);
  std::vector<int> h((size_t)H * Kh, 0);
  for (int a = 0; a < (M - K) / 2 + 1; ++a) S[(size_t)a * Kh] = g[a];
  for (int zz = 1; zz < Kh; ++zz)
    for (int a = zz; a < zz + (M - K) / 2 + 1; ++a) {
      double best = -std::numeric_limits<double>::max();
      int arg = 0;
      for (int ap = zz; ap <= a; ++ap) {
        const double t = S[(size_t)(ap - 1) * Kh + zz - 1] + g[(size_t)ap * H + a];
        if (t > best) { best = t; arg = ap; }
      }
      S[(size_t)a * Kh + zz] = best;
      h[(size_t)a * Kh + zz] = arg;
    }
  ivec astar(Kh + 1, 0);
  astar[Kh] = H;
  for (int kk = Kh - 1; kk > 0; --kk) astar[kk] = h[(size_t)(astar[kk + 1] - 1) * Kh + kk];
  int outlabel = 0;
  Q_out.assign(M_in, 0);
  for (int mm = 0; mm < M_in / 2; ++mm) {
    if (idx_sorted[mm + M_in / 2] - H >= astar[outlabel + 1]) ++outlabel;
    Q_out[idx_in[M_in / 2 + mm]] = Kh + outlabel;
    Q_out[idx_in[M_in / 2 - 1 - mm]] = Kh - 1 - outlabel;
  }
  p_out.assign(K, 0.0);
  for (int mm = 0; mm < M_in; ++mm) p_out[Q_out[mm]] += p_in[mm];
  return true;
}

// src/common.cpp:30-39
vec var_product_pmf(const std::vector<const vec*>& p_in) {
  vec p = *p_in.back();
  for (int ii = (int)p_in.size() - 2; ii >= 0; --ii) p = kron(p, *p_in[ii]);
  return p;
}
// src/common.cpp:193-228
int signed_to_unsigned_idx(int idx, const ivec& inres) {
  const int n = (int)inres.size();
  int out_max = 2;
  for (int r : inres) out_max *= r / 2;
  int parity = 0, idx_out = 0, base = 1, t = idx;
  for (int jj = 0; jj < n; ++jj) {
    const int in = t % inres[jj];
    t /= inres[jj];
    if (in < inres[jj] / 2) {
      parity ^= 1;
      idx_out += base * (inres[jj] / 2 - 1 - in);
    } else {
      idx_out += base * (in - inres[jj] / 2);
    }
    base *= inres[jj] / 2;
  }
  return parity == 0 ? out_max - 1 - idx_out : idx_out;
}
// src/common.cpp:41-70
vec chk_product_pmf(const std::vector<const vec*>& p_in) {
  const int n = (int)p_in.size();
  ivec res(n);
  for (int j = 0; j < n; ++j) res[j] = (int)p_in[j]->size();
  vec p0 = *p_in[n - 1], p1 = flip(*p_in[n - 1]);
  for (int ii = n - 2; ii >= 0; --ii) {
    const vec f = flip(*p_in[ii]);
    const vec a = kron(p0, *p_in[ii]), b = kron(p1, f), c = kron(p1, *p_in[ii]), d = kron(p0, f);
    vec n0(a.size()), n1(a.size());
    for (size_t i = 0; i < a.size(); ++i) {
      n0[i] = .5 * (a[i] + b[i]);
      n1[i] = .5 * (c[i] + d[i]);
    }
    p0 = n0;
    p1 = n1;
  }
  int out = 2;
  for (int r : res) out *= r / 2;
  vec comb(out, 0.0);
  for (int mm = 0; mm < (int)p0.size(); ++mm) comb[signed_to_unsigned_idx(mm, res)] += p0[mm];
  return comb;
}

// src/LDPC_DE.cpp:1061-1123 (pmf_plus / pmf_minus / pmf_join included)
vec chk_update_minsum(const vec& p_in, int dc) {
  const int N = (int)p_in.size(), Hn = N / 2;
  vec a_plus(Hn), a_minus(Hn);
  for (int nn = 0; nn < Hn; ++nn) {
    a_plus[nn] = p_in[Hn + nn] + p_in[Hn - 1 - nn];
    a_minus[nn] = p_in[Hn + nn] - p_in[Hn - 1 - nn];
  }
  const vec b_plus = a_plus, b_minus = a_minus;
  vec c_plus = a_plus, c_minus = a_minus;  // dc = 2: the reference reads unset memory; degree-2 checks pass the pmf through
  for (int dd = 1; dd < dc - 1; ++dd) {
    c_plus.assign(Hn, 0.0);
    c_minus.assign(Hn, 0.0);
    for (int ii = 0; ii < Hn; ++ii)
      for (int jj = 0; jj < Hn; ++jj) {
        const int kk = std::min(ii, jj);
        c_plus[kk] += a_plus[ii] * b_plus[jj];
        c_minus[kk] += a_minus[ii] * b_minus[jj];
      }
    a_plus = c_plus;
    a_minus = c_minus;
  }
  vec out(N);
  for (int nn = 0; nn < Hn; ++nn) {
    out[Hn + nn] = .5 * (c_plus[nn] + c_minus[nn]);
    out[Hn - 1 - nn] = .5 * (c_plus[nn] - c_minus[nn]);
  }
  return out;
}

// ------------------------------------------------------------------------------------------------ trees
enum { T_IM = 0, T_ROOT = 1, T_MSG = 2, T_CHA = 3 };        // LUT_Tree_Node::node_type_t, src/LUT_Tree.hpp:188-194
enum { VARTREE = 0, CHKTREE = 1, DECTREE = 2 };             // LUT_Tree::tree_type_t, src/LUT_Tree.hpp:50-55

struct DNode {
  int type = T_MSG, K = 0;
  ivec Q;
  vec p;
  std::vector<int> child;
};
struct DTree {
  int tree_type = VARTREE, num_leaves = 0;
  std::vector<DNode> n;  // n[0] is the root
  int add(int type) { n.emplace_back(); n.back().type = type; return (int)n.size() - 1; }
};

int count_leaves(const DTree& t, int i) {
  if (t.n[i].type == T_MSG || t.n[i].type == T_CHA) return 1;
  int c = 0;
  for (int ch : t.n[i].child) c += count_leaves(t, ch);
  return c;
}

// LUT_Tree_Node::gen_bin_balanced_tree, src/LUT_Tree.cpp:200-237: the message leaves merged pairwise through a queue,
// then ROOT[subtree, CHA] (var) or the last merged node as root (check trees)
DTree gen_bin_balanced(int num_leaves, bool var, int tree_type) {
  DTree t;
  t.tree_type = tree_type;
  t.num_leaves = num_leaves;
  const int root = t.add(T_ROOT);
  std::vector<int> q;
  for (int l = 0; l < num_leaves - (var ? 1 : 0); ++l) q.push_back(t.add(T_MSG));
  size_t head = 0;
  while (q.size() - head > 1) {
    const int l = q[head++], r = q[head++];
    const int im = t.add(T_IM);
    t.n[im].child = {l, r};
    q.push_back(im);
  }
  if (var) {
    const int cha = t.add(T_CHA);
    t.n[root].child = {q[head], cha};
  } else {  // the last node becomes the root
    t.n[root].child = t.n[q[head]].child;
    t.n[q[head]].child.clear();  // (left unused in the node array)
  }
  return t;
}

// LUT_Tree_Node::parse, src/LUT_Tree.cpp:167-198: pre-order over r i m c, '/' closes a child list, EOF closes everything
int parse_template(const std::string& s, size_t& pos, DTree& t, std::string& err) {
  if (pos >= s.size()) return -1;
  const char c = s[pos++];
  int type;
  switch (c) {
    case '/': return -1;
    case 'r': type = T_ROOT; break;
    case 'i': type = T_IM; break;
    case 'm': type = T_MSG; break;
    case 'c': type = T_CHA; break;
    default: err = std::string("tree template: unexpected character '") + c + "'"; return -2;
  }
  const int me = t.add(type);
  for (;;) {
    const int ch = parse_template(s, pos, t, err);
    if (ch == -2) return -2;
    if (ch < 0) break;
    t.n[me].child.push_back(ch);
  }
  return me;
}

void set_resolution(DTree& t, int Nq_in, int Nq_out, int Nq_cha) {  // src/LUT_Tree.cpp:296-306
  for (DNode& nd : t.n) nd.K = nd.type == T_ROOT ? Nq_out : nd.type == T_CHA ? Nq_cha : Nq_in;
}
void set_leaves(DTree& t, const vec& p_msg, const vec& p_cha) {  // :92-103
  for (DNode& nd : t.n) {
    if (nd.type == T_MSG) nd.p = p_msg;
    else if (nd.type == T_CHA) nd.p = p_cha;
  }
}

vec input_product_pmf(const DTree& t, int i) {  // LUT_Tree_Node::get_input_product_pmf, :536-560
  std::vector<const vec*> in;
  for (int ch : t.n[i].child) in.push_back(&t.n[ch].p);
  return t.tree_type == CHKTREE ? chk_product_pmf(in) : var_product_pmf(in);
}

// the quantiser of p_prod to Nq labels with zero-mass entries set aside (LUT_Tree::var_update, :709-742)
bool design_var_node(DNode& nd, const vec& p_prod, std::string& err) {
  const int M = (int)p_prod.size(), Nq = nd.K;
  std::vector<char> nz(M);
  vec sub;
  for (int mm = 0; mm < M; ++mm) {
    nz[mm] = (.5 * (p_prod[mm] + p_prod[M - 1 - mm])) != 0;
    if (nz[mm]) sub.push_back(p_prod[mm]);
  }
  ivec Qnz;
  vec p_out;
  if (!quant_mi_sym(p_out, Qnz, sub, Nq, false, err)) return false;
  ivec Q(M);
  for (int mm = 0; mm < M; ++mm) Q[mm] = mm < M / 2 ? Nq / 2 - 1 : Nq / 2;
  for (int mm = 0, k = 0; mm < M; ++mm)
    if (nz[mm]) Q[mm] = Qnz[k++];
  nd.Q.assign(Q.begin(), Q.begin() + M / 2);
  nd.p = vdiv(p_out, vsum(p_out));
  return true;
}
void reuse_node(DNode& nd, const vec& p_prod) {  // the reuse branch of var_update / chk_update
  const int M = (int)p_prod.size(), Nq = nd.K;
  vec p_out(Nq, 0.0);
  for (int mm = 0; mm < M; ++mm) {
    if (mm < M / 2) p_out[nd.Q[mm]] += p_prod[mm];
    else p_out[Nq - 1 - nd.Q[M - 1 - mm]] += p_prod[mm];
  }
  nd.p = vdiv(p_out, vsum(p_out));
}

// LUT_Tree_Node::tree_update, :114-129: children first, then this node (designs or re-applies its quantiser)
bool tree_update(DTree& t, int i, bool reuse, std::string& err) {
  DNode& nd = t.n[i];
  if (nd.type == T_MSG || nd.type == T_CHA) return true;
  for (int ch : nd.child)
    if (!tree_update(t, ch, reuse, err)) return false;
  const vec prod = input_product_pmf(t, i);
  if (reuse) { reuse_node(t.n[i], prod); return true; }
  if (t.tree_type == CHKTREE) {  // LUT_Tree::chk_update, :744-766
    vec p_out;
    ivec Q;
    if (!quant_mi_sym(p_out, Q, prod, t.n[i].K, false, err)) return false;
    t.n[i].Q.assign(Q.begin(), Q.begin() + Q.size() / 2);
    t.n[i].p = vdiv(p_out, vsum(p_out));
    return true;
  }
  return design_var_node(t.n[i], prod, err);
}

// level_lut_tree_update for the roots (joint_root_irr_lut_design, src/LDPC_DE.cpp:1345-1466): ONE quantiser for the
// root inputs of all degrees, weighted by the edge-degree distribution
bool joint_root_update(std::vector<DTree>& trees, const vec& degree_dist, std::string& err) {
  const int L = (int)trees.size();
  std::vector<vec> prod(L);
  int M_tot = 0;
  const int Nout = trees[0].n[0].K;
  for (int ll = 0; ll < L; ++ll) {
    prod[ll] = input_product_pmf(trees[ll], 0);
    M_tot += (int)prod[ll].size();
  }
  vec overall(M_tot, -1e9);
  int I = 0;
  for (int ll = 0; ll < L; ++ll) {
    // node_weights = leaves of the node / their sum over the level's nodes of this tree: one root, weight x / x
    const double nl = (double)count_leaves(trees[ll], 0), w = nl / nl;
    const int M = (int)prod[ll].size();
    for (int mm = 0; mm < M / 2; ++mm) {
      overall[I + mm] = w * degree_dist[ll] * prod[ll][mm];
      overall[M_tot - 1 - I - mm] = w * degree_dist[ll] * prod[ll][M - 1 - mm];
    }
    I += M / 2;
  }
  overall = vdiv(overall, vsum(overall));
  std::vector<char> nz(M_tot);
  vec sub;
  for (int mm = 0; mm < M_tot; ++mm) {
    nz[mm] = (.5 * (overall[mm] + overall[M_tot - 1 - mm])) != 0;
    if (nz[mm]) sub.push_back(overall[mm]);
  }
  ivec Qnz;
  vec p_out;
  if (!quant_mi_sym(p_out, Qnz, sub, Nout, false, err)) return false;
  ivec Q(M_tot);
  for (int mm = 0; mm < M_tot; ++mm) Q[mm] = mm < M_tot / 2 ? Nout / 2 - 1 : Nout / 2;
  for (int mm = 0, k = 0; mm < M_tot; ++mm)
    if (nz[mm]) Q[mm] = Qnz[k++];
  I = 0;
  for (int ll = 0; ll < L; ++ll) {
    const int M = (int)prod[ll].size();
    DNode& root = trees[ll].n[0];
    root.Q.assign(Q.begin() + I, Q.begin() + I + M / 2);
    I += M / 2;
    root.p.assign(Nout, 0.0);
    for (int mm = 0; mm < M; ++mm) {
      if (mm < M / 2) root.p[root.Q[mm]] += prod[ll][mm];
      else root.p[Nout - 1 - root.Q[M - 1 - mm]] += prod[ll][mm];
    }
  }
  return true;
}

// one irregular node update of evolve(): trees from the templates, individual design of every node, joint design of
// the roots, then the output pmfs with the final tables (var_update_irr / chk_update_irr, JOINT_ROOT)
bool irr_update(std::vector<DTree>& trees, const std::vector<DTree>& templ, bool reuse, const vec& p_msg, const vec& p_cha,
                int Nq_in, int Nq_out, int Nq_cha, const vec& dist, vec& p_avg, std::string& err) {
  const int L = (int)templ.size();
  if (!reuse) {
    trees = templ;
    for (DTree& t : trees) {
      set_leaves(t, p_msg, p_cha);
      set_resolution(t, Nq_in, Nq_out, Nq_cha);
    }
    for (DTree& t : trees)
      if (!tree_update(t, 0, false, err)) return false;
    if (!joint_root_update(trees, dist, err)) return false;
  } else {
    for (DTree& t : trees) set_leaves(t, p_msg, p_cha);
  }
  p_avg.assign(Nq_out, 0.0);
  for (int dd = 0; dd < L; ++dd) {
    if (!tree_update(trees[dd], 0, true, err)) return false;
    const vec& pt = trees[dd].n[0].p;
    for (int k = 0; k < Nq_out; ++k) p_avg[k] = p_avg[k] + dist[dd] * pt[k];
  }
  return true;
}

// text serialisation, src/LUT_Tree.cpp:488-517, 847-864
void serialize_node(const DTree& t, int i, std::ostringstream& os) {
  const DNode& nd = t.n[i];
  os << nd.child.size() << "\n" << nd.type << " " << nd.Q.size() << " " << nd.K << "\n";
  if (!nd.Q.empty()) {
    for (size_t k = 0; k + 1 < nd.Q.size(); ++k) os << nd.Q[k] << " ";
    os << nd.Q.back() << "\n";
  }
  for (int ch : nd.child) serialize_node(t, ch, os);
}
std::string serialize(const std::vector<std::vector<DTree>>& a) {
  std::ostringstream os;
  os << a.size() << "\n";
  for (const auto& slot : a) {
    os << slot.size() << "\n";
    for (const DTree& t : slot) {
      os << t.tree_type << " " << t.num_leaves << "\n";
      serialize_node(t, 0, os);
    }
  }
  return os.str();
}

// a minimal reader for the tree-template ini (sections var_iter_NNN / chk_iter_NNN / DT, keys var_deg_DDD / chk_deg_DDD)
bool read_ini(const std::string& path, std::map<std::string, std::map<std::string, std::string>>& ini, std::string& err) {
  std::ifstream f(path.c_str());
  if (!f.is_open()) { err = "cannot open tree template file " + path; return false; }
  auto trim = [](std::string s) {
    const size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
  };
  std::string line, sec;
  while (std::getline(f, line)) {
    line = trim(line);
    if (line.empty() || line[0] == ';' || line[0] == '#') continue;
    if (line[0] == '[') { sec = trim(line.substr(1, line.find(']') - 1)); continue; }
    const size_t eq = line.find('=');
    if (eq == std::string::npos) continue;
    ini[sec][trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
  }
  return true;
}
std::string key3(const char* prefix, int v) {
  char b[64];
  snprintf(b, sizeof b, "%s%03d", prefix, v);
  return b;
}

}  // namespace

bool design_luts(const Params& P, Result& R, std::string& err) {
  const int max_iters = P.max_iters;
  if (max_iters < 2 || (int)P.Nq_Msg.size() != max_iters) { err = "design: need max_iters >= 2 and Nq_Msg of that length"; return false; }
  if (P.Nq_Cha < 2 || P.Nq_Cha > 16 || (P.Nq_Cha & 1)) { err = "design: Nq_Cha must be even and <= 16"; return false; }
  for (int q : P.Nq_Msg) if (q < 2 || q > 16 || (q & 1)) { err = "design: Nq_Msg must be even and <= 16"; return false; }
  ivec reuse(max_iters, 0);
  if (!P.reuse_vec.empty()) {
    if ((int)P.reuse_vec.size() != max_iters) { err = "design: reuse_vec length"; return false; }
    reuse = P.reuse_vec;
  }
  if (reuse[0] || reuse[max_iters - 1]) { err = "design: reuse_vec must start and end with 0"; return false; }
  // ---- get_empirical_ensemble (src/LDPC_Ensemble.cpp:391-423): edge-perspective degree distributions
  const int max_degree = 200;
  vec ve(max_degree, 0.0), ce(max_degree, 0.0);
  for (int d : P.dv) { if (d < 1 || d > max_degree) { err = "design: variable degree out of range"; return false; } ve[d - 1] += d; }
  for (int d : P.dc) { if (d < 1 || d > max_degree) { err = "design: check degree out of range"; return false; } ce[d - 1] += d; }
  ve = vdiv(ve, vsum(ve));
  ce = vdiv(ce, vsum(ce));
  ivec deg_lam, deg_rho;
  vec lam, rho;
  for (int i = 0; i < max_degree; ++i) {
    if (ve[i] > 0) { deg_lam.push_back(i + 1); lam.push_back(ve[i]); }
    if (ce[i] > 0) { deg_rho.push_back(i + 1); rho.push_back(ce[i]); }
  }
  const int dv_act = (int)deg_lam.size(), dc_act = (int)deg_rho.size();
  // ---- get_lut_tree_templates (src/LDPC_DE.cpp:1124-1290)
  std::string tm = P.tree_method, filename;
  const size_t eq = tm.find('=');
  if (eq != std::string::npos) { filename = tm.substr(eq + 1); tm = tm.substr(0, eq); }
  std::vector<std::vector<DTree>> var_t(max_iters), chk_t;
  if (tm == "filename") {
    std::map<std::string, std::map<std::string, std::string>> ini;
    if (filename.empty() || !read_ini(filename, ini, err)) { if (err.empty()) err = "design: tree_method 'filename=<file>' needs a file"; return false; }
    auto load = [&](const std::string& sec, const char* kp, const ivec& degs, int tree_type, int leaves_off,
                    std::vector<DTree>& out) -> bool {
      out.clear();
      for (int d : degs) {
        auto it = ini[sec].find(key3(kp, d));
        if (it == ini[sec].end()) { err = "design: no tree for " + key3(kp, d) + " in [" + sec + "]"; return false; }
        DTree t;
        size_t pos = 0;
        const int root = parse_template(it->second, pos, t, err);
        if (root != 0) { if (err.empty()) err = "design: empty tree template"; return false; }
        t.tree_type = tree_type;
        t.num_leaves = count_leaves(t, 0);
        if (t.num_leaves != d + leaves_off) { err = "design: LUT tree does not match node degree"; return false; }
        if (tree_type != CHKTREE && it->second.find('c') == std::string::npos) { err = "design: tree needs a channel leaf"; return false; }
        out.push_back(t);
      }
      return true;
    };
    if (!ini.count("var_iter_000")) { err = "design: section [var_iter_000] missing in the tree template file"; return false; }
    if (!load("var_iter_000", "var_deg_", deg_lam, VARTREE, 0, var_t[0])) return false;
    for (int ii = 1; ii < max_iters - 1; ++ii) {
      const std::string sec = key3("var_iter_", ii);
      if (ini.count(sec)) { if (!load(sec, "var_deg_", deg_lam, VARTREE, 0, var_t[ii])) return false; }
      else var_t[ii] = var_t[ii - 1];
    }
    if (!ini.count("DT")) { err = "design: section [DT] missing in the tree template file"; return false; }
    if (!load("DT", "var_deg_", deg_lam, DECTREE, 1, var_t[max_iters - 1])) return false;
    if (!P.min_lut) {
      chk_t.resize(max_iters);
      if (!ini.count("chk_iter_000")) { err = "design: section [chk_iter_000] missing in the tree template file"; return false; }
      if (!load("chk_iter_000", "chk_deg_", deg_rho, CHKTREE, -1, chk_t[0])) return false;
      for (int ii = 1; ii < max_iters; ++ii) {
        const std::string sec = key3("chk_iter_", ii);
        if (ini.count(sec)) { if (!load(sec, "chk_deg_", deg_rho, CHKTREE, -1, chk_t[ii])) return false; }
        else chk_t[ii] = chk_t[ii - 1];
      }
    }
  } else if (tm == "auto_bin_balanced" && filename.empty()) {
    for (int ii = 0; ii < max_iters; ++ii)
      for (int d : deg_lam) {
        if (d < 2) { err = "design: a degree-1 variable node cannot be given a LUT tree"; return false; }
        var_t[ii].push_back(ii == max_iters - 1 ? gen_bin_balanced(d + 1, true, DECTREE) : gen_bin_balanced(d, true, VARTREE));
      }
    if (!P.min_lut) {
      chk_t.resize(max_iters);
      for (int ii = 0; ii < max_iters; ++ii)
        for (int d : deg_rho) {
          if (d < 3) { err = "design: LUT check nodes need degree >= 3"; return false; }
          chk_t[ii].push_back(gen_bin_balanced(d - 1, false, CHKTREE));
        }
    }
  } else {
    err = "design: tree_method must be auto_bin_balanced or filename=<tree template ini> (ber_sim treats the other auto "
          "modes as auto_bin_balanced, src/LDPC_BER_Sim.cpp:487-488)";
    return false;
  }
  // ---- thresholds (get_quant_bound, src/LDPC_DE.cpp:561-601) and channel pmfs (set_channel_pmf, :400-412)
  const double sig = std::sqrt(P.sigma2);
  const double LLR_max = 25.0;
  const int Nq_fine = 5000;
  const double delta = 2 * LLR_max / Nq_fine;
  const vec fine = gaussian_pmf(2 / (sig * sig), 2 / sig, Nq_fine, delta);
  vec pmf_cha, pmf_var2chk, pmf_chk2var;
  auto bounds = [&](int K, vec& pmf, vec& qb) -> bool {
    ivec Q;
    if (!quant_mi_sym(pmf, Q, fine, K, true, err)) return false;
    const int M = Nq_fine;
    vec half(K / 2 - 1, 0.0);
    int label = 0;
    for (int mm = 0; mm < M / 2; ++mm)
      if (Q[M / 2 + mm] - K / 2 > label) {
        if (label >= (int)half.size()) { err = "design: quantiser has more boundaries than labels"; return false; }
        half[label++] = mm * delta;
      }
    qb.clear();
    for (int i = (int)half.size() - 1; i >= 0; --i) qb.push_back(-half[i]);
    qb.push_back(0.0);
    qb.insert(qb.end(), half.begin(), half.end());
    return true;
  };
  if (!bounds(P.Nq_Cha, pmf_cha, R.qb_Cha)) return false;
  if (!bounds(P.Nq_Msg[0], pmf_var2chk, R.qb_Msg)) return false;
  // ---- evolve(save_luts = true), src/LDPC_DE.cpp:198-326
  ivec Nq = P.Nq_Msg;
  Nq.push_back(2);  // the output of the last (decision) update is binary
  std::vector<DTree> var_iter, chk_iter;
  std::vector<std::vector<DTree>> var_saved, chk_saved;
  for (int ii = 0; ii < max_iters; ++ii) {
    if (P.min_lut) {  // chk_update_irr, min-sum
      pmf_chk2var.assign(Nq[ii], 0.0);
      for (int dd = 0; dd < dc_act; ++dd) {
        const vec pt = chk_update_minsum(pmf_var2chk, deg_rho[dd]);
        for (int k = 0; k < Nq[ii]; ++k) pmf_chk2var[k] = pmf_chk2var[k] + rho[dd] * pt[k];
      }
    } else if (!irr_update(chk_iter, chk_t[ii], reuse[ii] != 0, pmf_var2chk, pmf_cha, Nq[ii], Nq[ii], P.Nq_Cha, rho, pmf_chk2var, err)) {
      return false;
    }
    if (!irr_update(var_iter, var_t[ii], reuse[ii] != 0, pmf_chk2var, pmf_cha, Nq[ii], Nq[ii + 1], P.Nq_Cha, lam, pmf_var2chk, err))
      return false;
    if (!reuse[ii]) {
      var_saved.push_back(var_iter);
      if (!P.min_lut) chk_saved.push_back(chk_iter);
    }
  }
  R.var_tree_string = serialize(var_saved);
  R.chk_tree_string = serialize(chk_saved);
  // ---- Nq_Cha_2_Nq_Msg_map (src/LDPC_Code_LUT.cpp:735-741): from a UNIFORMLY quantised channel pmf
  {
    const double d2 = 2 * LLR_max / P.Nq_Cha;
    const vec pc = gaussian_pmf(2 / (sig * sig), 2 / sig, P.Nq_Cha, d2);
    vec pm;
    ivec map;
    if (!quant_mi_sym(pm, map, pc, P.Nq_Msg[0], true, err)) return false;
    R.cha2msg.assign(map.begin(), map.end());
  }
  R.dv_act = dv_act;
  R.dc_act = dc_act;
  return true;
}

}  // namespace design
}  // namespace lutldpc
