// See graph.hpp.  Host-only C++.
#include "graph.hpp"

#include <algorithm>
#include <cstdio>
#include <fstream>
#include <sstream>

namespace lutldpc {

namespace {
// one line of the file as integers (the reference reads its alist line by line too: trailing zero padding of the
// index lists is simply never looked at)
bool ints_of_line(std::ifstream& f, std::vector<long>& out) {
  std::string line;
  if (!std::getline(f, line)) return false;
  out.clear();
  std::istringstream ss(line);
  long v;
  while (ss >> v) out.push_back(v);
  return true;
}
}  // namespace

bool load_alist_file(const char* path, ParityRows& H, std::string& err) {
  std::ifstream f(path);
  if (!f.is_open()) { err = std::string("cannot open alist file ") + path; return false; }
  std::vector<long> a;
  if (!ints_of_line(f, a) || a.size() < 2 || a[0] <= 0 || a[1] <= 0 || a[0] > (1 << 26) || a[1] > (1 << 26)) {
    err = "alist: bad first line (N M)"; return false;
  }
  const int N = (int)a[0], M = (int)a[1];
  if (!ints_of_line(f, a) || a.size() < 2 || a[0] < 0 || a[0] > M || a[1] < 0 || a[1] > N) {
    err = "alist: bad second line (max column weight, max row weight)"; return false;
  }
  std::vector<long> wn, wm;
  if (!ints_of_line(f, wn) || (int)wn.size() < N) { err = "alist: short column-weight line"; return false; }
  if (!ints_of_line(f, wm) || (int)wm.size() < M) { err = "alist: short row-weight line"; return false; }
  for (int i = 0; i < N; ++i) if (wn[i] < 0 || wn[i] > M) { err = "alist: column weight out of range"; return false; }
  for (int i = 0; i < M; ++i) if (wm[i] < 0 || wm[i] > N) { err = "alist: row weight out of range"; return false; }
  std::vector<std::vector<int32_t>> cols(N);
  for (int i = 0; i < N; ++i) {
    if (!ints_of_line(f, a) || (long)a.size() < wn[i]) { err = "alist: short column list"; return false; }
    for (long j = 0; j < wn[i]; ++j) {
      if (a[j] <= 0 || a[j] > M) { err = "alist: column list entry out of range"; return false; }
      cols[i].push_back((int32_t)a[j] - 1);
    }
    std::sort(cols[i].begin(), cols[i].end());
  }
  H = ParityRows();
  H.nvar = N; H.nchk = M;
  H.rows.assign(M, {});
  for (int i = 0; i < M; ++i) {
    if (!ints_of_line(f, a) || (long)a.size() < wm[i]) { err = "alist: short row list"; return false; }
    for (long j = 0; j < wm[i]; ++j) {
      if (a[j] <= 0 || a[j] > N) { err = "alist: row list entry out of range"; return false; }
      H.rows[i].push_back((int32_t)a[j] - 1);
    }
    std::sort(H.rows[i].begin(), H.rows[i].end());
    if (std::adjacent_find(H.rows[i].begin(), H.rows[i].end()) != H.rows[i].end()) {
      err = "alist: a row lists a column twice"; return false;
    }
  }
  // the two halves of the file must describe the same matrix
  std::vector<std::vector<int32_t>> cols2(N);
  for (int i = 0; i < M; ++i)
    for (int32_t v : H.rows[i]) cols2[v].push_back(i);
  if (cols2 != cols) { err = "alist: column lists and row lists disagree"; return false; }
  return true;
}

bool rows_from_graph(int nvar, int nchk, const std::vector<int32_t>& dv, const std::vector<int32_t>& dc,
                     const std::vector<int32_t>& cn_msg_idx, ParityRows& H, std::string& err) {
  if ((int)dv.size() != nvar || (int)dc.size() != nchk) { err = "graph: degree vectors have the wrong length"; return false; }
  std::vector<int32_t> edge_vn(cn_msg_idx.size());
  size_t e = 0;
  for (int v = 0; v < nvar; ++v)
    for (int k = 0; k < dv[v]; ++k) {
      if (e >= edge_vn.size()) { err = "graph: sum of variable degrees exceeds the edge count"; return false; }
      edge_vn[e++] = v;
    }
  if (e != edge_vn.size()) { err = "graph: sum of variable degrees differs from the edge count"; return false; }
  H = ParityRows();
  H.nvar = nvar; H.nchk = nchk;
  H.rows.assign(nchk, {});
  size_t at = 0;
  for (int c = 0; c < nchk; ++c) {
    for (int k = 0; k < dc[c]; ++k, ++at) {
      if (at >= cn_msg_idx.size() || cn_msg_idx[at] < 0 || cn_msg_idx[at] >= (int32_t)edge_vn.size()) {
        err = "graph: cn_msg_idx out of range"; return false;
      }
      H.rows[c].push_back(edge_vn[cn_msg_idx[at]]);
    }
    std::sort(H.rows[c].begin(), H.rows[c].end());
  }
  return true;
}

void flatten_graph(const ParityRows& H, FlatGraph& g) {
  g = FlatGraph();
  g.nvar = H.nvar; g.nchk = H.nchk;
  g.dv.assign(H.nvar, 0);
  g.dc.assign(H.nchk, 0);
  for (int c = 0; c < H.nchk; ++c) {
    g.dc[c] = (int32_t)H.rows[c].size();
    for (int32_t v : H.rows[c]) ++g.dv[v];
  }
  std::vector<int32_t> first(H.nvar + 1, 0);  // first edge of every VN
  for (int v = 0; v < H.nvar; ++v) first[v + 1] = first[v] + g.dv[v];
  g.num_edges = first[H.nvar];
  // Visiting the checks in ascending order hands every VN its edges in ascending check order, and within a check
  // the (sorted) row gives ascending VN order: exactly the two orders of decoder_parameterization (:515-527).
  std::vector<int32_t> next(first.begin(), first.end() - 1);
  g.cn_msg_idx.reserve(g.num_edges);
  for (int c = 0; c < H.nchk; ++c)
    for (int32_t v : H.rows[c]) g.cn_msg_idx.push_back(next[v]++);
}

void permute_columns(const ParityRows& H, const std::vector<int32_t>& perm, ParityRows& out) {
  out = ParityRows();
  out.nvar = H.nvar; out.nchk = H.nchk;
  out.rows.assign(H.nchk, {});
  for (int c = 0; c < H.nchk; ++c) {
    for (int32_t v : H.rows[c]) out.rows[c].push_back(perm[v]);
    std::sort(out.rows[c].begin(), out.rows[c].end());
  }
}

int gf2_rank_host(const ParityRows& H) {
  const size_t W = ((size_t)H.nvar + 63) / 64;
  std::vector<uint64_t> A((size_t)H.nchk * W, 0);
  for (int c = 0; c < H.nchk; ++c)
    for (int32_t v : H.rows[c]) A[(size_t)c * W + (v >> 6)] ^= 1ull << (v & 63);
  int rank = 0;
  for (int j = 0; j < H.nvar && rank < H.nchk; ++j) {
    const size_t w = (size_t)j >> 6;
    const uint64_t bit = 1ull << (j & 63);
    int p = -1;
    for (int i = rank; i < H.nchk; ++i)
      if (A[(size_t)i * W + w] & bit) { p = i; break; }
    if (p < 0) continue;
    if (p != rank) std::swap_ranges(A.begin() + (size_t)p * W, A.begin() + (size_t)(p + 1) * W, A.begin() + (size_t)rank * W);
    const uint64_t* pr = &A[(size_t)rank * W];
    for (int i = rank + 1; i < H.nchk; ++i) {
      uint64_t* ri = &A[(size_t)i * W];
      if (!(ri[w] & bit)) continue;
      for (size_t k = w; k < W; ++k) ri[k] ^= pr[k];  // columns before j are already zero in both rows
    }
    ++rank;
  }
  return rank;
}

}  // namespace lutldpc
