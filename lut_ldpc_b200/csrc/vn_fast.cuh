// Compile-time specialised variable-node pass for the tree shape every auto mode of the reference's
// ber_sim produces: LUT_Tree_Node::gen_bin_balanced_tree (src/LUT_Tree.cpp:200-237) -- the dv-1
// message leaves merged pairwise through a queue, then ROOT[subtree, CHA] -- with 4-bit inputs.
//
// The leave-one-out evaluation (LUT_Tree::var_msg_update, src/LUT_Tree.cpp:774-790) is unrolled at
// compile time into a straight-line list of two-input look-ups with common sub-expressions shared
// (the same algorithm as compile_program() in model.cpp, run in constexpr); all intermediate values
// live in registers and the table addresses are immediates.  Before a decoder uses this path the
// host checks that the program compiled from the decoder's actual trees is identical, op for op, to
// the constexpr one (fast_vn_matches); any other tree falls back to the generic interpreter k_node.
#pragma once
#include <cstdint>

#include "core.cuh"

namespace lutldpc {
namespace fastvn {

constexpr int kMaxDv = 24;
constexpr int kMaxOps = 192;

struct Op {
  int node, a, b;  // node = table index (order of first use); a, b = value ids
  bool root;
};
struct Prog {
  int dv = 0, n_ops = 0, n_tables = 0;
  Op ops[kMaxOps] = {};
  int out[kMaxDv] = {};  // value id of output ii
};

struct Builder {
  // tree: nodes [0, dv-1) message leaves, then inner nodes, one channel leaf, the root
  int left[4 * kMaxDv] = {}, right[4 * kMaxDv] = {}, kind[4 * kMaxDv] = {};  // kind 0 leaf, 1 inner
  int rank[4 * kMaxDv] = {};
  int n_nodes = 0, root = 0;
  Prog p;
  int seq[kMaxDv + 1] = {};
  int pos = 0;

  constexpr int eval(int n) {
    if (kind[n] == 0) return seq[pos++];
    int va = eval(left[n]);
    int vb = eval(right[n]);
    for (int i = 0; i < p.n_ops; ++i)
      if (p.ops[i].node == rank_of(n, false) && p.ops[i].a == va && p.ops[i].b == vb) return p.dv + 1 + i;
    int r = rank_of(n, true);
    p.ops[p.n_ops] = Op{r, va, vb, n == root};
    return p.dv + 1 + p.n_ops++;
  }
  constexpr int rank_of(int n, bool create) {
    if (rank[n] < 0 && create) rank[n] = p.n_tables++;
    return rank[n];
  }
};

constexpr Prog build(int dv) {
  Builder b;
  b.p.dv = dv;
  int q[2 * kMaxDv] = {};
  int head = 0, tail = 0;
  for (int i = 0; i < dv - 1; ++i) {
    b.kind[b.n_nodes] = 0;
    q[tail++] = b.n_nodes++;
  }
  while (tail - head > 1) {
    int l = q[head++], r = q[head++];
    b.kind[b.n_nodes] = 1;
    b.left[b.n_nodes] = l;
    b.right[b.n_nodes] = r;
    q[tail++] = b.n_nodes++;
  }
  int top = q[head];
  int cha = b.n_nodes++;
  b.kind[cha] = 0;
  b.root = b.n_nodes++;
  b.kind[b.root] = 1;
  b.left[b.root] = top;
  b.right[b.root] = cha;
  for (int i = 0; i < b.n_nodes; ++i) b.rank[i] = -1;
  for (int ii = 0; ii < dv; ++ii) {
    int n = 0;
    for (int k = 0; k < dv; ++k)
      if (k != ii) b.seq[n++] = k;
    b.seq[n++] = dv;
    b.pos = 0;
    b.p.out[ii] = b.eval(b.root);
  }
  return b.p;
}

template <int DV>
struct Static {
  static constexpr Prog P = build(DV);
};

#if defined(__CUDACC__)
// One two-input look-up for 8 frames: a, b hold 4-bit codes, tab is the 128-entry half table in
// shared memory indexed by a | mag(b) << 4 after folding b's sign out (model.hpp "sym").
__device__ __forceinline__ uint32_t lut2(uint32_t a, uint32_t b, const uint8_t* tab, bool root, int root_shift) {
  const uint32_t t = b & 0x88888888u;
  const uint32_t an = a ^ t, bn = b ^ t;
  const uint32_t X0 = (an & 0x0F0F0F0Fu) | ((bn << 4) & 0xF0F0F0F0u);
  const uint32_t X1 = ((an >> 4) & 0x0F0F0F0Fu) | (bn & 0xF0F0F0F0u);
  const uint32_t r0 = tab[X0 & 0xFF], r1 = tab[X1 & 0xFF];
  const uint32_t r2 = tab[__byte_perm(X0, 0, 0x4441)], r3 = tab[__byte_perm(X1, 0, 0x4441)];
  const uint32_t r4 = tab[__byte_perm(X0, 0, 0x4442)], r5 = tab[__byte_perm(X1, 0, 0x4442)];
  const uint32_t r6 = tab[X0 >> 24], r7 = tab[X1 >> 24];
  const uint32_t p0 = r1 * 16u + r0, p1 = r3 * 16u + r2, p2 = r5 * 16u + r4, p3 = r7 * 16u + r6;
  const uint32_t lo = p1 * 256u + p0, hi = p3 * 256u + p2;
  const uint32_t r = hi * 65536u + lo;
  return r ^ (root ? (t >> root_shift) : t);
}

template <int DV, int I>
struct Exec {
  static __device__ __forceinline__ void run(uint32_t (&v)[DV + 1 + Static<DV>::P.n_ops], const uint8_t* tabs,
                                             int root_shift) {
    if constexpr (I < Static<DV>::P.n_ops) {
      constexpr Op op = Static<DV>::P.ops[I];
      v[DV + 1 + I] = lut2(v[op.a], v[op.b], tabs + op.node * 128, op.root, root_shift);
      Exec<DV, I + 1>::run(v, tabs, root_shift);
    }
  }
};
#endif

}  // namespace fastvn
}  // namespace lutldpc
