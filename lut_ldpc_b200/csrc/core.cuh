// Per-thread arithmetic of the batched LUT decoder: one 32-bit word = the same edge (or node
// value) of 8 consecutive frames, one 4-bit sign/magnitude code per frame (model.hpp "storage code").
// Everything here is plain integer SWAR and compiles for host and device, so tests/host_emul.cpp can
// drive exactly the code the kernels run.
//
// Reference semantics restated here:
//   check node (min-sum)   LDPC_Code_LUT::chk_update_minsum   src/LDPC_Code_LUT.cpp:355-402
//   LUT node evaluation    LUT_Tree_Node::var_msg_update      src/LUT_Tree.cpp:402-418
//                          LUT_Tree_Node::chk_msg_update      src/LUT_Tree.cpp:420-445
#pragma once
#include <cstdint>

#include "model.hpp"

#if defined(__CUDACC__)
#define LUT_HD __host__ __device__ __forceinline__
#else
#define LUT_HD inline
#endif

namespace lutldpc {

constexpr uint32_t kNibOnes = 0x11111111u;

// Masks for a q-bit code replicated over the 8 nibbles of a word.
struct QMask {
  uint32_t H;  // sign bit of every nibble
  uint32_t M;  // magnitude bits of every nibble
  int s;       // q-1: position of the sign bit inside the nibble
};
LUT_HD QMask make_qmask(int q) {
  QMask m;
  m.s = q - 1;
  m.H = kNibOnes << (q - 1);
  m.M = m.H - kNibOnes;
  return m;
}

// Per-nibble mask (all magnitude bits set) where a >= b; a, b are magnitude words (sign bits clear).
// The cleared sign-bit position serves as the guard bit of the subtraction.
LUT_HD uint32_t ge_mask(uint32_t a, uint32_t b, const QMask& qm) {
  uint32_t t = ((a | qm.H) - b) & qm.H;
  return t - (t >> qm.s);
}

// Running (min1, min2, sign parity) of a check node over 8 frames.
struct CnAcc {
  uint32_t min1, min2, par;
};
// first two inputs (every check has degree >= 2; this replaces the reference's nz sentinel, which
// does not fit the magnitude field)
LUT_HD void cn_init(CnAcc& a, uint32_t w0, uint32_t w1, const QMask& qm) {
  uint32_t x0 = w0 & qm.M, x1 = w1 & qm.M;
  uint32_t m = ge_mask(x0, x1, qm);
  a.min1 = (x1 & m) | (x0 & ~m);
  a.min2 = (x0 & m) | (x1 & ~m);
  a.par = w0 ^ w1;
}
LUT_HD void cn_accum(CnAcc& a, uint32_t w, const QMask& qm) {
  uint32_t x = w & qm.M;
  uint32_t m = ge_mask(a.min1, x, qm);           // min1 >= x
  uint32_t hi = (a.min1 & m) | (x & ~m);         // max(min1, x)
  a.min1 = (x & m) | (a.min1 & ~m);
  uint32_t m2 = ge_mask(a.min2, hi, qm);
  a.min2 = (hi & m2) | (a.min2 & ~m2);
  a.par ^= w;
}
// Outgoing message on the edge that carried w: magnitude min2 if w holds the minimum (ties make
// min2 == min1, so "first arg-min" of the reference needs no index), else min1; sign = parity ^ own.
LUT_HD uint32_t cn_emit(const CnAcc& a, uint32_t w, const QMask& qm) {
  uint32_t x = w & qm.M;
  uint32_t me = ge_mask(a.min1, x, qm);          // x == min1 (min1 <= x always)
  uint32_t mag = (a.min2 & me) | (a.min1 & ~me);
  return mag | ((w ^ a.par) & qm.H);
}

// Guarded variant used by the vectorised check-node kernel: min1/min2 carry the guard bit (sign-bit
// position set) permanently, which saves the OR of every comparison; the selects below preserve it.
struct CnAccG {
  uint32_t min1g, min2g, par;
};
LUT_HD uint32_t ge_mask_g(uint32_t ag, uint32_t b, const QMask& qm) {  // ag has the guard bits set
  uint32_t t = (ag - b) & qm.H;
  return t - (t >> qm.s);
}
LUT_HD void cng_init(CnAccG& a, uint32_t w0, uint32_t w1, const QMask& qm) {
  uint32_t x0 = w0 & qm.M, x1 = w1 & qm.M;
  uint32_t m = ge_mask_g(x0 | qm.H, x1, qm);
  a.min1g = ((x1 & m) | (x0 & ~m)) | qm.H;
  a.min2g = ((x0 & m) | (x1 & ~m)) | qm.H;
  a.par = w0 ^ w1;
}
LUT_HD void cng_accum(CnAccG& a, uint32_t w, const QMask& qm) {
  uint32_t x = w & qm.M;
  uint32_t m = ge_mask_g(a.min1g, x, qm);        // min1 >= x
  uint32_t hi = (a.min1g & m) | (x & ~m);        // max(min1, x), guard bit clear (m covers magnitudes only)
  a.min1g = (x & m) | (a.min1g & ~m);            // guard bit kept from min1g
  uint32_t m2 = ge_mask_g(a.min2g, hi, qm);
  a.min2g = (hi & m2) | (a.min2g & ~m2);
  a.par ^= w;
}
LUT_HD uint32_t cng_emit(const CnAccG& a, uint32_t w, const QMask& qm) {
  uint32_t x = w & qm.M;
  uint32_t me = ge_mask_g(a.min1g, x, qm);       // x == min1
  uint32_t mag = (a.min2g & me) | (a.min1g & ~me);   // bit H set (guards) -> replaced by the sign below
  return (mag & qm.M) | ((w ^ a.par) & qm.H);
}

// Gathers bit `pos` of every nibble into the low byte (bit j = frame j).
LUT_HD uint32_t nib_bits_to_byte(uint32_t x, int pos) {
  x = (x >> pos) & kNibOnes;
  x = (x | (x >> 3)) & 0x03030303u;
  x = (x | (x >> 6)) & 0x000F000Fu;
  x = (x | (x >> 12)) & 0xFFu;
  return x;
}
// Spreads the low byte to bit 0 of every nibble.
LUT_HD uint32_t byte_to_nib_bits(uint32_t b) {
  b &= 0xFFu;
  b = (b | (b << 12)) & 0x000F000Fu;
  b = (b | (b << 6)) & 0x03030303u;
  b = (b | (b << 3)) & kNibOnes;
  return b;
}

// One LUT node for 8 frames.  `Vals` supplies get(slot); `tab` is the op's table (1 byte/entry).
template <class Vals>
LUT_HD uint32_t exec_op(const DevOp& op, const uint8_t* tab, const Vals& vals) {
  const int n = op.n_in;
  uint32_t t = 0, par = 0;
  if (op.sym) t = (vals.get(op.in_slot[n - 1]) >> (op.q[n - 1] - 1)) & kNibOnes;
  uint32_t X0 = 0, X1 = 0, X2 = 0, X3 = 0;
  const bool narrow = op.kind == OP_NARROW;
  for (int k = 0; k < n; ++k) {
    uint32_t x = vals.get(op.in_slot[k]);
    const int sk = op.q[k] - 1, sh = op.sh[k];
    if (op.chk) {
      par ^= x >> sk;
      x &= (kNibOnes << sk) - kNibOnes;
    } else {
      x ^= t << sk;  // sign normalisation; clears the last input's own sign bit
    }
    if (narrow) {
      X0 |= (x & 0x0F0F0F0Fu) << sh;         // frames 0,2,4,6 in byte lanes
      X1 |= ((x >> 4) & 0x0F0F0F0Fu) << sh;  // frames 1,3,5,7
    } else {
      X0 |= (x & 0x000F000Fu) << sh;          // frames 0,4 in 16-bit lanes
      X1 |= ((x >> 4) & 0x000F000Fu) << sh;   // frames 1,5
      X2 |= ((x >> 8) & 0x000F000Fu) << sh;   // frames 2,6
      X3 |= ((x >> 12) & 0x000F000Fu) << sh;  // frames 3,7
    }
  }
  uint32_t r;
  if (narrow) {
    r = (uint32_t)tab[X0 & 0xFF] | ((uint32_t)tab[X1 & 0xFF] << 4) |
        ((uint32_t)tab[(X0 >> 8) & 0xFF] << 8) | ((uint32_t)tab[(X1 >> 8) & 0xFF] << 12) |
        ((uint32_t)tab[(X0 >> 16) & 0xFF] << 16) | ((uint32_t)tab[(X1 >> 16) & 0xFF] << 20) |
        ((uint32_t)tab[X0 >> 24] << 24) | ((uint32_t)tab[X1 >> 24] << 28);
  } else {
    r = (uint32_t)tab[X0 & 0xFFFF] | ((uint32_t)tab[X1 & 0xFFFF] << 4) |
        ((uint32_t)tab[X2 & 0xFFFF] << 8) | ((uint32_t)tab[X3 & 0xFFFF] << 12) |
        ((uint32_t)tab[X0 >> 16] << 16) | ((uint32_t)tab[X1 >> 16] << 20) |
        ((uint32_t)tab[X2 >> 16] << 24) | ((uint32_t)tab[X3 >> 16] << 28);
  }
  const int so = op.q_out - 1;
  if (op.chk) r ^= (~par & kNibOnes) << so;  // parity 0 -> K-1-Q[label]  (src/LUT_Tree.cpp:440-443)
  else r ^= t << so;
  return r;
}

template <class Vals>
LUT_HD void run_program(uint32_t n_ops, const DevOp* ops, const uint8_t* tables, Vals& vals) {
  for (uint32_t i = 0; i < n_ops; ++i) {
    uint32_t r = exec_op(ops[i], tables + ops[i].table_off, vals);
    vals.set(ops[i].out_slot, r);
  }
}

// label <-> storage code for one nibble word, via 16-entry maps packed as 64-bit constants
// (entry i in bits 4i..4i+3)
LUT_HD uint32_t map_nibbles(uint32_t w, uint64_t map16) {
  uint32_t r = 0;
  for (int j = 0; j < 8; ++j) r |= (uint32_t)((map16 >> (((w >> (4 * j)) & 0xF) * 4)) & 0xF) << (4 * j);
  return r;
}

}  // namespace lutldpc
