// sm_100a kernels of the batched LUT-LDPC decoder.  One thread owns one 32-bit word = one edge (or
// node) of 8 consecutive frames; a warp covers 32 consecutive words of a row, i.e. one 128-byte line
// per edge, so every global access of the decode loop is a fully coalesced line.  Integer work only:
// no tensor cores (this is table look-up and compare, not a contraction).
#include "kernels.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "kdev.cuh"

namespace lutldpc {

// label -> storage code for one byte label (model.hpp)
__device__ __forceinline__ uint32_t lab2code(uint32_t m, int Nq, int q) {
  uint32_t nz = Nq >> 1;
  m = m < (uint32_t)Nq ? m : (uint32_t)Nq - 1;  // out-of-range labels are clamped, never indexed with
  return m < nz ? ((1u << (q - 1)) | (nz - 1 - m)) : (m - nz);
}

// ---------------------------------------------------------------------------------------------
// pack: frame-major byte labels -> frame-interleaved nibble words.
//   LLRout = (LLRin_cha < Nq_Cha/2)                      src/LDPC_Code_LUT.cpp:275
//   msgs[e] = LLRin_msg[v] for every edge e of v         src/LDPC_Code_LUT.cpp:284-289
// Tile = 32 variable nodes x 256 frames through shared memory so that reads are 32-byte sectors of a
// frame row and writes are 128-byte lines of an edge row.
// ---------------------------------------------------------------------------------------------
// PACKED: one byte per (frame, node) with the channel label in the low and the message label in the high nibble
// (cha_lab; msg_lab unused) instead of one byte per label: half the bytes across PCIe (lutldpc_decode_packed).
template <bool VEC4, bool PACKED>
__global__ void __launch_bounds__(256) k_pack(const uint8_t* __restrict__ cha_lab, const uint8_t* __restrict__ msg_lab,
                                              const int32_t* __restrict__ vn_edge0, int nvar, int F, int W,
                                              int Nq_cha, int q_cha, int Nq_msg, int q_msg,
                                              uint32_t* __restrict__ cha, uint32_t* __restrict__ msgs,
                                              uint8_t* __restrict__ hard, uint32_t* __restrict__ msg0) {
  // one byte per (frame, node): channel code in the low nibble, message code in the high nibble
  __shared__ uint8_t t[256][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int v0 = blockIdx.x * 32, w0 = blockIdx.y * 32, f0 = w0 * 8;
  if (VEC4) {  // nvar % 4 == 0 and 4-byte aligned label arrays: four nodes per load
    const int q4 = threadIdx.x & 7, fr = threadIdx.x >> 3;
    const int v = v0 + 4 * q4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int fl = fr + 32 * i, f = f0 + fl;
      uint32_t c4 = 0, m4 = 0;
      const bool in = f < F && v < nvar;
      if (in) {
        c4 = *reinterpret_cast<const uint32_t*>(cha_lab + (size_t)f * nvar + v);
        if (PACKED) {
          m4 = (c4 >> 4) & 0x0F0F0F0Fu;
          c4 &= 0x0F0F0F0Fu;
        } else {
          m4 = *reinterpret_cast<const uint32_t*>(msg_lab + (size_t)f * nvar + v);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        // padded frames carry the weakest positive label (code 0)
        const uint32_t c = in ? lab2code((c4 >> (8 * k)) & 0xFF, Nq_cha, q_cha) : 0u;
        const uint32_t m = in ? lab2code((m4 >> (8 * k)) & 0xFF, Nq_msg, q_msg) : 0u;
        t[fl][4 * q4 + k] = (uint8_t)(c | (m << 4));
      }
    }
  } else {
    for (int i = 0; i < 32; ++i) {
      int fl = ty + 8 * i, f = f0 + fl, v = v0 + tx;
      uint8_t c = 0, m = 0;
      bool in = f < F && v < nvar;
      if (in) {
        c = cha_lab[(size_t)f * nvar + v];
        if (PACKED) {
          m = c >> 4;
          c &= 0xF;
        } else {
          m = msg_lab[(size_t)f * nvar + v];
        }
      }
      t[fl][tx] = in ? (uint8_t)(lab2code(c, Nq_cha, q_cha) | (lab2code(m, Nq_msg, q_msg) << 4)) : 0;
    }
  }
  __syncthreads();
  const int w = w0 + tx;
  for (int vi = ty; vi < 32; vi += 8) {
    int v = v0 + vi;
    if (v >= nvar || w >= W) continue;
    uint32_t cw = 0, mw = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t x = t[8 * tx + j][vi];
      cw |= (x & 0xFu) << (4 * j);
      mw |= (x >> 4) << (4 * j);
    }
    cha[(size_t)v * W + w] = cw;
    hard[(size_t)v * W + w] = (uint8_t)nib_bits_to_byte(cw, q_cha - 1);
    if (msg0) {  // on-chip regime: the decode kernel broadcasts the word to the node's edges in shared memory
      msg0[(size_t)v * W + w] = mw;
      continue;
    }
    int e0 = vn_edge0[v], e1 = vn_edge0[v + 1];
    for (int e = e0; e < e1; ++e) msgs[(size_t)e * W + w] = mw;
  }
}

// bits [F][nvar] (0/1) -> hard[nvar][W]; used by the stand-alone syndrome check
__global__ void __launch_bounds__(256) k_pack_bits(const uint8_t* __restrict__ bits, int nvar, int F, int W,
                                                   uint8_t* __restrict__ hard) {
  __shared__ uint8_t tb[256][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int v0 = blockIdx.x * 32, w0 = blockIdx.y * 32, f0 = w0 * 8;
  for (int i = 0; i < 32; ++i) {
    int fl = ty + 8 * i, f = f0 + fl, v = v0 + tx;
    tb[fl][tx] = (f < F && v < nvar) ? (bits[(size_t)f * nvar + v] & 1) : 0;
  }
  __syncthreads();
  const int w = w0 + tx;
  for (int vi = ty; vi < 32; vi += 8) {
    int v = v0 + vi;
    if (v >= nvar || w >= W) continue;
    uint32_t b = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) b |= (uint32_t)tb[8 * tx + j][vi] << j;
    hard[(size_t)v * W + w] = (uint8_t)b;
  }
}

// done = padding frames; fail flags cleared
__global__ void k_reset_flags(uint32_t* done, uint32_t* vn_fail, uint32_t* cn_fail, uint32_t* syn_fail,
                              int32_t* iters, uint32_t* n_active, int W, int F) {
  int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w == 0) *n_active = (uint32_t)F;
  if (w >= W) return;
  uint32_t d = 0;
  for (int j = 0; j < 8; ++j) {
    int f = 8 * w + j;
    if (f >= F) d |= 1u << (4 * j);
    iters[f] = 0;
  }
  done[w] = d;
  vn_fail[w] = 0;
  cn_fail[w] = 0;
  syn_fail[w] = 0;
}

// ---------------------------------------------------------------------------------------------
// check-node pass, min-sum (LDPC_Code_LUT::chk_update_minsum, src/LDPC_Code_LUT.cpp:355-402).
// Thread = (check, word).  The sign parity of the incoming messages doubles as the syndrome bit of
// the previous variable-node pass (src/LDPC_Code_LUT.cpp:437-452: when every VN is unanimous, the
// sign on each edge IS the tentative bit), so the early-termination syndrome costs no extra pass.
// ---------------------------------------------------------------------------------------------
template <int MAXDC>
__global__ void __launch_bounds__(128) k_cn_minsum(const int32_t* __restrict__ cn_start,
                                                   const int32_t* __restrict__ cn_msg_idx, int nchk, int W, int q,
                                                   uint32_t* __restrict__ msgs, const uint32_t* __restrict__ done,
                                                   uint32_t* __restrict__ cn_fail, int want_fail) {
  const int c = blockIdx.x;
  const int w = blockIdx.y * blockDim.x + threadIdx.x;
  if (w >= W) return;
  if (done[w] == kNibOnes) return;  // all 8 frames of this word have finished
  const QMask qm = make_qmask(q);
  const int s = cn_start[c], dc = cn_start[c + 1] - s;
  const int32_t* idx = cn_msg_idx + s;
  uint32_t x[MAXDC];
#pragma unroll
  for (int k = 0; k < MAXDC; ++k)
    if (k < dc) x[k] = msgs[(size_t)idx[k] * W + w];
  CnAcc a;
  cn_init(a, x[0], x[1], qm);
#pragma unroll
  for (int k = 2; k < MAXDC; ++k)
    if (k < dc) cn_accum(a, x[k], qm);
#pragma unroll
  for (int k = 0; k < MAXDC; ++k)
    if (k < dc) msgs[(size_t)idx[k] * W + w] = cn_emit(a, x[k], qm);
  if (want_fail) {
    uint32_t f = (a.par >> qm.s) & kNibOnes;
    flag_or(&cn_fail[w], f);
  }
}

// Vectorised variant: each thread owns VEC consecutive words (8*VEC frames) of one check, so a warp
// moves 128*VEC contiguous bytes per edge row and keeps dc*VEC independent 4-byte loads in flight.

template <int MAXDC, int VEC>
__global__ void __launch_bounds__(128) k_cn_minsum_v(const int32_t* __restrict__ cn_start,
                                                     const int32_t* __restrict__ cn_msg_idx, int nchk, int W, int q,
                                                     uint32_t* __restrict__ msgs, const uint32_t* __restrict__ done,
                                                     uint32_t* __restrict__ cn_fail, int want_fail) {
  typedef typename WordVec<VEC>::T V;
  const int c = blockIdx.x;
  const int w = (blockIdx.y * blockDim.x + threadIdx.x) * VEC;
  if (w >= W) return;
  union U { V v; uint32_t u[VEC]; };
  U dn;
  dn.v = *reinterpret_cast<const V*>(done + w);
  bool all_done = true;
#pragma unroll
  for (int j = 0; j < VEC; ++j) all_done &= dn.u[j] == kNibOnes;
  if (all_done) return;
  const QMask qm = make_qmask(q);
  const int s = cn_start[c], dc = cn_start[c + 1] - s;
  const int32_t* idx = cn_msg_idx + s;
  int row[MAXDC];
#pragma unroll
  for (int k = 0; k < MAXDC; ++k)
    if (k < dc) row[k] = idx[k];
  U x[MAXDC];
#pragma unroll
  for (int k = 0; k < MAXDC; ++k)
    if (k < dc) x[k].v = __ldcs(reinterpret_cast<const V*>(msgs + (size_t)row[k] * W + w));
  CnAccG a[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) cng_init(a[j], x[0].u[j], x[1].u[j], qm);
#pragma unroll
  for (int k = 2; k < MAXDC; ++k)
    if (k < dc) {
#pragma unroll
      for (int j = 0; j < VEC; ++j) cng_accum(a[j], x[k].u[j], qm);
    }
#pragma unroll
  for (int k = 0; k < MAXDC; ++k)
    if (k < dc) {
      U o;
#pragma unroll
      for (int j = 0; j < VEC; ++j) o.u[j] = cng_emit(a[j], x[k].u[j], qm);
      __stcs(reinterpret_cast<V*>(msgs + (size_t)row[k] * W + w), o.v);
    }
  if (want_fail) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      uint32_t f = (a[j].par >> qm.s) & kNibOnes;
      flag_or(&cn_fail[w + j], f);
    }
  }
}

// Exact-degree variant: all checks of one launch have degree DC and their edge rows are stored back to
// back (rows[i*DC + k]), so the load chain is rows -> messages (no cn_start / predication) and the
// whole pass is straight-line code.
template <int DC, int VEC>
__global__ void __launch_bounds__(128) k_cn_exact(const int32_t* __restrict__ rows, int W, int q,
                                                  uint32_t* __restrict__ msgs, const uint32_t* __restrict__ done,
                                                  uint32_t* __restrict__ cn_fail, int want_fail,
                                                  const uint32_t* __restrict__ grp4,
                                                  const uint32_t* __restrict__ n_lists, int n_checks, int tx_log2,
                                                  int per) {
  typedef typename WordVec<VEC>::T V;
  // thread i of the pass owns VEC adjacent columns of live group grp4[i / (4 / VEC)]; a CTA covers
  // 2^tx_log2 such threads of 128 >> tx_log2 checks (narrow layouts: several checks per CTA)
  pdl_launch_dependents();
  // `per` checks in sequence per thread: fewer, longer-lived CTAs
  const int check0 = ((blockIdx.x * per) << (7 - tx_log2)) + (threadIdx.x >> tx_log2);
  const int i = (blockIdx.y << tx_log2) + (threadIdx.x & ((1 << tx_log2) - 1));
  pdl_wait();
  if (i >= (int)n_lists[3] * (4 / VEC)) return;
  const int w = (int)grp4[i / (4 / VEC)] * 4 + (i % (4 / VEC)) * VEC;
  union U { V v; uint32_t u[VEC]; };
  U dn;
  dn.v = *reinterpret_cast<const V*>(done + w);
  bool all_done = true;
#pragma unroll
  for (int j = 0; j < VEC; ++j) all_done &= dn.u[j] == kNibOnes;
  if (all_done) return;
  const QMask qm = make_qmask(q);
  uint32_t failacc[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) failacc[j] = 0;
  for (int cc = 0; cc < per; ++cc) {
  const int check = check0 + (cc << (7 - tx_log2));
  if (check >= n_checks) break;
  const int32_t* r = rows + (size_t)check * DC;
  uint32_t* p[DC];
#pragma unroll
  for (int k = 0; k < DC; ++k) p[k] = msgs + (size_t)r[k] * W + w;
  U x[DC];
#pragma unroll
  for (int k = 0; k < DC; ++k) x[k].v = __ldcs(reinterpret_cast<const V*>(p[k]));
  CnAccG a[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) cng_init(a[j], x[0].u[j], x[1].u[j], qm);
#pragma unroll
  for (int k = 2; k < DC; ++k) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) cng_accum(a[j], x[k].u[j], qm);
  }
#pragma unroll
  for (int k = 0; k < DC; ++k) {
    U o;
#pragma unroll
    for (int j = 0; j < VEC; ++j) o.u[j] = cng_emit(a[j], x[k].u[j], qm);
    __stcs(reinterpret_cast<V*>(p[k]), o.v);
  }
#pragma unroll
  for (int j = 0; j < VEC; ++j) failacc[j] |= a[j].par;
  }
  if (want_fail) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) flag_or(&cn_fail[w + j], (failacc[j] >> qm.s) & kNibOnes);
  }
}

// any degree (up to the reference's max_cnd = 200, src/LDPC_Code_LUT.hpp:363): second pass re-reads
__global__ void __launch_bounds__(128) k_cn_minsum_any(const int32_t* __restrict__ cn_start,
                                                       const int32_t* __restrict__ cn_msg_idx, int nchk, int W,
                                                       int q, uint32_t* __restrict__ msgs,
                                                       const uint32_t* __restrict__ done,
                                                       uint32_t* __restrict__ cn_fail, int want_fail) {
  const int c = blockIdx.x;
  const int w = blockIdx.y * blockDim.x + threadIdx.x;
  if (w >= W) return;
  if (done[w] == kNibOnes) return;
  const QMask qm = make_qmask(q);
  const int s = cn_start[c], dc = cn_start[c + 1] - s;
  const int32_t* idx = cn_msg_idx + s;
  CnAcc a;
  cn_init(a, msgs[(size_t)idx[0] * W + w], msgs[(size_t)idx[1] * W + w], qm);
  for (int k = 2; k < dc; ++k) cn_accum(a, msgs[(size_t)idx[k] * W + w], qm);
  for (int k = 0; k < dc; ++k) {
    size_t o = (size_t)idx[k] * W + w;
    msgs[o] = cn_emit(a, msgs[o], qm);
  }
  if (want_fail) {
    uint32_t f = (a.par >> qm.s) & kNibOnes;
    flag_or(&cn_fail[w], f);
  }
}

// ---------------------------------------------------------------------------------------------
// LUT node pass: variable-node update (LUT_Tree::var_msg_update, src/LUT_Tree.cpp:774-790), final
// decision (LUT_Tree::dec_update, :809-820) or LUT check-node update (LUT_Tree::chk_msg_update,
// :792-807).  All nodes of one launch share a degree and therefore one straight-line node program
// (model.hpp): the op list and its LUTs are staged into shared memory with one bulk copy, every
// thread then runs the program for its 8 frames with the intermediate values in a private column of
// shared memory (slot-major, so accesses are conflict-free).
// ---------------------------------------------------------------------------------------------
struct SmemVals {
  uint32_t* base;  // this thread's column
  __device__ __forceinline__ uint32_t get(int s) const { return base[s * kNodeThreads]; }
  __device__ __forceinline__ void set(int s, uint32_t v) { base[s * kNodeThreads] = v; }
};

template <int MODE>
__global__ void __launch_bounds__(kNodeThreads) k_node(const uint8_t* __restrict__ prog, uint32_t prog_bytes,
                                                        const int32_t* __restrict__ node_list, int n_nodes,
                                                        int nodes_per_cta, const int32_t* __restrict__ start,
                                                        const int32_t* __restrict__ cn_msg_idx, int W,
                                                        uint32_t* __restrict__ msgs, const uint32_t* __restrict__ cha,
                                                        uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
                                                        uint32_t* __restrict__ fail, int want_flags,
                                                        const uint32_t* __restrict__ cols,
                                                        const uint32_t* __restrict__ n_lists) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  const int n_cols = (int)n_lists[2];
  if ((int)(blockIdx.y * kNodeThreads) >= n_cols) return;  // whole CTA beyond the live columns
  stage_to_smem(smem, prog, prog_bytes, &bar);
  const DevProgHeader* hdr = reinterpret_cast<const DevProgHeader*>(smem);
  const DevOp* ops = reinterpret_cast<const DevOp*>(smem + hdr->ops_off);
  const uint8_t* outs = smem + hdr->outs_off;
  const uint8_t* tables = smem + hdr->tables_off;
  const uint32_t n_ops = hdr->n_ops;
  const int n_in = (int)hdr->n_in, n_out = (int)hdr->n_out;
  SmemVals vals{reinterpret_cast<uint32_t*>(smem + prog_bytes) + threadIdx.x};

  const int ci = blockIdx.y * kNodeThreads + threadIdx.x;
  if (ci >= n_cols) return;
  const int w = (int)cols[ci];
  const uint32_t dn = done[w];
  if (dn == kNibOnes) return;
  const uint32_t done_byte = nib_bits_to_byte(dn, 0);
  const int so = (int)hdr->q_out - 1, si = (int)hdr->q_in - 1;
  uint32_t failacc = 0;
  const int i0 = blockIdx.x * nodes_per_cta;
  const int i1 = min(i0 + nodes_per_cta, n_nodes);
  for (int i = i0; i < i1; ++i) {
    const int node = node_list[i];
    const int s = start[node];
    uint32_t par = 0;
    for (int k = 0; k < n_in; ++k) {
      const size_t row = MODE == NODE_CHK ? (size_t)cn_msg_idx[s + k] : (size_t)(s + k);
      uint32_t x = msgs[row * W + w];
      vals.set(k, x);
      if (MODE == NODE_CHK) par ^= x;
    }
    if (MODE != NODE_CHK) vals.set(n_in, cha[(size_t)node * W + w]);
    run_program(n_ops, ops, tables, vals);
    if (MODE == NODE_VAR) {
      uint32_t A = 0xFFFFFFFFu, O = 0, first = 0;
      for (int ii = 0; ii < n_out; ++ii) {
        uint32_t r = vals.get(outs[ii]);
        msgs[(size_t)(s + ii) * W + w] = r;
        A &= r;
        O |= r;
        if (ii == 0) first = r;
      }
      if (want_flags) {
        failacc |= (A ^ O) >> so;  // some outgoing signs differ (src/LDPC_Code_LUT.cpp:441-447)
        uint8_t* hp = hard + (size_t)node * W + w;
        uint32_t nb = nib_bits_to_byte(first, so);
        *hp = (uint8_t)((*hp & done_byte) | (nb & ~done_byte));
      }
    } else if (MODE == NODE_CHK) {
      for (int ii = 0; ii < n_out; ++ii)
        msgs[(size_t)cn_msg_idx[s + ii] * W + w] = vals.get(outs[ii]);
      failacc |= par >> si;
    } else {  // NODE_DEC: bit = (root output < 1), root alphabet 2 -> the sign bit (src/LDPC_Code_LUT.cpp:342)
      uint8_t* hp = hard + (size_t)node * W + w;
      uint32_t nb = nib_bits_to_byte(vals.get(outs[0]), so);
      *hp = (uint8_t)((*hp & done_byte) | (nb & ~done_byte));
    }
  }
  if (want_flags && MODE != NODE_DEC) {
    failacc &= kNibOnes;
    flag_or(&fail[w], failacc);
  }
}

// syndrome of the hard decisions (syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469)
__global__ void __launch_bounds__(128) k_syndrome(const int32_t* __restrict__ cn_start,
                                                  const int32_t* __restrict__ chk_equ_idx, int W,
                                                  const uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
                                                  uint32_t* __restrict__ syn_fail, const uint32_t* __restrict__ cols,
                                                  const uint32_t* __restrict__ n_lists) {
  const int c = blockIdx.x;
  const int ci = blockIdx.y * blockDim.x + threadIdx.x;
  if (ci >= (int)n_lists[2]) return;
  const int w = (int)cols[ci];
  if (done[w] == kNibOnes) return;
  const int s = cn_start[c], e = cn_start[c + 1];
  uint32_t x = 0;
  for (int k = s; k < e; ++k) x ^= hard[(size_t)chk_equ_idx[k] * W + w];
  if (x) flag_or(&syn_fail[w], byte_to_nib_bits(x));
}

// four word columns per thread (one 4-byte read per variable node of the check): thread i owns live group grp4[i]
__global__ void __launch_bounds__(128) k_syndrome4(const int32_t* __restrict__ cn_start,
                                                   const int32_t* __restrict__ chk_equ_idx, int W,
                                                   const uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
                                                   uint32_t* __restrict__ syn_fail, const uint32_t* __restrict__ grp4,
                                                   const uint32_t* __restrict__ n_lists, int nchk,
                                                   int checks_per_cta) {
  const int gi = blockIdx.y * blockDim.x + threadIdx.x;
  if (gi >= (int)n_lists[3]) return;
  const int w = (int)grp4[gi] * 4;
  uint32_t acc = 0;  // OR over the checks of this CTA: a frame fails if any check fails
  const int c1 = min(nchk, (int)(blockIdx.x + 1) * checks_per_cta);
  for (int c = blockIdx.x * checks_per_cta; c < c1; ++c) {
    const int s = cn_start[c], e = cn_start[c + 1];
    uint32_t x = 0;
    for (int k = s; k < e; ++k) x ^= *reinterpret_cast<const uint32_t*>(hard + (size_t)chk_equ_idx[k] * W + w);
    acc |= x;
  }
  if (!acc) return;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t b = (acc >> (8 * j)) & 0xFFu;
    if (b && done[w + j] != kNibOnes) flag_or(&syn_fail[w + j], byte_to_nib_bits(b));
  }
}

// Single-CTA (1024 threads) order-preserving compaction of the live columns: cols = words with an
// unfinished frame, grp4 = groups of 4 adjacent words with any such word; counts to n_active[2], [3].
__device__ void build_lists(int W, const uint32_t* done, uint32_t* cols, uint32_t* grp4, uint32_t* n_active) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t base_s;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int pass = 0; pass < 2; ++pass) {
    const int n = pass == 0 ? W : W / 4;
    uint32_t* list = pass == 0 ? cols : grp4;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (int c0 = 0; c0 < n; c0 += 1024) {
      const int i = c0 + threadIdx.x;
      bool live = false;
      if (i < n) {
        if (pass == 0) {
          live = done[i] != kNibOnes;
        } else {
          const uint4 d = *reinterpret_cast<const uint4*>(done + 4 * i);
          live = !(d.x == kNibOnes && d.y == kNibOnes && d.z == kNibOnes && d.w == kNibOnes);
        }
      }
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, live);
      if (lane == 0) wsum[wid] = __popc(bal);
      __syncthreads();
      uint32_t before = 0, total = 0;
      for (int j = 0; j < 32; ++j) {
        const uint32_t v = wsum[j];
        before += j < wid ? v : 0;
        total += v;
      }
      const uint32_t base = base_s;
      if (live) list[base + before + __popc(bal & ((1u << lane) - 1u))] = (uint32_t)i;
      __syncthreads();
      if (threadIdx.x == 0) base_s = base + total;
      __syncthreads();
    }
    if (threadIdx.x == 0) n_active[2 + pass] = base_s;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) k_build_lists(int W, const uint32_t* done, uint32_t* cols, uint32_t* grp4,
                                                      uint32_t* n_active) {
  build_lists(W, done, cols, grp4, n_active);
}

// Per-frame bookkeeping between passes (single CTA; W is at most a few thousand words).
//   FIN_INIT  pisc: frames whose channel hard decision is a codeword return 0     (:277)
//   FIN_ITER  psc : unanimous VNs + zero syndrome -> return `value` = ii+1        (:327)
//   FIN_LAST  after the decision pass: +-max_iters                                (:346-349)
// Covers columns [w0, w1); when that is the whole batch (W_lists > 0) the work lists are rebuilt.
__global__ void __launch_bounds__(1024) k_finalize(int mode, int value, int w0, int w1, uint32_t* done,
                                                   uint32_t* vn_fail, uint32_t* cn_fail, uint32_t* syn_fail,
                                                   int32_t* iters, uint32_t* n_active, int W_lists, uint32_t* cols,
                                                   uint32_t* grp4, uint32_t* n_lists, uint32_t* mirror) {
  __shared__ uint32_t red[32];
  pdl_launch_dependents();
  pdl_wait();
  uint32_t cnt = 0;
  for (int w = w0 + threadIdx.x; w < w1; w += blockDim.x) {
    uint32_t d = done[w], conv;
    if (mode == FIN_INIT) {
      conv = ~d & ~syn_fail[w] & kNibOnes;
      syn_fail[w] = 0;
    } else if (mode == FIN_ITER) {
      conv = ~d & ~(vn_fail[w] | cn_fail[w]) & kNibOnes;
      vn_fail[w] = 0;
      cn_fail[w] = 0;
    } else {
      conv = ~d & kNibOnes;
    }
    if (conv) {
      uint32_t sf = mode == FIN_LAST ? syn_fail[w] : 0;
      for (int j = 0; j < 8; ++j)
        if ((conv >> (4 * j)) & 1) iters[8 * w + j] = ((sf >> (4 * j)) & 1) ? -value : value;
    }
    if (mode == FIN_LAST) syn_fail[w] = 0;
    d |= conv;
    done[w] = d;
    cnt += __popc(~d & kNibOnes);
  }
  for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x < 32) {
    cnt = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0;
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
    if (threadIdx.x == 0) *n_active = cnt;
  }
  if (W_lists > 0) {
    __syncthreads();  // done[] of this CTA's writes is visible to the CTA
    build_lists(W_lists, done, cols, grp4, n_lists);
    if (mirror && threadIdx.x == 0) {  // build_lists ends with a barrier after thread 0 wrote the counts
      mirror[0] = n_lists[0];
      mirror[2] = n_lists[2];
      mirror[3] = n_lists[3];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Frame repacking.  The reference decodes one frame at a time (src/LDPC_Code_LUT.cpp:259-353), so a frame
// that returns early simply stops costing anything.  Here 8 frames share a word and a row of W words is
// the unit of every memory access: a finished frame keeps occupying its nibble, and a lone straggler
// keeps a whole 32-byte sector per edge alive.  When enough frames of a batch have returned, the
// unfinished ones are gathered -- order preserved -- into a narrower layout (W_new < W_old) and the
// remaining iterations run there.  Results of frames that finish in a repacked layout are moved back to
// the original layout (retire), so everything after the pass schedule still sees hard[nvar][W] / iters of
// the original batch.  Frames are independent: which word a frame sits in is not observable in its result.
//
// plan (single CTA): src[s'] = old slot of new slot s' (-1: padding), fid_new[s'] = original frame id,
// flags / work lists of the new layout.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_repack_plan(int W_old, const uint32_t* __restrict__ done_old,
                                                      const int32_t* __restrict__ fid_old, int W_new,
                                                      int32_t* __restrict__ src, int32_t* __restrict__ fid_new,
                                                      uint32_t* done_new, uint32_t* vn_fail, uint32_t* cn_fail,
                                                      uint32_t* syn_fail, int32_t* __restrict__ iters_new,
                                                      uint32_t* n_active_new, uint32_t* cols_new,
                                                      uint32_t* grp4_new) {
  __shared__ uint32_t psum[32];
  __shared__ uint32_t pbase;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S_new = 8 * W_new;
  if (threadIdx.x == 0) pbase = 0;
  __syncthreads();
  for (int c0 = 0; c0 < W_old; c0 += 1024) {
    const int w = c0 + threadIdx.x;
    const uint32_t lm = w < W_old ? (~done_old[w] & kNibOnes) : 0u;
    const uint32_t cnt = __popc(lm);
    uint32_t incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) psum[wid] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (int j = 0; j < 32; ++j) {
      const uint32_t v = psum[j];
      before += j < wid ? v : 0;
      total += v;
    }
    uint32_t at = pbase + before + incl - cnt;
    for (int j = 0; j < 8; ++j)
      if ((lm >> (4 * j)) & 1) {
        if (at < (uint32_t)S_new) {
          src[at] = 8 * w + j;
          fid_new[at] = fid_old ? fid_old[8 * w + j] : 8 * w + j;
        }
        ++at;
      }
    __syncthreads();
    if (threadIdx.x == 0) pbase += total;
    __syncthreads();
  }
  const int n_live = min((int)pbase, S_new);
  for (int s = n_live + threadIdx.x; s < S_new; s += 1024) {
    src[s] = -1;
    fid_new[s] = -1;
  }
  for (int s = threadIdx.x; s < S_new; s += 1024) iters_new[s] = 0;
  for (int w = threadIdx.x; w < W_new; w += 1024) {
    uint32_t d = 0;
    for (int j = 0; j < 8; ++j)
      if (8 * w + j >= n_live) d |= 1u << (4 * j);
    done_new[w] = d;
    vn_fail[w] = 0;
    cn_fail[w] = 0;
    syn_fail[w] = 0;
  }
  if (threadIdx.x == 0) {
    n_active_new[0] = (uint32_t)n_live;
    n_active_new[1] = 0;
  }
  __syncthreads();
  build_lists(W_new, done_new, cols_new, grp4_new, n_active_new);
}

// rows of msgs / cha: new word w' of row r = the nibbles of old slots src[8w' .. 8w'+7] of the same row.
// src is increasing, so the 8 sources of a thread lie in a few adjacent old words and neighbouring
// threads read neighbouring ranges: the old row is streamed once.
__global__ void __launch_bounds__(128) k_repack_rows(const int32_t* __restrict__ src,
                                                     const uint32_t* __restrict__ old_rows, int W_old,
                                                     uint32_t* __restrict__ new_rows, int W_new, int n_rows,
                                                     int rows_per_cta) {
  const int w = blockIdx.y * 128 + threadIdx.x;
  if (w >= W_new) return;
  const int4 sa = reinterpret_cast<const int4*>(src)[2 * w], sb = reinterpret_cast<const int4*>(src)[2 * w + 1];
  const int s[8] = {sa.x, sa.y, sa.z, sa.w, sb.x, sb.y, sb.z, sb.w};
  // padding slots (src < 0, only at the end of the layout) read word 0 and are masked out
  uint32_t off[8], sh[8], keep[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    off[j] = s[j] >= 0 ? (uint32_t)s[j] >> 3 : 0u;
    sh[j] = 4u * ((uint32_t)s[j] & 7u);
    keep[j] = s[j] >= 0 ? 0xFu : 0u;
  }
  const int r0 = blockIdx.x * rows_per_cta, r1 = min(n_rows, r0 + rows_per_cta);
  // Dense case (every thread of the warp finds its 8 sources inside 4 consecutive old words, i.e. at least a
  // quarter of the old slots is still in use): 4 loads per row and byte permutes.  The sources are the nibbles
  // p_j = s[j] - 8 * b0 of the 16-byte string x0 x1 x2 x3; PRMT fetches the bytes p_j >> 1 of the even (odd) output
  // nibbles into the four byte lanes, a masked shift moves each nibble to its half.  Old rows are read up to three
  // words past their end (the next row, or the slack the buffers carry) and those bytes are never selected.
  const uint32_t b0 = s[0] >= 0 ? (uint32_t)s[0] >> 3 : 0u;
  bool span_ok = true;
#pragma unroll
  for (int j = 0; j < 8; ++j) span_ok &= s[j] < 0 || ((uint32_t)s[j] >> 3) - b0 <= 3u;
  if (__all_sync(0xFFFFFFFFu, span_ok)) {
    uint32_t selA[2] = {0, 0}, selB[2] = {0, 0}, mB[2] = {0, 0}, mlo[2] = {0, 0}, mhi[2] = {0, 0};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (s[j] < 0) continue;
      const uint32_t pj = (uint32_t)s[j] - 8u * b0, q = pj >> 1, lane = (uint32_t)j >> 1, h = j & 1;
      if (q < 8) selA[h] |= q << (4 * lane);
      else {
        selB[h] |= (q - 8) << (4 * lane);
        mB[h] |= 0xFFu << (8 * lane);
      }
      // even outputs end in the low half of their byte lane, odd outputs in the high half
      const uint32_t want = h ? 0xF0u : 0x0Fu;
      if ((pj & 1) == (uint32_t)h) mlo[h] |= want << (8 * lane);  // already in the right half
      else mhi[h] |= want << (8 * lane);                           // needs the 4-bit shift
    }
    constexpr int R = 4;
    for (int r = r0; r < r1; r += R) {
      uint32_t x0[R], x1[R], x2[R], x3[R];
#pragma unroll
      for (int i = 0; i < R; ++i) {
        const uint32_t* p = old_rows + (size_t)min(r + i, r1 - 1) * W_old + b0;
        x0[i] = __ldg(p);
        x1[i] = __ldg(p + 1);
        x2[i] = __ldg(p + 2);
        x3[i] = __ldg(p + 3);
      }
#pragma unroll
      for (int i = 0; i < R; ++i) {
        const uint32_t Ae = __byte_perm(x0[i], x1[i], selA[0]), Be = __byte_perm(x2[i], x3[i], selB[0]);
        const uint32_t Ao = __byte_perm(x0[i], x1[i], selA[1]), Bo = __byte_perm(x2[i], x3[i], selB[1]);
        const uint32_t E = (Ae & ~mB[0]) | (Be & mB[0]), O = (Ao & ~mB[1]) | (Bo & mB[1]);
        const uint32_t out = (E & mlo[0]) | ((E >> 4) & mhi[0]) | (O & mlo[1]) | ((O << 4) & mhi[1]);
        if (r + i < r1) __stcs(new_rows + (size_t)(r + i) * W_new + w, out);
      }
    }
    return;
  }
  constexpr int R = 4;  // rows in flight per thread: 32 independent (mostly L1-resident) loads
  int r = r0;
  for (; r + R <= r1; r += R) {
    uint32_t x[R][8];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const uint32_t* row = old_rows + (size_t)(r + i) * W_old;
#pragma unroll
      for (int j = 0; j < 8; ++j) x[i][j] = __ldg(row + off[j]);
    }
#pragma unroll
    for (int i = 0; i < R; ++i) {
      uint32_t out = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) out |= ((x[i][j] >> sh[j]) & keep[j]) << (4 * j);
      __stcs(new_rows + (size_t)(r + i) * W_new + w, out);
    }
  }
  for (; r < r1; ++r) {
    const uint32_t* row = old_rows + (size_t)r * W_old;
    uint32_t out = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) out |= ((__ldg(row + off[j]) >> sh[j]) & keep[j]) << (4 * j);
    __stcs(new_rows + (size_t)r * W_new + w, out);
  }
}

// retire, step 1 (thread = slot of the repacked layout): lut_decode returns of the finished frames go to
// their original frame index; per ORIGINAL word w0 that has frames in this layout: the first slot they
// occupy (they are consecutive, order is preserved), which of its 8 frames are here (ow_in) and which of
// those have finished (ow_ret).  ow_ret is zero-filled before the launch.
__global__ void k_retire_plan(int W_c, const uint32_t* __restrict__ done_c, const int32_t* __restrict__ fid_c,
                              const int32_t* __restrict__ iters_c, int32_t* __restrict__ iters0,
                              int32_t* __restrict__ ow_first, uint8_t* __restrict__ ow_in,
                              uint8_t* __restrict__ ow_ret) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  const int S = 8 * W_c;
  if (s >= S) return;
  const int f = fid_c[s];
  if (f < 0) return;
  if ((done_c[s >> 3] >> (4 * (s & 7))) & 1) iters0[f] = iters_c[s];
  if (s > 0 && (fid_c[s - 1] >> 3) == (f >> 3)) return;
  const int w0 = f >> 3;
  uint32_t m_in = 0, m_ret = 0;
  for (int k = 0; k < 8 && s + k < S; ++k) {
    const int fk = fid_c[s + k];
    if (fk < 0 || (fk >> 3) != w0) break;
    const uint32_t bit = 1u << (fk & 7);
    m_in |= bit;
    if ((done_c[(s + k) >> 3] >> (4 * ((s + k) & 7))) & 1) m_ret |= bit;
  }
  ow_first[w0] = s;
  ow_in[w0] = (uint8_t)m_in;
  ow_ret[w0] = (uint8_t)m_ret;
}

// ordered list of the original words that have retiring frames (single CTA): neighbouring threads of the next
// kernel then touch neighbouring bytes of hard0
__global__ void __launch_bounds__(1024) k_retire_list(int W0, const uint8_t* __restrict__ ow_ret,
                                                      int32_t* __restrict__ ret_list, uint32_t* __restrict__ ret_count) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t base_s;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) base_s = 0;
  __syncthreads();
  for (int c0 = 0; c0 < W0; c0 += 1024) {
    const int i = c0 + threadIdx.x;
    const bool hit = i < W0 && ow_ret[i] != 0;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, hit);
    if (lane == 0) wsum[wid] = __popc(bal);
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (int j = 0; j < 32; ++j) {
      const uint32_t v = wsum[j];
      before += j < wid ? v : 0;
      total += v;
    }
    const uint32_t base = base_s;
    if (hit) ret_list[base + before + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (threadIdx.x == 0) base_s = base + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *ret_count = base_s;
}

// retire, step 2 (thread = (rows, listed original word)): hard bits of the finished frames back to hard0[nvar][W0]
__global__ void __launch_bounds__(128) k_retire_hard(int nvar, int W0, int W_c, const int32_t* __restrict__ ret_list,
                                                     const uint32_t* __restrict__ ret_count,
                                                     const int32_t* __restrict__ ow_first,
                                                     const uint8_t* __restrict__ ow_in,
                                                     const uint8_t* __restrict__ ow_ret,
                                                     const uint8_t* __restrict__ hard_c, uint8_t* __restrict__ hard0,
                                                     int rows_per_cta) {
  const int i = blockIdx.y * 128 + threadIdx.x;
  if (i >= (int)*ret_count) return;
  const int w0 = ret_list[i];
  const uint32_t m_ret = ow_ret[w0], m_in = ow_in[w0];
  const int s0 = ow_first[w0];
  const int byte0 = s0 >> 3, sh = s0 & 7;
  const int second = byte0 + 1 < W_c ? 1 : 0;
  // bit k of the shifted source = k-th frame of this word present in the layout -> bit position of that frame:
  // a bit-expand by m_in (Hacker's Delight 7-5), its three move masks are row-invariant
  uint32_t mv[3];
  {
    uint32_t m = m_in, mk = ~m << 1;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      uint32_t mp = mk ^ (mk << 1);
      mp ^= mp << 2;
      mp ^= mp << 4;
      mp ^= mp << 8;
      mp ^= mp << 16;
      mv[i] = mp & m;
      m = (m ^ mv[i]) | (mv[i] >> (1 << i));
      mk &= ~mp;
    }
  }
  const int v0 = blockIdx.x * rows_per_cta, v1 = min(nvar, v0 + rows_per_cta);
  constexpr int R = 8;
  for (int v = v0; v < v1; v += R) {
    uint32_t x[R], o[R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const int vv = min(v + k, v1 - 1);
      const uint8_t* hc = hard_c + (size_t)vv * W_c + byte0;
      x[k] = ((uint32_t)hc[0] | ((uint32_t)hc[second] << 8)) >> sh;
      o[k] = hard0[(size_t)vv * W0 + w0];
    }
#pragma unroll
    for (int k = 0; k < R; ++k) {
      uint32_t xx = x[k] & 0xFFu;
#pragma unroll
      for (int i = 2; i >= 0; --i) {
        const uint32_t t = xx << (1 << i);
        xx = (xx & ~mv[i]) | (t & mv[i]);
      }
      if (v + k < v1) hard0[(size_t)(v + k) * W0 + w0] = (uint8_t)((o[k] & ~m_ret) | (xx & m_ret));
    }
  }
}

// hard[nvar][W] -> out_bits[F][nvar]
template <bool VEC4>
__global__ void __launch_bounds__(256) k_unpack(const uint8_t* __restrict__ hard, int nvar, int F, int W,
                                                uint8_t* __restrict__ out) {
  __shared__ uint8_t t[32][36];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int v0 = blockIdx.x * 32, w0 = blockIdx.y * 32, f0 = w0 * 8;
  for (int vi = ty; vi < 32; vi += 8) {
    int v = v0 + vi, w = w0 + tx;
    t[vi][tx] = (v < nvar && w < W) ? hard[(size_t)v * W + w] : 0;
  }
  __syncthreads();
  if (VEC4) {  // nvar % 4 == 0 and a 4-byte aligned output: four nodes per store
    const int q4 = threadIdx.x & 7, fr = threadIdx.x >> 3;
    const int v = v0 + 4 * q4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int fl = fr + 32 * i, f = f0 + fl;
      if (f < F && v < nvar) {
        uint32_t x = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) x |= (uint32_t)((t[4 * q4 + k][fl >> 3] >> (fl & 7)) & 1) << (8 * k);
        *reinterpret_cast<uint32_t*>(out + (size_t)f * nvar + v) = x;
      }
    }
  } else {
    for (int i = 0; i < 32; ++i) {
      int fl = ty + 8 * i, f = f0 + fl, v = v0 + tx;
      if (f < F && v < nvar) out[(size_t)f * nvar + v] = (t[tx][fl >> 3] >> (fl & 7)) & 1;
    }
  }
}

// hard[nvar][W] -> out[F][NB] bit-packed rows, NB = ceil(nvar / 8), bit (v & 7) of byte v >> 3 = decision of node v
// (lutldpc_decode_packed: an eighth of the bytes of the one-byte-per-bit form).  Tile = 256 nodes x 32 words (256
// frames) through shared memory: reads are 32-byte sectors of a node row, writes 32-byte pieces of a frame row; in
// between every thread transposes 8 x 8 bit blocks (8 nodes x 8 frames).
__global__ void __launch_bounds__(256) k_unpack_bits(const uint8_t* __restrict__ hard, int nvar, int F, int W, int NB,
                                                     uint8_t* __restrict__ out) {
  __shared__ uint8_t t[256][33];  // [node of the tile][word of the tile]
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int v0 = blockIdx.x * 256, w0 = blockIdx.y * 32;
  for (int vi = ty; vi < 256; vi += 8) {
    const int v = v0 + vi, w = w0 + tx;
    t[vi][tx] = (v < nvar && w < W) ? hard[(size_t)v * W + w] : 0;
  }
  __syncthreads();
  // thread = (node group g of 8 nodes, word wl): 8 bytes in (one per node, bit j = frame j), 8 bytes out (one per frame)
  for (int item = threadIdx.x; item < 32 * 32; item += 256) {
    const int g = item & 31, wl = item >> 5;
    uint32_t lo = 0, hi = 0;  // nodes 0..3 and 4..7 of the group, one byte each
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      lo |= (uint32_t)t[8 * g + k][wl] << (8 * k);
      hi |= (uint32_t)t[8 * g + 4 + k][wl] << (8 * k);
    }
    // 8 x 8 bit transpose of the block m (bit 8r + c = node r, frame c): three masked swaps of ever larger sub-blocks
    unsigned long long m = (unsigned long long)lo | ((unsigned long long)hi << 32), tt;
    tt = 0x0F0F0F0F00000000ull & (m ^ (m << 28)); m ^= tt ^ (tt >> 28);
    tt = 0x3333000033330000ull & (m ^ (m << 14)); m ^= tt ^ (tt >> 14);
    tt = 0x5500550055005500ull & (m ^ (m << 7));  m ^= tt ^ (tt >> 7);
    // now byte j of m = frame j of the word: bit k = node k of the group
    const int gb = (v0 >> 3) + g;
    if (gb >= NB) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int f = 8 * (w0 + wl) + j;
      if (f < F) out[(size_t)f * NB + gb] = (uint8_t)(m >> (8 * j));
    }
  }
}

// all edge messages of one frame as reference labels (the verbosity>=2 dumps, :291-337)
__global__ void k_extract_frame(const uint32_t* __restrict__ msgs, int num_edges, int W, int frame, int Nq, int q,
                                uint8_t* __restrict__ out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= num_edges) return;
  uint32_t c = (msgs[(size_t)e * W + (frame >> 3)] >> (4 * (frame & 7))) & 0xF;
  uint32_t sign = (c >> (q - 1)) & 1, mag = c & ((1u << (q - 1)) - 1), nz = Nq >> 1;
  out[e] = (uint8_t)(sign ? nz - 1 - mag : nz + mag);
}

// quant_nonlin (src/common.cpp:120-138): number of leading thresholds with x > b[i]
__device__ __forceinline__ int quant_nonlin_dev(double x, const double* __restrict__ b, int nb) {
  int idx = 0;
  for (int i = 0; i < nb; ++i) {
    if (x > b[i]) ++idx;
    else break;
  }
  return idx;
}
// front end of LDPC_Code_LUT::decode (src/LDPC_Code_LUT.cpp:207-221)
__global__ void k_quantize(const double* __restrict__ llr, size_t n, const double* __restrict__ qb_cha, int n_qb_cha,
                           const double* __restrict__ qb_msg, int n_qb_msg, const int32_t* __restrict__ cha2msg,
                           int use_map, uint8_t* __restrict__ cha_lab, uint8_t* __restrict__ msg_lab) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double x = llr[i];
  int c = quant_nonlin_dev(x, qb_cha, n_qb_cha);
  cha_lab[i] = (uint8_t)c;
  msg_lab[i] = (uint8_t)(use_map ? cha2msg[c] : quant_nonlin_dev(x, qb_msg, n_qb_msg));
}

// ---------------------------------------------------------------------------------------------
// Specialised variable-node pass (vn_fast.cuh): balanced binary trees, 4-bit inputs.  Same contract
// as k_node<NODE_VAR>; the program is compile-time, values stay in registers, the LUTs of the class
// (128 B each, one per tree node) are staged into shared memory with one bulk copy.
// ---------------------------------------------------------------------------------------------
template <int DV>
__global__ void __launch_bounds__(kNodeThreads, (DV <= 8 ? 12 : 0)) k_vn_fast(const uint8_t* __restrict__ tables_g,
                                                           const int2* __restrict__ nodes, int n_nodes,
                                                           int nodes_per_cta, int W, uint32_t* __restrict__ msgs,
                                                           const uint32_t* __restrict__ cha, uint8_t* __restrict__ hard,
                                                           const uint32_t* __restrict__ done,
                                                           uint32_t* __restrict__ vn_fail, int want_flags, int q_out,
                                                           const uint32_t* __restrict__ cols,
                                                           const uint32_t* __restrict__ n_lists) {
  constexpr int NT = fastvn::Static<DV>::P.n_tables;
  __shared__ __align__(128) uint8_t tabs[NT * 128];
  __shared__ uint64_t bar;
  const int n_cols = (int)n_lists[2];
  if ((int)(blockIdx.y * kNodeThreads) >= n_cols) return;  // whole CTA beyond the live columns
  stage_to_smem(tabs, tables_g, NT * 128, &bar);
  const int i = blockIdx.y * kNodeThreads + threadIdx.x;
  if (i >= n_cols) return;
  const int w = (int)cols[i];
  const uint32_t dn = done[w];
  if (dn == kNibOnes) return;
  uint32_t failacc = 0;
  const int i0 = blockIdx.x * nodes_per_cta;
  vn_fast_body<DV>(tabs, nodes, i0, min(i0 + nodes_per_cta, n_nodes), W, w, msgs, cha, hard,
                   nib_bits_to_byte(dn, 0), want_flags, q_out, failacc);
  if (want_flags) flag_or(&vn_fail[w], failacc & kNibOnes);
}

static const fastvn::Prog* fast_prog(int dv) {
  switch (dv) {
#define X(D) case D: return &fastvn::Static<D>::P;
    LUTLDPC_FAST_DVS(X)
#undef X
  }
  return nullptr;
}

// All specialised degree classes of one VN pass in ONE launch.  A tile = a few nodes of one class; the
// host interleaves the tiles of the classes (FusedVnPlan), so CTAs that are bound by HBM latency
// (dv 2, 3) and CTAs that are bound by shared-memory look-ups (dv 8 ...) are resident on an SM together.
// Degrees up to kVnWideMaxDv run two word columns per thread (vn_fast_body2): thread i owns the column pair
// (i & 1) of live group grp4[i >> 1].
#ifndef VN_MIN_BLOCKS
#define VN_MIN_BLOCKS 8  // resident CTAs per SM the fused VN kernel is compiled for: 64 registers.  Measured on
                        // N=64800, 16384 frames per step: 12 (40 registers, 32 bytes of spills) 88.8 ms, 10 (48) 87.1,
                        // 8 (64) 86.2, 6 (80) 86.4, 5 (94) 87.1
#endif
constexpr int kVnWideMaxDv = 3;       // default: degrees that run two columns per thread
#ifndef VN_WIDE_COMPILED
#define VN_WIDE_COMPILED 3
#endif
constexpr int kVnWideCompiled = VN_WIDE_COMPILED;  // degrees the two-column body is compiled for
static int vn_wide_max_dv() {
  static const int v = [] {
    const char* e = getenv("LUTLDPC_VN_WIDE_MAXDV");
    int x = e ? atoi(e) : kVnWideMaxDv;
    return x < 0 ? 0 : (x > kVnWideCompiled ? kVnWideCompiled : x);
  }();
  return v;
}

template <int MAXDV>
__global__ void __launch_bounds__(kNodeThreads, (MAXDV <= 8 ? VN_MIN_BLOCKS : 0)) k_vn_fused(FusedVnParams P, const int4* __restrict__ tiles,
                                                            const int32_t* __restrict__ vn_edge0, int W,
                                                            uint32_t* __restrict__ msgs, const uint32_t* __restrict__ cha,
                                                            uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
                                                            uint32_t* __restrict__ vn_fail, int want_flags,
                                                            const uint32_t* __restrict__ cols,
                                                            const uint32_t* __restrict__ grp4,
                                                            const uint32_t* __restrict__ n_lists, int wide_max_dv) {
  __shared__ __align__(128) uint8_t tabs[(MAXDV - 1) * 128];
  __shared__ uint64_t bar;
  pdl_launch_dependents();
  const int4 tile = tiles[blockIdx.x];  // x = class, y = first node, z = node count
  const FusedVnClass cls = P.c[tile.x];
  const bool wide = cls.dv <= wide_max_dv;
  pdl_wait();
  const int n_items = wide ? 2 * (int)n_lists[3] : (int)n_lists[2];
  if ((int)(blockIdx.y * kNodeThreads) >= n_items) return;  // whole CTA beyond the live columns
  stage_to_smem(tabs, cls.tables, (uint32_t)(cls.dv - 1) * 128u, &bar);
  int i = blockIdx.y * kNodeThreads + threadIdx.x;
  int node0 = tile.y, node1 = tile.y + tile.z;
  if (n_items <= 64) {  // narrow layout: the thread groups of the CTA share the nodes of the tile
    const int tx_log2 = n_items <= 8 ? 3 : (n_items <= 16 ? 4 : (n_items <= 32 ? 5 : 6));
    const int nsub = kNodeThreads >> tx_log2, sub = threadIdx.x >> tx_log2;
    const int per = (tile.z + nsub - 1) / nsub;
    i = threadIdx.x & ((1 << tx_log2) - 1);
    node0 = tile.y + sub * per;
    node1 = min(tile.y + tile.z, node0 + per);
  }
  if (i >= n_items) return;
  if (wide) {
    const int w = (int)grp4[i >> 1] * 4 + (i & 1) * 2;
    const uint2 dn = *reinterpret_cast<const uint2*>(done + w);
    if (dn.x == kNibOnes && dn.y == kNibOnes) return;
    const uint32_t done_bytes = nib_bits_to_byte(dn.x, 0) | (nib_bits_to_byte(dn.y, 0) << 8);
    uint32_t f0 = 0, f1 = 0;
    switch (cls.dv) {
#define X(D)                                                                                                     \
  case D:                                                                                                        \
    if constexpr (D <= MAXDV && D <= kVnWideCompiled)                                                            \
      vn_fast_body2<D>(tabs, cls.nodes, node0, node1, W, w, msgs, cha, hard, done_bytes, want_flags, cls.q_out,  \
                       f0, f1);                                                                                  \
    break;
      LUTLDPC_FAST_DVS(X)
#undef X
    }
    if (want_flags) {
      flag_or(&vn_fail[w], f0 & kNibOnes);
      flag_or(&vn_fail[w + 1], f1 & kNibOnes);
    }
    return;
  }
  const int w = (int)cols[i];
  const uint32_t dn = done[w];
  if (dn == kNibOnes) return;
  const uint32_t done_byte = nib_bits_to_byte(dn, 0);
  uint32_t failacc = 0;
  switch (cls.dv) {
#define X(D)                                                                                                     \
  case D:                                                                                                        \
    if constexpr (D <= MAXDV)                                                                                    \
      vn_fast_body<D>(tabs, cls.nodes, node0, node1, W, w, msgs, cha, hard, done_byte,                          \
                      want_flags, cls.q_out, failacc);                                                                 \
    break;
    LUTLDPC_FAST_DVS(X)
#undef X
  }
  if (want_flags) flag_or(&vn_fail[w], failacc & kNibOnes);
}

// True when the program compiled from the decoder's trees is exactly the compile-time one.
bool fast_vn_matches(const NodeProgram& p) {
  const fastvn::Prog* P = fast_prog(p.n_in);
  if (!P || !p.has_cha || p.n_out != p.n_in || p.q_in != 4 || (int)p.ops.size() != P->n_ops) return false;
  if (p.q_out < 1 || p.q_out > 4) return false;
  for (int i = 0; i < P->n_ops; ++i) {
    const DevOp& d = p.ops[i];
    const fastvn::Op& f = P->ops[i];
    if (d.n_in != 2 || d.kind != OP_NARROW || !d.sym || d.chk || d.idx_bits != 7 || d.q[0] != 4 || d.q[1] != 4) return false;
    if (d.sh[0] != 0 || d.sh[1] != 4) return false;
    if (d.table_off != (uint32_t)f.node * 128u) return false;
    if (p.raw_in[i].size() != 2 || p.raw_in[i][0] != f.a || p.raw_in[i][1] != f.b) return false;
    if ((p.raw_node[i] == 0) != f.root) return false;
    if (d.q_out != (f.root ? p.q_out : 4)) return false;
  }
  for (int ii = 0; ii < p.n_in; ++ii)
    if (p.raw_out[ii] != P->out[ii]) return false;
  return (int)p.tables.size() >= P->n_tables * 128;
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
static constexpr int kMaxNodeSmem = 200 * 1024;

cudaError_t configure_kernels() {
  cudaError_t e;
  e = cudaFuncSetAttribute(k_node<NODE_VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxNodeSmem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_node<NODE_DEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxNodeSmem);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k_node<NODE_CHK>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxNodeSmem);
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// Launch with programmatic stream serialization (kdev.cuh: pdl_wait / pdl_launch_dependents).  LUTLDPC_PDL=0
// falls back to ordinary stream order.
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
  static const bool on = [] {
    const char* e = getenv("LUTLDPC_PDL");
    return !(e && e[0] == '0');
  }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

cudaError_t launch_pack(const uint8_t* d_cha_lab, const uint8_t* d_msg_lab, const GraphArrays& g,
                        const BatchBuffers& b, int Nq_cha, int Nq_msg, cudaStream_t st) {
  dim3 grid(cdiv(g.nvar, 32), cdiv(b.W, 32));
  const bool packed = d_msg_lab == nullptr;  // both labels of a symbol in one byte of d_cha_lab
  const bool vec4 = g.nvar % 4 == 0 && ((uintptr_t)d_cha_lab | (uintptr_t)d_msg_lab) % 4 == 0;
#define PACK_ARGS d_cha_lab, d_msg_lab, g.vn_edge0, g.nvar, b.F, b.W, Nq_cha, code_bits(Nq_cha), Nq_msg, code_bits(Nq_msg), \
                  b.cha, b.msgs, b.hard, b.msg0
  if (vec4 && packed) k_pack<true, true><<<grid, 256, 0, st>>>(PACK_ARGS);
  else if (vec4) k_pack<true, false><<<grid, 256, 0, st>>>(PACK_ARGS);
  else if (packed) k_pack<false, true><<<grid, 256, 0, st>>>(PACK_ARGS);
  else k_pack<false, false><<<grid, 256, 0, st>>>(PACK_ARGS);
#undef PACK_ARGS
  return cudaGetLastError();
}

cudaError_t launch_pack_bits(const uint8_t* d_bits, const GraphArrays& g, const BatchBuffers& b, cudaStream_t st) {
  dim3 grid(cdiv(g.nvar, 32), cdiv(b.W, 32));
  k_pack_bits<<<grid, 256, 0, st>>>(d_bits, g.nvar, b.F, b.W, b.hard);
  return cudaGetLastError();
}

cudaError_t launch_reset_flags(const BatchBuffers& b, cudaStream_t st) {
  k_reset_flags<<<cdiv(b.W, 128), 128, 0, st>>>(b.done, b.vn_fail, b.cn_fail, b.syn_fail, b.iters, b.n_active, b.W,
                                                 b.F);
  k_build_lists<<<1, 1024, 0, st>>>(b.W, b.done, b.cols, b.grp4, b.n_active);
  return cudaGetLastError();
}

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

cudaError_t launch_cn_minsum(const GraphArrays& g, const BatchBuffers& b, int q, int want_fail, cudaStream_t st) {
  static const int vec = env_int("LUTLDPC_CN_VEC", 2);  // words per thread (tuning knob, 0 = scalar kernel)
#define CN_ARGS g.cn_start, g.cn_msg_idx, g.nchk, b.W, q, b.msgs, b.done, b.cn_fail, want_fail
  if (vec == 2 || vec == 4) {
    dim3 grid(g.nchk, cdiv(b.W, 128 * vec));
#define CN_V(M)                                                        \
  if (vec == 2) k_cn_minsum_v<M, 2><<<grid, 128, 0, st>>>(CN_ARGS);    \
  else k_cn_minsum_v<M, 4><<<grid, 128, 0, st>>>(CN_ARGS);
    if (g.max_dc <= 8) { CN_V(8) return cudaGetLastError(); }
    if (g.max_dc <= 10) { CN_V(10) return cudaGetLastError(); }
    if (g.max_dc <= 16) { CN_V(16) return cudaGetLastError(); }
#undef CN_V
  }
  dim3 grid(g.nchk, cdiv(b.W, 128));
  if (g.max_dc <= 8) k_cn_minsum<8><<<grid, 128, 0, st>>>(CN_ARGS);
  else if (g.max_dc <= 12) k_cn_minsum<12><<<grid, 128, 0, st>>>(CN_ARGS);
  else if (g.max_dc <= 16) k_cn_minsum<16><<<grid, 128, 0, st>>>(CN_ARGS);
  else if (g.max_dc <= 32) k_cn_minsum<32><<<grid, 128, 0, st>>>(CN_ARGS);
  else k_cn_minsum_any<<<grid, 128, 0, st>>>(CN_ARGS);
#undef CN_ARGS
  return cudaGetLastError();
}

bool cn_exact_supported(int dc) { return dc >= 2 && dc <= 16; }

cudaError_t launch_cn_exact(int dc, const int32_t* d_rows, int n_checks, const BatchBuffers& b, int q, int want_fail,
                            cudaStream_t st) {
  if (n_checks == 0) return cudaSuccess;
  static const int vec = env_int("LUTLDPC_CN_VEC", 2) == 4 ? 4 : 2;
  if (b.host_grp4 <= 0) return cudaSuccess;
  const int items = b.host_grp4 * (4 / vec);
  const int tx_log2 = items <= 8 ? 3 : (items <= 16 ? 4 : (items <= 32 ? 5 : (items <= 64 ? 6 : 7)));
  // checks per thread in sequence.  One check per CTA column block means hundreds of thousands of CTAs that live
  // for a few microseconds; measured on N=64800, 16384 frames: 1 -> 0.705 ms per check pass, 2 -> 0.658, 4 -> 0.642
  // (0.94 of the measured HBM peak), 8 -> 0.640, 16 -> 0.650.  Narrow layouts keep 1 (their CTAs already share checks).
  static const int per_env = env_int("LUTLDPC_CN_PER", 4);
  const int per = items >= 128 ? (per_env < 1 ? 1 : per_env) : 1;
  dim3 grid(cdiv(n_checks, (128 >> tx_log2) * per), cdiv(items, 1 << tx_log2));
#define CN_X_ARGS d_rows, b.W, q, b.msgs, b.done, b.cn_fail, want_fail, b.grp4, b.n_active, n_checks, tx_log2, per
  cudaError_t le = cudaSuccess;
  switch (dc) {
#define X(D)                                                              \
  case D:                                                                 \
    if (vec == 2) le = launch_pdl(k_cn_exact<D, 2>, grid, dim3(128), st, CN_X_ARGS);      \
    else le = launch_pdl(k_cn_exact<D, 4>, grid, dim3(128), st, CN_X_ARGS);               \
    break;
    LUTLDPC_EXACT_DCS(X)
#undef X
    default:
      return cudaErrorInvalidValue;
  }
  return le != cudaSuccess ? le : cudaGetLastError();
}

cudaError_t launch_node(NodeMode mode, const NodeLaunch& nl, const GraphArrays& g, const BatchBuffers& b,
                        int want_flags, cudaStream_t st) {
  if (nl.n_nodes == 0) return cudaSuccess;
  size_t smem = (size_t)nl.prog_bytes + (size_t)nl.n_slots * kNodeThreads * sizeof(uint32_t);
  if (smem > (size_t)kMaxNodeSmem) return cudaErrorInvalidValue;
  if (b.host_cols <= 0) return cudaSuccess;
  int ycols = cdiv(b.host_cols, kNodeThreads);
  // enough CTAs to fill 148 SMs several times over, few enough that the program staging amortises
  int per = (int)(((long long)nl.n_nodes * ycols) / (148 * 24));
  per = per < 1 ? 1 : (per > 32 ? 32 : per);
  dim3 grid(cdiv(nl.n_nodes, per), ycols);
  uint32_t* fail = mode == NODE_CHK ? b.cn_fail : b.vn_fail;
  const int32_t* start = mode == NODE_CHK ? g.cn_start : g.vn_edge0;
#define NODE_ARGS nl.prog, nl.prog_bytes, nl.node_list, nl.n_nodes, per, start, g.cn_msg_idx, b.W, b.msgs, b.cha, \
                  b.hard, b.done, fail, want_flags, b.cols, b.n_active
  if (mode == NODE_VAR) k_node<NODE_VAR><<<grid, kNodeThreads, smem, st>>>(NODE_ARGS);
  else if (mode == NODE_DEC) k_node<NODE_DEC><<<grid, kNodeThreads, smem, st>>>(NODE_ARGS);
  else k_node<NODE_CHK><<<grid, kNodeThreads, smem, st>>>(NODE_ARGS);
#undef NODE_ARGS
  return cudaGetLastError();
}

cudaError_t launch_vn_fast(int dv, const uint8_t* d_tables, int q_out, const NodeLaunch& nl, const GraphArrays& g,
                           const BatchBuffers& b, int want_flags, cudaStream_t st) {
  if (nl.n_nodes == 0) return cudaSuccess;
  if (b.host_cols <= 0) return cudaSuccess;
  int ycols = cdiv(b.host_cols, kNodeThreads);
  int per = (int)(((long long)nl.n_nodes * ycols) / (148 * 24));
  per = per < 1 ? 1 : (per > 32 ? 32 : per);
  dim3 grid(cdiv(nl.n_nodes, per), ycols);
  switch (dv) {
#define X(D)                                                                                                       \
  case D:                                                                                                          \
    k_vn_fast<D><<<grid, kNodeThreads, 0, st>>>(d_tables, nl.node_pairs, nl.n_nodes, per, b.W, b.msgs,            \
                                                b.cha, b.hard, b.done, b.vn_fail, want_flags, q_out, b.cols,      \
                                                b.n_active);                                                       \
    break;
    LUTLDPC_FAST_DVS(X)
#undef X
    default:
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_vn_fused(const FusedVnParams& P, int max_dv, const int4* d_tiles, int n_tiles, const GraphArrays& g,
                            const BatchBuffers& b, int want_flags, cudaStream_t st) {
  if (n_tiles == 0 || b.host_cols <= 0) return cudaSuccess;
  const int wide = vn_wide_max_dv();
  dim3 grid(n_tiles, cdiv(std::max(b.host_cols, wide >= 2 ? 2 * b.host_grp4 : 0), kNodeThreads));
#define FUSED_ARGS P, d_tiles, g.vn_edge0, b.W, b.msgs, b.cha, b.hard, b.done, b.vn_fail, want_flags, b.cols, b.grp4, \
                   b.n_active, wide
  cudaError_t le;
  if (max_dv <= 4) le = launch_pdl(k_vn_fused<4>, grid, dim3(kNodeThreads), st, FUSED_ARGS);
  else if (max_dv <= 8) le = launch_pdl(k_vn_fused<8>, grid, dim3(kNodeThreads), st, FUSED_ARGS);
  else if (max_dv <= 12) le = launch_pdl(k_vn_fused<12>, grid, dim3(kNodeThreads), st, FUSED_ARGS);
  else if (max_dv <= 20) le = launch_pdl(k_vn_fused<20>, grid, dim3(kNodeThreads), st, FUSED_ARGS);
  else return cudaErrorInvalidValue;
#undef FUSED_ARGS
  return le != cudaSuccess ? le : cudaGetLastError();
}

cudaError_t launch_syndrome(const GraphArrays& g, const BatchBuffers& b, cudaStream_t st) {
  if (b.host_cols <= 0) return cudaSuccess;
  if (b.host_grp4 >= 32) {  // wide layouts: 4 columns per thread
    const int per = 8;  // checks per CTA
    dim3 grid4(cdiv(g.nchk, per), cdiv(b.host_grp4, 128));
    k_syndrome4<<<grid4, 128, 0, st>>>(g.cn_start, g.chk_equ_idx, b.W, b.hard, b.done, b.syn_fail, b.grp4, b.n_active,
                                       g.nchk, per);
    return cudaGetLastError();
  }
  dim3 grid(g.nchk, cdiv(b.host_cols, 128));
  k_syndrome<<<grid, 128, 0, st>>>(g.cn_start, g.chk_equ_idx, b.W, b.hard, b.done, b.syn_fail, b.cols, b.n_active);
  return cudaGetLastError();
}

cudaError_t launch_finalize(FinalizeMode mode, int value, const BatchBuffers& b, cudaStream_t st) {
  return launch_finalize_range(mode, value, b, 0, b.W, b.n_active, st);
}

cudaError_t launch_finalize_range(FinalizeMode mode, int value, const BatchBuffers& b, int w0, int w1,
                                  uint32_t* n_active, cudaStream_t st) {
  const int full = w0 == 0 && w1 == b.W;
  cudaError_t le = launch_pdl(k_finalize, dim3(1), dim3(1024), st, (int)mode, value, w0, w1, b.done, b.vn_fail,
                              b.cn_fail, b.syn_fail, b.iters, n_active, full ? b.W : 0, b.cols, b.grp4, b.n_active,
                              full ? b.mirror : (uint32_t*)nullptr);
  return le != cudaSuccess ? le : cudaGetLastError();
}

cudaError_t launch_repack(const GraphArrays& g, const BatchBuffers& from, const BatchBuffers& to, int32_t* d_src,
                          cudaStream_t st) {
  k_repack_plan<<<1, 1024, 0, st>>>(from.W, from.done, from.fid, to.W, d_src, to.fid, to.done, to.vn_fail, to.cn_fail,
                                    to.syn_fail, to.iters, to.n_active, to.cols, to.grp4);
  const int rows_per = 32;
  dim3 ge(cdiv(g.num_edges, rows_per), cdiv(to.W, 128));
  k_repack_rows<<<ge, 128, 0, st>>>(d_src, from.msgs, from.W, to.msgs, to.W, g.num_edges, rows_per);
  dim3 gn(cdiv(g.nvar, rows_per), cdiv(to.W, 128));
  k_repack_rows<<<gn, 128, 0, st>>>(d_src, from.cha, from.W, to.cha, to.W, g.nvar, rows_per);
  cudaError_t e = cudaMemsetAsync(to.hard, 0, (size_t)g.nvar * to.W, st);
  return e != cudaSuccess ? e : cudaGetLastError();
}

cudaError_t launch_retire(const GraphArrays& g, const BatchBuffers& from, const BatchBuffers& orig, int32_t* d_ow_first,
                          uint8_t* d_ow_mask, cudaStream_t st) {
  // d_ow_first: [2 * W0 + 1] ints = first slot per original word | list of original words | count
  // d_ow_mask:  [2 * W0] bytes    = frames present | frames retiring
  int32_t* d_list = d_ow_first + orig.W;
  uint32_t* d_count = reinterpret_cast<uint32_t*>(d_ow_first + 2 * orig.W);
  uint8_t *d_ow_in = d_ow_mask, *d_ow_ret = d_ow_mask + orig.W;
  cudaError_t e = cudaMemsetAsync(d_ow_ret, 0, (size_t)orig.W, st);
  if (e != cudaSuccess) return e;
  k_retire_plan<<<cdiv(8 * from.W, 256), 256, 0, st>>>(from.W, from.done, from.fid, from.iters, orig.iters, d_ow_first,
                                                      d_ow_in, d_ow_ret);
  k_retire_list<<<1, 1024, 0, st>>>(orig.W, d_ow_ret, d_list, d_count);
  const int rows_per = 64;
  const int max_words = std::min(orig.W, 8 * from.W);
  dim3 grid(cdiv(g.nvar, rows_per), cdiv(max_words, 128));
  k_retire_hard<<<grid, 128, 0, st>>>(g.nvar, orig.W, from.W, d_list, d_count, d_ow_first, d_ow_in, d_ow_ret, from.hard,
                                      orig.hard, rows_per);
  return cudaGetLastError();
}

cudaError_t launch_unpack(const GraphArrays& g, const BatchBuffers& b, uint8_t* d_out_bits, cudaStream_t st) {
  dim3 grid(cdiv(g.nvar, 32), cdiv(b.W, 32));
  if (g.nvar % 4 == 0 && (uintptr_t)d_out_bits % 4 == 0) k_unpack<true><<<grid, 256, 0, st>>>(b.hard, g.nvar, b.F, b.W, d_out_bits);
  else k_unpack<false><<<grid, 256, 0, st>>>(b.hard, g.nvar, b.F, b.W, d_out_bits);
  return cudaGetLastError();
}

cudaError_t launch_unpack_bits(const GraphArrays& g, const BatchBuffers& b, uint8_t* d_out_packed, cudaStream_t st) {
  dim3 grid(cdiv(g.nvar, 256), cdiv(b.W, 32));
  k_unpack_bits<<<grid, 256, 0, st>>>(b.hard, g.nvar, b.F, b.W, (g.nvar + 7) / 8, d_out_packed);
  return cudaGetLastError();
}

cudaError_t launch_extract_frame(const BatchBuffers& b, int num_edges, int frame, int Nq, uint8_t* d_out,
                                 cudaStream_t st) {
  k_extract_frame<<<cdiv(num_edges, 256), 256, 0, st>>>(b.msgs, num_edges, b.W, frame, Nq, code_bits(Nq), d_out);
  return cudaGetLastError();
}

cudaError_t launch_quantize(const double* d_llr, size_t n, const double* d_qb_cha, int n_qb_cha,
                            const double* d_qb_msg, int n_qb_msg, const int32_t* d_cha2msg, int use_map,
                            uint8_t* d_cha_lab, uint8_t* d_msg_lab, cudaStream_t st) {
  unsigned blocks = (unsigned)((n + 255) / 256);
  k_quantize<<<blocks, 256, 0, st>>>(d_llr, n, d_qb_cha, n_qb_cha, d_qb_msg, n_qb_msg, d_cha2msg, use_map, d_cha_lab,
                                     d_msg_lab);
  return cudaGetLastError();
}

}  // namespace lutldpc
