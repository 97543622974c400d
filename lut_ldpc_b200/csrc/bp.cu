// BP comparison decoder: what ber_sim runs for a [BP] section (src/LDPC_BER_Sim.cpp:74-78,157-244) -- IT++'s
// LDPC_Code::bp_decode (itpp/itpp/comm/ldpc.cpp:1398-1622) over integer LLRs with the LLR_calc_unit of the reference's
// vendored IT++ (itpp/itpp/comm/llr.h:120-330, llr.cpp:74-185) -- for a batch of frames.
//
// Layout: one int32 per edge and frame, msg[e][s] with the frame slot s contiguous, so a warp is 32 frames of one
// node and every access is a coalesced 128-byte row piece.  The reference keeps two arrays (mvc, mcv); every edge has
// exactly one of each and a pass reads one and writes the other, so both share one array that the passes update in
// place.  Edges are numbered VN-major like the LUT decoder's (variable i owns off[i]..off[i+1]-1 in ascending check
// order = mvc[i + k*nvar]); a check reads its edges in ascending variable order (= jind).  Hard decisions live as one
// bit per slot (hardw[i][s/32], written by warp ballot), which is also the layout of the bit-packed encoder.
//
// Frames that have returned stop being updated (done bit per slot); once at most half of the slots are still running
// the unfinished frames are gathered into a narrower batch and the schedule continues there.
//
// Arithmetic is bit-identical to the reference: same logexp table (built on the host with the same libm calls), same
// saturation, and per node degree the same association of Boxplus / qsum operations.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "decoder.hpp"
#include "graph.hpp"
#include "rng.cuh"

using namespace lutldpc;

namespace {

constexpr int kQllrMaxConst = INT32_MAX >> 4;  // QLLR_MAX, itpp/itpp/comm/llr.h:61
constexpr int kBpThreads = 128;
constexpr int kCnGroups = 4;   // groups of checks a block of the by-degree check kernel walks through
constexpr int kVnGroups = 1;   // the same for the by-degree variable kernel (more than one group measured no gain)
constexpr int kCnRegsMax = 32;  // check degrees up to this run out of registers, larger ones out of shared memory

struct Calc {
  int d1, d2, d3, qmax;  // Dint1..3, qllr_max = 2^(Dint4-1) - 1
  double scale, qmax_d;  // 2^Dint1, to_double(qllr_max)
  int four, two;         // 4 and 2 as run-time values: multiplications by them stay on the FMA pipe (see bp_logexp)
};

// LLR_calc_unit::to_qllr, llr.h:266-279 (positive overflow returns QLLR_MAX, not qllr_max)
__device__ __forceinline__ int bp_to_qllr(double l, const Calc& c) {
  if (l > c.qmax_d) return kQllrMaxConst;
  if (l < -c.qmax_d) return -c.qmax;
  return __double2int_rd(0.5 + c.scale * l);
}
__device__ __forceinline__ int bp_clamp(int v, int qmax) { return max(-qmax, min(qmax, v)); }
// qsum, llr.h:320-329
__device__ __forceinline__ int bp_qsum(int a, int b, int qmax) { return bp_clamp(a + b, qmax); }
// logexp, llr.h:282-298 (x >= 0): 0 beyond the table -- tab[d2] holds that 0, so the index is clamped, not tested.
// tab is the table's shared-memory address; the byte address is formed with a multiply-add (FMA pipe) because the
// integer ALU pipe is what bounds the check pass.
__device__ __forceinline__ int bp_logexp(int x, uint32_t tab, const Calc& c) {
  const int idx = min(x >> c.d3, c.d2);
  uint32_t addr;
  int v;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(addr) : "r"(idx), "r"(c.four), "r"(tab));
  asm("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
// Boxplus, llr.cpp:148-185: term1 + logexp(|a+b|) - logexp(|a-b|), saturated, with term1 = sign(a) sign(b) min(|a|,|b|)
// (the reference's nested selects give 0 when a or b is 0, like the product).  {|a+b|, |a-b|} = {|a|+|b|, ||a|-|b||},
// in this order when the signs agree and swapped when they differ, so the value is sign(a) sign(b) * r with
// r = min + logexp(|a|+|b|) - logexp(max-min); the saturation is symmetric, so it commutes with the sign.  For a or b
// = 0 both forms give logexp(max) - logexp(max) = 0.
// FAST: a table is in use and qllr_max >= table[0]; then r >= min - table[0] >= -qllr_max and the lower saturation
// cannot trigger.  The sign is applied by a multiplication (FMA pipe again).
template <bool FAST>
__device__ __forceinline__ int bp_boxplus(int a, int b, uint32_t tab, const Calc& c) {
  const int aa = abs(a), bb = abs(b);
  const int mn = min(aa, bb), mx = max(aa, bb);
  int r = mn;
  if (FAST || c.d2 != 0) r += bp_logexp(aa + bb, tab, c) - bp_logexp(mx - mn, tab, c);
  r = FAST ? min(r, c.qmax) : bp_clamp(r, c.qmax);
  const int sg = (a ^ b) >> 31;  // 0 or -1
  int sgn, res;
  asm("mad.lo.s32 %0, %1, %2, 1;" : "=r"(sgn) : "r"(sg), "r"(c.two));
  asm("mul.lo.s32 %0, %1, %2;" : "=r"(res) : "r"(r), "r"(sgn));
  return res;
}

__device__ __forceinline__ bool slot_done(const uint32_t* donew, size_t s) { return (donew[s >> 5] >> (s & 31)) & 1u; }

// ---- inputs: frame-major host layout [F][N] -> qin[N][S] (slot-contiguous), through a 32x32 shared-memory tile
template <class T>
__global__ void k_bp_load(const T* __restrict__ in, int F, int N, size_t S, int32_t* __restrict__ qin, Calc c) {
  __shared__ int32_t tile[32][33];
  const int i0 = blockIdx.x * 32, f0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
  for (int r = ty; r < 32; r += 8) {
    const int f = f0 + r, i = i0 + tx;
    int32_t v = 0;
    if (f < F && i < N) {
      if constexpr (sizeof(T) == 8) v = bp_to_qllr((double)in[(size_t)f * N + i], c);
      else v = (int32_t)in[(size_t)f * N + i];
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r;
    if (i < N) qin[(size_t)i * S + f0 + tx] = tile[tx][r];
  }
}

// initial step (ldpc.cpp:1422-1428): every edge of variable i carries LLRin(i); hard decision of the input for the
// initial syndrome check (:1408); soft output = input in case that check passes
__global__ void __launch_bounds__(kBpThreads) k_bp_init(const int32_t* __restrict__ vn_off, int N, size_t S,
                                                         const int32_t* __restrict__ qin, int32_t* __restrict__ msg,
                                                         uint32_t* __restrict__ hardw, int32_t* __restrict__ soft) {
  const int i = blockIdx.x;
  const size_t s = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  const int32_t q = qin[(size_t)i * S + s];
  for (int e = vn_off[i]; e < vn_off[i + 1]; ++e) msg[(size_t)e * S + s] = q;
  if (soft) soft[(size_t)i * S + s] = q;
  const uint32_t bal = __ballot_sync(0xFFFFFFFFu, q < 0);
  if ((threadIdx.x & 31) == 0) hardw[(size_t)i * (S >> 5) + (s >> 5)] = bal;
}

// ---- step 1, check to variable (ldpc.cpp:1436-1556).  Degrees 2..5 of the reference's switch compute exactly the
// general left/right recursion (:1529-1553); degree 6 (:1497-1521) pairs its inputs differently and is kept as is.
// With suf[k] = m[k] [+] m[k+1] [+] ... [+] m[dc-1] folded from the right (the reference's mr[dc-1-k]) and the running
// left fold l (its ml[k-1]): out[0] = suf[1], out[k] = l [+] suf[k+1], out[dc-1] = l.
// Checks are grouped by degree on the host (list = the checks of one degree), so that the common degrees run fully
// unrolled out of registers without a predicate per input.  Row e of the message array starts e * S4 bytes after the
// slot's column base: one 32 x 32 -> 64 bit multiply-add per access.
__device__ __forceinline__ int32_t* bp_row(int32_t* col, uint32_t e, uint32_t S4) {
  return reinterpret_cast<int32_t*>(reinterpret_cast<char*>(col) + (unsigned long long)e * S4);
}

// cls_edge = the edge ids of the class's checks, DC per check, back to back.  A block takes CPB checks: their
// CPB * DC edge ids are staged in shared memory with one coalesced load, then every thread has all CPB * DC message
// loads of its slot in flight before the first Boxplus -- the pass is bound by memory latency otherwise.
template <int DC, int CPB, bool FAST>
__global__ void __launch_bounds__(kBpThreads) k_bp_cn_deg(const int32_t* __restrict__ cls_edge, int n_list, uint32_t S4,
                                                           int32_t* __restrict__ msg, const uint32_t* __restrict__ donew, Calc c,
                                                           const int32_t* __restrict__ table) {
  extern __shared__ int s_tab[];
  __shared__ uint32_t s_idx_all[kCnGroups * CPB * DC];
  const int tb0 = blockIdx.x * (kCnGroups * CPB);
  for (int i = threadIdx.x; i < kCnGroups * CPB * DC; i += kBpThreads)
    s_idx_all[i] = tb0 * DC + i < n_list * DC ? (uint32_t)cls_edge[(size_t)tb0 * DC + i] : 0u;
  for (int i = threadIdx.x; i <= c.d2; i += kBpThreads) s_tab[i] = i < c.d2 ? table[i] : 0;
  __syncthreads();
  const uint32_t tb = (uint32_t)__cvta_generic_to_shared(s_tab);
  const size_t s = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  if (slot_done(donew, s)) return;
  int32_t* col = msg + s;
  for (int g = 0; g < kCnGroups; ++g) {  // the table and the edge ids are staged once for kCnGroups groups of CPB checks
  const int t0 = tb0 + g * CPB;
  if (t0 >= n_list) break;
  const uint32_t* s_idx = s_idx_all + g * CPB * DC;
  int m[CPB][DC];
#pragma unroll
  for (int q = 0; q < CPB; ++q)
#pragma unroll
    for (int k = 0; k < DC; ++k) m[q][k] = *bp_row(col, s_idx[q * DC + k], S4);  // tail checks read row 0: harmless
#pragma unroll
  for (int q = 0; q < CPB; ++q) {
    if (t0 + q >= n_list) break;
    const uint32_t* e = s_idx + q * DC;
    if constexpr (DC == 6) {
      const int m01 = bp_boxplus<FAST>(m[q][0], m[q][1], tb, c), m23 = bp_boxplus<FAST>(m[q][2], m[q][3], tb, c),
                m45 = bp_boxplus<FAST>(m[q][4], m[q][5], tb, c);
      const int m03 = bp_boxplus<FAST>(m01, m23, tb, c), m25 = bp_boxplus<FAST>(m23, m45, tb, c),
                m0145 = bp_boxplus<FAST>(m01, m45, tb, c);
      *bp_row(col, e[0], S4) = bp_boxplus<FAST>(m[q][1], m25, tb, c);
      *bp_row(col, e[1], S4) = bp_boxplus<FAST>(m[q][0], m25, tb, c);
      *bp_row(col, e[2], S4) = bp_boxplus<FAST>(m0145, m[q][3], tb, c);
      *bp_row(col, e[3], S4) = bp_boxplus<FAST>(m0145, m[q][2], tb, c);
      *bp_row(col, e[4], S4) = bp_boxplus<FAST>(m03, m[q][5], tb, c);
      *bp_row(col, e[5], S4) = bp_boxplus<FAST>(m03, m[q][4], tb, c);
    } else {
      int suf[DC];
      suf[DC - 1] = m[q][DC - 1];
#pragma unroll
      for (int k = DC - 2; k >= 1; --k) suf[k] = bp_boxplus<FAST>(suf[k + 1], m[q][k], tb, c);
      *bp_row(col, e[0], S4) = suf[1];
      int l = m[q][0];
#pragma unroll
      for (int k = 1; k <= DC - 2; ++k) {
        *bp_row(col, e[k], S4) = bp_boxplus<FAST>(l, suf[k + 1], tb, c);
        l = bp_boxplus<FAST>(l, m[q][k], tb, c);
      }
      *bp_row(col, e[DC - 1], S4) = l;
    }
  }
  }
}

// the same for any degree up to the reference's max_cnd = 200: inputs and right folds in shared memory
__global__ void __launch_bounds__(kBpThreads) k_bp_cn_big(const int32_t* __restrict__ list, int n_list, int cpb,
                                                           const int32_t* __restrict__ cn_off, const int32_t* __restrict__ cn_edge,
                                                           int maxdc, uint32_t S4, int32_t* __restrict__ msg,
                                                           const uint32_t* __restrict__ donew, Calc c,
                                                           const int32_t* __restrict__ table) {
  extern __shared__ int s_all[];
  int* s_tab = s_all;
  int* s_m = s_all + c.d2 + 1 + threadIdx.x;  // [k * blockDim.x]
  int* s_suf = s_m + (size_t)maxdc * blockDim.x;
  const int T = blockDim.x;
  for (int i = threadIdx.x; i <= c.d2; i += T) s_tab[i] = i < c.d2 ? table[i] : 0;
  __syncthreads();
  const uint32_t tb = (uint32_t)__cvta_generic_to_shared(s_tab);
  const size_t s = (size_t)blockIdx.y * T + threadIdx.x;
  if (slot_done(donew, s)) return;
  int32_t* col = msg + s;
  const int t1 = min(n_list, (int)(blockIdx.x + 1) * cpb);
  for (int t = blockIdx.x * cpb; t < t1; ++t) {
    const int j = list[t], e0 = cn_off[j], dc = cn_off[j + 1] - e0;
    const uint32_t* idx = reinterpret_cast<const uint32_t*>(cn_edge + e0);
    for (int k = 0; k < dc; ++k) s_m[k * T] = *bp_row(col, idx[k], S4);
    if (dc == 6) {
      const int m0 = s_m[0], m1 = s_m[T], m2 = s_m[2 * T], m3 = s_m[3 * T], m4 = s_m[4 * T], m5 = s_m[5 * T];
      const int m01 = bp_boxplus<false>(m0, m1, tb, c), m23 = bp_boxplus<false>(m2, m3, tb, c), m45 = bp_boxplus<false>(m4, m5, tb, c);
      const int m03 = bp_boxplus<false>(m01, m23, tb, c), m25 = bp_boxplus<false>(m23, m45, tb, c),
                m0145 = bp_boxplus<false>(m01, m45, tb, c);
      *bp_row(col, idx[0], S4) = bp_boxplus<false>(m1, m25, tb, c);
      *bp_row(col, idx[1], S4) = bp_boxplus<false>(m0, m25, tb, c);
      *bp_row(col, idx[2], S4) = bp_boxplus<false>(m0145, m3, tb, c);
      *bp_row(col, idx[3], S4) = bp_boxplus<false>(m0145, m2, tb, c);
      *bp_row(col, idx[4], S4) = bp_boxplus<false>(m03, m5, tb, c);
      *bp_row(col, idx[5], S4) = bp_boxplus<false>(m03, m4, tb, c);
      continue;
    }
    int r = s_m[(dc - 1) * T];
    s_suf[(dc - 1) * T] = r;
    for (int k = dc - 2; k >= 1; --k) {
      r = bp_boxplus<false>(r, s_m[k * T], tb, c);
      s_suf[k * T] = r;
    }
    *bp_row(col, idx[0], S4) = s_suf[T];
    int l = s_m[0];
    for (int k = 1; k <= dc - 2; ++k) {
      *bp_row(col, idx[k], S4) = bp_boxplus<false>(l, s_suf[(k + 1) * T], tb, c);
      l = bp_boxplus<false>(l, s_m[k * T], tb, c);
    }
    *bp_row(col, idx[dc - 1], S4) = l;
  }
}

// ---- step 2, variable to check (ldpc.cpp:1558-1611), soft output and its sign.  Degree 3 (:1583-1595) is the same
// arithmetic as the differential update (:1596-1609); degrees 1 and 2 are not and are kept as they are.
// Variables are grouped by degree like the checks (list = the variables of one degree).  A block takes VPB of them and
// has all their channel values, messages and old sign words requested before the first sum.
template <int DV, int VPB>
__global__ void __launch_bounds__(kBpThreads) k_bp_vn_deg(const int32_t* __restrict__ list, int n_list,
                                                           const int32_t* __restrict__ vn_off, uint32_t S4,
                                                           int32_t* __restrict__ msg, const int32_t* __restrict__ qin,
                                                           const uint32_t* __restrict__ donew, uint32_t* __restrict__ hardw,
                                                           int32_t* __restrict__ soft, int qmax) {
  __shared__ int s_i_all[kVnGroups * VPB], s_e0_all[kVnGroups * VPB];
  if (threadIdx.x < kVnGroups * VPB) {
    const int t = blockIdx.x * (kVnGroups * VPB) + threadIdx.x;
    const int i = t < n_list ? list[t] : -1;
    s_i_all[threadIdx.x] = i;
    s_e0_all[threadIdx.x] = i >= 0 ? vn_off[i] : 0;
  }
  __syncthreads();
  const size_t s = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  const uint32_t dmask = donew[s >> 5];
  const bool done = (dmask >> (s & 31)) & 1u;
  if (dmask == 0xFFFFFFFFu) return;  // the whole warp has returned
  const uint32_t S32 = S4 >> 7;
  const bool lane0 = (threadIdx.x & 31) == 0;
  for (int g = 0; g < kVnGroups; ++g) {  // the variable ids are staged once for kVnGroups groups of VPB variables
  const int* s_i = s_i_all + g * VPB;
  const int* s_e0 = s_e0_all + g * VPB;
  if (s_i[0] < 0) break;
  int L[VPB], m[VPB][DV];
  uint32_t old[VPB];
#pragma unroll
  for (int v = 0; v < VPB; ++v) {
    const int i = s_i[v];
    L[v] = 0;
    old[v] = 0;
#pragma unroll
    for (int k = 0; k < DV; ++k) m[v][k] = 0;
    if (i < 0) continue;
    if (!done) {
      L[v] = *bp_row(const_cast<int32_t*>(qin) + s, (uint32_t)i, S4);
      const char* e = reinterpret_cast<const char*>(bp_row(msg + s, (uint32_t)s_e0[v], S4));
#pragma unroll
      for (int k = 0; k < DV; ++k) m[v][k] = *reinterpret_cast<const int32_t*>(e + (size_t)k * S4);
    }
    if (dmask && lane0) old[v] = hardw[(size_t)i * S32 + (s >> 5)];
  }
#pragma unroll
  for (int v = 0; v < VPB; ++v) {
    const int i = s_i[v];
    if (i < 0) break;
    int out = 0;
    if (!done) {
      char* e = reinterpret_cast<char*>(bp_row(msg + s, (uint32_t)s_e0[v], S4));
      auto at = [&](int k) -> int32_t& { return *reinterpret_cast<int32_t*>(e + (size_t)k * S4); };
      if constexpr (DV == 1) {
        at(0) = L[v];
        out = bp_qsum(L[v], m[v][0], qmax);
      } else if constexpr (DV == 2) {
        const int o1 = bp_qsum(L[v], m[v][0], qmax);
        at(0) = bp_qsum(L[v], m[v][1], qmax);
        at(1) = o1;
        out = bp_qsum(o1, m[v][1], qmax);
      } else {
        int t = L[v];
#pragma unroll
        for (int k = 0; k < DV; ++k) t = bp_qsum(t, m[v][k], qmax);
        out = t;
#pragma unroll
        for (int k = 0; k < DV; ++k) at(k) = bp_qsum(t, -m[v][k], qmax);
      }
      if (soft) *bp_row(soft + s, (uint32_t)i, S4) = out;
    }
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, !done && out < 0);
    if (lane0) hardw[(size_t)i * S32 + (s >> 5)] = (old[v] & dmask) | (bal & ~dmask);
  }
  }
}

// any degree: two passes over the variable's messages (the second one finds them in L1)
__global__ void __launch_bounds__(kBpThreads) k_bp_vn_any(const int32_t* __restrict__ list, int n_list,
                                                           const int32_t* __restrict__ vn_off, uint32_t S4,
                                                           int32_t* __restrict__ msg, const int32_t* __restrict__ qin,
                                                           const uint32_t* __restrict__ donew, uint32_t* __restrict__ hardw,
                                                           int32_t* __restrict__ soft, int qmax) {
  const size_t s = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  const uint32_t dmask = donew[s >> 5];
  const bool done = (dmask >> (s & 31)) & 1u;
  if (dmask == 0xFFFFFFFFu) return;
  const uint32_t S32 = S4 >> 7;
  const int i = list[blockIdx.x], e0 = vn_off[i], dv = vn_off[i + 1] - e0;
  int out = 0;
  if (!done) {
    char* e = reinterpret_cast<char*>(bp_row(msg + s, (uint32_t)e0, S4));
    auto at = [&](int k) -> int32_t& { return *reinterpret_cast<int32_t*>(e + (size_t)k * S4); };
    int t = *bp_row(const_cast<int32_t*>(qin) + s, (uint32_t)i, S4);
#pragma unroll 4
    for (int k = 0; k < dv; ++k) t = bp_qsum(t, at(k), qmax);
    out = t;
#pragma unroll 4
    for (int k = 0; k < dv; ++k) at(k) = bp_qsum(t, -at(k), qmax);
    if (soft) *bp_row(soft + s, (uint32_t)i, S4) = out;
  }
  const uint32_t bal = __ballot_sync(0xFFFFFFFFu, !done && out < 0);
  if ((threadIdx.x & 31) == 0) {
    uint32_t* hw = hardw + (size_t)i * S32 + (s >> 5);
    *hw = dmask ? ((*hw & dmask) | (bal & ~dmask)) : bal;
  }
}

// syndrome_check(LLRout) (ldpc.cpp:1631-1652) for 32 slots at a time: XOR of the sign words of a check's variables
constexpr int kSynChecks = 8;  // checks per thread: their parities are OR-ed before the one atomic
__global__ void k_bp_syndrome(const int32_t* __restrict__ cn_off, const int32_t* __restrict__ cn_var, int M, int S32,
                              const uint32_t* __restrict__ hardw, const uint32_t* __restrict__ donew,
                              uint32_t* __restrict__ failw) {
  const int w = blockIdx.y * blockDim.x + threadIdx.x;
  const int j0 = (blockIdx.x * blockDim.y + threadIdx.y) * kSynChecks;
  if (w >= S32 || j0 >= M) return;
  if (donew[w] == 0xFFFFFFFFu) return;
  uint32_t fail = 0;
  const int j1 = min(M, j0 + kSynChecks);
  for (int j = j0; j < j1; ++j) {
    uint32_t x = 0;
    const int k1 = cn_off[j + 1];
#pragma unroll 4
    for (int k = cn_off[j]; k < k1; ++k) x ^= hardw[(size_t)cn_var[k] * S32 + w];
    fail |= x;
  }
  if (fail) atomicOr(&failw[w], fail);
}

// slots whose syndrome is zero return `iter` (:1613-1616, or 0 from the initial check :1408-1411)
__global__ void k_bp_mark(int iter, int S32, uint32_t* __restrict__ failw, uint32_t* __restrict__ donew,
                          const int32_t* __restrict__ fid, int32_t* __restrict__ iters) {
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < S32; w += gridDim.x * blockDim.x) {
    const uint32_t d = donew[w];
    uint32_t ok = ~failw[w] & ~d;
    failw[w] = 0;
    if (!ok) continue;
    donew[w] = d | ok;
    while (ok) {
      const int b = __ffs(ok) - 1;
      ok &= ok - 1;
      iters[fid[32 * w + b]] = iter;
    }
  }
}
// the rest returns -iter (:1621)
__global__ void k_bp_finish(int iter, int S32, const uint32_t* __restrict__ donew, const int32_t* __restrict__ fid,
                            int32_t* __restrict__ iters) {
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < S32; w += gridDim.x * blockDim.x) {
    uint32_t open = ~donew[w];
    while (open) {
      const int b = __ffs(open) - 1;
      open &= open - 1;
      iters[fid[32 * w + b]] = -iter;
    }
  }
}

// ---- results of the selected slots (sel == NULL: all with a frame) back to frame-major layouts
__global__ void k_bp_harvest_bits(const uint32_t* __restrict__ hardw, const uint32_t* __restrict__ sel,
                                  const int32_t* __restrict__ fid, int N, int S32, uint8_t* __restrict__ out) {
  const int i = blockIdx.x * 32 + threadIdx.x, w = blockIdx.y * blockDim.y + threadIdx.y;
  if (w >= S32) return;
  uint32_t m = sel ? sel[w] : 0xFFFFFFFFu;
  const uint32_t word = i < N ? hardw[(size_t)i * S32 + w] : 0u;
  while (m) {
    const int b = __ffs(m) - 1;
    m &= m - 1;
    const int f = fid[32 * w + b];
    if (f >= 0 && i < N) out[(size_t)f * N + i] = (word >> b) & 1u;
  }
}
__global__ void k_bp_soft_out(const int32_t* __restrict__ soft, const int32_t* __restrict__ fid, int N, size_t S,
                              int32_t* __restrict__ out) {
  __shared__ int32_t tile[32][33];
  const int i0 = blockIdx.x * 32, s0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
  for (int r = ty; r < 32; r += 8)
    if (i0 + r < N) tile[r][tx] = soft[(size_t)(i0 + r) * S + s0 + tx];
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int f = fid[s0 + r];
    if (f >= 0 && i0 + tx < N) out[(size_t)f * N + i0 + tx] = tile[tx][r];
  }
}

// ---- repack: rows of the unfinished slots into a narrower batch
__global__ void __launch_bounds__(kBpThreads) k_bp_repack(const int32_t* __restrict__ src, size_t S_old, int32_t* __restrict__ dst,
                                                           size_t S_new, const int32_t* __restrict__ map) {
  const size_t k = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  const int from = map[k];
  dst[(size_t)blockIdx.x * S_new + k] = from >= 0 ? src[(size_t)blockIdx.x * S_old + from] : 0;
}

// ---- Monte-Carlo front and back end (src/LDPC_BER_Sim.cpp:246-311)
// BPSK + AWGN + LLR = 4x/N0 + to_qllr; slicer errors.  Thread = (node quad, slot): one Philox block.
__global__ void __launch_bounds__(kBpThreads) k_bp_channel(const uint32_t* __restrict__ sent, int N, size_t S, int F,
                                                            uint64_t first_frame, uint32_t k0, uint32_t k1, double sigma,
                                                            double four_over_N0, Calc c, int32_t* __restrict__ qin,
                                                            unsigned long long* __restrict__ cnt_uncoded_err) {
  const size_t s = (size_t)blockIdx.y * kBpThreads + threadIdx.x;
  const int q = blockIdx.x, v0 = 4 * q;
  uint32_t nerr = 0;
  int32_t qv[4] = {0, 0, 0, 0};
  if (s < (size_t)F) {
    double nrm[4];
    normals4_for(first_frame + s, (uint32_t)q, k0, k1, nrm);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (v0 + i >= N) break;
      const uint32_t bit = sent ? (sent[(size_t)(v0 + i) * (S >> 5) + (s >> 5)] >> (s & 31)) & 1u : 0u;
      const double x = (bit ? -1.0 : 1.0) + sigma * nrm[i];
      qv[i] = bp_to_qllr(four_over_N0 * x, c);
      nerr += (x > 0.0 ? 0u : 1u) ^ bit;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (v0 + i < N) qin[(size_t)(v0 + i) * S + s] = qv[i];
  for (int o = 16; o; o >>= 1) nerr += __shfl_xor_sync(0xFFFFFFFFu, nerr, o);
  if ((threadIdx.x & 31) == 0 && nerr) atomicAdd(cnt_uncoded_err, (unsigned long long)nerr);
}

// BERC/BLERC over the first K symbols of the selected slots
__global__ void __launch_bounds__(kBpThreads) k_bp_count(const uint32_t* __restrict__ hardw, const uint32_t* __restrict__ sent,
                                                          const uint32_t* __restrict__ sel, int K, int S32, int rows_per_cta,
                                                          uint32_t* __restrict__ ferrw, unsigned long long* __restrict__ cnt_data_err) {
  const int w = blockIdx.y * kBpThreads + threadIdx.x;
  uint32_t nerr = 0, any = 0;
  if (w < S32) {
    const uint32_t m = sel[w];
    if (m) {
      const int k1 = min(K, (int)(blockIdx.x + 1) * rows_per_cta);
      for (int k = blockIdx.x * rows_per_cta; k < k1; ++k) {
        const uint32_t d = (hardw[(size_t)k * S32 + w] ^ (sent ? sent[(size_t)k * S32 + w] : 0u)) & m;
        nerr += __popc(d);
        any |= d;
      }
      if (any) atomicOr(&ferrw[w], any);
    }
  }
  for (int o = 16; o; o >>= 1) nerr += __shfl_xor_sync(0xFFFFFFFFu, nerr, o);
  if ((threadIdx.x & 31) == 0 && nerr) atomicAdd(cnt_data_err, (unsigned long long)nerr);
}
__global__ void k_bp_fold_ferr(uint32_t* __restrict__ ferrw, int S32, unsigned long long* __restrict__ cnt_ferr) {
  unsigned long long n = 0;
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < S32; w += gridDim.x * blockDim.x) {
    n += __popc(ferrw[w]);
    ferrw[w] = 0;
  }
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xFFFFFFFFu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(cnt_ferr, n);
}
// frames, bits, |return| statistics of a batch; counters laid out like lutldpc_sim_counters
__global__ void k_bp_count_iters(const int32_t* __restrict__ iters, int F, int K, int nvar, unsigned long long* __restrict__ cnt) {
  unsigned long long si = 0, nf = 0;
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
    int it = iters[f];
    nf += it < 0;
    it = it < 0 ? -it : it;
    si += it;
    atomicAdd(&cnt[8 + min(it, LUTLDPC_ITER_HIST_MAX - 1)], 1ull);
  }
  for (int o = 16; o; o >>= 1) {
    si += __shfl_xor_sync(0xFFFFFFFFu, si, o);
    nf += __shfl_xor_sync(0xFFFFFFFFu, nf, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (si) atomicAdd(&cnt[6], si);
    if (nf) atomicAdd(&cnt[7], nf);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt[0], (unsigned long long)F);
    atomicAdd(&cnt[2], (unsigned long long)F * K);
    atomicAdd(&cnt[4], (unsigned long long)F * nvar);
  }
}

// frame-major 0/1 bytes <-> one bit per slot rows (encoder interface)
__global__ void k_bp_bits_to_rows(const uint8_t* __restrict__ in, int K, int S32, int F, uint32_t* __restrict__ rows) {
  const int w = blockIdx.y * blockDim.x + threadIdx.x, k = blockIdx.x;
  if (w >= S32) return;
  uint32_t b = 0;
  for (int j = 0; j < 32; ++j) {
    const int f = 32 * w + j;
    if (f < F) b |= (uint32_t)(in[(size_t)f * K + k] & 1) << j;
  }
  rows[(size_t)k * S32 + w] = b;
}

thread_local std::string g_bp_create_error;

}  // namespace

struct lutldpc_bp {
  int device = 0;
  cudaStream_t stream = nullptr;
  FlatGraph g;
  int nchk_lin_indep = 0, max_dc = 0, max_dv = 0;
  lutldpc_bp_params par{};
  Calc calc{};
  bool repack_on = true;
  bool fast_boxplus = false;  // a table is in use and qllr_max >= table[0] (see bp_boxplus)
  std::vector<int32_t> table;
  DevBuf<int32_t> d_vn_off, d_cn_off, d_cn_edge, d_cn_var, d_table, d_cn_list, d_cls_edge;
  struct CnClass { int dc, n, off, edge_off; };
  DevBuf<int32_t> d_vn_list;
  std::vector<CnClass> vn_classes;  // variables grouped by degree: d_vn_list[off .. off + n) (dc = the degree)
  std::vector<CnClass> cn_classes;  // checks grouped by degree: d_cn_list[off .. off + n)
  uint32_t* d_P = nullptr;  // generator (gf2.cu)
  int P_wp = 0, gen_rows = 0, gen_cols = 0;
  // batch state
  DevBuf<int32_t> msgA, msgB, qinA, qinB, soft, iters, fid, map, stage_i32;
  DevBuf<double> stage_f64;
  DevBuf<uint32_t> hardw, failw, donew, validw, selw, ferrw, sent;
  DevBuf<uint8_t> bits_out, stage_u8;
  DevBuf<unsigned long long> cnt;
  uint32_t* h_words = nullptr;  // pinned: donew of the running batch
  size_t h_words_n = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int64_t launches = 0, stat_frame_iters = 0;
  bool have_stats = false;
  std::string err;

  ~lutldpc_bp() {
    if (d_P) cudaFree(d_P);
    if (h_words) cudaFreeHost(h_words);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
  }
  int ninfo() const { return g.nvar - nchk_lin_indep; }
  double rate() const { return 1.0 - (double)nchk_lin_indep / g.nvar; }  // LDPC_Code::get_rate, ldpc.h:938-940
  int fail(int code, const std::string& what) { err = what; return code; }
  int cuda_fail(cudaError_t e, const char* where) {
    err = std::string(where) + ": " + cudaGetErrorString(e);
    cudaGetLastError();
    return LUTLDPC_ERR_CUDA;
  }
  int max_batch(bool want_soft) const {  // slots that fit half of the free memory, at most 16384
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) return 256;
    const double per_slot = ((double)g.num_edges + g.nvar) * 4.0 * 1.5 + (want_soft ? 4.0 * g.nvar : 0.0) + 16.0 * g.nvar / 8.0;
    size_t mine = (msgA.n + msgB.n + qinA.n + qinB.n + soft.n) * 4;
    double slots = 0.5 * ((double)fr + (double)mine) / per_slot;
    int b = (int)std::min(16384.0, slots);
    return std::max(256, b / 256 * 256);
  }
};

#define CKB(call, where)                                   \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) return h->cuda_fail(e_, where); \
  } while (0)
#define CKBL(where) CKB(cudaGetLastError(), where)

namespace {

struct BatchIO {       // what a batch hands back
  bool want_soft = false;   // LLRout of every frame (no repacking then)
  uint8_t* d_bits = nullptr;  // [F][N] bytes on the device, or NULL
  int32_t* d_soft_out = nullptr;  // [F][N] on the device, or NULL
  bool count = false;       // Monte-Carlo counting into h->cnt
  const uint32_t* d_sent = nullptr;  // sent bits as slot rows (NULL: all-zero codeword)
};

template <int DC>
void launch_cn_deg(lutldpc_bp* h, const lutldpc_bp::CnClass& cl, size_t S, int32_t* msg) {
  // two checks per block up to degree 16: 56 registers at degrees 7-8, i.e. 32 warps per SM with 14-16 loads each in
  // flight (measured against 3 and 4 checks per block, which hold more loads per thread but 20 warps: 167 / 177 / 172 ms
  // for the N=64800 step)
  constexpr int CPB = DC <= 16 ? 2 : 1;
  dim3 grid((cl.n + kCnGroups * CPB - 1) / (kCnGroups * CPB), (unsigned)(S / kBpThreads));
  const size_t smem = ((size_t)h->calc.d2 + 1) * 4;
  if (h->fast_boxplus)
    k_bp_cn_deg<DC, CPB, true><<<grid, kBpThreads, smem, h->stream>>>(h->d_cls_edge.p + cl.edge_off, cl.n, (uint32_t)(S * 4), msg,
                                                                      h->donew.p, h->calc, h->d_table.p);
  else
    k_bp_cn_deg<DC, CPB, false><<<grid, kBpThreads, smem, h->stream>>>(h->d_cls_edge.p + cl.edge_off, cl.n, (uint32_t)(S * 4), msg,
                                                                       h->donew.p, h->calc, h->d_table.p);
}

int launch_cn(lutldpc_bp* h, size_t S, int32_t* msg) {
  for (const lutldpc_bp::CnClass& cl : h->cn_classes) {
    switch (cl.dc) {
#define BP_CN_CASE(D) case D: launch_cn_deg<D>(h, cl, S, msg); break;
      BP_CN_CASE(2) BP_CN_CASE(3) BP_CN_CASE(4) BP_CN_CASE(5) BP_CN_CASE(6) BP_CN_CASE(7) BP_CN_CASE(8) BP_CN_CASE(9)
      BP_CN_CASE(10) BP_CN_CASE(11) BP_CN_CASE(12) BP_CN_CASE(13) BP_CN_CASE(14) BP_CN_CASE(15) BP_CN_CASE(16)
      BP_CN_CASE(17) BP_CN_CASE(18) BP_CN_CASE(19) BP_CN_CASE(20) BP_CN_CASE(21) BP_CN_CASE(22) BP_CN_CASE(23) BP_CN_CASE(24)
      BP_CN_CASE(25) BP_CN_CASE(26) BP_CN_CASE(27) BP_CN_CASE(28) BP_CN_CASE(29) BP_CN_CASE(30) BP_CN_CASE(31) BP_CN_CASE(32)
#undef BP_CN_CASE
      default: {
        const int T = 64, cpb = 4;  // 2 * 200 * 64 * 4 = 100 KB at the reference's max_cnd
        const size_t smem = ((size_t)h->calc.d2 + 1) * 4 + (size_t)2 * h->max_dc * T * 4;
        dim3 g2((cl.n + cpb - 1) / cpb, (unsigned)(S / T));
        k_bp_cn_big<<<g2, T, smem, h->stream>>>(h->d_cn_list.p + cl.off, cl.n, cpb, h->d_cn_off.p, h->d_cn_edge.p, h->max_dc,
                                                 (uint32_t)(S * 4), msg, h->donew.p, h->calc, h->d_table.p);
      }
    }
    CKBL("bp check pass");
    ++h->launches;
  }
  return LUTLDPC_OK;
}

template <int DV>
void launch_vn_deg(lutldpc_bp* h, const lutldpc_bp::CnClass& cl, size_t S, int32_t* msg, const int32_t* qin, int32_t* soft) {
  constexpr int VPB = DV <= 2 ? 8 : DV <= 4 ? 4 : DV <= 8 ? 2 : 1;  // VPB * DV message registers per thread
  dim3 grid((cl.n + kVnGroups * VPB - 1) / (kVnGroups * VPB), (unsigned)(S / kBpThreads));
  k_bp_vn_deg<DV, VPB><<<grid, kBpThreads, 0, h->stream>>>(h->d_vn_list.p + cl.off, cl.n, h->d_vn_off.p, (uint32_t)(S * 4), msg, qin,
                                                          h->donew.p, h->hardw.p, soft, h->calc.qmax);
}

int launch_vn(lutldpc_bp* h, size_t S, int32_t* msg, const int32_t* qin, int32_t* soft) {
  for (const lutldpc_bp::CnClass& cl : h->vn_classes) {
    switch (cl.dc) {
#define BP_VN_CASE(D) case D: launch_vn_deg<D>(h, cl, S, msg, qin, soft); break;
      BP_VN_CASE(1) BP_VN_CASE(2) BP_VN_CASE(3) BP_VN_CASE(4) BP_VN_CASE(5) BP_VN_CASE(6) BP_VN_CASE(7) BP_VN_CASE(8)
      BP_VN_CASE(9) BP_VN_CASE(10) BP_VN_CASE(11) BP_VN_CASE(12) BP_VN_CASE(13) BP_VN_CASE(14) BP_VN_CASE(15) BP_VN_CASE(16)
#undef BP_VN_CASE
      default:
        k_bp_vn_any<<<dim3(cl.n, (unsigned)(S / kBpThreads)), kBpThreads, 0, h->stream>>>(h->d_vn_list.p + cl.off, cl.n, h->d_vn_off.p,
                                                                                        (uint32_t)(S * 4), msg, qin, h->donew.p,
                                                                                        h->hardw.p, soft, h->calc.qmax);
    }
    CKBL("bp variable pass");
    ++h->launches;
  }
  return LUTLDPC_OK;
}

// bp_decode for the F frames whose inputs are in h->qinA[N][S]
int run_batch(lutldpc_bp* h, int F, size_t S, const BatchIO& io) {
  cudaStream_t st = h->stream;
  const FlatGraph& g = h->g;
  const int N = g.nvar, M = g.nchk, E = g.num_edges, K = h->ninfo();
  int S32 = (int)(S >> 5);
  CKB(h->msgA.ensure((size_t)E * S), "alloc messages");
  CKB(h->hardw.ensure((size_t)N * S32), "alloc hard bits");
  CKB(h->failw.ensure(S32), "alloc"); CKB(h->donew.ensure(S32), "alloc"); CKB(h->validw.ensure(S32), "alloc");
  CKB(h->selw.ensure(S32), "alloc"); CKB(h->ferrw.ensure(S32), "alloc");
  CKB(h->iters.ensure(F), "alloc"); CKB(h->fid.ensure(S), "alloc"); CKB(h->map.ensure(S), "alloc");
  if (io.want_soft) CKB(h->soft.ensure((size_t)N * S), "alloc soft output");
  if (h->h_words_n < (size_t)S32) {
    if (h->h_words) cudaFreeHost(h->h_words);
    h->h_words = nullptr;
    CKB(cudaMallocHost((void**)&h->h_words, (size_t)S32 * 4), "alloc pinned");
    h->h_words_n = S32;
  }
  std::vector<int32_t> fid(S, -1);
  for (int f = 0; f < F; ++f) fid[f] = f;
  std::vector<uint32_t> done(S32, 0), valid(S32, 0);
  for (size_t s = 0; s < S; ++s) {
    if (fid[s] >= 0) valid[s >> 5] |= 1u << (s & 31);
    else done[s >> 5] |= 1u << (s & 31);
  }
  CKB(cudaMemcpyAsync(h->fid.p, fid.data(), S * 4, cudaMemcpyHostToDevice, st), "H2D");
  CKB(cudaMemcpyAsync(h->donew.p, done.data(), (size_t)S32 * 4, cudaMemcpyHostToDevice, st), "H2D");
  CKB(cudaMemcpyAsync(h->validw.p, valid.data(), (size_t)S32 * 4, cudaMemcpyHostToDevice, st), "H2D");
  CKB(cudaMemsetAsync(h->failw.p, 0, (size_t)S32 * 4, st), "memset");
  CKB(cudaMemsetAsync(h->ferrw.p, 0, (size_t)S32 * 4, st), "memset");
  CKB(cudaMemsetAsync(h->iters.p, 0, (size_t)F * 4, st), "memset");
  CKB(cudaStreamSynchronize(st), "sync");  // fid/done/valid are stack vectors

  int32_t* msg = h->msgA.p;
  int32_t* qin = h->qinA.p;
  bool in_A = true;
  int32_t* soft = io.want_soft ? h->soft.p : nullptr;
  k_bp_init<<<dim3(N, (unsigned)(S / kBpThreads)), kBpThreads, 0, st>>>(h->d_vn_off.p, N, S, qin, msg, h->hardw.p, soft);
  CKBL("bp init");
  ++h->launches;

  auto syndrome_and_mark = [&](int iter) -> int {
    k_bp_syndrome<<<dim3((M + 4 * kSynChecks - 1) / (4 * kSynChecks), (S32 + 31) / 32), dim3(32, 4), 0, st>>>(h->d_cn_off.p, h->d_cn_var.p, M, S32, h->hardw.p,
                                                                              h->donew.p, h->failw.p);
    CKBL("bp syndrome");
    k_bp_mark<<<std::max(1, std::min(64, S32 / 128)), 128, 0, st>>>(iter, S32, h->failw.p, h->donew.p, h->fid.p, h->iters.p);
    CKBL("bp mark");
    h->launches += 2;
    CKB(cudaMemcpyAsync(h->h_words, h->donew.p, (size_t)S32 * 4, cudaMemcpyDeviceToHost, st), "D2H done flags");
    CKB(cudaStreamSynchronize(st), "sync");
    return LUTLDPC_OK;
  };
  auto count_active = [&]() {
    size_t a = 0;
    for (int w = 0; w < S32; ++w) a += 32 - __builtin_popcount(h->h_words[w]);
    return a;
  };
  // results of the slots in `sel` (device words) leave the batch
  auto harvest = [&](const uint32_t* sel) -> int {
    if (io.d_bits) {
      k_bp_harvest_bits<<<dim3((N + 31) / 32, (S32 + 7) / 8), dim3(32, 8), 0, st>>>(h->hardw.p, sel, h->fid.p, N, S32, io.d_bits);
      CKBL("bp harvest");
      ++h->launches;
    }
    if (io.count) {
      const int rows_per = 64;
      k_bp_count<<<dim3((K + rows_per - 1) / rows_per, (S32 + kBpThreads - 1) / kBpThreads), kBpThreads, 0, st>>>(
          h->hardw.p, io.d_sent, sel, K, S32, rows_per, h->ferrw.p, h->cnt.p + 3);
      CKBL("bp count");
      k_bp_fold_ferr<<<1, 256, 0, st>>>(h->ferrw.p, S32, h->cnt.p + 1);
      CKBL("bp count");
      h->launches += 2;
    }
    return LUTLDPC_OK;
  };

  size_t active = F;
  int rc;
  if (h->par.pisc) {  // :1408-1411
    if ((rc = syndrome_and_mark(0))) return rc;
    active = count_active();
  }
  const int n_iter = std::max(1, (int)h->par.max_iters);  // do { } while (iter < max_iters): at least one pass
  int iter = 0;
  while (active > 0 && iter < n_iter) {
    ++iter;
    if ((rc = launch_cn(h, S, msg))) return rc;
    if ((rc = launch_vn(h, S, msg, qin, soft))) return rc;
    if (!h->par.psc) continue;
    if ((rc = syndrome_and_mark(iter))) return rc;
    active = count_active();
    // drop the frames that have returned once at most half of the slots are still running
    const size_t S_new = (active + 255) / 256 * 256;
    if (h->repack_on && !io.want_soft && !io.d_sent && active > 0 && iter < n_iter && S_new * 2 <= S) {
      // 1. results of the finished slots
      std::vector<uint32_t> sel(S32);
      std::vector<int32_t> fid_new(S_new, -1), map(S_new, -1);
      size_t k = 0;
      for (size_t s = 0; s < S; ++s) {
        const bool d = (h->h_words[s >> 5] >> (s & 31)) & 1u;
        if (!d) { map[k] = (int32_t)s; fid_new[k] = fid[s]; ++k; }
      }
      for (int w = 0; w < S32; ++w) sel[w] = h->h_words[w] & valid[w];
      CKB(cudaMemcpyAsync(h->selw.p, sel.data(), (size_t)S32 * 4, cudaMemcpyHostToDevice, st), "H2D");
      if ((rc = harvest(h->selw.p))) return rc;
      // 2. gather
      DevBuf<int32_t>& mdst = in_A ? h->msgB : h->msgA;
      DevBuf<int32_t>& qdst = in_A ? h->qinB : h->qinA;
      CKB(mdst.ensure((size_t)E * S_new), "alloc repack");
      CKB(qdst.ensure((size_t)N * S_new), "alloc repack");
      CKB(cudaMemcpyAsync(h->map.p, map.data(), S_new * 4, cudaMemcpyHostToDevice, st), "H2D");
      k_bp_repack<<<dim3(E, (unsigned)(S_new / kBpThreads)), kBpThreads, 0, st>>>(msg, S, mdst.p, S_new, h->map.p);
      CKBL("bp repack");
      k_bp_repack<<<dim3(N, (unsigned)(S_new / kBpThreads)), kBpThreads, 0, st>>>(qin, S, qdst.p, S_new, h->map.p);
      CKBL("bp repack");
      h->launches += 2;
      msg = mdst.p; qin = qdst.p; in_A = !in_A;
      // 3. new slot tables
      S = S_new; S32 = (int)(S >> 5);
      fid.swap(fid_new);
      done.assign(S32, 0); valid.assign(S32, 0);
      for (size_t s = 0; s < S; ++s) {
        if (fid[s] >= 0) valid[s >> 5] |= 1u << (s & 31);
        else done[s >> 5] |= 1u << (s & 31);
      }
      CKB(cudaMemcpyAsync(h->fid.p, fid.data(), S * 4, cudaMemcpyHostToDevice, st), "H2D");
      CKB(cudaMemcpyAsync(h->donew.p, done.data(), (size_t)S32 * 4, cudaMemcpyHostToDevice, st), "H2D");
      CKB(cudaMemcpyAsync(h->validw.p, valid.data(), (size_t)S32 * 4, cudaMemcpyHostToDevice, st), "H2D");
      CKB(cudaMemsetAsync(h->failw.p, 0, (size_t)S32 * 4, st), "memset");
      CKB(cudaStreamSynchronize(st), "sync");
    }
  }
  if (active > 0) {
    k_bp_finish<<<std::max(1, std::min(64, S32 / 128)), 128, 0, st>>>(iter, S32, h->donew.p, h->fid.p, h->iters.p);
    CKBL("bp finish");
    ++h->launches;
  }
  if ((rc = harvest(h->validw.p))) return rc;
  if (io.d_soft_out) {
    k_bp_soft_out<<<dim3((N + 31) / 32, (unsigned)(S / 32)), dim3(32, 8), 0, st>>>(h->soft.p, h->fid.p, N, S, io.d_soft_out);
    CKBL("bp soft out");
    ++h->launches;
  }
  if (io.count) {
    k_bp_count_iters<<<std::min(64, (F + 255) / 256), 256, 0, st>>>(h->iters.p, F, K, N, h->cnt.p);
    CKBL("bp count");
    ++h->launches;
  }
  return LUTLDPC_OK;
}

int begin_call(lutldpc_bp* h) {
  CKB(cudaSetDevice(h->device), "cudaSetDevice");
  if (!h->ev0) CKB(cudaEventCreate(&h->ev0), "event");
  if (!h->ev1) CKB(cudaEventCreate(&h->ev1), "event");
  h->launches = 0;
  h->have_stats = false;
  CKB(cudaEventRecord(h->ev0, h->stream), "event");
  return LUTLDPC_OK;
}

template <class T>
int decode_host(lutldpc_bp* h, int n_frames, const T* in, uint8_t* out_bits, int32_t* out_soft, int32_t* out_iters) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (n_frames < 0 || (n_frames && !in)) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  int rc = begin_call(h);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  const int N = h->g.nvar;
  const int maxb = h->max_batch(out_soft != nullptr);
  int64_t frame_iters = 0;
  std::vector<int32_t> its;
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    const int nf = std::min(maxb, n_frames - f0);
    const size_t S = ((size_t)nf + 255) / 256 * 256;
    const T* src = in + (size_t)f0 * N;
    const T* d_in;
    if constexpr (sizeof(T) == 8) {
      CKB(h->stage_f64.ensure((size_t)nf * N), "alloc staging");
      d_in = h->stage_f64.p;
    } else {
      CKB(h->stage_i32.ensure((size_t)nf * N), "alloc staging");
      d_in = h->stage_i32.p;
    }
    CKB(cudaMemcpyAsync((void*)d_in, src, (size_t)nf * N * sizeof(T), cudaMemcpyHostToDevice, st), "H2D inputs");
    CKB(h->qinA.ensure((size_t)N * S), "alloc inputs");
    k_bp_load<T><<<dim3((N + 31) / 32, (unsigned)(S / 32)), dim3(32, 8), 0, st>>>(d_in, nf, N, S, h->qinA.p, h->calc);
    CKBL("bp load");
    ++h->launches;
    BatchIO io;
    io.want_soft = out_soft != nullptr;
    if (out_bits) {
      CKB(h->bits_out.ensure((size_t)nf * N), "alloc bits");
      io.d_bits = h->bits_out.p;
    }
    if (out_soft) {  // the staging buffer of the inputs is free again after k_bp_load
      CKB(h->stage_i32.ensure((size_t)nf * N), "alloc staging");
      io.d_soft_out = h->stage_i32.p;
    }
    if ((rc = run_batch(h, nf, S, io))) return rc;
    if (out_bits) CKB(cudaMemcpyAsync(out_bits + (size_t)f0 * N, h->bits_out.p, (size_t)nf * N, cudaMemcpyDeviceToHost, st), "D2H bits");
    if (out_soft) CKB(cudaMemcpyAsync(out_soft + (size_t)f0 * N, h->stage_i32.p, (size_t)nf * N * 4, cudaMemcpyDeviceToHost, st), "D2H soft");
    its.resize(nf);
    CKB(cudaMemcpyAsync(its.data(), h->iters.p, (size_t)nf * 4, cudaMemcpyDeviceToHost, st), "D2H iters");
    CKB(cudaStreamSynchronize(st), "sync");
    for (int f = 0; f < nf; ++f) frame_iters += std::abs(its[f]);
    if (out_iters) memcpy(out_iters + f0, its.data(), (size_t)nf * 4);
  }
  CKB(cudaEventRecord(h->ev1, st), "event");
  CKB(cudaStreamSynchronize(st), "sync");
  h->stat_frame_iters = frame_iters;
  h->have_stats = n_frames > 0;
  return LUTLDPC_OK;
}

}  // namespace

extern "C" {

// the argument checks of LLR_calc_unit and set_exit_conditions
static int bp_check_params(const lutldpc_bp_params* p) {
  // LLR_calc_unit::init_llr_bounds asserts d4 <= 8*sizeof(QLLR)-4 (llr.cpp:74-80)
  if (p->qllr_total_res < 2 || p->qllr_total_res > 28 || p->qllr_scale_res < 0 || p->qllr_scale_res > 27 ||
      p->qllr_table_size < 0 || p->qllr_table_size > 8192 || p->qllr_spacing_res < 0 || p->qllr_spacing_res > 30 ||
      p->max_iters < 0) {
    g_bp_create_error = "LLR_calc_unit::init_llr_bounds(): Resolution too large! (or a negative [BP] parameter)";
    return LUTLDPC_ERR_ARG;
  }
  return LUTLDPC_OK;
}

// everything after the graph is in h->g: degree classes, LLR_calc_unit table, device copies
static int bp_finish(lutldpc_bp* h) {
  const FlatGraph& g = h->g;
  std::vector<int32_t> vn_off(g.nvar + 1, 0), cn_off(g.nchk + 1, 0), edge_var(g.num_edges), cn_var(g.num_edges);
  for (int i = 0; i < g.nvar; ++i) {
    if (g.dv[i] < 1) { g_bp_create_error = "LDPC_Code::bp_decode(): sumX1(i)=0"; return LUTLDPC_ERR_MODEL; }
    vn_off[i + 1] = vn_off[i] + g.dv[i];
    h->max_dv = std::max(h->max_dv, (int)g.dv[i]);
    for (int k = 0; k < g.dv[i]; ++k) edge_var[vn_off[i] + k] = i;
  }
  for (int j = 0; j < g.nchk; ++j) {
    if (g.dc[j] < 2) { g_bp_create_error = g.dc[j] == 0 ? "LDPC_Code::bp_decode(): sumX2(j)=0" : "LDPC_Code::bp_decode(): sumX2(j)=1"; return LUTLDPC_ERR_MODEL; }
    if (g.dc[j] > 200) { g_bp_create_error = "check node degrees >200 not supported in this version"; return LUTLDPC_ERR_MODEL; }
    cn_off[j + 1] = cn_off[j] + g.dc[j];
    h->max_dc = std::max(h->max_dc, (int)g.dc[j]);
  }
  for (int e = 0; e < g.num_edges; ++e) cn_var[e] = edge_var[g.cn_msg_idx[e]];
  std::vector<int32_t> cn_list, cls_edge;  // checks by degree, ascending index within a degree; their edge ids
  for (int d = 2; d <= h->max_dc; ++d) {
    lutldpc_bp::CnClass cl{d, 0, (int)cn_list.size(), (int)cls_edge.size()};
    for (int j = 0; j < g.nchk; ++j)
      if (g.dc[j] == d) {
        cn_list.push_back(j);
        ++cl.n;
        cls_edge.insert(cls_edge.end(), g.cn_msg_idx.begin() + cn_off[j], g.cn_msg_idx.begin() + cn_off[j + 1]);
      }
    if (cl.n) h->cn_classes.push_back(cl);
  }
  std::vector<int32_t> vn_list;
  for (int d = 1; d <= h->max_dv; ++d) {
    lutldpc_bp::CnClass cl{d, 0, (int)vn_list.size(), 0};
    for (int i = 0; i < g.nvar; ++i)
      if (g.dv[i] == d) { vn_list.push_back(i); ++cl.n; }
    if (cl.n) h->vn_classes.push_back(cl);
  }
  // LLR_calc_unit (llr.cpp:66-92)
  Calc& c = h->calc;
  c.d1 = h->par.qllr_scale_res; c.d2 = h->par.qllr_table_size; c.d3 = h->par.qllr_spacing_res;
  c.qmax = (1 << (h->par.qllr_total_res - 1)) - 1;
  c.scale = (double)(1 << c.d1);
  c.four = 4; c.two = 2;
  c.qmax_d = (double)c.qmax / (double)(1 << c.d1);
  h->table.assign(std::max(1, c.d2), 0);
  for (int i = 0; i < c.d2; ++i) {
    const double x = std::pow(2.0, (double)(c.d3 - c.d1)) * i;
    const double l = std::log(1 + std::exp(-x));
    h->table[i] = l > c.qmax_d ? kQllrMaxConst : l < -c.qmax_d ? -c.qmax : (int32_t)std::floor(0.5 + c.scale * l);
  }
  h->fast_boxplus = c.d2 > 0 && c.qmax >= h->table[0];
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = h->d_vn_off.upload(vn_off);
  if (e == cudaSuccess) e = h->d_cn_off.upload(cn_off);
  if (e == cudaSuccess) e = h->d_cn_edge.upload(g.cn_msg_idx);
  if (e == cudaSuccess) e = h->d_cn_var.upload(cn_var);
  if (e == cudaSuccess) e = h->d_table.upload(h->table);
  if (e == cudaSuccess) e = h->d_cn_list.upload(cn_list);
  if (e == cudaSuccess) e = h->d_cls_edge.upload(cls_edge);
  if (e == cudaSuccess) e = h->d_vn_list.upload(vn_list);
  if (e == cudaSuccess && h->max_dc > kCnRegsMax)
    e = cudaFuncSetAttribute(k_bp_cn_big, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(((size_t)c.d2 + 1) * 4 + (size_t)2 * h->max_dc * 64 * 4));
  if (e != cudaSuccess) {
    g_bp_create_error = std::string("device set-up: ") + cudaGetErrorString(e);
    cudaGetLastError();
    return LUTLDPC_ERR_CUDA;
  }
  return LUTLDPC_OK;
}

static lutldpc_bp* bp_new(const lutldpc_bp_params* p) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { g_bp_create_error = "no CUDA device: the BP decoder has no CPU path"; cudaGetLastError(); return nullptr; }
  lutldpc_bp* h = new (std::nothrow) lutldpc_bp();
  if (!h) { g_bp_create_error = "out of memory"; return nullptr; }
  h->device = dev;
  h->par = *p;
  return h;
}

int lutldpc_bp_create(const lutldpc_graph* gr, const lutldpc_bp_params* p, int with_generator, lutldpc_bp** out,
                      int32_t* col_perm) {
  if (!gr || !p || !out) { g_bp_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  *out = nullptr;
  if (gr->nvar <= 0 || gr->nchk <= 0 || gr->num_edges <= 0 || !gr->dv_vec || !gr->dc_vec || !gr->cn_msg_idx) {
    g_bp_create_error = "empty graph";
    return LUTLDPC_ERR_ARG;
  }
  int rc = bp_check_params(p);
  if (rc) return rc;
  lutldpc_bp* h = bp_new(p);
  if (!h) return LUTLDPC_ERR_CUDA;
  std::vector<int32_t> dv(gr->dv_vec, gr->dv_vec + gr->nvar), dc(gr->dc_vec, gr->dc_vec + gr->nchk),
      idx(gr->cn_msg_idx, gr->cn_msg_idx + gr->num_edges);
  ParityRows H;
  std::string err;
  if (!rows_from_graph(gr->nvar, gr->nchk, dv, dc, idx, H, err)) { g_bp_create_error = err; delete h; return LUTLDPC_ERR_MODEL; }
  if (with_generator) {  // LDPC_Generator_Systematic(H) permutes H in place, then LDPC_Code(H, G): src/LDPC_BER_Sim.cpp:171-192
    Gf2Result r;
    if (!gf2_systematic_device(H, r, true, nullptr, err)) { g_bp_create_error = err; delete h; return LUTLDPC_ERR_CUDA; }
    if (r.rank == 0 || r.rank >= gr->nvar) {
      if (r.d_P) cudaFree(r.d_P);
      g_bp_create_error = "parity-check matrix has no information columns";
      delete h;
      return LUTLDPC_ERR_MODEL;
    }
    ParityRows Hp;
    permute_columns(H, r.col_perm, Hp);
    flatten_graph(Hp, h->g);
    h->nchk_lin_indep = r.rank;
    h->d_P = r.d_P; h->P_wp = r.P_words_per_row; h->gen_rows = r.rank; h->gen_cols = gr->nvar - r.rank;
    if (col_perm) memcpy(col_perm, r.col_perm.data(), sizeof(int32_t) * gr->nvar);
  } else {
    flatten_graph(H, h->g);
    // LDPC_Parity::get_ncheck_lin_indep (itpp/itpp/comm/ldpc.h:138-147): rank below 5e4 columns, else ncheck
    h->nchk_lin_indep = gr->nchk_lin_indep >= 0 ? gr->nchk_lin_indep : (gr->nvar < 50000 ? gf2_rank_host(H) : gr->nchk);
    if (col_perm) for (int i = 0; i < gr->nvar; ++i) col_perm[i] = i;
  }
  if ((rc = bp_finish(h))) { delete h; return rc; }
  *out = h;
  return LUTLDPC_OK;
}

// the arrays LDPC_Code keeps for H (decoder_parameterization, itpp/itpp/comm/ldpc.cpp:1676-1745) from the flat graph
static void bp_codec_arrays(const FlatGraph& g, int nchk_lin_indep, BpCodecFile& f) {
  const int N = g.nvar, M = g.nchk;
  int cmax = 0, vmax = 0;
  for (int d : g.dv) cmax = std::max(cmax, d);
  for (int d : g.dc) vmax = std::max(vmax, d);
  std::vector<int32_t> vn_off(N + 1, 0), cn_off(M + 1, 0), edge_var(g.num_edges), edge_chk(g.num_edges), edge_pos(g.num_edges);
  for (int i = 0; i < N; ++i) {
    vn_off[i + 1] = vn_off[i] + g.dv[i];
    for (int k = 0; k < g.dv[i]; ++k) edge_var[vn_off[i] + k] = i;
  }
  for (int j = 0; j < M; ++j) {
    cn_off[j + 1] = cn_off[j] + g.dc[j];
    for (int k = 0; k < g.dc[j]; ++k) {
      edge_chk[g.cn_msg_idx[cn_off[j] + k]] = j;
      edge_pos[g.cn_msg_idx[cn_off[j] + k]] = k;
    }
  }
  f.nvar = N; f.ncheck = M; f.ncheck_lin_indep = nchk_lin_indep;
  f.sumX1.assign(g.dv.begin(), g.dv.end());
  f.sumX2.assign(g.dc.begin(), g.dc.end());
  f.C.assign((size_t)cmax * N, 0); f.iind.assign((size_t)N * cmax, 0);
  f.V.assign((size_t)M * vmax, 0); f.jind.assign((size_t)M * vmax, 0);
  for (int i = 0; i < N; ++i)
    for (int k = 0; k < g.dv[i]; ++k) {
      const int e = vn_off[i] + k;
      f.C[(size_t)k + (size_t)cmax * i] = edge_chk[e];                       // :1696-1701
      f.iind[(size_t)i + (size_t)k * N] = edge_chk[e] + edge_pos[e] * M;     // :1731-1741
    }
  for (int j = 0; j < M; ++j)
    for (int k = 0; k < g.dc[j]; ++k) {
      const int e = g.cn_msg_idx[cn_off[j] + k], v = edge_var[e];
      f.V[(size_t)j + (size_t)M * k] = v;                                    // :1705-1710
      f.jind[(size_t)j + (size_t)k * M] = v + (e - vn_off[v]) * N;           // :1714-1726
    }
}

int lutldpc_bp_convert_alist(const char* alist_path, const char* out_it_path) {
  if (!alist_path || !out_it_path) { g_bp_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  ParityRows H;
  std::string err;
  if (!load_alist_file(alist_path, H, err)) { g_bp_create_error = err; return LUTLDPC_ERR_MODEL; }
  FlatGraph g;
  flatten_graph(H, g);
  BpCodecFile f;
  bp_codec_arrays(g, H.nvar < 50000 ? gf2_rank_host(H) : H.nchk, f);  // LDPC_Parity::get_ncheck_lin_indep, ldpc.h:138-147
  if (!save_bp_codec_file(out_it_path, f, err)) { g_bp_create_error = err; return LUTLDPC_ERR_STATE; }
  return LUTLDPC_OK;
}

int lutldpc_bp_save_code(lutldpc_bp* h, const char* path) {
  if (!h || !path) return LUTLDPC_ERR_ARG;
  BpCodecFile f;
  bp_codec_arrays(h->g, h->nchk_lin_indep, f);
  if (h->d_P) {  // G: (nvar - ninfo) x ninfo, the first ceil(ninfo / 8) bytes of every bit-packed row
    CKB(cudaSetDevice(h->device), "cudaSetDevice");
    std::vector<uint32_t> P((size_t)h->gen_rows * h->P_wp);
    CKB(cudaMemcpy(P.data(), h->d_P, P.size() * 4, cudaMemcpyDeviceToHost), "D2H generator");
    const size_t nb = (size_t)(h->gen_cols + 7) / 8;
    f.gen_rows = h->gen_rows; f.gen_cols = h->gen_cols;
    f.gen_bytes.resize((size_t)h->gen_rows * nb);
    for (int r = 0; r < h->gen_rows; ++r) memcpy(&f.gen_bytes[r * nb], &P[(size_t)r * h->P_wp], nb);
  }
  std::string err;
  if (!save_bp_codec_file(path, f, err)) return h->fail(LUTLDPC_ERR_STATE, err);
  return LUTLDPC_OK;
}

int lutldpc_bp_create_from_file(const char* codec_path, const lutldpc_bp_params* p, lutldpc_bp** out) {
  if (!codec_path || !p || !out) { g_bp_create_error = "NULL argument"; return LUTLDPC_ERR_ARG; }
  *out = nullptr;
  int rc = bp_check_params(p);
  if (rc) return rc;
  BpCodecFile f;
  std::string err;
  if (!load_bp_codec_file(codec_path, f, err)) { g_bp_create_error = err; return LUTLDPC_ERR_MODEL; }
  // The graph exactly as the file holds it: the k-th edge of variable i is the check C(k + cmax*i), position ip of check
  // j is the variable V(j + ncheck*ip) and its edge jind(j + ncheck*ip) = v + k*nvar.  The orders are kept as stored
  // -- after LDPC_Generator_Systematic has permuted the columns of H, the lists of the reference's sparse matrix are
  // no longer sorted, and the order of a node's inputs is the order of its Boxplus / qsum operations.
  FlatGraph fg;
  fg.nvar = f.nvar; fg.nchk = f.ncheck;
  fg.dv.assign(f.sumX1.begin(), f.sumX1.end());
  fg.dc.assign(f.sumX2.begin(), f.sumX2.end());
  const size_t vmax = f.V.size() / (size_t)f.ncheck, cmax = f.C.size() / (size_t)f.nvar;
  std::vector<int64_t> vn_off(f.nvar + 1, 0);
  for (int i = 0; i < f.nvar; ++i) {
    if (f.sumX1[i] < 0 || (size_t)f.sumX1[i] > cmax) { g_bp_create_error = "bp codec file: sumX1 out of range"; return LUTLDPC_ERR_MODEL; }
    vn_off[i + 1] = vn_off[i] + f.sumX1[i];
  }
  if (vn_off[f.nvar] <= 0 || vn_off[f.nvar] > (1 << 28) || f.jind.size() != f.V.size() || f.iind.size() != f.C.size()) {
    g_bp_create_error = "bp codec file: inconsistent array sizes";
    return LUTLDPC_ERR_MODEL;
  }
  fg.num_edges = (int)vn_off[f.nvar];
  std::vector<uint8_t> seen(fg.num_edges, 0);
  for (int j = 0; j < f.ncheck; ++j) {
    if (f.sumX2[j] < 0 || (size_t)f.sumX2[j] > vmax) { g_bp_create_error = "bp codec file: sumX2 out of range"; return LUTLDPC_ERR_MODEL; }
    for (int ip = 0; ip < f.sumX2[j]; ++ip) {
      const size_t at = (size_t)j + (size_t)f.ncheck * ip;
      const int64_t ji = f.jind[at];
      const int64_t v = ji % f.nvar, k = ji / f.nvar;
      if (ji < 0 || v != f.V[at] || k >= f.sumX1[v] || f.C[(size_t)k + cmax * (size_t)v] != j || seen[vn_off[v] + k]) {
        g_bp_create_error = "bp codec file: C, V and jind do not describe one graph";
        return LUTLDPC_ERR_MODEL;
      }
      seen[vn_off[v] + k] = 1;
      fg.cn_msg_idx.push_back((int32_t)(vn_off[v] + k));
    }
  }
  if ((int)fg.cn_msg_idx.size() != fg.num_edges) { g_bp_create_error = "bp codec file: row and column weights disagree"; return LUTLDPC_ERR_MODEL; }
  lutldpc_bp* h = bp_new(p);
  if (!h) return LUTLDPC_ERR_CUDA;
  h->g = fg;
  h->nchk_lin_indep = f.ncheck_lin_indep;
  if (f.gen_rows > 0) {
    if (f.gen_rows != f.ncheck_lin_indep) { g_bp_create_error = "bp codec file: generator shape does not match the code"; delete h; return LUTLDPC_ERR_MODEL; }
    const int wp = (f.gen_cols + 31) / 32;
    const size_t nb = (size_t)(f.gen_cols + 7) / 8;
    std::vector<uint32_t> P((size_t)f.gen_rows * wp, 0u);
    for (int r = 0; r < f.gen_rows; ++r) memcpy(&P[(size_t)r * wp], &f.gen_bytes[r * nb], nb);
    cudaError_t e = cudaMalloc((void**)&h->d_P, P.size() * 4);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_P, P.data(), P.size() * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { g_bp_create_error = std::string("generator upload: ") + cudaGetErrorString(e); cudaGetLastError(); delete h; return LUTLDPC_ERR_CUDA; }
    h->P_wp = wp; h->gen_rows = f.gen_rows; h->gen_cols = f.gen_cols;
  }
  if ((rc = bp_finish(h))) { delete h; return rc; }
  *out = h;
  return LUTLDPC_OK;
}

void lutldpc_bp_destroy(lutldpc_bp* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  delete h;
}

const char* lutldpc_bp_last_error(const lutldpc_bp* h) { return h ? h->err.c_str() : g_bp_create_error.c_str(); }

int lutldpc_bp_get_info(const lutldpc_bp* h, lutldpc_info* info) {
  if (!h || !info) return LUTLDPC_ERR_ARG;
  memset(info, 0, sizeof *info);
  info->nvar = h->g.nvar; info->nchk = h->g.nchk; info->nchk_lin_indep = h->nchk_lin_indep; info->ninfo = h->ninfo();
  info->num_edges = h->g.num_edges; info->max_iters = h->par.max_iters; info->has_gen = h->d_P != nullptr;
  info->rate = h->rate();
  return LUTLDPC_OK;
}

int lutldpc_bp_set_repack(lutldpc_bp* h, int on) {
  if (!h) return LUTLDPC_ERR_ARG;
  h->repack_on = on != 0;
  return LUTLDPC_OK;
}

int lutldpc_bp_last_call_stats(const lutldpc_bp* h, double* device_ms, int64_t* kernel_launches, int64_t* frame_iterations) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (!h->have_stats) return LUTLDPC_ERR_STATE;
  float ms = 0;
  if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) { cudaGetLastError(); return LUTLDPC_ERR_CUDA; }
  if (device_ms) *device_ms = ms;
  if (kernel_launches) *kernel_launches = h->launches;
  if (frame_iterations) *frame_iterations = h->stat_frame_iters;
  return LUTLDPC_OK;
}

int lutldpc_bp_decode_llr(lutldpc_bp* h, int n_frames, const double* llr, uint8_t* out_bits, int32_t* out_iters) {
  return decode_host<double>(h, n_frames, llr, out_bits, nullptr, out_iters);
}

int lutldpc_bp_decode_qllr(lutldpc_bp* h, int n_frames, const int32_t* qllr_in, int32_t* qllr_out, int32_t* out_iters) {
  if (h && n_frames > 0 && !qllr_out) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  return decode_host<int32_t>(h, n_frames, qllr_in, nullptr, qllr_out, out_iters);
}

int lutldpc_bp_encode(lutldpc_bp* h, int n_frames, const uint8_t* info, uint8_t* codeword) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (n_frames < 0 || (n_frames && (!info || !codeword))) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  if (!h->d_P) return h->fail(LUTLDPC_ERR_STATE, "decoder has no generator matrix");
  CKB(cudaSetDevice(h->device), "cudaSetDevice");
  cudaStream_t st = h->stream;
  const int N = h->g.nvar, K = h->gen_cols, rows = h->gen_rows;
  const int maxb = 16384;
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    const int nf = std::min(maxb, n_frames - f0);
    const size_t S = ((size_t)nf + 255) / 256 * 256;
    const int S32 = (int)(S >> 5);
    CKB(h->sent.ensure((size_t)N * S32), "alloc cw");
    CKB(h->stage_u8.ensure((size_t)nf * K), "alloc");
    CKB(h->fid.ensure(S), "alloc");
    CKB(h->bits_out.ensure((size_t)nf * N), "alloc");
    std::vector<int32_t> fid(S, -1);
    for (int f = 0; f < nf; ++f) fid[f] = f;
    CKB(cudaMemcpyAsync(h->fid.p, fid.data(), S * 4, cudaMemcpyHostToDevice, st), "H2D");
    CKB(cudaMemcpyAsync(h->stage_u8.p, info + (size_t)f0 * K, (size_t)nf * K, cudaMemcpyHostToDevice, st), "H2D info");
    k_bp_bits_to_rows<<<dim3(K, (S32 + 127) / 128), 128, 0, st>>>(h->stage_u8.p, K, S32, nf, h->sent.p);
    CKBL("bits_to_rows");
    CKB(launch_encode_bits(h->d_P, h->P_wp, rows, K, (int)(S / 8), (uint8_t*)h->sent.p, st), "encode_bits");
    k_bp_harvest_bits<<<dim3((N + 31) / 32, (S32 + 7) / 8), dim3(32, 8), 0, st>>>(h->sent.p, nullptr, h->fid.p, N, S32, h->bits_out.p);
    CKBL("rows_to_bits");
    CKB(cudaMemcpyAsync(codeword + (size_t)f0 * N, h->bits_out.p, (size_t)nf * N, cudaMemcpyDeviceToHost, st), "D2H cw");
    CKB(cudaStreamSynchronize(st), "sync");
  }
  return LUTLDPC_OK;
}

int lutldpc_bp_simulate_snr(lutldpc_bp* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (!p || !acc || p->n_frames < 0) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  if (!p->zero_codeword && !h->d_P) return h->fail(LUTLDPC_ERR_STATE, "zero_codeword=0 needs a generator matrix in the decoder");
  int rc = begin_call(h);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  const int N = h->g.nvar, K = h->ninfo();
  int batch = p->batch_frames > 0 ? p->batch_frames : 8192;
  batch = std::min(batch, h->max_batch(false));
  batch = std::max(256, batch / 256 * 256);
  const size_t n_cnt = 8 + LUTLDPC_ITER_HIST_MAX;
  CKB(h->cnt.ensure(n_cnt), "alloc counters");
  CKB(cudaMemsetAsync(h->cnt.p, 0, n_cnt * 8, st), "memset counters");
  // N0 = 10^(-snr/10) / rate, sigma^2 = N0/2   (src/LDPC_BER_Sim.cpp:248-249)
  const double N0 = pow(10.0, -p->snr_db / 10.0) / h->rate();
  const double sigma = sqrt(N0 / 2.0), four_over_N0 = 4.0 / N0;
  const uint32_t k0 = (uint32_t)p->seed, k1 = (uint32_t)(p->seed >> 32);
  for (int64_t done = 0; done < p->n_frames; done += batch) {
    const int nf = (int)std::min<int64_t>(batch, p->n_frames - done);
    const size_t S = ((size_t)nf + 255) / 256 * 256;
    const int S32 = (int)(S >> 5);
    const uint64_t first = p->first_frame + (uint64_t)done;
    BatchIO io;
    io.count = true;
    if (!p->zero_codeword) {
      CKB(h->sent.ensure((size_t)N * S32), "alloc cw");
      k_info_bits<<<dim3(K, (unsigned)((S / 8 + 127) / 128)), 128, 0, st>>>((uint8_t*)h->sent.p, K, (int)(S / 8), nf, first, k0, k1);
      CKBL("info_bits");
      CKB(launch_encode_bits(h->d_P, h->P_wp, N - K, K, (int)(S / 8), (uint8_t*)h->sent.p, st), "encode_bits");
      h->launches += 2;
      io.d_sent = h->sent.p;
    }
    CKB(h->qinA.ensure((size_t)N * S), "alloc inputs");
    k_bp_channel<<<dim3((N + 3) / 4, (unsigned)(S / kBpThreads)), kBpThreads, 0, st>>>(io.d_sent, N, S, nf, first, k0, k1, sigma, four_over_N0,
                                                                                      h->calc, h->qinA.p, h->cnt.p + 5);
    CKBL("bp channel");
    ++h->launches;
    if ((rc = run_batch(h, nf, S, io))) return rc;
  }
  CKB(cudaEventRecord(h->ev1, st), "event");
  std::vector<unsigned long long> host(n_cnt);
  CKB(cudaMemcpyAsync(host.data(), h->cnt.p, n_cnt * 8, cudaMemcpyDeviceToHost, st), "D2H counters");
  CKB(cudaStreamSynchronize(st), "sync");
  h->have_stats = p->n_frames > 0;
  h->stat_frame_iters = (int64_t)host[6];
  acc->frames += host[0]; acc->frame_errors += host[1]; acc->data_bits += host[2]; acc->data_bit_errors += host[3];
  acc->uncoded_bits += host[4]; acc->uncoded_bit_errors += host[5]; acc->sum_iters += host[6];
  acc->decode_failures += host[7];
  for (int i = 0; i < LUTLDPC_ITER_HIST_MAX; ++i) acc->iter_hist[i] += host[8 + i];
  return LUTLDPC_OK;
}

}  // extern "C"
