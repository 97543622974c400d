// On-chip regime: LDPC_Code_LUT::lut_decode (src/LDPC_Code_LUT.cpp:259-353) for codes whose per-frame state fits
// shared memory (second regime: the (3,6) N=1000 code, the irregular N=500 code, the (6,32) N=2048 code).
//
// One CTA owns a group of C word columns (8 frames each) and keeps ALL of their decoder state -- E edge messages, N
// channel labels and N hard-decision bytes per column -- in shared memory for all iterations.  The check pass and
// the variable pass are phases of one persistent kernel separated by __syncthreads(); the per-iteration LUTs arrive
// by cp.async.bulk (TMA engine) while the check pass runs; the early-termination test (every VN unanimous + zero
// syndrome, src/LDPC_Code_LUT.cpp:437-452) is decided inside the CTA, per column: per-thread flags are OR-reduced
// across the warp (__reduce_or_sync) and merged in shared memory.  Within a phase the warps pull work units
// (32 nodes of one degree x one column, heaviest degree first) from a shared-memory counter, so an irregular code
// keeps all warps busy and a group whose columns have mostly finished spends all warps on the stragglers.  A group
// that has finished is replaced by the next one of the batch from a global queue.  HBM traffic is the inputs
// (2 nibbles per symbol) and the outputs (1 bit per symbol): no per-iteration launches, no finalize kernel, no host
// poll, no repacking.
//
// Same per-thread arithmetic as the HBM path (core.cuh SWAR min-sum, vn_fast.cuh leave-one-out programs, the node
// program interpreter for any other tree), so results are bit-identical to it and to the reference.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "kdev.cuh"
#include "onchip.hpp"

namespace lutldpc {

namespace {

// msgs in shared memory: word of edge e at e + (e >> 5).  The skew makes the variable pass -- lane l owns the dv
// consecutive edges starting at dv * l -- free of bank conflicts for even dv as well.
__device__ __forceinline__ uint32_t skew(uint32_t e) { return e + (e >> 5); }

struct OcVals {  // value slots of the node-program interpreter: this thread's column, slot-major
  uint32_t* base;
  int stride;
  __device__ __forceinline__ uint32_t get(int s) const { return base[s * stride]; }
  __device__ __forceinline__ void set(int s, uint32_t v) { base[s * stride] = v; }
};

__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int DV, int II>
struct OcStore {  // outputs of the compile-time program back to the node's edges; D = bits where outputs differ from the first
  static __device__ __forceinline__ void run(const uint32_t (&v)[DV + 1 + fastvn::Static<DV>::P.n_ops], uint32_t* msgs_s,
                                             const uint32_t (&ix)[DV], uint32_t r0, uint32_t& D) {
    if constexpr (II < DV) {
      constexpr int kOut = fastvn::Static<DV>::P.out[II];
      const uint32_t r = v[kOut];
      msgs_s[ix[II]] = r;
      if constexpr (II > 0) D |= r ^ r0;
      OcStore<DV, II + 1>::run(v, msgs_s, ix, r0, D);
    }
  }
};

// Flags of up to 8 columns in two words: column c owns bit (c & 3) of every nibble of word c >> 2 (a column's flags are
// one bit per frame, at bit 0 of the frame's nibble).
struct ColFlags {
  uint32_t w[2] = {0, 0};
  __device__ __forceinline__ void add(int col, uint32_t nib_bits) { w[col >> 2] |= (nib_bits & kNibOnes) << (col & 3); }
};
__device__ __forceinline__ uint32_t col_flags_of(const uint32_t* packed, int col) {
  return (packed[col >> 2] >> (col & 3)) & kNibOnes;
}
// OR over the warp, then one shared-memory atomic per word
__device__ __forceinline__ void warp_flags_or(uint32_t* s_packed, const ColFlags& f, int C) {
  uint32_t a = __reduce_or_sync(0xFFFFFFFFu, f.w[0]);
  if ((threadIdx.x & 31) == 0 && a) atomicOr(&s_packed[0], a);
  if (C > 4) {
    a = __reduce_or_sync(0xFFFFFFFFu, f.w[1]);
    if ((threadIdx.x & 31) == 0 && a) atomicOr(&s_packed[1], a);
  }
}

struct ColView {  // shared-memory state of the C columns of the group
  uint8_t* smem;
  uint32_t col_bytes, off_cha, off_hard;
  __device__ __forceinline__ uint32_t* msgs(int c) const { return reinterpret_cast<uint32_t*>(smem + (size_t)c * col_bytes); }
  __device__ __forceinline__ uint32_t* cha(int c) const { return reinterpret_cast<uint32_t*>(smem + (size_t)c * col_bytes + off_cha); }
  __device__ __forceinline__ uint8_t* hard(int c) const { return smem + (size_t)c * col_bytes + off_hard; }
};

// one variable node (this lane) in every live column: compile-time leave-one-out program, values in registers; the
// node's edge offsets and tables are column-invariant
template <int DV>
__device__ __forceinline__ void oc_vn_fast(const uint8_t* tabs, int2 nd, int q_out, const ColView& cv, int C, uint32_t live,
                                           ColFlags& fail) {
  constexpr int NOPS = fastvn::Static<DV>::P.n_ops;
  const int so = q_out - 1, root_shift = 4 - q_out;
  uint32_t ix[DV];
#pragma unroll
  for (int k = 0; k < DV; ++k) ix[k] = skew((uint32_t)nd.y + k);
  for (int col = 0; col < C; ++col) {
    if (!((live >> col) & 1)) continue;
    uint32_t* ms = cv.msgs(col);
    uint32_t v[DV + 1 + NOPS];
#pragma unroll
    for (int k = 0; k < DV; ++k) v[k] = ms[ix[k]];
    v[DV] = cv.cha(col)[nd.x];
    fastvn::Exec<DV, 0>::run(v, tabs, root_shift);
    constexpr int kOut0 = fastvn::Static<DV>::P.out[0];
    uint32_t D = 0;
    OcStore<DV, 0>::run(v, ms, ix, v[kOut0], D);
    fail.add(col, D >> so);  // some outgoing signs differ (:441-447)
  }
}

// one check (this lane) in every live column, exact degree: inputs in registers (chk_update_minsum, :355-402)
template <int DC>
__device__ __forceinline__ void oc_cn_exact(const int32_t* __restrict__ r, const ColView& cv, int C, uint32_t live,
                                            const QMask& qm, ColFlags& fail) {
  uint32_t ix[DC];
#pragma unroll
  for (int j = 0; j < DC; ++j) ix[j] = skew((uint32_t)__ldg(r + j));
  for (int col = 0; col < C; ++col) {
    if (!((live >> col) & 1)) continue;
    uint32_t* ms = cv.msgs(col);
    uint32_t x[DC];
#pragma unroll
    for (int j = 0; j < DC; ++j) x[j] = ms[ix[j]];
    CnAccG a;
    cng_init(a, x[0], x[1], qm);
#pragma unroll
    for (int j = 2; j < DC; ++j) cng_accum(a, x[j], qm);
#pragma unroll
    for (int j = 0; j < DC; ++j) ms[ix[j]] = cng_emit(a, x[j], qm);
    fail.add(col, a.par >> qm.s);
  }
}
__device__ __forceinline__ void oc_cn_any(const int32_t* __restrict__ r, int dc, const ColView& cv, int C, uint32_t live,
                                          const QMask& qm, ColFlags& fail) {
  for (int col = 0; col < C; ++col) {
    if (!((live >> col) & 1)) continue;
    uint32_t* ms = cv.msgs(col);
    CnAccG a;
    cng_init(a, ms[skew((uint32_t)__ldg(r))], ms[skew((uint32_t)__ldg(r + 1))], qm);
    for (int j = 2; j < dc; ++j) cng_accum(a, ms[skew((uint32_t)__ldg(r + j))], qm);
    for (int j = 0; j < dc; ++j) {
      const uint32_t ix = skew((uint32_t)__ldg(r + j));
      ms[ix] = cng_emit(a, ms[ix], qm);
    }
    fail.add(col, a.par >> qm.s);
  }
}

}  // namespace

#ifndef ONCHIP_MIN_BLOCKS
#define ONCHIP_MIN_BLOCKS 3  // resident CTAs per SM the kernel is compiled for (80 registers)
#endif
template <int MAXDV>
__global__ void __launch_bounds__(kOnchipThreads, (MAXDV <= 8 ? ONCHIP_MIN_BLOCKS : 1)) k_onchip(const OnchipParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* tab_s = smem + P.off_tab;
  uint32_t* slots_s = reinterpret_cast<uint32_t*>(smem + P.off_slots);
  __shared__ uint64_t bar;
  __shared__ uint32_t s_vn_fail[2], s_cn_fail[2], s_syn_fail[2];  // packed per column (ColFlags)
  __shared__ uint32_t s_done[kOnchipMaxCols], s_conv[kOnchipMaxCols];
  __shared__ int32_t s_iters[8 * kOnchipMaxCols];
  __shared__ int s_group;
  __shared__ unsigned int s_next[2];  // next work unit of the running phase; phases alternate between the two counters
  const int tid = threadIdx.x, lane = tid & 31, C = P.C;
  const ColView cv{smem, P.col_bytes, P.off_cha, P.off_hard};
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
    s_next[0] = s_next[1] = 0;
  }
  uint32_t bar_parity = 0;
  int staged_slot = -1;  // program images of this iteration slot are in tab_s
  bool copy_pending = false;
  __syncthreads();

  // program images of iteration slot `slot`: one bulk copy, completion on the mbarrier.  Called by all threads
  // after a __syncthreads() that ended every read of the previous slot's images.
  auto stage = [&](int slot) {
    if (slot == staged_slot) return;
    if (tid == 0) {
      fence_proxy_async();  // generic-proxy reads of tab_s are ordered before the async-proxy writes
      const uint32_t bytes = slot == P.dec_slot ? P.dec_bytes : P.slot_bytes;
      mbar_expect_tx(&bar, bytes);
      const uint8_t* src = P.slot_images + (size_t)slot * P.slot_stride;
      const uint32_t kPiece = 16384;
      for (uint32_t off = 0; off < bytes; off += kPiece) bulk_g2s(tab_s + off, src + off, min(kPiece, bytes - off), &bar);
    }
    staged_slot = slot;
    copy_pending = true;
  };
  auto stage_wait_all = [&]() {
    if (copy_pending) {
      mbar_wait(&bar, bar_parity);
      bar_parity ^= 1;
      copy_pending = false;
    }
  };
  // Work units of a phase (32 nodes or checks of one degree, in all live columns) are handed out by an atomic counter.
  // A phase uses counter (phase & 1) and thread 0 zeroes the other one meanwhile: the barrier that ends the phase orders
  // that store before the next phase's first fetch, and no warp can still be fetching from a counter when it is zeroed
  // (its phase ended two barriers ago).
  unsigned int phase = 0;
  auto begin_phase = [&]() {
    ++phase;
    if (tid == 0) s_next[(phase & 1) ^ 1] = 0;
  };
  auto next_unit = [&]() -> int {
    int u = 0;
    if (lane == 0) u = (int)atomicAdd(&s_next[phase & 1], 1u);
    return __shfl_sync(0xFFFFFFFFu, u, 0);
  };
  auto live_mask = [&]() -> uint32_t {  // columns with an unfinished frame
    uint32_t m = 0;
    for (int c = 0; c < C; ++c) m |= (s_done[c] != kNibOnes ? 1u : 0u) << c;
    return m;
  };

  // syndrome of the hard decisions over all checks of every live column (syndrome_check(const bvec&), :455-469)
  auto syndrome = [&](uint32_t live) {
    begin_phase();
    const int n_units = (P.nchk + 31) >> 5;
    for (int u = next_unit(); u < n_units; u = next_unit()) {
      const int c = u * 32 + lane;
      uint32_t x[kOnchipMaxCols];
#pragma unroll
      for (int col = 0; col < kOnchipMaxCols; ++col) x[col] = 0;
      if (c < P.nchk) {
        const int s = __ldg(P.cn_start + c), e = __ldg(P.cn_start + c + 1);
        for (int k = s; k < e; ++k) {
          const int v = __ldg(P.chk_equ_idx + k);
#pragma unroll
          for (int col = 0; col < kOnchipMaxCols; ++col)
            if (col < C && ((live >> col) & 1)) x[col] ^= cv.hard(col)[v];
        }
      }
      ColFlags f;
#pragma unroll
      for (int col = 0; col < kOnchipMaxCols; ++col)
        if (col < C) f.add(col, byte_to_nib_bits(x[col]));
      warp_flags_or(s_syn_fail, f, C);
    }
  };
  // one node through the program interpreter in one column: variable update of any tree shape, or the decision
  // (dec_update_lut, :340-344; hard bits of the frames in done_byte stay)
  auto interp_node = [&](const uint8_t* img, int2 nd, int col, uint32_t done_byte, bool decision, ColFlags& fail) {
    const DevProgHeader* hdr = reinterpret_cast<const DevProgHeader*>(img);
    const DevOp* ops = reinterpret_cast<const DevOp*>(img + hdr->ops_off);
    const uint8_t* outs = img + hdr->outs_off;
    const uint8_t* tables = img + hdr->tables_off;
    const int n_in = (int)hdr->n_in, n_out = (int)hdr->n_out, so = (int)hdr->q_out - 1;
    uint32_t* ms = cv.msgs(col);
    OcVals vals{slots_s + tid, P.slot_threads};  // callers have tid < slot_threads
    for (int k = 0; k < n_in; ++k) vals.set(k, ms[skew((uint32_t)nd.y + k)]);
    vals.set(n_in, cv.cha(col)[nd.x]);
    run_program(hdr->n_ops, ops, tables, vals);
    if (decision) {
      const uint32_t nb = nib_bits_to_byte(vals.get(outs[0]), so);  // bit = (root output < 1), :342
      uint8_t* hs = cv.hard(col);
      hs[nd.x] = (uint8_t)((hs[nd.x] & done_byte) | (nb & ~done_byte));
      return;
    }
    uint32_t A = 0xFFFFFFFFu, O = 0;
    for (int k = 0; k < n_out; ++k) {
      const uint32_t r = vals.get(outs[k]);
      ms[skew((uint32_t)nd.y + k)] = r;
      A &= r;
      O |= r;
    }
    fail.add(col, (A ^ O) >> so);
  };

  for (;;) {
    if (tid == 0) s_group = (int)atomicAdd(P.queue, 1u);
    __syncthreads();
    const int w0 = s_group * C;
    if (w0 >= P.W) break;
    // ---- load the group: LLRout = (cha < Nq_Cha/2) (:275), msgs[e] = msg_in[v(e)] (:284-289)
    for (int v = tid; v < P.nvar; v += kOnchipThreads) {
      const int e0 = __ldg(P.vn_edge0 + v), e1 = __ldg(P.vn_edge0 + v + 1);
      for (int col = 0; col < C; ++col) {
        const int w = w0 + col;
        if (w >= P.W) break;
        const uint32_t c = P.cha[(size_t)v * P.W + w], m0 = P.msg0[(size_t)v * P.W + w];
        cv.cha(col)[v] = c;
        cv.hard(col)[v] = (uint8_t)nib_bits_to_byte(c, P.q_cha - 1);
        uint32_t* ms = cv.msgs(col);
        for (int e = e0; e < e1; ++e) ms[skew((uint32_t)e)] = m0;
      }
    }
    if (tid < C) {
      // padding frames of a ragged batch are flagged done by the reset kernel; columns beyond the batch are all done
      s_done[tid] = w0 + tid < P.W ? P.done[w0 + tid] : kNibOnes;
    }
    if (tid < 2) s_vn_fail[tid] = s_cn_fail[tid] = s_syn_fail[tid] = 0;
    if (tid < 8 * C) s_iters[tid] = 0;
    __syncthreads();
    uint32_t live = live_mask();
    if (P.pisc && live) {  // :277
      syndrome(live);
      __syncthreads();
      if (tid == 0) {
        for (int c = 0; c < C; ++c) s_done[c] |= ~col_flags_of(s_syn_fail, c) & kNibOnes;  // these frames return 0
        s_syn_fail[0] = s_syn_fail[1] = 0;
      }
      __syncthreads();
      live = live_mask();
    }
    for (int ii = 0; ii < P.max_iters && live; ++ii) {
      const int want = P.psc && ii > 0;
      if (ii != P.max_iters - 1) stage(__ldg(P.iter_slot + ii));  // lands while the check pass runs
      const QMask qm = make_qmask(__ldg(P.iter_q + ii));
      // ---- check-node pass, min-sum (:304-310, :355-402); sign parities = syndrome of VN pass ii-1
      begin_phase();
      for (int u = next_unit(); u < P.n_cn_items; u = next_unit()) {
        const int4 it = __ldg(P.cn_items + u);  // (class, first check, count, -)
        const int dc = P.cc[it.x].dc;
        ColFlags f;
        if (lane < it.z) {
          const int32_t* r = P.cc[it.x].rows + (size_t)(it.y + lane) * dc;
          switch (dc) {
#define X(D)                                   \
  case D:                                      \
    oc_cn_exact<D>(r, cv, C, live, qm, f);     \
    break;
            LUTLDPC_EXACT_DCS(X)
#undef X
            default:
              oc_cn_any(r, dc, cv, C, live, qm, f);
          }
        }
        if (want) warp_flags_or(s_cn_fail, f, C);
      }
      __syncthreads();
      if (want) {  // psc: VN pass ii-1 unanimous and syndrome zero -> return ii (:327)
        if (tid == 0) {  // one thread: the flag words are shared by the columns and are cleared here, before the barrier
          for (int c = 0; c < C; ++c) {
            const uint32_t done = s_done[c];
            const uint32_t conv = ~done & ~(col_flags_of(s_vn_fail, c) | col_flags_of(s_cn_fail, c)) & kNibOnes;
            for (int j = 0; j < 8; ++j)
              if ((conv >> (4 * j)) & 1) s_iters[8 * c + j] = ii;
            s_done[c] = done | conv;
            s_conv[c] = conv;
          }
          s_vn_fail[0] = s_vn_fail[1] = s_cn_fail[0] = s_cn_fail[1] = 0;
        }
        __syncthreads();
        // Hard decisions of the frames that just returned.  Every check parity of such a frame is zero, so the sign
        // this check pass left on an edge (own sign ^ parity) is the sign the variable pass put there, and the VN was
        // unanimous: the sign of its first edge is the tentative bit syndrome_check stored in LLRout (:441-451).
        uint32_t any_conv = 0;
        for (int c = 0; c < C; ++c) any_conv |= s_conv[c];
        if (any_conv) {
          for (int v = tid; v < P.nvar; v += kOnchipThreads) {
            const uint32_t ix = skew((uint32_t)__ldg(P.vn_edge0 + v));
            for (int col = 0; col < C; ++col) {
              const uint32_t cb = nib_bits_to_byte(s_conv[col], 0);
              if (!cb) continue;
              uint8_t* hs = cv.hard(col);
              hs[v] = (uint8_t)((hs[v] & ~cb) | (nib_bits_to_byte(cv.msgs(col)[ix], qm.s) & cb));
            }
          }
          __syncthreads();  // the variable pass overwrites these messages
        }
        live = live_mask();
      }
      if (ii == P.max_iters - 1 || !live) break;
      // ---- variable-node pass (:320-325)
      stage_wait_all();
      begin_phase();
      for (int u = next_unit(); u < P.n_vn_items; u = next_unit()) {
        const int4 it = __ldg(P.vn_items + u);  // (class, first node, count, -)
        const OnchipClass& cl = P.vc[it.x];
        ColFlags f;
        if (lane < it.z) {
          const int2 nd = __ldg(cl.pairs + it.y + lane);  // (node, first edge)
          const uint8_t* img = tab_s + cl.img_off;
          const DevProgHeader* hdr = reinterpret_cast<const DevProgHeader*>(img);
          bool ran = false;
          if (cl.fast) {
            switch (cl.dv) {
#define X(D)                                                                                  \
  case D:                                                                                     \
    if constexpr (D <= MAXDV) {                                                               \
      oc_vn_fast<D>(img + hdr->tables_off, nd, (int)hdr->q_out, cv, C, live, f);              \
      ran = true;                                                                             \
    }                                                                                         \
    break;
              LUTLDPC_FAST_DVS(X)
#undef X
            }
          }
          if (!ran)
            for (int col = 0; col < C; ++col)
              if ((live >> col) & 1) interp_node(img, nd, col, 0, false, f);
        }
        if (P.psc) warp_flags_or(s_vn_fail, f, C);
      }
      __syncthreads();
    }
    if (live) {  // decision + final syndrome (:340-349); the warps that own value slots take the units
      stage(P.dec_slot);
      stage_wait_all();
      begin_phase();
      if (tid < P.slot_threads) {
        for (int u = next_unit(); u < P.n_vn_items; u = next_unit()) {
          const int4 it = __ldg(P.vn_items + u);
          const OnchipClass& cl = P.vc[it.x];
          ColFlags dummy;
          if (lane < it.z) {
            const int2 nd = __ldg(cl.pairs + it.y + lane);
            for (int col = 0; col < C; ++col)
              if ((live >> col) & 1) interp_node(tab_s + cl.dec_off, nd, col, nib_bits_to_byte(s_done[col], 0), true, dummy);
          }
        }
      }
      __syncthreads();
      syndrome(live);
      __syncthreads();
      if (tid < 8 * C) {
        const int col = tid >> 3, j = tid & 7;
        if (!((s_done[col] >> (4 * j)) & 1))
          s_iters[tid] = ((col_flags_of(s_syn_fail, col) >> (4 * j)) & 1) ? -P.max_iters : P.max_iters;
      }
    }
    stage_wait_all();  // a copy issued for an iteration that was not run must land before the buffer is reused
    __syncthreads();
    // ---- results of the group
    for (int v = tid; v < P.nvar; v += kOnchipThreads)
      for (int col = 0; col < C && w0 + col < P.W; ++col) P.hard[(size_t)v * P.W + w0 + col] = cv.hard(col)[v];
    if (tid < 8 * C && w0 + (tid >> 3) < P.W) P.iters[8 * w0 + tid] = s_iters[tid];
    if (tid < C && w0 + tid < P.W) P.done[w0 + tid] = kNibOnes;
    __syncthreads();  // state and flags are rewritten by the next group
  }
}

cudaError_t configure_onchip() {
  cudaError_t e;
#define X(M)                                                                                                \
  e = cudaFuncSetAttribute(k_onchip<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, kOnchipMaxSmem);       \
  if (e != cudaSuccess) return e;
  X(4) X(8) X(12) X(20)
#undef X
  return cudaSuccess;
}

cudaError_t launch_onchip(const OnchipParams& P, int max_fast_dv, int n_ctas, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(P.queue, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const int groups = (P.W + P.C - 1) / P.C;
  const int grid = std::max(1, std::min(n_ctas, groups));
  if (max_fast_dv <= 4) k_onchip<4><<<grid, kOnchipThreads, P.smem_bytes, st>>>(P);
  else if (max_fast_dv <= 8) k_onchip<8><<<grid, kOnchipThreads, P.smem_bytes, st>>>(P);
  else if (max_fast_dv <= 12) k_onchip<12><<<grid, kOnchipThreads, P.smem_bytes, st>>>(P);
  else k_onchip<20><<<grid, kOnchipThreads, P.smem_bytes, st>>>(P);
  return cudaGetLastError();
}

int onchip_ctas_per_sm(int max_fast_dv, size_t smem_bytes) {
  int n = 0;
  cudaError_t e;
  if (max_fast_dv <= 4) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<4>, kOnchipThreads, smem_bytes);
  else if (max_fast_dv <= 8) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<8>, kOnchipThreads, smem_bytes);
  else if (max_fast_dv <= 12) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<12>, kOnchipThreads, smem_bytes);
  else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<20>, kOnchipThreads, smem_bytes);
  return e == cudaSuccess ? n : 0;
}

}  // namespace lutldpc
