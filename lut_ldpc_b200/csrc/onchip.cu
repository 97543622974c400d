// On-chip regime: LDPC_Code_LUT::lut_decode (src/LDPC_Code_LUT.cpp:259-353) for codes whose per-frame state fits
// shared memory (second regime: the (3,6) N=1000 code, the irregular N=500 code, the (6,32) N=2048 code).
//
// One CTA owns one word column = 8 frames and keeps ALL of their decoder state -- the E edge messages, the N channel
// labels, the N hard-decision bytes -- in shared memory for all iterations: the check pass and the variable pass are
// phases of one persistent kernel separated by __syncthreads(), the per-iteration LUTs arrive by cp.async.bulk (TMA
// engine) while the check pass runs, and the early-termination test (every VN unanimous + zero syndrome,
// src/LDPC_Code_LUT.cpp:437-452) is decided inside the CTA: per-thread flags are OR-reduced across the warp
// (__reduce_or_sync) and merged in shared memory.  A column that has finished is replaced by the next one of the
// batch from a global queue.  HBM traffic is the inputs (2 nibbles per symbol) and the outputs (1 bit per symbol):
// no per-iteration launches, no finalize kernel, no host poll, no repacking.
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

// OR over the CTA of per-thread flag words: warp reduction, then one shared-memory atomic per warp
__device__ __forceinline__ void cta_flag_or(uint32_t* s_flag, uint32_t f) {
  f = __reduce_or_sync(0xFFFFFFFFu, f);
  if ((threadIdx.x & 31) == 0 && f) atomicOr(s_flag, f);
}

template <int DV, int II>
struct OcStore {  // outputs of the compile-time program back to the node's edges; D = bits where outputs differ from the first
  static __device__ __forceinline__ void run(const uint32_t (&v)[DV + 1 + fastvn::Static<DV>::P.n_ops], uint32_t* msgs_s,
                                             uint32_t e0, uint32_t r0, uint32_t& D) {
    if constexpr (II < DV) {
      constexpr int kOut = fastvn::Static<DV>::P.out[II];
      const uint32_t r = v[kOut];
      msgs_s[skew(e0 + II)] = r;
      if constexpr (II > 0) D |= r ^ r0;
      OcStore<DV, II + 1>::run(v, msgs_s, e0, r0, D);
    }
  }
};

template <int DV>
__device__ __forceinline__ void oc_vn_fast(const uint8_t* tabs, const int2* __restrict__ pairs, int n, int q_out,
                                           uint32_t* msgs_s, const uint32_t* cha_s, uint8_t* hard_s, uint32_t done_byte,
                                           int want_flags, uint32_t& failacc) {
  constexpr int NOPS = fastvn::Static<DV>::P.n_ops;
  const int so = q_out - 1, root_shift = 4 - q_out;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int2 nd = __ldg(pairs + i);  // (node, first edge)
    uint32_t v[DV + 1 + NOPS];
#pragma unroll
    for (int k = 0; k < DV; ++k) v[k] = msgs_s[skew((uint32_t)nd.y + k)];
    v[DV] = cha_s[nd.x];
    fastvn::Exec<DV, 0>::run(v, tabs, root_shift);
    constexpr int kOut0 = fastvn::Static<DV>::P.out[0];
    const uint32_t r0 = v[kOut0];
    uint32_t D = 0;
    OcStore<DV, 0>::run(v, msgs_s, (uint32_t)nd.y, r0, D);
    if (want_flags) {
      failacc |= D >> so;
      const uint32_t nb = nib_bits_to_byte(r0, so);
      hard_s[nd.x] = (uint8_t)((hard_s[nd.x] & done_byte) | (nb & ~done_byte));
    }
  }
}

}  // namespace

template <int MAXDV>
__global__ void __launch_bounds__(256) k_onchip(const OnchipParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint32_t* msgs_s = reinterpret_cast<uint32_t*>(smem);
  uint32_t* cha_s = reinterpret_cast<uint32_t*>(smem + P.off_cha);
  uint8_t* hard_s = smem + P.off_hard;
  uint8_t* tab_s = smem + P.off_tab;
  uint32_t* slots_s = reinterpret_cast<uint32_t*>(smem + P.off_slots);
  __shared__ uint64_t bar;
  __shared__ uint32_t s_vn_fail, s_cn_fail, s_syn_fail, s_done;
  __shared__ int s_col;
  const int T = blockDim.x, tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  uint32_t bar_parity = 0;
  int staged_slot = -1;   // program images of this iteration slot are in tab_s
  bool copy_pending = false;
  __syncthreads();

  // tables / programs of iteration slot `slot`: one bulk copy, completion on the mbarrier.  Called by all threads
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

  // syndrome of hard_s over all checks (syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469)
  auto syndrome = [&]() {
    uint32_t acc = 0;
    for (int c = tid; c < P.nchk; c += T) {
      const int s = __ldg(P.cn_start + c), e = __ldg(P.cn_start + c + 1);
      uint32_t x = 0;
      for (int k = s; k < e; ++k) x ^= hard_s[__ldg(P.chk_equ_idx + k)];
      acc |= x;
    }
    cta_flag_or(&s_syn_fail, byte_to_nib_bits(acc));
  };

  for (;;) {
    if (tid == 0) s_col = (int)atomicAdd(P.queue, 1u);
    __syncthreads();
    const int w = s_col;
    if (w >= P.W) break;
    // ---- load the column: LLRout = (cha < Nq_Cha/2) (:275), msgs[e] = msg_in[v(e)] (:284-289)
    for (int v = tid; v < P.nvar; v += T) {
      const uint32_t c = P.cha[(size_t)v * P.W + w], m0 = P.msg0[(size_t)v * P.W + w];
      cha_s[v] = c;
      hard_s[v] = (uint8_t)nib_bits_to_byte(c, P.q_cha - 1);
      const int e0 = __ldg(P.vn_edge0 + v), e1 = __ldg(P.vn_edge0 + v + 1);
      for (int e = e0; e < e1; ++e) msgs_s[skew((uint32_t)e)] = m0;
    }
    if (tid == 0) {
      s_done = P.done[w];  // padding frames of a ragged batch are flagged done by the reset kernel
      s_vn_fail = s_cn_fail = s_syn_fail = 0;
    }
    int32_t my_iter = 0;  // thread j < 8: lut_decode return of frame 8w + j
    __syncthreads();
    if (P.pisc) {  // :277
      syndrome();
      __syncthreads();
      if (tid == 0) {
        s_done |= ~s_syn_fail & kNibOnes;
        s_syn_fail = 0;
      }
      __syncthreads();  // frames that return here keep my_iter = 0
    }
    int ii = 0;
    for (; ii < P.max_iters; ++ii) {
      const uint32_t done = s_done;
      if (done == kNibOnes) break;
      const int want = P.psc && ii > 0;
      if (ii != P.max_iters - 1) stage(__ldg(P.iter_slot + ii));  // lands while the check pass runs
      // ---- check-node pass, min-sum (:304-310, :355-402); sign parities = syndrome of VN pass ii-1
      {
        const QMask qm = make_qmask(__ldg(P.iter_q + ii));
        uint32_t failacc = 0;
        for (int k = 0; k < P.n_cc; ++k) {
          const int dc = P.cc[k].dc, n = P.cc[k].n;
          const int32_t* rows = P.cc[k].rows;
          for (int c = tid; c < n; c += T) {
            const int32_t* r = rows + (size_t)c * dc;
            CnAccG a;
            const uint32_t i0 = skew((uint32_t)__ldg(r)), i1 = skew((uint32_t)__ldg(r + 1));
            cng_init(a, msgs_s[i0], msgs_s[i1], qm);
            for (int j = 2; j < dc; ++j) cng_accum(a, msgs_s[skew((uint32_t)__ldg(r + j))], qm);
            for (int j = 0; j < dc; ++j) {
              const uint32_t ix = skew((uint32_t)__ldg(r + j));
              msgs_s[ix] = cng_emit(a, msgs_s[ix], qm);
            }
            failacc |= a.par;
          }
        }
        if (want) cta_flag_or(&s_cn_fail, (failacc >> qm.s) & kNibOnes);
      }
      __syncthreads();
      if (want) {  // psc: VN pass ii-1 unanimous and syndrome zero -> return ii (:327)
        const uint32_t conv = ~done & ~(s_vn_fail | s_cn_fail) & kNibOnes;
        if (tid < 8 && ((conv >> (4 * tid)) & 1)) my_iter = ii;
        __syncthreads();
        if (tid == 0) {
          s_done = done | conv;
          s_vn_fail = s_cn_fail = 0;
        }
        __syncthreads();
        if ((done | conv) == kNibOnes) break;
      }
      if (ii == P.max_iters - 1) break;
      // ---- variable-node pass (:320-325)
      stage_wait_all();
      {
        const uint32_t done_now = s_done;
        const uint32_t done_byte = nib_bits_to_byte(done_now, 0);
        uint32_t failacc = 0;
        for (int c = 0; c < P.n_vc; ++c) {
          const OnchipClass& cl = P.vc[c];
          const uint8_t* img = tab_s + cl.img_off;
          const DevProgHeader* hdr = reinterpret_cast<const DevProgHeader*>(img);
          bool ran = false;
          if (cl.fast) {
            switch (cl.dv) {
#define X(D)                                                                                                         \
  case D:                                                                                                            \
    if constexpr (D <= MAXDV) {                                                                                      \
      oc_vn_fast<D>(img + hdr->tables_off, cl.pairs, cl.n, (int)hdr->q_out, msgs_s, cha_s, hard_s, done_byte, P.psc,  \
                    failacc);                                                                                        \
      ran = true;                                                                                                    \
    }                                                                                                                \
    break;
              LUTLDPC_FAST_DVS(X)
#undef X
            }
          }
          if (!ran) {  // any other tree: node program interpreter, values in this thread's shared-memory column
            const DevOp* ops = reinterpret_cast<const DevOp*>(img + hdr->ops_off);
            const uint8_t* outs = img + hdr->outs_off;
            const uint8_t* tables = img + hdr->tables_off;
            const int n_in = (int)hdr->n_in, n_out = (int)hdr->n_out, so = (int)hdr->q_out - 1;
            OcVals vals{slots_s + tid, T};
            for (int i = tid; i < cl.n; i += T) {
              const int2 nd = __ldg(cl.pairs + i);
              for (int k = 0; k < n_in; ++k) vals.set(k, msgs_s[skew((uint32_t)nd.y + k)]);
              vals.set(n_in, cha_s[nd.x]);
              run_program(hdr->n_ops, ops, tables, vals);
              uint32_t A = 0xFFFFFFFFu, O = 0, first = 0;
              for (int k = 0; k < n_out; ++k) {
                const uint32_t r = vals.get(outs[k]);
                msgs_s[skew((uint32_t)nd.y + k)] = r;
                A &= r;
                O |= r;
                if (k == 0) first = r;
              }
              if (P.psc) {
                failacc |= (A ^ O) >> so;
                const uint32_t nb = nib_bits_to_byte(first, so);
                hard_s[nd.x] = (uint8_t)((hard_s[nd.x] & done_byte) | (nb & ~done_byte));
              }
            }
          }
        }
        if (P.psc) cta_flag_or(&s_vn_fail, failacc & kNibOnes);
      }
      __syncthreads();
    }
    if (s_done != kNibOnes) {  // decision + final syndrome (:340-349)
      const uint32_t done = s_done;
      const uint32_t done_byte = nib_bits_to_byte(done, 0);
      stage(P.dec_slot);
      stage_wait_all();
      for (int c = 0; c < P.n_vc; ++c) {
        const OnchipClass& cl = P.vc[c];
        const uint8_t* img = tab_s + cl.dec_off;
        const DevProgHeader* hdr = reinterpret_cast<const DevProgHeader*>(img);
        const DevOp* ops = reinterpret_cast<const DevOp*>(img + hdr->ops_off);
        const uint8_t* outs = img + hdr->outs_off;
        const uint8_t* tables = img + hdr->tables_off;
        const int n_in = (int)hdr->n_in, so = (int)hdr->q_out - 1;
        OcVals vals{slots_s + tid, T};
        for (int i = tid; i < cl.n; i += T) {
          const int2 nd = __ldg(cl.pairs + i);
          for (int k = 0; k < n_in; ++k) vals.set(k, msgs_s[skew((uint32_t)nd.y + k)]);
          vals.set(n_in, cha_s[nd.x]);
          run_program(hdr->n_ops, ops, tables, vals);
          const uint32_t nb = nib_bits_to_byte(vals.get(outs[0]), so);  // bit = (root output < 1), :342
          hard_s[nd.x] = (uint8_t)((hard_s[nd.x] & done_byte) | (nb & ~done_byte));
        }
      }
      __syncthreads();
      syndrome();
      __syncthreads();
      if (tid < 8 && !((done >> (4 * tid)) & 1)) my_iter = ((s_syn_fail >> (4 * tid)) & 1) ? -P.max_iters : P.max_iters;
    }
    stage_wait_all();  // a copy issued for an iteration that was not run must land before the buffer is reused
    // ---- results of the column
    for (int v = tid; v < P.nvar; v += T) P.hard[(size_t)v * P.W + w] = hard_s[v];
    if (tid < 8) P.iters[8 * w + tid] = my_iter;
    if (tid == 0) P.done[w] = kNibOnes;
    __syncthreads();  // hard_s / flags are rewritten by the next column
  }
}

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e && *e ? atoi(e) : dflt;
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

cudaError_t launch_onchip(const OnchipParams& P, int max_fast_dv, int threads, int n_ctas, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(P.queue, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const int grid = std::max(1, std::min(n_ctas, P.W));
  if (max_fast_dv <= 4) k_onchip<4><<<grid, threads, P.smem_bytes, st>>>(P);
  else if (max_fast_dv <= 8) k_onchip<8><<<grid, threads, P.smem_bytes, st>>>(P);
  else if (max_fast_dv <= 12) k_onchip<12><<<grid, threads, P.smem_bytes, st>>>(P);
  else k_onchip<20><<<grid, threads, P.smem_bytes, st>>>(P);
  return cudaGetLastError();
}

int onchip_ctas_per_sm(int max_fast_dv, int threads, size_t smem_bytes) {
  int n = 0;
  cudaError_t e;
  if (max_fast_dv <= 4) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<4>, threads, smem_bytes);
  else if (max_fast_dv <= 8) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<8>, threads, smem_bytes);
  else if (max_fast_dv <= 12) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<12>, threads, smem_bytes);
  else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_onchip<20>, threads, smem_bytes);
  (void)env_int;
  return e == cudaSuccess ? n : 0;
}

}  // namespace lutldpc
