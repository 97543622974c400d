// Device-side helpers shared by the kernel translation units.
#pragma once
#include <cstdint>

#include "core.cuh"
#include "kernels.hpp"
#include "vn_fast.cuh"

namespace lutldpc {

// ---------------------------------------------------------------------------------------------
// bulk (TMA engine) global->shared copy of a contiguous, 16-byte aligned block, completion on an
// mbarrier.  Used to stage a node program (op list + LUTs) and index slices into shared memory.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// Whole-CTA helper: thread 0 issues the copy (in <= 16 KiB pieces), everyone waits for the bytes.
// `bytes` must be a multiple of 16.  The barrier is single-use (parity 0).
// stage_issue + stage_wait: the wait can be placed after independent work (e.g. the first message loads)
__device__ __forceinline__ void stage_issue(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, bytes);
    const uint32_t kPiece = 16384;
    for (uint32_t off = 0; off < bytes; off += kPiece) {
      uint32_t n = bytes - off < kPiece ? bytes - off : kPiece;
      bulk_g2s((char*)dst + off, (const char*)src + off, n, bar);
    }
  }
}
__device__ __forceinline__ void stage_wait(uint64_t* bar) { mbar_wait(bar, 0); }
__device__ __forceinline__ void stage_to_smem(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  stage_issue(dst, src, bytes, bar);
  stage_wait(bar);
}

// Programmatic dependent launch (launch_pdl in kernels.cu): a kernel of the decode loop lets its successor's CTAs
// become resident as soon as all of its own have started (pdl_launch_dependents, first statement), and waits for
// its predecessor's memory (pdl_wait) after its index arithmetic and constant-data reads.  The launch latency and
// CTA ramp-up of one pass then overlap the drain of the previous one.  (Moving more of the prologue -- LUT staging,
// row tables -- in front of the wait cost more in registers than it gained.)  Both are no-ops in a kernel
// that was launched without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// OR `f` into a per-word flag that thousands of threads of the same pass target: look first (L2) and
// only issue the atomic when it would set a new bit.  Flags only ever gain bits within a pass, so a
// stale read can cost an extra atomic but never lose one.  Without this, the 32400 checks of the
// N=64800 code all hit the same few hundred addresses and the pass is bound by L2 atomic serialisation.
__device__ __forceinline__ void flag_or(uint32_t* p, uint32_t f) {
  if (f) {
    const uint32_t cur = __ldcg(p);
    if (f & ~cur) atomicOr(p, f);
  }
}


template <int VEC> struct WordVec;
template <> struct WordVec<1> { typedef uint32_t T; };
template <> struct WordVec<2> { typedef uint2 T; };
template <> struct WordVec<4> { typedef uint4 T; };

// message accesses of the specialised VN pass: streaming (evict-first) like the check pass -- every line is
// touched once per pass and the batch is far larger than L2
#ifndef LUTLDPC_VN_PLAIN_MEM
#define VN_LD(p) __ldcs(p)
#define VN_ST(p, v) __stcs(p, v)
#else
#define VN_LD(p) (*(p))
#define VN_ST(p, v) (*(p) = (v))
#endif

#define LUTLDPC_EXACT_DCS(X) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16)
#define LUTLDPC_FAST_DVS(X) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) X(17) X(18) X(19) X(20)

// Offsets into msgs / cha / hard are 32-bit element indices: a batch is bounded so that num_edges * W < 2^32
// (lutldpc_decoder::max_batch_frames), which keeps the address arithmetic of the hot loop to one IMAD +
// one IMAD.WIDE per access.
template <int DV, int II>
struct FastStore {
  static __device__ __forceinline__ void run(const uint32_t (&v)[DV + 1 + fastvn::Static<DV>::P.n_ops],
                                             uint32_t* __restrict__ msgs, uint32_t off0, uint32_t W, uint32_t r0,
                                             uint32_t& D) {
    if constexpr (II < DV) {
      const uint32_t r = v[fastvn::Static<DV>::P.out[II]];
      VN_ST(msgs + off0 + (uint32_t)II * W, r);
      if constexpr (II > 0) D |= r ^ r0;  // bits where an outgoing message differs from the first one
      FastStore<DV, II + 1>::run(v, msgs, off0, W, r0, D);
    }
  }
};

// per-thread work of the specialised VN pass: nodes [i0, i1) of the class list, word column w.
// Software-pipelined two nodes deep: while node i is evaluated, the messages of node i+1 are already in
// flight and the (node, first edge) pair of node i+2 is being fetched, so the dependent chain
// list -> messages (two global round trips) never sits in front of the look-ups.
template <int DV>
__device__ __forceinline__ void vn_fast_body(const uint8_t* tabs, const int2* __restrict__ nodes, int i0, int i1,
                                             int W, int w, uint32_t* __restrict__ msgs,
                                             const uint32_t* __restrict__ cha, uint8_t* __restrict__ hard,
                                             uint32_t done_byte, int want_flags, int q_out, uint32_t& failacc) {
  constexpr int NOPS = fastvn::Static<DV>::P.n_ops;
  const int so = q_out - 1, root_shift = 4 - q_out;
  const uint32_t uW = (uint32_t)W, uw = (uint32_t)w;
  if (i0 >= i1) return;
  uint32_t nxt[DV + 1];
  int2 cur = nodes[i0];
  int2 ahead = nodes[min(i0 + 1, i1 - 1)];
  {
    const uint32_t off0 = (uint32_t)cur.y * uW + uw;
#pragma unroll
    for (int k = 0; k < DV; ++k) nxt[k] = VN_LD(msgs + off0 + (uint32_t)k * uW);
    nxt[DV] = cha[(uint32_t)cur.x * uW + uw];
  }
  for (int i = i0; i < i1; ++i) {
    const uint32_t off0 = (uint32_t)cur.y * uW + uw;
    const uint32_t offn = (uint32_t)cur.x * uW + uw;
    uint32_t v[DV + 1 + NOPS];
#pragma unroll
    for (int k = 0; k <= DV; ++k) v[k] = nxt[k];
    cur = ahead;
    if (i + 1 < i1) {  // node i+1: its pair arrived during the previous iteration
      const uint32_t o1 = (uint32_t)cur.y * uW + uw;
#pragma unroll
      for (int k = 0; k < DV; ++k) nxt[k] = VN_LD(msgs + o1 + (uint32_t)k * uW);
      nxt[DV] = cha[(uint32_t)cur.x * uW + uw];
      ahead = nodes[min(i + 2, i1 - 1)];
    }
    fastvn::Exec<DV, 0>::run(v, tabs, root_shift);
    const uint32_t r0 = v[fastvn::Static<DV>::P.out[0]];
    uint32_t D = 0;
    FastStore<DV, 0>::run(v, msgs, off0, uW, r0, D);
    if (want_flags) {
      failacc |= D >> so;  // sign position of the outputs; the caller masks with kNibOnes
      uint8_t* hp = hard + offn;
      const uint32_t nb = nib_bits_to_byte(r0, so);
      // frozen bits of frames that have returned are kept; with none in the word (the common case) no read
      if (done_byte) *hp = (uint8_t)((*hp & done_byte) | (nb & ~done_byte));
      else *hp = (uint8_t)nb;
    }
  }
}


template <int DV, int II>
struct FastStore2 {
  static __device__ __forceinline__ void run(const uint32_t (&va)[DV + 1 + fastvn::Static<DV>::P.n_ops],
                                             const uint32_t (&vb)[DV + 1 + fastvn::Static<DV>::P.n_ops],
                                             uint32_t* __restrict__ msgs, uint32_t off0, uint32_t W, uint32_t ra0,
                                             uint32_t rb0, uint32_t& Da, uint32_t& Db) {
    if constexpr (II < DV) {
      constexpr int kOut = fastvn::Static<DV>::P.out[II];
      const uint32_t ra = va[kOut], rb = vb[kOut];
      VN_ST(reinterpret_cast<uint2*>(msgs + off0 + (uint32_t)II * W), make_uint2(ra, rb));
      if constexpr (II > 0) {
        Da |= ra ^ ra0;
        Db |= rb ^ rb0;
      }
      FastStore2<DV, II + 1>::run(va, vb, msgs, off0, W, ra0, rb0, Da, Db);
    }
  }
};

// Two word columns (16 frames) per thread: the per-node work that does not depend on the number of frames
// (pair fetch, address arithmetic, loop control) is shared by twice the look-ups and the two independent
// look-up chains give the scheduler more to overlap.  Used for the low degrees, where that fixed part is as
// large as the look-ups themselves and the values still fit the register budget.  Columns w, w+1 (w even).
template <int DV>
__device__ __forceinline__ void vn_fast_body2(const uint8_t* tabs, const int2* __restrict__ nodes, int i0, int i1,
                                              int W, int w, uint32_t* __restrict__ msgs,
                                              const uint32_t* __restrict__ cha, uint8_t* __restrict__ hard,
                                              uint32_t done_bytes, int want_flags, int q_out, uint32_t& fail0,
                                              uint32_t& fail1) {
  constexpr int NOPS = fastvn::Static<DV>::P.n_ops;
  const int so = q_out - 1, root_shift = 4 - q_out;
  const uint32_t uW = (uint32_t)W, uw = (uint32_t)w;
  if (i0 >= i1) return;
  uint2 nxt[DV + 1];
  int2 cur = nodes[i0];
  int2 ahead = nodes[min(i0 + 1, i1 - 1)];
  {
    const uint32_t off0 = (uint32_t)cur.y * uW + uw;
#pragma unroll
    for (int k = 0; k < DV; ++k) nxt[k] = VN_LD(reinterpret_cast<const uint2*>(msgs + off0 + (uint32_t)k * uW));
    nxt[DV] = *reinterpret_cast<const uint2*>(cha + (uint32_t)cur.x * uW + uw);
  }
  for (int i = i0; i < i1; ++i) {
    const uint32_t off0 = (uint32_t)cur.y * uW + uw;
    const uint32_t offn = (uint32_t)cur.x * uW + uw;
    uint32_t va[DV + 1 + NOPS], vb[DV + 1 + NOPS];
#pragma unroll
    for (int k = 0; k <= DV; ++k) {
      va[k] = nxt[k].x;
      vb[k] = nxt[k].y;
    }
    cur = ahead;
    if (i + 1 < i1) {
      const uint32_t o1 = (uint32_t)cur.y * uW + uw;
#pragma unroll
      for (int k = 0; k < DV; ++k) nxt[k] = VN_LD(reinterpret_cast<const uint2*>(msgs + o1 + (uint32_t)k * uW));
      nxt[DV] = *reinterpret_cast<const uint2*>(cha + (uint32_t)cur.x * uW + uw);
      ahead = nodes[min(i + 2, i1 - 1)];
    }
    fastvn::Exec<DV, 0>::run(va, tabs, root_shift);
    fastvn::Exec<DV, 0>::run(vb, tabs, root_shift);
    constexpr int kOut0 = fastvn::Static<DV>::P.out[0];
    const uint32_t ra0 = va[kOut0], rb0 = vb[kOut0];
    uint32_t Da = 0, Db = 0;
    FastStore2<DV, 0>::run(va, vb, msgs, off0, uW, ra0, rb0, Da, Db);
    if (want_flags) {
      fail0 |= Da >> so;
      fail1 |= Db >> so;
      uint16_t* hp = reinterpret_cast<uint16_t*>(hard + offn);
      const uint32_t nb = nib_bits_to_byte(ra0, so) | (nib_bits_to_byte(rb0, so) << 8);
      if (done_bytes) *hp = (uint16_t)((*hp & done_bytes) | (nb & ~done_bytes));
      else *hp = (uint16_t)nb;
    }
  }
}

}  // namespace lutldpc
