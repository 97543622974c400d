// Staggered half-iteration kernel: the min-sum CHECK pass of one column range and the LUT VARIABLE
// pass of another column range of the same batch in ONE launch (see kernels.hpp, StaggerParams).
//
// Why: on the N=64800 code the check pass moves 2 GB per launch with ~25 instructions per word-edge
// (HBM bound, issue slots idle) while the variable pass spends ~31 issue slots and 8 LDS per 8
// look-ups (LSU/issue bound, HBM at 30 %).  Run back to back, each leaves the other's resource idle.
// The two halves of a batch are independent frames, so half A can be in its check pass while half B is
// in its variable pass; interleaving the CTAs of both kinds puts both instruction mixes on every SM.
//
// Reference semantics are unchanged: chk_update_minsum (src/LDPC_Code_LUT.cpp:355-402) and
// var_update_lut / LUT_Tree::var_msg_update (src/LDPC_Code_LUT.cpp:404-414, src/LUT_Tree.cpp:774-790)
// per frame; only the order in which independent frames are visited differs.
#include <cstdlib>

#include "kdev.cuh"

namespace lutldpc {

// checks [r0, r1) of an exact-degree row table, VEC consecutive word columns starting at w
template <int DC, int VEC>
__device__ __forceinline__ void cn_exact_rows(const int32_t* __restrict__ rows, int r0, int r1, int W, int w,
                                              const QMask qm, uint32_t* __restrict__ msgs, uint32_t (&fail)[VEC]) {
  typedef typename WordVec<VEC>::T V;
  union U { V v; uint32_t u[VEC]; };
  for (int i = r0; i < r1; ++i) {
    const int32_t* r = rows + (size_t)i * DC;
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
    for (int j = 0; j < VEC; ++j) fail[j] |= a[j].par;
  }
}

// any degree: accumulate, then re-read (the lines are in L1/L2) and emit; few registers
template <int VEC>
__device__ __forceinline__ void cn_any_rows(const int32_t* __restrict__ rows, int dc, int r0, int r1, int W, int w,
                                            const QMask qm, uint32_t* __restrict__ msgs, uint32_t (&fail)[VEC]) {
  typedef typename WordVec<VEC>::T V;
  union U { V v; uint32_t u[VEC]; };
  for (int i = r0; i < r1; ++i) {
    const int32_t* r = rows + (size_t)i * dc;
    U x0, x1;
    x0.v = *reinterpret_cast<const V*>(msgs + (size_t)r[0] * W + w);
    x1.v = *reinterpret_cast<const V*>(msgs + (size_t)r[1] * W + w);
    CnAccG a[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) cng_init(a[j], x0.u[j], x1.u[j], qm);
    for (int k = 2; k < dc; ++k) {
      U x;
      x.v = *reinterpret_cast<const V*>(msgs + (size_t)r[k] * W + w);
#pragma unroll
      for (int j = 0; j < VEC; ++j) cng_accum(a[j], x.u[j], qm);
    }
    for (int k = 0; k < dc; ++k) {
      V* p = reinterpret_cast<V*>(msgs + (size_t)r[k] * W + w);
      U x, o;
      x.v = *p;
#pragma unroll
      for (int j = 0; j < VEC; ++j) o.u[j] = cng_emit(a[j], x.u[j], qm);
      *p = o.v;
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) fail[j] |= a[j].par;
  }
}

template <int MAXDV, int MAXDC, int VEC>
__global__ void __launch_bounds__(kNodeThreads) k_stagger(const StaggerParams P, const int4* __restrict__ items,
                                                           const int32_t* __restrict__ vn_edge0,
                                                           uint32_t* __restrict__ msgs, const uint32_t* __restrict__ cha,
                                                           uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
                                                           uint32_t* __restrict__ vn_fail,
                                                           uint32_t* __restrict__ cn_fail) {
  __shared__ __align__(128) uint8_t tabs[(MAXDV - 1) * 128];
  __shared__ uint64_t bar;
  const int4 it = items[blockIdx.x];
  const int W = P.W;
  if (it.x >= 8) {
    // ---------------- check pass
    const int cls = it.x - 8;
    const int dc = P.cn_dc[cls];
    const int32_t* rows = P.cn_rows[cls];
    const QMask qm = make_qmask(P.cn_q);
    const int base = P.cn_w0 + blockIdx.y * kStaggerCols;
    typedef typename WordVec<VEC>::T V;
    union U { V v; uint32_t u[VEC]; };
#pragma unroll 1
    for (int s = 0; s < kStaggerCols / (kNodeThreads * VEC); ++s) {
      const int w = base + (s * kNodeThreads + threadIdx.x) * VEC;
      if (w >= P.cn_w1) break;
      U dn;
      dn.v = *reinterpret_cast<const V*>(done + w);
      bool all_done = true;
#pragma unroll
      for (int j = 0; j < VEC; ++j) all_done &= dn.u[j] == kNibOnes;
      if (all_done) continue;
      uint32_t fail[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) fail[j] = 0;
      switch (dc <= MAXDC ? dc : 0) {
#define X(D)                                                                                   \
  case D:                                                                                      \
    if constexpr (D <= MAXDC) cn_exact_rows<D, VEC>(rows, it.y, it.y + it.z, W, w, qm, msgs, fail); \
    break;
        LUTLDPC_EXACT_DCS(X)
#undef X
        default:
          cn_any_rows<VEC>(rows, dc, it.y, it.y + it.z, W, w, qm, msgs, fail);
      }
      if (P.cn_want) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) flag_or(&cn_fail[w + j], (fail[j] >> qm.s) & kNibOnes);
      }
    }
    return;
  }
  // ---------------- variable pass
  if (P.vn_w0 + (int)blockIdx.y * kStaggerCols >= P.vn_w1) return;
  const FusedVnClass cls = P.c[it.x];
  stage_to_smem(tabs, cls.tables, (uint32_t)(cls.dv - 1) * 128u, &bar);
#pragma unroll 1
  for (int s = 0; s < kStaggerCols / kNodeThreads; ++s) {
    const int w = P.vn_w0 + blockIdx.y * kStaggerCols + s * kNodeThreads + threadIdx.x;
    if (w >= P.vn_w1) break;
    const uint32_t dn = done[w];
    if (dn == kNibOnes) continue;
    const uint32_t done_byte = nib_bits_to_byte(dn, 0);
    uint32_t failacc = 0;
    switch (cls.dv) {
#define X(D)                                                                                                    \
  case D:                                                                                                       \
    if constexpr (D <= MAXDV)                                                                                   \
      vn_fast_body<D>(tabs, cls.node_list, it.y, it.y + it.z, vn_edge0, W, w, msgs, cha, hard, done_byte,       \
                      P.vn_want, cls.q_out, failacc);                                                           \
    break;
      LUTLDPC_FAST_DVS(X)
#undef X
    }
    if (P.vn_want) flag_or(&vn_fail[w], failacc & kNibOnes);
  }
}

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s && *s ? atoi(s) : dflt;
}

// exact-degree (fully unrolled) check code is instantiated up to kStaggerMaxExactDc; larger or rare degrees
// take the re-reading loop, which keeps the register count of the whole kernel low for the VN items
constexpr int kStaggerMaxExactDc = 8;
int stagger_max_exact_dc() { return kStaggerMaxExactDc; }
bool stagger_supported(int max_dv) { return max_dv >= 2 && max_dv <= 20; }

cudaError_t launch_stagger(const StaggerParams& P, int max_dv, const int4* d_items, int n_items, int n_colblocks,
                           const GraphArrays& g, const BatchBuffers& b, cudaStream_t st) {
  static const int vec = env_int("LUTLDPC_STAGGER_VEC", 2);
  dim3 grid(n_items, n_colblocks);
#define ARGS P, d_items, g.vn_edge0, b.msgs, b.cha, b.hard, b.done, b.vn_fail, b.cn_fail
  if (max_dv <= 8) {
    if (vec == 4) k_stagger<8, kStaggerMaxExactDc, 4><<<grid, kNodeThreads, 0, st>>>(ARGS);
    else k_stagger<8, kStaggerMaxExactDc, 2><<<grid, kNodeThreads, 0, st>>>(ARGS);
  } else {
    k_stagger<20, kStaggerMaxExactDc, 2><<<grid, kNodeThreads, 0, st>>>(ARGS);
  }
#undef ARGS
  return cudaGetLastError();
}

}  // namespace lutldpc
