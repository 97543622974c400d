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

// ---------------------------------------------------------------------------------------------
// Persistent, warp-specialised form.  2 CTAs per SM; every CTA keeps the LUTs of all degree classes of
// this iteration in shared memory.  Warps pull work from two global queues:
//   * VARIABLE tiles (a few nodes of one class x 128 word columns): vn_fast_body, values in registers;
//   * CHECK tiles (a few checks x 128 word columns, 4 columns per lane): the edge rows are fetched with
//     cp.async (LDGSTS, no registers held while in flight) into a per-warp ring of S rows, so two or
//     three check warps per CTA keep enough bytes in flight for the whole pass while the other ~22
//     warps spend the issue slots on look-ups.  Check warps join the variable queue when theirs is dry.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

constexpr int kRingBytes = 16384;             // S * dc * 512 B of edge rows per check warp
constexpr int kRingStride = kRingBytes + 512;  // + S * 32 edge indices

// checks [r0, r0+n) of a row table, 4 word columns per lane starting at w; any degree <= 32
template <int S>
__device__ __forceinline__ void cn_tile(const int32_t* __restrict__ rows, int dc, int r0, int n, int W, int w,
                                        bool active, const QMask qm, uint32_t* __restrict__ msgs, uint4* ring,
                                        int32_t* idx, uint32_t (&fail)[4]) {
  const int lane = threadIdx.x & 31;
  union U { uint4 v; uint32_t u[4]; };
  auto issue = [&](int i, int slot) {
    if (i < n) {
      if (lane < dc) idx[slot * 32 + lane] = rows[(size_t)(r0 + i) * dc + lane];
      __syncwarp();
      if (active) {
        for (int k = 0; k < dc; ++k)
          cp_async16(&ring[(slot * dc + k) * 32 + lane], msgs + (size_t)idx[slot * 32 + k] * W + w);
      }
    }
    cp_async_commit();
  };
#pragma unroll
  for (int s = 0; s < S; ++s) issue(s, s);
  for (int i = 0; i < n; ++i) {
    const int slot = i % S;
    cp_async_wait<S - 1>();
    if (active) {
      const uint4* rr = ring + (size_t)slot * dc * 32 + lane;
      U x0, x1;
      x0.v = rr[0];
      x1.v = rr[32];
      CnAccG a[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) cng_init(a[j], x0.u[j], x1.u[j], qm);
      for (int k = 2; k < dc; ++k) {
        U x;
        x.v = rr[k * 32];
#pragma unroll
        for (int j = 0; j < 4; ++j) cng_accum(a[j], x.u[j], qm);
      }
      for (int k = 0; k < dc; ++k) {
        U x, o;
        x.v = rr[k * 32];
#pragma unroll
        for (int j = 0; j < 4; ++j) o.u[j] = cng_emit(a[j], x.u[j], qm);
        __stcs(reinterpret_cast<uint4*>(msgs + (size_t)idx[slot * 32 + k] * W + w), o.v);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) fail[j] |= a[j].par;
    }
    __syncwarp();
    issue(i + S, slot);
  }
  cp_async_wait<0>();
}

__device__ __forceinline__ unsigned grab(unsigned int* ctr) {
  unsigned t = 0;
  if ((threadIdx.x & 31) == 0) t = atomicAdd(ctr, 1u);
  return __shfl_sync(0xFFFFFFFFu, t, 0);
}

template <int MAXDV>
struct StaggerCfg {
  static constexpr int kThreads = MAXDV <= 8 ? 640 : 512;  // x 2 CTAs per SM: 48 / 64 registers per thread
};

template <int MAXDV>
__global__ void __launch_bounds__(StaggerCfg<MAXDV>::kThreads, 2)
    k_stagger(const StaggerParams P, const int32_t* __restrict__ vn_edge0, uint32_t* __restrict__ msgs,
              const uint32_t* __restrict__ cha, uint8_t* __restrict__ hard, const uint32_t* __restrict__ done,
              uint32_t* __restrict__ vn_fail, uint32_t* __restrict__ cn_fail) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  const int W = P.W;
  const bool have_vn = P.vn_w1 > P.vn_w0, have_cn = P.cn_w1 > P.cn_w0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n_warps = blockDim.x >> 5;
  uint8_t* tabs = smem;
  uint8_t* rings = smem + ((P.tab_bytes + 127) & ~127);
  if (have_vn) {  // LUTs of every degree class of this iteration: bulk copies, one mbarrier
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_expect_tx(&bar, (uint32_t)P.tab_bytes);
      for (int j = 0; j < 8; ++j)
        if (P.c[j].dv) bulk_g2s(tabs + P.tab_off[j], P.c[j].tables, (uint32_t)(P.c[j].dv - 1) * 128u, &bar);
    }
  }
  // ---------------- check tiles (the last n_cn_warps warps of the CTA start here)
  if (have_cn && warp >= n_warps - P.n_cn_warps) {
    const int cw = warp - (n_warps - P.n_cn_warps);
    uint4* ring = reinterpret_cast<uint4*>(rings + (size_t)cw * kRingStride);
    int32_t* idx = reinterpret_cast<int32_t*>(rings + (size_t)cw * kRingStride + kRingBytes);
    const QMask qm = make_qmask(P.cn_q);
    const unsigned ncb = (unsigned)(P.cn_w1 - P.cn_w0 + 127) / 128u;
    const unsigned total = (unsigned)P.n_cn_items * ncb;
    for (;;) {
      const unsigned t = grab(P.ctr + 1);
      if (t >= total) break;
      const int4 it = P.cn_items[t / ncb];
      const int w = P.cn_w0 + (int)(t % ncb) * 128 + lane * 4;
      bool active = w < P.cn_w1;
      if (active) {
        const uint4 dn = *reinterpret_cast<const uint4*>(done + w);
        active = !(dn.x == kNibOnes && dn.y == kNibOnes && dn.z == kNibOnes && dn.w == kNibOnes);
      }
      uint32_t fail[4] = {0, 0, 0, 0};
      const int32_t* rows = P.cn_rows[it.x];
      const int dc = P.cn_dc[it.x];
      if (P.ring_rows == 4) cn_tile<4>(rows, dc, it.y, it.z, W, w, active, qm, msgs, ring, idx, fail);
      else if (P.ring_rows == 2) cn_tile<2>(rows, dc, it.y, it.z, W, w, active, qm, msgs, ring, idx, fail);
      else cn_tile<1>(rows, dc, it.y, it.z, W, w, active, qm, msgs, ring, idx, fail);
      if (P.cn_want && active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) flag_or(&cn_fail[w + j], (fail[j] >> qm.s) & kNibOnes);
      }
    }
  }
  if (!have_vn) return;
  // ---------------- variable tiles
  mbar_wait(&bar, 0);
  const unsigned ncb = (unsigned)(P.vn_w1 - P.vn_w0 + 127) / 128u;
  const unsigned total = (unsigned)P.n_vn_items * ncb;
  for (;;) {
    const unsigned t = grab(P.ctr);
    if (t >= total) break;
    const int4 it = P.vn_items[t / ncb];
    const FusedVnClass cls = P.c[it.x];
    const uint8_t* ctab = tabs + P.tab_off[it.x];
    const int wb = P.vn_w0 + (int)(t % ncb) * 128 + lane;
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
      const int w = wb + s * 32;
      if (w >= P.vn_w1) break;
      const uint32_t dn = done[w];
      if (dn == kNibOnes) continue;
      const uint32_t done_byte = nib_bits_to_byte(dn, 0);
      uint32_t failacc = 0;
      switch (cls.dv) {
#define X(D)                                                                                                 \
  case D:                                                                                                    \
    if constexpr (D <= MAXDV)                                                                                \
      vn_fast_body<D>(ctab, cls.nodes, it.y, it.y + it.z, W, w, msgs, cha, hard, done_byte,                 \
                      P.vn_want, cls.q_out, failacc);                                                        \
    break;
        LUTLDPC_FAST_DVS(X)
#undef X
      }
      if (P.vn_want) flag_or(&vn_fail[w], failacc & kNibOnes);
    }
  }
}

static constexpr int kStaggerMaxSmem = 160 * 1024;

bool stagger_supported(int max_dv, int max_dc) { return max_dv >= 2 && max_dv <= 20 && max_dc >= 2 && max_dc <= 32; }
int stagger_ring_rows(int max_dc) { return max_dc <= 8 ? 4 : max_dc <= 16 ? 2 : 1; }

cudaError_t configure_stagger() {
  cudaError_t e = cudaFuncSetAttribute(k_stagger<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStaggerMaxSmem);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k_stagger<20>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStaggerMaxSmem);
}

cudaError_t launch_stagger(const StaggerParams& P, int max_dv, int n_ctas, const GraphArrays& g, const BatchBuffers& b,
                           cudaStream_t st) {
  const size_t smem = (size_t)((P.tab_bytes + 127) & ~127) + (size_t)P.n_cn_warps * kRingStride;
  if (smem > (size_t)kStaggerMaxSmem) return cudaErrorInvalidConfiguration;
#define ARGS P, g.vn_edge0, b.msgs, b.cha, b.hard, b.done, b.vn_fail, b.cn_fail
  if (max_dv <= 8) k_stagger<8><<<n_ctas, StaggerCfg<8>::kThreads, smem, st>>>(ARGS);
  else k_stagger<20><<<n_ctas, StaggerCfg<20>::kThreads, smem, st>>>(ARGS);
#undef ARGS
  return cudaGetLastError();
}

}  // namespace lutldpc
