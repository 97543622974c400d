// Counter-based random numbers shared by the Monte-Carlo front ends (sim.cu: LUT decoder, bp.cu: BP comparison
// decoder): the same (seed, frame, symbol) gives the same noise sample and the same information bit in both, so the
// two decoders can be compared on identical channel realisations (the reference seeds both from the same IT++ RNG,
// src/LDPC_BER_Sim.cpp:129).
#pragma once
#include <cstdint>

namespace lutldpc {

// ---- Philox4x32-10 (Salmon et al., SC'11), counter = (c0..c3), key = (k0, k1)
static __device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
  uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
  uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
static __device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}
enum { STREAM_NOISE = 0, STREAM_INFO = 1 };

// Four standard normals for (frame, node quad q) = the noise of symbols 4q .. 4q+3 of that frame: one Philox block,
// two Box-Muller pairs (both outputs of each pair are used), all arithmetic in FP64.  The uniforms carry 32 bits:
// u = (c + 0.5) / 2^32 in (0,1), i.e. radii up to sqrt(-2 ln 2^-33) = 6.76 sigma on a grid far finer than the
// reference's own (IT++'s ziggurat spends 24 bits on the abscissa, itpp/itpp/base/random.cpp:690-720).
// A sample depends only on (seed, frame, symbol): results do not depend on batch size or on how frames are
// split over GPUs.
static __device__ __forceinline__ void normals4_for(uint64_t frame, uint32_t quad, uint32_t k0, uint32_t k1, double (&n)[4]) {
  uint32_t c[4] = {(uint32_t)frame, (uint32_t)(frame >> 32), quad, STREAM_NOISE};
  philox4x32_10(c, k0, k1);
  const double k2m32 = 1.0 / 4294967296.0;
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    const double u1 = ((double)c[2 * p] + 0.5) * k2m32;
    const double u2 = ((double)c[2 * p + 1] + 0.5) * k2m32;
    const double r = sqrt(-2.0 * log(u1));
    double sn, cs;
    sincospi(2.0 * u2, &sn, &cs);
    n[2 * p] = r * cs;
    n[2 * p + 1] = r * sn;
  }
}

// information bits: cw[k][w] bit j = randb for frame 8w+j, symbol k (k < K)
static __global__ void k_info_bits(uint8_t* __restrict__ cw, int K, int W, int F, uint64_t first_frame, uint32_t k0,
                            uint32_t k1) {
  int w = blockIdx.y * blockDim.x + threadIdx.x;
  int k = blockIdx.x;
  if (w >= W || k >= K) return;
  uint32_t byte = 0;
  for (int j = 0; j < 8; ++j) {
    int f = 8 * w + j;
    if (f >= F) break;
    uint64_t frame = first_frame + f;
    uint32_t c[4] = {(uint32_t)frame, (uint32_t)(frame >> 32), (uint32_t)(k >> 7), STREAM_INFO};
    philox4x32_10(c, k0, k1);
    byte |= ((c[(k >> 5) & 3] >> (k & 31)) & 1u) << j;
  }
  cw[(size_t)k * W + w] = (uint8_t)byte;
}

}  // namespace lutldpc
