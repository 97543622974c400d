// GF(2) set-up steps on the device (gf2.cu): rank, systematic generator, bit-packed encoder.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "graph.hpp"

namespace lutldpc {

struct Gf2Result {
  int rank = 0;
  std::vector<int32_t> piv_col;   // pivot (parity) columns in the order they were found
  std::vector<int32_t> col_perm;  // new index of old column v: information columns first, then the pivot columns
  uint32_t* d_P = nullptr;        // device, [rank][P_words_per_row]: parity t = XOR of the information bits k with bit k set
  int P_words_per_row = 0;        // (caller owns d_P: cudaFree)
};

// Gauss-Jordan elimination of H over GF(2) on the current device; with want_generator the parity part P of the
// systematic generator of the column-permuted code is left on the device
bool gf2_systematic_device(const ParityRows& H, Gf2Result& out, bool want_generator, cudaStream_t st, std::string& err);
// cw rows [nvar][W] bytes (bit j of byte w = frame 8w + j): rows [K, K + rows) = P * rows [0, K)
cudaError_t launch_encode_bits(const uint32_t* d_P, int Wp, int rows, int K, int W, uint8_t* d_cw, cudaStream_t st);

}  // namespace lutldpc
