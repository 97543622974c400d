// Parity-check matrices on the host: the alist reader, the graph flattening that defines the decoder's edge numbering,
// and the GF(2) rank.  What this mirrors in the reference (paths relative to the reference tree):
//   alist format / reader     GF2mat_sparse_alist::read, to_sparse       itpp/itpp/base/gf2mat.cpp:46-210
//   LDPC_Parity(alist)        LDPC_Parity::import_alist                  itpp/itpp/comm/ldpc.cpp:167-180
//   edge numbering            LDPC_Code_LUT::decoder_parameterization    src/LDPC_Code_LUT.cpp:488-541
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace lutldpc {

struct ParityRows {  // H as its rows: sorted variable-node indices of every check
  int nvar = 0, nchk = 0;
  std::vector<std::vector<int32_t>> rows;
};

struct FlatGraph {  // decoder_parameterization's arrays
  int nvar = 0, nchk = 0, num_edges = 0;
  std::vector<int32_t> dv, dc, cn_msg_idx;
};

// H from an alist file.  Like the reference (to_sparse uses mlist only, gf2mat.cpp:193-202) the ROW lists define
// the matrix; the column lists are parsed and checked for consistency with them.
bool load_alist_file(const char* path, ParityRows& H, std::string& err);
// rows of H back from a flattened graph (the VN that owns every listed edge)
bool rows_from_graph(int nvar, int nchk, const std::vector<int32_t>& dv, const std::vector<int32_t>& dc,
                     const std::vector<int32_t>& cn_msg_idx, ParityRows& H, std::string& err);
// edges numbered VN-major, within a VN by ascending check index; cn_msg_idx lists, check by check, the edge ids in
// ascending VN order (src/LDPC_Code_LUT.cpp:513-528)
void flatten_graph(const ParityRows& H, FlatGraph& g);
// columns renumbered: new index of old column v = perm[v]
void permute_columns(const ParityRows& H, const std::vector<int32_t>& perm, ParityRows& out);
// GF(2) row rank by bit-packed elimination on the host (decoder_parameterization computes it with a dense
// GF2mat::row_rank for nvar < 1e5, :493-499); O(nchk^2 * nvar / 64) word operations
int gf2_rank_host(const ParityRows& H);

}  // namespace lutldpc
