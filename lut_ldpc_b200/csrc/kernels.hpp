// Launch interface between the C-ABI object (capi.cu) and the sm_100a kernels (kernels.cu).
// HBM layout of one batch of F frames (W = words per row, 8 frames per 32-bit word, W % 32 == 0):
//   msgs  [num_edges][W] u32   edge messages, edge-major (reference edge order), frames interleaved
//   cha   [nvar][W]      u32   channel labels (storage code)
//   hard  [nvar][W]      u8    hard decisions, bit j of byte (v,w) = frame 8w+j
//   done / vn_fail / cn_fail / syn_fail [W] u32   per-frame flags at bit 0 of each nibble
//   iters [8W] i32             lut_decode return value per frame
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace lutldpc {

constexpr int kNodeThreads = 128;  // word columns per CTA of the node kernels

struct BatchBuffers {
  uint32_t* msgs = nullptr;
  uint32_t* cha = nullptr;
  uint32_t* msg0 = nullptr;      // on-chip regime only: [nvar][W] initial message code per node instead of msgs
  uint8_t* hard = nullptr;
  uint32_t* done = nullptr;
  uint32_t* vn_fail = nullptr;
  uint32_t* cn_fail = nullptr;
  uint32_t* syn_fail = nullptr;
  int32_t* iters = nullptr;
  uint32_t* n_active = nullptr;  // device counters written by finalize: [0] unfinished frames, [1] unused,
                                 // [2] entries of cols, [3] entries of grp4
  // Work lists rebuilt by every finalize: word columns that still hold an unfinished frame, and groups of 4
  // adjacent columns (vector accesses) with any such word.  The hot kernels map thread i to cols[i] / grp4[i],
  // so the issue slots follow the number of live words instead of the number of live warps.
  uint32_t* cols = nullptr;      // [W]
  uint32_t* grp4 = nullptr;      // [W/4]
  int host_cols = 0, host_grp4 = 0;  // host-side upper bounds of the list lengths (grid sizing)
  uint32_t* mirror = nullptr;    // host-mapped copy of n_active[0..3], written by every whole-batch finalize (polls
                                 // read it without a copy: a D2H memcpy would queue behind the result transfers)
  int32_t* fid = nullptr;        // repacked layouts: original frame index of every slot (-1 padding); nullptr = identity
  int W = 0;                     // words per row
  int F = 0;                     // real frames (<= 8W)
};

struct GraphArrays {  // device copies of the flattened graph
  const int32_t* vn_edge0 = nullptr;   // [nvar+1]
  const int32_t* dv = nullptr;         // [nvar]
  const int32_t* cn_start = nullptr;   // [nchk+1]
  const int32_t* cn_msg_idx = nullptr; // [E]
  const int32_t* chk_equ_idx = nullptr;// [E]
  int nvar = 0, nchk = 0, num_edges = 0, max_dc = 0;
};

struct NodeLaunch {       // one degree class of one pass
  const uint8_t* prog = nullptr;   // device image: DevProgHeader | ops | outs | tables
  uint32_t prog_bytes = 0, n_slots = 0, n_in = 0;
  const int32_t* node_list = nullptr;  // [n_nodes] VN (or check) indices of this class
  const int2* node_pairs = nullptr;    // [n_nodes] (VN index, first edge) of the same nodes (specialised VN pass)
  int n_nodes = 0;
};

enum NodeMode { NODE_VAR = 0, NODE_DEC = 1, NODE_CHK = 2 };
enum FinalizeMode { FIN_INIT = 0, FIN_ITER = 1, FIN_LAST = 2 };

// pack: frame-major labels -> cha/msgs/hard; q_cha/q_msg = code widths; Nq_* = alphabets.  d_msg_lab == nullptr:
// d_cha_lab holds both labels of a symbol in one byte (low nibble channel, high nibble message)
cudaError_t launch_pack(const uint8_t* d_cha_lab, const uint8_t* d_msg_lab, const GraphArrays& g,
                        const BatchBuffers& b, int Nq_cha, int Nq_msg, cudaStream_t st);
cudaError_t launch_reset_flags(const BatchBuffers& b, cudaStream_t st);
cudaError_t launch_cn_minsum(const GraphArrays& g, const BatchBuffers& b, int q, int want_fail, cudaStream_t st);
// exact-degree min-sum pass over the checks whose edge rows are listed back to back in d_rows
bool cn_exact_supported(int dc);
cudaError_t launch_cn_exact(int dc, const int32_t* d_rows, int n_checks, const BatchBuffers& b, int q, int want_fail,
                            cudaStream_t st);
cudaError_t launch_node(NodeMode mode, const NodeLaunch& nl, const GraphArrays& g, const BatchBuffers& b,
                        int want_flags, cudaStream_t st);
// compile-time specialised VN pass (vn_fast.cuh); d_tables = the program image's table blob
struct NodeProgram;
bool fast_vn_matches(const NodeProgram& p);
cudaError_t launch_vn_fast(int dv, const uint8_t* d_tables, int q_out, const NodeLaunch& nl, const GraphArrays& g,
                           const BatchBuffers& b, int want_flags, cudaStream_t st);
// all specialised classes of a VN pass in one launch; tiles[i] = (class, first node, node count, 0)
struct FusedVnClass {
  const uint8_t* tables;
  const int2* nodes;  // (VN index, first edge id) per node of the class: one load instead of two dependent ones
  int dv, q_out;
};
struct FusedVnParams {
  FusedVnClass c[8];
};
cudaError_t launch_vn_fused(const FusedVnParams& P, int max_dv, const int4* d_tiles, int n_tiles, const GraphArrays& g,
                            const BatchBuffers& b, int want_flags, cudaStream_t st);
// finalize restricted to columns [w0, w1); *n_active receives the unfinished frames of that range
cudaError_t launch_finalize_range(FinalizeMode mode, int value, const BatchBuffers& b, int w0, int w1,
                                  uint32_t* n_active, cudaStream_t st);
cudaError_t launch_syndrome(const GraphArrays& g, const BatchBuffers& b, cudaStream_t st);
cudaError_t launch_finalize(FinalizeMode mode, int value, const BatchBuffers& b, cudaStream_t st);
// frame repacking (kernels.cu): gather the unfinished frames of `from` into the narrower layout `to`
// (flags, work lists, frame ids, msgs and cha rows); d_src = scratch [8 * to.W]
cudaError_t launch_repack(const GraphArrays& g, const BatchBuffers& from, const BatchBuffers& to, int32_t* d_src,
                          cudaStream_t st);
// results (hard bits, lut_decode returns) of the frames that finished in the repacked layout `from` back to the
// original layout `orig`; scratch: d_ow_first [2 * orig.W + 1] ints, d_ow_mask [2 * orig.W] bytes
cudaError_t launch_retire(const GraphArrays& g, const BatchBuffers& from, const BatchBuffers& orig, int32_t* d_ow_first,
                          uint8_t* d_ow_mask, cudaStream_t st);
cudaError_t launch_unpack(const GraphArrays& g, const BatchBuffers& b, uint8_t* d_out_bits, cudaStream_t st);
// hard -> bit-packed frame rows [F][ceil(nvar/8)], bit (v & 7) of byte v >> 3 = decision of node v
cudaError_t launch_unpack_bits(const GraphArrays& g, const BatchBuffers& b, uint8_t* d_out_packed, cudaStream_t st);
cudaError_t launch_extract_frame(const BatchBuffers& b, int num_edges, int frame, int Nq, uint8_t* d_out,
                                 cudaStream_t st);
// quant_nonlin front end: llr [F][nvar] f64 -> labels
cudaError_t launch_quantize(const double* d_llr, size_t n, const double* d_qb_cha, int n_qb_cha,
                            const double* d_qb_msg, int n_qb_msg, const int32_t* d_cha2msg, int use_map,
                            uint8_t* d_cha_lab, uint8_t* d_msg_lab, cudaStream_t st);
// hard-bit pack for syndrome_check on caller-supplied words: bits [F][nvar] -> hard
cudaError_t launch_pack_bits(const uint8_t* d_bits, const GraphArrays& g, const BatchBuffers& b, cudaStream_t st);
cudaError_t configure_kernels();  // opt-in shared memory sizes; call once per device

}  // namespace lutldpc
