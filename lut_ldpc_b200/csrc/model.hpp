// Host-side model of a designed LUT decoder and the compiler that turns the reference's LUT trees
// into straight-line "node programs" the kernels execute.
//
// What this mirrors in the reference (paths relative to the reference tree):
//   graph flattening      LDPC_Code_LUT::decoder_parameterization   src/LDPC_Code_LUT.cpp:488-541
//   tree slot selection   LDPC_Code_LUT::set_trees                  src/LDPC_Code_LUT.cpp:120-169
//   tree text format      operator>>(Array<Array<LUT_Tree>>)        src/LUT_Tree.cpp:847-927, :448-535
//   node evaluation       LUT_Tree_Node::var_msg_update / chk_msg_update   src/LUT_Tree.cpp:402-445
//   leave-one-out         LUT_Tree::var_msg_update / chk_msg_update / dec_update   src/LUT_Tree.cpp:774-820
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace lutldpc {

enum NodeType { NODE_IM = 0, NODE_ROOT = 1, NODE_MSG = 2, NODE_CHA = 3 };  // src/LUT_Tree.hpp:188-194

struct TreeNode {
  int type = 0, K = 0;
  std::vector<int> Q;      // half table (var/dec trees) or magnitude table (check trees)
  std::vector<int> child;  // indices into Tree::nodes
};
struct Tree {
  int tree_type = 0, num_leaves = 0;
  std::vector<TreeNode> nodes;  // pre-order, nodes[0] is the root
};
struct TreeArray {  // Array<Array<LUT_Tree>>: [iteration slot][degree slot]
  int n_iter = 0, n_deg = 0;
  std::vector<Tree> t;
  const Tree& at(int it, int dg) const { return t[(size_t)it * n_deg + dg]; }
};

// Everything LDPC_Code_LUT::save_code stores that the decode path reads (src/LDPC_Code_LUT.cpp:643-697).
struct DecoderModel {
  int nvar = 0, nchk = 0, nchk_lin_indep = 0, num_edges = 0, max_iters = 0, Nq_Cha = 0;
  bool minLUT = true;
  std::vector<int32_t> dv, dc, cn_msg_idx, Nq_Msg, reuse, cha2msg;
  std::vector<double> qb_Cha, qb_Msg;
  TreeArray var_trees, chk_trees;
  std::string var_tree_text, chk_tree_text;  // the reference's text serialisation as it came in (written back by save)
  int gen_rows = 0, gen_cols = 0;
  std::vector<uint8_t> gen;  // dense parity part P, codeword = [u, P*u]
  // derived (set_trees / decoder_parameterization)
  std::vector<int32_t> var_tree_idx_iter, var_tree_idx_degree, chk_tree_idx_degree;
  std::vector<int32_t> vn_edge0;   // [nvar+1] first edge of each VN (edges are VN-major)
  std::vector<int32_t> cn_start;   // [nchk+1] offsets into cn_msg_idx
  std::vector<int32_t> chk_equ_idx;  // VN index of every entry of cn_msg_idx
  int ninfo() const { return nvar - nchk_lin_indep; }
  double rate() const { return double(ninfo()) / nvar; }
};

bool parse_tree_array(const char* txt, size_t len, TreeArray& out, std::string& err);
bool load_bundle_file(const char* path, DecoderModel& m, std::string& err);
// the reference's codec file written by LDPC_Code_LUT::save_code (IT++ it_file, src/LDPC_Code_LUT.cpp:643-697)
bool load_it_codec_file(const char* path, DecoderModel& m, std::string& err);
// LDPC_Code_LUT::save_code (src/LDPC_Code_LUT.cpp:643-697): the same variables in the same order, as an IT++ it_file the
// reference's load_code (and lut_ldpc_vhdl, scripts/) reads.  The generator is not written: a file to which
// LDPC_Generator_Systematic::save has appended (it replaces "Fileversion" by IT++'s own, 3) is refused by load_code (:578-579).
bool save_it_codec_file(const char* path, const DecoderModel& m, std::string& err);
// IT++'s LDPC_Code::save_code / load_code file (itpp/itpp/comm/ldpc.cpp:1250-1322), the bp_codec.it of ber_sim's [BP]
// path, with the block LDPC_Generator_Systematic::save appends (:1002-1009): the decoder parameterisation of H as
// LDPC_Code holds it (:1676-1745) and the parity part G of the systematic generator, rows packed 8 columns per byte.
struct BpCodecFile {
  int nvar = 0, ncheck = 0, ncheck_lin_indep = 0;
  std::vector<int32_t> C, V, sumX1, sumX2, iind, jind;
  int gen_rows = 0, gen_cols = 0;        // 0: no generator stored
  std::vector<uint8_t> gen_bytes;        // [gen_rows][(gen_cols + 7) / 8], bit (c & 7) of byte c >> 3 = column c
};
bool load_bp_codec_file(const char* path, BpCodecFile& f, std::string& err);
bool save_bp_codec_file(const char* path, const BpCodecFile& f, std::string& err);
// either of the two, told apart by the file magic
bool load_decoder_file(const char* path, DecoderModel& m, std::string& err);
// fills the derived members and validates; false + err on unsupported/inconsistent input
bool finalize_model(DecoderModel& m, std::string& err);

// ---- storage code ---------------------------------------------------------------------------
// A label m in [0,Nq) (reference convention: m < Nq/2 <=> negative LLR <=> bit 1) is stored as a
// q-bit sign/magnitude code  c = sign << (q-1) | mag,  q = ceil(log2 Nq):
//   m <  Nq/2 : sign = 1, mag = Nq/2 - 1 - m        m >= Nq/2 : sign = 0, mag = m - Nq/2
// (src/LDPC_Code_LUT.cpp:368-374 computes exactly this pair inside chk_update_minsum).
int code_bits(int Nq);
int label_to_code(int m, int Nq);
int code_to_label(int c, int Nq);  // -1 for a code no label maps to

// ---- node programs ----------------------------------------------------------------------------
constexpr int kMaxOpInputs = 16;
enum OpKind : uint8_t { OP_NARROW = 0, OP_WIDE = 1 };  // index fits a byte lane / a 16-bit lane

struct DevOp {  // 64 bytes, read by the kernels from shared memory
  uint32_t table_off;   // byte offset of this op's table inside the program's table blob
  uint8_t n_in, kind, out_slot, q_out;
  uint8_t sym;          // var/dec: 1 = half table, sign of the last input folded out
  uint8_t chk;          // 1 = check-node op (magnitude index + parity)
  uint8_t idx_bits, pad0;
  uint8_t in_slot[kMaxOpInputs];
  uint8_t sh[kMaxOpInputs];
  uint8_t q[kMaxOpInputs];
  uint8_t pad1[4];
};
static_assert(sizeof(DevOp) == 64, "DevOp layout");

struct DevProgHeader {  // 64 bytes
  uint32_t n_ops, n_slots, n_in, n_out;  // n_in = number of message inputs (dv or dc); +1 channel slot for var/dec
  uint32_t ops_off, outs_off, tables_off, total_bytes;  // byte offsets from the header start
  uint32_t table_bytes, has_cha, q_in, q_out;
  uint32_t pad[4];
};
static_assert(sizeof(DevProgHeader) == 64, "DevProgHeader layout");

struct NodeProgram {
  int n_in = 0, n_out = 0, n_slots = 0, n_lookups_naive = 0;
  bool has_cha = false;
  std::vector<DevOp> ops;
  std::vector<uint8_t> out_slot;  // slot holding output ii
  std::vector<uint8_t> tables;
  int q_in = 0, q_out = 0;
  // the program before slot allocation: value ids 0..n_in-1 = message inputs, n_in = channel label,
  // n_in+1+i = result of op i (used to match compile-time specialised kernels, vn_fast.cuh)
  std::vector<int> raw_node;              // tree node of op i
  std::vector<std::vector<int>> raw_in;   // operand value ids of op i
  std::vector<int> raw_out;               // value id of output ii
  // serialised image: header | ops | outs | tables, 16-byte aligned pieces
  std::vector<uint8_t> image() const;
};

enum ProgMode { PROG_VAR = 0, PROG_DEC = 1, PROG_CHK = 2 };
// Compiles tree `t` for a node of degree `deg`:
//   PROG_VAR: deg outputs, output ii = tree over (inputs without ii) ++ channel   (LUT_Tree.cpp:774-790)
//   PROG_DEC: 1 output, tree over all inputs ++ channel                           (LUT_Tree.cpp:809-820)
//   PROG_CHK: deg outputs, output ii = tree over (inputs without ii)              (LUT_Tree.cpp:792-807)
// Nq_in = alphabet of the message inputs, Nq_cha = alphabet of the channel label.
bool compile_program(const Tree& t, int deg, ProgMode mode, int Nq_in, int Nq_cha, NodeProgram& out,
                     std::string& err);

}  // namespace lutldpc
