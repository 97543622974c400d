// TEST INFRASTRUCTURE (includes reference headers) -- and at the same time the file a maintainer of
// mmeidlinger/lut_ldpc adds to use the B200 decoder: an itpp::Channel_Code subclass over the C ABI
// (include/lutldpc.h) that drops in wherever the reference holds its `Channel_Code* C`
// (src/LDPC_BER_Sim.hpp:159; call sites src/LDPC_BER_Sim.cpp:109,248,269,281).
//
// It is COMPILED here against the unmodified reference (oracle/Makefile -> _ref/b200_binding_test) and RUN by
// tests/test_gpu_binding.py: the reference's frame loop once with C = LDPC_Code_LUT*, once with
// C = LDPC_Code_LUT_B200*, same IT++ seed, bit-identical bitsout required.
//
// The designed decoder is private state of LDPC_Code_LUT (src/LDPC_Code_LUT.hpp:299-361).  In a maintained
// tree the export below would be a member `void LDPC_Code_LUT::export_b200(B200Export&) const` next to save_code
// (src/LDPC_Code_LUT.cpp:643-697); with the reference sources unmodified it goes through the friend
// declaration the class already carries (`friend class LDPC_BER_Sim_LUT`, src/LDPC_Code_LUT.hpp:68).
#pragma once
#include <itpp/itbase.h>
#include <itpp/itcomm.h>

#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "LDPC_Code_LUT.hpp"
#include "lutldpc.h"

namespace lut_ldpc {

struct B200Export {  // owns the arrays lutldpc_desc points into
  lutldpc_desc d;
  std::string var, chk;
  std::vector<int32_t> dv, dc, idx, nq, reuse, map;
  std::vector<double> qbc, qbm;
  std::vector<uint8_t> P;
};

// the friend of LDPC_Code_LUT (see header comment); a stand-in for src/LDPC_BER_Sim.cpp's class of that name, which
// this binary does not link
class LDPC_BER_Sim_LUT {
 public:
  static std::vector<int32_t> to32(const itpp::ivec& v) {
    std::vector<int32_t> o(v.size());
    for (int i = 0; i < v.size(); i++) o[i] = v(i);
    return o;
  }
  static void export_b200(const LDPC_Code_LUT& C, B200Export& x, bool with_generator) {
    std::ostringstream sv, sc;
    sv << C.var_trees;                    // Array<Array<LUT_Tree>>, src/LUT_Tree.cpp:855-864
    if (!C.minLUT) sc << C.chk_trees;
    x.var = sv.str();
    x.chk = sc.str();
    x.dv = to32(C.dv_vec);                // decoder_parameterization, src/LDPC_Code_LUT.cpp:501-528
    x.dc = to32(C.dc_vec);
    x.idx = to32(C.cn_msg_idx);
    x.nq = to32(C.Nq_Msg);
    x.map = to32(C.Nq_Cha_2_Nq_Msg_map);
    x.reuse.resize(C.reuse_vec.size());
    for (int i = 0; i < C.reuse_vec.size(); i++) x.reuse[i] = (int)C.reuse_vec(i);
    x.qbc.assign(C.qb_Cha._data(), C.qb_Cha._data() + C.qb_Cha.size());
    x.qbm.assign(C.qb_Msg._data(), C.qb_Msg._data() + C.qb_Msg.size());
    lutldpc_desc& d = x.d;
    std::memset(&d, 0, sizeof d);
    d.nvar = C.nvar; d.nchk = C.nchk; d.nchk_lin_indep = C.nchk_lin_indep; d.num_edges = C.num_edges;
    d.max_iters = C.max_iters; d.Nq_Cha = C.Nq_Cha; d.min_lut = C.minLUT;
    d.dv_vec = x.dv.data(); d.dc_vec = x.dc.data(); d.cn_msg_idx = x.idx.data();
    d.Nq_Msg = x.nq.data(); d.reuse_vec = x.reuse.data();
    d.Nq_Cha_2_Nq_Msg_map = (int)x.map.size() == C.Nq_Cha ? x.map.data() : nullptr;
    d.qb_Cha = x.qbc.data(); d.qb_Msg = x.qbm.data(); d.n_qb_Msg = (int32_t)x.qbm.size();
    d.var_tree_string = x.var.data(); d.var_tree_len = x.var.size();
    d.chk_tree_string = x.chk.empty() ? nullptr : x.chk.data(); d.chk_tree_len = x.chk.size();
    if (with_generator && C.G_defined) {  // dense parity part: column k = encode(e_k), itpp/itpp/comm/ldpc.cpp:1038
      const int K = C.nvar - C.nchk_lin_indep, rows = C.nvar - K;
      x.P.assign((size_t)rows * K, 0);
      itpp::bvec u = itpp::zeros_b(K), cw;
      for (int k = 0; k < K; k++) {
        u(k) = 1;
        C.G->encode(u, cw);
        for (int r = 0; r < rows; r++) x.P[(size_t)r * K + k] = (uint8_t)(int)cw(K + r);
        u(k) = 0;
      }
      d.gen_parity = x.P.data(); d.gen_rows = rows; d.gen_cols = K;
    }
  }
  // the mode enum is private to LDPC_Code_LUT; the reference sets it from this friend too (src/LDPC_BER_Sim.cpp:501-507)
  static bool qcha(const LDPC_Code_LUT& C) { return C.initial_message_mode == LDPC_Code_LUT::QCHA; }
  static void set_qcha(LDPC_Code_LUT& C) { C.set_initial_message_mode(LDPC_Code_LUT::QCHA); }
};

class LDPC_Code_LUT_B200 : public itpp::Channel_Code {
 public:
  // takes over the designed decoder of `cpu` (which must outlive this object: encode() stays on its generator)
  explicit LDPC_Code_LUT_B200(LDPC_Code_LUT& cpu) : cpu_(cpu), h_(nullptr) {
    B200Export x;
    LDPC_BER_Sim_LUT::export_b200(cpu, x, false);
    if (lutldpc_create(&x.d, &h_) != LUTLDPC_OK) it_error(lutldpc_last_error(nullptr));
    if (lutldpc_set_initial_message_mode(h_, LDPC_BER_Sim_LUT::qcha(cpu) ? LUTLDPC_INIT_QCHA : LUTLDPC_INIT_CONT) != LUTLDPC_OK)
      it_error(lutldpc_last_error(h_));
  }
  virtual ~LDPC_Code_LUT_B200() { lutldpc_destroy(h_); }

  // same signature and default values as LDPC_Code_LUT::set_exit_conditions (src/LDPC_Code_LUT.hpp:186-188)
  void set_exit_conditions(int max_iters, bool syndr_check_each_iter = true, bool syndr_check_at_start = false) {
    if (lutldpc_set_exit_conditions(h_, max_iters, syndr_check_each_iter, syndr_check_at_start) != LUTLDPC_OK)
      it_error(lutldpc_last_error(h_));
  }

  // itpp::Channel_Code (itpp/itpp/comm/channel_code.h:50-77)
  virtual void encode(const itpp::bvec& in, itpp::bvec& out) { cpu_.encode(in, out); }
  virtual itpp::bvec encode(const itpp::bvec& in) { itpp::bvec o; encode(in, o); return o; }
  virtual void decode(const itpp::bvec&, itpp::bvec&) { it_error("LDPC_Code_LUT::decode(): Hard input decoding not implemented"); }
  virtual itpp::bvec decode(const itpp::bvec&) { it_error("LDPC_Code_LUT::decode(): Hard input decoding not implemented"); return itpp::bvec(); }
  virtual void decode(const itpp::vec& llr, itpp::bvec& syst_bits) {  // src/LDPC_Code_LUT.cpp:204-239
    it_assert(llr.size() == cpu_.get_nvar(), "LDPC_Code_LUT::decode(): input length");
    std::vector<uint8_t> bits(llr.size());
    if (lutldpc_decode_llr(h_, 1, llr._data(), bits.data(), nullptr) != LUTLDPC_OK) it_error(lutldpc_last_error(h_));
    syst_bits.set_size(cpu_.get_ninfo());
    for (int i = 0; i < syst_bits.size(); i++) syst_bits(i) = bits[i];
  }
  virtual itpp::bvec decode(const itpp::vec& llr) { itpp::bvec o; decode(llr, o); return o; }
  virtual double get_rate() const { return cpu_.get_rate(); }

  // what a GPU is for: F frames per call, llr F x nvar row-major (frames as COLUMNS of an nvar x F itpp::mat)
  void decode_batch(int F, const double* llr, uint8_t* bits /*F x nvar*/, int32_t* iters /*F or NULL*/) {
    if (lutldpc_decode_llr(h_, F, llr, bits, iters) != LUTLDPC_OK) it_error(lutldpc_last_error(h_));
  }
  lutldpc_decoder* handle() { return h_; }

 private:
  LDPC_Code_LUT& cpu_;
  lutldpc_decoder* h_;
};

}  // namespace lut_ldpc
