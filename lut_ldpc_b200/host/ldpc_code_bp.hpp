// Host-side mirror of IT++'s LDPC_Code for the decode path (itpp/itpp/comm/ldpc.h:757-1000), the BP comparison
// decoder ber_sim builds for a [BP] section (src/LDPC_BER_Sim.cpp:157-244), over the C ABI (lutldpc_bp_*,
// include/lutldpc.h).  Same names and argument meaning; std::vector in place of the IT++ containers; errors as an
// IT++ built with ITPP_EXCEPTIONS reports them (std::runtime_error).  Header-only.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "ldpc_code_lut.hpp"

namespace lut_ldpc_b200 {

typedef std::vector<int32_t> QLLRvec;

// LLR_calc_unit(Dint1, Dint2, Dint3, Dint4), itpp/itpp/comm/llr.h:143
struct LLR_calc_unit {
  short Dint1 = 12, Dint2 = 300, Dint3 = 7, Dint4 = 28;
  LLR_calc_unit() {}
  LLR_calc_unit(short d1, short d2, short d3, short d4 = 28) : Dint1(d1), Dint2(d2), Dint3(d3), Dint4(d4) {}
};

class LDPC_Code : public Channel_Code {
 public:
  // LDPC_Code(const LDPC_Parity* H[, LDPC_Generator* G]) (ldpc.cpp:1200-1215) + set_exit_conditions + set_llrcalc,
  // which ber_sim calls right after construction (src/LDPC_BER_Sim.cpp:197-201).  H is named by its alist file
  // (LDPC_Parity(filename, "alist")); with_generator builds LDPC_Generator_Systematic(H) first.
  LDPC_Code(const std::string& alist_filename, bool with_generator, int max_iters = 50, bool syndr_check_each_iter = true,
            bool syndr_check_at_start = false, const LLR_calc_unit& llrcalc = LLR_calc_unit())
      : h_(nullptr) {
    lutldpc_graph gr;
    if (lutldpc_alist_graph(alist_filename.c_str(), 0, &gr) != LUTLDPC_OK)
      throw std::runtime_error(std::string("LDPC_Parity::load_alist(): ") + lutldpc_last_error(nullptr));
    lutldpc_bp_params p;
    p.max_iters = max_iters; p.psc = syndr_check_each_iter; p.pisc = syndr_check_at_start;
    p.qllr_scale_res = llrcalc.Dint1; p.qllr_table_size = llrcalc.Dint2; p.qllr_spacing_res = llrcalc.Dint3;
    p.qllr_total_res = llrcalc.Dint4;
    col_perm_.resize(gr.nvar);
    int rc = lutldpc_bp_create(&gr, &p, with_generator, &h_, col_perm_.data());
    lutldpc_graph_free(&gr);
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Code::LDPC_Code(): ") + lutldpc_bp_last_error(nullptr));
    ck(lutldpc_bp_get_info(h_, &info_), "get_info");
  }
  // LDPC_Code(const std::string& filename, LDPC_Generator* G) (ldpc.cpp:1217-1222): a stored bp_codec.it, with its
  // generator if one was saved
  struct from_file_t {};
  LDPC_Code(const std::string& codec_filename, from_file_t, int max_iters = 50, bool syndr_check_each_iter = true,
            bool syndr_check_at_start = false, const LLR_calc_unit& llrcalc = LLR_calc_unit())
      : h_(nullptr) {
    lutldpc_bp_params p;
    p.max_iters = max_iters; p.psc = syndr_check_each_iter; p.pisc = syndr_check_at_start;
    p.qllr_scale_res = llrcalc.Dint1; p.qllr_table_size = llrcalc.Dint2; p.qllr_spacing_res = llrcalc.Dint3;
    p.qllr_total_res = llrcalc.Dint4;
    if (lutldpc_bp_create_from_file(codec_filename.c_str(), &p, &h_) != LUTLDPC_OK)
      throw std::runtime_error(std::string("LDPC_Code::load_code(): ") + lutldpc_bp_last_error(nullptr));
    ck(lutldpc_bp_get_info(h_, &info_), "get_info");
  }
  // ldpc.cpp:1283-1322
  void save_code(const std::string& filename) const { ck(lutldpc_bp_save_code(h_, filename.c_str()), "save_code"); }
  ~LDPC_Code() override { lutldpc_bp_destroy(h_); }
  LDPC_Code(const LDPC_Code&) = delete;
  LDPC_Code& operator=(const LDPC_Code&) = delete;

  // ---- Channel_Code
  void encode(const bvec& input, bvec& output) override {  // ldpc.cpp:1351-1358
    if (!info_.has_gen) throw std::runtime_error("LDPC_Code::encode(): LDPC Generator is required for encoding");
    if ((int)input.size() != info_.ninfo) throw std::runtime_error("LDPC_Code::encode(): input length must equal ninfo");
    output.resize(info_.nvar);
    ck(lutldpc_bp_encode(h_, 1, input.data(), output.data()), "encode");
  }
  bvec encode(const bvec& input) override { bvec o; encode(input, o); return o; }
  void decode(const bvec&, bvec&) override { throw std::runtime_error("LDPC_Code::decode(): Not implemented"); }
  bvec decode(const bvec&) override { throw std::runtime_error("LDPC_Code::decode(): Not implemented"); }
  // ldpc.cpp:1367-1374: syst_bits = qllrout.left(nvar - ncheck_lin_indep) < 0
  void decode(const vec& llr_in, bvec& syst_bits) override {
    if ((int)llr_in.size() != info_.nvar) throw std::runtime_error("LDPC_Code::bp_decode(): Wrong input dimensions");
    bvec all(info_.nvar);
    ck(lutldpc_bp_decode_llr(h_, 1, llr_in.data(), all.data(), nullptr), "decode");
    syst_bits.assign(all.begin(), all.begin() + info_.ninfo);
  }
  bvec decode(const vec& llr_in) override { bvec o; decode(llr_in, o); return o; }
  double get_rate() const override { return info_.rate; }

  // ldpc.cpp:1398-1622; returns +iter (valid codeword) / -iter, 0 when the initial syndrome check passes
  int bp_decode(const QLLRvec& LLRin, QLLRvec& LLRout) {
    if ((int)LLRin.size() != info_.nvar) throw std::runtime_error("LDPC_Code::bp_decode(): Wrong input dimensions");
    LLRout.resize(info_.nvar);
    int32_t it = 0;
    ck(lutldpc_bp_decode_qllr(h_, 1, LLRin.data(), LLRout.data(), &it), "bp_decode");
    return it;
  }
  // ---- batched extensions
  void decode_batch(int n_frames, const double* llr, uint8_t* bits, int32_t* iters) {
    ck(lutldpc_bp_decode_llr(h_, n_frames, llr, bits, iters), "decode_batch");
  }
  void simulate_snr(const lutldpc_sim_params& p, lutldpc_sim_counters& acc) { ck(lutldpc_bp_simulate_snr(h_, &p, &acc), "simulate_snr"); }

  int get_nvar() const { return info_.nvar; }
  int get_ncheck() const { return info_.nchk; }
  int get_ninfo() const { return info_.ninfo; }
  int get_nrof_iterations() const { return info_.max_iters; }
  bool has_generator() const { return info_.has_gen != 0; }
  const std::vector<int32_t>& column_permutation() const { return col_perm_; }  // H as G permuted it
  lutldpc_bp* handle() { return h_; }

 private:
  void ck(int rc, const char* where) const {
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Code::") + where + "(): " + lutldpc_bp_last_error(h_));
  }
  lutldpc_bp* h_;
  lutldpc_info info_;
  std::vector<int32_t> col_perm_;
};

}  // namespace lut_ldpc_b200
