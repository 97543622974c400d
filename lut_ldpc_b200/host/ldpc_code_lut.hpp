// Host-side mirror of the reference's decoder class for the decode path, over the C ABI
// (include/lutldpc.h).  Same names, argument meaning and error behaviour as
// lut_ldpc::LDPC_Code_LUT (src/LDPC_Code_LUT.hpp:86-285) behind an itpp::Channel_Code-shaped base
// (itpp/itpp/comm/channel_code.h:50-77), with std::vector in place of the IT++ containers, plus the
// batched calls a GPU needs (decode_batch, simulate_snr).  Errors are reported the way an IT++ built
// with ITPP_EXCEPTIONS does (it_error -> std::runtime_error, itpp/itpp/base/itassert.cpp:61-82).
//
// Header-only; link against lut_ldpc_b200/liblutldpc_b200.so.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/lutldpc.h"

namespace lut_ldpc_b200 {

typedef std::vector<double> vec;
typedef std::vector<uint8_t> bvec;  // one 0/1 byte per bit, like itpp::bvec
typedef std::vector<int> ivec;

// itpp::Channel_Code (itpp/itpp/comm/channel_code.h:50-77)
class Channel_Code {
 public:
  virtual ~Channel_Code() {}
  virtual void encode(const bvec& uncoded_bits, bvec& coded_bits) = 0;
  virtual bvec encode(const bvec& uncoded_bits) = 0;
  virtual void decode(const bvec& codedbits, bvec& decoded_bits) = 0;
  virtual bvec decode(const bvec& coded_bits) = 0;
  virtual void decode(const vec& received_signal, bvec& decoded_bits) = 0;
  virtual bvec decode(const vec& received_signal) = 0;
  virtual double get_rate() const = 0;
};

class LDPC_Code_LUT : public Channel_Code {
 public:
  enum initial_message_mode_t { CONT = LUTLDPC_INIT_CONT, QCHA = LUTLDPC_INIT_QCHA };  // src/LDPC_Code_LUT.hpp:76-79

  // LDPC_Code_LUT(const std::string& filename, LDPC_Generator*) (src/LDPC_Code_LUT.cpp:60-66): a stored
  // codec -- the reference's own lut_codec.it (save_code, :643-697) or the flat bundle of oracle/ref_driver.cpp.
  explicit LDPC_Code_LUT(const std::string& codec_filename) : h_(nullptr) {
    int rc = lutldpc_create_from_bundle(codec_filename.c_str(), &h_);
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Code_LUT::load_code(): ") + lutldpc_last_error(nullptr));
    fetch_info();
  }
  // from the variables LDPC_Code_LUT::save_code writes (src/LDPC_Code_LUT.cpp:643-697)
  explicit LDPC_Code_LUT(const lutldpc_desc& desc) : h_(nullptr) {
    int rc = lutldpc_create(&desc, &h_);
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Code_LUT::set_code(): ") + lutldpc_last_error(nullptr));
    fetch_info();
  }
  // LDPC_Generator_Systematic(H) + LDPC_Code_LUT(H, G) for the code of `other` (src/LDPC_BER_Sim.cpp:447-473): a
  // decoder for the column-permuted code that carries the systematic generator (lutldpc_create_systematic)
  struct with_generator_t {};
  LDPC_Code_LUT(const LDPC_Code_LUT& other, with_generator_t) : h_(nullptr) {
    int rc = lutldpc_create_systematic(other.h_, &h_, nullptr);
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Generator_Systematic::construct(): ") + lutldpc_last_error(nullptr));
    fetch_info();
  }
  ~LDPC_Code_LUT() override { lutldpc_destroy(h_); }
  LDPC_Code_LUT(const LDPC_Code_LUT&) = delete;
  LDPC_Code_LUT& operator=(const LDPC_Code_LUT&) = delete;

  // src/LDPC_Code_LUT.cpp:176-185
  void set_exit_conditions(int max_iters, bool syndr_check_each_iter = true, bool syndr_check_at_start = false) {
    ck(lutldpc_set_exit_conditions(h_, max_iters, syndr_check_each_iter, syndr_check_at_start), "set_exit_conditions");
    fetch_info();
  }
  void set_initial_message_mode(initial_message_mode_t m) { ck(lutldpc_set_initial_message_mode(h_, (int)m), "set_initial_message_mode"); }

  // src/LDPC_Code_LUT.cpp:643-697: the reference's own codec file (without the generator, see include/lutldpc.h)
  void save_code(const std::string& filename) const { ck(lutldpc_save_code(h_, filename.c_str()), "save_code"); }

  // ---- Channel_Code
  void encode(const bvec& input, bvec& output) override {  // src/LDPC_Code_LUT.cpp:187-197
    if ((int)input.size() != info_.ninfo) throw std::runtime_error("LDPC_Code_LUT::encode(): input length must equal ninfo");
    output.resize(info_.nvar);
    ck(lutldpc_encode(h_, 1, input.data(), output.data()), "encode");
  }
  bvec encode(const bvec& input) override { bvec o; encode(input, o); return o; }
  // hard-input decoding is not implemented by the reference either (it_error, src/LDPC_Code_LUT.hpp:196-203)
  void decode(const bvec&, bvec&) override { throw std::runtime_error("LDPC_Code_LUT::decode(): Hard input decoding not implemented"); }
  bvec decode(const bvec&) override { throw std::runtime_error("LDPC_Code_LUT::decode(): Hard input decoding not implemented"); }
  // src/LDPC_Code_LUT.cpp:204-246: LLRs in, the ninfo systematic bits out
  void decode(const vec& llr_in, bvec& syst_bits) override {
    if ((int)llr_in.size() != info_.nvar) throw std::runtime_error("LDPC_Code_LUT::decode(): input length must equal nvar");
    bvec all(info_.nvar);
    ck(lutldpc_decode_llr(h_, 1, llr_in.data(), all.data(), nullptr), "decode");
    syst_bits.assign(all.begin(), all.begin() + info_.ninfo);
  }
  bvec decode(const vec& llr_in) override { bvec o; decode(llr_in, o); return o; }
  double get_rate() const override { return info_.rate; }  // src/LDPC_Code_LUT.hpp:246-248

  // src/LDPC_Code_LUT.cpp:259-353; returns 0 / +n / +-max_iters
  int lut_decode(const ivec& LLRin_cha, const ivec& LLRin_msg, bvec& LLRout) {
    if ((int)LLRin_cha.size() != info_.nvar || (int)LLRin_msg.size() != info_.nvar)
      throw std::runtime_error("LDPC_Code_LUT::lut_decode(): input length must equal nvar");  // :262-269
    std::vector<uint8_t> c(LLRin_cha.begin(), LLRin_cha.end()), m(LLRin_msg.begin(), LLRin_msg.end());
    LLRout.resize(info_.nvar);
    int32_t it = 0;
    ck(lutldpc_decode_quantized(h_, 1, c.data(), m.data(), LLRout.data(), &it), "lut_decode");
    return it;
  }
  bool syndrome_check(const bvec& b) {  // src/LDPC_Code_LUT.cpp:455-469
    if ((int)b.size() != info_.nvar) throw std::runtime_error("LDPC_Code_LUT::syndrome_check(): input length must equal nvar");
    uint8_t ok = 0;
    ck(lutldpc_syndrome_check(h_, 1, b.data(), &ok), "syndrome_check");
    return ok != 0;
  }

  // ---- batched extensions (what the GPU is for)
  // n_frames x nvar LLRs (frame-major) -> n_frames x nvar hard decisions and lut_decode returns
  void decode_batch(int n_frames, const double* llr, uint8_t* bits, int32_t* iters) {
    ck(lutldpc_decode_llr(h_, n_frames, llr, bits, iters), "decode_batch");
  }
  void lut_decode_batch(int n_frames, const uint8_t* cha, const uint8_t* msg, uint8_t* bits, int32_t* iters) {
    ck(lutldpc_decode_quantized(h_, n_frames, cha, msg, bits, iters), "lut_decode_batch");
  }
  // the frame loop of LDPC_BER_Sim::sim_snr_point on the device; counters are added to acc
  void simulate_snr(const lutldpc_sim_params& p, lutldpc_sim_counters& acc) { ck(lutldpc_simulate_snr(h_, &p, &acc), "simulate_snr"); }

  // one batch of the same step, with the per-frame pairs LDPC_Code_LUT::decode prints at output_verbosity > 0
  void simulate_snr_stimuli(const lutldpc_sim_params& p, lutldpc_sim_counters& acc, uint8_t* llr_cha, uint8_t* out_bits) {
    ck(lutldpc_simulate_snr_stimuli(h_, &p, &acc, llr_cha, out_bits), "simulate_snr_stimuli");
  }

  // getters, src/LDPC_Code_LUT.hpp:246-285
  int get_nvar() const { return info_.nvar; }
  int get_nchk() const { return info_.nchk; }
  int get_ncheck() const { return info_.nchk; }
  int get_nchk_lin_indep() const { return info_.nchk_lin_indep; }
  int get_ninfo() const { return info_.ninfo; }
  int get_nrof_iterations() const { return info_.max_iters; }
  bool is_min_lut() const { return info_.min_lut != 0; }
  bool has_generator() const { return info_.has_gen != 0; }
  lutldpc_decoder* handle() { return h_; }

 private:
  void ck(int rc, const char* where) const {
    if (rc != LUTLDPC_OK) throw std::runtime_error(std::string("LDPC_Code_LUT::") + where + "(): " + lutldpc_last_error(h_));
  }
  void fetch_info() { ck(lutldpc_get_info(h_, &info_), "get_info"); }
  lutldpc_decoder* h_;
  lutldpc_info info_;
};

}  // namespace lut_ldpc_b200
