// TEST INFRASTRUCTURE ONLY.  Runs the frame loop of LDPC_BER_Sim::sim_snr_point (src/LDPC_BER_Sim.cpp:260-290)
// twice with the same IT++ seed: once with the reference decoder behind `itpp::Channel_Code* C`, once with
// LDPC_Code_LUT_B200 (LDPC_Code_LUT_B200.hpp, the C ABI of liblutldpc_b200.so) behind the same pointer type, and
// requires bit-identical `bitsout` for every frame and every exit configuration.  The decoder is designed from an
// alist by the reference's own design step, as LDPC_BER_Sim_LUT::load does (src/LDPC_BER_Sim.cpp:437-500).
//
//   b200_binding_test <alist> --design-thr s | --design-snr dB  --max-iter I --qcha q --qmsg q [--gen]
//                     [--tree filename=<ini>] [--init-mode qcha] --snr dB --frames n --seed s
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>

#include "LDPC_Code_LUT_B200.hpp"

using namespace itpp;
using namespace lut_ldpc;

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: b200_binding_test <alist> [options]\n"); return 2; }
  std::map<std::string, std::string> kv;
  for (int i = 2; i < argc; i++) {
    std::string s = argv[i];
    if (s.rfind("--", 0) != 0) continue;
    if (i + 1 < argc && std::string(argv[i + 1]).rfind("--", 0) != 0) { kv[s.substr(2)] = argv[i + 1]; ++i; }
    else kv[s.substr(2)] = "1";
  }
  auto num = [&](const char* k, double d) { return kv.count(k) ? atof(kv[k].c_str()) : d; };
  LDPC_Parity* H = new LDPC_Parity(argv[1], "alist");
  LDPC_Generator_Systematic* G = kv.count("gen") ? new LDPC_Generator_Systematic(H) : nullptr;
  LDPC_Code_LUT* cpu = new LDPC_Code_LUT(H, G);
  const int max_iter = (int)num("max-iter", 30);
  const double sigma2 = kv.count("design-thr") ? sqr(num("design-thr", 0))
                                               : itpp::pow10(-num("design-snr", 0) / 10) / (2 * cpu->get_rate());
  cpu->design_luts(kv.count("tree") ? kv["tree"] : std::string("auto_bin_balanced"), get_empirical_ensemble(*H), true,
                   sigma2, max_iter, zeros_b(max_iter), 1 << (int)num("qcha", 4), ones_i(max_iter) * (1 << (int)num("qmsg", 4)));
  if (kv.count("init-mode") && kv["init-mode"] == "qcha") LDPC_BER_Sim_LUT::set_qcha(*cpu);
  LDPC_Code_LUT_B200* gpu = new LDPC_Code_LUT_B200(*cpu);

  const int N = cpu->get_nvar(), K = cpu->get_ninfo();
  const int n_frames = (int)num("frames", 20);
  const double snr = num("snr", 2.0);
  const unsigned seed = (unsigned)num("seed", 0);
  long mismatched_frames = 0, frames_run = 0, bit_errors_cpu = 0;
  const bool cfgs[3][2] = {{true, true}, {true, false}, {false, true}};  // (psc, pisc); the third runs every iteration
  for (int c = 0; c < 3; ++c) {
    // the maintainer-facing call has the reference's own parameter order (src/LDPC_Code_LUT.hpp:186-188)
    cpu->set_exit_conditions(max_iter, cfgs[c][0], cfgs[c][1]);
    gpu->set_exit_conditions(max_iter, cfgs[c][0], cfgs[c][1]);
    std::vector<bvec> got[2];
    Channel_Code* arms[2] = {cpu, gpu};
    for (int a = 0; a < 2; ++a) {
      Channel_Code* C = arms[a];  // what LDPC_BER_Sim holds (src/LDPC_BER_Sim.hpp:159)
      RNG_reset(seed);            // LDPC_BER_Sim::run, src/LDPC_BER_Sim.cpp:129
      const double N0 = itpp::pow10(-snr / 10.0) / C->get_rate();
      AWGN_Channel chan(N0 / 2);
      BPSK Mod;
      for (int ff = 0; ff < n_frames; ++ff) {
        bvec bitsin = G ? randb(K) : zeros_b(K);
        bvec coded = G ? C->encode(bitsin) : zeros_b(N);
        vec x = chan(Mod.modulate_bits(coded));
        vec softbits = Mod.demodulate_soft_bits(x, N0);
        bvec bitsout = C->decode(softbits);  // src/LDPC_BER_Sim.cpp:281
        got[a].push_back(bitsout);
        if (a == 0) for (int i = 0; i < K; ++i) bit_errors_cpu += bitsin(i) != bitsout(i);
      }
    }
    for (int ff = 0; ff < n_frames; ++ff) {
      ++frames_run;
      if (!(got[0][ff] == got[1][ff])) ++mismatched_frames;
    }
  }
  printf("{\"cmd\":\"binding_test\",\"nvar\":%d,\"ninfo\":%d,\"frames\":%ld,\"mismatched_frames\":%ld,"
         "\"cpu_data_bit_errors\":%ld,\"random_codewords\":%d}\n",
         N, K, frames_run, mismatched_frames, bit_errors_cpu, G ? 1 : 0);
  delete gpu;
  return mismatched_frames ? 1 : 0;
}
