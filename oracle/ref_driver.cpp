// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Driver around the UNMODIFIED reference classes (compiled where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/lutldpc_ref).  It is the
// parity checker and the timed CPU baseline ("kind": "reference"):
//
//   design   alist -> LDPC_Code_LUT::design_luts -> reference codec file (.it, via the
//            reference's own save_code) + flat decoder bundle (.bundle, format below)
//   bundle   codec.it -> flat decoder bundle
//   fixture  codec.it -> noisy frames through the reference decoder: quantised inputs,
//            hard decisions, iteration counts and per-iteration message states
//   sim      codec.it -> the frame loop of LDPC_BER_Sim::sim_snr_point
//            (src/LDPC_BER_Sim.cpp:246-311) restated around the reference decoder,
//            with wall-clock timing around the loop
//
// Private state of LDPC_Code_LUT is reached through the friend declaration the
// reference itself carries (src/LDPC_Code_LUT.hpp:68, "friend class LDPC_BER_Sim_LUT"):
// this file defines its own class of that name instead of linking LDPC_BER_Sim.cpp
// (which needs boost::filesystem).  No reference source is copied.
#include <itpp/itbase.h>
#include <itpp/itcomm.h>
#include "LDPC_Code_LUT.hpp"
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

using namespace itpp;
using namespace lut_ldpc;

static void put_i32(std::ostream& os, int32_t v) { os.write((const char*)&v, 4); }
static void put_f64(std::ostream& os, double v) { os.write((const char*)&v, 8); }
static void put_ivec(std::ostream& os, const ivec& v) {
  for (int i = 0; i < v.length(); i++) put_i32(os, v(i));
}
static void put_str(std::ostream& os, const std::string& s) {
  put_i32(os, (int32_t)s.size());
  os.write(s.data(), s.size());
}

namespace lut_ldpc {
// Friend of LDPC_Code_LUT (see header comment).
class LDPC_BER_Sim_LUT {
 public:
  // Flat decoder bundle, little endian:
  //   char[8] "LUTLDPC1"
  //   i32 nvar nchk nchk_lin_indep num_edges max_iters Nq_Cha minLUT n_qb_msg has_gen reserved
  //   i32 dv_vec[nvar]  dc_vec[nchk]  cn_msg_idx[E]
  //   i32 Nq_Msg[max_iters]  reuse_vec[max_iters]  Nq_Cha_2_Nq_Msg_map[Nq_Cha]
  //   f64 qb_Cha[Nq_Cha-1]  qb_Msg[n_qb_msg]
  //   str var_tree_string  str chk_tree_string   (i32 length + bytes; the reference's own
  //                                               text serialisation, src/LUT_Tree.cpp:847-864)
  //   if has_gen: i32 rows cols ; u8 parity_part[rows][cols] (dense, 0/1), codeword = [u, P*u]
  static void write_bundle(const LDPC_Code_LUT& C, const std::string& fn, bool with_gen) {
    std::ofstream os(fn.c_str(), std::ios::binary);
    os.write("LUTLDPC1", 8);
    int K = C.nvar - C.nchk_lin_indep;
    bool has_gen = with_gen && C.G_defined;
    put_i32(os, C.nvar); put_i32(os, C.nchk); put_i32(os, C.nchk_lin_indep);
    put_i32(os, C.num_edges); put_i32(os, C.max_iters); put_i32(os, C.Nq_Cha);
    put_i32(os, C.minLUT ? 1 : 0); put_i32(os, C.qb_Msg.length());
    put_i32(os, has_gen ? 1 : 0); put_i32(os, 0);
    put_ivec(os, C.dv_vec); put_ivec(os, C.dc_vec); put_ivec(os, C.cn_msg_idx);
    put_ivec(os, C.Nq_Msg);
    for (int i = 0; i < C.reuse_vec.length(); i++) put_i32(os, (int)C.reuse_vec(i));
    put_ivec(os, C.Nq_Cha_2_Nq_Msg_map);
    for (int i = 0; i < C.qb_Cha.length(); i++) put_f64(os, C.qb_Cha(i));
    for (int i = 0; i < C.qb_Msg.length(); i++) put_f64(os, C.qb_Msg(i));
    std::stringstream vs, cs;
    vs << C.var_trees;
    cs << C.chk_trees;
    put_str(os, vs.str()); put_str(os, cs.str());
    if (has_gen) {
      int rows = C.nvar - K;
      put_i32(os, rows); put_i32(os, K);
      std::vector<unsigned char> P((size_t)rows * K, 0);
      bvec u = zeros_b(K), cw;
      for (int k = 0; k < K; k++) {  // column k of the parity part = encode(e_k)
        u(k) = 1;
        C.G->encode(u, cw);
        for (int r = 0; r < rows; r++) P[(size_t)r * K + k] = (unsigned char)(int)cw(K + r);
        u(k) = 0;
      }
      os.write((const char*)P.data(), P.size());
    }
  }
  static void set_modes(LDPC_Code_LUT& C, const std::string& init_mode) {
    C.set_initial_message_mode(init_mode == "qcha" ? LDPC_Code_LUT::QCHA : LDPC_Code_LUT::CONT);
  }
  static int nvar(const LDPC_Code_LUT& C) { return C.nvar; }
  static int num_edges(const LDPC_Code_LUT& C) { return C.num_edges; }
  static int max_iters(const LDPC_Code_LUT& C) { return C.max_iters; }
  // Front end of LDPC_Code_LUT::decode (src/LDPC_Code_LUT.cpp:207-221), reproduced with
  // the reference's own quant_nonlin so the fixture holds the decoder's exact inputs.
  static void quantise(const LDPC_Code_LUT& C, const vec& llr, ivec& cha, ivec& msg) {
    cha = quant_nonlin(llr, C.qb_Cha);
    if (C.initial_message_mode == LDPC_Code_LUT::CONT) msg = quant_nonlin(llr, C.qb_Msg);
    else msg = C.Nq_Cha_2_Nq_Msg_map(cha);
  }
};
}  // namespace lut_ldpc
typedef lut_ldpc::LDPC_BER_Sim_LUT Peek;

struct Args {
  std::map<std::string, std::string> kv;
  std::vector<std::string> pos;
  std::string get(const std::string& k, const std::string& d = "") const {
    auto it = kv.find(k); return it == kv.end() ? d : it->second;
  }
  bool has(const std::string& k) const { return kv.count(k) > 0; }
  double num(const std::string& k, double d) const { return has(k) ? atof(get(k).c_str()) : d; }
};
static Args parse_args(int argc, char** argv) {
  Args a;
  for (int i = 2; i < argc; i++) {
    std::string s = argv[i];
    if (s.rfind("--", 0) == 0) {
      std::string k = s.substr(2);
      if (i + 1 < argc && std::string(argv[i + 1]).rfind("--", 0) != 0) a.kv[k] = argv[++i];
      else a.kv[k] = "1";
    } else a.pos.push_back(s);
  }
  return a;
}
static std::vector<double> parse_list(const std::string& s) {
  std::vector<double> v; std::stringstream ss(s); std::string t;
  while (std::getline(ss, t, ',')) if (!t.empty()) v.push_back(atof(t.c_str()));
  return v;
}
static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- design: the decoder-building part of LDPC_BER_Sim_LUT::load (src/LDPC_BER_Sim.cpp:434-520)
static LDPC_Code_LUT* build_from_alist(const Args& a, const std::string& alist,
                                       LDPC_Generator_Systematic** Gout, bool report) {
  double t0 = now_s();
  LDPC_Parity* H = new LDPC_Parity(alist, "alist");
  double t1 = now_s();
  LDPC_Generator_Systematic* G = nullptr;
  if (a.has("gen")) G = new LDPC_Generator_Systematic(H);  // permutes the columns of *H (ldpc.cpp:894-999)
  double t2 = now_s();
  LDPC_Code_LUT* C = new LDPC_Code_LUT(H, G);
  double t3 = now_s();
  int max_iter = (int)a.num("max-iter", 30);
  int q_cha = (int)a.num("qcha", 4), q_msg = (int)a.num("qmsg", 3);
  ivec Nq_Msg = ones_i(max_iter) * (1 << q_msg);
  if (a.has("qmsg-list")) {
    std::vector<double> l = parse_list(a.get("qmsg-list"));
    for (int i = 0; i < max_iter && i < (int)l.size(); i++) Nq_Msg(i) = 1 << (int)l[i];
  }
  bvec reuse = zeros_b(max_iter);
  if (a.has("reuse")) {
    std::vector<double> l = parse_list(a.get("reuse"));
    for (int i = 0; i < max_iter && i < (int)l.size(); i++) reuse(i) = (l[i] != 0);
  }
  double sigma2;
  if (a.has("design-thr")) sigma2 = sqr(a.num("design-thr", 0));
  else if (a.has("design-snr")) sigma2 = itpp::pow10(-a.num("design-snr", 0) / 10) / (2 * C->get_rate());
  else { fprintf(stderr, "need --design-thr or --design-snr\n"); exit(2); }
  std::string tree = a.get("tree", "auto_bin_balanced");  // or "filename=<abs path>"
  bool min_lut = a.num("min-lut", 1) != 0;
  C->design_luts(tree, get_empirical_ensemble(*H), min_lut, sigma2, max_iter, reuse, 1 << q_cha, Nq_Msg);
  C->set_exit_conditions(max_iter, a.num("psc", 1) != 0, a.num("pisc", 1) != 0);
  Peek::set_modes(*C, a.get("init-mode", "cont"));
  double t4 = now_s();
  if (report)
    printf("{\"cmd\":\"design\",\"nvar\":%d,\"ninfo\":%d,\"rate\":%.6f,\"sigma2\":%.9g,"
           "\"t_alist\":%.3f,\"t_gen\":%.3f,\"t_ctor_rank\":%.3f,\"t_design\":%.3f}\n",
           C->get_nvar(), C->get_ninfo(), C->get_rate(), sigma2, t1 - t0, t2 - t1, t3 - t2, t4 - t3);
  if (Gout) *Gout = G;
  return C;
}

static int cmd_design(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "design <alist> <out-prefix> [options]\n"); return 2; }
  LDPC_Code_LUT* C = build_from_alist(a, a.pos[0], nullptr, true);
  // NOTE: with a generator, G->save() overwrites "Fileversion" (ldpc.cpp) and the reference's own
  // load_code then rejects the file; codec files are therefore written without generator use in mind
  // and generator runs go through the alist path of fixture/sim.
  C->save_code(a.pos[1] + ".it");
  Peek::write_bundle(*C, a.pos[1] + ".bundle", a.has("gen"));
  return 0;
}

static LDPC_Code_LUT* load_codec(const Args& a, const std::string& fn, LDPC_Generator_Systematic** Gout) {
  if (fn.size() > 6 && fn.substr(fn.size() - 6) == ".alist") return build_from_alist(a, fn, Gout, false);
  LDPC_Generator_Systematic* G = new LDPC_Generator_Systematic();
  LDPC_Code_LUT* C = new LDPC_Code_LUT(fn, G);  // load_code, src/LDPC_Code_LUT.cpp:568-641
  if (!G->is_initialized()) { G = nullptr; }
  if (Gout) *Gout = G;
  int mi = a.has("max-iter") ? (int)a.num("max-iter", 0) : Peek::max_iters(*C);
  C->set_exit_conditions(mi, a.num("psc", 1) != 0, a.num("pisc", 1) != 0);
  Peek::set_modes(*C, a.get("init-mode", "cont"));
  return C;
}

static int cmd_bundle(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "bundle <codec.it> <out.bundle>\n"); return 2; }
  LDPC_Generator_Systematic* G = nullptr;
  LDPC_Code_LUT* C = load_codec(a, a.pos[0], &G);
  Peek::write_bundle(*C, a.pos[1], G != nullptr && !a.has("no-gen"));
  return 0;
}

// Parse the hex message dumps the reference prints at output_verbosity >= 2
// (src/LDPC_Code_LUT.cpp:291-298,311-317,331-337): the reference's own golden-vector hook.
struct TraceRec { int kind; int iter; std::vector<unsigned char> msgs; };
static std::vector<TraceRec> parse_trace(const std::string& txt, int E) {
  std::vector<TraceRec> out; std::istringstream is(txt); std::string line;
  while (std::getline(is, line)) {
    TraceRec r; r.iter = -1;
    if (line.rfind("Initial VN-to-CN", 0) == 0) r.kind = 0;
    else if (line.rfind("CN-to-VN", 0) == 0) r.kind = 1;
    else if (line.rfind("VN-to-CN", 0) == 0) r.kind = 2;
    else continue;
    if (r.kind) { size_t p = line.find("iteration "); r.iter = (int)strtol(line.c_str() + p + 10, nullptr, 16); }  // std::hex is sticky on cout
    std::string data; std::getline(is, data);
    std::istringstream ds(data); r.msgs.resize(E);
    for (int e = 0; e < E; e++) { unsigned v; ds >> std::hex >> v; r.msgs[e] = (unsigned char)v; }
    out.push_back(r);
  }
  return out;
}

// ---- fixture: frames through the reference decoder.
// File: char[8] "LUTFIX01"; i32 nvar E n_frames n_trace with_llr zero_cw; f64 snr_db; then per frame
//   [f64 llr[nvar] if with_llr] u8 cw[nvar] u8 cha[nvar] u8 msg[nvar] u8 out[nvar] i32 ret ;
// then for each traced frame (the first n_trace): i32 n_rec ; n_rec x { i32 kind, i32 iter, u8 msgs[E] }.
static int cmd_fixture(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "fixture <codec.it> <out.fix> [options]\n"); return 2; }
  LDPC_Generator_Systematic* G = nullptr;
  LDPC_Code_LUT* C = load_codec(a, a.pos[0], &G);
  int N = C->get_nvar(), K = C->get_ninfo(), E = Peek::num_edges(*C);
  int n_frames = (int)a.num("frames", 8), n_trace = (int)a.num("trace", 0);
  bool with_llr = a.num("with-llr", 1) != 0, zero_cw = (G == nullptr) || a.has("zero-cw");
  double snr = a.num("snr", 1.0);
  RNG_reset((unsigned)a.num("seed", 0));
  double N0 = itpp::pow10(-snr / 10.0) / C->get_rate();
  AWGN_Channel chan(N0 / 2); BPSK Mod;
  // --llr-from <file>: channel LLRs (f64 [n][nvar], all-zero codeword) chosen elsewhere -- e.g. frames the GPU search
  // tools/find_frames.py found to end in a rare return value -- instead of IT++'s own noise.  Quantisation, decoding
  // and every recorded output below are still the reference's.
  std::vector<double> llr_file;
  if (a.has("llr-from")) {
    std::ifstream is(a.get("llr-from").c_str(), std::ios::binary);
    is.seekg(0, std::ios::end);
    size_t bytes = (size_t)is.tellg();
    is.seekg(0);
    llr_file.resize(bytes / 8);
    is.read((char*)llr_file.data(), (std::streamsize)(llr_file.size() * 8));
    int avail = (int)(llr_file.size() / N), first = (int)a.num("llr-first", 0);
    if (!is || first + (a.has("frames") ? n_frames : 0) > avail) { fprintf(stderr, "--llr-from: short file\n"); return 2; }
    if (!a.has("frames")) n_frames = avail - first;
    zero_cw = true;
  }
  std::ofstream os(a.pos[1].c_str(), std::ios::binary);
  os.write("LUTFIX01", 8);
  put_i32(os, N); put_i32(os, E); put_i32(os, n_frames); put_i32(os, n_trace);
  put_i32(os, with_llr); put_i32(os, zero_cw); put_f64(os, snr);
  std::vector<std::vector<TraceRec> > traces;
  for (int ff = 0; ff < n_frames; ff++) {
    bvec bitsin = zero_cw ? zeros_b(K) : randb(K);
    bvec cw = zero_cw ? zeros_b(N) : C->encode(bitsin);
    vec llr;
    if (!llr_file.empty()) {
      llr.set_size(N);
      const double* src = llr_file.data() + (size_t)((int)a.num("llr-first", 0) + ff) * N;
      for (int i = 0; i < N; i++) llr(i) = src[i];
    } else {
      vec x = chan(Mod.modulate_bits(cw));
      llr = Mod.demodulate_soft_bits(x, N0);
    }
    ivec cha, msg; Peek::quantise(*C, llr, cha, msg);
    bvec out; int ret;
    if (ff < n_trace) {
      C->set_output_verbosity(3);
      std::stringstream cap; std::streambuf* old = std::cout.rdbuf(cap.rdbuf());
      ret = C->lut_decode(cha, msg, out);
      std::cout.rdbuf(old); C->set_output_verbosity(0);
      traces.push_back(parse_trace(cap.str(), E));
    } else ret = C->lut_decode(cha, msg, out);
    if (with_llr) for (int i = 0; i < N; i++) put_f64(os, llr(i));
    for (int i = 0; i < N; i++) os.put((char)(int)cw(i));
    for (int i = 0; i < N; i++) os.put((char)cha(i));
    for (int i = 0; i < N; i++) os.put((char)msg(i));
    for (int i = 0; i < N; i++) os.put((char)(int)out(i));
    put_i32(os, ret);
  }
  for (auto& t : traces) {
    put_i32(os, (int)t.size());
    for (auto& r : t) { put_i32(os, r.kind); put_i32(os, r.iter); os.write((const char*)r.msgs.data(), E); }
  }
  return 0;
}

// ---- sim: frame loop of LDPC_BER_Sim::sim_snr_point (src/LDPC_BER_Sim.cpp:246-311) around the
// reference decoder, using the same IT++ objects (BPSK, AWGN_Channel, BERC, BLERC).  One JSON line
// per SNR point; "wall_s" brackets the frame loop only (setup is reported by design/load).
static int cmd_sim(const Args& a) {
  if (a.pos.size() < 1) { fprintf(stderr, "sim <codec.it> --snr a,b,.. [options]\n"); return 2; }
  double tl0 = now_s();
  LDPC_Generator_Systematic* G = nullptr;
  LDPC_Code_LUT* C = load_codec(a, a.pos[0], &G);
  double tl1 = now_s();
  int N = C->get_nvar(), K = C->get_ninfo();
  bool zero_cw = (G == nullptr) || a.has("zero-cw");
  int64_t Nframes = (int64_t)a.num("frames", 100);
  int Nfers = (int)a.num("nfers", 1 << 30);
  double ber_min = a.num("ber-min", -1), fer_min = a.num("fer-min", -1);
  double max_seconds = a.num("max-seconds", 1e30);
  std::vector<double> snrs = parse_list(a.get("snr", "1.0"));
  RNG_reset((unsigned)a.num("seed", 0));  // LDPC_BER_Sim::run, src/LDPC_BER_Sim.cpp:129
  for (double snr : snrs) {
    double N0 = itpp::pow10(-snr / 10.0) / C->get_rate();
    AWGN_Channel chan(N0 / 2);
    BERC berc_data, berc_uncoded; BLERC ferc; BPSK Mod;
    ferc.set_blocksize(K);
    int64_t sum_iters = 0, n_fail = 0;
    double t0 = now_s();
    for (int64_t ff = 0; ff < Nframes; ff++) {
      bvec bitsin, bitsout, bitsin_coded, bitsout_slicer;
      if (zero_cw) { bitsin = zeros_b(K); bitsin_coded = zeros_b(N); }
      else { bitsin = randb(K); bitsin_coded = C->encode(bitsin); }
      vec s = Mod.modulate_bits(bitsin_coded);
      vec x = chan(s);
      vec softbits = Mod.demodulate_soft_bits(x, N0);
      bitsout_slicer = Mod.demodulate_bits(x);
      if (a.has("count-iters")) {  // same decode, but through lut_decode to see the return value
        ivec cha, msg; Peek::quantise(*C, softbits, cha, msg);
        bvec out; int ret = C->lut_decode(cha, msg, out);
        bitsout = out.left(K);
        sum_iters += ret < 0 ? -ret : ret; n_fail += ret < 0;
      } else bitsout = C->decode(softbits);
      berc_data.count(bitsin, bitsout);
      berc_uncoded.count(bitsin_coded, bitsout_slicer);
      ferc.count(bitsin, bitsout);
      if (ferc.get_errors() > Nfers) break;
      if (now_s() - t0 > max_seconds) break;
    }
    double wall = now_s() - t0;
    printf("{\"cmd\":\"sim\",\"snr_db\":%.6g,\"frames\":%.0f,\"data_bits\":%.0f,\"data_bit_errors\":%.0f,"
           "\"uncoded_bits\":%.0f,\"uncoded_bit_errors\":%.0f,\"frame_errors\":%.0f,\"sum_iters\":%lld,"
           "\"decode_failures\":%lld,\"wall_s\":%.6f,\"load_s\":%.3f,\"nvar\":%d,\"ninfo\":%d}\n",
           snr, ferc.get_total_blocks(), berc_data.get_total_bits(), berc_data.get_errors(),
           berc_uncoded.get_total_bits(), berc_uncoded.get_errors(), ferc.get_errors(),
           (long long)sum_iters, (long long)n_fail, wall, tl1 - tl0, N, K);
    fflush(stdout);
    if (ber_min >= 0 && (berc_data.get_errorrate() < ber_min || ferc.get_errorrate() < fer_min)) break;
  }
  return 0;
}

// results <file.it>: read a results file the way the reference's tooling does (IT++ it_ifile; variables of
// LDPC_BER_Sim_Results::write_itfile, src/LDPC_BER_Sim.cpp:342-361) and print it as one JSON line.  Used to
// check the results writer of lut_ldpc_b200/host/ber_sim.cpp against IT++'s own reader.
static int cmd_results(const Args& a) {
  if (a.pos.size() < 1) { fprintf(stderr, "results <file.it>\n"); return 2; }
  it_ifile f(a.pos[0]);
  const char* vecs[] = {"sim_SNRdB", "sim_Nframes", "sim_Ndatabits", "sim_frame_errors", "sim_data_bit_errors",
                        "sim_uncoded_bit_errors", "ldpc_nvar", "ldpc_nchk", "ldpc_code_rate"};
  printf("{");
  for (const char* n : vecs) {
    vec v;
    f >> Name(n) >> v;
    printf("\"%s\": [", n);
    for (int i = 0; i < v.size(); ++i) printf("%s%.17g", i ? ", " : "", v(i));
    printf("], ");
  }
  double runtime = 0;
  std::string git;
  f >> Name("runtime") >> runtime;
  f >> Name("gitversion") >> git;
  printf("\"runtime\": %.17g, \"gitversion\": \"%s\"}\n", runtime, git.c_str());
  return 0;
}


// ---- the BP comparison decoder of ber_sim's [BP] section (src/LDPC_BER_Sim.cpp:157-244): IT++'s LDPC_Code with the
// LLR_calc_unit the .ini names.  bpfixture records, per frame, the channel LLRs, to_qllr of them, bp_decode's
// soft output and its return value; bpsim is the frame loop of sim_snr_point around LDPC_Code::decode.
static LDPC_Code* load_bp(const Args& a, const std::string& alist, LDPC_Parity** Hout) {
  LDPC_Parity* H = nullptr;
  LDPC_Code* C;
  if (alist.size() > 3 && alist.substr(alist.size() - 3) == ".it") {  // a stored bp_codec.it (LDPC_Code(filename, G))
    it_ifile f(alist);
    bool g_def = false;
    f >> Name("G_defined") >> g_def;
    f.close();
    C = new LDPC_Code(alist, g_def ? new LDPC_Generator_Systematic() : nullptr);
  } else {
    H = new LDPC_Parity(alist, "alist");
    C = new LDPC_Code(H);
  }
  bool psc = a.num("psc", 1) != 0, pisc = a.num("pisc", a.num("psc", 1)) != 0;  // load(): (max_iter, pci, pci), :200
  C->set_exit_conditions((int)a.num("max-iter", 30), psc, pisc);
  C->set_llrcalc(LLR_calc_unit((short)a.num("d1", 12), (short)a.num("d2", 300), (short)a.num("d3", 7),
                               (short)a.num("d4", 8 * (int)sizeof(QLLR) - 4)));
  if (Hout) *Hout = H;
  return C;
}

static int cmd_bpfixture(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "bpfixture <code.alist> <out.fix> [options]\n"); return 2; }
  LDPC_Code* C = load_bp(a, a.pos[0], nullptr);
  int N = C->get_nvar();
  int n_frames = (int)a.num("frames", 8);
  double snr = a.num("snr", 1.0), scale = a.num("llr-scale", 1.0);
  RNG_reset((unsigned)a.num("seed", 0));
  double N0 = itpp::pow10(-snr / 10.0) / C->get_rate();
  AWGN_Channel chan(N0 / 2); BPSK Mod;
  LLR_calc_unit lc = C->get_llrcalc();
  std::ofstream os(a.pos[1].c_str(), std::ios::binary);
  os.write("BPFIX001", 8);
  put_i32(os, N); put_i32(os, n_frames); put_i32(os, C->get_nrof_iterations());
  put_i32(os, (int)(a.num("psc", 1) != 0)); put_i32(os, (int)(a.num("pisc", a.num("psc", 1)) != 0));
  ivec d = lc.get_Dint();
  for (int i = 0; i < 4; i++) put_i32(os, d(i));
  put_i32(os, C->get_ninfo());
  const bool with_llr = a.num("with-llr", 1) != 0;  // large codes: keep only to_qllr of the LLRs
  put_i32(os, (int)with_llr);
  put_f64(os, snr); put_f64(os, scale);
  for (int ff = 0; ff < n_frames; ff++) {
    vec x = chan(Mod.modulate_bits(zeros_b(N)));
    vec llr = Mod.demodulate_soft_bits(x, N0) * scale;  // scale > 1 drives the saturation branches
    QLLRvec qin = lc.to_qllr(llr), qout;
    int ret = C->bp_decode(qin, qout);
    bvec syst = C->decode(llr);  // the call ber_sim makes
    if (with_llr) for (int i = 0; i < N; i++) put_f64(os, llr(i));
    for (int i = 0; i < N; i++) put_i32(os, qin(i));
    for (int i = 0; i < N; i++) put_i32(os, qout(i));
    put_i32(os, ret);
    for (int i = 0; i < syst.length(); i++) os.put((char)(int)syst(i));
  }
  return 0;
}

static int cmd_bpsim(const Args& a) {
  if (a.pos.size() < 1) { fprintf(stderr, "bpsim <code.alist> --snr a,b,.. [options]\n"); return 2; }
  double tl0 = now_s();
  LDPC_Code* C = load_bp(a, a.pos[0], nullptr);
  double tl1 = now_s();
  int N = C->get_nvar(), K = C->get_ninfo();
  int64_t Nframes = (int64_t)a.num("frames", 100);
  int Nfers = (int)a.num("nfers", 1 << 30);
  double max_seconds = a.num("max-seconds", 1e30);
  std::vector<double> snrs = parse_list(a.get("snr", "1.0"));
  RNG_reset((unsigned)a.num("seed", 0));
  for (double snr : snrs) {
    double N0 = itpp::pow10(-snr / 10.0) / C->get_rate();
    AWGN_Channel chan(N0 / 2);
    BERC berc_data, berc_uncoded; BLERC ferc; BPSK Mod;
    ferc.set_blocksize(K);
    bvec bitsin = zeros_b(K), bitsin_coded = zeros_b(N);
    double t0 = now_s();
    for (int64_t ff = 0; ff < Nframes; ff++) {
      vec x = chan(Mod.modulate_bits(bitsin_coded));
      vec softbits = Mod.demodulate_soft_bits(x, N0);
      bvec bitsout_slicer = Mod.demodulate_bits(x);
      bvec bitsout = C->decode(softbits);
      berc_data.count(bitsin, bitsout);
      berc_uncoded.count(bitsin_coded, bitsout_slicer);
      ferc.count(bitsin, bitsout);
      if (ferc.get_errors() > Nfers) break;
      if (now_s() - t0 > max_seconds) break;
    }
    double wall = now_s() - t0;
    printf("{\"cmd\":\"bpsim\",\"snr_db\":%.6g,\"frames\":%.0f,\"data_bits\":%.0f,\"data_bit_errors\":%.0f,"
           "\"uncoded_bits\":%.0f,\"uncoded_bit_errors\":%.0f,\"frame_errors\":%.0f,"
           "\"wall_s\":%.6f,\"load_s\":%.3f,\"nvar\":%d,\"ninfo\":%d}\n",
           snr, ferc.get_total_blocks(), berc_data.get_total_bits(), berc_data.get_errors(),
           berc_uncoded.get_total_bits(), berc_uncoded.get_errors(), ferc.get_errors(), wall, tl1 - tl0, N, K);
    fflush(stdout);
  }
  return 0;
}

// ---- bp_codec.it: what LDPC_BER_Sim::load saves / loads for the BP decoder (src/LDPC_BER_Sim.cpp:207-216,219-239):
// IT++'s LDPC_Code::save_code / load_code (itpp/itpp/comm/ldpc.cpp:1250-1322) + LDPC_Generator_Systematic::save.
//   bpcodec <alist> <out.it> [--gen 1] [--pairs file --n K]   H [and G] -> LDPC_Code -> save_code; with --pairs also
//                                                             n (info, codeword) pairs of G->encode, one byte per bit
//   bpreload <in.it> <out.it> [--pairs file --n K]            load_code (with its generator if present) -> save_code
static void write_pairs(const Args& a, LDPC_Code& C) {
  if (!a.has("pairs")) return;
  std::ofstream os(a.get("pairs").c_str(), std::ios::binary);
  int n = (int)a.num("n", 4), K = C.get_ninfo(), N = C.get_nvar();
  put_i32(os, n); put_i32(os, K); put_i32(os, N);
  RNG_reset((unsigned)a.num("seed", 1));
  for (int f = 0; f < n; f++) {
    bvec u = randb(K), cw = C.encode(u);
    for (int i = 0; i < K; i++) os.put((char)(int)u(i));
    for (int i = 0; i < N; i++) os.put((char)(int)cw(i));
  }
}
static int cmd_bpcodec(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "bpcodec <code.alist> <out.it> [--gen 1]\n"); return 2; }
  RNG_reset((unsigned)a.num("seed", 0));
  LDPC_Parity* H = new LDPC_Parity(a.pos[0], "alist");
  LDPC_Code* C;
  if (a.num("gen", 0) != 0) {
    LDPC_Generator_Systematic* G = new LDPC_Generator_Systematic(H);  // permutes the columns of *H
    C = new LDPC_Code(H, G);
  } else C = new LDPC_Code(H);
  C->save_code(a.pos[1]);
  if (a.num("gen", 0) != 0) write_pairs(a, *C);
  printf("{\"cmd\":\"bpcodec\",\"nvar\":%d,\"ncheck\":%d,\"ninfo\":%d}\n", C->get_nvar(), C->get_ncheck(), C->get_ninfo());
  return 0;
}
static int cmd_bpreload(const Args& a) {
  if (a.pos.size() < 2) { fprintf(stderr, "bpreload <in.it> <out.it>\n"); return 2; }
  it_ifile f(a.pos[0]);
  bool g_def = false;
  f >> Name("G_defined") >> g_def;
  f.close();
  LDPC_Generator_Systematic* G = g_def ? new LDPC_Generator_Systematic() : nullptr;
  LDPC_Code C(a.pos[0], G);
  C.save_code(a.pos[1]);
  if (g_def) write_pairs(a, C);
  printf("{\"cmd\":\"bpreload\",\"nvar\":%d,\"ncheck\":%d,\"ninfo\":%d,\"generator\":%d}\n", C.get_nvar(), C.get_ncheck(),
         C.get_ninfo(), (int)g_def);
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: lutldpc_ref design|bundle|fixture|sim|results|bpfixture|bpsim|bpcodec|bpreload ...\n");
    return 2;
  }
  std::string cmd = argv[1];
  Args a = parse_args(argc, argv);
  if (cmd == "design") return cmd_design(a);
  if (cmd == "bundle") return cmd_bundle(a);
  if (cmd == "fixture") return cmd_fixture(a);
  if (cmd == "sim") return cmd_sim(a);
  if (cmd == "results") return cmd_results(a);
  if (cmd == "bpfixture") return cmd_bpfixture(a);
  if (cmd == "bpsim") return cmd_bpsim(a);
  if (cmd == "bpcodec") return cmd_bpcodec(a);
  if (cmd == "bpreload") return cmd_bpreload(a);
  fprintf(stderr, "unknown command %s\n", cmd.c_str());
  return 2;
}
