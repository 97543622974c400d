// ber_sim for the B200 decoder: same command line and params .ini as the reference's prog/ber_sim.cpp
// (:46-160) + LDPC_BER_Sim / LDPC_BER_Sim_LUT (src/LDPC_BER_Sim.cpp:42-155,246-311,376-550), with the
// frame loop of sim_snr_point running on the GPU(s) through the C ABI.
//
//   ber_sim -p params.ini [-b basedir] [-s seed] [-c custom-name] [--gpus G] [--batch F]
//
// Differences, all at the edges of the hot path:
//  * the decoder is either loaded -- Sim.codec_filename names the lut_codec.it a reference run saved
//    (LDPC_Code_LUT::save_code; looked up under Sim.codes_dir like the reference does, src/LDPC_BER_Sim.cpp:63,84-86,
//    524) or a decoder bundle -- or, as in the reference's load() (:434-520), DESIGNED from LDPC.parity_filename
//    (<codes_dir>/<name>.alist) and the [LUT] keys: alist import + graph flattening (lutldpc_alist_graph), LUT design
//    by density evolution on the host (lutldpc_design_luts, table-exact with the reference), and for
//    LDPC.zero_codeword = false the systematic generator by GF(2) elimination on the device
//    (lutldpc_create_systematic).  The shipped params/ber.ini.*.example files run as they are;
//  * results are written both as the reference's IT++ results file (<name>_rseedNNNN.it, the variables of
//    LDPC_BER_Sim_Results::write_itfile, src/LDPC_BER_Sim.cpp:342-361, readable by scripts/analyze_results.m)
//    and as JSON with the same variable names plus the counters only the GPU path keeps;
//  * with --gpus G the frames of every SNR point are split over G devices (one host thread per GPU,
//    disjoint Philox frame ranges) and the counters are summed; stop rules act on the sums.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <sys/stat.h>

#include "ldpc_code_bp.hpp"
#include "ldpc_code_lut.hpp"

// Minimal IT++ it_file (format version 3) writer: "IT++" 0x03, then per variable u64 hdr_bytes, data_bytes,
// block_bytes, name\0 type\0 desc\0, data -- little endian (itpp/itpp/base/itfile.h:98-125).  Only the three
// types LDPC_BER_Sim_Results::write_itfile uses: vec ("dvec": u64 count + f64), double ("float64"), string.
struct ItWriter {
  std::ofstream f;
  explicit ItWriter(const std::string& fn) : f(fn.c_str(), std::ios::binary) {
    const char h[5] = {'I', 'T', '+', '+', 3};
    f.write(h, 5);
  }
  void block(const std::string& name, const std::string& type, const std::string& data) {
    const uint64_t hdr = 24 + name.size() + 1 + type.size() + 1 + 1, db = data.size(), blk = hdr + db;
    f.write(reinterpret_cast<const char*>(&hdr), 8);
    f.write(reinterpret_cast<const char*>(&db), 8);
    f.write(reinterpret_cast<const char*>(&blk), 8);
    f.write(name.c_str(), name.size() + 1);
    f.write(type.c_str(), type.size() + 1);
    f.put('\0');  // empty description
    f.write(data.data(), data.size());
  }
  void vec(const std::string& name, const std::vector<double>& v) {
    std::string d(8 + 8 * v.size(), '\0');
    const uint64_t n = v.size();
    memcpy(&d[0], &n, 8);
    if (n) memcpy(&d[8], v.data(), 8 * v.size());
    block(name, "dvec", d);
  }
  void f64(const std::string& name, double x) { block(name, "float64", std::string(reinterpret_cast<const char*>(&x), 8)); }
  void str(const std::string& name, const std::string& s) {
    std::string d(8, '\0');
    const uint64_t n = s.size();
    memcpy(&d[0], &n, 8);
    block(name, "string", d + s);
  }
};

using namespace lut_ldpc_b200;

// ---------------------------------------------------------------- params .ini (boost::property_tree::ini_parser subset)
struct Ini {
  std::map<std::string, std::string> kv;  // "Section.key" -> value
  static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? "" : s.substr(a, b - a + 1);
  }
  bool read(const std::string& fn) {
    std::ifstream f(fn.c_str());
    if (!f) return false;
    std::string line, sec;
    while (std::getline(f, line)) {
      line = trim(line);
      if (line.empty() || line[0] == ';' || line[0] == '#') continue;
      if (line[0] == '[') { sec = trim(line.substr(1, line.find(']') - 1)); continue; }
      size_t eq = line.find('=');
      if (eq == std::string::npos) continue;
      kv[sec + "." + trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
    }
    return true;
  }
  bool has(const std::string& k) const { return kv.count(k) > 0; }
  bool has_section(const std::string& s) const {
    for (auto& e : kv) if (e.first.compare(0, s.size() + 1, s + ".") == 0) return true;
    return false;
  }
  std::string get(const std::string& k, const std::string& d) const { auto it = kv.find(k); return it == kv.end() ? d : it->second; }
  double num(const std::string& k, double d) const { return has(k) ? atof(get(k, "").c_str()) : d; }
  bool flag(const std::string& k, bool d) const {
    if (!has(k)) return d;
    std::string v = get(k, "");
    return v == "1" || v == "true" || v == "True";
  }
};

// IT++ vec string syntax as used for Sim.SNRdB: "a:step:b", "a:b", "[a b c]", "a,b,c"
static std::vector<double> parse_vec(std::string s) {
  for (char& c : s) if (c == '[' || c == ']' || c == ',' || c == ';') c = ' ';
  std::vector<double> out;
  std::istringstream is(s);
  std::string tok;
  while (is >> tok) {
    size_t c1 = tok.find(':');
    if (c1 == std::string::npos) { out.push_back(atof(tok.c_str())); continue; }
    size_t c2 = tok.find(':', c1 + 1);
    double a = atof(tok.substr(0, c1).c_str()), step = 1, b;
    if (c2 == std::string::npos) b = atof(tok.substr(c1 + 1).c_str());
    else { step = atof(tok.substr(c1 + 1, c2 - c1 - 1).c_str()); b = atof(tok.substr(c2 + 1).c_str()); }
    if (step == 0) { out.push_back(a); continue; }
    int n = (int)std::floor((b - a) / step + 1e-9);
    for (int i = 0; i <= n; ++i) out.push_back(a + i * step);
  }
  return out;
}

struct Point {
  double snr;
  lutldpc_sim_counters c;
};

static void usage() {
  std::cout << "Allowed options:\n"
               "  -h [ --help ]             produce help message\n"
               "  -p [ --params ] arg       parameter file (.ini, same keys as the reference's ber_sim)\n"
               "  -b [ --base-dir ] arg     base directory for relative paths (default: cwd)\n"
               "  -s [ --seed ] arg (=0)    random seed (frames are Philox streams keyed by seed + rand_seed_offset)\n"
               "  -c [ --custom-name ] arg  appended to the results name\n"
               "  --gpus arg (=1)           number of GPUs to split every SNR point over\n"
               "  --batch arg (=0)          frames in flight per GPU (0 = library default)\n";
}

int main(int argc, char** argv) {
  std::string params, basedir = ".", custom;
  uint64_t seed = 0;
  int gpus = 1, batch = 0;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "-h" || a == "--help") { usage(); return 0; }
    else if (a == "-p" || a == "--params") params = next();
    else if (a == "-b" || a == "--base-dir") basedir = next();
    else if (a == "-s" || a == "--seed") seed = strtoull(next().c_str(), nullptr, 10);
    else if (a == "-c" || a == "--custom-name") custom = next();
    else if (a == "--gpus") gpus = atoi(next().c_str());
    else if (a == "--batch") batch = atoi(next().c_str());
    else { std::cerr << "unknown option " << a << "\n"; usage(); return 1; }
  }
  if (params.empty()) { std::cerr << "Parameter file not specified\n"; usage(); return 1; }
  auto path = [&](const std::string& p) { return (!p.empty() && p[0] == '/') ? p : basedir + "/" + p; };
  Ini ini;
  if (!ini.read(path(params))) { std::cerr << "cannot read " << path(params) << "\n"; return 1; }
  // prog/ber_sim.cpp:134-147: a [LUT] section (or Sim.codec_type = LUT) selects the LUT decoder, else a [BP] section
  // (or codec_type = BP) the BP comparison decoder
  const bool use_lut = ini.has_section("LUT") || ini.get("Sim.codec_type", "none") == "LUT";
  const bool use_bp = !use_lut && (ini.has_section("BP") || ini.get("Sim.codec_type", "none") == "BP");
  if (!use_lut && !use_bp) {
    std::cerr << "You must specify the type of decoder in the params file by either\n * explicitly setting at least one "
                 "parameter in the [LUT] or [BP] section of the parameter file\n * loading a codec via its filename and "
                 "specifying its type\n";
    return 1;
  }
  try {
    // ---- [Sim] / [LDPC] / [LUT], defaults as src/LDPC_BER_Sim.cpp:53-78,386-413
    if (!ini.has("Sim.SNRdB")) throw std::runtime_error("LDPC_BER_Sim: Sim.SNRdB is required");
    std::vector<double> snrs = parse_vec(ini.get("Sim.SNRdB", ""));
    const int64_t Nframes = (int64_t)ini.num("Sim.Nframes", 1e5);
    const int64_t Nfers = (int64_t)ini.num("Sim.Nfers", 20);
    const double ber_min = ini.num("Sim.ber_min", 1e-7), fer_min = ini.num("Sim.fer_min", 1e-5);
    const uint64_t seed_off = (uint64_t)ini.num("Sim.rand_seed_offset", 0);
    const std::string results_prefix = ini.get("Sim.results_prefix", "RES"), results_dir = ini.get("Sim.results_dir", "results");
    custom = ini.get("Sim.custom_name", "") + custom;
    const bool zero_cw = ini.flag("LDPC.zero_codeword", true);
    const bool pci = ini.flag("LDPC.parity_check_iter", true);
    const std::string codec = ini.get("Sim.codec_filename", "");
    const std::string codes_dir = ini.get("Sim.codes_dir", "codes");
    const std::string imode = ini.get("LUT.initial_message_mode", "from_continuous_input");
    const int verbosity = (int)ini.num("LUT.output_verbosity", 0);
    const int ndev = lutldpc_device_count();
    if (ndev <= 0) throw std::runtime_error("no CUDA device (there is no CPU fallback)");
    if (gpus < 1 || gpus > ndev) throw std::runtime_error("--gpus exceeds the number of visible devices");
    std::vector<std::unique_ptr<LDPC_Code_LUT>> C(gpus);
    std::vector<std::unique_ptr<LDPC_Code>> Cbp(gpus);
    auto configure = [&](LDPC_Code_LUT& c) {
      const int mi = ini.has("LUT.max_iter") ? (int)ini.num("LUT.max_iter", 30) : c.get_nrof_iterations();
      c.set_exit_conditions(mi, pci, pci);  // psc = pisc = parity_check_iter, src/LDPC_BER_Sim.cpp:500
      c.set_initial_message_mode(imode == "from_quantized_channel_llrs" ? LDPC_Code_LUT::QCHA : LDPC_Code_LUT::CONT);
    };
    if (use_bp) {
      // ---- LDPC_BER_Sim::load, src/LDPC_BER_Sim.cpp:157-244: LDPC_Parity(alist) [+ LDPC_Generator_Systematic] ->
      // LDPC_Code, exit conditions (max_iter, pci, pci) and the LLR_calc_unit of the [BP] section
      LLR_calc_unit lc((short)ini.num("BP.qllr_scale_res", 12), (short)ini.num("BP.qllr_table_size", 300),
                       (short)ini.num("BP.qllr_spacing_res", 7), (short)ini.num("BP.qllr_total_res", 28));
      const int bp_iter = (int)ini.num("BP.max_iter", 30);
      struct stat sb;
      if (!codec.empty()) {  // LDPC_Code(codec_path, G), src/LDPC_BER_Sim.cpp:219-239
        std::string codec_file = path(codes_dir + "/" + codec);
        if (codec[0] == '/' || stat(codec_file.c_str(), &sb) != 0) codec_file = path(codec);
        if (stat(codec_file.c_str(), &sb) != 0) throw std::runtime_error("Codec file" + codec_file + " does not exist!");
        for (int g = 0; g < gpus; ++g) {
          if (lutldpc_set_device(g) != LUTLDPC_OK) throw std::runtime_error("cudaSetDevice failed");
          Cbp[g].reset(new LDPC_Code(codec_file, LDPC_Code::from_file_t(), bp_iter, pci, pci, lc));
        }
        if (!zero_cw && !Cbp[0]->has_generator())
          throw std::runtime_error("LDPC.zero_codeword=false needs a codec file that carries the generator");
      } else {
        const std::string alist = path(codes_dir + "/" + ini.get("LDPC.parity_filename", "") + ".alist");
        if (stat(alist.c_str(), &sb) != 0) throw std::runtime_error("Parity file" + alist + " does not exist!");  // :162-164
        for (int g = 0; g < gpus; ++g) {
          if (lutldpc_set_device(g) != LUTLDPC_OK) throw std::runtime_error("cudaSetDevice failed");
          Cbp[g].reset(new LDPC_Code(alist, !zero_cw, bp_iter, pci, pci, lc));
        }
      }
    } else if (!codec.empty()) {
      // codes_path / codec_filename (src/LDPC_BER_Sim.cpp:524), else relative to the base directory
      std::string codec_file = path(codes_dir + "/" + codec);
      struct stat sb;
      if (codec[0] == '/' || stat(codec_file.c_str(), &sb) != 0) codec_file = path(codec);
      for (int g = 0; g < gpus; ++g) {
        if (lutldpc_set_device(g) != LUTLDPC_OK) throw std::runtime_error("cudaSetDevice failed");
        C[g].reset(new LDPC_Code_LUT(codec_file));
      }
      if (!zero_cw && !C[0]->has_generator())
        throw std::runtime_error("LDPC.zero_codeword=false needs a decoder bundle that carries the generator");
    } else {
      // ---- LDPC_BER_Sim_LUT::load, src/LDPC_BER_Sim.cpp:434-507: alist -> graph -> design_luts
      if (!ini.has("LDPC.parity_filename"))
        throw std::runtime_error("LDPC_BER_Sim_LUT::load(): LDPC.parity_filename (or Sim.codec_filename) is required");
      const std::string alist = path(codes_dir + "/" + ini.get("LDPC.parity_filename", "") + ".alist");
      if (!ini.has("LUT.design_thr") && !ini.has("LUT.design_SNRdB"))
        throw std::runtime_error("LDPC_BER_Sim_LUT::load(): Either design_thr or design_SNRdB must be specified");  // :478-484
      const int max_iter = (int)ini.num("LUT.max_iter", 30);
      const int q_cha = (int)ini.num("LUT.qbits_channel", 4), q_msg = (int)ini.num("LUT.qbits_message_uniform", 3);
      std::vector<int32_t> nq(max_iter, 1 << q_msg), reuse(max_iter, 0);
      if (ini.has("LUT.qbits_messages")) {  // :398-400
        std::vector<double> v = parse_vec(ini.get("LUT.qbits_messages", ""));
        if ((int)v.size() != max_iter) throw std::runtime_error("LUT.qbits_messages must have max_iter entries");
        for (int i = 0; i < max_iter; ++i) nq[i] = 1 << (int)v[i];
      }
      if (ini.has("LUT.reuse_lut")) {
        std::vector<double> v = parse_vec(ini.get("LUT.reuse_lut", ""));
        if ((int)v.size() != max_iter) throw std::runtime_error("LUT.reuse_lut must have max_iter entries");
        for (int i = 0; i < max_iter; ++i) reuse[i] = v[i] != 0;
      }
      // auto_bin_high and root_only behave as auto_bin_balanced in the reference's ber_sim (:487-488)
      std::string method = "auto_bin_balanced";
      if (ini.get("LUT.tree_mode", "auto_bin_balanced") == "file")
        method = "filename=" + path(ini.get("LUT.trees_dir", "trees") + "/" + ini.get("LUT.trees_filename", ""));
      lutldpc_graph gr;
      // rank like decoder_parameterization (src/LDPC_Code_LUT.cpp:493-499); on the device for the large codes
      if (lutldpc_alist_graph(alist.c_str(), 0, &gr) != LUTLDPC_OK) throw std::runtime_error(lutldpc_last_error(nullptr));
      const int nvar0 = gr.nvar;
      lutldpc_graph_free(&gr);
      if (lutldpc_alist_graph(alist.c_str(), nvar0 >= 100000 ? 0 : (nvar0 > 20000 ? 2 : 1), &gr) != LUTLDPC_OK)
        throw std::runtime_error(lutldpc_last_error(nullptr));
      if (gr.nchk_lin_indep < 0) gr.nchk_lin_indep = gr.nchk;
      const double rate = 1.0 - (double)gr.nchk_lin_indep / gr.nvar;
      const double thr = ini.num("LUT.design_thr", 0);
      const double sigma2 = ini.has("LUT.design_thr") ? thr * thr : std::pow(10.0, -ini.num("LUT.design_SNRdB", 0) / 10) / (2 * rate);
      lutldpc_design_params dp;
      memset(&dp, 0, sizeof dp);
      dp.dv_vec = gr.dv_vec; dp.nvar = gr.nvar; dp.dc_vec = gr.dc_vec; dp.nchk = gr.nchk;
      dp.max_iters = max_iter; dp.Nq_Cha = 1 << q_cha; dp.min_lut = ini.flag("LUT.min_lut", true);
      dp.Nq_Msg = nq.data(); dp.reuse_vec = reuse.data(); dp.sigma2 = sigma2; dp.tree_method = method.c_str();
      lutldpc_design_result dr;
      if (lutldpc_design_luts(&dp, &dr) != LUTLDPC_OK) {
        std::string e = lutldpc_last_error(nullptr);
        lutldpc_graph_free(&gr);
        throw std::runtime_error("LDPC_Code_LUT::design_luts(): " + e);
      }
      lutldpc_desc d;
      memset(&d, 0, sizeof d);
      d.nvar = gr.nvar; d.nchk = gr.nchk; d.nchk_lin_indep = gr.nchk_lin_indep; d.num_edges = gr.num_edges;
      d.max_iters = max_iter; d.Nq_Cha = dp.Nq_Cha; d.min_lut = dp.min_lut;
      d.dv_vec = gr.dv_vec; d.dc_vec = gr.dc_vec; d.cn_msg_idx = gr.cn_msg_idx;
      d.Nq_Msg = nq.data(); d.reuse_vec = reuse.data(); d.Nq_Cha_2_Nq_Msg_map = dr.Nq_Cha_2_Nq_Msg_map;
      d.qb_Cha = dr.qb_Cha; d.qb_Msg = dr.qb_Msg; d.n_qb_Msg = dr.n_qb_Msg;
      d.var_tree_string = dr.var_tree_string; d.var_tree_len = dr.var_tree_len;
      d.chk_tree_string = dr.chk_tree_string; d.chk_tree_len = dr.chk_tree_len;
      try {
        for (int g = 0; g < gpus; ++g) {
          if (lutldpc_set_device(g) != LUTLDPC_OK) throw std::runtime_error("cudaSetDevice failed");
          C[g].reset(new LDPC_Code_LUT(d));
          // LDPC.zero_codeword = false: generator + column-permuted H, as load() builds G before the codec (:447-473)
          if (!zero_cw) C[g].reset(new LDPC_Code_LUT(*C[g], LDPC_Code_LUT::with_generator_t()));
        }
      } catch (...) {
        lutldpc_design_free(&dr);
        lutldpc_graph_free(&gr);
        throw;
      }
      lutldpc_design_free(&dr);
      lutldpc_graph_free(&gr);
      std::cout << "Designed LUT decoder for " << alist << " (sigma2 = " << sigma2 << ", " << max_iter << " iterations)" << std::endl;
    }
    for (int g = 0; g < gpus && use_lut; ++g) {
      if (lutldpc_set_device(g) != LUTLDPC_OK) throw std::runtime_error("cudaSetDevice failed");
      configure(*C[g]);
    }
    const int N = use_bp ? Cbp[0]->get_nvar() : C[0]->get_nvar(), K = use_bp ? Cbp[0]->get_ninfo() : C[0]->get_ninfo();
    const int nchk = use_bp ? Cbp[0]->get_ncheck() : C[0]->get_nchk();
    const double rate = use_bp ? Cbp[0]->get_rate() : C[0]->get_rate();
    std::ostringstream name;  // gen_filename, src/LDPC_BER_Sim.cpp:104-114 (BP), :553-568 (LUT)
    name << results_prefix << "_N" << N << "_R" << rate << "_maxIter" << (use_bp ? Cbp[0]->get_nrof_iterations() : C[0]->get_nrof_iterations())
         << "_zcw" << (int)zero_cw << "_frames" << Nframes;
    if (use_lut) name << (C[0]->is_min_lut() ? "_minLUT" : "_LUT");
    name << custom;
    if (use_bp && codec.empty() && (int64_t)seed == (int64_t)ini.num("Sim.save_codec", 0)) {  // :207-216
      mkdir(path(results_dir).c_str(), 0777);
      mkdir((path(results_dir) + "/" + name.str()).c_str(), 0777);
      Cbp[0]->save_code(path(results_dir) + "/" + name.str() + "/bp_codec.it");
    }
    // a designed decoder is saved when the seed equals Sim.save_codec (src/LDPC_BER_Sim.cpp:59,511-520)
    if (use_lut && codec.empty() && (int64_t)seed == (int64_t)ini.num("Sim.save_codec", 0)) {
      mkdir(path(results_dir).c_str(), 0777);
      mkdir((path(results_dir) + "/" + name.str()).c_str(), 0777);
      C[0]->save_code(path(results_dir) + "/" + name.str() + "/lut_codec.it");
    }

    // ---- run(), src/LDPC_BER_Sim.cpp:121-155
    auto t0 = std::chrono::steady_clock::now();
    std::vector<Point> points;
    uint64_t next_frame = 0;  // global Philox frame index: disjoint across SNR points, chunks and GPUs
    const int64_t chunk_per_gpu = batch > 0 ? batch : 8192;
    bool stop = false;
    for (double snr : snrs) {
      Point pt;
      pt.snr = snr;
      memset(&pt.c, 0, sizeof pt.c);
      if (!stop) {
        // sim_snr_point: frames until Nframes or more than Nfers frame errors (:260-291), at chunk granularity
        while ((int64_t)pt.c.frames < Nframes && (int64_t)pt.c.frame_errors <= Nfers) {
          // chunk size from the running FER estimate: the reference stops at the first frame with more than Nfers
          // frame errors (:289); a small first chunk, then ~1.25 x the frames expected to bring the missing errors,
          // so the overshoot is a fraction of a chunk instead of a whole maximal one
          int64_t want = chunk_per_gpu * gpus;
          if (Nfers < (1LL << 40)) {
            if (pt.c.frames == 0) want = std::min<int64_t>(want, std::max<int64_t>(1024, 64 * (Nfers + 1)));
            else {
              const double fer = std::max<double>((double)pt.c.frame_errors, 0.5) / (double)pt.c.frames;
              want = std::max<int64_t>(std::min<int64_t>(1024, want),
                                       std::min<int64_t>(want, (int64_t)(1.25 * (double)(Nfers + 1 - (int64_t)pt.c.frame_errors) / fer) + 1));
            }
          }
          int64_t todo = std::min<int64_t>(Nframes - (int64_t)pt.c.frames, want);
          std::vector<lutldpc_sim_counters> part(gpus);
          std::vector<std::thread> th;
          std::vector<std::string> errs(gpus);
          int64_t per = (todo + gpus - 1) / gpus, off = 0;
          for (int g = 0; g < gpus; ++g) {
            int64_t n = std::min(per, todo - off);
            memset(&part[g], 0, sizeof part[g]);
            if (n <= 0) continue;
            lutldpc_sim_params sp;
            sp.snr_db = snr; sp.n_frames = n; sp.seed = seed + seed_off; sp.first_frame = next_frame + (uint64_t)off;
            sp.zero_codeword = zero_cw; sp.batch_frames = batch;
            if (verbosity > 0 && use_lut) {
              // LUT.output_verbosity > 0: the "Stimuli Pair" print of LDPC_Code_LUT::decode (src/LDPC_Code_LUT.cpp:228-238),
              // frame by frame in frame order -- the golden vectors of the reference's VHDL flow
              std::vector<uint8_t> lab((size_t)n * N), bits((size_t)n * N);
              C[g]->simulate_snr_stimuli(sp, part[g], lab.data(), bits.data());
              for (int64_t f = 0; f < n; ++f) {
                std::cout << "Stimuli Pair (Quantized channel LLR decoder inputs in hex format and decoder output in binary format): "
                          << std::endl;
                char buf[16];
                for (int v = 0; v < N; ++v) { snprintf(buf, sizeof buf, "%08X  ", lab[(size_t)f * N + v]); std::cout << buf; }
                std::cout << std::endl;
                for (int v = 0; v < N; ++v) std::cout << (int)bits[(size_t)f * N + v] << "  ";
                std::cout << std::endl << std::endl;
              }
              off += n;
              continue;
            }
            th.emplace_back([&, g, sp]() {
              try {
                if (use_bp) Cbp[g]->simulate_snr(sp, part[g]);
                else C[g]->simulate_snr(sp, part[g]);
              } catch (const std::exception& e) { errs[g] = e.what(); }
            });
            off += n;
          }
          for (auto& t : th) t.join();
          for (auto& e : errs) if (!e.empty()) throw std::runtime_error(e);
          next_frame += (uint64_t)todo;
          for (int g = 0; g < gpus; ++g) {
            uint64_t* dst = reinterpret_cast<uint64_t*>(&pt.c);
            const uint64_t* src = reinterpret_cast<const uint64_t*>(&part[g]);
            for (size_t i = 0; i < sizeof(pt.c) / 8; ++i) dst[i] += src[i];
          }
        }
        const double ber = pt.c.data_bits ? (double)pt.c.data_bit_errors / pt.c.data_bits : 0;
        const double ubr = pt.c.uncoded_bits ? (double)pt.c.uncoded_bit_errors / pt.c.uncoded_bits : 0;
        const double fer = pt.c.frames ? (double)pt.c.frame_errors / pt.c.frames : 0;
        // the reference's line, src/LDPC_BER_Sim.cpp:294-300
        std::cout << "SNR = " << snr << "  Simulated " << pt.c.frames << " frames and " << pt.c.data_bits << " data bits. "
                  << "Obtained " << pt.c.data_bit_errors << " data bit errors. "
                  << " Data BER: " << ber << " Uncoded BER: " << ubr << " FER: " << fer << std::endl;
        if (ber < ber_min || fer < fer_min) stop = true;  // :307: the point is recorded, later points stay empty (:142-149)
      }
      points.push_back(pt);
    }
    const double runtime = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // ---- save(), src/LDPC_BER_Sim.cpp:317-361 (JSON with the reference's variable names)
    const std::string dir = path(results_dir) + "/" + name.str();
    mkdir(path(results_dir).c_str(), 0777);
    mkdir(dir.c_str(), 0777);
    char sfx[32];
    snprintf(sfx, sizeof sfx, "_rseed%04llu", (unsigned long long)(seed + seed_off));  // :327
    std::ofstream js((dir + "/" + name.str() + sfx + ".json").c_str());
    auto arr = [&](const char* key, uint64_t lutldpc_sim_counters::*m) {
      js << "  \"" << key << "\": [";
      for (size_t i = 0; i < points.size(); ++i) js << (i ? ", " : "") << points[i].c.*m;
      js << "],\n";
    };
    js << "{\n  \"sim_SNRdB\": [";
    for (size_t i = 0; i < points.size(); ++i) js << (i ? ", " : "") << points[i].snr;
    js << "],\n";
    arr("sim_Nframes", &lutldpc_sim_counters::frames);
    arr("sim_Ndatabits", &lutldpc_sim_counters::data_bits);
    arr("sim_frame_errors", &lutldpc_sim_counters::frame_errors);
    arr("sim_data_bit_errors", &lutldpc_sim_counters::data_bit_errors);
    arr("sim_uncoded_bit_errors", &lutldpc_sim_counters::uncoded_bit_errors);
    arr("sim_Nuncodedbits", &lutldpc_sim_counters::uncoded_bits);
    arr("sim_sum_iterations", &lutldpc_sim_counters::sum_iters);
    arr("sim_decode_failures", &lutldpc_sim_counters::decode_failures);
    js << "  \"ldpc_nvar\": " << N << ",\n  \"ldpc_nchk\": " << nchk << ",\n  \"ldpc_code_rate\": "
       << rate << ",\n  \"ldpc_ninfo\": " << K << ",\n  \"runtime\": " << runtime << ",\n  \"gpus\": " << gpus
       << ",\n  \"gitversion\": \"lut_ldpc_b200\"\n}\n";
    js.close();
    {  // LDPC_BER_Sim_Results::write_itfile, src/LDPC_BER_Sim.cpp:342-361
      ItWriter it(dir + "/" + name.str() + sfx + ".it");
      auto col = [&](uint64_t lutldpc_sim_counters::*m) {
        std::vector<double> v;
        for (auto& p : points) v.push_back((double)(p.c.*m));
        return v;
      };
      std::vector<double> snr_v;
      for (auto& p : points) snr_v.push_back(p.snr);
      it.vec("sim_SNRdB", snr_v);
      it.vec("sim_Nframes", col(&lutldpc_sim_counters::frames));
      it.vec("sim_Ndatabits", col(&lutldpc_sim_counters::data_bits));
      it.vec("sim_frame_errors", col(&lutldpc_sim_counters::frame_errors));
      it.vec("sim_data_bit_errors", col(&lutldpc_sim_counters::data_bit_errors));
      it.vec("sim_uncoded_bit_errors", col(&lutldpc_sim_counters::uncoded_bit_errors));
      it.vec("ldpc_nvar", {(double)N});
      it.vec("ldpc_nchk", {(double)nchk});
      it.vec("ldpc_code_rate", {rate});
      it.f64("runtime", runtime);
      it.str("gitversion", "lut_ldpc_b200");
    }
    // copy of the params file next to the results, like the reference (:331-338)
    std::ifstream src(path(params).c_str(), std::ios::binary);
    std::string pbase = params.substr(params.find_last_of('/') == std::string::npos ? 0 : params.find_last_of('/') + 1);
    std::ofstream dst((dir + "/" + pbase).c_str(), std::ios::binary);
    dst << src.rdbuf();
    std::cout << "Results written to " << dir << " (runtime " << runtime << " s)" << std::endl;
  } catch (const std::exception& e) {
    std::cerr << "ber_sim: " << e.what() << std::endl;
    return 1;
  }
  return 0;
}
