// LUT design on the host (design.cpp): LDPC_Code_LUT::design_luts restated, table-exact.
#pragma once
#include <string>
#include <vector>

namespace lutldpc {
namespace design {

struct Params {
  std::vector<int> dv, dc;        // node degrees of the code (the ensemble is the empirical one of H)
  int max_iters = 0, Nq_Cha = 0;
  std::vector<int> Nq_Msg;        // [max_iters]
  std::vector<int> reuse_vec;     // [max_iters] or empty (= zeros)
  bool min_lut = true;
  double sigma2 = 0;              // design noise variance: design_thr^2, or 10^(-design_SNRdB/10) / (2 rate)
  std::string tree_method = "auto_bin_balanced";  // or "filename=<tree template ini>"
};

struct Result {
  std::vector<double> qb_Cha, qb_Msg;
  std::vector<int> cha2msg;       // Nq_Cha_2_Nq_Msg_map
  std::string var_tree_string, chk_tree_string;  // the reference's text serialisation of Array<Array<LUT_Tree>>
  int dv_act = 0, dc_act = 0;
};

bool design_luts(const Params& p, Result& r, std::string& err);

}  // namespace design
}  // namespace lutldpc
