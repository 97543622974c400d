// TEST INFRASTRUCTURE: drives the C++ facade lut_ldpc_b200/host/ldpc_code_bp.hpp one frame per call, the way ber_sim
// uses IT++'s LDPC_Code, against a BPFIX001 fixture written by the compiled reference (layout: oracle/ref_driver.cpp
// cmd_bpfixture).  usage: facade_check_bp <code.alist> <fixture with LLRs>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iterator>
#include <vector>

#include "lut_ldpc_b200/host/ldpc_code_bp.hpp"

using namespace lut_ldpc_b200;

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::ifstream is(argv[2], std::ios::binary);
  std::vector<char> b((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
  if (b.size() < 68 || memcmp(b.data(), "BPFIX001", 8) != 0) { fprintf(stderr, "bad fixture\n"); return 2; }
  int32_t h[11];
  memcpy(h, b.data() + 8, 44);
  const int N = h[0], nf = h[1], K = h[9], with_llr = h[10];
  if (!with_llr) { fprintf(stderr, "fixture has no LLRs\n"); return 2; }
  long mismatches = 0, checked = 0;
  try {
    LDPC_Code C(argv[1], /*with_generator=*/false, h[2], h[3] != 0, h[4] != 0,
                LLR_calc_unit((short)h[5], (short)h[6], (short)h[7], (short)h[8]));
    mismatches += C.get_nvar() != N || C.get_ninfo() != K; ++checked;
    size_t off = 68;
    for (int f = 0; f < nf; ++f) {
      vec llr(N);
      memcpy(llr.data(), b.data() + off, 8 * (size_t)N); off += 8 * (size_t)N;
      QLLRvec qin(N), qout_ref(N);
      memcpy(qin.data(), b.data() + off, 4 * (size_t)N); off += 4 * (size_t)N;
      memcpy(qout_ref.data(), b.data() + off, 4 * (size_t)N); off += 4 * (size_t)N;
      int32_t ret; memcpy(&ret, b.data() + off, 4); off += 4;
      const uint8_t* syst = (const uint8_t*)b.data() + off; off += K;
      Channel_Code* cc = &C;
      bvec bits = cc->decode(llr);                                   // Channel_Code::decode(const vec&) -> first ninfo bits
      mismatches += (int)bits.size() != K || memcmp(bits.data(), syst, K) != 0; ++checked;
      QLLRvec qout;
      int r = C.bp_decode(qin, qout);                                // LDPC_Code::bp_decode: soft output + return value
      mismatches += r != ret; ++checked;
      mismatches += memcmp(qout.data(), qout_ref.data(), 4 * (size_t)N) != 0; ++checked;
    }
    bool threw = false;
    try { C.encode(bvec(K)); } catch (const std::runtime_error&) { threw = true; }
    mismatches += !threw; ++checked;                                 // no generator: encode is an error, like the reference
    LDPC_Code G(argv[1], /*with_generator=*/true, h[2], true, true);
    bvec u(G.get_ninfo());
    for (size_t i = 0; i < u.size(); ++i) u[i] = (uint8_t)((i * 7 + i / 3) & 1);
    bvec cw = G.encode(u);
    mismatches += memcmp(cw.data(), u.data(), u.size()) != 0; ++checked;  // systematic
    vec clean(cw.size());
    for (size_t i = 0; i < cw.size(); ++i) clean[i] = cw[i] ? -8.0 : 8.0;
    mismatches += G.decode(clean) != u; ++checked;
    printf("{\"frames\":%d,\"checked\":%ld,\"mismatches\":%ld}\n", nf, checked, mismatches);
  } catch (const std::exception& e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
  return mismatches ? 1 : 0;
}
