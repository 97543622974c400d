// TEST INFRASTRUCTURE: drives the C++ facade lut_ldpc_b200/host/ldpc_code_lut.hpp one frame per call, the way the
// reference's callers use LDPC_Code_LUT, against a LUTFIX01 fixture written by the compiled reference
// (layout: oracle/ref_driver.cpp).  usage: facade_check <bundle> <fixture with LLRs>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iterator>
#include <vector>

#include "lut_ldpc_b200/host/ldpc_code_lut.hpp"

using namespace lut_ldpc_b200;

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::ifstream is(argv[2], std::ios::binary);
  std::vector<char> b((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
  if (b.size() < 40 || memcmp(b.data(), "LUTFIX01", 8) != 0) { fprintf(stderr, "bad fixture\n"); return 2; }
  int32_t hdr[6];
  memcpy(hdr, b.data() + 8, 24);
  const int N = hdr[0], nf = hdr[2], with_llr = hdr[4], zero_cw = hdr[5];
  if (!with_llr) { fprintf(stderr, "fixture has no LLRs\n"); return 2; }
  long mismatches = 0, checked = 0;
  try {
    LDPC_Code_LUT C(argv[1]);
    C.set_exit_conditions(C.get_nrof_iterations(), true, true);  // (max_iters, psc, pisc) as src/LDPC_Code_LUT.hpp:186
    const int K = C.get_ninfo();
    size_t off = 40;
    for (int f = 0; f < nf; ++f) {
      vec llr(N);
      memcpy(llr.data(), b.data() + off, 8 * (size_t)N); off += 8 * (size_t)N;
      const uint8_t* cw = (const uint8_t*)b.data() + off; off += N;
      const uint8_t* cha = (const uint8_t*)b.data() + off; off += N;
      const uint8_t* msg = (const uint8_t*)b.data() + off; off += N;
      const uint8_t* out = (const uint8_t*)b.data() + off; off += N;
      int32_t ret; memcpy(&ret, b.data() + off, 4); off += 4;
      Channel_Code* cc = &C;
      bvec syst = cc->decode(llr);                                     // decode(const vec&) -> first ninfo bits
      mismatches += (int)syst.size() != K || memcmp(syst.data(), out, K) != 0; ++checked;
      bvec full;
      int r = C.lut_decode(ivec(cha, cha + N), ivec(msg, msg + N), full);  // lut_decode: all nvar bits + return value
      mismatches += r != ret; ++checked;
      mismatches += memcmp(full.data(), out, N) != 0; ++checked;
      mismatches += C.syndrome_check(full) != (ret >= 0); ++checked;   // a non-negative return is a valid codeword
      if (!zero_cw && C.has_generator()) {
        bvec enc = cc->encode(bvec(cw, cw + K));                       // systematic: [u, P*u]
        mismatches += memcmp(enc.data(), cw, N) != 0; ++checked;
        mismatches += !C.syndrome_check(enc); ++checked;
      }
    }
    bool threw = false;
    try { bvec dummy; C.decode(bvec(N), dummy); } catch (const std::runtime_error&) { threw = true; }
    mismatches += !threw; ++checked;                                   // hard-input decoding is an error, like the reference
    printf("{\"frames\":%d,\"checked\":%ld,\"mismatches\":%ld}\n", nf, checked, mismatches);
  } catch (const std::exception& e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
  return mismatches ? 1 : 0;
}
