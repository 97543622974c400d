/* TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
 *
 * Plain-C restatement of the reference's LUT decode path, used as the parity checker for the
 * CUDA kernels (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline leg only).
 *
 * Parity status: PINNED.  The reference ships no golden vectors for this path,
 * so this restatement is checked against outputs of the reference itself, compiled unmodified
 * into oracle/_ref/lutldpc_ref (oracle/Makefile) and run in the build container: the committed
 * fixtures under tests/golden/ (made by tests/golden/make_golden.py) hold the reference's
 * quantised inputs, hard decisions, lut_decode return values and per-iteration message states.
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference).
 */
#ifndef LUTLDPC_ORACLE_H
#define LUTLDPC_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_decoder orc_decoder;

enum { ORC_NVAR = 0, ORC_NCHK, ORC_NCHK_LIN_INDEP, ORC_NUM_EDGES, ORC_MAX_ITERS, ORC_NQ_CHA,
       ORC_MIN_LUT, ORC_NINFO, ORC_HAS_GEN, ORC_N_ITER_SLOTS, ORC_N_DEG_SLOTS };
enum { ORC_INIT_CONT = 0, ORC_INIT_QCHA = 1 };
enum { ORC_TRACE_INIT = 0, ORC_TRACE_CN = 1, ORC_TRACE_VN = 2 };

/* Reads a decoder bundle (format: oracle/ref_driver.cpp write_bundle). NULL + message on error. */
orc_decoder* orc_load_bundle(const char* path, char* err, int errlen);
void orc_free(orc_decoder* d);
int orc_info(const orc_decoder* d, int what);
/* copies of flat arrays, for tests (dst sized by the caller) */
void orc_get_dv(const orc_decoder* d, int32_t* dst);
void orc_get_dc(const orc_decoder* d, int32_t* dst);
void orc_get_cn_msg_idx(const orc_decoder* d, int32_t* dst);
void orc_get_nq_msg(const orc_decoder* d, int32_t* dst);
int orc_get_qb(const orc_decoder* d, int which /*0 cha, 1 msg*/, double* dst);

/* LDPC_Code_LUT::set_exit_conditions, src/LDPC_Code_LUT.cpp:176-185 */
void orc_set_exit_conditions(orc_decoder* d, int max_iters, int psc, int pisc);

/* quant_nonlin, src/common.cpp:120-138 */
int orc_quant_nonlin(double x, const double* boundaries, int nb);
/* front end of LDPC_Code_LUT::decode, src/LDPC_Code_LUT.cpp:204-221 */
void orc_quantise(const orc_decoder* d, const double* llr, int init_mode, uint8_t* cha, uint8_t* msg);

/* LDPC_Code_LUT::lut_decode, src/LDPC_Code_LUT.cpp:259-353.  Returns 0 / +n / +-max_iters.
 * Optional trace (the reference's verbosity>=2 dumps, :291-337): up to max_rec records of
 * num_edges bytes each are written to trace_msgs, with kind/iter per record. */
int orc_lut_decode(orc_decoder* d, const uint8_t* cha, const uint8_t* msg, uint8_t* out,
                   uint8_t* trace_msgs, int32_t* trace_kind, int32_t* trace_iter, int max_rec,
                   int32_t* n_rec);
/* n_frames independent frames, frame-major [n_frames][nvar] */
void orc_lut_decode_batch(orc_decoder* d, int n_frames, const uint8_t* cha, const uint8_t* msg,
                          uint8_t* out, int32_t* ret);
/* systematic encoder [u, P*u] (LDPC_Generator_Systematic::encode, itpp/itpp/comm/ldpc.cpp:1038);
 * returns -1 if the bundle holds no generator */
int orc_encode(const orc_decoder* d, const uint8_t* info, uint8_t* codeword);
/* H*b == 0 ?  (syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469) */
int orc_syndrome_ok(const orc_decoder* d, const uint8_t* bits);

#ifdef __cplusplus
}
#endif
#endif
