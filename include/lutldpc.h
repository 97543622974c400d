/* lutldpc.h -- C ABI of the B200-native batched LUT-LDPC decoder (liblutldpc_b200.so).
 *
 * The reference (mmeidlinger/lut_ldpc) has no FFI layer: its boundary for this path is the C++ class
 * lut_ldpc::LDPC_Code_LUT behind itpp::Channel_Code (src/LDPC_Code_LUT.hpp:86-285) called from
 * LDPC_BER_Sim::sim_snr_point (src/LDPC_BER_Sim.cpp:246-311).  Each entry point below names the
 * reference interface it replaces (paths relative to the reference tree).  INTEGRATION.md shows the
 * C++ stub a maintainer adds on the reference side.
 *
 * Conventions: plain pointers and sizes only; every call returns LUTLDPC_OK (0) or a negative
 * status and never throws/aborts across the ABI; lutldpc_last_error() gives the message of the last
 * failing call on that handle (or of the last failing create when handle == NULL).  A handle is bound
 * to the CUDA device that was current when it was created and is not thread-safe (the reference
 * object is not re-entrant either: lut_decode mutates the member `msgs`, src/LDPC_Code_LUT.hpp:311).
 * There is no CPU fallback: without a CUDA device create fails with LUTLDPC_ERR_CUDA.
 *
 * Labels: a quantised value m is the reference's integer in [0,Nq), m < Nq/2 <=> negative LLR <=>
 * bit 1 (src/LDPC_Code_LUT.cpp:260-261,275).  Frame-major buffers: element (f, v) at f*nvar + v.
 */
#ifndef LUTLDPC_H
#define LUTLDPC_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LUTLDPC_ABI_VERSION 1

enum {
  LUTLDPC_OK = 0,
  LUTLDPC_ERR_ARG = -1,     /* bad argument (NULL, size, range) */
  LUTLDPC_ERR_MODEL = -2,   /* decoder description inconsistent or unsupported */
  LUTLDPC_ERR_CUDA = -3,    /* CUDA runtime error / no device */
  LUTLDPC_ERR_STATE = -4    /* call not valid in the handle's current state */
};
enum { LUTLDPC_INIT_CONT = 0, LUTLDPC_INIT_QCHA = 1 };   /* LDPC_Code_LUT::initial_message_mode_t */
enum { LUTLDPC_TRACE_INIT = 0, LUTLDPC_TRACE_CN = 1, LUTLDPC_TRACE_VN = 2 };

typedef struct lutldpc_decoder lutldpc_decoder;

/* Flattened LDPC_Code_LUT: exactly the variables LDPC_Code_LUT::save_code writes
 * (src/LDPC_Code_LUT.cpp:643-697); trees as the reference's own text serialisation
 * (operator<< on Array<Array<LUT_Tree>>, src/LUT_Tree.cpp:847-864). */
typedef struct lutldpc_desc {
  int32_t nvar, nchk, nchk_lin_indep, num_edges, max_iters;
  int32_t Nq_Cha, min_lut;
  const int32_t* dv_vec;       /* [nvar]  */
  const int32_t* dc_vec;       /* [nchk]  */
  const int32_t* cn_msg_idx;   /* [num_edges] */
  const int32_t* Nq_Msg;       /* [max_iters] */
  const int32_t* reuse_vec;    /* [max_iters], 0/1 */
  const int32_t* Nq_Cha_2_Nq_Msg_map; /* [Nq_Cha] or NULL */
  const double* qb_Cha;        /* [Nq_Cha-1] */
  const double* qb_Msg;        /* [n_qb_Msg] */
  int32_t n_qb_Msg;
  const char* var_tree_string; size_t var_tree_len;
  const char* chk_tree_string; size_t chk_tree_len;   /* may be NULL/0 when min_lut != 0 */
  const uint8_t* gen_parity;   /* optional dense P [gen_rows][gen_cols], codeword = [u, P*u] */
  int32_t gen_rows, gen_cols;
} lutldpc_desc;

typedef struct lutldpc_info {
  int32_t nvar, nchk, nchk_lin_indep, ninfo, num_edges, max_iters, Nq_Cha, min_lut, has_gen;
  int32_t n_degree_classes;
  int64_t lookups_per_frame_iter;        /* LUT evaluations the compiled programs do per VN pass */
  int64_t lookups_per_frame_iter_naive;  /* what the reference's recursion does: sum dv*(nodes) */
  double rate;
  int32_t n_var_programs;       /* (iteration slot, degree class) programs of the VN pass */
  int32_t n_fast_var_programs;  /* ... of which run on the compile-time specialised kernel (vn_fast.cuh) */
  int32_t onchip;               /* 1: batches decode in the on-chip regime (lutldpc_set_onchip) */
  int32_t onchip_smem_bytes;    /* shared memory per CTA of that regime (0: the decoder does not qualify) */
} lutldpc_info;

int lutldpc_abi_version(void);
/* CUDA devices visible to the library, and the device the calling thread creates its next handles on
 * (cudaGetDeviceCount / cudaSetDevice, so that host code needs no CUDA headers) */
int lutldpc_device_count(void);
int lutldpc_set_device(int device);
/* message of the last failure; handle may be NULL for create errors */
const char* lutldpc_last_error(const lutldpc_decoder* h);

/* ---- construction: LDPC_Code_LUT ctor + set_code + set_trees (src/LDPC_Code_LUT.cpp:51-169,488-541) from a
 *      lutldpc_desc, or the from-file ctor / load_code (:60-66,568-641) from a file.  The file is either the
 *      reference's own codec file (the IT++ it_file that LDPC_Code_LUT::save_code writes, :643-697, e.g. the
 *      lut_codec.it that ber_sim leaves in its results directory and Sim.codec_filename names) or the flat
 *      "bundle" dump of oracle/ref_driver.cpp; the two are told apart by their magic bytes. */
int lutldpc_create(const lutldpc_desc* desc, lutldpc_decoder** out);
int lutldpc_create_from_bundle(const char* bundle_path, lutldpc_decoder** out);
void lutldpc_destroy(lutldpc_decoder* h);
/* LDPC_Code_LUT::save_code (src/LDPC_Code_LUT.cpp:643-697): the decoder as the reference's own codec file (IT++ it_file,
 * same variables in the same order), readable by the reference's load_code / Sim.codec_filename, lut_ldpc_vhdl and
 * scripts/.  The generator is not stored (G_defined = false): a file to which LDPC_Generator_Systematic::save has
 * appended carries IT++'s "Fileversion" 3 in place of the codec's 1 and the reference's load_code refuses it (:578-579).
 * lutldpc_convert_codec: bundle or codec file in, codec file out; pure host code (no device needed). */
int lutldpc_save_code(const lutldpc_decoder* h, const char* path);
int lutldpc_convert_codec(const char* in_path, const char* out_it_path);
int lutldpc_get_info(const lutldpc_decoder* h, lutldpc_info* info);

/* LDPC_Code_LUT::set_exit_conditions(max_iters, psc, pisc), src/LDPC_Code_LUT.cpp:176-185.
 * max_iters must equal the number of designed iterations: the decision trees belong to the last one. */
int lutldpc_set_exit_conditions(lutldpc_decoder* h, int max_iters, int psc, int pisc);
/* LDPC_Code_LUT::set_initial_message_mode, src/LDPC_Code_LUT.hpp:237-239 */
int lutldpc_set_initial_message_mode(lutldpc_decoder* h, int mode);
/* the stream all later work of this handle is enqueued on (a cudaStream_t; NULL = default stream) */
int lutldpc_set_stream(lutldpc_decoder* h, void* cuda_stream);
/* how often (in iterations) the host looks at the count of unfinished frames to stop a batch early;
 * 0 = never look (fully asynchronous: all max_iters iterations are enqueued, finished frames idle) */
int lutldpc_set_poll_interval(lutldpc_decoder* h, int every_n_iterations);

/* ---- decode.  n_frames independent frames per call (the reference decodes one).
 *
 * lutldpc_decode_quantized == LDPC_Code_LUT::lut_decode(LLRin_cha, LLRin_msg, LLRout)
 * (src/LDPC_Code_LUT.cpp:259-353), the bit-exact entry:
 *   llr_cha, llr_msg  [n_frames][nvar] labels in [0,Nq_Cha) / [0,Nq_Msg(0))
 *   out_bits          [n_frames][nvar] 0/1 (LLRout; the first ninfo of a frame are decode()'s result)
 *   out_iters         [n_frames] lut_decode's return: 0, +n, +max_iters or -max_iters
 * Host pointers; the copies run on the handle's stream and the call returns after they completed. */
int lutldpc_decode_quantized(lutldpc_decoder* h, int n_frames, const uint8_t* llr_cha, const uint8_t* llr_msg,
                             uint8_t* out_bits, int32_t* out_iters);
/* Same with DEVICE pointers; asynchronous on the handle's stream (no host synchronisation when
 * poll_every == 0: all max_iters iterations are enqueued, finished frames idle). */
int lutldpc_decode_quantized_device(lutldpc_decoder* h, int n_frames, const uint8_t* d_llr_cha,
                                    const uint8_t* d_llr_msg, uint8_t* d_out_bits, int32_t* d_out_iters);
/* Packed form of lutldpc_decode_quantized (same decoder, same results; no counterpart in the reference, whose
 * containers spend an int per label and a byte per bit): 73 KB instead of 194 KB per N=64800 frame across PCIe.
 *   cha_msg          [n_frames][nvar]  one byte per symbol: LLRin_cha label in the low nibble, LLRin_msg label in the high
 *   out_bits_packed  [n_frames][ceil(nvar/8)]  LLRout, bit (v & 7) of byte v >> 3 = decision of variable node v
 *   out_iters        [n_frames] lut_decode's return (may be NULL for the _device form)
 * Host pointers (pinned memory makes the copies asynchronous) / device pointers as for lutldpc_decode_quantized. */
int lutldpc_decode_packed(lutldpc_decoder* h, int n_frames, const uint8_t* cha_msg, uint8_t* out_bits_packed,
                          int32_t* out_iters);
int lutldpc_decode_packed_device(lutldpc_decoder* h, int n_frames, const uint8_t* d_cha_msg, uint8_t* d_out_bits_packed,
                                 int32_t* d_out_iters);
/* lutldpc_decode_llr == LDPC_Code_LUT::decode(const vec& llr_in, bvec& syst_bits)
 * (src/LDPC_Code_LUT.cpp:204-239): quant_nonlin front end (src/common.cpp:120-138) + lut_decode.
 *   llr [n_frames][nvar] doubles;  out_bits [n_frames][nvar];  out_iters may be NULL. */
int lutldpc_decode_llr(lutldpc_decoder* h, int n_frames, const double* llr, uint8_t* out_bits, int32_t* out_iters);
int lutldpc_decode_llr_device(lutldpc_decoder* h, int n_frames, const double* d_llr, uint8_t* d_out_bits,
                              int32_t* d_out_iters);
/* front end only: quant_nonlin(llr, qb_Cha) and quant_nonlin(llr, qb_Msg) | Nq_Cha_2_Nq_Msg_map */
int lutldpc_quantize(lutldpc_decoder* h, int n_frames, const double* llr, uint8_t* llr_cha, uint8_t* llr_msg);
int lutldpc_quantize_device(lutldpc_decoder* h, int n_frames, const double* d_llr, uint8_t* d_llr_cha,
                            uint8_t* d_llr_msg);

/* lut_decode with the reference's output_verbosity >= 2 message dumps for ONE frame of the batch
 * (src/LDPC_Code_LUT.cpp:291-298,311-317,331-337): record r holds all num_edges messages as labels,
 * kind/iter as in LUTLDPC_TRACE_*.  At most max_rec records are written. */
int lutldpc_decode_quantized_traced(lutldpc_decoder* h, int n_frames, const uint8_t* llr_cha,
                                    const uint8_t* llr_msg, uint8_t* out_bits, int32_t* out_iters,
                                    int trace_frame, uint8_t* trace_msgs, int32_t* trace_kind,
                                    int32_t* trace_iter, int max_rec, int32_t* n_rec);

/* LDPC_Generator_Systematic::encode (itpp/itpp/comm/ldpc.cpp:1038): codeword = [u, P*u].
 * info [n_frames][ninfo] -> codeword [n_frames][nvar]; LUTLDPC_ERR_STATE without a generator. */
int lutldpc_encode(lutldpc_decoder* h, int n_frames, const uint8_t* info, uint8_t* codeword);
/* syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469: ok[f] = 1 iff H*bits_f == 0 */
int lutldpc_syndrome_check(lutldpc_decoder* h, int n_frames, const uint8_t* bits, uint8_t* ok);

/* Allocation-free variant of lutldpc_syndrome_check with DEVICE pointers.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine): its scratch (bit rows, flag words, work lists) lives in a static __device__ array of
 * LUTLDPC_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Calls needing more fail
 * with LUTLDPC_ERR_ARG (lutldpc_noalloc_max_frames gives the limit for the handle's code: 4096 frames at
 * nvar = 64800, 127744 at nvar = 2048).  d_bits [n_frames][nvar] 0/1, d_ok [n_frames] = 1 iff H*bits_f == 0;
 * asynchronous on the handle's stream.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams or handles) race; the array keeps the last call's contents and each call clears what it
 * needs. */
#define LUTLDPC_NOALLOC_SCRATCH_BYTES (32u << 20)
int lutldpc_syndrome_check_device_noalloc(lutldpc_decoder* h, int n_frames, const uint8_t* d_bits, uint8_t* d_ok);
int lutldpc_noalloc_max_frames(const lutldpc_decoder* h);

/* ---- fused Monte-Carlo step == the frame loop of LDPC_BER_Sim::sim_snr_point
 * (src/LDPC_BER_Sim.cpp:246-311): source -> encode -> BPSK -> AWGN -> LLR = 4x/N0 -> quantise ->
 * lut_decode -> BERC/BLERC counting, all on the device; only the counters cross to the host. */
typedef struct lutldpc_sim_params {
  double snr_db;          /* Eb/N0 in dB; N0 = 10^(-snr/10)/rate, sigma^2 = N0/2 (:248-249) */
  int64_t n_frames;       /* frames to simulate in this call */
  uint64_t seed;          /* Philox4x32-10 key */
  uint64_t first_frame;   /* global index of the first frame (Philox counter); disjoint ranges per GPU/call */
  int32_t zero_codeword;  /* LDPC.zero_codeword; 0 needs a generator in the decoder */
  int32_t batch_frames;   /* frames in flight per launch (0 = default) */
} lutldpc_sim_params;

#define LUTLDPC_ITER_HIST_MAX 256
typedef struct lutldpc_sim_counters {
  uint64_t frames, frame_errors;             /* BLERC ferc: blocks / block errors over the ninfo data bits */
  uint64_t data_bits, data_bit_errors;       /* BERC berc_data */
  uint64_t uncoded_bits, uncoded_bit_errors; /* BERC berc_uncoded (slicer on all nvar symbols) */
  uint64_t sum_iters, decode_failures;       /* sum |lut_decode return|, # of negative returns */
  uint64_t iter_hist[LUTLDPC_ITER_HIST_MAX]; /* histogram of |return| (clamped) */
} lutldpc_sim_counters;
/* counters are ADDED to *acc (so successive calls / ranks accumulate) */
int lutldpc_simulate_snr(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc);

/* The same step for at most one batch of frames, handing back per frame what LDPC_Code_LUT::decode prints at
 * output_verbosity > 0 (the "Stimuli Pair" lines, src/LDPC_Code_LUT.cpp:228-238: the golden vectors of the reference's
 * VHDL flow): llr_cha [n_frames][nvar] = the quantised channel labels llr_in_cha, out_bits [n_frames][nvar] = LLRout. */
int lutldpc_simulate_snr_stimuli(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc,
                                 uint8_t* llr_cha, uint8_t* out_bits);

/* ---- per-kernel device times.  With profiling on, an event is recorded after every kernel launch
 * of the handle; lutldpc_get_profile returns, per kind, the summed device milliseconds and launch
 * counts since profiling was switched on.  Kinds: 0 reset, 1 pack, 2 channel, 3 min-sum check pass,
 * 4 syndrome, 5 finalize, 6 unpack, 7 count, 8 encode, 9 unused, 10 frame repack, 11 retire (results of repacked frames back), 12 on-chip lut_decode (whole batch), 16+c variable-node pass of degree class c,
 * 32+c decision pass, 48+c LUT check pass.  class_degree/class_nodes [16] describe the VN classes. */
#define LUTLDPC_PROFILE_KINDS 64
int lutldpc_set_profiling(lutldpc_decoder* h, int on);
int lutldpc_get_profile(lutldpc_decoder* h, double* ms, int64_t* launches, int32_t* class_degree,
                        int32_t* class_nodes);

/* ---- timing aid for bench.py: device-time (ms) of the last decode/simulate call's kernels, measured
 * with CUDA events on the handle's stream, and number of kernels it launched */
int lutldpc_last_call_stats(const lutldpc_decoder* h, double* device_ms, int64_t* kernel_launches,
                            int64_t* frame_iterations);

/* ---- frame repacking (no counterpart in the reference, which decodes one frame at a time).  A frame that
 * has returned keeps occupying its slot of the frame-interleaved batch; when fewer than `alpha` (0..1, default
 * 0.7) of the slots in the words that still hold an unfinished frame are in use, the unfinished frames are
 * gathered into a narrower layout and the remaining iterations run there.  Results are identical with
 * repacking on or off (frames are independent); it needs the per-iteration poll (lutldpc_set_poll_interval
 * >= 1).  on = 0 switches it off.  lutldpc_get_repack_count: repacks done by the handle so far. */
int lutldpc_set_repack(lutldpc_decoder* h, int on, double alpha);
int lutldpc_get_repack_count(const lutldpc_decoder* h, int64_t* repacks);

/* ---- parity-check matrices.
 * lutldpc_alist_graph == LDPC_Parity(alist) (itpp/itpp/base/gf2mat.cpp:46-210, itpp/itpp/comm/ldpc.cpp:167-180: the
 * row lists define H) followed by the graph flattening of LDPC_Code_LUT::decoder_parameterization
 * (src/LDPC_Code_LUT.cpp:488-541): dv_vec / dc_vec / cn_msg_idx with the reference's edge numbering, i.e. the graph
 * arrays of lutldpc_desc.  rank_mode: 0 = leave nchk_lin_indep at -1, 1 = GF(2) row rank on the host, 2 = on the
 * device (:493-499; the reference's dense elimination takes 528 s for N=64800, the device elimination seconds).
 * The arrays are malloc'ed by the library; release them with lutldpc_graph_free. */
typedef struct lutldpc_graph {
  int32_t nvar, nchk, num_edges, nchk_lin_indep;
  int32_t* dv_vec;      /* [nvar] */
  int32_t* dc_vec;      /* [nchk] */
  int32_t* cn_msg_idx;  /* [num_edges] */
} lutldpc_graph;
int lutldpc_alist_graph(const char* alist_path, int rank_mode, lutldpc_graph* out);
void lutldpc_graph_free(lutldpc_graph* g);
/* lutldpc_create_systematic == LDPC_Generator_Systematic::construct (itpp/itpp/comm/ldpc.cpp:894-999) for the
 * parity-check matrix of `src`, on the device: Gauss-Jordan elimination over GF(2) finds nchk_lin_indep independent
 * (parity) columns, the columns of H are renumbered to [information | parity] as construct() permutes *H in place
 * (:976-986), and *out is a decoder for that column-permuted code (same trees and thresholds) that carries the
 * generator: lutldpc_encode gives [u, P*u] and lutldpc_simulate_snr accepts zero_codeword = 0.  Parity columns are
 * taken in natural column order (construct()'s natural_ordering variant; the reference's default draws a random order,
 * so its permutation differs -- both are systematic generators of the same code).  col_perm (may be NULL) receives
 * the new index of every old column. */
int lutldpc_create_systematic(const lutldpc_decoder* src, lutldpc_decoder** out, int32_t* col_perm);

/* ---- LUT design on the host == LDPC_Code_LUT::design_luts (src/LDPC_Code_LUT.cpp:699-746): tree templates
 * (get_lut_tree_templates, src/LDPC_DE.cpp:1124-1290: "auto_bin_balanced" or "filename=<tree template ini>"),
 * discrete density evolution with LUT export (LDPC_DE_LUT::evolve, :198-326, joint_root strategy, min-sum or LUT
 * check nodes), quantiser thresholds (get_quant_bound, :561-601) and Nq_Cha_2_Nq_Msg_map, for the empirical ensemble
 * of the code (get_empirical_ensemble, src/LDPC_Ensemble.cpp:391-423).  Pure host code, no device needed.  The
 * result is table-exact: thresholds and trees equal the reference's element for element, the trees come back in the
 * reference's own text serialisation -- exactly the var_tree_string / chk_tree_string of lutldpc_desc, so
 * lutldpc_alist_graph + lutldpc_design_luts + lutldpc_create is LDPC_BER_Sim_LUT::load without a codec file.
 * sigma2 = design_thr^2 or 10^(-design_SNRdB/10) / (2 rate) (src/LDPC_BER_Sim.cpp:478-484).
 * Arrays of the result are malloc'ed by the library; release them with lutldpc_design_free. */
typedef struct lutldpc_design_params {
  const int32_t* dv_vec; int32_t nvar;      /* variable-node degrees */
  const int32_t* dc_vec; int32_t nchk;      /* check-node degrees */
  int32_t max_iters, Nq_Cha, min_lut;
  const int32_t* Nq_Msg;                    /* [max_iters] */
  const int32_t* reuse_vec;                 /* [max_iters] or NULL */
  double sigma2;
  const char* tree_method;                  /* NULL = "auto_bin_balanced" */
} lutldpc_design_params;
typedef struct lutldpc_design_result {
  double* qb_Cha; int32_t n_qb_Cha;
  double* qb_Msg; int32_t n_qb_Msg;
  int32_t* Nq_Cha_2_Nq_Msg_map; int32_t n_map;
  char* var_tree_string; size_t var_tree_len;
  char* chk_tree_string; size_t chk_tree_len;
} lutldpc_design_result;
int lutldpc_design_luts(const lutldpc_design_params* p, lutldpc_design_result* out);
void lutldpc_design_free(lutldpc_design_result* r);

/* ---- on-chip regime (no counterpart in the reference).  Decoders with min-sum checks whose per-column state
 * (E message words, N channel words, N hard bytes for 8 frames) fits shared memory at least twice per SM -- the
 * (3,6) N=1000, irregular N=500 and (6,32) N=2048 codes -- decode a batch in ONE persistent kernel: a CTA keeps the
 * 8 frames of a word column in shared memory for all iterations, decides their early termination itself and then
 * takes the next column.  Larger codes stream their messages through HBM once per half-iteration.  Results are
 * identical in both regimes.  on = 0 forces the HBM regime; on = 1 on a decoder that does not qualify is
 * LUTLDPC_ERR_STATE.  Traced decodes always take the HBM regime. */
int lutldpc_set_onchip(lutldpc_decoder* h, int on);

/* ---- BP comparison decoder == what ber_sim runs for a [BP] section (src/LDPC_BER_Sim.cpp:74-78,157-244,
 * prog/ber_sim.cpp:136-147): IT++'s LDPC_Code (itpp/itpp/comm/ldpc.h:757-1000) with an LLR_calc_unit
 * (itpp/itpp/comm/llr.h:120-330), batched.  Integer arithmetic on QLLR = int32 throughout, so results are bit-identical
 * to LDPC_Code::bp_decode (itpp/itpp/comm/ldpc.cpp:1398-1622): same Boxplus table, same saturation, same operation
 * order per node degree (incl. the hard-coded degree-6 check update whose association differs from the general one).
 *   lutldpc_bp_create           LDPC_Code(H[, G]) + set_exit_conditions + set_llrcalc   (src/LDPC_BER_Sim.cpp:197-201)
 *                               with_generator: LDPC_Generator_Systematic(H) first -- the decoder is then for the
 *                               column-permuted code [information | parity] and carries the encoder (col_perm as in
 *                               lutldpc_create_systematic)
 *   lutldpc_bp_decode_llr       LDPC_Code::decode(const vec&) for n frames: to_qllr, bp_decode, bits = qllrout < 0
 *   lutldpc_bp_decode_qllr      LDPC_Code::bp_decode for n frames: QLLR in, QLLR out, return values
 *   lutldpc_bp_simulate_snr     the frame loop of LDPC_BER_Sim::sim_snr_point around it (same noise as
 *                               lutldpc_simulate_snr for the same seed and frame index)
 * Buffers are host memory, frame-major.  Graph degrees the reference refuses at decode time (check degree < 2 or > 200,
 * variable degree 0: it_error, ldpc.cpp:1439-1443,1523-1527,1563) are refused at creation. */
typedef struct lutldpc_bp lutldpc_bp;
typedef struct lutldpc_bp_params {
  int32_t max_iters;               /* BP.max_iter, default 30 */
  int32_t psc, pisc;               /* ber_sim sets both to LDPC.parity_check_iter (src/LDPC_BER_Sim.cpp:200) */
  int32_t qllr_scale_res;          /* Dint1, default 12 */
  int32_t qllr_table_size;         /* Dint2, default 300; 0 = min-sum */
  int32_t qllr_spacing_res;        /* Dint3, default 7 */
  int32_t qllr_total_res;          /* Dint4, default 8*sizeof(QLLR)-4 = 28 */
} lutldpc_bp_params;
int lutldpc_bp_create(const lutldpc_graph* g, const lutldpc_bp_params* p, int with_generator, lutldpc_bp** out,
                      int32_t* col_perm);
/* bp_codec.it == IT++'s LDPC_Code::save_code / load_code file (itpp/itpp/comm/ldpc.cpp:1250-1322) with the block
 * LDPC_Generator_Systematic::save appends (:1002-1009): what ber_sim saves for a [BP] run when the seed equals
 * Sim.save_codec and loads through Sim.codec_filename + Sim.codec_type = BP (src/LDPC_BER_Sim.cpp:207-239).
 * lutldpc_bp_convert_alist: alist in, codec file (no generator) out; pure host code. */
int lutldpc_bp_create_from_file(const char* codec_path, const lutldpc_bp_params* p, lutldpc_bp** out);
int lutldpc_bp_save_code(lutldpc_bp* h, const char* path);
int lutldpc_bp_convert_alist(const char* alist_path, const char* out_it_path);
void lutldpc_bp_destroy(lutldpc_bp* h);
const char* lutldpc_bp_last_error(const lutldpc_bp* h);   /* h == NULL: error of the last failed lutldpc_bp_create */
int lutldpc_bp_get_info(const lutldpc_bp* h, lutldpc_info* info);   /* nvar, nchk, nchk_lin_indep, ninfo, num_edges,
                                                                        max_iters, has_gen, rate; rest 0 */
int lutldpc_bp_decode_llr(lutldpc_bp* h, int n_frames, const double* llr, uint8_t* out_bits /* [n][nvar] */,
                          int32_t* out_iters /* bp_decode's return values, may be NULL */);
int lutldpc_bp_decode_qllr(lutldpc_bp* h, int n_frames, const int32_t* qllr_in, int32_t* qllr_out, int32_t* out_iters);
int lutldpc_bp_encode(lutldpc_bp* h, int n_frames, const uint8_t* info, uint8_t* codeword);
int lutldpc_bp_simulate_snr(lutldpc_bp* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc);
/* device ms and kernel launches of the last decode/simulate call, like lutldpc_last_call_stats */
int lutldpc_bp_last_call_stats(const lutldpc_bp* h, double* device_ms, int64_t* kernel_launches, int64_t* frame_iterations);
/* frames that have returned are dropped from the batch once at most half of the slots are still running (the
 * analogue of lutldpc_set_repack); on = 0 keeps every frame in its slot.  Results do not depend on it. */
int lutldpc_bp_set_repack(lutldpc_bp* h, int on);

#ifdef __cplusplus
}
#endif
#endif
