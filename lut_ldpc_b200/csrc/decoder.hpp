// The object behind a lutldpc_decoder handle: host model, device copies, compiled node programs,
// batch buffers and the pass schedule of LDPC_Code_LUT::lut_decode (src/LDPC_Code_LUT.cpp:259-353).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/lutldpc.h"
#include "gf2.hpp"
#include "kernels.hpp"
#include "onchip.hpp"
#include "model.hpp"

namespace lutldpc {

enum ProfileKind {  // indices of lutldpc_get_profile
  KIND_RESET = 0, KIND_PACK = 1, KIND_CHANNEL = 2, KIND_CN = 3, KIND_SYNDROME = 4, KIND_FINALIZE = 5,
  KIND_UNPACK = 6, KIND_COUNT = 7, KIND_ENCODE = 8,
  KIND_REPACK = 10, KIND_RETIRE = 11,
  KIND_ONCHIP = 12,  // the whole lut_decode of a batch in one persistent kernel (onchip.cu)  // frame repacking: gather of the unfinished frames / results back
  KIND_VN0 = 16, KIND_DEC0 = 32, KIND_CHK0 = 48  // + degree-class index
};

template <class T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  ~DevBuf() { release(); }
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  cudaError_t ensure(size_t count) {  // grow-only
    if (count <= n) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
    if (e == cudaSuccess) n = count;
    else p = nullptr;
    return e;
  }
  cudaError_t upload(const std::vector<T>& v) {
    cudaError_t e = ensure(v.size() ? v.size() : 1);
    if (e != cudaSuccess || v.empty()) return e;
    return cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
  }
};

struct ProgRef {
  size_t off = 0;  // into the program blob
  uint32_t bytes = 0, n_slots = 0, n_in = 0;
  int64_t n_ops = 0, n_naive = 0;
  uint32_t tables_off = 0;  // of the table blob inside the image
  int fast_dv = 0, q_out = 0;  // fast_dv != 0: the compile-time specialised kernel applies
};

struct NodeClass {  // all variable (or check) nodes of one degree slot
  int degree = 0, n = 0;
  DevBuf<int32_t> list;
  DevBuf<int2> pairs;  // variable classes: (node, first edge id) in list order
};

// streams, events and double-buffered device staging of the pipelined host-buffer decode (capi.cu: decode_host)
struct HostPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_used[2] = {nullptr, nullptr}, ev_dec[2] = {nullptr, nullptr},
              ev_out[2] = {nullptr, nullptr};
  DevBuf<uint8_t> lab_cha[2], lab_msg[2], out[2];
  DevBuf<int32_t> its[2];
  DevBuf<double> llr[2];
  ~HostPipe() {
    for (int i = 0; i < 2; ++i)
      for (cudaEvent_t e : {ev_in[i], ev_used[i], ev_dec[i], ev_out[i]})
        if (e) cudaEventDestroy(e);
    if (s_in) cudaStreamDestroy(s_in);
    if (s_out) cudaStreamDestroy(s_out);
  }
};

// device buffers of one repacked layout (decoder.cu: run_passes); two of them alternate
struct RepackCtx {
  DevBuf<uint32_t> msgs, cha, flags, cols;
  DevBuf<uint8_t> hard;
  DevBuf<int32_t> iters, fid;
};

struct TraceSink {
  int frame = -1, max_rec = 0, n_rec = 0;
  uint8_t* msgs = nullptr;
  int32_t* kind = nullptr;
  int32_t* iter = nullptr;
  bool live = false;
};

}  // namespace lutldpc

struct lutldpc_decoder {
  lutldpc::DecoderModel m;
  int device = 0;
  cudaStream_t stream = nullptr;
  // psc/pisc: the reference constructs LDPC_Code_LUT with psc(true), pisc(false) (src/LDPC_Code_LUT.cpp:46-47)
  int max_iters = 0, psc = 1, pisc = 0, init_mode = LUTLDPC_INIT_CONT, poll_every = 1;
  std::string err;

  lutldpc::DevBuf<int32_t> d_vn_edge0, d_dv, d_cn_start, d_cn_msg_idx, d_chk_equ_idx, d_cha2msg;
  lutldpc::DevBuf<double> d_qb_cha, d_qb_msg;
  lutldpc::DevBuf<uint8_t> d_prog_blob, d_gen;
  // bit-packed parity part of a generator made on the device (lutldpc_create_systematic): [gen_rows][gen_bits_wp] words
  uint32_t* d_gen_bits = nullptr;
  int gen_bits_wp = 0, gen_bits_rows = 0, gen_bits_cols = 0;
  bool has_generator() const { return !m.gen.empty() || d_gen_bits != nullptr; }
  lutldpc::GraphArrays g;
  std::vector<lutldpc::NodeClass> vn_classes, cn_classes;
  std::vector<lutldpc::NodeClass> cn_exact;  // min-sum: checks grouped by degree, list = edge rows [n][degree]
  bool use_cn_exact = false;
  // fused VN launch: the classes that run on the specialised kernel, and their interleaved tiles
  std::vector<int> fused_classes;
  lutldpc::DevBuf<int4> d_fused_tiles;
  int n_fused_tiles = 0, fused_max_dv = 0;
  std::vector<lutldpc::ProgRef> var_progs;  // [iter slot][vn class]; last slot = decision trees
  std::vector<lutldpc::ProgRef> chk_progs;  // [iter slot][cn class]
  int n_iter_slots = 0;
  bool use_fast_vn = true;  // LUTLDPC_GENERIC_VN=1 in the environment forces the interpreter (tests)

  // on-chip regime (onchip.cu): codes whose per-column state fits shared memory decode in one persistent kernel
  bool onchip_ok = false;      // batches take the regime (the decoder qualifies and it is switched on)
  bool onchip_fits = false;    // the decoder qualifies
  bool batch_onchip = false;   // the current batch takes it (traced decodes never do)
  lutldpc::OnchipParams oc;    // static part; the per-batch pointers are filled at launch
  lutldpc::DevBuf<uint8_t> d_oc_images;
  lutldpc::DevBuf<int32_t> d_oc_iter_slot, d_oc_iter_q;
  lutldpc::DevBuf<unsigned int> d_oc_queue;
  lutldpc::DevBuf<int4> d_oc_vn_items, d_oc_cn_items;
  std::vector<lutldpc::NodeClass> oc_cn;  // checks by degree, list = edge rows [n][degree] (any degree)
  lutldpc::DevBuf<uint32_t> buf_msg0;
  int oc_ctas = 0, oc_ctas_per_sm = 0, oc_max_fast_dv = 0;
  size_t oc_state_bytes = 0;

  // batch state
  lutldpc::BatchBuffers b;
  lutldpc::DevBuf<uint32_t> buf_msgs, buf_cha, buf_flags, buf_cols;
  lutldpc::DevBuf<uint8_t> buf_hard, buf_lab_cha, buf_lab_msg, buf_out, buf_trace;
  lutldpc::HostPipe pipe;
  lutldpc::DevBuf<int32_t> buf_iters;
  lutldpc::DevBuf<double> buf_llr;
  lutldpc::DevBuf<uint8_t> buf_cw;                 // transmitted codeword bits, [nvar][W] like hard
  lutldpc::DevBuf<uint32_t> buf_ferr;              // per-word frame-error flags (byte form)
  lutldpc::DevBuf<unsigned long long> buf_cnt;     // device image of lutldpc_sim_counters
  uint32_t* h_active = nullptr;  // host-mapped pinned [8]: counters written by the finalize kernel
  uint32_t* d_active_mirror = nullptr;  // device view of h_active
  cudaEvent_t ev_lag = nullptr;
  bool lag_poll = true;
  int tail_lag_frames = 1024;  // lagged polls again once this few frames are unfinished
  // frame repacking: when fewer than repack_alpha of the slots in the live words hold an unfinished frame, the
  // unfinished frames are gathered into a narrower layout (LUTLDPC_REPACK=0 switches it off)
  lutldpc::RepackCtx rctx[2];
  lutldpc::DevBuf<int32_t> buf_rsrc, buf_ow_first;
  lutldpc::DevBuf<uint8_t> buf_ow_mask;
  bool repack_on = true;
  double repack_alpha = 0.7;
  int64_t repacks = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t ev_sim0 = nullptr, ev_sim1 = nullptr;  // lutldpc_simulate_snr: brackets of the whole call
  bool have_stats = false;
  int64_t launches = 0, last_frames = 0, stat_frame_iters = -1;

  // per-kernel device times (lutldpc_set_profiling)
  bool profiling = false;
  std::vector<cudaEvent_t> evpool;
  std::vector<int> evkind;
  int evn = 0;
  double prof_ms[LUTLDPC_PROFILE_KINDS] = {0};
  int64_t prof_n[LUTLDPC_PROFILE_KINDS] = {0};
  void launched(int kind);
  int flush_profile();

  ~lutldpc_decoder();
  int fail(int code, const std::string& msg) { err = msg; return code; }
  int cuda_fail(cudaError_t e, const char* where);
  int init_device();          // uploads + program compilation
  int init_onchip();          // on-chip regime set-up (part of init_device)
  int ensure_batch(int n_frames, bool onchip = false);
  // one batch, labels already on the device (frame-major); results left in b.hard / b.iters
  int run_batch(int n_frames, const uint8_t* d_cha_lab, const uint8_t* d_msg_lab, lutldpc::TraceSink* tr);
  int begin_batch(int n_frames, bool traced = false);  // buffers + flag reset + start event; picks the regime
  int run_decode(lutldpc::TraceSink* tr);   // lut_decode of the initialised batch in the regime begin_batch picked
  int run_onchip();                         // ... one persistent kernel, state in shared memory
  int run_passes(lutldpc::TraceSink* tr);   // ... all iterations + decision as separate passes over HBM
  // gathers the unfinished frames of `cur` (n_live of them) into rctx[slot]; `cur` becomes the new layout
  int repack_into(lutldpc::BatchBuffers& cur, int slot, int n_live);
  int retire_from(const lutldpc::BatchBuffers& cur);  // results of a repacked layout back into b
  int max_batch_frames() const;
};
