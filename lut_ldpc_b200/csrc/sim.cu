// Fused Monte-Carlo front/back end around the decoder: the frame loop of
// LDPC_BER_Sim::sim_snr_point (src/LDPC_BER_Sim.cpp:246-311) on the device.
//   source      bitsin = zeros | randb(K)                                   :263-268
//   encode      [u, P*u]  (LDPC_Generator_Systematic::encode)               itpp/itpp/comm/ldpc.cpp:1038
//   BPSK        0 -> +1, 1 -> -1                                            itpp/itpp/comm/modulator.cpp:351-357
//   AWGN        x = s + sqrt(N0/2) * n                                      itpp/itpp/comm/channel.cpp:1560-1568
//   LLR         4x/N0, slicer bit = (x > 0 ? 0 : 1)                         modulator.cpp:367-392
//   quantise    quant_nonlin(qb_Cha) and quant_nonlin(qb_Msg) | map         src/LDPC_Code_LUT.cpp:207-221
//   count       BERC (data, uncoded), BLERC                                 itpp/itpp/comm/error_counters.cpp
// The reference draws its normals from dSFMT + ziggurat; here Philox4x32-10 + Box-Muller, so parity
// with the reference is statistical (BER/FER inside the Monte-Carlo confidence interval), while the
// decode itself stays bit-exact.
#include <cstring>
#include <vector>

#include "core.cuh"
#include "decoder.hpp"
#include "rng.cuh"

using namespace lutldpc;

namespace {

__device__ __forceinline__ int quant_nonlin_s(double x, const double* b, int nb) {  // src/common.cpp:120-138
  int idx = 0;
  for (int i = 0; i < nb; ++i) {
    if (x > b[i]) ++idx;
    else break;
  }
  return idx;
}
// Same value for ascending thresholds (the designed quantisers are: src/LDPC_DE.cpp:561-601): the number of leading
// thresholds with x > b[i] is then the number of ALL thresholds below x -- a branch-free upper-bound search over the
// 16 padded entries (+inf beyond nb), 4 compares instead of up to 15.
__device__ __forceinline__ int quant_sorted16(double x, const double* b) {
  int idx = (x > b[7]) ? 8 : 0;
  idx += (x > b[idx + 3]) ? 4 : 0;
  idx += (x > b[idx + 1]) ? 2 : 0;
  idx += (x > b[idx]) ? 1 : 0;
  return idx;
}
__device__ __forceinline__ uint32_t lab2code_s(uint32_t m, int Nq, int q) {
  uint32_t nz = Nq >> 1;
  return m < nz ? ((1u << (q - 1)) | (nz - 1 - m)) : (m - nz);
}

// parity part: cw[K + r][w] = XOR_k P[r][k] & cw[k][w]   (dense GF(2) mat-vec, all frames of a word at once)
__global__ void __launch_bounds__(128) k_encode(const uint8_t* __restrict__ P, int rows, int K, int W,
                                                uint8_t* __restrict__ cw) {
  int w = blockIdx.y * blockDim.x + threadIdx.x;
  int r = blockIdx.x;
  if (w >= W || r >= rows) return;
  const uint8_t* row = P + (size_t)r * K;
  uint32_t acc = 0;
  for (int k = 0; k < K; ++k)
    if (row[k]) acc ^= cw[(size_t)k * W + w];  // row[k] is uniform over the CTA
  cw[(size_t)(K + r) * W + w] = (uint8_t)acc;
}

struct ChanParams {
  double sigma, four_over_N0;
  int n_qb_cha, n_qb_msg, use_map, Nq_cha, q_cha, Nq_msg, q_msg;
  int sorted;  // both threshold vectors ascend: binary search gives quant_nonlin's value
  int same_q;  // qb_Msg and qb_Cha are the same vector
};

// BPSK + AWGN + LLR + both quantisers -> channel labels, initial messages, channel hard decision,
// and the uncoded (slicer) error count.  Thread = (node quad 4q..4q+3, word): one Philox block per frame gives the
// four normals of the quad, so a thread draws 8 blocks for its 32 symbols.
__global__ void __launch_bounds__(128) k_channel(const uint8_t* __restrict__ cw, int zero_cw,
                                                 const int32_t* __restrict__ vn_edge0, int nvar, int W, int F,
                                                 uint64_t first_frame, uint32_t k0, uint32_t k1, ChanParams cp,
                                                 const double* __restrict__ qb_cha, const double* __restrict__ qb_msg,
                                                 const int32_t* __restrict__ cha2msg, uint32_t* __restrict__ cha,
                                                 uint32_t* __restrict__ msgs, uint8_t* __restrict__ hard,
                                                 uint32_t* __restrict__ msg0,
                                                 unsigned long long* __restrict__ cnt_uncoded_err) {
  __shared__ double s_qc[16], s_qm[16];
  __shared__ int s_map[16];
  if (threadIdx.x < 16) {
    const double inf = __longlong_as_double(0x7FF0000000000000LL);
    s_qc[threadIdx.x] = threadIdx.x < cp.n_qb_cha ? qb_cha[threadIdx.x] : inf;
    s_qm[threadIdx.x] = threadIdx.x < cp.n_qb_msg ? qb_msg[threadIdx.x] : inf;
    s_map[threadIdx.x] = (cp.use_map && threadIdx.x < cp.Nq_cha) ? cha2msg[threadIdx.x] : 0;
  }
  __syncthreads();
  const int w = blockIdx.y * blockDim.x + threadIdx.x;
  const int q = blockIdx.x, v0 = 4 * q;
  uint32_t nerr = 0;
  if (w < W) {
    uint32_t bits[4], cword[4] = {0, 0, 0, 0}, mword[4] = {0, 0, 0, 0}, slicer[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; ++i) bits[i] = (zero_cw || v0 + i >= nvar) ? 0u : cw[(size_t)(v0 + i) * W + w];
    for (int j = 0; j < 8; ++j) {
      const int f = 8 * w + j;
      if (f >= F) break;
      double nrm[4];
      normals4_for(first_frame + f, (uint32_t)q, k0, k1, nrm);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const double sgn = ((bits[i] >> j) & 1u) ? -1.0 : 1.0;
        const double x = sgn + cp.sigma * nrm[i];
        const double llr = cp.four_over_N0 * x;
        const int c = cp.sorted ? quant_sorted16(llr, s_qc) : quant_nonlin_s(llr, s_qc, cp.n_qb_cha);
        const int mm = cp.use_map ? s_map[c]
                       : cp.same_q ? c  // qb_Msg == qb_Cha element for element: the second quantiser repeats the first
                                   : (cp.sorted ? quant_sorted16(llr, s_qm) : quant_nonlin_s(llr, s_qm, cp.n_qb_msg));
        cword[i] |= lab2code_s((uint32_t)c, cp.Nq_cha, cp.q_cha) << (4 * j);
        mword[i] |= lab2code_s((uint32_t)mm, cp.Nq_msg, cp.q_msg) << (4 * j);
        slicer[i] |= (x > 0.0 ? 0u : 1u) << j;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int v = v0 + i;
      if (v >= nvar) break;
      cha[(size_t)v * W + w] = cword[i];
      hard[(size_t)v * W + w] = (uint8_t)nib_bits_to_byte(cword[i], cp.q_cha - 1);
      if (msg0) msg0[(size_t)v * W + w] = mword[i];  // on-chip regime
      else
        for (int e = vn_edge0[v]; e < vn_edge0[v + 1]; ++e) msgs[(size_t)e * W + w] = mword[i];
      nerr += __popc(slicer[i] ^ bits[i]);
    }
  }
  for (int o = 16; o; o >>= 1) nerr += __shfl_xor_sync(0xFFFFFFFFu, nerr, o);
  if ((threadIdx.x & 31) == 0 && nerr) atomicAdd(cnt_uncoded_err, (unsigned long long)nerr);
}

// data-bit errors over the first K symbols and per-frame error flags (BERC::count / BLERC::count)
__global__ void __launch_bounds__(128) k_count_data(const uint8_t* __restrict__ hard, const uint8_t* __restrict__ cw,
                                                    int zero_cw, int K, int W, int rows_per_cta,
                                                    uint32_t* __restrict__ ferr,
                                                    unsigned long long* __restrict__ cnt_data_err) {
  const int w = blockIdx.y * blockDim.x + threadIdx.x;
  uint32_t nerr = 0, any = 0;
  if (w < W) {
    int k0 = blockIdx.x * rows_per_cta, k1 = min(K, k0 + rows_per_cta);
    for (int k = k0; k < k1; ++k) {
      uint32_t d = hard[(size_t)k * W + w] ^ (zero_cw ? 0u : (uint32_t)cw[(size_t)k * W + w]);
      nerr += __popc(d);
      any |= d;
    }
    if (any) atomicOr(&ferr[w], any);
  }
  for (int o = 16; o; o >>= 1) nerr += __shfl_xor_sync(0xFFFFFFFFu, nerr, o);
  if ((threadIdx.x & 31) == 0 && nerr) atomicAdd(cnt_data_err, (unsigned long long)nerr);
}

// counters layout = lutldpc_sim_counters as u64[8 + LUTLDPC_ITER_HIST_MAX]
__global__ void k_count_frames(const uint32_t* __restrict__ ferr, const int32_t* __restrict__ iters, int F, int K,
                               int nvar, unsigned long long* __restrict__ cnt) {
  unsigned long long fe = 0, si = 0, nf = 0;
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
    fe += (ferr[f >> 3] >> (f & 7)) & 1u;
    int it = iters[f];
    nf += it < 0;
    it = it < 0 ? -it : it;
    si += it;
    atomicAdd(&cnt[8 + min(it, LUTLDPC_ITER_HIST_MAX - 1)], 1ull);
  }
  for (int o = 16; o; o >>= 1) {
    fe += __shfl_xor_sync(0xFFFFFFFFu, fe, o);
    si += __shfl_xor_sync(0xFFFFFFFFu, si, o);
    nf += __shfl_xor_sync(0xFFFFFFFFu, nf, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (fe) atomicAdd(&cnt[1], fe);
    if (si) atomicAdd(&cnt[6], si);
    if (nf) atomicAdd(&cnt[7], nf);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt[0], (unsigned long long)F);
    atomicAdd(&cnt[2], (unsigned long long)F * K);
    atomicAdd(&cnt[4], (unsigned long long)F * nvar);
  }
}

// channel codes [nvar][W] -> frame-major reference labels [F][nvar] (the llr_in_cha of the "Stimuli Pair" print)
__global__ void k_codes_to_labels(const uint32_t* __restrict__ cha, int N, int W, int F, int Nq, int q, uint8_t* out) {
  int w = blockIdx.y * blockDim.x + threadIdx.x, v = blockIdx.x;
  if (w >= W || v >= N) return;
  const uint32_t word = cha[(size_t)v * W + w], nz = Nq >> 1;
  for (int j = 0; j < 8; ++j) {
    int f = 8 * w + j;
    if (f >= F) break;
    const uint32_t c = (word >> (4 * j)) & 0xF, sign = (c >> (q - 1)) & 1, mag = c & ((1u << (q - 1)) - 1);
    out[(size_t)f * N + v] = (uint8_t)(sign ? nz - 1 - mag : nz + mag);
  }
}

// frame-major 0/1 bytes <-> bit-per-frame rows, for lutldpc_encode
__global__ void k_bits_to_rows(const uint8_t* __restrict__ in, int ncols_in, int K, int W, int F, uint8_t* rows) {
  int w = blockIdx.y * blockDim.x + threadIdx.x, k = blockIdx.x;
  if (w >= W || k >= K) return;
  uint32_t b = 0;
  for (int j = 0; j < 8; ++j) {
    int f = 8 * w + j;
    if (f < F) b |= (uint32_t)(in[(size_t)f * ncols_in + k] & 1) << j;
  }
  rows[(size_t)k * W + w] = (uint8_t)b;
}
__global__ void k_rows_to_bits(const uint8_t* __restrict__ rows, int N, int W, int F, uint8_t* out) {
  int w = blockIdx.y * blockDim.x + threadIdx.x, v = blockIdx.x;
  if (w >= W || v >= N) return;
  uint32_t b = rows[(size_t)v * W + w];
  for (int j = 0; j < 8; ++j) {
    int f = 8 * w + j;
    if (f < F) out[(size_t)f * N + v] = (b >> j) & 1;
  }
}

}  // namespace

#define CKH(call, where)                                   \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) return h->cuda_fail(e_, where); \
  } while (0)
#define CKL(where) CKH(cudaGetLastError(), where)

extern "C" {

int lutldpc_encode(lutldpc_decoder* h, int n_frames, const uint8_t* info, uint8_t* codeword) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (n_frames < 0 || (n_frames && (!info || !codeword))) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  if (!h->has_generator()) return h->fail(LUTLDPC_ERR_STATE, "decoder has no generator matrix");
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  const bool bits = h->d_gen_bits != nullptr;
  const int N = h->m.nvar, K = bits ? h->gen_bits_cols : h->m.gen_cols, rows = bits ? h->gen_bits_rows : h->m.gen_rows;
  if (K + rows != N) return h->fail(LUTLDPC_ERR_MODEL, "generator shape does not match nvar");
  cudaStream_t st = h->stream;
  const int maxb = h->max_batch_frames();
  for (int f0 = 0; f0 < n_frames; f0 += maxb) {
    int nf = std::min(maxb, n_frames - f0);
    int W = ((nf + 255) / 256) * 32;
    CKH(h->buf_cw.ensure((size_t)N * W), "alloc cw");
    CKH(h->buf_lab_cha.ensure((size_t)nf * N), "alloc");
    CKH(cudaMemcpyAsync(h->buf_lab_cha.p, info + (size_t)f0 * K, (size_t)nf * K, cudaMemcpyHostToDevice, st), "H2D info");
    dim3 gk(K, (W + 127) / 128), gr(rows, (W + 127) / 128), gn(N, (W + 127) / 128);
    k_bits_to_rows<<<gk, 128, 0, st>>>(h->buf_lab_cha.p, K, K, W, nf, h->buf_cw.p);
    CKL("bits_to_rows");
    if (bits) CKH(launch_encode_bits(h->d_gen_bits, h->gen_bits_wp, rows, K, W, h->buf_cw.p, st), "encode_bits");
    else k_encode<<<gr, 128, 0, st>>>(h->d_gen.p, rows, K, W, h->buf_cw.p);
    CKL("encode");
    CKH(h->buf_out.ensure((size_t)nf * N), "alloc out");
    k_rows_to_bits<<<gn, 128, 0, st>>>(h->buf_cw.p, N, W, nf, h->buf_out.p);
    CKL("rows_to_bits");
    CKH(cudaMemcpyAsync(codeword + (size_t)f0 * N, h->buf_out.p, (size_t)nf * N, cudaMemcpyDeviceToHost, st), "D2H cw");
    CKH(cudaStreamSynchronize(st), "sync");
  }
  return LUTLDPC_OK;
}

static int simulate_impl(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc, uint8_t* dump_cha,
                         uint8_t* dump_bits);

int lutldpc_simulate_snr(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc) {
  return simulate_impl(h, p, acc, nullptr, nullptr);
}

int lutldpc_simulate_snr_stimuli(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc,
                                 uint8_t* llr_cha, uint8_t* out_bits) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (!llr_cha || !out_bits) return h->fail(LUTLDPC_ERR_ARG, "NULL buffer");
  return simulate_impl(h, p, acc, llr_cha, out_bits);
}

static int simulate_impl(lutldpc_decoder* h, const lutldpc_sim_params* p, lutldpc_sim_counters* acc, uint8_t* dump_cha,
                         uint8_t* dump_bits) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (!p || !acc || p->n_frames < 0) return h->fail(LUTLDPC_ERR_ARG, "bad argument");
  const DecoderModel& m = h->m;
  if (!p->zero_codeword && !h->has_generator())
    return h->fail(LUTLDPC_ERR_STATE, "zero_codeword=0 needs a generator matrix in the decoder");
  if (h->init_mode == LUTLDPC_INIT_CONT && (int)m.qb_Msg.size() != m.Nq_Msg[0] - 1)
    return h->fail(LUTLDPC_ERR_STATE, "decoder has no message quantiser qb_Msg for continuous inputs");
  CKH(cudaSetDevice(h->device), "cudaSetDevice");
  cudaStream_t st = h->stream;
  const int N = m.nvar, K = m.ninfo();
  const bool gen_bits = h->d_gen_bits != nullptr;
  if (!p->zero_codeword && ((gen_bits ? h->gen_bits_cols : m.gen_cols) != K || (gen_bits ? h->gen_bits_rows : m.gen_rows) != N - K))
    return h->fail(LUTLDPC_ERR_MODEL, "generator shape does not match the code");
  // frames in flight: large batches run the decoder at its efficient point (N=64800: 32768 frames per batch decode
  // 5 % faster per frame than 8192: fewer repacks and tail iterations per frame)
  int batch = p->batch_frames > 0 ? p->batch_frames : 32768;
  batch = std::min(batch, h->max_batch_frames());
  batch = std::max(256, batch / 256 * 256);
  if (dump_cha && p->n_frames > batch) return h->fail(LUTLDPC_ERR_ARG, "lutldpc_simulate_snr_stimuli: n_frames exceeds one batch");
  const size_t n_cnt = 8 + LUTLDPC_ITER_HIST_MAX;
  CKH(h->buf_cnt.ensure(n_cnt), "alloc counters");
  CKH(cudaMemsetAsync(h->buf_cnt.p, 0, n_cnt * 8, st), "memset counters");
  // N0 = 10^(-snr/10) / rate, sigma^2 = N0/2   (src/LDPC_BER_Sim.cpp:248-249)
  const double N0 = pow(10.0, -p->snr_db / 10.0) / m.rate();
  ChanParams cp;
  cp.sigma = sqrt(N0 / 2.0);
  cp.four_over_N0 = 4.0 / N0;
  cp.n_qb_cha = (int)m.qb_Cha.size(); cp.n_qb_msg = (int)m.qb_Msg.size();
  cp.use_map = h->init_mode == LUTLDPC_INIT_QCHA;
  cp.Nq_cha = m.Nq_Cha; cp.q_cha = code_bits(m.Nq_Cha); cp.Nq_msg = m.Nq_Msg[0]; cp.q_msg = code_bits(m.Nq_Msg[0]);
  cp.sorted = 1;
  for (size_t i = 1; i < m.qb_Cha.size(); ++i) cp.sorted &= m.qb_Cha[i - 1] < m.qb_Cha[i];
  for (size_t i = 1; i < m.qb_Msg.size(); ++i) cp.sorted &= m.qb_Msg[i - 1] < m.qb_Msg[i];
  cp.same_q = m.qb_Cha == m.qb_Msg;
  const uint32_t k0 = (uint32_t)p->seed, k1 = (uint32_t)(p->seed >> 32);
  int64_t total_launches = 0;
  // whole-call events (all batches); owned by the handle, so no error path can leak them
  if (!h->ev_sim0) CKH(cudaEventCreate(&h->ev_sim0), "event");
  if (!h->ev_sim1) CKH(cudaEventCreate(&h->ev_sim1), "event");
  CKH(cudaEventRecord(h->ev_sim0, st), "event");
  for (int64_t done = 0; done < p->n_frames; done += batch) {
    int nf = (int)std::min<int64_t>(batch, p->n_frames - done);
    int rc = h->begin_batch(nf);
    if (rc) return rc;
    const int W = h->b.W;
    const uint64_t first = p->first_frame + (uint64_t)done;
    CKH(h->buf_ferr.ensure((size_t)W), "alloc ferr");
    CKH(cudaMemsetAsync(h->buf_ferr.p, 0, (size_t)W * 4, st), "memset ferr");
    dim3 gw((W + 127) / 128);
    if (!p->zero_codeword) {
      CKH(h->buf_cw.ensure((size_t)N * W), "alloc cw");
      k_info_bits<<<dim3(K, gw.x), 128, 0, st>>>(h->buf_cw.p, K, W, nf, first, k0, k1);
      CKL("info_bits");
      if (gen_bits) CKH(launch_encode_bits(h->d_gen_bits, h->gen_bits_wp, N - K, K, W, h->buf_cw.p, st), "encode_bits");
      else k_encode<<<dim3(N - K, gw.x), 128, 0, st>>>(h->d_gen.p, N - K, K, W, h->buf_cw.p);
      CKL("encode");
      h->launched(KIND_ENCODE);
    }
    k_channel<<<dim3((N + 3) / 4, gw.x), 128, 0, st>>>(h->buf_cw.p, p->zero_codeword != 0, h->g.vn_edge0, N, W, nf, first, k0, k1,
                                             cp, h->d_qb_cha.p, h->d_qb_msg.p, h->d_cha2msg.p, h->b.cha, h->b.msgs,
                                             h->b.hard, h->b.msg0, h->buf_cnt.p + 5);
    CKL("channel");
    h->launched(KIND_CHANNEL);
    if (dump_cha) {  // llr_in_cha of every frame, as labels
      CKH(h->buf_lab_cha.ensure((size_t)nf * N), "alloc labels");
      k_codes_to_labels<<<dim3(N, gw.x), 128, 0, st>>>(h->b.cha, N, W, nf, cp.Nq_cha, cp.q_cha, h->buf_lab_cha.p);
      CKL("codes_to_labels");
      CKH(cudaMemcpyAsync(dump_cha, h->buf_lab_cha.p, (size_t)nf * N, cudaMemcpyDeviceToHost, st), "D2H labels");
    }
    rc = h->run_decode(nullptr);
    if (rc) return rc;
    if (dump_bits) {  // LLRout of every frame
      CKH(h->buf_out.ensure((size_t)nf * N), "alloc out");
      CKH(launch_unpack(h->g, h->b, h->buf_out.p, st), "unpack");
      CKH(cudaMemcpyAsync(dump_bits, h->buf_out.p, (size_t)nf * N, cudaMemcpyDeviceToHost, st), "D2H bits");
    }
    int rows_per = 64;
    k_count_data<<<dim3((K + rows_per - 1) / rows_per, gw.x), 128, 0, st>>>(h->b.hard, h->buf_cw.p, p->zero_codeword != 0,
                                                                          K, W, rows_per, h->buf_ferr.p, h->buf_cnt.p + 3);
    CKL("count_data");
    k_count_frames<<<std::min(64, (nf + 255) / 256), 256, 0, st>>>(h->buf_ferr.p, h->b.iters, nf, K, N, h->buf_cnt.p);
    CKL("count_frames");
    h->launched(KIND_COUNT);
    total_launches += h->launches + 2;
  }
  CKH(cudaEventRecord(h->ev_sim1, st), "event");
  std::vector<unsigned long long> host(n_cnt);
  CKH(cudaMemcpyAsync(host.data(), h->buf_cnt.p, n_cnt * 8, cudaMemcpyDeviceToHost, st), "D2H counters");
  CKH(cudaStreamSynchronize(st), "sync");
  // report the whole call (all batches) through last_call_stats
  std::swap(h->ev0, h->ev_sim0);
  std::swap(h->ev1, h->ev_sim1);
  h->launches = total_launches;
  h->have_stats = p->n_frames > 0;
  h->stat_frame_iters = (int64_t)host[6];
  acc->frames += host[0]; acc->frame_errors += host[1]; acc->data_bits += host[2]; acc->data_bit_errors += host[3];
  acc->uncoded_bits += host[4]; acc->uncoded_bit_errors += host[5]; acc->sum_iters += host[6];
  acc->decode_failures += host[7];
  for (int i = 0; i < LUTLDPC_ITER_HIST_MAX; ++i) acc->iter_hist[i] += host[8 + i];
  return LUTLDPC_OK;
}

}  // extern "C"
