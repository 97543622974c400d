// GF(2) linear algebra on the device for the set-up steps that take the reference minutes on large codes:
//   rank of H                LDPC_Code_LUT::decoder_parameterization, GF2mat::row_rank   src/LDPC_Code_LUT.cpp:493-499
//                            (dense elimination on one core: 528 s for the N=64800 code)
//   systematic generator     LDPC_Generator_Systematic::construct                        itpp/itpp/comm/ldpc.cpp:894-999
//   encode                   LDPC_Generator_Systematic::encode, codeword = [u, G*u]      itpp/itpp/comm/ldpc.cpp:1027-1039
// H lives bit-packed in HBM ([nchk][ceil(nvar/64)] u64: 262 MB for N=64800); Gauss-Jordan elimination runs column by
// column with two small launches per column (pivot search by atomicMin, row XOR over the whole matrix) and no host
// synchronisation.  Pivot columns are taken in natural column order -- construct()'s `natural_ordering` variant; the
// reference's default draws a random order from IT++'s global RNG, any order gives a valid generator of the same code.
#include <algorithm>
#include <cstdio>
#include <vector>

#include "gf2.hpp"

namespace lutldpc {

namespace {

__global__ void k_gf2_fill(const int32_t* __restrict__ row_start, const int32_t* __restrict__ row_cols, int nchk,
                           size_t Wd, unsigned long long* __restrict__ A) {
  const int c = blockIdx.x;
  for (int k = row_start[c] + threadIdx.x; k < row_start[c + 1]; k += blockDim.x) {
    const int v = row_cols[k];
    atomicXor(&A[(size_t)c * Wd + (v >> 6)], 1ull << (v & 63));
  }
}

// state[0] = pivot row of the running column (INT_MAX: none), state[1] = number of pivots so far.
// colbit[i] = bit (i, j): a copy that stays put while the elimination rewrites the column.
__global__ void k_gf2_find(const unsigned long long* __restrict__ A, size_t Wd, int nchk, int j,
                           const uint8_t* __restrict__ used, uint8_t* __restrict__ colbit, int* __restrict__ state) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nchk) return;
  const uint8_t b = (uint8_t)((A[(size_t)i * Wd + (j >> 6)] >> (j & 63)) & 1ull);
  colbit[i] = b;
  if (b && !used[i]) atomicMin(&state[0], i);
}

// Gauss-Jordan step for column j: every other row that has the bit gets the pivot row XORed in.  A CTA walks rows
// blockIdx.x, blockIdx.x + gridDim.x, ...; its threads cover the words of a row.  The pivot row itself is left alone
// (everyone reads it); it is marked used by the bookkeeping kernel that follows, which also resets the search state.
__global__ void __launch_bounds__(256) k_gf2_eliminate(unsigned long long* __restrict__ A, size_t Wd, int nchk,
                                                       const uint8_t* __restrict__ colbit, const int* __restrict__ state) {
  const int p = state[0];
  if (p == 0x7FFFFFFF) return;  // dependent column
  const unsigned long long* rp = A + (size_t)p * Wd;
  for (int i = blockIdx.x; i < nchk; i += gridDim.x) {
    if (i == p || !colbit[i]) continue;
    unsigned long long* ri = A + (size_t)i * Wd;
    for (size_t w = threadIdx.x; w < Wd; w += blockDim.x) ri[w] ^= rp[w];
  }
}

__global__ void k_gf2_book(int j, uint8_t* __restrict__ used, int* __restrict__ state, int32_t* __restrict__ piv_row,
                           int32_t* __restrict__ piv_col) {
  const int p = state[0];
  if (p != 0x7FFFFFFF) {
    used[p] = 1;
    piv_row[state[1]] = p;
    piv_col[state[1]] = j;
    state[1] += 1;
  }
  state[0] = 0x7FFFFFFF;
}

// P[t][k] = A[piv_row[t]][info_col[k]]: bit k of row t, 32 columns per word
__global__ void k_gf2_gather_P(const unsigned long long* __restrict__ A, size_t Wd, const int32_t* __restrict__ piv_row,
                               const int32_t* __restrict__ info_col, int K, int Wp, uint32_t* __restrict__ P) {
  const int t = blockIdx.y;
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= Wp) return;
  const unsigned long long* r = A + (size_t)piv_row[t] * Wd;
  uint32_t out = 0;
  for (int b = 0; b < 32; ++b) {
    const int k = 32 * w + b;
    if (k >= K) break;
    const int c = info_col[k];
    out |= (uint32_t)((r[c >> 6] >> (c & 63)) & 1ull) << b;
  }
  P[(size_t)t * Wp + w] = out;
}

// parity rows of a batch: cw[K + t][.] = XOR over the set bits k of P[t] of cw[k][.]; rows are bytes of 8 frames, a
// thread owns 4 of them (32 frames).  CTA = kEncRows parity rows x 128 thread columns: the 32 information rows of one
// 32-bit piece of P are staged in shared memory once for all kEncRows parity rows (the information block is re-read
// rows / kEncRows times in all), and combined four at a time ("four Russians"): the 16 XOR combinations of every group
// of 4 staged rows are built once per piece, so a parity row costs one shared-memory load and one XOR per 4 columns of P
// whatever their bits are.
constexpr int kEncRows = 64;
__global__ void __launch_bounds__(128) k_encode_bits(const uint32_t* __restrict__ P, int Wp, int rows, int K, int W4,
                                                     uint32_t* __restrict__ cw /* [nvar][W4] words of 4 bytes */) {
  __shared__ uint32_t comb[4][16][128];  // [group of 4 information rows][subset][thread]: 32 KiB
  __shared__ uint32_t pw[kEncRows];
  const int tx = threadIdx.x, w = blockIdx.x * 128 + tx;
  const int t0 = blockIdx.y * kEncRows;
  uint32_t acc[kEncRows];
#pragma unroll
  for (int r = 0; r < kEncRows; ++r) acc[r] = 0;
  for (int kh = 0; kh < 2 * Wp; ++kh) {  // 16 columns of P (half a word) per stage
    const int kw = kh >> 1, sh = 16 * (kh & 1);
    __syncthreads();
    if (tx < kEncRows) pw[tx] = t0 + tx < rows ? (P[(size_t)(t0 + tx) * Wp + kw] >> sh) & 0xFFFFu : 0u;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint32_t u[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int k = 16 * kh + 4 * g + b;
        u[b] = (k < K && w < W4) ? cw[(size_t)k * W4 + w] : 0u;
      }
      uint32_t c[16];
      c[0] = 0;
#pragma unroll
      for (int m = 1; m < 16; ++m) {  // subset m = subset without its lowest row ^ that row
        const int low = m & -m, b = low == 1 ? 0 : low == 2 ? 1 : low == 4 ? 2 : 3;
        c[m] = c[m ^ low] ^ u[b];
      }
#pragma unroll
      for (int m = 0; m < 16; ++m) comb[g][m][tx] = c[m];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kEncRows; ++r) {
      const uint32_t m = pw[r];  // uniform over the CTA
#pragma unroll
      for (int g = 0; g < 4; ++g) acc[r] ^= comb[g][(m >> (4 * g)) & 15][tx];
    }
  }
  if (w < W4)
#pragma unroll
    for (int r = 0; r < kEncRows; ++r)
      if (t0 + r < rows) cw[(size_t)(K + t0 + r) * W4 + w] = acc[r];
}

}  // namespace

#define GCK(call)                          \
  do {                                     \
    cudaError_t e_ = (call);               \
    if (e_ != cudaSuccess) { err = std::string("CUDA error in gf2: ") + cudaGetErrorString(e_); return false; } \
  } while (0)

bool gf2_systematic_device(const ParityRows& H, Gf2Result& out, bool want_generator, cudaStream_t st, std::string& err) {
  const int M = H.nchk, N = H.nvar;
  const size_t Wd = ((size_t)N + 63) / 64;
  std::vector<int32_t> row_start(M + 1, 0), row_cols;
  for (int c = 0; c < M; ++c) {
    row_start[c + 1] = row_start[c] + (int32_t)H.rows[c].size();
    row_cols.insert(row_cols.end(), H.rows[c].begin(), H.rows[c].end());
  }
  int32_t *d_start = nullptr, *d_cols = nullptr, *d_piv_row = nullptr, *d_piv_col = nullptr, *d_info = nullptr;
  unsigned long long* A = nullptr;
  uint8_t *d_used = nullptr, *d_colbit = nullptr;
  int* d_state = nullptr;
  auto cleanup = [&]() {
    cudaFree(d_start); cudaFree(d_cols); cudaFree(d_piv_row); cudaFree(d_piv_col); cudaFree(d_info);
    cudaFree(A); cudaFree(d_used); cudaFree(d_colbit); cudaFree(d_state);
  };
  struct Guard { decltype(cleanup)& f; ~Guard() { f(); } } guard{cleanup};
  GCK(cudaMalloc((void**)&d_start, (M + 1) * sizeof(int32_t)));
  GCK(cudaMalloc((void**)&d_cols, std::max<size_t>(1, row_cols.size()) * sizeof(int32_t)));
  GCK(cudaMalloc((void**)&A, (size_t)M * Wd * 8));
  GCK(cudaMalloc((void**)&d_used, M));
  GCK(cudaMalloc((void**)&d_colbit, M));
  GCK(cudaMalloc((void**)&d_state, 2 * sizeof(int)));
  GCK(cudaMalloc((void**)&d_piv_row, M * sizeof(int32_t)));
  GCK(cudaMalloc((void**)&d_piv_col, M * sizeof(int32_t)));
  GCK(cudaMemcpyAsync(d_start, row_start.data(), (M + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  GCK(cudaMemcpyAsync(d_cols, row_cols.data(), row_cols.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  GCK(cudaMemsetAsync(A, 0, (size_t)M * Wd * 8, st));
  GCK(cudaMemsetAsync(d_used, 0, M, st));
  const int init[2] = {0x7FFFFFFF, 0};
  GCK(cudaMemcpyAsync(d_state, init, sizeof init, cudaMemcpyHostToDevice, st));
  k_gf2_fill<<<M, 64, 0, st>>>(d_start, d_cols, M, Wd, A);
  GCK(cudaGetLastError());
  const int ge = std::min(M, 148 * 8);
  for (int j = 0; j < N; ++j) {
    k_gf2_find<<<(M + 255) / 256, 256, 0, st>>>(A, Wd, M, j, d_used, d_colbit, d_state);
    k_gf2_eliminate<<<ge, 256, 0, st>>>(A, Wd, M, d_colbit, d_state);
    k_gf2_book<<<1, 1, 0, st>>>(j, d_used, d_state, d_piv_row, d_piv_col);
  }
  GCK(cudaGetLastError());
  int state[2];
  GCK(cudaMemcpyAsync(state, d_state, sizeof state, cudaMemcpyDeviceToHost, st));
  GCK(cudaStreamSynchronize(st));
  const int rank = state[1];
  out = Gf2Result();
  out.rank = rank;
  std::vector<int32_t> piv_col(rank);
  GCK(cudaMemcpy(piv_col.data(), d_piv_col, rank * sizeof(int32_t), cudaMemcpyDeviceToHost));
  out.piv_col = piv_col;
  // new column order as construct() leaves it: information columns first (natural order), then the pivot columns
  // in the order they were found (actual_ordering, ldpc.cpp:944-966)
  const int K = N - rank;
  std::vector<uint8_t> is_piv(N, 0);
  for (int c : piv_col) is_piv[c] = 1;
  out.col_perm.assign(N, -1);
  std::vector<int32_t> info_col;
  for (int v = 0; v < N; ++v)
    if (!is_piv[v]) { out.col_perm[v] = (int32_t)info_col.size(); info_col.push_back(v); }
  for (int t = 0; t < rank; ++t) out.col_perm[piv_col[t]] = K + t;
  if (want_generator && rank > 0 && K > 0) {
    const int Wp = (K + 31) / 32;
    GCK(cudaMalloc((void**)&d_info, K * sizeof(int32_t)));
    GCK(cudaMemcpyAsync(d_info, info_col.data(), K * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    uint32_t* dP = nullptr;
    GCK(cudaMalloc((void**)&dP, (size_t)rank * Wp * 4));
    k_gf2_gather_P<<<dim3((Wp + 127) / 128, rank), 128, 0, st>>>(A, Wd, d_piv_row, d_info, K, Wp, dP);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { cudaFree(dP); err = std::string("CUDA error in gf2 gather: ") + cudaGetErrorString(e); return false; }
    out.d_P = dP;
    out.P_words_per_row = Wp;
  }
  return true;
}

cudaError_t launch_encode_bits(const uint32_t* d_P, int Wp, int rows, int K, int W, uint8_t* d_cw, cudaStream_t st) {
  const int W4 = W / 4;  // W is a multiple of 32 bytes
  dim3 grid((W4 + 127) / 128, (rows + kEncRows - 1) / kEncRows);
  k_encode_bits<<<grid, 128, 0, st>>>(d_P, Wp, rows, K, W4, reinterpret_cast<uint32_t*>(d_cw));
  return cudaGetLastError();
}

}  // namespace lutldpc
