// Micro-benchmark (not product code): what limits the in-place row-gather pattern of the CN pass?
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#define CK(x)                                                         \
  do {                                                                \
    cudaError_t e = (x);                                              \
    if (e != cudaSuccess) {                                           \
      printf("err %s line %d\n", cudaGetErrorString(e), __LINE__);    \
      exit(1);                                                        \
    }                                                                 \
  } while (0)

template <int VEC> struct WV;
template <> struct WV<1> { typedef uint32_t T; };
template <> struct WV<2> { typedef uint2 T; };
template <> struct WV<4> { typedef uint4 T; };

// mode bit0: fake ALU work, bit1: no store
template <int DC, int VEC>
__global__ void __launch_bounds__(128) probe(const int* __restrict__ idx, int nchk, int W, int stride, uint32_t* msgs,
                                             int mode) {
  typedef typename WV<VEC>::T V;
  union U { V v; uint32_t u[VEC]; };
  int c = blockIdx.x;
  int w = (blockIdx.y * blockDim.x + threadIdx.x) * VEC;
  if (w >= W) return;
  int row[DC];
#pragma unroll
  for (int k = 0; k < DC; ++k) row[k] = idx[c * DC + k];
  U x[DC];
#pragma unroll
  for (int k = 0; k < DC; ++k) x[k].v = *reinterpret_cast<const V*>(msgs + (size_t)row[k] * stride + w);
  uint32_t acc[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) acc[j] = 0;
  if (mode & 1) {
#pragma unroll
    for (int k = 0; k < DC; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        uint32_t t = x[k].u[j];
        for (int r = 0; r < 5; ++r) {
          t = ((t | 0x88888888u) - acc[j]) & 0x88888888u;
          t = t - (t >> 3);
          acc[j] = (acc[j] & t) | (x[k].u[j] & ~t);
        }
      }
  }
  if (!(mode & 2)) {
#pragma unroll
    for (int k = 0; k < DC; ++k) {
      U o;
#pragma unroll
      for (int j = 0; j < VEC; ++j) o.u[j] = x[k].u[j] ^ acc[j];
      *reinterpret_cast<V*>(msgs + (size_t)row[k] * stride + w) = o.v;
    }
  } else {
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < DC; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) s ^= x[k].u[j] ^ acc[j];
    if (s == 0x12345678u) msgs[0] = s;
  }
}

template <int VEC>
float run(const int* d_idx, int nchk, int W, int stride, uint32_t* d, int mode) {
  dim3 grid(nchk, (W + 128 * VEC - 1) / (128 * VEC));
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  for (int i = 0; i < 2; ++i) probe<8, VEC><<<grid, 128>>>(d_idx, nchk, W, stride, d, mode);
  cudaEventRecord(a);
  for (int i = 0; i < 5; ++i) probe<8, VEC><<<grid, 128>>>(d_idx, nchk, W, stride, d, mode);
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms / 5;
}

int main() {
  int nchk = 30000, DC = 8;
  int E = nchk * DC;
  for (int W : {512, 1024, 2048})
    for (int pad : {0, 32}) {
      int stride = W + pad;
      uint32_t* d;
      CK(cudaMalloc(&d, (size_t)E * stride * 4));
      CK(cudaMemset(d, 1, (size_t)E * stride * 4));
      std::vector<int> perm(E);
      for (int i = 0; i < E; ++i) perm[i] = i;
      std::mt19937 g(1);
      int* d_idx;
      CK(cudaMalloc(&d_idx, E * 4));
      for (int shuffle = 0; shuffle < 2; ++shuffle) {
        if (shuffle) std::shuffle(perm.begin(), perm.end(), g);
        CK(cudaMemcpy(d_idx, perm.data(), E * 4, cudaMemcpyHostToDevice));
        double gb = 2.0 * E * W * 4 / 1e9;
        for (int mode : {0, 1, 2}) {
          float t1 = run<1>(d_idx, nchk, W, stride, d, mode), t2 = run<2>(d_idx, nchk, W, stride, d, mode),
                t4 = run<4>(d_idx, nchk, W, stride, d, mode);
          double f = (mode & 2) ? 0.5 : 1.0;
          printf("W=%d pad=%d %s mode=%d(%s%s): vec1 %.3f ms %.0f GB/s | vec2 %.3f ms %.0f GB/s | vec4 %.3f ms %.0f GB/s\n",
                 W, pad, shuffle ? "gather" : "seq   ", mode, (mode & 1) ? "alu" : "copy", (mode & 2) ? ",nostore" : "", t1,
                 f * gb / t1 * 1e3, t2, f * gb / t2 * 1e3, t4, f * gb / t4 * 1e3);
        }
      }
      cudaFree(d);
      cudaFree(d_idx);
    }
  return 0;
}
