// Allocation-free entry point: lutldpc_syndrome_check_device_noalloc (include/lutldpc.h).
//
// Some callers cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  This variant of the syndrome check (syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469) keeps the
// scratch the check needs -- the bit-per-frame rows hard[nvar][W], the per-frame flag words and the live-column
// lists -- in a file-scope __device__ array instead of the handle's grow-only cudaMalloc buffers.  The array has
// a fixed size (LUTLDPC_NOALLOC_SCRATCH_BYTES); every GPU that loads the library holds it until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the device address is fetched once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernels as a plain pointer.
// All calls on a device share the array and it keeps the last call's contents: what has to start at zero is
// cleared by the call itself, and calls that overlap on the GPU (different streams) race.
#include <cstdio>
#include <mutex>

#include "decoder.hpp"

using namespace lutldpc;

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[LUTLDPC_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_ptr[kMaxDevices];  // device address of g_noalloc_scratch, per device (nullptr: not fetched)
std::mutex g_scratch_mutex;

// once per device, with that device current
cudaError_t scratch_for_device(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[device];
  return cudaSuccess;
}

inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Layout {
  size_t hard, flags, cols, iters, total;
};
// scratch of one call: W words per row (8 frames per word, W % 32 == 0)
Layout layout_for(int nvar, int W) {
  Layout l;
  l.hard = 0;
  l.flags = up256((size_t)nvar * W);
  l.cols = l.flags + up256(((size_t)4 * W + 32) * 4);
  l.iters = l.cols + up256(((size_t)W + W / 4) * 4);
  l.total = l.iters + up256((size_t)8 * W * 4);
  return l;
}

// ok[f] = 1 iff no check of frame f failed (flag at bit 0 of nibble f & 7 of word f >> 3)
__global__ void k_flags_to_ok(const uint32_t* __restrict__ syn_fail, int F, uint8_t* __restrict__ ok) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < F) ok[f] = (uint8_t)(((syn_fail[f >> 3] >> (4 * (f & 7))) & 1u) ^ 1u);
}

}  // namespace

extern "C" int lutldpc_noalloc_max_frames(const lutldpc_decoder* h) {
  if (!h) return LUTLDPC_ERR_ARG;
  int lo = 0, hi = 1 << 20;  // largest n with layout(32 * ceil(n / 256)).total <= the static scratch
  while (lo < hi) {
    const int mid = lo + (hi - lo + 1) / 2;
    if (layout_for(h->m.nvar, ((mid + 255) / 256) * 32).total <= (size_t)LUTLDPC_NOALLOC_SCRATCH_BYTES) lo = mid;
    else hi = mid - 1;
  }
  return lo;
}

extern "C" int lutldpc_syndrome_check_device_noalloc(lutldpc_decoder* h, int n_frames, const uint8_t* d_bits,
                                                     uint8_t* d_ok) {
  if (!h) return LUTLDPC_ERR_ARG;
  if (n_frames < 0) return h->fail(LUTLDPC_ERR_ARG, "n_frames < 0");
  if (n_frames == 0) return LUTLDPC_OK;
  if (!d_bits || !d_ok) return h->fail(LUTLDPC_ERR_ARG, "NULL buffer");
  const int W = ((n_frames + 255) / 256) * 32;
  const Layout l = layout_for(h->m.nvar, W);
  if (l.total > (size_t)LUTLDPC_NOALLOC_SCRATCH_BYTES) {
    char buf[200];
    snprintf(buf, sizeof buf,
             "lutldpc_syndrome_check_device_noalloc: %d frames need %zu bytes of scratch, the static array holds %u "
             "(at most %d frames per call for this code)",
             n_frames, l.total, (unsigned)LUTLDPC_NOALLOC_SCRATCH_BYTES, lutldpc_noalloc_max_frames(h));
    h->err = buf;
    return LUTLDPC_ERR_ARG;
  }
  cudaError_t e = cudaSetDevice(h->device);
  if (e != cudaSuccess) return h->cuda_fail(e, "cudaSetDevice");
  unsigned char* s = nullptr;
  e = scratch_for_device(h->device, &s);
  if (e != cudaSuccess) return h->cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
  BatchBuffers b;
  b.hard = s + l.hard;
  uint32_t* flags = reinterpret_cast<uint32_t*>(s + l.flags);
  b.done = flags; b.vn_fail = flags + W; b.cn_fail = flags + 2 * W; b.syn_fail = flags + 3 * W;
  b.n_active = flags + 4 * W;
  b.cols = reinterpret_cast<uint32_t*>(s + l.cols);
  b.grp4 = b.cols + W;
  b.iters = reinterpret_cast<int32_t*>(s + l.iters);
  b.host_cols = W; b.host_grp4 = W / 4;
  b.W = W; b.F = n_frames;
  cudaStream_t st = h->stream;
  // the array keeps the previous call's contents: flags and lists are rebuilt, hard is rewritten for every (v, w)
  if ((e = launch_reset_flags(b, st)) != cudaSuccess) return h->cuda_fail(e, "reset_flags");
  if ((e = launch_pack_bits(d_bits, h->g, b, st)) != cudaSuccess) return h->cuda_fail(e, "pack_bits");
  if ((e = launch_syndrome(h->g, b, st)) != cudaSuccess) return h->cuda_fail(e, "syndrome");
  k_flags_to_ok<<<(n_frames + 255) / 256, 256, 0, st>>>(b.syn_fail, n_frames, d_ok);
  if ((e = cudaGetLastError()) != cudaSuccess) return h->cuda_fail(e, "flags_to_ok");
  return LUTLDPC_OK;
}
