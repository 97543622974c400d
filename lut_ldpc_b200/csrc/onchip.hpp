// Launch interface of the on-chip regime (onchip.cu): one CTA = one word column (8 frames) with its whole decoder
// state in shared memory for all iterations.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace lutldpc {

constexpr int kOnchipMaxSmem = 227 * 1024 - 1024;  // dynamic shared memory a CTA may ask for (static part: < 1 KiB)
constexpr int kOnchipMaxClasses = 8;

struct OnchipClass {  // variable nodes of one degree slot
  const int2* pairs;   // (VN index, first edge id)
  int n, dv;
  int fast;            // 1: the compile-time leave-one-out program (vn_fast.cuh) applies in every iteration slot
  uint32_t img_off;    // of this class's program image inside the image of an iteration slot
  uint32_t dec_off;    // ... inside the image of the decision slot
};
struct OnchipCn {       // checks of one degree (min-sum)
  const int32_t* rows;  // [n][dc] edge ids
  int n, dc;
};

struct OnchipParams {
  OnchipClass vc[kOnchipMaxClasses];
  OnchipCn cc[kOnchipMaxClasses];
  int n_vc, n_cc;
  // program images: slot s at slot_images + s * slot_stride, slot_bytes each (16-byte multiple): the images of the
  // classes back to back (DevProgHeader | ops | outs | tables, model.hpp), the decision trees in slot dec_slot
  const uint8_t* slot_images;
  uint32_t slot_stride, slot_bytes, dec_bytes;  // bytes to stage for an iteration slot / for the decision slot
  const int32_t* iter_slot;  // [max_iters] var_tree_idx_iter
  const int32_t* iter_q;     // [max_iters] code width of Nq_Msg(ii)
  int max_iters, psc, pisc, dec_slot, q_cha;
  int nvar, nchk, num_edges, W;
  const int32_t* vn_edge0;
  const int32_t* cn_start;
  const int32_t* chk_equ_idx;
  const uint32_t* cha;   // [nvar][W] channel codes
  const uint32_t* msg0;  // [nvar][W] initial message code of every node (its edges all start with it)
  uint8_t* hard;         // [nvar][W] out
  int32_t* iters;        // [8W] out
  uint32_t* done;        // [W] in: padding frames; out: all ones
  unsigned int* queue;   // next column to hand out
  uint32_t off_cha, off_hard, off_tab, off_slots, smem_bytes;  // dynamic shared memory layout (bytes)
};

cudaError_t configure_onchip();
int onchip_ctas_per_sm(int max_fast_dv, int threads, size_t smem_bytes);
cudaError_t launch_onchip(const OnchipParams& P, int max_fast_dv, int threads, int n_ctas, cudaStream_t st);

}  // namespace lutldpc
