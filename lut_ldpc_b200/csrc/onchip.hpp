// Launch interface of the on-chip regime (onchip.cu): one CTA = a group of C word columns (8 frames each) with their
// whole decoder state in shared memory for all iterations.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace lutldpc {

constexpr int kOnchipMaxSmem = 227 * 1024 - 2048;  // dynamic shared memory a CTA may ask for (static part: < 2 KiB)
constexpr int kOnchipMaxClasses = 8;
constexpr int kOnchipThreads = 256;
constexpr int kOnchipMaxCols = 8;  // columns per CTA

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
  // work items of a phase: up to 32 nodes of one class, heaviest class first; a unit = item x column
  const int4* vn_items;  // (class, first node of the class list, count, 0)
  const int4* cn_items;  // (check class, first check, count, 0)
  int n_vn_items, n_cn_items;
  // program images: slot s at slot_images + s * slot_stride: the images of the classes back to back
  // (DevProgHeader | ops | outs | tables, model.hpp), the decision trees in slot dec_slot
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
  unsigned int* queue;   // next column group to hand out
  // dynamic shared memory: C column blocks of col_bytes (msgs | cha at off_cha | hard at off_hard), then the
  // program images of one slot at off_tab, then the interpreter's value slots [n_slots][slot_threads] at off_slots
  int C, slot_threads;
  uint32_t col_bytes, off_cha, off_hard, off_tab, off_slots, smem_bytes;
};

cudaError_t configure_onchip();
int onchip_ctas_per_sm(int max_fast_dv, size_t smem_bytes);
cudaError_t launch_onchip(const OnchipParams& P, int max_fast_dv, int n_ctas, cudaStream_t st);

}  // namespace lutldpc
