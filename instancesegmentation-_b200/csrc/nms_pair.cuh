// One pair of torchvision's greedy NMS (boxes.py:35 -> torch.ops.torchvision.nms, CPU kernel semantics), shared by the
// suppression-matrix kernels of nms.cu and the fused Mask R-CNN pipelines of rcnn_pipeline.cu.
#pragma once
#include "isb_common.cuh"

namespace isb {

// boxes.max() of boxes.py:78 propagates NaN (torch.max); fmaxf would drop it
__device__ __forceinline__ float nan_max(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b); }

// Does the higher-ranked box c suppress the row box a?   inter / ((area_a + area_c) - inter) > thr, all in float32.
// torchvision's CPU kernel: std::max(i, j) = (i < j) ? j : i and std::min(i, j) = (j < i) ? j : i with i = the
// higher-ranked box (c), j = the row box (a) -- this decides what a NaN coordinate does; max(0, d) = (0 < d) ? d : 0
// turns a NaN extent into 0.  Disjoint pairs (the vast majority) skip the division: 0 / union is +-0 for every finite
// non-zero union, and the slow path of the IEEE division for a zero numerator is the most expensive thing in the loop.
__device__ __forceinline__ bool nms_pair_hit(const float4 a, const float area_a, const float4 c, const float area_c, const float thr,
                                             const bool zero_hits /* 0.0f > thr */) {
  const float xx1 = (c.x < a.x) ? a.x : c.x, yy1 = (c.y < a.y) ? a.y : c.y;
  const float xx2 = (a.z < c.z) ? a.z : c.z, yy2 = (a.w < c.w) ? a.w : c.w;
  const float dw = xx2 - xx1, dh = yy2 - yy1;
  const float w = (0.0f < dw) ? dw : 0.0f;
  const float hgt = (0.0f < dh) ? dh : 0.0f;
  const float inter = w * hgt;
  const float uni = (area_a + area_c) - inter;
  if (inter == 0.0f && uni != 0.0f && fabsf(uni) <= 3.402823466e38f) return zero_hits;
  return div_rn_guarded(inter, uni) > thr;
}

}  // namespace isb
