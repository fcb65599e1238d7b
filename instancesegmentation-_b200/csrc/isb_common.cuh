// Internal shared declarations for libisb.so (not part of the public ABI).
#pragma once
#include <stdlib.h>

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "isb.h"

namespace isb {

// ---------------------------------------------------------------------------------------
// Candidate key layout (64 bit).  Ascending unsigned order of `cand key` is
//   class ascending, score descending, row (anchor) ascending
// and of `det key` is score descending, class ascending, row ascending -- the order the
// reference produces when its two sorts (utils_bbox.py:61, :94) are stable.
// ---------------------------------------------------------------------------------------
constexpr int kRowBits = 20;    // < 1 Mi anchors / rows
constexpr int kClassBits = 12;  // < 4096 foreground classes
constexpr uint64_t kRowMask = (1ull << kRowBits) - 1;
constexpr uint64_t kClassMask = (1ull << kClassBits) - 1;
constexpr uint64_t kPadKey = ~0ull;

__host__ __device__ inline uint32_t float_desc_bits(float f) {
  // monotone map float -> uint32 such that larger floats give SMALLER integers
  uint32_t u;
#ifdef __CUDA_ARCH__
  u = __float_as_uint(f);
#else
  memcpy(&u, &f, 4);
#endif
  if ((u & 0x7fffffffu) > 0x7f800000u) u = 0x7fc00000u;  // any NaN ranks first, before +inf, as in torch.sort(descending=True)
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending orderable
  return ~u;
}
__host__ __device__ inline float float_from_desc_bits(uint32_t d) {
  uint32_t u = ~d;
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}
__host__ __device__ inline uint64_t make_cand_key(uint32_t cls, float score, uint32_t row) {
  return ((uint64_t)cls << (32 + kRowBits)) | ((uint64_t)float_desc_bits(score) << kRowBits) | row;
}
__host__ __device__ inline uint32_t cand_class(uint64_t k) { return (uint32_t)(k >> (32 + kRowBits)); }
__host__ __device__ inline uint32_t cand_row(uint64_t k) { return (uint32_t)(k & kRowMask); }
__host__ __device__ inline uint32_t cand_score_bits(uint64_t k) { return (uint32_t)(k >> kRowBits); }
__host__ __device__ inline uint64_t det_key_from_cand(uint64_t k) {
  return ((uint64_t)cand_score_bits(k) << 32) | ((uint64_t)cand_class(k) << kRowBits) | cand_row(k);
}
__host__ __device__ inline uint32_t det_class(uint64_t k) { return (uint32_t)((k >> kRowBits) & kClassMask); }
__host__ __device__ inline uint32_t det_row(uint64_t k) { return (uint32_t)(k & kRowMask); }
__host__ __device__ inline float det_score(uint64_t k) { return float_from_desc_bits((uint32_t)(k >> 32)); }

// Per-detection record the mask kernel consumes (written by the NMS kernel).
struct DetBounds {
  int32_t xlo, xhi, ylo, yhi;  // half-open integer pixel ranges equivalent to the float crop tests
};

// Per-detection record the mask compose kernel consumes (written by the sigmoid-map kernel):
// crop ranges, offset of the detection's sigmoid map in the per-image scratch and its prototype
// footprint (first column / row, width / height).
struct __align__(16) DetRec {
  int32_t xlo, xhi, ylo, yhi;
  int32_t soff;
  int16_t pa, pya, fw, fh;
  int32_t j;
};
static_assert(sizeof(DetRec) == 32, "DetRec is two 16-byte words");

// ---------------------------------------------------------------------------------------
// fp32 building blocks shared by the fused path and the stage entry points, so both give
// the same bits.  The library is compiled with -fmad=false: every +,-,* below is one
// IEEE-rounded operation, in the reference's order.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float4 decode_box(float4 p, float4 a, float var0, float var1) {
  // utils_bbox.py:17-24
  float cx = a.x + (p.x * var0) * a.z;
  float cy = a.y + (p.y * var0) * a.w;
  float w = a.z * expf(p.z * var1);
  float h = a.w * expf(p.w * var1);
  float x1 = cx - w / 2.0f;
  float y1 = cy - h / 2.0f;
  return make_float4(x1, y1, w + x1, h + y1);
}

// IEEE fp32 division the compiler may not speculate.  A plain `a / b` behind a branch is if-converted and hoisted above the
// test that was meant to skip it, and its slow path (taken for a ZERO numerator -- every disjoint box pair) then runs for
// every pair: 45 of the 80 instructions per pair in the NMS loops (ncu, round 2).
__device__ __forceinline__ float div_rn_guarded(float a, float b) {
  float q;
  asm volatile("div.rn.f32 %0, %1, %2;" : "=f"(q) : "f"(a), "f"(b));
  return q;
}

__device__ __forceinline__ float box_area(float4 b) { return (b.z - b.x) * (b.w - b.y); }

__device__ __forceinline__ float iou_pair(float4 a, float area_a, float4 b, float area_b) {
  // utils_bbox.py:38-47: inter / ((area_a + area_b) - inter), no eps
  float dx = fmaxf(fminf(a.z, b.z) - fmaxf(a.x, b.x), 0.0f);
  float dy = fmaxf(fminf(a.w, b.w) - fmaxf(a.y, b.y), 0.0f);
  float inter = dx * dy;
  return inter / ((area_a + area_b) - inter);
}

__device__ __forceinline__ float4 correct_box(float4 b, float wf, float hf) {
  // utils_bbox.py:147-151 (note :149 uses the already-updated x1y1)
  float x1 = b.x * wf, y1 = b.y * hf, x2 = b.z * wf, y2 = b.w * hf;
  x1 = fminf(x1, x2);
  y1 = fminf(y1, y2);
  x2 = fmaxf(x1, x2);
  y2 = fmaxf(y1, y2);
  x1 = fmaxf(x1, 0.0f);
  y1 = fmaxf(y1, 0.0f);
  x2 = fminf(x2, wf);
  y2 = fminf(y2, hf);
  return make_float4(x1, y1, x2, y2);
}

// Integer pixel range [lo, hi) equivalent to the float tests (col >= v1) && (col < v2) of
// utils_bbox.py:162-165 for integer-valued col in [0, size).
__device__ __forceinline__ void crop_range(float v1, float v2, int size, int& lo, int& hi) {
  float l = ceilf(v1), u = ceilf(v2);
  lo = (l <= 0.0f) ? 0 : (l >= (float)size ? size : (int)l);
  hi = (u <= 0.0f) ? 0 : (u >= (float)size ? size : (int)u);
  if (!(v1 == v1) || !(v2 == v2)) { lo = 0; hi = 0; }  // NaN compares false everywhere
}

__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.0f / (1.0f + expf(-x)); }

// Bilinear source index / weight, aten area_pixel_compute_source_index +
// guard_index_and_lambda with align_corners=False (reached from utils_bbox.py:228).  The source coordinate is ONE
// fused multiply-add, as in torch's own binaries (vectorised CPU kernel and CUDA kernel alike): with two roundings it
// is off by an ulp of ~in_size, i.e. the weights by up to 1e-4 on a 1000-pixel axis (oracle/yolact_np.py).
__device__ __forceinline__ void bilinear_src(int dst, float scale, int in_size, int& i0, int& i1, float& l0, float& l1) {
  float src = fmaf(scale, (float)dst + 0.5f, -0.5f);
  if (src < 0.0f) src = 0.0f;
  int f = (int)floorf(src);
  i0 = f < in_size - 1 ? f : in_size - 1;
  float lam = src - (float)i0;
  lam = fminf(fmaxf(lam, 0.0f), 1.0f);
  l1 = lam;
  l0 = 1.0f - lam;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
}

// prototype footprint of a crop box: first column/row and extent (0 x 0 for an empty box)
__device__ __forceinline__ void det_footprint(const DetBounds& bd, float sy, float sx, int ph, int pw, int& pa, int& pya,
                                              int& fw, int& fh) {
  pa = pya = fw = fh = 0;
  if (bd.xlo >= bd.xhi || bd.ylo >= bd.yhi) return;
  int t0, t1, pb, pyb;
  float f0, f1;
  bilinear_src(bd.xlo, sx, pw, pa, t1, f0, f1);
  bilinear_src(bd.xhi - 1, sx, pw, t0, pb, f0, f1);
  bilinear_src(bd.ylo, sy, ph, pya, t1, f0, f1);
  bilinear_src(bd.yhi - 1, sy, ph, t0, pyb, f0, f1);
  fw = pb - pa + 1;
  fh = pyb - pya + 1;
}

// Eight elements of sigmoid maps, computed by 8 cooperating lanes (q = lane & 7): element i of the group is
// S = sigmoid(proto_vec_i . coef_vec_i) (utils_bbox.py:223 restricted to the prototype pixels a crop box can
// reach).  Every lane passes the 128-byte prototype / coefficient vectors of ITS element (pp, cf); for each of
// the 8 elements the vector indices are broadcast and lane q loads channels 4q..4q+3, so a vector is read as one
// coalesced line.  The 8 x 8 partial sums are then transpose-reduced with a fixed xor butterfly (4, 2, 1):
// lane q returns the complete dot product of element q -- same summation order everywhere this is used.
// The vectors are named by their index in units of 128 bytes (prototype pixel / coefficient row of the image): one
// 32-bit shuffle each instead of a 64-bit pointer (two).
__device__ __forceinline__ float smap_dot8(const float4* proto_img, const float4* coef_img, int pix, int det, int q) {
  float part[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int src = (threadIdx.x & 24) | i;
    const float4* pi = proto_img + (size_t)__shfl_sync(0xffffffffu, pix, src) * (ISB_NUM_COEF / 4);
    const float4* ci = coef_img + (size_t)__shfl_sync(0xffffffffu, det, src) * (ISB_NUM_COEF / 4);
    const float4 a = __ldg(pi + q), c = __ldg(ci + q);
    float acc = a.x * c.x;
    acc = fmaf(a.y, c.y, acc);
    acc = fmaf(a.z, c.z, acc);
    acc = fmaf(a.w, c.w, acc);
    part[i] = acc;
  }
  float r4[4], r2[2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool up = (q & 4) != 0;
    const float recv = __shfl_xor_sync(0xffffffffu, up ? part[j] : part[j + 4], 4);
    r4[j] = (up ? part[j + 4] : part[j]) + recv;
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const bool up = (q & 2) != 0;
    const float recv = __shfl_xor_sync(0xffffffffu, up ? r4[j] : r4[j + 2], 2);
    r2[j] = (up ? r4[j + 2] : r4[j]) + recv;
  }
  const bool up = (q & 1) != 0;
  const float recv = __shfl_xor_sync(0xffffffffu, up ? r2[0] : r2[1], 1);
  return (up ? r2[1] : r2[0]) + recv;
}

// Exclusive prefix sum of one int per thread over a CTA of up to 1024 threads (warp shuffles + one
// shared word per warp); every thread must call it.  Returns the exclusive sum, *total the CTA sum.
__device__ __forceinline__ int cta_exclusive_scan(int v, int* s_warp /* [32] */, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  int off = 0, tot = 0;
  if (nw <= 8) {  // few warps: a short loop of broadcast reads beats seven dependent shuffles
    for (int w = 0; w < nw; ++w) {
      const int c = s_warp[w];
      if (w < warp) off += c;
      tot += c;
    }
  } else {  // every warp scans the (<= 32) warp totals with shuffles: lane w holds warp w's total
    const int wt = lane < nw ? s_warp[lane] : 0;
    int winc = wt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += u;
    }
    tot = __shfl_sync(0xffffffffu, winc, 31);
    off = __shfl_sync(0xffffffffu, winc - wt, warp);
  }
  __syncthreads();
  *total = tot;
  return off + incl - v;
}

// ---------------------------------------------------------------------------------------
// sm_100a async-copy helpers (1-D bulk TMA + mbarrier); SASS: UBLKCP / SYNCS.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {  // release at CTA scope: the thread's earlier accesses are ordered before it
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// the same for data that is read exactly once: the lines enter the L2 as evict-first, so they do not push out what the
// kernels around this one keep there (policy from l2_evict_first_policy(), one createpolicy per thread that issues copies)
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_stream(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}
// shared -> global bulk store (bulk async-group completion).
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// orders async-proxy accesses (bulk stores) against generic-proxy accesses, all state spaces
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Programmatic dependent launch inside ONE entry point: a kernel that calls pdl_enter() first thing may be launched with
// launch_chained(..., chained = true) right behind another kernel of the same call -- it is scheduled while its predecessor
// drains and starts the moment that one has completed (griddepcontrol.wait), which hides the launch latency between
// short kernels.  Launched normally, pdl_enter() does nothing.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__host__ __device__ inline int next_pow2(int v) {
#ifdef __CUDA_ARCH__
  return v <= 1 ? 1 : 1 << (32 - __clz(v - 1));
#else
  int p = 1;
  while (p < v) p <<= 1;
  return p;
#endif
}

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool chained,
                                  Args... args) {
  static const bool no_pdl = getenv("ISB_NO_PDL") != nullptr;
  cudaLaunchConfig_t lc;
  memset(&lc, 0, sizeof(lc));
  lc.gridDim = grid;
  lc.blockDim = block;
  lc.dynamicSmemBytes = smem;
  lc.stream = s;
  cudaLaunchAttribute attr;
  memset(&attr, 0, sizeof(attr));
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = &attr;
  lc.numAttrs = (chained && !no_pdl) ? 1 : 0;
  return cudaLaunchKernelEx(&lc, kernel, static_cast<KArgs>(args)...);
}
#endif

}  // namespace isb

// ---------------------------------------------------------------------------------------
// Handle
// ---------------------------------------------------------------------------------------
// A captured launch sequence of one multi-kernel entry point, replayed while the call's arguments (pointers included)
// and the workspace it touches stay the same.
struct GraphSlot {
  unsigned char key[512];
  int key_len;
  unsigned char pending[512];
  int pending_len;
  cudaGraphExec_t exec;
  int nodes;  // kernel launches inside the graph (for isb_launch_count)
};

struct isb_handle {
  int device;
  int num_sms;
  isb_limits lim;
  int64_t cand_cap_pow2;  // per-image capacity of the key arrays (power of two)
  char err[512];
  int64_t launches;
  size_t ws_bytes;
  // workspace (device)
  GraphSlot graphs[2];   // 0 = isb_rpn_filter, 1 = isb_box_postprocess
  cudaStream_t capture_stream;
  int graphs_off;        // set when a capture failed once: stay eager
  int32_t* fnms_cls;     // wide Fast NMS: class-start table + kept counter
  void* rle_scratch;     // run-length encoding scratch (grown on demand)
  size_t rle_bytes;
  int32_t* cand_count;   // [max_batch]
  int32_t* work_counter; // [1] dynamic work counter of the mask kernel
  uint64_t* cand_keys;   // [max_batch][cand_cap_pow2]
  uint64_t* det_keys;    // [max_batch][cand_cap_pow2]   kept keys (global fallback of the NMS kernel)
  float4* cand_box;      // [max_batch][cand_cap_pow2]   decoded boxes (global fallback)
  float* det_coef;       // [max_batch][max_det][K]
  isb::DetRec* det_rec;        // [max_batch][max_det]
  int32_t* smap_total;         // [max_batch] sigmoid-map elements per image
  float* smap;                 // sigmoid-map scratch of the mask kernels, grown on demand
  size_t smap_floats;
  // device staging for the *_host entry points (allocated on first use)
  void* stage;
  size_t stage_bytes;
  // sub-batch pipelining (isb_yolact_detect_host): image offset into the per-image workspace arrays and
  // the lane whose work counter / stream the next launches use
  int32_t ws_b0, ws_lane;
  int32_t ws_images;       // images the per-image workspace must hold while a sub-batch pipeline runs (0: the call's batch)
  int32_t in_host_pipe;    // set while isb_*_host drives sub-batches through the device entry
  cudaStream_t lane_stream[3];
  cudaEvent_t lane_event[3];
  cudaEvent_t fork_event;
  // fused count exchange (isb_counts_*): this rank's ring, the peers' mappings of theirs
  unsigned long long* cx_ring;        // [cx_depth][cx_world * cx_local] on this device (cudaMalloc, exported through CUDA IPC)
  unsigned long long* cx_peer[8];     // ring of rank r as mapped into this process (cx_peer[cx_rank] == cx_ring)
  int32_t cx_world, cx_rank, cx_local, cx_depth, cx_attached;
  uint32_t cx_epoch;                  // epoch of the last isb_yolact_detect call that published counts
  int32_t* cx_status;                 // device word the wait kernel reports into
  void* host_pipe;         // HostPipe (isb_api.cu): worker pool, pinned landing buffer, copy streams / events of the host delivery
  int32_t host_threads;    // isb_set_host_threads: -1 = default (host_default_threads()), 0 = copy engine only
  int64_t host_stats[8];   // last *_host call: images by copy engine, images by host threads, D2H bytes, H2D bytes, ...
  // Mask R-CNN side: NMS workspace, grown on demand (ensure_nms_ws)
  int64_t nms_cap;         // boxes the arrays below hold
  uint64_t* nms_keys;      // [pow2(nms_cap)]
  uint64_t* nms_keys2;     // [nms_cap]  rank-sorted keys
  int32_t* nms_rank;       // [nms_cap]  rank scratch of the rank sort
  int32_t* nms_slabel;     // [nms_cap]  class per sorted position (per-class NMS without offsets)
  float4* nms_sbox;        // [nms_cap]  boxes in score order (offset by class for batched NMS)
  int32_t* nms_order;      // [nms_cap]  original index per sorted position
  uint64_t* nms_mask;      // [nms_cap][ceil(nms_cap/64)] suppression bits
  uint32_t* nms_scalars;   // [8] max-coordinate (orderable bits), counters
  // generic scratch for the fused RPN / box-head pipelines (ensure_scratch)
  void* scratch;
  size_t scratch_bytes;
  int roi_configured;      // ms_roi_align_kernel's dynamic shared memory limit has been raised
  int fused_grid[2];       // CTAs of the one-launch rpn_filter / box_postprocess kernels (0: not asked yet, -1: unavailable)
  void* prep;              // pre-processing state (preprocess.cu): coefficient tables, intermediate image, value table
  // optional per-stage event ring (isb_profile_begin / isb_profile_read)
  cudaEvent_t* prof_events;  // [prof_cap][5]
  cudaEvent_t prof_mid;      // set while a profiled call is inside launch_masks
  int nms_follows_cand;      // fast_nms_kernel right behind candidates_kernel on the stream (fused path): programmatic dependent
  int smap_follows_nms;      // same for the sigmoid-map kernel right behind fast_nms_kernel
  int compose_follows_smap;  // set by the fused path around its compose launch: the sigmoid-map kernel of the same call is the
                             // stream's previous operation, so the compose kernel may be launched as its programmatic dependent
  int32_t prof_cap, prof_used;
};

#define ISB_FAIL(h, code, ...)                          \
  do {                                                  \
    snprintf((h)->err, sizeof((h)->err), __VA_ARGS__);  \
    return (code);                                      \
  } while (0)

#define ISB_CUDA(h, expr)                                                                    \
  do {                                                                                       \
    cudaError_t e_ = (expr);                                                                 \
    if (e_ != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

// Makes the handle's device current for the rest of the calling scope and puts the caller's device back on the way out:
// an entry point must not leave the runtime's current device changed behind a caller that drives several GPUs from one process.
struct IsbDeviceGuard {
  int prev = -1;
  bool changed = false;
  cudaError_t set(int dev) {
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return e;
    if (prev == dev) return cudaSuccess;
    e = cudaSetDevice(dev);
    changed = e == cudaSuccess;
    return e;
  }
  ~IsbDeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};
#define ISB_SET_DEVICE(h)    \
  IsbDeviceGuard dev_guard_; \
  ISB_CUDA(h, dev_guard_.set((h)->device))

#define ISB_LAUNCH_CHECK(h, name)                                                                 \
  do {                                                                                            \
    cudaError_t e_ = cudaGetLastError();                                                          \
    if (e_ != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch %s: %s", name, cudaGetErrorString(e_)); \
    (h)->launches++;                                                                              \
  } while (0)

// Kernel launchers implemented in the other translation units.
namespace isb {

struct CandParams {
  const float* scores;   // [B, rows, ncols]
  int32_t rows, ncols;   // ncols includes skipped leading columns
  int32_t col0;          // first column that is a foreground class (1 for conf with background, 0 for stage call)
  int32_t use_threshold; // 0: every entry is a candidate
  float threshold;
  int32_t batch;
  int32_t logits;        // 1: scores are raw logits, softmax over the ncols columns is fused into the filter
};
int launch_candidates(isb_handle* h, const CandParams& p, cudaStream_t s);

struct NmsParams {
  int32_t batch;
  // box source: either raw regressions + anchors (decode) or ready boxes
  const float4* pred_box;  // [B, rows] (decode mode) or nullptr
  const float4* anchors;   // [rows]
  const float4* boxes;     // [B, rows] ready boxes (stage mode) or nullptr
  const float* coef;       // [B, rows, K]
  int32_t rows;
  float nms_iou;
  int32_t top_k, max_det;
  int32_t correct;         // 1: scale/clamp to pixels and emit crop bounds (fused path)
  int32_t out_h, out_w;
  int32_t proto_h, proto_w; // fused path: prototype size (footprints of the sigmoid maps)
  float* out_box;          // [B, max_det, 4]
  float* out_score;        // [B, max_det]
  int64_t* out_class;      // [B, max_det]
  float* out_coef;         // [B, max_det, K]
  int32_t* out_count;      // [B]
  // fused count exchange (counts_exchange.cu): with n_peers > 0 the kernel also stores epoch << 32 | count of image b into
  // peer_ring[r][peer_index0 + b] for every rank r (peer-mapped device memory)
  unsigned long long* peer_ring[8];
  int32_t n_peers;
  int32_t peer_index0;
  uint32_t peer_epoch;
};
int launch_fast_nms(isb_handle* h, const NmsParams& p, cudaStream_t s);
// one image, n candidates known on the host, spread over the GPU (n > fast_nms_smem_candidates())
int launch_fast_nms_wide(isb_handle* h, const NmsParams& p, int n, cudaStream_t s);
int fast_nms_smem_candidates();

struct MaskParams {
  int32_t batch;
  const float* proto;        // [B, Ph, Pw, K]
  const float* coef;         // [B, max_det_stride, K]
  const int32_t* counts;     // [B]   device; or nullptr when m_fixed >= 0
  int32_t m_fixed;
  int32_t det_stride;        // rows between images in coef / bounds
  int32_t m_cap;             // upper bound of the detections per image in THIS call (sizes shared memory and the sigmoid-map
                             // scratch); 0 = det_stride
  int32_t proto_h, proto_w, out_h, out_w;
  int64_t* masks_arg;        // [B, H, W]
  uint8_t* masks;            // B slots
  int64_t mask_slot_bytes;
  // optional per-pixel class map (yolact.py:167-169): class_ids[arg] + 1 where the arg-max mask is set, else 0
  uint8_t* class_map;        // [B, H, W] or nullptr
  const int64_t* class_ids;  // [B, class_stride], needed with class_map
  int32_t class_stride;      // rows between images in class_ids (the CALLER's array: in->max_det, not the workspace stride)
  int32_t map_only;          // 1: masks / masks_arg are not produced at all
  // optional bit-packed copy of the masks (transport format of the *_host path): slot b holds the image's H*W*m_b
  // mask bits, bit i = element i of the dense [H,W,m_b] stream, in little-endian 32-bit words
  uint8_t* packed;           // B slots of packed_slot_bytes, or nullptr
  int64_t packed_slot_bytes; // >= 4 * ceil(H*W*max_det / 32), multiple of 16
  int32_t skip_dense;        // 1 with `packed`: the dense masks are not written (masks may be nullptr); masks_arg still is
  // fused count exchange, deferred from the NMS kernel (NmsParams below): the sigmoid-map kernel stores epoch << 32 | count
  // of every image into peer_ring[r][peer_index0 + b] at its START, so the NVLink stores drain under its own 38 us instead
  // of holding up the end of the NMS kernel (a kernel is not complete before its remote stores are: +4 us at 8 GPUs)
  unsigned long long* peer_ring[8];
  int32_t n_peers;
  int32_t peer_index0;
  uint32_t peer_epoch;
};
// which: bit 0 = sigmoid maps (K3a), bit 1 = compose (K3b); the two may be launched for different image ranges / streams
int launch_masks(isb_handle* h, const MaskParams& p, cudaStream_t s, int which = 3);
bool masks_tc_smap_forced();  // ISB_TC_SMAP measurement switch: the sigmoid maps come from the tensor-core kernel
// grows the sigmoid-map scratch for `images` images of at most m_cap detections; returns floats per image in *s_cap
int ensure_mask_scratch(isb_handle* h, int proto_h, int proto_w, int m_cap, int64_t images, int64_t* s_cap);

struct RoiLevels {
  const float* feat[8];
  int h[8], w[8];
  float scale[8];
  int num_levels;
  int k_min, k_max;
  float canonical_scale, canonical_level, eps;
};

int launch_box_decode(isb_handle* h, const float* rel, const float* boxes, int64_t n, int k, const float w[4], float clip,
                      float* out, cudaStream_t s);
int launch_mask_select_sigmoid(isb_handle* h, const float* logits, const int64_t* labels, int m, int num_classes, int plane,
                               float* out, cudaStream_t s);
int launch_resize_boxes(isb_handle* h, const float* in, int64_t m, float rw, float rh, float* out, cudaStream_t s);
int launch_paste_masks(isb_handle* h, const float* masks, const float* boxes, int m, int S, int pad, int im_h, int im_w,
                       float* out, cudaStream_t s);
int launch_ms_roi_align(isb_handle* h, const RoiLevels& L, const float* boxes, const int32_t* batch_idx, int num_rois,
                        int channels, int oh, int ow, int sr, float* out, cudaStream_t s);
int configure_nms(isb_handle* h);
int configure_mask_gemm_tc(isb_handle* h);
int launch_mask_logits_tc(isb_handle* h, const float* proto, const float* coef, int64_t npix, int m, float* out,
                          cudaStream_t s);
// batched tcgen05 product (mask_gemm_tc.cu): mode 0 logits, 1 sigmoid -> dense out [B, npix, m]; mode 2 sigmoid scattered
// into the packed footprint maps (rec / smap / s_cap), m read per image from counts
struct TcBatchArgs {
  const float* proto; const float* coef; int64_t coef_rows; const int32_t* counts; int32_t m_fixed;
  int32_t batch, npix, proto_w, mode; float* out; const DetRec* rec; int32_t rec_stride; float* smap; int64_t s_cap;
};
int launch_mask_logits_batch(isb_handle* h, const TcBatchArgs& a, cudaStream_t s);
constexpr int kTcBatchMaxDet = 128;

int launch_counts_wait(isb_handle* h, const unsigned long long* slot, int n, unsigned epoch, int32_t* counts, int64_t* offsets,
                       int32_t* status, long long timeout_ns, cudaStream_t s);
int launch_sort_u64(isb_handle* h, uint64_t* keys, int n_pow2, cudaStream_t s);
int launch_rank_sort_u64(isb_handle* h, const uint64_t* keys, int n_max, const int32_t* n_dev, int32_t* rank, uint64_t* out,
                         cudaStream_t s, bool rank_is_zero = false);
constexpr int kRankSortMax = 32768;  // above this the O(n^2) rank sort loses to the bitonic network
int ensure_nms_ws(isb_handle* h, int64_t n);
int ensure_scratch(isb_handle* h, size_t bytes);

// Greedy NMS (torchvision semantics) over boxes/scores on the device.
struct NmsCall {
  const float4* boxes;     // [n]
  const float* scores;     // [n]
  const int64_t* idxs64;   // batched NMS categories as int64 (boxes.py:38) or nullptr
  const int32_t* idxs32;   // ... or as int32 (internal pipelines) or nullptr
  const int32_t* labels32; // per-class NMS on the UNSHIFTED boxes (utils_bbox.py:102-140): classes < 4096, n < 2^20
  int32_t n_max;           // upper bound of n known on the host (sizes grids and workspace)
  const int32_t* n_dev;    // actual n on the device, or nullptr when n == n_max
  double iou_threshold;
  int32_t max_keep;        // stop after this many kept boxes (<= 0: no limit)
  int64_t* keep;           // [n_max] kept original indices, score-descending
  int32_t* count;          // [1]
};
int launch_nms(isb_handle* h, const NmsCall& c, cudaStream_t s);

}  // namespace isb
