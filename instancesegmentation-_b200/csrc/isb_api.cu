// C-ABI entry points of libisb.so (see include/isb.h): argument validation, workspace
// management and kernel sequencing.  No torch, no exceptions, plain pointers and sizes.
#include <chrono>
#include <deque>
#include <new>
#include <thread>
#include <vector>

#include "host_expand.h"
#include "isb_common.cuh"

namespace isb {
int configure_fast_nms(isb_handle* h);
int configure_masks(isb_handle* h);
int configure_rcnn_ops(isb_handle* h);
void destroy_preprocess(isb_handle* h);
int configure_candidates(isb_handle* h);
int launch_decode_boxes(isb_handle*, const float*, const float*, int64_t, float, float, float*, cudaStream_t);
int launch_jaccard(isb_handle*, const float*, const float*, int, int, int, int, float*, cudaStream_t);
int launch_correct_boxes(isb_handle*, const float*, int64_t, int, int, float*, cudaStream_t);
int launch_crop(isb_handle*, const float*, const float*, int, int, int, float*, cudaStream_t);
int launch_det_layout(isb_handle*, const float*, int, int, int, int, int, DetRec*, int32_t*, cudaStream_t);
int launch_mask_logits(isb_handle*, const float*, const float*, int64_t, int, float*, cudaStream_t);
int launch_row_argmax(isb_handle*, const float*, int, int, int32_t*, cudaStream_t);
int launch_trad_gather(isb_handle*, const int64_t*, const int32_t*, int, const float*, const float*, const int32_t*, const float*,
                       float*, float*, int64_t*, float*, int32_t*, cudaStream_t);
}  // namespace isb

using namespace isb;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool aligned4(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 3) == 0; }

extern "C" int isb_version(void) { return ISB_VERSION; }

extern "C" const char* isb_last_error(const isb_handle* h) { return h ? h->err : "null handle"; }
extern "C" size_t isb_workspace_bytes(const isb_handle* h) { return h ? h->ws_bytes : 0; }
extern "C" int64_t isb_launch_count(const isb_handle* h) { return h ? h->launches : 0; }

static void free_profile(isb_handle* h) {
  if (h->prof_events) {
    for (int i = 0; i < h->prof_cap * 5; ++i) cudaEventDestroy(h->prof_events[i]);
    delete[] h->prof_events;
  }
  h->prof_events = nullptr;
  h->prof_cap = h->prof_used = 0;
}

extern "C" int isb_profile_begin(isb_handle* h, int32_t max_calls) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (max_calls < 0 || max_calls > (1 << 16)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "profile_begin: bad max_calls");
  ISB_SET_DEVICE(h);
  free_profile(h);
  if (max_calls == 0) return ISB_OK;
  h->prof_events = new (std::nothrow) cudaEvent_t[(size_t)max_calls * 5];
  if (!h->prof_events) ISB_FAIL(h, ISB_ERR_CUDA, "profile_begin: out of host memory");
  for (int i = 0; i < max_calls * 5; ++i) ISB_CUDA(h, cudaEventCreate(&h->prof_events[i]));
  h->prof_cap = max_calls;
  return ISB_OK;
}

extern "C" int isb_profile_read(isb_handle* h, double stage_ms[4], int32_t* calls) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!stage_ms || !calls) ISB_FAIL(h, ISB_ERR_BAD_ARG, "profile_read: null output");
  stage_ms[0] = stage_ms[1] = stage_ms[2] = stage_ms[3] = 0.0;
  *calls = h->prof_used;
  if (h->prof_used == 0) return ISB_OK;
  ISB_SET_DEVICE(h);
  ISB_CUDA(h, cudaEventSynchronize(h->prof_events[(size_t)(h->prof_used - 1) * 5 + 4]));
  for (int c = 0; c < h->prof_used; ++c)
    for (int st = 0; st < 4; ++st) {
      float ms = 0.f;
      ISB_CUDA(h, cudaEventElapsedTime(&ms, h->prof_events[(size_t)c * 5 + st], h->prof_events[(size_t)c * 5 + st + 1]));
      stage_ms[st] += ms;
    }
  h->prof_used = 0;
  return ISB_OK;
}

static void destroy_host_pipe(isb_handle* h);
static void cx_release(isb_handle* h);

static void free_ws(isb_handle* h) {
  free_profile(h);
  destroy_host_pipe(h);
  cx_release(h);
  for (int l = 0; l < 3; ++l) {
    if (h->lane_stream[l]) cudaStreamDestroy(h->lane_stream[l]);
    if (h->lane_event[l]) cudaEventDestroy(h->lane_event[l]);
  }
  if (h->fork_event) cudaEventDestroy(h->fork_event);
  cudaFree(h->rle_scratch);
  cudaFree(h->fnms_cls);
  for (int g = 0; g < 2; ++g)
    if (h->graphs[g].exec) cudaGraphExecDestroy(h->graphs[g].exec);
  if (h->capture_stream) cudaStreamDestroy(h->capture_stream);
  cudaFree(h->cand_count);
  cudaFree(h->work_counter);
  cudaFree(h->cand_keys);
  cudaFree(h->det_keys);
  cudaFree(h->cand_box);
  cudaFree(h->det_coef);
  cudaFree(h->det_rec);
  cudaFree(h->smap_total);
  cudaFree(h->smap);
  cudaFree(h->stage);
  cudaFree(h->nms_keys);
  cudaFree(h->nms_keys2);
  cudaFree(h->nms_rank);
  cudaFree(h->nms_slabel);
  cudaFree(h->nms_sbox);
  cudaFree(h->nms_order);
  cudaFree(h->nms_mask);
  cudaFree(h->nms_scalars);
  cudaFree(h->scratch);
}

extern "C" void isb_destroy(isb_handle* h) {
  if (!h) return;
  IsbDeviceGuard guard;
  guard.set(h->device);
  destroy_preprocess(h);
  free_ws(h);
  delete h;
}

extern "C" int isb_create(int device, const isb_limits* limits, isb_handle** out) {
  if (!out) return ISB_ERR_BAD_ARG;
  *out = nullptr;
  isb_handle* h = new (std::nothrow) isb_handle();
  if (!h) return ISB_ERR_CUDA;
  memset(h, 0, sizeof(*h));
  *out = h;  // returned even on failure so that isb_last_error() can be read; caller destroys it
  h->device = device;
  h->host_threads = -1;
  isb_limits L;
  memset(&L, 0, sizeof(L));
  if (limits) L = *limits;
  if (L.max_batch <= 0) L.max_batch = 64;
  if (L.max_anchors <= 0) L.max_anchors = 19248;
  if (L.max_classes <= 0) L.max_classes = 80;
  if (L.max_det <= 0) L.max_det = 100;
  if (L.max_top_k <= 0) L.max_top_k = 200;
  if (L.cand_capacity <= 0) L.cand_capacity = 16384;
  if (L.max_det > ISB_MAX_DET_LIMIT) ISB_FAIL(h, ISB_ERR_LIMIT, "max_det %d > %d", L.max_det, ISB_MAX_DET_LIMIT);
  if (L.max_anchors > (1 << kRowBits)) ISB_FAIL(h, ISB_ERR_LIMIT, "max_anchors %d > 2^%d", L.max_anchors, kRowBits);
  if (L.max_classes >= (1 << kClassBits)) ISB_FAIL(h, ISB_ERR_LIMIT, "max_classes %d >= 2^%d", L.max_classes, kClassBits);
  if (L.cand_capacity > (1ll << 30)) ISB_FAIL(h, ISB_ERR_LIMIT, "cand_capacity too large");
  h->lim = L;
  h->cand_cap_pow2 = next_pow2((int)L.cand_capacity) < 64 ? 64 : next_pow2((int)L.cand_capacity);
  IsbDeviceGuard guard;
  ISB_CUDA(h, guard.set(device));
  cudaDeviceProp prop;
  ISB_CUDA(h, cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) ISB_FAIL(h, ISB_ERR_CUDA, "libisb is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
  h->num_sms = prop.multiProcessorCount;
  const size_t B = (size_t)L.max_batch, cap = (size_t)h->cand_cap_pow2;
  size_t total = 0;
#define WS_ALLOC(field, bytes)                                   \
  do {                                                           \
    ISB_CUDA(h, cudaMalloc((void**)&h->field, (bytes)));         \
    total += (bytes);                                            \
  } while (0)
  WS_ALLOC(cand_count, B * sizeof(int32_t));
  WS_ALLOC(work_counter, 256);
  WS_ALLOC(cand_keys, B * cap * sizeof(uint64_t));
  WS_ALLOC(det_keys, B * cap * sizeof(uint64_t));
  WS_ALLOC(cand_box, B * cap * sizeof(float4));
  WS_ALLOC(det_coef, B * (size_t)L.max_det * ISB_NUM_COEF * sizeof(float));
  WS_ALLOC(det_rec, B * (size_t)L.max_det * sizeof(DetRec));
  WS_ALLOC(smap_total, B * sizeof(int32_t));
#undef WS_ALLOC
  h->ws_bytes = total;
  int rc;
  if ((rc = configure_candidates(h)) != ISB_OK) return rc;
  if ((rc = configure_fast_nms(h)) != ISB_OK) return rc;
  if ((rc = configure_masks(h)) != ISB_OK) return rc;
  if ((rc = configure_rcnn_ops(h)) != ISB_OK) return rc;
  if ((rc = configure_nms(h)) != ISB_OK) return rc;
  if ((rc = configure_mask_gemm_tc(h)) != ISB_OK) return rc;
  h->err[0] = 0;
  return ISB_OK;
}

// ----------------------------------------------------------------------------------- YOLACT
// Both public output structs (full masks / class map only) funnel into this one.
struct YolactOutAll {
  float* boxes;
  float* scores;
  int64_t* class_ids;
  int64_t* masks_arg;
  uint8_t* masks;
  int64_t mask_slot_bytes;
  int32_t* counts;
  uint8_t* class_map;
  int map_only;
  uint8_t* packed;            // host pipeline only: bit-packed copy of the masks (device), or nullptr
  int64_t packed_slot_bytes;
};

static YolactOutAll widen(const isb_yolact_out* o) {
  YolactOutAll a;
  memset(&a, 0, sizeof(a));
  if (o) {
    a.boxes = o->boxes; a.scores = o->scores; a.class_ids = o->class_ids; a.masks_arg = o->masks_arg;
    a.masks = o->masks; a.mask_slot_bytes = o->mask_slot_bytes; a.counts = o->counts;
  }
  return a;
}

static YolactOutAll widen(const isb_yolact_map_out* o) {
  YolactOutAll a;
  memset(&a, 0, sizeof(a));
  if (o) {
    a.boxes = o->boxes; a.scores = o->scores; a.class_ids = o->class_ids; a.counts = o->counts;
    a.class_map = o->class_map;
  }
  a.map_only = 1;
  return a;
}

static int check_yolact(isb_handle* h, const isb_yolact_in* in, const YolactOutAll* out, bool have_out) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!in || !have_out) ISB_FAIL(h, ISB_ERR_BAD_ARG, "null in/out struct");
  if (!in->box || !in->conf || !in->coef || !in->proto || !in->anchors) ISB_FAIL(h, ISB_ERR_BAD_ARG, "null input tensor");
  if (!out->boxes || !out->scores || !out->class_ids || !out->counts) ISB_FAIL(h, ISB_ERR_BAD_ARG, "null output tensor");
  if (out->map_only) {
    if (!out->class_map) ISB_FAIL(h, ISB_ERR_BAD_ARG, "null class_map");
    if (in->num_classes > 255) ISB_FAIL(h, ISB_ERR_LIMIT, "class map is uint8: at most 254 foreground classes, got %d", in->num_classes - 1);
  } else if (!out->masks_arg || !out->masks) {
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "null output tensor");
  }
  if (in->batch <= 0 || in->num_anchors <= 0 || in->num_classes < 2 || in->proto_h <= 0 || in->proto_w <= 0 ||
      in->out_h <= 0 || in->out_w <= 0 || in->top_k <= 0 || in->max_det <= 0)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "non-positive dimension");
  if (in->num_coef != ISB_NUM_COEF) ISB_FAIL(h, ISB_ERR_BAD_ARG, "num_coef must be %d, got %d", ISB_NUM_COEF, in->num_coef);
  if (in->batch > h->lim.max_batch) ISB_FAIL(h, ISB_ERR_LIMIT, "batch %d > max_batch %d", in->batch, h->lim.max_batch);
  if (in->num_anchors > h->lim.max_anchors) ISB_FAIL(h, ISB_ERR_LIMIT, "num_anchors %d > max_anchors %d", in->num_anchors, h->lim.max_anchors);
  if (in->num_classes - 1 > h->lim.max_classes) ISB_FAIL(h, ISB_ERR_LIMIT, "classes %d > max_classes %d", in->num_classes - 1, h->lim.max_classes);
  if (in->max_det > h->lim.max_det) ISB_FAIL(h, ISB_ERR_LIMIT, "max_det %d > limit %d", in->max_det, h->lim.max_det);
  if (in->top_k > h->lim.max_top_k) ISB_FAIL(h, ISB_ERR_LIMIT, "top_k %d > limit %d", in->top_k, h->lim.max_top_k);
  if (!out->map_only && (out->mask_slot_bytes < (int64_t)in->out_h * in->out_w * in->max_det || (out->mask_slot_bytes & 15)))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_slot_bytes must be >= H*W*max_det and a multiple of 16");
  return ISB_OK;
}

// One image range [b0, b0 + nb) of a call through K1 (score filter), K2 (sort + Fast NMS) and K3a (sigmoid maps) on stream
// `st`; the workspace is addressed by absolute image index (base + b0).
struct DetectRange {
  isb_handle* h;
  const isb_yolact_in* in;
  const YolactOutAll* out;
  int base;          // h->ws_b0 of the call (host pipeline: the sub-batch's first image)
  bool publish;      // fused count exchange: this call's counts also go to every rank's ring
  cudaEvent_t* ev;   // profiling events of a sequential call, or nullptr
};

static MaskParams mask_params(const DetectRange& d, int b0, int nb) {
  const isb_yolact_in* in = d.in;
  const YolactOutAll* out = d.out;
  const size_t HW = (size_t)in->out_h * in->out_w;
  MaskParams mp;
  memset(&mp, 0, sizeof(mp));
  mp.batch = nb;
  mp.proto = in->proto + (size_t)b0 * in->proto_h * in->proto_w * ISB_NUM_COEF;
  mp.coef = d.h->det_coef + (size_t)(d.base + b0) * d.h->lim.max_det * ISB_NUM_COEF;
  mp.counts = out->counts + b0;
  mp.m_fixed = -1;
  mp.det_stride = d.h->lim.max_det;
  mp.m_cap = in->max_det;
  mp.proto_h = in->proto_h;
  mp.proto_w = in->proto_w;
  mp.out_h = in->out_h;
  mp.out_w = in->out_w;
  mp.masks_arg = out->masks_arg ? out->masks_arg + (size_t)b0 * HW : nullptr;
  mp.masks = out->masks ? out->masks + (size_t)b0 * out->mask_slot_bytes : nullptr;
  mp.mask_slot_bytes = out->mask_slot_bytes;
  mp.class_map = out->class_map ? out->class_map + (size_t)b0 * HW : nullptr;
  mp.class_ids = out->class_ids + (size_t)b0 * in->max_det;
  mp.class_stride = in->max_det;  // fnms_emit wrote the caller's [B, in->max_det] array
  mp.map_only = out->map_only;
  mp.packed = out->packed ? out->packed + (size_t)b0 * out->packed_slot_bytes : nullptr;
  mp.packed_slot_bytes = out->packed_slot_bytes;
  return mp;
}

static int detect_k1(const DetectRange& d, int b0, int nb, cudaStream_t st) {
  isb_handle* h = d.h;
  const isb_yolact_in* in = d.in;
  CandParams cp;
  cp.scores = in->conf + (size_t)b0 * in->num_anchors * in->num_classes;
  cp.rows = in->num_anchors;
  cp.ncols = in->num_classes;
  cp.col0 = 1;  // utils_bbox.py:193 drops the background column
  cp.use_threshold = 1;
  cp.threshold = in->confidence;
  cp.batch = nb;
  cp.logits = in->conf_is_logits ? 1 : 0;
  if (cp.logits && !aligned16(cp.scores)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "conf must be 16-byte aligned when conf_is_logits is set");
  h->ws_b0 = d.base + b0;
  const int rc = launch_candidates(h, cp, st);
  h->ws_b0 = d.base;
  return rc;
}

static int detect_k2_k3a(const DetectRange& d, int b0, int nb, cudaStream_t st) {
  isb_handle* h = d.h;
  const isb_yolact_in* in = d.in;
  const YolactOutAll* out = d.out;
  NmsParams np;
  memset(&np, 0, sizeof(np));
  np.batch = nb;
  np.pred_box = reinterpret_cast<const float4*>(in->box) + (size_t)b0 * in->num_anchors;
  np.anchors = reinterpret_cast<const float4*>(in->anchors);
  np.boxes = nullptr;
  np.coef = in->coef + (size_t)b0 * in->num_anchors * ISB_NUM_COEF;
  np.rows = in->num_anchors;
  np.nms_iou = in->nms_iou;
  np.top_k = in->top_k;
  np.max_det = in->max_det;
  np.correct = 1;
  np.out_h = in->out_h;
  np.out_w = in->out_w;
  np.proto_h = in->proto_h;
  np.proto_w = in->proto_w;
  np.out_box = out->boxes + (size_t)b0 * in->max_det * 4;
  np.out_score = out->scores + (size_t)b0 * in->max_det;
  np.out_class = out->class_ids + (size_t)b0 * in->max_det;
  np.out_coef = nullptr;  // coefficients of kept detections go to the workspace
  np.out_count = out->counts + b0;
  // the counts reach the peers from the sigmoid-map kernel that follows (MaskParams); from this kernel only when the
  // measurement switch replaces that kernel
  const bool publish_here = d.publish && masks_tc_smap_forced();
  if (publish_here) {
    np.n_peers = h->cx_world;
    for (int r = 0; r < h->cx_world; ++r) np.peer_ring[r] = h->cx_peer[r];
    np.peer_index0 = (int32_t)((h->cx_epoch % (uint32_t)h->cx_depth) * (uint32_t)(h->cx_world * h->cx_local) + h->cx_rank * h->cx_local + b0);
    np.peer_epoch = h->cx_epoch;
  }
  h->ws_b0 = d.base + b0;
  h->nms_follows_cand = d.ev ? 0 : 1;  // (a profiled call records an event in between)
  int rc = launch_fast_nms(h, np, st);
  h->nms_follows_cand = 0;
  if (rc == ISB_OK) {
    if (d.ev) {
      const cudaError_t e = cudaEventRecord(d.ev[2], st);
      if (e != cudaSuccess) { h->ws_b0 = d.base; ISB_FAIL(h, ISB_ERR_CUDA, "cudaEventRecord: %s", cudaGetErrorString(e)); }
    }
    MaskParams mp = mask_params(d, b0, nb);
    if (d.publish && !publish_here) {
      mp.n_peers = h->cx_world;
      for (int r = 0; r < h->cx_world; ++r) mp.peer_ring[r] = h->cx_peer[r];
      mp.peer_index0 = (int32_t)((h->cx_epoch % (uint32_t)h->cx_depth) * (uint32_t)(h->cx_world * h->cx_local) + h->cx_rank * h->cx_local + b0);
      mp.peer_epoch = h->cx_epoch;
    }
    h->smap_follows_nms = d.ev ? 0 : 1;  // (a profiled call records an event in between)
    rc = launch_masks(h, mp, st, 1);
    h->smap_follows_nms = 0;
  }
  h->ws_b0 = d.base;
  return rc;
}

static int yolact_detect_device(isb_handle* h, const isb_yolact_in* in, const YolactOutAll* out, cudaStream_t s) {
  if (!aligned16(in->box) || !aligned16(in->coef) || !aligned16(in->proto) || !aligned16(in->anchors) ||
      !aligned4(in->conf) || !aligned16(out->boxes) || (!out->map_only && !aligned16(out->masks)))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "box/coef/proto/anchors/boxes/masks must be 16-byte aligned, conf 4-byte aligned");
  int rc;
  cudaEvent_t* ev = (h->prof_events && h->prof_used < h->prof_cap) ? h->prof_events + (size_t)h->prof_used * 5 : nullptr;
  DetectRange d;
  d.h = h; d.in = in; d.out = out; d.base = h->ws_b0; d.ev = ev;
  d.publish = h->cx_attached && in->batch == h->cx_local && !h->in_host_pipe;
  if (d.publish) h->cx_epoch += 1;
  const int B = in->batch;
  // the sigmoid-map scratch must hold the whole call before any range is launched (no reallocation under running kernels)
  const int32_t ws_images_saved = h->ws_images;
  if (h->ws_images < d.base + B) h->ws_images = d.base + B;
  // The four kernels run back to back on the caller's stream.  (Measured and rejected in round 2: running the first half's
  // K2 + K3a on a second stream under the second half's K1 -- K2 is latency-bound, ~37 us whatever the image count, so every
  // extra K2 launch that is not fully hidden costs more than the overlap saves: 0.4765 ms with the two-half overlap against
  // 0.4669 ms for the plain sequence, profiles/r02_split_overlap.json.)
  if (ev) ISB_CUDA(h, cudaEventRecord(ev[0], s));
  rc = detect_k1(d, 0, B, s);
  if (rc == ISB_OK && ev) ISB_CUDA(h, cudaEventRecord(ev[1], s));
  h->prof_mid = ev ? ev[3] : nullptr;  // recorded between the sigmoid-map and the compose kernel
  if (rc == ISB_OK) rc = detect_k2_k3a(d, 0, B, s);
  h->prof_mid = nullptr;
  if (rc == ISB_OK) {
    const MaskParams mp = mask_params(d, 0, B);
    h->compose_follows_smap = ev ? 0 : 1;  // (a profiled call records an event in between)
    rc = launch_masks(h, mp, s, 2);
    h->compose_follows_smap = 0;
  }
  h->ws_images = ws_images_saved;
  if (rc != ISB_OK) return rc;
  if (ev) {
    ISB_CUDA(h, cudaEventRecord(ev[4], s));
    h->prof_used++;
  }
  return ISB_OK;
}

extern "C" int isb_yolact_detect(isb_handle* h, const isb_yolact_in* in, const isb_yolact_out* out, isb_stream stream) {
  const YolactOutAll all = widen(out);
  int rc = check_yolact(h, in, &all, out != nullptr);
  if (rc != ISB_OK) return rc;
  ISB_SET_DEVICE(h);
  return yolact_detect_device(h, in, &all, (cudaStream_t)stream);
}

extern "C" int isb_yolact_class_map(isb_handle* h, const isb_yolact_in* in, const isb_yolact_map_out* out, isb_stream stream) {
  const YolactOutAll all = widen(out);
  int rc = check_yolact(h, in, &all, out != nullptr);
  if (rc != ISB_OK) return rc;
  ISB_SET_DEVICE(h);
  return yolact_detect_device(h, in, &all, (cudaStream_t)stream);
}

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

static int yolact_host(isb_handle* h, const isb_yolact_in* in, const YolactOutAll* out, isb_stream stream);

extern "C" int isb_yolact_detect_host(isb_handle* h, const isb_yolact_in* in, const isb_yolact_out* out,
                                      isb_stream stream) {
  const YolactOutAll all = widen(out);
  int rc = check_yolact(h, in, &all, out != nullptr);
  if (rc != ISB_OK) return rc;
  return yolact_host(h, in, &all, stream);
}

extern "C" int isb_yolact_class_map_host(isb_handle* h, const isb_yolact_in* in, const isb_yolact_map_out* out,
                                         isb_stream stream) {
  const YolactOutAll all = widen(out);
  int rc = check_yolact(h, in, &all, out != nullptr);
  if (rc != ISB_OK) return rc;
  return yolact_host(h, in, &all, stream);
}

// Host delivery state of the *_host entry points (created on first use, owned by the handle).
//
// Streams by ROLE, so that no transfer ever waits behind one of another kind:
//   up     host -> device only (one stream: two alternating ones were slower): every chunk's inputs, enqueued back to back at call start (the upload never stalls);
//   comp   kernels only: chunk c waits for upload event c, records kernel event c;
//   cnt    the counts and per-detection records of a chunk (a few KB): what tells the host which images are ready;
//   fast   the packed mask bits of the images the host threads take (3.8 MB per image);
//   dense  bulk device -> host traffic in pieces of kPieceBytes with at most kDensePieces outstanding: the arg-max maps
//          and the dense masks of the images the copy engine takes -- so the other two streams' copies interleave with it
//          on the device-to-host link instead of queueing behind 30 MB.
// Every device-to-host copy is issued by the host only once its source is READY: nothing blocked ever sits in a copy queue.
// (Round 1 and the first version of this pipeline put a chunk's upload, kernels and downloads on one of three lane
// streams: the next upload of a lane sat behind that lane's mask downloads, and the small copies shared hardware queues
// with blocked work -- measured: packed copies issued at 1.7 ms completed at 25.8 ms, chunks 5-7 became ready at 25 ms.)
struct HostPipe {
  static constexpr int kDensePieces = 2;       // dense pieces in flight
  static constexpr size_t kPieceBytes = 4u << 20;
  static constexpr int kPackedSlots = 6;       // images being copied as bits / expanded by the worker pool
  isb::HostPool* pool = nullptr;
  int pool_threads = 0;
  uint8_t* land = nullptr;  // pinned landing buffer of the packed bits
  size_t land_bytes = 0;
  cudaStream_t up = nullptr, comp = nullptr, cnt = nullptr, fast = nullptr, dense = nullptr;
  std::vector<cudaEvent_t> up_ev, k_ev, cnt_ev;  // per chunk: inputs uploaded, kernels done, counts on the host
  cudaEvent_t piece_ev[kDensePieces] = {};
  bool piece_busy[kDensePieces] = {};
  cudaEvent_t join_ev = nullptr;
  struct PackedSlot {
    cudaEvent_t ev = nullptr;
    int img = -1;
    int state = 0;  // 0 free, 1 bits on their way to the landing buffer, 2 being expanded
    isb::ExpandJob job;
  } packed[kPackedSlots];
};

static void destroy_host_pipe(isb_handle* h) {
  HostPipe* hp = static_cast<HostPipe*>(h->host_pipe);
  if (!hp) return;
  if (hp->pool) isb::host_pool_destroy(hp->pool);
  if (hp->land) cudaFreeHost(hp->land);
  for (cudaStream_t st : {hp->up, hp->comp, hp->cnt, hp->fast, hp->dense})
    if (st) cudaStreamDestroy(st);
  for (auto* v : {&hp->up_ev, &hp->k_ev, &hp->cnt_ev})
    for (cudaEvent_t e : *v) cudaEventDestroy(e);
  for (int i = 0; i < HostPipe::kDensePieces; ++i)
    if (hp->piece_ev[i]) cudaEventDestroy(hp->piece_ev[i]);
  if (hp->join_ev) cudaEventDestroy(hp->join_ev);
  for (int i = 0; i < HostPipe::kPackedSlots; ++i)
    if (hp->packed[i].ev) cudaEventDestroy(hp->packed[i].ev);
  delete hp;
  h->host_pipe = nullptr;
}

extern "C" int isb_set_host_threads(isb_handle* h, int32_t threads) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (threads > 256) ISB_FAIL(h, ISB_ERR_BAD_ARG, "set_host_threads: at most 256 threads");
  h->host_threads = threads < 0 ? -1 : threads;
  return ISB_OK;
}

extern "C" int isb_host_stats(const isb_handle* h, int64_t stats[8]) {
  if (!h || !stats) return ISB_ERR_BAD_ARG;
  memcpy(stats, h->host_stats, sizeof(h->host_stats));
  return ISB_OK;
}

static int get_host_pipe(isb_handle* h, int nchunks, size_t land_bytes, HostPipe** out) {
  HostPipe* hp = static_cast<HostPipe*>(h->host_pipe);
  if (!hp) {
    hp = new (std::nothrow) HostPipe();
    if (!hp) ISB_FAIL(h, ISB_ERR_CUDA, "host pipeline: out of host memory");
    h->host_pipe = hp;
    for (cudaStream_t* st : {&hp->up, &hp->comp, &hp->cnt, &hp->fast, &hp->dense}) ISB_CUDA(h, cudaStreamCreateWithFlags(st, cudaStreamNonBlocking));
    for (int i = 0; i < HostPipe::kDensePieces; ++i) ISB_CUDA(h, cudaEventCreateWithFlags(&hp->piece_ev[i], cudaEventDisableTiming));
    for (int i = 0; i < HostPipe::kPackedSlots; ++i) ISB_CUDA(h, cudaEventCreateWithFlags(&hp->packed[i].ev, cudaEventDisableTiming));
    ISB_CUDA(h, cudaEventCreateWithFlags(&hp->join_ev, cudaEventDisableTiming));
  }
  const int want = h->host_threads < 0 ? isb::host_default_threads() : h->host_threads;
  if (want != hp->pool_threads) {
    if (hp->pool) isb::host_pool_destroy(hp->pool);
    hp->pool = want > 0 ? isb::host_pool_create(want) : nullptr;
    hp->pool_threads = want;
  }
  for (auto* v : {&hp->up_ev, &hp->k_ev, &hp->cnt_ev})
    while ((int)v->size() < nchunks) {
      cudaEvent_t e;
      ISB_CUDA(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      v->push_back(e);
    }
  if (hp->pool && land_bytes > hp->land_bytes) {
    if (hp->land) cudaFreeHost(hp->land);
    hp->land = nullptr;
    hp->land_bytes = 0;
    ISB_CUDA(h, cudaHostAlloc((void**)&hp->land, land_bytes, cudaHostAllocDefault));
    hp->land_bytes = land_bytes;
  }
  *out = hp;
  return ISB_OK;
}

// cudaEventQuery as a tri-state: 1 done, 0 not yet, < 0 error (message set)
static int event_done(isb_handle* h, cudaEvent_t e) {
  const cudaError_t r = cudaEventQuery(e);
  if (r == cudaSuccess) return 1;
  if (r == cudaErrorNotReady) return 0;
  ISB_FAIL(h, ISB_ERR_CUDA, "cudaEventQuery: %s", cudaGetErrorString(r));
}

static int yolact_host(isb_handle* h, const isb_yolact_in* in, const YolactOutAll* out, isb_stream stream) {
  int rc;
  const bool map_only = out->map_only != 0;
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t B = in->batch, A = in->num_anchors, C = in->num_classes, K = ISB_NUM_COEF;
  const size_t HW = (size_t)in->out_h * in->out_w, D = in->max_det;
  // sub-batches: small enough that the pipeline fills quickly (the first masks leave after one chunk's upload) and
  // drains quickly (the last chunk's delivery is all that remains once the upload ends)
  const int chunk = (int)(B >= 32 ? 4 : (B >= 8 ? 2 : 1));
  const int nchunks = (int)((B + chunk - 1) / chunk);
  // packed transport: one slot per image, 64-byte multiple, 16 bytes of slack for the expansion's look-ahead word
  const size_t pslot = map_only ? 0 : ((4 * ((HW * D + 31) / 32) + 16 + 63) & ~(size_t)63);
  HostPipe* hp = nullptr;
  // the landing buffer has one slot per in-flight packed image (not per image of the batch): a few MB that are written by the
  // copy engine and read by the worker threads straight away, i.e. out of the last-level cache rather than DRAM
  if ((rc = get_host_pipe(h, nchunks, (size_t)HostPipe::kPackedSlots * pslot, &hp)) != ISB_OK) return rc;
  const bool use_pool = !map_only && hp->pool != nullptr;
  const size_t n_box = B * A * 4 * 4, n_conf = B * A * C * 4, n_coef = B * A * K * 4,
               n_proto = B * (size_t)in->proto_h * in->proto_w * K * 4, n_anc = A * 4 * 4;
  const size_t o_box = B * D * 16, o_score = B * D * 4, o_cls = B * D * 8, o_arg = map_only ? 0 : B * HW * 8,
               o_mask = map_only ? 0 : B * (size_t)out->mask_slot_bytes, o_cnt = B * 4, o_map = map_only ? B * HW : 0,
               o_pack = use_pool ? B * pslot : 0;
  const size_t need = up256(n_box) + up256(n_conf) + up256(n_coef) + up256(n_proto) + up256(n_anc) + up256(o_box) +
                      up256(o_score) + up256(o_cls) + up256(o_arg) + up256(o_mask) + up256(o_cnt) + up256(o_map) + up256(o_pack);
  if (need > h->stage_bytes) {
    ISB_CUDA(h, cudaStreamSynchronize(s));
    cudaFree(h->stage);
    h->stage = nullptr;
    h->stage_bytes = 0;
    ISB_CUDA(h, cudaMalloc(&h->stage, need));
    h->stage_bytes = need;
  }
  char* p = static_cast<char*>(h->stage);
  auto take = [&](size_t bytes) { char* r = p; p += up256(bytes); return r; };
  isb_yolact_in din = *in;
  YolactOutAll dout = *out;
  din.box = (const float*)take(n_box);
  din.conf = (const float*)take(n_conf);
  din.coef = (const float*)take(n_coef);
  din.proto = (const float*)take(n_proto);
  din.anchors = (const float*)take(n_anc);
  dout.boxes = (float*)take(o_box);
  dout.scores = (float*)take(o_score);
  dout.class_ids = (int64_t*)take(o_cls);
  dout.masks_arg = (int64_t*)take(o_arg);
  dout.masks = (uint8_t*)take(o_mask);
  dout.counts = (int32_t*)take(o_cnt);
  dout.class_map = map_only ? (uint8_t*)take(o_map) : nullptr;
  dout.packed = use_pool ? (uint8_t*)take(o_pack) : nullptr;
  dout.packed_slot_bytes = (int64_t)pslot;
  if (!h->fork_event) ISB_CUDA(h, cudaEventCreateWithFlags(&h->fork_event, cudaEventDisableTiming));
  int64_t d2h_bytes = 0, h2d_bytes = 0, n_dense = 0, n_pool = 0;
  ISB_CUDA(h, cudaEventRecord(h->fork_event, s));
  for (cudaStream_t st : {hp->up, hp->comp, hp->cnt, hp->fast, hp->dense}) ISB_CUDA(h, cudaStreamWaitEvent(st, h->fork_event, 0));
  const size_t slot = (size_t)out->mask_slot_bytes, PP = (size_t)in->proto_h * in->proto_w * K;
  static const bool trace = getenv("ISB_HOST_TRACE") != nullptr;  // per-image delivery timeline on stderr
  const auto t_call = std::chrono::steady_clock::now();
  auto now_us = [&]() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_call).count(); };
  struct TraceRec { int img; char path; double t_ready, t_issue, t_copied, t_done; };
  std::vector<TraceRec> tr;
  if (trace) tr.resize(B, TraceRec{-1, '-', 0, 0, 0, 0});
  // ---- 1. enqueue every chunk's upload (stream `up`) and kernels (stream `comp`) ------------------------------------
  ISB_CUDA(h, cudaMemcpyAsync((void*)din.anchors, in->anchors, n_anc, cudaMemcpyHostToDevice, hp->up));
  h2d_bytes += (int64_t)n_anc;
  rc = ISB_OK;
  h->ws_images = (int32_t)B;  // the per-image workspace is addressed by absolute image index (ws_b0)
  h->in_host_pipe = 1;
  int issued = 0;
  for (int c = 0; c < nchunks && rc == ISB_OK; ++c) {
    const size_t b0 = (size_t)c * chunk, nb = (b0 + chunk < B ? b0 + chunk : B) - b0;
    cudaStream_t upc = hp->up;  // (two alternating upload streams were measured: 26.9 ms against 24.2 ms for one)
    ISB_CUDA(h, cudaMemcpyAsync((char*)din.box + b0 * A * 16, (const char*)in->box + b0 * A * 16, nb * A * 16, cudaMemcpyHostToDevice, upc));
    ISB_CUDA(h, cudaMemcpyAsync((char*)din.conf + b0 * A * C * 4, (const char*)in->conf + b0 * A * C * 4, nb * A * C * 4, cudaMemcpyHostToDevice, upc));
    ISB_CUDA(h, cudaMemcpyAsync((char*)din.coef + b0 * A * K * 4, (const char*)in->coef + b0 * A * K * 4, nb * A * K * 4, cudaMemcpyHostToDevice, upc));
    ISB_CUDA(h, cudaMemcpyAsync((char*)din.proto + b0 * PP * 4, (const char*)in->proto + b0 * PP * 4, nb * PP * 4, cudaMemcpyHostToDevice, upc));
    ISB_CUDA(h, cudaEventRecord(hp->up_ev[c], upc));
    h2d_bytes += (int64_t)(nb * (A * 16 + A * C * 4 + A * K * 4 + PP * 4));
    ISB_CUDA(h, cudaStreamWaitEvent(hp->comp, hp->up_ev[c], 0));
    isb_yolact_in cin = din;
    YolactOutAll cout = dout;
    cin.batch = (int32_t)nb;
    cin.box = din.box + b0 * A * 4;
    cin.conf = din.conf + b0 * A * C;
    cin.coef = din.coef + b0 * A * K;
    cin.proto = din.proto + b0 * PP;
    cout.boxes = dout.boxes + b0 * D * 4;
    cout.scores = dout.scores + b0 * D;
    cout.class_ids = dout.class_ids + b0 * D;
    cout.masks_arg = map_only ? nullptr : dout.masks_arg + b0 * HW;
    cout.masks = map_only ? nullptr : dout.masks + b0 * slot;
    cout.counts = dout.counts + b0;
    cout.class_map = map_only ? dout.class_map + b0 * HW : nullptr;
    cout.packed = use_pool ? dout.packed + b0 * pslot : nullptr;
    h->ws_b0 = (int32_t)b0;
    rc = yolact_detect_device(h, &cin, &cout, hp->comp);
    h->ws_b0 = 0;
    if (rc != ISB_OK) break;
    ISB_CUDA(h, cudaEventRecord(hp->k_ev[c], hp->comp));
    issued = c + 1;
  }
  h->ws_images = 0;
  h->in_host_pipe = 0;
  const double t_enqueued = now_us();
  // ---- 2. results down, as they become ready.  The host issues every device-to-host copy itself, only once its source
  // is complete, so the copy queues never hold blocked work:
  //   kernels of chunk c done  -> counts, records, arg-max (or class map) of the chunk on `fast`;
  //   counts of chunk c here   -> its images join the ready list, which two consumers drain:
  //       copy engine: dense bool [H,W,m] in pieces of kPieceBytes on `dense` (newest ready image first);
  //       host threads: the image's packed bits on `fast` into the landing buffer, then expanded into the caller's
  //       buffer (oldest ready image first).
  // Whichever consumer has a free slot takes the next image, so the split follows the measured rates of this box.
  bool overflow = false;
  if (rc == ISB_OK) {
    std::deque<int> ready;
    struct BulkJob { char* dst; const char* src; size_t bytes; };
    std::deque<BulkJob> bulk;   // arg-max / class maps waiting for the piece-wise copier (taken before any new dense image)
    int next_k = 0, next_cnt = 0, busy = 0;
    int cur_img = -1;           // image the copy engine is working through (-2: a bulk job)
    char* cur_dst = nullptr;
    const char* cur_src = nullptr;
    size_t cur_off = 0, cur_bytes = 0;
    for (int i = 0; i < HostPipe::kDensePieces; ++i) hp->piece_busy[i] = false;
    for (int i = 0; i < HostPipe::kPackedSlots; ++i) { hp->packed[i].state = 0; hp->packed[i].img = -1; }
    static const int env_slots = getenv("ISB_HOST_POOL_SLOTS") ? atoi(getenv("ISB_HOST_POOL_SLOTS")) : 0;  // tuning knob
    const int pool_slots = env_slots > 0 && env_slots < HostPipe::kPackedSlots ? env_slots : HostPipe::kPackedSlots;
    auto cuda_fail = [&](const char* what) {
      snprintf(h->err, sizeof(h->err), "host delivery (%s): %s", what, cudaGetErrorString(cudaGetLastError()));
      rc = ISB_ERR_CUDA;
    };
    std::vector<double> t_up(trace ? issued : 0, 0.0), t_kd(trace ? issued : 0, 0.0), t_cnt(trace ? issued : 0, 0.0);
    int next_up = 0;
    while (rc == ISB_OK && (next_cnt < issued || !ready.empty() || !bulk.empty() || cur_img != -1 || busy > 0)) {
      bool progress = false;
      if (trace && next_up < issued && cudaEventQuery(hp->up_ev[next_up]) == cudaSuccess) t_up[next_up++] = now_us();
      if (next_k < issued) {
        const int d = event_done(h, hp->k_ev[next_k]);
        if (d < 0) { rc = d; break; }
        if (d) {
          const size_t b0 = (size_t)next_k * chunk, nb = (b0 + chunk < B ? b0 + chunk : B) - b0;
          bool ok = cudaMemcpyAsync(out->counts + b0, dout.counts + b0, nb * 4, cudaMemcpyDeviceToHost, hp->cnt) == cudaSuccess;
          ok = ok && cudaEventRecord(hp->cnt_ev[next_k], hp->cnt) == cudaSuccess;
          ok = ok && cudaMemcpyAsync(out->boxes + b0 * D * 4, dout.boxes + b0 * D * 4, nb * D * 16, cudaMemcpyDeviceToHost, hp->cnt) == cudaSuccess;
          ok = ok && cudaMemcpyAsync(out->scores + b0 * D, dout.scores + b0 * D, nb * D * 4, cudaMemcpyDeviceToHost, hp->cnt) == cudaSuccess;
          ok = ok && cudaMemcpyAsync(out->class_ids + b0 * D, dout.class_ids + b0 * D, nb * D * 8, cudaMemcpyDeviceToHost, hp->cnt) == cudaSuccess;
          // the chunk's arg-max maps (or class maps) are bulk traffic: queued for the piece-wise copier below
          if (map_only) bulk.push_back(BulkJob{(char*)(out->class_map + b0 * HW), (const char*)(dout.class_map + b0 * HW), nb * HW});
          else bulk.push_back(BulkJob{(char*)(out->masks_arg + b0 * HW), (const char*)(dout.masks_arg + b0 * HW), nb * HW * 8});
          if (!ok) { cuda_fail("result copies"); break; }
          d2h_bytes += (int64_t)(nb * (4 + D * 28));
          if (trace) t_kd[next_k] = now_us();
          ++next_k;
          progress = true;
        }
      }
      if (next_cnt < next_k) {
        const int d = event_done(h, hp->cnt_ev[next_cnt]);
        if (d < 0) { rc = d; break; }
        if (d) {
          const size_t b0 = (size_t)next_cnt * chunk, b1 = b0 + chunk < B ? b0 + chunk : B;
          for (size_t b = b0; b < b1; ++b) {
            const int m = out->counts[b];
            if (m < 0) overflow = true;
            if (m > 0 && !map_only) ready.push_back((int)b);
            if (trace) { tr[b].img = (int)b; tr[b].t_ready = now_us(); }
          }
          if (trace) t_cnt[next_cnt] = now_us();
          ++next_cnt;
          progress = true;
        }
      }
      // ---- copy engine: dense pieces ---------------------------------------------------------------------------------
      for (int i = 0; i < HostPipe::kDensePieces && rc == ISB_OK; ++i) {
        if (hp->piece_busy[i]) {
          const int d = event_done(h, hp->piece_ev[i]);
          if (d < 0) { rc = d; break; }
          if (d) { hp->piece_busy[i] = false; --busy; progress = true; }
        }
        if (!hp->piece_busy[i]) {
          if (cur_img == -1 && !bulk.empty()) {
            const BulkJob j = bulk.front();
            bulk.pop_front();
            cur_img = -2;
            cur_dst = j.dst; cur_src = j.src; cur_off = 0; cur_bytes = j.bytes;
          }
          if (cur_img == -1 && !ready.empty()) {
            cur_img = ready.back();
            ready.pop_back();
            cur_dst = (char*)out->masks + (size_t)cur_img * slot;
            cur_src = (const char*)dout.masks + (size_t)cur_img * slot;
            cur_off = 0;
            cur_bytes = HW * (size_t)out->counts[cur_img];
            ++n_dense;
            if (trace) { tr[cur_img].path = 'D'; tr[cur_img].t_issue = now_us(); }
          }
          if (cur_img != -1) {
            const size_t nbytes = cur_bytes - cur_off < HostPipe::kPieceBytes ? cur_bytes - cur_off : HostPipe::kPieceBytes;
            if (cudaMemcpyAsync(cur_dst + cur_off, cur_src + cur_off, nbytes, cudaMemcpyDeviceToHost, hp->dense) != cudaSuccess ||
                cudaEventRecord(hp->piece_ev[i], hp->dense) != cudaSuccess) { cuda_fail("bulk copy"); break; }
            hp->piece_busy[i] = true;
            ++busy;
            d2h_bytes += (int64_t)nbytes;
            cur_off += nbytes;
            if (cur_off >= cur_bytes) {
              if (trace && cur_img >= 0) tr[cur_img].t_done = now_us();  // (last piece issued)
              cur_img = -1;
            }
            progress = true;
          }
        }
      }
      // ---- host threads: packed bits, then expansion --------------------------------------------------------------------
      for (int i = 0; use_pool && i < pool_slots && rc == ISB_OK; ++i) {
        HostPipe::PackedSlot& ps = hp->packed[i];
        if (ps.state == 1) {
          const int d = event_done(h, ps.ev);
          if (d < 0) { rc = d; break; }
          if (d) {
            ps.job.src = reinterpret_cast<const uint64_t*>(hp->land + (size_t)i * pslot);
            ps.job.dst = out->masks + (size_t)ps.img * slot;
            ps.job.out_bytes = HW * (size_t)out->counts[ps.img];
            isb::host_pool_submit(hp->pool, &ps.job);
            if (trace) tr[ps.img].t_copied = now_us();
            ps.state = 2;
            progress = true;
          }
        } else if (ps.state == 2 && ps.job.done()) {
          if (trace) tr[ps.img].t_done = now_us();
          ps.state = 0;
          --busy;
          progress = true;
        }
        if (ps.state == 0 && !ready.empty()) {
          const int b = ready.front();
          ready.pop_front();
          const size_t bytes = 4 * ((HW * (size_t)out->counts[b] + 31) / 32);
          if (cudaMemcpyAsync(hp->land + (size_t)i * pslot, dout.packed + (size_t)b * pslot, bytes, cudaMemcpyDeviceToHost, hp->fast) != cudaSuccess ||
              cudaEventRecord(ps.ev, hp->fast) != cudaSuccess) { cuda_fail("packed copy"); break; }
          ps.img = b;
          ps.state = 1;
          if (trace) { tr[b].path = 'P'; tr[b].t_issue = now_us(); }
          ++busy; ++n_pool;
          d2h_bytes += (int64_t)bytes;
          progress = true;
        }
      }
      if (!progress) std::this_thread::yield();
    }
    if (trace)
      for (int c = 0; c < issued; ++c)
        fprintf(stderr, "[isb host] chunk %2d uploaded %7.0f kernels done %7.0f counts here %7.0f\n", c, t_up[c], t_kd[c], t_cnt[c]);
    if (rc != ISB_OK)  // wait for whatever is still running before the caller's buffers can be released
      for (int i = 0; use_pool && i < HostPipe::kPackedSlots; ++i)
        while (hp->packed[i].state == 2 && !hp->packed[i].job.done()) std::this_thread::yield();
  }
  // join (also on the error path, so nothing of this call is in flight when we return)
  for (cudaStream_t st : {hp->up, hp->comp, hp->cnt, hp->fast, hp->dense}) {
    const cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess && rc == ISB_OK) {
      snprintf(h->err, sizeof(h->err), "host delivery (join): %s", cudaGetErrorString(e));
      rc = ISB_ERR_CUDA;
    }
  }
  if (trace) {
    fprintf(stderr, "[isb host] B=%zu chunks=%d enqueued at %.0f us, total %.0f us, dense %lld, pool %lld (threads %d)\n", B, nchunks,
            t_enqueued, now_us(), (long long)n_dense, (long long)n_pool, use_pool ? hp->pool_threads : 0);
    for (size_t b = 0; b < tr.size(); ++b)
      if (tr[b].img >= 0)
        fprintf(stderr, "[isb host] img %2zu %c ready %7.0f issue %7.0f copied %7.0f done %7.0f\n", b, tr[b].path, tr[b].t_ready,
                tr[b].t_issue, tr[b].t_copied, tr[b].t_done);
  }
  h->host_stats[0] = n_dense; h->host_stats[1] = n_pool; h->host_stats[2] = d2h_bytes; h->host_stats[3] = h2d_bytes;
  h->host_stats[4] = use_pool ? hp->pool_threads : 0;
  if (rc != ISB_OK) return rc;
  if (overflow) ISB_FAIL(h, ISB_ERR_OVERFLOW, "candidate buffer overflow: raise isb_limits.cand_capacity (now %lld)", (long long)h->lim.cand_capacity);
  return ISB_OK;
}

// ----------------------------------------------------------------------------------- fused count exchange
static void cx_release(isb_handle* h) {
  for (int r = 0; r < 8; ++r) {
    if (h->cx_peer[r] && r != h->cx_rank && h->cx_attached) cudaIpcCloseMemHandle(h->cx_peer[r]);
    h->cx_peer[r] = nullptr;
  }
  cudaFree(h->cx_ring);
  cudaFree(h->cx_status);
  h->cx_ring = nullptr;
  h->cx_status = nullptr;
  h->cx_attached = 0;
  h->cx_world = h->cx_local = h->cx_depth = 0;
}

extern "C" int isb_counts_export(isb_handle* h, int32_t n_local, int32_t world, int32_t depth, unsigned char handle_out[64]) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!handle_out || n_local <= 0 || world < 1 || world > 8 || depth < 2 || depth > 64)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_export: need n_local > 0, 1 <= world <= 8, 2 <= depth <= 64");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  ISB_SET_DEVICE(h);
  ISB_CUDA(h, cudaDeviceSynchronize());
  cx_release(h);
  const size_t words = (size_t)depth * world * n_local;
  ISB_CUDA(h, cudaMalloc((void**)&h->cx_ring, words * sizeof(unsigned long long)));
  ISB_CUDA(h, cudaMemset(h->cx_ring, 0, words * sizeof(unsigned long long)));  // epoch 0 = nothing published yet
  ISB_CUDA(h, cudaMalloc((void**)&h->cx_status, sizeof(int32_t)));
  ISB_CUDA(h, cudaDeviceSynchronize());
  cudaIpcMemHandle_t ipc;
  ISB_CUDA(h, cudaIpcGetMemHandle(&ipc, h->cx_ring));
  memcpy(handle_out, &ipc, 64);
  h->cx_world = world;
  h->cx_local = n_local;
  h->cx_depth = depth;
  h->cx_epoch = 0;
  return ISB_OK;
}

extern "C" int isb_counts_attach(isb_handle* h, int32_t rank, const unsigned char* handles) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!h->cx_ring) ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_attach: call isb_counts_export first");
  if (!handles || rank < 0 || rank >= h->cx_world) ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_attach: bad rank or null handles");
  ISB_SET_DEVICE(h);
  h->cx_rank = rank;
  for (int r = 0; r < h->cx_world; ++r) {
    if (r == rank) { h->cx_peer[r] = h->cx_ring; continue; }
    cudaIpcMemHandle_t ipc;
    memcpy(&ipc, handles + (size_t)r * 64, 64);
    void* ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, ipc, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int q = 0; q < r; ++q)
        if (q != rank && h->cx_peer[q]) { cudaIpcCloseMemHandle(h->cx_peer[q]); h->cx_peer[q] = nullptr; }
      ISB_FAIL(h, ISB_ERR_CUDA, "counts_attach: cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
    }
    h->cx_peer[r] = static_cast<unsigned long long*>(ptr);
  }
  h->cx_attached = 1;
  return ISB_OK;
}

extern "C" int isb_counts_detach(isb_handle* h) {
  if (!h) return ISB_ERR_BAD_ARG;
  ISB_SET_DEVICE(h);
  ISB_CUDA(h, cudaDeviceSynchronize());
  cx_release(h);
  return ISB_OK;
}

extern "C" int64_t isb_counts_epoch(const isb_handle* h) { return h ? (int64_t)h->cx_epoch : -1; }

extern "C" int isb_counts_wait(isb_handle* h, int64_t epoch, int32_t* counts, int64_t* offsets, int32_t* status,
                               int32_t timeout_ms, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!h->cx_attached) ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_wait: no count exchange attached");
  if (!counts) ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_wait: null counts");
  if (epoch < 0) epoch = h->cx_epoch;
  if (epoch == 0 || epoch > (int64_t)h->cx_epoch) ISB_FAIL(h, ISB_ERR_BAD_ARG, "counts_wait: epoch %lld has not been published by this rank", (long long)epoch);
  if ((int64_t)h->cx_epoch - epoch >= h->cx_depth) ISB_FAIL(h, ISB_ERR_LIMIT, "counts_wait: epoch %lld left the ring (depth %d)", (long long)epoch, h->cx_depth);
  ISB_SET_DEVICE(h);
  const int n = h->cx_world * h->cx_local;
  const unsigned long long* slot = h->cx_ring + (size_t)((uint32_t)epoch % (uint32_t)h->cx_depth) * n;
  if (timeout_ms <= 0) timeout_ms = 2000;
  return launch_counts_wait(h, slot, n, (unsigned)epoch, counts, offsets, status ? status : h->cx_status,
                            (long long)timeout_ms * 1000000ll, (cudaStream_t)stream);
}

// ----------------------------------------------------------------------------------- stages
extern "C" int isb_decode_boxes(isb_handle* h, const float* pred, const float* anchors, int64_t n, float var0, float var1,
                                float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (n < 0 || !out || (n > 0 && (!pred || !anchors))) ISB_FAIL(h, ISB_ERR_BAD_ARG, "decode_boxes: bad argument");
  if (!aligned16(pred) || !aligned16(anchors) || !aligned16(out)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "decode_boxes: pointers must be 16-byte aligned");
  if (n == 0) return ISB_OK;
  ISB_SET_DEVICE(h);
  return launch_decode_boxes(h, pred, anchors, n, var0, var1, out, (cudaStream_t)stream);
}

extern "C" int isb_jaccard(isb_handle* h, const float* a, const float* b, int32_t batch, int32_t na, int32_t nb,
                           int32_t iscrowd, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (batch < 0 || na < 0 || nb < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "jaccard: negative size");
  if ((int64_t)batch * na * nb == 0) return ISB_OK;
  if (!a || !b || !out || !aligned16(a) || !aligned16(b)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "jaccard: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return launch_jaccard(h, a, b, batch, na, nb, iscrowd, out, (cudaStream_t)stream);
}

extern "C" int isb_fast_nms(isb_handle* h, const float* box, const float* scores, const float* coef, int32_t n,
                            int32_t num_fg_classes, int32_t num_coef, float nms_iou, int32_t top_k, int32_t max_det,
                            float* out_box, float* out_score, int64_t* out_class, float* out_coef, int32_t* count,
                            isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!count || !out_box || !out_score || !out_class || !out_coef) ISB_FAIL(h, ISB_ERR_BAD_ARG, "fast_nms: null output");
  if (n < 0 || num_fg_classes <= 0 || top_k <= 0 || max_det <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "fast_nms: bad size");
  if (num_coef != ISB_NUM_COEF) ISB_FAIL(h, ISB_ERR_BAD_ARG, "fast_nms: num_coef must be %d", ISB_NUM_COEF);
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) {
    ISB_CUDA(h, cudaMemsetAsync(count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  if (!box || !scores || !coef || !aligned16(box) || !aligned16(coef) || !aligned16(out_box) || !aligned16(out_coef))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "fast_nms: null or misaligned pointer");
  if (n > h->lim.max_anchors) ISB_FAIL(h, ISB_ERR_LIMIT, "fast_nms: n %d > max_anchors %d", n, h->lim.max_anchors);
  if (num_fg_classes > h->lim.max_classes) ISB_FAIL(h, ISB_ERR_LIMIT, "fast_nms: classes %d > max_classes %d", num_fg_classes, h->lim.max_classes);
  if (max_det > h->lim.max_det) ISB_FAIL(h, ISB_ERR_LIMIT, "fast_nms: max_det %d > limit %d", max_det, h->lim.max_det);
  if (top_k > h->lim.max_top_k) ISB_FAIL(h, ISB_ERR_LIMIT, "fast_nms: top_k %d > limit %d", top_k, h->lim.max_top_k);
  if ((int64_t)n * num_fg_classes > h->lim.cand_capacity)
    ISB_FAIL(h, ISB_ERR_LIMIT, "fast_nms: n*classes = %lld > cand_capacity %lld", (long long)n * num_fg_classes, (long long)h->lim.cand_capacity);
  int rc;
  CandParams cp;
  cp.scores = scores;
  cp.rows = n;
  cp.ncols = num_fg_classes;
  cp.col0 = 0;
  cp.use_threshold = 0;  // the stand-alone method ranks every entry (utils_bbox.py:55-64)
  cp.threshold = 0.0f;
  cp.batch = 1;
  cp.logits = 0;
  if ((rc = launch_candidates(h, cp, s)) != ISB_OK) return rc;
  NmsParams np;
  memset(&np, 0, sizeof(np));
  np.batch = 1;
  np.boxes = reinterpret_cast<const float4*>(box);
  np.coef = coef;
  np.rows = n;
  np.nms_iou = nms_iou;
  np.top_k = top_k;
  np.max_det = max_det;
  np.correct = 0;
  np.out_box = out_box;
  np.out_score = out_score;
  np.out_class = out_class;
  np.out_coef = out_coef;
  np.out_count = count;
  // one image whose candidates do not fit the single-CTA kernel's shared memory: spread the work over the GPU
  if ((int64_t)n * num_fg_classes > fast_nms_smem_candidates()) return launch_fast_nms_wide(h, np, n * num_fg_classes, s);
  return launch_fast_nms(h, np, s);
}

extern "C" int isb_traditional_nms(isb_handle* h, const float* box, const float* class_scores, const float* coef,
                                   const float* max_scores, int32_t n, int32_t num_fg_classes, int32_t num_coef,
                                   double nms_iou, int32_t max_det, float* out_box, float* out_score, int64_t* out_class,
                                   float* out_coef, int32_t* count, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!count || !out_box || !out_score || !out_class || !out_coef) ISB_FAIL(h, ISB_ERR_BAD_ARG, "traditional_nms: null output");
  if (n < 0 || num_fg_classes <= 0 || max_det <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "traditional_nms: bad size");
  if (num_coef != ISB_NUM_COEF) ISB_FAIL(h, ISB_ERR_BAD_ARG, "traditional_nms: num_coef must be %d", ISB_NUM_COEF);
  if (num_fg_classes >= (1 << kClassBits)) ISB_FAIL(h, ISB_ERR_LIMIT, "traditional_nms: classes %d >= 2^%d", num_fg_classes, kClassBits);
  if (n >= (1 << kRowBits)) ISB_FAIL(h, ISB_ERR_LIMIT, "traditional_nms: n %d >= 2^%d", n, kRowBits);
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) {
    ISB_CUDA(h, cudaMemsetAsync(count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  if (!box || !class_scores || !coef || !max_scores || !aligned16(box) || !aligned16(coef) || !aligned16(out_box) ||
      !aligned16(out_coef))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "traditional_nms: null or misaligned pointer");
  const size_t b_lab = ((size_t)n * 4 + 255) & ~(size_t)255, b_keep = ((size_t)n * 8 + 255) & ~(size_t)255;
  int rc = ensure_scratch(h, b_lab + b_keep + 256);
  if (rc != ISB_OK) return rc;
  int32_t* labels = reinterpret_cast<int32_t*>(h->scratch);
  int64_t* keep = reinterpret_cast<int64_t*>(static_cast<char*>(h->scratch) + b_lab);
  int32_t* kept = reinterpret_cast<int32_t*>(static_cast<char*>(h->scratch) + b_lab + b_keep);
  if ((rc = launch_row_argmax(h, class_scores, n, num_fg_classes, labels, s)) != ISB_OK) return rc;
  NmsCall c;
  memset(&c, 0, sizeof(c));
  c.boxes = reinterpret_cast<const float4*>(box);
  c.scores = max_scores;
  c.labels32 = labels;
  c.n_max = n;
  c.iou_threshold = nms_iou;
  c.max_keep = max_det;  // the global order is by score, so the first max_det kept boxes are the final answer (:136)
  c.keep = keep;
  c.count = kept;
  if ((rc = launch_nms(h, c, s)) != ISB_OK) return rc;
  return launch_trad_gather(h, keep, kept, max_det, box, max_scores, labels, coef, out_box, out_score, out_class, out_coef,
                            count, s);
}

extern "C" int isb_correct_boxes(isb_handle* h, const float* boxes, int64_t m, int32_t out_h, int32_t out_w, float* out,
                                 isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (m < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "correct_boxes: negative m");
  if (m == 0) return ISB_OK;
  if (!boxes || !out || !aligned16(boxes) || !aligned16(out)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "correct_boxes: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return launch_correct_boxes(h, boxes, m, out_h, out_w, out, (cudaStream_t)stream);
}

extern "C" int isb_crop(isb_handle* h, const float* masks, const float* boxes, int32_t hgt, int32_t wid, int32_t m,
                        float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (hgt < 0 || wid < 0 || m < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "crop: negative size");
  if ((int64_t)hgt * wid * m == 0) return ISB_OK;
  if (!masks || !boxes || !out || !aligned16(boxes)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "crop: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return launch_crop(h, masks, boxes, hgt, wid, m, out, (cudaStream_t)stream);
}

extern "C" int isb_mask_logits(isb_handle* h, const float* proto, const float* coef, int32_t proto_h, int32_t proto_w,
                               int32_t m, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (proto_h < 0 || proto_w < 0 || m < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_logits: negative size");
  if ((int64_t)proto_h * proto_w * m == 0) return ISB_OK;
  if (!proto || !coef || !out || !aligned16(proto) || !aligned16(coef)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_logits: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return launch_mask_logits_tc(h, proto, coef, (int64_t)proto_h * proto_w, m, out, (cudaStream_t)stream);
}

extern "C" int isb_mask_logits_batch(isb_handle* h, const float* proto, const float* coef, int32_t batch, int32_t proto_h,
                                     int32_t proto_w, int32_t m, int32_t apply_sigmoid, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (batch < 0 || proto_h < 0 || proto_w < 0 || m < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_logits_batch: negative size");
  if ((int64_t)batch * proto_h * proto_w * m == 0) return ISB_OK;
  if (m > kTcBatchMaxDet) ISB_FAIL(h, ISB_ERR_LIMIT, "mask_logits_batch: m %d > %d (use isb_mask_logits per picture)", m, kTcBatchMaxDet);
  if ((int64_t)proto_h * proto_w >= (1ll << 31)) ISB_FAIL(h, ISB_ERR_LIMIT, "mask_logits_batch: prototype map too large");
  if (!proto || !coef || !out || !aligned16(proto) || !aligned16(coef) || !aligned16(out))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_logits_batch: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  TcBatchArgs a;
  memset(&a, 0, sizeof(a));
  a.proto = proto; a.coef = coef; a.coef_rows = m; a.counts = nullptr; a.m_fixed = m; a.batch = batch;
  a.npix = proto_h * proto_w; a.proto_w = proto_w; a.mode = apply_sigmoid ? 1 : 0; a.out = out;
  return launch_mask_logits_batch(h, a, (cudaStream_t)stream);
}

extern "C" int isb_mask_assemble(isb_handle* h, const float* proto, const float* coef, const float* boxes_px,
                                 int32_t proto_h, int32_t proto_w, int32_t m, int32_t out_h, int32_t out_w,
                                 int64_t* masks_arg, uint8_t* masks, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (proto_h <= 0 || proto_w <= 0 || out_h <= 0 || out_w <= 0 || m < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_assemble: bad size");
  if (m == 0) return ISB_OK;
  if (m > h->lim.max_det) ISB_FAIL(h, ISB_ERR_LIMIT, "mask_assemble: m %d > max_det %d", m, h->lim.max_det);
  if (!proto || !coef || !boxes_px || !masks_arg || !masks || !aligned16(proto) || !aligned16(coef) ||
      !aligned16(boxes_px) || !aligned16(masks))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_assemble: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = launch_det_layout(h, boxes_px, m, out_h, out_w, proto_h, proto_w, h->det_rec, h->smap_total, s);
  if (rc != ISB_OK) return rc;
  MaskParams mp;
  memset(&mp, 0, sizeof(mp));
  mp.batch = 1;
  mp.proto = proto;
  mp.coef = coef;
  mp.counts = nullptr;
  mp.m_fixed = m;
  mp.det_stride = m;
  mp.proto_h = proto_h;
  mp.proto_w = proto_w;
  mp.out_h = out_h;
  mp.out_w = out_w;
  mp.masks_arg = masks_arg;
  mp.masks = masks;
  mp.mask_slot_bytes = (((int64_t)out_h * out_w * m) + 15) & ~15ll;
  return launch_masks(h, mp, s);
}
