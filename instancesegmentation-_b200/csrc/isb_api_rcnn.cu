// C-ABI entry points of the Mask R-CNN side (see include/isb.h).
#include "isb_common.cuh"

using namespace isb;

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int isb_box_decode(isb_handle* h, const float* rel_codes, const float* boxes, int64_t n, int32_t k, float wx,
                              float wy, float ww, float wh, float clip, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (n < 0 || k <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_decode: bad size");
  if (n == 0) return ISB_OK;
  if (!rel_codes || !boxes || !out || !al16(rel_codes) || !al16(boxes) || !al16(out))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_decode: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  const float w[4] = {wx, wy, ww, wh};
  return launch_box_decode(h, rel_codes, boxes, n, k, w, clip, out, (cudaStream_t)stream);
}

static int nms_common(isb_handle* h, const float* boxes, const float* scores, const int64_t* idxs, int32_t n,
                      double thr, int64_t* keep, int32_t* count, isb_stream stream, const char* what) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (n < 0 || !count) ISB_FAIL(h, ISB_ERR_BAD_ARG, "%s: bad size or null count", what);
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) {  // boxes.py:66-67: empty input -> empty int64 result
    ISB_CUDA(h, cudaMemsetAsync(count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  if (!boxes || !scores || !keep || !al16(boxes)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "%s: null or misaligned pointer", what);
  NmsCall c;
  memset(&c, 0, sizeof(c));
  c.boxes = reinterpret_cast<const float4*>(boxes);
  c.scores = scores;
  c.idxs64 = idxs;
  c.n_max = n;
  c.iou_threshold = thr;
  c.max_keep = 0;
  c.keep = keep;
  c.count = count;
  return launch_nms(h, c, s);
}

extern "C" int isb_nms(isb_handle* h, const float* boxes, const float* scores, int32_t n, double iou_threshold,
                       int64_t* keep, int32_t* count, isb_stream stream) {
  return nms_common(h, boxes, scores, nullptr, n, iou_threshold, keep, count, stream, "nms");
}

extern "C" int isb_batched_nms(isb_handle* h, const float* boxes, const float* scores, const int64_t* idxs, int32_t n,
                               double iou_threshold, int64_t* keep, int32_t* count, isb_stream stream) {
  if (h && n > 0 && !idxs) ISB_FAIL(h, ISB_ERR_BAD_ARG, "batched_nms: null idxs");
  return nms_common(h, boxes, scores, idxs, n, iou_threshold, keep, count, stream, "batched_nms");
}

extern "C" int isb_ms_roi_align(isb_handle* h, const isb_roi_align_cfg* cfg, const float* boxes, const int32_t* batch_idx,
                                int32_t num_rois, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!cfg || num_rois < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: null cfg or negative roi count");
  if (cfg->num_levels < 1 || cfg->num_levels > 8) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: 1..8 levels supported");
  if (cfg->sampling_ratio <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: sampling_ratio must be > 0 (the reference passes 2)");
  if (cfg->out_h <= 0 || cfg->out_w <= 0 || cfg->channels <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: bad output size");
  if (cfg->out_h * cfg->sampling_ratio > 128 || cfg->out_w * cfg->sampling_ratio > 128)
    ISB_FAIL(h, ISB_ERR_LIMIT, "ms_roi_align: out*sampling_ratio must be <= 128 per axis");
  if (num_rois == 0) return ISB_OK;
  if (!boxes || !out || !al16(boxes)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: null or misaligned pointer");
  RoiLevels L;
  memset(&L, 0, sizeof(L));
  L.num_levels = cfg->num_levels;
  for (int i = 0; i < cfg->num_levels; ++i) {
    if (!cfg->feat[i] || cfg->feat_h[i] <= 0 || cfg->feat_w[i] <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "ms_roi_align: bad level %d", i);
    L.feat[i] = cfg->feat[i];
    L.h[i] = cfg->feat_h[i];
    L.w[i] = cfg->feat_w[i];
    L.scale[i] = cfg->scale[i];
  }
  L.k_min = cfg->k_min;
  L.k_max = cfg->k_max;
  L.canonical_scale = (float)cfg->canonical_scale;
  L.canonical_level = (float)cfg->canonical_level;
  L.eps = 1e-6f;
  ISB_SET_DEVICE(h);
  return launch_ms_roi_align(h, L, boxes, batch_idx, num_rois, cfg->channels, cfg->out_h, cfg->out_w, cfg->sampling_ratio,
                             out, (cudaStream_t)stream);
}

extern "C" int isb_mask_select_sigmoid(isb_handle* h, const float* logits, const int64_t* labels, int32_t m,
                                       int32_t num_classes, int32_t mask_size, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (m < 0 || num_classes <= 0 || mask_size <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_select_sigmoid: bad size");
  if (m == 0) return ISB_OK;
  if (!logits || !labels || !out) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask_select_sigmoid: null pointer");
  ISB_SET_DEVICE(h);
  return launch_mask_select_sigmoid(h, logits, labels, m, num_classes, mask_size * mask_size, out, (cudaStream_t)stream);
}

extern "C" int isb_resize_boxes(isb_handle* h, const float* boxes, int64_t m, int32_t orig_h, int32_t orig_w, int32_t new_h,
                                int32_t new_w, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (m < 0 || orig_h <= 0 || orig_w <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "resize_boxes: bad size");
  if (m == 0) return ISB_OK;
  if (!boxes || !out || !al16(boxes) || !al16(out)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "resize_boxes: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  // transform.py:471-475: float32(new) / float32(orig) per axis
  const float rh = (float)new_h / (float)orig_h, rw = (float)new_w / (float)orig_w;
  return launch_resize_boxes(h, boxes, m, rw, rh, out, (cudaStream_t)stream);
}

extern "C" int isb_paste_masks(isb_handle* h, const float* masks, const float* boxes, int32_t m, int32_t mask_size,
                               int32_t padding, int32_t im_h, int32_t im_w, float* out, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (m < 0 || mask_size <= 0 || padding < 0 || im_h <= 0 || im_w <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "paste_masks: bad size");
  if (m == 0) return ISB_OK;
  if (!masks || !boxes || !out || !al16(boxes) || !al16(out)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "paste_masks: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return launch_paste_masks(h, masks, boxes, m, mask_size, padding, im_h, im_w, out, (cudaStream_t)stream);
}

namespace isb {
int launch_paste_class_map(isb_handle* h, const float* masks, const float* boxes, const int64_t* labels, const float* scores, int m,
                           int S, int pad, int paste_h, int paste_w, int out_h, int out_w, float confidence, uint8_t* class_map,
                           int64_t* masks_arg, int32_t* kept_count, cudaStream_t s);
}

extern "C" int isb_paste_class_map(isb_handle* h, const float* masks, const float* boxes, const int64_t* labels,
                                   const float* scores, int32_t m, int32_t mask_size, int32_t padding, int32_t paste_h,
                                   int32_t paste_w, int32_t out_h, int32_t out_w, float confidence, uint8_t* class_map,
                                   int64_t* masks_arg, int32_t* kept_count, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (m < 0 || mask_size <= 0 || padding < 0 || paste_h <= 0 || paste_w <= 0 || out_h <= 0 || out_w <= 0)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "paste_class_map: bad size");
  if ((int64_t)out_h * out_w >= (1ll << 31)) ISB_FAIL(h, ISB_ERR_LIMIT, "paste_class_map: output image too large");
  if (!class_map) ISB_FAIL(h, ISB_ERR_BAD_ARG, "paste_class_map: null class_map");
  if (m > 0 && (!masks || !boxes || !labels || !scores || !al16(boxes)))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "paste_class_map: null or misaligned pointer");
  ISB_SET_DEVICE(h);
  return isb::launch_paste_class_map(h, masks, boxes, labels, scores, m, mask_size, padding, paste_h, paste_w, out_h, out_w,
                                     confidence, class_map, masks_arg, kept_count, (cudaStream_t)stream);
}

namespace isb {
int launch_rcnn_transform(isb_handle* h, const float* img, int in_h, int in_w, int out_h, int out_w, const float mean[3],
                          const float stdv[3], float* canvas, int can_h, int can_w, cudaStream_t s);
}

extern "C" int isb_rcnn_transform(isb_handle* h, const float* image, int32_t in_h, int32_t in_w, int32_t out_h, int32_t out_w,
                                  const float* mean, const float* std, float* canvas, int32_t canvas_h, int32_t canvas_w,
                                  isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || canvas_h < out_h || canvas_w < out_w)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "rcnn_transform: bad size (the canvas must hold the resized image)");
  if (canvas_h > 65535) ISB_FAIL(h, ISB_ERR_LIMIT, "rcnn_transform: canvas higher than 65535 rows");
  if (!image || !mean || !std || !canvas) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rcnn_transform: null pointer");
  ISB_SET_DEVICE(h);
  return isb::launch_rcnn_transform(h, image, in_h, in_w, out_h, out_w, mean, std, canvas, canvas_h, canvas_w, (cudaStream_t)stream);
}
