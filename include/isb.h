/* isb.h -- C ABI of libisb.so: B200 (sm_100a) kernels for the instance-segmentation
 * post-processing path of Leyan529/InstanceSegmentation-.
 *
 * The reference has no FFI layer; its boundary for this path is a set of Python call
 * signatures.  Each entry point below names the reference
 * function(s) it replaces as file:line under the reference tree.  All pointers are plain
 * DEVICE pointers unless the name says `_host`; tensors are dense row-major float32 /
 * int64 / uint8 exactly as the reference lays them out.  Nothing here depends on torch.
 *
 * Conventions
 *   - every call returns an isb_status (0 = ok, < 0 = error); the message is available
 *     from isb_last_error(handle).  "No detections" is a value (count 0), not an error.
 *   - all kernels are enqueued on the caller's stream; no call synchronises the device
 *     except the `_host` variants (which must, to hand back host buffers).
 *   - a handle is bound to one device and is not thread-safe: one handle per (device,
 *     stream) in flight.
 */
#ifndef ISB_H_
#define ISB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ISB_API __attribute__((visibility("default")))

#define ISB_VERSION 100 /* 0.1.0 */
#define ISB_NUM_COEF 32 /* prototype / coefficient channels (reference nets/yolact.py:70-111) */
#define ISB_MAX_DET_LIMIT 1024

typedef struct isb_handle isb_handle;
typedef void* isb_stream; /* cudaStream_t */

typedef enum {
  ISB_OK = 0,
  ISB_ERR_BAD_ARG = -1,  /* null pointer, bad shape, misaligned pointer                 */
  ISB_ERR_CUDA = -2,     /* a CUDA runtime call failed                                  */
  ISB_ERR_LIMIT = -3,    /* request exceeds the isb_limits the handle was created with  */
  ISB_ERR_OVERFLOW = -4  /* reported by a `_host` call: candidate buffer too small      */
} isb_status;

/* Upper bounds the workspace is sized for.  Zero fields take the defaults in brackets. */
typedef struct {
  int32_t max_batch;      /* images per call                                    [64]    */
  int32_t max_anchors;    /* anchors / rows per image                           [19248] */
  int32_t max_classes;    /* foreground classes                                 [80]    */
  int32_t max_det;        /* detections kept per image, <= ISB_MAX_DET_LIMIT    [100]   */
  int32_t max_top_k;      /* per-class Fast-NMS window                          [200]   */
  int64_t cand_capacity;  /* (anchor,class) candidates buffered per image       [16384] */
} isb_limits;

ISB_API int isb_version(void);
ISB_API int isb_create(int device, const isb_limits* limits, isb_handle** out);
ISB_API void isb_destroy(isb_handle* h);
ISB_API const char* isb_last_error(const isb_handle* h);
ISB_API size_t isb_workspace_bytes(const isb_handle* h);
/* Number of kernels this library has launched through the handle (bench `gpu_launches`). */
ISB_API int64_t isb_launch_count(const isb_handle* h);

/* Per-stage device timing of isb_yolact_detect, taken with CUDA events recorded on the launch
 * stream between the kernels (no synchronisation while recording).  isb_profile_begin arms a
 * ring of `max_calls` slots (0 disables); every following isb_yolact_detect call fills one slot;
 * isb_profile_read synchronises on the last slot and returns, per stage (0 = candidate filter,
 * 1 = sort + Fast NMS, 2 = sigmoid maps, 3 = mask compose), the summed milliseconds over the recorded calls. */
ISB_API int isb_profile_begin(isb_handle* h, int32_t max_calls);
ISB_API int isb_profile_read(isb_handle* h, double stage_ms[4], int32_t* calls);

/* ------------------------------------------------------------------------------------
 * YOLACT Detect + mask assembly, fused, for a batch of images.
 * Replaces BBoxUtility.decode_nms, inst_model/yolact/utils/utils_bbox.py:170-243
 * (traditional_nms=False branch), i.e. decode_boxes :11, score filter :193-204,
 * fast_non_max_suppression :50-100, second threshold :208-212, yolact_correct_boxes :142,
 * matmul+sigmoid :223, bilinear upsample :227-229, crop :154-168, argmax :236, >0.5 :241.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  int32_t batch;        /* B                                                            */
  int32_t num_anchors;  /* A                                                            */
  int32_t num_classes;  /* C, INCLUDING the background column 0                         */
  int32_t num_coef;     /* K, must equal ISB_NUM_COEF                                   */
  int32_t proto_h, proto_w;
  int32_t out_h, out_w; /* image_shape (H, W) the masks are upsampled to                */
  const float* box;     /* [B,A,4]  raw box regressions                                 */
  const float* conf;    /* [B,A,C]  softmaxed class scores                              */
  const float* coef;    /* [B,A,K]  mask coefficients                                   */
  const float* proto;   /* [B,Ph,Pw,K] prototypes, channels last                        */
  const float* anchors; /* [A,4]    priors (cx,cy,w,h), shared by the batch             */
  float confidence;
  float nms_iou;
  int32_t top_k;        /* reference default 200 (utils_bbox.py:50)                     */
  int32_t max_det;      /* reference default 100                                        */
  int32_t conf_is_logits; /* 0: conf is softmaxed (the reference's input).  1: conf holds RAW class logits and
                             the softmax of nets/yolact.py:204 (F.softmax(pred_classes, -1)) is fused into the
                             score filter -- the [B,A,C] probabilities are never written or re-read (SURVEY 8f
                             rank 2).  conf must then be 16-byte aligned.                              */
} isb_yolact_in;

typedef struct {
  float* boxes;          /* [B,max_det,4] pixel corners, rows >= counts[b] untouched    */
  float* scores;         /* [B,max_det]                                                 */
  int64_t* class_ids;    /* [B,max_det]  0-based foreground class                       */
  int64_t* masks_arg;    /* [B,H,W]      per-pixel argmax over detections               */
  uint8_t* masks;        /* B slots of mask_slot_bytes; slot b holds bool [H,W,m_b]
                            densely packed with m_b = counts[b] innermost               */
  int64_t mask_slot_bytes; /* >= H*W*max_det, multiple of 16                            */
  int32_t* counts;       /* [B] m_b; 0 = no detection (reference returns 5x None);
                            -1 = candidate buffer overflow (raise cand_capacity)        */
} isb_yolact_out;

ISB_API int isb_yolact_detect(isb_handle* h, const isb_yolact_in* in, const isb_yolact_out* out, isb_stream stream);

/* Same computation with HOST buffers (pinned recommended): copies the five inputs to the
 * device, runs isb_yolact_detect, copies every output back and synchronises.  All
 * pointers in `in` / `out` are host pointers.  This is the end-to-end path bench.py
 * reports as `e2e`.  Returns ISB_ERR_OVERFLOW if any counts[b] is -1. */
ISB_API int isb_yolact_detect_host(isb_handle* h, const isb_yolact_in* in, const isb_yolact_out* out, isb_stream stream);

/* How isb_yolact_detect_host delivers bool [H,W,m] to the host.  97 % of those bytes are zeros and one byte per bit is
 * the reference's format (utils_bbox.py:241), so the call is bound by the device-to-host link; the device therefore ALSO
 * writes every image's masks one bit per element, and two consumers drain the batch concurrently: the copy engine moves
 * images densely (as before), and `threads` host worker threads take other images as packed bits (an eighth of the PCIe
 * bytes) and expand them into the caller's buffer with non-temporal stores.  Which image goes which way is decided at
 * run time by whichever consumer is free, so the split follows the box (link rate, cores, memory bandwidth).  The
 * caller's contract does not change: every byte of out->masks[b][0 .. H*W*m_b) is written, identical to the device path.
 *   threads < 0: default = min(8, CPUs of this process / LOCAL_WORLD_SIZE - 1) (ISB_HOST_THREADS overrides);  0: copy engine only. */
ISB_API int isb_set_host_threads(isb_handle* h, int32_t threads);

/* Counters of the last *_host call: stats[0] images delivered by the copy engine, [1] images expanded by host threads,
 * [2] device-to-host bytes copied, [3] host-to-device bytes copied, [4] worker threads used, [5..7] reserved (0). */
ISB_API int isb_host_stats(const isb_handle* h, int64_t stats[8]);

/* ------------------------------------------------------------------------------------
 * Multi-GPU: the path's one exchange -- every rank learns every image's detection count (the reference has
 * no multi-GPU inference path: utils_bbox.py:177-180 handles one image) -- fused into isb_yolact_detect over NVLink peer
 * memory instead of a collective launched after it.
 *   isb_counts_export  allocates this rank's ring [depth][world * n_local] of 64-bit words and returns its 64-byte CUDA IPC
 *                      handle; the caller exchanges the handles of all ranks (any side channel, once);
 *   isb_counts_attach  maps every peer's ring (handles = world x 64 bytes, rank-major).  From then on every
 *                      isb_yolact_detect call with batch == n_local advances the epoch and its NMS kernel stores
 *                      epoch << 32 | count of each of its images into slot (epoch mod depth) of EVERY rank's ring -- no
 *                      extra launch, no collective, no cross-rank stream dependency;
 *   isb_counts_wait    enqueues a one-CTA kernel that waits (bounded by timeout_ms, default 2000) until all world * n_local
 *                      words of `epoch` (-1 = this rank's latest) have arrived, then writes counts int32 [world * n_local]
 *                      (rank-major, -1 = candidate overflow) and, if not NULL, offsets int64 = exclusive scan of
 *                      max(count, 0).  *status (device int32, may be NULL): 0 ok, 1 timed out, 2 the slot was overwritten
 *                      by a newer epoch (a rank ran more than `depth` calls ahead of this read);
 *   isb_counts_epoch   epoch of this rank's latest publishing call;  isb_counts_detach unmaps and frees.
 * All ranks must issue the same sequence of publishing calls.
 * ---------------------------------------------------------------------------------- */
ISB_API int isb_counts_export(isb_handle* h, int32_t n_local, int32_t world, int32_t depth, unsigned char handle_out[64]);
ISB_API int isb_counts_attach(isb_handle* h, int32_t rank, const unsigned char* handles);
ISB_API int isb_counts_wait(isb_handle* h, int64_t epoch, int32_t* counts, int64_t* offsets, int32_t* status,
                            int32_t timeout_ms, isb_stream stream);
ISB_API int64_t isb_counts_epoch(const isb_handle* h);
ISB_API int isb_counts_detach(isb_handle* h);

/* ------------------------------------------------------------------------------------
 * Detect + per-pixel CLASS MAP, for callers that draw / score a class image and never look at
 * the per-instance masks.  Replaces decode_nms (utils_bbox.py:170-243) TOGETHER WITH the
 * wrapper tail inst_model/yolact/yolact.py:162-170:
 *     masks_class = masks_sigmoid * (class_ids[None, None, :] + 1)
 *     masks_class = masks_class.reshape(-1, m)[arange(H*W), masks_arg.reshape(-1)].reshape(H, W)
 * i.e. class_map[y, x] = class_ids[masks_arg[y, x]] + 1 where that detection's mask is set at
 * (y, x), else 0.  bool [H,W,m] and int64 [H,W] are never written (the wrapper copies 30 MB
 * of them to the host per image, yolact.py:162); the map is uint8 [H,W].
 * Images without detections (counts[b] <= 0) get an all-zero map.  num_classes <= 255.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  float* boxes;          /* [B,max_det,4]                                               */
  float* scores;         /* [B,max_det]                                                 */
  int64_t* class_ids;    /* [B,max_det]                                                 */
  uint8_t* class_map;    /* [B,H,W]                                                     */
  int32_t* counts;       /* [B], as in isb_yolact_out                                   */
} isb_yolact_map_out;

ISB_API int isb_yolact_class_map(isb_handle* h, const isb_yolact_in* in, const isb_yolact_map_out* out, isb_stream stream);

/* Host-buffer version (same pipeline as isb_yolact_detect_host; D2H is 0.3 MB instead of 32 MB per image). */
ISB_API int isb_yolact_class_map_host(isb_handle* h, const isb_yolact_in* in, const isb_yolact_map_out* out, isb_stream stream);

/* ------------------------------------------------------------------------------------
 * Wrapper pre-processing, for a batch of same-sized pictures.
 * Replaces inst_model/yolact/yolact.py:137-148 and inst_model/Mask_RCNN/mask_rcnn.py:140-148:
 *     resize_image(image, (W_in, H_in))          utils/utils.py:19-22, PIL Image.resize(.., Image.BICUBIC)
 *     preprocess_input(np.array(.., float32))    utils/utils.py:32-36, (x - mean) / std in float64
 *     transpose(2, 0, 1)[None] -> FloatTensor    yolact.py:145-148
 * The resize follows Pillow's 8-bit fixed-point resampling (libImaging/Resample.c) bit for bit.
 *   images:  device uint8 [B, in_h, in_w, 3] (RGB, HWC);
 *   mean, std: 3 doubles each (utils.py:33-34: 123.68, 116.78, 103.94 / 58.40, 57.12, 57.38);
 *   resized: device uint8 [B, out_h, out_w, 3] = the PIL image resize_image returns, or NULL;
 *   input:   device float32 [B, 3, out_h, out_w] = the tensor the wrapper feeds the network, or NULL.
 * ---------------------------------------------------------------------------------- */
ISB_API int isb_preprocess(isb_handle* h, const uint8_t* images, int32_t batch, int32_t in_h, int32_t in_w, int32_t out_h,
                   int32_t out_w, const double* mean, const double* std, uint8_t* resized, float* input, isb_stream stream);

/* ------------------------------------------------------------------------------------
 * COCO run-length encoding of the instance masks on the device.
 * Replaces the per-instance host call inst_model/yolact/utils/utils_map.py:33-34
 *     pycocotools.mask.encode(np.asfortranarray(segmentation.astype(np.uint8)))
 * (driven by prep_metrics :56-62) -- third-party cocoapi maskApi.c rleEncode: column-major scan,
 * `counts` = lengths of the alternating runs starting with a (possibly empty) run of zeros.
 *   masks: the layout isb_yolact_detect writes (B slots, bool [H,W,m_b], m innermost);
 *   counts: device [B] m_b (as written by detect), or NULL and m_fixed for every image;
 *   rle_counts: flat device buffer of `capacity` uint32; detection (b, j) owns
 *     rle_counts[run_offset[b*max_det+j] .. + run_count[b*max_det+j]);  run_count 0 = no such detection;
 *   totals: device int64[2] = {boundaries, runs}; totals[1] > capacity means the buffer was too small
 *     (nothing is written past it) -- call again with a larger one.
 * ---------------------------------------------------------------------------------- */
ISB_API int isb_rle_encode(isb_handle* h, const uint8_t* masks, int64_t mask_slot_bytes, const int32_t* counts, int32_t m_fixed,
                           int32_t max_det, int32_t batch, int32_t hgt, int32_t wid, uint32_t* rle_counts, int64_t capacity,
                           int32_t* run_count, int64_t* run_offset, int64_t* totals, isb_stream stream);

/* HOST helper (no GPU work): the ASCII string pycocotools stores in rle['counts'] (maskApi.c rleToString) for one
 * instance's `n` counts.  Returns the length written, -1 on bad arguments, -2 if out_cap is too small. */
ISB_API int64_t isb_rle_to_string(const uint32_t* counts, int64_t n, char* out, int64_t out_cap);

/* ------------------------------------------------------------------------------------
 * Stage entry points (known-answer tests call these one by one; BBoxUtility's public
 * helper methods map onto them).
 * ---------------------------------------------------------------------------------- */

/* utils_bbox.py:11-25 decode_boxes(pred_box[n,4], anchors[n,4], variances) -> [n,4] */
ISB_API int isb_decode_boxes(isb_handle* h, const float* pred, const float* anchors, int64_t n, float var0, float var1,
                     float* out, isb_stream stream);

/* utils_bbox.py:27-48 jaccard: a [nb,na,4], b [nb,nbx,4] -> iou [nb,na,nbx] */
ISB_API int isb_jaccard(isb_handle* h, const float* a, const float* b, int32_t batch, int32_t na, int32_t nb, int32_t iscrowd,
                float* out, isb_stream stream);

/* utils_bbox.py:50-100 fast_non_max_suppression(box[n,4], scores[n,C], coef[n,K]).
 * Outputs sized for max_det rows; *count (device int32) receives the number of rows. */
ISB_API int isb_fast_nms(isb_handle* h, const float* box, const float* scores, const float* coef, int32_t n,
                 int32_t num_fg_classes, int32_t num_coef, float nms_iou, int32_t top_k, int32_t max_det,
                 float* out_box, float* out_score, int64_t* out_class, float* out_coef, int32_t* count,
                 isb_stream stream);

/* utils_bbox.py:102-140 traditional_non_max_suppression(box[n,4], class_scores[n,C], coef[n,K], max_scores[n]):
 * class = argmax over the C foreground scores, greedy per-class NMS (torchvision semantics, IoU > nms_iou
 * suppresses, unshifted boxes) on max_scores, concatenation ordered by score, first max_det rows.
 * Outputs sized for max_det rows; *count (device int32) receives the number of rows. */
ISB_API int isb_traditional_nms(isb_handle* h, const float* box, const float* class_scores, const float* coef,
                        const float* max_scores, int32_t n, int32_t num_fg_classes, int32_t num_coef, double nms_iou,
                        int32_t max_det, float* out_box, float* out_score, int64_t* out_class, float* out_coef,
                        int32_t* count, isb_stream stream);

/* utils_bbox.py:142-152 yolact_correct_boxes(boxes[m,4], (H,W)) -> [m,4] */
ISB_API int isb_correct_boxes(isb_handle* h, const float* boxes, int64_t m, int32_t out_h, int32_t out_w, float* out,
                      isb_stream stream);

/* utils_bbox.py:154-168 crop(masks[H,W,m] float32, boxes[m,4]) -> [H,W,m] float32 */
ISB_API int isb_crop(isb_handle* h, const float* masks, const float* boxes, int32_t hgt, int32_t wid, int32_t m, float* out,
             isb_stream stream);

/* utils_bbox.py:223 inner product only: proto[Ph,Pw,K] @ coef[m,K]^T -> logits [Ph,Pw,m] */
ISB_API int isb_mask_logits(isb_handle* h, const float* proto, const float* coef, int32_t proto_h, int32_t proto_w, int32_t m,
                    float* out, isb_stream stream);

/* utils_bbox.py:223 for a BATCH of pictures in one persistent launch on the tensor cores: proto [B,Ph,Pw,K], coef [B,m,K]
 * (m <= 128) -> out [B,Ph,Pw,m] = proto @ coef^T, or sigmoid of it (apply_sigmoid != 0: the whole of :223).  tcgen05.mma
 * kind::tf32 with every fp32 operand split exactly into three tf32 parts (six partial products, fp32 accumulation in tensor
 * memory: the result meets the 1e-5 bar), prototype tiles fetched by bulk TMA two deep, one bulk store per 128-pixel tile. */
ISB_API int isb_mask_logits_batch(isb_handle* h, const float* proto, const float* coef, int32_t batch, int32_t proto_h,
                                  int32_t proto_w, int32_t m, int32_t apply_sigmoid, float* out, isb_stream stream);

/* utils_bbox.py:223-241 for one image with m known on the host:
 * sigmoid(proto@coef^T) -> bilinear to (H,W) -> crop by boxes_px -> argmax, >0.5.
 * masks: bool [H,W,m] (uint8), masks_arg: int64 [H,W]. */
ISB_API int isb_mask_assemble(isb_handle* h, const float* proto, const float* coef, const float* boxes_px, int32_t proto_h,
                      int32_t proto_w, int32_t m, int32_t out_h, int32_t out_w, int64_t* masks_arg, uint8_t* masks,
                      isb_stream stream);

/* ------------------------------------------------------------------------------------
 * Mask R-CNN side (inst_model/Mask_RCNN/net/network_files/...)
 * ---------------------------------------------------------------------------------- */

/* det_utils.py:224-267 BoxCoder.decode_single: rel_codes [n,4k], boxes [n,4] -> out [n,4k];
 * weights (wx,wy,ww,wh) = (1,1,1,1) for the RPN (rpn_function.py:352), (10,10,5,5) for the box head
 * (roi_head.py:187); clip = log(1000/16) (det_utils.py:146). */
ISB_API int isb_box_decode(isb_handle* h, const float* rel_codes, const float* boxes, int64_t n, int32_t k, float wx, float wy,
                   float ww, float wh, float clip, float* out, isb_stream stream);

/* boxes.py:7-35 nms -> torch.ops.torchvision.nms: keep [n] int64 receives the kept indices in
 * decreasing score order, *count (device int32) their number. */
ISB_API int isb_nms(isb_handle* h, const float* boxes, const float* scores, int32_t n, double iou_threshold, int64_t* keep,
            int32_t* count, isb_stream stream);

/* boxes.py:38-83 batched_nms: idxs [n] int64 categories; coordinates are offset by
 * idx * (boxes.max() + 1) in float32 before the IoU, exactly as the reference does. */
ISB_API int isb_batched_nms(isb_handle* h, const float* boxes, const float* scores, const int64_t* idxs, int32_t n,
                    double iou_threshold, int64_t* keep, int32_t* count, isb_stream stream);

/* isb_nms without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as isb_nms, but no handle: the scratch of the pipeline (sort keys, ranks, boxes in score
 * order, the suppression bit matrix) is a file-scope `static __device__` array of 2 260 992 bytes that every GPU which
 * loads the library holds until the process exits.  Runs on the CURRENT device and on `stream`; calls no cudaMalloc,
 * cudaFree or other allocating routine (the first call on a device looks the array's address up once with
 * cudaGetSymbolAddress -- make that first call outside a graph capture).
 *   n > ISB_NMS_STATIC_MAX  ->  ISB_ERR_LIMIT (use isb_nms, whose workspace grows).
 *   All calls on one GPU share the array, which keeps the previous call's contents (a call clears what it needs at zero
 *   on `stream`); calls that OVERLAP on the GPU -- different streams, different host threads -- race on it.
 * `err` (may be NULL) receives the error text, at most err_len bytes including the terminator. */
#define ISB_NMS_STATIC_MAX 4096
ISB_API int isb_nms_static(const float* boxes, const float* scores, int32_t n, double iou_threshold, int64_t* keep,
                           int32_t* count, isb_stream stream, char* err, int32_t err_len);

/* torchvision MultiScaleRoIAlign.forward (constructed at faster_rcnn_framework.py:326-329 and
 * mask_rcnn.py:150, called at roi_head.py:491,540). */
typedef struct {
  int32_t num_levels;       /* 1..8 feature maps, finest first                              */
  const float* feat[8];     /* [N,C,H_l,W_l] float32                                         */
  int32_t feat_h[8];
  int32_t feat_w[8];
  float scale[8];           /* spatial scale of each level, 2^-k (poolers.py _infer_scale)   */
  int32_t k_min, k_max;     /* LevelMapper bounds: -log2(scale[0]), -log2(scale[last])       */
  int32_t canonical_scale;  /* 224 */
  int32_t canonical_level;  /* 4   */
  int32_t channels;
  int32_t out_h, out_w;
  int32_t sampling_ratio;   /* > 0 (the reference always passes 2)                            */
} isb_roi_align_cfg;
/* boxes [K,4] in image coordinates, batch_idx [K] int32 image index of each box (NULL = all 0),
 * out [K,C,out_h,out_w]. */
ISB_API int isb_ms_roi_align(isb_handle* h, const isb_roi_align_cfg* cfg, const float* boxes, const int32_t* batch_idx,
                     int32_t num_rois, float* out, isb_stream stream);

/* roi_head.py:59-91 maskrcnn_inference: logits [m,C,S,S], labels [m] int64 -> out [m,1,S,S] =
 * sigmoid(logits[i, labels[i]]). */
ISB_API int isb_mask_select_sigmoid(isb_handle* h, const float* logits, const int64_t* labels, int32_t m, int32_t num_classes,
                            int32_t mask_size, float* out, isb_stream stream);

/* transform.py:462-484 resize_boxes: boxes [m,4] scaled by new/orig per axis. */
ISB_API int isb_resize_boxes(isb_handle* h, const float* boxes, int64_t m, int32_t orig_h, int32_t orig_w, int32_t new_h,
                     int32_t new_w, float* out, isb_stream stream);

/* transform.py:212-230 paste_masks_in_image: masks [m,1,S,S] probabilities, boxes [m,4] ->
 * out float32 [m,1,im_h,im_w]. */
ISB_API int isb_paste_masks(isb_handle* h, const float* masks, const float* boxes, int32_t m, int32_t mask_size,
                    int32_t padding, int32_t im_h, int32_t im_w, float* out, isb_stream stream);

/* GeneralizedRCNNTransform.forward for inference, one image (transform.py:436-459): normalize :264-270, resize
 * :284-315 (F.interpolate bilinear, align_corners=False, to the size the host computed with the reference's rule,
 * transform.py:88-110) and the image's slot of batch_images :352-386 in one pass.
 *   image: device float32 [3, in_h, in_w];  mean, std: 3 floats each (host);
 *   canvas: device float32 [3, canvas_h, canvas_w], this image's slot of the batch: the resized image at the top-left,
 *   zeros elsewhere (every element is written). */
ISB_API int isb_rcnn_transform(isb_handle* h, const float* image, int32_t in_h, int32_t in_w, int32_t out_h, int32_t out_w,
                       const float* mean, const float* std, float* canvas, int32_t canvas_h, int32_t canvas_w,
                       isb_stream stream);

/* Wrapper tail inst_model/Mask_RCNN/mask_rcnn.py:174-198 fused with paste_masks_in_image (transform.py:212-230):
 * keep detections with score > confidence, paste their SxS probability masks into a (paste_h, paste_w) frame,
 * F.interpolate(bilinear, align_corners=False) to (out_h, out_w), arg-max over the kept detections, > 0.5, and
 *     class_map[y, x] = labels[kept[arg]] + 1 where that mask is set, else 0          (uint8 [out_h, out_w]).
 * The [m,1,Hp,Wp] pasted floats and the [H,W,m] resized masks are never materialised.  masks_arg (int64 [H,W],
 * index into the KEPT detections) and kept_count (int32) are optional (NULL = not written). */
ISB_API int isb_paste_class_map(isb_handle* h, const float* masks, const float* boxes, const int64_t* labels,
                                const float* scores, int32_t m, int32_t mask_size, int32_t padding, int32_t paste_h,
                                int32_t paste_w, int32_t out_h, int32_t out_w, float confidence, uint8_t* class_map,
                                int64_t* masks_arg, int32_t* kept_count, isb_stream stream);

/* rpn_function.py:434-458 _get_top_n_idx + :460-529 filter_proposals, for ONE image:
 * proposals [A,4] (already decoded, rpn_function.py:624), objectness [A] logits, anchors_per_level
 * (HOST array of num_levels ints).  Per-level top pre_nms_top_n by logit, sigmoid, clip to the image,
 * drop boxes with a side < min_size and scores < score_thresh, batched NMS per level, keep
 * post_nms_top_n.  out_boxes [post,4], out_scores [post], *out_count (device int32). */
ISB_API int isb_rpn_filter(isb_handle* h, const float* proposals, const float* objectness, const int32_t* anchors_per_level,
                   int32_t num_levels, int32_t img_h, int32_t img_w, int32_t pre_nms_top_n, int32_t post_nms_top_n,
                   double nms_thresh, float score_thresh, float min_size, float* out_boxes, float* out_scores,
                   int32_t* out_count, isb_stream stream);

/* roi_head.py:366-457 postprocess_detections for ONE image: class_logits [R,C] (softmax applied here,
 * :401) OR class_probs [R,C] already softmaxed (pass the other as NULL), box_regression [R,4C],
 * proposals [R,4].  Decode (weights wx..wh, clip), clip to the image, drop background, score >
 * score_thresh, drop boxes with a side < 1, batched NMS per class, keep detections_per_img.
 * out_boxes [D,4], out_scores [D], out_labels [D] int64, *out_count (device int32; -1 if class_probs
 * rows do not sum to 1 and the candidate bound R*floor(1/score_thresh) was exceeded). */
ISB_API int isb_box_postprocess(isb_handle* h, const float* class_logits, const float* class_probs, const float* box_regression,
                        const float* proposals, int32_t num_rois, int32_t num_classes, int32_t img_h, int32_t img_w,
                        float wx, float wy, float ww, float wh, float clip, float score_thresh, double nms_thresh,
                        int32_t detections_per_img, float* out_boxes, float* out_scores, int64_t* out_labels,
                        int32_t* out_count, isb_stream stream);

/* isb_box_postprocess for a pipeline that never reads a count on the host (roi_head.py:366-457 on the PADDED proposals of
 * one image): the rows are [max_rois, ...] as isb_rpn_filter leaves them (post_nms_top_n rows, valid ones first), and
 * *num_rois_dev (device memory, e.g. isb_rpn_filter's out_count) says how many are valid; rows past it are ignored.
 * Results are identical to isb_box_postprocess(num_rois = *num_rois_dev). */
ISB_API int isb_box_postprocess_rows(isb_handle* h, const float* class_logits, const float* class_probs, const float* box_regression,
                             const float* proposals, int32_t max_rois, const int32_t* num_rois_dev, int32_t num_classes,
                             int32_t img_h, int32_t img_w, float wx, float wy, float ww, float wh, float clip,
                             float score_thresh, double nms_thresh, int32_t detections_per_img, float* out_boxes,
                             float* out_scores, int64_t* out_labels, int32_t* out_count, isb_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* ISB_H_ */
