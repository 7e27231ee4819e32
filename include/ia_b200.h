/*
 * ia_b200.h — C ABI of libia_b200.so: the B200 (sm_100a) implementation of
 * Inference Arena's `shared.processing` hot path.
 *
 * Conventions
 *   - Plain pointers and sizes only.  Pointers named *_dev are device pointers on the
 *     CURRENT CUDA device; `stream` is a cudaStream_t passed as void* (NULL = default
 *     stream).  Calls are stream-ordered, never synchronise and never allocate unless
 *     the comment says otherwise.  Safe to call concurrently on different streams as
 *     long as the buffers (including `workspace`) are distinct.
 *   - Return value: 0 = success; > 0 = a cudaError_t from the launch; < 0 = IA_ERR_*.
 *   - All float arithmetic that the reference does in numpy float32 is done in IEEE
 *     float32 with round-to-nearest and no FMA contraction; OpenCV's 8-bit INTER_LINEAR
 *     is reproduced in its own 11-bit fixed point, so results are bit-identical to the
 *     reference's CPU path.
 *   - Every declaration cites the reference interface it replaces
 *     (paths relative to the reference repository root).
 */
#ifndef IA_B200_H
#define IA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IA_OK 0
#define IA_ERR_INVALID_ARGUMENT (-1)
#define IA_ERR_WORKSPACE_TOO_SMALL (-2)
#define IA_ERR_UNSUPPORTED (-3)
#define IA_ERR_CAPACITY (-4)

#define IA_HEAD_LAYOUT_V5 0 /* (N, A, 5+C): cx,cy,w,h,obj,cls...   curator.py:289-307 */
#define IA_HEAD_LAYOUT_V8 1 /* (N, 4+C, A): channel-major, no obj   postprocess.py:33-41 */

/* Library / build identification ("ia_b200 <version> sm_100a"). */
const char* ia_version(void);
/* Human-readable text for a return value of any function below. */
const char* ia_error_string(int code);

/* ---------------------------------------------------------------------------------
 * Letterbox geometry — host scalars of letterbox(), src/shared/processing/
 * transforms.py:150-157,174-175: scale = min(T/h, T/w) in float64, new = int(dim*scale)
 * (truncation), pad = (T - new) // 2.  Pure host arithmetic, no CUDA.
 * ------------------------------------------------------------------------------- */
typedef struct ia_letterbox_geom {
  int32_t src_h, src_w, target;
  int32_t new_h, new_w;
  int32_t pad_h, pad_w;
  double scale;
} ia_letterbox_geom;
int ia_letterbox_geometry(int src_h, int src_w, int target, ia_letterbox_geom* out);

/* ---------------------------------------------------------------------------------
 * K1 — fused letterbox + /divisor + HWC->CHW for a batch of equally-sized frames.
 * Replaces YOLOPreprocessor.preprocess, src/shared/processing/yolo_preprocess.py:128-174
 * (letterbox transforms.py:118-180, astype(f32)/255.0 :158, transpose :161, contiguous
 * :167) applied to each of the n frames.
 *   frames_dev  uint8 [n, src_h, src_w, 3]
 *   out_dev     float [n, 3, target, target]
 * ------------------------------------------------------------------------------- */
int ia_yolo_preprocess(const uint8_t* frames_dev, int n, int src_h, int src_w, int target,
                       float divisor, float* out_dev, void* stream);

/* Ragged batches: YOLOPreprocessor.preprocess (yolo_preprocess.py:128-174) for
 * n frames of DIFFERENT sizes in one call — what a service sees with real COCO frames.
 *   frames_host  HOST array of n descriptors (device pointer + height + width per frame)
 *   out_dev      float [n, 3, target, target]; image i comes from frames_host[i]
 * Frames are grouped by shape on the host; the frame pointers and each shape's geometry travel
 * in the kernel parameters (no device scratch, no copies, no synchronisation).  Up to 32 shapes
 * and 256 frames are ONE launch (a CTA looks its shape up from its block index); beyond that,
 * one launch per shape. */
typedef struct ia_frame_desc {
  const uint8_t* data_dev; /* uint8 [h, w, 3], contiguous */
  int32_t h, w;
} ia_frame_desc;
int ia_yolo_preprocess_ragged(const ia_frame_desc* frames_host, int n, int target, float divisor,
                              float* out_dev, void* stream);

/* letterbox() itself, transforms.py:118-180: uint8 [n, target, target, 3] HWC with the
 * padding colour `pad_rgb` (LETTERBOX_COLOR, transforms.py:37). */
int ia_letterbox_u8(const uint8_t* frames_dev, int n, int src_h, int src_w, int target,
                    const uint8_t pad_rgb[3], uint8_t* out_dev, void* stream);

/* ---------------------------------------------------------------------------------
 * K2 + K3 — YOLO head decode, confidence filter + stream compaction, per-class greedy
 * NMS and row assembly, for a batch of n images, entirely on the device.
 * Replaces, per image: DetectionCounter._parse_yolov5_raw_output, src/shared/data/
 * curator.py:289-307 (conf = obj * max(cls), class = argmax) or the v8 parse,
 * architectures/monolithic/app/postprocess.py:33-41; apply_nms, postprocess.py:76-160
 * (== curator.py:323-400); and the output assembly postprocess.py:52-73.
 *   head_dev         float, layout per `layout` (IA_HEAD_LAYOUT_*), n images
 *   max_candidates   capacity of the per-image candidate list (anchors passing the
 *                    confidence filter); `anchors` is always sufficient
 *   max_det          rows available per image in dets_dev / keep_dev
 *   dets_dev         float [n, max_det, 6]  rows [x1,y1,x2,y2,conf,class], class-ascending
 *                    then confidence-descending — the order apply_nms returns
 *   keep_dev         int32 [n, max_det]     original anchor index of each row
 *   count_dev        int32 [n]              detections kept per image
 *   status_dev       int32 [n]              0 ok; bit0 = candidate list overflowed
 *                    (> max_candidates), bit1 = more than max_det kept.  Rows are exact
 *                    whenever the status is 0.
 *   workspace        device scratch of at least ia_decode_nms_workspace_bytes(...)
 * Equal confidences inside one class (order undefined in the reference, postprocess.py:
 * 129) are ordered larger-anchor-index first.
 * ------------------------------------------------------------------------------- */
size_t ia_decode_nms_workspace_bytes(int n, int anchors, int max_candidates);
int ia_decode_nms(const float* head_dev, int layout, int n, int anchors, int num_classes,
                  float conf_threshold, float iou_threshold, int max_candidates, int max_det,
                  float* dets_dev, int32_t* keep_dev, int32_t* count_dev, int32_t* status_dev,
                  void* workspace_dev, size_t workspace_bytes, void* stream);

/* ia_decode_nms without a workspace argument and without any allocation.  This entry point
 * has no counterpart in the reference; it exists for callers that cannot obtain device
 * memory when they call: processes whose HBM is wholly owned by another allocator (an
 * inference server's pre-reserved arena, an exhausted MPS/MIG quota) and latency-critical
 * threads that must stay out of cudaMalloc/cudaFree, which lock and synchronise the device.
 * The candidate lists live in a fixed 524 544-byte array inside the library (held on every
 * GPU that loads it until the process exits), which bounds one call to
 * IA_STATIC_MAX_IMAGES images and IA_STATIC_MAX_CANDIDATES candidates per image:
 * n above the first, or an effective max_candidates (max_candidates, or `anchors` when it
 * is < 1 or > anchors) above the second, returns IA_ERR_CAPACITY and launches nothing.
 * Inputs, outputs and results are exactly those of ia_decode_nms.
 *   - The first call on each device looks the array's address up (a one-time host-side
 *     set-up); make that call outside of, and before, any stream capture.
 *   - All calls on a device share the one array and it keeps the previous call's contents;
 *     the counters are re-zeroed in stream order by every call.  Two calls whose kernels
 *     can overlap on the GPU (different streams, no dependency) race on it: serialise them. */
#define IA_STATIC_MAX_IMAGES 64
#define IA_STATIC_MAX_CANDIDATES 1024
int ia_decode_nms_static(const float* head_dev, int layout, int n, int anchors, int num_classes,
                         float conf_threshold, float iou_threshold, int max_candidates,
                         int max_det, float* dets_dev, int32_t* keep_dev, int32_t* count_dev,
                         int32_t* status_dev, void* stream);

/* apply_nms() on already-decoded arrays, postprocess.py:76-160: boxes [m,4] cx,cy,w,h,
 * scores [m], class ids int32 [m] -> keep_dev int32 [max_det] (original indices),
 * count_dev [1], status_dev [1].  Workspace: ia_decode_nms_workspace_bytes(1, m, m).
 * Class ids must lie in [0, 1023): the sort key holds a class in 10 bits (the Python wrapper
 * renumbers arbitrary ids densely, as np.unique does in the reference, before it calls this). */
int ia_nms(const float* boxes_dev, const float* scores_dev, const int32_t* class_ids_dev, int m,
           float conf_threshold, float iou_threshold, int max_det, int32_t* keep_dev,
           int32_t* count_dev, int32_t* status_dev, void* workspace_dev, size_t workspace_bytes,
           void* stream);

/* ---------------------------------------------------------------------------------
 * scale_boxes(), src/shared/processing/transforms.py:183-230 (also
 * YOLOPreprocessResult.scale_boxes_to_original, yolo_preprocess.py:60-69): un-pad,
 * divide by float32(scale), clip to the original frame; columns >= 4 are copied.
 *   boxes_dev float [m, cols] -> out_dev float [m, cols]   (cols >= 4)
 * ------------------------------------------------------------------------------- */
int ia_scale_boxes(const float* boxes_dev, int m, int cols, float scale, int pad_w, int pad_h,
                   int orig_h, int orig_w, float* out_dev, void* stream);

/* The same for float64 boxes (numpy keeps the array's precision: all arithmetic in binary64). */
int ia_scale_boxes_f64(const double* boxes_dev, int m, int cols, double scale, int pad_w, int pad_h,
                       int orig_h, int orig_w, double* out_dev, void* stream);

/* ---------------------------------------------------------------------------------
 * K4 — fan-out: for every kept detection, scale_boxes -> extract_crop -> bilinear resize
 * to out_size x out_size -> ImageNet normalise -> CHW, reading the NMS output on the
 * device.  Replaces the per-detection loop of architectures/monolithic/app/inference.py:
 * 188-193: scale_boxes_to_original (yolo_preprocess.py:60-69), extract_crop
 * (src/shared/processing/mobilenet_preprocess.py:236-269: truncating int(), clamp,
 * empty -> 1x1 zeros), MobileNetPreprocessor.preprocess (mobilenet_preprocess.py:114-160)
 * and imagenet_normalize (transforms.py:238-272).
 *   frames_dev      uint8 [n, src_h, src_w, 3]  original-resolution frames
 *   dets_dev        float [n, max_det, 6]       letterbox-space rows from ia_decode_nms
 *   count_dev       int32 [n]
 *   scale,pad_*     letterbox geometry of the frames (ia_letterbox_geometry)
 *   mean,std        float[3] each (experiment.yaml:244-245)
 *   crops_dev       float [roi_capacity, 3, out_size, out_size]; ROI r of image i, k-th
 *                   detection, is r = roi_offsets[i] + k
 *   roi_offsets_dev int32 [n+1] exclusive scan of min(count, max_det); [n] = total ROIs
 *   rois_dev        int32 [roi_capacity, 8] scratch/out: {image, x1, y1, x2, y2, empty,0,0}
 *                   (the integer crop rectangle extract_crop would use).  While the call runs,
 *                   the two padding words of record 0 are its work counter (tiles are handed
 *                   out one at a time so that crops of very different size finish together);
 *                   they are 0 again when it completes.
 * ROIs beyond roi_capacity are not written (roi_offsets still reports the true total).
 * Two staged rows of the widest possible crop must fit the SM's shared memory next to the
 * look-up table: frames wider than about 10 800 pixels (10 816 on B200) return
 * IA_ERR_UNSUPPORTED before anything is launched; the same bound applies to max_src_w of the
 * ragged form and to src_w of ia_resize_normalize_rois.
 * ------------------------------------------------------------------------------- */
int ia_crop_resize_normalize(const uint8_t* frames_dev, int n, int src_h, int src_w,
                             const float* dets_dev, const int32_t* count_dev, int max_det,
                             float scale, int pad_w, int pad_h, int out_size,
                             const float mean[3], const float std[3], float* crops_dev,
                             int roi_capacity, int32_t* roi_offsets_dev, int32_t* rois_dev,
                             void* stream);

/* K4 on a ragged batch: as ia_crop_resize_normalize, but every image has its
 * own size.  frames_dev is a DEVICE array of n descriptors; the letterbox geometry of each
 * image (scale, padding for `letterbox_target`) is derived on the device in the reference's
 * float64 arithmetic (transforms.py:153-157), max_src_w bounds the frame widths (it sizes the
 * shared-memory staging). */
int ia_crop_resize_normalize_ragged(const ia_frame_desc* frames_dev, int n, int max_src_w,
                                    const float* dets_dev, const int32_t* count_dev, int max_det,
                                    int letterbox_target, int out_size, const float mean[3],
                                    const float std[3], float* crops_dev, int roi_capacity,
                                    int32_t* roi_offsets_dev, int32_t* rois_dev, void* stream);

/* The same resize+normalise+CHW on explicit rectangles (no detection input):
 * rois_dev int32 [num_rois, 8] as above.  With one ROI covering a whole image this is
 * MobileNetPreprocessor.preprocess (mobilenet_preprocess.py:114-160) on a uint8 crop. */
int ia_resize_normalize_rois(const uint8_t* frames_dev, int n, int src_h, int src_w,
                             const int32_t* rois_dev, int num_rois, int out_size,
                             const float mean[3], const float std[3], float* crops_dev,
                             void* stream);

/* extract_crop()'s rectangle, src/shared/processing/mobilenet_preprocess.py:256-267, for m
 * boxes already in original-frame coordinates: int() truncation toward zero, clamp to the
 * frame, empty test.  boxes_dev float [m, cols>=4] -> rois_dev int32 [m, 8]
 * {image_index, x1, y1, x2, y2, empty, 0, 0} (the record ia_resize_normalize_rois takes). */
int ia_extract_crop_rects(const float* boxes_dev, int m, int cols, int image_index, int src_h,
                          int src_w, int32_t* rois_dev, void* stream);

/* imagenet_normalize() on uint8 HWC, transforms.py:238-272 -> float HWC (same layout). */
int ia_imagenet_normalize_u8(const uint8_t* image_dev, size_t pixels, const float mean[3],
                             const float std[3], float* out_dev, void* stream);

/* ---------------------------------------------------------------------------------
 * float32-input branch.  MobileNetPreprocessor.preprocess on a float32 crop,
 * src/shared/processing/mobilenet_preprocess.py:114-160 with :204-206 (float32 accepted) and
 * imagenet_normalize's float branch, transforms.py:264-267 (divide by 255 only when the
 * maximum exceeds 1).  crop_dev float [h, w, 3] -> out_chw_dev float [3, out_size, out_size].
 * cv2.resize on float data depends on the OpenCV build, so the resize comes in two flavours, each
 * bit-identical to its build: IA_F32_RESIZE_IPP — the opencv-python wheels the reference pins
 * (uv.lock:2549-2550) hand float resizes to Intel IPP; IA_F32_RESIZE_OPENCV — OpenCV's own
 * generic float INTER_LINEAR, what cv2 computes with IPP off or absent.  (On [0,1] data the two
 * differ by up to ~3e-5.)  ia_resize_normalize_f32 is the OPENCV flavour.
 * scratch_dev: 16-byte aligned, at least ia_f32_scratch_bytes(out_size) bytes.
 * ------------------------------------------------------------------------------- */
#define IA_F32_RESIZE_OPENCV 0
#define IA_F32_RESIZE_IPP 1
size_t ia_f32_scratch_bytes(int out_size);
int ia_resize_normalize_f32(const float* crop_dev, int h, int w, int out_size, const float mean[3],
                            const float std[3], float* out_chw_dev, void* scratch_dev,
                            size_t scratch_bytes, void* stream);
int ia_resize_normalize_f32_as(const float* crop_dev, int h, int w, int out_size, int flavour,
                               const float mean[3], const float std[3], float* out_chw_dev,
                               void* scratch_dev, size_t scratch_bytes, void* stream);

/* imagenet_normalize() on float32 HWC, transforms.py:264-270 -> float HWC.  scratch_dev: 8 bytes. */
int ia_imagenet_normalize_f32(const float* image_dev, size_t pixels, const float mean[3],
                              const float std[3], float* out_dev, void* scratch_dev,
                              size_t scratch_bytes, void* stream);

/* ---------------------------------------------------------------------------------
 * Classifier logits (the step after the path).
 *   softmax != 0: ClassificationInference.classify,
 *     architectures/microservices/classification/app/inference.py:112-132 --
 *     probs = exp(s - max(s)) / sum, the k largest in argsort(probs)[::-1] order
 *     (low_index_first = 0: among equal values the larger index comes first, what a
 *     stable ascending sort reversed gives; the reference's unstable sort leaves ties open).
 *   softmax == 0, k = 1, low_index_first = 1: np.argmax + the raw value,
 *     architectures/monolithic/app/inference.py:201-202, triton/gateway/app/pipeline.py:181-182.
 * logits_dev float [rows, classes] -> ids_dev int32 [rows, k], values_dev float [rows, k],
 * status_dev int32 [rows] (may be NULL; bit 0 = the row holds a NaN, every probability is
 * then NaN as in numpy).  NaN orders above +inf (numpy's sort order).  1 <= k <= classes;
 * classes * 4 bytes must fit one CTA's shared memory (<= 51200 classes), else IA_ERR_UNSUPPORTED.
 * ------------------------------------------------------------------------------- */
int ia_classify_topk(const float* logits_dev, int rows, int classes, int k, int softmax,
                     int low_index_first, int32_t* ids_dev, float* values_dev,
                     int32_t* status_dev, void* stream);

/* ---------------------------------------------------------------------------------
 * JPEG decode on the device (the step before the path).  Replaces
 * load_image_from_bytes, src/shared/processing/transforms.py:77-110 —
 * cv2.imdecode(buf, IMREAD_COLOR) + cvtColor(BGR2RGB) — for baseline / extended-sequential
 * Huffman JPEGs (8-bit; 1 component, or 3 components YCbCr in one interleaved scan;
 * sampling 4:4:4, 4:2:2, 4:2:0, 4:4:0, 4:1:1; restart intervals; EXIF orientation).
 * Bit-identical to libjpeg-turbo's default decoder as OpenCV drives it (islow IDCT, fancy
 * upsampling, 16-bit fixed-point YCbCr->RGB), output directly in RGB.  Progressive,
 * arithmetic-coded, 12-bit, CMYK and multi-scan files return IA_ERR_UNSUPPORTED from
 * ia_jpeg_parse (they stay on the host decoder).  Only the compressed bytes cross PCIe.
 *
 *   1. ia_jpeg_parse       host: headers of one file -> ia_jpeg_job (no CUDA)
 *   2. ia_jpeg_workspace_bytes / ia_jpeg_plan   host: carve the device workspace, fill the
 *                          device pointers of every job
 *   3. caller copies the file bytes and the job array to the device
 *   4. ia_jpeg_decode      device: unstuff -> parallel Huffman -> IDCT -> upsample+colour
 * ------------------------------------------------------------------------------- */
#define IA_JPEG_MAX_COMPONENTS 3
#define IA_JPEG_MAX_BLOCKS_PER_MCU 10

typedef struct ia_jpeg_job {
  /* --- filled by ia_jpeg_parse --- */
  int32_t width, height;         /* as coded */
  int32_t out_width, out_height; /* after the EXIF orientation (what imdecode returns) */
  int32_t orientation;           /* EXIF tag 0x0112, 1..8 */
  int32_t ncomp;
  int32_t h[IA_JPEG_MAX_COMPONENTS], v[IA_JPEG_MAX_COMPONENTS]; /* sampling factors */
  int32_t hmax, vmax;
  /* per component: real (subsampled) size, row pitch of its plane, upsampling factors */
  int32_t comp_w[IA_JPEG_MAX_COMPONENTS], comp_h[IA_JPEG_MAX_COMPONENTS];
  int32_t comp_pitch[IA_JPEG_MAX_COMPONENTS];
  int32_t hexp[IA_JPEG_MAX_COMPONENTS], vexp[IA_JPEG_MAX_COMPONENTS];
  int32_t mcux, mcuy, blocks_per_mcu, total_blocks;
  int32_t restart_interval; /* MCUs, 0 = none */
  int32_t num_segments;     /* restart segments = ceil(mcus / interval), 1 without */
  uint32_t scan_offset;     /* first entropy-coded byte within the file */
  uint32_t file_bytes;
  uint8_t block_comp[IA_JPEG_MAX_BLOCKS_PER_MCU + 2]; /* component of block k of an MCU */
  uint8_t block_bx[IA_JPEG_MAX_BLOCKS_PER_MCU + 2], block_by[IA_JPEG_MAX_BLOCKS_PER_MCU + 2];
  uint8_t comp_dc[IA_JPEG_MAX_COMPONENTS + 1], comp_ac[IA_JPEG_MAX_COMPONENTS + 1]; /* table ids */
  uint16_t quant[IA_JPEG_MAX_COMPONENTS][64]; /* natural order */
  uint8_t dc_counts[4][16], dc_vals[4][16];
  uint8_t ac_counts[4][16], ac_vals[4][256];
  /* --- filled by ia_jpeg_plan (device pointers) --- */
  const uint8_t* file_dev; /* the whole file; read in aligned 16-byte units, so the bytes up to the next
                              16-byte boundary after its end must be readable (any cudaMalloc'ed buffer is) */
  uint8_t* clean_dev;      /* entropy-coded bytes without stuffing / RST markers */
  uint32_t* segment_dev;   /* [0] clean bytes, [1] markers found inside the data, [2] flags (1 = no marker
                              behind the data), [3..] clean byte offsets of the regions between the
                              markers, then one marker code / restart segment per region */
  int16_t* coef_dev;       /* [total_blocks, 64] natural order, DC absolute, not dequantised */
  uint8_t* plane_dev[IA_JPEG_MAX_COMPONENTS]; /* per component, padded to whole MCUs */
  uint8_t* out_dev;        /* uint8 [out_height, out_width, 3] RGB */
  uint32_t* chain_dev;     /* per 64 KB chunk of the clean stream: hand-over record (zeroed per call) */
  uint32_t* tile_dev;      /* per 8 KB tile of the scan: unstuffing counts */
  uint32_t* ticket_dev;    /* one counter per call (shared by all jobs of the batch) */
} ia_jpeg_job;

/* sizeof(ia_jpeg_job) as the library was compiled, for bindings that mirror the struct. */
size_t ia_jpeg_job_size(void);
/* Host only.  IA_OK, IA_ERR_INVALID_ARGUMENT (whatever makes libjpeg give up on the headers,
 * so that cv2.imdecode returns None and the reference raises "Failed to decode image from
 * bytes": no FF D8 FF signature, truncated or mis-sized segments, unknown markers, a second
 * SOI / SOF, sampling factors outside 1..4, missing quantisation tables, Huffman tables whose
 * codes overflow or whose DC symbols exceed 15, more than 10 blocks per MCU) or
 * IA_ERR_UNSUPPORTED.  Huffman table ids 0 and 1 default to the standard tables of T.81
 * Annex K when the file defines none, as libjpeg does for Motion-JPEG frames. */
int ia_jpeg_parse(const uint8_t* file_host, size_t file_bytes, ia_jpeg_job* job);
/* Device scratch for n parsed jobs (clean streams, coefficients, planes). */
size_t ia_jpeg_workspace_bytes(const ia_jpeg_job* jobs_host, int n);
/* Host only: points job i at file_dev[i] / out_dev[i] and at its slices of the workspace
 * (256-byte aligned base).  IA_ERR_WORKSPACE_TOO_SMALL if it does not fit. */
int ia_jpeg_plan(ia_jpeg_job* jobs_host, int n, const uint8_t* const* file_dev,
                 uint8_t* const* out_dev, void* workspace_dev, size_t workspace_bytes);
/* jobs_dev: the planned job array copied to device memory by the caller (jobs_host is read
 * for launch geometry only).  status_dev int32 [n]: 0 = the pixels are libjpeg's (one known
 * exception, damaged data only: an 8x8 block whose coefficients drive the 16-bit IDCT lanes
 * to their limit);
 * bit 0 = a damaged stream (too few or too many blocks, restart markers out of place): it is
 * decoded and concealed the way libjpeg does it — cv2.imdecode warns "Corrupt JPEG data" and
 * returns these very pixels;
 * bit 1 = in addition the data simply ends, without a marker behind it, before the last
 * block — there libjpeg runs out of input, cv2.imdecode returns None and the reference
 * raises "Failed to decode image from bytes" (shared/processing/jpeg.py raises it too). */
int ia_jpeg_decode(const ia_jpeg_job* jobs_host, const ia_jpeg_job* jobs_dev, int n,
                   int32_t* status_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IA_B200_H */
