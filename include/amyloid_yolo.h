/*
 * amyloid_yolo.h — C ABI of the B200 (sm_100a) tile-detection engine.
 *
 * The reference (keiserlab/amyloid-yolo-paper) is pure Python on PyTorch and has
 * no FFI of its own; its boundary for this path is the Python API
 *   Darknet(config_path).forward(x)              models.py:228-255
 *   non_max_suppression(pred, conf, nms)         utils/utils.py:235-273
 *   rescale_boxes(boxes, dim, shape)             utils/utils.py:36-50
 *   mergeDetections(detections)                  core.py:366-423
 * The entry points below are what a ctypes shim bound to those functions calls
 * (see INTEGRATION.md for the stub a maintainer would add to the reference).
 *
 * Conventions
 *   - every function returns 0 on success, a negative AY_E* code on failure;
 *     ay_last_error() returns a thread-local message for the last failure;
 *   - nothing throws across the boundary; the caller owns every buffer;
 *   - pointers are DEVICE pointers unless the name ends in _h (host);
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it
 *     unless stated otherwise;
 *   - one ay_ctx per GPU; a ctx is not thread-safe, distinct ctxs are independent.
 */
#ifndef AMYLOID_YOLO_H_
#define AMYLOID_YOLO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AY_OK            0
#define AY_EINVAL       -1   /* bad argument / unsupported configuration      */
#define AY_ECUDA        -2   /* a CUDA runtime or driver call failed          */
#define AY_ENOMEM       -3
#define AY_ESTATE       -4   /* call sequence violated (e.g. forward before finalize) */
#define AY_ECAPACITY    -5   /* a per-tile capacity was exceeded; see message */

typedef struct ay_ctx ay_ctx;

/* layer kinds, one per cfg block type handled by models.py:16-83 */
enum { AY_CONV = 0, AY_MAXPOOL = 1, AY_UPSAMPLE = 2, AY_ROUTE = 3, AY_SHORTCUT = 4, AY_YOLO = 5,
       AY_LINEAR = 6 };      /* fully connected OUTPUT layer over the flattened previous block (core.py:199-205, the
                              * CAA filter network): `filters` outputs; a network ends in [yolo] blocks or in one
                              * AY_LINEAR, whose [B, filters] fp32 result is what ay_forward writes to out_decoded */

/* input kinds accepted by ay_forward */
enum { AY_IN_F32_NCHW = 0,    /* [B,3,S,S] float, already /255 and resized (Darknet.forward contract) */
       AY_IN_U8_HWC   = 1 };  /* [B,H,W,3] uint8 tiles; ToTensor + nearest Resize (utils/transforms.py:89-111) fused into the first conv */

/* activation storage type: AY_BF16 / AY_FP16 run the tcgen05 tensor-core path (2e-2 bar on the raw heads);
 * AY_F32 is the accuracy mode (1e-4 bar) — fp32 storage, one plain kernel per cfg block, no fusion; its convolutions
 * run on the tensor cores as a 3-pass tf32 split (A_hi W_hi + A_lo W_hi + A_hi W_lo, fp32 accumulation in TMEM) where
 * the channel counts are multiples of 32 and as an FFMA kernel on the CUDA cores elsewhere (first layer, heads). */
enum { AY_BF16 = 0, AY_FP16 = 1, AY_F32 = 2 };

typedef struct ay_layer_desc {
  int32_t type;             /* AY_CONV ...                                              */
  int32_t filters;          /* conv: output channels                                    */
  int32_t size;             /* conv / maxpool: kernel size (conv pad = (size-1)/2)      */
  int32_t stride;           /* conv / maxpool / upsample                                */
  int32_t leaky;            /* conv: after bias 0 = linear, 1 = LeakyReLU(0.1), 2 = ReLU */
  int32_t n_src;            /* route: number of sources; shortcut: 1                    */
  int32_t src[4];           /* ABSOLUTE layer indices (route layers / shortcut from)    */
  int32_t n_anchors;        /* yolo                                                     */
  int32_t classes;          /* yolo                                                     */
  float   anchors[16];      /* yolo: (w,h) pairs already selected by mask, in pixels    */
} ay_layer_desc;

/* ---- context ------------------------------------------------------------------ */
int  ay_create(ay_ctx** out, int device);
int  ay_destroy(ay_ctx* ctx);
const char* ay_last_error(void);
int  ay_abi_version(void);

/* ---- network (replaces models.py:16-83 create_modules + 237-255 forward) -------- */
/* Describe the network; allocates activations for up to max_batch tiles.          */
int  ay_net_build(ay_ctx* ctx, const ay_layer_desc* layers_h, int n_layers,
                  int in_channels, int img_size, int max_batch, int act_dtype);
/* Weights of conv layer `layer` (index into layers_h): OIHW fp32 with BatchNorm
 * already folded by the caller (w' = w*gamma/sqrt(var+eps), b' = beta-mean*gamma/sqrt(var+eps)),
 * bias_h has `filters` entries.  For an AY_LINEAR layer: nn.Linear's weight [filters][C*H*W] over the flattened NCHW
 * block (core.py:203-204) and its bias.  Synchronous (host buffers may be freed on return). */
int  ay_net_set_conv(ay_ctx* ctx, int layer, const float* w_oihw_h, const float* bias_h);
int  ay_net_finalize(ay_ctx* ctx);
/* rows (sum A*G*G) and columns (5+C) of the decoded output per tile                */
int  ay_net_output_shape(ay_ctx* ctx, int* rows, int* cols);
/* Forward + YOLO decode.  out_decoded: [B, rows, cols] fp32 (xywh, as models.py:161-169). */
int  ay_forward(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                float* out_decoded, void* stream);
/* ay_forward with the confidence threshold of non_max_suppression (utils/utils.py:248, `conf >= conf_thres` on the
 * fp32 value) fused into the head convolutions' epilogue: besides the dense decoded tensor, the decoded ROW INDEX of
 * every passing row is appended to its tile's list cand_idx[b, 0..cand_count[b]) (order arbitrary; cand_count is
 * zeroed here).  cand_count[b] > cand_cap means tile b's list is incomplete - ay_nms_candidates then scans that
 * tile's dense rows instead.  bf16 / fp16 networks only.                                         */
int  ay_forward_candidates(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                           float* out_decoded, float conf_thres,
                           int32_t* cand_idx, int32_t* cand_count, int cand_cap, void* stream);
/* Debug/test helper: copy the stored output of layer `layer` (conv, maxpool, upsample,
 * shortcut, route) of the last forward into out_nchw [B,C,H,W] fp32.                */
int  ay_net_layer_output(ay_ctx* ctx, int layer, int batch, float* out_nchw, void* stream);
int  ay_net_layer_shape(ay_ctx* ctx, int layer, int* channels, int* height, int* width);
/* Which convolution implementation ran for `layer` in the last forward:
 * 0 = none, 1 = tcgen05 implicit GEMM, 2 = first-layer direct kernel, 3 = CUDA-core
 * cross-check over the 16-bit buffers, 4 = fp32 CUDA-core kernel, 5 = 3 x tf32 tcgen05 GEMM (AY_F32 networks).  */
int  ay_net_layer_impl(ay_ctx* ctx, int layer);
/* number of kernels one ay_forward launches for the built network                  */
int  ay_net_launch_count(ay_ctx* ctx);

/* ---- post-processing ------------------------------------------------------------ */
/* utils/utils.py:235-273.  pred: [B,N,5+C] fp32 decoded (xywh); columns 0..3 are
 * rewritten to xyxy IN PLACE (the reference mutates its argument, utils.py:244).
 * out_rows [B,cap,7] (x1,y1,x2,y2,conf,cls_conf,cls_pred), out_keep [B,cap] original
 * row index of each kept box, out_counts [B].  cap >= N guarantees no overflow;
 * tiles with more than `cap` kept boxes set counts to -1 and the call reports
 * AY_ECAPACITY at the next ay_check().                                              */
int  ay_nms(ay_ctx* ctx, float* pred, int batch, int n_rows, int n_classes,
            float conf_thres, float nms_thres,
            float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream);
/* ay_nms starting from the candidate lists of ay_forward_candidates (same conf_thres): only the listed rows of
 * `decoded` are read, and `decoded` is NOT rewritten to xyxy (it is the pipeline's internal buffer, not a caller's
 * tensor).  Results are bit-identical to ay_nms on the same decoded tensor.                          */
int  ay_nms_candidates(ay_ctx* ctx, float* decoded, const int32_t* cand_idx, const int32_t* cand_count, int cand_cap,
                       int batch, int n_rows, int n_classes, float conf_thres, float nms_thres,
                       float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream);
/* ay_nms with a CUDA event between its four kernels (prepare, mask, sweep, cluster): synchronises `stream` and
 * writes their milliseconds to ms4_h[4].  Measurement aid of bench.py --config c5.          */
int  ay_nms_timed(ay_ctx* ctx, float* pred, int batch, int n_rows, int n_classes,
                  float conf_thres, float nms_thres,
                  float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream,
                  float* ms4_h);
/* utils/utils.py:36-50 followed (merge != 0) by core.py:366-423, per tile.
 * rows [B,cap,7] / counts [B] as produced by ay_nms; results in out_rows [B,cap,7]
 * (row order = the reference's CPython-3.12 set order) and out_counts [B].          */
int  ay_rescale_merge(ay_ctx* ctx, const float* rows, const int32_t* counts, int batch, int cap,
                      int current_dim, int orig_h, int orig_w, int merge,
                      float* out_rows, int32_t* out_counts, void* stream);
/* crop.py:13-25 (dzsave, layout google, tile_size 1536, depth one) followed by the nearest Resize of
 * utils/transforms.py:104-111, for a slide (or slide region) already in device memory: wsi is [height, width, 3] uint8
 * with rows pitch_bytes apart; tile (row, col) covers slide pixels [row*tile_size, +tile_size) x [col*tile_size, ...),
 * padded with `background` (dzsave pads edge tiles white, 255) where it leaves the slide.  Writes the out_size^2
 * pixels the resize samples from each of the n_rows x n_cols tiles starting at (row0, col0), row-major, as
 * out [n_rows*n_cols, out_size, out_size, 3] uint8 - the AY_IN_U8_HWC input of ay_forward.  The JPEG (Q=90) round trip of
 * the reference's tile files is not reproduced.                                        */
int  ay_wsi_gather(ay_ctx* ctx, const uint8_t* wsi, int height, int width, long long pitch_bytes, int tile_size,
                   int out_size, int row0, int col0, int n_rows, int n_cols, uint8_t* out, int background, void* stream);
/* Inputs of the CAA filter network for all boxes of one tile: get256Img (core.py:109-159) + ToTensor + Normalize
 * (core.py:214-217, 435).  tile: [height, width, 3] uint8 RGB in device memory; windows: device int32 [n, 4] =
 * (x_left, x_right, y_top, y_bottom) per box, as get256Img slices them (the branch logic stays on the host,
 * caa.crop_window); mean3_h / std3_h: host fp32.  out: device fp32 [n, 3, crop, crop] with the bits of the reference's transforms
 * (x / 255, - mean, / std, one fp32 rounding each) - the AY_IN_F32_NCHW input of the network built from AY_LINEAR.  */
int  ay_crop_normalize(ay_ctx* ctx, const uint8_t* tile, int height, int width, long long pitch_bytes,
                       const int32_t* windows, int n, int crop, const float* mean3_h, const float* std3_h, float* out,
                       void* stream);
/* bbox_iou (utils/utils.py:202-232, x1y1x2y2=True): iou[i] of box1[i or 0] vs box2[i]. */
int  ay_bbox_iou(ay_ctx* ctx, const float* box1, int n1, const float* box2, int n2,
                 float* out_iou, void* stream);
/* ay_nms without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments and results as ay_nms (bit-identical; `pred` is rewritten to xyxy in place), but no context: it runs on
 * the CURRENT device, and the scratch of its four kernels is one file-scope `static __device__` array of 32 MiB
 * (csrc/nms.cu) that every GPU loading the library holds until the process exits.  ay_nms_static_capacity() returns
 * the usable bytes: 2 tiles x 10 647 rows (yolov3 at 416^2), 9 tiles x 6 300 rows (320^2), 36 tiles x 2 535 rows
 * (yolov3-tiny).  A larger batch x n_rows is rejected with AY_ECAPACITY before anything is launched.  It calls no
 * cudaMalloc / cudaFree / allocating library routine.  All calls share the array: two calls that overlap on one GPU
 * (different streams or threads) race - serialise them.  With no context there is no ay_check(): a tile with more
 * than `cap` kept boxes reports out_counts[b] = -1 (cap >= n_rows cannot overflow).                    */
size_t ay_nms_static_capacity(void);
int  ay_nms_static(float* pred, int batch, int n_rows, int n_classes, float conf_thres, float nms_thres,
                   float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream);
/* Synchronise `stream` and report asynchronous errors / capacity overflows.         */
int  ay_check(ay_ctx* ctx, void* stream);

/* ---- test hooks ------------------------------------------------------------------ */
/* Plain CUDA-core direct convolution over the same padded NHWC buffers (no tensor
 * cores); used by tests to cross-check the tcgen05 kernel at full size on the GPU.  */
int  ay_net_use_reference_conv(ay_ctx* ctx, int enable);
/* Profiling: ay_forward with a CUDA event before every cfg block; synchronises `stream` and
 * writes the milliseconds each block's kernels took into ms_per_layer_h[n_layers].   */
int  ay_forward_profile(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                        float* out_decoded, void* stream, float* ms_per_layer_h, int n_layers);
/* Ablation switches of the convolution kernel (tools/layer_times.py): 1 no MMA, 2 no operand
 * loads, 4 no output stores, 8 no residual.  Results are meaningless while any bit is set. */
int  ay_debug_set(ay_ctx* ctx, int bits);
/* Cycle counters of the last ay_forward_profile: 16 per layer (meaning: conv_gemm_sm100.cu). */
int  ay_debug_counters(ay_ctx* ctx, unsigned long long* out_h, int n_layers);
/* Attribution of one forward WITHOUT events between the launches (so launch gaps and the overlap of chained
 * launches are what a normal ay_forward sees): every convolution CTA stamps %globaltimer (ns) at entry, at the
 * end of its prologue and at exit.  stamps_h: [n_layers][ay_prof_words()] — words 0..15 as ay_debug_counters,
 * then 4 blocks of 160 words (entry / prologue end / exit / wait for the previous layer returned, indexed by
 * CTA; 0 = CTA not launched).  Synchronises. */
int  ay_prof_words(void);
int  ay_forward_stamps(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                       float* out_decoded, void* stream, unsigned long long* stamps_h, int n_layers);

#ifdef __cplusplus
}
#endif
#endif /* AMYLOID_YOLO_H_ */
