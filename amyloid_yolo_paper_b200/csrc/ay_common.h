// Shared declarations of the sm_100a tile-detection engine (internal; the public
// C ABI is include/amyloid_yolo.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/amyloid_yolo.h"

namespace ay {

void set_error(const char* fmt, ...);

// negative-side slope of a convolution's activation (ay_layer_desc.leaky): 1 LeakyReLU(0.1), 2 ReLU
__host__ __device__ inline float act_slope(int leaky) { return leaky == 2 ? 0.0f : 0.1f; }

#define AY_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ay::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,     \
                    __LINE__);                                                            \
      return AY_ECUDA;                                                                    \
    }                                                                                     \
  } while (0)

// ---------------------------------------------------------------------------
// Activations: zero-haloed NHWC, 16-bit.  A tensor of logical shape [B,C,H,W]
// is stored as rows of `pitch` elements, one row per pixel of a PADDED grid of
// Hp x Wp = (H + kPadLo + 1) x (W + kPadLo + 1) pixels per image:
//   row(n, y, x) = (n*Hp + y + kPadLo)*Wp + x + kPadLo,   y in [0,H), x in [0,W)
// with every other pixel of the grid, and channels >= C, held at zero.  A 3x3/stride-1
// tap (kh,kw) of output row m is then input row m + (kh-1)*Wp + (kw-1): every
// convolution becomes a GEMM over shifted rows.
//
// kPadLo = 0 (default), the SHARED halo: one zero column after each image row and
// one zero row after each image.  The left neighbour of a row's first pixel is the
// zero column that ends the previous row, the row above an image's first row is the
// zero row that ends the previous image (before image 0 the TMA's out-of-bounds fill
// supplies the zeros), so the frame is paid once instead of twice: the GEMM's M
// dimension shrinks from (H+2)(W+2) to (H+1)(W+1) rows per image - 13 % fewer MMA
// rows at 13x13, 7 % at 26x26.  kPadLo = 1 (-DAY_PAD_LO=1) is the classic one-pixel
// frame of round 1, kept for A/B measurements.
// ---------------------------------------------------------------------------
#ifndef AY_PAD_LO
#define AY_PAD_LO 0
#endif
constexpr int kPadLo = AY_PAD_LO;
__host__ __device__ constexpr int padded(int n) { return n + kPadLo + 1; }
struct ActView {
  void* ptr = nullptr;   // element (0,0,0, channel 0) of this view
  int C = 0;             // logical channels of the view
  int Cp = 0;            // channels rounded up to a multiple of 32 (GEMM K granularity)
  int H = 0, W = 0;      // unpadded spatial size
  int pitch = 0;         // elements between consecutive pixels
  long long rows_alloc = 0;  // pixel rows allocated behind ptr (for max_batch)
};

// Parameters of one CTA-pair implicit-GEMM launch (conv_gemm_sm100.cu).  K is processed in 64-channel
// blocks.  A "group" is one A box (a_rows rows x 64 channels) together with the G weight tiles of the
// taps that read it (G = 3: the kw taps of one kh of a 3x3/stride-1 layer; G = 1 otherwise).
struct ConvGemmParams {
  int P;                 // output rows to compute: batch * (H+2) * (W+2)
  int num_m_tiles;       // ceil(P / (256 * m_batch)): one tile per CTA pair
  int num_n_tiles;       // Cout_padded / BN
  int n_groups;          // A boxes per k-block
  int kb_per_tap;        // ceil(Cin / 64)
  int k16_all, k16_last; // K=16 MMA steps of a 64-channel k-block that hold data: all taps of a group / its last tap (0 = 4)
  int a_row_off[27];     // row offset of group g's A box
  int b_tap0[27];        // first weight tap of group g (K layout is tap-major)
  // fp32 accuracy mode on the tensor cores (kind::tf32, 3-pass split): a group is (pass, tap group); pass 0 contracts
  // A_lo x W_hi, pass 1 A_hi x W_lo, pass 2 A_hi x W_hi (small terms first, see engine.cu).  The activation rows hold [hi | lo] side by side and the weight
  // rows [W_hi | W_lo], so a pass is a column offset into each (elements)
  int a_col_off[27], b_col_off[27];
  int tf32_na;           // tf32 mode: accumulators per tile, summed in the epilogue (1, 2 or 4)
  int tf32;              // 1: fp32 operands (32 channels per 128-byte k-block), plain fp32 row store epilogue
  int pad_lo;            // tf32 mode: zero pixels before each row / column of the OUTPUT grid (the fp32 path keeps a full frame)
  int a_rows;            // rows of the A box (128, or 130 with tap reuse)
  int BN;                // N of the pair tile (each CTA loads BN/2 weight rows)
  int num_stages;
  int m_batch;           // 256-row blocks per tile (row batching; > 1 only with BN = 32, no residual, not a head)
  int units_per_stage;   // (group, k-block) units loaded and consumed per pipeline stage (divides n_groups*kb_per_tap)
  int num_acc;           // TMEM accumulator buffers (512 / BN, at most 16)
  int b_resident;        // this CTA's half of ALL weight tiles stays in shared memory for the whole launch
  int b_res_tiles;       // number of [BN/2 x 64] weight tiles (= K / 64)
  int a_mode;            // 0: 128B-swizzled [a_rows x 64 ch] boxes; 1: first layer, 8-channel pixels read unswizzled
  int nbuf;              // epilogue staging buffers per warp (1 or 2)
  int Hp, Wp;            // padded output size (halo test)
  unsigned div_hw_mul, div_wp_mul; int div_hw_shift, div_wp_shift;   // filled by launch_conv_gemm: row / (Hp*Wp), rem / Wp
  int epi_dr, epi_qrows;                                             // filled by launch_conv_gemm: tile step of an epilogue warp
  const float* bias;     // [num_n_tiles*BN]
  int has_res;           // residual tile added after the activation (shortcut fused)
  void* out_ptr; int out_pitch;          // output rows (16-bit elements), channel 0 of this layer's slice
  const void* res_ptr; int res_pitch;    // residual rows
  int leaky;
  int dbg;               // ablation bits (AY_DBG): 1 no MMA, 2 no operand TMA, 4 no output store, 8 no residual
  // parity-plane output for a stride-2 consumer (see s2_gather_kernel for the layout):
  // 0 = rows only, 1 = rows and planes, 2 = planes only
  int plane_mode;
  void* planes;
  long long plane_rows;
  int plane_pitch;
  int up_mode;           // the only consumer is a x2 nearest upsample: write the 2x2 block into ITS buffer (out_ptr/out_pitch)
  int plane_packed;      // 32-channel producer: the column-parity pair shares one 64-channel row of a ROW-parity plane
  // fused YOLO decode (head convolutions)
  int head;
  int head_fast;         // 3 anchors x 7 columns in one chunk: compile-time channel roles, MUFU exp / reciprocal
  int head_stage_bytes;  // per epilogue warp: n_anchors * 32 * dec_cols floats when the head fits one 32-column chunk, else 0
  float* dec_out;        // [B, dec_rows_total, dec_cols]
  int dec_rows_total, dec_row_off, dec_cols, n_anchors, cout;
  float stride;
  float anchor_w[8], anchor_h[8];   // already divided by stride (models.py:122)
  // confidence-threshold compaction (heads; nullptr = off): rows with conf >= conf_thres are appended to the tile's list
  int32_t* cand_idx;     // [B, cand_cap] decoded row indices, arbitrary order
  int32_t* cand_count;   // [B], zeroed by the caller before the forward; counts past cand_cap mean "list incomplete"
  int cand_cap;
  float conf_thres;
  unsigned int* err_flag;
  unsigned long long* prof;   // kProfWords counters (profiling runs only), else nullptr: [0..15] cycle counters of CTA 0,
                              // then %globaltimer stamps of every CTA: entry, end of prologue, exit, dependency resolved
  int pdl;                    // launch with programmatic stream serialization (prologue overlaps the previous layer's tail)
};
constexpr int kProfStampBase = 16;
constexpr int kProfStampCtas = 160;                       // >= CTAs per launch (148)
constexpr int kProfWords = kProfStampBase + 4 * kProfStampCtas;
int launch_conv_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvGemmParams& p, int G, int is_fp16,
                     int num_sms, cudaStream_t stream);   // is_fp16: 0 bf16, 1 fp16, 2 tf32 (p.tf32 set)
int conv_gemm_stage_bytes(int BN, int G, int a_rows, int resident);
void conv_gemm_pick_smem(int BN, int G, int a_rows, int resident_bytes, int units_per_stage, int extra_bytes, int* stages, int* nbuf);
#ifndef AY_EPI_PARTS
#define AY_EPI_PARTS 2
#endif
constexpr int kConvEpilogueParts = AY_EPI_PARTS;            // epilogue warps per TMEM lane quarter (2 or 4)
constexpr int kConvEpilogueWarps = 4 * kConvEpilogueParts;

// aux_kernels.cu ------------------------------------------------------------
struct FirstConvParams {
  const void* in; int in_kind; int batch; int S; int src_h, src_w;
  const int* idx_y; const int* idx_x;      // nearest-resize source indices (u8 input)
  const float* w;                           // [Cout][3][3][3] (kh,kw,c) fp32
  const float* bias; int cout; int leaky;
  void* out; int out_pitch; int Cp; int is_fp16;
};
// first convolution with the A operand built in tensor memory (first_conv_ts.cu)
struct FirstConvTsParams {
  const void* in; int in_kind; int batch; int S; int src_h, src_w;
  const int* idx_y; const int* idx_x;      // nearest-resize source indices (u8 input)
  const void* w_ts;                         // [32 x 32] 16-bit, K-major, canonical unswizzled core-matrix layout (2 KB)
  const float* bias;                        // [32]
  int leaky, is_fp16;
  void* out_ptr; int out_pitch;             // output rows (unless plane_mode == 2)
  int plane_mode, plane_packed; void* planes; long long plane_rows; int plane_pitch;   // as ConvGemmParams
  unsigned int* err_flag;
};
int launch_first_conv_ts(const FirstConvTsParams& p, int num_sms, cudaStream_t s);
int launch_first_conv(const FirstConvParams& p, cudaStream_t s);
int launch_prep_input(const FirstConvParams& p, cudaStream_t s);   // p.out = [B,S+2,S+2,8] 16-bit
int launch_s2_gather(const void* in, int in_pitch, int C8, int batch, int H, int W,
                     void* planes, long long plane_rows, cudaStream_t s);
int launch_upsample2(const void* in, int in_pitch, int Cp, int batch, int H, int W,
                     void* out, int out_pitch, cudaStream_t s);
int launch_maxpool(const void* in, int in_pitch, int Cp, int batch, int H, int W, int size, int stride,
                   void* out, int out_pitch, int is_fp16, cudaStream_t s);
int launch_add(const void* a, int a_pitch, const void* b, int b_pitch, int Cp, long long rows,
               void* out, int out_pitch, int is_fp16, cudaStream_t s);
int launch_copy_channels(const void* in, int in_pitch, int Cp, long long rows, void* out, int out_pitch,
                         cudaStream_t s);
int launch_dump_nchw(const void* in, int in_pitch, int C, int batch, int H, int W, int is_fp16,
                     float* out, cudaStream_t s);
int launch_dump_planes_nchw(const void* planes, long long plane_rows, int pitch, int C, int batch, int H, int W,
                            int is_fp16, int packed, float* out, cudaStream_t s);
struct RefConvParams {
  const void* in; int in_pitch; int Cin_p; int Cin_k; int H, W;   // input (unpadded H,W); Cin_k: weight columns per tap
  const void* w; int ktot;                                     // [Cout_p][taps*Cin_k] 16-bit
  const float* bias; int cout_p; int size, stride; int leaky;
  const void* residual; int res_pitch;
  void* out; int out_pitch; int Ho, Wo; int batch; int is_fp16;
  float* raw_f32;                                              // head: [P_out][cout_p] fp32 instead of `out`
};
int launch_ref_conv(const RefConvParams& p, cudaStream_t s);
struct DecodeParams {
  const float* raw; int raw_pitch; int batch; int G; int n_anchors; int dec_cols;
  float stride; float anchor_w[8], anchor_h[8];
  float* dec_out; int dec_rows_total, dec_row_off;
  int lo, hi;            // zero pixels before / after each row and column of `raw`'s padded grid
};
int launch_decode(const DecodeParams& p, cudaStream_t s);
int launch_crop_normalize(const uint8_t* tile, int height, int width, long long pitch, const int32_t* windows, int n, int crop,
                          const float* mean3, const float* std3, float* out, cudaStream_t s);
int launch_wsi_gather(const uint8_t* wsi, int height, int width, long long pitch, int T, int S, int row0, int col0, int n_rows,
                      int n_cols, uint8_t* out, int background, cudaStream_t s);
// fully connected output layer over a zero-haloed NHWC activation (16-bit: elem_bytes 2, fp32: 4); w [O][H*W*C] in
// (y, x, c) order, out [batch][O]
int launch_linear(const void* act, int elem_bytes, int is_fp16, int pitch, int lo, int hi, int batch, int C, int H, int W,
                  const float* w, const float* bias, int O, float* out, cudaStream_t s);

// fp32_path.cu (act_dtype = AY_F32: fp32 storage, CUDA-core FFMA accumulation) ------------------
struct F32PrepParams {
  const void* in; int in_kind; int batch; int S; int src_h, src_w;
  const int* idx_y; const int* idx_x;
  float* out;                                 // [B,S+2,S+2,3]
};
struct F32ConvParams {
  const float* in; int in_pitch; int Cin; int H, W;   // zero-haloed NHWC input (unpadded H,W)
  const float* w;                                     // [size*size*Cin][Cout], K order (kh,kw,c)
  const float* bias; int Cout; int size, stride, leaky;
  float* out; int out_pitch; int Ho, Wo; int batch;
};
int launch_f32_prep(const F32PrepParams& p, cudaStream_t s);
// v -> [hi | lo] with hi = v rounded to tf32 (10-bit mantissa, round to nearest) and lo = v - hi (exact): rows of C floats
// in, rows of 2C floats out
int launch_f32_split(const float* in, int in_pitch, int C, long long rows, float* out, cudaStream_t s);
// parity planes (classic one-pixel frame) of a [B,H,W] fp32 tensor with `pitch` floats per pixel, for stride-2 convolutions
int launch_f32_split_s2_gather(const float* in, int in_pitch, int C, int batch, int H, int W, float* planes,
                               long long plane_rows, cudaStream_t s);
int launch_f32_s2_gather(const float* in, int pitch, int batch, int H, int W, float* planes, long long plane_rows,
                         cudaStream_t s);
int launch_f32_conv(const F32ConvParams& p, cudaStream_t s);
int launch_f32_maxpool(const float* in, int C, int batch, int H, int W, int size, int stride, float* out, cudaStream_t s);
int launch_f32_upsample2(const float* in, int C, int batch, int H, int W, float* out, cudaStream_t s);
int launch_f32_add(const float* a, const float* b, long long n, float* out, cudaStream_t s);
int launch_f32_copy_channels(const float* in, int C, long long rows, float* out, int out_pitch, cudaStream_t s);
int launch_f32_dump_nchw(const float* in, int C, int batch, int H, int W, float* out, cudaStream_t s);

// nms.cu / merge.cu -----------------------------------------------------------
struct NmsScratch {
  void* buf = nullptr; size_t bytes = 0;
};
size_t nms_scratch_bytes(int batch, int n_rows);
int launch_nms(float* pred, int batch, int n_rows, int n_classes, float conf_thres, float nms_thres,
               float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap,
               void* scratch, size_t scratch_bytes, unsigned int* err_flag, cudaStream_t s, cudaEvent_t* ev = nullptr,
               const int32_t* cand_idx = nullptr, const int32_t* cand_count = nullptr, int cand_cap = 0);
size_t merge_scratch_bytes(int batch, int cap);
int launch_rescale_merge(const float* rows, const int32_t* counts, int batch, int cap,
                         float pad_x_half, float pad_y_half, float unpad_w, float unpad_h,
                         float orig_w, float orig_h, int merge,
                         float* out_rows, int32_t* out_counts, void* scratch, size_t scratch_bytes,
                         unsigned int* err_flag, cudaStream_t s);
int launch_bbox_iou(const float* b1, int n1, const float* b2, int n2, float* out, cudaStream_t s);

}  // namespace ay
