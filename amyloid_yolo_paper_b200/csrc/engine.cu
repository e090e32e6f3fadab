// Host side of the engine: network plan, buffers, tensor maps, per-layer launches
// and the extern "C" boundary declared in include/amyloid_yolo.h.
//
// Plan = the reference's Darknet.forward layer loop (models.py:237-255) resolved at
// build time: every cfg block gets an output "view" (zero-haloed NHWC rows, see
// ay_common.h); shortcuts are folded into the producing convolution's epilogue,
// single-source routes alias their source, multi-source routes become a shared
// concat buffer that the producers write into directly, [yolo] blocks become the
// decode epilogue of the preceding head convolution.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "ay_common.h"

namespace ay {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int round_up(int v, int m) { return (v + m - 1) / m * m; }

// 2-D K-major map over `rows` rows of `inner` 16-bit elements, row pitch `pitch` elements.
int make_map_2d(CUtensorMap* map, void* base, int inner, long long rows, int pitch, int box_inner, int box_rows,
                int is_fp16, bool swizzle = true) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return AY_ECUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
  const int esz = is_fp16 == 2 ? 4 : 2;                    // is_fp16: 0 bf16, 1 fp16, 2 fp32 (tf32 operands)
  cuuint64_t strides[1] = {(cuuint64_t)pitch * esz};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = !swizzle ? CU_TENSOR_MAP_SWIZZLE_NONE
                                : (box_inner * esz == 128) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  CUresult r = enc(map, is_fp16 == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                        : is_fp16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): inner=%d rows=%lld pitch=%d box=%dx%d base=%p", (int)r, inner, rows,
              pitch, box_inner, box_rows, base);
    return AY_ECUDA;
  }
  return AY_OK;
}

uint16_t to16(float v, int is_fp16) {
  if (is_fp16) {
    __half h = __float2half_rn(v);
    uint16_t u;
    memcpy(&u, &h, 2);
    return u;
  }
  __nv_bfloat16 b = __float2bfloat16_rn(v);
  uint16_t u;
  memcpy(&u, &b, 2);
  return u;
}

float from16(uint16_t u, int is_fp16) {
  if (is_fp16) {
    __half h;
    memcpy(&h, &u, 2);
    return __half2float(h);
  }
  uint32_t w = (uint32_t)u << 16;
  float f;
  memcpy(&f, &w, 4);
  return f;
}

struct Layer {
  ay_layer_desc d;
  int C = 0, H = 0, W = 0;          // logical output shape
  ActView out;                      // where the output lives (ptr == nullptr: not materialised)
  bool owns = false;                // this layer allocated out.ptr
  // convolution
  int cin = 0, cin_p = 0, cin_k = 0, cout_rows = 0, ktot = 0;   // cin_p: channels stored (x32); cin_k: K per tap (x64)
  int BN = 0, G = 1, a_rows = 128, stages = 0, nbuf = 1;
  int units_per_stage = 1;          // (group, k-block) units per pipeline stage
  int m_batch = 1;                  // 256-row blocks per tile (first layer: 4)
  int head_stage_bytes = 0;         // head with <= 32 output channels: shared-memory staging of the decoded rows, per warp
  int b_resident = 0;               // weights stay in shared memory for the whole launch (single N tile, small K)
  int impl = 0;                     // 0 none, 1 tcgen05 gemm
  bool first = false;               // 3-channel first layer: GEMM over the staged 8-channel input x8
  void* x8 = nullptr;               // [B, S+2, S+2, 8] 16-bit (first layer only)
  void* w_ts = nullptr;             // first layer, A-in-TMEM kernel (first_conv_ts.cu): [32 x 32] 16-bit weights, core-matrix layout
  bool l0_ts = false;
  void* w16 = nullptr;              // [cout_rows][ktot] 16-bit (gemm)
  float* w32 = nullptr;             // [cout][27] fp32 (first layer)
  float* bias = nullptr;            // [cout_rows]
  bool weights_set = false;
  int fused_shortcut = -1;          // index of the [shortcut] folded into this conv's epilogue
  int residual_src = -1;
  bool is_head = false;             // followed by [yolo]: decode fused
  int yolo_layer = -1;
  void* planes = nullptr;           // stride-2: parity planes [4][plane_rows][cin_p]
  long long plane_rows = 0;
  bool gather_fused = false;        // stride-2: the producer of the input writes the planes itself
  bool s2_packed = false;           // stride-2 over 32 channels: column-parity pairs packed into 64-channel rows (see pass 3c)
  void* w16_ref = nullptr;          // packed layers: weights in the plain tap-major layout for the CUDA-core reference
  int plane_consumer = -1;          // conv: index of the stride-2 conv whose planes this conv's epilogue fills
  int up_consumer = -1;             // conv: index of the [upsample] block whose output this conv's epilogue writes directly
  bool fused_into_producer = false; // upsample: written by the producing conv
  int plane_mode = 0;               // conv: 1 = rows and planes, 2 = planes only (nothing else reads the rows)
  float* raw_f32 = nullptr;         // reference path scratch for heads
  // tmA: a_rows x 64-channel box of the input (or parity planes); tmB: BN/2 x 64 box of the weights
  CUtensorMap tmA, tmB;
  // yolo
  int dec_row_off = 0;
  // route copy fallback
  std::vector<int> copy_srcs;       // sources that must be copied into the concat buffer
  std::vector<int> copy_offs;
  bool fused_into_prev = false;     // shortcut handled by the previous conv
  // fp32 accuracy mode (AY_F32, fp32_path.cu): zero-haloed NHWC fp32 rows, pitch = C; weights [K][Cout]
  float* f_out = nullptr;
  float* f_w = nullptr;
  float* lin_w32 = nullptr;         // AY_LINEAR: [filters][H*W*C] fp32, columns in (y, x, c) order
  int lin_c = 0, lin_h = 0, lin_w = 0;   // AY_LINEAR: shape of the block it flattens
  float* f_w2 = nullptr;            // tensor-core variant (impl 5): [Cout][2 * taps * Cin] = [W_hi | W_lo], K order (tap, c)
};

}  // namespace
}  // namespace ay

struct ay_ctx {
  int device = 0;
  int num_sms = 148;
  unsigned int* err_flag = nullptr;   // device
  // network
  std::vector<ay::Layer> layers;
  int in_channels = 3, img_size = 416, max_batch = 0, is_fp16 = 0;
  bool built = false, finalized = false, use_ref = false;
  bool is_f32 = false;                // AY_F32 network: every block runs the fp32 kernels of fp32_path.cu
  float* f_in = nullptr;              // AY_F32: staged input [B,S+2,S+2,3]
  float* f_split = nullptr;           // AY_F32, tensor-core convolutions: [hi | lo] copy of the current conv's input rows
  float* f_planes = nullptr;          // ... and its parity planes when the conv has stride 2
  size_t f_split_floats = 0, f_planes_floats = 0;
  int dbg = 0;                        // AY_DBG ablation bits (see ConvGemmParams::dbg)
  int tf32_na = 2;                    // AY_F32 tensor-core convolutions: accumulators per tile (AY_TF32_NA = 1 | 2 | 4)
  bool no_head_fast = false;          // AY_NO_HEADFAST=1: generic decode epilogue for every head
  bool pdl = true;                    // chain the convolution launches with programmatic dependent launch (AY_NO_PDL=1: off)
  int dec_rows = 0, dec_cols = 0;
  int* idx_y = nullptr;               // nearest-resize tables (device), rebuilt when the source size changes
  int* idx_x = nullptr;
  int idx_src_h = -1, idx_src_w = -1;
  std::vector<void*> allocs;
  int launches = 0;
  // per-layer timing (ay_forward_profile): one event before every cfg block and one after the last
  bool profiling = false;
  // ay_forward_candidates: where the head epilogues append the rows that clear the confidence threshold (one forward)
  int32_t* cand_idx = nullptr; int32_t* cand_count = nullptr; int cand_cap = 0; float cand_thres = 0.f;
  bool stamping = false;              // ay_forward_stamps: per-CTA %globaltimer stamps, no events between the launches
  std::vector<cudaEvent_t> prof_events;
  unsigned long long* prof_counters = nullptr;   // [n_layers][16], device
  // post-processing scratch
  void* nms_scratch = nullptr; size_t nms_bytes = 0;
  void* merge_scratch = nullptr; size_t merge_bytes = 0;
};

namespace ay {
namespace {

int dev_alloc(ay_ctx* ctx, void** p, size_t bytes, bool zero) {
  if (bytes == 0) bytes = 256;
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    return AY_ENOMEM;
  }
  ctx->allocs.push_back(*p);
  if (zero) AY_CUDA(cudaMemset(*p, 0, bytes));
  return AY_OK;
}

void free_network(ay_ctx* ctx) {
  for (void* p : ctx->allocs) cudaFree(p);
  ctx->allocs.clear();
  ctx->layers.clear();
  ctx->built = ctx->finalized = false;
  ctx->idx_y = ctx->idx_x = nullptr;
  ctx->idx_src_h = ctx->idx_src_w = -1;
  ctx->f_split = ctx->f_planes = nullptr;
  ctx->f_split_floats = ctx->f_planes_floats = 0;
}

long long padded_rows(int batch, int H, int W) { return (long long)batch * padded(H) * padded(W); }
// the fp32 accuracy mode keeps the classic one-pixel frame (fp32_path.cu)
long long padded_rows_f32(int batch, int H, int W) { return (long long)batch * (H + 2) * (W + 2); }

int alloc_view(ay_ctx* ctx, ActView* v, int C, int H, int W) {
  v->C = C;
  v->Cp = round_up(C, 32);
  v->H = H;
  v->W = W;
  v->pitch = v->Cp;
  v->rows_alloc = padded_rows(ctx->max_batch, H, W);
  // zero-filled so that channel padding and never-written rows hold finite zeros
  return dev_alloc(ctx, &v->ptr, (size_t)v->rows_alloc * v->pitch * 2, true);
}

}  // namespace
}  // namespace ay

// =============================================================================
// C ABI
// =============================================================================
using ay::Layer;
using ay::ActView;
using ay::set_error;

#define AY_ENTER(ctx)                                   \
  do {                                                  \
    if (!(ctx)) {                                       \
      set_error("null context");                        \
      return AY_EINVAL;                                 \
    }                                                   \
    AY_CUDA(cudaSetDevice((ctx)->device));              \
  } while (0)

#define AY_TRY(expr)               \
  do {                             \
    int _r = (expr);               \
    if (_r != AY_OK) return _r;    \
  } while (0)

extern "C" const char* ay_last_error(void) { return ay::g_err; }
extern "C" int ay_abi_version(void) { return 1; }

extern "C" int ay_create(ay_ctx** out, int device) {
  if (!out) { set_error("ay_create: null output pointer"); return AY_EINVAL; }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    set_error("no CUDA device available (%s); this engine has no CPU fallback",
              e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return AY_ECUDA;
  }
  if (device < 0 || device >= count) { set_error("device %d out of range (have %d)", device, count); return AY_EINVAL; }
  AY_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  AY_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major, prop.minor);
    return AY_EINVAL;
  }
  ay_ctx* ctx = new ay_ctx();
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  if (cudaMalloc(&ctx->err_flag, sizeof(unsigned int)) != cudaSuccess) {
    delete ctx;
    set_error("cudaMalloc failed");
    return AY_ENOMEM;
  }
  cudaMemset(ctx->err_flag, 0, sizeof(unsigned int));
  if (const char* e = getenv("AY_DBG")) ctx->dbg = atoi(e);
  if (getenv("AY_NO_PDL")) ctx->pdl = false;
  if (getenv("AY_NO_HEADFAST")) ctx->no_head_fast = true;
  if (const char* e = getenv("AY_TF32_NA")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4) ctx->tf32_na = v; }
  *out = ctx;
  return AY_OK;
}

extern "C" int ay_destroy(ay_ctx* ctx) {
  if (!ctx) return AY_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  ay::free_network(ctx);
  for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
  if (ctx->prof_counters) cudaFree(ctx->prof_counters);
  if (ctx->nms_scratch) cudaFree(ctx->nms_scratch);
  if (ctx->merge_scratch) cudaFree(ctx->merge_scratch);
  if (ctx->err_flag) cudaFree(ctx->err_flag);
  delete ctx;
  return AY_OK;
}

extern "C" int ay_check(ay_ctx* ctx, void* stream) {
  AY_ENTER(ctx);
  cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) { set_error("asynchronous CUDA failure: %s", cudaGetErrorString(e)); return AY_ECUDA; }
  unsigned int flag = 0;
  AY_CUDA(cudaMemcpy(&flag, ctx->err_flag, sizeof(flag), cudaMemcpyDeviceToHost));
  if (flag) {
    AY_CUDA(cudaMemset(ctx->err_flag, 0, sizeof(flag)));
    if ((flag & 0xFFFF0000u) == 0xCA9A0000u) {
      if ((flag & 0xFFFF) == 4) {
        set_error("merge: a detection holds a non-finite value (the reference raises OverflowError/ValueError at int(), "
                  "core.py:330-337)");
        return AY_EINVAL;
      }
      set_error("per-tile capacity exceeded in %s", (flag & 0xFFFF) == 1 ? "nms (more kept boxes than cap)"
                                                    : (flag & 0xFFFF) == 2 ? "merge (more than 512 detections in a tile)"
                                                                           : "merge (entry pool)");
      return AY_ECAPACITY;
    }
    set_error("device-side failure flag 0x%08x (pipeline barrier timeout)", flag);
    return AY_ECUDA;
  }
  return AY_OK;
}

extern "C" int ay_net_use_reference_conv(ay_ctx* ctx, int enable) {
  if (!ctx) { set_error("null context"); return AY_EINVAL; }
  ctx->use_ref = enable != 0;
  return AY_OK;
}

extern "C" int ay_net_build(ay_ctx* ctx, const ay_layer_desc* descs, int n_layers, int in_channels, int img_size,
                            int max_batch, int act_dtype) {
  AY_ENTER(ctx);
  if (!descs || n_layers <= 0 || max_batch <= 0 || img_size <= 0 || img_size % 32 != 0 ||
      (act_dtype != AY_BF16 && act_dtype != AY_FP16 && act_dtype != AY_F32)) {
    set_error("ay_net_build: bad arguments (n_layers=%d max_batch=%d img_size=%d dtype=%d)", n_layers, max_batch,
              img_size, act_dtype);
    return AY_EINVAL;
  }
  AY_CUDA(cudaDeviceSynchronize());
  ay::free_network(ctx);
  ctx->in_channels = in_channels;
  ctx->img_size = img_size;
  ctx->max_batch = max_batch;
  ctx->is_fp16 = act_dtype == AY_FP16;
  ctx->is_f32 = act_dtype == AY_F32;
  ctx->f_in = nullptr;
  ctx->layers.resize(n_layers);
  std::vector<Layer>& Ls = ctx->layers;

  // ---- pass 1: shapes (models.py:16-83 bookkeeping of output_filters, plus spatial sizes)
  int pc = in_channels, ph = img_size, pw = img_size;
  std::vector<std::vector<int>> explicit_refs(n_layers);
  ctx->dec_rows = 0;
  ctx->dec_cols = 0;
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = Ls[i];
    L.d = descs[i];
    const ay_layer_desc& d = L.d;
    switch (d.type) {
      case AY_CONV: {
        if (d.size != 1 && d.size != 3) { set_error("layer %d: conv size %d unsupported", i, d.size); return AY_EINVAL; }
        if (d.stride != 1 && d.stride != 2) { set_error("layer %d: conv stride %d unsupported", i, d.stride); return AY_EINVAL; }
        const int pad = (d.size - 1) / 2;
        L.cin = pc;
        L.C = d.filters;
        L.H = (ph + 2 * pad - d.size) / d.stride + 1;
        L.W = (pw + 2 * pad - d.size) / d.stride + 1;
        break;
      }
      case AY_MAXPOOL:
        L.C = pc;
        L.H = d.stride == 2 ? ph / 2 : ph;
        L.W = d.stride == 2 ? pw / 2 : pw;
        break;
      case AY_UPSAMPLE:
        if (d.stride != 2) { set_error("layer %d: upsample stride %d unsupported", i, d.stride); return AY_EINVAL; }
        L.C = pc; L.H = ph * 2; L.W = pw * 2;
        break;
      case AY_ROUTE: {
        if (d.n_src < 1 || d.n_src > 4) { set_error("layer %d: route with %d sources", i, d.n_src); return AY_EINVAL; }
        int c = 0;
        for (int k = 0; k < d.n_src; ++k) {
          const int s = d.src[k];
          if (s < 0 || s >= i) { set_error("layer %d: route source %d out of range", i, s); return AY_EINVAL; }
          if (k && (Ls[s].H != Ls[d.src[0]].H || Ls[s].W != Ls[d.src[0]].W)) {
            set_error("layer %d: route sources differ in size", i); return AY_EINVAL;
          }
          c += Ls[s].C;
          explicit_refs[s].push_back(i);
        }
        L.C = c; L.H = Ls[d.src[0]].H; L.W = Ls[d.src[0]].W;
        break;
      }
      case AY_SHORTCUT: {
        const int s = d.src[0];
        if (i == 0 || s < 0 || s >= i || Ls[s].C != pc || Ls[s].H != ph || Ls[s].W != pw) {
          set_error("layer %d: shortcut source %d has a different shape", i, s); return AY_EINVAL;
        }
        explicit_refs[s].push_back(i);
        L.C = pc; L.H = ph; L.W = pw;
        break;
      }
      case AY_YOLO: {
        if (i == 0 || Ls[i - 1].d.type != AY_CONV || d.n_anchors < 1 || d.n_anchors > 8 ||
            pc != d.n_anchors * (5 + d.classes) || ph != pw) {
          set_error("layer %d: yolo block must follow a conv with %d*(5+%d) filters", i, d.n_anchors, d.classes);
          return AY_EINVAL;
        }
        if (ctx->dec_cols && ctx->dec_cols != 5 + d.classes) { set_error("yolo blocks disagree on classes"); return AY_EINVAL; }
        ctx->dec_cols = 5 + d.classes;
        L.dec_row_off = ctx->dec_rows;
        ctx->dec_rows += d.n_anchors * ph * pw;
        L.C = pc; L.H = ph; L.W = pw;
        break;
      }
      case AY_LINEAR: {
        if (i == 0 || i != n_layers - 1 || d.filters < 1 || ctx->dec_rows != 0) {
          set_error("layer %d: a fully connected layer must be the last block of a network without [yolo] blocks", i);
          return AY_EINVAL;
        }
        L.lin_c = pc; L.lin_h = ph; L.lin_w = pw;
        ctx->dec_rows = 1;
        ctx->dec_cols = d.filters;
        L.C = d.filters; L.H = 1; L.W = 1;
        break;
      }
      default:
        set_error("layer %d: unknown type %d", i, d.type);
        return AY_EINVAL;
    }
    pc = L.C; ph = L.H; pw = L.W;
  }
  if (ctx->dec_rows == 0) { set_error("network has no [yolo] block and no fully connected output layer"); return AY_EINVAL; }

  // ---- fp32 accuracy mode: one buffer per block, no fusion
  if (ctx->is_f32) {
    if (in_channels != 3) { set_error("fp32 mode: the input must have 3 channels"); return AY_EINVAL; }
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->f_in),
                         sizeof(float) * (size_t)ay::padded_rows_f32(max_batch, img_size, img_size) * 3, true));
    for (int i = 0; i < n_layers; ++i) {
      Layer& L = Ls[i];
      const size_t elems = (size_t)ay::padded_rows_f32(max_batch, L.H, L.W) * L.C;
      switch (L.d.type) {
        case AY_CONV:
          L.impl = 4;
          // Tensor-core variant (3 x tf32): every convolution whose channel counts are multiples of 32 - all but the
          // first layer and the heads of the shipped cfgs, > 99.5 % of the FLOPs.  AY_F32_FFMA=1 keeps the CUDA-core
          // kernel everywhere (cross-check and A/B).
          if (L.cin % 32 == 0 && L.C % 32 == 0 && !getenv("AY_F32_FFMA")) {
            L.impl = 5;
            const int taps = L.d.size * L.d.size;
            L.ktot = taps * L.cin;
            // N tile of at most 128 columns, so that a tile can spread its MMAs over several accumulators (tf32_na)
            const int bn_cap = getenv("AY_TF32_BN") ? atoi(getenv("AY_TF32_BN")) : 128;
            L.BN = (L.C % 128 == 0 && bn_cap >= 128) ? 128 : (L.C % 64 == 0 && bn_cap >= 64) ? 64 : 32;
            L.cout_rows = L.C;
            L.G = (L.d.size == 3 && L.d.stride == 1) ? 3 : 1;
            L.a_rows = L.G == 3 ? 130 : 128;
            L.units_per_stage = 1;
            ay::conv_gemm_pick_smem(L.BN, L.G, L.a_rows, 0, 1, 0, &L.stages, &L.nbuf);
            if (L.G == 1 && !getenv("AY_ONE_UNIT")) {
              // 1x1 and stride-2 layers have one tap per unit (four K=8 MMAs): three consecutive units per stage amortise
              // the stage's barrier round trip over twelve MMAs, as the three taps of a stride-1 3x3 unit do
              int st3 = 0, nb3 = 0;
              ay::conv_gemm_pick_smem(L.BN, L.G, L.a_rows, 0, 3, 0, &st3, &nb3);
              if (st3 >= 3) { L.units_per_stage = 3; L.stages = st3; L.nbuf = nb3; }
            }
            if (L.stages < 2) { set_error("layer %d: tf32 tile does not fit in shared memory", i); return AY_EINVAL; }
            AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.f_w2), sizeof(float) * (size_t)L.C * 2 * L.ktot, true));
            const int iH = i ? Ls[i - 1].H : img_size, iW = i ? Ls[i - 1].W : img_size;
            ctx->f_split_floats = std::max(ctx->f_split_floats, (size_t)ay::padded_rows_f32(max_batch, iH, iW) * 2 * L.cin);
            if (L.d.stride == 2) {
              if (L.d.size != 3 || (iH & 1) || (iW & 1)) { set_error("layer %d: stride-2 conv needs size 3 and even input", i); return AY_EINVAL; }
              L.plane_rows = ay::padded_rows_f32(max_batch, L.H, L.W);
              ctx->f_planes_floats = std::max(ctx->f_planes_floats, (size_t)(4 * L.plane_rows * 2 * L.cin));
            }
          }
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.f_w), sizeof(float) * (size_t)L.d.size * L.d.size * L.cin * L.C, false));
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), sizeof(float) * L.C, true));
          // tensor-core convolution followed by a [shortcut] and read by nothing else: the add moves into its epilogue and
          // the convolution's own output is never stored (AY_F32_NO_RESFUSE=1: separate add, every block stored)
          if (L.impl == 5 && i + 1 < n_layers && Ls[i + 1].d.type == AY_SHORTCUT && explicit_refs[i].empty() &&
              Ls[i + 1].d.src[0] != i && !getenv("AY_F32_NO_RESFUSE")) {
            L.fused_shortcut = i + 1;
            L.residual_src = Ls[i + 1].d.src[0];
            Ls[i + 1].fused_into_prev = true;
            break;
          }
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.f_out), sizeof(float) * elems, true));
          break;
        case AY_ROUTE:
          if (L.d.n_src == 1) { L.f_out = Ls[L.d.src[0]].f_out; break; }
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.f_out), sizeof(float) * elems, true));
          break;
        case AY_YOLO:
          L.f_out = Ls[i - 1].f_out;
          break;
        case AY_LINEAR:
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.lin_w32), sizeof(float) * (size_t)L.C * L.lin_c * L.lin_h * L.lin_w, true));
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), sizeof(float) * L.C, true));
          break;
        default:
          AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.f_out), sizeof(float) * elems, true));
          break;
      }
    }
    if (ctx->f_split_floats) AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->f_split), sizeof(float) * ctx->f_split_floats, true));
    if (ctx->f_planes_floats) AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->f_planes), sizeof(float) * ctx->f_planes_floats, true));
    for (int i = 0; i < n_layers; ++i) {
      Layer& L = Ls[i];
      if (L.d.type != AY_CONV || L.impl != 5) continue;
      const int iH = i ? Ls[i - 1].H : img_size, iW = i ? Ls[i - 1].W : img_size;
      AY_TRY(ay::make_map_2d(&L.tmB, L.f_w2, 2 * L.ktot, L.C, 2 * L.ktot, 32, L.BN / 2, 2));
      if (L.d.stride == 2)
        AY_TRY(ay::make_map_2d(&L.tmA, ctx->f_planes, 2 * L.cin, 4 * L.plane_rows, 2 * L.cin, 32, L.a_rows, 2));
      else
        AY_TRY(ay::make_map_2d(&L.tmA, ctx->f_split, 2 * L.cin, ay::padded_rows_f32(max_batch, iH, iW), 2 * L.cin, 32, L.a_rows, 2));
    }
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->idx_y), sizeof(int) * img_size, true));
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->idx_x), sizeof(int) * img_size, true));
    AY_CUDA(cudaDeviceSynchronize());
    ctx->built = true;
    return AY_OK;
  }

  // ---- pass 2: fusion decisions
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = Ls[i];
    if (L.d.type == AY_CONV) {
      const bool first = (i == 0 && L.cin == in_channels && L.cin < 8);
      if (first) {
        if (L.d.size != 3 || L.d.stride != 1 || in_channels != 3) {
          set_error("first convolution must be 3x3 stride 1 over 3 channels"); return AY_EINVAL;
        }
        L.impl = 1;
        L.first = true;
      } else {
        L.impl = 1;
        if (L.d.stride == 2 && (L.d.size != 3 || (Ls[i - 1].H & 1) || (Ls[i - 1].W & 1))) {
          set_error("layer %d: stride-2 conv needs size 3 and even input", i); return AY_EINVAL;
        }
      }
      if (i + 1 < n_layers && Ls[i + 1].d.type == AY_YOLO) {
        if (L.impl != 1 || !explicit_refs[i].empty()) { set_error("layer %d: unsupported head layout", i); return AY_EINVAL; }
        L.is_head = true;
        L.yolo_layer = i + 1;
      } else if (i + 1 < n_layers && Ls[i + 1].d.type == AY_SHORTCUT && L.impl == 1 && explicit_refs[i].empty()) {
        L.fused_shortcut = i + 1;
        L.residual_src = Ls[i + 1].d.src[0];
        Ls[i + 1].fused_into_prev = true;
      }
    }
  }

  // ---- pass 3a: multi-source routes own a concat buffer; sources write into it when possible
  std::vector<char> placed(n_layers, 0);
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = Ls[i];
    if (L.d.type != AY_ROUTE || L.d.n_src < 2) continue;
    for (int k = 0; k < L.d.n_src; ++k)
      if (Ls[L.d.src[k]].C % 32 != 0) { set_error("layer %d: concat of a %d-channel tensor unsupported", i, Ls[L.d.src[k]].C); return AY_EINVAL; }
    AY_TRY(ay::alloc_view(ctx, &L.out, L.C, L.H, L.W));
    L.owns = true;
    placed[i] = 1;
    int off = 0;
    for (int k = 0; k < L.d.n_src; ++k) {
      const int s = L.d.src[k];
      Layer& S = Ls[s];
      const bool materialises = (S.d.type == AY_CONV && !S.is_head && S.fused_shortcut < 0) || S.d.type == AY_MAXPOOL ||
                                S.d.type == AY_UPSAMPLE || S.d.type == AY_SHORTCUT;
      if (materialises && !placed[s]) {
        S.out = L.out;
        S.out.ptr = static_cast<uint16_t*>(L.out.ptr) + off;
        S.out.C = S.C;
        S.out.Cp = S.C;
        placed[s] = 1;
      } else {
        L.copy_srcs.push_back(s);
        L.copy_offs.push_back(off);
      }
      off += S.C;
    }
  }
  // ---- pass 3b: everything else
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = Ls[i];
    const int t = L.d.type;
    const bool materialises = (t == AY_CONV && !L.is_head && L.fused_shortcut < 0) || t == AY_MAXPOOL ||
                              t == AY_UPSAMPLE || t == AY_SHORTCUT;
    if (materialises && !placed[i]) {
      AY_TRY(ay::alloc_view(ctx, &L.out, L.C, L.H, L.W));
      L.owns = true;
      placed[i] = 1;
    } else if (t == AY_ROUTE && L.d.n_src == 1) {
      L.out = Ls[L.d.src[0]].out;
      if (!L.out.ptr) { set_error("layer %d: route source %d is not materialised", i, L.d.src[0]); return AY_EINVAL; }
    }
  }

  // ---- pass 3c: parity planes of stride-2 convolutions; the GEMM that produces their input fills them directly
  for (int i = 1; i < n_layers; ++i) {
    Layer& L = Ls[i];
    if (L.d.type != AY_CONV || L.impl != 1 || L.d.stride != 2) continue;
    const ActView& in = Ls[i - 1].out;
    if (!in.ptr) { set_error("layer %d: input of the convolution is not materialised", i); return AY_EINVAL; }
    L.plane_rows = ay::padded_rows(max_batch, L.H, L.W);
    AY_TRY(ay::dev_alloc(ctx, &L.planes, (size_t)4 * L.plane_rows * in.Cp * 2, true));
    int pc = -1;
    if (Ls[i - 1].d.type == AY_CONV && Ls[i - 1].impl == 1 && !Ls[i - 1].is_head) pc = i - 1;
    else if (Ls[i - 1].d.type == AY_SHORTCUT && Ls[i - 1].fused_into_prev) pc = i - 2;
    if (pc >= 0 && Ls[pc].plane_consumer < 0) {      // (the planes have their own pitch: the input may live in a concat slice)
      Ls[pc].plane_consumer = i;
      Ls[pc].plane_mode = explicit_refs[i - 1].empty() ? 2 : 1;
      L.gather_fused = true;
      // 32-channel input (64-byte pixels): the producer interleaves the two column-parity planes, so that a 128-byte
      // row of a ROW-parity plane holds the pixel pair (x even | x odd).  Taps kw = 0,1 of a filter row are then one
      // 64-channel k-block of box row r, and tap kw = 2 is the first half of box row r + 1: three 129-row A boxes per
      // tile instead of nine half-empty ones, and 6 k-blocks of MMAs instead of 9.
      if (in.Cp == 32 && !getenv("AY_NO_PACKED")) L.s2_packed = true;
    }
  }

  // ---- pass 3d: conv -> [upsample] with no other reader of the conv: the epilogue writes the 2x2 blocks itself
  for (int i = 0; i + 1 < n_layers; ++i) {
    Layer& L = Ls[i];
    if (L.d.type != AY_CONV || L.impl != 1 || L.is_head || L.fused_shortcut >= 0 || L.plane_consumer >= 0) continue;
    if (Ls[i + 1].d.type != AY_UPSAMPLE || !explicit_refs[i].empty() || getenv("AY_NO_UPFUSE")) continue;
    if (i + 2 < n_layers && Ls[i + 2].d.type != AY_ROUTE) continue;      // (the conv's own rows would be needed otherwise)
    L.up_consumer = i + 1;
    Ls[i + 1].fused_into_producer = true;
  }

  const int ph0 = img_size, pw0 = img_size;
  // ---- pass 4: convolution resources
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = Ls[i];
    if (L.d.type != AY_CONV) continue;
    if (L.first) {
      // First layer as a GEMM with K = 3 (kh) x 64: x8 holds 8 channels (3 used) per padded pixel, 16 bytes, so the
      // three pixels (y+kh-1, x-1..x+1) of a filter row are 24 contiguous elements.  The kernel loads 144 consecutive
      // pixels per filter row and reads them as overlapping unswizzled core matrices (conv_gemm_sm100.cu, a_mode 1).
      // Weight column kh*64 + kw*8 + c.
      AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.w32), sizeof(float) * L.C * 27, true));   // reference path
      const long long rows8 = (ay::padded_rows(max_batch, ph0, pw0) + 7) / 8 * 8;
      AY_TRY(ay::dev_alloc(ctx, &L.x8, (size_t)rows8 * 8 * 2, true));
      L.cin_p = 8;
      L.cin_k = 64;
      L.ktot = 3 * 64;
    }
    static const ActView no_view;
    const ActView& in = L.first ? no_view : Ls[i - 1].out;
    if (!L.first && !in.ptr) { set_error("layer %d: input of the convolution is not materialised", i); return AY_EINVAL; }
    const int taps = L.d.size * L.d.size;
    if (!L.first) {
      L.cin_p = in.Cp;
      L.cin_k = ay::round_up(L.cin_p, 64);
      L.ktot = taps * L.cin_k;
      if (L.s2_packed) L.ktot = 6 * 64;       // per filter row: [kw0 | kw1] and [kw2 | zeros]
    }
    const int cout_p = ay::round_up(L.C, 32);
    if (L.is_head) L.BN = cout_p <= 256 ? cout_p : 256;
    else L.BN = (cout_p % 256 == 0) ? 256 : (cout_p % 128 == 0) ? 128 : (cout_p % 64 == 0) ? 64 : 32;
    L.cout_rows = ay::round_up(cout_p, L.BN);
    if (L.is_head && cout_p == 32 && !getenv("AY_NO_HEADSTAGE")) {
      const ay_layer_desc& Y = Ls[L.yolo_layer].d;
      L.head_stage_bytes = (Y.n_anchors * 32 * (5 + Y.classes) * 4 + 15) & ~15;
    } else if (L.is_head && !getenv("AY_NO_HEADSTAGE")) {
      L.head_stage_bytes = 32 * 33 * 4;         // wide heads (yolov3-tiny: 255 channels): a 32 x 32 transpose per warp and chunk
    }
    L.G = L.s2_packed ? 2 : (L.d.size == 3 && L.d.stride == 1 && !L.first) ? 3 : 1;
    // first layer: four 256-row blocks per tile (K is only 3 x 32: one barrier round and three loads per 1024 rows);
    // each CTA's box is its 512 consecutive pixels + shift + taps = 66 rows of 8 pixels
    if (L.first && !getenv("AY_NO_MBATCH")) L.m_batch = 4;
    L.a_rows = L.first ? (128 * L.m_batch + 16) / 8 : L.G == 3 ? 130 : L.G == 2 ? 129 : 128;
    {
      // resident weights when this CTA's half of the matrix leaves room for the A ring;
      // units per stage: the largest divisor of the tile's unit count that keeps a stage <= 56 KB and >= 3 stages
      const int res_bytes = (L.BN / 2) * L.ktot * 2;
      const int n_groups = L.first ? 3 : L.d.size == 1 ? 1 : (L.d.stride == 1 || L.s2_packed) ? 3 : 9;
      const int units = n_groups * (L.s2_packed ? 1 : L.cin_k / 64);
      // 128 KB admits the 1x1 512 -> 256 layers (11 at 26^2, one at 13^2): cold they gain nothing (the resident load is
      // exposed at the start of the launch), but the power-capped forward is 1.2 % faster with half the L2 -> SM weight
      // traffic of those layers gone (A/B on the B200: 18 293 / 18 341 us against 18 049 / 18 122)
      const int res_cap = (getenv("AY_RES_KB") ? atoi(getenv("AY_RES_KB")) : 128) * 1024;
      const bool try_res = L.cout_rows == L.BN && res_bytes <= res_cap && !getenv("AY_NO_RESIDENT");
      L.units_per_stage = 1;
      const int stage_cap = (getenv("AY_STAGE_KB") ? atoi(getenv("AY_STAGE_KB")) : 56) * 1024;
      for (int pass = try_res ? 0 : 1; pass < 2 && L.stages == 0; ++pass) {
        const int rb = pass == 0 ? res_bytes : 0;
        const int unit_bytes = ay::conv_gemm_stage_bytes(L.BN, L.G, L.a_rows, rb > 0);
        int best_u = 0, best_st = 0, best_nb = 0;
        for (int u = 1; u <= units; ++u) {
          if (units % u != 0 || (u > 1 && (u * unit_bytes > stage_cap || getenv("AY_ONE_UNIT")))) continue;
          int st = 0, nb = 0;
          ay::conv_gemm_pick_smem(L.BN, L.G, L.a_rows, rb, u, ay::kConvEpilogueWarps * L.head_stage_bytes, &st, &nb);
          if (st >= 3 && (u == 1 || nb == 2)) { best_u = u; best_st = st; best_nb = nb; }
        }
        if (best_u == 0 && pass == 1) {
          best_u = 1;
          ay::conv_gemm_pick_smem(L.BN, L.G, L.a_rows, 0, 1, ay::kConvEpilogueWarps * L.head_stage_bytes, &best_st, &best_nb);
        }
        if (best_u > 0 && (pass == 1 || best_st >= 3)) {
          L.b_resident = pass == 0;
          L.units_per_stage = best_u;
          L.stages = best_st;
          L.nbuf = best_nb;
        }
      }
    }
    if (L.stages < 2) { set_error("layer %d: tile does not fit in shared memory", i); return AY_EINVAL; }
    // first layer with <= 32 filters and a plain output, opt-in (AY_L0_TS=1): im2col rows built in tensor memory
    // (first_conv_ts.cu).  Bit-compatible in the tests, but instruction-bound - 1 434 us against 1 123 for the staged
    // path at B = 256 (profiles/r2_ab_l0_tmem.txt) - so the staged path stays the default.
    if (L.first && cout_p == 32 && !L.is_head && L.fused_shortcut < 0 && L.up_consumer < 0 && (img_size & 1) == 0 &&
        getenv("AY_L0_TS")) {
      L.l0_ts = true;
      AY_TRY(ay::dev_alloc(ctx, &L.w_ts, 32 * 32 * 2, true));
    }
    AY_TRY(ay::dev_alloc(ctx, &L.w16, (size_t)L.cout_rows * L.ktot * 2, true));
    if (L.s2_packed) AY_TRY(ay::dev_alloc(ctx, &L.w16_ref, (size_t)L.cout_rows * 9 * L.cin_k * 2, true));
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), sizeof(float) * L.cout_rows, true));
    AY_TRY(ay::make_map_2d(&L.tmB, L.w16, L.ktot, L.cout_rows, L.ktot, 64, L.BN / 2, ctx->is_fp16));
    // A boxes are always 64 channels wide; when the tensor stores only 32 (cin_p = 32) the upper half of the
    // box lies outside the tensor and the TMA zero-fills it (the matching weight columns are zero as well)
    if (L.first) {
      // x8 viewed as rows of 8 pixels x 8 channels (128 contiguous bytes): an 18-row box is 144 consecutive pixels
      AY_TRY(ay::make_map_2d(&L.tmA, L.x8, 64, (ay::padded_rows(max_batch, ph0, pw0) + 7) / 8, 64, 64, L.a_rows,
                             ctx->is_fp16, /*swizzle=*/false));
    } else if (L.d.stride == 2) {
      if (L.s2_packed) AY_TRY(ay::make_map_2d(&L.tmA, L.planes, 64, 2 * L.plane_rows, 64, 64, L.a_rows, ctx->is_fp16));
      else AY_TRY(ay::make_map_2d(&L.tmA, L.planes, L.cin_p, 4 * L.plane_rows, L.cin_p, 64, L.a_rows, ctx->is_fp16));
    } else {
      AY_TRY(ay::make_map_2d(&L.tmA, in.ptr, L.cin_p, in.rows_alloc, in.pitch, 64, L.a_rows, ctx->is_fp16));
    }
    if (L.is_head)
      AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.raw_f32),
                           sizeof(float) * (size_t)ay::padded_rows(max_batch, L.H, L.W) * L.cout_rows, true));
  }
  if (Ls[n_layers - 1].d.type == AY_LINEAR) {
    Layer& L = Ls[n_layers - 1];
    if (!Ls[n_layers - 2].out.ptr) { set_error("layer %d: input of the fully connected layer is not materialised", n_layers - 1); return AY_EINVAL; }
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.lin_w32), sizeof(float) * (size_t)L.C * L.lin_c * L.lin_h * L.lin_w, true));
    AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), sizeof(float) * L.C, true));
  }
  AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->idx_y), sizeof(int) * img_size, true));
  AY_TRY(ay::dev_alloc(ctx, reinterpret_cast<void**>(&ctx->idx_x), sizeof(int) * img_size, true));
  AY_CUDA(cudaDeviceSynchronize());
  ctx->built = true;
  return AY_OK;
}

extern "C" int ay_net_set_conv(ay_ctx* ctx, int layer, const float* w, const float* bias) {
  AY_ENTER(ctx);
  if (!ctx->built || layer < 0 || layer >= (int)ctx->layers.size() || (ctx->layers[layer].d.type != AY_CONV && ctx->layers[layer].d.type != AY_LINEAR) || !w || !bias) {
    set_error("ay_net_set_conv: bad layer %d", layer);
    return AY_EINVAL;
  }
  Layer& L = ctx->layers[layer];
  if (L.d.type == AY_LINEAR) {
    // nn.Linear weight [O][C*H*W] over x.view(B, -1) of an NCHW tensor (core.py:203-204) -> columns in (y, x, c) order
    const int C = L.lin_c, HW = L.lin_h * L.lin_w;
    std::vector<float> r((size_t)L.C * C * HW);
    for (int o = 0; o < L.C; ++o)
      for (int c = 0; c < C; ++c)
        for (int px = 0; px < HW; ++px) r[((size_t)o * HW + px) * C + c] = w[((size_t)o * C + c) * HW + px];
    AY_CUDA(cudaMemcpy(L.lin_w32, r.data(), r.size() * sizeof(float), cudaMemcpyHostToDevice));
    AY_CUDA(cudaMemcpy(L.bias, bias, sizeof(float) * L.C, cudaMemcpyHostToDevice));
    L.weights_set = true;
    ctx->finalized = false;
    return AY_OK;
  }
  const int k = L.d.size, taps = k * k, cout = L.C, cin = L.cin;
  if (ctx->is_f32) {
    std::vector<float> r((size_t)taps * cin * cout);          // OIHW -> [(kh,kw,c)][o]
    for (int o = 0; o < cout; ++o)
      for (int c = 0; c < cin; ++c)
        for (int t = 0; t < taps; ++t) r[((size_t)t * cin + c) * cout + o] = w[((size_t)o * cin + c) * taps + t];
    AY_CUDA(cudaMemcpy(L.f_w, r.data(), r.size() * sizeof(float), cudaMemcpyHostToDevice));
    AY_CUDA(cudaMemcpy(L.bias, bias, sizeof(float) * cout, cudaMemcpyHostToDevice));
    if (L.impl == 5) {
      // [W_hi | W_lo] per output channel, K order (tap, c): hi = w rounded to tf32, lo = w - hi (exact in fp32).
      //
      // De-biasing of the accumulation.  The tensor core adds each K=8 MMA into the fp32 accumulator with TRUNCATION: the
      // running sum loses on average half an ulp of itself per MMA, E = 0.5 * 0.7213 * 2^-23 = 4.3e-8 relative (0.7213 =
      // mean of 1/mantissa).  The term contributed by the q-th of the n = K/8 MMAs of the final (A_hi x W_hi) pass is part
      // of the sum for the remaining n - q truncations, so it arrives shrunk by (n - q) * E: measured as a coherent
      // 5e-5 shrink of the deepest head, 5e-4 on small elements that are differences of large partial sums.  The expected
      // loss is linear in the terms, so it is restored exactly (in expectation, for every element) by inflating each
      // weight by its own factor (1 + (n - q) * E).  W_hi cannot carry a 1e-5 change (10-bit mantissa), W_lo can: the
      // inflation is added there, no extra pass.  q follows the kernel's issue order (tap group, k-block, tap, K=8 step).
      // The factor 0.75 is calibrated on the B200 (sweep of AY_TF32_KAPPA over 0 .. 2 against the float64 oracle: the
      // deepest raw head of yolov3-custom is off by 5.0e-5 of its range without the ramp, 3.8e-6 at 0.75, 1.1e-5 at 1.0):
      // the adder keeps a little more than the 24 stored bits before it truncates.
      const double E = 0.5 * 0.7213 / 8388608.0 * (getenv("AY_TF32_KAPPA") ? atof(getenv("AY_TF32_KAPPA")) : 0.75);
      const int Gt = L.G, kbpt = cin / 32;
      // units (tap group, k-block) of a pass; the final pass's units are U = 2*units .. 3*units-1 of the tile and unit U
      // accumulates into accumulator U % NA: position of each unit among ITS accumulator's final-pass units
      const int units = (taps / Gt) * kbpt, NAw = ctx->tf32_na;
      std::vector<int> upos(units), ucnt(NAw, 0);
      for (int i2 = 0; i2 < units; ++i2) upos[i2] = ucnt[(2 * units + i2) % NAw]++;
      std::vector<float> r2((size_t)cout * 2 * L.ktot);
      for (int o = 0; o < cout; ++o)
        for (int c = 0; c < cin; ++c)
          for (int t = 0; t < taps; ++t) {
            const float v = w[((size_t)o * cin + c) * taps + t];
            uint32_t u;
            memcpy(&u, &v, 4);
            u = (u + 0x1000u) & 0xFFFFE000u;
            float hi;
            memcpy(&hi, &u, 4);
            const int unit = (t / Gt) * kbpt + c / 32;
            const int q = (upos[unit] * Gt + t % Gt) * 4 + (c % 32) / 8;                    // MMAs before this one in its accumulator
            const int n_mma = ucnt[(2 * units + unit) % NAw] * Gt * 4;
            r2[(size_t)o * 2 * L.ktot + (size_t)t * cin + c] = hi;
            r2[(size_t)o * 2 * L.ktot + L.ktot + (size_t)t * cin + c] = (float)((double)v * (1.0 + (double)(n_mma - q) * E) - (double)hi);
          }
      AY_CUDA(cudaMemcpy(L.f_w2, r2.data(), r2.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    L.weights_set = true;
    ctx->finalized = false;
    return AY_OK;
  }
  if (L.first) {
    std::vector<float> r((size_t)cout * 27);
    std::vector<uint16_t> r16((size_t)L.cout_rows * L.ktot, 0);
    for (int o = 0; o < cout; ++o)
      for (int c = 0; c < 3; ++c)
        for (int t = 0; t < 9; ++t) {
          // the CUDA-core reference of this layer uses the SAME rounded operands as the tensor-core path
          r[(size_t)o * 27 + t * 3 + c] =
              ay::from16(ay::to16((float)((double)w[((size_t)o * 3 + c) * 9 + t] / 255.0), ctx->is_fp16), ctx->is_fp16);
          // the staged input holds pixel values 0..255 (exact in 16 bits): ToTensor's 1/255 lives in the weights
          r16[(size_t)o * L.ktot + (t / 3) * 64 + (t % 3) * 8 + c] =
              ay::to16((float)((double)w[((size_t)o * 3 + c) * 9 + t] / 255.0), ctx->is_fp16);
        }
    std::vector<float> b(L.cout_rows, 0.f);
    memcpy(b.data(), bias, sizeof(float) * cout);
    AY_CUDA(cudaMemcpy(L.w32, r.data(), r.size() * sizeof(float), cudaMemcpyHostToDevice));
    AY_CUDA(cudaMemcpy(L.w16, r16.data(), r16.size() * 2, cudaMemcpyHostToDevice));
    if (L.l0_ts) {
      // [filter n][k = (kh*3 + kw)*3 + c], K-major, as unswizzled 8 x 16-byte core matrices: K chunk k/8 of filter group n/8
      // at ((n/8)*4 + k/8) * 128 bytes, row n%8 at 16 bytes, element k%8 at 2 bytes
      std::vector<uint16_t> ts(32 * 32, 0);
      for (int o = 0; o < cout; ++o)
        for (int c = 0; c < 3; ++c)
          for (int t = 0; t < 9; ++t) {
            const int k = t * 3 + c;
            ts[(size_t)(((o / 8) * 4 + k / 8) * 64 + (o % 8) * 8 + k % 8)] =
                ay::to16((float)((double)w[((size_t)o * 3 + c) * 9 + t] / 255.0), ctx->is_fp16);
          }
      AY_CUDA(cudaMemcpy(L.w_ts, ts.data(), ts.size() * 2, cudaMemcpyHostToDevice));
    }
    AY_CUDA(cudaMemcpy(L.bias, b.data(), sizeof(float) * L.cout_rows, cudaMemcpyHostToDevice));
  } else {
    std::vector<uint16_t> r((size_t)L.cout_rows * L.ktot, 0);
    if (L.s2_packed) {
      // k-block (kh, 0) = [kw0 | kw1], k-block (kh, 1) = [kw2 | zeros], 32 channels each half
      const int kref = 9 * L.cin_k;
      std::vector<uint16_t> rr((size_t)L.cout_rows * kref, 0);
      for (int o = 0; o < cout; ++o)
        for (int c = 0; c < cin; ++c)
          for (int t = 0; t < 9; ++t) {
            const int kh = t / 3, kw = t % 3;
            const uint16_t v = ay::to16(w[((size_t)o * cin + c) * 9 + t], ctx->is_fp16);
            r[(size_t)o * L.ktot + (size_t)(kh * 2 + (kw == 2)) * 64 + (kw == 1 ? 32 : 0) + c] = v;
            rr[(size_t)o * kref + (size_t)t * L.cin_k + c] = v;
          }
      AY_CUDA(cudaMemcpy(L.w16_ref, rr.data(), rr.size() * 2, cudaMemcpyHostToDevice));
    } else {
      for (int o = 0; o < cout; ++o)
        for (int c = 0; c < cin; ++c)
          for (int t = 0; t < taps; ++t)
            r[(size_t)o * L.ktot + (size_t)t * L.cin_k + c] = ay::to16(w[((size_t)o * cin + c) * taps + t], ctx->is_fp16);
    }
    std::vector<float> b(L.cout_rows, 0.f);
    memcpy(b.data(), bias, sizeof(float) * cout);
    AY_CUDA(cudaMemcpy(L.w16, r.data(), r.size() * 2, cudaMemcpyHostToDevice));
    AY_CUDA(cudaMemcpy(L.bias, b.data(), sizeof(float) * L.cout_rows, cudaMemcpyHostToDevice));
  }
  L.weights_set = true;
  ctx->finalized = false;
  return AY_OK;
}

extern "C" int ay_net_finalize(ay_ctx* ctx) {
  AY_ENTER(ctx);
  if (!ctx->built) { set_error("ay_net_finalize before ay_net_build"); return AY_ESTATE; }
  for (size_t i = 0; i < ctx->layers.size(); ++i)
    if ((ctx->layers[i].d.type == AY_CONV || ctx->layers[i].d.type == AY_LINEAR) && !ctx->layers[i].weights_set) {
      set_error("layer %zu has no weights", i);
      return AY_ESTATE;
    }
  ctx->finalized = true;
  return AY_OK;
}

extern "C" int ay_net_output_shape(ay_ctx* ctx, int* rows, int* cols) {
  if (!ctx || !ctx->built) { set_error("network not built"); return AY_ESTATE; }
  if (rows) *rows = ctx->dec_rows;
  if (cols) *cols = ctx->dec_cols;
  return AY_OK;
}

extern "C" int ay_net_layer_shape(ay_ctx* ctx, int layer, int* c, int* h, int* w) {
  if (!ctx || !ctx->built || layer < 0 || layer >= (int)ctx->layers.size()) { set_error("bad layer"); return AY_EINVAL; }
  if (c) *c = ctx->layers[layer].C;
  if (h) *h = ctx->layers[layer].H;
  if (w) *w = ctx->layers[layer].W;
  return AY_OK;
}

extern "C" int ay_net_layer_impl(ay_ctx* ctx, int layer) {
  if (!ctx || !ctx->built || layer < 0 || layer >= (int)ctx->layers.size()) return 0;
  const Layer& L = ctx->layers[layer];
  if (L.d.type != AY_CONV) return 0;
  return (L.impl == 1 && ctx->use_ref) ? 3 : L.impl;
}

extern "C" int ay_net_launch_count(ay_ctx* ctx) { return ctx ? ctx->launches : 0; }

namespace {

// ATen's nearest-neighbour source index (UpSample.h nearest_neighbor_compute_source_index):
// min(floor(dst * float(src)/float(dst)), src-1), all in fp32.
void nearest_table(int dst, int src, std::vector<int>* out) {
  out->resize(dst);
  const float scale = (float)src / (float)dst;
  for (int i = 0; i < dst; ++i) {
    int s = (int)floorf((float)i * scale);
    (*out)[i] = s < src - 1 ? s : src - 1;
  }
}

// AY_F32 networks: the cfg blocks one by one (models.py:237-255), every tensor in fp32
int forward_f32(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w, float* dec_out,
                cudaStream_t st) {
  std::vector<Layer>& Ls = ctx->layers;
  ay::F32PrepParams pp{};
  pp.in = input; pp.in_kind = in_kind; pp.batch = batch; pp.S = ctx->img_size; pp.src_h = src_h; pp.src_w = src_w;
  pp.idx_y = ctx->idx_y; pp.idx_x = ctx->idx_x; pp.out = ctx->f_in;
  AY_TRY(ay::launch_f32_prep(pp, st));
  ++ctx->launches;
  for (int i = 0; i < (int)Ls.size(); ++i) {
    Layer& L = Ls[i];
    if (ctx->profiling) AY_CUDA(cudaEventRecord(ctx->prof_events[i], st));
    const float* prev = i ? Ls[i - 1].f_out : ctx->f_in;
    const int pC = i ? Ls[i - 1].C : 3, pH = i ? Ls[i - 1].H : ctx->img_size, pW = i ? Ls[i - 1].W : ctx->img_size;
    switch (L.d.type) {
      case AY_CONV: {
        if (L.impl == 5) {
          // [hi | lo] copy of the input rows, parity planes for stride 2, then the 3-pass tf32 GEMM
          if (L.d.stride == 2 && !getenv("AY_F32_TWO_PASS")) {
            AY_TRY(ay::launch_f32_split_s2_gather(prev, pC, L.cin, batch, pH, pW, ctx->f_planes, L.plane_rows, st));
            ++ctx->launches;
          } else {
            AY_TRY(ay::launch_f32_split(prev, pC, L.cin, ay::padded_rows_f32(batch, pH, pW), ctx->f_split, st));
            ++ctx->launches;
            if (L.d.stride == 2) {
              AY_TRY(ay::launch_f32_s2_gather(ctx->f_split, 2 * L.cin, batch, pH, pW, ctx->f_planes, L.plane_rows, st));
              ++ctx->launches;
            }
          }
          const int Wq = L.W + 2, taps = L.d.size * L.d.size;
          ay::ConvGemmParams g{};
          g.tf32 = 1; g.pad_lo = 1;
          g.P = (int)ay::padded_rows_f32(batch, L.H, L.W);
          g.m_batch = 1;
          g.num_m_tiles = (g.P + 255) / 256;
          g.num_n_tiles = L.C / L.BN;
          g.a_rows = L.a_rows;
          g.kb_per_tap = L.cin / 32;
          g.k16_all = g.k16_last = 4;
          const int tg = L.d.size == 1 ? 1 : L.d.stride == 1 ? 3 : 9;        // tap groups (A boxes) per pass
          g.n_groups = 3 * tg;
          for (int pass = 0; pass < 3; ++pass)
            for (int t = 0; t < tg; ++t) {
              const int gi = pass * tg + t;
              long long off = 0;
              int tap0 = 0;
              if (L.d.size == 3 && L.d.stride == 1) { off = (long long)(t - 1) * Wq - 1; tap0 = t * 3; }
              else if (L.d.size == 3) {
                const int kh = t / 3, kw = t % 3;
                off = (long long)((kh & 1) * 2 + (kw & 1)) * L.plane_rows + (kh >> 1) * Wq + (kw >> 1);
                tap0 = t;
              }
              if (off > 0x7fffffffLL - (1LL << 26)) { set_error("layer %d: batch too large for 32-bit row offsets", i); return AY_EINVAL; }
              g.a_row_off[gi] = (int)off;
              g.b_tap0[gi] = tap0;
              // Small terms first: the tensor core adds into its fp32 accumulator with truncation, i.e. every MMA shrinks
              // the running sum by about half an ulp OF THE SUM.  While only the two correction passes (2^-11 of the
              // result) have been added that is negligible; the A_hi x W_hi pass comes last, so a third of the MMAs
              // truncate at full magnitude instead of all of them.
              g.a_col_off[gi] = pass == 0 ? L.cin : 0;              // pass 0: A_lo x W_hi
              g.b_col_off[gi] = pass == 1 ? L.ktot : 0;             // pass 1: A_hi x W_lo; pass 2: A_hi x W_hi
            }
          (void)taps;
          g.BN = L.BN; g.num_stages = L.stages; g.nbuf = L.nbuf; g.units_per_stage = L.units_per_stage;
          g.tf32_na = ctx->tf32_na;
          g.num_acc = std::min(16, 512 / (L.BN * g.tf32_na));
          if (g.num_acc > 1) g.num_acc &= ~1;
          g.dbg = ctx->dbg;
          g.Hp = L.H + 2; g.Wp = Wq;
          g.bias = L.bias; g.leaky = L.d.leaky; g.cout = L.C;
          g.out_ptr = L.f_out; g.out_pitch = L.C;
          if (L.fused_shortcut >= 0) {
            const float* res = Ls[L.residual_src].f_out;
            if (!res) { set_error("layer %d: residual source %d has no stored output", i, L.residual_src); return AY_ESTATE; }
            g.out_ptr = Ls[L.fused_shortcut].f_out;
            g.has_res = 1; g.res_ptr = res; g.res_pitch = L.C;
          }
          g.err_flag = ctx->err_flag;
          g.pdl = 0;
          AY_TRY(ay::launch_conv_gemm(L.tmA, L.tmB, g, L.G, 2, ctx->num_sms, st));
          ++ctx->launches;
          break;
        }
        ay::F32ConvParams q{};
        q.in = prev; q.in_pitch = pC; q.Cin = L.cin; q.H = pH; q.W = pW;
        q.w = L.f_w; q.bias = L.bias; q.Cout = L.C; q.size = L.d.size; q.stride = L.d.stride; q.leaky = L.d.leaky;
        q.out = L.f_out; q.out_pitch = L.C; q.Ho = L.H; q.Wo = L.W; q.batch = batch;
        AY_TRY(ay::launch_f32_conv(q, st));
        ++ctx->launches;
        break;
      }
      case AY_MAXPOOL:
        AY_TRY(ay::launch_f32_maxpool(prev, pC, batch, pH, pW, L.d.size, L.d.stride, L.f_out, st));
        ++ctx->launches;
        break;
      case AY_UPSAMPLE:
        AY_TRY(ay::launch_f32_upsample2(prev, pC, batch, pH, pW, L.f_out, st));
        ++ctx->launches;
        break;
      case AY_ROUTE:
        if (L.d.n_src > 1) {
          int off = 0;
          for (int k = 0; k < L.d.n_src; ++k) {
            const Layer& S = Ls[L.d.src[k]];
            AY_TRY(ay::launch_f32_copy_channels(S.f_out, S.C, ay::padded_rows_f32(batch, S.H, S.W), L.f_out + off, L.C, st));
            ++ctx->launches;
            off += S.C;
          }
        }
        break;
      case AY_SHORTCUT:
        if (L.fused_into_prev) break;                       // written by the previous convolution's epilogue
        AY_TRY(ay::launch_f32_add(prev, Ls[L.d.src[0]].f_out, ay::padded_rows_f32(batch, L.H, L.W) * L.C, L.f_out, st));
        ++ctx->launches;
        break;
      case AY_YOLO: {
        ay::DecodeParams d{};
        d.raw = prev; d.raw_pitch = pC; d.batch = batch; d.G = L.H; d.n_anchors = L.d.n_anchors;
        d.dec_cols = ctx->dec_cols;
        const double stride = (double)ctx->img_size / (double)L.H;      // models.py:119
        d.stride = (float)stride;
        for (int a = 0; a < L.d.n_anchors; ++a) {
          d.anchor_w[a] = (float)((double)L.d.anchors[2 * a] / stride);  // models.py:122-124
          d.anchor_h[a] = (float)((double)L.d.anchors[2 * a + 1] / stride);
        }
        d.dec_out = dec_out; d.dec_rows_total = ctx->dec_rows; d.dec_row_off = L.dec_row_off;
        d.lo = 1; d.hi = 1;
        AY_TRY(ay::launch_decode(d, st));
        ++ctx->launches;
        break;
      }
      case AY_LINEAR:
        AY_TRY(ay::launch_linear(prev, 4, 0, pC, 1, 1, batch, L.lin_c, L.lin_h, L.lin_w, L.lin_w32, L.bias, L.C, dec_out, st));
        ++ctx->launches;
        break;
      default:
        break;
    }
  }
  if (ctx->profiling) AY_CUDA(cudaEventRecord(ctx->prof_events[Ls.size()], st));
  return AY_OK;
}

int run_conv(ay_ctx* ctx, int i, const void* input, int in_kind, int batch, int src_h, int src_w, float* dec_out,
             cudaStream_t st) {
  std::vector<Layer>& Ls = ctx->layers;
  Layer& L = Ls[i];
  const ActView* target = L.fused_shortcut >= 0 ? &Ls[L.fused_shortcut].out : &L.out;
  if (L.first) {
    ay::FirstConvParams p{};
    p.in = input; p.in_kind = in_kind; p.batch = batch; p.S = ctx->img_size; p.src_h = src_h; p.src_w = src_w;
    p.idx_y = ctx->idx_y; p.idx_x = ctx->idx_x; p.w = L.w32; p.bias = L.bias; p.cout = L.C; p.leaky = L.d.leaky;
    p.out = target->ptr; p.out_pitch = target->pitch; p.Cp = target->Cp; p.is_fp16 = ctx->is_fp16;
    ++ctx->launches;
    if (ctx->use_ref) return ay::launch_first_conv(p, st);   // fp32 CUDA-core reference of this layer
    if (L.l0_ts) {
      ay::FirstConvTsParams t{};
      t.in = input; t.in_kind = in_kind; t.batch = batch; t.S = ctx->img_size; t.src_h = src_h; t.src_w = src_w;
      t.idx_y = ctx->idx_y; t.idx_x = ctx->idx_x; t.w_ts = L.w_ts; t.bias = L.bias; t.leaky = L.d.leaky; t.is_fp16 = ctx->is_fp16;
      t.out_ptr = target->ptr; t.out_pitch = target->pitch;
      if (L.plane_consumer >= 0) {
        const Layer& Cn = Ls[L.plane_consumer];
        t.plane_mode = L.plane_mode; t.planes = Cn.planes; t.plane_rows = Cn.plane_rows;
        t.plane_packed = Cn.s2_packed ? 1 : 0; t.plane_pitch = target->Cp;
      }
      t.err_flag = ctx->err_flag;
      return ay::launch_first_conv_ts(t, ctx->num_sms, st);
    }
    p.out = L.x8;
    AY_TRY(ay::launch_prep_input(p, st));
  }
  static const ActView no_view;
  const ActView& in = L.first ? no_view : Ls[i - 1].out;
  const int Hq = ay::padded(L.H), Wq = ay::padded(L.W);
  const ActView* res = L.residual_src >= 0 ? &Ls[L.residual_src].out : nullptr;
  const Layer* Y = L.is_head ? &Ls[L.yolo_layer] : nullptr;

  if (ctx->use_ref) {
    ay::RefConvParams r{};
    r.in = in.ptr; r.in_pitch = in.pitch; r.Cin_p = L.cin_p; r.Cin_k = L.cin_k; r.H = in.H; r.W = in.W;
    r.w = L.s2_packed ? L.w16_ref : L.w16; r.ktot = L.s2_packed ? 9 * L.cin_k : L.ktot; r.bias = L.bias; r.cout_p = L.is_head ? L.cout_rows : target->Cp;
    r.size = L.d.size; r.stride = L.d.stride; r.leaky = L.d.leaky;
    r.residual = res ? res->ptr : nullptr; r.res_pitch = res ? res->pitch : 0;
    r.out = L.is_head ? nullptr : target->ptr; r.out_pitch = L.is_head ? 0 : target->pitch;
    r.Ho = L.H; r.Wo = L.W; r.batch = batch; r.is_fp16 = ctx->is_fp16;
    r.raw_f32 = L.is_head ? L.raw_f32 : nullptr;
    ++ctx->launches;
    AY_TRY(ay::launch_ref_conv(r, st));
    if (L.is_head) {
      ay::DecodeParams d{};
      d.raw = L.raw_f32; d.raw_pitch = L.cout_rows; d.batch = batch; d.G = L.H; d.n_anchors = Y->d.n_anchors;
      d.dec_cols = ctx->dec_cols; d.stride = (float)((double)ctx->img_size / (double)L.H);
      for (int a = 0; a < Y->d.n_anchors; ++a) {
        d.anchor_w[a] = (float)((double)Y->d.anchors[2 * a] / ((double)ctx->img_size / (double)L.H));
        d.anchor_h[a] = (float)((double)Y->d.anchors[2 * a + 1] / ((double)ctx->img_size / (double)L.H));
      }
      d.dec_out = dec_out; d.dec_rows_total = ctx->dec_rows; d.dec_row_off = Y->dec_row_off;
      d.lo = ay::kPadLo; d.hi = 1;
      ++ctx->launches;
      AY_TRY(ay::launch_decode(d, st));
    }
    return AY_OK;
  }

  ay::ConvGemmParams q{};
  q.P = (int)ay::padded_rows(batch, L.H, L.W);
  q.m_batch = L.m_batch;
  q.num_m_tiles = (q.P + 256 * L.m_batch - 1) / (256 * L.m_batch);
  q.num_n_tiles = L.cout_rows / L.BN;
  q.a_rows = L.a_rows;
  q.kb_per_tap = L.cin_k / 64;
  q.k16_all = q.k16_last = 4;
  if (!getenv("AY_FULL_K")) {
    if (!L.first && !L.s2_packed && L.cin_p == 32) q.k16_all = q.k16_last = 2;   // upper half of the k-block: zero weights
    if (L.s2_packed) q.k16_last = 2;                                             // [kw2 | 0]
  }
  if (L.first) {
    q.n_groups = 3;                       // one group per kh; K of a group = the 3 pixels x 8 channels of that row
    q.kb_per_tap = 1;
    for (int kh = 0; kh < 3; ++kh) {
      q.a_row_off[kh] = (kh - 1) * Wq - 1;
      q.b_tap0[kh] = kh;
    }
  } else if (L.d.size == 1) {
    q.n_groups = 1;
    q.a_row_off[0] = 0;
    q.b_tap0[0] = 0;
  } else if (L.d.stride == 1) {
    q.n_groups = 3;                       // one group per kh: the three kw taps share a 130-row A box
    for (int kh = 0; kh < 3; ++kh) {
      q.a_row_off[kh] = (kh - 1) * Wq - 1;
      q.b_tap0[kh] = kh * 3;
    }
  } else if (L.s2_packed) {
    q.n_groups = 3;                       // one group per kh: box row r = taps kw 0,1; box row r + 1 = tap kw 2
    q.kb_per_tap = 1;
    for (int kh = 0; kh < 3; ++kh) {
      const long long off = (long long)(kh & 1) * L.plane_rows + (kh >> 1) * Wq;
      if (off > 0x7fffffffLL - (1LL << 26)) { set_error("layer %d: batch too large for 32-bit row offsets", i); return AY_EINVAL; }
      q.a_row_off[kh] = (int)off;
      q.b_tap0[kh] = kh * 2;
    }
  } else {
    if (!L.gather_fused) {
      AY_TRY(ay::launch_s2_gather(in.ptr, in.pitch, L.cin_p / 8, batch, in.H, in.W, L.planes, L.plane_rows, st));
      ++ctx->launches;
    }
    q.n_groups = 9;
    for (int kh = 0; kh < 3; ++kh)
      for (int kw = 0; kw < 3; ++kw) {
        const long long off = (long long)((kh & 1) * 2 + (kw & 1)) * L.plane_rows + (kh >> 1) * Wq + (kw >> 1);
        if (off > 0x7fffffffLL - (1LL << 26)) { set_error("layer %d: batch too large for 32-bit row offsets", i); return AY_EINVAL; }
        q.a_row_off[kh * 3 + kw] = (int)off;
        q.b_tap0[kh * 3 + kw] = kh * 3 + kw;
      }
  }
  q.BN = L.BN; q.num_stages = L.stages; q.nbuf = L.nbuf; q.units_per_stage = L.units_per_stage;
  q.num_acc = std::min(16, 512 / (L.BN * L.m_batch));
  q.b_resident = L.b_resident;
  q.b_res_tiles = L.ktot / 64;
  q.a_mode = L.first ? 1 : 0;
  q.dbg = ctx->dbg;
  q.Hp = Hq; q.Wp = Wq;
  q.bias = L.bias; q.leaky = L.d.leaky;
  q.err_flag = ctx->err_flag;
  q.prof = ((ctx->profiling || ctx->stamping) && ctx->prof_counters) ? ctx->prof_counters + (size_t)i * ay::kProfWords : nullptr;
  q.pdl = ctx->pdl ? 1 : 0;
  if (L.is_head) {
    q.head = 1;
    q.head_stage_bytes = L.head_stage_bytes;
    q.cand_idx = ctx->cand_idx; q.cand_count = ctx->cand_count; q.cand_cap = ctx->cand_cap; q.conf_thres = ctx->cand_thres;
    q.head_fast = (ctx->dec_cols == 7 && Y->d.n_anchors == 3 && L.C == 21 && L.BN == 32 && L.cout_rows == 32 &&
                   L.m_batch == 1 && !ctx->no_head_fast) ? 1 : 0;
    q.dec_out = dec_out; q.dec_rows_total = ctx->dec_rows; q.dec_row_off = Y->dec_row_off;
    q.dec_cols = ctx->dec_cols; q.n_anchors = Y->d.n_anchors; q.cout = L.C;
    const double stride = (double)ctx->img_size / (double)L.H;   // models.py:119
    q.stride = (float)stride;
    for (int a = 0; a < Y->d.n_anchors; ++a) {
      q.anchor_w[a] = (float)((double)Y->d.anchors[2 * a] / stride);      // models.py:122-124
      q.anchor_h[a] = (float)((double)Y->d.anchors[2 * a + 1] / stride);
    }
  } else {
    q.has_res = res ? 1 : 0;
    q.cout = L.C;
    q.out_ptr = target->ptr; q.out_pitch = target->pitch;
    if (L.up_consumer >= 0) {
      const ActView& up = Ls[L.up_consumer].out;
      q.up_mode = 1;
      q.out_ptr = up.ptr; q.out_pitch = up.pitch;
    }
    q.res_ptr = res ? res->ptr : nullptr; q.res_pitch = res ? res->pitch : 0;
    if (L.plane_consumer >= 0) {
      const Layer& Cn = Ls[L.plane_consumer];
      q.plane_mode = L.plane_mode;
      q.planes = Cn.planes;
      q.plane_rows = Cn.plane_rows;
      q.plane_packed = Cn.s2_packed ? 1 : 0;
      q.plane_pitch = target->Cp;
    }
  }
  ++ctx->launches;
  return ay::launch_conv_gemm(L.tmA, L.tmB, q, L.G, ctx->is_fp16, ctx->num_sms, st);
}

}  // namespace

extern "C" int ay_forward(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                          float* out_decoded, void* stream) {
  AY_ENTER(ctx);
  if (!ctx->finalized) { set_error("ay_forward before ay_net_finalize"); return AY_ESTATE; }
  if (!input || !out_decoded || batch < 1 || batch > ctx->max_batch) {
    set_error("ay_forward: bad arguments (batch=%d, capacity %d)", batch, ctx->max_batch);
    return AY_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (in_kind == AY_IN_F32_NCHW) {
    if (src_h != ctx->img_size || src_w != ctx->img_size) {
      set_error("ay_forward: float input must be %dx%d (got %dx%d)", ctx->img_size, ctx->img_size, src_h, src_w);
      return AY_EINVAL;
    }
  } else if (in_kind == AY_IN_U8_HWC) {
    if (src_h != src_w || src_h < 1) { set_error("ay_forward: uint8 tiles must be square (got %dx%d)", src_h, src_w); return AY_EINVAL; }
    if (src_h != ctx->idx_src_h || src_w != ctx->idx_src_w) {
      std::vector<int> ty, tx;
      nearest_table(ctx->img_size, src_h, &ty);
      nearest_table(ctx->img_size, src_w, &tx);
      AY_CUDA(cudaStreamSynchronize(st));
      AY_CUDA(cudaMemcpy(ctx->idx_y, ty.data(), sizeof(int) * ty.size(), cudaMemcpyHostToDevice));
      AY_CUDA(cudaMemcpy(ctx->idx_x, tx.data(), sizeof(int) * tx.size(), cudaMemcpyHostToDevice));
      ctx->idx_src_h = src_h;
      ctx->idx_src_w = src_w;
    }
  } else {
    set_error("ay_forward: unknown input kind %d", in_kind);
    return AY_EINVAL;
  }
  ctx->launches = 0;
  if (ctx->is_f32) return forward_f32(ctx, input, in_kind, batch, src_h, src_w, out_decoded, st);
  std::vector<Layer>& Ls = ctx->layers;
  for (int i = 0; i < (int)Ls.size(); ++i) {
    Layer& L = Ls[i];
    if (ctx->profiling) AY_CUDA(cudaEventRecord(ctx->prof_events[i], st));
    switch (L.d.type) {
      case AY_CONV:
        AY_TRY(run_conv(ctx, i, input, in_kind, batch, src_h, src_w, out_decoded, st));
        break;
      case AY_MAXPOOL: {
        const ActView& in = Ls[i - 1].out;
        AY_TRY(ay::launch_maxpool(in.ptr, in.pitch, in.Cp, batch, in.H, in.W, L.d.size, L.d.stride, L.out.ptr,
                                  L.out.pitch, ctx->is_fp16, st));
        ++ctx->launches;
        break;
      }
      case AY_UPSAMPLE: {
        if (L.fused_into_producer && !ctx->use_ref) break;
        const ActView& in = Ls[i - 1].out;
        AY_TRY(ay::launch_upsample2(in.ptr, in.pitch, in.Cp, batch, in.H, in.W, L.out.ptr, L.out.pitch, st));
        ++ctx->launches;
        break;
      }
      case AY_ROUTE:
        for (size_t k = 0; k < L.copy_srcs.size(); ++k) {
          const ActView& s = Ls[L.copy_srcs[k]].out;
          AY_TRY(ay::launch_copy_channels(s.ptr, s.pitch, s.Cp, ay::padded_rows(batch, s.H, s.W),
                                          static_cast<uint16_t*>(L.out.ptr) + L.copy_offs[k], L.out.pitch, st));
          ++ctx->launches;
        }
        break;
      case AY_LINEAR: {
        const ActView& in = Ls[i - 1].out;
        AY_TRY(ay::launch_linear(in.ptr, 2, ctx->is_fp16, in.pitch, ay::kPadLo, 1, batch, L.lin_c, L.lin_h, L.lin_w, L.lin_w32,
                                 L.bias, L.C, out_decoded, st));
        ++ctx->launches;
        break;
      }
      case AY_SHORTCUT:
        if (!L.fused_into_prev) {
          const ActView& a = Ls[i - 1].out;
          const ActView& b = Ls[L.d.src[0]].out;
          AY_TRY(ay::launch_add(a.ptr, a.pitch, b.ptr, b.pitch, L.out.Cp, ay::padded_rows(batch, L.H, L.W), L.out.ptr,
                                L.out.pitch, ctx->is_fp16, st));
          ++ctx->launches;
        }
        break;
      default:
        break;
    }
  }
  if (ctx->profiling) AY_CUDA(cudaEventRecord(ctx->prof_events[Ls.size()], st));
  return AY_OK;
}

static int ensure_scratch(void** buf, size_t* have, size_t need) {
  if (*have >= need) return AY_OK;
  AY_CUDA(cudaDeviceSynchronize());
  if (*buf) cudaFree(*buf);
  *buf = nullptr;
  *have = 0;
  cudaError_t e = cudaMalloc(buf, need);
  if (e != cudaSuccess) { set_error("cudaMalloc(%zu bytes of scratch) failed: %s", need, cudaGetErrorString(e)); return AY_ENOMEM; }
  *have = need;
  return AY_OK;
}

extern "C" int ay_forward_candidates(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                                     float* out_decoded, float conf_thres, int32_t* cand_idx, int32_t* cand_count,
                                     int cand_cap, void* stream) {
  AY_ENTER(ctx);
  if (!cand_idx || !cand_count || cand_cap < 1) { set_error("ay_forward_candidates: bad candidate buffers"); return AY_EINVAL; }
  if (ctx->is_f32 || ctx->use_ref) {
    set_error("ay_forward_candidates: the compaction is part of the tensor-core head epilogue (bf16 / fp16 networks)");
    return AY_EINVAL;
  }
  if (batch >= 1) AY_CUDA(cudaMemsetAsync(cand_count, 0, sizeof(int32_t) * (size_t)batch, static_cast<cudaStream_t>(stream)));
  ctx->cand_idx = cand_idx; ctx->cand_count = cand_count; ctx->cand_cap = cand_cap; ctx->cand_thres = conf_thres;
  const int rc = ay_forward(ctx, input, in_kind, batch, src_h, src_w, out_decoded, stream);
  ctx->cand_idx = nullptr; ctx->cand_count = nullptr; ctx->cand_cap = 0;
  if (rc == AY_OK) ++ctx->launches;                       // the memset node
  return rc;
}

extern "C" int ay_nms_candidates(ay_ctx* ctx, float* decoded, const int32_t* cand_idx, const int32_t* cand_count,
                                 int cand_cap, int batch, int n_rows, int n_classes, float conf_thres, float nms_thres,
                                 float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream) {
  AY_ENTER(ctx);
  if (!decoded || !cand_idx || !cand_count || cand_cap < 1 || !out_rows || !out_keep || !out_counts || batch < 1 || n_rows < 1) {
    set_error("ay_nms_candidates: bad arguments");
    return AY_EINVAL;
  }
  AY_TRY(ensure_scratch(&ctx->nms_scratch, &ctx->nms_bytes, ay::nms_scratch_bytes(batch, n_rows)));
  return ay::launch_nms(decoded, batch, n_rows, n_classes, conf_thres, nms_thres, out_rows, out_keep, out_counts, cap,
                        ctx->nms_scratch, ctx->nms_bytes, ctx->err_flag, static_cast<cudaStream_t>(stream), nullptr,
                        cand_idx, cand_count, cand_cap < n_rows ? cand_cap : n_rows);
}

extern "C" int ay_debug_set(ay_ctx* ctx, int bits) {
  if (!ctx) { set_error("null context"); return AY_EINVAL; }
  ctx->dbg = bits;
  return AY_OK;
}

extern "C" int ay_forward_profile(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                                  float* out_decoded, void* stream, float* ms_per_layer_h, int n_layers) {
  AY_ENTER(ctx);
  if (!ms_per_layer_h || n_layers != (int)ctx->layers.size()) {
    set_error("ay_forward_profile: expected %d layer slots", (int)ctx->layers.size());
    return AY_EINVAL;
  }
  while (ctx->prof_events.size() < ctx->layers.size() + 1) {
    cudaEvent_t e;
    AY_CUDA(cudaEventCreate(&e));
    ctx->prof_events.push_back(e);
  }
  const size_t prof_bytes = sizeof(unsigned long long) * ay::kProfWords * ctx->layers.size();
  if (!ctx->prof_counters) AY_CUDA(cudaMalloc(&ctx->prof_counters, prof_bytes));
  AY_CUDA(cudaMemsetAsync(ctx->prof_counters, 0, prof_bytes, static_cast<cudaStream_t>(stream)));
  ctx->profiling = true;
  const int rc = ay_forward(ctx, input, in_kind, batch, src_h, src_w, out_decoded, stream);
  ctx->profiling = false;
  if (rc != AY_OK) return rc;
  AY_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  for (int i = 0; i < n_layers; ++i)
    AY_CUDA(cudaEventElapsedTime(&ms_per_layer_h[i], ctx->prof_events[i], ctx->prof_events[i + 1]));
  return AY_OK;
}

extern "C" int ay_prof_words(void) { return ay::kProfWords; }

extern "C" int ay_forward_stamps(ay_ctx* ctx, const void* input, int in_kind, int batch, int src_h, int src_w,
                                 float* out_decoded, void* stream, unsigned long long* stamps_h, int n_layers) {
  AY_ENTER(ctx);
  if (!stamps_h || n_layers != (int)ctx->layers.size()) {
    set_error("ay_forward_stamps: expected %d layer slots", (int)ctx->layers.size());
    return AY_EINVAL;
  }
  const size_t prof_bytes = sizeof(unsigned long long) * ay::kProfWords * ctx->layers.size();
  if (!ctx->prof_counters) AY_CUDA(cudaMalloc(&ctx->prof_counters, prof_bytes));
  AY_CUDA(cudaMemsetAsync(ctx->prof_counters, 0, prof_bytes, static_cast<cudaStream_t>(stream)));
  ctx->stamping = true;
  const int rc = ay_forward(ctx, input, in_kind, batch, src_h, src_w, out_decoded, stream);
  ctx->stamping = false;
  if (rc != AY_OK) return rc;
  AY_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  AY_CUDA(cudaMemcpy(stamps_h, ctx->prof_counters, prof_bytes, cudaMemcpyDeviceToHost));
  return AY_OK;
}

extern "C" int ay_debug_counters(ay_ctx* ctx, unsigned long long* out_h, int n_layers) {
  AY_ENTER(ctx);
  if (!out_h || n_layers != (int)ctx->layers.size() || !ctx->prof_counters) {
    set_error("ay_debug_counters: call ay_forward_profile first");
    return AY_ESTATE;
  }
  for (int i = 0; i < n_layers; ++i)
    AY_CUDA(cudaMemcpy(out_h + (size_t)i * 16, ctx->prof_counters + (size_t)i * ay::kProfWords, sizeof(unsigned long long) * 16,
                       cudaMemcpyDeviceToHost));
  return AY_OK;
}

extern "C" int ay_net_layer_output(ay_ctx* ctx, int layer, int batch, float* out_nchw, void* stream) {
  AY_ENTER(ctx);
  if (!ctx->built || layer < 0 || layer >= (int)ctx->layers.size() || batch < 1 || batch > ctx->max_batch) {
    set_error("ay_net_layer_output: bad arguments");
    return AY_EINVAL;
  }
  const Layer& L = ctx->layers[layer];
  if (ctx->is_f32) {
    if (!L.f_out) { set_error("layer %d has no stored output", layer); return AY_EINVAL; }
    return ay::launch_f32_dump_nchw(L.f_out, L.C, batch, L.H, L.W, out_nchw, static_cast<cudaStream_t>(stream));
  }
  if (!L.out.ptr || (!ctx->use_ref && L.up_consumer >= 0)) {
    set_error("layer %d is fused into its consumer and has no stored output", layer);
    return AY_EINVAL;
  }
  // a tensor read only by a stride-2 convolution is stored as that convolution's parity planes
  int producer = -1;
  if (L.d.type == AY_CONV && L.fused_shortcut < 0) producer = layer;
  else if (L.d.type == AY_SHORTCUT && L.fused_into_prev) producer = layer - 1;
  if (!ctx->use_ref && producer >= 0 && ctx->layers[producer].plane_mode == 2) {
    const Layer& Cn = ctx->layers[ctx->layers[producer].plane_consumer];
    return ay::launch_dump_planes_nchw(Cn.planes, Cn.plane_rows, L.out.Cp, L.C, batch, L.H, L.W, ctx->is_fp16, Cn.s2_packed ? 1 : 0, out_nchw,
                                       static_cast<cudaStream_t>(stream));
  }
  return ay::launch_dump_nchw(L.out.ptr, L.out.pitch, L.C, batch, L.H, L.W, ctx->is_fp16, out_nchw,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int ay_nms(ay_ctx* ctx, float* pred, int batch, int n_rows, int n_classes, float conf_thres,
                      float nms_thres, float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream) {
  AY_ENTER(ctx);
  if (!pred || !out_rows || !out_keep || !out_counts || batch < 1 || n_rows < 1) {
    set_error("ay_nms: bad arguments");
    return AY_EINVAL;
  }
  AY_TRY(ensure_scratch(&ctx->nms_scratch, &ctx->nms_bytes, ay::nms_scratch_bytes(batch, n_rows)));
  return ay::launch_nms(pred, batch, n_rows, n_classes, conf_thres, nms_thres, out_rows, out_keep, out_counts, cap,
                        ctx->nms_scratch, ctx->nms_bytes, ctx->err_flag, static_cast<cudaStream_t>(stream));
}

extern "C" int ay_nms_timed(ay_ctx* ctx, float* pred, int batch, int n_rows, int n_classes, float conf_thres,
                            float nms_thres, float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream,
                            float* ms4_h) {
  AY_ENTER(ctx);
  if (!pred || !out_rows || !out_keep || !out_counts || !ms4_h || batch < 1 || n_rows < 1) {
    set_error("ay_nms_timed: bad arguments");
    return AY_EINVAL;
  }
  AY_TRY(ensure_scratch(&ctx->nms_scratch, &ctx->nms_bytes, ay::nms_scratch_bytes(batch, n_rows)));
  cudaEvent_t ev[5];
  for (int i = 0; i < 5; ++i) AY_CUDA(cudaEventCreate(&ev[i]));
  int rc = ay::launch_nms(pred, batch, n_rows, n_classes, conf_thres, nms_thres, out_rows, out_keep, out_counts, cap,
                          ctx->nms_scratch, ctx->nms_bytes, ctx->err_flag, static_cast<cudaStream_t>(stream), ev);
  if (rc == AY_OK && cudaStreamSynchronize(static_cast<cudaStream_t>(stream)) != cudaSuccess) {
    set_error("ay_nms_timed: stream failed");
    rc = AY_ECUDA;
  }
  for (int i = 0; rc == AY_OK && i < 4; ++i)
    if (cudaEventElapsedTime(&ms4_h[i], ev[i], ev[i + 1]) != cudaSuccess) { set_error("ay_nms_timed: event timing failed"); rc = AY_ECUDA; }
  for (int i = 0; i < 5; ++i) cudaEventDestroy(ev[i]);
  return rc;
}

extern "C" int ay_rescale_merge(ay_ctx* ctx, const float* rows, const int32_t* counts, int batch, int cap,
                                int current_dim, int orig_h, int orig_w, int merge, float* out_rows,
                                int32_t* out_counts, void* stream) {
  AY_ENTER(ctx);
  if (!rows || !counts || !out_rows || !out_counts || batch < 1 || cap < 1 || current_dim < 1 || orig_h < 1 || orig_w < 1) {
    set_error("ay_rescale_merge: bad arguments");
    return AY_EINVAL;
  }
  AY_TRY(ensure_scratch(&ctx->merge_scratch, &ctx->merge_bytes, ay::merge_scratch_bytes(batch, cap)));
  // utils/utils.py:38-45 in double (python floats), rounded to fp32 where they meet the tensor
  const double mx = (double)std::max(orig_h, orig_w);
  const double pad_x = (double)std::max(orig_h - orig_w, 0) * ((double)current_dim / mx);
  const double pad_y = (double)std::max(orig_w - orig_h, 0) * ((double)current_dim / mx);
  const double unpad_h = (double)current_dim - pad_y, unpad_w = (double)current_dim - pad_x;
  return ay::launch_rescale_merge(rows, counts, batch, cap, (float)floor(pad_x / 2.0), (float)floor(pad_y / 2.0),
                                  (float)unpad_w, (float)unpad_h, (float)orig_w, (float)orig_h, merge, out_rows,
                                  out_counts, ctx->merge_scratch, ctx->merge_bytes, ctx->err_flag,
                                  static_cast<cudaStream_t>(stream));
}

extern "C" int ay_crop_normalize(ay_ctx* ctx, const uint8_t* tile, int height, int width, long long pitch_bytes,
                                 const int32_t* windows, int n, int crop, const float* mean3_h, const float* std3_h, float* out,
                                 void* stream) {
  AY_ENTER(ctx);
  if (!tile || !windows || !mean3_h || !std3_h || !out || height < 1 || width < 1 || pitch_bytes < 3LL * width || n < 1 || crop < 1) {
    set_error("ay_crop_normalize: bad arguments (tile %dx%d pitch %lld, %d windows of %d)", height, width, pitch_bytes, n, crop);
    return AY_EINVAL;
  }
  for (int c = 0; c < 3; ++c)
    if (!(std3_h[c] != 0.0f)) { set_error("ay_crop_normalize: std[%d] is zero", c); return AY_EINVAL; }
  return ay::launch_crop_normalize(tile, height, width, pitch_bytes, windows, n, crop, mean3_h, std3_h, out,
                                   static_cast<cudaStream_t>(stream));
}

extern "C" int ay_wsi_gather(ay_ctx* ctx, const uint8_t* wsi, int height, int width, long long pitch_bytes, int tile_size,
                             int out_size, int row0, int col0, int n_rows, int n_cols, uint8_t* out, int background,
                             void* stream) {
  AY_ENTER(ctx);
  if (!wsi || !out || height < 1 || width < 1 || pitch_bytes < 3LL * width || tile_size < 1 || out_size < 1 ||
      out_size > tile_size || row0 < 0 || col0 < 0 || n_rows < 1 || n_cols < 1 || background < 0 || background > 255) {
    set_error("ay_wsi_gather: bad arguments (slide %dx%d pitch %lld, tile %d -> %d, tiles %d..+%d x %d..+%d)", height, width,
              pitch_bytes, tile_size, out_size, row0, n_rows, col0, n_cols);
    return AY_EINVAL;
  }
  return ay::launch_wsi_gather(wsi, height, width, pitch_bytes, tile_size, out_size, row0, col0, n_rows, n_cols, out, background,
                               static_cast<cudaStream_t>(stream));
}

extern "C" int ay_bbox_iou(ay_ctx* ctx, const float* box1, int n1, const float* box2, int n2, float* out_iou,
                           void* stream) {
  AY_ENTER(ctx);
  if (!box1 || !box2 || !out_iou || (n1 != 1 && n1 != n2)) { set_error("ay_bbox_iou: bad arguments"); return AY_EINVAL; }
  return ay::launch_bbox_iou(box1, n1, box2, n2, out_iou, static_cast<cudaStream_t>(stream));
}
