// Bandwidth-bound helper kernels around the tensor-core convolution:
//   first_conv      Cin=3 first layer, with ToTensor(/255) + nearest Resize fused
//                   (utils/transforms.py:89-111, models.py:26-45 layer 0)
//   s2_gather       parity planes feeding stride-2 3x3 convolutions
//   upsample2       nearest x2 into a channel slice of a concat buffer (models.py:86-96, 244-245)
//   maxpool         2x2 s2 and 2x2 s1 (+ZeroPad right/bottom) for yolov3-tiny (models.py:47-53)
//   add             stand-alone shortcut (models.py:246-248) when it cannot be fused
//   dump_nchw       test helper: padded NHWC 16-bit -> NCHW fp32
//   ref_conv        CUDA-core direct convolution, test cross-check of the tcgen05 kernel
//   decode          stand-alone YOLO decode (models.py:137-169) used with ref_conv
//   linear          fully connected output layer of the CAA filter network (core.py:161-205)
//   crop_normalize  256^2 crops of a tile + ToTensor + Normalize for the CAA filter (core.py:109-159, 214-217)
//   wsi_gather      slide in HBM -> the 416^2 detector inputs of its 1536^2 tiles (crop.py:13-25 + the resize)
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ay_common.h"

namespace ay {
namespace {

__device__ __forceinline__ float ld16(const void* base, size_t idx, int is_fp16) {
  if (is_fp16) return __half2float(static_cast<const __half*>(base)[idx]);
  return __bfloat162float(static_cast<const __nv_bfloat16*>(base)[idx]);
}

__device__ __forceinline__ uint16_t cvt16(float v, int is_fp16) {
  if (is_fp16) {
    __half h = __float2half_rn(v);
    return *reinterpret_cast<uint16_t*>(&h);
  }
  __nv_bfloat16 b = __float2bfloat16_rn(v);
  return *reinterpret_cast<uint16_t*>(&b);
}

__device__ __forceinline__ uint32_t pack16(float lo, float hi, int is_fp16) {
  return (uint32_t)cvt16(lo, is_fp16) | ((uint32_t)cvt16(hi, is_fp16) << 16);
}

// ---------------------------------------------------------------- first conv
// One thread per padded output pixel; all output channels in registers, 32 at a time.
// Test reference of the first layer: fp32 arithmetic on the same 16-bit-rounded operands as the tensor-core path
// (pixel values 0..255, weights / 255).
__global__ void first_conv_kernel(FirstConvParams p) {
  extern __shared__ float sw[];  // [cout][27] then bias[cout]
  const int nw = p.cout * 27;
  for (int i = threadIdx.x; i < nw + p.cout; i += blockDim.x) sw[i] = (i < nw) ? p.w[i] : p.bias[i - nw];
  __syncthreads();
  const int Sp = padded(p.S);
  const long long total = (long long)p.batch * Sp * Sp;
  const long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= total) return;
  const int n = (int)(m / ((long long)Sp * Sp));
  const int pix = (int)(m - (long long)n * Sp * Sp);
  const int y = pix / Sp - kPadLo, x = pix - (pix / Sp) * Sp - kPadLo;      // unpadded output pixel
  uint16_t* orow = static_cast<uint16_t*>(p.out) + (size_t)m * p.out_pitch;
  const bool halo = (y < 0) | (y >= p.S) | (x < 0) | (x >= p.S);
  if (halo) {
    for (int c = 0; c < p.Cp; c += 8) *reinterpret_cast<uint4*>(orow + c) = make_uint4(0, 0, 0, 0);
    return;
  }
  float in[27];
#pragma unroll
  for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
    for (int kw = 0; kw < 3; ++kw) {
      const int iy = y + kh - 1, ix = x + kw - 1;
      const bool ok = (iy >= 0) & (iy < p.S) & (ix >= 0) & (ix < p.S);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float v = 0.f;
        if (ok) {
          if (p.in_kind == AY_IN_F32_NCHW) {
            v = 255.0f * __ldg(static_cast<const float*>(p.in) + (((size_t)n * 3 + c) * p.S + iy) * p.S + ix);
            const uint16_t r16 = cvt16(v, p.is_fp16);          // same rounding as the staged 16-bit input of the GEMM path
            v = ld16(&r16, 0, p.is_fp16);
          } else {
            const int sy = __ldg(p.idx_y + iy), sx = __ldg(p.idx_x + ix);
            const uint8_t u = __ldg(static_cast<const uint8_t*>(p.in) +
                                    (((size_t)n * p.src_h + sy) * p.src_w + sx) * 3 + c);
            v = (float)u;                                      // 1/255 is folded into the weights
          }
        }
        in[(kh * 3 + kw) * 3 + c] = v;
      }
    }
  }
  for (int oc0 = 0; oc0 < p.Cp; oc0 += 8) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int oc = oc0 + j;
      float a = 0.f;
      if (oc < p.cout) {
        const float* w = sw + oc * 27;
#pragma unroll
        for (int k = 0; k < 27; ++k) a = fmaf(in[k], w[k], a);
        a += sw[nw + oc];
        if (p.leaky) a = a > 0.f ? a : act_slope(p.leaky) * a;
      }
      acc[j] = a;
    }
    *reinterpret_cast<uint4*>(orow + oc0) =
        make_uint4(pack16(acc[0], acc[1], p.is_fp16), pack16(acc[2], acc[3], p.is_fp16),
                   pack16(acc[4], acc[5], p.is_fp16), pack16(acc[6], acc[7], p.is_fp16));
  }
}

// ---------------------------------------------------------------- input staging for the first convolution
// ToTensor (/255) + nearest Resize (utils/transforms.py:89-111) into a zero-haloed NHWC tensor with 8 channels
// (3 used) per pixel: 16 bytes per pixel, so that three horizontally adjacent pixels are 24 contiguous
// elements and one TMA row of the first layer's A operand (see engine.cu, first layer).
__global__ void prep_input_kernel(FirstConvParams p) {
  // grid: (x blocks, padded rows, images) - no division per pixel
  const int Sp = padded(p.S);
  const int xp = blockIdx.x * blockDim.x + threadIdx.x;
  if (xp >= Sp) return;
  const int n = blockIdx.z;
  const int y = (int)blockIdx.y - kPadLo, x = xp - kPadLo;
  const long long m = ((long long)n * Sp + blockIdx.y) * Sp + xp;
  float v[3] = {0.f, 0.f, 0.f};
  if (y >= 0 && y < p.S && x >= 0 && x < p.S) {
    if (p.in_kind == AY_IN_F32_NCHW) {
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = 255.0f * __ldg(static_cast<const float*>(p.in) + (((size_t)n * 3 + c) * p.S + y) * p.S + x);
    } else {
      const int sy = __ldg(p.idx_y + y), sx = __ldg(p.idx_x + x);
      const uint8_t* src = static_cast<const uint8_t*>(p.in) + (((size_t)n * p.src_h + sy) * p.src_w + sx) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = (float)__ldg(src + c);   // 0..255 is exact in bf16; the 1/255 of ToTensor is folded into the weights
    }
  }
  reinterpret_cast<uint4*>(p.out)[m] = make_uint4(pack16(v[0], v[1], p.is_fp16), pack16(v[2], 0.f, p.is_fp16), 0u, 0u);
}

// ---------------------------------------------------------------- stride-2 gather
// Plane pl = ph*2+pw on the consumer's padded OUTPUT grid (Hq x Wq): position (a, b) holds the input pixel with
// v = (unpadded coordinate) + 1 = (2(a - kPadLo) + ph, 2(b - kPadLo) + pw), zero when that is outside the image.
// Output tap (kh,kw) of padded output row m is then row  pl*plane_rows + m + (kh>>1)*Wq + (kw>>1).
__global__ void s2_gather_kernel(const uint4* __restrict__ in, int in_pitch8, int C8, int batch, int H, int W,
                                 uint4* __restrict__ planes, long long plane_rows) {
  const int Ho = H / 2, Wo = W / 2;
  const int Hq = padded(Ho), Wq = padded(Wo), Hp = padded(H), Wp = padded(W);
  const long long per_plane = (long long)batch * Hq * Wq * C8;
  const long long total = 4 * per_plane;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int pl = (int)(i / per_plane);
    long long r = i - (long long)pl * per_plane;
    const int c = (int)(r % C8);
    r /= C8;
    const int b = (int)(r % Wq);
    r /= Wq;
    const int a = (int)(r % Hq);
    const int n = (int)(r / Hq);
    const int iy = 2 * (a - kPadLo) + (pl >> 1) - 1, ix = 2 * (b - kPadLo) + (pl & 1) - 1;     // unpadded input pixel
    uint4 v = make_uint4(0, 0, 0, 0);
    if (iy >= 0 && iy < H && ix >= 0 && ix < W)
      v = __ldg(in + ((size_t)((size_t)n * Hp + iy + kPadLo) * Wp + ix + kPadLo) * in_pitch8 + c);
    planes[((size_t)pl * plane_rows + ((size_t)n * Hq + a) * Wq + b) * C8 + c] = v;
  }
}

// ---------------------------------------------------------------- upsample x2 (nearest)
__global__ void upsample2_kernel(const uint4* __restrict__ in, int in_pitch8, int C8, int batch, int H, int W,
                                 uint4* __restrict__ out, int out_pitch8) {
  const int Ho = 2 * H, Wo = 2 * W;
  const int Hq = padded(Ho), Wq = padded(Wo);
  const long long total = (long long)batch * Hq * Wq * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int x = (int)(r % Wq) - kPadLo;
    r /= Wq;
    const int y = (int)(r % Hq) - kPadLo;
    const int n = (int)(r / Hq);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < Ho && x >= 0 && x < Wo)
      v = __ldg(in + ((size_t)((size_t)n * padded(H) + (y >> 1) + kPadLo) * padded(W) + (x >> 1) + kPadLo) * in_pitch8 + c);
    out[((size_t)((size_t)n * Hq + y + kPadLo) * Wq + x + kPadLo) * out_pitch8 + c] = v;
  }
}

// ---------------------------------------------------------------- maxpool 2x2
// Eight channels (16 bytes) per thread: four 128-bit loads, packed 16-bit maxima (exact: a maximum needs no rounding; NaN
// propagates as in torch's max_pool2d), one 128-bit store.  The scalar form this replaces ran at a quarter of the HBM
// rate and was 46 % of the yolov3-tiny forward.
template <bool FP16>
__device__ __forceinline__ uint32_t max16x2(uint32_t a, uint32_t b) {
  if (FP16) {
    const __half2 r = __hmax2_nan(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  const __nv_bfloat162 r = __hmax2_nan(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}

template <bool FP16>
__global__ void __launch_bounds__(256)
maxpool_kernel(const uint4* __restrict__ in, int in_pitch8, int C8, int batch, int H, int W, int stride, uint4* __restrict__ out,
               int out_pitch8) {
  const int Ho = (stride == 2) ? H / 2 : H, Wo = (stride == 2) ? W / 2 : W;
  const int Hq = padded(Ho), Wq = padded(Wo), Wp = padded(W);
  const long long total = (long long)batch * Hq * Wq * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C8);
    long long r = i / C8;
    const int x = (int)(r % Wq) - kPadLo;
    r /= Wq;
    const int y = (int)(r % Hq) - kPadLo;
    const int n = (int)(r / Hq);
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (y >= 0 && y < Ho && x >= 0 && x < Wo) {
      // top-left element of the window; the zero pixel after each row / image supplies the
      // ZeroPad2d((0,1,0,1)) of the stride-1 variant (models.py:50-51)
      const int iy = y * stride, ix = x * stride;
      const size_t base = ((size_t)n * padded(H) + iy + kPadLo) * Wp + ix + kPadLo;
      const uint4 a = __ldg(in + base * in_pitch8 + c), b = __ldg(in + (base + 1) * in_pitch8 + c);
      const uint4 d = __ldg(in + (base + Wp) * in_pitch8 + c), e = __ldg(in + (base + Wp + 1) * in_pitch8 + c);
      v.x = max16x2<FP16>(max16x2<FP16>(a.x, b.x), max16x2<FP16>(d.x, e.x));
      v.y = max16x2<FP16>(max16x2<FP16>(a.y, b.y), max16x2<FP16>(d.y, e.y));
      v.z = max16x2<FP16>(max16x2<FP16>(a.z, b.z), max16x2<FP16>(d.z, e.z));
      v.w = max16x2<FP16>(max16x2<FP16>(a.w, b.w), max16x2<FP16>(d.w, e.w));
    }
    out[(((size_t)n * Hq + y + kPadLo) * Wq + x + kPadLo) * out_pitch8 + c] = v;
  }
}

// ---------------------------------------------------------------- add / copy
__global__ void add_kernel(const void* __restrict__ a, int a_pitch, const void* __restrict__ b, int b_pitch, int Cp,
                           long long rows, void* __restrict__ out, int out_pitch, int is_fp16) {
  const long long total = rows * Cp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / Cp;
    const int c = (int)(i - r * Cp);
    const float v = ld16(a, (size_t)r * a_pitch + c, is_fp16) + ld16(b, (size_t)r * b_pitch + c, is_fp16);
    static_cast<uint16_t*>(out)[(size_t)r * out_pitch + c] = cvt16(v, is_fp16);
  }
}

__global__ void copy_channels_kernel(const uint4* __restrict__ in, int in_pitch8, int C8, long long rows,
                                     uint4* __restrict__ out, int out_pitch8) {
  const long long total = rows * C8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / C8;
    const int c = (int)(i - r * C8);
    out[(size_t)r * out_pitch8 + c] = __ldg(in + (size_t)r * in_pitch8 + c);
  }
}

// ---------------------------------------------------------------- dump
__global__ void dump_nchw_kernel(const void* __restrict__ in, int in_pitch, int C, int batch, int H, int W, int is_fp16,
                                 float* __restrict__ out) {
  const long long total = (long long)batch * C * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    long long r = i / W;
    const int y = (int)(r % H);
    r /= H;
    const int c = (int)(r % C);
    const int n = (int)(r / C);
    out[i] = ld16(in, (((size_t)n * padded(H) + y + kPadLo) * padded(W) + x + kPadLo) * in_pitch + c, is_fp16);
  }
}

// same, reading the parity planes a stride-2 consumer uses (a producer may store only those)
// packed != 0: two ROW-parity planes whose rows hold the column-parity pair side by side (2 x pitch channels)
__global__ void dump_planes_nchw_kernel(const void* __restrict__ planes, long long plane_rows, int pitch, int C, int batch,
                                        int H, int W, int is_fp16, int packed, float* __restrict__ out) {
  const long long total = (long long)batch * C * H * W;
  const int Hq = padded(H / 2), Wq = padded(W / 2);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    long long r = i / W;
    const int y = (int)(r % H);
    r /= H;
    const int c = (int)(r % C);
    const int n = (int)(r / C);
    const int yp = y + 1, xp = x + 1;                       // v = unpadded coordinate + 1 (see s2_gather_kernel)
    const size_t pos = ((size_t)n * Hq + (yp >> 1) + kPadLo) * Wq + (xp >> 1) + kPadLo;
    if (packed) {
      out[i] = ld16(planes, ((size_t)(yp & 1) * plane_rows + pos) * (2 * pitch) + (xp & 1) * pitch + c, is_fp16);
    } else {
      const size_t row = (size_t)((yp & 1) * 2 + (xp & 1)) * plane_rows + pos;
      out[i] = ld16(planes, row * pitch + c, is_fp16);
    }
  }
}

// ---------------------------------------------------------------- reference conv (CUDA cores)
__global__ void ref_conv_kernel(RefConvParams p) {
  const int Hq = padded(p.Ho), Wq = padded(p.Wo), Hp = padded(p.H), Wp = padded(p.W);
  const long long total = (long long)p.batch * Hq * Wq * p.cout_p;
  const int pad = (p.size - 1) / 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int oc = (int)(i % p.cout_p);
    long long m = i / p.cout_p;
    const int x = (int)(m % Wq) - kPadLo;                 // unpadded output pixel
    const int y = (int)((m / Wq) % Hq) - kPadLo;
    const int n = (int)(m / ((long long)Wq * Hq));
    const bool halo = (y < 0) | (y >= p.Ho) | (x < 0) | (x >= p.Wo);
    float acc = 0.f;
    if (!halo) {
      for (int kh = 0; kh < p.size; ++kh)
        for (int kw = 0; kw < p.size; ++kw) {
          const int iy = y * p.stride + kh - pad, ix = x * p.stride + kw - pad;      // unpadded input pixel
          if (iy < 0 || iy >= p.H || ix < 0 || ix >= p.W) continue;
          const size_t irow = ((size_t)n * Hp + iy + kPadLo) * Wp + ix + kPadLo;
          const size_t wrow = (size_t)oc * p.ktot + (size_t)(kh * p.size + kw) * p.Cin_k;
          for (int c = 0; c < p.Cin_p; ++c)
            acc = fmaf(ld16(p.in, irow * p.in_pitch + c, p.is_fp16), ld16(p.w, wrow + c, p.is_fp16), acc);
        }
      acc += p.bias[oc];
      if (p.leaky) acc = acc > 0.f ? acc : act_slope(p.leaky) * acc;
      if (p.residual) acc += ld16(p.residual, (size_t)m * p.res_pitch + oc, p.is_fp16);
    }
    if (p.raw_f32) p.raw_f32[(size_t)m * p.cout_p + oc] = acc;
    else static_cast<uint16_t*>(p.out)[(size_t)m * p.out_pitch + oc] = cvt16(acc, p.is_fp16);
  }
}

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void decode_kernel(DecodeParams p) {
  const int Gq = p.G + p.lo + p.hi;
  const int cout = p.n_anchors * p.dec_cols;
  const long long total = (long long)p.batch * p.G * p.G * cout;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(i % cout);
    long long r = i / cout;
    const int gx = (int)(r % p.G);
    const int gy = (int)((r / p.G) % p.G);
    const int n = (int)(r / ((long long)p.G * p.G));
    const size_t m = ((size_t)n * Gq + gy + p.lo) * Gq + gx + p.lo;
    const float t = p.raw[m * p.raw_pitch + ch];
    const int a = ch / p.dec_cols, f = ch - a * p.dec_cols;
    float v;
    if (f == 0) v = __fmul_rn(__fadd_rn(sigmoidf_acc(t), (float)gx), p.stride);
    else if (f == 1) v = __fmul_rn(__fadd_rn(sigmoidf_acc(t), (float)gy), p.stride);
    else if (f == 2) v = __fmul_rn(__fmul_rn(expf(t), p.anchor_w[a]), p.stride);
    else if (f == 3) v = __fmul_rn(__fmul_rn(expf(t), p.anchor_h[a]), p.stride);
    else v = sigmoidf_acc(t);
    p.dec_out[(((size_t)n * p.dec_rows_total + p.dec_row_off) + (size_t)a * p.G * p.G + (size_t)gy * p.G + gx) *
                  p.dec_cols + f] = v;
  }
}

inline int grid_for(long long total, int block = 256) {
  long long g = (total + block - 1) / block;
  const long long cap = 148LL * 32;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

int launch_first_conv(const FirstConvParams& p, cudaStream_t s) {
  const long long total = (long long)p.batch * padded(p.S) * padded(p.S);
  const int block = 128;
  const long long grid = (total + block - 1) / block;
  const size_t smem = (size_t)(p.cout * 28) * sizeof(float);
  first_conv_kernel<<<(unsigned)grid, block, smem, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_prep_input(const FirstConvParams& p, cudaStream_t s) {
  const int Sp = padded(p.S);
  if (p.batch > 65535 || Sp > 65535) { set_error("prep_input: batch %d / size %d beyond the launch grid", p.batch, p.S); return AY_EINVAL; }
  prep_input_kernel<<<dim3((Sp + 127) / 128, Sp, p.batch), 128, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_s2_gather(const void* in, int in_pitch, int C8, int batch, int H, int W, void* planes,
                     long long plane_rows, cudaStream_t s) {
  const long long total = 4LL * batch * padded(H / 2) * padded(W / 2) * C8;
  s2_gather_kernel<<<grid_for(total), 256, 0, s>>>(static_cast<const uint4*>(in), in_pitch / 8, C8, batch, H, W,
                                                   static_cast<uint4*>(planes), plane_rows);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_upsample2(const void* in, int in_pitch, int Cp, int batch, int H, int W, void* out, int out_pitch,
                     cudaStream_t s) {
  const long long total = (long long)batch * padded(2 * H) * padded(2 * W) * (Cp / 8);
  upsample2_kernel<<<grid_for(total), 256, 0, s>>>(static_cast<const uint4*>(in), in_pitch / 8, Cp / 8, batch, H, W,
                                                   static_cast<uint4*>(out), out_pitch / 8);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_maxpool(const void* in, int in_pitch, int Cp, int batch, int H, int W, int size, int stride, void* out,
                   int out_pitch, int is_fp16, cudaStream_t s) {
  if (size != 2 || (stride != 1 && stride != 2) || (stride == 2 && ((H | W) & 1))) {
    set_error("maxpool: only 2x2 stride 1/2 on even sizes is supported (size=%d stride=%d %dx%d)", size, stride, H, W);
    return AY_EINVAL;
  }
  const int Ho = stride == 2 ? H / 2 : H, Wo = stride == 2 ? W / 2 : W;
  if ((Cp | in_pitch | out_pitch) & 7) { set_error("maxpool: channel counts and pitches must be multiples of 8"); return AY_EINVAL; }
  const long long total = (long long)batch * padded(Ho) * padded(Wo) * (Cp / 8);
  if (is_fp16)
    maxpool_kernel<true><<<grid_for(total), 256, 0, s>>>(static_cast<const uint4*>(in), in_pitch / 8, Cp / 8, batch, H, W, stride,
                                                         static_cast<uint4*>(out), out_pitch / 8);
  else
    maxpool_kernel<false><<<grid_for(total), 256, 0, s>>>(static_cast<const uint4*>(in), in_pitch / 8, Cp / 8, batch, H, W, stride,
                                                          static_cast<uint4*>(out), out_pitch / 8);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_add(const void* a, int a_pitch, const void* b, int b_pitch, int Cp, long long rows, void* out,
               int out_pitch, int is_fp16, cudaStream_t s) {
  add_kernel<<<grid_for(rows * Cp), 256, 0, s>>>(a, a_pitch, b, b_pitch, Cp, rows, out, out_pitch, is_fp16);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_copy_channels(const void* in, int in_pitch, int Cp, long long rows, void* out, int out_pitch,
                         cudaStream_t s) {
  copy_channels_kernel<<<grid_for(rows * (Cp / 8)), 256, 0, s>>>(static_cast<const uint4*>(in), in_pitch / 8, Cp / 8,
                                                                 rows, static_cast<uint4*>(out), out_pitch / 8);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_dump_nchw(const void* in, int in_pitch, int C, int batch, int H, int W, int is_fp16, float* out,
                     cudaStream_t s) {
  dump_nchw_kernel<<<grid_for((long long)batch * C * H * W), 256, 0, s>>>(in, in_pitch, C, batch, H, W, is_fp16, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_dump_planes_nchw(const void* planes, long long plane_rows, int pitch, int C, int batch, int H, int W,
                            int is_fp16, int packed, float* out, cudaStream_t s) {
  dump_planes_nchw_kernel<<<grid_for((long long)batch * C * H * W), 256, 0, s>>>(planes, plane_rows, pitch, C, batch, H, W,
                                                                               is_fp16, packed, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_ref_conv(const RefConvParams& p, cudaStream_t s) {
  const long long total = (long long)p.batch * padded(p.Ho) * padded(p.Wo) * p.cout_p;
  ref_conv_kernel<<<grid_for(total), 256, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

// ---------------------------------------------------------------- fully connected output layer (AY_LINEAR)
// out[b, o] = bias[o] + sum_{y,x,c} act[b, y, x, c] * w[o, (y*W + x)*C + c]   (core.py:161-205: x.view(B, -1) + nn.Linear,
// the weight columns reordered from (c, y, x) to (y, x, c) by ay_net_set_conv).  One block per (tile, output); the
// activation is a zero-haloed NHWC tensor of 16-bit (elem_bytes 2) or fp32 (4) elements with `lo` halo pixels before
// and `hi` after every row / image.  A few thousand multiply-adds per output: the time of a 6 KB read, not a GEMM.
namespace {
__global__ void __launch_bounds__(128)
linear_kernel(const void* __restrict__ act, int elem_bytes, int is_fp16, int pitch, int lo, int hi, int C, int H, int W,
              const float* __restrict__ w, const float* __restrict__ bias, int O, float* __restrict__ out) {
  const int b = blockIdx.y, o = blockIdx.x;
  const int Hp = H + lo + hi, Wp = W + lo + hi;
  const int n = H * W * C;
  const float* wr = w + (size_t)o * n;
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int c = i % C, px = i / C;
    const int y = px / W, x = px - y * W;
    const size_t row = ((size_t)b * Hp + y + lo) * Wp + x + lo;
    const float a = elem_bytes == 4 ? static_cast<const float*>(act)[row * pitch + c] : ld16(act, row * pitch + c, is_fp16);
    acc = fmaf(a, wr[i], acc);
  }
  __shared__ float part[4];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) out[(size_t)b * O + o] = ((part[0] + part[1]) + (part[2] + part[3])) + bias[o];
}
}  // namespace

int launch_linear(const void* act, int elem_bytes, int is_fp16, int pitch, int lo, int hi, int batch, int C, int H, int W,
                  const float* w, const float* bias, int O, float* out, cudaStream_t s) {
  if (batch < 1 || O < 1 || C < 1 || H < 1 || W < 1) { set_error("linear: bad shape"); return AY_EINVAL; }
  linear_kernel<<<dim3(O, batch), 128, 0, s>>>(act, elem_bytes, is_fp16, pitch, lo, hi, C, H, W, w, bias, O, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

// ---------------------------------------------------------------- slide -> detector input
// The reference cuts a slide into 1536^2 JPEG tiles (crop.py:13-25, dzsave layout google: tile (row, col) starts at
// (row*T, col*T), edge tiles padded with the white background) and later resizes every tile to 416^2 with nearest
// sampling (utils/transforms.py:104-111).  With the slide resident in HBM both steps are one gather: output pixel
// (y, x) of tile (row, col) is slide pixel (row*T + idx(y), col*T + idx(x)), idx(i) = min(floor(i * float(T)/float(S)), T-1)
// (ATen's nearest index, as engine.cu nearest_table), `background` outside the slide.  Reads (S/T)^2 = 7 % of the slide.
namespace {
__global__ void __launch_bounds__(256)
wsi_gather_kernel(const uint8_t* __restrict__ wsi, int height, int width, long long pitch, int T, int S, int row0, int col0,
                  int n_rows, int n_cols, uint8_t* __restrict__ out, int background) {
  const long long total = (long long)n_rows * n_cols * S * S;
  const float scale = (float)T / (float)S;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % S), y = (int)((i / S) % S);
    const int t = (int)(i / ((long long)S * S));
    const int row = row0 + t / n_cols, col = col0 + t % n_cols;
    int sy = (int)floorf(__fmul_rn((float)y, scale)), sx = (int)floorf(__fmul_rn((float)x, scale));
    sy = sy < T - 1 ? sy : T - 1;
    sx = sx < T - 1 ? sx : T - 1;
    const long long gy = (long long)row * T + sy, gx = (long long)col * T + sx;
    uint8_t r = (uint8_t)background, g = r, b = r;
    if (gy < height && gx < width) {
      const uint8_t* src = wsi + gy * pitch + gx * 3;
      r = __ldg(src); g = __ldg(src + 1); b = __ldg(src + 2);
    }
    uint8_t* o = out + i * 3;
    o[0] = r; o[1] = g; o[2] = b;
  }
}
}  // namespace

int launch_wsi_gather(const uint8_t* wsi, int height, int width, long long pitch, int T, int S, int row0, int col0, int n_rows,
                      int n_cols, uint8_t* out, int background, cudaStream_t s) {
  const long long total = (long long)n_rows * n_cols * S * S;
  wsi_gather_kernel<<<grid_for(total), 256, 0, s>>>(wsi, height, width, pitch, T, S, row0, col0, n_rows, n_cols, out, background);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

// ---------------------------------------------------------------- CAA filter input (core.py:109-159, 214-217, 435)
// get256Img + ToTensor + Normalize for all boxes of a tile at once: window w = (x_left, x_right, y_top, y_bottom) of the
// uint8 RGB tile -> out[w, c, y, x] = (tile[y_top + y, x_left + x, c] / 255 - mean[c]) / std[c], every operation one fp32
// rounding in torchvision's order (div, sub, div), so the network sees the reference's bits.
namespace {
__global__ void __launch_bounds__(256)
crop_normalize_kernel(const uint8_t* __restrict__ tile, int height, int width, long long pitch, const int32_t* __restrict__ windows,
                      int n, int crop, float m0, float m1, float m2, float s0, float s1, float s2, float* __restrict__ out) {
  const long long total = (long long)n * crop * crop;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % crop), y = (int)((i / crop) % crop), w = (int)(i / ((long long)crop * crop));
    const int gx = __ldg(windows + 4 * w) + x, gy = __ldg(windows + 4 * w + 2) + y;
    float r = 0.f, g = 0.f, b = 0.f;
    if (gx >= 0 && gx < width && gy >= 0 && gy < height) {
      const uint8_t* src = tile + gy * pitch + (long long)gx * 3;
      r = (float)__ldg(src); g = (float)__ldg(src + 1); b = (float)__ldg(src + 2);
    }
    float* o = out + (size_t)w * 3 * crop * crop + (size_t)y * crop + x;
    o[0] = __fdiv_rn(__fsub_rn(__fdiv_rn(r, 255.0f), m0), s0);
    o[(size_t)crop * crop] = __fdiv_rn(__fsub_rn(__fdiv_rn(g, 255.0f), m1), s1);
    o[(size_t)2 * crop * crop] = __fdiv_rn(__fsub_rn(__fdiv_rn(b, 255.0f), m2), s2);
  }
}
}  // namespace

int launch_crop_normalize(const uint8_t* tile, int height, int width, long long pitch, const int32_t* windows, int n, int crop,
                          const float* mean3, const float* std3, float* out, cudaStream_t s) {
  const long long total = (long long)n * crop * crop;
  crop_normalize_kernel<<<grid_for(total), 256, 0, s>>>(tile, height, width, pitch, windows, n, crop, mean3[0], mean3[1], mean3[2],
                                                        std3[0], std3[1], std3[2], out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_decode(const DecodeParams& p, cudaStream_t s) {
  const long long total = (long long)p.batch * p.G * p.G * p.n_anchors * p.dec_cols;
  decode_kernel<<<grid_for(total), 256, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
