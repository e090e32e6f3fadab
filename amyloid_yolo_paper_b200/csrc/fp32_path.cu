// fp32 accuracy mode of the tile-detection engine (act_dtype = AY_F32).
//
// The tensor-core path stores activations in 16 bits, which bounds its agreement with the reference at the
// 2e-2 bar.  This mode keeps every activation and every weight in fp32 and accumulates with FFMA on the CUDA
// cores, for the 1e-4 bar on the raw head outputs (the reference computes models.py:33-45, 86-96, 244-248 in
// fp32 through ATen).  It is the accuracy mode, not the headline path: no fusion, one kernel per cfg block.
//
// Layout: the same zero-haloed NHWC rows as the 16-bit path (ay_common.h), fp32 elements, pitch = C (no channel
// padding):  row(n, y, x) = (n*(H+2) + y)*(W+2) + x.  Kernels write interior pixels only; the halo stays at the
// zeros the buffers were allocated with, so 3x3 taps and the ZeroPad2d of the stride-1 maxpool read zeros.
#include "ay_common.h"

#include <cstdlib>

namespace ay {
namespace {

constexpr int kBM = 64, kBN = 64, kBK = 16;

inline int grid_for32(long long total, int block = 256) {
  long long g = (total + block - 1) / block;
  const long long cap = 148LL * 32;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ToTensor (u8 / 255, utils/transforms.py:89-101) + nearest Resize (transforms.py:104-111), or a copy of the
// caller's [B,3,S,S] floats, into [B,S+2,S+2,3].
__global__ void f32_prep_kernel(F32PrepParams p) {
  const long long total = (long long)p.batch * p.S * p.S;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % p.S);
    const int y = (int)((i / p.S) % p.S);
    const int n = (int)(i / ((long long)p.S * p.S));
    float v[3];
    if (p.in_kind == AY_IN_F32_NCHW) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        v[c] = __ldg(static_cast<const float*>(p.in) + (((size_t)n * 3 + c) * p.S + y) * p.S + x);
    } else {
      const int sy = __ldg(p.idx_y + y), sx = __ldg(p.idx_x + x);
      const uint8_t* src = static_cast<const uint8_t*>(p.in) + (((size_t)n * p.src_h + sy) * p.src_w + sx) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = __fdiv_rn((float)__ldg(src + c), 255.0f);
    }
    float* o = p.out + (((size_t)n * (p.S + 2) + y + 1) * (p.S + 2) + x + 1) * 3;
    o[0] = v[0]; o[1] = v[1]; o[2] = v[2];
  }
}

// First layer of the fp32 mode (3 input channels, 3x3, stride 1, up to 64 filters): one thread per output pixel, its 27
// inputs in registers, the [27][Cout] weights in shared memory read four filters at a time as broadcasts.  The generic
// kernel below spends this layer on scalar gathers with a division per element and on the unused half of its 64-filter tile
// (3.8 ms of a 35 ms forward at B = 64).  Each output is one 27-term FMA chain plus the bias: K is too short to need the
// compensated two-level sum.
__global__ void __launch_bounds__(128) f32_first_conv_kernel(F32ConvParams p) {
  __shared__ __align__(16) float ws[27 * 64];
  __shared__ float bs[64];
  for (int i = threadIdx.x; i < 27 * p.Cout; i += blockDim.x) ws[i] = __ldg(p.w + i);
  for (int i = threadIdx.x; i < p.Cout; i += blockDim.x) bs[i] = __ldg(p.bias + i);
  __syncthreads();
  const int ox = blockIdx.x * blockDim.x + threadIdx.x;
  if (ox >= p.Wo) return;
  const int oy = blockIdx.y, n = blockIdx.z;
  const int Wq = p.W + 2;
  float in[27];
#pragma unroll
  for (int kh = 0; kh < 3; ++kh) {
    const float* row = p.in + (((size_t)n * (p.H + 2) + oy + kh) * Wq + ox) * p.in_pitch;   // padded (oy + kh, ox + kw)
#pragma unroll
    for (int j = 0; j < 9; ++j) in[kh * 9 + j] = __ldg(row + (j / 3) * p.in_pitch + (j % 3));
  }
  float* o = p.out + (((size_t)n * (p.Ho + 2) + oy + 1) * (p.Wo + 2) + ox + 1) * p.out_pitch;
  for (int c = 0; c < p.Cout; c += 4) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      const float4 w = *reinterpret_cast<const float4*>(&ws[k * p.Cout + c]);
      a0 = fmaf(in[k], w.x, a0); a1 = fmaf(in[k], w.y, a1); a2 = fmaf(in[k], w.z, a2); a3 = fmaf(in[k], w.w, a3);
    }
    float v[4] = {__fadd_rn(a0, bs[c]), __fadd_rn(a1, bs[c + 1]), __fadd_rn(a2, bs[c + 2]), __fadd_rn(a3, bs[c + 3])};
    if (p.leaky) {
#pragma unroll
      for (int e = 0; e < 4; ++e) v[e] = v[e] > 0.f ? v[e] : __fmul_rn(v[e], act_slope(p.leaky));
    }
    *reinterpret_cast<float4*>(o + c) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// Implicit GEMM on the CUDA cores: a 64-pixel x 64-filter tile per CTA, 4 x 4 outputs per thread, K = taps * Cin
// in steps of 16 through shared memory.  K order is (kh, kw, c), the weights are [K][Cout].
__global__ void __launch_bounds__(256) f32_conv_kernel(F32ConvParams p) {
  __shared__ __align__(16) float As[kBK][kBM + 4];
  __shared__ __align__(16) float Bs[kBK][kBN + 4];
  const int tid = threadIdx.x;
  const int Wq = p.W + 2;
  const int pad = (p.size - 1) / 2;
  const long long total = (long long)p.batch * p.Ho * p.Wo;
  const long long m0 = (long long)blockIdx.x * kBM;
  const int n0 = blockIdx.y * kBN;
  const int K = p.size * p.size * p.Cin;

  // operand loads: this thread fetches 4 consecutive k of one pixel and 4 consecutive filters of one k
  const int a_pix = tid >> 2, a_k = (tid & 3) * 4;
  long long am = m0 + a_pix;
  if (am >= total) am = total - 1;
  const int a_ox = (int)(am % p.Wo);
  const int a_oy = (int)((am / p.Wo) % p.Ho);
  const int a_n = (int)(am / ((long long)p.Wo * p.Ho));
  const size_t a_base = ((size_t)a_n * (p.H + 2) + a_oy * p.stride + 1 - pad) * Wq + a_ox * p.stride + 1 - pad;
  const int b_k = tid >> 4, b_c = (tid & 15) * 4;
  const bool a_vec = (p.Cin & 3) == 0 && (p.in_pitch & 3) == 0;
  const bool b_vec = (p.Cout & 3) == 0;

  const int ty = tid >> 4, tx = tid & 15;
  // Two-level summation: the 16 products of a K step are chained with FMA from zero, the K/16 step sums are added
  // with Kahan compensation — the result carries the rounding of a 16-term sum instead of a K-term one (K up to 4608),
  // which is what keeps this mode inside the 1e-4 elementwise bar against a float64 evaluation.
  float acc[4][4], comp[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = comp[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += kBK) {
    {
      const int k = k0 + a_k;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (a_vec) {
        if (k < K) {
          const int tap = k / p.Cin, c = k - tap * p.Cin;
          const int kh = tap / p.size, kw = tap - kh * p.size;
          const float4 q = __ldg(reinterpret_cast<const float4*>(p.in + (a_base + (size_t)kh * Wq + kw) * p.in_pitch + c));
          v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        }
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int kk = k + e;
          if (kk < K) {
            const int tap = kk / p.Cin, c = kk - tap * p.Cin;
            const int kh = tap / p.size, kw = tap - kh * p.size;
            v[e] = __ldg(p.in + (a_base + (size_t)kh * Wq + kw) * p.in_pitch + c);
          }
        }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) As[a_k + e][a_pix] = v[e];
    }
    {
      const int k = k0 + b_k, c = n0 + b_c;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (k < K) {
        if (b_vec && c + 3 < p.Cout) {
          const float4 q = __ldg(reinterpret_cast<const float4*>(p.w + (size_t)k * p.Cout + c));
          v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (c + e < p.Cout) v[e] = __ldg(p.w + (size_t)k * p.Cout + c + e);
        }
      }
      *reinterpret_cast<float4*>(&Bs[b_k][b_c]) = make_float4(v[0], v[1], v[2], v[3]);
    }
    __syncthreads();
    float part[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) part[i][j] = 0.f;
#pragma unroll
    for (int kk = 0; kk < kBK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) part[i][j] = fmaf(av[i], bv[j], part[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float y = __fsub_rn(part[i][j], comp[i][j]);
        const float t = __fadd_rn(acc[i][j], y);
        comp[i][j] = __fsub_rn(__fsub_rn(t, acc[i][j]), y);
        acc[i][j] = t;
      }
    __syncthreads();
  }

  // bias, LeakyReLU(0.1) (models.py:45), interior rows of the padded output
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= total) continue;
    const int ox = (int)(m % p.Wo);
    const int oy = (int)((m / p.Wo) % p.Ho);
    const int n = (int)(m / ((long long)p.Wo * p.Ho));
    float* o = p.out + (((size_t)n * (p.Ho + 2) + oy + 1) * (p.Wo + 2) + ox + 1) * p.out_pitch;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + tx * 4 + j;
      if (c >= p.Cout) continue;
      float v = __fadd_rn(acc[i][j], __ldg(p.bias + c));
      if (p.leaky) v = v > 0.f ? v : __fmul_rn(v, act_slope(p.leaky));
      o[c] = v;
    }
  }
}

// 2x2 max pooling, stride 2, or stride 1 behind ZeroPad2d((0,1,0,1)) (models.py:47-53): the halo is that padding
__global__ void f32_maxpool_kernel(const float* __restrict__ in, int C, int batch, int H, int W, int stride,
                                   float* __restrict__ out) {
  const int Ho = stride == 2 ? H / 2 : H, Wo = stride == 2 ? W / 2 : W;
  const long long total = (long long)batch * Ho * Wo * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long r = i / C;
    const int x = (int)(r % Wo);
    r /= Wo;
    const int y = (int)(r % Ho);
    const int n = (int)(r / Ho);
    const size_t base = ((size_t)n * (H + 2) + y * stride + 1) * (W + 2) + x * stride + 1;
    const float a = in[base * C + c], b = in[(base + 1) * C + c];
    const float d = in[(base + W + 2) * C + c], e = in[(base + W + 3) * C + c];
    out[(((size_t)n * (Ho + 2) + y + 1) * (Wo + 2) + x + 1) * C + c] = fmaxf(fmaxf(a, b), fmaxf(d, e));
  }
}

// nearest x2 (models.py:86-96): out[y][x] = in[y >> 1][x >> 1]
__global__ void f32_upsample2_kernel(const float* __restrict__ in, int C, int batch, int H, int W,
                                     float* __restrict__ out) {
  const int Ho = 2 * H, Wo = 2 * W;
  const long long total = (long long)batch * Ho * Wo * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long r = i / C;
    const int x = (int)(r % Wo);
    r /= Wo;
    const int y = (int)(r % Ho);
    const int n = (int)(r / Ho);
    out[(((size_t)n * (Ho + 2) + y + 1) * (Wo + 2) + x + 1) * C + c] =
        in[(((size_t)n * (H + 2) + (y >> 1) + 1) * (W + 2) + (x >> 1) + 1) * C + c];
  }
}

// shortcut (models.py:246-248): out = a + b over whole padded rows (halo: 0 + 0)
__global__ void f32_add_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n,
                               float* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = __fadd_rn(a[i], b[i]);
}

// route (models.py:244-245): source channels into their slice of the concatenated rows
__global__ void f32_copy_channels_kernel(const float* __restrict__ in, int C, long long rows, float* __restrict__ out,
                                         int out_pitch) {
  const long long total = rows * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / C;
    out[r * out_pitch + (i - r * C)] = in[i];
  }
}

// Operand split for the tensor-core convolution of this mode (conv_gemm_sm100.cu, TF32): hi = v rounded to tf32
// (10-bit mantissa, round to nearest, ties away), lo = v - hi, which is exact in fp32.  The MMA ignores the low 13
// mantissa bits of its operands, so hi passes through unchanged and lo loses only its own 2^-11 tail: hi*W_hi + lo*W_hi
// + hi*W_lo reproduces the fp32 product to ~2^-22 relative.  One row of C floats in, [hi | lo] = 2C floats out.
__device__ __forceinline__ float tf32_round(float v) {
  return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u);
}
__global__ void f32_split_kernel(const float4* __restrict__ in, int in_pitch4, int C4, long long rows,
                                 float4* __restrict__ out) {
  const long long total = rows * C4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / C4;
    const int c = (int)(i - r * C4);
    const float4 v = __ldg(in + r * in_pitch4 + c);
    float4 h, l;
    h.x = tf32_round(v.x); h.y = tf32_round(v.y); h.z = tf32_round(v.z); h.w = tf32_round(v.w);
    l.x = __fsub_rn(v.x, h.x); l.y = __fsub_rn(v.y, h.y); l.z = __fsub_rn(v.z, h.z); l.w = __fsub_rn(v.w, h.w);
    out[r * 2 * C4 + c] = h;
    out[r * 2 * C4 + C4 + c] = l;
  }
}

// Parity planes for the stride-2 tensor-core convolutions of this mode (classic one-pixel frame): plane pl = ph*2+pw,
// position (a, b) of the output's padded grid holds padded input pixel (2(a-1)+ph, 2(b-1)+pw), zero outside.
__global__ void f32_s2_gather_kernel(const float4* __restrict__ in, int P4, int batch, int H, int W,
                                     float4* __restrict__ planes, long long plane_rows) {
  const int Ho = H / 2, Wo = W / 2;
  const int Hq = Ho + 2, Wq = Wo + 2, Hp = H + 2, Wp = W + 2;
  const long long per_plane = (long long)batch * Hq * Wq * P4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < 4 * per_plane;
       i += (long long)gridDim.x * blockDim.x) {
    const int pl = (int)(i / per_plane);
    long long r = i - (long long)pl * per_plane;
    const int c = (int)(r % P4);
    r /= P4;
    const int b = (int)(r % Wq);
    r /= Wq;
    const int a = (int)(r % Hq);
    const int n = (int)(r / Hq);
    const int iy = 2 * (a - 1) + (pl >> 1), ix = 2 * (b - 1) + (pl & 1);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (iy >= 0 && iy < Hp && ix >= 0 && ix < Wp) v = __ldg(in + ((size_t)((size_t)n * Hp + iy) * Wp + ix) * P4 + c);
    planes[((size_t)pl * plane_rows + ((size_t)n * Hq + a) * Wq + b) * P4 + c] = v;
  }
}

// The two passes above in one for stride-2 layers: raw fp32 rows in, [hi | lo] parity planes out (one read of the input
// and one write of the planes instead of a split copy in between).
__global__ void f32_split_s2_gather_kernel(const float4* __restrict__ in, int in_pitch4, int C4, int batch, int H, int W,
                                           float4* __restrict__ planes, long long plane_rows) {
  const int Ho = H / 2, Wo = W / 2;
  const int Hq = Ho + 2, Wq = Wo + 2, Hp = H + 2, Wp = W + 2;
  const long long per_plane = (long long)batch * Hq * Wq * C4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < 4 * per_plane;
       i += (long long)gridDim.x * blockDim.x) {
    const int pl = (int)(i / per_plane);
    long long r = i - (long long)pl * per_plane;
    const int c = (int)(r % C4);
    r /= C4;
    const int b = (int)(r % Wq);
    r /= Wq;
    const int a = (int)(r % Hq);
    const int n = (int)(r / Hq);
    const int iy = 2 * (a - 1) + (pl >> 1), ix = 2 * (b - 1) + (pl & 1);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (iy >= 0 && iy < Hp && ix >= 0 && ix < Wp) v = __ldg(in + ((size_t)((size_t)n * Hp + iy) * Wp + ix) * in_pitch4 + c);
    float4 h, l;
    h.x = tf32_round(v.x); h.y = tf32_round(v.y); h.z = tf32_round(v.z); h.w = tf32_round(v.w);
    l.x = __fsub_rn(v.x, h.x); l.y = __fsub_rn(v.y, h.y); l.z = __fsub_rn(v.z, h.z); l.w = __fsub_rn(v.w, h.w);
    float4* o = planes + ((size_t)pl * plane_rows + ((size_t)n * Hq + a) * Wq + b) * 2 * C4;
    o[c] = h;
    o[C4 + c] = l;
  }
}

__global__ void f32_dump_nchw_kernel(const float* __restrict__ in, int C, int batch, int H, int W,
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
    out[i] = in[(((size_t)n * (H + 2) + y + 1) * (W + 2) + x + 1) * C + c];
  }
}

}  // namespace

int launch_f32_prep(const F32PrepParams& p, cudaStream_t s) {
  f32_prep_kernel<<<grid_for32((long long)p.batch * p.S * p.S), 256, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_conv(const F32ConvParams& p, cudaStream_t s) {
  if (p.Cin == 3 && p.in_pitch == 3 && p.size == 3 && p.stride == 1 && p.Cout <= 64 && (p.Cout & 3) == 0 && (p.out_pitch & 3) == 0 &&
      p.Ho <= 65535 && p.batch <= 65535 && !getenv("AY_F32_GENERIC_L0")) {
    f32_first_conv_kernel<<<dim3((p.Wo + 127) / 128, p.Ho, p.batch), 128, 0, s>>>(p);
    AY_CUDA(cudaGetLastError());
    return AY_OK;
  }
  const long long total = (long long)p.batch * p.Ho * p.Wo;
  const long long gx = (total + kBM - 1) / kBM;
  if (gx > 0x7fffffffLL) { set_error("fp32 conv: batch too large"); return AY_EINVAL; }
  dim3 grid((unsigned)gx, (unsigned)((p.Cout + kBN - 1) / kBN));
  f32_conv_kernel<<<grid, 256, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_maxpool(const float* in, int C, int batch, int H, int W, int size, int stride, float* out,
                       cudaStream_t s) {
  if (size != 2 || (stride != 1 && stride != 2) || (stride == 2 && ((H | W) & 1))) {
    set_error("maxpool: only 2x2 stride 1/2 on even sizes is supported (size=%d stride=%d %dx%d)", size, stride, H, W);
    return AY_EINVAL;
  }
  const long long total = (long long)batch * (stride == 2 ? H / 2 : H) * (stride == 2 ? W / 2 : W) * C;
  f32_maxpool_kernel<<<grid_for32(total), 256, 0, s>>>(in, C, batch, H, W, stride, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_upsample2(const float* in, int C, int batch, int H, int W, float* out, cudaStream_t s) {
  f32_upsample2_kernel<<<grid_for32(4LL * batch * H * W * C), 256, 0, s>>>(in, C, batch, H, W, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_add(const float* a, const float* b, long long n, float* out, cudaStream_t s) {
  f32_add_kernel<<<grid_for32(n), 256, 0, s>>>(a, b, n, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_copy_channels(const float* in, int C, long long rows, float* out, int out_pitch, cudaStream_t s) {
  f32_copy_channels_kernel<<<grid_for32(rows * C), 256, 0, s>>>(in, C, rows, out, out_pitch);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_split(const float* in, int in_pitch, int C, long long rows, float* out, cudaStream_t s) {
  if ((C & 3) || (in_pitch & 3)) { set_error("f32 split: channels must be a multiple of 4"); return AY_EINVAL; }
  f32_split_kernel<<<grid_for32(rows * (C / 4)), 256, 0, s>>>(reinterpret_cast<const float4*>(in), in_pitch / 4, C / 4, rows,
                                                              reinterpret_cast<float4*>(out));
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_s2_gather(const float* in, int pitch, int batch, int H, int W, float* planes, long long plane_rows,
                         cudaStream_t s) {
  const long long total = 4LL * batch * (H / 2 + 2) * (W / 2 + 2) * (pitch / 4);
  f32_s2_gather_kernel<<<grid_for32(total), 256, 0, s>>>(reinterpret_cast<const float4*>(in), pitch / 4, batch, H, W,
                                                         reinterpret_cast<float4*>(planes), plane_rows);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_split_s2_gather(const float* in, int in_pitch, int C, int batch, int H, int W, float* planes,
                               long long plane_rows, cudaStream_t s) {
  if ((C & 3) || (in_pitch & 3)) { set_error("f32 split: channels must be a multiple of 4"); return AY_EINVAL; }
  const long long total = 4LL * batch * (H / 2 + 2) * (W / 2 + 2) * (C / 4);
  f32_split_s2_gather_kernel<<<grid_for32(total), 256, 0, s>>>(reinterpret_cast<const float4*>(in), in_pitch / 4, C / 4, batch, H, W,
                                                               reinterpret_cast<float4*>(planes), plane_rows);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

int launch_f32_dump_nchw(const float* in, int C, int batch, int H, int W, float* out, cudaStream_t s) {
  f32_dump_nchw_kernel<<<grid_for32((long long)batch * C * H * W), 256, 0, s>>>(in, C, batch, H, W, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
