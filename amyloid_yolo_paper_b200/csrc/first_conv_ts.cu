// First convolution (3 input channels, 3x3, stride 1, pad 1, <= 32 filters) with the A operand built IN TENSOR MEMORY.
//
// Replaces models.py:26-45 for layer 0 together with ToTensor + nearest Resize (utils/transforms.py:89-111), like
// prep_input_kernel + the a_mode-1 variant of conv_gemm_kernel, which stay the DEFAULT: this kernel is opt-in (AY_L0_TS=1,
// measured slower, see the end of this comment) and cross-checked against them in the tests.  Why a second formulation: with both operands in shared memory an MMA instruction costs
// the 128 rows x 32 bytes of its A operand whatever N is (~70 cycles), the staged-pixel layout needs six
// of them per 128 output rows (K = 3 x 32 for 27 real taps) and a staging pass of 16 bytes per pixel in front.  Here
//   * eight builder warps (two per TMEM lane quarter, alternate blocks) gather each output pixel's 27 inputs straight from the uint8 tile (or the caller's floats),
//     pack the im2col row [27 -> 32] into sixteen 32-bit words and write it to TMEM with tcgen05.st: no staging buffer;
//   * the issuer multiplies it from TMEM (tcgen05.mma, A in tensor memory, K = 32: two instructions per 128 rows,
//     8-16 cycles each) with the [32 x 32] weight matrix resident in shared memory;
//   * four epilogue warps add the bias, apply the activation and store 64 bytes per pixel.
// One CTA per SM, persistent over blocks of 128 consecutive rows of the zero-haloed output grid; A and D are four-deep
// rings in TMEM, handed between the roles through mbarriers.
// Measured (B = 256, one B200): 1 434 us against 1 123 us for staging + GEMM.  The tensor pipe and HBM are idle most of the
// time; the bound is instruction issue - about 550 warp instructions per 128 pixels between the builders (index arithmetic,
// predicated loads, 18 shuffles, conversions, packing) and the epilogue, on four schedulers.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ay_common.h"

namespace ay {
namespace {

constexpr int kTsBuildGroups = 2;       // builder warps per TMEM lane quarter (they take alternate blocks)
constexpr int kTsEpiGroups = 2;         // epilogue warps per TMEM lane quarter (alternate blocks as well)
constexpr int kTsThreads = 32 * (1 + 4 * kTsBuildGroups + 4 * kTsEpiGroups);   // warp 0: TMEM owner + MMA issuer, builders, epilogue
constexpr int kTsRing = 4;              // A and D buffers in flight
constexpr int kTsACols = 16;            // 32 bf16 of K = 16 words per row
constexpr int kTsDCols = 32;            // fp32 accumulator columns (N = 32)
constexpr unsigned long long kTsSpinLimit = 6000000000ull;

__device__ __forceinline__ uint32_t ts_smem(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void ts_bar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void ts_bar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ts_bar_wait(uint32_t bar, uint32_t parity, unsigned int* err_flag, uint32_t code) {
  uint32_t done = 0, spins = 0;
  unsigned long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if ((++spins & 0xFFF) == 0) {
      const unsigned long long now = clock64();
      if (t0 == 0) t0 = now;
      if (now - t0 > kTsSpinLimit) {
        if (err_flag) atomicExch(err_flag, 0xDEAD0200u | code);
        __trap();
      }
    }
  }
}
__device__ __forceinline__ bool ts_elect() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void ts_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void ts_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem descriptor]   (M = 128, one CTA)
__device__ __forceinline__ void ts_mma(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void ts_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ts_store16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void ts_load32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void ts_stg256(void* ptr, const uint32_t* v) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"l"(ptr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
template <bool FP16>
__device__ __forceinline__ uint32_t ts_pack(float lo, float hi) {
  if (FP16) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&b);
}

// Inputs as the 16-bit values the staged path feeds the MMA: pixel values 0..255 (exact in bf16 / fp16; ToTensor's 1/255 is
// folded into the weights), floats as 255 * x (rounded when packed).  col[kh*3 + c] = channel c of image pixel
// (y + kh - 1, x), zero outside the image - kept RAW here (byte value or float bits; converted in ts_build) and loaded
// without branches, so that nothing in this function waits for a load: the values are consumed one block later.
__device__ __forceinline__ void ts_load_column(const FirstConvTsParams& p, bool in_batch, int n, int y, int x, uint32_t (&col)[9]) {
  const bool xok = in_batch && x >= 0 && x < p.S;
  if (p.in_kind == AY_IN_U8_HWC) {
    const int sx = xok ? __ldg(p.idx_x + x) * 3 : 0;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int iy = y + kh - 1;
      const bool ok = xok && iy >= 0 && iy < p.S;
      const int sy = ok ? __ldg(p.idx_y + iy) : 0;
      const uint8_t* px = static_cast<const uint8_t*>(p.in) + ((size_t)n * p.src_h + sy) * p.src_w * 3 + sx;
#pragma unroll
      for (int c = 0; c < 3; ++c) col[kh * 3 + c] = ok ? (uint32_t)__ldg(px + c) : 0u;
    }
  } else {
    const float* img = static_cast<const float*>(p.in) + (size_t)n * 3 * p.S * p.S;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int iy = y + kh - 1;
      const bool ok = xok && iy >= 0 && iy < p.S;
#pragma unroll
      for (int c = 0; c < 3; ++c) col[kh * 3 + c] = ok ? __float_as_uint(__ldg(img + ((size_t)c * p.S + iy) * p.S + x)) : 0u;
    }
  }
}

// One warp = 32 consecutive pixels of the padded grid.  Every lane loads the three rows of ITS column only; the columns
// x - 1 and x + 1 come from the neighbouring lanes (the zero pixel that ends every padded row makes this right across row
// ends as well: its own column is all zero), lanes 0 and 31 load their outer neighbour themselves (`ext`).
template <bool FP16>
__device__ __forceinline__ void ts_build(const uint32_t (&own)[9], const uint32_t (&ext)[9], int lane, bool is_u8, uint32_t (&w)[16]) {
  float v[28];
  v[27] = 0.f;
#pragma unroll
  for (int kh = 0; kh < 3; ++kh)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uint32_t raw = own[kh * 3 + c];
      const float mine = is_u8 ? (float)raw : 255.0f * __uint_as_float(raw);
      const uint32_t eraw = ext[kh * 3 + c];
      const float outer = is_u8 ? (float)eraw : 255.0f * __uint_as_float(eraw);
      float left = __shfl_up_sync(0xffffffffu, mine, 1), right = __shfl_down_sync(0xffffffffu, mine, 1);
      if (lane == 0) left = outer;
      if (lane == 31) right = outer;
      v[(kh * 3 + 0) * 3 + c] = left;
      v[(kh * 3 + 1) * 3 + c] = mine;
      v[(kh * 3 + 2) * 3 + c] = right;
    }
#pragma unroll
  for (int j = 0; j < 14; ++j) w[j] = ts_pack<FP16>(v[2 * j], v[2 * j + 1]);     // K element 2j in the low half
  w[14] = 0u;
  w[15] = 0u;
}

template <bool FP16>
__global__ void __launch_bounds__(kTsThreads, 1) first_conv_ts_kernel(const __grid_constant__ FirstConvTsParams p) {
  __shared__ __align__(128) uint16_t s_w[32 * 32];
  __shared__ float s_bias[32];
  __shared__ __align__(8) unsigned long long s_bar[4 * kTsRing];
  __shared__ uint32_t s_tmem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Sp = padded(p.S);
  const long long P = (long long)p.batch * Sp * Sp;
  const int n_blocks = (int)((P + 127) / 128);
  // a CTA owns a CONTIGUOUS range of blocks: the image rows above and below a block's pixels are 3.26 blocks away, so two
  // of the three source rows a builder reads were loaded by this SM a moment ago and hit its L1 (with blocks dealt round
  // robin every source sector crossed L2 -> SM three times and that traffic was the bound)
  const int chunk = (n_blocks + (int)gridDim.x - 1) / (int)gridDim.x;
  const int b_begin = (int)blockIdx.x * chunk, b_end = min(n_blocks, b_begin + chunk);
  const uint32_t bar0 = ts_smem(s_bar);
  auto a_full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto a_empty = [&](int s) { return bar0 + 8u * (uint32_t)(kTsRing + s); };
  auto d_full = [&](int s) { return bar0 + 8u * (uint32_t)(2 * kTsRing + s); };
  auto d_empty = [&](int s) { return bar0 + 8u * (uint32_t)(3 * kTsRing + s); };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTsRing; ++s) {
      ts_bar_init(a_full(s), 4);      // one arrival per builder warp of the block (its four lane quarters)
      ts_bar_init(a_empty(s), 1);     // tcgen05.commit
      ts_bar_init(d_full(s), 1);      // tcgen05.commit
      ts_bar_init(d_empty(s), 4);     // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ts_smem(&s_tmem)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < 32 * 32 / 8; i += kTsThreads)
    reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(p.w_ts) + i);
  if (threadIdx.x < 32) s_bias[threadIdx.x] = __ldg(p.bias + threadIdx.x);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the weights are read by the tensor core
  ts_fence_before();
  __syncthreads();
  ts_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t a_col0 = tmem, d_col0 = tmem + (uint32_t)(kTsRing * kTsACols);

  if (warp == 0) {
    // ---------------- issuer: two K = 16 MMAs per block, A from tensor memory ----------------
    constexpr uint32_t fmt = FP16 ? 0u : 1u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    // B: [32 filters x 32 K] unswizzled K-major core matrices (8 rows x 16 bytes, 128 bytes each): K chunk j of filter
    // group g at (g * 4 + j) * 128 bytes -> LBO (next K chunk) = 128 B, SBO (next 8 filters) = 512 B
    const uint64_t desc = (0ull << 61) | (1ull << 46) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)(128 >> 4) << 16);
    const uint32_t wb = ts_smem(s_w);
    const uint64_t b0 = desc | (uint64_t)((wb >> 4) & 0x3FFFu), b1 = desc | (uint64_t)(((wb + 256u) >> 4) & 0x3FFFu);
    int i = 0;
    for (int b = b_begin; b < b_end; ++b, ++i) {
      const int s = i % kTsRing;
      const uint32_t ph = (uint32_t)(i / kTsRing) & 1u;
      ts_bar_wait(a_full(s), ph, p.err_flag, 1);
      ts_bar_wait(d_empty(s), ph ^ 1u, p.err_flag, 2);
      __syncwarp();
      ts_fence_after();
      if (ts_elect()) {
        const uint32_t a = a_col0 + (uint32_t)(s * kTsACols), d = d_col0 + (uint32_t)(s * kTsDCols);
        ts_mma(d, a, b0, idesc, 0u);
        ts_mma(d, a + 8u, b1, idesc, 1u);
        ts_commit(a_empty(s));
        ts_commit(d_full(s));
      }
      __syncwarp();
    }
  } else if (warp <= 4 * kTsBuildGroups) {
    // ---------------- builders: im2col rows into tensor memory ----------------
    const int group = (warp - 1) >> 2, quarter = warp & 3;
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    const unsigned spsp = (unsigned)(Sp * Sp);
    // loads of one block: this lane's column (3 rows x 3 channels) and, for lanes 0 / 31, the column outside the warp
    auto load_block = [&](int b, uint32_t (&own)[9], uint32_t (&ext)[9]) {
      const unsigned m = (unsigned)b * 128u + (unsigned)(quarter * 32 + lane);
      const unsigned n = m / spsp, pix = m - n * spsp;
      const int py = (int)(pix / (unsigned)Sp), px = (int)(pix - (unsigned)py * (unsigned)Sp);
      const bool in_batch = (long long)m < P;
      ts_load_column(p, in_batch, (int)n, py - kPadLo, px - kPadLo, own);
      // the outer neighbour is the previous / next pixel of the same padded row; past a row end it is a zero pixel
      const int ex = lane == 0 ? px - 1 : px + 1;
      const bool need = (lane == 0 || lane == 31) && ex >= 0 && ex < Sp;
      ts_load_column(p, in_batch && need, (int)n, py - kPadLo, ex - kPadLo, ext);
    };
    uint32_t own[9], ext[9];
    const bool is_u8 = p.in_kind == AY_IN_U8_HWC;
    int i = group;
    int b = b_begin + group;
    if (b < b_end) load_block(b, own, ext);
    for (; b < b_end; b += kTsBuildGroups, i += kTsBuildGroups) {
      const int s = i % kTsRing;
      const uint32_t ph = (uint32_t)(i / kTsRing) & 1u;
      uint32_t cur[9], cex[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) { cur[k] = own[k]; cex[k] = ext[k]; }
      const int nb = b + kTsBuildGroups;
      if (nb < b_end) load_block(nb, own, ext);          // the next block's loads are in flight while this one is packed
      uint32_t w[16];
      ts_build<FP16>(cur, cex, lane, is_u8, w);
      ts_bar_wait(a_empty(s), ph ^ 1u, p.err_flag, 3);
      __syncwarp();
      ts_fence_after();
      ts_store16(a_col0 + (uint32_t)(s * kTsACols) + lane_base, w);
      ts_fence_before();
      __syncwarp();
      if (lane == 0) ts_bar_arrive(a_full(s));
    }
  } else {
    // ---------------- epilogue: bias, activation, 64 bytes per output pixel ----------------
    const int group = (warp - 1 - 4 * kTsBuildGroups) >> 2, quarter = warp & 3;
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    const unsigned spsp = (unsigned)(Sp * Sp);
    const int S2 = p.S / 2;
    const uint32_t Hq2 = (uint32_t)padded(S2), Wq2 = Hq2;
    const uint32_t pitch_b = (uint32_t)p.plane_pitch * 2u;
    int i = group;
    for (int b = b_begin + group; b < b_end; b += kTsEpiGroups, i += kTsEpiGroups) {
      const int s = i % kTsRing;
      const uint32_t ph = (uint32_t)(i / kTsRing) & 1u;
      const unsigned m = (unsigned)b * 128u + (unsigned)(quarter * 32 + lane);
      const unsigned n = m / spsp, pix = m - n * spsp;
      const int py = (int)(pix / (unsigned)Sp), px = (int)(pix - (unsigned)py * (unsigned)Sp);
      const bool live = (long long)m < P && py >= kPadLo && py < Sp - 1 && px >= kPadLo && px < Sp - 1;
      ts_bar_wait(d_full(s), ph, p.err_flag, 4);
      __syncwarp();
      ts_fence_after();
      uint32_t v[32];
      ts_load32(d_col0 + (uint32_t)(s * kTsDCols) + lane_base, v);
      ts_fence_before();
      __syncwarp();
      if (lane == 0) ts_bar_arrive(d_empty(s));
      if (!live) continue;            // halo pixels and rows past the batch keep the zeros the buffers were allocated with
      uint32_t o[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        float y0 = __uint_as_float(v[2 * e]) + s_bias[2 * e], y1 = __uint_as_float(v[2 * e + 1]) + s_bias[2 * e + 1];
        if (p.leaky == 1) { y0 = fmaxf(y0, 0.1f * y0); y1 = fmaxf(y1, 0.1f * y1); }
        else if (p.leaky == 2) { y0 = fmaxf(y0, 0.f); y1 = fmaxf(y1, 0.f); }
        o[e] = ts_pack<FP16>(y0, y1);
      }
      if (p.plane_mode != 2) {
        uint16_t* dst = static_cast<uint16_t*>(p.out_ptr) + (size_t)m * p.out_pitch;
        ts_stg256(dst, &o[0]);
        ts_stg256(dst + 16, &o[8]);
      }
      if (p.plane_mode != 0) {
        // parity planes of the stride-2 consumer (conv_gemm_sm100.cu, row_target): pixel v = unpadded coordinate + 1
        const int vy = py + 1 - kPadLo, vx = px + 1 - kPadLo;
        const uint32_t pos = ((uint32_t)n * Hq2 + (uint32_t)((vy >> 1) + kPadLo)) * Wq2 + (uint32_t)((vx >> 1) + kPadLo);
        char* dst;
        if (p.plane_packed)
          dst = static_cast<char*>(p.planes) + ((unsigned long long)((uint32_t)(vy & 1) * (uint32_t)p.plane_rows + pos) * (2u * pitch_b) +
                                                (uint32_t)(vx & 1) * pitch_b);
        else
          dst = static_cast<char*>(p.planes) + (unsigned long long)((uint32_t)((vy & 1) * 2 + (vx & 1)) * (uint32_t)p.plane_rows + pos) * pitch_b;
        ts_stg256(dst, &o[0]);
        ts_stg256(dst + 32, &o[8]);
      }
    }
  }
  ts_fence_before();
  __syncthreads();
  if (warp == 0) {
    ts_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
  }
}

}  // namespace

int launch_first_conv_ts(const FirstConvTsParams& p, int num_sms, cudaStream_t s) {
  const long long P = (long long)p.batch * padded(p.S) * padded(p.S);
  if (P <= 0 || P > 0x7fffff00LL || (p.S & 1) || p.plane_rows > 0x3fffffffLL) {
    set_error("first_conv_ts: unsupported size (batch %d, S %d)", p.batch, p.S);
    return AY_EINVAL;
  }
  const long long blocks = (P + 127) / 128;
  const int grid = (int)(blocks < num_sms ? blocks : num_sms);
  if (p.is_fp16) first_conv_ts_kernel<true><<<grid, kTsThreads, 0, s>>>(p);
  else first_conv_ts_kernel<false><<<grid, kTsThreads, 0, s>>>(p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
