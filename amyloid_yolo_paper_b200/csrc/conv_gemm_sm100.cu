// Implicit-GEMM convolution for sm_100a: TMA -> shared memory -> tcgen05.mma -> TMEM
// -> fused epilogue (bias, LeakyReLU, residual add, halo zeroing, 16-bit store, or
// the YOLO decode for head convolutions).
//
// Replaces, for the whole network, the reference's per-block
//   nn.Conv2d + nn.BatchNorm2d + nn.LeakyReLU            (models.py:26-45)
//   shortcut add                                         (models.py:246-248)
//   YOLOLayer inference decode                           (models.py:137-169)
//
// Formulation.  Activations are zero-haloed NHWC rows (ay_common.h).  For output
// row m (a padded pixel) and filter tap t, the input row is m + a_row_off[t]; so
//   out[m, n] = epi( sum_t sum_c A[m + a_row_off[t], c] * W[n, t*Cin + c] )
// which is a GEMM whose A tile for k-block (t, c0) is the 128 x BK box of the
// activation matrix at (row m0 + a_row_off[t], column c0): one 2-D TMA load, the
// hardware zero-filling rows outside the tensor.  Halo rows of the output are
// computed like any other row and overwritten with zeros in the epilogue, which
// keeps the invariant the next layer relies on.
//
// Kernel shape: persistent, one CTA per SM, 192 threads:
//   warp 0   : TMA producer (one lane)           full/empty mbarrier ring, `stages` deep
//   warp 1   : tcgen05.mma issuer (one lane)     UMMA 128 x BN x 16, fp32 accumulate in TMEM
//   warps 2-5: epilogue, one TMEM lane quarter each (tcgen05.ld 32x32b.x32)
// TMEM holds two BN-column accumulators so the epilogue of tile i overlaps the
// main loop of tile i+1.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ay_common.h"

namespace ay {
namespace {

constexpr int kBlockM = 128;
constexpr int kThreads = 192;
constexpr unsigned long long kSpinLimitCycles = 6000000000ull;  // ~3 s: trap instead of hanging the GPU

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned int* err_flag) {
  uint32_t done = 0;
  unsigned long long t0 = 0;
  uint32_t spins = 0;
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
      unsigned long long now = clock64();
      if (t0 == 0) t0 = now;
      if (now - t0 > kSpinLimitCycles) {
        if (err_flag) atomicExch(err_flag, 0xDEAD0001u);
        __trap();
      }
    }
  }
}

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t bar, uint32_t dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
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

// K-major operand tile in shared memory, rows of BK 16-bit elements, hardware
// swizzle equal to the row size (128 B for BK=64, 64 B for BK=32); 8-row groups
// are contiguous (SBO = 8 * row bytes).  sm_100 descriptor version 1.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t row_bytes) {
  const uint64_t layout = (row_bytes == 128) ? 2ull : 4ull;  // SWIZZLE_128B : SWIZZLE_64B
  const uint64_t sbo = (8u * row_bytes) >> 4;
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi, int is_fp16) {
  if (is_fp16) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&b);
}

__device__ __forceinline__ float2 unpack2(uint32_t w, int is_fp16) {
  if (is_fp16) return __half22float2(*reinterpret_cast<__half2*>(&w));
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
}

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ ConvGemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = kBlockM * p.BK * 2;
  const uint32_t b_bytes = p.BN * p.BK * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const int S = p.num_stages;
  const uint32_t bar_base = smem_base + S * stage_bytes;
  // barrier i at bar_base + 8*i : full[S], empty[S], tmem_full[2], tmem_empty[2]; then tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (S + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * S + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * S + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * S + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t tmem_cols = (p.BN <= 16) ? 32u : (p.BN <= 32) ? 64u : (p.BN <= 64) ? 128u : (p.BN <= 128) ? 256u : 512u;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  const int total_tiles = p.num_m_tiles * p.num_n_tiles;
  const int k_iters = p.n_taps * p.kb_per_tap;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m0 = (tile / p.num_n_tiles) * kBlockM;
        const int n0 = (tile % p.num_n_tiles) * p.BN;
        for (int t = 0; t < p.n_taps; ++t) {
          const int arow = m0 + p.a_row_off[t];
          for (int kb = 0; kb < p.kb_per_tap; ++kb) {
            mbar_wait(empty_bar(stage), phase ^ 1u, p.err_flag);
            mbar_expect_tx(full_bar(stage), stage_bytes);
            const uint32_t sa = smem_base + stage * stage_bytes;
            tma_load_2d(&tmA, full_bar(stage), sa, kb * p.BK, arow);
            tma_load_2d(&tmB, full_bar(stage), sa + a_bytes, (t * p.kb_per_tap + kb) * p.BK, n0);
            if (++stage == S) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor: D=f32, A/B = bf16|f16, both K-major, N = BN, M = 128
      const uint32_t fmt = p.is_fp16 ? 0u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(p.BN >> 3) << 17) |
                             ((uint32_t)(kBlockM >> 4) << 24);
      const uint32_t row_bytes = p.BK * 2;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u, p.err_flag);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.BN);
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait(full_bar(stage), phase, p.err_flag);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint64_t adesc = make_smem_desc(sa, row_bytes);
          const uint64_t bdesc = make_smem_desc(sa + a_bytes, row_bytes);
          const int ksteps = p.BK >> 4;
          for (int k = 0; k < ksteps; ++k) {
            // advance 16 elements (32 bytes) along K inside the swizzle row: +2 in 16-byte units
            umma_bf16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                         (uint32_t)((it | k) != 0));
          }
          umma_commit(empty_bar(stage));  // frees the smem slot once these MMAs have read it
          if (++stage == S) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(acc));  // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ---------------- epilogue: warps 2..5, TMEM lane quarter = warp % 4 ----------------
    const int quarter = warp & 3;
    const int row_in_tile = quarter * 32 + lane;
    const uint32_t lane_addr = ((uint32_t)(quarter * 32)) << 16;
    const int hw = p.Hp * p.Wp;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int m0 = (tile / p.num_n_tiles) * kBlockM;
      const int n0 = (tile % p.num_n_tiles) * p.BN;
      const int m = m0 + row_in_tile;
      const bool in_range = m < p.P;
      const int img = m / hw;
      const int pix = m - img * hw;
      const int yy = pix / p.Wp;
      const int xx = pix - yy * p.Wp;
      const bool halo = (yy == 0) | (yy == p.Hp - 1) | (xx == 0) | (xx == p.Wp - 1);

      mbar_wait(tfull_bar(acc), acc_phase, p.err_flag);
      tcgen05_fence_after();
      const uint32_t t_acc = tmem_base + (uint32_t)(acc * p.BN) + lane_addr;

      for (int c0 = 0; c0 < p.BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(t_acc + (uint32_t)c0, v);
        const int nbase = n0 + c0;
        if (!p.head) {
          const float4* bias4 = reinterpret_cast<const float4*>(p.bias + nbase);
          uint32_t packed[16];
          uint4 res[4];
          const bool use_res = (p.residual != nullptr) && in_range && !halo;
          if (use_res) {
            const uint4* rp = reinterpret_cast<const uint4*>(
                static_cast<const uint16_t*>(p.residual) + (size_t)m * p.res_pitch + nbase);
#pragma unroll
            for (int q = 0; q < 4; ++q) res[q] = __ldg(rp + q);
          }
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b = __ldg(bias4 + q);
            float f0 = __uint_as_float(v[4 * q + 0]) + b.x;
            float f1 = __uint_as_float(v[4 * q + 1]) + b.y;
            float f2 = __uint_as_float(v[4 * q + 2]) + b.z;
            float f3 = __uint_as_float(v[4 * q + 3]) + b.w;
            if (p.leaky) {
              f0 = f0 > 0.f ? f0 : 0.1f * f0;
              f1 = f1 > 0.f ? f1 : 0.1f * f1;
              f2 = f2 > 0.f ? f2 : 0.1f * f2;
              f3 = f3 > 0.f ? f3 : 0.1f * f3;
            }
            if (use_res) {
              const uint32_t* rw = reinterpret_cast<const uint32_t*>(&res[q >> 1]);
              const float2 r01 = unpack2(rw[(q & 1) * 2 + 0], p.is_fp16);
              const float2 r23 = unpack2(rw[(q & 1) * 2 + 1], p.is_fp16);
              f0 += r01.x; f1 += r01.y; f2 += r23.x; f3 += r23.y;
            }
            if (halo) { f0 = f1 = f2 = f3 = 0.f; }
            packed[2 * q + 0] = pack2(f0, f1, p.is_fp16);
            packed[2 * q + 1] = pack2(f2, f3, p.is_fp16);
          }
          if (in_range) {
            uint4* op = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out) + (size_t)m * p.out_pitch + nbase);
#pragma unroll
            for (int q = 0; q < 4; ++q)
              op[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
          }
        } else if (in_range && !halo) {
          // fused YOLOLayer decode (models.py:137-169): channel ch = a*(5+C) + f
          const int G = p.Wp - 2;
          const int gx = xx - 1, gy = yy - 1;
          float* obase = p.dec_out + ((size_t)img * p.dec_rows_total + p.dec_row_off) * p.dec_cols;
#pragma unroll 1
          for (int j = 0; j < 32; ++j) {
            const int ch = nbase + j;
            if (ch >= p.cout) break;
            const int a = ch / p.dec_cols;
            const int f = ch - a * p.dec_cols;
            const float t = __uint_as_float(v[j]) + __ldg(p.bias + ch);
            float r;
            if (f == 0) r = __fmul_rn(__fadd_rn(sigmoidf_acc(t), (float)gx), p.stride);
            else if (f == 1) r = __fmul_rn(__fadd_rn(sigmoidf_acc(t), (float)gy), p.stride);
            else if (f == 2) r = __fmul_rn(__fmul_rn(expf(t), p.anchor_w[a]), p.stride);
            else if (f == 3) r = __fmul_rn(__fmul_rn(expf(t), p.anchor_h[a]), p.stride);
            else r = sigmoidf_acc(t);
            obase[((size_t)a * G * G + (size_t)gy * G + gx) * p.dec_cols + f] = r;
          }
        }
      }
      tcgen05_fence_before();
      mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

}  // namespace

int conv_gemm_smem_bytes(int BN, int BK, int stages) {
  return stages * (kBlockM * BK * 2 + BN * BK * 2) + 8 * (2 * stages + 4) + 16 + 1024;
}

int conv_gemm_pick_stages(int BN, int BK) {
  const int stage = kBlockM * BK * 2 + BN * BK * 2;
  int s = (200 * 1024) / stage;
  if (s > 8) s = 8;
  if (s < 2) s = 2;
  return s;
}

int launch_conv_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvGemmParams& p, int num_sms,
                     cudaStream_t stream) {
  if (p.BN % 32 != 0 || p.BN < 32 || p.BN > 256 || (p.BK != 32 && p.BK != 64)) {
    set_error("conv_gemm: unsupported tile BN=%d BK=%d", p.BN, p.BK);
    return AY_EINVAL;
  }
  static int configured_smem = 0;
  const int smem = conv_gemm_smem_bytes(p.BN, p.BK, p.num_stages);
  if (smem > configured_smem) {
    AY_CUDA(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured_smem = 227 * 1024;
  }
  const int total = p.num_m_tiles * p.num_n_tiles;
  const int grid = total < num_sms ? total : num_sms;
  conv_gemm_kernel<<<grid, kThreads, smem, stream>>>(tmA, tmB, p);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
