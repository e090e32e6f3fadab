// Implicit-GEMM convolution for sm_100a, CTA-pair version:
//   TMA -> shared memory -> tcgen05.mma.cta_group::2 (UMMA 256 x BN x 16) -> TMEM -> fused epilogue.
//
// Replaces, for the whole network, the reference's per-block
//   nn.Conv2d + nn.BatchNorm2d + nn.LeakyReLU            (models.py:26-45)
//   shortcut add                                         (models.py:246-248)
//   YOLOLayer inference decode                           (models.py:137-169)
//
// Formulation (see ay_common.h): activations are zero-haloed NHWC rows, so for output
// row m and filter tap t the input row is m + a_row_off[t] and the convolution is
//   out[m, n] = epi( sum_t sum_c A[m + a_row_off[t], c] * W[n, t*Cin + c] ).
//
// Why a CTA pair: with one CTA per 128 x 256 tile every 64-channel k-block moves 48 KB
// from L2 into shared memory for 512 tensor-pipe cycles (94 B/clk/SM, x148 SMs = twice what
// the L2 fabric delivers) and the MMA reads A and B out of the same shared memory the TMA is
// filling.  Here two CTAs of a cluster own one 256 x BN tile: each loads its own 128 rows of A
// and HALF of the weight tile, and the leader CTA issues cta_group::2 MMAs that read both halves.
//
// Tap reuse: for a 3x3/stride-1 layer the three kw taps of one kh are the same A rows shifted
// by one row.  One TMA box of 130 rows serves all three: the MMA descriptor of tap kw starts kw
// rows (kw*128 B) into the box (the 128-byte swizzle is a function of the absolute shared-memory
// address, so a row-shifted start address reads the right data).  Per (kh, 64-channel block) a CTA
// loads 130 rows of A once and three half weight tiles, instead of three full A and weight tiles.
// K is always processed in 64-channel blocks; a 32-channel input is read through a 64-wide TMA box
// whose upper half is out of bounds and therefore zero-filled by the hardware.
//
// Kernel shape: persistent, one CTA per SM, clusters of 2, 320 threads per CTA:
//   warp 0   : TMA producer (all lanes converged, the copies predicated by elect.sync, in both CTAs; every load
//              signals the LEADER's full barrier)
//   warp 1   : tcgen05.mma issuer (leader CTA only, same converged form); tcgen05.commit multicast to both CTAs
//   warps 2-9: epilogue, two warps per TMEM lane quarter (tcgen05.ld 32x32b.x32).  Tiles of up to four 32-column
//              chunks are taken whole by alternate warps of a quarter (two tiles in flight per quarter), wider
//              tiles are split by alternate chunks.
// TMEM holds 512 / (BN * m_batch) accumulators, so the issuer runs that many tiles ahead of the epilogue.
// The epilogue is thread = output row: residual in and result out are 256-bit ld/st.global straight from/to
// registers (one 32-byte sector per thread and instruction), the math runs on fp32 pairs (FADD2/FMUL2), halo rows
// are never written (the plan zero-fills every buffer once).  An earlier version staged both directions through
// TMA boxes in shared memory; its load round trip, store drain and proxy fence per chunk were the bound of the
// 3x3 layers.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ay_common.h"

namespace ay {
namespace {

constexpr int kRowsPerCta = 128;
constexpr int kEpiParts = kConvEpilogueParts;   // epilogue warps per TMEM lane quarter
constexpr int kEpiWarps = kConvEpilogueWarps;
constexpr int kThreads2 = 32 * (2 + kEpiWarps);   // producer, MMA issuer, 8 epilogue warps
constexpr unsigned long long kSpinLimit = 6000000000ull;  // ~3 s: trap instead of hanging the GPU

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// one lane of a converged warp.  The producer and MMA warps run their loops with all 32 lanes and predicate only the
// TMA / tcgen05 instructions on this: under `if (lane == 0)` every operand counts as divergent and ptxas wraps each
// UTCHMMA / UTMALDG in a VOTEU-ELECT-R2UR.BROADCAST loop (~115 cycles per MMA, measured), which made the issuing
// thread the bound of every layer with N < 256.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_arrive_remote(uint32_t cluster_bar) {
  // default semantics (release at CTA scope): a .release.cluster here compiles to MEMBAR.ALL.GPU + ERRBAR,
  // which cost 16 % of all stall samples; the TMEM hand-over only needs tcgen05.fence::before_thread_sync
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity, unsigned int* err_flag, uint32_t code) {
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
      if (now - t0 > kSpinLimit) {
        if (err_flag) atomicExch(err_flag, 0xDEAD0100u | code);
        __trap();
      }
    }
  }
}
// 2-D tiled TMA load issued by either CTA of the pair; completion bytes go to `cluster_bar`
// (an mbarrier of the leader CTA, shared::cluster address).
__device__ __forceinline__ void tma_load_pair(const CUtensorMap* map, uint32_t cluster_bar, uint32_t dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(cluster_bar), "r"(c0), "r"(c1)
      : "memory");
}
// pixel (py, px) of a padded Hp x Wp grid is one of the zero pixels (kPadLo before, one after each row and column)
__device__ __forceinline__ bool is_halo(int py, int px, int Hp, int Wp) {
  return (py < kPadLo) | (py >= Hp - 1) | (px < kPadLo) | (px >= Wp - 1);
}
__device__ __forceinline__ unsigned long long global_timer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// PDL: block until every grid this launch depends on has completed and its writes are visible (no-op for a launch
// without the programmatic-serialization attribute)
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma2_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
// fp32 accuracy mode: the same instruction on tf32 operands (32-bit elements, the low 13 mantissa bits ignored), K = 8
__device__ __forceinline__ void umma2_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
// arrive on the mbarrier at the same shared-memory offset in both CTAs once all prior MMAs retire
__device__ __forceinline__ void umma2_commit(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tmem_load32(uint32_t taddr, uint32_t (&v)[32]) {
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

// exp and reciprocal as single MUFU instructions (head decode fast path)
__device__ __forceinline__ float fast_exp(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
  return r;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
template <bool FP16>
__device__ __forceinline__ uint32_t pack_pair(float lo, float hi) {
  if (FP16) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&b);
}
template <bool FP16>
__device__ __forceinline__ float2 unpack_pair(uint32_t w) {
  if (FP16) return __half22float2(*reinterpret_cast<__half2*>(&w));
  // bf16 -> fp32 is a 16-bit shift: two ALU ops per pair (the library conversion compiles to PRMT + two IMAD.U32)
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  return *reinterpret_cast<float2*>(&r);
}

constexpr int kMaxChunksPerWarp = 8 / kEpiParts;     // BN <= 256: 8 chunks of 32 columns, two warps per TMEM lane quarter
constexpr int kMaxAcc = 16;             // TMEM accumulators: 512 columns / BN
constexpr int kBiasFloats = 1024;
constexpr int kBK = 64;                 // channels per k-block: one 128-byte swizzle row
constexpr uint32_t kRowBytes = 128;

// 256-bit global accesses (sm_100): a thread moves one whole 32-byte sector per instruction
__device__ __forceinline__ void ldg256(const void* ptr, uint32_t* v) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "l"(ptr));
}
__device__ __forceinline__ void stg256(void* ptr, const uint32_t* v) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"l"(ptr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
// Confidence-threshold compaction fused into the head epilogue (utils/utils.py:248: image_pred[image_pred[:, 4] >=
// conf_thres]): the decoded row index of every passing row is appended to its tile's candidate list.  One atomic per
// warp and anchor when the warp's rows belong to one image (always, except across an image boundary); the order of
// the list is arbitrary - NMS sorts candidates by (score, row) before anything depends on it.  Counts beyond cand_cap
// are still counted, so the consumer sees the overflow and falls back to scanning the dense tensor for that tile.
__device__ __forceinline__ void append_candidates(const ConvGemmParams& p, bool pass, bool one_image, int img,
                                                  int img_first, int row, int lane) {
  if (one_image) {
    const unsigned m = __ballot_sync(0xffffffffu, pass);
    if (m == 0u) return;
    int base = 0;
    const int leader = __ffs(m) - 1;
    if (lane == leader) base = atomicAdd(p.cand_count + img_first, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pass) {
      const int slot = base + __popc(m & ((1u << lane) - 1u));
      if (slot < p.cand_cap) p.cand_idx[(size_t)img_first * p.cand_cap + slot] = row;
    }
  } else if (pass) {
    const int slot = atomicAdd(p.cand_count + img, 1);
    if (slot < p.cand_cap) p.cand_idx[(size_t)img * p.cand_cap + slot] = row;
  }
}

// G = taps per pipeline stage (3: the kw taps of one kh share one 130-row A box; 1: one tap per stage)
// TF32: fp32 accuracy mode (precision = "fp32" on the tensor cores).  Operands are fp32 rows ([hi | lo] activations,
// [W_hi | W_lo] weights), a 128-byte k-block is 32 channels, and a "group" is (pass, tap group) with the three passes
// A_hi W_hi + A_lo W_hi + A_hi W_lo accumulated in the same fp32 TMEM tile: the product carries ~22 mantissa bits.  The
// epilogue is a plain bias + LeakyReLU + fp32 row store (the fp32 network keeps one kernel per cfg block, fp32_path.cu).
template <int G, bool FP16, bool SPLIT, bool TF32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ ConvGemmParams p) {
  // shared memory: [S pipeline stages][epilogue staging][mbarriers][tmem slot][bias]
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_region = (uint32_t)((p.a_rows + 7) & ~7) * kRowBytes;    // multiple of 1024
  const uint32_t b_tile = (uint32_t)(p.BN / 2) * kRowBytes;
  // resident weights: when this CTA's half of the whole weight matrix fits beside the A ring it is loaded once per
  // launch and the ring holds A boxes only (early layers: hundreds of tiles per CTA re-reading the same few KB)
  const uint32_t res_bytes = p.b_resident ? (uint32_t)p.b_res_tiles * b_tile : 0u;
  const uint32_t ring_base = smem_base + res_bytes;
  const int U = p.units_per_stage;                    // (group, k-block) units per pipeline stage
  // row batching (small-N layers): a tile is R consecutive 256-row blocks, each with its own BN accumulator columns,
  // loaded, multiplied and handed over together - one barrier round per R blocks instead of one per block
  const int R = p.m_batch;
  const int tile_rows = 2 * kRowsPerCta * R;
  // tf32 mode: a tile accumulates round-robin into NA accumulators (unit u -> accumulator u % NA) that the epilogue adds
  // with round-to-nearest: the tensor core's truncating accumulation then acts on NA shorter, smaller partial sums
  const int NA = TF32 ? p.tf32_na : 1;
  const int acc_cols = p.BN * R * NA;
  // tiles of <= 4 chunks (32 accumulator columns each): the two epilogue warps of a TMEM lane quarter take alternate tiles
  // (SPLIT: chosen by launch_conv_gemm - a single chunk, or up to kMaxChunksPerWarp chunks of a non-head layer)
  constexpr bool tile_split = SPLIT;
  const uint32_t unit_bytes = a_region + (p.b_resident ? 0u : G * b_tile);
  const uint32_t stage_bytes = (uint32_t)U * unit_bytes;
  const uint32_t tx_bytes = (uint32_t)U * ((uint32_t)p.a_rows * kRowBytes + (p.b_resident ? 0u : G * b_tile));   // one CTA's loads per stage
  const int S = p.num_stages;
  const uint32_t staging_base = ring_base + S * stage_bytes;
  // head convolutions with a single chunk stage their decoded rows here (one buffer per epilogue warp)
  const uint32_t bar_base = staging_base + (uint32_t)kEpiWarps * (uint32_t)p.head_stage_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (S + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * S + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * S + kMaxAcc + a); };
  const uint32_t dummy_bar = bar_base + 8u * (2 * S + 2 * kMaxAcc);
  const uint32_t bres_bar = dummy_bar + 8u;
  const uint32_t tmem_slot = bres_bar + 8u;
  // whole-layer bias staged once: with this much shared memory the L1 is ~1 KB, so a global bias load
  // would be an L2 round trip in the middle of the epilogue's dependency chain
  float* s_bias = reinterpret_cast<float*>(smem_raw + (((tmem_slot + 4u + 15u) & ~15u) - smem_addr(smem_raw)));

  constexpr int kBKe = TF32 ? 32 : kBK;            // channels per 128-byte k-block
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // attribution stamps (profiling runs only): %globaltimer at entry / after the prologue / at exit of every CTA, and
  // when the producer warp's wait for the previous layer returned
  if (p.prof != nullptr && threadIdx.x == 0) p.prof[kProfStampBase + blockIdx.x] = global_timer();
  const uint32_t rank = blockIdx.x & 1u;            // == %cluster_ctarank for __cluster_dims__(2,1,1); warp-uniform
  const bool leader = rank == 0;
  // all 512 TMEM columns (one CTA per SM): 512/BN accumulators, so that the MMA issuer runs that many tiles ahead
  // of the epilogue and the commit -> epilogue -> arrive round trip is off the critical path of small tiles
  const uint32_t tmem_cols = 512u;
  const int num_acc = (p.dbg & 32) ? kEpiParts : p.num_acc;
  const bool has_res = p.has_res && !(p.dbg & 8);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < S; ++s) {
      // leader's copy is the one in use: 1 arrive.expect_tx + bytes of both CTAs (ablation without loads: an
      // arrive of each CTA's producer, so that the pair stays in step)
      bar_init(full_bar(s), (p.dbg & 2) ? 2 : 1);
      bar_init(empty_bar(s), 1);    // tcgen05.commit multicast
    }
    bar_init(dummy_bar, 1);
    bar_init(bres_bar, 1);
    for (int a = 0; a < kMaxAcc; ++a) {
      bar_init(tfull_bar(a), 1);                // tcgen05.commit multicast
      // leader's copy: the epilogue warps of both CTAs that work on a tile (all 8, or 4 when BN = 32: see `split`)
      bar_init(tempty_bar(a), tile_split ? 8 : 2 * kEpiWarps);   // arrivals per accumulator, both CTAs
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.num_n_tiles * p.BN; i += kThreads2) s_bias[i] = __ldg(p.bias + i);
  __syncthreads();
  fence_before();
  cluster_sync_all();
  fence_after();
  // Programmatic dependent launch: everything above touched only constants (tensor maps, bias) and this CTA's own
  // shared / tensor memory, so it ran while the previous layer's last CTAs were still finishing (this CTA became
  // resident as soon as an SM was free).  Let the NEXT layer's CTAs do the same, then wait for the previous layer's
  // results before the first dependent access (A loads, residual loads, every store).
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (p.prof != nullptr && threadIdx.x == 0) p.prof[kProfStampBase + kProfStampCtas + blockIdx.x] = global_timer();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  if (tmem_base != 0u) {       // all 512 columns of an otherwise idle SM: the MMA warp relies on base 0 (uniform operand)
    if (p.err_flag) atomicExch(p.err_flag, 0xDEAD0200u);
    __trap();
  }

  // cycle counters of one representative CTA (tools/layer_times.py --counters): [0] producer waiting for a free
  // stage, [1] producer issuing loads, [2] MMA thread waiting for a free accumulator, [3] waiting for operands,
  // [4] issuing MMAs + commit, [5] MMA thread total, [6] tiles, [7] stages, [8] epilogue warp 0 busy+waiting, [9] its tiles
  const bool prof = p.prof != nullptr && blockIdx.x == 0;
  long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = prof ? clock64() : 0;
  const int total_tiles = p.num_m_tiles * p.num_n_tiles;
  const int n_clusters = gridDim.x >> 1;
  const int first_tile = blockIdx.x >> 1;
  const int stages_per_tile = p.n_groups * p.kb_per_tap / U;

  // shared::cluster address of the LEADER's copy of a barrier: the CTA rank sits in bit 24 of the window address
  constexpr uint32_t kLeaderMask = 0xFEFFFFFFu;
  if (warp == 0) {
    {
      int stage = 0;
      uint32_t phase = 0;
      if (p.b_resident && elect_one()) {
        const uint32_t lead_bres = bres_bar & kLeaderMask;
        if (leader) bar_expect_tx(bres_bar, 2u * res_bytes);
        for (int t = 0; t < p.b_res_tiles; ++t)
          tma_load_pair(&tmB, lead_bres, smem_base + (uint32_t)t * b_tile, t * kBK, (int)rank * (p.BN / 2));
      }
      __syncwarp();
      grid_dependency_wait();                     // (the resident weights above are constants: loaded before the wait)
      if (p.prof != nullptr && lane == 0) p.prof[kProfStampBase + 3 * kProfStampCtas + blockIdx.x] = global_timer();
      for (int tile = first_tile; tile < total_tiles; tile += n_clusters) {
        const int m0 = (tile / p.num_n_tiles) * tile_rows + (int)rank * kRowsPerCta * R;   // this CTA's rows are contiguous
        const int n0 = (tile % p.num_n_tiles) * p.BN + (int)rank * (p.BN / 2);
        // a pipeline stage holds U consecutive (row block, group, k-block) units: one barrier round trip per stage
        int g = 0, kb = 0;
        for (int st = 0; st < stages_per_tile; ++st) {
          const long long c0 = prof ? clock64() : 0;
          bar_wait(empty_bar(stage), phase ^ 1u, p.err_flag, 1);
          __syncwarp();
          const long long c1 = prof ? clock64() : 0;
          const uint32_t lead_full = full_bar(stage) & kLeaderMask;
          const bool el = elect_one();
          const bool load = el && !(p.dbg & 2);
          if (el && (p.dbg & 2)) bar_arrive_remote(lead_full);          // ablation: no operand loads
          if (load && leader) bar_expect_tx(full_bar(stage), 2u * tx_bytes);
          uint32_t sa = ring_base + stage * stage_bytes;
          for (int u = 0; u < U; ++u, sa += unit_bytes) {
            if (load) {
              const int arow = m0 + p.a_row_off[g];
              if (p.a_mode == 1) tma_load_pair(&tmA, lead_full, sa, 0, arow >> 3);   // 8 pixels per 128-byte row
              else tma_load_pair(&tmA, lead_full, sa, kb * kBKe + (TF32 ? p.a_col_off[g] : 0), arow);
              if (!p.b_resident) {
                // weight taps of group g are g*G .. g*G+G-1 (tf32: first tap and pass column from the tables)
                const int bcol0 = TF32 ? p.b_tap0[g] * p.kb_per_tap * kBKe + p.b_col_off[g] : g * G * p.kb_per_tap * kBKe;
#pragma unroll
                for (int j = 0; j < G; ++j)
                  tma_load_pair(&tmB, lead_full, sa + a_region + j * b_tile, bcol0 + (j * p.kb_per_tap + kb) * kBKe, n0);
              }
            }
            if (++kb == p.kb_per_tap) { kb = 0; ++g; }
          }
          __syncwarp();
          if (prof) { pc[0] += c1 - c0; pc[1] += clock64() - c1; }
          if (++stage == S) { stage = 0; phase ^= 1u; }
        }
      }
      if (prof && lane == 0) { atomicAdd(p.prof + 0, (unsigned long long)pc[0]); atomicAdd(p.prof + 1, (unsigned long long)pc[1]); }
    }
  } else if (warp == 1) {
    if (leader) {
      // instruction descriptor: D=f32, A/B = bf16|f16, both K-major, N = BN, M = 256 (pair)
      constexpr uint32_t fmt = TF32 ? 2u : FP16 ? 0u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(p.BN >> 3) << 17) |
                             ((uint32_t)((2 * kRowsPerCta) >> 4) << 24);
      // descriptor constants: SWIZZLE_128B (2 << 61), version 1 (1 << 46), SBO = 1024 B (64 << 32).  Tap j of a group
      // starts j rows (j*128 B) into the A box: the swizzle is a function of the absolute shared-memory address, so
      // the base-offset field stays 0 (measured on B200: setting it to j gives wrong results).
      constexpr uint64_t desc_hi = (2ull << 61) | (1ull << 46) | (64ull << 32);
      // first layer (a_mode 1): the A box is 18 rows of 8 pixels x 8 channels, i.e. 144 consecutive pixels of 16 bytes,
      // read WITHOUT swizzle as overlapping core matrices: 8 consecutive pixels are one 8-row x 16-byte core matrix,
      // the next 8 rows are 128 B further (SBO) and the next 8-element K chunk is the NEXT pixel, 16 B further (LBO).
      // One K=16 MMA therefore contracts taps (kw0,kw1) of a filter row, a second one (kw2, +1 pixel x zero weights).
      constexpr uint64_t desc_a1 = (0ull << 61) | (1ull << 46) | (8ull << 32) | (1ull << 16);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t b_step = p.b_resident ? (uint32_t)p.kb_per_tap * (b_tile >> 4) : (b_tile >> 4);   // tap j+1 of a group
      const int k16_all = p.k16_all > 0 ? p.k16_all : kBK / 16, k16_last = p.k16_last > 0 ? p.k16_last : kBK / 16;
      if (p.b_resident) {
        bar_wait(bres_bar, 0u, p.err_flag, 6);
        __syncwarp();
        fence_after();
      }
      for (int tile = first_tile; tile < total_tiles; tile += n_clusters) {
        const long long c0 = prof ? clock64() : 0;
        bar_wait(tempty_bar(acc), acc_phase ^ 1u, p.err_flag, 2);
        __syncwarp();
        fence_after();
        if (prof) { pc[2] += clock64() - c0; pc[6] += 1; }
        const uint32_t d_tmem = (uint32_t)(acc * acc_cols);      // tmem_base == 0 (checked after the allocation)
        const int m0 = (tile / p.num_n_tiles) * tile_rows;       // (m0 & 7) == 0, also for the peer CTA's first row
        // groups and k-blocks as nested loops with running addresses: the issuing warp has nothing to hide latency
        // with, so every dependent scalar instruction between two barriers (a division, an indexed constant load)
        // shows up in the layer time
        uint32_t first = 0u;                               // 0 for the very first MMA of the tile (overwrite D)
        int g = 0, kb = 0;
        int uix = 0;                                       // tf32: running unit index of the tile (accumulator = uix % NA)
        for (int st = 0; st < stages_per_tile; ++st) {
          const long long c1 = prof ? clock64() : 0;
          bar_wait(full_bar(stage), phase, p.err_flag, 3);
          __syncwarp();
          fence_after();
          const long long c2 = prof ? clock64() : 0;
          const bool el = elect_one();
          const bool issue = el && !(p.dbg & 1);
          uint32_t sa = ring_base + stage * stage_bytes;
          for (int u = 0; u < U; ++u, sa += unit_bytes) {
            if (issue) {
              // weight taps of group g are g*G + j; resident tile index (g*G + j) * kb_per_tap + kb
              const uint32_t sb = p.b_resident ? smem_base + (uint32_t)(g * G * p.kb_per_tap + kb) * b_tile : sa + a_region;
              const uint64_t b0 = desc_hi | (uint64_t)((sb >> 4) & 0x3FFFu);
              if (p.a_mode == 1) {
                // the box holds this CTA's 128*R consecutive rows (+ shift and taps): row block rb starts rb*128 pixels in
                // and accumulates into its own BN columns
                const uint32_t sub16 = ((uint32_t)((m0 + (g - 1) * p.Wp - 1) & 7)) * 16u;
                const uint64_t a0 = desc_a1 | (uint64_t)(((sa + sub16) >> 4) & 0x3FFFu);
                const uint32_t acc_flag = g != 0 ? 1u : 0u;
#pragma unroll 4
                for (int rb = 0; rb < R; ++rb) {
#pragma unroll
                  for (int k = 0; k < 2; ++k)
                    umma2_f16(d_tmem + (uint32_t)(rb * p.BN), a0 + (uint64_t)(rb * kRowsPerCta + 2 * k), b0 + (uint64_t)(2 * k), idesc,
                              acc_flag | (uint32_t)k);
                }
              } else {
                const uint64_t a0 = desc_hi | (uint64_t)((sa >> 4) & 0x3FFFu);
                // a 32-channel input fills only the first two K=16 steps of its 64-channel block (the rest is the TMA's
                // zero fill against zero weights): those MMAs are not issued
#pragma unroll
                for (int j = 0; j < G; ++j) {
                  const int ks = j == G - 1 ? k16_last : k16_all;
#pragma unroll
                  for (int k = 0; k < kBK / 16; ++k)
                    if (k < ks) {
                      if (TF32) umma2_tf32(d_tmem + (uint32_t)((uix % NA) * p.BN), a0 + (uint64_t)(j * (kRowBytes >> 4) + 2 * k),
                                           b0 + (uint64_t)(j * b_step + 2 * k), idesc, (uix >= NA ? 1u : 0u) | (uint32_t)(j | k));
                      else umma2_f16(d_tmem, a0 + (uint64_t)(j * (kRowBytes >> 4) + 2 * k), b0 + (uint64_t)(j * b_step + 2 * k), idesc,
                                     first | (uint32_t)(j | k));
                    }
                }
              }
            }
            first = 1u;
            ++uix;
            if (++kb == p.kb_per_tap) { kb = 0; ++g; }
          }
          if (el) umma2_commit(empty_bar(stage));   // frees this stage in both CTAs
          if (el && (p.dbg & 128)) umma2_commit(dummy_bar);   // ablation: what one more commit costs the issuing thread
          __syncwarp();
          if (prof) { pc[3] += c2 - c1; pc[4] += clock64() - c2; pc[7] += 1; }
          if (++stage == S) { stage = 0; phase ^= 1u; }
        }
        if (elect_one()) umma2_commit(tfull_bar(acc));       // accumulator complete -> both epilogues
        __syncwarp();
        if (++acc == num_acc) { acc = 0; acc_phase ^= 1u; }
      }
      if (prof && lane == 0) {
        pc[5] = clock64() - t_start;
        for (int i = 2; i < 8; ++i) atomicAdd(p.prof + i, (unsigned long long)pc[i]);
      }
    }
  } else {
    // ---------------- epilogue: warps 2..9; TMEM lane quarter = warp % 4, chunk parity = (warp-2)/4 ----------------
    grid_dependency_wait();
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int part = ew >> 2;
    const int row_in_tile = quarter * 32 + lane;
    const uint32_t lane_addr = ((uint32_t)(quarter * 32)) << 16;
    const int hw = p.Hp * p.Wp;
    const int nchunks = (p.BN >> 5) * R;                    // (tf32: per accumulator; the NA accumulators are summed)
    // A tile with a single 32-column chunk would leave half of the epilogue warps idle and make the per-tile
    // latency chain (wait, tcgen05.ld, math, store, arrive) the bound; there the two warps of a quarter take
    // alternate TILES instead of alternate chunks (tile parity = accumulator buffer = `half`).
    // The epilogue of one warp is a long dependent instruction stream (~550 instructions per tile and chunk at IPC ~0.2),
    // so tiles of up to four chunks are not shared between the two warps of a quarter: they take alternate TILES and
    // all chunks of their quarter, which keeps two tiles in flight per quarter and halves the per-tile fixed work.
    constexpr bool split = SPLIT;
    const int c_first = split ? 0 : part;
    constexpr int c_step = split ? 1 : kEpiParts;
    // Tile sequence of this warp and the row it owns in each tile.
    const int t_step = n_clusters * (split ? kEpiParts : 1);
    constexpr int it_step = split ? kEpiParts : 1;
    const int nnt = p.num_n_tiles;
    int tile = first_tile + (split ? part * n_clusters : 0);
    int acc = (split ? part : 0) % num_acc;
    uint32_t acc_phase = 0;
    int ntile = tile % nnt;
    int m = (tile / nnt) * tile_rows + (int)rank * kRowsPerCta * R + row_in_tile;
    // tile += t_step  ==  (m tile, n tile) += (dq, dr) with carry; qrows = dq * tile_rows (host-computed: kept in the
    // constant bank instead of registers that ptxas would spill across the tile loop)
    // row -> (image, y, x) of the padded grid by multiply-shift with host-computed reciprocals (exact for every row of the
    // launch, see magic_div): six instructions per tile and no carried state (the incremental form kept nine more
    // registers live across the tile loop, which ptxas spilled to local memory on this path)
    auto split_row = [&](int row, int& image, int& py, int& px) {
      image = (int)(((unsigned long long)(unsigned)row * p.div_hw_mul) >> p.div_hw_shift);
      const int rem = row - image * hw;
      py = (int)(((unsigned long long)(unsigned)rem * p.div_wp_mul) >> p.div_wp_shift);
      px = rem - py * p.Wp;
    };
    for (; tile < total_tiles; tile += t_step) {
      const int n0 = ntile * p.BN;
      const bool in_range = m < p.P;
      int img, yy, xx;
      split_row(m, img, yy, xx);
      const bool halo = is_halo(yy, xx, p.Hp, p.Wp);
      bool released = false;                                 // accumulator already handed back (row epilogue, last chunk)

      if (p.dbg & 64) {                                          // ablation: no epilogue work at all
        bar_wait(tfull_bar(acc), acc_phase, p.err_flag, 4);
        fence_after();
      } else if (TF32) {
        // fp32 accuracy mode: bias + LeakyReLU in fp32 (one rounding each, as fp32_path.cu), 128 bytes per row and chunk
        const bool store = in_range && !((yy < p.pad_lo) | (yy >= p.Hp - 1) | (xx < p.pad_lo) | (xx >= p.Wp - 1));
        bar_wait(tfull_bar(acc), acc_phase, p.err_flag, 4);
        fence_after();
        const uint32_t t_acc = tmem_base + (uint32_t)(acc * acc_cols) + lane_addr;
        float* orow = static_cast<float*>(p.out_ptr) + (size_t)m * p.out_pitch + n0;
#pragma unroll
        for (int i = 0; i < kMaxChunksPerWarp; ++i) {
          const int c = c_first + c_step * i;
          if (c < nchunks) {
            uint32_t v[32];
            tmem_load32(t_acc + (uint32_t)(c * 32), v);
            for (int sub = 1; sub < NA; ++sub) {
              uint32_t w[32];
              tmem_load32(t_acc + (uint32_t)(sub * p.BN + c * 32), w);
#pragma unroll
              for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__fadd_rn(__uint_as_float(v[e]), __uint_as_float(w[e])));
            }
            const float* bias = s_bias + n0 + c * 32;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              float y = __fadd_rn(__uint_as_float(v[e]), bias[e]);
              if (p.leaky) y = y > 0.f ? y : __fmul_rn(y, act_slope(p.leaky));
              v[e] = __float_as_uint(y);
            }
            if (store && !(p.dbg & 4)) {
              if (has_res) {
                // fused [shortcut] of the fp32 mode (models.py:246-248): one more fp32 rounding, as the stand-alone add
                const float* rrow = static_cast<const float*>(p.res_ptr) + (size_t)m * p.res_pitch + n0 + c * 32;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  uint32_t r8[8];
                  ldg256(rrow + q * 8, r8);
#pragma unroll
                  for (int e = 0; e < 8; ++e)
                    v[q * 8 + e] = __float_as_uint(__fadd_rn(__uint_as_float(v[q * 8 + e]), __uint_as_float(r8[e])));
                }
              }
#pragma unroll
              for (int q = 0; q < 4; ++q) stg256(orow + c * 32 + q * 8, &v[q * 8]);
            }
          }
        }
      } else if (!p.head) {
        // Thread = output row.  All global traffic of the epilogue is 256-bit per thread (one full 32-byte sector, two per
        // 32-channel chunk): the residual is read straight into registers BEFORE waiting for the accumulator, so its
        // latency hides behind the tile's MMAs, and the result is stored straight from registers.  (The earlier TMA
        // staging - residual box in, result box out, one buffer per warp - serialised a load round trip, a store drain
        // and a proxy fence per chunk and made the epilogue, not the tensor pipe, the bound of the 3x3 layers.)
        // Chunk c of a tile is 32 accumulator columns: channels c*32.. of the tile's rows, or - for row-batched tiles
        // (m_batch > 1, BN = 32) - all 32 channels of the rows of sub-block c.
        struct RowTarget { bool store, up; uint16_t* out; uint16_t* plane; };
        // Addresses: 32-bit row indices (every buffer has < 2^31 rows, checked by the plan) and ONE 32x32->64-bit multiply
        // by the row pitch in bytes per target; size_t arithmetic throughout cost ~90 instructions per call, which the
        // row-batched first layer pays per chunk.
        auto at = [](void* base, uint32_t row, uint32_t row_bytes, uint32_t extra_bytes) {
          return reinterpret_cast<uint16_t*>(static_cast<char*>(base) + ((unsigned long long)row * row_bytes + extra_bytes));
        };
        const uint32_t n0b = (uint32_t)n0 * 2u;
        auto row_target = [&](int mr, int image, int py, int px, bool is_halo) {
          RowTarget t;
          // halo rows are never written: every output buffer (rows and planes) is zero-filled when the plan allocates it
          // and each padded pixel owns its row, so the halo stays zero without a select per value and a store per row
          t.store = (mr < p.P) & !is_halo;
          t.plane = nullptr;
          t.up = false;
          if (p.up_mode) {
            // nearest x2 upsample fused (models.py:86-96): pixel (y, x) fills the 2x2 block whose top-left pixel is
            // (2y, 2x) of the consumer's grid, i.e. padded (2py - kPadLo, 2px - kPadLo); halo rows write nothing
            // (their index below may wrap and is never dereferenced)
            const uint32_t Wt = (uint32_t)padded(2 * (p.Wp - kPadLo - 1)), Ht = (uint32_t)padded(2 * (p.Hp - kPadLo - 1));
            t.up = t.store;
            const uint32_t row = ((uint32_t)image * Ht + (uint32_t)(2 * py - kPadLo)) * Wt + (uint32_t)(2 * px - kPadLo);
            t.out = at(p.out_ptr, row, (uint32_t)p.out_pitch * 2u, n0b);
          } else {
            t.out = at(p.out_ptr, (uint32_t)mr, (uint32_t)p.out_pitch * 2u, n0b);
          }
          if (p.plane_mode != 0) {
            // Parity planes of the stride-2 consumer, laid out on ITS padded output grid so that its taps are row shifts:
            // with v = (unpadded coordinate) + 1 (v = 0 is the zero pixel before the row / column), pixel (vy, vx) goes
            // to plane (vy&1, vx&1) at position (vy/2 + kPadLo, vx/2 + kPadLo); the consumer's tap (kh, kw) of output
            // row m is then row m + (kh>>1)*Wq + (kw>>1) of plane (kh&1, kw&1).
            const int vy = py + 1 - kPadLo, vx = px + 1 - kPadLo;
            const uint32_t Hq2 = (uint32_t)padded((p.Hp - kPadLo - 1) / 2), Wq2 = (uint32_t)padded((p.Wp - kPadLo - 1) / 2);
            const uint32_t pos = ((uint32_t)image * Hq2 + (uint32_t)((vy >> 1) + kPadLo)) * Wq2 + (uint32_t)((vx >> 1) + kPadLo);
            const uint32_t prow_n = (uint32_t)p.plane_rows, pitch_b = (uint32_t)p.plane_pitch * 2u;
            if (p.plane_packed) {
              // row-parity plane, 2 x pitch channels per row: [pixel with even vx | pixel with odd vx]
              t.plane = at(p.planes, (uint32_t)(vy & 1) * prow_n + pos, 2u * pitch_b, (uint32_t)(vx & 1) * pitch_b + n0b);
            } else {
              t.plane = at(p.planes, (uint32_t)((vy & 1) * 2 + (vx & 1)) * prow_n + pos, pitch_b, n0b);
            }
          }
          return t;
        };
        const RowTarget t0 = row_target(m, img, yy, xx, halo);
        if (has_res) {
          // pull this thread's residual row of its tile two steps ahead into L2 (single-N-tile layers: the small tiles
          // whose MMAs are too short to hide an HBM round trip)
          const long long mp = (long long)m + 2LL * p.epi_qrows;
          if (nnt == 1 && mp < p.P) {
            const uint16_t* pf = at(const_cast<void*>(p.res_ptr), (uint32_t)mp, (uint32_t)p.res_pitch * 2u, 0u);
            for (int l = split ? 0 : part; l < (p.BN >> 6); l += c_step)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + l * 64) : "memory");
          }
        }
        // The residual of chunk i is loaded two chunks ahead into a ring of two register buffers (the first two before
        // the accumulator wait, so their latency hides behind the tile's MMAs): holding all four chunks' residuals
        // at once (64 registers) made ptxas spill loop state of this hot path to local memory.
        uint32_t res[2][16];
        const uint16_t* res_row = at(const_cast<void*>(p.res_ptr), (uint32_t)m, (uint32_t)p.res_pitch * 2u, n0b);
        const bool load_res = has_res && in_range && !halo;
        auto load_residual = [&](int i) {
          const int c = c_first + c_step * i;
          if (load_res && c < nchunks) {
            ldg256(res_row + c * 32, &res[i & 1][0]);
            ldg256(res_row + c * 32 + 16, &res[i & 1][8]);
          }
        };
        load_residual(0);
        if (kMaxChunksPerWarp > 1) load_residual(1);
        bar_wait(tfull_bar(acc), acc_phase, p.err_flag, 4);
        fence_after();
        const uint32_t t_acc = tmem_base + (uint32_t)(acc * acc_cols) + lane_addr;
#pragma unroll
        for (int i = 0; i < kMaxChunksPerWarp; ++i) {
          const int c = c_first + c_step * i;
          if (c < nchunks) {
            uint32_t v[32];
            tmem_load32(t_acc + (uint32_t)(c * 32), v);
            if (c + c_step >= nchunks) {
              // that was this warp's last read of the accumulator: hand it back to the MMA issuer NOW, before the last
              // chunk's arithmetic and stores (wide tiles have two accumulators only, and the issuer was waiting 16 % of a
              // 52^2 residual layer's time for the epilogue to let go)
              fence_before();
              __syncwarp();
              if (lane == 0) bar_arrive_remote(tempty_bar(acc) & kLeaderMask);
              released = true;
            }
            RowTarget t = t0;
            int ch = c * 32;
            if (R > 1) {                                         // sub-block c of a row-batched tile: 128 rows further per c
              const int mr = m + c * kRowsPerCta;
              int px, py, image;
              split_row(mr, image, py, px);
              t = row_target(mr, image, py, px, is_halo(py, px, p.Hp, p.Wp));
              ch = 0;
            }
            // fp32 pairs (add/mul.f32x2: FADD2 / FMUL2, same IEEE results as the scalar forms, half the instructions)
            const float4* bias4 = reinterpret_cast<const float4*>(s_bias + n0 + ch);
            float2 f[16];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 bb = bias4[q];
              f[2 * q + 0] = add2(make_float2(__uint_as_float(v[4 * q + 0]), __uint_as_float(v[4 * q + 1])), make_float2(bb.x, bb.y));
              f[2 * q + 1] = add2(make_float2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])), make_float2(bb.z, bb.w));
            }
            // (two branches with immediate operands: a slope held in registers across this block made ptxas spill loop
            // state of the epilogue - 28 bytes instead of 4 - and cost layers 0-3 10-20 %)
            if (p.leaky == 1) {
#pragma unroll
              for (int e = 0; e < 16; ++e) {                     // max(x, 0.1x) == (x > 0 ? x : 0.1x) for every x
                const float2 s = mul2(f[e], make_float2(0.1f, 0.1f));
                f[e].x = fmaxf(f[e].x, s.x);
                f[e].y = fmaxf(f[e].y, s.y);
              }
            } else if (p.leaky == 2) {                           // ReLU (the CAA filter network, core.py:165-198)
#pragma unroll
              for (int e = 0; e < 16; ++e) {
                f[e].x = fmaxf(f[e].x, 0.f);
                f[e].y = fmaxf(f[e].y, 0.f);
              }
            }
            uint32_t o[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              float2 y = f[e];
              if (has_res) y = add2(y, unpack_pair<FP16>(res[i & 1][e]));
              o[e] = pack_pair<FP16>(y.x, y.y);
            }
            if (i + 2 < kMaxChunksPerWarp) load_residual(i + 2);
            if (p.up_mode) {
              if (t.up && !(p.dbg & 4)) {
                const size_t down = (size_t)padded(2 * (p.Wp - kPadLo - 1)) * p.out_pitch;      // one row of the upsampled grid
#pragma unroll
                for (int d = 0; d < 4; ++d) {
                  uint16_t* dst = t.out + (d >> 1) * down + (d & 1) * p.out_pitch + ch;
                  stg256(dst, &o[0]);
                  stg256(dst + 16, &o[8]);
                }
              }
            } else if (t.store && !(p.dbg & 4)) {
              if (p.plane_mode != 2) {
                stg256(t.out + ch, &o[0]);
                stg256(t.out + ch + 16, &o[8]);
              }
              if (p.plane_mode != 0) {
                stg256(t.plane + ch, &o[0]);
                stg256(t.plane + ch + 16, &o[8]);
              }
            }
          }
        }
      } else {
        // fused YOLOLayer decode (models.py:137-169): channel ch = a*(5+C) + f
        bar_wait(tfull_bar(acc), acc_phase, p.err_flag, 4);
        fence_after();
        const uint32_t t_acc = tmem_base + (uint32_t)(acc * acc_cols) + lane_addr;
        const int Gd = p.Wp - kPadLo - 1;
        const int gx = xx - kPadLo, gy = yy - kPadLo;
        float* obase = p.dec_out + ((size_t)img * p.dec_rows_total + p.dec_row_off) * p.dec_cols +
                       ((size_t)gy * Gd + gx) * p.dec_cols;
        const size_t a_step = (size_t)Gd * Gd * p.dec_cols;
        // Decoded rows of consecutive pixels are consecutive in memory (dec_cols floats each, halo pixels skipped), so the
        // valid rows of this warp form ONE contiguous span per anchor: staged in shared memory and written with full
        // 128-byte store instructions instead of dec_cols scalar stores per row at a 28-byte stride.
        const bool valid = in_range && !halo;
        const unsigned vmask = __ballot_sync(0xffffffffu, valid);
        const int rank_v = __popc(vmask & ((1u << lane) - 1u));
        const int nvalid = __popc(vmask);
        const int lane0 = vmask ? __ffs(vmask) - 1 : 0;
        const int img_first = __shfl_sync(0xffffffffu, img, lane0);     // (every lane takes part: no short-circuit around it)
        const bool one_image = __all_sync(0xffffffffu, !valid || img == img_first) != 0;
        const bool staged = p.head_stage_bytes > 0 && one_image;
        const bool chunk_staged = staged && nchunks > 1;              // wide heads: 32 x 33 floats per warp, chunk by chunk
        float* hs = reinterpret_cast<float*>(smem_raw + (staging_base - smem_addr(smem_raw)) + (uint32_t)ew * (uint32_t)p.head_stage_bytes);
        if (p.head_fast) {
          // The shipped detector's head: 3 anchors x (5 + 2 classes) = 21 channels in ONE 32-column chunk.  Anchor and
          // field of every channel are compile-time constants here (no select chain per channel), and the two
          // transcendentals are single MUFU instructions: ex2.approx on t*log2(e) and rcp.approx, each within 2 ulp -
          // four orders of magnitude below the 16-bit activations' own error on these outputs (the stand-alone
          // decode_kernel of the test path keeps expf and the IEEE division).
          uint32_t v[32];
          tmem_load32(t_acc, v);
          const float fgx = (float)gx, fgy = (float)gy;
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            float r[7];
#pragma unroll
            for (int f = 0; f < 7; ++f) {
              const float t = __uint_as_float(v[a * 7 + f]) + s_bias[a * 7 + f];
              if (f == 2 || f == 3) {
                r[f] = __fmul_rn(__fmul_rn(fast_exp(t), f == 2 ? p.anchor_w[a] : p.anchor_h[a]), p.stride);
              } else {
                const float sg = fast_rcp(1.0f + fast_exp(-t));
                r[f] = f < 2 ? __fmul_rn(__fadd_rn(sg, f == 0 ? fgx : fgy), p.stride) : sg;
              }
            }
            if (valid && !(p.dbg & 256)) {
              float* dst = staged ? hs + (a * 32 + rank_v) * 7 : obase + (size_t)a * a_step;
#pragma unroll
              for (int f = 0; f < 7; ++f) dst[f] = r[f];
            }
            if (p.cand_count != nullptr)
              append_candidates(p, valid && r[4] >= p.conf_thres, one_image, img, img_first,
                                p.dec_row_off + a * Gd * Gd + gy * Gd + gx, lane);
          }
        } else
        for (int c = c_first; c < nchunks; c += c_step) {
          uint32_t v[32];
          tmem_load32(t_acc + (uint32_t)(c * 32), v);
          const int ch0 = n0 + c * 32;
          // branch-free per channel, so that the 32 independent exp / reciprocal chains overlap (with an if-chain on the
          // field index every channel waited for the previous one's MUFU latencies)
          float r[32];
          int aj = ch0 / p.dec_cols, fj = ch0 - aj * p.dec_cols;  // running (anchor, field) of channel ch0 + j
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (ch0 + j >= p.cout) break;                         // (warp-uniform: channels past cout do not exist)
            const int ch = ch0 + j;
            const bool wh = (fj == 2) | (fj == 3);
            const float t = __uint_as_float(v[j]) + s_bias[ch];
            // single MUFU instructions (ex2.approx, rcp.approx: 2 ulp), as in the fast path above: with expf and the IEEE
            // division the 255-channel heads of yolov3-tiny spent 25 k cycles per chunk here
            const float e = (p.dbg & 512) ? t : fast_exp(wh ? t : -t);      // (ablation 512: no transcendental)
            const float sg = fast_rcp(1.0f + e);                                  // sigmoid(t)
            const float anchor = (fj == 2) ? p.anchor_w[aj] : p.anchor_h[aj];
            const float off = (fj == 0) ? (float)gx : (float)gy;
            const float box = wh ? __fmul_rn(__fmul_rn(e, anchor), p.stride) : __fmul_rn(__fadd_rn(sg, off), p.stride);
            r[j] = (fj < 4) ? box : sg;
            // objectness channel: rows that clear the confidence threshold are appended to the tile's candidate list
            // (utils/utils.py:248 `>=` on the fp32 value that is stored)
            if (p.cand_count != nullptr && fj == 4 && valid && sg >= p.conf_thres) {
              const int slot = atomicAdd(p.cand_count + img, 1);
              if (slot < p.cand_cap) p.cand_idx[(size_t)img * p.cand_cap + slot] = p.dec_row_off + aj * Gd * Gd + gy * Gd + gx;
            }
            if (++fj == p.dec_cols) { fj = 0; aj = min(aj + 1, p.n_anchors - 1); }
          }
          if (chunk_staged) {
            // Wide heads (255 channels: 3 anchors x 85): the chunk's 32 x 32 values are transposed through shared memory
            // (row pitch 33 floats: conflict-free both ways) and leave as one 128-byte run per valid row - lane j carries
            // channel ch0 + j, whose (anchor, field) is fixed for the chunk - instead of 32 scalar stores per lane at a
            // 340-byte stride.
            if (valid) {
#pragma unroll
              for (int j = 0; j < 32; ++j) hs[rank_v * 33 + j] = r[j];
            }
            __syncwarp();
            const int chl = ch0 + lane;
            const int al = chl / p.dec_cols, fl = chl - al * p.dec_cols;
            // (the shuffles are executed by every lane, outside the per-lane condition)
            const unsigned long long first = (unsigned long long)obase;
            const unsigned long long b64 = ((unsigned long long)__shfl_sync(0xffffffffu, (unsigned)(first >> 32), lane0) << 32) |
                                           (unsigned long long)__shfl_sync(0xffffffffu, (unsigned)first, lane0);
            if (chl < p.cout && !(p.dbg & 256)) {
              float* dst = reinterpret_cast<float*>(b64) + (size_t)al * a_step + fl;
              for (int rr = 0; rr < nvalid; ++rr) dst[(size_t)rr * p.dec_cols] = hs[rr * 33 + lane];
            }
            __syncwarp();
            continue;
          }
          int a = ch0 / p.dec_cols, fi = ch0 - a * p.dec_cols;   // running (anchor, field) of channel ch
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (ch0 + j < p.cout && valid && !(p.dbg & 256)) {                 // (ablation 256: no decoded stores)
              if (staged) hs[(a * 32 + rank_v) * p.dec_cols + fi] = r[j];
              else obase[(size_t)a * a_step + fi] = r[j];
            }
            if (++fi == p.dec_cols) { fi = 0; ++a; }
          }
        }
        if (staged && !chunk_staged && nvalid > 0 && !(p.dbg & 256)) {
          __syncwarp();
          const unsigned long long first = (unsigned long long)obase;
          const unsigned long long b64 = ((unsigned long long)__shfl_sync(0xffffffffu, (unsigned)(first >> 32), lane0) << 32) |
                                         (unsigned long long)__shfl_sync(0xffffffffu, (unsigned)first, lane0);
          float* span = reinterpret_cast<float*>(b64);
          const int nfl = nvalid * p.dec_cols;
          for (int a = 0; a < p.n_anchors; ++a)
            for (int i = lane; i < nfl; i += 32) span[(size_t)a * a_step + i] = hs[a * 32 * p.dec_cols + i];
          __syncwarp();
        }
      }
      // this warp is done reading the accumulator: one arrival per warp on the LEADER's barrier
      if (!released) {
        fence_before();
        __syncwarp();
        if (lane == 0) bar_arrive_remote(tempty_bar(acc) & kLeaderMask);
      }
      // next tile of this warp
      acc += it_step;
      if (acc >= num_acc) { acc -= num_acc; acc_phase ^= 1u; }
      ntile += p.epi_dr;
      m += p.epi_qrows;
      if (ntile >= nnt) {
        ntile -= nnt;
        m += tile_rows;
      }
    }
  }

  // both CTAs must stay resident until the pair's last MMA, commit and remote arrive have landed
  fence_before();
  cluster_sync_all();
  if (warp == 2) {
    fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
  if (p.prof != nullptr && threadIdx.x == 0) p.prof[kProfStampBase + 2 * kProfStampCtas + blockIdx.x] = global_timer();
}

}  // namespace

// bytes of one (group, k-block) unit: an A box and, unless the weights are resident, its G weight tiles
int conv_gemm_stage_bytes(int BN, int G, int a_rows, int resident) {
  return ((a_rows + 7) & ~7) * (int)kRowBytes + (resident ? 0 : G * (BN / 2) * (int)kRowBytes);
}

static int conv_gemm_fixed_smem(int num_stages, int nbuf, int extra_bytes = 0) {
  (void)nbuf;
  return extra_bytes + 8 * (2 * num_stages + 2 * kMaxAcc + 2) + 32 +
         4 * kBiasFloats + 1024 /*alignment*/;
}

// -> stages (2..8) and staging buffers per epilogue warp (2 when they fit beside >= 3 stages, else 1)
void conv_gemm_pick_smem(int BN, int G, int a_rows, int resident_bytes, int units_per_stage, int extra_bytes, int* stages, int* nbuf) {
  const int sb = units_per_stage * conv_gemm_stage_bytes(BN, G, a_rows, resident_bytes > 0);
  const int limit = 227 * 1024 - resident_bytes;
  int s = (limit - conv_gemm_fixed_smem(8, 2, extra_bytes)) / sb;
  if (s > 8) s = 8;
  *stages = s;
  *nbuf = 2;      // kept for the plan's bookkeeping: the epilogue no longer stages through shared memory
}

// floor(n / d) == (n * mul) >> shift for every n < 2^nbits:  shift = nbits + ceil(log2 d), mul = ceil(2^shift / d)
// (mul * d - 2^shift < d <= 2^ceil(log2 d), so the error term n * (mul*d - 2^shift) / (d * 2^shift) stays below 1/d)
static void magic_div(unsigned d, int nbits, unsigned* mul, int* shift) {
  int lg = 0;
  while ((1ull << lg) < d) ++lg;
  *shift = nbits + lg;
  *mul = (unsigned)(((1ull << *shift) + d - 1) / d);
}

int launch_conv_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvGemmParams& p_in, int G, int is_fp16,
                     int num_sms, cudaStream_t stream) {
  ConvGemmParams p = p_in;
  {
    // rows reach P rounded up to whole tiles (plus the row batch of the last one)
    const unsigned long long max_row = (unsigned long long)p.num_m_tiles * 256ull * (unsigned)p.m_batch + 1024ull;
    int nbits = 1;
    while ((1ull << nbits) <= max_row) ++nbits;
    const unsigned hw = (unsigned)p.Hp * (unsigned)p.Wp;
    int lg_hw = 0;
    while ((1ull << lg_hw) < hw) ++lg_hw;
    if (nbits > 30) { set_error("conv_gemm: %llu rows exceed the row-index range", max_row); return AY_EINVAL; }
    magic_div(hw, nbits, &p.div_hw_mul, &p.div_hw_shift);
    magic_div((unsigned)p.Wp, lg_hw, &p.div_wp_mul, &p.div_wp_shift);     // remainder < Hp*Wp <= 2^lg_hw
  }
  if (p.BN % 32 != 0 || p.BN < 32 || p.BN > 256 || p.num_stages < 2 || p.num_stages > 8 || (G != 1 && G != 2 && G != 3) ||
      (p.nbuf != 1 && p.nbuf != 2) || p.num_n_tiles * p.BN > kBiasFloats || p.units_per_stage < 1 ||
      p.m_batch < 1 || (p.m_batch > 1 && (p.BN != 32 || p.head || p.has_res || p.m_batch > 8)) ||
      (p.m_batch > 1 && p.a_mode != 1) || (p.n_groups * p.kb_per_tap) % p.units_per_stage != 0) {
    set_error("conv_gemm: unsupported tile BN=%d G=%d stages=%d nbuf=%d n_tiles=%d", p.BN, G, p.num_stages, p.nbuf,
              p.num_n_tiles);
    return AY_EINVAL;
  }
  const int smem = (p.b_resident ? p.b_res_tiles * (p.BN / 2) * (int)kRowBytes : 0) +
                   p.num_stages * p.units_per_stage * conv_gemm_stage_bytes(p.BN, G, p.a_rows, p.b_resident) +
                   conv_gemm_fixed_smem(p.num_stages, p.nbuf, kEpiWarps * p.head_stage_bytes);
  if (smem > 227 * 1024 || (p.b_resident && p.num_n_tiles != 1)) {
    set_error("conv_gemm: shared memory plan does not fit (%d bytes, resident=%d)", smem, p.b_resident);
    return AY_EINVAL;
  }
  const int total = p.num_m_tiles * p.num_n_tiles;
  const int pairs = num_sms / 2;
  const int grid = 2 * (total < pairs ? total : pairs);
  // epilogue schedule: tiles of one chunk, or of up to kMaxChunksPerWarp chunks in a non-head layer, are taken whole
  // by alternate warps of a TMEM lane quarter (SPLIT); wider tiles are shared chunk by chunk
  const int chunks = (p.BN * p.m_batch) >> 5;
  const bool split = chunks == 1 || (chunks <= kMaxChunksPerWarp && !p.head);
  if (split && (p.num_acc % kEpiParts) != 0) { set_error("conv_gemm: %d accumulators cannot be shared by tile", p.num_acc); return AY_EINVAL; }
  {
    const int t_step = (grid / 2) * (split ? kEpiParts : 1);
    const int dq = t_step / p.num_n_tiles;
    p.epi_dr = t_step - dq * p.num_n_tiles;
    p.epi_qrows = dq * 256 * p.m_batch;
  }
  // Programmatic dependent launch (chained layers): this grid's CTAs may become resident, and run their prologue,
  // while the previous kernel in the stream drains; the kernel itself waits (griddepcontrol.wait) before touching
  // anything the previous kernel wrote.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kThreads2);
  cfg.dynamicSmemBytes = (size_t)smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.pdl ? 1 : 0;
#define AY_LAUNCH4(GG, FF, SS, TT)                                                                                            \
  do {                                                                                                                        \
    AY_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<GG, FF, SS, TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)); \
    AY_CUDA(cudaLaunchKernelEx(&cfg, conv_gemm_kernel<GG, FF, SS, TT>, tmA, tmB, p));                                         \
  } while (0)
#define AY_LAUNCH3(GG, FF, SS) AY_LAUNCH4(GG, FF, SS, false)
  if (p.tf32) {
    if (p.head || p.b_resident || p.a_mode != 0 || p.m_batch != 1 || (G != 1 && G != 3) || p.tf32_na < 1 ||
        p.BN * p.tf32_na * p.num_acc > 512) {
      set_error("conv_gemm: tf32 mode supports plain 1x1 / 3x3 tiles only");
      return AY_EINVAL;
    }
    if (G == 3) { if (split) AY_LAUNCH4(3, false, true, true); else AY_LAUNCH4(3, false, false, true); }
    else        { if (split) AY_LAUNCH4(1, false, true, true); else AY_LAUNCH4(1, false, false, true); }
    AY_CUDA(cudaGetLastError());
    return AY_OK;
  }
#define AY_LAUNCH(GG, FF) do { if (split) AY_LAUNCH3(GG, FF, true); else AY_LAUNCH3(GG, FF, false); } while (0)
  if (G == 3) { if (is_fp16) AY_LAUNCH(3, true); else AY_LAUNCH(3, false); }
  else if (G == 2) { if (is_fp16) AY_LAUNCH(2, true); else AY_LAUNCH(2, false); }
  else        { if (is_fp16) AY_LAUNCH(1, true); else AY_LAUNCH(1, false); }
#undef AY_LAUNCH
#undef AY_LAUNCH3
#undef AY_LAUNCH4
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
