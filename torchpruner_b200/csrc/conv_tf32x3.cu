// fp32-accurate conv3x3 / linear + per-channel affine (+ReLU) on the 5th-gen tensor cores.
//
// This is the downstream-layer engine of batched Shapley scoring (reference
// shapley_values.py:57 -> attributions.py:85: every marginal evaluation re-runs all layers after
// the evaluation module).  Scores are sums of differences of losses, so TF32/BF16 rounding of the
// operands is not admissible; the layers must be fp32-accurate.  Each fp32 operand
// x is therefore carried as two TF32-exact terms, hi = rn_tf32(x) and lo = rn_tf32(x - hi)
// (x = hi + lo up to 2^-23 relative), and a product is formed as hi*hi + hi*lo + lo*hi on
// tcgen05.mma.kind::tf32 with fp32 accumulation in TMEM ("3xTF32"; the dropped lo*lo term is
// 2^-22 relative).
//
// Formulation: implicit GEMM over NHWC activations.  D[m, n] = sum_k A[m, k] * B[n, k] with
//   m = output pixel (image, y, x), n = output channel, k = (tap r, s; input channel c).
// One CTA computes a 128-pixel x BLOCK_N-channel tile.  A 128-pixel tile is a rectangular box of
// the [N, H, W] pixel grid (bn images x bh rows x all W columns), so for tap (r, s) its A operand
// is ONE 4-D TMA tile load at coordinates (c0, s-1, h0+r-1, n0): out-of-range coordinates are
// zero-filled by the TMA unit, which is exactly the conv's zero padding.  Weights are stored
// [Cout][taps*Cin] (K-major) and arrive by 2-D TMA.  Both land in shared memory in the 128-byte
// swizzled K-major layout the UMMA shared-memory descriptors expect.
//
// Accumulation.  The tensor core adds into its fp32 accumulator with truncation, not
// round-to-nearest: left alone, the error of a K-long dot product grows linearly with the number
// of MMA steps (measured 4e-6 relative at K = 576, 3.5e-5 at K = 4608 — far from fp32).  So the
// TMEM accumulator is only trusted over `promote` k-blocks (64 inputs by default): after that the
// partial sum is read out by the epilogue warps and added, with IEEE rounding, into per-thread
// fp32 registers while the next partial sum accumulates in a second TMEM buffer.
//
// Warp roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM allocator + MMA
// issuer (one lane), warps 2-5 = epilogue (TMEM -> registers -> scale/shift/ReLU -> hi/lo split
// -> global).  smem ring of 3-4 stages with full/empty mbarriers; tcgen05.commit releases a stage
// when the MMAs that read it have finished, and signals the epilogue after the last k-block.
#include "common.cuh"

#include <cuda.h>
#include <cuda_fp16.h>

namespace tp {

constexpr int kBlockM = 128;
constexpr int kBlockK = 32;  // fp32 per k-block: 128 bytes = one swizzle row
constexpr int kConvThreads = 192;

struct ConvGeom {
  int N, H, W, Cin, Cout;
  int taps;      // 9: 3x3 stride 1 pad 1;  1: 1x1 / linear
  int bw, bh, bn;  // pixel box of an M tile, bw*bh*bn == 128, bw == W
  int tiles_h;   // M tiles per image (bn == 1) else 1
  int tiles_w;   // tiles along W (1 unless position-major: bw == bh == 1, bn == 128)
  int num_kb;    // taps * Cin / 32
  int relu;      // activation: 0 none, 1 ReLU, 2 LeakyReLU(slope)
  float slope;
  float out_scale;  // fp16 split outputs hold y * out_scale (a power of two)
  int split;     // 1: y_hi/y_lo get the TF32 split of the result; 0: y_hi gets the raw fp32
  int promote;   // k-blocks accumulated in TMEM before the partial sum is folded into registers
  int pool;      // persistent kernel, position-major fp16 split outputs: MaxPool2d(2) fused into
                 // the epilogue (tiles_h x tiles_w is then the POOLED grid; every tile index
                 // stands for the four output pixels of one window, computed back to back).
                 // 2: a linear layer whose 4*Cout output features are four groups of Cout (the
                 // dense form of a 3x3 convolution on a 2x2 image: one group per output pixel);
                 // the four groups of a tile are computed back to back from the same rows of
                 // the input and B rows sub*Cout.., and their maximum is stored
  long long m_total;
};

template <int BLOCK_N>
struct ConvCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 4;
  static constexpr int kBBytes = BLOCK_N * kBlockK * 4;
  static constexpr int kStageBytes = 2 * kABytes + 2 * kBBytes;
  static constexpr int kStages = (BLOCK_N >= 128) ? 3 : 4;
  static constexpr int kTmemCols = 2 * BLOCK_N < 32 ? 32 : 2 * BLOCK_N;  // two accumulators
  static constexpr int kSmemBytes =
      kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/ + 2 * BLOCK_N * 4;
};

// ---- PTX wrappers ---------------------------------------------------------------------
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, int c0, int c1,
                                            int c2, int c3, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// One lane of a fully active warp (always the same one for the same mask).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
               : "memory");
}
// tcgen05.commit: the mbarrier gets one arrival when every MMA issued so far by this thread has
// completed (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], TF32 inputs, fp32 accumulate.
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Same with fp16 inputs (K = 16 per instruction, twice the TF32 rate).
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
template <bool F16>
__device__ __forceinline__ void umma_x(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                       uint32_t idesc, uint32_t accumulate) {
  if (F16) {
    umma_f16(tmem_d, desc_a, desc_b, idesc, accumulate);
  } else {
    umma_tf32(tmem_d, desc_a, desc_b, idesc, accumulate);
  }
}
// Instruction descriptor: D = F32 (bits 4-5 = 1); A = B = TF32 (2) or F16 (0) in bits 7-9 / 10-12;
// both K-major; N>>3 at bit 17, M>>4 at bit 24 (M = 256 for a CTA pair: 128 rows per CTA).
template <int BLOCK_N, bool F16, bool PAIR = false>
__device__ __forceinline__ constexpr uint32_t make_idesc() {
  return (1u << 4) | ((F16 ? 0u : 2u) << 7) | ((F16 ? 0u : 2u) << 10) |
         ((uint32_t)(BLOCK_N >> 3) << 17) | ((uint32_t)((PAIR ? 2 * kBlockM : kBlockM) >> 4) << 24);
}

// ---- CTA pairs (cta_group::2) -----------------------------------------------------------
// Two CTAs of a cluster, on the two SMs of one TPC, compute one 256-row x BLOCK_N tile: each
// holds its own 128 rows of A and HALF of the B tile in shared memory, the leader (cluster rank
// 0) issues tcgen05.mma.cta_group::2, which reads both CTAs' shared memory and writes both
// CTAs' TMEM.  Per SM and per FLOP a quarter of the operand bytes (half of B) no longer has to
// come in from L2 — the single-CTA kernel was bound by exactly that: 61 B/clk/SM of L2->SM
// traffic at 74 % tensor-pipe activity (profiles/r01_ncu_conv_f16x3_summary.txt).
// All tensor maps of this file describe buffers in the memory of the GPU that runs the kernel.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;"
               ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA tile loads of a CTA pair: data lands in the executing CTA's shared memory, completion bytes
// are counted on the LEADER's mbarrier (`bar_cluster` = its shared::cluster address).
__device__ __forceinline__ void tma_load_4d_pair(void* smem_dst, const CUtensorMap* map, int c0,
                                                 int c1, int c2, int c3, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int c0,
                                                 int c1, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
template <bool F16>
__device__ __forceinline__ void umma_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  if (F16) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// One arrival on the mbarrier at this shared-memory offset in BOTH CTAs of the pair when every
// MMA issued so far has completed.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
               : "memory");
}
// Elements per 128-byte k-block row.
template <bool F16>
struct KBlock {
  static constexpr int kElems = F16 ? 64 : 32;
};

// Shared-memory matrix descriptor, K-major, 128-byte swizzle: rows of 128 bytes, 8-row swizzle
// atoms 1024 bytes apart (SBO), version 1 (Blackwell), layout type 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_sw128(const void* smem_tile) {
  const uint64_t addr = (uint64_t)((smem_u32(smem_tile) & 0x3FFFFu) >> 4);
  return addr | (1ull << 16) /*LBO (ignored for swizzled K-major)*/ |
         ((uint64_t)(1024 >> 4) << 32) /*SBO*/ | (1ull << 46) /*version*/ | (2ull << 61);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
      "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 consecutive fp32 columns of this thread's TMEM lane <- registers (the running maximum of a
// pooling window lives in TMEM between the four tiles that make it up)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
      "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]),
      "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]),
      "f"(v[16]), "f"(v[17]), "f"(v[18]), "f"(v[19]), "f"(v[20]), "f"(v[21]), "f"(v[22]),
      "f"(v[23]), "f"(v[24]), "f"(v[25]), "f"(v[26]), "f"(v[27]), "f"(v[28]), "f"(v[29]),
      "f"(v[30]), "f"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Round-to-nearest TF32 (10 explicit mantissa bits, low 13 bits of the fp32 pattern zero).
// Both terms of the split are rounded to nearest: truncation would bias every product the same
// way and the bias grows like K instead of sqrt(K) over a K-long dot product (measured: 5e-6
// relative with truncation vs < 1e-6 with rounding at K = 576).
__device__ __forceinline__ float tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float tf32_lo(float x, float hi) { return tf32_hi(x - hi); }
// 0: identity, 1: ReLU, 2: LeakyReLU (x > 0 ? x : x * slope, as ATen computes it)
__device__ __forceinline__ float apply_act(float t, int act, float slope) {
  if (act == 1) return nan_relu(t);
  if (act == 2) return t > 0.f ? t : t * slope;
  return t;
}

// ---- the kernel ------------------------------------------------------------------------
template <int BLOCK_N, bool F16>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_tf32x3_kernel(const __grid_constant__ CUtensorMap tm_a_hi,
                   const __grid_constant__ CUtensorMap tm_a_lo,
                   const __grid_constant__ CUtensorMap tm_b_hi,
                   const __grid_constant__ CUtensorMap tm_b_lo, const ConvGeom g,
                   const float* __restrict__ scale, const float* __restrict__ shift,
                   float* __restrict__ y_hi, float* __restrict__ y_lo) {
  using Cfg = ConvCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  // 128-byte swizzle atoms must sit on 1024-byte boundaries of the shared address space.
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* acc_full = empty + Cfg::kStages;   // [2] MMA -> epilogue: partial sum complete
  uint64_t* acc_empty = acc_full + 2;          // [2] epilogue -> MMA: TMEM buffer drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* s_scale = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + 256);
  float* s_shift = s_scale + BLOCK_N;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int mt = blockIdx.x;
  const int nt = blockIdx.y;
  int n0, h0;
  if (g.bn == 1) {
    n0 = mt / g.tiles_h;
    h0 = (mt - n0 * g.tiles_h) * g.bh;
  } else {
    n0 = mt * g.bn;
    h0 = 0;
  }

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tm_a_hi);
    prefetch_tensormap(&tm_a_lo);
    prefetch_tensormap(&tm_b_hi);
    prefetch_tensormap(&tm_b_lo);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 4);  // one arrival per epilogue warp
    }
    fence_mbar_init();
  } else if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
  } else if (warp >= 2) {
    for (int j = threadIdx.x - 64; j < BLOCK_N; j += kConvThreads - 64) {
      const int col = nt * BLOCK_N + j;
      s_scale[j] = col < g.Cout ? scale[col] : 0.f;
      s_shift[j] = col < g.Cout ? shift[col] : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      constexpr int kKE = KBlock<F16>::kElems;
      const int kc = (g.Cin + kKE - 1) / kKE;
      int stage = 0;
      uint32_t phase = 0;
      int tap = 0, cb = 0;
      for (int kb = 0; kb < g.num_kb; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1u);
        mbar_arrive_expect_tx(&full[stage], Cfg::kStageBytes);
        uint8_t* st = smem + stage * Cfg::kStageBytes;
        const int r = (g.taps == 9) ? tap / 3 : 1;
        const int s = (g.taps == 9) ? tap - r * 3 : 1;
        tma_load_4d(st, &tm_a_hi, cb * kKE, s - 1, h0 + r - 1, n0, &full[stage]);
        tma_load_4d(st + Cfg::kABytes, &tm_a_lo, cb * kKE, s - 1, h0 + r - 1, n0, &full[stage]);
        tma_load_2d(st + 2 * Cfg::kABytes, &tm_b_hi, kb * kKE, nt * BLOCK_N, &full[stage]);
        tma_load_2d(st + 2 * Cfg::kABytes + Cfg::kBBytes, &tm_b_lo, kb * kKE, nt * BLOCK_N,
                    &full[stage]);
        if (++cb == kc) {
          cb = 0;
          ++tap;
        }
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer: the warp walks the loop, one elected lane issues (see the persistent
    // kernel: under `if (lane == 0)` the loop was 250 instructions per k-block) =====
    constexpr uint32_t idesc = make_idesc<BLOCK_N, F16>();
    int stage = 0;
    uint32_t phase = 0;
    int chunk = 0, in_chunk = 0;
    for (int kb = 0; kb < g.num_kb; ++kb) {
      const int buf = chunk & 1;
      if (in_chunk == 0) {  // this TMEM buffer must have been drained (2 chunks ago)
        mbar_wait(&acc_empty[buf], (uint32_t)(((chunk >> 1) & 1) ^ 1));
        tc_fence_after();
      }
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      uint8_t* st = smem + stage * Cfg::kStageBytes;
      const uint64_t a_hi = umma_desc_sw128(st);
      const uint64_t a_lo = umma_desc_sw128(st + Cfg::kABytes);
      const uint64_t b_hi = umma_desc_sw128(st + 2 * Cfg::kABytes);
      const uint64_t b_lo = umma_desc_sw128(st + 2 * Cfg::kABytes + Cfg::kBBytes);
      const uint32_t tmem_d = tmem_base + (uint32_t)(buf * BLOCK_N);
      const bool chunk_done = (in_chunk + 1 == g.promote) || (kb == g.num_kb - 1);
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < kBlockK / 8; ++kk) {
          const uint64_t off = (uint64_t)(kk * 32 >> 4);  // 8 tf32 = 32 bytes inside the swizzle row
          // small cross terms first, the dominant hi*hi term last
          umma_x<F16>(tmem_d, a_lo + off, b_hi + off, idesc, (in_chunk | kk) != 0);
          umma_x<F16>(tmem_d, a_hi + off, b_lo + off, idesc, 1u);
          umma_x<F16>(tmem_d, a_hi + off, b_hi + off, idesc, 1u);
        }
        umma_commit(&empty[stage]);  // stage reusable once these MMAs have read it
        if (chunk_done) umma_commit(&acc_full[buf]);  // partial sum of this chunk complete
      }
      __syncwarp();
      if (++stage == Cfg::kStages) {
        stage = 0;
        phase ^= 1u;
      }
      ++in_chunk;
      if (chunk_done) {
        in_chunk = 0;
        ++chunk;
      }
    }
  } else {
    // ===== epilogue: fold partial sums (TMEM -> registers), then affine / ReLU / split =====
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;
    const long long m = (long long)mt * kBlockM + row;
    const bool live = m < g.m_total;
    constexpr int kChunk = BLOCK_N >= 32 ? 32 : 16;
    float acc[BLOCK_N];
#pragma unroll
    for (int j = 0; j < BLOCK_N; ++j) acc[j] = 0.f;
    const int n_chunks = (g.num_kb + g.promote - 1) / g.promote;
    for (int chunk = 0; chunk < n_chunks; ++chunk) {
      const int buf = chunk & 1;
      mbar_wait_unbounded(&acc_full[buf], (uint32_t)((chunk >> 1) & 1));
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BLOCK_N);
#pragma unroll
      for (int c0 = 0; c0 < BLOCK_N; c0 += kChunk) {
        uint32_t v[32];
        if (kChunk == 32) {
          tmem_ld32(taddr + c0, v);
        } else {
          tmem_ld16(taddr + c0, v);
        }
#pragma unroll
        for (int j = 0; j < kChunk; ++j) acc[c0 + j] += __uint_as_float(v[j]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
    float* out_hi = y_hi + m * g.Cout + (long long)nt * BLOCK_N;
    float* out_lo = y_lo ? y_lo + m * g.Cout + (long long)nt * BLOCK_N : nullptr;
    const bool vec_ok = (g.Cout % 4 == 0);
#pragma unroll
    for (int c0 = 0; c0 < BLOCK_N; c0 += kChunk) {
      float y[kChunk];
#pragma unroll
      for (int j = 0; j < kChunk; ++j) {
        float t = fmaf(acc[c0 + j], s_scale[c0 + j], s_shift[c0 + j]);
        t = apply_act(t, g.relu, g.slope);
        y[j] = t;
      }
      if (live) {
        const int col0 = nt * BLOCK_N + c0;
        if (vec_ok && col0 + kChunk <= g.Cout) {
#pragma unroll
          for (int j = 0; j < kChunk; j += 4) {
            if (g.split) {
              const float4 hi = make_float4(tf32_hi(y[j]), tf32_hi(y[j + 1]), tf32_hi(y[j + 2]),
                                            tf32_hi(y[j + 3]));
              const float4 lo = make_float4(tf32_lo(y[j], hi.x), tf32_lo(y[j + 1], hi.y),
                                            tf32_lo(y[j + 2], hi.z), tf32_lo(y[j + 3], hi.w));
              *reinterpret_cast<float4*>(out_hi + c0 + j) = hi;
              *reinterpret_cast<float4*>(out_lo + c0 + j) = lo;
            } else {
              *reinterpret_cast<float4*>(out_hi + c0 + j) =
                  make_float4(y[j], y[j + 1], y[j + 2], y[j + 3]);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < kChunk; ++j) {
            if (col0 + j < g.Cout) {
              if (g.split) {
                const float hi = tf32_hi(y[j]);
                out_hi[c0 + j] = hi;
                out_lo[c0 + j] = tf32_lo(y[j], hi);
              } else {
                out_hi[c0 + j] = y[j];
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}


// =====================================================================================
// Persistent variant (default): one CTA per SM walks output tiles (n-tile fastest, so the CTAs
// that share an activation tile run at the same time and hit it in L2); four TMEM accumulators
// let the MMA issuer run up to three partial sums ahead of the epilogue warps, so the epilogue
// of tile i — scale/shift/ReLU/split, staged through swizzled shared memory and written with
// TMA stores (full 128-byte lines instead of one scattered row per thread) — overlaps the MMAs
// of tile i+1.
// =====================================================================================
template <int BLOCK_N, bool PAIR = false>
struct ConvCfgP {
  static constexpr int kABytes = kBlockM * kBlockK * 4;
  static constexpr int kBRows = PAIR ? BLOCK_N / 2 : BLOCK_N;   // B rows held by one CTA
  static constexpr int kBBytes = kBRows * kBlockK * 4;
  static constexpr int kStageBytes = 2 * kABytes + 2 * kBBytes;
  static constexpr int kStages = (kStageBytes > 48 * 1024) ? 3 : 4;
  static constexpr int kNumAcc = 4;
  static constexpr int kTmemCols = kNumAcc * BLOCK_N;  // 512 / 256
  // Epilogue teams: with 128 output columns, two groups of four warps take 64 columns each.  One
  // team could not keep up: per tile it drains the TMEM partial sums (4 tcgen05.ld per warp and
  // chunk), adds them up, and at the end scales / activates / splits / stages / stores all 128
  // columns, ~90 % of the tile's MMA time — the MMA issuer was seen waiting for a free
  // accumulator 4.7 polls per chunk while it never waited for operands (profiles/
  // r02_ncu_conv_pair_summary.txt), and a stripped kernel issues the same MMAs at 100 % of the
  // pipe (profiles/micro/mma_pace.cu).
  static constexpr int kTeams = BLOCK_N >= 128 ? 2 : 1;
  static constexpr int kTeamCols = BLOCK_N / kTeams;
  static constexpr int kThreads = 64 + 128 * kTeams;
  static constexpr int kStoreBytes = kBlockM * 32 * 4;  // one [128 rows][32 cols] staging tile
  static constexpr int kSmemBytes = kStages * kStageBytes + 2 * kStoreBytes + 1024 /*align*/ +
                                    512 /*barriers*/ + 2 * BLOCK_N * 4;
};

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0,
                                             int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* smem_src, int c0,
                                             int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(map)),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// Tile -> pixel box origin.  Tiles of one image block are consecutive (all positions, then the
// next image block) so the activations they share stay in L2.
__device__ __forceinline__ void tile_origin(const ConvGeom& g, int mt, int& n0, int& h0, int& w0) {
  const int per_block = g.tiles_h * g.tiles_w;
  const int nb = mt / per_block;
  const int rem = mt - nb * per_block;
  const int ht = rem / g.tiles_w;
  n0 = nb * g.bn;
  h0 = ht * g.bh;
  w0 = (rem - ht * g.tiles_w) * g.bw;
}
// A tap contributes to a tile only if its shifted pixel box meets the image: position-major
// tiles (one output pixel x 128 images) skip every tap that falls in the zero padding.
__device__ __forceinline__ bool tap_meets_image(const ConvGeom& g, int h0, int w0, int r, int s) {
  const int hs = h0 + r - 1, ws = w0 + s - 1;
  return hs + g.bh > 0 && hs < g.H && ws + g.bw > 0 && ws < g.W;
}
__device__ __forceinline__ int count_kblocks(const ConvGeom& g, int h0, int w0) {
  if (g.taps != 9) return g.num_kb;
  int n = 0;
#pragma unroll
  for (int t = 0; t < 9; ++t) n += tap_meets_image(g, h0, w0, t / 3, t % 3) ? 1 : 0;
  return n * (g.num_kb / 9);
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// named barriers of the epilogue: 1 + team for the 128 threads of one team, 3 for all teams
__device__ __forceinline__ void epi_barrier(int team) {
  asm volatile("bar.sync %0, 128;" ::"r"(1 + team) : "memory");
}
template <int THREADS>
__device__ __forceinline__ void epi_barrier_all() {
  asm volatile("bar.sync 3, %0;" ::"n"(THREADS) : "memory");
}

// Tile -> pixel box origin of CTA `rank` of a pair: the two CTAs take the SAME position (so they
// walk the same taps) in two adjacent image blocks.  An odd image-block count leaves the last
// pair's second CTA outside the tensor: its loads are zero-filled and its stores clipped.
__device__ __forceinline__ void tile_origin_pair(const ConvGeom& g, int mp, int rank, int& n0,
                                                 int& h0, int& w0) {
  const int per_block = g.tiles_h * g.tiles_w;
  const int nbp = mp / per_block;
  const int rem = mp - nbp * per_block;
  const int ht = rem / g.tiles_w;
  n0 = (2 * nbp + rank) * g.bn;
  h0 = ht * g.bh;
  w0 = (rem - ht * g.tiles_w) * g.bw;
}

template <int BLOCK_N, bool F16, bool PAIR>
__global__ void __launch_bounds__(ConvCfgP<BLOCK_N, PAIR>::kThreads, 1)
conv_tf32x3_persistent(const __grid_constant__ CUtensorMap tm_a_hi,
                       const __grid_constant__ CUtensorMap tm_a_lo,
                       const __grid_constant__ CUtensorMap tm_b_hi,
                       const __grid_constant__ CUtensorMap tm_b_lo,
                       const __grid_constant__ CUtensorMap tm_y_hi,
                       const __grid_constant__ CUtensorMap tm_y_lo, const ConvGeom g,
                       const float* __restrict__ scale, const float* __restrict__ shift,
                       int m_tiles, int n_tiles) {
  // PAIR: m_tiles counts 256-row pair tiles; blockIdx.x >> 1 is the pair, cluster rank the half
  using Cfg = ConvCfgP<BLOCK_N, PAIR>;
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int first_tile = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* store_buf = smem + Cfg::kStages * Cfg::kStageBytes;  // 2 x 16 KB, 1024-aligned
  uint64_t* full = reinterpret_cast<uint64_t*>(store_buf + 2 * Cfg::kStoreBytes);
  uint64_t* empty = full + Cfg::kStages;
  uint64_t* acc_full = empty + Cfg::kStages;
  uint64_t* acc_empty = acc_full + Cfg::kNumAcc;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + Cfg::kNumAcc);
  float* s_scale = reinterpret_cast<float*>(store_buf + 2 * Cfg::kStoreBytes + 512);
  float* s_shift = s_scale + BLOCK_N;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = m_tiles * n_tiles;
  // fused pooling: four output pixels per tile index, and the last accumulator's TMEM columns
  // hold the running maximum instead of a partial sum
  const int subs = g.pool ? 4 : 1;
  const int num_acc = g.pool ? Cfg::kNumAcc - 1 : Cfg::kNumAcc;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tm_a_hi);
    prefetch_tensormap(&tm_a_lo);
    prefetch_tensormap(&tm_b_hi);
    prefetch_tensormap(&tm_b_lo);
    prefetch_tensormap(&tm_y_hi);
    prefetch_tensormap(&tm_y_lo);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full[s], PAIR ? 2 : 1);   // pair: the leader's own arrival + the peer producer's
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < Cfg::kNumAcc; ++b) {
      mbar_init(&acc_full[b], 1);
      // one arrival per epilogue warp (pair: of both CTAs)
      mbar_init(&acc_empty[b], (PAIR ? 8 : 4) * Cfg::kTeams);
    }
    fence_mbar_init();
  } else if (warp == 1) {
    if (PAIR) {
      tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
    } else {
      tmem_alloc(tmem_slot, Cfg::kTmemCols);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();   // the peer's barriers exist before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      constexpr int kKE = KBlock<F16>::kElems;
      const int kc = (g.Cin + kKE - 1) / kKE;
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
        const int mt = tile / n_tiles, nt = tile - mt * n_tiles;
        int n0, hb, wb;
        if (PAIR) {
          tile_origin_pair(g, mt, rank, n0, hb, wb);
        } else {
          tile_origin(g, mt, n0, hb, wb);
        }
        for (int sub = 0; sub < subs; ++sub) {
          const int h0 = g.pool == 1 ? 2 * hb + (sub >> 1) : hb;
          const int w0 = g.pool == 1 ? 2 * wb + (sub & 1) : wb;
          const int bsub = g.pool == 2 ? sub * g.Cout : 0;   // B rows of this output group
          for (int tap = 0; tap < g.taps; ++tap) {
            const int r = (g.taps == 9) ? tap / 3 : 1;
            const int s = (g.taps == 9) ? tap - r * 3 : 1;
            if (g.taps == 9 && !tap_meets_image(g, h0, w0, r, s)) continue;
            for (int cb = 0; cb < kc; ++cb) {
              const int kb = tap * kc + cb;  // k-block of the weight matrix
              mbar_wait(&empty[stage], phase ^ 1u);
              uint8_t* st = smem + stage * Cfg::kStageBytes;
              if (PAIR) {
                // both CTAs' bytes are counted on the leader's barrier
                const uint32_t bar = mapa_shared(smem_u32(&full[stage]), 0);
                if (rank == 0) {
                  mbar_arrive_expect_tx(&full[stage], 2 * Cfg::kStageBytes);
                } else {
                  mbar_arrive_cluster(bar);
                }
                const int brow = bsub + nt * BLOCK_N + rank * Cfg::kBRows;
                tma_load_4d_pair(st, &tm_a_hi, cb * kKE, w0 + s - 1, h0 + r - 1, n0, bar);
                tma_load_4d_pair(st + Cfg::kABytes, &tm_a_lo, cb * kKE, w0 + s - 1, h0 + r - 1, n0,
                                 bar);
                tma_load_2d_pair(st + 2 * Cfg::kABytes, &tm_b_hi, kb * kKE, brow, bar);
                tma_load_2d_pair(st + 2 * Cfg::kABytes + Cfg::kBBytes, &tm_b_lo, kb * kKE, brow,
                                 bar);
              } else {
                mbar_arrive_expect_tx(&full[stage], Cfg::kStageBytes);
                tma_load_4d(st, &tm_a_hi, cb * kKE, w0 + s - 1, h0 + r - 1, n0, &full[stage]);
                tma_load_4d(st + Cfg::kABytes, &tm_a_lo, cb * kKE, w0 + s - 1, h0 + r - 1, n0,
                            &full[stage]);
                tma_load_2d(st + 2 * Cfg::kABytes, &tm_b_hi, kb * kKE, bsub + nt * BLOCK_N,
                            &full[stage]);
                tma_load_2d(st + 2 * Cfg::kABytes + Cfg::kBBytes, &tm_b_lo, kb * kKE,
                            bsub + nt * BLOCK_N, &full[stage]);
              }
              if (++stage == Cfg::kStages) {
                stage = 0;
                phase ^= 1u;
              }
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (the leader CTA of a pair) =====
    // The whole warp walks the loop and ONE elected lane issues.  With the loop under
    // `if (lane == 0)` every descriptor and address is a per-thread value that has to be moved
    // to the uniform registers tcgen05.mma reads, one guarded move at a time: 250 instructions
    // per k-block, longer than the 12 MMAs they issue take on the pipe (768 clocks) — the pipe
    // idled 39 % of the time with operands always ready.  Converged, the same arithmetic stays
    // on the uniform datapath.
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc<BLOCK_N, F16, PAIR>();
      int stage = 0;
      uint32_t phase = 0;
      int buf = 0;
      uint32_t acc_phase = 0;  // flips every time buf wraps
      for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
        int n0, hb, wb;
        if (PAIR) {
          tile_origin_pair(g, tile / n_tiles, 0, n0, hb, wb);
        } else {
          tile_origin(g, tile / n_tiles, n0, hb, wb);
        }
        for (int sub = 0; sub < subs; ++sub) {
          const int h0 = g.pool == 1 ? 2 * hb + (sub >> 1) : hb;
          const int w0 = g.pool == 1 ? 2 * wb + (sub & 1) : wb;
          const int nk = count_kblocks(g, h0, w0);
          int in_chunk = 0;
          for (int kb = 0; kb < nk; ++kb) {
            if (in_chunk == 0) {
              mbar_wait(&acc_empty[buf], acc_phase ^ 1u);
              tc_fence_after();
            }
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            uint8_t* st = smem + stage * Cfg::kStageBytes;
            const uint64_t a_hi = umma_desc_sw128(st);
            const uint64_t a_lo = umma_desc_sw128(st + Cfg::kABytes);
            const uint64_t b_hi = umma_desc_sw128(st + 2 * Cfg::kABytes);
            const uint64_t b_lo = umma_desc_sw128(st + 2 * Cfg::kABytes + Cfg::kBBytes);
            const uint32_t tmem_d = tmem_base + (uint32_t)(buf * BLOCK_N);
            const bool chunk_done = (in_chunk + 1 == g.promote) || (kb == nk - 1);
            if (elect_one()) {
#pragma unroll
              for (int kk = 0; kk < kBlockK / 8; ++kk) {
                const uint64_t off = (uint64_t)(kk * 32 >> 4);
                if (PAIR) {
                  umma_pair<F16>(tmem_d, a_lo + off, b_hi + off, idesc, (in_chunk | kk) != 0);
                  umma_pair<F16>(tmem_d, a_hi + off, b_lo + off, idesc, 1u);
                  umma_pair<F16>(tmem_d, a_hi + off, b_hi + off, idesc, 1u);
                } else {
                  umma_x<F16>(tmem_d, a_lo + off, b_hi + off, idesc, (in_chunk | kk) != 0);
                  umma_x<F16>(tmem_d, a_hi + off, b_lo + off, idesc, 1u);
                  umma_x<F16>(tmem_d, a_hi + off, b_hi + off, idesc, 1u);
                }
              }
              if (PAIR) {
                umma_commit_pair(&empty[stage]);   // frees the stage in both CTAs
                if (chunk_done) umma_commit_pair(&acc_full[buf]);   // wakes both CTAs' epilogues
              } else {
                umma_commit(&empty[stage]);
                if (chunk_done) umma_commit(&acc_full[buf]);
              }
            }
            __syncwarp();
            if (++stage == Cfg::kStages) {
              stage = 0;
              phase ^= 1u;
            }
            ++in_chunk;
            if (chunk_done) {
              in_chunk = 0;
              if (++buf == num_acc) {
                buf = 0;
                acc_phase ^= 1u;
              }
            }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue warps: kTeams groups of four, each owning kTeamCols output columns =====
    constexpr int kTeams = Cfg::kTeams, kCols = Cfg::kTeamCols;
    const int et = threadIdx.x - 64;  // 0 .. 128 * kTeams - 1
    const int team = et >> 7;
    const int cb = team * kCols;      // first column of this team inside the tile
    const int q = warp & 3;           // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;    // row of the tile == TMEM lane
    const bool issuer = ((et & 127) == 0);
    // staging tiles: two teams have one each, a single team double-buffers
    constexpr int kWaitRead = kTeams == 2 ? 0 : 1;
    int buf = 0;
    uint32_t acc_phase = 0;
    uint32_t stores = 0;
    for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
      const int mt = tile / n_tiles, nt = tile - mt * n_tiles;
      int n0, hb, wb;
      if (PAIR) {
        tile_origin_pair(g, mt, rank, n0, hb, wb);
      } else {
        tile_origin(g, mt, n0, hb, wb);
      }
      for (int sub = 0; sub < subs; ++sub) {
        const int h0 = g.pool == 1 ? 2 * hb + (sub >> 1) : hb;
        const int w0 = g.pool == 1 ? 2 * wb + (sub & 1) : wb;
        const int nk = count_kblocks(g, h0, w0);
        const int chunks_this_tile = (nk + g.promote - 1) / g.promote;
        // per-tile scale/shift (everyone finished reading the previous tile's at its last barrier)
        if (sub == 0) {
          epi_barrier_all<128 * kTeams>();
          for (int j = et; j < BLOCK_N; j += 128 * kTeams) {
            const int col = nt * BLOCK_N + j;
            s_scale[j] = col < g.Cout ? scale[col] : 0.f;
            s_shift[j] = col < g.Cout ? shift[col] : 0.f;
          }
        }
        float acc[kCols];
#pragma unroll
        for (int j = 0; j < kCols; ++j) acc[j] = 0.f;
        for (int chunk = 0; chunk < chunks_this_tile; ++chunk) {
          mbar_wait_unbounded(&acc_full[buf], acc_phase);
          tc_fence_after();
          const uint32_t taddr =
              tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BLOCK_N + cb);
#pragma unroll
          for (int c0 = 0; c0 < kCols; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
#pragma unroll
            for (int j = 0; j < 32; ++j) acc[c0 + j] += __uint_as_float(v[j]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (PAIR) {   // the MMA issuer lives in the leader CTA: tell it there
              mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[buf]), 0));
            } else {
              mbar_arrive(&acc_empty[buf]);
            }
          }
          if (++buf == num_acc) {
            buf = 0;
            acc_phase ^= 1u;
          }
        }
        epi_barrier_all<128 * kTeams>();  // scale/shift of this tile visible to all epilogue threads
        if (F16 && g.split) {
          // fp16 split outputs: staging rows are 64 halves (128 bytes), so 64 columns per store
#pragma unroll
          for (int c0 = 0; c0 < kCols; c0 += 64) {
            float y[64];
#pragma unroll
            for (int j = 0; j < 64; ++j) {
              float t = fmaf(acc[c0 + j], s_scale[cb + c0 + j], s_shift[cb + c0 + j]);
              y[j] = apply_act(t, g.relu, g.slope) * g.out_scale;
            }
            if (g.pool) {
              // running maximum of the 2x2 window in the TMEM columns behind the accumulators
              // (this thread's lane, so no other thread ever touches it)
              const uint32_t paddr = tmem_base + ((uint32_t)(q * 32) << 16) +
                                     (uint32_t)(num_acc * BLOCK_N + cb + c0);
              if (sub > 0) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                  uint32_t v[32];
                  tmem_ld32(paddr + 32 * h, v);
#pragma unroll
                  for (int j = 0; j < 32; ++j) y[32 * h + j] = nan_max(y[32 * h + j], __uint_as_float(v[j]));
                }
              }
              if (sub < 3) {
                tmem_st32(paddr, y);
                tmem_st32(paddr + 32, y + 32);
                tmem_wait_st();
                continue;
              }
            }
            for (int pass = 0; pass < 2; ++pass) {
              uint8_t* sbuf = store_buf + (kTeams == 2 ? (uint32_t)team : (stores & 1u)) * Cfg::kStoreBytes;
              if (issuer) bulk_wait_read<kWaitRead>();
              epi_barrier(team);
              uint8_t* rowp = sbuf + row * 128;
#pragma unroll
              for (int c = 0; c < 8; ++c) {
                uint32_t w[4];
#pragma unroll
                for (int h2 = 0; h2 < 4; ++h2) {
                  const float v0 = y[8 * c + 2 * h2], v1 = y[8 * c + 2 * h2 + 1];
                  __half a0 = __float2half_rn(v0), a1 = __float2half_rn(v1);
                  if (pass == 1) {
                    a0 = __float2half_rn(v0 - __half2float(a0));
                    a1 = __float2half_rn(v1 - __half2float(a1));
                  }
                  w[h2] = (uint32_t)__half_as_ushort(a0) | ((uint32_t)__half_as_ushort(a1) << 16);
                }
                *reinterpret_cast<uint4*>(rowp + ((c ^ (row & 7)) << 4)) =
                    make_uint4(w[0], w[1], w[2], w[3]);
              }
              fence_proxy_async();
              epi_barrier(team);
              if (issuer) {
                tma_store_4d(pass == 0 ? &tm_y_hi : &tm_y_lo, sbuf, nt * BLOCK_N + cb + c0,
                             g.pool ? wb : w0, g.pool ? hb : h0, n0);
                bulk_commit();
              }
              ++stores;
            }
          }
          continue;
        }
        const int passes = g.split ? 2 : 1;
#pragma unroll
        for (int c0 = 0; c0 < kCols; c0 += 32) {
          float y[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float t = fmaf(acc[c0 + j], s_scale[cb + c0 + j], s_shift[cb + c0 + j]);
            t = apply_act(t, g.relu, g.slope);
            y[j] = t;
          }
          for (int pass = 0; pass < passes; ++pass) {
            uint8_t* sbuf = store_buf + (kTeams == 2 ? (uint32_t)team : (stores & 1u)) * Cfg::kStoreBytes;
            // the TMA store that used this staging tile before must have read it out
            if (issuer) bulk_wait_read<kWaitRead>();
            epi_barrier(team);
            uint8_t* rowp = sbuf + row * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              float4 o;
              if (g.split) {
                const float h0 = tf32_hi(y[4 * c]), h1 = tf32_hi(y[4 * c + 1]),
                            h2 = tf32_hi(y[4 * c + 2]), h3 = tf32_hi(y[4 * c + 3]);
                if (pass == 0) {
                  o = make_float4(h0, h1, h2, h3);
                } else {
                  o = make_float4(tf32_lo(y[4 * c], h0), tf32_lo(y[4 * c + 1], h1),
                                  tf32_lo(y[4 * c + 2], h2), tf32_lo(y[4 * c + 3], h3));
                }
              } else {
                o = make_float4(y[4 * c], y[4 * c + 1], y[4 * c + 2], y[4 * c + 3]);
              }
              *reinterpret_cast<float4*>(rowp + ((c ^ (row & 7)) << 4)) = o;  // 128B swizzle
            }
            fence_proxy_async();
            epi_barrier(team);
            if (issuer) {
              tma_store_4d(pass == 0 ? &tm_y_hi : &tm_y_lo, sbuf, nt * BLOCK_N + cb + c0, w0, h0, n0);
              bulk_commit();
            }
            ++stores;
          }
        }
      }
    }
    if (issuer) bulk_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) {
    // neither CTA may leave (or free its TMEM) while the other can still read its shared
    // memory through an MMA or signal one of its barriers
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
  } else {
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---- small elementwise companions ------------------------------------------------------
// Split storage: either the TF32 pair in fp32 containers, or the fp16 pair of v * scale
// (scale a power of two chosen per tensor so the values sit mid-range of fp16; an fp16 number
// has the same 11 significant bits as a TF32 one, so hi + lo carries the same 2^-23 precision).
template <bool HALF>
struct SplitStore {
  using T = float;
  __device__ __forceinline__ static void put(T* hi, T* lo, long long i, float v, float) {
    const float h = tf32_hi(v);
    hi[i] = h;
    lo[i] = tf32_lo(v, h);
  }
};
template <>
struct SplitStore<true> {
  using T = __half;
  __device__ __forceinline__ static void put(T* hi, T* lo, long long i, float v, float scale) {
    const float vs = v * scale;
    const __half h = __float2half_rn(vs);
    hi[i] = h;
    lo[i] = __float2half_rn(vs - __half2float(h));
  }
};

template <bool HALF>
__global__ void __launch_bounds__(256)
split_kernel(const float* __restrict__ x, long long n, float scale,
             typename SplitStore<HALF>::T* __restrict__ hi,
             typename SplitStore<HALF>::T* __restrict__ lo) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += gridDim.x * (long long)blockDim.x)
    SplitStore<HALF>::put(hi, lo, i, x[i], scale);
}

// z: NCHW [B, C, H, W].  out: NHWC [P*B, Ho, Wo, Cp] hi/lo of the (optionally 2x2 max-pooled)
// activation with the first (k0+p+1) channels of the permutation zeroed.  Pooling commutes with
// the per-channel mask, so it is applied once, before the P-fold replication.  Cp >= C: channels
// C..Cp-1 are written as zeros (the engine pads channel counts to a multiple of 64).
template <bool HALF>
__global__ void __launch_bounds__(256)
mask_stack_nhwc_kernel(const float* __restrict__ z, const int32_t* __restrict__ inv_perm,
                       long long k0, int P, int B, int C, int H, int W, int pool, float scale, int Cp,
                       typename SplitStore<HALF>::T* __restrict__ out_hi,
                       typename SplitStore<HALF>::T* __restrict__ out_lo) {
  const int Ho = pool ? H / 2 : H, Wo = pool ? W / 2 : W;
  const long long per_copy = (long long)B * Ho * Wo * Cp;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < per_copy;
       e += gridDim.x * (long long)blockDim.x) {
    const int c = (int)(e % Cp);
    long long t = e / Cp;
    const int wo = (int)(t % Wo);
    t /= Wo;
    const int ho = (int)(t % Ho);
    const int b = (int)(t / Ho);
    float v = 0.f;
    long long pos = -1;   // padding channels: "masked" in every copy
    if (c < C) {
      const float* src = z + (((long long)b * C + c) * H) * W;
      if (pool) {
        const float* p0 = src + (2 * ho) * W + 2 * wo;
        v = nan_max(nan_max(p0[0], p0[1]), nan_max(p0[W], p0[W + 1]));
      } else {
        v = src[ho * W + wo];
      }
      pos = (long long)inv_perm[c] - k0;  // masked for every p >= pos
    }
    for (int p = 0; p < P; ++p)
      SplitStore<HALF>::put(out_hi, out_lo, p * per_copy + e, p >= pos ? 0.f : v, scale);
  }
}

// x: NHWC raw fp32 [N, H, W, C] -> 2x2 max pool -> hi/lo [N, H/2, W/2, C].  C % 4 == 0.
template <bool HALF>
__global__ void __launch_bounds__(256)
maxpool2_split_nhwc_kernel(const float* __restrict__ x, long long N, int H, int W, int C,
                           float scale, typename SplitStore<HALF>::T* __restrict__ hi,
                           typename SplitStore<HALF>::T* __restrict__ lo) {
  const int Ho = H / 2, Wo = W / 2, C4 = C / 4;
  const long long total = N * Ho * Wo * C4;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += gridDim.x * (long long)blockDim.x) {
    const int c4 = (int)(e % C4);
    long long t = e / C4;
    const int wo = (int)(t % Wo);
    t /= Wo;
    const int ho = (int)(t % Ho);
    const long long n = t / Ho;
    const long long base = ((n * H + 2 * ho) * W + 2 * wo) * C4 + c4;
    const float4 a = x4[base], b = x4[base + C4], c = x4[base + (long long)W * C4],
                 d = x4[base + (long long)W * C4 + C4];
    const float m0 = nan_max(nan_max(a.x, b.x), nan_max(c.x, d.x));
    const float m1 = nan_max(nan_max(a.y, b.y), nan_max(c.y, d.y));
    const float m2 = nan_max(nan_max(a.z, b.z), nan_max(c.z, d.z));
    const float m3 = nan_max(nan_max(a.w, b.w), nan_max(c.w, d.w));
    SplitStore<HALF>::put(hi, lo, 4 * e, m0, scale);
    SplitStore<HALF>::put(hi, lo, 4 * e + 1, m1, scale);
    SplitStore<HALF>::put(hi, lo, 4 * e + 2, m2, scale);
    SplitStore<HALF>::put(hi, lo, 4 * e + 3, m3, scale);
  }
}

// ---- host side ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static int make_map(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                    const cuuint32_t* box, bool half = false) {
  EncodeTiledFn fn = encode_tiled_fn();
  TP_REQUIRE(fn != nullptr, TP_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
  cuuint64_t strides[4];
  cuuint64_t acc = half ? 2 : 4;
  for (int i = 0; i < rank - 1; ++i) {
    acc *= dims[i];
    strides[i] = acc;
  }
  const cuuint32_t ones[4] = {1, 1, 1, 1};
  const CUresult rc =
      fn(map, half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
         (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TP_REQUIRE(rc == CUDA_SUCCESS, TP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d",
             (int)rc);
  return TP_OK;
}

template <int BLOCK_N, bool F16>
static int launch_conv(const CUtensorMap& a_hi, const CUtensorMap& a_lo, const void* w_hi,
                       const void* w_lo, const ConvGeom& g, const float* scale,
                       const float* shift, float* y_hi, float* y_lo, cudaStream_t st) {
  using Cfg = ConvCfg<BLOCK_N>;
  CUtensorMap b_hi, b_lo;
  const cuuint64_t bdims[2] = {(cuuint64_t)g.taps * g.Cin, (cuuint64_t)g.Cout};
  const cuuint32_t bbox[2] = {(cuuint32_t)KBlock<F16>::kElems, BLOCK_N};
  int rc = make_map(&b_hi, w_hi, 2, bdims, bbox, F16);
  if (rc != TP_OK) return rc;
  rc = make_map(&b_lo, w_lo, 2, bdims, bbox, F16);
  if (rc != TP_OK) return rc;
  auto kern = conv_tf32x3_kernel<BLOCK_N, F16>;
  TP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  const long long m_tiles = (g.m_total + kBlockM - 1) / kBlockM;
  const int n_tiles = (g.Cout + BLOCK_N - 1) / BLOCK_N;
  dim3 grid((unsigned)m_tiles, (unsigned)n_tiles);
  kern<<<grid, kConvThreads, Cfg::kSmemBytes, st>>>(a_hi, a_lo, b_hi, b_lo, g, scale, shift, y_hi,
                                                    y_lo);
  TP_LAUNCH_CHECK("conv_tf32x3_kernel");
  return TP_OK;
}


template <int BLOCK_N, bool F16, bool PAIR>
static int launch_conv_persistent(const CUtensorMap& a_hi, const CUtensorMap& a_lo,
                                  const void* w_hi, const void* w_lo, const ConvGeom& g,
                                  const float* scale, const float* shift, void* y_hi, void* y_lo,
                                  cudaStream_t st) {
  using Cfg = ConvCfgP<BLOCK_N, PAIR>;
  CUtensorMap b_hi, b_lo, o_hi, o_lo;
  const cuuint64_t bdims[2] = {(cuuint64_t)g.taps * g.Cin,
                               (cuuint64_t)g.Cout * (g.pool == 2 ? 4 : 1)};
  const cuuint32_t bbox[2] = {(cuuint32_t)KBlock<F16>::kElems, (cuuint32_t)Cfg::kBRows};
  int rc = make_map(&b_hi, w_hi, 2, bdims, bbox, F16);
  if (rc != TP_OK) return rc;
  rc = make_map(&b_lo, w_lo, 2, bdims, bbox, F16);
  if (rc != TP_OK) return rc;
  // fp16 split outputs are stored 64 halves (128 bytes) per staging row; everything else is fp32
  const bool half_out = F16 && g.split;
  const cuuint64_t ydims[4] = {(cuuint64_t)g.Cout, (cuuint64_t)(g.pool == 1 ? g.W / 2 : g.W),
                               (cuuint64_t)(g.pool == 1 ? g.H / 2 : g.H), (cuuint64_t)g.N};
  const cuuint32_t ybox[4] = {half_out ? 64u : 32u, (cuuint32_t)g.bw, (cuuint32_t)g.bh,
                              (cuuint32_t)g.bn};
  rc = make_map(&o_hi, y_hi, 4, ydims, ybox, half_out);
  if (rc != TP_OK) return rc;
  rc = make_map(&o_lo, y_lo ? y_lo : y_hi, 4, ydims, ybox, half_out);
  if (rc != TP_OK) return rc;
  auto kern = conv_tf32x3_persistent<BLOCK_N, F16, PAIR>;
  TP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  long long img_blocks = (g.N + g.bn - 1) / g.bn;
  if (PAIR) img_blocks = (img_blocks + 1) / 2;   // pairs of image blocks
  const long long m_tiles = img_blocks * g.tiles_h * g.tiles_w;
  const int n_tiles = (g.Cout + BLOCK_N - 1) / BLOCK_N;
  TP_REQUIRE(m_tiles * n_tiles < (1ll << 30), TP_ERR_LIMIT, "tp_conv_tf32x3: too many tiles");
  long long grid = m_tiles * n_tiles * (PAIR ? 2 : 1);
  const long long cap = PAIR ? (sm_count() & ~1) : sm_count();
  if (grid > cap) grid = cap;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(Cfg::kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = PAIR ? 2 : 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TP_CUDA(cudaLaunchKernelEx(&cfg, kern, a_hi, a_lo, b_hi, b_lo, o_hi, o_lo, g, scale, shift,
                             (int)m_tiles, n_tiles));
  return TP_OK;
}

template <bool F16>
static int conv_x3_impl(const void* x_hi, const void* x_lo, int64_t N, int64_t H, int64_t W,
                        int64_t Cin, const void* w_hi, const void* w_lo, int taps, int64_t Cout,
                        const float* scale, const float* shift, int act, float leaky_slope,
                        float out_scale, void* y_hi, void* y_lo, tp_stream_t stream,
                        int pool = 0) {
  constexpr int kKE = KBlock<F16>::kElems;
  TP_REQUIRE(N >= 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0, TP_ERR_INVALID,
             "tp_conv_tf32x3: bad dimension");
  TP_REQUIRE(taps == 1 || taps == 9, TP_ERR_INVALID, "tp_conv_tf32x3: taps must be 1 or 9");
  TP_REQUIRE(act >= 0 && act <= 2, TP_ERR_INVALID, "tp_conv_tf32x3: act must be 0, 1 or 2");
  // A linear layer's last k-block may hang over the end of K: the TMA unit zero-fills both
  // operands there.  A convolution's k index is tap*Cin + c, so its Cin must fill whole blocks.
  TP_REQUIRE(taps == 1 ? (Cin % (F16 ? 8 : 4) == 0) : (Cin % kKE == 0), TP_ERR_LIMIT,
             "tp_conv_%s: Cin must be a multiple of %d (conv) / %d (linear)",
             F16 ? "f16x3" : "tf32x3", kKE, F16 ? 8 : 4);
  TP_REQUIRE(W <= 128 && (128 % W) == 0, TP_ERR_LIMIT,
             "tp_conv_tf32x3: W must divide 128 (got %lld)", (long long)W);
  if (N == 0) return TP_OK;
  TP_REQUIRE(x_hi && x_lo && w_hi && w_lo && scale && shift && y_hi, TP_ERR_INVALID,
             "tp_conv_tf32x3: NULL pointer");
  TP_REQUIRE(N * H * W < (1ll << 31) && N < (1ll << 31), TP_ERR_LIMIT,
             "tp_conv_tf32x3: too many pixels");
  ConvGeom g;
  g.N = (int)N;
  g.H = (int)H;
  g.W = (int)W;
  g.Cin = (int)Cin;
  g.Cout = (int)Cout;
  g.taps = taps;
  g.bw = (int)W;
  const int rows = 128 / (int)W;  // pixel rows per tile
  if (rows <= H) {
    TP_REQUIRE(H % rows == 0, TP_ERR_LIMIT, "tp_conv_tf32x3: H must be a multiple of %d", rows);
    g.bh = rows;
    g.bn = 1;
    g.tiles_h = (int)H / rows;
  } else {
    TP_REQUIRE(rows % H == 0, TP_ERR_LIMIT, "tp_conv_tf32x3: H*W must divide 128");
    g.bh = (int)H;
    g.bn = rows / (int)H;
    g.tiles_h = 1;
  }
  g.tiles_w = 1;
  // Position-major tiles for small images: one output pixel x 128 images per tile, so the taps
  // that fall entirely in the zero padding are skipped (31 % of the MMAs at 4x4, 16 % at 8x8).
  // The persistent kernel writes through TMA stores: row pitch must be a multiple of 16 bytes and
  // the 32-column store box must not straddle the end of a row.
  const bool persistent = (Cout > 16) && (Cout % (F16 ? 64 : 32) == 0) &&
                          tuning().conv_variant != 1 && ((uintptr_t)y_hi % 16 == 0) &&
                          (!y_lo || (uintptr_t)y_lo % 16 == 0);
  TP_REQUIRE(!F16 || persistent || y_lo == nullptr, TP_ERR_LIMIT,
             "tp_conv_f16x3: split fp16 outputs need Cout %% 64 == 0");
  const bool pos_major = persistent && taps == 9 && tuning().conv_pos_major != 1 &&
                         H * W <= tuning().conv_pos_major_max_hw && H * W >= 4 && N >= 128;
  if (pos_major) {
    g.bw = 1;
    g.bh = 1;
    g.bn = 128;
    g.tiles_h = (int)H;
    g.tiles_w = (int)W;
  }
  g.pool = 0;
  if (pool == 2) {
    TP_REQUIRE(F16 && persistent && taps == 1 && H == 1 && W == 1 && y_lo != nullptr, TP_ERR_LIMIT,
               "tp_linear_f16x3_pool4: needs Cout %% 64 == 0, Cout > 16 and split outputs");
    g.pool = 2;
  } else if (pool) {
    TP_REQUIRE(F16 && pos_major && y_lo != nullptr && H % 2 == 0 && W % 2 == 0, TP_ERR_LIMIT,
               "tp_conv_f16x3_pool2: needs position-major tiles (3x3, H*W <= %d, N >= 128), even "
               "H and W and split outputs; ask tp_conv_pool2_fusable first",
               tuning().conv_pos_major_max_hw);
    g.pool = 1;
    g.tiles_h = (int)H / 2;
    g.tiles_w = (int)W / 2;
  }
  g.num_kb = taps * (int)((Cin + kKE - 1) / kKE);
  g.relu = act;
  g.slope = leaky_slope;
  g.out_scale = out_scale;
  g.split = (y_lo != nullptr);
  g.promote = tuning().conv_promote;
  g.m_total = N * H * W;
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  CUtensorMap a_hi, a_lo;
  const cuuint64_t adims[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  const cuuint32_t abox[4] = {(cuuint32_t)kKE, (cuuint32_t)g.bw, (cuuint32_t)g.bh,
                              (cuuint32_t)g.bn};
  int rc = make_map(&a_hi, x_hi, 4, adims, abox, F16);
  if (rc != TP_OK) return rc;
  rc = make_map(&a_lo, x_lo, 4, adims, abox, F16);
  if (rc != TP_OK) return rc;
  float* yf_hi = static_cast<float*>(y_hi);
  float* yf_lo = static_cast<float*>(y_lo);
  if (Cout <= 16)
    return launch_conv<16, F16>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, yf_hi, yf_lo, st);
  if (persistent) {
    // CTA pairs (256-row tiles, half of B per SM) wherever two image blocks exist.  With the MMA
    // issue loop on the uniform datapath the single-CTA kernel is bound by what the M128xN128
    // MMA reads from shared memory (1 090-1 170 TFLOP/s executed); pairs read a quarter less and
    // reach 1 290-1 510 on every 128-wide layer (profiles/r02_conv_pair_ab.txt).  64-wide tiles
    // gain nothing (719 vs 739) and stay single.  conv_pair: 1 = never, 2 = wherever possible.
    const bool pair_wins = Cout > 64;
    const bool pair = tuning().conv_pair != 1 && (N + g.bn - 1) / g.bn >= 2 &&
                      (pair_wins || tuning().conv_pair == 2);
    if (Cout <= 64) {
      if (pair)
        return launch_conv_persistent<64, F16, true>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, y_hi,
                                                     y_lo, st);
      return launch_conv_persistent<64, F16, false>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, y_hi,
                                                    y_lo, st);
    }
    if (pair)
      return launch_conv_persistent<128, F16, true>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, y_hi,
                                                    y_lo, st);
    return launch_conv_persistent<128, F16, false>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, y_hi,
                                                   y_lo, st);
  }
  if (Cout <= 64)
    return launch_conv<64, F16>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, yf_hi, yf_lo, st);
  return launch_conv<128, F16>(a_hi, a_lo, w_hi, w_lo, g, scale, shift, yf_hi, yf_lo, st);
}


static inline unsigned ew_grid(long long items) {
  long long blocks = (items + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace tp

extern "C" {

int tp_conv_tf32x3(const float* x_hi, const float* x_lo, int64_t N, int64_t H, int64_t W,
                   int64_t Cin, const float* w_hi, const float* w_lo, int taps, int64_t Cout,
                   const float* scale, const float* shift, int act, float leaky_slope,
                   float* y_hi, float* y_lo, tp_stream_t stream) {
  return tp::conv_x3_impl<false>(x_hi, x_lo, N, H, W, Cin, w_hi, w_lo, taps, Cout, scale, shift,
                                 act, leaky_slope, 1.0f, y_hi, y_lo, stream);
}

int tp_conv_f16x3(const void* x_hi, const void* x_lo, int64_t N, int64_t H, int64_t W, int64_t Cin,
                  const void* w_hi, const void* w_lo, int taps, int64_t Cout, const float* scale,
                  const float* shift, int act, float leaky_slope, float out_scale, void* y_hi,
                  void* y_lo, tp_stream_t stream) {
  return tp::conv_x3_impl<true>(x_hi, x_lo, N, H, W, Cin, w_hi, w_lo, taps, Cout, scale, shift,
                                act, leaky_slope, out_scale, y_hi, y_lo, stream);
}

int tp_conv_f16x3_pool2(const void* x_hi, const void* x_lo, int64_t N, int64_t H, int64_t W,
                        int64_t Cin, const void* w_hi, const void* w_lo, int64_t Cout,
                        const float* scale, const float* shift, int act, float leaky_slope,
                        float out_scale, void* y_hi, void* y_lo, tp_stream_t stream) {
  return tp::conv_x3_impl<true>(x_hi, x_lo, N, H, W, Cin, w_hi, w_lo, 9, Cout, scale, shift, act,
                                leaky_slope, out_scale, y_hi, y_lo, stream, 1);
}

int tp_linear_f16x3_pool4(const void* x_hi, const void* x_lo, int64_t N, int64_t Cin,
                          const void* w_hi, const void* w_lo, int64_t Cout, const float* scale,
                          const float* shift, int act, float leaky_slope, float out_scale,
                          void* y_hi, void* y_lo, tp_stream_t stream) {
  TP_REQUIRE(tp::tuning().conv_variant != 1, TP_ERR_LIMIT,
             "tp_linear_f16x3_pool4: needs the persistent kernel (conv_variant 0)");
  return tp::conv_x3_impl<true>(x_hi, x_lo, N, 1, 1, Cin, w_hi, w_lo, 1, Cout, scale, shift, act,
                                leaky_slope, out_scale, y_hi, y_lo, stream, 2);
}

int tp_conv_pool2_fusable(int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout) {
  if (N < 128 || H <= 0 || W <= 0 || (H % 2) || (W % 2) || Cin <= 0 || Cin % 64 || Cout <= 16 ||
      Cout % 64 || W > 128 || 128 % W)
    return 0;
  const tp::Tuning& t = tp::tuning();
  return t.conv_variant != 1 && t.conv_pos_major != 1 && H * W <= t.conv_pos_major_max_hw &&
                 H * W >= 4
             ? 1
             : 0;
}

int64_t tp_conv_tf32x3_flops(int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int taps) {
  // 2 * multiply-accumulates the kernel executes per fp32 product term (x3 on the tensor pipe):
  // position-major tiles skip the taps that fall entirely in the zero padding.
  if (N <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cout <= 0) return 0;
  const bool persistent = (Cout > 16) && (Cout % 32 == 0) && tp::tuning().conv_variant != 1;
  const bool pos_major = persistent && taps == 9 && tp::tuning().conv_pos_major != 1 &&
                         H * W <= tp::tuning().conv_pos_major_max_hw && H * W >= 4 && N >= 128;
  if (!pos_major) return 2 * N * H * W * Cout * (int64_t)taps * Cin;
  const int64_t vr = H >= 2 ? 3 * H - 2 : 1, vs = W >= 2 ? 3 * W - 2 : 1;  // valid (tap, pixel) pairs
  return 2 * N * Cout * Cin * vr * vs;
}

int tp_tf32_split(const float* x, int64_t n, float* hi, float* lo, tp_stream_t stream) {
  TP_REQUIRE(n >= 0, TP_ERR_INVALID, "tp_tf32_split: negative size");
  if (n == 0) return TP_OK;
  TP_REQUIRE(x && hi && lo, TP_ERR_INVALID, "tp_tf32_split: NULL pointer");
  tp::split_kernel<false><<<tp::ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, 1.f,
                                                                                       hi, lo);
  TP_LAUNCH_CHECK("split_kernel");
  return TP_OK;
}

int tp_f16_split(const float* x, int64_t n, float scale, void* hi, void* lo, tp_stream_t stream) {
  TP_REQUIRE(n >= 0, TP_ERR_INVALID, "tp_f16_split: negative size");
  if (n == 0) return TP_OK;
  TP_REQUIRE(x && hi && lo, TP_ERR_INVALID, "tp_f16_split: NULL pointer");
  tp::split_kernel<true><<<tp::ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, n, scale, static_cast<__half*>(hi), static_cast<__half*>(lo));
  TP_LAUNCH_CHECK("split_kernel");
  return TP_OK;
}

static int mask_stack_nhwc_impl(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P,
                                int64_t B, int64_t C, int64_t H, int64_t W, int pool, float scale,
                                int64_t C_pad, void* out_hi, void* out_lo, bool half,
                                tp_stream_t stream) {
  TP_REQUIRE(P >= 0 && B >= 0 && C >= 0 && H >= 0 && W >= 0 && k0 >= -1, TP_ERR_INVALID,
             "tp_mask_stack_nhwc: negative dimension");
  TP_REQUIRE(C_pad >= C && C_pad < (1ll << 31), TP_ERR_INVALID,
             "tp_mask_stack_nhwc: C_pad must be at least C");
  TP_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), TP_ERR_INVALID,
             "tp_mask_stack_nhwc: pooling needs even H and W");
  const long long per_copy = B * (pool ? H / 2 : H) * (pool ? W / 2 : W) * C_pad;
  if (per_copy == 0 || P == 0) return TP_OK;
  TP_REQUIRE(z && inv_perm && out_hi && out_lo, TP_ERR_INVALID, "tp_mask_stack_nhwc: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (half) {
    tp::mask_stack_nhwc_kernel<true><<<tp::ew_grid(per_copy), 256, 0, st>>>(
        z, inv_perm, k0, (int)P, (int)B, (int)C, (int)H, (int)W, pool, scale, (int)C_pad,
        static_cast<__half*>(out_hi), static_cast<__half*>(out_lo));
  } else {
    tp::mask_stack_nhwc_kernel<false><<<tp::ew_grid(per_copy), 256, 0, st>>>(
        z, inv_perm, k0, (int)P, (int)B, (int)C, (int)H, (int)W, pool, 1.f, (int)C_pad,
        static_cast<float*>(out_hi), static_cast<float*>(out_lo));
  }
  TP_LAUNCH_CHECK("mask_stack_nhwc_kernel");
  return TP_OK;
}

int tp_mask_stack_nhwc(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P, int64_t B,
                       int64_t C, int64_t H, int64_t W, int pool, int64_t C_pad, float* out_hi,
                       float* out_lo, tp_stream_t stream) {
  return mask_stack_nhwc_impl(z, inv_perm, k0, P, B, C, H, W, pool, 1.f, C_pad, out_hi, out_lo,
                              false, stream);
}

int tp_mask_stack_nhwc_f16(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P,
                           int64_t B, int64_t C, int64_t H, int64_t W, int pool, float scale,
                           int64_t C_pad, void* out_hi, void* out_lo, tp_stream_t stream) {
  return mask_stack_nhwc_impl(z, inv_perm, k0, P, B, C, H, W, pool, scale, C_pad, out_hi, out_lo,
                              true, stream);
}

static int maxpool2_split_impl(const float* x, int64_t N, int64_t H, int64_t W, int64_t C,
                               float scale, void* hi, void* lo, bool half, tp_stream_t stream) {
  TP_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0, TP_ERR_INVALID,
             "tp_maxpool2_split_nhwc: negative dimension");
  TP_REQUIRE(H % 2 == 0 && W % 2 == 0 && C % 4 == 0, TP_ERR_LIMIT,
             "tp_maxpool2_split_nhwc: needs even H, W and C %% 4 == 0");
  const long long total = N * (H / 2) * (W / 2) * (C / 4);
  if (total == 0) return TP_OK;
  TP_REQUIRE(x && hi && lo, TP_ERR_INVALID, "tp_maxpool2_split_nhwc: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (half) {
    tp::maxpool2_split_nhwc_kernel<true><<<tp::ew_grid(total), 256, 0, st>>>(
        x, N, (int)H, (int)W, (int)C, scale, static_cast<__half*>(hi), static_cast<__half*>(lo));
  } else {
    tp::maxpool2_split_nhwc_kernel<false><<<tp::ew_grid(total), 256, 0, st>>>(
        x, N, (int)H, (int)W, (int)C, 1.f, static_cast<float*>(hi), static_cast<float*>(lo));
  }
  TP_LAUNCH_CHECK("maxpool2_split_nhwc_kernel");
  return TP_OK;
}

int tp_maxpool2_split_nhwc(const float* x, int64_t N, int64_t H, int64_t W, int64_t C, float* hi,
                           float* lo, tp_stream_t stream) {
  return maxpool2_split_impl(x, N, H, W, C, 1.f, hi, lo, false, stream);
}

int tp_maxpool2_split_nhwc_f16(const float* x, int64_t N, int64_t H, int64_t W, int64_t C,
                               float scale, void* hi, void* lo, tp_stream_t stream) {
  return maxpool2_split_impl(x, N, H, W, C, scale, hi, lo, true, stream);
}

}  // extern "C"
