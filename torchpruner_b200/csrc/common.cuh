// Shared helpers for the torchpruner_b200 CUDA library (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/torchpruner_b200.h"

namespace tp {

// ---- error plumbing -----------------------------------------------------------------
void set_error(const char* fmt, ...);

#define TP_REQUIRE(cond, code, ...)       \
  do {                                    \
    if (!(cond)) {                        \
      ::tp::set_error(__VA_ARGS__);       \
      return (code);                      \
    }                                     \
  } while (0)

#define TP_CUDA(expr)                                                               \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      ::tp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),       \
                      __FILE__, __LINE__);                                          \
      return TP_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

#define TP_LAUNCH_CHECK(name)                                                       \
  do {                                                                              \
    cudaError_t _e = cudaGetLastError();                                            \
    if (_e != cudaSuccess) {                                                        \
      ::tp::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));     \
      return TP_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

// Cached per-process device facts (SM count of the current device).
int sm_count();

// ---- tuning knobs (tp_set_tuning) ---------------------------------------------------
struct Tuning {
  int reduce_variant = 0;        // 0 auto, 1 direct loads, 2 TMA bulk tiles
  int reduce_stages = 4;         // smem ring depth of the TMA variant
  int reduce_tile_floats = 4096; // floats per stream per stage
  int reduce_ctas_per_sm = 2;    // persistent CTAs per SM (TMA variant)
  int reduce_tma_min_elems = 1 << 20;
  int reduce_ldg_waves = 4;      // grid cap of the direct-load kernel, in waves of 8 CTAs/SM
  int conv_promote = 2;          // k-blocks (of 32 inputs) per TMEM partial sum
  int conv_variant = 0;          // 0 auto (persistent + TMA stores), 1 one tile per CTA
  int conv_pos_major = 0;        // 0 auto (position-major tiles for small images), 1 never
  int conv_pos_major_max_hw = 256;  // largest H*W that gets position-major tiles
  int conv_pair = 0;             // CTA pairs (cta_group::2): 0 where measured faster, 1 never, 2 always
  int prefix_variant = 0;        // stored-form prefix sweep: 0 auto (cp.async ring), 1 register loads, 2 ring
  int prefix_threads = 0;        // threads per CTA of the ring variant: 0 auto (even waves), else forced
  int prefix_ring_cfg = 0;       // ring depth / CTAs per SM: 0 auto, 1.. a fixed configuration
};
Tuning& tuning();

// ---- device helpers -----------------------------------------------------------------
// max and ReLU as ATen defines them: a NaN operand gives NaN (fmaxf would drop it).  The engine
// relies on this to notice an fp16 overflow: inf - inf = NaN in the lo half, which must reach
// the logits instead of being flattened to 0 by the next ReLU or pooling window.
// (A/B on one box, fmaxf forms against these: 1252.7 vs 1256.3 ms per bench step — noise.)
__device__ __forceinline__ float nan_max(float a, float b) { return (a > b || a != a) ? a : b; }
__device__ __forceinline__ float nan_relu(float t) { return t < 0.f ? 0.f : t; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

// Make mbarrier initialisation visible to the async (TMA) proxy.
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
      "selp.u32 %0, 1, 0, P1;\n"
      "}"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}

// Bounded wait.  Every legitimate wait in this library ends within microseconds (a TMA tile
// landing, an MMA group retiring, an epilogue draining a TMEM buffer).  A phase that never
// completes — a tensor map that does not describe the buffer it is used on, a byte count that
// does not match the copy — would otherwise spin until the watchdog, if there is one, kills the
// process: after 2^26 failed polls (seconds) the kernel traps instead, the launch fails with a
// sticky CUDA error and the next entry point returns TP_ERR_CUDA.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
#ifdef TP_MBAR_UNBOUNDED   // A/B builds only (profiles/): the round-1 wait, which spins forever
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
  return;
#endif
  if (mbar_try_wait(bar, parity)) return;
  uint32_t polls = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++polls == (1u << 26)) __trap();
  }
}

// Unbounded form for waiters that hold many live registers (the conv epilogue warps keep a
// 128-float accumulator row each): the poll counter of the bounded form costs them a spill.  It
// may only wait for an event that a BOUNDED waiter is also (transitively) waiting for — the
// epilogue waits for MMAs whose issuer waits, bounded, for TMA data — so a dead pipeline still
// ends in a trap.
__device__ __forceinline__ void mbar_wait_unbounded(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// Streaming 128-bit load that does not allocate in L1 (data is read exactly once).
__device__ __forceinline__ float4 ldg_stream4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_stream1(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

}  // namespace tp
