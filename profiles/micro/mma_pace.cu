// What paces tcgen05.mma in the conv engine?  A stripped kernel: operands resident in shared
// memory (no loads in the loop), one thread issuing the engine's MMA pattern (per 64-half k-block:
// 4 K-steps x 3 products, M = 128, N = 128 or 256, kind::f16, fp32 accumulate in TMEM), optionally
// while the four epilogue warps drain a TMEM accumulator with tcgen05.ld (as the partial-sum
// promotion does) and/or while 1-D bulk copies keep writing 64 KB "stages" into shared memory (as
// the TMA producer does).  Prints SM clocks per MMA.
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o profiles/micro/mma_pace profiles/micro/mma_pace.cu
#include "../../torchpruner_b200/csrc/common.cuh"
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>

namespace tp { void set_error(const char*, ...) {} }
using namespace tp;

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t a, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(a), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc_sw128(const void* p) {
  const uint64_t addr = (uint64_t)((smem_u32(p) & 0x3FFFFu) >> 4);
  return addr | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
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

// mode bit 0: epilogue warps drain TMEM; bit 1: bulk copies write shared memory meanwhile;
// bit 2: all 12 MMAs of a k-block alternate between TWO accumulators (independent chains)
template <int N>
__global__ void __launch_bounds__(192, 1)
pace_kernel(int kblocks, int mode, const float* __restrict__ src, long long* out_clk,
            unsigned* out_drains) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
  constexpr int A = 128 * 128, B = N * 128;             // bytes per operand tile (64 halves per row)
  uint8_t* a_hi = smem;
  uint8_t* a_lo = a_hi + A;
  uint8_t* b_hi = a_lo + A;
  uint8_t* b_lo = b_hi + B;
  uint8_t* sink = b_lo + B;                              // 64 KB that the bulk copies overwrite
  uint64_t* bars = reinterpret_cast<uint64_t*>(sink + 65536);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 8);
  volatile int* stop = reinterpret_cast<volatile int*>(slot + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (2 * A + 2 * B) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    *stop = 0;
  }
  if (warp == 1) tmem_alloc(slot, 512);
  fence_proxy_async();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (warp == 1 && lane == 0) {
    const uint64_t ah = desc_sw128(a_hi), al = desc_sw128(a_lo), bh = desc_sw128(b_hi), bl = desc_sw128(b_lo);
    const long long t0 = clock64();
    for (int kb = 0; kb < kblocks; ++kb) {
      const uint32_t d0 = tmem + (uint32_t)(((kb >> 1) % (512 / N)) * N);
      const uint32_t d1 = (mode & 4) ? tmem + (uint32_t)((((kb >> 1) + 1) % (512 / N)) * N) : d0;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t off = (uint64_t)(kk * 32 >> 4);
        umma_f16(d0, al + off, bh + off, idesc, (kb | kk) != 0);
        umma_f16(d1, ah + off, bl + off, idesc, 1u);
        umma_f16(d0, ah + off, bh + off, idesc, 1u);
      }
      umma_commit(&bars[kb & 3]);
    }
    umma_commit(&bars[4]);
    mbar_wait(&bars[4], 0);
    const long long t1 = clock64();
    out_clk[blockIdx.x] = t1 - t0;
    *stop = 1;
  } else if (warp >= 2 && (mode & 1)) {
    const int q = warp & 3;
    unsigned n = 0;
    uint32_t keep = 0;
    while (!*stop) {
      uint32_t v[32];
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + 384 + c, v);   // the last accumulator's columns
        keep += v[lane & 31];
      }
      ++n;
    }
    if (lane == 0) out_drains[blockIdx.x * 4 + q] = n + (keep == 0xdeadbeef);
  } else if (warp == 0 && lane == 0 && (mode & 2)) {
    uint32_t phase = 0;
    while (!*stop) {   // 64 KB per round, like one stage of the engine
      mbar_arrive_expect_tx(&bars[5], 65536);
      for (int c = 0; c < 4; ++c) bulk_g2s(sink + c * 16384, src + (size_t)blockIdx.x * 16384 + c * 4096, 16384, &bars[5]);
      mbar_wait(&bars[5], phase);
      phase ^= 1u;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

template <int N>
static void run(int mode, int kblocks, const float* src) {
  long long* clk;
  unsigned* drains;
  cudaMalloc(&clk, 148 * 8);
  cudaMalloc(&drains, 148 * 16);
  cudaMemset(drains, 0, 148 * 16);
  const int smem = 2 * 128 * 128 + 2 * N * 128 + 65536 + 1024 + 256;
  cudaFuncSetAttribute(pace_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) pace_kernel<N><<<148, 192, smem>>>(kblocks, mode, src, clk, drains);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("N=%d mode=%d: %s\n", N, mode, cudaGetErrorString(e)); exit(1); }
  long long h[148];
  unsigned d[148 * 4];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  cudaMemcpy(d, drains, sizeof(d), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  const double per_mma = avg / (12.0 * kblocks);
  printf("N=%3d  drain=%d  smem-writes=%d  two-chains=%d : %7.1f clk per MMA (floor %d) -> %5.1f %% of the pipe;"
         " TMEM drains of 64 KB per k-block pair: %.2f\n",
         N, mode & 1, (mode >> 1) & 1, (mode >> 2) & 1, per_mma, N / 2, 100.0 * (N / 2) / per_mma,
         (mode & 1) ? 2.0 * d[0] / kblocks : 0.0);
  cudaFree(clk);
  cudaFree(drains);
}

int main() {
  float* src;
  cudaMalloc(&src, 148ull * 65536);
  cudaMemset(src, 0, 148ull * 65536);
  const int kb = 4096;
  for (int mode = 0; mode < 8; ++mode) run<128>(mode, kb, src);
  for (int mode = 0; mode < 8; ++mode) run<256>(mode, kb, src);
  return 0;
}
