// K8: detection side of the pruner's "NaN trick" (reference pruner.py:146-168).
//
// The reference sums a cascading module's input over every dim but dim 1 (successive sum(-1),
// then sum(0)), tests isnan, and calls nonzero() + .cpu() — four reduction launches, a compaction
// and a device->host sync per module.  Here: one pass classifies every element (NaN / +inf /
// -inf) into a per-channel flag word, a second tiny kernel compacts the flagged channels into an
// ascending index list.  A fp32 sum is NaN iff a NaN is present or infinities of both signs are
// (overflow of finite partial sums to opposite infinities is not modelled; it cannot occur for
// the 2-sample probe batch of pruner.py:178 on a finite network).
#include "common.cuh"

#include <mutex>

namespace tp {

// Scratch of tp_nan_any_dim1_static: one flag word per channel, resident in every GPU context
// that loads this library (TP_NAN_STATIC_MAX_C * 4 bytes = 256 KiB) until the process exits.
static __device__ __align__(256) int32_t g_static_nan_flags[TP_NAN_STATIC_MAX_C];

enum { FLAG_NAN = 1, FLAG_PINF = 2, FLAG_NINF = 4 };

__device__ __forceinline__ int classify(float v) {
  const uint32_t u = __float_as_uint(v);
  if ((u & 0x7f800000u) != 0x7f800000u) return 0;  // finite: the overwhelmingly common case
  if (u & 0x007fffffu) return FLAG_NAN;
  return (u >> 31) ? FLAG_NINF : FLAG_PINF;
}

// x: [B, C, inner].  One warp per (b, c) row; lanes stride over `inner`.
__global__ void __launch_bounds__(256)
nan_flags_kernel(const float* __restrict__ x, int64_t rows, int64_t C, int64_t inner,
                 int32_t* __restrict__ flags) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (gridDim.x * (int64_t)blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float* p = x + r * inner;
    int f = 0;
    for (int64_t i = lane; i < inner; i += 32) f |= classify(p[i]);
    f = __reduce_or_sync(0xffffffffu, f);
    if (lane == 0 && f) atomicOr(&flags[r % C], f);
  }
}

// Single CTA: ordered compaction of flagged channels (ascending, like nonzero()).
__global__ void __launch_bounds__(1024)
nan_compact_kernel(const int32_t* __restrict__ flags, int64_t C, int64_t* __restrict__ out_idx,
                   int64_t* __restrict__ out_count) {
  __shared__ int warp_tot[32];
  __shared__ int64_t running;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) running = 0;
  __syncthreads();
  for (int64_t c0 = 0; c0 < C; c0 += blockDim.x) {
    const int64_t c = c0 + tid;
    int hit = 0;
    if (c < C) {
      const int f = flags[c];
      hit = (f & FLAG_NAN) || ((f & (FLAG_PINF | FLAG_NINF)) == (FLAG_PINF | FLAG_NINF));
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    const int before = __popc(m & ((1u << lane) - 1));
    if (lane == 0) warp_tot[wid] = __popc(m);
    __syncthreads();
    int woff = 0, tot = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      const int t = warp_tot[w];
      if (w < wid) woff += t;
      tot += t;
    }
    const int64_t base = running;
    if (hit) out_idx[base + woff + before] = c;
    __syncthreads();
    if (tid == 0) running = base + tot;
    __syncthreads();
  }
  if (tid == 0) *out_count = running;
}

int nan_any_dim1_impl(const float* x, int64_t B, int64_t C, int64_t inner, int32_t* flags,
                      int64_t* out_idx, int64_t* out_count, cudaStream_t st) {
  // Flags must start at zero: they are OR-accumulated.
  TP_CUDA(cudaMemsetAsync(flags, 0, (size_t)C * sizeof(int32_t), st));
  const int64_t rows = B * C;
  if (rows > 0 && inner > 0) {
    int64_t blocks = (rows + 7) / 8;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    nan_flags_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, rows, C, inner, flags);
    TP_LAUNCH_CHECK("nan_flags_kernel");
  }
  nan_compact_kernel<<<1, 1024, 0, st>>>(flags, C, out_idx, out_count);
  TP_LAUNCH_CHECK("nan_compact_kernel");
  return TP_OK;
}

// Device address of g_static_nan_flags on the CURRENT device, resolved once per device.  In host
// code the symbol name is only a handle; the kernels need the address cudaGetSymbolAddress
// returns for the device that is current when it is called.
static int static_flags_address(int32_t** out) {
  static std::mutex mu;
  static int32_t* addr[64] = {nullptr};
  int dev = -1;
  TP_CUDA(cudaGetDevice(&dev));
  TP_REQUIRE(dev >= 0 && dev < 64, TP_ERR_LIMIT, "device ordinal %d is out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    TP_CUDA(cudaGetSymbolAddress(&p, g_static_nan_flags));
    TP_REQUIRE(p != nullptr, TP_ERR_CUDA, "cudaGetSymbolAddress returned a NULL address");
    addr[dev] = static_cast<int32_t*>(p);
  }
  *out = addr[dev];
  return TP_OK;
}

}  // namespace tp

extern "C" {

int tp_nan_any_dim1_static(const float* x, int64_t B, int64_t C, int64_t inner, int64_t* out_idx,
                           int64_t* out_count, tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && inner >= 0, TP_ERR_INVALID,
             "tp_nan_any_dim1_static: negative dimension");
  TP_REQUIRE(C <= TP_NAN_STATIC_MAX_C, TP_ERR_LIMIT,
             "tp_nan_any_dim1_static: C = %lld exceeds the built-in scratch of %d channels; use "
             "tp_nan_any_dim1 with caller-provided scratch instead",
             (long long)C, (int)TP_NAN_STATIC_MAX_C);
  TP_REQUIRE(out_count != nullptr, TP_ERR_INVALID, "tp_nan_any_dim1_static: out_count is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 0) {
    TP_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int64_t), st));
    return TP_OK;
  }
  TP_REQUIRE(out_idx != nullptr, TP_ERR_INVALID, "tp_nan_any_dim1_static: out_idx is NULL");
  TP_REQUIRE(B * inner == 0 || x != nullptr, TP_ERR_INVALID, "tp_nan_any_dim1_static: x is NULL");
  // The scratch lives on the CURRENT device; buffers that belong to another GPU would make the
  // kernels dereference a foreign symbol address.  (Attribute query only: nothing is allocated.)
  {
    int dev = -1;
    TP_CUDA(cudaGetDevice(&dev));
    cudaPointerAttributes attr;
    TP_CUDA(cudaPointerGetAttributes(&attr, out_count));
    TP_REQUIRE((attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) &&
                   (attr.type == cudaMemoryTypeManaged || attr.device == dev),
               TP_ERR_INVALID,
               "tp_nan_any_dim1_static: the buffers are not device memory of the current device "
               "(%d); make the device that owns them current before the call", dev);
  }
  int32_t* flags = nullptr;
  const int rc = tp::static_flags_address(&flags);
  if (rc != TP_OK) return rc;
  // The shared array still holds the previous call's flags; the implementation clears the C
  // words it is about to OR into (stream-ordered memset, no allocation) before the first kernel.
  return tp::nan_any_dim1_impl(x, B, C, inner, flags, out_idx, out_count, st);
}

int tp_nan_any_dim1(const float* x, int64_t B, int64_t C, int64_t inner, int32_t* flags_scratch,
                    int64_t* out_idx, int64_t* out_count, tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && inner >= 0, TP_ERR_INVALID, "tp_nan_any_dim1: negative dimension");
  TP_REQUIRE(out_count != nullptr, TP_ERR_INVALID, "tp_nan_any_dim1: out_count is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 0) {
    TP_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int64_t), st));
    return TP_OK;
  }
  TP_REQUIRE(flags_scratch && out_idx, TP_ERR_INVALID, "tp_nan_any_dim1: NULL pointer");
  TP_REQUIRE(B * inner == 0 || x != nullptr, TP_ERR_INVALID, "tp_nan_any_dim1: x is NULL");
  return tp::nan_any_dim1_impl(x, B, C, inner, flags_scratch, out_idx, out_count, st);
}

}  // extern "C"
