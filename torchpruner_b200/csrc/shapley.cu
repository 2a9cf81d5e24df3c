// K5 / K6: device side of Monte-Carlo Shapley scoring (reference shapley_values.py:28-64).
//
//  * prefix_mask_stack  builds the batch of P cumulatively-masked copies of the cached
//    activation in one launch (write-bound: reads z once, writes P*|z|).
//  * index_fill_dim1    in-place channel fill (slow path shapley_values.py:95-97, and the NaN
//    fill of pruner.py:137-142).
//  * loss_delta_scatter turns the P*B logits rows (or ready-made per-sample losses) into
//    adjacent-prefix loss differences and adds them to the fp64 score table sv[d, C] — the
//    reference's per-evaluation device->host sync and host `+=` (shapley_values.py:58-60)
//    become one launch per chunk of P evaluations and no sync at all.
#include "common.cuh"

#include <math.h>

namespace tp {

// out[p, e] = masked(z[e]) for p in [0,P): thread owns one 16-byte (or 4-byte) unit of z and
// stores it P times, P-stride apart, so every store instruction of a warp is a full 512-byte
// (128-byte for scalars) contiguous run.
template <bool VEC>
__global__ void __launch_bounds__(256)
prefix_mask_stack_kernel(const float* __restrict__ z, const int32_t* __restrict__ inv_perm,
                         int64_t k0, int64_t P, int64_t C, int64_t HWu, int64_t units,
                         float* __restrict__ out) {
  // units = B*C*HWu where HWu = HW/4 (VEC) or HW.
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < units;
       e += gridDim.x * (int64_t)blockDim.x) {
    const int64_t c = (e / HWu) % C;
    const int64_t pos = (int64_t)inv_perm[c] - k0;  // masked in every p >= pos
    if (VEC) {
      const float4 v = reinterpret_cast<const float4*>(z)[e];
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      float4* o = reinterpret_cast<float4*>(out) + e;
#pragma unroll 4
      for (int64_t p = 0; p < P; ++p) o[p * units] = (p >= pos) ? zero : v;
    } else {
      const float v = z[e];
      float* o = out + e;
#pragma unroll 4
      for (int64_t p = 0; p < P; ++p) o[p * units] = (p >= pos) ? 0.0f : v;
    }
  }
}

__global__ void __launch_bounds__(256)
index_fill_dim1_kernel(float* __restrict__ x, int64_t C, int64_t inner,
                       const int64_t* __restrict__ idx, int64_t n_idx, int64_t total,
                       uint32_t value_bits) {
  // total = B * n_idx * inner
  const float value = __uint_as_float(value_bits);
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total;
       e += gridDim.x * (int64_t)blockDim.x) {
    const int64_t i = e % inner;
    const int64_t j = (e / inner) % n_idx;
    const int64_t b = e / (inner * n_idx);
    const int64_t c = idx[j];
    if (c >= 0 && c < C) x[(b * C + c) * inner + i] = value;
  }
}

// Cross entropy of one logits row, computed the way ATen's CPU log_softmax + nll_loss do:
// m = max_k x_k ; lse = log(sum_k exp(x_k - m)) ; loss = -((x_t - m) - lse)   (all fp32).
__device__ __forceinline__ float ce_row(const float* __restrict__ x, int64_t K, int64_t t) {
  if ((uint64_t)t >= (uint64_t)K) return NAN;  // torch raises here; make the misuse visible
  float m = -INFINITY;
  for (int64_t k = 0; k < K; ++k) m = fmaxf(m, x[k]);
  float s = 0.f;
  for (int64_t k = 0; k < K; ++k) s += expf(x[k] - m);
  const float lse = logf(s);
  return -((x[t] - m) - lse);
}

template <int KIND>
__global__ void __launch_bounds__(256)
loss_delta_scatter_kernel(const float* __restrict__ src, const int64_t* __restrict__ target,
                          const float* __restrict__ prev_in, float* __restrict__ prev_out,
                          const int64_t* __restrict__ perm, int64_t k0, int64_t P, int64_t B,
                          int64_t K, float sv_samples, double* __restrict__ sv, int64_t ld_sv,
                          int64_t row0, int64_t scatter_from, float* __restrict__ losses_out) {
  const int64_t total = P * B;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total;
       e += gridDim.x * (int64_t)blockDim.x) {
    const int64_t p = e / B;
    const int64_t b = e % B;
    float cur, prev;
    if (KIND == TP_LOSS_CE) {
      const int64_t t = target[b];
      cur = ce_row(src + e * K, K, t);
      // Recompute the neighbour instead of exchanging it: K is tiny (10 logits) and this keeps
      // the kernel free of inter-thread dependencies and scratch memory.
      prev = (p == 0) ? prev_in[b] : ce_row(src + (e - B) * K, K, t);
    } else {
      cur = src[e];
      prev = (p == 0) ? prev_in[b] : src[e - B];
    }
    if (p >= scatter_from) {
      const float delta = __fsub_rn(cur, prev);
      const float contrib = __fdiv_rn(delta, sv_samples);
      const int64_t ch = perm[k0 + p];
      sv[(row0 + b) * ld_sv + ch] += (double)contrib;
    }
    if (p == P - 1) prev_out[b] = cur;
    if (losses_out) losses_out[e] = cur;
  }
}

// Ablation curve (reference experiments "Layerwise robustness" notebook, cell 8, lines
// 1272-1277): for each of P cumulatively masked copies, the MEAN cross entropy over the batch and
// the number of correct top-1 predictions (first maximum wins ties, like torch.argmax).
// One CTA per mask; fixed-order fp32 summation (per-thread partials, shuffle tree, 8 warp sums).
__global__ void __launch_bounds__(256)
ce_mean_acc_kernel(const float* __restrict__ logits, const int64_t* __restrict__ target, int64_t P,
                   int64_t B, int64_t K, float* __restrict__ mean_loss,
                   int32_t* __restrict__ correct) {
  __shared__ float s_sum[8];
  __shared__ int s_cnt[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    float sum = 0.f;
    int cnt = 0;
    for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
      const float* x = logits + (p * B + b) * K;
      const int64_t t = target[b];
      sum += ce_row(x, K, t);
      int64_t best = 0;
      float bv = x[0];
      for (int64_t k = 1; k < K; ++k) {
        if (x[k] > bv) {
          bv = x[k];
          best = k;
        }
      }
      cnt += (best == t);
    }
    for (int off = 16; off > 0; off >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
      cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    }
    if (lane == 0) {
      s_sum[wid] = sum;
      s_cnt[wid] = cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float tot = 0.f;
      int c = 0;
      for (int w = 0; w < 8; ++w) {
        tot += s_sum[w];
        c += s_cnt[w];
      }
      mean_loss[p] = tot / (float)B;
      correct[p] = c;
    }
    __syncthreads();
  }
}

static inline unsigned grid_for(int64_t work_items, int threads, int waves = 8) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)sm_count() * waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace tp

extern "C" {

int tp_prefix_mask_stack(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P,
                         int64_t B, int64_t C, int64_t HW, float* out, tp_stream_t stream) {
  TP_REQUIRE(P >= 0 && B >= 0 && C >= 0 && HW >= 0 && k0 >= 0, TP_ERR_INVALID,
             "tp_prefix_mask_stack: negative dimension");
  const int64_t n = B * C * HW;
  if (n == 0 || P == 0) return TP_OK;
  TP_REQUIRE(z && inv_perm && out, TP_ERR_INVALID, "tp_prefix_mask_stack: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = (HW % 4 == 0) && ((uintptr_t)z % 16 == 0) && ((uintptr_t)out % 16 == 0);
  if (vec) {
    const int64_t units = n / 4;
    tp::prefix_mask_stack_kernel<true><<<tp::grid_for(units, 256), 256, 0, st>>>(
        z, inv_perm, k0, P, C, HW / 4, units, out);
  } else {
    tp::prefix_mask_stack_kernel<false><<<tp::grid_for(n, 256), 256, 0, st>>>(
        z, inv_perm, k0, P, C, HW, n, out);
  }
  TP_LAUNCH_CHECK("prefix_mask_stack_kernel");
  return TP_OK;
}

int tp_index_fill_dim1(float* x, int64_t B, int64_t C, int64_t inner, const int64_t* idx,
                       int64_t n_idx, uint32_t value_bits, tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && inner >= 0 && n_idx >= 0, TP_ERR_INVALID,
             "tp_index_fill_dim1: negative dimension");
  const int64_t total = B * n_idx * inner;
  if (total == 0) return TP_OK;
  TP_REQUIRE(x && idx, TP_ERR_INVALID, "tp_index_fill_dim1: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  tp::index_fill_dim1_kernel<<<tp::grid_for(total, 256), 256, 0, st>>>(x, C, inner, idx, n_idx,
                                                                        total, value_bits);
  TP_LAUNCH_CHECK("index_fill_dim1_kernel");
  return TP_OK;
}

int tp_loss_delta_scatter(const float* src, const int64_t* target, int kind, const float* prev_in,
                          float* prev_out, const int64_t* perm, int64_t k0, int64_t P, int64_t B,
                          int64_t K, float sv_samples, double* sv, int64_t ld_sv, int64_t row0,
                          int64_t scatter_from, float* losses_out, tp_stream_t stream) {
  TP_REQUIRE(P >= 0 && B >= 0 && k0 >= 0 && row0 >= 0 && ld_sv >= 0 && scatter_from >= 0,
             TP_ERR_INVALID, "tp_loss_delta_scatter: negative size");
  TP_REQUIRE(kind == TP_LOSS_GIVEN || kind == TP_LOSS_CE, TP_ERR_INVALID,
             "tp_loss_delta_scatter: unknown kind %d", kind);
  if (P == 0 || B == 0) return TP_OK;
  TP_REQUIRE(src && prev_in && prev_out && perm && sv, TP_ERR_INVALID,
             "tp_loss_delta_scatter: NULL pointer");
  TP_REQUIRE(prev_in != prev_out, TP_ERR_INVALID,
             "tp_loss_delta_scatter: prev_out must not alias prev_in");
  TP_REQUIRE(kind != TP_LOSS_CE || (target && K >= 1), TP_ERR_INVALID,
             "tp_loss_delta_scatter: cross entropy needs targets and K >= 1");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned grid = tp::grid_for(P * B, 256);
  if (kind == TP_LOSS_CE) {
    tp::loss_delta_scatter_kernel<TP_LOSS_CE><<<grid, 256, 0, st>>>(
        src, target, prev_in, prev_out, perm, k0, P, B, K, sv_samples, sv, ld_sv, row0,
        scatter_from, losses_out);
  } else {
    tp::loss_delta_scatter_kernel<TP_LOSS_GIVEN><<<grid, 256, 0, st>>>(
        src, target, prev_in, prev_out, perm, k0, P, B, K, sv_samples, sv, ld_sv, row0,
        scatter_from, losses_out);
  }
  TP_LAUNCH_CHECK("loss_delta_scatter_kernel");
  return TP_OK;
}

int tp_ce_mean_acc(const float* logits, const int64_t* target, int64_t P, int64_t B, int64_t K,
                   float* mean_loss, int32_t* correct, tp_stream_t stream) {
  TP_REQUIRE(P >= 0 && B >= 0 && K >= 1, TP_ERR_INVALID, "tp_ce_mean_acc: bad dimension");
  if (P == 0) return TP_OK;
  TP_REQUIRE(B > 0, TP_ERR_INVALID, "tp_ce_mean_acc: empty batch");
  TP_REQUIRE(logits && target && mean_loss && correct, TP_ERR_INVALID,
             "tp_ce_mean_acc: NULL pointer");
  int64_t blocks = P;
  const int64_t cap = (int64_t)tp::sm_count() * 8;
  if (blocks > cap) blocks = cap;
  tp::ce_mean_acc_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, target, P, B, K, mean_loss, correct);
  TP_LAUNCH_CHECK("ce_mean_acc_kernel");
  return TP_OK;
}

}  // extern "C"
