// Incremental evaluation of the FIRST layer after the evaluation module.
//
// Along one Shapley permutation the masks are nested: mask k removes channels pi_0..pi_k of the
// cached activation z.  The first downstream layer (conv3x3 or linear) is linear in z, so its
// pre-activation under mask k is
//        S_k = S_all - sum_{j<=k} T_{pi_j},      T_c = layer(z restricted to channel c),
// and the whole ladder S_0..S_{C-1} costs ONE layer evaluation (the C per-channel contributions
// T_c, computed once per batch in exact fp32 FMA arithmetic) plus a memory-bound prefix sweep,
// instead of C tensor-core evaluations of the layer — 26 % of all FLOPs of the VGG16 job
// (reference shapley_values.py:55-57 re-runs the full layer for every k).
//
//  * channel_contrib_kernel   T[c][pixel][o]  = sum_taps z'(pixel+tap, c) * Wt[c][tap][o]
//  * contrib_sum_kernel       S_all[e]        = sum_c T[c][e]            (float64)
//  * prefix_contrib_stack     for p < P:  run -= T[perm[k0+p]];  out[p] = split(act(run*a+b))
//                             with the running sum carried in float64 between chunks, the optional
//                             MaxPool2d(2) that follows the layer folded in, and an
//                             accumulate-only mode (no output) to position a sharded rank at the
//                             start of its span.
#include "common.cuh"

#include <cuda_fp16.h>

namespace tp {

__device__ __forceinline__ float inc_tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// streaming 16-byte load of a contribution (read once per mask, never re-used by this CTA)
__device__ __forceinline__ float4 ld_nc_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ void st_cs_u2(void* p, uint2 v) {
  asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// z: NCHW [B, C, H, W]; pool != 0 => z' = MaxPool2d(2)(z), spatial Hp x Wp.
__device__ __forceinline__ float zprime(const float* __restrict__ z, int b, int c, int h, int w,
                                        int C, int H, int W, int pool) {
  const float* src = z + (((long long)b * C + c) * H) * W;
  if (!pool) return src[h * W + w];
  const float* p0 = src + (2 * h) * W + 2 * w;
  return nan_max(nan_max(p0[0], p0[1]), nan_max(p0[W], p0[W + 1]));
}

// mode 0: conv3x3 pad 1 over the Hp x Wp grid, 9 taps, output pixel = (b, h, w)
// mode 1: dense over all Hp*Wp positions (taps = Hp*Wp), output "pixel" = b
// wt: [C][taps][C1]; T: [C][pixels][C1].  One thread per (pixel, 4 outputs).
__global__ void __launch_bounds__(256)
channel_contrib_kernel(const float* __restrict__ z, const float* __restrict__ wt,
                       float* __restrict__ T, int B, int C, int H, int W, int pool, int mode,
                       int C1, long long pixels) {
  const int c = blockIdx.y;
  const int Hp = pool ? H / 2 : H, Wp = pool ? W / 2 : W;
  const int taps = mode == 0 ? 9 : Hp * Wp;
  const unsigned opp = (unsigned)C1 >> 2;  // float4 groups per pixel
  // per-channel work fits 32-bit indices (checked on the host): 64-bit divisions per item would
  // cost more than the 36 FMAs the item performs
  const unsigned items = (unsigned)pixels * opp;
  const float4* w4 = reinterpret_cast<const float4*>(wt) + (long long)c * taps * opp;
  float4* out4 = reinterpret_cast<float4*>(T) + (long long)c * pixels * opp;
  for (unsigned it = blockIdx.x * blockDim.x + threadIdx.x; it < items; it += gridDim.x * blockDim.x) {
    const unsigned pix = it / opp;
    const unsigned o4 = it - pix * opp;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (mode == 0) {
      const int w = (int)(pix % (unsigned)Wp);
      const unsigned t = pix / (unsigned)Wp;
      const int h = (int)(t % (unsigned)Hp);
      const int b = (int)(t / (unsigned)Hp);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int hh = h + r - 1;
        if (hh < 0 || hh >= Hp) continue;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int ww = w + s - 1;
          if (ww < 0 || ww >= Wp) continue;
          const float zv = zprime(z, b, c, hh, ww, C, H, W, pool);
          const float4 wv = w4[(r * 3 + s) * opp + o4];
          acc.x = fmaf(zv, wv.x, acc.x);
          acc.y = fmaf(zv, wv.y, acc.y);
          acc.z = fmaf(zv, wv.z, acc.z);
          acc.w = fmaf(zv, wv.w, acc.w);
        }
      }
    } else {
      const int b = (int)pix;
      for (int tp = 0; tp < taps; ++tp) {
        const float zv = zprime(z, b, c, tp / Wp, tp % Wp, C, H, W, pool);
        const float4 wv = w4[tp * opp + o4];
        acc.x = fmaf(zv, wv.x, acc.x);
        acc.y = fmaf(zv, wv.y, acc.y);
        acc.z = fmaf(zv, wv.z, acc.z);
        acc.w = fmaf(zv, wv.w, acc.w);
      }
    }
    out4[it] = acc;
  }
}

// conv3x3 specialisation: a thread keeps the 9 x 4 weights of its output group in registers and
// walks pixels, so per pixel it issues 9 (warp-broadcast) activation loads, 36 FMAs and one
// 16-byte store.  (The generic kernel above re-reads the 18 KB weight slab of the channel from L1
// for every pixel and is L1-bandwidth bound.)  Requires C1/4 to be a power of two <= 256.
__global__ void __launch_bounds__(256)
channel_contrib_conv_kernel(const float* __restrict__ z, const float* __restrict__ wt,
                            float* __restrict__ T, int B, int C, int H, int W, int pool, int C1,
                            unsigned pixels) {
  const int c = blockIdx.y;
  const int Hp = pool ? H / 2 : H, Wp = pool ? W / 2 : W;
  const unsigned opp = (unsigned)C1 >> 2;
  const unsigned o4 = threadIdx.x % opp;
  const unsigned pl = threadIdx.x / opp;
  const unsigned ppb = blockDim.x / opp;
  const float4* w4 = reinterpret_cast<const float4*>(wt) + (long long)c * 9 * opp;
  float4 wreg[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wreg[t] = w4[t * opp + o4];
  float4* out4 = reinterpret_cast<float4*>(T) + (long long)c * pixels * opp;
  for (unsigned pix = blockIdx.x * ppb + pl; pix < pixels; pix += gridDim.x * ppb) {
    const int w = (int)(pix % (unsigned)Wp);
    const unsigned t = pix / (unsigned)Wp;
    const int h = (int)(t % (unsigned)Hp);
    const int b = (int)(t / (unsigned)Hp);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = h + r - 1;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = w + s - 1;
        const bool in = hh >= 0 && hh < Hp && ww >= 0 && ww < Wp;
        const float zv = in ? zprime(z, b, c, hh, ww, C, H, W, pool) : 0.f;
        const float4 wv = wreg[r * 3 + s];
        acc.x = fmaf(zv, wv.x, acc.x);
        acc.y = fmaf(zv, wv.y, acc.y);
        acc.z = fmaf(zv, wv.z, acc.z);
        acc.w = fmaf(zv, wv.w, acc.w);
      }
    }
    out4[(unsigned long long)pix * opp + o4] = acc;
  }
}

// Same, a thread owning a 2x2 block of output pixels: the 4x4 activation patch is loaded once
// (16 values instead of 36 for the four pixels, and with a pool in front of the layer each value
// is itself four loads and three max), the FMAs per pixel and their order are unchanged.
// Requires even Hp, Wp.
__global__ void __launch_bounds__(256)
channel_contrib_conv2x2_kernel(const float* __restrict__ z, const float* __restrict__ wt,
                               float* __restrict__ T, int B, int C, int H, int W, int pool,
                               int C1, unsigned blocks) {
  const int c = blockIdx.y;
  const int Hp = pool ? H / 2 : H, Wp = pool ? W / 2 : W;
  const int Hb = Hp / 2, Wb = Wp / 2;
  const unsigned opp = (unsigned)C1 >> 2;
  const unsigned o4 = threadIdx.x % opp;
  const unsigned bl = threadIdx.x / opp;
  const unsigned bpb = blockDim.x / opp;
  const float4* w4 = reinterpret_cast<const float4*>(wt) + (long long)c * 9 * opp;
  float4 wreg[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wreg[t] = w4[t * opp + o4];
  float4* out4 = reinterpret_cast<float4*>(T) + (long long)c * ((long long)B * Hp * Wp) * opp;
  for (unsigned blk = blockIdx.x * bpb + bl; blk < blocks; blk += gridDim.x * bpb) {
    const int wb = (int)(blk % (unsigned)Wb);
    const unsigned t = blk / (unsigned)Wb;
    const int hb = (int)(t % (unsigned)Hb);
    const int b = (int)(t / (unsigned)Hb);
    const int h0 = 2 * hb, w0 = 2 * wb;
    float patch[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int hh = h0 + i - 1;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int ww = w0 + j - 1;
        const bool in = hh >= 0 && hh < Hp && ww >= 0 && ww < Wp;
        patch[i][j] = in ? zprime(z, b, c, hh, ww, C, H, W, pool) : 0.f;
      }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const float zv = patch[s / 2 + r][s % 2 + q];
          const float4 wv = wreg[r * 3 + q];
          acc.x = fmaf(zv, wv.x, acc.x);
          acc.y = fmaf(zv, wv.y, acc.y);
          acc.z = fmaf(zv, wv.z, acc.z);
          acc.w = fmaf(zv, wv.w, acc.w);
        }
      const unsigned long long pix =
          ((unsigned long long)b * Hp + h0 + s / 2) * Wp + w0 + s % 2;
      out4[pix * opp + o4] = acc;
    }
  }
}

__global__ void __launch_bounds__(256)
contrib_sum_kernel(const float* __restrict__ T, int C, long long M1, double* __restrict__ s_all) {
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < M1;
       e += gridDim.x * (long long)blockDim.x) {
    double s = 0.0;
    for (int c = 0; c < C; ++c) s += (double)T[(long long)c * M1 + e];
    s_all[e] = s;
  }
}

struct PrefixGeom {
  long long M1;      // elements of one contribution / of the running sum: pixels * C1
  int C1;
  int pool;          // MaxPool2d(2) after the activation: output [B, Hp/2, Wp/2, C1]
  int Hp, Wp;        // spatial size of the layer's output (before that pool)
  int relu;          // activation: 0 none, 1 ReLU, 2 LeakyReLU(slope)
  float slope;
  int emit;          // 0: accumulate only (advance the running sum), 1: write out_hi / out_lo
  float out_scale;   // fp16 outputs hold y * out_scale
};

__device__ __forceinline__ float affine_act(double run, float a, float b, int act, float slope) {
  float t = fmaf((float)run, a, b);
  if (act == 1) return nan_relu(t);
  if (act == 2) return t > 0.f ? t : t * slope;
  return t;
}

// One thread owns 4 consecutive output channels of one output position (after the optional pool)
// and walks the P masks of the chunk with its running sums in float64 registers.
template <bool POOL, bool HALF>
__global__ void __launch_bounds__(256)
prefix_contrib_stack_kernel(const float* __restrict__ T, const int64_t* __restrict__ perm,
                            long long k0, int P, double* __restrict__ base,
                            const float* __restrict__ scale, const float* __restrict__ shift,
                            PrefixGeom g, void* __restrict__ out_hi_v, void* __restrict__ out_lo_v) {
  constexpr int NSRC = POOL ? 4 : 1;
  const int opp = g.C1 >> 2;
  const long long out_pos = POOL ? (g.M1 / g.C1) / 4 : (g.M1 / g.C1);  // output positions
  const long long items = out_pos * opp;
  const long long m_out = out_pos * g.C1;
  for (long long it = blockIdx.x * (long long)blockDim.x + threadIdx.x; it < items;
       it += gridDim.x * (long long)blockDim.x) {
    const long long pos = it / opp;
    const int o4 = (int)(it - pos * opp);
    const float4 a = reinterpret_cast<const float4*>(scale)[o4];
    const float4 b = reinterpret_cast<const float4*>(shift)[o4];
    long long src[NSRC];
    if (POOL) {
      const int Wo = g.Wp / 2, Ho = g.Hp / 2;
      const int wo = (int)(pos % Wo);
      const long long t = pos / Wo;
      const int ho = (int)(t % Ho);
      const long long bi = t / Ho;
      const long long p00 = (bi * g.Hp + 2 * ho) * g.Wp + 2 * wo;
      src[0] = p00;
      src[NSRC > 1 ? 1 : 0] = p00 + 1;
      src[NSRC > 1 ? 2 : 0] = p00 + g.Wp;
      src[NSRC > 1 ? 3 : 0] = p00 + g.Wp + 1;
      src[0] = p00;
    } else {
      src[0] = pos;
    }
    double run[NSRC][4];
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      const double* bp = base + src[s] * g.C1 + o4 * 4;
      run[s][0] = bp[0];
      run[s][1] = bp[1];
      run[s][2] = bp[2];
      run[s][3] = bp[3];
    }
    // U masks per trip: all their contribution loads are issued before the first dependent
    // subtraction, so a thread keeps U x NSRC 16-byte loads in flight instead of NSRC
    constexpr int U = POOL ? 3 : 8;   // (pooled: 3 x 4 loads fit the same 128 registers as 2 x 4, 3-5 % faster)
    for (int p0 = 0; p0 < P; p0 += U) {
      float4 tv[U][NSRC];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long c = perm[k0 + min(p0 + u, P - 1)];
        const float* Tc = T + c * g.M1;
#pragma unroll
        for (int s = 0; s < NSRC; ++s)
          tv[u][s] = ld_nc_f4(Tc + src[s] * g.C1 + o4 * 4);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int p = p0 + u;
        if (p >= P) break;
        float4 y = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
        for (int s = 0; s < NSRC; ++s) {
          const float4 t = tv[u][s];
          run[s][0] -= (double)t.x;
          run[s][1] -= (double)t.y;
          run[s][2] -= (double)t.z;
          run[s][3] -= (double)t.w;
          if (g.emit) {
            y.x = nan_max(y.x, affine_act(run[s][0], a.x, b.x, g.relu, g.slope));
            y.y = nan_max(y.y, affine_act(run[s][1], a.y, b.y, g.relu, g.slope));
            y.z = nan_max(y.z, affine_act(run[s][2], a.z, b.z, g.relu, g.slope));
            y.w = nan_max(y.w, affine_act(run[s][3], a.w, b.w, g.relu, g.slope));
          }
        }
        if (g.emit) {
          const long long o = (long long)p * m_out + pos * g.C1 + o4 * 4;
          if (HALF) {  // fp16 pair of y * out_scale (same 11 significant bits as a TF32 number)
            const float v[4] = {y.x * g.out_scale, y.y * g.out_scale, y.z * g.out_scale,
                                y.w * g.out_scale};
            __half h[4], l[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              h[q] = __float2half_rn(v[q]);
              l[q] = __float2half_rn(v[q] - __half2float(h[q]));
            }
            st_cs_u2(static_cast<__half*>(out_hi_v) + o, *reinterpret_cast<const uint2*>(h));
            st_cs_u2(static_cast<__half*>(out_lo_v) + o, *reinterpret_cast<const uint2*>(l));
          } else {
            const float4 hi = make_float4(inc_tf32_hi(y.x), inc_tf32_hi(y.y), inc_tf32_hi(y.z),
                                          inc_tf32_hi(y.w));
            const float4 lo = make_float4(inc_tf32_hi(y.x - hi.x), inc_tf32_hi(y.y - hi.y),
                                          inc_tf32_hi(y.z - hi.z), inc_tf32_hi(y.w - hi.w));
            *reinterpret_cast<float4*>(static_cast<float*>(out_hi_v) + o) = hi;
            *reinterpret_cast<float4*>(static_cast<float*>(out_lo_v) + o) = lo;
          }
        }
      }
    }
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      double* bp = base + src[s] * g.C1 + o4 * 4;
      bp[0] = run[s][0];
      bp[1] = run[s][1];
      bp[2] = run[s][2];
      bp[3] = run[s][3];
    }
  }
}


// =====================================================================================
// Direct form: the contributions are never stored.  T_c is 9 FMAs per output element from data
// that sits in L1/L2 (one channel of the activation and one 9 x C1 weight slab), which is cheaper
// than writing C*pixels*C1 floats once per batch and reading them back once per permutation (the
// stored form moves 8 bytes per output element per mask, this one 4 — the output itself — and
// needs no C-fold scratch).  The arithmetic is the stored form's, FMA for FMA, so both produce
// the same bits.
//
//   zc : channel-major activation [C][B][Hp][Wp] (after the optional pool in front of the layer)
//   conv mode : a thread owns a BLK x BLK block of output pixels and 4 output channels; per mask
//               it loads the (BLK+2)^2 activation patch of the removed channel and that channel's
//               9 float4 weights, forms the BLK^2 contributions and updates its float64 sums.
//   dense mode: a thread owns one example and 4 output features; taps = Hp*Wp positions.
//   perm == NULL walks the channels in order (used with sign = +1, emit = 0 to form S_all).
// =====================================================================================
__global__ void __launch_bounds__(256)
channel_major_kernel(const float* __restrict__ z, int B, int C, int H, int W, int pool,
                     float* __restrict__ zc) {
  const int Hp = pool ? H / 2 : H, Wp = pool ? W / 2 : W;
  const long long total = (long long)C * B * Hp * Wp;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += gridDim.x * (long long)blockDim.x) {
    const int w = (int)(e % Wp);
    long long t = e / Wp;
    const int h = (int)(t % Hp);
    t /= Hp;
    const int b = (int)(t % B);
    const int c = (int)(t / B);
    zc[e] = zprime(z, b, c, h, w, C, H, W, pool);
  }
}

struct DirectGeom {
  int B, C, Hp, Wp, C1;
  int pool, relu, emit;
  float slope, out_scale, sign;
};

template <bool HALF>
__device__ __forceinline__ void emit_split4(float4 y, float out_scale, void* out_hi_v,
                                            void* out_lo_v, long long o) {
  if (HALF) {
    const float v[4] = {y.x * out_scale, y.y * out_scale, y.z * out_scale, y.w * out_scale};
    __half h[4], l[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      h[q] = __float2half_rn(v[q]);
      l[q] = __float2half_rn(v[q] - __half2float(h[q]));
    }
    st_cs_u2(static_cast<__half*>(out_hi_v) + o, *reinterpret_cast<const uint2*>(h));
    st_cs_u2(static_cast<__half*>(out_lo_v) + o, *reinterpret_cast<const uint2*>(l));
  } else {
    const float4 hi = make_float4(inc_tf32_hi(y.x), inc_tf32_hi(y.y), inc_tf32_hi(y.z),
                                  inc_tf32_hi(y.w));
    const float4 lo = make_float4(inc_tf32_hi(y.x - hi.x), inc_tf32_hi(y.y - hi.y),
                                  inc_tf32_hi(y.z - hi.z), inc_tf32_hi(y.w - hi.w));
    *reinterpret_cast<float4*>(static_cast<float*>(out_hi_v) + o) = hi;
    *reinterpret_cast<float4*>(static_cast<float*>(out_lo_v) + o) = lo;
  }
}

// ---- stored form, contributions staged through a shared-memory ring -------------------------
// Same items, same arithmetic in the same order as prefix_contrib_stack_kernel (bit-identical
// results), but a thread's contribution loads are cp.async copies into its own slots of a D-deep
// ring instead of register loads: D-1 masks x NSRC 16-byte copies stay in flight per thread at all
// times (the register version drains to zero at the end of every U-mask trip), and without the
// U x NSRC float4 staging registers three (pooled) or four CTAs fit an SM instead of two.
// A thread reads back only what it copied itself, so cp.async.wait_group is the only
// synchronisation; slot (p+D-1)%D is refilled one iteration after the same thread read it.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <bool POOL, bool HALF, int D, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
prefix_contrib_ring_kernel(const float* __restrict__ T, const int64_t* __restrict__ perm,
                           long long k0, int P, double* __restrict__ base,
                           const float* __restrict__ scale, const float* __restrict__ shift,
                           PrefixGeom g, void* __restrict__ out_hi_v, void* __restrict__ out_lo_v) {
  constexpr int NSRC = POOL ? 4 : 1;
  extern __shared__ __align__(16) unsigned char ring_raw[];
  float4* ring = reinterpret_cast<float4*>(ring_raw);          // [D][NSRC][blockDim.x]
  const unsigned nthr = blockDim.x;
  const uint32_t ring_u32 = smem_u32(ring) + threadIdx.x * 16u;
  const int opp = g.C1 >> 2;
  const long long out_pos = POOL ? (g.M1 / g.C1) / 4 : (g.M1 / g.C1);
  const long long items = out_pos * opp;
  const long long m_out = out_pos * g.C1;
  // element offsets of the 2x2 window's other three pixels from the first
  const int d1 = g.C1, d2 = g.Wp * g.C1, d3 = d2 + g.C1;
  for (long long it = blockIdx.x * (long long)nthr + threadIdx.x; it < items;
       it += gridDim.x * (long long)nthr) {
    const long long pos = it / opp;
    const int o4 = (int)(it - pos * opp);
    long long off0;
    if (POOL) {
      const int Wo = g.Wp / 2, Ho = g.Hp / 2;
      const int wo = (int)(pos % Wo);
      const long long t = pos / Wo;
      const int ho = (int)(t % Ho);
      const long long bi = t / Ho;
      off0 = ((bi * g.Hp + 2 * ho) * g.Wp + 2 * wo) * g.C1 + o4 * 4;
    } else {
      off0 = pos * g.C1 + o4 * 4;
    }
    auto off = [&](int s) -> long long {
      return off0 + (s == 0 ? 0 : s == 1 ? d1 : s == 2 ? d2 : d3);
    };
    double run[NSRC][4];
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      const double* bp = base + off(s);
      run[s][0] = bp[0];
      run[s][1] = bp[1];
      run[s][2] = bp[2];
      run[s][3] = bp[3];
    }
    const uint32_t stage_bytes = NSRC * nthr * 16u;
    auto issue = [&](int p, int stage) {
      const float* Tc = T + perm[k0 + p] * g.M1;
      const uint32_t slot = ring_u32 + (uint32_t)stage * stage_bytes;
#pragma unroll
      for (int s = 0; s < NSRC; ++s) cp_async16(slot + s * nthr * 16u, Tc + off(s));
    };
#pragma unroll
    for (int d = 0; d < D - 1; ++d) {
      if (d < P) issue(d, d);
      cp_async_commit();
    }
    int fill = D - 1, drain = 0;     // ring stages of mask p+D-1 and of mask p
    for (int p = 0; p < P; ++p) {
      if (p + D - 1 < P) issue(p + D - 1, fill);
      cp_async_commit();
      cp_async_wait<D - 1>();
      const float4* slot = ring + (size_t)drain * NSRC * nthr + threadIdx.x;
      fill = fill + 1 == D ? 0 : fill + 1;
      drain = drain + 1 == D ? 0 : drain + 1;
      float4 y = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
      for (int s = 0; s < NSRC; ++s) {
        const float4 t = slot[s * nthr];
        run[s][0] -= (double)t.x;
        run[s][1] -= (double)t.y;
        run[s][2] -= (double)t.z;
        run[s][3] -= (double)t.w;
        if (g.emit) {
          const float4 a = reinterpret_cast<const float4*>(scale)[o4];
          const float4 b = reinterpret_cast<const float4*>(shift)[o4];
          y.x = nan_max(y.x, affine_act(run[s][0], a.x, b.x, g.relu, g.slope));
          y.y = nan_max(y.y, affine_act(run[s][1], a.y, b.y, g.relu, g.slope));
          y.z = nan_max(y.z, affine_act(run[s][2], a.z, b.z, g.relu, g.slope));
          y.w = nan_max(y.w, affine_act(run[s][3], a.w, b.w, g.relu, g.slope));
        }
      }
      if (g.emit)
        emit_split4<HALF>(y, g.out_scale, out_hi_v, out_lo_v,
                          (long long)p * m_out + pos * g.C1 + o4 * 4);
    }
    cp_async_wait<0>();
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      double* bp = base + off(s);
      bp[0] = run[s][0];
      bp[1] = run[s][1];
      bp[2] = run[s][2];
      bp[3] = run[s][3];
    }
  }
}

template <int BLK, bool HALF>
__global__ void __launch_bounds__(256, 2)
prefix_direct_conv_kernel(const float* __restrict__ zc, const float* __restrict__ wt,
                          const int64_t* __restrict__ perm, long long k0, int P,
                          double* __restrict__ base, const float* __restrict__ scale,
                          const float* __restrict__ shift, DirectGeom g,
                          void* __restrict__ out_hi_v, void* __restrict__ out_lo_v) {
  constexpr int NSRC = BLK * BLK;
  constexpr int PW = BLK + 2;
  const unsigned opp = (unsigned)g.C1 >> 2;
  const int Hb = g.Hp / BLK, Wb = g.Wp / BLK;
  const long long blocks = (long long)g.B * Hb * Wb;
  const long long items = blocks * opp;
  const long long pixels = (long long)g.B * g.Hp * g.Wp;
  const long long m_out = (g.pool ? pixels / 4 : pixels) * g.C1;
  const long long img = (long long)g.Hp * g.Wp;
  const float4* w4 = reinterpret_cast<const float4*>(wt);
  for (long long it = blockIdx.x * (long long)blockDim.x + threadIdx.x; it < items;
       it += gridDim.x * (long long)blockDim.x) {
    const long long blk = it / opp;
    const unsigned o4 = (unsigned)(it - blk * opp);
    const int wb = (int)(blk % Wb);
    const long long t = blk / Wb;
    const int hb = (int)(t % Hb);
    const int b = (int)(t / Hb);
    const int h0 = hb * BLK, w0 = wb * BLK;
    const float4 a = reinterpret_cast<const float4*>(scale)[o4];
    const float4 bsh = reinterpret_cast<const float4*>(shift)[o4];
    long long pix[NSRC];
    double run[NSRC][4];
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      pix[s] = ((long long)b * g.Hp + h0 + s / BLK) * g.Wp + w0 + s % BLK;
      const double* bp = base + pix[s] * g.C1 + o4 * 4;
      run[s][0] = bp[0];
      run[s][1] = bp[1];
      run[s][2] = bp[2];
      run[s][3] = bp[3];
    }
    // which patch entries lie inside the image (the others are the conv's zero padding)
    bool rin[PW], cin[PW];
#pragma unroll
    for (int i = 0; i < PW; ++i) {
      rin[i] = (h0 + i - 1) >= 0 && (h0 + i - 1) < g.Hp;
      cin[i] = (w0 + i - 1) >= 0 && (w0 + i - 1) < g.Wp;
    }
    for (int p = 0; p < P; ++p) {
      const long long c = perm ? perm[k0 + p] : (k0 + p);
      const float* zp = zc + (c * g.B + b) * img + (long long)(h0 - 1) * g.Wp + (w0 - 1);
      float patch[PW][PW];
#pragma unroll
      for (int i = 0; i < PW; ++i)
#pragma unroll
        for (int j = 0; j < PW; ++j)
          patch[i][j] = (rin[i] && cin[j]) ? __ldg(zp + i * g.Wp + j) : 0.f;
      float4 wv[9];
#pragma unroll
      for (int tp = 0; tp < 9; ++tp) wv[tp] = __ldg(w4 + (c * 9 + tp) * opp + o4);
      float4 y = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
      for (int s = 0; s < NSRC; ++s) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            const float zv = patch[s / BLK + r][s % BLK + q];
            const float4 w = wv[r * 3 + q];
            acc.x = fmaf(zv, w.x, acc.x);
            acc.y = fmaf(zv, w.y, acc.y);
            acc.z = fmaf(zv, w.z, acc.z);
            acc.w = fmaf(zv, w.w, acc.w);
          }
        run[s][0] += (double)(g.sign * acc.x);
        run[s][1] += (double)(g.sign * acc.y);
        run[s][2] += (double)(g.sign * acc.z);
        run[s][3] += (double)(g.sign * acc.w);
        if (g.emit) {
          float4 v;
          v.x = affine_act(run[s][0], a.x, bsh.x, g.relu, g.slope);
          v.y = affine_act(run[s][1], a.y, bsh.y, g.relu, g.slope);
          v.z = affine_act(run[s][2], a.z, bsh.z, g.relu, g.slope);
          v.w = affine_act(run[s][3], a.w, bsh.w, g.relu, g.slope);
          if (BLK == 2 && g.pool) {
            y.x = nan_max(y.x, v.x);
            y.y = nan_max(y.y, v.y);
            y.z = nan_max(y.z, v.z);
            y.w = nan_max(y.w, v.w);
          } else {
            emit_split4<HALF>(v, g.out_scale, out_hi_v, out_lo_v,
                              (long long)p * m_out + pix[s] * g.C1 + o4 * 4);
          }
        }
      }
      if (g.emit && BLK == 2 && g.pool)
        emit_split4<HALF>(y, g.out_scale, out_hi_v, out_lo_v,
                          (long long)p * m_out + blk * g.C1 + o4 * 4);
    }
#pragma unroll
    for (int s = 0; s < NSRC; ++s) {
      double* bp = base + pix[s] * g.C1 + o4 * 4;
      bp[0] = run[s][0];
      bp[1] = run[s][1];
      bp[2] = run[s][2];
      bp[3] = run[s][3];
    }
  }
}

// dense mode: wt [C][S][C1], zc [C][B][S]; output row = example
template <bool HALF>
__global__ void __launch_bounds__(256)
prefix_direct_dense_kernel(const float* __restrict__ zc, const float* __restrict__ wt,
                           const int64_t* __restrict__ perm, long long k0, int P,
                           double* __restrict__ base, const float* __restrict__ scale,
                           const float* __restrict__ shift, DirectGeom g,
                           void* __restrict__ out_hi_v, void* __restrict__ out_lo_v) {
  const unsigned opp = (unsigned)g.C1 >> 2;
  const int S = g.Hp * g.Wp;
  const long long items = (long long)g.B * opp;
  const long long m_out = (long long)g.B * g.C1;
  const float4* w4 = reinterpret_cast<const float4*>(wt);
  for (long long it = blockIdx.x * (long long)blockDim.x + threadIdx.x; it < items;
       it += gridDim.x * (long long)blockDim.x) {
    const long long b = it / opp;
    const unsigned o4 = (unsigned)(it - b * opp);
    const float4 a = reinterpret_cast<const float4*>(scale)[o4];
    const float4 bsh = reinterpret_cast<const float4*>(shift)[o4];
    double* bp = base + b * g.C1 + o4 * 4;
    double run[4] = {bp[0], bp[1], bp[2], bp[3]};
    for (int p = 0; p < P; ++p) {
      const long long c = perm ? perm[k0 + p] : (k0 + p);
      const float* zp = zc + (c * g.B + b) * S;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int tp = 0; tp < S; ++tp) {
        const float zv = __ldg(zp + tp);
        const float4 w = __ldg(w4 + (c * S + tp) * opp + o4);
        acc.x = fmaf(zv, w.x, acc.x);
        acc.y = fmaf(zv, w.y, acc.y);
        acc.z = fmaf(zv, w.z, acc.z);
        acc.w = fmaf(zv, w.w, acc.w);
      }
      run[0] += (double)(g.sign * acc.x);
      run[1] += (double)(g.sign * acc.y);
      run[2] += (double)(g.sign * acc.z);
      run[3] += (double)(g.sign * acc.w);
      if (g.emit) {
        float4 v;
        v.x = affine_act(run[0], a.x, bsh.x, g.relu, g.slope);
        v.y = affine_act(run[1], a.y, bsh.y, g.relu, g.slope);
        v.z = affine_act(run[2], a.z, bsh.z, g.relu, g.slope);
        v.w = affine_act(run[3], a.w, bsh.w, g.relu, g.slope);
        emit_split4<HALF>(v, g.out_scale, out_hi_v, out_lo_v,
                          (long long)p * m_out + b * g.C1 + o4 * 4);
      }
    }
    bp[0] = run[0];
    bp[1] = run[1];
    bp[2] = run[2];
    bp[3] = run[3];
  }
}

static inline unsigned inc_grid(long long items, int waves = 16) {
  long long blocks = (items + 255) / 256;
  const long long cap = (long long)sm_count() * waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

// items are walked one per thread for all P masks, so a launch is as long as its number of waves:
// the CTA size is the smallest that keeps that number, which evens the waves out
template <bool POOL, bool HALF, int D, int MAXT, int MINB>
static int launch_prefix_ring(const float* T, const int64_t* perm, long long k0, int P, double* base,
                              const float* scale, const float* shift, const PrefixGeom& g,
                              void* out_hi, void* out_lo, long long items, cudaStream_t st) {
  const long long slots = (long long)sm_count() * MINB;
  const long long waves = (items + slots * MAXT - 1) / (slots * MAXT);
  long long threads = (items + slots * waves - 1) / (slots * waves);
  threads = (threads + 31) / 32 * 32;
  if (threads > MAXT) threads = MAXT;
  if (tuning().prefix_threads) threads = std::min<long long>(tuning().prefix_threads, MAXT);
  const size_t smem = (size_t)D * (POOL ? 4 : 1) * threads * 16;
  long long blocks = (items + threads - 1) / threads;
  const long long cap = slots * 16;
  if (blocks > cap) blocks = cap;
  auto kern = prefix_contrib_ring_kernel<POOL, HALF, D, MAXT, MINB>;
  TP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, (unsigned)threads, smem, st>>>(T, perm, k0, P, base, scale, shift, g,
                                                         out_hi, out_lo);
  TP_LAUNCH_CHECK("prefix_contrib_ring_kernel");
  return TP_OK;
}

// ring configurations: 1..3 pooled (depth 4 x 3 CTAs of 256, depth 3 x 4 CTAs of 224, depth 8 x one
// CTA), 1..2 unpooled (depth 8 x 4 CTAs of 256, depth 16 x 2 CTAs)
template <bool HALF>
static int launch_prefix_ring_cfg(int cfg, bool pool, const float* T, const int64_t* perm,
                                  long long k0, int P, double* base, const float* scale,
                                  const float* shift, const PrefixGeom& g, void* out_hi,
                                  void* out_lo, long long items, cudaStream_t st) {
#define TP_RING(POOL_, D_, MAXT_, MINB_)                                                       \
  return launch_prefix_ring<POOL_, HALF, D_, MAXT_, MINB_>(T, perm, k0, P, base, scale, shift, g, \
                                                           out_hi, out_lo, items, st)
  if (pool) {
    if (cfg == 2) TP_RING(true, 3, 224, 4);
    if (cfg == 3) TP_RING(true, 8, 256, 1);
    TP_RING(true, 4, 256, 3);
  }
  if (cfg == 2) TP_RING(false, 16, 256, 2);
  TP_RING(false, 8, 256, 4);
#undef TP_RING
}

}  // namespace tp

extern "C" {

int tp_channel_contrib(const float* z, const float* wt, int64_t B, int64_t C, int64_t H, int64_t W,
                       int pool, int mode, int64_t C1, float* T, tp_stream_t stream) {
  using namespace tp;
  TP_REQUIRE(B >= 0 && C >= 0 && H > 0 && W > 0 && C1 > 0, TP_ERR_INVALID,
             "tp_channel_contrib: bad dimension");
  TP_REQUIRE(mode == 0 || mode == 1, TP_ERR_INVALID, "tp_channel_contrib: mode must be 0 or 1");
  TP_REQUIRE(C1 % 4 == 0, TP_ERR_LIMIT, "tp_channel_contrib: C1 must be a multiple of 4");
  TP_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), TP_ERR_INVALID,
             "tp_channel_contrib: pooling needs even H and W");
  TP_REQUIRE(C <= 65535, TP_ERR_LIMIT, "tp_channel_contrib: too many channels");
  if (B == 0 || C == 0) return TP_OK;
  TP_REQUIRE(z && wt && T, TP_ERR_INVALID, "tp_channel_contrib: NULL pointer");
  const int Hp = pool ? (int)H / 2 : (int)H, Wp = pool ? (int)W / 2 : (int)W;
  const long long pixels = mode == 0 ? (long long)B * Hp * Wp : (long long)B;
  TP_REQUIRE(pixels * (C1 / 4) < (1ll << 31), TP_ERR_LIMIT,
             "tp_channel_contrib: more than 2^31 items per channel");
  const long long opp = C1 / 4;
  if (mode == 0 && opp <= 256 && (opp & (opp - 1)) == 0 && Hp % 2 == 0 && Wp % 2 == 0) {
    const long long blocks = pixels / 4;
    const long long bpb = 256 / opp;
    long long bx = (blocks + bpb - 1) / bpb;
    const long long cap = ((long long)sm_count() * 8 + C - 1) / C;  // ~8 CTAs/SM over all channels
    if (bx > cap) bx = cap < 1 ? 1 : cap;
    dim3 grid((unsigned)bx, (unsigned)C);
    channel_contrib_conv2x2_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        z, wt, T, (int)B, (int)C, (int)H, (int)W, pool, (int)C1, (unsigned)blocks);
    TP_LAUNCH_CHECK("channel_contrib_conv2x2_kernel");
    return TP_OK;
  }
  if (mode == 0 && opp <= 256 && (opp & (opp - 1)) == 0) {
    const long long ppb = 256 / opp;
    long long bx = (pixels + ppb - 1) / ppb;
    const long long cap = ((long long)sm_count() * 8 + C - 1) / C;  // ~8 CTAs/SM over all channels
    if (bx > cap) bx = cap < 1 ? 1 : cap;
    dim3 grid((unsigned)bx, (unsigned)C);
    channel_contrib_conv_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        z, wt, T, (int)B, (int)C, (int)H, (int)W, pool, (int)C1, (unsigned)pixels);
    TP_LAUNCH_CHECK("channel_contrib_conv_kernel");
    return TP_OK;
  }
  dim3 grid(inc_grid(pixels * (C1 / 4), 4), (unsigned)C);
  channel_contrib_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      z, wt, T, (int)B, (int)C, (int)H, (int)W, pool, mode, (int)C1, pixels);
  TP_LAUNCH_CHECK("channel_contrib_kernel");
  return TP_OK;
}

int tp_contrib_sum(const float* T, int64_t C, int64_t M1, double* s_all, tp_stream_t stream) {
  TP_REQUIRE(C >= 0 && M1 >= 0, TP_ERR_INVALID, "tp_contrib_sum: negative size");
  if (M1 == 0) return TP_OK;
  TP_REQUIRE(T && s_all, TP_ERR_INVALID, "tp_contrib_sum: NULL pointer");
  tp::contrib_sum_kernel<<<tp::inc_grid(M1), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      T, (int)C, M1, s_all);
  TP_LAUNCH_CHECK("contrib_sum_kernel");
  return TP_OK;
}

static int prefix_contrib_stack_impl(const float* T, const int64_t* perm, int64_t k0, int64_t P,
                                     double* base, const float* scale, const float* shift,
                                     int64_t pixels, int64_t C1, int64_t Hp, int64_t Wp, int pool,
                                     int act, float leaky_slope, float out_scale, bool half,
                                     void* out_hi, void* out_lo, tp_stream_t stream) {
  using namespace tp;
  TP_REQUIRE(P >= 0 && k0 >= 0 && pixels >= 0 && C1 > 0, TP_ERR_INVALID,
             "tp_prefix_contrib_stack: bad dimension");
  TP_REQUIRE(C1 % 4 == 0, TP_ERR_LIMIT, "tp_prefix_contrib_stack: C1 must be a multiple of 4");
  TP_REQUIRE(!pool || (Hp % 2 == 0 && Wp % 2 == 0 && Hp > 0 && Wp > 0 && pixels % (Hp * Wp) == 0),
             TP_ERR_INVALID, "tp_prefix_contrib_stack: pooling needs even Hp, Wp dividing pixels");
  TP_REQUIRE((out_hi == nullptr) == (out_lo == nullptr), TP_ERR_INVALID,
             "tp_prefix_contrib_stack: out_hi and out_lo must both be given or both be NULL");
  if (P == 0 || pixels == 0) return TP_OK;
  TP_REQUIRE(T && perm && base && scale && shift, TP_ERR_INVALID,
             "tp_prefix_contrib_stack: NULL pointer");
  // 16-byte vector loads / cp.async of the contributions and the per-channel vectors, vector stores
  TP_REQUIRE((uintptr_t)T % 16 == 0 && (uintptr_t)scale % 16 == 0 && (uintptr_t)shift % 16 == 0 &&
                 (uintptr_t)base % 8 == 0 && (uintptr_t)out_hi % (half ? 8 : 16) == 0 &&
                 (uintptr_t)out_lo % (half ? 8 : 16) == 0,
             TP_ERR_INVALID, "tp_prefix_contrib_stack: T, scale, shift and the outputs must be "
                             "16-byte aligned (fp16 outputs: 8)");
  PrefixGeom g;
  g.M1 = pixels * C1;
  g.C1 = (int)C1;
  g.pool = pool;
  g.Hp = (int)Hp;
  g.Wp = (int)Wp;
  g.relu = act;
  g.slope = leaky_slope;
  g.emit = out_hi != nullptr;
  g.out_scale = out_scale;
  const long long items = (pool ? pixels / 4 : pixels) * (C1 / 4);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Which kernel (profiles/r02_prefix_ring_ab.txt): below ~512 items per SM a launch is one thin
  // wave either way and the register-load kernel is as fast or faster; above, the ring, and for the
  // pooled sweep the 4 x 224-thread configuration where it saves a wave (10 % better filled).
  const long long ipsm = (items + sm_count() - 1) / sm_count();
  const int variant = tuning().prefix_variant;
  if (variant == 2 || (variant == 0 && ipsm >= 512)) {
    int cfg = tuning().prefix_ring_cfg;
    if (cfg == 0) {
      auto fill = [&](long long cap) { return (double)ipsm / (double)((ipsm + cap - 1) / cap * cap); };
      cfg = (pool && fill(4 * 224) > fill(3 * 256) + 0.1) ? 2 : 1;
    }
    return half ? launch_prefix_ring_cfg<true>(cfg, pool != 0, T, perm, k0, (int)P, base, scale,
                                               shift, g, out_hi, out_lo, items, st)
                : launch_prefix_ring_cfg<false>(cfg, pool != 0, T, perm, k0, (int)P, base, scale,
                                                shift, g, out_hi, out_lo, items, st);
  }
  const unsigned grid = inc_grid(items);
  if (pool && half) {
    prefix_contrib_stack_kernel<true, true><<<grid, 256, 0, st>>>(T, perm, k0, (int)P, base, scale,
                                                                  shift, g, out_hi, out_lo);
  } else if (pool) {
    prefix_contrib_stack_kernel<true, false><<<grid, 256, 0, st>>>(T, perm, k0, (int)P, base, scale,
                                                                   shift, g, out_hi, out_lo);
  } else if (half) {
    prefix_contrib_stack_kernel<false, true><<<grid, 256, 0, st>>>(T, perm, k0, (int)P, base, scale,
                                                                   shift, g, out_hi, out_lo);
  } else {
    prefix_contrib_stack_kernel<false, false><<<grid, 256, 0, st>>>(T, perm, k0, (int)P, base,
                                                                    scale, shift, g, out_hi, out_lo);
  }
  TP_LAUNCH_CHECK("prefix_contrib_stack_kernel");
  return TP_OK;
}

int tp_prefix_contrib_stack(const float* T, const int64_t* perm, int64_t k0, int64_t P,
                            double* base, const float* scale, const float* shift, int64_t pixels,
                            int64_t C1, int64_t Hp, int64_t Wp, int pool, int act,
                            float leaky_slope, float* out_hi, float* out_lo, tp_stream_t stream) {
  return prefix_contrib_stack_impl(T, perm, k0, P, base, scale, shift, pixels, C1, Hp, Wp, pool, act,
                                   leaky_slope, 1.f, false, out_hi, out_lo, stream);
}

int tp_prefix_contrib_stack_f16(const float* T, const int64_t* perm, int64_t k0, int64_t P,
                                double* base, const float* scale, const float* shift,
                                int64_t pixels, int64_t C1, int64_t Hp, int64_t Wp, int pool,
                                int act, float leaky_slope, float out_scale, void* out_hi,
                                void* out_lo, tp_stream_t stream) {
  return prefix_contrib_stack_impl(T, perm, k0, P, base, scale, shift, pixels, C1, Hp, Wp, pool, act,
                                   leaky_slope, out_scale, true, out_hi, out_lo, stream);
}

int tp_channel_major(const float* z, int64_t B, int64_t C, int64_t H, int64_t W, int pool, float* zc,
                     tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && H > 0 && W > 0, TP_ERR_INVALID, "tp_channel_major: bad dimension");
  TP_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), TP_ERR_INVALID,
             "tp_channel_major: pooling needs even H and W");
  if (B == 0 || C == 0) return TP_OK;
  TP_REQUIRE(z && zc, TP_ERR_INVALID, "tp_channel_major: NULL pointer");
  const long long total = B * C * (pool ? H / 2 : H) * (pool ? W / 2 : W);
  tp::channel_major_kernel<<<tp::inc_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      z, (int)B, (int)C, (int)H, (int)W, pool, zc);
  TP_LAUNCH_CHECK("channel_major_kernel");
  return TP_OK;
}

int tp_prefix_contrib_direct(const float* zc, const float* wt, const int64_t* perm, int64_t k0,
                             int64_t P, double* base, const float* scale, const float* shift,
                             int64_t B, int64_t C, int64_t Hp, int64_t Wp, int mode, int64_t C1,
                             int pool, int act, float leaky_slope, float sign, float out_scale,
                             int out_f16, void* out_hi, void* out_lo, tp_stream_t stream) {
  using namespace tp;
  TP_REQUIRE(P >= 0 && k0 >= 0 && B >= 0 && C >= 0 && Hp > 0 && Wp > 0 && C1 > 0, TP_ERR_INVALID,
             "tp_prefix_contrib_direct: bad dimension");
  TP_REQUIRE(mode == 0 || mode == 1, TP_ERR_INVALID, "tp_prefix_contrib_direct: mode must be 0 or 1");
  TP_REQUIRE(C1 % 4 == 0, TP_ERR_LIMIT, "tp_prefix_contrib_direct: C1 must be a multiple of 4");
  TP_REQUIRE(k0 + P <= C, TP_ERR_INVALID, "tp_prefix_contrib_direct: k0 + P exceeds the channels");
  TP_REQUIRE(!pool || (mode == 0 && Hp % 2 == 0 && Wp % 2 == 0), TP_ERR_INVALID,
             "tp_prefix_contrib_direct: pooling needs the conv mode and even Hp, Wp");
  TP_REQUIRE((out_hi == nullptr) == (out_lo == nullptr), TP_ERR_INVALID,
             "tp_prefix_contrib_direct: out_hi and out_lo must both be given or both be NULL");
  TP_REQUIRE(sign == 1.f || sign == -1.f, TP_ERR_INVALID,
             "tp_prefix_contrib_direct: sign must be +1 or -1");
  if (P == 0 || B == 0) return TP_OK;
  TP_REQUIRE(zc && wt && base && scale && shift, TP_ERR_INVALID,
             "tp_prefix_contrib_direct: NULL pointer");
  DirectGeom g;
  g.B = (int)B;
  g.C = (int)C;
  g.Hp = (int)Hp;
  g.Wp = (int)Wp;
  g.C1 = (int)C1;
  g.pool = pool;
  g.relu = act;
  g.emit = out_hi != nullptr;
  g.slope = leaky_slope;
  g.out_scale = out_scale;
  g.sign = sign;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool half = out_f16 != 0;
  if (mode == 1) {
    const unsigned grid = inc_grid(B * (C1 / 4));
    if (half)
      prefix_direct_dense_kernel<true><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                            shift, g, out_hi, out_lo);
    else
      prefix_direct_dense_kernel<false><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                             shift, g, out_hi, out_lo);
    TP_LAUNCH_CHECK("prefix_direct_dense_kernel");
    return TP_OK;
  }
  const bool blk2 = Hp % 2 == 0 && Wp % 2 == 0;
  const long long items = (blk2 ? B * (Hp / 2) * (Wp / 2) : B * Hp * Wp) * (C1 / 4);
  const unsigned grid = inc_grid(items);
  if (blk2 && half)
    prefix_direct_conv_kernel<2, true><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                            shift, g, out_hi, out_lo);
  else if (blk2)
    prefix_direct_conv_kernel<2, false><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                             shift, g, out_hi, out_lo);
  else if (half)
    prefix_direct_conv_kernel<1, true><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                            shift, g, out_hi, out_lo);
  else
    prefix_direct_conv_kernel<1, false><<<grid, 256, 0, st>>>(zc, wt, perm, k0, (int)P, base, scale,
                                                             shift, g, out_hi, out_lo);
  TP_LAUNCH_CHECK("prefix_direct_conv_kernel");
  return TP_OK;
}

}  // extern "C"
