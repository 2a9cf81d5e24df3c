// K1-K4: single-pass per-row reductions over a flat [R, L] fp32 view of an NCHW tensor
// (R = B*C rows, L = H*W contiguous elements per row), one fp32 result per row.
//
//   Taylor      out[r] = | - sum_l g[r,l]*z[r,l] |      reference taylor.py:40-44
//   APoZ        out[r] = #{l : z[r,l] > 0}              reference apoz.py:31-33
//   Sensitivity out[r] = sum_l |g[r,l]|                 reference sensitivity.py:27-29
//   WeightNorm  out[o] = sum_k |w[o,k]|                 reference weight_norm.py:15-18
//
// All four are HBM-bound streaming reads (algorithmic bytes: 8*R*L+4*R for Taylor, 4*R*L+4*R
// for the others).  Two implementations share the per-element op and the reduction tree:
//
//  * reduce_rows_tma  — persistent CTAs (a multiple of the SM count).  Each CTA walks tiles of
//    whole rows (or 16 KB chunks of one long row); tiles are brought into a shared-memory ring
//    by 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx) issued `stages-1` tiles
//    ahead, so the bytes in flight per SM do not depend on register count or occupancy.  Rows
//    whose length is not a multiple of 4 floats (ResNet 7x7 = 49) still move as 16-byte-aligned
//    tiles because a tile always starts on a multiple of 4 rows.
//  * reduce_rows_ldg  — direct 128-bit (or scalar, for unaligned views) streaming loads with
//    `LPR` lanes cooperating on a row; used for small problems, unaligned pointers and the
//    ragged tail the TMA kernel cannot express.
//
// Summation is fp32 in a fixed order (per-lane strided partials, then a shuffle tree), so the
// result is deterministic run to run; it differs from ATen's order only by fp32 rounding.
#include "common.cuh"

#include <type_traits>

namespace tp {

enum RowOp { OP_TAYLOR = 0, OP_APOZ = 1, OP_ABSSUM = 2 };

template <int OP>
struct RowOpTraits {
  using acc_t = float;
  static constexpr int kStreams = (OP == OP_TAYLOR) ? 2 : 1;
  __device__ __forceinline__ static void add(acc_t& acc, float a, float b) {
    if (OP == OP_TAYLOR) {
      // -(g*z) with the product rounded to fp32 first (no FMA contraction), as torch's mul is.
      acc = __fsub_rn(acc, __fmul_rn(a, b));
    } else {
      acc = __fadd_rn(acc, fabsf(a));
    }
  }
  __device__ __forceinline__ static float finalize(acc_t acc, int is_signed) {
    if (OP == OP_TAYLOR) return is_signed ? acc : fabsf(acc);
    return acc;
  }
};

template <>
struct RowOpTraits<OP_APOZ> {
  using acc_t = int;
  static constexpr int kStreams = 1;
  // one compare and one predicated add per element (false for NaN and -0.0, as the reference's
  // `output > 0`).  Written as `acc += (a > 0.0f)` the compiler emits three: compare, add into a
  // copy, predicated move back — and the consumer warps of the small-row shapes are issue-bound.
  __device__ __forceinline__ static void add(acc_t& acc, float a, float) {
    asm("{\n\t.reg .pred p;\n\tsetp.gt.f32 p, %1, 0f00000000;\n\t@p add.s32 %0, %0, 1;\n\t}"
        : "+r"(acc) : "f"(a));
  }
  __device__ __forceinline__ static float finalize(acc_t acc, int) { return (float)acc; }
};

template <typename T>
__device__ __forceinline__ T group_reduce(T v, int lpr) {
  // Sum over aligned groups of `lpr` lanes (lpr a power of two <= 32).  Every lane takes part.
  for (int off = lpr >> 1; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

template <int OP>
__device__ __forceinline__ void add4(typename RowOpTraits<OP>::acc_t& acc, const float4& a,
                                     const float4& b) {
  RowOpTraits<OP>::add(acc, a.x, b.x);
  RowOpTraits<OP>::add(acc, a.y, b.y);
  RowOpTraits<OP>::add(acc, a.z, b.z);
  RowOpTraits<OP>::add(acc, a.w, b.w);
}

// =====================================================================================
// Direct-load kernel.  LPR lanes cooperate on a row (unit = float4 if VEC else float).
//   LOOP == false: a row has n <= LPR units: one unit per lane, 32/LPR rows per warp pass and 4
//                  passes in flight per warp (row-level memory parallelism for short rows).
//   LOOP == true : n > LPR: each lane strides over its row's units, 4 loads in flight.
// LPR is chosen on the host so that a lane sums ~4-8 units before the shuffle tree starts (the
// tree costs as much as the loads for short rows).
// =====================================================================================
template <int OP, int LPR, bool VEC, bool LOOP>
__global__ void __launch_bounds__(256)
reduce_rows_ldg(const float* __restrict__ a, const float* __restrict__ b,
                float* __restrict__ out, int64_t R, int64_t L, int is_signed) {
  using T = RowOpTraits<OP>;
  using acc_t = typename T::acc_t;
  constexpr bool kTwo = (T::kStreams == 2);
  constexpr int RPW = 32 / LPR;
  // loads in flight per lane: 4 x 16 B vectors per stream for two streams, 8 for one stream, or
  // 8 x 4 B scalars (unaligned / odd-length rows need the deeper unroll to keep enough bytes in
  // flight)
  constexpr int U = (VEC && kTwo) ? 4 : 8;   // one stream: twice the loads to have the same bytes in flight
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPR;
  const int slot = lane / LPR;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (gridDim.x * (int64_t)blockDim.x) >> 5;
  const int n = (int)(VEC ? (L >> 2) : L);

  if (!LOOP) {
    for (int64_t base = warp * (RPW * U); base < R; base += nwarps * (RPW * U)) {
      acc_t acc[U];
      if (VEC) {
        float4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t row = base + u * RPW + slot;
          va[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          vb[u] = va[u];
          if (row < R && sub < n) {
            va[u] = ldg_stream4(reinterpret_cast<const float4*>(a) + row * n + sub);
            if (kTwo) vb[u] = ldg_stream4(reinterpret_cast<const float4*>(b) + row * n + sub);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u] = 0;
          add4<OP>(acc[u], va[u], vb[u]);
        }
      } else {
        float va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t row = base + u * RPW + slot;
          va[u] = 0.f;
          vb[u] = 0.f;
          if (row < R && sub < n) {
            va[u] = ldg_stream1(a + row * n + sub);
            if (kTwo) vb[u] = ldg_stream1(b + row * n + sub);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u] = 0;
          T::add(acc[u], va[u], vb[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const acc_t s = group_reduce(acc[u], LPR);
        const int64_t row = base + u * RPW + slot;
        if (sub == 0 && row < R) out[row] = T::finalize(s, is_signed);
      }
    }
  } else {
    for (int64_t base = warp * RPW; base < R; base += nwarps * RPW) {
      const int64_t row = base + slot;
      const bool live = row < R;
      acc_t acc = 0;
      if (VEC) {
        const float4* pa = reinterpret_cast<const float4*>(a) + row * n;
        const float4* pb = reinterpret_cast<const float4*>(b) + row * n;
        for (int i = sub; i < n; i += LPR * U) {
          float4 va[U], vb[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int j = i + LPR * u;
            va[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            vb[u] = va[u];
            if (live && j < n) {
              va[u] = ldg_stream4(pa + j);
              if (kTwo) vb[u] = ldg_stream4(pb + j);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) add4<OP>(acc, va[u], vb[u]);
        }
      } else {
        const float* pa = a + row * n;
        const float* pb = b + row * n;
        for (int i = sub; i < n; i += LPR * U) {
          float va[U], vb[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int j = i + LPR * u;
            va[u] = 0.f;
            vb[u] = 0.f;
            if (live && j < n) {
              va[u] = ldg_stream1(pa + j);
              if (kTwo) vb[u] = ldg_stream1(pb + j);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) T::add(acc, va[u], vb[u]);
        }
      }
      const acc_t s = group_reduce(acc, LPR);
      if (sub == 0 && live) out[row] = T::finalize(s, is_signed);
    }
  }
}

// =====================================================================================
// TMA-bulk tiled kernel (persistent).  Shared memory: [stages][streams][tile_floats] floats,
// then `stages` mbarriers, then 8 floats/ints of block-reduction scratch.
//   chunk_mode == 0: tile = `tile_rows` whole rows (host guarantees tile_rows*L % 4 == 0 and
//                    the last, shorter tile is also a multiple of 16 bytes).
//   chunk_mode == 1: L > tile_floats and L % 4 == 0; a CTA owns whole rows and streams each
//                    one through the ring in chunks of tile_floats.
// =====================================================================================
constexpr int kTmaThreads = 256;

template <int OP, bool VEC>
__global__ void __launch_bounds__(kTmaThreads)
reduce_rows_tma(const float* __restrict__ a, const float* __restrict__ b,
                float* __restrict__ out, int64_t R, int64_t L, int is_signed, int tile_rows,
                int chunk_mode, int stages, int tile_floats, int lpr) {
  using T = RowOpTraits<OP>;
  using acc_t = typename T::acc_t;
  constexpr int kStreams = T::kStreams;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* ring = reinterpret_cast<float*>(smem_raw);
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(smem_raw + (size_t)stages * kStreams * tile_floats * 4);
  acc_t* red = reinterpret_cast<acc_t*>(bars + stages);

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int wid = tid >> 5;
  const int64_t G = gridDim.x;
  const int64_t n_tiles = chunk_mode ? R : (R + tile_rows - 1) / tile_rows;  // units CTAs stride over
  const int n_chunks = chunk_mode ? (int)((L + tile_floats - 1) / tile_floats) : 1;

  // A cursor over this CTA's tiles.  All index arithmetic is incremental: 64-bit divisions on
  // the per-tile critical path cost more than the tile's loads.
  struct Cursor {
    int64_t unit;  // tile index (row-tile mode) or row index (chunk mode)
    int ch;        // chunk within the row (chunk mode)
  };
  auto valid = [&](const Cursor& c) { return c.unit < n_tiles; };
  auto advance = [&](Cursor& c) {
    if (chunk_mode && ++c.ch < n_chunks) return;
    c.ch = 0;
    c.unit += G;
  };
  auto extent = [&](const Cursor& c, int64_t& first_elem, int& n_floats) {
    if (chunk_mode) {
      const int off = c.ch * tile_floats;
      const int64_t left = L - off;
      first_elem = c.unit * L + off;
      n_floats = left < tile_floats ? (int)left : tile_floats;
    } else {
      const int64_t r0 = c.unit * tile_rows;
      const int64_t rows = (R - r0) < tile_rows ? (R - r0) : tile_rows;
      first_elem = r0 * L;
      n_floats = (int)(rows * L);
    }
  };

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
  }
  __syncthreads();

  Cursor prod{(int64_t)blockIdx.x, 0};  // next tile to request (thread 0 only)
  int prod_stage = 0;
  auto issue = [&]() {
    int64_t first;
    int nf;
    extent(prod, first, nf);
    const uint32_t bytes = (uint32_t)nf * 4u;
    mbar_arrive_expect_tx(&bars[prod_stage], bytes * kStreams);
    float* dst = ring + (size_t)prod_stage * kStreams * tile_floats;
    bulk_g2s(dst, a + first, bytes, &bars[prod_stage]);
    if (kStreams == 2) bulk_g2s(dst + tile_floats, b + first, bytes, &bars[prod_stage]);
    advance(prod);
    if (++prod_stage == stages) prod_stage = 0;
  };
  if (tid == 0) {
    for (int s = 0; s < stages - 1 && valid(prod); ++s) issue();
  }

  Cursor cons{(int64_t)blockIdx.x, 0};
  int stage = 0;
  uint32_t phase = 0;
  acc_t row_acc = 0;  // chunk mode: partial of the row in flight
  const int rpw = 32 / lpr;
  const int sub = lane % lpr;
  const int slot = lane / lpr;
  const int n = VEC ? (int)(L >> 2) : (int)L;

  for (; valid(cons); advance(cons)) {
    // The stage consumed in the previous iteration was released by the barrier that ended it,
    // so it can be refilled now: the ring always has stages-1 tiles in flight.
    if (tid == 0 && valid(prod)) issue();
    mbar_wait(&bars[stage], phase);

    const float* sa = ring + (size_t)stage * kStreams * tile_floats;
    const float* sb = sa + tile_floats;
    int64_t first;
    int nf;
    extent(cons, first, nf);

    if (chunk_mode) {
      const int nv = nf >> 2;
      const float4* sa4 = reinterpret_cast<const float4*>(sa);
      const float4* sb4 = reinterpret_cast<const float4*>(sb);
      for (int v = tid; v < nv; v += kTmaThreads) {
        const float4 va = sa4[v];
        const float4 vb = (kStreams == 2) ? sb4[v] : va;
        add4<OP>(row_acc, va, vb);
      }
      if (cons.ch == n_chunks - 1) {
        const acc_t w = group_reduce(row_acc, 32);
        if (lane == 0) red[wid] = w;
        __syncthreads();
        if (tid == 0) {
          acc_t tot = 0;
          for (int k = 0; k < kTmaThreads / 32; ++k) tot += red[k];
          out[cons.unit] = T::finalize(tot, is_signed);
        }
        row_acc = 0;
      }
    } else {
      const int64_t r0 = cons.unit * tile_rows;
      const int rows = (int)((R - r0) < tile_rows ? (R - r0) : tile_rows);
      // Two row groups per warp iteration: independent chains hide the LDS + shuffle latency.
      for (int rb = wid * rpw; rb < rows; rb += (kTmaThreads / 32) * rpw * 2) {
        const int r_0 = rb + slot;
        const int r_1 = r_0 + (kTmaThreads / 32) * rpw;
        acc_t acc0 = 0, acc1 = 0;
        if (VEC) {
          const float4* pa0 = reinterpret_cast<const float4*>(sa) + (size_t)r_0 * n;
          const float4* pb0 = reinterpret_cast<const float4*>(sb) + (size_t)r_0 * n;
          const float4* pa1 = reinterpret_cast<const float4*>(sa) + (size_t)r_1 * n;
          const float4* pb1 = reinterpret_cast<const float4*>(sb) + (size_t)r_1 * n;
          const bool l0 = r_0 < rows, l1 = r_1 < rows;
          for (int j = sub; j < n; j += lpr) {
            if (l0) {
              const float4 va = pa0[j];
              const float4 vb = (kStreams == 2) ? pb0[j] : va;
              add4<OP>(acc0, va, vb);
            }
            if (l1) {
              const float4 va = pa1[j];
              const float4 vb = (kStreams == 2) ? pb1[j] : va;
              add4<OP>(acc1, va, vb);
            }
          }
        } else {
          const float* pa0 = sa + (size_t)r_0 * n;
          const float* pb0 = sb + (size_t)r_0 * n;
          const float* pa1 = sa + (size_t)r_1 * n;
          const float* pb1 = sb + (size_t)r_1 * n;
          const bool l0 = r_0 < rows, l1 = r_1 < rows;
          for (int j = sub; j < n; j += lpr) {
            if (l0) {
              const float va = pa0[j];
              const float vb = (kStreams == 2) ? pb0[j] : va;
              T::add(acc0, va, vb);
            }
            if (l1) {
              const float va = pa1[j];
              const float vb = (kStreams == 2) ? pb1[j] : va;
              T::add(acc1, va, vb);
            }
          }
        }
        acc0 = group_reduce(acc0, lpr);
        acc1 = group_reduce(acc1, lpr);
        if (sub == 0 && r_0 < rows) out[r0 + r_0] = T::finalize(acc0, is_signed);
        if (sub == 0 && r_1 < rows) out[r0 + r_1] = T::finalize(acc1, is_signed);
      }
    }
    __syncthreads();  // stage fully consumed (and `red` free again)
    if (++stage == stages) {
      stage = 0;
      phase ^= 1u;
    }
  }
}

// =====================================================================================
// Several tensors, one launch.  The per-layer tensors of a VGG-sized model are 2-134 MB: a
// launch per tensor spends as long starting and draining (~4 us of launch gap, first-byte
// latency and tail) as streaming, and the 13-layer sweep of config 2 sat at 0.42-0.56 of the
// HBM peak.  Here the tensors are cut into tiles of whole rows (<= tile_floats floats) that form
// ONE list; persistent CTAs stride over it with the same TMA-bulk ring as reduce_rows_tma, so
// the start-up and the tail are paid once per sweep and tiles of the next tensor are already in
// flight while the last tiles of the previous one are reduced.
// =====================================================================================
struct MultiDesc {
  const float* a;
  const float* b;
  float* out;
  long long R;
  long long end_tile;   // first tile index of the NEXT tensor
  int L;
  int tile_rows;
  int lpr;
  int pad_;
};
struct MultiTable {
  int n;
  int pad_;
  MultiDesc d[TP_REDUCE_MAX_DESCS];
};

template <int OP>
__global__ void __launch_bounds__(kTmaThreads)
reduce_rows_multi_kernel(const __grid_constant__ MultiTable t, int is_signed, int stages,
                         int tile_floats) {
  using T = RowOpTraits<OP>;
  using acc_t = typename T::acc_t;
  constexpr int kStreams = T::kStreams;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* ring = reinterpret_cast<float*>(smem_raw);
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(smem_raw + (size_t)stages * kStreams * tile_floats * 4);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const long long G = gridDim.x;
  const long long total = t.d[t.n - 1].end_tile;

  // (tile, tensor) cursors; tiles only move forward, so the tensor index does too
  struct Cursor {
    long long tile;
    int di;
  };
  auto seek = [&](Cursor& c) {
    while (c.tile < total && c.tile >= t.d[c.di].end_tile) ++c.di;
  };
  auto extent = [&](const Cursor& c, long long& r0, int& rows) {
    const MultiDesc& d = t.d[c.di];
    const long long first = c.di ? t.d[c.di - 1].end_tile : 0;
    r0 = (c.tile - first) * d.tile_rows;
    const long long left = d.R - r0;
    rows = left < d.tile_rows ? (int)left : d.tile_rows;
  };

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
  }
  __syncthreads();

  Cursor prod{(long long)blockIdx.x, 0};
  seek(prod);
  int prod_stage = 0;
  auto issue = [&]() {
    long long r0;
    int rows;
    extent(prod, r0, rows);
    const MultiDesc& d = t.d[prod.di];
    const uint32_t bytes = (uint32_t)rows * (uint32_t)d.L * 4u;
    mbar_arrive_expect_tx(&bars[prod_stage], bytes * kStreams);
    float* dst = ring + (size_t)prod_stage * kStreams * tile_floats;
    bulk_g2s(dst, d.a + r0 * d.L, bytes, &bars[prod_stage]);
    if (kStreams == 2) bulk_g2s(dst + tile_floats, d.b + r0 * d.L, bytes, &bars[prod_stage]);
    prod.tile += G;
    seek(prod);
    if (++prod_stage == stages) prod_stage = 0;
  };
  if (tid == 0) {
    for (int s = 0; s < stages - 1 && prod.tile < total; ++s) issue();
  }

  Cursor cons{(long long)blockIdx.x, 0};
  seek(cons);
  int stage = 0;
  uint32_t phase = 0;
  for (; cons.tile < total; cons.tile += G, seek(cons)) {
    if (tid == 0 && prod.tile < total) issue();
    mbar_wait(&bars[stage], phase);
    const MultiDesc& d = t.d[cons.di];
    long long r0;
    int rows;
    extent(cons, r0, rows);
    const int lpr = d.lpr, rpw = 32 / lpr, sub = lane % lpr, slot = lane / lpr;
    const int n = d.L >> 2;
    const float4* sa = reinterpret_cast<const float4*>(ring + (size_t)stage * kStreams * tile_floats);
    const float4* sb = sa + (tile_floats >> 2);
    float* out = d.out + r0;
    for (int rb = wid * rpw; rb < rows; rb += (kTmaThreads / 32) * rpw * 2) {
      const int r_0 = rb + slot;
      const int r_1 = r_0 + (kTmaThreads / 32) * rpw;
      const bool l0 = r_0 < rows, l1 = r_1 < rows;
      const float4* pa0 = sa + (size_t)r_0 * n;
      const float4* pb0 = sb + (size_t)r_0 * n;
      const float4* pa1 = sa + (size_t)r_1 * n;
      const float4* pb1 = sb + (size_t)r_1 * n;
      acc_t acc0 = 0, acc1 = 0;
      for (int j = sub; j < n; j += lpr) {
        if (l0) {
          const float4 va = pa0[j];
          const float4 vb = (kStreams == 2) ? pb0[j] : va;
          add4<OP>(acc0, va, vb);
        }
        if (l1) {
          const float4 va = pa1[j];
          const float4 vb = (kStreams == 2) ? pb1[j] : va;
          add4<OP>(acc1, va, vb);
        }
      }
      acc0 = group_reduce(acc0, lpr);
      acc1 = group_reduce(acc1, lpr);
      if (sub == 0 && l0) out[r_0] = T::finalize(acc0, is_signed);
      if (sub == 0 && l1) out[r_1] = T::finalize(acc1, is_signed);
    }
    __syncthreads();  // stage fully consumed
    if (++stage == stages) {
      stage = 0;
      phase ^= 1u;
    }
  }
}

// ---- host-side dispatch ---------------------------------------------------------------

static inline int pow2_ceil(int64_t v) {
  int p = 1;
  while (p < v && p < 32) p <<= 1;
  return p;
}

// Lanes per row for a row of n units (float4 if vec else float).  Short rows: one unit per
// lane.  Longer rows: each lane sums several units serially so that the shuffle tree (log2 LPR
// steps, as expensive as the loads for short rows) is amortised; at least 8 lanes per row keeps
// 128-bit shared-memory reads of a quarter-warp inside one row (bank-conflict free).
static inline int lanes_per_row(int64_t n, bool vec) {
  if (n <= 8) return pow2_ceil(n);
  if (n <= 64) return 8;
  if (n <= (vec ? 128 : 256)) return 16;
  return 32;
}

static inline int64_t gcd64(int64_t x, int64_t y) {
  while (y) {
    int64_t t = x % y;
    x = y;
    y = t;
  }
  return x;
}

template <int OP, int LPR, bool VEC, bool LOOP>
static int launch_ldg_inst(const float* a, const float* b, float* out, int64_t R, int64_t L,
                           int is_signed, cudaStream_t st) {
  constexpr int RPW = 32 / LPR;
  const int64_t rows_per_warp =
      LOOP ? RPW : RPW * ((VEC && RowOpTraits<OP>::kStreams == 2) ? 4 : 8);
  const int64_t warps = (R + rows_per_warp - 1) / rows_per_warp;
  int64_t blocks = (warps + 7) / 8;
  const int64_t cap = (int64_t)sm_count() * 8 * tuning().reduce_ldg_waves;  // 8 resident CTAs/SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  reduce_rows_ldg<OP, LPR, VEC, LOOP><<<(unsigned)blocks, 256, 0, st>>>(a, b, out, R, L, is_signed);
  TP_LAUNCH_CHECK("reduce_rows_ldg");
  return TP_OK;
}

template <int OP, bool VEC>
static int launch_ldg_vec(const float* a, const float* b, float* out, int64_t R, int64_t L,
                          int is_signed, cudaStream_t st) {
  const int64_t n = VEC ? (L >> 2) : L;
  const int lpr = lanes_per_row(n, VEC);
  if (n > lpr) {
    switch (lpr) {
      case 8: return launch_ldg_inst<OP, 8, VEC, true>(a, b, out, R, L, is_signed, st);
      case 16: return launch_ldg_inst<OP, 16, VEC, true>(a, b, out, R, L, is_signed, st);
      default: return launch_ldg_inst<OP, 32, VEC, true>(a, b, out, R, L, is_signed, st);
    }
  }
  switch (lpr) {
    case 1: return launch_ldg_inst<OP, 1, VEC, false>(a, b, out, R, L, is_signed, st);
    case 2: return launch_ldg_inst<OP, 2, VEC, false>(a, b, out, R, L, is_signed, st);
    case 4: return launch_ldg_inst<OP, 4, VEC, false>(a, b, out, R, L, is_signed, st);
    default: return launch_ldg_inst<OP, 8, VEC, false>(a, b, out, R, L, is_signed, st);
  }
}

template <int OP>
static int launch_ldg(const float* a, const float* b, float* out, int64_t R, int64_t L,
                      int is_signed, cudaStream_t st) {
  const bool two = RowOpTraits<OP>::kStreams == 2;
  const bool aligned =
      ((uintptr_t)a % 16 == 0) && (!two || (uintptr_t)b % 16 == 0) && (L % 4 == 0);
  if (aligned) return launch_ldg_vec<OP, true>(a, b, out, R, L, is_signed, st);
  return launch_ldg_vec<OP, false>(a, b, out, R, L, is_signed, st);
}

template <int OP, bool VEC>
static int launch_tma_inst(const float* a, const float* b, float* out, int64_t R, int64_t L,
                           int is_signed, int tile_rows, int chunk_mode, int tile_floats,
                           cudaStream_t st) {
  const Tuning& t = tuning();
  constexpr int kStreams = RowOpTraits<OP>::kStreams;
  int ctas = t.reduce_ctas_per_sm;
  const size_t per_stage = (size_t)kStreams * tile_floats * 4;
  // Fewer resident CTAs rather than a ring shallower than two stages.
  while (ctas > 1 && ((size_t)(220 * 1024) / ctas - 256) < 2 * per_stage) --ctas;
  const size_t budget = (size_t)(220 * 1024) / ctas - 256;
  int stages = t.reduce_stages;
  if ((size_t)stages * per_stage > budget) stages = (int)(budget / per_stage);
  TP_REQUIRE(stages >= 2, TP_ERR_INVALID,
             "reduce tuning leaves room for < 2 stages (tile_floats=%d, ctas_per_sm=%d)",
             tile_floats, ctas);
  const size_t smem = (size_t)stages * per_stage + (size_t)stages * 8 + 64;
  const int64_t units = chunk_mode ? R : (R + tile_rows - 1) / tile_rows;
  int64_t grid = (int64_t)sm_count() * ctas;
  if (grid > units) grid = units;
  auto kern = reduce_rows_tma<OP, VEC>;
  TP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int lpr = lanes_per_row(VEC ? (L >> 2) : L, VEC);
  kern<<<(unsigned)grid, kTmaThreads, smem, st>>>(a, b, out, R, L, is_signed, tile_rows, chunk_mode,
                                                  stages, tile_floats, lpr);
  TP_LAUNCH_CHECK("reduce_rows_tma");
  return TP_OK;
}

template <int OP>
static int launch_reduce(const float* a, const float* b, float* out, int64_t R, int64_t L,
                         int is_signed, tp_stream_t stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool two = RowOpTraits<OP>::kStreams == 2;
  TP_REQUIRE(R >= 0 && L >= 0, TP_ERR_INVALID, "negative size (rows=%lld, len=%lld)",
             (long long)R, (long long)L);
  if (R == 0) return TP_OK;
  TP_REQUIRE(out != nullptr, TP_ERR_INVALID, "out is NULL");
  if (L == 0) {  // empty sum
    TP_CUDA(cudaMemsetAsync(out, 0, (size_t)R * 4, st));
    return TP_OK;
  }
  TP_REQUIRE(a != nullptr && (!two || b != nullptr), TP_ERR_INVALID, "input pointer is NULL");
  TP_REQUIRE(L < (1ll << 31) && R < (1ll << 40), TP_ERR_LIMIT, "shape too large");

  const Tuning& t = tuning();
  const bool ptr_aligned = ((uintptr_t)a % 16 == 0) && (!two || (uintptr_t)b % 16 == 0);
  // Auto policy (profiles/r01_tune_reduce_v2.txt): the direct-load kernel is at full HBM speed
  // for rows up to a tile; rows longer than a tile stream fastest through the TMA ring with the
  // whole CTA on one row (1.05-1.09 of the measured copy bandwidth on 50 KB rows).
  bool use_tma = ptr_aligned &&
                 (t.reduce_variant == 2 ||
                  (t.reduce_variant == 0 && L > t.reduce_tile_floats && L % 4 == 0 &&
                   R * L >= t.reduce_tma_min_elems));
  if (use_tma) {
    const int tile_floats = t.reduce_tile_floats;
    if (L <= tile_floats) {
      const int64_t ar = 4 / gcd64(L, 4);  // tiles must start on a 16-byte boundary
      const int64_t tr = (tile_floats / L) / ar * ar;
      if (tr >= 1) {
        const int64_t n_full = R / tr, rem = R % tr;
        const int64_t rows_tma = n_full * tr + (((rem * L) % 4 == 0) ? rem : 0);
        if (rows_tma > 0) {
          int rc = (L % 4 == 0)
                       ? launch_tma_inst<OP, true>(a, b, out, rows_tma, L, is_signed, (int)tr, 0,
                                                   tile_floats, st)
                       : launch_tma_inst<OP, false>(a, b, out, rows_tma, L, is_signed, (int)tr, 0,
                                                    tile_floats, st);
          if (rc != TP_OK) return rc;
        }
        if (rows_tma < R)
          return launch_ldg<OP>(a + rows_tma * L, two ? b + rows_tma * L : b, out + rows_tma,
                                R - rows_tma, L, is_signed, st);
        return TP_OK;
      }
    } else if (L % 4 == 0) {
      return launch_tma_inst<OP, true>(a, b, out, R, L, is_signed, 1, 1, tile_floats, st);
    }
  }
  return launch_ldg<OP>(a, b, out, R, L, is_signed, st);
}

template <int OP>
static int launch_multi(const tp_reduce_desc* descs, int64_t n, int is_signed, cudaStream_t st) {
  constexpr int kStreams = RowOpTraits<OP>::kStreams;
  const Tuning& tn = tuning();
  // per-tile costs (barrier, cursor arithmetic, one thread issuing the copies) are the same for
  // one stream as for two: single-stream ops take tiles of twice the floats, i.e. the same bytes
  const int tile_floats = tn.reduce_tile_floats * (kStreams == 1 ? 2 : 1);
  MultiTable t;
  t.n = 0;
  long long tiles = 0;
  for (int64_t i = 0; i < n; ++i) {
    const tp_reduce_desc& d = descs[i];
    TP_REQUIRE(d.rows >= 0 && d.len >= 0, TP_ERR_INVALID, "tp_reduce_rows_multi: negative size");
    if (d.rows == 0) continue;
    TP_REQUIRE(d.out != nullptr, TP_ERR_INVALID, "tp_reduce_rows_multi: out is NULL");
    if (d.len == 0) {
      TP_CUDA(cudaMemsetAsync(d.out, 0, (size_t)d.rows * 4, st));
      continue;
    }
    TP_REQUIRE(d.a != nullptr && (kStreams == 1 || d.b != nullptr), TP_ERR_INVALID,
               "tp_reduce_rows_multi: input pointer is NULL");
    TP_REQUIRE(d.len % 4 == 0 && d.len <= tile_floats && (uintptr_t)d.a % 16 == 0 &&
                   (kStreams == 1 || (uintptr_t)d.b % 16 == 0),
               TP_ERR_LIMIT,
               "tp_reduce_rows_multi: tensor %lld has rows of %lld floats; rows must be a multiple "
               "of 4 floats, at most %d, in 16-byte aligned tensors (use the single-tensor entry "
               "points otherwise)", (long long)i, (long long)d.len, tile_floats);
    TP_REQUIRE(d.rows < (1ll << 40), TP_ERR_LIMIT, "tp_reduce_rows_multi: too many rows");
    MultiDesc& m = t.d[t.n++];
    m.a = d.a;
    m.b = d.b;
    m.out = d.out;
    m.R = d.rows;
    m.L = (int)d.len;
    m.tile_rows = (int)(tile_floats / d.len);
    m.lpr = lanes_per_row(d.len >> 2, true);
    tiles += (d.rows + m.tile_rows - 1) / m.tile_rows;
    m.end_tile = tiles;
  }
  if (t.n == 0) return TP_OK;
  int ctas = tn.reduce_ctas_per_sm;
  const size_t per_stage = (size_t)kStreams * tile_floats * 4;
  while (ctas > 1 && ((size_t)(220 * 1024) / ctas - 256) < 3 * per_stage) --ctas;
  const size_t budget = (size_t)(220 * 1024) / ctas - 256;
  int stages = 8;
  if ((size_t)stages * per_stage > budget) stages = (int)(budget / per_stage);
  TP_REQUIRE(stages >= 2, TP_ERR_INVALID, "tp_reduce_rows_multi: reduce_tile_floats leaves < 2 stages");
  const size_t smem = (size_t)stages * per_stage + (size_t)stages * 8 + 64;
  long long grid = (long long)sm_count() * ctas;
  if (grid > tiles) grid = tiles;
  auto kern = reduce_rows_multi_kernel<OP>;
  TP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)grid, kTmaThreads, smem, st>>>(t, is_signed, stages, tile_floats);
  TP_LAUNCH_CHECK("reduce_rows_multi_kernel");
  return TP_OK;
}

}  // namespace tp

extern "C" {

int tp_reduce_rows_multi(int op, const tp_reduce_desc* descs, int64_t n, int is_signed,
                         tp_stream_t stream) {
  TP_REQUIRE(n >= 0 && n <= TP_REDUCE_MAX_DESCS, TP_ERR_LIMIT,
             "tp_reduce_rows_multi: at most %d tensors per launch", (int)TP_REDUCE_MAX_DESCS);
  if (n == 0) return TP_OK;
  TP_REQUIRE(descs != nullptr, TP_ERR_INVALID, "tp_reduce_rows_multi: descs is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (op) {
    case TP_REDUCE_TAYLOR: return tp::launch_multi<tp::OP_TAYLOR>(descs, n, is_signed, st);
    case TP_REDUCE_APOZ:
      for (int64_t i = 0; i < n; ++i)
        TP_REQUIRE(descs[i].len < (1 << 24), TP_ERR_LIMIT,
                   "tp_reduce_rows_multi: rows must be shorter than 2^24 for exact counts");
      return tp::launch_multi<tp::OP_APOZ>(descs, n, 0, st);
    case TP_REDUCE_ABSSUM: return tp::launch_multi<tp::OP_ABSSUM>(descs, n, 0, st);
    default:
      tp::set_error("tp_reduce_rows_multi: unknown op %d", op);
      return TP_ERR_INVALID;
  }
}

int tp_taylor_reduce(const float* z, const float* g, int64_t B, int64_t C, int64_t HW,
                     int is_signed, float* out, tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && HW >= 0, TP_ERR_INVALID, "tp_taylor_reduce: negative dimension");
  // a = g, b = z: the op computes -(a*b); multiplication commutes exactly.
  return tp::launch_reduce<tp::OP_TAYLOR>(g, z, out, B * C, HW, is_signed, stream);
}

int tp_apoz_count(const float* z, int64_t B, int64_t C, int64_t HW, float* out,
                  tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && HW >= 0, TP_ERR_INVALID, "tp_apoz_count: negative dimension");
  TP_REQUIRE(HW < (1 << 24), TP_ERR_LIMIT, "tp_apoz_count: HW must be < 2^24 for exact fp32");
  return tp::launch_reduce<tp::OP_APOZ>(z, nullptr, out, B * C, HW, 0, stream);
}

int tp_abs_sum(const float* g, int64_t B, int64_t C, int64_t HW, float* out, tp_stream_t stream) {
  TP_REQUIRE(B >= 0 && C >= 0 && HW >= 0, TP_ERR_INVALID, "tp_abs_sum: negative dimension");
  return tp::launch_reduce<tp::OP_ABSSUM>(g, nullptr, out, B * C, HW, 0, stream);
}

int tp_weight_abs_sum(const float* w, int64_t Cout, int64_t K, float* out, tp_stream_t stream) {
  TP_REQUIRE(Cout >= 0 && K >= 0, TP_ERR_INVALID, "tp_weight_abs_sum: negative dimension");
  return tp::launch_reduce<tp::OP_ABSSUM>(w, nullptr, out, Cout, K, 0, stream);
}

}  // extern "C"
