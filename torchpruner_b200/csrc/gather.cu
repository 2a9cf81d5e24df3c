// K7: fused multi-tensor index_select (structured-pruning slicing).
//
// The reference slices one tensor per kernel launch, each preceded by a host->device copy of
// the keep indices (pruner.py:105-112 for param.data / param.grad, opt_pruner.py:16-19 for the
// SGD momentum buffer): ~10-20 launches + as many H2D copies per pruned layer.  Here every
// tensor touched by one prune_model() call is a descriptor in a table that travels in the
// kernel's parameter space, and ONE launch performs all copies.  The copies are bit-exact
// (raw 1/2/4/8/16-byte moves, no arithmetic).  Bandwidth-wise this is a gather that reads
// n_keep/n_src of each source and writes each destination once (2 x bytes kept).
#include "common.cuh"

namespace tp {

struct GatherItem {
  const unsigned char* src;
  unsigned char* dst;
  const int64_t* keep;   // device pointer to this tensor's keep list
  int64_t n_src;
  int64_t n_keep;
  int64_t inner_units;   // units of `1 << unit_log2` bytes per (outer, index)
  int64_t total_units;   // outer * n_keep * inner_units
  int64_t block_first;   // first block of this item in the launch
  int unit_log2;
};

struct GatherBatch {
  GatherItem item[TP_GATHER_MAX_DESCS];
  int n;
};

constexpr int kGatherThreads = 256;
constexpr int kGatherUnroll = 4;

template <typename U>
__device__ __forceinline__ void gather_units(const GatherItem& it, int64_t local_block) {
  const U* __restrict__ src = reinterpret_cast<const U*>(it.src);
  U* __restrict__ dst = reinterpret_cast<U*>(it.dst);
  const int64_t base = local_block * (kGatherThreads * kGatherUnroll) + threadIdx.x;
  const int64_t row_units = it.n_keep * it.inner_units;
  U v[kGatherUnroll];
  bool ok[kGatherUnroll];
#pragma unroll
  for (int k = 0; k < kGatherUnroll; ++k) {
    const int64_t e = base + (int64_t)k * kGatherThreads;
    ok[k] = e < it.total_units;
    if (ok[k]) {
      const int64_t o = e / row_units;
      const int64_t r = e - o * row_units;
      const int64_t j = r / it.inner_units;
      const int64_t i = r - j * it.inner_units;
      const int64_t s = it.keep[j];
      v[k] = src[(o * it.n_src + s) * it.inner_units + i];
    }
  }
#pragma unroll
  for (int k = 0; k < kGatherUnroll; ++k) {
    const int64_t e = base + (int64_t)k * kGatherThreads;
    if (ok[k]) dst[e] = v[k];
  }
}

__global__ void __launch_bounds__(kGatherThreads)
multi_gather_kernel(const __grid_constant__ GatherBatch batch) {
  // Block -> descriptor: linear scan of at most 32 prefix entries held in parameter space.
  int d = 0;
  const int64_t blk = blockIdx.x;
  while (d + 1 < batch.n && blk >= batch.item[d + 1].block_first) ++d;
  const GatherItem& it = batch.item[d];
  const int64_t local_block = blk - it.block_first;
  switch (it.unit_log2) {
    case 4: gather_units<uint4>(it, local_block); break;
    case 3: gather_units<uint2>(it, local_block); break;
    case 2: gather_units<uint32_t>(it, local_block); break;
    case 1: gather_units<uint16_t>(it, local_block); break;
    default: gather_units<uint8_t>(it, local_block); break;
  }
}

}  // namespace tp

extern "C" {

int tp_multi_gather(const tp_gather_desc* descs, int64_t n_desc, const int64_t* keep,
                    int64_t keep_len, tp_stream_t stream) {
  TP_REQUIRE(n_desc >= 0 && keep_len >= 0, TP_ERR_INVALID, "tp_multi_gather: negative count");
  if (n_desc == 0) return TP_OK;
  TP_REQUIRE(descs != nullptr, TP_ERR_INVALID, "tp_multi_gather: descs is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  // Validate the whole table before the first launch so a bad descriptor cannot leave the
  // caller with a half-sliced model.
  for (int64_t i = 0; i < n_desc; ++i) {
    const tp_gather_desc& d = descs[i];
    TP_REQUIRE(d.outer >= 0 && d.n_src >= 0 && d.n_keep >= 0 && d.inner_bytes >= 0 &&
                   d.keep_offset >= 0,
               TP_ERR_INVALID, "tp_multi_gather: descriptor %lld has a negative field",
               (long long)i);
    TP_REQUIRE(d.keep_offset + d.n_keep <= keep_len, TP_ERR_INVALID,
               "tp_multi_gather: descriptor %lld reads past the keep list", (long long)i);
    const int64_t bytes = d.outer * d.n_keep * d.inner_bytes;
    TP_REQUIRE(bytes == 0 || (d.src && d.dst && keep), TP_ERR_INVALID,
               "tp_multi_gather: descriptor %lld has a NULL pointer", (long long)i);
    TP_REQUIRE(bytes == 0 || d.n_src > 0, TP_ERR_INVALID,
               "tp_multi_gather: descriptor %lld gathers from an empty axis", (long long)i);
  }

  int64_t i = 0;
  while (i < n_desc) {
    tp::GatherBatch batch;
    batch.n = 0;
    int64_t blocks = 0;
    while (i < n_desc && batch.n < TP_GATHER_MAX_DESCS) {
      const tp_gather_desc& d = descs[i++];
      const int64_t bytes = d.outer * d.n_keep * d.inner_bytes;
      if (bytes == 0) continue;
      int ul = 4;
      while (ul > 0 && ((d.inner_bytes & ((1ll << ul) - 1)) || ((uintptr_t)d.src & ((1u << ul) - 1)) ||
                        ((uintptr_t)d.dst & ((1u << ul) - 1))))
        --ul;
      tp::GatherItem& it = batch.item[batch.n++];
      it.src = static_cast<const unsigned char*>(d.src);
      it.dst = static_cast<unsigned char*>(d.dst);
      it.keep = keep + d.keep_offset;
      it.n_src = d.n_src;
      it.n_keep = d.n_keep;
      it.inner_units = d.inner_bytes >> ul;
      it.total_units = bytes >> ul;
      it.block_first = blocks;
      it.unit_log2 = ul;
      const int64_t per_block = tp::kGatherThreads * tp::kGatherUnroll;
      blocks += (it.total_units + per_block - 1) / per_block;
    }
    if (batch.n == 0) continue;
    TP_REQUIRE(blocks < (1ll << 31), TP_ERR_LIMIT, "tp_multi_gather: too much work in one launch");
    tp::multi_gather_kernel<<<(unsigned)blocks, tp::kGatherThreads, 0, st>>>(batch);
    TP_LAUNCH_CHECK("multi_gather_kernel");
  }
  return TP_OK;
}

}  // extern "C"
