// Library-level entry points: version, error string, device check, tuning knobs.
#include "common.cuh"

#include <string.h>

#include <atomic>

namespace tp {

static thread_local char g_err[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  // One entry per device ordinal; filled lazily.  cudaDeviceGetAttribute is cheap but this is
  // on the launch path of microsecond kernels.
  // (atomics: entry points may be called from several host threads — autograd runs backward
  // hooks on its own — and two threads filling the same slot write the same value)
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// Process-wide knobs for experiments (tp_set_tuning).  Plain ints read on every launch: set them
// before other threads launch, not concurrently with launches.
Tuning& tuning() {
  static Tuning t;
  return t;
}

}  // namespace tp

extern "C" {

int tp_version(void) { return 100; /* 0.1.0 */ }

const char* tp_last_error(void) { return tp::g_err; }

int tp_device_check(int device) {
  int major = 0, minor = 0;
  TP_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  TP_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  TP_REQUIRE(major == 10, TP_ERR_ARCH,
             "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, major,
             minor);
  return TP_OK;
}

int tp_set_tuning(const char* key, int64_t value) {
  TP_REQUIRE(key != nullptr, TP_ERR_INVALID, "tp_set_tuning: key is NULL");
  tp::Tuning& t = tp::tuning();
  if (!strcmp(key, "reduce_variant")) {
    TP_REQUIRE(value >= 0 && value <= 2, TP_ERR_INVALID, "reduce_variant must be 0, 1 or 2");
    t.reduce_variant = (int)value;
  } else if (!strcmp(key, "reduce_stages")) {
    TP_REQUIRE(value >= 2 && value <= 8, TP_ERR_INVALID, "reduce_stages must be in [2,8]");
    t.reduce_stages = (int)value;
  } else if (!strcmp(key, "reduce_tile_floats")) {
    TP_REQUIRE(value >= 256 && value <= 16384 && value % 256 == 0, TP_ERR_INVALID,
               "reduce_tile_floats must be a multiple of 256 in [256,16384]");
    t.reduce_tile_floats = (int)value;
  } else if (!strcmp(key, "reduce_ctas_per_sm")) {
    TP_REQUIRE(value >= 1 && value <= 8, TP_ERR_INVALID, "reduce_ctas_per_sm must be in [1,8]");
    t.reduce_ctas_per_sm = (int)value;
  } else if (!strcmp(key, "reduce_tma_min_elems")) {
    TP_REQUIRE(value >= 0, TP_ERR_INVALID, "reduce_tma_min_elems must be >= 0");
    t.reduce_tma_min_elems = (int)(value > (1ll << 30) ? (1ll << 30) : value);
  } else if (!strcmp(key, "reduce_ldg_waves")) {
    TP_REQUIRE(value >= 1 && value <= 64, TP_ERR_INVALID, "reduce_ldg_waves must be in [1,64]");
    t.reduce_ldg_waves = (int)value;
  } else if (!strcmp(key, "conv_pos_major_max_hw")) {
    TP_REQUIRE(value >= 1 && value <= 16384, TP_ERR_INVALID, "conv_pos_major_max_hw out of range");
    t.conv_pos_major_max_hw = (int)value;
  } else if (!strcmp(key, "conv_pos_major")) {
    TP_REQUIRE(value == 0 || value == 1, TP_ERR_INVALID, "conv_pos_major must be 0 or 1");
    t.conv_pos_major = (int)value;
  } else if (!strcmp(key, "conv_variant")) {
    TP_REQUIRE(value == 0 || value == 1, TP_ERR_INVALID, "conv_variant must be 0 or 1");
    t.conv_variant = (int)value;
  } else if (!strcmp(key, "conv_pair")) {
    TP_REQUIRE(value >= 0 && value <= 2, TP_ERR_INVALID, "conv_pair must be 0, 1 or 2");
    t.conv_pair = (int)value;
  } else if (!strcmp(key, "prefix_variant")) {
    TP_REQUIRE(value >= 0 && value <= 2, TP_ERR_INVALID, "prefix_variant must be 0, 1 or 2");
    t.prefix_variant = (int)value;
  } else if (!strcmp(key, "prefix_threads")) {
    TP_REQUIRE(value >= 0 && value <= 256 && value % 32 == 0, TP_ERR_INVALID,
               "prefix_threads must be 0 or a multiple of 32 up to 256");
    t.prefix_threads = (int)value;
  } else if (!strcmp(key, "prefix_ring_cfg")) {
    TP_REQUIRE(value >= 0 && value <= 3, TP_ERR_INVALID, "prefix_ring_cfg must be in [0,3]");
    t.prefix_ring_cfg = (int)value;
  } else if (!strcmp(key, "conv_promote")) {
    TP_REQUIRE(value >= 1 && value <= 4096, TP_ERR_INVALID, "conv_promote must be in [1,4096]");
    t.conv_promote = (int)value;
  } else {
    tp::set_error("tp_set_tuning: unknown key '%s'", key);
    return TP_ERR_INVALID;
  }
  return TP_OK;
}

}  // extern "C"
