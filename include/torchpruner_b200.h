/*
 * torchpruner_b200 — C-ABI of the B200 (sm_100a) attribution-scoring / slicing hot path.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream handle
 * (a cudaStream_t passed as void*; NULL = the legacy default stream).  No torch
 * types cross this boundary.  Unless stated otherwise an entry point
 *   - allocates nothing and never synchronises the stream (work is enqueued),
 *   - expects contiguous row-major fp32 tensors,
 *   - returns TP_OK (0) or a negative TP_ERR_* code; tp_last_error() then holds a
 *     thread-local human-readable message.
 *
 * "Reference" citations are file:line under the upstream TorchPruner tree
 * (torchpruner/...), i.e. the eager-PyTorch op chain each kernel replaces.
 */
#ifndef TORCHPRUNER_B200_H
#define TORCHPRUNER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* tp_stream_t; /* cudaStream_t */

enum {
  TP_OK = 0,
  TP_ERR_INVALID = -1, /* bad argument (null pointer, negative size, misuse) */
  TP_ERR_CUDA = -2,    /* a CUDA runtime call / launch failed */
  TP_ERR_ARCH = -3,    /* device is not compute capability 10.x */
  TP_ERR_LIMIT = -4    /* size beyond a documented limit */
};

/* Library version (major*10000 + minor*100 + patch). */
int tp_version(void);

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* tp_last_error(void);

/* TP_OK iff `device` (a CUDA ordinal) is an sm_100-class GPU this library was built for. */
int tp_device_check(int device);

/* Tuning / test hook (process-wide; not part of the reference surface).  Keys:
 *   reduce_variant (0 auto, 1 direct-load, 2 TMA-bulk tiles), reduce_stages (2..8),
 *   reduce_tile_floats (multiple of 256), reduce_ctas_per_sm, reduce_tma_min_elems,
 *   reduce_ldg_waves; conv_variant (0 persistent, 1 one tile per CTA), conv_promote (k-blocks per
 *   TMEM partial sum), conv_pos_major (0 auto, 1 never), conv_pos_major_max_hw, conv_pair (256-row
 *   tiles on CTA pairs: 0 where measured faster, 1 never, 2 wherever possible); prefix_variant
 *   (tp_prefix_contrib_stack: 0 by size, 1 register-load kernel, 2 cp.async ring kernel),
 *   prefix_ring_cfg (0 by size, 1..3 a fixed ring depth / CTAs per SM), prefix_threads (0 even
 *   waves, else a forced CTA size).  Every setting of these gives the same bits.
 * Returns TP_ERR_INVALID for an unknown key or an out-of-range value. */
int tp_set_tuning(const char* key, int64_t value);

/* ---------------------------------------------------------------- K1
 * Taylor: out[b,c] = | - sum_hw g[b,c,hw] * z[b,c,hw] |   (abs skipped when is_signed != 0)
 * Replaces taylor.py:40-44 (grad_output[0] * activation, *-1, flatten(2).sum(-1), abs).
 * The product is rounded to fp32 before summation, as the reference's elementwise mul is.
 * z, g: [B,C,HW] fp32; out: [B,C] fp32.  HW == 1 covers Linear outputs ([B,C]).  */
int tp_taylor_reduce(const float* z, const float* g, int64_t B, int64_t C, int64_t HW,
                     int is_signed, float* out, tp_stream_t stream);

/* ---------------------------------------------------------------- K2
 * APoZ: out[b,c] = #{hw : z[b,c,hw] > 0} as fp32 (exact integer, HW < 2^24).
 * Replaces apoz.py:31-33 ((output > 0).float() summed over every trailing dim). */
int tp_apoz_count(const float* z, int64_t B, int64_t C, int64_t HW, float* out,
                  tp_stream_t stream);

/* ---------------------------------------------------------------- K3
 * Sensitivity: out[b,c] = sum_hw |g[b,c,hw]|.  Replaces sensitivity.py:27-29. */
int tp_abs_sum(const float* g, int64_t B, int64_t C, int64_t HW, float* out,
               tp_stream_t stream);

/* ---------------------------------------------------------------- K4
 * WeightNorm: out[o] = sum_k |w[o,k]|, k = Cin*kh*kw.  Replaces weight_norm.py:15-18
 * (device->host copy of the weight, numpy abs, successive sum(-1)). */
int tp_weight_abs_sum(const float* w, int64_t Cout, int64_t K, float* out, tp_stream_t stream);

/* ---------------------------------------------------------------- K1-K4, several tensors
 * The same per-row reductions for up to TP_REDUCE_MAX_DESCS tensors in ONE launch.  Replaces the
 * per-layer hook arithmetic of taylor.py:40-44 / apoz.py:31-33 / sensitivity.py:27-29 /
 * weight_norm.py:15-18 when several layers are scored from one data pass (the reference runs the
 * whole model once per layer): small activations (2-134 MB on VGG16) are launch-latency bound
 * one at a time.  For tensor i: out[r] = reduce over len of (a[r,:], b[r,:]) with
 *   TP_REDUCE_TAYLOR  |-sum a*b| (a = gradient, b = activation; signed when is_signed != 0)
 *   TP_REDUCE_APOZ    #{a > 0}         (b ignored)
 *   TP_REDUCE_ABSSUM  sum |a|          (b ignored)
 * APoZ counts are exact; the float sums equal the single-tensor entry points' up to the order
 * of the fp32 additions within a row (fixed, so deterministic run to run).  Rows must be a multiple of 4 floats, at most reduce_tile_floats (4096 by
 * default) long, in 16-byte aligned tensors; otherwise TP_ERR_LIMIT and nothing is launched. */
enum { TP_REDUCE_TAYLOR = 0, TP_REDUCE_APOZ = 1, TP_REDUCE_ABSSUM = 2 };
#define TP_REDUCE_MAX_DESCS 64
typedef struct {
  const float* a;
  const float* b;
  float* out;     /* [rows] fp32 */
  int64_t rows;
  int64_t len;
} tp_reduce_desc;
int tp_reduce_rows_multi(int op, const tp_reduce_desc* descs, int64_t n, int is_signed,
                         tp_stream_t stream);

/* ---------------------------------------------------------------- K5
 * Stack of P prefix masks over a cached activation.  For p in [0,P):
 *   out[p,b,c,:] = (inv_perm[c] <= k0 + p) ? +0.0f : z[b,c,:]
 * where inv_perm[c] is the position of channel c in the permutation.  out viewed as
 * [P*B, C, HW] is the batch handed to the downstream layers.  Replaces the clone at
 * shapley_values.py:53 plus P successive index_fill_ calls (:56), incl. the 1-element
 * host->device copy each of them makes.  z: [B,C,HW]; inv_perm: [C] int32 (device). */
int tp_prefix_mask_stack(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P,
                         int64_t B, int64_t C, int64_t HW, float* out, tp_stream_t stream);

/* In-place x[:, idx[j], ...] = value for every j.  x: [B,C,inner]; idx: [n_idx] int64 on the
 * device, each in [0,C).  Replaces index_fill_ at shapley_values.py:95-97 (value 0) and
 * pruner.py:137-142 (value NaN).  `value_bits` is the fp32 bit pattern to store. */
int tp_index_fill_dim1(float* x, int64_t B, int64_t C, int64_t inner, const int64_t* idx,
                       int64_t n_idx, uint32_t value_bits, tp_stream_t stream);

/* ---------------------------------------------------------------- K6
 * Per-sample loss differences of adjacent prefixes, accumulated into the fp64 score table.
 *   L[p,b]   = kind==TP_LOSS_GIVEN ? src[p*B+b]
 *            : cross_entropy(src[(p*B+b)*K .. +K], target[b])     (kind==TP_LOSS_CE)
 *   prev     = p==0 ? prev_in[b] : L[p-1,b]
 *   if p >= scatter_from:
 *     sv[(row0+b)*ld_sv + perm[k0+p]] += (double)((L[p,b]-prev) / sv_samples)   (fp32 sub+div)
 *   prev_out[b] = L[P-1,b]                 (prev_out must not alias prev_in)
 *   if losses_out: losses_out[p*B+b] = L[p,b]
 * Replaces shapley_values.py:57-61 (criterion(reduction="none"), sub, div, .cpu().numpy(),
 * host += per evaluation) and the criterion call of attributions.py:86-87 for cross entropy.
 * Within one call every (row, channel) cell is touched once, so no atomics are used and the
 * accumulation order over permutations is the launch order, as in the reference.
 * perm: [>= k0+P] int64 device; target: [B] int64 device (ignored for TP_LOSS_GIVEN). */
enum { TP_LOSS_GIVEN = 0, TP_LOSS_CE = 1 };
int tp_loss_delta_scatter(const float* src, const int64_t* target, int kind,
                          const float* prev_in, float* prev_out, const int64_t* perm,
                          int64_t k0, int64_t P, int64_t B, int64_t K, float sv_samples,
                          double* sv, int64_t ld_sv, int64_t row0, int64_t scatter_from,
                          float* losses_out, tp_stream_t stream);

/* Ablation / robustness curve (reference experiments notebook "CIFAR-10 - VGG16 - Layerwise
 * robustness", cell 8: per removed unit a forward_partial, a mean cross entropy, an argmax
 * accuracy and two .item() syncs).  For each of P masked copies: mean_loss[p] = mean over the
 * batch of cross_entropy(logits[p,b,:], target[b]); correct[p] = #{b : argmax == target[b]}
 * (first maximum wins ties).  logits: [P,B,K] fp32; target: [B] int64; outputs on the device. */
int tp_ce_mean_acc(const float* logits, const int64_t* target, int64_t P, int64_t B, int64_t K,
                   float* mean_loss, int32_t* correct, tp_stream_t stream);

/* ---------------------------------------------------------------- K7
 * Fused index_select of many tensors in one launch.  Descriptor i copies
 *   dst[o, j, :] = src[o, keep[keep_offset + j], :]   o < outer, j < n_keep
 * with `inner_bytes` contiguous bytes per (o, j).  Replaces one index_select kernel plus one
 * host->device index copy per tensor at pruner.py:105-112 and opt_pruner.py:16-19
 * (weight, bias, grad, BatchNorm running stats, SGD momentum).  Copies are bit-exact.
 * `descs` is a HOST array (it is passed to the kernel by value, <= TP_GATHER_MAX_DESCS per
 * launch; longer tables are split into several launches); `keep` is a DEVICE int64 array. */
typedef struct tp_gather_desc {
  const void* src;
  void* dst;
  int64_t outer;        /* product of dims before the sliced axis            */
  int64_t n_src;        /* length of the sliced axis in src                  */
  int64_t n_keep;       /* length of the sliced axis in dst                  */
  int64_t inner_bytes;  /* bytes per index along the axis (dims after * esz) */
  int64_t keep_offset;  /* first element of this tensor's keep list in keep  */
} tp_gather_desc;
#define TP_GATHER_MAX_DESCS 32
int tp_multi_gather(const tp_gather_desc* descs, int64_t n_desc, const int64_t* keep,
                    int64_t keep_len, tp_stream_t stream);

/* ---------------------------------------------------------------- K8
 * NaN trick, detection side.  For x: [B,C,inner] lists (ascending) every c for which
 * sum over b and inner of x[b,c,:] is NaN: a NaN is present, or both +inf and -inf are.
 * Replaces pruner.py:153-167 (sum(-1)... sum(0), isnan, nonzero).  out_idx: [C] int64,
 * out_count: [1] int64, flags_scratch: [C] int32 work area (contents undefined on return);
 * all three on the device.  The caller reads out_count/out_idx back when it needs them. */
int tp_nan_any_dim1(const float* x, int64_t B, int64_t C, int64_t inner,
                    int32_t* flags_scratch, int64_t* out_idx, int64_t* out_count,
                    tp_stream_t stream);

/* ---------------------------------------------------------------- downstream-layer engine
 * fp32-accurate layers for the batched Shapley evaluations (reference shapley_values.py:57 ->
 * attributions.py:85 re-runs every layer after the evaluation module once per evaluation; for
 * recognised conv3x3/BatchNorm/ReLU/MaxPool/Linear tails those layers run here instead of in
 * cuDNN/cuBLAS).  Activations are NHWC and carried as a TF32 split pair: hi = rn_tf32(x),
 * lo = rn_tf32(x - hi); products are hi*hi + hi*lo + lo*hi on tcgen05 (3xTF32).
 *
 * y = act(conv(x, w) * scale[c] + shift[c]),  conv = 3x3 stride 1 pad 1 (taps == 9) or 1x1 /
 * linear (taps == 1).  x_hi/x_lo: [N,H,W,Cin]; w_hi/w_lo: [Cout][taps*Cin] with k = tap*Cin + c,
 * tap = r*3 + s; scale/shift: [Cout] (BatchNorm and bias folded by the caller); y_hi/y_lo:
 * [N,H,W,Cout].  y_lo == NULL writes the unsplit fp32 result to y_hi (logits, or the input of a
 * pooling step).  act: 0 none, 1 ReLU, 2 LeakyReLU(leaky_slope).  Limits: Cin % 32 == 0 for
 * taps == 9, Cin % 4 == 0 for taps == 1 (the K tail is zero-filled by the TMA unit); W divides
 * 128 and H*W divides 128 or 128/W divides H. */
int tp_conv_tf32x3(const float* x_hi, const float* x_lo, int64_t N, int64_t H, int64_t W,
                   int64_t Cin, const float* w_hi, const float* w_lo, int taps, int64_t Cout,
                   const float* scale, const float* shift, int act, float leaky_slope,
                   float* y_hi, float* y_lo, tp_stream_t stream);

/* 2 x the multiply-accumulates tp_conv_tf32x3 executes for this shape (per product term; the
 * tensor pipe does 3x that).  Small images use position-major tiles that skip taps lying wholly in
 * the zero padding, so this can be less than 2*N*H*W*Cout*taps*Cin.  For roofline accounting. */
int64_t tp_conv_tf32x3_flops(int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int taps);

/* hi[i] = rn_tf32(x[i]), lo[i] = rn_tf32(x[i] - hi[i]); hi + lo == x up to 2^-23 relative. */
int tp_tf32_split(const float* x, int64_t n, float* hi, float* lo, tp_stream_t stream);

/* K5 for the engine above: z is NCHW [B,C,H,W]; out_hi/out_lo are NHWC [P*B, H', W', C_pad]
 * holding the TF32 split of the (optionally 2x2 max-pooled, pool != 0) activation with channels c
 * such that inv_perm[c] <= k0+p zeroed in copy p (k0 = -1 makes copy 0 the unmasked activation).
 * C_pad >= C: channels C..C_pad-1 are zeros (the engine pads channel counts to a multiple of 64
 * so that layers whose width pruning left at, say, 48 or 445 still fit its tiles).
 * Replaces shapley_values.py:53-56 like
 * tp_prefix_mask_stack does, plus the layout change and a leading MaxPool2d(2). */
int tp_mask_stack_nhwc(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P, int64_t B,
                       int64_t C, int64_t H, int64_t W, int pool, int64_t C_pad, float* out_hi,
                       float* out_lo, tp_stream_t stream);

/* MaxPool2d(2) over NHWC raw fp32 [N,H,W,C] followed by the TF32 split -> [N,H/2,W/2,C]. */
int tp_maxpool2_split_nhwc(const float* x, int64_t N, int64_t H, int64_t W, int64_t C, float* hi,
                           float* lo, tp_stream_t stream);

/* fp16 flavour of the engine ("fp16x3").  A TF32 number and an fp16 number both carry 11
 * significant bits, so with a per-tensor power-of-two scale that centres the values in fp16's
 * exponent range the same split (hi = rn_fp16(x*s), lo = rn_fp16(x*s - hi)) and the same three
 * products give the same 2^-22 accuracy — on tcgen05.mma.kind::f16, which runs at twice the TF32
 * rate, from operands that take half the bytes.  x_*, w_* are fp16; `scale` must already contain
 * 1/(s_x*s_w); split outputs (y_lo != NULL) are fp16 pairs of y*out_scale, unsplit output is
 * fp32.  Limits: Cin % 64 == 0 (conv) / % 8 (linear); split outputs need Cout % 64 == 0.
 * The host engine falls back to the TF32 flavour when a tensor leaves fp16's range. */
int tp_conv_f16x3(const void* x_hi, const void* x_lo, int64_t N, int64_t H, int64_t W, int64_t Cin,
                  const void* w_hi, const void* w_lo, int taps, int64_t Cout, const float* scale,
                  const float* shift, int act, float leaky_slope, float out_scale, void* y_hi,
                  void* y_lo, tp_stream_t stream);
/* The same 3x3 layer with the MaxPool2d(2) that follows it (cifar10.py:62-77: every VGG block ends
 * conv-BN-ReLU-MaxPool) fused into the epilogue: y_* are [N, H/2, W/2, Cout] fp16 pairs.  Tiles
 * are position-major (one output pixel x 128 images); the four pixels of a window are computed
 * back to back by the same CTA and their running maximum stays in TMEM, so the full-resolution
 * output never reaches HBM.  Only where tp_conv_pool2_fusable() says 1 (3x3, H*W <= 256, even H
 * and W, N >= 128, Cin % 64 == 0, Cout % 64 == 0); TP_ERR_LIMIT otherwise — the caller then runs
 * tp_conv_f16x3 + tp_maxpool2_split_nhwc_f16. */
int tp_conv_f16x3_pool2(const void* x_hi, const void* x_lo, int64_t N, int64_t H, int64_t W,
                        int64_t Cin, const void* w_hi, const void* w_lo, int64_t Cout,
                        const float* scale, const float* shift, int act, float leaky_slope,
                        float out_scale, void* y_hi, void* y_lo, tp_stream_t stream);
int tp_conv_pool2_fusable(int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout);
/* The last VGG block's conv-BN-ReLU-MaxPool on a 2x2 image (cifar10.py:62-77) in its dense form: a
 * linear layer x [N, Cin] -> four groups of Cout features (one per output pixel; w_* are
 * [4*Cout, Cin], group-major), scale/shift [Cout] shared by the groups, and the maximum over the
 * four groups stored: y_* [N, Cout] fp16 pairs of max_g act(scale*(x.W_g) + shift) * out_scale.
 * The four groups of a tile are computed back to back by the same CTA with their running maximum
 * in TMEM, like tp_conv_f16x3_pool2; same bits as tp_conv_f16x3 on [4*Cout, Cin] followed by
 * tp_maxpool2_split_nhwc_f16 on the [N, 2, 2, Cout] view.  Cout % 64 == 0, Cin % 8 == 0. */
int tp_linear_f16x3_pool4(const void* x_hi, const void* x_lo, int64_t N, int64_t Cin,
                          const void* w_hi, const void* w_lo, int64_t Cout, const float* scale,
                          const float* shift, int act, float leaky_slope, float out_scale,
                          void* y_hi, void* y_lo, tp_stream_t stream);
int tp_f16_split(const float* x, int64_t n, float scale, void* hi, void* lo, tp_stream_t stream);
int tp_mask_stack_nhwc_f16(const float* z, const int32_t* inv_perm, int64_t k0, int64_t P,
                           int64_t B, int64_t C, int64_t H, int64_t W, int pool, float scale,
                           int64_t C_pad, void* out_hi, void* out_lo, tp_stream_t stream);
int tp_maxpool2_split_nhwc_f16(const float* x, int64_t N, int64_t H, int64_t W, int64_t C,
                               float scale, void* hi, void* lo, tp_stream_t stream);
int tp_prefix_contrib_stack_f16(const float* T, const int64_t* perm, int64_t k0, int64_t P,
                                double* base, const float* scale, const float* shift,
                                int64_t pixels, int64_t C1, int64_t Hp, int64_t Wp, int pool,
                                int act, float leaky_slope, float out_scale, void* out_hi,
                                void* out_lo, tp_stream_t stream);

/* ---------------------------------------------------------------- incremental first layer
 * Along one permutation the Shapley masks are nested and the first layer after the evaluation
 * module is linear in the activation, so its pre-activation under mask k is
 * S_k = S_all - sum_{j<=k} T_{perm[j]} with T_c the layer applied to channel c alone.  The
 * reference re-runs the whole layer for every k (shapley_values.py:55-57).
 *
 * tp_channel_contrib: T[c][pixel][o] for every channel c.  z: NCHW [B,C,H,W] (pool != 0: take
 *   MaxPool2d(2) of it first); wt: [C][taps][C1]; mode 0 = conv3x3 stride 1 pad 1 (taps = 9,
 *   pixel = (b,h,w)), mode 1 = dense over all spatial positions (taps = H'*W', pixel = b).
 * tp_contrib_sum: s_all[e] = sum_c T[c][e] in float64 (M1 = pixels*C1 elements).
 * tp_prefix_contrib_stack: for p in [0,P): base -= T[perm[k0+p]] (float64, carried in `base`
 *   between calls); out[p] = tf32_split(act(base*scale + shift)), optionally max-pooled 2x2
 *   (pool != 0, spatial Hp x Wp).  out_hi == out_lo == NULL only advances `base`. */
int tp_channel_contrib(const float* z, const float* wt, int64_t B, int64_t C, int64_t H, int64_t W,
                       int pool, int mode, int64_t C1, float* T, tp_stream_t stream);
int tp_contrib_sum(const float* T, int64_t C, int64_t M1, double* s_all, tp_stream_t stream);
int tp_prefix_contrib_stack(const float* T, const int64_t* perm, int64_t k0, int64_t P,
                            double* base, const float* scale, const float* shift, int64_t pixels,
                            int64_t C1, int64_t Hp, int64_t Wp, int pool, int act,
                            float leaky_slope, float* out_hi, float* out_lo, tp_stream_t stream);

/* Direct form of the same ladder: the contributions are recomputed where they are consumed
 * instead of being stored once per batch (C * pixels * C1 floats, several GB on VGG16) and read
 * back once per permutation — 9 FMAs per output element from L1/L2-resident data against 4 bytes
 * of HBM read per element per mask.  Same arithmetic, FMA for FMA, so the same bits.
 *
 * tp_channel_major: zc[c][b][h][w] = z[b][c][h][w] (pool != 0: MaxPool2d(2) of it), the layout in
 *   which one removed channel's patch is contiguous.
 * tp_prefix_contrib_direct: for p in [0,P): c = perm ? perm[k0+p] : k0+p;
 *   base += sign * T_c (float64, sign = -1 to walk a permutation, +1 with perm == NULL, P == C and
 *   a zeroed base to form S_all); out[p] = split(act(base*scale + shift) [* out_scale]) as in
 *   tp_prefix_contrib_stack (out_f16 != 0: fp16 pairs), optionally max-pooled 2x2 (conv mode).
 *   mode 0: conv3x3 pad 1 over Hp x Wp (wt [C][9][C1]); mode 1: dense over the Hp*Wp positions
 *   (wt [C][Hp*Wp][C1], one output row per example). */
int tp_channel_major(const float* z, int64_t B, int64_t C, int64_t H, int64_t W, int pool, float* zc,
                     tp_stream_t stream);
int tp_prefix_contrib_direct(const float* zc, const float* wt, const int64_t* perm, int64_t k0,
                             int64_t P, double* base, const float* scale, const float* shift,
                             int64_t B, int64_t C, int64_t Hp, int64_t Wp, int mode, int64_t C1,
                             int pool, int act, float leaky_slope, float sign, float out_scale,
                             int out_f16, void* out_hi, void* out_lo, tp_stream_t stream);

/* ---------------------------------------------------------------- no-allocation variant
 * No counterpart in the reference: this entry point exists for callers of THIS library that
 * cannot allocate device memory around the call.  Typical cases: a forward hook that runs in the
 * middle of somebody else's model, where a cudaMalloc/cudaFree pair would synchronise the whole
 * device and stall every other stream; a process whose GPU memory has been handed in full to a
 * framework's pool, so a raw cudaMalloc fails; embedded or real-time hosts whose policy forbids
 * allocation after start-up.  tp_nan_any_dim1 makes such a caller bring a [C] int32 work area;
 * this variant needs none.
 *
 * Same result as tp_nan_any_dim1 (ascending channel list in out_idx, its length in out_count),
 * but the per-channel flag words live in a file-scope `static __device__` array inside the
 * library instead of caller memory.  The caller still provides x, out_idx ([C] int64) and
 * out_count ([1] int64), all on the device.  The call performs no cudaMalloc, no cudaFree and no
 * other allocating routine.
 *
 * Limits and rules:
 *  - C <= TP_NAN_STATIC_MAX_C (65536 channels, 256 KiB of device memory held by every GPU that
 *    loads the library, until the process exits).  Larger C returns TP_ERR_LIMIT and touches
 *    nothing.
 *  - The first call on each device resolves the array's device address (cudaGetSymbolAddress,
 *    with that device current, result cached per device).  Make that first call during set-up,
 *    before any stream capture begins.
 *  - Every call on a device uses the same array.  It keeps the previous call's contents, so each
 *    call clears the C words it needs before use; two calls that overlap on the GPU (different
 *    streams, or several host threads without ordering) race on it and give undefined results.
 *    Calls ordered on one stream are safe.
 *  - The array belongs to the device that is current when the call is made, so that device must
 *    be the one that owns x, out_idx, out_count and the stream.  The call checks it (a pointer
 *    attribute query on out_count, no allocation) and returns TP_ERR_INVALID for buffers of
 *    another GPU or of the host instead of launching on a foreign address. */
#define TP_NAN_STATIC_MAX_C 65536
int tp_nan_any_dim1_static(const float* x, int64_t B, int64_t C, int64_t inner, int64_t* out_idx,
                           int64_t* out_count, tp_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TORCHPRUNER_B200_H */
