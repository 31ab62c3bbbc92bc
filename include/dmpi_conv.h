/* C ABI of the tcgen05 implicit-GEMM engine (csrc/kernels/conv).
 *
 * One descriptor drives three kernels:
 *   dmpi_igemm_*  : D[pixels, n_out] = sum_taps A_tap[pixels, Cin] * B[n_out, tap*Cin..]^T   (K-major operands)
 *                   = Conv2d fprop, Conv2d dgrad (stride 1 or parity-decomposed stride 2), Linear fwd/dgrad,
 *                   ConvTranspose2d(k=2,s=2) fwd/dgrad.   Replaces the cuDNN/cuBLAS calls behind
 *                   ref:pytorch/unet/model.py:8-14,37-38,68 and torchvision resnet conv/fc
 *                   (K1, K2, K4, K10).
 *   dmpi_wgrad_*  : dW[k, tap, c] += sum_pixels dY[pixels, k] * X_tap[pixels, c]            (MN-major operands)
 *                   = Conv2d / Linear / ConvTranspose2d weight gradients, accumulated with
 *                   fp32 red.global straight into the gradient-bucket slot of the parameter.
 *
 * Activations are NHWC bf16.  Every operand is described as a rank-4 strided tensor (c, w, h, n); the
 * kernels move tiles with TMA (zero fill outside the tensor gives the convolution halo).
 */
#ifndef DMPI_CONV_H_
#define DMPI_CONV_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DMPI_MAX_TAPS 32
#define DMPI_MAX_MAPS 4

typedef struct {
  int64_t dims[4];    /* c, w, h, n (elements) */
  int64_t strides[3]; /* element strides of w, h, n */
  int64_t offset;     /* element offset from the base pointer passed at run time */
} dmpi_tensor4;

typedef struct {
  /* A operand: up to 4 views of the same base pointer (stride-2 convs use one view per parity class). */
  int a_nviews;
  dmpi_tensor4 a_view[DMPI_MAX_MAPS];
  /* pixel box of one M tile: rows = box_w*box_h*box_n (<= 128 for igemm, == 64 recommended for wgrad). */
  int box_w, box_h, box_n;
  /* output-pixel grid that the tiles cover (igemm: output dims; wgrad: dY dims) */
  int grid_w, grid_h, grid_n;
  /* filter taps: for tap t the A box origin is (cb*64, w0+dw, h0+dh, n0) in view tap_view[t]; the matching
   * B columns start at tap_koff[t] (elements along B's K axis).  cblocks = Cin/64 K-blocks per tap. */
  int ntaps, cblocks;
  int tap_view[DMPI_MAX_TAPS];
  int tap_dw[DMPI_MAX_TAPS];
  int tap_dh[DMPI_MAX_TAPS];
  int tap_koff[DMPI_MAX_TAPS];
  /* B operand (igemm): [b_rows, b_k] bf16, K contiguous. Rows past b_rows read as zero. */
  int64_t b_rows, b_k;
  /* outputs (igemm): up to 4 views of the output base pointer; N tile j writes view j / nblk_per_oview at
   * channel (j % nblk_per_oview) * bn. */
  int o_nviews, nblk_per_oview;
  dmpi_tensor4 o_view[DMPI_MAX_MAPS];
  int n_out; /* number of valid output columns */
  int bn;    /* N tile: 64, 128 or 256 */
  /* epilogue */
  int bias_mod;  /* bias index = column % bias_mod (0: no bias) */
  int stat_mod;  /* per-channel sum / sum-of-squares of the stored bf16 values at column % stat_mod (0: off) */
  int relu;
  int out_mode;  /* 0: bf16 NHWC through TMA; 1: fp32 row-major [m, out_ld] direct stores (1x1/Linear only) */
  int64_t out_ld;
  /* wgrad only: dW is fp32 [k_out, ntaps * tap_ld] with the tap slice at column tap_woff[t]; c_total = Cin */
  int k_out, c_total;
  int64_t w_ld;
  int tap_woff[DMPI_MAX_TAPS];
  int splits; /* wgrad: split-K factor over pixel tiles (0: auto).  igemm: > 1 = split-K over the taps x channel blocks
               * (needs dmpi_conv_plan_set_splitk_scratch; per-tap loads, cta_group 1, bf16 output) */
  int cta_group; /* igemm: 2 = CTA pairs (tcgen05.mma.cta_group::2, 256-row tiles; bn must be 128 or 256), else 1 */
  /* igemm patch mode (stride-1 R x S convolutions, halo = pad): the M tile is box_h output rows x box_w output columns;
   * ONE shared-memory patch of (box_h + 2*halo) x (box_w + 2*halo) input pixels per 64-channel block serves every
   * filter tap through row-shifted tcgen05 descriptors (tap (dh, dw) = shift of ((dh+halo)*(box_w+2*halo) + dw+halo)
   * pixel rows).  Needs (box_h-1)*(box_w+2*halo) + box_w <= 128, box_n == 1, one A view, cta_group 1. */
  int patch, halo;
} dmpi_conv_desc;

typedef struct dmpi_conv_plan dmpi_conv_plan;

/* Plans cache the encoded tensor maps; they are re-encoded only when a base pointer changes. */
dmpi_conv_plan* dmpi_conv_plan_create(const dmpi_conv_desc* desc);
void dmpi_conv_plan_destroy(dmpi_conv_plan* plan);

/* Deterministic BatchNorm statistics for dmpi_igemm_run*: zero-initialised device scratch of at least
 * (CTAs * 2 * bn + 64) * 4 bytes (512 KiB covers every configuration on 148 SMs) that no concurrently running launch
 * shares.  With it the per-channel sums are added in a fixed order (bit-reproducible); without it fp32 atomics. */
void dmpi_conv_plan_set_stat_scratch(dmpi_conv_plan* plan, void* scratch, long long bytes);

/* Split-K scratch of igemm plans with desc.splits > 1: 4 KiB of tickets (zero at first use, left zero by the kernel)
 * followed by splits * tiles * 128 * bn floats of partial tiles; not shared with a concurrently running launch. */
void dmpi_conv_plan_set_splitk_scratch(dmpi_conv_plan* plan, void* scratch, long long bytes);

/* D = A (*) B with fused bias / ReLU / BatchNorm statistics.  Returns 0 or a negative error. */
int dmpi_igemm_run(dmpi_conv_plan* plan, const void* a, const void* b, void* out, const float* bias,
                   float* stat_sum, float* stat_sqsum, void* stream);

/* Same, plus `addend` (bf16, element strides of its n / h / w axes, channel stride 1, may be NULL): a tensor on the
 * output's pixel grid that is added to the result before rounding — the residual-branch gradient in a dgrad. */
int dmpi_igemm_run_add(dmpi_conv_plan* plan, const void* a, const void* b, void* out, const float* bias,
                       float* stat_sum, float* stat_sqsum, const void* addend, long long add_sn, long long add_sh,
                       long long add_sw, void* stream);

/* dW += dY^T (*) X.  `dw` must already hold the value to accumulate onto (zero for a fresh gradient). */
int dmpi_wgrad_run(dmpi_conv_plan* plan, const void* dy, const void* x, float* dw, void* stream);

/* dW += ... for 3x3 stride-1 convolutions from one shared-memory input patch per pixel tile (taps = row-shifted
 * descriptors, two taps per M = 128 operand); desc as for dmpi_wgrad_run plus patch-mode box_w / box_h / halo. */
int dmpi_wgrad_patch_run(dmpi_conv_plan* plan, const void* dy, const void* x, float* dw, void* stream);

/* number of kernels the last call launched (for launch accounting) */
int dmpi_conv_last_error(char* buf, int buflen);

#ifdef __cplusplus
}
#endif
#endif
