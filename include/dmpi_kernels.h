/* C ABI of the non-GEMM sm_100a kernels (csrc/kernels/{norm,pool,loss,optim,misc}).
 *
 * Conventions: activations are NHWC bf16 with a uniform pixel stride `ld*` (elements), per-channel vectors are fp32,
 * `stream` is a cudaStream_t passed as void*, every function returns 0 on success or a negative code, and nothing
 * here allocates: outputs and scratch are supplied by the caller.
 */
#ifndef DMPI_KERNELS_H_
#define DMPI_KERNELS_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- BatchNorm2d (+residual) (+ReLU)   (ref:pytorch/unet/model.py:10-11,13-14; torchvision resnet bn/relu) */
/* out = relu?( (y - mean) * invstd * gamma + beta [+ res] ).  training != 0: mean/var come from sum/sqsum over the M
 * rows (as produced by the igemm epilogue or dmpi_bn_stats), save_mean/save_invstd are written, running stats and
 * num_batches_tracked are updated like torch (momentum EMA, unbiased running_var).  training == 0: running stats. */
int dmpi_bn_forward(const void* y, const void* res, void* out, long long M, int C, long long ldy, long long ldr,
                    long long ldo, const float* sum, const float* sqsum, const float* gamma, const float* beta,
                    float* save_mean, float* save_invstd, float* running_mean, float* running_var,
                    long long* num_batches_tracked, float eps, float momentum, int training, int relu, void* stream);
/* dy = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)), dz = dout * (act > 0) (act == NULL: no ReLU);
 * dres (optional) = dz; dgamma += sum(dz*xhat); dbeta += sum(dz).
 * `acc` is caller-provided scratch of 2*C floats that MUST be zero on entry (it is left dirty). */
int dmpi_bn_backward(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M, int C,
                     long long ld_dout, long long ldy, long long ld_act, long long ld_dy, long long ld_dres,
                     const float* mean, const float* invstd, const float* gamma, float* acc, float* dgamma,
                     float* dbeta, void* stream);
/* Extended form: relu_from_y != 0 with act == NULL recomputes the ReLU gate as (y*gamma*invstd + beta - mean*gamma*invstd > 0)
 * - exact for BatchNorm+ReLU without a residual, and one tensor read less per pass. */
int dmpi_bn_backward_ex(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M, int C,
                        long long ld_dout, long long ldy, long long ld_act, long long ld_dy, long long ld_dres,
                        const float* mean, const float* invstd, const float* gamma, const float* beta, int relu_from_y,
                        float* acc, float* dgamma, float* dbeta, void* stream);
/* ReLU bit mask variants: the forward pass also writes one bit per output element (set where the stored value is > 0,
 * [M, C/8] bytes) and the backward pass takes its ReLU gate from those bits instead of re-reading the activation. */
int dmpi_bn_forward_mask(const void* y, const void* res, void* out, long long M, int C, long long ldy, long long ldr,
                         long long ldo, const float* sum, const float* sqsum, const float* gamma, const float* beta,
                         float* save_mean, float* save_invstd, float* running_mean, float* running_var,
                         long long* num_batches_tracked, float eps, float momentum, int training, int relu,
                         void* relu_mask, void* stream);
int dmpi_bn_backward_mask(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M, int C,
                          long long ld_dout, long long ldy, long long ld_act, long long ld_dy, long long ld_dres,
                          const float* mean, const float* invstd, const float* gamma, const float* beta,
                          int relu_from_y, const void* relu_mask, float* acc, float* dgamma, float* dbeta,
                          void* stream);
/* Same with deterministic channel reductions: `det_scratch` is zero-initialised device memory (1 MiB covers every
 * shape) that no concurrently running launch shares; the partial sums of the thread blocks are added in a fixed order
 * instead of with fp32 atomics, and `acc` need not be zeroed. */
int dmpi_bn_backward_det(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M, int C,
                         long long ld_dout, long long ldy, long long ld_act, long long ld_dy, long long ld_dres,
                         const float* mean, const float* invstd, const float* gamma, const float* beta,
                         int relu_from_y, const void* relu_mask, float* acc, float* dgamma, float* dbeta,
                         void* det_scratch, long long det_bytes, void* stream);
/* Same computation for callers that cannot allocate the scratch (GPU memory fully owned by the host application's
 * pool, latency-critical paths where cudaMalloc/cudaFree would stall the device, plug-ins barred from any allocator).
 * This entry point has no counterpart in the reference; it exists for such embedders only.
 * The 2*C floats live in a 16 KiB __device__ array inside libdmpi_cuda.so (held by every GPU that loads the library
 * until the process exits), hence the limit C <= 2048: larger C returns -4 with a message on stderr.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress): make it during set-up, with that
 * device current, before any stream capture.  Each call clears the floats it uses and leaves them dirty.  All calls
 * on a GPU share the array, so calls whose kernels may overlap on the GPU race: keep them stream-ordered.
 * Allocates nothing (no cudaMalloc / cudaFree / library workspace).  Codes: 0, -1 bad C, -2 launch, -4 too large,
 * -5 / -6 device or symbol lookup failed. */
int dmpi_bn_backward_noalloc(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M, int C,
                             long long ld_dout, long long ldy, long long ld_act, long long ld_dy, long long ld_dres,
                             const float* mean, const float* invstd, const float* gamma, float* dgamma, float* dbeta,
                             void* stream);
/* sum[c] += sum_rows y, sqsum[c] += sum_rows y*y (sqsum may be NULL: channel sums only, e.g. a bias gradient) */
int dmpi_bn_stats(const void* y, long long M, int C, long long ld, float* sum, float* sqsum, void* stream);

/* ---- pooling (ref:pytorch/unet/model.py:25; torchvision resnet maxpool/avgpool) */
int dmpi_maxpool_forward(const void* x, void* y, void* idx /* uint8 arg-max tap or NULL */, int N, int H, int W, int C,
                         int OH, int OW, int k, int s, int p, long long xsn, long long xsh, long long xsw,
                         long long ysn, long long ysh, long long ysw, void* stream);
int dmpi_maxpool_backward(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int OH, int OW, int k,
                          int s, int p, long long dsn, long long dsh, long long dsw, long long xsn, long long xsh,
                          long long xsw, void* stream);
/* Same, plus an optional second gradient of the pooled tensor (bf16, element strides of n / h / w, channel stride 1)
 * added in the same pass: the UNet skip-connection gradient, so no separate add kernel runs. */
int dmpi_maxpool_backward_add(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int OH, int OW,
                              int k, int s, int p, long long dsn, long long dsh, long long dsw, long long xsn,
                              long long xsh, long long xsw, const void* addend, long long asn, long long ash,
                              long long asw, void* stream);
int dmpi_avgpool_forward(const void* x, void* y, int N, int HW, int C, void* stream);
int dmpi_avgpool_backward(const void* dy, void* dx, int N, int HW, int C, void* stream);

/* ---- bilinear 2x up-sampling, align_corners = True (nn.Upsample in ref:pytorch/unet/model.py:39-40).
 * x / dx: NHWC bf16 [N, H, W, C]; y / dy: [N, 2H, 2W, C]; element strides of n / h / w (channel stride 1, C % 8 == 0,
 * strides multiples of 8), so y may be a channel slice of the decoder's concat buffer.  Backward is a gather (no
 * atomics, deterministic). */
int dmpi_upsample2x_bilinear_forward(const void* x, void* y, int N, int H, int W, int C, long long xsn, long long xsh,
                                     long long xsw, long long ysn, long long ysh, long long ysw, void* stream);
int dmpi_upsample2x_bilinear_backward(const void* dy, void* dx, int N, int H, int W, int C, long long dsn, long long dsh,
                                      long long dsw, long long xsn, long long xsh, long long xsw, void* stream);

/* ---- losses, forward value + input gradient in one pass (ref:pytorch/resnet/main.py:113; ref:pytorch/unet/train.py:162) */
int dmpi_softmax_xent(const float* logits, const long long* labels, void* dlogits_f32, float* loss_sum, int M, int N,
                      long long ld, long long ldd, int ldd_cols, float grad_scale, void* stream);
int dmpi_bce_logits(const float* x, const float* t, float* dx, float* loss_sum, long long M, float grad_scale,
                    void* stream);
/* dst = src * (*scale_dev): applies the incoming gradient of the scalar loss (a device scalar) to the stored dlogits. */
int dmpi_scale_dev(const float* src, float* dst, long long n, const float* scale_dev, void* stream);
/* Backward-pass preparation of nn.Linear (ref:pytorch/resnet/main.py:41): dy fp32 [M, N] (row stride ld) -> bf16 copy
 * zero-padded to Npad columns (operand of the tcgen05 wgrad / dgrad GEMMs) and dbias[c] += sum_rows dy (NULL: skip). */
int dmpi_linear_bwd_prep(const float* dy, long long ld, int M, int N, void* dy_bf16, int Npad, float* dbias,
                         void* stream);
/* narrow 1x1 conv head, K <= 8 output classes (ref:pytorch/unet/model.py:68) */
int dmpi_head_forward(const void* x, const float* w, const float* b, float* y, long long M, int C, int K, long long ldx,
                      void* stream);
int dmpi_head_backward(const float* dy, const void* x, const float* w, void* dx, float* dw, float* db, long long M,
                       int C, int K, long long ldx, long long ld_dx, void* stream);

/* ---- flat-arena optimizers (ref:pytorch/resnet/main.py:114; ref:pytorch/unet/train.py:160-161,194) */
int dmpi_sgd_step(float* p, const float* g, float* mom, void* bf16_shadow, long long n, float lr, float momentum,
                  float dampening, float wd, int first_step, const float* clip_sumsq, float max_norm, void* stream);
int dmpi_adam_step(float* p, const float* g, float* m, float* v, void* bf16_shadow, long long n, float lr, float beta1,
                   float beta2, float eps, float wd, float bias_corr1, float bias_corr2, const float* clip_sumsq,
                   float max_norm, void* stream);
/* Adam with the step count t read from device memory (bias corrections computed in-kernel): CUDA-graph safe. */
int dmpi_adam_step_dev(float* p, const float* g, float* m, float* v, void* bf16_shadow, long long n, float lr,
                       float beta1, float beta2, float eps, float wd, const float* step_dev, const float* clip_sumsq,
                       float max_norm, void* stream);
int dmpi_sumsq(const float* g, long long n, float* out, void* stream);
/* one-element device bookkeeping: op 0: *dst = *src + arg (Adam step counter); op 1: *dst = sqrt(*src) (gradient norm) */
int dmpi_scalar_op(float* dst, const float* src, int op, float arg, void* stream);
int dmpi_scale_by_clip(float* g, long long n, const float* sumsq, float max_norm, void* stream);
int dmpi_cast_bf16(const float* src, void* dst, long long n, void* stream);
int dmpi_cast_f32(const void* src_bf16, float* dst, long long n, void* stream); /* n % 4 == 0 */

/* ---- layout / staging (ref:pytorch/resnet/main.py:126 input .to(device)) */
/* Fills [p, p + bytes) on `stream`; value 0 runs a store kernel (a memset NODE in a replayed CUDA graph delays the next
 * kernel by ~16 us), other values go to cudaMemsetAsync. */
int dmpi_memset_async(void* p, int value, long long bytes, void* stream);
int dmpi_nchw_to_nhwc8_padded(const float* x, void* out, int N, int C, int H, int W, int pad, int Hp, int Wp,
                              void* stream);
int dmpi_nchw_to_nhwc(const float* x, void* out, int N, int C, int HW, void* stream);
int dmpi_nhwc_to_nchw(const void* x, float* out, int N, int C, int HW, long long ldx, void* stream);
int dmpi_transpose_table(const void* table, int n_entries, int total_tiles, void* stream);
int dmpi_stem_pack(const float* w, void* out, int K, int R, int S, int C, void* stream);
int dmpi_stem_unpack_grad(const float* dwin, float* dw, int K, int R, int S, int C, void* stream);

#ifdef __cplusplus
}
#endif
#endif
