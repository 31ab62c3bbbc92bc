// Fused flat-arena optimizers.
//
// Parameters, gradients and optimizer state live in contiguous fp32 arenas (the gradient arena IS the
// DDP bucket memory), so one launch updates the whole model:
//   * SGD with momentum / dampening / weight decay  == torch.optim.SGD   (ref:pytorch/resnet/main.py:114)
//   * Adam (bias-corrected, eps outside sqrt)        == torch.optim.Adam  (ref:pytorch/unet/train.py:160-161)
// Both optionally apply the clip_grad_norm_ coefficient computed on the device
// (ref:pytorch/unet/train.py:194) and write the bf16 mirror of the parameters that the tcgen05
// kernels consume, so no separate cast pass exists.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

__device__ __forceinline__ float clip_coef(const float* sumsq, float max_norm) {
  if (!sumsq) return 1.f;
  const float norm = sqrtf(*sumsq);
  const float c = max_norm / (norm + 1e-6f);
  return c < 1.f ? c : 1.f;
}

__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                  float* __restrict__ mom, __nv_bfloat16* __restrict__ shadow,
                                                  long long n, float lr, float momentum, float dampening, float wd,
                                                  int first_step, const float* sumsq, float max_norm) {
  const float coef = clip_coef(sumsq, max_norm);
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float pp[4] = {pv.x, pv.y, pv.z, pv.w};
    const float gg[4] = {gv.x, gv.y, gv.z, gv.w};
    float mm[4] = {0.f, 0.f, 0.f, 0.f};
    if (mom && !first_step) {
      const float4 mv = reinterpret_cast<float4*>(mom)[i];
      mm[0] = mv.x; mm[1] = mv.y; mm[2] = mv.z; mm[3] = mv.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float d = gg[j] * coef + wd * pp[j];
      if (mom) {
        mm[j] = first_step ? d : momentum * mm[j] + (1.f - dampening) * d;
        d = mm[j];
      }
      pp[j] -= lr * d;
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    if (mom) reinterpret_cast<float4*>(mom)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    if (shadow) {
      uint2 o;
      o.x = pack_bf16x2(pp[0], pp[1]);
      o.y = pack_bf16x2(pp[2], pp[3]);
      reinterpret_cast<uint2*>(shadow)[i] = o;
    }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v,
                                                   __nv_bfloat16* __restrict__ shadow, long long n, float lr,
                                                   float beta1, float beta2, float eps, float wd, float bc1, float bc2,
                                                   const float* sumsq, float max_norm, const float* step_dev) {
  if (step_dev) {  // step count kept on the device (CUDA-graph replay): bias corrections computed here
    const float t = *step_dev;
    bc1 = 1.f - powf(beta1, t);
    bc2 = 1.f - powf(beta2, t);
  }
  const float coef = clip_coef(sumsq, max_norm);
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pp[4] = {pv.x, pv.y, pv.z, pv.w};
    const float gg[4] = {gv.x, gv.y, gv.z, gv.w};
    float mm[4] = {mv.x, mv.y, mv.z, mv.w};
    float v2[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d = gg[j] * coef + wd * pp[j];
      mm[j] = beta1 * mm[j] + (1.f - beta1) * d;
      v2[j] = beta2 * v2[j] + (1.f - beta2) * d * d;
      const float denom = sqrtf(v2[j]) * inv_sqrt_bc2 + eps;
      pp[j] -= step_size * (mm[j] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(v2[0], v2[1], v2[2], v2[3]);
    if (shadow) {
      uint2 o;
      o.x = pack_bf16x2(pp[0], pp[1]);
      o.y = pack_bf16x2(pp[2], pp[3]);
      reinterpret_cast<uint2*>(shadow)[i] = o;
    }
  }
}

// out[0] += sum(g^2) (out must be zeroed by the caller when a fresh norm is wanted)
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, long long n, float* out) {
  float acc = 0.f;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(g)[i];
    acc = fmaf(v.x, v.x, acc);
    acc = fmaf(v.y, v.y, acc);
    acc = fmaf(v.z, v.z, acc);
    acc = fmaf(v.w, v.w, acc);
  }
  acc = warp_sum(acc);
  __shared__ float ws[8];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 8) {
    float v = ws[threadIdx.x];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
    if (threadIdx.x == 0) atomicAdd(out, v);
  }
}

// g *= min(1, max_norm / (sqrt(sumsq) + 1e-6))   — the in-place form of clip_grad_norm_
__global__ void __launch_bounds__(256) scale_by_clip_kernel(float* __restrict__ g, long long n, const float* sumsq,
                                                            float max_norm) {
  const float coef = clip_coef(sumsq, max_norm);
  if (coef >= 1.f) return;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 v = reinterpret_cast<float4*>(g)[i];
    v.x *= coef; v.y *= coef; v.z *= coef; v.w *= coef;
    reinterpret_cast<float4*>(g)[i] = v;
  }
}

__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                        long long n) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    reinterpret_cast<uint2*>(dst)[i] = o;
  }
}

__global__ void __launch_bounds__(256) cast_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                                                       long long n) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint2 v = reinterpret_cast<const uint2*>(src)[i];
    reinterpret_cast<float4*>(dst)[i] = make_float4(bf16lo(v.x), bf16hi(v.x), bf16lo(v.y), bf16hi(v.y));
  }
}

// One-element bookkeeping on the device (Adam's step counter, the norm returned by clip_grad_norm_): keeps the training
// step free of library kernels and graph-capturable.  op 0: dst = src + arg;  op 1: dst = sqrt(src).
__global__ void scalar_op_kernel(float* dst, const float* src, int op, float arg) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *dst = op == 0 ? *src + arg : sqrtf(*src);
}

static int flat_grid(long long n) {
  long long b = (n / 4 + 255) / 256;
  if (b > 148LL * 8) b = 148LL * 8;
  return b < 1 ? 1 : static_cast<int>(b);
}

}  // namespace dmpi

using namespace dmpi;

// All arenas must be 16-byte aligned and n a multiple of 4 (the arena allocator pads every tensor).
extern "C" int dmpi_sgd_step(float* p, const float* g, float* mom, void* shadow, long long n, float lr, float momentum,
                             float dampening, float wd, int first_step, const float* sumsq, float max_norm,
                             void* stream) {
  if (n % 4) return -1;
  sgd_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, mom, static_cast<__nv_bfloat16*>(shadow), n, lr, momentum, dampening, wd, first_step, sumsq, max_norm);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_adam_step(float* p, const float* g, float* m, float* v, void* shadow, long long n, float lr,
                              float beta1, float beta2, float eps, float wd, float bias_corr1, float bias_corr2,
                              const float* sumsq, float max_norm, void* stream) {
  if (n % 4) return -1;
  adam_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, static_cast<__nv_bfloat16*>(shadow), n, lr, beta1, beta2, eps, wd, bias_corr1, bias_corr2, sumsq,
      max_norm, nullptr);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

// Same update with the step count t read from device memory (bias corrections 1 - beta^t computed in the kernel), so
// that a captured CUDA graph of the training step stays correct across replays.
extern "C" int dmpi_adam_step_dev(float* p, const float* g, float* m, float* v, void* shadow, long long n, float lr,
                                  float beta1, float beta2, float eps, float wd, const float* step_dev,
                                  const float* sumsq, float max_norm, void* stream) {
  if (n % 4 || !step_dev) return -1;
  adam_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, static_cast<__nv_bfloat16*>(shadow), n, lr, beta1, beta2, eps, wd, 1.f, 1.f, sumsq, max_norm, step_dev);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_sumsq(const float* g, long long n, float* out, void* stream) {
  if (n % 4) return -1;
  sumsq_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(g, n, out);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_scalar_op(float* dst, const float* src, int op, float arg, void* stream) {
  scalar_op_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(dst, src, op, arg);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_scale_by_clip(float* g, long long n, const float* sumsq, float max_norm, void* stream) {
  if (n % 4) return -1;
  scale_by_clip_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(g, n, sumsq, max_norm);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_cast_f32(const void* src, float* dst, long long n, void* stream) {
  if (n % 4) return -1;
  cast_f32_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(src), dst, n);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_cast_bf16(const float* src, void* dst, long long n, void* stream) {
  if (n % 4) return -1;
  cast_bf16_kernel<<<flat_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, static_cast<__nv_bfloat16*>(dst), n);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
