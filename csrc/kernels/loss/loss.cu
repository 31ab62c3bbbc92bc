// Fused loss kernels: forward value and input gradient in one pass.
//   * softmax cross-entropy with mean reduction  (nn.CrossEntropyLoss, ref:pytorch/resnet/main.py:113)
//   * binary cross-entropy with logits, mean     (nn.BCEWithLogitsLoss, ref:pytorch/unet/train.py:162)
// plus the 1x1 "segmentation head" (Conv2d(64, out_classes, 1), ref:pytorch/unet/model.py:68) whose
// out_classes is 1..8: far too narrow for a tensor-core tile, so it is a bandwidth-bound dot product.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

// One warp per row.  logits fp32 [M, ld]; labels int64; dlogits fp32 [M, ldd] (columns N..ldd_cols zeroed);
// loss_sum accumulates sum_i (-log softmax_i[label_i]) / M.
__global__ void __launch_bounds__(256) xent_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                   float* __restrict__ dlogits, float* __restrict__ loss_sum,
                                                   int M, int N, long long ld, long long ldd, int ldd_cols,
                                                   float grad_scale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= M) return;
  const float* row = logits + static_cast<long long>(warp) * ld;
  float mx = -INFINITY;
  for (int j = lane; j < N; j += 32) mx = fmaxf(mx, row[j]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float se = 0.f;
  for (int j = lane; j < N; j += 32) se += __expf(row[j] - mx);
  se = warp_sum(se);
  const float lse = mx + __logf(se);
  const int label = static_cast<int>(labels[warp]);
  const float invM = 1.f / static_cast<float>(M);
  if (lane == 0) atomicAdd(loss_sum, (lse - row[label]) * invM);
  float* drow = dlogits + static_cast<long long>(warp) * ldd;
  const float gs = grad_scale * invM;
  for (int j = lane; j < ldd_cols; j += 32) {
    float g = 0.f;
    if (j < N) g = (__expf(row[j] - lse) - (j == label ? 1.f : 0.f)) * gs;
    drow[j] = g;
  }
}

// logits fp32 [M], targets fp32 [M]; dlogits fp32 [M]; numerically stable form max(x,0) - x*t + log1p(exp(-|x|)).
__global__ void __launch_bounds__(256) bce_kernel(const float* __restrict__ x, const float* __restrict__ t,
                                                  float* __restrict__ dx, float* __restrict__ loss_sum, long long M,
                                                  float grad_scale) {
  const float invM = 1.f / static_cast<float>(M);
  float local = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < M;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float xv = x[i], tv = t[i];
    local += fmaxf(xv, 0.f) - xv * tv + log1pf(__expf(-fabsf(xv)));
    const float sg = 1.f / (1.f + __expf(-xv));
    dx[i] = (sg - tv) * invM * grad_scale;
  }
  local = warp_sum(local);
  __shared__ float ws[8];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x < 8) {
    float v = ws[threadIdx.x];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
    if (threadIdx.x == 0) atomicAdd(loss_sum, v * invM);
  }
}

// ---- narrow 1x1 head: y[p, k] = sum_c x[p, c] * w[k, c] + b[k],  k < K <= 8, C % 8 == 0, C <= 512 ----
constexpr int kHeadMaxK = 8;

// KT: compile-time bound on K (1 for the UNet's single-class head).  Four pixels per thread in flight.
template <int KT>
__global__ void __launch_bounds__(256) head_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                       const float* __restrict__ b, float* __restrict__ y, long long M,
                                                       int C, int K, long long ldx) {
  constexpr int U = KT == 1 ? 4 : 2;
  extern __shared__ float sw[];  // [K][C]
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  // 8 threads cooperate on one pixel (each takes C/8 channels in 8-wide vectors), shuffle-reduce.
  const int sub = threadIdx.x & 7;
  const int lane = threadIdx.x & 31;
  const long long warp0 = ((blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5) * 4;
  const long long wstride = (static_cast<long long>(gridDim.x) * blockDim.x) >> 3;  // pixels per grid step
  for (long long p0 = warp0; p0 < M; p0 += U * wstride) {
    // warp-uniform loop (the shuffles below need every lane); lanes past the end clamp to the last pixel
    long long pix[U];
    bool pvalid[U];
    float acc[U][KT];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long pix_raw = p0 + u * wstride + (lane >> 3);
      pvalid[u] = pix_raw < M;
      pix[u] = pvalid[u] ? pix_raw : M - 1;
#pragma unroll
      for (int k = 0; k < KT; ++k) acc[u][k] = 0.f;
    }
    for (int c0 = sub * 8; c0 < C; c0 += 64) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = *reinterpret_cast<const uint4*>(x + pix[u] * ldx + c0);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t vi[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
        float xv[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          xv[2 * j] = bf16lo(vi[j]);
          xv[2 * j + 1] = bf16hi(vi[j]);
        }
#pragma unroll
        for (int k = 0; k < KT; ++k) {
          if (k < K) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[u][k] = fmaf(xv[j], sw[k * C + c0 + j], acc[u][k]);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        if (k < K) {
          float v = acc[u][k];
          v += __shfl_xor_sync(0xffffffffu, v, 1);
          v += __shfl_xor_sync(0xffffffffu, v, 2);
          v += __shfl_xor_sync(0xffffffffu, v, 4);
          if (sub == 0 && pvalid[u]) y[pix[u] * K + k] = v + (b ? b[k] : 0.f);
        }
      }
  }
}

// dx[p, c] = sum_k dy[p, k] * w[k, c]  (bf16 out);  dw[k, c] += sum_p dy[p,k] x[p,c];  db[k] += sum_p dy[p,k]
// KT: compile-time bound on K (1 for the UNet's single-class head: the weights of the thread's 8 channels then live in
// registers and the loop is one load, 16 FMAs and one store per pixel; kHeadMaxK otherwise).  Four pixels per
// iteration with every load issued up front: the kernel is a pure stream (1.07 GB for the 512x512 UNet) and one or
// two 16-byte loads per thread left the memory system idle (1.9 TB/s).
template <int KT>
__global__ void __launch_bounds__(256) head_bwd_kernel(const float* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                       const float* __restrict__ w, __nv_bfloat16* __restrict__ dx,
                                                       float* __restrict__ dw, float* __restrict__ db, long long M,
                                                       int C, int K, long long ldx, long long ld_dx) {
  constexpr int U = KT == 1 ? 4 : 2;
  extern __shared__ float sm[];  // w[K][C], then accumulators dw[K][C] + db[K]
  float* sw = sm;
  float* sdw = sm + K * C;
  float* sdb = sdw + K * C;
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) {
    sw[i] = w[i];
    sdw[i] = 0.f;
  }
  if (threadIdx.x < K) sdb[threadIdx.x] = 0.f;
  __syncthreads();
  const int vecC = C >> 3;
  const int cv = threadIdx.x % vecC;  // host picks blockDim multiple of vecC
  const int slab = threadIdx.x / vecC, nslab = blockDim.x / vecC;
  float adw[KT][8];
  float adb[KT];
  float wr[KT == 1 ? 8 : 1];  // K = 1: this thread's weights
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    adb[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) adw[k][j] = 0.f;
  }
  if (KT == 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[j] = sw[cv * 8 + j];
  }
  const long long stride = static_cast<long long>(gridDim.x) * nslab;
  for (long long pix0 = blockIdx.x * static_cast<long long>(nslab) + slab; pix0 < M; pix0 += U * stride) {
    uint4 vv[U];
    float dd[U][KT];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long pix = pix0 + u * stride;
      if (pix < M) {
        vv[u] = *reinterpret_cast<const uint4*>(x + pix * ldx + cv * 8);
#pragma unroll
        for (int k = 0; k < KT; ++k) dd[u][k] = k < K ? dy[pix * K + k] : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long pix = pix0 + u * stride;
      if (pix >= M) break;
      const uint32_t vi[4] = {vv[u].x, vv[u].y, vv[u].z, vv[u].w};
      float xv[8], g[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        xv[2 * j] = bf16lo(vi[j]);
        xv[2 * j + 1] = bf16hi(vi[j]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        if (k < K) {
          const float d = dd[u][k];
          if (cv == 0) adb[k] += d;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            g[j] = fmaf(d, KT == 1 ? wr[KT == 1 ? j : 0] : sw[k * C + cv * 8 + j], g[j]);
            adw[k][j] = fmaf(d, xv[j], adw[k][j]);
          }
        }
      }
      uint4 o;
      o.x = pack_bf16x2(g[0], g[1]);
      o.y = pack_bf16x2(g[2], g[3]);
      o.z = pack_bf16x2(g[4], g[5]);
      o.w = pack_bf16x2(g[6], g[7]);
      *reinterpret_cast<uint4*>(dx + pix * ld_dx + cv * 8) = o;
    }
  }
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    if (k < K) {
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&sdw[k * C + cv * 8 + j], adw[k][j]);
      if (cv == 0) atomicAdd(&sdb[k], adb[k]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) atomicAdd(dw + i, sdw[i]);
  if (threadIdx.x < K && db) atomicAdd(db + threadIdx.x, sdb[threadIdx.x]);
}

}  // namespace dmpi

using namespace dmpi;

namespace dmpi {

// dst[i] = src[i] * (*scale): the incoming gradient of a scalar loss (a device scalar, usually 1) applied to the
// gradient the fused forward pass already produced, without a host round trip.
__global__ void __launch_bounds__(256) scale_dev_kernel(const float4* __restrict__ src, float4* __restrict__ dst,
                                                        long long nvec, const float* __restrict__ src_tail,
                                                        float* __restrict__ dst_tail, int ntail,
                                                        const float* __restrict__ scale) {
  const float s = __ldg(scale);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec; i += stride) {
    float4 v = src[i];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    dst[i] = v;
  }
  if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < ntail) dst_tail[threadIdx.x] = src_tail[threadIdx.x] * s;
}

// Backward-pass preparation of the fully connected layer (nn.Linear of ref:pytorch/resnet/main.py:41): one pass over
// the fp32 logit gradient dy [M, N] that (a) writes the bf16 copy zero-padded to Npad columns which the tcgen05
// weight-/data-gradient GEMMs consume and (b) adds the column sums (= bias gradient) into dbias.  One block owns 32
// columns, so every dbias element has a single writer and a fixed summation order.
__global__ void __launch_bounds__(256) linear_bwd_prep_kernel(const float* __restrict__ dy, long long ld, int M, int N,
                                                              __nv_bfloat16* __restrict__ out, int Npad,
                                                              float* __restrict__ dbias) {
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rg = threadIdx.x >> 5;  // 8 row groups
  float acc = 0.f;
  if (col < Npad) {
    for (int r = rg; r < M; r += 8) {
      const float v = col < N ? dy[static_cast<long long>(r) * ld + col] : 0.f;
      out[static_cast<long long>(r) * Npad + col] = __float2bfloat16(v);
      acc += v;
    }
  }
  __shared__ float red[8][32];
  red[rg][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rg == 0 && col < N && dbias != nullptr) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    dbias[col] += t;
  }
}

}  // namespace dmpi

extern "C" int dmpi_scale_dev(const float* src, float* dst, long long n, const float* scale_dev, void* stream) {
  if (n <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) return -1;
  const long long nvec = n / 4;
  const int ntail = static_cast<int>(n - nvec * 4);
  long long blocks = (nvec + 256 * 4 - 1) / (256 * 4);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  dmpi::scale_dev_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(src), reinterpret_cast<float4*>(dst), nvec, src + nvec * 4, dst + nvec * 4, ntail,
      scale_dev);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_linear_bwd_prep(const float* dy, long long ld, int M, int N, void* dy_bf16, int Npad, float* dbias,
                                    void* stream) {
  if (M <= 0 || N <= 0 || Npad < N) return -1;
  dmpi::linear_bwd_prep_kernel<<<(Npad + 31) / 32, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dy, ld, M, N, static_cast<__nv_bfloat16*>(dy_bf16), Npad, dbias);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_softmax_xent(const float* logits, const long long* labels, void* dlogits, float* loss_sum, int M,
                                 int N, long long ld, long long ldd, int ldd_cols, float grad_scale, void* stream) {
  const int warps_per_block = 8;
  const int grid = (M + warps_per_block - 1) / warps_per_block;
  xent_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(logits, labels, static_cast<float*>(dlogits),
                                                                   loss_sum, M, N, ld, ldd, ldd_cols, grad_scale);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_bce_logits(const float* x, const float* t, float* dx, float* loss_sum, long long M,
                               float grad_scale, void* stream) {
  long long blocks = (M + 256 * 4 - 1) / (256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  bce_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, t, dx, loss_sum, M, grad_scale);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_head_forward(const void* x, const float* w, const float* b, float* y, long long M, int C, int K,
                                 long long ldx, void* stream) {
  if (K < 1 || K > kHeadMaxK || C % 64 != 0) return -1;
  long long blocks = (M * 8 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  auto kern = K == 1 ? head_fwd_kernel<1> : head_fwd_kernel<kHeadMaxK>;
  kern<<<static_cast<int>(blocks), 256, K * C * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), w, b, y, M, C, K, ldx);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_head_backward(const float* dy, const void* x, const float* w, void* dx, float* dw, float* db,
                                  long long M, int C, int K, long long ldx, long long ld_dx, void* stream) {
  if (K < 1 || K > kHeadMaxK || C % 8 != 0 || C / 8 > 256) return -1;
  const int vecC = C / 8;
  const int threads = (256 / vecC) * vecC;
  const int nslab = threads / vecC;
  long long blocks = (M + nslab * 16 - 1) / (nslab * 16);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  const size_t smem = (2 * K * C + K) * sizeof(float);
  auto kern = K == 1 ? head_bwd_kernel<1> : head_bwd_kernel<kHeadMaxK>;
  kern<<<static_cast<int>(blocks), threads, smem, static_cast<cudaStream_t>(stream)>>>(
      dy, static_cast<const __nv_bfloat16*>(x), w, static_cast<__nv_bfloat16*>(dx), dw, db, M, C, K, ldx, ld_dx);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
