// Bilinear 2x up-sampling, align_corners = True, on NHWC bf16 — the `up_sample_mode="bilinear"` branch of the UNet
// decoder (nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True), ref:pytorch/unet/model.py:39-40).
//
// Forward : one thread per (output pixel, 8-channel chunk): four 16-byte loads, fp32 lerp, one 16-byte store.  The
//           output may be a channel slice of a wider buffer (the decoder's concat buffer), so it has its own strides.
// Backward: gather form, one thread per (INPUT pixel, 8-channel chunk) summing the few output pixels that touched it
//           (the source coordinate of output row i is i * (H-1)/(2H-1), so an input row receives from at most five
//           output rows): no atomics, a fixed summation order, bit-reproducible.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = bf16lo(w[j]);
    f[2 * j + 1] = bf16hi(w[j]);
  }
}

__global__ void __launch_bounds__(256) upsample2x_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                             int N, int H, int W, int C8, long long xsn, long long xsh,
                                                             long long xsw, long long ysn, long long ysh, long long ysw,
                                                             float rh, float rw) {
  const int OH = 2 * H, OW = 2 * W;
  const long long total = static_cast<long long>(N) * OH * OW * C8;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(t % C8);
    long long r = t / C8;
    const int ox = static_cast<int>(r % OW);
    r /= OW;
    const int oy = static_cast<int>(r % OH);
    const int n = static_cast<int>(r / OH);
    const float sy = oy * rh, sx = ox * rw;
    const int y0 = min(static_cast<int>(sy), H - 1), x0 = min(static_cast<int>(sx), W - 1);
    const int y1 = min(y0 + 1, H - 1), x1 = min(x0 + 1, W - 1);
    const float wy = sy - y0, wx = sx - x0;
    const __nv_bfloat16* base = x + n * xsn + c8 * 8;
    float a[8], b[8], c[8], d[8];
    unpack8(*reinterpret_cast<const uint4*>(base + y0 * xsh + x0 * xsw), a);
    unpack8(*reinterpret_cast<const uint4*>(base + y0 * xsh + x1 * xsw), b);
    unpack8(*reinterpret_cast<const uint4*>(base + y1 * xsh + x0 * xsw), c);
    unpack8(*reinterpret_cast<const uint4*>(base + y1 * xsh + x1 * xsw), d);
    const float w00 = (1.f - wy) * (1.f - wx), w01 = (1.f - wy) * wx, w10 = wy * (1.f - wx), w11 = wy * wx;
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float v0 = w00 * a[2 * j] + w01 * b[2 * j] + w10 * c[2 * j] + w11 * d[2 * j];
      const float v1 = w00 * a[2 * j + 1] + w01 * b[2 * j + 1] + w10 * c[2 * j + 1] + w11 * d[2 * j + 1];
      o[j] = pack_bf16x2(v0, v1);
    }
    *reinterpret_cast<uint4*>(y + n * ysn + oy * ysh + ox * ysw + c8 * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// weight with which output index `o` (source coordinate o * r) touches input index `i` of an axis of length L
__device__ __forceinline__ float tap_weight(int o, int i, int L, float r) {
  const float s = o * r;
  const int i0 = min(static_cast<int>(s), L - 1);
  const int i1 = min(i0 + 1, L - 1);
  const float w1 = s - i0;
  float w = 0.f;
  if (i0 == i) w += 1.f - w1;
  if (i1 == i) w += w1;
  return w;
}

__global__ void __launch_bounds__(256) upsample2x_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                                             int N, int H, int W, int C8, long long dsn, long long dsh,
                                                             long long dsw, long long xsn, long long xsh, long long xsw,
                                                             float rh, float rw) {
  const int OH = 2 * H, OW = 2 * W;
  const long long total = static_cast<long long>(N) * H * W * C8;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(t % C8);
    long long r = t / C8;
    const int ix = static_cast<int>(r % W);
    r /= W;
    const int iy = static_cast<int>(r % H);
    const int n = static_cast<int>(r / H);
    // output rows / columns whose two source taps can include (iy, ix): source coordinate in (i - 1, i + 1)
    const int oy_lo = max(0, 2 * iy - 3), oy_hi = min(OH - 1, 2 * iy + 3);
    const int ox_lo = max(0, 2 * ix - 3), ox_hi = min(OW - 1, 2 * ix + 3);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const __nv_bfloat16* base = dy + n * dsn + c8 * 8;
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      const float wy = tap_weight(oy, iy, H, rh);
      if (wy == 0.f) continue;
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        const float w = wy * tap_weight(ox, ix, W, rw);
        if (w == 0.f) continue;
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(base + oy * dsh + ox * dsw), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(w, g[j], acc[j]);
      }
    }
    *reinterpret_cast<uint4*>(dx + n * xsn + iy * xsh + ix * xsw + c8 * 8) =
        make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]),
                   pack_bf16x2(acc[6], acc[7]));
  }
}

static int grid_for(long long total) {
  long long b = (total + 255) / 256;
  if (b > 148LL * 16) b = 148LL * 16;
  return static_cast<int>(b < 1 ? 1 : b);
}

}  // namespace dmpi

extern "C" int dmpi_upsample2x_bilinear_forward(const void* x, void* y, int N, int H, int W, int C, long long xsn,
                                                long long xsh, long long xsw, long long ysn, long long ysh,
                                                long long ysw, void* stream) {
  if (C % 8 != 0 || H < 1 || W < 1) return -1;
  if (((xsn | xsh | xsw | ysn | ysh | ysw) & 7) != 0) return -1;  // 16-byte accesses
  const float rh = H > 1 ? static_cast<float>(H - 1) / static_cast<float>(2 * H - 1) : 0.f;
  const float rw = W > 1 ? static_cast<float>(W - 1) / static_cast<float>(2 * W - 1) : 0.f;
  const long long total = 4LL * N * H * W * (C / 8);
  dmpi::upsample2x_fwd_kernel<<<dmpi::grid_for(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), N, H, W, C / 8, xsn, xsh, xsw, ysn, ysh, ysw,
      rh, rw);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_upsample2x_bilinear_backward(const void* dy, void* dx, int N, int H, int W, int C, long long dsn,
                                                 long long dsh, long long dsw, long long xsn, long long xsh,
                                                 long long xsw, void* stream) {
  if (C % 8 != 0 || H < 1 || W < 1) return -1;
  if (((xsn | xsh | xsw | dsn | dsh | dsw) & 7) != 0) return -1;
  const float rh = H > 1 ? static_cast<float>(H - 1) / static_cast<float>(2 * H - 1) : 0.f;
  const float rw = W > 1 ? static_cast<float>(W - 1) / static_cast<float>(2 * W - 1) : 0.f;
  const long long total = 1LL * N * H * W * (C / 8);
  dmpi::upsample2x_bwd_kernel<<<dmpi::grid_for(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), N, H, W, C / 8, dsn, dsh, dsw, xsn, xsh,
      xsw, rh, rw);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
