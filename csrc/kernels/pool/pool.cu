// Pooling kernels, NHWC bf16.
//   * MaxPool2d(kernel k, stride s, pad p): ResNet stem 3/2/1 (torchvision resnet.py maxpool) and UNet 2/2/0
//     (ref:pytorch/unet/model.py:25).  The forward stores the arg-max tap (uint8) so the backward routes
//     the gradient exactly like PyTorch (first maximum in scan order), without atomics.
//   * Global average pool [N,H,W,C] -> [N,C] and its backward (ResNet avgpool before fc).
// All are bandwidth-bound: 16-byte vector accesses, one thread per (pixel, 8 channels).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                          __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx,
                                                          int N, int H, int W, int C, int OH, int OW, int k, int s,
                                                          int p, long long xsn, long long xsh, long long xsw,
                                                          long long ysn, long long ysh, long long ysw) {
  // grid: x covers (ow, channel vector) of one output row, y = n * OH + oh  (32-bit index math only)
  const int vecC = C >> 3;
  const int n = blockIdx.y / OH, oh = blockIdx.y - n * OH;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < OW * vecC; i += gridDim.x * blockDim.x) {
    const int ow = i / vecC, cv = i - ow * vecC;
    float best[8];
    int bidx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      best[j] = -INFINITY;
      bidx[j] = 0;
    }
    for (int r = 0; r < k; ++r) {
      const int ih = oh * s + r - p;
      if (ih < 0 || ih >= H) continue;
      for (int q = 0; q < k; ++q) {
        const int iw = ow * s + q - p;
        if (iw < 0 || iw >= W) continue;
        const uint4 v = *reinterpret_cast<const uint4*>(x + n * xsn + ih * xsh + iw * xsw + cv * 8);
        const uint32_t vi[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float a = bf16lo(vi[j]), b = bf16hi(vi[j]);
          if (a > best[2 * j] || a != a) {
            best[2 * j] = a;
            bidx[2 * j] = r * k + q;
          }
          if (b > best[2 * j + 1] || b != b) {
            best[2 * j + 1] = b;
            bidx[2 * j + 1] = r * k + q;
          }
        }
      }
    }
    uint4 o;
    o.x = pack_bf16x2(best[0], best[1]);
    o.y = pack_bf16x2(best[2], best[3]);
    o.z = pack_bf16x2(best[4], best[5]);
    o.w = pack_bf16x2(best[6], best[7]);
    *reinterpret_cast<uint4*>(y + n * ysn + oh * ysh + ow * ysw + cv * 8) = o;
    if (idx) {
      uint2 pk;
      pk.x = bidx[0] | (bidx[1] << 8) | (bidx[2] << 16) | (bidx[3] << 24);
      pk.y = bidx[4] | (bidx[5] << 8) | (bidx[6] << 16) | (bidx[7] << 24);
      *reinterpret_cast<uint2*>(idx + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + cv * 8) = pk;
    }
  }
}

// Compile-time (K, S, P) variant of the forward pass: the K*K loads of a window are issued before the first compare
// (the generic loop above waits for each load in turn: 2.7 TB/s on the ResNet stem).
template <int K, int S, int P>
__global__ void __launch_bounds__(256) maxpool_fwd_fixed_kernel(const __nv_bfloat16* __restrict__ x,
                                                                __nv_bfloat16* __restrict__ y,
                                                                uint8_t* __restrict__ idx, int N, int H, int W, int C,
                                                                int OH, int OW, long long xsn, long long xsh,
                                                                long long xsw, long long ysn, long long ysh,
                                                                long long ysw) {
  const int vecC = C >> 3;
  const int n = blockIdx.y / OH, oh = blockIdx.y - n * OH;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < OW * vecC; i += gridDim.x * blockDim.x) {
    const int ow = i / vecC, cv = i - ow * vecC;
    uint4 v[K][K];
    bool ok[K][K];
#pragma unroll
    for (int r = 0; r < K; ++r)
#pragma unroll
      for (int q = 0; q < K; ++q) {
        const int ih = oh * S + r - P, iw = ow * S + q - P;
        ok[r][q] = ih >= 0 && ih < H && iw >= 0 && iw < W;
        if (ok[r][q]) v[r][q] = *reinterpret_cast<const uint4*>(x + n * xsn + ih * xsh + iw * xsw + cv * 8);
      }
    float best[8];
    int bidx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      best[j] = -INFINITY;
      bidx[j] = 0;
    }
#pragma unroll
    for (int r = 0; r < K; ++r)
#pragma unroll
      for (int q = 0; q < K; ++q) {
        if (!ok[r][q]) continue;
        const uint32_t vi[4] = {v[r][q].x, v[r][q].y, v[r][q].z, v[r][q].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float a = bf16lo(vi[j]), b = bf16hi(vi[j]);
          if (a > best[2 * j] || a != a) {
            best[2 * j] = a;
            bidx[2 * j] = r * K + q;
          }
          if (b > best[2 * j + 1] || b != b) {
            best[2 * j + 1] = b;
            bidx[2 * j + 1] = r * K + q;
          }
        }
      }
    uint4 o;
    o.x = pack_bf16x2(best[0], best[1]);
    o.y = pack_bf16x2(best[2], best[3]);
    o.z = pack_bf16x2(best[4], best[5]);
    o.w = pack_bf16x2(best[6], best[7]);
    *reinterpret_cast<uint4*>(y + n * ysn + oh * ysh + ow * ysw + cv * 8) = o;
    if (idx) {
      uint2 pk;
      pk.x = bidx[0] | (bidx[1] << 8) | (bidx[2] << 16) | (bidx[3] << 24);
      pk.y = bidx[4] | (bidx[5] << 8) | (bidx[6] << 16) | (bidx[7] << 24);
      *reinterpret_cast<uint2*>(idx + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + cv * 8) = pk;
    }
  }
}

// One thread per (input pixel, 8 channels): gathers from every pooling window that covers the pixel.
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
                                                          const uint8_t* __restrict__ idx,
                                                          __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C,
                                                          int OH, int OW, int k, int s, int p, long long dsn,
                                                          long long dsh, long long dsw, long long xsn, long long xsh,
                                                          long long xsw, const __nv_bfloat16* __restrict__ addend,
                                                          long long asn, long long ash, long long asw) {
  // `addend` (optional): a second gradient of the pooled tensor (the UNet skip connection) added in the same pass
  // grid: x covers (iw, channel vector) of one input row, y = n * H + ih  (32-bit index math only)
  const int vecC = C >> 3;
  const int n = blockIdx.y / H, ih = blockIdx.y - n * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < W * vecC; i += gridDim.x * blockDim.x) {
    const int iw = i / vecC, cv = i - iw * vecC;
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = 0.f;
    // windows oh with oh*s - p <= ih <= oh*s - p + k - 1
    const int oh_lo = max(0, (ih + p - k + s) / s), oh_hi = min(OH - 1, (ih + p) / s);
    const int ow_lo = max(0, (iw + p - k + s) / s), ow_hi = min(OW - 1, (iw + p) / s);
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int r = ih - (oh * s - p);
      if (r < 0 || r >= k) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int q = iw - (ow * s - p);
        if (q < 0 || q >= k) continue;
        const int tap = r * k + q;
        const uint2 id = *reinterpret_cast<const uint2*>(idx + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + cv * 8);
        const uint4 d = *reinterpret_cast<const uint4*>(dy + n * dsn + oh * dsh + ow * dsw + cv * 8);
        const uint32_t di[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int tj = ((j < 4 ? id.x : id.y) >> ((j & 3) * 8)) & 0xFF;
          if (tj == tap) g[j] += (j & 1) ? bf16hi(di[j >> 1]) : bf16lo(di[j >> 1]);
        }
      }
    }
    if (addend != nullptr) {
      const uint4 a = *reinterpret_cast<const uint4*>(addend + n * asn + ih * ash + iw * asw + cv * 8);
      const uint32_t ai[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[2 * j] += bf16lo(ai[j]);
        g[2 * j + 1] += bf16hi(ai[j]);
      }
    }
    uint4 o;
    o.x = pack_bf16x2(g[0], g[1]);
    o.y = pack_bf16x2(g[2], g[3]);
    o.z = pack_bf16x2(g[4], g[5]);
    o.w = pack_bf16x2(g[6], g[7]);
    *reinterpret_cast<uint4*>(dx + n * xsn + ih * xsh + iw * xsw + cv * 8) = o;
  }
}

// Compile-time (K, S, P) variant of the backward pass: the generic kernel above is issue-bound (~500 instructions per
// 8-channel item: runtime divisions, per-channel byte extraction); here the candidate windows are an unrolled K x K
// loop with constant tap ids, the "is this pixel the arg-max" test is a 4-byte SIMD compare and the masking of the
// bf16 pairs is two byte permutes per 32-bit word.
template <int K, int S, int P>
__global__ void __launch_bounds__(256) maxpool_bwd_fixed_kernel(const __nv_bfloat16* __restrict__ dy,
                                                                const uint8_t* __restrict__ idx,
                                                                __nv_bfloat16* __restrict__ dx, int N, int H, int W,
                                                                int C, int OH, int OW, long long dsn, long long dsh,
                                                                long long dsw, long long xsn, long long xsh,
                                                                long long xsw, const __nv_bfloat16* __restrict__ addend,
                                                                long long asn, long long ash, long long asw) {
  const int vecC = C >> 3;
  const int n = blockIdx.y / H, ih = blockIdx.y - n * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < W * vecC; i += gridDim.x * blockDim.x) {
    const int iw = i / vecC, cv = i - iw * vecC;
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
    for (int r = 0; r < K; ++r) {
      const int th = ih + P - r;  // = oh * S for the window whose row r covers this pixel
      if (th < 0 || th % S != 0 || th / S >= OH) continue;
      const int oh = th / S;
#pragma unroll
      for (int q = 0; q < K; ++q) {
        const int tw = iw + P - q;
        if (tw < 0 || tw % S != 0 || tw / S >= OW) continue;
        const int ow = tw / S;
        const uint2 id = *reinterpret_cast<const uint2*>(idx + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + cv * 8);
        const uint4 d = *reinterpret_cast<const uint4*>(dy + n * dsn + oh * dsh + ow * dsw + cv * 8);
        const uint32_t tapv = static_cast<uint32_t>(r * K + q) * 0x01010101u;  // constant after unrolling
        const uint32_t m0 = __vcmpeq4(id.x, tapv), m1 = __vcmpeq4(id.y, tapv);  // 0xFF per matching channel byte
        const uint32_t w0 = d.x & __byte_perm(m0, 0, 0x1100), w1 = d.y & __byte_perm(m0, 0, 0x3322);
        const uint32_t w2 = d.z & __byte_perm(m1, 0, 0x1100), w3 = d.w & __byte_perm(m1, 0, 0x3322);
        g[0] += bf16lo(w0);
        g[1] += bf16hi(w0);
        g[2] += bf16lo(w1);
        g[3] += bf16hi(w1);
        g[4] += bf16lo(w2);
        g[5] += bf16hi(w2);
        g[6] += bf16lo(w3);
        g[7] += bf16hi(w3);
      }
    }
    if (addend != nullptr) {
      const uint4 a = *reinterpret_cast<const uint4*>(addend + n * asn + ih * ash + iw * asw + cv * 8);
      const uint32_t ai[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[2 * j] += bf16lo(ai[j]);
        g[2 * j + 1] += bf16hi(ai[j]);
      }
    }
    uint4 o;
    o.x = pack_bf16x2(g[0], g[1]);
    o.y = pack_bf16x2(g[2], g[3]);
    o.z = pack_bf16x2(g[4], g[5]);
    o.w = pack_bf16x2(g[6], g[7]);
    *reinterpret_cast<uint4*>(dx + n * xsn + ih * xsh + iw * xsw + cv * 8) = o;
  }
}

// Stride-2 pooling on even images (H = 2*OH, W = 2*OW; K = 3 with pad 1, or K = 2 without pad): one thread per 2x2
// input quad and 8 channels.  The quad's four pixels draw on the same <= 4 windows (K = 3) or on one window (K = 2), so
// every load is issued up front and unconditionally, nothing diverges inside a warp, and each window is fetched once per
// quad instead of once per pixel.  (The per-pixel kernel above serialises up to four dependent load round trips under
// divergent branches: 405 us for the ResNet stem, 1.4 TB/s.)
template <int K>
__global__ void __launch_bounds__(256) maxpool_bwd_quad_kernel(const __nv_bfloat16* __restrict__ dy,
                                                               const uint8_t* __restrict__ idx,
                                                               __nv_bfloat16* __restrict__ dx, int N, int C, int OH,
                                                               int OW, long long dsn, long long dsh, long long dsw,
                                                               long long xsn, long long xsh, long long xsw,
                                                               const __nv_bfloat16* __restrict__ addend, long long asn,
                                                               long long ash, long long asw) {
  constexpr int P = K == 3 ? 1 : 0;
  constexpr int NW = K == 3 ? 2 : 1;  // windows per dimension that touch the quad: m, m + 1 (K = 3) or m (K = 2)
  const int vecC = C >> 3;
  const int n = blockIdx.y / OH, m = blockIdx.y - n * OH;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < OW * vecC; i += gridDim.x * blockDim.x) {
    const int qw = i / vecC, cv = i - qw * vecC;
    uint4 d[NW][NW];
    uint2 id[NW][NW];
#pragma unroll
    for (int a = 0; a < NW; ++a)
#pragma unroll
      for (int b = 0; b < NW; ++b) {
        d[a][b] = make_uint4(0, 0, 0, 0);
        id[a][b] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);  // matches no tap
        if (m + a < OH && qw + b < OW) {
          id[a][b] = *reinterpret_cast<const uint2*>(idx + ((static_cast<long long>(n) * OH + m + a) * OW + qw + b) * C + cv * 8);
          d[a][b] = *reinterpret_cast<const uint4*>(dy + n * dsn + (m + a) * dsh + (qw + b) * dsw + cv * 8);
        }
      }
    uint4 ad[2][2];
    if (addend != nullptr) {
#pragma unroll
      for (int dh = 0; dh < 2; ++dh)
#pragma unroll
        for (int dw = 0; dw < 2; ++dw)
          ad[dh][dw] = *reinterpret_cast<const uint4*>(addend + n * asn + (2 * m + dh) * ash + (2 * qw + dw) * asw + cv * 8);
    }
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        float g[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
        // window (m + a, qw + b) sees pixel (2m + dh, 2qw + dw) at tap (dh + P - 2a, dw + P - 2b); ascending tap order
#pragma unroll
        for (int a = NW - 1; a >= 0; --a)
#pragma unroll
          for (int b = NW - 1; b >= 0; --b) {
            const int r = dh + P - 2 * a, q = dw + P - 2 * b;
            if (r < 0 || r >= K || q < 0 || q >= K) continue;  // resolved at compile time
            const uint32_t tapv = static_cast<uint32_t>(r * K + q) * 0x01010101u;
            const uint32_t m0 = __vcmpeq4(id[a][b].x, tapv), m1 = __vcmpeq4(id[a][b].y, tapv);
            const uint32_t w0 = d[a][b].x & __byte_perm(m0, 0, 0x1100), w1 = d[a][b].y & __byte_perm(m0, 0, 0x3322);
            const uint32_t w2 = d[a][b].z & __byte_perm(m1, 0, 0x1100), w3 = d[a][b].w & __byte_perm(m1, 0, 0x3322);
            g[0] += bf16lo(w0);
            g[1] += bf16hi(w0);
            g[2] += bf16lo(w1);
            g[3] += bf16hi(w1);
            g[4] += bf16lo(w2);
            g[5] += bf16hi(w2);
            g[6] += bf16lo(w3);
            g[7] += bf16hi(w3);
          }
        if (addend != nullptr) {
          const uint32_t ai[4] = {ad[dh][dw].x, ad[dh][dw].y, ad[dh][dw].z, ad[dh][dw].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            g[2 * j] += bf16lo(ai[j]);
            g[2 * j + 1] += bf16hi(ai[j]);
          }
        }
        uint4 o;
        o.x = pack_bf16x2(g[0], g[1]);
        o.y = pack_bf16x2(g[2], g[3]);
        o.z = pack_bf16x2(g[4], g[5]);
        o.w = pack_bf16x2(g[6], g[7]);
        *reinterpret_cast<uint4*>(dx + n * xsn + (2 * m + dh) * xsh + (2 * qw + dw) * xsw + cv * 8) = o;
      }
  }
}

// [N, HW, C] -> [N, C]: block = (C/8 vector lanes) x row slabs, tree-free accumulate then smem reduce.
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                          __nv_bfloat16* __restrict__ y, int HW, int C) {
  const int n = blockIdx.y;
  const int c0 = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c0 >= C) return;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  const __nv_bfloat16* base = x + static_cast<long long>(n) * HW * C + c0;
  for (int i = 0; i < HW; ++i) {
    const uint4 v = *reinterpret_cast<const uint4*>(base + static_cast<long long>(i) * C);
    const uint32_t vi[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      s[2 * j] += bf16lo(vi[j]);
      s[2 * j + 1] += bf16hi(vi[j]);
    }
  }
  const float inv = 1.f / static_cast<float>(HW);
  uint4 o;
  o.x = pack_bf16x2(s[0] * inv, s[1] * inv);
  o.y = pack_bf16x2(s[2] * inv, s[3] * inv);
  o.z = pack_bf16x2(s[4] * inv, s[5] * inv);
  o.w = pack_bf16x2(s[6] * inv, s[7] * inv);
  *reinterpret_cast<uint4*>(y + static_cast<long long>(n) * C + c0) = o;
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
                                                          __nv_bfloat16* __restrict__ dx, int N, int HW, int C) {
  const int vecC = C >> 3;
  const long long total = static_cast<long long>(N) * HW * vecC;
  const float inv = 1.f / static_cast<float>(HW);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % vecC);
    const long long n = i / (static_cast<long long>(HW) * vecC);
    const uint4 d = *reinterpret_cast<const uint4*>(dy + n * C + cv * 8);
    uint4 o;
    o.x = pack_bf16x2(bf16lo(d.x) * inv, bf16hi(d.x) * inv);
    o.y = pack_bf16x2(bf16lo(d.y) * inv, bf16hi(d.y) * inv);
    o.z = pack_bf16x2(bf16lo(d.z) * inv, bf16hi(d.z) * inv);
    o.w = pack_bf16x2(bf16lo(d.w) * inv, bf16hi(d.w) * inv);
    *reinterpret_cast<uint4*>(dx + i * 8) = o;
  }
}

static int ew_grid(long long items) {
  long long b = (items + 255) / 256;
  if (b > 148LL * 16) b = 148LL * 16;
  return b < 1 ? 1 : static_cast<int>(b);
}

}  // namespace dmpi

using namespace dmpi;

extern "C" int dmpi_maxpool_forward(const void* x, void* y, void* idx, int N, int H, int W, int C, int OH, int OW,
                                    int k, int s, int p, long long xsn, long long xsh, long long xsw, long long ysn,
                                    long long ysh, long long ysw, void* stream) {
  if (C % 8 || k * k > 255) return -1;
  const dim3 grid((OW * (C / 8) + 255) / 256, N * OH);
  if ((k == 3 && s == 2 && p == 1) || (k == 2 && s == 2 && p == 0)) {
    auto kern = k == 3 ? maxpool_fwd_fixed_kernel<3, 2, 1> : maxpool_fwd_fixed_kernel<2, 2, 0>;
    kern<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), N, H, W, C, OH,
        OW, xsn, xsh, xsw, ysn, ysh, ysw);
    return cudaGetLastError() == cudaSuccess ? 0 : -2;
  }
  maxpool_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), N, H, W, C, OH,
      OW, k, s, p, xsn, xsh, xsw, ysn, ysh, ysw);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_maxpool_backward_add(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int OH,
                                         int OW, int k, int s, int p, long long dsn, long long dsh, long long dsw,
                                         long long xsn, long long xsh, long long xsw, const void* addend,
                                         long long asn, long long ash, long long asw, void* stream) {
  if (C % 8) return -1;
  if (addend != nullptr && ((reinterpret_cast<uintptr_t>(addend) & 15) || asn % 8 || ash % 8 || asw % 8)) return -3;
  if (((k == 3 && s == 2 && p == 1) || (k == 2 && s == 2 && p == 0)) && H == 2 * OH && W == 2 * OW) {
    const dim3 qgrid((OW * (C / 8) + 255) / 256, N * OH);
    auto kern = k == 3 ? maxpool_bwd_quad_kernel<3> : maxpool_bwd_quad_kernel<2>;
    kern<<<qgrid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), N, C,
        OH, OW, dsn, dsh, dsw, xsn, xsh, xsw, static_cast<const __nv_bfloat16*>(addend), asn, ash, asw);
    return cudaGetLastError() == cudaSuccess ? 0 : -2;
  }
  const dim3 grid((W * (C / 8) + 255) / 256, N * H);
  if ((k == 3 && s == 2 && p == 1) || (k == 2 && s == 2 && p == 0)) {
    auto kern = k == 3 ? maxpool_bwd_fixed_kernel<3, 2, 1> : maxpool_bwd_fixed_kernel<2, 2, 0>;
    kern<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), N, H,
        W, C, OH, OW, dsn, dsh, dsw, xsn, xsh, xsw, static_cast<const __nv_bfloat16*>(addend), asn, ash, asw);
    return cudaGetLastError() == cudaSuccess ? 0 : -2;
  }
  maxpool_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), N, H,
      W, C, OH, OW, k, s, p, dsn, dsh, dsw, xsn, xsh, xsw, static_cast<const __nv_bfloat16*>(addend), asn, ash, asw);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_maxpool_backward(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int OH,
                                     int OW, int k, int s, int p, long long dsn, long long dsh, long long dsw,
                                     long long xsn, long long xsh, long long xsw, void* stream) {
  return dmpi_maxpool_backward_add(dy, idx, dx, N, H, W, C, OH, OW, k, s, p, dsn, dsh, dsw, xsn, xsh, xsw, nullptr, 0,
                                   0, 0, stream);
}

extern "C" int dmpi_avgpool_forward(const void* x, void* y, int N, int HW, int C, void* stream) {
  if (C % 8) return -1;
  const int threads = 64;
  dim3 grid((C / 8 + threads - 1) / threads, N);
  avgpool_fwd_kernel<<<grid, threads, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x),
                                                                              static_cast<__nv_bfloat16*>(y), HW, C);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_avgpool_backward(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  if (C % 8) return -1;
  const long long items = static_cast<long long>(N) * HW * (C / 8);
  avgpool_bwd_kernel<<<ew_grid(items), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), N, HW, C);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
