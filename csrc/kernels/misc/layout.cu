// Layout / staging kernels.
//   * input staging  : NCHW fp32 (what a torch DataLoader yields, ref:pytorch/resnet/main.py:126,
//                      ref:pytorch/unet/train.py:175) -> zero-padded NHWC8 bf16 for the RGB stem.
//   * weight shadows : the optimizer keeps fp32 master weights in PyTorch-compatible layouts (checkpoint parity,
//                      SURVEY.md N25) and a bf16 mirror; dgrad / ConvTranspose need the *transposed* K-major
//                      form, produced here for every layer by ONE table-driven launch.
//   * stem repack    : [K, R, S, 3] <-> the [K, R, 8, 8] window layout used by smallc_fprop / smallc_wgrad.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

__global__ void __launch_bounds__(256) nchw_to_nhwc8_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                            int N, int C, int H, int W, int pad, int Hp, int Wp) {
  const long long total = static_cast<long long>(N) * Hp * Wp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int wp = static_cast<int>(i % Wp);
    const long long t = i / Wp;
    const int hp = static_cast<int>(t % Hp);
    const int n = static_cast<int>(t / Hp);
    const int h = hp - pad, w = wp - pad;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (h >= 0 && h < H && w >= 0 && w < W) {
      const float* src = x + (static_cast<long long>(n) * C * H + h) * W + w;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c < C) v[c] = __ldg(src + static_cast<long long>(c) * H * W);
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    reinterpret_cast<uint4*>(out)[i] = o;
  }
}

// generic NCHW fp32 -> NHWC bf16 (C % 8 == 0) and back (fp32 out), used at framework boundaries only
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                           int N, int C, int HW) {
  const long long total = static_cast<long long>(N) * HW * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const long long t = i / C;
    const int p = static_cast<int>(t % HW);
    const long long n = t / HW;
    out[i] = __float2bfloat16(x[(n * C + c) * HW + p]);
  }
}
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                           int N, int C, int HW, long long ldx) {
  const long long total = static_cast<long long>(N) * HW * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int p = static_cast<int>(i % HW);
    const long long t = i / HW;
    const int c = static_cast<int>(t % C);
    const long long n = t / C;
    out[i] = __bfloat162float(x[(n * HW + p) * ldx + c]);
  }
}

struct TransposeEntry {
  const __nv_bfloat16* src;  // [A][T][B]
  __nv_bfloat16* dst;        // mode 0: [B][T][A] with row stride ld;  mode 1: [T][B][A]
  int A, T, B, mode;
  long long ld;
  int tile_begin, tiles_a, tiles_b, pad_;
};

__global__ void __launch_bounds__(256) transpose_table_kernel(const TransposeEntry* __restrict__ table, int n_entries) {
  __shared__ __nv_bfloat16 tile[32][33];
  // locate the entry of this block (binary search on tile_begin)
  int lo = 0, hi = n_entries - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (table[mid].tile_begin <= static_cast<int>(blockIdx.x)) lo = mid;
    else hi = mid - 1;
  }
  const TransposeEntry e = table[lo];
  int local = blockIdx.x - e.tile_begin;
  const int tb = local % e.tiles_b;
  local /= e.tiles_b;
  const int ta = local % e.tiles_a;
  const int t = local / e.tiles_a;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int r = ty; r < 32; r += 8) {
    const int a = ta * 32 + r, b = tb * 32 + tx;
    if (a < e.A && b < e.B) tile[r][tx] = e.src[(static_cast<long long>(a) * e.T + t) * e.B + b];
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int b = tb * 32 + r, a = ta * 32 + tx;
    if (a < e.A && b < e.B) {
      const long long o = e.mode == 0 ? static_cast<long long>(b) * e.ld + static_cast<long long>(t) * e.A + a
                                      : (static_cast<long long>(t) * e.B + b) * e.A + a;
      e.dst[o] = tile[tx][r];
    }
  }
}

// master fp32 [K][R][S][C] (C <= 8) -> bf16 [K][R][8][8] (zero where s >= S or c >= C)
__global__ void stem_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int K, int R, int S, int C) {
  const int total = K * R * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i & 7, s = (i >> 3) & 7, kr = i >> 6;
    float v = 0.f;
    if (s < S && c < C) v = w[(static_cast<long long>(kr) * S + s) * C + c];
    out[i] = __float2bfloat16(v);
  }
}
// fp32 window gradient [K][R][8][8] -> accumulate into master-layout gradient [K][R][S][C]
__global__ void stem_unpack_grad_kernel(const float* __restrict__ dwin, float* __restrict__ dw, int K, int R, int S, int C) {
  const int total = K * R * S * C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i % C;
    const int s = (i / C) % S;
    const int kr = i / (C * S);
    dw[i] += dwin[kr * 64 + s * 8 + c];
  }
}

}  // namespace dmpi

using namespace dmpi;

static int lgrid(long long items) {
  long long b = (items + 255) / 256;
  if (b > 148LL * 16) b = 148LL * 16;
  return b < 1 ? 1 : static_cast<int>(b);
}

// Clears a byte range (gradient arena, BatchNorm / loss accumulators - everything a step has to start from zero) with
// a plain store kernel.  Not cudaMemsetAsync: inside a replayed CUDA graph a memset NODE delays the kernel behind it by
// ~16 us on this driver (measured with bench/timeline.py: 4-byte memset -> next kernel), a kernel node by ~0.3 us.
namespace dmpi {
__global__ void __launch_bounds__(256) zero_fill_kernel(uint4* __restrict__ p16, long long nvec, unsigned char* __restrict__ tail,
                                                        int ntail, unsigned char* __restrict__ head, int nhead) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec; i += stride) p16[i] = z;
  if (blockIdx.x == 0) {
    if (static_cast<int>(threadIdx.x) < ntail) tail[threadIdx.x] = 0;
    if (static_cast<int>(threadIdx.x) < nhead) head[threadIdx.x] = 0;
  }
}
}  // namespace dmpi

extern "C" int dmpi_memset_async(void* p, int value, long long bytes, void* stream) {
  if (bytes <= 0) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (value != 0)
    return cudaMemsetAsync(p, value, static_cast<size_t>(bytes), s) == cudaSuccess ? 0 : -2;
  unsigned char* b = static_cast<unsigned char*>(p);
  const int nhead = static_cast<int>((16 - (reinterpret_cast<uintptr_t>(b) & 15)) & 15) < bytes
                        ? static_cast<int>((16 - (reinterpret_cast<uintptr_t>(b) & 15)) & 15) : static_cast<int>(bytes);
  const long long body = bytes - nhead;
  const long long nvec = body / 16;
  const int ntail = static_cast<int>(body - nvec * 16);
  long long blocks = (nvec + 256 * 8 - 1) / (256 * 8);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  dmpi::zero_fill_kernel<<<static_cast<int>(blocks), 256, 0, s>>>(reinterpret_cast<uint4*>(b + nhead), nvec,
                                                                  b + nhead + nvec * 16, ntail, b, nhead);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_nchw_to_nhwc8_padded(const float* x, void* out, int N, int C, int H, int W, int pad, int Hp, int Wp,
                                         void* stream) {
  if (C > 8) return -1;
  nchw_to_nhwc8_kernel<<<lgrid(static_cast<long long>(N) * Hp * Wp), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, static_cast<__nv_bfloat16*>(out), N, C, H, W, pad, Hp, Wp);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_nchw_to_nhwc(const float* x, void* out, int N, int C, int HW, void* stream) {
  nchw_to_nhwc_kernel<<<lgrid(static_cast<long long>(N) * C * HW), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, static_cast<__nv_bfloat16*>(out), N, C, HW);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_nhwc_to_nchw(const void* x, float* out, int N, int C, int HW, long long ldx, void* stream) {
  nhwc_to_nchw_kernel<<<lgrid(static_cast<long long>(N) * C * HW), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), out, N, C, HW, ldx);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

// `table` is a device array of n_entries records laid out as dmpi::TransposeEntry (see python side).
extern "C" int dmpi_transpose_table(const void* table, int n_entries, int total_tiles, void* stream) {
  if (n_entries <= 0 || total_tiles <= 0) return 0;
  transpose_table_kernel<<<total_tiles, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const TransposeEntry*>(table), n_entries);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_stem_pack(const float* w, void* out, int K, int R, int S, int C, void* stream) {
  if (S > 8 || C > 8) return -1;
  stem_pack_kernel<<<lgrid(K * R * 64), 256, 0, static_cast<cudaStream_t>(stream)>>>(w, static_cast<__nv_bfloat16*>(out), K,
                                                                                   R, S, C);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_stem_unpack_grad(const float* dwin, float* dw, int K, int R, int S, int C, void* stream) {
  stem_unpack_grad_kernel<<<lgrid(K * R * S * C), 256, 0, static_cast<cudaStream_t>(stream)>>>(dwin, dw, K, R, S, C);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
