// Hardware probe (test-only): does a tcgen05 shared-memory descriptor whose start address is shifted by whole 128-byte
// rows INSIDE a 1024-byte swizzle atom read the rows a TMA load put there?  This is what lets one shared-memory patch
// of an NHWC image serve all filter taps of a 3x3 convolution (tap = pixel shift = 128-byte row shift for C = 64).
//
//   mode bit 0: put (shift & 7) into the descriptor's base-offset field (bits 49..51) instead of leaving it 0
//   mode bit 1: MN-major variant (wgrad operand): D[ch_y, ch_x] = sum_px Y[px, ch_y] * X[px + shift, ch_x]
//               else K-major (fprop operand):      D[px, n]      = sum_c  X[px + shift, c] * W[n, c]
//
// One CTA, 128 threads.  X is [rows, 64] bf16, loaded whole (rows <= 192) by one TMA box with 128B swizzle.
#include "../../common/ptx.cuh"
#include "../../common/tensormap.h"

namespace dmpi {

struct ProbeParams {
  CUtensorMap x_map;  // [rows, 64]
  CUtensorMap w_map;  // K-major: W [64, 64];  MN-major: Y [64 px, 128 ch] (two 64-channel boxes)
  float* out;         // K-major: [128, 64];   MN-major: [128, 64]
  int rows, shift, mode;
};

__device__ __forceinline__ uint64_t probe_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t base_off) {
  uint64_t d = umma_smem_desc(saddr, lbo, sbo);
  d |= static_cast<uint64_t>(base_off & 7u) << 49;
  return d;
}

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const __grid_constant__ ProbeParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* xs = smem;                 // rows * 128 bytes (<= 24 KB)
  uint8_t* ws = smem + 24 * 1024;     // 8 KB (W) or 16 KB (Y: two atoms)
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 40 * 1024);
  uint64_t* done = bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool mn = (p.mode & 2) != 0;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) {
    const uint32_t bytes = static_cast<uint32_t>(p.rows) * 128u + (mn ? 16384u : 8192u);
    mbar_arrive_expect_tx(bar, bytes);
    tma_load_2d(xs, &p.x_map, bar, 0, 0);
    if (mn) {
      tma_load_2d(ws, &p.w_map, bar, 0, 0);
      tma_load_2d(ws + 8192, &p.w_map, bar, 64, 0);
    } else {
      tma_load_2d(ws, &p.w_map, bar, 0, 0);
    }
    mbar_wait(bar, 0);
    tc_fence_after_sync();
    const uint32_t boff = (p.mode & 1) ? static_cast<uint32_t>(p.shift & 7) : 0u;
    const uint32_t x_addr = smem_u32(xs) + static_cast<uint32_t>(p.shift) * 128u;
    const uint32_t w_addr = smem_u32(ws);
    if (!mn) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t adesc = probe_desc(x_addr + k * 32, 0, 1024, boff);
        const uint64_t bdesc = umma_smem_desc(w_addr + k * 32, 0, 1024);
        umma_bf16(tmem_base, adesc, bdesc, idesc, k != 0 ? 1u : 0u);
      }
    } else {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t adesc = umma_smem_desc(w_addr + k * 2048, 8192, 1024);
        const uint64_t bdesc = probe_desc(x_addr + k * 2048, 8192, 1024, boff);
        umma_bf16(tmem_base, adesc, bdesc, idesc, k != 0 ? 1u : 0u);
      }
    }
    umma_commit(done);
  }
  __syncwarp();
  mbar_wait(done, 0);
  tc_fence_after_sync();
  const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
  const int row = warp * 32 + lane;
#pragma unroll 1
  for (int chunk = 0; chunk < 2; ++chunk) {
    uint32_t v[32];
    tmem_ld_32x32(t_addr + chunk * 32, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) p.out[row * 64 + chunk * 32 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc<64>(tmem_base);
}

}  // namespace dmpi

using namespace dmpi;

// x: [rows, 64] bf16;  w: K-major W [64, 64] bf16, or (mode & 2) Y [64, 128] bf16;  out: [128, 64] fp32.
extern "C" int dmpi_umma_probe(const void* x, const void* w, float* out, int rows, int shift, int mode, void* stream) {
  if (rows > 192 || rows < 1) return -1;
  ProbeParams p;
  {
    int64_t dims[2] = {64, rows};
    int64_t strides[1] = {64};
    int box[2] = {64, rows};
    if (encode_bf16_map(&p.x_map, x, 2, dims, strides, box)) return -2;
  }
  if (mode & 2) {
    int64_t dims[2] = {128, 64};
    int64_t strides[1] = {128};
    int box[2] = {64, 64};
    if (encode_bf16_map(&p.w_map, w, 2, dims, strides, box)) return -3;
  } else {
    int64_t dims[2] = {64, 64};
    int64_t strides[1] = {64};
    int box[2] = {64, 64};
    if (encode_bf16_map(&p.w_map, w, 2, dims, strides, box)) return -3;
  }
  p.out = out;
  p.rows = rows;
  p.shift = shift;
  p.mode = mode;
  const int smem = 42 * 1024 + 1024;
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -4;
    attr = true;
  }
  umma_probe_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}
