// Weight gradient of 3x3 (R x S, stride 1, halo = pad) convolutions from ONE shared-memory input patch per pixel tile.
//
//   dW[k, tap, c] += sum over output pixels p  dY[p, k] * X[p + tap, c]
//
// wgrad.cu streams a separately shifted X tile per tap (nine loads of the same pixels) and, for C = K = 64, wastes half
// of every M = 128 MMA.  Here the roles are swapped and the taps come from row-shifted descriptors:
//   * per pipeline stage one TMA box brings the (box_h + 2) x (box_w + 2) pixel patch of X (64 channels) and one box
//     brings the matching dY rows (64 filters); pixels are the GEMM K axis, both operands MN-major;
//   * M side = X channels of TWO taps: the two 64-channel "atoms" of an M = 128 operand are the same patch at two
//     row shifts, expressed through the descriptor's leading byte offset (LBO = shift difference * 128 B);
//   * N side = 64 dY channels; five accumulators (tap pairs 0-1, 2-3, 4-5, 6-7, 8-8) of 64 TMEM columns each.
// K runs over the flattened patch rows (row-major over the padded width), so the dY tile must be zero at the padded
// positions: full-width tiles get that from TMA's out-of-bounds fill, row-segment tiles (box_h == 1) have no padded
// positions inside K.  The tails that K's rounding to 16 and the row shifts read beyond the boxes are zeroed once.
// Split-K over pixel tiles; partial sums are added into the fp32 gradient (the DDP bucket) with red.global.add.
//
// Relies on the hardware behaviour pinned by tests/test_conv_gpu.py::test_umma_row_shifted_descriptors.
#include <stdio.h>

#include "../../common/ptx.cuh"
#include "../../common/tensormap.h"
#include "conv_common.cuh"

namespace dmpi {

constexpr int kWpThreads = 256;
constexpr int kWpGroups = 5;      // tap pairs of a 3x3 filter
constexpr int kWpTmemCols = 512;  // 5 x 64 columns, rounded up to a power of two

struct WgradPatchParams {
  CUtensorMap x_map;   // X view, box (64, pw, box_h + 2*halo, 1)
  CUtensorMap y_map;   // dY view, box (64, yw, box_h, 1)   (yw = pw for full-width tiles, box_w for row segments)
  float* dw;
  long long w_ld;
  int tap_woff[9];
  int shift[9];        // patch row shift of each tap, in pixels
  int box_w, box_h, halo, pw;
  int tiles_w, tiles_h, tiles_n, num_m_tiles;
  int k_steps;         // K / 16 per stage
  int x_bytes, y_bytes, x_stride, y_stride, stages;
  int k_tiles, c_tiles, splits;
  int k_out, c_total;
  float* partials;  // deterministic split-K: [tile][split][5][128][64] floats, added in split order by the reduce kernel
};

__global__ void __launch_bounds__(kWpThreads, 1) wgrad_patch_kernel(const __grid_constant__ WgradPatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.stages;
  const int stage_bytes = p.x_stride + p.y_stride;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + 8;
  uint64_t* tfull = bars + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int c_blk = blockIdx.x % p.c_tiles;
  const int k_blk = blockIdx.x / p.c_tiles;
  const int split = blockIdx.y;
  const int pt_begin = static_cast<int>((static_cast<long long>(p.num_m_tiles) * split) / p.splits);
  const int pt_end = static_cast<int>((static_cast<long long>(p.num_m_tiles) * (split + 1)) / p.splits);
  const int num_it = pt_end - pt_begin;

  // zero, once, the regions beyond the TMA boxes: they are read (K rounding, row shifts) but never written
  for (int st = 0; st < S; ++st) {
    uint8_t* sx = smem + st * stage_bytes;
    for (int i = p.x_bytes + threadIdx.x * 16; i < p.x_stride; i += kWpThreads * 16)
      *reinterpret_cast<uint4*>(sx + i) = make_uint4(0, 0, 0, 0);
    for (int i = p.x_stride + p.y_bytes + threadIdx.x * 16; i < stage_bytes; i += kWpThreads * 16)
      *reinterpret_cast<uint4*>(sx + i) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.x_map);
    tma_prefetch_desc(&p.y_map);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<kWpTmemCols>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (num_it > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const uint32_t tx = static_cast<uint32_t>(p.x_bytes + p.y_bytes);
        int stage = 0;
        uint32_t phase = 0;
        for (int it = 0; it < num_it; ++it) {
          const int m_blk = pt_begin + it;
          const int tw = m_blk % p.tiles_w;
          const int th = (m_blk / p.tiles_w) % p.tiles_h;
          const int tn = m_blk / (p.tiles_w * p.tiles_h);
          const int w0 = tw * p.box_w, h0 = th * p.box_h;
          mbar_wait_spin(&empty[stage], phase ^ 1u);
          mbar_arrive_expect_tx(&full[stage], tx);
          uint8_t* sx = smem + stage * stage_bytes;
          tma_load_4d(sx, &p.x_map, &full[stage], c_blk * 64, w0 - p.halo, h0 - p.halo, tn);
          tma_load_4d(sx + p.x_stride, &p.y_map, &full[stage], k_blk * 64, w0, h0, tn);
          if (++stage == S) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
      __syncwarp();
    } else if (warp == 1) {
      const bool leader = elect_one();
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
      const uint32_t hi = umma_desc_hi(1024);
      // A descriptors of the five tap pairs: low word without the start address (LBO = shift difference of the pair)
      uint32_t a_off[kWpGroups], a_lbo[kWpGroups];
#pragma unroll
      for (int g = 0; g < kWpGroups; ++g) {
        const int t0 = 2 * g, t1 = (2 * g + 1 < 9) ? 2 * g + 1 : 2 * g;
        a_off[g] = static_cast<uint32_t>(p.shift[t0]) * 8u;                                   // 128 B >> 4
        a_lbo[g] = (static_cast<uint32_t>(p.shift[t1] - p.shift[t0]) * 8u & 0x3FFFu) << 16;    // LBO field
      }
      const uint32_t ring = (smem_u32(smem) & 0x3FFFFu) >> 4;
      const uint32_t stage_step = static_cast<uint32_t>(stage_bytes) >> 4;
      const uint32_t y_off = static_cast<uint32_t>(p.x_stride) >> 4;
      const int k_steps = p.k_steps;
      int stage = 0;
      uint32_t phase = 0;
      uint32_t accum = 0;
      for (int it = 0; it < num_it; ++it) {
        mbar_wait_spin(&full[stage], phase);
        tc_fence_after_sync();
        const uint32_t x_lo = ring + static_cast<uint32_t>(stage) * stage_step;
        const uint32_t y_lo = x_lo + y_off;
        if (leader) {
#pragma unroll 1
          for (int kk = 0; kk < k_steps; ++kk) {
            const uint32_t kofs = static_cast<uint32_t>(kk) * 128u;  // 16 pixel rows = 2048 B
#pragma unroll
            for (int g = 0; g < kWpGroups; ++g)
              umma_bf16_lohi(tmem_base + g * 64, (x_lo + a_off[g] + kofs) | a_lbo[g], hi, y_lo + kofs, hi, idesc,
                             kk == 0 ? accum : 1u);
          }
          umma_commit(&empty[stage]);
        }
        accum = 1u;
        if (++stage == S) {
          stage = 0;
          phase ^= 1u;
        }
      }
      if (leader) umma_commit(tfull);
      __syncwarp();
    } else if (warp >= 4) {
      // accumulator row m of group g: tap 2g + m/64, input channel c_blk*64 + m%64; columns: 64 filters
      const int q = warp - 4;
      const int m = q * 32 + lane;
      const int c = c_blk * 64 + (m & 63);
      mbar_wait(tfull, 0);
      tc_fence_after_sync();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int g = 0; g < kWpGroups; ++g) {
        const int tap = 2 * g + (m >> 6);
        const bool live = tap < 9 && c < p.c_total;
#pragma unroll 1
        for (int chunk = 0; chunk < 2; ++chunk) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + g * 64 + chunk * 32, v);
          tmem_ld_wait();
          if (p.partials != nullptr) {
            float* part = p.partials + ((static_cast<size_t>(blockIdx.x) * gridDim.y + blockIdx.y) * kWpGroups + g) * (128 * 64) +
                          m * 64 + chunk * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<uint4*>(part + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else if (live) {
            // lanes hold consecutive input channels: every red below is one coalesced 128-byte request per filter
            float* dst = p.dw + static_cast<long long>(k_blk * 64 + chunk * 32) * p.w_ld + p.tap_woff[tap] + c;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (k_blk * 64 + chunk * 32 + j < p.k_out)
                asm volatile("red.global.add.f32 [%0], %1;" ::"l"(dst + static_cast<long long>(j) * p.w_ld),
                             "f"(__uint_as_float(v[j]))
                             : "memory");
            }
          }
        }
      }
    }
  }
  else if (warp >= 4 && p.partials != nullptr) {
    // no pixel tiles for this split: its partial must still read as zero
    const int m = (warp - 4) * 32 + lane;
    for (int g = 0; g < kWpGroups; ++g) {
      float* part = p.partials + ((static_cast<size_t>(blockIdx.x) * gridDim.y + blockIdx.y) * kWpGroups + g) * (128 * 64) + m * 64;
      for (int j = 0; j < 64; j += 4) *reinterpret_cast<uint4*>(part + j) = make_uint4(0, 0, 0, 0);
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<kWpTmemCols>(tmem_base);
}

// Deterministic split-K: grid = (tiles, 5 tap pairs); adds the partial tiles in split order onto the gradient.
__global__ void __launch_bounds__(256) wgrad_patch_reduce_kernel(const __grid_constant__ WgradPatchParams p) {
  const int tile = blockIdx.x, g = blockIdx.y;
  const int c_blk = tile % p.c_tiles, k_blk = tile / p.c_tiles;
  const size_t split_stride = static_cast<size_t>(kWpGroups) * 128 * 64;
  const float* base = p.partials + (static_cast<size_t>(tile) * p.splits * kWpGroups + g) * (128 * 64);
  for (int i = threadIdx.x; i < 128 * 16; i += blockDim.x) {
    const int m = i >> 4, j = (i & 15) * 4;
    const int tap = 2 * g + (m >> 6), c = c_blk * 64 + (m & 63);
    if (tap >= 9 || c >= p.c_total) continue;
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* src = base + m * 64 + j;
    int sidx = 0;
    for (; sidx + 4 <= p.splits; sidx += 4) {
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = __ldcg(reinterpret_cast<const float4*>(src + (sidx + k) * split_stride));
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        t.x += v[k].x; t.y += v[k].y; t.z += v[k].z; t.w += v[k].w;
      }
    }
    for (; sidx < p.splits; ++sidx) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + sidx * split_stride));
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    const float tv[4] = {t.x, t.y, t.z, t.w};
    float* dst = p.dw + static_cast<long long>(k_blk * 64 + j) * p.w_ld + p.tap_woff[tap] + c;
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (k_blk * 64 + j + e < p.k_out) dst[static_cast<long long>(e) * p.w_ld] += tv[e];
  }
}

}  // namespace dmpi

using namespace dmpi;

// desc: a_view[0] = X (NHWC view), o_view[0] = dY, box_w x box_h = pixel tile (box_h == 1 for row segments, else
// box_w == image width), halo = pad, ntaps == 9 with tap_dh / tap_dw / tap_woff, grid_w/h/n = output dims,
// k_out, c_total, w_ld as for dmpi_wgrad_run.
extern "C" int dmpi_wgrad_patch_run(dmpi_conv_plan* plan, const void* dy, const void* x, float* dw, void* stream) {
  const dmpi_conv_desc& d = plan->desc;
  ConvKernelParams& kp = plan->params;
  if (d.ntaps != 9 || d.halo < 1 || d.box_n != 1) return -1;
  const int pw = d.box_w + 2 * d.halo;
  const bool segment = d.box_h == 1 && d.box_w < d.grid_w;
  if (!segment && d.box_w != d.grid_w) return -2;  // multi-row tiles need the zero fill right of the image
  const int yw = segment ? d.box_w : pw;
  const int K = yw * d.box_h;
  const int k_steps = (K + 15) / 16;
  if (pw > 256 || yw > 256 || d.box_h + 2 * d.halo > 256) return -3;
  WgradPatchParams p;
  if (!plan->maps_valid || plan->cached_a != x || plan->cached_b != dy) {
    const int xbox[4] = {64, pw, d.box_h + 2 * d.halo, 1};
    const int ybox[4] = {64, yw, d.box_h, 1};
    if (encode_bf16_map(&kp.a_map[0], static_cast<const uint8_t*>(x) + d.a_view[0].offset * 2, 4, d.a_view[0].dims,
                        d.a_view[0].strides, xbox))
      return -4;
    if (encode_bf16_map(&kp.b_map, static_cast<const uint8_t*>(dy) + d.o_view[0].offset * 2, 4, d.o_view[0].dims,
                        d.o_view[0].strides, ybox))
      return -5;
    plan->cached_a = x;
    plan->cached_b = dy;
    plan->maps_valid = true;
  }
  p.x_map = kp.a_map[0];
  p.y_map = kp.b_map;
  p.dw = dw;
  p.w_ld = d.w_ld;
  int max_shift = 0;
  for (int t = 0; t < 9; ++t) {
    p.tap_woff[t] = d.tap_woff[t];
    p.shift[t] = (d.tap_dh[t] + d.halo) * pw + d.tap_dw[t] + d.halo;
    if (p.shift[t] < 0) return -6;
    if (p.shift[t] > max_shift) max_shift = p.shift[t];
  }
  for (int g = 0; g < 4; ++g)
    if (p.shift[2 * g + 1] < p.shift[2 * g]) return -6;  // LBO is an unsigned stride
  p.box_w = d.box_w;
  p.box_h = d.box_h;
  p.halo = d.halo;
  p.pw = pw;
  p.tiles_w = (d.grid_w + d.box_w - 1) / d.box_w;
  p.tiles_h = (d.grid_h + d.box_h - 1) / d.box_h;
  p.tiles_n = d.grid_n;
  p.num_m_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
  p.k_steps = k_steps;
  p.x_bytes = pw * (d.box_h + 2 * d.halo) * 128;
  p.y_bytes = yw * d.box_h * 128;
  const int x_rows = 16 * k_steps + max_shift;
  const int x_need = x_rows * 128 > p.x_bytes ? x_rows * 128 : p.x_bytes;
  p.x_stride = ((x_need + 1023) / 1024) * 1024;
  p.y_stride = ((16 * k_steps * 128 + 1023) / 1024) * 1024;
  const int stage_bytes = p.x_stride + p.y_stride;
  int stages = (220 * 1024) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return -7;
  p.stages = stages;
  p.k_tiles = (d.k_out + 63) / 64;
  p.c_tiles = (d.c_total + 63) / 64;
  p.k_out = d.k_out;
  p.c_total = d.c_total;
  if (plan->num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&plan->num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (plan->num_sms <= 0) plan->num_sms = 148;
  }
  const int tiles = p.k_tiles * p.c_tiles;
  // one full wave: every CTA pays a fixed prologue (ring zeroing, TMEM allocation) and an epilogue of scalar reds over
  // the whole 9 x 64 x 64 slice, so more, shorter CTAs lose (84 us at 148 splits vs 100 us at 296 on 56x56x64)
  int splits = d.splits > 0 ? d.splits : (tiles >= plan->num_sms ? 1 : plan->num_sms / tiles);
  const int max_splits = p.num_m_tiles / 2 > 0 ? p.num_m_tiles / 2 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  p.splits = splits;
  const int smem = stages * stage_bytes + 1024 + 256;
  static int attr_smem = 0;
  if (smem > attr_smem) {
    if (cudaFuncSetAttribute(wgrad_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
      return -10;
    attr_smem = 227 * 1024;
  }
  dim3 grid(tiles, splits, 1);
  const long long need = static_cast<long long>(tiles) * splits * kWpGroups * 128 * 64 * 4;
  p.partials = (plan->stat_scratch != nullptr && need <= plan->stat_scratch_bytes) ? static_cast<float*>(plan->stat_scratch)
                                                                                      : nullptr;
  wgrad_patch_kernel<<<grid, kWpThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  if (p.partials != nullptr)
    wgrad_patch_reduce_kernel<<<dim3(tiles, kWpGroups), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -11;
}
