// Weight-gradient kernel for sm_100a:  dW[k, tap, c] += sum over pixels  dY[pixel, k] * X[pixel + tap, c].
//
// The reduction runs over pixels, which is the *strided* axis of both NHWC operands, so both UMMA
// operands are MN-major: a TMA box of 64 pixels x 64 channels lands in shared memory as 64 rows of
// 128 bytes (128B swizzle) and is consumed directly as a [K=64 pixels] x [MN=64 channels] atom.
//   M side: dY channels (k), 2 atoms = 128 rows.     N side: X channels (c), BN/64 atoms per tap.
// Each CTA owns one (k tile, c tile, tap GROUP) and one slice of the pixel range (split-K).  A group of G filter
// taps shares every dY tile: one dY load feeds G accumulators (G*BN TMEM columns), which divides the L2 -> SM
// operand traffic of the narrow-channel layers (C = 64, 128) that are otherwise bound by it.
// Partial results are added into the fp32 gradient with red.global.add.v4.f32, i.e. straight into
// the parameter's slot of the gradient bucket that the all-reduce kernel later reads.
//
// Replaces cuDNN wgrad / cuBLAS dW GEMMs.
#include <stdio.h>
#include <stdlib.h>

#include "../../common/ptx.cuh"
#include "../../common/tensormap.h"
#include "conv_common.cuh"

namespace dmpi {

constexpr int kWThreads = 256;
constexpr int kPix = 64;                    // pixels per pipeline stage (GEMM K)
constexpr int kAtomBytes = kPix * 128;      // 64 pixels x 64 bf16
constexpr int kWABytes = 2 * kAtomBytes;    // M = 128

constexpr int pow2_cols(int c) { return c <= 32 ? 32 : c <= 64 ? 64 : c <= 128 ? 128 : c <= 256 ? 256 : 512; }

// CG = 2: a CTA pair (tcgen05.mma.cta_group::2) covers TWO 128-filter tiles of the same (tap group, channel tile,
// pixel slice); each CTA loads its own dY tile and only half of the X tile, which both tensor cores share.
template <int BN, int G, int CG = 1>
struct WgradCfg {
  static constexpr int kBAtoms = BN / CG / 64;  // 64-channel X atoms this CTA loads per tap
  static constexpr int kBTapBytes = kBAtoms * kAtomBytes;
  static constexpr int kStageBytes = kWABytes + G * kBTapBytes;
  static constexpr int kStages = (220 * 1024) / kStageBytes > 8 ? 8 : (220 * 1024) / kStageBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
  static constexpr int kTmemCols = pow2_cols(G * BN);
  static_assert(G * BN <= 512, "accumulators exceed TMEM");
  static_assert(kStages >= 2, "pipeline too shallow");
  static_assert(CG == 1 || (G == 1 && BN >= 128), "CTA pairs: single-tap wide tiles only");
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int BN, int G, int CG>
__global__ void __launch_bounds__(kWThreads, 1) wgrad_kernel(const __grid_constant__ ConvKernelParams p) {
  using Cfg = WgradCfg<BN, G, CG>;
  constexpr int S = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * Cfg::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tfull = bars + 2 * S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // tile decode: blockIdx.x = ((tap_group * num_k_tiles + k_blk) * num_c_tiles + c_blk), blockIdx.y = split;
  // with CTA pairs k_blk is the fastest index so that the two CTAs of a cluster own filter tiles 2j and 2j+1
  const int rank = CG == 2 ? static_cast<int>(cluster_ctarank()) : 0;
  const int c_blk = CG == 2 ? (blockIdx.x / p.num_k_tiles) % p.num_c_tiles : blockIdx.x % p.num_c_tiles;
  const int k_blk = CG == 2 ? blockIdx.x % p.num_k_tiles : (blockIdx.x / p.num_c_tiles) % p.num_k_tiles;
  const int tap0 = (blockIdx.x / (p.num_c_tiles * p.num_k_tiles)) * G;
  const int ntap = (p.ntaps - tap0) < G ? (p.ntaps - tap0) : G;
  const int split = blockIdx.y;
  const int pt_total = p.num_m_tiles;
  const int pt_begin = static_cast<int>((static_cast<long long>(pt_total) * split) / p.splits);
  const int pt_end = static_cast<int>((static_cast<long long>(pt_total) * (split + 1)) / p.splits);
  const int num_it = pt_end - pt_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.a_map[p.tap_view[tap0]]);
    tma_prefetch_desc(&p.b_map);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], CG);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (CG == 2) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    else tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (num_it > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const uint32_t tx_bytes = kWABytes + static_cast<uint32_t>(ntap) * Cfg::kBTapBytes;
        int stage = 0;
        uint32_t phase = 0;
        for (int it = 0; it < num_it; ++it) {
          const int m_blk = pt_begin + it;
          const int tw = m_blk % p.tiles_w;
          const int th = (m_blk / p.tiles_w) % p.tiles_h;
          const int tn = m_blk / (p.tiles_w * p.tiles_h);
          const int w0 = tw * p.box_w, h0 = th * p.box_h, n0 = tn * p.box_n;
          mbar_wait_spin(&empty[stage], phase ^ 1u);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + kWABytes;
          if (CG == 2) {
            // both CTAs report to the leader's barrier; the leader registers the bytes of the pair
            const uint32_t lead_full = mapa_u32(smem_u32(&full[stage]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * tx_bytes);
            else mbar_arrive_cluster(lead_full);
#pragma unroll
            for (int j = 0; j < 2; ++j)
              tma_load_4d_pair(sa + j * kAtomBytes, &p.b_map, lead_full, k_blk * 128 + j * 64, w0, h0, n0);
            const CUtensorMap* xmap = &p.a_map[p.tap_view[tap0]];
#pragma unroll
            for (int j = 0; j < Cfg::kBAtoms; ++j)
              tma_load_4d_pair(sb + j * kAtomBytes, xmap, lead_full, c_blk * BN + (rank * Cfg::kBAtoms + j) * 64,
                               w0 + p.tap_dw[tap0], h0 + p.tap_dh[tap0], n0);
          } else {
            mbar_arrive_expect_tx(&full[stage], tx_bytes);
#pragma unroll
            for (int j = 0; j < 2; ++j)
              tma_load_4d(sa + j * kAtomBytes, &p.b_map, &full[stage], k_blk * 128 + j * 64, w0, h0, n0);
#pragma unroll
            for (int g = 0; g < G; ++g) {
              if (g < ntap) {
                const int t = tap0 + g;
                const CUtensorMap* xmap = &p.a_map[p.tap_view[t]];
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  tma_load_4d(sb + g * Cfg::kBTapBytes + j * kAtomBytes, xmap, &full[stage], c_blk * BN + j * 64,
                              w0 + p.tap_dw[t], h0 + p.tap_dh[t], n0);
              }
            }
          }
          if (++stage == S) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
      __syncwarp();
    } else if (warp == 1) {
      // warp-uniform loop, one elected lane issues (see elect_one() in ptx.cuh); split descriptors: one add per MMA
      const bool leader = elect_one() && rank == 0;  // of a pair only the leader CTA issues
      constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, BN, 1, 1);
      const uint32_t hi = umma_desc_hi(1024);
      // MN-major, 128B swizzle: LBO = stride between 64-channel atoms, SBO = stride between 8-pixel groups;
      // one MMA consumes 16 pixels = 2048 bytes (128 descriptor units) of every atom.
      const uint32_t a_lo_ring = umma_desc_lo(smem_u32(smem), kAtomBytes);
      constexpr uint32_t stage_step = static_cast<uint32_t>(Cfg::kStageBytes) >> 4;
      constexpr uint32_t b_off = static_cast<uint32_t>(kWABytes) >> 4;
      constexpr uint32_t tap_step = static_cast<uint32_t>(Cfg::kBTapBytes) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      uint32_t accum = 0;
      if (CG == 1 || rank == 0) {
        for (int it = 0; it < num_it; ++it) {
          mbar_wait_spin(&full[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_lo = a_lo_ring + static_cast<uint32_t>(stage) * stage_step;
          if (leader) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
              if (g < ntap) {
#pragma unroll
                for (int k = 0; k < kPix / 16; ++k) {
                  if (CG == 2)
                    umma_bf16_pair_lohi(tmem_base + g * BN, a_lo + k * 128, hi, a_lo + b_off + g * tap_step + k * 128,
                                        hi, idesc, k == 0 ? accum : 1u);
                  else
                    umma_bf16_lohi(tmem_base + g * BN, a_lo + k * 128, hi, a_lo + b_off + g * tap_step + k * 128, hi,
                                   idesc, k == 0 ? accum : 1u);
                }
              }
            }
            if (CG == 2) umma_commit_pair(&empty[stage], 0b11);
            else umma_commit(&empty[stage]);
          }
          accum = 1u;
          if (++stage == S) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (leader) {
          if (CG == 2) umma_commit_pair(tfull, 0b11);
          else umma_commit(tfull);
        }
      }
      __syncwarp();
    } else if (warp >= 4) {
      const int q = warp - 4;
      const int krow = k_blk * 128 + q * 32 + lane;
      mbar_wait(tfull, 0);
      tc_fence_after_sync();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
      // deterministic mode: the partial tile goes to scratch [tile][split][128][G*BN]; wgrad_reduce_kernel adds the
      // splits in order.  Otherwise red.add straight into the gradient.
      float* part = p.stat_scratch == nullptr ? nullptr
                        : p.stat_scratch + (static_cast<size_t>(blockIdx.x) * gridDim.y + blockIdx.y) * (128 * G * BN) +
                              static_cast<size_t>(q * 32 + lane) * (G * BN);
#pragma unroll 1
      for (int g = 0; g < (part ? G : ntap); ++g) {
        float* dst_row = p.dw + static_cast<long long>(krow) * p.w_ld + p.tap_woff[tap0 + (g < ntap ? g : 0)] + c_blk * BN;
#pragma unroll 1
        for (int chunk = 0; chunk < BN / 32; ++chunk) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + g * BN + chunk * 32, v);
          tmem_ld_wait();
          if (part != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<uint4*>(part + g * BN + chunk * 32 + j) =
                  g < ntap ? make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]) : make_uint4(0, 0, 0, 0);
          } else if (krow < p.k_out) {
            const int c0 = c_blk * BN + chunk * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (c0 + j < p.c_total)
                red_add_v4(dst_row + chunk * 32 + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                           __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
            }
          }
        }
      }
    }
  } else if (warp >= 4 && p.stat_scratch != nullptr) {
    // no pixel tiles for this split: its partial must still read as zero
    float* part = p.stat_scratch + (static_cast<size_t>(blockIdx.x) * gridDim.y + blockIdx.y) * (128 * G * BN) +
                  static_cast<size_t>((warp - 4) * 32 + lane) * (G * BN);
    for (int j = 0; j < G * BN; j += 4) *reinterpret_cast<uint4*>(part + j) = make_uint4(0, 0, 0, 0);
  }
  tc_fence_before_sync();
  if (CG == 2) {
    cluster_sync_all();
    if (warp == 2) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
  } else {
    __syncthreads();
    if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// Deterministic split-K: adds the partial tiles of wgrad_kernel in split order and accumulates onto the gradient.
// grid = (tiles, 16): a block owns 8 accumulator rows of one tile.
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const __grid_constant__ ConvKernelParams p, int G, int BN,
                                                           int CG, int splits) {
  const int tile = blockIdx.x;
  const int c_blk = CG == 2 ? (tile / p.num_k_tiles) % p.num_c_tiles : tile % p.num_c_tiles;
  const int k_blk = CG == 2 ? tile % p.num_k_tiles : (tile / p.num_c_tiles) % p.num_k_tiles;
  const int tap0 = (tile / (p.num_c_tiles * p.num_k_tiles)) * G;
  const int ntap = (p.ntaps - tap0) < G ? (p.ntaps - tap0) : G;
  const int cols = G * BN, q4 = cols >> 2;
  const float* base = p.stat_scratch + static_cast<size_t>(tile) * splits * (128 * cols);
  for (int i = threadIdx.x; i < 8 * q4; i += blockDim.x) {
    const int row = blockIdx.y * 8 + i / q4, col = (i % q4) * 4;
    const int g = col / BN, cc = col - g * BN;
    const int krow = k_blk * 128 + row, c = c_blk * BN + cc;
    if (g >= ntap || krow >= p.k_out || c >= p.c_total) continue;
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* src = base + static_cast<size_t>(row) * cols + col;
    int sidx = 0;
    for (; sidx + 4 <= splits; sidx += 4) {  // four splits in flight, added in split order
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = __ldcg(reinterpret_cast<const float4*>(src + static_cast<size_t>(sidx + k) * (128 * cols)));
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        t.x += v[k].x; t.y += v[k].y; t.z += v[k].z; t.w += v[k].w;
      }
    }
    for (; sidx < splits; ++sidx) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + static_cast<size_t>(sidx) * (128 * cols)));
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    float4* dst = reinterpret_cast<float4*>(p.dw + static_cast<long long>(krow) * p.w_ld + p.tap_woff[tap0 + g] + c);
    float4 d = *dst;
    d.x += t.x; d.y += t.y; d.z += t.z; d.w += t.w;
    *dst = d;
  }
}

template <int BN, int G, int CG = 1>
static int launch_wgrad(const ConvKernelParams& kp, cudaStream_t stream) {
  using Cfg = WgradCfg<BN, G, CG>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e =
        cudaFuncSetAttribute(wgrad_kernel<BN, G, CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return -10;
    attr_set = true;
  }
  const int groups = (kp.ntaps + G - 1) / G;
  dim3 grid(groups * kp.num_k_tiles * kp.num_c_tiles, kp.splits, 1);
  if (CG == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;  // k_blk is the fastest index and num_k_tiles is even: clusters pair filter tiles 2j, 2j+1
    cfg.blockDim = dim3(kWThreads, 1, 1);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, wgrad_kernel<BN, G, CG>, kp) != cudaSuccess) return -11;
  } else {
    wgrad_kernel<BN, G, CG><<<grid, kWThreads, Cfg::kSmemBytes, stream>>>(kp);
  }
  if (kp.stat_scratch != nullptr)
    wgrad_reduce_kernel<<<dim3(grid.x, 16), 256, 0, stream>>>(kp, G, BN, CG, kp.splits);
  return cudaGetLastError() == cudaSuccess ? 0 : -11;
}

}  // namespace dmpi

using namespace dmpi;

// M-side operand (dY) is described by desc.o_view[0]; the tapped N-side operand (X) by desc.a_view[] / taps.
extern "C" int dmpi_wgrad_run(dmpi_conv_plan* plan, const void* dy, const void* x, float* dw, void* stream) {
  const dmpi_conv_desc& d = plan->desc;
  ConvKernelParams& kp = plan->params;
  if (kp.rows != kPix) return -1;
  if (d.bn != 64 && d.bn != 128 && d.bn != 256) return -2;
  if (!plan->maps_valid || plan->cached_a != x || plan->cached_b != dy) {
    int box[4] = {64, d.box_w, d.box_h, d.box_n};
    for (int i = 0; i < d.a_nviews; ++i) {
      const uint8_t* ptr = static_cast<const uint8_t*>(x) + d.a_view[i].offset * 2;
      if (encode_bf16_map(&kp.a_map[i], ptr, 4, d.a_view[i].dims, d.a_view[i].strides, box)) return -4;
    }
    for (int i = d.a_nviews; i < DMPI_MAX_MAPS; ++i) kp.a_map[i] = kp.a_map[0];
    const uint8_t* yptr = static_cast<const uint8_t*>(dy) + d.o_view[0].offset * 2;
    if (encode_bf16_map(&kp.b_map, yptr, 4, d.o_view[0].dims, d.o_view[0].strides, box)) return -5;
    plan->cached_a = x;
    plan->cached_b = dy;
    plan->maps_valid = true;
  }
  kp.dw = dw;
  kp.num_k_tiles = (d.k_out + 127) / 128;
  kp.num_c_tiles = (d.c_total + d.bn - 1) / d.bn;
  if (plan->num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&plan->num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (plan->num_sms <= 0) plan->num_sms = 148;
  }
  // taps per CTA: narrow N tiles share the dY tile between several taps
  int G = 1;
  if (kp.ntaps > 1) {
    if (d.bn == 64) G = (kp.ntaps % 3 == 0) ? 3 : 4;
    else if (d.bn == 128) G = 3;
  }
  const int groups = (kp.ntaps + G - 1) / G;
  int splits = d.splits;
  if (splits <= 0) {
    const int tiles = groups * kp.num_k_tiles * kp.num_c_tiles;
    // ONE full wave of CTAs: every CTA pays a fixed prologue (TMEM allocation, pipeline fill) and an epilogue that adds
    // a whole 128 x BN fp32 tile into the gradient, so fewer, longer CTAs win (14x14 256->1024: 51 us at two waves,
    // 42 us at one, 60 us at three), and never a partial extra wave.
    static const int waves = getenv("DMPI_WGRAD_WAVES") ? atoi(getenv("DMPI_WGRAD_WAVES")) : 1;
    splits = (waves * plan->num_sms) / tiles;
    const int max_splits = kp.num_m_tiles / 4 > 0 ? kp.num_m_tiles / 4 : 1;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  if (splits > kp.num_m_tiles) splits = kp.num_m_tiles > 0 ? kp.num_m_tiles : 1;
  kp.splits = splits;
  {
    // deterministic split-K needs one partial tile per CTA in the scratch the caller attached to the plan
    const long long need = static_cast<long long>(groups) * kp.num_k_tiles * kp.num_c_tiles * splits * 128LL * G * d.bn * 4;
    const bool aligned = d.c_total % 4 == 0 && d.w_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0;
    kp.stat_scratch = (plan->stat_scratch != nullptr && need <= plan->stat_scratch_bytes && aligned)
                          ? static_cast<float*>(plan->stat_scratch) : nullptr;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (d.bn == 64) {
    if (G == 3) return launch_wgrad<64, 3>(kp, s);
    if (G == 4) return launch_wgrad<64, 4>(kp, s);
    return launch_wgrad<64, 1>(kp, s);
  }
  // CTA pairs for the single-tap wide tiles when the filter tiles pair up (L2 -> SM traffic per FLOP drops by a third)
  static const int cg_env = getenv("DMPI_WGRAD_CG") ? atoi(getenv("DMPI_WGRAD_CG")) : 2;
  const bool pair = cg_env == 2 && G == 1 && kp.num_k_tiles % 2 == 0;
  if (d.bn == 128) {
    if (G == 3) return launch_wgrad<128, 3>(kp, s);
    return pair ? launch_wgrad<128, 1, 2>(kp, s) : launch_wgrad<128, 1>(kp, s);
  }
  return pair ? launch_wgrad<256, 1, 2>(kp, s) : launch_wgrad<256, 1>(kp, s);
}
