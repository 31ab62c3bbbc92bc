// Persistent, warp-specialised implicit-GEMM kernel for sm_100a.
//
//   D[pixels, n] = sum over taps t, channel blocks cb of  A_t[pixels, cb*64 .. +64] * B[n, koff_t + cb*64 .. +64]^T
//
// * A tiles (128 output pixels x 64 channels, bf16) are fetched by TMA from a rank-4 NHWC view; the
//   filter tap is a coordinate offset and the halo is TMA's out-of-bounds zero fill.  Stride-2
//   convolutions use one strided view per input parity class, so every tap is still a dense box.
// * B tiles (BN filters x 64 channels) come from the K-major bf16 weight shadow.
// * tcgen05.mma (M=128, N=BN, K=16; cta_group::1, or cta_group::2 with 256-row tiles, see IgemmCfg) accumulates in
//   TMEM; two accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
// * Patch mode (stride-1 3x3): one shared-memory patch of input pixels per (tile, channel block) serves all nine
//   taps through row-shifted descriptors; the filter taps stream through a second ring or stay resident.
// * The epilogue is a two-stage pipeline of its own, in sub-tiles of min(BN, 128) columns:
//     converter warps   read TMEM (tcgen05.ld), add bias / a residual tensor / ReLU, round to bf16 and stage the
//                       sub-tile in 128B-swizzled shared memory (two buffers);
//     store/stat warps  issue the TMA store into the (possibly strided / channel-sliced) output view and accumulate
//                       the per-channel sum and sum of squares of the stored values for BatchNorm (published with
//                       atomics, or, with a scratch attached, as per-CTA partials that the last CTA adds in order).
//   The two groups hand buffers over through mbarriers only, so conversion of sub-tile i+1 overlaps the store and
//   the statistics of sub-tile i.  (ncu on the K=64 1x1 layers showed the old single-group epilogue, not the main
//   loop, pacing the kernel: ~740 dependent warp instructions per tile on two warps per scheduler.)
//
// Roles: warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator, warps 4..11 = converters,
//        warps 12..15 = store + statistics.
//
// Replaces cuDNN fprop/dgrad and cuBLAS GEMM behind ref:pytorch/unet/model.py:8-14,37-38,68 and
// torchvision's resnet conv/fc.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>

#include "../../common/ptx.cuh"
#include "../../common/tensormap.h"
#include "conv_common.cuh"

// Ablation experiments (skip the loads / MMAs / stores / statistics of the kernel to see what paces it) exist only in
// builds made with -DDMPI_ABLATE (DMPI_NVCC_EXTRA="-DDMPI_ABLATE" python build.py); the shipped kernel has no such
// switch, so no environment variable can turn it into a no-op.
#ifdef DMPI_ABLATE
#define DMPI_DBG(p, bit) (((p).dbg & (bit)) != 0)
#else
#define DMPI_DBG(p, bit) false
#endif

namespace dmpi {

constexpr int kThreads = 512;       // 4 control warps + 8 converter warps + 4 store/statistics warps
constexpr int kCvtThreads = 256;
constexpr int kCvtWarps = 8;
constexpr int kStatThreads = 128;
constexpr int kStatWarps = 4;
constexpr int kTileM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kABytes = kTileM * kBlockK * 2;
constexpr int kDetInFlight = 4;  // partial-sum rows a thread of the tail reduction keeps in flight
constexpr int kTailCtas = 8;     // CTAs that share the deterministic statistics tail
constexpr int kStatHdr = 64;  // floats reserved at the head of the statistics scratch (ticket counter)
constexpr int kSkHeaderBytes = 4096;  // split-K scratch: one ticket per tile (<= 1024 tiles) ahead of the partial tiles

// CG = 1: one CTA per tile.  CG = 2: a CTA pair (cluster of two SMs of one TPC) works on a 256-row tile with
// tcgen05.mma.cta_group::2 — each CTA loads its own 128 A rows and only HALF of the B tile (the tensor cores of both
// SMs read both halves), which cuts the L2 -> SM operand traffic per FLOP by a third (BN=256) that bounds the
// deep layers (~6300 B/clk chip-wide).
template <int BN, int CG>
struct IgemmCfg {
  // Pipeline depth is what hides the ~1000-cycle L2/HBM latency of the TMA loads: one stage is only 128 (BN=64),
  // 256 (BN=128) or 512 (BN=256) tensor-core cycles of work, so the narrow tiles need the most stages.
  static constexpr int kStages = CG == 2 ? (BN == 256 ? 5 : (BN == 128 ? 6 : 9)) : (BN == 256 ? 3 : (BN == 128 ? 5 : 8));
  static constexpr int kBRows = BN / CG;              // B rows (filters) this CTA loads per K block
  static constexpr int kBBytes = kBRows * kBlockK * 2;
  static constexpr int kSubN = BN < 128 ? BN : 128;   // columns per staged sub-tile
  static constexpr int kSubs = BN / kSubN;            // sub-tiles per tile: 1, 1, 2
  static constexpr int kOutBytes = kTileM * kSubN * 2;
  static constexpr int kOutBufs = 2;
  static constexpr int kSmemBytes =
      kStages * (kABytes + kBBytes) + kOutBufs * kOutBytes + 1024 /*align slack*/ + 512 /*barriers*/ + 128 /*tap table*/;
  static constexpr int kRingBytes = kStages * (kABytes + kBBytes);  // patch mode re-carves this region (2 rings)
  static constexpr int kTmemCols = 2 * BN;
  // (BN = 64 with pairs: each CTA holds 32 filters of every tile - for narrow layers whose filter is too large to stay
  // resident, e.g. the UNet decoder's 192 -> 64 at 512x512, where the streamed filter taps were the L2 -> SM bottleneck)
};

// Walks the persistent tile sequence without a division per tile: the grid step is decomposed once into
// (n_blk, tile_w, tile_h, tile_n) digits and added with carries.  With CTA pairs the sequence runs over pair tiles
// (n_blk, m_pair) per cluster and the CTA of rank r owns M tile 2 * m_pair + r (possibly one past the end: such a
// tile only loads zero fill and stores nothing).
struct TileWalk {
  int n_blk, tw, th, tn;
  int s_nb, s_tw, s_th, s_tn;
  int remaining;
  int cg;
  __device__ __forceinline__ void init(const ConvKernelParams& p, int cg_, int rank) {
    cg = cg_;
    const int m_units = (p.num_m_tiles + cg - 1) / cg;  // M tiles or M-tile pairs
    const int total = m_units * p.num_n_tiles;
    // split-K: the grid holds `splits` CTAs per tile, CTA b takes tile b % total (exactly one work item each)
    const bool sk = p.splits > 1;
    const int first = sk ? static_cast<int>(blockIdx.x) % total : static_cast<int>(blockIdx.x) / cg;
    const int step = sk ? total : static_cast<int>(gridDim.x) / cg;
    remaining = first < total ? (total - first + step - 1) / step : 0;
    n_blk = first % p.num_n_tiles;
    int m = (first / p.num_n_tiles) * cg + rank;
    tw = m % p.tiles_w;
    m /= p.tiles_w;
    th = m % p.tiles_h;
    tn = m / p.tiles_h;
    s_nb = step % p.num_n_tiles;
    int sm = (step / p.num_n_tiles) * cg;
    s_tw = sm % p.tiles_w;
    sm /= p.tiles_w;
    s_th = sm % p.tiles_h;
    s_tn = sm / p.tiles_h;
  }
  __device__ __forceinline__ void next(const ConvKernelParams& p) {
    --remaining;
    n_blk += s_nb;
    tw += s_tw;
    if (n_blk >= p.num_n_tiles) {
      n_blk -= p.num_n_tiles;
      tw += cg;
    }
    th += s_th;
    tn += s_tn;
    while (tw >= p.tiles_w) {
      tw -= p.tiles_w;
      ++th;
    }
    while (th >= p.tiles_h) {
      th -= p.tiles_h;
      ++tn;
    }
  }
};

// SK: the split-K variant (a separate instantiation so that its extra registers do not touch the main kernels).
template <int BN, int CG, bool SK = false>
__global__ void __launch_bounds__(kThreads, 1) igemm_kernel(const __grid_constant__ ConvKernelParams p) {
  using Cfg = IgemmCfg<BN, CG>;
  constexpr int S = Cfg::kStages;
  constexpr int SN = Cfg::kSubN;
  constexpr int kSubs = Cfg::kSubs;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + S * kABytes;
  uint8_t* smem_o = smem_b + S * Cfg::kBBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_o + Cfg::kOutBufs * Cfg::kOutBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tfull = bars + 2 * S;
  uint64_t* tempty = bars + 2 * S + 2;
  uint64_t* ofull = bars + 2 * S + 4;   // converters -> store/stat warps: sub-tile staged
  uint64_t* ofree = bars + 2 * S + 6;   // store/stat warps -> converters: buffer may be overwritten
  uint64_t* aload = bars + 2 * S + 8;   // addend sub-tile landed in the staging buffer (replaces ofree when used)
  // patch mode: separate rings for input patches (<= kMaxPA stages) and filter-tap tiles (<= kMaxPB stages)
  constexpr int kMaxPA = 4, kMaxPB = 12;
  uint64_t* pa_full = bars + 2 * S + 10;
  uint64_t* pa_empty = pa_full + kMaxPA;
  uint64_t* pb_full = pa_empty + kMaxPA;
  uint64_t* pb_empty = pb_full + kMaxPB;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pb_empty + kMaxPB);
  uint32_t* last_flag = reinterpret_cast<uint32_t*>(smem_o + Cfg::kOutBufs * Cfg::kOutBytes + 512 + 124);  // see end
  static_assert((2 * S + 10 + 2 * kMaxPA + 2 * kMaxPB) * 8 + 8 <= 512, "barrier region too small");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = CG == 2 ? static_cast<int>(cluster_ctarank()) : 0;  // rank 0 of a pair is the MMA leader

  if (warp == 0 && lane == 0) {
    *last_flag = 0u;
    for (int i = 0; i < DMPI_MAX_MAPS; ++i) tma_prefetch_desc(&p.a_map[i]);
    tma_prefetch_desc(&p.b_map);
    tma_prefetch_desc(&p.o_map[0]);
    if (p.has_add) tma_prefetch_desc(&p.add_map);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], CG);  // pair: the leader's expect_tx arrival + the peer's plain arrival
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], CG * kCvtWarps);  // pair: the converters of both CTAs release the leader's barrier
      mbar_init(&ofull[i], kCvtWarps);
      mbar_init(&ofree[i], kStatWarps);
      mbar_init(&aload[i], 1);
    }
    if (p.patch) {
      for (int i = 0; i < kMaxPA; ++i) {
        mbar_init(&pa_full[i], CG);  // pair: the leader's expect_tx arrival + the peer's plain arrival
        mbar_init(&pa_empty[i], 1);
      }
      for (int i = 0; i < kMaxPB; ++i) {
        mbar_init(&pb_full[i], CG);
        mbar_init(&pb_empty[i], 1);
      }
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if (CG == 2) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    else tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all();  // the peer's barriers must be initialised before anything signals them
  else __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  const int num_kb = p.ntaps * p.cblocks;
  // split-K slice of this CTA (the whole range without it)
  const bool splitk = SK && p.splits > 1;
  const int sk_units = p.num_m_tiles * p.num_n_tiles;
  const int sk_split = splitk ? static_cast<int>(blockIdx.x) / sk_units : 0;
  const int sk_unit = splitk ? static_cast<int>(blockIdx.x) % sk_units : 0;
  const int kb_per = splitk ? (num_kb + p.splits - 1) / p.splits : num_kb;
  const int kb_lo = sk_split * kb_per;
  const int kb_hi = kb_lo + kb_per < num_kb ? kb_lo + kb_per : num_kb;
  uint32_t* sk_flag = reinterpret_cast<uint32_t*>(smem_o + Cfg::kOutBufs * Cfg::kOutBytes + 512 + 120);  // 1: this CTA finalises

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && p.patch) {
      // Patch mode: one input patch per (tile, channel block) into the A ring, one filter tile per tap into the B ring
      // (loaded once and kept when the whole filter fits: b_resident).
      uint8_t* ring_a = smem;
      uint8_t* ring_b = smem + p.pa_stages * p.pa_stride;
      int sa = 0, sb = 0;
      uint32_t pha = 0, phb = 0;
      bool first = true;
      TileWalk tk;
      for (tk.init(p, CG, rank); tk.remaining > 0; tk.next(p)) {
        const int w0 = tk.tw * p.box_w, h0 = tk.th * p.box_h, n0 = tk.tn * p.box_n;
        for (int cb = 0; cb < p.cblocks; ++cb) {
          mbar_wait_spin(&pa_empty[sa], pha ^ 1u);
          if (CG == 1) {
            mbar_arrive_expect_tx(&pa_full[sa], static_cast<uint32_t>(p.pa_bytes));
            tma_load_4d(ring_a + sa * p.pa_stride, &p.a_map[0], &pa_full[sa], cb * kBlockK, w0 - p.halo, h0 - p.halo, n0);
          } else {
            // CTA pair: each CTA brings the patch of its own M tile; both report to the leader's barrier
            const uint32_t lead = mapa_u32(smem_u32(&pa_full[sa]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&pa_full[sa], 2u * static_cast<uint32_t>(p.pa_bytes));
            else mbar_arrive_cluster(lead);
            tma_load_4d_pair(ring_a + sa * p.pa_stride, &p.a_map[0], lead, cb * kBlockK, w0 - p.halo, h0 - p.halo, n0);
          }
          if (++sa == p.pa_stages) {
            sa = 0;
            pha ^= 1u;
          }
          if (p.b_resident && !first) continue;
          for (int t = 0; t < p.ntaps; ++t) {
            const int slot = p.b_resident ? cb * p.ntaps + t : sb;
            if (!p.b_resident) mbar_wait_spin(&pb_empty[sb], phb ^ 1u);
            if (CG == 1) {
              mbar_arrive_expect_tx(&pb_full[slot], Cfg::kBBytes);
              tma_load_2d(ring_b + slot * Cfg::kBBytes, &p.b_map, &pb_full[slot], p.tap_koff[t] + cb * kBlockK,
                          tk.n_blk * BN);
            } else {
              // each CTA of the pair loads its half of the filter-tap tile
              const uint32_t lead = mapa_u32(smem_u32(&pb_full[slot]), 0);
              if (rank == 0) mbar_arrive_expect_tx(&pb_full[slot], 2u * Cfg::kBBytes);
              else mbar_arrive_cluster(lead);
              tma_load_2d_pair(ring_b + slot * Cfg::kBBytes, &p.b_map, lead, p.tap_koff[t] + cb * kBlockK,
                               tk.n_blk * BN + rank * Cfg::kBRows);
            }
            if (!p.b_resident && ++sb == p.pb_stages) {
              sb = 0;
              phb ^= 1u;
            }
          }
        }
        first = false;
      }
    } else if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = static_cast<uint32_t>(p.rows) * 128u + Cfg::kBBytes;
      TileWalk tk;
      for (tk.init(p, CG, rank); tk.remaining > 0; tk.next(p)) {
        const int w0 = tk.tw * p.box_w, h0 = tk.th * p.box_h, n0 = tk.tn * p.box_n;
        int t = kb_lo / p.cblocks, cb = kb_lo - t * p.cblocks;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          const CUtensorMap* amap = &p.a_map[p.tap_view[t]];
          const int cw = w0 + p.tap_dw[t], ch = h0 + p.tap_dh[t], koff = p.tap_koff[t];
          mbar_wait_spin(&empty[stage], phase ^ 1u);
          if (DMPI_DBG(p, 4)) {  // experiment: no loads, barriers only (MMA runs on stale shared memory)
            if (CG == 1 || rank == 0) mbar_arrive(&full[stage]);
            else mbar_arrive_cluster(mapa_u32(smem_u32(&full[stage]), 0));
          } else if (CG == 1) {
            mbar_arrive_expect_tx(&full[stage], tx_bytes);
            tma_load_4d(smem_a + stage * kABytes, amap, &full[stage], cb * kBlockK, cw, ch, n0);
            tma_load_2d(smem_b + stage * Cfg::kBBytes, &p.b_map, &full[stage], koff + cb * kBlockK, tk.n_blk * BN);
          } else {
            // both CTAs report to the leader's barrier; the leader registers the bytes of the pair
            const uint32_t lead_full = mapa_u32(smem_u32(&full[stage]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * tx_bytes);
            else mbar_arrive_cluster(lead_full);
            tma_load_4d_pair(smem_a + stage * kABytes, amap, lead_full, cb * kBlockK, cw, ch, n0);
            tma_load_2d_pair(smem_b + stage * Cfg::kBBytes, &p.b_map, lead_full, koff + cb * kBlockK,
                             tk.n_blk * BN + rank * Cfg::kBRows);
          }
          if (++stage == S) {
            stage = 0;
            phase ^= 1u;
          }
          if (++cb == p.cblocks) {
            cb = 0;
            ++t;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // The whole warp runs the loops (warp-uniform control flow, every lane polls the barriers); one elected lane
    // issues the tcgen05 instructions.  Descriptors are split into a constant high word and a low word that advances
    // by (bytes >> 4), so a K step or a tap shift is one 32-bit add.
    const bool leader = elect_one();
    const uint32_t hi = umma_desc_hi(1024);
    if (p.patch) {
      // (CTA pairs: only the leader CTA issues; both CTAs' patches sit at the same shared-memory offsets, each CTA's
      // filter-tap slot holds its half of the tile)
      constexpr uint32_t idesc = umma_idesc_bf16(kTileM * CG, BN, 0, 0);
      const bool issuer = leader && rank == 0;
      auto mma4 = [&](uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t accum0) {
        if (CG == 2) {
          umma_bf16_pair_lohi(d, a_lo, hi, b_lo, hi, idesc, accum0);
          umma_bf16_pair_lohi(d, a_lo + 2, hi, b_lo + 2, hi, idesc, 1u);
          umma_bf16_pair_lohi(d, a_lo + 4, hi, b_lo + 4, hi, idesc, 1u);
          umma_bf16_pair_lohi(d, a_lo + 6, hi, b_lo + 6, hi, idesc, 1u);
        } else {
          umma_bf16_lohi(d, a_lo, hi, b_lo, hi, idesc, accum0);
          umma_bf16_lohi(d, a_lo + 2, hi, b_lo + 2, hi, idesc, 1u);
          umma_bf16_lohi(d, a_lo + 4, hi, b_lo + 4, hi, idesc, 1u);
          umma_bf16_lohi(d, a_lo + 6, hi, b_lo + 6, hi, idesc, 1u);
        }
      };
      auto commit = [&](uint64_t* bar) {
        if (CG == 2) umma_commit_pair(bar, 0b11);
        else umma_commit(bar);
      };
      const uint32_t ring_a = smem_u32(smem);
      const uint32_t ring_b = ring_a + p.pa_stages * p.pa_stride;
      const uint32_t a_lo_ring = umma_desc_lo(ring_a, 0);
      const uint32_t b_lo0 = umma_desc_lo(ring_b, 0);
      const uint32_t pa_step = static_cast<uint32_t>(p.pa_stride) >> 4;
      constexpr uint32_t pb_step = static_cast<uint32_t>(Cfg::kBBytes) >> 4;
      // per-tap row shift inside the patch, in descriptor units (128 B >> 4 = 8 per pixel row)
      uint32_t* tap_shift = reinterpret_cast<uint32_t*>(smem_o + Cfg::kOutBufs * Cfg::kOutBytes + 512);  // 32 words
      if (lane < p.ntaps)
        tap_shift[lane] = static_cast<uint32_t>((p.tap_dh[lane] + p.halo) * p.pw + p.tap_dw[lane] + p.halo) * 8u;
      __syncwarp();
      const int ntaps = p.ntaps, cblocks = p.cblocks, pb_stages = p.pb_stages, pa_stages = p.pa_stages;
      const bool resident = p.b_resident != 0;
      // 3x3 filters (every caller today): tap shifts live in registers and the tap loop is fully unrolled, so the
      // steady state is ~3 instructions per MMA
      uint32_t ts[9];
#pragma unroll
      for (int t = 0; t < 9; ++t) ts[t] = t < ntaps ? tap_shift[t] : 0u;
      int sa = 0, sb = 0;
      uint32_t pha = 0, phb = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      bool first = true;
      TileWalk tk;
      tk.init(p, CG, rank);
      if (rank != 0) tk.remaining = 0;  // the peer CTA's MMA warp issues nothing (its barriers live in the leader)
      for (; tk.remaining > 0; --tk.remaining) {
        mbar_wait_spin(&tempty[acc], acc_phase ^ 1u);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        uint32_t accum = 0;
        for (int cb = 0; cb < cblocks; ++cb) {
          mbar_wait_spin(&pa_full[sa], pha);
          tc_fence_after_sync();
          const uint32_t a_lo0 = a_lo_ring + static_cast<uint32_t>(sa) * pa_step;
          if (ntaps == 9 && resident && !first) {
            // whole filter resident: no barrier traffic inside the tile
            const uint32_t b_cb = b_lo0 + static_cast<uint32_t>(cb * 9) * pb_step;
            if (issuer) {
#pragma unroll
              for (int t = 0; t < 9; ++t) {
                const uint32_t a_lo = a_lo0 + ts[t];
                const uint32_t b_lo = b_cb + static_cast<uint32_t>(t) * pb_step;
                mma4(d_tmem, a_lo, b_lo, t == 0 ? accum : 1u);
              }
            }
            accum = 1u;
          } else if (ntaps == 9 && !resident) {
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              mbar_wait_spin(&pb_full[sb], phb);
              tc_fence_after_sync();
              if (issuer) {
                const uint32_t a_lo = a_lo0 + ts[t];
                const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(sb) * pb_step;
                mma4(d_tmem, a_lo, b_lo, accum);
                commit(&pb_empty[sb]);
              }
              accum = 1u;
              if (++sb == pb_stages) {
                sb = 0;
                phb ^= 1u;
              }
            }
          } else {
            // generic path: any tap count, first tile of the resident mode
#pragma unroll 1
            for (int t = 0; t < ntaps; ++t) {
              const int slot = resident ? cb * ntaps + t : sb;
              if (!resident || first) {
                mbar_wait_spin(&pb_full[slot], resident ? 0u : phb);
                tc_fence_after_sync();
              }
              const uint32_t a_lo = a_lo0 + tap_shift[t];
              const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(slot) * pb_step;
              if (issuer) {
                mma4(d_tmem, a_lo, b_lo, accum);
                if (!resident) commit(&pb_empty[sb]);
              }
              accum = 1u;
              if (!resident && ++sb == pb_stages) {
                sb = 0;
                phb ^= 1u;
              }
            }
          }
          if (issuer) commit(&pa_empty[sa]);
          if (++sa == pa_stages) {
            sa = 0;
            pha ^= 1u;
          }
        }
        first = false;
        if (issuer) commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    } else if (rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kTileM * CG, BN, 0, 0);
      const uint32_t a_lo_ring = umma_desc_lo(smem_u32(smem_a), 0);
      const uint32_t b_lo_ring = umma_desc_lo(smem_u32(smem_b), 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      TileWalk tk;
      for (tk.init(p, CG, rank); tk.remaining > 0; --tk.remaining) {  // only the count matters here
        mbar_wait_spin(&tempty[acc], acc_phase ^ 1u);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        uint32_t accum = 0;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait_spin(&full[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_lo = a_lo_ring + static_cast<uint32_t>(stage) * (kABytes >> 4);
          const uint32_t b_lo = b_lo_ring + static_cast<uint32_t>(stage) * (Cfg::kBBytes >> 4);
          if (leader) {
            if (!DMPI_DBG(p, 8)) {  // (experiment bit 8: no MMAs, commits only)
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {
                if (CG == 2) umma_bf16_pair_lohi(d_tmem, a_lo + 2 * k, hi, b_lo + 2 * k, hi, idesc, k == 0 ? accum : 1u);
                else umma_bf16_lohi(d_tmem, a_lo + 2 * k, hi, b_lo + 2 * k, hi, idesc, k == 0 ? accum : 1u);
              }
            }
            // frees the stage in both CTAs of a pair (each producer waits on its own barrier)
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
          if (CG == 2) umma_commit_pair(&tfull[acc], 0b11);
          else umma_commit(&tfull[acc]);
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
    __syncwarp();
  }
  else if (warp >= 4 && warp < 4 + kCvtWarps) {
    // ------------------------------------------------------------------ converters (8 warps)
    // TMEM lane quadrant q = warp % 4 (hardware rule); the two warps of a quadrant split the columns of a sub-tile.
    constexpr int CHs = SN / 64;  // 32-column chunks per warp and sub-tile: 1 (BN=64) or 2
    const int q = warp & 3;
    const int hf = (warp - 4) >> 2;
    const int row = q * 32 + lane;  // accumulator row (TMEM lane)
    // (wi, hi, ni): pixel of this row inside the tile;  srow: its row in the dense staged tile
    int wi, hi, ni, srow;
    bool row_ok;
    if (p.patch) {
      // accumulator rows run over the padded patch width: row = hi * pw + wi; columns >= box_w and rows >= box_h are
      // halo positions whose results are discarded (they take the unused tail rows of the staged tile, zeroed)
      wi = row % p.pw;
      hi = row / p.pw;
      ni = 0;
      row_ok = wi < p.box_w && hi < p.box_h;
      const int below = (hi < p.box_h ? hi : p.box_h) * p.box_w + (hi < p.box_h ? (wi < p.box_w ? wi : p.box_w) : 0);
      srow = row_ok ? hi * p.box_w + wi : p.box_w * p.box_h + (row - below);
    } else {
      wi = row % p.box_w;
      hi = (row / p.box_w) % p.box_h;
      ni = row / (p.box_w * p.box_h);
      row_ok = row < p.rows;
      srow = row;
    }
    const uint32_t obase = smem_u32(smem_o);
    const bool has_add = p.has_add != 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t u = 0;  // staged sub-tile counter: buffer u & 1, barrier phase (u >> 1) & 1
    const uint32_t lead_tempty = CG == 2 ? mapa_u32(smem_u32(&tempty[0]), 0) : 0u;
    TileWalk tk;
    for (tk.init(p, CG, rank); tk.remaining > 0; tk.next(p)) {
      const int w0 = tk.tw * p.box_w, h0 = tk.th * p.box_h, n0 = tk.tn * p.box_n;
      const bool valid = row_ok && (w0 + wi) < p.grid_w && (h0 + hi) < p.grid_h && (n0 + ni) < p.grid_n;

      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after_sync();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BN);
      const size_t sk_tile = static_cast<size_t>(kTileM) * BN;  // floats per partial tile
      if (splitk) {
        // split-K, phase A: park this CTA's fp32 partial tile (each thread its accumulator row), hand the accumulator
        // back, then draw the tile's ticket; only the CTA that draws the last one carries on into the epilogue
        float* mine = p.sk_part + (static_cast<size_t>(sk_split) * sk_units + sk_unit) * sk_tile + static_cast<size_t>(row) * BN;
#pragma unroll
        for (int sub = 0; sub < kSubs; ++sub) {
#pragma unroll
          for (int ci = 0; ci < CHs; ++ci) {
            const int col_t = sub * SN + (hf * CHs + ci) * 32;
            uint32_t v[32];
            tmem_ld_32x32(t_addr + col_t, v);
            tmem_ld_wait();
            float4* dst = reinterpret_cast<float4*>(mine + col_t);
#pragma unroll
            for (int i = 0; i < 8; ++i)
              dst[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                                   __uint_as_float(v[4 * i + 3]));
          }
        }
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        __threadfence();
        named_bar_sync(3, kCvtThreads);
        if (threadIdx.x == 4 * 32) {
          const uint32_t t = atomicAdd(&p.sk_tickets[sk_unit], 1u);
          const bool last = t == static_cast<uint32_t>(p.splits - 1);
          if (last) {
            __threadfence();
            p.sk_tickets[sk_unit] = 0u;  // re-armed for the next launch (every split of this tile has drawn)
          }
          *sk_flag = last ? 1u : 0u;
        }
        named_bar_sync(4, kCvtThreads + kStatThreads);  // the store / statistics warps learn the outcome here too
        if (*sk_flag == 0u) {
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1u;
          continue;
        }
      }
#pragma unroll
      for (int sub = 0; sub < kSubs; ++sub) {
        const uint32_t b = u & 1u;
        if (p.out_mode == 0) {
          // the buffer is free (and, with a residual addend, already holds the addend sub-tile)
          if (has_add) mbar_wait(&aload[b], (u >> 1) & 1u);
          else mbar_wait(&ofree[b], ((u >> 1) & 1u) ^ 1u);
        }
        const uint32_t obuf = obase + b * Cfg::kOutBytes;
#pragma unroll
        for (int ci = 0; ci < CHs; ++ci) {
          const int cs = hf * CHs + ci;             // chunk inside the sub-tile
          const int col_t = sub * SN + cs * 32;     // column inside the tile
          const int col0 = tk.n_blk * BN + col_t;   // global output column
          uint32_t v[32];
          if (!splitk) tmem_ld_32x32(t_addr + col_t, v);
          const uint32_t half_base = obuf + (cs >> 1) * (kTileM * 128) + srow * 128;
          const int c16_0 = (cs & 1) * 4;
          uint4 ad[4];
          if (has_add) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ad[i] = lds128(half_base + (((c16_0 + i) ^ (srow & 7)) * 16));
          }
          float f[32];
          if (!splitk) {
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          } else {
            // split-K, phase B: the partial tiles of all splits added in split order (the same order on every launch)
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = 0.f;
            const float* src0 = p.sk_part + static_cast<size_t>(sk_unit) * sk_tile + static_cast<size_t>(row) * BN + col_t;
#pragma unroll 1
            for (int sp = 0; sp < p.splits; ++sp) {
              const float4* src = reinterpret_cast<const float4*>(src0 + static_cast<size_t>(sp) * sk_units * sk_tile);
              float4 qv[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) qv[i] = __ldcg(src + i);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                f[4 * i] += qv[i].x;
                f[4 * i + 1] += qv[i].y;
                f[4 * i + 2] += qv[i].z;
                f[4 * i + 3] += qv[i].w;
              }
            }
          }
          if (p.bias_mod > 0) {
            const int b0 = col0 % p.bias_mod;  // chunks never straddle a bias period (bias_mod % 32 == 0 or col < n_out)
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (col0 + j < p.n_out) f[j] += __ldg(p.bias + ((b0 + j) < p.bias_mod ? (b0 + j) : (b0 + j - p.bias_mod)));
            }
          }
          if (has_add) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t w[4] = {ad[i].x, ad[i].y, ad[i].z, ad[i].w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                f[i * 8 + 2 * j] += bf16lo(w[j]);
                f[i * 8 + 2 * j + 1] += bf16hi(w[j]);
              }
            }
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          if (p.out_mode == 0) {
            // 128-byte swizzle: 16-byte chunk index XOR (row & 7) inside each 64-channel half.
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint32_t pk[4];
#pragma unroll
              for (int j = 0; j < 4; ++j)
                pk[j] = valid ? pack_bf16x2(f[i * 8 + 2 * j], f[i * 8 + 2 * j + 1]) : 0u;
              const int c16 = (c16_0 + i) ^ (srow & 7);
              sts128(half_base + c16 * 16, pk[0], pk[1], pk[2], pk[3]);
            }
          } else if (valid) {
            float* dst = p.out_f32 + static_cast<long long>(w0 + wi) * p.out_ld + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.n_out) dst[j] = f[j];
          }
        }
        if (sub == kSubs - 1 && !splitk) {
          // every TMEM read of this accumulator has completed: hand it back to the MMA issuer
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) {
            if (CG == 2) mbar_arrive_cluster(lead_tempty + acc * 8);
            else mbar_arrive(&tempty[acc]);
          }
        }
        if (p.out_mode == 0) {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&ofull[b]);
        }
        ++u;
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  } else if (warp >= 4 + kCvtWarps) {
    // ------------------------------------------------------------------ TMA store + BatchNorm statistics (4 warps)
    // thread -> (8-column chunk, row slab) of a staged sub-tile; running sums stay in registers across tiles and
    // are reduced over the slabs through shared memory once per CTA.
    constexpr int kC8 = SN / 8;                   // 16-byte column chunks per sub-tile row: 8 or 16
    constexpr int kSlabs = kStatThreads / kC8;    // 16 or 8
    constexpr int kSlabRows = kTileM / kSlabs;    // 8 or 16
    const int st = threadIdx.x - (4 + kCvtWarps) * 32;
    const int st_c8 = st % kC8;
    const int st_slab = st / kC8;
    const bool do_stats = p.stat_mod > 0 && !DMPI_DBG(p, 2);
    float st_sum[kSubs][8], st_sq[kSubs][8];
#pragma unroll
    for (int s = 0; s < kSubs; ++s)
#pragma unroll
      for (int j = 0; j < 8; ++j) st_sum[s][j] = st_sq[s][j] = 0.f;
    int st_nblk = -1;
    const uint32_t obase = smem_u32(smem_o);
    if (p.out_mode == 0) {
      uint32_t u = 0;
      // Residual addend: thread 0 walks a second cursor two sub-tiles ahead and TMA-loads the addend sub-tile into
      // the staging buffer as soon as that buffer has been stored and read for the statistics.
      const bool has_add = p.has_add != 0;
      TileWalk ahead;
      uint32_t ua = 0;
      auto prefetch_addend = [&]() {  // loads sub-tile `ua` (if any) into buffer ua & 1
        if (ahead.remaining <= 0) return;
        const int sub = static_cast<int>(ua % kSubs);
        const uint32_t b = ua & 1u;
        const int aw0 = ahead.tw * p.box_w, ah0 = ahead.th * p.box_h, an0 = ahead.tn * p.box_n;
        const int ch0 = ahead.n_blk * BN + sub * SN;
        int halves = 0;
#pragma unroll
        for (int h = 0; h < SN / 64; ++h) halves += (ch0 + h * 64 < p.n_out) ? 1 : 0;
        mbar_arrive_expect_tx(&aload[b], static_cast<uint32_t>(halves * p.rows) * 128u);
#pragma unroll
        for (int h = 0; h < SN / 64; ++h) {
          if (ch0 + h * 64 < p.n_out)
            tma_load_4d(smem_o + b * Cfg::kOutBytes + h * (kTileM * 128), &p.add_map, &aload[b], ch0 + h * 64, aw0, ah0,
                        an0);
        }
        ++ua;
        if (ua % kSubs == 0) ahead.next(p);
      };
      // split-K: only the CTA that drew a tile's last ticket stores anything; the others still take part in the
      // statistics bookkeeping below (with zero sums), so that every CTA of the grid draws its statistics ticket
      bool sk_skip = false;
      if (splitk) {
        named_bar_sync(4, kCvtThreads + kStatThreads);  // the converters have decided (see phase A there)
        sk_skip = *sk_flag == 0u;
      }
      if (has_add && st == 0 && !sk_skip) {
        ahead.init(p, CG, rank);
        prefetch_addend();
        prefetch_addend();
      }
      TileWalk tk;
      for (tk.init(p, CG, rank); tk.remaining > 0; tk.next(p)) {
        const int w0 = tk.tw * p.box_w, h0 = tk.th * p.box_h, n0 = tk.tn * p.box_n;
        if (do_stats && tk.n_blk != st_nblk) {
          // the CTA moves to another N tile (rare: the grid is a multiple of the N-tile count): flush directly
          if (st_nblk >= 0) {
#pragma unroll
            for (int s = 0; s < kSubs; ++s)
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const int gcol = st_nblk * BN + s * SN + st_c8 * 8 + j;
                if (gcol < p.n_out) {
                  atomicAdd(p.stat_sum + (gcol % p.stat_mod), st_sum[s][j]);
                  atomicAdd(p.stat_sq + (gcol % p.stat_mod), st_sq[s][j]);
                }
                st_sum[s][j] = st_sq[s][j] = 0.f;
              }
          }
          st_nblk = tk.n_blk;
        }
        if (sk_skip) continue;
#pragma unroll
        for (int sub = 0; sub < kSubs; ++sub) {
          const uint32_t b = u & 1u;
          const uint32_t obuf = obase + b * Cfg::kOutBytes;
          mbar_wait(&ofull[b], (u >> 1) & 1u);
          if (st == 0 && !DMPI_DBG(p, 1)) {
            const CUtensorMap* omap = &p.o_map[tk.n_blk / p.nblk_per_oview];
            const int ch0 = (tk.n_blk % p.nblk_per_oview) * BN + sub * SN;
#pragma unroll
            for (int h = 0; h < SN / 64; ++h) {
              if (tk.n_blk * BN + sub * SN + h * 64 < p.n_out)
                tma_store_4d(omap, smem_o + b * Cfg::kOutBytes + h * (kTileM * 128), ch0 + h * 64, w0, h0, n0);
            }
            tma_store_commit();
          }
          if (do_stats) {
            // column sums of the bf16 sub-tile just staged (exactly the values that reach memory), 16 bytes per load
            const uint32_t hb = obuf + (st_c8 >> 3) * (kTileM * 128);
            const int c16 = st_c8 & 7;
#pragma unroll
            for (int rr = 0; rr < kSlabRows; ++rr) {
              const int r = st_slab * kSlabRows + rr;
              const uint4 v = lds128(hb + r * 128 + ((c16 ^ (r & 7)) * 16));
              const uint32_t vi[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float x0 = bf16lo(vi[j]), x1 = bf16hi(vi[j]);
                st_sum[sub][2 * j] += x0;
                st_sum[sub][2 * j + 1] += x1;
                st_sq[sub][2 * j] = fmaf(x0, x0, st_sq[sub][2 * j]);
                st_sq[sub][2 * j + 1] = fmaf(x1, x1, st_sq[sub][2 * j + 1]);
              }
            }
          }
          if (st == 0) tma_store_wait_read0();  // the store has finished reading the buffer
          if (has_add) {
            // nobody may still be reading the buffer when the next addend sub-tile lands in it
            if (do_stats) named_bar_sync(2, kStatThreads);
            if (st == 0) prefetch_addend();
          } else {
            __syncwarp();
            if (lane == 0) mbar_arrive(&ofree[b]);
          }
          ++u;
        }
      }
      if (do_stats && st_nblk >= 0) {
        // every buffer has been consumed and drained: reuse the staging area as reduction scratch
        named_bar_sync(2, kStatThreads);
        float* red = reinterpret_cast<float*>(smem_o);  // [2][kSubs][kSlabs][SN] floats (<= 16 KB)
#pragma unroll
        for (int s = 0; s < kSubs; ++s)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            red[((0 * kSubs + s) * kSlabs + st_slab) * SN + st_c8 * 8 + j] = st_sum[s][j];
            red[((1 * kSubs + s) * kSlabs + st_slab) * SN + st_c8 * 8 + j] = st_sq[s][j];
          }
        named_bar_sync(2, kStatThreads);
        for (int c = st; c < BN; c += kStatThreads) {
          const int s = c / SN, cc = c % SN;
          float t1 = 0.f, t2 = 0.f;
#pragma unroll
          for (int sl = 0; sl < kSlabs; ++sl) {
            t1 += red[((0 * kSubs + s) * kSlabs + sl) * SN + cc];
            t2 += red[((1 * kSubs + s) * kSlabs + sl) * SN + cc];
          }
          const int gcol = st_nblk * BN + c;
          if (p.stat_scratch != nullptr) {
            // deterministic path: park the partial sums; the last CTAs add all of them up in CTA order (below)
            if (!splitk) {
              p.stat_scratch[kStatHdr + (blockIdx.x * 2 + 0) * BN + c] = t1;
              p.stat_scratch[kStatHdr + (blockIdx.x * 2 + 1) * BN + c] = t2;
            } else {
              // split-K: WHICH CTA finalises a tile changes from launch to launch, so its sums go to the slot of the
              // tile's first CTA (index sk_unit) and every other slot holds zeros - a fixed summation order again
              if (!sk_skip) {
                p.stat_scratch[kStatHdr + (sk_unit * 2 + 0) * BN + c] = t1;
                p.stat_scratch[kStatHdr + (sk_unit * 2 + 1) * BN + c] = t2;
              }
              if (static_cast<int>(blockIdx.x) >= sk_units) {
                p.stat_scratch[kStatHdr + (blockIdx.x * 2 + 0) * BN + c] = 0.f;
                p.stat_scratch[kStatHdr + (blockIdx.x * 2 + 1) * BN + c] = 0.f;
              }
            }
          } else if (gcol < p.n_out) {
            atomicAdd(p.stat_sum + (gcol % p.stat_mod), t1);
            atomicAdd(p.stat_sq + (gcol % p.stat_mod), t2);
          }
        }
        if (p.stat_scratch != nullptr) {
          // release: the barrier orders every thread's partial stores before thread 0's fence + ticket (one fence per
          // CTA instead of 128); acquire on the other side: fence after drawing the last ticket, then the CTA barrier
          named_bar_sync(2, kStatThreads);
          if (st == 0 && !DMPI_DBG(p, 32)) {
            __threadfence();
            uint32_t* ticket = reinterpret_cast<uint32_t*>(p.stat_scratch);  // fixed place: launches of every shape share it
            const uint32_t t = atomicAdd(ticket, 1u);
            const uint32_t tail = gridDim.x < static_cast<uint32_t>(kTailCtas) ? gridDim.x : static_cast<uint32_t>(kTailCtas);
            // the last `tail` CTAs to finish share the final summation (role = order of arrival among them)
            *last_flag = t >= gridDim.x - tail ? t - (gridDim.x - tail) + 1u : 0u;
          }
        }
      }
      if (st == 0) tma_store_wait_all0();
    }
  }

  tc_fence_before_sync();
  if (CG == 2) {
    cluster_sync_all();  // the leader's MMAs read the peer's shared memory and write its TMEM until the last commit
    if (warp == 2) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
  } else {
    __syncthreads();
    if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }

  // Deterministic BatchNorm statistics: the last kTailCtas CTAs to finish add the partial sums of all CTAs in CTA
  // order, each for its own slice of the channels (role r takes quads [nq*r/R, nq*(r+1)/R)).  L lanes share a quad of
  // 4 channels (each takes every L-th CTA of that N tile's list, ascending) and combine with a fixed shuffle tree, so
  // the summation order is a function of the launch geometry only — not of which CTA ends up in which role.
  // The tail CTAs wait (bounded spin) until every CTA of the grid has published its partials; at most kTailCtas SMs
  // are held while waiting, so CTAs that have not been scheduled yet always find room.
  // This code runs once per launch from a cold instruction cache (a single CTA walking an unrolled 7 KB version of it
  // cost 10-12 us per launch): it is kept small, rolled and parallel on purpose.
  if (p.stat_scratch != nullptr && *last_flag != 0u && !DMPI_DBG(p, 16)) {
    const int role = static_cast<int>(*last_flag) - 1;
    const int grid = static_cast<int>(gridDim.x);
    const int R = grid < kTailCtas ? grid : kTailCtas;
    uint32_t* ctr = reinterpret_cast<uint32_t*>(p.stat_scratch);  // [0] finished CTAs, [1] finished tail roles
    if (threadIdx.x == 0) {
      uint64_t t0 = 0;
      uint32_t spins = 0, seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
        if ((++spins & 0xFFFu) == 0) {
          const uint64_t now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          else if (DMPI_MBAR_TIMEOUT_NS != 0ull && now - t0 > DMPI_MBAR_TIMEOUT_NS) __trap();
        }
      } while (seen < static_cast<uint32_t>(grid));
    }
    __syncthreads();
    const int nnt = p.num_n_tiles;
    const int cpt = ((grid / CG + nnt - 1) / nnt) * CG;  // CTAs per N tile (upper bound)
    const int nq = p.stat_mod >> 2;
    const int q_lo = static_cast<int>(static_cast<long long>(nq) * role / R);
    const int q_hi = static_cast<int>(static_cast<long long>(nq) * (role + 1) / R);
    int L = 32;  // lanes per quad: as many as still cover this role's quads in one pass
    while (L > 1 && (kThreads / L < q_hi - q_lo || L > cpt)) L >>= 1;
    const int l = threadIdx.x & (L - 1);
    const int per_pass = kThreads / L;
    for (int q0 = q_lo; q0 < q_hi; q0 += per_pass) {
      const int quad = q0 + static_cast<int>(threadIdx.x) / L;
      float4 t1 = make_float4(0.f, 0.f, 0.f, 0.f), t2 = t1;
      if (quad < q_hi) {
        for (int gcol = quad * 4; gcol < p.n_out; gcol += p.stat_mod) {  // column groups that share the channels (convT)
          const int n_blk = gcol / BN, cc = gcol - n_blk * BN;
          const float* base = p.stat_scratch + kStatHdr + cc;
          // list entry idx of this tile: CTA blockIdx = CG * (n_blk + (idx / CG) * nnt) + idx % CG
#pragma unroll 1
          for (int idx = l; idx < cpt; idx += kDetInFlight * L) {
            float4 a1[kDetInFlight], a2[kDetInFlight];
#pragma unroll
            for (int k = 0; k < kDetInFlight; ++k) {  // independent loads in flight; entries past the list add zeros
              const int id = idx + k * L;
              const int b = CG * (n_blk + (id / CG) * nnt) + (id % CG);
              a1[k] = a2[k] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (id < cpt && b < grid) {
                a1[k] = __ldcg(reinterpret_cast<const float4*>(base + static_cast<size_t>(b * 2 + 0) * BN));
                a2[k] = __ldcg(reinterpret_cast<const float4*>(base + static_cast<size_t>(b * 2 + 1) * BN));
              }
            }
#pragma unroll
            for (int k = 0; k < kDetInFlight; ++k) {
              t1.x += a1[k].x; t1.y += a1[k].y; t1.z += a1[k].z; t1.w += a1[k].w;
              t2.x += a2[k].x; t2.y += a2[k].y; t2.z += a2[k].z; t2.w += a2[k].w;
            }
          }
        }
      }
#pragma unroll 1
      for (int o = 1; o < L; o <<= 1) {
        t1.x += __shfl_xor_sync(0xffffffffu, t1.x, o); t1.y += __shfl_xor_sync(0xffffffffu, t1.y, o);
        t1.z += __shfl_xor_sync(0xffffffffu, t1.z, o); t1.w += __shfl_xor_sync(0xffffffffu, t1.w, o);
        t2.x += __shfl_xor_sync(0xffffffffu, t2.x, o); t2.y += __shfl_xor_sync(0xffffffffu, t2.y, o);
        t2.z += __shfl_xor_sync(0xffffffffu, t2.z, o); t2.w += __shfl_xor_sync(0xffffffffu, t2.w, o);
      }
      if (l == 0 && quad < q_hi) {
        float4* s1 = reinterpret_cast<float4*>(p.stat_sum + quad * 4);
        float4* s2 = reinterpret_cast<float4*>(p.stat_sq + quad * 4);
        float4 v1 = *s1, v2 = *s2;
        v1.x += t1.x; v1.y += t1.y; v1.z += t1.z; v1.w += t1.w;
        v2.x += t2.x; v2.y += t2.y; v2.z += t2.z; v2.w += t2.w;
        *s1 = v1;
        *s2 = v2;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(&ctr[1], 1u) == static_cast<uint32_t>(R - 1)) {
      ctr[1] = 0u;  // the last role re-arms both counters for the next launch (stream order makes this visible)
      ctr[0] = 0u;
    }
  }
}

// ------------------------------------------------------------------------------------------ host
static thread_local char g_err[512] = {0};
static void set_err(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }

static int device_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

static int encode_view(CUtensorMap* map, const void* base, const dmpi_tensor4& v, int box_c, int bw, int bh, int bn) {
  const uint8_t* ptr = static_cast<const uint8_t*>(base) + v.offset * 2;
  int box[4] = {box_c, bw, bh, bn};
  // TMA box extents may not exceed 256 and need not be clipped to the tensor; they must be >= 1.
  return encode_bf16_map(map, ptr, 4, v.dims, v.strides, box);
}

template <int BN, int CG, bool SK = false>
static int launch_igemm(const ConvKernelParams& kp, int num_sms, cudaStream_t stream, long long scratch_bytes) {
  using Cfg = IgemmCfg<BN, CG>;
  static bool attr_set = false;
  static int max_clusters = 0;
  if (!attr_set) {
    cudaError_t e =
        cudaFuncSetAttribute(igemm_kernel<BN, CG, SK>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (getenv("DMPI_TRACE")) fprintf(stderr, "[dmpi] cudaFuncSetAttribute: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) {
      set_err(cudaGetErrorString(e));
      return -10;
    }
    if (CG == 2) {
      // how many CTA pairs the device can hold at once (a pair occupies the two SMs of one TPC)
      cudaLaunchConfig_t qc = {};
      qc.gridDim = dim3(num_sms, 1, 1);
      qc.blockDim = dim3(kThreads, 1, 1);
      qc.dynamicSmemBytes = Cfg::kSmemBytes;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = 2;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      qc.attrs = qa;
      qc.numAttrs = 1;
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, igemm_kernel<BN, CG, SK>, &qc);
      if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = num_sms / 2;
      }
      max_clusters = n < num_sms / 2 ? n : num_sms / 2;
      if (getenv("DMPI_TRACE")) fprintf(stderr, "[dmpi] max active CTA pairs: %d\n", max_clusters);
    }
    attr_set = true;
  }
  const int units = (CG == 2 ? (kp.num_m_tiles + 1) / 2 : kp.num_m_tiles) * kp.num_n_tiles;  // tiles or pair tiles
  const int slots = CG == 2 ? max_clusters : num_sms;
  int grid = units < slots ? units : slots;
  if (kp.splits > 1) {
    if (CG != 1 || units * kp.splits > slots) {
      set_err("igemm: split-K needs single-CTA tiles and units * splits <= number of SMs");
      return -13;
    }
    grid = units * kp.splits;  // every CTA exactly one (tile, K slice); all resident at once
  }
  // A grid that is a multiple of the N-tile count keeps n_blk fixed per CTA, so the BatchNorm partial sums stay
  // in registers for the whole kernel (one flush per CTA).
  if (kp.splits <= 1 && grid == slots && kp.num_n_tiles > 1 && kp.num_n_tiles <= slots)
    grid = (slots / kp.num_n_tiles) * kp.num_n_tiles;
  // deterministic statistics need every CTA to stay on one N tile and the scratch to hold one partial per CTA
  const bool one_n_tile_per_cta = grid == units || grid % kp.num_n_tiles == 0;  // (split-K: grid = units * splits)
  grid *= CG;
  ConvKernelParams kp2;
  const ConvKernelParams* kpp = &kp;
  if (kp.stat_scratch != nullptr &&
      (!one_n_tile_per_cta || scratch_bytes < (static_cast<long long>(grid) * 2 * BN + kStatHdr) * 4 || kp.stat_mod <= 0)) {
    kp2 = kp;
    kp2.stat_scratch = nullptr;  // fall back to atomics for this launch
    kpp = &kp2;
  }
  if (getenv("DMPI_TRACE")) fprintf(stderr, "[dmpi] launch grid=%d smem=%d params=%zu bytes\n", grid, Cfg::kSmemBytes, sizeof(kp));
  cudaError_t e;
  if (CG == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, igemm_kernel<BN, CG, SK>, *kpp);
  } else {
    igemm_kernel<BN, CG, SK><<<grid, kThreads, Cfg::kSmemBytes, stream>>>(*kpp);
    e = cudaGetLastError();
  }
  if (getenv("DMPI_TRACE")) {
    fprintf(stderr, "[dmpi] launched: %s\n", cudaGetErrorString(e));
    cudaError_t se = cudaDeviceSynchronize();
    fprintf(stderr, "[dmpi] synced: %s\n", cudaGetErrorString(se));
  }
  if (e != cudaSuccess) {
    set_err(cudaGetErrorString(e));
    return -11;
  }
  return 0;
}

}  // namespace dmpi

using namespace dmpi;

extern "C" dmpi_conv_plan* dmpi_conv_plan_create(const dmpi_conv_desc* desc) {
  dmpi_conv_plan* plan = new (std::nothrow) dmpi_conv_plan();
  if (!plan) return nullptr;
  memset(&plan->params, 0, sizeof(plan->params));
  plan->desc = *desc;
  plan->cached_a = plan->cached_b = plan->cached_o = plan->cached_add = nullptr;
  plan->stat_scratch = nullptr;
  plan->stat_scratch_bytes = 0;
  plan->sk_scratch = nullptr;
  plan->sk_scratch_bytes = 0;
  plan->maps_valid = false;
  plan->num_sms = 0;
  ConvKernelParams& kp = plan->params;
  const dmpi_conv_desc& d = plan->desc;
  kp.ntaps = d.ntaps;
  kp.cblocks = d.cblocks;
  for (int t = 0; t < DMPI_MAX_TAPS; ++t) {
    kp.tap_view[t] = d.tap_view[t];
    kp.tap_dw[t] = d.tap_dw[t];
    kp.tap_dh[t] = d.tap_dh[t];
    kp.tap_koff[t] = d.tap_koff[t];
    kp.tap_woff[t] = d.tap_woff[t];
  }
  kp.box_w = d.box_w;
  kp.box_h = d.box_h;
  kp.box_n = d.box_n;
  kp.rows = d.box_w * d.box_h * d.box_n;
  kp.grid_w = d.grid_w;
  kp.grid_h = d.grid_h;
  kp.grid_n = d.grid_n;
  kp.tiles_w = (d.grid_w + d.box_w - 1) / d.box_w;
  kp.tiles_h = (d.grid_h + d.box_h - 1) / d.box_h;
  kp.tiles_n = (d.grid_n + d.box_n - 1) / d.box_n;
  kp.num_m_tiles = kp.tiles_w * kp.tiles_h * kp.tiles_n;
  kp.num_n_tiles = d.bn > 0 ? (d.n_out + d.bn - 1) / d.bn : 0;
  kp.nblk_per_oview = d.nblk_per_oview > 0 ? d.nblk_per_oview : (kp.num_n_tiles > 0 ? kp.num_n_tiles : 1);
  kp.n_out = d.n_out;
  kp.bias_mod = d.bias_mod;
  kp.stat_mod = d.stat_mod;
  kp.relu = d.relu;
  kp.out_mode = d.out_mode;
  kp.out_ld = d.out_ld;
  kp.k_out = d.k_out;
  kp.c_total = d.c_total;
  kp.w_ld = d.w_ld;
  kp.splits = d.splits;
  kp.patch = d.patch;
  kp.halo = d.halo;
  kp.pw = d.box_w + 2 * d.halo;
  return plan;
}

extern "C" void dmpi_conv_plan_destroy(dmpi_conv_plan* plan) { delete plan; }

// Scratch for deterministic BatchNorm statistics (see ConvKernelParams::stat_scratch): zero-initialised device memory
// of at least (CTAs * 2 * bn + 1) * 4 bytes that no other stream uses concurrently; null disables (atomics).
extern "C" void dmpi_conv_plan_set_stat_scratch(dmpi_conv_plan* plan, void* scratch, long long bytes) {
  plan->stat_scratch = scratch;
  plan->stat_scratch_bytes = scratch ? bytes : 0;
}

// Split-K scratch for dmpi_igemm_run* plans created with desc.splits > 1: device memory whose first 4 KiB (the tile
// tickets) are ZERO at the first use - the kernel leaves them zero - followed by room for splits * tiles partial tiles of
// 128 x bn floats; not shared with a concurrently running launch.
extern "C" void dmpi_conv_plan_set_splitk_scratch(dmpi_conv_plan* plan, void* scratch, long long bytes) {
  plan->sk_scratch = scratch;
  plan->sk_scratch_bytes = scratch ? bytes : 0;
}

extern "C" int dmpi_conv_last_error(char* buf, int buflen) {
  snprintf(buf, buflen, "%s", g_err);
  return static_cast<int>(strlen(g_err));
}

extern "C" int dmpi_igemm_run(dmpi_conv_plan* plan, const void* a, const void* b, void* out, const float* bias,
                              float* stat_sum, float* stat_sqsum, void* stream) {
  return dmpi_igemm_run_add(plan, a, b, out, bias, stat_sum, stat_sqsum, nullptr, 0, 0, 0, stream);
}

extern "C" int dmpi_igemm_run_add(dmpi_conv_plan* plan, const void* a, const void* b, void* out, const float* bias,
                                  float* stat_sum, float* stat_sqsum, const void* addend, long long add_sn,
                                  long long add_sh, long long add_sw, void* stream) {
  const dmpi_conv_desc& d = plan->desc;
  ConvKernelParams& kp = plan->params;
  static const bool trace = getenv("DMPI_TRACE") != nullptr;
  if (trace) fprintf(stderr, "[dmpi] igemm_run plan=%p a=%p b=%p out=%p bn=%d rows=%d\n", (void*)plan, a, b, out, d.bn, kp.rows);
  if (kp.rows < 1 || kp.rows > kTileM) {
    set_err("igemm: pixel box must hold 1..128 rows");
    return -1;
  }
  if (d.bn != 64 && d.bn != 128 && d.bn != 256) {
    set_err("igemm: bn must be 64, 128 or 256");
    return -2;
  }
  if (d.out_mode == 1 && !(d.box_h == 1 && d.box_n == 1 && d.grid_h == 1 && d.grid_n == 1)) {
    set_err("igemm: fp32 direct output needs the flat (GEMM) pixel layout");
    return -3;
  }
  if (d.patch) {
    // ring geometry: as many filter-tap stages as fit next to 2-3 patch stages; the whole filter stays resident when
    // it fits and every tile of the launch uses the same N tile
    const int ph = d.box_h + 2 * d.halo;
    if (d.box_n != 1 || d.a_nviews != 1 || (d.box_h - 1) * kp.pw + d.box_w > kTileM) {
      set_err("igemm: illegal patch-mode geometry");
      return -12;
    }
    const int read_rows = kTileM + 2 * d.halo * kp.pw + 2 * d.halo;  // furthest row a (discarded) accumulator row reads
    const int rows_alloc = read_rows > kp.pw * ph ? read_rows : kp.pw * ph;
    kp.pa_bytes = kp.pw * ph * 128;
    kp.pa_stride = ((rows_alloc * 128 + 1023) / 1024) * 1024;
    const bool pair = d.cta_group == 2;
    const int ring = d.bn == 64 ? (pair ? IgemmCfg<64, 2>::kRingBytes : IgemmCfg<64, 1>::kRingBytes)
                                : (d.bn == 128 ? (pair ? IgemmCfg<128, 2>::kRingBytes : IgemmCfg<128, 1>::kRingBytes)
                                               : (pair ? IgemmCfg<256, 2>::kRingBytes : IgemmCfg<256, 1>::kRingBytes));
    const int bbytes = (pair ? d.bn / 2 : d.bn) * kBlockK * 2;  // a CTA of a pair holds half of every filter-tap tile
    const int taps_total = d.ntaps * d.cblocks;
    // prefer a configuration in which the whole filter stays resident (then only patches stream); 3 patch stages
    // if they fit beside it, else 2
    int pa = 3;
    const bool can_reside = !pair && kp.num_n_tiles == 1 && taps_total <= 12;
    if (can_reside && ring - 3 * kp.pa_stride < taps_total * bbytes && ring - 2 * kp.pa_stride >= taps_total * bbytes) pa = 2;
    if (ring - pa * kp.pa_stride < 4 * bbytes) pa = 2;
    int pb = (ring - pa * kp.pa_stride) / bbytes;
    if (pb > 12) pb = 12;
    if (pb < 2) {
      set_err("igemm: patch does not fit in shared memory");
      return -12;
    }
    kp.pa_stages = pa;
    kp.pb_stages = pb;
    kp.b_resident = (!pair && kp.num_n_tiles == 1 && taps_total <= pb) ? 1 : 0;
  }
  if (!plan->maps_valid || plan->cached_a != a || plan->cached_b != b || plan->cached_o != out) {
    if (d.patch) {
      if (encode_view(&kp.a_map[0], a, d.a_view[0], kBlockK, kp.pw, d.box_h + 2 * d.halo, 1)) {
        set_err("igemm: patch tensor map encode failed");
        return -4;
      }
    } else
    for (int i = 0; i < d.a_nviews; ++i)
      if (encode_view(&kp.a_map[i], a, d.a_view[i], kBlockK, d.box_w, d.box_h, d.box_n)) {
        set_err("igemm: A tensor map encode failed");
        return -4;
      }
    for (int i = d.patch ? 1 : d.a_nviews; i < DMPI_MAX_MAPS; ++i) kp.a_map[i] = kp.a_map[0];
    {
      int64_t dims[2] = {d.b_k, d.b_rows};
      int64_t strides[1] = {d.b_k};
      int box[2] = {kBlockK, d.cta_group == 2 ? d.bn / 2 : d.bn};  // a CTA of a pair loads half of the B tile
      if (encode_bf16_map(&kp.b_map, b, 2, dims, strides, box)) {
        set_err("igemm: B tensor map encode failed");
        return -5;
      }
    }
    if (d.out_mode == 0) {
      for (int i = 0; i < d.o_nviews; ++i)
        if (encode_view(&kp.o_map[i], out, d.o_view[i], 64, d.box_w, d.box_h, d.box_n)) {
          set_err("igemm: output tensor map encode failed");
          return -6;
        }
      for (int i = d.o_nviews; i < DMPI_MAX_MAPS; ++i) kp.o_map[i] = kp.o_map[0];
    } else {
      for (int i = 0; i < DMPI_MAX_MAPS; ++i) kp.o_map[i] = kp.a_map[0];
    }
    plan->cached_a = a;
    plan->cached_b = b;
    plan->cached_o = out;
    plan->maps_valid = true;
  }
#ifdef DMPI_ABLATE
  static const int dbg = getenv("DMPI_IGEMM_DBG") ? atoi(getenv("DMPI_IGEMM_DBG")) : 0;
  kp.dbg = dbg;
#else
  kp.dbg = 0;
#endif
  kp.has_add = 0;
  if (addend != nullptr) {
    if (d.out_mode != 0 || d.o_nviews != 1) {
      set_err("igemm: the residual addend needs the bf16 output mode and a single output view");
      return -9;
    }
    if (plan->cached_add != addend || plan->cached_add_strides[0] != add_sn || plan->cached_add_strides[1] != add_sh ||
        plan->cached_add_strides[2] != add_sw) {
      dmpi_tensor4 av = d.o_view[0];  // same pixel grid and channel count as the output, its own strides
      av.offset = 0;
      av.strides[0] = add_sw;
      av.strides[1] = add_sh;
      av.strides[2] = add_sn;
      if (encode_view(&kp.add_map, addend, av, 64, d.box_w, d.box_h, d.box_n)) {
        set_err("igemm: addend tensor map encode failed");
        return -9;
      }
      plan->cached_add = addend;
      plan->cached_add_strides[0] = add_sn;
      plan->cached_add_strides[1] = add_sh;
      plan->cached_add_strides[2] = add_sw;
    }
    kp.has_add = 1;
  }
  kp.splits = d.splits > 1 ? d.splits : 0;
  kp.sk_part = nullptr;
  kp.sk_tickets = nullptr;
  if (kp.splits > 1) {
    const int num_kb = d.ntaps * d.cblocks;
    const int per = (num_kb + kp.splits - 1) / kp.splits;
    const long long units = static_cast<long long>(kp.num_m_tiles) * kp.num_n_tiles;
    const long long need = kSkHeaderBytes + static_cast<long long>(kp.splits) * units * kTileM * d.bn * 4;
    if (d.patch || d.cta_group == 2 || d.out_mode != 0 || per * (kp.splits - 1) >= num_kb || units > kSkHeaderBytes / 4) {
      set_err("igemm: split-K needs per-tap loads, one CTA per tile, bf16 output and a K range for every split");
      return -13;
    }
    if (plan->sk_scratch == nullptr || plan->sk_scratch_bytes < need) {
      set_err("igemm: split-K scratch missing or too small (dmpi_conv_plan_set_splitk_scratch)");
      return -13;
    }
    kp.sk_tickets = static_cast<unsigned int*>(plan->sk_scratch);
    kp.sk_part = reinterpret_cast<float*>(static_cast<char*>(plan->sk_scratch) + kSkHeaderBytes);
  }
  kp.stat_scratch = (d.stat_mod > 0 && stat_sum != nullptr) ? static_cast<float*>(plan->stat_scratch) : nullptr;
  kp.bias = bias;
  kp.stat_sum = stat_sum;
  kp.stat_sq = stat_sqsum;
  kp.out_f32 = d.out_mode == 1 ? static_cast<float*>(out) : nullptr;
  if (d.bias_mod > 0 && !bias) {
    set_err("igemm: bias_mod set but bias pointer is null");
    return -7;
  }
  if (d.stat_mod > 0 && (!stat_sum || !stat_sqsum)) {
    set_err("igemm: stat_mod set but statistics pointers are null");
    return -8;
  }
  if (trace) fprintf(stderr, "[dmpi] maps encoded\n");
  if (plan->num_sms == 0) plan->num_sms = device_sm_count();
  if (trace) fprintf(stderr, "[dmpi] sms=%d launching\n", plan->num_sms);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (d.cta_group == 2) {
    if (d.bn == 128) return launch_igemm<128, 2>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    if (d.bn == 256) return launch_igemm<256, 2>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    if (d.bn == 64) return launch_igemm<64, 2>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    set_err("igemm: bn must be 64, 128 or 256");
    return -2;
  }
  if (kp.splits > 1) {
    switch (d.bn) {
      case 64: return launch_igemm<64, 1, true>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
      case 128: return launch_igemm<128, 1, true>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
      default: return launch_igemm<256, 1, true>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    }
  }
  switch (d.bn) {
    case 64: return launch_igemm<64, 1>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    case 128: return launch_igemm<128, 1>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
    default: return launch_igemm<256, 1>(kp, plan->num_sms, s, plan->stat_scratch_bytes);
  }
}
