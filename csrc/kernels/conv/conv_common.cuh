// Shared between the K-major implicit-GEMM kernel (igemm.cu) and the MN-major weight-gradient
// kernel (wgrad.cu): kernel parameter block and the host-side plan that owns the tensor maps.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../../include/dmpi_conv.h"

namespace dmpi {

struct alignas(64) ConvKernelParams {
  CUtensorMap a_map[DMPI_MAX_MAPS];  // igemm: A views;  wgrad: X (N-side) views
  CUtensorMap b_map;                 // igemm: weights;  wgrad: dY (M-side)
  CUtensorMap o_map[DMPI_MAX_MAPS];  // igemm: output views
  int ntaps, cblocks;
  int tap_view[DMPI_MAX_TAPS];
  int tap_dw[DMPI_MAX_TAPS];
  int tap_dh[DMPI_MAX_TAPS];
  int tap_koff[DMPI_MAX_TAPS];
  int tap_woff[DMPI_MAX_TAPS];
  int box_w, box_h, box_n, rows;
  int tiles_w, tiles_h, tiles_n;
  int grid_w, grid_h, grid_n;
  int num_m_tiles, num_n_tiles, nblk_per_oview;
  int n_out;
  int bias_mod, stat_mod, relu, out_mode;
  long long out_ld;
  const float* bias;
  float* stat_sum;
  float* stat_sq;
  float* out_f32;
  // wgrad
  float* dw;
  long long w_ld;
  int k_out, c_total, splits, num_k_tiles, num_c_tiles;
  // igemm: optional tensor added to the result in the epilogue (bf16, same pixel grid and box as the output view);
  // its sub-tiles are TMA-prefetched into the staging buffers two sub-tiles ahead of the converters
  CUtensorMap add_map;
  int has_add;
  // igemm patch mode (see dmpi_conv_desc.patch): patch width in pixels, bytes per patch TMA box, shared-memory stride of
  // one patch stage, ring depths, and whether the whole filter stays resident in the B ring
  int patch, halo, pw, pa_bytes, pa_stride, pa_stages, pb_stages, b_resident;
  // igemm: deterministic BatchNorm statistics.  Every CTA stores its per-column partial sums in stat_scratch
  // (a 64-float header holding the uint32 ticket, then [gridDim.x][2][BN] floats) and the last CTA to finish adds them up in CTA order, so the
  // result does not depend on the order in which CTAs retire (fp32 atomics made whole-network gradients differ by
  // ~20 % run to run through ReLU gate flips).  Null: atomics.
  float* stat_scratch;
  // igemm split-K (splits > 1, one CTA per tile, no patch mode): CTA b works on tile b % units over K slice b / units,
  // parks its fp32 partial tile in sk_part[split][unit][128][BN]; the CTA that draws the last ticket of a tile adds the
  // partials in split order (deterministic) and runs the normal epilogue.  For layers with only a handful of tiles,
  // where every CTA otherwise streams its whole K at the bandwidth of a single SM.
  float* sk_part;
  unsigned int* sk_tickets;
  int dbg;  // -DDMPI_ABLATE builds only (otherwise always 0): 1 no TMA store, 2 no statistics, 4 no loads, 8 no MMAs
};

}  // namespace dmpi

struct dmpi_conv_plan {
  dmpi_conv_desc desc;
  dmpi::ConvKernelParams params;
  const void* cached_a;
  const void* cached_b;
  const void* cached_o;
  const void* cached_add;
  long long cached_add_strides[3];  // (sn, sh, sw) the addend map was encoded with: they are per-call arguments
  void* stat_scratch;
  long long stat_scratch_bytes;
  void* sk_scratch;          // split-K: tickets (first kSkHeaderBytes, zero-initialised) + partial tiles
  long long sk_scratch_bytes;
  bool maps_valid;
  int num_sms;
};
