// Fused BatchNorm2d (+ residual add) (+ ReLU), training and eval, forward and backward, NHWC bf16.
//
// The per-channel batch statistics (sum, sum of squares) are produced by the convolution epilogue
// (igemm.cu), so the forward is ONE pass: normalise + affine (+ residual) (+ ReLU), plus the
// running-stat update done by block 0.  The backward is two passes (reduce, apply) with the ReLU
// mask folded in.  Semantics match torch.nn.BatchNorm2d in training mode (biased variance for
// normalisation, unbiased for running_var, momentum EMA, num_batches_tracked += 1) followed by
// ReLU(inplace) — ref:pytorch/unet/model.py:10-11,13-14 and torchvision resnet bn/relu
// (K5, K6, K7).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../common/ptx.cuh"

namespace dmpi {

constexpr int kBnThreads = 256;
constexpr int kDetHdr = 64;  // floats at the head of the deterministic-reduction scratch (ticket counters)
constexpr int kMaxC = 2048;

struct BnFwdArgs {
  const __nv_bfloat16* y;    // conv output [M, C] (row stride ldy)
  const __nv_bfloat16* res;  // optional residual [M, C] (row stride ldr)
  __nv_bfloat16* out;        // [M, C] (row stride ldo)
  long long M;
  int C;
  long long ldy, ldr, ldo;
  const float* sum;    // [C] batch sum of y      (training)
  const float* sqsum;  // [C] batch sum of y*y    (training)
  const float* gamma;
  const float* beta;
  float* save_mean;    // [C] out (training)
  float* save_invstd;  // [C] out (training)
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;
  float eps, momentum;
  int training, relu;
  uint8_t* relu_mask;  // optional [M, C/8]: bit j of byte (row, c/8) = (out[row, c + j] > 0); lets the backward pass
                       // read 1/16 of the bytes of `out` for the ReLU gate of a residual block
};

// kRes: a residual tensor is added; U rows per thread are in flight (all loads issued before the first use).
template <bool kRes, int U>
__global__ void __launch_bounds__(kBnThreads) bn_fwd_kernel(const BnFwdArgs a) {
  extern __shared__ float sm[];  // scale[C], shift[C]
  float* scale = sm;
  float* shift = sm + a.C;
  // Thread (tx, ty): tx owns 8 consecutive channels (16 bytes) for the whole kernel -> scale/shift in registers,
  // ty strides the rows; U rows per iteration with all loads issued up front.
  const int vecC = a.C >> 3;
  const int tx = threadIdx.x % vecC;
  const int ty = threadIdx.x / vecC;
  const int rows_per_block = blockDim.x / vecC;
  const bool active = ty < rows_per_block;
  const int c0 = tx << 3;
  const long long stride = static_cast<long long>(gridDim.x) * rows_per_block;
  long long row0 = blockIdx.x * static_cast<long long>(rows_per_block) + ty;
  // The first batch of rows is requested BEFORE the per-channel constants are derived: the two are independent, and
  // with the constants first (then a block barrier, then the rows) every launch paid two memory round trips in a row -
  // a third of the run time of the small late-layer launches.
  uint4 v[U], r[kRes ? U : 1];
  auto load_rows = [&](long long base) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = base + u * stride;
      if (row < a.M) {
        v[u] = *reinterpret_cast<const uint4*>(a.y + row * a.ldy + c0);
        if (kRes) r[u] = *reinterpret_cast<const uint4*>(a.res + row * a.ldr + c0);
      }
    }
  };
  if (active) load_rows(row0);

  const float invM = 1.f / static_cast<float>(a.M);
  for (int c = threadIdx.x; c < a.C; c += blockDim.x) {
    float mean, invstd;
    if (a.training) {
      mean = a.sum[c] * invM;
      float var = fmaxf(a.sqsum[c] * invM - mean * mean, 0.f);
      invstd = rsqrtf(var + a.eps);
      if (blockIdx.x == 0) {
        a.save_mean[c] = mean;
        a.save_invstd[c] = invstd;
        if (a.running_mean) {
          const float unbiased = a.M > 1 ? var * (static_cast<float>(a.M) / static_cast<float>(a.M - 1)) : var;
          a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * mean;
          a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * unbiased;
        }
      }
    } else {
      mean = a.running_mean[c];
      invstd = rsqrtf(a.running_var[c] + a.eps);
    }
    const float g = a.gamma ? a.gamma[c] : 1.f;
    const float b = a.beta ? a.beta[c] : 0.f;
    scale[c] = g * invstd;
    shift[c] = b - mean * g * invstd;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.training && a.num_batches_tracked) *a.num_batches_tracked += 1;
  __syncthreads();
  if (!active) return;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale[c0 + j];
    sh[j] = shift[c0 + j];
  }
  while (row0 < a.M) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * stride;
      if (row >= a.M) break;
      const uint4 rr = kRes ? r[kRes ? u : 0] : make_uint4(0, 0, 0, 0);
      const uint32_t vi[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
      const uint32_t ri[4] = {rr.x, rr.y, rr.z, rr.w};
      uint32_t oi[4];
      uint32_t bits = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo = fmaf(bf16lo(vi[j]), sc[2 * j], sh[2 * j]) + bf16lo(ri[j]);
        float hi = fmaf(bf16hi(vi[j]), sc[2 * j + 1], sh[2 * j + 1]) + bf16hi(ri[j]);
        if (a.relu) {
          lo = fmaxf(lo, 0.f);
          hi = fmaxf(hi, 0.f);
        }
        oi[j] = pack_bf16x2(lo, hi);
        // the gate the backward pass uses is (stored bf16 value > 0)
        bits |= ((oi[j] & 0x7FFFu) != 0u && !(oi[j] & 0x8000u) ? 1u : 0u) << (2 * j);
        bits |= ((oi[j] & 0x7FFF0000u) != 0u && !(oi[j] & 0x80000000u) ? 1u : 0u) << (2 * j + 1);
      }
      *reinterpret_cast<uint4*>(a.out + row * a.ldo + c0) = make_uint4(oi[0], oi[1], oi[2], oi[3]);
      if (a.relu_mask) a.relu_mask[row * vecC + tx] = static_cast<uint8_t>(bits);
    }
    row0 += U * stride;
    if (row0 < a.M) load_rows(row0);
  }
}

// Expands the 8 gate bits of one (row, 8-channel group) into the bf16x2 layout of an activation vector (1.0 / 0.0),
// so that the gate tests below work unchanged on either source.
__device__ __forceinline__ uint4 expand_relu_bits(uint32_t b) {
  uint4 v;
  v.x = ((b & 1u) ? 0x3f80u : 0u) | ((b & 2u) ? 0x3f800000u : 0u);
  v.y = ((b & 4u) ? 0x3f80u : 0u) | ((b & 8u) ? 0x3f800000u : 0u);
  v.z = ((b & 16u) ? 0x3f80u : 0u) | ((b & 32u) ? 0x3f800000u : 0u);
  v.w = ((b & 64u) ? 0x3f80u : 0u) | ((b & 128u) ? 0x3f800000u : 0u);
  return v;
}

struct BnBwdArgs {
  const __nv_bfloat16* dout;  // grad wrt the block output (post ReLU) [M, C]
  const __nv_bfloat16* y;     // saved conv output (pre BN)
  const __nv_bfloat16* act;   // saved block output (post ReLU); ReLU mask = act > 0 (null: no ReLU)
  __nv_bfloat16* dy;          // grad wrt conv output
  __nv_bfloat16* dres;        // optional: grad wrt the residual input (= masked dout)
  long long M;
  int C;
  long long ld_dout, ldy, ld_act, ld_dy, ld_dres;
  const float* mean;
  const float* invstd;
  const float* gamma;
  float* acc;     // scratch [2*C]: sum(dz), sum(dz * xhat); must be zero on entry of the reduce pass
  float* dgamma;  // parameter gradients, accumulated (+=) by the apply pass
  float* dbeta;
  const float* beta;  // only for relu_from_y
  const uint8_t* relu_mask;  // bit mask written by the forward pass (replaces `act` as the gate source when set)
  int relu_from_y;    // ReLU mask recomputed as (y*scale + shift > 0) instead of being read from `act`: exact for a
                      // BatchNorm+ReLU without residual (same fp32 fma as the forward) and saves one tensor read per pass
};

// Where the ReLU gate of the backward pass comes from (compile-time: the variants need different registers).
constexpr int kGateNone = 0;  // no ReLU
constexpr int kGateY = 1;     // recomputed as (y*scale + shift > 0)
constexpr int kGateBits = 2;  // the bit mask written by the forward pass
constexpr int kGateAct = 3;   // the saved forward output (> 0)

// Pass 1: per-channel reductions.  Thread (tx, ty): tx owns 8 consecutive channels, ty strides rows.
// kDet selects the deterministic variant: grid = (C/64 channel slabs, row blocks), a block covers 64 channels x 32 rows
// per step, every block parks its 128 partial sums in `partials` and the last block of a slab (ticket counter) adds
// the slab's partials in row-block order — no floating-point atomics, so the result is bit-reproducible.
// U rows per thread are in flight: a read-only pass needs ~100 KB of loads in flight per SM to approach the HBM rate
// (loaded latency ~1.7 us); two rows per thread on four resident blocks gave 32-64 KB and 3.3-5 of 6.5 TB/s.
template <int kGate, bool kDet, int U>
__global__ void __launch_bounds__(kBnThreads, 2) bn_bwd_reduce_kernel(const BnBwdArgs a, float* partials) {
  const int vecC = a.C >> 3;
  // tx: global 8-channel group of this thread; ty: its row inside the block's row step
  const int tx = kDet ? static_cast<int>(blockIdx.x) * 8 + (threadIdx.x & 7) : threadIdx.x % vecC;
  const int ty = kDet ? (threadIdx.x >> 3) : threadIdx.x / vecC;
  const int rows_per_block = kDet ? 32 : blockDim.x / vecC;
  const int c0 = tx << 3;
  // s2 accumulates dz * (y - mean); the common factor invstd is applied once per block after the reduction.
  float s1[8], s2[8], mean[8], sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s1[j] = s2[j] = 0.f;
    mean[j] = a.mean[c0 + j];
    sc[j] = sh[j] = 0.f;
    if (kGate == kGateY) {
      const float g = a.gamma ? a.gamma[c0 + j] : 1.f;
      const float b = a.beta ? a.beta[c0 + j] : 0.f;
      const float istd = a.invstd[c0 + j];
      sc[j] = g * istd;
      sh[j] = b - mean[j] * g * istd;
    }
  }
  if (ty < rows_per_block) {
    const long long stride = static_cast<long long>(kDet ? gridDim.y : gridDim.x) * rows_per_block;
    // The masked gradient is the residual branch's gradient; writing it here lets the apply pass read it back
    // instead of dout + gate (measured: storing it from the apply pass instead is 6 % slower overall).
    const bool write_dres = (kGate == kGateBits || kGate == kGateAct) && a.dres != nullptr;
    for (long long row0 = (kDet ? blockIdx.y : blockIdx.x) * static_cast<long long>(rows_per_block) + ty; row0 < a.M;
         row0 += U * stride) {
      uint4 g[U], yv[U], av[kGate == kGateAct ? U : 1];
      uint32_t mb[kGate == kGateBits ? U : 1];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = row0 + u * stride;
        if (row < a.M) {
          g[u] = *reinterpret_cast<const uint4*>(a.dout + row * a.ld_dout + c0);
          yv[u] = *reinterpret_cast<const uint4*>(a.y + row * a.ldy + c0);
          if (kGate == kGateBits) mb[u] = a.relu_mask[row * vecC + tx];
          if (kGate == kGateAct) av[u] = *reinterpret_cast<const uint4*>(a.act + row * a.ld_act + c0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = row0 + u * stride;
        if (row >= a.M) break;
        uint4 gate = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
        if (kGate == kGateBits) gate = expand_relu_bits(mb[u]);
        if (kGate == kGateAct) gate = av[u];
        const uint32_t gi[4] = {g[u].x, g[u].y, g[u].z, g[u].w}, yi[4] = {yv[u].x, yv[u].y, yv[u].z, yv[u].w},
                       ai[4] = {gate.x, gate.y, gate.z, gate.w};
        uint32_t ri[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool m0 = kGate == kGateY ? fmaf(bf16lo(yi[j]), sc[2 * j], sh[2 * j]) > 0.f : bf16lo(ai[j]) > 0.f;
          const bool m1 = kGate == kGateY ? fmaf(bf16hi(yi[j]), sc[2 * j + 1], sh[2 * j + 1]) > 0.f : bf16hi(ai[j]) > 0.f;
          const float d0 = m0 ? bf16lo(gi[j]) : 0.f;
          const float d1 = m1 ? bf16hi(gi[j]) : 0.f;
          s1[2 * j] += d0;
          s1[2 * j + 1] += d1;
          s2[2 * j] = fmaf(d0, bf16lo(yi[j]) - mean[2 * j], s2[2 * j]);
          s2[2 * j + 1] = fmaf(d1, bf16hi(yi[j]) - mean[2 * j + 1], s2[2 * j + 1]);
          ri[j] = (m0 ? (gi[j] & 0x0000FFFFu) : 0u) | (m1 ? (gi[j] & 0xFFFF0000u) : 0u);
        }
        if (write_dres) *reinterpret_cast<uint4*>(a.dres + row * a.ld_dres + c0) = make_uint4(ri[0], ri[1], ri[2], ri[3]);
      }
    }
  }
  // block reduction over ty through shared memory
  extern __shared__ float red[];  // [rows_per_block][C][2] would be large; reduce in 8-channel slabs
  float* r1 = red;                 // [blockDim.x * 8]
  float* r2 = red + blockDim.x * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    r1[threadIdx.x * 8 + j] = s1[j];
    r2[threadIdx.x * 8 + j] = s2[j];
  }
  __syncthreads();
  if (kDet) {
    // thread q < 128: quantity q / 64 (sum dz | sum dz*(y-mean)) of slab channel q % 64, rows added in ty order
    __shared__ uint32_t is_last;
    const int slab = blockIdx.x, RB = gridDim.y;
    uint32_t* tickets = reinterpret_cast<uint32_t*>(partials);          // [kDetHdr] header: one ticket per slab
    float* part = partials + kDetHdr + static_cast<size_t>(slab) * RB * 128;
    if (threadIdx.x < 128) {
      const float* src = threadIdx.x < 64 ? r1 : r2;
      const int ch = threadIdx.x & 63;
      float t = 0.f;
#pragma unroll 8
      for (int r = 0; r < 32; ++r) t += src[(r * 8 + (ch >> 3)) * 8 + (ch & 7)];
      part[blockIdx.y * 128 + threadIdx.x] = t;
    }
    __syncthreads();  // orders every thread's partial store before thread 0's fence + ticket (release)
    if (threadIdx.x == 0) {
      __threadfence();
      const bool last = atomicAdd(&tickets[slab], 1u) == static_cast<uint32_t>(RB - 1);
      if (last) __threadfence();  // acquire
      is_last = last ? 1u : 0u;
    }
    __syncthreads();
    if (is_last) {
      // 8 warps x 32 lanes: a lane owns four of the slab's 128 quantities (one 16-byte load per row block), warp w adds
      // row blocks w, w + 8, ... in ascending order with eight loads in flight; the eight warp sums are then added in
      // warp order.  The order depends on the launch geometry only.
      const int quad = threadIdx.x & 31, grp = threadIdx.x >> 5;
      const float4* src = reinterpret_cast<const float4*>(part) + quad;  // row block stride: 32 float4
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      int rb = grp;
      for (; rb + 56 < RB; rb += 64) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = __ldcg(src + static_cast<size_t>(rb + 8 * k) * 32);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          t.x += v[k].x; t.y += v[k].y; t.z += v[k].z; t.w += v[k].w;
        }
      }
      for (; rb < RB; rb += 8) {
        const float4 v = __ldcg(src + static_cast<size_t>(rb) * 32);
        t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
      }
      reinterpret_cast<float4*>(r1)[grp * 32 + quad] = t;  // (r1 / r2 were last read before the ticket barrier)
      __syncthreads();
      if (threadIdx.x < 128) {
        float total = 0.f;
#pragma unroll
        for (int w = 0; w < kBnThreads / 32; ++w) total += r1[w * 128 + threadIdx.x];
        const int c = slab * 64 + (threadIdx.x & 63);
        if (threadIdx.x < 64) a.acc[c] = total;
        else a.acc[a.C + c] = total * a.invstd[c];
      }
      if (threadIdx.x == 0) tickets[slab] = 0u;  // ready for the next launch
    }
    return;
  }
  // thread t < C handles channel t (loop if C > blockDim)
  for (int c = threadIdx.x; c < a.C; c += blockDim.x) {
    const int ctx = c >> 3, cj = c & 7;
    float t1 = 0.f, t2 = 0.f;
    for (int r = 0; r < rows_per_block; ++r) {
      t1 += r1[(r * vecC + ctx) * 8 + cj];
      t2 += r2[(r * vecC + ctx) * 8 + cj];
    }
    atomicAdd(a.acc + c, t1);
    atomicAdd(a.acc + a.C + c, t2 * a.invstd[c]);
  }
}

// Pass 2: dy = gamma * invstd * (dz - mean(dz) - xhat * mean(dz * xhat)); block 0 also emits dgamma/dbeta.
// kGate: gate applied to `dout` here; kGateNone also covers the case where the reduce pass already stored the gated
// gradient in dres (kFromDres: read that back instead of dout).  U rows per thread in flight.
template <int kGate, bool kFromDres, int U>
__global__ void __launch_bounds__(kBnThreads, U >= 4 ? 2 : 3) bn_bwd_apply_kernel(const BnBwdArgs a) {
  // Thread (tx, ty): tx owns 8 consecutive channels for the whole kernel, so the per-channel constants live in
  // registers and the streaming loop is pure load / FMA / store:  dy = A*dz + B*y + D  with
  //   A = gamma*invstd,  B = -A*invstd*mean(dz*xhat),  D = -A*mean(dz) - B*mean.
  const int vecC = a.C >> 3;
  const int tx = threadIdx.x % vecC;
  const int ty = threadIdx.x / vecC;
  const int rows_per_block = blockDim.x / vecC;
  const int c0 = tx << 3;
  const float invM = 1.f / static_cast<float>(a.M);
  float kA[8], kB[8], kD[8], sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + j;
    const float sdz = a.acc[c], sdzx = a.acc[a.C + c];
    const float g = a.gamma ? a.gamma[c] : 1.f;
    const float istd = a.invstd[c], mean = a.mean[c];
    sc[j] = g * istd;
    sh[j] = (kGate == kGateY && a.beta ? a.beta[c] : 0.f) - mean * g * istd;
    kA[j] = g * istd;
    kB[j] = -kA[j] * istd * (sdzx * invM);
    kD[j] = -kA[j] * (sdz * invM) - kB[j] * mean;
    if (blockIdx.x == 0 && ty == 0) {
      if (a.dgamma) a.dgamma[c] += sdzx;
      if (a.dbeta) a.dbeta[c] += sdz;
    }
  }
  if (ty >= rows_per_block) return;
  const long long stride = static_cast<long long>(gridDim.x) * rows_per_block;
  const __nv_bfloat16* gsrc = kFromDres ? a.dres : a.dout;
  const long long ldg = kFromDres ? a.ld_dres : a.ld_dout;
  const bool write_dres = !kFromDres && a.dres != nullptr;
  for (long long row0 = blockIdx.x * static_cast<long long>(rows_per_block) + ty; row0 < a.M; row0 += U * stride) {
    // all loads of the U rows are issued before the first use
    uint4 g[U], yv[U], av[kGate == kGateAct ? U : 1];
    uint32_t mb[kGate == kGateBits ? U : 1];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * stride;
      if (row < a.M) {
        g[u] = *reinterpret_cast<const uint4*>(gsrc + row * ldg + c0);
        yv[u] = *reinterpret_cast<const uint4*>(a.y + row * a.ldy + c0);
        if (kGate == kGateBits) mb[u] = a.relu_mask[row * vecC + tx];
        if (kGate == kGateAct) av[u] = *reinterpret_cast<const uint4*>(a.act + row * a.ld_act + c0);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long r = row0 + u * stride;
      if (r >= a.M) break;
      uint4 gate = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
      if (kGate == kGateBits) gate = expand_relu_bits(mb[u]);
      if (kGate == kGateAct) gate = av[u];
      const uint32_t gi[4] = {g[u].x, g[u].y, g[u].z, g[u].w}, yi[4] = {yv[u].x, yv[u].y, yv[u].z, yv[u].w},
                     ai[4] = {gate.x, gate.y, gate.z, gate.w};
      uint32_t oi[4], ri[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool m0 = kGate == kGateY ? fmaf(bf16lo(yi[j]), sc[2 * j], sh[2 * j]) > 0.f : bf16lo(ai[j]) > 0.f;
        const bool m1 = kGate == kGateY ? fmaf(bf16hi(yi[j]), sc[2 * j + 1], sh[2 * j + 1]) > 0.f : bf16hi(ai[j]) > 0.f;
        const float d0 = m0 ? bf16lo(gi[j]) : 0.f;
        const float d1 = m1 ? bf16hi(gi[j]) : 0.f;
        oi[j] = pack_bf16x2(fmaf(kA[2 * j], d0, fmaf(kB[2 * j], bf16lo(yi[j]), kD[2 * j])),
                            fmaf(kA[2 * j + 1], d1, fmaf(kB[2 * j + 1], bf16hi(yi[j]), kD[2 * j + 1])));
        ri[j] = pack_bf16x2(d0, d1);
      }
      *reinterpret_cast<uint4*>(a.dy + r * a.ld_dy + c0) = make_uint4(oi[0], oi[1], oi[2], oi[3]);
      if (write_dres) *reinterpret_cast<uint4*>(a.dres + r * a.ld_dres + c0) = make_uint4(ri[0], ri[1], ri[2], ri[3]);
    }
  }
}

// Standalone per-channel statistics (for inputs that did not come out of the igemm epilogue).
__global__ void __launch_bounds__(kBnThreads) bn_stats_kernel(const __nv_bfloat16* y, long long M, int C, long long ld,
                                                              float* sum, float* sqsum) {
  const int vecC = C >> 3;
  const int tx = threadIdx.x % vecC, ty = threadIdx.x / vecC;
  const int rows_per_block = blockDim.x / vecC;
  const int c0 = tx << 3;
  float s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
  if (ty < rows_per_block) {
    constexpr int U = 4;  // rows in flight per thread (a read-only stream: see bn_bwd_reduce_kernel)
    const long long stride = static_cast<long long>(gridDim.x) * rows_per_block;
    for (long long row0 = blockIdx.x * static_cast<long long>(rows_per_block) + ty; row0 < M; row0 += U * stride) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (row0 + u * stride < M) v[u] = *reinterpret_cast<const uint4*>(y + (row0 + u * stride) * ld + c0);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (row0 + u * stride >= M) break;
        const uint32_t vi[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float a0 = bf16lo(vi[j]), a1 = bf16hi(vi[j]);
          s1[2 * j] += a0;
          s1[2 * j + 1] += a1;
          s2[2 * j] = fmaf(a0, a0, s2[2 * j]);
          s2[2 * j + 1] = fmaf(a1, a1, s2[2 * j + 1]);
        }
      }
    }
  }
  extern __shared__ float red[];
  float* r1 = red;
  float* r2 = red + blockDim.x * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    r1[threadIdx.x * 8 + j] = s1[j];
    r2[threadIdx.x * 8 + j] = s2[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int ctx = c >> 3, cj = c & 7;
    float t1 = 0.f, t2 = 0.f;
    for (int r = 0; r < rows_per_block; ++r) {
      t1 += r1[(r * vecC + ctx) * 8 + cj];
      t2 += r2[(r * vecC + ctx) * 8 + cj];
    }
    atomicAdd(sum + c, t1);
    if (sqsum != nullptr) atomicAdd(sqsum + c, t2);
  }
}

// Row-streaming kernels run as ONE wave of resident blocks that stride over the rows (`per_sm` = resident blocks per
// SM at the kernel's register count), with at least 8 rows per thread: the per-thread prologue (per-channel constants
// from ~5 small arrays) costs as much as streaming several rows, which made the small late-layer launches
// latency-bound.
static int wave_grid(long long rows, int rows_per_block, int per_sm) {
  long long blocks = (rows + 8LL * rows_per_block - 1) / (8LL * rows_per_block);
  const long long cap = 148LL * per_sm;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : static_cast<int>(blocks);
}

}  // namespace dmpi

using namespace dmpi;


static int gate_kind(const dmpi::BnBwdArgs& a) {
  if (a.relu_from_y) return dmpi::kGateY;
  if (a.relu_mask) return dmpi::kGateBits;
  if (a.act) return dmpi::kGateAct;
  return dmpi::kGateNone;
}

// Four rows in flight per thread in both passes (measured per ResNet-50 shape with bench/bn_bwd_shapes.py: 2 rows
// -12 %, 8 rows in the reduce pass -6 %; the act-gated apply variant keeps 2 rows for its third load stream).
template <bool kDet>
static void launch_bwd_reduce(const dmpi::BnBwdArgs& a, float* part, dim3 grid, int threads, cudaStream_t s) {
  using namespace dmpi;
  const size_t smem = static_cast<size_t>(threads) * 16 * sizeof(float);
  switch (gate_kind(a)) {
    case kGateY: bn_bwd_reduce_kernel<kGateY, kDet, 4><<<grid, threads, smem, s>>>(a, part); break;
    case kGateBits: bn_bwd_reduce_kernel<kGateBits, kDet, 4><<<grid, threads, smem, s>>>(a, part); break;
    case kGateAct: bn_bwd_reduce_kernel<kGateAct, kDet, 4><<<grid, threads, smem, s>>>(a, part); break;
    default: bn_bwd_reduce_kernel<kGateNone, kDet, 4><<<grid, threads, smem, s>>>(a, part); break;
  }
}

static void launch_bwd_apply(const dmpi::BnBwdArgs& a, int rows_per_block, int threads, cudaStream_t s) {
  using namespace dmpi;
  const int gate = gate_kind(a);
  // the reduce pass already gated dout into dres: read that back, no gate needed here
  const bool from_dres = a.dres != nullptr && (gate == kGateBits || gate == kGateAct);
  // one wave: every block is resident (2 per SM for the 4-row variants, 3 for the 2-row one) and strides over the rows
  const int grid = wave_grid(a.M, rows_per_block, (gate == kGateAct && !from_dres) ? 3 : 2);
  if (from_dres) bn_bwd_apply_kernel<kGateNone, true, 4><<<grid, threads, 0, s>>>(a);
  else if (gate == kGateY) bn_bwd_apply_kernel<kGateY, false, 4><<<grid, threads, 0, s>>>(a);
  else if (gate == kGateBits) bn_bwd_apply_kernel<kGateBits, false, 4><<<grid, threads, 0, s>>>(a);
  else if (gate == kGateAct) bn_bwd_apply_kernel<kGateAct, false, 2><<<grid, threads, 0, s>>>(a);
  else bn_bwd_apply_kernel<kGateNone, false, 4><<<grid, threads, 0, s>>>(a);
}

extern "C" int dmpi_bn_forward_mask(const void* y, const void* res, void* out, long long M, int C, long long ldy,
                                    long long ldr, long long ldo, const float* sum, const float* sqsum,
                                    const float* gamma, const float* beta, float* save_mean, float* save_invstd,
                                    float* running_mean, float* running_var, long long* num_batches_tracked, float eps,
                                    float momentum, int training, int relu, void* relu_mask, void* stream);

extern "C" int dmpi_bn_forward(const void* y, const void* res, void* out, long long M, int C, long long ldy,
                               long long ldr, long long ldo, const float* sum, const float* sqsum, const float* gamma,
                               const float* beta, float* save_mean, float* save_invstd, float* running_mean,
                               float* running_var, long long* num_batches_tracked, float eps, float momentum,
                               int training, int relu, void* stream) {
  return dmpi_bn_forward_mask(y, res, out, M, C, ldy, ldr, ldo, sum, sqsum, gamma, beta, save_mean, save_invstd,
                              running_mean, running_var, num_batches_tracked, eps, momentum, training, relu, nullptr,
                              stream);
}

// `relu_mask` (optional, [M, C/8] bytes): receives one bit per output element, set where the stored value is > 0.
extern "C" int dmpi_bn_forward_mask(const void* y, const void* res, void* out, long long M, int C, long long ldy,
                                    long long ldr, long long ldo, const float* sum, const float* sqsum,
                                    const float* gamma, const float* beta, float* save_mean, float* save_invstd,
                                    float* running_mean, float* running_var, long long* num_batches_tracked, float eps,
                                    float momentum, int training, int relu, void* relu_mask, void* stream) {
  if (C % 8 != 0 || C > kMaxC) return -1;
  BnFwdArgs a;
  a.relu_mask = static_cast<uint8_t*>(relu_mask);
  a.y = static_cast<const __nv_bfloat16*>(y);
  a.res = static_cast<const __nv_bfloat16*>(res);
  a.out = static_cast<__nv_bfloat16*>(out);
  a.M = M;
  a.C = C;
  a.ldy = ldy;
  a.ldr = ldr;
  a.ldo = ldo;
  a.sum = sum;
  a.sqsum = sqsum;
  a.gamma = gamma;
  a.beta = beta;
  a.save_mean = save_mean;
  a.save_invstd = save_invstd;
  a.running_mean = running_mean;
  a.running_var = running_var;
  a.num_batches_tracked = num_batches_tracked;
  a.eps = eps;
  a.momentum = momentum;
  a.training = training;
  a.relu = relu;
  const int vecC = C / 8;
  if (vecC > kBnThreads) return -3;
  const int threads = (kBnThreads / vecC) * vecC;  // a whole number of rows per block
  const int rows_per_block = threads / vecC;
  // one wave of resident blocks (3 per SM at 72-80 registers), at least 8 rows per thread, 4 rows in flight
  // (bench/bn_fwd_shapes.py: 2 rows x 8 blocks per SM and 4 rows x 6 are within 3 % of this)
  const int grid = wave_grid(M, rows_per_block, 3);
  const size_t smem = 2 * static_cast<size_t>(C) * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.res) bn_fwd_kernel<true, 4><<<grid, threads, smem, st>>>(a);
  else bn_fwd_kernel<false, 4><<<grid, threads, smem, st>>>(a);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

// Launches the reduce and apply passes.  `acc` is scratch of 2*C floats that must be zero on entry.
// Extended form: with relu_from_y != 0 (and act == NULL) the ReLU mask is recomputed from y, gamma, beta and the saved
// statistics instead of being read from the forward output.
extern "C" int dmpi_bn_backward_ex(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                   int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                   long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                   const float* beta, int relu_from_y, float* acc, float* dgamma, float* dbeta,
                                   void* stream);

extern "C" int dmpi_bn_backward(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                float* acc, float* dgamma, float* dbeta, void* stream) {
  return dmpi_bn_backward_ex(dout, y, act, dy, dres, M, C, ld_dout, ldy, ld_act, ld_dy, ld_dres, mean, invstd, gamma,
                             nullptr, 0, acc, dgamma, dbeta, stream);
}

extern "C" int dmpi_bn_backward_mask(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                     int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                     long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                     const float* beta, int relu_from_y, const void* relu_mask, float* acc,
                                     float* dgamma, float* dbeta, void* stream);

extern "C" int dmpi_bn_backward_ex(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                   int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                   long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                   const float* beta, int relu_from_y, float* acc, float* dgamma, float* dbeta,
                                   void* stream) {
  return dmpi_bn_backward_mask(dout, y, act, dy, dres, M, C, ld_dout, ldy, ld_act, ld_dy, ld_dres, mean, invstd, gamma,
                               beta, relu_from_y, nullptr, acc, dgamma, dbeta, stream);
}

extern "C" int dmpi_bn_backward_det(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                    int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                    long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                    const float* beta, int relu_from_y, const void* relu_mask, float* acc,
                                    float* dgamma, float* dbeta, void* det_scratch, long long det_bytes, void* stream);

// `relu_mask` (optional): the bit mask dmpi_bn_forward_mask wrote; used as the ReLU gate instead of `act`.
extern "C" int dmpi_bn_backward_mask(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                     int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                     long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                     const float* beta, int relu_from_y, const void* relu_mask, float* acc,
                                     float* dgamma, float* dbeta, void* stream) {
  return dmpi_bn_backward_det(dout, y, act, dy, dres, M, C, ld_dout, ldy, ld_act, ld_dy, ld_dres, mean, invstd, gamma,
                              beta, relu_from_y, relu_mask, acc, dgamma, dbeta, nullptr, 0, stream);
}

// `det_scratch` (optional, zero-initialised, not shared with a concurrently running launch; 1 MiB covers every shape):
// with it the channel reductions are added in a fixed order (bit-reproducible) and `acc` need not be zeroed; without it
// they use fp32 atomics onto the zeroed `acc`.
extern "C" int dmpi_bn_backward_det(const void* dout, const void* y, const void* act, void* dy, void* dres, long long M,
                                    int C, long long ld_dout, long long ldy, long long ld_act, long long ld_dy,
                                    long long ld_dres, const float* mean, const float* invstd, const float* gamma,
                                    const float* beta, int relu_from_y, const void* relu_mask, float* acc,
                                    float* dgamma, float* dbeta, void* det_scratch, long long det_bytes, void* stream) {
  if (C % 8 != 0) return -1;
  const int vecC = C / 8;
  BnBwdArgs a;
  a.dout = static_cast<const __nv_bfloat16*>(dout);
  a.y = static_cast<const __nv_bfloat16*>(y);
  a.act = static_cast<const __nv_bfloat16*>(act);
  a.dy = static_cast<__nv_bfloat16*>(dy);
  a.dres = static_cast<__nv_bfloat16*>(dres);
  a.M = M;
  a.C = C;
  a.ld_dout = ld_dout;
  a.ldy = ldy;
  a.ld_act = ld_act;
  a.ld_dy = ld_dy;
  a.ld_dres = ld_dres;
  a.mean = mean;
  a.invstd = invstd;
  a.gamma = gamma;
  a.acc = acc;
  a.dgamma = dgamma;
  a.dbeta = dbeta;
  a.beta = beta;
  a.relu_mask = static_cast<const uint8_t*>(relu_mask);
  if (a.relu_mask) a.act = nullptr;
  a.relu_from_y = (relu_from_y && act == nullptr && relu_mask == nullptr) ? 1 : 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // reduce: block = vecC * rows_per_block threads (vecC <= 256 for C <= 2048); two blocks of 256 threads are resident
  // per SM (launch bounds), and the grid is one such wave at most so that no block waits for a second wave
  const long long wave = 148LL * 2;
  int threads = kBnThreads;
  if (vecC > threads) return -3;
  threads = (threads / vecC) * vecC;
  const int rows_per_block = threads / vecC;
  long long blocks = (M + rows_per_block * 8 - 1) / (rows_per_block * 8);
  if (blocks > wave) blocks = wave;
  if (blocks < 1) blocks = 1;
  bool det = false;
  if (det_scratch != nullptr && C % 64 == 0) {
    const int slabs = C / 64;
    long long rb = (M + 32 * 8 - 1) / (32 * 8);  // >= 8 rows per thread
    long long cap = wave / slabs;
    if (cap < 1) cap = 1;
    if (rb > cap) rb = cap;
    if (rb < 1) rb = 1;
    if ((kDetHdr + static_cast<long long>(slabs) * rb * 128) * 4 <= det_bytes && slabs <= kDetHdr) {
      det = true;
      const dim3 grid(slabs, static_cast<unsigned>(rb));
      launch_bwd_reduce<true>(a, static_cast<float*>(det_scratch), grid, kBnThreads, s);
    }
  }
  if (!det) launch_bwd_reduce<false>(a, nullptr, dim3(static_cast<unsigned>(blocks)), threads, s);
  launch_bwd_apply(a, rows_per_block, threads, s);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_bn_stats(const void* y, long long M, int C, long long ld, float* sum, float* sqsum, void* stream) {
  if (C % 8 != 0) return -1;
  const int vecC = C / 8;
  int threads = kBnThreads;
  if (vecC > threads) return -3;
  threads = (threads / vecC) * vecC;
  const int rows_per_block = threads / vecC;
  bn_stats_kernel<<<wave_grid(M, rows_per_block, 4), threads, threads * 16 * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), M, C, ld, sum, sqsum);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

// ------------------------------------------------------------------------------------------------------------
// BatchNorm backward for callers that must not allocate.
//
// dmpi_bn_backward() above needs 2*C floats of zeroed scratch from its caller.  Some callers have nowhere to get it:
//   * a host program that gave the whole GPU to its own memory pool at start-up, so a later cudaMalloc simply fails;
//   * code on a latency-critical path, where cudaMalloc / cudaFree would stall every stream of the device;
//   * a plug-in running inside somebody else's callback that may neither call the CUDA allocator nor borrow the host
//     framework's one (so no torch tensors either).
// For them the scratch lives inside the library image instead: one file-scope __device__ array, 16 KiB, present on
// every GPU that loads libdmpi_cuda.so for as long as the process lives.  The limit that follows is C <= 2048
// channels (the largest BatchNorm in ResNet-50/101/152 and in the UNet has 2048).
//
// Things a caller has to know:
//   * The first call on each device looks the array's device address up (cudaGetSymbolAddress) and remembers it.
//     Make that first call during set-up, with the right device current and before you start capturing any stream.
//   * Every call clears the 2*C floats it is about to use (an async memset on `stream`) and leaves them dirty.
//   * All calls on one GPU share the one array.  Two calls whose kernels can overlap on the GPU (different streams
//     without an ordering between them, or different host threads doing the same) corrupt each other's sums; keep
//     calls on a device stream-ordered.
//   * Nothing is allocated or freed: no cudaMalloc / cudaFree / cudaMallocAsync, no library workspace.
// Return codes: 0 ok; -1 C not a positive multiple of 8; -2 launch failure; -4 C above the 2048-channel limit;
// -5 / -6 the device or the scratch symbol could not be resolved (a message is also written to stderr).
#include <stdio.h>

#include <mutex>

namespace dmpi {
constexpr int kNoAllocMaxChannels = 2048;
constexpr int kNoAllocMaxDevices = 64;
// Host code must never use this name as a pointer: on the host it is only a handle for cudaGetSymbolAddress.
static __device__ __align__(256) float g_bn_bwd_static_scratch[2 * kNoAllocMaxChannels];
static std::mutex g_noalloc_mutex;
static float* g_noalloc_scratch_addr[kNoAllocMaxDevices] = {nullptr};
}  // namespace dmpi

extern "C" int dmpi_bn_backward_noalloc(const void* dout, const void* y, const void* act, void* dy, void* dres,
                                        long long M, int C, long long ld_dout, long long ldy, long long ld_act,
                                        long long ld_dy, long long ld_dres, const float* mean, const float* invstd,
                                        const float* gamma, float* dgamma, float* dbeta, void* stream) {
  if (C <= 0 || C % 8 != 0) return -1;
  if (C > dmpi::kNoAllocMaxChannels) {
    fprintf(stderr, "[dmpi] dmpi_bn_backward_noalloc: C=%d exceeds the built-in scratch (max %d channels); "
                    "use dmpi_bn_backward with caller-provided scratch\n", C, dmpi::kNoAllocMaxChannels);
    return -4;
  }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= dmpi::kNoAllocMaxDevices) {
    fprintf(stderr, "[dmpi] dmpi_bn_backward_noalloc: cannot identify the current device\n");
    return -5;
  }
  float* acc = nullptr;
  {
    std::lock_guard<std::mutex> lock(dmpi::g_noalloc_mutex);
    acc = dmpi::g_noalloc_scratch_addr[dev];
    if (acc == nullptr) {  // first call on this device: resolve the device address of the static array
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, dmpi::g_bn_bwd_static_scratch);
      if (e != cudaSuccess || p == nullptr) {
        fprintf(stderr, "[dmpi] dmpi_bn_backward_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev,
                cudaGetErrorString(e));
        return -6;
      }
      acc = static_cast<float*>(p);
      dmpi::g_noalloc_scratch_addr[dev] = acc;
    }
  }
  // the reduce pass accumulates with atomics, so the part of the shared array used by this call must start at zero
  if (cudaMemsetAsync(acc, 0, 2 * static_cast<size_t>(C) * sizeof(float), static_cast<cudaStream_t>(stream)) !=
      cudaSuccess)
    return -2;
  return dmpi_bn_backward(dout, y, act, dy, dres, M, C, ld_dout, ldy, ld_act, ld_dy, ld_dres, mean, invstd, gamma, acc,
                          dgamma, dbeta, stream);
}
