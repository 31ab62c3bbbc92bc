// NVLink peer-to-peer collectives for one NVSwitch domain.
//
// Every rank owns a "symmetric arena" (csrc/p2p/arena.cu) that all peers map; the kernels below read and
// write peers' arenas directly with 16-byte loads/stores.  No NCCL anywhere on these paths
// (the reference reaches ncclAllReduce / ncclBroadcast / ncclSend through torch.distributed:
//  ref:pytorch/resnet/main.py:44-46,131; ref:pytorch/hello_world/hello_world.py:26,29).
//
//   allreduce_twoshot : in place.  Rank r pulls slice r of every peer's bucket, sums in fixed rank order (so all
//                       ranks end bit-identical), applies the fused scale (1/world_size), and pushes the reduced
//                       slice into every peer's bucket.  2 barriers, 2*(W-1)/W*S bytes per direction.
//   allreduce_oneshot : latency path.  Every rank pushes its whole vector into a per-source staging slot on every
//                       peer, ONE barrier, then each rank reduces its own W local slots.  Staging is double
//                       buffered by barrier-epoch parity, so no trailing barrier is needed.
//   broadcast         : root pushes to every peer (wrap-time parameter sync, per-iteration BatchNorm buffer sync).
//   send / recv       : put-with-flag into a per-pair mailbox, acknowledged by the receiver.
//
// Synchronisation: per-block monotonically increasing epochs on a signal pad; flag stores are
// st.release.sys after a CTA barrier, waits are ld.acquire.sys.  Block b of rank r only ever
// synchronises with block b of the peers, so no grid-wide barrier is required.  Waits are bounded
// (trap after DMPI_COMM_TIMEOUT_S seconds, default 600, 0 = never) so a dead peer cannot hang the GPU forever
//  while a peer that is legitimately late — rank 0 alone evaluating and checkpointing between epochs,
// ref:pytorch/resnet/main.py:136-139 — is simply waited for.
//
// Single-GPU emulation (tests): blockIdx.y selects a virtual rank; all W "ranks" are buffers of one device
// and run inside ONE launch, as required when ranks must wait on each other on a single GPU.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../common/ptx.cuh"
#include "../../../include/dmpi_comm.h"

namespace dmpi {

constexpr int kCommThreads = 512;
constexpr int kInFlight = 4;  // 16-byte vectors per thread and peer issued before the first use (NVLink latency ~2 us)

__device__ unsigned long long g_comm_timeout_ns = 600ull * 1000000000ull;  // 0: wait forever

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float4* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 ld_peer_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_u4(uint4* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ void spin_until_ge(const uint32_t* flag, uint32_t target) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (static_cast<int32_t>(ld_acquire_sys(flag) - target) < 0) {
    if ((++spins & 1023u) == 0) {
      const uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (g_comm_timeout_ns != 0ull && now - t0 > g_comm_timeout_ns) __trap();
    }
  }
}

// All threads of the block call this.  Signals epoch `e` to block `blk` of every rank and waits for all of them.
__device__ __forceinline__ void block_barrier(const dmpi_comm_ptrs& p, int rank, int blk, uint32_t e) {
  __syncthreads();
  if (static_cast<int>(threadIdx.x) < p.world) {
    const int peer = threadIdx.x;
    st_release_sys(p.sig[peer] + blk * DMPI_MAX_RANKS + rank, e);
    spin_until_ge(p.sig[rank] + blk * DMPI_MAX_RANKS + peer, e);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------------ all-reduce
// MAXW bounds the world size at compile time so that the loads of kInFlightBytes / MAXW vectors per peer can be held
// in registers at once: a peer load takes ~2 us over NVLink, and with one vector per thread in flight (the first
// version of this kernel) a W = 2 launch of 32 blocks kept only 0.5 MB on the wire — latency-bound at 0.8x NCCL.
template <bool kBf16, int MAXW>
__global__ void __launch_bounds__(kCommThreads) allreduce_twoshot_kernel(const dmpi_comm_ptrs p, long long off_bytes,
                                                                         long long nvec, float scale) {
  constexpr int U = 16 / MAXW;  // vectors per thread and peer in flight: 8 (W <= 2), 4 (W <= 4), 2 (W <= 8)
  const int rank = p.rank_base + blockIdx.y;
  const int blk = blockIdx.x;
  const int W = p.world;
  uint32_t e = p.epoch[rank][blk];
  block_barrier(p, rank, blk, e + 1);

  const long long v_begin = nvec * rank / W, v_end = nvec * (rank + 1) / W;
  uint4* bufs[MAXW];
#pragma unroll
  for (int q = 0; q < MAXW; ++q) bufs[q] = reinterpret_cast<uint4*>(static_cast<char*>(p.data[q < W ? q : 0]) + off_bytes);

  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = v_begin + static_cast<long long>(blk) * blockDim.x + threadIdx.x; i0 < v_end; i0 += stride * U) {
    uint4 in[U][MAXW];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < v_end) {
#pragma unroll
        for (int q = 0; q < MAXW; ++q)
          if (q < W) in[u][q] = ld_peer_u4(bufs[q] + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= v_end) break;
      uint4 out;
      if (kBf16) {
        float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int q = 0; q < MAXW; ++q)
          if (q < W) {
            const uint32_t w4[4] = {in[u][q].x, in[u][q].y, in[u][q].z, in[u][q].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              a[2 * j] += bf16lo(w4[j]);
              a[2 * j + 1] += bf16hi(w4[j]);
            }
          }
        out.x = pack_bf16x2(a[0] * scale, a[1] * scale);
        out.y = pack_bf16x2(a[2] * scale, a[3] * scale);
        out.z = pack_bf16x2(a[4] * scale, a[5] * scale);
        out.w = pack_bf16x2(a[6] * scale, a[7] * scale);
      } else {
        float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int q = 0; q < MAXW; ++q)
          if (q < W) {  // fixed rank order: every rank ends with bit-identical sums
            a[0] += __uint_as_float(in[u][q].x);
            a[1] += __uint_as_float(in[u][q].y);
            a[2] += __uint_as_float(in[u][q].z);
            a[3] += __uint_as_float(in[u][q].w);
          }
        out.x = __float_as_uint(a[0] * scale);
        out.y = __float_as_uint(a[1] * scale);
        out.z = __float_as_uint(a[2] * scale);
        out.w = __float_as_uint(a[3] * scale);
      }
#pragma unroll
      for (int q = 0; q < MAXW; ++q)
        if (q < W) st_peer_u4(bufs[q] + i, out);
    }
  }
  __threadfence_system();
  block_barrier(p, rank, blk, e + 2);
  if (threadIdx.x == 0) p.epoch[rank][blk] = e + 2;
}

// fp32 data; the wire/staging format is fp32 (kWireBf16 = false) or bf16 (fused cast, half the NVLink bytes).
template <bool kWireBf16>
__global__ void __launch_bounds__(kCommThreads) allreduce_oneshot_kernel(const dmpi_comm_ptrs p, long long off_bytes,
                                                                         long long nvec, float scale,
                                                                         long long slot_bytes) {
  const int rank = p.rank_base + blockIdx.y;
  const int blk = blockIdx.x;
  const int W = p.world;
  const uint32_t e = p.epoch[rank][blk];
  // Staging parity comes from a counter of one-shot launches, identical in every block and on every rank (launches
  // on an arena are stream-ordered), NOT from the per-block epoch: blocks that sat out a smaller earlier call would
  // otherwise disagree with their neighbours about which half a given element was staged in, and a fast rank's next
  // push could land in the half a slower peer is still reducing.
  const uint32_t calls = p.epoch[rank][DMPI_EPOCH_ONESHOT_CALLS];
  const int parity = calls & 1u;
  float4* mine = reinterpret_cast<float4*>(static_cast<char*>(p.data[rank]) + off_bytes);
  // staging layout per rank: [parity][source rank][slot_bytes]
  const long long slot_off = (static_cast<long long>(parity) * DMPI_MAX_RANKS + rank) * slot_bytes;

  for (long long i = static_cast<long long>(blk) * blockDim.x + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = mine[i];
    if (kWireBf16) {
      uint2 w;
      w.x = pack_bf16x2(v.x, v.y);
      w.y = pack_bf16x2(v.z, v.w);
#pragma unroll
      for (int q = 0; q < DMPI_MAX_RANKS; ++q)
        if (q < W) {
          uint2* dst = reinterpret_cast<uint2*>(static_cast<char*>(p.stage[q]) + slot_off) + i;
          asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(w.x), "r"(w.y) : "memory");
        }
    } else {
#pragma unroll
      for (int q = 0; q < DMPI_MAX_RANKS; ++q)
        if (q < W) st_peer_f4(reinterpret_cast<float4*>(static_cast<char*>(p.stage[q]) + slot_off) + i, v);
    }
  }
  __threadfence_system();
  block_barrier(p, rank, blk, e + 1);

  const char* stage_mine = static_cast<const char*>(p.stage[rank]) +
                           static_cast<long long>(parity) * DMPI_MAX_RANKS * slot_bytes;
  for (long long i = static_cast<long long>(blk) * blockDim.x + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < DMPI_MAX_RANKS; ++q)
      if (q < W) {
        if (kWireBf16) {
          uint2 w;
          asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];"
                       : "=r"(w.x), "=r"(w.y)
                       : "l"(reinterpret_cast<const uint2*>(stage_mine + q * slot_bytes) + i)
                       : "memory");
          a[0] += bf16lo(w.x);
          a[1] += bf16hi(w.x);
          a[2] += bf16lo(w.y);
          a[3] += bf16hi(w.y);
        } else {
          const float4 w = ld_peer_f4(reinterpret_cast<const float4*>(stage_mine + q * slot_bytes) + i);
          a[0] += w.x;
          a[1] += w.y;
          a[2] += w.z;
          a[3] += w.w;
        }
      }
    mine[i] = make_float4(a[0] * scale, a[1] * scale, a[2] * scale, a[3] * scale);
  }
  if (threadIdx.x == 0) {
    p.epoch[rank][blk] = e + 1;
    // the block that finishes last advances the call counter (a block that has not started yet must still read the
    // old value), and re-arms the ticket for the next launch
    if (atomicAdd(&p.epoch[rank][DMPI_EPOCH_ONESHOT_DONE], 1u) == gridDim.x - 1) {
      p.epoch[rank][DMPI_EPOCH_ONESHOT_DONE] = 0u;
      p.epoch[rank][DMPI_EPOCH_ONESHOT_CALLS] = calls + 1;
    }
  }
}

// NVLS: the reduction happens inside the NVSwitch.  `mc_data` is the multicast mapping of the data region; a
// multimem.ld_reduce on it returns the sum over every rank's copy, a multimem.st writes every rank's copy.  Rank r
// handles slice r, so each element crosses this GPU's link once inbound (already reduced) and once outbound.
template <bool kBf16>
__global__ void __launch_bounds__(kCommThreads) allreduce_nvls_kernel(const dmpi_comm_ptrs p, long long off_bytes,
                                                                      long long nvec, float scale) {
  const int rank = p.rank_base;
  const int blk = blockIdx.x;
  const int W = p.world;
  const uint32_t e = p.epoch[rank][blk];
  block_barrier(p, rank, blk, e + 1);
  char* mc = static_cast<char*>(p.mc_data) + off_bytes;
  const long long v_begin = nvec * rank / W, v_end = nvec * (rank + 1) / W;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  // kInFlight in-switch reductions per thread are issued before the first result is used (a multimem.ld_reduce is a
  // round trip through the NVSwitch, ~3 us)
  for (long long i0 = v_begin + static_cast<long long>(blk) * blockDim.x + threadIdx.x; i0 < v_end;
       i0 += stride * kInFlight) {
    if (kBf16) {
      uint32_t r[kInFlight][4];
#pragma unroll
      for (int u = 0; u < kInFlight; ++u) {
        const long long i = i0 + u * stride;
        if (i < v_end)
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                       : "=r"(r[u][0]), "=r"(r[u][1]), "=r"(r[u][2]), "=r"(r[u][3])
                       : "l"(mc + i * 16)
                       : "memory");
      }
#pragma unroll
      for (int u = 0; u < kInFlight; ++u) {
        const long long i = i0 + u * stride;
        if (i >= v_end) break;
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = pack_bf16x2(bf16lo(r[u][j]) * scale, bf16hi(r[u][j]) * scale);
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + i * 16),
                     "f"(__uint_as_float(o[0])), "f"(__uint_as_float(o[1])), "f"(__uint_as_float(o[2])),
                     "f"(__uint_as_float(o[3]))
                     : "memory");
      }
    } else {
      float r[kInFlight][4];
#pragma unroll
      for (int u = 0; u < kInFlight; ++u) {
        const long long i = i0 + u * stride;
        if (i < v_end)
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(r[u][0]), "=f"(r[u][1]), "=f"(r[u][2]), "=f"(r[u][3])
                       : "l"(mc + i * 16)
                       : "memory");
      }
#pragma unroll
      for (int u = 0; u < kInFlight; ++u) {
        const long long i = i0 + u * stride;
        if (i >= v_end) break;
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + i * 16),
                     "f"(r[u][0] * scale), "f"(r[u][1] * scale), "f"(r[u][2] * scale), "f"(r[u][3] * scale)
                     : "memory");
      }
    }
  }
  __threadfence_system();
  block_barrier(p, rank, blk, e + 2);
  if (threadIdx.x == 0) p.epoch[rank][blk] = e + 2;
}

// ------------------------------------------------------------------------------------------------ broadcast
__global__ void __launch_bounds__(kCommThreads) broadcast_kernel(const dmpi_comm_ptrs p, long long off_bytes,
                                                                 long long nvec, int root) {
  const int rank = p.rank_base + blockIdx.y;
  const int blk = blockIdx.x;
  const int W = p.world;
  const uint32_t e = p.epoch[rank][blk];
  block_barrier(p, rank, blk, e + 1);  // every receiver is past its last use of the destination
  if (rank == root) {
    const uint4* src = reinterpret_cast<const uint4*>(static_cast<const char*>(p.data[root]) + off_bytes);
    for (long long i = static_cast<long long>(blk) * blockDim.x + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
      const uint4 v = src[i];
#pragma unroll
      for (int q = 0; q < DMPI_MAX_RANKS; ++q)
        if (q < W && q != root) st_peer_u4(reinterpret_cast<uint4*>(static_cast<char*>(p.data[q]) + off_bytes) + i, v);
    }
    __threadfence_system();
  }
  block_barrier(p, rank, blk, e + 2);
  if (threadIdx.x == 0) p.epoch[rank][blk] = e + 2;
}

// ------------------------------------------------------------------------------------------------ send / recv
// Mailbox of rank d: [source rank][mailbox_bytes]; flags: seq[d][s] written by s, ack[s][d] written by d.
__global__ void __launch_bounds__(kCommThreads) send_kernel(const uint4* __restrict__ src, uint4* peer_mailbox,
                                                            long long nvec, uint32_t* peer_seq, const uint32_t* my_ack,
                                                            uint32_t seq) {
  if (threadIdx.x == 0) spin_until_ge(my_ack, seq - 1);  // receiver consumed the previous message
  __syncthreads();
  for (long long i = threadIdx.x; i < nvec; i += blockDim.x) st_peer_u4(peer_mailbox + i, src[i]);
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) st_release_sys(peer_seq, seq);
}

__global__ void __launch_bounds__(kCommThreads) recv_kernel(uint4* __restrict__ dst, const uint4* my_mailbox,
                                                            long long nvec, const uint32_t* my_seq, uint32_t* peer_ack,
                                                            uint32_t seq) {
  if (threadIdx.x == 0) spin_until_ge(my_seq, seq);
  __syncthreads();
  for (long long i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = ld_peer_u4(my_mailbox + i);
  __syncthreads();
  if (threadIdx.x == 0) st_release_sys(peer_ack, seq);
}

// Plain grid-stride copy with kInFlight vectors per thread in flight: the measuring stick for the NVLink roofline
// (src or dst is a mapped peer pointer: peer read or peer write bandwidth; both local: the HBM copy rate).
__global__ void __launch_bounds__(kCommThreads) peer_copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                                 long long nvec) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < nvec; i0 += stride * kInFlight) {
    uint4 v[kInFlight];
#pragma unroll
    for (int u = 0; u < kInFlight; ++u)
      if (i0 + u * stride < nvec) v[u] = ld_peer_u4(src + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < kInFlight; ++u)
      if (i0 + u * stride < nvec) st_peer_u4(dst + i0 + u * stride, v[u]);
  }
}

// Emulated pair for single-GPU tests: one launch, block 0 sends and block 1 receives.
__global__ void __launch_bounds__(kCommThreads) sendrecv_emulated_kernel(const uint4* src, uint4* mailbox, uint4* dst,
                                                                         long long nvec, uint32_t* seq_flag,
                                                                         uint32_t* ack_flag, uint32_t seq) {
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) spin_until_ge(ack_flag, seq - 1);
    __syncthreads();
    for (long long i = threadIdx.x; i < nvec; i += blockDim.x) st_peer_u4(mailbox + i, src[i]);
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(seq_flag, seq);
  } else {
    if (threadIdx.x == 0) spin_until_ge(seq_flag, seq);
    __syncthreads();
    for (long long i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = ld_peer_u4(mailbox + i);
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(ack_flag, seq);
  }
}

}  // namespace dmpi

using namespace dmpi;

// DMPI_COMM_TIMEOUT_S is applied once, by the first collective launched in this process (outside stream capture:
// every captured step is preceded by eager warm-up steps).
static void apply_timeout_env() {
  static const bool once = [] {
    if (const char* e = getenv("DMPI_COMM_TIMEOUT_S")) dmpi_comm_set_timeout(atof(e));
    return true;
  }();
  (void)once;
}

extern "C" int dmpi_comm_set_timeout(double seconds) {
  const unsigned long long ns = seconds <= 0.0 ? 0ull : static_cast<unsigned long long>(seconds * 1e9);
  return cudaMemcpyToSymbol(g_comm_timeout_ns, &ns, sizeof(ns)) == cudaSuccess ? 0 : -1;
}

static int clamp_blocks(int nblocks, long long nvec) {
  if (nblocks < 1) nblocks = 1;
  if (nblocks > DMPI_MAX_COMM_BLOCKS) nblocks = DMPI_MAX_COMM_BLOCKS;
  const long long need = (nvec + kCommThreads - 1) / kCommThreads;
  if (need < nblocks) nblocks = need < 1 ? 1 : static_cast<int>(need);
  return nblocks;
}

extern "C" int dmpi_allreduce_twoshot(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int dtype,
                                      float scale, int nblocks, int emulate_ranks, void* stream) {
  if (nbytes % 16 || off_bytes % 16) return -1;
  const long long nvec = nbytes / 16;
  // every rank must use the same block count: it is derived from the message size only
  const int nb = clamp_blocks(nblocks, (nvec + ptrs->world - 1) / ptrs->world);
  dim3 grid(nb, emulate_ranks > 0 ? emulate_ranks : 1);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  apply_timeout_env();
  const bool bf = dtype == DMPI_BF16;
#define DMPI_TWOSHOT(MAXW)                                                                                  \
  do {                                                                                                      \
    if (bf) allreduce_twoshot_kernel<true, MAXW><<<grid, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale); \
    else allreduce_twoshot_kernel<false, MAXW><<<grid, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale);   \
  } while (0)
  if (ptrs->world <= 2) DMPI_TWOSHOT(2);
  else if (ptrs->world <= 4) DMPI_TWOSHOT(4);
  else DMPI_TWOSHOT(8);
#undef DMPI_TWOSHOT
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_allreduce_nvls(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int dtype, float scale,
                                   int nblocks, void* stream) {
  if (nbytes % 16 || off_bytes % 16) return -1;
  if (!ptrs->mc_data) return -4;  // no multicast mapping on this host
  apply_timeout_env();
  const long long nvec = nbytes / 16;
  const int nb = clamp_blocks(nblocks, (nvec + ptrs->world - 1) / ptrs->world);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == DMPI_BF16) allreduce_nvls_kernel<true><<<nb, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale);
  else allreduce_nvls_kernel<false><<<nb, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_allreduce_oneshot(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int wire_bf16,
                                      float scale, long long slot_bytes, int nblocks, int emulate_ranks,
                                      void* stream) {
  if (nbytes % 16 || off_bytes % 16) return -1;
  if (nbytes > slot_bytes) return -3;
  apply_timeout_env();
  const long long nvec = nbytes / 16;
  const int nb = clamp_blocks(nblocks, nvec);
  dim3 grid(nb, emulate_ranks > 0 ? emulate_ranks : 1);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (wire_bf16) allreduce_oneshot_kernel<true><<<grid, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale, slot_bytes);
  else allreduce_oneshot_kernel<false><<<grid, kCommThreads, 0, s>>>(*ptrs, off_bytes, nvec, scale, slot_bytes);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_broadcast(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int root, int nblocks,
                              int emulate_ranks, void* stream) {
  if (nbytes % 16 || off_bytes % 16) return -1;
  const long long nvec = nbytes / 16;
  apply_timeout_env();
  const int nb = clamp_blocks(nblocks, nvec);
  dim3 grid(nb, emulate_ranks > 0 ? emulate_ranks : 1);
  broadcast_kernel<<<grid, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(*ptrs, off_bytes, nvec, root);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_peer_copy(void* dst, const void* src, long long nbytes, int nblocks, void* stream) {
  if (nbytes % 16) return -1;
  if (nblocks < 1) nblocks = 1;
  peer_copy_kernel<<<nblocks, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4*>(dst), static_cast<const uint4*>(src), nbytes / 16);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_send(const void* src, void* peer_mailbox, long long nbytes, uint32_t* peer_seq,
                         const uint32_t* my_ack, uint32_t seq, void* stream) {
  if (nbytes % 16) return -1;
  apply_timeout_env();
  send_kernel<<<1, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(src), static_cast<uint4*>(peer_mailbox), nbytes / 16, peer_seq, my_ack, seq);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_recv(void* dst, const void* my_mailbox, long long nbytes, const uint32_t* my_seq,
                         uint32_t* peer_ack, uint32_t seq, void* stream) {
  if (nbytes % 16) return -1;
  apply_timeout_env();
  recv_kernel<<<1, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4*>(dst), static_cast<const uint4*>(my_mailbox), nbytes / 16, my_seq, peer_ack, seq);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

extern "C" int dmpi_sendrecv_emulated(const void* src, void* mailbox, void* dst, long long nbytes, uint32_t* seq_flag,
                                      uint32_t* ack_flag, uint32_t seq, void* stream) {
  if (nbytes % 16) return -1;
  sendrecv_emulated_kernel<<<2, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(src), static_cast<uint4*>(mailbox), static_cast<uint4*>(dst), nbytes / 16, seq_flag,
      ack_flag, seq);
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
