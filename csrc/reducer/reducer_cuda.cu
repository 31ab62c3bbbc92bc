// CUDA side of the reducer: the launch callback that reducer.cpp invokes when a bucket
// (and all earlier buckets) is complete.  Pure stream/event choreography plus one kernel launch:
//
//   compute stream --record(ready[b])--> comm stream waits --> all-reduce kernel for bucket b
//   after the last bucket: record(done) on the comm stream;  finalize(): compute stream waits on done.
//
// The all-reduce variant is chosen per bucket size (one-shot push below the threshold measured by
// bench/allreduce_sweep.py; above it the NVSwitch in-switch reduction (multimem) when the arena has a multicast
// mapping and world > 2, else two-shot), with the 1/world_size scale fused into the kernel.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <new>
#include <vector>

#include "../../include/dmpi_comm.h"
#include "../../include/dmpi_kernels.h"

struct ReducerCuda {
  dmpi_comm_ptrs ptrs;
  int world = 1;
  std::vector<long long> off, bytes;
  long long off_base = 0;  // offset of the gradient region (= of the lowest bucket)
  std::vector<cudaEvent_t> ready;
  cudaEvent_t done = nullptr;
  cudaEvent_t fwd_done = nullptr, bcast_done = nullptr;  // buffer broadcast on the comm stream (see side_broadcast)
  bool bcast_pending = false;
  bool bcast_captured = false;  // the pending broadcast was recorded while its stream was being captured into a graph
  cudaStream_t compute = nullptr, comm = nullptr;
  bool own_comm = false;
  long long oneshot_max_bytes = 0, slot_bytes = 0;
  int nblocks = 32;
  int wire_bf16 = 0;
  long long wire_off = -1;  // byte offset (in the data region) of the bf16 wire buffer; -1: none
  bool use_nvls = false;
  float scale = 1.f;
  int launched = 0;
  int last_error = 0;
};

extern "C" void* dmpi_reducer_cuda_create(const dmpi_comm_ptrs* ptrs, int n_buckets, const long long* off_bytes,
                                          const long long* nbytes, float scale, long long oneshot_max_bytes,
                                          long long slot_bytes, int nblocks, int wire_bf16) {
  ReducerCuda* r = new (std::nothrow) ReducerCuda();
  if (!r) return nullptr;
  r->ptrs = *ptrs;
  r->world = ptrs->world;
  r->off.assign(off_bytes, off_bytes + n_buckets);
  r->bytes.assign(nbytes, nbytes + n_buckets);
  r->off_base = n_buckets > 0 ? r->off[0] : 0;
  for (long long o : r->off) r->off_base = o < r->off_base ? o : r->off_base;
  r->ready.resize(n_buckets);
  for (auto& e : r->ready) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&r->done, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&r->fwd_done, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&r->bcast_done, cudaEventDisableTiming);
  int lo = 0, hi = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStreamCreateWithPriority(&r->comm, cudaStreamNonBlocking, hi);  // highest priority: comm blocks get SMs first
  r->own_comm = true;
  r->scale = scale;
  r->oneshot_max_bytes = oneshot_max_bytes;
  r->slot_bytes = slot_bytes;
  r->nblocks = nblocks;
  if (const char* e = getenv("DMPI_COMM_BLOCKS")) r->nblocks = atoi(e) > 0 ? atoi(e) : nblocks;
  r->wire_bf16 = wire_bf16;
  // profiles/allreduce_sweep_w8.json: NVLS beats two-shot at every size above the one-shot range at W=8, loses at W=2
  r->use_nvls = ptrs->mc_data != nullptr && ptrs->world > 2;
  if (const char* e = getenv("DMPI_NVLS")) r->use_nvls = ptrs->mc_data != nullptr && atoi(e) != 0;
  // profiles/allreduce_sweep_w8.json (round 2, 4 reductions in flight per thread): the NVLS kernel is as fast with 16
  // blocks as with 32 or 64 at every bucket size (9.4 MB: 62 / 65 / 67 us, 25.3 MB: 103 / 109 / 115 us), so it takes
  // half as many SMs away from the persistent convolution kernels it overlaps with; two-shot still wants 32
  if (r->use_nvls && getenv("DMPI_COMM_BLOCKS") == nullptr && r->nblocks > 16) r->nblocks = 16;
  return r;
}

// bf16 wire buffer: `bytes` >= half the gradient region, 16-byte aligned offset inside the symmetric data region.
extern "C" int dmpi_reducer_cuda_set_wire_buffer(void* h, long long off_bytes, long long bytes) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  long long need = 0;
  for (size_t b = 0; b < r->off.size(); ++b) {
    const long long end = (r->off[b] - r->off_base + r->bytes[b]) / 2;
    need = end > need ? end : need;
  }
  if (off_bytes % 16 || bytes < need) return -1;
  r->wire_off = off_bytes;
  return 0;
}

extern "C" void dmpi_reducer_cuda_destroy(void* h) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  if (!r) return;
  for (auto& e : r->ready) cudaEventDestroy(e);
  cudaEventDestroy(r->done);
  cudaEventDestroy(r->fwd_done);
  cudaEventDestroy(r->bcast_done);
  if (r->own_comm) cudaStreamDestroy(r->comm);
  delete r;
}

// Called at the start of every synchronising iteration with the stream the backward will run on.
extern "C" void dmpi_reducer_cuda_begin(void* h, void* compute_stream) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  r->compute = static_cast<cudaStream_t>(compute_stream);
  r->launched = 0;
}

// The dmpi_launch_cb installed into the reducer core.
extern "C" void dmpi_reducer_cuda_launch(void* h, int bucket) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  cudaEventRecord(r->ready[bucket], r->compute);
  cudaStreamWaitEvent(r->comm, r->ready[bucket], 0);
  int rc = 0;
  if (r->world > 1 && r->wire_bf16 && r->wire_off >= 0 && r->bytes[bucket] > r->oneshot_max_bytes) {
    // bf16 wire format for the large buckets: cast the fp32 bucket into the (symmetric, multicast-mapped) wire buffer,
    // all-reduce that with fp32 accumulation (half the NVLink bytes), cast the mean back.  Three launches on the comm
    // stream instead of one; worth it where the link, not the launch, bounds the bucket.
    const long long n = r->bytes[bucket] / 4;
    const long long woff = r->wire_off + (r->off[bucket] - r->off_base) / 2;
    char* base = static_cast<char*>(r->ptrs.data[r->ptrs.rank_base]);
    rc = dmpi_cast_bf16(reinterpret_cast<const float*>(base + r->off[bucket]), base + woff, n, r->comm);
    if (rc == 0) {
      if (r->use_nvls) rc = dmpi_allreduce_nvls(&r->ptrs, woff, n * 2, DMPI_BF16, r->scale, r->nblocks, r->comm);
      else rc = dmpi_allreduce_twoshot(&r->ptrs, woff, n * 2, DMPI_BF16, r->scale, r->nblocks, 0, r->comm);
    }
    if (rc == 0) rc = dmpi_cast_f32(base + woff, reinterpret_cast<float*>(base + r->off[bucket]), n, r->comm);
  } else if (r->world > 1) {
    if (r->bytes[bucket] <= r->oneshot_max_bytes && r->bytes[bucket] <= r->slot_bytes)
      rc = dmpi_allreduce_oneshot(&r->ptrs, r->off[bucket], r->bytes[bucket], r->wire_bf16, r->scale, r->slot_bytes,
                                  r->nblocks, 0, r->comm);
    else if (r->use_nvls)
      rc = dmpi_allreduce_nvls(&r->ptrs, r->off[bucket], r->bytes[bucket], DMPI_F32, r->scale, r->nblocks, r->comm);
    else
      rc = dmpi_allreduce_twoshot(&r->ptrs, r->off[bucket], r->bytes[bucket], DMPI_F32, r->scale, r->nblocks, 0,
                                  r->comm);
  }
  if (rc != 0) r->last_error = rc;
  r->launched++;
  if (r->launched == static_cast<int>(r->ready.size())) cudaEventRecord(r->done, r->comm);
}

// End of backward: later work on the compute stream (clip, optimizer) waits for the last all-reduce.
extern "C" int dmpi_reducer_cuda_finalize(void* h) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  if (r->launched != static_cast<int>(r->ready.size())) return -1;
  cudaStreamWaitEvent(r->compute, r->done, 0);
  r->bcast_pending = false;  // `done` was recorded on the comm stream after the broadcast: already joined
  const int e = r->last_error;
  r->last_error = 0;
  return e;
}

// Per-iteration buffer broadcast (BatchNorm running statistics, rank 0 -> all) OFF the forward's critical
// path: torch DDP broadcasts at the head of forward i+1 what rank 0 holds after forward i; nothing reads or writes the
// buffers between the end of forward i and the BatchNorm kernels of forward i+1, so the same bytes can travel any time
// in between.  Called right after the forward pass was enqueued: the comm stream waits for the forward, runs the
// broadcast kernel (its two cross-GPU barriers overlap the loss and the first backward kernels) and the bucket
// all-reduces queue up behind it.  The compute stream picks the result up in dmpi_reducer_cuda_join_broadcast (head of
// the next forward) or, earlier, through the `done` event of finalize (same stream, later).
extern "C" int dmpi_reducer_cuda_side_broadcast(void* h, long long off_bytes, long long nbytes, int root, int nblocks,
                                                void* compute_stream) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  if (r->world <= 1 || nbytes <= 0) return 0;
  cudaStream_t cs = static_cast<cudaStream_t>(compute_stream);
  cudaEventRecord(r->fwd_done, cs);
  cudaStreamWaitEvent(r->comm, r->fwd_done, 0);
  const int rc = dmpi_broadcast(&r->ptrs, off_bytes, nbytes, root, nblocks, 0, r->comm);
  cudaEventRecord(r->bcast_done, r->comm);
  r->bcast_pending = true;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(cs, &cap);
  r->bcast_captured = cap == cudaStreamCaptureStatusActive;
  return rc;
}

extern "C" void dmpi_reducer_cuda_join_broadcast(void* h, void* stream) {
  ReducerCuda* r = static_cast<ReducerCuda*>(h);
  if (!r->bcast_pending) return;
  r->bcast_pending = false;
  // An event may only be waited for on the side of the capture boundary it was recorded on.  Crossing it needs no
  // wait: stream capture begins and ends with the device quiesced (torch.cuda.graph synchronises before capturing, and
  // a captured step joins the comm stream in its own finalize).
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(static_cast<cudaStream_t>(stream), &cap);
  if ((cap == cudaStreamCaptureStatusActive) != r->bcast_captured) return;
  cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), r->bcast_done, 0);
}

extern "C" void* dmpi_reducer_cuda_comm_stream(void* h) { return static_cast<ReducerCuda*>(h)->comm; }
extern "C" int dmpi_reducer_cuda_launched(void* h) { return static_cast<ReducerCuda*>(h)->launched; }
extern "C" void* dmpi_reducer_cuda_launch_fn(void) { return reinterpret_cast<void*>(&dmpi_reducer_cuda_launch); }
