// Symmetric arena: one cudaMalloc block per rank with an identical layout on every rank, exported with
// cudaIpcGetMemHandle and mapped by every peer (cudaIpcOpenMemHandle enables peer access lazily).
// The 64-byte handles travel through the framework's own TCP store (csrc/store) during
// init_process_group — this is the role ncclCommInitRank + the TCPStore exchange of the NCCL unique id play
// for the reference.
//
// Layout (all regions 256-byte aligned, zero-initialised):
//   [ data_bytes ]                                   gradient buckets | params | buffers (symmetric offsets)
//   [ 2 * MAX_RANKS * stage_slot_bytes ]             one-shot all-reduce staging (double buffered)
//   [ MAX_RANKS * mailbox_bytes ]                    send/recv mailboxes, one per source rank
//   [ MAX_COMM_BLOCKS * MAX_RANKS * 4 ]              signal pad
//   [ (MAX_COMM_BLOCKS + 64) * 4 ]                   per-block epochs + launch-wide counters (local use)
//   [ MAX_RANKS * 4 ] [ MAX_RANKS * 4 ]              send/recv sequence and acknowledge flags
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <new>

#include "../../include/dmpi_comm.h"

thread_local char g_arena_err[256] = {0};
#define ARENA_CHECK(expr)                                                                              \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess) {                                                                           \
      snprintf(g_arena_err, sizeof(g_arena_err), "%s failed: %s", #expr, cudaGetErrorString(_e));      \
      return -1;                                                                                       \
    }                                                                                                  \
  } while (0)

#include "arena_internal.h"

void dmpi_arena_layout(dmpi_arena* a) {
  a->off_stage = align256(a->data_bytes);
  a->off_mail = a->off_stage + align256(2LL * DMPI_MAX_RANKS * a->stage_slot_bytes);
  a->off_sig = a->off_mail + align256(1LL * DMPI_MAX_RANKS * a->mailbox_bytes);
  a->off_epoch = a->off_sig + align256(4LL * DMPI_MAX_COMM_BLOCKS * DMPI_MAX_RANKS);
  a->off_seq = a->off_epoch + align256(4LL * (DMPI_MAX_COMM_BLOCKS + 64));  // + launch-wide counters (DMPI_EPOCH_*)
  a->off_ack = a->off_seq + align256(4LL * DMPI_MAX_RANKS);
  a->total = a->off_ack + align256(4LL * DMPI_MAX_RANKS);
}

void dmpi_arena_fill_ptrs(dmpi_arena* a) {
  memset(&a->ptrs, 0, sizeof(a->ptrs));
  a->ptrs.world = a->world;
  a->ptrs.rank_base = a->emulated ? 0 : a->rank;
  for (int r = 0; r < a->world; ++r) {
    char* b = static_cast<char*>(a->base[r]);
    if (!b) continue;
    a->ptrs.data[r] = b;
    a->ptrs.stage[r] = b + a->off_stage;
    a->ptrs.sig[r] = reinterpret_cast<uint32_t*>(b + a->off_sig);
    a->ptrs.epoch[r] = reinterpret_cast<uint32_t*>(b + a->off_epoch);
  }
  a->ptrs.mc_data = (a->vmm && a->mc_va) ? reinterpret_cast<void*>(a->mc_va) : nullptr;
}

static dmpi_arena* make(int rank, int world, long long data_bytes, long long slot, long long mailbox, bool emulated) {
  if (world < 1 || world > DMPI_MAX_RANKS) {
    snprintf(g_arena_err, sizeof(g_arena_err), "world size %d outside 1..%d", world, DMPI_MAX_RANKS);
    return nullptr;
  }
  dmpi_arena* a = new (std::nothrow) dmpi_arena();
  if (!a) return nullptr;
  memset(a, 0, sizeof(*a));
  a->rank = rank;
  a->world = world;
  a->emulated = emulated;
  a->data_bytes = align256(data_bytes);
  a->stage_slot_bytes = align256(slot);
  a->mailbox_bytes = align256(mailbox);
  dmpi_arena_layout(a);
  const int n_local = emulated ? world : 1;
  for (int i = 0; i < n_local; ++i) {
    const int r = emulated ? i : rank;
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, a->total);
    if (e == cudaSuccess) e = cudaMemset(p, 0, a->total);
    if (e != cudaSuccess) {
      snprintf(g_arena_err, sizeof(g_arena_err), "arena allocation of %lld bytes failed: %s", a->total,
               cudaGetErrorString(e));
      for (int j = 0; j < DMPI_MAX_RANKS; ++j)
        if (a->base[j]) cudaFree(a->base[j]);
      delete a;
      return nullptr;
    }
    a->base[r] = p;
  }
  dmpi_arena_fill_ptrs(a);
  return a;
}

extern "C" dmpi_arena* dmpi_arena_create(int rank, int world, long long data_bytes, long long stage_slot_bytes,
                                         long long mailbox_bytes) {
  return make(rank, world, data_bytes, stage_slot_bytes, mailbox_bytes, false);
}

extern "C" dmpi_arena* dmpi_arena_create_emulated(int world, long long data_bytes, long long stage_slot_bytes,
                                                  long long mailbox_bytes) {
  return make(0, world, data_bytes, stage_slot_bytes, mailbox_bytes, true);
}

void dmpi_arena_vmm_destroy(dmpi_arena* a);

extern "C" void dmpi_arena_destroy(dmpi_arena* a) {
  if (!a) return;
  if (a->vmm) {
    dmpi_arena_vmm_destroy(a);
    delete a;
    return;
  }
  for (int r = 0; r < DMPI_MAX_RANKS; ++r) {
    if (!a->base[r]) continue;
    if (a->opened[r]) cudaIpcCloseMemHandle(a->base[r]);
    else cudaFree(a->base[r]);
  }
  delete a;
}

extern "C" int dmpi_arena_ipc_handle(dmpi_arena* a, void* out64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == DMPI_IPC_HANDLE_BYTES, "unexpected IPC handle size");
  cudaIpcMemHandle_t h;
  ARENA_CHECK(cudaIpcGetMemHandle(&h, a->base[a->rank]));
  memcpy(out64, &h, sizeof(h));
  return 0;
}

extern "C" int dmpi_arena_open_peers(dmpi_arena* a, const void* handles) {
  const char* hb = static_cast<const char*>(handles);
  for (int r = 0; r < a->world; ++r) {
    if (r == a->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, hb + static_cast<size_t>(r) * DMPI_IPC_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    ARENA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    a->base[r] = p;
    a->opened[r] = true;
  }
  dmpi_arena_fill_ptrs(a);
  return 0;
}

extern "C" const dmpi_comm_ptrs* dmpi_arena_ptrs(dmpi_arena* a) { return &a->ptrs; }
extern "C" void* dmpi_arena_data(dmpi_arena* a, int rank) { return a->base[rank]; }
extern "C" long long dmpi_arena_stage_slot_bytes(dmpi_arena* a) { return a->stage_slot_bytes; }
extern "C" long long dmpi_arena_mailbox_bytes(dmpi_arena* a) { return a->mailbox_bytes; }
extern "C" void* dmpi_arena_mailbox(dmpi_arena* a, int owner_rank, int source_rank) {
  return static_cast<char*>(a->base[owner_rank]) + a->off_mail + static_cast<long long>(source_rank) * a->mailbox_bytes;
}
extern "C" uint32_t* dmpi_arena_seq_flag(dmpi_arena* a, int owner_rank, int source_rank) {
  return reinterpret_cast<uint32_t*>(static_cast<char*>(a->base[owner_rank]) + a->off_seq) + source_rank;
}
extern "C" uint32_t* dmpi_arena_ack_flag(dmpi_arena* a, int owner_rank, int dest_rank) {
  return reinterpret_cast<uint32_t*>(static_cast<char*>(a->base[owner_rank]) + a->off_ack) + dest_rank;
}
extern "C" const char* dmpi_arena_last_error(void) { return g_arena_err; }
