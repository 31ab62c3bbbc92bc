// Internal definition of the symmetric arena shared by arena.cu (cudaMalloc + cudaIpc mode) and vmm.cu
// (cuMemCreate + POSIX-fd sharing + NVSwitch multicast mode).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dmpi_comm.h"

extern thread_local char g_arena_err[256];

static inline long long align256(long long v) { return (v + 255) & ~255LL; }

struct dmpi_arena {
  int rank, world;
  bool emulated;
  long long data_bytes, stage_slot_bytes, mailbox_bytes;
  long long off_stage, off_mail, off_sig, off_epoch, off_seq, off_ack, total;
  void* base[DMPI_MAX_RANKS];  // mapped base of every rank's arena (own = local allocation)
  bool opened[DMPI_MAX_RANKS];
  dmpi_comm_ptrs ptrs;
  // VMM / NVLS mode
  bool vmm;
  size_t map_size;
  CUmemGenericAllocationHandle h_mem[DMPI_MAX_RANKS];
  CUmemGenericAllocationHandle h_mc;
  bool has_mc;
  CUdeviceptr mc_va;
};


void dmpi_arena_layout(dmpi_arena* a);
void dmpi_arena_fill_ptrs(dmpi_arena* a);
void dmpi_arena_vmm_destroy(dmpi_arena* a);
