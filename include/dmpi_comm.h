/* C ABI of the NVLink peer-to-peer communication layer (csrc/p2p, csrc/kernels/comm).
 *
 * Replaces, on a single NVSwitch node, the NCCL calls that the reference reaches through torch.distributed:
 * all-reduce of gradient buckets (ref:pytorch/resnet/main.py:131 via DDP), parameter/buffer broadcast at
 * wrap time and per forward (ref:pytorch/resnet/main.py:44-46,128), and send/recv
 * (ref:pytorch/hello_world/hello_world.py:26,29).   SURVEY.md N2, N4, N6, N10, N11.
 */
#ifndef DMPI_COMM_H_
#define DMPI_COMM_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DMPI_MAX_RANKS 8
#define DMPI_MAX_COMM_BLOCKS 64
/* slots of the epoch array past the per-block epochs (a rank only touches its own) */
#define DMPI_EPOCH_ONESHOT_CALLS DMPI_MAX_COMM_BLOCKS      /* one-shot all-reduce launches so far: staging parity */
#define DMPI_EPOCH_ONESHOT_DONE (DMPI_MAX_COMM_BLOCKS + 1) /* blocks of the running one-shot launch that have finished */
#define DMPI_F32 0
#define DMPI_BF16 1
#define DMPI_IPC_HANDLE_BYTES 64

/* Peer pointer table handed to every kernel (by value).  Index = rank. */
typedef struct {
  void* data[DMPI_MAX_RANKS];      /* symmetric data region (gradient buckets, params, buffers) */
  uint32_t* sig[DMPI_MAX_RANKS];   /* signal pad: [DMPI_MAX_COMM_BLOCKS][DMPI_MAX_RANKS] epochs */
  void* stage[DMPI_MAX_RANKS];     /* one-shot staging: [2 parities][DMPI_MAX_RANKS][slot_bytes] */
  uint32_t* epoch[DMPI_MAX_RANKS]; /* per-block epoch counters, then the DMPI_EPOCH_* launch-wide counters */
  int world;
  int rank_base;                   /* this process' rank (0 in single-GPU emulation) */
  void* mc_data;                   /* NVLS: multicast address of the data region (NULL when the host exposes none) */
} dmpi_comm_ptrs;

typedef struct dmpi_arena dmpi_arena;

/* Allocates this rank's arena on the current device:  data | staging | mailboxes | signal pads (all zeroed). */
dmpi_arena* dmpi_arena_create(int rank, int world, long long data_bytes, long long stage_slot_bytes,
                              long long mailbox_bytes);
/* Single-GPU emulation: `world` arenas on the current device sharing one pointer table. */
dmpi_arena* dmpi_arena_create_emulated(int world, long long data_bytes, long long stage_slot_bytes,
                                       long long mailbox_bytes);
void dmpi_arena_destroy(dmpi_arena* a);
/* 64-byte cudaIpcMemHandle of the local allocation, to be exchanged through the store. */
int dmpi_arena_ipc_handle(dmpi_arena* a, void* out64);
/* handles: world * 64 bytes ordered by rank (own entry ignored).  Maps every peer's arena. */
int dmpi_arena_open_peers(dmpi_arena* a, const void* handles);
/* ---- VMM / NVLS mode: the arena is a cuMemCreate allocation shared through POSIX file descriptors, which (unlike
 * cudaMalloc + cudaIpc memory) can also be bound to an NVSwitch multicast object.  Set-up sequence, all ranks:
 *   a = dmpi_arena_create_vmm(...); fd = dmpi_arena_export_fd(a);            exchange fds (dmpi_fd_* below)
 *   dmpi_arena_import_peer_fd(a, r, fd_r) for every peer r
 *   rank 0: mcfd = dmpi_arena_mc_create(a); others: dmpi_arena_mc_import(a, mcfd)
 *   dmpi_arena_mc_add_device(a); <barrier>; dmpi_arena_mc_bind_map(a); <barrier>
 * Every step returns 0 or a negative code (dmpi_arena_last_error()); multicast being unavailable is reported by
 * dmpi_arena_multicast_supported() == 0 and is not an error: the kernels then use the peer-pointer algorithms. */
dmpi_arena* dmpi_arena_create_vmm(int rank, int world, long long data_bytes, long long stage_slot_bytes,
                                  long long mailbox_bytes);
int dmpi_arena_multicast_supported(void);
int dmpi_arena_export_fd(dmpi_arena* a);
int dmpi_arena_import_peer_fd(dmpi_arena* a, int peer_rank, int fd);
int dmpi_arena_mc_create(dmpi_arena* a);
int dmpi_arena_mc_import(dmpi_arena* a, int fd);
int dmpi_arena_mc_add_device(dmpi_arena* a);
int dmpi_arena_mc_bind_map(dmpi_arena* a);
/* File-descriptor passing between the ranks of one node (SCM_RIGHTS over abstract unix sockets). */
int dmpi_fd_listen(const char* name);
int dmpi_fd_send(const char* peer_name, int fd, int tag);
int dmpi_fd_recv(int sock, int* tag, int timeout_ms);
void dmpi_fd_close(int sock);
/* In-place all-reduce through the switch: multimem.ld_reduce of slice r, fused scale, multimem.st to all ranks. */
int dmpi_allreduce_nvls(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int dtype, float scale,
                        int nblocks, void* stream);
const dmpi_comm_ptrs* dmpi_arena_ptrs(dmpi_arena* a);
void* dmpi_arena_data(dmpi_arena* a, int rank);
long long dmpi_arena_stage_slot_bytes(dmpi_arena* a);
/* mailbox / flag addresses for send(src -> dst): see comm.cu */
void* dmpi_arena_mailbox(dmpi_arena* a, int owner_rank, int source_rank);
uint32_t* dmpi_arena_seq_flag(dmpi_arena* a, int owner_rank, int source_rank);
uint32_t* dmpi_arena_ack_flag(dmpi_arena* a, int owner_rank, int dest_rank);
long long dmpi_arena_mailbox_bytes(dmpi_arena* a);
const char* dmpi_arena_last_error(void);

/* In-place all-reduce of [off_bytes, off_bytes + nbytes) of the data region, result = scale * sum over ranks.
 * dtype: DMPI_F32 or DMPI_BF16 (fp32 accumulation).  emulate_ranks > 0: single-launch emulation of that many ranks. */
int dmpi_allreduce_twoshot(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int dtype, float scale,
                           int nblocks, int emulate_ranks, void* stream);
/* Latency path for small fp32 messages (nbytes <= slot_bytes); wire_bf16 halves the NVLink traffic.  Staging is
 * double buffered by the parity of a launch-wide call counter, so consecutive calls may differ in size and nblocks;
 * calls on one arena must be stream-ordered on every rank (as for every collective here). */
int dmpi_allreduce_oneshot(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int wire_bf16,
                           float scale, long long slot_bytes, int nblocks, int emulate_ranks, void* stream);
/* Copies [off_bytes, +nbytes) of root's data region into the same range of every rank. */
int dmpi_broadcast(const dmpi_comm_ptrs* ptrs, long long off_bytes, long long nbytes, int root, int nblocks,
                   int emulate_ranks, void* stream);
/* Grid-stride 16-byte copy (either pointer may be a mapped peer address): NVLink read / write bandwidth probe. */
int dmpi_peer_copy(void* dst, const void* src, long long nbytes, int nblocks, void* stream);
/* Point-to-point: message number `seq` (1, 2, ...) of the (source, destination) pair. */
int dmpi_send(const void* src, void* peer_mailbox, long long nbytes, uint32_t* peer_seq, const uint32_t* my_ack,
              uint32_t seq, void* stream);
int dmpi_recv(void* dst, const void* my_mailbox, long long nbytes, const uint32_t* my_seq, uint32_t* peer_ack,
              uint32_t seq, void* stream);
int dmpi_sendrecv_emulated(const void* src, void* mailbox, void* dst, long long nbytes, uint32_t* seq_flag,
                           uint32_t* ack_flag, uint32_t seq, void* stream);

/* Upper bound, in seconds, of every device-side wait on a peer's flag (barriers, send/recv); a wait that exceeds it
 * traps, which surfaces on the host as a launch failure instead of a hung GPU.  Default 600 (the order of NCCL's
 * watchdog: rank 0 may legitimately spend minutes in a rank-0-only evaluation + checkpoint while its peers already
 * wait in the next step's first collective); 0 disables the bound.  Also read once from DMPI_COMM_TIMEOUT_S.
 * Synchronous (cudaMemcpyToSymbol): call it outside stream capture. */
int dmpi_comm_set_timeout(double seconds);

#ifdef __cplusplus
}
#endif
#endif
