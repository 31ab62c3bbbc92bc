// VMM / NVLS mode of the symmetric arena.
//
// cudaMalloc + cudaIpc memory (arena.cu) cannot be bound to an NVSwitch multicast object; memory created with
// cuMemCreate and shared as a POSIX file descriptor can.  In this mode every rank
//   1. creates its arena with cuMemCreate(requestedHandleTypes = POSIX_FILE_DESCRIPTOR) and maps it,
//   2. exports it as an fd, passes the fd to every peer over an abstract unix socket (SCM_RIGHTS) and maps the peers'
//      allocations (cuMemImportFromShareableHandle + cuMemMap + cuMemSetAccess): the peer pointer table,
//   3. joins the multicast object created by rank 0 (cuMulticastCreate / AddDevice / BindMem) and maps it: one virtual
//      address whose loads reduce across all GPUs in the switch (multimem.ld_reduce) and whose stores are broadcast
//      (multimem.st) — the "NVLS" path of NCCL, here driven by our own kernel (comm.cu: allreduce_nvls_kernel).
// All driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the library still loads on a build box
// without libcuda.  Multicast is optional: when CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 the arena works as a plain
// peer-mapped arena.  
#include <errno.h>
#include <poll.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <new>

#include "arena_internal.h"

namespace {

struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok;
};

template <typename F>
bool resolve(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    snprintf(g_arena_err, sizeof(g_arena_err), "driver entry point %s unavailable", name);
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

DriverApi* api() {
  static DriverApi d = {};
  static bool tried = false;
  if (!tried) {
    tried = true;
    cudaFree(0);  // make sure the primary context exists
    d.ok = resolve(d.MemCreate, "cuMemCreate") && resolve(d.MemRelease, "cuMemRelease") &&
           resolve(d.MemAddressReserve, "cuMemAddressReserve") && resolve(d.MemAddressFree, "cuMemAddressFree") &&
           resolve(d.MemMap, "cuMemMap") && resolve(d.MemUnmap, "cuMemUnmap") &&
           resolve(d.MemSetAccess, "cuMemSetAccess") &&
           resolve(d.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
           resolve(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
           resolve(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
           resolve(d.MulticastCreate, "cuMulticastCreate") && resolve(d.MulticastAddDevice, "cuMulticastAddDevice") &&
           resolve(d.MulticastBindMem, "cuMulticastBindMem") &&
           resolve(d.MulticastGetGranularity, "cuMulticastGetGranularity") && resolve(d.DeviceGet, "cuDeviceGet") &&
           resolve(d.DeviceGetAttribute, "cuDeviceGetAttribute");
  }
  return d.ok ? &d : nullptr;
}

#define CU_TRY(call)                                                                                    \
  do {                                                                                                  \
    CUresult _r = (call);                                                                               \
    if (_r != CUDA_SUCCESS) {                                                                           \
      snprintf(g_arena_err, sizeof(g_arena_err), "%s failed with CUresult %d", #call, static_cast<int>(_r)); \
      return -1;                                                                                        \
    }                                                                                                   \
  } while (0)

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

CUmemAllocationProp mem_prop(int dev) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_handle(DriverApi* d, CUmemGenericAllocationHandle h, size_t size, int dev, CUdeviceptr* out) {
  CUdeviceptr va = 0;
  CU_TRY(d->MemAddressReserve(&va, size, 0, 0, 0));
  CU_TRY(d->MemMap(va, size, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_TRY(d->MemSetAccess(va, size, &acc, 1));
  *out = va;
  return 0;
}

}  // namespace

extern "C" int dmpi_arena_multicast_supported(void) {
  DriverApi* d = api();
  if (!d) return 0;
  CUdevice cd;
  if (d->DeviceGet(&cd, current_device()) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v;
}

extern "C" dmpi_arena* dmpi_arena_create_vmm(int rank, int world, long long data_bytes, long long stage_slot_bytes,
                                             long long mailbox_bytes) {
  DriverApi* d = api();
  if (!d) return nullptr;
  if (world < 1 || world > DMPI_MAX_RANKS) {
    snprintf(g_arena_err, sizeof(g_arena_err), "world size %d outside 1..%d", world, DMPI_MAX_RANKS);
    return nullptr;
  }
  dmpi_arena* a = new (std::nothrow) dmpi_arena();
  if (!a) return nullptr;
  memset(a, 0, sizeof(*a));
  a->rank = rank;
  a->world = world;
  a->vmm = true;
  a->data_bytes = align256(data_bytes);
  a->stage_slot_bytes = align256(stage_slot_bytes);
  a->mailbox_bytes = align256(mailbox_bytes);
  dmpi_arena_layout(a);
  const int dev = current_device();
  CUmemAllocationProp prop = mem_prop(dev);
  size_t gran = 0;
  if (d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0)
    gran = 2u << 20;
  if (dmpi_arena_multicast_supported()) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = static_cast<size_t>(a->total);
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
  }
  a->map_size = (static_cast<size_t>(a->total) + gran - 1) / gran * gran;
  CUresult r = d->MemCreate(&a->h_mem[rank], a->map_size, &prop, 0);
  if (r != CUDA_SUCCESS) {
    snprintf(g_arena_err, sizeof(g_arena_err), "cuMemCreate(%zu bytes) failed with CUresult %d", a->map_size, (int)r);
    delete a;
    return nullptr;
  }
  CUdeviceptr va = 0;
  if (map_handle(d, a->h_mem[rank], a->map_size, dev, &va) != 0) {
    d->MemRelease(a->h_mem[rank]);
    delete a;
    return nullptr;
  }
  a->base[rank] = reinterpret_cast<void*>(va);
  cudaMemset(a->base[rank], 0, a->map_size);
  cudaDeviceSynchronize();
  dmpi_arena_fill_ptrs(a);
  return a;
}

extern "C" int dmpi_arena_export_fd(dmpi_arena* a) {
  DriverApi* d = api();
  if (!d || !a->vmm) return -1;
  int fd = -1;
  CU_TRY(d->MemExportToShareableHandle(&fd, a->h_mem[a->rank], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return fd;
}

extern "C" int dmpi_arena_import_peer_fd(dmpi_arena* a, int peer_rank, int fd) {
  DriverApi* d = api();
  if (!d || !a->vmm || peer_rank < 0 || peer_rank >= a->world || peer_rank == a->rank) return -1;
  CU_TRY(d->MemImportFromShareableHandle(&a->h_mem[peer_rank], reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  CUdeviceptr va = 0;
  if (map_handle(d, a->h_mem[peer_rank], a->map_size, current_device(), &va) != 0) return -1;
  a->base[peer_rank] = reinterpret_cast<void*>(va);
  a->opened[peer_rank] = true;
  ::close(fd);
  dmpi_arena_fill_ptrs(a);
  return 0;
}

// rank 0: create the multicast object and return an fd for the other ranks
extern "C" int dmpi_arena_mc_create(dmpi_arena* a) {
  DriverApi* d = api();
  if (!d || !a->vmm) return -1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = static_cast<unsigned>(a->world);
  mp.size = a->map_size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CU_TRY(d->MulticastCreate(&a->h_mc, &mp));
  a->has_mc = true;
  int fd = -1;
  CU_TRY(d->MemExportToShareableHandle(&fd, a->h_mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return fd;
}

extern "C" int dmpi_arena_mc_import(dmpi_arena* a, int fd) {
  DriverApi* d = api();
  if (!d || !a->vmm) return -1;
  CU_TRY(d->MemImportFromShareableHandle(&a->h_mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  a->has_mc = true;
  ::close(fd);
  return 0;
}

extern "C" int dmpi_arena_mc_add_device(dmpi_arena* a) {
  DriverApi* d = api();
  if (!d || !a->has_mc) return -1;
  CUdevice cd;
  CU_TRY(d->DeviceGet(&cd, current_device()));
  CU_TRY(d->MulticastAddDevice(a->h_mc, cd));
  return 0;
}

// after every rank has added its device: bind the local allocation and map the multicast address range
extern "C" int dmpi_arena_mc_bind_map(dmpi_arena* a) {
  DriverApi* d = api();
  if (!d || !a->has_mc) return -1;
  CU_TRY(d->MulticastBindMem(a->h_mc, 0, a->h_mem[a->rank], 0, a->map_size, 0));
  CUdeviceptr va = 0;
  if (map_handle(d, a->h_mc, a->map_size, current_device(), &va) != 0) return -1;
  a->mc_va = va;
  dmpi_arena_fill_ptrs(a);
  return 0;
}

void dmpi_arena_vmm_destroy(dmpi_arena* a) {
  DriverApi* d = api();
  if (!d) return;
  cudaDeviceSynchronize();
  if (a->mc_va) {
    d->MemUnmap(a->mc_va, a->map_size);
    d->MemAddressFree(a->mc_va, a->map_size);
  }
  for (int r = 0; r < DMPI_MAX_RANKS; ++r) {
    if (!a->base[r]) continue;
    const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(a->base[r]);
    d->MemUnmap(va, a->map_size);
    d->MemAddressFree(va, a->map_size);
    d->MemRelease(a->h_mem[r]);
  }
  if (a->has_mc) d->MemRelease(a->h_mc);
}

// ------------------------------------------------------------------------------------------ fd passing
static int make_addr(const char* name, sockaddr_un* addr, socklen_t* len) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  const size_t n = strlen(name);
  if (n + 2 > sizeof(addr->sun_path)) return -1;
  addr->sun_path[0] = '\0';  // abstract namespace: no file system entry, vanishes with the process
  memcpy(addr->sun_path + 1, name, n);
  *len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
  return 0;
}

extern "C" int dmpi_fd_listen(const char* name) {
  const int s = ::socket(AF_UNIX, SOCK_DGRAM, 0);
  if (s < 0) return -1;
  sockaddr_un addr;
  socklen_t len;
  if (make_addr(name, &addr, &len) != 0 || ::bind(s, reinterpret_cast<sockaddr*>(&addr), len) != 0) {
    ::close(s);
    return -1;
  }
  return s;
}

extern "C" int dmpi_fd_send(const char* peer_name, int fd, int tag) {
  const int s = ::socket(AF_UNIX, SOCK_DGRAM, 0);
  if (s < 0) return -1;
  sockaddr_un addr;
  socklen_t len;
  if (make_addr(peer_name, &addr, &len) != 0) {
    ::close(s);
    return -1;
  }
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec iov;
  iov.iov_base = &tag;
  iov.iov_len = sizeof(tag);
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_name = &addr;
  msg.msg_namelen = len;
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  int rc = -1;
  for (int attempt = 0; attempt < 200; ++attempt) {  // the peer may not have bound its socket yet
    if (::sendmsg(s, &msg, 0) >= 0) {
      rc = 0;
      break;
    }
    if (errno != ECONNREFUSED && errno != ENOENT) break;
    usleep(20000);
  }
  ::close(s);
  return rc;
}

extern "C" int dmpi_fd_recv(int sock, int* tag, int timeout_ms) {
  pollfd p;
  p.fd = sock;
  p.events = POLLIN;
  if (::poll(&p, 1, timeout_ms) <= 0) return -1;
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  int t = 0;
  iovec iov;
  iov.iov_base = &t;
  iov.iov_len = sizeof(t);
  char ctrl[CMSG_SPACE(sizeof(int))];
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (::recvmsg(sock, &msg, 0) < 0) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) return -1;
  int fd = -1;
  memcpy(&fd, CMSG_DATA(c), sizeof(int));
  if (tag) *tag = t;
  return fd;
}

extern "C" void dmpi_fd_close(int sock) { ::close(sock); }
