"""Process-group front-end with the ``torch.distributed`` names the reference scripts call
(SURVEY.md 2.7): ``init_process_group / destroy_process_group / send / recv / get_rank / get_world_size /
barrier / all_reduce / broadcast``.

Backends
--------
``"nccl"`` (the reference's default, ref:pytorch/resnet/main.py:158, ref:pytorch/unet/train.py:284) and ``"p2p"``
    On one NVSwitch node (world_size <= 8, every rank on this host) the group is served by this framework's
    own NVLink peer-to-peer kernels over a symmetric arena (csrc/p2p, csrc/kernels/comm) — NCCL is not
    touched.  Across nodes (or with ``DMPI_FORCE_NCCL=1``) it degrades to torch.distributed's NCCL group.
``"gloo"`` (ref:pytorch/hello_world/hello_world.py:44)
    CPU tensors over torch.distributed's gloo group, so the plumbing runs without a GPU.

Rendezvous follows the torchrun env contract (RANK, LOCAL_RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT).  Rank 0
hosts the framework's own C++ TCP store; under torchrun, whose agent already owns MASTER_PORT, the store's
port is published through the agent's store.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from .. import _native as N
from .store import TCPStore

_MAILBOX_BYTES = 1 << 20
_SCRATCH_BYTES = 64 << 20
_STAGE_SLOT_BYTES = 1 << 20


class P2PComm:
    """One symmetric arena mapped by every rank + the kernels that operate on it."""

    def __init__(self, state: "GroupState", data_bytes: int, tag: str):
        lib = N.cuda()
        self.lib, self.state = lib, state
        self.rank, self.world = state.rank, state.world
        self.has_multicast = False
        self.handle = None
        if self.world > 1 and os.environ.get("DMPI_VMM", "1") == "1":
            self._create_vmm(int(data_bytes), tag)
        if not self.handle:
            self._create_ipc(int(data_bytes), tag)
        self.ptrs = lib.dmpi_arena_ptrs(self.handle)
        self.data_ptr = lib.dmpi_arena_data(self.handle, self.rank)
        self.data_bytes = int(data_bytes)
        self.slot_bytes = lib.dmpi_arena_stage_slot_bytes(self.handle)
        self.mailbox_bytes = lib.dmpi_arena_mailbox_bytes(self.handle)
        self._send_seq: Dict[int, int] = {}
        self._recv_seq: Dict[int, int] = {}

    def _err(self) -> str:
        return (self.lib.dmpi_arena_last_error() or b"").decode()

    def _create_ipc(self, data_bytes: int, tag: str) -> None:
        """cudaMalloc arena shared through cudaIpc handles (exchanged via the TCP store)."""
        lib, state = self.lib, self.state
        self.handle = lib.dmpi_arena_create(self.rank, self.world, data_bytes, _STAGE_SLOT_BYTES, _MAILBOX_BYTES)
        if not self.handle:
            raise RuntimeError("arena allocation failed: " + self._err())
        if self.world > 1:
            h = C.create_string_buffer(64)
            if lib.dmpi_arena_ipc_handle(self.handle, h) != 0:
                raise RuntimeError("cudaIpcGetMemHandle failed: " + self._err())
            all_h = state.store.all_gather(f"ipc/{tag}", self.rank, self.world, h.raw)
            blob = C.create_string_buffer(b"".join(all_h), 64 * self.world)
            if lib.dmpi_arena_open_peers(self.handle, blob) != 0:
                raise RuntimeError("cudaIpcOpenMemHandle failed: " + self._err())
            state.store.barrier(f"ipc_open/{tag}", self.rank, self.world)

    def _create_vmm(self, data_bytes: int, tag: str) -> None:
        """cuMemCreate arena shared as POSIX fds (SCM_RIGHTS) and, where the host exposes it, bound to an NVSwitch
        multicast object (NVLS).  Every rank takes the same decisions (agreed through the store); any failure falls
        back, on all ranks together, to the cudaIpc arena."""
        lib, st = self.lib, self.state
        store, rank, world = st.store, self.rank, self.world
        uid = store.get("__run_uid__").decode() if store.check("__run_uid__") else "0"
        name = lambda r: f"dmpi-{uid}-{tag}-{r}".encode()  # noqa: E731
        handle = lib.dmpi_arena_create_vmm(rank, world, data_bytes, _STAGE_SLOT_BYTES, _MAILBOX_BYTES)
        sock = lib.dmpi_fd_listen(name(rank)) if handle else -1
        ok = bool(handle) and sock >= 0
        votes = store.all_gather(f"vmm_ok/{tag}", rank, world, b"1" if ok else b"0")
        if not all(v == b"1" for v in votes):
            if sock >= 0:
                lib.dmpi_fd_close(sock)
            if handle:
                lib.dmpi_arena_destroy(handle)
            return
        try:
            fd = lib.dmpi_arena_export_fd(handle)
            ok = fd >= 0
            if ok:
                for r in range(world):
                    if r != rank and lib.dmpi_fd_send(name(r), fd, rank) != 0:
                        ok = False
                os.close(fd)
                for _ in range(world - 1):
                    t = C.c_int(-1)
                    pfd = lib.dmpi_fd_recv(sock, C.byref(t), 30000)
                    if pfd < 0 or lib.dmpi_arena_import_peer_fd(handle, t.value, pfd) != 0:
                        ok = False
                        break
            votes = store.all_gather(f"vmm_mapped/{tag}", rank, world, b"1" if ok else b"0")
            if not all(v == b"1" for v in votes):
                raise RuntimeError("peer mapping failed: " + self._err())
            # ---- optional multicast object
            mc = lib.dmpi_arena_multicast_supported() == 1 and os.environ.get("DMPI_NVLS", "1") == "1"
            votes = store.all_gather(f"mc_cap/{tag}", rank, world, b"1" if mc else b"0")
            if all(v == b"1" for v in votes):
                ok = True
                if rank == 0:
                    mfd = lib.dmpi_arena_mc_create(handle)
                    ok = mfd >= 0
                    if ok:
                        for r in range(1, world):
                            ok = ok and lib.dmpi_fd_send(name(r), mfd, 1000) == 0
                        os.close(mfd)
                votes0 = store.all_gather(f"mc_created/{tag}", rank, world, b"1" if ok else b"0")
                if votes0[0] == b"1":
                    if rank != 0:
                        t = C.c_int(-1)
                        mfd = lib.dmpi_fd_recv(sock, C.byref(t), 30000)
                        ok = mfd >= 0 and lib.dmpi_arena_mc_import(handle, mfd) == 0
                    ok = ok and lib.dmpi_arena_mc_add_device(handle) == 0
                    votes = store.all_gather(f"mc_added/{tag}", rank, world, b"1" if ok else b"0")
                    if all(v == b"1" for v in votes):
                        ok = lib.dmpi_arena_mc_bind_map(handle) == 0
                        votes = store.all_gather(f"mc_bound/{tag}", rank, world, b"1" if ok else b"0")
                        self.has_multicast = all(v == b"1" for v in votes)
                        if not self.has_multicast and ok:
                            raise RuntimeError("multicast bind failed on a peer")  # mixed state: start over with IPC
            self.handle = handle
        except RuntimeError:
            lib.dmpi_arena_destroy(handle)
            self.handle = None
        finally:
            lib.dmpi_fd_close(sock)
        store.barrier(f"vmm_done/{tag}", rank, world)

    def tensor(self, offset: int, nbytes: int) -> torch.Tensor:
        from ..utils.arena import tensor_from_ptr

        return tensor_from_ptr(self.data_ptr + offset, nbytes, self.state.device)

    # ---- collectives on byte ranges of the data region (all ranks must call in the same order)
    def allreduce(self, off: int, nbytes: int, scale: float, dtype: int = 0, algo: str = "auto",
                  nblocks: int = 32, stream: Optional[int] = None) -> None:
        s = N.stream_ptr() if stream is None else stream
        if self.world == 1:
            return
        if algo == "auto":
            if dtype == 0 and nbytes <= self.state.oneshot_max_bytes:
                algo = "oneshot"
            else:
                # measured (profiles/allreduce_sweep_w8.json): in-switch reduction wins from the one-shot crossover up at
                # W=8 (1 GB: 835 vs 724 GB/s NCCL); at W=2 there is nothing for the switch to save and two-shot is faster
                nvls = self.has_multicast and self.world > 2 and nbytes >= self.state.nvls_min_bytes
                algo = "nvls" if nvls else "twoshot"
        if algo == "nvls":
            rc = self.lib.dmpi_allreduce_nvls(self.ptrs, off, nbytes, dtype, scale, nblocks, s)
        elif algo == "oneshot":
            rc = self.lib.dmpi_allreduce_oneshot(self.ptrs, off, nbytes, 0, scale, self.slot_bytes, nblocks, 0, s)
        elif algo == "oneshot_bf16":
            rc = self.lib.dmpi_allreduce_oneshot(self.ptrs, off, nbytes, 1, scale, self.slot_bytes, nblocks, 0, s)
        else:
            rc = self.lib.dmpi_allreduce_twoshot(self.ptrs, off, nbytes, dtype, scale, nblocks, 0, s)
        N.count_launch()
        if rc != 0:
            raise RuntimeError(f"all-reduce launch failed ({rc})")

    def broadcast(self, off: int, nbytes: int, root: int, nblocks: int = 16, stream: Optional[int] = None) -> None:
        if self.world == 1 or nbytes == 0:
            return
        s = N.stream_ptr() if stream is None else stream
        rc = self.lib.dmpi_broadcast(self.ptrs, off, nbytes, root, nblocks, 0, s)
        N.count_launch()
        if rc != 0:
            raise RuntimeError(f"broadcast launch failed ({rc})")

    def send(self, src: torch.Tensor, dst_rank: int) -> None:
        lib = self.lib
        seq = self._send_seq.get(dst_rank, 0) + 1
        self._send_seq[dst_rank] = seq
        nbytes = src.numel() * src.element_size()
        rc = lib.dmpi_send(src.data_ptr(), lib.dmpi_arena_mailbox(self.handle, dst_rank, self.rank), nbytes,
                           lib.dmpi_arena_seq_flag(self.handle, dst_rank, self.rank),
                           lib.dmpi_arena_ack_flag(self.handle, self.rank, dst_rank), seq, N.stream_ptr())
        N.count_launch()
        if rc != 0:
            raise RuntimeError(f"send launch failed ({rc})")

    def recv(self, dst: torch.Tensor, src_rank: int) -> None:
        lib = self.lib
        seq = self._recv_seq.get(src_rank, 0) + 1
        self._recv_seq[src_rank] = seq
        nbytes = dst.numel() * dst.element_size()
        rc = lib.dmpi_recv(dst.data_ptr(), lib.dmpi_arena_mailbox(self.handle, self.rank, src_rank), nbytes,
                           lib.dmpi_arena_seq_flag(self.handle, self.rank, src_rank),
                           lib.dmpi_arena_ack_flag(self.handle, src_rank, self.rank), seq, N.stream_ptr())
        N.count_launch()
        if rc != 0:
            raise RuntimeError(f"recv launch failed ({rc})")

    def destroy(self) -> None:
        if self.handle:
            self.lib.dmpi_arena_destroy(self.handle)
            self.handle = None


@dataclass
class GroupState:
    kind: str                      # 'p2p' | 'nccl' | 'gloo'
    backend: str                   # name the user passed
    rank: int
    world: int
    local_rank: int
    device: torch.device
    store: Optional[TCPStore] = None
    comms: list = field(default_factory=list)
    scratch: Optional[P2PComm] = None
    n_comms: int = 0
    barrier_count: int = 0
    oneshot_max_bytes: int = 256 << 10   # crossover one-shot -> two-shot; bench/allreduce_sweep.py measures it
    nvls_min_bytes: int = 256 << 10      # crossover two-shot -> NVLS multicast for world > 2 (from the W=8 sweep)

    def new_comm(self, data_bytes: int) -> P2PComm:
        tag = f"comm{self.n_comms}"
        self.n_comms += 1
        c = P2PComm(self, data_bytes, tag)
        self.comms.append(c)
        return c

    def scratch_comm(self) -> P2PComm:
        """Arena behind dist.send / recv / all_reduce / broadcast.  Creating an arena is a collective over the whole
        world (handle exchange through the store), so it happens inside init_process_group while every rank is known to
        be present — not lazily at the first call, which only a subset of ranks may ever make (send/recv between ranks
        0 and 1 of a 4-rank job).  DMPI_LAZY_SCRATCH=1 restores the lazy behaviour (saves 64 MB + the mailboxes per
        rank when the job only ever uses DistributedDataParallel)."""
        if self.scratch is None:
            self.scratch = self.new_comm(_SCRATCH_BYTES)
        return self.scratch


_state: Optional[GroupState] = None
_init_count = 0  # process groups created by this process so far (every rank creates them in the same order)


# ------------------------------------------------------------------------------------------ public API
def is_initialized() -> bool:
    return _state is not None


def _require() -> GroupState:
    if _state is None:
        raise RuntimeError("Default process group has not been initialized, please make sure to call init_process_group.")
    return _state


def get_rank() -> int:
    return _require().rank


def get_world_size() -> int:
    return _require().world


def get_backend() -> str:
    return _require().kind


def _env_int(name: str, default: Optional[int] = None) -> int:
    v = os.environ.get(name)
    if v is None:
        if default is None:
            raise KeyError(f"environment variable {name} is not set (launch with torchrun or export it)")
        return default
    return int(v)


def _bootstrap_store(rank: int, world: int, timeout_s: float) -> TCPStore:
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = _env_int("MASTER_PORT", 29500)
    if os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "False") == "True":
        # torchrun's agent already listens on MASTER_PORT: host our store on an ephemeral port and publish it there
        import datetime

        import torch.distributed as tdist

        agent = tdist.TCPStore(addr, port, world, False, timeout=datetime.timedelta(seconds=timeout_s))
        run_id = os.environ.get("TORCHELASTIC_RUN_ID", "none") + "/" + os.environ.get("TORCHELASTIC_RESTART_COUNT", "0")
        key = f"dmpi/store_port/{run_id}/{_init_count}"  # a re-created group must not read the previous store's port
        if rank == 0:
            store = TCPStore(addr, 0, True, timeout_s)
            agent.set(key, str(store.port))
            return store
        my_port = int(agent.get(key).decode())
        return TCPStore(addr, my_port, False, timeout_s)
    return TCPStore(addr, port, rank == 0, timeout_s)


def init_process_group(backend: str = "nccl", init_method: Optional[str] = None, rank: Optional[int] = None,
                       world_size: Optional[int] = None, timeout_s: float = 300.0, timeout=None, **_ignored) -> None:
    """Same call shapes as the reference: ``init_process_group(backend)`` and
    ``init_process_group(backend, rank=WORLD_RANK, world_size=WORLD_SIZE)``.  ``timeout`` (a ``datetime.timedelta``, as
    in torch.distributed) or ``timeout_s`` bounds the rendezvous and, on the gloo path, every collective."""
    global _state, _init_count
    if _state is not None:
        raise RuntimeError("trying to initialize the default process group twice!")
    _init_count += 1
    if timeout is not None:
        timeout_s = timeout.total_seconds() if hasattr(timeout, "total_seconds") else float(timeout)
    backend = (backend or "nccl").lower()
    rank = _env_int("RANK") if rank is None or rank < 0 else int(rank)
    world = _env_int("WORLD_SIZE") if world_size is None or world_size < 0 else int(world_size)
    local_rank = _env_int("LOCAL_RANK", rank)
    local_world = _env_int("LOCAL_WORLD_SIZE", world)

    if backend == "gloo" or not torch.cuda.is_available():
        if backend != "gloo":
            raise RuntimeError(f"backend '{backend}' needs CUDA; use --backend gloo on CPU-only machines")
        import torch.distributed as tdist

        if not tdist.is_initialized():
            import datetime

            tdist.init_process_group("gloo", rank=rank, world_size=world,
                                     timeout=datetime.timedelta(seconds=timeout_s))
        _state = GroupState("gloo", backend, rank, world, local_rank, torch.device("cpu"))
        return

    if backend not in ("nccl", "p2p"):
        raise ValueError(f"unknown backend '{backend}' (expected nccl, p2p or gloo)")
    ndev = torch.cuda.device_count()
    dev_index = local_rank % max(ndev, 1)
    torch.cuda.set_device(dev_index)  # the reference never calls set_device; we do it here
    device = torch.device("cuda", dev_index)
    single_node = local_world == world
    use_p2p = single_node and world <= 8 and os.environ.get("DMPI_FORCE_NCCL", "0") != "1"
    if backend == "p2p" and not use_p2p:
        raise RuntimeError("backend 'p2p' needs all ranks on one NVSwitch node (world_size <= 8)")
    if use_p2p and world > 1:
        for peer in range(ndev):
            if peer != dev_index and peer < world and not torch.cuda.can_device_access_peer(dev_index, peer):
                use_p2p = False
    if not use_p2p:
        import torch.distributed as tdist

        if not tdist.is_initialized():
            tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
        _state = GroupState("nccl", backend, rank, world, local_rank, device)
        return
    N.cuda()  # fail loudly now if the native library is missing
    store = _bootstrap_store(rank, world, timeout_s) if world > 1 else None
    _state = GroupState("p2p", backend, rank, world, local_rank, device, store)
    thr = os.environ.get("DMPI_ONESHOT_MAX_BYTES")
    if thr:
        _state.oneshot_max_bytes = int(thr)
    if world > 1:
        if rank == 0:
            store.set("__run_uid__", os.urandom(6).hex().encode())  # namespaces the abstract unix sockets of this job
        store.barrier("init", rank, world)
        if os.environ.get("DMPI_LAZY_SCRATCH", "0") != "1":
            _state.scratch_comm()


def destroy_process_group() -> None:
    global _state
    st = _state
    if st is None:
        return
    if st.kind == "p2p":
        torch.cuda.synchronize()
        if st.store is not None:
            try:
                st.store.barrier("destroy", st.rank, st.world, timeout_s=30)
            except Exception:
                pass
        for c in st.comms:
            c.destroy()
        if st.store is not None:
            if st.rank == 0:
                # give the other ranks a moment to read the barrier key before the server goes away
                import time

                time.sleep(0.2)
            st.store.close()
    else:
        import torch.distributed as tdist

        if tdist.is_initialized():
            tdist.destroy_process_group()
    _state = None


def barrier() -> None:
    st = _require()
    if st.kind != "p2p":
        import torch.distributed as tdist

        tdist.barrier()
        return
    torch.cuda.synchronize()
    if st.world > 1:
        st.barrier_count += 1
        st.store.barrier(f"user{st.barrier_count}", st.rank, st.world)


def _as_bytes_view(t: torch.Tensor) -> torch.Tensor:
    return t.contiguous().view(-1).view(torch.uint8)


def send(tensor: torch.Tensor, dst: int, **_ignored) -> None:
    """Blocking point-to-point send (ref:pytorch/hello_world/hello_world.py:26)."""
    st = _require()
    if st.kind != "p2p":
        import torch.distributed as tdist

        tdist.send(tensor=tensor, dst=dst)
        return
    comm = st.scratch_comm()
    src = tensor.to(st.device)
    raw = _as_bytes_view(src)
    chunk = comm.mailbox_bytes
    for o in range(0, raw.numel(), chunk):
        part = raw[o: o + chunk]
        if part.numel() % 16:
            pad = torch.zeros(16 - part.numel() % 16, dtype=torch.uint8, device=st.device)
            part = torch.cat([part, pad])
        comm.send(part, dst)
    torch.cuda.current_stream().synchronize()


def recv(tensor: torch.Tensor, src: int, **_ignored) -> None:
    """Blocking point-to-point receive into ``tensor`` (ref:pytorch/hello_world/hello_world.py:29)."""
    st = _require()
    if st.kind != "p2p":
        import torch.distributed as tdist

        tdist.recv(tensor=tensor, src=src)
        return
    comm = st.scratch_comm()
    nbytes = tensor.numel() * tensor.element_size()
    out = torch.empty((nbytes + 15) // 16 * 16, dtype=torch.uint8, device=st.device)
    chunk = comm.mailbox_bytes
    for o in range(0, nbytes, chunk):
        n = min(chunk, nbytes - o)
        comm.recv(out[o: o + (n + 15) // 16 * 16], src)
    torch.cuda.current_stream().synchronize()
    tensor.copy_(out[:nbytes].view(tensor.dtype).view(tensor.shape))


def all_reduce(tensor: torch.Tensor, average: bool = False) -> None:
    """In-place sum (or mean) over ranks of an arbitrary tensor.  P2P path: staged through the scratch arena."""
    st = _require()
    if st.kind != "p2p":
        import torch.distributed as tdist

        tdist.all_reduce(tensor)
        if average:
            tensor.div_(st.world)
        return
    if st.world == 1:
        return
    comm = st.scratch_comm()
    flat = tensor.contiguous().view(-1)
    work = flat.float() if flat.dtype != torch.float32 else flat
    cap = comm.data_bytes // 4
    stage = comm.tensor(0, comm.data_bytes).view(torch.float32)
    scale = 1.0 / st.world if average else 1.0
    for o in range(0, work.numel(), cap):
        n = min(cap, work.numel() - o)
        npad = (n + 3) // 4 * 4
        stage[:n].copy_(work[o: o + n])
        if npad > n:
            stage[n:npad].zero_()
        comm.allreduce(0, npad * 4, scale)
        work[o: o + n].copy_(stage[:n])
    if work is not flat:
        flat.copy_(work)
    if flat.data_ptr() != tensor.data_ptr():
        tensor.copy_(flat.view(tensor.shape))


def broadcast(tensor: torch.Tensor, src: int) -> None:
    st = _require()
    if st.kind != "p2p":
        import torch.distributed as tdist

        tdist.broadcast(tensor, src)
        return
    if st.world == 1:
        return
    comm = st.scratch_comm()
    raw = _as_bytes_view(tensor)
    stage = comm.tensor(0, comm.data_bytes)
    cap = comm.data_bytes
    for o in range(0, raw.numel(), cap):
        n = min(cap, raw.numel() - o)
        npad = (n + 15) // 16 * 16
        if st.rank == src:
            stage[:n].copy_(raw[o: o + n])
        comm.broadcast(0, npad, src)
        if st.rank != src:
            raw[o: o + n].copy_(stage[:n])
    if raw.data_ptr() != tensor.data_ptr():
        tensor.copy_(raw.view(tensor.dtype).view(tensor.shape))


def broadcast_object(obj, src: int = 0):
    """Returns rank ``src``'s picklable ``obj`` on every rank (small control-plane data: a palette, a config dict).
    P2P groups move it through the rendezvous store, the others through torch.distributed."""
    st = _require()
    if st.world == 1:
        return obj
    if st.kind != "p2p":
        import torch.distributed as tdist

        box = [obj if st.rank == src else None]
        tdist.broadcast_object_list(box, src=src)
        return box[0]
    import pickle

    st.barrier_count += 1
    key = f"bcast_obj{st.barrier_count}"
    if st.rank == src:
        st.store.set(key, pickle.dumps(obj))
        return obj
    return pickle.loads(st.store.get(key))
