"""DistributedDataParallel-compatible wrapper.

    ddp_model = DistributedDataParallel(model, device_ids=[LOCAL_RANK], output_device=LOCAL_RANK)

is the exact call of ref:pytorch/resnet/main.py:44-46 and ref:pytorch/unet/train.py:68-70; ``.train() / .eval() /
.parameters() / .state_dict()`` (``module.``-prefixed keys) / ``.load_state_dict()`` / ``__call__`` behave the same.

What happens underneath is different from c10d:

* wrap time: parameters, gradients and buffers are re-homed into ONE peer-mapped symmetric arena
  (``[grads | params | buffers]``); shapes are verified across ranks by exchanging a layout hash through the store;
  rank 0's parameters and buffers are pushed to every peer by the NVLink broadcast kernel.
* buckets are contiguous byte ranges of the gradient arena in reverse registration order (1 MiB first bucket,
  then ``bucket_cap_mb``).  Gradients are produced *in* the bucket (wgrad kernels accumulate into their slot), the
  C++ reducer counts readiness and launches the fused (sum * 1/world) all-reduce kernel for bucket b on a
  high-priority comm stream as soon as buckets 0..b are complete — overlapping the rest of backward.
* every training forward re-broadcasts rank 0's buffers (BatchNorm running statistics) like
  ``broadcast_buffers=True`` does — with ONE kernel over the contiguous buffer range, issued on the comm stream right
  after the forward pass was enqueued (the bytes torch would send at the head of the next forward), so its two
  cross-GPU barriers overlap the backward pass instead of delaying the next forward.  Forwards under
  ``torch.no_grad()`` never communicate, so a rank-0-only evaluation between epochs
  (ref:pytorch/resnet/main.py:136-139) cannot dead-lock the other ranks.
* ``no_sync()`` accumulates locally.  On gloo / NCCL-fallback groups the same reducer drives
  ``torch.distributed.all_reduce`` on the flat bucket views.
* after the first complete backward pass the bucket LAUNCH ORDER is re-derived from the observed gradient-ready order
  (c10d rebuilds its buckets at the same point) and cross-checked between the ranks.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
from typing import List, Optional

import torch
import torch.nn as nn

from .. import _native as N
from ..utils.arena import ParamArena, _round_up, ensure_arena
from ..utils.runtime import FaultInjector
from . import dist as D

_MIB = 1 << 20


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids: Optional[List[int]] = None, output_device=None,
                 broadcast_buffers: bool = True, bucket_cap_mb: float = 25.0, first_bucket_mb: float = 1.0,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = True,
                 grad_wire_dtype: Optional[str] = None, **_ignored):
        super().__init__()
        st = D._require()
        self.module = module
        self.broadcast_buffers = broadcast_buffers
        self.require_backward_grad_sync = True
        self._st = st
        self.world, self.rank = st.world, st.rank
        self._fault = FaultInjector()  # DMPI_FAULT test hook; inactive unless the variable is set
        # gradient wire format of the large buckets: None / "fp32" = exact fp32 sums (default), "bf16" = each rank's
        # bucket is rounded to bf16 for the transfer and summed in fp32 (half the NVLink bytes; DMPI_WIRE_BF16=1)
        if grad_wire_dtype is None and os.environ.get("DMPI_WIRE_BF16", "0") == "1":
            grad_wire_dtype = "bf16"
        if grad_wire_dtype not in (None, "fp32", "bf16"):
            raise ValueError("grad_wire_dtype must be None, 'fp32' or 'bf16'")
        self.grad_wire_dtype = grad_wire_dtype or "fp32"
        if find_unused_parameters:
            raise NotImplementedError("find_unused_parameters is not used by the reference and not supported")
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter "
                               "that requires a gradient.")
        self._is_cuda = params[0].is_cuda
        if st.kind == "p2p" and not self._is_cuda:
            raise RuntimeError("the p2p/nccl backend needs the module on its CUDA device before wrapping "
                               "(model.to(device), ref:pytorch/resnet/main.py:42)")
        # ---- re-home params / grads / buffers
        self._comm = None
        if st.kind == "p2p":
            total = sum(_round_up(p.numel()) for p in module.parameters())
            buf_bytes = sum(b.numel() * b.element_size() + 16 for b in module.buffers()) + 512
            g_bytes = _round_up(total * 4, 256)
            p_bytes = g_bytes
            b_bytes = _round_up(buf_bytes, 256)
            w_bytes = _round_up(g_bytes // 2, 256) if self.grad_wire_dtype == "bf16" else 0
            self._comm = st.new_comm(g_bytes + p_bytes + b_bytes + w_bytes)
            self._off = {"grads": 0, "params": g_bytes, "buffers": g_bytes + p_bytes, "wire": g_bytes + p_bytes + b_bytes}
            self._sizes = {"grads": g_bytes, "params": p_bytes, "buffers": b_bytes, "wire": w_bytes}
            comm = self._comm

            def alloc(name: str, nbytes: int) -> torch.Tensor:
                assert nbytes <= self._sizes[name], (name, nbytes, self._sizes[name])
                return comm.tensor(self._off[name], nbytes)

            self.arena: ParamArena = ensure_arena(module, alloc)
        else:
            self.arena = ensure_arena(module)
        arena = self.arena
        # ---- bucket plan over the trainable slots (already in bucket-fill order; frozen parameters sit behind them
        #      in the arena and take no part in gradient reduction)
        slots = arena.slots[: arena.n_trainable]
        self._slots = slots
        sizes = (C.c_int64 * len(slots))(*[_round_up(s.numel) * 4 for s in slots])
        bucket_of = (C.c_int * len(slots))()
        host = N.host()
        self.n_buckets = host.dmpi_plan_buckets(sizes, len(slots), int(first_bucket_mb * _MIB),
                                                int(bucket_cap_mb * _MIB), bucket_of)
        self.bucket_of = list(bucket_of)
        self.bucket_ranges = []  # (element offset, element count) per bucket
        for b in range(self.n_buckets):
            idx = [i for i, bb in enumerate(self.bucket_of) if bb == b]
            lo = slots[idx[0]].offset
            hi = slots[idx[-1]].offset + _round_up(slots[idx[-1]].numel)
            self.bucket_ranges.append((lo, hi - lo))
        # ---- verify the layout on every rank (N5), then sync state from rank 0 (N6)
        logical = (C.c_int64 * len(slots))(*[s.numel for s in slots])
        self.plan_hash = int(host.dmpi_plan_hash(logical, bucket_of, len(slots)))
        self._verify_across_ranks()
        self._sync_module_states()
        # ---- reducer
        self._reducer = host.dmpi_reducer_create(len(slots), bucket_of, self.n_buckets)
        self._cuda_red = None
        if st.kind == "p2p":
            lib = N.cuda()
            offs = (C.c_int64 * self.n_buckets)(*[self._off["grads"] + lo * 4 for lo, _ in self.bucket_ranges])
            nbytes = (C.c_int64 * self.n_buckets)(*[n * 4 for _, n in self.bucket_ranges])
            self._cuda_red = lib.dmpi_reducer_cuda_create(self._comm.ptrs, self.n_buckets, offs, nbytes,
                                                          1.0 / self.world, st.oneshot_max_bytes,
                                                          self._comm.slot_bytes, 32,
                                                          1 if self.grad_wire_dtype == "bf16" else 0)
            if self.grad_wire_dtype == "bf16":
                if lib.dmpi_reducer_cuda_set_wire_buffer(self._cuda_red, self._off["wire"], self._sizes["wire"]) != 0:
                    raise RuntimeError("DDP: bf16 wire buffer could not be attached")
            host.dmpi_reducer_set_launch_cb(self._reducer, lib.dmpi_reducer_cuda_launch_fn(), self._cuda_red)
        self._out_buckets = (C.c_int * max(self.n_buckets, 1))()
        self._pending_work = []
        self._armed = False
        self._finalize_queued = False
        self._marked = bytearray(len(slots))
        self._uid = f"{st.n_comms}_{self.plan_hash & 0xffffffff:x}"  # same on every rank: comms are created collectively
        self._order_final = os.environ.get("DMPI_BUCKET_REBUILD", "1") != "1"   # launch order not yet rebuilt
        self._bcast_inline = os.environ.get("DMPI_BCAST_INLINE", "0") == "1"    # A/B: broadcast at the head of forward
        self.launch_order = list(range(self.n_buckets))
        self._install_hooks()

    # ------------------------------------------------------------------ construction helpers
    def _verify_across_ranks(self) -> None:
        st = self._st
        if self.world == 1:
            return
        if st.kind == "p2p":
            tag = f"verify{st.n_comms}"
            all_h = st.store.all_gather(tag, self.rank, self.world, str(self.plan_hash).encode())
            if any(h != all_h[0] for h in all_h):
                raise RuntimeError("DDP: parameter shapes / bucket plan differ across ranks "
                                   f"(hashes {[h.decode() for h in all_h]})")
        else:
            import torch.distributed as tdist

            h = torch.tensor([self.plan_hash & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64,
                             device=self.arena.device if self._is_cuda else "cpu")
            parts = [torch.zeros_like(h) for _ in range(self.world)]
            tdist.all_gather(parts, h)  # symmetric: every rank sees every hash, so all ranks raise together
            if any(int(p.item()) != int(parts[0].item()) for p in parts):
                raise RuntimeError("DDP: parameter shapes / bucket plan differ across ranks")

    def _sync_module_states(self) -> None:
        """Wrap-time broadcast of parameters and buffers from rank 0 (c10d: _sync_module_states)."""
        if self.world == 1:
            return
        st = self._st
        if st.kind == "p2p":
            self._comm.broadcast(self._off["params"], self._sizes["params"] + self._sizes["buffers"], 0, nblocks=32)
            torch.cuda.current_stream().synchronize()
            self.arena.refresh_shadows(force=True)
        else:
            import torch.distributed as tdist

            with torch.no_grad():
                tdist.broadcast(self.arena.flat_p, 0)
            self._broadcast_buffers_torch()

    def _install_hooks(self) -> None:
        ext = N.torch_hooks() if self._cuda_red is not None else None
        self._hx = None
        if ext is not None:
            # C++ path (csrc/hooks/autograd_hooks.cpp): post-accumulate-grad hooks and the finalize callback are C++
            # objects; the fused kernels' wrappers report readiness through a pybind call that drops the GIL
            from ..utils.arena import _logical_view

            host, lib = N.host(), N.cuda()
            self._ext = ext
            self._hx = ext.create(self._reducer, self._cuda_red, N.fn_addr(host, "dmpi_reducer_mark_ready"),
                                  N.fn_addr(host, "dmpi_reducer_pending"), N.fn_addr(lib, "dmpi_reducer_cuda_finalize"),
                                  self.n_buckets)
            flags = torch.zeros(len(self._slots), dtype=torch.uint8)
            for i, slot in enumerate(self._slots):
                flags[i] = 1 if slot.fused else 0
            self.arena.fused_flags = flags
            ext.install(self._hx, [s.param for s in self._slots],
                        [_logical_view(self.arena.flat_g, s) for s in self._slots], [s.name for s in self._slots], flags)
            for i, slot in enumerate(self._slots):
                slot.ready_cb = (lambda idx=i: ext.mark_ready(self._hx, idx))
            return
        for i, slot in enumerate(self._slots):
            slot.ready_cb = (lambda idx=i: self._mark_ready(idx))
            # parameters whose gradient comes from stock autograd (not from our fused backward)
            slot.param.register_post_accumulate_grad_hook(lambda p, idx=i: self._autograd_hook(idx))

    # ------------------------------------------------------------------ per-iteration machinery
    def _autograd_hook(self, idx: int) -> None:
        slot = self.arena.slots[idx]
        if getattr(slot, "fused", False):
            # written by a fused backward kernel (possibly still running on the weight-gradient side stream): only
            # arena.grad_ready(), called once that kernel is ordered before the compute stream, may publish it
            return
        g = slot.param.grad
        base = self.arena.flat_g.data_ptr() + slot.offset * 4
        if g is not None and g.data_ptr() != base:
            # autograd installed a fresh tensor (e.g. after zero_grad(set_to_none=True)): copy into the bucket view
            from ..utils.arena import _logical_view

            view = _logical_view(self.arena.flat_g, slot)
            view.copy_(g)
            slot.param.grad = view
        if self._marked[idx]:
            return  # the fused backward already published this slot in this iteration
        self._mark_ready(idx)

    def _mark_ready(self, idx: int) -> None:
        if self._hx is not None:
            self._ext.mark_ready(self._hx, idx)
            return
        if not self._armed:
            return
        if not self._finalize_queued:
            self._finalize_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
        self._marked[idx] = 1
        n = N.host().dmpi_reducer_mark_ready(self._reducer, idx, self._out_buckets, self.n_buckets)
        if n == -1:
            raise RuntimeError(f"DDP: parameter '{self.arena.slots[idx].name}' was marked ready twice in one "
                               "backward pass (shared module used twice?)")
        if n > 0:
            if self._cuda_red is not None:
                N.count_launch(n if self.world > 1 else 0)
            else:
                self._launch_torch_buckets([self._out_buckets[i] for i in range(n)])

    def _launch_torch_buckets(self, buckets: List[int]) -> None:
        import torch.distributed as tdist

        for b in buckets:
            lo, n = self.bucket_ranges[b]
            view = self.arena.flat_g[lo: lo + n]
            if self.world > 1:
                view.div_(self.world)
                self._pending_work.append(tdist.all_reduce(view, async_op=True))

    def _finalize_backward(self) -> None:
        self._finalize_queued = False
        if not self._armed:
            return
        self._armed = False
        pending = N.host().dmpi_reducer_pending(self._reducer)
        if pending != 0:
            raise RuntimeError(f"DDP: backward finished with {pending} gradient bucket(s) incomplete — some "
                               "parameters did not receive a gradient (unused parameters are not supported)")
        if self._cuda_red is not None:
            rc = N.cuda().dmpi_reducer_cuda_finalize(self._cuda_red)
            if rc != 0:
                raise RuntimeError(f"DDP: all-reduce launch failed ({rc})")
        for w in self._pending_work:
            w.wait()
        self._pending_work.clear()
        if not self._order_final:
            self._rebuild_launch_order()

    def _rebuild_launch_order(self) -> None:
        """Once, after the first complete backward pass: launch buckets in the order in which they became complete
        (every rank observed the same autograd graph; the result is still compared across ranks, and on any
        disagreement all ranks keep the initial order)."""
        self._order_final = True
        host = N.host()
        if host.dmpi_reducer_rebuild(self._reducer) < 0:
            return
        buf = (C.c_int * max(self.n_buckets, 1))()
        host.dmpi_reducer_launch_order(self._reducer, buf, self.n_buckets)
        order = [buf[i] for i in range(self.n_buckets)]
        if self.world > 1:
            st = self._st
            if st.kind == "p2p":
                tag = f"order{self._uid}"
                mine = ",".join(map(str, order)).encode()
                agree = all(o == mine for o in st.store.all_gather(tag, self.rank, self.world, mine))
            else:
                import torch.distributed as tdist

                t = torch.tensor(order, dtype=torch.int64, device=self.arena.device if self._is_cuda else "cpu")
                parts = [torch.zeros_like(t) for _ in range(self.world)]
                tdist.all_gather(parts, t)
                agree = all(torch.equal(p, parts[0]) for p in parts)
            if not agree:
                ident = (C.c_int * self.n_buckets)(*range(self.n_buckets))
                host.dmpi_reducer_set_launch_order(self._reducer, ident, self.n_buckets)
                order = list(range(self.n_buckets))
        self.launch_order = order

    def _sync_buffers(self) -> None:
        if self.world == 1 or not self.broadcast_buffers or self.arena.buffer_bytes == 0:
            return
        if self._st.kind == "p2p":
            self._comm.broadcast(self._off["buffers"], _round_up(self.arena.buffer_bytes, 256), 0, nblocks=4)
        else:
            self._broadcast_buffers_torch()

    def _broadcast_buffers_torch(self) -> None:
        """gloo / NCCL-fallback groups: coalesce per dtype, broadcast from rank 0, scatter back."""
        import torch.distributed as tdist

        arena = self.arena
        with torch.no_grad():
            if arena.flat_b is not None:
                tdist.broadcast(arena.flat_b, 0)
                return
            for dtype in (torch.float32, torch.int64):
                group = [b for b in arena.buffer_list if b.dtype == dtype]
                if not group:
                    continue
                flat = torch.cat([b.reshape(-1) for b in group])
                tdist.broadcast(flat, 0)
                o = 0
                for b in group:
                    b.copy_(flat[o: o + b.numel()].view(b.shape))
                    o += b.numel()

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def forward(self, *inputs, **kwargs):
        sync = torch.is_grad_enabled() and self.training
        side = (sync and self._cuda_red is not None and self.world > 1 and self.broadcast_buffers
                and self.arena.buffer_bytes > 0 and not self._bcast_inline)
        if self._cuda_red is not None:
            # a buffer broadcast issued after the previous forward must have landed before BatchNorm touches the buffers
            N.cuda().dmpi_reducer_cuda_join_broadcast(self._cuda_red, N.stream_ptr())
        if sync:
            if self._fault.active and self.require_backward_grad_sync:
                self._fault.tick(self.rank)
            if not side:
                self._sync_buffers()
            if self.require_backward_grad_sync:
                self._arm()
            else:
                self._armed = False
                if self._hx is not None:
                    self._ext.arm(self._hx, False)
        out = self.module(*inputs, **kwargs)
        if side:
            rc = N.cuda().dmpi_reducer_cuda_side_broadcast(self._cuda_red, self._off["buffers"],
                                                           _round_up(self.arena.buffer_bytes, 256), 0, 4, N.stream_ptr())
            N.count_launch()
            if rc != 0:
                raise RuntimeError(f"DDP: buffer broadcast launch failed ({rc})")
        return out

    def _arm(self) -> None:
        """Prepare the reducer for one synchronising backward pass (c10d: prepare_for_backward)."""
        if self._hx is not None:
            if self.world > 1:
                N.count_launch(self._ext.take_launched(self._hx))
            if not self._order_final and self._ext.completed(self._hx) > 0:
                self._rebuild_launch_order()  # needs the ready order of the finished pass: before the reset below
        N.host().dmpi_reducer_reset(self._reducer)
        if self._cuda_red is not None:
            N.cuda().dmpi_reducer_cuda_begin(self._cuda_red, N.stream_ptr())
        self._armed = True
        self._marked = bytearray(len(self._slots))
        if self._hx is not None:
            self._ext.arm(self._hx, True)

    def __del__(self):  # pragma: no cover
        try:
            if getattr(self, "_hx", None) is not None:
                # the C++ hooks stay installed on the parameters: disarmed, they never touch the reducer again
                self._ext.arm(self._hx, False)
        except Exception:
            pass
        try:
            if getattr(self, "_reducer", None):
                N.host().dmpi_reducer_destroy(self._reducer)
            if getattr(self, "_cuda_red", None):
                N.cuda().dmpi_reducer_cuda_destroy(self._cuda_red)
        except Exception:
            pass
