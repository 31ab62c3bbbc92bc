"""ctypes bindings to the in-tree native libraries (built by ``build.py``).

``libdmpi_host.so``  — TCP key/value store and the reducer core; pure C++, loads anywhere.
``libdmpi_cuda.so``  — every sm_100a kernel, the P2P arena and the CUDA side of the reducer.

The CUDA library is mandatory on a GPU machine: if CUDA is available and the library is missing or
fails to load we raise instead of silently falling back to eager PyTorch.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent / "_lib"

MAX_TAPS = 32
MAX_MAPS = 4


class Tensor4(C.Structure):
    _fields_ = [("dims", C.c_int64 * 4), ("strides", C.c_int64 * 3), ("offset", C.c_int64)]


class ConvDesc(C.Structure):
    """Mirror of ``dmpi_conv_desc`` (include/dmpi_conv.h)."""

    _fields_ = [
        ("a_nviews", C.c_int),
        ("a_view", Tensor4 * MAX_MAPS),
        ("box_w", C.c_int), ("box_h", C.c_int), ("box_n", C.c_int),
        ("grid_w", C.c_int), ("grid_h", C.c_int), ("grid_n", C.c_int),
        ("ntaps", C.c_int), ("cblocks", C.c_int),
        ("tap_view", C.c_int * MAX_TAPS),
        ("tap_dw", C.c_int * MAX_TAPS),
        ("tap_dh", C.c_int * MAX_TAPS),
        ("tap_koff", C.c_int * MAX_TAPS),
        ("b_rows", C.c_int64), ("b_k", C.c_int64),
        ("o_nviews", C.c_int), ("nblk_per_oview", C.c_int),
        ("o_view", Tensor4 * MAX_MAPS),
        ("n_out", C.c_int), ("bn", C.c_int),
        ("bias_mod", C.c_int), ("stat_mod", C.c_int), ("relu", C.c_int), ("out_mode", C.c_int),
        ("out_ld", C.c_int64),
        ("k_out", C.c_int), ("c_total", C.c_int),
        ("w_ld", C.c_int64),
        ("tap_woff", C.c_int * MAX_TAPS),
        ("splits", C.c_int),
        ("cta_group", C.c_int),
        ("patch", C.c_int), ("halo", C.c_int),
    ]


_host = None
_cuda = None
_launch_count = 0


def count_launch(n: int = 1) -> None:
    """Launch accounting for bench.py's ``gpu_launches``: every wrapper that starts one of OUR kernels calls this."""
    global _launch_count
    _launch_count += n


def launches() -> int:
    return _launch_count


def _load(name: str):
    path = _LIB_DIR / name
    if not path.exists():
        raise FileNotFoundError(f"{path} not found — run `python build.py` (or __graft_entry__.build()) first")
    return C.CDLL(str(path))  # RTLD_LOCAL: never interpose symbols of other libraries


def host():
    global _host
    if _host is None:
        lib = _load("libdmpi_host.so")
        _bind_host(lib)
        _host = lib
    return _host


def cuda():
    global _cuda
    if _cuda is None:
        lib = _load("libdmpi_cuda.so")
        _bind_cuda(lib)
        _cuda = lib
    return _cuda


_torch_hooks = None


def torch_hooks():
    """The pybind11 module with the C++ autograd hooks (csrc/hooks/autograd_hooks.cpp), or None when it was not built
    or DMPI_PY_HOOKS=1 asks for the Python hooks."""
    global _torch_hooks
    import os

    if os.environ.get("DMPI_PY_HOOKS", "0") == "1":
        return None
    if _torch_hooks is None:
        path = _LIB_DIR / "dmpi_torch_hooks.so"
        if not path.exists():
            _torch_hooks = False
        else:
            import importlib.util

            import torch  # noqa: F401  (libtorch must be loaded before the extension)

            spec = importlib.util.spec_from_file_location("dmpi_torch_hooks", str(path))
            mod = importlib.util.module_from_spec(spec)
            try:
                spec.loader.exec_module(mod)
                _torch_hooks = mod
            except Exception as e:  # ABI mismatch after a torch upgrade: rebuild with `python build.py --force`
                import warnings

                warnings.warn(f"dmpi_torch_hooks.so could not be loaded ({e}); using the Python hooks")
                _torch_hooks = False
    return _torch_hooks or None


def fn_addr(lib, name: str) -> int:
    """Address of an exported C function (handed to the C++ hooks as a plain function pointer)."""
    return C.cast(getattr(lib, name), C.c_void_p).value


def cuda_available() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:  # pragma: no cover
        return False


def _sig(lib, name, restype, argtypes):
    fn = getattr(lib, name, None)
    if fn is None:
        return
    fn.restype = restype
    fn.argtypes = argtypes


VP = C.c_void_p
I = C.c_int
I64 = C.c_int64
F = C.c_float


def _bind_host(lib) -> None:
    _sig(lib, "dmpi_store_server_start", VP, [I, C.POINTER(I)])
    _sig(lib, "dmpi_store_server_stop", None, [VP])
    _sig(lib, "dmpi_store_connect", VP, [C.c_char_p, I, I])
    _sig(lib, "dmpi_store_close", None, [VP])
    _sig(lib, "dmpi_store_set", I, [VP, C.c_char_p, VP, I64])
    _sig(lib, "dmpi_store_get", I64, [VP, C.c_char_p, VP, I64, I])
    _sig(lib, "dmpi_store_add", I64, [VP, C.c_char_p, I64])
    _sig(lib, "dmpi_store_check", I, [VP, C.c_char_p])
    _sig(lib, "dmpi_plan_buckets", I, [C.POINTER(I64), I, I64, I64, C.POINTER(I)])
    _sig(lib, "dmpi_reducer_create", VP, [I, C.POINTER(I), I])
    _sig(lib, "dmpi_reducer_destroy", None, [VP])
    _sig(lib, "dmpi_reducer_reset", None, [VP])
    _sig(lib, "dmpi_reducer_mark_ready", I, [VP, I, C.POINTER(I), I])
    _sig(lib, "dmpi_reducer_num_buckets", I, [VP])
    _sig(lib, "dmpi_reducer_pending", I, [VP])
    _sig(lib, "dmpi_reducer_ready_order", I, [VP, C.POINTER(I), I])
    _sig(lib, "dmpi_reducer_set_launch_cb", None, [VP, VP, VP])
    _sig(lib, "dmpi_reducer_rebuild", I, [VP])
    _sig(lib, "dmpi_reducer_set_launch_order", I, [VP, C.POINTER(I), I])
    _sig(lib, "dmpi_reducer_launch_order", I, [VP, C.POINTER(I), I])
    _sig(lib, "dmpi_plan_hash", C.c_uint64, [C.POINTER(I64), C.POINTER(I), I])


def _bind_cuda(lib) -> None:
    _sig(lib, "dmpi_conv_plan_create", VP, [C.POINTER(ConvDesc)])
    _sig(lib, "dmpi_conv_plan_destroy", None, [VP])
    _sig(lib, "dmpi_conv_plan_set_stat_scratch", None, [VP, VP, I64])
    _sig(lib, "dmpi_conv_plan_set_splitk_scratch", None, [VP, VP, I64])
    _sig(lib, "dmpi_igemm_run", I, [VP, VP, VP, VP, VP, VP, VP, VP])
    _sig(lib, "dmpi_igemm_run_add", I, [VP, VP, VP, VP, VP, VP, VP, VP, I64, I64, I64, VP])
    _sig(lib, "dmpi_wgrad_run", I, [VP, VP, VP, VP, VP])
    _sig(lib, "dmpi_wgrad_patch_run", I, [VP, VP, VP, VP, VP])
    _sig(lib, "dmpi_umma_probe", I, [VP, VP, VP, I, I, I, VP])
    _sig(lib, "dmpi_conv_last_error", I, [C.c_char_p, I])
    # norm
    _sig(lib, "dmpi_bn_forward", I, [VP, VP, VP, I64, I, I64, I64, I64, VP, VP, VP, VP, VP, VP, VP, VP, VP, F, F, I, I, VP])
    _sig(lib, "dmpi_bn_backward", I, [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, VP, VP, VP])
    _sig(lib, "dmpi_bn_backward_ex", I, [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, I, VP, VP, VP, VP])
    _sig(lib, "dmpi_bn_forward_mask", I, [VP, VP, VP, I64, I, I64, I64, I64, VP, VP, VP, VP, VP, VP, VP, VP, VP, F, F, I, I, VP, VP])
    _sig(lib, "dmpi_bn_backward_mask", I, [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, I, VP, VP, VP, VP, VP])
    _sig(lib, "dmpi_bn_backward_det", I, [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, I, VP, VP, VP, VP, VP, I64, VP])
    _sig(lib, "dmpi_bn_stats", I, [VP, I64, I, I64, VP, VP, VP])
    # pool
    _sig(lib, "dmpi_maxpool_forward", I, [VP, VP, VP, I, I, I, I, I, I, I, I, I, I64, I64, I64, I64, I64, I64, VP])
    _sig(lib, "dmpi_maxpool_backward", I, [VP, VP, VP, I, I, I, I, I, I, I, I, I, I64, I64, I64, I64, I64, I64, VP])
    _sig(lib, "dmpi_maxpool_backward_add", I, [VP, VP, VP, I, I, I, I, I, I, I, I, I, I64, I64, I64, I64, I64, I64, VP,
                                               I64, I64, I64, VP])
    _sig(lib, "dmpi_avgpool_forward", I, [VP, VP, I, I, I, VP])
    _sig(lib, "dmpi_avgpool_backward", I, [VP, VP, I, I, I, VP])
    _sig(lib, "dmpi_upsample2x_bilinear_forward", I, [VP, VP, I, I, I, I, I64, I64, I64, I64, I64, I64, VP])
    _sig(lib, "dmpi_upsample2x_bilinear_backward", I, [VP, VP, I, I, I, I, I64, I64, I64, I64, I64, I64, VP])
    # loss / head
    _sig(lib, "dmpi_softmax_xent", I, [VP, VP, VP, VP, I, I, I64, I64, I, F, VP])
    _sig(lib, "dmpi_bce_logits", I, [VP, VP, VP, VP, I64, F, VP])
    _sig(lib, "dmpi_scale_dev", I, [VP, VP, I64, VP, VP])
    _sig(lib, "dmpi_linear_bwd_prep", I, [VP, I64, I, I, VP, I, VP, VP])
    _sig(lib, "dmpi_head_forward", I, [VP, VP, VP, VP, I64, I, I, I64, VP])
    _sig(lib, "dmpi_head_backward", I, [VP, VP, VP, VP, VP, VP, I64, I, I, I64, I64, VP])
    # optim
    _sig(lib, "dmpi_sgd_step", I, [VP, VP, VP, VP, I64, F, F, F, F, I, VP, F, VP])
    _sig(lib, "dmpi_adam_step", I, [VP, VP, VP, VP, VP, I64, F, F, F, F, F, F, F, VP, F, VP])
    _sig(lib, "dmpi_adam_step_dev", I, [VP, VP, VP, VP, VP, I64, F, F, F, F, F, VP, VP, F, VP])
    _sig(lib, "dmpi_sumsq", I, [VP, I64, VP, VP])
    _sig(lib, "dmpi_scalar_op", I, [VP, VP, I, F, VP])
    _sig(lib, "dmpi_scale_by_clip", I, [VP, I64, VP, F, VP])
    _sig(lib, "dmpi_cast_bf16", I, [VP, VP, I64, VP])
    _sig(lib, "dmpi_cast_f32", I, [VP, VP, I64, VP])
    # layout
    _sig(lib, "dmpi_memset_async", I, [VP, I, I64, VP])
    _sig(lib, "dmpi_nchw_to_nhwc8_padded", I, [VP, VP, I, I, I, I, I, I, I, VP])
    _sig(lib, "dmpi_nchw_to_nhwc", I, [VP, VP, I, I, I, VP])
    _sig(lib, "dmpi_nhwc_to_nchw", I, [VP, VP, I, I, I, I64, VP])
    _sig(lib, "dmpi_transpose_table", I, [VP, I, I, VP])
    _sig(lib, "dmpi_stem_pack", I, [VP, VP, I, I, I, I, VP])
    _sig(lib, "dmpi_stem_unpack_grad", I, [VP, VP, I, I, I, I, VP])
    # p2p arena + collectives
    _sig(lib, "dmpi_arena_create", VP, [I, I, I64, I64, I64])
    _sig(lib, "dmpi_arena_create_emulated", VP, [I, I64, I64, I64])
    _sig(lib, "dmpi_arena_destroy", None, [VP])
    _sig(lib, "dmpi_arena_create_vmm", VP, [I, I, I64, I64, I64])
    _sig(lib, "dmpi_arena_multicast_supported", I, [])
    _sig(lib, "dmpi_arena_export_fd", I, [VP])
    _sig(lib, "dmpi_arena_import_peer_fd", I, [VP, I, I])
    _sig(lib, "dmpi_arena_mc_create", I, [VP])
    _sig(lib, "dmpi_arena_mc_import", I, [VP, I])
    _sig(lib, "dmpi_arena_mc_add_device", I, [VP])
    _sig(lib, "dmpi_arena_mc_bind_map", I, [VP])
    _sig(lib, "dmpi_fd_listen", I, [C.c_char_p])
    _sig(lib, "dmpi_fd_send", I, [C.c_char_p, I, I])
    _sig(lib, "dmpi_fd_recv", I, [I, C.POINTER(I), I])
    _sig(lib, "dmpi_fd_close", None, [I])
    _sig(lib, "dmpi_allreduce_nvls", I, [VP, I64, I64, I, F, I, VP])
    _sig(lib, "dmpi_arena_ipc_handle", I, [VP, VP])
    _sig(lib, "dmpi_arena_open_peers", I, [VP, VP])
    _sig(lib, "dmpi_arena_ptrs", VP, [VP])
    _sig(lib, "dmpi_arena_data", VP, [VP, I])
    _sig(lib, "dmpi_arena_stage_slot_bytes", I64, [VP])
    _sig(lib, "dmpi_arena_mailbox", VP, [VP, I, I])
    _sig(lib, "dmpi_arena_seq_flag", VP, [VP, I, I])
    _sig(lib, "dmpi_arena_ack_flag", VP, [VP, I, I])
    _sig(lib, "dmpi_arena_mailbox_bytes", I64, [VP])
    _sig(lib, "dmpi_arena_last_error", C.c_char_p, [])
    _sig(lib, "dmpi_allreduce_twoshot", I, [VP, I64, I64, I, F, I, I, VP])
    _sig(lib, "dmpi_allreduce_oneshot", I, [VP, I64, I64, I, F, I64, I, I, VP])
    _sig(lib, "dmpi_broadcast", I, [VP, I64, I64, I, I, I, VP])
    _sig(lib, "dmpi_comm_set_timeout", I, [C.c_double])
    _sig(lib, "dmpi_peer_copy", I, [VP, VP, I64, I, VP])
    _sig(lib, "dmpi_send", I, [VP, VP, I64, VP, VP, C.c_uint32, VP])
    _sig(lib, "dmpi_recv", I, [VP, VP, I64, VP, VP, C.c_uint32, VP])
    _sig(lib, "dmpi_sendrecv_emulated", I, [VP, VP, VP, I64, VP, VP, C.c_uint32, VP])
    # reducer (CUDA glue)
    _sig(lib, "dmpi_reducer_cuda_create", VP, [VP, I, C.POINTER(I64), C.POINTER(I64), F, I64, I64, I, I])
    _sig(lib, "dmpi_reducer_cuda_destroy", None, [VP])
    _sig(lib, "dmpi_reducer_cuda_set_wire_buffer", I, [VP, I64, I64])
    _sig(lib, "dmpi_reducer_cuda_begin", None, [VP, VP])
    _sig(lib, "dmpi_reducer_cuda_launch", None, [VP, I])
    _sig(lib, "dmpi_reducer_cuda_finalize", I, [VP])
    _sig(lib, "dmpi_reducer_cuda_side_broadcast", I, [VP, I64, I64, I, I, VP])
    _sig(lib, "dmpi_reducer_cuda_join_broadcast", None, [VP, VP])
    _sig(lib, "dmpi_reducer_cuda_comm_stream", VP, [VP])
    _sig(lib, "dmpi_reducer_cuda_launched", I, [VP])
    _sig(lib, "dmpi_reducer_cuda_launch_fn", VP, [])


def conv_last_error() -> str:
    buf = C.create_string_buffer(512)
    cuda().dmpi_conv_last_error(buf, 512)
    return buf.value.decode()


def stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream
