"""ctypes binding of libb200ctc.so (the C ABI declared in include/b200ctc.h).

The library is loaded lazily.  There is NO fallback: if the shared library is
missing, or a compute call is made without a CUDA device, a ``RuntimeError`` is
raised - the reference's torch path is never used silently instead.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ctc.so")
_lib = None

DTYPE_CODE = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}

_P = c_void_p
# name -> (restype, argtypes); mirrors include/b200ctc.h one to one
SIGNATURES = {
    "b200ctc_abi_version": (c_int, []),
    "b200ctc_last_error": (c_char_p, []),
    "b200ctc_launch_count": (c_int64, []),
    "b200ctc_reset_launch_count": (None, []),
    "b200ctc_profile_enable": (None, [c_int]),
    "b200ctc_profile_read": (c_int, [POINTER(c_float), POINTER(c_int), c_int]),
    "b200ctc_device_info": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "b200ctc_softmax_fwd": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, _P]),
    "b200ctc_softmax_fwd_bias": (c_int, [_P, _P, _P, c_int64, c_int, c_int, c_int, _P]),
    "b200ctc_softmax_bwd": (c_int, [_P, _P, _P, c_int64, c_int, c_int, c_int, _P]),
    "b200ctc_argmax": (c_int, [_P, _P, c_int64, c_int, c_int, _P]),
    "b200ctc_argmax_bias": (c_int, [_P, _P, _P, c_int64, c_int, c_int, _P]),
    "b200ctc_greedy_decode": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, _P]),
    "b200ctc_loss_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200ctc_loss_fwd": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                 c_int, c_int, _P, _P, c_float, _P, c_size_t, _P]),
    "b200ctc_loss_bwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P,
                                 c_int, c_float, _P, _P, c_size_t, _P]),
    "b200ctc_loss_fwd_grad": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                      c_int, c_float, _P, _P, c_float, _P, _P, c_size_t, _P]),
    "b200ctc_loss_fwd_grad_sharded": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                              c_int, c_int, c_float, _P, _P, c_float, _P, _P,
                                              c_size_t, _P, _P, _P]),
    "b200ctc_loss_fwd_grad_hints": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                            c_int, _P, _P, _P, _P, c_size_t, _P]),
    "b200ctc_grad_rescale_hints": (c_int, [_P, c_int, _P, c_int, c_float, _P, c_int, c_int, c_int, _P]),
    "b200ctc_peer_create": (c_int, [c_int, c_int, _P, _P]),
    "b200ctc_peer_connect": (c_int, [_P, _P]),
    "b200ctc_peer_destroy": (c_int, [_P]),
    "b200ctc_peer_error": (c_int, [_P]),
    "b200ctc_peer_exchange": (c_int, [_P, _P, _P, _P]),
    "b200ctc_peer_configure": (c_int, [_P, c_int, c_int]),
    "b200ctc_peer_wait": (c_int, [_P, _P, _P]),
    "b200ctc_grad_rescale": (c_int, [_P, c_int, _P, c_int, c_float, c_float, c_int, c_int, c_int,
                                     _P]),
    "b200ctc_align_workspace_bytes": (c_size_t, [c_int, c_int]),
    "b200ctc_forced_align": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P, _P, c_size_t, _P]),
    "b200ctc_forced_align_static": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P, _P]),
    "b200ctc_prefix_pad": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_float, _P]),
    "b200ctc_prefix_init_state": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, c_float, _P, _P]),
    "b200ctc_prefix_window": (c_int, [_P, c_int, c_int, _P, _P]),
    "b200ctc_prefix_transpose": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "b200ctc_prefix_step": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                    c_int, c_int, c_int, c_int, c_int, c_float, _P, _P, _P, _P,
                                    _P, c_int, _P, _P]),
    "b200ctc_prefix_window_utt": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P]),
    "b200ctc_prefix_select": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                      _P, _P, _P]),
    "b200ctc_beam_prebeam": (c_int, [_P, _P, _P, c_int, c_int64, c_int, c_int, _P, _P]),
    "b200ctc_beam_select": (c_int, [_P, _P, _P, c_int, _P, _P, _P, c_int64, c_int, _P, _P, c_int,
                                    c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P, c_int64,
                                    _P, _P]),
    "b200ctc_beam_plan_bytes": (c_size_t, []),
    "b200ctc_beam_position": (c_int, [_P, _P, _P, c_int, c_int, _P]),
    "b200ctc_loss_fwd_bwd_host": (c_int, [_P, c_int, _P, _P, _P, c_int, c_int, c_int, c_int,
                                          c_int, c_float, _P, _P, POINTER(c_void_p), c_int]),
    "b200ctc_host_pipeline_release": (None, []),
}


class BeamPlan(ctypes.Structure):
    """``b200ctc_beam_plan`` (include/b200ctc.h): the per-decode buffers of
    ``b200ctc_beam_position``.  ``lib()`` checks the size against the library's own."""

    _fields_ = [
        ("U", ctypes.c_int32), ("W", ctypes.c_int32), ("V", ctypes.c_int32), ("T", ctypes.c_int32),
        ("snum", ctypes.c_int32), ("K_full", ctypes.c_int32), ("prebeam_key", ctypes.c_int32),
        ("blank", ctypes.c_int32), ("eos", ctypes.c_int32), ("logzero", c_float),
        ("weights", c_float * 8),
        ("lp", _P), ("lp_t", _P), ("lp_t_ld", c_int64), ("end_frames", _P),
        ("r_state", _P * 2), ("s_state", _P * 2), ("r", _P), ("log_psi", _P), ("ctc_scores", _P),
        ("idmap", _P), ("rsum", _P), ("part_ids", _P),
        ("run_score", _P * 2), ("cum", _P * 2), ("yseq", _P * 2), ("yseq_stride", c_int64),
        ("active", _P), ("maxlen", _P), ("out_val", _P), ("out_flat", _P),
        ("pack", _P), ("pack_host", _P),
    ]


def lib():
    """Load (once) and return the ctypes handle; raises if the .so is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m espnet1_b200.build` "
                "(there is no CPU / PyTorch fallback for the CTC hot path)")
        h = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        if h.b200ctc_beam_plan_bytes() != ctypes.sizeof(BeamPlan):
            raise RuntimeError("b200ctc_beam_plan: the ctypes mirror and the library disagree on the layout "
                               f"({ctypes.sizeof(BeamPlan)} vs {h.b200ctc_beam_plan_bytes()} bytes)")
        _lib = h
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200ctc_last_error().decode(errors="replace")
        raise RuntimeError(f"libb200ctc {what} failed (status {rc}): {msg}")


def _find_cuda_device(args):
    for a in args:
        if isinstance(a, torch.Tensor):
            if a.is_cuda:
                return a.device
        elif isinstance(a, (list, tuple)) and a and isinstance(a[0], torch.Tensor) and a[0].is_cuda:
            return a[0].device
        else:
            d = getattr(a, "device", None)  # a scorer object (self) that remembers its device
            if isinstance(d, torch.device) and d.type == "cuda":
                return d
    return None


def on_tensor_device(fn):
    """Run ``fn`` with the CUDA device of its first CUDA tensor argument (or of ``self.device``)
    current.  The library launches on the process's CURRENT device and takes the current stream
    of the tensor's device: with tensors on cuda:1 and cuda:0 current (a model moved with
    ``.to('cuda:1')`` and no ``set_device``) that would be device-1 pointers in a device-0
    launch.  The previous device is restored on return."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        dev = _find_cuda_device(args)
        if dev is not None and dev.index is not None and dev.index != torch.cuda.current_device():
            with torch.cuda.device(dev):
                return fn(*args, **kwargs)
        return fn(*args, **kwargs)

    return wrapper


def stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"{name} must be a CUDA tensor: the B200 CTC path has no CPU fallback "
            f"(got device {t.device})")


def dtype_code(t: torch.Tensor) -> int:
    try:
        return DTYPE_CODE[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}; expected fp32/fp16/bf16") from None


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def launch_count() -> int:
    return int(lib().b200ctc_launch_count())


def reset_launch_count() -> None:
    lib().b200ctc_reset_launch_count()
