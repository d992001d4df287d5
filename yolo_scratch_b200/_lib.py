"""ctypes binding of libys_b200.so (include/ys_b200.h).

There is no fallback: if the shared library is missing or a symbol is absent this
module raises, and every product entry point raises with it.  The library is built
in-tree by `python -m yolo_scratch_b200.build` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os
import re
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libys_b200.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "ys_b200.h")

_lock = threading.Lock()
_lib = None

c_void_p, c_int, c_float, c_size_t = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes)
PROTOTYPES = {
    "ys_version": (c_int, []),
    "ys_last_error": (C.c_char_p, []),
    "ys_device_info": (c_int, [C.POINTER(c_int)] * 3),
    "ys_detection_loss_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "ys_detection_loss": (c_int, [c_int, C.POINTER(c_void_p), C.POINTER(c_void_p), C.POINTER(c_int), C.POINTER(c_int),
                                  C.POINTER(c_float), c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                  c_float, c_float, c_float, c_float, c_float, c_float, c_float, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "ys_scale_inplace": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p]),
    "ys_assign_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "ys_assign": (c_int, [c_int, c_int, c_int, c_int] + [c_void_p] * 6 + [c_float] * 3 + [c_void_p] * 6 +
                  [c_void_p, c_size_t, c_void_p]),
    "ys_assign_ex_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "ys_assign_ex": (c_int, [c_int, c_int, c_int, c_int] + [c_void_p] * 6 + [c_float] * 3 + [c_int, c_int] + [c_void_p] * 6 +
                     [c_void_p, c_size_t, c_void_p]),
    "ys_decode_infer": (c_int, [c_int, C.POINTER(c_void_p), C.POINTER(c_int), C.POINTER(c_int), C.POINTER(c_float),
                                c_int, c_int, c_int, c_void_p, c_int, c_void_p]),
    "ys_nms_workspace_bytes": (c_size_t, [c_int, c_int]),
    "ys_nms_batched": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_float, c_int] +
                       [c_void_p] * 5 + [c_void_p, c_size_t, c_void_p]),
    "ys_nms_single": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p, c_size_t,
                              c_void_p]),
    "ys_pairwise_iou": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "ys_match_detections": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_float,
                                    C.POINTER(C.c_double), c_int, c_void_p, c_void_p, c_void_p]),
    "ys_cda_nms": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, C.c_double, c_void_p, c_void_p, c_void_p]),
    "ys_bbox_loss_workspace_bytes": (c_size_t, [c_int]),
    "ys_bbox_loss": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p] * 3 + [c_void_p, c_size_t, c_void_p]),
    "ys_bbox_loss_static_max_rows": (c_int, []),
    "ys_bbox_loss_static": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p] * 3 + [c_void_p]),
    "ys_bbox_iou": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
}


def header_symbols():
    """Every function name include/ys_b200.h declares."""
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ys_[a-z0-9_]+)\s*\(", text)))


class YsError(RuntimeError):
    pass


def lib():
    """Load (once) and return the CDLL.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(SO_PATH):
            raise YsError(f"{SO_PATH} is missing: build it with `python -m yolo_scratch_b200.build` "
                          "(there is no CPU or PyTorch fallback for the hot path)")
        dll = C.CDLL(SO_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(dll, name)     # AttributeError if the symbol is absent: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = dll
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().ys_last_error().decode(errors="replace")
        raise YsError(f"{what} failed (rc={rc}): {msg}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise YsError(f"{name} must be a CUDA tensor: the B200 hot path has no CPU fallback")


class on_device:
    """`with on_device(t0, t1, ...):` -- every tensor must live on ONE CUDA device, which becomes the current device
    for the ctypes launches inside (the library launches on the process's current device; a tensor on another GPU
    would otherwise give illegal addresses or a stream of the wrong device)."""

    def __init__(self, *tensors):
        devs = {t.device for t in tensors if t is not None}
        if len(devs) != 1:
            raise YsError(f"all tensors of one call must be on the same CUDA device, got {sorted(map(str, devs))}")
        (self.dev,) = devs
        if self.dev.type != "cuda":
            raise YsError("the B200 hot path has no CPU fallback: tensors must be CUDA tensors")
        self.ctx = torch.cuda.device(self.dev)

    def __enter__(self):
        self.ctx.__enter__()
        return self.dev

    def __exit__(self, *exc):
        return self.ctx.__exit__(*exc)


class Workspace:
    """Grow-only byte buffer on one device, 256-byte aligned (torch CUDA allocations are 512-aligned)."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != torch.device(device):
            self.buf = torch.zeros(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=device)
        return self.buf
