"""CPU: the C-ABI library loads, exports every symbol include/ys_b200.h declares, and validates
arguments before touching the GPU (no compute launches here)."""
import ctypes as C

import pytest

from yolo_scratch_b200 import _lib


def test_exports_every_declared_symbol():
    lib = _lib.lib()
    declared = _lib.header_symbols()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ys_b200.h but not exported"
        assert name in _lib.PROTOTYPES, f"{name} has no ctypes prototype"
    assert lib.ys_version() >= 100


def test_workspace_sizes_grow():
    lib = _lib.lib()
    a = lib.ys_detection_loss_workspace_bytes(2, 8400, 40, 80)
    b = lib.ys_detection_loss_workspace_bytes(64, 8400, 1280, 80)
    assert 0 < a < b
    assert lib.ys_assign_workspace_bytes(2, 8400, 20) > 0
    assert lib.ys_nms_workspace_bytes(256, 8400) > 256 * 8400 * 8


def test_bad_arguments_are_rejected_without_a_launch():
    lib = _lib.lib()
    # n_levels = 0
    rc = lib.ys_detection_loss(0, None, None, None, None, None, 2, 80, 16, None, None, None, 0, 640.0, 640.0, 7.5, 0.5,
                               1.5, 0.5, 4.0, C.c_void_p(16), None, None, None, 0, C.c_void_p(256), 1 << 20, None)
    assert rc == -1 and b"n_levels" in lib.ys_last_error()
    # reg_max != 16 is unsupported
    xs = (C.c_void_p * 1)(256)
    hs = (C.c_int * 1)(4)
    ws = (C.c_int * 1)(4)
    st = (C.c_float * 1)(8.0)
    rc = lib.ys_detection_loss(1, xs, None, hs, ws, st, 2, 80, 8, None, None, None, 0, 32.0, 32.0, 7.5, 0.5, 1.5, 0.5, 4.0,
                               C.c_void_p(16), None, None, None, 0, C.c_void_p(256), 1 << 20, None)
    assert rc == -3
    # workspace too small
    rc = lib.ys_detection_loss(1, xs, None, hs, ws, st, 2, 80, 16, None, None, None, 0, 32.0, 32.0, 7.5, 0.5, 1.5, 0.5,
                               4.0, C.c_void_p(16), None, None, None, 0, C.c_void_p(256), 16, None)
    assert rc == -2 and b"workspace" in lib.ys_last_error()
    # level pointer that is not even 4-byte aligned (16-byte alignment is no longer required: such levels take the
    # plain-load loaders)
    xs_bad = (C.c_void_p * 1)(258)
    rc = lib.ys_detection_loss(1, xs_bad, None, hs, ws, st, 2, 80, 16, None, None, None, 0, 32.0, 32.0, 7.5, 0.5, 1.5,
                               0.5, 4.0, C.c_void_p(16), None, None, None, 0, C.c_void_p(256), 1 << 20, None)
    assert rc == -1
    assert lib.ys_nms_batched(None, 1, 1, 1, 1, 0.1, 0.5, 640.0, 0, None, None, None, None, None, None, 0, None) == -1
    assert lib.ys_assign(0, 1, 1, 1, None, None, None, None, None, None, 0.5, 4.0, 1e-9, None, None, None, None, None,
                         None, None, 0, None) == -1
    assert lib.ys_pairwise_iou(None, 4, None, 4, 0, None, None) == -1
    assert lib.ys_bbox_iou(None, None, 4, 0, 7, 1e-7, None, None, None, None) == -1


def test_product_refuses_cpu_tensors():
    import torch
    from yolo_scratch_b200.train_yolo11_cuda import nms_simple
    from yolo_scratch_b200.utils import calculate_iou_batch
    with pytest.raises(_lib.YsError):
        nms_simple(torch.zeros(3, 4), torch.zeros(3), 0.5)
    with pytest.raises(_lib.YsError):
        calculate_iou_batch(torch.zeros(3, 4), torch.zeros(3, 4))
