"""ORACLE (test infrastructure): ctypes wrapper of oracle/scalar_spec.c."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_c.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "scalar_spec.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_iou_assign.restype = C.c_float
        _lib.orc_iou_nms.restype = C.c_float
    return _lib


def _p(a, ty=C.c_float):
    return a.ctypes.data_as(C.POINTER(ty))


def _f32(x):
    return np.ascontiguousarray(x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x, dtype=np.float32)


def assign(pd_scores, pd_bboxes, anc_points, gt_labels, gt_bboxes, mask_gt, nc, alpha=0.5, beta=4.0, eps=1e-9):
    ps, pb, an, gl, gb, mg = map(_f32, (pd_scores, pd_bboxes, anc_points, gt_labels, gt_bboxes, mask_gt))
    B, A = ps.shape[:2]
    G = gb.shape[1]
    tgi = np.zeros((B, A), np.int32)
    fg = np.zeros((B, A), np.uint8)
    norm = np.zeros((B, A), np.float32)
    lab = np.zeros((B, A), np.float32)
    tb = np.zeros((B, A, 4), np.float32)
    rc = lib().orc_assign(B, A, G, int(nc), _p(ps), _p(pb), _p(an), _p(gl), _p(gb), _p(mg), C.c_float(alpha),
                          C.c_float(beta), C.c_float(eps), _p(tgi, C.c_int32), _p(fg, C.c_uint8), _p(norm), _p(lab),
                          _p(tb))
    assert rc == 0, rc
    return {"target_gt_idx": torch.from_numpy(tgi), "fg_mask": torch.from_numpy(fg).bool(),
            "norm": torch.from_numpy(norm), "target_labels": torch.from_numpy(lab),
            "target_bboxes": torch.from_numpy(tb)}


def nms(boxes, scores, thr, labels=None):
    b, s = _f32(boxes), _f32(scores)
    n = s.shape[0]
    keep = np.zeros(max(n, 1), np.int32)
    lp = None
    if labels is not None:
        l = np.ascontiguousarray(labels.detach().cpu().numpy() if isinstance(labels, torch.Tensor) else labels,
                                 dtype=np.int32)
        lp = _p(l, C.c_int32)
    k = lib().orc_nms(n, _p(b), _p(s), lp, C.c_float(thr), _p(keep, C.c_int32))
    return keep[:k].tolist()


def filter_candidates(pred_img, conf):
    p = _f32(pred_img)
    A, w = p.shape
    nc = w - 4
    xyxy = np.zeros((A, 4), np.float32)
    sc = np.zeros(A, np.float32)
    cl = np.zeros(A, np.int32)
    src = np.zeros(A, np.int32)
    n = lib().orc_filter_candidates(A, nc, _p(p), C.c_float(conf), _p(xyxy), _p(sc), _p(cl, C.c_int32),
                                    _p(src, C.c_int32))
    return xyxy[:n], sc[:n], cl[:n], src[:n]


def iou_matrix(b1, b2):
    a, b = _f32(b1), _f32(b2)
    out = np.zeros((a.shape[0], b.shape[0]), np.float32)
    lib().orc_iou_matrix(a.shape[0], b.shape[0], _p(a), _p(b), _p(out))
    return torch.from_numpy(out)
