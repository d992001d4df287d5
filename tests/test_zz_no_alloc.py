"""ys_bbox_loss_static: the allocation-free variant of ys_bbox_loss (scratch in a file-scope __device__ array whose
address the library takes with cudaGetSymbolAddress).  No counterpart in the reference.
Called through the C ABI and compared with ys_bbox_loss (caller workspace) on the same inputs, two calls in a
row with different sizes so that the second call meets the first one's leftovers in the shared scratch.  Same kernel,
same grid, fixed summation order: the outputs must be identical, not merely close.

Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, A, nc, R, seed, p_fg=0.1):
    g = torch.Generator().manual_seed(seed)
    n = B * A
    side = max(int(A ** 0.5), 1)
    idx = torch.arange(A)
    anchors = torch.stack([(idx % side).float() + 0.5, (idx // side).float() + 0.5], 1)
    ap = anchors.repeat(B, 1)
    lt = torch.rand(n, 2, generator=g) * 6 + 0.2
    rb = torch.rand(n, 2, generator=g) * 6 + 0.2
    pred_bboxes = torch.cat([ap - lt, ap + rb], 1)
    lt2 = torch.rand(n, 2, generator=g) * 7 + 0.1
    rb2 = torch.rand(n, 2, generator=g) * 7 + 0.1
    target_bboxes = torch.cat([ap - lt2, ap + rb2], 1)
    fg = (torch.rand(n, generator=g) < p_fg)
    target_scores = torch.zeros(n, nc)
    cls = torch.randint(0, nc, (n,), generator=g)
    target_scores[torch.arange(n), cls] = torch.rand(n, generator=g) * fg.float()
    pred_dist = torch.randn(n, 4 * R, generator=g) * 2
    tss = target_scores.sum().clamp(min=1.0).reshape(1)
    dev = "cuda"
    return dict(pred_dist=pred_dist.to(dev), pred_bboxes=pred_bboxes.to(dev), anchor_points=anchors.to(dev),
                target_bboxes=target_bboxes.to(dev), target_scores=target_scores.to(dev),
                fg_mask=fg.to(torch.uint8).to(dev), tss=tss.to(dev), n=n, A=A, nc=nc, R=R)


def _run(lib, x, static):
    from yolo_scratch_b200._lib import check, stream_ptr
    out = torch.full((2,), float("nan"), device="cuda")
    g_pb = torch.full((x["n"], 4), float("nan"), device="cuda")
    g_pd = torch.full((x["n"], 4 * x["R"]), float("nan"), device="cuda")
    head = [x[k].data_ptr() for k in ("pred_dist", "pred_bboxes", "anchor_points", "target_bboxes", "target_scores",
                                      "fg_mask", "tss")]
    tail = [x["n"], x["A"], x["nc"], x["R"], out.data_ptr(), g_pb.data_ptr(), g_pd.data_ptr()]
    if static:
        before = torch.cuda.memory_allocated()
        check(lib.ys_bbox_loss_static(*head, *tail, stream_ptr()), "ys_bbox_loss_static")
        assert torch.cuda.memory_allocated() == before
    else:
        ws = torch.zeros(lib.ys_bbox_loss_workspace_bytes(x["n"]), dtype=torch.uint8, device="cuda")
        check(lib.ys_bbox_loss(*head, *tail, ws.data_ptr(), ws.numel(), stream_ptr()), "ys_bbox_loss")
    torch.cuda.synchronize()
    return out, g_pb, g_pd


def test_bbox_loss_static_matches_workspace_entry_twice(cuda_lib):
    lib = cuda_lib
    assert lib.ys_bbox_loss_static_max_rows() == 1048576
    # call 1: 16 images x 8400 anchors (525 blocks of partial sums); call 2: a smaller, different problem that meets
    # call 1's leftovers in the shared scratch; call 3: the first one again (run-to-run identical)
    cases = [_inputs(16, 8400, 80, 16, seed=501), _inputs(3, 2100, 7, 16, seed=502, p_fg=0.4),
             _inputs(16, 8400, 80, 16, seed=501)]
    results = []
    for x in cases:
        ref = _run(lib, x, static=False)
        got = _run(lib, x, static=True)
        assert torch.isfinite(ref[0]).all() and float(ref[0][0]) > 0 and float(ref[0][1]) > 0
        for a, b, what in zip(got, ref, ("loss_out", "grad_pred_bboxes", "grad_pred_dist")):
            assert torch.equal(a, b), what
        results.append(got)
    for a, b in zip(results[0], results[2]):
        assert torch.equal(a, b)


def test_bbox_loss_static_rejects_oversize(cuda_lib):
    from yolo_scratch_b200._lib import stream_ptr
    YS_ERR_UNSUPPORTED = -3          # include/ys_b200.h
    lib = cuda_lib
    x = _inputs(1, 64, 4, 16, seed=503)
    out = torch.zeros(2, device="cuda")
    head = [x[k].data_ptr() for k in ("pred_dist", "pred_bboxes", "anchor_points", "target_bboxes", "target_scores",
                                      "fg_mask", "tss")]
    n_big = lib.ys_bbox_loss_static_max_rows() + 64          # rejected on the host before anything is read
    rc = lib.ys_bbox_loss_static(*head, n_big, 64, 4, 16, out.data_ptr(), 0, 0, stream_ptr())
    assert rc == YS_ERR_UNSUPPORTED
    msg = lib.ys_last_error().decode()
    assert "exceeds" in msg and "1048576" in msg
    # and the entry point still works afterwards
    rc = lib.ys_bbox_loss_static(*head, x["n"], 64, 4, 16, out.data_ptr(), 0, 0, stream_ptr())
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
