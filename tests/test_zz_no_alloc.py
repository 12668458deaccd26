"""GPU: hf_box3d_iou_static, the one entry point of libhfops.so that takes no workspace and allocates nothing.

Who needs it: a host that cannot get device memory at the moment it calls the op.  Examples are a plugin inside another
framework's executor that forbids cudaMalloc on its worker threads (an allocation synchronises the device and stalls
the framework's other streams), and a real-time loop whose memory budget is frozen after start-up.  hf_box3d_iou asks
such a caller for a staging buffer it cannot provide; hf_box3d_iou_static stages the two BEV box lists in a fixed
320 KiB array that lives inside the library's device image instead (include/hfops.h explains the limit, the per-device
set-up on the first call, and why overlapping calls race).

What is checked: called straight through the C-ABI (ctypes, raw pointers), the entry gives the answers of the existing
hf_box3d_iou on the same boxes, within the IoU tolerance of tests/test_gpu_glue.py (rtol 1e-5, atol 1e-6), on two
calls in a row.  The second call is smaller than the first, so the array still holds the first call's boxes behind
the part the second call rewrites: stale contents must not leak into the result.  Sizes past the limit are refused.

The file sorts last on purpose: the driver runs `pytest -x`, so a failure here cannot hide any other test.  It passes
alone in a fresh process (the first call then does the per-device address look-up) and after the rest of the suite.
"""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import _lib, synthetic
from heterofusionrcnn_b200._tensors import stream_ptr
from heterofusionrcnn_b200.compute_iou import box3d_iou_tf

pytestmark = pytest.mark.gpu

MAX_BOXES = 8192  # HF_BOX3D_IOU_STATIC_MAX_BOXES in include/hfops.h


def boxes3d(seed, num, clusters=25):
    """(num,7) boxes scattered around a few cluster centres, so that many pairs overlap."""
    rng = np.random.default_rng(seed)
    xyz, _ = synthetic.make_scan(seed, 4096)
    base = synthetic.make_boxes3d(seed, xyz, clusters)
    b = base[rng.integers(0, clusters, num)].copy()
    b[:, [0, 2]] += rng.normal(0, 0.3, (num, 2)).astype(np.float32)
    b[:, 6] += rng.normal(0, 0.1, num).astype(np.float32)
    return np.ascontiguousarray(b.astype(np.float32))


def call_static(lib, a, b):
    """The C-ABI call a non-allocating host would make: it owns inputs and outputs, and passes nothing else."""
    n, m = a.shape[0], b.shape[0]
    # poisoned outputs: every element must be written by the call
    iou_3d = torch.full((n, m), float("nan"), dtype=torch.float32, device=a.device)
    iou_2d = torch.full((n, m), float("nan"), dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        rc = lib.hf_box3d_iou_static(n, a.data_ptr(), m, b.data_ptr(), iou_3d.data_ptr(), iou_2d.data_ptr(),
                                     stream_ptr(a.device))
    _lib.check(rc, "box3d_iou_static")
    return iou_3d, iou_2d


def test_static_entry_matches_workspace_entry_on_consecutive_calls(cuda):
    lib = _lib.load()
    launches_before = _lib.launch_count()
    # first call large, second small and different: the array then holds leftovers of the first call
    for seed, n, m in [(1, 1500, 700), (5, 300, 40)]:
        a = torch.from_numpy(boxes3d(seed, n)).to(cuda)
        b = torch.from_numpy(boxes3d(seed + 1, m)).to(cuda)
        b[:5] = a[:5]  # identical boxes: IoU 1
        got3, got2 = call_static(lib, a, b)
        want3, want2 = box3d_iou_tf(a, b)  # hf_box3d_iou with a torch-allocated workspace
        torch.cuda.synchronize()
        assert torch.isfinite(got3).all() and torch.isfinite(got2).all()
        assert np.allclose(got2.cpu().numpy(), want2.cpu().numpy(), rtol=1e-5, atol=1e-6), (n, m)
        assert np.allclose(got3.cpu().numpy(), want3.cpu().numpy(), rtol=1e-5, atol=1e-6), (n, m)
        assert np.allclose(np.diag(got3.cpu().numpy()[:5, :5]), 1.0, atol=1e-4) and int((got3 > 0).sum()) > 100
    assert _lib.launch_count() - launches_before == 2 * (4 + 4)  # four launches per call, either entry


def test_static_entry_accepts_the_limit_and_refuses_more(cuda):
    lib = _lib.load()
    a = torch.from_numpy(boxes3d(3, MAX_BOXES)).to(cuda)
    b = torch.from_numpy(boxes3d(4, 16)).to(cuda)
    got3, got2 = call_static(lib, a, b)  # exactly the limit on one side
    want3, want2 = box3d_iou_tf(a, b)
    assert np.allclose(got3.cpu().numpy(), want3.cpu().numpy(), rtol=1e-5, atol=1e-6)
    assert np.allclose(got2.cpu().numpy(), want2.cpu().numpy(), rtol=1e-5, atol=1e-6)

    out = torch.zeros((4,), dtype=torch.float32, device=cuda)
    launches = _lib.launch_count()
    for n, m in [(MAX_BOXES + 1, 1), (1, MAX_BOXES + 1)]:  # refused before any pointer is touched
        rc = lib.hf_box3d_iou_static(n, a.data_ptr(), m, b.data_ptr(), out.data_ptr(), out.data_ptr(), stream_ptr(cuda))
        assert rc == _lib.HF_ERR_UNSUPPORTED
        assert b"at most 8192 boxes per side" in lib.hf_last_error()
    assert _lib.launch_count() == launches  # nothing was launched
    assert lib.hf_box3d_iou_static(0, None, 5, None, None, None, None) == _lib.HF_OK  # empty side: nothing to do
