"""Drive the UNMODIFIED reference CUDA kernels (oracle/_ref/libref_cuda.so, built by
oracle/Makefile from /root/reference/*/*_g.cu for sm_100a) on the GPU box.  Test-only:
a second, GPU-side oracle — the reference itself, run here."""
import ctypes
import os
import subprocess

import torch

SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_cuda.so")
_i, _f, _p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p
_lib = None
_syms = {}


def available():
    return os.path.exists(SO)


def _fn(name):
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(SO)
        out = subprocess.run(["nm", "-D", "--defined-only", SO], check=True, stdout=subprocess.PIPE, text=True).stdout
        for line in out.splitlines():
            sym = line.split()[-1]
            if sym.startswith("_Z") and "__device_stub__" not in sym:
                digits = ""
                rest = sym[2:]
                while rest and rest[0].isdigit():
                    digits += rest[0]
                    rest = rest[1:]
                if digits:
                    _syms.setdefault(rest[: int(digits)], sym)
    f = getattr(_lib, _syms[name])
    f.restype = None
    return f


def _d(t):
    return _p(t.data_ptr())


def fps(npoint, inp):
    b, n, _ = inp.shape
    temp = torch.empty((32, n), dtype=torch.float32, device=inp.device)
    out = torch.zeros((b, npoint), dtype=torch.int32, device=inp.device)
    torch.cuda.synchronize()
    _fn("farthestpointsamplingLauncher")(_i(b), _i(n), _i(npoint), _d(inp), _d(temp), _d(out))
    torch.cuda.synchronize()
    return out


def gather_point(inp, idx):
    b, n, _ = inp.shape
    m = idx.shape[1]
    out = torch.zeros((b, m, 3), dtype=torch.float32, device=inp.device)
    torch.cuda.synchronize()
    _fn("gatherpointLauncher")(_i(b), _i(n), _i(m), _d(inp), _d(idx), _d(out))
    torch.cuda.synchronize()
    return out


def query_ball_point(radius, nsample, xyz1, xyz2):
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    idx = torch.zeros((b, m, nsample), dtype=torch.int32, device=xyz1.device)
    cnt = torch.zeros((b, m), dtype=torch.int32, device=xyz1.device)
    torch.cuda.synchronize()
    _fn("queryBallPointLauncher")(_i(b), _i(n), _i(m), _f(radius), _i(nsample), _d(xyz1), _d(xyz2), _d(idx), _d(cnt))
    torch.cuda.synchronize()
    return idx, cnt


def group_point(points, idx):
    b, n, c = points.shape
    m, ns = idx.shape[1], idx.shape[2]
    out = torch.zeros((b, m, ns, c), dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    _fn("groupPointLauncher")(_i(b), _i(n), _i(c), _i(m), _i(ns), _d(points), _d(idx), _d(out))
    torch.cuda.synchronize()
    return out


def three_nn(xyz1, xyz2):
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    dist = torch.zeros((b, n, 3), dtype=torch.float32, device=xyz1.device)
    idx = torch.zeros((b, n, 3), dtype=torch.int32, device=xyz1.device)
    torch.cuda.synchronize()
    _fn("three_nn_gpu")(_i(b), _i(n), _i(m), _d(xyz1), _d(xyz2), _d(dist), _d(idx))
    torch.cuda.synchronize()
    return dist, idx


def three_interpolate_channel_first(points, idx, weight):
    b, c, m = points.shape
    n = idx.shape[1]
    out = torch.zeros((b, c, n), dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    _fn("three_interpolate_gpu")(_i(b), _i(c), _i(m), _i(n), _d(points), _d(idx), _d(weight), _d(out))
    torch.cuda.synchronize()
    return out


def pc_crop_and_sample(pts, fts, intensities, mask, boxes, box_ind, resize):
    B, P, C = fts.shape
    IC = intensities.shape[2]
    N, R = boxes.shape[0], int(resize)
    dev = pts.device
    crop_pts = torch.zeros((N, R, 3), dtype=torch.float32, device=dev)
    crop_fts = torch.zeros((N, R, C), dtype=torch.float32, device=dev)
    crop_int = torch.zeros((N, R, IC), dtype=torch.float32, device=dev)
    crop_mask = torch.zeros((N, R), dtype=torch.bool, device=dev)
    crop_ind = torch.zeros((N, R), dtype=torch.int32, device=dev)
    non_empty = torch.ones((N,), dtype=torch.bool, device=dev)  # wrapper memsets, tf_cropping.cpp:171-176
    torch.cuda.synchronize()
    _fn("pccropandsample_gpu")(_d(pts), _d(fts), _d(intensities), _d(mask), _d(boxes), _d(box_ind), _i(N), _i(B), _i(P),
                               _i(R), _i(C), _i(IC), _d(crop_pts), _d(crop_fts), _d(crop_int), _d(crop_mask),
                               _d(crop_ind), _d(non_empty))
    torch.cuda.synchronize()
    return crop_pts, crop_fts, crop_int, crop_mask, crop_ind, non_empty


def compute_bev_iou(a, g):
    n, m = a.shape[0], g.shape[0]
    overlap = torch.zeros((n, m), dtype=torch.float32, device=a.device)
    iou = torch.zeros((n, m), dtype=torch.float32, device=a.device)
    torch.cuda.synchronize()
    _fn("compute_bev_iou_gpu")(_i(n), _d(a), _i(m), _d(g), _d(overlap), _d(iou))
    torch.cuda.synchronize()
    return overlap, iou


def nms_mask(boxes, thresh):
    n = boxes.shape[0]
    mask = torch.zeros((n, (n + 63) // 64), dtype=torch.int64, device=boxes.device)
    torch.cuda.synchronize()
    _fn("oriented_nms_gpu")(_d(boxes), _d(mask), _i(n), _f(thresh))
    torch.cuda.synchronize()
    return mask


# ---- gradients: the wrappers memset the output first (tf_sampling.cpp:174, tf_interpolate.cpp:178) ----
def gather_point_grad(inp, idx, out_g):
    b, n, _ = inp.shape
    m = idx.shape[1]
    inp_g = torch.zeros((b, n, 3), dtype=torch.float32, device=inp.device)
    torch.cuda.synchronize()
    _fn("scatteraddpointLauncher")(_i(b), _i(n), _i(m), _d(out_g), _d(idx), _d(inp_g))
    torch.cuda.synchronize()
    return inp_g


def group_point_grad(points, idx, grad_out):
    b, n, c = points.shape
    m, ns = idx.shape[1], idx.shape[2]
    g = torch.zeros((b, n, c), dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    _fn("groupPointGradLauncher")(_i(b), _i(n), _i(c), _i(m), _i(ns), _d(grad_out), _d(idx), _d(g))
    torch.cuda.synchronize()
    return g


def three_interpolate_grad_channel_first(points, idx, weight, grad_out):
    b, c, m = points.shape
    n = idx.shape[1]
    g = torch.zeros((b, c, m), dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    _fn("three_interpolate_grad_gpu")(_i(b), _i(c), _i(n), _i(m), _d(grad_out), _d(idx), _d(weight), _d(g))
    torch.cuda.synchronize()
    return g


def pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, grad_crop_fts):
    B, P, C = fts.shape
    N, R = crop_ind.shape
    g = torch.zeros((B, P, C), dtype=torch.float32, device=fts.device)
    torch.cuda.synchronize()
    _fn("pccropandsamplegradfts_gpu")(_d(box_ind), _d(crop_ind), _d(grad_crop_fts), _i(N), _i(P), _i(R), _i(C), _d(g))
    torch.cuda.synchronize()
    return g


def select_top_k(k, dist):
    b, m, n = dist.shape
    outi = torch.zeros((b, m, n), dtype=torch.int32, device=dist.device)
    out = torch.zeros((b, m, n), dtype=torch.float32, device=dist.device)
    torch.cuda.synchronize()
    _fn("selectionSortLauncher")(_i(b), _i(n), _i(m), _i(int(k)), _d(dist), _d(outi), _d(out))
    torch.cuda.synchronize()
    return outi, out


def prob_sample(inp, inpr):
    b, n = inp.shape
    m = inpr.shape[1]
    temp = torch.zeros((b, n), dtype=torch.float32, device=inp.device)
    out = torch.zeros((b, m), dtype=torch.int32, device=inp.device)
    torch.cuda.synchronize()
    _fn("probsampleLauncher")(_i(b), _i(n), _i(m), _d(inp), _d(inpr), _d(temp), _d(out))
    torch.cuda.synchronize()
    return out, temp
