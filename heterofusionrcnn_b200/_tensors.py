"""Argument validation shared by the op modules: torch CUDA tensors used as device buffers."""
import torch

from ._lib import InvalidArgumentError


def require(cond, msg):
    if not cond:
        raise InvalidArgumentError(msg)


def dev_tensor(t, dtype, name):
    """Return `t` as a contiguous CUDA tensor of `dtype` (no copies across devices, no CPU path)."""
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor on a CUDA device, got %s" % (name, type(t).__name__))
    if not t.is_cuda:
        raise RuntimeError(
            "%s is on %s: the hfops operators run on CUDA tensors only (there is no CPU fallback)" % (name, t.device)
        )
    if t.dtype != dtype:
        raise TypeError("%s must have dtype %s, got %s" % (name, dtype, t.dtype))
    return t if t.is_contiguous() else t.contiguous()


def stream_ptr(device):
    return torch.cuda.current_stream(device).cuda_stream


def scatter_workspace(lib, scenes, dst_rows, entries, device):
    """Scratch of the deterministic scatter-add gradients (csrc/backward.cu)."""
    nbytes = int(lib.hf_scatter_workspace_bytes(int(scenes), int(dst_rows), int(entries)))
    require(nbytes > 0, "gradient: more than 2^31 destination rows or entries")
    return torch.empty((nbytes,), dtype=torch.uint8, device=device), nbytes
