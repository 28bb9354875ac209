"""Tensor plumbing shared by the mirror modules (no arithmetic on grid-sized data)."""
from __future__ import annotations

import numpy as np
import torch

F64 = torch.float64


def device():
    if not torch.cuda.is_available():
        from .._lib import VdpError
        raise VdpError("vdp_b200 needs a CUDA device (there is no CPU path)")
    return torch.device("cuda", torch.cuda.current_device())


def as_dev(x, dtype=F64):
    """numpy / list / tensor -> contiguous CUDA tensor of ``dtype``."""
    if isinstance(x, torch.Tensor):
        t = x.to(device=device(), dtype=dtype)
    else:
        t = torch.as_tensor(np.asarray(x), dtype=dtype).to(device())
    return t.contiguous()


def as_np(x):
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    return np.asarray(x, dtype=np.float64)


def batched(t, core_dims):
    """Give ``t`` exactly one leading batch axis.  Returns (tensor, had_batch)."""
    if t.dim() == core_dims:
        return t.unsqueeze(0).contiguous(), False
    if t.dim() == core_dims + 1:
        return t.contiguous(), True
    lead = t.shape[:t.dim() - core_dims]
    return t.reshape((-1,) + tuple(t.shape[t.dim() - core_dims:])).contiguous(), tuple(lead)


def unbatched(t, had_batch):
    if had_batch is False:
        return t[0]
    if had_batch is True:
        return t
    return t.reshape(tuple(had_batch) + tuple(t.shape[1:]))
