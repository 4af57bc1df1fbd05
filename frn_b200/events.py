"""Event voxelisation on the GPU: drop-in for DSECDet.preprocess_events
(retinanet/data/dsec_data.py:347-387) and normalize_event_data (:78-93).

    preprocess_events(events, device="cuda") -> {'img': Tensor(5, 480, 640)}      # same return value
    normalize_event_data(img_event, method='tanh')

`events` is the reference's dict of equal-length 1-D arrays 'x', 'y', 't', 'p' (NumPy arrays as read from the
DSEC h5 file, or torch tensors; CUDA tensors are used in place).  No CPU fallback."""
from __future__ import annotations


import numpy as np
import torch

from . import _lib
from .functional import _bytes, _ptr, _stream

_NORMALIZE = {None: 0, "none": 0, "tanh": 1, "clip": 2}


def _to_device(a, device, dtype):
    t = torch.from_numpy(np.ascontiguousarray(a)) if isinstance(a, np.ndarray) else a
    if dtype == "xy":
        if t.dtype == torch.uint16:
            pass                                           # DSEC h5 layout, used as is
        elif t.dtype in (torch.int16, torch.uint8):
            t = t.to(torch.int32)
        elif t.dtype != torch.int32:
            t = t.to(torch.int64).clamp(-2 ** 31, 2 ** 31 - 1).to(torch.int32)
    elif dtype == "t":
        t = t if t.dtype == torch.int64 else t.to(torch.int64)
    elif dtype == "p":
        t = t if t.dtype == torch.uint8 else (t > 0).to(torch.uint8)
    return t.to(device, non_blocking=True).contiguous()


def voxelize_events(events, C_bins=5, H=480, W=640, normalize=None, device="cuda"):
    """(C_bins, H, W) fp32 voxel grid of one sample on `device`; normalize in {None, 'tanh', 'clip'}."""
    if normalize not in _NORMALIZE:
        raise ValueError(f"frn_b200.voxelize_events: normalize must be one of {sorted(map(str, _NORMALIZE))}")
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("frn_b200: CUDA device required (there is no CPU fallback)")
    lib = _lib.load()
    x, y = _to_device(events["x"], device, "xy"), _to_device(events["y"], device, "xy")
    if x.dtype != y.dtype:
        x, y = x.to(torch.int32), y.to(torch.int32)
    t, p = _to_device(events["t"], device, "t"), _to_device(events["p"], device, "p")
    n = int(x.numel())
    assert y.numel() == n and t.numel() == n and p.numel() == n
    voxel = torch.empty((C_bins, H, W), dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        ws_bytes = lib.frn_voxelize_workspace_bytes()
        ws = _bytes(ws_bytes, device)
        _lib.check(lib.frn_voxelize_events(_stream(), n, x.element_size(), _ptr(x), _ptr(y), _ptr(t), _ptr(p),
                                           C_bins, H, W, _NORMALIZE[normalize], _ptr(voxel), _ptr(ws), ws_bytes),
                   "voxelize_events")
    return voxel


def preprocess_events(events, device="cuda"):
    """dsec_data.py:347-387: H, W = 480, 640 and C = 5 are fixed there."""
    return {"img": voxelize_events(events, 5, 480, 640, None, device)}


def normalize_event_data(img_event, method="tanh"):
    """dsec_data.py:78-93 for a grid already on the GPU ('tanh' and 'clip'; 'minmax' is not on the reference's
    default path and is refused)."""
    if method not in ("tanh", "clip"):
        raise NotImplementedError("frn_b200.normalize_event_data: method 'tanh' or 'clip'")
    if not img_event.is_cuda:
        raise RuntimeError("frn_b200: CUDA tensors required (there is no CPU fallback)")
    out = img_event.detach().to(torch.float32).contiguous().clone()
    with torch.cuda.device(out.device):
        _lib.check(_lib.load().frn_voxel_normalize(_stream(), out.numel(), 1 if method == "tanh" else 2, _ptr(out)),
                   "voxel_normalize")
    return out
