"""Rank-0 ``state_dict`` checkpoints.

The reference never saves anything; the convention kept here is the PyTorch one its users
expect: global rank 0 writes ``torch.save(ddp.module.state_dict(), path)`` -- keys are the
UNWRAPPED module's (``conv1.weight`` ... ``fc1.bias`` for launch_dist.py's ConvNet), so the
file loads into the reference's own model with stock ``torch.load`` -- then everyone meets at
a barrier.  ``load_checkpoint`` reads on every rank (map_location = the module's device).
Optimizer state uses ``torch.optim.SGD.state_dict()`` layout."""
from __future__ import annotations

import os

import torch

from ..parallel import distributed as dist


def _unwrap(model):
    return model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model


def save_checkpoint(model, path, optimizer=None, epoch=None, extra=None, group=None):
    """Write on global rank 0 only, then barrier.  Returns True on the writing rank."""
    wrote = False
    if not dist.is_initialized() or dist.get_rank() == 0:
        state = {k: v.detach().to("cpu", copy=True).contiguous() for k, v in _unwrap(model).state_dict().items()}
        if optimizer is None and epoch is None and extra is None:
            payload = state  # plain state_dict: loadable with model.load_state_dict(torch.load(path))
        else:
            payload = {"model": state}
            if optimizer is not None:
                payload["optimizer"] = optimizer.state_dict()
            if epoch is not None:
                payload["epoch"] = int(epoch)
            if extra is not None:
                payload["extra"] = extra
        tmp = path + ".tmp"
        torch.save(payload, tmp)
        os.replace(tmp, path)
        wrote = True
    if dist.is_initialized():
        dist.barrier(group)
    return wrote


def load_checkpoint(model, path, optimizer=None, strict=True, map_location=None):
    """Load on every rank.  Returns the payload's extra fields (epoch, extra) as a dict."""
    target = _unwrap(model)
    if map_location is None:
        try:
            map_location = next(target.parameters()).device
        except StopIteration:
            map_location = "cpu"
    payload = torch.load(path, map_location=map_location, weights_only=False)
    if isinstance(payload, dict) and "model" in payload and isinstance(payload["model"], dict):
        state = payload["model"]
        info = {k: payload[k] for k in ("epoch", "extra") if k in payload}
        if optimizer is not None and "optimizer" in payload:
            optimizer.load_state_dict(payload["optimizer"])
    else:
        state, info = payload, {}
    state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state.items()}
    with torch.no_grad():
        own = target.state_dict()
        missing = [k for k in own if k not in state]
        unexpected = [k for k in state if k not in own]
        if strict and (missing or unexpected):
            raise RuntimeError("checkpoint does not match the model: missing %s, unexpected %s" % (missing, unexpected))
        for k, v in state.items():
            if k in own:
                own[k].copy_(v)  # in place: parameters may be views of a flat arena
    return info
