"""torch-compatible checkpoints (SURVEY T10 and §5 "Checkpoint / resume").

Reference behaviour: `torch.save(network.state_dict(), 'results/model.pth')` and
`torch.save(optimizer.state_dict(), 'results/optimizer.pth')` every log interval
(src/train.py:84-85); rank 0 saves the unwrapped module's state_dict to 'model.pt'
(src/train_dist.py:163-164). There is no load path in the reference; `load_checkpoint`
and `resume` are the natural completion. Files written here load straight into the
reference's `Net().load_state_dict(...)` and into `torch.optim.SGD.load_state_dict`.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch


def _cpu_state(sd):
    return {k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in sd.items()}


def model_state_dict(module: torch.nn.Module) -> Dict[str, torch.Tensor]:
    """state_dict of the unwrapped module (no `module.` prefix), tensors copied to CPU."""
    inner = getattr(module, "module", module)
    return _cpu_state(inner.state_dict())


def sgd_state_dict(momentum_buffers, lr: float, momentum: float, weight_decay: float = 0.0,
                   nesterov: bool = False) -> dict:
    """A dict shaped like torch.optim.SGD.state_dict() for one param group."""
    n = len(momentum_buffers)
    state = {i: {"momentum_buffer": b.detach().cpu().clone()} for i, b in
             enumerate(momentum_buffers) if b is not None}
    group = {"lr": lr, "momentum": momentum, "dampening": 0, "weight_decay": weight_decay,
             "nesterov": nesterov, "maximize": False, "foreach": None, "differentiable": False,
             "fused": None, "params": list(range(n))}
    return {"state": state, "param_groups": [group]}


def save_checkpoint(path: str, module: torch.nn.Module, rank: int = 0) -> Optional[str]:
    """Rank-0-only model save (src/train_dist.py:163-164)."""
    if rank != 0:
        return None
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)  # the reference crashes if results/ is missing
    tmp = path + ".tmp"
    torch.save(model_state_dict(module), tmp)
    os.replace(tmp, path)
    return path


def save_optimizer(path: str, optim_state: dict, rank: int = 0) -> Optional[str]:
    if rank != 0:
        return None
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    tmp = path + ".tmp"
    torch.save(optim_state, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path: str, module: torch.nn.Module, strict: bool = True):
    """Load a model state_dict in place: parameters that are views of a flat buffer keep
    their storage (copy_, not rebind), so engines built on the module see the new values."""
    sd = torch.load(path, map_location="cpu", weights_only=True)
    inner = getattr(module, "module", module)
    own = inner.state_dict()
    missing = [k for k in own if k not in sd]
    unexpected = [k for k in sd if k not in own]
    if strict and (missing or unexpected):
        raise KeyError(f"checkpoint mismatch: missing {missing}, unexpected {unexpected}")
    with torch.no_grad():
        for k, v in sd.items():
            if k in own:
                own[k].copy_(v)
    return sd


def load_momentum(path: str, momentum_views) -> dict:
    """Restore momentum buffers from a torch-SGD-shaped optimizer checkpoint."""
    sd = torch.load(path, map_location="cpu", weights_only=False)
    with torch.no_grad():
        for i, view in enumerate(momentum_views):
            st = sd["state"].get(i)
            if st is not None and st.get("momentum_buffer") is not None:
                view.copy_(st["momentum_buffer"])
    return sd
