"""NCCL baseline group for the comparison scripts. The product's process group is gloo-only
(bootstrap); a script that wants to time torch/NCCL creates its own group here."""
import torch.distributed as dist

_G = None


def nccl_group():
    global _G
    if _G is None and dist.is_initialized() and dist.get_world_size() > 1:
        _G = dist.new_group(backend="nccl")
    return _G
