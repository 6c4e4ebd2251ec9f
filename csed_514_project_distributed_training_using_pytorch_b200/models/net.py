"""The MNIST network of the reference (src/model.py:4-22), B200-native.

Same layers, same parameter names and shapes, same default initialisation (the
parameter holders are stock `nn.Conv2d` / `nn.Linear`, so `torch.manual_seed(1); Net()`
gives the reference's initial weights) and the same `state_dict()` keys
(T10). What differs is execution: on CUDA tensors the forward and backward run the
hand-written sm_100a kernels of csrc/net_kernels.cu through ops.net_ops; on CPU tensors
they run the PyTorch oracle (the reference's own semantics).

`Net.forward(x)` returns log-probabilities [B,10] like the reference. Training at full
speed goes through `ops.fused.FusedNetEngine`, which runs forward + loss + backward +
optimizer as two kernels per step without materialising activations in HBM.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from ..ops import net_ops


class Net(nn.Module):
    """2-conv + 2-FC MNIST CNN returning log-probabilities (src/model.py:4-22)."""

    PARAM_ORDER = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
                   "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")

    def __init__(self, dropout_seed: Optional[int] = None, conv2_impl: str = "simt"):
        super().__init__()
        # conv2 of the layered CUDA path: "simt" (fp32-exact), "tc" (TF32 on tcgen05: implicit
        # GEMM forward, data gradient and weight gradient, csrc/conv2_tc.cu), "auto" (the measured
        # mix: tensor-core backward, tensor-core forward from batch 2048), "tcgen05" (the first
        # tensor-core forward with an in-shared-memory im2col, kept for comparison)
        if conv2_impl not in ("simt", "tc", "auto", "tcgen05"):
            raise ValueError(f"conv2_impl must be simt, tc, auto or tcgen05, not {conv2_impl!r}")
        self.conv2_impl = conv2_impl
        # Parameter holders, registered in the reference's order (src/model.py:9-13) so that
        # the state_dict keys and the values drawn under a given torch seed are the same.
        for name, layer in (("conv1", nn.Conv2d(1, 10, 5)), ("conv2", nn.Conv2d(10, 20, 5)),
                            ("conv2_drop", nn.Dropout2d(p=0.5)),
                            ("fc1", nn.Linear(20 * 4 * 4, 50)), ("fc2", nn.Linear(50, 10))):
            self.add_module(name, layer)
        # Philox key/counter of the CUDA dropout masks (K4, K7): keyed by the torch seed
        # in force at construction unless given, advanced once per training forward.
        self.dropout_seed = int(torch.initial_seed() if dropout_seed is None else dropout_seed) \
            & 0xFFFFFFFFFFFFFFFF
        self.dropout_step = 0

    def param_list(self):
        return [self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias,
                self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias]

    def forward(self, x: torch.Tensor,
                masks: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> torch.Tensor:
        """x: [B,1,28,28] fp32 (normalised). masks=(m2 [B,20], m3 [B,50]) injects dropout
        masks (values 0 or 1/(1-p)); otherwise training mode draws them, eval uses none."""
        if not x.is_cuda:
            return self._forward_cpu(x, masks)
        x = x.reshape(-1, 1, 28, 28)
        if x.dtype != torch.float32:
            x = x.float()
        if masks is not None:
            mode, m2, m3 = 2, masks[0].contiguous().float(), masks[1].contiguous().float()
        elif self.training:
            mode, m2, m3 = 1, None, None
            self.dropout_step += 1
        else:
            mode, m2, m3 = 0, None, None
        seed, step = self.dropout_seed, self.dropout_step
        a1 = net_ops.conv1_pool_relu(x, self.conv1.weight, self.conv1.bias)
        a2 = net_ops.conv2_drop_pool_relu(a1, self.conv2.weight, self.conv2.bias, m2, mode,
                                          self.conv2_drop.p, seed, step, self.conv2_impl)
        h1 = net_ops.fc1_relu_drop(a2.view(-1, 320), self.fc1.weight, self.fc1.bias, m3, mode,
                                   0.5, seed, step)
        return net_ops.fc2_log_softmax(h1, self.fc2.weight, self.fc2.bias)

    def _forward_cpu(self, x, masks):
        """CPU tensors: the functional oracle (models/oracle.py). With no masks given, training
        mode applies torch's own dropout ops (so the random stream is the one a stock torch
        model would consume); eval mode applies none."""
        from .oracle import net_forward
        x = x.reshape(-1, 1, 28, 28)
        if masks is not None:
            return net_forward(self.param_list(), x, masks[0], masks[1])
        if not self.training:
            return net_forward(self.param_list(), x)
        return net_forward(self.param_list(), x, self.conv2_drop,
                           lambda h: torch.nn.functional.dropout(h, 0.5, True))
