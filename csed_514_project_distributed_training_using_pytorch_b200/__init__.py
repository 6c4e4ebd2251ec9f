"""B200-native data-parallel training framework.

Same capabilities as abhishekiitm/CSED_514_Project_Distributed_Training_using_PyTorch
(MNIST `Net`, single-process and multi-worker SGD-momentum training, DistributedSampler
sharding, DDP-style averaged gradients, send/recv demo, torch state_dict checkpoints),
rebuilt for one 8xB200 NVLink/NVSwitch box: hand-written sm_100a kernels behind a C ABI,
a fused all-reduce + SGD reducer over peer memory, CUDA-graph steps.

Layout: models/ (Net, wide MLP, CPU oracle), ops/ (kernel bindings + autograd),
parallel/ (process group, symmetric memory, collectives, reducer, DataParallel,
launcher), utils/ (config, sampler, data, checkpoint, metrics, timing), cli/.
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401
