"""dps_b200 -- a Blackwell-native (B200 / sm_100a) parameter-server training framework.

Same capabilities as Jjjing2023/Distributed-Parameter-Server-for-ML-Training (ParameterServer /
Worker roles, the Register / PushGradrients / FetchParameters / JobFinished contract, sync and
async modes, fp16 gradient compression, disjoint data partitions, JSON metrics, state_dict
checkpoints), re-designed for one 8xB200 NVSwitch box: the store is sharded across the GPUs and
the data path is hand-written sm_100a kernels.

Importing this package never touches CUDA or the native library; device code loads lazily.
"""
from .config import Config  # noqa: F401

__version__ = "0.1.0"

_LAZY = {
    "ParameterServer": ("dps_b200.api", "ParameterServer"),
    "Worker": ("dps_b200.api", "Worker"),
    "Job": ("dps_b200.api", "Job"),
    "ParameterServerServicer": ("dps_b200.rpc.servicer", "ParameterServerServicer"),
    "DistributedWorker": ("dps_b200.rpc.client", "DistributedWorker"),
    "serve": ("dps_b200.rpc.servicer", "serve"),
    "ResNet18": ("dps_b200.models.resnet", "ResNet18"),
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        mod, attr = _LAZY[name]
        return getattr(importlib.import_module(mod), attr)
    raise AttributeError(name)
