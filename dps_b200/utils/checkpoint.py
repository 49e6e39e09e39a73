"""Checkpoint / resume.

The reference has no checkpointing (future work, reference: DEPLOYMENT.md:307-309);
its parameters live only in an in-memory dict (src/parameter_server/server.py:96).
The natural format is implied by that dict: the 122 keys of ``ResNet18().state_dict()``.
``save`` writes exactly that (``torch.save({name: tensor})``, contiguous NCHW-shaped
tensors) so the file loads into the reference model with
``load_state_dict(strict=True)``; a sidecar ``<path>.resume`` carries what a restart
needs beyond the weights: global step, per-shard version counters, momentum buffers,
RNG state and the config.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Mapping, Optional, Tuple

import numpy as np
import torch


def _to_tensor(v: Any) -> torch.Tensor:
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().contiguous().clone()
    return torch.from_numpy(np.array(v, copy=True))


def save(path: str, state: Mapping[str, Any], *, global_step: int = 0,
         momentum: Optional[Mapping[str, Any]] = None, shard_versions: Optional[list] = None,
         config: Optional[dict] = None, extra: Optional[dict] = None) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    sd = {k: _to_tensor(v) for k, v in state.items()}
    tmp = path + ".tmp"
    torch.save(sd, tmp)
    os.replace(tmp, path)
    side = {
        "global_step": int(global_step),
        "shard_versions": list(shard_versions or []),
        "config": config or {},
        "extra": extra or {},
        "torch_rng": torch.get_rng_state(),
        "momentum": {k: _to_tensor(v) for k, v in (momentum or {}).items()},
    }
    torch.save(side, path + ".resume.tmp")
    os.replace(path + ".resume.tmp", path + ".resume")
    with open(path + ".meta.json", "w") as f:
        json.dump({"global_step": int(global_step), "num_tensors": len(sd),
                   "num_elements": int(sum(t.numel() for t in sd.values()))}, f)


def load(path: str, restore_rng: bool = False) -> Tuple[Dict[str, torch.Tensor], dict]:
    sd = torch.load(path, map_location="cpu", weights_only=True)
    side: dict = {}
    if os.path.exists(path + ".resume"):
        side = torch.load(path + ".resume", map_location="cpu", weights_only=True)
        if restore_rng and "torch_rng" in side:
            torch.set_rng_state(side["torch_rng"])
    return sd, side


def to_numpy_state(sd: Mapping[str, torch.Tensor]) -> Dict[str, np.ndarray]:
    return {k: v.detach().cpu().numpy() for k, v in sd.items()}
