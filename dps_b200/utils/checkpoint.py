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
import threading
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
    """Both files are written under names unique to this writer and renamed into place, sidecar FIRST: a reader (or
    a crash between the two renames) then sees either the old pair or new weights with the new sidecar, never new
    weights with an old step / momentum.  The weights file also carries its step in `<path>.meta.json` and the
    sidecar repeats it (``load`` checks that they agree)."""
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    sd = {k: _to_tensor(v) for k, v in state.items()}
    tag = f".tmp.{os.getpid()}.{threading.get_ident()}.{int(global_step)}"
    tmp = path + tag
    torch.save(sd, tmp)
    side = {
        "global_step": int(global_step),
        "shard_versions": list(shard_versions or []),
        "config": config or {},
        "extra": extra or {},
        "torch_rng": torch.get_rng_state(),
        "momentum": {k: _to_tensor(v) for k, v in (momentum or {}).items()},
    }
    side_tmp = path + ".resume" + tag
    torch.save(side, side_tmp)
    meta_tmp = path + ".meta.json" + tag
    with open(meta_tmp, "w") as f:
        json.dump({"global_step": int(global_step), "num_tensors": len(sd),
                   "num_elements": int(sum(t.numel() for t in sd.values()))}, f)
    os.replace(side_tmp, path + ".resume")
    os.replace(tmp, path)
    os.replace(meta_tmp, path + ".meta.json")


def load(path: str, restore_rng: bool = False) -> Tuple[Dict[str, torch.Tensor], dict]:
    sd = torch.load(path, map_location="cpu", weights_only=True)
    side: dict = {}
    if os.path.exists(path + ".resume"):
        side = torch.load(path + ".resume", map_location="cpu", weights_only=True)
        try:                                   # a crash between the renames leaves a sidecar newer than the weights
            with open(path + ".meta.json") as f:
                meta_step = int(json.load(f).get("global_step", -1))
            if meta_step >= 0 and meta_step != int(side.get("global_step", meta_step)):
                side["step_mismatch"] = {"weights": meta_step, "sidecar": int(side.get("global_step", 0))}
        except (OSError, ValueError):
            pass
        if restore_rng and "torch_rng" in side:
            torch.set_rng_state(side["torch_rng"])
    return sd, side


def to_numpy_state(sd: Mapping[str, torch.Tensor]) -> Dict[str, np.ndarray]:
    return {k: v.detach().cpu().numpy() for k, v in sd.items()}
