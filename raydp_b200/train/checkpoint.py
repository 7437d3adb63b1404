"""Directory checkpoints.

Parity: ``Checkpoint.from_directory`` / ``to_directory`` and
``TorchCheckpoint(dir).get_model(model)`` as used by reference ``torch/estimator.py:263-270,
392-396``.  The layout is part of the compatibility surface: a directory holding ``model.pt`` =
``torch.save(state_dict)`` of the *unwrapped* module, written by world rank 0 only.
"""
from __future__ import annotations

import contextlib
import os
import shutil
import tempfile
import uuid
from typing import Any, Dict, Optional

MODEL_KEY = "model.pt"


class Checkpoint:
    def __init__(self, path: str):
        self.path = os.path.abspath(path)

    @classmethod
    def from_directory(cls, path: str) -> "Checkpoint":
        return cls(path)

    def to_directory(self, path: Optional[str] = None) -> str:
        if path is None:
            path = os.path.join(tempfile.gettempdir(), f"raydp_b200_ckpt_{uuid.uuid4().hex[:12]}")
        if os.path.abspath(path) != self.path:
            shutil.copytree(self.path, path, dirs_exist_ok=True)
        return path

    @contextlib.contextmanager
    def as_directory(self):
        yield self.path

    def _persist(self, storage_dir: str) -> "Checkpoint":
        os.makedirs(storage_dir, exist_ok=True)
        dest = os.path.join(storage_dir, f"checkpoint_{uuid.uuid4().hex[:8]}")
        shutil.copytree(self.path, dest)
        return Checkpoint(dest)

    def __repr__(self):
        return f"Checkpoint(path={self.path})"


class TorchCheckpoint(Checkpoint):
    """Loads ``model.pt`` (a state_dict, or a whole pickled module) into a model."""

    def get_model(self, model=None):
        import torch
        payload = torch.load(os.path.join(self.path, MODEL_KEY), map_location="cpu",
                             weights_only=False)
        if isinstance(payload, torch.nn.Module):
            return payload
        if model is None:
            raise ValueError("the checkpoint holds a state_dict; pass the module to load it into")
        model.load_state_dict(payload)
        return model

    @classmethod
    def from_model(cls, model, path: Optional[str] = None) -> "TorchCheckpoint":
        import torch
        path = path or tempfile.mkdtemp(prefix="raydp_b200_ckpt_")
        os.makedirs(path, exist_ok=True)
        torch.save(model.state_dict(), os.path.join(path, MODEL_KEY))
        return cls(path)

    @classmethod
    def from_state_dict(cls, state: Dict[str, Any], path: Optional[str] = None) -> "TorchCheckpoint":
        import torch
        path = path or tempfile.mkdtemp(prefix="raydp_b200_ckpt_")
        os.makedirs(path, exist_ok=True)
        torch.save(state, os.path.join(path, MODEL_KEY))
        return cls(path)
