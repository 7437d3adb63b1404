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

    # dict checkpoints of the AIR generation: one pickled mapping inside the directory
    _DICT_FILE = "dict_checkpoint.pkl"

    @classmethod
    def from_dict(cls, data: Dict[str, Any]) -> "Checkpoint":
        import cloudpickle
        path = tempfile.mkdtemp(prefix="raydp_b200_ckpt_")
        with open(os.path.join(path, cls._DICT_FILE), "wb") as fh:
            cloudpickle.dump(dict(data), fh)
        return cls(path)

    def to_dict(self) -> Dict[str, Any]:
        import pickle
        target = os.path.join(self.path, self._DICT_FILE)
        if os.path.exists(target):
            with open(target, "rb") as fh:
                return pickle.load(fh)
        out: Dict[str, Any] = {}
        for name in sorted(os.listdir(self.path)):          # a directory checkpoint: file name -> bytes
            full = os.path.join(self.path, name)
            if os.path.isfile(full):
                with open(full, "rb") as fh:
                    out[name] = fh.read()
        return out

    def get_metadata(self) -> Dict[str, Any]:
        import json
        meta = os.path.join(self.path, ".metadata.json")
        if not os.path.exists(meta):
            return {}
        with open(meta, "r", encoding="utf-8") as fh:
            return json.load(fh)

    def set_metadata(self, metadata: Dict[str, Any]) -> None:
        import json
        with open(os.path.join(self.path, ".metadata.json"), "w", encoding="utf-8") as fh:
            json.dump(metadata, fh)

    def update_metadata(self, metadata: Dict[str, Any]) -> None:
        merged = self.get_metadata()
        merged.update(metadata)
        self.set_metadata(merged)

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
