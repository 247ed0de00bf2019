"""Checkpoint directory management with tf.train.Saver's conventions.

``log_root/checkpoint`` is the text CheckpointState proto the reference's evaluators poll through
``tf.train.get_checkpoint_state`` (resnet_cifar_eval.py:101-109):

    model_checkpoint_path: "model.ckpt-1234"
    all_model_checkpoint_paths: "model.ckpt-1000"
    all_model_checkpoint_paths: "model.ckpt-1234"

Checkpoints are named ``model.ckpt-<global_step>``, the last ``keep_max`` (5) are kept, and the
state file is replaced atomically after the bundle files exist.
"""
from __future__ import annotations

import glob
import os
import re
import time
from typing import Dict, List, Optional

import numpy as np

from .tensorbundle import BundleReader, write_bundle

STATE_FILE = "checkpoint"


class CheckpointState:
    def __init__(self, model_checkpoint_path: str = "", all_paths: Optional[List[str]] = None):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = list(all_paths or [])

    def to_text(self) -> str:
        lines = [f'model_checkpoint_path: "{self.model_checkpoint_path}"']
        lines += [f'all_model_checkpoint_paths: "{p}"' for p in self.all_model_checkpoint_paths]
        return "\n".join(lines) + "\n"

    @staticmethod
    def from_text(text: str) -> "CheckpointState":
        st = CheckpointState()
        for m in re.finditer(r'^\s*(\w+)\s*:\s*"((?:[^"\\]|\\.)*)"', text, re.M):
            key, val = m.group(1), m.group(2)
            if key == "model_checkpoint_path":
                st.model_checkpoint_path = val
            elif key == "all_model_checkpoint_paths":
                st.all_model_checkpoint_paths.append(val)
        return st


def get_checkpoint_state(log_root: str) -> Optional[CheckpointState]:
    path = os.path.join(log_root, STATE_FILE)
    try:
        with open(path, "r") as f:
            st = CheckpointState.from_text(f.read())
    except OSError:
        return None
    return st if st.model_checkpoint_path else None


def _abs(log_root: str, p: str) -> str:
    return p if os.path.isabs(p) else os.path.join(log_root, p)


def latest_checkpoint(log_root: str) -> Optional[str]:
    st = get_checkpoint_state(log_root)
    if st is None:
        return None
    prefix = _abs(log_root, st.model_checkpoint_path)
    return prefix if os.path.exists(prefix + ".index") else None


class Saver:
    def __init__(self, log_root: str, keep_max: int = 5, basename: str = "model.ckpt"):
        self.log_root = log_root
        self.keep_max = keep_max
        self.basename = basename
        self.last_save_time = time.time()
        self.last_save_step = -1

    def save(self, variables: Dict[str, np.ndarray], global_step: int) -> str:
        os.makedirs(self.log_root, exist_ok=True)
        name = f"{self.basename}-{int(global_step)}"
        prefix = os.path.join(self.log_root, name)
        write_bundle(prefix, variables)
        st = get_checkpoint_state(self.log_root) or CheckpointState()
        paths = [p for p in st.all_model_checkpoint_paths if p != name]
        paths.append(name)
        while self.keep_max > 0 and len(paths) > self.keep_max:
            old = paths.pop(0)
            for f in glob.glob(glob.escape(_abs(self.log_root, old)) + ".*"):
                if re.search(r"\.(index|meta|data-\d{5}-of-\d{5})$", f):
                    try:
                        os.remove(f)
                    except OSError:
                        pass
        new = CheckpointState(name, paths)
        tmp = os.path.join(self.log_root, STATE_FILE + ".tmp")
        with open(tmp, "w") as f:
            f.write(new.to_text())
        os.replace(tmp, os.path.join(self.log_root, STATE_FILE))
        self.last_save_time = time.time()
        self.last_save_step = int(global_step)
        return prefix

    @staticmethod
    def restore(prefix: str, names: Optional[List[str]] = None) -> Dict[str, np.ndarray]:
        """Reads the named variables (all if None). The evaluator only needs the non-Momentum ones
        (its graph has no optimizer, resnet_model.py:65-66)."""
        r = BundleReader(prefix)
        keys = r.keys() if names is None else names
        missing = [k for k in keys if not r.has_tensor(k)]
        if missing:
            raise KeyError(f"checkpoint {prefix} lacks {missing[:5]}{'...' if len(missing) > 5 else ''}")
        return {k: r.get_tensor(k) for k in keys}
