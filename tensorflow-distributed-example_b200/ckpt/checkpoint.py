"""Checkpoint / resume / export.

Reference behaviour: ``RunConfig(model_dir, save_checkpoints_steps=500)``
(mnist_keras_distributed.py:245-248) makes the Estimator write ``model.ckpt-<step>`` files plus a
``checkpoint`` state file, keep the last 5, and AUTO-RESUME from the latest checkpoint found in
``model_dir``; ``FinalExporter('exporter', serving_input_fn)`` (mnist_keras_distributed.py:264) writes a
SavedModel under ``<model_dir>/export/exporter/<timestamp>/`` with a ``[None, 784]`` float32 signature
(mnist_keras_distributed.py:151-162).

On-disk format here (TensorFlow / h5py are not available): one ``.npz`` per checkpoint keyed by the
Keras variable names with Keras shapes/layouts (``conv2d/kernel`` HWIO, ``dense/kernel`` (in,out),
``batch_normalization/{beta,moving_mean,moving_variance}``) -- the kernel-native layout, so no
transposes in either direction -- plus ``global_step``, optimizer slots (gathered from their owner
ranks when the fused two-shot all-reduce shards them), the RNG seed and the dataset position.
``.safetensors`` is written instead when the path ends with that suffix. Only the chief writes.
"""
from __future__ import annotations

import json
import os
import re
import time
from typing import Dict, List, Optional

import numpy as np
import torch

STATE_FILE = "checkpoint"
PREFIX = "model.ckpt"
_META_KEYS = ("global_step", "_seed", "_dataset_position", "_optimizer")


# ----------------------------------------------------------------------------------------------
# plain weights
# ----------------------------------------------------------------------------------------------
def _write(path: str, arrays: Dict[str, np.ndarray]):
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    if path.endswith(".safetensors"):
        from safetensors.numpy import save_file
        save_file({k: np.ascontiguousarray(v) for k, v in arrays.items()}, path)
        return path
    if not path.endswith(".npz"):
        path = path + ".npz"
    tmp = path + ".tmp.npz"
    np.savez(tmp, **arrays)
    os.replace(tmp, path)   # atomic: an evaluator polling the directory never sees a torn file
    return path


def _read(path: str) -> Dict[str, np.ndarray]:
    if path.endswith(".safetensors"):
        from safetensors.numpy import load_file
        return load_file(path)
    if not path.endswith(".npz") and os.path.exists(path + ".npz"):
        path = path + ".npz"
    with np.load(path, allow_pickle=False) as f:
        return {k: f[k] for k in f.files}


def save_weights(model, path: str) -> str:
    return _write(path, model.named_weights())


def load_weights(model, path: str, strict: bool = True):
    arrays = _read(path)
    missing = []
    for v in model.variables:
        if v.name not in arrays:
            missing.append(v.name)
            continue
        a = np.asarray(arrays[v.name], dtype=np.float32)
        if tuple(a.shape) != v.shape:
            raise ValueError(f"{v.name}: checkpoint shape {a.shape} != variable shape {v.shape}")
        v.value.copy_(torch.from_numpy(a).to(v.value.device))
    if missing and strict:
        raise KeyError(f"variables missing from {path}: {missing}")
    return arrays


# ----------------------------------------------------------------------------------------------
# optimizer slots (possibly sharded by owner rank)
# ----------------------------------------------------------------------------------------------
def _gather_slot(model, slot: Optional[torch.Tensor]) -> Optional[np.ndarray]:
    """Returns the full-length slot (arena.n floats) on every rank."""
    if slot is None:
        return None
    a = model.arena
    st = model.strategy
    if not getattr(a, "state_sharded", False) or st.num_replicas_in_sync == 1:
        return slot[:a.n].detach().cpu().numpy().copy()
    import torch.distributed as dist
    parts: List[Optional[np.ndarray]] = [None] * st.num_replicas_in_sync
    dist.all_gather_object(parts, slot.detach().cpu().numpy())
    return np.concatenate(parts)[:a.n].copy()


def _scatter_slot(model, full: np.ndarray, slot: torch.Tensor):
    a = model.arena
    full_t = torch.zeros(max(a.n_total, len(full)), dtype=torch.float32)
    full_t[:len(full)] = torch.from_numpy(np.asarray(full, dtype=np.float32))
    if getattr(a, "state_sharded", False) and model.strategy.num_replicas_in_sync > 1:
        off, ln = a.state_slice
        piece = torch.zeros(ln, dtype=torch.float32)
        avail = max(0, min(ln, full_t.numel() - off))
        piece[:avail] = full_t[off:off + avail]
        slot.copy_(piece.to(slot.device))
    else:
        slot.copy_(full_t[:slot.numel()].to(slot.device))


# ----------------------------------------------------------------------------------------------
# training checkpoints
# ----------------------------------------------------------------------------------------------
def checkpoint_path(model_dir: str, step: int) -> str:
    return os.path.join(model_dir, f"{PREFIX}-{int(step)}.npz")


def _read_state(model_dir: str) -> Dict:
    p = os.path.join(model_dir, STATE_FILE)
    if not os.path.exists(p):
        return {"model_checkpoint_path": None, "all_model_checkpoint_paths": []}
    with open(p) as f:
        return json.load(f)


def latest_checkpoint(model_dir: str) -> Optional[str]:
    """tf.train.latest_checkpoint: path named by the ``checkpoint`` state file (None if absent)."""
    st = _read_state(model_dir)
    name = st.get("model_checkpoint_path")
    if name:
        p = os.path.join(model_dir, name)
        if os.path.exists(p):
            return p
    # fall back to scanning (state file lost)
    best, best_step = None, -1
    if os.path.isdir(model_dir):
        for f in os.listdir(model_dir):
            m = re.fullmatch(rf"{re.escape(PREFIX)}-(\d+)\.npz", f)
            if m and int(m.group(1)) > best_step:
                best, best_step = os.path.join(model_dir, f), int(m.group(1))
    return best


def step_of(path: str) -> int:
    m = re.search(rf"{re.escape(PREFIX)}-(\d+)", os.path.basename(path))
    return int(m.group(1)) if m else -1


def save_checkpoint(model, model_dir: str, step: int, max_to_keep: int = 5,
                    dataset_position: int = 0) -> Optional[str]:
    """Collective when optimizer state is sharded (every rank must call); only the chief writes."""
    model.sync()
    opt = model.optimizer
    s1 = _gather_slot(model, model.arena.s1) if opt is not None and opt.num_slots >= 1 else None
    s2 = _gather_slot(model, model.arena.s2) if opt is not None and opt.num_slots >= 2 else None
    if not model.strategy.is_chief:
        return None
    arrays = dict(model.named_weights())
    arrays["global_step"] = np.asarray(step, dtype=np.int64)
    arrays["_seed"] = np.asarray(model.seed, dtype=np.int64)
    arrays["_dataset_position"] = np.asarray(dataset_position, dtype=np.int64)
    if opt is not None:
        arrays["_optimizer"] = np.frombuffer(json.dumps(opt.get_config()).encode(), dtype=np.uint8)
        if s1 is not None:
            arrays["_slot1"] = s1
        if s2 is not None:
            arrays["_slot2"] = s2
    path = _write(checkpoint_path(model_dir, step), arrays)
    st = _read_state(model_dir)
    allp = [p for p in st.get("all_model_checkpoint_paths", []) if os.path.exists(os.path.join(model_dir, p))]
    name = os.path.basename(path)
    if name in allp:
        allp.remove(name)
    allp.append(name)
    while max_to_keep and len(allp) > max_to_keep:
        old = allp.pop(0)
        try:
            os.remove(os.path.join(model_dir, old))
        except OSError:
            pass
    tmp = os.path.join(model_dir, STATE_FILE + ".tmp")
    with open(tmp, "w") as f:
        json.dump({"model_checkpoint_path": name, "all_model_checkpoint_paths": allp,
                   "last_preserved_timestamp": time.time()}, f)
    os.replace(tmp, os.path.join(model_dir, STATE_FILE))
    return path


def restore_checkpoint(model, path: str) -> int:
    """Restores weights, optimizer slots and step; returns the global step."""
    arrays = load_weights(model, path, strict=True)
    step = int(arrays.get("global_step", 0))
    opt = model.optimizer
    if opt is not None:
        if getattr(opt, "_bound_to", None) is not model:
            opt.iterations = step
            model._bind_optimizer()
        opt.iterations = step
        if opt.hparams is not None:
            from ..ops import native
            native.hparams_set(opt.hparams, step=step)
            if opt.kind == "adam":
                import struct
                for idx, val in ((native.HP_B1POW, opt.beta_1 ** step), (native.HP_B2POW, opt.beta_2 ** step)):
                    opt.hparams[idx].fill_(struct.unpack("<i", struct.pack("<f", float(val)))[0])
        if "_slot1" in arrays and model.arena.s1 is not None:
            _scatter_slot(model, arrays["_slot1"], model.arena.s1)
        if "_slot2" in arrays and model.arena.s2 is not None:
            _scatter_slot(model, arrays["_slot2"], model.arena.s2)
    model.restored_dataset_position = int(arrays.get("_dataset_position", 0))
    return step


# ----------------------------------------------------------------------------------------------
# export for serving
# ----------------------------------------------------------------------------------------------
def export_saved_model(model, export_dir_base: str, serving_input_receiver=None,
                       timestamp: Optional[int] = None) -> str:
    """Writes ``<export_dir_base>/<timestamp>/{saved_model.json, variables.npz}``. The JSON carries the
    layer configs and the serving signature (input name/shape/dtype from the ServingInputReceiver:
    [None, 784] float32 in the reference)."""
    ts = int(time.time()) if timestamp is None else int(timestamp)
    out = os.path.join(export_dir_base, str(ts))
    while os.path.exists(out):
        ts += 1
        out = os.path.join(export_dir_base, str(ts))
    os.makedirs(out)
    sig_in = {"name": "input", "shape": [None] + list(model.input_shape), "dtype": "float32"}
    if serving_input_receiver is not None:
        sig_in = serving_input_receiver.signature()
    spec = {
        "format": "tfde_b200.saved_model.v1",
        "model_name": model.name,
        "input_shape": list(model.input_shape),
        "layers": [{"class": type(l).__name__, "config": l.get_config()} for l in model.layers],
        "signature": {"serving_default": {"inputs": sig_in,
                                           "outputs": {"name": "output", "shape": [None] + list(model.output_shape),
                                                       "dtype": "float32"}}},
    }
    with open(os.path.join(out, "saved_model.json"), "w") as f:
        json.dump(spec, f, indent=1)
    _write(os.path.join(out, "variables.npz"), model.named_weights())
    return out


def load_saved_model(path: str, device: Optional[str] = None):
    """Rebuilds the exported Sequential and returns it (``model.predict(x)`` serves the signature)."""
    from ..keras import layers as L
    from ..keras.models import Sequential
    from ..parallel.strategy import Strategy
    with open(os.path.join(path, "saved_model.json")) as f:
        spec = json.load(f)
    with Strategy(device=device).scope():
        model = Sequential([L.layer_from_config(d["class"], d["config"]) for d in spec["layers"]],
                           name=spec.get("model_name"))
        model.build(tuple(spec["input_shape"]))
    load_weights(model, os.path.join(path, "variables.npz"))
    model.signature = spec["signature"]
    return model
