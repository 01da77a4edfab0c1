"""Checkpoints keyed by the reference's TensorFlow variable names.

What ``tf.train.Saver`` + ``Supervisor`` do for the reference:

* files ``<train_dir>/model.ckpt-<global_step>`` plus the text-proto state file ``<train_dir>/checkpoint``
  (``model_checkpoint_path: "model.ckpt-N"`` / ``all_model_checkpoint_paths: ...``), which is how the eval
  scripts find the newest checkpoint and parse the step from the suffix (cnn/cifar10_eval.py:73-80);
* every variable under its full logical name in **TensorFlow layout** (conv kernels HWIO, FC ``[in, out]``),
  independent of how it was sharded over parameter servers;
* optimizer slots ``<var>/RMSProp`` (ms) and ``<var>/RMSProp_1`` (mom), EMA shadows
  ``<var>/ExponentialMovingAverage``, and the global step under the trainer's name for it;
* keep-N policy (``max_to_keep``: 5 by default, 1 for ResNet — resnet/cifar10_resnet_bsp.py:107).

Two containers: a NumPy ``.npz`` (one array per name; the default) and TensorFlow's own V2 bundle
(``model.ckpt-N.index`` + ``model.ckpt-N.data-00000-of-00001``, ``ckpt/tf_bundle.py``; ``--ckpt_format tf``).
``load`` / ``exists`` / the keep-N policy recognise both, so a run can resume from, evaluate or fine-tune on either.
"""
from __future__ import annotations

import os
import re
import shutil
import time
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch

from ..slim.variables import Variable, VariableStore
from . import tf_bundle

STATE_FILE = "checkpoint"
PREFIX = "model.ckpt"
EMA_SUFFIX = "/ExponentialMovingAverage"


class CheckpointState:
    def __init__(self, model_checkpoint_path: str, all_model_checkpoint_paths: List[str]):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = all_model_checkpoint_paths


def _write_state(train_dir: str, paths: List[str]) -> None:
    tmp = os.path.join(train_dir, STATE_FILE + ".tmp")
    with open(tmp, "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % paths[-1])
        for p in paths:
            f.write('all_model_checkpoint_paths: "%s"\n' % p)
    os.replace(tmp, os.path.join(train_dir, STATE_FILE))


def get_checkpoint_state(checkpoint_dir: str) -> Optional[CheckpointState]:
    """``tf.train.get_checkpoint_state``: parse ``<dir>/checkpoint`` (None when absent)."""
    path = os.path.join(checkpoint_dir, STATE_FILE)
    if not os.path.exists(path):
        return None
    latest, allp = None, []
    with open(path) as f:
        for line in f:
            m = re.match(r'\s*(\w+):\s*"(.*)"', line)
            if not m:
                continue
            if m.group(1) == "model_checkpoint_path":
                latest = m.group(2)
            elif m.group(1) == "all_model_checkpoint_paths":
                allp.append(m.group(2))
    if latest is None:
        return None

    def absolutise(p):
        return p if os.path.isabs(p) else os.path.join(checkpoint_dir, p)
    return CheckpointState(absolutise(latest), [absolutise(p) for p in allp])


def latest_checkpoint(checkpoint_dir: str) -> Optional[str]:
    st = get_checkpoint_state(checkpoint_dir)
    return st.model_checkpoint_path if st else None


def step_of(path: str) -> int:
    """global step from ``.../model.ckpt-1234`` (cnn/cifar10_eval.py:80)."""
    return int(path.split("/")[-1].split("-")[-1].split(".")[0])


def _file(path: str) -> str:
    return path if path.endswith(".npz") else path + ".npz"


def exists(path: str) -> bool:
    """A checkpoint with this prefix exists in either container."""
    return bool(path) and (os.path.exists(_file(path)) or tf_bundle.is_bundle(path))


def _remove(path: str) -> None:
    for f in [_file(path)] + tf_bundle.bundle_files(path):
        try:
            os.remove(f)
        except OSError:
            pass


def collect_tensors(store: VariableStore, slot_flats: Optional[Dict[str, torch.Tensor]] = None,
                    layout=None, global_step: int = 0, global_step_name: str = "global_step") -> Dict[str, np.ndarray]:
    """name -> array in TF layout.  ``slot_flats``: full-length flat tensors per slot
    ('RMSProp', 'RMSProp_1', 'ExponentialMovingAverage') gathered from the PS shards."""
    out: Dict[str, np.ndarray] = {}
    for v in store.all():
        if v.name == "global_step":
            continue
        if v.tf_layout == "minmax":          # quantisation range: TensorFlow keeps two scalars, .../min and .../max
            base, t = v.name.rsplit("/", 1)[0], v.tensor.detach().cpu().numpy()
            out[base + "/min"], out[base + "/max"] = np.asarray(t[0]), np.asarray(t[1])
            continue
        out[v.name] = v.to_tf().cpu().numpy()
    if slot_flats and layout is not None:
        for slot, flat in slot_flats.items():
            for e in layout.entries:
                v = store.get(e.name)
                t = flat[e.offset:e.offset + e.numel].view(e.shape)
                tmp = Variable(e.name, t, True, tf_layout=v.tf_layout)
                out[f"{e.name}/{slot}"] = tmp.to_tf().cpu().numpy()
    out[global_step_name] = np.asarray(global_step, dtype=np.int64)
    return out


def save(train_dir: str, tensors: Dict[str, np.ndarray], global_step: int, max_to_keep: int = 5,
         fmt: str = "npz") -> str:
    """``fmt``: 'npz' or 'tf' (TensorFlow V2 bundle: ``.index`` + ``.data-00000-of-00001``)."""
    os.makedirs(train_dir, exist_ok=True)
    name = f"{PREFIX}-{int(global_step)}"
    if fmt == "tf":
        tf_bundle.write_bundle(os.path.join(train_dir, name), tensors)
    elif fmt == "npz":
        tmp = os.path.join(train_dir, name + ".tmp.npz")
        np.savez(tmp, **{k.replace("/", "|"): v for k, v in tensors.items()})
        os.replace(tmp, os.path.join(train_dir, name + ".npz"))
    else:
        raise ValueError("checkpoint format must be 'npz' or 'tf', not %r" % fmt)
    st = get_checkpoint_state(train_dir)
    paths = [os.path.basename(p) for p in (st.all_model_checkpoint_paths if st else [])]
    paths = [p for p in paths if p != name and exists(os.path.join(train_dir, p))] + [name]
    while max_to_keep and len(paths) > max_to_keep:
        _remove(os.path.join(train_dir, paths.pop(0)))
    _write_state(train_dir, paths)
    return os.path.join(train_dir, name)


def load(path: str) -> Dict[str, np.ndarray]:
    if not os.path.exists(_file(path)) and tf_bundle.is_bundle(path):
        return tf_bundle.read_bundle(path)
    with np.load(_file(path)) as z:
        return {k.replace("|", "/"): z[k] for k in z.files}


def restore(store: VariableStore, tensors: Dict[str, np.ndarray], use_ema: bool = False,
            strict: bool = True) -> List[str]:
    """Assign checkpoint values to the store's variables.  ``use_ema``: prefer the shadow value
    ``<name>/ExponentialMovingAverage`` when present (``variable_averages.variables_to_restore()``,
    cnn/cifar10_eval.py:132-135).  Returns the names restored."""
    done = []
    for v in store.get_variables_to_restore():
        if v.name == "global_step":
            continue
        if v.tf_layout == "minmax":
            base = v.name.rsplit("/", 1)[0]
            if base + "/min" in tensors and base + "/max" in tensors:
                with torch.no_grad():
                    v.tensor.copy_(torch.tensor([float(tensors[base + "/min"]), float(tensors[base + "/max"])]))
                done.append(v.name)
            elif strict:
                raise KeyError(f"checkpoint has no tensors named {base}/min, {base}/max")
            continue
        key = v.name
        if use_ema and (v.name + EMA_SUFFIX) in tensors:
            key = v.name + EMA_SUFFIX
        if key not in tensors:
            if strict:
                raise KeyError(f"checkpoint has no tensor named {key}")
            continue
        v.from_tf(torch.from_numpy(np.asarray(tensors[key])))
        done.append(v.name)
    return done


SLOT_NAMES = ("RMSProp", "RMSProp_1", "ExponentialMovingAverage")


def slot_flats_from(store: VariableStore, tensors: Dict[str, np.ndarray], layout) -> Dict[str, torch.Tensor]:
    """Inverse of the slot part of :func:`collect_tensors`: full-length flat tensors (parameter order of
    ``layout``) for every optimiser slot / EMA shadow the checkpoint holds for *all* trainable variables —
    ``Supervisor`` / ``Saver`` restore these together with the variables (inception/imagenet_inception_bsp.py:156-167).
    A slot that is missing for some variable is left out altogether."""
    out: Dict[str, torch.Tensor] = {}
    for slot in SLOT_NAMES:
        if not all(f"{e.name}/{slot}" in tensors for e in layout.entries):
            continue
        flat = torch.zeros(layout.total, dtype=torch.float32)
        if slot == "RMSProp":
            flat.fill_(1.0)                       # padding between tensors keeps the slot's initial value
        for e in layout.entries:
            v = store.get(e.name)
            dst = flat[e.offset:e.offset + e.numel].view(e.shape)
            tmp = Variable(e.name, dst, True, tf_layout=v.tf_layout)
            tmp.from_tf(torch.from_numpy(np.asarray(tensors[f"{e.name}/{slot}"])))
        out[slot] = flat
    return out


def restore_training_state(trainer, tensors: Dict[str, np.ndarray]) -> List[str]:
    """Variables, optimiser slots and EMA shadows of a checkpoint into a live trainer (resume).  Without shadows in
    the checkpoint the EMA restarts from the restored weights rather than from the random initial values."""
    done = restore(trainer.store, tensors, strict=False)
    comm = trainer.comm
    flats = slot_flats_from(trainer.store, tensors, trainer.layout)
    if flats:
        comm.load_slot_flats(flats)
        done += sorted(flats)
    if "ExponentialMovingAverage" not in flats:
        comm.reset_ema_to_params()
    return done


def wipe(train_dir: str) -> None:
    """CIFAR chiefs delete and recreate train_dir at start (cifarnet/cifar10_cifarnet_bsp.py:45-48)."""
    if os.path.isdir(train_dir):
        shutil.rmtree(train_dir)
    os.makedirs(train_dir, exist_ok=True)


def wait_for_checkpoint(checkpoint_dir: str, poll_secs: float = 1.0, timeout: Optional[float] = None) -> Optional[str]:
    """Poll until a checkpoint exists (re-reading the state each time — the reference's wait loops do not,
    cifarnet/cifar10_cifarnet_eval.py:75-82)."""
    t0 = time.time()
    while True:
        p = latest_checkpoint(checkpoint_dir)
        if p and exists(p):
            return p
        if timeout is not None and time.time() - t0 > timeout:
            return None
        time.sleep(poll_secs)
