"""Variable store with TensorFlow-style names, collections and parameter-server placement.

Role of the reference's ``inception/slim/variables.py`` (``variable`` :249-290, ``global_step`` :222-245,
``VariableDeviceChooser`` :175-206, ``variable_device`` :210-218, collection helpers :97-172) plus the
TF-core pieces the trainers lean on (``tf.variable_scope`` name nesting and default-name uniquification,
``tf.train.replica_device_setter``, ``tf.fixed_size_partitioner``).

There is no graph: a :class:`VariableStore` is an ``nn.Module`` that owns every variable of one model.  The
model function runs eagerly inside ``with store.as_default():``; the first pass *creates* variables, later
passes look the same names up again (``begin_forward`` resets the name uniquifier so ``Conv``, ``Conv_1`` …
resolve identically each time).  Names are exactly the reference checkpoint keys; storage
layout is the compute layout (OHWI conv kernels, ``[out, in]`` FC) and ``Variable.tf_layout`` tells the
checkpoint layer how to convert to TensorFlow's HWIO / ``[in, out]``.
"""
from __future__ import annotations

import contextlib
import math
import threading
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch
from torch import nn

from .scopes import add_arg_scope

# collection names (reference: variables.py:91-94, ops.py:42, losses.py)
MODEL_VARIABLES = "_model_variables_"
VARIABLES_TO_RESTORE = "_variables_to_restore_"
UPDATE_OPS_COLLECTION = "_update_ops_"
LOSSES_COLLECTION = "_losses"
REGULARIZATION_LOSSES = "regularization_losses"
GLOBAL_VARIABLES = "variables"
TRAINABLE_VARIABLES = "trainable_variables"
MOVING_AVERAGE_VARIABLES = "moving_average_variables"


# ----------------------------------------------------------------------------------------------
# initializers: callables (shape, generator, fan_in, fan_out) -> tensor
# ----------------------------------------------------------------------------------------------
def truncated_normal_initializer(stddev: float = 1.0, mean: float = 0.0) -> Callable:
    def init(shape, gen, fan_in=None, fan_out=None):
        t = torch.empty(tuple(shape), dtype=torch.float32)
        if t.numel():
            nn.init.trunc_normal_(t, mean=mean, std=stddev, a=mean - 2 * stddev, b=mean + 2 * stddev, generator=gen)
        return t
    return init


def random_normal_initializer(mean: float = 0.0, stddev: float = 1.0) -> Callable:
    """tf.random_normal_initializer (un-truncated) — the GAN zoo's weight init (vgg/nets/pix2pix.py:59)."""
    def init(shape, gen, fan_in=None, fan_out=None):
        return torch.randn(tuple(shape), generator=gen, dtype=torch.float32) * stddev + mean
    return init


def constant_initializer(value: float = 0.0) -> Callable:
    def init(shape, gen, fan_in=None, fan_out=None):
        return torch.full(tuple(shape), float(value), dtype=torch.float32)
    return init


def zeros_initializer() -> Callable:
    return constant_initializer(0.0)


def ones_initializer() -> Callable:
    return constant_initializer(1.0)


def xavier_initializer() -> Callable:
    """tf.contrib.layers.xavier_initializer (uniform): limit = sqrt(6 / (fan_in + fan_out)) — the slim
    default that AlexNet/CifarNet end up with because their arg_scopes are never applied."""
    def init(shape, gen, fan_in=None, fan_out=None):
        lim = math.sqrt(6.0 / max(1, (fan_in or 1) + (fan_out or 1)))
        return (torch.rand(tuple(shape), generator=gen, dtype=torch.float32) * 2 - 1) * lim
    return init


def variance_scaling_initializer(scale: float = 1.0) -> Callable:
    """tf.variance_scaling_initializer(): truncated normal, fan_in mode (resnet/resnet_model.py:90)."""
    def init(shape, gen, fan_in=None, fan_out=None):
        std = math.sqrt(scale / max(1, fan_in or 1)) / 0.87962566103423978
        return truncated_normal_initializer(std)(shape, gen)
    return init


def contrib_variance_scaling_initializer(factor: float = 2.0, mode: str = "fan_in") -> Callable:
    """tf.contrib.layers.variance_scaling_initializer(factor, mode, uniform=False): truncated normal with
    stddev = sqrt(1.3 * factor / n), n = fan_in | fan_out | their mean (vgg/nets/nasnet/nasnet.py:117-118)."""
    def init(shape, gen, fan_in=None, fan_out=None):
        fi, fo = max(1, fan_in or 1), max(1, fan_out or 1)
        n = {"fan_in": fi, "fan_out": fo, "fan_avg": (fi + fo) / 2.0}[mode.lower()]
        return truncated_normal_initializer(math.sqrt(1.3 * factor / n))(shape, gen)
    return init


# ----------------------------------------------------------------------------------------------
# placement helpers
# ----------------------------------------------------------------------------------------------
class VariableDeviceChooser:
    """Round-robin variables over parameter servers (reference variables.py:175-206):
    variable k goes to ``/job:ps/task:(k mod P)/<placement>``; with P == 0 everything stays local."""

    def __init__(self, num_parameter_servers: int = 0, ps_device: str = "/job:ps", placement: str = "CPU:0"):
        self._num_ps = num_parameter_servers
        self._ps_device = ps_device
        self._placement = placement if num_parameter_servers == 0 else "CPU:0"
        self._next_task_id = 0

    def __call__(self, op_name: str = "") -> str:
        device_string = ""
        if self._num_ps > 0:
            task_id = self._next_task_id
            self._next_task_id = (self._next_task_id + 1) % self._num_ps
            device_string = "%s/task:%d" % (self._ps_device, task_id)
        device_string += "/%s" % self._placement
        return device_string


def replica_device_setter(ps_tasks: int) -> Callable[[str], str]:
    """tf.train.replica_device_setter(ps_tasks=P): variables round-robin onto ``/job:ps/task:k``
    in creation order (cnn/cifar10_cnn_bsp.py:42)."""
    state = {"next": 0}

    def choose(op_name: str = "") -> str:
        if ps_tasks <= 0:
            return ""
        k = state["next"]
        state["next"] = (k + 1) % ps_tasks
        return "/job:ps/task:%d" % k
    return choose


def fixed_size_partitioner(num_shards: int, axis: int = 0) -> Callable[[Sequence[int]], List[int]]:
    """tf.fixed_size_partitioner: split ``axis`` into min(num_shards, dim) near-equal contiguous slices
    (cifarnet/cifar10_cifarnet_bsp.py:52-53).  Returns the list of slice lengths along ``axis``
    (on the *TensorFlow* shape of the variable)."""
    def part(tf_shape: Sequence[int]) -> List[int]:
        if not tf_shape:
            return [1]
        dim = int(tf_shape[axis])
        n = max(1, min(num_shards, dim))
        base, extra = divmod(dim, n)
        return [base + (1 if i < extra else 0) for i in range(n)]
    part.axis = axis  # type: ignore[attr-defined]
    return part


def variable_device(device, name: str) -> str:
    """Resolve a device spec that may be a callable chooser (reference variables.py:210-218)."""
    if callable(device):
        return device(name) or ""
    return device or ""


def ps_task_of(device: str) -> Optional[int]:
    """``'/job:ps/task:3/CPU:0'`` -> 3 (None when the variable is not on a PS)."""
    if not device or "/job:ps" not in device:
        return None
    for part in device.split("/"):
        if part.startswith("task:"):
            return int(part[5:])
    return 0


# ----------------------------------------------------------------------------------------------
# the store
# ----------------------------------------------------------------------------------------------
@dataclass
class Variable:
    name: str
    tensor: torch.Tensor                 # nn.Parameter when trainable, plain buffer otherwise
    trainable: bool
    device: str = ""                     # TF-style placement string ('' = local)
    weight_decay: float = 0.0            # l2 regulariser weight: loss += wd * sum(w^2)/2
    l1: float = 0.0
    tf_layout: Optional[str] = None      # 'hwio' (stored OHWI), 'hwoi' (transposed conv, stored OHWI),
                                         # 'io' (stored [out,in]) or None (as is)
    tf_shape: Tuple[int, ...] = ()
    restore: bool = True
    partitions: Optional[List[int]] = None   # slice lengths along axis 0 of tf_shape (axis-0 partitioner)
    collections: List[str] = field(default_factory=list)

    @property
    def ps_task(self) -> Optional[int]:
        return ps_task_of(self.device)

    def to_tf(self) -> torch.Tensor:
        """Value in TensorFlow layout (HWIO / [in,out])."""
        t = self.tensor.detach()
        if self.tf_layout == "hwio":
            return t.permute(1, 2, 3, 0).contiguous()
        if self.tf_layout == "io":
            return t.t().contiguous()
        if self.tf_layout == "hwoi":          # transposed convolution: stored [C_out, R, S, C_in]
            return t.permute(1, 2, 0, 3).contiguous()
        return t.clone()

    def from_tf(self, value: torch.Tensor) -> None:
        with torch.no_grad():
            v = value.to(self.tensor.dtype)
            if self.tf_layout == "hwio":
                v = v.permute(3, 0, 1, 2)
            elif self.tf_layout == "io":
                v = v.t()
            elif self.tf_layout == "hwoi":
                v = v.permute(2, 0, 1, 3)
            self.tensor.copy_(v.reshape(self.tensor.shape))


_tls = threading.local()


def current_store() -> "VariableStore":
    st = getattr(_tls, "stores", None)
    if not st:
        raise RuntimeError("no VariableStore is active: wrap the model call in `with store.as_default():`")
    return st[-1]


class VariableStore(nn.Module):
    """Owns the variables of one replica; see the module docstring."""

    def __init__(self, seed: int = 0, device_fn: Optional[Callable[[str], str]] = None,
                 partitioner: Optional[Callable] = None):
        super().__init__()
        self._vars: "OrderedDict[str, Variable]" = OrderedDict()
        self._collections: Dict[str, List[str]] = {}
        self._scope: List[str] = []
        self._uniq: Dict[str, int] = {}
        self._gen = torch.Generator().manual_seed(seed)
        self.device_fn = device_fn
        self.partitioner = partitioner
        self._partition_scopes: List[bool] = []
        self.frozen = False                 # after the build pass: creating a new variable is an error
        self.losses: List[torch.Tensor] = []        # LOSSES_COLLECTION of the current forward pass
        self.update_ops: List[Callable[[], None]] = []
        self.end_points: Dict[str, torch.Tensor] = {}
        self._params = nn.ParameterDict()
        self._n_buffers = 0
        self.fetch_hook: Optional[Callable[["Variable"], None]] = None   # called on every look-up of an existing variable

    # ---- activation -------------------------------------------------------------------------
    @contextlib.contextmanager
    def as_default(self):
        st = getattr(_tls, "stores", None)
        if st is None:
            st = []
            _tls.stores = st
        st.append(self)
        try:
            yield self
        finally:
            st.pop()

    def begin_forward(self) -> None:
        """Reset per-pass state so op scopes get the same unique names as in the build pass."""
        self._uniq.clear()
        self._scope.clear()
        self.losses = []
        self.update_ops = []
        self.end_points = {}
        if self.device_fn is not None and hasattr(self.device_fn, "_next_task_id") and not self.frozen:
            pass

    def freeze(self) -> None:
        self.frozen = True

    # ---- name scopes ------------------------------------------------------------------------
    def _prefix(self) -> str:
        return "/".join(self._scope)

    @contextlib.contextmanager
    def scope(self, name: Optional[str], default_name: Optional[str] = None, partitioned: bool = False):
        """``tf.variable_scope(name, default_name)``: an explicit name is used verbatim (re-entering
        re-uses it); a default name is uniquified within the parent (``Conv``, ``Conv_1`` …)."""
        if name is None or name == "":
            if default_name is None:
                yield self._prefix()
                return
            key = (self._prefix() + "/" if self._scope else "") + default_name
            idx = self._uniq.get(key, 0)
            self._uniq[key] = idx + 1
            name = default_name if idx == 0 else f"{default_name}_{idx}"
        parts = [p for p in name.split("/") if p]
        self._scope.extend(parts)
        self._partition_scopes.append(partitioned)
        try:
            yield self._prefix()
        finally:
            self._partition_scopes.pop()
            del self._scope[len(self._scope) - len(parts):]

    def unique_name(self, base: str) -> str:
        """tf.layers-style auto names: ``conv2d``, ``conv2d_1`` … counted per store and pass."""
        key = "@" + self._prefix() + "/" + base
        idx = self._uniq.get(key, 0)
        self._uniq[key] = idx + 1
        return base if idx == 0 else f"{base}_{idx}"

    # ---- variables --------------------------------------------------------------------------
    def has(self, full_name: str) -> bool:
        return full_name in self._vars

    def get(self, full_name: str) -> Variable:
        return self._vars[full_name]

    def full_name(self, name: str) -> str:
        pre = self._prefix()
        return f"{pre}/{name}" if pre else name

    def variable(self, name: str, shape: Sequence[int], initializer: Optional[Callable] = None,
                 trainable: bool = True, weight_decay: float = 0.0, l1: float = 0.0,
                 collections: Optional[Iterable[str]] = None, device=None, restore: bool = True,
                 tf_layout: Optional[str] = None, tf_shape: Optional[Sequence[int]] = None,
                 fan_in: Optional[int] = None, fan_out: Optional[int] = None,
                 dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """Create (build pass) or fetch (later passes) the variable ``<scope>/<name>``."""
        full = self.full_name(name)
        v = self._vars.get(full)
        if v is not None:
            if tuple(v.tensor.shape) != tuple(shape):
                raise ValueError(f"variable {full} exists with shape {tuple(v.tensor.shape)}, asked {tuple(shape)}")
            if self.fetch_hook is not None:
                self.fetch_hook(v)
            return v.tensor
        if self.frozen:
            raise RuntimeError(f"variable {full} requested after the store was frozen")
        init = initializer or zeros_initializer()
        value = init(shape, self._gen, fan_in, fan_out).to(dtype)
        tf_shape = tuple(tf_shape) if tf_shape is not None else tuple(shape)
        dev = variable_device(device if device is not None else self.device_fn, full)
        if trainable:
            tensor = nn.Parameter(value, requires_grad=True)
            self._params[full.replace(".", "_").replace("/", "__")] = tensor
        else:
            bname = f"_buf{self._n_buffers}"
            self._n_buffers += 1
            self.register_buffer(bname, value)
            tensor = getattr(self, bname)
        colls = [GLOBAL_VARIABLES, MODEL_VARIABLES]
        if trainable:
            colls.append(TRAINABLE_VARIABLES)
        if restore:
            colls.append(VARIABLES_TO_RESTORE)
        for c in (collections or []):
            if c not in colls:
                colls.append(c)
        parts = None
        if self.partitioner is not None and any(self._partition_scopes) and len(tf_shape) > 0:
            parts = self.partitioner(tf_shape)
        var = Variable(full, tensor, trainable, dev, weight_decay, l1, tf_layout, tf_shape, restore, parts, colls)
        if not trainable:
            var._buffer_name = bname  # type: ignore[attr-defined]
        self._vars[full] = var
        for c in colls:
            self._collections.setdefault(c, []).append(full)
        return tensor

    def rebind(self, full_name: str, tensor: torch.Tensor) -> None:
        """Point a variable at new storage (a view of the flat parameter-server buffer)."""
        v = self._vars[full_name]
        if v.trainable:
            v.tensor.data = tensor
        else:
            setattr(self, v._buffer_name, tensor)  # type: ignore[attr-defined]
            v.tensor = tensor

    def _apply(self, fn, *a, **k):  # keep Variable.tensor pointing at moved buffers after .to()/.cuda()
        out = super()._apply(fn, *a, **k)
        for v in self._vars.values():
            if not v.trainable:
                v.tensor = getattr(self, v._buffer_name)  # type: ignore[attr-defined]
        return out

    # ---- collections ------------------------------------------------------------------------
    def all(self) -> List[Variable]:
        return list(self._vars.values())

    def add_to_collection(self, coll: str, full_name: str) -> None:
        self._collections.setdefault(coll, [])
        if full_name not in self._collections[coll]:
            self._collections[coll].append(full_name)
            self._vars[full_name].collections.append(coll)

    def get_collection(self, coll: str) -> List[Variable]:
        return [self._vars[n] for n in self._collections.get(coll, [])]

    def get_variables(self, scope: str = "", suffix: str = "") -> List[Variable]:
        """reference variables.py:103-118: model variables filtered by scope prefix and name suffix."""
        out = []
        for v in self.get_collection(MODEL_VARIABLES):
            if scope and not v.name.startswith(scope):
                continue
            if suffix and not v.name.endswith(suffix):
                continue
            out.append(v)
        return out

    def get_variables_to_restore(self) -> List[Variable]:
        return self.get_collection(VARIABLES_TO_RESTORE)

    def get_variables_by_name(self, given_name: str, scope: str = "") -> List[Variable]:
        """Variables whose last path component is ``given_name`` (reference variables.py:131-143)."""
        return [v for v in self.get_variables(scope) if v.name.split("/")[-1] == given_name]

    def get_unique_variable(self, name: str) -> Variable:
        cands = [v for v in self.all() if v.name == name]
        if not cands:
            raise ValueError(f"Couldnt find variable {name}")
        if len(cands) > 1:
            raise ValueError(f"Variable {name} does not uniquely identify a variable")
        return cands[0]

    def trainable_variables(self) -> List[Variable]:
        return [v for v in self._vars.values() if v.trainable]

    def regularization_loss(self) -> torch.Tensor:
        """sum over variables of  wd * sum(w^2)/2  (+ l1 * sum|w|) — reference losses.py:56-71."""
        total = None
        for v in self._vars.values():
            if v.weight_decay:
                t = v.tensor.pow(2).sum() * (0.5 * v.weight_decay)
                total = t if total is None else total + t
            if v.l1:
                t = v.tensor.abs().sum() * v.l1
                total = t if total is None else total + t
        if total is None:
            any_t = next(iter(self._vars.values())).tensor
            return torch.zeros((), device=any_t.device)
        return total

    def num_regularizers(self) -> int:
        return sum(1 for v in self._vars.values() if v.weight_decay or v.l1)


# ----------------------------------------------------------------------------------------------
# module-level API mirroring the reference's functions (operate on the current store)
# ----------------------------------------------------------------------------------------------
@add_arg_scope
def variable(name, shape=None, dtype=torch.float32, initializer=None, regularizer=None, trainable=True,
             collections=None, device="", restore=True, tf_layout=None, tf_shape=None, fan_in=None,
             fan_out=None):
    """reference variables.py:249-290.  ``regularizer`` is an ``l2_regularizer``/``l1_regularizer``
    object from ``slim.losses`` (its weight is recorded on the variable and applied by the fused update)."""
    store = current_store()
    wd = getattr(regularizer, "l2", 0.0) if regularizer is not None else 0.0
    l1 = getattr(regularizer, "l1", 0.0) if regularizer is not None else 0.0
    return store.variable(name, shape, initializer, trainable, wd, l1, collections, device or None, restore,
                          tf_layout, tf_shape, fan_in, fan_out, dtype)


def global_step(device: str = "") -> torch.Tensor:
    """int64 scalar named ``global_step`` (reference variables.py:222-245)."""
    store = current_store()
    if store.has("global_step"):
        return store.get("global_step").tensor
    saved, store._scope = store._scope, []
    try:
        t = store.variable("global_step", (), zeros_initializer(), trainable=False, device=device or None,
                           dtype=torch.int64)
    finally:
        store._scope = saved
    return t


def get_variables(scope: str = "", suffix: str = ""):
    return current_store().get_variables(scope, suffix)


def get_variables_to_restore():
    return current_store().get_variables_to_restore()


def get_variables_by_name(given_name: str, scope: str = ""):
    return current_store().get_variables_by_name(given_name, scope)


def get_unique_variable(name: str):
    return current_store().get_unique_variable(name)


def add_variable(full_name: str, restore: bool = True) -> None:
    """Put an existing variable into the model (and restore) collections (reference :97-100)."""
    st = current_store()
    st.add_to_collection(MODEL_VARIABLES, full_name)
    if restore:
        st.add_to_collection(VARIABLES_TO_RESTORE, full_name)
