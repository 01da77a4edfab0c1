"""The worker-side training engine: model replica + flat PS layout + consistency mode.

This is what the body of every reference trainer script builds by hand (SURVEY §2.2 "common skeleton",
call stack §3.2): variables placed on parameter servers, model, loss, staircase learning-rate decay,
optimiser (wrapped in SyncReplicasOptimizer for BSP), then a loop of ``sess.run(train_op)``.  Here:

    trainer = Trainer(TrainSetup(model='resnet', mode='bsp', ...), comm, device)
    out = trainer.train_step(images, labels)        # pull -> fwd/bwd -> push (+fused update) -> clock

One step =  forward + backward on this rank's batch  ->  ``comm.bsp_step`` (mean of N gradients, one
update per global step, EMA on the owner)  or  ``comm.async_step`` (one-sided apply, global_step += 1 per
worker step)  ->  for SSP ``comm.ssp_wait(step, bound)``.

On CUDA the forward/backward (+ single-GPU update) is captured once into a CUDA graph and replayed, so the
~hundreds of small kernels of a CIFAR-sized step cost one launch.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, Optional, Tuple

import torch

from ..config import Preset, preset_for
from ..models import alexnet, cifarnet, inception_v3, plain_cnn, registry, resnet
from ..ops import functional as Fn
from ..optim import EmaConfig, OptimizerConfig, exponential_decay
from ..slim import variables as V
from ..slim.variables import VariableStore
from .comm import Comm, LocalComm
from .layout import FlatLayout


@dataclass
class TrainSetup:
    model: str = "resnet"
    mode: str = "bsp"                 # bsp | asp | ssp
    batch_size: int = 128             # per worker (FLAGS.batch_size)
    num_workers: int = 1
    staleness: int = 5
    resnet_size: int = 32
    ref_compat: bool = False
    placement: str = "flat"
    seed: int = 0
    initial_lr: Optional[float] = None
    epochs_per_decay: Optional[float] = None
    decay_factor: Optional[float] = None
    fused_l2: bool = True             # apply wd*w inside the PS update instead of autograd (same gradient)
    cuda_graph: bool = True
    fuse_bn: bool = True              # native: BN statistics from conv epilogues, BN+ReLU applied on load
    preset: Optional[Preset] = None
    branch_streams: bool = True       # CUDA: independent branches of Inception-style blocks on side streams
    image_size: Optional[int] = None  # zoo networks: input resolution / classes / arg-scope weight decay
    num_classes: Optional[int] = None
    weight_decay: Optional[float] = None
    quantize: bool = False            # fake-quantised weights / activations with folded batch norm (slim/quantize.py)
    quant_delay: int = 0              # global step at which the fake-quant ops switch on
    precision: str = "tf32"           # tensor-core multiplies: 'tf32' | 'tf32x3' (split operands, ~fp32; 3x the work)

    def resolved(self) -> "TrainSetup":
        p = self.preset or preset_for(self.model, self.mode)
        self.preset = p
        if self.initial_lr is None:
            self.initial_lr = p.initial_lr
        if self.epochs_per_decay is None:
            self.epochs_per_decay = p.epochs_per_decay
        if self.decay_factor is None:
            self.decay_factor = p.decay_factor
        return self


# ---- per-model forward + loss ---------------------------------------------------------------------
def build_forward(setup: TrainSetup) -> Callable[[torch.Tensor, torch.Tensor, bool], Tuple[torch.Tensor, torch.Tensor]]:
    """-> fn(images NHWC, labels, is_training) -> (data loss, logits).  Runs inside ``store.as_default()``."""
    m = setup.model
    if m == "cnn":
        def fn(x, y, tr):
            logits = plain_cnn.inference(x)
            return plain_cnn.loss(logits, y), logits
    elif m == "cifarnet":
        def fn(x, y, tr):
            out = cifarnet.cifarnet(x, 10, is_training=tr, return_logits=not setup.ref_compat)
            # reference: tf.losses.softmax_cross_entropy(onehot, logits=<softmax output>) — Q4 under ref_compat
            return Fn.softmax_cross_entropy(out, y.long()), out
    elif m == "alexnet":
        def fn(x, y, tr):
            logits, _ = alexnet.alexnet_v2(x, 10, is_training=tr)
            return Fn.softmax_cross_entropy(logits, y.long()), logits
    elif m == "vgg":
        net = registry.get_network_fn("vgg_16", num_classes=10)      # is_training=False, wd=0 (Q5)
        def fn(x, y, tr):
            with V.current_store().scope("root"):
                logits, _ = net(x)
            return Fn.softmax_cross_entropy(logits, y.long()), logits
    elif m == "resnet":
        net = resnet.cifar10_resnet_v2_generator(setup.resnet_size, 10)
        def fn(x, y, tr):
            logits = net(x, tr)
            return Fn.softmax_cross_entropy(logits, y.long()), logits
    elif m == "inception":
        def fn(x, y, tr):
            logits, aux = inception_v3.inference(x, 1001, for_training=tr)
            return inception_v3.loss((logits, aux), y.long()), logits
    elif registry.is_zoo_model(m):
        # any other network of the zoo, built through the factory under its own arg scope; loss as in the zoo's
        # trainer (softmax cross-entropy on the logits, + 0.4 x the auxiliary head where the network has one,
        # + the arg scope's L2 regularisers, which the PS update applies as wd * w)
        spec = registry.spec_for(m, setup.image_size, setup.num_classes)
        nets = {}

        def fn(x, y, tr):
            if tr not in nets:
                # weight_decay=None: the L2 weight of the network's own arg scope
                nets[tr] = registry.get_network_fn(m, spec.num_classes, weight_decay=setup.weight_decay,
                                                   is_training=tr)
            out, end_points = nets[tr](x)
            logits = out[0] if isinstance(out, tuple) else out
            loss = Fn.softmax_cross_entropy(logits, y.long())
            aux = end_points.get("AuxLogits") if isinstance(end_points, dict) else None
            if aux is not None and tr:
                loss = loss + Fn.softmax_cross_entropy(aux, y.long(), weight=0.4)
            return loss, logits
    else:
        raise ValueError(m)
    return fn


_SCOPE_PREFIX = {"cifarnet": "partitioned_space", "alexnet": "partitioned_space"}


class Trainer:
    def __init__(self, setup: TrainSetup, comm: Optional[Comm] = None, device: Optional[torch.device] = None):
        self.setup = setup.resolved()
        self.comm = comm or LocalComm()
        self.device = device or torch.device("cpu")
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.p = self.setup.preset
        spec = registry.spec_for(self.setup.model, self.setup.image_size, self.setup.num_classes)
        self.spec = spec
        self.store = VariableStore(seed=self.setup.seed)
        if self.setup.precision not in ("tf32", "tf32x3"):
            raise ValueError("precision must be 'tf32' or 'tf32x3'")
        # the fused batch-norm apply on load is not linear in the operand: 3xTF32 runs the unfused layers
        self.store.fuse_bn = bool(self.setup.fuse_bn) and self.setup.precision == "tf32"
        self.forward = build_forward(self.setup)
        self.prefix = _SCOPE_PREFIX.get(self.setup.model, "")
        if self.setup.quantize:
            from ..slim.quantize import QuantState
            self.store.quant = QuantState(quant_delay=int(self.setup.quant_delay))
        # ---- build pass: create variables (CPU, deterministic from the seed), then move ----
        with torch.no_grad():
            x0 = torch.zeros(2, spec.image_size, spec.image_size, spec.channels)
            y0 = torch.zeros(2, dtype=torch.int64)
            self._run_forward(x0, y0, True, build=True)
        self.store.freeze()
        self.store.to(self.device)
        if self.setup.quantize and not any(v.tf_layout == "minmax" for v in self.store.all()):
            raise ValueError("--quantize: %r builds its layers without the slim layer functions that carry the "
                             "fake-quantisation hooks (plain CNN and the tf.layers ResNet); nothing would be "
                             "quantised" % self.setup.model)
        if self.setup.quantize:
            # the quant_delay switch lives in device memory: flipping it needs no re-capture of the step graph
            self.store.quant.gate = torch.zeros((), dtype=torch.float32, device=self.device)
        # ---- PS layout + optimiser -------------------------------------------------------------
        wd = self.p.weight_decay if self.p.decay_all_trainables else None
        owners = getattr(self.comm, "owners", self.comm.world)     # PS tasks (== workers in the SPMD transports)
        self.layout = FlatLayout(self.store, owners, self.setup.placement, self.device,
                                 default_weight_decay=wd)
        ema = EmaConfig(self.p.ema_decay) if self.p.ema_decay else None
        self.opt_cfg = OptimizerConfig(kind=self.p.optimizer, weight_decay=self.layout.weight_decay, ema=ema)
        self.comm.mode = self.setup.mode          # transports that shard differently per consistency mode
        self.comm.setup(self.layout, self.opt_cfg)
        if hasattr(self.comm, "on_fetch"):
            # parameter-server pull gates: the transport learns where each variable is first used (parallel/symm.py)
            self.store.fetch_hook = self.comm.on_fetch
            if self.setup.quantize:
                self.comm.no_fused_gate = True    # folded / fake-quantised weights are read by other kernels first
        if hasattr(self.comm, "grad_ready"):
            for v in self.store.trainable_variables():
                v.tensor.register_post_accumulate_grad_hook(self.comm.grad_ready)
        # non-trainable state (BN moving statistics): replicated per rank, chief's values at start (C6/C7)
        self.comm.broadcast_state([v.tensor for v in self.store.all() if not v.trainable], src=0)
        self.local_step = 0
        self._graph = None
        self._wt_batch = None
        self._static = None
        self.graph_kernel_launches = 0      # native kernels inside one replay of the captured step
        self.last_timing: Dict[str, float] = {}

    # ---- helpers ---------------------------------------------------------------------------------
    def _run_forward(self, x, y, training, build=False):
        from ..ops import streams
        streams.set_branch_streams(self.setup.branch_streams and self.device.type == "cuda" and not build)
        if self.device.type == "cuda":
            from ..ops import conv_native
            conv_native.set_precision(self.setup.precision)
        if self.setup.quantize:
            self.store.quant.is_training = bool(training)
        with self.store.as_default():
            self.store.begin_forward()
            if self.prefix:
                with self.store.scope(self.prefix):
                    return self.forward(x, y, training)
            return self.forward(x, y, training)

    def num_workers(self) -> int:
        return max(1, self.setup.num_workers)

    def decay_steps(self) -> int:
        p, s = self.p, self.setup
        steps = p.examples_per_epoch * s.epochs_per_decay / s.batch_size
        if p.decay_steps_divides_by_workers:
            steps = steps / self.num_workers()
        return max(1, int(steps))

    def learning_rate(self, global_step: int) -> float:
        s = self.setup
        lr0 = s.initial_lr * (self.num_workers() if (s.mode == "bsp" and self.p.lr_times_workers_in_bsp) else 1)
        return exponential_decay(lr0, global_step, self.decay_steps(), s.decay_factor, staircase=True)

    def l2_term(self) -> torch.Tensor:
        """wd * sum(w^2)/2 over the decayed range (the loss term the reference adds explicitly)."""
        L = self.layout
        if not L.weight_decay or L.wd_boundary == 0:
            return torch.zeros((), device=self.device)
        w = L.params[: L.wd_boundary]
        return 0.5 * L.weight_decay * torch.dot(w, w)

    def prepare_inputs(self, images: torch.Tensor) -> torch.Tensor:
        """NHWC fp32 on the compute device (the native stem conv pads 3 -> 4 channels internally)."""
        return images.to(self.device, non_blocking=True)

    # ---- one step --------------------------------------------------------------------------------
    def _fwd_bwd(self, x, y):
        native = self.device.type == "cuda" and Fn.get_backend() != "torch"
        if native:
            from ..ops import conv_native, native_ops
            native_ops.begin_step(self.device)
            native_ops.advance_dropout_seed(self.device)
            if self._wt_batch is None:
                self._wt_batch = conv_native.WeightTransposeBatch()
            conv_native.set_wt_batch(self._wt_batch)      # dgrad operands of all layers: one launch per step
            if hasattr(self.comm, "grad_ready"):
                native_ops.set_grad_listener(self.comm.grad_ready)
        # --trace: device-timed phase rows (only outside graph capture: a captured step is one replay)
        tr = getattr(self, "trace", None)
        if tr is not None and torch.cuda.is_current_stream_capturing():
            tr = None
        try:
            m0 = tr.mark() if tr else None
            if tr:
                torch.cuda.nvtx.range_push("forward")      # NVTX ranges for external profilers, same phases
            self.comm.begin_step()                 # zero the gradients (bucketed push: behind the previous step's flags)
            loss, logits = self._run_forward(x, y, True)
            total = loss
            if not self.setup.fused_l2:
                total = loss + self.l2_term()
            m1 = tr.mark() if tr else None
            if tr:
                torch.cuda.nvtx.range_pop()
                torch.cuda.nvtx.range_push("backward")
            self.comm.before_backward()
            total.backward()
            if tr:
                torch.cuda.nvtx.range_pop()
            if native:
                native_ops.join_wgrad_stream(self.device)       # weight gradients issued on their side stream
            m2 = tr.mark() if tr else None
            self.comm.end_step()                   # bucketed push: join the push stream
            if tr:
                tr.add_device("forward", m0, m1, step=self.local_step)
                tr.add_device("backward (+ overlapped push)", m1, m2, step=self.local_step)
                tr.add_device("push tail / join", m2, tr.mark(), step=self.local_step)
        finally:
            if native:
                conv_native.set_wt_batch(None)
                native_ops.set_grad_listener(None)
                native_ops.nat.lib().dtf_reserve_sms(0)      # (a push transport may have kept SMs out of the backward grids)
        return loss.detach(), logits.detach()

    def _fwd_bwd_graphed(self, x, y):
        if self._graph is None:
            self._static = (torch.empty_like(x), torch.empty_like(y))
            self._static[0].copy_(x)
            self._static[1].copy_(y)
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            snap = self._snapshot_side_state()
            self.comm.push_enabled = False               # warm-up passes must not update the parameters
            with torch.cuda.stream(s):
                for _ in range(2):                       # warm-up outside capture (allocator, tensor maps)
                    self._fwd_bwd(*self._static)
                self._undo_warmup_side_effects(snap)
            self.comm.push_enabled = True
            torch.cuda.current_stream().wait_stream(s)
            from .. import _native as nat
            before = nat.launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._out = self._fwd_bwd(*self._static)
            self.graph_kernel_launches = nat.launch_count() - before
            self._graph = g
        self._static[0].copy_(x, non_blocking=True)
        self._static[1].copy_(y, non_blocking=True)
        self._graph.replay()
        return self._out

    def _snapshot_side_state(self):
        """Everything a forward/backward pass changes besides the gradients: batch-norm moving statistics,
        quantisation ranges (non-trainable variables) and the dropout seed."""
        snap = [(v.tensor, v.tensor.detach().clone()) for v in self.store.all() if not v.trainable]
        if self.device.type == "cuda" and Fn.get_backend() != "torch":
            from ..ops import native_ops
            seed = native_ops.dropout_seed(self.device)
            snap.append((seed, seed.clone()))
        return snap

    @torch.no_grad()
    def _undo_warmup_side_effects(self, snap):
        """The warm-up passes before the graph capture ran the real step body: put back what they moved, so the
        first captured step starts from the same state as an eager first step would."""
        for t, saved in snap:
            t.copy_(saved)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor, batch_size: Optional[int] = None) -> Dict:
        """One worker step.  ``batch_size``: this step's actual batch (gradient re-weighting hook,
        cnn/cifar10_cnn_bsp.py:47,66)."""
        s = self.setup
        x = self.prepare_inputs(images)
        gs_before = self.comm.global_step()
        lr = self.learning_rate(gs_before)
        if s.quantize:
            self.store.quant.set_step(gs_before)
        scale = 1.0 if batch_size is None else float(batch_size) / float(s.batch_size)
        self.comm.prepare_step(lr, scale)          # transports that push from inside the backward pass
        use_graph = s.cuda_graph and self.device.type == "cuda" and x.shape[0] == s.batch_size
        loss, logits = (self._fwd_bwd_graphed if use_graph else self._fwd_bwd)(x, labels)
        if s.mode == "bsp":
            gs = self.comm.bsp_step(lr, scale)
        else:
            gs = self.comm.async_step(lr, scale)
        self.local_step += 1
        waited = 0.0
        if s.mode == "ssp":
            waited = self.comm.ssp_wait(self.local_step, s.staleness)
        return {"loss": loss, "logits": logits, "global_step": gs, "lr": lr, "ssp_wait_s": waited}

    @torch.no_grad()
    def evaluate_batch(self, images: torch.Tensor, labels: torch.Tensor, training_bn: bool = False):
        x = self.prepare_inputs(images)
        self.comm.sync_replica()
        loss, logits = self._run_forward(x, labels, training_bn)
        return loss, logits
