"""The trainer program behind every ``<model> x <mode>`` entry point.

One parametrised implementation of the skeleton that each reference trainer script repeats (SURVEY §2.2,
call stack §3.2): parse ``--ps_hosts/--worker_hosts/--job_name/--task_id`` -> bring up the cluster -> the
chief (worker 0) wipes ``train_dir`` (CIFAR models) or resumes (Inception) -> build model + PS layout ->
wait for the chief's initial values -> loop ``max_steps`` times: batch, step, log line -> periodic
checkpoints by the chief (``save_model_secs`` 60 / ``save_interval_secs`` 600) -> NaN guard (Inception) ->
train-batch accuracy every 200 steps (ResNet).

Roles and transports
--------------------
* GPU box: workers are SPMD ranks that also own the PS shards (``--backend symm``: our NVLink push/pull
  kernels; ``nccl``: collective baseline).  ``--job_name=ps`` processes have nothing to host and exit at once.
* CPU / multi-host plumbing (``--backend gloo``): dedicated PS processes as in the reference
  (``parallel/ps_process.py``), SSP through the TCP clock service on PS 0's host.
* ``--backend local``: single process.
"""
from __future__ import annotations

import math
import os
import sys
import time
from typing import Optional

import torch

from .. import config as cfg
from ..ckpt import store as ckpt
from ..data.synthetic import SyntheticBatches
from ..models.registry import MODEL_SPECS
from ..ops import functional as Fn
from ..parallel.comm import Comm, LocalComm
from ..parallel.engine import Trainer, TrainSetup
from ..utils.logs import ChromeTrace, MetricsWriter, format_step, log
from ..utils.pipeline import DevicePrefetcher, ScalarReadback
from ..utils.tfevents import EventFileWriter

GLOBAL_STEP_NAME = {"cnn": "Variable", "cifarnet": "partitioned_space/Variable",
                    "alexnet": "partitioned_space/Variable", "vgg": "root/Variable", "resnet": "Variable",
                    "inception": "global_step"}


def _host_port(hp: str, default_port: int):
    if ":" in hp:
        h, p = hp.rsplit(":", 1)
        return h, int(p)
    return hp, default_port


def resolve_backend(args) -> str:
    if args.backend != "auto":
        return args.backend
    nw = max(1, len(args.worker_host_list))
    if torch.cuda.is_available():
        return "symm" if (nw > 1 or int(os.environ.get("WORLD_SIZE", "1")) > 1) else "local"
    if nw > 1 or args.ps_host_list:
        return "gloo"
    return "local"


def make_setup(args, num_workers: int) -> TrainSetup:
    return TrainSetup(model=args.model, mode=args.mode, batch_size=args.batch_size, num_workers=num_workers,
                      staleness=args.staleness, resnet_size=args.resnet_size, ref_compat=args.ref_compat,
                      placement=args.placement, seed=args.seed, initial_lr=args.initial_learning_rate,
                      epochs_per_decay=args.num_epochs_per_decay, decay_factor=args.learning_rate_decay_factor,
                      cuda_graph=bool(args.cuda_graph), preset=args.preset, image_size=args.image_size,
                      num_classes=args.num_classes, weight_decay=args.weight_decay,
                      quantize=bool(args.quantize), quant_delay=int(args.quant_delay),
                      precision=getattr(args, "precision", "tf32"))


def make_batches(args, spec, rank: int, world: int, device=None):
    """Real data when present (and not ``--synthetic``), synthetic batches of the model's shape otherwise."""
    use_synth = args.synthetic
    if spec.dataset == "cifar10":
        from ..data import cifar10
        if use_synth is None:
            use_synth = not cifar10.maybe_download_and_extract(args.data_dir)
        if not use_synth and getattr(args, "input_pipeline", "host") == "device" and spec.image_size <= 32:
            # records resident on the training device, augmentation as one kernel per batch
            return cifar10.DeviceCifar10Batches(args.data_dir, args.batch_size, spec.image_size,
                                                device=device if device is not None else "cpu", rank=rank,
                                                world=world, seed=args.seed)
        if not use_synth:
            return cifar10.distorted_inputs(args.data_dir, args.batch_size, spec.image_size, rank=rank, world=world,
                                            seed=args.seed)
    else:
        from ..data import imagenet
        if use_synth is None:
            use_synth = not imagenet.ImagenetData(args.subset, args.data_dir).available()
        if not use_synth:
            return imagenet.distorted_inputs(imagenet.ImagenetData(args.subset, args.data_dir), args.batch_size,
                                             image_size=spec.image_size, num_preprocess_threads=args.num_preprocess_threads,
                                             rank=rank, world=world, seed=args.seed)
    if spec.dataset == "cifar10":
        # synthetic 32x32 records run through the real distortion pipeline (random crop to 24 for the plain
        # CNN, flip, brightness, contrast, standardisation) — BASELINE.json config 1 / SURVEY Q10
        from ..data.synthetic import SyntheticCifarBatches
        return SyntheticCifarBatches(args.batch_size, spec.image_size, seed=args.seed + rank)
    return SyntheticBatches(args.batch_size, spec.image_size, spec.num_classes, seed=args.seed + rank)


def _init_gloo(args, backend_rank: int, world: int) -> None:
    import torch.distributed as dist
    host, port = _host_port(args.ps_host_list[0] if args.ps_host_list else args.worker_host_list[0], cfg.PS_PORT)
    if host in ("localhost",):
        host = "127.0.0.1"
    dist.init_process_group("gloo", init_method=f"tcp://{host}:{port + 1000}", rank=backend_rank, world_size=world)


def run_ps(args) -> int:
    """``--job_name=ps``."""
    backend = resolve_backend(args)
    if backend != "gloo":
        log("ps task %d: parameter-server shards live on the worker GPUs in this mode (backend=%s); nothing to host"
            % (args.task_id, backend))
        return 0
    from ..parallel.layout import FlatLayout
    from ..parallel.ps_process import ParameterServer
    P, W = len(args.ps_host_list), len(args.worker_host_list)
    _init_gloo(args, args.task_id, P + W)
    # build the same variables as the workers to learn the layout (values come from the chief)
    t = Trainer(make_setup(args, W), _LayoutOnlyComm(P), torch.device("cpu"))
    host, _ = _host_port(args.ps_host_list[0], cfg.PS_PORT)
    ps = ParameterServer(args.task_id, P, W, t.layout, t.opt_cfg, mode=args.mode, ssp_host="0.0.0.0",
                         ssp_port=cfg.SSP_PORT, staleness=args.staleness)
    log("ps task %d serving %d workers (%s, %s)" % (args.task_id, W, args.mode, t.layout.describe()))
    steps = ps.serve()
    log("ps task %d done after %d updates" % (args.task_id, steps))
    import torch.distributed as dist
    dist.destroy_process_group()
    return 0


class _LayoutOnlyComm(LocalComm):
    """Used by PS processes: they need the model's layout, not a worker transport."""

    def __init__(self, owners: int):
        super().__init__()
        self.owners = owners

    def setup(self, layout, opt_cfg):
        self.layout = layout
        self.opts = []


def make_comm(args, backend: str, num_workers: int):
    """-> (comm, device, rank, world)"""
    if backend == "local":
        dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
        return LocalComm(), dev, 0, 1
    if backend == "gloo":
        from ..parallel.ps_process import PsWorkerComm
        P, W = len(args.ps_host_list), len(args.worker_host_list)
        if P == 0:
            raise SystemExit("--backend gloo needs --ps_hosts (dedicated parameter servers)")
        _init_gloo(args, P + args.task_id, P + W)
        host, _ = _host_port(args.ps_host_list[0], cfg.PS_PORT)
        if host == "localhost":
            host = "127.0.0.1"
        comm = PsWorkerComm(P, W, args.task_id, ssp_host=host, ssp_port=cfg.SSP_PORT, mode=args.mode,
                            staleness=args.staleness)
        return comm, torch.device("cpu"), args.task_id, W
    # GPU SPMD transports
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", str(num_workers)))
    rank = int(os.environ.get("RANK", str(args.task_id)))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank % max(1, torch.cuda.device_count()))))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if not dist.is_initialized():
        if "MASTER_ADDR" not in os.environ:
            host, port = _host_port(args.worker_host_list[0], cfg.WORKER_PORT)
            os.environ["MASTER_ADDR"] = "127.0.0.1" if host == "localhost" else host
            os.environ["MASTER_PORT"] = str(port + 1000)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    if backend == "symm":
        from ..parallel.symm import SymmComm
        return SymmComm(timeout_s=getattr(args, "ps_timeout_secs", None),
                        ssp_timeout_s=getattr(args, "ssp_timeout_secs", None)), dev, rank, world
    from ..parallel.comm import TorchDistComm
    return TorchDistComm("rs_ag"), dev, rank, world


def save_checkpoint(trainer: Trainer, train_dir: str, max_to_keep: int, fmt: str = "npz") -> Optional[str]:
    """Chief-side save (all ranks must call: gathering the sharded slots is collective)."""
    comm = trainer.comm
    comm.sync_replica()                    # the chief's replica must hold every owner's last update
    slots = comm.gather_slot_flats()
    path = None
    if comm.rank == 0:
        gs = comm.global_step()
        tensors = ckpt.collect_tensors(trainer.store, slots, trainer.layout, gs, GLOBAL_STEP_NAME.get(trainer.setup.model, "global_step"))
        path = ckpt.save(train_dir, tensors, gs, max_to_keep, fmt=fmt)
    # nobody runs ahead into the next step's flag waits while the chief is still writing (seconds for a large
    # model): a peer's wait would otherwise count the save against its timeout
    comm.barrier()
    return path


def run_worker(args) -> int:
    backend = resolve_backend(args)
    num_workers = max(1, len(args.worker_host_list)) if backend in ("gloo", "local") else \
        int(os.environ.get("WORLD_SIZE", str(max(1, len(args.worker_host_list)))))
    Fn.set_backend("torch" if args.kernels == "torch" else "auto")
    comm, device, rank, world = make_comm(args, backend, num_workers)
    p = args.preset
    is_chief = rank == 0
    # ---- chief: train_dir policy ----
    if is_chief and p.save_secs is not None:
        if p.wipe_train_dir and not args.resume:
            ckpt.wipe(args.train_dir)
        else:
            os.makedirs(args.train_dir, exist_ok=True)
    trainer = Trainer(make_setup(args, world), comm, device)
    spec = trainer.spec
    log("worker %d/%d on %s: %s, mode=%s backend=%s" % (rank, world, device, trainer.layout.describe(), args.mode, backend))
    # ---- resume ----
    if (args.resume or not p.wipe_train_dir) and p.save_secs is not None:
        latest = ckpt.latest_checkpoint(args.train_dir)
        if latest and ckpt.exists(latest):
            tensors = ckpt.load(latest)
            ckpt.restore_training_state(trainer, tensors)      # variables + RMSProp slots + EMA shadows
            comm.set_global_step(ckpt.step_of(latest))
            log("restored %s" % latest)
    if args.init_checkpoint:
        src = args.init_checkpoint
        if os.path.isdir(src):
            src = ckpt.latest_checkpoint(src)
        if not src or not ckpt.exists(src):
            raise SystemExit("--init_checkpoint: no checkpoint at %r" % args.init_checkpoint)
        ckpt.restore(trainer.store, ckpt.load(src), strict=False)
        comm.reset_ema_to_params()                             # fine-tuning: shadows start at the loaded weights
        comm.set_global_step(0)
        log("initialised from %s (global step reset to 0)" % src)
    batches = DevicePrefetcher(make_batches(args, spec, rank, world, device), device)
    losses_rb = ScalarReadback(device)
    metrics = MetricsWriter(args.metrics_jsonl if is_chief else "")
    # chief: scalar summaries into train_dir every save_summaries_secs (Supervisor's summary thread; the reference
    # passes summary_op=None, so its event files hold global_step/sec only — loss and learning rate are extras)
    events = EventFileWriter(args.train_dir) if (is_chief and p.save_secs is not None and args.save_summaries_secs) \
        else None
    last_summary, last_summary_step = time.time(), 0
    from ..utils.summaries import LossAverages, activation_summaries, variable_summaries
    loss_averages = LossAverages(0.9)              # the chief's moving average of the loss (…_bsp.py:116-120)
    last_batch = None
    trace = ChromeTrace(args.trace, pid=rank, device=device)
    trainer.trace = trace if trace.on_device else None     # eager steps also get forward / backward / push rows
    last_save = time.time()
    t_run0 = time.time()
    for step in range(0, args.max_steps):          # init_global_step := 0
        t0 = time.time()
        images, labels = next(batches)
        t1 = time.time()
        ev0 = trace.mark()
        if events is not None and args.summarize_activations:
            last_batch = (images, labels)
        out = trainer.train_step(images, labels, batch_size=args.batch_size)      # already on the device (prefetcher)
        trace.add_device("train_step", ev0, trace.mark(), t_run0, step=step)
        want_log = (step % max(1, args.log_every) == 0) or step == args.max_steps - 1
        if want_log or p.nan_guard:
            # the loss is read back one step late (after the next step has been queued) so that the device never
            # drains; it is still every logged step's own loss, and the NaN guard (inception…bsp.py:191) sees all
            losses_rb.push(out["loss"], tag=(step, out["global_step"], out["lr"], out["ssp_wait_s"], t0, want_log))
        t2 = time.time()
        trace.add("input", t0 - t_run0, t1 - t0, step=step)
        trace.add("train_step", t1 - t_run0, t2 - t1, step=step)
        while len(losses_rb) > (0 if step == args.max_steps - 1 else 1):
            loss_value, (lstep, lgs, llr, lwait, lt0, llog) = losses_rb.pop()
            if p.nan_guard:
                assert not math.isnan(loss_value), "Model diverged with loss = NaN"
            averaged = loss_averages.update(total_loss=loss_value) if is_chief else {}
            if events is not None and time.time() - last_summary >= args.save_summaries_secs:
                now = time.time()
                scalars = {"global_step/sec": (lstep - last_summary_step) / max(now - last_summary, 1e-9),
                           "learning_rate": llr}
                scalars.update(averaged)           # 'total_loss (raw)' and its moving average 'total_loss'
                histos = None
                if args.summarize_activations and last_batch is not None:
                    # one extra forward pass outside the captured step: its intermediate tensors are readable
                    snap = trainer._snapshot_side_state()          # (moving statistics, dropout seed: put back after)
                    trainer.evaluate_batch(last_batch[0], last_batch[1], training_bn=True)
                    trainer._undo_warmup_side_effects(snap)
                    act_scalars, histos = activation_summaries(trainer.store.end_points)
                    scalars.update(act_scalars)
                    histos.update(variable_summaries(trainer.store))
                events.add_summaries(int(lgs), scalars, histos)
                events.flush()
                last_summary, last_summary_step = now, lstep
            if llog:
                t3 = time.time()
                duration = max((t3 - lt0) / max(1, step - lstep + 1), 1e-9)
                log(format_step(p.log_style, lstep, lgs, loss_value, args.batch_size / duration, duration))
                metrics.write(step=lstep, global_step=lgs, loss=loss_value, lr=llr,
                              examples_per_sec=args.batch_size / duration, sec_per_batch=duration,
                              ssp_wait_s=lwait, time=t3)
        if p.accuracy_every and step % p.accuracy_every == 0:
            acc = float((out["logits"].argmax(1) == labels.to(device)).float().mean())
            log("step %d, train-batch accuracy = %.4f" % (step, acc))
        if p.save_secs is not None and args.save_interval_secs and backend != "gloo":
            # time-based like Supervisor(save_model_secs); every rank takes the same decision from the chief's clock
            due = time.time() - last_save >= args.save_interval_secs
            if world > 1:
                flag = torch.tensor([1.0 if (due and is_chief) else 0.0], device=device)
                comm.broadcast_state([flag], src=0)
                due = bool(flag.item() > 0)
            if due:
                path = save_checkpoint(trainer, args.train_dir, p.max_to_keep, args.ckpt_format)
                last_save = time.time()
                if path:
                    log("saved %s" % path)
    # final checkpoint + shutdown
    if p.save_secs is not None and backend != "gloo":
        path = save_checkpoint(trainer, args.train_dir, p.max_to_keep, args.ckpt_format)
        if path:
            log("saved %s" % path)
    elif p.save_secs is not None and is_chief:
        # dedicated-PS mode: the chief's replica is fresh after its last pull; slots live on the PS processes
        tensors = ckpt.collect_tensors(trainer.store, None, None, comm.global_step(), GLOBAL_STEP_NAME.get(args.model, "global_step"))
        log("saved %s" % ckpt.save(args.train_dir, tensors, comm.global_step(), p.max_to_keep, fmt=args.ckpt_format))
    comm.finish()
    if events is not None:
        events.close()
    trace.save()
    metrics.close()
    if hasattr(batches, "close"):
        batches.close()
    if backend in ("gloo", "symm", "nccl"):
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    return 0


def main(argv=None, model: Optional[str] = None, mode: Optional[str] = None) -> int:
    args = cfg.parse_args(argv, model, mode)
    if args.job_name == "ps":
        return run_ps(args)
    return run_worker(args)


if __name__ == "__main__":
    sys.exit(main())
