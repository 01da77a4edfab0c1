"""Cluster launcher — the role of the reference's ``train.sh`` / ``clear.sh``.

    export ps='h1 h2' workers='h3 h4 h5'
    bash train.sh <model> <bsp|asp|ssp>            # model: cnn cifarnet alexnet vgg resnet inception
    bash clear.sh

Same contract as the reference: hosts come from the ``ps`` / ``workers`` environment variables
(space-separated); PS tasks get port 22222 and workers 22223 (train.sh:15,23); one process per PS task and
per worker, ``--job_name/--task_id/--ps_hosts/--worker_hosts`` on the command line (train.sh:49,56); PS
processes run with ``CUDA_VISIBLE_DEVICES=`` (train.sh:33); stdout+stderr of each process go to
``ps_<i>.log`` / ``worker_<j>.log`` and the evaluator started next to worker 0 after 20 s goes to
``model_eval.log`` (train.sh:57-59).  All 18 model x mode pairs are accepted (the reference ships 9), and so is any
network name of the model-zoo registry (``bash train.sh mobilenet_v1 bsp``), which runs the generic trainer.

Differences, on purpose: processes on ``localhost`` are spawned directly instead of through ssh; workers on
the same host get distinct ports and, on a GPU box, one GPU each (``LOCAL_RANK``) with a shared
``MASTER_ADDR/PORT`` rendezvous; the eval script name is mapped per model (``cnn`` -> ``cifar10_eval``, the
reference's launcher gets this wrong, Q14) and ``vgg`` gets an evaluator too; every PID is recorded in
``<log_dir>/dtf_pids.txt`` so that ``clear.sh`` kills exactly what was started — on workers *and* PS hosts
(the reference's ``pkill python`` skips the PS hosts, Q15, and would take down unrelated programs).
"""
from __future__ import annotations

import argparse
import os
import shlex
import signal
import subprocess
import sys
import time
from typing import Dict, List

from . import config as cfg

PKG = "distributed_tensorflow_models_b200"
TRAINER = {"cnn": "cifar10_cnn", "cifarnet": "cifar10_cifarnet", "alexnet": "cifar10_alexnet", "vgg": "cifar10_vgg",
           "resnet": "cifar10_resnet", "inception": "imagenet_inception"}
EVALUATOR = {"cnn": "cifar10_eval", "cifarnet": "cifar10_cifarnet_eval", "alexnet": "cifar10_alexnet_eval",
             "vgg": "cifar10_vgg_eval", "resnet": "cifar10_resnet_eval", "inception": "inception_eval"}
LOCAL = ("localhost", "127.0.0.1", "")
PIDFILE = "dtf_pids.txt"


def _is_local(host: str) -> bool:
    return host in LOCAL or host == os.uname().nodename


def host_lists(ps: List[str], workers: List[str]):
    """'h' -> 'h:22222' / 'h:22223'; repeated hosts get consecutive ports so several tasks can share a machine."""
    def assign(hosts, base):
        seen: Dict[str, int] = {}
        out = []
        for h in hosts:
            k = seen.get(h, 0)
            seen[h] = k + 1
            out.append(f"{h}:{base + 10 * k}")
        return out
    return assign(ps, cfg.PS_PORT), assign(workers, cfg.WORKER_PORT)


def _spawn(host: str, cmd: List[str], env: Dict[str, str], log_path: str, delay: float = 0.0) -> subprocess.Popen:
    logf = open(log_path, "w")
    if _is_local(host):
        full_env = dict(os.environ)
        full_env.update(env)
        if delay:
            cmd = ["bash", "-c", f"sleep {delay}; exec " + " ".join(shlex.quote(c) for c in cmd)]
        return subprocess.Popen(cmd, stdout=logf, stderr=subprocess.STDOUT, env=full_env, start_new_session=True)
    envs = " ".join(f"{k}={shlex.quote(v)}" for k, v in env.items())
    remote = (f"sleep {delay}; " if delay else "") + f"cd {shlex.quote(os.getcwd())} && {envs} " + \
        " ".join(shlex.quote(c) for c in cmd)
    return subprocess.Popen(["ssh", host, remote], stdout=logf, stderr=subprocess.STDOUT, start_new_session=True)


def local_thread_shares(ps: List[str], workers: List[str], cores: int = 0):
    """-> (env for a local PS process, env for a local worker): the cores of this host divided among the processes
    started on it.  Every process would otherwise open an OpenMP pool as wide as the machine; with 1 PS + 2 workers on
    an 8-thread host the pools fight each other and a CPU step takes 6x longer than alone.  A PS applies elementwise
    updates and gets one thread; the workers share the rest.  Nothing is set when the user already exported
    ``OMP_NUM_THREADS``."""
    if os.environ.get("OMP_NUM_THREADS"):
        return {}, {}
    cores = cores or os.cpu_count() or 1
    n_ps = sum(1 for h in ps if _is_local(h))
    n_w = sum(1 for h in workers if _is_local(h))
    if n_ps + n_w <= 1:
        return {}, {}
    per_worker = max(1, (cores - n_ps) // max(1, n_w))

    def env(n):
        return {"OMP_NUM_THREADS": str(n), "MKL_NUM_THREADS": str(n)}
    return env(1), env(per_worker)


def launch(model: str, mode: str, ps: List[str], workers: List[str], log_dir: str, extra: List[str],
           with_eval: bool = True, eval_delay: float = 20.0, python: str = sys.executable) -> List[subprocess.Popen]:
    os.makedirs(log_dir, exist_ok=True)
    ps_hosts, worker_hosts = host_lists(ps, workers)
    common = [f"--ps_hosts={','.join(ps_hosts)}", f"--worker_hosts={','.join(worker_hosts)}"] + extra
    if model in TRAINER:
        mod_args = [f"{PKG}.trainers.{TRAINER[model]}_{mode}"]
    else:       # any other network of the model zoo: the generic trainer, selected by registry name
        mod_args = [f"{PKG}.trainers.train", f"--model={model}", f"--mode={mode}"]
    procs: List[subprocess.Popen] = []
    master_port = str(cfg.WORKER_PORT + 1000)
    ps_threads, worker_threads = local_thread_shares(ps, workers)
    for i, h in enumerate(ps):
        cmd = [python, "-m"] + mod_args + ["--job_name=ps", f"--task_id={i}"] + common
        env = {"CUDA_VISIBLE_DEVICES": ""}
        if _is_local(h):
            env.update(ps_threads)
        procs.append(_spawn(h, cmd, env, os.path.join(log_dir, f"ps_{i}.log")))
    local_slot = 0
    for j, h in enumerate(workers):
        env = {"RANK": str(j), "WORLD_SIZE": str(len(workers)), "MASTER_PORT": master_port,
               "MASTER_ADDR": "127.0.0.1" if _is_local(workers[0]) else workers[0]}
        if _is_local(h):
            env.update(worker_threads)
        if _is_local(h):
            env["LOCAL_RANK"] = str(local_slot)
            local_slot += 1
        cmd = [python, "-m"] + mod_args + ["--job_name=worker", f"--task_id={j}"] + common
        procs.append(_spawn(h, cmd, env, os.path.join(log_dir, f"worker_{j}.log")))
        if j == 0 and with_eval and model in EVALUATOR:
            ecmd = [python, "-m", f"{PKG}.eval.{EVALUATOR[model]}"] + [a for a in extra if a.startswith(
                ("--train_dir", "--data_dir", "--checkpoint_dir", "--eval_dir", "--run_once", "--eval_interval_secs",
                 "--num_examples", "--synthetic", "--resnet_size", "--batch_size"))]
            procs.append(_spawn(h, ecmd, {"CUDA_VISIBLE_DEVICES": ""}, os.path.join(log_dir, "model_eval.log"),
                                delay=eval_delay))
    with open(os.path.join(log_dir, PIDFILE), "a") as f:
        for p in procs:
            f.write(f"{p.pid}\n")
    return procs


def clear(log_dir: str, ps: List[str], workers: List[str]) -> int:
    """Kill exactly the process groups this launcher started (local pidfile), then do the same on every
    remote worker *and* PS host."""
    n = 0
    path = os.path.join(log_dir, PIDFILE)
    if os.path.exists(path):
        for line in open(path):
            try:
                pid = int(line.strip())
                os.killpg(pid, signal.SIGTERM)
                n += 1
            except (ValueError, ProcessLookupError, PermissionError):
                pass
        os.remove(path)
    for h in sorted(set(ps + workers)):
        if not _is_local(h):
            subprocess.call(["ssh", h, f"cd {shlex.quote(os.getcwd())} && {sys.executable} -m {PKG}.launcher --clear "
                                      f"--log_dir {shlex.quote(log_dir)}"])
    return n


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(add_help=True)
    ap.add_argument("model", nargs="?")
    ap.add_argument("mode", nargs="?")
    ap.add_argument("--clear", action="store_true")
    ap.add_argument("--log_dir", default=os.environ.get("DTF_LOG_DIR", os.getcwd()))
    ap.add_argument("--no_eval", action="store_true")
    ap.add_argument("--eval_delay", type=float, default=20.0)
    ap.add_argument("--wait", action="store_true", help="wait for the workers and return their worst exit code")
    args, extra = ap.parse_known_args(argv)
    ps = os.environ.get("ps", "").split()
    workers = os.environ.get("workers", "").split()
    if args.clear:
        print(f"killed {clear(args.log_dir, ps, workers)} process group(s)")
        return 0
    if not args.model or not args.mode:
        print("usage: export ps='h1 ...' workers='h2 h3 ...'; bash train.sh <model> <bsp|asp|ssp>   "
              f"(models: {' '.join(cfg.MODELS)})")
        return 1
    known = args.model in cfg.MODELS
    if not known:
        from .models import registry
        known = registry.is_zoo_model(args.model)
    if not known or args.mode not in cfg.MODES:
        print(f"unknown model/mode {args.model}/{args.mode}")
        return 1
    if not workers:
        print("the `workers` environment variable is empty (space-separated host names)")
        return 1
    procs = launch(args.model, args.mode, ps, workers, args.log_dir, extra, with_eval=not args.no_eval,
                   eval_delay=args.eval_delay)
    print(f"started {len(procs)} processes; logs in {args.log_dir} (ps_<i>.log, worker_<j>.log, model_eval.log)")
    if args.wait:
        rc = 0
        n_ps = len(ps)
        for p in procs[n_ps:n_ps + len(workers) + 1]:
            rc = max(rc, p.wait() or 0)
        for p in procs:
            if p.poll() is None:
                try:
                    p.wait(timeout=30)
                except subprocess.TimeoutExpired:
                    os.killpg(p.pid, signal.SIGTERM)
        return rc
    return 0


if __name__ == "__main__":
    sys.exit(main())
