"""Product transport on >= 2 GPUs: `SymmComm` (peer-mapped HBM + our push/update/pull, async-apply and clock
kernels over NVLink) must reproduce the NCCL reduce-scatter/all-gather baseline for BSP, count one global step
per worker step for ASP, and respect the SSP bound.  Runs `scripts/multi_gpu_check.py` under torchrun."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_symm_transport_matches_nccl_baseline_and_async_modes():
    n = min(torch.cuda.device_count(), 8)
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.join(ROOT, "scripts", "multi_gpu_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, (r.stdout[-2000:], r.stderr[-2000:])
    res = json.loads(lines[-1])
    assert res["ok"], res
    for case in ("bsp_cnn_sgd_ema_graph", "bsp_cnn_rmsprop_eager", "bsp_cnn_rmsprop_no_overlap"):
        assert res[case]["param_rel_diff"] < 1e-4 and res[case]["replicas_identical"], res[case]
        assert res[case]["buckets"] >= 1
    assert res["bsp_cnn_sgd_ema_graph"]["buckets"] > 1          # the overlapped path really ran bucket by bucket
    assert res["asp"]["global_step_on_rank0"] == res["asp"]["expected"]
    assert res["ssp"]["max_lead_after_release"] <= res["ssp"]["bound"]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_flag_after_data_ordering_litmus():
    """No reader that has seen an owner's `updated` flag ever reads a value of the previous epoch
    (scripts/litmus_push_order.py: multimem.st / P2P st, fence, st.release.sys against ld.acquire.sys + loads)."""
    n = 2 if torch.cuda.device_count() < 8 else 8
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "scripts", "litmus_push_order.py"), "200", "cifarnet"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, (r.stdout[-2000:], r.stderr[-2000:])
    res = json.loads(lines[-1])
    assert res["ok"], res
