"""Parameter-server protocol tests on the in-process fake cluster (threads sharing CPU tensors) and on real
multi-process gloo — the tests the reference never had: BSP = mean of N / one update per global
step, ASP = one global step per worker step, SSP staleness bound as an invariant, placement policies, and
convergence parity of the PS step with a plain single-process SGD run."""
import os
import subprocess
import sys
import time

import pytest
import torch

from distributed_tensorflow_models_b200.optim import EmaConfig, OptimizerConfig, ShardOptimizer, exponential_decay
from distributed_tensorflow_models_b200.optim.schedule import ema_decay
from distributed_tensorflow_models_b200.parallel.comm import LocalComm, ThreadCluster
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.parallel.layout import FlatLayout
from distributed_tensorflow_models_b200.slim import VariableStore
from distributed_tensorflow_models_b200.slim import variables as V

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _data(model, bs, seed):
    from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
    sp = MODEL_SPECS[model]
    g = torch.Generator().manual_seed(seed)
    return torch.randn(bs, sp.image_size, sp.image_size, 3, generator=g), torch.randint(0, sp.num_classes, (bs,), generator=g)


# ---- optimiser oracles ------------------------------------------------------------------------------------------
def test_exponential_decay_staircase_and_ema_decay():
    assert exponential_decay(0.1, 0, 100, 0.1, True) == pytest.approx(0.1)
    assert exponential_decay(0.1, 99, 100, 0.1, True) == pytest.approx(0.1)
    assert exponential_decay(0.1, 100, 100, 0.1, True) == pytest.approx(0.01)
    assert exponential_decay(0.1, 50, 100, 0.1, False) == pytest.approx(0.1 * 0.1 ** 0.5)
    assert ema_decay(0.9999, 0) == pytest.approx(0.1) and ema_decay(0.9999, 10 ** 9) == pytest.approx(0.9999)
    assert ema_decay(0.9, None) == 0.9


def test_tf1_rmsprop_formula_and_sgd_with_decay_and_ema():
    w0 = torch.tensor([1.0, -2.0, 3.0, 0.5])
    g = torch.tensor([0.1, -0.2, 0.3, 0.4])
    # RMSProp as TensorFlow 1.x: ms starts at ONE, epsilon inside the square root
    o = ShardOptimizer(OptimizerConfig(kind="rmsprop", rms_decay=0.9, momentum=0.9, epsilon=1.0), w0.clone(), 0, 4, 0)
    o.apply_reference(g, lr=0.045)
    ms = 0.9 * 1.0 + 0.1 * g * g
    mom = 0.045 * g / torch.sqrt(ms + 1.0)
    assert torch.allclose(o.slots["RMSProp"], ms) and torch.allclose(o.slots["RMSProp_1"], mom)
    assert torch.allclose(o.w, w0 - mom)
    o.apply_reference(g, lr=0.045)
    ms2 = 0.9 * ms + 0.1 * g * g
    mom2 = 0.9 * mom + 0.045 * g / torch.sqrt(ms2 + 1.0)
    assert torch.allclose(o.w, w0 - mom - mom2, atol=1e-7)
    # SGD + L2 on the first two elements only + EMA with num_updates
    o = ShardOptimizer(OptimizerConfig(kind="sgd", weight_decay=0.1, ema=EmaConfig(0.9999)), w0.clone(), 0, 4, 2)
    o.apply_reference(g, lr=0.5, grad_scale=2.0, num_updates=0)
    gs = 2.0 * g
    expect = w0 - 0.5 * (gs + torch.tensor([0.1 * 1.0, 0.1 * -2.0, 0.0, 0.0]))
    assert torch.allclose(o.w, expect)
    assert torch.allclose(o.ema, w0 - (1 - 0.1) * (w0 - expect))          # d = min(.9999, 1/10) at step 0


# ---- layout / placement ---------------------------------------------------------------------------------------------
def _toy_store():
    st = VariableStore(device_fn=V.replica_device_setter(2))
    with st.as_default():
        st.variable("a/weights", (6, 3, 3, 4), V.truncated_normal_initializer(0.1), weight_decay=1e-3)
        st.variable("a/biases", (6,), V.zeros_initializer())
        st.variable("b/weights", (5, 6), V.truncated_normal_initializer(0.1), weight_decay=1e-3)
        st.variable("b/biases", (5,), V.zeros_initializer())
    return st


def test_flat_layout_views_alignment_and_decay_boundary():
    st = _toy_store()
    before = {v.name: v.tensor.detach().clone() for v in st.all()}
    L = FlatLayout(st, world_size=4)
    assert [e.name for e in L.entries] == ["a/weights", "b/weights", "a/biases", "b/biases"]   # decayed first
    assert all(e.offset % 32 == 0 for e in L.entries) and L.total % (4 * 32) == 0
    assert L.wd_boundary == 256 and L.weight_decay == 1e-3
    for v in st.all():
        assert torch.equal(v.tensor, before[v.name])                    # values preserved through re-binding
        assert v.tensor.data_ptr() == L.view(v.name).data_ptr()         # ... and they ARE the flat buffer
        assert v.tensor.grad.data_ptr() == L.view(v.name, L.grads).data_ptr()
    assert [(lo, hi) for lo, hi, _ in L.segments] == [(i * L.total // 4, (i + 1) * L.total // 4) for i in range(4)]


def test_round_robin_and_axis0_placement():
    L = FlatLayout(_toy_store(), world_size=2, placement="round_robin")
    owner = {e.name: [o for lo, hi, o in L.segments if lo == e.offset][0] for e in L.entries}
    assert owner == {"a/weights": 0, "a/biases": 1, "b/weights": 0, "b/biases": 1}     # creation order k mod P
    L = FlatLayout(_toy_store(), world_size=2, placement="axis0")
    e = L.by_name["a/weights"]
    segs = [(lo, hi, o) for lo, hi, o in L.segments if e.offset <= lo < e.offset + e.numel]
    assert [hi - lo for lo, hi, _ in segs][0] == 3 * 36 and len(segs) == 2            # split along the leading axis
    covered = sorted((lo, hi) for lo, hi, _ in L.segments)
    assert covered[0][0] == 0 and covered[-1][1] == L.total
    assert all(a[1] == b[0] for a, b in zip(covered, covered[1:]))                     # exact cover, no overlap
    assert sum(L.owner_bytes()) == L.total * 4


# ---- BSP ------------------------------------------------------------------------------------------------------
def test_bsp_is_mean_of_n_one_update_per_global_step():
    W = 3
    cl = ThreadCluster(W)

    def work(comm):
        t = Trainer(TrainSetup(model="cnn", batch_size=4, num_workers=W, seed=comm.rank), comm)
        p0 = t.layout.params.clone()
        x, y = _data("cnn", 4, 10 + comm.rank)
        out = t.train_step(x, y)
        return p0, t.layout.grads.clone(), t.layout.params.clone(), out["global_step"], out["lr"], t.layout

    res = cl.run(work)
    p0, L = res[0][0], res[0][5]
    assert all(torch.equal(r[0], p0) for r in res)                       # chief's initial values everywhere
    assert all(r[3] == 1 for r in res)                                   # ONE global step for N gradients
    assert all(torch.equal(r[2], res[0][2]) for r in res)                # replicas identical after the pull
    gmean = sum(r[1] for r in res) / W
    lr = res[0][4]
    assert lr == pytest.approx(0.32)                                     # plain CNN: LR not multiplied by N
    expect = p0.clone()
    gm = gmean.clone()
    gm[: L.wd_boundary] += L.weight_decay * p0[: L.wd_boundary]
    expect -= lr * gm
    assert torch.allclose(res[0][2], expect, atol=1e-6)


def test_bsp_lr_scaling_by_workers_and_decay_steps():
    t = Trainer(TrainSetup(model="resnet", batch_size=128, num_workers=8))
    assert t.learning_rate(0) == pytest.approx(0.08 * 8)
    assert t.decay_steps() == int(50000 * 32.0 / 128)
    assert t.learning_rate(t.decay_steps()) == pytest.approx(0.064)
    t = Trainer(TrainSetup(model="resnet", mode="asp", batch_size=128, num_workers=8))
    assert t.learning_rate(0) == pytest.approx(0.08)                     # ASP/SSP: no xN


def test_bsp_two_workers_equals_single_process_on_concatenated_batch():
    """Convergence parity: 2-worker BSP (mean of gradients) == 1 process on the union batch (mean loss)."""
    xa, ya = _data("cnn", 4, 1)
    xb, yb = _data("cnn", 4, 2)
    cl = ThreadCluster(2)

    def work(comm):
        t = Trainer(TrainSetup(model="cnn", batch_size=4, num_workers=2, seed=0), comm)
        x, y = (xa, ya) if comm.rank == 0 else (xb, yb)
        for _ in range(3):
            t.train_step(x, y)
        return t.layout.params[: t.layout.used].clone()

    dist_params = cl.run(work)[0]
    t = Trainer(TrainSetup(model="cnn", batch_size=8, num_workers=1, seed=0), LocalComm())
    x, y = torch.cat([xa, xb]), torch.cat([ya, yb])
    for _ in range(3):
        t.train_step(x, y)
    assert torch.allclose(dist_params, t.layout.params[: t.layout.used], atol=2e-5)


# ---- ASP / SSP ---------------------------------------------------------------------------------------------------
def test_asp_global_step_counts_every_worker_step():
    W, steps = 3, 4
    cl = ThreadCluster(W)

    def work(comm):
        t = Trainer(TrainSetup(model="cnn", mode="asp", batch_size=4, num_workers=W), comm)
        x, y = _data("cnn", 4, comm.rank)
        for _ in range(steps):
            out = t.train_step(x, y)
        return out["global_step"]

    cl.run(work)
    assert cl.gs == W * steps


@pytest.mark.parametrize("bound", [0, 2])
def test_ssp_staleness_bound_is_never_exceeded(bound):
    W, steps = 3, 10
    cl = ThreadCluster(W)
    seen = []

    def work(comm):
        t = Trainer(TrainSetup(model="cnn", mode="ssp", staleness=bound, batch_size=2, num_workers=W), comm)
        x, y = _data("cnn", 2, comm.rank)
        for s in range(steps):
            t.train_step(x, y)
            seen.append(t.local_step - min(cl.clock))                   # lead right after being released
            if comm.rank == W - 1:
                time.sleep(0.01)                                        # a straggler
        return t.local_step

    assert cl.run(work) == [steps] * W
    assert cl.max_lead_seen <= bound
    assert max(seen) <= bound + 1        # +1: another worker may already have advanced past the release point


def test_dynamic_batch_size_rescales_gradient():
    t1 = Trainer(TrainSetup(model="cnn", batch_size=8, seed=3))
    t2 = Trainer(TrainSetup(model="cnn", batch_size=8, seed=3))
    x, y = _data("cnn", 4, 5)
    p0 = t1.layout.params.clone()
    t1.train_step(x, y, batch_size=4)              # half-size batch: gradient weighted by 4/8
    t2.train_step(x, y)
    d1, d2 = t1.layout.params - p0, t2.layout.params - p0
    nd = slice(t1.layout.wd_boundary, None)         # undecayed part: pure gradient
    assert torch.allclose(d1[nd] * 2, d2[nd], atol=1e-7)


# ---- dedicated PS processes over gloo (BASELINE.json config 1) ----------------------------------------------------------
@pytest.mark.parametrize("mode", ["bsp", "ssp"])
def test_train_sh_cnn_one_ps_two_workers_localhost(tmp_path, mode):
    env = dict(os.environ, ps="localhost", workers="localhost localhost", DTF_LOG_DIR=str(tmp_path))
    env["PYTHONPATH"] = ROOT
    cmd = ["bash", os.path.join(ROOT, "train.sh"), "cnn", mode, "--wait", "--no_eval", "--max_steps=4", "--batch_size=4",
           f"--train_dir={tmp_path}/train", "--staleness=1"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    logs = {f: open(os.path.join(tmp_path, f)).read() for f in os.listdir(tmp_path) if f.endswith(".log")}
    assert r.returncode == 0, (r.stdout, r.stderr, logs)
    assert set(logs) == {"ps_0.log", "worker_0.log", "worker_1.log"}
    expected_updates = 4 if mode == "bsp" else 8
    assert f"done after {expected_updates} updates" in logs["ps_0.log"], logs["ps_0.log"]
    for w in ("worker_0.log", "worker_1.log"):
        assert logs[w].count("examples/sec") == 4 and "loss = " in logs[w]
    assert os.path.exists(tmp_path / "train" / "checkpoint")


def test_ssp_command_line_probe(capsys):
    """The manual probe (role of inception/ssp/CheckStaleness-remote) talks to a live service."""
    from distributed_tensorflow_models_b200.parallel import ssp_manager
    mgr = ssp_manager.SspManager(2, 3, port=0)
    server = mgr.create_rpc_server("127.0.0.1")
    server.serve_in_background()
    try:
        assert ssp_manager.main(["-h", "127.0.0.1:%d" % server.port, "check_staleness", "0", "2"]) == 0
        assert capsys.readouterr().out.strip() == "1"
        assert mgr.state.local_step_list == [2, 0]
    finally:
        server.stop()


def test_device_prefetcher_and_scalar_readback_cpu_path():
    """The host pipelining helpers degrade to plain iteration / immediate reads without a GPU."""
    from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher, ScalarReadback
    batches = [(torch.full((2, 3), float(i)), torch.tensor([i, i])) for i in range(5)]
    got = list(DevicePrefetcher(iter(batches), torch.device("cpu")))
    assert len(got) == 5 and all(float(x[0, 0]) == i and int(y[0]) == i for i, (x, y) in enumerate(got))
    lazy = DevicePrefetcher(iter(batches), torch.device("cpu"), lazy=True)
    assert not lazy.started and float(next(lazy)[0][0, 0]) == 0.0 and lazy.started
    rb = ScalarReadback(torch.device("cpu"), depth=2)
    rb.push(torch.tensor(1.5), tag="a")
    rb.push(torch.tensor(2.5), tag="b")
    with pytest.raises(RuntimeError):
        rb.push(torch.tensor(3.5))
    assert rb.pop() == (1.5, "a") and len(rb) == 1 and rb.pop() == (2.5, "b")


def test_branch_streams_fall_back_to_call_order_on_cpu():
    from distributed_tensorflow_models_b200.ops import streams
    order = []

    def thunk(i):
        def run():
            order.append(i)
            return torch.full((1,), float(i))
        return run

    streams.set_branch_streams(True)
    try:
        outs = streams.run_branches([thunk(0), thunk(1), thunk(2)], torch.zeros(1))
        aux = streams.Detached(thunk(7), torch.zeros(1))
        assert [float(o) for o in outs] == [0.0, 1.0, 2.0] and order == [0, 1, 2, 7] and float(aux.join()) == 7.0
    finally:
        streams.set_branch_streams(False)


def test_collective_baseline_arm_emulates_asp_over_gloo(tmp_path):
    """TorchDistComm.async_step (the NCCL / gloo baseline arm of the one-sided push): every worker's own gradient is
    applied as a separate optimiser step, global_step advances by W per round, replicas stay identical — checked
    against the two gradients applied one after the other by hand."""
    script = tmp_path / "asp_gloo.py"
    script.write_text(
        "import os, sys, json, torch, torch.distributed as dist\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "from distributed_tensorflow_models_b200.parallel.comm import TorchDistComm\n"
        "from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "g = torch.Generator().manual_seed(100 + r)\n"
        "x, y = torch.randn(4, 24, 24, 3, generator=g), torch.randint(0, 10, (4,), generator=g)\n"
        "t = Trainer(TrainSetup(model='cnn', mode='asp', batch_size=4, num_workers=w, seed=0), TorchDistComm('rs_ag'))\n"
        "gs = [t.train_step(x, y)['global_step']]\n"
        "sum1 = float(t.layout.params[:t.layout.used].double().sum())\n"
        "gs += [t.train_step(x, y)['global_step'] for _ in range(2)]\n"
        "p = t.layout.params[:t.layout.used]\n"
        "chk = torch.stack([p.double().sum(), p.double().abs().sum()])\n"
        "allc = [torch.zeros_like(chk) for _ in range(w)]\n"
        "dist.all_gather(allc, chk)\n"
        "if r == 0: print(json.dumps({'gs': gs, 'same': all(torch.equal(c, allc[0]) for c in allc), 'sum1': sum1}))\n"
        "dist.destroy_process_group()\n")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29577", str(script)],
                       capture_output=True, text=True, timeout=300, env=env)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, (r.stdout[-1500:], r.stderr[-1500:])
    import json
    res = json.loads(lines[-1])
    assert res["gs"] == [2, 4, 6] and res["same"]

    # oracle for the first round: both workers' gradients taken at the initial parameters, applied one after the other
    t = Trainer(TrainSetup(model="cnn", mode="asp", batch_size=4, num_workers=2, seed=0))
    grads = []
    for r_ in range(2):
        g = torch.Generator().manual_seed(100 + r_)
        x, y = torch.randn(4, 24, 24, 3, generator=g), torch.randint(0, 10, (4,), generator=g)
        t._fwd_bwd(x, y)
        grads.append(t.layout.grads.clone())
    o = t.comm.owned_optimizers()[0]
    for q, gq in enumerate(grads):
        o.apply(gq[o.lo:o.hi], t.learning_rate(0), 1.0, num_updates=q)
    p = t.layout.params[:t.layout.used]
    assert float(p.double().sum()) == pytest.approx(res["sum1"], rel=1e-6)
