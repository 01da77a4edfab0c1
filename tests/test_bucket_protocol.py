"""Host-side logic of the bucketed push (parallel/symm.py, parallel/layout.py) and the protocol checker of the fake
backend (parallel/protocol_model.py) — everything here runs without a GPU."""
import pytest
import torch

from distributed_tensorflow_models_b200.parallel.layout import Bucket, Entry, plan_buckets
from distributed_tensorflow_models_b200.parallel.protocol_model import BucketProtocolModel
from distributed_tensorflow_models_b200.parallel.symm import SymmComm


def _entries(sizes):
    ents, off = [], 0
    for i, n in enumerate(sizes):
        ents.append(Entry(f"v{i}", (n,), n, off, 0.0, i))
        off += (n + 31) // 32 * 32
    return ents, off


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_plan_buckets_covers_the_buffer_and_splits_evenly(world):
    sizes = [432, 16, 16, 2304, 2304, 36864, 73728, 147456, 100000, 64, 10]
    ents, used = _entries(sizes)
    total = (used + world * 32 - 1) // (world * 32) * (world * 32)
    bks = plan_buckets(ents, total, world, target_floats=60000, head_floats=2048)
    assert bks[0].lo == 0 and bks[-1].hi == total
    assert all(a.hi == b.lo for a, b in zip(bks, bks[1:]))                   # contiguous, no gaps
    assert sorted(n for b in bks for n in b.names) == sorted(e.name for e in ents)    # every tensor in one bucket
    starts = {e.offset for e in ents}
    assert all(b.lo in starts for b in bks)                                    # cuts at tensor starts only
    assert bks[0].length < 60000                                               # the head bucket is small
    slot = 0
    for b in bks:
        assert b.chunk % 4 == 0 and b.chunk * world >= b.length and b.slot_off == slot
        parts = [b.part(r) for r in range(world)]
        assert parts[0][0] == b.lo and parts[-1][1] == b.hi
        assert all(p[1] == q[0] for p, q in zip(parts, parts[1:]))            # the owners' parts tile the bucket
        slot += b.chunk


def _fake_comm(world, rank, buckets, total):
    c = object.__new__(SymmComm)
    c.world, c.rank, c.buckets, c.bucketed = world, rank, buckets, True
    c.n_slot = sum(b.chunk for b in buckets)

    class L:
        pass
    c.layout = L()
    c.layout.total = total
    return c


def test_bucket_ownership_round_trip():
    """gather_slot_flats / load_slot_flats: [W, S] slot arrays <-> full-length flat tensor in parameter order."""
    world = 3
    ents, used = _entries([1000, 5000, 64, 7000, 300])
    total = (used + 95) // 96 * 96
    bks = plan_buckets(ents, total, world, target_floats=4000)
    flat = torch.arange(total, dtype=torch.float32)
    rows = []
    for r in range(world):
        c = _fake_comm(world, r, bks, total)
        dst = torch.full((c.n_slot,), -1.0)
        c._owned_copy(flat, dst)
        rows.append(dst)
    back = _fake_comm(world, 0, bks, total)._unshard(torch.stack(rows))
    assert torch.equal(back, flat)


def test_buckets_are_issued_in_the_agreed_order_whatever_order_they_complete_in():
    world = 2
    ents, used = _entries([64] * 6)
    total = 384
    bks = plan_buckets(ents, total, world, target_floats=128)
    assert len(bks) == 3
    c = _fake_comm(world, 0, bks, total)
    grads = torch.zeros(total)
    c.layout.grads = grads
    params = []
    for e in ents:
        p = torch.nn.Parameter(torch.zeros(e.numel))
        p.grad = grads[e.offset:e.offset + e.numel]
        params.append(p)
    c._bucket_lo = [b.lo for b in bks]
    c._learned, c._in_step, c.push_enabled, c.overlap, c.inline = False, True, True, True, False
    c._need, c._need_total, c._order = [], [], None
    c._reset_step_state()
    launched = []
    c._launch_bucket = lambda b, stream=None: launched.append(b)

    class _Ev:
        def record(self, s):
            pass

    class _Stream:
        cuda_stream = 7

        def wait_event(self, ev):
            pass
    import distributed_tensorflow_models_b200.parallel.symm as symm
    real_event, real_cur = symm.torch.cuda.Event, symm.torch.cuda.current_stream
    symm.torch.cuda.Event, symm.torch.cuda.current_stream = _Ev, lambda: _Stream()
    c.comm_stream = _Stream()
    try:
        # learning pass: parameters report in backward order (last tensor first); bucket 2 completes first
        for p in reversed(params):
            c.grad_ready(p)
        c.dist = None
        c.world = 1                                   # (no broadcast of the order in this single-process test)
        c._finish_learning()
        assert c._order == [2, 1, 0]
        # a later step in which bucket 0 happens to complete before bucket 1: nothing overtakes the agreed order
        c._reset_step_state()
        for i in (5, 4, 0, 1):
            c.grad_ready(params[i])
        assert launched == [2]
        for i in (3, 2):
            c.grad_ready(params[i])
        assert launched == [2, 1, 0]
    finally:
        symm.torch.cuda.Event, symm.torch.cuda.current_stream = real_event, real_cur


@pytest.mark.parametrize("world,buckets", [(2, 1), (3, 4), (8, 3)])
def test_flag_protocol_is_clean_under_random_interleavings(world, buckets):
    for seed in range(3):
        errs = BucketProtocolModel(world, buckets, steps=6, seed=seed).run()
        assert errs == [], errs


@pytest.mark.parametrize("rule,needle", [
    ("zero_without_gate", "zeros"),               # gradients zeroed while an owner of the previous step still reads them
    ("forward_without_gate", "forward of step"),  # stale parameters consumed
    ("reduce_without_ready", "reduces bucket"),   # an owner reads a gradient that is not complete yet
])
def test_the_checker_catches_each_race_when_its_rule_is_dropped(rule, needle):
    found = []
    for seed in range(24):
        found += BucketProtocolModel(4, 3, steps=8, seed=seed, jitter=5e-4, break_rule=rule).run(timeout=20)
        if found:
            break
    assert found and any(needle in m for m in found), found


def test_committed_sass_listings_match_the_sources():
    """profiles/sass/*.sass are evidence: each carries the hash of the source it was generated from, and this test
    fails when a kernel was edited without `python scripts/regen_sass.py` (round 1 shipped a listing of the push
    kernel that predated its NVLS rewrite)."""
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "scripts" / "regen_sass.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.parametrize("model,world,jitter", [("cnn", 2, 0.0), ("cnn", 3, 2e-3), ("resnet", 2, 1e-3)])
def test_bucketed_push_on_the_thread_cluster_equals_whole_buffer_bsp(model, world, jitter):
    """The engine's step hooks (pull gates at variable look-up, per-bucket push from post-accumulate hooks, gated
    zeroing) drive the CPU twin of the bucketed protocol with real numbers: parameters, slots and EMA shadows after
    a few steps equal whole-buffer BSP on the same data, under random interleavings."""
    import dataclasses

    from distributed_tensorflow_models_b200 import config as cfg
    from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
    from distributed_tensorflow_models_b200.parallel.comm import BucketedThreadComm, ThreadCluster
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    preset = None
    if model == "cnn":
        preset = dataclasses.replace(cfg.PRESETS[model], optimizer="rmsprop", initial_lr=0.01, ema_decay=0.9999)
    sp = MODEL_SPECS[model]
    bs = 4

    def data(rank):
        g = torch.Generator().manual_seed(50 + rank)
        return torch.randn(bs, sp.image_size, sp.image_size, 3, generator=g), torch.randint(0, 10, (bs,), generator=g)

    def run(make_comm):
        def work(c):
            comm = make_comm(c)
            t = Trainer(TrainSetup(model=model, batch_size=bs, num_workers=world, preset=preset, seed=0,
                                   resnet_size=8), comm)
            x, y = data(comm.rank)
            losses = [float(t.train_step(x, y)["loss"]) for _ in range(4)]
            comm.sync_replica()
            comm.barrier()
            slots = {k: v.clone() for k, v in comm.gather_slot_flats().items()}
            return losses, t.layout.params[: t.layout.used].clone(), slots, t.layout
        return ThreadCluster(world).run(work)

    ref = run(lambda c: c)
    got = run(lambda c: BucketedThreadComm(c.c, c.rank, bucket_floats=1 << 14, head_floats=512, jitter=jitter,
                                           seed=3))
    layout = ref[0][3]
    live = torch.zeros(layout.total, dtype=torch.bool)
    for e in layout.entries:
        live[e.offset:e.offset + e.numel] = True
    for (l0, p0, s0, _), (l1, p1, s1, _) in zip(ref, got):
        assert l0 == pytest.approx(l1, rel=1e-5)
        assert torch.allclose(p0, p1, rtol=1e-5, atol=1e-7)
        assert set(s0) == set(s1)
        for k in s0:
            assert torch.allclose(s0[k][live], s1[k][live], rtol=1e-5, atol=1e-7), k
    assert all(torch.equal(got[0][1], g[1]) for g in got[1:])          # replicas identical across ranks


def test_the_thread_twin_shows_what_the_pull_gate_is_for():
    """Same run with a slow owner: with the gates at variable look-up the result is exactly whole-buffer BSP; without
    them a fast worker's next forward pass reads the slow owner's part of a bucket before it has landed."""
    from distributed_tensorflow_models_b200.parallel.comm import BucketedThreadComm, ThreadCluster
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup

    def data(rank):
        g = torch.Generator().manual_seed(50 + rank)
        return torch.randn(4, 32, 32, 3, generator=g), torch.randint(0, 10, (4,), generator=g)

    def run(make):
        def work(c):
            comm = make(c)
            t = Trainer(TrainSetup(model="resnet", resnet_size=8, batch_size=4, num_workers=2, seed=0), comm)
            x, y = data(comm.rank)
            for _ in range(3):
                t.train_step(x, y)
            comm.sync_replica()
            comm.barrier()
            return t.layout.params[: t.layout.used].clone()
        return ThreadCluster(2).run(work)

    class NoGate(BucketedThreadComm):
        def on_fetch(self, var):
            pass

    ref = run(lambda c: c)
    kw = dict(bucket_floats=1 << 12, head_floats=512, owner_delay=0.05)
    good = run(lambda c: BucketedThreadComm(c.c, c.rank, **kw))
    assert max(float((a - b).abs().max()) for a, b in zip(ref, good)) < 1e-6
    worst = 0.0
    for delay in (0.05, 0.25):          # (a second, slower owner if a loaded machine hid the race the first time)
        kw["owner_delay"] = delay
        bad = run(lambda c: NoGate(c.c, c.rank, **kw))
        worst = max(float((a - b).abs().max()) for a, b in zip(ref, bad))
        if worst > 1e-4:
            break
    assert worst > 1e-4


def test_symm_transport_refuses_reference_placements_with_a_readable_message():
    """`--placement round_robin|axis0` belongs to the transports that keep whole tensors on parameter servers; the
    NVLink transport says so instead of tripping an assert."""
    from distributed_tensorflow_models_b200.parallel.layout import FlatLayout
    from distributed_tensorflow_models_b200.slim import VariableStore
    from distributed_tensorflow_models_b200.slim import variables as V
    st = VariableStore()
    with st.as_default():
        st.variable("a/weights", (6, 3, 3, 4), V.truncated_normal_initializer(0.1))
        st.variable("b/weights", (5, 6), V.truncated_normal_initializer(0.1))
    layout = FlatLayout(st, world_size=2, placement="round_robin")
    comm = object.__new__(SymmComm)
    comm.world, comm.rank = 2, 0
    with pytest.raises(ValueError, match="--placement round_robin"):
        comm.setup(layout, None)


def test_concat_view_function_hands_each_branch_its_gradient_slice():
    """K11 host side: branches that wrote channel slices of one buffer are 'concatenated' for free, and the backward
    pass gives every branch the matching slice of the incoming gradient as a view (no copy)."""
    from distributed_tensorflow_models_b200.ops import native_ops
    buf = torch.zeros(2, 3, 3, 10)
    a = torch.randn(2, 3, 3, 4, requires_grad=True)
    b = torch.randn(2, 3, 3, 6, requires_grad=True)

    class WriteInto(torch.autograd.Function):          # stands in for a kernel that takes the pending output slot
        @staticmethod
        def forward(ctx, x):
            slot = native_ops.take_output_slot(x.shape)
            slot.data.copy_(x)            # like a native kernel: writes through the pointer, no version bump
            return slot

        @staticmethod
        def backward(ctx, g):
            assert g.stride(-1) == 1 and g.stride(2) == 10      # a strided view of the block's gradient
            return g.clone()
    native_ops.set_output_slot(buf[..., :4])
    oa = WriteInto.apply(a)
    native_ops.set_output_slot(buf[..., 4:])
    ob = WriteInto.apply(b)
    y = native_ops.concat_views(buf, [oa, ob])
    assert y.shape == buf.shape and torch.equal(y[..., :4], a.detach()) and torch.equal(y[..., 4:], b.detach())
    w = torch.randn_like(buf)
    (y * w).sum().backward()
    assert torch.equal(a.grad, w[..., :4]) and torch.equal(b.grad, w[..., 4:])
    # output-slot hand-over: taken only by an op whose output has exactly the slot's shape, and only once
    native_ops.set_output_slot(buf[..., :4])
    assert native_ops.take_output_slot((2, 3, 3, 6)) is None and native_ops.output_slot_pending()
    assert native_ops.take_output_slot((2, 3, 3, 4)) is not None and not native_ops.output_slot_pending()


def test_tf32_operand_split_is_exact():
    """--precision tf32x3: a = hi + lo exactly, hi has no bits below TF32's 10-bit mantissa, lo is at most 2^-11 |a|."""
    from distributed_tensorflow_models_b200.ops import conv_native as cn
    x = torch.randn(4096) * torch.logspace(-6, 6, 4096)
    hi, lo = cn._split_tf32(x)
    assert torch.equal(hi + lo, x)
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert bool((lo.abs() <= x.abs() * 2.0 ** -10).all())
    with pytest.raises(ValueError):
        cn.set_precision("fp16")
