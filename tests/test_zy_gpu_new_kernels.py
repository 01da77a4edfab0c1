"""GPU tests of what was added after the last GPU session of round 1 — atrous convolution on the dense kernels,
the fake-quantisation kernels and the quantised training step, the device-resident CIFAR-10 distortion kernel.
Their CPU counterparts are in test_zoo.py, test_quantize.py and test_input_native.py.  The file sorts behind every
other GPU test (and before test_zz_no_alloc.py) on purpose: these have not met a B200 yet, and a run that stops at
the first failure should have been through everything that has."""
import numpy as np
import pytest
import torch

from distributed_tensorflow_models_b200 import _native as nat
from distributed_tensorflow_models_b200.data import cifar10
from distributed_tensorflow_models_b200.ops import functional as Fn

pytestmark = pytest.mark.gpu

TF32_TOL = 5e-3
FP32_TOL = 3e-5


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _both(fn, inputs, tol):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    outs, grads = [], []
    g = torch.Generator(device="cpu").manual_seed(11)
    dy = None
    launches = 0
    for backend in ("auto", "torch"):
        Fn.set_backend(backend)
        try:
            ins = [t.detach().clone().requires_grad_(True) if isinstance(t, torch.Tensor) else t for t in inputs]
            before = nat.launch_count()
            y = fn(*ins)
            if dy is None:
                dy = torch.randn(y.shape, generator=g).to(y.device)
            y.backward(dy)
            if backend == "auto":
                launches = nat.launch_count() - before
            outs.append(y.detach())
            grads.append([t.grad for t in ins if isinstance(t, torch.Tensor)])
        finally:
            Fn.set_backend("auto")
    assert launches >= 1, "the native kernels did not run"
    assert outs[0].shape == outs[1].shape
    assert _rel(outs[0], outs[1]) < tol, f"forward mismatch {_rel(outs[0], outs[1])}"
    for i, (a, b) in enumerate(zip(grads[0], grads[1])):
        assert _rel(a, b) < tol, f"grad {i} mismatch {_rel(a, b)}"


def _r(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("cfg", [
    # n, h, w, c, c_out, k, rate, padding
    (2, 28, 28, 32, 64, 3, 2, "SAME"),       # ResNet / MobileNet output_stride=16: last stage dilated by 2
    (2, 33, 35, 16, 32, 3, 4, "SAME"),       # output_stride=8; extents that are not multiples of the rate
    (2, 21, 23, 16, 16, 3, 3, "VALID"),
])
def test_atrous_conv2d_and_depthwise(cfg):
    """rate > 1: sampling phases stacked into the batch, then the dense tcgen05 / depthwise kernels."""
    n, h, w, c, co, k, rate, pad = cfg
    x, wt, b = _r(n, h, w, c, seed=1), _r(co, k, k, c, seed=2, scale=0.1), _r(co, seed=3)
    _both(lambda x_, w_, b_: Fn.atrous_conv2d(x_, w_, b_, rate, pad), [x, wt, b], TF32_TOL)
    dw = _r(k, k, c, 1, seed=4, scale=0.3)
    _both(lambda x_, w_: Fn.depthwise_conv2d(x_, w_, None, 1, pad, rate=rate), [x, dw], FP32_TOL)


@pytest.mark.parametrize("cfg", [
    # shape, [min, max], bits, narrow, gate
    ((4, 28, 28, 32), (-6.0, 6.0), 8, False, None),
    ((3, 7, 5, 13), (-0.37, 2.9), 8, False, 1.0),           # odd element count: scalar tail; range nudged onto zero
    ((64, 3, 3, 32), (-0.21, 0.17), 8, True, 1.0),          # weights: narrow range
    ((2, 16, 16, 24), (0.4, 5.0), 4, False, 1.0),           # all-positive range, few levels
    ((2, 16, 16, 24), (-6.0, 6.0), 8, False, 0.0),          # closed gate: identity
    ((2, 16, 16, 24), (1.0, 1.0), 8, False, 1.0),           # degenerate range: identity
])
def test_fake_quant(cfg):
    """dtf_fake_quant_fwd / _bwd against the PyTorch statement of FakeQuantWithMinMaxVars (--quantize)."""
    shape, (mn, mx), bits, narrow, gate = cfg
    x = _r(*shape, seed=5, scale=3.0)
    rng = torch.tensor([mn, mx], device="cuda")
    gate_t = None if gate is None else torch.full((), gate, device="cuda")
    dy = _r(*shape, seed=6)
    res = []
    for backend in ("auto", "torch"):
        Fn.set_backend(backend)
        try:
            xi = x.clone().requires_grad_(True)
            before = nat.launch_count()
            y = Fn.fake_quant(xi, rng, gate_t, bits, narrow)
            y.backward(dy)
            if backend == "auto":
                assert nat.launch_count() - before == 2, "the native kernels did not run"
            res.append((y.detach(), xi.grad))
        finally:
            Fn.set_backend("auto")
    (y_n, g_n), (y_t, g_t) = res
    identity = gate == 0.0 or mx <= mn
    if identity:
        assert torch.equal(y_n, x) and torch.equal(y_t, x) and torch.equal(g_n, dy)
        return
    step = float(Fn.nudged_range(rng[0], rng[1], bits, narrow)[2])
    # a value that sits on a half step may round to the neighbouring level under a different fused-multiply-add choice
    diff = (y_n - y_t).abs()
    assert float(diff.max()) <= step * 1.001
    assert float((diff > 1e-6).float().mean()) < 1e-3
    assert float((g_n != g_t).float().mean()) < 1e-3
    assert len(torch.unique(y_n)) <= 2 ** bits


@pytest.mark.parametrize("n", [5, 4096, 3 * 224 * 224 + 3])
def test_quant_range_update(n):
    """Batch min / max through float-ordered integer atomics, then last-value or moving-average range update."""
    for sign in (1.0, -1.0, 0.0):                       # all-positive, all-negative and mixed-sign tensors
        base = _r(n, seed=7, scale=2.0)
        x = base if sign == 0.0 else sign * (base.abs() + 0.25)
        mm = torch.tensor([float("inf"), float("-inf")], device="cuda")
        for decay in (-1.0, 0.9):
            rng_n = torch.tensor([-6.0, 6.0], device="cuda")
            rng_t = rng_n.clone()
            before = nat.launch_count()
            Fn.quant_range_update(x, rng_n, decay, mm)
            assert nat.launch_count() - before == 1
            Fn.set_backend("torch")
            try:
                Fn.quant_range_update(x, rng_t, decay)
            finally:
                Fn.set_backend("auto")
            assert torch.allclose(rng_n, rng_t, rtol=1e-6, atol=1e-7), (sign, decay, rng_n, rng_t)
            assert mm.tolist() == [float("inf"), float("-inf")]      # scratch handed back ready for the next call


def _records(n=600, seed=0):
    g = torch.Generator().manual_seed(seed)
    imgs = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, generator=g)
    imgs[3] = 77                        # a constant image: the standard deviation floor 1/sqrt(n) applies
    return imgs


@pytest.mark.parametrize("size", [24, 32])
def test_device_kernel_matches_the_host_pass(size):
    """dtf_cifar_distort (one block per image over HBM-resident records) against dtf_cifar_distort_host and the
    step-by-step PyTorch statement, through the prefetcher's copy stream."""
    from distributed_tensorflow_models_b200 import _native as nat
    dev = torch.device("cuda", 0)
    imgs = _records(2000, seed=3)

    class Source(cifar10.DeviceCifar10Batches):
        def __init__(self):                                  # records from memory instead of files
            self.device, self.images, self.size, self.pin = dev, imgs, size, True
            self.images_dev = imgs.to(dev)

    src = Source()
    for b, seed in ((128, 1), (7, 2), (512, 3)):
        idx = torch.randperm(2000, generator=torch.Generator().manual_seed(seed))[:b]
        draws = cifar10.draw_distortion(b, size, torch.Generator().manual_seed(seed + 10))
        want = cifar10.distort_records_native(imgs, idx, *draws, size)
        ref = cifar10.distort_batch_reference(imgs[idx], size, torch.Generator().manual_seed(seed + 10))
        before = nat.launch_count()
        got = cifar10.RecordBatch(src, idx, *draws).to(dev)
        assert nat.launch_count() == before + 1, "the native kernel did not run"
        torch.cuda.synchronize()
        assert torch.allclose(got.cpu(), want, atol=2e-6), float((got.cpu() - want).abs().max())
        keep = idx != 3
        assert torch.allclose(got.cpu()[keep], ref[keep], atol=2e-5)



def test_device_resident_pipeline_feeds_the_same_batches_as_the_host_pipeline(tmp_path):
    """DeviceCifar10Batches behind the prefetcher (upload of 21 bytes per image + kernel on the copy stream, event
    ordered against the consumer) next to the host pipeline behind the same prefetcher."""
    from distributed_tensorflow_models_b200.data import synthetic
    from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher
    dev = torch.device("cuda", 0)
    synthetic.write_cifar10_binary(str(tmp_path), train_per_file=40, test=8)
    host = cifar10.Cifar10Batches(str(tmp_path), 16, 24, seed=3)
    resident = cifar10.DeviceCifar10Batches(str(tmp_path), 16, 24, device=dev, seed=3)
    a, b = DevicePrefetcher(host, dev), DevicePrefetcher(resident, dev)
    try:
        for k in range(30):                                  # several passes over the three device buffers
            (xa, ya), (xb, yb) = next(a), next(b)
            assert xb.is_cuda and xb.shape == (16, 24, 24, 3) and torch.equal(ya, yb), k
            assert torch.allclose(xa, xb, atol=2e-6), (k, float((xa - xb).abs().max()))
            xb.add_(1.0)                                     # a consumer may scribble over its batch
    finally:
        host.close(), resident.close()


def test_quantised_trainer_on_the_native_kernels_under_a_cuda_graph():
    """The whole quantised step (range kernels, fake-quant forward / backward, folded batch norm) inside the captured
    CUDA graph: the delay switch flips between replays without a re-capture, the ranges keep moving, training learns."""
    from distributed_tensorflow_models_b200 import _native as nat
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(8, 64, 64, 3, generator=g).to(dev), torch.randint(0, 5, (8,), generator=g).to(dev)
    t = Trainer(TrainSetup(model="mobilenet_v1_025", batch_size=8, image_size=64, num_classes=5, seed=0,
                           initial_lr=0.02, cuda_graph=True, quantize=True, quant_delay=2), device=dev)
    before = nat.launch_count()
    losses, gates, ranges = [], [], []
    for _ in range(16):
        losses.append(float(t.train_step(x, y)["loss"]))
        gates.append(float(t.store.quant.gate))
        ranges.append(t.store.get("MobilenetV1/Conv2d_0/act_quant/min_max").tensor.clone())
    assert nat.launch_count() > before and t.graph_kernel_launches > 0
    assert gates[:3] == [0.0, 0.0, 1.0] and all(v == 1.0 for v in gates[3:])
    assert all(np.isfinite(losses)) and min(losses[-3:]) < 0.9 * losses[0]
    assert not torch.equal(ranges[0], ranges[-1]) and float(ranges[-1][0]) > -6.0
    w_rng = t.store.get("MobilenetV1/Conv2d_0/weights_quant/min_max").tensor
    assert float(w_rng[0]) < 0 < float(w_rng[1]) and float(w_rng[1] - w_rng[0]) < 12
    # evaluation with the learned ranges: every logit finite, nothing moves
    snap = {v.name: v.tensor.clone() for v in t.store.all() if "_quant" in v.name}
    _, logits = t.evaluate_batch(x, y)
    assert torch.isfinite(logits).all()
    for name, val in snap.items():
        assert torch.equal(val, t.store.get(name).tensor), name
