"""End-to-end GPU checks: every model trains through the native kernels and tracks the PyTorch-op
reference run (same seed, same data) step for step; the fused ResNet path and CUDA-graph replay agree with
the unfused eager path."""
import pytest
import torch

from distributed_tensorflow_models_b200.ops import functional as Fn
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup

pytestmark = pytest.mark.gpu


def _data(model, bs, seed=0):
    from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
    sp = MODEL_SPECS[model]
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(bs, sp.image_size, sp.image_size, 3, generator=g).cuda()
    y = torch.randint(0, sp.num_classes, (bs,), generator=g).cuda()
    return x, y


def _run(model, bs, steps, backend, **kw):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    Fn.set_backend(backend)
    try:
        t = Trainer(TrainSetup(model=model, batch_size=bs, **kw), device=torch.device("cuda"))
        x, y = _data(model, bs)
        losses = []
        for _ in range(steps):
            losses.append(float(t.train_step(x, y)["loss"]))
        return losses, t
    finally:
        Fn.set_backend("auto")


@pytest.mark.parametrize("model,bs", [("cnn", 16), ("cifarnet", 16), ("alexnet", 8), ("vgg", 4), ("resnet", 16)])
def test_native_tracks_reference(model, bs):
    # dropout-free comparison (masks differ between generators): alexnet/cifarnet use dropout in training
    ln, _ = _run(model, bs, 4, "auto", cuda_graph=False, fuse_bn=False)
    lr_, _ = _run(model, bs, 4, "torch", cuda_graph=False)
    if model in ("alexnet", "cifarnet"):
        assert abs(ln[0] - lr_[0]) / abs(lr_[0]) < 0.05
    else:
        # tf32 multiplies + ReLU-mask sensitivity: trajectories agree closely at first and drift slowly
        assert abs(ln[0] - lr_[0]) / abs(lr_[0]) < 2e-3, (ln, lr_)
        for a, b in zip(ln, lr_):
            assert abs(a - b) / max(abs(b), 1e-3) < 6e-2, (ln, lr_)
    assert ln[-1] == ln[-1]    # not NaN


def test_resnet_fused_path_matches_unfused():
    lf, tf_ = _run("resnet", 16, 4, "auto", cuda_graph=False, fuse_bn=True)
    lu, tu = _run("resnet", 16, 4, "auto", cuda_graph=False, fuse_bn=False)
    for a, b in zip(lf, lu):
        assert abs(a - b) / max(abs(b), 1e-3) < 6e-2, (lf, lu)
    # moving statistics are updated on both paths
    mf = [v.tensor for v in tf_.store.all() if v.name.endswith("moving_mean")]
    mu = [v.tensor for v in tu.store.all() if v.name.endswith("moving_mean")]
    assert float((mf[3] - mu[3]).abs().max()) < 1e-2 and float(mf[3].abs().max()) > 0


def test_resnet_cuda_graph_matches_eager():
    lg, _ = _run("resnet", 16, 5, "auto", cuda_graph=True, fuse_bn=True)
    le, _ = _run("resnet", 16, 5, "auto", cuda_graph=False, fuse_bn=True)
    for a, b in zip(lg, le):
        assert abs(a - b) / max(abs(b), 1e-3) < 6e-2, (lg, le)


def test_inception_native_step():
    ln, _ = _run("inception", 2, 2, "auto", cuda_graph=False)
    lr_, _ = _run("inception", 2, 2, "torch", cuda_graph=False)
    assert abs(ln[0] - lr_[0]) / abs(lr_[0]) < 0.05, (ln, lr_)
    assert ln[1] == ln[1]
