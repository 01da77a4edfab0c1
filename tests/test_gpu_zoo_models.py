"""Zoo networks on the native sm_100a kernels: forward and backward of whole networks agree with the same network
on plain PyTorch ops (fp32, TF32 off) from identical initial values, and the native library really ran."""
import contextlib
import functools

import pytest
import torch

from distributed_tensorflow_models_b200 import _native as nat
from distributed_tensorflow_models_b200.models import gan, nasnet
from distributed_tensorflow_models_b200.models.registry import get_network_fn
from distributed_tensorflow_models_b200.ops import functional as Fn
from distributed_tensorflow_models_b200.slim import VariableStore

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _run_both(build, x_cpu, fuse_bn=False):
    """build(x) -> tensor (called inside store.as_default()).  Runs the network three times from identical initial
    values: native kernels, PyTorch ops in strict fp32, PyTorch ops with TF32 allowed (cuDNN / cuBLAS tensor cores).
    The third run measures how much TF32 rounding this particular network amplifies (tiny-batch batch norm does, a
    lot), which is the yardstick the native run — also TF32 multiplies — is held to."""
    res = {}
    for backend in ("auto", "torch", "torch_tf32"):
        torch.backends.cudnn.allow_tf32 = backend == "torch_tf32"
        torch.backends.cuda.matmul.allow_tf32 = backend == "torch_tf32"
        st = VariableStore(seed=3)
        with torch.no_grad(), st.as_default():          # build pass on the CPU: identical initial values
            st.begin_forward()
            build(x_cpu)
        st.freeze()
        st.to("cuda")
        st.fuse_bn = fuse_bn and backend == "auto"
        Fn.set_backend("auto" if backend == "auto" else "torch")
        try:
            before = nat.launch_count()
            if backend == "auto":
                from distributed_tensorflow_models_b200.ops import native_ops
                native_ops.begin_step(torch.device("cuda", torch.cuda.current_device()))
            with st.as_default():
                st.begin_forward()
                out = build(x_cpu.cuda())
            g = torch.Generator().manual_seed(5)
            out.backward(torch.randn(out.shape, generator=g).cuda() / out.numel() ** 0.5)
            launches = nat.launch_count() - before
        finally:
            Fn.set_backend("auto")
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False
        tv = st.trainable_variables()
        res[backend] = (out.detach(), {v.name: v.tensor.grad.detach().clone() for v in tv if v.tensor.grad is not None},
                        launches)
    assert res["auto"][2] > 10, "native kernels did not run"
    assert res["torch"][2] == 0
    return res


def _check(res, out_tol, grad_tol, tag=""):
    """native vs fp32 must be within ``tol`` or within 4x of what PyTorch's own TF32 path deviates from fp32."""
    (on, gn, _), (ot, gt, _), (o32, g32, _) = res["auto"], res["torch"], res["torch_tf32"]
    assert torch.isfinite(on).all()
    e_n, e_t = _rel(on, ot), _rel(o32, ot)
    assert set(gn) == set(gt)
    names = [n for n in gt if float(gt[n].norm()) > 1e-6]
    worst_n = max((_rel(gn[n], gt[n]), n) for n in names)
    worst_t = max((_rel(g32[n], gt[n]), n) for n in names)
    excess = max((_rel(gn[n], gt[n]) - 4 * _rel(g32[n], gt[n]), n) for n in names)
    try:
        import os
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/zoo_models_diag.log", "a") as f:
            f.write(f"{tag}: out native {e_n:.4f} / torch-tf32 {e_t:.4f}; worst grad native {worst_n[0]:.4f} "
                    f"({worst_n[1]}) / torch-tf32 {worst_t[0]:.4f} ({worst_t[1]}); excess {excess[0]:.4f}\n")
    except OSError:
        pass
    assert e_n < max(out_tol, 4 * e_t), f"output mismatch {e_n} (torch tf32: {e_t})"
    assert excess[0] < grad_tol, f"gradient mismatch {worst_n} (torch tf32: {worst_t}; excess {excess})"


@pytest.mark.parametrize("name,size,bs,fuse_bn", [
    ("mobilenet_v1_025", 64, 4, False),
    ("mobilenet_v1_025", 64, 4, True),        # conv + BN statistics epilogue, ReLU6 applied after
    ("mobilenet_v2", 64, 4, False),
    ("resnet_v1_50", 64, 2, False),
    ("resnet_v2_50_slim", 64, 2, False),
    ("inception_v1", 96, 2, False),
    ("inception_v2", 96, 2, False),
    ("inception_v3_slim", 64, 2, False),
    ("inception_resnet_v2", 107, 2, False),
    ("nasnet_cifar", 32, 4, False),
])
def test_zoo_classifier_native_matches_torch(name, size, bs, fuse_bn):
    fn = get_network_fn(name, num_classes=12, weight_decay=0.0, is_training=True)
    kw = {}
    if name.startswith("mobilenet_v1") or name.startswith("inception"):
        kw["dropout_keep_prob"] = 1.0            # masks come from different generators on the two back-ends
    if name == "nasnet_cifar":
        cfg = nasnet.cifar_config()
        cfg.num_cells, cfg.drop_path_keep_prob, cfg.use_aux_head = 6, 1.0, 0
        kw["config"] = cfg
    if name == "mobilenet_v2":
        kw["dropout_keep_prob"] = 1.0
    if name.startswith("inception_v3") or name.startswith("inception_resnet"):
        kw["create_aux_logits"] = False
    g = torch.Generator().manual_seed(1)
    x = torch.randn(bs, size, size, 3, generator=g)

    def build(inp):
        out, _ = fn(inp, **kw)
        return out

    _check(_run_both(build, x, fuse_bn), 1e-2, 6e-2, f"{name} fuse_bn={fuse_bn}")


def test_dcgan_native_matches_torch():
    g = torch.Generator().manual_seed(2)
    z = torch.randn(8, 64, generator=g)

    def build(noise):
        img, _ = gan.dcgan_generator(noise, depth=32, final_size=32)
        logits, _ = gan.dcgan_discriminator(Fn.tanh(img), depth=32)
        return logits

    _check(_run_both(build, z), 1e-2, 6e-2, "dcgan")


@pytest.mark.parametrize("method", ["conv2d_transpose", "nn_upsample_conv", "bilinear_upsample_conv"])
def test_cyclegan_native_matches_torch(method):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 32, 32, 3, generator=g)
    up = functools.partial(gan.cyclegan_upsample, method=method)

    def build(img):
        pred, _ = gan.cyclegan_generator_resnet(img, num_resnet_blocks=2, num_filters=16, upsample_fn=up)
        return pred

    _check(_run_both(build, x), 1e-2, 6e-2, f"cyclegan {method}")


@pytest.mark.parametrize("method", ["nn_upsample_conv", "conv2d_transpose"])
def test_pix2pix_native_matches_torch(method):
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 64, 64, 3, generator=g)
    blocks = [gan.Pix2PixBlock(16, 0), gan.Pix2PixBlock(32, 0), gan.Pix2PixBlock(64, 0), gan.Pix2PixBlock(64, 0)]

    def build(img):
        with gan.pix2pix_arg_scope():
            fake, _ = gan.pix2pix_generator(img, 3, blocks=blocks, upsample_method=method)
            with VariableStoreScope("disc"):
                logits, _ = gan.pix2pix_discriminator(torch.cat([img, Fn.tanh(fake)], dim=3), [16, 32, 64])
        return logits

    _check(_run_both(build, x), 1e-2, 6e-2, f"pix2pix {method}")


@contextlib.contextmanager
def VariableStoreScope(name):
    from distributed_tensorflow_models_b200.slim import variables as V
    with V.current_store().scope(name):
        yield
