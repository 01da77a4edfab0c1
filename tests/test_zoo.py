"""The vendored model zoo on this framework's layer library: parameter censuses, variable names,
end points and shapes — the properties the reference pins in vgg/nets/*_test.py (e.g. mobilenet_v1_test.py:36-120,
resnet_v1_test.py:270-330, inception_v4_test.py:28-120, nets_factory_test.py:30-58, dcgan_test.py, pix2pix_test.py,
cyclegan_test.py)."""
import contextlib
import functools

import pytest
import torch

from distributed_tensorflow_models_b200.models import gan, inception_zoo, mobilenet, nasnet, registry, resnet_slim
from distributed_tensorflow_models_b200.models.registry import get_network_fn
from distributed_tensorflow_models_b200.ops import functional as Fn
from distributed_tensorflow_models_b200.slim import VariableStore, ops
from distributed_tensorflow_models_b200.slim import variables as V


def _build(fn, cm, *args, grad=False, **kw):
    st = VariableStore()
    with (contextlib.nullcontext() if grad else torch.no_grad()), st.as_default():
        st.begin_forward()
        with cm:
            out, end_points = fn(*args, **kw)
    return st, out, end_points


def _nparams(st, exclude=None):
    return sum(v.tensor.numel() for v in st.trainable_variables() if not (exclude and exclude in v.name))


# ---- censuses -----------------------------------------------------------------------------------------------
def test_mobilenet_v1_census_and_names():
    st, out, ep = _build(mobilenet.mobilenet_v1, mobilenet.mobilenet_v1_arg_scope(), torch.randn(1, 224, 224, 3),
                         num_classes=1001)
    assert tuple(out.shape) == (1, 1001)
    assert _nparams(st) == 4233001                     # 4.2 M: the published MobileNet-224 1.0 size
    names = [v.name for v in st.all()]
    assert names[0] == "MobilenetV1/Conv2d_0/weights"
    assert "MobilenetV1/Conv2d_1_depthwise/depthwise_weights" in names
    assert "MobilenetV1/Conv2d_13_pointwise/BatchNorm/moving_variance" in names
    assert names[-2:] == ["MobilenetV1/Logits/Conv2d_1c_1x1/weights", "MobilenetV1/Logits/Conv2d_1c_1x1/biases"]
    assert tuple(ep["Conv2d_13_pointwise"].shape) == (1, 7, 7, 1024)
    assert tuple(st.get("MobilenetV1/Conv2d_1_depthwise/depthwise_weights").tensor.shape) == (3, 3, 32, 1)
    # depthwise filters are not regularised, 1x1 / stem convolutions and the logits are: 1 + 13 + 1
    assert st.num_regularizers() == 15


@pytest.mark.parametrize("fn,params", [(mobilenet.mobilenet_v1_075, None), (mobilenet.mobilenet_v1_050, None),
                                       (mobilenet.mobilenet_v1_025, 470329)])
def test_mobilenet_v1_fractions(fn, params):
    st, out, ep = _build(fn, mobilenet.mobilenet_v1_arg_scope(), torch.randn(1, 128, 128, 3), num_classes=1001)
    assert tuple(out.shape) == (1, 1001)
    if params:
        assert _nparams(st) == params


def test_mobilenet_v1_base_endpoints_and_output_stride():
    with pytest.raises(ValueError):
        _build(mobilenet.mobilenet_v1_base, contextlib.nullcontext(), torch.randn(1, 64, 64, 3), final_endpoint="nope")
    st, net, ep = _build(mobilenet.mobilenet_v1_base, mobilenet.mobilenet_v1_arg_scope(), torch.randn(1, 64, 64, 3),
                         final_endpoint="Conv2d_5_pointwise")
    assert tuple(net.shape) == (1, 8, 8, 256) and "Conv2d_6_depthwise" not in ep
    st, net, ep = _build(mobilenet.mobilenet_v1_base, mobilenet.mobilenet_v1_arg_scope(), torch.randn(1, 64, 64, 3),
                         output_stride=8)
    assert tuple(net.shape) == (1, 8, 8, 1024)        # strides beyond 8 became dilation


def test_mobilenet_v2_census_and_names():
    st, out, ep = _build(mobilenet.mobilenet_v2, mobilenet.mobilenet_v2_training_scope(), torch.randn(1, 224, 224, 3),
                         num_classes=1001)
    assert tuple(out.shape) == (1, 1001)
    assert _nparams(st) == 3506153                     # 3.5 M
    names = [v.name for v in st.all()]
    assert names[0] == "MobilenetV2/Conv/weights"
    assert "MobilenetV2/expanded_conv/depthwise/depthwise_weights" in names
    assert "MobilenetV2/expanded_conv/expand/weights" not in names      # first block has expansion 1
    assert "MobilenetV2/expanded_conv_16/project/BatchNorm/gamma" in names
    assert tuple(ep["layer_19"].shape) == (1, 7, 7, 1280)
    assert tuple(ep["layer_4"].shape) == (1, 56, 56, 24)
    # multiplier 0.5 keeps channel counts divisible by 8; fine-grain mode keeps the 1280 head
    st, out, ep = _build(mobilenet.mobilenet_v2, contextlib.nullcontext(), torch.randn(1, 96, 96, 3), num_classes=11,
                         depth_multiplier=0.5, finegrain_classification_mode=True)
    assert ep["layer_19"].shape[-1] == 1280 and ep["layer_2"].shape[-1] == 8


@pytest.mark.parametrize("fn,nc,params,nvars", [
    (resnet_slim.resnet_v1_50, 1000, 25557032, 267),       # == the canonical ResNet-50 / -101 sizes
    (resnet_slim.resnet_v1_101, 1000, 44549160, 522),
    (resnet_slim.resnet_v2_50, 1001, 25570409, 272),
])
def test_slim_resnet_census(fn, nc, params, nvars):
    st, out, ep = _build(fn, resnet_slim.resnet_arg_scope(), torch.randn(1, 64, 64, 3), num_classes=nc)
    assert tuple(out.shape) == (1, nc)
    assert _nparams(st) == params and len(st.all()) == nvars


def test_slim_resnet_names_endpoints_and_dense_mode():
    st, out, ep = _build(resnet_slim.resnet_v1_50, resnet_slim.resnet_arg_scope(), torch.randn(1, 65, 65, 3),
                         num_classes=None, global_pool=False, output_stride=8)
    assert tuple(out.shape) == (1, 9, 9, 2048)         # (65 - 1) / 8 + 1
    names = [v.name for v in st.all()]
    assert names[0] == "resnet_v1_50/conv1/weights"
    assert "resnet_v1_50/block1/unit_1/bottleneck_v1/shortcut/weights" in names
    assert "resnet_v1_50/block1/unit_2/bottleneck_v1/shortcut/weights" not in names
    assert "resnet_v1_50/block4/unit_3/bottleneck_v1/conv3/BatchNorm/gamma" in names
    assert tuple(ep["resnet_v1_50/block1"].shape) == (1, 9, 9, 256)
    st, out, ep = _build(resnet_slim.resnet_v2_50, resnet_slim.resnet_arg_scope(), torch.randn(1, 64, 64, 3),
                         num_classes=10)
    names = [v.name for v in st.all()]
    assert names[:2] == ["resnet_v2_50/conv1/weights", "resnet_v2_50/conv1/biases"]
    assert "resnet_v2_50/block1/unit_1/bottleneck_v2/preact/gamma" in names
    assert "resnet_v2_50/postnorm/moving_mean" in names
    assert names[-2:] == ["resnet_v2_50/logits/weights", "resnet_v2_50/logits/biases"]
    with pytest.raises(ValueError):
        _build(resnet_slim.resnet_v1_50, resnet_slim.resnet_arg_scope(), torch.randn(1, 64, 64, 3), output_stride=6)


@pytest.mark.parametrize("fn,scope,size,total,no_aux,nvars,last,last_shape", [
    (inception_zoo.inception_v1, inception_zoo.inception_arg_scope, 224, 6618649, 6618649, 230, "Mixed_5c", (7, 7, 1024)),
    (inception_zoo.inception_v2, inception_zoo.inception_arg_scope, 224, 11179361, 11179361, 277, "Mixed_5c", (7, 7, 1024)),
    (inception_zoo.inception_v4, inception_zoo.inception_v4_arg_scope, 299, 46009106, 42649769, 608, "Mixed_7d", (8, 8, 1536)),
    (inception_zoo.inception_resnet_v2, inception_zoo.inception_resnet_v2_arg_scope, 299, 59182258, 55814729, 908,
     "Conv2d_7b_1x1", (8, 8, 1536)),
])
def test_inception_zoo_census(fn, scope, size, total, no_aux, nvars, last, last_shape):
    st, out, ep = _build(fn, scope(), torch.randn(1, size, size, 3), num_classes=1001)
    assert tuple(out.shape) == (1, 1001)
    assert _nparams(st) == total and _nparams(st, exclude="AuxLogits") == no_aux
    assert len(st.all()) == nvars
    assert tuple(ep[last].shape[1:]) == last_shape
    assert torch.allclose(ep["Predictions"].sum(1), torch.ones(1), atol=1e-5)


def test_inception_v3_slim_edition_matches_trained_v3_census_and_survives_cifar_inputs():
    # same architecture as the in-tree model the reference trains: identical 27,145,970 parameters / 388 variables
    st, out, ep = _build(inception_zoo.inception_v3, inception_zoo.inception_v3_arg_scope(),
                         torch.randn(1, 299, 299, 3), num_classes=1001)
    assert _nparams(st) == 27145970 and len(st.all()) == 388
    assert tuple(ep["Mixed_7c"].shape) == (1, 10, 10, 2048)      # SAME padding everywhere: 299 -> 10, not 8
    st, out, ep = _build(inception_zoo.inception_v3, inception_zoo.inception_v3_arg_scope(),
                         torch.randn(2, 32, 32, 3), num_classes=1001)
    assert tuple(out.shape) == (2, 1001) and tuple(ep["AuxLogits"].shape) == (2, 1001)   # the point of the edits
    names = [v.name for v in st.all()]
    assert "InceptionV3/Mixed_5c/Branch_1/Conv_1_0c_5x5/weights" in names               # irregular names kept
    assert "InceptionV3/Mixed_6a/Branch_0/Conv2d_1a_1x1/weights" in names
    assert "InceptionV3/Mixed_7b/Branch_1/Conv2d_0b_3x1/weights" in names
    assert "InceptionV3/Mixed_7c/Branch_1/Conv2d_0c_3x1/weights" in names
    assert not any("gamma" in n for n in names)


def test_inception_names_quirks_and_final_endpoints():
    st, net, ep = _build(inception_zoo.inception_v1_base, inception_zoo.inception_arg_scope(),
                         torch.randn(1, 224, 224, 3), final_endpoint="Mixed_5b")
    assert "InceptionV1/Mixed_5b/Branch_2/Conv2d_0a_3x3/weights" in [v.name for v in st.all()]
    assert tuple(net.shape) == (1, 7, 7, 832) and "Mixed_5c" not in ep
    st, net, ep = _build(inception_zoo.inception_v2_base, inception_zoo.inception_arg_scope(),
                         torch.randn(1, 224, 224, 3), final_endpoint="Mixed_4a")
    names = [v.name for v in st.all()]
    assert names[:3] == ["InceptionV2/Conv2d_1a_7x7/depthwise_weights", "InceptionV2/Conv2d_1a_7x7/pointwise_weights",
                         "InceptionV2/Conv2d_1a_7x7/biases"]
    assert tuple(st.get(names[0]).tensor.shape) == (7, 7, 3, 8)
    assert tuple(net.shape) == (1, 14, 14, 576)
    with pytest.raises(ValueError):
        _build(inception_zoo.inception_v4_base, inception_zoo.inception_arg_scope(), torch.randn(1, 75, 75, 3),
               final_endpoint="Mixed_9z")
    st, net, ep = _build(inception_zoo.inception_resnet_v2_base, inception_zoo.inception_resnet_v2_arg_scope(),
                         torch.randn(1, 149, 149, 3), final_endpoint="PreAuxLogits", output_stride=8,
                         align_feature_maps=True)
    assert tuple(net.shape) == (1, 19, 19, 1088)
    names = [v.name for v in st.all()]
    assert "InceptionResnetV2/Repeat/block35_10/Conv2d_1x1/biases" in names
    assert "InceptionResnetV2/Repeat_1/block17_20/Branch_1/Conv2d_0c_7x1/weights" in names


@pytest.mark.parametrize("fn,scope,size,nc,train,params,endpoints", [
    # 5,291,035 / 88,757,183 are the published NASNet-A mobile / large sizes (+ the background class of 1001)
    (nasnet.build_nasnet_mobile, nasnet.nasnet_mobile_arg_scope, 224, 1001, False, 5291035,
     {"Stem": (28, 28, 88), "Cell_0": (28, 28, 264), "Reduction_Cell_1": (7, 7, 704), "Cell_11": (7, 7, 1056)}),
    (nasnet.build_nasnet_mobile, nasnet.nasnet_mobile_arg_scope, 224, 1001, True, 7703044, {"AuxLogits": (1001,)}),
    (nasnet.build_nasnet_cifar, nasnet.nasnet_cifar_arg_scope, 32, 10, True, 4978036,
     {"Stem": (32, 32, 96), "Cell_17": (8, 8, 768), "AuxLogits": (10,)}),
    (nasnet.build_nasnet_large, nasnet.nasnet_large_arg_scope, 331, 1001, False, 88757183,
     {"Cell_0": (42, 42, 1008), "Cell_17": (11, 11, 4032)}),
    (nasnet.build_pnasnet_large, nasnet.pnasnet_large_arg_scope, 331, 1001, False, 86061989,
     {"Stem": (42, 42, 540), "Cell_11": (11, 11, 4320)}),
])
def test_nasnet_census(fn, scope, size, nc, train, params, endpoints):
    st, out, ep = _build(fn, scope(), torch.randn(1, size, size, 3), num_classes=nc, is_training=train)
    assert tuple(out.shape) == (1, nc) and _nparams(st) == params
    for k, shape in endpoints.items():
        assert tuple(ep[k].shape[1:]) == shape, k
    names = {v.name for v in st.all()}
    assert not any(n.endswith("biases") and "FC" not in n for n in names)      # bare convolutions
    if fn is not nasnet.build_nasnet_cifar:
        assert {"conv0/weights", "conv0_bn/gamma", "cell_stem_0/1x1/weights", "cell_stem_1/path1_conv/weights",
                "cell_stem_1/final_path_bn/beta", "cell_0/path1_conv/weights", "cell_1/prev_1x1/weights",
                "final_layer/FC/weights"} <= names
    if fn is nasnet.build_nasnet_mobile:
        assert {"cell_0/comb_iter_0/left/separable_5x5_1/depthwise_weights",
                "cell_0/comb_iter_0/left/bn_sep_5x5_2/moving_variance",
                "reduction_cell_0/comb_iter_1/right/separable_7x7_2/pointwise_weights",
                "reduction_cell_1/comb_iter_0/left/separable_5x5_1/depthwise_weights"} <= names
        assert ("aux_7/aux_logits/proj/weights" in names) == train


def test_nasnet_genotype_helpers_and_drop_path_schedule():
    assert nasnet.calc_reduction_layers(18, 2) == [6, 12] and nasnet.calc_reduction_layers(12, 2) == [4, 8]
    cell = nasnet.Cell(nasnet.NASNET_A_NORMAL, 8, 0.6, 20, 1000, True, current_step=0)
    x = torch.ones(4, 2, 2, 8)
    assert torch.equal(cell._drop_path(x, 3), x)                 # ramp starts at keep = 1
    cell.current_step = 1000
    torch.manual_seed(0)
    vals = cell._drop_path(torch.ones(256, 1, 1, 8), 19).unique().tolist()
    assert len(vals) == 2 and vals[0] == 0.0 and abs(vals[1] - 1 / 0.6) < 1e-5      # last cell, end of training
    st = VariableStore()
    with torch.no_grad(), st.as_default(), nasnet.nasnet_cifar_arg_scope():
        y = nasnet.factorized_reduction(torch.randn(1, 7, 7, 8), 12, 2)
    assert tuple(y.shape) == (1, 4, 4, 12)
    assert [v.name for v in st.trainable_variables()] == ["path1_conv/weights", "path2_conv/weights",
                                                          "final_path_bn/beta", "final_path_bn/gamma"]


# ---- the factory over everything it registers ---------------------------------------------------------------
# the reference's local edits made alexnet_v2 (10 classes), the VGGs and the nets-edition Inception-v3 CIFAR-sized nets
_SMALL = {"cnn": 24, "cifarnet": 32, "lenet": 28, "resnet_cifar": 32, "inception_v3_slim": 32, "alexnet_v2": 32,
          "vgg_a": 32, "vgg_16": 32, "vgg_19": 32}


@pytest.mark.parametrize("name", sorted(registry.networks_map))
def test_nets_factory_builds_every_registered_network(name):
    if name in ("resnet_v2_152", "resnet_v2_200", "resnet_v1_152", "resnet_v1_200", "resnet_v2_152_slim",
                "resnet_v2_200_slim", "resnet_v2_101", "resnet_v2_101_slim", "resnet_v1_101", "nasnet_large",
                "pnasnet_large"):
        pytest.skip("deeper copies of a tested architecture (CPU time)")
    fn = get_network_fn(name, num_classes=10, is_training=False)
    size = _SMALL.get(name, getattr(fn, "default_image_size", 224))
    st = VariableStore()
    with torch.no_grad(), st.as_default():
        st.begin_forward()
        out, end_points = fn(torch.randn(1, size, size, 3))
    logits = out[0] if isinstance(out, tuple) else out
    assert tuple(logits.shape) == (1, 10), name
    assert torch.isfinite(logits).all()


# ---- GANs --------------------------------------------------------------------------------------------------
def test_dcgan_shapes_names_and_validation():
    st, logits, ep = _build(gan.dcgan_discriminator, contextlib.nullcontext(), torch.randn(2, 32, 32, 3))
    assert tuple(logits.shape) == (2, 1) and _nparams(st) == 11147201
    names = [v.name for v in st.all()]
    assert names[:2] == ["Discriminator/conv1/weights", "Discriminator/conv1/biases"]    # no BN on the first layer
    assert "Discriminator/conv2/BatchNorm/beta" in names and names[-1] == "Discriminator/Conv/biases"
    assert tuple(ep["conv5"].shape) == (2, 1, 1, 1024)
    for bad in (torch.randn(2, 32, 16, 3), torch.randn(2, 24, 24, 3)):
        with pytest.raises(ValueError):
            _build(gan.dcgan_discriminator, contextlib.nullcontext(), bad)
    st, img, ep = _build(gan.dcgan_generator, contextlib.nullcontext(), torch.randn(2, 64), final_size=32, grad=True)
    assert tuple(img.shape) == (2, 32, 32, 3) and tuple(ep["deconv1"].shape) == (2, 4, 4, 512)
    w = st.get("Generator/deconv2/weights")
    assert tuple(w.tensor.shape) == (256, 4, 4, 512) and tuple(w.to_tf().shape) == (4, 4, 256, 512)   # [kh, kw, out, in]
    img.square().mean().backward()
    assert all(v.tensor.grad is not None and torch.isfinite(v.tensor.grad).all() for v in st.trainable_variables())
    for size in (24, 4):
        with pytest.raises(ValueError):
            _build(gan.dcgan_generator, contextlib.nullcontext(), torch.randn(2, 64), final_size=size)


@pytest.mark.parametrize("method", ["nn_upsample_conv", "conv2d_transpose"])
def test_pix2pix_generator_and_discriminator(method):
    x = torch.randn(1, 128, 128, 3)
    blocks = [gan.Pix2PixBlock(16, 0.5), gan.Pix2PixBlock(32, 0.5), gan.Pix2PixBlock(64, 0), gan.Pix2PixBlock(64, 0)]
    st, logits, ep = _build(gan.pix2pix_generator, gan.pix2pix_arg_scope(), x, 3, blocks=blocks, upsample_method=method,
                            grad=True)
    assert tuple(logits.shape) == (1, 128, 128, 3)
    assert tuple(ep["encoder3"].shape) == (1, 8, 8, 64) and tuple(ep["decoder0"].shape) == (1, 16, 16, 64)
    assert float(ep["predictions"].detach().abs().max()) <= 1.0
    names = [v.name for v in st.all()]
    assert names[:2] == ["encoder/Conv/weights", "encoder/Conv/biases"]
    assert "encoder/Conv_1/InstanceNorm/gamma" in names
    assert ("decoder/Conv2d_transpose/weights" in names) == (method == "conv2d_transpose")
    logits.mean().backward()
    assert all(v.tensor.grad is not None for v in st.trainable_variables())
    with pytest.raises(ValueError):
        _build(gan.pix2pix_generator, gan.pix2pix_arg_scope(), x, 3, blocks=blocks, upsample_method="bogus")
    st, logits, ep = _build(gan.pix2pix_discriminator, gan.pix2pix_arg_scope(), torch.randn(1, 256, 256, 6),
                            [64, 128, 256, 512])
    assert tuple(logits.shape) == (1, 35, 35, 1) and _nparams(st) == 2768705     # the 70x70 PatchGAN
    assert tuple(ep["conv0"].shape) == (1, 129, 129, 64)


@pytest.mark.parametrize("method,params", [("conv2d_transpose", 7841219), ("nn_upsample_conv", 7841411),
                                           ("bilinear_upsample_conv", 7841411)])
def test_cyclegan_generator(method, params):
    up = functools.partial(gan.cyclegan_upsample, method=method)
    st, pred, ep = _build(gan.cyclegan_generator_resnet, contextlib.nullcontext(), torch.randn(1, 40, 52, 3),
                          upsample_fn=up, grad=True)
    assert tuple(pred.shape) == (1, 40, 52, 3) and _nparams(st) == params
    assert tuple(ep["encoder_2"].shape) == (1, 10, 13, 256) and tuple(ep["decoder1"].shape) == (1, 20, 26, 128)
    assert "residual_blocks/block_5/Conv_1/InstanceNorm/beta" in [v.name for v in st.all()]
    pred.mean().backward()
    assert all(v.tensor.grad is not None for v in st.trainable_variables())
    with pytest.raises(ValueError):
        _build(gan.cyclegan_generator_resnet, contextlib.nullcontext(), torch.randn(1, 42, 52, 3))


# ---- the op kinds that only the zoo uses (CPU oracle side; the CUDA kernels are checked in test_gpu_zoo_ops) -----
def test_depthwise_and_separable_conv_against_grouped_convolution():
    torch.manual_seed(0)
    x, w = torch.randn(2, 9, 9, 6), torch.randn(3, 3, 6, 2)
    y = Fn.depthwise_conv2d(x, w, None, 2, "SAME")
    ref = torch.zeros(2, 5, 5, 12)
    xp = torch.nn.functional.pad(x, (0, 0, 1, 1, 1, 1))
    for oh in range(5):
        for ow in range(5):
            ref[:, oh, ow] = torch.einsum("nrsc,rscm->ncm", xp[:, 2 * oh:2 * oh + 3, 2 * ow:2 * ow + 3], w).reshape(2, 12)
    assert torch.allclose(y, ref, atol=1e-5)
    st = VariableStore()
    with st.as_default():
        out = ops.separable_convolution2d(x, 7, [3, 3], depth_multiplier=2, scope="sep")
    assert tuple(out.shape) == (2, 9, 9, 7)
    assert [(v.name, tuple(v.tf_shape)) for v in st.all()] == [
        ("sep/depthwise_weights", (3, 3, 6, 2)), ("sep/pointwise_weights", (1, 1, 12, 7)), ("sep/biases", (7,))]


@pytest.mark.parametrize("cfg", [(2, 17, 19, 8, 12, 3, 2, "SAME"), (1, 16, 16, 4, 4, 3, 4, "SAME"),
                                 (2, 21, 13, 4, 6, 3, 3, "VALID"), (2, 9, 9, 4, 6, 1, 2, "SAME")])
def test_dilated_convolution_as_one_dense_convolution_over_sampling_phases(cfg):
    """The native path of slim.conv2d(rate=) / separable_conv2d(rate=): phases into the batch, dense VALID
    convolution, phases back — equal to the dilated convolution in value and in both gradients."""
    n, h, w, c, co, k, rate, pad = cfg
    torch.manual_seed(3)
    x = torch.randn(n, h, w, c, requires_grad=True)
    wt, b, dw = torch.randn(co, k, k, c, requires_grad=True), torch.randn(co), torch.randn(k, k, c, 2)
    er = (k - 1) * rate + 1
    oh, ow, ph, pw = Fn.resolve_padding(h, w, er, er, (1, 1), pad)
    y0 = Fn.atrous_conv2d(x, wt, b, rate, pad)
    y1 = Fn._by_phases(Fn.conv2d, x, wt, b, rate, ph, pw, oh, ow)
    assert y0.shape == y1.shape == (n, oh, ow, co)
    assert torch.allclose(y0, y1, atol=1e-5)
    g = torch.randn_like(y0)
    for a, e in zip(torch.autograd.grad(y1, [x, wt], g), torch.autograd.grad(y0, [x, wt], g)):
        assert torch.allclose(a, e, atol=1e-3, rtol=1e-4)
    d0 = Fn.depthwise_conv2d(x, dw, None, 1, pad, rate=rate)
    d1 = Fn._by_phases(Fn.depthwise_conv2d, x, dw, None, rate, ph, pw, oh, ow)
    assert torch.allclose(d0, d1, atol=1e-5)
    # round trip of the two permutations alone
    xb = Fn._space_to_batch(x, rate, (0, 0), (0, 0))
    assert xb.shape[0] == n * rate * rate
    assert torch.equal(Fn._batch_to_space(xb, rate, n, h, w), x)


@pytest.mark.parametrize("k,s,pad,h", [(4, 2, "SAME", 5), (3, 2, "SAME", 5), (4, 1, "VALID", 1), (3, 2, "VALID", 4)])
def test_conv2d_transpose_is_the_adjoint_of_conv2d(k, s, pad, h):
    torch.manual_seed(1)
    x, w = torch.randn(2, h, h, 3), torch.randn(5, k, k, 3)
    y = Fn.conv2d_transpose(x, w, None, s, pad)
    assert y.shape[1] == Fn.conv2d_transpose_output_size(h, k, s, pad)
    g = torch.randn_like(y)
    back = Fn.conv2d(g, w.permute(3, 1, 2, 0).contiguous(), None, s, pad)     # the convolution being transposed
    assert abs(float((back * x).sum() - (g * y).sum())) < 1e-3


def test_instance_norm_resize_pad_activations():
    torch.manual_seed(2)
    x = torch.randn(2, 6, 5, 4) * 3 + 1
    y = Fn.instance_norm(x, None, None, 1e-6)
    assert torch.allclose(y.mean((1, 2)), torch.zeros(2, 4), atol=1e-5)
    assert torch.allclose(y.var((1, 2), unbiased=False), torch.ones(2, 4), atol=1e-3)
    up = Fn.resize_nearest(x, (12, 10))
    assert torch.equal(up[:, ::2, ::2], x) and torch.equal(up[:, 1::2, 1::2], x)
    bl = Fn.resize_bilinear(x, (11, 9), align_corners=True)
    ref = torch.nn.functional.interpolate(x.permute(0, 3, 1, 2), size=(11, 9), mode="bilinear", align_corners=True)
    assert torch.allclose(bl, ref.permute(0, 2, 3, 1), atol=1e-5)
    p = Fn.pad_spatial(x, (2, 1, 0, 3), "reflect")
    assert tuple(p.shape) == (2, 9, 8, 4) and torch.equal(p[:, 0], p[:, 4]) and torch.equal(p[:, :, -1], p[:, :, -7])
    assert float(Fn.relu6(torch.tensor([-1.0, 3.0, 9.0])).sum()) == 9.0
    assert torch.allclose(Fn.leaky_relu(torch.tensor([-2.0, 2.0])), torch.tensor([-0.4, 2.0]))
    d = Fn.drop_path(torch.ones(64, 2, 2, 3), 0.5, True, generator=torch.Generator().manual_seed(0))
    assert set(d.unique().tolist()) <= {0.0, 2.0} and torch.equal(Fn.drop_path(x, 0.5, False), x)


# ---- zoo networks through the training engine and the zoo's own train / eval entry points ----------------------
def test_engine_trains_any_registered_network():
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    t = Trainer(TrainSetup(model="mobilenet_v2", batch_size=8, image_size=32, num_classes=10, initial_lr=0.02,
                           cuda_graph=False))
    assert t.spec.image_size == 32 and t.spec.num_classes == 10
    # the arg scope's L2 (4e-5) sits on the convolution weights only: depthwise filters / BN parameters are not decayed
    assert t.layout.weight_decay == pytest.approx(4e-5) and 0 < t.layout.wd_boundary < t.layout.params.numel()
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(8, 32, 32, 3, generator=g), torch.randint(0, 10, (8,), generator=g)
    losses_ = [float(t.train_step(x, y)["loss"]) for _ in range(10)]
    assert losses_[-1] < 0.5 * losses_[0]            # memorises the batch (dropout + tiny-batch BN make it noisy)
    # a network with an auxiliary head: its loss joins with weight 0.4 in training, not in evaluation
    t = Trainer(TrainSetup(model="inception_v3_slim", batch_size=2, image_size=32, num_classes=10, cuda_graph=False))
    x, y = torch.randn(2, 32, 32, 3, generator=g), torch.randint(0, 10, (2,), generator=g)
    eval_loss, logits = t.evaluate_batch(x, y)
    assert tuple(logits.shape) == (2, 10) and torch.isfinite(logits).all() and torch.isfinite(eval_loss)
    out = t.train_step(x, y)
    main_only = float(Fn.softmax_cross_entropy(out["logits"], y))
    assert float(out["loss"]) > main_only + 0.3      # + 0.4 x the auxiliary head's cross-entropy (~ln 10 at init)


def test_mobilenet_v1_train_eval_and_fine_tune_entry_points(tmp_path):
    from distributed_tensorflow_models_b200.ckpt import store as ckpt
    from distributed_tensorflow_models_b200.eval import mobilenet_v1_eval
    from distributed_tensorflow_models_b200.trainers import mobilenet_v1_train
    common = ["--depth_multiplier=0.25", "--image_size=64", "--batch_size=2", "--num_classes=7", "--synthetic"]
    d1, d2 = str(tmp_path / "train"), str(tmp_path / "ft")
    assert mobilenet_v1_train.main(common + ["--checkpoint_dir=" + d1, "--number_of_steps=2"]) == 0
    latest = ckpt.latest_checkpoint(d1)
    assert latest.endswith("model.ckpt-2")
    tensors = ckpt.load(latest)
    assert tensors["MobilenetV1/Conv2d_1_depthwise/depthwise_weights"].shape == (3, 3, 8, 1)
    assert tensors["MobilenetV1/Logits/Conv2d_1c_1x1/weights"].shape == (1, 1, 256, 7)       # HWIO on disk
    assert int(tensors["global_step"]) == 2
    assert mobilenet_v1_eval.main(common + ["--checkpoint_dir=" + d1, "--num_examples=4",
                                            "--eval_dir=" + str(tmp_path / "eval")]) == 0
    assert (tmp_path / "eval" / "eval.jsonl").read_text().count("recall_at_5") == 1
    # fine-tuning: variables from the first run, global step back to 0, learning rate 1e-4
    assert mobilenet_v1_train.main(common + ["--checkpoint_dir=" + d2, "--fine_tune_checkpoint=" + d1,
                                             "--number_of_steps=1"]) == 0
    assert ckpt.latest_checkpoint(d2).endswith("model.ckpt-1")
    with pytest.raises(SystemExit):
        mobilenet_v1_train.main(["--depth_multiplier=0.3"])


def test_zoo_network_under_two_worker_bsp_on_the_thread_cluster():
    """A zoo network behind the parameter-server protocol: two workers (threads over shared CPU tensors) in BSP take
    ONE global step per pair of gradients, end with identical replicas, and apply the arg scope's L2 on the owner."""
    from distributed_tensorflow_models_b200.parallel.comm import ThreadCluster
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    cl = ThreadCluster(2)

    def work(comm):
        t = Trainer(TrainSetup(model="mobilenet_v1_025", batch_size=2, num_workers=2, image_size=32, num_classes=5,
                               seed=0, cuda_graph=False), comm)
        g = torch.Generator().manual_seed(10 + comm.rank)
        x, y = torch.randn(2, 32, 32, 3, generator=g), torch.randint(0, 5, (2,), generator=g)
        p0 = t.layout.params.clone()
        out = t.train_step(x, y)
        return p0, t.layout.grads.clone(), t.layout.params.clone(), out["global_step"], out["lr"], t.layout

    res = cl.run(work)
    p0, L, lr = res[0][0], res[0][5], res[0][4]
    assert all(r[3] == 1 for r in res) and torch.equal(res[0][2], res[1][2])
    assert lr == pytest.approx(0.045)                       # the zoo trainer's constant, not multiplied by N
    gmean = (res[0][1] + res[1][1]) / 2
    gmean[: L.wd_boundary] += L.weight_decay * p0[: L.wd_boundary]
    assert torch.allclose(res[0][2], p0 - lr * gmean, atol=1e-6)
