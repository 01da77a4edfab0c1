"""Variable-census and shape oracles (the reference pins these in inception/slim/collections_test.py:35-92,
inception_test.py:27-129, vgg/nets/alexnet_test.py:107-132, vgg/nets/vgg_test.py:292-332)."""
import pytest
import torch

from distributed_tensorflow_models_b200.models import inception_v3, registry, resnet
from distributed_tensorflow_models_b200.models.registry import get_network_fn
from distributed_tensorflow_models_b200.slim import VariableStore


def _build(name, shape, nc, is_training=True, **kw):
    st = VariableStore()
    fn = get_network_fn(name, nc, is_training=is_training)
    with torch.no_grad(), st.as_default():
        st.begin_forward()
        out, end_points = fn(torch.randn(*shape), **kw)
    return st, out, end_points


def _nparams(st):
    return sum(v.tensor.numel() for v in st.trainable_variables())


@pytest.mark.parametrize("name,shape,params,ntensors", [
    ("cnn", (2, 24, 24, 3), 1068298, 10),
    ("cifarnet", (2, 32, 32, 3), 1756426, 10),
    ("alexnet_v2", (2, 32, 32, 3), 46247882, 16),
    ("vgg_16", (2, 32, 32, 3), 134301514, 32),
    ("resnet_cifar", (2, 32, 32, 3), 466970, 98),
])
def test_parameter_counts(name, shape, params, ntensors):
    st, out, _ = _build(name, shape, 10)
    assert tuple(out.shape) == (2, 10)
    assert _nparams(st) == params
    assert len(st.trainable_variables()) == ntensors


def test_resnet110_params_and_names():
    st, out, _ = _build("resnet_cifar", (1, 32, 32, 3), 10, resnet_size=110)
    assert _nparams(st) == 1730778
    names = [v.name for v in st.all()]
    assert names[:6] == ["conv2d/kernel", "batch_normalization/gamma", "batch_normalization/beta",
                         "batch_normalization/moving_mean", "batch_normalization/moving_variance", "conv2d_1/kernel"]
    assert names[-2:] == ["dense/kernel", "dense/bias"]
    with pytest.raises(ValueError):
        resnet.cifar10_resnet_v2_generator(33, 10)


def test_resnet32_has_62_moving_stats_and_unprefixed_names():
    st, _, _ = _build("resnet_cifar", (1, 32, 32, 3), 10)
    moving = [v for v in st.all() if not v.trainable]
    assert len(moving) == 62
    assert all("/" in v.name and not v.name.startswith("root/") for v in st.all())


def test_plain_cnn_names_and_decay():
    st, _, _ = _build("cnn", (1, 24, 24, 3), 10)
    names = [v.name for v in st.all()]
    assert names == [f"{s}/{k}" for s in ("conv1", "conv2", "local3", "local4", "softmax_linear")
                     for k in ("weights", "biases")]
    decayed = {v.name: v.weight_decay for v in st.all() if v.weight_decay}
    assert decayed == {"local3/weights": 0.004, "local4/weights": 0.004}
    assert st.get("local3/weights").tf_shape == (2304, 384)


def test_alexnet_variable_names():
    st, _, _ = _build("alexnet_v2", (1, 32, 32, 3), 10)
    expect = [f"alexnet_v2/{l}/{k}" for l in ("conv1", "conv2", "conv3", "conv4", "conv5", "fc6", "fc7", "fc8")
              for k in ("weights", "biases")]
    assert [v.name for v in st.all()] == expect
    assert st.get("alexnet_v2/fc6/weights").tf_shape == (5, 5, 256, 4096)
    assert st.get("alexnet_v2/conv1/weights").tf_shape == (11, 11, 3, 64)


def test_vgg16_variable_names_and_fc6():
    st, _, eps = _build("vgg_16", (1, 32, 32, 3), 10)
    names = [v.name for v in st.all()]
    assert names[0] == "vgg_16/conv1/conv1_1/weights" and names[3] == "vgg_16/conv1/conv1_2/biases"
    assert "vgg_16/conv5/conv5_3/weights" in names and names[-2:] == ["vgg_16/fc8/weights", "vgg_16/fc8/biases"]
    assert st.get("vgg_16/fc6/weights").tf_shape == (7, 7, 512, 4096)     # 102.8 M parameters in one tensor
    assert tuple(eps["pool5"].shape) == (1, 1, 1, 512)


def test_vgg_variants_and_imagenet_resnets_build():
    st, out, _ = _build("vgg_a", (1, 32, 32, 3), 10)
    assert tuple(out.shape) == (1, 10)
    st, out, _ = _build("resnet_v2_18", (1, 64, 64, 3), 7)
    assert tuple(out.shape) == (1, 7)
    with pytest.raises(ValueError):
        get_network_fn("no_such_net", 10)


def test_cifarnet_returns_probabilities_by_default():
    st, out, eps = _build("cifarnet", (3, 32, 32, 3), 10)
    assert torch.allclose(out.sum(1), torch.ones(3), atol=1e-5)           # quirk Q4: softmax output
    assert [v.name for v in st.all()][:2] == ["CifarNet/conv1/weights", "CifarNet/conv1/biases"]


# ---- Inception-v3: the 388-variable census of collections_test.py -----------------------------------------------
@pytest.fixture(scope="module")
def inception_store():
    st = VariableStore()
    with torch.no_grad(), st.as_default():
        st.begin_forward()
        logits, aux = inception_v3.inference(torch.randn(2, 299, 299, 3), 1001, for_training=True)
    return st, logits, aux


def test_inception_census(inception_store):
    st, logits, aux = inception_store
    names = [v.name for v in st.all()]
    assert tuple(logits.shape) == (2, 1001) and tuple(aux.shape) == (2, 1001)
    assert len(names) == 388
    assert sum(n.endswith("weights") for n in names) == 98
    assert sum(n.endswith("biases") for n in names) == 2
    assert sum(n.endswith("beta") for n in names) == 96
    assert sum(n.endswith("gamma") for n in names) == 0
    assert sum(n.endswith("moving_mean") for n in names) == 96
    assert sum(n.endswith("moving_variance") for n in names) == 96
    assert _nparams(st) == 27145970
    assert st.num_regularizers() == 98


def test_inception_per_block_counts(inception_store):
    st, _, _ = inception_store
    expect = {"conv0": 4, "conv1": 4, "conv2": 4, "conv3": 4, "conv4": 4, "mixed_35x35x256a": 28,
              "mixed_35x35x288a": 28, "mixed_35x35x288b": 28, "mixed_17x17x768a": 16, "mixed_17x17x768b": 40,
              "mixed_17x17x768c": 40, "mixed_17x17x768d": 40, "mixed_17x17x768e": 40, "aux_logits": 10,
              "mixed_17x17x1280a": 24, "mixed_8x8x2048a": 36, "mixed_8x8x2048b": 36, "logits": 2}
    for scope, n in expect.items():
        assert len(st.get_variables(scope + "/")) == n, scope
    assert "mixed_35x35x256a/branch5x5/Conv_1/BatchNorm/moving_mean" in [v.name for v in st.all()]
    assert "aux_logits/FC/weights" in [v.name for v in st.all()] and "logits/logits/biases" in [v.name for v in st.all()]


def test_inception_endpoints_and_half_size():
    st = VariableStore()
    with torch.no_grad(), st.as_default():
        st.begin_forward()
        with inception_v3.inception_v3_parameters():
            _, eps = inception_v3.inception_v3(torch.randn(1, 299, 299, 3), num_classes=1000, is_training=False)
    assert tuple(eps["mixed_8x8x2048b"].shape) == (1, 8, 8, 2048)
    assert tuple(eps["mixed_35x35x256a"].shape) == (1, 35, 35, 256)
    assert tuple(eps["mixed_17x17x1280a"].shape) == (1, 8, 8, 1280)
    assert tuple(eps["logits"].shape) == (1, 1000)
    # same variables serve a second (eval) pass: reuse, no growth
    n = len(st.all())
    with torch.no_grad(), st.as_default():
        st.begin_forward()
        with inception_v3.inception_v3_parameters():
            inception_v3.inception_v3(torch.randn(1, 299, 299, 3), num_classes=1000, is_training=False)
    assert len(st.all()) == n


def test_model_specs_cover_reference_defaults():
    s = registry.MODEL_SPECS
    assert {k: v.batch_size for k, v in s.items()} == {"cnn": 512, "cifarnet": 512, "alexnet": 128, "vgg": 72,
                                                       "resnet": 128, "inception": 32}
    assert s["cnn"].image_size == 24 and s["inception"].num_classes == 1001


def test_zoo_lenet_and_overfeat():
    """Small zoo members reachable through the factory (vgg/nets/nets_factory.py:39-100)."""
    st, out, eps = _build("lenet", (2, 28, 28, 1), 10)
    assert tuple(out.shape) == (2, 10)
    names = [v.name for v in st.all()]
    assert names == ["LeNet/conv1/weights", "LeNet/conv1/biases", "LeNet/conv2/weights", "LeNet/conv2/biases",
                     "LeNet/fc3/weights", "LeNet/fc3/biases", "LeNet/fc4/weights", "LeNet/fc4/biases"]
    # 5*5*1*32+32 + 5*5*32*64+64 + 7*7*64*1024+1024 + 1024*10+10
    assert _nparams(st) == 832 + 51264 + 3212288 + 10250
    assert tuple(eps["pool2"].shape) == (2, 7, 7, 64)
    st, out, eps = _build("overfeat", (1, 231, 231, 3), 1000)
    assert tuple(out.shape) == (1, 1000)
    assert tuple(eps["pool5"].shape) == (1, 6, 6, 1024) and tuple(eps["fc7"].shape) == (1, 1, 1, 4096)
    assert st.get("overfeat/fc6/weights").tf_shape == (6, 6, 1024, 3072)
    assert get_network_fn("overfeat", 1000).default_image_size == 231
