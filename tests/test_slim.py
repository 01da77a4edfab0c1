"""slim library semantics (modelled on inception/slim/{scopes,variables,ops,losses}_test.py)."""
import math

import pytest
import torch

from distributed_tensorflow_models_b200.slim import VariableStore, losses, ops, scopes
from distributed_tensorflow_models_b200.slim import variables as V


# ---- arg_scope ----------------------------------------------------------------------------------------------
@scopes.add_arg_scope
def _f1(*args, **kwargs):
    return args, kwargs


@scopes.add_arg_scope
def _f2(*args, **kwargs):
    return args, kwargs


def test_arg_scope_defaults_override_and_nesting():
    assert scopes.has_arg_scope(_f1)
    with scopes.arg_scope([_f1], a=1, b=None):
        assert _f1(0)[1] == {"a": 1, "b": None}
        assert _f1(0, a=2)[1] == {"a": 2, "b": None}             # call-site value wins
        with scopes.arg_scope([_f1], b=5, c=[1]):
            assert _f1(0)[1] == {"a": 1, "b": 5, "c": [1]}     # nested scopes merge
        assert _f1(0)[1] == {"a": 1, "b": None}                 # and unwind
        assert _f2(0)[1] == {}                                  # other ops untouched
    assert _f1(0)[1] == {}


def test_arg_scope_reuse_by_dict_and_errors():
    with scopes.arg_scope([_f1, _f2], a=1) as sc:
        pass
    with scopes.arg_scope(sc):
        assert _f1(0)[1] == {"a": 1} and _f2(0)[1] == {"a": 1}
    with pytest.raises(ValueError):
        with scopes.arg_scope(sc, a=2):
            pass
    with pytest.raises(TypeError):
        with scopes.arg_scope(_f1, a=1):
            pass

    def undecorated():
        pass
    with pytest.raises(ValueError):
        with scopes.arg_scope([undecorated], a=1):
            pass


# ---- variables: naming, collections, placement ---------------------------------------------------------------
def test_variable_naming_and_unique_default_scopes():
    st = VariableStore()
    with st.as_default():
        st.begin_forward()
        x = torch.randn(2, 8, 8, 4)
        with st.scope("tower"):
            ops.conv2d(x, 8, [3, 3])
            ops.conv2d(x, 8, [3, 3])
            ops.conv2d(x, 8, [3, 3], scope="named")
    names = [v.name for v in st.all()]
    assert "tower/Conv/weights" in names and "tower/Conv_1/weights" in names and "tower/named/biases" in names
    n = len(names)
    with st.as_default():                         # second pass resolves to the same names
        st.begin_forward()
        with st.scope("tower"):
            ops.conv2d(x, 8, [3, 3])
            ops.conv2d(x, 8, [3, 3])
    assert len(st.all()) == n
    st.freeze()
    with st.as_default(), pytest.raises(RuntimeError):
        st.begin_forward()
        ops.conv2d(x, 8, [3, 3], scope="brand_new")


def test_collections_and_restore_lists():
    st = VariableStore()
    with st.as_default():
        a = V.variable("a", (3,))
        b = V.variable("b", (3,), restore=False)
        c = V.variable("c", (3,), trainable=False, collections=["mine"])
        gs = V.global_step()
    assert [v.name for v in st.get_variables()] == ["a", "b", "c", "global_step"]
    assert [v.name for v in st.get_variables_to_restore()] == ["a", "c", "global_step"]
    assert [v.name for v in st.get_collection("mine")] == ["c"]
    assert [v.name for v in st.trainable_variables()] == ["a", "b"]
    assert gs.dtype == torch.int64 and st.get_unique_variable("a").tensor is a
    assert [v.name for v in st.get_variables_by_name("b")] == ["b"]
    with pytest.raises(ValueError):
        st.get_unique_variable("nope")


def test_variable_device_chooser_round_robin():
    """variables_test.py:167-213: a,b,d,e alternate over /job:ps/task:0|1/cpu:0."""
    st = VariableStore(device_fn=V.VariableDeviceChooser(num_parameter_servers=2))
    with st.as_default():
        for n in "abde":
            V.variable(n, (1,))
    assert [st.get(n).device for n in "abde"] == ["/job:ps/task:0/CPU:0", "/job:ps/task:1/CPU:0"] * 2
    assert [st.get(n).ps_task for n in "abde"] == [0, 1, 0, 1]
    assert V.VariableDeviceChooser()("x") == "/CPU:0"            # no PS: stays local
    setter = V.replica_device_setter(3)
    assert [setter(), setter(), setter(), setter()] == ["/job:ps/task:0", "/job:ps/task:1", "/job:ps/task:2",
                                                         "/job:ps/task:0"]


def test_fixed_size_partitioner():
    part = V.fixed_size_partitioner(2, axis=0)
    assert part((7, 7, 512, 4096)) == [4, 3]          # VGG fc6 with two PS tasks
    assert part((1, 1, 4096, 10)) == [1]              # 1x1 kernels cannot be split
    assert V.fixed_size_partitioner(4)((10,)) == [3, 3, 2, 2]


# ---- ops ------------------------------------------------------------------------------------------------------
def test_conv_fc_shapes_bias_rule_and_bn_variables():
    st = VariableStore()
    x = torch.randn(5, 9, 9, 3)
    with st.as_default():
        st.begin_forward()
        y = ops.conv2d(x, 32, [3, 3], stride=2, padding="SAME", scope="c1")
        assert tuple(y.shape) == (5, 5, 5, 32)
        y = ops.conv2d(x, 32, [3, 3], padding="VALID", batch_norm_params={"decay": 0.9}, scope="c2")
        assert tuple(y.shape) == (5, 7, 7, 32)
        z = ops.fc(ops.flatten(y), 11, scope="f")
        assert tuple(z.shape) == (5, 11)
    names = [v.name for v in st.all()]
    assert "c1/biases" in names and "c2/biases" not in names           # bias only without batch-norm
    assert {"c2/BatchNorm/beta", "c2/BatchNorm/moving_mean", "c2/BatchNorm/moving_variance"} <= set(names)
    assert "c2/BatchNorm/gamma" not in names                            # scale=False by default
    assert st.get("c1/weights").tf_shape == (3, 3, 3, 32) and st.get("f/weights").tf_shape == (1568, 11)


def test_batch_norm_moving_stats_follow_batch_stats():
    """ops_test.py testComputeMovingVars / testEvalMovingVars."""
    st = VariableStore()
    x = torch.randn(64, 3, 3, 4) * 3.0 + 2.0
    with st.as_default():
        for _ in range(10):
            st.begin_forward()
            y = ops.batch_norm(x, decay=0.1, scope="bn")
    mean, var = x.reshape(-1, 4).mean(0), x.reshape(-1, 4).var(0, unbiased=False)
    assert torch.allclose(st.get("bn/moving_mean").tensor, mean, atol=1e-3)
    assert torch.allclose(st.get("bn/moving_variance").tensor, var, rtol=1e-3)
    assert torch.allclose(y.reshape(-1, 4).mean(0), torch.zeros(4), atol=1e-5)
    with st.as_default():
        st.begin_forward()
        ye = ops.batch_norm(x, decay=0.1, is_training=False, scope="bn")
    assert torch.allclose(ye, y, atol=2e-3)                               # eval uses the moving statistics


def test_pool_dropout_onehot_flatten_repeat():
    x = torch.randn(2, 7, 7, 8)
    assert tuple(ops.max_pool(x, [3, 3]).shape) == (2, 3, 3, 8)           # stride 2, VALID by default
    assert tuple(ops.max_pool(x, [3, 3], padding="SAME").shape) == (2, 4, 4, 8)
    assert tuple(ops.avg_pool(x, [3, 3], stride=1, padding="SAME").shape) == (2, 7, 7, 8)
    ones = torch.ones(1, 4, 4, 4)
    assert torch.allclose(ops.avg_pool(ones, [3, 3], stride=1, padding="SAME"), ones)   # padding not averaged in
    assert ops.dropout(x, 0.5, is_training=False) is x
    d = ops.dropout(torch.ones(10000), 0.8, is_training=True)
    assert abs(float((d > 0).float().mean()) - 0.8) < 0.03 and abs(float(d.max()) - 1.25) < 1e-6
    oh = ops.one_hot_encoding(torch.tensor([0, 2]), 3)
    assert oh.tolist() == [[1, 0, 0], [0, 0, 1]]
    assert tuple(ops.flatten(x).shape) == (2, 392)
    st = VariableStore()
    with st.as_default():
        st.begin_forward()
        y = ops.repeat_op(3, torch.randn(1, 5, 5, 4), ops.conv2d, 4, [3, 3], scope="rep")
    assert [v.name for v in st.all() if v.name.endswith("weights")] == ["rep/Conv/weights", "rep/Conv_1/weights",
                                                                        "rep/Conv_2/weights"]


def test_same_padding_arithmetic_matches_tensorflow():
    from distributed_tensorflow_models_b200.ops.functional import same_pad_1d
    assert same_pad_1d(24, 3, 2) == (12, 0, 1)       # extra pixel goes after
    assert same_pad_1d(299, 3, 2) == (150, 1, 1)
    assert same_pad_1d(32, 5, 1) == (32, 2, 2)
    assert same_pad_1d(1, 7, 1) == (1, 3, 3)


def test_lrn_matches_definition():
    from distributed_tensorflow_models_b200.ops import functional as Fn
    x = torch.randn(2, 3, 3, 16)
    y = Fn.lrn(x, 4, 1.0, 0.001 / 9.0, 0.75)
    c = 7
    ss = (x[..., c - 4:c + 5] ** 2).sum(-1)
    assert torch.allclose(y[..., c], x[..., c] / (1.0 + 0.001 / 9.0 * ss) ** 0.75, atol=1e-6)


# ---- losses ---------------------------------------------------------------------------------------------------
def test_regularizers_and_losses():
    w = torch.tensor([[1.0, -2.0], [3.0, -4.0]])
    assert float(losses.l1_regularizer(0.5)(w)) == pytest.approx(5.0)
    assert float(losses.l2_regularizer(0.5)(w)) == pytest.approx(0.5 * 30.0 / 2)
    assert float(losses.l1_l2_regularizer(1.0, 1.0)(w)) == pytest.approx(10.0 + 15.0)
    assert float(losses.l2_loss(w, 2.0)) == pytest.approx(30.0)


def test_cross_entropy_all_correct_all_wrong_weight_and_smoothing():
    logits = torch.tensor([[10.0, 0, 0], [0, 10.0, 0], [0, 0, 10.0]])
    right = torch.eye(3)
    wrong = torch.tensor([[0, 0, 1.0], [1.0, 0, 0], [0, 1.0, 0]])
    st = VariableStore()
    with st.as_default():
        st.begin_forward()
        assert float(losses.cross_entropy_loss(logits, right)) == pytest.approx(0.0, abs=1e-3)
        assert float(losses.cross_entropy_loss(logits, wrong)) == pytest.approx(10.0, abs=1e-3)
        assert float(losses.cross_entropy_loss(logits, wrong, weight=0.4)) == pytest.approx(4.0, abs=1e-3)
        assert len(st.losses) == 3                                           # the `_losses` collection
        sm = losses.cross_entropy_loss(logits, right, label_smoothing=0.1)
        target = right * 0.9 + 0.1 / 3
        ref = -(target * torch.log_softmax(logits, 1)).sum(1).mean()
        assert float(sm) == pytest.approx(float(ref), rel=1e-5)
        idx = losses.cross_entropy_loss(logits, torch.tensor([0, 1, 2]), label_smoothing=0.1)
        assert float(idx) == pytest.approx(float(ref), rel=1e-5)             # sparse labels = dense one-hot
    with pytest.raises(ValueError):
        losses.cross_entropy_loss(logits, torch.zeros(3, 4))


def test_inception_loss_weights():
    from distributed_tensorflow_models_b200.models import inception_v3
    logits, aux = torch.randn(4, 1001), torch.randn(4, 1001)
    labels = torch.tensor([1, 5, 7, 1000])
    total = inception_v3.loss((logits, aux), labels)
    from distributed_tensorflow_models_b200.ops.functional import softmax_cross_entropy as xe
    assert float(total) == pytest.approx(float(xe(logits, labels, 0.1, 1.0) + xe(aux, labels, 0.1, 0.4)), rel=1e-6)


def test_contrib_conv_honours_batch_norm_scope_bias_switch_and_separable_variables():
    """slim.conv2d(normalizer_fn=slim.batch_norm) picks `is_training` / decay / scale up from an enclosing
    arg_scope([batch_norm], ...) — the way every zoo network switches between training and inference."""
    import torch
    from distributed_tensorflow_models_b200.slim import VariableStore, ops, scopes
    x = torch.randn(4, 6, 6, 3)
    st = VariableStore()
    with st.as_default(), scopes.arg_scope([ops.batch_norm], is_training=False, scale=True, decay=0.5):
        y = ops.convolution2d(x, 5, [3, 3], normalizer_fn=ops.batch_norm, scope="c")
    names = [v.name for v in st.all()]
    assert names == ["c/weights", "c/BatchNorm/beta", "c/BatchNorm/gamma", "c/BatchNorm/moving_mean",
                     "c/BatchNorm/moving_variance"]
    # inference mode: moving statistics (mean 0, variance 1) were used and left untouched
    assert float(st.get("c/BatchNorm/moving_mean").tensor.abs().max()) == 0.0
    ref = torch.relu(torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), st.get("c/weights").tensor.permute(0, 3, 1, 2),
                                                padding=1) / (1 + 1e-3) ** 0.5).permute(0, 2, 3, 1)
    assert torch.allclose(y, ref, atol=1e-5)
    with st.as_default(), scopes.arg_scope([ops.batch_norm], is_training=True, scale=True, decay=0.5):
        st.begin_forward()
        ops.convolution2d(x, 5, [3, 3], normalizer_fn=ops.batch_norm, scope="c")
    assert float(st.get("c/BatchNorm/moving_mean").tensor.abs().max()) > 0.0          # training mode updates them
    st2 = VariableStore()
    with st2.as_default():
        ops.convolution2d(x, 5, [1, 1], use_bias=False, scope="bare")
        ops.convolution2d(x, 5, [1, 1], scope="biased")
        ops.separable_convolution2d(x, None, [3, 3], depth_multiplier=2, normalizer_fn=ops.batch_norm, scope="dw")
        up = ops.convolution2d_transpose(x, 4, [4, 4], stride=2, scope="up")
        inorm = ops.instance_norm(x, scope="in")
    assert [v.name for v in st2.all()][:3] == ["bare/weights", "biased/weights", "biased/biases"]
    assert tuple(st2.get("dw/depthwise_weights").tensor.shape) == (3, 3, 3, 2) and not st2.has("dw/pointwise_weights")
    assert tuple(up.shape) == (4, 12, 12, 4) and tuple(st2.get("up/weights").to_tf().shape) == (4, 4, 4, 3)
    assert {"in/beta", "in/gamma"} <= {v.name for v in st2.all()} and tuple(inorm.shape) == tuple(x.shape)


def test_docs_examples_run():
    """Every python block of docs/SLIM.md executes, in order, in one namespace."""
    import os
    import re
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "docs", "SLIM.md")
    blocks = re.findall(r"```python\n(.*?)```", open(path).read(), re.S)
    assert len(blocks) >= 4
    ns: dict = {}
    for code in blocks:
        exec(compile(code, path, "exec"), ns)
