"""Quantisation-aware training (``--quantize``; the reference applies tf.contrib.quantize's graph rewrites in
vgg/nets/mobilenet_v1_train.py:135-139 and mobilenet_v1_eval.py:126): fake-quant arithmetic against
TensorFlow's documented formulas, batch-norm folding, the delay switch, checkpoint names, entry points."""
import numpy as np
import pytest
import torch

from distributed_tensorflow_models_b200.ops import functional as Fn
from distributed_tensorflow_models_b200.slim import ops, quantize as Q
from distributed_tensorflow_models_b200.slim.variables import VariableStore


def test_fake_quant_matches_the_min_max_vars_formulas():
    torch.manual_seed(0)
    x = (torch.randn(4000) * 3).requires_grad_(True)
    # [-6, 6], 8 bits: step 12/255, zero point 127.5 -> 128, so the range moves to [-128, 127] * step
    rng = torch.tensor([-6.0, 6.0])
    nmin, nmax, step = Fn.nudged_range(rng[0], rng[1], 8, False)
    assert float(step) == pytest.approx(12 / 255) and float(nmin) == pytest.approx(-128 * 12 / 255)
    assert float(nmax) == pytest.approx(127 * 12 / 255)
    y = Fn.fake_quant(x, rng, None, 8, False)
    ref = torch.fake_quantize_per_tensor_affine(x.detach(), 12 / 255, 128, 0, 255)
    assert torch.allclose(y, ref, atol=1e-6)
    assert len(torch.unique(y.detach())) <= 256
    # every output is a whole number of steps and 0.0 is one of them
    assert torch.allclose(y.detach() / step, torch.round(y.detach() / step), atol=1e-4)
    assert float(Fn.fake_quant(torch.zeros(1), rng)) == 0.0
    # straight-through gradient inside the nudged range, none outside
    y.sum().backward()
    inside = (x.detach() >= nmin) & (x.detach() <= nmax)
    assert torch.equal(x.grad, inside.float())
    # narrow range drops the lowest step: 255 levels, symmetric for a symmetric range
    nmin_n, nmax_n, _ = Fn.nudged_range(rng[0], rng[1], 8, True)
    assert float(nmin_n) == pytest.approx(-float(nmax_n))
    # an all-positive range is nudged down to start at zero
    lo, hi, _ = Fn.nudged_range(torch.tensor(0.3), torch.tensor(5.0), 8, False)
    assert float(lo) == 0.0 and float(hi) == pytest.approx(4.7)
    # closed gate / degenerate range: identity, gradient one everywhere
    x2 = x.detach().clone().requires_grad_(True)
    y2 = Fn.fake_quant(x2, rng, torch.zeros(()), 8, False)
    assert torch.equal(y2, x2)
    y2.sum().backward()
    assert torch.equal(x2.grad, torch.ones_like(x2))
    assert torch.equal(Fn.fake_quant(x.detach(), torch.tensor([1.0, 1.0])), x.detach())


def test_quant_range_tracking_last_value_and_moving_average():
    x = torch.tensor([0.5, 2.0, 3.0])
    rng = torch.tensor([-6.0, 6.0])
    Fn.quant_range_update(x, rng, -1.0)
    assert rng.tolist() == [0.5, 3.0]
    rng = torch.tensor([-6.0, 6.0])
    Fn.quant_range_update(x, rng, 0.9)             # batch range widened to contain zero: [0, 3]
    assert rng.tolist() == pytest.approx([-5.4, 5.7])


def _net(x, is_training, quant_store=None):
    bn = {"is_training": is_training, "scale": True, "decay": 0.9, "epsilon": 1e-3}
    y = ops.convolution2d(x, 8, [3, 3], normalizer_fn=ops.batch_norm, normalizer_params=bn, activation_fn=ops.relu6,
                          scope="c1")
    y = ops.separable_convolution2d(y, None, [3, 3], normalizer_fn=ops.batch_norm, normalizer_params=bn,
                                    activation_fn=ops.relu6, scope="dw")
    y = ops.separable_convolution2d(y, 12, [3, 3], stride=2, activation_fn=None, scope="sep")
    y = y.mean(dim=(1, 2))
    return ops.fully_connected(y, 5, activation_fn=None, scope="fc")


def _run(store, x, is_training):
    with store.as_default():
        store.begin_forward()
        if getattr(store, "quant", None) is not None:
            store.quant.is_training = is_training
        return _net(x, is_training)


def test_quantised_layers_fold_batch_norm_and_keep_the_float_network_within_quantisation_noise():
    torch.manual_seed(1)
    x = torch.randn(4, 12, 12, 3)
    plain, quant = VariableStore(seed=3), VariableStore(seed=3)
    quant.quant = Q.QuantState(gate=torch.ones(()))
    y_plain = _run(plain, x, True)
    y_quant = _run(quant, x, True)
    # same float variables in the same places; the quantised store has one range per weight tensor / layer output
    names_p, names_q = [v.name for v in plain.all()], [v.name for v in quant.all()]
    assert [n for n in names_q if "_quant" not in n] == names_p
    assert [n for n in names_q if "_quant" in n] == [
        "c1/weights_quant/min_max", "c1/act_quant/min_max", "dw/weights_quant/min_max", "dw/act_quant/min_max",
        "sep/depthwise_weights_quant/min_max", "sep/pointwise_weights_quant/min_max", "sep/act_quant/min_max",
        "fc/weights_quant/min_max", "fc/act_quant/min_max"]
    for n in names_p:
        if "moving" not in n:
            assert torch.equal(plain.get(n).tensor, quant.get(n).tensor)
    # 8-bit rounding perturbs, it does not change the function
    assert float((y_quant - y_plain).detach().abs().max()) < 0.25 * float(y_plain.detach().abs().max()) + 0.05
    assert not torch.equal(y_quant, y_plain)
    # weight range = min / max of the FOLDED weights (gamma / sqrt(moving_var + eps), moving_var taken before the step)
    w, g = quant.get("c1/weights").tensor, quant.get("c1/BatchNorm/gamma").tensor
    folded = w * (g * torch.rsqrt(torch.ones(8) + 1e-3)).reshape(8, 1, 1, 1)
    assert quant.get("c1/weights_quant/min_max").tensor.tolist() == pytest.approx(
        [float(folded.detach().min()), float(folded.detach().max())], rel=1e-5)
    # activation range: one moving-average step from [-6, 6] towards the batch range of a relu6 output
    lo, hi = quant.get("c1/act_quant/min_max").tensor.tolist()
    assert lo == pytest.approx(-6 * 0.999) and 5.99 < hi < 6.0
    # gradients reach every trainable variable through the rounding (straight-through) and the fold
    y_quant.square().mean().backward()
    for v in quant.trainable_variables():
        assert v.tensor.grad is not None and torch.isfinite(v.tensor.grad).all(), v.name
        assert float(v.tensor.grad.abs().sum()) > 0, v.name


def test_quantised_evaluation_is_the_folded_inference_form():
    """In evaluation mode conv(x, Q(w * s)) / s through batch norm equals conv(x, Q(w * s)) + (beta - s * mean):
    what an integer engine computes from the folded, quantised weights."""
    torch.manual_seed(2)
    x = torch.randn(2, 8, 8, 3)
    st = VariableStore(seed=5)
    st.quant = Q.QuantState()
    for _ in range(3):
        _run(st, x, True)                        # move the statistics and ranges away from their initial values
    before = {v.name: v.tensor.clone() for v in st.all()}
    with torch.no_grad():
        with st.as_default():
            st.begin_forward()
            st.quant.is_training = False
            bn = {"is_training": False, "scale": True, "decay": 0.9, "epsilon": 1e-3}
            y = ops.convolution2d(x, 8, [3, 3], normalizer_fn=ops.batch_norm, normalizer_params=bn,
                                  activation_fn=None, scope="c1")
    for v in st.all():                           # evaluation moves nothing
        assert torch.equal(before[v.name], v.tensor), v.name
    g = st.get("c1/BatchNorm/gamma").tensor.detach()
    beta = st.get("c1/BatchNorm/beta").tensor.detach()
    mm, mv = st.get("c1/BatchNorm/moving_mean").tensor, st.get("c1/BatchNorm/moving_variance").tensor
    s = g * torch.rsqrt(mv + 1e-3)
    wq = Fn.fake_quant(st.get("c1/weights").tensor.detach() * s.reshape(8, 1, 1, 1),
                       st.get("c1/weights_quant/min_max").tensor, None, 8, True)
    folded = Fn.conv2d(x, wq, beta - s * mm, 1, "SAME")
    expect = Fn.fake_quant(folded, st.get("c1/act_quant/min_max").tensor, None, 8, False)
    step = float(Fn.nudged_range(*st.get("c1/act_quant/min_max").tensor, 8, False)[2])
    # equal up to one quantisation step where float rounding puts a value on the other side of a half step
    assert float((y - expect).abs().max()) <= step * 1.001
    assert float(((y - expect).abs() > 1e-5).float().mean()) < 0.02


def test_trainer_quant_delay_checkpoint_names_and_entry_points(tmp_path):
    from distributed_tensorflow_models_b200.ckpt import store as ckpt
    from distributed_tensorflow_models_b200.eval import mobilenet_v1_eval
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    from distributed_tensorflow_models_b200.trainers import mobilenet_v1_train
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(8, 64, 64, 3, generator=g), torch.randint(0, 5, (8,), generator=g)

    def trainer(**kw):
        return Trainer(TrainSetup(model="mobilenet_v1_025", batch_size=8, image_size=64, num_classes=5, seed=0,
                                  initial_lr=0.02, cuda_graph=False, **kw))
    plain, delayed, on = trainer(), trainer(quantize=True, quant_delay=2), trainer(quantize=True)
    assert len(on.store.all()) == len(plain.store.all()) + 2 * 28
    # same trainable layout: the ranges are state, not parameters
    assert on.layout.params.numel() == plain.layout.params.numel()
    lp = [float(plain.train_step(x, y)["loss"]) for _ in range(3)]
    ld, gates = [], []
    for _ in range(3):
        ld.append(float(delayed.train_step(x, y)["loss"]))
        gates.append(float(delayed.store.quant.gate))
    lo = [float(on.train_step(x, y)["loss"]) for _ in range(3)]
    # the switch opens when the global step reaches the delay; before that the quantised trainer is the float
    # trainer up to the rounding of w * s / s ...
    assert gates == [0.0, 0.0, 1.0]
    assert ld[0] == pytest.approx(lp[0], rel=5e-3)
    # ... while its ranges already follow the data
    assert delayed.store.get("MobilenetV1/Conv2d_0/act_quant/min_max").tensor[0] > -6.0
    assert abs(lo[0] - lp[0]) > abs(ld[0] - lp[0]) and all(np.isfinite(lo))
    rng = on.store.get("MobilenetV1/Conv2d_0/weights_quant/min_max").tensor
    assert float(rng[0]) < 0 < float(rng[1]) and float(rng[1] - rng[0]) < 12          # no longer the initial [-6, 6]
    # quantised training still learns
    more = [float(on.train_step(x, y)["loss"]) for _ in range(12)]
    assert min(more[-3:]) < 0.7 * lo[0]
    # checkpoints carry TensorFlow's two scalars per range and restore into the 2-vectors
    tensors = ckpt.collect_tensors(on.store, global_step=15)
    assert tensors["MobilenetV1/Conv2d_0/weights_quant/min"].shape == ()
    assert float(tensors["MobilenetV1/Conv2d_0/act_quant/max"]) > 0
    assert not any(k.endswith("min_max") for k in tensors)
    fresh = trainer(quantize=True)
    ckpt.restore(fresh.store, tensors)
    for v in on.store.all():
        assert torch.equal(v.tensor, fresh.store.get(v.name).tensor), v.name
    # a float checkpoint initialises a quantised run (fine-tuning); strict restore names what is missing
    float_tensors = ckpt.collect_tensors(plain.store, global_step=3)
    assert len(ckpt.restore(trainer(quantize=True).store, float_tensors, strict=False)) == len(plain.store.all())
    with pytest.raises(KeyError):
        ckpt.restore(trainer(quantize=True).store, float_tensors, strict=True)
    # the reference's command lines: --quantize on the MobileNet scripts, with the script's own delay rule
    argv, net, _ = mobilenet_v1_train.translate(["--quantize=True", "--depth_multiplier=0.25"])
    assert net == "mobilenet_v1_025" and argv[argv.index("--quant_delay") + 1] == "250000"
    argv, _, _ = mobilenet_v1_train.translate(["--quantize", "--fine_tune_checkpoint=/x"])
    assert argv[argv.index("--quant_delay") + 1] == "0"
    assert "--quantize" not in mobilenet_v1_train.translate([])[0]
    common = ["--depth_multiplier=0.25", "--image_size=64", "--batch_size=2", "--num_classes=5", "--synthetic",
              "--quantize=True"]
    d = str(tmp_path / "q")
    assert mobilenet_v1_train.main(common + ["--checkpoint_dir=" + d, "--number_of_steps=2", "--quant_delay=1"]) == 0
    saved = ckpt.load(ckpt.latest_checkpoint(d))
    assert "MobilenetV1/Logits/Conv2d_1c_1x1/act_quant/min" in saved
    assert mobilenet_v1_eval.main(common + ["--checkpoint_dir=" + d, "--num_examples=4",
                                            "--eval_dir=" + str(tmp_path / "eval")]) == 0
    assert (tmp_path / "eval" / "eval.jsonl").read_text().count("precision_at_1") == 1


@pytest.mark.parametrize("model,adds", [("mobilenet_v2", 10), ("resnet_v1_50", 16), ("resnet_v2_50_slim", 16)])
def test_residual_additions_get_a_range_of_their_own(model, adds):
    """TensorFlow's rewriter quantises the output of every bypass addition; here the models ask for it themselves."""
    from distributed_tensorflow_models_b200.models import registry
    net = registry.get_network_fn(model, 7, is_training=True)
    size = 64
    plain, quant = VariableStore(seed=1), VariableStore(seed=1)
    quant.quant = Q.QuantState()
    x = torch.randn(2, size, size, 3, generator=torch.Generator().manual_seed(0))
    outs = []
    for st in (plain, quant):
        with st.as_default():
            st.begin_forward()
            out = net(x)[0]
            outs.append(out[0] if isinstance(out, tuple) else out)
    assert outs[0].shape == outs[1].shape and torch.isfinite(outs[1]).all()
    names = [v.name for v in quant.all()]
    assert sum(n.endswith("add_quant/min_max") for n in names) == adds
    assert [n for n in names if "_quant" not in n] == [v.name for v in plain.all()]


def test_first_class_models_quantise_through_the_old_slim_layers_or_say_that_they_cannot():
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(4, 32, 32, 3, generator=g), torch.randint(0, 10, (4,), generator=g)
    t = Trainer(TrainSetup(model="alexnet", batch_size=4, cuda_graph=False, quantize=True))
    names = [v.name for v in t.store.all() if "_quant" in v.name]
    assert len(names) == 16 and names[0] == "partitioned_space/alexnet_v2/conv1/weights_quant/min_max"
    losses_ = [float(t.train_step(x, y)["loss"]) for _ in range(4)]
    assert all(np.isfinite(losses_)) and losses_[-1] < losses_[0]
    for model in ("cnn", "resnet"):
        with pytest.raises(ValueError, match="--quantize"):
            Trainer(TrainSetup(model=model, batch_size=4, cuda_graph=False, quantize=True))
