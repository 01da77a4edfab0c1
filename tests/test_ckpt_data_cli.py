"""Checkpoint store, eval loop, input pipelines, flags/presets and launcher helpers."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from distributed_tensorflow_models_b200 import config as cfg
from distributed_tensorflow_models_b200.ckpt import store as ckpt
from distributed_tensorflow_models_b200.data import cifar10, imagenet, synthetic, tfrecord
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.trainers import common
from distributed_tensorflow_models_b200.utils import logs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- checkpoints ------------------------------------------------------------------------------------------------
def test_checkpoint_names_layouts_slots_and_state_file(tmp_path):
    t = Trainer(TrainSetup(model="cifarnet", batch_size=4))
    x = torch.randn(4, 32, 32, 3)
    y = torch.randint(0, 10, (4,))
    t.train_step(x, y)
    path = common.save_checkpoint(t, str(tmp_path), max_to_keep=2)
    assert path.endswith("model.ckpt-1") and os.path.exists(path + ".npz")
    tensors = ckpt.load(path)
    # TF names under the trainer's variable scope, TF layouts (HWIO / [in, out])
    assert tensors["partitioned_space/CifarNet/conv1/weights"].shape == (5, 5, 3, 64)
    assert tensors["partitioned_space/CifarNet/fc3/weights"].shape == (4096, 384)
    assert tensors["partitioned_space/CifarNet/conv1/weights/ExponentialMovingAverage"].shape == (5, 5, 3, 64)
    assert int(tensors["partitioned_space/Variable"]) == 1
    st = ckpt.get_checkpoint_state(str(tmp_path))
    assert st.model_checkpoint_path == path and ckpt.step_of(path) == 1
    assert 'model_checkpoint_path: "model.ckpt-1"' in open(tmp_path / "checkpoint").read()
    # keep-N
    for _ in range(3):
        t.train_step(x, y)
        common.save_checkpoint(t, str(tmp_path), max_to_keep=2)
    kept = sorted(f for f in os.listdir(tmp_path) if f.endswith(".npz"))
    assert kept == ["model.ckpt-3.npz", "model.ckpt-4.npz"]
    # round trip into a fresh replica, raw and EMA-restored
    t2 = Trainer(TrainSetup(model="cifarnet", batch_size=4, seed=99))
    tensors = ckpt.load(ckpt.latest_checkpoint(str(tmp_path)))
    ckpt.restore(t2.store, tensors)
    assert torch.equal(t2.layout.params[: t.layout.used], t.layout.params[: t.layout.used])
    ckpt.restore(t2.store, tensors, use_ema=True)
    ema = t.comm.opts[0].ema
    assert torch.allclose(t2.layout.params[: t.layout.used], ema[: t.layout.used])


def test_rmsprop_slot_names_and_inception_global_step(tmp_path):
    t = Trainer(TrainSetup(model="cnn", batch_size=2, preset=cfg.PRESETS["cnn"].__class__(
        **{**cfg.PRESETS["cnn"].__dict__, "optimizer": "rmsprop"})))
    t.train_step(torch.randn(2, 24, 24, 3), torch.randint(0, 10, (2,)))
    slots = t.comm.gather_slot_flats()
    tensors = ckpt.collect_tensors(t.store, slots, t.layout, 7, "global_step")
    assert "conv1/weights/RMSProp" in tensors and "conv1/weights/RMSProp_1" in tensors
    assert tensors["conv1/weights/RMSProp"].shape == (5, 5, 3, 64) and int(tensors["global_step"]) == 7


def test_eval_once_on_saved_checkpoint(tmp_path):
    rc = subprocess.run([sys.executable, "-m", "distributed_tensorflow_models_b200.trainers.cifar10_resnet_bsp",
                         "--max_steps=2", "--batch_size=4", f"--train_dir={tmp_path}/t", "--synthetic"],
                        cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert rc.returncode == 0, rc.stderr
    assert "(gs 2), loss=" in rc.stderr and "samples/s" in rc.stderr               # resnet log format
    assert "train-batch accuracy" in rc.stderr
    rc = subprocess.run([sys.executable, "-m", "distributed_tensorflow_models_b200.eval.cifar10_resnet_eval",
                         f"--checkpoint_dir={tmp_path}/t", f"--eval_dir={tmp_path}/e", "--run_once", "--num_examples=16",
                         "--synthetic"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert rc.returncode == 0, rc.stderr
    assert "precision @ 1 = " in rc.stderr and "global_step: 2" in rc.stderr
    assert os.path.exists(tmp_path / "e" / "eval.jsonl")


# ---- CIFAR-10 -----------------------------------------------------------------------------------------------------
def test_cifar10_binary_reader_and_pipeline(tmp_path):
    synthetic.write_cifar10_binary(str(tmp_path), train_per_file=20, test=12)
    labels, images = cifar10.read_records(cifar10.data_files(str(tmp_path), eval_data=False))
    assert labels.shape == (100,) and images.shape == (100, 32, 32, 3) and images.dtype == np.uint8
    raw = np.fromfile(tmp_path / "cifar-10-batches-bin" / "data_batch_1.bin", dtype=np.uint8).reshape(-1, 3073)
    assert labels[0] == raw[0, 0]
    assert np.array_equal(images[0, :, :, 0], raw[0, 1:1025].reshape(32, 32))      # CHW on disk -> HWC
    it = cifar10.distorted_inputs(str(tmp_path), 8, image_size=24, seed=1)
    x, y = next(it)
    it.close()
    assert tuple(x.shape) == (8, 24, 24, 3) and x.dtype == torch.float32 and y.dtype == torch.int64
    assert torch.allclose(x.reshape(8, -1).mean(1), torch.zeros(8), atol=1e-4)       # per-image standardisation
    assert torch.allclose(x.reshape(8, -1).std(1, unbiased=False), torch.ones(8), atol=1e-3)
    ev = cifar10.inputs(True, str(tmp_path), 5, image_size=32)
    assert sum(b[0].shape[0] for b in ev) == 12
    with pytest.raises(ValueError):
        cifar10.data_files(str(tmp_path / "nope"), False)
    assert not cifar10.maybe_download_and_extract(str(tmp_path / "nope"))
    # disjoint per-rank shards
    a = cifar10.Cifar10Batches(str(tmp_path), 4, 32, rank=0, world=2, shuffle=False)
    b = cifar10.Cifar10Batches(str(tmp_path), 4, 32, rank=1, world=2, shuffle=False)
    assert a.labels.numel() + b.labels.numel() == 100
    a.close(), b.close()


# ---- TFRecords / ImageNet ----------------------------------------------------------------------------------------
def test_tfrecord_and_example_roundtrip(tmp_path):
    assert tfrecord.crc32c(b"123456789") == 0xE3069283
    feats = {"image/encoded": [b"\xff\xd8abc"], "image/class/label": [7], "image/class/text": [b"tench"],
             "image/object/bbox/xmin": [0.1, 0.25], "neg": [-3, 1 << 40]}
    rec = tfrecord.encode_example(feats)
    out = tfrecord.decode_example(rec)
    assert out["image/encoded"] == [b"\xff\xd8abc"] and out["image/class/label"] == [7]
    assert out["image/object/bbox/xmin"] == pytest.approx([0.1, 0.25]) and out["neg"] == [-3, 1 << 40]
    p = str(tmp_path / "x-00000-of-00001")
    with tfrecord.TFRecordWriter(p) as w:
        w.write(rec)
        w.write(b"second")
    assert list(tfrecord.read_records(p, check_crc=True)) == [rec, b"second"]
    try:
        from google.protobuf import descriptor  # noqa: F401
    except Exception:
        pass


def test_imagenet_pipeline_and_builders(tmp_path):
    synthetic.write_imagenet_tfrecords(str(tmp_path), "train", 10, 2, size=40)
    ds = imagenet.ImagenetData("train", str(tmp_path))
    assert ds.available() and ds.num_classes() == 1000 and ds.num_examples_per_epoch() == 1281167
    it = imagenet.distorted_inputs(ds, 4, image_size=32, num_preprocess_threads=2, shuffle_buffer=2)
    x, y = next(it)
    it.close()
    assert tuple(x.shape) == (4, 32, 32, 3) and float(x.min()) >= -1.0 and float(x.max()) <= 1.0
    assert all(1 <= int(v) <= 1000 for v in y)
    import random
    assert imagenet.sample_distorted_bounding_box(100, 80, np.zeros((0, 4), np.float32), random.Random(0))[2] <= 100
    img = torch.rand(50, 60, 3)
    assert tuple(imagenet.eval_image(img, 32, 32).shape) == (32, 32, 3)
    hsv = imagenet.rgb_to_hsv(img)
    assert torch.allclose(imagenet.hsv_to_rgb(hsv), img, atol=1e-5)
    # folder tree -> TFRecords -> reader
    from PIL import Image
    from distributed_tensorflow_models_b200.data import build_image_data as bid
    for lab in ("daisy", "rose"):
        os.makedirs(tmp_path / "raw" / lab)
        for i in range(3):
            Image.fromarray(np.random.randint(0, 255, (20, 24, 3), np.uint8)).save(tmp_path / "raw" / lab / f"{i}.png")
    f, t, l = bid.find_image_files(str(tmp_path / "raw"), ["daisy", "rose"])
    assert sorted(set(l)) == [1, 2] and len(f) == 6
    outs = bid.process_image_files("train", f, t, l, 2, str(tmp_path / "out"), 2)
    recs = [r for p in outs for r in tfrecord.read_records(p)]
    assert len(recs) == 6
    image, label, bbox, text = imagenet.parse_example_proto(recs[0])
    assert image[:3] == b"\xff\xd8\xff" and text in (b"daisy", b"rose")              # PNG re-encoded as JPEG


def test_bounding_box_xml_to_csv(tmp_path):
    from distributed_tensorflow_models_b200.data import process_bounding_boxes as pbb
    d = tmp_path / "n01440764"
    d.mkdir()
    (d / "n01440764_1.xml").write_text(
        "<annotation><size><width>200</width><height>100</height></size>"
        "<object><name>n01440764</name><bndbox><xmin>20</xmin><ymin>10</ymin><xmax>220</xmax><ymax>50</ymax></bndbox></object>"
        "<object><name>n99999999</name><bndbox><xmin>1</xmin><ymin>1</ymin><xmax>2</xmax><ymax>2</ymax></bndbox></object>"
        "</annotation>")
    assert list(pbb.process(str(tmp_path))) == ["n01440764_1.JPEG,0.1000,0.1000,1.0000,0.5000"]


# ---- flags / presets / launcher ------------------------------------------------------------------------------------
def test_flags_and_presets_match_reference_defaults():
    a = cfg.parse_args(["--job_name=worker", "--task_id=1", "--ps_hosts=a:22222", "--worker_hosts=b:22223,c:22223"],
                       model="inception", mode="ssp")
    assert (a.batch_size, a.max_steps, a.initial_learning_rate, a.staleness) == (32, 1000000, 0.045, 5)
    assert a.learning_rate_decay_factor == 0.94 and a.num_epochs_per_decay == 2.0 and a.save_interval_secs == 600.0
    assert a.worker_host_list == ["b:22223", "c:22223"] and a.task_id == 1
    assert a.preset.optimizer == "rmsprop" and a.preset.ema_decay is None and not a.preset.lr_times_workers_in_bsp
    p = cfg.PRESETS
    assert (p["cnn"].initial_lr, p["cnn"].lr_times_workers_in_bsp) == (0.32, False)
    assert (p["resnet"].weight_decay, p["resnet"].epochs_per_decay, p["resnet"].max_to_keep) == (2e-3, 32.0, 1)
    assert p["vgg"].save_secs is None and p["vgg"].ema_decay is None
    assert cfg.preset_for("vgg", "asp").ema_decay == 0.9999
    assert cfg.parse_args(["--run_once=True", "--synthetic"], model="cnn").run_once is True
    assert (cfg.PS_PORT, cfg.WORKER_PORT, cfg.SSP_PORT) == (22222, 22223, 8000)


def test_log_line_formats():
    s = logs.format_step("cifar", 3, 7, 2.345, 1234.56, 0.1037, now=0.0)
    assert s.startswith("time: 0.0; ") and s.endswith(": step 3 (global_step 7), loss = 2.35 (1234.6 examples/sec; 0.104 sec/batch)")
    assert "step 3 (gs 7), loss= 2.35 (1234.6 samples/s; 0.104 s/batch)" in logs.format_step("resnet", 3, 7, 2.345, 1234.56, 0.1037)
    assert not logs.format_step("cnn", 1, 1, 1.0, 1.0, 1.0).startswith("time:")
    assert logs.format_precision(0.8312, 99).endswith("precision @ 1 = 0.831, global_step: 99")


def test_launcher_host_lists_usage_and_all_pairs():
    from distributed_tensorflow_models_b200 import launcher
    ps, wk = launcher.host_lists(["h1", "h2"], ["h3", "h3", "h4"])
    assert ps == ["h1:22222", "h2:22222"] and wk == ["h3:22223", "h3:22233", "h4:22223"]
    r = subprocess.run(["bash", os.path.join(ROOT, "train.sh")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "please specify" in r.stdout
    import importlib
    for m, base in launcher.TRAINER.items():
        for mode in cfg.MODES:
            assert hasattr(importlib.import_module(f"distributed_tensorflow_models_b200.trainers.{base}_{mode}"), "main")
    for mod in launcher.EVALUATOR.values():
        importlib.import_module(f"distributed_tensorflow_models_b200.eval.{mod}")


def test_launcher_divides_the_cores_of_a_host_among_its_local_processes(monkeypatch):
    from distributed_tensorflow_models_b200 import launcher
    monkeypatch.delenv("OMP_NUM_THREADS", raising=False)
    ps_env, w_env = launcher.local_thread_shares(["localhost"], ["localhost", "localhost"], cores=8)
    assert ps_env["OMP_NUM_THREADS"] == "1" and w_env == {"OMP_NUM_THREADS": "3", "MKL_NUM_THREADS": "3"}
    assert launcher.local_thread_shares([], ["localhost"] * 8, cores=192)[1]["OMP_NUM_THREADS"] == "24"
    assert launcher.local_thread_shares(["localhost"], ["localhost"] * 16, cores=4)[1]["OMP_NUM_THREADS"] == "1"
    # remote hosts do not count, a single local process keeps the machine, and the user's own setting wins
    assert launcher.local_thread_shares(["far-away"], ["localhost"], cores=8) == ({}, {})
    monkeypatch.setenv("OMP_NUM_THREADS", "5")
    assert launcher.local_thread_shares(["localhost"], ["localhost", "localhost"], cores=8) == ({}, {})


def test_launcher_accepts_zoo_networks(tmp_path, monkeypatch):
    """`train.sh <registry name> <mode>` runs the generic trainer for networks outside the six trained families."""
    from distributed_tensorflow_models_b200 import launcher
    spawned = []
    monkeypatch.setattr(launcher, "_spawn", lambda host, cmd, env, log, delay=0.0: spawned.append(cmd) or
                        type("P", (), {"pid": 0})())
    monkeypatch.setenv("workers", "localhost localhost")
    monkeypatch.setenv("ps", "")
    assert launcher.main(["mobilenet_v2", "asp", "--log_dir", str(tmp_path), "--no_eval", "--synthetic"]) == 0
    assert len(spawned) == 2
    assert spawned[0][2].endswith(".trainers.train") and "--model=mobilenet_v2" in spawned[0] and "--mode=asp" in spawned[0]
    assert launcher.main(["no_such_net", "bsp", "--log_dir", str(tmp_path)]) == 1


# ---- TensorFlow V2 checkpoint bundles (ckpt/tf_bundle.py) -------------------------------------------------------
def test_tf_bundle_round_trip_multi_block_and_listing(tmp_path):
    import numpy as np
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    rng = np.random.default_rng(0)
    tensors = {f"scope_{i // 7}/layer_{i}/weights": rng.standard_normal((3, 3, 4, 5)).astype(np.float32)
               for i in range(60)}
    tensors.update({"global_step": np.asarray(1234, dtype=np.int64), "empty": np.zeros((0, 3), np.float32),
                    "conv1/biases": rng.standard_normal(7).astype(np.float32), "flags": np.array([True, False]),
                    "scope_0/layer_0/weights/RMSProp": np.ones((2, 2), np.float64),
                    "ids": np.arange(5, dtype=np.int32)})
    prefix = str(tmp_path / "model.ckpt-1234")
    tb.write_bundle(prefix, tensors, block_size=512)          # tiny blocks: many data blocks + a real index block
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-1234.data-00000-of-00001", "model.ckpt-1234.index"]
    back = tb.read_bundle(prefix)
    assert set(back) == set(tensors)
    for k, v in tensors.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape and np.array_equal(back[k], v), k
    listing = tb.list_bundle(prefix)
    assert listing["global_step"] == (np.dtype("int64"), ()) and listing["empty"] == (np.dtype("float32"), (0, 3))
    assert tb.read_bundle(prefix, names=["ids"]).keys() == {"ids"}
    with pytest.raises(KeyError):
        tb.read_bundle(prefix, names=["nope"])
    # the table really is sorted, prefix-compressed and split into several blocks
    raw = open(prefix + ".index", "rb").read()
    keys = [k for k, _ in tb.read_table(raw)]
    assert keys[0] == b"" and keys == sorted(keys) and len(keys) == len(tensors) + 1
    assert raw.count(b"scope_0/layer_0/weights") <= 2        # later keys store only their unshared suffix


def test_tf_bundle_byte_layout_matches_the_published_format(tmp_path):
    """A one-tensor bundle, checked byte by byte against the table / bundle format written out by hand."""
    import struct
    import numpy as np
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    from distributed_tensorflow_models_b200.data.tfrecord import crc32c, mask_crc
    prefix = str(tmp_path / "m")
    value = np.array([1.0, 2.0], dtype=np.float32)
    tb.write_bundle(prefix, {"w": value})
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    assert data == struct.pack("<2f", 1.0, 2.0)
    header = bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])                  # num_shards=1, version{producer=1}
    entry = (bytes([0x08, 0x01,                                           # dtype = DT_FLOAT
                    0x12, 0x04, 0x12, 0x02, 0x08, 0x02,                   # shape { dim { size: 2 } }
                    0x28, 0x08,                                           # size = 8   (offset 0, shard 0 omitted)
                    0x35]) + struct.pack("<I", mask_crc(crc32c(data))))   # crc32c (fixed32, masked)
    block = (bytes([0, 0, len(header)]) + header +                        # shared, unshared, value_len, key "", value
             bytes([0, 1, len(entry)]) + b"w" + entry +
             struct.pack("<II", 0, 1))                                    # restart[0] = 0, num_restarts = 1
    trailer = lambda b: b"\x00" + struct.pack("<I", mask_crc(crc32c(b + b"\x00")))
    meta = struct.pack("<II", 0, 1)
    meta_off = len(block) + 5
    index_block = (bytes([0, 1, 2]) + b"w" + bytes([0, len(block)]) + struct.pack("<II", 0, 1))
    index_off = meta_off + len(meta) + 5
    footer = bytes([meta_off, len(meta), index_off, len(index_block)])
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", 0xDB4775248B80FB57)
    expected = block + trailer(block) + meta + trailer(meta) + index_block + trailer(index_block) + footer
    assert open(prefix + ".index", "rb").read() == expected
    assert crc32c(b"123456789") == 0xE3069283                              # the CRC-32C check value


def test_tf_bundle_detects_corruption_and_unsupported_input(tmp_path):
    import numpy as np
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    prefix = str(tmp_path / "c")
    tb.write_bundle(prefix, {"a/b": np.arange(100, dtype=np.float32), "z": np.ones(3, np.int64)})
    raw = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    raw[17] ^= 0x40
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        tb.read_bundle(prefix)
    assert tb.read_bundle(prefix, verify=False)["z"].tolist() == [1, 1, 1]
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[3] ^= 0x01
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError, match="checksum"):
        tb.read_bundle(prefix)
    open(prefix + ".index", "wb").write(b"not a table at all, but long enough to have a footer ..........")
    with pytest.raises(ValueError, match="magic"):
        tb.read_bundle(prefix)
    with pytest.raises(TypeError):
        tb.write_bundle(str(tmp_path / "d"), {"c": np.zeros(2, np.complex64)})
    with pytest.raises(ValueError):
        w = tb.TableWriter(open(tmp_path / "t", "wb"))
        w.add(b"b", b"1")
        w.add(b"a", b"2")


def test_trainer_saves_resumes_and_evaluates_tf_bundles(tmp_path):
    """--ckpt_format tf: the chief writes model.ckpt-N.{index,data-*} + the `checkpoint` state file; eval, resume and
    the keep-N policy work on them exactly as on the .npz container."""
    from distributed_tensorflow_models_b200.ckpt import store as ckpt
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    from distributed_tensorflow_models_b200.eval import cifar10_eval
    from distributed_tensorflow_models_b200.trainers import cifar10_cnn_bsp
    d = str(tmp_path / "train")
    base = ["--synthetic", "--batch_size=4", "--train_dir=" + d, "--ckpt_format=tf", "--backend=local"]
    assert cifar10_cnn_bsp.main(base + ["--max_steps=2"]) == 0
    latest = ckpt.latest_checkpoint(d)
    assert latest.endswith("model.ckpt-2") and tb.is_bundle(latest) and not os.path.exists(latest + ".npz")
    listing = tb.list_bundle(latest)
    assert listing["conv1/weights"][1] == (5, 5, 3, 64) and listing["local3/weights"][1] == (2304, 384)   # HWIO / [in,out]
    assert "conv1/weights/ExponentialMovingAverage" in listing and listing["Variable"][1] == ()
    assert cifar10_eval.main(["--checkpoint_dir=" + d, "--run_once", "--synthetic", "--num_examples=8",
                              "--batch_size=4", "--eval_dir=" + str(tmp_path / "eval")]) == 0
    assert cifar10_cnn_bsp.main(base + ["--max_steps=1", "--resume"]) == 0
    assert ckpt.latest_checkpoint(d).endswith("model.ckpt-3")
    tensors = ckpt.load(ckpt.latest_checkpoint(d))
    assert int(tensors["Variable"]) == 3
    for step in range(4, 11):                                  # keep-N: old bundles (both files) are deleted
        ckpt.save(d, {"x": tensors["Variable"]}, step, max_to_keep=5, fmt="tf")
    kept = sorted(f for f in os.listdir(d) if f.endswith(".index"))
    assert len(kept) == 5 and "model.ckpt-2.index" not in kept
    assert len([f for f in os.listdir(d) if ".data-" in f]) == 5


def test_checkpoint_inspect_and_convert_cli(tmp_path, capsys):
    import numpy as np
    from distributed_tensorflow_models_b200.ckpt import inspect as ck_inspect
    from distributed_tensorflow_models_b200.ckpt import store as ckpt
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    d = str(tmp_path)
    ckpt.save(d, {"a/weights": np.arange(6, dtype=np.float32).reshape(2, 3), "global_step": np.asarray(7)}, 7)
    assert ck_inspect.main([d]) == 0
    out = capsys.readouterr().out
    assert "model.ckpt-7" in out and "a/weights" in out and "(2, 3)" in out and "2 tensors, 7 values" in out
    assert ck_inspect.main([os.path.join(d, "model.ckpt-7.npz"), "--tensor", "a/weights"]) == 0
    assert "float32 (2, 3)" in capsys.readouterr().out
    out_prefix = str(tmp_path / "conv" / "model.ckpt-7")
    assert ck_inspect.main([d, "--convert", "tf", "--out", out_prefix]) == 0
    assert np.array_equal(tb.read_bundle(out_prefix)["a/weights"], np.arange(6, dtype=np.float32).reshape(2, 3))
    assert ck_inspect.main([out_prefix + ".index"]) == 0 and "tf bundle" in capsys.readouterr().out
    with pytest.raises(SystemExit):
        ck_inspect.main([str(tmp_path / "missing")])


def test_tf_bundle_partitioned_variables_are_reassembled(tmp_path):
    """The reference's fixed_size_partitioner trainers save variables as axis-0 slices (one slice-listing entry + one
    data entry per slice under an ordered-code key); reading gives back the whole tensor under its logical name."""
    import numpy as np
    from distributed_tensorflow_models_b200.ckpt import tf_bundle as tb
    rng = np.random.default_rng(1)
    tensors = {"partitioned_space/CifarNet/conv1/weights": rng.standard_normal((5, 5, 3, 64)).astype(np.float32),
               "partitioned_space/CifarNet/fc3/weights": rng.standard_normal((4096, 384)).astype(np.float32),
               "partitioned_space/CifarNet/fc3/biases": rng.standard_normal(384).astype(np.float32),
               "partitioned_space/Variable": np.asarray(9, np.int64)}
    parts = {"partitioned_space/CifarNet/conv1/weights": [3, 2],            # fixed_size_partitioner(2, axis=0)
             "partitioned_space/CifarNet/fc3/weights": [1366, 1365, 1365],   # three parameter servers
             "partitioned_space/CifarNet/fc3/biases": [192, 192]}
    prefix = str(tmp_path / "model.ckpt-9")
    tb.write_bundle(prefix, tensors, partitions=parts)
    keys = [k for k, _ in tb.read_table(open(prefix + ".index", "rb").read())]
    assert keys == sorted(keys) and sum(k.startswith(b"\x00") for k in keys) == 7       # the seven slice data entries
    k0 = tb.encode_slice_key("partitioned_space/CifarNet/fc3/weights", [(1366, 1365), (0, None)])
    assert k0 in keys and tb.decode_slice_key_name(k0) == "partitioned_space/CifarNet/fc3/weights"
    back = tb.read_bundle(prefix)
    assert set(back) == set(tensors)
    for k, v in tensors.items():
        assert back[k].shape == v.shape and np.array_equal(back[k], v), k
    assert tb.list_bundle(prefix)["partitioned_space/CifarNet/fc3/weights"] == (np.dtype("float32"), (4096, 384))
    # ordered code: order-preserving signed integers, escaped strings
    vals = [-70000, -8193, -8192, -65, -64, -1, 0, 1, 63, 64, 8191, 8192, 1 << 20, 1 << 40]
    enc = [tb._oc_signed_increasing(v) for v in vals]
    assert enc == sorted(enc) and enc[5] == b"\x7f" and enc[6] == b"\x80" and enc[9] == b"\xc0\x40"
    assert tb.decode_slice_key_name(tb.encode_slice_key("a\x00b", [(0, None)])) == "a\x00b"
    with pytest.raises(ValueError):
        tb.write_bundle(str(tmp_path / "bad"), {"w": np.zeros((4, 2), np.float32)}, partitions={"w": [3, 2]})


def test_tensorboard_event_files(tmp_path):
    """Scalar summaries in TensorBoard's event-file format: evaluators write `Precision @ 1` into eval_dir, the chief
    writes global_step/sec (+ loss, learning rate) into train_dir every --save_summaries_secs."""
    import glob
    import struct
    from distributed_tensorflow_models_b200.eval import cifar10_eval
    from distributed_tensorflow_models_b200.trainers import cifar10_cnn_bsp
    from distributed_tensorflow_models_b200.utils import tfevents
    w = tfevents.EventFileWriter(str(tmp_path / "unit"))
    w.add_scalars(12, {"Precision @ 1": 0.75, "Recall @ 5": 1.0}, wall_time=1234.5)
    w.close()
    evs = tfevents.read_events(w.path)                        # CRC-checked TFRecord framing
    assert os.path.basename(w.path).startswith("events.out.tfevents.")
    assert evs[0]["file_version"] == "brain.Event:2" and evs[1]["step"] == 12 and evs[1]["wall_time"] == 1234.5
    assert evs[1]["scalars"] == {"Precision @ 1": 0.75, "Recall @ 5": 1.0}
    # Event{wall_time=1(fixed64), step=2, summary=5{value=1{tag=1, simple_value=2(fixed32)}}}
    raw = tfevents.encode_event(2.0, 3, scalars={"a": 0.5})
    assert raw == (b"\x09" + struct.pack("<d", 2.0) + b"\x10\x03" + b"\x2a\x0a" + b"\x0a\x08" + b"\x0a\x01a" +
                   b"\x15" + struct.pack("<f", 0.5))
    d = str(tmp_path / "train")
    assert cifar10_cnn_bsp.main(["--synthetic", "--batch_size=4", "--train_dir=" + d, "--max_steps=3",
                                 "--backend=local", "--save_summaries_secs=0.000001"]) == 0
    files = glob.glob(os.path.join(d, "events.out.tfevents.*"))
    assert len(files) == 1
    evs = tfevents.read_events(files[0])
    assert len(evs) >= 2 and {"global_step/sec", "total_loss", "learning_rate"} <= set(evs[-1]["scalars"])
    e = str(tmp_path / "eval")
    assert cifar10_eval.main(["--checkpoint_dir=" + d, "--run_once", "--synthetic", "--num_examples=8",
                              "--batch_size=4", "--eval_dir=" + e]) == 0
    evs = tfevents.read_events(glob.glob(os.path.join(e, "events.out.tfevents.*"))[0])
    assert evs[-1]["step"] == 3 and "Precision @ 1" in evs[-1]["scalars"]


def test_cifar_fused_distortion_matches_stepwise_and_stream_is_thread_count_independent(tmp_path):
    """distort_batch (crop + flip as one uint8 gather, brightness / contrast folded into the standardisation) equals
    the step-by-step reference distortion; the batch stream is a pure function of (seed, rank, batch number)."""
    g = torch.Generator().manual_seed(0)
    raw = torch.randint(0, 256, (16, 32, 32, 3), dtype=torch.uint8, generator=g)
    for size in (24, 32):
        a = cifar10.distort_batch_reference(raw, size, torch.Generator().manual_seed(5))
        b = cifar10.distort_batch(raw, size, torch.Generator().manual_seed(5))
        assert a.shape == b.shape and float((a - b).abs().max()) < 1e-5
    flat = torch.full((2, 32, 32, 3), 7, dtype=torch.uint8)              # constant image: the 1/sqrt(n) floor applies
    assert torch.isfinite(cifar10.distort_batch(flat, 32, torch.Generator().manual_seed(1))).all()
    synthetic.write_cifar10_binary(str(tmp_path), train_per_file=40, test=8)
    one = cifar10.distorted_inputs(str(tmp_path), 16, 32, seed=3, num_threads=1)
    three = cifar10.distorted_inputs(str(tmp_path), 16, 32, seed=3, num_threads=3)
    try:
        for _ in range(30):                                               # crosses an epoch boundary (200 / 16)
            (xa, ya), (xb, yb) = next(one), next(three)
            assert torch.equal(xa, xb) and torch.equal(ya, yb)
    finally:
        one.close()
        three.close()
    other = cifar10.distorted_inputs(str(tmp_path), 16, 32, seed=4, num_threads=1)
    assert not torch.equal(next(other)[1], ya)
    other.close()


def test_imagenet_pipeline_process_workers_and_fast_colour_math(tmp_path):
    """Process workers hand images over through the shared-memory ring: a finite pass delivers every image exactly
    once, shutting down mid-stream leaves no worker behind; the planar NumPy colour distortion equals the tensor
    formulas it replaced."""
    import random
    from distributed_tensorflow_models_b200.data import imagenet
    synthetic.write_imagenet_tfrecords(str(tmp_path), "train", num_images=40, num_shards=4, size=48)
    ds = imagenet.ImagenetData("train", str(tmp_path))
    it = imagenet.ImagenetBatches(ds, 16, True, 32, 4, repeat=False, workers="process", seed=1)
    assert it.mode == "process"
    batches = list(it)
    it.close()
    assert sum(b[0].shape[0] for b in batches) == 40 and batches[0][0].shape == (16, 32, 32, 3)
    assert all(float(b[0].abs().max()) <= 1.0 for b in batches)
    it = imagenet.distorted_inputs(ds, 8, image_size=32, num_preprocess_threads=4, workers="process")
    next(it)
    it.close()
    assert not any(w.is_alive() for w in it.workers)
    with pytest.raises(ValueError):
        imagenet.ImagenetBatches(ds, 8, True, 32, 2, workers="fibers")
    # colour math: NumPy planes vs the torch HSV formulas, both orderings
    def stepwise(img, tid, rng):
        def brightness(x): return x + rng.uniform(-32.0 / 255.0, 32.0 / 255.0)
        def saturation(x):
            hsv = imagenet.rgb_to_hsv(x.clamp(0, 1)); hsv[..., 1] = (hsv[..., 1] * rng.uniform(0.5, 1.5)).clamp(0, 1)
            return imagenet.hsv_to_rgb(hsv)
        def hue(x):
            hsv = imagenet.rgb_to_hsv(x.clamp(0, 1)); hsv[..., 0] = (hsv[..., 0] + rng.uniform(-0.2, 0.2)) % 1.0
            return imagenet.hsv_to_rgb(hsv)
        def contrast(x):
            m = x.mean(dim=(0, 1), keepdim=True); return (x - m) * rng.uniform(0.5, 1.5) + m
        for fn in ((brightness, saturation, hue, contrast) if tid % 2 == 0 else (brightness, contrast, saturation, hue)):
            img = fn(img)
        return img.clamp(0.0, 1.0)
    img = torch.rand(37, 41, 3, generator=torch.Generator().manual_seed(0))
    for tid in (0, 1):
        a = stepwise(img.clone(), tid, random.Random(7))
        b = imagenet.distort_color(img.clone(), tid, random.Random(7))
        assert float((a - b).abs().max()) < 1e-5


def test_build_imagenet_data_without_a_synsets_file(tmp_path, capsys):
    """The ILSVRC synsets file is the sorted list of the class directories: the builder reads the classes off the
    training tree when the file is not there, and labels come out 1..N in that order (0 = background)."""
    from PIL import Image
    from distributed_tensorflow_models_b200.data import build_imagenet_data as bimd
    rs = np.random.RandomState(0)
    for split, per in (("train", 3), ("val", 1)):
        for syn in ("n02000000", "n01000000"):
            os.makedirs(tmp_path / split / syn)
            for i in range(per):
                Image.fromarray(rs.randint(0, 255, (20, 24, 3), dtype=np.uint8)).save(
                    tmp_path / split / syn / f"{syn}_{i}.JPEG")
    out = tmp_path / "records"
    os.makedirs(out)
    assert bimd.main(["--train_directory", str(tmp_path / "train"), "--validation_directory", str(tmp_path / "val"),
                      "--output_directory", str(out), "--train_shards", "2", "--validation_shards", "1",
                      "--num_threads", "1", "--labels_file", str(tmp_path / "missing.txt"),
                      "--imagenet_metadata_file", str(tmp_path / "missing_too.txt"),
                      "--bounding_box_file", str(tmp_path / "none.csv")]) == 0
    assert "2 classes taken from the directories" in capsys.readouterr().out
    got = {}
    for p in sorted(os.listdir(out)):
        for rec in tfrecord.read_records(str(out / p)):
            image, label, bbox, text = imagenet.parse_example_proto(rec)
            got.setdefault(p.split("-")[0], []).append((int(label), text))
    assert sorted(got["train"]) == [(1, b"n01000000")] * 3 + [(2, b"n02000000")] * 3
    assert sorted(got["validation"]) == [(1, b"n01000000"), (2, b"n02000000")]
    with pytest.raises(SystemExit):
        bimd.main(["--train_directory", str(out), "--labels_file", str(tmp_path / "missing.txt")])


def test_resume_restores_optimizer_slots_and_ema_shadows(tmp_path):
    """Save -> new process -> resume: RMSProp's ms / mom and the EMA shadows come back with the variables (the
    reference's Supervisor / Saver restore all of them, inception/imagenet_inception_bsp.py:156-167); a resumed run
    continues exactly like the uninterrupted one."""
    import dataclasses

    from distributed_tensorflow_models_b200 import config as cfg
    from distributed_tensorflow_models_b200.ckpt import store as ckpt
    from distributed_tensorflow_models_b200.parallel.comm import ThreadCluster
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    preset = dataclasses.replace(cfg.PRESETS["cnn"], optimizer="rmsprop", initial_lr=0.01, ema_decay=0.9999)
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(4, 24, 24, 3, generator=g), torch.randint(0, 10, (4,), generator=g)

    def mk(seed):
        return TrainSetup(model="cnn", batch_size=4, num_workers=2, preset=preset, seed=seed)

    saved = {}

    def first(comm):
        t = Trainer(mk(0), comm)
        for _ in range(3):
            t.train_step(x, y)
        slots = comm.gather_slot_flats()
        if comm.rank == 0:
            tensors = ckpt.collect_tensors(t.store, slots, t.layout, comm.global_step(), "Variable")
            saved["path"] = ckpt.save(str(tmp_path), tensors, comm.global_step())
        comm.barrier()
        out = t.train_step(x, y)                       # the uninterrupted run's 4th step
        return float(out["loss"]), t.layout.params.clone(), {k: v.clone() for k, v in comm.gather_slot_flats().items()}

    ref = ThreadCluster(2).run(first)
    names = set(ckpt.load(saved["path"]))
    assert {"conv1/weights/RMSProp", "conv1/weights/RMSProp_1", "conv1/weights/ExponentialMovingAverage"} <= names

    def resumed(comm):
        t = Trainer(mk(5), comm)                       # different initial values: everything must come from the file
        done = ckpt.restore_training_state(t, ckpt.load(saved["path"]))
        comm.set_global_step(ckpt.step_of(saved["path"]))
        assert {"RMSProp", "RMSProp_1", "ExponentialMovingAverage"} <= set(done)
        out = t.train_step(x, y)
        return float(out["loss"]), t.layout.params.clone(), {k: v.clone() for k, v in comm.gather_slot_flats().items()}

    got = ThreadCluster(2).run(resumed)
    probe = Trainer(dataclasses.replace(mk(0), num_workers=1))
    live = torch.zeros(ref[0][1].numel(), dtype=torch.bool)
    for e in probe.layout.entries:
        live[e.offset:e.offset + e.numel] = True
    for (l0, p0, s0), (l1, p1, s1) in zip(ref, got):
        assert l0 == pytest.approx(l1, rel=1e-6)
        assert torch.allclose(p0, p1, rtol=1e-6, atol=1e-8)
        for k in s0:                                   # (the alignment padding between tensors is not state)
            assert torch.allclose(s0[k][live], s1[k][live], rtol=1e-6, atol=1e-8), k

    # a checkpoint without shadows (e.g. written by a run without EMA): the EMA restarts from the restored weights
    tensors = {k: v for k, v in ckpt.load(saved["path"]).items() if not k.endswith("/ExponentialMovingAverage")}
    t = Trainer(dataclasses.replace(mk(7), num_workers=1))
    ckpt.restore_training_state(t, tensors)
    o = t.comm.owned_optimizers()[0]
    assert torch.equal(o.ema, o.w)


def test_cifar10_download_helper_fetches_and_unpacks(tmp_path, monkeypatch):
    """maybe_download_and_extract (cnn/cifar10.py:394-411): missing files -> fetch the tarball -> unpack it; an
    unreachable source is not an error (the trainers then use synthetic data)."""
    import tarfile

    from distributed_tensorflow_models_b200.data import cifar10, synthetic
    src = tmp_path / "src"
    synthetic.write_cifar10_binary(str(src), train_per_file=6, test=4)
    tarball = tmp_path / "cifar-10-binary.tar.gz"
    with tarfile.open(tarball, "w:gz") as tf:
        tf.add(str(src / "cifar-10-batches-bin"), arcname="cifar-10-batches-bin")
    monkeypatch.delenv("DTF_NO_DOWNLOAD", raising=False)
    dst = tmp_path / "data"
    assert cifar10.maybe_download_and_extract(str(dst), url="file://" + str(tarball))
    assert len(cifar10.data_files(str(dst), False)) == 5 and os.path.exists(dst / "cifar-10-binary.tar.gz")
    assert cifar10.maybe_download_and_extract(str(dst), url="file:///nonexistent")        # present: nothing to do
    assert not cifar10.maybe_download_and_extract(str(tmp_path / "other"), url="file:///nonexistent/x.tar.gz")
    assert not os.path.exists(tmp_path / "other" / "x.tar.gz.part")


def test_download_and_preprocess_shell_drivers(tmp_path):
    """inception/data/download_and_preprocess_flowers.sh and download_imagenet.sh: fetch (here: file:// URLs and
    archives already in place), unpack into the tree the builders expect, shard to TFRecords."""
    import glob
    import tarfile

    from PIL import Image
    data = os.path.join(ROOT, "distributed_tensorflow_models_b200", "data")
    # ---- flowers: tarball behind a URL -> train / validation shards ----
    src = tmp_path / "src" / "flower_photos"
    for lab in ("daisy", "roses"):
        (src / lab).mkdir(parents=True)
        for i in range(4):
            Image.fromarray(np.random.randint(0, 255, (20, 24, 3), np.uint8)).save(src / lab / f"{i}.jpg")
    (src / "LICENSE.txt").write_text("not a label")
    tgz = tmp_path / "flower_photos.tgz"
    with tarfile.open(tgz, "w:gz") as tf:
        tf.add(str(src), arcname="flower_photos")
    out = tmp_path / "flowers"
    env = dict(os.environ, FLOWERS_URL="file://" + str(tgz), PYTHONPATH=ROOT)
    r = subprocess.run(["bash", os.path.join(data, "download_and_preprocess_flowers.sh"), str(out), "1"],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert (out / "flower_photos.tgz").exists() and open(out / "raw-data" / "labels.txt").read().split() == ["daisy", "roses"]
    assert glob.glob(str(out / "train-*")) and glob.glob(str(out / "validation-*"))
    assert len(os.listdir(out / "raw-data" / "validation" / "daisy")) == 1
    # ---- ImageNet: the three official archives already in <out_dir> (no credentials needed, none stored) ----
    raw = tmp_path / "inet"
    raw.mkdir()
    work = tmp_path / "work"
    for syn in ("n01440764", "n01443537"):
        (work / syn).mkdir(parents=True)
        for i in range(2):
            Image.fromarray(np.random.randint(0, 255, (16, 16, 3), np.uint8)).save(work / syn / f"{syn}_{i}.JPEG")
        with tarfile.open(work / f"{syn}.tar", "w") as tf:
            for f in sorted(os.listdir(work / syn)):
                tf.add(str(work / syn / f), arcname=f)
    with tarfile.open(raw / "ILSVRC2012_img_train.tar", "w") as tf:
        for syn in ("n01440764", "n01443537"):
            tf.add(str(work / f"{syn}.tar"), arcname=f"{syn}.tar")
    (work / "val").mkdir()
    Image.fromarray(np.zeros((8, 8, 3), np.uint8)).save(work / "val" / "ILSVRC2012_val_00000001.JPEG")
    with tarfile.open(raw / "ILSVRC2012_img_val.tar", "w") as tf:
        tf.add(str(work / "val" / "ILSVRC2012_val_00000001.JPEG"), arcname="ILSVRC2012_val_00000001.JPEG")
    (work / "bb" / "n01440764").mkdir(parents=True)
    (work / "bb" / "n01440764" / "n01440764_0.xml").write_text("<annotation/>")
    with tarfile.open(raw / "ILSVRC2012_bbox_train_v2.tar.gz", "w:gz") as tf:
        tf.add(str(work / "bb" / "n01440764"), arcname="n01440764")
    env2 = {k: v for k, v in os.environ.items() if not k.startswith("IMAGENET_")}
    r = subprocess.run(["bash", os.path.join(data, "download_imagenet.sh"), str(raw)], capture_output=True, text=True,
                       env=env2, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert sorted(os.listdir(raw / "train")) == ["n01440764", "n01443537"]
    assert len(os.listdir(raw / "train" / "n01443537")) == 2 and len(os.listdir(raw / "validation")) == 1
    assert (raw / "bounding_boxes" / "n01440764" / "n01440764_0.xml").exists()
    # a missing archive without credentials: a clear message, exit code 2, nothing fetched
    os.remove(raw / "ILSVRC2012_img_val.tar")
    r = subprocess.run(["bash", os.path.join(data, "download_imagenet.sh"), str(raw)], capture_output=True, text=True,
                       env=env2, timeout=60)
    assert r.returncode == 2 and "IMAGENET_USERNAME" in r.stdout
    # no credentials anywhere in the shipped drivers
    for f in glob.glob(os.path.join(data, "*.sh")):
        assert "ACCESS_KEY=" not in open(f).read().replace('"$IMAGENET_ACCESS_KEY"', "")


def test_activation_loss_and_variable_summaries(tmp_path):
    """The summaries the reference's model files define (cnn/cifar10.py:104-121,309-335,376-382): per end point a
    histogram `<name>/activations` and a scalar `<name>/sparsity`, `<loss> (raw)` next to its moving average,
    histograms of every trainable variable and of its gradient — written by the chief on request."""
    import glob
    import struct

    from distributed_tensorflow_models_b200.trainers import cifar10_cnn_bsp
    from distributed_tensorflow_models_b200.utils import tfevents
    from distributed_tensorflow_models_b200.utils.summaries import LossAverages, activation_summaries
    # histogram encoding: TensorFlow's buckets, counts add up, moments exact
    x = torch.cat([torch.zeros(6), torch.tensor([1.0, -2.0, 3.5])])
    h = tfevents._decode_histogram(tfevents.encode_histogram(x))
    assert (h["min"], h["max"], h["num"], h["sum"], h["sum_squares"]) == (-2.0, 3.5, 9.0, 2.5, 17.25)
    assert sum(h["bucket"]) == 9 and len(h["bucket"]) == len(h["bucket_limit"])
    lim = tfevents.default_bucket_limits()
    assert lim[len(lim) // 2 - 1] == 0.0 or 0.0 in lim
    assert all(b > a for a, b in zip(lim, lim[1:]))
    zero_bucket = [c for l, c in zip(h["bucket_limit"], h["bucket"]) if l == 0.0]
    assert zero_bucket == [6.0]
    sc, hs = activation_summaries({"tower_0/conv1": torch.tensor([0.0, 0.0, 1.0, 2.0])})
    assert sc == {"conv1/sparsity": 0.5} and list(hs) == ["conv1/activations"]
    la = LossAverages(0.9)
    assert la.update(total_loss=2.0) == {"total_loss (raw)": 2.0, "total_loss": 2.0}
    assert la.update(total_loss=1.0)["total_loss"] == pytest.approx(1.9)
    # through the trainer
    d = str(tmp_path / "train")
    assert cifar10_cnn_bsp.main(["--synthetic", "--batch_size=4", "--train_dir=" + d, "--max_steps=3",
                                 "--backend=local", "--save_summaries_secs=0.000001", "--summarize_activations"]) == 0
    evs = tfevents.read_events(glob.glob(os.path.join(d, "events.out.tfevents.*"))[0])
    last = evs[-1]
    assert {"total_loss (raw)", "total_loss", "learning_rate", "conv1/sparsity", "local4/sparsity",
            "softmax_linear/sparsity"} <= set(last["scalars"])
    assert {"conv1/activations", "conv2/activations", "local3/activations", "conv1/weights", "conv1/weights/gradients",
            "softmax_linear/biases"} <= set(last["histograms"])
    assert last["histograms"]["conv1/weights"]["num"] == 64 * 5 * 5 * 3
