"""The native CIFAR-10 input path (csrc/input_pipeline.cu): one pass from the uint8 records to the distorted float
batch, on host threads and as a kernel over records resident on the device — against the step-by-step statement of
the reference's distortion (cnn/cifar10_input.py:176-190, :242-247)."""
import pytest
import torch

from distributed_tensorflow_models_b200.data import cifar10, synthetic
from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher

needs_lib = pytest.mark.skipif(cifar10._native_lib() is None, reason="native library not built")


def _records(n=600, seed=0):
    g = torch.Generator().manual_seed(seed)
    imgs = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, generator=g)
    imgs[3] = 77                        # a constant image: the standard deviation floor 1/sqrt(n) applies
    return imgs


@needs_lib
@pytest.mark.parametrize("size", [24, 32, 17])
def test_native_host_pass_matches_the_step_by_step_distortion(size):
    imgs = _records()
    idx = torch.randperm(600, generator=torch.Generator().manual_seed(1))[:64]
    ref = cifar10.distort_batch_reference(imgs[idx], size, torch.Generator().manual_seed(9))
    draws = cifar10.draw_distortion(64, size, torch.Generator().manual_seed(9))
    for threads in (1, 3):
        out = cifar10.distort_records_native(imgs, idx, *draws, size, threads=threads)
        assert out.shape == ref.shape and torch.allclose(out, ref, atol=2e-5)
    # the constant image comes out as exact zeros (the float formulas leave rounding noise there)
    const = cifar10.distort_records_native(imgs, torch.tensor([3]), *cifar10.draw_distortion(1, size,
                                           torch.Generator().manual_seed(2)), size)
    assert torch.equal(const, torch.zeros_like(const))
    # evaluation preprocessing = the same pass with a central window, no flip, contrast 1
    o = torch.full((64,), (32 - size) // 2)
    ev = cifar10.distort_records_native(imgs, idx, o, o, torch.zeros(64, dtype=torch.bool), torch.ones(64), size)
    keep = idx != 3
    assert torch.allclose(ev[keep], cifar10.center_batch(imgs[idx], size)[keep], atol=2e-5)
    # bad arguments are refused, not read
    with pytest.raises(RuntimeError):
        cifar10.distort_records_native(imgs, torch.tensor([600]), o[:1], o[:1], torch.zeros(1), torch.ones(1), size)
    with pytest.raises(RuntimeError):
        cifar10.distort_records_native(imgs, torch.tensor([0]), torch.tensor([33 - size]), o[:1], torch.zeros(1),
                                       torch.ones(1), size)


@needs_lib
def test_host_and_device_resident_pipelines_deliver_the_same_batches(tmp_path):
    synthetic.write_cifar10_binary(str(tmp_path), train_per_file=20, test=12)
    kw = dict(seed=4, rank=1, world=2)
    plain = cifar10.Cifar10Batches(str(tmp_path), 8, 24, native=False, **kw)        # torch formulas
    fast = cifar10.Cifar10Batches(str(tmp_path), 8, 24, native=True, num_threads=3, **kw)
    resident = cifar10.DeviceCifar10Batches(str(tmp_path), 8, 24, device="cpu", **kw)
    assert fast.native and not plain.native
    fed = DevicePrefetcher(resident, torch.device("cpu"))
    try:
        for k in range(9):                       # crosses an epoch boundary (50 records per rank, 7 batches)
            (xa, ya), (xb, yb), (xc, yc) = next(plain), next(fast), next(fed)
            assert torch.equal(ya, yb) and torch.equal(ya, yc), k
            assert torch.allclose(xa, xb, atol=2e-5) and torch.equal(xb, xc), k
            assert isinstance(xc, torch.Tensor) and tuple(xc.shape) == (8, 24, 24, 3)
    finally:
        plain.close(), fast.close(), resident.close()
    # what crosses to the device per image: 21 bytes of record number and draws instead of 6912 bytes of pixels
    item = resident._make(0)[0]
    assert item.stage.numel() == 8 * cifar10.RecordBatch.BYTES_PER_IMAGE == 8 * 21
    idx, oy, ox, flip, factor = item.fields(item.stage)
    assert idx.dtype == torch.int64 and int(oy.max()) <= 8 and float(factor.min()) >= 0.2
    # evaluation: every record once, central crop
    ev_a = cifar10.inputs(True, str(tmp_path), 5, image_size=24, native=False)
    ev_b = cifar10.DeviceCifar10Batches(str(tmp_path), 5, 24, device="cpu", train=False, shuffle=False, repeat=False)
    got = [(x.to("cpu"), y) for x, y in ev_b]
    want = list(ev_a)
    assert len(got) == len(want) == 3 and sum(x.shape[0] for x, _ in got) == 12
    for (xa, ya), (xb, yb) in zip(want, got):
        assert torch.equal(ya, yb) and torch.allclose(xa, xb, atol=2e-5)
    with pytest.raises(ValueError):
        cifar10.DeviceCifar10Batches(str(tmp_path), 5, 40, device="cpu")


@needs_lib
def test_trainer_entry_point_with_the_device_resident_pipeline(tmp_path):
    from distributed_tensorflow_models_b200.trainers import cifar10_cnn_bsp
    data, train = str(tmp_path / "data"), str(tmp_path / "train")
    synthetic.write_cifar10_binary(data, train_per_file=16, test=8)
    assert cifar10_cnn_bsp.main(["--data_dir", data, "--train_dir", train, "--max_steps", "3", "--batch_size", "8",
                                 "--input_pipeline", "device", "--synthetic", "False"]) == 0


@needs_lib
@pytest.mark.parametrize("shape", [(299, 299, 3), (37, 41, 3), (1, 3, 3)])
def test_native_colour_distortion_matches_the_numpy_planes(shape):
    """dtf_distort_color_host (two sweeps, four pixels per SSE2 iteration + scalar tail) against the NumPy statement
    of inception/image_processing.py:164-195, for both orderings; greys, blacks, whites and primaries included."""
    import random

    from distributed_tensorflow_models_b200.data import imagenet
    img = torch.rand(*shape, generator=torch.Generator().manual_seed(0))
    if shape[0] > 80:
        img[:20, :20] = 0.5            # grey: hue undefined (d == 0)
        img[20:40, :20] = 0.0
        img[40:60, :20] = 1.0
        img[60:80, :20, 1:] = 0.0      # pure red: maximum shared by one channel only
    for thread_id in (0, 1, 2, 3):
        for seed in range(4):
            a = imagenet.distort_color(img, thread_id, random.Random(seed))
            b = imagenet.distort_color_numpy(img, thread_id, random.Random(seed))
            assert a.shape == b.shape and a.dtype == torch.float32
            assert float((a - b).abs().max()) < 1e-5, (thread_id, seed)
            assert float(a.min()) >= 0.0 and float(a.max()) <= 1.0
    # the input is not modified and both paths leave the generator in the same state
    before = img.clone()
    r1, r2 = random.Random(5), random.Random(5)
    imagenet.distort_color(img, 1, r1), imagenet.distort_color_numpy(img, 1, r2)
    assert torch.equal(img, before) and r1.random() == r2.random()


@needs_lib
@pytest.mark.parametrize("geom", [(20, 50, 300, 350, 299, 299), (0, 0, 375, 500, 299, 299), (100, 100, 60, 90, 299, 299),
                                  (3, 7, 301, 299, 224, 160), (0, 0, 10, 10, 8, 8), (5, 5, 1, 1, 4, 4)])
def test_native_crop_resize_matches_interpolate(geom):
    """dtf_crop_resize_host (separable taps straight from the uint8 window) against torch.nn.functional.interpolate
    on the float crop, for the four resize methods the ImageNet distortion cycles through, plain and mirrored."""
    import numpy as np

    from distributed_tensorflow_models_b200.data import imagenet
    y0, x0, ch, cw, oh, ow = geom
    img = np.random.default_rng(1).integers(0, 256, (375, 500, 3), dtype=np.uint8)
    for method in range(4):
        for flip in (False, True):
            ref = imagenet.resize(imagenet._to_float(img[y0:y0 + ch, x0:x0 + cw]), oh, ow, method)
            ref = ref.flip(1) if flip else ref
            out = imagenet.crop_resize_native(img, y0, x0, ch, cw, oh, ow, method, flip)
            assert out.shape == ref.shape and out.dtype == torch.float32
            assert float((out - ref).abs().max()) < 3e-5, (method, flip)        # a hundredth of one grey level
    with pytest.raises(RuntimeError):
        imagenet.crop_resize_native(img, 300, 0, 100, 100, 8, 8, 0, False)      # window leaves the image
    assert imagenet.crop_resize_native(img.astype(np.float32), 0, 0, 8, 8, 4, 4, 0, False) is None


@needs_lib
def test_imagenet_preprocessing_native_and_fallback_agree(monkeypatch):
    """Whole per-image path (decode, distorted crop, resize, flip, colour, scale to [-1, 1]) with and without the
    native passes: same draws, same image."""
    import io
    import random

    import numpy as np
    from PIL import Image

    from distributed_tensorflow_models_b200.data import cifar10 as c10
    from distributed_tensorflow_models_b200.data import imagenet
    yy, xx = np.mgrid[0:120, 0:160]
    arr = np.stack([(np.sin(xx / 20) + 1) * 100 + 30, (np.cos(yy / 15) + 1) * 90 + 40, (xx + yy) % 255], -1)
    buf = io.BytesIO()
    Image.fromarray(arr.astype(np.uint8)).save(buf, format="JPEG", quality=90)
    bbox = np.asarray([[0.1, 0.1, 0.9, 0.9]], dtype=np.float32)
    for train in (True, False):
        for tid in range(4):
            a = imagenet.image_preprocessing(buf.getvalue(), bbox, train, tid, random.Random(tid), image_size=64)
            monkeypatch.setattr(c10, "_native_lib", lambda: None)
            b = imagenet.image_preprocessing(buf.getvalue(), bbox, train, tid, random.Random(tid), image_size=64)
            monkeypatch.undo()
            assert a.shape == (64, 64, 3) and float(a.min()) >= -1.0 and float(a.max()) <= 1.0
            assert float((a - b).abs().max()) < 1e-4, (train, tid)
