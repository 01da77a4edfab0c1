"""Driver for compute-sanitizer: every kernel of csrc/zoo_ops.cu (and the transposed convolution on the tcgen05 dgrad
kernel) once, forward and backward, on small tensors with odd extents so that partial tiles, halos and channel tails
are all exercised.  Run through scripts/sanitize.sh."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from distributed_tensorflow_models_b200.ops import functional as Fn  # noqa: E402

g = torch.Generator().manual_seed(0)
dev = torch.device("cuda", 0)


def r(*s):
    return torch.randn(*s, generator=g).to(dev).requires_grad_(True)


def fb(y):
    y.backward(torch.ones_like(y))
    torch.cuda.synchronize()


n = 0
for (h, w, c, mult, k, s, pad) in [(9, 7, 8, 1, 3, 1, "SAME"), (9, 7, 8, 1, 3, 2, "SAME"), (11, 13, 12, 1, 5, 1, "SAME"),
                                   (11, 13, 12, 1, 5, 2, "VALID"), (9, 9, 36, 1, 7, 1, "SAME"), (9, 9, 36, 1, 7, 2, "SAME"),
                                   (6, 5, 6, 2, 3, 1, "SAME"), (6, 5, 5, 1, 3, 2, "VALID"), (6, 5, 8, 1, 4, 2, "SAME")]:
    fb(Fn.depthwise_conv2d(r(2, h, w, c), r(k, k, c, mult), r(c * mult), s, pad)); n += 1
for kind in ("relu6", "leaky_relu", "tanh", "sigmoid"):
    fb(Fn._act(r(3, 5, 7, 13), kind, 0.2)); n += 1
fb(Fn.instance_norm(r(2, 9, 7, 20), r(20), r(20), 1e-3)); n += 1
fb(Fn.instance_norm(r(1, 40, 40, 36), None, None, 1e-3)); n += 1
for bil in (False, True):
    for al in (False, True):
        f = Fn.resize_bilinear if bil else Fn.resize_nearest
        fb(f(r(2, 7, 5, 12), (10, 13), al)); fb(f(r(1, 8, 8, 8), (4, 4), al)); n += 2
for mode in ("constant", "reflect"):
    fb(Fn.pad_spatial(r(2, 9, 11, 12), (3, 0, 1, 2), mode)); n += 1
for (h, ci, co, k, s, pad) in [(5, 8, 12, 4, 2, "SAME"), (5, 16, 32, 3, 2, "SAME"), (1, 16, 24, 4, 1, "VALID"), (6, 8, 3, 4, 2, "SAME")]:
    fb(Fn.conv2d_transpose(r(2, h, h, ci), r(co, k, k, ci), r(co), s, pad)); n += 1
print(f"sanitize driver: {n} op cases ran forward + backward")
