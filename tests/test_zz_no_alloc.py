"""`dtf_bn_forward_train_static` — the batch-norm training forward for callers that cannot allocate.

Some users of the kernel library are not allowed to create device memory at the point where they call us: a
serving / framework plug-in that may only enqueue work on the stream it was handed, a process whose memory
pool is already fully owned by its host framework, or a tight loop where `cudaMalloc` (a device-wide
synchronisation) is unacceptable.  For them the per-channel sum scratch of the BN forward lives in a static
`__device__` array inside the library instead of a caller-provided buffer.

The test drives the entry point through the C ABI (ctypes, raw pointers) exactly as such a caller would, and
checks it against the ordinary entry point (`dtf_bn_forward_train`, caller-provided scratch) on two calls in a
row with different shapes — the second call would see stale sums if the shared array were not cleared —
plus the documented refusal of channel counts beyond the built-in maximum.  Inputs/outputs are torch tensors
only for convenience; the entry point itself allocates nothing.  Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_C = 4096      # DTF_BN_STATIC_MAX_CHANNELS in include/dtf_b200.h


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _call(lib, static, x, gamma, beta, mm, mv, relu, coef, y):
    m, c = x.numel() // x.shape[-1], x.shape[-1]
    vp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    common = [vp(x), C.c_longlong(m), c, c, vp(gamma), vp(beta), C.c_float(1e-3), C.c_float(0.9), 1, 1,
              vp(mm), vp(mv), int(relu)]
    if static:
        return lib.dtf_bn_forward_train_static(*common, vp(coef), vp(y), c, stream)
    scratch = torch.empty(2 * c, device=x.device)
    return lib.dtf_bn_forward_train(*common, vp(scratch), vp(coef), vp(y), c, stream)


def _run_pair(lib, shape, relu, seed):
    g = torch.Generator().manual_seed(seed)
    c = shape[-1]
    x = (torch.randn(*shape, generator=g) * 2 + 0.3).cuda()
    gamma = (torch.rand(c, generator=g) + 0.5).cuda()
    beta = (torch.randn(c, generator=g) * 0.1).cuda()
    outs = []
    for static in (False, True):
        mm, mv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
        coef, y = torch.empty(4 * c, device="cuda"), torch.empty_like(x)
        rc = _call(lib, static, x, gamma, beta, mm, mv, relu, coef, y)
        assert rc == 0, lib.dtf_last_error()
        torch.cuda.synchronize()
        outs.append((y, coef, mm, mv))
    for a, b in zip(outs[1], outs[0]):
        assert _rel(a, b) < 2e-4
    # and against the plain PyTorch definition of the op
    xf = x.reshape(-1, c)
    ref = (xf - xf.mean(0)) * torch.rsqrt(xf.var(0, unbiased=False) + 1e-3) * gamma + beta
    if relu:
        ref = torch.relu(ref)
    assert _rel(outs[1][0].reshape(-1, c), ref) < 2e-4


def test_static_scratch_variant_matches_regular_entry_point_twice_in_a_row():
    from distributed_tensorflow_models_b200 import _native as nat
    lib = nat.lib()
    lib.dtf_last_error.restype = C.c_char_p
    # first call: wide tensor (fills a large part of the shared array) ...
    _run_pair(lib, (4, 6, 6, 448), relu=True, seed=1)
    # ... second call straight after it, narrower and with different data: stale sums would show up here
    _run_pair(lib, (8, 9, 9, 64), relu=False, seed=2)


def test_static_scratch_variant_rejects_oversize():
    from distributed_tensorflow_models_b200 import _native as nat
    lib = nat.lib()
    lib.dtf_last_error.restype = C.c_char_p
    c = MAX_C + 4
    x = torch.zeros(2, 1, 1, c, device="cuda")
    mm, mv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    coef, y = torch.empty(4 * c, device="cuda"), torch.empty_like(x)
    rc = _call(lib, True, x, None, None, mm, mv, False, coef, y)
    assert rc != 0
    assert b"exceeds" in lib.dtf_last_error()
    # the regular entry point accepts the same size
    assert _call(lib, False, x, None, None, mm, mv, False, coef, y) == 0
    torch.cuda.synchronize()
