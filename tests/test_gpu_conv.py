"""GPU: one convolution layer through the C ABI (idn_conv2d_nchw) against torch's fp32 conv2d on
the same bf16-rounded operands.  Both back-ends: the tcgen05 kernel (product) and the CUDA-core
cross-check, which must also agree with each other."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from idunno_b200 import conv2d_nchw

pytestmark = pytest.mark.gpu


def _bf16_round(a):
    return torch.from_numpy(np.asarray(a, np.float32)).to(torch.bfloat16).float()


def _case(n, cin, h, w, cout, r, stride, pad, residual, relu, seed):
    g = torch.Generator().manual_seed(seed)
    x = _bf16_round(torch.randn(n, cin, h, w, generator=g))
    wt = _bf16_round(torch.randn(cout, cin, r, r, generator=g) / np.sqrt(cin * r * r))
    scale = torch.empty(cout).uniform_(0.5, 1.5, generator=g)
    shift = torch.randn(cout, generator=g) * 0.1
    y = F.conv2d(x, wt, None, stride=stride, padding=pad)
    res = _bf16_round(torch.randn(y.shape, generator=g)) if residual else None
    y = y * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
    if res is not None:
        y = y + res
    if relu:
        y = F.relu(y)
    return x.numpy(), wt.numpy(), scale.numpy(), shift.numpy(), (res.numpy() if residual else None), y.numpy()


# (n, cin, h, w, cout, r, stride, pad, residual, relu)
SHAPES = [
    (2, 64, 8, 8, 64, 3, 1, 1, False, True),        # smallest: one k-pipeline wrap at most
    (3, 64, 56, 56, 64, 3, 1, 1, True, True),       # ResNet18 layer1
    (2, 64, 56, 56, 128, 3, 2, 1, False, True),     # layer2.0.conv1 (parity planes)
    (2, 64, 56, 56, 128, 1, 2, 0, False, False),    # layer2.0.downsample
    (5, 128, 28, 28, 128, 3, 1, 1, True, True),     # layer2
    (9, 256, 14, 14, 256, 3, 1, 1, True, True),     # layer3, BLOCK_N up to 256
    (20, 512, 7, 7, 512, 3, 1, 1, True, True),      # layer4
    (4, 64, 27, 27, 192, 5, 1, 2, False, True),     # AlexNet features.3 (25 taps, N=192)
    (6, 192, 13, 13, 384, 3, 1, 1, False, True),    # AlexNet features.6
    (2, 3, 224, 224, 64, 7, 2, 3, False, True),     # ResNet18 stem (5-D overlapping-stride TMA)
    (2, 3, 224, 224, 64, 11, 4, 2, False, True),    # AlexNet stem
]


def _err(got, want):
    scale = float(np.sqrt((want.astype(np.float64) ** 2).mean())) + 1e-6
    return float(np.abs(got - want).max()) / scale


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_conv_layer_matches_torch(built_library, shape, impl):
    n, cin, h, w, cout, r, stride, pad, residual, relu = shape
    x, wt, sc, sh, res, want = _case(*shape, seed=hash(shape) % 10000)
    got = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl=impl)
    assert got.shape == want.shape
    e = _err(got, want)
    # bf16 output rounding is 2^-9 relative; 2e-2 of the output RMS leaves room for large values
    assert e < 2e-2, f"{impl} {shape}: max err / rms = {e:.4f}"


def test_fc_shaped_gemm(built_library):
    """1x1 'convolution' on a 1x1 image == the FC layers (K = 4096, several N tiles, M tail)."""
    g = torch.Generator().manual_seed(5)
    n, k, m = 300, 1024, 512
    x = _bf16_round(torch.randn(n, k, 1, 1, generator=g))
    wt = _bf16_round(torch.randn(m, k, 1, 1, generator=g) / 32.0)
    b = torch.randn(m, generator=g)
    want = F.relu(F.conv2d(x, wt, b)).numpy()
    for impl in ("simt", "tc"):
        got = conv2d_nchw(x.numpy(), wt.numpy(), None, b.numpy(), None, relu=True, impl=impl)
        assert _err(got, want) < 2e-2, impl
