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
    (2, 3, 224, 224, 64, 7, 2, 3, False, True),     # ResNet18 stem (overlapping no-swizzle descriptor)
    (5, 3, 224, 224, 64, 7, 2, 3, False, False),    # same, more tiles than SMs, no ReLU
]


def _err(got, want):
    """max over elements of |got - want| / (|want| + rms(want)).  The output is stored as bf16
    (half an ulp = 2^-9 = 0.2 % of the value); fp32 accumulation-order noise is far below that."""
    rms = float(np.sqrt((want.astype(np.float64) ** 2).mean())) + 1e-6
    return float((np.abs(got - want) / (np.abs(want) + rms)).max())


TOL = 6e-3


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_conv_layer_matches_torch(built_library, shape, impl):
    n, cin, h, w, cout, r, stride, pad, residual, relu = shape
    x, wt, sc, sh, res, want = _case(*shape, seed=hash(shape) % 10000)
    got = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl=impl)
    assert got.shape == want.shape
    e = _err(got, want)
    assert e < TOL, f"{impl} {shape}: max err = {e:.4f}"


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
        assert _err(got, want) < TOL, impl


@pytest.mark.parametrize("block_n,shape", [
    (256, (9, 256, 14, 14, 256, 3, 1, 1, True, True)),
    (256, (20, 512, 7, 7, 512, 3, 1, 1, True, True)),
    (192, (4, 64, 27, 27, 192, 5, 1, 2, False, True)),
    (192, (6, 192, 13, 13, 384, 3, 1, 1, False, True)),
    (64, (5, 128, 28, 28, 128, 3, 1, 1, True, True)),
], ids=lambda v: str(v).replace(" ", ""))
def test_every_n_tile_width(built_library, monkeypatch, block_n, shape):
    """Small problems pick narrow N tiles by themselves; force the wide instantiations
    (BLOCK_N = 192 / 256: 5- and 4-stage rings, 512 TMEM columns) that batch-400 plans use."""
    monkeypatch.setenv("IDN_BLOCK_N", str(block_n))
    monkeypatch.setenv("IDN_NO_HALO", "1")       # keep the layer on conv_tc_kernel
    n, cin, h, w, cout, r, stride, pad, residual, relu = shape
    x, wt, sc, sh, res, want = _case(*shape, seed=block_n + n)
    got = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl="tc")
    assert _err(got, want) < TOL


@pytest.mark.parametrize("shape", [SHAPES[1], SHAPES[4], SHAPES[5], SHAPES[7], SHAPES[8]],
                         ids=lambda s: "x".join(map(str, s)))
def test_halo_and_per_tap_kernels_agree(built_library, monkeypatch, shape):
    """conv_halo_kernel (one staged A tile for all taps, resident / reused weights) and
    conv_tc_kernel (one TMA box per tap) differ only in accumulation order."""
    n, cin, h, w, cout, r, stride, pad, residual, relu = shape
    monkeypatch.setenv("IDN_HALO_WIDE", "1")     # also the streamed-weights instantiations
    x, wt, sc, sh, res, want = _case(*shape, seed=11)
    a = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl="tc")
    monkeypatch.setenv("IDN_NO_HALO", "1")
    b = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl="tc")
    assert _err(a, want) < TOL and _err(b, want) < TOL
    assert _err(a, b) < TOL


def test_many_tiles_per_cta(built_library):
    """More M tiles than SMs: every CTA loops over several tiles, so both TMEM accumulator stages
    and several wraps of the smem ring are exercised (persistent scheduling)."""
    shape = (40, 64, 56, 56, 64, 3, 1, 1, True, True)     # 40*57*57/128 = 1016 tiles
    x, wt, sc, sh, res, want = _case(*shape, seed=77)
    got = conv2d_nchw(x, wt, sc, sh, res, stride=1, pad=1, relu=True, impl="tc")
    assert _err(got, want) < TOL


# Product layers at sizes that select the kernels the batch-400 plans use (round-1 review: these ran only inside
# whole-model tests of a few images, where the planner picks narrower tiles).
PRODUCT_SHAPES = [
    (96, 128, 28, 28, 256, 3, 2, 1, False, True),    # layer3.0.conv1: stride-2 3x3, conv_tc_pair_kernel<256>
    (150, 256, 14, 14, 512, 3, 2, 1, False, True),   # layer4.0.conv1: the same kernel, two N tiles
    (8, 128, 28, 28, 256, 1, 2, 0, False, False),    # layer3.0.downsample
    (8, 256, 14, 14, 512, 1, 2, 0, False, False),    # layer4.0.downsample
    (6, 384, 13, 13, 256, 3, 1, 1, False, True),     # AlexNet features.8: conv_pair_kernel<256, 3>, Cin 384
    (6, 256, 13, 13, 256, 3, 1, 1, False, True),     # AlexNet features.10
]


@pytest.mark.parametrize("shape", PRODUCT_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_product_layer_shapes(built_library, shape):
    n, cin, h, w, cout, r, stride, pad, residual, relu = shape
    x, wt, sc, sh, res, want = _case(*shape, seed=(n * 131 + cin) % 10000)
    got = conv2d_nchw(x, wt, sc, sh, res, stride=stride, pad=pad, relu=relu, impl="tc")
    assert got.shape == want.shape
    assert _err(got, want) < TOL


def test_fc6_shape_on_cta_pairs(built_library):
    """AlexNet classifier.1 as the batch-400 plan runs it: M = 400, K = 9216, N = 4096 on conv_tc_pair_kernel<128>
    (144 k-blocks through the ring, 32 N tiles, two pair items per N tile)."""
    g = torch.Generator().manual_seed(9)
    n, k, m = 400, 9216, 4096
    x = _bf16_round(torch.randn(n, k, 1, 1, generator=g))
    wt = _bf16_round(torch.randn(m, k, 1, 1, generator=g) / 96.0)
    b = torch.randn(m, generator=g)
    want = F.relu(F.conv2d(x, wt, b)).numpy()
    got = conv2d_nchw(x.numpy(), wt.numpy(), None, b.numpy(), None, relu=True, impl="tc")
    assert _err(got, want) < TOL


def test_alexnet_stem_as_space_to_depth_conv(built_library):
    """AlexNet features.0 (11x11, stride 4, pad 2) the way the plan computes it: a 3x3 stride-1 convolution over
    the 4x4 space-to-depth image (57 x 57 x 48(+16), block origin at pixel (-2, -2); engine.cu build_alexnet).
    Built here on the host from an NCHW image and the 11x11 weights, run through the library's 3x3 kernel with
    'same' padding, and compared on the 55 x 55 interior with torch's strided convolution."""
    g = torch.Generator().manual_seed(13)
    n = 5
    x = _bf16_round(torch.randn(n, 3, 224, 224, generator=g))
    w11 = _bf16_round(torch.randn(64, 3, 11, 11, generator=g) / np.sqrt(363.0))
    bias = torch.randn(64, generator=g) * 0.1
    want = F.relu(F.conv2d(x, w11, bias, stride=4, padding=2)).numpy()                     # [n, 64, 55, 55]
    xp = F.pad(x, (2, 2, 2, 2))                                                             # 228 = 57 * 4
    s2d = xp.view(n, 3, 57, 4, 57, 4).permute(0, 3, 5, 1, 2, 4).reshape(n, 48, 57, 57)      # channel (dy*4+dx)*3 + c
    s2d = torch.cat([s2d, torch.zeros(n, 16, 57, 57)], dim=1)
    w3 = torch.zeros(64, 64, 3, 3)
    for r in range(11):
        for s in range(11):
            a, dy, b_, dx = r // 4, r % 4, s // 4, s % 4
            w3[:, (dy * 4 + dx) * 3:(dy * 4 + dx) * 3 + 3, a, b_] = w11[:, :, r, s]
    got = conv2d_nchw(s2d.numpy(), w3.numpy(), None, bias.numpy(), None, stride=1, pad=1, relu=True, impl="tc")
    # 'same' output pixel (i, j) = valid output (i - 1, j - 1): valid rows/cols 0..54 sit at 1..55
    assert _err(got[:, :, 1:56, 1:56], want) < TOL
