"""GPU: the preprocess kernel against the oracle, through the C ABI (idn_preprocess_u8hwc).
Bar: bit-exact.  The uint8 resize/crop stage is integer work; the normalised value is
round_to_bf16 of the oracle's fp32 value."""
import numpy as np
import pytest
import torch

from idunno_b200 import preprocess
from idunno_b200.synth import MIXED_SIZES, noise_images, synth_image
from oracle import pil_resize

pytestmark = pytest.mark.gpu


def _bf16_bits(x: np.ndarray) -> np.ndarray:
    return torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)


def _check(images):
    got = preprocess(images)
    for i, img in enumerate(images):
        want = _bf16_bits(pil_resize.preprocess(img))
        bad = int((got[i] != want).sum())
        assert bad == 0, f"image {i} {img.shape}: {bad} of {want.size} bf16 values differ"


def test_preprocess_224_sources(built_library):
    _check([synth_image(i) for i in range(6)])


def test_preprocess_mixed_sizes_ragged_batch(built_library):
    _check([synth_image(i, mixed=True) for i in range(len(MIXED_SIZES))])


def test_preprocess_extreme_shapes(built_library):
    sizes = [(224, 224), (256, 256), (225, 1000), (1000, 225), (32, 32), (257, 255), (1, 1), (3, 700)]
    _check([synth_image(100 + i, size=s) for i, s in enumerate(sizes)])


def test_preprocess_saturated_and_noise(built_library):
    imgs = [np.zeros((224, 224, 3), np.uint8), np.full((300, 200, 3), 255, np.uint8)]
    imgs += list(noise_images(3, seed=3, h=240, w=320))
    _check(imgs)
