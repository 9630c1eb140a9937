"""Seeded synthetic inputs for the inference hot path.

There are no ImageNet JPEGs or pretrained checkpoints offline, so every
test, fixture and bench run uses:

* images  : ``synth_image(i)`` -- a structured uint8 HWC RGB picture whose content and size
            depend only on the image index ``i`` (seed = 1000 + i);
* weights : ``make_state_dict(model, seed, variant)`` -- torchvision's default initialisation
            under ``torch.manual_seed(seed)`` ("default"), or the same followed by a seeded
            randomisation of every BatchNorm parameter/buffer and a rescaled classifier
            ("stress"), which is what makes a wrong BN fold or a wrong flatten order visible.

Reference sites these stand in for: ``alexnet_resnet.py:18,21`` (pretrained checkpoints) and
``alexnet_resnet.py:49-50`` (test_<i>.JPEG files).
"""
from __future__ import annotations

import numpy as np

# Source sizes cycled through by ``synth_image(i, size=None)``: (H, W).
MIXED_SIZES = [(224, 224), (375, 500), (500, 375), (500, 333), (256, 256), (64, 96),
               (231, 487), (300, 224), (224, 299), (480, 640)]


def synth_size(i: int, mixed: bool) -> tuple[int, int]:
    if not mixed:
        return (224, 224)
    return MIXED_SIZES[i % len(MIXED_SIZES)]


def synth_image(i: int, size: tuple[int, int] | None = None, mixed: bool = False) -> np.ndarray:
    """Structured uint8 [H, W, 3] image for index ``i`` (low-frequency fields, a colour cast,
    a contrast change, a few rectangles and a little noise)."""
    h, w = size if size is not None else synth_size(i, mixed)
    rng = np.random.default_rng(1000 + int(i))
    yy = np.linspace(0.0, 1.0, h, dtype=np.float64)[:, None]
    xx = np.linspace(0.0, 1.0, w, dtype=np.float64)[None, :]
    img = np.zeros((h, w, 3), dtype=np.float64)
    for c in range(3):
        acc = np.zeros((h, w), dtype=np.float64)
        for _ in range(3):
            fy, fx = rng.uniform(0.5, 6.0, size=2)
            py, px = rng.uniform(0.0, 2 * np.pi, size=2)
            amp = rng.uniform(0.2, 1.0)
            acc += amp * np.sin(2 * np.pi * fy * yy + py) * np.cos(2 * np.pi * fx * xx + px)
        img[:, :, c] = acc
    img = img / (np.abs(img).max() + 1e-9)
    contrast = rng.uniform(0.3, 1.0)
    cast = rng.uniform(-0.35, 0.35, size=3)
    img = 0.5 + 0.5 * contrast * img + cast[None, None, :] * 0.5
    for _ in range(int(rng.integers(2, 6))):
        y0 = int(rng.integers(0, max(1, h - 4)))
        x0 = int(rng.integers(0, max(1, w - 4)))
        y1 = int(min(h, y0 + rng.integers(4, max(5, h // 2))))
        x1 = int(min(w, x0 + rng.integers(4, max(5, w // 2))))
        img[y0:y1, x0:x1, :] = rng.uniform(0.0, 1.0, size=3)[None, None, :]
    img += rng.normal(0.0, 0.02, size=img.shape)
    return np.clip(np.rint(img * 255.0), 0, 255).astype(np.uint8)


def noise_images(n: int, seed: int = 7, h: int = 224, w: int = 224) -> np.ndarray:
    """Uniform-noise uint8 [n, h, w, 3]; throughput runs only."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)


def make_state_dict(model: str, seed: int = 0, variant: str = "default"):
    """Seed-fixed random-init weights with torchvision's own constructors/initialisers
    (tv:models/alexnet.py:17-52, tv:models/resnet.py:166-285)."""
    import torch
    import torchvision

    name = "alexnet" if model == "alexnet" else "resnet18"
    torch.manual_seed(seed)
    net = getattr(torchvision.models, name)(weights=None)
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    if variant == "default":
        return sd
    if variant != "stress":
        raise ValueError(f"unknown weight variant {variant!r}")
    g = torch.Generator().manual_seed(seed + 12345)
    for k in list(sd.keys()):
        if k.endswith("running_mean"):
            base = k[: -len("running_mean")]
            n = sd[k].numel()
            sd[base + "weight"] = torch.empty(n).uniform_(0.6, 1.4, generator=g)
            sd[base + "bias"] = torch.empty(n).normal_(0.0, 0.15, generator=g)
            sd[base + "running_mean"] = torch.empty(n).normal_(0.0, 0.2, generator=g)
            sd[base + "running_var"] = torch.empty(n).uniform_(0.5, 1.5, generator=g)
    if name == "alexnet":
        # default-init AlexNet logits are bias dominated (+-0.03): give the biases some spread.
        for k in sd:
            if k.endswith("bias"):
                sd[k] = torch.empty(sd[k].shape).normal_(0.0, 0.05, generator=g)
        head_w, head_b = "classifier.6.weight", "classifier.6.bias"
    else:
        head_w, head_b = "fc.weight", "fc.bias"
    # Centre and rescale the classifier on a fixed calibration set so that logits depend on the
    # image (random features are dominated by a common mode, which would make every image map to
    # one class -- SURVEY.md 7.3-B) and softmax does not saturate.
    net.load_state_dict(sd)
    net.eval()
    feats = []
    hook = (net.classifier[6] if name == "alexnet" else net.fc).register_forward_hook(
        lambda m, i, o: feats.append(i[0].detach().clone()))
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    calib = np.stack([synth_image(900000 + j) for j in range(6)])
    x = (torch.from_numpy(calib).permute(0, 3, 1, 2).float() / 255.0 - mean) / std
    with torch.no_grad():
        net(x)
    hook.remove()
    f = feats[0]
    fbar = f.mean(dim=0)
    w = sd[head_w]
    centred = (f - fbar) @ w.t()
    scale = 2.5 / float(centred.std() + 1e-12)
    sd[head_w] = w * scale
    sd[head_b] = -(sd[head_w] @ fbar) + torch.empty(w.shape[0]).normal_(0.0, 0.3, generator=g)
    return sd
