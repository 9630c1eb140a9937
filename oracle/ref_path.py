"""TEST INFRASTRUCTURE ONLY -- CPU restatement of ``alexnet_resnet.deeplearning``.

Follows /root/reference/alexnet_resnet.py:12-92 line by line (file discovery loop, inclusive
range, per-image batch of one, softmax, top-1) but takes the seeded state_dict explicitly instead
of ``torch.hub.load(..., pretrained=True)`` (``alexnet_resnet.py:18,21`` -- needs the network).

The network arithmetic itself is third-party and not under /root/reference: ``pytorch/vision``
pinned by tag v0.10.0 through torch.hub (``alexnet_resnet.py:18,21``).  Its published
architectures are restated below with ``torch.nn.functional`` in fp32, eval mode:
AlexNet   tv:models/alexnet.py:21-52, ResNet18 tv:models/resnet.py:59-105,166-285.

Pinning: the reference has NO golden vectors, known-answer tests or fixtures for this path
(SURVEY.md 8c).  ``oracle/make_golden.py`` runs the UNMODIFIED reference file in the build
container (four offline shims) and writes tests/golden/*.npz; tests/test_oracle.py checks this
restatement against those fixtures and against torchvision's own model classes.
"""
from __future__ import annotations

import glob
import os
import time

import numpy as np
import torch
import torch.nn.functional as F

from . import pil_resize


def alexnet_forward(sd: dict, x: torch.Tensor) -> torch.Tensor:
    """tv:models/alexnet.py:21-52.  Dropout is inactive in eval (``alexnet_resnet.py:19``);
    AdaptiveAvgPool2d((6, 6)) is the identity for 224x224 inputs but is kept for other sizes."""
    x = F.relu(F.conv2d(x, sd["features.0.weight"], sd["features.0.bias"], stride=4, padding=2))
    x = F.max_pool2d(x, 3, 2)
    x = F.relu(F.conv2d(x, sd["features.3.weight"], sd["features.3.bias"], padding=2))
    x = F.max_pool2d(x, 3, 2)
    x = F.relu(F.conv2d(x, sd["features.6.weight"], sd["features.6.bias"], padding=1))
    x = F.relu(F.conv2d(x, sd["features.8.weight"], sd["features.8.bias"], padding=1))
    x = F.relu(F.conv2d(x, sd["features.10.weight"], sd["features.10.bias"], padding=1))
    x = F.max_pool2d(x, 3, 2)
    x = F.adaptive_avg_pool2d(x, (6, 6))
    x = torch.flatten(x, 1)
    x = F.relu(F.linear(x, sd["classifier.1.weight"], sd["classifier.1.bias"]))
    x = F.relu(F.linear(x, sd["classifier.4.weight"], sd["classifier.4.bias"]))
    return F.linear(x, sd["classifier.6.weight"], sd["classifier.6.bias"])


def _bn(sd: dict, p: str, x: torch.Tensor) -> torch.Tensor:
    # eval-mode BatchNorm2d, eps=1e-5 (tv:models/resnet.py:197; torch default eps)
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"],
                        sd[p + ".weight"], sd[p + ".bias"], training=False, eps=1e-5)


def _basic_block(sd: dict, p: str, x: torch.Tensor, stride: int) -> torch.Tensor:
    """tv:models/resnet.py:89-105 (BasicBlock.forward)."""
    out = F.relu(_bn(sd, p + ".bn1", F.conv2d(x, sd[p + ".conv1.weight"], None, stride=stride, padding=1)))
    out = _bn(sd, p + ".bn2", F.conv2d(out, sd[p + ".conv2.weight"], None, stride=1, padding=1))
    if (p + ".downsample.0.weight") in sd:
        idn = _bn(sd, p + ".downsample.1", F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride=stride))
    else:
        idn = x
    return F.relu(out + idn)


def resnet18_forward(sd: dict, x: torch.Tensor) -> torch.Tensor:
    """tv:models/resnet.py:266-282 (_forward_impl)."""
    x = F.relu(_bn(sd, "bn1", F.conv2d(x, sd["conv1.weight"], None, stride=2, padding=3)))
    x = F.max_pool2d(x, 3, 2, 1)
    for li, stride in ((1, 1), (2, 2), (3, 2), (4, 2)):
        x = _basic_block(sd, f"layer{li}.0", x, stride)
        x = _basic_block(sd, f"layer{li}.1", x, 1)
    x = torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)
    return F.linear(x, sd["fc.weight"], sd["fc.bias"])


def forward(modelname: str, sd: dict, x: torch.Tensor) -> torch.Tensor:
    # any modelname other than "alexnet" selects ResNet18 (alexnet_resnet.py:17-22)
    with torch.no_grad():
        return alexnet_forward(sd, x) if modelname == "alexnet" else resnet18_forward(sd, x)


def preprocess_library(img: np.ndarray) -> np.ndarray:
    """The reference's own preprocess call chain (alexnet_resnet.py:57-66): PIL image through
    torchvision's Compose.  Bit-identical to ``pil_resize.preprocess`` (tests/test_oracle.py) and
    much faster (Pillow's C resample); used where the CPU path is TIMED."""
    from PIL import Image
    from torchvision import transforms
    tf = transforms.Compose([                                               # :57-62, rebuilt per image
        transforms.Resize(256), transforms.CenterCrop(224), transforms.ToTensor(),
        transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    return tf(Image.fromarray(img)).numpy()


def infer_arrays(modelname: str, sd: dict, images: list[np.ndarray], library_preprocess: bool = False):
    """Hot path on in-memory uint8 HWC images, one image per forward as the reference does
    (``alexnet_resnet.py:66-84``).  Returns dict(pre[n,3,224,224] f32, logits[n,1000] f32,
    idx[n] int64, prob[n] f32)."""
    pre, logits, idx, prob = [], [], [], []
    prep = preprocess_library if library_preprocess else pil_resize.preprocess
    for img in images:
        t = torch.from_numpy(prep(img)).unsqueeze(0)                         # :66-67
        out = forward(modelname, sd, t)                                      # :74-75
        p = F.softmax(out[0], dim=0)                                         # :80
        top_prob, top_catid = torch.topk(p, 1)                               # :84
        pre.append(t[0].numpy()); logits.append(out[0].numpy())
        idx.append(int(top_catid[0])); prob.append(float(top_prob[0]))
    return dict(pre=np.stack(pre) if pre else np.zeros((0, 3, 224, 224), np.float32),
                logits=np.stack(logits) if logits else np.zeros((0, 1000), np.float32),
                idx=np.asarray(idx, dtype=np.int64), prob=np.asarray(prob, dtype=np.float32))


def load_image_file(path: str) -> np.ndarray:
    """``Image.open`` + RGB conversion (alexnet_resnet.py:50-52); the reference also rewrites a
    non-RGB file in place (:53-54), reproduced here."""
    from PIL import Image
    im = Image.open(path)
    if im.mode != "RGB":
        im = im.convert("RGB")
        os.remove(path)
        im.save(path)
    return np.asarray(im, dtype=np.uint8)


def deeplearning(filename, modelname, start_image, end_image, sd: dict, categories: list[str],
                 tap: dict | None = None):
    """Same loop structure and result format as alexnet_resnet.py:12-92.  ``tap`` (optional)
    receives the per-image arrays of ``infer_arrays`` for parity tests."""
    start_image = str(start_image); end_image = str(end_image)            # :13-14
    inference_result = []
    time_start = time.time()                                               # :16
    file_list = glob.glob("./" + filename)                                 # :24
    current_image = start_image
    taps = []
    for folder in file_list:                                               # :46
        while int(current_image) <= int(end_image):                        # :48 inclusive
            for f in glob.glob(folder + "/test_" + str(current_image) + ".JPEG"):   # :49
                img = load_image_file(f)
                r = infer_arrays(modelname, sd, [img])
                taps.append(r)
                image_name = "test_" + str(current_image) + ".JPEG"      # :86
                inference_result.append((image_name, categories[int(r["idx"][0])], float(r["prob"][0])))
            current_image = int(current_image) + 1                         # :90
    time_end = time.time()
    if tap is not None and taps:
        for k in ("pre", "logits", "idx", "prob"):
            tap[k] = np.concatenate([t[k] for t in taps])
    return inference_result, (time_end - time_start)                       # :92


def imagenet_categories() -> list[str]:
    """The 1000 lines of imagenet_classes.txt (alexnet_resnet.py:29-42).  The file is downloaded
    by the reference; offline the same names come from torchvision's metadata."""
    from torchvision.models._meta import _IMAGENET_CATEGORIES
    return list(_IMAGENET_CATEGORIES)


def forward_taps(modelname: str, sd: dict, x: torch.Tensor) -> dict:
    """Intermediate activations of the restated forward, keyed by the plan-step names of the CUDA
    engine (csrc/engine.cu), for layer-by-layer parity tests.  fp32 NCHW."""
    taps = {}
    with torch.no_grad():
        if modelname == "alexnet":
            x = F.relu(F.conv2d(x, sd["features.0.weight"], sd["features.0.bias"], stride=4, padding=2))
            taps["features.0+relu"] = x
            x = F.max_pool2d(x, 3, 2); taps["features.2 maxpool"] = x
            x = F.relu(F.conv2d(x, sd["features.3.weight"], sd["features.3.bias"], padding=2))
            taps["features.3+relu"] = x
            x = F.max_pool2d(x, 3, 2); taps["features.5 maxpool"] = x
            for k in ("6", "8", "10"):
                x = F.relu(F.conv2d(x, sd[f"features.{k}.weight"], sd[f"features.{k}.bias"], padding=1))
                taps[f"features.{k}+relu"] = x
            x = F.max_pool2d(x, 3, 2); taps["features.12 maxpool"] = x
            x = torch.flatten(F.adaptive_avg_pool2d(x, (6, 6)), 1)
            x = F.relu(F.linear(x, sd["classifier.1.weight"], sd["classifier.1.bias"]))
            taps["classifier.1+relu"] = x[:, :, None, None]
            x = F.relu(F.linear(x, sd["classifier.4.weight"], sd["classifier.4.bias"]))
            taps["classifier.4+relu"] = x[:, :, None, None]
            x = F.linear(x, sd["classifier.6.weight"], sd["classifier.6.bias"])
            taps["classifier.6"] = x[:, :, None, None]
            return taps
        x = F.relu(_bn(sd, "bn1", F.conv2d(x, sd["conv1.weight"], None, stride=2, padding=3)))
        taps["conv1+bn1+relu"] = x
        x = F.max_pool2d(x, 3, 2, 1); taps["maxpool"] = x
        for li, stride in ((1, 1), (2, 2), (3, 2), (4, 2)):
            for bi, st in ((0, stride), (1, 1)):
                p = f"layer{li}.{bi}"
                out = F.relu(_bn(sd, p + ".bn1", F.conv2d(x, sd[p + ".conv1.weight"], None, stride=st, padding=1)))
                taps[p + ".conv1"] = out
                out = _bn(sd, p + ".bn2", F.conv2d(out, sd[p + ".conv2.weight"], None, padding=1))
                if (p + ".downsample.0.weight") in sd:
                    idn = _bn(sd, p + ".downsample.1", F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride=st))
                    taps[p + ".downsample"] = idn
                else:
                    idn = x
                x = F.relu(out + idn)
                taps[p + ".conv2+add"] = x
    return taps


# ---------------------------------------------------------------------------------------------
# Precision contract of the CUDA path, restated on the CPU: bf16 operands (weights and every
# stored activation rounded to bf16, round-to-nearest-even), fp32 accumulation, BatchNorm / bias
# applied in fp32 as a per-channel scale/shift after the convolution, residual added in fp32,
# ResNet18's last FC in fp32, AlexNet's last FC with bf16 weights and fp32 output.  Used by the
# tests to state the bf16-vs-fp32 tolerance: |this - fp32 oracle| is what bf16 arithmetic costs on
# a given weight set, independently of any kernel.
# ---------------------------------------------------------------------------------------------
def _q(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def _fold(sd, p):
    s = sd[p + ".weight"] / torch.sqrt(sd[p + ".running_var"] + 1e-5)
    return s, sd[p + ".bias"] - sd[p + ".running_mean"] * s


def forward_bf16_contract(modelname: str, sd: dict, x: torch.Tensor) -> torch.Tensor:
    with torch.no_grad():
        x = _q(x)
        if modelname == "alexnet":
            def cv(x, k, **kw):
                return _q(F.relu(F.conv2d(x, _q(sd[f"features.{k}.weight"]), None, **kw)
                                 + sd[f"features.{k}.bias"].view(1, -1, 1, 1)))
            x = F.max_pool2d(cv(x, 0, stride=4, padding=2), 3, 2)
            x = F.max_pool2d(cv(x, 3, padding=2), 3, 2)
            x = cv(cv(cv(x, 6, padding=1), 8, padding=1), 10, padding=1)
            x = torch.flatten(F.max_pool2d(x, 3, 2), 1)
            x = _q(F.relu(F.linear(x, _q(sd["classifier.1.weight"])) + sd["classifier.1.bias"]))
            x = _q(F.relu(F.linear(x, _q(sd["classifier.4.weight"])) + sd["classifier.4.bias"]))
            return F.linear(x, _q(sd["classifier.6.weight"])) + sd["classifier.6.bias"]

        def cbn(x, conv, bn, stride=1, padding=0):
            s, b = _fold(sd, bn)
            return F.conv2d(x, _q(sd[conv + ".weight"]), None, stride=stride, padding=padding) \
                * s.view(1, -1, 1, 1) + b.view(1, -1, 1, 1)
        x = F.max_pool2d(_q(F.relu(cbn(x, "conv1", "bn1", 2, 3))), 3, 2, 1)
        for li, stride in ((1, 1), (2, 2), (3, 2), (4, 2)):
            for bi, st in ((0, stride), (1, 1)):
                p = f"layer{li}.{bi}"
                out = _q(F.relu(cbn(x, p + ".conv1", p + ".bn1", st, 1)))
                idn = _q(cbn(x, p + ".downsample.0", p + ".downsample.1", st, 0)) \
                    if (p + ".downsample.0.weight") in sd else x
                x = _q(F.relu(cbn(out, p + ".conv2", p + ".bn2", 1, 1) + idn))
        x = torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)
        return F.linear(x, sd["fc.weight"], sd["fc.bias"])
