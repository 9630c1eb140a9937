"""Host-side image staging: folder of test_<i>.JPEG files -> uint8 HWC arrays.

Reproduces the discovery loop and decode of the reference (alexnet_resnet.py:24,46-54): the folder
is a CWD-relative glob, the index range is inclusive, a missing file is skipped silently, and a
non-RGB file is converted and rewritten in place.  JPEG decode stays on the host.
"""
from __future__ import annotations

import glob
import os

import numpy as np


def load_range(filename: str, start_image, end_image):
    """Returns (indices, images) for every existing ./<filename>/test_<i>.JPEG, i in [start, end]."""
    from PIL import Image
    indices, images = [], []
    current = int(start_image)
    for folder in glob.glob("./" + str(filename)):
        while current <= int(end_image):
            for f in glob.glob(folder + "/test_" + str(current) + ".JPEG"):
                im = Image.open(f)
                if im.mode != "RGB":
                    im = im.convert("RGB")
                    os.remove(f)
                    im.save(f)
                indices.append(current)
                images.append(np.asarray(im, dtype=np.uint8))
            current += 1
    return indices, images


def categories():
    """imagenet_classes.txt in the CWD (alexnet_resnet.py:29-42).  The reference downloads the
    file when absent; offline it is written from torchvision's metadata (same 1000 names)."""
    path = "imagenet_classes.txt"
    if not os.path.exists(path):
        from torchvision.models._meta import _IMAGENET_CATEGORIES
        with open(path, "w") as f:
            f.write("\n".join(_IMAGENET_CATEGORIES) + "\n")
    with open(path, "r") as f:
        return [s.strip() for s in f.readlines()]
