"""TEST / BASELINE INFRASTRUCTURE ONLY -- runs the reference's own `deeplearning()` (oracle/_ref, the
byte-compiled unmodified /root/reference/alexnet_resnet.py; see build_ref.py) on synthetic inputs with the
four offline shims of SURVEY.md 8c: stub `matplotlib.pyplot` / `wget` (imported, unused), `torch.hub.load`
-> seeded torchvision model, `imagenet_classes.txt` from torchvision's metadata, images written losslessly
as PNG bytes under ./<model>/test_<i>.JPEG.  Used by bench.py (reference arm, cpu_baseline) only."""
from __future__ import annotations

import importlib.machinery
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
PYC = os.path.join(HERE, "_ref", "alexnet_resnet.bin")


def available() -> bool:
    return os.path.exists(PYC)


def load_reference():
    """The reference module, imported from its byte code (never from this repository's drop-in of the same name)."""
    for m in ("matplotlib", "matplotlib.pyplot", "wget"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    loader = importlib.machinery.SourcelessFileLoader("reference_alexnet_resnet", PYC)
    spec = importlib.util.spec_from_loader("reference_alexnet_resnet", loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod


class ReferenceRun:
    """Working directory with ./<model>/test_<i>.JPEG and imagenet_classes.txt; `deeplearning(model, s, e)`
    calls the reference with `torch.hub.load` answering from the seeded state_dict (model construction and
    weight load happen inside every call, as in the reference: alexnet_resnet.py:16-22)."""

    def __init__(self, workdir: str, state_dicts: dict):
        self.workdir, self.sds, self.ref = workdir, state_dicts, load_reference()
        from torchvision.models._meta import _IMAGENET_CATEGORIES
        with open(os.path.join(workdir, "imagenet_classes.txt"), "w") as f:
            f.write("\n".join(_IMAGENET_CATEGORIES) + "\n")

    def write_images(self, model: str, images: dict):
        from PIL import Image
        os.makedirs(os.path.join(self.workdir, model), exist_ok=True)
        for i, arr in images.items():
            Image.fromarray(arr).save(os.path.join(self.workdir, model, f"test_{i}.JPEG"), format="PNG")

    def deeplearning(self, model: str, start: int, end: int):
        import torch
        import torchvision

        def fake_hub_load(repo, name, pretrained=True, **kw):
            net = getattr(torchvision.models, name)(weights=None)
            net.load_state_dict(self.sds[name])
            return net

        cwd, orig = os.getcwd(), torch.hub.load
        os.chdir(self.workdir)
        torch.hub.load = fake_hub_load
        try:
            return self.ref.deeplearning(model, model, str(start), str(end))
        finally:
            torch.hub.load = orig
            os.chdir(cwd)
