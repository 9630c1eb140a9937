"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python oracle/make_golden.py

Imports /root/reference/alexnet_resnet.py as is, with the four offline shims of SURVEY.md 8c:
 1. empty ``matplotlib.pyplot`` / ``wget`` modules (imported at alexnet_resnet.py:6-7, unused);
 2. ``torch.hub.load`` -> torchvision model loaded with the seeded state_dict
    (``idunno_b200.synth.make_state_dict``), with a forward hook that records the model's
    input tensor and logits;
 3. ``imagenet_classes.txt`` written from torchvision's metadata;
 4. synthetic images saved losslessly as PNG bytes under ``./<folder>/test_<i>.JPEG``
    (``PIL.Image.open`` sniffs content, so the reference sees the exact uint8 array).
"""
from __future__ import annotations

import hashlib
import os
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))

REF = "/root/reference"
CASES = [  # (fixture name, model, weight variant, first index, last index, mixed sizes)
    ("alexnet_default", "alexnet", "default", 0, 7, True),
    ("alexnet_stress", "alexnet", "stress", 0, 11, True),
    ("resnet18_default", "resnet18", "default", 0, 7, True),
    ("resnet18_stress", "resnet18", "stress", 0, 11, True),
    # BASELINE.json configs[0] at full size: range 0-399, every file present, "diverse" weights (a stand-in
    # for a trained network: 400 distinct classes, margins far above the bf16 error; idunno_b200/synth.py)
    ("alexnet_diverse400", "alexnet", "diverse", 0, 399, True),
    ("resnet18_diverse400", "resnet18", "diverse", 0, 399, True),
]


def run_reference(model: str, variant: str, first: int, last: int, mixed: bool, skip=()):
    import torch
    import torchvision
    from PIL import Image
    from idunno_b200.synth import make_state_dict, synth_image

    for m in ("matplotlib", "matplotlib.pyplot", "wget"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import alexnet_resnet as ref  # the reference file, unmodified

    rec = {"pre": [], "logits": []}
    sd = make_state_dict(model, 0, variant)

    def fake_hub_load(repo, name, pretrained=True, **kw):
        net = getattr(torchvision.models, name)(weights=None)
        net.load_state_dict(sd)
        def hook(mod, inp, out):
            rec["pre"].append(inp[0][0].detach().numpy().copy())
            rec["logits"].append(out[0].detach().numpy().copy())

        net.register_forward_hook(hook)
        return net

    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            from torchvision.models._meta import _IMAGENET_CATEGORIES
            with open("imagenet_classes.txt", "w") as f:
                f.write("\n".join(_IMAGENET_CATEGORIES) + "\n")
            os.makedirs(model)
            for i in range(first, last + 1):
                if i in skip:
                    continue
                Image.fromarray(synth_image(i, mixed=mixed)).save(
                    os.path.join(model, f"test_{i}.JPEG"), format="PNG")
            orig = torch.hub.load
            torch.hub.load = fake_hub_load
            try:
                result, elapsed = ref.deeplearning(model, model, str(first), str(last))
            finally:
                torch.hub.load = orig
        finally:
            os.chdir(cwd)
    return result, elapsed, rec, sd


def main():
    from torchvision.models._meta import _IMAGENET_CATEGORIES
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    only = sys.argv[1:]
    for name, model, variant, first, last, mixed in CASES:
        if only and name not in only:
            continue
        big = last - first >= 100
        skip = () if big else (first + 2,)  # one missing file: the reference skips it silently (:49)
        result, elapsed, rec, sd = run_reference(model, variant, first, last, mixed, skip)
        names = [r[0] for r in result]
        cats = [r[1] for r in result]
        probs = np.asarray([r[2] for r in result], dtype=np.float64)
        idx = np.asarray([_IMAGENET_CATEGORIES.index(c) for c in cats], dtype=np.int64)
        logits = np.stack(rec["logits"]).astype(np.float32)
        pre = np.stack(rec["pre"]).astype(np.float32)
        pre_sha = [hashlib.sha256(p.tobytes()).hexdigest() for p in pre]
        w_sha = hashlib.sha256(b"".join(v.numpy().tobytes() for v in sd.values())).hexdigest()
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            model=model, variant=variant, first=first, last=last, mixed=mixed,
            skipped=np.asarray(skip, dtype=np.int64), names=np.asarray(names), categories=np.asarray(cats),
            idx=idx, prob=probs, logits=logits, pre_sha256=np.asarray(pre_sha),
            pre_sample=pre[:8, :, ::16, ::16].copy(), weights_sha256=w_sha)
        srt = np.sort(logits, axis=1)
        print(f"{name}: {len(result)} results, {elapsed:.2f}s, {len(set(idx.tolist()))} distinct classes, "
              f"median top-1/top-2 margin {np.median(srt[:, -1] - srt[:, -2]):.4f}"
              + ("" if big else f", classes={sorted(set(idx.tolist()))}"))


if __name__ == "__main__":
    main()
