"""Drop-in replacement of the reference's worker module ``alexnet_resnet`` (same module name, same
``deeplearning`` signature and result format; reference: alexnet_resnet.py:12-92, called from
mp4_machinelearning.py:596 and :765).

What changes underneath: images of the range are decoded on the host exactly as before, then the
whole range goes through the B200 engine (idunno_b200, C ABI in include/idunno_b200.h) in tensor
batches instead of one ``model(input_batch)`` per image; the model is built once per process and
kept resident instead of ``torch.hub.load`` per call.

Weights: the reference downloads torchvision's pretrained checkpoints (``pretrained=True``,
:18,21).  This module loads ``<modelname>.pth`` (a torchvision state_dict) from the directory in
``IDUNNO_WEIGHTS_DIR`` (default: CWD).  A missing checkpoint is an ERROR, as a failed download is in
the reference: a worker with a wrong weights directory must not answer with well-formed but
meaningless tuples.  Only with ``IDUNNO_ALLOW_RANDOM_WEIGHTS=1`` (tests, benchmarks, offline demos)
does it fall back to the seed-fixed synthetic weights of ``idunno_b200.synth`` and say so on stderr.
"""
from __future__ import annotations

import os
import sys
import time

_PKG = os.path.dirname(os.path.abspath(__file__))
if _PKG not in sys.path:
    sys.path.insert(0, _PKG)

from idunno_b200 import Engine, staging, synth  # noqa: E402

DEBUG_MODE = False
TENSOR_BATCH = int(os.environ.get("IDUNNO_TENSOR_BATCH", "400"))
_engines: dict = {}


def _state_dict(name: str):
    path = os.path.join(os.environ.get("IDUNNO_WEIGHTS_DIR", "."), name + ".pth")
    if os.path.exists(path):
        import torch
        return torch.load(path, map_location="cpu")
    if os.environ.get("IDUNNO_ALLOW_RANDOM_WEIGHTS") != "1":
        raise FileNotFoundError(
            f"{path} not found: put the torchvision state_dict of {name} there (IDUNNO_WEIGHTS_DIR selects the "
            "directory) or install weights with alexnet_resnet.set_weights(); set IDUNNO_ALLOW_RANDOM_WEIGHTS=1 "
            "to classify with seeded synthetic weights instead")
    seed = int(os.environ.get("IDUNNO_WEIGHTS_SEED", "0"))
    variant = os.environ.get("IDUNNO_WEIGHTS_VARIANT", "default")
    sys.stderr.write(f"[alexnet_resnet] {path} not found and IDUNNO_ALLOW_RANDOM_WEIGHTS=1: "
                     f"using seed-{seed} synthetic '{variant}' {name} weights\n")
    return synth.make_state_dict(name, seed, variant)


def set_weights(modelname: str, state_dict, device: int | None = None):
    """Install a state_dict for a model (tests, or a host that manages checkpoints itself)."""
    name = "alexnet" if modelname == "alexnet" else "resnet18"
    dev = int(os.environ.get("LOCAL_RANK", "0")) if device is None else device
    old = _engines.pop(name, None)
    if old is not None:
        old.close()
    _engines[name] = Engine(name, state_dict, device=dev, max_batch=TENSOR_BATCH)


def _engine(modelname: str) -> Engine:
    name = "alexnet" if modelname == "alexnet" else "resnet18"   # :17-22: anything else = ResNet18
    if name not in _engines:
        set_weights(name, _state_dict(name))
    return _engines[name]


def deeplearning(filename, modelname, start_image, end_image):
    """-> (list[(image_name, category_name, top1_probability)], elapsed_seconds), one tuple per
    existing ``./<filename>/test_<i>.JPEG`` with i in [start_image, end_image] (inclusive; str or
    int), ascending.  A missing file is skipped; a CUDA failure raises (the caller's
    ``except Exception: print(e)`` at mp4_machinelearning.py:615 then drops the shard and the
    worker-loss path reassigns it)."""
    start_image = str(start_image)
    end_image = str(end_image)
    time_start = time.time()
    eng = _engine(modelname)
    categories = staging.categories()
    indices, idx, prob = staging.classify_range(eng, filename, start_image, end_image)
    inference_result = []
    for i, k, p in zip(indices, idx, prob):
        inference_result.append(("test_" + str(i) + ".JPEG", categories[int(k)], float(p)))
    time_end = time.time()
    return inference_result, (time_end - time_start)
