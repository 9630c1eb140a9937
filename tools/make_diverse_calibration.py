"""Writes idunno_b200/synth_calib/<model>_diverse_seed0.npz: the calibrated per-channel statistics
and the prototype classifier of the "diverse" weight variant (idunno_b200/synth.py).  CPU, about a
minute per model.  Run once in the build container; the files are committed so that every machine
builds bit-identical weights."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))

from idunno_b200 import synth  # noqa: E402

for model in ("resnet18", "alexnet"):
    calib = synth.compute_diverse_calibration(model, 0)
    path = synth._calib_path(model, 0)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **calib)
    print(model, path, os.path.getsize(path), "bytes")
