"""Experiment (CPU): how diverse / well-conditioned is a candidate weight set on the 400-image
BASELINE query?  Prints distinct classes, margin statistics and the bf16-contract error."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))

from idunno_b200.synth import make_state_dict, synth_image  # noqa: E402
from oracle import ref_path  # noqa: E402


def main():
    variant = sys.argv[1] if len(sys.argv) > 1 else "diverse"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 400
    t0 = time.time()
    imgs = [synth_image(i, mixed=True) for i in range(n)]
    x = torch.from_numpy(np.stack([ref_path.preprocess_library(im) for im in imgs]))
    print(f"images + preprocess: {time.time() - t0:.1f}s")
    for model in ("resnet18", "alexnet"):
        t0 = time.time()
        sd = make_state_dict(model, 0, variant)
        t1 = time.time()
        logits = torch.cat([ref_path.forward(model, sd, x[i:i + 50]) for i in range(0, n, 50)]).numpy()
        contract = torch.cat([ref_path.forward_bf16_contract(model, sd, x[i:i + 50]) for i in range(0, n, 50)]).numpy()
        srt = np.sort(logits, axis=1)
        margin = srt[:, -1] - srt[:, -2]
        idx = logits.argmax(1)
        err = np.abs(contract - logits).max()
        err_img = np.abs(contract - logits).max(axis=1)
        rng = np.abs(logits).max()
        flips = int((contract.argmax(1) != idx).sum())
        p = torch.softmax(torch.from_numpy(logits), 1).max(1).values.numpy()
        print(f"{model}/{variant}: weights {t1 - t0:.1f}s fwd {time.time() - t1:.1f}s | classes {len(set(idx.tolist()))} "
              f"| |logit|max {rng:.3f} std {logits.std():.3f} | margin median {np.median(margin):.4f} min {margin.min():.5f} "
              f"| contract err max {err:.5f} ({err / rng:.3%} of range) median-per-image {np.median(err_img):.5f} "
              f"| median margin / err = {np.median(margin) / err:.1f} | safe(margin>2err) {int((margin > 2 * err).sum())}/{n} "
              f"| contract flips {flips} | top prob median {np.median(p):.3f}")


if __name__ == "__main__":
    main()
