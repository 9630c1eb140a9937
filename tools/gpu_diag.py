"""Bring-up diagnostics for the tcgen05 kernel: structured GEMM/conv cases whose outputs reveal
operand-layout mistakes; arrays are dumped to gpurun_out/diag_*.npz for offline inspection."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from idunno_b200 import conv2d_nchw  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def bf(a):
    return torch.from_numpy(np.asarray(a, np.float32)).to(torch.bfloat16).float().numpy()


def run(name, x, w, stride=1, pad=0, **kw):
    want = F.conv2d(torch.from_numpy(x), torch.from_numpy(w), None, stride=stride, padding=pad).numpy()
    res = {"x": x, "w": w, "want": want}
    for impl in ("simt", "tc"):
        try:
            got = conv2d_nchw(x, w, stride=stride, pad=pad, impl=impl, **kw)
            err = float(np.abs(got - want).max())
            print(f"{name:28s} {impl:4s} max abs err {err:.5f}  (ref max {np.abs(want).max():.3f})", flush=True)
            res["got_" + impl] = got
        except Exception as e:  # noqa: BLE001
            print(f"{name:28s} {impl:4s} FAILED: {e}", flush=True)
    np.savez_compressed(os.path.join(OUT, f"diag_{name}.npz"), **res)


rng = np.random.default_rng(0)
# a. FC-shaped GEMM, one k-block, identity weights: output must equal the input exactly
x = bf(rng.standard_normal((128, 64, 1, 1)))
run("gemm_k64_identity", x, np.eye(64, dtype=np.float32).reshape(64, 64, 1, 1))
# b. two k-blocks, weights [I | 2I]
w = np.concatenate([np.eye(64), 2 * np.eye(64)], axis=1).astype(np.float32).reshape(64, 128, 1, 1)
run("gemm_k128_i_2i", bf(rng.standard_normal((200, 128, 1, 1))), w)
# c. N = 128 / 256 tiles, random weights, K long enough to wrap the smem ring several times
for cout in (128, 256):
    run(f"gemm_k1024_n{cout}", bf(rng.standard_normal((300, 1024, 1, 1))),
        bf(rng.standard_normal((cout, 1024, 1, 1)) / 32))
# d. 3x3 convolution with a single non-zero tap at a time (checks the tap row offsets)
xc = bf(rng.standard_normal((2, 64, 8, 8)))
for (r, s) in ((0, 0), (1, 1), (2, 1), (1, 2)):
    wc = np.zeros((64, 64, 3, 3), np.float32)
    wc[:, :, r, s] = np.eye(64)
    run(f"conv3x3_tap{r}{s}", xc, wc, pad=1)
# e. stride-2 taps through the parity planes
for (r, s) in ((0, 0), (1, 1), (2, 2)):
    wc = np.zeros((64, 64, 3, 3), np.float32)
    wc[:, :, r, s] = np.eye(64)
    run(f"conv3x3s2_tap{r}{s}", xc, wc, stride=2, pad=1)
# f. stems with one non-zero tap
xs = bf(rng.standard_normal((1, 3, 224, 224)))
for (R, st, pd) in ((7, 2, 3),):
    for (r, s) in ((0, 0), (R // 2, R // 2), (R - 1, R - 1)):
        ws = np.zeros((64, 3, R, R), np.float32)
        ws[0, 0, r, s] = 1.0; ws[1, 1, r, s] = 1.0; ws[2, 2, r, s] = 1.0
        run(f"stem{R}_tap{r}_{s}", xs, ws, stride=st, pad=pd)
print("diag done")
