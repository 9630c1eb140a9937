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


# ---------------------------------------------------------------------------------------------
# Staging pipeline: decode on host threads -> pinned ragged buffer -> asynchronous tensor batches
# ---------------------------------------------------------------------------------------------
ALIGN = 16        # image starts are 16-byte aligned so the preprocess kernel's 128-bit loads line up


def discover(filename: str, start_image, end_image):
    """The reference's discovery loop (alexnet_resnet.py:24,46-49,90) without the decode:
    [(index, path)] for every existing ./<filename>/test_<i>.JPEG, ascending."""
    found = []
    current = int(start_image)
    for folder in glob.glob("./" + str(filename)):
        while current <= int(end_image):
            for f in glob.glob(folder + "/test_" + str(current) + ".JPEG"):
                found.append((current, f))
            current += 1
    return found


def _open_rgb(path: str):
    """PIL open + the reference's RGB rewrite (alexnet_resnet.py:50-54).  Image.open only reads the
    header, so the size is known before the pixels are decoded."""
    from PIL import Image
    im = Image.open(path)
    if im.mode != "RGB":
        im = im.convert("RGB")
        os.remove(path)
        im.save(path)
    return im


class _Chunk:
    __slots__ = ("indices", "offsets", "hw", "nbytes", "buf")


def stage_chunk(files, buf_for, pool):
    """Decodes ``files`` = [(index, path)] into one ragged uint8 HWC buffer.  ``buf_for(nbytes)``
    returns a writable uint8 array of at least that size (pinned in the product); ``pool`` is a
    concurrent.futures executor -- Pillow releases the GIL while it decodes, so the threads scale."""
    ims = list(pool.map(lambda f: _open_rgb(f[1]), files))
    c = _Chunk()
    c.indices = np.asarray([f[0] for f in files], dtype=np.int64)
    c.hw = np.asarray([[im.size[1], im.size[0]] for im in ims], dtype=np.int32).reshape(-1, 2)
    sizes = c.hw[:, 0].astype(np.int64) * c.hw[:, 1] * 3
    padded = (sizes + ALIGN - 1) // ALIGN * ALIGN
    c.offsets = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.int64)
    c.nbytes = int(padded.sum())
    c.buf = buf_for(max(c.nbytes, ALIGN))
    arr = c.buf if isinstance(c.buf, np.ndarray) else c.buf.array

    def decode(k):
        h, w = int(c.hw[k, 0]), int(c.hw[k, 1])
        o = int(c.offsets[k])
        np.copyto(arr[o:o + h * w * 3].reshape(h, w, 3), np.asarray(ims[k], dtype=np.uint8))
    list(pool.map(decode, range(len(ims))))
    return c


def classify_range(eng, filename: str, start_image, end_image, threads: int = 8, device: int | None = None):
    """Folder range -> (indices, top1_idx, top1_prob), tensor batches of ``eng.max_batch`` files.
    While the GPU works on batch k (and the DMA engine copies it), the host threads decode batch
    k+1 into the next pinned buffer; four buffers rotate (three submissions in flight + one being
    filled, kept on the engine between calls).  Missing files are skipped exactly like the reference does.
    The page-locked ring is allocated in the context of the engine's own GPU.  If staging or a collect
    raises (a corrupt file, a file deleted between discovery and open), the submissions already in flight
    are drained before the exception leaves, so the engine stays usable for the next shard."""
    from concurrent.futures import ThreadPoolExecutor
    from .engine import MAX_IN_FLIGHT, PinnedBuffer
    files = discover(filename, start_image, end_image)
    n = len(files)
    indices = np.asarray([f[0] for f in files], dtype=np.int64)
    idx = np.zeros(n, dtype=np.int32)
    prob = np.zeros(n, dtype=np.float32)
    if n == 0:
        return indices, idx, prob
    if device is None:
        device = getattr(eng, "device", 0)
    if not hasattr(eng, "_stage_ring"):               # page-locked memory is slow to allocate: keep it
        eng._stage_ring = [None] * (MAX_IN_FLIGHT + 1)
    ring = eng._stage_ring

    def buf_for(slot):
        def get(nbytes):
            if ring[slot] is None or ring[slot].nbytes < nbytes:
                if ring[slot] is not None:
                    ring[slot].close()
                ring[slot] = PinnedBuffer(max(nbytes, 1 << 20) * 5 // 4, device)
            return ring[slot]
        return get

    B = eng.max_batch
    chunks = [files[i:i + B] for i in range(0, n, B)]
    pending = []                                      # (ticket, first position, chunk)
    try:
        with ThreadPoolExecutor(max_workers=threads) as pool:
            for k, chunk_files in enumerate(chunks):
                # this buffer last held chunk k-4, which was collected before chunk k-1 was submitted
                c = stage_chunk(chunk_files, buf_for(k % (MAX_IN_FLIGHT + 1)), pool)
                if len(pending) == MAX_IN_FLIGHT:
                    t, at, _c = pending.pop(0)
                    i, p, _ = eng.collect(t)
                    idx[at:at + len(i)], prob[at:at + len(i)] = i, p
                pending.append((eng.submit_packed(c.offsets, c.hw, c.buf.ptr), k * B, c))
            while pending:
                t, at, _c = pending.pop(0)
                i, p, _ = eng.collect(t)
                idx[at:at + len(i)], prob[at:at + len(i)] = i, p
    except BaseException:
        # tickets nobody will collect would keep their slots busy for ever ("three submissions are already
        # in flight" on every later call); wait for the GPU and free them, then let the error out
        try:
            eng.abort()
        except Exception:
            pass
        raise
    return indices, idx, prob
