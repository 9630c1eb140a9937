"""CPU: host-side logic (sharding, packing, weights blob) and the C-ABI library surface."""
import ctypes
import os
import re

import numpy as np
import pytest

from idunno_b200 import _lib, pack_images, sharding, weights
from idunno_b200.synth import make_state_dict, synth_image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_split_range_matches_reference_examples():
    # SURVEY.md 3.1 step 5 / mp4_machinelearning.py:523-536
    assert [(s, e) for _, s, e in sharding.split_range(0, 399, list(range(5)))] == \
        [(0, 79), (80, 159), (160, 239), (240, 319), (320, 399)]
    assert [(s, e) for _, s, e in sharding.split_range(0, 399, list(range(8)))] == \
        [(50 * i, 50 * i + 49) for i in range(8)]
    assert [(s, e) for _, s, e in sharding.split_range(0, 399, list(range(3)))] == \
        [(0, 132), (133, 266), (267, 399)]
    assert [(s, e) for _, s, e in sharding.split_range(0, 79999, list(range(8)))] == \
        [(10000 * i, 10000 * i + 9999) for i in range(8)]


@pytest.mark.parametrize("start,end,k", [(0, 0, 1), (5, 11, 3), (0, 2, 5), (7, 406, 8), (0, 1000, 7)])
def test_split_range_covers_range_exactly_once(start, end, k):
    shards = sharding.split_range(start, end, list(range(k)))
    seen = []
    for _, s, e in shards:
        seen.extend(range(s, e + 1))
    assert seen == list(range(start, end + 1))


def test_split_queries_and_fair_share():
    assert sharding.split_queries(0, 399) == [(0, 399)]
    assert sharding.split_queries(0, 3999)[:2] == [(0, 399), (400, 799)]
    assert len(sharding.split_queries(0, 3999)) == 10
    # the reference's constant 100/100 average times give 5 workers per query
    assert sharding.fair_share("alexnet", {"alexnet": 100, "resnet18": 100}, 10) == 5
    assert sharding.fair_share("resnet18", {"alexnet": 100, "resnet18": 100}, 10) == 5
    # AlexNet queries twice as fast -> ResNet18 gets the larger share, clamped to the box
    assert sharding.fair_share("alexnet", {"alexnet": 1.0, "resnet18": 2.0}, 8) == 3
    assert sharding.fair_share("resnet18", {"alexnet": 1.0, "resnet18": 2.0}, 8) == 7
    assert sharding.fair_share("resnet18", {"alexnet": 1.0, "resnet18": 0}, 8) == 5


def test_reassign_moves_whole_subrange_to_ring_successor():
    shards = sharding.split_range(0, 399, list(range(4)))
    out = sharding.reassign(shards, dead=2, alive={0, 1, 3}, world=4)
    assert out == [(0, 0, 99), (1, 100, 199), (3, 200, 299), (3, 300, 399)]
    # successor of the last rank is the coordinator (0), which is skipped (:720) -> rank 1
    out = sharding.reassign(shards, dead=3, alive={0, 1, 2}, world=4)
    assert out[3] == (1, 300, 399)
    # only the coordinator left
    out = sharding.reassign([(1, 0, 9)], dead=1, alive={0}, world=2)
    assert out == [(0, 0, 9)]


def test_pack_images_ragged_and_aligned():
    imgs = [synth_image(i, mixed=True) for i in range(5)]
    packed, offsets, hw = pack_images(imgs)
    assert all(int(o) % 16 == 0 for o in offsets)
    for im, o, (h, w) in zip(imgs, offsets, hw):
        assert (h, w) == im.shape[:2]
        assert np.array_equal(packed[int(o):int(o) + im.size].reshape(im.shape), im)
    p0, o0, hw0 = pack_images([])
    assert o0.shape == (0,) and hw0.shape == (0, 2)
    with pytest.raises(ValueError):
        pack_images([np.zeros((4, 4), dtype=np.uint8)])


def test_weight_blob_sizes_match_library():
    lib = _lib.load_library()
    for model, mid in (("alexnet", 0), ("resnet18", 1)):
        blob = weights.pack_state_dict(make_state_dict(model, 0))
        assert blob.dtype == np.float32 and blob.size == lib.idn_weight_count(mid)
    assert lib.idn_weight_count(7) == -1


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "idunno_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(idn_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in _lib.load_library().idn_version()


def test_rows_sent_to_the_gpu_are_exactly_the_rows_the_crop_reads():
    """The host-buffer path copies only source rows [first, end) of equally sized images (idn_debug_needed_rows, host
    arithmetic, no GPU).  Against the oracle's restatement of Pillow's coefficient bounds: the union of the rows the 224
    crop rows read must be exactly that range -- one row less would classify stale pixels -- over upscaled, downscaled,
    square, portrait and landscape sources, including the smallest and some odd sizes."""
    from oracle import pil_resize
    lib = _lib.load_library()
    sizes = [(224, 224), (225, 224), (1, 1), (2, 3), (3, 700), (37, 41), (255, 255), (256, 256), (257, 300), (300, 260),
             (231, 517), (517, 231), (375, 500), (500, 375), (480, 640), (900, 640), (640, 900), (1000, 225), (1200, 1600)]
    rng = np.random.default_rng(5)
    sizes += [(int(a), int(b)) for a, b in rng.integers(1, 1500, (40, 2))]
    for h, w in sizes:
        nh, nw = pil_resize.resized_size(h, w)
        top, _ = pil_resize.crop_offsets(nh, nw)
        if nh == h:
            lo, hi = top, top + 224                               # Pillow skips the vertical pass
        else:
            xmin, xcnt, _ = pil_resize.precompute_coeffs(h, nh)
            lo = int(min(xmin[top + y] for y in range(224)))
            hi = int(max(xmin[top + y] + xcnt[top + y] for y in range(224)))
        a, b = ctypes.c_int(), ctypes.c_int()
        assert lib.idn_debug_needed_rows(h, w, ctypes.byref(a), ctypes.byref(b)) == 0
        assert (a.value, b.value) == (lo, hi), (h, w, a.value, b.value, lo, hi)
        assert 0 <= a.value < b.value <= h


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _lib.load_library()
    blob = np.zeros(lib.idn_weight_count(1), dtype=np.float32)
    h = ctypes.c_void_p()
    rc = lib.idn_create(0, 1, blob.ctypes.data, blob.nbytes, 8, ctypes.byref(h))
    assert rc != 0 and not h.value
    assert b"no CUDA device" in lib.idn_last_error() or b"failed" in lib.idn_last_error()
    out = np.zeros((1, 3, 224, 224), dtype=np.uint16)
    packed, offsets, hw = pack_images([synth_image(0)])
    assert lib.idn_preprocess_u8hwc(0, packed.ctypes.data, offsets.ctypes.data, hw.ctypes.data, 1,
                                    out.ctypes.data) != 0
    # bad arguments are rejected before any device work
    assert lib.idn_create(0, 5, blob.ctypes.data, blob.nbytes, 8, ctypes.byref(h)) == -1
    assert lib.idn_create(0, 1, blob.ctypes.data, 16, 8, ctypes.byref(h)) == -4


def test_stage_chunk_layout_and_reference_file_semantics(tmp_path, monkeypatch):
    """Decode threads -> ragged buffer: 16-byte aligned starts, exact pixels, missing files skipped,
    non-RGB files converted and rewritten in place (alexnet_resnet.py:46-54)."""
    from concurrent.futures import ThreadPoolExecutor
    from PIL import Image
    from idunno_b200 import staging
    from idunno_b200.synth import synth_image
    monkeypatch.chdir(tmp_path)
    os.makedirs("alexnet")
    sizes = {3: (224, 224), 4: (37, 53), 6: (300, 200), 7: (64, 96)}          # index 5 is missing
    want = {}
    for i, hw in sizes.items():
        a = synth_image(i, hw)
        if i == 6:
            a = a[:, :, 0]                                                     # grayscale file
            Image.fromarray(a, mode="L").save(f"alexnet/test_{i}.JPEG", format="PNG")
            want[i] = np.repeat(a[:, :, None], 3, axis=2)
        else:
            Image.fromarray(a).save(f"alexnet/test_{i}.JPEG", format="PNG")
            want[i] = a
    files = staging.discover("alexnet", "3", 8)
    assert [f[0] for f in files] == [3, 4, 6, 7]
    with ThreadPoolExecutor(max_workers=3) as pool:
        c = staging.stage_chunk(files, lambda n: np.zeros(n, dtype=np.uint8), pool)
    assert list(c.indices) == [3, 4, 6, 7]
    assert all(int(o) % staging.ALIGN == 0 for o in c.offsets)
    for k, i in enumerate(c.indices):
        h, w = c.hw[k]
        assert (h, w) == sizes[int(i)]
        o = int(c.offsets[k])
        assert np.array_equal(c.buf[o:o + h * w * 3].reshape(h, w, 3), want[int(i)])
    assert Image.open("alexnet/test_6.JPEG").mode == "RGB"                    # rewritten like the reference
    assert staging.discover("alexnet", 100, 105) == [] and staging.discover("nosuchfolder", 0, 5) == []


def test_missing_checkpoint_is_an_error_unless_opted_in(tmp_path, monkeypatch):
    """The reference loads pretrained checkpoints (alexnet_resnet.py:18,21) and fails when it cannot; the
    drop-in must not classify with synthetic weights just because <model>.pth is missing."""
    import alexnet_resnet as shim
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("IDUNNO_WEIGHTS_DIR", str(tmp_path))
    monkeypatch.delenv("IDUNNO_ALLOW_RANDOM_WEIGHTS", raising=False)
    with pytest.raises(FileNotFoundError, match="IDUNNO_ALLOW_RANDOM_WEIGHTS"):
        shim._state_dict("resnet18")
    monkeypatch.setenv("IDUNNO_ALLOW_RANDOM_WEIGHTS", "1")
    sd = shim._state_dict("resnet18")
    assert "fc.weight" in sd
    import torch
    torch.save(make_state_dict("alexnet", 3), tmp_path / "alexnet.pth")
    monkeypatch.delenv("IDUNNO_ALLOW_RANDOM_WEIGHTS")
    assert torch.equal(shim._state_dict("alexnet")["classifier.6.bias"], make_state_dict("alexnet", 3)["classifier.6.bias"])


def test_classify_range_drains_tickets_when_staging_fails(tmp_path, monkeypatch):
    """A file that cannot be decoded for batch k+1 while batch k is in flight: the exception must leave
    classify_range, and the engine's submission slots must have been released (Engine.abort)."""
    from PIL import Image
    from idunno_b200 import staging
    monkeypatch.chdir(tmp_path)
    os.makedirs("alexnet")
    for i in range(5):
        Image.fromarray(synth_image(i, (32, 40))).save(f"alexnet/test_{i}.JPEG", format="PNG")
    with open("alexnet/test_5.JPEG", "wb") as f:
        f.write(b"this is not an image")

    class FakeEngine:
        max_batch, device = 2, 0

        def __init__(self):
            self.in_flight, self.aborted, self._stage_ring = 0, 0, [None] * 4

        def submit_packed(self, offsets, hw, ptr):
            assert self.in_flight < 3
            self.in_flight += 1
            return (self.in_flight, len(offsets), False)

        def collect(self, ticket):
            self.in_flight -= 1
            return np.zeros(ticket[1], np.int32), np.ones(ticket[1], np.float32), None

        def abort(self):
            self.aborted += 1
            self.in_flight = 0

    class HostBuf:                                      # stands in for the page-locked buffer (no GPU here)
        def __init__(self, n, device=0):
            self.nbytes, self.array, self.ptr = n, np.zeros(n, np.uint8), 1

        def close(self):
            pass

    import idunno_b200.engine as engine_mod
    monkeypatch.setattr(engine_mod, "PinnedBuffer", HostBuf)
    eng = FakeEngine()
    with pytest.raises(Exception):
        staging.classify_range(eng, "alexnet", 0, 5, threads=2)
    assert eng.aborted == 1 and eng.in_flight == 0
    os.remove("alexnet/test_5.JPEG")
    indices, idx, prob = staging.classify_range(eng, "alexnet", 0, 5, threads=2)
    assert list(indices) == [0, 1, 2, 3, 4] and eng.in_flight == 0 and eng.aborted == 1
