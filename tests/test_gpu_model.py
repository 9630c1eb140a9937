"""GPU: whole hot path through the C ABI against the oracle and the reference-generated golden
fixtures.

Stated tolerance.  The CUDA path computes with bf16 operands and fp32 accumulation; the oracle is
fp32.  What that costs depends on the weights (random-init networks are dominated by a common mode,
so the classifier amplifies rounding noise: 0.25 % of the logit range on default-init weights,
3-7 % on the "stress" weights whose classifier is centred on that common mode).  The tolerance is
therefore stated per weight set and measured, not guessed:

    tau = max(TOL_FLOOR_FRAC * max|logit|,  TOL_FACTOR * max|bf16_contract - fp32|)

where ``bf16_contract`` is ``oracle.ref_path.forward_bf16_contract``: the same network evaluated on
the CPU with every weight and stored activation rounded to bf16 (a kernel-independent statement of
what bf16 costs).  Logits must satisfy |gpu - fp32| <= tau; top-1 must be identical wherever the
oracle's top-1/top-2 margin exceeds twice the measured error; every near-tie flip is printed."""
import os

import numpy as np
import pytest
import torch

from idunno_b200 import Engine, pack_images
from idunno_b200.synth import make_state_dict, noise_images, synth_image
from oracle import pil_resize, ref_path

pytestmark = pytest.mark.gpu
TOL_FLOOR_FRAC = 0.005
TOL_FACTOR = 2.0
PROB_TOL_FACTOR = 2.5      # |dprob| <= factor * tau * prob  (d softmax / d logit <= prob)


def _oracle(model, sd, imgs):
    want = ref_path.infer_arrays(model, sd, imgs)
    x = torch.from_numpy(want["pre"])
    want["contract"] = ref_path.forward_bf16_contract(model, sd, x).numpy() if len(imgs) else want["logits"]
    return want


def _compare(name, got_idx, got_prob, got_logits, want):
    rng = float(np.abs(want["logits"]).max())
    err = float(np.abs(got_logits - want["logits"]).max())
    proxy = float(np.abs(want["contract"] - want["logits"]).max())
    vs_contract = float(np.abs(got_logits - want["contract"]).max())
    tau = max(TOL_FLOOR_FRAC * rng, TOL_FACTOR * proxy)
    srt = np.sort(want["logits"], axis=1)
    margin = srt[:, -1] - srt[:, -2]
    safe = margin > 2 * err
    agree = got_idx == want["idx"]
    print(f"\n[{name}] max|gpu-fp32|={err:.5f} ({err / rng:.3%} of range {rng:.3f}; rel to |logit|max) "
          f"bf16-contract proxy={proxy:.5f} tau={tau:.5f} max|gpu-contract|={vs_contract:.5f}; "
          f"top-1 agree all {int(agree.sum())}/{len(agree)}, margin>2*err {int((agree & safe).sum())}/{int(safe.sum())}")
    for i in np.nonzero(~agree)[0]:
        print(f"   near-tie image {i}: oracle class {want['idx'][i]} (margin {margin[i]:.5f}) vs gpu {got_idx[i]}"
              f" (gpu logits {got_logits[i, want['idx'][i]]:.5f} / {got_logits[i, got_idx[i]]:.5f})")
    assert err <= tau, f"{name}: |gpu - fp32| = {err} exceeds the stated tolerance {tau}"
    assert vs_contract <= tau, f"{name}: |gpu - bf16 contract| = {vs_contract} exceeds {tau}"
    assert bool(np.all(agree[safe])), f"{name}: top-1 differs outside the tolerance band"
    assert np.all(np.abs(got_prob[agree] - want["prob"][agree])
                  <= PROB_TOL_FACTOR * tau * np.maximum(want["prob"][agree], got_prob[agree]) + 1e-6), \
        f"{name}: top-1 probability"
    return err, tau


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("case", ["alexnet_default", "alexnet_stress", "resnet18_default", "resnet18_stress"])
def test_against_reference_fixture(built_library, golden_dir, case, impl):
    """Logits / classes / probabilities recorded from the UNMODIFIED reference (oracle/make_golden.py)."""
    g = np.load(os.path.join(golden_dir, case + ".npz"))
    model, variant = str(g["model"]), str(g["variant"])
    skipped = set(int(s) for s in g["skipped"])
    idxs = [i for i in range(int(g["first"]), int(g["last"]) + 1) if i not in skipped]
    imgs = [synth_image(i, mixed=bool(g["mixed"])) for i in idxs]
    sd = make_state_dict(model, 0, variant)
    with Engine(model, sd, max_batch=16, impl=impl) as eng:
        idx, prob, logits = eng.infer(imgs, want_logits=True)
    x = torch.from_numpy(np.stack([pil_resize.preprocess(im) for im in imgs]))
    want = dict(logits=g["logits"], idx=g["idx"], prob=g["prob"].astype(np.float32),
                contract=ref_path.forward_bf16_contract(model, sd, x).numpy())
    _compare(f"{case}/{impl}", idx, prob, logits, want)


@pytest.mark.parametrize("model", ["alexnet", "resnet18"])
def test_layer_by_layer_against_oracle(built_library, model):
    """Every plan step's activation, read back right after the step ran (idn_debug_stop_after),
    against the fp32 oracle's tensor at the same point.  Bound: the error a chain of bf16 layers
    accumulates (about 0.4 % of the RMS per layer, measured 3 % at the deepest layer)."""
    n = 3
    imgs = [synth_image(40 + i, mixed=True) for i in range(n)]
    sd = make_state_dict(model, 0, "stress")
    x = torch.from_numpy(np.stack([pil_resize.preprocess(im) for im in imgs]))
    taps = ref_path.forward_taps(model, sd, x)
    report = []
    with Engine(model, sd, max_batch=4) as eng:
        for si, (name, chw) in enumerate(eng.steps()):
            if name not in taps or chw[0] == 0:
                continue
            eng.stop_after(si)
            eng.infer_packed(*pack_images(imgs))
            got = eng.read_activation(si, n)
            want = taps[name].numpy()
            assert got.shape == want.shape, (name, got.shape, want.shape)
            rms = float(np.sqrt((want ** 2).mean())) + 1e-9
            report.append((name, float(np.abs(got - want).max()) / rms,
                           float(np.sqrt(((got - want) ** 2).mean())) / rms))
        eng.stop_after(-1)
    print("\n" + "\n".join(f"   {model} {nm:28s} max err / rms = {e:.4f}   rms err / rms = {r:.5f}"
                           for nm, e, r in report))
    assert len(report) >= 8, "too few comparable steps"
    for nm, e, r in report:
        if nm == "classifier.6":      # the centred "stress" classifier: bounded by the logit tolerance tests
            continue
        assert r < 0.02, f"{model} {nm}: rms error {r}"
        assert e < 0.25, f"{model} {nm}: max error {e}"


@pytest.mark.parametrize("model", ["alexnet", "resnet18"])
def test_chunking_order_and_duplicates_are_bit_exact(built_library, model):
    """Size-independent properties: per-image results do not depend on the tensor batch an image
    falls in, on its position, or on its neighbours (empty and ragged inputs included)."""
    sd = make_state_dict(model, 0, "stress")
    imgs = [synth_image(i, mixed=(i % 3 == 0)) for i in range(23)]
    with Engine(model, sd, max_batch=23) as big, Engine(model, sd, max_batch=5) as small:
        i1, p1, l1 = big.infer(imgs, want_logits=True)
        i2, p2, l2 = small.infer(imgs, want_logits=True)            # 5 chunks, last one of 3
        assert np.array_equal(i1, i2) and np.array_equal(p1, p2) and np.array_equal(l1, l2)
        perm = np.random.default_rng(0).permutation(len(imgs))
        i3, p3, l3 = big.infer([imgs[j] for j in perm], want_logits=True)
        assert np.array_equal(i3, i1[perm]) and np.array_equal(l3, l1[perm])
        i4, _, l4 = big.infer([imgs[4]] * 7, want_logits=True)
        assert np.array_equal(l4, np.repeat(l1[4:5], 7, axis=0))
        i5, p5, _ = big.infer([])
        assert i5.shape == (0,) and p5.shape == (0,)
        # softmax consistency: prob == 1 / sum exp(l - max), argmax == first maximum
        ref_p = 1.0 / np.exp(l1 - l1.max(axis=1, keepdims=True)).sum(axis=1)
        assert np.allclose(p1, ref_p, rtol=1e-4) and np.array_equal(i1, l1.argmax(axis=1))


def test_tc_and_simt_backends_agree(built_library):
    """Two independent kernels over the same layouts: they differ only by accumulation order, which
    flips individual bf16 roundings; the difference must stay inside the bf16 tolerance."""
    sd = make_state_dict("resnet18", 0, "stress")
    imgs = [synth_image(70 + i) for i in range(6)]
    want = _oracle("resnet18", sd, imgs)
    with Engine("resnet18", sd, max_batch=8) as eng:
        _, _, a = eng.infer(imgs, want_logits=True)
        eng.set_impl("simt")
        _, _, b = eng.infer(imgs, want_logits=True)
    tau = TOL_FACTOR * float(np.abs(want["contract"] - want["logits"]).max())
    assert float(np.abs(a - b).max()) <= tau


@pytest.mark.parametrize("model", ["alexnet", "resnet18"])
def test_full_query_of_400_is_consistent_with_small_batches(built_library, model):
    """BASELINE size (one 400-image query in one tensor batch, wide N tiles, every CTA busy) against
    the same images pushed through a 16-image engine (narrow tiles): the per-image arithmetic is the
    same, so the results must be bit-identical; the structured tail is also checked against the oracle."""
    sd = make_state_dict(model, 0, "default")
    imgs = list(noise_images(380, seed=11)) + [synth_image(i, mixed=True) for i in range(20)]
    with Engine(model, sd, max_batch=400) as big:
        i1, p1, l1 = big.infer(imgs, want_logits=True)
    with Engine(model, sd, max_batch=16) as small:
        i2, p2, l2 = small.infer(imgs, want_logits=True)
    d = float(np.abs(l1 - l2).max())
    print(f"\n[{model}] batch-400 vs batch-16 engines: max|dlogit| = {d:.3e}, bit-exact = {np.array_equal(l1, l2)}")
    assert np.array_equal(l1, l2) and np.array_equal(i1, i2) and np.array_equal(p1, p2)
    want = _oracle(model, sd, imgs[380:])
    _compare(f"{model}/batch400-tail", i1[380:], p1[380:], l1[380:], want)


@pytest.mark.parametrize("shape", [(224, 224), (300, 260), (231, 517), (900, 640)])
def test_only_the_rows_the_crop_reads_are_copied(built_library, monkeypatch, shape):
    """A batch of equally sized images at a constant stride goes to the GPU as one pitched copy of the source rows
    Resize(256) + CenterCrop(224) reads.  The staging buffers are reused, so the rows that are skipped hold the
    previous batch's pixels: results must still equal an engine that copies whole images (IDN_FULL_H2D=1) bit for
    bit, for upscaled, downscaled, portrait and landscape sources, and fewer bytes must have been queued."""
    h, w = shape
    rng = np.random.default_rng(h * 1000 + w)
    a = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(6)]
    b = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(6)]
    sd = make_state_dict("resnet18", 0, "default")
    with Engine("resnet18", sd, max_batch=6) as eng:
        for _ in range(3):                                   # every slot's staging buffer has held batch a
            eng.infer(a)
        k0 = eng.h2d_bytes
        i1, p1, l1 = eng.infer(b, want_logits=True)
        copied = eng.h2d_bytes - k0
    monkeypatch.setenv("IDN_FULL_H2D", "1")
    with Engine("resnet18", sd, max_batch=6) as eng:
        k0 = eng.h2d_bytes
        i2, p2, l2 = eng.infer(b, want_logits=True)
        full = eng.h2d_bytes - k0
    assert np.array_equal(l1, l2) and np.array_equal(i1, i2) and np.array_equal(p1, p2)
    short = min(h, w)
    print(f"\n[{h}x{w}] copied {copied} of {full} bytes ({copied / full:.3f})")
    assert full >= 6 * h * w * 3
    if h <= w:                                               # the rows are the shorter side: 224/256 of them (+ filter support)
        assert copied < 0.93 * full, (copied, full)
    else:
        assert copied <= full


def test_alexnet_fc8_partials_summed_by_the_softmax_kernel_are_bit_identical(built_library, monkeypatch):
    """FC8 splits K eight ways.  In the product path it only writes its partial tiles and softmax_top1_splitk_kernel adds
    them; with IDN_NO_SPLIT_DEFER=1 (and for activation taps of FC8) the FC kernel's own CTAs wait for one another and
    add them.  Same split order, same fmaf with the bias: logits, classes and probabilities must be the same bits, for
    a ragged batch (37 on an engine of 48: second 128-row half empty) and one of 300 (both row pairs)."""
    sd = make_state_dict("alexnet", 0, "stress")
    imgs = [synth_image(i, mixed=True) for i in range(37)]
    big = list(noise_images(300, seed=3))
    with Engine("alexnet", sd, max_batch=48) as eng:
        assert dict(eng.step_kernels())["softmax+top1"] == "softmax_top1_splitk_kernel"
        i1, p1, l1 = eng.infer(imgs, want_logits=True)
    with Engine("alexnet", sd, max_batch=300) as eng:
        i3, p3, l3 = eng.infer(big, want_logits=True)
    monkeypatch.setenv("IDN_NO_SPLIT_DEFER", "1")
    with Engine("alexnet", sd, max_batch=48) as eng:
        assert dict(eng.step_kernels())["softmax+top1"] == "softmax_top1_kernel"
        i2, p2, l2 = eng.infer(imgs, want_logits=True)
    with Engine("alexnet", sd, max_batch=300) as eng:
        i4, p4, l4 = eng.infer(big, want_logits=True)
    assert np.array_equal(l1, l2) and np.array_equal(i1, i2) and np.array_equal(p1, p2)
    assert np.array_equal(l3, l4) and np.array_equal(i3, i4) and np.array_equal(p3, p4)


def test_alexnet_tensor_batch_of_1200_matches_the_batch_400_engine(built_library):
    """A tensor batch larger than one wave of work items: with max_batch = 1200 FC8's split-K has 160 items for 74
    CTA pairs (a pair runs several rounds and waits, inside a round, for splits its neighbours are still computing)
    and features.0's pooled runs are 219 rows long.  The partial sums are added in split order whoever arrives last,
    so the results must equal the batch-400 engine's bit for bit -- and the call must come back at all."""
    sd = make_state_dict("alexnet", 0, "default")
    imgs = list(noise_images(1190, seed=23)) + [synth_image(i, mixed=True) for i in range(10)]
    with Engine("alexnet", sd, max_batch=1200) as big:
        i1, p1, l1 = big.infer(imgs, want_logits=True)
        i1b, _, l1b = big.infer(imgs, want_logits=True)
    with Engine("alexnet", sd, max_batch=400) as ref:
        i2, p2, l2 = ref.infer(imgs, want_logits=True)           # 400 + 400 + 400
    assert np.array_equal(l1, l1b) and np.array_equal(i1, i1b)
    assert np.array_equal(l1, l2) and np.array_equal(i1, i2) and np.array_equal(p1, p2)


@pytest.mark.parametrize("model", ["alexnet", "resnet18"])
def test_baseline_query_top1_against_reference(built_library, golden_dir, model):
    """BASELINE.json configs[0] at full size: images 0-399 in ONE 400-image tensor batch against the
    logits / classes / probabilities the UNMODIFIED reference produced for them (oracle/make_golden.py,
    "diverse" weights: 400 distinct known classes, median margin > 10x the bf16-contract error).
    This is the top-1 bar of the north star in a form that can fail: the set on which top-1 is provable
    (margin > 2 x measured error) must hold at least 360 of the 400 images, and on it -- and here on all
    400 -- the class must be the reference's."""
    from idunno_b200.synth import diverse_class_of
    g = np.load(os.path.join(golden_dir, f"{model}_diverse400.npz"))
    n = int(g["last"]) - int(g["first"]) + 1
    imgs = [synth_image(i, mixed=True) for i in range(int(g["first"]), int(g["last"]) + 1)]
    sd = make_state_dict(model, 0, "diverse")
    with Engine(model, sd, max_batch=400) as eng:
        idx, prob, logits = eng.infer(imgs, want_logits=True)
    x = torch.from_numpy(np.stack([pil_resize.preprocess(im) for im in imgs]))
    contract = torch.cat([ref_path.forward_bf16_contract(model, sd, x[i:i + 50]) for i in range(0, n, 50)]).numpy()
    want = dict(logits=g["logits"], idx=g["idx"], prob=g["prob"].astype(np.float32), contract=contract)
    err, tau = _compare(f"{model}/diverse400", idx, prob, logits, want)
    srt = np.sort(want["logits"], axis=1)
    margin = srt[:, -1] - srt[:, -2]
    safe = margin > 2 * err
    agree = idx == want["idx"]
    a, b, c = int(agree.sum()), int((agree & safe).sum()), int(safe.sum())
    print(f"[{model}/diverse400] top-1 agree {a}/{n}, margin>2err {b}/{c}; median margin / err = {np.median(margin) / err:.1f}; "
          f"error {err / float(np.abs(want['logits']).max()):.3%} of the logit range")
    assert c > 0.5 * n, "the provable set must not be vacuous"
    assert c >= 360 and b == c
    assert a == n, f"{n - a} top-1 flips against the reference"
    assert np.array_equal(idx, np.asarray([diverse_class_of(model, i) for i in range(n)]))     # the known answer


def test_deeplearning_dropin_matches_oracle(built_library, tmp_path, monkeypatch):
    """The reference-facing call: same signature, same result format (alexnet_resnet.py:12,85-92)."""
    from PIL import Image
    import alexnet_resnet as shim
    monkeypatch.chdir(tmp_path)
    for model in ("alexnet", "resnet18"):
        os.makedirs(model)
        present = (0, 1, 2, 4, 5)                       # test_3 is missing: skipped silently
        for i in present:
            Image.fromarray(synth_image(i, mixed=True)).save(f"{model}/test_{i}.JPEG", format="PNG")
        Image.fromarray(synth_image(6)[:, :, 0]).save(f"{model}/test_6.JPEG", format="PNG")   # mode L
        sd = make_state_dict(model, 0, "default")
        shim.set_weights(model, sd, device=0)
        res, secs = shim.deeplearning(model, model, "0", 6)      # str and int arguments
        cats = ref_path.imagenet_categories()
        tap = {}
        want, _ = ref_path.deeplearning(model, model, 0, "6", sd, cats, tap)
        assert [r[0] for r in res] == [w[0] for w in want] == [f"test_{i}.JPEG" for i in present + (6,)]
        assert isinstance(res[0], tuple) and isinstance(res[0][1], str) and isinstance(res[0][2], float)
        srt = np.sort(tap["logits"], axis=1)
        rng = float(np.abs(tap["logits"]).max())
        safe = (srt[:, -1] - srt[:, -2]) > 4 * TOL_FLOOR_FRAC * rng
        for k in range(len(res)):
            if safe[k]:
                assert res[k][1] == want[k][1], (model, k)
                assert abs(res[k][2] - want[k][2]) <= 0.02 * want[k][2]
        assert int(safe.sum()) >= 1
        assert isinstance(secs, float) and secs > 0
        assert shim.deeplearning(model, model, 100, 105)[0] == []          # nothing found
        assert Image.open(f"{model}/test_6.JPEG").mode == "RGB"            # rewritten in place (:51-54)


def test_submit_collect_overlaps_and_matches_infer(built_library):
    """idn_submit_u8hwc / idn_collect: tensor batches in flight give the same results as the
    one-call form, a fourth submission is refused, and long inputs (which use the same slots inside
    idn_infer_u8hwc) come back in order."""
    import torch
    from idunno_b200 import IdunnoError
    sd = make_state_dict("resnet18", 0, "default")
    a = [synth_image(i, mixed=True) for i in range(9)]
    b = [synth_image(50 + i) for i in range(5)]
    with Engine("resnet18", sd, max_batch=12) as eng:
        ia, pa, la = eng.infer(a, want_logits=True)
        ib, pb, _ = eng.infer(b)
        pk_a, off_a, hw_a = pack_images(a)
        pk_b, off_b, hw_b = pack_images(b)
        ta = eng.submit_packed(off_a, hw_a, pk_a.ctypes.data, want_logits=True)
        tb = eng.submit_packed(off_b, hw_b, pk_b.ctypes.data)
        tc = eng.submit_packed(off_a, hw_a, pk_a.ctypes.data)
        with pytest.raises(IdunnoError):                     # IDN_MAX_IN_FLIGHT = 3
            eng.submit_packed(off_b, hw_b, pk_b.ctypes.data)
        ga = eng.collect(ta)
        gb = eng.collect(tb)
        gc = eng.collect(tc)
        assert np.array_equal(gc[0], ia) and np.array_equal(gc[1], pa)
        assert np.array_equal(ga[0], ia) and np.array_equal(ga[1], pa) and np.array_equal(ga[2], la)
        assert np.array_equal(gb[0], ib) and np.array_equal(gb[1], pb) and gb[2] is None
        # idn_abort: tickets nobody collects are voided, all three slots are free again
        eng.submit_packed(off_a, hw_a, pk_a.ctypes.data)
        eng.submit_packed(off_b, hw_b, pk_b.ctypes.data)
        eng.abort()
        ts = [eng.submit_packed(off_a, hw_a, pk_a.ctypes.data) for _ in range(3)]
        for t in ts:
            assert np.array_equal(eng.collect(t)[0], ia)
        many = a + b + a + b + a + b + a             # 51 images through a 12-image engine: 5 batches
        im, pm, _ = eng.infer(many)
        assert np.array_equal(im, np.concatenate([ia, ib, ia, ib, ia, ib, ia]))
        assert np.array_equal(pm, np.concatenate([pa, pb, pa, pb, pa, pb, pa]))


def test_resnet_stem_pool_fusion_is_bit_identical(built_library, monkeypatch):
    """The product path folds ResNet18's max-pool into the stem's epilogue (partial window maxima with
    red.max into a zeroed buffer); IDN_NO_POOL_FUSION=1 keeps the two kernels.  max is exact and order
    independent, so logits must be identical bit for bit -- also across repeated passes (the pooled
    buffer is re-zeroed on a side stream every pass) and for a ragged batch."""
    sd = make_state_dict("resnet18", 0, "stress")
    imgs = [synth_image(i, mixed=True) for i in range(21)]
    with Engine("resnet18", sd, max_batch=16) as eng:
        i1, p1, l1 = eng.infer(imgs, want_logits=True)          # two tensor batches: 16 + 5
        i2, p2, l2 = eng.infer(imgs[:7], want_logits=True)
        i3, p3, l3 = eng.infer(imgs, want_logits=True)
    monkeypatch.setenv("IDN_NO_POOL_FUSION", "1")
    with Engine("resnet18", sd, max_batch=16) as eng:
        j1, q1, m1 = eng.infer(imgs, want_logits=True)
    assert np.array_equal(l1, m1) and np.array_equal(i1, j1) and np.array_equal(p1, q1)
    assert np.array_equal(l3, l1) and np.array_equal(l2, l1[:7])


def test_alexnet_first_pool_in_the_conv_epilogue_is_bit_identical(built_library, monkeypatch):
    """AlexNet's features.2 max-pool runs in the epilogue of features.0 (conv rows kept in a shared-memory ring by
    the CTA that owns a contiguous run of pooled rows); IDN_NO_POOL_FUSION=1 keeps conv + maxpool3x3s2_kernel.
    max over the same nine bf16 values is exact, so the pooled activation and everything after it must be
    identical bit for bit: 16 images (432 pooled rows over 148 CTAs: runs start and end inside images), a ragged
    second batch, a single image, and repeated passes."""
    sd = make_state_dict("alexnet", 0, "stress")
    imgs = [synth_image(i, mixed=True) for i in range(21)]
    with Engine("alexnet", sd, max_batch=16) as eng:
        names = [n for n, _ in eng.steps()]
        pool_step = names.index("features.2 maxpool")
        assert dict(eng.step_kernels())["features.2 maxpool"].startswith("(folded")
        i1, p1, l1 = eng.infer(imgs, want_logits=True)          # two tensor batches: 16 + 5
        i2, p2, l2 = eng.infer(imgs[:1], want_logits=True)
        i3, p3, l3 = eng.infer(imgs, want_logits=True)
        eng.infer(imgs[:16])
        k0 = eng.kernel_launches
        eng.infer(imgs[:16])
        fused_launches = eng.kernel_launches - k0
        eng.stop_after(pool_step)
        eng.infer(imgs[:16])
        a1 = eng.read_activation(pool_step, 16)
    monkeypatch.setenv("IDN_NO_POOL_FUSION", "1")
    with Engine("alexnet", sd, max_batch=16) as eng:
        assert dict(eng.step_kernels())["features.2 maxpool"] == "maxpool3x3s2_kernel"
        j1, q1, m1 = eng.infer(imgs, want_logits=True)
        eng.infer(imgs[:16])
        k0 = eng.kernel_launches
        eng.infer(imgs[:16])
        plain_launches = eng.kernel_launches - k0
        eng.stop_after(pool_step)
        eng.infer(imgs[:16])
        b1 = eng.read_activation(pool_step, 16)
    assert np.array_equal(a1, b1), f"pooled activation differs in {int((a1 != b1).sum())} of {a1.size} values"
    assert np.array_equal(l1, m1) and np.array_equal(i1, j1) and np.array_equal(p1, q1)
    assert np.array_equal(l3, l1) and np.array_equal(l2, l1[:1])
    assert fused_launches == plain_launches - 2, (fused_launches, plain_launches)   # features.2 and features.12 ride along


def test_alexnet_later_pools_on_the_pair_kernel_are_bit_identical(built_library, monkeypatch):
    """features.12 runs inside features.10 (CTA-pair kernel: both CTAs write their conv rows to the pair's ring in
    global memory, hand over through an mbarrier per item and pool alternate units); features.5 inside features.3 is
    the same code at N = 192 but does not pay and is off by default -- here it is switched on as well.  Against
    IDN_NO_POOL_FUSION=1: pooled activations and logits identical bit for bit, for 37 images on an engine of 48 (runs
    that start inside images, ragged last item), a single image, repeated passes; three launches fewer per pass."""
    sd = make_state_dict("alexnet", 0, "stress")
    imgs = [synth_image(i, mixed=True) for i in range(37)]
    with Engine("alexnet", sd, max_batch=48) as eng:         # the product default: features.12 only
        kernels = dict(eng.step_kernels())
        assert kernels["features.12 maxpool"].startswith("(folded") and kernels["features.5 maxpool"] == "maxpool3x3s2_kernel"
    monkeypatch.setenv("IDN_FUSE_FEATURES5", "1")            # the 192-wide pair kernel's pooled form is exercised too
    with Engine("alexnet", sd, max_batch=48) as eng:
        names = [n for n, _ in eng.steps()]
        kernels = dict(eng.step_kernels())
        assert kernels["features.5 maxpool"].startswith("(folded") and kernels["features.12 maxpool"].startswith("(folded")
        i1, p1, l1 = eng.infer(imgs, want_logits=True)
        i2, p2, l2 = eng.infer(imgs[:1], want_logits=True)
        i3, p3, l3 = eng.infer(imgs, want_logits=True)
        k0 = eng.kernel_launches
        eng.infer(imgs)
        fused_launches = eng.kernel_launches - k0
        taps = {}
        for nm in ("features.5 maxpool", "features.12 maxpool"):
            eng.stop_after(names.index(nm))
            eng.infer(imgs)
            taps[nm] = eng.read_activation(names.index(nm), len(imgs))
    monkeypatch.setenv("IDN_NO_POOL_FUSION", "1")
    with Engine("alexnet", sd, max_batch=48) as eng:
        j1, q1, m1 = eng.infer(imgs, want_logits=True)
        k0 = eng.kernel_launches
        eng.infer(imgs)
        plain_launches = eng.kernel_launches - k0
        for nm in ("features.5 maxpool", "features.12 maxpool"):
            eng.stop_after(names.index(nm))
            eng.infer(imgs)
            b = eng.read_activation(names.index(nm), len(imgs))
            assert np.array_equal(taps[nm], b), f"{nm}: {int((taps[nm] != b).sum())} of {b.size} values differ"
    assert np.array_equal(l1, m1) and np.array_equal(i1, j1) and np.array_equal(p1, q1)
    assert np.array_equal(l3, l1) and np.array_equal(l2, l1[:1])
    assert fused_launches == plain_launches - 3, (fused_launches, plain_launches)


def test_resnet_head_fusion_is_bit_identical(built_library, monkeypatch):
    """avgpool + fp32 FC + softmax/top-1 as ONE kernel (clusters of 4 CTAs, 11 images each) against the three
    separate kernels (IDN_NO_HEAD_FUSION=1): the fused kernel keeps their arithmetic and its order (pixels h, w
    ascending; k ascending with fmaf; the same warp-shuffle softmax), so logits, classes and probabilities
    must be identical bit for bit -- for a full cluster, a ragged tail (23 = 2 x 11 + 1) and a single image."""
    sd = make_state_dict("resnet18", 0, "stress")
    imgs = [synth_image(i, mixed=True) for i in range(23)]
    with Engine("resnet18", sd, max_batch=23) as eng:
        i1, p1, l1 = eng.infer(imgs, want_logits=True)
        i2, p2, _ = eng.infer(imgs)                      # without the logits copy
        i3, p3, l3 = eng.infer(imgs[:1], want_logits=True)
    monkeypatch.setenv("IDN_NO_HEAD_FUSION", "1")
    with Engine("resnet18", sd, max_batch=23) as eng:
        j1, q1, m1 = eng.infer(imgs, want_logits=True)
    assert np.array_equal(l1, m1) and np.array_equal(i1, j1) and np.array_equal(p1, q1)
    assert np.array_equal(i2, i1) and np.array_equal(p2, p1)
    assert np.array_equal(l3, l1[:1]) and np.array_equal(p3, p1[:1])


def test_downsample_riding_in_the_stride2_launch_is_bit_identical(built_library, monkeypatch):
    """ResNet18's three 1x1 / stride-2 downsample convolutions run as a second problem inside the launch of
    their 3x3 / stride-2 partner (same parity planes, same output tiling); IDN_NO_DS_FUSION=1 launches them on
    their own.  The arithmetic is the same, so every logit must be identical -- also layer by layer for the
    downsample outputs themselves (read back after the merged launch), for a batch that uses the CTA-pair
    flavour of the kernel (96 images) and for a small one."""
    sd = make_state_dict("resnet18", 0, "stress")
    imgs = [synth_image(i, mixed=(i % 4 == 0)) for i in range(96)]
    with Engine("resnet18", sd, max_batch=96) as eng:
        names = [n for n, _ in eng.steps()]
        kern = dict(eng.step_kernels())
        assert all("rides in the launch" in kern[f"layer{l}.0.downsample"] for l in (2, 3, 4))
        i1, p1, l1 = eng.infer(imgs, want_logits=True)
        launches0 = eng.kernel_launches                   # second pass: the resample tables are on the device
        eng.infer(imgs)
        per_pass = eng.kernel_launches - launches0
        i2, p2, l2 = eng.infer(imgs[:5], want_logits=True)
        taps = {}
        for l in (2, 3, 4):
            si = names.index(f"layer{l}.0.downsample")
            eng.stop_after(si)
            eng.infer(imgs[:7])
            taps[l] = eng.read_activation(si, 7)
        eng.stop_after(-1)
    monkeypatch.setenv("IDN_NO_DS_FUSION", "1")
    with Engine("resnet18", sd, max_batch=96) as eng:
        j1, q1, m1 = eng.infer(imgs, want_logits=True)
        launches0 = eng.kernel_launches
        eng.infer(imgs)
        per_pass_unfused = eng.kernel_launches - launches0
        for l in (2, 3, 4):
            si = names.index(f"layer{l}.0.downsample")
            eng.stop_after(si)
            eng.infer(imgs[:7])
            assert np.array_equal(eng.read_activation(si, 7), taps[l]), f"layer{l}.0.downsample"
        eng.stop_after(-1)
    assert np.array_equal(l1, m1) and np.array_equal(i1, j1) and np.array_equal(p1, q1)
    assert np.array_equal(l2, l1[:5])
    print(f"\n[resnet18] kernel launches per pass: {per_pass} (downsamples riding along) vs {per_pass_unfused}")
    assert per_pass == per_pass_unfused - 3 and per_pass <= 20, (per_pass, per_pass_unfused)
