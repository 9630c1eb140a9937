"""CPU: pin the oracle (oracle/) against Pillow/torchvision and against the golden fixtures that
oracle/make_golden.py produced by running the UNMODIFIED reference file."""
import os

import numpy as np
import pytest
import torch

from idunno_b200.synth import MIXED_SIZES, make_state_dict, synth_image
from oracle import pil_resize, ref_path

CASES = ["alexnet_default", "alexnet_stress", "resnet18_default", "resnet18_stress",
         "alexnet_diverse400", "resnet18_diverse400"]


def test_resize_matches_pillow_bit_exact():
    from PIL import Image
    for i, (h, w) in enumerate(MIXED_SIZES + [(1000, 130), (97, 1201)]):
        img = synth_image(i, size=(h, w))
        nh, nw = pil_resize.resized_size(h, w)
        want = np.asarray(Image.fromarray(img).resize((nw, nh), Image.BILINEAR))
        got = pil_resize.resize_bilinear_u8(img, nh, nw) if (nh, nw) != (h, w) else img
        assert np.array_equal(got, want), (h, w)


def test_preprocess_matches_torchvision_bit_exact():
    from PIL import Image
    from torchvision import transforms
    tf = transforms.Compose([transforms.Resize(256), transforms.CenterCrop(224), transforms.ToTensor(),
                             transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    for i in range(len(MIXED_SIZES)):
        img = synth_image(i, mixed=True)
        assert np.array_equal(pil_resize.preprocess(img), tf(Image.fromarray(img)).numpy())


@pytest.mark.parametrize("model", ["alexnet", "resnet18"])
def test_functional_forward_matches_torchvision_module(model):
    import torchvision
    sd = make_state_dict(model, 0, "stress")
    net = getattr(torchvision.models, model)(weights=None)
    net.load_state_dict(sd)
    net.eval()
    x = torch.from_numpy(np.stack([pil_resize.preprocess(synth_image(i)) for i in range(3)]))
    with torch.no_grad():
        want = net(x)
    got = ref_path.forward(model, sd, x)
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-4)
    taps = ref_path.forward_taps(model, sd, x)
    last = "classifier.6" if model == "alexnet" else "layer4.1.conv2+add"
    assert last in taps


@pytest.mark.parametrize("case", CASES)
def test_restatement_matches_reference_fixture(case, golden_dir):
    g = np.load(os.path.join(golden_dir, case + ".npz"))
    model, variant = str(g["model"]), str(g["variant"])
    first, last, mixed = int(g["first"]), int(g["last"]), bool(g["mixed"])
    skipped = set(int(s) for s in g["skipped"])
    sd = make_state_dict(model, 0, variant)
    import hashlib
    w_sha = hashlib.sha256(b"".join(v.numpy().tobytes() for v in sd.values())).hexdigest()
    assert w_sha == str(g["weights_sha256"]), "seeded weights differ from the fixture's"
    idxs = [i for i in range(first, last + 1) if i not in skipped]
    imgs = [synth_image(i, mixed=mixed) for i in idxs]
    r = ref_path.infer_arrays(model, sd, imgs)
    if variant == "diverse":
        _check_diverse_properties(model, sd, idxs, r, g)
    assert [f"test_{i}.JPEG" for i in idxs] == [str(n) for n in g["names"]]
    for k, p in enumerate(r["pre"]):
        assert hashlib.sha256(p.tobytes()).hexdigest() == str(g["pre_sha256"][k])
    np.testing.assert_allclose(r["logits"], g["logits"], rtol=1e-4, atol=1e-5)
    assert np.array_equal(r["idx"], g["idx"])
    np.testing.assert_allclose(r["prob"], g["prob"], rtol=1e-4)
    cats = ref_path.imagenet_categories()
    assert [cats[i] for i in r["idx"]] == [str(c) for c in g["categories"]]


def _check_diverse_properties(model, sd, idxs, r, g):
    """What makes the BASELINE-size fixture a parity set that can fail: the
    unmodified reference predicts the KNOWN class of every picture (400 distinct classes), and the margins
    are far above what bf16 arithmetic costs on these weights."""
    from idunno_b200.synth import diverse_class_of
    want_cls = np.asarray([diverse_class_of(model, i) for i in idxs])
    assert np.array_equal(g["idx"], want_cls), "the reference does not predict the classes the head has learnt"
    assert len(set(g["idx"].tolist())) >= 50
    srt = np.sort(g["logits"], axis=1)
    margin = srt[:, -1] - srt[:, -2]
    x = torch.from_numpy(r["pre"])
    contract = torch.cat([ref_path.forward_bf16_contract(model, sd, x[i:i + 50]) for i in range(0, len(x), 50)]).numpy()
    err = float(np.abs(contract - g["logits"]).max())
    print(f"\n[{model}/diverse400] median margin {np.median(margin):.3f}, bf16-contract error {err:.4f} "
          f"({err / float(np.abs(g['logits']).max()):.3%} of the logit range), ratio {np.median(margin) / err:.1f}, "
          f"margin > 2 err on {int((margin > 2 * err).sum())}/{len(margin)}")
    assert np.median(margin) >= 10 * err
    assert int((margin > 2 * err).sum()) >= 0.9 * len(margin)
    assert np.array_equal(contract.argmax(axis=1), g["idx"])


def test_deeplearning_restatement_file_loop(tmp_path, monkeypatch):
    """The restated loop: inclusive range, str/int arguments, missing files skipped, any model name
    other than "alexnet" means ResNet18 (alexnet_resnet.py:13-22,48-49)."""
    from PIL import Image
    monkeypatch.chdir(tmp_path)
    os.makedirs("resnet18")
    for i in (3, 4, 6):
        Image.fromarray(synth_image(i)).save(f"resnet18/test_{i}.JPEG", format="PNG")
    sd = make_state_dict("resnet18", 0, "stress")
    cats = ref_path.imagenet_categories()
    tap = {}
    res, secs = ref_path.deeplearning("resnet18", "something-else", "3", 6, sd, cats, tap)
    assert [r[0] for r in res] == ["test_3.JPEG", "test_4.JPEG", "test_6.JPEG"]
    assert tap["logits"].shape == (3, 1000) and secs > 0
    assert ref_path.deeplearning("resnet18", "resnet18", 7, 9, sd, cats)[0] == []
