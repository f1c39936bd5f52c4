"""Local counterparts of the reference's AWS tooling (tools/prepare_imagenet.py, tools/launch_tensorboard.py)."""
import importlib.util
import os
import pickle
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "tools", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _tiny_imagenet(root, sizes=((80, 60), (40, 90), (64, 64))):
    from PIL import Image
    for split, per_class in (("train", 3), ("validation", 2)):
        for c in ("n01", "n02"):
            d = os.path.join(root, split, c)
            os.makedirs(d)
            for i in range(per_class):
                w, h = sizes[i % len(sizes)]
                Image.new("RGB", (w, h), (40 * i, 10, 200)).save(os.path.join(d, f"img{i}.JPEG"))


def test_resize_index_replicate(tmp_path):
    prep = _load("prepare_imagenet")
    from PIL import Image
    root = str(tmp_path / "imagenet")
    _tiny_imagenet(root)
    made = prep.resize(root, [32], workers=1)
    assert made[32] == {"images": 10, "written": 10}
    out = os.path.join(root + "-sz", "32")
    files = prep.list_images(out)
    assert len(files) == 10 and all(f.endswith(".jpg") for f in files)
    for f in files:
        with Image.open(os.path.join(out, f)) as im:
            assert min(im.size) == 32                      # shorter side scaled down, aspect ratio kept
    assert prep.resize(root, [32], workers=1)[32]["written"] == 0        # idempotent
    # aspect-ratio index next to the validation directory, sorted by ratio
    idx = prep.index(root)
    cache = os.path.join(root, "sorted_idxar.p")
    assert os.path.exists(cache) and len(idx) == 4
    with open(cache, "rb") as fh:
        assert len(pickle.load(fh)) == 4
    ratios = [ar for _, ar in idx]
    assert ratios == sorted(ratios)
    # replicas are verified against the manifest of the source
    t1, t2 = str(tmp_path / "disk0"), str(tmp_path / "disk1")
    rep = prep.replicate(root, [t1, t2], checksum=True)
    assert rep[t1]["missing"] == [] and rep[t2]["mismatched"] == []
    assert prep.manifest(t1) == prep.manifest(root)
    prep.replicate(root, [t1, t2], delete=True)
    assert not os.path.exists(t1) and not os.path.exists(t2)


def test_replicate_detects_a_broken_copy(tmp_path, monkeypatch):
    prep = _load("prepare_imagenet")
    root = str(tmp_path / "src")
    _tiny_imagenet(root)
    real = prep._copytree

    def lossy(src, dst):
        real(src, dst)
        os.remove(os.path.join(dst, "train", "n01", "img0.JPEG"))
    monkeypatch.setattr(prep, "_copytree", lossy)
    with pytest.raises(RuntimeError, match="1 missing"):
        prep.replicate(root, [str(tmp_path / "bad")])


def test_launch_tensorboard_dry_run(tmp_path, capsys):
    tb = _load("launch_tensorboard")
    cmd = tb.main([str(tmp_path), "--port", "6123", "--dry-run"])
    assert "--logdir" in cmd and str(tmp_path) in cmd and "6123" in cmd
    assert "Tensorboard at http://" in capsys.readouterr().out


def test_batch_norm_studies_smoke():
    """Script counterparts of the reference's batch-norm notebook run end to end on CPU (tiny settings)."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bn_studies", os.path.join(root, "examples", "batch_norm_studies.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    res = mod.main(["all", "--steps", "4", "--samples", "256", "--width", "8", "--lanczos", "5", "--device", "cpu"])
    assert set(res) == {"lr-stress", "channel-stats", "hessian", "lars"}
    assert len(res["lr-stress"]) == 7 and all(len(r) == 3 for r in res["lr-stress"])
    ev_bn, ev_plain = res["hessian"][True], res["hessian"][False]
    assert len(ev_bn) == 3 and ev_bn[0] >= ev_bn[1] >= ev_bn[2] and len(ev_plain) == 3
    assert all(len(v) == 4 for v in res["lars"].values())
