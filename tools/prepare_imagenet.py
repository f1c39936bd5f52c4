#!/usr/bin/env python
"""Dataset preparation for the progressive-resizing ImageNet recipe, on local disks.

Local counterpart of the reference's AWS tooling (``IMAGENET/tools/replicate_imagenet.py`` clones an EBS snapshot of
the prepared data set into one volume per machine, ``create_imagenet_snapshot.py`` is the recipe that made the
snapshot, ``setup.sh`` downloads the validation aspect-ratio index): the same three jobs for directories.

    # 1. down-scaled copies the early phases read (``trndir: "-sz/160"`` -> <root>-sz/160/{train,validation})
    python tools/prepare_imagenet.py resize  /data/imagenet --sizes 160 352 --workers 32
    # 2. the aspect-ratio index rectangular validation sorts by (<root>/sorted_idxar.p)
    python tools/prepare_imagenet.py index   /data/imagenet
    # 3. one replica per node-local disk, verified against a manifest
    python tools/prepare_imagenet.py replicate /data/imagenet /nvme0/imagenet /nvme1/imagenet

Layout throughout: ``<root>/{train,validation}/<class>/<image>`` (torchvision ``ImageFolder``).
"""
import argparse
import hashlib
import json
import os
import shutil
import sys
from concurrent.futures import ProcessPoolExecutor

IMG_EXT = (".jpg", ".jpeg", ".png", ".bmp", ".webp")
SPLITS = ("train", "validation")


def list_images(root: str):
    """Relative paths of every image below ``root`` (sorted: the order is part of the manifest)."""
    out = []
    for d, _, files in os.walk(root):
        for f in files:
            if f.lower().endswith(IMG_EXT):
                out.append(os.path.relpath(os.path.join(d, f), root))
    return sorted(out)


# ---- resize -------------------------------------------------------------------------------------------------
def _resize_one(job):
    src, dst, size = job
    from PIL import Image
    if os.path.exists(dst):
        return 0
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    with Image.open(src) as im:
        im = im.convert("RGB")
        w, h = im.size
        if min(w, h) > size:                       # shorter side -> size, never up-scale
            s = size / min(w, h)
            im = im.resize((max(1, round(w * s)), max(1, round(h * s))), Image.BILINEAR)
        tmp = dst + ".tmp"
        im.save(tmp, format="JPEG", quality=90)
        os.replace(tmp, dst)
    return 1


def resize(root: str, sizes, workers: int = 8, splits=SPLITS):
    """``<root>-sz/<size>/<split>/...``: copies whose shorter side is ``size`` (JPEG, quality 90)."""
    root = root.rstrip("/")
    made = {}
    for size in sizes:
        jobs = []
        for split in splits:
            sdir = os.path.join(root, split)
            if not os.path.isdir(sdir):
                continue
            for rel in list_images(sdir):
                dst = os.path.join(f"{root}-sz", str(size), split, os.path.splitext(rel)[0] + ".jpg")
                jobs.append((os.path.join(sdir, rel), dst, size))
        if workers > 1 and len(jobs) > 64:
            with ProcessPoolExecutor(workers) as ex:
                n = sum(ex.map(_resize_one, jobs, chunksize=256))
        else:
            n = sum(_resize_one(j) for j in jobs)
        made[size] = {"images": len(jobs), "written": n}
        print(f"size {size}: {len(jobs)} images, {n} written -> {root}-sz/{size}")
    return made


# ---- aspect-ratio index -----------------------------------------------------------------------------------------
def index(root: str):
    """Writes ``<root>/sorted_idxar.p`` (what ``data.imagenet.sort_ar`` caches; the reference downloads it)."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import layerwise_b200  # noqa: F401
    from layerwise_b200.data.imagenet import sort_ar
    valdir = os.path.join(root, "validation")
    idx = sort_ar(valdir)
    print(f"{len(idx)} validation images indexed by aspect ratio -> {os.path.join(root, 'sorted_idxar.p')}")
    return idx


# ---- replication ------------------------------------------------------------------------------------------------
def manifest(root: str, checksum: bool = False):
    """``{relative path: size}`` (or ``[size, sha1]``) of every file below ``root``."""
    out = {}
    for d, _, files in os.walk(root):
        for f in sorted(files):
            if f == "MANIFEST.json":
                continue
            path = os.path.join(d, f)
            rel = os.path.relpath(path, root)
            size = os.path.getsize(path)
            if checksum:
                h = hashlib.sha1()
                with open(path, "rb") as fh:
                    for blk in iter(lambda: fh.read(1 << 20), b""):
                        h.update(blk)
                out[rel] = [size, h.hexdigest()]
            else:
                out[rel] = size
    return out


def _copytree(src: str, dst: str):
    shutil.copytree(src, dst, dirs_exist_ok=True)


def replicate(src: str, targets, checksum: bool = False, delete: bool = False):
    """Copy the prepared tree to every target (a node-local disk per replica) and verify it against the manifest of the
    source; ``delete`` removes the replicas instead (the reference tool's ``--delete``)."""
    if delete:
        for t in targets:
            shutil.rmtree(t, ignore_errors=True)
            print(f"removed {t}")
        return {}
    want = manifest(src, checksum)
    report = {}
    for t in targets:
        _copytree(src, t)
        got = manifest(t, checksum)
        missing = sorted(set(want) - set(got))
        wrong = sorted(k for k in want if k in got and got[k] != want[k])
        with open(os.path.join(t, "MANIFEST.json"), "w") as fh:
            json.dump({"source": os.path.abspath(src), "files": len(want), "checksum": checksum}, fh)
        report[t] = {"files": len(got), "missing": missing, "mismatched": wrong}
        status = "ok" if not missing and not wrong else f"{len(missing)} missing, {len(wrong)} mismatched"
        print(f"{t}: {len(got)} files, {status}")
        if missing or wrong:
            raise RuntimeError(f"replica {t} differs from {src}: {status}")
    return report


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    r = sub.add_parser("resize")
    r.add_argument("root")
    r.add_argument("--sizes", type=int, nargs="+", default=[160, 352])
    r.add_argument("--workers", type=int, default=8)
    i = sub.add_parser("index")
    i.add_argument("root")
    c = sub.add_parser("replicate")
    c.add_argument("src")
    c.add_argument("targets", nargs="+")
    c.add_argument("--checksum", action="store_true", help="compare SHA-1 digests, not only sizes")
    c.add_argument("--delete", action="store_true", help="remove the replicas instead of creating them")
    args = ap.parse_args(argv)
    if args.cmd == "resize":
        return resize(args.root, args.sizes, args.workers)
    if args.cmd == "index":
        return index(args.root)
    return replicate(args.src, args.targets, args.checksum, args.delete)


if __name__ == "__main__":
    main()
