"""Prepare the Oxford-IIIT Pet images for `pytorch_oxford_ddp.py` - the data cells of 2_training_oxford-pet_ddp.ipynb (:250-268
corrupt-file list, :378 `splitfolders.ratio(..., seed=1337, ratio=(.8, .1, .1))`) without the notebook or the splitfolders package.

    python tools/prepare_oxford_pet.py --images /data/oxford-iiit-pet/images --output /data/oxford_pet_dataset

Input: the flat `images/` directory of the dataset (`Abyssinian_1.jpg`, `american_bulldog_12.jpg`, ...: class = file name up to the
last underscore).  Output: `<output>/{train,val,test}/<class>/<file>` (copies, or symlinks with --symlink), the layout the
training script's ImageFolder-style dataset reads (`<data_dir>/{train,val}/<class>/*.jpg`, pytorch_oxford_ddp.py:59-89,168,196).
Files that do not decode as RGB images are dropped - the notebook hard-codes the 12 known corrupt JPEGs; here every file is
checked.  Per class: files sorted, shuffled with `random.Random(seed)`, first 80 % train, next 10 % val, rest test (the
splitfolders rule).
"""
from __future__ import annotations

import argparse
import os
import random
import shutil
import sys
from collections import defaultdict

KNOWN_CORRUPT = {"Egyptian_Mau_14.jpg", "Egyptian_Mau_139.jpg", "Egyptian_Mau_145.jpg", "Egyptian_Mau_156.jpg", "Egyptian_Mau_167.jpg",
                 "Egyptian_Mau_177.jpg", "Egyptian_Mau_186.jpg", "Egyptian_Mau_191.jpg", "Abyssinian_5.jpg", "Abyssinian_34.jpg",
                 "chihuahua_121.jpg", "beagle_116.jpg"}


def decodes(path: str) -> bool:
    try:
        from PIL import Image
        with Image.open(path) as im:
            im.verify()
        with Image.open(path) as im:
            im.convert("RGB").load()
        return True
    except Exception:
        return False


def split_class(files, seed: int, ratio=(0.8, 0.1, 0.1)):
    files = sorted(files)
    random.Random(seed).shuffle(files)
    n = len(files)
    a = int(ratio[0] * n)
    b = a + int(ratio[1] * n)
    return {"train": files[:a], "val": files[a:b], "test": files[b:]}


def prepare(images: str, output: str, seed: int = 1337, ratio=(0.8, 0.1, 0.1), symlink: bool = False, check: bool = True):
    by_class = defaultdict(list)
    dropped = []
    for f in sorted(os.listdir(images)):
        if not f.lower().endswith((".jpg", ".jpeg", ".png")) or "_" not in f:
            continue
        if f in KNOWN_CORRUPT or (check and not decodes(os.path.join(images, f))):
            dropped.append(f)
            continue
        by_class[f.rsplit("_", 1)[0]].append(f)
    counts = {"train": 0, "val": 0, "test": 0}
    for cls, files in sorted(by_class.items()):
        for part, fs in split_class(files, seed, ratio).items():
            d = os.path.join(output, part, cls)
            os.makedirs(d, exist_ok=True)
            for f in fs:
                src, dst = os.path.abspath(os.path.join(images, f)), os.path.join(d, f)
                if os.path.lexists(dst):
                    os.remove(dst)
                os.symlink(src, dst) if symlink else shutil.copyfile(src, dst)
                counts[part] += 1
    return {"classes": len(by_class), "dropped": dropped, **counts}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--images", required=True)
    ap.add_argument("--output", required=True)
    ap.add_argument("--seed", type=int, default=1337)
    ap.add_argument("--symlink", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip decoding every file (only the notebook's list of corrupt files is dropped)")
    a = ap.parse_args(argv)
    r = prepare(a.images, a.output, a.seed, symlink=a.symlink, check=not a.no_check)
    print(f"{r['classes']} classes: train {r['train']}, val {r['val']}, test {r['test']}; dropped {len(r['dropped'])}: {r['dropped'][:12]}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
