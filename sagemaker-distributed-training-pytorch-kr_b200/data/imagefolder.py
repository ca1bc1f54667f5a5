"""Image-folder dataset with OpenCV decoding + augmentation (reference: AlbumentationImageDataset,
pytorch_oxford_ddp.py:37-89, and its albumentations pipelines :142-162 / :188-194), plus the synthetic stand-in
of the same shapes that BASELINE.json asks for.  albumentations is not installed here, so the train pipeline is
re-expressed with OpenCV/numpy: RandomResizedCrop(scale 0.8-1), GaussNoise, VerticalFlip, one-of blur family,
one-of CLAHE / sharpen / brightness-contrast, HueSaturationValue, Normalize(ImageNet)."""
from __future__ import annotations

import os
import random
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")
MEAN = np.array([0.485, 0.456, 0.406], dtype=np.float32)
STD = np.array([0.229, 0.224, 0.225], dtype=np.float32)


def find_classes(directory: str) -> Tuple[List[str], Dict[str, int]]:
    classes = sorted(e.name for e in os.scandir(directory) if e.is_dir())
    if not classes:
        raise FileNotFoundError(f"Couldn't find any class folder in {directory}.")
    return classes, {c: i for i, c in enumerate(classes)}


def make_dataset(directory: str, class_to_idx: Dict[str, int]) -> List[Tuple[str, int]]:
    items = []
    for cls in sorted(class_to_idx):
        d = os.path.join(directory, cls)
        if not os.path.isdir(d):
            continue
        for root, _, fnames in sorted(os.walk(d, followlinks=True)):
            for fn in sorted(fnames):
                if fn.lower().endswith(IMG_EXTENSIONS):
                    items.append((os.path.join(root, fn), class_to_idx[cls]))
    return items


class TrainAugment:
    def __init__(self, height: int, width: int, seed: Optional[int] = None):
        self.h, self.w = height, width
        self.rng = random.Random(seed)

    def __call__(self, img: np.ndarray) -> np.ndarray:
        import cv2
        r = self.rng
        H, W = img.shape[:2]
        scale = r.uniform(0.8, 1.0)
        ratio = r.uniform(3 / 4, 4 / 3)
        ch = min(H, int(round((H * W * scale / ratio) ** 0.5)))
        cw = min(W, int(round((H * W * scale * ratio) ** 0.5)))
        y0, x0 = r.randint(0, H - ch), r.randint(0, W - cw)
        img = cv2.resize(img[y0:y0 + ch, x0:x0 + cw], (self.w, self.h), interpolation=cv2.INTER_LINEAR)
        if r.random() < 0.5:
            noise = np.random.normal(0, r.uniform(3, 7), img.shape).astype(np.float32)
            img = np.clip(img.astype(np.float32) + noise, 0, 255).astype(np.uint8)
        if r.random() < 0.5:
            img = img[::-1].copy()
        if r.random() < 0.2:
            k = r.choice([3, 5])
            img = r.choice([lambda a: cv2.blur(a, (k, k)), lambda a: cv2.medianBlur(a, k), lambda a: cv2.GaussianBlur(a, (k, k), 0)])(img)
        if r.random() < 0.3:
            which = r.randint(0, 2)
            if which == 0:
                lab = cv2.cvtColor(img, cv2.COLOR_RGB2LAB)
                lab[..., 0] = cv2.createCLAHE(clipLimit=2.0).apply(lab[..., 0])
                img = cv2.cvtColor(lab, cv2.COLOR_LAB2RGB)
            elif which == 1:
                img = cv2.filter2D(img, -1, np.array([[0, -1, 0], [-1, 5, -1], [0, -1, 0]], dtype=np.float32))
            else:
                a, b = 1 + r.uniform(-0.2, 0.2), 255 * r.uniform(-0.2, 0.2)
                img = np.clip(img.astype(np.float32) * a + b, 0, 255).astype(np.uint8)
        if r.random() < 0.3:
            hsv = cv2.cvtColor(img, cv2.COLOR_RGB2HSV).astype(np.int16)
            hsv[..., 0] = (hsv[..., 0] + r.randint(-20, 20)) % 180
            hsv[..., 1] = np.clip(hsv[..., 1] + r.randint(-30, 30), 0, 255)
            hsv[..., 2] = np.clip(hsv[..., 2] + r.randint(-20, 20), 0, 255)
            img = cv2.cvtColor(hsv.astype(np.uint8), cv2.COLOR_HSV2RGB)
        return (img.astype(np.float32) / 255.0 - MEAN) / STD


class EvalTransform:
    def __init__(self, height: int, width: int):
        self.h, self.w = height, width

    def __call__(self, img):
        import cv2
        img = cv2.resize(img, (self.w, self.h), interpolation=cv2.INTER_LINEAR)
        return (img.astype(np.float32) / 255.0 - MEAN) / STD


class ImageFolderDataset(torch.utils.data.Dataset):
    """<root>/<class>/*.jpg ; cv2.imread + BGR->RGB ; HWC -> CHW float32 (pytorch_oxford_ddp.py:49-57)."""

    def __init__(self, root: str, transform: Optional[Callable] = None, raw_uint8: bool = False):
        self.classes, self.class_to_idx = find_classes(root)
        self.samples = make_dataset(root, self.class_to_idx)
        self.transform = transform
        self.raw_uint8 = raw_uint8          # keep the image as uint8 HWC (the device-side pipeline of data/gpu_augment.py takes it from there)

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        import cv2
        path, target = self.samples[idx]
        raw = cv2.imread(path)
        if raw is None:         # the reference fails inside cvtColor here; the notebook deletes the 12 known corrupt files beforehand
            raise RuntimeError(f"{path} does not decode as an image (tools/prepare_oxford_pet.py drops such files)")
        img = cv2.cvtColor(raw, cv2.COLOR_BGR2RGB)
        if self.transform is not None:
            img = self.transform(img)
        if self.raw_uint8:
            return torch.from_numpy(np.ascontiguousarray(img, dtype=np.uint8)), target
        return torch.from_numpy(np.ascontiguousarray(img.transpose(2, 0, 1), dtype=np.float32)), target


class SyntheticImageDataset(torch.utils.data.Dataset):
    """Random normalised images of the named shape with class-dependent means (so loss can fall)."""

    def __init__(self, n: int, num_classes: int, height: int, width: int, seed: int = 0):
        self.n, self.nc, self.h, self.w, self.seed = n, num_classes, height, width, seed
        self.classes = [f"class_{i:03d}" for i in range(num_classes)]
        self.class_to_idx = {c: i for i, c in enumerate(self.classes)}

    def __len__(self):
        return self.n

    def __getitem__(self, idx):
        g = torch.Generator().manual_seed(self.seed * 1000003 + idx)
        y = idx % self.nc
        x = torch.randn(3, self.h, self.w, generator=g) * 0.5 + (y / max(self.nc - 1, 1) - 0.5)
        return x, y
