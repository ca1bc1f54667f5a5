"""Training-time image augmentation as batched tensor ops on the device the model runs on.

The reference augments on the host, one image at a time, in two DataLoader workers (albumentations over OpenCV,
2_training_oxford-pet_ddp/pytorch_oxford_ddp.py:142-162, loader `:173-181`).  At a few thousand images per second and GPU that
pipeline is the bottleneck long before the model is, and it ships fp32 CHW tensors over PCIe (12 bytes per pixel).  Here the
workers only decode and resize to a fixed pre-crop size; the batch crosses PCIe as uint8 HWC (3 bytes per pixel) and the random
part runs on the GPU for the whole batch at once:

    RandomResizedCrop(scale 0.8-1, ratio 3/4-4/3) + VerticalFlip(0.5)   one affine grid + bilinear grid_sample
    GaussNoise(0.5)                                                     sigma 3-7 grey levels per image
    one of box / Gaussian blur, k in {3, 5} (0.2)                        depthwise convolutions, selected per image
    one of sharpen / brightness-contrast (0.3)                           (CLAHE and median blur have no cheap batched form: their
                                                                         probability mass goes to their siblings)
    HueSaturationValue(0.3)                                             RGB -> HSV shift -> RGB
    Normalize(ImageNet mean / std)

Every random draw comes from one torch.Generator on the device, so a seed reproduces a run.  The same code runs on the CPU
(tests/test_vision_ops_cpu.py)."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

MEAN = (0.485, 0.456, 0.406)
STD = (0.229, 0.224, 0.225)


def rgb_to_hsv(x: torch.Tensor) -> torch.Tensor:
    """x [B, 3, H, W] in [0, 1] -> h in [0, 1), s, v in [0, 1]"""
    r, g, b = x[:, 0], x[:, 1], x[:, 2]
    mx, mn = x.amax(1), x.amin(1)
    d = mx - mn
    safe = torch.where(d > 0, d, torch.ones_like(d))
    h = torch.where(mx == r, ((g - b) / safe) % 6.0, torch.where(mx == g, (b - r) / safe + 2.0, (r - g) / safe + 4.0))
    h = torch.where(d > 0, h / 6.0, torch.zeros_like(h))
    s = torch.where(mx > 0, d / torch.where(mx > 0, mx, torch.ones_like(mx)), torch.zeros_like(mx))
    return torch.stack([h, s, mx], 1)


def hsv_to_rgb(x: torch.Tensor) -> torch.Tensor:
    h, s, v = x[:, 0] * 6.0, x[:, 1], x[:, 2]
    i = torch.floor(h)
    f = h - i
    p, q, t = v * (1 - s), v * (1 - s * f), v * (1 - s * (1 - f))
    i = i.long() % 6
    r = torch.stack([v, q, p, p, t, v], 1).gather(1, i[:, None])[:, 0]
    g = torch.stack([t, v, v, q, p, p], 1).gather(1, i[:, None])[:, 0]
    b = torch.stack([p, p, t, v, v, q], 1).gather(1, i[:, None])[:, 0]
    return torch.stack([r, g, b], 1)


def _depthwise(x: torch.Tensor, k2d: torch.Tensor) -> torch.Tensor:
    k = k2d.shape[-1]
    w = k2d.to(x.dtype).expand(x.shape[1], 1, k, k).contiguous()
    return F.conv2d(F.pad(x, (k // 2,) * 4, mode="replicate"), w, groups=x.shape[1])


def _gauss_kernel(k: int, device) -> torch.Tensor:
    sigma = 0.3 * ((k - 1) * 0.5 - 1) + 0.8                    # OpenCV's rule for sigma = 0
    a = torch.arange(k, device=device, dtype=torch.float32) - (k - 1) / 2
    g = torch.exp(-a * a / (2 * sigma * sigma))
    g = g / g.sum()
    return g[:, None] * g[None, :]


class GpuAugment:
    """uint8 [B, h0, w0, 3] (any fixed pre-crop size) -> normalised float [B, 3, height, width] in `dtype`."""

    def __init__(self, height: int, width: int, seed: Optional[int] = None, dtype: torch.dtype = torch.float32,
                 p_noise: float = 0.5, p_flip: float = 0.5, p_blur: float = 0.2, p_tone: float = 0.3, p_hsv: float = 0.3,
                 scale=(0.8, 1.0), ratio=(3 / 4, 4 / 3)):
        self.h, self.w, self.seed, self.dtype = height, width, seed, dtype
        self.p_noise, self.p_flip, self.p_blur, self.p_tone, self.p_hsv = p_noise, p_flip, p_blur, p_tone, p_hsv
        self.scale, self.ratio = scale, ratio
        self._gen = None

    def _g(self, device):
        if self._gen is None or self._gen.device != device:
            self._gen = torch.Generator(device=device)
            if self.seed is not None:
                self._gen.manual_seed(self.seed)
        return self._gen

    def __call__(self, batch_u8: torch.Tensor) -> torch.Tensor:
        assert batch_u8.dtype == torch.uint8 and batch_u8.dim() == 4 and batch_u8.shape[-1] == 3
        dev, B = batch_u8.device, batch_u8.shape[0]
        g = self._g(dev)

        def u(lo, hi, n=B):
            return torch.rand(n, device=dev, generator=g) * (hi - lo) + lo

        def coin(p):
            return torch.rand(B, device=dev, generator=g) < p
        x = batch_u8.permute(0, 3, 1, 2).float() / 255.0                         # [B, 3, h0, w0] in [0, 1]
        # --- RandomResizedCrop + VerticalFlip as one sampling grid (normalised coordinates: the crop is a scaled, shifted box)
        area, rat = u(*self.scale), torch.exp(u(math.log(self.ratio[0]), math.log(self.ratio[1])))
        sw = torch.sqrt(area * rat).clamp(max=1.0)                               # crop extent as a fraction of the source width / height
        sh = torch.sqrt(area / rat).clamp(max=1.0)
        cx = (u(0.0, 1.0) * 2 - 1) * (1 - sw)                                    # crop centre keeps the box inside the image
        cy = (u(0.0, 1.0) * 2 - 1) * (1 - sh)
        flip = torch.where(coin(self.p_flip), -1.0, 1.0)
        theta = torch.zeros(B, 2, 3, device=dev)
        theta[:, 0, 0], theta[:, 0, 2] = sw, cx
        theta[:, 1, 1], theta[:, 1, 2] = sh * flip, cy
        grid = F.affine_grid(theta, (B, 3, self.h, self.w), align_corners=False)
        x = F.grid_sample(x, grid, mode="bilinear", padding_mode="border", align_corners=False)
        # --- GaussNoise
        sigma = torch.where(coin(self.p_noise), u(3.0, 7.0), torch.zeros(B, device=dev)) / 255.0
        x = (x + torch.randn(x.shape, device=dev, generator=g) * sigma[:, None, None, None]).clamp_(0, 1)
        # --- blur: box or Gaussian, 3 or 5 taps, chosen per image
        if self.p_blur > 0:
            on, kind, five = coin(self.p_blur), coin(0.5), coin(0.5)
            out = x
            for k in (3, 5):
                for gauss in (False, True):
                    sel = on & (kind == gauss) & (five == (k == 5))          # no host-side test of `sel`: nothing here synchronises
                    ker = _gauss_kernel(k, dev) if gauss else torch.full((k, k), 1.0 / (k * k), device=dev)
                    out = torch.where(sel[:, None, None, None], _depthwise(x, ker), out)
            x = out
        # --- sharpen or brightness / contrast
        if self.p_tone > 0:
            on, sharp = coin(self.p_tone), coin(0.5)
            ker = torch.tensor([[0.0, -1.0, 0.0], [-1.0, 5.0, -1.0], [0.0, -1.0, 0.0]], device=dev)
            a, b = 1 + u(-0.2, 0.2), u(-0.2, 0.2)
            toned = torch.where(sharp[:, None, None, None], _depthwise(x, ker), x * a[:, None, None, None] + b[:, None, None, None])
            x = torch.where(on[:, None, None, None], toned, x).clamp_(0, 1)
        # --- HueSaturationValue (OpenCV units: hue +-20 of 180, saturation +-30 of 255, value +-20 of 255)
        if self.p_hsv > 0:
            on = coin(self.p_hsv)
            hsv = rgb_to_hsv(x)
            dh, ds, dv = u(-20.0, 20.0) / 180.0, u(-30.0, 30.0) / 255.0, u(-20.0, 20.0) / 255.0
            hsv = torch.stack([(hsv[:, 0] + dh[:, None, None]) % 1.0, (hsv[:, 1] + ds[:, None, None]).clamp(0, 1),
                               (hsv[:, 2] + dv[:, None, None]).clamp(0, 1)], 1)
            x = torch.where(on[:, None, None, None], hsv_to_rgb(hsv), x)
        mean = torch.tensor(MEAN, device=dev)[None, :, None, None]
        std = torch.tensor(STD, device=dev)[None, :, None, None]
        return ((x - mean) / std).to(self.dtype)


class DecodeResize:
    """Host side of the GPU pipeline (a Dataset transform): RGB uint8 image -> uint8 [h0, w0, 3] at a fixed pre-crop size, so that
    batches collate; the random crop is taken from it on the device."""

    def __init__(self, height: int, width: int, margin: float = 1.15):
        self.h0, self.w0 = int(round(height * margin)), int(round(width * margin))

    def __call__(self, img):
        import cv2
        return cv2.resize(img, (self.w0, self.h0), interpolation=cv2.INTER_LINEAR)
