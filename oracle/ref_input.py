"""Input pipeline step in front of the path — CPU restatement.  TEST INFRASTRUCTURE ONLY.

The reference normalises every image with albumentations before the model sees it
(datasets/cityscapes.py:122-123: `AT.Normalize(mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225))`
followed by `ToTensorV2()`), then `images.to(device, dtype=torch.float32)` (core/seg_trainer.py:42).

albumentations is a third-party dependency of the reference (requirements.txt, unpinned) and is NOT installed in
this image, so this function restates its published algorithm (albumentations.augmentations.functional.normalize):

    mean = np.array(mean, dtype=np.float32) * max_pixel_value        # float32 array times 255.0
    std  = np.array(std,  dtype=np.float32) * max_pixel_value
    denominator = np.reciprocal(std, dtype=np.float32)
    img = img.astype(np.float32); img -= mean; img *= denominator

PARITY UNPINNED for this one function: there is no albumentations here to run and the reference holds no golden
vector for it.  It is cross-checked in tests/test_oracle_golden.py against the mathematically equivalent
torchvision-style formula ((x/255 - mean)/std in float64) to within float32 rounding.
"""
from __future__ import annotations

import numpy as np
import torch

CITYSCAPES_MEAN = (0.485, 0.456, 0.406)   # reference datasets/cityscapes.py:122
CITYSCAPES_STD = (0.229, 0.224, 0.225)


def normalize_u8(img_u8, mean=CITYSCAPES_MEAN, std=CITYSCAPES_STD) -> torch.Tensor:
    """uint8 [N, H, W, C] (numpy or torch) -> float32 NCHW tensor, as AT.Normalize + ToTensorV2 produce it."""
    img = img_u8.numpy() if isinstance(img_u8, torch.Tensor) else np.asarray(img_u8)
    assert img.dtype == np.uint8 and img.ndim == 4
    c = img.shape[-1]
    m = np.array(mean[:c], dtype=np.float32)
    m *= 255.0
    s = np.array(std[:c], dtype=np.float32)
    s *= 255.0
    den = np.reciprocal(s, dtype=np.float32)
    out = img.astype(np.float32)
    out -= m
    out *= den
    return torch.from_numpy(np.ascontiguousarray(out.transpose(0, 3, 1, 2)))   # ToTensorV2: HWC -> CHW


def make_image_u8(seed: int, n: int, h: int, w: int, c: int = 3) -> torch.Tensor:
    """Seeded synthetic decoded images: uint8 [N, H, W, C]."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, h, w, c), generator=g, dtype=torch.uint8)
