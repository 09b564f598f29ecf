"""AutoAugment, the fixed ImageNet policy (Cubuk et al. 2018, appendix table 9): 25 sub-policies
of two (operation, probability, strength 0..10) steps; one sub-policy is drawn per image.
API parity with megatron/data/autoaugment.py:49-320 (``ImageNetPolicy``, ``SubPolicy``); like
there it is CPU/PIL-only and no model in the tree consumes it.
"""
from __future__ import annotations

import random

import numpy as np

_LEVELS = 11

# strength index -> magnitude, per operation (None: the operation takes no magnitude)
_RANGES = {
    "shearX": np.linspace(0, 0.3, _LEVELS), "shearY": np.linspace(0, 0.3, _LEVELS),
    "translateX": np.linspace(0, 150 / 331, _LEVELS), "translateY": np.linspace(0, 150 / 331, _LEVELS),
    "rotate": np.linspace(0, 30, _LEVELS),
    "color": np.linspace(0.0, 0.9, _LEVELS), "contrast": np.linspace(0.0, 0.9, _LEVELS),
    "sharpness": np.linspace(0.0, 0.9, _LEVELS), "brightness": np.linspace(0.0, 0.9, _LEVELS),
    "posterize": np.round(np.linspace(8, 4, _LEVELS)).astype(int),
    "solarize": np.linspace(256, 0, _LEVELS),
    "autocontrast": None, "equalize": None, "invert": None,
}

_IMAGENET_TABLE = [
    ("posterize", 0.4, 8, "rotate", 0.6, 9), ("solarize", 0.6, 5, "autocontrast", 0.6, 5),
    ("equalize", 0.8, 8, "equalize", 0.6, 3), ("posterize", 0.6, 7, "posterize", 0.6, 6),
    ("equalize", 0.4, 7, "solarize", 0.2, 4), ("equalize", 0.4, 4, "rotate", 0.8, 8),
    ("solarize", 0.6, 3, "equalize", 0.6, 7), ("posterize", 0.8, 5, "equalize", 1.0, 2),
    ("rotate", 0.2, 3, "solarize", 0.6, 8), ("equalize", 0.6, 8, "posterize", 0.4, 6),
    ("rotate", 0.8, 8, "color", 0.4, 0), ("rotate", 0.4, 9, "equalize", 0.6, 2),
    ("equalize", 0.0, 7, "equalize", 0.8, 8), ("invert", 0.6, 4, "equalize", 1.0, 8),
    ("color", 0.6, 4, "contrast", 1.0, 8), ("rotate", 0.8, 8, "color", 1.0, 2),
    ("color", 0.8, 8, "solarize", 0.8, 7), ("sharpness", 0.4, 7, "invert", 0.6, 8),
    ("shearX", 0.6, 5, "equalize", 1.0, 9), ("color", 0.4, 0, "equalize", 0.6, 3),
    ("equalize", 0.4, 7, "solarize", 0.2, 4), ("solarize", 0.6, 5, "autocontrast", 0.6, 5),
    ("invert", 0.6, 4, "equalize", 1.0, 8), ("color", 0.6, 4, "contrast", 1.0, 8),
    ("equalize", 0.8, 8, "equalize", 0.6, 3),
]


def _sign():
    return random.choice((-1, 1))


def _apply(op: str, img, mag, fill):
    from PIL import Image, ImageEnhance, ImageOps
    if op in ("shearX", "shearY", "translateX", "translateY"):
        m = mag * _sign()
        coeffs = {"shearX": (1, m, 0, 0, 1, 0), "shearY": (1, 0, 0, m, 1, 0),
                  "translateX": (1, 0, m * img.size[0], 0, 1, 0),
                  "translateY": (1, 0, 0, 0, 1, m * img.size[1])}[op]
        return img.transform(img.size, Image.AFFINE, coeffs, Image.BICUBIC, fillcolor=fill)
    if op == "rotate":
        # rotate on an alpha canvas, then lay it over the fill colour so the corners are not black
        rot = img.convert("RGBA").rotate(mag * _sign())
        canvas = Image.new("RGBA", rot.size, tuple(fill) + (255,))
        return Image.composite(rot, canvas, rot).convert(img.mode)
    if op in ("color", "contrast", "sharpness", "brightness"):
        enh = {"color": ImageEnhance.Color, "contrast": ImageEnhance.Contrast,
               "sharpness": ImageEnhance.Sharpness, "brightness": ImageEnhance.Brightness}[op]
        return enh(img).enhance(1 + mag * _sign())
    if op == "posterize":
        return ImageOps.posterize(img, int(mag))
    if op == "solarize":
        return ImageOps.solarize(img, mag)
    if op == "autocontrast":
        return ImageOps.autocontrast(img)
    if op == "equalize":
        return ImageOps.equalize(img)
    if op == "invert":
        return ImageOps.invert(img)
    raise ValueError(op)


class SubPolicy:
    def __init__(self, operation1, probability1, magnitude_idx1, operation2, probability2, magnitude_idx2,
                 fillcolor=(128, 128, 128)):
        self.steps = []
        for op, p, idx in ((operation1, probability1, magnitude_idx1), (operation2, probability2, magnitude_idx2)):
            if op not in _RANGES:
                raise ValueError(f"unsupported AutoAugment operation {op!r}")
            if not (0.0 <= p <= 1.0) or not (isinstance(idx, int) and 0 <= idx < _LEVELS):
                raise ValueError(f"bad probability / strength for {op}: {p}, {idx}")
            rng = _RANGES[op]
            self.steps.append((op, p, None if rng is None else rng[idx]))
        self.fillcolor = tuple(fillcolor)

    def __call__(self, img):
        for op, p, mag in self.steps:
            if random.random() < p:
                img = _apply(op, img, mag, self.fillcolor)
        return img


class ImageNetPolicy:
    """Usable as a torchvision-style transform: ``Compose([Resize(256), ImageNetPolicy(), ToTensor()])``."""

    def __init__(self, fillcolor=(128, 128, 128)):
        self.policies = [SubPolicy(*row, fillcolor=fillcolor) for row in _IMAGENET_TABLE]

    def __call__(self, img):
        return random.choice(self.policies)(img)

    def __repr__(self):
        return "ImageNetPolicy"
