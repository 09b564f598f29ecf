"""Class-per-directory image dataset (``root/<class>/<file>``), kept for API parity with
megatron/data/image_folder.py:114-302 (``DatasetFolder`` / ``ImageFolder``). No vision model ships
with either framework; the class is a plain CPU dataset that hands PIL images to a transform.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

from torch.utils.data import Dataset

IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")


def has_file_allowed_extension(filename: str, extensions: Sequence[str]) -> bool:
    return filename.lower().endswith(tuple(extensions))


def is_image_file(filename: str) -> bool:
    return has_file_allowed_extension(filename, IMG_EXTENSIONS)


def find_classes(root: str) -> Tuple[List[str], Dict[str, int]]:
    """Sub-directories of ``root``, sorted, and their indices."""
    classes = sorted(e.name for e in os.scandir(root) if e.is_dir())
    if not classes:
        raise FileNotFoundError(f"no class directories under {root}")
    return classes, {c: i for i, c in enumerate(classes)}


def make_dataset(directory: str, class_to_idx: Dict[str, int], extensions: Optional[Sequence[str]] = None,
                 is_valid_file: Optional[Callable[[str], bool]] = None,
                 data_per_class_fraction: float = 1.0) -> List[Tuple[str, int]]:
    """[(path, class index)] in (class, path) order. Exactly one of ``extensions`` /
    ``is_valid_file`` selects the files; ``data_per_class_fraction`` keeps the first share of each
    class (the few-shot switch of image_folder.py:100-108)."""
    if (extensions is None) == (is_valid_file is None):
        raise ValueError("pass exactly one of extensions and is_valid_file")
    accept = is_valid_file or (lambda p: has_file_allowed_extension(p, extensions))
    samples: List[Tuple[str, int]] = []
    for cls in sorted(class_to_idx):
        cdir = os.path.join(os.path.expanduser(directory), cls)
        if not os.path.isdir(cdir):
            continue
        found = []
        for dirpath, _, files in sorted(os.walk(cdir, followlinks=True)):
            found += [os.path.join(dirpath, f) for f in sorted(files) if accept(os.path.join(dirpath, f))]
        keep = int(len(found) * data_per_class_fraction)
        samples += [(p, class_to_idx[cls]) for p in found[:keep]]
    return samples


def pil_loader(path: str):
    from PIL import Image
    with open(path, "rb") as f:
        return Image.open(f).convert("RGB")


default_loader = pil_loader


class DatasetFolder(Dataset):
    """``loader(path)`` -> sample, optional ``transform`` / ``target_transform``.
    ``classes_fraction`` keeps the first share of the (sorted) classes."""

    def __init__(self, root: str, loader: Callable[[str], Any] = default_loader,
                 extensions: Optional[Sequence[str]] = None, transform: Optional[Callable] = None,
                 target_transform: Optional[Callable] = None, classes_fraction: float = 1.0,
                 data_per_class_fraction: float = 1.0, is_valid_file: Optional[Callable[[str], bool]] = None):
        self.root = root
        self.loader, self.extensions = loader, extensions
        self.transform, self.target_transform = transform, target_transform
        classes, class_to_idx = find_classes(root)
        classes = classes[:max(1, int(len(classes) * classes_fraction))]
        class_to_idx = {c: class_to_idx[c] for c in classes}
        self.samples = make_dataset(root, class_to_idx, extensions, is_valid_file, data_per_class_fraction)
        if not self.samples:
            raise FileNotFoundError(f"no valid files under {root}")
        self.classes, self.class_to_idx = classes, class_to_idx
        self.targets = [t for _, t in self.samples]
        self.total = len(self.samples)

    def __len__(self) -> int:
        return self.total

    def __getitem__(self, index: int):
        # an unreadable file is replaced by the next one instead of killing the epoch
        for probe in range(self.total):
            path, target = self.samples[(index + probe) % self.total]
            try:
                sample = self.loader(path)
                break
            except OSError:
                continue
        else:
            raise OSError(f"no readable sample in {self.root}")
        if self.transform is not None:
            sample = self.transform(sample)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return sample, target


class ImageFolder(DatasetFolder):
    def __init__(self, root: str, transform: Optional[Callable] = None, target_transform: Optional[Callable] = None,
                 classes_fraction: float = 1.0, data_per_class_fraction: float = 1.0,
                 loader: Callable[[str], Any] = default_loader, is_valid_file: Optional[Callable[[str], bool]] = None):
        super().__init__(root, loader, None if is_valid_file else IMG_EXTENSIONS, transform, target_transform,
                         classes_fraction, data_per_class_fraction, is_valid_file)
        self.imgs = self.samples
