"""The data format either side of the training step: models/dataset.py:28-53, 206-305.

The reference stores every 3 s chunk as a grayscale PNG (128 x 130) under ``<root>/<label>/`` and decodes it with
PIL inside ``__getitem__`` -- per sample, ``num_workers=0`` (train.py:311) -- so the loader, not the GPU, paces its
training loop.  Here the same images live in ONE packed uint8 tensor per label:

  PackedSpectrograms          {label: uint8 [n, H, W]} + a flat device pool; built once from the PNG folders
                              (``from_folder``: the reference's ImageFolder order = sorted file names) or from arrays,
                              saved / loaded as a single ``.pt`` file (the on-disk format)
  SpectrogramPairDataset      same constructor, ``__len__`` / ``__getitem__`` contract and CSV pairing file as the
                              reference class (dataset.py:206-241): element = ((img1, label1), (img2, label2)) with
                              img = crop(0, 0, 128, 128) / 255 as float32 [1, 128, 128]
      .generate_pairings      dataset.py:263-303: the same RandomState(42) draw -> the same CSV
      .loader(batch_size)     batches for ``LDMTrainer.train_epoch``: ONE gather kernel per batch reads the paired
                              images straight from the device pool (crop + /255 fused, mstk_u8_gather_crop); no PIL,
                              no per-sample Python, no host->device copy of pixels in the loop
  SpectrogramDataset          dataset.py:28-53 (autoencoder stage), same packed storage; prepare_dataset(config)

``__getitem__`` is host-side format access (torch CPU indexing, for code that iterates the dataset like the
reference's); the training path is ``loader``.  PNG decoding needs PIL and happens only in ``from_folder``.
"""
from __future__ import annotations

import csv
import os
from typing import Dict, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from .ops import MstkError

CROP = 128          # dataset.py:51: img.crop((0, 0, 128, 128))
IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")   # torchvision's list (dataset.py:119-200)


class PackedSpectrograms:
    def __init__(self, images: Dict[str, torch.Tensor]):
        if not images:
            raise MstkError("PackedSpectrograms: no labels")
        self.labels = sorted(images)
        shapes = {tuple(images[k].shape[1:]) for k in self.labels}
        if len(shapes) != 1 or any(images[k].dtype != torch.uint8 or images[k].dim() != 3 for k in self.labels):
            raise MstkError("PackedSpectrograms: every label needs a uint8 [n, H, W] tensor of one common image size")
        self.images = {k: images[k].contiguous() for k in self.labels}
        self.height, self.width = next(iter(shapes))
        if self.height < CROP or self.width < CROP:
            raise MstkError(f"PackedSpectrograms: images must be at least {CROP}x{CROP} (dataset.py:51), got {self.height}x{self.width}")
        self.offsets, off = {}, 0
        for k in self.labels:
            self.offsets[k] = off
            off += self.images[k].shape[0]
        self.total = off
        self._pool = {}

    # -- construction / persistence ------------------------------------------------------------
    @classmethod
    def from_folder(cls, root_folder: str) -> "PackedSpectrograms":
        """<root>/<label>/*.png in the reference's order: labels sorted (dataset.py:229), files sorted by name as
        torchvision's make_dataset does (dataset.py:119-200).  Grayscale value = the PNG's L channel."""
        from PIL import Image
        out = {}
        for label in sorted(os.listdir(root_folder)):
            path = os.path.join(root_folder, label)
            if not os.path.isdir(path):
                continue
            files = sorted(f for f in os.listdir(path) if f.lower().endswith(IMG_EXTENSIONS))
            if not files:
                continue
            arrs = [np.asarray(Image.open(os.path.join(path, f)).convert("RGB").convert("L"), dtype=np.uint8) for f in files]
            out[label] = torch.from_numpy(np.stack(arrs))
        return cls(out)

    def save(self, path: str) -> None:
        torch.save({"format": "mstk-packed-spectrograms-v1", "images": self.images}, path)

    @classmethod
    def load(cls, path: str) -> "PackedSpectrograms":
        blob = torch.load(path, map_location="cpu")
        if not isinstance(blob, dict) or blob.get("format") != "mstk-packed-spectrograms-v1":
            raise MstkError(f"{path}: not a packed spectrogram file")
        return cls(blob["images"])

    # -- access ---------------------------------------------------------------------------------
    def __len__(self) -> int:
        return self.total

    def count(self, label: str) -> int:
        return self.images[label].shape[0]

    def item(self, label: str, idx: int) -> torch.Tensor:
        """float32 [1, 128, 128] in [0, 1]: crop(0, 0, 128, 128), grayscale, ToTensor (dataset.py:51-53)"""
        return self.images[label][idx, :CROP, :CROP].to(torch.float32).div_(255.0).unsqueeze(0)

    def flat_index(self, label: str, idx: int) -> int:
        n = self.count(label)
        if not 0 <= idx < n:
            raise IndexError(f"index {idx} outside label '{label}' with {n} images")
        return self.offsets[label] + idx

    def pool(self, device) -> torch.Tensor:
        """all images as one uint8 [total, H, W] tensor on ``device`` (16.6 KB per 128x130 image: the reference's
        whole 4-instrument dataset is < 1 GB of HBM)"""
        device = torch.device(device)
        p = self._pool.get(device)
        if p is None:
            if device.type != "cuda":
                raise MstkError("PackedSpectrograms.pool: the batch path runs on sm_100a kernels only")
            p = torch.cat([self.images[k] for k in self.labels]).to(device)
            self._pool[device] = p
        return p


class _PairLoader:
    """iterable of ((content [B,1,128,128], labels), (style [B,1,128,128], labels)) batches, the element layout
    LDMTrainer.train_epoch unpacks (train.py:220-227); one gather kernel per side and batch"""

    def __init__(self, ds: "SpectrogramPairDataset", batch_size: int, device, shuffle: bool, drop_last: bool, generator):
        self.ds, self.batch_size, self.device, self.shuffle, self.drop_last, self.gen = ds, int(batch_size), torch.device(device), shuffle, drop_last, generator
        self.pool = ds.store.pool(self.device)
        self.idx1 = torch.tensor([ds.store.flat_index(a, i) for a, i, _, _ in ds.pairs], dtype=torch.int64, device=self.device)
        self.idx2 = torch.tensor([ds.store.flat_index(b, j) for _, _, b, j in ds.pairs], dtype=torch.int64, device=self.device)

    def __len__(self) -> int:
        n = len(self.ds)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self) -> Iterator:
        n = len(self.ds)
        order = torch.randperm(n, generator=self.gen) if self.shuffle else torch.arange(n)
        order_d = order.to(self.device)
        for b in range(len(self)):
            sel = order[b * self.batch_size:(b + 1) * self.batch_size]
            sel_d = order_d[b * self.batch_size:(b + 1) * self.batch_size]
            x = ops.u8_gather_crop(self.pool, self.idx1[sel_d].contiguous(), CROP, CROP)
            s = ops.u8_gather_crop(self.pool, self.idx2[sel_d].contiguous(), CROP, CROP)
            l1 = [self.ds.pairs[i][0] for i in sel.tolist()]
            l2 = [self.ds.pairs[i][2] for i in sel.tolist()]
            yield (x, l1), (s, l2)


class SpectrogramPairDataset(torch.utils.data.Dataset):
    """models/dataset.py:206-260.  ``root_folder``: the PNG root (packed on construction), a packed ``.pt`` file, or a
    PackedSpectrograms; ``pairing_file``: CSV rows ``label1,idx1,label2,idx2``.  ``transform`` is accepted for
    signature compatibility and must be None: the crop / grayscale / ToTensor chain of dataset.py:243-260 is built in."""

    def __init__(self, root_folder, pairing_file, transform=None):
        if transform is not None:
            raise MstkError("SpectrogramPairDataset: custom transforms are not supported (the reference's crop(0,0,128,128) / "
                            "Grayscale / ToTensor chain is fused into the batch kernel)")
        self.root_folder, self.pairing_file = root_folder, pairing_file
        if isinstance(root_folder, PackedSpectrograms):
            self.store = root_folder
        elif os.path.isdir(root_folder):
            self.store = PackedSpectrograms.from_folder(root_folder)
        else:
            self.store = PackedSpectrograms.load(root_folder)
        self.pairs: List[Tuple[str, int, str, int]] = []
        with open(pairing_file, "r") as f:
            for row in csv.reader(f):
                self.pairs.append((row[0], int(row[1]), row[2], int(row[3])))

    def __len__(self):
        return len(self.pairs)

    def __getitem__(self, index):
        label1, idx1, label2, idx2 = self.pairs[index]
        return (self.store.item(label1, idx1), label1), (self.store.item(label2, idx2), label2)

    def loader(self, batch_size: int, device="cuda", shuffle: bool = True, drop_last: bool = False, generator=None) -> _PairLoader:
        return _PairLoader(self, batch_size, device, shuffle, drop_last, generator)

    @classmethod
    def generate_pairings(cls, root_folder, output_file_path="spectrogram_pair_dataset_pairings.csv", num_pairs=15000):
        """dataset.py:263-303: RandomState(42); two distinct labels, then one index in each."""
        store = root_folder if isinstance(root_folder, PackedSpectrograms) else (
            PackedSpectrograms.from_folder(root_folder) if os.path.isdir(root_folder) else PackedSpectrograms.load(root_folder))
        labels = store.labels
        if len(labels) < 2:
            raise ValueError("Need at least two classes to form pairs.")
        rng = np.random.RandomState(42)
        pairs = []
        for _ in range(num_pairs):
            label1, label2 = rng.choice(labels, size=2, replace=False)
            idx1 = rng.randint(0, store.count(label1))
            idx2 = rng.randint(0, store.count(label2))
            pairs.append((label1, idx1, label2, idx2))
        with open(output_file_path, "w", newline="") as f:
            w = csv.writer(f)
            for p in pairs:
                w.writerow(p)
        print(f"Pairings saved to {output_file_path}")


class SpectrogramDataset(torch.utils.data.Dataset):
    """models/dataset.py:28-53: every image of every label in ImageFolder order, element (img [1,128,128], class index)."""

    def __init__(self, config):
        src = config["processed_spectograms_dataset_folderpath"]
        self.store = src if isinstance(src, PackedSpectrograms) else (
            PackedSpectrograms.from_folder(src) if os.path.isdir(src) else PackedSpectrograms.load(src))
        self.index = [(k, i, c) for c, k in enumerate(self.store.labels) for i in range(self.store.count(k))]

    def __len__(self):
        return len(self.index)

    def __getitem__(self, idx):
        k, i, c = self.index[idx]
        return self.store.item(k, i), c

    def loader(self, batch_size: int, indices: Optional[Sequence[int]] = None, device="cuda", shuffle: bool = False, generator=None):
        """batches (img [B,1,128,128] on the device, class indices) for train_autoencoder (train.py:62-66)"""
        ds, pool = self, self.store.pool(device)
        base = torch.tensor([self.store.flat_index(k, i) for k, i, _ in self.index], dtype=torch.int64)
        cls_ = torch.tensor([c for _, _, c in self.index], dtype=torch.int64)
        sub = torch.arange(len(self.index)) if indices is None else torch.as_tensor(list(indices), dtype=torch.int64)

        class _L:
            def __len__(self_l):
                return (len(sub) + batch_size - 1) // batch_size

            def __iter__(self_l):
                order = sub[torch.randperm(len(sub), generator=generator)] if shuffle else sub
                for b in range(len(self_l)):
                    sel = order[b * batch_size:(b + 1) * batch_size]
                    yield ops.u8_gather_crop(pool, base[sel].to(pool.device), CROP, CROP), cls_[sel]

        return _L()


def prepare_dataset(config, device="cuda"):
    """dataset.py:306-316: 80 / 20 random split, train shuffled; returns (train_loader, test_loader) of device batches"""
    ds = SpectrogramDataset(config)
    n_train = int(0.8 * len(ds))
    perm = torch.randperm(len(ds)).tolist()
    return (ds.loader(config["batch_size"], perm[:n_train], device=device, shuffle=True),
            ds.loader(config["batch_size"], perm[n_train:], device=device, shuffle=False))
