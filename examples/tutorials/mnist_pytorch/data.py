"""MNIST loader with a synthetic fallback (no network in the build image)."""
import gzip
import os
import pathlib
import struct

import numpy as np
import torch


def _read_idx(path: pathlib.Path) -> np.ndarray:
    with gzip.open(path, "rb") as f:
        _, _, dims = struct.unpack(">HBB", f.read(4))
        shape = struct.unpack(">" + "I" * dims, f.read(4 * dims))
        return np.frombuffer(f.read(), dtype=np.uint8).reshape(shape)


def get_dataset(train: bool, n_synthetic: int = 4096) -> torch.utils.data.Dataset:
    root = pathlib.Path(os.environ.get("DET_DATA_DIR", "/data")) / "MNIST" / "raw"
    prefix = "train" if train else "t10k"
    imgs, labels = root / f"{prefix}-images-idx3-ubyte.gz", root / f"{prefix}-labels-idx1-ubyte.gz"
    if imgs.exists() and labels.exists():
        x = torch.from_numpy(_read_idx(imgs).copy()).float().div_(255).sub_(0.1307).div_(0.3081).unsqueeze(1)
        y = torch.from_numpy(_read_idx(labels).copy()).long()
    else:
        g = torch.Generator().manual_seed(0 if train else 1)
        y = torch.randint(0, 10, (n_synthetic,), generator=g)
        # class-dependent blobs so that the loss can actually go down
        x = torch.randn(n_synthetic, 1, 28, 28, generator=g) * 0.3
        for c in range(10):
            x[y == c, :, 2 * c: 2 * c + 8, 2 * c: 2 * c + 8] += 1.5
    return torch.utils.data.TensorDataset(x, y)
