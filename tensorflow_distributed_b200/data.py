"""Dataset input.

Stand-in for ``tensorflow.examples.tutorials.mnist.input_data`` as the reference
uses it: ``read_data_sets(dir, one_hot=True)`` giving ``.train`` /
``.validation`` / ``.test`` splits with ``next_batch(n)`` returning
``([n,784] float32 in [0,1], [n,10] one-hot float32)``
(``/root/reference/mnist_python_m.py:46,133,291,313``;
``/root/reference/mnist_single.py:14-15,110,127``).

There is no network on the box, so the default source is a *synthetic*
MNIST-shaped dataset (class templates + noise: learnable, deterministic).  If
the four idx(.gz) files are present in ``data_dir`` they are parsed instead.
``download_only`` (m:52-54,133-135) becomes "prepare/verify the data, then exit".
"""
from __future__ import annotations

import gzip
import os
import struct
from typing import Optional, Tuple

import numpy as np

from .models.spec import N_CLASSES, N_INPUT

_FILES = {
    "train_images": "train-images-idx3-ubyte",
    "train_labels": "train-labels-idx1-ubyte",
    "test_images": "t10k-images-idx3-ubyte",
    "test_labels": "t10k-labels-idx1-ubyte",
}
VALIDATION_SIZE = 5000  # the tutorial helper's split: 55000 / 5000 / 10000


def _open(path: str):
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return open(path, "rb")


def read_idx_images(path: str) -> np.ndarray:
    with _open(path) as f:
        magic, n, rows, cols = struct.unpack(">IIII", f.read(16))
        if magic != 2051:
            raise ValueError("bad magic %d in image file %s" % (magic, path))
        data = np.frombuffer(f.read(n * rows * cols), dtype=np.uint8)
    return data.reshape(n, rows * cols).astype(np.float32) / 255.0


def read_idx_labels(path: str) -> np.ndarray:
    with _open(path) as f:
        magic, n = struct.unpack(">II", f.read(8))
        if magic != 2049:
            raise ValueError("bad magic %d in label file %s" % (magic, path))
        data = np.frombuffer(f.read(n), dtype=np.uint8)
    return data.astype(np.int64)


def write_idx(path_images: str, path_labels: str, images_u8: np.ndarray, labels: np.ndarray, gz: bool = True) -> None:
    """Writer used by tests (round-trips the reader)."""
    n = images_u8.shape[0]
    op = (lambda p: gzip.open(p + ".gz", "wb")) if gz else (lambda p: open(p, "wb"))
    with op(path_images) as f:
        f.write(struct.pack(">IIII", 2051, n, 28, 28))
        f.write(images_u8.astype(np.uint8).tobytes())
    with op(path_labels) as f:
        f.write(struct.pack(">II", 2049, n))
        f.write(labels.astype(np.uint8).tobytes())


def one_hot(labels: np.ndarray, n: int = N_CLASSES) -> np.ndarray:
    out = np.zeros((labels.shape[0], n), dtype=np.float32)
    out[np.arange(labels.shape[0]), labels] = 1.0
    return out


class DataSet:
    """In-memory split with the tutorial helper's epoch/shuffle behaviour."""

    def __init__(self, images: np.ndarray, labels_onehot: np.ndarray, seed: int = 0):
        assert images.shape[0] == labels_onehot.shape[0]
        self._images, self._labels = images, labels_onehot
        self._rng = np.random.RandomState(seed)
        self._index = 0
        self.epochs_completed = 0

    @property
    def num_examples(self) -> int:
        return self._images.shape[0]

    @property
    def images(self) -> np.ndarray:
        return self._images

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    def next_batch(self, batch_size: int) -> Tuple[np.ndarray, np.ndarray]:
        n = self.num_examples
        if batch_size > n:
            raise ValueError("batch larger than the split")
        if self._index + batch_size > n:
            self.epochs_completed += 1
            perm = self._rng.permutation(n)
            self._images, self._labels = self._images[perm], self._labels[perm]
            self._index = 0
        s = self._index
        self._index += batch_size
        return self._images[s:s + batch_size], self._labels[s:s + batch_size]


class SyntheticSplit:
    """Endless MNIST-shaped stream: x = clip(template[class] + noise), generated per batch."""

    def __init__(self, templates: np.ndarray, seed: int, num_examples: int, noise: float = 0.35):
        self._templates = templates
        self._rng = np.random.RandomState(seed)
        self._noise = noise
        self.num_examples = num_examples
        self.epochs_completed = 0
        self._served = 0

    def next_batch(self, batch_size: int) -> Tuple[np.ndarray, np.ndarray]:
        labels = self._rng.randint(0, N_CLASSES, size=batch_size)
        x = self._templates[labels] + self._noise * self._rng.standard_normal((batch_size, N_INPUT)).astype(np.float32)
        np.clip(x, 0.0, 1.0, out=x)
        self._served += batch_size
        self.epochs_completed = self._served // self.num_examples
        return x.astype(np.float32), one_hot(labels)


class DataSets:
    def __init__(self, train, validation, test, source: str):
        self.train, self.validation, self.test, self.source = train, validation, test, source


def synthetic_templates(seed: int = 1234) -> np.ndarray:
    rng = np.random.RandomState(seed)
    t = np.zeros((N_CLASSES, 28, 28), dtype=np.float32)
    for c in range(N_CLASSES):
        # a few random strokes per class: sparse, MNIST-like intensity statistics
        for _ in range(6):
            r0, c0 = rng.randint(4, 24, size=2)
            dr, dc = rng.randint(-1, 2, size=2)
            for k in range(10):
                r, cc = r0 + dr * k, c0 + dc * k
                if 0 <= r < 28 and 0 <= cc < 28:
                    t[c, r, cc] = 1.0
    return t.reshape(N_CLASSES, N_INPUT)


def idx_files_present(data_dir: str) -> bool:
    return all(os.path.exists(os.path.join(data_dir, f)) or os.path.exists(os.path.join(data_dir, f + ".gz"))
               for f in _FILES.values())


def read_data_sets(data_dir: str, one_hot_labels: bool = True, synthetic: Optional[bool] = None,
                   seed: int = 0) -> DataSets:
    """``input_data.read_data_sets`` analogue.  ``synthetic=None`` => use idx files if present."""
    assert one_hot_labels, "the reference always asks for one_hot=True"
    have = idx_files_present(data_dir) if data_dir else False
    if synthetic is False and not have:
        raise FileNotFoundError(
            "MNIST idx files not found in %r and there is no network to download them; "
            "use --synthetic" % data_dir)
    if have and synthetic is not True:
        tr_x = read_idx_images(os.path.join(data_dir, _FILES["train_images"]))
        tr_y = one_hot(read_idx_labels(os.path.join(data_dir, _FILES["train_labels"])))
        te_x = read_idx_images(os.path.join(data_dir, _FILES["test_images"]))
        te_y = one_hot(read_idx_labels(os.path.join(data_dir, _FILES["test_labels"])))
        nv = min(VALIDATION_SIZE, tr_x.shape[0] // 2)
        return DataSets(DataSet(tr_x[nv:], tr_y[nv:], seed), DataSet(tr_x[:nv], tr_y[:nv], seed + 1),
                        DataSet(te_x, te_y, seed + 2), "idx:" + data_dir)
    t = synthetic_templates()
    return DataSets(SyntheticSplit(t, seed * 3 + 11, 55000), SyntheticSplit(t, seed * 3 + 12, 5000),
                    SyntheticSplit(t, seed * 3 + 13, 10000), "synthetic")
