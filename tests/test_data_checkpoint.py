"""Dataset loader (C14), download_only analogue (C3), checkpoint names."""
import os

import numpy as np
import pytest
import torch

from tensorflow_distributed_b200 import checkpoint, data
from tensorflow_distributed_b200.models.spec import cnn_spec
from tensorflow_distributed_b200.ops.oracle import TFAdam, init_params


def test_synthetic_shapes_and_determinism():
    d1 = data.read_data_sets("", synthetic=True, seed=3)
    d2 = data.read_data_sets("", synthetic=True, seed=3)
    x, y = d1.train.next_batch(128)
    x2, y2 = d2.train.next_batch(128)
    assert x.shape == (128, 784) and y.shape == (128, 10) and x.dtype == np.float32
    assert np.array_equal(x, x2) and np.array_equal(y, y2)
    assert np.all(y.sum(1) == 1) and x.min() >= 0 and x.max() <= 1
    assert d1.validation.next_batch(1000)[0].shape == (1000, 784)


def test_idx_reader_roundtrip(tmp_path):
    rng = np.random.RandomState(0)
    for split, n in (("train", 600), ("t10k", 100)):
        data.write_idx(str(tmp_path / ("%s-images-idx3-ubyte" % split)), str(tmp_path / ("%s-labels-idx1-ubyte" % split)),
                       rng.randint(0, 256, size=(n, 784)), rng.randint(0, 10, size=n))
    ds = data.read_data_sets(str(tmp_path), synthetic=None)
    assert ds.source.startswith("idx:")
    assert ds.train.num_examples == 300 and ds.validation.num_examples == 300 and ds.test.num_examples == 100
    x, y = ds.train.next_batch(256)
    assert x.shape == (256, 784) and y.shape == (256, 10)
    ds.train.next_batch(256)  # wraps an epoch
    assert ds.train.epochs_completed == 1
    with pytest.raises(FileNotFoundError):
        data.read_data_sets(str(tmp_path / "none"), synthetic=False)


def test_checkpoint_reference_names_and_roundtrip(tmp_path):
    spec = cnn_spec()
    p = init_params(spec, 0)
    adam = TFAdam(0.01)
    adam.apply(p, {k: torch.ones_like(v) for k, v in p.items()})
    path = checkpoint.save(str(tmp_path), spec, 7, p, adam.state_dict())
    assert os.path.basename(path) == "model.ckpt-7.pt"
    assert checkpoint.latest_checkpoint(str(tmp_path)) == path
    blob = checkpoint.load(path, spec)
    assert blob["global_step"] == 7
    assert set(blob["variables"]) == {"weights/wc1", "weights/wc2", "weights/wd1", "weights/out",
                                      "biases/bc1", "biases/bc2", "biases/bd1", "biases/out"}
    assert tuple(blob["variables"]["weights/wc2"].shape) == (5, 5, 32, 64)  # HWIO
    assert tuple(blob["variables"]["weights/wd1"].shape) == (3136, 1024)
    assert blob["tf_alias"]["weights/wd1"] == "Variable_2" and blob["adam"]["t"] == 1
    assert torch.equal(blob["variables"]["weights/out"], p["weights/out"])
    checkpoint.save(str(tmp_path), spec, 9, p)
    assert checkpoint.latest_checkpoint(str(tmp_path)).endswith("model.ckpt-9.pt")
