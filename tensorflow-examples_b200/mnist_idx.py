"""MNIST IDX files -> the arrays `input_data.read_data_sets(..., one_hot=True)` hands the reference
(distributed/distributed.py:52-55): float32 images scaled by 1/255, one-hot float32 labels, the first
5000 training rows held out as validation (55000 / 5000 / 10000 split).  There is no downloader: the
four files must already be in the directory (the box has no network)."""
import gzip
import os
import struct

import numpy as np

FILES = {"train_x": "train-images-idx3-ubyte", "train_y": "train-labels-idx1-ubyte",
         "test_x": "t10k-images-idx3-ubyte", "test_y": "t10k-labels-idx1-ubyte"}


def _open(path):
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return open(path, "rb")


def read_idx(path):
    with _open(path) as f:
        magic = struct.unpack(">I", f.read(4))[0]
        ndim = magic & 0xFF
        if (magic >> 8) != 0x08:
            raise ValueError("%s: not an unsigned-byte IDX file (magic 0x%08x)" % (path, magic))
        shape = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        data = np.frombuffer(f.read(), dtype=np.uint8)
    return data.reshape(shape)


def available(data_dir):
    return all(os.path.exists(os.path.join(data_dir, n)) or os.path.exists(os.path.join(data_dir, n + ".gz"))
               for n in FILES.values())


def load(data_dir, validation_size=5000, classes=10):
    def images(name):
        a = read_idx(os.path.join(data_dir, FILES[name]))
        return (a.reshape(a.shape[0], -1).astype(np.float32) * np.float32(1.0 / 255.0))

    def labels(name):
        a = read_idx(os.path.join(data_dir, FILES[name]))
        y = np.zeros((a.shape[0], classes), dtype=np.float32)
        y[np.arange(a.shape[0]), a] = 1.0
        return y
    tx, ty = images("train_x"), labels("train_y")
    return {"train": (tx[validation_size:], ty[validation_size:]),
            "validation": (tx[:validation_size], ty[:validation_size]),
            "test": (images("test_x"), labels("test_y"))}
