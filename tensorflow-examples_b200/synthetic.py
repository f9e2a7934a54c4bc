"""Synthetic data of the examples' shapes, and the initial weights every arm of a comparison shares.

Replaces the MNIST loader + feed_dict path (distributed/distributed.py:52-55,146) with a pool that
lives on the device: x ~ U[0,1) float32 [N,784] (MNIST's value range), labels uniform over 10 classes
as one-hot float32 [N,10] (the reference's y_ layout, distributed.py:80).  Generators are fixed
 so the CPU oracle and the GPU path see identical bytes.
"""
import numpy as np

POOL_SEED = 1234
WEIGHT_SEED = 2  # distributed.py:83 tf.set_random_seed(2); TF's own Philox stream is not reproduced


def make_pool(rows, in_dim=784, classes=10, seed=POOL_SEED):
    rng = np.random.default_rng(seed)
    x = rng.random((rows, in_dim), dtype=np.float32)
    labels = rng.integers(0, classes, rows)
    y = np.zeros((rows, classes), dtype=np.float32)
    y[np.arange(rows), labels] = 1.0
    return x, y


def make_weights(dims, scaled, seed=WEIGHT_SEED):
    """W ~ N(0,1) as tf.random_normal (distributed.py:85-86) when scaled is False; N(0, 1/fan_in)
    when True (needed where stddev 1 would underflow the reference's un-fused log(softmax))."""
    rng = np.random.default_rng(seed)
    Ws, bs = [], []
    for i in range(len(dims) - 1):
        W = rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32)
        if scaled:
            W = (W / np.float32(np.sqrt(dims[i]))).astype(np.float32)
        Ws.append(W)
        bs.append(np.zeros(dims[i + 1], dtype=np.float32))  # distributed.py:90-91
    return Ws, bs
