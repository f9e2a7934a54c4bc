"""B200-native training step for the two scripts of manigoswami/tensorflow-examples.

Only what the hot path needs lives here:
  csrc/        hand-written sm_100a CUDA kernels + the C ABI (include/b2mlp.h) -> csrc/libb2mlp.so
  build.py     the nvcc recipe
  _ffi.py      ctypes binding of the C ABI
  session.py   host-side mirror of the reference's `sess.run(train_op, feed_dict)` interface
  synthetic.py device-resident synthetic batches of the examples' shapes (replaces feed_dict input)
  cluster.py   ps/worker flags -> rank / world size (replaces ClusterSpec / Server / replica_device_setter)
There is no CPU fallback: every compute call fails loudly without the CUDA library and a B200.
"""
from . import build  # noqa: F401
from ._ffi import B2Error, lib_path, load_library  # noqa: F401
from .session import (ACT_IDENTITY, ACT_SIGMOID, LOSS_SOFTMAX_CE, LOSS_SSE, DenseTrainer)  # noqa: F401

__all__ = ["B2Error", "DenseTrainer", "load_library", "lib_path", "build",
           "ACT_IDENTITY", "ACT_SIGMOID", "LOSS_SSE", "LOSS_SOFTMAX_CE"]
