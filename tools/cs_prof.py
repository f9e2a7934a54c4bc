#!/usr/bin/env python
"""Per-phase SM clocks of the cluster kernel (B2_CS_PROF=1) for a few (model, batch) points."""
import os
import sys

os.environ["B2_CS_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

dev = torch.device("cuda", 0)
hx, hy = synthetic.make_pool(55000)
x, y = torch.from_numpy(hx).to(dev), torch.from_numpy(hy).to(dev)
for dims, act, scaled, B in [([784, 10], 0, True, 16), ([784, 10], 0, True, 256), ([784, 100, 10], 1, False, 16),
                             ([784, 100, 10], 1, False, 100)]:
    tr = b2.DenseTrainer(dims, act, b2.LOSS_SOFTMAX_CE, max_local_batch=B)
    tr.set_params(*synthetic.make_weights(dims, scaled))
    tr.train_steps(x, y, B, 200)
    torch.cuda.synchronize()
    tr.close()
