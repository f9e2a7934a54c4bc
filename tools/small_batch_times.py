#!/usr/bin/env python
"""us/step of the reference's small-batch configurations (distributed.py: 784-100-10 at 100 rows, distributed.py:46) and of
784->10, with the whole loop in one launch (cluster kernel) and with one launch sequence per step."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

dev = torch.device("cuda", 0)
hx, hy = synthetic.make_pool(55000)
x, y = torch.from_numpy(hx).to(dev), torch.from_numpy(hy).to(dev)
cases = [([784, 10], 0, True, [16, 64, 100, 256]), ([784, 100, 10], 1, False, [16, 32, 64, 100, 128, 200])]
for dims, act, scaled, batches in cases:
    for B in batches:
        for persistent in (True, False):
            tr = b2.DenseTrainer(dims, act, b2.LOSS_SOFTMAX_CE, max_local_batch=B)
            tr.set_params(*synthetic.make_weights(dims, scaled))
            tr.persistent_steps(persistent)
            n = 550 if persistent else 200
            tr.train_steps(x, y, B, 50)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tr.train_steps(x, y, B, n)
            e1.record()
            torch.cuda.synchronize()
            print("%-14s batch %4d %-10s %8.2f us/step" % ("-".join(map(str, dims)), B, "one-launch" if persistent else "per-step",
                                                          1e3 * e0.elapsed_time(e1) / n), flush=True)
            tr.close()
