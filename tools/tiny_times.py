#!/usr/bin/env python
"""ms per whole training run of TINY models in the one-warp kernel (tiny_steps.cu): simple.py as shipped ([1,1], four points,
1005 steps, simple.py:25-35) and two small stacks, each as ONE launch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402

dev = torch.device("cuda", 0)
rng = np.random.RandomState(3)
cases = [("simple.py [1,1] 4 rows SSE", [1, 1], b2.ACT_IDENTITY, b2.LOSS_SSE, 4, 4, 1005, 0.01),
         ("[1,1] 4 of 64 rows", [1, 1], b2.ACT_IDENTITY, b2.LOSS_SSE, 4, 64, 1005, 0.01),
         ("[8,16,4] 32 rows softmax", [8, 16, 4], b2.ACT_SIGMOID, b2.LOSS_SOFTMAX_CE, 32, 256, 1005, 0.05),
         ("[16,16,16,16,10] 32 rows", [16, 16, 16, 16, 10], b2.ACT_SIGMOID, b2.LOSS_SOFTMAX_CE, 32, 256, 1005, 0.05)]
for name, dims, act, loss, B, pool, steps, lr in cases:
    if dims == [1, 1]:
        hx = np.tile(np.array([[1.0], [2.0], [3.0], [4.0]], np.float32), (pool // 4, 1))
        hy = np.tile(np.array([[0.0], [-1.0], [-2.0], [-3.0]], np.float32), (pool // 4, 1))
        w = [(np.array([[0.3]], np.float32), np.array([-0.3], np.float32))]
    else:
        hx = rng.rand(pool, dims[0]).astype(np.float32)
        hy = np.eye(dims[-1], dtype=np.float32)[rng.randint(0, dims[-1], pool)]
        w = [(rng.randn(dims[i], dims[i + 1]).astype(np.float32) * 0.3, np.zeros(dims[i + 1], np.float32)) for i in range(len(dims) - 1)]
    x, y = torch.from_numpy(hx).to(dev), torch.from_numpy(hy).to(dev)
    tr = b2.DenseTrainer(dims, act, loss, max_local_batch=B)
    tr.set_params([a for a, _ in w], [b for _, b in w])
    tr.train_steps(x, y, B, 50, lr=lr)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tr.train_steps(x, y, B, steps, lr=lr)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("%-30s %5d steps  %7.3f ms  (%.2f us/step, %d launches)" % (name, steps, best, 1e3 * best / steps, tr.launch_count), flush=True)
    tr.close()
