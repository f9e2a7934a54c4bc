#!/usr/bin/env python
"""Small driver for compute-sanitizer: a few training steps of the 784 -> 10 model on a ragged batch (1000 rows), through the
per-step kernels (fused step + reduction) and through the multi-step kernel, plus simple.py's tiny model.
usage: compute-sanitizer --tool memcheck|racecheck python tools/sanitize_case.py [per-step|multi|tiny]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "per-step"
dev = torch.device("cuda", 0)
if which == "tiny":
    tr = b2.DenseTrainer([1, 1], b2.ACT_IDENTITY, b2.LOSS_SSE, max_local_batch=4)
    tr.set_params([np.array([[.3]], np.float32)], [np.array([-.3], np.float32)])
    x = torch.tensor([[1], [2], [3], [4]], dtype=torch.float32, device=dev)
    y = torch.tensor([[0], [-1], [-2], [-3]], dtype=torch.float32, device=dev)
    lt, _ = tr.train_steps(x, y, 4, 50, lr=0.01)
else:
    batch, steps = 1000, 3
    xh, yh = synthetic.make_pool(batch * steps)
    tr = b2.DenseTrainer([784, 10], b2.ACT_IDENTITY, b2.LOSS_SOFTMAX_CE, max_local_batch=batch)
    tr.set_params(*synthetic.make_weights([784, 10], True))
    tr.multistep_fused(which == "multi")
    lt, _ = tr.train_steps(torch.from_numpy(xh).to(dev), torch.from_numpy(yh).to(dev), batch, steps, lr=0.001)
torch.cuda.synchronize()
print(which, "losses", lt.cpu().numpy()[:3], "launches", tr.launch_count)
tr.close()
