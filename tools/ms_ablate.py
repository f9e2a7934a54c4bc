#!/usr/bin/env python
"""What each part of the multi-step kernel's step boundary costs: per-step time with parts left out (B2_MS_DEBUG mask:
1 no partial loads, 2 no apply, 4 no W reload, 8 no grid barriers, 16 no partial write).  Results are wrong on purpose."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

dev = torch.device("cuda", 0)
hx, hy = synthetic.make_pool(1 << 17)
x = torch.from_numpy(hx).to(dev).repeat(4, 1).contiguous()
y = torch.from_numpy(hy).to(dev).repeat(4, 1).contiguous()
for B in (256, 4096, 65536):
    for mask in (0, 1, 2, 3, 4, 8, 16, 31):
        os.environ["B2_MS_DEBUG"] = str(mask)
        tr = b2.DenseTrainer([784, 10], b2.ACT_IDENTITY, b2.LOSS_SOFTMAX_CE, max_local_batch=B)
        tr.set_params(*synthetic.make_weights([784, 10], True))
        tr.multistep_fused(True)
        tr.train_steps(x, y, B, 20)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tr.train_steps(x, y, B, 200)
        e1.record()
        torch.cuda.synchronize()
        print("batch %6d mask %2d  %8.2f us/step" % (B, mask, 1e3 * e0.elapsed_time(e1) / 200), flush=True)
        tr.close()
