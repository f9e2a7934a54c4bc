#!/usr/bin/env python
"""Timeline of the multi-step kernel's step boundary (B2_MS_STAMPS=1 makes the launcher print CTA 0's %globaltimer stamps)
and per-step times of both paths at a few batch sizes.  usage: python tools/ms_stamps.py [batches...]"""
import os
import sys

os.environ["B2_MS_STAMPS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

batches = [int(a) for a in sys.argv[1:]] or [256, 4096, 65536]
dev = torch.device("cuda", 0)
hx, hy = synthetic.make_pool(1 << 17)
x = torch.from_numpy(hx).to(dev).repeat(4, 1).contiguous()
y = torch.from_numpy(hy).to(dev).repeat(4, 1).contiguous()
for B in batches:
    for persistent in (True, False):
        tr = b2.DenseTrainer([784, 10], b2.ACT_IDENTITY, b2.LOSS_SOFTMAX_CE, max_local_batch=B)
        tr.set_params(*synthetic.make_weights([784, 10], True))
        tr.multistep_fused(persistent)
        os.environ["B2_MS_STAMPS"] = "0"
        tr.train_steps(x, y, B, 20)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tr.train_steps(x, y, B, 200)
        e1.record()
        torch.cuda.synchronize()
        print("batch %6d %-10s %8.2f us/step" % (B, "one-launch" if persistent else "per-step", 1e3 * e0.elapsed_time(e1) / 200), flush=True)
        if persistent:
            os.environ["B2_MS_STAMPS"] = "1"
            tr.train_steps(x, y, B, 40)
            torch.cuda.synchronize()
        tr.close()
