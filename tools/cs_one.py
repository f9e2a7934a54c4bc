#!/usr/bin/env python
"""One call of the cluster kernel for ncu: 784-100-10 at 100 rows (or argv: model batch steps)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "c2"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 100
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 200
dims, act, scaled = ([784, 100, 10], 1, False) if model == "c2" else ([784, 10], 0, True)
dev = torch.device("cuda", 0)
hx, hy = synthetic.make_pool(55000)
x, y = torch.from_numpy(hx).to(dev), torch.from_numpy(hy).to(dev)
tr = b2.DenseTrainer(dims, act, b2.LOSS_SOFTMAX_CE, max_local_batch=B)
tr.set_params(*synthetic.make_weights(dims, scaled))
lt, _ = tr.train_steps(x, y, B, steps)
torch.cuda.synchronize()
print("loss", float(lt[0]), float(lt[steps - 1]))
tr.close()
