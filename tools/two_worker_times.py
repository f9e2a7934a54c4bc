#!/usr/bin/env python
"""distributed.py's canonical set-up -- several workers, 100 rows each (distributed.py:11-13,46) -- us/step with the whole
loop in one launch (cluster kernel + in-kernel NVLink exchange) and with one launch sequence per step.
usage: torchrun --nproc-per-node N tools/two_worker_times.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo")
dev = torch.device("cuda", rank)
x, y = synthetic.make_pool(55000)
xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
for dims, act, scaled in (([784, 100, 10], 1, False), ([784, 10], 0, True)):
    for persistent in (True, False):
        tr = b2.DenseTrainer(dims, act, b2.LOSS_SOFTMAX_CE, max_local_batch=100, device=rank)
        tr.set_params(*synthetic.make_weights(dims, scaled))
        uid = [b2.DenseTrainer.new_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        tr.init_comm(rank, world, uid[0])
        hs = [None] * world
        dist.all_gather_object(hs, tr.peer_export())
        tr.peer_import(hs)
        tr.persistent_steps(persistent)
        tr.train_steps(xd, yd, 100, 50)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 55000 // (100 * world) // 2
        e0.record()
        lt, ct = tr.train_steps(xd, yd, 100, n, first_row=0)
        e1.record()
        torch.cuda.synchronize()
        if rank == 0:
            print("%-12s %d workers x 100 rows %-10s %7.2f us/step  loss %.7f" % ("-".join(map(str, dims)), world,
                  "one-launch" if persistent else "per-step", 1e3 * e0.elapsed_time(e1) / n, float(lt[n - 1])), flush=True)
        dist.barrier()
        tr.close()
dist.destroy_process_group()
