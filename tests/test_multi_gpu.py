"""Data-parallel path on real GPUs (needs >= 2; run under `gpurun --gpus 2`): N ranks with an NCCL
all-reduce inside the step must reproduce the single-process trace at the same global batch, and the
reference's 1 ps + 2 workers command lines (distributed.py:11-13) must run as shipped."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, json
sys.path.insert(0, %(root)r)
import numpy as np, torch
import torch.distributed as dist
import tensorflow_examples_b200 as b2
from tensorflow_examples_b200 import synthetic
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo")
dims, act, scaled, lb, steps = %(dims)r, %(act)d, %(scaled)r, %(lb)d, %(steps)d
gb = lb * world
x, y = synthetic.make_pool(gb * steps)
Ws, bs = synthetic.make_weights(dims, scaled)
tr = b2.DenseTrainer(dims, act, b2.LOSS_SOFTMAX_CE, max_local_batch=lb, device=rank, precision=%(prec)r)
tr.set_params(Ws, bs)
uid = [b2.DenseTrainer.new_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
tr.init_comm(rank, world, uid[0])
if %(async_k)d:
    tr.set_async_period(%(async_k)d)
if %(sharded)d:
    tr.sharded_apply(True)
if %(peer)d:
    hs = [None] * world
    dist.all_gather_object(hs, tr.peer_export())
    tr.peer_import(hs)
dev = torch.device("cuda", rank)
lt, ct = tr.train_steps(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), lb, steps, lr=0.001)
torch.cuda.synchronize()
W, b = tr.get_params()
np.savez(os.path.join(%(out)r, "rank%%d.npz" %% rank), loss=lt.cpu().numpy(), correct=ct.cpu().numpy(), W0=W[0], b0=b[0],
         gs=np.float32(tr.global_step))
tr.close()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("peer", [1, 0], ids=["peer-memory", "nccl"])
@pytest.mark.parametrize("dims,act,scaled,prec", [([784, 10], 0, True, "fp32"), ([784, 100, 10], 1, False, "fp32"),
                                                  ([784, 256, 128, 10], 1, True, "bf16")])
def test_n_rank_trace_equals_single_process_oracle(cuda_device, tmp_path, dims, act, scaled, prec, peer):
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    lb, steps = 96, 12
    script = tmp_path / "worker.py"
    script.write_text(WORKER % dict(root=ROOT, dims=dims, act=act, scaled=scaled, lb=lb, steps=steps, out=str(tmp_path), peer=peer, prec=prec, async_k=0, sharded=0))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
                        "--master-addr", "127.0.0.1", "--master-port", "29611", str(script)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    gb = lb * world
    x, y = synthetic.make_pool(gb * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, oracle.LOSS_SOFTMAX_CE, Ws, bs, precision=prec)
    want = [g.train_step(x[s * gb:(s + 1) * gb], y[s * gb:(s + 1) * gb], 0.001) for s in range(steps)]
    outs = [np.load(tmp_path / ("rank%d.npz" % r_)) for r_ in range(world)]
    ltol, wtol, ctol = (1e-5, 1e-5, 1) if prec == "fp32" else (2e-3, 2e-3, 4)
    for o in outs:
        for s in range(steps):
            assert abs(float(o["loss"][s]) - float(want[s][0])) <= ltol * max(1.0, abs(float(want[s][0])))
            assert abs(int(o["correct"][s]) - want[s][1]) <= ctol
        assert np.abs(o["W0"] - g.W[0]).max() <= wtol * max(1.0, float(np.abs(g.W[0]).max()))
        assert o["gs"] == np.float32(steps)
        np.testing.assert_array_equal(o["W0"], outs[0]["W0"])      # replicas stay bit-identical
        np.testing.assert_array_equal(o["loss"], outs[0]["loss"])


def test_reference_command_lines_1ps_2workers(cuda_device):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    script = os.path.join(ROOT, "distributed", "distributed.py")
    common = ["--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223,127.0.0.1:2224"]     # distributed.py:11-13
    extra = ["--train_rows=4000", "--test_rows=1000", "--training_epochs=2", "--noshuffle", "--data_dir=/nonexistent"]
    ps = subprocess.Popen([sys.executable, script] + common + ["--job_name=ps", "--task_index=0"], stdout=subprocess.PIPE,
                          stderr=subprocess.PIPE, text=True)
    ws = [subprocess.Popen([sys.executable, script] + common + ["--job_name=worker", "--task_index=%d" % i] + extra,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for i in range(2)]
    outs = [w.communicate(timeout=900) for w in ws]
    ps_out = ps.communicate(timeout=60)
    assert ps.returncode == 0 and "nothing to serve" in ps_out[0]
    for w, (so, se) in zip(ws, outs):
        assert w.returncode == 0, se[-3000:]
        assert so.strip().splitlines()[-1] == "done with training"
    # synchronous data parallelism: both workers print the same costs, equal to the single-process oracle at batch 200
    costs = [re.findall(r"Cost now: ([0-9.]+),", so) for so, _ in outs]
    assert costs[0] == costs[1] and len(costs[0]) == 2
    x, y = synthetic.make_pool(5000)
    Ws, bs = synthetic.make_weights([784, 100, 10], False)
    g = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    want = []
    for e in range(2):
        for i in range(20):
            lo, _ = g.train_step(x[i * 200:(i + 1) * 200], y[i * 200:(i + 1) * 200], 0.001)
        want.append(float(lo))
    for got, w in zip(costs[0], want):
        assert abs(float(got) - w) <= 1e-4 + 1e-5 * w


@pytest.mark.parametrize("period", [1, 4])
def test_async_emulation_mode(cuda_device, tmp_path, period):
    """distributed.py:107-113 trains asynchronously (SyncReplicas is commented out).  The emulation lets every rank apply
    its own gradients and averages the replicas every `period` steps: with period 1 that IS the synchronous update
    (mean of the per-rank mean gradients), with a longer period the replicas drift but re-converge at every average."""
    world = min(torch.cuda.device_count(), 2)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    dims, act, scaled, lb, steps = [784, 100, 10], 1, False, 64, 8
    script = tmp_path / "worker.py"
    script.write_text(WORKER % dict(root=ROOT, dims=dims, act=act, scaled=scaled, lb=lb, steps=steps, out=str(tmp_path),
                                    peer=0, prec="fp32", async_k=period, sharded=0))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
                        "--master-addr", "127.0.0.1", "--master-port", "29613", str(script)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    outs = [np.load(tmp_path / ("rank%d.npz" % r_)) for r_ in range(world)]
    np.testing.assert_array_equal(outs[0]["W0"], outs[1]["W0"])       # steps is a multiple of the period: just averaged
    assert outs[0]["gs"] == np.float32(steps * world)                  # global_step counts every worker's update
    gb = lb * world
    x, y = synthetic.make_pool(gb * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    for s in range(steps):
        g.train_step(x[s * gb:(s + 1) * gb], y[s * gb:(s + 1) * gb], 0.001)
    diff = np.abs(outs[0]["W0"] - g.W[0]).max()
    if period == 1:
        assert diff <= 1e-5 * max(1.0, float(np.abs(g.W[0]).max()))    # same update as synchronous data parallelism
    else:
        assert 0 < diff < 1e-2                                          # drifted, but still the same training run


@pytest.mark.parametrize("dims,act,scaled,prec", [([784, 100, 10], 1, False, "fp32"), ([784, 256, 128, 10], 1, True, "bf16")])
def test_sharded_apply_equals_single_process_oracle(cuda_device, tmp_path, dims, act, scaled, prec):
    """SURVEY 8(f) row 3, second half (distributed.py:63-65: replica_device_setter spreads the variables over k parameter
    servers): every rank owns one slice of the parameter vector -- reduce-scatter of the gradients, SGD apply on the own
    slice, all-gather of the parameters.  The N-rank trace must equal the single-process oracle at the global batch, and
    the replicas must stay bit-identical."""
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    lb, steps = 96, 12
    script = tmp_path / "worker.py"
    script.write_text(WORKER % dict(root=ROOT, dims=dims, act=act, scaled=scaled, lb=lb, steps=steps, out=str(tmp_path), peer=0,
                                    prec=prec, async_k=0, sharded=1))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
                        "--master-addr", "127.0.0.1", "--master-port", "29615", str(script)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    gb = lb * world
    x, y = synthetic.make_pool(gb * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, oracle.LOSS_SOFTMAX_CE, Ws, bs, precision=prec)
    want = [g.train_step(x[s * gb:(s + 1) * gb], y[s * gb:(s + 1) * gb], 0.001) for s in range(steps)]
    outs = [np.load(tmp_path / ("rank%d.npz" % r_)) for r_ in range(world)]
    ltol, wtol, ctol = (1e-5, 1e-5, 1) if prec == "fp32" else (2e-3, 2e-3, 4)
    for o in outs:
        for s in range(steps):
            assert abs(float(o["loss"][s]) - float(want[s][0])) <= ltol * max(1.0, abs(float(want[s][0])))
            assert abs(int(o["correct"][s]) - want[s][1]) <= ctol
        assert np.abs(o["W0"] - g.W[0]).max() <= wtol * max(1.0, float(np.abs(g.W[0]).max()))
        assert np.abs(o["b0"] - g.b[0]).max() <= wtol * max(1.0, float(np.abs(g.W[0]).max()))
        assert o["gs"] == np.float32(steps)
        np.testing.assert_array_equal(o["W0"], outs[0]["W0"])      # replicas stay bit-identical
        np.testing.assert_array_equal(o["loss"], outs[0]["loss"])
