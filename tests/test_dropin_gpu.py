"""GPU tests of the drop-in scripts: same command lines, same printed lines as the reference,
numbers checked against the CPU oracle."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_simple_py_prints_the_reference_line(cuda_device):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "simple", "simple.py")], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = r.stdout.strip().splitlines()[-1]
    m = re.fullmatch(r"W: \[(-?[0-9.e+-]+)\] b: \[(-?[0-9.e+-]+)\] loss ([0-9.e+-]+)", line)   # simple.py:39
    assert m, line
    W, b, loss = (float(g) for g in m.groups())
    Wo, bo, lo, _ = oracle.simple_py()
    assert abs(W - float(Wo[0])) <= 2e-6 and abs(b - float(bo[0])) <= 2e-6 and loss < 1e-9
    assert abs(W + 1) < 1e-5 and abs(b - 1) < 1e-5


def _run_worker(args, timeout=900):
    script = os.path.join(ROOT, "distributed", "distributed.py")
    return subprocess.run([sys.executable, script] + args, capture_output=True, text=True, timeout=timeout)


def test_distributed_py_single_worker_matches_oracle(cuda_device):
    rows, test_rows, bs, epochs = 5500, 1000, 100, 2
    r = _run_worker(["--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--task_index=0",
                     "--train_rows=%d" % rows, "--test_rows=%d" % test_rows, "--training_epochs=%d" % epochs,
                     "--noshuffle", "--data_dir=/nonexistent"])
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout.strip().splitlines()
    assert out[0] == "Variables initialized ..."                                   # distributed.py:127
    assert out[-1] == "done with training"                                         # distributed.py:169
    prog = [l for l in out if l.startswith("Step so far")]
    pat = re.compile(r"Step so far: (\d+),  Epoch so far: +(\d+),  Batch used: +(\d+) of +(\d+),  Cost now: ([0-9.]+),"
                     r"  Time spent \(delta\): +[0-9.]+ms")
    # the oracle on the same synthetic pool and initial weights
    x, y = synthetic.make_pool(rows + test_rows)
    Ws, bsz = synthetic.make_weights([784, 100, 10], False)
    g = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bsz)
    per_epoch = rows // bs
    costs = {}
    step = 0
    for e in range(epochs):
        for i in range(per_epoch):
            lo, _ = g.train_step(x[i * bs:(i + 1) * bs], y[i * bs:(i + 1) * bs], 0.001)
            step += 1
            costs[step] = float(lo)
    assert len(prog) == epochs * 1                                                 # 55 batches: one line per epoch
    for l in prog:
        m = pat.fullmatch(l)
        assert m, l
        st, ep, ba, bc, cost = int(m[1]), int(m[2]), int(m[3]), int(m[4]), float(m[5])
        assert bc == per_epoch and ba == per_epoch and st == ep * per_epoch
        assert abs(cost - costs[st]) <= 1e-4 + 1e-5 * costs[st], (l, costs[st])
    acc_line = [l for l in out if l.startswith("Acc: ")][0]
    _, correct = g.evaluate(x[rows:], y[rows:])
    assert acc_line == "Acc: %2.2f" % (correct / float(test_rows))                 # distributed.py:164
    assert any(re.fullmatch(r"Time Taken: [0-9.]+s", l) for l in out)
    assert [l for l in out if l.startswith("Final Computed Cost: ")][0] == "Final Computed Cost: %.2f" % costs[step]


def test_distributed_py_strict_bug_mode_raises_like_the_reference(cuda_device):
    r = _run_worker(["--ps_hosts=a:1", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--train_rows=1000",
                     "--test_rows=200", "--training_epochs=1", "--strict_reference_bugs", "--data_dir=/nonexistent"])
    assert r.returncode != 0 and "TypeError: not all arguments converted during string formatting" in r.stderr
    assert "Acc: " in r.stdout                                                      # distributed.py:164 ran first


def test_distributed_py_writes_tensorboard_scalars(cuda_device, tmp_path):
    """distributed.py:121-125,138,151: cost and accuracy scalars, one event per step."""
    pytest.importorskip("tensorboard")
    r = _run_worker(["--ps_hosts=a:1", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--train_rows=1000",
                     "--test_rows=200", "--training_epochs=1", "--data_dir=/nonexistent", "--logs_path=%s" % tmp_path])
    assert r.returncode == 0, r.stderr[-2000:]
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader
    files = [f for f in os.listdir(tmp_path) if "tfevents" in f]
    assert files
    tags = {}
    for ev in EventFileLoader(os.path.join(tmp_path, files[0])).Load():
        for v in ev.summary.value:
            tags.setdefault(v.tag, []).append((ev.step, v.simple_value if v.HasField("simple_value") else float(v.tensor.float_val[0])))
    assert len(tags["cost"]) == 10 and len(tags["accuracy"]) == 10          # 1000 rows / batch 100
    assert [s for s, _ in tags["cost"]] == list(range(10))
    assert all(0.0 <= a <= 1.0 for _, a in tags["accuracy"])


def test_distributed_py_tf_op_seeds_save_restore(cuda_device, tmp_path):
    """--tf_op_seeds draws W1, W2 with the restated tf.random_normal stream (distributed.py:83-86); --save_path /
    --restore_path carry the variables from one run into the next (the reference keeps no checkpoint)."""
    from oracle import tf1_random
    rows, test_rows, bs = 1100, 500, 100
    common = ["--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--task_index=0",
              "--train_rows=%d" % rows, "--test_rows=%d" % test_rows, "--noshuffle", "--data_dir=/nonexistent"]
    ckpt = str(tmp_path / "vars.npz")
    r1 = _run_worker(common + ["--training_epochs=1", "--tf_op_seeds=11,17", "--save_path=" + ckpt])
    assert r1.returncode == 0, r1.stderr[-2000:]
    # the oracle from the same generator stream
    x, y = synthetic.make_pool(rows + test_rows)
    Ws = [tf1_random.random_normal((784, 100), 2, 11), tf1_random.random_normal((100, 10), 2, 17)]
    bsz = [np.zeros(100, np.float32), np.zeros(10, np.float32)]
    g = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bsz)
    cost = None
    for i in range(rows // bs):
        cost, _ = g.train_step(x[i * bs:(i + 1) * bs], y[i * bs:(i + 1) * bs], 0.001)
    final1 = [l for l in r1.stdout.splitlines() if l.startswith("Final Computed Cost: ")][0]
    assert abs(float(final1.split(": ")[1]) - float(cost)) <= 0.011, (final1, float(cost))
    with np.load(ckpt) as d:
        assert np.abs(d["W0"] - g.W[0]).max() <= 1e-4 and np.abs(d["W1"] - g.W[1]).max() <= 1e-4
        assert float(d["global_step"]) == float(rows // bs)
    # second run continues from the checkpoint: its first epoch is the oracle's second
    r2 = _run_worker(common + ["--training_epochs=1", "--restore_path=" + ckpt])
    assert r2.returncode == 0, r2.stderr[-2000:]
    for i in range(rows // bs):
        cost, _ = g.train_step(x[i * bs:(i + 1) * bs], y[i * bs:(i + 1) * bs], 0.001)
    final2 = [l for l in r2.stdout.splitlines() if l.startswith("Final Computed Cost: ")][0]
    assert abs(float(final2.split(": ")[1]) - float(cost)) <= 0.011, (final2, float(cost))
