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
    """distributed.py:121-125,138,151: cost and accuracy scalars, one event per step -- and the VALUES are the step's
    pre-update cost and training-batch accuracy, i.e. the oracle's trace."""
    pytest.importorskip("tensorboard")
    rows, bs = 1000, 100
    r = _run_worker(["--ps_hosts=a:1", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--train_rows=%d" % rows,
                     "--test_rows=200", "--training_epochs=1", "--data_dir=/nonexistent", "--noshuffle",
                     "--logs_path=%s" % tmp_path])
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
    x, y = synthetic.make_pool(rows + 200)
    Ws, bsz = synthetic.make_weights([784, 100, 10], False)
    g = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bsz)
    for i in range(rows // bs):
        lo, co = g.train_step(x[i * bs:(i + 1) * bs], y[i * bs:(i + 1) * bs], 0.001)
        assert abs(tags["cost"][i][1] - float(lo)) <= 1e-5 * max(1.0, abs(float(lo))), (i, tags["cost"][i], float(lo))
        assert abs(tags["accuracy"][i][1] - co / float(bs)) <= 1e-6, (i, tags["accuracy"][i], co)


def _write_idx(path, arr):
    import struct
    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    with open(path, "wb") as f:
        f.write(struct.pack(">I", 0x0800 | arr.ndim))
        f.write(struct.pack(">" + "I" * arr.ndim, *arr.shape))
        f.write(arr.tobytes())


def test_distributed_py_from_idx_files_with_shuffling(cuda_device, tmp_path):
    """SURVEY 8(f) row 1 (distributed.py:52-55,142,146,164): the drop-in reads the four MNIST IDX files, holds out the
    first 5000 training rows, reshuffles every epoch like next_batch, and evaluates the test set.  Four small files are
    written here; the permutations the run drew are exported (--dump_perm) and the oracle replays them."""
    rng = np.random.default_rng(11)
    n_train, n_val, n_test, bs, epochs = 1200, 5000, 700, 100, 3
    tr_img = rng.integers(0, 256, (n_val + n_train, 28, 28), dtype=np.uint8)
    tr_lab = rng.integers(0, 10, n_val + n_train).astype(np.uint8)
    te_img = rng.integers(0, 256, (n_test, 28, 28), dtype=np.uint8)
    te_lab = rng.integers(0, 10, n_test).astype(np.uint8)
    d = tmp_path / "MNIST_data"
    d.mkdir()
    _write_idx(d / "train-images-idx3-ubyte", tr_img)
    _write_idx(d / "train-labels-idx1-ubyte", tr_lab)
    _write_idx(d / "t10k-images-idx3-ubyte", te_img)
    _write_idx(d / "t10k-labels-idx1-ubyte", te_lab)
    perm_path = str(tmp_path / "perm.npy")
    r = _run_worker(["--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223", "--job_name=worker", "--task_index=0",
                     "--training_epochs=%d" % epochs, "--data_dir=%s" % d, "--shuffle_seed=3", "--dump_perm=" + perm_path])
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout.strip().splitlines()
    perms = np.load(perm_path)
    assert perms.shape == (epochs, n_train)
    for e in range(epochs):
        assert sorted(perms[e].tolist()) == list(range(n_train))           # a permutation of the 1200 training rows
    assert not np.array_equal(perms[0], perms[1])                           # reshuffled every epoch

    # what input_data.read_data_sets(one_hot=True) hands the reference: float32 pixel / 255, one-hot labels, 5000 held out
    def onehot(lab):
        yy = np.zeros((len(lab), 10), np.float32)
        yy[np.arange(len(lab)), lab] = 1.0
        return yy
    tx = tr_img[n_val:].reshape(n_train, 784).astype(np.float32) * np.float32(1.0 / 255.0)
    ty = onehot(tr_lab[n_val:])
    ex = te_img.reshape(n_test, 784).astype(np.float32) * np.float32(1.0 / 255.0)
    ey = onehot(te_lab)
    Ws, bsz = synthetic.make_weights([784, 100, 10], False)
    g = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bsz)
    per_epoch = n_train // bs
    want = []
    for e in range(epochs):
        px, py = tx[perms[e]], ty[perms[e]]
        for i in range(per_epoch):
            lo, _ = g.train_step(px[i * bs:(i + 1) * bs], py[i * bs:(i + 1) * bs], 0.001)
        want.append(float(lo))                                              # the epoch's last batch is the printed one
    prog = [ln for ln in out if ln.startswith("Step so far")]
    assert len(prog) == epochs
    for e, ln in enumerate(prog):
        m = re.search(r"Step so far: (\d+),  Epoch so far: +(\d+),  Batch used: +(\d+) of +(\d+),  Cost now: ([0-9.]+),", ln)
        assert m, ln
        assert int(m[1]) == (e + 1) * per_epoch and int(m[2]) == e + 1 and int(m[3]) == int(m[4]) == per_epoch
        assert abs(float(m[5]) - want[e]) <= 1e-4 + 1e-4 * want[e], (ln, want[e])
    _, correct = g.evaluate(ex, ey)
    g64 = oracle.TF1Graph([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, g.W, g.b, dtype=np.float64)
    _, z = g64.forward(ex)
    top2 = np.sort(z, axis=1)[:, -2:]
    amb = int(((top2[:, 1] - top2[:, 0]) < 1e-3).sum())
    acc_line = [ln for ln in out if ln.startswith("Acc: ")][0]              # distributed.py:164 on the 700 test rows
    got_acc = float(acc_line.split(": ")[1])
    assert abs(got_acc - correct / float(n_test)) <= 0.005 + amb / float(n_test), (acc_line, correct)
    assert out[-1] == "done with training"


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
