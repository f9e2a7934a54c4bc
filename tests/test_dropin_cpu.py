"""CPU tests of the host-side logic around the hot path: flags -> ranks, printed lines, the MNIST IDX
reader, and the N>1 data-parallel arithmetic (two gloo processes standing in for two workers)."""
import gzip
import os
import socket
import struct
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import cluster, mnist_idx, report, synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


# ---- printed lines ---------------------------------------------------------
def test_progress_line_format():
    # distributed.py:157-161 -- five print() arguments: single-space separator + the literal leading space
    line = report.progress_line(np.float32(99), 0, 99, 550, 12.34567, 0.32, 100)
    assert line == "Step so far: 100,  Epoch so far:  1,  Batch used: 100 of 550,  Cost now: 12.3457,  Time spent (delta): 3.2ms"
    assert report.accuracy_line(0.9312) == "Acc: 0.93"                       # distributed.py:164: a fraction
    assert report.final_cost_line(0.1234) == "Final Computed Cost: 0.12"     # distributed.py:166
    assert report.time_taken_line(3.5) == "Time Taken: 3.500000s"
    with pytest.raises(TypeError):                                           # distributed.py:165 as shipped
        report.time_taken_line(3.5, strict_reference_bugs=True)


def test_simple_report_matches_reference_format():
    # simple.py:39 with numpy float32 values
    line = report.simple_report(np.array([-0.9999971], np.float32), np.array([0.9999914], np.float32),
                                np.float32(4.9244164e-11))
    assert line == "W: [-0.9999971] b: [0.9999914] loss 4.9244164e-11"


# ---- flags -> cluster -> rank ------------------------------------------------------------------------
def test_cluster_roles():
    c = cluster.parse_cluster("127.0.0.1:2222", "127.0.0.1:2223,127.0.0.1:2224")   # distributed.py:11-13
    assert c.ps == ["127.0.0.1:2222"] and c.worker == ["127.0.0.1:2223", "127.0.0.1:2224"]
    w1 = cluster.resolve_role(c, "worker", 1)
    assert (w1.rank, w1.world, w1.is_chief, w1.store_host, w1.store_port) == (1, 2, False, "127.0.0.1", 2223)
    assert cluster.resolve_role(c, "worker", 0).is_chief
    assert cluster.resolve_role(c, "ps", 0).rank == -1
    with pytest.raises(ValueError):
        cluster.resolve_role(c, "worker", 2)
    with pytest.raises(AttributeError):          # distributed.py:37: a missing host flag fails in .split
        cluster.parse_cluster(None, "a:1")
    assert cluster.steps_per_epoch(55000, 100, 1) == 550     # distributed.py:142
    assert cluster.steps_per_epoch(55000, 100, 2) == 275


def test_distributed_script_flags_and_ps_role():
    script = os.path.join(ROOT, "distributed", "distributed.py")
    r = subprocess.run([sys.executable, script, "--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223,127.0.0.1:2224",
                        "--job_name=ps", "--task_index=0"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "nothing to serve" in r.stdout
    r = subprocess.run([sys.executable, script, "--job_name=worker"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "AttributeError" in r.stderr
    # names, types, defaults and help strings of distributed.py:24-32
    import importlib.util
    from absl import flags
    spec = importlib.util.spec_from_file_location("b2_distributed_script", script)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)            # defines the flags; main() is not run
    F = flags.FLAGS
    assert F["job_name"].default == "" and F["job_name"].help == "Either 'ps' or 'worker'"
    assert F["task_index"].default == 0 and F["task_index"].help == "Index of task within the job"
    assert F["worker_hosts"].default is None and "worker url list" in F["worker_hosts"].help
    assert F["ps_hosts"].default is None and "ps url list" in F["ps_hosts"].help
    assert (F["batch_size"].default, F["learning_rate"].default, F["training_epochs"].default) == (100, 0.001, 50)


# ---- MNIST IDX reader ---------------------------------------------------------------------------------
def test_mnist_idx_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    def write(name, arr, gz):
        hdr = struct.pack(">I", 0x0800 | arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape)
        path = os.path.join(tmp_path, name + (".gz" if gz else ""))
        with (gzip.open(path, "wb") if gz else open(path, "wb")) as f:
            f.write(hdr + arr.tobytes())
    tx = rng.integers(0, 256, (60, 28, 28), dtype=np.uint8)
    ty = rng.integers(0, 10, 60, dtype=np.uint8)
    ex = rng.integers(0, 256, (20, 28, 28), dtype=np.uint8)
    ey = rng.integers(0, 10, 20, dtype=np.uint8)
    write("train-images-idx3-ubyte", tx, True); write("train-labels-idx1-ubyte", ty, True)
    write("t10k-images-idx3-ubyte", ex, False); write("t10k-labels-idx1-ubyte", ey, False)
    assert mnist_idx.available(str(tmp_path))
    d = mnist_idx.load(str(tmp_path), validation_size=10)
    x, y = d["train"]
    assert x.shape == (50, 784) and y.shape == (50, 10) and x.dtype == np.float32
    np.testing.assert_allclose(x[0], tx[10].reshape(-1) / 255.0, rtol=1e-6)
    assert (y.argmax(1) == ty[10:]).all() and (y.sum(1) == 1).all()
    assert d["test"][0].shape == (20, 784) and d["validation"][0].shape == (10, 784)


# ---- N>1: two gloo processes play two workers -----------------------------------------------------------
def _worker(rank, world, port, store_port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c = cluster.parse_cluster("127.0.0.1:1", ",".join("127.0.0.1:%d" % (store_port + i) for i in range(world)))
    role = cluster.resolve_role(c, "worker", rank)
    uid, store = cluster.exchange_unique_id(role, lambda: bytes(range(128)))
    cluster.store_barrier(store, role, "ready")
    dims, local_batch, steps, lr = [784, 100, 10], 32, 3, 0.001
    gb = local_batch * world
    x, y = synthetic.make_pool(gb * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    g = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, dtype=np.float64)
    losses = []
    for s in range(steps):
        # the row slice b2_train_steps gives this rank: start_s + rank*local_batch (include/b2mlp.h)
        lo_row = s * gb + rank * local_batch
        loss, correct, dW, db = g.gradients(x[lo_row:lo_row + local_batch], y[lo_row:lo_row + local_batch], global_batch=gb)
        flat = torch.from_numpy(np.concatenate([a.ravel() for a in dW + db] + [np.array([loss, correct], np.float64)]))
        dist.all_reduce(flat)          # the one exchange step: [grads..., loss, correct]
        flat = flat.numpy()
        off = 0
        new_dW, new_db = [], []
        for a in dW:
            new_dW.append(flat[off:off + a.size].reshape(a.shape)); off += a.size
        for a in db:
            new_db.append(flat[off:off + a.size].reshape(a.shape)); off += a.size
        g.apply(new_dW, new_db, lr)
        losses.append((flat[off], flat[off + 1]))
    q.put((rank, uid, losses, g.W[0].copy()))
    cluster.store_barrier(store, role, "done")
    dist.destroy_process_group()


def test_two_worker_data_parallel_equals_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, store_port = _free_port(), _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, store_port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=180) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert out[0][1] == out[1][1] == bytes(range(128))            # every worker got the chief's id
    dims, local_batch, steps, lr = [784, 100, 10], 32, 3, 0.001
    gb = local_batch * world
    x, y = synthetic.make_pool(gb * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    ref = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, dtype=np.float64)
    for s in range(steps):
        lo, co = ref.train_step(x[s * gb:(s + 1) * gb], y[s * gb:(s + 1) * gb], lr)
        for r in range(world):
            assert abs(out[r][2][s][0] - float(lo)) < 1e-12 and int(round(out[r][2][s][1])) == co
    np.testing.assert_allclose(out[0][3], ref.W[0], rtol=1e-12, atol=1e-15)
    np.testing.assert_array_equal(out[0][3], out[1][3])           # replicas stay identical


def test_peer_exchange_only_when_all_workers_share_the_node():
    """The NVLink peer-memory exchange maps the other workers' buffers with CUDA IPC: it is only tried when every worker
    address names this node (distributed.py:11-13 uses 127.0.0.1 throughout); otherwise all workers stay on NCCL."""
    assert cluster.single_node(cluster.parse_cluster("127.0.0.1:2222", "127.0.0.1:2223,127.0.0.1:2224"))
    assert cluster.single_node(cluster.parse_cluster("a:1", "localhost:2223,127.0.0.1:2224"))
    assert cluster.single_node(cluster.parse_cluster("a:1", "node7:2223,node7:2224"))
    assert not cluster.single_node(cluster.parse_cluster("a:1", "node7:2223,node8:2223"))
    assert not cluster.single_node(cluster.parse_cluster("a:1", "10.0.0.1:2223,127.0.0.1:2224"))


def test_run_train_keeps_converted_feeds_alive():
    """sess.run takes Python lists (simple.py:25-26) and float64 arrays: both feeds are converted to float32 copies, and BOTH
    copies must stay alive until the C call has read them (the first used to be dropped when the second was converted)."""
    from tensorflow_examples_b200 import session
    x = [[1.0], [2.0], [3.0], [4.0]]                       # a list, as in simple.py
    y = np.array([[0.0], [-1.0], [-2.0], [-3.0]])          # float64
    px, rows_x, keep_x = session._host_ptr(x, 1)
    py, rows_y, keep_y = session._host_ptr(y, 1)
    assert rows_x == rows_y == 4 and keep_x.dtype == np.float32 and keep_y.dtype == np.float32
    assert px.value == keep_x.ctypes.data and py.value == keep_y.ctypes.data and px.value != py.value
    np.testing.assert_array_equal(keep_x.ravel(), [1, 2, 3, 4])      # x's copy is intact after y was converted
    np.testing.assert_array_equal(keep_y.ravel(), [0, -1, -2, -3])
    t = torch.arange(6, dtype=torch.float64).reshape(3, 2)[:, :1]    # non-contiguous float64 tensor
    pt, rows_t, keep_t = session._host_ptr(t, 1)
    assert rows_t == 3 and keep_t.dtype == torch.float32 and keep_t.is_contiguous() and pt.value == keep_t.data_ptr()


def test_native_event_writer_is_read_by_tensorboard(tmp_path, monkeypatch):
    """distributed.py:121-125,138,151 without the tensorboard package: the native writer's file -- TFRecord framing, masked
    CRC-32C, hand-encoded Event / Summary messages -- is what tensorboard's own loader (which verifies both checksums of
    every record) reads back: the file-version event first, then one event per step with the two scalars."""
    from tensorflow_examples_b200 import summaries
    assert summaries.crc32c(b"123456789") == 0xE3069283                      # the CRC-32C check value
    assert summaries.crc32c(b"") == 0
    monkeypatch.setenv("B2_NATIVE_SUMMARIES", "1")
    w = summaries.SummaryWriter(str(tmp_path))
    assert w.enabled and w._native is not None
    vals = [(0, 2.3025851, 0.09), (1, 1.5, 0.25), (300, 1e-9, 1.0), (2 ** 40, -3.75, 0.0)]
    for step, cost, acc in vals:
        w.add_scalars(step, cost, acc)
    w.close()
    files = [f for f in os.listdir(tmp_path) if "tfevents" in f]
    assert len(files) == 1
    path = os.path.join(tmp_path, files[0])
    # (1) a parser of our own: framing and checksums
    import struct
    raw = open(path, "rb").read()
    pos, records = 0, []
    while pos < len(raw):
        (n,) = struct.unpack_from("<Q", raw, pos)
        assert struct.unpack_from("<I", raw, pos + 8)[0] == summaries.masked_crc32c(raw[pos:pos + 8])
        data = raw[pos + 12:pos + 12 + n]
        assert struct.unpack_from("<I", raw, pos + 12 + n)[0] == summaries.masked_crc32c(data)
        records.append(data)
        pos += 16 + n
    assert pos == len(raw) and len(records) == 1 + len(vals)
    # (2) tensorboard's loader, when the package is there
    tb = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    events = list(tb.EventFileLoader(path).Load())
    assert events[0].file_version == "brain.Event:2"
    got = [(ev.step, {v.tag: (v.simple_value if v.HasField("simple_value") else float(v.tensor.float_val[0]))
                      for v in ev.summary.value}) for ev in events[1:]]
    assert [s for s, _ in got] == [s for s, _, _ in vals]
    for (_, d), (_, cost, acc) in zip(got, vals):
        assert d["cost"] == np.float32(cost) and d["accuracy"] == np.float32(acc)
