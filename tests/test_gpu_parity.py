"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs, against the committed fixtures, and through size-independent properties at full size.

Stated tolerances (fp32 path):
  per-step loss            |gpu - oracle| <= 1e-5 * max(1, |oracle|)
  weights after the trace  max|gpu - oracle| <= 1e-5 * max(1, max|W|)
  correct-row count        identical, except rows whose top-2 logit margin (float64) is below 1e-4,
                           where fp32 summation order may legitimately flip the argmax
"""
import os

import numpy as np
import pytest
import torch

import tensorflow_examples_b200 as pkg
from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOSS_RTOL = 1e-5
W_RTOL = 1e-5


def _ambiguous(g64, x, margin=1e-4):
    _, z = g64.forward(x)
    top2 = np.sort(z, axis=1)[:, -2:]
    return int(((top2[:, 1] - top2[:, 0]) < margin).sum())


def _run_case(dev, dims, act, loss, Ws, bs, x, y, batch, steps, lr, force_layered=False):
    tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=batch, device=0)
    if force_layered:
        tr.force_layered(True)
    tr.set_params(Ws, bs)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    lt, ct = tr.train_steps(xd, yd, batch, steps, lr=lr)
    torch.cuda.synchronize()
    gW, gb = tr.get_params()
    gs = tr.global_step
    tr.close()
    return lt.cpu().numpy(), ct.cpu().numpy(), gW, gb, gs


def _check_against_oracle(dims, act, loss, Ws, bs, x, y, batch, steps, lr, got):
    lt, ct, gW, gb, gs = got
    g32 = oracle.TF1Graph(dims, act, loss, Ws, bs)
    g64 = oracle.TF1Graph(dims, act, loss, Ws, bs, dtype=np.float64)
    nb = len(x) // batch
    for s in range(steps):
        sl = slice((s % nb) * batch, (s % nb) * batch + batch)
        amb = _ambiguous(g64, x[sl]) if loss == oracle.LOSS_SOFTMAX_CE else 0
        lo, co = g32.train_step(x[sl], y[sl], lr)
        g64.train_step(x[sl], y[sl], lr)
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), (s, lt[s], lo)
        if loss == oracle.LOSS_SOFTMAX_CE:
            assert abs(int(ct[s]) - co) <= amb, (s, int(ct[s]), co, amb)
    for l in range(len(dims) - 1):
        tol = W_RTOL * max(1.0, float(np.abs(g32.W[l]).max()))
        assert np.abs(gW[l] - g32.W[l]).max() <= tol, ("W", l, np.abs(gW[l] - g32.W[l]).max(), tol)
        assert np.abs(gb[l] - g32.b[l]).max() <= tol, ("b", l)
    assert gs == np.float32(steps)
    return g32


# ---- config 1: simple/simple.py as shipped -----------------------------------------------------
def test_simple_py_1005_steps(cuda_device):
    x = np.array([[1], [2], [3], [4]], np.float32)           # simple.py:25
    y = np.array([[0], [-1], [-2], [-3]], np.float32)        # simple.py:26
    tr = pkg.DenseTrainer([1, 1], pkg.ACT_IDENTITY, pkg.LOSS_SSE, max_local_batch=4)
    tr.set_params([np.array([[0.3]], np.float32)], [np.array([-0.3], np.float32)])   # simple.py:11-12
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    lt, _ = tr.train_steps(xd, yd, 4, 1005, lr=0.01)          # simple.py:22,34-35
    final, _ = tr.evaluate(xd, yd)                            # simple.py:37
    gW, gb = tr.get_params()
    W, b, loss, trace = oracle.simple_py()
    lt = lt.cpu().numpy()
    assert abs(float(gW[0][0, 0]) - float(W[0])) <= 2e-6 and abs(float(gb[0][0]) - float(b[0])) <= 2e-6
    assert float(final) < 1e-9
    for s in (0, 1, 2, 10, 100):
        assert abs(float(lt[s]) - float(trace[s])) <= 1e-5 * max(1.0, float(trace[s])), s
    f = np.load(os.path.join(GOLDEN, "simple_py.npz"))
    assert abs(float(gW[0][0, 0]) - float(f["W"][0])) <= 2e-6
    assert tr.global_step == np.float32(1005)
    tr.close()


# ---- config 2 shape: distributed.py's 784-100-10 at its shipped batch size --------------------------
def test_c2_784_100_10_b100_trace(cuda_device):
    dims, batch, steps, lr = [784, 100, 10], 100, 100, 0.001   # distributed.py:46-47,85-86
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=False)
    got = _run_case(cuda_device, dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, got)
    # the committed fixture (20 steps over a 20-batch pool: the label stream depends on the pool size)
    f = np.load(os.path.join(GOLDEN, "c2_784_100_10_b100.npz"))
    n = len(f["loss32"])
    x, y = synthetic.make_pool(batch * n)
    got = _run_case(cuda_device, dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, n, lr)
    np.testing.assert_allclose(got[0], f["loss32"], rtol=LOSS_RTOL)
    np.testing.assert_array_equal(got[1], f["correct32"])
    np.testing.assert_allclose(got[3][0], f["b0"], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(got[2][0][::16], f["W0_rows16"], rtol=1e-4, atol=1e-6)


# ---- config 3: 784 -> 10, fused single-kernel step ----------------------------------------------------
def test_c3_784_10_b256_trace_fused_and_layered(cuda_device):
    dims, batch, steps, lr = [784, 10], 256, 100, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    fused = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, fused)
    layered = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr,
                        force_layered=True)
    _check_against_oracle(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, layered)
    f = np.load(os.path.join(GOLDEN, "c3_784_10_b256.npz"))
    n = len(f["loss32"])
    x, y = synthetic.make_pool(batch * n)
    fused = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, n, lr)
    np.testing.assert_allclose(fused[0], f["loss32"], rtol=LOSS_RTOL)
    np.testing.assert_array_equal(fused[1], f["correct32"])
    np.testing.assert_allclose(fused[2][0][::16], f["W0_rows16"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("batch", [1, 7, 8, 31, 32, 33, 100, 255, 1000, 4736 + 5])
def test_c3_ragged_batches(cuda_device, batch):
    dims, steps, lr = [784, 10], 3, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    got = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, got)


def test_c3_full_size_b65536(cuda_device):
    """BASELINE.json's largest single-layer batch: oracle comparison (numpy finishes in seconds) plus
    properties: fused and layered paths agree, and a repeated run is bit-identical."""
    dims, batch, steps, lr = [784, 10], 65536, 3, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    a = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, a)
    b = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[2][0], b[2][0])            # no float atomics: bit-reproducible
    c = _run_case(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr,
                  force_layered=True)
    np.testing.assert_allclose(a[0], c[0], rtol=LOSS_RTOL)
    assert np.abs(a[2][0] - c[2][0]).max() <= W_RTOL
    assert np.abs(a[1] - c[1]).max() <= 2


def test_c2_model_large_batch(cuda_device):
    dims, batch, steps, lr = [784, 100, 10], 8192, 3, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=False)
    got = _run_case(cuda_device, dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr, got)


# ---- generic shapes / heads -------------------------------------------------------------------------
@pytest.mark.parametrize("dims,act,loss,batch", [
    ([7, 3], oracle.ACT_IDENTITY, oracle.LOSS_SSE, 5),
    ([20, 16, 12, 5], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, 77),
    ([6, 5, 2], oracle.ACT_SIGMOID, oracle.LOSS_SSE, 64),
    ([130, 70, 33], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, 129),
])
def test_generic_shapes(cuda_device, dims, act, loss, batch):
    steps, lr = 5, 0.001
    x, y = synthetic.make_pool(batch * steps, dims[0], dims[-1])
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    got = _run_case(cuda_device, dims, act, loss, Ws, bs, x, y, batch, steps, lr)
    _check_against_oracle(dims, act, loss, Ws, bs, x, y, batch, steps, lr, got)


# ---- the other entry points --------------------------------------------------------------------------
def test_eval_matches_oracle(cuda_device):
    dims = [784, 100, 10]
    x, y = synthetic.make_pool(10000)                           # distributed.py:164 evaluates 10000 test rows
    Ws, bs = synthetic.make_weights(dims, scaled=False)
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=4096)
    tr.set_params(Ws, bs)
    loss, correct = tr.evaluate(torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device))
    g = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    lo, co = g.evaluate(x, y)
    g64 = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, dtype=np.float64)
    assert abs(float(loss) - float(lo)) <= LOSS_RTOL * abs(float(lo))
    assert abs(correct - co) <= _ambiguous(g64, x)
    tr.close()


@pytest.mark.parametrize("dims,act,scaled", [([784, 10], 0, True), ([784, 100, 10], 1, False)])
def test_session_run_with_host_feed_equals_device_pool_path(cuda_device, dims, act, scaled):
    batch, steps, lr = 128, 4, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    a = _run_case(cuda_device, dims, act, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, lr)
    tr = pkg.DenseTrainer(dims, act, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch)
    tr.set_params(Ws, bs)
    for s in range(steps):
        lo, co = tr.run_train(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)
        assert np.float32(lo) == a[0][s] and co == int(a[1][s])
    gW, _ = tr.get_params()
    np.testing.assert_array_equal(gW[0], a[2][0])
    tr.close()


def test_argument_errors_surface_as_exceptions(cuda_device):
    tr = pkg.DenseTrainer([784, 10], pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, max_local_batch=64)
    x = torch.zeros(128, 784, device=cuda_device)
    y = torch.zeros(128, 10, device=cuda_device)
    with pytest.raises(pkg.B2Error):
        tr.train_steps(x, y, 128, 1)             # local_batch > max_local_batch
    with pytest.raises(pkg.B2Error):
        tr.train_steps(x, y, 64, 1, first_row=3)  # not a multiple of the batch
    tr.close()


# ---- distributed.py:99,102: the un-fused log(softmax) yields NaN when a probability underflows ---------------------
@pytest.mark.parametrize("path", ["cluster-784-10", "fused-per-step", "fused-multistep", "layered-784-10", "cluster-784-100-10", "narrow-head-784-100-10", "tiny-4-3",
                                  "layered-20-3"])
def test_softmax_underflow_gives_nan_like_the_reference(cuda_device, path):
    """The reference writes tf.log(tf.nn.softmax(z)) un-fused.  When one class's probability underflows to 0 the row's
    term is 0 * log(0) = 0 * -inf = NaN and so is the fetched cost; every GPU path keeps that, exactly like the oracle
    (tests/test_oracle.py::test_unfused_log_softmax_nan_is_reproduced)."""
    if path in ("cluster-784-10", "fused-per-step", "fused-multistep", "layered-784-10"):
        dims, act = [784, 10], pkg.ACT_IDENTITY
    elif path in ("narrow-head-784-100-10", "cluster-784-100-10"):
        dims, act = [784, 100, 10], pkg.ACT_SIGMOID
    elif path == "tiny-4-3":
        dims, act = [4, 3], pkg.ACT_IDENTITY
    else:
        dims, act = [20, 3], pkg.ACT_IDENTITY
    batch = 48 if dims[0] > 16 else 8
    rng = np.random.default_rng(5)
    x = rng.random((batch, dims[0]), dtype=np.float32)
    C = dims[-1]
    y = np.zeros((batch, C), np.float32)
    y[np.arange(batch), rng.integers(0, C, batch)] = 1.0
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    bs[-1][0] = 300.0                         # class 0's logit towers over the rest: exp(z - max) underflows to 0 for them
    g = oracle.TF1Graph(dims, act, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    want, _ = g.evaluate(x, y)
    assert np.isnan(want)
    tr = pkg.DenseTrainer(dims, act, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch)
    if path.startswith("layered"):
        tr.force_layered(True)
    if path == "fused-multistep":
        tr.multistep_fused(True)
    if path in ("fused-per-step", "narrow-head-784-100-10"):
        tr.persistent_steps(False)
    tr.set_params(Ws, bs)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    loss_eval, _ = tr.evaluate(xd, yd)
    lt, _ = tr.train_steps(xd, yd, batch, 1, lr=0.001)
    torch.cuda.synchronize()
    assert np.isnan(float(loss_eval)), (path, loss_eval)
    assert np.isnan(float(lt[0].item())), (path, float(lt[0].item()))
    tr.close()


def test_misaligned_rows_are_refused_not_trapped(cuda_device):
    """The fused step loads x rows with cp.async.bulk, which needs a 16-byte aligned source: a pool that starts at an odd
    float offset (a view into a larger buffer) must come back as B2_ERR_ARG with a message, from every entry point that
    takes device rows -- not as a misaligned-address fault that kills the context."""
    dims, rows = [784, 10], 64
    x, y = synthetic.make_pool(rows, dims[0], dims[-1])
    flat = torch.zeros(rows * dims[0] + 8, dtype=torch.float32, device=cuda_device)
    xd = flat[1:1 + rows * dims[0]].view(rows, dims[0])             # contiguous, 4 bytes off a 16-byte boundary
    xd.copy_(torch.from_numpy(x))
    yd = torch.from_numpy(y).to(cuda_device)
    assert xd.is_contiguous() and xd.data_ptr() % 16 == 4
    tr = pkg.DenseTrainer(dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, max_local_batch=rows)
    tr.set_params(*synthetic.make_weights(dims, True))
    with pytest.raises(pkg.B2Error, match="16-byte aligned"):
        tr.train_steps(xd, yd, rows, 1, lr=0.001)
    with pytest.raises(pkg.B2Error, match="16-byte aligned"):
        tr.evaluate(xd, yd)
    good = torch.from_numpy(x).to(cuda_device)                      # the context is alive and the aligned pool works
    lt, _ = tr.train_steps(good, yd, rows, 1, lr=0.001)
    torch.cuda.synchronize()
    assert np.isfinite(float(lt[0]))
    tr.close()
