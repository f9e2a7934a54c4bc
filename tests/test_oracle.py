"""CPU tests of the oracle (oracle/tf1_graph.py).

The reference has no tests or golden vectors of its own, so the oracle is pinned against
(1) the analytic / hand-computed answers its source implies, (2) an independent autodiff of the same
forward graph (torch.autograd, float64), (3) committed regression fixtures (tests/golden/).
"""
import os

import numpy as np
import pytest
import torch

from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- simple/simple.py --------------------------------------------------------------------------
def test_simple_py_step0_by_hand():
    # simple.py:11-12,20,25-26: pred=[0,.3,.6,.9], d=[0,1.3,2.6,3.9], loss=23.66, dW=52.0, db=15.6
    g = oracle.TF1Graph([1, 1], oracle.ACT_IDENTITY, oracle.LOSS_SSE, [[0.3]], [[-0.3]])
    loss, _, dW, db = g.gradients(np.array([1, 2, 3, 4], np.float32), np.array([0, -1, -2, -3], np.float32))
    assert abs(float(loss) - 23.66) < 1e-5
    assert abs(float(dW[0][0, 0]) - 52.0) < 1e-4
    assert abs(float(db[0][0]) - 15.6) < 1e-5
    g.apply(dW, db, 0.01)
    assert abs(float(g.W[0][0, 0]) - (-0.22)) < 1e-6
    assert abs(float(g.b[0][0]) - (-0.456)) < 1e-6


def test_simple_py_known_answer():
    # the line simple.py:39 prints after 1005 fp32 steps
    assert oracle.simple_py_report() == "W: [-0.9999971] b: [0.9999914] loss 4.9244164e-11"
    W, b, loss, trace = oracle.simple_py()
    assert abs(float(trace[0]) - 23.66) < 1e-5
    assert abs(float(trace[1]) - 4.018144) < 1e-5
    assert abs(float(trace[2]) - 1.819874) < 1e-5
    # analytic optimum W=-1, b=1, loss=0 (data lie on y = -x + 1)
    assert abs(float(W[0]) + 1) < 1e-5 and abs(float(b[0]) - 1) < 1e-5 and float(loss) < 1e-9
    W64, b64, loss64, _ = oracle.simple_py(dtype=np.float64)
    assert abs(W64[0] - (-0.99999707)) < 1e-8 and abs(b64[0] - 0.99999139) < 1e-8


def test_simple_py_1000_steps_equals_the_line_tensorflow_publishes():
    """An EXTERNAL pin: TensorFlow 1.x's own "Getting Started" guide lists this very program (the source of
    simple/simple.py:11-39) run for 1000 steps and prints
        W: [-0.9999969] b: [ 0.99999082] loss: 5.69997e-11
    (numpy 1.x spacing / 8 significant digits).  The restatement gives the same float32 values digit for digit."""
    assert oracle.simple_py_report(1000) == "W: [-0.9999969] b: [0.9999908] loss 5.6999738e-11"
    W, b, loss, _ = oracle.simple_py(1000)
    assert "%.7f" % float(W[0]) == "-0.9999969"
    assert "%.8f" % float(b[0]) == "0.99999082"           # the guide's 8-digit rendering of b
    assert "%.5e" % float(loss) == "5.69997e-11"           # the guide's 6-digit rendering of the loss


def test_simple_py_forward_and_loss_lines_of_the_guide():
    """The same guide evaluates the untrained model first (W = .3, b = -.3, the initial values of simple.py:11-12) and prints
        [ 0.          0.30000001  0.60000002  0.90000004]      sess.run(linear_model, {x: [1, 2, 3, 4]})
        23.66                                                  sess.run(loss, {x: [1, 2, 3, 4], y: [0, -1, -2, -3]})
        0.0                                                    the loss after assigning W = -1, b = 1
    (float32 values at numpy 1.x's 8 significant digits).  The restatement's forward pass gives those float32 values."""
    x, y = np.array([1, 2, 3, 4], np.float32), np.array([0, -1, -2, -3], np.float32)
    g = oracle.TF1Graph([1, 1], oracle.ACT_IDENTITY, oracle.LOSS_SSE, [[0.3]], [[-0.3]])
    _, pred = g.forward(x.reshape(-1, 1))
    assert ["%.8g" % v for v in np.asarray(pred, np.float32).ravel()] == ["0", "0.30000001", "0.60000002", "0.90000004"]
    loss, _ = g.evaluate(x, y)
    assert "%.4g" % float(loss) == "23.66"
    g = oracle.TF1Graph([1, 1], oracle.ACT_IDENTITY, oracle.LOSS_SSE, [[-1.0]], [[1.0]])
    loss, _ = g.evaluate(x, y)
    assert float(loss) == 0.0


def test_simple_py_fixture():
    f = np.load(os.path.join(GOLDEN, "simple_py.npz"))
    W, b, loss, trace = oracle.simple_py()
    np.testing.assert_array_equal(W, f["W"])
    np.testing.assert_array_equal(b, f["b"])
    assert np.float32(loss) == f["loss"]
    np.testing.assert_array_equal(np.array(trace, np.float32), f["trace"])
    assert str(f["report"]) == oracle.simple_py_report()


# ---- gradient compositions vs an independent autodiff -------------------------------------------------
def _autograd_reference(dims, act, loss, Ws, bs, x, y):
    Wt = [torch.tensor(w, dtype=torch.float64, requires_grad=True) for w in Ws]
    bt = [torch.tensor(b, dtype=torch.float64, requires_grad=True) for b in bs]
    a = torch.tensor(x, dtype=torch.float64).reshape(-1, dims[0])
    yt = torch.tensor(y, dtype=torch.float64).reshape(-1, dims[-1])
    for i in range(len(Wt)):
        z = a @ Wt[i] + bt[i]
        a = torch.sigmoid(z) if (act == oracle.ACT_SIGMOID and i < len(Wt) - 1) else z
    if loss == oracle.LOSS_SSE:
        L = ((z - yt) ** 2).sum()
    else:
        L = (-(yt * torch.log(torch.softmax(z, dim=1))).sum(dim=1)).mean()   # distributed.py:99,102
    L.backward()
    return float(L.detach()), [w.grad.numpy() for w in Wt], [b.grad.numpy() for b in bt]


@pytest.mark.parametrize("dims,act,loss,scaled", [
    ([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, False),
    ([784, 10], oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, True),
    ([20, 16, 12, 5], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, True),
    ([7, 3], oracle.ACT_IDENTITY, oracle.LOSS_SSE, True),
    ([6, 5, 2], oracle.ACT_SIGMOID, oracle.LOSS_SSE, True),
])
def test_gradients_match_autograd(dims, act, loss, scaled):
    rows = 64
    x, y = synthetic.make_pool(rows, dims[0], dims[-1])
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, loss, Ws, bs, dtype=np.float64)
    L, _, dW, db = g.gradients(x, y)
    L_ref, dW_ref, db_ref = _autograd_reference(dims, act, loss, Ws, bs, x, y)
    assert abs(float(L) - L_ref) <= 1e-12 * max(1, abs(L_ref))
    for a, b in zip(dW + db, dW_ref + db_ref):
        np.testing.assert_allclose(a, b, rtol=1e-9, atol=1e-13)


def test_fp32_tracks_fp64_twin():
    f = np.load(os.path.join(GOLDEN, "c2_784_100_10_b100.npz"))
    np.testing.assert_allclose(f["loss32"], f["loss64"], rtol=2e-6)
    np.testing.assert_array_equal(f["correct32"], f["correct64"])


# ---- regression fixtures ------------------------------------------------------------------------
@pytest.mark.parametrize("name,dims,act,scaled,batch", [
    ("c2_784_100_10_b100", [784, 100, 10], oracle.ACT_SIGMOID, False, 100),
    ("c3_784_10_b256", [784, 10], oracle.ACT_IDENTITY, True, 256),
])
def test_fixture_traces(name, dims, act, scaled, batch):
    f = np.load(os.path.join(GOLDEN, name + ".npz"))
    steps = len(f["loss32"])
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    for s in range(steps):
        lo, co = g.train_step(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], 0.001)
        # BLAS summation order may differ between hosts: tolerance, not bit equality
        assert abs(float(lo) - float(f["loss32"][s])) <= 2e-6 * abs(float(f["loss32"][s]))
        assert co == int(f["correct32"][s])
    for l in range(len(dims) - 1):
        np.testing.assert_allclose(g.b[l], f["b%d" % l], rtol=1e-4, atol=1e-7)
        np.testing.assert_allclose(g.W[l][::16], f["W%d_rows16" % l], rtol=1e-4, atol=1e-6)
    assert g.global_step == np.float32(steps)   # distributed.py:68-70,108
    # initial regime the survey quotes for the reference model: cost ~ 9-13, accuracy ~ 0.1
    if name.startswith("c2"):
        assert 8.0 < float(f["loss32"][0]) < 14.0 and 0 <= int(f["correct32"][0]) <= 25


# ---- data-parallel form -----------------------------------------------------------------------------
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_sharded_step_equals_single_process(nranks):
    dims, batch = [784, 100, 10], 64
    x, y = synthetic.make_pool(batch * 3)
    Ws, bs = synthetic.make_weights(dims, False)
    a = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, dtype=np.float64)
    b = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, dtype=np.float64)
    for s in range(3):
        xs, ys = x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch]
        la, ca = a.train_step(xs, ys, 0.001)
        lb, cb = b.train_step_sharded(xs, ys, 0.001, nranks)
        assert abs(float(la) - float(lb)) < 1e-12 and ca == cb
    for l in range(2):
        np.testing.assert_allclose(a.W[l], b.W[l], rtol=1e-11, atol=1e-14)


def test_bf16_round():
    a = np.array([1.0, 1.00390625, 1.005859375, -3.1415927, 0.0, 65504.0], np.float32)
    got = oracle.bf16_round(a)
    want = torch.tensor(a).to(torch.bfloat16).to(torch.float32).numpy()
    np.testing.assert_array_equal(got, want)


def test_unfused_log_softmax_nan_is_reproduced():
    # distributed.py:99,102 computes log(softmax(z)) un-fused: an underflowing probability gives 0*(-inf) = NaN
    g = oracle.TF1Graph([1, 2], oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, [[[0.0, 200.0]]], [[0.0, 0.0]])
    loss, _ = g.evaluate(np.array([[1.0]], np.float32), np.array([[0.0, 1.0]], np.float32))
    assert np.isnan(loss)


def test_torch_cpu_baseline_matches_numpy_oracle():
    from oracle.torch_cpu import TorchCPUGraph
    dims, batch, steps = [784, 100, 10], 100, 5
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    a = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    b = TorchCPUGraph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    xt, yt = torch.from_numpy(x), torch.from_numpy(y)
    for s in range(steps):
        sl = slice(s * batch, (s + 1) * batch)
        la, ca = a.train_step(x[sl], y[sl], 0.001)
        lb, cb = b.train_step(xt[sl], yt[sl], 0.001)
        assert abs(float(la) - lb) <= 1e-5 * abs(float(la)) and ca == cb
    np.testing.assert_allclose(a.W[0], b.W[0].numpy(), rtol=1e-4, atol=1e-5)


# ---- tf.random_normal restatement (oracle/tf1_random.py) ---------------------------------------------------------
def test_philox4x32_10_known_answers():
    """The block function behind tf.random_normal against the Random123 known-answer vectors."""
    from oracle import tf1_random
    for ctr, key, want in tf1_random.KAT:
        got = tf1_random.philox4x32_10(np.array([ctr], dtype=np.uint32), key)[0]
        assert tuple(int(v) for v in got) == want


def test_random_normal_stream_properties():
    from oracle import tf1_random
    a = tf1_random.random_normal((784, 100), 2, 11)
    assert a.dtype == np.float32 and a.shape == (784, 100)
    assert abs(float(a.mean())) < 0.02 and abs(float(a.std()) - 1.0) < 0.02 and float(np.abs(a).max()) < 6.0
    # deterministic; a different op seed is a different stream; the first n values do not depend on n (4 per block)
    np.testing.assert_array_equal(a, tf1_random.random_normal((784, 100), 2, 11))
    assert not np.array_equal(a, tf1_random.random_normal((784, 100), 2, 12))
    np.testing.assert_array_equal(a.reshape(-1)[:1001], tf1_random.random_normal((1001,), 2, 11))
    b = tf1_random.random_normal((1000,), 2, 11, mean=3.0, stddev=0.5)
    np.testing.assert_allclose(b, a.reshape(-1)[:1000] * np.float32(0.5) + np.float32(3.0), rtol=0, atol=1e-6)
