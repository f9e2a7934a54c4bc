"""Persistent multi-step kernels: b2_train_steps runs all n_steps inside ONE launch (simple.py:34-35 and
distributed.py:145-150 are Python loops around sess.run in the reference).  They must give what the per-step launch
sequence gives -- bit for bit for the 784 -> 10 kernel, whose in-kernel reduction keeps the association of the
separate reduction kernel -- and what the CPU oracle gives within the stated fp32 tolerances."""
import numpy as np
import pytest
import torch

import tensorflow_examples_b200 as pkg
from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic
from tensorflow_examples_b200.report import simple_report

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-5
W_RTOL = 1e-5


def _run(dev, dims, act, loss, Ws, bs, x, y, batch, steps, lr, persistent, first_row=0):
    tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=batch, device=0)
    tr.persistent_steps(persistent)
    tr.multistep_fused(persistent)        # 784 -> 10: the in-kernel step boundary is an option (off by default)
    tr.set_params(Ws, bs)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    l0 = tr.launch_count
    lt, ct = tr.train_steps(xd, yd, batch, steps, first_row=first_row, lr=lr)
    torch.cuda.synchronize()
    launches = tr.launch_count - l0
    gW, gb = tr.get_params()
    gs = tr.global_step
    tr.close()
    return lt.cpu().numpy(), ct.cpu().numpy(), gW, gb, gs, launches


@pytest.mark.parametrize("batch,steps", [(1, 5), (16, 7), (100, 40), (256, 300), (1000, 9), (4741, 6), (16384, 5), (65536, 4)])
@pytest.mark.parametrize("loss", [pkg.LOSS_SOFTMAX_CE, pkg.LOSS_SSE], ids=["softmax-ce", "sse"])
def test_c3_one_launch_equals_per_step_launches_bit_for_bit(cuda_device, batch, steps, loss):
    if loss == pkg.LOSS_SSE and batch not in (100, 4741):
        pytest.skip("the SSE head shares everything but the head warp's branch")
    dims = [784, 10]
    lr = 0.001 if loss == pkg.LOSS_SOFTMAX_CE else 1e-6
    pool = batch * min(steps, 7) + (batch if batch < 1000 else 0)       # fewer batches than steps: the pool wraps
    x, y = synthetic.make_pool(pool)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    one = _run(cuda_device, dims, pkg.ACT_IDENTITY, loss, Ws, bs, x, y, batch, steps, lr, True, first_row=batch)
    per = _run(cuda_device, dims, pkg.ACT_IDENTITY, loss, Ws, bs, x, y, batch, steps, lr, False, first_row=batch)
    assert one[5] == 1 and per[5] == 2 * steps                          # launches: the whole loop in one kernel
    np.testing.assert_array_equal(one[0], per[0])
    np.testing.assert_array_equal(one[1], per[1])
    np.testing.assert_array_equal(one[2][0], per[2][0])
    np.testing.assert_array_equal(one[3][0], per[3][0])
    assert one[4] == per[4] == np.float32(steps)


def test_c3_one_launch_vs_oracle_long_trace(cuda_device):
    dims, batch, steps, lr = [784, 10], 256, 100, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    lt, ct, gW, gb, gs, launches = _run(cuda_device, dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps,
                                        lr, True)
    assert launches == 1
    g = oracle.TF1Graph(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs)
    for s in range(steps):
        lo, co = g.train_step(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), s
        assert abs(int(ct[s]) - co) <= 1, s
    assert np.abs(gW[0] - g.W[0]).max() <= W_RTOL * max(1.0, float(np.abs(g.W[0]).max()))
    assert np.abs(gb[0] - g.b[0]).max() <= W_RTOL
    assert gs == np.float32(steps)


# ---- tiny models: one warp runs the whole loop --------------------------------------------------------------------
@pytest.mark.parametrize("steps,line", [(1000, "W: [-0.9999969] b: [0.9999908] loss 5.6999738e-11"),   # TensorFlow's guide
                                        (1005, "W: [-0.9999971] b: [0.9999914] loss 4.9244164e-11")])  # simple.py:34
def test_simple_py_known_answer_in_one_launch(cuda_device, steps, line):
    """simple/simple.py:11-39 on the GPU: the printed line equals the oracle's -- and, at 1000 steps, the line TensorFlow's
    own guide publishes for this program -- digit for digit, and training is ONE kernel launch."""
    x = torch.tensor([[1], [2], [3], [4]], dtype=torch.float32, device=cuda_device)      # simple.py:25
    y = torch.tensor([[0], [-1], [-2], [-3]], dtype=torch.float32, device=cuda_device)   # simple.py:26
    tr = pkg.DenseTrainer([1, 1], pkg.ACT_IDENTITY, pkg.LOSS_SSE, max_local_batch=4)
    tr.set_params([np.array([[.3]], np.float32)], [np.array([-.3], np.float32)])        # simple.py:11-12
    l0 = tr.launch_count
    lt, _ = tr.train_steps(x, y, 4, steps, lr=0.01)
    torch.cuda.synchronize()
    assert tr.launch_count - l0 == 1
    loss, _ = tr.evaluate(x, y)
    Ws, bs = tr.get_params()
    assert simple_report(Ws[0].reshape(1), bs[0].reshape(1), loss) == line
    assert oracle.simple_py_report(steps) == line
    _, _, _, trace = oracle.simple_py(steps)
    got, want = lt.cpu().numpy(), np.asarray(trace, dtype=np.float32)                       # every pre-update loss
    # numpy's BLAS dot and the kernel's fused multiply-adds round the 4-term sums differently: 1e-5 relative while the
    # loss is a loss, 1e-8 absolute once it is the rounding noise of the residuals (it ends near 5e-11)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-8)
    assert tr.global_step == np.float32(steps)
    tr.close()


@pytest.mark.parametrize("dims,act,loss,batch", [
    ([7, 3], oracle.ACT_IDENTITY, oracle.LOSS_SSE, 5),
    ([6, 5, 2], oracle.ACT_SIGMOID, oracle.LOSS_SSE, 32),
    ([16, 16, 12, 5], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, 31),
    ([12, 10], oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, 17),
])
def test_tiny_models_vs_oracle_and_vs_layer_kernels(cuda_device, dims, act, loss, batch):
    steps, lr = 25, 0.01
    x, y = synthetic.make_pool(batch * 5, dims[0], dims[-1])
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    one = _run(cuda_device, dims, act, loss, Ws, bs, x, y, batch, steps, lr, True)
    per = _run(cuda_device, dims, act, loss, Ws, bs, x, y, batch, steps, lr, False)
    assert one[5] == 1 and per[5] > steps
    g = oracle.TF1Graph(dims, act, loss, Ws, bs)
    for s in range(steps):
        sl = slice((s % 5) * batch, (s % 5) * batch + batch)
        lo, co = g.train_step(x[sl], y[sl], lr)
        for got in (one, per):
            assert abs(float(got[0][s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), s
            if loss == oracle.LOSS_SOFTMAX_CE:
                assert abs(int(got[1][s]) - co) <= 1, s
    for l in range(len(dims) - 1):
        tol = W_RTOL * max(1.0, float(np.abs(g.W[l]).max()))
        assert np.abs(one[2][l] - g.W[l]).max() <= tol and np.abs(one[3][l] - g.b[l]).max() <= tol
        assert np.abs(one[2][l] - per[2][l]).max() <= tol
    assert one[4] == np.float32(steps)


# ---- small batches of the 784-wide models: one cluster of 14 CTAs runs the loop (cluster_steps.cu) ------------------
def _ambiguous(g64, x, margin=1e-4):
    _, z = g64.forward(x)
    top2 = np.sort(z, axis=1)[:, -2:]
    return int(((top2[:, 1] - top2[:, 0]) < margin).sum())


@pytest.mark.parametrize("dims,act,scaled,loss,batch,steps", [
    ([784, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SOFTMAX_CE, 128, 60),       # the largest batch it takes for 784 -> 10
    ([784, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SOFTMAX_CE, 1, 5),
    ([784, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SOFTMAX_CE, 13, 9),         # fewer rows than CTAs
    ([784, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SOFTMAX_CE, 127, 7),
    ([784, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SSE, 100, 7),
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 100, 120),  # distributed.py as shipped (:46,85-86)
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 1, 5),
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 37, 11),
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 99, 6),
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 200, 8),   # the largest batch it takes
    ([784, 100, 10], oracle.ACT_SIGMOID, False, oracle.LOSS_SOFTMAX_CE, 173, 5),
    ([784, 100, 10], oracle.ACT_IDENTITY, True, oracle.LOSS_SOFTMAX_CE, 64, 6),
    ([784, 100, 10], oracle.ACT_SIGMOID, True, oracle.LOSS_SSE, 50, 6),
])
def test_cluster_kernel_vs_oracle_and_vs_per_step_kernels(cuda_device, dims, act, scaled, loss, batch, steps):
    lr = 1e-4 if loss == oracle.LOSS_SSE else 0.001
    nb = min(steps, 9)
    x, y = synthetic.make_pool(batch * nb)
    Ws, bs = synthetic.make_weights(dims, scaled)
    one = _run(cuda_device, dims, act, loss, Ws, bs, x, y, batch, steps, lr, True)
    # (persistent=True also switches the optional fused multi-step path on; it only applies to 784 -> 10, where it would
    #  take precedence -- so ask for the cluster kernel explicitly)
    tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=batch, device=0)
    tr.set_params(Ws, bs)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    l0 = tr.launch_count
    lt, ct = tr.train_steps(xd, yd, batch, steps, lr=lr)
    torch.cuda.synchronize()
    assert tr.launch_count - l0 == 1                     # the whole loop is one launch
    gW, gb = tr.get_params()
    gs = tr.global_step
    tr.close()
    lt, ct = lt.cpu().numpy(), ct.cpu().numpy()
    per = _run(cuda_device, dims, act, loss, Ws, bs, x, y, batch, steps, lr, False)
    g = oracle.TF1Graph(dims, act, loss, Ws, bs)
    g64 = oracle.TF1Graph(dims, act, loss, Ws, bs, dtype=np.float64)
    for s in range(steps):
        sl = slice((s % nb) * batch, (s % nb) * batch + batch)
        amb = _ambiguous(g64, x[sl]) if loss == oracle.LOSS_SOFTMAX_CE else 0
        lo, co = g.train_step(x[sl], y[sl], lr)
        g64.train_step(x[sl], y[sl], lr)
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), (s, float(lt[s]), float(lo))
        assert abs(float(per[0][s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), s
        if loss == oracle.LOSS_SOFTMAX_CE:
            assert abs(int(ct[s]) - co) <= amb, (s, int(ct[s]), co, amb)
    for l in range(len(dims) - 1):
        tol = W_RTOL * max(1.0, float(np.abs(g.W[l]).max()))
        assert np.abs(gW[l] - g.W[l]).max() <= tol, ("W", l, float(np.abs(gW[l] - g.W[l]).max()), tol)
        assert np.abs(gb[l] - g.b[l]).max() <= tol, ("b", l)
        assert np.abs(gW[l] - per[2][l]).max() <= tol
    assert gs == np.float32(steps)
    del one


def test_cluster_kernel_is_bit_reproducible_and_matches_session_run(cuda_device):
    """Two runs give the same bits (fixed-order sums, no atomics), and the sess.run-shaped entry point, which runs the same
    kernel one step at a time, gives the same bits as the one-launch loop."""
    dims, batch, steps, lr = [784, 100, 10], 100, 12, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    runs = []
    for _ in range(2):
        tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, device=0)
        tr.set_params(Ws, bs)
        lt, ct = tr.train_steps(torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device), batch, steps, lr=lr)
        torch.cuda.synchronize()
        runs.append((lt.cpu().numpy(), ct.cpu().numpy(), tr.get_params()))
        tr.close()
    np.testing.assert_array_equal(runs[0][0], runs[1][0])
    np.testing.assert_array_equal(runs[0][2][0][0], runs[1][2][0][0])
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, device=0)
    tr.set_params(Ws, bs)
    for s in range(steps):
        lo, co = tr.run_train(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)
        assert np.float32(lo) == runs[0][0][s] and co == int(runs[0][1][s])
    W, b = tr.get_params()
    np.testing.assert_array_equal(W[0], runs[0][2][0][0])
    np.testing.assert_array_equal(W[1], runs[0][2][0][1])
    tr.close()
