"""GPU parity of the two kernels the fp32 multi-layer path gained for large batches (all through the C ABI):

  * gemm_split3.cu -- a wide fp32 layer (>= 256 inputs, >= 32 outputs) on the tensor cores as six bf16 products over
    three-plane splits of the operands (the layer input is split inside the kernel);
  * narrow_head.cu -- the output layer (<= 128 inputs, <= 16 classes): forward, loss head and its whole backward pass.

Both are checked against the CPU oracle with the fp32 tolerances of test_gpu_parity.py, against the CUDA-core path
they replace (B2_OPT_NO_SPLIT3 / B2_OPT_NO_NARROW_HEAD), on ragged batch sizes, and on evaluation.
"""
import numpy as np
import pytest
import torch

import tensorflow_examples_b200 as pkg
from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-5
W_RTOL = 1e-5


def _train(dims, act, loss, Ws, bs, x, y, batch, steps, lr, split3=True, narrow=True):
    tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=batch, device=0)
    tr.tensor_core_fp32(split3)
    tr.fused_output_layer(narrow)
    tr.set_params(Ws, bs)
    dev = torch.device("cuda", 0)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    lt, ct = tr.train_steps(xd, yd, batch, steps, lr=lr)
    ev = tr.evaluate(xd[:batch], yd[:batch])
    torch.cuda.synchronize()
    W, b = tr.get_params()
    gs = tr.global_step
    tr.close()
    return lt.cpu().numpy(), ct.cpu().numpy(), W, b, gs, ev


def _ambiguous(g64, x, margin=1e-4):
    """Rows whose top-2 logit margin (float64) is below `margin`: only there may fp32 summation order flip the argmax."""
    _, z = g64.forward(x)
    top2 = np.sort(z, axis=1)[:, -2:]
    return int(((top2[:, 1] - top2[:, 0]) < margin).sum())


def _oracle(dims, act, loss, Ws, bs, x, y, batch, steps, lr):
    """fp32 oracle trace; next to each step the number of rows the float64 twin finds ambiguous (north_star: argmax
    bit-exact -- the correct-row count may differ from the oracle's by at most that many rows, usually zero)."""
    g = oracle.TF1Graph(dims, act, loss, Ws, bs)
    g64 = oracle.TF1Graph(dims, act, loss, Ws, bs, dtype=np.float64)
    tr, amb = [], []
    for s in range(steps):
        xs, ys = x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch]
        amb.append(_ambiguous(g64, xs) if loss == oracle.LOSS_SOFTMAX_CE else 0)
        tr.append(g.train_step(xs, ys, lr))
        g64.train_step(xs, ys, lr)
    ev = g.evaluate(x[:batch], y[:batch])
    amb.append(_ambiguous(g64, x[:batch]) if loss == oracle.LOSS_SOFTMAX_CE else 0)
    return g, tr, ev, amb


def _check(got, want):
    lt, ct, W, b, gs, ev = got
    g, tr, ev_o, amb = want
    for s, (lo, co) in enumerate(tr):
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), (s, lt[s], lo)
        assert abs(int(ct[s]) - co) <= amb[s], (s, int(ct[s]), co, amb[s])
    for l in range(len(W)):
        tol = W_RTOL * max(1.0, float(np.abs(g.W[l]).max()))
        assert np.abs(W[l] - g.W[l]).max() <= tol, ("W", l, float(np.abs(W[l] - g.W[l]).max()), tol)
        assert np.abs(b[l] - g.b[l]).max() <= tol, ("b", l, float(np.abs(b[l] - g.b[l]).max()), tol)
    assert gs == np.float32(len(tr))
    assert abs(float(ev[0]) - float(ev_o[0])) <= LOSS_RTOL * max(1.0, abs(float(ev_o[0])))
    assert abs(int(ev[1]) - ev_o[1]) <= amb[-1]


@pytest.mark.parametrize("batch,steps", [(1024, 5), (4136, 3), (30000, 2), (65536, 2)],
                         ids=["1024", "ragged-4136", "30000", "65536-bench-config"])
def test_distributed_py_model_large_batch(cuda_device, batch, steps):
    """784-100-10 (distributed.py:83-99) at batches where layer 1 runs as the three-plane tensor-core GEMM."""
    dims = [784, 100, 10]
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    want = _oracle(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.001)
    got = _train(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.001)
    _check(got, want)
    # the same steps with both new kernels switched off: the CUDA-core kernels they replace
    ref = _train(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.001, split3=False, narrow=False)
    _check(ref, want)
    np.testing.assert_allclose(got[0], ref[0], rtol=LOSS_RTOL)
    assert np.abs(got[2][0] - ref[2][0]).max() <= W_RTOL * float(np.abs(ref[2][0]).max())


@pytest.mark.parametrize("dims,act", [([512, 64, 10], pkg.ACT_SIGMOID), ([784, 128, 16], pkg.ACT_IDENTITY),
                                      ([260, 36, 3], pkg.ACT_SIGMOID), ([784, 300, 100, 10], pkg.ACT_SIGMOID)],
                         ids=["512-64-10", "784-128-16-identity", "260-36-3", "784-300-100-10"])
def test_other_wide_shapes(cuda_device, dims, act):
    """Shapes around the limits: 128 hidden units / 16 classes (the narrow-head maximum), widths that are not multiples
    of the tile sizes, an identity hidden layer, and a deeper model whose second layer (300 -> 100) is wide as well."""
    batch, steps = 2048 + 24, 3
    x, y10 = synthetic.make_pool(batch * steps)
    x = np.ascontiguousarray(x[:, :dims[0]])
    C = dims[-1]
    lab = np.argmax(y10, axis=1) % C
    y = np.zeros((len(x), C), np.float32)
    y[np.arange(len(x)), lab] = 1.0
    Ws, bs = synthetic.make_weights(dims, True)
    oact = oracle.ACT_SIGMOID if act == pkg.ACT_SIGMOID else oracle.ACT_IDENTITY
    want = _oracle(dims, oact, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.01)
    got = _train(dims, act, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.01)
    _check(got, want)


def test_sse_head_through_narrow_kernel(cuda_device):
    """simple.py's loss (sum of squares, simple.py:20) on a two-layer model: the SSE branch of the narrow-head kernel."""
    dims, batch, steps = [784, 100, 10], 1500, 3
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, True)
    want = _oracle(dims, oracle.ACT_SIGMOID, oracle.LOSS_SSE, Ws, bs, x, y, batch, steps, 1e-6)
    got = _train(dims, pkg.ACT_SIGMOID, pkg.LOSS_SSE, Ws, bs, x, y, batch, steps, 1e-6)
    lt, _, W, b, _, ev = got
    g, tr, ev_o, _ = want
    for s, (lo, _) in enumerate(tr):
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo)))
    for l in range(2):
        assert np.abs(W[l] - g.W[l]).max() <= W_RTOL * max(1.0, float(np.abs(g.W[l]).max()))
        assert np.abs(b[l] - g.b[l]).max() <= W_RTOL * max(1.0, float(np.abs(g.W[l]).max()))
    assert abs(float(ev[0]) - float(ev_o[0])) <= LOSS_RTOL * max(1.0, abs(float(ev_o[0])))


def test_small_batch_uses_narrow_head_only(cuda_device):
    """The reference's shipped batch of 100 (distributed.py:46): below the tensor-core threshold, narrow head on."""
    dims, batch, steps = [784, 100, 10], 100, 20
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, False)
    want = _oracle(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.001)
    _check(_train(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, Ws, bs, x, y, batch, steps, 0.001), want)


def test_sse_head_on_the_fused_single_layer_kernel(cuda_device):
    """784 -> 10 with simple.py's sum-of-squares loss: the SSE branch of the fused single-kernel step, against the
    oracle and against the layered kernels (B2_OPT_FORCE_LAYERED)."""
    dims, batch, steps, lr = [784, 10], 3000, 4, 1e-6
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, True)
    g = oracle.TF1Graph(dims, oracle.ACT_IDENTITY, oracle.LOSS_SSE, Ws, bs)
    want = [g.train_step(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)[0] for s in range(steps)]
    outs = []
    for layered in (False, True):
        tr = pkg.DenseTrainer(dims, pkg.ACT_IDENTITY, pkg.LOSS_SSE, max_local_batch=batch, device=0)
        tr.force_layered(layered)
        tr.set_params(Ws, bs)
        lt, _ = tr.train_steps(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), batch, steps, lr=lr)
        torch.cuda.synchronize()
        W, b = tr.get_params()
        tr.close()
        lt = lt.cpu().numpy()
        for s in range(steps):
            assert abs(float(lt[s]) - float(want[s])) <= LOSS_RTOL * max(1.0, abs(float(want[s])))
        assert np.abs(W[0] - g.W[0]).max() <= W_RTOL * max(1.0, float(np.abs(g.W[0]).max()))
        assert np.abs(b[0] - g.b[0]).max() <= W_RTOL * max(1.0, float(np.abs(g.W[0]).max()))
        outs.append(lt)
    np.testing.assert_allclose(outs[0], outs[1], rtol=LOSS_RTOL)


def test_device_random_normal_matches_the_oracle(cuda_device, tmp_path):
    """b2_random_normal (tf.random_normal restated, distributed.py:85-86) against oracle/tf1_random.py: the Philox
    integers are exact, the Box-Muller floats differ only by the device's logf / sincosf rounding; then save / restore."""
    from oracle import tf1_random
    dims = [784, 100, 10]
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=128, device=0)
    tr.init_random_normal(2, [7, 1234567890123])
    W, b = tr.get_params()
    for l, op_seed in enumerate([7, 1234567890123]):
        want = tf1_random.random_normal((dims[l], dims[l + 1]), 2, op_seed)
        assert np.abs(W[l] - want).max() <= 2e-6, float(np.abs(W[l] - want).max())
        assert not b[l].any()
    p = str(tmp_path / "vars.npz")
    tr.save(p)
    tr2 = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=128, device=0)
    assert tr2.restore(p) == np.float32(0)
    W2, b2 = tr2.get_params()
    for l in range(2):
        np.testing.assert_array_equal(W[l], W2[l])
        np.testing.assert_array_equal(b[l], b2[l])
    tr.close()
    tr2.close()


def test_random_normal_writes_exactly_n_values(cuda_device):
    """A ragged count (not a multiple of the four values per Philox block) must not write past the n-th element."""
    from oracle import tf1_random
    from tensorflow_examples_b200 import _ffi
    lib = _ffi.load_library()
    buf = torch.full((1024,), 7.0, dtype=torch.float32, device="cuda")
    for n in (1, 2, 3, 5, 1001):
        buf.fill_(7.0)
        _ffi.check(lib, lib.b2_random_normal(buf.data_ptr(), n, 2, 99, 0.0, 1.0, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        got = buf.cpu().numpy()
        assert np.abs(got[:n] - tf1_random.random_normal((n,), 2, 99)).max() <= 2e-6
        assert (got[n:] == 7.0).all()
