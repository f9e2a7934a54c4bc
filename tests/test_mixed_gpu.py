"""GPU tests of the bf16-operand / fp32-accumulate path (tcgen05 GEMMs for the hidden layers), the
configuration BASELINE.json configs[4] names.  Oracle: oracle/tf1_graph.py with precision="bf16" (every
GEMM operand rounded to bf16, fp32 accumulation, fp32 master weights).

Stated tolerances: per-step loss within 2e-2 relative of the bf16-rounded
oracle (observed ~1e-4); correct-row count identical except rows whose top-2 logit margin is below 2e-2;
weights after the trace within 2e-3 * max|W| (a bf16 ulp of an activation moves a gradient by ~0.4 %).
"""
import numpy as np
import pytest
import torch

import tensorflow_examples_b200 as pkg
from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
LOSS_RTOL = 2e-2


def _ambiguous(g, x, margin):
    _, z = g.forward(x)
    top2 = np.sort(z, axis=1)[:, -2:]
    return int(((top2[:, 1] - top2[:, 0]) < margin).sum())


@pytest.mark.parametrize("dims,batch", [([784, 512, 256, 10], 1024), ([64, 128, 10], 200), ([784, 256, 256, 256, 10], 384)])
def test_bf16_path_matches_bf16_oracle(cuda_device, dims, batch):
    steps, lr = 5, 0.001
    x, y = synthetic.make_pool(batch * steps, dims[0], dims[-1])
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, precision="bf16")
    tr.set_params(Ws, bs)
    lt, ct = tr.train_steps(torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device), batch, steps, lr=lr)
    torch.cuda.synchronize()
    gW, gb = tr.get_params()
    lt, ct = lt.cpu().numpy(), ct.cpu().numpy()
    g = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, precision="bf16")
    for s in range(steps):
        sl = slice(s * batch, (s + 1) * batch)
        amb = _ambiguous(g, x[sl], 2e-2)
        lo, co = g.train_step(x[sl], y[sl], lr)
        assert abs(float(lt[s]) - float(lo)) <= LOSS_RTOL * abs(float(lo)), (s, lt[s], lo)
        assert abs(float(lt[s]) - float(lo)) <= 2e-3 * abs(float(lo)), ("observed accuracy regressed", s, lt[s], lo)
        assert abs(int(ct[s]) - co) <= amb, (s, int(ct[s]), co, amb)
    for l in range(len(dims) - 1):
        tol = 2e-3 * float(np.abs(g.W[l]).max())
        assert np.abs(gW[l] - g.W[l]).max() <= tol, ("W", l, np.abs(gW[l] - g.W[l]).max(), tol)
        assert np.abs(gb[l] - g.b[l]).max() <= tol + 1e-6, ("b", l)
    assert tr.global_step == np.float32(steps)
    tr.close()


def test_bf16_path_tracks_fp32_path(cuda_device):
    dims, batch, steps, lr = [784, 1024, 1024, 10], 4096, 4, 0.001
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    out = {}
    for prec in ("fp32", "bf16"):
        tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, precision=prec)
        tr.set_params(Ws, bs)
        lt, ct = tr.train_steps(xd, yd, batch, steps, lr=lr)
        ev = tr.evaluate(xd[:batch], yd[:batch])
        torch.cuda.synchronize()
        out[prec] = (lt.cpu().numpy(), ct.cpu().numpy(), ev)
        tr.close()
    np.testing.assert_allclose(out["bf16"][0], out["fp32"][0], rtol=LOSS_RTOL)
    assert abs(out["bf16"][2][0] - out["fp32"][2][0]) <= LOSS_RTOL * abs(out["fp32"][2][0])
    assert np.abs(out["bf16"][1] - out["fp32"][1]).max() <= 0.02 * batch


def test_c5_full_size_runs_and_matches_oracle_forward(cuda_device):
    """784-4096-4096-10 at batch 65536 (BASELINE.json configs[4]): two training steps run, the loss is finite
    and moves down, and the forward pass agrees with the bf16 oracle on an 2048-row sample (the CPU oracle needs
    ~1 s per 1000 rows at this width)."""
    dims, batch = [784, 4096, 4096, 10], 65536
    x, y = synthetic.make_pool(batch)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, precision="bf16")
    tr.set_params(Ws, bs)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    n = 2048
    loss, correct = tr.evaluate(xd[:n].contiguous(), yd[:n].contiguous())
    g = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, precision="bf16")
    lo, co = g.evaluate(x[:n], y[:n])
    assert abs(float(loss) - float(lo)) <= 2e-3 * abs(float(lo)), (loss, lo)
    assert abs(correct - co) <= _ambiguous(g, x[:n], 2e-2)
    lt, ct = tr.train_steps(xd, yd, batch, 3, lr=0.001)
    torch.cuda.synchronize()
    lt = lt.cpu().numpy()
    assert np.isfinite(lt).all() and lt[2] < lt[0], lt
    assert abs(float(lt[0]) - float(np.log(10))) < 0.5          # scaled init: near-uniform predictions
    tr.close()


def test_c5_full_width_training_step_matches_bf16_oracle(cuda_device):
    """One TRAINING step of 784-4096-4096-10 (full widths: the 4096-wide tcgen05 wgrad / dgrad tiles) at batch 2048
    against the bf16 oracle, including every weight matrix afterwards.  lr is large so that the update (lr * dW) stands
    well clear of the fp32 rounding of W - lr*dW: the comparison is then a comparison of the gradients themselves."""
    dims, batch, lr = [784, 4096, 4096, 10], 2048, 4.0
    x, y = synthetic.make_pool(batch)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, precision="bf16")
    tr.set_params(Ws, bs)
    lt, ct = tr.train_steps(torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device), batch, 1, lr=lr)
    torch.cuda.synchronize()
    gW, gb = tr.get_params()
    tr.close()
    g = oracle.TF1Graph(dims, oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, Ws, bs, precision="bf16")
    amb = _ambiguous(g, x, 2e-2)
    lo, co = g.train_step(x, y, lr)
    assert abs(float(lt[0].item()) - float(lo)) <= 2e-3 * abs(float(lo))
    assert abs(int(ct[0].item()) - co) <= amb
    for l in range(3):
        upd_o, upd_g = g.W[l] - Ws[l], gW[l] - Ws[l]                 # -lr * dW
        scale = float(np.abs(upd_o).max())
        assert scale > 0
        # bf16 operands: one ulp of an activation or of a dz moves a product by 0.4 %; sums of 2048 of them agree far better
        assert np.abs(upd_g - upd_o).max() <= 2e-2 * scale, ("dW", l, float(np.abs(upd_g - upd_o).max()), scale)
        rel = np.linalg.norm(upd_g - upd_o) / np.linalg.norm(upd_o)
        assert rel <= 3e-3, ("dW", l, float(rel))
        bscale = float(np.abs(g.b[l] - bs[l]).max())
        assert np.abs((gb[l] - bs[l]) - (g.b[l] - bs[l])).max() <= 2e-2 * bscale + 1e-7, ("db", l)


def test_c5_full_size_gradient_is_the_sum_of_its_sub_batch_gradients(cuda_device):
    """A size-independent property at BASELINE.json's full size (65536 rows), where the CPU oracle is too slow: the mean
    gradient over the batch equals the mean of the 16 mean gradients of its 4096-row pieces, all taken from the same
    initial weights.  The divisors are powers of two, so every bf16-rounded dz of the big step is exactly 1/16 of the
    piece's; only the fp32 accumulation order (batch as the K dimension of the wgrad GEMM: 65536 vs 4096 deep) differs."""
    dims, batch, pieces, lr = [784, 4096, 4096, 10], 65536, 16, 64.0
    x, y = synthetic.make_pool(batch)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    tr = pkg.DenseTrainer(dims, pkg.ACT_SIGMOID, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch, precision="bf16")
    W0 = [torch.from_numpy(w).to(cuda_device) for w in Ws]

    def update(rows, first_row):
        tr.set_params(Ws, bs)
        lt, _ = tr.train_steps(xd, yd, rows, 1, first_row=first_row, lr=lr)
        torch.cuda.synchronize()
        out = []
        for l in range(3):
            W = torch.empty_like(W0[l])
            b = torch.empty(dims[l + 1], dtype=torch.float32, device=cuda_device)
            pkg._ffi.check(tr._lib, tr._lib.b2_params_get(tr._h, l, W.data_ptr(), b.data_ptr(), tr._stream()))
            torch.cuda.synchronize()
            out.append((W - W0[l]).double())
        return out, float(lt[0].item())

    big, loss_big = update(batch, 0)
    acc, loss_acc = [torch.zeros_like(b) for b in big], 0.0
    sub = batch // pieces
    for i in range(pieces):
        u, lo = update(sub, i * sub)
        loss_acc += lo
        for l in range(3):
            acc[l] += u[l]
    assert abs(loss_big - loss_acc / pieces) <= 1e-5 * abs(loss_big)
    for l in range(3):
        mean = acc[l] / pieces
        scale = float(big[l].abs().max())
        diff = float((big[l] - mean).abs().max())
        assert scale > 0 and diff <= 2e-3 * scale, (l, diff, scale)
    tr.close()
