"""GPU test of b2_linear_step_noalloc, the entry point for callers that cannot allocate device memory
(plugins inside another framework's callback, latency-critical loops, hosts whose pool owns the GPU).

It is called through the C ABI exactly as such a caller would -- raw device pointers, nothing created
by the library -- and checked against the allocating path (b2_create + b2_train_steps) on the same
inputs, for two calls in a row: the second call reuses the library's built-in scratch array with the
first call's leftovers still in it, which is the case that would expose a stale-scratch bug.
Tolerances are the ones tests/test_gpu_parity.py states (loss 1e-5 relative, weights 1e-5 of max|W|);
the two paths run the same step kernel, so in practice they agree bit for bit.

Named test_zz_* so pytest collects it last: a failure here never hides another test's result.
"""
import ctypes

import numpy as np
import pytest
import torch

import tensorflow_examples_b200 as pkg
from oracle import tf1_graph as oracle
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-5
W_RTOL = 1e-5
MAX_ROWS = 1 << 20


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.parametrize("batch", [256, 1000])
def test_noalloc_step_matches_allocating_path_twice_in_a_row(cuda_device, batch):
    lib = pkg.load_library()
    dims, lr, steps = [784, 10], 0.001, 2
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled=True)
    xd, yd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(y).to(cuda_device)

    # the existing entry points: context-owned parameters and scratch
    ref = pkg.DenseTrainer(dims, pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, max_local_batch=batch)
    ref.set_params(Ws, bs)
    cpu = oracle.TF1Graph(dims, oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, Ws, bs)     # and the CPU oracle beside both

    # the caller-owned buffers of the no-allocation path
    W = torch.from_numpy(Ws[0]).to(cuda_device).contiguous()
    b = torch.from_numpy(bs[0]).to(cuda_device).contiguous()
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_device)
    correct = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda_device).cuda_stream)

    for s in range(steps):
        lt, ct = ref.train_steps(xd, yd, batch, 1, first_row=s * batch, lr=lr)
        xs, ys = xd[s * batch:(s + 1) * batch], yd[s * batch:(s + 1) * batch]
        rc = lib.b2_linear_step_noalloc(_ptr(xs), _ptr(ys), batch, batch, lr, _ptr(W), _ptr(b), _ptr(loss),
                                        _ptr(correct), stream)
        assert rc == 0, lib.b2_last_error()
        torch.cuda.synchronize(cuda_device)
        want_loss, want_correct = float(lt[0].item()), int(ct[0].item())
        assert abs(float(loss.item()) - want_loss) <= LOSS_RTOL * max(1.0, abs(want_loss)), (s, loss.item(), want_loss)
        assert int(correct.item()) == want_correct, (s, int(correct.item()), want_correct)
        rW, rb = ref.get_params()
        tol = W_RTOL * max(1.0, float(np.abs(rW[0]).max()))
        assert np.abs(W.cpu().numpy() - rW[0]).max() <= tol, s
        assert np.abs(b.cpu().numpy() - rb[0]).max() <= tol, s
        lo, co = cpu.train_step(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)
        assert abs(float(loss.item()) - float(lo)) <= LOSS_RTOL * max(1.0, abs(float(lo))), (s, loss.item(), float(lo))
        assert abs(int(correct.item()) - co) <= 1, (s, int(correct.item()), co)
        assert np.abs(W.cpu().numpy() - cpu.W[0]).max() <= tol and np.abs(b.cpu().numpy() - cpu.b[0]).max() <= tol, s
    ref.close()


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(cuda_device):
    lib = pkg.load_library()
    t = torch.zeros(16, device=cuda_device)
    rc = lib.b2_linear_step_noalloc(_ptr(t), _ptr(t), MAX_ROWS + 1, MAX_ROWS + 1, 0.001, _ptr(t), _ptr(t), None,
                                    None, None)
    assert rc == 1                                             # B2_ERR_ARG, nothing was launched
    assert b"exceeds the fixed-scratch limit" in lib.b2_last_error()
    rc = lib.b2_linear_step_noalloc(None, _ptr(t), 8, 8, 0.001, _ptr(t), _ptr(t), None, None, None)
    assert rc == 1 and b"NULL" in lib.b2_last_error()
    # a view at an odd float offset: the rows are fetched with 16-byte bulk copies, so this is refused, not run
    big = torch.zeros(8 * 784 + 8, device=cuda_device)
    rc = lib.b2_linear_step_noalloc(ctypes.c_void_p(big.data_ptr() + 4), _ptr(t), 8, 8, 0.001, _ptr(t), _ptr(t), None,
                                    None, None)
    assert rc == 1 and b"16-byte aligned" in lib.b2_last_error()
    torch.cuda.synchronize(cuda_device)
