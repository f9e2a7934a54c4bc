"""GPU tests of the tcgen05 bf16 GEMM (b2_matmul_bf16) against torch on the same bf16-rounded operands.
Tolerance: fp32 accumulation order differs, so |gpu - ref| <= 2e-3 * sqrt(K)-scaled magnitude for fp32 outputs and
one bf16 ulp-ish (1/128 relative) for bf16 outputs."""
import ctypes

import pytest
import torch

import tensorflow_examples_b200 as pkg

pytestmark = pytest.mark.gpu


def _call(lib, A, a_mn, B, b_mn, M, N, K, epi, bias, act, out):
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b2_matmul_bf16(p(A), a_mn, p(B), b_mn, M, N, K, epi, p(bias), p(act), p(out), stream)
    assert rc == 0, lib.b2_last_error()
    torch.cuda.synchronize()


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (1000, 768, 784), (784, 4096, 1024), (4096, 4096, 512)])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 1), (0, 0), (1, 1)], ids=["fwd", "dgrad", "wgrad"])
def test_matmul_layouts_fp32_out(cuda_device, M, N, K, a_mn, b_mn):
    lib = pkg.load_library()
    g = torch.Generator(device=cuda_device).manual_seed(1)
    A = torch.randn(M, K, device=cuda_device, generator=g).to(torch.bfloat16)
    B = torch.randn(K, N, device=cuda_device, generator=g).to(torch.bfloat16)
    ref = A.float() @ B.float()
    Ag = A.t().contiguous() if a_mn else A.contiguous()          # [K,M] when MN-major
    Bg = B.contiguous() if b_mn else B.t().contiguous()          # [K,N] when MN-major, [N,K] when K-major
    out = torch.empty(M, N, dtype=torch.float32, device=cuda_device)
    _call(lib, Ag, a_mn, Bg, b_mn, M, N, K, 2, None, None, out)
    err = (out - ref).abs().max().item()
    assert err <= 2e-3 * (K ** 0.5), (err, ref.abs().max().item())


def test_matmul_epilogues(cuda_device):
    lib = pkg.load_library()
    M, N, K = 512, 1024, 784
    g = torch.Generator(device=cuda_device).manual_seed(2)
    A = (torch.randn(M, K, device=cuda_device, generator=g) * 0.1).to(torch.bfloat16)
    W = (torch.randn(K, N, device=cuda_device, generator=g) * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device=cuda_device, generator=g)
    z = A.float() @ W.float() + bias
    # 0: sigmoid(acc + bias) -> bf16
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    _call(lib, A, 0, W, 1, M, N, K, 0, bias, None, out)
    assert (out.float() - torch.sigmoid(z)).abs().max().item() <= 1.0 / 128
    # 3: acc + bias -> fp32
    out32 = torch.empty(M, N, dtype=torch.float32, device=cuda_device)
    _call(lib, A, 0, W, 1, M, N, K, 3, bias, None, out32)
    assert (out32 - z).abs().max().item() <= 2e-3
    # 1: acc * a * (1 - a) -> bf16      (dA = dZ . W^T with dZ [M,N], W [K,N] K-major over N)
    dZ = (torch.randn(M, N, device=cuda_device, generator=g) * 0.1).to(torch.bfloat16)
    act = torch.rand(M, K, device=cuda_device, generator=g).to(torch.bfloat16)
    ref = (dZ.float() @ W.float().t()) * act.float() * (1 - act.float())
    dA = torch.empty(M, K, dtype=torch.bfloat16, device=cuda_device)
    _call(lib, dZ, 0, W, 0, M, K, N, 1, None, act, dA)
    assert (dA.float() - ref).abs().max().item() <= ref.abs().max().item() / 64 + 1e-3


def test_matmul_throughput_report(cuda_device):
    """Not a pass/fail on speed: prints the achieved TFLOP/s of the three layer products at the scaled-up
    configuration's shapes (one launch each, CUDA events) so a GPU run leaves the numbers in its log."""
    lib = pkg.load_library()
    M, N, K = 65536, 4096, 4096
    A = torch.randn(M, K, device=cuda_device).to(torch.bfloat16)
    W = torch.randn(K, N, device=cuda_device).to(torch.bfloat16)
    dZ = torch.randn(M, N, device=cuda_device).to(torch.bfloat16)
    bias = torch.zeros(N, device=cuda_device)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    dW = torch.empty(K, N, dtype=torch.float32, device=cuda_device)
    cases = [("fwd", (A, 0, W, 1, M, N, K, 0, bias, None, out)),
             ("dgrad", (dZ, 0, W, 0, M, K, N, 1, None, A, out)),
             ("wgrad", (A, 1, dZ, 1, K, N, M, 2, None, None, dW))]
    for name, args in cases:
        _call(lib, *args)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            _call(lib, *args)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print("tcgen05 gemm %s: %.3f ms, %.0f TFLOP/s" % (name, ms, 2.0 * M * N * K / ms / 1e9))
