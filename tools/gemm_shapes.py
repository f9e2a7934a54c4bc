import ctypes, sys, torch
sys.path.insert(0, '.')
import tensorflow_examples_b200 as pkg
lib = pkg.load_library()
dev = torch.device('cuda')
def call(A, a_mn, B, b_mn, M, N, K, epi, bias, act, out):
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    rc = lib.b2_matmul_bf16(p(A), a_mn, p(B), b_mn, M, N, K, epi, p(bias), p(act), p(out), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.b2_last_error()
def bench(name, args, flops):
    call(*args); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): call(*args)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("%-28s %.3f ms  %.0f TFLOP/s" % (name, ms, flops / ms / 1e9))
B = 65536
x = torch.randn(B, 784, device=dev).to(torch.bfloat16)
W1 = torch.randn(784, 4096, device=dev).to(torch.bfloat16)
a1 = torch.rand(B, 4096, device=dev).to(torch.bfloat16)
dz = torch.randn(B, 4096, device=dev).to(torch.bfloat16)
bias = torch.zeros(4096, device=dev)
out = torch.empty(B, 4096, dtype=torch.bfloat16, device=dev)
dW1 = torch.empty(784, 4096, dtype=torch.float32, device=dev)
bench("fwd1 65536x4096x784", (x, 0, W1, 1, B, 4096, 784, 0, bias, None, out), 2.0 * B * 4096 * 784)
bench("wgrad1 784x4096x65536", (x, 1, dz, 1, 784, 4096, B, 2, None, None, dW1), 2.0 * B * 4096 * 784)
W3 = torch.randn(4096, 64, device=dev).to(torch.bfloat16)
dz3 = torch.randn(B, 64, device=dev).to(torch.bfloat16)
bench("dzprev 65536x4096x64", (dz3, 0, W3, 0, B, 4096, 64, 1, None, a1, out), 2.0 * B * 4096 * 64)
lg = torch.empty(B, 64, dtype=torch.float32, device=dev)
b64 = torch.zeros(64, device=dev)
bench("fwd3 65536x64x4096", (a1, 0, W3, 1, B, 64, 4096, 3, b64, None, lg), 2.0 * B * 4096 * 64)
