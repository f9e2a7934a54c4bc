"""Generates tests/golden/*.npz from the CPU oracle (oracle/tf1_graph.py).

The reference cannot run here (TensorFlow 1.x is not installable in this image), so these are
regression pins of the restatement, not outputs of the reference: PARITY UNPINNED, see the oracle
header.  Inputs are regenerated from fixed seeds (tensorflow_examples_b200.synthetic); only the
outputs are stored.  Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import tf1_graph as oracle  # noqa: E402
from tensorflow_examples_b200 import synthetic  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (dims, act, loss, scaled_init, batch, steps, lr)
    "c2_784_100_10_b100": ([784, 100, 10], oracle.ACT_SIGMOID, oracle.LOSS_SOFTMAX_CE, False, 100, 20, 0.001),
    "c3_784_10_b256": ([784, 10], oracle.ACT_IDENTITY, oracle.LOSS_SOFTMAX_CE, True, 256, 20, 0.001),
}


def run_case(dims, act, loss, scaled, batch, steps, lr, dtype):
    x, y = synthetic.make_pool(batch * steps)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = oracle.TF1Graph(dims, act, loss, Ws, bs, dtype=dtype)
    losses, corrects = [], []
    for s in range(steps):
        lo, co = g.train_step(x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch], lr)
        losses.append(lo)
        corrects.append(co)
    return g, np.array(losses, dtype=dtype), np.array(corrects, dtype=np.int64)


def main():
    W, b, final, trace = oracle.simple_py()
    W64, b64, final64, _ = oracle.simple_py(dtype=np.float64)
    np.savez(os.path.join(HERE, "simple_py.npz"), W=W, b=b, loss=np.float32(final),
             trace=np.array(trace, np.float32), W64=W64, b64=b64, loss64=final64,
             report=np.array(oracle.simple_py_report()))
    for name, (dims, act, loss, scaled, batch, steps, lr) in CASES.items():
        g32, l32, c32 = run_case(dims, act, loss, scaled, batch, steps, lr, np.float32)
        g64, l64, c64 = run_case(dims, act, loss, scaled, batch, steps, lr, np.float64)
        out = {"loss32": l32, "correct32": c32, "loss64": l64, "correct64": c64}
        for l in range(len(dims) - 1):
            out["b%d" % l] = g32.b[l]
            out["W%d_rows16" % l] = g32.W[l][::16].copy()
            out["W%d_sum64" % l] = np.float64(g64.W[l].sum())
        np.savez(os.path.join(HERE, name + ".npz"), **out)
        print(name, l32[:3], c32[:3])


if __name__ == "__main__":
    main()
