"""Pins the oracle against TensorFlow itself -- only where TensorFlow is importable (it is not in this image, so the
whole module is skipped here; SURVEY.md section 8(c): "If TF turns up on the GPU box")."""
import numpy as np
import pytest

from oracle import tf1_graph as oracle
from oracle import tf_live
from tensorflow_examples_b200 import synthetic

pytestmark = pytest.mark.skipif(not tf_live.available(), reason="TensorFlow is not installed (parity with TF stays unpinned)")


@pytest.mark.parametrize("dims,act,loss,scaled", [([784, 100, 10], 1, 1, False), ([784, 10], 0, 1, True), ([1, 1], 0, 0, True)])
def test_oracle_matches_live_tensorflow(dims, act, loss, scaled):
    batch, steps, lr = 100, 10, 0.001
    if dims == [1, 1]:
        x = np.array([[1], [2], [3], [4]], np.float32)            # simple.py:25-26
        y = np.array([[0], [-1], [-2], [-3]], np.float32)
        Ws, bs, batch, lr = [np.array([[0.3]], np.float32)], [np.array([-0.3], np.float32)], 4, 0.01
        x, y = np.tile(x, (steps, 1)), np.tile(y, (steps, 1))
    else:
        x, y = synthetic.make_pool(batch * steps)
        Ws, bs = synthetic.make_weights(dims, scaled)
    live = tf_live.LiveGraph(dims, act, loss, Ws, bs, lr)
    g = oracle.TF1Graph(dims, act, loss, Ws, bs)
    for s in range(steps):
        sl = slice(s * batch, (s + 1) * batch)
        lc, lk = live.train_step(x[sl], y[sl])
        oc, ok = g.train_step(x[sl], y[sl], lr)
        assert abs(float(lc) - float(oc)) <= 1e-5 * max(1.0, abs(float(oc)))
        assert lk == ok or loss == 0
    W, b = live.variables()
    for l in range(len(W)):
        assert np.abs(W[l] - g.W[l]).max() <= 1e-5 * max(1.0, float(np.abs(g.W[l]).max()))
        assert np.abs(b[l] - g.b[l]).max() <= 1e-5 * max(1.0, float(np.abs(g.W[l]).max()))
