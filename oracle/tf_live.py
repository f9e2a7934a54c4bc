"""The reference's own graph on a live TensorFlow, when one is importable (last row).

TEST INFRASTRUCTURE ONLY.  TensorFlow is NOT installed in this image (and cannot be: no network, no TF 1.x wheel for
this Python), so everything here is normally skipped; it exists so that a box that does have TensorFlow (1.x, or 2.x
through tf.compat.v1 with v2 behaviour disabled) pins oracle/tf1_graph.py against the real thing and times it as the
CPU baseline (`bench.py --impl reference` then reports kind "reference" instead of "port").

The graph is the one distributed/distributed.py:77-108,117-118 builds (and simple/simple.py:11-23 for the SSE head),
verbatim op for op, except that the initial values are injected (tf.Variable(initial_value=...)) instead of drawn
(distributed.py:83-86) -- the same injection north_star prescribes for both sides.
"""
from __future__ import annotations

import numpy as np


def available():
    try:
        import tensorflow  # noqa: F401
        return True
    except Exception:
        return False


class LiveGraph:
    def __init__(self, dims, act, loss, Ws, bs, lr):
        import tensorflow as tf
        tf1 = tf.compat.v1 if hasattr(tf, "compat") and hasattr(tf.compat, "v1") else tf
        if hasattr(tf1, "disable_v2_behavior"):
            tf1.disable_v2_behavior()
        self.tf1 = tf1
        self.graph = tf1.Graph()
        with self.graph.as_default():
            self.x = tf1.placeholder(tf1.float32, shape=[None, dims[0]])              # distributed.py:77
            self.y_ = tf1.placeholder(tf1.float32, shape=[None, dims[-1]])            # distributed.py:80
            self.W = [tf1.Variable(np.asarray(w, np.float32)) for w in Ws]             # :85-86 (values injected)
            self.b = [tf1.Variable(np.asarray(b, np.float32)) for b in bs]             # :90-91
            a = self.x
            n = len(Ws)
            for i in range(n):
                z = tf1.add(tf1.matmul(a, self.W[i]), self.b[i])                       # :96, :98
                if i < n - 1:
                    a = tf1.nn.sigmoid(z) if act == 1 else z                           # :97
            if loss == 1:
                y = tf1.nn.softmax(z)                                                  # :99
                self.cost = tf1.reduce_mean(-tf1.reduce_sum(self.y_ * tf1.log(y), reduction_indices=[1]))   # :102
                hit = tf1.equal(tf1.argmax(y, 1), tf1.argmax(self.y_, 1))              # :117
                self.correct = tf1.reduce_sum(tf1.cast(hit, tf1.float32))
            else:
                self.cost = tf1.reduce_sum(tf1.square(z - self.y_))                    # simple.py:20
                self.correct = tf1.constant(0.0)
            self.train = tf1.train.GradientDescentOptimizer(lr).minimize(self.cost)   # :107-108 / simple.py:22-23
            self.sess = tf1.Session(graph=self.graph)
            self.sess.run(tf1.global_variables_initializer())

    def train_step(self, x, y):
        _, c, k = self.sess.run([self.train, self.cost, self.correct], feed_dict={self.x: x, self.y_: y})   # :148-150
        return np.float32(c), int(k + 0.5)

    def variables(self):
        return self.sess.run(self.W), self.sess.run(self.b)
