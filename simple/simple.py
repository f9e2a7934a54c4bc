'''
Drop-in for simple/simple.py of manigoswami/tensorflow-examples: the same scalar linear regression
(W*x + b, sum-of-squares loss, plain gradient descent, 1005 steps, same printed line), with the TF1
session replaced by hand-written sm_100a CUDA kernels behind the C ABI in include/b2mlp.h.

    python simple/simple.py            # exactly as the reference is run

Reference lines are cited as simple.py:N.
'''
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402  (device memory + stream handles only)

import tensorflow_examples_b200 as b2  # noqa: E402
from tensorflow_examples_b200.report import simple_report  # noqa: E402


def main(argv):
    steps = 1005                                              # simple.py:34
    for a in argv[1:]:                                        # optional, the reference takes no arguments
        if a.startswith("--steps="):
            steps = int(a.split("=", 1)[1])
    dev = torch.device("cuda", 0)
    # W = tf.Variable([.3]); b = tf.Variable([-.3]); linear_model = W * x + b      simple.py:11-16
    # loss = reduce_sum(square(linear_model - y)); GradientDescentOptimizer(0.01)  simple.py:20-23
    model = b2.DenseTrainer([1, 1], b2.ACT_IDENTITY, b2.LOSS_SSE, max_local_batch=4)
    model.set_params([np.array([[.3]], np.float32)], [np.array([-.3], np.float32)])
    x_train = torch.tensor([[1], [2], [3], [4]], dtype=torch.float32, device=dev)      # simple.py:25
    y_train = torch.tensor([[0], [-1], [-2], [-3]], dtype=torch.float32, device=dev)   # simple.py:26
    # for i in range(1005): sess.run(train, {x: x_train, y: y_train})             simple.py:34-35
    model.train_steps(x_train, y_train, 4, steps, lr=0.01)
    # curr_W, curr_b, curr_loss = sess.run([W, b, loss], ...)                      simple.py:37
    curr_loss, _ = model.evaluate(x_train, y_train)
    Ws, bs = model.get_params()
    print(simple_report(Ws[0].reshape(1), bs[0].reshape(1), curr_loss))              # simple.py:39
    model.close()


if __name__ == "__main__":
    main(sys.argv)
