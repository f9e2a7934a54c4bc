"""CPU oracle: a numpy restatement of the TF 1.x graphs the reference scripts build.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the package, the C-ABI
library, the drop-in scripts) may import this module.  It is used by `tests/`,
by `__graft_entry__.smoke()` and by `bench.py`'s `cpu_baseline` / `--impl
reference` legs, as the checker and as the timed CPU baseline.

PARITY UNPINNED.  The reference (`/root/reference`) holds no arithmetic of its
own: every number comes out of TensorFlow 1.x, which is neither vendored nor
pinned (`README.md:6-8`) and cannot be installed here (no network, no TF1 wheel
for Python 3.12).  The reference ships no tests, golden vectors or fixtures
either.  What this oracle IS pinned against (tests/test_oracle.py):
  * the analytic optimum and the hand-computed step-0 gradient of simple.py;
  * an independent autodiff (torch.autograd on CPU, float64) of the same
    forward graph, which checks every hand-written gradient composition below;
  * its own committed fixtures under tests/golden/ (regression pins).

What is restated, by reference line:
  simple/simple.py:11-12   W=[.3], b=[-.3]                     -> caller passes them in
  simple/simple.py:16      linear_model = W * x + b            -> dense 1->1, identity
  simple/simple.py:20      loss = reduce_sum(square(. - y))    -> head "sse"
  simple/simple.py:22-23   GradientDescentOptimizer(0.01)      -> train_step(lr=0.01)
  distributed/distributed.py:96-97  a2 = sigmoid(x@W1 + b1)    -> act "sigmoid"
  distributed/distributed.py:98     z3 = a2@W2 + b2
  distributed/distributed.py:99     y  = softmax(z3)           -> exp(z-max) * (1/sum)   [TF1]
  distributed/distributed.py:102    mean(-sum(y_*log(y),1))    -> head "softmax_ce", un-fused log(softmax)
  distributed/distributed.py:107-108 SGD lr, global_step += 1  -> train_step
  distributed/distributed.py:117-118 accuracy, first-index argmax ties
The gradient compositions are the ones TF1's Python gradient functions build
(SURVEY.md section 2.2): Mean -> 1/B, Neg, Sum -> tile, Mul, Log -> g*(1/y),
SoftmaxGrad (g - sum(g*y))*y, BiasAdd -> column sums, MatMul -> a^T@g and g@W^T,
SigmoidGrad g*a*(1-a).  dx of the first layer is pruned, as in TF1.

`precision="bf16"` restates the scaled-up configuration (BASELINE.json configs[4]):
every GEMM operand (inputs, activations, weights, output-side gradients) is
rounded to bfloat16 (round-to-nearest-even) before the product, accumulation and
everything else stay float32, master weights stay float32.  Hidden activations and the
gradients w.r.t. hidden pre-activations are STORED in bf16 (so their bias gradients sum the
rounded values); the output layer's dz stays float32 for its bias gradient.
"""
from __future__ import annotations

import numpy as np

ACT_IDENTITY = 0
ACT_SIGMOID = 1
LOSS_SSE = 0
LOSS_SOFTMAX_CE = 1


def bf16_round(a: np.ndarray) -> np.ndarray:
    """Round float32 values to the nearest bfloat16 (ties to even), returned as float32."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    u = a.view(np.uint32).astype(np.uint64)
    rounded = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return rounded.astype(np.uint32).view(np.float32).reshape(a.shape)


def _sigmoid(z):
    # distributed.py:97 tf.nn.sigmoid.  1/(1+exp(-z)) evaluated in the working dtype.
    one = z.dtype.type(1)
    return one / (one + np.exp(-z))


class TF1Graph:
    """The trainable model + optimizer state of either reference script.

    dims  : [in, hidden..., out]        (simple.py: [1, 1]; distributed.py: [784, 100, 10])
    act   : activation of every hidden layer (the last layer is always linear)
    loss  : LOSS_SSE (simple.py:20) or LOSS_SOFTMAX_CE (distributed.py:99-102)
    Ws[i] : [dims[i], dims[i+1]] row-major, the reference's [in, out] layout (distributed.py:85-86)
    bs[i] : [dims[i+1]]
    """

    def __init__(self, dims, act, loss, Ws, bs, dtype=np.float32, precision="fp32"):
        self.dims = list(dims)
        self.act = act
        self.loss = loss
        self.dtype = np.dtype(dtype)
        self.precision = precision
        assert len(Ws) == len(bs) == len(self.dims) - 1
        self.W = [np.array(w, dtype=self.dtype).reshape(self.dims[i], self.dims[i + 1]) for i, w in enumerate(Ws)]
        self.b = [np.array(b, dtype=self.dtype).reshape(self.dims[i + 1]) for i, b in enumerate(bs)]
        # distributed.py:68-70 -- global_step is a float32 scalar variable.
        self.global_step = np.float32(0)

    # ---- helpers -------------------------------------------------------------
    def _q(self, a):
        return bf16_round(a) if self.precision == "bf16" else a

    def _as2d(self, x, width):
        x = np.asarray(x, dtype=self.dtype)
        return x.reshape(-1, width)

    # ---- forward (distributed.py:96-99, simple.py:16) ---------------------------
    def forward(self, x):
        """Returns (activations, z_last): activations[0] = x, activations[i] = input of layer i."""
        a = self._as2d(x, self.dims[0])
        acts = [a]
        n = len(self.W)
        z = None
        for i in range(n):
            z = self._q(a) @ self._q(self.W[i]) + self.b[i]
            if i < n - 1:
                a = _sigmoid(z) if self.act == ACT_SIGMOID else z
                if self.precision == "bf16":
                    a = bf16_round(a)  # activations are stored in bf16 between layers
                acts.append(a)
        return acts, z

    @staticmethod
    def softmax(z):
        # tf.nn.softmax CPU kernel [TF1]: shifted exp, then multiply by the reciprocal of the row sum.
        e = np.exp(z - z.max(axis=1, keepdims=True))
        return e * (z.dtype.type(1) / e.sum(axis=1, keepdims=True))

    def evaluate(self, x, y_):
        """loss and number of correct rows without updating anything (distributed.py:164, simple.py:37)."""
        _, z = self.forward(x)
        y_ = self._as2d(y_, self.dims[-1])
        if self.loss == LOSS_SSE:
            d = z - y_
            return (d * d).sum(dtype=self.dtype), 0
        y = self.softmax(z)
        with np.errstate(divide="ignore", invalid="ignore"):
            per_row = -(y_ * np.log(y)).sum(axis=1)
        loss = per_row.mean(dtype=self.dtype)
        correct = int((np.argmax(y, axis=1) == np.argmax(y_, axis=1)).sum())
        return loss, correct

    # ---- gradients (TF1 gradient compositions) -------------------
    def gradients(self, x, y_, global_batch=None):
        """Returns (loss, correct, dWs, dbs) for the rows given.

        global_batch: the B of reduce_mean when these rows are one shard of a larger
        synchronous batch (data-parallel form).  `loss` is then this shard's
        contribution sum(per_row)/global_batch, so shard losses and gradients add up
        to the single-process values.
        """
        acts, z = self.forward(x)
        y_ = self._as2d(y_, self.dims[-1])
        rows = z.shape[0]
        B = self.dtype.type(global_batch if global_batch is not None else rows)
        if self.loss == LOSS_SSE:
            d = z - y_
            loss = (d * d).sum(dtype=self.dtype)
            correct = 0
            dz = self.dtype.type(2) * d  # Square grad: 2*(pred-y) * upstream(=1)
        else:
            y = self.softmax(z)
            with np.errstate(divide="ignore", invalid="ignore"):
                logy = np.log(y)
                per_row = -(y_ * logy).sum(axis=1)
                loss = per_row.sum(dtype=self.dtype) / B
                correct = int((np.argmax(y, axis=1) == np.argmax(y_, axis=1)).sum())
                g_log = -(np.ones_like(y) / B) * y_            # Mean, Neg, Sum(tile), Mul
                g_y = g_log * (self.dtype.type(1) / y)          # Log grad: g * reciprocal(y)
                dz = (g_y - (g_y * y).sum(axis=1, keepdims=True)) * y   # SoftmaxGrad
        n = len(self.W)
        dWs = [None] * n
        dbs = [None] * n
        for i in range(n - 1, -1, -1):
            a_in = acts[i]
            dbs[i] = dz.sum(axis=0, dtype=self.dtype)                       # BiasAdd grad
            dWs[i] = self._q(a_in).T @ self._q(dz)                          # MatMul grad (transpose_a)
            if i > 0:
                da = self._q(dz) @ self._q(self.W[i]).T                     # MatMul grad (transpose_b)
                if self.act == ACT_SIGMOID:
                    dz = da * a_in * (self.dtype.type(1) - a_in)            # SigmoidGrad
                else:
                    dz = da
                if self.precision == "bf16":
                    dz = bf16_round(dz)   # hidden-layer gradients are stored in bf16, like the activations
        return loss, correct, dWs, dbs

    def apply(self, dWs, dbs, lr):
        # ApplyGradientDescent: var -= lr * grad, then global_step += 1 (distributed.py:107-108)
        lr = self.dtype.type(lr)
        for i in range(len(self.W)):
            self.W[i] -= lr * dWs[i].astype(self.dtype)
            self.b[i] -= lr * dbs[i].astype(self.dtype)
        self.global_step = np.float32(self.global_step + np.float32(1))

    def train_step(self, x, y_, lr):
        """One sess.run(train_op, ...) (simple.py:35, distributed.py:148-150).

        Returns the PRE-update loss and correct count, as the reference's fetches do.
        """
        loss, correct, dWs, dbs = self.gradients(x, y_)
        self.apply(dWs, dbs, lr)
        return loss, correct

    def train_step_sharded(self, x, y_, lr, nranks):
        """The synchronous data-parallel form: rows split evenly over `nranks`, shard
        gradients summed in rank order (what the allreduce produces), one update."""
        x = self._as2d(x, self.dims[0])
        y_ = self._as2d(y_, self.dims[-1])
        rows = x.shape[0]
        assert rows % nranks == 0
        per = rows // nranks
        tot = None
        for r in range(nranks):
            part = self.gradients(x[r * per:(r + 1) * per], y_[r * per:(r + 1) * per], global_batch=rows)
            if tot is None:
                tot = [part[0], part[1], list(part[2]), list(part[3])]
            else:
                tot[0] = tot[0] + part[0]
                tot[1] += part[1]
                tot[2] = [a + b for a, b in zip(tot[2], part[2])]
                tot[3] = [a + b for a, b in zip(tot[3], part[3])]
        self.apply(tot[2], tot[3], lr)
        return tot[0], tot[1]


# ---- the two scripts as shipped ---------------------------------------------------
def simple_py(steps=1005, dtype=np.float32):
    """simple/simple.py:11-39 end to end.  Returns (W, b, final_loss, pre_update_loss_trace)."""
    g = TF1Graph([1, 1], ACT_IDENTITY, LOSS_SSE, [[0.3]], [[-0.3]], dtype=dtype)
    x = np.array([1, 2, 3, 4], dtype=dtype)          # simple.py:25
    y = np.array([0, -1, -2, -3], dtype=dtype)       # simple.py:26
    trace = []
    for _ in range(steps):                           # simple.py:34-35
        loss, _ = g.train_step(x, y, 0.01)
        trace.append(loss)
    final, _ = g.evaluate(x, y)                      # simple.py:37
    return g.W[0].reshape(1), g.b[0].reshape(1), final, trace


def simple_py_report(steps=1005):
    """The line simple/simple.py:39 prints."""
    W, b, loss, _ = simple_py(steps)
    return "W: %s b: %s loss %s" % (W, b, loss)
