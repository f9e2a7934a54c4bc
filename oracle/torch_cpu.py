"""The oracle's graph again, on torch CPU tensors, so the timed CPU baseline can use every host core
(MKL GEMMs and multi-threaded element-wise ops, as TF1's Eigen thread pool would).

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/tf1_graph.py for the rules and the reference lines it
follows; PARITY UNPINNED applies here too).  tests/test_oracle.py checks it against the numpy oracle.
"""
import torch

from .tf1_graph import ACT_SIGMOID, LOSS_SSE


class TorchCPUGraph:
    def __init__(self, dims, act, loss, Ws, bs, threads=None):
        if threads:
            torch.set_num_threads(int(threads))
        self.dims, self.act, self.loss = list(dims), act, loss
        self.W = [torch.tensor(w, dtype=torch.float32).reshape(dims[i], dims[i + 1]).clone() for i, w in enumerate(Ws)]
        self.b = [torch.tensor(b, dtype=torch.float32).reshape(dims[i + 1]).clone() for i, b in enumerate(bs)]
        self.global_step = 0.0

    @torch.no_grad()
    def train_step(self, x, y_, lr):
        """One sess.run(train_op) (distributed.py:148-150): returns (pre-update loss, correct rows)."""
        a = x.reshape(-1, self.dims[0])
        y_ = y_.reshape(-1, self.dims[-1])
        acts = [a]
        n = len(self.W)
        for i in range(n):
            z = torch.addmm(self.b[i], a, self.W[i])                       # distributed.py:96,98
            if i < n - 1:
                a = torch.sigmoid(z) if self.act == ACT_SIGMOID else z      # distributed.py:97
                acts.append(a)
        B = z.shape[0]
        if self.loss == LOSS_SSE:
            d = z - y_
            loss = (d * d).sum()                                            # simple.py:20
            correct = 0
            dz = 2.0 * d
        else:
            e = torch.exp(z - z.max(dim=1, keepdim=True).values)
            y = e * (1.0 / e.sum(dim=1, keepdim=True))                      # distributed.py:99
            loss = (-(y_ * torch.log(y)).sum(dim=1)).mean()                 # distributed.py:102
            correct = int((y.argmax(dim=1) == y_.argmax(dim=1)).sum())      # distributed.py:117
            g_y = (-(1.0 / B) * y_) * (1.0 / y)
            dz = (g_y - (g_y * y).sum(dim=1, keepdim=True)) * y
        for i in range(n - 1, -1, -1):
            a_in = acts[i]
            db = dz.sum(dim=0)
            dW = a_in.t() @ dz
            if i > 0:
                da = dz @ self.W[i].t()
                dz = da * a_in * (1.0 - a_in) if self.act == ACT_SIGMOID else da
            self.W[i].sub_(dW, alpha=lr)                                    # distributed.py:107-108
            self.b[i].sub_(db, alpha=lr)
        self.global_step += 1.0
        return float(loss), correct
