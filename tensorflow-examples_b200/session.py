"""Host-side mirror of the reference's training interface.

The reference drives its model through `sess.run(fetches, feed_dict)` (simple/simple.py:35,37;
distributed/distributed.py:148-150,164).  `DenseTrainer` offers the same three operations --
a training step that returns the pre-update cost, an evaluation that returns cost/accuracy, and a
fetch of the variables -- on top of the C ABI (include/b2mlp.h).  torch is used for device memory
and the current stream handle only.
"""
import ctypes
import os

import numpy as np

from . import _ffi

ACT_IDENTITY, ACT_SIGMOID = 0, 1
LOSS_SSE, LOSS_SOFTMAX_CE = 0, 1
_PREC = {"fp32": 0, "bf16": 1}


def _torch():
    import torch
    return torch


class DenseTrainer:
    """Model + plain gradient-descent optimizer state on one GPU.

    dims : [in, hidden..., out]; act: activation of the hidden layers; loss: LOSS_SSE (simple.py:20)
    or LOSS_SOFTMAX_CE (distributed.py:99-102).  Weights use the reference's [in, out] layout.
    """

    def __init__(self, dims, act, loss, max_local_batch, device=0, precision="fp32"):
        torch = _torch()
        if not torch.cuda.is_available():
            raise RuntimeError("DenseTrainer needs a CUDA device (B200); there is no CPU fallback")
        self._lib = _ffi.load_library()
        self.dims = [int(d) for d in dims]
        self.n_layers = len(self.dims) - 1
        self.act, self.loss = int(act), int(loss)
        self.device = int(device)
        self.torch_device = torch.device("cuda", self.device)
        self.max_local_batch = int(max_local_batch)
        self.rank, self.nranks = 0, 1
        arr = (ctypes.c_int * len(self.dims))(*self.dims)
        h = ctypes.c_void_p()
        _ffi.check(self._lib, self._lib.b2_create(ctypes.byref(h), self.device, self.n_layers, arr, self.act,
                                                  self.loss, _PREC[precision], self.max_local_batch))
        self._h = h

    # -- life cycle --------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.b2_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        torch = _torch()
        return ctypes.c_void_p(torch.cuda.current_stream(self.torch_device).cuda_stream)

    def set_option(self, option, value):
        _ffi.check(self._lib, self._lib.b2_set_option(self._h, int(option), int(value)))

    def force_layered(self, on=True):
        """Use the per-layer kernels even where a fused single-kernel step exists (for cross-checks)."""
        self.set_option(_ffi.OPT_FORCE_LAYERED, 1 if on else 0)

    def kernel_timing(self, on=True):
        """Bracket every kernel launch with CUDA events (read them back with kernel_times())."""
        self.set_option(_ffi.OPT_KERNEL_TIMING, 1 if on else 0)

    def kernel_times(self):
        """{kernel class: (total ms, launches)} since the last call.  Synchronises."""
        ms = (ctypes.c_float * _ffi.KERNEL_CLASSES)()
        n = (ctypes.c_long * _ffi.KERNEL_CLASSES)()
        _ffi.check(self._lib, self._lib.b2_kernel_times(self._h, ms, n))
        return {self._lib.b2_kernel_class_name(k).decode(): (float(ms[k]), int(n[k]))
                for k in range(_ffi.KERNEL_CLASSES) if n[k]}

    @property
    def param_count(self):
        return int(self._lib.b2_param_count(self._h))

    @property
    def exchange_floats(self):
        """Floats one gradient exchange carries: the flat parameter layout (every W / b segment padded to 16 bytes)
        plus loss and correct count.  The peer-memory exchange takes at most 2**18 of them."""
        r4 = lambda v: (v + 3) & ~3
        return sum(r4(self.dims[l] * self.dims[l + 1]) + r4(self.dims[l + 1]) for l in range(self.n_layers)) + 4

    @property
    def global_step(self):
        """distributed.py:68-70: a float32 counter of applied updates."""
        return np.float32(self._lib.b2_global_step(self._h))

    @property
    def launch_count(self):
        return int(self._lib.b2_launch_count(self._h))

    # -- variables ----------------------------------------------------------------------------
    def set_params(self, Ws, bs):
        """Load initial values (replaces sess.run(init_op), distributed.py:126,135; simple.py:32)."""
        torch = _torch()
        keep = []
        for l in range(self.n_layers):
            W = torch.as_tensor(np.asarray(Ws[l], dtype=np.float32).reshape(self.dims[l], self.dims[l + 1])
                                if not isinstance(Ws[l], torch.Tensor) else Ws[l], dtype=torch.float32)
            b = torch.as_tensor(np.asarray(bs[l], dtype=np.float32).reshape(self.dims[l + 1])
                                if not isinstance(bs[l], torch.Tensor) else bs[l], dtype=torch.float32)
            W = W.to(self.torch_device).contiguous()
            b = b.to(self.torch_device).contiguous()
            keep += [W, b]
            _ffi.check(self._lib, self._lib.b2_params_set(self._h, l, W.data_ptr(), b.data_ptr(), self._stream()))
        torch.cuda.synchronize(self.torch_device)

    def get_params(self):
        """Fetch the variables (simple.py:37 sess.run([W, b, ...])).  Returns numpy float32 arrays."""
        torch = _torch()
        Ws, bs = [], []
        for l in range(self.n_layers):
            W = torch.empty(self.dims[l], self.dims[l + 1], dtype=torch.float32, device=self.torch_device)
            b = torch.empty(self.dims[l + 1], dtype=torch.float32, device=self.torch_device)
            _ffi.check(self._lib, self._lib.b2_params_get(self._h, l, W.data_ptr(), b.data_ptr(), self._stream()))
            Ws.append(W)
            bs.append(b)
        torch.cuda.synchronize(self.torch_device)
        return [w.cpu().numpy() for w in Ws], [b.cpu().numpy() for b in bs]

    def init_random_normal(self, seed, op_seeds, stddev=1.0):
        """Initialise the variables the way distributed.py:83-91 does: every W[l] = tf.random_normal(shape, stddev) drawn
        on the device with TensorFlow 1.x's generator for op attributes (seed, op_seeds[l]) -- seed is the graph seed
        of tf.set_random_seed (2 in the reference), op_seeds[l] the op counter TensorFlow assigned to that
        tf.random_normal (version-dependent, hence explicit) -- and every bias zero (distributed.py:90-91)."""
        torch = _torch()
        assert len(op_seeds) == self.n_layers
        Ws, bs = [], []
        for l in range(self.n_layers):
            W = torch.empty(self.dims[l], self.dims[l + 1], dtype=torch.float32, device=self.torch_device)
            _ffi.check(self._lib, self._lib.b2_random_normal(W.data_ptr(), W.numel(), int(seed), int(op_seeds[l]), 0.0,
                                                             float(stddev), self._stream()))
            b = torch.zeros(self.dims[l + 1], dtype=torch.float32, device=self.torch_device)
            _ffi.check(self._lib, self._lib.b2_params_set(self._h, l, W.data_ptr(), b.data_ptr(), self._stream()))
            Ws.append(W)
            bs.append(b)
        torch.cuda.synchronize(self.torch_device)

    # -- save / restore (the reference builds no Saver: Supervisor(logdir=None), distributed.py:129-131) ------------
    def save(self, path):
        """Write the variables and global_step to a .npz (keys W0, b0, W1, ... and global_step)."""
        Ws, bs = self.get_params()
        arrays = {"global_step": np.float32(self.global_step), "dims": np.asarray(self.dims, dtype=np.int64)}
        for l in range(self.n_layers):
            arrays["W%d" % l] = Ws[l]
            arrays["b%d" % l] = bs[l]
        with open(path, "wb") as f:
            np.savez(f, **arrays)

    def restore(self, path):
        """Load variables written by save().  global_step restarts at 0 (it lives on the device and counts the steps of
        this context); the saved value is returned so that a caller can offset its own step display."""
        with np.load(path) as d:
            if list(d["dims"]) != list(self.dims):
                raise ValueError("checkpoint %s holds dims %s, this model is %s" % (path, list(d["dims"]), list(self.dims)))
            self.set_params([d["W%d" % l] for l in range(self.n_layers)], [d["b%d" % l] for l in range(self.n_layers)])
            return np.float32(d["global_step"])

    # -- data parallel ------------------------------------------------------------------------
    @staticmethod
    def new_unique_id():
        lib = _ffi.load_library()
        hint = _ffi.nccl_library_hint()
        if hint and not os.environ.get("B2_NCCL_LIB"):
            os.environ["B2_NCCL_LIB"] = hint
        buf = ctypes.create_string_buffer(_ffi.UNIQUE_ID_BYTES)
        _ffi.check(lib, lib.b2_get_unique_id(buf))
        return buf.raw

    def init_comm(self, rank, nranks, unique_id):
        """rank = worker task_index, nranks = number of workers (distributed.py:37-43)."""
        hint = _ffi.nccl_library_hint()
        if hint and not os.environ.get("B2_NCCL_LIB"):
            os.environ["B2_NCCL_LIB"] = hint
        buf = ctypes.create_string_buffer(bytes(unique_id), _ffi.UNIQUE_ID_BYTES) if unique_id else None
        _ffi.check(self._lib, self._lib.b2_comm_init(self._h, int(rank), int(nranks), buf))
        self.rank, self.nranks = int(rank), int(nranks)

    def peer_export(self):
        """64-byte CUDA IPC handle of this rank's gradient-exchange buffer (see include/b2mlp.h)."""
        buf = ctypes.create_string_buffer(_ffi.PEER_HANDLE_BYTES)
        _ffi.check(self._lib, self._lib.b2_peer_export(self._h, buf))
        return buf.raw

    def peer_import(self, handles):
        """handles: the exported handles of all ranks, in rank order (list of bytes or one bytes object)."""
        blob = b"".join(handles) if isinstance(handles, (list, tuple)) else bytes(handles)
        assert len(blob) == self.nranks * _ffi.PEER_HANDLE_BYTES
        _ffi.check(self._lib, self._lib.b2_peer_import(self._h, ctypes.create_string_buffer(blob, len(blob))))

    def set_async_period(self, k):
        """k > 0: emulate the reference's asynchronous parameter-server training (local updates, replicas averaged
        every k steps; distributed.py:107-113).  0: synchronous all-reduce (default)."""
        self.set_option(_ffi.OPT_ASYNC_PERIOD, int(k))

    def sharded_apply(self, on=True):
        """Several ranks, NCCL exchange: every rank owns one slice of the parameter vector (reduce-scatter of the gradients,
        SGD apply on the own slice, all-gather of the parameters) -- what k parameter servers are in the reference
        (distributed.py:63-65) -- instead of an all-reduce and a full apply on every rank."""
        self.set_option(_ffi.OPT_SHARDED_APPLY, 1 if on else 0)

    def use_peer_exchange(self, on=True):
        self.set_option(_ffi.OPT_NO_PEER_EXCHANGE, 0 if on else 1)

    def tensor_core_fp32(self, on=True):
        """fp32 models: run the wide layers' GEMMs on the tensor cores as fp32-accurate three-plane bf16 splits
        (default for batches >= 1024), or keep them on the CUDA cores."""
        self.set_option(_ffi.OPT_NO_SPLIT3, 0 if on else 1)

    def fused_output_layer(self, on=True):
        """fp32 models with >= 2 layers: output layer forward + loss head + backward as one kernel (default) or as the
        separate layer kernels."""
        self.set_option(_ffi.OPT_NO_NARROW_HEAD, 0 if on else 1)

    def persistent_steps(self, on=True):
        """train_steps() runs all its steps inside ONE kernel launch where a persistent kernel exists for the model
        (tiny models such as simple.py's; small batches of the 784-wide models) -- the default -- or enqueues every
        step's kernels separately."""
        self.set_option(_ffi.OPT_NO_PERSISTENT, 0 if on else 1)

    def multistep_fused(self, on=True):
        """784 -> 10 only: all steps of train_steps() inside one launch of the fused step kernel (in-kernel grid barriers
        around the reduction / apply).  Same bits as the default path, measured slower: an option, off by default."""
        self.set_option(_ffi.OPT_MULTISTEP_FUSED, 1 if on else 0)

    def overlap_exchange(self, on=True):
        """bf16 path, several ranks: all-reduce each wide layer's weight gradient on a second stream while the backward
        pass of the layers below it runs (default), or send everything in one all-reduce after the backward pass."""
        self.set_option(_ffi.OPT_NO_BUCKET_OVERLAP, 0 if on else 1)

    # -- the hot path ---------------------------------------------------------------------------
    def train_steps(self, x_pool, y_pool, local_batch, n_steps, first_row=0, lr=0.001,
                    loss_trace=None, correct_trace=None):
        """n_steps training steps over device-resident pools (torch CUDA tensors, float32, row-major).

        Asynchronous: returns (loss_trace, correct_trace) device tensors that are valid once the
        current stream has drained.  loss_trace[s] is the PRE-update cost of step s.
        """
        torch = _torch()
        assert x_pool.is_cuda and y_pool.is_cuda and x_pool.dtype == torch.float32 and y_pool.dtype == torch.float32
        assert x_pool.is_contiguous() and y_pool.is_contiguous()
        pool_rows = x_pool.shape[0]
        if loss_trace is None:
            loss_trace = torch.empty(max(n_steps, 1), dtype=torch.float32, device=self.torch_device)
        if correct_trace is None:
            correct_trace = torch.empty(max(n_steps, 1), dtype=torch.int32, device=self.torch_device)
        _ffi.check(self._lib, self._lib.b2_train_steps(
            self._h, x_pool.data_ptr(), y_pool.data_ptr(), pool_rows, int(local_batch),
            int(local_batch) * self.nranks, int(n_steps), int(first_row), float(lr),
            loss_trace.data_ptr(), correct_trace.data_ptr(), self._stream()))
        return loss_trace, correct_trace

    def run_train(self, x_host, y_host, lr):
        """sess.run([train_op, cross_entropy, accuracy-count], feed_dict={x: x_host, y_: y_host}).

        x_host / y_host: host float32 arrays (numpy, or pinned torch CPU tensors).  Synchronous.
        Returns (pre-update cost as np.float32, number of correct rows).
        """
        # the third element keeps a converted copy (python list, float64, non-contiguous view) alive until the C call,
        # which reads the host buffers synchronously, has returned
        xp, rows_x, keep_x = _host_ptr(x_host, self.dims[0])
        yp, rows_y, keep_y = _host_ptr(y_host, self.dims[-1])
        assert rows_x == rows_y, "x and y_ feeds disagree on the number of rows"
        loss = ctypes.c_float()
        correct = ctypes.c_int()
        _ffi.check(self._lib, self._lib.b2_session_run_train(self._h, xp, yp, rows_x, float(lr),
                                                             ctypes.byref(loss), ctypes.byref(correct)))
        del keep_x, keep_y
        return np.float32(loss.value), int(correct.value)

    def evaluate(self, x_dev, y_dev):
        """sess.run([cost, accuracy]) without a train op (distributed.py:164; simple.py:37)."""
        torch = _torch()
        assert x_dev.is_cuda and y_dev.is_cuda
        rows = x_dev.shape[0] if x_dev.dim() > 1 else x_dev.numel() // self.dims[0]
        loss = torch.empty(1, dtype=torch.float32, device=self.torch_device)
        correct = torch.empty(1, dtype=torch.int64, device=self.torch_device)
        _ffi.check(self._lib, self._lib.b2_eval(self._h, x_dev.data_ptr(), y_dev.data_ptr(), int(rows),
                                                loss.data_ptr(), correct.data_ptr(), self._stream()))
        torch.cuda.synchronize(self.torch_device)
        return np.float32(loss.item()), int(correct.item())


def _host_ptr(a, width):
    """(pointer, rows, keepalive) of a host float32 buffer given as numpy array, torch CPU tensor or nested list.
    `keepalive` owns the memory the pointer refers to; the caller holds it for as long as the pointer is in use."""
    torch = _torch()
    if isinstance(a, torch.Tensor):
        assert not a.is_cuda, "feeds are host buffers"
        if a.dtype != torch.float32 or not a.is_contiguous():
            a = a.to(torch.float32).contiguous()
        return ctypes.c_void_p(a.data_ptr()), a.numel() // width, a
    a = np.ascontiguousarray(a, dtype=np.float32)      # no copy when it already is contiguous float32
    return ctypes.c_void_p(a.ctypes.data), a.size // width, a
