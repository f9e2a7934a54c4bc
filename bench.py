#!/usr/bin/env python
"""Headline benchmark: training samples/sec of the hot path (BASELINE.json `metric`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2] [--batch B]
  python bench.py --impl reference ...      # the reference's CPU implementation of the same step

One JSON line on stdout (rank 0).  A "step" is one pass of the training step (forward, loss,
accuracy, backward, [all-reduce], SGD apply) over one batch of synthetic MNIST-shaped rows.
  value   whole-job samples/s with the data pool resident in HBM (CUDA events, max over ranks)
  e2e     the same through the sess.run-shaped C-ABI call with HOST feeds (H2D + D2H inside the timing)
  roofline / cpu_baseline / clocks / gpu_launches   see DESIGN.md "Measurement"
Multi-GPU: launched by torchrun, one rank per GPU; weak scaling (per-GPU batch fixed).
"""
import argparse
import datetime
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (dims, act, loss, scaled_init, default local batch, description, precision)
    "c3": ([784, 10], 0, 1, True, 65536,
           "784->10 dense + softmax-CE (simple.py-style single layer on MNIST-shaped rows), fp32", "fp32"),
    "c2": ([784, 100, 10], 1, 1, False, 65536,
           "784-100-10 sigmoid MLP + softmax-CE (distributed.py model), fp32", "fp32"),
    "c5": ([784, 4096, 4096, 10], 1, 1, True, 65536,
           "784-4096-4096-10 sigmoid MLP + softmax-CE, bf16 operands / fp32 accumulate (tcgen05 GEMMs)", "bf16"),
}
CPU_SAMPLE_BATCH = {"c5": 2048}        # rows per CPU step where a full batch would take minutes on the host
BYTES_PER_SAMPLE = 784 * 4 + 10 * 4   # x row + one-hot y row, each read once
LR = 0.001
HOST_POOL_ROWS = 1 << 17               # generated on the host with the fixed seed, tiled on the device


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic_bytes(workload):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(workload)
        except Exception:
            return None
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for nme, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


_CPU_KIND = {"port": "torch-CPU restatement of the TF1 graph (TensorFlow is not importable on this box)",
             "reference": "the reference's own graph on the TensorFlow found on this box (tf.compat.v1 session, feed_dict)"}


def cpu_reference_run(dims, act, loss, scaled, batch, steps, warmup, budget_s=None, lr=LR):
    """The reference's CPU implementation of the step on all host threads: the verbatim TF graph when TensorFlow is
    importable on this box (kind "reference"), else its restatement oracle/torch_cpu.py (kind "port")."""
    import torch
    from oracle import tf_live
    from oracle.torch_cpu import TorchCPUGraph
    from tensorflow_examples_b200 import synthetic
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rows = min(HOST_POOL_ROWS, max(batch, 2 * batch))
    rows = max(rows, batch)
    x, y = synthetic.make_pool(rows)
    Ws, bs = synthetic.make_weights(dims, scaled)
    kind = "port"
    g = None
    if tf_live.available():
        try:
            live = tf_live.LiveGraph(dims, act, loss, Ws, bs, lr)

            class _Live:                      # same call shape as TorchCPUGraph.train_step
                def train_step(self, xb, yb, _lr):
                    return live.train_step(xb.numpy(), yb.numpy())
            g, kind = _Live(), "reference"
        except Exception:
            g = None
    if g is None:
        g = TorchCPUGraph(dims, act, loss, Ws, bs, threads=threads)
    xt, yt = torch.from_numpy(x), torch.from_numpy(y)
    nb = rows // batch
    losses = []
    for s in range(warmup):
        sl = slice((s % nb) * batch, (s % nb) * batch + batch)
        losses.append(g.train_step(xt[sl], yt[sl], lr)[0])
    t0 = time.perf_counter()
    done = 0
    for s in range(warmup, warmup + steps):
        sl = slice((s % nb) * batch, (s % nb) * batch + batch)
        losses.append(g.train_step(xt[sl], yt[sl], lr)[0])
        done += 1
        if budget_s and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return {"value": batch * done / dt, "steps": done, "seconds": dt, "threads": threads, "kind": kind,
            "first_loss": losses[0], "ms_per_step": 1e3 * dt / done}


# ---- batch-size sweep / strong scaling: `bench.py --sweep` --------------------
SWEEP_POINTS = {
    # name: (dims, act, loss, scaled init, precision, flop/sample, learning rate)
    "c3": ([784, 10], 0, 1, True, "fp32", 31360, LR),
    # simple.py's SSE-SUM head on the same layer; its gradient grows with the batch, so the step is 1e-9
    "c3-sse": ([784, 10], 0, 0, True, "fp32", 31360, 1e-9),
    "c2": ([784, 100, 10], 1, 1, False, "fp32", 319600, LR),
}
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.92e9 / 1e12               # 148 SMs x 128 FMA lanes x 2 flop x 1.92 GHz (spec-derived)


def cpu_numpy_run(dims, act, loss, scaled, batch, budget_s, lr=LR):
    from oracle.tf1_graph import TF1Graph
    from tensorflow_examples_b200 import synthetic
    x, y = synthetic.make_pool(2 * batch)
    Ws, bs = synthetic.make_weights(dims, scaled)
    g = TF1Graph(dims, act, loss, Ws, bs)
    g.train_step(x[:batch], y[:batch], lr)
    t0, done = time.perf_counter(), 0
    while done < 50 and time.perf_counter() - t0 < budget_s:
        o = (done % 2) * batch
        g.train_step(x[o:o + batch], y[o:o + batch], lr)
        done += 1
    return batch * done / (time.perf_counter() - t0), done


def sweep_main(args):
    """For every (workload, batch) point: `steps` training steps over a device-resident pool, CUDA events on the launch
    stream after `warmup` untimed steps, max over ranks; samples/s, the fraction of the HBM ceiling (3176 B/sample over
    the measured copy bandwidth) and of the spec-derived CUDA-core fp32 ceiling.  --sweep-cpu times the torch-CPU and
    numpy restatements beside it (rank 0).  Points below B ~ 16384 are launch-latency bound."""
    import torch
    import tensorflow_examples_b200 as pkg
    from tensorflow_examples_b200 import synthetic
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=600))   # a lost rank ends the run, not hangs it
    peak_gbs, _ = _peaks()
    batches = [int(b) for b in args.sweep_batches.split(",")]
    hx, hy = synthetic.make_pool(HOST_POOL_ROWS)
    pool_rows = max(2 * max(batches) * (1 if args.strong else world), 1 << 19)
    reps = (pool_rows + HOST_POOL_ROWS - 1) // HOST_POOL_ROWS
    x_pool = torch.from_numpy(hx).to(dev).repeat(reps, 1)[:pool_rows].contiguous()
    y_pool = torch.from_numpy(hy).to(dev).repeat(reps, 1)[:pool_rows].contiguous()
    rows_out = []
    for name in args.sweep_points.split(","):
        dims, act, loss, scaled, prec, flops, lr = SWEEP_POINTS[name]
        for B in batches:
            local = B // world if args.strong else B
            if local < 1 or (args.strong and B % world):
                continue
            gb = local * world
            tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=local, device=local_rank, precision=prec)
            tr.set_params(*synthetic.make_weights(dims, scaled))
            if world > 1:
                uid = [pkg.DenseTrainer.new_unique_id() if rank == 0 else None]
                dist.broadcast_object_list(uid, src=0)
                tr.init_comm(rank, world, uid[0])
                hs = [None] * world
                dist.all_gather_object(hs, tr.peer_export())
                tr.peer_import(hs)
            lt, _ = tr.train_steps(x_pool, y_pool, local, args.warmup, first_row=0, lr=lr)
            torch.cuda.synchronize(dev)
            if dist is not None:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            lt, _ = tr.train_steps(x_pool, y_pool, local, args.steps, first_row=(args.warmup * gb) % pool_rows, lr=lr)
            e1.record()
            torch.cuda.synchronize(dev)
            ms = e0.elapsed_time(e1)
            if dist is not None:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            sps = gb * args.steps / (ms / 1e3)
            row = {"point": name, "dims": dims, "head": "sse-sum" if loss == 0 else "softmax-ce-mean", "n_gpus": world,
                   "scaling": "strong" if args.strong else "weak", "global_batch": gb, "local_batch": local,
                   "us_per_step": 1e3 * ms / args.steps, "samples_per_s": sps,
                   "hbm_frac": sps / world * BYTES_PER_SAMPLE / (peak_gbs * 1e9),
                   "fp32_frac": sps / world * flops / (FP32_PEAK_TFLOPS * 1e12), "last_loss": float(lt[args.steps - 1].item())}
            if args.sweep_cpu and rank == 0:
                r = cpu_reference_run(dims, act, loss, scaled, gb, 50, 2, budget_s=4.0, lr=lr)
                row["cpu_torch_samples_per_s"], row["cpu_threads"] = r["value"], r["threads"]
                row["cpu_numpy_samples_per_s"], _ = cpu_numpy_run(dims, act, loss, scaled, gb, 4.0, lr)
            tr.close()
            if rank == 0:
                rows_out.append(row)
                print(json.dumps(row), flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(args.sweep_out)), exist_ok=True)
        json.dump({"hbm_peak_gbs": peak_gbs, "fp32_peak_tflops_spec": FP32_PEAK_TFLOPS, "steps": args.steps,
                   "warmup": args.warmup, "cpu_count": os.cpu_count(), "rows": rows_out}, open(args.sweep_out, "w"), indent=1)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def make_config(workload, batch, world, no_peer=False, no_overlap=False):
    """The `config` object of the JSON line -- built the same way by both arms, so the driver sees one config."""
    dims, act, loss, scaled, dbatch, desc, precision = WORKLOADS[workload]
    global_batch = batch * world
    pool_rows = max(2 * global_batch, 1 << 19)
    pool_rows = ((pool_rows + global_batch - 1) // global_batch) * global_batch
    exchange_floats = sum(((dims[i] * dims[i + 1] + 3) & ~3) + ((dims[i + 1] + 3) & ~3) for i in range(len(dims) - 1)) + 4
    if world == 1:
        gx = "none (one GPU)"
    elif precision == "bf16":
        gx = ("nccl all-reduce after the backward pass + apply" if no_overlap else
              "nccl all-reduce per layer bucket on a second stream, overlapped with the backward pass, + apply")
    elif no_peer or exchange_floats > (1 << 18):
        gx = "nccl all-reduce + apply"
    else:
        gx = "pushed over NVLink peer memory inside the reduction step ({value, step} cells), fused with the SGD apply"
    return {"workload": "%s: %s" % (workload, desc), "dims": dims, "local_batch": batch, "global_batch": global_batch,
            "lr": LR, "parallelism": "dp%d" % world, "pool_rows": pool_rows,
            "l2": "inputs larger than L2: %.2f GB pool, consecutive batches" % (pool_rows * BYTES_PER_SAMPLE / 1e9),
            "gradient_exchange": gx}


class Dist:
    """rank / world plumbing shared by every workload of one bench.py process."""

    def __init__(self):
        import torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            # a rank that died leaves the others in a collective: bound that wait so the run ends with an error
            dist.init_process_group("nccl", device_id=self.dev, timeout=datetime.timedelta(seconds=600))
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def max_over_ranks(self, v):
        if self.dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def make_trainer(D, workload, batch, args):
    """Model on this rank's GPU with the shared initial weights; communicator + peer mapping when there are several ranks."""
    import tensorflow_examples_b200 as pkg
    from tensorflow_examples_b200 import synthetic
    dims, act, loss, scaled, _, _, precision = WORKLOADS[workload]
    tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=batch, device=D.local_rank, precision=precision)
    tr.set_params(*synthetic.make_weights(dims, scaled))
    if args.no_persistent:
        tr.persistent_steps(False)
    if args.multistep_fused:
        tr.multistep_fused(True)
    if D.world > 1:
        uid = [pkg.DenseTrainer.new_unique_id() if D.rank == 0 else None]
        D.dist.broadcast_object_list(uid, src=0)
        tr.init_comm(D.rank, D.world, uid[0])
        if not args.no_peer and tr.exchange_floats <= (1 << 18):
            # map every rank's gradient cells over NVLink: the exchange + SGD apply happen inside the step's own kernel
            hs = [None] * D.world
            D.dist.all_gather_object(hs, tr.peer_export())
            tr.peer_import(hs)
        if precision == "bf16":
            tr.overlap_exchange(not args.no_overlap)
        if args.sharded_apply:
            tr.sharded_apply(True)
    return tr


def parity_check(D, tr, workload, batch, x_host, y_host, loss0, correct0):
    """Outside every timed region: step 0 of this run against the CPU oracle at the GLOBAL batch (rank 0), and -- with
    several ranks -- a bit-level comparison of the replicas' parameters after all the steps run so far."""
    import torch
    dims, act, loss, scaled, _, _, precision = WORKLOADS[workload]
    out = {}
    gb = batch * D.world
    if D.rank == 0 and precision == "bf16" and gb > 2 * HOST_POOL_ROWS:
        out["oracle"] = "skipped: the 4096-wide forward of %d rows takes minutes in numpy" % gb
    elif D.rank == 0:
        from oracle.tf1_graph import TF1Graph
        from tensorflow_examples_b200 import synthetic
        Ws, bs = synthetic.make_weights(dims, scaled)
        g = TF1Graph(dims, act, loss, Ws, bs, precision="bf16" if precision == "bf16" else "fp32")
        idx = np.arange(gb) % len(x_host)                      # the device pool is the host pool tiled
        lo, co = g.evaluate(x_host[idx], y_host[idx])          # forward + loss head: the pre-update loss of step 0
        out["oracle"] = "numpy restatement of the TF1 graph, one process, global batch %d" % gb
        out["loss_rel"] = abs(loss0 - float(lo)) / max(1.0, abs(float(lo)))
        out["correct_diff"] = int(correct0) - int(co)
    Ws, bs = tr.get_params()
    flat = np.concatenate([w.ravel() for w in Ws] + [b.ravel() for b in bs]).view(np.uint32).astype(np.uint64)
    chk = int((flat * (np.arange(flat.size, dtype=np.uint64) % np.uint64(65521) + np.uint64(1))).sum() % np.uint64(1 << 62))
    if D.dist is not None:
        t = torch.tensor([chk], dtype=torch.int64, device=D.dev)
        all_t = [torch.zeros_like(t) for _ in range(D.world)]
        D.dist.all_gather(all_t, t)
        out["replicas_identical"] = bool(all(int(v.item()) == chk for v in all_t))
    else:
        out["replicas_identical"] = True
    out["param_checksum"] = chk
    return out


def run_b200(D, args, workload, batch, steps, warmup, hx, hy, with_e2e, with_cpu, with_parity=True):
    """One workload on this process's GPU (all ranks call it together).  Returns the JSON line's fields (rank 0) or None."""
    import torch
    dims, act, loss, scaled, _, desc, precision = WORKLOADS[workload]
    rank, world, dev = D.rank, D.world, D.dev
    global_batch = batch * world
    cfg = make_config(workload, batch, world, args.no_peer, args.no_overlap)
    pool_rows = cfg["pool_rows"]
    # data: fixed-seed host pool, tiled on the device to >= 2 global batches and >= 1.6 GB (>> 126 MB L2)
    reps = (pool_rows + HOST_POOL_ROWS - 1) // HOST_POOL_ROWS
    x_pool = torch.from_numpy(hx).to(dev).repeat(reps, 1)[:pool_rows].contiguous()
    y_pool = torch.from_numpy(hy).to(dev).repeat(reps, 1)[:pool_rows].contiguous()
    tr = make_trainer(D, workload, batch, args)

    # ---- value: device-resident pool -----------------------------------------------------
    # the clock sampler is started BEFORE the warm-up steps, so that nothing idles between warm-up and the timed region
    sampler = ClockSampler(D.local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    D.barrier()
    lt, ct = tr.train_steps(x_pool, y_pool, batch, warmup, first_row=0, lr=LR)
    torch.cuda.synchronize(dev)
    first = torch.stack([lt[0], ct[0].float()]).cpu()               # one small read; the values are used after the timing
    # W warm-up steps of a 60 us step are 0.3 ms: not enough for the GPU to reach its steady clocks.  Keep warming up (untimed,
    # same steps) until >= 20 ms of steps have run; the line reports the number of warm-up steps actually run.  The step
    # time is estimated from a second, already warm round of W steps (the first one pays module loading).
    t_w = time.perf_counter()
    tr.train_steps(x_pool, y_pool, batch, warmup, first_row=(warmup * global_batch) % pool_rows, lr=LR)
    torch.cuda.synchronize(dev)
    round_s = time.perf_counter() - t_w
    extra = warmup
    more = 0
    if round_s < 0.020:
        more = int(min(4000, (0.020 - round_s) / max(round_s / warmup, 1e-6)))
    if D.dist is not None:
        # the same count on every rank, and EVERY rank takes part whatever its own clock said: a round that takes 20 ms on
        # one rank and 19.9 ms on another must not leave one of them alone in the collective
        tt = torch.tensor([more], dtype=torch.int64, device=dev)
        D.dist.all_reduce(tt, op=D.dist.ReduceOp.MAX)
        more = int(tt.item())
    if more > 0:
        tr.train_steps(x_pool, y_pool, batch, more, first_row=(2 * warmup * global_batch) % pool_rows, lr=LR)
        torch.cuda.synchronize(dev)
    extra += more
    warm_total = warmup + extra
    D.barrier()
    torch.cuda.synchronize(dev)
    l0 = tr.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    lt, ct = tr.train_steps(x_pool, y_pool, batch, steps, first_row=(warm_total * global_batch) % pool_rows, lr=LR)
    ev1.record()
    torch.cuda.synchronize(dev)
    D.barrier()
    ms = D.max_over_ranks(ev0.elapsed_time(ev1))
    launches = tr.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    last_loss = float(lt[steps - 1].item())
    first_loss, first_correct = float(first[0].item()), int(first[1].item())
    value = global_batch * steps / (ms / 1e3)

    # ---- roofline: per-launch device times over the same steps (events on the launch stream) ----
    tr.kernel_timing(True)
    tsteps = 200 if precision == "fp32" else min(max(steps, 3), 20)   # enough launches for a stable mean (13 ms of fp32 steps)
    tr.train_steps(x_pool, y_pool, batch, tsteps, first_row=0, lr=LR)
    kt = tr.kernel_times()
    tr.kernel_timing(False)
    total_ms = sum(v[0] for v in kt.values())
    if precision == "bf16":
        # tensor-core bound: flops of the three GEMM classes (2*in*out per layer and product)
        hid = [(dims[i], dims[i + 1]) for i in range(len(dims) - 2)]
        fl = {"gemm_fwd_tcgen05": sum(2.0 * a * b for a, b in hid) * batch,
              "gemm_wgrad_tcgen05": sum(2.0 * a * b for a, b in hid) * batch,
              "gemm_dgrad_tcgen05": sum(2.0 * a * b for a, b in hid[1:]) * batch}
        gemm_ms = sum(kt[k][0] for k in fl if k in kt) / tsteps
        gemm_flops = sum(fl.values())
        peak_tf, src = 1406.5, "fallback"
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak_tf, src = float(pk["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained, cuBLAS loop)"
        except Exception:
            pass
        achieved = gemm_flops / (gemm_ms / 1e3) / 1e12
        roofline = {"bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf,
                    "traffic": _traffic_bytes(workload), "kernel": "gemm_bf16_kernel (tcgen05; fwd+dgrad+wgrad launches)",
                    "kernel_ms": gemm_ms, "kernel_share_of_step": gemm_ms * tsteps / total_ms, "peak_source": "of " + src,
                    "algorithmic_flops_per_step": gemm_flops,
                    "per_class_tflops": {k: fl[k] / (kt[k][0] / tsteps / 1e3) / 1e12 for k in fl if k in kt},
                    "kernel_ms_by_class": {k: v[0] / tsteps for k, v in kt.items()}}
    else:
        dom = max(kt.items(), key=lambda kv: kv[1][0])
        # steps one launch of the dominant kernel covers: a persistent kernel runs the whole train_steps call
        steps_per_launch = tsteps // dom[1][1] if dom[0] == "persistent_steps" else 1
        dom_ms = dom[1][0] / dom[1][1]
        peak, peak_src = _peaks()
        bytes_per_launch = steps_per_launch * batch * BYTES_PER_SAMPLE
        achieved = bytes_per_launch / (dom_ms / 1e3) / 1e9
        traffic = _traffic_bytes(workload)
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic * steps_per_launch if (traffic and steps_per_launch > 1) else traffic,
                    "kernel": dom[0], "kernel_ms": dom_ms, "steps_per_launch": steps_per_launch,
                    "kernel_share_of_step": dom[1][0] / total_ms, "peak_source": "of " + peak_src,
                    "algorithmic_bytes_per_launch": bytes_per_launch,
                    "kernel_ms_by_class": {k: v[0] / v[1] for k, v in kt.items()}}
        if workload == "c2":
            # 784-100-10 in fp32 is not HBM-bound: its work is the two 65536x784x100 products, run as six
            # bf16 tensor-core products each; report the fp32-equivalent rate next to the HBM fraction
            roofline["fp32_equivalent_tflops"] = value / world * 319600 / 1e12
            roofline["cuda_core_fp32_peak_tflops_spec"] = FP32_PEAK_TFLOPS
            roofline["note"] = ("compute-bound model: wide layer = 6 bf16 tcgen05 products per fp32 product (three-plane "
                                "split, fp32-accurate); the HBM fraction is reported for completeness, it is not the bound")

    # ---- parity (outside the timed regions): oracle at the global batch, replicas bit-identical ------------------
    parity = parity_check(D, tr, workload, batch, hx, hy, first_loss, first_correct) if with_parity else None

    # ---- e2e: host feeds through the sess.run-shaped call ------------------------------------
    e2e = None
    if with_e2e:
        in_b, out_b = batch * dims[0] * 4, batch * dims[-1] * 4
        feeds = []
        for i in range(2):
            fx = torch.empty(batch, dims[0], dtype=torch.float32).pin_memory()
            fy = torch.empty(batch, dims[-1], dtype=torch.float32).pin_memory()
            idx = (np.arange(batch) + i * batch) % HOST_POOL_ROWS
            fx.copy_(torch.from_numpy(hx[idx]))
            fy.copy_(torch.from_numpy(hy[idx]))
            feeds.append((fx, fy))
        esteps = max(5, min(steps, 50))
        for i in range(3):
            tr.run_train(feeds[i % 2][0], feeds[i % 2][1], LR)
        D.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for i in range(esteps):
            tr.run_train(feeds[i % 2][0], feeds[i % 2][1], LR)
        torch.cuda.synchronize(dev)
        dt = D.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": global_batch * esteps / dt, "unit": "samples/s", "h2d_bytes_per_step": in_b + out_b,
               "d2h_bytes_per_step": 8, "steps": esteps, "ms_per_step": 1e3 * dt / esteps,
               "api": "DenseTrainer.run_train -> b2_session_run_train (pinned host feeds, synchronous)"}

    # ---- cpu_baseline: rank 0, N=1 only, bounded sample -----------------------------------------
    cpu = None
    if with_cpu and rank == 0 and world == 1:
        cbatch = min(batch, CPU_SAMPLE_BATCH.get(workload, batch))
        r = cpu_reference_run(dims, act, loss, scaled, cbatch, 1200, 2, budget_s=15.0)   # ~10 s of CPU work at 8-9 ms per 65536-row step
        cpu = {"value": r["value"], "unit": "samples/s", "cores": r["threads"], "kind": r["kind"],
               "sample": "%d steps of %d rows in %.1f s, %s, all host threads"
                         % (r["steps"], cbatch, r["seconds"], _CPU_KIND[r["kind"]]),
               "first_loss_cpu": r["first_loss"], "first_loss_gpu": first_loss}
    tr.close()
    del x_pool, y_pool
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    line = {"metric": "training samples/sec", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": steps, "warmup": warm_total, "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if precision == "bf16" else "f32",
            "data": "synthetic", "config": cfg, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
            "gpu_launches": launches, "loss_first": first_loss, "loss_last": last_loss}
    if parity is not None:
        line["parity"] = parity
    return line


def embedded_sweep(D, args, hx, hy, steps=100, warmup=10):
    """BASELINE.json configs[2] asks for a batch sweep 256-65536: five points for 784->10 and for distributed.py's
    784-100-10, device-resident pool, CUDA events around `steps` steps after `warmup`, one GPU.  Points below ~16384 rows
    are latency-bound (their HBM fraction is printed for completeness only)."""
    import torch
    import tensorflow_examples_b200 as pkg
    from tensorflow_examples_b200 import synthetic
    peak_gbs, _ = _peaks()
    pool_rows = 1 << 19
    reps = pool_rows // HOST_POOL_ROWS
    x_pool = torch.from_numpy(hx).to(D.dev).repeat(reps, 1).contiguous()
    y_pool = torch.from_numpy(hy).to(D.dev).repeat(reps, 1).contiguous()
    rows = []
    for name in ("c3", "c2"):
        dims, act, loss, scaled, prec, flops, lr = SWEEP_POINTS[name]
        for B in (256, 1024, 4096, 16384, 65536):
            tr = pkg.DenseTrainer(dims, act, loss, max_local_batch=B, device=D.local_rank, precision=prec)
            tr.set_params(*synthetic.make_weights(dims, scaled))
            if args.no_persistent:
                tr.persistent_steps(False)
            tr.train_steps(x_pool, y_pool, B, warmup, first_row=0, lr=lr)
            torch.cuda.synchronize(D.dev)
            t_w = time.perf_counter()                                # ... and on until 20 ms of steps have run (see run_b200)
            tr.train_steps(x_pool, y_pool, B, warmup, first_row=0, lr=lr)
            torch.cuda.synchronize(D.dev)
            round_s = time.perf_counter() - t_w
            if round_s < 0.020:
                tr.train_steps(x_pool, y_pool, B, int(min(4000, (0.020 - round_s) / max(round_s / warmup, 1e-6))) + 1, first_row=0, lr=lr)
                torch.cuda.synchronize(D.dev)
            l0 = tr.launch_count
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tr.train_steps(x_pool, y_pool, B, steps, first_row=(warmup * B) % pool_rows, lr=lr)
            e1.record()
            torch.cuda.synchronize(D.dev)
            ms = e0.elapsed_time(e1)
            sps = B * steps / (ms / 1e3)
            rows.append({"point": name, "dims": dims, "global_batch": B, "us_per_step": 1e3 * ms / steps,
                         "samples_per_s": sps, "hbm_frac": sps * BYTES_PER_SAMPLE / (peak_gbs * 1e9),
                         "launches_per_step": (tr.launch_count - l0) / steps})
            tr.close()
    del x_pool, y_pool
    torch.cuda.empty_cache()
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=None, help="rows per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the extra lines a default run embeds: `secondary` (scaled-up MLP, c5), `distributed_py_model` "
                         "(784-100-10, c2) and, on one GPU, the batch `sweep`")
    ap.add_argument("--no-peer", action="store_true", help="multi-GPU: use the NCCL all-reduce instead of the peer-memory kernel")
    ap.add_argument("--no-overlap", action="store_true",
                    help="multi-GPU bf16 path: one all-reduce after the backward pass instead of overlapped per-layer buckets")
    ap.add_argument("--no-persistent", action="store_true", help="one launch sequence per step instead of the multi-step kernels")
    ap.add_argument("--sharded-apply", action="store_true",
                    help="multi-GPU, NCCL exchange: reduce-scatter + apply on the own slice + all-gather instead of all-reduce + full apply")
    ap.add_argument("--multistep-fused", action="store_true", help="784->10: all steps in one launch of the fused kernel (optional path)")
    ap.add_argument("--sweep", action="store_true", help="batch-size sweep of c3 / c3 with the SSE head / 784-100-10 instead of the headline line")
    ap.add_argument("--sweep-out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--sweep-points", default="c3,c3-sse,c2")
    ap.add_argument("--sweep-batches", default="256,1024,4096,16384,65536")
    ap.add_argument("--sweep-cpu", action="store_true", help="sweep: also time the CPU restatements on rank 0")
    ap.add_argument("--strong", action="store_true", help="sweep: the batch is the GLOBAL batch, split over the ranks")
    args = ap.parse_args()
    if args.sweep:
        return sweep_main(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dims, act, loss, scaled, dbatch, desc, precision = WORKLOADS[args.workload]
    batch = args.batch or dbatch
    warmup = max(3, args.warmup)
    cfg = make_config(args.workload, batch, world, args.no_peer, args.no_overlap)

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        cbatch = min(batch, CPU_SAMPLE_BATCH.get(args.workload, batch))
        r = cpu_reference_run(dims, act, loss, scaled, cbatch, args.steps, min(warmup, 5), budget_s=90.0)
        line = {"impl": "reference", "metric": "training samples/sec", "value": r["value"], "unit": "samples/s",
                "n_gpus": args.gpus, "steps": r["steps"], "warmup": min(warmup, 5), "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": r["value"], "unit": "samples/s", "cores": r["threads"], "kind": r["kind"],
                                 "sample": "%d steps of %d rows each (one GPU's batch), %s" % (r["steps"], cbatch, _CPU_KIND[r["kind"]])},
                "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------ B200 arm
    import torch
    from tensorflow_examples_b200 import synthetic
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    D = Dist()
    hx, hy = synthetic.make_pool(HOST_POOL_ROWS)
    line = run_b200(D, args, args.workload, batch, args.steps, warmup, hx, hy, with_e2e=not args.no_e2e,
                    with_cpu=not args.no_cpu_baseline)
    if args.workload == "c3" and args.batch is None and not args.no_secondary:
        # BASELINE.json quotes the metric on two models; the headline line is the simple.py-style 784->10 model
        # (HBM-bound), the scaled-up 784-4096-4096-10 bf16 MLP (tensor-core-bound) rides along as `secondary`, and
        # distributed.py's own 784-100-10 in fp32 as `distributed_py_model` -- at every N, so the scaling run has them.
        keep = ("metric", "value", "unit", "n_gpus", "ms_per_step", "steps", "warmup", "dtype", "config", "roofline",
                "gpu_launches", "loss_first", "loss_last", "parity", "scaling")
        extra = {}
        for key, wl, st, wu in (("secondary", "c5", 20, 3), ("distributed_py_model", "c2", 100, 10)):
            try:
                d = run_b200(D, args, wl, WORKLOADS[wl][4], st, wu, hx, hy, with_e2e=False, with_cpu=False)
                if d is not None:
                    extra[key] = {k: d[k] for k in keep if k in d}
            except Exception as e:   # the headline line must not depend on the extra ones
                if world > 1:
                    raise            # ... but a rank that dropped out of a collective cannot be papered over
                extra[key] = {"error": repr(e)[:300]}
        if world == 1:
            try:
                extra["sweep"] = embedded_sweep(D, args, hx, hy)
            except Exception as e:
                extra["sweep"] = {"error": repr(e)[:300]}
        if line is not None:
            line.update(extra)
    if rank == 0:
        print(json.dumps(line))
    D.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
