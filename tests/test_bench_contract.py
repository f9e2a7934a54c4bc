"""The reference arm of bench.py runs on the CPU: check the JSON line the driver parses (no GPU needed)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ref_line(extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "3",
                        "--batch", "256"] + extra, capture_output=True, text=True, timeout=600, env=e)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.strip().splitlines()


def test_reference_arm_line_has_the_contract_keys():
    lines = _ref_line([])
    assert len(lines) == 1                      # exactly ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "training samples/sec" and d["unit"] == "samples/s"
    for key in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
                "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["value"] > 0 and d["steps"] == 2
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_print_nothing():
    # under torchrun only rank 0 runs the CPU arm; the other ranks exit 0 without work
    assert _ref_line(["--gpus", "2"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_both_arms_build_the_same_config_object():
    """The driver compares the two arms' `config` objects (`same_config`): both come from bench.make_config, and every key
    is a function of (workload, batch, world, exchange flags) only -- nothing an arm could fill in differently."""
    sys.path.insert(0, ROOT)
    import bench
    d = json.loads(_ref_line([])[0])
    assert d["config"] == bench.make_config("c3", 256, 1)
    for world in (1, 2, 8):
        cfg = bench.make_config("c3", 65536, world)
        assert cfg["global_batch"] == 65536 * world and cfg["parallelism"] == "dp%d" % world
        assert cfg["pool_rows"] % cfg["global_batch"] == 0 and cfg["pool_rows"] >= 2 * cfg["global_batch"]
        assert ("NVLink" in cfg["gradient_exchange"]) == (world > 1)
    assert "nccl" in bench.make_config("c5", 65536, 8)["gradient_exchange"]
    assert "nccl" in bench.make_config("c3", 65536, 8, no_peer=True)["gradient_exchange"]
