"""CPU-only: the reference arm of bench.py (`--impl reference`) -- the reference's own shader source compiled as C++ (oracle/_ref) or
the oracle port, timed on the host cores -- prints ONE JSON line with the contract's keys on the same metric / unit / config as
the GPU arm, and under a multi-rank launch only rank 0 works."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(extra_env, *args):
    env = {**os.environ, **extra_env}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--cpu-seconds", "1", *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    return [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_prints_one_contract_line():
    lines = run_bench({})
    assert len(lines) == 1
    d = lines[0]
    assert d["impl"] == "reference" and d["metric"] == "cluster-light tests/s" and d["unit"] == "tests/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["config"]["workload"] == "C16M" and d["config"]["lights"] == 1 << 24 and d["config"]["grid"] == [64, 36, 96]
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_runs_on_rank_0_only():
    env = {"WORLD_SIZE": "2", "LOCAL_RANK": "1", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29577"}
    assert run_bench({**env, "RANK": "1"}, "--gpus", "2") == []          # the other ranks leave without work and without output
    lines = run_bench({**env, "RANK": "0", "LOCAL_RANK": "0"}, "--gpus", "2")
    assert len(lines) == 1 and lines[0]["impl"] == "reference" and lines[0]["n_gpus"] == 2
