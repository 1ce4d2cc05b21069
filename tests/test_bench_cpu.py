"""bench.py contract that can be checked without a GPU: the reference arm (`--impl reference`, the oracle port on the host
cores) prints one JSON line with the contract's keys, and every rank but 0 exits 0 without work."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = dict(os.environ); env.update(env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env, capture_output=True,
                          text=True, timeout=600)


def test_reference_arm_line():
    # a launcher's OMP_NUM_THREADS=1 (torchrun) must not turn the baseline into a one-core run
    r = _run({"OMP_NUM_THREADS": "1"}, "--impl", "reference", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["unit"] == "ms" and j["higher_is_better"] is False
    assert j["metric"].startswith("ms/MD-step") and j["value"] > 0 and j["ms_per_step"] == j["value"]
    assert j["steps"] == 1 and j["warmup"] == 0 and j["dtype"] == "f64" and j["data"] == "synthetic"
    assert "au4096" in j["config"]["workload"]
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == j["value"] and cb["unit"] == "ms" and cb["sample"]
    assert cb["cores"] == os.cpu_count() == j["cores"]
    assert j["sampled"] is True and j["extrapolated"] is True and cb["scale_factors"]["density"] == 4096 / 256
    assert all(cb["scale_factors"][k] == 1.0 for k in ("neighbours", "radial", "assembly", "forces"))
    assert cb["variant_reference_dense_dgemm"]["density_ms"] > 0
    assert j["e2e"] == {"value": j["value"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_do_nothing():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0 and r.stdout.strip() == ""
