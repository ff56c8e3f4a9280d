"""bench.py's contract as far as it can be checked without a GPU: the reference arm (the CPU port of the reference's path, the
one leg of bench.py that may execute oracle/) prints ONE JSON line with the keys the driver reads; the CUDA arm refuses to run
without a device instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, cwd=ROOT,
                          timeout=600)


def test_reference_arm_prints_the_contract_line_for_c1():
    r = _run("--impl", "reference", "--config", "c1", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "atmospheric TOD detector-samples/sec" and j["unit"] == "detector-samples/s"
    assert j["higher_is_better"] is True and j["vs_baseline"] is None and j["n_gpus"] == 1 and j["steps"] == 1
    assert j["value"] > 1e5 and j["ms_per_step"] > 0
    # C1 runs at full size on the CPU arm: 217 detectors x 30 000 samples per step
    assert abs(j["value"] * j["ms_per_step"] * 1e-3 - 217 * 30000) < 1.0
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == j["value"] and "217 of 217 detectors" in cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["gpu_launches"] == 0 and "217 detectors (217 on this rank" in j["config"]["workload"]


def test_cuda_arm_has_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        return  # on a GPU box the arm runs; its output is checked by the driver
    r = _run("--steps", "1", "--warmup", "3")
    assert r.returncode != 0 and "no CPU fallback" in r.stderr and not any(ln.startswith("{") for ln in r.stdout.splitlines())
