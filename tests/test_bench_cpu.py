"""bench.py's CPU-side contract: the reference arm (oracle port on the host cores) prints one JSON
line with the keys the driver reads, and the device arm refuses to run without a GPU instead of
falling back to anything."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT,
                          capture_output=True, text=True, timeout=600, env=e)


def test_reference_arm_json_line():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "pairs/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    # default workload = the full refine loop the metric is quoted on (BASELINE.json configs[4])
    assert d["config"]["workload"].startswith("cfg5")
    assert "rasteriser" in d["cpu_baseline"]["sample"]


def test_reference_arm_cfg2():
    p = _run("--impl", "reference", "--config", "2", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][0])
    assert d["config"]["workload"].startswith("cfg2") and d["value"] > 0


def test_reference_arm_other_ranks_exit_quietly():
    p = _run("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
             env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert p.returncode == 0 and not [l for l in p.stdout.splitlines() if l.startswith("{")]


def test_device_arm_needs_a_gpu():
    import torch

    if torch.cuda.is_available():
        return
    p = _run("--steps", "1", "--warmup", "1", "--no-cpu-baseline")
    assert p.returncode != 0
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]
