"""CPU checks of bench.py's contract: the reference arm's JSON line, its rank / thread policy under torchrun's
environment, and that the B200 arm refuses to run without a device (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(args, env_extra=None, timeout=300):
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "TORCHELASTIC_RUN_ID", "OMP_NUM_THREADS"):
        env.pop(k, None)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=timeout)


def test_reference_arm_prints_the_contract_line():
    p = run(["--impl", "reference", "--steps", "2", "--warmup", "1", "--no-extra"])
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Msamples/s" and d["unit"] == "Msamples/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"]
    assert d["cpu_baseline"]["cores"] >= 1 and "1920x1080" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "configs[1]" in d["config"]["workload"]
    # the same workload config the B200 arm prints: real size, the workload's OWN spp per step, seed, jitter
    import bench
    assert d["config"] == bench.workload_config("c2", bench.WORKLOADS["c2"], 1, "p2p")
    assert d["config"]["width"] == 1920 and d["config"]["height"] == 1080 and d["config"]["spp_per_step_total"] == 64
    assert "64 of 64 spp per step" in d["cpu_baseline"]["sample"] and d["cpu_baseline"]["resolve_ms"] > 0


def test_reference_arm_weak_scaled_config_follows_the_gpu_count():
    import bench
    c1 = bench.workload_config("c2", bench.WORKLOADS["c2"], 1, "p2p")
    c8 = bench.workload_config("c2", bench.WORKLOADS["c2"], 8, "p2p")
    assert c1["spp_per_step_total"] == 64 and c8["spp_per_step_total"] == 512 and c8["n_gpus"] == 8
    assert bench.workload_config("c5", bench.WORKLOADS["c5"], 8, "p2p")["spp_per_step_total"] == 1024      # strong


def test_reference_arm_under_torchrun_environment():
    """Only rank 0 works and prints; it asks OpenMP for every core although torchrun exports OMP_NUM_THREADS=1."""
    tr = {"WORLD_SIZE": "2", "TORCHELASTIC_RUN_ID": "t", "OMP_NUM_THREADS": "1", "MASTER_ADDR": "127.0.0.1"}
    p = run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--no-extra"], dict(tr, RANK="1", LOCAL_RANK="1"))
    assert p.returncode == 0 and p.stdout.strip() == ""
    p = run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--no-extra"], dict(tr, RANK="0", LOCAL_RANK="0"))
    assert p.returncode == 0, p.stderr
    d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][0])
    assert d["n_gpus"] == 2 and d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert d["config"]["spp_per_step_total"] == 128 and "128 of 128 spp" in d["cpu_baseline"]["sample"]


def test_b200_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a device is present")
    p = run(["--steps", "1", "--warmup", "0"])
    assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)
