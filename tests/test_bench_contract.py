"""bench.py's JSON contract, as far as it can be checked without a GPU: the reference arm (CPU) prints ONE
line with the keys the driver reads, ranks other than 0 print nothing, and the GPU arm refuses to run
without a CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")


def _run(args, env_extra=None, timeout=240):
    env = dict(os.environ)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, BENCH] + args, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_reference_arm_prints_one_contract_line():
    r = _run(["--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "1", "--cpu-seconds", "2"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "8x8 env steps/sec (lockstep batch)" and d["unit"] == "env_steps/s"
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1 and d["value"] > 0
    assert d["config"]["workload"].startswith("8x8 lockstep batch")
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    from oracle import make_ref
    assert cb["kind"] == ("reference" if make_ref.available() else "port")
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


@pytest.mark.parametrize("config,unit", [("sizes", "env_steps/s"), ("ppo_feed", "env_steps/s"),
                                         ("tournament", "games/s")])
def test_reference_arm_of_the_other_configs(config, unit):
    from oracle import make_ref
    if config == "tournament" and not make_ref.available():
        pytest.skip("the reference's policies are needed (oracle/_ref not built)")
    r = _run(["--impl", "reference", "--config", config, "--steps", "1", "--warmup", "0", "--cpu-seconds", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == unit and d["value"] > 0 and d["e2e"]["value"] == d["value"]
    assert d["cpu_baseline"]["cores"] >= 1 and "configs[" in d["config"]["workload"]


def test_reference_arm_is_silent_on_other_ranks():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"], {"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_gpu_arm_fails_loudly_without_a_device():
    r = _run(["--steps", "1", "--warmup", "1"], timeout=120)
    assert r.returncode != 0 and r.stdout.strip() == ""
    assert "CUDA" in (r.stderr + r.stdout)
