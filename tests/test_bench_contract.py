"""bench.py's line contract, on the CPU: the reference arm (`--impl reference`: the reference's own ThreadRender on the host
cores, oracle/_ref/qaref_perf) prints exactly ONE line on stdout, a JSON object with the keys the driver reads, and the same
`config` keys as the B200 arm; the B200 arm refuses to run without a CUDA device (there is no CPU path to fall back to)."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ASSETS, ROOT


def run_bench(args, timeout=600):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                          timeout=timeout, cwd=ROOT)


def test_reference_arm_prints_one_json_line():
    from oracle import refapi
    if not os.path.exists(refapi.REF_PERF):
        pytest.skip("oracle/_ref/qaref_perf missing (run `make -C oracle ref` where /root/reference exists)")
    if not os.path.exists(os.path.join(ASSETS, "STANDIN_ASSETS.txt")):
        pytest.skip("stand-in assets missing")
    r = run_bench(["--impl", "reference", "--scene", "example_project3_sphere", "--spp", "4", "--steps", "2", "--warmup", "1", "--gpus", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "stdout must carry the JSON line only: %r" % r.stdout[:500]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "spp" in cb["sample"]
    cfg = d["config"]
    assert cfg["scene"] == "example_project3_sphere" and cfg["spp_per_gpu"] == 4 and cfg["max_bounce"] == 5 and "workload" in cfg
    assert "model" not in cfg


def test_reference_arm_ranks_other_than_zero_do_nothing():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=120, cwd=ROOT, env=dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"))
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_b200_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = run_bench(["--steps", "1", "--warmup", "0", "--no-cpu-baseline"], timeout=300)
    assert r.returncode != 0 and r.stdout.strip() == ""
    assert "no CUDA device" in r.stderr
