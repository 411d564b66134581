"""CPU tests of bench.py's contract pieces that need no GPU: the reference arm (`--impl reference`: the unmodified reference, oracle/_ref, on
the host cores; one JSON line on stdout, rank 0 only under torchrun) and the loud failure of the GPU arm without a device."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env_extra=None):
    env = dict(os.environ)
    env.pop("RANK", None); env.pop("WORLD_SIZE", None); env.pop("LOCAL_RANK", None)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)


def test_reference_arm_prints_one_contract_line():
    from oracle import refpy
    if not refpy.available("sse2"):
        pytest.skip("oracle/_ref not built")
    r = _run(["--impl", "reference", "--steps", "2", "--warmup", "3", "--ref-sources", "128"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "sources" and d["vs_baseline"] is None
    assert d["steps"] == 2 and d["warmup"] == 3 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_exit_quietly():
    r = _run(["--impl", "reference", "--steps", "2", "--warmup", "3", "--ref-sources", "128"], {"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_gpu_arm_fails_loudly_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--steps", "1", "--warmup", "3", "--no-e2e", "--no-cpu-baseline"])
    assert r.returncode != 0
    assert "no CUDA device" in (r.stderr + r.stdout)
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]      # no metric line from a CPU fallback


def test_reference_arm_never_maps_the_product_library():
    """the CPU arm must time the reference alone: nothing of jplspatial_b200 (libjplg.so) may be loaded into its process"""
    from oracle import refpy
    if not refpy.available("sse2"):
        pytest.skip("oracle/_ref not built")
    code = ("import sys, os; sys.path.insert(0, %r); import bench; r = bench.best_reference(64, 1); assert r and r['best']['ms_per_block_all_threads'] > 0; "
            "maps = open('/proc/self/maps').read(); assert 'libjplg' not in maps, 'libjplg.so mapped'; assert 'jplspatial_b200' not in sys.modules; "
            "assert 'libjplref' in maps; print('clean', r['best']['variant'])") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "clean" in r.stdout
