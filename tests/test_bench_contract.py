"""bench.py contract checks that need no GPU: the reference arm (CPU oracle port on the host cores) prints
ONE JSON line with the keys the driver reads; the product arm refuses to run without a CUDA device."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _run(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, env=e,
                          timeout=600, cwd=ROOT)


def test_reference_arm_prints_the_contract_line():
    r = _run("--impl", "reference", "--n-spheres", "20000", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s" and d["unit"] == "Mrays/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert "workload" in d["config"]


def test_reference_arm_uses_all_cores_and_never_loads_the_product():
    """torchrun exports OMP_NUM_THREADS=1 to its workers: the arm must still use every host core (the
    N>1 reference ratios were inflated ~20x by a 1-thread arm), and it must not load librtb200.so (RTB200_LIB points
    at a file that does not exist: any call into the product would raise)."""
    r = _run("--impl", "reference", "--n-spheres", "20000", "--steps", "1", "--warmup", "0",
             env={"OMP_NUM_THREADS": "1", "RTB200_LIB": "/nonexistent/librtb200.so"})
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == os.cpu_count() or d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert list(d["config"].keys()) == ["workload"]  # nothing the product arm's config does not have


def test_terrain_ramp_restatement_matches_the_reference_table():
    """bench.py's numpy ramp (reference arm: no product, no oracle/_ref) against ColorRamp.cuh's own bake (golden)."""
    import numpy as np
    sys.path.insert(0, ROOT)
    import bench
    gold = np.load(os.path.join(HERE, "golden", "ref_color_ramp.npz"))
    for n in (1, 2, 7, 500):
        assert np.allclose(bench.terrain_ramp_numpy(n), gold[f"terrain_{n}"], rtol=0, atol=2e-7), n


def test_reference_arm_other_ranks_exit_quietly():
    r = _run("--impl", "reference", "--gpus", "2", "--n-spheres", "20000", "--steps", "1", "--warmup", "0",
             env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and not [l for l in r.stdout.splitlines() if l.startswith("{")]


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a CUDA device is present")
    r = _run("--n-spheres", "1000", "--steps", "1", "--warmup", "0")
    assert r.returncode != 0
    assert "no CUDA device" in (r.stderr + r.stdout)


def test_nested_step_functions_are_called_with_their_own_keywords():
    """run_ours and run_c5 each define frame()/timed(); a call must only use keywords its own definition accepts
    (a search-and-replace across both once broke the C5 arm, which no CPU test executes)."""
    import ast
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    for fn in [n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name in ("run_ours", "run_c5")]:
        defs = {n.name: [a.arg for a in n.args.args] for n in ast.walk(fn) if isinstance(n, ast.FunctionDef)}
        for c in ast.walk(fn):
            if isinstance(c, ast.Call) and isinstance(c.func, ast.Name) and c.func.id in defs:
                for k in c.keywords:
                    assert k.arg in defs[c.func.id], (fn.name, c.func.id, k.arg, c.lineno)
