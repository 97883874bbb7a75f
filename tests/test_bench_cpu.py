"""CPU tests of the measurement contract: `bench.py --impl reference` (the CPU arm: oracle port on the host cores) prints
one JSON line with the contract's keys; the roofline byte models are the DESIGN.md §4 formulas; the default arm refuses
to run without a GPU instead of falling back."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=240):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def test_reference_arm_prints_the_contract_line():
    r = _run("--impl", "reference", "--workload", "ladybug49", "--steps", "2", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "ba_lm_iterations_per_s" and d["unit"] == "LM iterations/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["config"]["workload"] == "ladybug49" and d["config"]["N"] == 31843 and d["data"] == "synthetic"
    assert d["value"] > 0 and abs(d["value"] - 1e3 / d["ms_per_step"]) < 1e-6 * d["value"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["vs_baseline"] is None


def test_reference_arm_under_torchrun_env_only_rank0_prints():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29977")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--workload", "ladybug49",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]   # ranks > 0 exit 0 without work


def test_default_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    r = _run("--workload", "ladybug49", "--steps", "1", "--warmup", "0", "--no-pose", "--no-cpu-baseline", timeout=120)
    assert r.returncode != 0
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]


def test_bytes_models_are_the_design_formulas():
    sys.path.insert(0, ROOT)
    import bench
    C, P, N = 1778, 993923, 5001946
    m = bench.bytes_models(C, P, N, 4, tiled=False, fused=False)
    assert m["resjac"] == 96 * N + 32 * P + 88 * C          # N·(16+4+4+8·jsz+8+32) + 32·P + 88·C
    assert m["mv_pt"] == 32 * N + 52 * P                    # chunk kernel: N·(3·jsz + 4 + 8 + 8) + 52·P
    assert m["mv_cam"] == 48 * N + (2 * 576 + 1024 + 160) * C
    assert m["cost"] == 20 * N + 32 * P + 88 * C
    assert m["eval"] == 32 * N + 32 * P + 96 * C
    mt = bench.bytes_models(C, P, N, 4, tiled=True, fused=True)
    assert mt["mv_pt"] == 34 * N + 52 * P                   # tiled kernel: N·(3·jsz + 2 + 4 + 8 + 8) + 52·P
    assert mt["resjac"] == 88 * N + 32 * P + (88 + 576) * C  # fused: no r_B array, U_c / g_c written by the same launch
    m8 = bench.bytes_models(C, P, N, 8, tiled=False, fused=False)
    assert m8["mv_cam"] - m["mv_cam"] == 32 * N and m8["mv_pt"] - m["mv_pt"] == 12 * N


def test_both_arms_share_one_config_dict():
    """The driver compares `config` of the two arms byte for byte."""
    sys.path.insert(0, ROOT)
    import argparse

    import bench

    class S:
        C, P, N = 3, 8, 20
    assert bench.workload_config(argparse.Namespace(workload="tiny"), S) == {"workload": "tiny", "C": 3, "P": 8, "N": 20}
