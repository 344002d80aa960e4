"""CPU: bench.py's reference arm (the reference's CPU path, timed on the host cores) prints the contract's
JSON line.  The native arm needs a GPU and must refuse to run without one (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tiny",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True,
                         timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "diloco_outer_sync_ms" and d["unit"] == "ms"
    assert d["higher_is_better"] is False and d["n_gpus"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # measured on the whole tensor list, never scaled from a sample
    assert d["cpu_baseline"]["whole_workload"] is True and d["cpu_baseline"]["extrapolated"] is False
    assert d["cpu_baseline"]["sample_elems"] == 9_766_528 and d["partial_workload"] is False


def test_reference_arm_on_a_capped_sample_says_so_and_does_not_scale():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tiny",
                          "--steps", "2", "--warmup", "1", "--cpu-max-elems", "5000000"], capture_output=True,
                         text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][0])
    cb = d["cpu_baseline"]
    assert cb["whole_workload"] is False and cb["extrapolated"] is False and cb["sample_elems"] < 5_000_001
    assert d["partial_workload"] is True and d["config"]["workload"].startswith("SAMPLE of tiny")
    assert d["value"] == cb["value"]                  # the sample's own time


def test_native_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode != 0 and "no CUDA device" in (out.stderr + out.stdout)


def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_cpu_sample_is_a_bounded_prefix_of_the_tensor_list():
    b = _bench_module()
    shapes = b.workload_shapes("llama1b")
    assert len(shapes) == 201 and sum(b.numel(s) for s in shapes) == 1_261_529_088
    sample, n = b.cpu_sample_shapes(shapes, 256_000_000)
    assert sample == shapes[:len(sample)] and n <= 256_000_000 and n > 100_000_000
    one, n1 = b.cpu_sample_shapes(shapes, 1)            # never empty: at least the first tensor
    assert len(one) == 1 and n1 == 32000 * 2048


def test_workload_config_names_the_workload():
    import argparse
    b = _bench_module()
    cfg = b.workload_config(argparse.Namespace(workload="llama1b", replicas=2, dtype="bf16"), 1_261_529_088, 201, 8)
    assert "llama1b" in cfg["workload"] and "4 worker(s), 2 GPU(s) per worker" in cfg["workload"]
    assert cfg["parallelism"] == "diloco-dp4x2" and "model" not in cfg


def test_clock_sampler_degrades_without_nvml():
    b = _bench_module()
    s = b.ClockSampler(0)
    s.start(); s.stop(); s.close()
    out = s.summary()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons"}
