"""The driver's contract for bench.py, checked where it can be without a GPU: the reference (CPU) arm prints ONE JSON line with
the keys the driver reads, on all host threads even when torch.distributed.run has exported OMP_NUM_THREADS=1, and with the same
`config.workload` string the GPU arm prints; ranks other than 0 exit 0 without output; the GPU arm refuses to run without a device
(no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=e, timeout=600)


def test_reference_arm_line():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-seconds", "0.5"], {"OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    sys.path.insert(0, ROOT)
    import bench
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == "samples/s" and d["higher_is_better"] is True
    assert d["config"]["workload"] == bench.workload_name(bench.SAMPLES_PER_STEP)
    assert d["value"] > 1e4 and d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] and cb["cores"] == len(os.sched_getaffinity(0)) and ("spp" in cb["sample"] or "tiles" in cb["sample"])
    assert d["gpu_launches"] == 0 and d["steps"] == 1


def test_reference_arm_other_ranks_are_silent():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--cpu-seconds", "0.5"], {"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_gpu_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    r = _run(["--steps", "1", "--warmup", "0"])
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


def test_roofline_traffic_is_refused_when_the_capture_is_from_other_sources(monkeypatch):
    """roofline.traffic comes from a committed ncu capture (ncu cannot run inside the timed region); the capture records the
    SHA-256 of the NIF kernel's sources and bench.py must answer (None, why) instead of a stale constant when they differ."""
    sys.path.insert(0, ROOT)
    import bench
    value, source = bench.nif_traffic_bytes()
    if value is not None:                      # a matching capture is committed: it must be the newest round's and sane
        assert source.startswith("profiles/r") and 1e8 < value < 1e9
    monkeypatch.setattr(bench, "nif_source_hashes", lambda: {"nif_tc.cu": "edited"})
    value, why = bench.nif_traffic_bytes()
    assert value is None and "re-capture" in why


def test_multi_gpu_check_is_wired_into_the_gpu_arm():
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert 'extras["multi_gpu_check"]' in src and "def multi_gpu_check(" in src and '"bit_equal"' in src and '"counters_equal"' in src
    assert '"OMP_NUM_THREADS"' in src            # the C++ host run must not inherit torchrun's single-thread setting
