"""bench.py pieces that run without a GPU: the reference arm's 'unavailable' line, the clock sampler's window
filtering, and the argument contract the driver relies on."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_prints_unavailable_json_and_exits_zero():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1",
                        "--steps", "2", "--warmup", "3"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and d["unavailable"]


def test_clock_sampler_reports_only_samples_inside_the_timed_windows():
    import bench
    s = bench.ClockSampler(0)
    s._nvml = object()                       # pretend NVML is up: exercise the aggregation only
    s._t = type("T", (), {"join": lambda self, timeout=None: None})()
    s.max_mhz = 1965.0
    t0 = time.perf_counter()
    s.samples = [(t0 + 0.0, 1000.0, 0), (t0 + 1.0, 1965.0, 0x4), (t0 + 2.0, 1950.0, 0), (t0 + 3.0, 500.0, 0x8)]
    s.windows = [(t0 + 0.5, t0 + 2.5)]
    out = s.stop()
    assert out["samples_inside_timed_region"] == 2 and out["sm_mhz"] == 1957.5
    assert out["reasons"] == ["sw_power_cap"]          # hw_slowdown happened outside the window
    assert out["sm_max_mhz"] == 1965.0 and out["source"] == "nvml"


def test_clock_sampler_without_nvml_or_nvidia_smi_degrades():
    import bench
    s = bench.ClockSampler(0)
    s.start(); s.begin(); s.end()
    out = s.stop()
    assert "reasons" in out and "sm_mhz" in out


def test_bench_defaults_match_the_driver_contract():
    import bench
    old = sys.argv
    sys.argv = ["bench.py"]
    try:
        a = bench.parse_args()
    finally:
        sys.argv = old
    assert a.gpus == 1 and a.warmup >= 3 and a.steps > 0 and a.impl == "ours" and a.l2 == "inputs"
