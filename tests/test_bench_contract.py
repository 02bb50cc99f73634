"""bench.py's one JSON line: the keys the driver reads, identical workload ``config`` in every arm, clock samples
attributed to the timed regions. Runs without a GPU (the measuring code itself is covered by tests/test_gpu_multi.py)."""
import argparse
import datetime
import io
import json
import os
import sys
import time
from contextlib import redirect_stdout

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

REQUIRED = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "impl"}


def _line(impl, extra):
    args = argparse.Namespace(steps=20, warmup=5)
    buf = io.StringIO()
    with redirect_stdout(buf):
        bench.finish(args, impl, 0, 1, 2.5, 2.6, 11000, 8, 1700, {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "reasons": []}, extra)
    lines = [l for l in buf.getvalue().splitlines() if l.startswith("{")]
    assert len(lines) == 1
    return json.loads(lines[0])


def test_one_json_line_with_the_contract_keys():
    d = _line("ours", {"transport": "local"})
    assert REQUIRED <= set(d)
    assert d["value"] == round(50 / 2.5e-3, 2) and d["higher_is_better"] is True and d["scaling"] == "weak"
    assert d["e2e"]["value"] == round(50 / 2.6e-3, 2)
    assert d["e2e"]["h2d_bytes_per_step"] == 11000 and d["e2e"]["d2h_bytes_per_step"] == 8
    assert d["steps"] == 20 and d["warmup"] == 5 and d["gpu_launches"] == 1700
    assert "l2_policy" in d["config"] and d["details"]["transport"] == "local"


def test_config_is_the_same_in_every_arm():
    ours, ref, lib = _line("ours", {"transport": "nvlink"}), _line("reference", {"note": "x"}), _line("torch", {})
    assert ours["config"] == ref["config"] == lib["config"]
    assert ref["dtype"] == "fp32" and ours["dtype"] == "bf16"
    assert "no clip" in ref["details"]["privacy"] and "clip C=2" in ours["details"]["privacy"]


def test_clock_samples_are_attributed_to_the_timed_regions():
    s = bench.ClockSampler(0)

    class _Proc:
        def terminate(self): pass
        def wait(self, t): pass
    s.proc = _Proc()
    now = time.time()

    def ts(t):
        return datetime.datetime.fromtimestamp(t).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
    idle = "Not Active, Not Active, Not Active, Not Active"
    with open(s.path, "w") as f:
        f.write(f"{ts(now - 1.0)}, 1965, 1965, 200.1, 0x0, {idle}\n")
        f.write(f"{ts(now + 0.05)}, 1800, 1965, 900.1, 0x4, Not Active, Not Active, Not Active, Active\n")
        f.write(f"{ts(now + 0.06)}, 1700, 1965, 900.1, 0x4, Not Active, Not Active, Not Active, Active\n")
        f.write(f"{ts(now + 5.0)}, 1965, 1965, 200.1, 0x0, {idle}\n")
    s.windows.append((now, now + 0.1))
    c = s.stop()
    assert c["samples"] == 2 and c["samples_total"] == 4 and c["sm_mhz"] in (1700.0, 1800.0)
    assert c["reasons"] == ["sw_power_cap"] and c["sm_max_mhz"] == 1965.0


def test_reference_arm_answers_without_a_gpu():
    import subprocess
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(bench.__file__), "bench.py"), "--impl", "reference",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert d["impl"] == "reference" and ("unavailable" in d or "value" in d)
