"""Host-side helpers of bench.py that the GPU box cannot be used to debug: the parser of the out-of-process clock
sampler and the rule that stdout carries nothing but the JSON line."""
import io
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


class _FakeProc:
    def __init__(self, text):
        self.stdout = io.StringIO(text)


def test_clock_rows_are_parsed_and_windowed():
    s = object.__new__(bench.ClockSampler)
    s.rows, s.h, s.nv, s.wall, s.alive = [], None, None, True, True
    now = time.time()

    def stamp(t):
        lt = time.localtime(t)
        return time.strftime("%Y/%m/%d %H:%M:%S", lt) + ".%03d" % int((t % 1) * 1000)
    lines = [f"{stamp(now - 5.0)}, 1200, 1965, 300.5, Not Active, Not Active, Not Active, Not Active",
             f"{stamp(now - 1.0)}, 1965, 1965, 700.25, Not Active, Not Active, Not Active, Active",
             f"{stamp(now - 0.9)}, 1950, 1965, 710.00, Not Active, Not Active, Not Active, Not Active",
             "garbage line",
             f"{stamp(now + 5.0)}, 900, 1965, 100.0, Active, Not Active, Not Active, Not Active"]
    s.proc = _FakeProc("\n".join(lines) + "\n")
    s.t0, s.t1 = now - 2.0, now
    s._drain()
    assert len(s.rows) == 4
    out = s._summary()
    assert out["samples"] == 2 and out["sm_min_mhz"] == 1950.0 and out["sm_max_mhz"] == 1965.0
    assert out["reasons"] == ["sw_power_cap"] and abs(out["power_w_max"] - 710.0) < 1e-6
    assert "nvidia-smi" in out["source"]


def test_stdout_carries_only_the_json_line():
    code = ("import os, sys; sys.path.insert(0, %r); import bench; bench.claim_stdout(); "
            "os.write(1, b'NCCL version banner\\n'); print('library chatter'); bench.emit({'metric': 'x', 'value': 1})") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert json.loads(r.stdout) == {"metric": "x", "value": 1}
    assert "NCCL version banner" in r.stderr and "library chatter" in r.stderr
