"""Host-side pieces of bench.py that can be checked without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_nccl_log_capture_keeps_stdout_to_one_json_line():
    """bench.py at N > 1 leaves NCCL's INFO log on: whatever the C runtime writes to stdout (NCCL's default sink) must
    land in the capture file and be relayed to stderr, and only the JSON line may reach the real stdout."""
    code = r'''
import sys, ctypes
sys.path.insert(0, %r)
import bench
n = bench.NcclLog()
libc = ctypes.CDLL(None)
libc.printf(b"box:7:7 [0] NCCL INFO NCCL version 2.28.9+cuda12.9\n")
libc.printf(b"box:7:7 [0] NCCL INFO comm 0xabc rank 0 nranks 2 cudaDev 0 nvmlDev 0 busId 1000 - Init COMPLETE\n")
print("stray python print")
r = n.relay(2)
n.emit(__import__("json").dumps({"nccl": r}))
''' % ROOT
    env = dict(os.environ)
    env.pop("NCCL_DEBUG_FILE", None)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, p.stdout
    got = json.loads(lines[0])["nccl"]
    assert got["nranks"] == 2 and got["comm_nranks_ok"] is True and got["lines"] == 2
    assert "nranks 2" in p.stderr and "stray python print" not in p.stdout


def test_both_arms_describe_the_same_config():
    sys.path.insert(0, ROOT)
    import bench
    a = bench.bench_config(1, 65536)
    b = bench.bench_config(1, 65536)
    assert a == b and "workload" in a
