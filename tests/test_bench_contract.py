"""CPU: the parts of bench.py's contract that need no GPU -- the flags the driver passes exist, and nothing
but the ONE JSON line can reach stdout (library chatter on file descriptor 1 is sent to stderr)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_accepts_the_drivers_flags():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in r.stdout
    assert "reference" in r.stdout


def test_only_the_json_line_reaches_stdout():
    code = ("import os, sys; sys.path.insert(0, %r); import bench; bench._only_the_line_on_stdout(); "
            "os.write(1, b'NCCL version x\\n'); os.system('echo chatter'); print('{\"metric\": 1}')") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0
    assert r.stdout == '{"metric": 1}\n'
    assert "NCCL version x" in r.stderr and "chatter" in r.stderr
