"""Guard mode (PGM_GUARD=1, pgm_api.cu ensure_/guard_check): every device buffer exact-sized between 4 KiB guard zones
filled with 0xA5, checked after each download. compute-sanitizer is closed on this GPU pool
(profiles/sanitizer_r02_unavailable.md), so this is the out-of-bounds-write evidence of SURVEY.md section 5:
the parity tests themselves run again under the guards (a write past a buffer fails the call; data a kernel reads
without writing it first is 0xA5 garbage instead of a fresh allocation's zeros, so it fails parity)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, guard, timeout=1500):
    env = dict(os.environ, PGM_GUARD=guard)
    return subprocess.run([sys.executable] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)


def test_guard_sees_a_write_behind_a_buffer():
    p = _run(["profiles/micro/sanitize_run.py"], "2")
    assert p.returncode != 0
    assert "PGM_GUARD" in (p.stdout + p.stderr) and "byte 3 behind the buffer" in (p.stdout + p.stderr)


def test_every_kernel_family_under_guards():
    p = _run(["profiles/micro/sanitize_run.py"], "1")
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert "sanitize_run: all cases ran" in p.stdout


def test_parity_suites_under_guards():
    files = ["tests/test_gpu_parity.py", "tests/test_gpu_golden.py", "tests/test_gpu_text.py", "tests/test_gpu_svb.py", "tests/test_gpu_model.py",
             "tests/test_gpu_pipeline.py"]
    p = _run(["-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider"] + files, "1")
    assert p.returncode == 0, (p.stdout + p.stderr)[-4000:]
