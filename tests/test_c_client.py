"""GPU: a plain C99 program (no CUDA headers, no Python) drives the library through include/gradavg.h
with host buffers and checks the result bit for bit against its own rank-order fold."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_plain_c_client(built, tmp_path):
    libdir = os.path.join(ROOT, "pytorch-distributed-example_b200", "csrc")
    exe = str(tmp_path / "host_client")
    subprocess.check_call(["gcc", "-std=c99", "-O1", "-Wall", "-Werror", "-ffp-contract=off",
                           "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "c_client", "host_client.c"),
                           "-L", libdir, "-lgradavg", "-Wl,-rpath," + libdir, "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120,
                         env=dict(os.environ, GA_TIMEOUT_MS="8000"))
    assert out.returncode == 0, (out.stdout + out.stderr)[-1500:]
    assert "C client OK" in out.stdout
