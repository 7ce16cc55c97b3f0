"""The C ABI from a plain-C caller (examples/c_api/demo.c): the header is valid C99, the demo links against
the shared library, fails loudly without a GPU and, on a B200, scores its synthetic pair."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    import __graft_entry__ as ge
    ge.build()
    exe = str(tmp_path / "demo")
    lib = os.path.join(ROOT, "cgnn_b200")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "c_api", "demo.c"), "-o", exe, "-L", lib, "-lcgnn_b200", "-lm",
                        "-Wl,-rpath," + lib], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_demo_compiles_and_fails_loudly_without_gpu(tmp_path):
    import torch
    exe = _build(tmp_path)
    if not torch.cuda.is_available():
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode == 2 and "cgnn_b200_create failed" in r.stderr


@pytest.mark.gpu
def test_c_demo_runs_on_gpu(tmp_path):
    r = subprocess.run([_build(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "causal score" in r.stdout and "kernel" in r.stdout
