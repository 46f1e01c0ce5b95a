"""Builds and runs the C++ host layer's own test program (include/ashcuda_host.hpp over the C ABI)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "ash_renderer_b200")


def _build(tmp_path):
    exe = str(tmp_path / "test_host")
    subprocess.run(["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "test_host.cpp"), "-o", exe, "-L", LIBDIR, "-lashcuda",
                    f"-Wl,-rpath,{LIBDIR}"], check=True)
    return exe


def test_cpp_host_layer(tmp_path):
    import torch
    args = ["--gpu"] if torch.cuda.is_available() else []
    out = subprocess.run([_build(tmp_path), *args], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "host tests ok" in out.stdout


@pytest.mark.gpu
def test_cpp_feature_runs_a_view_on_the_gpu(tmp_path):
    out = subprocess.run([_build(tmp_path), "--gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "host tests ok (gpu:" in out.stdout
