"""Host-side arithmetic of the samples variant (opengl_raytracer_b200/csrc/rt_spec.cuh), checked without a GPU:
nvcc compiles a small host program that includes the kernel header and compares the LCG jump tables with
RandomValue()'s state update (fragment.glsl:206-208) stepped one draw at a time, plus the size-class ladder."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="needs nvcc (host code only, no GPU)")
def test_jump_tables_and_size_classes(tmp_path):
    exe = tmp_path / "spec_host_check"
    src = os.path.join(ROOT, "tests", "native", "spec_host_check.cu")
    build = subprocess.run(["nvcc", "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false",
                            "-o", str(exe), src], capture_output=True, text=True, timeout=600)
    assert build.returncode == 0, build.stdout + build.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0 and run.stdout.startswith("ok"), run.stdout + run.stderr
