"""The reference's call sequence written in plain C (examples/c_abi_demo.c), built with gcc against
include/miraculix_b200.h and the in-tree library, run on the GPU: a C / Fortran / Julia caller needs
nothing but the header's six reference symbols."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu


def test_c_demo_runs_on_the_gpu(gpu_lib):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    out = os.path.join(root, "miraculix_b200", "build", "c_abi_demo_gpu")
    libdir = os.path.join(root, "miraculix_b200")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    r = subprocess.run([gcc, "-std=c99", "-O2", "-I", os.path.join(root, "include"),
                        os.path.join(root, "examples", "c_abi_demo.c"), "-o", out, "-L", libdir,
                        "-lmiraculix_b200", f"-Wl,-rpath,{libdir}", "-lm"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for env_extra in ({}, {"MB200_SMALL_LIMIT": "0"}, {"CUDA_DEVICE": "0,0", "MB200_SMALL_LIMIT": "0"}):
        env = dict(os.environ, **env_extra)   # latency path, tcgen05 path, two shards behind the same calls
        run = subprocess.run([out], capture_output=True, text=True, timeout=300, env=env)
        assert run.returncode == 0 and run.stdout.strip().endswith("ok"), (env_extra, run.stdout, run.stderr)
