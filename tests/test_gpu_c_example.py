"""The plain-C example against the built library on a real GPU: five Adam steps through the C ABI alone."""
import os
import shutil
import subprocess

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "tensorflow_distributed_b200", "_native")


@pytest.mark.skipif(not torch.cuda.is_available() or shutil.which("gcc") is None
                    or not os.path.isdir("/usr/local/cuda/include"), reason="needs a GPU, gcc and the CUDA headers")
def test_c_example_runs_five_adam_steps(tmp_path):
    exe = str(tmp_path / "dense_step")
    subprocess.run(["gcc", "-Wall", "-std=c99", os.path.join(ROOT, "examples", "c_api_dense_step.c"),
                    "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", "-L" + LIBDIR, "-ltfd_b200",
                    "-L/usr/local/cuda/lib64", "-lcudart", "-lm", "-o", exe], check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=LIBDIR + ":/usr/local/cuda/lib64:" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "adam t=5 version=5 global_step=5" in r.stdout
