"""Multi-GPU parameter server (needs >= 2 GPUs; skipped on a 1-GPU box): launched like the driver launches bench.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(nproc, args, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "cluster_check.py")] + args
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


needs2 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")


@needs2
@pytest.mark.parametrize("layout,model,pull", [("colocated", "cnn", "fused"), ("colocated", "cnn", "prefetch"),
                                               ("dedicated", "cnn", "auto"), ("colocated", "mlp", "fused"),
                                               ("colocated", "mlp", "prefetch")])
def test_two_gpu_sync_ps(layout, model, pull):
    out = _torchrun(2, ["--layout", layout, "--model", model, "--pull", pull], 29611)
    assert "OK" in out and "FAIL" not in out


@needs2
def test_two_gpu_async_apply_on_arrival_counts_every_update():
    out = _torchrun(2, ["--layout", "dedicated", "--async_mode"], 29613)
    assert "OK" in out and "FAIL" not in out and "async" in out
