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
@pytest.mark.parametrize("precision", ["tf32", "bf16"])
@pytest.mark.parametrize("layout,model,pull", [("colocated", "cnn", "fused"), ("colocated", "cnn", "prefetch"),
                                               ("dedicated", "cnn", "auto"), ("colocated", "mlp", "fused"),
                                               ("colocated", "mlp", "prefetch")])
def test_two_gpu_sync_ps(layout, model, pull, precision):
    out = _torchrun(2, ["--layout", layout, "--model", model, "--pull", pull, "--precision", precision], 29611)
    assert "OK" in out and "FAIL" not in out and precision in out


@needs2
def test_two_gpu_fused_pull_is_bit_identical_to_copy_then_gemm():
    """the pull-fused GEMM (weight tiles TMA-loaded from the owner GPUs inside the kernel) against
    copy-then-unfused (the weights copied to a local mirror first, same GEMM on local memory).  No kernel of either
    model uses floating-point atomics, so after 12 synchronous steps parameters and Adam slots must agree bit for
    bit -- for the MLP and for the CNN."""
    import re
    for model in ("mlp", "cnn"):
        digests = []
        for pull in ("fused", "prefetch"):
            out = _torchrun(2, ["--layout", "colocated", "--model", model, "--pull", pull, "--digest"], 29617)
            assert "OK" in out and "FAIL" not in out
            digests.append(re.search(r"DIGEST ([0-9a-f]{64})", out).group(1))
        assert digests[0] == digests[1], model


@needs2
def test_two_gpu_async_apply_on_arrival_counts_every_update():
    out = _torchrun(2, ["--layout", "dedicated", "--async_mode"], 29613)
    assert "OK" in out and "FAIL" not in out and "async" in out


needs3 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 3, reason="needs 3 GPUs")


def _run_scripts(extra, tmp_path, port):
    """ps on gpu:0, chief on gpu:1, worker 1 on gpu:2 -- three OS processes started like m / w1 / w2."""
    common = ["--ps_hosts=gpu:0", "--worker_hosts=gpu:1,gpu:2", "--train_steps=6", "--init_stddev=0.05",
              "--validation_batch_size=500", "--validation_batches=2", "--step_timeout_secs=120"] + extra
    env = dict(os.environ, PYTHONPATH=ROOT, TFD_MASTER_PORT=str(port))
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, s)] + common, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)
             for s in ("mnist_python_m.py", "mnist_python_w1.py", "mnist_python_w2.py")]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=280)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    return outs


@needs3
def test_gpu_cluster_scripts_sync_and_resume(tmp_path):
    ps, w1, w2 = _run_scripts(["--train_dir=%s" % tmp_path], tmp_path, 29700)
    assert "job name = ps" in ps
    for o, t in ((w1, 0), (w2, 1)):
        assert "Worker %d: Session initialization complete." % t in o
        assert o.count(": training step") == 6 and "(global step: 6)" in o
        assert "Mean Accuracy :" in o and "Training elapsed time:" in o
    from tensorflow_distributed_b200 import checkpoint
    from tensorflow_distributed_b200.models.spec import cnn_spec
    blob = checkpoint.load(checkpoint.latest_checkpoint(str(tmp_path)), cnn_spec())
    assert blob["global_step"] == 6 and blob["adam"]["t"] == 6
    assert tuple(blob["variables"]["weights/wd1"].shape) == (3136, 1024)
    # resume: the checkpoint already satisfies train_steps, so no further step is taken
    ps, w1, w2 = _run_scripts(["--train_dir=%s" % tmp_path], tmp_path, 29710)
    assert w1.count(": training step") == 0 and "Mean Accuracy :" in w1


@needs3
def test_gpu_cluster_scripts_async(tmp_path):
    ps, w1, w2 = _run_scripts(["--sync_replicas=False"], tmp_path, 29720)
    n = w1.count(": training step") + w2.count(": training step")
    assert n >= 6  # global_step counts worker updates (m:301)
    assert "Mean Accuracy :" in w1 and "Mean Accuracy :" in w2


@needs2
@pytest.mark.parametrize("model", ["cnn", "mlp"])
def test_two_gpu_bf16_gradient_push(model):
    out = _torchrun(2, ["--layout", "colocated", "--model", model, "--grad_dtype", "bf16"], 29615)
    assert "OK" in out and "FAIL" not in out and "bf16-grad-push" in out


@needs2
def test_single_box_launcher_colocated(tmp_path):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29640", "-m", "tensorflow_distributed_b200.launch", "--layout=colocated",
           "--train_steps=20", "--init_stddev=0.05", "--validation_batch_size=500", "--validation_batches=2",
           "--train_dir=%s" % tmp_path]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("training step 20 done (global step: 20)") == 2
    accs = [float(l.split(":")[1]) for l in r.stdout.splitlines() if l.startswith("Mean Accuracy")]
    assert len(accs) == 2 and min(accs) > 0.9
    assert os.path.exists(os.path.join(str(tmp_path), "model.ckpt-20.pt"))


needs4 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 4, reason="needs 4 GPUs")


@needs4
def test_gpu_backup_workers_drop_a_stragglers_stale_gradients(tmp_path):
    """R-of-N on the GPU ps (m:62-65): 1 ps GPU + 3 worker GPUs, aggregate the first 2 gradients of every update;
    worker 2 is stalled on the host before every step, so most of its gradients arrive stale and are never read."""
    common = ["--ps_hosts=gpu:0", "--worker_hosts=gpu:1,gpu:2,gpu:3", "--train_steps=12", "--init_stddev=0.05",
              "--replicas_to_aggregate=2", "--validation_batch_size=500", "--validation_batches=2",
              "--step_timeout_secs=120", "--inject_fault=stall:2:0.05"]
    env = dict(os.environ, PYTHONPATH=ROOT, TFD_MASTER_PORT="29730")
    cmds = [("mnist_python_m.py", []), ("mnist_python_w1.py", []), ("mnist_python_w2.py", []),
            ("mnist_python_w2.py", ["--task_index=2"])]
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, s)] + common + e, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT) for s, e in cmds]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=280)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    steps = [o.count(": training step") for o in outs[1:]]
    assert steps[2] < min(steps[0], steps[1])          # the straggler did fewer local steps
    assert max(steps[0], steps[1]) >= 12                # the fast workers carried every update
    for o in outs[1:]:
        assert "Mean Accuracy :" in o
    assert any("(global step: 12)" in o for o in outs[1:])


def _multicast_supported() -> bool:
    """NVSwitch multicast (NVLS) on device 0, asked of the driver (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED)."""
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        _, dev = cu.cuDeviceGet(0)
        _, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(v)
    except Exception:  # noqa: BLE001
        return False


@needs2
@pytest.mark.parametrize("layout,model,precision", [("colocated", "cnn", "tf32"), ("dedicated", "cnn", "tf32"),
                                                    ("colocated", "cnn", "bf16"), ("colocated", "mlp", "tf32")])
def test_two_gpu_nvls_transport_matches_the_peer_memory_transport(layout, model, precision):
    """The NVSwitch-multicast transport (gradient sum in the switch by multimem.ld_reduce, one multimem.st publish)
    against the peer-memory transport: same seed, same batches, parameters / Adam slots / loss trajectory agree."""
    if not _multicast_supported():
        pytest.skip("this box exposes no NVSwitch multicast (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 0)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "scripts", "nvls_check.py"), "--layout", layout,
           "--model", model, "--precision", precision]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "OK" in r.stdout and "FAIL" not in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a GPU")
def test_cpu_ps_with_native_gpu_workers(tmp_path):
    """The reference's literal deployment (ps on a CPU host over the network transport, worker compute on a GPU):
    host:port cluster over gloo, --backend=cuda on the workers."""
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    common = ["--ps_hosts=127.0.0.1:%d" % port, "--worker_hosts=127.0.0.1:%d,127.0.0.1:%d" % (port + 1, port + 2),
              "--train_steps=6", "--validation_batches=1", "--validation_batch_size=200", "--init_stddev=0.05",
              "--num_gpus=1"]
    env = dict(os.environ, PYTHONPATH=ROOT)
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, sc)] + common + extra, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)
             for sc, extra in (("mnist_python_m.py", ["--backend=cpu"]), ("mnist_python_w1.py", ["--backend=cuda"]),
                               ("mnist_python_w2.py", ["--backend=cuda"]))]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "training step" in outs[1] and "Mean Accuracy" in outs[1] and "Mean Accuracy" in outs[2]
