"""Config 2: 1 ps + 2 workers as three OS processes on CPU/gloo, launched like m / w1 / w2."""
import os
import re
import subprocess
import sys

import pytest
import torch

from tensorflow_distributed_b200 import checkpoint, data
from tensorflow_distributed_b200.models.spec import make_spec
from tensorflow_distributed_b200.ops import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(script, port, extra, env_extra=None):
    args = [sys.executable, os.path.join(ROOT, script), "--ps_hosts=127.0.0.1:%d" % port,
            "--worker_hosts=127.0.0.1:%d,127.0.0.1:%d" % (port + 1, port + 2), "--backend=cpu",
            "--step_timeout_secs=90"] + extra
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    env.update(env_extra or {})
    return subprocess.Popen(args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)


def _run3(port, extra, timeout=240):
    procs = [_launch(s, port, extra) for s in ("mnist_python_m.py", "mnist_python_w1.py", "mnist_python_w2.py")]
    outs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(o)
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
    return outs


@pytest.mark.timeout(300)
def test_sync_two_workers_equals_large_batch_single_process(tmp_path, free_port):
    steps = 3
    extra = ["--model=mlp", "--hidden_units=32", "--init_stddev=0.1", "--train_steps=%d" % steps, "--keep_prob=1.0",
             "--sync_init_tokens=0", "--batch_size=64", "--validation_batch_size=200", "--train_dir=%s" % tmp_path,
             "--log_device_placement", "--trace_file=%s" % (tmp_path / "tr.json"), "--trace_start=2", "--trace_steps=1"]
    ps_out, w1_out, w2_out = _run3(free_port, extra)
    for t in (0, 1):  # every worker traced its own step 2 (pull, forward/backward on the oracle, push)
        assert os.path.getsize(str(tmp_path / ("tr.json.worker%d" % t))) > 0
        assert "trace: steps 2..2 written to" in (w1_out, w2_out)[t]
    # log lines of the reference
    assert "job name = ps" in ps_out and "task index = 0" in ps_out
    assert "Worker 0: Initializing session..." in w1_out
    assert "Worker 1: Waiting for session to be initialized..." in w2_out
    for o, t in ((w1_out, 0), (w2_out, 1)):
        assert "Worker %d: Session initialization complete." % t in o
        assert "Training begins @" in o and "Training elapsed time:" in o and "Mean Accuracy :" in o
        assert o.count(": training step") == steps
        assert "training step %d done (global step: %d)" % (steps, steps) in o
    assert "weights/h1" in w1_out  # placement table
    # oracle: one process, both workers' batches, averaged gradient, one TF-Adam apply per global step
    spec = make_spec("mlp", 32, 1)
    p = oracle.init_params(spec, 0, 0.1)
    adam = oracle.TFAdam(0.01)
    streams = [data.read_data_sets("", synthetic=True, seed=101 * w).train for w in range(2)]
    for _ in range(steps):
        gs = []
        for s in streams:
            x, y = s.next_batch(64)
            _, g, _ = oracle.loss_and_grads("mlp", torch.from_numpy(x), torch.from_numpy(y), p)
            gs.append(g)
        adam.apply(p, {k: (gs[0][k] + gs[1][k]) / 2 for k in p})
    blob = checkpoint.load(checkpoint.latest_checkpoint(str(tmp_path)), spec)
    assert blob["global_step"] == steps and blob["adam"]["t"] == steps
    for k in p:
        assert torch.allclose(blob["variables"][k], p[k], atol=1e-6), k


@pytest.mark.timeout(300)
def test_async_two_workers_counts_every_update(free_port):
    steps = 6
    extra = ["--model=mlp", "--hidden_units=16", "--init_stddev=0.1", "--train_steps=%d" % steps,
             "--sync_replicas=False", "--batch_size=32", "--validation_batch_size=100"]
    ps_out, w1_out, w2_out = _run3(free_port, extra)
    n = w1_out.count(": training step") + w2_out.count(": training step")
    # global_step counts worker updates (m:301): total applies >= train_steps, and nobody waits for anybody
    assert n >= steps
    assert "global step: %d" % n in (w1_out + w2_out) or n > steps


@pytest.mark.timeout(300)
def test_cnn_default_tokens_two_ps_shards(free_port):
    """The reference's CNN with TF's default token preload, wd1 split across two ps tasks."""
    port = free_port
    common = ["--ps_hosts=127.0.0.1:%d,127.0.0.1:%d" % (port, port + 3), "--train_steps=2", "--batch_size=16",
              "--validation_batch_size=50", "--validation_batches=1"]
    scripts = [("mnist_python_m.py", []), ("mnist_python_m.py", ["--task_index=1"]),
               ("mnist_python_w1.py", []), ("mnist_python_w2.py", [])]
    procs = []
    for s, e in scripts:
        args = [sys.executable, os.path.join(ROOT, s), "--worker_hosts=127.0.0.1:%d,127.0.0.1:%d" % (port + 1, port + 2),
                "--backend=cpu", "--step_timeout_secs=90"] + common + e
        procs.append(subprocess.Popen(args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                      env=dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2"), cwd=ROOT))
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
    assert "Mean Accuracy :" in outs[2] and "Mean Accuracy :" in outs[3]


@pytest.mark.timeout(300)
def test_backup_workers_tolerate_a_straggler(free_port):
    """R-of-N (m:62-65): 3 workers, aggregate the fastest 2; worker 2 is stalled by fault injection.
    Training still reaches the global step; the straggler's stale gradients are dropped, it rejoins via tokens."""
    port = free_port
    hosts = ["--ps_hosts=127.0.0.1:%d" % port,
             "--worker_hosts=127.0.0.1:%d,127.0.0.1:%d,127.0.0.1:%d" % (port + 1, port + 2, port + 3)]
    common = hosts + ["--backend=cpu", "--step_timeout_secs=120", "--model=mlp", "--hidden_units=16", "--init_stddev=0.1",
                      "--train_steps=4", "--replicas_to_aggregate=2", "--batch_size=32", "--validation_batch_size=100",
                      "--validation_batches=1", "--inject_fault=stall:2:1.5"]
    cmds = [("mnist_python_m.py", []), ("mnist_python_w1.py", []), ("mnist_python_w2.py", []),
            ("mnist_python_w2.py", ["--task_index=2"])]
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, s)] + common + e, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True, env=dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2"),
                              cwd=ROOT) for s, e in cmds]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
    fast = outs[1].count(": training step") + outs[2].count(": training step")
    slow = outs[3].count(": training step")
    assert fast >= 2 * 3 and slow < outs[1].count(": training step")  # the stalled worker did fewer steps
    for o in outs[1:]:
        assert "Mean Accuracy :" in o


@pytest.mark.timeout(300)
def test_kill_and_resume_from_periodic_checkpoint(tmp_path, free_port):
    """the chief saves every 2 global steps; the worker process dies at its 4th step; the ps notices
    the dead peer and exits with an error instead of hanging; a relaunch restores step 2's checkpoint and finishes."""
    def launch(port, extra):
        base = ["--ps_hosts=127.0.0.1:%d" % port, "--worker_hosts=127.0.0.1:%d" % (port + 1), "--backend=cpu",
                "--step_timeout_secs=60", "--model=mlp", "--hidden_units=16", "--init_stddev=0.1", "--train_steps=5",
                "--replicas_to_aggregate=1", "--batch_size=32", "--validation_batch_size=100", "--validation_batches=1",
                "--train_dir=%s" % tmp_path, "--save_checkpoint_steps=2"] + extra
        env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
        return [subprocess.Popen([sys.executable, os.path.join(ROOT, s)] + base, stdout=subprocess.PIPE,
                                 stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)
                for s in ("mnist_python_m.py", "mnist_python_w1.py")]
    ps, w = launch(free_port, ["--inject_fault=kill:0:3"])
    wout = w.communicate(timeout=120)[0]
    psout = ps.communicate(timeout=120)[0]
    assert w.returncode == 17 and "fault injection" in wout
    assert ps.returncode != 0  # failure detection: the ps does not sit in join() forever once its only worker is gone
    latest = checkpoint.latest_checkpoint(str(tmp_path))
    assert latest is not None and latest.endswith("model.ckpt-2.pt")
    # relaunch (new ports): restore, continue from global step 2, finish at 5
    ps, w = launch(free_port + 7, [])
    wout = w.communicate(timeout=120)[0]
    psout = ps.communicate(timeout=120)[0]
    assert w.returncode == 0 and ps.returncode == 0, wout + psout
    assert "restored variables from" in wout and "model.ckpt-2.pt" in wout
    assert "training step 1 done (global step: 3)" in wout and "(global step: 5)" in wout
    assert wout.count(": training step") in (3, 4)  # the TF token preload may let the worker run one step ahead
    assert checkpoint.latest_checkpoint(str(tmp_path)).endswith("model.ckpt-5.pt")


@pytest.mark.timeout(400)
def test_existing_servers_attach_to_a_server_that_outlives_its_clients(free_port):
    """m:76-80 / m:268-273: with --existing_servers a worker does not bring a server up, it attaches to one that is
    already running.  The ps task here serves two successive sets of worker processes and keeps variables, Adam slots
    and global_step in between: the second set finds an initialised, trained model and continues from its step."""
    common = ["--model=mlp", "--hidden_units=16", "--init_stddev=0.1", "--keep_prob=1.0", "--batch_size=32",
              "--validation_batch_size=100", "--validation_batches=1"]
    ps = _launch("mnist_python_m.py", free_port, common + ["--server_generations=2", "--train_steps=3"])
    first = [_launch(s, free_port, common + ["--train_steps=3"]) for s in ("mnist_python_w1.py", "mnist_python_w2.py")]
    outs1 = [p.communicate(timeout=200)[0] for p in first]
    assert all(p.returncode == 0 for p in first), "\n".join(outs1)
    assert "Worker 0: Initializing session..." in outs1[0] and "(global step: 3)" in outs1[0]
    assert ps.poll() is None  # the server is still there
    second = [_launch(s, free_port, common + ["--train_steps=6", "--existing_servers=True"])
              for s in ("mnist_python_w1.py", "mnist_python_w2.py")]
    outs2 = [p.communicate(timeout=200)[0] for p in second]
    assert all(p.returncode == 0 for p in second), "\n".join(outs2)
    ps_out = ps.communicate(timeout=60)[0]
    assert ps.returncode == 0, ps_out
    assert "ps 0: serving client generation 1 from global step 3" in ps_out
    for t, o in enumerate(outs2):
        assert "Using existing server at: grpc://127.0.0.1:%d" % (free_port + 1 + t) in o
        assert "Worker %d: attached to running servers at global step 3" % t in o
        # the counter a worker reads after its first push may already hold the other worker's push, so only the
        # order-independent facts are checked
        first_gs = int(re.search(r"training step 1 done \(global step: (\d+)\)", o).group(1))
        assert first_gs in (3, 4, 5), o  # 3: ran on a preloaded token before the first update landed
        assert "After 6 training step(s)" in o
        assert 3 <= o.count(": training step") <= 6  # (preloaded tokens let a worker run ahead of the update)
