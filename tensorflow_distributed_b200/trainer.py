"""Training drivers: the single-process trainer and the ps/worker cluster program.

``run_single`` reproduces ``/root/reference/mnist_single.py:90-134`` (and the
notebook): train while ``step*batch_size < training_iters``, print the minibatch
loss with keep_prob=1 on the *same* batch every ``display_step`` steps, then
validate ``nt`` batches and print mean accuracy and the two timings.  The
printed lines are the reference's.

``run_cluster`` is the one program behind ``mnist_python_{m,w1,w2}.py``
(m:132-320): role/task validation, cluster spec, ps that serves until released,
chief init / non-chief wait, sync or async replica loop ending on the *global*
step, wall-clock timing, 5x1000 validation.
"""
from __future__ import annotations

import sys
import time
from typing import Callable, List, Optional

import torch

from . import cluster as cl
from . import data as data_mod
from .config import Flags
from .engine import OracleEngine, make_engine
from .models.spec import make_spec
from .parallel.placement import Placement
from .supervisor import Supervisor
from .utils.trace import StepTracer


# --------------------------------------------------------------------------
# single process (mnist_single.py / convolutional_network.ipynb)
# --------------------------------------------------------------------------
def run_single(*, model: str = "cnn", learning_rate: float = 0.01, training_iters: int = 4000,
               batch_size: int = 128, display_step: int = 10, dropout: float = 0.75, train_flag: int = 1,
               nt: int = 50, val_batch: int = 100, data_dir: str = "MNIST_data/", synthetic: Optional[bool] = None,
               backend: str = "auto", seed: int = 0, hidden_units: int = 100, mlp_layers: int = 1,
               train_dir: str = "", init_stddev: float = 1.0, trace_file: str = "", trace_start: int = 2,
               trace_steps: int = 3, precision: Optional[str] = None, log: Callable[[str], None] = print) -> dict:
    mnist = data_mod.read_data_sets(data_dir, True, synthetic, seed)
    tracer = StepTracer(trace_file, trace_start, trace_steps, log)
    spec = make_spec(model, hidden_units, mlp_layers)
    start_time = time.time()  # the reference starts its timer before session creation (single:102)
    eng = make_engine(spec, backend, learning_rate=learning_rate, seed=seed, init_stddev=init_stddev,
                      precision=precision)
    sv = Supervisor(True, spec, train_dir, log=log) if train_dir else None
    if sv is not None:
        def _restore(blob):
            eng.set_params(blob["variables"])
            if "adam" in blob:
                eng.set_adam_state(blob["adam"])
            eng.global_step = blob["global_step"]
        sv.prepare_or_wait(init_fn=lambda: None, restore_fn=_restore, wait_fn=lambda s: None)
    losses: List[float] = []
    step = 1 + (eng.global_step if sv is not None else 0)
    if train_flag == 1:
        while step * batch_size < training_iters:
            batch_x, batch_y = mnist.train.next_batch(batch_size)
            tracer.before(step)
            eng.train_step(batch_x, batch_y, dropout)
            tracer.after(step)
            if step % display_step == 0:
                loss = eng.eval_loss(batch_x, batch_y)
                losses.append(loss)
                log("Iter " + str(step * batch_size) + ", Minibatch Loss= " + "{:.6f}".format(loss))
            step += 1
        eng.synchronize()
        tracer.close()
        log("Optimization Finished!")
        training_end_time = time.time()
        log("--- %s seconds Time for Training ---" % (training_end_time - start_time))
        if sv is not None:
            sv.save(eng.global_step, eng.get_params(), eng.get_adam_state())
    else:
        training_end_time = start_time
    accuracy_arr = []
    for _ in range(1, nt + 1):
        val_x, val_y = mnist.validation.next_batch(val_batch)
        accuracy_arr.append(eng.eval_accuracy(val_x, val_y))
    mean_accuracy = sum(accuracy_arr) / len(accuracy_arr)
    log("Mean Accuracy : %f" % mean_accuracy)
    testing_end_time = time.time()
    log("--- %s seconds Time for Inference ---" % (testing_end_time - training_end_time))
    return {"losses": losses, "mean_accuracy": mean_accuracy, "steps": step - 1,
            "train_seconds": training_end_time - start_time,
            "inference_seconds": testing_end_time - training_end_time, "engine": eng, "data_source": mnist.source,
            "trace": tracer.written}


# --------------------------------------------------------------------------
# cluster program (mnist_python_m.py / _w1.py / _w2.py)
# --------------------------------------------------------------------------
def _maybe_inject_fault(flags: Flags, local_step: int) -> None:
    if not flags.inject_fault:
        return
    kind, task, arg = flags.inject_fault.split(":")
    if int(task) != flags.task_index:
        return
    if kind == "stall":
        time.sleep(float(arg))
    elif kind == "kill" and local_step == int(arg):
        print("fault injection: worker %d exits at local step %d" % (flags.task_index, local_step))
        sys.stdout.flush()
        import os
        os._exit(17)


def run_cluster(flags: Flags, log: Callable[[str], None] = print) -> dict:
    mnist = data_mod.read_data_sets(flags.data_dir, True, True if flags.synthetic else None,
                                    flags.seed + 101 * flags.task_index * (flags.job_name == "worker"))
    if flags.download_only:
        raise SystemExit(0)
    cl.validate_role(flags)
    log("job name = %s" % flags.job_name)
    log("task index = %d" % flags.task_index)
    cluster = cl.ClusterSpec.from_flags(flags)
    if cluster.all_gpu:
        from .parallel.ps_gpu import run_gpu_cluster_task
        return run_gpu_cluster_task(flags, cluster, mnist, log)
    return _run_cpu_cluster_task(flags, cluster, mnist, log)


def _run_cpu_cluster_task(flags: Flags, cluster: cl.ClusterSpec, mnist, log) -> dict:
    from .parallel import ps_cpu

    num_workers = cluster.num_workers
    spec = make_spec(flags.model, flags.hidden_units, flags.mlp_layers)
    shards = flags.ps_shards or cluster.num_ps
    if shards != cluster.num_ps:
        raise ValueError("CPU transport runs one shard per ps task: ps_shards must equal len(ps_hosts)")
    placement = Placement(spec, shards)
    R = cl.effective_replicas_to_aggregate(flags, num_workers)
    if flags.log_device_placement:
        log(placement.table())

    if flags.job_name == "ps":
        # tf.train.Server + server.join() (m:158-161).  The server is long-lived: it serves `server_generations`
        # successive sets of client processes (0 = until SHUTDOWN) and keeps variables, Adam slots and global_step
        # between them -- what lets a worker started with --existing_servers=True attach to servers that outlive their
        # clients (m:76-80, m:268-273).  One gloo group per generation, same endpoint.
        server = ps_cpu.PSServer(cluster, spec, placement, flags.task_index, flags.sync_replicas, R,
                                 flags.learning_rate)
        generation, why = 0, "done"
        while True:
            ps_cpu.init_process_group(cluster, flags.job_name, flags.task_index,
                                      flags.step_timeout_secs if generation == 0 else 24 * 3600.0)
            if generation:
                server.new_generation()
                log("ps %d: serving client generation %d from global step %d" % (flags.task_index, generation,
                                                                               server.global_step))
            why = server.serve_forever()
            torch.distributed.destroy_process_group()
            generation += 1
            if why == "shutdown" or (flags.server_generations and generation >= flags.server_generations):
                break
        return {"role": "ps", "global_step": server.global_step, "served": server.served, "generations": generation}

    if flags.existing_servers:
        host, port = cluster.task_address("worker", flags.task_index).host_port
        log("Using existing server at: %s" % ("grpc://%s:%d" % (host, port)))
    ps_cpu.init_process_group(cluster, flags.job_name, flags.task_index, flags.step_timeout_secs)

    is_chief = cl.is_chief(flags)
    device = cl.worker_device(flags, cluster)
    native_worker = flags.backend == "cuda"
    if native_worker:
        # The reference's literal deployment: ps tasks on CPU hosts, worker compute on a GPU (m:164-177).  Parameters
        # arrive over gloo, go H2D, the sm_100a kernels compute the gradients, which go D2H and back over gloo.
        from .engine_native import CTRL_T, NativeEngine
        dev = cl.torch_device_of(device)
        eng = NativeEngine(spec, flags.learning_rate, flags.seed + 31 * flags.task_index, flags.init_stddev,
                           device=dev if dev != "cpu" else "cuda:0", precision=flags.precision)
    else:
        eng = OracleEngine(spec, flags.learning_rate, flags.seed, flags.init_stddev, cl.torch_device_of(device)
                           if torch.cuda.is_available() else "cpu")
        # per-worker dropout stream
        eng._gen.manual_seed(flags.seed + 7919 + 31 * flags.task_index)
    client = ps_cpu.PSClient(cluster, spec, placement, flags.task_index, device_filters=cl.device_filters(flags))
    sv = Supervisor(is_chief, spec, flags.train_dir, recovery_wait_secs=1.0, log=log)

    if is_chief:
        log("Worker %d: Initializing session..." % flags.task_index)
    else:
        log("Worker %d: Waiting for session to be initialized..." % flags.task_index)

    def _restore(blob):
        ad = blob.get("adam")
        zeros = {k: torch.zeros_like(v) for k, v in blob["variables"].items()}
        client.set_state(blob["global_step"], ad["t"] if ad else 0, blob["variables"],
                         ad["m"] if ad else zeros, ad["v"] if ad else zeros)

    if flags.existing_servers and client.ready():
        # the servers outlived their previous clients: their variables are initialised and trained; attach as they are
        log("Worker %d: attached to running servers at global step %d" % (flags.task_index, client.global_step()))
    else:
        sv.prepare_or_wait(init_fn=lambda: client.init_variables(eng.get_params()), restore_fn=_restore,
                           wait_fn=lambda s: client.wait_ready(s, flags.step_timeout_secs))
    if sv.restored_from:
        log("Worker %d: restored variables from %s" % (flags.task_index, sv.restored_from))
    log("Worker %d: Session initialization complete." % flags.task_index)

    if flags.sync_replicas and is_chief:
        # sess.run(sync_init_op); sv.start_queue_runners (m:279-282).  Aggregation itself runs on the ps.
        client.init_tokens(flags.sync_init_tokens)

    time_begin = time.time()
    log("Training begins @ %f" % time_begin)
    local_step = 0
    last_saved = 0
    token = client.global_step()  # local_step_init_op: local step starts at the global step
    step = token
    tracer = StepTracer(flags.trace_file and "%s.worker%d" % (flags.trace_file, flags.task_index), flags.trace_start,
                        flags.trace_steps, log)
    while True:
        batch_xs, batch_ys = mnist.train.next_batch(flags.batch_size)
        tracer.before(local_step + 1)
        params = client.pull()
        _maybe_inject_fault(flags, local_step)
        if native_worker:
            eng.set_params(params)
            eng.shard.ctrl[CTRL_T] = local_step  # the dropout stream is indexed by this word; no Adam runs here
            _, grads, _ = eng.loss_and_grads(batch_xs, batch_ys, flags.keep_prob)
        else:
            _, grads, _ = eng.loss_and_grads(batch_xs, batch_ys, flags.keep_prob, params=params)
        if flags.sync_replicas:
            client.push(grads, token, sync=True)
            token = client.dequeue_token()
            step = client.global_step()
        else:
            step = client.push(grads, 0, sync=False)
        local_step += 1
        tracer.after(local_step)
        now = time.time()
        log("%f: Worker %d: training step %d done (global step: %d)" % (now, flags.task_index, local_step, step))
        if is_chief and flags.train_dir and flags.save_checkpoint_steps > 0 and \
                step // flags.save_checkpoint_steps > last_saved // flags.save_checkpoint_steps:
            # the Supervisor's periodic saver (every 600 s in TF, m:239-253); here every N global steps
            st = client.get_state()
            sv.save(st["global_step"], st["params"], {"t": st["t"], "m": st["m"], "v": st["v"],
                                                      "beta1_power": 0.9 ** st["t"], "beta2_power": 0.999 ** st["t"]})
            last_saved = step
        if step >= flags.train_steps:
            break
    tracer.close()
    time_end = time.time()
    log("Training ends @ %f" % time_end)
    training_time = time_end - time_begin
    log("Training elapsed time: %f s" % training_time)

    saved = None
    if is_chief and flags.train_dir:
        st = client.get_state()
        saved = sv.save(st["global_step"], st["params"],
                        {"t": st["t"], "m": st["m"], "v": st["v"],
                         "beta1_power": 0.9 ** st["t"], "beta2_power": 0.999 ** st["t"]})

    eng.set_params(client.pull())
    accuracy_arr = []
    for _ in range(1, flags.validation_batches + 1):
        val_x, val_y = mnist.validation.next_batch(flags.validation_batch_size)
        acc = eng.eval_accuracy(val_x, val_y)
        log("After %d training step(s)" % flags.train_steps)  # the reference's print never formats (m:316)
        log("Accuracy : %f" % acc)
        accuracy_arr.append(acc)
    mean_accuracy = sum(accuracy_arr) / len(accuracy_arr)
    log("Mean Accuracy : %f" % mean_accuracy)
    final_params = eng.get_params()
    client.done()
    return {"role": "worker", "task_index": flags.task_index, "local_steps": local_step, "global_step": step,
            "mean_accuracy": mean_accuracy, "training_time": training_time, "params": final_params,
            "checkpoint": saved}
