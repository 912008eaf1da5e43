"""Multi-GPU ps/worker check, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/cluster_check.py --layout colocated

Every rank trains a few synchronous steps through GpuClusterEngine; rank 0 then checks the pushed gradients
against the CPU oracle, the Adam update against the average of the pushed gradients, and the counters."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tensorflow_distributed_b200.models.spec import make_spec  # noqa: E402
from tensorflow_distributed_b200.ops import oracle  # noqa: E402
from tensorflow_distributed_b200.parallel.ps_gpu import GpuClusterEngine, Topology  # noqa: E402


def rel(a, b):
    """Relative Frobenius error."""
    return (a.float() - b.float()).norm().item() / (b.float().norm().item() + 1e-20)


def relmax(a, b):
    """Largest element error relative to the largest element of the reference."""
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layout", default="colocated")
    ap.add_argument("--ps", type=int, default=0)
    ap.add_argument("--model", default="cnn")
    ap.add_argument("--hidden", type=int, default=512)
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--async_mode", action="store_true")
    ap.add_argument("--pull", default="auto")
    ap.add_argument("--grad_dtype", default="fp32")
    ap.add_argument("--replicas", type=int, default=0)
    ap.add_argument("--precision", default=None, choices=["tf32", "bf16"])
    ap.add_argument("--digest", action="store_true", help="rank 0 prints a SHA-256 of the final parameters and Adam slots")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda:%d" % local)
    dist.init_process_group("nccl", device_id=device)
    spec = make_spec(a.model, a.hidden, a.layers)
    topo = Topology.make(world, a.layout, a.ps)
    eng = GpuClusterEngine(spec, topo, rank, device, learning_rate=0.01, seed=3, init_stddev=0.05,
                           async_mode=a.async_mode, pull_mode=a.pull, grad_dtype=a.grad_dtype,
                           replicas_to_aggregate=a.replicas, precision=a.precision)
    B = 128
    keep = 1.0
    w = topo.worker_of(rank)
    from tensorflow_distributed_b200 import data
    stream = data.read_data_sets("", synthetic=True, seed=100 + (w if w is not None else 99)).train
    batches = [stream.next_batch(B) for _ in range(a.steps)]  # learnable synthetic digits, one stream per worker
    xs, ys = [b[0] for b in batches], [b[1] for b in batches]
    p0 = eng.get_params()
    dist.barrier()
    # ---- step 1 (eager) ----
    l1 = eng.train_step(xs[0], ys[0], keep, fetch_loss=True)
    torch.cuda.synchronize()
    dist.barrier()
    ok = True
    msgs = []
    if rank == 0 and not a.async_mode:
        # Every worker's pushed gradient against the PURE fp32 oracle on that worker's batch.  Budget per tensor = what
        # quantising the tensor-core operands alone does to the oracle (oracle.quantizer), x3 for the second rounding
        # the kernels add (dy operands, stored activations, bf16 slots), checked in the Frobenius norm AND element-wise.
        slots = [eng.get_slot_grads(i) for i in range(topo.num_workers)]
        xs_all, ys_all = [], []
        for i in range(topo.num_workers):
            bx, by = data.read_data_sets("", synthetic=True, seed=100 + i).train.next_batch(B)
            x, y = torch.from_numpy(bx), torch.from_numpy(by)
            xs_all.append(x); ys_all.append(y)
            _, og, _ = oracle.loss_and_grads(a.model, x, y, p0)
            _, oq, _ = oracle.loss_and_grads(a.model, x, y, p0, precision=eng.precision)
            for k in og:
                budget = 3.0 * rel(oq[k], og[k]) + (2e-2 if a.grad_dtype == "bf16" else 2e-3)
                e, em = rel(slots[i][k], og[k]), relmax(slots[i][k], og[k])
                if e > budget or em > 2.0 * budget + 1e-2:
                    ok = False
                    msgs.append("worker %d grad %s frob %.4f max %.4f (budget %.4f)" % (i, k, e, em, budget))
        # SURVEY section 4, "Distributed": the N-worker synchronous update equals the single-process update on the
        # concatenated N*B batch (mean over workers of per-worker mean gradients == large-batch mean gradient)
        _, og_big, _ = oracle.loss_and_grads(a.model, torch.cat(xs_all), torch.cat(ys_all), p0)
        _, oq_big, _ = oracle.loss_and_grads(a.model, torch.cat(xs_all), torch.cat(ys_all), p0, precision=eng.precision)
        avg_chk = {k: sum(s[k] for s in slots) / topo.num_workers for k in slots[0]}
        for k in og_big:
            budget = 3.0 * rel(oq_big[k], og_big[k]) + (2e-2 if a.grad_dtype == "bf16" else 2e-3)
            if rel(avg_chk[k], og_big[k]) > budget:
                ok = False
                msgs.append("large-batch equivalence %s frob %.4f (budget %.4f)" % (k, rel(avg_chk[k], og_big[k]), budget))
        avg = {k: sum(s[k] for s in slots) / topo.num_workers for k in slots[0]}
        ref = {k: v.clone() for k, v in p0.items()}
        oracle.TFAdam(0.01).apply(ref, avg)
        p1 = eng.get_params()
        for k in ref:
            if not torch.allclose(p1[k], ref[k], atol=2e-5 if a.grad_dtype == "fp32" else 2e-4, rtol=1e-4):
                ok = False
                msgs.append("adam mismatch %s max diff %.3e" % (k, (p1[k] - ref[k]).abs().max().item()))
        if eng.global_step != 1:
            ok = False
            msgs.append("global_step %d != 1" % eng.global_step)
    dist.barrier()
    # ---- more steps (second is eager warm-up of the graph path, then captured graphs) ----
    losses = [l1]
    for s in range(1, a.steps):
        losses.append(eng.train_step(xs[s], ys[s], keep, fetch_loss=True))
    torch.cuda.synchronize()
    dist.barrier()
    gs = eng.global_step
    acc = eng.eval_accuracy(xs[0], ys[0]) if eng.is_worker else 0.0
    if rank == 0:
        # sync: one apply per global step.  async: every pushed gradient applied exactly once (no lost update)
        want = a.steps * topo.num_workers if a.async_mode else a.steps
        if gs != want:
            ok = False
            msgs.append("global_step %d != %d" % (gs, want))
    if eng.is_worker and not (losses[-1] < losses[0]):
        ok = False
        msgs.append("rank %d loss did not decrease: %.4f -> %.4f" % (rank, losses[0], losses[-1]))
    if a.digest:
        import hashlib
        final = eng.get_params()
        adam = eng.get_adam_state()
        if rank == 0:
            h = hashlib.sha256()
            for k in sorted(final):
                h.update(final[k].detach().float().cpu().contiguous().numpy().tobytes())
            for slot in ("m", "v"):
                for k in sorted(adam[slot]):
                    h.update(adam[slot][k].detach().float().cpu().contiguous().numpy().tobytes())
            print("DIGEST %s" % h.hexdigest(), flush=True)
    flag = torch.tensor([0 if ok else 1], device=device)
    dist.all_reduce(flag)
    print("rank %d %s %s losses %.4f -> %.4f acc %.3f gs %d %s" % (rank, eng.describe(), "OK" if ok else "FAIL",
                                                                 losses[0], losses[-1], acc, gs, "; ".join(msgs)), flush=True)
    eng.shutdown()
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
