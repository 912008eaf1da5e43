"""End-to-end GPU engine vs the CPU oracle: gradients of a full step, Adam trajectory, eval, trainer."""
import os

import numpy as np

import pytest
import torch

from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.ops import oracle

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


def _batch(B, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 784, generator=g)
    y = torch.eye(10)[torch.randint(0, 10, (B,), generator=g)]
    return x, y


def _engine(spec, **kw):
    from tensorflow_distributed_b200.engine_native import NativeEngine
    return NativeEngine(spec, **kw)


def _frob(a, b):
    return (a.float() - b.float()).norm().item() / (b.float().norm().item() + 1e-20)


ARMS = [("tf32", False), ("bf16", False), ("bf16", True)]  # (precision, conv2 on the CUDA-core validation kernels)


def _entitled(g_fp32, g_quant, k):
    """Error budget of a precision arm for gradient k: what quantising the tensor-core operands alone does to the fp32
    oracle (oracle.quantizer), times a factor for the second rounding the kernels add (dy operands of the backward
    GEMMs, bf16-stored activations), plus accumulation-order noise."""
    return 3.0 * _frob(g_quant[k], g_fp32[k]) + 2e-3


@pytest.mark.parametrize("precision,conv2_ref", ARMS)
@pytest.mark.parametrize("keep_prob", [1.0, 0.75])
def test_cnn_step_gradients_match_oracle(precision, conv2_ref, keep_prob):
    """Whole-step gradients against the PURE fp32 oracle (its weights are not rounded to what the tensor cores see):
    the tolerance is what the arm's operand quantisation is entitled to, measured on the oracle itself."""
    spec = cnn_spec()
    eng = _engine(spec, init_stddev=0.05, seed=1, conv2_ref=conv2_ref, use_graphs=False, precision=precision)
    x, y = _batch(128, 1)
    loss, grads, mask = eng.loss_and_grads(x, y, keep_prob)
    p = eng.get_params()
    km = None if mask is None else mask.bool().cpu()
    oloss, g32, _ = oracle.loss_and_grads("cnn", x, y, p, keep_prob, km)
    _, gq, _ = oracle.loss_and_grads("cnn", x, y, p, keep_prob, km, precision=precision)
    assert abs(loss - oloss.item()) < (2e-3 if precision == "tf32" else 2e-2) * max(1.0, abs(oloss.item()))
    for k in g32:
        assert _frob(grads[k], g32[k]) < _entitled(g32, gq, k), (k, _frob(grads[k], g32[k]), _frob(gq[k], g32[k]))
        # no single element is far off either (max-abs error relative to the largest gradient of the tensor)
        assert _rel(grads[k], g32[k]) < 2.0 * _entitled(g32, gq, k) + 1e-2, (k, _rel(grads[k], g32[k]))


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
def test_cnn_training_trajectory_tracks_oracle_and_learns(precision):
    from tensorflow_distributed_b200 import data
    spec = cnn_spec()
    eng = _engine(spec, init_stddev=0.05, seed=2, precision=precision)
    ds = data.read_data_sets("", synthetic=True, seed=0)
    first = last = None
    for i in range(30):
        x, y = ds.train.next_batch(128)
        l = eng.train_step(x, y, 0.75, fetch_loss=True)
        first = l if first is None else first
        last = l
    assert eng.global_step == 30
    assert last < first
    vx, vy = ds.validation.next_batch(1000)
    assert eng.eval_accuracy(vx, vy) > 0.9
    assert eng.eval_loss(vx, vy) >= 0.0


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
def test_twenty_steps_at_the_reference_init_against_the_fp32_oracle(precision):
    """The regime the reference and bench.py actually run in: every parameter ~ N(0, 1) (m:186-195), so fc1
    activations are in the hundreds and the first losses in the thousands.  20 training steps (dropout off so both
    sides see the same function) from the same weights and batches: the loss trajectory and the final weights of each
    precision arm against the fp32 CPU oracle.  Bounds = a multiple of what the arm's operand quantisation does to the
    ORACLE's own trajectory (measured here, not guessed)."""
    spec = cnn_spec()
    eng = _engine(spec, init_stddev=1.0, seed=11, precision=precision)
    p0 = eng.get_params()
    batches = [_batch(128, 100 + i) for i in range(20)]

    def run_oracle(prec):
        params = {k: v.clone() for k, v in p0.items()}
        adam, losses = oracle.TFAdam(0.01), []
        for x, y in batches:
            l, g, _ = oracle.loss_and_grads("cnn", x, y, params, precision=prec)
            adam.apply(params, g)
            losses.append(float(l))
        return losses, params

    l32, w32 = run_oracle(None)
    lq, wq = run_oracle(precision)
    le = [eng.train_step(x, y, 1.0, fetch_loss=True) for x, y in batches]
    we = eng.get_params()
    rel_loss = max(abs(a - b) / max(abs(b), 1.0) for a, b in zip(le, l32))
    rel_loss_q = max(abs(a - b) / max(abs(b), 1.0) for a, b in zip(lq, l32))
    upd = lambda w, k: w[k] - p0[k]  # noqa: E731  (20 Adam steps move a weight by at most 0.2: compare the UPDATE)
    print("precision %s: max rel loss error %.3e (oracle-with-quantised-operands: %.3e)" % (precision, rel_loss, rel_loss_q))
    assert le[0] > 100.0 and all(torch.isfinite(torch.tensor(le)))
    assert rel_loss < 3.0 * rel_loss_q + (5e-3 if precision == "tf32" else 2e-2)
    for k in w32:
        e, eq = _frob(upd(we, k), upd(w32, k)), _frob(upd(wq, k), upd(w32, k))
        print("  %-12s update error vs fp32 oracle %.3e (entitled %.3e)" % (k, e, eq))
        assert e < 3.0 * eq + 5e-2, (k, e, eq)


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
def test_adam_step_matches_tf_adam_on_engine_grads(precision):
    spec = cnn_spec()
    eng = _engine(spec, init_stddev=0.05, seed=3, use_graphs=False, precision=precision)
    x, y = _batch(128, 3)
    p0 = eng.get_params()
    eng.train_step(x, y, 1.0)
    g = eng.get_grads()
    p1 = eng.get_params()
    adam = oracle.TFAdam(0.01)
    ref = {k: v.clone() for k, v in p0.items()}
    adam.apply(ref, g)
    for k in ref:
        assert torch.allclose(p1[k], ref[k], atol=2e-5, rtol=1e-4), k
    if precision == "bf16":  # the published bf16 shadow is the rounded master copy
        assert torch.equal(eng.S("weights/wd1").cpu(), p1["weights/wd1"].to(torch.bfloat16))
    else:                    # tf32: no shadow at all, the tensor cores read the master
        assert eng.S("weights/wd1").data_ptr() == eng.P("weights/wd1").data_ptr()


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
@pytest.mark.parametrize("hidden,layers", [(100, 1), (256, 3)])
def test_mlp_step_gradients_match_oracle(hidden, layers, precision):
    spec = mlp_spec(hidden, layers)
    eng = _engine(spec, init_stddev=0.05, seed=4, use_graphs=False, precision=precision)
    x, y = _batch(128, 4)
    loss, grads, _ = eng.loss_and_grads(x, y, 1.0)
    p = eng.get_params()
    oloss, g32, _ = oracle.loss_and_grads("mlp", x, y, p)
    _, gq, _ = oracle.loss_and_grads("mlp", x, y, p, precision=precision)
    assert abs(loss - oloss.item()) < (2e-3 if precision == "tf32" else 2e-2) * max(1.0, abs(oloss.item()))
    for k in g32:
        assert _frob(grads[k], g32[k]) < _entitled(g32, gq, k), (k, _frob(grads[k], g32[k]), _frob(gq[k], g32[k]))


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
def test_graph_replay_equals_eager_bit_for_bit(precision):
    """Every reduction of the step has a fixed order (split-K planes, per-block partials summed by the last block,
    the conv2 wgrad plane reduction), so replaying the captured graph and launching the kernels one by one give the
    same bits."""
    spec = cnn_spec()
    x, y = _batch(128, 5)
    outs = []
    for graphs in (False, True):
        eng = _engine(spec, init_stddev=0.05, seed=5, use_graphs=graphs, precision=precision)
        for _ in range(5):
            eng.train_step(x, y, 0.75)
        outs.append((eng.get_params(), eng.get_adam_state()))
    for k in outs[0][0]:
        assert torch.equal(outs[0][0][k], outs[1][0][k]), k
        assert torch.equal(outs[0][1]["m"][k], outs[1][1]["m"][k]), k


def test_run_single_on_gpu_prints_reference_lines():
    from tensorflow_distributed_b200.trainer import run_single
    lines = []
    out = run_single(backend="cuda", synthetic=True, nt=5, val_batch=100, init_stddev=0.05, log=lines.append)
    assert out["steps"] == 31 and out["engine"].name == "native-sm100a"
    assert sum(l.startswith("Iter ") for l in lines) == 3 and "Optimization Finished!" in lines
    assert out["mean_accuracy"] > 0.5


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
def test_checkpoint_resume_gives_the_same_next_step(tmp_path, precision):
    """save -> (process would die here) -> restore into a fresh engine -> identical next step (bit for
    bit: the reductions are deterministic)."""
    from tensorflow_distributed_b200 import checkpoint
    spec = cnn_spec()
    x, y = _batch(128, 7)
    a = _engine(spec, init_stddev=0.05, seed=7, use_graphs=False, precision=precision)
    for _ in range(3):
        a.train_step(x, y, 1.0)
    path = checkpoint.save(str(tmp_path), spec, a.global_step, a.get_params(), a.get_adam_state())
    blob = checkpoint.load(checkpoint.latest_checkpoint(str(tmp_path)), spec)
    b = _engine(spec, init_stddev=0.05, seed=99, use_graphs=False, precision=precision)  # everything comes from the file
    b.set_params(blob["variables"])
    b.set_adam_state(blob["adam"])
    b.global_step = blob["global_step"]
    assert b.global_step == 3 and b.get_adam_state()["t"] == 3
    a.train_step(x, y, 1.0)
    b.train_step(x, y, 1.0)
    pa, pb = a.get_params(), b.get_params()
    for k in pa:
        assert torch.equal(pa[k], pb[k]), k
    assert a.global_step == b.global_step == 4


def test_native_predict_proba_matches_the_oracle():
    import numpy as np
    spec = cnn_spec()
    eng = _engine(spec, init_stddev=0.05, seed=4)
    rng = np.random.RandomState(0)
    x = rng.rand(50, 784).astype(np.float32)
    p = eng.predict_proba(x)
    params = eng.get_params()
    want = torch.softmax(oracle.forward("cnn", torch.from_numpy(x), params), dim=1).numpy()
    assert p.shape == (50, 10) and np.allclose(p.sum(axis=1), 1.0, atol=1e-4)
    assert np.abs(p - want).max() < 0.03 and (p.argmax(1) == want.argmax(1)).mean() > 0.9


def test_train_step_back_to_back_feeds_the_batch_it_was_given():
    """train_step(fetch_loss=False) does not synchronise, so the host runs ahead of the ~0.1 ms device step; the pinned
    staging buffers are double-buffered behind events and must never be overwritten while an H2D copy that reads
    them is still queued (round-1 advisor finding: torn / duplicated batches)."""
    from tensorflow_distributed_b200.engine_native import NativeEngine
    from tensorflow_distributed_b200.models.spec import cnn_spec
    eng = NativeEngine(cnn_spec(), learning_rate=0.0, seed=1, init_stddev=0.05, device="cuda:0")
    B, n = 128, 40
    rng = np.random.RandomState(5)
    xs = [np.full((B, 784), float(i), dtype=np.float32) + rng.rand(B, 784).astype(np.float32) * 0.25 for i in range(n)]
    ys = [np.eye(10, dtype=np.float32)[np.full(B, i % 10)] for i in range(n)]
    ws = eng.workspace(B)
    seen = []
    for i in range(n):
        eng.train_step(xs[i], ys[i], 0.75, fetch_loss=False)
        seen.append((ws.x.clone(), ws.y.clone()))  # stream-ordered snapshot of what step i actually consumed
    torch.cuda.synchronize()
    for i, (x, y) in enumerate(seen):
        assert torch.equal(x.cpu(), torch.from_numpy(xs[i])), "step %d trained on another batch" % i
        assert torch.equal(y.cpu(), torch.from_numpy(ys[i]))


def test_phase_stamps_of_the_last_step_are_ordered():
    """The Adam and conv1-forward kernels leave globaltimer stamps of their phases (diagnostics for the multi-GPU
    protocol: how long a step waits for arrivals / for the parameter version); on one GPU they must be monotone and the
    Adam kernel must come after the conv1 forward of the same step."""
    from tensorflow_distributed_b200.ops import native as nat
    eng = _engine(cnn_spec(), seed=0, init_stddev=0.05, device="cuda:0")
    x, y = _batch(128)
    for _ in range(3):
        eng.train_step(x.numpy(), y.numpy(), 0.75)
    adam, conv1 = nat.debug_stamps()
    assert 0 < conv1[0] <= conv1[2] < adam[0] <= adam[1] <= adam[2] <= adam[3]
    assert adam[4] >= adam[2] and (adam[4] - adam[0]) < 5_000_000  # an Adam kernel of ~10 us, not milliseconds
