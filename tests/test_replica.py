"""Sync / async replica state machines vs the TF semantics in SURVEY 2.6."""
import pytest
import torch

from tensorflow_distributed_b200.parallel.replica import AsyncReplicas, ConditionalAccumulator, SyncReplicas


def _mk(R, N):
    applied = []
    s = SyncReplicas(["a", "b"], R, N, lambda g: applied.append({k: v.clone() for k, v in g.items()}))
    return s, applied


def g(v):
    return {"a": torch.tensor([float(v)]), "b": torch.tensor([2.0 * v])}


def test_accumulator_stale_drop():
    a = ConditionalAccumulator("x")
    a.set_global_step(3)
    assert not a.apply_grad(torch.ones(1), 2) and a.dropped == 1 and a.count == 0
    assert a.apply_grad(torch.ones(1), 3) and a.apply_grad(3 * torch.ones(1), 4)
    assert a.ready(2) and torch.allclose(a.take_grad(2), torch.tensor([2.0]))
    assert a.count == 0


def test_sync_R_equals_N_average_and_tokens():
    s, applied = _mk(2, 2)
    s.init_tokens(0)
    s.push(g(1), 0)
    assert not s.try_aggregate() and not s.token_available()
    s.push(g(3), 0)
    assert s.try_aggregate()
    assert s.global_step == 1 and torch.allclose(applied[0]["a"], torch.tensor([2.0]))
    assert torch.allclose(applied[0]["b"], torch.tensor([4.0]))
    assert len(s.tokens) == 2 and s.dequeue_token() == 1 and s.dequeue_token() == 1


def test_default_token_preload_lets_workers_run_ahead():
    s, _ = _mk(2, 2)
    s.init_tokens()  # R tokens of step 0 (m:233, m:281)
    assert len(s.tokens) == 2 and s.dequeue_token() == 0


def test_backup_workers_R_less_than_N_drops_stale():
    s, applied = _mk(2, 3)
    s.init_tokens(1)
    s.push(g(1), 0)
    s.push(g(2), 0)
    assert s.try_aggregate() and s.global_step == 1
    # the straggler's step-0 gradient arrives late: dropped by every accumulator
    assert s.push(g(100), 0) == 0 and s.dropped == 2
    assert not s.try_aggregate()
    # averaging is over R (what arrived), and N tokens were released
    assert torch.allclose(applied[0]["a"], torch.tensor([1.5]))
    assert len(s.tokens) == 1 + 3
    with pytest.raises(ValueError):
        _mk(1, 3)[0].init_tokens(1)  # needs >= N-R tokens


def test_async_ordered_apply_counts_every_update():
    seen = []
    a = AsyncReplicas(lambda gr: seen.append(float(gr["a"])))
    t0, t1 = a.take_ticket(), a.take_ticket()
    with pytest.raises(RuntimeError):
        a.push(1, g(5), ticket=t1)  # out of order
    assert a.push(0, g(7), ticket=t0) == 1
    assert a.push(1, g(5), ticket=t1) == 2
    assert seen == [7.0, 5.0] and a.order == [0, 1] and a.global_step == 2
