"""Property tests (hypothesis) of the host-side logic every transport relies on: placement covers every parameter
element exactly once on tile-aligned shard boundaries, and the sync / async replica protocols of
/root/reference/mnist_python_m.py:210-233 keep their invariants under arbitrary worker interleavings."""
import torch
from hypothesis import given, settings, strategies as st

from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.parallel.placement import ARENA_ALIGN, Placement
from tensorflow_distributed_b200.parallel.replica import AsyncReplicas, SyncReplicas

SPECS = st.one_of(st.just(cnn_spec()),
                  st.builds(mlp_spec, st.sampled_from([8, 100, 512, 1024, 4096]), st.integers(1, 4)))


@settings(max_examples=60, deadline=None)
@given(spec=SPECS, shards=st.integers(1, 8))
def test_placement_partitions_every_variable_on_aligned_boundaries(spec, shards):
    pl = Placement(spec, shards)
    used = [[] for _ in range(shards)]
    for v in spec.variables:
        pieces = sorted(pl.pieces[v.name], key=lambda p: p.row_start)
        rows = v.shape[0] if len(v.shape) == 1 else v.rows  # filters are viewed as [kh*kw*cin, cout]
        assert pieces[0].row_start == 0 and pieces[-1].row_end == rows
        for a, b in zip(pieces, pieces[1:]):
            assert a.row_end == b.row_start and a.row_end % 128 == 0   # contiguous, split on tile rows only
        assert sum(p.numel for p in pieces) == v.numel
        assert len({p.shard for p in pieces}) == len(pieces)            # at most one piece per shard
        for p in pieces:
            assert p.offset % ARENA_ALIGN == 0 and p.numel > 0
            used[p.shard].append((p.offset, p.offset + p.numel))
    for s in range(shards):
        spans = sorted(used[s])
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 <= b0                                             # pieces never overlap inside an arena
        assert not spans or spans[-1][1] <= pl.shard_sizes[s]


@settings(max_examples=25, deadline=None)
@given(hidden=st.sampled_from([8, 64, 512]), layers=st.integers(1, 3), shards=st.integers(1, 5), seed=st.integers(0, 99))
def test_scatter_then_gather_is_the_identity(hidden, layers, shards, seed):
    spec = mlp_spec(hidden, layers)
    pl = Placement(spec, shards)
    g = torch.Generator().manual_seed(seed)
    params = {v.name: torch.randn(*v.shape, generator=g) for v in spec.variables}
    arenas = [torch.zeros(n) for n in pl.shard_sizes]
    pl.scatter(params, arenas)
    back = pl.gather(arenas)
    assert all(torch.equal(back[k], params[k]) for k in params)


@settings(max_examples=80, deadline=None)
@given(N=st.integers(1, 6), data=st.data())
def test_sync_replicas_apply_once_per_R_fresh_gradients_whatever_the_interleaving(N, data):
    R = data.draw(st.integers(1, N))
    applied = []
    sr = SyncReplicas(["w"], R, N, lambda g: applied.append(g["w"].clone()))
    sr.init_tokens(N)  # enough tokens for every worker to start (m:233)
    local = [sr.dequeue_token() for _ in range(N)]
    accepted_since_apply, value_sum = 0, 0.0
    for _ in range(data.draw(st.integers(1, 40))):
        w = data.draw(st.integers(0, N - 1))
        val = float(data.draw(st.integers(-5, 5)))
        ok = sr.push({"w": torch.tensor([val])}, local[w])
        assert ok == (1 if local[w] >= sr.global_step else 0)           # stale gradients are dropped, fresh kept
        if ok:
            accepted_since_apply += 1
            value_sum += val
        before = sr.global_step
        if sr.try_aggregate():
            assert accepted_since_apply >= R and sr.global_step == before + 1
            assert abs(applied[-1].item() - value_sum / accepted_since_apply) < 1e-6  # the average of what arrived
            accepted_since_apply, value_sum = 0, 0.0
        else:
            assert accepted_since_apply < R
        if sr.token_available():
            local[w] = sr.dequeue_token()                               # the worker adopts the token as local step
            assert local[w] <= sr.global_step
    assert sr.applied == len(applied) == sr.global_step


@settings(max_examples=50, deadline=None)
@given(workers=st.lists(st.integers(0, 7), min_size=1, max_size=60))
def test_async_replicas_apply_every_gradient_exactly_once_in_ticket_order(workers):
    total = []
    ar = AsyncReplicas(lambda g: total.append(g["w"].item()))
    tickets = [(ar.take_ticket(), w) for w in workers]
    for t, w in tickets:
        assert ar.push(w, {"w": torch.tensor([float(t)])}, ticket=t) == t + 1
    assert total == [float(t) for t, _ in tickets] and ar.order == workers and ar.global_step == len(workers)
