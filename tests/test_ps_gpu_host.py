"""Host-side logic of the GPU parameter server that needs no GPU: topology, heap layout, shard layout."""
import pytest

from tensorflow_distributed_b200.engine_native import ShardState, needs_shadow
from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.parallel.placement import Placement
from tensorflow_distributed_b200.parallel.ps_gpu import HeapLayout, Topology


def test_topology_layouts():
    t = Topology.make(8, "colocated")
    assert t.ps_ranks == list(range(8)) and t.worker_ranks == list(range(8)) and t.describe() == "ps8+w8 colocated"
    t = Topology.make(8, "dedicated", 1)  # BASELINE config 3
    assert t.ps_ranks == [0] and t.worker_ranks == list(range(1, 8)) and t.num_workers == 7
    assert t.shard_of(0) == 0 and t.shard_of(3) is None and t.worker_of(3) == 2 and t.worker_of(0) is None
    t = Topology.make(8, "dedicated", 2)  # BASELINE config 4
    assert t.num_shards == 2 and t.num_workers == 6
    with pytest.raises(ValueError):
        Topology.make(2, "dedicated", 2)
    with pytest.raises(ValueError):
        Topology.make(4, "ring")


def test_shadow_only_for_tensor_core_operands():
    spec = cnn_spec()
    assert [v.name for v in spec.variables if needs_shadow(spec, v.name)] == ["weights/wc2", "weights/wd1"]
    m = mlp_spec(100, 2)
    assert [v.name for v in m.variables if needs_shadow(m, v.name)] == ["weights/h1", "weights/h2"]


@pytest.mark.parametrize("S,W,gb", [(1, 1, 4), (2, 6, 4), (8, 8, 2)])
def test_heap_layout_is_disjoint_aligned_and_identical_on_every_rank(S, W, gb):
    spec = cnn_spec()
    pl = Placement(spec, S)
    for s in range(S):
        a = HeapLayout(spec, pl, s, W, gb)
        b = HeapLayout(spec, pl, s, W, gb)
        assert vars(a) == vars(b)
        regions = [(a.ctrl, 1024), (a.w, a.n * 4), (a.m, a.n * 4), (a.v, a.n * 4), (a.shadow, a.shadow_n * 2)]
        regions += [(a.slot(w), a.n * gb) for w in range(W)]
        regions.sort()
        for (o0, n0), (o1, _) in zip(regions, regions[1:]):
            assert o0 % 1024 == 0 and o0 + n0 <= o1
        assert regions[-1][0] + regions[-1][1] <= a.total
    none = HeapLayout(spec, pl, None, W, gb)  # worker-only rank: control words only
    assert none.n == 0 and none.total >= 1024


def test_shard_layout_pads_shadow_rows_to_16_bytes():
    spec = mlp_spec(100, 1)
    pl = Placement(spec, 1)
    pieces, off, ld, n = ShardState.layout(spec, pl, 0)
    assert ld["weights/h1"] == 104 and off["weights/h1"] == 0 and n >= 784 * 104
    by = {p.arena_off: p for p in pieces}
    assert any(p.ld_shadow == 104 and p.cols == 100 for p in pieces)
    assert sum(1 for p in pieces if p.ld_shadow == 0) == 3  # out weights and both biases have no shadow
    cnn = cnn_spec()
    pl8 = Placement(cnn, 8)
    rows = [(p.row_start, p.row_end) for p in pl8.pieces["weights/wd1"]]
    assert all(r0 % 128 == 0 for r0, _ in rows) and rows[-1][1] == 3136


def test_copy_engine_paths_are_chosen_only_where_they_were_measured_faster(monkeypatch):
    """Prefetch / push through cudaMemcpyAsync nodes: faster with one remote piece per matrix (2 colocated GPUs:
    0.153 -> 0.1324 ms/step), slower with seven (8 GPUs: 0.2106 against 0.1777 ms) -- `auto` must pick accordingly, and
    the environment can force either."""
    from types import SimpleNamespace
    from tensorflow_distributed_b200.parallel.ps_gpu import GpuClusterEngine
    use = GpuClusterEngine._use_ce
    colocated = SimpleNamespace(v2=True, is_ps=True, is_worker=True)
    for env in ("TFD_PREFETCH", "TFD_PUSH"):
        monkeypatch.delenv(env, raising=False)
        assert use(colocated, env, 1) and not use(colocated, env, 3) and not use(colocated, env, 7)
        assert not use(SimpleNamespace(v2=True, is_ps=False, is_worker=True), env, 1)   # dedicated layout: not measured
        assert not use(SimpleNamespace(v2=False, is_ps=True, is_worker=True), env, 1)  # async / R<N / bf16 slots
        monkeypatch.setenv(env, "ce")
        assert use(colocated, env, 7)
        monkeypatch.setenv(env, "kernel")
        assert not use(colocated, env, 1)
