"""Host-side arithmetic of the experimental NVLS transport (parallel/ps_nvls.py): the symmetric heap is laid out
identically on every rank, and the shards' piece tables tile the single-shard mirror / gradient arenas exactly once."""
import pytest

from tensorflow_distributed_b200.engine_native import ShardState, needs_shadow
from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.ops import native as nat
from tensorflow_distributed_b200.parallel.ps_nvls import NvlsLayout

pytestmark = pytest.mark.skipif(not nat.available(), reason="needs the built native library (ctypes structs)")


@pytest.mark.parametrize("spec", [cnn_spec(), mlp_spec(512, 3), mlp_spec(8192, 4)], ids=["cnn", "mlp512x3", "mlp8192x4"])
@pytest.mark.parametrize("shards", [1, 2, 3, 8])
def test_piece_tables_tile_the_symmetric_arenas_exactly_once(spec, shards):
    L = NvlsLayout(spec, shards)
    regions = [L.ctrl, L.octrl, L.mirror_w, L.mirror_shadow, L.grad, L.own_w, L.own_m, L.own_v, L.total]
    assert regions == sorted(regions) and all(r % 1024 == 0 for r in regions)
    assert L.grad - L.mirror_shadow >= L.shadow_n1 * 2 and L.total - L.own_v >= L.nmax * 4
    covered_flat, covered_shadow = {}, {}
    for s in range(shards):
        owner_spans = []
        for pc in L.pieces[s]:
            assert pc.owner_off % 4 == 0 and pc.grad_off % 4 == 0 and pc.mirror_off == pc.grad_off
            assert pc.owner_off + pc.numel <= L.nmax and pc.grad_off + pc.numel <= L.n1
            owner_spans.append((pc.owner_off, pc.owner_off + pc.numel))
            covered_flat[pc.grad_off] = pc.numel
            if pc.ld_shadow:
                assert pc.ld_shadow == pc.cols and pc.shadow_off % 4 == 0
                assert pc.shadow_off + pc.numel <= L.shadow_n1
                covered_shadow[pc.shadow_off] = pc.numel
        owner_spans.sort()
        assert all(a[1] <= b[0] for a, b in zip(owner_spans, owner_spans[1:]))
    # every variable's elements appear exactly once across the shards, contiguously in the single-shard layout
    for v in spec.variables:
        (q,) = L.placement1.pieces[v.name]
        mine = sorted(o for o in covered_flat if q.offset <= o < q.offset + v.numel)
        assert mine and mine[0] == q.offset
        for a, b in zip(mine, mine[1:]):
            assert a + covered_flat[a] == b
        assert sum(covered_flat.pop(o) for o in mine) == v.numel
        if needs_shadow(spec, v.name):
            s0 = L.shadow_off1[v.name]
            mine = sorted(o for o in covered_shadow if s0 <= o < s0 + v.numel)
            assert mine and mine[0] == s0
            for a, b in zip(mine, mine[1:]):
                assert a + covered_shadow[a] == b
            assert sum(covered_shadow.pop(o) for o in mine) == v.numel
    assert not covered_flat and not covered_shadow


def test_mirror_layout_is_the_single_gpu_engines_layout():
    spec = cnn_spec()
    L = NvlsLayout(spec, 4)
    pieces, shadow_off, shadow_ld, shadow_n = ShardState.layout(spec, L.placement1, 0)
    assert (shadow_off, shadow_ld, shadow_n) == (L.shadow_off1, L.shadow_ld1, L.shadow_n1)
    assert L.n1 == L.placement1.shard_sizes[0]


def test_ragged_tensor_core_operands_are_rejected_up_front():
    with pytest.raises(ValueError, match="multiple of 8"):
        NvlsLayout(mlp_spec(100, 1), 2)  # 100 columns: the bf16 rows would be padded, multimem.st moves >= 32 bits
