"""Variable placement (C9): TF round robin vs intra-tensor sharding."""
import torch

from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.ops.oracle import init_params
from tensorflow_distributed_b200.parallel.placement import Placement, tf_round_robin


def test_param_counts():
    assert cnn_spec().num_params == 3274634
    assert mlp_spec(8192, 4).num_params == 207863818
    assert [v.tf_auto_name for v in cnn_spec().variables][:3] == ["Variable", "Variable_1", "Variable_2"]


def test_tf_round_robin_puts_wd1_on_one_ps():
    rr = tf_round_robin(cnn_spec(), 2)
    assert rr["global_step"] == 0 and rr["weights/wc1"] == 1 and rr["weights/wd1"] == 1


def test_sharded_placement_splits_wd1_on_k_blocks():
    for S in (1, 2, 8):
        pl = Placement(cnn_spec(), S)
        rows = pl.owner_rows("weights/wd1")
        assert len(rows) == S
        assert rows[0][1] == 0 and rows[-1][2] == 3136
        for (s, r0, r1) in rows:
            assert r0 % 128 == 0
        if S > 1:
            sizes = [r1 - r0 for _, r0, r1 in rows]
            assert max(sizes) - min(sizes) < 2 * 128  # one 128-row block, plus a ragged last block
        for p in pl.shard_pieces(0):
            assert p.offset % 64 == 0
        assert "weights/wd1" in pl.table()


def test_scatter_gather_roundtrip():
    spec = cnn_spec()
    params = init_params(spec, 3)
    pl = Placement(spec, 3)
    arenas = [torch.zeros(n) for n in pl.shard_sizes]
    pl.scatter(params, arenas)
    back = pl.gather(arenas)
    for k in params:
        assert torch.equal(back[k], params[k])
