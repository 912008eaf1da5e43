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


# ---------------------------------------------------------------------------------------------------------------
# Protocol simulation on the CPU: the heaps are byte tensors, multimem.ld_reduce is a sum over the ranks' copies and
# multimem.st a store into every copy.  What is checked is everything the kernels take from NvlsLayout -- the piece
# tables' owner / gradient / mirror / shadow offsets and the counters -- against the oracle's large-batch update.
# ---------------------------------------------------------------------------------------------------------------
import torch  # noqa: E402

from tensorflow_distributed_b200.ops import oracle  # noqa: E402
from tensorflow_distributed_b200.parallel.ps_nvls import W_ARRIVE, W_STEP, W_VERSION  # noqa: E402


class _SimRank:
    def __init__(self, L):
        self.L = L
        self.heap = torch.zeros(L.total, dtype=torch.uint8)
        self.ctrl = self.heap[L.ctrl:L.ctrl + 256].view(torch.int32)

    def f32(self, off, n):
        return self.heap[off:off + 4 * n].view(torch.float32)

    def bf16(self, off, n):
        return self.heap[off:off + 2 * n].view(torch.bfloat16)

    def mirror_params(self, spec):
        """What the worker's kernels read: fp32 small tensors from the mirror, bf16 shadows upcast."""
        L, out = self.L, {}
        for v in spec.variables:
            (q,) = L.placement1.pieces[v.name]
            if needs_shadow(spec, v.name):
                ld = L.shadow_ld1[v.name]
                rows = v.numel // q.cols
                out[v.name] = self.bf16(L.mirror_shadow, L.shadow_n1)[L.shadow_off1[v.name]:L.shadow_off1[v.name] + rows * ld] \
                    .view(rows, ld)[:, :q.cols].float().reshape(v.shape)
            else:
                out[v.name] = self.f32(L.mirror_w, L.n1)[q.offset:q.offset + v.numel].reshape(v.shape).clone()
        return out

    def write_grads(self, spec, grads):
        g = self.f32(self.L.grad, self.L.n1)
        for v in spec.variables:
            (q,) = self.L.placement1.pieces[v.name]
            g[q.offset:q.offset + v.numel] = grads[v.name].reshape(-1)


def _apply(ranks, owner_rank, pieces, L, adam, t, scale, lr=0.01, b1=0.9, b2=0.999, eps=1e-8):
    own = ranks[owner_rank]
    w, m, v = own.f32(L.own_w, L.nmax), own.f32(L.own_m, L.nmax), own.f32(L.own_v, L.nmax)
    for pc in pieces:
        o, n = pc.owner_off, pc.numel
        if adam:
            g = sum(r.f32(L.grad, L.n1)[pc.grad_off:pc.grad_off + n] for r in ranks) * scale  # multimem.ld_reduce
            lr_t = lr * (1 - b2 ** t) ** 0.5 / (1 - b1 ** t)
            m[o:o + n] = b1 * m[o:o + n] + (1 - b1) * g
            v[o:o + n] = b2 * v[o:o + n] + (1 - b2) * g * g
            w[o:o + n] -= lr_t * m[o:o + n] / (v[o:o + n].sqrt() + eps)
        for r in ranks:  # multimem.st
            if pc.ld_shadow == 0:
                r.f32(L.mirror_w, L.n1)[pc.mirror_off:pc.mirror_off + n] = w[o:o + n]
            else:
                r.bf16(L.mirror_shadow, L.shadow_n1)[pc.shadow_off:pc.shadow_off + n] = w[o:o + n].to(torch.bfloat16)
    if adam:
        for r in ranks:  # multimem.red on the version word
            r.ctrl[W_VERSION] += 1


@pytest.mark.parametrize("kind,shards", [("mlp", 1), ("mlp", 3), ("cnn", 2), ("cnn", 4)])
def test_simulated_protocol_equals_the_large_batch_update(kind, shards):
    torch.manual_seed(0)
    spec = mlp_spec(1024, 2) if kind == "mlp" else cnn_spec()
    L = NvlsLayout(spec, shards)
    W = shards  # colocated: every rank is a worker and owns one shard
    ranks = [_SimRank(L) for _ in range(W)]
    p0 = oracle.init_params(spec, 1, 0.05)
    for s in range(shards):  # owners take their rows
        w = ranks[s].f32(L.own_w, L.nmax)
        for v in spec.variables:
            for p in L.placement.pieces[v.name]:
                if p.shard == s:
                    w[p.offset:p.offset + p.numel] = p0[v.name].reshape(-1, p.cols)[p.row_start:p.row_end].reshape(-1)
        _apply(ranks, s, L.pieces[s], L, adam=False, t=0, scale=0.0)
    want = {k: (v.to(torch.bfloat16).float() if needs_shadow(spec, k) else v) for k, v in p0.items()}
    for r in ranks:  # every mirror holds every parameter, tensor-core operands rounded to bf16
        got = r.mirror_params(spec)
        assert all(torch.equal(got[k], want[k]) for k in want)
    ref, adam = {k: v.clone() for k, v in p0.items()}, oracle.TFAdam(0.01)
    B = 4
    for step in range(1, 3):
        gsum = None
        for wi, r in enumerate(ranks):
            assert int(r.ctrl[W_VERSION]) >= shards * int(r.ctrl[W_STEP])        # the gate of this step is open
            x = torch.rand(B, 784)
            y = torch.eye(10)[torch.randint(0, 10, (B,))]
            _, g, _ = oracle.loss_and_grads(kind, x, y, r.mirror_params(spec))
            r.write_grads(spec, g)
            gsum = g if gsum is None else {k: gsum[k] + g[k] for k in g}
            for q in ranks:  # multimem.red on every arrival counter, then the local step word
                q.ctrl[W_ARRIVE] += 1
            r.ctrl[W_STEP] += 1
        for s in range(shards):
            assert int(ranks[s].ctrl[W_ARRIVE]) >= W * step                         # the apply's wait is satisfied
            _apply(ranks, s, L.pieces[s], L, adam=True, t=step, scale=1.0 / W)
        adam.apply(ref, {k: gsum[k] / W for k in gsum})
        # owners' master weights == one TF-Adam step on the average gradient; mirrors == their (rounded) copy
        for v in spec.variables:
            parts = [ranks[p.shard].f32(L.own_w, L.nmax)[p.offset:p.offset + p.numel].view(-1, p.cols)
                     for p in sorted(L.placement.pieces[v.name], key=lambda q: q.row_start)]
            master = torch.cat(parts, 0).reshape(v.shape)
            assert torch.allclose(master, ref[v.name], atol=1e-6, rtol=1e-5), v.name
            mir = ranks[-1].mirror_params(spec)[v.name]
            assert torch.equal(mir, master.to(torch.bfloat16).float() if needs_shadow(spec, v.name) else master), v.name
    assert all(int(r.ctrl[W_VERSION]) == 2 * shards and int(r.ctrl[W_STEP]) == 2 for r in ranks)
