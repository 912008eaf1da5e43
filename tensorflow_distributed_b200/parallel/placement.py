"""Variable placement onto parameter-server shards.

The reference places whole variables on ps tasks round-robin in creation order
via ``tf.train.replica_device_setter`` (``/root/reference/mnist_python_m.py:
177``).  That puts 98% of the CNN's bytes (``wd1``) on one ps.
This module keeps whole-variable round-robin for small tensors and splits big
tensors *inside* the tensor by rows (the GEMM K dimension of the forward pass),
in units of ``row_align`` rows (128 = one tcgen05 M tile = two K blocks) so every shard boundary is a tile
boundary of the pull-fused (forward/dgrad) and push-fused (wgrad) GEMMs.  Each shard owns one flat fp32 arena (master weights, Adam
m/v alongside) -- one allocation per ps GPU.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

from ..models.spec import ModelSpec

ARENA_ALIGN = 64  # elements; keeps every piece 256-byte aligned in the fp32 arena


@dataclass(frozen=True)
class Piece:
    var: str
    shard: int
    row_start: int
    row_end: int
    cols: int
    offset: int  # element offset inside the shard's arena

    @property
    def numel(self) -> int:
        return (self.row_end - self.row_start) * self.cols


def _align(n: int, a: int = ARENA_ALIGN) -> int:
    return (n + a - 1) // a * a


class Placement:
    def __init__(self, spec: ModelSpec, num_shards: int, split_threshold: int = 1 << 18, row_align: int = 128):
        if num_shards < 1:
            raise ValueError("need at least one ps shard")
        self.spec = spec
        self.num_shards = num_shards
        self.pieces: Dict[str, List[Piece]] = {}
        sizes = [0] * num_shards
        rr = 1  # global_step takes slot 0 of the round robin (created first, m:178)
        self.global_step_shard = 0
        for v in spec.variables:
            rows, cols = v.rows, v.cols
            if len(v.shape) == 1:
                rows, cols = v.shape[0], 1
            if num_shards > 1 and v.numel >= split_threshold and rows >= num_shards * row_align:
                blocks = (rows + row_align - 1) // row_align
                base, extra = divmod(blocks, num_shards)
                plist, r0 = [], 0
                for s in range(num_shards):
                    nb = base + (1 if s < extra else 0)
                    r1 = min(rows, r0 + nb * row_align)
                    plist.append(Piece(v.name, s, r0, r1, cols, sizes[s]))
                    sizes[s] += _align((r1 - r0) * cols)
                    r0 = r1
                assert r0 == rows
                self.pieces[v.name] = plist
            else:
                s = rr % num_shards
                rr += 1
                self.pieces[v.name] = [Piece(v.name, s, 0, rows, cols, sizes[s])]
                sizes[s] += _align(rows * cols)
        self.shard_sizes = [max(_align(s), ARENA_ALIGN) for s in sizes]

    def shard_pieces(self, shard: int) -> List[Piece]:
        return [p for v in self.spec.variables for p in self.pieces[v.name] if p.shard == shard]

    def owner_rows(self, var: str) -> List[Tuple[int, int, int]]:
        return [(p.shard, p.row_start, p.row_end) for p in self.pieces[var]]

    def table(self) -> str:
        """Human-readable placement table (what log_device_placement would show)."""
        lines = ["variable            shape               shard  rows            arena_offset  elems"]
        lines.append("global_step         ()                  %-5d  -               -             1" % self.global_step_shard)
        for v in self.spec.variables:
            for p in self.pieces[v.name]:
                lines.append("%-19s %-19s %-5d  %-15s %-13d %d" % (
                    v.name, "x".join(map(str, v.shape)), p.shard,
                    "[%d,%d)" % (p.row_start, p.row_end), p.offset, p.numel))
        for s, n in enumerate(self.shard_sizes):
            lines.append("shard %d: %d elements (%.2f MB fp32)" % (s, n, n * 4 / 1e6))
        return "\n".join(lines)

    # ---- flat <-> dict helpers (CPU path and checkpoint gather) ----
    def scatter(self, params, arenas: Sequence) -> None:
        """Copy a name->tensor dict into per-shard flat arenas."""
        for v in self.spec.variables:
            t = params[v.name].reshape(-1, self.pieces[v.name][0].cols)
            for p in self.pieces[v.name]:
                arenas[p.shard][p.offset:p.offset + p.numel].copy_(t[p.row_start:p.row_end].reshape(-1))

    def gather(self, arenas: Sequence) -> dict:
        import torch

        out = {}
        for v in self.spec.variables:
            parts = [arenas[p.shard][p.offset:p.offset + p.numel].reshape(p.row_end - p.row_start, p.cols)
                     for p in self.pieces[v.name]]
            out[v.name] = torch.cat([x.cpu() for x in parts], dim=0).reshape(v.shape).clone()
        return out


def tf_round_robin(spec: ModelSpec, num_ps: int) -> Dict[str, int]:
    """What replica_device_setter would do: whole variables, creation order, global_step first."""
    out = {"global_step": 0}
    for i, v in enumerate(spec.variables):
        out[v.name] = (i + 1) % num_ps
    return out
