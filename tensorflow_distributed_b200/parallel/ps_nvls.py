"""Parameter server over NVLS multicast -- EXPERIMENTAL transport (``TFD_TRANSPORT=nvls``).

Status: the layout arithmetic below is unit-tested on the CPU (``tests/test_nvls_layout.py``) and the kernels
compile for sm_100a (``csrc/nvls_kernels.cu``), but the engine has not been run on hardware yet (the round's GPU
budget was spent); the default transport stays ``parallel/ps_gpu.py``.  It is the next step named in DESIGN.md
("What the measurements say to build next", items 3 and 4).

Idea.  ``ps_gpu`` keeps the reference's star topology (``/root/reference/mnist_python_m.py:177``: every parameter
and gradient byte crosses the ps task's link once per worker).  With one ps GPU and seven workers that is 7 x 13 MB
in and 7 x 6.5 MB out of one NVLink port per step, which is what bounds BASELINE config 3.  Here every rank maps one
*symmetric* heap (``torch.distributed._symmetric_memory``) at a unicast and at a multicast address:

* workers run exactly the single-GPU kernel sequence on LOCAL memory: parameters come from a local mirror,
  gradients go to a local arena (no pull kernels, no peer TMA loads, no remote epilogue stores);
* after the backward pass a worker adds 1 to every shard's arrival counter with one ``multimem.red``;
* the owner shard reads the gradient SUM with ``multimem.ld_reduce`` (the switch fetches and adds the workers'
  copies: ingress 1x), applies TF-form Adam to its private master weights / slots, and stores the new values
  into every mirror with ``multimem.st`` (egress 1x); the version counter of every GPU is bumped the same way;
* the next step's first kernel waits on the LOCAL version word.

The owner shard alone still holds master weights and Adam slots and applies the update once per global step from
the average of all workers' gradients: the semantics of the reference's synchronous replicas (m:210-233).
Sync mode with R = N only; async and backup workers stay on ``ps_gpu``.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..engine_native import NativeEngine, ShardState, Workspace, needs_shadow
from ..models.spec import ModelSpec
from ..ops import native as nat
from .placement import Placement

# words of the symmetric control block (every rank has one; workers and shards use their own copy)
W_VERSION, W_ARRIVE, W_STEP = 2, 4, 8
# words of the owner control block (second kilobyte of the heap; meaningful on ps ranks)
O_T, O_BLOCKS, O_GSTEP = 0, 1, 3


def _al(n: int, a: int = 1024) -> int:
    return (n + a - 1) // a * a


class NvlsLayout:
    """Byte offsets of the symmetric heap (identical on every rank) and the per-shard piece tables.

    ``[ctrl | owner ctrl | mirror fp32 | mirror bf16 shadow | gradient | owner w | owner m | owner v]``: the mirror,
    shadow and gradient arenas use the single-shard layout of ``NativeEngine`` (so a worker is a NativeEngine on
    local memory); the owner arenas use the ``num_shards`` placement and are sized for the largest shard.
    """

    def __init__(self, spec: ModelSpec, num_shards: int, precision: str = "bf16"):
        self.spec, self.num_shards, self.precision = spec, num_shards, precision
        self.placement1 = Placement(spec, 1)
        self.placement = Placement(spec, num_shards)
        _, self.shadow_off1, self.shadow_ld1, self.shadow_n1 = ShardState.layout(spec, self.placement1, 0, precision)
        self.n1 = self.placement1.shard_sizes[0]
        self.nmax = max(self.placement.shard_sizes)
        off = 0
        self.ctrl = off; off += 1024
        self.octrl = off; off += 1024
        self.mirror_w = off; off += _al(self.n1 * 4)
        self.mirror_shadow = off; off += _al(self.shadow_n1 * 2)
        self.grad = off; off += _al(self.n1 * 4)
        self.own_w = off; off += _al(self.nmax * 4)
        self.own_m = off; off += _al(self.nmax * 4)
        self.own_v = off; off += _al(self.nmax * 4)
        self.total = off
        self.pieces: List[List[nat.NvlsPiece]] = [self._pieces_of(s) for s in range(num_shards)]

    def _pieces_of(self, shard: int) -> List[nat.NvlsPiece]:
        out = []
        for p in self.placement.shard_pieces(shard):
            (q,) = self.placement1.pieces[p.var]  # where the whole variable sits in the single-shard arenas
            flat = q.offset + p.row_start * p.cols
            ld, soff = 0, 0
            if needs_shadow(self.spec, p.var, self.precision):
                ld = self.shadow_ld1[p.var]
                if ld != p.cols:
                    raise ValueError("NVLS transport: %s has %d columns; bf16-published tensors need a multiple of 8"
                                     % (p.var, p.cols))
                soff = self.shadow_off1[p.var] + p.row_start * ld
            out.append(nat.NvlsPiece(p.offset, p.numel, p.cols, ld, flat, flat, soff))
        return out


class NvlsClusterEngine(NativeEngine):
    """One rank of the ps/worker cluster on the NVLS transport.  Same public API as NativeEngine / GpuClusterEngine."""

    name = "native-sm100a-nvls"
    _prio_default = "1"

    def __init__(self, spec: ModelSpec, topo, rank: int, device, learning_rate: float = 0.01, seed: int = 0,
                 init_stddev: float = 1.0, use_graphs: bool = True, conv2_ref: bool = False,
                 precision: Optional[str] = None):
        if not dist.is_initialized():
            raise RuntimeError("NvlsClusterEngine needs torch.distributed (bootstrap only) to be initialised")
        nat.lib()
        import torch.distributed._symmetric_memory as symm
        self.spec, self.topo, self.rank = spec, topo, rank
        self._set_precision(precision, conv2_ref)  # tf32: fp32 mirrors only; bf16: fp32 mirrors + bf16 shadow mirrors
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        nat.set_device(self.device.index or 0)
        nat.set_slot_dtype(False)
        self.lr, self.seed = learning_rate, seed
        self.dropout_seed = seed + 31 * (topo.worker_of(rank) or 0)  # per-worker dropout stream, shared init seed
        self.use_graphs, self.conv2_ref = use_graphs, conv2_ref
        self.ws: Dict[int, Workspace] = {}
        self.my_shard, self.my_worker = topo.shard_of(rank), topo.worker_of(rank)
        self.is_ps, self.is_worker = self.my_shard is not None, self.my_worker is not None
        self.num_workers, self.num_shards = topo.num_workers, topo.num_shards
        self.async_mode, self.backup = False, False
        L = self.layout = NvlsLayout(spec, topo.num_shards, self.precision)
        # ---- symmetric heap: unicast view of every rank + one multicast alias ----
        self._buf = symm.empty(L.total, dtype=torch.uint8, device=self.device)
        self._buf.zero_()
        torch.cuda.synchronize(self.device)
        self._symm = symm.rendezvous(self._buf, dist.group.WORLD)
        self.heap_ptr = self._buf.data_ptr()
        self.peer_base = {r: int(self._symm.buffer_ptrs[r]) for r in range(topo.world)}
        self.mc_base = int(self._symm.multicast_ptr) if getattr(self._symm, "has_multicast_support", False) else 0
        if not self.mc_base:
            raise RuntimeError("NVLS transport: this box / driver exposes no multicast address for symmetric memory")
        dev = self.device
        here = lambda off, n, dt: nat.tensor_from_ptr(self.heap_ptr + off, n, dt, dev)  # noqa: E731
        # ---- worker side: a NativeEngine whose parameter arena, shadow and gradient arena live in the heap ----
        offs = {"w": L.mirror_w, "shadow": L.mirror_shadow, "ctrl": L.ctrl}

        def alloc(name, numel, dtype):
            if name in offs:
                return here(offs[name], numel, dtype)
            return torch.zeros(64, dtype=dtype, device=dev)  # Adam slots exist on the owner shards only

        self.placement = L.placement1
        self.shard = ShardState(spec, L.placement1, 0, dev, alloc=alloc, precision=self.precision)
        self.grad = here(L.grad, L.n1, torch.float32)
        self.ctrl = self.shard.ctrl
        # ---- owner side ----
        self.octrl = here(L.octrl, 64, torch.int32)
        if self.is_ps:
            self.own_w, self.own_m, self.own_v = (here(o, L.nmax, torch.float32) for o in (L.own_w, L.own_m, L.own_v))
            for i, v in enumerate(spec.variables):  # tf.global_variables_initializer() for the rows this shard owns
                for p in L.placement.pieces[v.name]:
                    if p.shard == self.my_shard:
                        nat.random_normal(self.own_w[p.offset:p.offset + p.numel], (seed << 8) + i, init_stddev,
                                          p.row_start * p.cols)
        torch.cuda.synchronize(dev)
        dist.barrier()
        self._publish()

    # ---- protocol pieces --------------------------------------------------------------------
    def _mc(self, off: int) -> int:
        return self.mc_base + off

    def _launch_apply(self, adam: bool) -> None:
        L = self.layout
        nat.nvls_apply(self.own_w, self.own_m, self.own_v, self._mc(L.grad), self._mc(L.mirror_w), self._mc(L.mirror_shadow),
                       L.pieces[self.my_shard], 1.0 / self.num_workers, self.lr, self.octrl[O_T:], self.ctrl[W_ARRIVE:],
                       self.num_workers, self._mc(L.ctrl + 4 * W_VERSION), self.octrl[O_GSTEP:], adam=adam)

    def _publish(self) -> None:
        """Owners store their current parameters into every mirror (initialisation / restore); collective."""
        if self.is_ps:
            self._launch_apply(adam=False)
        torch.cuda.synchronize(self.device)
        dist.barrier()

    def step_ptr(self):
        return self.ctrl[W_STEP:]

    def _pre_step(self, ws, training: bool) -> None:
        # every shard adds 1 to the version per applied step: step k may start once it reads num_shards * k.  The CNN's
        # first kernel does the waiting itself (conv1_fwd below); the MLP has a one-block gate kernel.
        if self.spec.kind != "cnn":
            nat.nvls_gate(self.ctrl[W_VERSION:], self.step_ptr(), self.num_shards, 0)

    def conv1_fwd(self, ws, training: bool = True) -> None:
        nat.conv1_fwd(ws.x, self.P("weights/wc1"), self.P("biases/bc1"), ws.pool1, ws.am1, ws.B, tf32=self.tf32,
                      gate_flags=[self.ctrl[W_VERSION:]], want_ptr=self.step_ptr(), want_mul=self.num_shards)

    def _post_backward(self, ws) -> None:
        if self._training_apply:  # "grads" mode computes gradients without taking part in an update
            nat.nvls_signal(self._mc(self.layout.ctrl + 4 * W_ARRIVE), self.step_ptr())

    def _apply(self) -> None:
        if self.is_ps:
            self._launch_apply(adam=True)

    def _sequence(self, ws, mode: str, keep_prob: float, mask_out=None) -> None:
        if self.is_worker:
            return super()._sequence(ws, mode, keep_prob, mask_out)
        if mode == "train":  # dedicated ps GPU: nothing but the apply, like the reference's passive ps (m:160-161)
            self._training_apply = True
            self._apply()

    # ---- public API -------------------------------------------------------------------------
    def workspace(self, B: int) -> Workspace:
        if B not in self.ws:
            self.ws[B] = Workspace(self.spec, B if self.is_worker else 1, self.device, self.precision)
            self.ws[B].B = B if self.is_worker else 1
        return self.ws[B]

    def load_batch(self, ws, x, y) -> None:
        if self.is_worker:
            super().load_batch(ws, x, y)

    def describe(self) -> str:
        return self.topo.describe() + " " + self.precision + " sync nvls"

    def _owner_word(self, shard: int, word: int) -> torch.Tensor:
        r = self.topo.ps_ranks[shard]
        return nat.tensor_from_ptr(self.peer_base[r] + self.layout.octrl + 4 * word, 1, torch.int32, self.device)

    @property
    def global_step(self) -> int:
        return int(self._owner_word(0, O_GSTEP).item())

    def _read_owner(self, which: str) -> Dict[str, torch.Tensor]:
        """name -> full tensor, read from the owners' private arenas through the unicast peer mapping."""
        torch.cuda.synchronize(self.device)
        L, out = self.layout, {}
        for v in self.spec.variables:
            parts = []
            for p in sorted(L.placement.pieces[v.name], key=lambda q: q.row_start):
                base = self.peer_base[self.topo.ps_ranks[p.shard]] + getattr(L, which) + 4 * p.offset
                parts.append(nat.tensor_from_ptr(base, p.numel, torch.float32, self.device).cpu()
                             .view(p.row_end - p.row_start, p.cols))
            out[v.name] = torch.cat(parts, 0).reshape(v.shape).clone()
        return out

    def get_params(self) -> Dict[str, torch.Tensor]:
        return self._read_owner("own_w")

    def get_adam_state(self) -> dict:
        t = int(self._owner_word(0, O_T).item())
        return {"t": t, "beta1_power": 0.9 ** t, "beta2_power": 0.999 ** t, "m": self._read_owner("own_m"),
                "v": self._read_owner("own_v")}

    def get_grads(self) -> Dict[str, torch.Tensor]:
        """This worker's last gradient (its local arena)."""
        return super().get_grads()

    def _copy_owned(self, arena: torch.Tensor, full: Dict[str, torch.Tensor]) -> None:
        for v in self.spec.variables:
            for p in self.layout.placement.pieces[v.name]:
                if p.shard == self.my_shard:
                    src = full[v.name].to(torch.float32).reshape(-1, p.cols)[p.row_start:p.row_end]
                    arena[p.offset:p.offset + p.numel].copy_(src.reshape(-1).to(self.device))

    def set_params(self, params: Dict[str, torch.Tensor]) -> None:
        """Restore: every owner copies its rows, then republishes them into all mirrors (call on all ranks)."""
        if self.is_ps:
            self._copy_owned(self.own_w, params)
        self._publish()

    def set_adam_state(self, sd: dict) -> None:
        T = int(sd["t"])
        if self.is_ps:
            self._copy_owned(self.own_m, sd["m"])
            self._copy_owned(self.own_v, sd["v"])
            self.octrl[O_T] = T
        # every protocol counter restarts at the restored update index T
        self.ctrl[W_VERSION] = T * self.num_shards
        self.ctrl[W_ARRIVE] = T * self.num_workers
        self.ctrl[W_STEP] = T
        torch.cuda.synchronize(self.device)
        dist.barrier()

    def set_global_step(self, step: int) -> None:
        if self.is_ps:
            self.octrl[O_GSTEP] = int(step)
        torch.cuda.synchronize(self.device)

    def shutdown(self) -> None:
        torch.cuda.synchronize(self.device)
        dist.barrier()

    @classmethod
    def from_bench_args(cls, spec, a, rank: int, world: int, device) -> "NvlsClusterEngine":
        from .ps_gpu import Topology
        return cls(spec, Topology.make(world, a.layout, a.ps), rank, device, learning_rate=0.01, seed=0,
                   init_stddev=a.init_stddev, precision=getattr(a, "precision", None))
