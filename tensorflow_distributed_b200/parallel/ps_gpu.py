"""GPU parameter server: ps shards in HBM, reached by the workers' kernels over NVLink peer memory.

What TF's gRPC runtime does implicitly for the reference (parameter pull X1, gradient push X2/X3,
sync token X4, global_step X5, init/readiness X6; ``/root/reference/mnist_python_m.py:158-161,177,210-233``)
becomes explicit device-side traffic on one 8xB200 box:

* a **ps task is a GPU** that owns one shard: fp32 master arena, Adam m/v, a bf16 shadow of the tensor-core
  operands, one gradient staging slot per worker, and a few control words -- all in ONE cudaIpc allocation that
  every other rank maps (symmetric heap; ``torch.distributed`` only carries the 64-byte handles at start-up);
* **pull**: small tensors are copied by one kernel gated on each shard's ``version`` word; the big sharded
  matrices are never copied -- the forward/dgrad GEMMs TMA-load their B tiles straight from the owner GPUs;
* **push**: the wgrad GEMM epilogue stores its tiles straight into this worker's slot on the owner GPU; the small
  gradients follow in one kernel that then adds 1 to every shard's ``arrive`` word (release, system scope);
* **apply**: each shard's Adam kernel spins until ``arrive >= workers * t``, sums the slots in a fixed order,
  scales by 1/R, updates w/m/v, republishes the bf16 shadow and stores ``version = t`` (release).
  ``version``/``arrive`` are the token queue of SyncReplicasOptimizer reduced to two monotonic counters.

Layouts: ``colocated`` (default) -- every GPU is a worker and hosts a shard; ``dedicated`` -- the first S GPUs are
ps-only and never run a model kernel, like the reference's ps process that sits in ``server.join()``.
"""
from __future__ import annotations

import os

from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..engine_native import (CTRL_ARRIVE, CTRL_GSTEP, CTRL_T, CTRL_VERSION, NativeEngine, ShardState, Workspace,
                             needs_shadow, tc_operand, _pad8)
from ..models.spec import ModelSpec
from ..ops import native as nat
from .placement import Placement

WCTRL_STEP, WCTRL_PUSHBLK = 8, 9   # worker-local control words (inside the rank's own ctrl block)
CTRL_TICKET, CTRL_SHUTDOWN = 6, 7  # async: ticket dispenser / release word of a shard
CTRL_RING, RING_LEN = 16, 32       # async: arrival ring (worker id + 1 per served-in-order ticket)
CTRL_STAMP, CTRL_MASK = 16, 24     # backup workers (sync, R < N): per-worker stamps and the decided slot mask
WCTRL_TAG_CUR, WCTRL_TAG_NEXT = 10, 11  # backup workers: tag of my last push / of my next push (worker-local)
CTRL_DONE = 56                     # async: shard 0 holds one 'worker finished' word per worker (56..63)
WCTRL_MYTICKET = 48                # async: worker-local, one word per shard: version that serves my last push
DIRECT_THRESHOLD = 1 << 18         # tensors at least this big are consumed in place over NVLink, never mirrored
WCTRL_VER = 64                     # worker-local: version word of shard s at WCTRL_VER + s, stored by the owner shard
CTRL_WORDS = 128


@dataclass
class Topology:
    world: int
    ps_ranks: List[int]
    worker_ranks: List[int]
    layout: str

    @classmethod
    def make(cls, world: int, layout: str = "colocated", num_ps: int = 0) -> "Topology":
        if world < 1:
            raise ValueError("world size must be >= 1")
        if layout == "colocated":
            s = num_ps or world
            if not 1 <= s <= world:
                raise ValueError("colocated layout: need 1 <= ps shards <= GPUs")
            return cls(world, list(range(s)), list(range(world)), layout)
        if layout == "dedicated":
            s = num_ps or 1
            if not 1 <= s < world:
                raise ValueError("dedicated layout needs at least one ps GPU and one worker GPU")
            return cls(world, list(range(s)), list(range(s, world)), layout)
        raise ValueError("unknown layout %r" % layout)

    @property
    def num_shards(self) -> int:
        return len(self.ps_ranks)

    @property
    def num_workers(self) -> int:
        return len(self.worker_ranks)

    def shard_of(self, rank: int) -> Optional[int]:
        return self.ps_ranks.index(rank) if rank in self.ps_ranks else None

    def worker_of(self, rank: int) -> Optional[int]:
        return self.worker_ranks.index(rank) if rank in self.worker_ranks else None

    def describe(self) -> str:
        return "ps%d+w%d %s" % (self.num_shards, self.num_workers, self.layout)


def multicast_supported(device_index: int = 0) -> bool:
    """Does the driver expose NVSwitch multicast (NVLS) for this GPU?  (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED)"""
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        _, dev = cu.cuDeviceGet(device_index)
        _, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(v)
    except Exception:  # noqa: BLE001
        return False


def choose_transport(requested: Optional[str], topo: "Topology", sync: bool, full_aggregation: bool,
                     grad_dtype: str = "fp32", device_index: int = 0) -> str:
    """'peer' (cudaIpc peer memory, unicast) or 'nvls' (NVSwitch multicast).

    auto: the multicast transport where it pays and applies -- strict synchronous replicas on DEDICATED ps GPUs, where
    the star topology of the reference (every byte crosses the ps task's link once per worker, m:177) makes one
    NVLink port carry workers-x the parameter and gradient bytes (measured, 1 ps + 7 workers, tf32: 0.348 ms per step
    on the peer transport, 0.216 ms with the gradient sum done in the switch and a one-store publish).  Everything
    else -- colocated shards, async, backup workers, bf16 gradient slots -- uses the peer transport."""
    requested = requested or os.environ.get("TFD_TRANSPORT", "auto")
    if requested not in ("auto", "peer", "nvls"):
        raise ValueError("transport must be auto, peer or nvls")
    applicable = sync and full_aggregation and grad_dtype == "fp32"
    if requested == "nvls":
        if not applicable:
            raise ValueError("the nvls transport implements strict synchronous replicas with fp32 gradients only")
        return "nvls"
    if requested == "auto" and applicable and topo.layout == "dedicated" and multicast_supported(device_index):
        return "nvls"
    return "peer"


def _al(n: int, a: int = 1024) -> int:
    return (n + a - 1) // a * a


def direct_vars(spec: ModelSpec):
    """Tensor-core weights big enough to be consumed in place over NVLink (fused pull / push), never mirrored."""
    return {v.name for v in spec.variables if tc_operand(spec, v.name) and v.numel >= DIRECT_THRESHOLD}


class HeapLayout:
    """Byte offsets inside one rank's peer-mapped allocation.  Identical arithmetic on every rank.

    ``[ctrl | mirrors | w | m | v | shadow | slots]``.  The mirror region has the SAME offsets on every rank (it only
    depends on the model): one fp32 mirror per small variable and, on the bf16 arm, a bf16 mirror of the small
    tensor-core weights.  The owner shards store into the workers' mirrors from inside their Adam kernel."""

    def __init__(self, spec: ModelSpec, placement: Placement, shard: Optional[int], num_workers: int,
                 grad_bytes: int = 4, precision: str = "bf16"):
        self.grad_bytes = grad_bytes
        off = 0
        self.ctrl = off; off += 1024
        direct = direct_vars(spec)
        self.mw: Dict[str, int] = {}
        self.ms: Dict[str, int] = {}
        for v in spec.variables:
            if v.name in direct:
                continue
            self.mw[v.name] = off; off += _al((v.numel + 3) // 4 * 16, 256)
            if needs_shadow(spec, v.name, precision):
                rows = v.numel // v.cols
                self.ms[v.name] = off; off += _al(rows * _pad8(v.cols) * 2 + 16, 256)
        off = _al(off)
        self.n = self.shadow_n = 0
        self.w = self.m = self.v = self.shadow = self.slots = off
        if shard is not None:
            self.n = placement.shard_sizes[shard]
            _, _, _, self.shadow_n = ShardState.layout(spec, placement, shard, precision)
            self.w = off; off += _al(self.n * 4)
            self.m = off; off += _al(self.n * 4)
            self.v = off; off += _al(self.n * 4)
            self.shadow = off; off += _al(self.shadow_n * 2)
            self.slots = off; off += num_workers * _al(self.n * grad_bytes)
        self.slot_stride = _al(self.n * grad_bytes)
        self.total = max(off, 4096)

    def slot(self, worker: int) -> int:
        return self.slots + worker * self.slot_stride


class GpuClusterEngine(NativeEngine):
    """One rank (= one B200) of the ps/worker cluster.  Same public API as NativeEngine."""

    name = "native-sm100a-ps"
    # Critical-path kernels on a high-priority capture stream: with the pull/push kernels in the step the graph's
    # branches line up differently and conv2-wgrad (side branch) used to take the SMs ahead of conv2-dgrad, which
    # cannot be co-resident with it.  2 GPUs: 0.1524 -> 0.1429 ms per step.  TFD_PRIO=0 turns it off.
    _prio_default = "1"

    def __init__(self, spec: ModelSpec, topo: Topology, rank: int, device, learning_rate: float = 0.01,
                 seed: int = 0, init_stddev: float = 1.0, use_graphs: bool = True, conv2_ref: bool = False,
                 async_mode: bool = False, pull_mode: str = "auto", prefetch_max_numel: int = 8 << 20,
                 grad_dtype: str = "fp32", replicas_to_aggregate: int = 0, precision: Optional[str] = None):
        # backup workers: sync mode that aggregates only the first R of N gradients per update and drops stale ones
        self.R = replicas_to_aggregate if 0 < replicas_to_aggregate < topo.num_workers else topo.num_workers
        self.backup = (not async_mode) and self.R < topo.num_workers
        if self.backup and topo.layout != "dedicated":
            # a colocated rank applies its shard in its own worker stream: a straggling worker would stall its shard
            # and with it everybody -- only ps-only GPUs can run ahead of stragglers
            raise ValueError("replicas_to_aggregate < workers needs dedicated ps GPUs (layout='dedicated')")
        if grad_dtype not in ("fp32", "bf16"):
            raise ValueError("grad_dtype must be fp32 or bf16")
        self.grad_dtype, self.grad_bytes = grad_dtype, (2 if grad_dtype == "bf16" else 4)
        self.pull_mode, self.prefetch_max_numel = pull_mode, prefetch_max_numel
        if async_mode and topo.layout != "dedicated":
            # A colocated rank runs its worker kernels and its shard's apply in ONE in-order stream.  With
            # apply-on-arrival a worker blocked on "my ticket has been served" would sit in front of the very apply
            # kernel another rank is waiting for (A waits on B's shard, B on A's): only ps-only GPUs can serve
            # tickets unconditionally -- which is also the reference's topology (separate ps processes).
            raise ValueError("async replicas need dedicated ps GPUs (layout='dedicated')")
        self.async_mode = async_mode
        # Strict sync with fp32 slots runs the protocol WITHOUT stand-alone rendezvous kernels: gradients are written
        # straight into the owners' slots by the kernels that produce them, the arrival signal and the parameter
        # publish live inside the shard's Adam kernel, and workers wait on version words in their own memory.
        # Async, backup-worker and bf16-slot modes keep the pull / push kernels.
        self.v2 = (not async_mode) and (not self.backup) and grad_dtype == "fp32"
        if not dist.is_initialized():
            raise RuntimeError("GpuClusterEngine needs torch.distributed (bootstrap only) to be initialised")
        nat.lib()
        self.spec, self.topo, self.rank = spec, topo, rank
        self._set_precision(precision, conv2_ref)
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        nat.set_device(self.device.index or 0)
        self.lr, self.seed = learning_rate, seed
        self.use_graphs, self.conv2_ref = use_graphs, conv2_ref
        self.ws: Dict[int, Workspace] = {}
        self.my_shard = topo.shard_of(rank)
        self.my_worker = topo.worker_of(rank)
        # the initialisation seed is shared (every shard generates its slice of one Philox stream); the dropout stream
        # is per worker, as each of the reference's workers draws its own tf.nn.dropout mask (m:124)
        self.dropout_seed = seed + 31 * (self.my_worker or 0)
        self.is_ps, self.is_worker = self.my_shard is not None, self.my_worker is not None
        self.num_workers = topo.num_workers
        S = topo.num_shards
        self.placement = Placement(spec, S)
        self.layouts = [HeapLayout(spec, self.placement, topo.shard_of(r), topo.num_workers, self.grad_bytes,
                                   self.precision) for r in range(topo.world)]
        nat.set_slot_dtype(self.grad_dtype == "bf16")
        # ---- symmetric heap: allocate, exchange the 64-byte cudaIpc handles, map every ps rank's heap ----
        self.mc_base = 0  # multicast (NVLS) alias of the heap; only the experimental symmetric-memory heap has one
        self._opened: List[int] = []
        if os.environ.get("TFD_HEAP", "ipc") == "symm":
            # Experimental (not yet run on hardware): the heap comes from torch's symmetric-memory allocator (VMM
            # allocations exchanged over the process group), which also yields the multicast address the NVLS
            # plan in DESIGN.md needs.  Every rank allocates the same size; peers are addressed as before.
            import torch.distributed._symmetric_memory as symm
            total = max(l.total for l in self.layouts)
            self._symm_buf = symm.empty(total, dtype=torch.uint8, device=self.device)
            self._symm_buf.zero_()
            torch.cuda.synchronize(self.device)
            self._symm = symm.rendezvous(self._symm_buf, dist.group.WORLD)
            self.heap_ptr = self._symm_buf.data_ptr()
            self.peer_base: Dict[int, int] = {r: int(self._symm.buffer_ptrs[r]) for r in range(topo.world)}
            if getattr(self._symm, "has_multicast_support", False):
                self.mc_base = int(self._symm.multicast_ptr)
        else:
            self.heap_ptr, handle = nat.ipc_alloc(self.layouts[rank].total)
            handles: List[Optional[bytes]] = [None] * topo.world
            dist.all_gather_object(handles, handle)
            self.peer_base = {rank: self.heap_ptr}
            for r in range(topo.world):  # every heap: workers read the shards', shards write the workers' mirrors
                if r != rank and (r in topo.ps_ranks or (self.is_ps and self.v2)):
                    self.peer_base[r] = nat.ipc_open(handles[r])
                    self._opened.append(self.peer_base[r])
        L = self.layouts[rank]
        self.ctrl = nat.tensor_from_ptr(self.heap_ptr + L.ctrl, CTRL_WORDS, torch.int32, self.device)
        # ---- ps role ----
        self.shard = None
        self._shard_layouts = [ShardState.layout(spec, self.placement, s, self.precision) for s in range(S)]
        if self.is_ps:
            offs = {"w": L.w, "m": L.m, "v": L.v, "shadow": L.shadow, "ctrl": L.ctrl}
            self.shard = ShardState(spec, self.placement, self.my_shard, self.device,
                                    alloc=lambda name, numel, dtype: nat.tensor_from_ptr(
                                        self.heap_ptr + offs[name], numel, dtype, self.device),
                                    precision=self.precision)
            self.slot_ptrs = [self.heap_ptr + L.slot(w) for w in range(topo.num_workers)]
            self.shard.init_variables(seed, init_stddev)
        # ---- worker role ----
        # Big tensor-core operands are never pushed through a local gradient copy (push-fused wgrad).  For the pull
        # there are two strategies: "fused" -- forward and dgrad GEMMs TMA-load B from the owner GPUs (bandwidth-bound
        # layers: 97% of the NVLink roofline on 8192x8192); "prefetch" -- a few-MB matrix such as the CNN's wd1 is
        # copied once per step on a side stream *under* the conv layers and both GEMMs read the local copy, which
        # takes two latency-bound NVLink reads off the critical path (2 GPUs: 154.9 vs 157.7 us/step;
        # 1 ps + 7 workers: 222 vs 317 us/step).  "auto" therefore prefetches matrices up to prefetch_max_numel
        # elements and fuses the pull of everything bigger; "fused" / "prefetch" force one strategy.
        self.direct = direct_vars(spec)
        self.prefetched = {n for n in self.direct
                           if pull_mode in ("prefetch", "auto") and spec.by_name()[n].numel <= prefetch_max_numel}
        if self.is_worker:
            self._build_worker_tables()
        self._pub = self._build_pub() if (self.is_ps and self.v2) else None
        torch.cuda.synchronize(self.device)
        dist.barrier()  # X6: everybody waits until every shard has been initialised
        self._publish_host(0)

    # ---- addressing helpers ----------------------------------------------------------------
    def _ps_rank(self, shard: int) -> int:
        return self.topo.ps_ranks[shard]

    def _ctrl_ptr(self, shard: int, word: int) -> int:
        r = self._ps_rank(shard)
        return self.peer_base[r] + self.layouts[r].ctrl + 4 * word

    def _w_ptr(self, shard: int, elem_off: int) -> int:
        r = self._ps_rank(shard)
        return self.peer_base[r] + self.layouts[r].w + 4 * elem_off

    def _tc_ptr(self, shard: int, var: str) -> int:
        """What the tensor cores read of `var` on its owner shard: the bf16 shadow, or (tf32 arm) the fp32 master."""
        r = self._ps_rank(shard)
        if self.tf32:
            (p,) = [q for q in self.placement.pieces[var] if q.shard == shard]
            return self.peer_base[r] + self.layouts[r].w + 4 * p.offset
        return self.peer_base[r] + self.layouts[r].shadow + 2 * self._shard_layouts[shard][1][var]

    def _tc_ld(self, cols: int) -> int:
        return cols if self.tf32 else _pad8(cols)

    def _slot_ptr(self, shard: int, elem_off: int) -> int:
        r = self._ps_rank(shard)
        return self.peer_base[r] + self.layouts[r].slot(self.my_worker) + self.grad_bytes * elem_off

    def _wver_ptr(self, rank: int, shard: int) -> int:
        """Version word of `shard` in the heap of worker rank `rank` (the owner stores it, the worker polls it)."""
        return self.peer_base[rank] + self.layouts[rank].ctrl + 4 * (WCTRL_VER + shard)

    def _build_worker_tables(self) -> None:
        dev = self.device
        L = self.layouts[self.rank]
        S = self.topo.num_shards
        tcdt = torch.float32 if self.tf32 else torch.bfloat16
        esz = 4 if self.tf32 else 2
        self.mirror_w: Dict[str, torch.Tensor] = {}
        self.mirror_s: Dict[str, torch.Tensor] = {}
        self.local_g: Dict[str, torch.Tensor] = {}
        pull, push = [], []
        seen_shards = set()
        # version word a consumer waits on for shard s: v2 -- my own copy, stored by the owner; legacy -- the owner's
        ver_of = (lambda s: self._wver_ptr(self.rank, s)) if self.v2 else (lambda s: self._ctrl_ptr(s, CTRL_VERSION))
        self._gate = [self._wver_ptr(self.rank, s) for s in range(S)]
        for v in self.spec.variables:
            if v.name in self.direct:
                continue
            (p,) = self.placement.pieces[v.name]  # small tensors live whole on one shard
            rows, cols = p.row_end - p.row_start, p.cols
            if self.v2:
                # parameters: mirror inside my heap, written by the owner's Adam kernel.  gradients: the kernels write
                # straight into my slot on the owner GPU (a torch view over peer-mapped memory)
                self.mirror_w[v.name] = nat.tensor_from_ptr(self.heap_ptr + L.mw[v.name], p.numel, torch.float32,
                                                            dev).view(rows, cols)
                self.local_g[v.name] = nat.tensor_from_ptr(self._slot_ptr(p.shard, p.offset), p.numel, torch.float32,
                                                           dev).view(rows, cols)
                if needs_shadow(self.spec, v.name, self.precision):
                    ld = _pad8(cols)
                    self.mirror_s[v.name] = nat.tensor_from_ptr(self.heap_ptr + L.ms[v.name], rows * ld, torch.bfloat16,
                                                                dev).view(rows, ld)
                continue
            self.mirror_w[v.name] = torch.zeros(rows, cols, dtype=torch.float32, device=dev)
            self.local_g[v.name] = torch.zeros(rows, cols, dtype=torch.float32, device=dev)
            ver = ver_of(p.shard)
            pull.append((self._w_ptr(p.shard, p.offset), self.mirror_w[v.name].data_ptr(), p.numel, ver, p.shard))
            push.append((self.local_g[v.name].data_ptr(), self._slot_ptr(p.shard, p.offset), p.numel))
            seen_shards.add(p.shard)
            if needs_shadow(self.spec, v.name, self.precision):
                ld = _pad8(cols)
                self.mirror_s[v.name] = torch.zeros(rows, ld, dtype=torch.bfloat16, device=dev)
                pull.append((self._tc_ptr(p.shard, v.name), self.mirror_s[v.name].data_ptr(), rows * ld // 2, ver, p.shard))
        if not self.v2:
            for s in range(S):  # wait for EVERY shard's version (its slots are then free again)
                if s not in seen_shards:
                    pull.append((self._w_ptr(s, 0), self.ctrl.data_ptr() + 4 * 60, 0, ver_of(s), s))
        self._pull, self._push = pull, push
        self._arrive = [self._ctrl_ptr(s, CTRL_ARRIVE) for s in range(S)]
        # per direct variable: shard pieces in row order
        self._direct_info = {}
        self._wg_stage, self._push_branches, self._push_used = {}, [], 0
        self._prefetch = []
        for name in self.direct:
            pcs = sorted(self.placement.pieces[name], key=lambda q: q.row_start)
            ld = self._tc_ld(pcs[0].cols)
            if name in self.prefetched:
                self.mirror_s[name] = torch.zeros(pcs[-1].row_end, ld, dtype=tcdt, device=dev)
                for q in pcs:
                    self._prefetch.append((self._tc_ptr(q.shard, name),
                                           self.mirror_s[name].data_ptr() + esz * q.row_start * ld,
                                           (q.row_end - q.row_start) * ld * esz // 4, ver_of(q.shard), q.shard))
            self._direct_info[name] = {
                "rows": [q.row_start for q in pcs] + [pcs[-1].row_end], "cols": pcs[0].cols, "ld": ld,
                "shadow": [self._tc_ptr(q.shard, name) for q in pcs],
                "slot": [self._slot_ptr(q.shard, q.offset) for q in pcs], "shard": [q.shard for q in pcs],
                "ver": None if (self.async_mode or self.backup) else [ver_of(q.shard) for q in pcs]}

    def _build_pub(self):
        """What my shard's Adam kernel publishes: the new values of every small variable I own go into EVERY worker's
        mirror (same offsets in every heap), then every worker's version word for my shard is released; on a
        colocated rank the kernel also sends this rank's own arrival signals first."""
        L = self.layouts[self.rank]
        pieces = []
        for p in self.placement.shard_pieces(self.my_shard):
            if p.var in self.direct:
                pieces.append((-1, -1, 0))
            else:
                pieces.append((L.mw[p.var], L.ms.get(p.var, -1), _pad8(p.cols)))
        wr = self.topo.worker_ranks
        return nat.make_pub(pieces, [self.peer_base[r] for r in wr], [self._wver_ptr(r, self.my_shard) for r in wr],
                            signal_arrive=self._arrive if self.is_worker else (),
                            local_step=(self.ctrl.data_ptr() + 4 * WCTRL_STEP) if self.is_worker else None)

    def _publish_host(self, version: Optional[int] = None) -> None:
        """Initialisation / restore: every shard copies its small variables into every worker's mirror and sets every
        worker's version word for itself, from the host (steady state is done by the Adam kernel).  Collective."""
        if not self.v2:
            return
        if self.is_ps:
            s = self.shard
            if version is None:
                version = int(s.ctrl[CTRL_VERSION].item())
            for r in self.topo.worker_ranks:
                Lr = self.layouts[r]
                for p in self.placement.shard_pieces(self.my_shard):
                    if p.var in self.direct:
                        continue
                    nat.tensor_from_ptr(self.peer_base[r] + Lr.mw[p.var], p.numel, torch.float32, self.device).copy_(
                        s.w[p.offset:p.offset + p.numel])
                    if p.var in Lr.ms:
                        rows, ld = p.row_end - p.row_start, _pad8(p.cols)
                        nat.tensor_from_ptr(self.peer_base[r] + Lr.ms[p.var], rows * ld, torch.bfloat16,
                                            self.device).copy_(s.shadow_view(p.var).reshape(-1))
                nat.tensor_from_ptr(self._wver_ptr(r, self.my_shard), 1, torch.int32, self.device).fill_(int(version))
        torch.cuda.synchronize(self.device)
        dist.barrier()

    # ---- parameter / gradient views used by the kernel sequences -----------------------------
    def P(self, name: str) -> torch.Tensor:
        return self.mirror_w[name]

    def S(self, name: str) -> torch.Tensor:
        """What the tensor cores read locally: a bf16 mirror, a prefetched copy of a big matrix, or (tf32 arm, small
        tensor-core weights such as wc2) the fp32 mirror itself."""
        if name in self.mirror_s:
            return self.mirror_s[name]
        return self.mirror_w[name]

    def G(self, name: str) -> torch.Tensor:
        return self.local_g[name]

    def step_ptr(self):
        return self.ctrl[WCTRL_STEP:]

    def gemm_fwd(self, wname, A, M, out, *, mode, bn, splits=1, split_stride=0, bias=None, ld_out=None):
        if wname in self.prefetched:
            self._join_prefetch()
        if wname not in self.direct or wname in self.prefetched:
            return super().gemm_fwd(wname, A, M, out, mode=mode, bn=bn, splits=splits, split_stride=split_stride,
                                    bias=bias, ld_out=ld_out)
        d = self._direct_info[wname]
        nat.gemm_bf16_seg(A, 0, d["shadow"], 1, M, d["cols"], d["rows"][-1], out, mode=mode, lda=A.stride(0),
                          ldb=d["ld"], ld_out=ld_out if ld_out is not None else out.stride(0), seg_mode=nat.SEG_B_K,
                          seg_rows=d["rows"], wait_flags=d["ver"], wait_value_ptr=self.step_ptr(), bn=bn,
                          splits=splits, split_stride=split_stride, bias=bias, tf32=self.tf32)

    def gemm_dgrad(self, wname, dY, M, out, *, mode, bn, splits=1, split_stride=0, relu_mask=None, ld_out=None):
        if wname in self.prefetched:
            self._join_prefetch()
        if wname not in self.direct or wname in self.prefetched:
            return super().gemm_dgrad(wname, dY, M, out, mode=mode, bn=bn, splits=splits, split_stride=split_stride,
                                      relu_mask=relu_mask, ld_out=ld_out)
        d = self._direct_info[wname]
        nat.gemm_bf16_seg(dY, 0, d["shadow"], 0, M, d["rows"][-1], d["cols"], out, mode=mode, lda=dY.stride(0),
                          ldb=d["ld"], ld_out=ld_out if ld_out is not None else out.stride(0), seg_mode=nat.SEG_B_N,
                          seg_rows=d["rows"], wait_flags=d["ver"], wait_value_ptr=self.step_ptr(), bn=bn,
                          splits=splits, split_stride=split_stride, relu_mask=relu_mask,
                          ld_mask=relu_mask.stride(0) if relu_mask is not None else 0, tf32=self.tf32)

    def gemm_wgrad(self, wname, X, dY, Bsz, *, bn):
        if wname not in self.direct:
            return super().gemm_wgrad(wname, X, dY, Bsz, bn=bn)
        d = self._direct_info[wname]
        nseg = len(d["slot"])
        out_ptrs, copies = d["slot"], []
        # (`auto`: only for the few-MB matrices that are also prefetched -- the CNN's wd1, the measured case; the
        # epilogue push of an 8192 x 8192 MLP layer already runs at 0.93 of the NVLink roofline inside the GEMM)
        forced = os.environ.get("TFD_PUSH", "auto") == "ce"
        if (forced or wname in self.prefetched) and self._use_ce(
                "TFD_PUSH", len([x for x in d["shard"] if not (self.is_ps and x == self.my_shard)])):
            # Strict sync: the GEMM stores only MY shard's rows straight into their slot; the other shards' rows go
            # into a local staging matrix and from there to their owners by copy engine, one branch per owner, under
            # the conv backward kernels.  Stored from the epilogue over NVLink they kept the GEMM's CTAs resident
            # 6-10 us longer (2-8 GPUs) and slowed the kernels that share its SMs -- fc1 dgrad, pool2 backward and the
            # head -- by as much, on the critical path.
            st = self._wg_stage.get(wname)
            if st is None:
                st = self._wg_stage[wname] = torch.empty(d["rows"][-1], d["cols"], dtype=torch.float32,
                                                         device=self.device)
            mine = self.my_shard if self.is_ps else -1
            out_ptrs = []
            for i in range(nseg):
                if d["shard"][i] == mine:
                    out_ptrs.append(d["slot"][i])
                else:
                    sp = st.data_ptr() + 4 * d["rows"][i] * d["cols"]
                    out_ptrs.append(sp)
                    copies.append((d["slot"][i], sp, 4 * (d["rows"][i + 1] - d["rows"][i]) * d["cols"]))
        nat.gemm_bf16_seg(X, 1, [dY.data_ptr()] + [0] * (nseg - 1), 1, d["rows"][-1], d["cols"], Bsz, None,
                          mode=nat.EPI_STORE_BF16 if self.grad_dtype == "bf16" else nat.EPI_STORE_F32, lda=X.stride(0),
                          ldb=dY.stride(0), ld_out=d["cols"], seg_mode=nat.SEG_OUT_M, seg_rows=d["rows"],
                          out_ptrs=out_ptrs, bn=bn, tf32=self.tf32)
        if copies:
            done = torch.cuda.current_stream().record_event()
            while len(self._push_branches) < len(copies):
                self._push_branches.append(torch.cuda.Stream(device=self.device))
            for br, (dst, src, nbytes) in zip(self._push_branches, copies):
                br.wait_event(done)
                with torch.cuda.stream(br):
                    nat.memcpy_async(dst, src, nbytes)
            self._push_used = max(self._push_used, len(copies))

    # ---- step pieces -----------------------------------------------------------------------
    def conv1_fwd(self, ws, training: bool = True) -> None:
        if not self.v2:
            return super().conv1_fwd(ws, training)
        # the step's first kernel waits -- on version words in THIS GPU's memory -- until every shard has applied the
        # update this step builds on (which also means my slots have been consumed and may be overwritten)
        nat.conv1_fwd(ws.x, self.P("weights/wc1"), self.P("biases/bc1"), ws.pool1, ws.am1, ws.B, tf32=self.tf32,
                      gate_flags=self._gate, want_ptr=self.step_ptr())
        if getattr(self, "_pf_local", None):  # the device-local part of the prefetch goes behind conv1, under conv2
            main = torch.cuda.current_stream()
            self._pf_stream.wait_stream(main)
            with torch.cuda.stream(self._pf_stream):
                self._issue_local_prefetch()

    def _pre_step(self, ws, training: bool) -> None:
        """Pull (X1).  v2: nothing to copy -- the owners have already stored the small parameters into my mirror; only
        the wait remains, folded into conv1 (CNN) or a one-block gate kernel (MLP).  Legacy modes: small tensors are
        copied once the owning shard's version has caught up with this worker's step counter; the same kernel thereby
        waits for every shard (slots are free to overwrite)."""
        if self._prefetch:
            self._launch_prefetch(training)
        if self.v2:
            if self.spec.kind != "cnn":
                nat.gate_flags(self._gate, self.step_ptr())
            return
        p = self._pull
        if self.backup and not training:
            # evaluation after training: my last push may carry a tag the ps will never apply (it has already done
            # its train_steps updates), so read whatever is published now
            nat.pull_multi([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p], 0)
            return
        if self.backup:  # wait until my last tagged push has been applied or overtaken; derive the next tag
            tagcur = self.ctrl.data_ptr() + 4 * WCTRL_TAG_CUR
            seg = next(i for i, q in enumerate(p) if q[4] == self.placement.global_step_shard)
            nat.pull_multi_tag([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p],
                               [tagcur] * len(p), seg, self.ctrl[WCTRL_TAG_NEXT:])
            return
        if self.async_mode:  # wait (per shard) until my previous gradient has been applied: read-your-writes
            nat.pull_multi_async([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p],
                                 [self.ctrl.data_ptr() + 4 * (WCTRL_MYTICKET + s[4]) for s in p])
            return
        nat.pull_multi([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p], 0,
                       want_ptr=self.step_ptr())

    def _launch_prefetch(self, training: bool = True) -> None:
        """Copy the prefetched matrices' bf16 shadow from the owner GPUs on a side stream (joined by the first GEMM
        that reads them), gated on the same parameter version as everything else."""
        main = torch.cuda.current_stream()
        if not hasattr(self, "_pf_stream"):
            self._pf_stream = torch.cuda.Stream(device=self.device)
            self._pf_branches, self._pf_used, self._pf_local = [], 0, []
        self._pf_stream.wait_stream(main)
        p = self._prefetch
        with torch.cuda.stream(self._pf_stream):
            if self.backup and not training:
                nat.pull_multi([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p], 0)
            elif self.backup:
                nat.pull_multi_async([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p],
                                     [self.ctrl.data_ptr() + 4 * WCTRL_TAG_CUR] * len(p))
            elif self.async_mode:
                nat.pull_multi_async([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p],
                                     [self.ctrl.data_ptr() + 4 * (WCTRL_MYTICKET + s[4]) for s in p])
            elif self._use_ce("TFD_PREFETCH", len([s for s in p if not (self.is_ps and s[4] == self.my_shard)])):
                # strict sync: a one-block gate on the LOCAL version words of the owner shards, then one copy-engine
                # transfer per remote shard piece, each on its own graph branch (a peer copy takes ~7 us to start; in
                # one stream seven of them would queue up).  A copy kernel lost its SMs to conv2 (one CTA per SM) and
                # finished late (8 GPUs: 42 us, on the critical path of the fc1 GEMM); copy engines need no SM.
                nat.gate_flags(sorted({s[3] for s in p}), self.step_ptr())
                gated = self._pf_stream.record_event()
                mine = self.my_shard if self.is_ps else -1
                remote = [s for s in p if s[4] != mine]
                while len(self._pf_branches) < len(remote):
                    self._pf_branches.append(torch.cuda.Stream(device=self.device))
                for br, (src, dst, n, _ver, _shard) in zip(self._pf_branches, remote):
                    br.wait_event(gated)
                    with torch.cuda.stream(br):
                        nat.memcpy_async(dst, src, n * 4)  # piece lengths are counted in fp32 words
                self._pf_used = len(remote)
                # my own shard's piece is a device-local copy, which the driver runs as a kernel: next to conv1 it
                # took that kernel's SMs (conv1 17 us instead of 9), so the CNN issues it behind conv1 (conv1_fwd)
                self._pf_local = [s for s in p if s[4] == mine]
                if self.spec.kind != "cnn":
                    self._issue_local_prefetch()
            else:
                nat.pull_multi([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], [s[3] for s in p], 0,
                               want_ptr=self.step_ptr())
        self._pf_pending = True

    def _use_ce(self, env: str, n_remote: int) -> bool:
        """Copy engines instead of SM kernels / epilogue stores for the big-matrix prefetch (env TFD_PREFETCH) and
        gradient push (TFD_PUSH) of strict-sync mode: `ce`, `kernel`, or `auto` (default).  Measured: with ONE remote
        piece per matrix (2 colocated GPUs) the copy-engine paths are faster; with
        seven (8 GPUs) the memcpy nodes of one GPU queue behind each other (~5 us start latency each) and the step got
        30 us SLOWER than with the kernel paths, so `auto` uses the copy engines only in the measured-good case: a
        colocated rank with exactly one remote piece."""
        if not self.v2:
            return False
        mode = os.environ.get(env, "auto")
        if mode in ("ce", "kernel"):
            return mode == "ce"
        return self.is_ps and self.is_worker and n_remote == 1

    def _issue_local_prefetch(self) -> None:
        for src, dst, n, _ver, _shard in self._pf_local:
            nat.memcpy_async(dst, src, n * 4)
        self._pf_local = []

    def _join_prefetch(self) -> None:
        if getattr(self, "_pf_pending", False):
            main = torch.cuda.current_stream()
            main.wait_stream(self._pf_stream)
            for br in self._pf_branches[:self._pf_used]:
                main.wait_stream(br)
            self._pf_used = 0
            self._pf_pending = False

    def _post_backward(self, ws) -> None:
        """Push (X2): small gradients into this worker's slots, then arrive += 1 on every shard.  v2: the gradient
        kernels have already written the slots; a colocated rank signals from inside its Adam kernel, a worker on a
        GPU without a shard with one tiny block."""
        if self.v2:
            main = torch.cuda.current_stream()
            for br in self._push_branches[:self._push_used]:  # copy-engine pushes of the big gradients
                main.wait_stream(br)
            self._push_used = 0
            if not self.is_ps:
                nat.signal_arrive(self._arrive, self.ctrl[WCTRL_STEP:])
            return
        p = self._push
        if self.backup:
            S = self.topo.num_shards
            nat.push_multi_stamp([s[0] for s in p], [s[1] for s in p], [s[2] for s in p],
                                 [self._ctrl_ptr(s, CTRL_STAMP + self.my_worker) for s in range(S)],
                                 self.ctrl[WCTRL_TAG_NEXT:], self.ctrl[WCTRL_TAG_CUR:], self.ctrl[WCTRL_PUSHBLK:],
                                 local_step=self.ctrl[WCTRL_STEP:])
            return
        if self.async_mode:
            S = self.topo.num_shards
            nat.push_multi_async([s[0] for s in p], [s[1] for s in p], [s[2] for s in p],
                                 [self._ctrl_ptr(s, CTRL_TICKET) for s in range(S)],
                                 [self._ctrl_ptr(s, CTRL_RING) for s in range(S)], RING_LEN, self.my_worker,
                                 self.ctrl[WCTRL_MYTICKET:], self.ctrl[WCTRL_PUSHBLK:], local_step=self.ctrl[WCTRL_STEP:])
            return
        nat.push_multi([s[0] for s in p], [s[1] for s in p], [s[2] for s in p], self._arrive, 1.0,
                       self.ctrl[WCTRL_PUSHBLK:], local_step=self.ctrl[WCTRL_STEP:])

    def _apply(self) -> None:
        """Shard-side reduce + 1/R + Adam + shadow publish, gated on the arrivals of this step."""
        if not self.is_ps:
            return
        s = self.shard
        if self.backup:  # first R of N gradients stamped for this update, averaged; stale stamps never match
            nat.adam_apply_backup(s.w, s.m, s.v, None if self.tf32 else s.shadow, self.slot_ptrs, self.num_workers, self.R, s.n, s.pieces,
                                  self.lr, s.ctrl[CTRL_T:], s.ctrl[CTRL_STAMP:], s.ctrl[CTRL_MASK:],
                                  s.ctrl[CTRL_SHUTDOWN:], s.ctrl[CTRL_VERSION:], s.ctrl[CTRL_GSTEP:])
            return
        if self.async_mode:  # one arriving gradient, in ticket order, scale 1 (apply-on-arrival)
            nat.adam_serve(s.w, s.m, s.v, None if self.tf32 else s.shadow, self.slot_ptrs, self.num_workers, s.n, s.pieces, self.lr,
                           s.ctrl[CTRL_T:], s.ctrl[CTRL_RING:], RING_LEN, s.ctrl[CTRL_SHUTDOWN:], s.ctrl[CTRL_VERSION:],
                           s.ctrl[CTRL_GSTEP:])
            return
        nat.adam_apply(s.w, s.m, s.v, None if self.tf32 else s.shadow, self.slot_ptrs, self.num_workers,
                       1.0 / self.num_workers, s.n, s.pieces, self.lr, s.ctrl[CTRL_T:], s.ctrl[CTRL_ARRIVE:],
                       s.ctrl[CTRL_VERSION:], s.ctrl[CTRL_GSTEP:], pub=self._pub)

    def _sequence(self, ws, mode: str, keep_prob: float, mask_out=None) -> None:
        if self.is_worker:
            return super()._sequence(ws, mode, keep_prob, mask_out)
        if mode == "train":  # dedicated ps: nothing but the apply, like the reference's passive ps (m:160-161)
            for _ in range(self.num_workers if self.async_mode else 1):  # async: one serve per worker push
                self._apply()

    # ---- public API ------------------------------------------------------------------------
    def workspace(self, B: int) -> Workspace:
        if B not in self.ws:
            self.ws[B] = Workspace(self.spec, B if self.is_worker else 1, self.device, self.precision)
            self.ws[B].B = B if self.is_worker else 1
        return self.ws[B]

    def load_batch(self, ws, x, y) -> None:
        if self.is_worker:
            super().load_batch(ws, x, y)

    @property
    def global_step(self) -> int:
        t = nat.tensor_from_ptr(self._ctrl_ptr(self.placement.global_step_shard, CTRL_GSTEP), 1, torch.int32, self.device)
        return int(t.item())

    def describe(self) -> str:
        return self.topo.describe() + " " + self.precision + (" async" if self.async_mode else " sync") + \
            (" R=%d-of-%d" % (self.R, self.num_workers) if self.backup else "") + \
            (" bf16-grad-push" if self.grad_dtype == "bf16" else "")

    def _read_arena(self, which: str) -> Dict[str, torch.Tensor]:
        """Gather a full name->tensor dict by reading every shard's arena through the peer mapping."""
        torch.cuda.synchronize(self.device)
        out = {}
        for v in self.spec.variables:
            parts = []
            for p in sorted(self.placement.pieces[v.name], key=lambda q: q.row_start):
                r = self._ps_rank(p.shard)
                base = self.peer_base[r] + getattr(self.layouts[r], which) + 4 * p.offset
                parts.append(nat.tensor_from_ptr(base, p.numel, torch.float32, self.device).cpu()
                             .view(p.row_end - p.row_start, p.cols))
            out[v.name] = torch.cat(parts, 0).reshape(v.shape).clone()
        return out

    def get_params(self) -> Dict[str, torch.Tensor]:
        return self._read_arena("w")

    def get_adam_state(self) -> dict:
        r = self._ps_rank(0)
        t = int(nat.tensor_from_ptr(self._ctrl_ptr(0, CTRL_T), 1, torch.int32, self.device).item())
        return {"t": t, "beta1_power": 0.9 ** t, "beta2_power": 0.999 ** t, "m": self._read_arena("m"),
                "v": self._read_arena("v")}

    def get_grads(self):
        raise NotImplementedError("gradients live in the ps staging slots; use get_slot_grads()")

    def get_slot_grads(self, worker: int) -> Dict[str, torch.Tensor]:
        """What worker `worker` pushed last (read back from the shards' staging slots): tests."""
        torch.cuda.synchronize(self.device)
        out = {}
        for v in self.spec.variables:
            parts = []
            for p in sorted(self.placement.pieces[v.name], key=lambda q: q.row_start):
                r = self._ps_rank(p.shard)
                base = self.peer_base[r] + self.layouts[r].slot(worker) + self.grad_bytes * p.offset
                dt = torch.bfloat16 if self.grad_dtype == "bf16" else torch.float32
                parts.append(nat.tensor_from_ptr(base, p.numel, dt, self.device).float().cpu()
                             .view(p.row_end - p.row_start, p.cols))
            out[v.name] = torch.cat(parts, 0).reshape(v.shape).clone()
        return out

    def set_params(self, params: Dict[str, torch.Tensor]) -> None:
        """Chief-side restore: every ps rank copies its own pieces (call on all ranks with the same dict)."""
        if self.is_ps:
            s = self.shard
            for v in self.spec.variables:
                p = s.piece_of(v.name)
                if p is not None:
                    full = params[v.name].to(torch.float32).reshape(-1, p.cols)
                    s.view(s.w, v.name).copy_(full[p.row_start:p.row_end].to(self.device))
            s.publish_shadow()
        torch.cuda.synchronize(self.device)
        dist.barrier()
        self._publish_host()

    def shutdown(self) -> None:
        torch.cuda.synchronize(self.device)
        dist.barrier()
        for p in self._opened:
            nat.ipc_close(p)
        self._opened = []

    @classmethod
    def from_bench_args(cls, spec, a, rank: int, world: int, device) -> "GpuClusterEngine":
        topo = Topology.make(world, a.layout, a.ps)
        return cls(spec, topo, rank, device, learning_rate=0.01, seed=0, init_stddev=a.init_stddev,
                   async_mode=getattr(a, "async_mode", False), pull_mode=getattr(a, "pull", "auto"),
                   grad_dtype=getattr(a, "grad_dtype", "fp32"), precision=getattr(a, "precision", None))

    # ---- restore / release helpers used by the cluster program -------------------------------
    def set_adam_state(self, sd: dict) -> None:
        if self.is_ps:
            s = self.shard
            for v in self.spec.variables:
                p = s.piece_of(v.name)
                if p is not None:
                    for arena, key in ((s.m, "m"), (s.v, "v")):
                        full = sd[key][v.name].to(torch.float32).reshape(-1, p.cols)
                        s.view(arena, v.name).copy_(full[p.row_start:p.row_end].to(self.device))
            # every protocol counter restarts at the restored update index T: version == applies so far, the
            # arrival counter has seen workers*T pushes, the ticket dispenser has handed out T tickets
            T = int(sd["t"])
            s.ctrl[CTRL_T] = T
            s.ctrl[CTRL_VERSION] = T
            s.ctrl[CTRL_ARRIVE] = T * self.num_workers
            s.ctrl[CTRL_TICKET] = T
        if self.is_worker:
            # sync workers wait for version >= their step counter; in every mode the dropout stream is indexed by it,
            # so a resumed run continues the stream instead of replaying the masks of step 0
            self.ctrl[WCTRL_STEP] = int(sd["t"])
        torch.cuda.synchronize(self.device)
        dist.barrier()
        self._publish_host(int(sd["t"]))

    def set_global_step(self, step: int) -> None:
        if self.is_ps:
            self.shard.ctrl[CTRL_GSTEP] = int(step)
        torch.cuda.synchronize(self.device)

    def request_shutdown(self) -> None:
        """Release ps GPUs that are waiting for a ticket (async): the clean replacement for killing the ps process."""
        for s in range(self.topo.num_shards):
            nat.tensor_from_ptr(self._ctrl_ptr(s, CTRL_SHUTDOWN), 1, torch.int32, self.device).fill_(1)
        torch.cuda.synchronize(self.device)

    def mark_done(self) -> None:
        """Async: tell the chief this worker will push no more gradients (its own word in shard 0's ctrl block)."""
        nat.tensor_from_ptr(self._ctrl_ptr(0, CTRL_DONE + self.my_worker), 1, torch.int32, self.device).fill_(1)
        torch.cuda.synchronize(self.device)

    def all_workers_done(self) -> bool:
        t = nat.tensor_from_ptr(self._ctrl_ptr(0, CTRL_DONE), self.num_workers, torch.int32, self.device)
        return bool(int(t.sum().item()) == self.num_workers)

    def shutdown_requested(self) -> bool:
        return bool(self.is_ps and int(self.shard.ctrl[CTRL_SHUTDOWN].item()) != 0)


def run_gpu_cluster_task(flags, cluster, mnist, log) -> dict:
    """The cluster program (mnist_python_{m,w1,w2}.py) when ps_hosts / worker_hosts name GPUs of one box
    (``gpu:<ordinal>``): one OS process per task, ps tasks on their own GPUs (``/root/reference/mnist_python_m.py:
    145-320``).  ps and worker tasks must name different GPUs here: two processes cannot share one GPU while their
    kernels wait on each other; the colocated layout is reached through ``python -m tensorflow_distributed_b200.launch``
    (one process per GPU that is both worker and ps shard) and bench.py."""
    addr = cluster.task_address(flags.job_name, flags.task_index)
    gpus = [cluster.task_address(j, i).gpu_ordinal for j in ("ps", "worker") for i in range(cluster.num_tasks(j))]
    if len(set(gpus)) != len(gpus):
        raise ValueError("gpu cluster: every ps and worker task needs its own GPU (got %s)" % gpus)
    if not torch.cuda.is_available() or torch.cuda.device_count() <= max(gpus):
        raise RuntimeError("ps_hosts/worker_hosts name GPUs %s but this process sees %d CUDA device(s); use host:port "
                           "entries with --backend=cpu for the CPU/gloo transport"
                           % (sorted(gpus), torch.cuda.device_count() if torch.cuda.is_available() else 0))
    device = torch.device("cuda:%d" % addr.gpu_ordinal)
    torch.cuda.set_device(device)
    rank, world = cluster.rank_of(flags.job_name, flags.task_index), cluster.world_size
    host, port = cluster.master_endpoint()
    import datetime
    dist.init_process_group("nccl", init_method="tcp://%s:%d" % (host, port), rank=rank, world_size=world,
                            device_id=device, timeout=datetime.timedelta(seconds=flags.step_timeout_secs))
    topo = Topology(world, list(range(cluster.num_ps)), list(range(cluster.num_ps, world)), "dedicated")
    return _gpu_task(flags, topo, rank, device, mnist, log)


def _gpu_task(flags, topo: Topology, rank: int, device, mnist, log) -> dict:
    """One rank of the GPU cluster program, whatever decided the topology (host lists or the single-box launcher)."""
    import time

    from .. import checkpoint
    from .. import cluster as cl
    from ..models.spec import make_spec
    from ..supervisor import Supervisor
    from ..trainer import _maybe_inject_fault

    spec = make_spec(flags.model, flags.hidden_units, flags.mlp_layers)
    R = cl.effective_replicas_to_aggregate(flags, topo.num_workers)
    if flags.sync_replicas and R != topo.num_workers and topo.layout != "dedicated":
        log("note: colocated ps shards aggregate all %d workers per step (replicas_to_aggregate=%d needs dedicated "
            "ps GPUs)" % (topo.num_workers, R))
    widx = topo.worker_of(rank)
    is_worker = widx is not None
    is_chief = is_worker and widx == 0
    if is_worker:
        log("Worker %d: %s" % (widx, "Initializing session..." if is_chief
                               else "Waiting for session to be initialized..."))
    transport = choose_transport(flags.transport, topo, flags.sync_replicas, R == topo.num_workers, flags.grad_dtype,
                                 device.index or 0)
    if transport == "nvls":
        from .ps_nvls import NvlsClusterEngine  # NVSwitch multicast (strict sync only)
        eng = NvlsClusterEngine(spec, topo, rank, device, learning_rate=flags.learning_rate, seed=flags.seed,
                                init_stddev=flags.init_stddev, precision=flags.precision)
    else:
        eng = GpuClusterEngine(spec, topo, rank, device, learning_rate=flags.learning_rate, seed=flags.seed,
                               init_stddev=flags.init_stddev, async_mode=not flags.sync_replicas,
                               grad_dtype=flags.grad_dtype, precision=flags.precision,
                               replicas_to_aggregate=R if topo.layout == "dedicated" else 0)
    if flags.log_device_placement and (is_chief or not is_worker):
        log(eng.placement.table())
    sv = Supervisor(is_chief, spec, flags.train_dir, recovery_wait_secs=1.0, log=log)
    start_step = 0
    path = checkpoint.latest_checkpoint(flags.train_dir) if flags.train_dir else None
    if path is not None:  # every task reads the same file; each ps restores the rows it owns
        blob = checkpoint.load(path, spec)
        eng.set_params(blob["variables"])
        if "adam" in blob:
            eng.set_adam_state(blob["adam"])
        eng.set_global_step(blob["global_step"])
        start_step = blob["global_step"]
        dist.barrier()
    if is_worker:
        log("Worker %d: Session initialization complete." % widx)
    steps_left = max(0, flags.train_steps - start_step)
    B = flags.batch_size
    if not is_worker:
        # server.join() (m:160-161): the ps only ever runs its shard's apply kernel until it is released
        if flags.sync_replicas:
            ws = eng.workspace(B)
            for _ in range(steps_left):
                eng.train_step_resident(ws, flags.keep_prob)
            eng.synchronize()
        else:
            ws = eng.workspace(B)
            while not eng.shutdown_requested():
                eng._apply()
                eng.synchronize()
        dist.barrier()
        out = {"role": "ps", "global_step": eng.global_step}
        eng.shutdown()
        return out
    time_begin = time.time()
    log("Training begins @ %f" % time_begin)
    local_step, step = 0, start_step
    last_saved = start_step
    every = flags.save_checkpoint_steps if (flags.train_dir and flags.save_checkpoint_steps > 0) else 0
    from ..utils.trace import StepTracer
    tracer = StepTracer(flags.trace_file and "%s.worker%d" % (flags.trace_file, widx), flags.trace_start,
                        flags.trace_steps, log)  # a no-op unless --trace_file is given
    while step < flags.train_steps:
        batch_xs, batch_ys = mnist.train.next_batch(B)
        _maybe_inject_fault(flags, local_step)
        tracer.before(local_step + 1)
        eng.train_step(batch_xs, batch_ys, flags.keep_prob)
        local_step += 1
        tracer.after(local_step)
        # X5: strict sync knows the global step without asking; async / backup workers read it from shard 0
        step = (start_step + local_step) if (flags.sync_replicas and not eng.backup) else eng.global_step
        log("%f: Worker %d: training step %d done (global step: %d)" % (time.time(), widx, local_step, step))
        if every and is_chief and step // every > last_saved // every and step < flags.train_steps:
            # the Supervisor's periodic saver (m:239-253; every 600 s in TF, here every N global steps).  The chief reads
            # the shards through the peer mapping while the other workers may already be one step further: like TF's
            # saver it snapshots a live model, so a tensor may be one update ahead of another (never torn inside a
            # step on strict sync: the chief's own next push has not happened, so no shard can apply update step+1)
            eng.synchronize()
            sv.save(step, eng.get_params(), eng.get_adam_state())
            last_saved = step
    eng.synchronize()
    tracer.close()
    time_end = time.time()
    log("Training ends @ %f" % time_end)
    log("Training elapsed time: %f s" % (time_end - time_begin))
    accuracy_arr = []
    for _ in range(flags.validation_batches):
        val_x, val_y = mnist.validation.next_batch(flags.validation_batch_size)
        acc = eng.eval_accuracy(val_x, val_y)
        log("After %d training step(s)" % flags.train_steps)
        log("Accuracy : %f" % acc)
        accuracy_arr.append(acc)
    mean_accuracy = sum(accuracy_arr) / len(accuracy_arr)
    log("Mean Accuracy : %f" % mean_accuracy)
    saved = None
    if is_chief and flags.train_dir:
        saved = sv.save(eng.global_step, eng.get_params(), eng.get_adam_state())
    if not flags.sync_replicas:
        # no NCCL while ps GPUs may still be spinning in a serve kernel: finish through peer-memory words
        eng.mark_done()
        if is_chief:
            t0 = time.time()
            while not eng.all_workers_done():
                if time.time() - t0 > flags.step_timeout_secs:
                    raise TimeoutError("chief: workers did not finish within %.0f s" % flags.step_timeout_secs)
                time.sleep(0.01)
            eng.request_shutdown()
    dist.barrier()
    out = {"role": "worker", "task_index": widx, "local_steps": local_step, "global_step": eng.global_step,
           "mean_accuracy": mean_accuracy, "checkpoint": saved}
    eng.shutdown()
    return out
