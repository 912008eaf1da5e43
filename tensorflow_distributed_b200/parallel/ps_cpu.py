"""CPU parameter server over ``torch.distributed`` (gloo): one OS process per task.

This is the transport for the reference's default topology -- a ps process and
worker processes started separately with ``--job_name/--task_index`` and host
lists (``/root/reference/mnist_python_m.py:145-161``) -- on machines without a
GPU (BASELINE.json configs 1-2).  The reference's transport is TF's gRPC
runtime with implicit Send/Recv; here the pulls, pushes and
tokens are explicit messages:

    worker -> ps   header int64[4] = [op, worker_task, local_step, aux]
    PULL           ps replies  [global_step] + shard arena           (X1)
    PUSH           worker sends shard gradient arena                 (X2 / X3)
    TOKEN          ps replies a token when one is available          (X4)
    STEP           ps replies global_step                            (X5)
    INIT/READY     chief initialises, non-chiefs poll once a second  (X6, m:245)

The ps is passive and single-threaded like ``server.join()`` (m:160-161): it
holds the master weights, Adam slots, accumulators and the token queue and
serves requests until every worker has said DONE or someone sends SHUTDOWN.
"""
from __future__ import annotations

import datetime
import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..cluster import ClusterSpec
from ..models.spec import ModelSpec
from ..ops.oracle import TFAdam
from .placement import Placement
from .replica import AsyncReplicas, SyncReplicas

OP_PULL, OP_PUSH, OP_TOKEN, OP_STEP, OP_INIT, OP_READY, OP_GET_STATE, OP_SET_STATE, OP_DONE, OP_SHUTDOWN, \
    OP_INIT_TOKENS = range(11)


def init_process_group(cluster: ClusterSpec, job: str, index: int, timeout_s: float = 120.0) -> int:
    host, port = cluster.master_endpoint()
    rank = cluster.rank_of(job, index)
    dist.init_process_group("gloo", init_method="tcp://%s:%d" % (host, port), rank=rank,
                            world_size=cluster.world_size, timeout=datetime.timedelta(seconds=timeout_s))
    return rank


class PSServer:
    """State and service loop of one ps task (one shard)."""

    def __init__(self, cluster: ClusterSpec, spec: ModelSpec, placement: Placement, shard: int,
                 sync: bool, replicas_to_aggregate: int, learning_rate: float):
        self.cluster, self.spec, self.placement, self.shard = cluster, spec, placement, shard
        n = placement.shard_sizes[shard]
        self.arena = torch.zeros(n)
        self.adam = TFAdam(learning_rate)
        self.adam.m["arena"] = torch.zeros(n)
        self.adam.v["arena"] = torch.zeros(n)
        self.initialised = False
        self.sync = sync
        self.N = cluster.num_workers
        self.R = replicas_to_aggregate
        self.pending_tokens: List[int] = []  # ranks waiting for a token
        self.done = 0
        self.served = {"pull": 0, "push": 0, "token": 0}
        self._make_protocol(0)

    def _apply(self, grads: Dict[str, torch.Tensor]) -> None:
        self.adam.apply({"arena": self.arena}, grads)

    def _make_protocol(self, global_step: int) -> None:
        if self.sync:
            self.proto = SyncReplicas(["arena"], self.R, self.N, self._apply, global_step)
        else:
            self.proto = AsyncReplicas(self._apply, global_step)

    @property
    def global_step(self) -> int:
        return self.proto.global_step

    def _serve_tokens(self) -> None:
        while self.pending_tokens and self.proto.token_available():
            rank = self.pending_tokens.pop(0)
            dist.send(torch.tensor([self.proto.dequeue_token()], dtype=torch.int64), dst=rank)
            self.served["token"] += 1

    def new_generation(self) -> None:
        """A new set of client processes attaches to this (long-lived) server: variables, Adam slots and global_step
        stay; per-client protocol state (waiting token requests, DONE count, unconsumed tokens) starts afresh."""
        self.pending_tokens = []
        self.done = 0
        self._make_protocol(self.global_step)

    def serve_forever(self) -> str:
        """Serve until every worker of this generation has said DONE ("done") or someone sent SHUTDOWN ("shutdown")."""
        hdr = torch.zeros(4, dtype=torch.int64)
        while True:
            src = dist.recv(hdr)
            op, task, local_step, aux = (int(v) for v in hdr)
            if op == OP_PULL:
                dist.send(torch.tensor([self.global_step], dtype=torch.int64), dst=src)
                dist.send(self.arena, dst=src)
                self.served["pull"] += 1
            elif op == OP_PUSH:
                g = torch.empty_like(self.arena)
                dist.recv(g, src=src)
                self.served["push"] += 1
                if self.sync:
                    self.proto.push({"arena": g}, local_step)
                    while self.proto.try_aggregate():
                        pass
                    self._serve_tokens()
                else:
                    step = self.proto.push(task, {"arena": g})
                    dist.send(torch.tensor([step], dtype=torch.int64), dst=src)
            elif op == OP_TOKEN:
                self.pending_tokens.append(src)
                self._serve_tokens()
            elif op == OP_INIT_TOKENS:
                self.proto.init_tokens(aux)
                self._serve_tokens()
            elif op == OP_STEP:
                dist.send(torch.tensor([self.global_step], dtype=torch.int64), dst=src)
            elif op == OP_INIT:
                dist.recv(self.arena, src=src)
                self.initialised = True
            elif op == OP_READY:
                dist.send(torch.tensor([int(self.initialised)], dtype=torch.int64), dst=src)
            elif op == OP_GET_STATE:
                dist.send(torch.tensor([self.global_step, self.adam.t], dtype=torch.int64), dst=src)
                for t in (self.arena, self.adam.m["arena"], self.adam.v["arena"]):
                    dist.send(t, dst=src)
            elif op == OP_SET_STATE:
                meta = torch.zeros(2, dtype=torch.int64)
                dist.recv(meta, src=src)
                for t in (self.arena, self.adam.m["arena"], self.adam.v["arena"]):
                    dist.recv(t, src=src)
                self.adam.t = int(meta[1])
                self._make_protocol(int(meta[0]))
                self.initialised = True
            elif op == OP_DONE:
                self.done += 1
                if self.done >= self.N:
                    return "done"
            elif op == OP_SHUTDOWN:
                return "shutdown"
            else:
                raise RuntimeError("ps: unknown op %d from rank %d" % (op, src))


class PSClient:
    """Worker-side handle on all ps shards."""

    def __init__(self, cluster: ClusterSpec, spec: ModelSpec, placement: Placement, task_index: int,
                 device_filters: Optional[List[str]] = None):
        self.cluster, self.spec, self.placement, self.task = cluster, spec, placement, task_index
        self.S = placement.num_shards
        self._arenas = [torch.zeros(n) for n in placement.shard_sizes]
        # the session's device filters (m:258): every peer this client talks to is resolved through peer_rank, which
        # refuses a task outside the filters -- a worker cannot address another worker
        self.device_filters = device_filters

    def peer_rank(self, job: str, task_index: int) -> int:
        from ..cluster import filter_allows
        if self.device_filters is not None and not filter_allows(self.device_filters, job, task_index):
            raise PermissionError("/job:%s/task:%d is outside this session's device filters %s"
                                  % (job, task_index, self.device_filters))
        return self.cluster.rank_of(job, task_index)

    def _hdr(self, shard: int, op: int, local_step: int = 0, aux: int = 0) -> None:
        dist.send(torch.tensor([op, self.task, local_step, aux], dtype=torch.int64), dst=self.peer_rank("ps", shard))

    def _recv_i64(self, shard: int, n: int = 1) -> List[int]:
        t = torch.zeros(n, dtype=torch.int64)
        dist.recv(t, src=self.peer_rank("ps", shard))
        return [int(v) for v in t]

    # X1
    def pull(self) -> Dict[str, torch.Tensor]:
        # one shard at a time, request then reply: gloo sends complete only once the peer has posted the
        # matching recv, so a worker must never hold requests open on two ps tasks at once (deadlock cycle
        # worker->ps1->chief->ps0->worker otherwise)
        for s in range(self.S):
            self._hdr(s, OP_PULL)
            self._recv_i64(s)
            dist.recv(self._arenas[s], src=self.peer_rank("ps", s))
        return self.placement.gather(self._arenas)

    # X2 / X3
    def push(self, grads: Dict[str, torch.Tensor], local_step: int, sync: bool) -> Optional[int]:
        flat = [torch.zeros(n) for n in self.placement.shard_sizes]
        self.placement.scatter(grads, flat)
        steps = []
        for s in range(self.S):
            self._hdr(s, OP_PUSH, local_step)
            dist.send(flat[s], dst=self.peer_rank("ps", s))
            if not sync:
                steps.append(self._recv_i64(s)[0])
        return None if sync else steps[self.placement.global_step_shard]

    # X4
    def dequeue_token(self) -> int:
        s = self.placement.global_step_shard
        self._hdr(s, OP_TOKEN)
        return self._recv_i64(s)[0]

    # X5
    def global_step(self) -> int:
        s = self.placement.global_step_shard
        self._hdr(s, OP_STEP)
        return self._recv_i64(s)[0]

    # X6
    def init_variables(self, params: Dict[str, torch.Tensor]) -> None:
        flat = [torch.zeros(n) for n in self.placement.shard_sizes]
        self.placement.scatter(params, flat)
        for s in range(self.S):
            self._hdr(s, OP_INIT)
            dist.send(flat[s], dst=self.peer_rank("ps", s))

    def init_tokens(self, num_tokens: int) -> None:
        self._hdr(self.placement.global_step_shard, OP_INIT_TOKENS, aux=num_tokens)

    def ready(self) -> bool:
        ok = True
        for s in range(self.S):
            self._hdr(s, OP_READY)
            ok = (self._recv_i64(s)[0] == 1) and ok
        return ok

    def wait_ready(self, recovery_wait_secs: float = 1.0, timeout_s: float = 120.0) -> None:
        t0 = time.time()
        while not self.ready():
            if time.time() - t0 > timeout_s:
                raise TimeoutError("worker %d: model not initialised after %.0fs" % (self.task, timeout_s))
            time.sleep(recovery_wait_secs)

    def get_state(self) -> dict:
        arenas, ms, vs, meta = [], [], [], []
        for s in range(self.S):
            self._hdr(s, OP_GET_STATE)
            meta.append(self._recv_i64(s, 2))
            trio = [torch.zeros(self.placement.shard_sizes[s]) for _ in range(3)]
            for t in trio:
                dist.recv(t, src=self.peer_rank("ps", s))
            arenas.append(trio[0]); ms.append(trio[1]); vs.append(trio[2])
        g = self.placement.global_step_shard
        return {"global_step": meta[g][0], "t": meta[g][1], "params": self.placement.gather(arenas),
                "m": self.placement.gather(ms), "v": self.placement.gather(vs)}

    def set_state(self, global_step: int, t: int, params, m, v) -> None:
        trios = []
        for src in (params, m, v):
            flat = [torch.zeros(n) for n in self.placement.shard_sizes]
            self.placement.scatter(src, flat)
            trios.append(flat)
        for s in range(self.S):
            self._hdr(s, OP_SET_STATE)
            dist.send(torch.tensor([global_step, t], dtype=torch.int64), dst=self.peer_rank("ps", s))
            for k in range(3):
                dist.send(trios[k][s], dst=self.peer_rank("ps", s))

    def done(self) -> None:
        for s in range(self.S):
            self._hdr(s, OP_DONE)

    def shutdown(self) -> None:
        for s in range(self.S):
            self._hdr(s, OP_SHUTDOWN)
