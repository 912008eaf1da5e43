"""NativeEngine: the sm_100a execution engine (single GPU: one ps shard + one worker on the same device).

What the reference delegates to ``tf.Session.run`` on a graph TF differentiated for it
(``/root/reference/mnist_single.py:92-112``, ``/root/reference/mnist_python_m.py:202-222``) is spelled out here
as an explicit forward + hand-written backward over the kernels of ``libtfd_b200.so``, captured once per
(batch, mode) in a CUDA graph and replayed: one graph launch per training step, no host round trip inside.

Precision arms (the reference is fp32 end to end, m:198-200).  Both keep fp32 master weights, Adam
slots, gradients and accumulators (TMEM) on the ps shard; they differ in what the tensor cores read:

* ``tf32`` (default): activations and weights stay fp32 in memory and are read as tf32 by ``tcgen05 kind::tf32``
  (10-bit mantissa products, fp32 accumulation); no weight shadow exists.  This is the arm whose numbers may be set
  against an fp32 reference.
* ``bf16``: bf16 activations and a bf16 shadow of the tensor-core weights that the ps republishes from inside the
  Adam kernel: half the operand bytes.
"""
from __future__ import annotations

import os

from typing import Dict, List, Optional

import numpy as np
import torch

from .models.spec import ModelSpec
from .ops import native as nat
from .parallel.placement import Placement

CTRL_T, CTRL_SCRATCH, CTRL_VERSION, CTRL_GSTEP, CTRL_ARRIVE, CTRL_PUSHBLK = 0, 1, 2, 3, 4, 5


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


PRECISIONS = ("tf32", "bf16")


def default_precision() -> str:
    p = os.environ.get("TFD_PRECISION", "tf32")
    if p not in PRECISIONS:
        raise ValueError("TFD_PRECISION must be one of %s" % (PRECISIONS,))
    return p


def tc_operand(spec: ModelSpec, name: str) -> bool:
    """Weights consumed by tensor-core kernels (everything else is read in fp32 by CUDA-core kernels)."""
    if spec.kind == "cnn":
        return name in ("weights/wc2", "weights/wd1")
    return name.startswith("weights/h")


def needs_shadow(spec: ModelSpec, name: str, precision: str = "bf16") -> bool:
    """bf16 arm: tensor-core weights get a bf16 shadow.  tf32 arm: the kernels read the fp32 master, no shadow."""
    return precision == "bf16" and tc_operand(spec, name)


class ShardState:
    """One parameter-server shard in HBM: fp32 arena + Adam slots + bf16 shadow + control words.

    ``alloc(name, numel, dtype)`` lets the multi-GPU runtime carve w / m / v / shadow / ctrl out of a
    peer-mapped (cudaIpc) heap; by default they are ordinary torch allocations.
    """

    def __init__(self, spec: ModelSpec, placement: Placement, shard: int, device, alloc=None, precision: str = "bf16"):
        self.spec, self.placement, self.shard, self.device = spec, placement, shard, device
        self.precision = precision
        self.n = placement.shard_sizes[shard]
        self.pieces, self.shadow_off, self.shadow_ld, self.shadow_n = self.layout(spec, placement, shard, precision)
        alloc = alloc or (lambda name, numel, dtype: torch.zeros(numel, dtype=dtype, device=device))
        self.w = alloc("w", self.n, torch.float32)
        self.m = alloc("m", self.n, torch.float32)
        self.v = alloc("v", self.n, torch.float32)
        self.shadow = alloc("shadow", self.shadow_n, torch.bfloat16)
        self.ctrl = alloc("ctrl", 64, torch.int32)

    @staticmethod
    def layout(spec: ModelSpec, placement: Placement, shard: int, precision: str = "bf16"):
        pieces: List[nat.Piece] = []
        shadow_off: Dict[str, int] = {}
        shadow_ld: Dict[str, int] = {}
        off = 0
        for p in placement.shard_pieces(shard):
            if needs_shadow(spec, p.var, precision):
                ld = _pad8(p.cols)
                pieces.append(nat.Piece(p.offset, p.numel, p.cols, ld, off))
                shadow_off[p.var], shadow_ld[p.var] = off, ld
                off += (p.row_end - p.row_start) * ld
                off = (off + 63) // 64 * 64
            else:
                pieces.append(nat.Piece(p.offset, p.numel, p.cols, 0, 0))
        return pieces, shadow_off, shadow_ld, max(off, 64)

    def piece_of(self, var: str):
        for p in self.placement.pieces[var]:
            if p.shard == self.shard:
                return p
        return None

    def view(self, arena: torch.Tensor, var: str) -> torch.Tensor:
        p = self.piece_of(var)
        return arena[p.offset:p.offset + p.numel].view(p.row_end - p.row_start, p.cols)

    def shadow_view(self, var: str) -> torch.Tensor:
        p = self.piece_of(var)
        ld = self.shadow_ld[var]
        o = self.shadow_off[var]
        return self.shadow[o:o + (p.row_end - p.row_start) * ld].view(p.row_end - p.row_start, ld)

    def publish_shadow(self) -> None:
        """(Re)build the bf16 shadow from the fp32 arena (init / restore; steady state is done by the Adam kernel)."""
        for var in self.shadow_off:
            p = self.piece_of(var)
            nat.cast_bf16(self.view(self.w, var), self.shadow_view(var), p.row_end - p.row_start, p.cols,
                          p.cols, self.shadow_ld[var])

    def init_variables(self, seed: int, stddev: float) -> None:
        """tf.global_variables_initializer() for the pieces this shard owns: every parameter ~ N(0, stddev).
        The Philox stream is indexed by (variable, global element), so the values do not depend on the sharding."""
        for i, v in enumerate(self.spec.variables):
            p = self.piece_of(v.name)
            if p is not None:
                nat.random_normal(self.view(self.w, v.name), (seed << 8) + i, stddev, p.row_start * p.cols)
        self.m.zero_(); self.v.zero_(); self.ctrl.zero_()
        self.publish_shadow()


class Workspace:
    """Activations / gradients-of-activations for one batch size (all static: graph-capturable)."""

    def __init__(self, spec: ModelSpec, B: int, device, precision: str = "bf16"):
        self.B = B
        tf32 = precision == "tf32"
        adt = torch.float32 if tf32 else torch.bfloat16  # activation storage type of the precision arm
        z = lambda *s, dtype=torch.float32: torch.zeros(*s, dtype=dtype, device=device)  # noqa: E731
        self.x = z(B, 784)
        self.y = z(B, 10)
        self.stage = [(torch.zeros(B, 784, dtype=torch.float32).pin_memory(),
                       torch.zeros(B, 10, dtype=torch.float32).pin_memory(), torch.cuda.Event()) for _ in range(2)]
        self.stage_i = 0
        self.scalars = z(2)
        self.scalars_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        self.dlogits, self.loss_row = z(B, 10), z(B)
        self.hit = z(B, dtype=torch.uint8)
        if spec.kind == "cnn":
            self.pool1 = z(B, 14, 14, 32, dtype=adt)
            self.am1 = z(B, 14, 14, 32, dtype=torch.uint8)
            self.pool2 = z(B, 7, 7, 64, dtype=adt)
            self.am2 = z(B, 7, 7, 64, dtype=torch.uint8)
            kb = 3136 // (32 if tf32 else 64)  # K blocks of the fc1 forward GEMM
            self.fc1_splits = max(1, min(8, kb, 148 // (((B + 127) // 128) * 16)))
            self.fc1acc = z(self.fc1_splits, B, 1024)   # split-K partial planes (summed by the head kernel)
            self.hid = z(B, 1024, dtype=adt)
            self.dhid = z(B, 1024, dtype=adt)
            self.dg_splits = max(1, min(3, 148 // (((B + 127) // 128) * 49)))
            self.dpool2 = z(self.dg_splits, B, 3136)    # split-K partial planes (summed by pool2_bwd)
            self.dy2 = z(B, 14, 14, 64, dtype=adt)
            self.dp1_splits = 2 if tf32 else 1          # tf32 conv2 dgrad: one partial plane per half of dy's channels
            self.dpool1 = z(self.dp1_splits, B, 14, 14, 32)
            # reduction scratch of the deterministic gradient kernels (zeroed once; the kernels own the counters)
            self.c1w_scratch = z(nat.scratch_floats("conv1_wgrad", B))
            self.c2w_scratch = z(nat.scratch_floats("conv2_wgrad", B))
            self.p2b_scratch = z(nat.scratch_floats("pool2_bwd", B))
            self.keep_mask = None
        else:
            dims = [784] + list(spec.hidden)
            # tf32: the fp32 input batch IS the first GEMM's A operand; bf16: it is cast once per step
            self.act = [self.x if tf32 else z(B, _pad8(784), dtype=adt)] + [z(B, _pad8(d), dtype=adt) for d in dims[1:]]
            self.dact = [None] + [z(B, _pad8(d), dtype=adt) for d in dims[1:]]
        self.graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}


class NativeEngine:
    """Single-GPU engine: parameters live in one ShardState on the same device as the worker kernels."""

    name = "native-sm100a"

    def __init__(self, spec: ModelSpec, learning_rate: float = 0.01, seed: int = 0, init_stddev: float = 1.0,
                 device: Optional[str] = None, use_graphs: bool = True, conv2_ref: bool = False,
                 precision: Optional[str] = None):
        if not torch.cuda.is_available():
            raise nat.NativeError("NativeEngine needs a CUDA device (backend=cpu selects the oracle engine)")
        nat.lib()
        self.spec = spec
        self._set_precision(precision, conv2_ref)
        self.device = torch.device(device or ("cuda:%d" % torch.cuda.current_device()))
        torch.cuda.set_device(self.device)
        nat.set_device(self.device.index or 0)
        nat.set_slot_dtype(False)
        self.lr, self.seed = learning_rate, seed
        self.dropout_seed = seed
        self.use_graphs, self.conv2_ref = use_graphs, conv2_ref
        self.placement = Placement(spec, 1)
        self.shard = ShardState(spec, self.placement, 0, self.device, precision=self.precision)
        self.grad = torch.zeros(self.shard.n, dtype=torch.float32, device=self.device)
        self.slot_ptrs = [self.grad.data_ptr()]
        self.ws: Dict[int, Workspace] = {}
        self.init_variables(seed, init_stddev)

    def _set_precision(self, precision: Optional[str], conv2_ref: bool = False) -> None:
        precision = precision or default_precision()
        if precision not in PRECISIONS:
            raise ValueError("precision must be one of %s, got %r" % (PRECISIONS, precision))
        if conv2_ref and precision != "bf16":
            raise ValueError("conv2_ref (CUDA-core validation kernels) exists for the bf16 arm only")
        self.precision, self.tf32 = precision, precision == "tf32"
        self.dtype_name = "tf32" if self.tf32 else "bf16"
        if self.tf32:
            for v in self.spec.variables:
                if tc_operand(self.spec, v.name) and v.cols % 4:
                    raise ValueError("tf32 arm: %s has %d columns; fp32 tensor-core operands need a multiple of 4"
                                     % (v.name, v.cols))

    # ---- variables -------------------------------------------------------------------------
    def init_variables(self, seed: int, stddev: float) -> None:
        """tf.global_variables_initializer(): every parameter ~ N(0, stddev) from the device Philox stream."""
        self.shard.init_variables(seed, stddev)
        torch.cuda.synchronize(self.device)

    @property
    def global_step(self) -> int:
        return int(self.shard.ctrl[CTRL_GSTEP].item())

    @global_step.setter
    def global_step(self, v: int) -> None:
        self.shard.ctrl[CTRL_GSTEP] = int(v)

    def P(self, name: str) -> torch.Tensor:
        return self.shard.view(self.shard.w, name)

    def S(self, name: str) -> torch.Tensor:
        """What the tensor cores read for weight `name`: its bf16 shadow, or (tf32 arm) the fp32 master itself."""
        return self.P(name) if self.tf32 else self.shard.shadow_view(name)

    def G(self, name: str) -> torch.Tensor:
        return self.shard.view(self.grad, name)

    def get_params(self) -> Dict[str, torch.Tensor]:
        return {v.name: self.P(v.name).detach().cpu().reshape(v.shape).clone() for v in self.spec.variables}

    def set_params(self, params: Dict[str, torch.Tensor]) -> None:
        for v in self.spec.variables:
            self.P(v.name).copy_(params[v.name].to(self.device, torch.float32).reshape(self.P(v.name).shape))
        self.shard.publish_shadow()
        torch.cuda.synchronize(self.device)

    def get_grads(self) -> Dict[str, torch.Tensor]:
        return {v.name: self.G(v.name).detach().cpu().reshape(v.shape).clone() for v in self.spec.variables}

    def get_adam_state(self) -> dict:
        s = self.shard
        t = int(s.ctrl[CTRL_T].item())
        return {"t": t, "beta1_power": 0.9 ** t, "beta2_power": 0.999 ** t,
                "m": {v.name: s.view(s.m, v.name).cpu().reshape(v.shape).clone() for v in self.spec.variables},
                "v": {v.name: s.view(s.v, v.name).cpu().reshape(v.shape).clone() for v in self.spec.variables}}

    def set_adam_state(self, sd: dict) -> None:
        s = self.shard
        for v in self.spec.variables:
            s.view(s.m, v.name).copy_(sd["m"][v.name].to(self.device).reshape(s.view(s.m, v.name).shape))
            s.view(s.v, v.name).copy_(sd["v"][v.name].to(self.device).reshape(s.view(s.v, v.name).shape))
        s.ctrl[CTRL_T] = int(sd["t"])

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)

    def workspace(self, B: int) -> Workspace:
        if B not in self.ws:
            self.ws[B] = Workspace(self.spec, B, self.device, self.precision)
        return self.ws[B]

    # ---- GEMM hooks: the multi-GPU worker overrides these to address ps shards over NVLink ----------
    def gemm_fwd(self, wname: str, A, M: int, out, *, mode: int, bn: int, splits: int = 1, split_stride: int = 0,
                 bias=None, ld_out: Optional[int] = None) -> None:
        """out[M,N] = A[M,K] W[K,N]  (A K-major, W MN-major as stored)."""
        W = self.S(wname)
        K, N = W.shape[0], self.spec.by_name()[wname].cols
        nat.gemm_bf16(A, 0, W, 1, M, N, K, out, mode=mode, bn=bn, splits=splits, split_stride=split_stride, bias=bias,
                      ld_out=ld_out, tf32=self.tf32)

    def gemm_dgrad(self, wname: str, dY, M: int, out, *, mode: int, bn: int, splits: int = 1, split_stride: int = 0,
                   relu_mask=None, ld_out: Optional[int] = None) -> None:
        """out[M,K] = dY[M,N] W^T  (W [K,N] read in place as the K-major B operand [N'=K, K'=N])."""
        W = self.S(wname)
        K, N = W.shape[0], self.spec.by_name()[wname].cols
        nat.gemm_bf16(dY, 0, W, 0, M, K, N, out, mode=mode, bn=bn, splits=splits, split_stride=split_stride,
                      relu_mask=relu_mask, ld_out=ld_out, tf32=self.tf32)

    def gemm_wgrad(self, wname: str, X, dY, Bsz: int, *, bn: int) -> None:
        """dW[K,N] = X^T dY  (both MN-major, contraction over the batch); written where the gradient lives."""
        G = self.G(wname)
        nat.gemm_bf16(X, 1, dY, 1, G.shape[0], G.shape[1], Bsz, G, mode=nat.EPI_STORE_F32, bn=bn, tf32=self.tf32)

    def step_ptr(self):
        """Device word that advances once per training step (dropout stream, version gates)."""
        return self.shard.ctrl[CTRL_T:]

    def _pre_step(self, ws, training: bool) -> None:
        pass  # multi-GPU worker: pull small parameters / wait for the parameter version

    def _post_backward(self, ws) -> None:
        pass  # multi-GPU worker: push small gradients, signal arrival

    # ---- kernel sequences ------------------------------------------------------------------
    def _splits(self, mtiles: int, ntiles: int, kblocks: int, cap: int = 16) -> int:
        return max(1, min(cap, kblocks, 148 // max(1, mtiles * ntiles)))

    def _cnn_forward(self, ws: Workspace, training: bool, keep_prob: float, mask_out=None) -> None:
        from .models import cnn
        cnn.forward(self, ws, training, keep_prob, mask_out)

    _prio_default = "1"  # critical path (conv2 dgrad) ahead of the weight-gradient branches when both are ready

    def conv1_fwd(self, ws: "Workspace", training: bool = True) -> None:
        nat.conv1_fwd(ws.x, self.P("weights/wc1"), self.P("biases/bc1"), ws.pool1, ws.am1, ws.B, tf32=self.tf32)

    def _capture_stream(self):
        """Stream the step graph is captured on.  With TFD_PRIO=1 it has the highest priority, so the kernels of the
        critical path (everything on the main branch) get SMs before the weight-gradient kernels on the side
        branches (default priority) whenever both are ready and cannot be co-resident."""
        if not hasattr(self, "_cap"):
            prio = -5 if os.environ.get("TFD_PRIO", self._prio_default) == "1" else 0
            self._cap = torch.cuda.Stream(device=self.device, priority=prio)
        return self._cap

    def _side_streams(self):
        if not hasattr(self, "_sides"):
            self._sides = [torch.cuda.Stream(device=self.device) for _ in range(3)]
        return self._sides

    def _cnn_backward(self, ws: Workspace) -> None:
        from .models import cnn
        cnn.backward(self, ws)

    def _mlp_forward(self, ws: Workspace, training: bool) -> None:
        from .models import mlp
        mlp.forward(self, ws, training)

    def _mlp_backward(self, ws: Workspace) -> None:
        from .models import mlp
        mlp.backward(self, ws)

    def _apply(self) -> None:
        s = self.shard
        nat.adam_apply(s.w, s.m, s.v, s.shadow if not self.tf32 else None, self.slot_ptrs, 1, 1.0, s.n, s.pieces, self.lr,
                       s.ctrl[CTRL_T:], None, s.ctrl[CTRL_VERSION:], s.ctrl[CTRL_GSTEP:])

    def _sequence(self, ws: Workspace, mode: str, keep_prob: float, mask_out=None) -> None:
        training = mode in ("train", "grads")
        self._training_apply = (mode == "train")  # "grads" computes gradients without taking part in an update
        self._pre_step(ws, training)
        if self.spec.kind == "cnn":
            self._cnn_forward(ws, training, keep_prob, mask_out)
            if training:
                self._cnn_backward(ws)
        else:
            self._mlp_forward(ws, training)
            if training:
                self._mlp_backward(ws)
        if training:
            self._post_backward(ws)
        if not training:
            nat.head_cols(None, None, 0, ws.dlogits, ws.loss_row, ws.hit, ws.B, 1, None, None, None, ws.scalars,
                          tf32=self.tf32)
        if mode == "train":
            self._apply()

    def _run(self, ws: Workspace, mode: str, keep_prob: float = 1.0, mask_out=None) -> None:
        """Launch one step: eager the first time (sets kernel attributes), then capture + replay a CUDA graph."""
        key = (mode, round(float(keep_prob), 6))
        if not self.use_graphs or mask_out is not None:
            self._sequence(ws, mode, keep_prob, mask_out)
            return
        g = ws.graphs.get(key)
        if g is None:
            if key not in getattr(ws, "_warm", set()):
                ws._warm = getattr(ws, "_warm", set()) | {key}
                self._sequence(ws, mode, keep_prob)
                return
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            before = nat.launch_count
            with torch.cuda.graph(g, stream=self._capture_stream()):
                self._sequence(ws, mode, keep_prob)
            ws.graphs[key] = g
            ws.graph_kernels = getattr(ws, "graph_kernels", {})
            ws.graph_kernels[key] = nat.launch_count - before
            nat.launch_count = before  # capture records, it does not execute
        g.replay()
        nat.launch_count += ws.graph_kernels[key]

    # ---- public API: the fetches the reference asks a session for ---------------------------
    def load_batch(self, ws: Workspace, x, y) -> None:
        """feed_dict: host numpy -> pinned staging -> device (X7 in SURVEY.md 2.4)."""
        if isinstance(x, torch.Tensor) and x.is_cuda:
            ws.x.copy_(x.reshape(ws.B, 784)); ws.y.copy_(y.reshape(ws.B, 10))
            return
        # Two pinned staging pairs, each guarded by an event recorded behind its H2D copies: the host may run ahead of
        # the device (train_step does not sync unless a loss is fetched), so a pair is only overwritten once the
        # copies that last read it have completed.  Without this a fast caller tears or duplicates batches.
        k = ws.stage_i
        ws.stage_i ^= 1
        xh, yh, ev = ws.stage[k]
        ev.synchronize()
        xh.copy_(torch.as_tensor(np.ascontiguousarray(x) if isinstance(x, np.ndarray) else x).reshape(ws.B, 784))
        yh.copy_(torch.as_tensor(np.ascontiguousarray(y) if isinstance(y, np.ndarray) else y).reshape(ws.B, 10))
        ws.x.copy_(xh, non_blocking=True)
        ws.y.copy_(yh, non_blocking=True)
        ev.record(torch.cuda.current_stream())

    def _fetch_scalars(self, ws: Workspace):
        ws.scalars_host.copy_(ws.scalars, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(ws.scalars_host[0]), float(ws.scalars_host[1])

    def train_step(self, x, y, keep_prob: float = 0.75, fetch_loss: bool = False):
        """sess.run(train_op): forward, backward, Adam apply.  Returns the loss only when asked (a D2H sync)."""
        ws = self.workspace(len(x))
        self.load_batch(ws, x, y)
        self._run(ws, "train", keep_prob)
        return self._fetch_scalars(ws)[0] if fetch_loss else None

    def train_step_resident(self, ws: Workspace, keep_prob: float = 0.75) -> None:
        """Step on the batch already resident in ws.x / ws.y (device-timed benchmark loop)."""
        self._run(ws, "train", keep_prob)

    def eval_loss(self, x, y) -> float:
        ws = self.workspace(len(x))
        self.load_batch(ws, x, y)
        self._run(ws, "eval", 1.0)
        return self._fetch_scalars(ws)[0]

    def eval_accuracy(self, x, y) -> float:
        ws = self.workspace(len(x))
        self.load_batch(ws, x, y)
        self._run(ws, "eval", 1.0)
        return self._fetch_scalars(ws)[1]

    def predict_proba(self, x) -> np.ndarray:
        """Serving call: class probabilities softmax(model(x)) for a host batch, [B, 10] float32 (dropout off).

        Runs the captured inference graph (forward + fused head) with a dummy one-hot label: the head kernel leaves
        ``(softmax * sum(y) - y) / B`` in ``dlogits``, from which the probabilities follow without another kernel."""
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float32).reshape(-1, 784))
        B = len(x)
        y = np.zeros((B, 10), dtype=np.float32)
        y[:, 0] = 1.0
        ws = self.workspace(B)
        self.load_batch(ws, x, y)
        self._run(ws, "eval", 1.0)
        return (ws.dlogits * float(B) + ws.y).cpu().numpy()  # the D2H copy synchronises the stream

    def predict(self, x) -> np.ndarray:
        """argmax class per image (``tf.argmax(pred, 1)``, m:206)."""
        return self.predict_proba(x).argmax(axis=1)

    def loss_and_grads(self, x, y, keep_prob: float = 1.0):
        """Forward + backward without the apply (tests; cluster workers).  Returns (loss, grads, keep_mask)."""
        ws = self.workspace(len(x))
        self.load_batch(ws, x, y)
        mask = None
        if self.spec.kind == "cnn" and keep_prob < 1.0:
            mask = torch.ones(ws.B, 1024, dtype=torch.uint8, device=self.device)
        self._sequence(ws, "grads", keep_prob, mask)
        loss, _ = self._fetch_scalars(ws)
        return loss, self.get_grads(), mask

    def train_loop(self, batches, keep_prob: float = 0.75, fetch_loss: bool = True) -> List[float]:
        """Public streaming API: run one training step per (x, y) host batch, end to end.

        Every step's inputs go pinned-host -> device (H2D on a copy stream, double-buffered so the copy of step i+1
        runs under the compute of step i) and every step's loss comes back device -> pinned host (D2H, read by the
        host two steps late so it never stalls the device).  Returns the per-step losses."""
        batches = iter(batches)
        first = next(batches, None)
        if first is None:
            return []
        B = len(first[0])
        if not getattr(self, "is_worker", True):  # ps-only GPU of a dedicated layout: one shard apply per step
            ws, n = self.workspace(B), 1
            self._run(ws, "train", keep_prob)
            for _ in batches:
                self._run(ws, "train", keep_prob)
                n += 1
            return [0.0] * n
        ws = self.workspace(B)
        dev = self.device
        if not hasattr(ws, "_slots"):
            ws._slots = [(torch.zeros(B, 784, device=dev), torch.zeros(B, 10, device=dev),
                          torch.zeros(B, 784).pin_memory(), torch.zeros(B, 10).pin_memory()) for _ in range(2)]
            ws._copy_stream = torch.cuda.Stream(device=dev)
            ws._copied = [torch.cuda.Event() for _ in range(2)]
            ws._consumed = [torch.cuda.Event() for _ in range(2)]
            ws._loss_ring = torch.zeros(4, 2).pin_memory()
            ws._loss_ev = [torch.cuda.Event() for _ in range(4)]
        main = torch.cuda.current_stream()
        losses: List[float] = []
        pending: List[int] = []
        i, item = 0, first
        for k in range(2):
            ws._consumed[k].record(main)
        while item is not None:
            x, y = item
            xd, yd, xh, yh = ws._slots[i & 1]
            # host -> pinned staging (the slot is free once its previous H2D has been issued and consumed)
            ws._consumed[i & 1].synchronize()
            xh.copy_(torch.as_tensor(np.ascontiguousarray(x) if isinstance(x, np.ndarray) else x).reshape(B, 784))
            yh.copy_(torch.as_tensor(np.ascontiguousarray(y) if isinstance(y, np.ndarray) else y).reshape(B, 10))
            with torch.cuda.stream(ws._copy_stream):
                xd.copy_(xh, non_blocking=True)
                yd.copy_(yh, non_blocking=True)
                ws._copied[i & 1].record(ws._copy_stream)
            main.wait_event(ws._copied[i & 1])
            ws.x.copy_(xd, non_blocking=True)
            ws.y.copy_(yd, non_blocking=True)
            ws._consumed[i & 1].record(main)
            self._run(ws, "train", keep_prob)
            if fetch_loss:
                if len(pending) == 4:  # ring full: the oldest loss has long been written
                    j = pending.pop(0)
                    ws._loss_ev[j & 3].synchronize()
                    losses.append(float(ws._loss_ring[j & 3, 0]))
                ws._loss_ring[i & 3].copy_(ws.scalars, non_blocking=True)
                ws._loss_ev[i & 3].record(main)
                pending.append(i)
            i += 1
            item = next(batches, None)
        for j in pending:
            ws._loss_ev[j & 3].synchronize()
            losses.append(float(ws._loss_ring[j & 3, 0]))
        return losses
