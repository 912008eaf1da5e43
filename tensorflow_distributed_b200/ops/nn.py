"""Tensor-in / tensor-out front end of the sm_100a kernels, named after the TF ops the reference calls
(``/root/reference/mnist_python_m.py:93-128,205-208``).  Each function allocates its outputs with torch and launches
one hand-written kernel on the current stream; layouts are the reference's (NHWC activations, HWIO filters,
``[in, out]`` dense weights).  The engines use ``ops/native.py`` directly with preallocated buffers; this module is
the convenient surface for users who want single ops.

    conv2d + bias_add + relu + max_pool (m:93-101)  ->  conv1_block / conv2_block
    matmul (+ add, relu) (m:121-122,127)           ->  dense, dense_grad_input, dense_grad_weight
    softmax_cross_entropy_with_logits + mean (m:205), argmax/equal/mean (m:206-207), dropout (m:124)
                                                    ->  classifier_head
    random_normal (m:186-195)                       ->  random_normal
    AdamOptimizer apply (m:208)                     ->  adam_update

Precision arm = the dtype of what you pass: fp32 tensors run the tf32 arm (fp32 in memory, ``tcgen05 kind::tf32``,
fp32 activations out), bf16 tensors the bf16 arm.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import native as nat


def _need(cond: bool, msg: str) -> None:
    if not cond:
        raise ValueError(msg)


def _cuda(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    _need(t.is_cuda, "%s must be a CUDA tensor" % name)
    _need(t.dtype == dtype, "%s must be %s, got %s" % (name, dtype, t.dtype))
    return t.contiguous()


def _arm(*tensors) -> bool:
    """True = tf32 arm (all fp32), False = bf16 arm (all bf16)."""
    dts = {t.dtype for t in tensors}
    _need(len(dts) == 1 and dts <= {torch.float32, torch.bfloat16},
          "tensor-core operands must all be fp32 (tf32 arm) or all be bf16, got %s" % sorted(map(str, dts)))
    return dts == {torch.float32}


def random_normal(shape, stddev: float = 1.0, seed: int = 0, device="cuda") -> torch.Tensor:
    """tf.random_normal(shape, stddev=stddev): Philox4x32-10 + Box-Muller, deterministic in ``seed``."""
    out = torch.empty(shape, dtype=torch.float32, device=device)
    nat.random_normal(out, seed, stddev)
    return out


def conv1_block(x: torch.Tensor, w_hwio: torch.Tensor, b: torch.Tensor,
                out_dtype=torch.bfloat16) -> Tuple[torch.Tensor, torch.Tensor]:
    """maxpool2d(relu(conv2d(x, w) + b)) for the 5x5, 1 -> 32 layer (CUDA cores, fp32 math).  x fp32 [B,784] or
    [B,28,28,1]; returns (pooled [B,14,14,32] in ``out_dtype`` -- fp32 feeds the tf32 arm --, argmax u8)."""
    x = _cuda(x, torch.float32, "x").reshape(-1, 784)
    w = _cuda(w_hwio, torch.float32, "w").reshape(25, 32)
    b = _cuda(b, torch.float32, "b")
    B = x.shape[0]
    _need(out_dtype in (torch.bfloat16, torch.float32), "out_dtype must be bfloat16 or float32")
    pool = torch.empty(B, 14, 14, 32, dtype=out_dtype, device=x.device)
    am = torch.empty(B, 14, 14, 32, dtype=torch.uint8, device=x.device)
    nat.conv1_fwd(x, w, b, pool, am, B, tf32=out_dtype == torch.float32)
    return pool, am


def conv2_block(x: torch.Tensor, w_hwio: torch.Tensor, b: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """maxpool2d(relu(conv2d(x, w) + b)) for the 5x5, 32 -> 64 layer on tcgen05 (implicit GEMM, TMA im2col).
    x [B,14,14,32], w [5,5,32,64], both bf16 or both fp32; returns (pooled [B,7,7,64] in that type, argmax u8)."""
    tf = _arm(x, w_hwio)
    x = _cuda(x, x.dtype, "x")
    w = _cuda(w_hwio, x.dtype, "w").reshape(800, 64)
    b = _cuda(b, torch.float32, "b")
    _need(tuple(x.shape[1:]) == (14, 14, 32), "x must be [B,14,14,32]")
    B = x.shape[0]
    pool = torch.empty(B, 7, 7, 64, dtype=x.dtype, device=x.device)
    am = torch.empty(B, 7, 7, 64, dtype=torch.uint8, device=x.device)
    nat.conv2_fwd(x, w, b, pool, am, B, tf32=tf)
    return pool, am


def dense(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
          out_dtype=torch.float32) -> torch.Tensor:
    """tf.matmul(x, w) (+ bias) (+ relu) on tcgen05.  x [M,K], w [K,N] ([in,out], read in place as the MN-major
    operand), both bf16 or both fp32 (tf32 arm: the output is always fp32).  bf16 arm: out fp32 (plain) or bf16
    (``relu=True`` or ``out_dtype=torch.bfloat16``)."""
    tf = _arm(x, w)
    x, w = _cuda(x, x.dtype, "x"), _cuda(w, x.dtype, "w")
    _need(x.shape[1] == w.shape[0], "inner dimensions differ: %s x %s" % (tuple(x.shape), tuple(w.shape)))
    _need(x.shape[1] % 8 == 0 and w.shape[1] % 8 == 0, "K and N must be multiples of 8 (16-byte rows for TMA)")
    M, K, N = x.shape[0], x.shape[1], w.shape[1]
    if tf:
        out = torch.empty(M, N, dtype=torch.float32, device=x.device)
        nat.gemm_bf16(x, 0, w, 1, M, N, K, out, mode=nat.EPI_BIAS_RELU_F32 if relu else nat.EPI_STORE_F32,
                      bn=128 if N >= 128 else 64, bias=None if bias is None else _cuda(bias, torch.float32, "bias"),
                      tf32=True)
        return out
    if relu or out_dtype == torch.bfloat16:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
        mode = nat.EPI_BIAS_RELU_BF16 if relu else nat.EPI_STORE_BF16
    else:
        out = torch.empty(M, N, dtype=torch.float32, device=x.device)
        mode = nat.EPI_STORE_F32
    nat.gemm_bf16(x, 0, w, 1, M, N, K, out, mode=mode, bn=128 if N >= 128 else 64,
                  bias=None if bias is None else _cuda(bias, torch.float32, "bias"))
    return out


def dense_grad_input(dy: torch.Tensor, w: torch.Tensor, relu_of: Optional[torch.Tensor] = None) -> torch.Tensor:
    """d(x) of y = x w: dy w^T (w read K-major in place).  With ``relu_of`` (the bf16 activation that fed the layer)
    the gradient is masked by relu'(activation) and returned in bf16; otherwise fp32."""
    tf = _arm(dy, w)
    dy, w = _cuda(dy, dy.dtype, "dy"), _cuda(w, dy.dtype, "w")
    M, N, K = dy.shape[0], dy.shape[1], w.shape[0]
    _need(w.shape[1] == N, "dy [M,N] and w [K,N] disagree on N")
    if tf:
        out = torch.empty(M, K, dtype=torch.float32, device=dy.device)
        if relu_of is not None:
            nat.gemm_bf16(dy, 0, w, 0, M, K, N, out, mode=nat.EPI_RELUMASK_F32, bn=64,
                          relu_mask=_cuda(relu_of, torch.float32, "relu_of"), tf32=True)
        else:
            nat.gemm_bf16(dy, 0, w, 0, M, K, N, out, mode=nat.EPI_STORE_F32, bn=64, tf32=True)
        return out
    if relu_of is not None:
        out = torch.empty(M, K, dtype=torch.bfloat16, device=dy.device)
        nat.gemm_bf16(dy, 0, w, 0, M, K, N, out, mode=nat.EPI_RELUMASK_BF16, bn=64,
                      relu_mask=_cuda(relu_of, torch.bfloat16, "relu_of"))
    else:
        out = torch.empty(M, K, dtype=torch.float32, device=dy.device)
        nat.gemm_bf16(dy, 0, w, 0, M, K, N, out, mode=nat.EPI_STORE_F32, bn=64)
    return out


def dense_grad_weight(x: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    """d(w) of y = x w: x^T dy, contraction over the batch, both operands read MN-major in place.  fp32 [K,N]."""
    tf = _arm(x, dy)
    x, dy = _cuda(x, x.dtype, "x"), _cuda(dy, x.dtype, "dy")
    _need(x.shape[0] == dy.shape[0], "x and dy must have the same batch")
    K, N, Bsz = x.shape[1], dy.shape[1], x.shape[0]
    out = torch.empty(K, N, dtype=torch.float32, device=x.device)
    nat.gemm_bf16(x, 1, dy, 1, K, N, Bsz, out, mode=nat.EPI_STORE_F32, bn=128 if N >= 128 else 64, tf32=tf)
    return out


def classifier_head(pre: torch.Tensor, hid_bias: torch.Tensor, out_w: torch.Tensor, out_b: torch.Tensor,
                    labels: torch.Tensor, keep_prob: float = 1.0, seed: int = 0, step: int = 0, training: bool = True,
                    act_dtype=torch.bfloat16):
    """relu(pre + hid_bias) -> dropout(keep_prob) -> @ out_w + out_b -> mean softmax cross entropy with dense labels
    and accuracy, plus (training) every gradient of that tail.  pre fp32 [B,H]; the hidden activation and its gradient
    come back in ``act_dtype`` (fp32 for the tf32 arm).  Returns a dict."""
    pre = _cuda(pre, torch.float32, "pre")
    _need(act_dtype in (torch.bfloat16, torch.float32), "act_dtype must be bfloat16 or float32")
    tf = act_dtype == torch.float32
    B, H = pre.shape
    dev = pre.device
    hid = torch.empty(B, H, dtype=act_dtype, device=dev)
    dhid = torch.empty(B, H, dtype=act_dtype, device=dev) if training else None
    dlogits = torch.empty(B, 10, device=dev)
    loss_row = torch.empty(B, device=dev)
    hit = torch.empty(B, dtype=torch.uint8, device=dev)
    mask = torch.ones(B, H, dtype=torch.uint8, device=dev)
    nat.head_rows(pre, None, _cuda(hid_bias, torch.float32, "hid_bias"), _cuda(out_w, torch.float32, "out_w"),
                  _cuda(out_b, torch.float32, "out_b"), _cuda(labels, torch.float32, "labels"), B, H, H, keep_prob, seed,
                  step, training, hid, dhid, dlogits, loss_row, hit, mask, tf32=tf)
    scalars = torch.empty(2, device=dev)
    out = {"hidden": hid, "keep_mask": mask, "dlogits": dlogits}
    if training:
        d_out_w, d_hid_bias, d_out_b = torch.empty(H, 10, device=dev), torch.empty(H, device=dev), torch.empty(10, device=dev)
        nat.head_cols(hid, dhid, H, dlogits, loss_row, hit, B, H, d_out_w, d_hid_bias, d_out_b, scalars, tf32=tf)
        out.update(d_hidden=dhid, d_out_w=d_out_w, d_hid_bias=d_hid_bias, d_out_b=d_out_b)
    else:
        nat.head_cols(None, None, H, dlogits, loss_row, hit, B, H, None, None, None, scalars, tf32=tf)
    out["loss"], out["accuracy"] = scalars[0], scalars[1]
    return out


class AdamState:
    """Slots of tf.train.AdamOptimizer for one flat fp32 parameter tensor (length a multiple of 4)."""

    def __init__(self, param: torch.Tensor):
        _need(param.is_cuda and param.dtype == torch.float32 and param.is_contiguous(), "param must be contiguous CUDA fp32")
        _need(param.numel() % 4 == 0, "param length must be a multiple of 4")
        self.m, self.v = torch.zeros_like(param), torch.zeros_like(param)
        self.ctrl = torch.zeros(8, dtype=torch.int32, device=param.device)  # [0] = t

    @property
    def t(self) -> int:
        return int(self.ctrl[0].item())


def adam_update(param: torch.Tensor, grad: torch.Tensor, state: AdamState, lr: float = 0.01, beta1: float = 0.9,
                beta2: float = 0.999, eps: float = 1e-8) -> None:
    """One TF-form Adam step in place (epsilon outside the bias correction, m:208)."""
    _need(grad.shape == param.shape and grad.is_cuda and grad.dtype == torch.float32, "grad must match param")
    n = param.numel()
    nat.set_slot_dtype(False)
    nat.adam_apply(param, state.m, state.v, None, [grad.contiguous().data_ptr()], 1, 1.0, n, [nat.Piece(0, n, 1, 0, 0)], lr,
                   state.ctrl[0:], None, None, None, beta1, beta2, eps)
