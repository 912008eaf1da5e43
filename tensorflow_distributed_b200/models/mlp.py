"""The north-star MLP (784 -> hidden x L -> 10; the reference only keeps its docstring m:18 and the dead flag
``hidden_units`` m:66-67) as explicit kernel sequences.  Every hidden layer is one tcgen05 GEMM with a fused
bias+ReLU epilogue; the backward pass runs each layer's wgrad (push-fused on a cluster) on a graph branch while the
main branch continues with the dgrad through the ReLU mask.
"""
from __future__ import annotations

import torch

from ..ops import native as nat


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def forward(eng, ws: Workspace, training: bool) -> None:
    B, dims, tf = ws.B, [784] + list(eng.spec.hidden), eng.tf32
    L = len(dims) - 1
    if not tf:  # tf32: ws.act[0] IS ws.x (the fp32 batch feeds the first GEMM directly)
        nat.cast_bf16(ws.x, ws.act[0], B, 784, 784, 784)
    for l in range(1, L + 1):
        H, K = dims[l], dims[l - 1]
        eng.gemm_fwd("weights/h%d" % l, ws.act[l - 1], B, ws.act[l],
                      mode=nat.EPI_BIAS_RELU_F32 if tf else nat.EPI_BIAS_RELU_BF16, bn=64, bias=eng.P("biases/b%d" % l))
    H = dims[L]
    nat.head_rows(None, ws.act[L], None, eng.P("weights/out"), eng.P("biases/out"), ws.y, B, H, _pad8(H), 1.0,
                  eng.dropout_seed, 0, training, None, ws.dact[L] if training else None, ws.dlogits, ws.loss_row, ws.hit,
                  tf32=tf)


def backward(eng, ws: Workspace) -> None:
    B, dims, tf = ws.B, [784] + list(eng.spec.hidden), eng.tf32
    L = len(dims) - 1
    H = dims[L]
    nat.head_cols(ws.act[L], ws.dact[L], _pad8(H), ws.dlogits, ws.loss_row, ws.hit, B, H, eng.G("weights/out"),
                  eng.G("biases/b%d" % L), eng.G("biases/out"), ws.scalars, tf32=tf)
    # wgrad_l only needs dact[l] and act[l-1]; it runs on a side stream (a graph branch) while the main stream
    # continues with dgrad_l.  On a multi-GPU run that also overlaps the two NVLink directions: wgrad pushes
    # (egress) while dgrad pulls (ingress).
    main = torch.cuda.current_stream()
    s1 = eng._side_streams()[0]
    for l in range(L, 0, -1):
        H, K = dims[l], dims[l - 1]
        s1.wait_stream(main)
        with torch.cuda.stream(s1):
            eng.gemm_wgrad("weights/h%d" % l, ws.act[l - 1], ws.dact[l], B, bn=128 if H >= 128 else 64)
        if l > 1:
            eng.gemm_dgrad("weights/h%d" % l, ws.dact[l], B, ws.dact[l - 1],
                            mode=nat.EPI_RELUMASK_F32 if tf else nat.EPI_RELUMASK_BF16, bn=64, relu_mask=ws.act[l - 1])
            nat.colsum(ws.dact[l - 1], _pad8(K), B, K, eng.G("biases/b%d" % (l - 1)), tf32=tf)
    main.wait_stream(s1)
