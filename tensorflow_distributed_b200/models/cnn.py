"""The reference's CNN (``conv_net``, ``/root/reference/mnist_python_m.py:104-128``) as explicit kernel sequences:
reshape -> conv5x5(1->32)+bias+ReLU -> pool2 -> conv5x5(32->64)+bias+ReLU -> pool2 -> flatten(3136, NHWC order) ->
FC1024+ReLU -> dropout -> FC10, the loss of m:205 and the hand-written backward pass TF's autodiff generates at
m:222.  ``eng`` supplies parameters (``P`` fp32, ``S`` what the tensor cores read: the fp32 master on the tf32 arm, a
bf16 shadow on the bf16 arm), gradient destinations (``G``) and the three GEMM hooks, so the same sequence runs on one
GPU and on a ps/worker cluster (where the hooks address ps shards and ``G`` may be a staging slot on another GPU:
every gradient kernel WRITES its result -- fixed-order reductions, no atomics, nothing to zero).
"""
from __future__ import annotations

import torch

from ..ops import native as nat


def forward(eng, ws: Workspace, training: bool, keep_prob: float, mask_out=None) -> None:
    B, tf = ws.B, eng.tf32
    eng.conv1_fwd(ws, training)  # (on a ps cluster this kernel also waits for the parameter version, locally)
    nat.conv2_fwd(ws.pool1, eng.S("weights/wc2"), eng.P("biases/bc2"), ws.pool2, ws.am2, B, ref=eng.conv2_ref, tf32=tf)
    eng.gemm_fwd("weights/wd1", ws.pool2.view(B, 3136), B, ws.fc1acc, mode=nat.EPI_STORE_F32, bn=64,
                  splits=ws.fc1_splits, split_stride=B * 1024, ld_out=1024)
    nat.head_rows(ws.fc1acc, None, eng.P("biases/bd1"), eng.P("weights/out"), eng.P("biases/out"), ws.y, B,
                  1024, 1024, keep_prob if training else 1.0, eng.dropout_seed, 0, training, ws.hid,
                  ws.dhid if training else None, ws.dlogits, ws.loss_row, ws.hit, mask_out,
                  step_ptr=eng.step_ptr() if training else None, pre_nsplit=ws.fc1_splits,
                  pre_split_stride=B * 1024, tf32=tf)


def backward(eng, ws: Workspace) -> None:
    """Backward as a DAG: the weight-gradient kernels are leaves (nothing downstream reads them before the
    apply), so they run on side streams -- graph branches after capture -- next to the critical path
    head -> fc1 dgrad -> pool2 bwd -> conv2 dgrad -> conv1 wgrad."""
    B, tf = ws.B, eng.tf32
    main = torch.cuda.current_stream()
    s1, s2, s3 = eng._side_streams()
    # Two independent leaves right behind the head's row phase, each on its own branch: the fc1 weight gradient (200
    # CTAs) runs under fc1 dgrad / pool2 backward and is gone before conv2 dgrad needs whole SMs -- queued behind the
    # 14 us head_cols kernel it used to start just in front of conv2 dgrad and delay it by 5 us.
    ready = main.record_event()  # head_rows done: dhid / hid / dlogits exist
    # fc1 dgrad FIRST in capture order (critical path; dpool2[B,3136] = dhid wd1^T, wd1 read K-major in place, split-K
    # partial planes summed by the pool2 / ReLU backward kernel), then the two leaves on their branches
    eng.gemm_dgrad("weights/wd1", ws.dhid, B, ws.dpool2, mode=nat.EPI_STORE_F32, bn=64, splits=ws.dg_splits,
                   split_stride=B * 3136, ld_out=3136)
    s1.wait_event(ready)
    s3.wait_event(ready)
    with torch.cuda.stream(s1):
        # fc1 wgrad: dW[3136,1024] = pool2^T dhid   (both operands MN-major, K = batch)
        eng.gemm_wgrad("weights/wd1", ws.pool2.view(B, 3136), ws.dhid, B, bn=128)
    with torch.cuda.stream(s3):
        nat.head_cols(ws.hid, ws.dhid, 1024, ws.dlogits, ws.loss_row, ws.hit, B, 1024, eng.G("weights/out"),
                      eng.G("biases/bd1"), eng.G("biases/out"), ws.scalars, tf32=tf)
    nat.pool2_bwd(ws.dpool2, ws.pool2, ws.am2, ws.dy2, eng.G("biases/bc2"), ws.p2b_scratch, B, nsplit=ws.dg_splits,
                  split_stride=B * 3136, tf32=tf)
    s2.wait_stream(main)
    with torch.cuda.stream(s2):
        nat.conv2_wgrad(ws.pool1, ws.dy2, eng.G("weights/wc2"), ws.c2w_scratch, B, ref=eng.conv2_ref, tf32=tf)
    nat.conv2_dgrad(ws.dy2, eng.S("weights/wc2"), ws.dpool1, B, ref=eng.conv2_ref, tf32=tf)
    nat.conv1_wgrad(ws.x, ws.dpool1, ws.pool1, ws.am1, eng.G("weights/wc1"), eng.G("biases/bc1"), ws.c1w_scratch, B,
                    tf32=tf, nsplit=ws.dp1_splits, split_stride=B * 14 * 14 * 32)
    main.wait_stream(s1)
    main.wait_stream(s2)
    main.wait_stream(s3)
