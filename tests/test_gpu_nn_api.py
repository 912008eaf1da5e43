"""The tensor-in/tensor-out op front end (ops/nn.py) against the PyTorch fp32 oracle."""
import pytest
import torch

from tensorflow_distributed_b200.ops import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def bf(x):
    return x.to(torch.bfloat16)


def rel(a, b):
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


def test_cnn_forward_composed_from_single_ops_matches_oracle():
    from tensorflow_distributed_b200.ops import nn
    torch.manual_seed(0)
    B = 32
    x = torch.rand(B, 784, device=DEV)
    p = {"weights/wc1": nn.random_normal((5, 5, 1, 32), 0.1, 1), "biases/bc1": nn.random_normal((32,), 0.1, 2),
         "weights/wc2": nn.random_normal((5, 5, 32, 64), 0.05, 3), "biases/bc2": nn.random_normal((64,), 0.1, 4),
         "weights/wd1": nn.random_normal((3136, 1024), 0.02, 5), "biases/bd1": nn.random_normal((1024,), 0.1, 6),
         "weights/out": nn.random_normal((1024, 10), 0.05, 7), "biases/out": nn.random_normal((10,), 0.1, 8)}
    y = torch.eye(10, device=DEV)[torch.randint(0, 10, (B,), device=DEV)]
    h1, _ = nn.conv1_block(x, p["weights/wc1"], p["biases/bc1"])
    h2, _ = nn.conv2_block(h1, bf(p["weights/wc2"]), p["biases/bc2"])
    pre = nn.dense(h2.reshape(B, 3136), bf(p["weights/wd1"]))
    head = nn.classifier_head(pre, p["biases/bd1"], p["weights/out"], p["biases/out"], y, keep_prob=1.0, training=True)
    torch.cuda.synchronize()
    pr = dict(p)
    pr["weights/wc2"], pr["weights/wd1"] = bf(p["weights/wc2"]).float(), bf(p["weights/wd1"]).float()
    logits = oracle.cnn_forward(x, pr)
    loss = oracle.softmax_xent_mean(logits, y)
    assert abs(head["loss"].item() - loss.item()) < 3e-2 * max(1.0, loss.item())
    assert abs(head["accuracy"].item() - oracle.accuracy(logits, y).item()) <= 2.0 / B


def test_dense_family_and_adam_update():
    from tensorflow_distributed_b200.ops import nn
    torch.manual_seed(1)
    x, w = bf(torch.randn(128, 256, device=DEV)), bf(torch.randn(256, 512, device=DEV) * 0.1)
    b = torch.randn(512, device=DEV)
    assert rel(nn.dense(x, w), x.float() @ w.float()) < 2e-3
    act = nn.dense(x, w, bias=b, relu=True)
    assert act.dtype == torch.bfloat16 and rel(act, torch.relu(x.float() @ w.float() + b)) < 1e-2
    dy = bf(torch.randn(128, 512, device=DEV))
    assert rel(nn.dense_grad_weight(x, dy), x.float().t() @ dy.float()) < 2e-3
    assert rel(nn.dense_grad_input(dy, w), dy.float() @ w.float().t()) < 2e-3
    masked = nn.dense_grad_input(dy, w, relu_of=x)
    assert rel(masked, (dy.float() @ w.float().t()) * (x.float() > 0)) < 1e-2
    with pytest.raises(ValueError):
        nn.dense(x, bf(torch.randn(100, 64, device=DEV)))
    # Adam
    prm = torch.randn(1024, device=DEV)
    ref = {"w": prm.clone().cpu()}
    st, opt = nn.AdamState(prm), oracle.TFAdam(0.01)
    for i in range(3):
        g = torch.randn(1024, device=DEV)
        nn.adam_update(prm, g, st, lr=0.01)
        opt.apply(ref, {"w": g.cpu()})
    torch.cuda.synchronize()
    assert st.t == 3 and torch.allclose(prm.cpu(), ref["w"], atol=1e-5, rtol=1e-5)


def test_tf32_arm_of_the_single_op_api_matches_the_fp32_oracle():
    """fp32 tensors in -> tf32 arm: the forward pass composed from single ops tracks the PURE fp32 oracle to tf32
    accuracy (the bf16 arm above needs the oracle's weights rounded first), and the dense family returns fp32."""
    from tensorflow_distributed_b200.ops import nn
    torch.manual_seed(2)
    B = 32
    x = torch.rand(B, 784, device=DEV)
    p = {"weights/wc1": nn.random_normal((5, 5, 1, 32), 0.1, 1), "biases/bc1": nn.random_normal((32,), 0.1, 2),
         "weights/wc2": nn.random_normal((5, 5, 32, 64), 0.05, 3), "biases/bc2": nn.random_normal((64,), 0.1, 4),
         "weights/wd1": nn.random_normal((3136, 1024), 0.02, 5), "biases/bd1": nn.random_normal((1024,), 0.1, 6),
         "weights/out": nn.random_normal((1024, 10), 0.05, 7), "biases/out": nn.random_normal((10,), 0.1, 8)}
    y = torch.eye(10, device=DEV)[torch.randint(0, 10, (B,), device=DEV)]
    h1, _ = nn.conv1_block(x, p["weights/wc1"], p["biases/bc1"], out_dtype=torch.float32)
    h2, _ = nn.conv2_block(h1, p["weights/wc2"], p["biases/bc2"])
    pre = nn.dense(h2.reshape(B, 3136), p["weights/wd1"])
    head = nn.classifier_head(pre, p["biases/bd1"], p["weights/out"], p["biases/out"], y, keep_prob=1.0, training=True,
                              act_dtype=torch.float32)
    torch.cuda.synchronize()
    assert h2.dtype == torch.float32 and pre.dtype == torch.float32 and head["hidden"].dtype == torch.float32
    logits = oracle.cnn_forward(x, p)
    loss = oracle.softmax_xent_mean(logits, y)
    assert abs(head["loss"].item() - loss.item()) < 3e-3 * max(1.0, loss.item())
    xw, w = torch.randn(128, 256, device=DEV), torch.randn(256, 512, device=DEV) * 0.1
    dy = torch.randn(128, 512, device=DEV)
    assert rel(nn.dense(xw, w), xw.double() @ w.double()) < 3e-3
    assert rel(nn.dense_grad_weight(xw, dy), xw.double().t() @ dy.double()) < 3e-3
    assert rel(nn.dense_grad_input(dy, w, relu_of=xw), (dy.double() @ w.double().t()) * (xw > 0)) < 3e-3
    with pytest.raises(ValueError):
        nn.dense(xw, bf(w))  # mixed arms
