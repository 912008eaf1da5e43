"""Oracle ops keep TF semantics."""
import math

import torch

from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.ops import oracle


def test_conv_same_matches_naive():
    torch.manual_seed(0)
    x = torch.randn(2, 6, 6, 3)
    w = torch.randn(5, 5, 3, 4)
    y = oracle.conv2d_same_nhwc(x, w)
    ref = torch.zeros(2, 6, 6, 4)
    xp = torch.nn.functional.pad(x, (0, 0, 2, 2, 2, 2))
    for kh in range(5):
        for kw in range(5):
            ref += xp[:, kh:kh + 6, kw:kw + 6, :] @ w[kh, kw]
    assert torch.allclose(y, ref, atol=1e-4)


def test_cnn_shapes_and_flatten_order():
    spec = cnn_spec()
    p = oracle.init_params(spec, 0)
    assert p["weights/wd1"].std().item() > 0.9  # stddev 1 init
    logits = oracle.cnn_forward(torch.rand(4, 784), p)
    assert logits.shape == (4, 10)


def test_dropout_inverted():
    x = torch.ones(2, 4)
    m = torch.tensor([[1, 0, 1, 1], [0, 0, 1, 1]], dtype=torch.bool)
    y = oracle.dropout_with_mask(x, m, 0.75)
    assert torch.allclose(y, m.float() / 0.75)
    assert torch.equal(oracle.dropout_with_mask(x, m, 1.0), x)


def test_tf_adam_formula():
    p = {"w": torch.tensor([1.0, -2.0])}
    g = {"w": torch.tensor([0.5, 0.25])}
    opt = oracle.TFAdam(0.01)
    opt.apply(p, g)
    lr_t = 0.01 * math.sqrt(1 - 0.999) / (1 - 0.9)
    m = 0.1 * g["w"]
    v = 0.001 * g["w"] ** 2
    exp = torch.tensor([1.0, -2.0]) - lr_t * m / (v.sqrt() + 1e-8)
    assert torch.allclose(p["w"], exp, atol=1e-7)
    opt.apply(p, g)
    assert opt.t == 2


def test_softmax_xent_dense_labels_and_accuracy():
    logits = torch.tensor([[2.0, 1.0, 0.0], [0.0, 0.0, 5.0]])
    y = torch.tensor([[1.0, 0, 0], [0, 1.0, 0]])
    ref = torch.nn.functional.cross_entropy(logits, torch.tensor([0, 1]))
    assert torch.allclose(oracle.softmax_xent_mean(logits, y), ref)
    assert oracle.accuracy(logits, y).item() == 0.5


def test_mlp_forward_and_grads():
    spec = mlp_spec(16, 2)
    p = oracle.init_params(spec, 1, 0.1)
    x, y = torch.rand(8, 784), torch.eye(10)[torch.arange(8) % 10]
    loss, grads, logits = oracle.loss_and_grads("mlp", x, y, p)
    assert set(grads) == set(p) and logits.shape == (8, 10) and loss.item() > 0
