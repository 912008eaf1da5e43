"""export.py: the reference's layouts (NHWC / HWIO / [in,out] / h,w,c flatten, m:106-127) <-> stock torch modules."""
import pytest
import torch

from tensorflow_distributed_b200.export import from_torch_module, to_torch_module
from tensorflow_distributed_b200.models.spec import cnn_spec, mlp_spec
from tensorflow_distributed_b200.ops import oracle


@pytest.mark.parametrize("spec", [cnn_spec(), mlp_spec(64, 3)], ids=["cnn", "mlp"])
def test_torch_module_reproduces_the_oracle_logits_and_round_trips(spec):
    p = oracle.init_params(spec, 5, 0.1)
    x = torch.rand(6, 784)
    want = oracle.forward(spec.kind, x, p)
    mod = to_torch_module(spec, p)
    with torch.no_grad():
        assert torch.allclose(mod(x), want, atol=2e-5, rtol=1e-5)
        if spec.kind == "cnn":  # image-shaped inputs are accepted too
            assert torch.allclose(mod(x.reshape(6, 28, 28, 1)), want, atol=2e-5, rtol=1e-5)
    back = from_torch_module(spec, mod)
    assert set(back) == set(p) and all(torch.equal(back[k], p[k]) and back[k].shape == p[k].shape for k in p)


def test_a_torch_trained_module_can_be_loaded_into_an_engine():
    from tensorflow_distributed_b200.engine import make_engine
    spec = mlp_spec(32, 1)
    mod = to_torch_module(spec, oracle.init_params(spec, 1, 0.1)).train()
    opt = torch.optim.SGD(mod.parameters(), lr=0.1)
    x, y = torch.rand(16, 784), torch.eye(10)[torch.randint(0, 10, (16,))]
    for _ in range(3):
        opt.zero_grad()
        oracle.softmax_xent_mean(mod(x), y).backward()
        opt.step()
    eng = make_engine(spec, "cpu", learning_rate=0.01, seed=0, init_stddev=0.1)
    eng.set_params(from_torch_module(spec, mod))
    with torch.no_grad():
        want = float(oracle.softmax_xent_mean(mod.eval()(x), y))
    assert abs(eng.eval_loss(x.numpy(), y.numpy()) - want) < 1e-5
