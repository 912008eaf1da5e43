"""Config 1: single-process trainer on CPU, synthetic batch, the reference's print lines."""
import re

from tensorflow_distributed_b200.trainer import run_single


def test_single_cnn_31_steps_loss_decreases_and_log_lines():
    lines = []
    out = run_single(backend="cpu", synthetic=True, nt=5, val_batch=100, log=lines.append)
    assert out["steps"] == 31  # while step*128 < 4000 (single:109)
    iters = [l for l in lines if l.startswith("Iter ")]
    assert [re.match(r"Iter (\d+), Minibatch Loss= -?\d+\.\d{6}$", l).group(1) for l in iters] == ["1280", "2560", "3840"]
    assert out["losses"][-1] < out["losses"][0]
    assert "Optimization Finished!" in lines
    assert any(re.match(r"--- .* seconds Time for Training ---", l) for l in lines)
    assert any(re.match(r"Mean Accuracy : \d\.\d+", l) for l in lines)
    assert any(re.match(r"--- .* seconds Time for Inference ---", l) for l in lines)


def test_train_flag_zero_is_inference_only_at_chance():
    lines = []
    out = run_single(backend="cpu", synthetic=True, train_flag=0, nt=10, val_batch=100, log=lines.append)
    assert out["steps"] == 0 and "Optimization Finished!" not in lines
    assert out["mean_accuracy"] < 0.35  # random-init weights: about chance (nb:180)


def test_single_mlp_learns_and_resumes(tmp_path):
    kw = dict(model="mlp", backend="cpu", synthetic=True, hidden_units=100, init_stddev=0.05, nt=5,
              training_iters=128 * 41, train_dir=str(tmp_path), log=lambda s: None)
    out = run_single(**kw)
    assert out["mean_accuracy"] > 0.8
    assert out["engine"].global_step == 40
    # resume: picks up global_step from the checkpoint and does no further training steps
    out2 = run_single(**kw)
    assert out2["engine"].global_step == 40 and out2["steps"] == 40


def test_predict_is_consistent_with_eval_accuracy_on_the_oracle_engine():
    import numpy as np
    from tensorflow_distributed_b200 import data
    from tensorflow_distributed_b200.engine import make_engine
    from tensorflow_distributed_b200.models.spec import mlp_spec
    ds = data.read_data_sets("", synthetic=True, seed=3)
    eng = make_engine(mlp_spec(32, 1), "cpu", learning_rate=0.01, seed=0, init_stddev=0.1)
    for _ in range(15):
        eng.train_step(*ds.train.next_batch(64), 1.0)
    x, y = ds.validation.next_batch(200)
    p = eng.predict_proba(x)
    assert p.shape == (200, 10) and np.allclose(p.sum(axis=1), 1.0, atol=1e-5) and (p >= 0).all()
    assert abs(float((eng.predict(x) == y.argmax(axis=1)).mean()) - eng.eval_accuracy(x, y)) < 1e-6
