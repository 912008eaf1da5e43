"""convolutional_network.ipynb (counterpart of /root/reference/convolutional_network.ipynb): valid nbformat-4 JSON,
no stale outputs, and its code cells run end to end on the CPU oracle with the reference's printed lines."""
import contextlib
import io
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NB = os.path.join(ROOT, "convolutional_network.ipynb")


def _load():
    with open(NB) as f:
        return json.load(f)


def test_notebook_is_clean_nbformat4():
    nb = _load()
    assert nb["nbformat"] == 4
    kinds = [c["cell_type"] for c in nb["cells"]]
    assert kinds.count("code") >= 5 and "markdown" in kinds
    for c in nb["cells"]:
        if c["cell_type"] == "code":
            assert c["outputs"] == [] and c["execution_count"] is None  # the reference ships a stale output (nb:180)


def test_notebook_matches_generator():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_notebook", os.path.join(ROOT, "scripts", "make_notebook.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.NB == _load()


def _run(train_flag: int, iters: int):
    src = []
    for c in _load()["cells"]:
        if c["cell_type"] == "code":
            src.append("".join(c["source"]))
    prog = "\n\n".join(src)
    prog = prog.replace("training_iters = 200000", "training_iters = %d" % iters)
    prog = prog.replace('backend = "auto"', 'backend = "cpu"')
    prog = prog.replace("train_flag = 1 ", "train_flag = %d " % train_flag)
    assert "training_iters = %d" % iters in prog and 'backend = "cpu"' in prog
    out = io.StringIO()
    ns = {"__name__": "__notebook__"}
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        with contextlib.redirect_stdout(out):
            exec(compile(prog, NB, "exec"), ns)
    finally:
        os.chdir(cwd)
    return out.getvalue(), ns


def test_notebook_trains_and_validates_on_cpu():
    text, ns = _run(train_flag=1, iters=128 * 21)
    losses = [float(m) for m in re.findall(r"Iter \d+, Minibatch Loss= (-?\d+\.\d{6})", text)]
    assert len(losses) == 2  # steps 10 and 20
    assert "Optimization Finished!" in text
    assert re.search(r"--- \S+ seconds Time for Training ---", text)
    assert len(re.findall(r"^Accuracy : \d\.\d{6}$", text, flags=re.M)) == 5
    assert re.search(r"^Mean Accuracy : \d\.\d{6}$", text, flags=re.M)
    assert re.search(r"--- \S+ seconds Time for Inference ---", text)
    assert "total parameters: 3274634" in text
    assert ns["mean_accuracy"] > 0.2  # 20 Adam steps on the synthetic set are already far above chance


def test_notebook_train_flag_zero_validates_random_init():
    text, ns = _run(train_flag=0, iters=128 * 21)
    assert "Optimization Finished!" not in text
    assert ns["mean_accuracy"] < 0.3
