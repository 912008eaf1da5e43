"""Writes convolutional_network.ipynb (nbformat 4): the interactive counterpart of mnist_single.py, like
/root/reference/convolutional_network.ipynb is for the reference's single-process trainer (nb:61-69 parameters,
nb:187-229 train / validate / timing).  The cells drive this framework's engine API step by step instead of a TF
graph; tests/test_notebook.py executes them on the CPU oracle."""
import json
import os

CELLS = []


def md(text):
    CELLS.append({"cell_type": "markdown", "metadata": {}, "source": text.strip("\n").splitlines(True)})


def code(text):
    CELLS.append({"cell_type": "code", "execution_count": None, "metadata": {}, "outputs": [],
                  "source": text.strip("\n").splitlines(True)})


md("""
# MNIST convolutional network on the B200 parameter-server engine

Two 5x5 convolutions (1 -> 32 -> 64 channels, ReLU, 2x2 max-pool each), a 1024-unit dense layer with dropout and a
10-way softmax, trained with TF-form Adam. With `backend = "cuda"` every op below is a hand-written sm_100a kernel and
the whole step is one CUDA-graph launch; with `backend = "cpu"` the same API runs the fp32 PyTorch oracle.
""")

code("""
import time

from tensorflow_distributed_b200 import data
from tensorflow_distributed_b200.engine import make_engine
from tensorflow_distributed_b200.models.spec import cnn_spec

# idx-gz MNIST files in this directory are used when present; otherwise a deterministic MNIST-shaped synthetic set
mnist = data.read_data_sets("MNIST_data/", one_hot_labels=True)
print("data source:", mnist.source)
""")

code("""
# Training parameters
learning_rate = 0.01
training_iters = 200000      # images seen, not steps
batch_size = 128
display_step = 10
dropout = 0.75               # probability to KEEP a unit
train_flag = 1               # 0: skip training and validate the random initialisation (expect ~10 %)
backend = "auto"             # "cuda" (sm_100a kernels), "cpu" (oracle) or "auto"
""")

code("""
# Variables: the reference's eight tensors, N(0, 1) initialised, in TF layouts (HWIO filters, [in, out] dense)
spec = cnn_spec()
for v in spec.variables:
    print("%-12s %-18s %9d" % (v.name, tuple(v.shape), v.numel))
print("total parameters:", sum(v.numel for v in spec.variables))

engine = make_engine(spec, backend, learning_rate=learning_rate, seed=0, init_stddev=1.0)
print("engine:", type(engine).__name__)
""")

code("""
start_time = time.time()
step = 1
if train_flag == 1:
    while step * batch_size < training_iters:
        batch_x, batch_y = mnist.train.next_batch(batch_size)
        engine.train_step(batch_x, batch_y, dropout)          # forward + backward + Adam
        if step % display_step == 0:
            loss = engine.eval_loss(batch_x, batch_y)          # same batch, keep_prob = 1
            print("Iter " + str(step * batch_size) + ", Minibatch Loss= " + "{:.6f}".format(loss))
        step += 1
    engine.synchronize()
    print("Optimization Finished!")
    training_end_time = time.time()
    print("--- %s seconds Time for Training ---" % (training_end_time - start_time))
else:
    training_end_time = start_time
""")

code("""
# Validation: 5 batches of 1000 held-out images, dropout off
nt = 5
accuracy_arr = []
for counter in range(1, nt + 1):
    val_x, val_y = mnist.validation.next_batch(1000)
    acc = engine.eval_accuracy(val_x, val_y)
    accuracy_arr.append(acc)
    print("Accuracy : %f" % acc)
mean_accuracy = sum(accuracy_arr) / len(accuracy_arr)
print("Mean Accuracy : %f" % mean_accuracy)
engine.synchronize()
testing_end_time = time.time()
print("--- %s seconds Time for Inference ---" % (testing_end_time - training_end_time))
""")

code("""
# Trained parameters, by the checkpoint names the cluster scripts use
params = engine.get_params()
for name in sorted(params):
    p = params[name].float().cpu().numpy()
    print("%-12s mean %+.4f  std %.4f" % (name, p.mean(), p.std()))
""")

NB = {
    "cells": CELLS,
    "metadata": {
        "kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
        "language_info": {"name": "python", "version": "3.12"},
    },
    "nbformat": 4,
    "nbformat_minor": 4,
}

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "convolutional_network.ipynb")
    with open(out, "w") as f:
        json.dump(NB, f, indent=1)
        f.write("\n")
    print("wrote", out)
