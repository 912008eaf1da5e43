"""Cluster program, defaults of the reference's mnist_python_w2.py: worker task 1.

Same flags as /root/reference/mnist_python_w2.py:49-87; every default can be overridden.
"""
from tensorflow_distributed_b200.config import ROLE_DEFAULTS, parse_flags
from tensorflow_distributed_b200.trainer import run_cluster

if __name__ == "__main__":
    run_cluster(parse_flags(defaults=ROLE_DEFAULTS["w2"]))
