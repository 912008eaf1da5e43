"""tensorflow_distributed_b200 -- a B200-native parameter-server training framework.

Capabilities of PranjalSahu/tensorflow_distributed (single-process MNIST CNN
trainer; ps/worker cluster with sync and async replica updates), rebuilt for
Blackwell: PyTorch host, hand-written sm_100a CUDA kernels (tcgen05/TMEM/TMA),
parameter-server shards in GPU HBM reached over NVLink peer memory.
"""
from .config import Flags, parse_flags, ROLE_DEFAULTS  # noqa: F401
from .cluster import ClusterSpec  # noqa: F401
from .models.spec import cnn_spec, mlp_spec, make_spec  # noqa: F401

__version__ = "0.1.0"
