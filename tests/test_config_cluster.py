"""Flags, role validation, cluster spec, chief, device strings (C1-C8)."""
import pytest

from tensorflow_distributed_b200 import cluster as cl
from tensorflow_distributed_b200.config import Flags, ROLE_DEFAULTS, parse_flags, help_text, reference_flag_names


def test_reference_defaults():
    f = Flags()
    assert (f.data_dir, f.download_only, f.task_index, f.num_gpus) == ("/tmp/mnist-data", False, 0, 0)
    assert (f.replicas_to_aggregate, f.hidden_units, f.train_steps, f.batch_size) == (2, 100, 4, 128)
    assert f.learning_rate == 0.01 and f.sync_replicas is True and f.existing_servers is False
    assert f.ps_hosts == "10.0.1.3:2222" and f.worker_hosts == "10.0.1.6:2223,10.0.1.2:2224"
    assert f.job_name == "ps"
    assert len(reference_flag_names()) == 14
    for n in reference_flag_names():
        assert "--" + n in help_text()


def test_role_defaults_match_three_scripts():
    assert parse_flags([], ROLE_DEFAULTS["m"]).job_name == "ps"
    w1 = parse_flags([], ROLE_DEFAULTS["w1"])
    w2 = parse_flags([], ROLE_DEFAULTS["w2"])
    assert (w1.job_name, w1.task_index) == ("worker", 0)
    assert (w2.job_name, w2.task_index) == ("worker", 1)


def test_parse_forms():
    f = parse_flags(["--job_name=worker", "--task_index", "3", "--sync_replicas=False", "--learning_rate=0.5",
                     "--nodownload_only", "--existing_servers", "--replicas_to_aggregate=None"])
    assert f.job_name == "worker" and f.task_index == 3 and f.sync_replicas is False
    assert f.learning_rate == 0.5 and f.download_only is False and f.existing_servers is True
    assert f.replicas_to_aggregate is None
    assert parse_flags(["--sync_replicas", "false"]).sync_replicas is False
    with pytest.raises(SystemExit):
        parse_flags(["--nope=1"])


def test_validate_role():
    cl.validate_role(Flags(job_name="worker"))
    with pytest.raises(ValueError):
        cl.validate_role(Flags(job_name=""))
    with pytest.raises(ValueError):
        cl.validate_role(Flags(job_name="chief"))
    with pytest.raises(ValueError):
        cl.validate_role(Flags(job_name="worker", num_gpus=-1))
    with pytest.raises(ValueError):
        cl.validate_role(Flags(job_name="ps", existing_servers=True))


def test_cluster_spec_and_ranks():
    f = Flags(ps_hosts="a:1,b:2", worker_hosts="c:3,d:4,e:5")
    c = cl.ClusterSpec.from_flags(f)
    assert c.num_ps == 2 and c.num_workers == 3 and c.world_size == 5
    assert c.as_dict() == {"ps": ["a:1", "b:2"], "worker": ["c:3", "d:4", "e:5"]}
    assert c.rank_of("ps", 1) == 1 and c.rank_of("worker", 0) == 2
    assert c.task_of_rank(4) == ("worker", 2)
    assert c.master_endpoint() == ("a", 1)
    with pytest.raises(ValueError):
        c.task_address("worker", 3)
    with pytest.raises(ValueError):
        cl.ClusterSpec({"ps": [""], "worker": ["x:1"]})


def test_chief_and_devices():
    assert cl.is_chief(Flags(task_index=0)) and not cl.is_chief(Flags(task_index=1))
    assert cl.worker_device(Flags(job_name="worker", task_index=5, num_gpus=4)) == "/job:worker/task:5/gpu:1"
    assert cl.worker_device(Flags(job_name="worker", task_index=2, num_gpus=0)) == "/job:worker/task:2/cpu:0"
    g = Flags(job_name="worker", task_index=1, ps_hosts="gpu:0", worker_hosts="gpu:1,gpu:2")
    c = cl.ClusterSpec.from_flags(g)
    assert c.all_gpu and cl.worker_device(g, c) == "/job:worker/task:1/gpu:2"
    assert cl.torch_device_of("/job:worker/task:1/gpu:2") == "cuda:2"
    assert cl.device_filters(g) == ["/job:ps", "/job:worker/task:1"]
    f = cl.device_filters(g)
    assert cl.filter_allows(f, "ps", 0) and cl.filter_allows(f, "ps", 3) and cl.filter_allows(f, "worker", 1)
    assert not cl.filter_allows(f, "worker", 0) and not cl.filter_allows(f, "worker", 10)


def test_ps_client_refuses_a_peer_outside_its_device_filters():
    """m:258: a worker's session sees the ps job and itself.  The CPU transport resolves every peer through the
    filters, so a client cannot address another worker."""
    from tensorflow_distributed_b200.models.spec import mlp_spec
    from tensorflow_distributed_b200.parallel.placement import Placement
    from tensorflow_distributed_b200.parallel.ps_cpu import PSClient
    g = Flags(job_name="worker", task_index=1, ps_hosts="h:1,h:2", worker_hosts="h:3,h:4")
    c = cl.ClusterSpec.from_flags(g)
    spec = mlp_spec(16, 1)
    client = PSClient(c, spec, Placement(spec, 2), 1, device_filters=cl.device_filters(g))
    assert client.peer_rank("ps", 0) == c.rank_of("ps", 0) and client.peer_rank("ps", 1) == c.rank_of("ps", 1)
    assert client.peer_rank("worker", 1) == c.rank_of("worker", 1)
    with pytest.raises(PermissionError):
        client.peer_rank("worker", 0)


def test_replicas_to_aggregate_documented_default():
    assert cl.effective_replicas_to_aggregate(Flags(replicas_to_aggregate=None), 7) == 7
    assert cl.effective_replicas_to_aggregate(Flags(replicas_to_aggregate=0), 7) == 7
    assert cl.effective_replicas_to_aggregate(Flags(), 7) == 2


def test_gpu_hosts_without_gpus_fail_with_a_clear_message():
    """`gpu:<n>` host entries on a box without (enough) CUDA devices: a readable error, not a driver stack trace."""
    import torch
    if torch.cuda.is_available() and torch.cuda.device_count() >= 2:
        pytest.skip("box has the GPUs the host lists name")
    from tensorflow_distributed_b200.trainer import run_cluster
    flags = parse_flags(["--ps_hosts=gpu:0", "--worker_hosts=gpu:1", "--job_name=worker", "--task_index=0"])
    with pytest.raises(RuntimeError, match="name GPUs"):
        run_cluster(flags, log=lambda s: None)


def test_every_flag_has_help_text():
    from tensorflow_distributed_b200.config import _HELP, _HELP_EXTRA, flag_names
    assert set(flag_names()) == set(_HELP) | set(_HELP_EXTRA)
    assert not set(_HELP) & set(_HELP_EXTRA)
    assert all(len(v) > 8 for v in list(_HELP.values()) + list(_HELP_EXTRA.values()))
