"""TF_CONFIG parsing (README.md:322-327) and the sparklyr synthesis rule (README.md:180-183)."""
import json

import pytest

from distributed_b200.parallel.cluster import ClusterSpec, ClusterSpecError, local_cluster

WORKERS = ["172.17.0.6:10090", "172.17.0.3:10088", "172.17.0.4:10087", "172.17.0.5:10089"]


@pytest.mark.parametrize("index", range(4))
def test_readme_four_worker_spec(index):
    env = {"TF_CONFIG": json.dumps({"cluster": {"worker": WORKERS}, "task": {"type": "worker", "index": index}})}
    spec = ClusterSpec.from_env(env)
    assert spec.num_workers == 4 and spec.task_index == index and spec.task_type == "worker"
    assert spec.workers == tuple(WORKERS)
    assert spec.is_chief == (index == 0)
    assert spec.host(1) == "172.17.0.3" and spec.port(1) == 10088
    assert ClusterSpec.from_json(spec.to_json()) == spec


def test_malformed_json_and_bad_fields():
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_env({"TF_CONFIG": "{not json"})
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_dict({"cluster": {"worker": WORKERS}, "task": {"type": "worker", "index": 4}})
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_dict({"cluster": {"worker": WORKERS}, "task": {"type": "ps", "index": 0}})
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_dict({"cluster": {"worker": ["nohostport"]}, "task": {"type": "worker", "index": 0}})
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_dict({"cluster": {}, "task": {"type": "worker", "index": 0}})
    with pytest.raises(ClusterSpecError):
        ClusterSpec.from_dict({"task": {"type": "worker", "index": 0}})


def test_r_style_unboxed_scalars_as_lists():
    # jsonlite::toJSON without auto_unbox wraps scalars in arrays (README.md:84-89 uses auto_unbox)
    spec = ClusterSpec.from_dict({"cluster": {"worker": WORKERS}, "task": {"type": ["worker"], "index": [2]}})
    assert spec.task_index == 2


def test_barrier_synthesis_matches_sparklyr_recipe():
    # paste(gsub(":[0-9]+$", "", barrier$address), 8000 + seq_along(barrier$address), sep = ":")
    spec = ClusterSpec.from_barrier(["10.0.0.1:41234", "10.0.0.2:51234", "10.0.0.3:61234"], partition=1)
    assert spec.workers == ("10.0.0.1:8001", "10.0.0.2:8002", "10.0.0.3:8003")
    assert spec.task_index == 1


def test_torchrun_env_and_default():
    spec = ClusterSpec.from_env({"RANK": "3", "WORLD_SIZE": "8", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29511"})
    assert spec.num_workers == 8 and spec.task_index == 3 and spec.single_host()
    assert ClusterSpec.from_env({}).num_workers == 1


def test_local_cluster_is_single_host_with_distinct_ports():
    specs = local_cluster(3)
    assert len({s.port() for s in specs}) == 3 and all(s.single_host() for s in specs)
    assert len({s.job_id() for s in specs}) == 1
