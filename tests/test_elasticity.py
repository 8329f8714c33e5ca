"""Elastic batch planning: properties of the planner plus the small documented example of the reference."""
import pytest

from shuffle_exchange_b200.elasticity import ElasticityConfigError, ElasticityIncompatibleWorldSize, compute_elastic_config
from shuffle_exchange_b200.elasticity.elasticity import HCN_LIST, get_valid_gpus

BASE = {"elasticity": {"enabled": True, "max_train_batch_size": 10000, "micro_batch_sizes": [8, 12, 16, 17],
                       "min_gpus": 32, "max_gpus": 1500, "min_time": 20, "version": 0.1}}


def test_hcn_generation():
    assert HCN_LIST[:12] == [1, 2, 4, 6, 12, 24, 36, 48, 60, 120, 180, 240] and HCN_LIST[-1] == 720720


def test_basic_10k():
    bs, gpus = compute_elastic_config(BASE, target_deepspeed_version="0.3.8")
    assert bs <= 10000 and bs > 5000
    assert len(gpus) >= 20
    for g in gpus:
        assert any(bs % (mb * g) == 0 for mb in BASE["elasticity"]["micro_batch_sizes"])
        assert 32 <= g <= 1500


def test_world_size_and_microbatch():
    cfg = {"elasticity": {"enabled": True, "max_train_batch_size": 4, "micro_batch_sizes": [1, 2, 3, 4], "min_gpus": 1,
                          "max_gpus": 4, "min_time": 20, "version": 0.1}}
    bs, gpus, mb = compute_elastic_config(cfg, "0.3.8", world_size=2)
    # the LCM base (12) exceeds the cap and is kept as its own candidate, as in the reference (elasticity.py:31-33)
    assert bs == 12 and gpus == [1, 2, 3, 4] and (bs // 2) % mb == 0 and mb == 3
    cfg2 = {"elasticity": dict(cfg["elasticity"], micro_batch_sizes=[1, 2, 4])}
    with pytest.raises(ElasticityIncompatibleWorldSize):
        compute_elastic_config(cfg2, "0.3.8", world_size=3)
    # reference unit test (tests/unit/elasticity/test_elastic.py: test_proper_mbsz)
    ref = {"elasticity": {"enabled": True, "max_train_batch_size": 32, "micro_batch_sizes": [1, 2, 3, 7], "min_gpus": 1,
                          "version": 0.1}}
    _, _, m = compute_elastic_config(ref, "0.3.8", world_size=7)
    assert m == 3


def test_disabled_and_missing():
    with pytest.raises(ElasticityConfigError):
        compute_elastic_config({"elasticity": {"enabled": False}}, "0.3.8")
    with pytest.raises(ElasticityConfigError):
        compute_elastic_config({}, "0.3.8")


def test_v02_node_granularity():
    cfg = {"elasticity": {"enabled": True, "max_train_batch_size": 512, "micro_batch_sizes": [2, 4], "min_gpus": 8,
                          "max_gpus": 64, "num_gpus_per_node": 8, "model_parallel_size": 2, "version": 0.2}}
    bs, gpus, mb = compute_elastic_config(cfg, "0.3.8", world_size=16, return_microbatch=True)
    assert bs % 4 == 0 and all(g % 4 == 0 for g in gpus) and mb in (2, 4)


def test_valid_gpus_bruteforce():
    assert get_valid_gpus(24, [2, 3], 1, 100) == sorted({g for mb in (2, 3) for g in range(1, 25) if (24 // mb) % g == 0})
