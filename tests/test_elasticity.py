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


def test_elastic_agent_restarts_failed_group_with_launcher_env(tmp_path):
    """Two local workers under DSElasticAgent: one dies on the first attempt, the agent restarts the group, the second
    attempt succeeds; the launcher's environment reaches every attempt; membership shrink reads as unhealthy."""
    import os
    import subprocess
    import sys
    import textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    train = tmp_path / "train.py"
    train.write_text(textwrap.dedent("""
        import os, sys
        marker = sys.argv[1]
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        assert os.environ.get("SXB_TEST_ENV") == "from-launcher"
        attempt = int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"))
        open(f"{marker}.rank{rank}.attempt{attempt}", "w").write(str(world))
        if attempt == 0 and rank == 1:
            sys.exit(3)
    """))
    driver = tmp_path / "driver.py"
    driver.write_text(textwrap.dedent(f"""
        import socket, sys, types
        sys.path.insert(0, {root!r})
        from shuffle_exchange_b200.elasticity.elastic_agent import run_elastic
        s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
        args = types.SimpleNamespace(no_python=False, module=False, training_script={str(train)!r},
                                     training_script_args=[sys.argv[1]], min_elastic_nodes=1, max_elastic_nodes=1,
                                     master_addr="127.0.0.1", master_port=port, elastic_monitor_interval=0.5,
                                     max_elastic_restarts=2)
        sys.exit(run_elastic(args, {{"SXB_TEST_ENV": "from-launcher"}}, [0, 1], 2))
    """))
    marker = str(tmp_path / "m")
    r = subprocess.run([sys.executable, str(driver), marker], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    made = sorted(f for f in os.listdir(tmp_path) if f.startswith("m."))
    assert made == ["m.rank0.attempt0", "m.rank0.attempt1", "m.rank1.attempt0", "m.rank1.attempt1"]

    # scale-down rule, on a stub rendezvous: fewer participants than at the last start -> unhealthy
    import types
    from torch.distributed.elastic.agent.server.api import RunResult, WorkerState
    from shuffle_exchange_b200.elasticity import elastic_agent as ea
    handler = types.SimpleNamespace(_state_holder=types.SimpleNamespace(state=types.SimpleNamespace(participants={"a": 0, "b": 1})))
    assert ea._participants(handler) == 2 and ea._participants(object()) == 0
    agent = ea.DSElasticAgent.__new__(ea.DSElasticAgent)
    agent._participants_at_start = 2
    wg = types.SimpleNamespace(spec=types.SimpleNamespace(rdzv_handler=handler, role="trainer"))
    healthy = RunResult(state=WorkerState.HEALTHY)
    base = ea.LocalElasticAgent._monitor_workers
    ea.LocalElasticAgent._monitor_workers = lambda self, g: healthy
    try:
        assert agent._monitor_workers(wg).state == WorkerState.HEALTHY
        handler._state_holder.state.participants.pop("b")
        assert agent._monitor_workers(wg).state == WorkerState.UNHEALTHY
    finally:
        ea.LocalElasticAgent._monitor_workers = base
