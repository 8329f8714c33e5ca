"""Launcher: resource parsing of the front end (hostfile, include / exclude filters, node counts), multi-node command
lines, and the per-node launcher actually starting a 2-process gloo job (and tearing it down when one rank fails)."""
import collections
import os
import subprocess
import sys
import textwrap
import types

import pytest

from shuffle_exchange_b200.launcher import launch, multinode_runner as mnr, runner

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hostfile_and_resource_filters(tmp_path):
    hf = tmp_path / "hostfile"
    hf.write_text("# cluster\nworker-0 slots=4\nworker-1 slots=4\n\nworker-2 slots=2\n")
    pool = runner.fetch_hostfile(str(hf))
    assert pool == collections.OrderedDict([("worker-0", 4), ("worker-1", 4), ("worker-2", 2)])
    assert runner.fetch_hostfile(str(tmp_path / "absent")) is None
    for bad in ("worker-0 slots=4\nworker-0 slots=2\n", "worker-0 gpus=4\n", "\n"):
        with pytest.raises(ValueError):
            runner._parse_hostfile(bad.splitlines())
    inc = runner.parse_inclusion_exclusion(pool, "worker-0:0,2@worker-2", "")
    assert inc == collections.OrderedDict([("worker-0", [0, 2]), ("worker-2", [0, 1])])
    exc = runner.parse_inclusion_exclusion(pool, "", "worker-1@worker-0:3")
    assert exc == collections.OrderedDict([("worker-0", [0, 1, 2]), ("worker-2", [0, 1])])
    with pytest.raises(ValueError):
        runner.parse_inclusion_exclusion(pool, "worker-0", "worker-1")
    with pytest.raises(ValueError):
        runner.parse_inclusion_exclusion(pool, "worker-9", "")
    with pytest.raises(ValueError):
        runner.parse_inclusion_exclusion(pool, "worker-2:5", "")
    assert runner.parse_num_nodes("3", False) == (3, -1) and runner.parse_num_nodes("2:4", True) == (2, 4)
    with pytest.raises(RuntimeError):
        runner.parse_num_nodes("4:2", True)
    assert runner.parse_node_config_list(["a:0,1", "b"]) == collections.OrderedDict([("a", [0, 1]), ("b", [])])
    info = runner.encode_world_info(inc)
    assert launch.decode_world_info(info) == {"worker-0": [0, 2], "worker-2": [0, 1]}
    mapping, world = launch.global_rank_mapping(launch.decode_world_info(info))
    assert dict(mapping) == {"worker-0": [0, 1], "worker-2": [2, 3]} and world == 4


def _args(**kw):
    base = dict(user_script="train.py", user_args=["--lr", "1e-3", "--name", "a b"], master_addr="10.0.0.1",
                master_port=29500, launcher_args="", hostfile="/job/hostfile", no_python=False, module=False,
                no_local_rank=False, save_pid=False, elastic_training=False, include="", exclude="", num_nodes=-1,
                num_gpus=-1, ssh_port=None, comment="")
    base.update(kw)
    return types.SimpleNamespace(**base)


def test_multinode_command_lines():
    pool = collections.OrderedDict([("w0", 8), ("w1", 8)])
    active = collections.OrderedDict((h, list(range(n))) for h, n in pool.items())
    info = runner.encode_world_info(active)
    r = mnr.PDSHRunner(_args(ssh_port=2222), info)
    r.add_export("NCCL_DEBUG", "INFO")
    env = {}
    cmd, kill_cmd, env = r.get_cmd(env, active)
    assert cmd[:6] == ["pdsh", "-S", "-f", str(mnr.PDSH_MAX_FAN_OUT), "-w", "w0,w1"]
    assert env["PDSH_RCMD_TYPE"] == "ssh" and "-p 2222" in env["PDSH_SSH_ARGS_APPEND"]
    joined = " ".join(cmd)
    assert "export NCCL_DEBUG=INFO;" in joined and "shuffle_exchange_b200.launcher.launch" in joined
    assert f"--world_info={info}" in cmd and "--node_rank=%n" in cmd and "--master_addr=10.0.0.1" in cmd
    assert cmd[-5:] == ["train.py", "--lr", "1e-3", "--name", "'a b'"]        # user arguments are shell-quoted

    o = mnr.OpenMPIRunner(_args(), info, pool)
    o.add_export("FOO", "bar baz")
    c = o.get_cmd({}, active)
    assert c[:5] == ["mpirun", "-n", "16", "-hostfile", "/job/hostfile"] and "UCX_TLS=tcp" in c and "FOO='bar baz'" in c
    assert c[c.index("train.py") - 2:c.index("train.py")] == [sys.executable, "-u"]
    with pytest.raises(ValueError):
        mnr.OpenMPIRunner(_args(include="w0"), info, pool).validate_args()

    m = mnr.MPICHRunner(_args(), info, pool).get_cmd({}, active)
    assert m[:5] == ["mpirun", "-n", "16", "-ppn", "8"] and m[m.index("-hosts") + 1] == "w0,w1"
    assert m[m.index("WORLD_SIZE") + 1] == "16"

    s = mnr.SlurmRunner(_args(num_nodes=2, comment="exp1"), info, pool)
    s.add_export("A", "1")
    sc = s.get_cmd({}, active)
    assert sc[:3] == ["srun", "-n", "16"] and "--nodes" in sc and "--export=ALL,A=1" in sc and "exp1" in sc


def _run_launch(tmp_path, body, extra=()):
    script = tmp_path / "job.py"
    script.write_text(textwrap.dedent(body))
    info = runner.encode_world_info({"localhost": [0, 1]})
    from conftest import free_port
    cmd = [sys.executable, "-u", "-m", "shuffle_exchange_b200.launcher.launch", f"--world_info={info}",
           "--master_addr=127.0.0.1", f"--master_port={free_port()}", *extra, str(script), str(tmp_path)]
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)


def test_node_launcher_runs_a_two_process_job(tmp_path):
    r = _run_launch(tmp_path, """
        import argparse, os, sys, torch, torch.distributed as dist
        ap = argparse.ArgumentParser(); ap.add_argument("--local_rank", type=int); ap.add_argument("out")
        a = ap.parse_args()
        assert a.local_rank == int(os.environ["LOCAL_RANK"]) and os.environ["CUDA_VISIBLE_DEVICES"] == "0,1"
        assert os.environ["LOCAL_SIZE"] == "2" and os.environ["CROSS_SIZE"] == "1"
        dist.init_process_group("gloo")
        t = torch.tensor([float(dist.get_rank() + 1)])
        dist.all_reduce(t)
        print("sum", t.item(), flush=True)
        open(os.path.join(a.out, f"done{dist.get_rank()}"), "w").write(str(t.item()))
        dist.destroy_process_group()
    """, extra=(f"--enable_each_rank_log={tmp_path}/logs",))
    assert r.returncode == 0, r.stderr[-2000:]
    assert open(tmp_path / "done0").read() == open(tmp_path / "done1").read() == "3.0"
    assert "sum 3.0" in open(tmp_path / "logs" / "rank1.log").read()


def test_node_launcher_takes_the_job_down_when_a_rank_fails(tmp_path):
    r = _run_launch(tmp_path, """
        import os, sys, time
        if os.environ["RANK"] == "1":
            sys.exit(7)
        time.sleep(120)           # rank 0 would hang forever waiting for its peer
    """)
    assert r.returncode == 7, (r.returncode, r.stderr[-1500:])
