"""Schedule compiler: plan IR, list scheduling under a memory budget, passes, pass schedule, engine integration."""
import pytest
import torch

from conftest import run_dist
from shuffle_exchange_b200.compile import backend as cb
from shuffle_exchange_b200.compile.list_schedule import fast_free_schedule, simple_prefetch
from shuffle_exchange_b200.compile.plan import COMPUTE, GATHER, RELEASE, build_plan
from shuffle_exchange_b200.compile.profilers import create_predictor


def _plan(n=6, size=100):
    return build_plan(range(n), {i: size for i in range(n)}, fwd_us={i: 50.0 for i in range(n)},
                      bwd_us={i: 100.0 for i in range(n)}, gather_us_per_byte=0.4, reduce_us_per_byte=0.4)


def test_naive_plan_peaks_at_one_unit_and_exposes_all_comm():
    p = _plan()
    p.validate()
    sim = p.simulate()
    assert sim["peak_bytes"] == 100
    # forward: 6 gathers fully exposed; backward: every gather also queues behind the previous unit's reduce
    assert sim["exposed_comm_us"] == pytest.approx(6 * 40.0 + 40.0 + 5 * 80.0)


def test_simple_prefetch_respects_budget_and_hides_comm():
    p = _plan()
    for budget, max_live in ((100, 100), (250, 200), (10 ** 6, 600)):
        q = simple_prefetch(p, budget)
        q.validate()
        sim = q.simulate()
        assert sim["peak_bytes"] <= max(budget, 100) and sim["peak_bytes"] <= max_live
        order = [n.unit for n in q.nodes if n.kind == GATHER and n.phase == "fwd"]
        assert order == sorted(order)                       # gathers never reorder
    assert simple_prefetch(p, 300).simulate()["exposed_comm_us"] < p.simulate()["exposed_comm_us"]
    table = simple_prefetch(p, 300).prefetch_table()
    assert any(len(v) > 0 for v in table.values())
    assert all(u not in v for (u, _), v in table.items())


def test_validate_rejects_use_before_gather():
    p = _plan(3)
    p.nodes = [n for n in p.nodes if not (n.kind == GATHER and n.unit == 1 and n.phase == "fwd")]
    with pytest.raises(ValueError):
        p.validate()


def test_selective_gather_and_offload_passes():
    p = fast_free_schedule(_plan(4, 100), 0)
    p.unit_bytes[3] = 10            # small unit: best time-saved per byte... equal times, smaller size wins
    ctx = cb.PassContext(available_mem=100 + 10 + 100, total_mem=10 ** 9, world_size=4)
    q = cb.opt_passes["selective_gather"](p, ctx)
    assert 3 in q.persistent and len(q.persistent) >= 1
    q.validate()
    r = cb.opt_passes["offload_parameters"](p, ctx)
    assert r.meta["offload_parameters"] == [0, 1, 2, 3]
    assert sum(1 for n in r.nodes if n.kind == "reload") == 4
    s = cb.opt_passes["offload_adam_states_sync"](p, cb.PassContext(total_mem=10, opt_state_bytes=1200))
    assert s.meta["offload_adam_states"] and s.nodes[0].kind == "offload" and s.nodes[-1].kind == "reload"
    t = cb.opt_passes["offload_activation"](p, cb.PassContext(total_mem=1000, activation_bytes={0: 600, 1: 600}))
    assert t.meta["offload_activation"] == [0]


def test_schedule_registration_and_validation():
    cb.register_compile_pass("noop", lambda plan, ctx: plan)
    cb.init_schedule([(0, ["zero3_compile"]), (3, ["noop", "prefetch"])])
    with pytest.raises(AssertionError):
        cb.init_schedule([(2, ["zero3_compile"]), (1, ["prefetch"])])
    with pytest.raises(AssertionError):
        cb.init_schedule([(0, ["does_not_exist"])])
    f = create_predictor([(100, 10.0), (1000, 55.0)])
    assert f(100) == pytest.approx(10.0) and f(550) == pytest.approx(32.5) and f(2000) > 55.0


def _engine_worker(rank, world):
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.models import GPT2LMHeadModel, gpt2_config
    torch.manual_seed(0)
    model = GPT2LMHeadModel(gpt2_config("gpt2-tiny"))
    cfg = {"train_micro_batch_size_per_gpu": 2, "steps_per_print": 10 ** 9, "bf16": {"enabled": True},
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0},
           "compile": {"deepcompile": True, "debug_log": True}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg, model_parameters=model.parameters())
    assert engine.is_deepcompile_enabled()
    engine.compile()
    assert engine.is_compiled
    g = torch.Generator().manual_seed(rank)
    losses = []
    for step in range(5):
        ids = torch.randint(0, 512, (2, 16), generator=g)
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    opt = engine.optimizer
    plan = getattr(opt, "_compiled_plan", None)
    assert plan is not None, "compile passes never ran"
    plan.validate()
    assert getattr(opt, "_compiled_prefetch", None)
    assert all(l == l for l in losses) and losses[-1] < losses[0] + 1.0


def test_engine_deepcompile_zero3_two_ranks():
    run_dist(_engine_worker, world=2)


def test_native_scheduler_core_agrees_with_the_python_oracle():
    """csrc/host/plan_sched.cpp: simulation and gather hoisting must reproduce the Python implementation exactly on
    random plans (random unit sizes, times, persistent sets, budgets), and be much faster on a large plan."""
    import random
    import time
    from shuffle_exchange_b200.compile import native_sched
    from shuffle_exchange_b200.compile.list_schedule import simple_prefetch
    from shuffle_exchange_b200.compile.plan import build_plan
    assert native_sched.available()
    rng = random.Random(0)
    for trial in range(25):
        n_units = rng.randint(2, 14)
        order = list(range(n_units))
        ub = {u: rng.choice([1, 2, 4, 8]) * 1024 for u in order}
        fwd = {u: rng.uniform(10, 200) for u in order}
        bwd = {u: 2 * fwd[u] for u in order}
        pers = set(rng.sample(order, rng.randint(0, max(0, n_units // 4))))
        plan = build_plan(order, ub, persistent=pers, fwd_us=fwd, bwd_us=bwd, gather_us_per_byte=0.01,
                          reduce_us_per_byte=0.02)
        a, b = plan.simulate(base_bytes=123), native_sched.simulate(plan, base_bytes=123)
        assert a["peak_bytes"] == b["peak_bytes"] and abs(a["time_us"] - b["time_us"]) < 1e-6 \
            and abs(a["exposed_comm_us"] - b["exposed_comm_us"]) < 1e-6
        budget = rng.choice([max(ub.values()), 2 * max(ub.values()), sum(ub.values())])
        py = simple_prefetch(plan, budget, use_native=False)
        nat = simple_prefetch(plan, budget, use_native=True)
        assert [n.key() for n in py.nodes] == [n.key() for n in nat.nodes], trial
        assert nat.simulate()["peak_bytes"] <= max(budget, py.simulate()["peak_bytes"])
    # scaling: 120 units
    order = list(range(120))
    ub = {u: 4096 for u in order}
    plan = build_plan(order, ub, fwd_us={u: 50.0 for u in order}, bwd_us={u: 100.0 for u in order},
                      gather_us_per_byte=0.01)
    t0 = time.time(); nat = simple_prefetch(plan, 3 * 4096, use_native=True); t_nat = time.time() - t0
    t0 = time.time(); py = simple_prefetch(plan, 3 * 4096, use_native=False); t_py = time.time() - t0
    assert [n.key() for n in py.nodes] == [n.key() for n in nat.nodes]
    assert t_nat < t_py, (t_nat, t_py)
    print(f"hoisting 120 units: python {t_py:.3f}s, native {t_nat:.4f}s")
