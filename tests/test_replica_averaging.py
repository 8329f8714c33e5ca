"""The fork's decentralised averaging: schedule determinism and method semantics."""
import torch

from conftest import run_dist
from shuffle_exchange_b200.runtime.zero.replica_averaging import ReplicaAverager, ShuffleSchedule


def test_schedule_is_deterministic_partition():
    a, b = ShuffleSchedule(8, 4, seed=7), ShuffleSchedule(8, 4, seed=7)
    assert a.table == [[0, 1], [2, 3], [4, 5], [6, 7]]          # initial rings are contiguous
    seen = set()
    for _ in range(20):
        ta, tb = a.reshuffle(), b.reshuffle()
        assert ta == tb                                          # same seed => same rings everywhere
        flat = sorted(x for r in ta for x in r)
        assert flat == list(range(8)) and all(len(r) == 2 for r in ta)
        seen.add(tuple(map(tuple, ta)))
    assert len(seen) > 5                                         # rings really change
    c = ShuffleSchedule(7, 3, seed=1)                            # world % rings != 0 is handled
    assert sorted(len(r) for r in c.reshuffle()) == [2, 2, 3]
    d = a.gossip_destinations(3)
    assert d == b.gossip_destinations(3) and all(x != i for i, x in enumerate(d))


def _avg_worker(rank, world, method):
    av = ReplicaAverager(method=method, rings=2, shuffle_step=1, slice_count=1, seed=3)
    x = torch.full((16,), float(rank + 1))
    if method == "Gossip":
        total_before = torch.tensor([av.alpha * (rank + 1)])
        av.average([x])
        m = [x.clone()]
        av.gossip_merge(m)
        # push-sum invariant: sum_r alpha_r * w_r is conserved (messages in flight counted at the receiver)
        mass = torch.tensor([av.alpha * float(m[0][0])])
        torch.distributed.all_reduce(mass)
        torch.distributed.all_reduce(total_before)
        assert abs(float(mass) - float(total_before)) < 1e-5, (float(mass), float(total_before))
        alpha = torch.tensor([av.alpha])
        torch.distributed.all_reduce(alpha)
        assert abs(float(alpha) - 1.0) < 1e-6
        return
    av.average(x)
    if method in ("RR", "H-RR"):
        assert torch.allclose(x, torch.full((16,), (world + 1) / 2.0)), x
    else:
        ring = av.schedule.ring_of(av.replica)
        assert torch.allclose(x, torch.full((16,), sum(r + 1 for r in ring) / len(ring))), (ring, x)
        av.shuffle_exchange()                                     # shuffle_step=1 -> new rings
        y = torch.full((16,), float(rank + 1))
        av.average(y)
        ring = av.schedule.ring_of(av.replica)
        assert torch.allclose(y, torch.full((16,), sum(r + 1 for r in ring) / len(ring)))


def test_rr():
    run_dist(_avg_worker, 4, ("RR",))


def test_hierarchical_rr_equals_global_mean():
    run_dist(_avg_worker, 4, ("H-RR",))


def test_shuffle_rings():
    run_dist(_avg_worker, 4, ("shuffle",))


def test_gossip_push_sum_conserves_mass():
    run_dist(_avg_worker, 4, ("Gossip",))
