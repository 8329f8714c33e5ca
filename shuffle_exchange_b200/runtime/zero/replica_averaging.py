"""Shuffle-exchange replica averaging — the feature the reference fork is named after.

Reference behaviour (deepspeed/runtime/zero/stage_1_and_2.py:163-240, 692-734, 2092-2108, 2185-2250):
the world is cut into consecutive *slices* of ``slice_count`` ranks; ZeRO partitioning and gradient
reduction happen only inside a slice, so there are ``R = world / slice_count`` independent model
replicas doing local SGD.  After every optimizer step each rank averages its updated bit16
partition with the *same partition* of other replicas, chosen by ``method``:

    RR       all R replicas                                   (plain all-reduce mean)
    shuffle  the replicas of my ring; rings are re-drawn as a uniform random partition of the
             replicas into ``rings`` equal groups every ``shuffle_step`` batches
    H-RR     two-level: reduce inside each of 2 rings to its head, all-reduce between heads,
             broadcast back (numerically the global mean)
    Gossip   push-sum: every replica halves its weight alpha and pushes (alpha, partition) to one random
             peer; receivers merge weighted by alpha at the start of their next step

The reference implements ring changes by destroying and re-creating NCCL process groups and relies
on every rank having the same *implicit* CPU RNG state.  Here the schedule is **data**: a seeded
generator (identical on all ranks by construction) produces a membership table; the exchange runs
over ONE fixed "column" group (all replicas holding the same partition index) either as

  * a fused symmetric-memory kernel that reads the ring members' partitions straight over
    NVLink using the table as its peer-address schedule (``comm/symm.py``; used when available), or
  * batched ``isend/irecv`` inside the column group (NCCL or gloo) — the baseline path, also what the
    CPU tests exercise.

No communicator is ever rebuilt, ``world % rings != 0`` is handled (rings differ in size by at most
one; the reference only logs an error, stage_1_and_2.py:699-700), and Gossip needs no unmatched
blocking ``send``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from ...comm import comm as dist
from ...utils.logging import logger

METHODS = ("RR", "shuffle", "H-RR", "Gossip")


class ShuffleSchedule:
    """Deterministic ring membership: ``rings_at(epoch)`` is a uniform random partition of
    ``range(n_replicas)`` into ``rings`` groups, identical on every rank for a given seed."""

    def __init__(self, n_replicas: int, rings: int, seed: int = 1234):
        self.n = n_replicas
        self.rings = max(1, min(rings, n_replicas))
        self.seed = seed
        self.epoch = 0
        self._table = self._contiguous()

    def _split(self, order: List[int]) -> List[List[int]]:
        base, extra = divmod(self.n, self.rings)
        out, cur = [], 0
        for r in range(self.rings):
            size = base + (1 if r < extra else 0)
            out.append(sorted(order[cur: cur + size]))
            cur += size
        return out

    def _contiguous(self) -> List[List[int]]:
        return self._split(list(range(self.n)))

    def reshuffle(self) -> List[List[int]]:
        self.epoch += 1
        g = torch.Generator(device="cpu")
        g.manual_seed(self.seed * 1000003 + self.epoch)
        self._table = self._split(torch.randperm(self.n, generator=g).tolist())
        return self._table

    def reset(self, rings: int):
        self.rings = max(1, min(rings, self.n))
        return self.reshuffle()

    @property
    def table(self) -> List[List[int]]:
        return self._table

    def ring_of(self, replica: int) -> List[int]:
        for ring in self._table:
            if replica in ring:
                return ring
        raise ValueError(replica)

    def gossip_destinations(self, step: int) -> List[int]:
        """dest[r] != r drawn uniformly; same list on every rank."""
        if self.n == 1:
            return [0]
        g = torch.Generator(device="cpu")
        g.manual_seed(self.seed * 7919 + 13 * step + 1)
        d = torch.randint(0, self.n - 1, (self.n,), generator=g).tolist()
        return [x if x < r else x + 1 for r, x in enumerate(d)]

    def state_dict(self):
        return {"epoch": self.epoch, "rings": self.rings, "seed": self.seed, "table": self._table}

    def load_state_dict(self, sd):
        self.epoch, self.rings, self.seed = sd["epoch"], sd["rings"], sd["seed"]
        self._table = [list(r) for r in sd["table"]]


class ReplicaAverager:
    """Owns the slice / column process groups and performs the per-step averaging."""

    def __init__(self, method: str = "RR", rings: int = 8, shuffle_step: int = 50, slice_count: int = 2,
                 seed: int = 1234, fused: Optional[bool] = None):
        if method not in METHODS:
            raise ValueError(f"replica averaging method must be one of {METHODS}, got {method}")
        self.method = method
        self.real_world = dist.get_world_size()
        self.real_rank = dist.get_rank()
        slice_count = max(1, min(slice_count, self.real_world))
        if self.real_world % slice_count != 0:
            raise ValueError(f"world size {self.real_world} is not divisible by slice_count {slice_count}")
        self.slice_count = slice_count
        self.n_replicas = self.real_world // slice_count
        self.replica = self.real_rank // slice_count
        self.slice_partition = self.real_rank % slice_count
        if method == "H-RR":
            rings = 2
        self.schedule = ShuffleSchedule(self.n_replicas, rings, seed)
        self.shuffle_step = max(1, shuffle_step)
        self.batch_count = 0
        self.step_count = 0
        self.fused = fused
        self._symm = None
        self.alpha = 1.0 / self.n_replicas       # push-sum weight (Gossip)
        self._inbox: List[Tuple[float, List[torch.Tensor]]] = []
        self.slice_pg = None
        self.column_pg = None
        self.column_ranks = [r * slice_count + self.slice_partition for r in range(self.n_replicas)]
        self._build_groups()

    # ----------------------------------------------------------------------------------- groups
    def _build_groups(self):
        if not dist.is_initialized() or self.real_world == 1:
            return
        for s in range(self.n_replicas):          # one slice group per replica
            ranks = list(range(s * self.slice_count, (s + 1) * self.slice_count))
            g = dist.new_group(ranks)
            if self.real_rank in ranks:
                self.slice_pg = g
        for c in range(self.slice_count):         # one column group per partition index
            ranks = [r * self.slice_count + c for r in range(self.n_replicas)]
            g = dist.new_group(ranks)
            if self.real_rank in ranks:
                self.column_pg = g

    # -------------------------------------------------------------------------- user-facing API
    def shuffle_exchange(self) -> bool:
        """Call once per batch; re-draws the rings every ``shuffle_step`` batches (``shuffle`` only)."""
        if self.method != "shuffle":
            return False
        self.batch_count += 1
        if self.batch_count % self.shuffle_step == 0:
            self.schedule.reshuffle()
            return True
        return False

    def reset_rings(self, rings: int):
        if self.method != "shuffle":
            return
        self.schedule.reset(rings)
        self.batch_count = 0

    def synchronization(self, flat_buffers: List[torch.Tensor]):
        """Global mean of whole flat bit16 buffers over the REAL world (``shuffle`` / ``Gossip``)."""
        if self.real_world == 1 or self.method not in ("shuffle", "Gossip"):
            return
        for t in flat_buffers:
            t.div_(self.real_world)
            dist.all_reduce(t)

    # -------------------------------------------------------------------------------- averaging
    def peers(self) -> List[int]:
        """Replica ids this replica averages with in the current step (including itself)."""
        if self.method == "RR" or self.method == "H-RR":
            return list(range(self.n_replicas))
        if self.method == "shuffle":
            return self.schedule.ring_of(self.replica)
        return [self.replica]

    def average(self, shards):
        """In-place averaging of this rank's updated partition(s) (one tensor per param group)."""
        if torch.is_tensor(shards):
            shards = [shards]
        self.step_count += 1
        if self.n_replicas == 1:
            return
        if self.method == "Gossip":
            return self._gossip_push(shards)
        for shard in shards:
            if self.method == "RR":
                # A mean over ALL replicas of the fixed column group is a plain all-reduce: NCCL's in-switch (NVLS)
                # algorithm beat the peer-load kernel on it (8 GPUs, 4 replicas x 512 MB: 1.47 vs 2.04 ms,
                # profiles/replica_avg_bench_8gpu_r2.json), so the kernel serves RR only when asked for (fused=True).
                if not (self.fused is True and self._try_fused(shard, list(range(self.n_replicas)))):
                    shard.div_(self.n_replicas)
                    dist.all_reduce(shard, group=self.column_pg)
            elif self.method == "H-RR":
                # reduce-to-head / all-reduce between heads / broadcast sums every replica once and divides by their
                # number: numerically the global mean.  On a switch there is no hierarchy to exploit, so the one-shot
                # kernel reads every replica directly; the three-hop p2p form remains the NCCL / gloo path.
                if not self._try_fused(shard, list(range(self.n_replicas))):
                    self._hierarchical(shard)
            else:
                self._subset_mean(shard, self.schedule.ring_of(self.replica))

    def _subset_mean(self, shard: torch.Tensor, members: List[int]):
        """Mean over ``members`` (replica ids) using p2p inside the fixed column group."""
        if len(members) <= 1:
            return
        if self._try_fused(shard, members):
            return
        others = [m for m in members if m != self.replica]
        recv = [torch.empty_like(shard) for _ in others]
        ops = []
        for buf, m in zip(recv, others):
            peer = self.column_ranks[m]
            ops.append(torch.distributed.P2POp(torch.distributed.irecv, buf, peer, self.column_pg))
            ops.append(torch.distributed.P2POp(torch.distributed.isend, shard, peer, self.column_pg))
        # order deterministically by (peer, kind) so every pair matches on gloo as well
        for w in torch.distributed.batch_isend_irecv(ops):
            w.wait()
        acc = shard.float()
        for buf in recv:
            acc += buf.float()
        shard.copy_((acc / len(members)).to(shard.dtype))

    def _hierarchical(self, shard: torch.Tensor):
        """reduce -> inter-head all-reduce -> broadcast, all as p2p on the column group."""
        table = self.schedule.table
        heads = [ring[0] for ring in table]
        ring = self.schedule.ring_of(self.replica)
        head = ring[0]
        shard.div_(self.n_replicas)
        P2P = torch.distributed.P2POp
        # 1) members -> head (sum)
        if self.replica == head:
            bufs = [torch.empty_like(shard) for _ in ring[1:]]
            ops = [P2P(torch.distributed.irecv, b, self.column_ranks[m], self.column_pg) for b, m in zip(bufs, ring[1:])]
            if ops:
                for w in torch.distributed.batch_isend_irecv(ops):
                    w.wait()
            acc = shard.float()
            for b in bufs:
                acc += b.float()
            # 2) heads exchange partial sums
            hb = [torch.empty_like(acc) for h in heads if h != head]
            ops = []
            for b, h in zip(hb, [h for h in heads if h != head]):
                ops.append(P2P(torch.distributed.irecv, b, self.column_ranks[h], self.column_pg))
                ops.append(P2P(torch.distributed.isend, acc, self.column_ranks[h], self.column_pg))
            if ops:
                for w in torch.distributed.batch_isend_irecv(ops):
                    w.wait()
            tot = acc.clone()
            for b in hb:
                tot += b
            shard.copy_(tot.to(shard.dtype))
            # 3) head -> members
            ops = [P2P(torch.distributed.isend, shard, self.column_ranks[m], self.column_pg) for m in ring[1:]]
            if ops:
                for w in torch.distributed.batch_isend_irecv(ops):
                    w.wait()
        else:
            for w in torch.distributed.batch_isend_irecv([P2P(torch.distributed.isend, shard, self.column_ranks[head], self.column_pg)]):
                w.wait()
            for w in torch.distributed.batch_isend_irecv([P2P(torch.distributed.irecv, shard, self.column_ranks[head], self.column_pg)]):
                w.wait()

    # ----------------------------------------------------------------------------------- gossip
    def _gossip_push(self, shards: List[torch.Tensor]):
        dests = self.schedule.gossip_destinations(self.step_count)
        senders_to_me = [r for r, d in enumerate(dests) if d == self.replica and r != self.replica]
        # Push-sum weights depend on the (shared, seeded) schedule only, never on data: every rank tracks the weight of
        # EVERY replica, so no alpha travels with the payload and nothing is read back from the device.
        if not hasattr(self, "_alphas"):
            self._alphas = [1.0 / self.n_replicas] * self.n_replicas
            self._pending_alpha = [0.0] * self.n_replicas
        for r in range(self.n_replicas):           # what merge() of the previous step added, then this step's halving
            self._alphas[r] += self._pending_alpha[r]
            self._pending_alpha[r] = 0.0
        sent = [a * 0.5 for a in self._alphas]
        self._alphas = list(sent)
        for r, d in enumerate(dests):
            self._pending_alpha[d] += sent[r]
        self.alpha = self._alphas[self.replica]
        P2P = torch.distributed.P2POp
        ops, recv = [], []
        dst = self.column_ranks[dests[self.replica]]
        for sh in shards:
            ops.append(P2P(torch.distributed.isend, sh, dst, self.column_pg))
        for s in senders_to_me:
            bufs = [torch.empty_like(sh) for sh in shards]
            recv.append((sent[s], bufs))
            for b in bufs:
                ops.append(P2P(torch.distributed.irecv, b, self.column_ranks[s], self.column_pg))
        for w in torch.distributed.batch_isend_irecv(ops):
            w.wait()
        self._inbox.extend(recv)

    def gossip_merge(self, masters: List[torch.Tensor]):
        """Start-of-step merge of queued messages: w <- (alpha w + sum alpha' w') / (alpha + sum alpha')."""
        if self.method != "Gossip" or not self._inbox:
            return
        tot = self.alpha + sum(a for a, _ in self._inbox)
        for gi, m in enumerate(masters):
            acc = m.float() * self.alpha
            for a, bufs in self._inbox:
                acc += a * bufs[gi].to(device=m.device, dtype=torch.float32)
            m.copy_((acc / tot).to(m.dtype))
        self.alpha = tot
        self._inbox = []

    # ------------------------------------------------------------------------------------ fused
    def _try_fused(self, shard, members, weights=None) -> bool:
        if self.fused is False or not shard.is_cuda:
            return False
        try:
            from ...comm import symm
        except Exception:
            return False
        if self._symm is None:
            self._symm = symm.ReplicaAverageKernel.maybe_create(self.column_pg, shard)
        if self._symm is None or not self._symm.fits(shard):
            if self.fused:
                raise RuntimeError("fused replica averaging requested but symmetric memory is unavailable")
            return False
        # replica id == rank inside the column group.  Without explicit weights every member of the list computes the
        # same mean (RR, H-RR: all replicas; shuffle: the rings partition the replicas), so the two-phase kernel applies:
        # each member reduces one slice, the reduced slices are exchanged.
        self._symm.run(shard, list(members), weights, shared=weights is None)
        return True

    def state_dict(self):
        return {"schedule": self.schedule.state_dict(), "batch_count": self.batch_count, "step_count": self.step_count,
                "alpha": self.alpha, "method": self.method}

    def load_state_dict(self, sd):
        self.schedule.load_state_dict(sd["schedule"])
        self.batch_count, self.step_count, self.alpha = sd["batch_count"], sd["step_count"], sd["alpha"]
