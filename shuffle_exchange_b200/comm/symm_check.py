"""Flag-protocol checker for the peer-memory collectives (the reference has no race tooling; its
defences are stream choreography and asserts).

Every symmetric buffer owns a signal pad laid out ``[channel][16 peers]`` of ``uint32`` epochs: peer ``p`` announces
"I reached epoch ``e`` on channel ``c``" by storing ``e`` into slot ``[c][p]`` of EVERY rank's pad, and a rank proceeds
when all slots of a channel reached its own epoch.  The protocol is correct only while these invariants hold at every
point where the host can observe the pad:

* **no regress** — a slot never decreases between two snapshots (epochs are monotonic; a smaller value means a stale or
  misdirected store, e.g. two buffers sharing a pad);
* **bounded lead** — a peer is never more than ``max_lead`` epochs ahead of this rank's own counter for the channel
  (handshake channels need this rank's participation for the next epoch, so the lead is at most 1);
* **completed waits** — after this rank *passed* the wait for epoch ``e`` on a channel, every participating slot is
  ``>= e``;
* **quiescent agreement** — at a global barrier all participating slots of a channel are equal.

``check_pad`` evaluates them on a host snapshot; ``PadMonitor`` keeps the previous snapshot per buffer and can be
called from a training loop (``SXB_SYMM_CHECK=1`` makes ``FusedZ3Comm`` do so once per step)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

MAX_PEERS = 16
CHANNEL_NAMES = ("grads_ready", "grads_read", "avg_ready", "avg_read", "ag_ready", "a2a_free", "a2a_landed")


@dataclass
class Violation:
    rule: str
    channel: int
    peer: int
    value: int
    expected: str

    def __str__(self):
        name = CHANNEL_NAMES[self.channel] if self.channel < len(CHANNEL_NAMES) else str(self.channel)
        return f"{self.rule}: channel {name} peer {self.peer} holds {self.value}, expected {self.expected}"


def _signed_diff(a: int, b: int) -> int:
    """``a - b`` in the wrap-around arithmetic the kernels use (``(int32)(a - b)``)."""
    d = (a - b) & 0xFFFFFFFF
    return d - (1 << 32) if d & 0x80000000 else d


def check_pad(pad: Sequence[int], epochs: Sequence[int], world: int, prev: Optional[Sequence[int]] = None,
              max_lead: int = 1, passed_wait: bool = True, quiescent: bool = False,
              channels: Optional[Sequence[int]] = None) -> List[Violation]:
    """``pad``: flat ``[channels * 16]`` host copy of this rank's signal pad; ``epochs``: this rank's per-channel
    counters (``SymmBuffer.epochs``).  Returns the violated invariants (empty list = consistent)."""
    out: List[Violation] = []
    n_ch = len(pad) // MAX_PEERS
    for c in (channels if channels is not None else range(min(n_ch, len(epochs)))):
        mine = int(epochs[c])
        row = [int(pad[c * MAX_PEERS + p]) for p in range(world)]
        if prev is not None:
            for p, v in enumerate(row):
                before = int(prev[c * MAX_PEERS + p])
                if _signed_diff(v, before) < 0:
                    out.append(Violation("no regress", c, p, v, f">= {before}"))
        if mine == 0 and not any(row):
            continue                                   # channel never used on this buffer
        for p, v in enumerate(row):
            lead = _signed_diff(v, mine)
            if lead > max_lead:
                out.append(Violation("bounded lead", c, p, v, f"<= {mine + max_lead}"))
            if passed_wait and lead < 0:
                out.append(Violation("completed waits", c, p, v, f">= {mine}"))
        if quiescent and len(set(row)) > 1:
            p = min(range(world), key=lambda i: row[i])
            out.append(Violation("quiescent agreement", c, p, row[p], f"== {max(row)}"))
    return out


class PadMonitor:
    """Remembers the last snapshot of every buffer's pad; ``check`` raises (or returns) on the first inconsistency."""

    def __init__(self, raise_on_violation: bool = True):
        self._prev: Dict[int, List[int]] = {}
        self.raise_on_violation = raise_on_violation
        self.checked = 0

    @staticmethod
    def snapshot(buf) -> List[int]:
        """Host copy of ``buf``'s local signal pad (synchronises the current stream)."""
        import torch
        pad = buf.handle.get_signal_pad(buf.rank, (len(buf.epochs) * MAX_PEERS,), torch.uint32) \
            if hasattr(buf.handle, "get_signal_pad") else None
        if pad is None:
            raise RuntimeError("this torch build does not expose the symmetric-memory signal pad")
        torch.cuda.current_stream().synchronize()
        return [int(v) for v in pad.to(torch.int64).cpu().tolist()]

    def check(self, buf, quiescent: bool = False, pad: Optional[Sequence[int]] = None) -> List[Violation]:
        snap = list(pad) if pad is not None else self.snapshot(buf)
        v = check_pad(snap, buf.epochs, buf.world, prev=self._prev.get(id(buf)), quiescent=quiescent)
        self._prev[id(buf)] = snap
        self.checked += 1
        if v and self.raise_on_violation:
            raise AssertionError("symmetric-memory flag protocol violated:\n  " + "\n  ".join(map(str, v)))
        return v
