"""``NvmeShardStore``: parameter shards in files behind a bounded set of host buffers (ZeRO-Infinity parameter offload)."""
import os

import torch

from shuffle_exchange_b200.runtime.zero.nvme_shards import NvmeShardStore


def test_store_round_trips_bounded_and_writes_back_only_what_changed(tmp_path):
    st = NvmeShardStore(str(tmp_path), torch.bfloat16, buffer_count=2)
    shards = {f"g{i}": torch.randn(1000 + 37 * i).to(torch.bfloat16) for i in range(5)}      # sizes not block-aligned
    for k, t in shards.items():
        st.put(k, t)
        assert os.path.getsize(st.path(k)) % 4096 == 0        # whole blocks on disk
    assert st.resident_count == 0
    w0 = st.writes
    for k, t in shards.items():                                 # read-only pass: nothing is written back
        assert torch.equal(st.get(k), t)
        assert st.resident_count <= 2
    assert st.writes == w0 and st.reads == 5
    a = st.get("g0")
    a.mul_(2)
    st.written("g0")
    for k in ("g1", "g2", "g3"):                                # pushes g0 out: it must be written first
        st.get(k)
    assert st.writes == w0 + 1
    assert torch.equal(st.get("g0"), shards["g0"] * 2)
    b = st.get("g4")
    b.zero_()
    st.written("g4")
    st.release_all()                                            # flush + drop
    assert st.resident_count == 0 and torch.count_nonzero(st.get("g4")) == 0
    st.put("g1", torch.ones(10, dtype=torch.bfloat16))          # replacing a shard changes its length
    assert st.get("g1").numel() == 10 and float(st.get("g1").sum()) == 10.0


def test_max_in_cpu_lets_more_shards_stay_resident(tmp_path):
    """``offload_param.max_in_cpu``: beyond the first ``buffer_count`` shards, shards stay in host memory while the
    resident elements fit the budget; the budget is a hard bound for those extra shards."""
    st = NvmeShardStore(str(tmp_path), torch.float32, buffer_count=1, max_in_cpu=2500)
    for i in range(4):
        st.put(f"g{i}", torch.full((1000,), float(i)))
    for i in range(4):
        st.get(f"g{i}")
        assert st.resident_numel <= 2500 or st.resident_count == 1
    assert st.resident_count == 2 and st.reads == 4              # two 1000-element shards fit 2500, a third does not
    r0 = st.reads
    assert float(st.get("g3")[0]) == 3.0 and float(st.get("g2")[0]) == 2.0 and st.reads == r0     # both served from memory
    big = NvmeShardStore(str(tmp_path / "b"), torch.float32, buffer_count=1, max_in_cpu=10 ** 9)
    for i in range(4):
        big.put(f"g{i}", torch.zeros(1000))
        big.get(f"g{i}")
    assert big.resident_count == 4                                # the reference's default budget keeps small models in RAM
