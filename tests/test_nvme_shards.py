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
