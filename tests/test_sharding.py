"""Multi-rank host logic on CPU: shard ranges, key packing, and a world_size-2 gloo run in which
every rank costs its slice with the oracle and one MIN all-reduce merges the argmin — the result
must equal the single-rank answer."""
import os
import socket

import numpy as np
import pytest

from doapp_b200 import sharding
from tests import synth


def test_shard_ranges_cover_exactly():
    for mg in (0, 1, 7, 1024, 1000003):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(mg, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == mg
            for (o0, c0), (o1, _) in zip(spans, spans[1:]):
                assert o0 + c0 == o1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def test_key_order_matches_host_argmin():
    rng = np.random.default_rng(0)
    s = rng.uniform(0, 50, (4, 300)).astype(np.float32)
    s[:, 17] = s[:, 200] = s.min(axis=1)  # ties: the lowest index must win
    for mode in (0, 1):
        keys = sharding.pack_keys(s, 1000, mode)
        for c in range(4):
            hi, idx = sharding.unpack_key(keys[c])
            ref = np.ceil(s[c]) if mode == 0 else s[c]
            assert idx - 1000 == int(np.argmin(ref))  # np.argmin: first lowest, like strict <
    assert sharding.key_hi(np.array([np.nan, 3e9], np.float32), 0).tolist() == [2147483647] * 2


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    k, mg, n = 2, 101, 32  # ragged: 101 samples over 2 ranks
    off, cnt = sharding.shard_range(mg, rank, world)
    orc = oracle.Oracle(k=k, m=cnt, n=n, sigma=0.1, sample_offset=off, m_global=mg, select_mode=1)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    base = np.stack([oracle.plan_lerp(synth.START, synth.GOALS[c], n) for c in range(k)])
    r = orc.cost_eval(base, synth.START, None, iteration=2)
    keys = sharding.pack_keys(r["scores"], off, 1)
    t = torch.from_numpy(keys.astype(np.int64))  # scores >= 0: keys < 2^63, order preserved
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    q.put((rank, t.numpy().astype(np.uint64).tolist(), r["scores"].tolist(), off))
    dist.destroy_process_group()


def test_world2_gloo_argmin_merge(oracle_mod):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1]  # every rank knows the winner
    # single-rank answer over all 101 samples
    k, mg, n = 2, 101, 32
    orc = oracle_mod.Oracle(k=k, m=mg, n=n, sigma=0.1, select_mode=1)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    base = np.stack([oracle_mod.plan_lerp(synth.START, synth.GOALS[c], n) for c in range(k)])
    full = orc.cost_eval(base, synth.START, None, iteration=2)
    merged = np.concatenate([np.array(res[0][2]), np.array(res[1][2])], axis=1).astype(np.float32)
    assert np.array_equal(merged.view(np.uint32), full["scores"].view(np.uint32))  # shard-invariant noise
    for c in range(k):
        hi, idx = sharding.unpack_key(res[0][1][c])
        assert idx == int(np.argmin(full["scores"][c]))
        assert np.uint32(hi).view(np.float32) == full["scores"][c].min()
    best_c = int(np.argmin([np.uint32(sharding.unpack_key(x)[0]).view(np.float32) for x in res[0][1]]))
    assert full["argmin"] == best_c * mg + sharding.unpack_key(res[0][1][best_c])[1]
