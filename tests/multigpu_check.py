"""Run under torchrun with N ranks (one per GPU): the sharded planner (NCCL MIN all-reduce of the
per-climber argmin keys, grid broadcast from rank 0) must give bit-identical results to the
single-GPU run over the same global sample set and to the oracle.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multigpu_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from doapp_b200 import capi, sharding
    from tests import synth
    import oracle

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    # (update rule, peer mailboxes mapped, shape): NCCL per iteration; the persistent planner
    # merging keys in-kernel; the weighted rule; and a problem too large for the persistent
    # planner, whose cost launches trade keys through the mailboxes (xchg_send_kernel)
    small, large = (2, 1000, 64, 8), (2, 40001, 32, 7)
    for mode_kw, peers, shape in ((dict(), False, small), (dict(), True, small),
                                  (dict(update_mode=1, update_gain=1e-3), False, small),
                                  (dict(), True, large), (dict(), False, large)):
        check(rank, world, lr, dev, mode_kw, peers, shape)
        if peers:  # twice more: the mailbox sets alternate per plan
            check(rank, world, lr, dev, mode_kw, peers, shape, plans=3)
    dist.barrier()
    dist.destroy_process_group()


def check(rank, world, lr, dev, mode_kw, peers, shape, plans=1):
    """keep-best (8*k-byte MIN all-reduce per iteration) or weighted update (plus one exact
    int64 SUM all-reduce of the fixed-point update-vector sums)."""
    import torch
    import torch.distributed as dist
    from doapp_b200 import capi, sharding
    from tests import synth
    import oracle
    k, mg, n, iters = shape
    off, cnt = sharding.shard_range(mg, rank, world)
    ctx = capi.Context(k=k, m=cnt, n=n, device=lr, sigma=0.05, select_mode=1, sample_offset=off,
                       global_samples_per_climber=mg, **mode_kw)
    uid = torch.zeros(capi.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
    if rank == 0:
        uid = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).to(dev)
    dist.broadcast(uid, 0)
    ctx.comm_init(uid.cpu().numpy().tobytes(), rank, world)
    grid = synth.box_grid(128, 0.01)
    if rank == 0:
        ctx.set_grid(grid, 0.01)
    else:
        ctx.grid_alloc(capi.GRID_U8, capi.FRAME_CDG, 128, 128, 128, 0.01)
    ctx.grid_broadcast(0)
    peer_ok = sharding.connect_peers(ctx, rank, world, dev) if peers else False
    assert np.array_equal(ctx.grid_download((128, 128, 128)), grid), "grid broadcast differs"
    start, goal = synth.START, synth.GOALS[1]
    init = capi.plan_lerp(start, goal, n)
    for _ in range(plans):
        r = ctx.plan(start, goal, init, iters)
    st = ctx.plan_state()
    mode = ctx.plan_mode()
    ctx.close()
    # every rank holds the same kept trajectories
    t = torch.from_numpy(st["trajectories"].copy()).to(dev)
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi), "ranks disagree on the kept trajectories"
    if rank == 0:
        one = capi.Context(k=k, m=mg, n=n, device=lr, sigma=0.05, select_mode=1, **mode_kw)
        one.set_grid(grid, 0.01)
        r1 = one.plan(start, goal, init, iters)
        s1 = one.plan_state()
        one.close()
        assert np.array_equal(s1["trajectories"].view(np.uint32), st["trajectories"].view(np.uint32))
        assert np.float32(r1["best_score"]) == np.float32(r["best_score"])
        orc = oracle.Oracle(k=k, m=mg, n=n, sigma=0.05, select_mode=1, **mode_kw)
        orc.set_grid(grid, 0.01)
        ro = orc.plan(start, init, iters, threads=8)
        assert np.array_equal(ro["trajectories"].view(np.uint32), st["trajectories"].view(np.uint32))
        assert np.array_equal(ro["best_traj"].view(np.uint32), r["best_traj"].view(np.uint32))
        print(f"multigpu_check ok: world={world} shape={shape} plans={plans} mode={mode_kw} peer_mailboxes={peer_ok} plan_mode={mode} best_score={r['best_score']:.6f} "
              f"found_better={st['found_better'].tolist()}", flush=True)
    dist.barrier()


if __name__ == "__main__":
    main()
