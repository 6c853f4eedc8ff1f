"""C2 on N ranks (torchrun): persistent planner with the peer mailboxes at several speculation
depths; prints us/iteration and the first CTA's phase stamps (DOAPP_P2_TRACE=1)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from doapp_b200 import capi, sharding
    from tests import synth
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    g = synth.box_grid(128, 0.01)
    for depth in sys.argv[1:] or ["6", "10", "15"]:
        os.environ["DOAPP_PERSIST_DEPTH"] = depth
        ctx = capi.Context(k=1, m=1024, n=64, device=lr, sigma=0.05, select_mode=1,
                           sample_offset=rank * 1024, global_samples_per_climber=1024 * world)
        uid = torch.zeros(capi.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(uid, 0)
        ctx.comm_init(uid.cpu().numpy().tobytes(), rank, world)
        ctx.set_grid(g, 0.01)
        sharding.connect_peers(ctx, rank, world, dev)
        init = capi.plan_lerp(synth.START, synth.GOALS[1], 64)
        ctx.plan(synth.START, synth.GOALS[1], init, 100)
        for _ in range(3):
            ctx.plan_bench(100, 1)
        dist.barrier()
        ms = min(ctx.plan_bench(100, 1) for _ in range(10))
        if rank == 0:
            print("world", world, "depth", depth, "us/iter", ms * 10, "mode", ctx.plan_mode(), flush=True)
            st = ctx.plan_stats()
            print("  stats", st, flush=True)
        else:
            ctx.plan_stats() if False else None
        dist.barrier()
        ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
