"""Host-side sharding arithmetic of the multi-GPU path.

Samples of every climber are split into contiguous ranges, one per rank; sample identity
(the Philox counter and the argmin index) is the GLOBAL index, so results do not depend on
the number of ranks.  Per iteration the ranks merge one packed 64-bit key per climber with a
MIN all-reduce: key = (order-preserving bits of the score) << 32 | global sample index, so
the lowest score wins and ties go to the lowest index (planning_state.cu:116-122: strict <).
"""
from __future__ import annotations

import numpy as np

KEY_NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def shard_range(m_global: int, rank: int, world: int) -> tuple[int, int]:
    """[offset, offset + count) of `rank`; the first m_global % world ranks get one extra."""
    if not (0 <= rank < world) or m_global < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(m_global, world)
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def key_hi(scores: np.ndarray, select_mode: int) -> np.ndarray:
    """CEIL_INT (0): int(ceilf(score)) saturating (generate_trajectories.cu:109);
    FLOAT (1): the IEEE bits of the (non-negative) score; NaN sorts last."""
    s = np.asarray(scores, np.float32)
    if select_mode == 0:
        c = np.ceil(s.astype(np.float64))
        hi = np.where(s < 2147483648.0, c, 2147483647.0)
        hi = np.where(np.isnan(s), 2147483647.0, hi)
        return hi.astype(np.int64).astype(np.uint32)
    return np.where(np.isnan(s), np.uint32(0xFFFFFFFF), s.view(np.uint32))


def pack_keys(scores: np.ndarray, sample_offset: int, select_mode: int) -> np.ndarray:
    """scores [k][m_local] -> per-climber minimum key [k] (uint64)."""
    s = np.atleast_2d(np.asarray(scores, np.float32))
    idx = np.arange(s.shape[1], dtype=np.uint64) + np.uint64(sample_offset)
    keys = (key_hi(s, select_mode).astype(np.uint64) << np.uint64(32)) | idx[None, :]
    return keys.min(axis=1) if s.shape[1] else np.full(s.shape[0], KEY_NONE, np.uint64)


def unpack_key(key) -> tuple[int, int]:
    key = int(key)
    return key >> 32, key & 0xFFFFFFFF


def connect_peers(ctx, rank: int, world: int, device) -> bool:
    """Maps the ranks' argmin mailboxes into each other (doapp_peer_export / doapp_peer_import):
    gathers the CUDA IPC handles over torch.distributed and synchronises once.  Returns False
    (the NCCL per-iteration path stays in use) if peer mapping is not possible on this box."""
    import torch
    import torch.distributed as dist
    ok = 1
    try:
        handle = ctx.peer_export()
    except Exception:  # noqa: BLE001 - any failure means "no peer path"
        handle, ok = bytes(64), 0
    mine = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(device)
    table = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(table, mine)
    flag = torch.tensor([ok], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 1:
        try:
            ctx.peer_import(b"".join(bytes(t.cpu().numpy().tobytes()) for t in table), rank, world)
        except Exception:  # noqa: BLE001
            ok = 0
    else:
        ok = 0
    flag = torch.tensor([ok], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)  # all ranks take the same path, and a barrier
    if int(flag.item()) != 1 and ok == 1:
        ctx.peer_disable()
    return int(flag.item()) == 1
