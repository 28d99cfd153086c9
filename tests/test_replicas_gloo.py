"""CPU, world_size 2 over gloo: the replicated-world exchanges of voxel_game_b200/replicas.py.
 * sharded update + replicas: each rank updates its x-slab with the oracle (halo exchange as in
   test_sharding_gloo), then every rank's changed chunks are gathered into every replica; after every
   tick both replicas must equal the single-world state.
 * sharded collision batch: each rank runs the oracle's compute_projectile on its half of a batch against
   its replica, the voxel damage it added is exchanged and replayed; both replicas must equal the
   single-world result of the whole batch."""
import os
import socket

import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import CpuWorld

RS, SP = (4, 2, 3), (100, 50, 70)
TICKS = 24


def _global_world(seed):
    """A mostly settled world (saturated air) with a few stones and water cells at chunk corners, faces and on
    the slab boundary: distance fields and water then change a few chunks per tick, often only the neighbour
    of the work item that made the change — what the change-flag dilation has to get right."""
    n = RS[0] * RS[1] * RS[2]
    w = CpuWorld(RS, SP, 2 + n)
    chunks = w.load_all_chunks()
    w.pool[2:] = 0x00FFFFFF
    grid = w.pool[2:].reshape(RS[0], RS[1], RS[2], 16, 16, 16)
    stone, water = np.uint32(1 << 24), np.uint32(7 << 24)
    grid[1, 0, 1, 15, 15, 15] = stone   # corner: touches 8 chunks, on the slab boundary (x = 1 | 2)
    grid[2, 1, 0, 0, 0, 0] = stone      # first voxel of rank 1's first column
    grid[0, 0, 2, 8, 15, 0] = stone
    grid[3, 1, 2, 15, 3, 15] = stone
    grid[1, 1, 1, 15, 8, 7] = water     # falls and spreads across the slab boundary
    grid[2, 1, 1, 0, 9, 7] = water
    grid[0, 1, 0, 5, 0, 5] = water      # falls into the chunk below
    rng = np.random.RandomState(seed)
    for _ in range(6):
        c = tuple(rng.randint(0, r) for r in RS)
        grid[c][tuple(rng.randint(0, 16, 3))] = stone
    return w, chunks


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _batch(seed, n=600):
    lo = (16 * SP[0] + 2, 16 * SP[1] + 2, 16 * SP[2] + 2)
    hi = (16 * (SP[0] + RS[0]) - 2, 16 * (SP[1] + RS[1]) - 2, 16 * (SP[2] + RS[2]) - 2)
    pr = scenes.random_projectiles(n, lo, hi, seed=seed, bounce_frac=0.0)
    pr["damage"] = np.array([1, 2, 5], np.float32)[np.arange(n) % 3]
    return pr


def _worker(rank, world_size, port, out_dir):
    import torch
    import torch.distributed as dist

    from voxel_game_b200.replicas import changed_chunk_mask, exchange_damage, exchange_damage_fixed, exchange_ghost_mask, gather_changed_chunks
    from voxel_game_b200.sharding import SlabShard, halo_exchange

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        g, chunks = _global_world(5)
        n_total = len(chunks)
        replica = CpuWorld(RS, SP, 2 + n_total)
        replica.load_all_chunks()
        replica.pool[:] = g.pool
        sh = SlabShard(RS, SP, rank, world_size)
        w = CpuWorld(sh.local_render_size, sh.local_start_pos, 2 + sh.n_local)
        local_chunks = w.load_all_chunks()
        pc = sh.plane_chunks
        w.pool[2:] = g.pool[2 + sh.x0 * pc:2 + (sh.x0 + sh.cols + sh.ghost) * pc]
        pool_t = torch.from_numpy(w.pool.view(np.int32))
        first, ghost = sh.views(pool_t)
        sendbuf = torch.empty((pc, 256), dtype=torch.int32)
        recvbuf = torch.empty((pc, 256), dtype=torch.int32)
        lst = scenes.update_list(local_chunks[:sh.n_own])
        own_pool = pool_t[2:2 + sh.n_own]
        replica_t = torch.from_numpy(replica.pool.view(np.int32))[2:]
        idx_buf = torch.empty(sh.n_own, dtype=torch.int64)
        chunk_buf = torch.empty((sh.n_own, 4096), dtype=torch.int32)
        states, moved = [], 0
        for t in range(TICKS):
            off = scenes.PHASES[t % 8]
            if off[0] == 1:
                halo_exchange(dist, sh, first, ghost, sendbuf, recvbuf, pre=True)
            flags = w.voxel_update(lst, off)
            if off[0] == 1:
                halo_exchange(dist, sh, first, ghost, sendbuf, recvbuf, pre=False)
            mask, ghost_mask = changed_chunk_mask(torch, torch.from_numpy(flags.astype(np.int32)), sh.cols, RS[1], RS[2], off, bool(sh.ghost))
            if off[0] == 1:
                got = exchange_ghost_mask(dist, torch, ghost_mask, rank, world_size, pc, torch.device("cpu"))
                if got is not None:
                    mask[:pc] |= got
            moved += gather_changed_chunks(dist, torch, mask, own_pool, sh.x0 * pc, replica_t, idx_buf, chunk_buf, rank, world_size)
            states.append(replica.pool[2:].copy())

        # ---- sharded collision batch on the (now identical) replicas ----
        pr = _batch(17)
        half = len(pr) // world_size
        mine = pr[rank * half:(rank + 1) * half]
        before = replica.pool.copy()
        out_pr, _, _ = replica.projectiles(mine, np.zeros(0, mine.dtype) if False else scenes.players_grid(1, (0, 0, 0), 1.0, -1000.0, seed=1)[:0], capacity=16)
        diff = np.argwhere(replica.pool != before)
        rec = np.zeros((max(len(diff), 1) + 64, 4), np.int32)
        for i, (s, k) in enumerate(diff):
            rec[i] = (s, k, 0, np.int64(replica.pool[s, k]) - np.int64(before[s, k]))

        def apply_fn(rows):
            for s, k, _, amt in rows.numpy():
                replica.pool[s, k] = np.uint32((int(replica.pool[s, k]) + int(amt)) & 0xFFFFFFFF)

        applied = exchange_damage(dist, torch, torch.from_numpy(rec), len(diff), rank, world_size, apply_fn)
        # the fixed-capacity variant (counts stay "on the device") must replay the same rows: undo, redo, compare
        after = replica.pool.copy()
        replica.pool[:] = before
        for s_, k_, _, amt in rec[:len(diff)]:
            replica.pool[s_, k_] = np.uint32((int(replica.pool[s_, k_]) + int(amt)) & 0xFFFFFFFF)
        exchange_damage_fixed(dist, torch, torch.from_numpy(rec), torch.tensor([len(diff)], dtype=torch.int32), rank, world_size,
                              lambda rows, cnt: apply_fn(rows[:int(cnt[0])]))
        assert np.array_equal(replica.pool, after)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), states=np.stack(states), moved=moved, final=replica.pool, applied=applied,
                 n_rec=len(diff), health=out_pr["health"])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_replicas_follow_the_single_world(tmp_path):
    import torch.multiprocessing as mp

    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    g, chunks = _global_world(5)
    lst = scenes.update_list(chunks)
    r0 = np.load(os.path.join(str(tmp_path), "rank0.npz"))
    r1 = np.load(os.path.join(str(tmp_path), "rank1.npz"))
    changed_total = 0
    for t in range(TICKS):
        f = g.voxel_update(lst, scenes.PHASES[t % 8])
        changed_total += int((f & 1).sum())
        assert np.array_equal(r0["states"][t], g.pool[2:]), "replica 0 differs after tick %d" % t
        assert np.array_equal(r1["states"][t], g.pool[2:]), "replica 1 differs after tick %d" % t
    # the refresh is selective: it moved at least the flagged chunks and not everything
    assert int(r0["moved"]) == int(r1["moved"]) and 0 < changed_total <= int(r0["moved"]) < TICKS * len(chunks)
    # the whole batch on the single world
    pr = _batch(17)
    out_pr, _, _ = g.projectiles(pr, scenes.players_grid(1, (0, 0, 0), 1.0, -1000.0, seed=1)[:0], capacity=16)
    assert np.array_equal(r0["final"], g.pool) and np.array_equal(r1["final"], g.pool)
    assert int(r0["applied"]) == int(r1["n_rec"]) and int(r1["applied"]) == int(r0["n_rec"]) and int(r0["n_rec"]) + int(r1["n_rec"]) > 0
    assert np.array_equal(np.concatenate([r0["health"], r1["health"]]), out_pr["health"])
