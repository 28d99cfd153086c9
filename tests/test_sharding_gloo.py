"""CPU, world_size 2 over gloo: the x-slab sharding + halo exchange protocol of
voxel_game_b200/sharding.py gives exactly the single-world result. Each rank drives the oracle on
its slab (own columns + one ghost column); the bench uses the same module with CUDA tensors / NCCL."""
import os
import socket

import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import CpuWorld

RS, SP = (4, 2, 3), (100, 50, 70)
TICKS = 16


def _global_world(seed):
    n = RS[0] * RS[1] * RS[2]
    w = CpuWorld(RS, SP, 2 + n)
    chunks = w.load_all_chunks()
    w.pool[2:] = scenes.random_pool(2 + n, seed, water=0.2, air=0.5)
    return w, chunks


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world_size, port, out_dir):
    import torch
    import torch.distributed as dist

    from voxel_game_b200.sharding import SlabShard, halo_exchange

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        g, _ = _global_world(11)
        sh = SlabShard(RS, SP, rank, world_size)
        w = CpuWorld(sh.local_render_size, sh.local_start_pos, 2 + sh.n_local)
        local_chunks = w.load_all_chunks()
        pc = sh.plane_chunks
        # global slots are in (x,y,z) order too: local column c is global column x0 + c
        w.pool[2:] = g.pool[2 + sh.x0 * pc:2 + (sh.x0 + sh.cols + sh.ghost) * pc]
        pool_t = torch.from_numpy(w.pool.view(np.int32))
        first, ghost = sh.views(pool_t)
        sendbuf = torch.empty((pc, 256), dtype=torch.int32)
        recvbuf = torch.empty((pc, 256), dtype=torch.int32)
        lst = scenes.update_list(local_chunks[:sh.n_own])
        all_flags = []
        for t in range(TICKS):
            off = scenes.PHASES[t % 8]
            if off[0] == 1:
                halo_exchange(dist, sh, first, ghost, sendbuf, recvbuf, pre=True)
            all_flags.append(w.voxel_update(lst, off))
            if off[0] == 1:
                halo_exchange(dist, sh, first, ghost, sendbuf, recvbuf, pre=False)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), pool=w.pool[2:2 + sh.n_own], flags=np.stack(all_flags))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_halo_exchange_equals_single_world(tmp_path):
    import torch.multiprocessing as mp

    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    g, chunks = _global_world(11)
    lst = scenes.update_list(chunks)
    ref_flags = []
    for t in range(TICKS):
        ref_flags.append(g.voxel_update(lst, scenes.PHASES[t % 8]))
    ref_flags = np.stack(ref_flags)
    r0 = np.load(os.path.join(str(tmp_path), "rank0.npz"))
    r1 = np.load(os.path.join(str(tmp_path), "rank1.npz"))
    n_half = len(chunks) // 2
    assert np.array_equal(np.concatenate([r0["pool"], r1["pool"]]), g.pool[2:])
    assert np.array_equal(np.concatenate([r0["flags"], r1["flags"]], axis=1), ref_flags)
    assert r0["pool"].shape[0] == n_half and (ref_flags & 1).any()


def test_slab_geometry():
    from voxel_game_b200.sharding import SlabShard

    s = [SlabShard((256, 16, 256), (497, 105, 497), r, 8) for r in range(8)]
    assert [x.cols for x in s] == [32] * 8 and [x.ghost for x in s] == [1] * 7 + [0]
    assert s[3].local_start_pos == (497 + 96, 105, 497) and s[3].local_render_size == (33, 16, 256)
    assert s[0].plane_chunks == 4096 and s[7].n_local == s[7].n_own == 32 * 4096
    with pytest.raises(ValueError):
        SlabShard((10, 1, 1), (0, 0, 0), 0, 4)
