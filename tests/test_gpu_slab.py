"""vxb_slab_* / vxb_group_*: the update step sharded by x-slab with a peer-mapped ghost column (csrc/slab.cu)
gives exactly the single-world result — against the oracle on random worlds, against a single ctx on a
generated world, in one process (vxb_group) and over two processes (file descriptors over a unix socket).

Slabs may share a device, so everything here also runs on a one-GPU box; with two GPUs the in-process
group is spread over both."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import CpuWorld

pytestmark = pytest.mark.gpu

TICKS = 16


def _devices(n):
    import torch

    k = torch.cuda.device_count()
    return [r % k for r in range(n)]


def _oracle_world(rs, sp, seed):
    n = rs[0] * rs[1] * rs[2]
    w = CpuWorld(rs, sp, 2 + n)
    chunks = w.load_all_chunks()
    w.pool[2:] = scenes.random_pool(2 + n, seed, water=0.2, air=0.5)
    return w, chunks


def _fill_slab(slab, oracle_pool, rs, x0):
    """Table: every chunk of the local window loaded at its positional slot; pool: the oracle's chunks of the own columns."""
    pc = rs[1] * rs[2]
    lrs = slab.local_render_size
    table = np.zeros(lrs, np.uint32)
    lsp = slab.local_start_pos
    for lx in range(lrs[0]):
        for y in range(rs[1]):
            for z in range(rs[2]):
                table[(lsp[0] + lx) % lrs[0], (lsp[1] + y) % lrs[1], (lsp[2] + z) % lrs[2]] = slab.slot(lx, y, z)
    slab.write_table(table)
    own = oracle_pool[2 + x0 * pc:2 + (x0 + slab.cols) * pc]
    slab.write_pool(int(slab.info.slot_first), own[:pc])
    if slab.cols > 1:
        slab.write_pool(int(slab.info.slot_rest), own[pc:])
    chunks = [(lsp[0] + lx, lsp[1] + y, lsp[2] + z) for lx in range(slab.cols) for y in range(rs[1]) for z in range(rs[2])]
    slab.update_list_upload(scenes.update_list(chunks))


def _read_own(slab):
    pc = slab.plane_chunks
    parts = [slab.read_pool(int(slab.info.slot_first), pc)]
    if slab.cols > 1:
        parts.append(slab.read_pool(int(slab.info.slot_rest), (slab.cols - 1) * pc))
    return np.concatenate(parts)


@pytest.mark.parametrize("splits,variant", [([0, 2, 4], 2), ([0, 2, 3, 5], 0), ([0, 1, 2, 3, 5], 2)])
def test_group_matches_the_oracle_on_random_worlds(splits, variant):
    """Uneven slabs, including one-column slabs whose first column is also their last."""
    import voxel_game_b200 as vg

    rs, sp = (splits[-1], 2, 3), (100, 50, 70)
    n_slabs = len(splits) - 1
    g, chunks = _oracle_world(rs, sp, 11 + n_slabs)
    gs = vg.GameSettings(render_size=rs, max_update_rate=rs[0] * rs[1] * rs[2])
    group = vg.SlabGroup(gs, sp, _devices(n_slabs), col_splits=splits)
    for r, slab in enumerate(group.slabs):
        assert slab.cols == splits[r + 1] - splits[r] and slab.ghost == (1 if r + 1 < n_slabs else 0)
        slab.set_update_kernel(variant)
        _fill_slab(slab, g.pool, rs, splits[r])
    group.synchronize()
    lst = scenes.update_list(chunks)
    pc = rs[1] * rs[2]
    for t in range(TICKS):
        off = scenes.PHASES[t % 8] if t < 12 else scenes.PHASES[[1, 3, 0, 2][t - 12]]  # also x == 1 twice and x == 0 twice in a row
        group.update(off)
        ref_flags = g.voxel_update(lst, off)
        if t % 5 == 4:
            group.synchronize()
            got = np.concatenate([s.update_list_flags() for s in group.slabs])
            assert np.array_equal(got, ref_flags), "flags, tick %d" % t
    group.synchronize()
    got = np.concatenate([_read_own(s) for s in group.slabs])
    assert np.array_equal(got, g.pool[2:])
    assert len(got) == rs[0] * pc
    group.close()


def test_group_populate_matches_a_single_ctx():
    """Natural world (forced = false): all-air chunks collapse to table entry 1, so the ghost column's table has to
    follow the neighbour's (publish / sync)."""
    import voxel_game_b200 as vg

    rs = (6, 4, 3)
    sp = vg.initial_start_pos((10000, 1812, 10000), rs)
    n = rs[0] * rs[1] * rs[2]
    gs = vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * n, max_update_rate=n)
    single = vg.VoxelComputePipeline(gs)
    chunks = [(sp[0] + x, sp[1] + y, sp[2] + z) for x in range(rs[0]) for y in range(rs[1]) for z in range(rs[2])]
    from tests.oracle_lib import chunk_loads_for

    single.dispatch_worldgen(sp, chunk_loads_for(chunks, range(2, 2 + n)))
    table = single.read_table()
    loaded = [c for c in chunks if table[c[0] % rs[0], c[1] % rs[1], c[2] % rs[2]] >= 2]
    assert 0 < len(loaded) < n, "the window must straddle the surface for this test to mean anything"
    single.update_list_upload(scenes.update_list(loaded))
    group = vg.SlabGroup(gs, sp, _devices(3))
    group.populate(forced=False)
    for t in range(TICKS):
        single.update_list_run(sp, scenes.PHASES[t % 8])
        group.update(scenes.PHASES[t % 8])
    group.synchronize()
    for r, slab in enumerate(group.slabs):
        lt = slab.read_table()
        lrs, lsp = slab.local_render_size, slab.local_start_pos
        for lx in range(slab.cols + slab.ghost):
            for y in range(rs[1]):
                for z in range(rs[2]):
                    c = (lsp[0] + lx, lsp[1] + y, lsp[2] + z)
                    e_single = int(table[c[0] % rs[0], c[1] % rs[1], c[2] % rs[2]])
                    e_slab = int(lt[c[0] % lrs[0], c[1] % lrs[1], c[2] % lrs[2]])
                    assert min(e_single, 2) == min(e_slab, 2), (r, lx, y, z)
                    if e_slab >= 2:
                        assert e_slab == slab.slot(lx, y, z)
                        if lx < slab.cols:
                            assert np.array_equal(slab.read_pool(e_slab, 1), single.read_pool(e_single, 1)), (r, lx, y, z)
    group.close()
    single.close()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
rank, n, port, out_dir = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
import torch, torch.distributed as dist
import voxel_game_b200 as vg
from voxel_game_b200.sharding import exchange_slab_fds
from tests import scenes
from tests.test_gpu_slab import _oracle_world, _fill_slab, _read_own, TICKS
os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", port
dist.init_process_group("gloo", rank=rank, world_size=n)
rs, sp = (4, 2, 3), (100, 50, 70)
g, _ = _oracle_world(rs, sp, 23)
dev = rank % torch.cuda.device_count()
slab = vg.SlabPipeline(vg.GameSettings(render_size=rs, max_update_rate=64), sp, dev, rank, n)
_fill_slab(slab, g.pool, rs, rank * slab.cols)
exchange_slab_fds(slab, port)
slab.synchronize()
dist.barrier()
flags = []
for t in range(TICKS):
    slab.update(scenes.PHASES[t % 8])
    flags.append(slab.update_list_flags())
slab.status()
dist.barrier()
np.savez(os.path.join(out_dir, "rank%d.npz" % rank), pool=_read_own(slab), flags=np.stack(flags))
dist.barrier()
slab.close()
dist.destroy_process_group()
"""


@pytest.mark.timeout(600)
def test_two_processes_over_file_descriptors(tmp_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = str(_free_port())
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    procs = [subprocess.Popen([sys.executable, str(script), root, str(r), "2", port, str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=540)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    rs, sp = (4, 2, 3), (100, 50, 70)
    g, chunks = _oracle_world(rs, sp, 23)
    lst = scenes.update_list(chunks)
    ref_flags = np.stack([g.voxel_update(lst, scenes.PHASES[t % 8]) for t in range(TICKS)])
    r0, r1 = (np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(2))
    assert np.array_equal(np.concatenate([r0["pool"], r1["pool"]]), g.pool[2:])
    assert np.array_equal(np.concatenate([r0["flags"], r1["flags"]], axis=1), ref_flags)


def test_two_devices_in_one_process():
    """the update kernel's one-time set-up (shared-memory opt-in, grid size) is per device."""
    import torch

    import voxel_game_b200 as vg

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    rs, sp = (2, 2, 2), (100, 50, 70)
    g, chunks = _oracle_world(rs, sp, 5)
    lst = scenes.update_list(chunks)
    ctxs = [vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=10, max_update_rate=8), device=d) for d in (0, 1)]
    for c in ctxs:
        c.write_table(g.table)
        c.write_pool(2, g.pool[2:])
    for off in scenes.PHASES:
        ref = g.voxel_update(lst, off)
        for c in ctxs:
            assert np.array_equal(c.dispatch_voxel_update(sp, lst, off), ref)
    for c in ctxs:
        assert np.array_equal(c.read_pool(0, 10), g.pool)
        c.close()
