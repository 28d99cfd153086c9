"""GPU: size-independent properties at sizes the oracle cannot reach, through the C ABI.
 - two independent kernels (tiled/pipelined vs reference-shaped gather) agree on a large world,
 - the update is independent of work-list order (checksum of checksums),
 - worldgen is deterministic, a settled waterless world is a fixed point with quiet flags,
 - the x-slab sharding with ghost columns (two ctxs on one GPU, halo planes copied directly)
   reproduces the single-ctx world."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import chunk_loads_for

pytestmark = pytest.mark.gpu


def _grid(rs, sp):
    return [(sp[0] + x, sp[1] + y, sp[2] + z) for x in range(rs[0]) for y in range(rs[1]) for z in range(rs[2])]


def _make(rs, sp, forced=True, world_gen=0, seed=0):
    import voxel_game_b200 as vg

    n = rs[0] * rs[1] * rs[2]
    gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * n, max_update_rate=n,
                                                   world_gen=vg.WorldGen(world_gen), world_gen_seed=seed))
    chunks = _grid(rs, sp)
    gpu.dispatch_worldgen(sp, chunk_loads_for(chunks, range(2, 2 + n), forced=forced), want_results=False)
    return gpu, chunks


def test_two_update_kernels_agree_on_a_large_world():
    rs, sp = (24, 8, 24), (613, 108, 613)  # 4608 chunks, 18.9 M voxels around the spawn
    a, chunks = _make(rs, sp)
    b, _ = _make(rs, sp)
    assert a.world_checksum(sp) == b.world_checksum(sp)  # worldgen is deterministic
    b.set_update_kernel(1)
    lst = scenes.update_list(chunks)
    a.update_list_upload(lst)
    b.update_list_upload(lst)
    seen = set()
    for t in range(24):
        off = scenes.PHASES[t % 8]
        a.update_list_run(sp, off)
        b.update_list_run(sp, off)
        assert np.array_equal(a.update_list_flags(), b.update_list_flags()), "tick %d" % t
        ca, cb = a.world_checksum(sp), b.world_checksum(sp)
        assert ca == cb, "tick %d" % t
        seen.add(ca)
    assert len(seen) > 8  # the world really evolves
    n = len(chunks)
    assert np.array_equal(a.read_pool(2, n), b.read_pool(2, n))
    a.close(); b.close()


def test_update_is_independent_of_list_order_at_scale():
    rs, sp = (16, 8, 16), (617, 108, 617)
    a, chunks = _make(rs, sp)
    b, _ = _make(rs, sp)
    lst = scenes.update_list(chunks)
    perm = np.random.RandomState(0).permutation(len(lst))
    for t in range(8):
        off = scenes.PHASES[t]
        fa = a.dispatch_voxel_update(sp, lst, off)
        fb = b.dispatch_voxel_update(sp, lst[perm], off)
        assert np.array_equal(fa[perm], fb)
        assert a.world_checksum(sp) == b.world_checksum(sp)
    a.close(); b.close()


def test_control_map_settles_to_a_fixed_point():
    # the control arena has no water: after enough ticks nothing changes any more
    rs, sp = (12, 4, 12), (619, 111, 619)
    g, chunks = _make(rs, sp, world_gen=1, seed=3646)
    lst = scenes.update_list(chunks)
    g.update_list_upload(lst)
    for t in range(80):
        g.update_list_run(sp, scenes.PHASES[t % 8])
    c0 = g.world_checksum(sp)
    for t in range(8):
        g.update_list_run(sp, scenes.PHASES[t])
        assert (g.update_list_flags() == 0).all()
    assert g.world_checksum(sp) == c0
    g.set_update_kernel(2)  # store elision must not change a settled world either
    for t in range(8):
        g.update_list_run(sp, scenes.PHASES[t])
    assert g.world_checksum(sp) == c0
    g.close()


def test_slab_sharding_with_ghost_columns_matches_single_ctx():
    import voxel_game_b200 as vg
    from voxel_game_b200.sharding import SlabShard

    torch = pytest.importorskip("torch")
    rs, sp = (8, 4, 6), (621, 110, 622)
    single, chunks = _make(rs, sp)
    lst = scenes.update_list(chunks)
    shards = [SlabShard(rs, sp, r, 2) for r in range(2)]
    ctxs, views, lists = [], [], []

    class DevView:
        def __init__(self, ptr, n_words):
            self.__cuda_array_interface__ = {"shape": (int(n_words),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}

    for sh in shards:
        g, lc = _make(sh.local_render_size, sh.local_start_pos)
        ctxs.append(g)
        lists.append(scenes.update_list(lc[:sh.n_own]))
        pool_t = torch.as_tensor(DevView(g.device_ptrs().voxel_pool, (2 + sh.n_local) * 4096), device="cuda").view(2 + sh.n_local, 4096)
        views.append(sh.views(pool_t))
    for t in range(16):
        off = scenes.PHASES[t % 8]
        if off[0] == 1:
            for g in ctxs:
                g.synchronize()
            views[0][1].copy_(views[1][0])  # pre: rank 1's first column -> rank 0's ghost
            torch.cuda.synchronize()
        single.dispatch_voxel_update(sp, lst, off, want_flags=False)
        flags = [g.dispatch_voxel_update(sh.local_start_pos, l, off) for g, sh, l in zip(ctxs, shards, lists)]
        if off[0] == 1:
            views[1][0].copy_(views[0][1])  # post: ghost -> owner
            torch.cuda.synchronize()
        assert len(flags[0]) + len(flags[1]) == len(lst)
    n0 = shards[0].n_own
    ref = single.read_pool(2, len(chunks))
    assert np.array_equal(ctxs[0].read_pool(2, n0), ref[:n0])
    assert np.array_equal(ctxs[1].read_pool(2, shards[1].n_own), ref[n0:])
    for g in ctxs + [single]:
        g.close()


@pytest.mark.timeout(900)
def test_config3_full_size_three_kernels_agree():
    """BASELINE configs[2] at its full size (256x16x256 chunks, 4.3 G voxels, 17 GB per world): the default kernel
    (store elision, runs of list neighbours with the z hand-over picked automatically at this list length), the
    same kernel with unconditional stores, and the reference-shaped gather kernel walk the same 16 ticks on
    three copies of the world; change flags and the world checksum must agree after every tick."""
    import torch

    import bench
    import voxel_game_b200 as vg

    free, _ = torch.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip("needs 3 x 17 GB of device memory")
    rs, sp = bench.C3_RS, bench.C3_SP
    n = rs[0] * rs[1] * rs[2]
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)
    worlds = []
    for variant in (2, 0, 1):
        g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
        bench.generate_world(g, sp, chunks, 2, True, torch)
        g.update_list_upload(lst)
        g.set_update_kernel(variant)
        worlds.append(g)
    c0 = [g.world_checksum(sp) for g in worlds]
    assert c0[0] == c0[1] == c0[2]
    seen, changed = set(), 0
    for t in range(16):
        off = scenes.PHASES[t % 8]
        flags, cks = [], []
        for g in worlds:
            g.update_list_run(sp, off)
            flags.append(g.update_list_flags())
            cks.append(g.world_checksum(sp))
        assert np.array_equal(flags[0], flags[1]) and np.array_equal(flags[0], flags[2]), "flags differ at tick %d" % t
        assert cks[0] == cks[1] == cks[2], "world differs at tick %d" % t
        seen.add(cks[0])
        changed += int((flags[0] & 1).sum())
    assert len(seen) == 16 and changed > n  # the world really evolves, everywhere
    for g in worlds:
        g.close()


@pytest.mark.timeout(600)
def test_pipelined_dispatch_equals_resident_list_run():
    """vxb_dispatch_voxel_update cuts a long host list into segments (upload, kernel and flag download overlap) while
    vxb_update_list_run sweeps the resident list in one launch: same flags, same world, tick by tick, on a list that does
    not divide evenly into segments and with both store policies."""
    import torch

    import bench
    import voxel_game_b200 as vg

    rs, sp = (72, 16, 256), bench.C3_SP  # 294 912 chunks: three segments
    n = rs[0] * rs[1] * rs[2]
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)[: n - 777]
    worlds = []
    for _ in range(2):
        g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
        bench.generate_world(g, sp, chunks, 2, True, torch)
        worlds.append(g)
    a, b = worlds
    b.update_list_upload(lst)
    changed = 0
    for t in range(12):
        variant = 2 if t < 8 else 0
        a.set_update_kernel(variant)
        b.set_update_kernel(variant)
        off = scenes.PHASES[t % 8]
        fa = a.dispatch_voxel_update(sp, lst, off)
        b.update_list_run(sp, off)
        fb = b.update_list_flags()
        assert np.array_equal(fa, fb), "flags differ at tick %d" % t
        assert a.world_checksum(sp) == b.world_checksum(sp), "world differs at tick %d" % t
        changed += int((fa & 1).sum())
    assert changed > n
    a.close(); b.close()


def test_segmented_resident_list_equals_one_launch():
    """A resident list of 65 536 entries or more is uploaded in segments and its first sweep is launched per segment
    (vxb_update_list_run); vxb_update_list_run_range waits for the whole list and sweeps it in ONE launch. Same flags, same world,
    for a fresh upload every tick (what a multi-GPU host does) and both store policies."""
    import torch

    import bench
    import voxel_game_b200 as vg

    rs, sp = (17, 16, 256), bench.C3_SP  # 69 632 chunks
    n = rs[0] * rs[1] * rs[2]
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)[: n - 5]
    worlds = []
    for _ in range(2):
        g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
        bench.generate_world(g, sp, chunks, 2, True, torch)
        worlds.append(g)
    a, b = worlds
    changed = 0
    for t in range(10):
        variant = 2 if t < 6 else 0
        off = scenes.PHASES[(t * 3) % 8]
        for g in (a, b):
            g.set_update_kernel(variant)
            g.update_list_upload(lst)
        a.update_list_run(sp, off)                      # per segment, as the entries arrive
        b.update_list_run_range(sp, off, 0, len(lst))   # joined, one launch
        fa, fb = a.update_list_flags(), b.update_list_flags()
        assert np.array_equal(fa, fb), "flags differ at tick %d" % t
        assert a.world_checksum(sp) == b.world_checksum(sp), "world differs at tick %d" % t
        changed += int((fa & 1).sum())
    assert changed > n
    a.close(); b.close()


def _config2_world(torch):
    """BASELINE configs[1] world as the benchmark builds it: 64x16x64 chunks, non-forced worldgen, 64 ticks."""
    import bench
    import voxel_game_b200 as vg

    rs, sp = bench.C2_RS, bench.C2_SP
    n = rs[0] * rs[1] * rs[2]
    g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024,
                                                max_collisions=1 << 20))
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    bench.generate_world(g, sp, chunks, 2, False, torch)
    table = g.read_table()
    g.update_list_upload(bench._loaded_list(table, chunks, rs))
    for t in range(64):
        g.update_list_run(sp, scenes.PHASES[t % 8])
    cx, cz = 16 * sp[0] + 8 * rs[0], 16 * sp[2] + 8 * rs[2]
    sy = bench._surface_from_gpu(g, table, rs, sp, cx, cz)
    return g, rs, sp, (cx, sy, cz)


@pytest.mark.timeout(900)
def test_config2_full_size_primary_rays_equal_explicit_rays():
    """BASELINE configs[1] at full size (1024x256x1024 voxels, 1920x1080): the per-pixel kernel that builds its
    rays from the camera and writes layers as it goes, and the explicit-ray kernel fed with the same rays built
    on the host in float32, must return the same layers, depths and iteration total — two code paths, 2 M rays."""
    import torch

    import voxel_game_b200 as vg

    g, rs, sp, (cx, sy, cz) = _config2_world(torch)
    w, h = 1920, 1080
    cam = scenes.camera((cx + 0.5, sy + 4.5, cz + 0.5), 0.7, -0.2)
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    g.raymarch(sp, cam, w, h, caps)
    a = g.ray_download()
    f = np.float32
    px, py = np.meshgrid(np.arange(w, dtype=f), np.arange(h, dtype=f))
    aspect = f(w) / f(h)
    sx = ((px + f(0.5)) / f(w) * f(2.0) - f(1.0)) * aspect
    syy = ((py + f(0.5)) / f(h) * f(2.0) - f(1.0)) * f(1.0)
    d, r, u = (cam[k][0, :3].astype(f) for k in ("dir", "right", "up"))
    v = (d[None, None, :] + sx[..., None] * r[None, None, :]) - syy[..., None] * u[None, None, :]
    ln = np.sqrt((v[..., 0] * v[..., 0] + v[..., 1] * v[..., 1]) + v[..., 2] * v[..., 2])
    rays = (v / ln[..., None]).reshape(-1, 3).astype(f)
    pos = np.broadcast_to(cam["pos"][0, :3].astype(f), rays.shape)
    g.raymarch_rays(sp, pos, rays, 0, 100)
    b = g.ray_download()
    assert a["total_iterations"] == b["total_iterations"] > 5 * w * h
    assert np.array_equal(a["layer_count"], b["layer_count"]) and (a["layer_count"] > 0).mean() > 0.5
    assert np.array_equal(a["dist"].view(np.uint32), b["dist"].view(np.uint32))
    lc = a["layer_count"]
    for l in range(5):
        m = lc > l
        if not m.any():
            continue
        for k in ("voxel", "leaving"):
            assert np.array_equal(a[k][l][m], b[k][l][m]), (k, l)
        for k in ("pos", "normal"):
            assert np.array_equal(a[k][l][:, m].view(np.uint32), b[k][l][:, m].view(np.uint32)), (k, l)
        assert np.array_equal(a["ldist"][l][m].view(np.uint32), b["ldist"][l][m].view(np.uint32))
    g.close()


@pytest.mark.timeout(900)
def test_config2_full_size_wavefront_render_equals_fused_render():
    """BASELINE configs[1]'s world at 1920x1080 through the whole fragment shader: the default wavefront of small kernels
    (per-layer rounds over compacted pixel lists) and the fused per-pixel kernel must write the same colour bits and walk the
    same number of DDA iterations, for a camera over land and one that looks across water and leaves (multi-layer pixels)."""
    import torch

    import voxel_game_b200 as vg

    g, rs, sp, (cx, sy, cz) = _config2_world(torch)
    w, h = 1920, 1080
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    for yaw, pitch, t in ((0.7, -0.2, 0.0), (3.9, -0.05, 2.5)):
        cam = scenes.camera((cx + 0.5, sy + 4.5, cz + 0.5), yaw, pitch)
        out = []
        for variant in (0, 1):
            g.set_render_kernel(variant)
            g.render(sp, cam, w, h, caps, time=t)
            out.append((g.color_download()[0].copy(), g.ray_download()["total_iterations"]))
        (a, ia), (b, ib) = out
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) and ia == ib > 30 * w * h
        assert np.isfinite(a).all() and (a[..., 3] == 1.0).all() and a[..., :3].std() > 0.02
    g.set_render_kernel(0)
    # the wavefront leaves the primary layers of the frame in the ray buffers
    lc = g.ray_download()["layer_count"]
    assert lc.shape == (w * h,) and (lc > 1).mean() > 0.01
    g.close()


@pytest.mark.timeout(900)
def test_config4_full_size_sharded_batch_equals_single_batch():
    """BASELINE configs[3] at full size (1 048 576 projectiles + 64 hulls against the config-2 world): the batch split
    over two replicas with the voxel damage exchanged (the multi-GPU protocol, here two contexts on one GPU) gives the
    projectiles, the collision set and the world of the unsplit batch."""
    import torch

    g_s, rs, sp, (cx, sy, cz) = _config2_world(torch)
    g_a = _config2_world(torch)[0]
    g_b = _config2_world(torch)[0]
    assert g_s.world_checksum(sp) == g_a.world_checksum(sp) == g_b.world_checksum(sp)
    n = 1 << 20
    lo = (16 * sp[0] + 8, 16 * sp[1] + 8, 16 * sp[2] + 8)
    hi = (16 * (sp[0] + rs[0]) - 8, 16 * (sp[1] + rs[1]) - 8, 16 * (sp[2] + rs[2]) - 8)
    pr = scenes.c4_projectiles(n, lo, hi)
    pl = scenes.players_grid(8, (cx, sy, cz), 48.0, sy + 4.0, seed=64)
    pr["pos"][:4000, :3] = pl["pos"][np.arange(4000) % 64, :3] + np.random.RandomState(3).randn(4000, 3).astype(np.float32) * 1.5 + (0, -1.5, 0)
    half = n // 2
    for g in (g_a, g_b):
        g.damage_log_enable(half)
    pa, _, ca, na = g_a.dispatch_entities(sp, projectiles=pr[:half], players=pl, run_players=False)
    pb, _, cb, nb = g_b.dispatch_entities(sp, projectiles=pr[half:], players=pl, run_players=False)
    ra, rb = g_a.damage_log_read(half), g_b.damage_log_read(half)
    g_a.damage_apply(sp, records=rb)
    g_b.damage_apply(sp, records=ra)
    ps, _, cs, ns = g_s.dispatch_entities(sp, projectiles=pr, players=pl, run_players=False)
    assert len(ra) + len(rb) > 50000 and ns > 100
    assert g_s.world_checksum(sp) == g_a.world_checksum(sp) == g_b.world_checksum(sp)
    assert np.array_equal(np.concatenate([pa, pb]).view(np.uint8), ps.view(np.uint8))
    cb = cb.copy()
    cb["id1"] += half
    merged = np.concatenate([ca, cb])
    merged = merged[np.lexsort((merged["id2"], merged["id1"]))]
    assert na + nb == ns and np.array_equal(merged[["id1", "id2", "properties"]], cs[["id1", "id2", "properties"]])
    for g in (g_s, g_a, g_b):
        g.close()


def test_device_replica_refresh_follows_the_world():
    """replicas.DeviceReplicaRefresh (mask, pack, unpack kernels; nothing read back) keeps a second copy of a world
    equal to the first through 24 ticks of sparse activity; its mask kernel equals the tensor version of the same
    rule, with and without a ghost column; an undersized buffer reports its overflow through the record count."""
    import torch

    import voxel_game_b200 as vg
    from voxel_game_b200.replicas import DeviceReplicaRefresh, changed_chunk_mask
    from voxel_game_b200.sharding import SlabShard

    rs, sp = (6, 4, 5), (300, 40, 70)
    n = rs[0] * rs[1] * rs[2]

    # world A (updated) and replica B: saturated air with a few stones / water cells near chunk corners and faces
    grid = np.full((rs[0], rs[1], rs[2], 16, 16, 16), 0x00FFFFFF, np.uint32)
    stone, water = np.uint32(1 << 24), np.uint32(7 << 24)
    grid[1, 0, 1, 15, 15, 15] = stone
    grid[2, 1, 0, 0, 0, 0] = stone
    grid[0, 0, 2, 8, 15, 0] = stone
    grid[3, 1, 2, 15, 3, 15] = stone
    grid[5, 3, 4, 15, 15, 15] = stone
    grid[1, 1, 1, 15, 8, 7] = water
    grid[2, 1, 1, 0, 9, 7] = water
    grid[0, 1, 0, 5, 0, 5] = water
    chunks = _grid(rs, sp)
    table = np.zeros(rs, np.uint32)
    for i, c in enumerate(chunks):
        table[c[0] % rs[0], c[1] % rs[1], c[2] % rs[2]] = 2 + i
    worlds = []
    for _ in range(3):
        g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_update_rate=n))
        g.write_table(table)
        g.write_pool(2, grid.reshape(n, 4096))
        worlds.append(g)
    a, b, c = worlds  # b follows a through pack / unpack, c through the push into a peer replica
    a.replica_peers_add([c])
    st = torch.cuda.Stream()
    for g in worlds:
        g.set_stream(st.cuda_stream)
    lst = scenes.update_list(chunks)
    a.update_list_upload(lst)
    shard = SlabShard(rs, sp, 0, 1)
    with torch.cuda.stream(st):
        ref = DeviceReplicaRefresh(torch, None, a, b, shard, sp, sp, capacity=n)
        small = DeviceReplicaRefresh(torch, None, a, b, shard, sp, sp, capacity=2)
        pushed = DeviceReplicaRefresh(torch, None, a, c, shard, sp, sp, capacity=n, push=True)
        moved = 0
        for t in range(24):
            off = scenes.PHASES[t % 8]
            a.update_list_run(sp, off)
            if t == 0:  # an undersized buffer: the record count tells the caller that it overflowed
                small.refresh(off)
                st.synchronize()
                assert int(small.counts()[0].item()) > 2
            ref.refresh(off)
            pushed.refresh(off)
            st.synchronize()
            assert int(pushed.counts()[0].item()) == int(ref.counts()[0].item())
            flags = torch.from_numpy(a.update_list_flags().astype(np.int32)).cuda()
            want, _ = changed_chunk_mask(torch, flags, rs[0], rs[1], rs[2], off, False)
            assert torch.equal(ref.mask.bool(), want), "mask differs at tick %d" % t
            moved += int(ref.counts()[0].item())
            assert np.array_equal(b.read_pool(0, 2 + n), a.read_pool(0, 2 + n)), "replica differs after tick %d" % t
            assert np.array_equal(c.read_pool(0, 2 + n), a.read_pool(0, 2 + n)), "pushed replica differs after tick %d" % t
        assert 0 < moved < 24 * n
    for g in worlds:
        g.close()

    # ghost column: a slab of 3 own columns + 1 ghost, x-shifted phase; kernel mask / ghost mask against the tensor rule
    shard2 = SlabShard(rs, sp, 0, 2)
    lrs, lsp = shard2.local_render_size, shard2.local_start_pos
    nl = lrs[0] * lrs[1] * lrs[2]
    s = vg.VoxelComputePipeline(vg.GameSettings(render_size=lrs, max_loaded_chunks=2 + nl, max_update_rate=nl))
    lchunks = _grid(lrs, lsp)
    ltable = np.zeros(lrs, np.uint32)
    for i, c in enumerate(lchunks):
        ltable[c[0] % lrs[0], c[1] % lrs[1], c[2] % lrs[2]] = 2 + i
    s.write_table(ltable)
    s.write_pool(2, scenes.random_pool(2 + nl, 5, water=0.1, air=0.7))
    s.update_list_upload(scenes.update_list(lchunks[:shard2.n_own]))
    mask = torch.zeros(shard2.n_own, dtype=torch.uint8, device="cuda")
    ghost = torch.zeros(shard2.plane_chunks, dtype=torch.uint8, device="cuda")
    for off in scenes.PHASES:
        s.update_list_run(lsp, off)
        s.changed_chunk_mask(shard2.cols, rs[1], rs[2], off, mask.data_ptr(), ghost.data_ptr())
        s.synchronize()
        flags = torch.from_numpy(s.update_list_flags().astype(np.int32)).cuda()
        want, want_ghost = changed_chunk_mask(torch, flags, shard2.cols, rs[1], rs[2], off, True)
        assert torch.equal(mask.bool(), want) and torch.equal(ghost.bool(), want_ghost), off
    s.close()
