"""GPU parity: every kernel, called through the C ABI (libvoxelb200.so via ctypes), against the CPU
oracle on the same seeded inputs. Bit-exact for voxel words, tables, flags, hit voxels, layer
counts and the collision set; floats are compared bit-exact too where both sides use the same
operation order (expected), with the north_star tolerance 1e-4 relative as the stated bar."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import MAT, CpuWorld, chunk_loads_for

pytestmark = pytest.mark.gpu

RTOL = 1e-4  # north_star: "within a stated tolerance (e.g. 1e-4 relative) for hit distances and normals"


def make_pair(rs, sp, n_slots, world_gen=0, seed=0, **kw):
    import voxel_game_b200 as vg

    gs = vg.GameSettings(render_size=rs, max_loaded_chunks=n_slots, max_worldgen_rate=8 * n_slots, max_update_rate=max(n_slots, 1),
                         world_gen=vg.WorldGen(world_gen), world_gen_seed=seed, **kw)
    gpu = vg.VoxelComputePipeline(gs)
    cpu = CpuWorld(rs, sp, n_slots, dt=gs.delta_time, seed=seed)
    return gpu, cpu


def sync_world_to_gpu(gpu, cpu):
    gpu.write_table(cpu.table)
    gpu.write_pool(2, cpu.pool[2:])


def assert_world_equal(gpu, cpu):
    t = gpu.read_table()
    assert np.array_equal(t, cpu.table)
    pool = gpu.read_pool(0, cpu.pool.shape[0])
    bad = np.argwhere(pool != cpu.pool)
    assert bad.size == 0, "first mismatch slot %d idx %d: gpu %08x cpu %08x (%d mismatches)" % (
        bad[0][0], bad[0][1], pool[tuple(bad[0])], cpu.pool[tuple(bad[0])], len(bad))


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("seed", [1, 2])
def test_voxel_update_random(variant, seed):
    rs, sp = (3, 2, 3), (100, 50, 70)
    n_slots = 2 + 18
    gpu, cpu = make_pair(rs, sp, n_slots)
    table, chunks, _ = scenes.full_table(rs, sp, holes=0.15, seed=seed)
    cpu.table[:] = table
    cpu.pool[2:] = scenes.random_pool(n_slots, seed)
    sync_world_to_gpu(gpu, cpu)
    gpu.set_update_kernel(variant)
    lst = scenes.update_list(chunks)
    for tick, off in enumerate(scenes.PHASES * 2):
        f_gpu = gpu.dispatch_voxel_update(sp, lst, off)
        f_cpu = cpu.voxel_update(lst, off)
        assert np.array_equal(f_gpu, f_cpu), "flags differ at tick %d phase %s" % (tick, off)
        assert_world_equal(gpu, cpu)
    gpu.close()


@pytest.mark.parametrize("variant", [0 | (3 << 8), 2 | (3 << 8), 2 | (5 << 8), 0 | (2 << 8)])
def test_voxel_update_z_runs(variant):
    """Thread blocks that visit runs of consecutive list entries hand the z == 16 voxels of one entry over to
    the next when that is its +z neighbour (offset.z == 1). Forced run lengths 2 / 4 / 16 on a world that is
    long in z; the list is z-fastest for its first part and shuffled for the rest (links and non-links)."""
    rs, sp = (2, 2, 9), (200, 30, 41)
    n_slots = 2 + 36
    gpu, cpu = make_pair(rs, sp, n_slots)
    table, chunks, _ = scenes.full_table(rs, sp, holes=0.12, seed=11)
    cpu.table[:] = table
    cpu.pool[2:] = scenes.random_pool(n_slots, 11)
    sync_world_to_gpu(gpu, cpu)
    gpu.set_update_kernel(variant)
    chunks = np.asarray(chunks)
    order = np.lexsort((chunks[:, 2], chunks[:, 1], chunks[:, 0]))  # z fastest
    chunks = chunks[order]
    rng = np.random.RandomState(3)
    cut = 2 * len(chunks) // 3
    chunks = np.concatenate([chunks[:cut], chunks[cut:][rng.permutation(len(chunks) - cut)]])
    lst = scenes.update_list(chunks)
    for tick, off in enumerate(scenes.PHASES * 2):
        f_gpu = gpu.dispatch_voxel_update(sp, lst, off)
        f_cpu = cpu.voxel_update(lst, off)
        assert np.array_equal(f_gpu, f_cpu), "flags differ at tick %d phase %s" % (tick, off)
        assert_world_equal(gpu, cpu)
    gpu.close()


def test_update_list_run_range_equals_full_sweep():
    """vxb_update_list_run_range over two halves of the resident list (the second on another stream, as the
    sharded step does for its boundary column) gives the world and flags of one full dispatch."""
    import torch

    rs, sp = (3, 2, 4), (100, 50, 70)
    n_slots = 2 + 24
    gpu, cpu = make_pair(rs, sp, n_slots)
    table, chunks, _ = scenes.full_table(rs, sp, holes=0.1, seed=4)
    cpu.table[:] = table
    cpu.pool[2:] = scenes.random_pool(n_slots, 4)
    sync_world_to_gpu(gpu, cpu)
    lst = scenes.update_list(chunks)
    gpu.update_list_upload(lst)
    side = torch.cuda.Stream()
    half = len(lst) // 2
    for off in scenes.PHASES:
        gpu.update_list_run_range(sp, off, 0, half)
        gpu.update_list_run_range(sp, off, half, len(lst) - half, cuda_stream=side.cuda_stream)
        side.synchronize()
        f_gpu = gpu.update_list_flags()
        f_cpu = cpu.voxel_update(lst, off)
        assert np.array_equal(f_gpu, f_cpu), off
        assert_world_equal(gpu, cpu)
    with pytest.raises(Exception):
        gpu.update_list_run_range(sp, (0, 0, 0), half, len(lst))
    gpu.close()


def test_voxel_update_clustered_toroidal():
    # start_pos not a multiple of render_size: the table is addressed modulo render_size
    rs, sp = (4, 3, 2), (497, 106, 501)
    n_slots = 2 + 24
    gpu, cpu = make_pair(rs, sp, n_slots)
    table, chunks, _ = scenes.full_table(rs, sp, holes=0.1, seed=7)
    cpu.table[:] = table
    cpu.pool[2:] = scenes.clustered_pool(n_slots, 7)
    sync_world_to_gpu(gpu, cpu)
    lst = scenes.update_list(chunks)
    total_flags = 0
    for off in scenes.PHASES * 3:
        f_gpu = gpu.dispatch_voxel_update(sp, lst, off)
        f_cpu = cpu.voxel_update(lst, off)
        assert np.array_equal(f_gpu, f_cpu)
        total_flags += int((f_cpu & 1).sum())
    assert total_flags > 0
    assert_world_equal(gpu, cpu)
    gpu.close()


def test_voxel_update_partial_list_and_empty():
    rs, sp = (2, 2, 2), (10, 10, 10)
    gpu, cpu = make_pair(rs, sp, 10)
    table, chunks, _ = scenes.full_table(rs, sp)
    cpu.table[:] = table
    cpu.pool[2:] = scenes.random_pool(10, 3)
    sync_world_to_gpu(gpu, cpu)
    assert gpu.dispatch_voxel_update(sp, np.zeros((0, 4), np.uint32), (1, 1, 1)).size == 0
    lst = scenes.update_list(chunks[1::2])
    for off in [(1, 1, 1), (0, 1, 0), (1, 0, 1)]:
        assert np.array_equal(gpu.dispatch_voxel_update(sp, lst, off), cpu.voxel_update(lst, off))
    assert_world_equal(gpu, cpu)
    gpu.close()


@pytest.mark.parametrize("world_gen,seed,center", [(0, 0, (625, 112, 625)), (1, 3646, (625, 113, 625)), (1, 3645, (630, 112, 622))])
def test_worldgen(world_gen, seed, center):
    rs = (2, 2, 2)
    sp = tuple(c - 1 for c in center)
    gpu, cpu = make_pair(rs, sp, 2 + 8, world_gen=world_gen, seed=seed)
    chunks = [(sp[0] + x, sp[1] + y, sp[2] + z) for x in range(2) for y in range(2) for z in range(2)]
    slots = list(range(2, 10))
    loads = chunk_loads_for(chunks, slots, forced=False)
    r_gpu = gpu.dispatch_worldgen(sp, loads, complete=True)
    r_cpu = cpu.worldgen(loads, world_gen, complete=True)
    assert np.array_equal(r_gpu, r_cpu)
    assert_world_equal(gpu, cpu)
    # forced: negative slot ids keep the slot even for all-air chunks
    loads_f = chunk_loads_for(chunks[:2], slots[:2], forced=True)
    assert np.array_equal(gpu.dispatch_worldgen(sp, loads_f, complete=True), cpu.worldgen(loads_f, world_gen, complete=True))
    assert_world_equal(gpu, cpu)
    gpu.close()


def test_worldgen_scattered_chunks():
    # chunks far apart (different lattice cells, large coordinates): exercises the per-block candidate
    # culling of the open-world kernel against the oracle's full 5x5x5 loops
    rng = np.random.RandomState(12)
    n = 24
    for world_gen, seed in ((0, 0), (1, 77)):
        for k in range(n // 8):
            base = (int(rng.randint(8, 4000)), int(rng.randint(108, 116)), int(rng.randint(8, 4000)))
            if world_gen == 1:
                base = (int(rng.randint(610, 640)), int(rng.randint(110, 114)), int(rng.randint(618, 632)))
            rs = (2, 2, 2)
            gpu, cpu = make_pair(rs, base, 2 + 8, world_gen=world_gen, seed=seed)
            chunks = [(base[0] + x, base[1] + y, base[2] + z) for x in range(2) for y in range(2) for z in range(2)]
            loads = chunk_loads_for(chunks, range(2, 10), forced=True)
            assert np.array_equal(gpu.dispatch_worldgen(base, loads), cpu.worldgen(loads, world_gen))
            assert_world_equal(gpu, cpu)
            gpu.close()


def test_worldgen_sky_chunk_is_all_air():
    sp = (625, 125, 625)  # y = 2000: far above SEA_LEVEL
    gpu, cpu = make_pair((1, 1, 1), sp, 4)
    loads = chunk_loads_for([sp], [2])
    r_gpu = gpu.dispatch_worldgen(sp, loads)
    r_cpu = cpu.worldgen(loads, 0)
    assert r_cpu[0] == 1 and r_gpu[0] == 1  # stays the host's pre-set 1 -> table entry 1
    assert_world_equal(gpu, cpu)
    gpu.close()


def test_write_voxels_and_unload():
    rs, sp = (16, 2, 16), (32, 8, 48)
    n_slots = 2 + 16 * 2 * 16
    gpu, cpu = make_pair(rs, sp, n_slots)
    cpu.load_all_chunks()
    sync_world_to_gpu(gpu, cpu)
    rng = np.random.RandomState(5)
    writes = np.stack([rng.randint(16 * sp[0], 16 * (sp[0] + rs[0]), 300), rng.randint(16 * sp[1], 16 * (sp[1] + rs[1]), 300),
                       rng.randint(16 * sp[2], 16 * (sp[2] + rs[2]), 300), rng.randint(0, 13, 300) << 24], 1).astype(np.uint32)
    writes = writes[np.unique(writes[:, :3], axis=0, return_index=True)[1]]  # distinct targets: order-free
    gpu.dispatch_write_voxels(sp, writes)
    cpu.write_voxels(writes)
    for comp, idx in [(0, sp[0] + 3), (2, sp[2] + 15), (1, sp[1])]:
        gpu.dispatch_unload(idx, comp)
        cpu.unload(idx, comp)
    assert_world_equal(gpu, cpu)
    gpu.close()


def _generated_world(rs, sp, world_gen=0, seed=0, update_ticks=8):
    n_chunks = rs[0] * rs[1] * rs[2]
    gpu, cpu = make_pair(rs, sp, 2 + n_chunks, world_gen=world_gen, seed=seed)
    chunks = [(sp[0] + x, sp[1] + y, sp[2] + z) for x in range(rs[0]) for y in range(rs[1]) for z in range(rs[2])]
    loads = chunk_loads_for(chunks, range(2, 2 + n_chunks))
    gpu.dispatch_worldgen(sp, loads)
    table = gpu.read_table()
    loaded = [c for c in chunks if table[c[0] % rs[0], c[1] % rs[1], c[2] % rs[2]] >= 2]
    lst = scenes.update_list(loaded)
    for off in (scenes.PHASES * ((update_ticks + 7) // 8))[:update_ticks]:
        gpu.dispatch_voxel_update(sp, lst, off, want_flags=False)
    cpu.table[:] = gpu.read_table()
    cpu.pool[:] = gpu.read_pool(0, 2 + n_chunks)
    return gpu, cpu


def _surface_y(cpu, x, z, y_hi):
    y = y_hi
    while y > 16 * cpu.sp[1] and (cpu.get_voxel(x, y, z) >> 24) in (MAT["AIR"], MAT["AIR_OOB"]):
        y -= 1
    return y


def _compare_layers(g, c, n):
    """g: product download ([5,3,n] SoA), c: oracle ([n,5,...] AoS)."""
    assert np.array_equal(g["layer_count"], c["layer_count"])
    lc = c["layer_count"]
    for l in range(5):
        m = lc > l
        if not m.any():
            continue
        assert np.array_equal(g["voxel"][l][m], c["voxel"][m, l])
        assert np.array_equal(g["leaving"][l][m], c["leaving"][m, l])
        np.testing.assert_allclose(g["ldist"][l][m], c["ldist"][m, l], rtol=RTOL, atol=0)
        for k in range(3):
            np.testing.assert_allclose(g["pos"][l, k][m], c["pos"][m, l, k], rtol=RTOL, atol=0)
            np.testing.assert_allclose(g["normal"][l, k][m], c["normal"][m, l, k], rtol=RTOL, atol=1e-7)
        # hit voxel index floor(pos) is bit-exact
        assert np.array_equal(np.floor(g["pos"][l][:, m]), np.floor(c["pos"][m, l].T))
    np.testing.assert_allclose(g["dist"], c["dist"], rtol=RTOL, atol=0)


@pytest.mark.parametrize("world_gen,seed,yaw,pitch", [(0, 0, 0.3, -0.2), (0, 0, 2.5, 0.1), (1, 3646, 1.2, -0.35)])
def test_raymarch_primary_image(world_gen, seed, yaw, pitch):
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820 if world_gen == 0 else 1816, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, world_gen, seed, update_ticks=16)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    cam = scenes.camera((center[0] + 0.5, sy + 4.3, center[2] + 0.5), yaw, pitch)
    w, h = 160, 90
    projectiles = scenes.random_projectiles(6, (center[0] - 8, sy + 2, center[2] - 8), (center[0] + 8, sy + 8, center[2] + 8), 11)
    gpu.dispatch_entities(sp, projectiles=projectiles, players=cam, run_projectiles=False, run_players=False)  # upload only
    caps = vg.RayCaps(100, 40, 20, 15, 6)
    gpu.raymarch(sp, cam, w, h, caps)
    g = gpu.ray_download()
    c = cpu.raycast_image(cam, w, h, 100, projectiles=projectiles)
    assert g["total_iterations"] == c["total_iterations"]
    _compare_layers(g, c, w * h)
    assert (c["layer_count"] > 1).any()  # the scene exercises multi-layer rays or projectile layers
    gpu.close()


def test_raymarch_depth_clip_and_views():
    import voxel_game_b200 as vg

    rs = (4, 3, 4)
    center = (10000, 1812, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=8)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    cams = np.concatenate([scenes.camera((center[0] + 0.5, sy + 3.1, center[2] + 0.5), k * np.pi / 2, -0.25) for k in range(3)])
    w, h = 64, 48
    rng = np.random.RandomState(3)
    depth = np.where(rng.rand(3, h, w) < 0.3, rng.rand(3, h, w) * 20 + 0.5, 0).astype(np.float32)
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    gpu.raymarch(sp, cams, w, h, caps, raster_depth=depth)
    g = gpu.ray_download()
    for v in range(3):
        c = cpu.raycast_image(cams[v:v + 1], w, h, 100, raster_depth=depth[v])
        sl = slice(v * w * h, (v + 1) * w * h)
        gv = {k: (a[..., sl] if isinstance(a, np.ndarray) else a) for k, a in g.items()}
        _compare_layers(gv, c, w * h)
    gpu.close()


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_ray_batches(kind):
    import voxel_game_b200 as vg

    rs = (4, 3, 4)
    center = (10000, 1812, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=8)
    rng = np.random.RandomState(kind)
    n = 5000
    lo = np.array([16 * sp[0] + 2, 16 * sp[1] + 2, 16 * sp[2] + 2], np.float32)
    hi = np.array([16 * (sp[0] + rs[0]) - 2, 16 * (sp[1] + rs[1]) - 2, 16 * (sp[2] + rs[2]) - 2], np.float32)
    pos = (lo + rng.rand(n, 3) * (hi - lo)).astype(np.float32)
    d = rng.randn(n, 3).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:50, 0] = 0.0  # axis-parallel rays: division by zero paths
    d = d.astype(np.float32)
    projectiles = scenes.random_projectiles(5, lo + 20, hi - 20, 4)
    cam = scenes.camera(center, 0, 0)
    gpu.dispatch_entities(sp, projectiles=projectiles, players=cam, run_projectiles=False, run_players=False)
    iters = {0: 100, 1: 20, 2: 15, 3: 40}[kind]
    gpu.raymarch_rays(sp, pos, d, kind, iters, projectile_count=5)
    g = gpu.ray_download()
    if kind == 0:
        c = cpu.raycast_batch(pos, d, iters, projectiles=projectiles)
        assert g["total_iterations"] == c["total_iterations"]
        _compare_layers(g, c, n)
    else:
        c = cpu.secondary_batch(kind, pos, d, iters, projectiles=projectiles)
        assert np.array_equal(g["voxel"][0], c["voxel"])
        np.testing.assert_allclose(g["ldist"][0], c["dist"], rtol=RTOL, atol=0)
        np.testing.assert_allclose(g["pos"][0].T, c["pos"], rtol=RTOL, atol=0)
        np.testing.assert_allclose(g["normal"][0].T, c["normal"], rtol=RTOL, atol=1e-7)
        if kind >= 2:
            np.testing.assert_allclose(g["dist"], c["light"], rtol=RTOL, atol=0)
    gpu.close()


def test_projectiles_and_players():
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=8)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    players = scenes.players_grid(4, center, 9.0, sy + 5.0, seed=2)
    # grounded players, a no-collision player, a falling one
    players["collision_vec"][::3] = (0, 1, 0)
    players["has_world_collision"][5] = 0
    players["vel"][7, :3] = (3.0, -25.0, 1.0)
    lo = (16 * sp[0] + 8, sy - 6, 16 * sp[2] + 8)
    hi = (16 * (sp[0] + rs[0]) - 8, sy + 14, 16 * (sp[2] + rs[2]) - 8)
    pr = scenes.random_projectiles(4000, lo, hi, seed=9)
    # some projectiles right on top of players so that hitsphere + LOS paths run
    for k in range(0, 400):
        pr["pos"][k, :3] = players["pos"][k % 16, :3] + np.random.RandomState(k).randn(3).astype(np.float32) * 1.5 + (0, -1.5, 0)
    pr["should_collide_with_terrain"][::7] = 0
    g_pr, g_pl, g_co, g_n = gpu.dispatch_entities(sp, projectiles=pr, players=players)
    # oracle: projectiles first, then players, on the same world (dispatch order of voxel_sim_manager.rs:944-957)
    c_pr, c_co, counters = cpu.projectiles(pr, players, capacity=1024)
    c_pl = cpu.entities(players)
    assert g_n == len(c_co) and g_n > 0
    assert np.array_equal(g_co[["id1", "id2", "properties"]], c_co[["id1", "id2", "properties"]])
    np.testing.assert_allclose(g_co["when"], c_co["when"], rtol=RTOL, atol=1e-6)
    for f in ("chunk_update_pos", "health", "wall_bounce", "owner"):
        assert np.array_equal(g_pr[f], c_pr[f]), f
    np.testing.assert_allclose(g_pr["pos"], c_pr["pos"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(g_pr["dir"], c_pr["dir"], rtol=RTOL, atol=1e-7)
    assert (c_pr["health"] == 0).any() and (c_pr["dir"] != pr["dir"]).any()  # kills and bounces happened
    assert np.array_equal(g_pl["collision_vec"], c_pl["collision_vec"])
    np.testing.assert_allclose(g_pl["pos"], c_pl["pos"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(g_pl["vel"], c_pl["vel"], rtol=RTOL, atol=1e-7)
    # voxel damage written by projectiles (imageAtomicAdd): world state equal
    assert_world_equal(gpu, cpu)
    gpu.close()


def test_projectile_lattice_edge_cases():
    """The terrain lattice of compute_projectile.glsl:181-243 at its corners: lattices that straddle the
    window border or lie outside it, sit in all-air chunks, span chunk corners, have zero / negative /
    NaN extents, and one too large for the warp-cooperative search (falls back to the per-thread loop)."""
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=4)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    players = scenes.players_grid(2, center, 9.0, sy + 5.0, seed=2)
    wlo = np.array([16 * sp[0], 16 * sp[1], 16 * sp[2]], np.float32)
    whi = np.array([16 * (sp[0] + rs[0]), 16 * (sp[1] + rs[1]), 16 * (sp[2] + rs[2])], np.float32)
    n = 3000
    pr = scenes.random_projectiles(n, wlo - 6, whi + 6, seed=77)  # in, across and outside the window
    rng = np.random.RandomState(5)
    # chunk corners: lattices that touch 8 chunks
    k = np.arange(0, 600)
    pr["pos"][k, :3] = (np.round(pr["pos"][k, :3] / 16) * 16 + rng.randn(600, 3) * 0.4).astype(np.float32)
    # the sky (all-air chunks) and just above the surface
    pr["pos"][600:900, 1] = whi[1] - 1 - rng.rand(300).astype(np.float32) * 30
    pr["pos"][900:1200, 1] = sy + rng.rand(300).astype(np.float32) * 8 - 2
    pr["pos"][900:1200, 0] = center[0] + rng.rand(300).astype(np.float32) * 40 - 20
    pr["pos"][900:1200, 2] = center[2] + rng.rand(300).astype(np.float32) * 40 - 20
    # degenerate extents
    pr["size"][1200:1230, 0] = 0.0
    pr["size"][1230:1260, 1] = -0.5
    pr["size"][1260:1290, 2] = -3.0
    pr["size"][1290:1300, 0] = np.nan
    pr["vel"][1300:1320] = -400.0
    pr["pos"][1320:1330, 0] = np.nan
    pr["dir"][1330:1360] *= 1.7  # non-unit quaternions: no player culling, sheared lattice
    # long, fast boxes (many batches), and one lattice beyond 2^20 samples
    pr["size"][1360:1400, :3] = (3.0, 2.5, 6.0)
    pr["vel"][1360:1400] = 300.0
    pr["size"][1400, :3] = (38.0, 38.0, 38.0)
    pr["pos"][1400, :3] = ((whi[0] + wlo[0]) / 2, whi[1] - 20, (whi[2] + wlo[2]) / 2)
    pr["pos"][1401:1500, :3] = players["pos"][rng.randint(0, 4, 99), :3] + rng.randn(99, 3).astype(np.float32) * 1.2 + (0, -1.5, 0)
    g_pr, _, g_co, g_n = gpu.dispatch_entities(sp, projectiles=pr, players=players, run_players=False)
    c_pr, c_co, _ = cpu.projectiles(pr, players, capacity=1024)
    assert g_n == len(c_co)
    order_g = np.lexsort((g_co["id2"], g_co["id1"]))
    order_c = np.lexsort((c_co["id2"], c_co["id1"]))
    assert np.array_equal(g_co[["id1", "id2", "properties"]][order_g], c_co[["id1", "id2", "properties"]][order_c])
    for f in ("chunk_update_pos", "wall_bounce", "owner"):
        assert np.array_equal(g_pr[f], c_pr[f]), f
    assert np.array_equal(g_pr["health"], c_pr["health"], equal_nan=True)
    np.testing.assert_allclose(g_pr["pos"], c_pr["pos"], rtol=RTOL, atol=0, equal_nan=True)
    np.testing.assert_allclose(g_pr["dir"], c_pr["dir"], rtol=RTOL, atol=1e-7, equal_nan=True)
    assert (c_pr["health"] == 0).sum() > 200 and (c_pr["chunk_update_pos"][:, 3] == 1).sum() > 100
    assert_world_equal(gpu, cpu)
    gpu.close()


def test_sharded_collision_batch_damage_log():
    """Two replicas of one world each run half of a projectile batch with the voxel damage log on, then
    replay the other's log (one through a host array, one straight from the device log): both pools must
    equal a single context's run of the whole batch and the oracle's."""
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu_a, cpu = _generated_world(rs, sp, 0, 0, update_ticks=4)
    gpu_b, _ = _generated_world(rs, sp, 0, 0, update_ticks=4)
    gpu_s, _ = _generated_world(rs, sp, 0, 0, update_ticks=4)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    players = scenes.players_grid(2, center, 9.0, sy + 5.0, seed=2)
    lo = (16 * sp[0] + 8, sy - 8, 16 * sp[2] + 8)
    hi = (16 * (sp[0] + rs[0]) - 8, sy + 6, 16 * (sp[2] + rs[2]) - 8)
    pr = scenes.random_projectiles(4000, lo, hi, seed=21, bounce_frac=0.05)
    pr["damage"] = np.array([0, 1, 5, 7], np.float32)[np.arange(len(pr)) % 4]
    half = len(pr) // 2
    for g in (gpu_a, gpu_b):
        g.damage_log_enable(half)
    pa, _, ca, na = gpu_a.dispatch_entities(sp, projectiles=pr[:half], players=players, run_players=False)
    pb, _, cb, nb = gpu_b.dispatch_entities(sp, projectiles=pr[half:], players=players, run_players=False)
    rec_a = gpu_a.damage_log_read(half)
    rec_b = gpu_b.damage_log_read(half)
    assert len(rec_a) > 50 and len(rec_b) > 50
    gpu_a.damage_apply(sp, records=rec_b)
    ptr_a, cnt_a, cap = gpu_a.damage_log_device()
    assert cap == half
    gpu_b.damage_apply_counted(sp, ptr_a, cnt_a, cap)  # count read on the device, as the multi-GPU exchange does
    gpu_b.synchronize()
    ps, _, cs, ns = gpu_s.dispatch_entities(sp, projectiles=pr, players=players, run_players=False)
    c_pr, c_co, _ = cpu.projectiles(pr, players, capacity=1024)
    assert_world_equal(gpu_s, cpu)
    assert_world_equal(gpu_a, cpu)
    assert_world_equal(gpu_b, cpu)
    both = np.concatenate([pa, pb])
    for f in ("chunk_update_pos", "health", "wall_bounce"):
        assert np.array_equal(both[f], ps[f]) and np.array_equal(both[f], c_pr[f]), f
    cb = cb.copy()
    cb["id1"] += half  # the second rank's projectile indices are local to its range
    merged = np.concatenate([ca, cb])
    merged = merged[np.lexsort((merged["id2"], merged["id1"]))]
    assert na + nb == ns == len(c_co)
    assert np.array_equal(merged[["id1", "id2", "properties"]], cs[["id1", "id2", "properties"]])
    # an overflowing log is an error, not a silent truncation
    gpu_a.damage_log_enable(4)
    gpu_a.dispatch_entities(sp, projectiles=pr[:half], players=players, run_players=False)
    with pytest.raises(vg.VxbError):
        gpu_a.damage_log_read(4)
    for g in (gpu_a, gpu_b, gpu_s):
        g.close()


@pytest.mark.parametrize("devices,inbox", [((0, 0, 0), True), ((0, 0, 0), False), ((0, 1, 0), True)])
def test_sharded_collision_batch_peer_replicas(devices, inbox):
    """Three replicas of one world (on one GPU, or spread over two when the box has them) each run a third of a
    projectile batch knowing the other two as peers. With inboxes k_projectiles stores its damage records into its region
    of each peer's inbox and every replica replays what it received; without, it adds straight into the peers' pools.
    No log, no all-gather — and every pool equals a single context's run of the whole batch and the oracle's."""
    import torch

    import voxel_game_b200 as vg

    if max(devices) >= torch.cuda.device_count():
        pytest.skip("needs two GPUs")
    rs = (6, 4, 6)
    center = (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs)
    single, cpu = _generated_world(rs, sp, 0, 0, update_ticks=4)
    n_slots = cpu.pool.shape[0]
    reps = []
    for d in devices:  # copies of the same world (table and pool) on the devices asked for
        g = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=n_slots, max_projectiles=4096, max_collisions=4096), device=d)
        g.write_table(cpu.table)
        g.write_pool(0, cpu.pool)
        if inbox:
            g.replica_inbox_enable(3, 2048)
        reps.append(g)
    for k, g in enumerate(reps):
        g.replica_peers_add([o for o in reps if o is not g], source_index=k)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    players = scenes.players_grid(2, center, 9.0, sy + 5.0, seed=2)
    lo = (16 * sp[0] + 8, sy - 8, 16 * sp[2] + 8)
    hi = (16 * (sp[0] + rs[0]) - 8, sy + 6, 16 * (sp[2] + rs[2]) - 8)
    ref_pool, ticks = None, (2 if inbox else 1)  # two ticks: the inbox parity flips
    for tick in range(ticks):
        pr = scenes.random_projectiles(3999, lo, hi, seed=21 + tick, bounce_frac=0.05)
        pr["damage"] = np.array([0, 1, 5, 7], np.float32)[np.arange(len(pr)) % 4]
        third = len(pr) // 3
        outs = [g.dispatch_entities(sp, projectiles=pr[k * third:(k + 1) * third], players=players, run_players=False) for k, g in enumerate(reps)]
        for g in reps:
            g.synchronize()  # the barrier: every replica's dispatch has finished
        if inbox:
            for k, g in enumerate(reps):
                counts = g.replica_inbox_counts()
                assert counts[k] == 0 and (np.delete(counts, k) > 20).all() and counts.max() <= 2048
                g.replica_inbox_apply(sp)
        ps, _, cs, ns = single.dispatch_entities(sp, projectiles=pr, players=players, run_players=False)
        c_pr, c_co, _ = cpu.projectiles(pr, players, capacity=4096)
        assert (c_pr["chunk_update_pos"][:, 3] == 1).sum() > 100  # damage really landed
        ref_pool = single.read_pool(0, n_slots)
        assert np.array_equal(ref_pool, cpu.pool)
        for g in reps:
            assert np.array_equal(g.read_pool(0, n_slots), ref_pool), "tick %d" % tick
        both = np.concatenate([o[0] for o in outs])
        assert np.array_equal(both.view(np.uint8), ps.view(np.uint8))
        assert sum(o[3] for o in outs) == ns == len(c_co)
    # peers can be dropped again: a dispatch then only touches its own world
    reps[0].replica_peers_clear()
    reps[0].dispatch_entities(sp, projectiles=pr[:third], players=players, run_players=False)
    assert np.array_equal(reps[1].read_pool(0, n_slots), ref_pool) and not np.array_equal(reps[0].read_pool(0, n_slots), ref_pool)
    for g in reps + [single]:
        g.close()


def test_replica_api_rejects_misuse():
    """Error behaviour of the peer-replica entry points: clear messages, nothing half-registered."""
    import voxel_game_b200 as vg

    gs = vg.GameSettings(render_size=(2, 2, 2), max_loaded_chunks=10)
    a, b = vg.VoxelComputePipeline(gs), vg.VoxelComputePipeline(gs)
    with pytest.raises(vg.VxbError, match="no inbox"):
        a.replica_inbox_apply((0, 0, 0))
    with pytest.raises(vg.VxbError, match="1..16 sources"):
        a.replica_inbox_enable(0, 16)
    with pytest.raises(vg.VxbError, match="1..16 sources"):
        a.replica_inbox_enable(17, 16)
    b.replica_inbox_enable(2, 16)
    with pytest.raises(vg.VxbError, match="source_index beyond"):
        a.replica_peers_add([b], source_index=2)
    with pytest.raises(vg.VxbError, match="NULL or the ctx itself"):
        a.replica_peers_add([a], source_index=0)
    a.replica_peers_add([b], source_index=1)
    with pytest.raises(vg.VxbError, match="one source_index per ctx"):
        a.replica_peers_add([b], source_index=0)
    with pytest.raises(vg.VxbError, match="no projectile dispatch yet"):
        b.replica_inbox_apply((0, 0, 0))
    with pytest.raises(vg.VxbError, match="no peer replicas"):
        b.update_list_upload(np.zeros((1, 4), np.uint32))
        b.update_list_push((0, 0, 0), 1, (0, 0, 0), 1, 4)
    h = a.replica_export()
    assert h.shape == (224,) and list(h[192:204].view(np.uint32)) == [2, 2, 2] and not h[128:192].any()  # no inbox: third handle zero
    assert list(b.replica_export()[204:212].view(np.uint32)) == [2, 16]
    a.replica_peers_clear()
    a.replica_peers_add([b] * 15, source_index=0)
    with pytest.raises(vg.VxbError, match="at most 15 peers"):
        a.replica_peers_add([b], source_index=0)
    a.close()
    b.close()


@pytest.mark.parametrize("world_gen,seed,yaw,pitch,t", [(0, 0, 0.3, -0.2, 0.0), (0, 0, 2.5, 0.15, 3.25), (1, 3646, 1.2, -0.35, 1.5)])
def test_render_full_fragment_shader(world_gen, seed, yaw, pitch, t):
    """frag.glsl main = primary DDA + get_color. One transcendental (depth_transparency ^ dist) and long
    float chains: colours are compared at the north-star tolerance, not bit-exactly."""
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820 if world_gen == 0 else 1816, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, world_gen, seed, update_ticks=16)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    cam = scenes.camera((center[0] + 0.5, sy + 4.3, center[2] + 0.5), yaw, pitch)
    w, h = 128, 72
    projectiles = scenes.random_projectiles(4, (center[0] - 8, sy + 2, center[2] - 8), (center[0] + 8, sy + 8, center[2] + 8), 21)
    gpu.dispatch_entities(sp, projectiles=projectiles, players=cam, run_projectiles=False, run_players=False)
    caps = vg.RayCaps(100, 40, 20, 15, 4)
    game = vg.Game(gpu, vg.GameState(start_pos=list(sp)), cam)
    vg.PointLightingSystem().draw((w, h), None, None, None, None, game, vg.GraphicsSettings(), rendered_projectile_count=4, time=t)
    g = gpu.color_download()[0]
    it_wave = gpu.ray_download()["total_iterations"]
    # the default is a wavefront of small kernels (shade_wave.cuh); the fused per-pixel kernel must give the same bits and
    # walk the same number of DDA iterations (primary + secondary)
    gpu.set_render_kernel(1)
    vg.PointLightingSystem().draw((w, h), None, None, None, None, game, vg.GraphicsSettings(), rendered_projectile_count=4, time=t)
    fused = gpu.color_download()[0]
    assert np.array_equal(g.view(np.uint32), fused.view(np.uint32)) and it_wave == gpu.ray_download()["total_iterations"] > 20 * w * h
    gpu.set_render_kernel(0)
    c = cpu.render_image(cam, w, h, caps=(100, 40, 20, 15), time=t, projectiles=projectiles)
    assert np.isfinite(c).all() and g.shape == c.shape
    close = np.isclose(g, c, rtol=RTOL, atol=1e-5).all(axis=2)
    bad = np.argwhere(~close)
    report = "; ".join("(%d,%d) gpu %s cpu %s" % (y, x, np.round(g[y, x, :3], 5).tolist(), np.round(c[y, x, :3], 5).tolist()) for y, x in bad[:6])
    # a SECONDARY ray that grazes a voxel face can flip a branch under a 1-ulp difference of the pow: allow 0.1 % of pixels ...
    assert close.mean() > 0.999, "only %.4f of the pixels agree: %s" % (close.mean(), report)
    # ... but nothing else: the primary DDA of the same frame (layers, hit voxels, projectile boxes) is exact on EVERY pixel, so a
    # pixel outside the tolerance can only come from the shading of identical layers
    gpu.raymarch(sp, cam, w, h, caps)
    gr = gpu.ray_download()
    cr = cpu.raycast_image(cam, w, h, 100, projectiles=projectiles)
    assert np.array_equal(gr["layer_count"], cr["layer_count"]) and gr["total_iterations"] == cr["total_iterations"]
    for l in range(5):
        m = cr["layer_count"] > l
        assert np.array_equal(gr["voxel"][l][m], cr["voxel"][m, l]), "layer %d" % l
    assert c[..., :3].std() > 0.02  # a real picture, not a constant
    gpu.close()


def _raster_scene(w, h, yaw_cam=0.3):
    """Camera matrices the way main.rs builds them (:473-481, :614-617) and a ring of players around it
    (shared with the CPU golden test: tests/raster_scene.py)."""
    from tests.raster_scene import raster_scene

    return raster_scene(yaw_cam)


@pytest.mark.parametrize("w,h", [(160, 90), (97, 131)])
def test_raster_gbuffer_matches_the_oracle(w, h):
    """RasterizerSystem::draw as CUDA kernels against oracle/raster.h: depth, normals (through float16), diffuse
    and the ray pass's max_depth (frag.glsl:716-728), bit for bit, on a capsule mesh instanced around the camera."""
    import voxel_game_b200 as vg
    from tests import oracle_lib

    gpu, _ = make_pair((2, 2, 2), (100, 50, 70), 2 + 8)
    pos, nrm = scenes.capsule_mesh()
    rs = vg.RasterizerSystem(gpu, pos, nrm)
    cam_pos, view, s2w, players, facing, health = _raster_scene(w, h)
    inst = rs.instances(players, facing, health)
    assert inst.size == 12  # player 0 (the camera) and the dead one are skipped
    rs.draw((w, h), view, players, facing, health)
    gpu.raster_max_depth(s2w, w / h, cam_pos)
    g = gpu.raster_download()
    c = oracle_lib.raster_draw(w, h, view, pos, nrm, inst)
    covered = c["depth"] < 1.0
    assert 0.02 < covered.mean() < 0.9, covered.mean()
    assert np.array_equal(g["depth"].view(np.uint32), c["depth"].view(np.uint32))
    assert np.array_equal(g["diffuse_x"], c["diffuse_x"])
    assert np.array_equal(g["normal"].view(np.uint32), c["normal"].view(np.uint32))
    cm = oracle_lib.raster_max_depth(c["depth"], s2w, w / h, cam_pos)
    assert np.array_equal(g["max_depth"].view(np.uint32), cm.view(np.uint32))
    assert (cm[covered] > 0).all() and (cm[~covered] == 0).all()
    # an empty draw leaves the clear values
    rs.draw((w, h), view, players[:1])
    e = gpu.raster_download(want_max_depth=False)
    assert (e["depth"] == 1.0).all() and (e["normal"] == 0).all() and (e["diffuse_x"] == 0).all()
    gpu.close()


def test_ray_pass_reads_the_gbuffer():
    """frag.glsl:428-437, 725-728: with the G-buffer bound, the primary ray stops at the rasterised depth and the
    extra layer carries the rasterised normal and MAT_PLAYER; layers equal the oracle's fed with the same G-buffer."""
    import voxel_game_b200 as vg
    from tests import oracle_lib

    rs_, center = (6, 4, 6), (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs_)
    gpu, cpu = _generated_world(rs_, sp, 0, 0, update_ticks=8)
    w, h = 128, 72
    pos, nrm = scenes.capsule_mesh()
    rast = vg.RasterizerSystem(gpu, pos, nrm)
    cam_pos, view, s2w, players, facing, health = _raster_scene(w, h)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs_[1]) - 1)
    shift = np.array([0, sy + 6 - cam_pos[1], 0], np.float32)
    players["pos"][:, :3] += shift
    cam_pos = cam_pos + shift
    cam = scenes.camera(tuple(cam_pos), 0.3, -0.1)
    view = np.linalg.inv((vg.translation(-cam_pos) @ vg.quat_to_mat4((0, np.sin(0.15), 0, np.cos(0.15)))).astype(np.float64)).astype(np.float32)
    rast.draw((w, h), view, players, facing, health)
    gpu.raster_max_depth(s2w, w / h, cam_pos)
    gpu.raster_bind(True)
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    gpu.raymarch(sp, cam, w, h, caps)
    g = gpu.ray_download()
    gb = gpu.raster_download()
    c = cpu.raycast_image(cam, w, h, 100, raster_depth=gb["max_depth"], raster_normal=gb["normal"], raster_diffuse_x=gb["diffuse_x"])
    _compare_layers(g, c, w * h)
    player_layers = (c["voxel"] >> 24) == 8
    assert player_layers.any()
    # and the shaded frame
    gpu.render(sp, cam, w, h, caps, time=0.5)
    col = gpu.color_download().reshape(h, w, 4)
    ref = cpu.render_image(cam, w, h, (100, 40, 20, 15), 0.5, raster_depth=gb["max_depth"], raster_normal=gb["normal"], raster_diffuse_x=gb["diffuse_x"])
    close = np.isclose(col, ref, rtol=1e-4, atol=1e-4).all(axis=2)
    assert close.mean() > 0.999, close.mean()
    gpu.raster_bind(False)
    gpu.close()


def test_resident_cameras_give_the_same_frames():
    """vxb_cams_upload + vxb_raymarch_resident (no host traffic per frame) against vxb_raymarch with host cameras."""
    import voxel_game_b200 as vg

    rs_, center = (6, 4, 6), (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs_)
    gpu, cpu = _generated_world(rs_, sp, 0, 0, update_ticks=8)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs_[1]) - 1)
    cams = np.concatenate([scenes.camera((center[0] + 0.5, sy + 4.5, center[2] + 0.5), k * 0.9, -0.2) for k in range(4)])
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    w, h = 96, 54
    gpu.raymarch(sp, cams[1:3], w, h, caps)
    a = gpu.ray_download()
    gpu.cams_upload(cams)
    gpu.raymarch_resident(sp, 1, 2, w, h, caps)
    b = gpu.ray_download()
    assert a["total_iterations"] == b["total_iterations"]
    for k in ("layer_count", "voxel", "leaving"):
        assert np.array_equal(a[k], b[k]), k
    for k in ("pos", "normal", "ldist", "dist"):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), k
    with pytest.raises(vg.VxbError):
        gpu.raymarch_resident(sp, 3, 2, w, h, caps)  # beyond the uploaded cameras
    gpu.raymarch(sp, cams[:1], w, h, caps)           # replaces the resident cameras
    with pytest.raises(vg.VxbError):
        gpu.raymarch_resident(sp, 0, 1, w, h, caps)
    gpu.close()
