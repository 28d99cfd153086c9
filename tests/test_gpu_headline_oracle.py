"""GPU: the oracle on BASELINE's own configurations (not on small stand-ins).
 - configs[2]: a 4-column slab of the 256x16x256-chunk world, generated as bench.py generates it (forced worldgen at C3_SP), walked
   through the benchmark's 96 settle ticks and 8 more on the GPU and on the oracle: change flags every tick, the world at three points;
   and a 16-column slab (65 536 chunks, the list length from which the launcher claims runs) while it is still changing everywhere;
 - configs[1]: one whole 1920x1080 primary-ray frame of the 1024x256x1024-voxel world against the oracle's raycast on the
   downloaded world: layers, hit voxels, iteration total exact, distances to 1e-4;
 - configs[3]: the first 65 536 projectiles of the benchmark's batch + the 64 hulls against the oracle on the same world."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import CpuWorld

pytestmark = pytest.mark.gpu
RTOL = 1e-4  # north_star: hit distances and normals within 1e-4 relative


@pytest.mark.timeout(900)
@pytest.mark.parametrize("first_col", [0, 126])
def test_config3_slab_walks_with_the_oracle(first_col):
    import torch

    import bench
    import voxel_game_b200 as vg

    rs = (4, bench.C3_RS[1], bench.C3_RS[2])
    sp = (bench.C3_SP[0] + first_col, bench.C3_SP[1], bench.C3_SP[2])
    n = rs[0] * rs[1] * rs[2]
    gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    bench.generate_world(gpu, sp, chunks, 2, True, torch)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)
    gpu.update_list_upload(lst)
    cpu = CpuWorld(rs, sp, 2 + n)
    cpu.table[:] = gpu.read_table()
    cpu.pool[:] = gpu.read_pool(0, 2 + n)
    assert (cpu.table >= 2).all()  # forced: every chunk owns a slot
    changed = 0
    for t in range(96 + 8):
        off = scenes.PHASES[t % 8]
        gpu.update_list_run(sp, off)
        g_flags = gpu.update_list_flags()
        c_flags = cpu.voxel_update(lst, off)
        assert np.array_equal(g_flags, c_flags), "flags differ at tick %d" % t
        changed += int((c_flags & 1).sum())
        if t in (7, 95, 103):
            assert np.array_equal(gpu.read_pool(0, 2 + n), cpu.pool), "world differs after tick %d" % t
    assert changed > n  # the slab really evolved
    gpu.close()


@pytest.mark.timeout(900)
def test_config3_sixteen_columns_with_the_launchers_run_lengths():
    """65 536 chunks of the config-3 world: long enough for the launcher to pick what it picks at full size (claims of two entries
    with stores elided; runs of eight with the z hand-over for unconditional stores on z-shifted phases), walked 16 ticks on the GPU
    alone (distance fields still building: most chunks change), then 8 ticks of each store policy against the oracle: flags every
    tick, the world after each policy."""
    import torch

    import bench
    import voxel_game_b200 as vg

    rs = (16, bench.C3_RS[1], bench.C3_RS[2])
    sp = (bench.C3_SP[0] + 64, bench.C3_SP[1], bench.C3_SP[2])
    n = rs[0] * rs[1] * rs[2]
    gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
    chunks = bench.chunk_grid(0, rs[0], rs, sp)
    bench.generate_world(gpu, sp, chunks, 2, True, torch)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)
    gpu.update_list_upload(lst)
    for t in range(16):
        gpu.update_list_run(sp, scenes.PHASES[t % 8])
    cpu = CpuWorld(rs, sp, 2 + n)
    cpu.table[:] = gpu.read_table()
    cpu.pool[:] = gpu.read_pool(0, 2 + n)
    changed = 0
    for variant in (2, 0):
        gpu.set_update_kernel(variant)
        for t in range(8):
            off = scenes.PHASES[t]
            gpu.update_list_run(sp, off)
            c_flags = cpu.voxel_update(lst, off)
            assert np.array_equal(gpu.update_list_flags(), c_flags), "flags differ, variant %d phase %d" % (variant, t)
            changed += int((c_flags & 1).sum())
        assert np.array_equal(gpu.read_pool(0, 2 + n), cpu.pool), "world differs after variant %d" % variant
    assert changed > n  # the world is still building its distance fields: the stores and the hand-over carry real changes
    gpu.close()


@pytest.mark.timeout(900)
def test_config2_full_frame_against_the_oracle():
    import torch

    import voxel_game_b200 as vg
    from tests.test_gpu_scale import _config2_world

    g, rs, sp, (cx, sy, cz) = _config2_world(torch)
    n = rs[0] * rs[1] * rs[2]
    cpu = CpuWorld(rs, sp, 2 + n)
    cpu.table[:] = g.read_table()
    cpu.pool[:] = g.read_pool(0, 2 + n)
    w, h = 1920, 1080
    for yaw in (0.7, 3.9):
        cam = scenes.camera((cx + 0.5, sy + 4.5, cz + 0.5), yaw, -0.2)
        g.raymarch(sp, cam, w, h, vg.RayCaps(100, 40, 20, 15, 0))
        a = g.ray_download()
        c = cpu.raycast_image(cam, w, h, 100)
        assert a["total_iterations"] == c["total_iterations"] > 5 * w * h
        assert np.array_equal(a["layer_count"], c["layer_count"]) and (c["layer_count"] > 0).mean() > 0.5
        np.testing.assert_allclose(a["dist"], c["dist"], rtol=RTOL, atol=0)
        for l in range(5):
            m = c["layer_count"] > l
            if not m.any():
                continue
            assert np.array_equal(a["voxel"][l][m], c["voxel"][m, l]), l
            assert np.array_equal(a["leaving"][l][m], c["leaving"][m, l]), l
            np.testing.assert_allclose(a["ldist"][l][m], c["ldist"][m, l], rtol=RTOL, atol=0)
            np.testing.assert_allclose(a["pos"][l][:, m].T, c["pos"][m, l], rtol=RTOL, atol=0)
            np.testing.assert_allclose(a["normal"][l][:, m].T, c["normal"][m, l], rtol=RTOL, atol=1e-7)
    g.close()


@pytest.mark.timeout(900)
def test_config4_projectile_slice_against_the_oracle():
    import torch

    from tests.test_gpu_scale import _config2_world

    g, rs, sp, (cx, sy, cz) = _config2_world(torch)
    n = rs[0] * rs[1] * rs[2]
    cpu = CpuWorld(rs, sp, 2 + n)
    cpu.table[:] = g.read_table()
    cpu.pool[:] = g.read_pool(0, 2 + n)
    lo = (16 * sp[0] + 8, 16 * sp[1] + 8, 16 * sp[2] + 8)
    hi = (16 * (sp[0] + rs[0]) - 8, 16 * (sp[1] + rs[1]) - 8, 16 * (sp[2] + rs[2]) - 8)
    pr = scenes.c4_projectiles(65536, lo, hi)  # bench.py's batch, first 64 k
    pl = scenes.players_grid(8, (cx, sy, cz), 48.0, sy + 4.0, seed=64)
    pr["pos"][:2000, :3] = pl["pos"][np.arange(2000) % 64, :3] + np.random.RandomState(3).randn(2000, 3).astype(np.float32) * 1.5 + (0, -1.5, 0)
    g_pr, g_pl, g_co, g_n = g.dispatch_entities(sp, projectiles=pr, players=pl)
    c_pr, c_co, counters = cpu.projectiles(pr, pl, capacity=1 << 16)
    c_pl = cpu.entities(pl)
    assert g_n == len(c_co) and g_n > 50
    assert np.array_equal(g_co[["id1", "id2", "properties"]], c_co[["id1", "id2", "properties"]])
    np.testing.assert_allclose(g_co["when"], c_co["when"], rtol=RTOL, atol=1e-6)
    for f in ("chunk_update_pos", "health", "wall_bounce", "owner"):
        assert np.array_equal(g_pr[f], c_pr[f]), f
    np.testing.assert_allclose(g_pr["pos"], c_pr["pos"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(g_pr["dir"], c_pr["dir"], rtol=RTOL, atol=1e-7)
    assert (c_pr["health"] == 0).sum() > 1000 and int(counters[1]) > 20 * len(pr)  # terrain hits; the lattice really ran
    assert np.array_equal(g_pl["collision_vec"], c_pl["collision_vec"])
    np.testing.assert_allclose(g_pl["pos"], c_pl["pos"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(g_pl["vel"], c_pl["vel"], rtol=RTOL, atol=1e-7)
    assert np.array_equal(g.read_pool(0, 2 + n), cpu.pool)  # voxel damage landed on the same voxels
    g.close()
